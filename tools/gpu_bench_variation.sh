q() { nvidia-smi --query-gpu=temperature.gpu,clocks.sm,power.draw,clocks_event_reasons.active --format=csv,noheader; }
B="python bench.py --no-cpu-baseline --no-alt --no-small"
for s in 20 20 100 20 20; do
  q; $B --steps $s --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('steps',d['steps'],'ms',round(d['ms_per_step'],3),'kernel_ms',round(d['roofline']['kernel_ms'],3),'median',round(d['roofline']['kernel_ms_median'],3),d['clocks'])"
done
q
