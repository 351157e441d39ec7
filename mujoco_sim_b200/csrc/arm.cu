// arm.cu -- host side of the articulated (warp-per-env) engine.
#include <cstdlib>
#include <vector>

#include "arm_sim.cuh"

template <int NOBJ>
static void layout_sizes(int& sd, int& si) { sd = ArmLayout<NOBJ>::SD; si = ArmLayout<NOBJ>::SI; }

static int dispatch(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first = 0, int count = 0) {
  return a->precision == 0 ? arm_launch_f64(a, which, io, st, first, count) : arm_launch_f32(a, which, io, st, first, count);
}

ArmSim* arm_create(const MjsModel& m, const MjsTask& t, int N, unsigned long long seed, unsigned int env_offset, int precision, double* d_stats) {
  ArmSim* a = new ArmSim();
  a->N = N; a->seed = seed; a->env_offset = env_offset; a->precision = precision; a->d_stats = d_stats;
  if (build_arm_model<float>(m, t, 0, a->mf) || build_arm_model<double>(m, t, 1, a->md)) { delete a; return nullptr; }
  a->nobj = a->mf.nobj;
  if (a->mf.npair + ARM_NLINK > 32) { mjs_set_error("arm engine: npair + 6 limit lanes exceed one warp"); delete a; return nullptr; }
  if (a->nobj == 0) layout_sizes<0>(a->SD, a->SI);
  else if (a->nobj == 1) layout_sizes<1>(a->SD, a->SI);
  else if (a->nobj == 2) layout_sizes<2>(a->SD, a->SI);
  else if (a->nobj == 3) layout_sizes<3>(a->SD, a->SI);
  else { mjs_set_error("arm engine: built for 0 to 3 props, got %d", a->nobj); delete a; return nullptr; }
  if (cudaMalloc(&a->sd, sizeof(double) * (size_t)a->SD * N) != cudaSuccess || cudaMalloc(&a->si, sizeof(int) * (size_t)a->SI * N) != cudaSuccess) {
    mjs_set_error("arm engine: cudaMalloc failed"); delete a; return nullptr;
  }
  cudaMalloc(&a->d_debug, sizeof(double) * 1024);
  if (cudaMalloc(&a->queue, 64) != cudaSuccess) { mjs_set_error("arm engine: cudaMalloc (work queue) failed"); delete a; return nullptr; }
  if (a->mf.settle_steps > 0) {
    if (cudaMalloc(&a->slot_sd, sizeof(double) * (size_t)a->SD * N) != cudaSuccess || cudaMalloc(&a->slot_si, sizeof(int) * (size_t)a->SI * N) != cudaSuccess ||
        cudaMalloc(&a->slot_obs, sizeof(double) * 16 * (size_t)N) != cudaSuccess || cudaMalloc(&a->slot_ready, sizeof(int) * (size_t)N) != cudaSuccess) {
      mjs_set_error("arm engine: cudaMalloc (reset slots) failed"); delete a; return nullptr;
    }
    cudaMemset(a->slot_sd, 0, sizeof(double) * (size_t)a->SD * N);
    cudaMemset(a->slot_si, 0, sizeof(int) * (size_t)a->SI * N);
    cudaMemset(a->slot_ready, 0, sizeof(int) * (size_t)N);
    // settle jobs in flight ~ N * (settle_steps / n_sub) / episode length (1.2 % of N for Planar Push): room for 3 %
    // (rounded up to whole blocks at launch); a backlog only delays slots, which are needed an episode later
    a->job_cap = N * 3 / 100 + 1;
    if (cudaMalloc(&a->slot_prog, sizeof(int) * (size_t)N) != cudaSuccess || cudaMalloc(&a->jobs, sizeof(int) * ((size_t)a->job_cap + 64 + 2)) != cudaSuccess) {
      mjs_set_error("arm engine: cudaMalloc (settle jobs) failed"); delete a; return nullptr;
    }
    cudaMemset(a->slot_prog, 0, sizeof(int) * (size_t)N);
    cudaMemset(a->jobs, 0, sizeof(int) * ((size_t)a->job_cap + 64 + 2));
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    cudaStreamCreateWithPriority(&a->side, cudaStreamNonBlocking, lo);  // lowest priority: fills gaps
    cudaEventCreateWithFlags(&a->ev_step, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&a->ev_refill, cudaEventDisableTiming);
  }
  cudaMemset(a->sd, 0, sizeof(double) * (size_t)a->SD * N);
  cudaMemset(a->si, 0, sizeof(int) * (size_t)a->SI * N);
  return a;
}

void arm_destroy(ArmSim* a) {
  if (!a) return;
  arm_forget_f32(a); arm_forget_f64(a);
  cudaFree(a->sd); cudaFree(a->si); cudaFree(a->d_debug); cudaFree(a->d_times); cudaFree(a->queue);
  if (a->slot_ready) {
    cudaStreamSynchronize(a->side);
    cudaFree(a->slot_sd); cudaFree(a->slot_si); cudaFree(a->slot_obs); cudaFree(a->slot_ready); cudaFree(a->slot_prog); cudaFree(a->jobs);
    cudaStreamDestroy(a->side); cudaEventDestroy(a->ev_step); cudaEventDestroy(a->ev_refill);
  }
  delete a;
}

static ArmIO make_io(ArmSim* a) {
  ArmIO io;
  memset(&io, 0, sizeof(io));
  // Lock-step of the step kernel's warp groups (shared instruction fetch): 0 none; 1 barrier at substep start; 2 + before the
  // solve and a group vote per Newton iteration; 3 + every stage; +8: levels 2/3 without the vote (the data-dependent Newton
  // loop runs free), +16 (P - 1): the substep-start barrier every P-th substep.  Default measured on B200
  // (profiles/r1/sync_modes.txt, exp9): level 1 for both precisions.
  { const char* lv = getenv("MJS_SYNC_LEVEL"); io.sync_level = lv ? atoi(lv) : 1; }
  io.sd = a->sd; io.si = a->si; io.stats = a->d_stats;
  io.obs64 = a->obs64; io.reward64 = a->reward64; io.final_obs64 = a->final_obs64;
  io.slot_sd = a->slot_sd; io.slot_si = a->slot_si; io.slot_obs = a->slot_obs; io.slot_ready = a->slot_ready; io.slot_prog = a->slot_prog; io.times = a->d_times; io.N = a->N; io.seed = a->seed; io.env_offset = a->env_offset;
  return io;
}

// Launch the refill kernel on the low-priority side stream once the work already
// queued on `st` is done, unless the previous refill is still running.
static int maybe_refill(ArmSim* a, cudaStream_t st) {
  if (!a->slot_ready) return 0;
  if (a->refill_pending) {
    cudaError_t q = cudaEventQuery(a->ev_refill);
    if (q == cudaErrorNotReady) return 0;
    if (q != cudaSuccess) { mjs_set_error("refill stream: %s", cudaGetErrorString(q)); return -1; }
    a->refill_pending = 0;
  }
  MJS_CUDA_CHECK(cudaEventRecord(a->ev_step, st));
  MJS_CUDA_CHECK(cudaStreamWaitEvent(a->side, a->ev_step, 0));
  ArmIO io = make_io(a);
  if (dispatch(a, 1, io, a->side, 1, 0) < 0) return -1;
  MJS_CUDA_CHECK(cudaEventRecord(a->ev_refill, a->side));
  a->refill_pending = 1;
  a->extra_launches++;
  return 1;
}

int arm_reset(ArmSim* a, const unsigned char* d_mask, float* d_obs, cudaStream_t st) {
  ArmIO io = make_io(a);
  io.mask = d_mask; io.obs = d_obs;
  int r = dispatch(a, 1, io, st, 0, 0);
  if (r < 0) return r;
  int q = maybe_refill(a, st);
  return q < 0 ? q : r + q;
}

int arm_step(ArmSim* a, const float* d_actions, float* d_obs, float* d_reward, unsigned char* d_flags, float* d_final_obs, int autoreset,
             unsigned int step_index, cudaStream_t st) {
  ArmIO io = make_io(a);
  io.actions = d_actions; io.obs = d_obs; io.reward = d_reward; io.flags = d_flags; io.final_obs = d_final_obs;
  io.autoreset = autoreset; io.step_index = step_index;
  if (!autoreset) { io.slot_ready = nullptr; }
  static const bool side_refill = getenv("MJS_SIDE_REFILL") && atoi(getenv("MJS_SIDE_REFILL"));  // A/B: the co-running refill kernel
  if (io.slot_ready && side_refill) {
    int r = dispatch(a, 0, io, st);
    if (r < 0) return r;
    int q = maybe_refill(a, st);
    return q < 0 ? q : r + q;
  }
  if (io.slot_ready) {
    // a refill kernel started by mjs_reset must be done before job warps touch the slots
    if (a->refill_pending) { MJS_CUDA_CHECK(cudaStreamWaitEvent(st, a->ev_refill, 0)); a->refill_pending = 0; }
    int* counters = a->jobs + a->job_cap + 64;  // the launch rounds job_cap up to whole blocks (< 64 warps)
    io.jobs = a->jobs; io.job_cap = a->job_cap;
    io.njobs = counters + (a->step_launches & 1); io.njobs_next = counters + ((a->step_launches + 1) & 1);
    a->step_launches++;
  }
  // Work distribution of the step kernel: MJS_QUEUE=1 -> persistent grid, warps pull envs from an atomic counter
  // (no lock-step: sync_level 0); MJS_QUEUE=0 -> one block per 12/16 envs with lock-step stage barriers.
  // Default per precision (profiles/r2/sync_queue_sweep.txt): f64 -> queue (2.32 M vs 2.20 M env-steps/s), FP32 -> lock-step blocks
  // of 16 warps (2.94 M vs 1.60 M: the FP32 kernel's 16 free-running warps do not share instruction fetches).
  static const int queue_env = getenv("MJS_QUEUE") ? atoi(getenv("MJS_QUEUE")) : -1;
  const int use_queue = queue_env >= 0 ? queue_env : (a->precision == 0 ? MJS_QUEUE_DEFAULT : 0);
  if (use_queue) {
    io.sync_level = 0;
    io.queue = a->queue;
    MJS_CUDA_CHECK(cudaMemsetAsync(a->queue, 0, sizeof(int), st));
  }
  return dispatch(a, 0, io, st);
}

// ---- mjs_step_noalloc: the step launch with its scratch (the settle-job list of Planar Push) in a file-scope device array
// instead of the simulator's cudaMalloc'ed one.  Added for callers that cannot
// allocate (CUDA graph capture, serving loops); the reference has no counterpart.  The array is shared by every simulator
// and every call on a device and keeps the last call's contents: the two counters are cleared on the stream before each
// launch, the list entries are written before they are read.  Calls that overlap on the GPU race on it.
#define MJS_NOALLOC_MAX_JOBS 16384  // list entries; a simulator needs num_envs * 3 / 100 + 1 of them (num_envs <= 546 100)
static __device__ __align__(256) int g_noalloc_jobs[MJS_NOALLOC_MAX_JOBS + 64 + 2];
static int* g_noalloc_ptr[64];  // device address of the array, per device (the symbol's host-side name is not an address)

int arm_step_noalloc(ArmSim* a, const float* d_actions, float* d_obs, float* d_reward, unsigned char* d_flags, float* d_final_obs, int autoreset,
                     unsigned int step_index, cudaStream_t st) {
  if (!(a->slot_ready && autoreset)) return arm_step(a, d_actions, d_obs, d_reward, d_flags, d_final_obs, autoreset, step_index, st);  // no scratch on this path
  if (a->job_cap > MJS_NOALLOC_MAX_JOBS) {
    mjs_set_error("mjs_step_noalloc: %d envs need %d settle-job entries, the static scratch holds %d (num_envs <= %d)", a->N, a->job_cap,
                  MJS_NOALLOC_MAX_JOBS, (MJS_NOALLOC_MAX_JOBS - 1) * 100 / 3);
    return -1;
  }
  int dev = 0;
  MJS_CUDA_CHECK(cudaGetDevice(&dev));
  if (!g_noalloc_ptr[dev & 63]) {  // set-up, once per device, with that device current
    int* p = nullptr;
    MJS_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_noalloc_jobs));
    g_noalloc_ptr[dev & 63] = p;
  }
  int* base = g_noalloc_ptr[dev & 63];
  ArmIO io = make_io(a);
  io.actions = d_actions; io.obs = d_obs; io.reward = d_reward; io.flags = d_flags; io.final_obs = d_final_obs;
  io.autoreset = autoreset; io.step_index = step_index;
  if (a->refill_pending) { MJS_CUDA_CHECK(cudaStreamWaitEvent(st, a->ev_refill, 0)); a->refill_pending = 0; }
  int* counters = base + MJS_NOALLOC_MAX_JOBS + 64;
  MJS_CUDA_CHECK(cudaMemsetAsync(counters, 0, 2 * sizeof(int), st));  // another simulator's call may have left them non-zero
  io.jobs = base; io.job_cap = a->job_cap;
  io.njobs = counters; io.njobs_next = counters + 1;
  return dispatch(a, 0, io, st);
}

int arm_poses(ArmSim* a, float* d_gpose, float* d_aux, cudaStream_t st) {
  ArmIO io = make_io(a);
  io.gpose = d_gpose; io.aux = d_aux;
  return dispatch(a, 6, io, st);
}

// mjs_status: the status column of the env-major integer records, gathered into the caller's buffer on its stream
int arm_status(ArmSim* a, int* d_status, cudaStream_t st) {
  const int I_STATUS = 5;  // ArmLayout<NOBJ>::I_STATUS (same index for every NOBJ)
  static_assert(ArmLayout<0>::I_STATUS == 5 && ArmLayout<3>::I_STATUS == 5, "status column moved");
  MJS_CUDA_CHECK(cudaMemcpy2DAsync(d_status, sizeof(int), a->si + I_STATUS, sizeof(int) * (size_t)a->SI, sizeof(int), (size_t)a->N, cudaMemcpyDeviceToDevice, st));
  return 0;
}

void arm_set_output64(ArmSim* a, double* d_obs64, double* d_reward64, double* d_final_obs64) {
  a->obs64 = d_obs64; a->reward64 = d_reward64; a->final_obs64 = d_final_obs64;
}

// mjs_reseed: new Philox key, episode counters back to 0 (so a reset after reseeding with the same seed repeats the same
// episodes, dmc2gym.py:126-131,157-163), every pre-settled slot of the old key discarded
int arm_reseed(ArmSim* a, unsigned long long seed) {
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  a->seed = seed;
  MJS_CUDA_CHECK(cudaMemset2D(a->si + 6 /* I_EPISODE */, sizeof(int) * a->SI, 0, sizeof(int), a->N));
  if (a->slot_ready) {
    MJS_CUDA_CHECK(cudaMemset(a->slot_ready, 0, sizeof(int) * (size_t)a->N));
    MJS_CUDA_CHECK(cudaMemset(a->slot_prog, 0, sizeof(int) * (size_t)a->N));
    MJS_CUDA_CHECK(cudaMemset(a->slot_si, 0, sizeof(int) * (size_t)a->SI * a->N));
    MJS_CUDA_CHECK(cudaMemset(a->jobs, 0, sizeof(int) * ((size_t)a->job_cap + 64 + 2)));
    a->refill_pending = 0;
  }
  return 0;
}

// make `st` wait for the refill work in flight (bench: close the timed region)
int arm_join_aux(ArmSim* a, cudaStream_t st) {
  if (a->slot_ready && a->refill_pending) MJS_CUDA_CHECK(cudaStreamWaitEvent(st, a->ev_refill, 0));
  return 0;
}

template <int NOBJ>
static void pack(const MjsEnvState& x, double* s, int* si) {
  using L = ArmLayout<NOBJ>;
  for (int k = 0; k < L::NQ; k++) s[L::QPOS + k] = x.qpos[k];
  for (int k = 0; k < L::NV; k++) { s[L::QVEL + k] = x.qvel[k]; s[L::WARM + k] = x.qacc_warm[k]; }
  for (int k = 0; k < 6; k++) { s[L::CTRL + k] = x.ctrl[k]; s[L::Q0 + k] = x.traj_q0[k]; s[L::Q1 + k] = x.traj_q1[k]; }
  s[L::TIME] = x.time; s[L::T0] = x.traj_t0; s[L::RET] = x.ep_return;
  for (int k = 0; k < 3; k++) { s[L::TARGET + k] = x.target[k]; s[L::OVR + k] = x.ovr_pos[k]; }
  si[L::I_STEP] = x.episode_step; si[L::I_TRAJ] = x.traj_active; si[L::I_SWA] = x.sw_active; si[L::I_SWP] = x.sw_pressed;
  si[L::I_SWC] = x.sw_count; si[L::I_STATUS] = x.status; si[L::I_EPISODE] = x.episode_id; si[L::I_LEN] = 0;
}
template <int NOBJ>
static void unpack(MjsEnvState& x, const double* s, const int* si) {
  using L = ArmLayout<NOBJ>;
  memset(&x, 0, sizeof(x));
  for (int k = 0; k < L::NQ; k++) x.qpos[k] = s[L::QPOS + k];
  for (int k = 0; k < L::NV; k++) { x.qvel[k] = s[L::QVEL + k]; x.qacc_warm[k] = s[L::WARM + k]; }
  for (int k = 0; k < 6; k++) { x.ctrl[k] = s[L::CTRL + k]; x.traj_q0[k] = s[L::Q0 + k]; x.traj_q1[k] = s[L::Q1 + k]; }
  x.time = s[L::TIME]; x.traj_t0 = s[L::T0]; x.ep_return = s[L::RET];
  for (int k = 0; k < 3; k++) { x.target[k] = s[L::TARGET + k]; x.ovr_pos[k] = s[L::OVR + k]; }
  x.episode_step = si[L::I_STEP]; x.traj_active = si[L::I_TRAJ]; x.sw_active = si[L::I_SWA]; x.sw_pressed = si[L::I_SWP];
  x.sw_count = si[L::I_SWC]; x.status = si[L::I_STATUS]; x.episode_id = si[L::I_EPISODE];
}

int arm_set_state(ArmSim* a, const MjsEnvState* h, int first, int count, int forward) {
  std::vector<double> sd((size_t)a->SD * count);
  std::vector<int> si((size_t)a->SI * count);
  for (int i = 0; i < count; i++) {
    double* d = sd.data() + (size_t)i * a->SD; int* q = si.data() + (size_t)i * a->SI;
    switch (a->nobj) { case 0: pack<0>(h[i], d, q); break; case 1: pack<1>(h[i], d, q); break; case 2: pack<2>(h[i], d, q); break; default: pack<3>(h[i], d, q); }
  }
  MJS_CUDA_CHECK(cudaMemcpy(a->sd + (size_t)first * a->SD, sd.data(), sd.size() * sizeof(double), cudaMemcpyHostToDevice));
  MJS_CUDA_CHECK(cudaMemcpy(a->si + (size_t)first * a->SI, si.data(), si.size() * sizeof(int), cudaMemcpyHostToDevice));
  if (forward) {
    ArmIO io = make_io(a);
    if (dispatch(a, 2, io, 0, first, count) < 0) return -1;
    MJS_CUDA_CHECK(cudaDeviceSynchronize());
  }
  return 0;
}

int arm_get_state(ArmSim* a, MjsEnvState* h, int first, int count) {
  std::vector<double> sd((size_t)a->SD * count);
  std::vector<int> si((size_t)a->SI * count);
  MJS_CUDA_CHECK(cudaMemcpy(sd.data(), a->sd + (size_t)first * a->SD, sd.size() * sizeof(double), cudaMemcpyDeviceToHost));
  MJS_CUDA_CHECK(cudaMemcpy(si.data(), a->si + (size_t)first * a->SI, si.size() * sizeof(int), cudaMemcpyDeviceToHost));
  for (int i = 0; i < count; i++) {
    const double* d = sd.data() + (size_t)i * a->SD; const int* q = si.data() + (size_t)i * a->SI;
    switch (a->nobj) { case 0: unpack<0>(h[i], d, q); break; case 1: unpack<1>(h[i], d, q); break; case 2: unpack<2>(h[i], d, q); break; default: unpack<3>(h[i], d, q); }
  }
  return 0;
}

int arm_demo_policy(ArmSim* a, double noise, unsigned int step_index, float* d_actions, int* d_ok, cudaStream_t st) {
  if (a->mf.kind == MJS_TASK_PUSH) { mjs_set_error("mjs_demo_policy: Planar Push has no scripted demonstrator (keyboard tele-operation in the reference)"); return -1; }
  ArmIO io = make_io(a);
  io.step_index = step_index;
  a->pol_actions = d_actions; a->pol_ok = d_ok; a->pol_noise = noise;
  return dispatch(a, 5, io, st);
}

// first call arms the per-env timers of the step kernel; later calls read them back
int arm_debug_times(ArmSim* a, long long* h_out, int n) {
  size_t words = MJS_TIMES_WORDS * (size_t)a->N;
  if (!a->d_times) {
    MJS_CUDA_CHECK(cudaMalloc(&a->d_times, words * sizeof(long long)));
    MJS_CUDA_CHECK(cudaMemset(a->d_times, 0, words * sizeof(long long)));
    return 0;
  }
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  MJS_CUDA_CHECK(cudaMemcpy(h_out, a->d_times, sizeof(long long) * ((size_t)n < words ? (size_t)n : words), cudaMemcpyDeviceToHost));
  return 0;
}

int arm_debug_substeps(ArmSim* a, int k) {
  ArmIO io = make_io(a);
  if (dispatch(a, 4, io, 0, 0, k) < 0) return -1;
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  return 0;
}

int arm_debug_forward(ArmSim* a, int env, double* h_out, int n) {
  ArmIO io = make_io(a);
  const char* pre = getenv("MJS_DEBUG_PRE");  // test hook: in-kernel substeps before the dumped forward pass
  if (dispatch(a, 3, io, 0, env, pre ? atoi(pre) : 0) < 0) return -1;
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  MJS_CUDA_CHECK(cudaMemcpy(h_out, a->d_debug, sizeof(double) * (n < 1024 ? n : 1024), cudaMemcpyDeviceToHost));
  return 0;
}
