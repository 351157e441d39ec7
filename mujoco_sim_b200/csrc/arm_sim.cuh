// arm_sim.cuh -- host-side simulator record of the articulated engine, shared by arm.cu (state
// management) and the per-precision launch units arm_f32.cu / arm_f64.cu (compiled in parallel).
#pragma once
#include "arm_api.cuh"
#include "arm_kernel.cuh"

#define ARM_WPB 4  // warps (= envs) per thread block (reset / forward / debug kernels)
// Step kernel: one block of ARM_STEP_WPB warps per SM running the substep stages in
// lock-step (barriers between stages) so the warps share their instruction fetches.
#ifndef ARM_STEP_WPB
#define ARM_STEP_WPB 16
#endif
#ifndef ARM_STEP_WPB_F64
#define ARM_STEP_WPB_F64 12  // 12 x 19.3 KB of FP64 WarpMem fill the 227 KB of one SM; 168 regs x 384 threads fit the RF
#endif
#ifndef MJS_QUEUE_DEFAULT
#define MJS_QUEUE_DEFAULT 1  // persistent grid + atomic work queue: measured faster than the lock-step blocks since the round-2 code shrink (profiles/r2/sync_queue_sweep.txt)
#endif
#ifndef ARM_STEP_SYNC
#define ARM_STEP_SYNC true
#endif

struct ArmSim {
  int N, nobj, precision;
  unsigned long long seed;
  unsigned int env_offset;
  double* d_stats;
  double* sd;
  int* si;
  int SD, SI;
  ArmModel<float> mf;
  ArmModel<double> md;
  double* d_debug;
  int* queue;     // work-item counter of the persistent step grid (queue mode)
  // pre-settled reset slots + side stream (tasks with settle steps)
  double* slot_sd;
  int* slot_si;
  double* slot_obs;
  double *obs64, *reward64, *final_obs64;  // registered by mjs_set_output64 (may be null)
  int* slot_ready;
  int* slot_prog;
  int* jobs;      // job_cap env ids + the two ping-pong counters behind them
  int job_cap;
  unsigned int step_launches;
  long long* d_times;
  float* pol_actions; int* pol_ok; double pol_noise;  // arguments of the pending arm_policy_kernel launch
  cudaStream_t side;
  cudaEvent_t ev_step, ev_refill;
  int refill_pending;
  long long extra_launches;
};


// per-precision launch units: run kernel `which` (0 step, 1 reset/refill, 2 forward, 3 debug, 4 substeps, 5 demo policy);
// return the number of kernels launched or a negative error
int arm_launch_f32(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count);
int arm_launch_f64(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count);
void arm_forget_f32(ArmSim* a);
void arm_forget_f64(ArmSim* a);
