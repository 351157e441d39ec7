// arm_launch.cuh -- kernel launch code, instantiated once per precision (arm_f32.cu / arm_f64.cu).
#pragma once
#include <cstdlib>
#include "arm_sim.cuh"

// The kernels read the model from the __constant__ block c_arm_model (uniform fields) and its global-memory
// twin g_arm_model (lane-indexed tables).  Both are
// re-uploaded (after a device sync) whenever a different simulator or precision
// launches on this device, so several simulators may coexist in a process.
static const void* g_const_owner[64];
template <typename real>
static int bind_model(const ArmModel<real>& M) {
  int dev = 0;
  MJS_CUDA_CHECK(cudaGetDevice(&dev));
  if (g_const_owner[dev & 63] == (const void*)&M) return 0;
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  MJS_CUDA_CHECK(cudaMemcpyToSymbol(c_arm_model, &M, sizeof(M)));
  MJS_CUDA_CHECK(cudaMemcpyToSymbol(g_arm_model, &M, sizeof(M)));
  g_const_owner[dev & 63] = (const void*)&M;
  return 0;
}

template <typename real, int NOBJ>
static int launch(ArmSim* a, const ArmModel<real>& M, int which, ArmIO io, cudaStream_t st, int first = 0, int count = 0) {
  if (bind_model<real>(M)) return -1;
  size_t smem = sizeof(WarpMem<real, NOBJ>) * ARM_WPB;
  int n = which == 2 ? count : a->N;
  int grid = (n + ARM_WPB - 1) / ARM_WPB;
  if (which == 0) {
    // doubles need twice the shared memory: half the warps per block
    constexpr int WPBS = sizeof(real) == 8 ? ARM_STEP_WPB_F64 : ARM_STEP_WPB;
    constexpr int SG = ArmEngine<real, NOBJ>::SYNC_GROUP;
    static_assert(WPBS % SG == 0 && WPBS / SG <= 15, "lock-step groups must tile the block");
    auto k = arm_step_kernel<real, NOBJ, WPBS, ARM_STEP_SYNC>;
    size_t smem_step = sizeof(WarpMem<real, NOBJ>) * WPBS;
    MJS_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_step));
    // settle jobs ride in the first job_blocks blocks of the same grid (see arm_settle_job)
    io.job_blocks = 0;
    if (io.slot_ready && io.jobs) {
      io.job_blocks = (io.job_cap + WPBS - 1) / WPBS;
      io.job_cap = io.job_blocks * WPBS;
      arm_jobs_kernel<<<(n + 255) / 256, 256, 0, st>>>(io);
    }
    int blocks = (n + WPBS - 1) / WPBS + io.job_blocks;
    if (io.queue) {  // persistent grid: as many blocks as are resident at once (one per SM: shared memory is full)
      static int resident_of[64];  // per device (this function is one template instantiation = one kernel)
      int dev = 0;
      MJS_CUDA_CHECK(cudaGetDevice(&dev));
      int& resident = resident_of[dev & 63];
      if (!resident) {
        int sms = 0, per_sm = 0;
        MJS_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        MJS_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, WPBS * 32, smem_step));
        resident = sms * (per_sm > 0 ? per_sm : 1);
      }
      if (blocks > resident) blocks = resident;
    }
    k<<<blocks, WPBS * 32, smem_step, st>>>(io);
    MJS_CUDA_CHECK(cudaGetLastError());
    return io.job_blocks ? 2 : 1;
  } else if (which == 1 && first == 1) {
    // refill (side stream, low priority): ONE warp per block so a refill block's shared memory fits next to
    // a resident step block instead of taking a whole SM away from the step kernel for 150 substeps
    auto k = arm_reset_kernel<real, NOBJ, 1>;
    size_t smem1 = sizeof(WarpMem<real, NOBJ>);
    MJS_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    k<<<n, 32, smem1, st>>>(io, first);
  } else if (which == 1) {
    auto k = arm_reset_kernel<real, NOBJ, ARM_WPB>;
    MJS_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, ARM_WPB * 32, smem, st>>>(io, first);
  } else if (which == 4) {
    auto k = arm_substeps_kernel<real, NOBJ, ARM_WPB>;
    MJS_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, ARM_WPB * 32, smem, st>>>(io, count);
  } else if (which == 5) {
    auto k = arm_policy_kernel<real, NOBJ, ARM_WPB>;
    MJS_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, ARM_WPB * 32, smem, st>>>(io, a->pol_actions, a->pol_ok, a->pol_noise);
  } else if (which == 6) {
    auto k = arm_pose_kernel<real, NOBJ, ARM_WPB>;
    MJS_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, ARM_WPB * 32, smem, st>>>(io);
  } else if (which == 3) {
    auto k = arm_debug_kernel<real, NOBJ, ARM_WPB>;
    MJS_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<1, ARM_WPB * 32, smem, st>>>(io, first, a->d_debug, count);
  } else {
    auto k = arm_forward_kernel<real, NOBJ, ARM_WPB>;
    MJS_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, ARM_WPB * 32, smem, st>>>(io, first, count);
  }
  MJS_CUDA_CHECK(cudaGetLastError());
  return 1;
}

