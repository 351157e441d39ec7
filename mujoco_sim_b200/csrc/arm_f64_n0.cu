// arm_f64_n0.cu -- double instantiation for scenes without props (Robot Reach, Button Push).
#include "arm_launch.cuh"

int arm_launch_f64_n0(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count) {
  return launch<double, 0>(a, a->md, which, io, st, first, count);
}

void arm_forget_f64_n0(ArmSim* a) {
  for (int d = 0; d < 64; d++) if (g_const_owner[d] == (const void*)&a->md) g_const_owner[d] = nullptr;
}
