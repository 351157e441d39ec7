// arm_f64_n2.cu -- double instantiation for Planar Push with 2 props (RobotPushConfig.n_objects, robot_planar_push.py:61).
#include "arm_launch.cuh"

int arm_launch_f64_n2(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count) {
  return launch<double, 2>(a, a->md, which, io, st, first, count);
}

void arm_forget_f64_n2(ArmSim* a) {
  for (int d = 0; d < 64; d++) if (g_const_owner[d] == (const void*)&a->md) g_const_owner[d] = nullptr;
}
