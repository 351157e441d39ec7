// arm_f32_n2.cu -- float instantiation for Planar Push with 2 props (RobotPushConfig.n_objects, robot_planar_push.py:61).
#include "arm_launch.cuh"

int arm_launch_f32_n2(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count) {
  return launch<float, 2>(a, a->mf, which, io, st, first, count);
}

void arm_forget_f32_n2(ArmSim* a) {
  for (int d = 0; d < 64; d++) if (g_const_owner[d] == (const void*)&a->mf) g_const_owner[d] = nullptr;
}
