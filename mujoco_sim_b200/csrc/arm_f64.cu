// arm_f64.cu -- double instantiation of the articulated engine's kernels for 3 props (Planar Push).
#include "arm_launch.cuh"

int arm_launch_f64_n0(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count);
void arm_forget_f64_n0(ArmSim* a);
int arm_launch_f64_n1(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count);
int arm_launch_f64_n2(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count);
void arm_forget_f64_n1(ArmSim* a);
void arm_forget_f64_n2(ArmSim* a);

int arm_launch_f64(ArmSim* a, int which, ArmIO io, cudaStream_t st, int first, int count) {
  switch (a->nobj) {
    case 0: return arm_launch_f64_n0(a, which, io, st, first, count);
    case 1: return arm_launch_f64_n1(a, which, io, st, first, count);
    case 2: return arm_launch_f64_n2(a, which, io, st, first, count);
    case 3: return launch<double, 3>(a, a->md, which, io, st, first, count);
    default: mjs_set_error("arm engine: built for 0 to 3 props, got %d", a->nobj); return -1;
  }
}

void arm_forget_f64(ArmSim* a) {
  for (int d = 0; d < 64; d++) if (g_const_owner[d] == (const void*)&a->md) g_const_owner[d] = nullptr;
  arm_forget_f64_n0(a);
  arm_forget_f64_n1(a);
  arm_forget_f64_n2(a);
}
