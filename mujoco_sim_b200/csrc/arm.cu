// arm.cu -- placeholder until the articulated engine lands.
#include "arm_api.cuh"
ArmSim* arm_create(const MjsModel&, const MjsTask&, int, unsigned long long, unsigned int, int, double*) { mjs_set_error("articulated engine not built"); return nullptr; }
void arm_destroy(ArmSim*) {}
int arm_reset(ArmSim*, const unsigned char*, float*, cudaStream_t) { return -1; }
int arm_step(ArmSim*, const float*, float*, float*, unsigned char*, float*, int, unsigned int, cudaStream_t) { return -1; }
int arm_set_state(ArmSim*, const MjsEnvState*, int, int, int) { return -1; }
int arm_get_state(ArmSim*, MjsEnvState*, int, int) { return -1; }
