// arm_api.cuh -- host interface of the articulated (warp-per-env) engine.
#pragma once
#include "mjs_common.cuh"
struct ArmSim;
ArmSim* arm_create(const MjsModel& m, const MjsTask& t, int N, unsigned long long seed, unsigned int env_offset, int precision, double* d_stats);
void arm_destroy(ArmSim* a);
// each returns the number of kernels launched (>=0) or a negative error
int arm_reset(ArmSim* a, const unsigned char* d_mask, float* d_obs, cudaStream_t st);
int arm_step(ArmSim* a, const float* d_actions, float* d_obs, float* d_reward, unsigned char* d_flags, float* d_final_obs, int autoreset,
             unsigned int step_index, cudaStream_t st);
int arm_set_state(ArmSim* a, const MjsEnvState* h, int first, int count, int forward);
int arm_get_state(ArmSim* a, MjsEnvState* h, int first, int count);
int arm_debug_forward(ArmSim* a, int env, double* h_out, int n);
int arm_join_aux(ArmSim* a, cudaStream_t st);
int arm_debug_substeps(ArmSim* a, int k);
int arm_debug_times(ArmSim* a, long long* h_out, int n);
int arm_demo_policy(ArmSim* a, double noise, unsigned int step_index, float* d_actions, int* d_ok, cudaStream_t st);
int arm_step_noalloc(ArmSim* a, const float* d_actions, float* d_obs, float* d_reward, unsigned char* d_flags, float* d_final_obs, int autoreset,
                     unsigned int step_index, cudaStream_t st);
void arm_set_output64(ArmSim* a, double* d_obs64, double* d_reward64, double* d_final_obs64);
int arm_reseed(ArmSim* a, unsigned long long seed);
int arm_poses(ArmSim* a, float* d_gpose, float* d_aux, cudaStream_t st);
int arm_status(ArmSim* a, int* d_status, cudaStream_t st);
