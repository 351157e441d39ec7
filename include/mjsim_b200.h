/* mjsim_b200.h -- C ABI of libmjsim_b200.so, the drop-in boundary.
 *
 * The reference has no FFI for this path: its hot path is the Python class
 *   DMCEnvironmentAdapter            /root/reference/mujoco_sim/environments/dmc2gym.py:78-168
 * calling dm_control.composer.Environment.step/reset, which calls MuJoCo's C
 * engine.  Each entry point below names the reference interface it replaces;
 * INTEGRATION.md shows the ctypes stub a maintainer would add to dmc2gym.py.
 *
 * Conventions: plain pointers and sizes; no torch types.  "d_" pointers are
 * CUDA device pointers, "h_" pointers are host pointers.  All functions return
 * 0 on success and a negative value on error (mjs_last_error() has the text).
 * `stream` is a cudaStream_t passed as void* (NULL = default stream); no call
 * synchronises the host unless it takes host buffers.
 */
#ifndef MJSIM_B200_H_
#define MJSIM_B200_H_

#include "mjs_model.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct MjsSim MjsSim;

/* Precision of the arithmetic inside the step kernels. */
#define MJS_PRECISION_F64 0 /* all-double (parity runs)                        */
#define MJS_PRECISION_MIXED 1 /* FP64 state integration, FP32 dynamics/solver  */

/* Replaces: task construction + `Environment(task, ...)` + MJCF compilation
 * (mujoco_sim/__init__.py:19-23).  The flattened model/task are copied.
 * env_offset = global id of this shard's env 0 (multi-GPU: rank * num_envs). */
int mjs_create(const MjsModel* model, const MjsTask* task, int num_envs, int device, unsigned long long seed,
               unsigned int env_offset, int precision, MjsSim** out);
void mjs_destroy(MjsSim* sim);

/* Replaces: DMCEnvironmentAdapter.reset (dmc2gym.py:157-163) for the envs whose
 * d_mask byte is non-zero (NULL = all).  Samples the task's initial-state boxes
 * with Philox keyed by (seed, global env id, episode id).  d_obs may be NULL. */
int mjs_reset(MjsSim* sim, const unsigned char* d_mask, float* d_obs, void* stream);

/* Replaces: DMCEnvironmentAdapter.seed (dmc2gym.py:126-131), which installs RandomState(seed) on a LIVE env so that the
 * next reset draws from it (reset(seed=s) calls it first, dmc2gym.py:157-160).  Re-keys the Philox streams, sets every env's
 * episode counter and the action-stream index back to 0 and discards the next-episode slots prepared under the old key: two
 * resets after mjs_reseed(sim, s) with the same s produce identical episodes (test/test_gym_envs.py:21-36).  Synchronises. */
int mjs_reseed(MjsSim* sim, unsigned long long seed);

/* Replaces: the float64 observation / reward arrays the reference returns (dm_env specs are float64: dmc2gym.py:66-75,
 * reward is a Python float).  Registers caller-owned DEVICE buffers obs64 [N][O], reward64 [N], final_obs64 [N][O] (any may
 * be NULL; all NULL = off) that every later mjs_reset / mjs_step* fills with the unrounded double values next to the float32
 * outputs.  The buffers must stay valid until they are replaced or the simulator is destroyed. */
int mjs_set_output64(MjsSim* sim, double* d_obs64, double* d_reward64, double* d_final_obs64);

/* Replaces: DMCEnvironmentAdapter.step (dmc2gym.py:133-155) for all envs.
 * d_actions [N][A] float32 (NULL: synthetic rollout, actions ~ U(action box)
 * drawn on device), d_obs [N][O], d_reward [N], d_flags [N][4] bytes
 * (terminated, truncated, is_success, done), d_final_obs [N][O] or NULL: the
 * observation of the terminal state when autoreset replaces it in d_obs. */
int mjs_step(MjsSim* sim, const float* d_actions, float* d_obs, float* d_reward, unsigned char* d_flags,
             float* d_final_obs, int autoreset, void* stream);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments and results as mjs_step,
 * but the launch's scratch (Planar Push's settle-job list) lives in a file-scope static __device__ array of
 * MJS_NOALLOC_MAX_JOBS = 16384 entries (66 KB, held by every GPU that loads the library until exit) instead of memory
 * allocated for the simulator; the call runs no cudaMalloc / cudaFree.  Limit: num_envs * 3 / 100 + 1 <= 16384
 * (num_envs <= 546 100); larger simulators get a clear error (-1).  The array is shared by all simulators and calls on a
 * device: calls that overlap on the GPU (different streams, no ordering) race on it.  Tasks whose step uses no scratch
 * (PointMass, Reach, Button Push, autoreset == 0) are forwarded to mjs_step, which allocates nothing either. */
int mjs_step_noalloc(MjsSim* sim, const float* d_actions, float* d_obs, float* d_reward, unsigned char* d_flags,
                     float* d_final_obs, int autoreset, void* stream);

/* Same call with HOST buffers: copies actions H2D, steps, copies results D2H and
 * synchronises.  This is the end-to-end call the reference-side stub uses. */
int mjs_step_host(MjsSim* sim, const float* h_actions, float* h_obs, float* h_reward, unsigned char* h_flags,
                  float* h_final_obs, int autoreset);
int mjs_reset_host(MjsSim* sim, const unsigned char* h_mask, float* h_obs);

/* Replaces: direct writes to physics.data.qpos/qvel/mocap_pos/ctrl through
 * physics.bind()/named.data (e.g. pointmass.py:100-130, robot.py:185-189).
 * AoS MjsEnvState records on the host, envs [first, first+count).
 * mjs_set_state also re-runs the reset-time mj_forward (warm start) when
 * `forward` is non-zero. */
int mjs_set_state(MjsSim* sim, const MjsEnvState* h_states, int first, int count, int forward);
int mjs_get_state(MjsSim* sim, MjsEnvState* h_states, int first, int count);

/* Episode statistics accumulated by the step kernel since the last call
 * (SURVEY.md 8e): {return sum, length sum, episodes, successes, truncations,
 * errors, physics substeps, Newton iterations}.  d_out may be fed to an NCCL all-reduce directly. */
int mjs_stats_device(MjsSim* sim, double** d_out);
int mjs_stats(MjsSim* sim, double* h_out8, int reset);

/* Per-env status words (MJS_STATUS_* bits of include/mjs_model.h: non-finite state, IK unreachable -- the reference's
 * ValueError("IK failed"), robots/robot.py:221-224 --, contact overflow, unsupported geom pair) copied into the caller's
 * device buffer d_status[num_envs] on `stream`; no host synchronisation.  Python exceptions of the reference become these
 * bits: device code never throws. */
int mjs_status(MjsSim* sim, int* d_status, void* stream);

/* Planar Push prepares the next episode of finished envs (sampling + 150 settle
 * steps) on an internal low-priority stream.  mjs_join_aux makes `stream` wait for
 * whatever of that work is in flight (used to close a timed region). */
int mjs_join_aux(MjsSim* sim, void* stream);

/* Test hook: mj_forward for one env of an articulated task, dumping intermediate
 * stages (layout documented at arm_debug_kernel). */
int mjs_debug_forward(MjsSim* sim, int env, double* h_out, int n);

/* Test hook: continue the current control step of every env by k physics substeps (no before_step
 * hook, no outputs), so tests can split a control step at a substep boundary. */
int mjs_debug_substeps(MjsSim* sim, int k);
/* The task's scripted demonstrator evaluated on the current device state of every env: actions[N][A] float32 (device),
 * ok[N] int32 (device, may be NULL; 0 where Button Push's IK found no solution and the joints are held).
 * Replaces create_demonstration_policy of point_reach.py:227-240, robot_reach.py:222-253 (noise = noise_level) and
 * robot_push_button.py:231-300; Planar Push's demonstrator is keyboard tele-operation and returns an error. */
int mjs_demo_policy(MjsSim* sim, double noise, float* d_actions, int* d_ok, void* stream);
/* Test/profiling hook: the first call arms per-env timers in the step kernel; later calls copy, per env, 16 words: the 4 words
 * {start ns, end of control step ns, end ns, newton_iterations << 32 | max_iterations << 16 | smid} of the last launch. */
int mjs_debug_times(MjsSim* sim, long long* h_out, int n);

/* Replaces: Camera.get_rgb_image -> physics.render(height, width, camera_id) (entities/camera.py:94-103; the RGB observables
 * of point_reach.py:22,95-98, robot_reach.py:52,116-118, robot_planar_push.py:42-43, robot_push_button.py:87-96).  Ray-casts
 * the current state of every env from a fixed camera: d_rgb [N][height][width][3] uint8, channel-last like the reference.
 * The visible primitives are given by the caller (model geoms by index, markers that follow the env's target / switch);
 * shading model and limits are stated in csrc/render.cuh.  MuJoCo's OpenGL output is unavailable on this box and is not
 * reproduced bit for bit.  Allocates its pose scratch on the first call. */
int mjs_render(MjsSim* sim, const MjsCamera* cam, const MjsVisual* vis, int nvis, int width, int height, unsigned char* d_rgb, void* stream);

/* Number of kernels this library launched since creation (bench `gpu_launches`). */
long long mjs_launch_count(const MjsSim* sim);
int mjs_num_envs(const MjsSim* sim);
int mjs_obs_dim(const MjsSim* sim);
int mjs_action_dim(const MjsSim* sim);
const char* mjs_last_error(void);
const char* mjs_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MJSIM_B200_H_ */
