// api.cu -- C ABI of libmjsim_b200.so (include/mjsim_b200.h).
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/mjsim_b200.h"
#include "mjs_common.cuh"
#include "pointmass.cuh"
#include "render.cuh"
#include "arm_api.cuh"

static thread_local char g_err[1024] = "";
void mjs_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

struct MjsSim {
  MjsModel model;
  MjsTask task;
  int N, device, precision;
  unsigned long long seed;
  unsigned int env_offset, step_index;
  long long launches;
  double* d_stats;
  // host-call staging (pinned host + device)
  float *d_act, *d_obs, *d_rew, *d_fobs;
  float *d_gpose, *d_aux;  // mjs_render scratch, allocated at the first call
  double *obs64, *rew64, *fobs64;  // caller-owned float64 outputs registered by mjs_set_output64 (may be null)
  unsigned char *d_flags, *d_mask;
  cudaStream_t hstream;
  // PointMass
  double* pm_sd;
  int* pm_si;
  PmParams<double> pm_pd;
  PmParams<float> pm_pf;
  // articulated tasks
  ArmSim* arm;
};

// ------------------------------------------------------------ host-side helpers
static void h_quat2mat(double* m, const double* q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
static void h_make_frame(double* f) {
  auto nrm = [](double* a) { double n = std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); a[0] /= n; a[1] /= n; a[2] /= n; };
  nrm(f);
  f[3] = f[4] = f[5] = 0;
  if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1;
  double t = f[0] * f[3] + f[1] * f[4] + f[2] * f[5];
  for (int i = 0; i < 3; i++) f[3 + i] -= t * f[i];
  nrm(f + 3);
  f[6] = f[1] * f[5] - f[2] * f[4]; f[7] = f[2] * f[3] - f[0] * f[5]; f[8] = f[0] * f[4] - f[1] * f[3];
}

// Specialise the flat model into the PointMass kernel's constant block.  The
// kernel assumes the scene shape of point_reach.py:75-98: one body with two
// slide joints along world x/y carrying one sphere, static planes, one weld to
// a mocap body.  Anything else is rejected loudly.
template <typename real>
static int build_pm_params(const MjsModel& m, const MjsTask& t, PmParams<real>& P) {
  memset(&P, 0, sizeof(P));
  if (m.nv != 2 || m.njnt != 2 || m.neq != 1 || m.nmocap != 1) { mjs_set_error("pointmass kernel: unexpected model shape (nv=%d neq=%d)", m.nv, m.neq); return -1; }
  int b = t.pm_body;
  for (int j = 0; j < 2; j++) {
    const double* ax = m.jnt_axis + 3 * j;
    if (m.jnt_type[j] != MJS_JNT_SLIDE || m.jnt_body[j] != b || std::fabs(ax[j] - 1.0) > 0 || m.body_parent[b] != 0) { mjs_set_error("pointmass kernel: joints must be world-aligned x/y slides"); return -1; }
  }
  const double* bq = m.body_quat + 4 * b;
  if (bq[0] != 1.0) { mjs_set_error("pointmass kernel: body must be unrotated"); return -1; }
  double mass = m.body_mass[b];
  P.m00 = (real)(mass + m.dof_armature[0]); P.m11 = (real)(mass + m.dof_armature[1]); P.m01 = 0;
  P.timestep = (real)m.timestep;
  P.n_sub = t.n_sub; P.reward_type = t.reward_type;
  P.scale = (real)(1.0 / (m.meaninertia * 2));
  P.max_iter = m.solver_iterations;
  // mjOption.tolerance as given (parity tests pass 1e-12 to both sides); FP32 stops at what FP32 can resolve
  P.tol = sizeof(real) == 8 ? (real)m.tolerance : (real)(m.tolerance > 1e-7 ? m.tolerance : 1e-7);
  P.body_x = (real)m.body_pos[3 * b]; P.body_y = (real)m.body_pos[3 * b + 1];
  double bz = m.body_pos[3 * b + 2];
  P.weld_relx = (real)m.eq_data[3]; P.weld_rely = (real)m.eq_data[4];
  P.weld_diag = (real)(m.body_invweight0[2 * m.eq_body1[0]] + m.body_invweight0[2 * m.eq_body2[0]]);
  for (int i = 0; i < 2; i++) P.weld_solref[i] = (real)m.eq_solref[i];
  for (int i = 0; i < 5; i++) P.weld_solimp[i] = (real)m.eq_solimp[i];
  P.impratio = (real)m.impratio;
  int sphere = -1;
  for (int g = 0; g < m.ngeom; g++) if (m.geom_body[g] == b && m.geom_type[g] == MJS_GEOM_SPHERE) sphere = g;
  if (sphere < 0) { mjs_set_error("pointmass kernel: no sphere geom"); return -1; }
  double r = m.geom_size[3 * sphere];
  for (int p = 0; p < m.npair; p++) {
    int g1 = m.pair_geom1[p], g2 = m.pair_geom2[p];
    if (g2 != sphere || m.geom_type[g1] != MJS_GEOM_PLANE) { mjs_set_error("pointmass kernel: unsupported collision pair"); return -1; }
    if (P.nplane >= PM_MAXPLANE) { mjs_set_error("pointmass kernel: too many planes"); return -1; }
    int k = P.nplane++;
    double R[9], f[9];
    h_quat2mat(R, m.geom_quat + 4 * g1);
    double n[3] = {R[2], R[5], R[8]};
    const double* p0 = m.geom_pos + 3 * g1;
    // oracle: dist = n.(c - p0) - r  with c = (bx+qx, by+qy, bz)
    P.pl_nx[k] = (real)n[0]; P.pl_ny[k] = (real)n[1];
    P.pl_c0[k] = (real)(n[2] * (bz - p0[2]) - n[0] * p0[0] - n[1] * p0[1] - r);
    f[0] = n[0]; f[1] = n[1]; f[2] = n[2];
    h_make_frame(f);
    int dim = m.geom_condim[g1] > m.geom_condim[g2] ? m.geom_condim[g1] : m.geom_condim[g2];
    if (dim != 3 || m.geom_priority[g1] != m.geom_priority[g2]) { mjs_set_error("pointmass kernel: condim 3 / equal priority only"); return -1; }
    double mu = std::fmax(m.geom_friction[3 * g1], m.geom_friction[3 * g2]);
    mu = std::fmax(mu, MJS_MINMU);
    // body2 = sphere body: Jp = [ex ey]; contact-frame Jacobian = frame[:, 0:2]
    for (int e = 0; e < 4; e++) {
      int axis = 1 + e / 2;
      double sgn = (e % 2 == 0) ? 1.0 : -1.0;
      for (int c = 0; c < 2; c++) P.pl_J[k][e][c] = (real)(f[c] + sgn * mu * f[3 * axis + c]);
    }
    double tran = m.body_invweight0[2 * b] + m.body_invweight0[0];
    P.pl_diag[k] = (real)(tran + mu * mu * tran);
    P.pl_mu[k] = (real)mu;
    P.pl_margin[k] = (real)(std::fmax(m.geom_margin[g1], m.geom_margin[g2]) - std::fmax(m.geom_gap[g1], m.geom_gap[g2]));
    double x1 = m.geom_solmix[g1], x2 = m.geom_solmix[g2];
    double mix = (x1 >= MJS_MINVAL && x2 >= MJS_MINVAL) ? x1 / (x1 + x2) : ((x1 < MJS_MINVAL && x2 < MJS_MINVAL) ? 0.5 : (x1 < MJS_MINVAL ? 0.0 : 1.0));
    const double *r1 = m.geom_solref + 2 * g1, *r2 = m.geom_solref + 2 * g2;
    for (int i = 0; i < 2; i++) P.pl_solref[k][i] = (real)((r1[0] > 0 && r2[0] > 0) ? mix * r1[i] + (1 - mix) * r2[i] : std::fmin(r1[i], r2[i]));
    for (int i = 0; i < 5; i++) P.pl_solimp[k][i] = (real)(mix * m.geom_solimp[5 * g1 + i] + (1 - mix) * m.geom_solimp[5 * g2 + i]);
  }
  P.lo = (real)t.arena_range[0]; P.hi = (real)t.arena_range[1];
  P.goal_threshold = (real)t.goal_threshold; P.time_limit = (real)t.time_limit;
  for (int i = 0; i < 2; i++) {
    P.spawn_lo[i] = (real)t.spawn_lo[i]; P.spawn_hi[i] = (real)t.spawn_hi[i];
    P.target_lo[i] = (real)t.target_lo[i]; P.target_hi[i] = (real)t.target_hi[i];
    P.action_lo[i] = (real)t.action_lo[i]; P.action_hi[i] = (real)t.action_hi[i];
  }
  return 0;
}

static PmIO pm_io(MjsSim* s, const float* act, float* obs, float* rew, unsigned char* flags, float* fobs, int autoreset) {
  PmIO io;
  io.sd = s->pm_sd; io.si = s->pm_si; io.actions = act; io.obs = obs; io.reward = rew; io.flags = flags; io.final_obs = fobs;
  io.obs64 = s->obs64; io.reward64 = s->rew64; io.final_obs64 = s->fobs64;
  io.stats = s->d_stats; io.N = s->N; io.autoreset = autoreset; io.seed = s->seed; io.env_offset = s->env_offset; io.step_index = s->step_index;
  return io;
}

// Every entry point runs with the simulator's device current and leaves the caller's current device as it found it
// (a simulator created on cuda:1 may be stepped while the caller's current device is cuda:0).
struct DeviceGuard {
  int prev = -1, changed = 0;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) { cudaSetDevice(dev); changed = 1; }
  }
  ~DeviceGuard() { if (changed) cudaSetDevice(prev); }
};

extern "C" {

const char* mjs_last_error(void) { return g_err; }
const char* mjs_version(void) { return "mjsim_b200 0.1 (sm_100a)"; }
int mjs_num_envs(const MjsSim* s) { return s->N; }
int mjs_obs_dim(const MjsSim* s) { return s->task.obs_dim; }
int mjs_action_dim(const MjsSim* s) { return s->task.action_dim; }
long long mjs_launch_count(const MjsSim* s) { return s->launches; }

int mjs_create(const MjsModel* model, const MjsTask* task, int num_envs, int device, unsigned long long seed, unsigned int env_offset,
               int precision, MjsSim** out) {
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    mjs_set_error("mjs_create: no CUDA device -- this library has no CPU fallback");
    return -2;
  }
  if (device < 0 || device >= ndev) { mjs_set_error("mjs_create: device %d out of range (%d visible)", device, ndev); return -1; }
  DeviceGuard guard(device);
  MjsSim* s = new MjsSim();
  memset(s, 0, sizeof(*s));
  s->model = *model; s->task = *task; s->N = num_envs; s->device = device; s->seed = seed; s->env_offset = env_offset; s->precision = precision;
  int N = num_envs, O = task->obs_dim, A = task->action_dim;
  MJS_CUDA_CHECK(cudaMalloc(&s->d_stats, 8 * sizeof(double)));
  MJS_CUDA_CHECK(cudaMemset(s->d_stats, 0, 8 * sizeof(double)));
  MJS_CUDA_CHECK(cudaMalloc(&s->d_act, sizeof(float) * N * A));
  MJS_CUDA_CHECK(cudaMalloc(&s->d_obs, sizeof(float) * N * O));
  MJS_CUDA_CHECK(cudaMalloc(&s->d_fobs, sizeof(float) * N * O));
  MJS_CUDA_CHECK(cudaMalloc(&s->d_rew, sizeof(float) * N));
  MJS_CUDA_CHECK(cudaMalloc(&s->d_flags, 4 * N));
  MJS_CUDA_CHECK(cudaMalloc(&s->d_mask, N));
  MJS_CUDA_CHECK(cudaStreamCreateWithFlags(&s->hstream, cudaStreamNonBlocking));
  if (task->kind == MJS_TASK_POINTMASS) {
    if (build_pm_params<double>(*model, *task, s->pm_pd) || build_pm_params<float>(*model, *task, s->pm_pf)) { delete s; return -1; }
    MJS_CUDA_CHECK(cudaMalloc(&s->pm_sd, sizeof(double) * PM_NFIELD * N));
    MJS_CUDA_CHECK(cudaMemset(s->pm_sd, 0, sizeof(double) * PM_NFIELD * N));
    MJS_CUDA_CHECK(cudaMalloc(&s->pm_si, sizeof(int) * PM_I_NFIELD * N));
    MJS_CUDA_CHECK(cudaMemset(s->pm_si, 0, sizeof(int) * PM_I_NFIELD * N));
  } else {
    s->arm = arm_create(*model, *task, N, seed, env_offset, precision, s->d_stats);
    if (!s->arm) { delete s; return -1; }
  }
  *out = s;
  return 0;
}

void mjs_destroy(MjsSim* s) {
  if (!s) return;
  DeviceGuard guard(s->device);
  cudaFree(s->d_stats); cudaFree(s->d_act); cudaFree(s->d_obs); cudaFree(s->d_fobs); cudaFree(s->d_rew); cudaFree(s->d_flags); cudaFree(s->d_mask);
  cudaFree(s->pm_sd); cudaFree(s->pm_si); cudaFree(s->d_gpose); cudaFree(s->d_aux);
  if (s->arm) arm_destroy(s->arm);
  cudaStreamDestroy(s->hstream);
  delete s;
}

int mjs_reset(MjsSim* s, const unsigned char* d_mask, float* d_obs, void* stream) {
  DeviceGuard guard(s->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (s->task.kind == MJS_TASK_POINTMASS) {
    PmIO io = pm_io(s, nullptr, d_obs, nullptr, nullptr, nullptr, 0);
    int grid = (s->N + 127) / 128;
    if (s->precision == MJS_PRECISION_F64) pm_reset_kernel<double><<<grid, 128, 0, st>>>(s->pm_pd, io, d_mask);
    else pm_reset_kernel<float><<<grid, 128, 0, st>>>(s->pm_pf, io, d_mask);
    s->launches++;
  } else {
    int r = arm_reset(s->arm, d_mask, d_obs, st);
    if (r < 0) return r;
    s->launches += r;
  }
  MJS_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int mjs_step(MjsSim* s, const float* d_actions, float* d_obs, float* d_reward, unsigned char* d_flags, float* d_final_obs, int autoreset, void* stream) {
  DeviceGuard guard(s->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (s->task.kind == MJS_TASK_POINTMASS) {
    PmIO io = pm_io(s, d_actions, d_obs, d_reward, d_flags, d_final_obs, autoreset);
    // one warp per block: 65 536 envs are 2048 blocks = 13.8 per SM (blocks of 128 threads: 3.5 per SM, a 15 % imbalance between SMs)
    int grid = (s->N + 31) / 32;
    if (s->precision == MJS_PRECISION_F64) pm_step_kernel<double><<<grid, 32, 0, st>>>(s->pm_pd, io);
    else pm_step_kernel<float><<<grid, 32, 0, st>>>(s->pm_pf, io);
    s->launches++;
  } else {
    int r = arm_step(s->arm, d_actions, d_obs, d_reward, d_flags, d_final_obs, autoreset, s->step_index, st);
    if (r < 0) return r;
    s->launches += r;
  }
  s->step_index++;
  MJS_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int mjs_step_noalloc(MjsSim* s, const float* d_actions, float* d_obs, float* d_reward, unsigned char* d_flags, float* d_final_obs, int autoreset, void* stream) {
  DeviceGuard guard(s->device);
  if (!s->arm) return mjs_step(s, d_actions, d_obs, d_reward, d_flags, d_final_obs, autoreset, stream);  // PointMass: the step uses no scratch and allocates nothing
  int r = arm_step_noalloc(s->arm, d_actions, d_obs, d_reward, d_flags, d_final_obs, autoreset, s->step_index, (cudaStream_t)stream);
  if (r < 0) return r;
  s->launches += r;
  s->step_index++;
  MJS_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int mjs_step_host(MjsSim* s, const float* h_actions, float* h_obs, float* h_reward, unsigned char* h_flags, float* h_final_obs, int autoreset) {
  DeviceGuard guard(s->device);
  int N = s->N, A = s->task.action_dim, O = s->task.obs_dim;
  cudaStream_t st = s->hstream;
  if (h_actions) MJS_CUDA_CHECK(cudaMemcpyAsync(s->d_act, h_actions, sizeof(float) * N * A, cudaMemcpyHostToDevice, st));
  int r = mjs_step(s, h_actions ? s->d_act : nullptr, s->d_obs, s->d_rew, s->d_flags, h_final_obs ? s->d_fobs : nullptr, autoreset, st);
  if (r) return r;
  MJS_CUDA_CHECK(cudaMemcpyAsync(h_obs, s->d_obs, sizeof(float) * N * O, cudaMemcpyDeviceToHost, st));
  MJS_CUDA_CHECK(cudaMemcpyAsync(h_reward, s->d_rew, sizeof(float) * N, cudaMemcpyDeviceToHost, st));
  MJS_CUDA_CHECK(cudaMemcpyAsync(h_flags, s->d_flags, 4 * N, cudaMemcpyDeviceToHost, st));
  if (h_final_obs) MJS_CUDA_CHECK(cudaMemcpyAsync(h_final_obs, s->d_fobs, sizeof(float) * N * O, cudaMemcpyDeviceToHost, st));
  MJS_CUDA_CHECK(cudaStreamSynchronize(st));
  return 0;
}

int mjs_reset_host(MjsSim* s, const unsigned char* h_mask, float* h_obs) {
  DeviceGuard guard(s->device);
  cudaStream_t st = s->hstream;
  if (h_mask) MJS_CUDA_CHECK(cudaMemcpyAsync(s->d_mask, h_mask, s->N, cudaMemcpyHostToDevice, st));
  int r = mjs_reset(s, h_mask ? s->d_mask : nullptr, s->d_obs, st);
  if (r) return r;
  if (h_obs) MJS_CUDA_CHECK(cudaMemcpyAsync(h_obs, s->d_obs, sizeof(float) * s->N * s->task.obs_dim, cudaMemcpyDeviceToHost, st));
  MJS_CUDA_CHECK(cudaStreamSynchronize(st));
  return 0;
}

int mjs_set_state(MjsSim* s, const MjsEnvState* h, int first, int count, int forward) {
  DeviceGuard guard(s->device);
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  int N = s->N;
  if (first < 0 || count < 0 || first + count > N) { mjs_set_error("mjs_set_state: range out of bounds"); return -1; }
  if (s->task.kind == MJS_TASK_POINTMASS) {
    // only the columns [first, first + count) of the field-major arrays travel (pitch = N elements)
    const size_t C = (size_t)count;
    if (C == 0) return 0;
    std::vector<double> sd((size_t)PM_NFIELD * C);
    std::vector<int> si((size_t)PM_I_NFIELD * C);
    MJS_CUDA_CHECK(cudaMemcpy2D(sd.data(), C * sizeof(double), s->pm_sd + first, (size_t)N * sizeof(double), C * sizeof(double), PM_NFIELD, cudaMemcpyDeviceToHost));
    MJS_CUDA_CHECK(cudaMemcpy2D(si.data(), C * sizeof(int), s->pm_si + first, (size_t)N * sizeof(int), C * sizeof(int), PM_I_NFIELD, cudaMemcpyDeviceToHost));
    for (size_t e = 0; e < C; e++) {
      const MjsEnvState& x = h[e];
      sd[PM_QX * C + e] = x.qpos[0]; sd[PM_QY * C + e] = x.qpos[1]; sd[PM_VX * C + e] = x.qvel[0]; sd[PM_VY * C + e] = x.qvel[1];
      sd[PM_MX * C + e] = x.mocap_pos[0]; sd[PM_MY * C + e] = x.mocap_pos[1]; sd[PM_WX * C + e] = x.qacc_warm[0]; sd[PM_WY * C + e] = x.qacc_warm[1];
      sd[PM_TIME * C + e] = x.time; sd[PM_TX * C + e] = x.target[0]; sd[PM_TY * C + e] = x.target[1];
      sd[PM_DIST * C + e] = x.dist; sd[PM_PDIST * C + e] = x.prev_dist; sd[PM_RET * C + e] = x.ep_return;
      si[PM_I_EPISODE * C + e] = x.episode_id; si[PM_I_STATUS * C + e] = x.status; si[PM_I_STEPS * C + e] = x.episode_step;
    }
    MJS_CUDA_CHECK(cudaMemcpy2D(s->pm_sd + first, (size_t)N * sizeof(double), sd.data(), C * sizeof(double), C * sizeof(double), PM_NFIELD, cudaMemcpyHostToDevice));
    MJS_CUDA_CHECK(cudaMemcpy2D(s->pm_si + first, (size_t)N * sizeof(int), si.data(), C * sizeof(int), C * sizeof(int), PM_I_NFIELD, cudaMemcpyHostToDevice));
    (void)forward;  // callers pass the warm start explicitly for PointMass
    return 0;
  }
  return arm_set_state(s->arm, h, first, count, forward);
}

int mjs_get_state(MjsSim* s, MjsEnvState* h, int first, int count) {
  DeviceGuard guard(s->device);
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  int N = s->N;
  if (first < 0 || count < 0 || first + count > N) { mjs_set_error("mjs_get_state: range out of bounds"); return -1; }
  if (s->task.kind == MJS_TASK_POINTMASS) {
    const size_t C = (size_t)count;
    if (C == 0) return 0;
    std::vector<double> sd((size_t)PM_NFIELD * C);
    std::vector<int> si((size_t)PM_I_NFIELD * C);
    MJS_CUDA_CHECK(cudaMemcpy2D(sd.data(), C * sizeof(double), s->pm_sd + first, (size_t)N * sizeof(double), C * sizeof(double), PM_NFIELD, cudaMemcpyDeviceToHost));
    MJS_CUDA_CHECK(cudaMemcpy2D(si.data(), C * sizeof(int), s->pm_si + first, (size_t)N * sizeof(int), C * sizeof(int), PM_I_NFIELD, cudaMemcpyDeviceToHost));
    for (size_t e = 0; e < C; e++) {
      MjsEnvState& x = h[e];
      memset(&x, 0, sizeof(x));
      x.qpos[0] = sd[PM_QX * C + e]; x.qpos[1] = sd[PM_QY * C + e]; x.qvel[0] = sd[PM_VX * C + e]; x.qvel[1] = sd[PM_VY * C + e];
      x.mocap_pos[0] = sd[PM_MX * C + e]; x.mocap_pos[1] = sd[PM_MY * C + e]; x.qacc_warm[0] = sd[PM_WX * C + e]; x.qacc_warm[1] = sd[PM_WY * C + e];
      x.time = sd[PM_TIME * C + e]; x.target[0] = sd[PM_TX * C + e]; x.target[1] = sd[PM_TY * C + e];
      x.target[2] = s->task.target_lo[2];
      x.dist = sd[PM_DIST * C + e]; x.prev_dist = sd[PM_PDIST * C + e]; x.ep_return = sd[PM_RET * C + e];
      x.episode_id = si[PM_I_EPISODE * C + e]; x.status = si[PM_I_STATUS * C + e]; x.episode_step = si[PM_I_STEPS * C + e];
    }
    return 0;
  }
  return arm_get_state(s->arm, h, first, count);
}

int mjs_join_aux(MjsSim* s, void* stream) {
  DeviceGuard guard(s->device);
  if (!s->arm) return 0;
  return arm_join_aux(s->arm, (cudaStream_t)stream);
}

int mjs_debug_forward(MjsSim* s, int env, double* h_out, int n) {
  if (!s->arm) { mjs_set_error("mjs_debug_forward: articulated tasks only"); return -1; }
  DeviceGuard guard(s->device);
  return arm_debug_forward(s->arm, env, h_out, n);
}

int mjs_debug_substeps(MjsSim* s, int k) {
  if (!s->arm) { mjs_set_error("mjs_debug_substeps: articulated tasks only"); return -1; }
  DeviceGuard guard(s->device);
  return arm_debug_substeps(s->arm, k);
}

int mjs_demo_policy(MjsSim* s, double noise, float* d_actions, int* d_ok, void* stream) {
  DeviceGuard guard(s->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (!d_actions) { mjs_set_error("mjs_demo_policy: d_actions is null"); return -1; }
  if (s->task.kind == MJS_TASK_POINTMASS) {
    PmIO io = pm_io(s, nullptr, nullptr, nullptr, nullptr, nullptr, 0);
    int grid = (s->N + 127) / 128;
    if (s->precision == MJS_PRECISION_F64) pm_policy_kernel<double><<<grid, 128, 0, st>>>(s->pm_pd, io, d_actions, d_ok, noise);
    else pm_policy_kernel<float><<<grid, 128, 0, st>>>(s->pm_pf, io, d_actions, d_ok, noise);
    s->launches++;
  } else {
    int r = arm_demo_policy(s->arm, noise, s->step_index, d_actions, d_ok, st);
    if (r < 0) return r;
    s->launches += r;
  }
  MJS_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int mjs_debug_times(MjsSim* s, long long* h_out, int n) {
  if (!s->arm) { mjs_set_error("mjs_debug_times: articulated tasks only"); return -1; }
  DeviceGuard guard(s->device);
  return arm_debug_times(s->arm, h_out, n);
}

static void h_mulquat(double* r, const double* a, const double* b) {
  r[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3]; r[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  r[2] = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1]; r[3] = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
}

int mjs_render(MjsSim* s, const MjsCamera* cam, const MjsVisual* vis, int nvis, int width, int height, unsigned char* d_rgb, void* stream) {
  DeviceGuard guard(s->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (nvis < 0 || nvis > MJS_MAXVIS) { mjs_set_error("mjs_render: %d visuals exceed MJS_MAXVIS = %d", nvis, MJS_MAXVIS); return -1; }
  if (width <= 0 || height <= 0 || width > 4096 || height > 4096) { mjs_set_error("mjs_render: bad image size %d x %d", width, height); return -1; }
  if (!d_rgb || !cam || (!vis && nvis)) { mjs_set_error("mjs_render: null argument"); return -1; }
  const int N = s->N, ng = s->model.ngeom;
  RenderScene S;
  memset(&S, 0, sizeof(S));
  S.nvis = nvis; S.width = width; S.height = height;
  auto norm_quat_mat = [](const double* q, float* m) {
    double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]), u[4] = {1, 0, 0, 0}, md[9];
    if (n > 0) for (int k = 0; k < 4; k++) u[k] = q[k] / n;
    h_quat2mat(md, u);
    for (int k = 0; k < 9; k++) m[k] = (float)md[k];
  };
  for (int i = 0; i < nvis; i++) {
    const MjsVisual& v = vis[i];
    RenderVis& r = S.vis[i];
    r.type = v.type; r.geom = v.geom; r.anchor = v.anchor; r.use_active = v.use_active;
    if ((v.anchor == MJS_ANCHOR_GEOM || v.anchor == MJS_ANCHOR_GEOMFRAME) && (v.geom < 0 || v.geom >= ng)) { mjs_set_error("mjs_render: visual %d names geom %d of %d", i, v.geom, ng); return -1; }
    if (v.type != MJS_GEOM_PLANE && v.type != MJS_GEOM_SPHERE && v.type != MJS_GEOM_CAPSULE && v.type != MJS_GEOM_CYLINDER && v.type != MJS_GEOM_BOX) { mjs_set_error("mjs_render: visual %d has unsupported type %d", i, v.type); return -1; }
    for (int k = 0; k < 3; k++) { r.pos[k] = (float)v.pos[k]; r.size[k] = (float)v.size[k]; r.rgb[k] = (float)v.rgba[k]; r.rgb_active[k] = (float)v.rgba_active[k]; }
    norm_quat_mat(v.quat, r.mat);
  }
  S.cam_anchor = cam->anchor; S.cam_geom = cam->geom;
  if (cam->anchor == MJS_ANCHOR_GEOMFRAME && (cam->geom < 0 || cam->geom >= ng)) { mjs_set_error("mjs_render: camera names geom %d of %d", cam->geom, ng); return -1; }
  for (int k = 0; k < 3; k++) S.cam_pos[k] = (float)cam->pos[k];
  norm_quat_mat(cam->quat, S.cam_mat);
  S.tan_half_fovy = (float)std::tan(0.5 * cam->fovy * 3.14159265358979323846 / 180.0);
  if (!s->d_gpose) {
    MJS_CUDA_CHECK(cudaMalloc(&s->d_gpose, sizeof(float) * 12 * MJS_MAXGEOM * (size_t)N));
    MJS_CUDA_CHECK(cudaMalloc(&s->d_aux, sizeof(float) * 8 * (size_t)N));
  }
  if (s->arm) {
    int r = arm_poses(s->arm, s->d_gpose, s->d_aux, st);
    if (r < 0) return r;
    s->launches += r;
  } else {
    // PointMass: world pose of every geom at q = 0 (bodies are children of the world, unrotated: build_pm_params checks)
    PmPoseTable T;
    memset(&T, 0, sizeof(T));
    T.ngeom = ng;
    for (int g = 0; g < ng; g++) {
      int b = s->model.geom_body[g];
      double q[4], m[9];
      h_mulquat(q, s->model.body_quat + 4 * b, s->model.geom_quat + 4 * g);
      h_quat2mat(m, q);
      double bm[9], p[3];
      h_quat2mat(bm, s->model.body_quat + 4 * b);
      for (int k = 0; k < 3; k++) p[k] = s->model.body_pos[3 * b + k] + bm[3 * k] * s->model.geom_pos[3 * g] + bm[3 * k + 1] * s->model.geom_pos[3 * g + 1] + bm[3 * k + 2] * s->model.geom_pos[3 * g + 2];
      for (int k = 0; k < 3; k++) T.base[g][k] = (float)p[k];
      for (int k = 0; k < 9; k++) T.base[g][3 + k] = (float)m[k];
      T.moves[g] = b == s->task.pm_body;
    }
    pm_pose_kernel<<<(N + 127) / 128, 128, 0, st>>>(T, s->pm_sd, N, s->d_gpose, s->d_aux);
    s->launches++;
  }
  dim3 grid(N, (width * height + 127) / 128);
  render_kernel<<<grid, 128, 0, st>>>(S, s->d_gpose, s->d_aux, d_rgb);
  s->launches++;
  MJS_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int mjs_set_output64(MjsSim* s, double* d_obs64, double* d_reward64, double* d_final_obs64) {
  s->obs64 = d_obs64; s->rew64 = d_reward64; s->fobs64 = d_final_obs64;
  if (s->arm) arm_set_output64(s->arm, d_obs64, d_reward64, d_final_obs64);
  return 0;
}

int mjs_reseed(MjsSim* s, unsigned long long seed) {
  DeviceGuard guard(s->device);
  s->seed = seed;
  s->step_index = 0;
  if (s->arm) return arm_reseed(s->arm, seed);
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  MJS_CUDA_CHECK(cudaMemset(s->pm_si + (size_t)PM_I_EPISODE * s->N, 0, sizeof(int) * (size_t)s->N));
  return 0;
}

int mjs_stats_device(MjsSim* s, double** d_out) { *d_out = s->d_stats; return 0; }

int mjs_status(MjsSim* s, int* d_status, void* stream) {
  DeviceGuard guard(s->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (!d_status) { mjs_set_error("mjs_status: d_status is NULL"); return -1; }
  if (s->task.kind == MJS_TASK_POINTMASS) {  // field-major: the status column is contiguous
    MJS_CUDA_CHECK(cudaMemcpyAsync(d_status, s->pm_si + (size_t)PM_I_STATUS * s->N, sizeof(int) * (size_t)s->N, cudaMemcpyDeviceToDevice, st));
    return 0;
  }
  return arm_status(s->arm, d_status, st);
}

int mjs_stats(MjsSim* s, double* h_out8, int reset) {
  DeviceGuard guard(s->device);
  MJS_CUDA_CHECK(cudaDeviceSynchronize());
  MJS_CUDA_CHECK(cudaMemcpy(h_out8, s->d_stats, 8 * sizeof(double), cudaMemcpyDeviceToHost));
  if (reset) MJS_CUDA_CHECK(cudaMemset(s->d_stats, 0, 8 * sizeof(double)));
  return 0;
}

}  // extern "C"
