// pointmass.cuh -- PointMass Reach control-step kernel: ONE THREAD PER ENV.
//
// nv = 2, so a warp per env would idle 30 lanes; all
// n_sub physics substeps, the composer hooks (point_reach.py:150-202), episode
// bookkeeping, auto-reset and observation assembly run in one launch with the
// whole env state in registers.  State in HBM is SoA (field-major, env index
// fastest) so every load/store of a warp is one coalesced 128/256-byte line.
//
// Physics restated per substep ("legacy_step" order):
//   rows(q, v, mocap)   weld x/y rows + pyramidal rows of every active wall
//   Newton solve        2x2 Hessian, exact piecewise-quadratic line search
//   semi-implicit Euler
#pragma once
#include "mjs_common.cuh"

#define PM_MAXPLANE 8
#define PM_MAXROW (2 + 4 * PM_MAXPLANE)

// SoA field indices of the PointMass state (double words)
enum { PM_QX = 0, PM_QY, PM_VX, PM_VY, PM_MX, PM_MY, PM_WX, PM_WY, PM_TIME, PM_TX, PM_TY, PM_DIST, PM_PDIST, PM_RET, PM_NFIELD };
// int fields
enum { PM_I_EPISODE = 0, PM_I_STATUS, PM_I_STEPS, PM_I_NFIELD };

template <typename real>
struct PmParams {
  int nplane, n_sub, reward_type, pad;
  real m00, m01, m11;            // joint-space inertia (constant)
  real timestep, tol, scale;     // solver tolerance and 1/(meaninertia*nv)
  int max_iter, pad2;
  real body_x, body_y;           // body offset: xpos = body + q
  // weld (mocap <-> body): residual = (mocap + relpos) - xpos, J = -I
  real weld_relx, weld_rely, weld_diag, weld_solref[2], weld_solimp[5];
  // planes: signed distance of the sphere centre = n.(c) - d0 - r
  real pl_nx[PM_MAXPLANE], pl_ny[PM_MAXPLANE], pl_c0[PM_MAXPLANE];  // c0 = nz*cz - d0 - r
  real pl_J[PM_MAXPLANE][4][2];  // pyramid edge Jacobians (x,y dofs)
  real pl_diag[PM_MAXPLANE], pl_mu[PM_MAXPLANE], pl_margin[PM_MAXPLANE], pl_solref[PM_MAXPLANE][2], pl_solimp[PM_MAXPLANE][5];
  real impratio;
  // task
  real lo, hi, goal_threshold, time_limit;
  real spawn_lo[2], spawn_hi[2], target_lo[2], target_hi[2], action_lo[2], action_hi[2];
};

// Constraint rows of one env, sized for what the scene can activate: the 2 weld rows and the pyramids of at most PM_MAXACT
// planes at once (a sphere in the convex arena touches at most two walls; the ground plane is exactly tangent, dist == 0, and
// mj_collision drops dist >= margin).  Every array is indexed with compile-time constants (unrolled slot loops with
// `s < nact` guards), so the rows, their residuals and the line-search scratch live in registers: the previous layout
// (J[34][2] + D[34] + aref[34] + jar[34] + jd[34] per thread) sat in local memory -- 2.9 M local loads per launch (ncu, r1).
#define PM_MAXACT 2
template <typename real>
struct PmRows {
  real wD[2], waref[2];                                           // weld rows: J = -e_i
  real pD[PM_MAXACT], pJ[PM_MAXACT][4][2], paref[PM_MAXACT][4];  // active planes, in plane order
  int nact, overflow;
};

// rows from (q, v, mocap): mj_makeConstraint + mj_makeImpedance + mj_referenceConstraint
template <typename real>
__device__ __forceinline__ void pm_build_rows(const PmParams<real>& P, real qx, real qy, real vx, real vy, real mx, real my, PmRows<real>& R) {
  real k, b;
  // weld rows (equality): J = -I on (x, y)
  mjs_kb<real>(P.weld_solref, P.weld_solimp, P.timestep, &k, &b);
  real res[2] = {(mx + P.weld_relx) - (P.body_x + qx), (my + P.weld_rely) - (P.body_y + qy)};
#pragma unroll
  for (int i = 0; i < 2; i++) {
    real imp = mjs_impedance<real>(P.weld_solimp, res[i], (real)0);
    real Rr = fmax((real)MJS_MINVAL, (1 - imp) * P.weld_diag / imp);
    R.wD[i] = 1 / Rr;
    real vel = i == 0 ? -vx : -vy;
    R.waref[i] = -b * vel - k * imp * res[i];
  }
  R.nact = 0; R.overflow = 0;
#pragma unroll
  for (int s = 0; s < PM_MAXACT; s++) {
    R.pD[s] = 0;
#pragma unroll
    for (int e = 0; e < 4; e++) { R.pJ[s][e][0] = 0; R.pJ[s][e][1] = 0; R.paref[s][e] = 0; }
  }
  for (int p = 0; p < P.nplane; p++) {
    real dist = P.pl_nx[p] * (P.body_x + qx) + P.pl_ny[p] * (P.body_y + qy) + P.pl_c0[p];
    // narrowphase keeps dist <= margin; mj_collision excludes dist >= margin - gap (gap = 0)
    if (dist >= P.pl_margin[p]) continue;
    if (R.nact >= PM_MAXACT) { R.overflow = 1; continue; }
    mjs_kb<real>(P.pl_solref[p], P.pl_solimp[p], P.timestep, &k, &b);
    real imp = mjs_impedance<real>(P.pl_solimp[p], dist, P.pl_margin[p]);
    real R0 = fmax((real)MJS_MINVAL, (1 - imp) * P.pl_diag[p] / imp);
    real mu = P.pl_mu[p] / sqrt(P.impratio);
    real D = 1 / (2 * mu * mu * R0);
    const int slot = R.nact++;
#pragma unroll
    for (int s = 0; s < PM_MAXACT; s++) {
      if (s == slot) {
        R.pD[s] = D;
#pragma unroll
        for (int e = 0; e < 4; e++) {
          real j0 = P.pl_J[p][e][0], j1 = P.pl_J[p][e][1];
          R.pJ[s][e][0] = j0; R.pJ[s][e][1] = j1;
          R.paref[s][e] = -b * (j0 * vx + j1 * vy) - k * imp * (dist - P.pl_margin[p]);
        }
      }
    }
  }
}

// residuals jar = J a - aref of all rows
template <typename real>
struct PmJar { real w[2], p[PM_MAXACT][4]; };

template <typename real>
__device__ __forceinline__ void pm_jar(const PmRows<real>& R, real ax, real ay, PmJar<real>& j) {
  j.w[0] = -ax - R.waref[0]; j.w[1] = -ay - R.waref[1];
#pragma unroll
  for (int s = 0; s < PM_MAXACT; s++)
#pragma unroll
    for (int e = 0; e < 4; e++) j.p[s][e] = R.pJ[s][e][0] * ax + R.pJ[s][e][1] * ay - R.paref[s][e];
}

template <typename real>
__device__ __forceinline__ real pm_cost(const PmParams<real>& P, const PmRows<real>& R, real ax, real ay) {
  // qacc_smooth = 0 (no smooth forces act on the two slide dofs)
  real c = (real)0.5 * (P.m00 * ax * ax + 2 * P.m01 * ax * ay + P.m11 * ay * ay);
  PmJar<real> j;
  pm_jar(R, ax, ay, j);
  c += (real)0.5 * R.wD[0] * j.w[0] * j.w[0];
  c += (real)0.5 * R.wD[1] * j.w[1] * j.w[1];
#pragma unroll
  for (int s = 0; s < PM_MAXACT; s++)
#pragma unroll
    for (int e = 0; e < 4; e++)
      if (s < R.nact && j.p[s][e] < 0) c += (real)0.5 * R.pD[s] * j.p[s][e] * j.p[s][e];
  return c;
}

// primal Newton with exact line search (same algorithm as oracle solve_constraints)
template <typename real>
__device__ __forceinline__ void pm_solve(const PmParams<real>& P, const PmRows<real>& R, real& ax, real& ay) {
  // warm start: the cheaper of qacc_warmstart and qacc_smooth (= 0)
  real cw = pm_cost(P, R, ax, ay), cs = pm_cost(P, R, (real)0, (real)0);
  if (!(cw < cs)) { ax = 0; ay = 0; cw = cs; }
  real cost = cw;
  PmJar<real> jar, jd;
  pm_jar(R, ax, ay, jar);
  const int nact = R.nact;
  for (int it = 0; it < P.max_iter; it++) {
    real gx = P.m00 * ax + P.m01 * ay, gy = P.m01 * ax + P.m11 * ay;
    real h00 = P.m00, h01 = P.m01, h11 = P.m11;
    // weld rows: J = -e_i, always active
    gx += -R.wD[0] * jar.w[0]; h00 += R.wD[0];
    gy += -R.wD[1] * jar.w[1]; h11 += R.wD[1];
#pragma unroll
    for (int s = 0; s < PM_MAXACT; s++)
#pragma unroll
      for (int e = 0; e < 4; e++)
        if (s < nact && jar.p[s][e] < 0) {
          real dj0 = R.pD[s] * R.pJ[s][e][0], dj1 = R.pD[s] * R.pJ[s][e][1];
          gx += dj0 * jar.p[s][e]; gy += dj1 * jar.p[s][e];
          h00 += dj0 * R.pJ[s][e][0]; h01 += dj0 * R.pJ[s][e][1]; h11 += dj1 * R.pJ[s][e][1];
        }
    if (sqrt(gx * gx + gy * gy) * P.scale < P.tol) break;
    real det = h00 * h11 - h01 * h01;
    real dx = -(h11 * gx - h01 * gy) / det, dy = -(-h01 * gx + h00 * gy) / det;
    // phi'(0), phi''(0+)
    real Mdx = P.m00 * dx + P.m01 * dy, Mdy = P.m01 * dx + P.m11 * dy;
    real sl = dx * (P.m00 * ax + P.m01 * ay) + dy * (P.m01 * ax + P.m11 * ay);
    real cu = dx * Mdx + dy * Mdy;
    jd.w[0] = -dx; jd.w[1] = -dy;
    sl += R.wD[0] * jar.w[0] * jd.w[0]; cu += R.wD[0] * jd.w[0] * jd.w[0];
    sl += R.wD[1] * jar.w[1] * jd.w[1]; cu += R.wD[1] * jd.w[1] * jd.w[1];
#pragma unroll
    for (int s = 0; s < PM_MAXACT; s++)
#pragma unroll
      for (int e = 0; e < 4; e++) {
        real t = R.pJ[s][e][0] * dx + R.pJ[s][e][1] * dy;
        jd.p[s][e] = t;
        bool act = s < nact && (jar.p[s][e] < 0 || (jar.p[s][e] == 0 && t < 0));
        if (act) { sl += R.pD[s] * jar.p[s][e] * t; cu += R.pD[s] * t * t; }
      }
    // walk the breakpoints alpha_r = -jar/jd in ascending order
    real alpha = 0, best = -1;
    for (int guard = 0; guard <= 4 * nact; guard++) {
      if (sl >= 0) { best = alpha; break; }
      real next = (real)1e30;
      bool any = false;
#pragma unroll
      for (int s = 0; s < PM_MAXACT; s++)
#pragma unroll
        for (int e = 0; e < 4; e++) {
          if (s < nact && jd.p[s][e] != 0) {
            real al = -jar.p[s][e] / jd.p[s][e];
            if (al > alpha && al < next) { next = al; any = true; }
          }
        }
      real astar = alpha - sl / cu;
      if (astar <= next || !any) { best = astar; break; }
      sl += cu * (next - alpha);
      alpha = next;
      // all rows sharing this breakpoint switch together
#pragma unroll
      for (int s = 0; s < PM_MAXACT; s++)
#pragma unroll
        for (int e = 0; e < 4; e++) {
          if (s < nact && jd.p[s][e] != 0) {
            real al = -jar.p[s][e] / jd.p[s][e];
            if (al == next) { real q = R.pD[s] * jd.p[s][e] * jd.p[s][e]; cu += (jd.p[s][e] > 0) ? -q : q; }
          }
        }
    }
    if (best < 0) best = alpha;
    ax += best * dx; ay += best * dy;
    pm_jar(R, ax, ay, jar);
    real nc = pm_cost(P, R, ax, ay);
    real improvement = cost - nc;
    cost = nc;
    if (improvement * P.scale < P.tol) break;
  }
}

struct PmIO {
  double* sd;          // [PM_NFIELD][N] doubles
  int* si;             // [PM_I_NFIELD][N]
  const float* actions;  // [N][2] (nullptr: draw U(action box) from Philox)
  float* obs;          // [N][4]
  float* reward;       // [N]
  unsigned char* flags;  // [N][4]
  float* final_obs;    // [N][4] observation of the terminal state (before auto-reset), may be null
  double *obs64, *reward64, *final_obs64;  // optional float64 twins (mjs_set_output64), may be null
  double* stats;       // [8] accumulators: return sum, length sum, episodes, successes, truncations, errors, env-steps
  int N, autoreset;
  unsigned long long seed;
  unsigned int env_offset, step_index;
};

template <typename real>
__device__ __forceinline__ void pm_sample_reset(const PmParams<real>& P, unsigned long long seed, unsigned int genv, unsigned int episode,
                                                double& qx, double& qy, double& tx, double& ty) {
  // point_reach.py:125-144: goal_x, goal_y, start_x, start_y (in this order)
  Philox4 a = philox4x32_10(genv, episode, MJS_RNG_RESET, 0, (uint32_t)seed, (uint32_t)(seed >> 32));
  Philox4 b = philox4x32_10(genv, episode, MJS_RNG_RESET, 1, (uint32_t)seed, (uint32_t)(seed >> 32));
  tx = (double)P.target_lo[0] + ((double)P.target_hi[0] - (double)P.target_lo[0]) * u53(a.v[0], a.v[1]);
  ty = (double)P.target_lo[1] + ((double)P.target_hi[1] - (double)P.target_lo[1]) * u53(a.v[2], a.v[3]);
  qx = (double)P.spawn_lo[0] + ((double)P.spawn_hi[0] - (double)P.spawn_lo[0]) * u53(b.v[0], b.v[1]);
  qy = (double)P.spawn_lo[1] + ((double)P.spawn_hi[1] - (double)P.spawn_lo[1]) * u53(b.v[2], b.v[3]);
}

// mj_forward at the reset state (actuation irrelevant): leaves the warm start
template <typename real>
__device__ __forceinline__ void pm_forward_reset(const PmParams<real>& P, real qx, real qy, real mx, real my, real& wx, real& wy) {
  PmRows<real> R;
  pm_build_rows(P, qx, qy, (real)0, (real)0, mx, my, R);
  wx = 0; wy = 0;
  pm_solve(P, R, wx, wy);
}

// Scripted demonstrator of point_reach.py:227-240: step towards the goal, |largest component| = MAX_STEP_SIZE;
// noise scales each component by 1 + N(0, noise).  One thread per env.
template <typename real>
__global__ void __launch_bounds__(128) pm_policy_kernel(PmParams<real> P, PmIO io, float* actions, int* ok_out, double noise) {
  int e = blockIdx.x * blockDim.x + threadIdx.x, N = io.N;
  if (e >= N) return;
  const double* s = io.sd;
  double a[2] = {s[PM_TX * N + e] - ((double)P.body_x + s[PM_QX * N + e]), s[PM_TY * N + e] - ((double)P.body_y + s[PM_QY * N + e])};
  if (noise > 0) {
    Philox4 r = philox4x32_10(io.env_offset + e, io.step_index, MJS_RNG_POLICY, 0, (uint32_t)io.seed, (uint32_t)(io.seed >> 32));
    double u1 = fmax(u53(r.v[0], r.v[1]), 1e-300), u2 = u53(r.v[2], r.v[3]), rad = sqrt(-2.0 * log(u1));
    a[0] *= 1.0 + noise * rad * cos(6.283185307179586 * u2);
    a[1] *= 1.0 + noise * rad * sin(6.283185307179586 * u2);
  }
  double m = fmax(fabs(a[0]), fabs(a[1]));
  double k = m > 0 ? (double)P.action_hi[0] / m : 0.0;  // MAX_STEP_SIZE is the action bound (point_reach.py:208)
  actions[2 * (size_t)e] = (float)(a[0] * k); actions[2 * (size_t)e + 1] = (float)(a[1] * k);
  if (ok_out) ok_out[e] = 1;
}

template <typename real>
__global__ void __launch_bounds__(128) pm_reset_kernel(PmParams<real> P, PmIO io, const unsigned char* mask) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= io.N) return;
  if (mask && !mask[e]) return;
  int N = io.N;
  unsigned int episode = (unsigned int)io.si[PM_I_EPISODE * N + e] + 1u;
  double qx, qy, tx, ty;
  pm_sample_reset(P, io.seed, io.env_offset + e, episode, qx, qy, tx, ty);
  real wx, wy;
  pm_forward_reset<real>(P, (real)qx, (real)qy, (real)qx, (real)qy, wx, wy);
  double* s = io.sd;
  s[PM_QX * N + e] = qx; s[PM_QY * N + e] = qy; s[PM_VX * N + e] = 0; s[PM_VY * N + e] = 0;
  s[PM_MX * N + e] = qx; s[PM_MY * N + e] = qy; s[PM_WX * N + e] = wx; s[PM_WY * N + e] = wy;
  s[PM_TIME * N + e] = 0; s[PM_TX * N + e] = tx; s[PM_TY * N + e] = ty;
  // distance_to_target is NOT reset by the reference (point_reach.py:112-113): first episode starts at 1.0
  if (episode == 1u) { s[PM_DIST * N + e] = 1.0; s[PM_PDIST * N + e] = 1.0; }
  s[PM_RET * N + e] = 0;
  io.si[PM_I_EPISODE * N + e] = (int)episode;
  io.si[PM_I_STEPS * N + e] = 0;
  if (io.obs) {
    float4 o = make_float4((float)((double)P.body_x + qx), (float)((double)P.body_y + qy), (float)tx, (float)ty);
    reinterpret_cast<float4*>(io.obs)[e] = o;
  }
  if (io.obs64) { double* o = io.obs64 + 4 * (size_t)e; o[0] = (double)P.body_x + qx; o[1] = (double)P.body_y + qy; o[2] = tx; o[3] = ty; }
}

template <typename real>
__global__ void __launch_bounds__(128) pm_step_kernel(PmParams<real> P, PmIO io) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  int N = io.N;
  double ret_sum = 0, len_sum = 0, n_ep = 0, n_succ = 0, n_trunc = 0, n_err = 0;
  if (e < N) {
    double* s = io.sd;
    double qx = s[PM_QX * N + e], qy = s[PM_QY * N + e], vx = s[PM_VX * N + e], vy = s[PM_VY * N + e];
    double mx = s[PM_MX * N + e], my = s[PM_MY * N + e];
    real wx = (real)s[PM_WX * N + e], wy = (real)s[PM_WY * N + e];
    double time = s[PM_TIME * N + e], tx = s[PM_TX * N + e], ty = s[PM_TY * N + e];
    double dist = s[PM_DIST * N + e], ret = s[PM_RET * N + e];
    int steps = io.si[PM_I_STEPS * N + e];
    float a0, a1;
    if (io.actions) {
      float2 a = reinterpret_cast<const float2*>(io.actions)[e];
      a0 = a.x; a1 = a.y;
    } else {
      Philox4 r = philox4x32_10(io.env_offset + e, io.step_index, MJS_RNG_ACTION, 0, (uint32_t)io.seed, (uint32_t)(io.seed >> 32));
      a0 = (float)((double)P.action_lo[0] + ((double)P.action_hi[0] - (double)P.action_lo[0]) * u53(r.v[0], r.v[1]));
      a1 = (float)((double)P.action_lo[1] + ((double)P.action_hi[1] - (double)P.action_lo[1]) * u53(r.v[2], r.v[3]));
    }
    // rows left behind by the previous mj_step1: built with the OLD mocap
    PmRows<real> R;
    pm_build_rows<real>(P, (real)qx, (real)qy, (real)vx, (real)vy, (real)mx, (real)my, R);
    int overflow = R.overflow;
    // before_step: mocap = clip(xpos + action) (point_reach.py:150-163; both axes use X_RANGE)
    mx = fmin(fmax((double)P.body_x + qx + (double)a0, (double)P.lo), (double)P.hi);
    my = fmin(fmax((double)P.body_y + qy + (double)a1, (double)P.lo), (double)P.hi);
    for (int k = 0; k < P.n_sub; k++) {
      pm_solve<real>(P, R, wx, wy);                       // mj_step2: constraint solve
      vx += (double)P.timestep * (double)wx; vy += (double)P.timestep * (double)wy;  // Euler
      qx += (double)P.timestep * vx; qy += (double)P.timestep * vy;
      time += (double)P.timestep;
      pm_build_rows<real>(P, (real)qx, (real)qy, (real)vx, (real)vy, (real)mx, (real)my, R);  // mj_step1
      overflow |= R.overflow;
    }
    // after_step + reward/termination (point_reach.py:165-202)
    double px = (double)P.body_x + qx, py = (double)P.body_y + qy;
    double pdist = dist;
    dist = sqrt((px - tx) * (px - tx) + (py - ty) * (py - ty));
    double reward;
    if (P.reward_type == 0) reward = dist < (double)P.goal_threshold ? 1.0 : 0.0;
    else if (P.reward_type == 1) reward = -dist;
    else if (P.reward_type == 2) reward = -dist + 0.5;
    else reward = pdist - dist;
    bool goal = dist < (double)P.goal_threshold;
    bool timeup = (double)P.time_limit > 0 && time >= (double)P.time_limit;
    bool last = goal || timeup;
    bool bad = !(isfinite(qx) && isfinite(qy) && isfinite(vx) && isfinite(vy));
    if (bad) last = true;
    ret += reward;
    steps += 1;
    float4 o = make_float4((float)px, (float)py, (float)tx, (float)ty);
    uchar4 f = make_uchar4(last && goal, last && !goal, goal, last);
    io.reward[e] = (float)reward;
    if (io.reward64) io.reward64[e] = reward;
    reinterpret_cast<uchar4*>(io.flags)[e] = f;
    if (io.final_obs) reinterpret_cast<float4*>(io.final_obs)[e] = o;
    if (io.final_obs64) { double* d = io.final_obs64 + 4 * (size_t)e; d[0] = px; d[1] = py; d[2] = tx; d[3] = ty; }
    double o64[4] = {px, py, tx, ty};
    int episode = io.si[PM_I_EPISODE * N + e];
    int status = io.si[PM_I_STATUS * N + e] | (bad ? MJS_STATUS_NAN : 0) | (overflow ? MJS_STATUS_CONTACT_OVERFLOW : 0);
    if (last) {
      ret_sum = ret; len_sum = steps; n_ep = 1; n_succ = goal; n_trunc = !goal; n_err = bad;
    }
    if (last && io.autoreset) {
      episode += 1;
      pm_sample_reset(P, io.seed, io.env_offset + e, (unsigned int)episode, qx, qy, tx, ty);
      mx = qx; my = qy; vx = 0; vy = 0; time = 0; ret = 0; steps = 0;
      pm_forward_reset<real>(P, (real)qx, (real)qy, (real)mx, (real)my, wx, wy);
      o = make_float4((float)((double)P.body_x + qx), (float)((double)P.body_y + qy), (float)tx, (float)ty);
      o64[0] = (double)P.body_x + qx; o64[1] = (double)P.body_y + qy; o64[2] = tx; o64[3] = ty;
    }
    reinterpret_cast<float4*>(io.obs)[e] = o;
    if (io.obs64) { double* d = io.obs64 + 4 * (size_t)e; d[0] = o64[0]; d[1] = o64[1]; d[2] = o64[2]; d[3] = o64[3]; }
    s[PM_QX * N + e] = qx; s[PM_QY * N + e] = qy; s[PM_VX * N + e] = vx; s[PM_VY * N + e] = vy;
    s[PM_MX * N + e] = mx; s[PM_MY * N + e] = my; s[PM_WX * N + e] = (double)wx; s[PM_WY * N + e] = (double)wy;
    s[PM_TIME * N + e] = time; s[PM_TX * N + e] = tx; s[PM_TY * N + e] = ty;
    s[PM_DIST * N + e] = dist; s[PM_PDIST * N + e] = pdist; s[PM_RET * N + e] = ret;
    io.si[PM_I_EPISODE * N + e] = episode;
    io.si[PM_I_STATUS * N + e] = status;
    io.si[PM_I_STEPS * N + e] = steps;
  }
  // episode statistics: warp-reduce, one atomic per warp that finished an episode
  if (io.stats && __any_sync(0xffffffffu, n_ep > 0)) {  // warp-uniform: most steps no env of the warp finishes an episode
    unsigned full = 0xffffffffu;
    for (int off = 16; off > 0; off >>= 1) {
      ret_sum += __shfl_down_sync(full, ret_sum, off); len_sum += __shfl_down_sync(full, len_sum, off);
      n_ep += __shfl_down_sync(full, n_ep, off); n_succ += __shfl_down_sync(full, n_succ, off);
      n_trunc += __shfl_down_sync(full, n_trunc, off); n_err += __shfl_down_sync(full, n_err, off);
    }
    if ((threadIdx.x & 31) == 0 && n_ep > 0) {
      atomicAdd(io.stats + 0, ret_sum); atomicAdd(io.stats + 1, len_sum); atomicAdd(io.stats + 2, n_ep);
      atomicAdd(io.stats + 3, n_succ); atomicAdd(io.stats + 4, n_trunc); atomicAdd(io.stats + 5, n_err);
    }
  }
}
