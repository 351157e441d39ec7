// arm_kernel.cuh -- ONE WARP PER ENVIRONMENT control-step engine for the
// "UR arm + free props + static geoms" scenes (Robot Reach, Planar Push,
// Button Push).  All n_sub physics substeps, the composer hooks, the analytic
// UR IK, reward/termination/observation assembly and auto-reset run inside one
// launch; the env state is loaded from HBM once and stored once.
//
// Lane roles change per stage (SURVEY.md 8a-2 stage ids in brackets):
//   [P1] kinematics    lanes 0..5 = links (local transforms in parallel, chain
//                      composed through shuffles), lane k = free body k, lane g = geom g
//   [P2,P6] CRB / RNE  lanes 0..5 = links: prefix/suffix scans over the chain
//   [P4] collision     lane p = candidate geom pair p; warp-scan compaction
//   [P5] rows          lane c = contact c (contact-frame Jacobian blocks in smem)
//   [P9] Newton        lane i = dof i for vectors / Hessian rows / Cholesky rows,
//                      lane c = contact c for cone forces and the line search
//   [P11] integrator   lane i = dof i
// No tensor cores: nv <= 30, the work is latency/issue bound FP32 (FP64 for the
// state integration), see DESIGN.md.
#pragma once
#include "arm_model.cuh"
#include "../../include/mjs_convex.h"

#define FULLMASK 0xffffffffu
// warps per lock-step group (must divide the warps per block of the step kernel).  Measured on B200,
// Planar Push: FP64 12 warps/block: group 12 -> 1.30 M env-steps/s kernel-only, 6 -> 1.36 M, 4/3/2 -> 1.36-1.37 M but
// worse once the refill kernel co-runs; FP32 16 warps/block: 16 -> 1.59 M, 4 -> 1.42 M, 2 -> 1.30 M.
#ifndef ARM_SYNC_GROUP_F64
#define ARM_SYNC_GROUP_F64 6
#endif
#ifndef ARM_SYNC_GROUP_F32
#define ARM_SYNC_GROUP_F32 16
#endif

// The specialised model lives in constant memory: every stage reads it as
// constant-bank operands, and the heavy per-substep code can be one
// non-inlined function (small instruction footprint) without passing it around.
__constant__ __align__(16) unsigned char c_arm_model[sizeof(ArmModel<double>)];
template <typename real>
__device__ __forceinline__ const ArmModel<real>& cmodel() { return *reinterpret_cast<const ArmModel<real>*>(c_arm_model); }
// Second copy of the model in global memory for the tables that are indexed by lane (link / geom / pair / joint
// records): a constant-cache load with a different address per lane is replayed once per address, an L1-cached global
// load is one coalesced request.  Measured +3.5 % (f64) / +12 % (FP32) on Planar Push.  Uniform scalars stay in c[].
__device__ __align__(16) unsigned char g_arm_model[sizeof(ArmModel<double>)];
#define GM() (*reinterpret_cast<const ArmModel<real>*>(g_arm_model))

#ifdef MJS_STAGE_CLOCKS_REPEAT
#define MJS_NCLK 14  // + second (instruction-cache-hot) runs of dynamics and collision
#else
#define MJS_NCLK 12
#endif
template <typename real, int NOBJ>
struct WarpMem {
  static constexpr int NV = ARM_NLINK + 6 * NOBJ;
  static constexpr int NQ = ARM_NLINK + 7 * NOBJ;
  static constexpr int HS = NV + 1;
  static constexpr int CS = ARM_MAXC + 1;
  double qpos[NQ + 1];
  double qvel[NV];
  double ctrl[ARM_NLINK];
  real xpos[ARM_NSLOT][3];
  real xmat[ARM_NSLOT][9];
  real gpos[ARM_MAXGEOM][3];
  real gmat[ARM_MAXGEOM][9];
  real S[ARM_NLINK][6];         // arm motion axes about the world origin (ang, lin)
  real Marm[ARM_NLINK][ARM_NLINK];
  real H[NV][HS];               // Newton Hessian / Cholesky factor (solver order: props first, arm last)
  real a[NV], dir[NV], qv[NV], fsm[NV], fcon[NV];
  real cJ[48][CS];              // contact-frame Jacobian blocks: [(side*4 + k)*6 + j][contact]
  real cf[4][CS];               // contact-space force of the current iterate
  real cW[7][CS];               // arrow Hessian weights: W00, W01..W03, W11..W33
  real craw[10][CS];            // dist, pos[3], normal[3], yhint[3]
  // small integer tables are narrow on purpose: 12 FP64 warps must fit in 227 KB of shared memory
  short cpair[ARM_MAXC];
  signed char cblk[2][ARM_MAXC];        // dof block of side 0/1: -1 none, 0 arm, 1+k prop k
  unsigned char ccols[2][ARM_MAXC];     // number of arm columns (link+1) when the side is the arm
  int bcount[1 + ARM_MAXOBJ];
  unsigned short blist[1 + ARM_MAXOBJ][ARM_MAXCB];  // contact | side<<8
  int ncon;
  int status;
  int sync;     // lock-step mode: all warps of the block pass the stage barriers together
  int period;   // the substep-start barrier is taken every `period` substeps (1: every substep)
  int novote;   // lock-step without the Newton loop: the stages before the solve stay aligned, the data-dependent iterations run free
  int coup;     // bit b: the arm block shares a contact with prop b
  int objcoup;  // non-zero: two props share a contact (dense factorisation path)
#ifdef MJS_CONVEX_COUNT
  int ncvx, ncvx_hit;  // profiling build: calls of the general convex routine / calls that found a contact, per control step
#endif
#ifdef MJS_STAGE_CLOCKS
  long long clk[MJS_NCLK];  // profiling build: cycles per stage, accumulated by lane 0
#endif
};

#ifdef MJS_STAGE_CLOCKS
#define STAGE_CLK(W, k, t) do { long long _n = clock64(); if (lane == 0) (W).clk[k] += _n - (t); (t) = _n; } while (0)
#else
#define STAGE_CLK(W, k, t) do { } while (0)
#endif
// ------------------------------------------------------------------ warp helpers
// Double shuffles split the value with __double2loint / __double2hiint (register-pair aliases, no instructions)
// instead of the `asm volatile("mov.b64 ...")` of cuda's sm_30_intrinsics.hpp, which costs four real moves per shuffle
// (ncu, profiles/r1: 300 of the 6.6 k instructions executed once per substep were those moves).
__device__ __forceinline__ float shfl(float v, int src) { return __shfl_sync(FULLMASK, v, src); }
__device__ __forceinline__ int shfl(int v, int src) { return __shfl_sync(FULLMASK, v, src); }
__device__ __forceinline__ float shfl_up(float v, int d) { return __shfl_up_sync(FULLMASK, v, d); }
__device__ __forceinline__ float shfl_down(float v, int d) { return __shfl_down_sync(FULLMASK, v, d); }
__device__ __forceinline__ float shfl_xor(float v, int m) { return __shfl_xor_sync(FULLMASK, v, m); }
#define MJS_SHFL64(NAME, OP)                                                          \
  __device__ __forceinline__ double NAME(double v, int a) {                           \
    int lo = OP(FULLMASK, __double2loint(v), a), hi = OP(FULLMASK, __double2hiint(v), a); \
    return __hiloint2double(hi, lo);                                                  \
  }
MJS_SHFL64(shfl, __shfl_sync)
MJS_SHFL64(shfl_up, __shfl_up_sync)
MJS_SHFL64(shfl_down, __shfl_down_sync)
MJS_SHFL64(shfl_xor, __shfl_xor_sync)
#undef MJS_SHFL64
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += shfl_xor(v, o);
  return v;
}

template <typename real> __device__ __forceinline__ void sincos_r(real x, real* s, real* c);
template <> __device__ __forceinline__ void sincos_r<float>(float x, float* s, float* c) { sincosf(x, s, c); }
// sin / cos of a joint half-angle or a quaternion increment in ~35 instructions instead of the ~160 of sincos(): two-constant
// Cody-Waite reduction by pi/2 (exact enough for |x| < 1e5; the arguments here are bounded by the joint ranges) and the
// fdlibm kernel polynomials on [-pi/4, pi/4].  Max error against libm over [-7, 7]: 1.1e-16 (tools/check_sincos.py).
template <> __device__ __forceinline__ void sincos_r<double>(double x, double* s, double* c) {
  double n = rint(x * 0.63661977236758134308);
  double r = fma(-n, 1.5707963267948966, x);
  r = fma(-n, 6.123233995736766e-17, r);
  double z = r * r;
  double ps = fma(z, fma(z, fma(z, fma(z, fma(z, 1.58969099521155010221e-10, -2.50507602534068634195e-08), 2.75573137070700676789e-06),
                                -1.98412698298579493134e-04), 8.33333333332248946124e-03), -1.66666666666666324348e-01);
  double pc = fma(z, fma(z, fma(z, fma(z, fma(z, -1.13596475577881948265e-11, 2.08757232129817482790e-09), -2.75573143513906633035e-07),
                                2.48015872894767294178e-05), -1.38888888888741095749e-03), 4.16666666666666019037e-02);
  double sr = fma(r * z, ps, r), cr = fma(z, fma(z, pc, -0.5), 1.0);
  int k = (int)n;
  double ss = (k & 1) ? cr : sr, cc = (k & 1) ? sr : cr;
  *s = (k & 2) ? -ss : ss;
  *c = ((k + 1) & 2) ? -cc : cc;
}
template <typename real> __device__ __forceinline__ real rsqrt_r(real x);
template <> __device__ __forceinline__ float rsqrt_r<float>(float x) { return rsqrtf(x); }
// rsqrt(double): MUFU.RSQ64H seed + Newton steps (~1 ulp), a quarter of the instructions of 1.0 / sqrt(x)
#ifdef MJS_RSQRT_NOINLINE  // A/B: one shared copy of the ~25-instruction sequence instead of one per call site (hot code -3.7 KB)
static __device__ __noinline__ double rsqrt_shared(double x) { return rsqrt(x); }
template <> __device__ __forceinline__ double rsqrt_r<double>(double x) { return rsqrt_shared(x); }
#else
template <> __device__ __forceinline__ double rsqrt_r<double>(double x) { return rsqrt(x); }
#endif

template <typename real>
__device__ __forceinline__ void cross3(real* r, const real* a, const real* b) {
  real x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename real>
__device__ __forceinline__ real dot3(const real* a, const real* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename real>
__device__ __forceinline__ void quat_mul(real* r, const real* a, const real* b) {
  real w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  real x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  real y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  real z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
template <typename real>
__device__ __forceinline__ void quat_norm(real* q) {
  real n = rsqrt_r<real>(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  q[0] *= n; q[1] *= n; q[2] *= n; q[3] *= n;
}
template <typename real>
__device__ __forceinline__ void quat_to_mat(real* m, const real* q) {
  real w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
template <typename real>
__device__ __forceinline__ void mat_vec(real* r, const real* m, const real* v) {
  real x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2], z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename real>
__device__ __forceinline__ void matT_vec(real* r, const real* m, const real* v) {
  real x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2], y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2], z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename real>
__device__ __forceinline__ void cross_motion(real* r, const real* vel, const real* v) {
  real a[3], b[3], c[3];
  cross3(a, vel, v); cross3(b, vel, v + 3); cross3(c, vel + 3, v);
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
template <typename real>
__device__ __forceinline__ void cross_force(real* r, const real* vel, const real* f) {
  real a[3], b[3], c[3];
  cross3(a, vel, f); cross3(b, vel + 3, f + 3); cross3(c, vel, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}
// spatial inertia about the world origin: I[0]=m, I[1..3]=m*com, I[4..9]= xx yy zz xy xz yz
template <typename real>
__device__ __forceinline__ void spi_mul(real* f, const real* I, const real* v) {
  real c[3], t[3];
  cross3(c, v, I + 1);
  f[3] = I[0] * v[3] + c[0]; f[4] = I[0] * v[4] + c[1]; f[5] = I[0] * v[5] + c[2];
  cross3(t, I + 1, v + 3);
  f[0] = I[4] * v[0] + I[7] * v[1] + I[8] * v[2] + t[0];
  f[1] = I[7] * v[0] + I[5] * v[1] + I[9] * v[2] + t[1];
  f[2] = I[8] * v[0] + I[9] * v[1] + I[6] * v[2] + t[2];
}

// General convex pair (mjc_Convex restated by include/mjs_convex.h, the same source the oracle compiles): MPR penetration,
// one contact.  Deliberately not inlined: it runs only for the rare pairs without an analytic routine (a prop's cap against
// another prop or the pusher, cylinder x box, capsule x cylinder / box) and must stay out of the substep's hot code.
// Cheap, sound rejection before the general convex routine: two cylinders are disjoint if their projections on some axis
// do not overlap (separating-axis test on four candidate axes: both cylinder axes, their cross product and `extra`, the
// direction between the closest points of the axis segments or between the centres).  Extent of a cylinder (radius r,
// half-height h, axis a) along a unit axis d: h |a.d| + r sqrt(1 - (a.d)^2).  The oracle runs the same test (cyl_separated).
template <typename real>
__device__ __forceinline__ bool cyl_separated(const real* dif, const real* a1, real r1, real h1, const real* a2, real r2, real h2, const real* extra) {
  real cx[3];
  cross3(cx, a1, a2);
  bool sep = false;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    real d[3];
#pragma unroll
    for (int i = 0; i < 3; i++) d[i] = k == 0 ? a1[i] : k == 1 ? a2[i] : k == 2 ? cx[i] : extra[i];
    real n2 = dot3(d, d);
    if (n2 > (real)1e-20) {
      real inv = rsqrt_r<real>(n2);
      real c1 = dot3(a1, d) * inv, c2 = dot3(a2, d) * inv;
      real e1 = h1 * fabs(c1) + r1 * sqrt(fmax((real)0, 1 - c1 * c1)), e2 = h2 * fabs(c2) + r2 * sqrt(fmax((real)0, 1 - c2 * c2));
      if (fabs(dot3(dif, d)) * inv > e1 + e2) sep = true;
    }
  }
  return sep;
}

// separating axis of this lane's pair found by the last miss of the general convex routine (zero: none); lives in the pair
// lane's registers for the rest of the control step
#ifdef MJS_NO_SEPCACHE   // A/B build: no cache (nothing carried through the substep)
template <typename real>
struct SepAxis { static constexpr float v[3] = {0.f, 0.f, 0.f}; };
#define MJS_SEP_SET(S, H, N) do { } while (0)
#define MJS_SEP_ZERO {}
#else
template <typename real>
struct SepAxis { float v[3]; };  // FP32 on purpose: any axis gives a sound test, and it lives in registers across the whole substep
#define MJS_SEP_SET(S, H, N) do { (S).v[0] = (H) ? 0.f : (float)(N)[0]; (S).v[1] = (H) ? 0.f : (float)(N)[1]; (S).v[2] = (H) ? 0.f : (float)(N)[2]; } while (0)
#define MJS_SEP_ZERO {{0, 0, 0}}
#endif
template <typename real>
__device__ __forceinline__ bool cyl_axis_separates(const real* dif, const real* a1, real r1, real h1, const real* a2, real r2, real h2, const real* d) {
  real n2 = dot3(d, d);
  if (!(n2 > (real)1e-20)) return false;
  real inv = rsqrt_r<real>(n2);
  real c1 = dot3(a1, d) * inv, c2 = dot3(a2, d) * inv;
  real e1 = h1 * fabs(c1) + r1 * sqrt(fmax((real)0, 1 - c1 * c1)), e2 = h2 * fabs(c2) + r2 * sqrt(fmax((real)0, 1 - c2 * c2));
  return fabs(dot3(dif, d)) * inv > (e1 + e2) * (1 + (real)1e-6);  // the cached axis is FP32: keep a relative safety margin
}

template <typename real>
struct ConvexHit { real dist, n[3], p[3]; int hit; };
// Arguments are geom indices and shared-memory poses only, the result comes back by value: nothing of the caller's register
// state (cd / cp / nrm / sizes in collide) becomes addressable, so the hot path keeps its registers.
#ifndef MJS_CONVEX_INLINE
#define MJS_CONVEX_INLINE __forceinline__   // one call site, inside the non-inlined convex_fixup
#endif
// The portal search runs in the kernel's own precision (FP64 for precision f64: parity with the oracle to rounding).  One lane
// walks a dependent chain of ~15 support evaluations while its warp -- and, through the substep barrier, its lock-step
// group -- waits: ~4 substeps of SM time per call in FP64 (profiles/r2/convex_pairs.txt).  Most calls are persistent
// near-misses of resting props, so the separating axis of a miss is cached per pair lane for the rest of the control step
// (collide, `sep`) and tested first: sound, and 8x fewer calls.
template <typename real>
__device__ MJS_CONVEX_INLINE ConvexHit<real> convex_pair_dev(int g1, int g2, const real* p1, const real* m1, const real* p2, const real* m2) {
  MjsShape<real> A, B;
  A.type = GM().geom_type[g1]; B.type = GM().geom_type[g2];
#pragma unroll
  for (int k = 0; k < 3; k++) { A.size[k] = GM().geom_size[g1][k]; A.pos[k] = 0; B.size[k] = GM().geom_size[g2][k]; B.pos[k] = p2[k] - p1[k]; }
#pragma unroll
  for (int k = 0; k < 9; k++) { A.mat[k] = m1[k]; B.mat[k] = m2[k]; }
  ConvexHit<real> r;
  real depth = 0;
#if defined(MJS_CONVEX_MODE) && MJS_CONVEX_MODE == 1   // A/B build: the call without the algorithm
  r.hit = 0; r.n[0] = r.n[1] = r.n[2] = 0; r.p[0] = r.p[1] = r.p[2] = 0; r.dist = A.size[0] + B.size[0]; return r;
#endif
  r.hit = mjs_mpr_penetration<real>(A, B, &depth, r.n, r.p);  // on a miss r.n is the separating axis found (or zero)
#if defined(MJS_CONVEX_MODE) && MJS_CONVEX_MODE == 2   // A/B build: the algorithm runs, its contact is dropped
  r.hit = 0;
#endif
  r.dist = -depth;
#pragma unroll
  for (int k = 0; k < 3; k++) r.p[k] += p1[k];
  return r;
}

template <typename real, int NOBJ>
struct ArmEngine {
  using WM = WarpMem<real, NOBJ>;
  using AM = ArmModel<real>;
  static constexpr int NV = WM::NV;
  static constexpr int NQ = WM::NQ;
  // Unroll factor of the outer loops whose bodies index no register array (chain walk, scan steps, contact sides).
  // Rolled: measured +3 % on Planar Push and +6 % on Button Push in f64 against full unrolling (less straight-line code
  // to fetch, same schedule).  Rolling the Cholesky column loops as well measured 4 % (f64) to 19 % (FP32) slower.
#ifndef ARM_CHOL_UNROLL
#define ARM_CHOL_UNROLL 6
#endif
#ifndef ARM_ROLL_FACTOR
#define ARM_ROLL_FACTOR 1
#endif
  static constexpr int ROLL = ARM_ROLL_FACTOR;
  static constexpr int CHOLU = ARM_CHOL_UNROLL;  // unroll factor of the 6-step column loops of the block Cholesky (6 = straight-line code)

  // solver ordering: props first, arm last (no fill-in between props)
  __device__ static __forceinline__ int sidx(int dof) { return dof >= ARM_NLINK ? dof - ARM_NLINK : dof + 6 * NOBJ; }
  __device__ static __forceinline__ int dof_of_sidx(int s) { return s < 6 * NOBJ ? s + ARM_NLINK : s - 6 * NOBJ; }
  __device__ static __forceinline__ int blk_base(int blk) { return blk == 0 ? 0 : ARM_NLINK + 6 * (blk - 1); }

  // Lock-step mode: the per-substep code is far larger than the instruction cache
  // (L1.5 = 32 KB) and has almost no temporal reuse inside a warp, so warps that
  // drift apart all stream their own instructions from L2.  Barriers between the
  // stages keep the warps of a block inside the same code window.
  // Barriers are per lock-step GROUP of ARM_SYNC_GROUP warps (named barriers 1..15), not per block: warps
  // that wait for the slowest member of their group lose less when the group is smaller, at the price of
  // sharing instruction fetches among fewer warps.
  static constexpr int SYNC_GROUP = sizeof(real) == 8 ? ARM_SYNC_GROUP_F64 : ARM_SYNC_GROUP_F32;
  __device__ static __forceinline__ int group_id() { return 1 + (threadIdx.x >> 5) / SYNC_GROUP; }
  __device__ static __forceinline__ void group_sync() {
    int id = group_id();
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(SYNC_GROUP * 32) : "memory");
  }
  __device__ static __forceinline__ bool group_or(bool pred) {
    int id = group_id(), r;
    asm volatile("{ .reg .pred p, q; setp.ne.s32 p, %1, 0; bar.red.or.pred q, %2, %3, p; selp.s32 %0, 1, 0, q; }"
                 : "=r"(r) : "r"((int)pred), "r"(id), "r"(SYNC_GROUP * 32) : "memory");
    return r != 0;
  }
  __device__ static __forceinline__ void stage_sync(const WM& W, int level = 2) { if (W.sync >= level) group_sync(); }

  // ------------------------------------------------------------- [P1] kinematics
  // Restates mj_kinematics for the chain/free-body structure.  Leaves xpos/xmat of
  // every slot, S (motion axes), geom poses in shared memory.
  __device__ static void kinematics(const AM& A, WM& W, int lane) {
    real lq[4] = {1, 0, 0, 0}, lp[3] = {0, 0, 0};
    if (lane < ARM_NLINK) {
      real q = (real)W.qpos[lane], s, c;
      sincos_r<real>(q * (real)0.5, &s, &c);
      real qr[4] = {c, GM().link_axis[lane][0] * s, GM().link_axis[lane][1] * s, GM().link_axis[lane][2] * s};
      real bq[4] = {GM().link_quat[lane][0], GM().link_quat[lane][1], GM().link_quat[lane][2], GM().link_quat[lane][3]};
      quat_mul(lq, bq, qr);
      lp[0] = GM().link_pos[lane][0]; lp[1] = GM().link_pos[lane][1]; lp[2] = GM().link_pos[lane][2];
    }
    // World frames of the chain by an inclusive scan over the link-local transforms (3 shuffle steps instead of
    // a 6-step serial walk): X_i = X_0 o ... o X_i with (qa,pa) o (qb,pb) = (qa*qb, pa + R(qa) pb).  mj_kinematics
    // normalises the quaternion after every body; here once per scan step -- the frames agree to ~1e-16.
    real mq[4] = {lq[0], lq[1], lq[2], lq[3]}, mp[3] = {lp[0], lp[1], lp[2]};   // this lane's link frame
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      real pq[4], pp[3];
#pragma unroll
      for (int k = 0; k < 4; k++) pq[k] = shfl_up(mq[k], o);
#pragma unroll
      for (int k = 0; k < 3; k++) pp[k] = shfl_up(mp[k], o);
      if (lane >= o && lane < ARM_NLINK) {
        real Rm[9], v[3], nq[4];
        quat_to_mat(Rm, pq);
        mat_vec(v, Rm, mp);
        mp[0] = pp[0] + v[0]; mp[1] = pp[1] + v[1]; mp[2] = pp[2] + v[2];
        quat_mul(nq, pq, mq);
        quat_norm(nq);
#pragma unroll
        for (int k = 0; k < 4; k++) mq[k] = nq[k];
      }
    }
    if (lane < ARM_NLINK) {
      real Rm[9], ax[3], al[3] = {GM().link_axis[lane][0], GM().link_axis[lane][1], GM().link_axis[lane][2]}, lin[3];
      quat_to_mat(Rm, mq);
      mat_vec(ax, Rm, al);
      cross3(lin, mp, ax);
#pragma unroll
      for (int k = 0; k < 3; k++) { W.xpos[SLOT_LINK0 + lane][k] = mp[k]; W.S[lane][k] = ax[k]; W.S[lane][3 + k] = lin[k]; }
#pragma unroll
      for (int k = 0; k < 9; k++) W.xmat[SLOT_LINK0 + lane][k] = Rm[k];
    }
    if (lane < NOBJ) {
      int qa = ARM_NLINK + 7 * lane;
      real q4[4] = {(real)W.qpos[qa + 3], (real)W.qpos[qa + 4], (real)W.qpos[qa + 5], (real)W.qpos[qa + 6]}, Rm[9];
      quat_norm(q4);
      quat_to_mat(Rm, q4);
#pragma unroll
      for (int k = 0; k < 3; k++) W.xpos[SLOT_OBJ0 + lane][k] = (real)W.qpos[qa + k];
#pragma unroll
      for (int k = 0; k < 9; k++) W.xmat[SLOT_OBJ0 + lane][k] = Rm[k];
    }
    __syncwarp();
    if (lane < A.ngeom) {
      int sl = GM().geom_slot[lane];
      real gp[3] = {GM().geom_pos[lane][0], GM().geom_pos[lane][1], GM().geom_pos[lane][2]}, v[3];
      const real* R = W.xmat[sl];
      mat_vec(v, R, gp);
#pragma unroll
      for (int k = 0; k < 3; k++) W.gpos[lane][k] = W.xpos[sl][k] + v[k];
#pragma unroll
      for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++)
          W.gmat[lane][3 * r + c] = R[3 * r] * GM().geom_mat[lane][c] + R[3 * r + 1] * GM().geom_mat[lane][3 + c] + R[3 * r + 2] * GM().geom_mat[lane][6 + c];
    }
    __syncwarp();
  }

  // spatial inertia about the world origin and inertial + gravity force of inertial body `a` riding on link `mylink`
  __device__ static __forceinline__ void body_inertia_force(const AM& A, const WM& W, int a, int mylink, const real* vb, const real* ab, real* I, real* f) {
    const real* R = W.xmat[SLOT_LINK0 + mylink];
    real cl[3] = {GM().ab_com[a][0], GM().ab_com[a][1], GM().ab_com[a][2]}, h[3];
    mat_vec(h, R, cl);
    h[0] += W.xpos[SLOT_LINK0 + mylink][0]; h[1] += W.xpos[SLOT_LINK0 + mylink][1]; h[2] += W.xpos[SLOT_LINK0 + mylink][2];
    // Ic_world = R * Ilocal * R^T
    real Il[9] = {GM().ab_I[a][0], GM().ab_I[a][3], GM().ab_I[a][4], GM().ab_I[a][3], GM().ab_I[a][1], GM().ab_I[a][5], GM().ab_I[a][4], GM().ab_I[a][5], GM().ab_I[a][2]};
    real T[9];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
      for (int c = 0; c < 3; c++) T[3 * r + c] = R[3 * r] * Il[c] + R[3 * r + 1] * Il[3 + c] + R[3 * r + 2] * Il[6 + c];
    real Ic[6];  // xx yy zz xy xz yz
    Ic[0] = T[0] * R[0] + T[1] * R[1] + T[2] * R[2];
    Ic[1] = T[3] * R[3] + T[4] * R[4] + T[5] * R[5];
    Ic[2] = T[6] * R[6] + T[7] * R[7] + T[8] * R[8];
    Ic[3] = T[0] * R[3] + T[1] * R[4] + T[2] * R[5];
    Ic[4] = T[0] * R[6] + T[1] * R[7] + T[2] * R[8];
    Ic[5] = T[3] * R[6] + T[4] * R[7] + T[5] * R[8];
    real mass = GM().ab_mass[a], hh = dot3(h, h);
    I[0] = mass; I[1] = mass * h[0]; I[2] = mass * h[1]; I[3] = mass * h[2];
    I[4] = Ic[0] + mass * (hh - h[0] * h[0]); I[5] = Ic[1] + mass * (hh - h[1] * h[1]); I[6] = Ic[2] + mass * (hh - h[2] * h[2]);
    I[7] = Ic[3] - mass * h[0] * h[1]; I[8] = Ic[4] - mass * h[0] * h[2]; I[9] = Ic[5] - mass * h[1] * h[2];
    real fa[6], iv[6], g[6];
    spi_mul(fa, I, ab);
    spi_mul(iv, I, vb);
    cross_force(g, vb, iv);
    // net gravity (after compensation) as a spatial force at the com
    real Fg[3] = {GM().ab_gfac[a] * mass * A.gravity[0], GM().ab_gfac[a] * mass * A.gravity[1], GM().ab_gfac[a] * mass * A.gravity[2]}, tg[3];
    cross3(tg, h, Fg);
#pragma unroll
    for (int k = 0; k < 3; k++) { f[k] = fa[k] + g[k] - tg[k]; f[3 + k] = fa[3 + k] + g[3 + k] - Fg[k]; }
  }

  // -------------------------------------------- [P2,P6,P7] CRB, RNE bias, actuation
  // Leaves Marm (joint-space inertia of the chain) and fsm = qfrc_smooth =
  // qfrc_passive - qfrc_bias + qfrc_actuator for every dof.  Gravity compensation
  // (robot.py:81-82) is folded into the RNE gravity term: each body feels
  // (1 - gravcomp) * m * g.  Returns the actuator "force is clamped" flag (lane j).
  // Every inertial-body lane (6 links + <= 2 fixed attachments) walks the chain up to its own link in registers -- motion axes and
  // joint rates come from shared memory -- instead of prefix scans through shuffles; the bodies' spatial inertias and forces go
  // through a shared-memory scratch (the contact Jacobian block, free until make_rows) and every link lane sums its sub-chain.
  // No shuffle in the stage: 10.2 k -> see profiles/r2/stage_cycles.txt.
  __device__ static bool dynamics(const AM& A, WM& W, int lane) {
    real Sj[6] = {0, 0, 0, 0, 0, 0};
    const int mylink = lane < A.nab ? GM().ab_link[lane] : -1;
    real vb[6] = {0, 0, 0, 0, 0, 0}, ab[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll 1
    for (int j = 0; j <= mylink; j++) {
      real s[6], sd[6];
      const real qd = (real)W.qvel[j];
#pragma unroll
      for (int k = 0; k < 6; k++) s[k] = W.S[j][k];
      cross_motion(sd, vb, s);  // uses the velocity of the parent link
#pragma unroll
      for (int k = 0; k < 6; k++) { ab[k] += sd[k] * qd; vb[k] += s[k] * qd; }
    }
    if (lane < ARM_NLINK) {
#pragma unroll
      for (int k = 0; k < 6; k++) Sj[k] = W.S[lane][k];
    }
    real I[10], f[6];
    if (lane < A.nab) {
      body_inertia_force(A, W, lane, mylink, vb, ab, I, f);
#pragma unroll
      for (int k = 0; k < 10; k++) W.cJ[lane][k] = I[k];
#pragma unroll
      for (int k = 0; k < 6; k++) W.cJ[lane][10 + k] = f[k];
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 10; k++) I[k] = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) f[k] = 0;
    if (lane < ARM_NLINK) {
#pragma unroll 1
      for (int a = 0; a < A.nab; a++) {
        if (GM().ab_link[a] >= lane) {
#pragma unroll
          for (int k = 0; k < 10; k++) I[k] += W.cJ[a][k];
#pragma unroll
          for (int k = 0; k < 6; k++) f[k] += W.cJ[a][10 + k];
        }
      }
    }
    bool clamped = false;
    if (lane < ARM_NLINK) {
      real tau = 0;
#pragma unroll
      for (int k = 0; k < 6; k++) tau += Sj[k] * f[k];
      real y[6];
      spi_mul(y, I, Sj);
#pragma unroll ROLL
      for (int j = 0; j < ARM_NLINK; j++) {
        real d = 0;
#pragma unroll
        for (int k = 0; k < 6; k++) d += W.S[j][k] * y[k];
        if (j <= lane) {
          if (j == lane) d += GM().armature[lane];
          W.Marm[lane][j] = d;
          W.Marm[j][lane] = d;
        }
      }
      // [P7] position actuator in double: kp*(ctrl - q) - kv*qdot, clamped to forcerange
      double ctrl = W.ctrl[lane];
      if (GM().ctrllimited[lane]) ctrl = fmin(fmax(ctrl, GM().ctrl_lo[lane]), GM().ctrl_hi[lane]);
      double frc = GM().kp[lane] * (ctrl - W.qpos[lane]) - GM().kv[lane] * W.qvel[lane];
      if (GM().forcelimited[lane]) {
        if (frc <= GM().frc_lo[lane]) { frc = GM().frc_lo[lane]; clamped = true; }
        if (frc >= GM().frc_hi[lane]) { frc = GM().frc_hi[lane]; clamped = true; }
      }
      W.fsm[lane] = (real)frc - tau;
    } else if (lane < NV) {
      int k = (lane - ARM_NLINK) / 6, d = (lane - ARM_NLINK) % 6;
      if (d < 3) W.fsm[lane] = GM().obj_mass[k] * GM().gravity[d];
      else {
        int b = ARM_NLINK + 6 * k + 3;
        real w[3] = {(real)W.qvel[b], (real)W.qvel[b + 1], (real)W.qvel[b + 2]};
        real Iw[3] = {GM().obj_I[k][0] * w[0], GM().obj_I[k][1] * w[1], GM().obj_I[k][2] * w[2]}, c[3];
        cross3(c, w, Iw);
        W.fsm[lane] = -c[d - 3];
      }
    }
    if (lane < NV) W.qv[lane] = (real)W.qvel[lane];
    __syncwarp();
    return clamped;
  }

  // ------------------------------------------------------------- [P4] collision
  // One lane per candidate pair: bounding-sphere broadphase, analytic narrowphase
  // (same routines as oracle/mjs_oracle.cpp), warp-scan compaction into W.craw.
  // Joint-limit rows ride along as pseudo-contacts (cpair = -1 - joint).
  // Contact slot N (0..3, a run-time value) is written through a select chain over static indices so that cd[] / cp[][]
  // stay in registers: a dynamically indexed array lives in the thread's stack frame, and the stack traffic of 384
  // threads does not fit the 24 KB of L1 left beside 231 KB of shared memory.
#define MJS_CPUT(N, D, X, Y, Z)                                                                       \
  do {                                                                                                \
    const real _d = (D), _x = (X), _y = (Y), _z = (Z);                                                \
    _Pragma("unroll") for (int _s = 0; _s < 4; _s++)                                                  \
      if (_s == (N)) { cd[_s] = _d; cp[_s][0] = _x; cp[_s][1] = _y; cp[_s][2] = _z; }                 \
  } while (0)
  __device__ static void collide(const AM& A, WM& W, int lane, const SepAxis<real>& sep, bool& cvx) {
    cvx = false;
    int n = 0;
    // (zero-initialised on purpose: ptxas keeps the arrays in registers only when every slot has a definite value)
    real cd[4] = {0, 0, 0, 0}, cp[4][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, nrm[3] = {0, 0, 1}, yh[3] = {0, 0, 0};
    int code = lane;  // pair index; limits use -1 - j
    unsigned valid = 0;  // slots that become contacts (bit k = slot k); their order is kept
    if (lane < A.npair) {
      int g1 = GM().pair_g1[lane], g2 = GM().pair_g2[lane];
      int t1 = GM().geom_type[g1], t2 = GM().geom_type[g2];
      real margin = GM().pair_margin[lane];
      const real *p1 = W.gpos[g1], *p2 = W.gpos[g2], *m1 = W.gmat[g1], *m2 = W.gmat[g2];
      real s2[3] = {GM().geom_size[g2][0], GM().geom_size[g2][1], GM().geom_size[g2][2]};
      real dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
      bool pass;
      if (t1 == MJS_GEOM_PLANE) {
        nrm[0] = m1[2]; nrm[1] = m1[5]; nrm[2] = m1[8];
        pass = dot3(dif, nrm) - GM().geom_rbound[g2] <= margin;
      } else {
        pass = sqrt(dot3(dif, dif)) <= GM().geom_rbound[g1] + GM().geom_rbound[g2] + margin;
      }
      if (pass) {
        if (t1 == MJS_GEOM_PLANE && t2 == MJS_GEOM_SPHERE) {
          real dist = dot3(dif, nrm) - s2[0];
          if (dist <= margin) { cd[0] = dist; for (int k = 0; k < 3; k++) cp[0][k] = p2[k] - nrm[k] * (s2[0] + dist * (real)0.5); n = 1; }
        } else if (t1 == MJS_GEOM_PLANE && t2 == MJS_GEOM_CAPSULE) {
          real ax[3] = {m2[2], m2[5], m2[8]};
          yh[0] = ax[0]; yh[1] = ax[1]; yh[2] = ax[2];
#pragma unroll
          for (int e = 0; e < 2; e++) {
            real sg = e == 0 ? s2[1] : -s2[1];
            real end[3] = {p2[0] + ax[0] * sg, p2[1] + ax[1] * sg, p2[2] + ax[2] * sg};
            real d3[3] = {end[0] - p1[0], end[1] - p1[1], end[2] - p1[2]};
            real dist = dot3(d3, nrm) - s2[0];
            if (dist <= margin) { const real hh = s2[0] + dist * (real)0.5; MJS_CPUT(n, dist, end[0] - nrm[0] * hh, end[1] - nrm[1] * hh, end[2] - nrm[2] * hh); n++; }
          }
        } else if (t1 == MJS_GEOM_PLANE && t2 == MJS_GEOM_CYLINDER) {
          real ax[3] = {m2[2], m2[5], m2[8]};
          real prjaxis = dot3(nrm, ax);
          if (prjaxis > 0) { ax[0] = -ax[0]; ax[1] = -ax[1]; ax[2] = -ax[2]; prjaxis = -prjaxis; }
          real dist0 = dot3(dif, nrm);
          real vec[3] = {ax[0] * prjaxis - nrm[0], ax[1] * prjaxis - nrm[1], ax[2] * prjaxis - nrm[2]};
          real len_sqr = dot3(vec, vec);
          if (len_sqr >= (real)(MJS_CYL_FLAT_TILT * MJS_CYL_FLAT_TILT)) {
            real s = s2[0] / sqrt(len_sqr);
            vec[0] *= s; vec[1] *= s; vec[2] *= s;
          } else {
            vec[0] = m2[0] * s2[0]; vec[1] = m2[3] * s2[0]; vec[2] = m2[6] * s2[0];
          }
          real prjvec = dot3(vec, nrm);
          ax[0] *= s2[1]; ax[1] *= s2[1]; ax[2] *= s2[1];
          prjaxis *= s2[1];
          if (dist0 + prjaxis + prjvec <= margin) {
            real d = dist0 + prjaxis + prjvec;
            cd[0] = d;
            for (int k = 0; k < 3; k++) cp[0][k] = p2[k] + vec[k] + ax[k] - nrm[k] * d * (real)0.5;
            n = 1;
            if (dist0 - prjaxis + prjvec <= margin) {
              d = dist0 - prjaxis + prjvec;
              cd[1] = d;
              for (int k = 0; k < 3; k++) cp[1][k] = p2[k] + vec[k] - ax[k] - nrm[k] * d * (real)0.5;
              n = 2;
            }
            real prjvec1 = -prjvec * (real)0.5;
            if (dist0 + prjaxis + prjvec1 <= margin && n <= 2) {
              real v1[3];
              cross3(v1, vec, ax);
              real l = sqrt(dot3(v1, v1));
              if (l < (real)MJS_MINVAL) { v1[0] = 1; v1[1] = 0; v1[2] = 0; } else { v1[0] /= l; v1[1] /= l; v1[2] /= l; }
              real s = s2[0] * (real)0.8660254037844386;  // sqrt(3)/2
              d = dist0 + prjaxis + prjvec1;
              real pa[3], pb[3];
#pragma unroll
              for (int k = 0; k < 3; k++) {
                pa[k] = p2[k] + v1[k] * s + ax[k] - vec[k] * (real)0.5 - nrm[k] * d * (real)0.5;
                pb[k] = p2[k] + v1[k] * -s + ax[k] - vec[k] * (real)0.5 - nrm[k] * d * (real)0.5;
              }
              if (n == 1) { cd[1] = d; cd[2] = d; for (int k = 0; k < 3; k++) { cp[1][k] = pa[k]; cp[2][k] = pb[k]; } }
              else { cd[2] = d; cd[3] = d; for (int k = 0; k < 3; k++) { cp[2][k] = pa[k]; cp[3][k] = pb[k]; } }
              n += 2;
            }
          }
        } else if (t1 == MJS_GEOM_PLANE && t2 == MJS_GEOM_BOX) {
          real dist = dot3(dif, nrm);
          for (int i = 0; i < 8 && n < 4; i++) {
            real vloc[3] = {(i & 1) ? s2[0] : -s2[0], (i & 2) ? s2[1] : -s2[1], (i & 4) ? s2[2] : -s2[2]}, corner[3];
            mat_vec(corner, m2, vloc);
            real ldist = dot3(nrm, corner);
            if (dist + ldist > margin || ldist > 0) continue;
            const real dc = dist + ldist;
            MJS_CPUT(n, dc, corner[0] + p2[0] - nrm[0] * dc * (real)0.5, corner[1] + p2[1] - nrm[1] * dc * (real)0.5, corner[2] + p2[2] - nrm[2] * dc * (real)0.5);
            n++;
          }
        } else {
          // one-contact pairs: reduce to "sphere(q1,r1) vs sphere(q2,r2)" or an explicit (dist, normal)
          real q1[3] = {p1[0], p1[1], p1[2]}, q2[3] = {p2[0], p2[1], p2[2]}, r1 = 0, r2 = 0;
          real s1[3] = {GM().geom_size[g1][0], GM().geom_size[g1][1], GM().geom_size[g1][2]};
          int mode = 1;  // 1: spheres, 2: explicit result in (cd[0], nrm, cp[0]), 0: nothing
          bool needs_convex = false;  // the pair goes to the general convex routine (rare; called once, below)
          if (t1 == MJS_GEOM_SPHERE && t2 == MJS_GEOM_SPHERE) { r1 = s1[0]; r2 = s2[0]; }
          else if (t1 == MJS_GEOM_SPHERE && t2 == MJS_GEOM_CAPSULE) {
            real ax[3] = {m2[2], m2[5], m2[8]}, vv[3] = {-dif[0], -dif[1], -dif[2]};
            real x = fmin(fmax(dot3(ax, vv), -s2[1]), s2[1]);
            for (int k = 0; k < 3; k++) q2[k] = p2[k] + ax[k] * x;
            r1 = s1[0]; r2 = s2[0];
          } else if (t1 == MJS_GEOM_CYLINDER && t2 == MJS_GEOM_CYLINDER) {
            real a1[3] = {m1[2], m1[5], m1[8]}, a2[3] = {m2[2], m2[5], m2[8]};
            const real sepd[3] = {(real)sep.v[0], (real)sep.v[1], (real)sep.v[2]};
            real tc = dot3(a1, dif), ta = fabs(dot3(a1, a2) * s2[1]);
            real lo = fmax(-s1[1], tc - ta), hi = fmin(s1[1], tc + ta);
            if (lo > hi) {
              // no axial overlap of the axis segments: only a cap configuration can touch -> general convex routine
              needs_convex = !cyl_separated<real>(dif, a1, s1[0], s1[1], a2, s2[0], s2[1], dif) &&
                             !cyl_axis_separates<real>(dif, a1, s1[0], s1[1], a2, s2[0], s2[1], sepd);
              mode = 0;
            } else {
              real sm = (real)0.5 * (lo + hi);
              for (int k = 0; k < 3; k++) q1[k] = p1[k] + a1[k] * sm;
              real w3[3] = {q1[0] - p2[0], q1[1] - p2[1], q1[2] - p2[2]};
              real u = fmin(fmax(dot3(a2, w3), -s2[1]), s2[1]);
              for (int k = 0; k < 3; k++) q2[k] = p2[k] + a2[k] * u;
              r1 = s1[0]; r2 = s2[0];
              // side-side is only right while the radial direction is the axis of least penetration (oracle cylinder_cylinder)
              real dq[3] = {q2[0] - q1[0], q2[1] - q1[1], q2[2] - q1[2]};
              real radial_pen = r1 + r2 - sqrt(dot3(dq, dq));
              if (radial_pen > 0) {  // (rare: only for pairs in contact)
                // extent of cylinder 2 along axis 1 (axis segment + rim), so that a prop lying on its side still overlaps axially
                real ca = dot3(a1, a2), ext = fabs(ca) * s2[1] + s2[0] * sqrt(fmax((real)0, 1 - ca * ca));
                real axial_pen = fmin(s1[1], tc + ext) - fmax(-s1[1], tc - ext);
                if (axial_pen < radial_pen) {  // cap-cap / cap-rim / cap-side: the general convex routine
                  needs_convex = !cyl_separated<real>(dif, a1, s1[0], s1[1], a2, s2[0], s2[1], dq) &&
                                 !cyl_axis_separates<real>(dif, a1, s1[0], s1[1], a2, s2[0], s2[1], sepd);
                  mode = 0;
                }
              }
            }
          } else if (t1 == MJS_GEOM_SPHERE && t2 == MJS_GEOM_CYLINDER) {
            real ax[3] = {m2[2], m2[5], m2[8]}, vv[3] = {-dif[0], -dif[1], -dif[2]};
            real x = dot3(vv, ax);
            real perp[3] = {vv[0] - ax[0] * x, vv[1] - ax[1] * x, vv[2] - ax[2] * x};
            real plen = sqrt(dot3(perp, perp));
            bool inside_h = fabs(x) <= s2[1], inside_r = plen <= s2[0];
            real sgn = x >= 0 ? (real)1 : (real)-1;
            r1 = s1[0];
            if (inside_h && !inside_r) {
              for (int k = 0; k < 3; k++) q2[k] = p2[k] + ax[k] * x;
              r2 = s2[0];
            } else if (inside_r) {
              real dist = fabs(x) - s2[1] - r1;
              mode = 0;
              if (dist <= margin) {
                mode = 2; cd[0] = dist;
                for (int k = 0; k < 3; k++) { nrm[k] = -sgn * ax[k]; cp[0][k] = p1[k] + nrm[k] * (r1 + (real)0.5 * dist); }
              }
            } else {
              for (int k = 0; k < 3; k++) q2[k] = p2[k] + ax[k] * sgn * s2[1] + perp[k] * (s2[0] / plen);
              r2 = 0;
            }
          } else if (t1 == MJS_GEOM_SPHERE && t2 == MJS_GEOM_BOX) {
            real vv[3] = {-dif[0], -dif[1], -dif[2]}, loc[3], cl[3];
            matT_vec(loc, m2, vv);
            bool inside = true;
            for (int k = 0; k < 3; k++) { cl[k] = fmin(fmax(loc[k], -s2[k]), s2[k]); if (cl[k] != loc[k]) inside = false; }
            r1 = s1[0];
            if (!inside) {
              mat_vec(q2, m2, cl);
              for (int k = 0; k < 3; k++) q2[k] += p2[k];
              r2 = 0;
            } else {
              int best = 0; real bd = (real)1e30;
              for (int k = 0; k < 3; k++) { real dd = s2[k] - fabs(loc[k]); if (dd < bd) { bd = dd; best = k; } }
              real nl[3] = {0, 0, 0};
              nl[best] = loc[best] >= 0 ? (real)-1 : (real)1;
              mat_vec(nrm, m2, nl);
              mode = 2; cd[0] = -bd - r1;
              for (int k = 0; k < 3; k++) cp[0][k] = p1[k] + nrm[k] * (r1 + (real)0.5 * cd[0]);
            }
          } else if (t1 == MJS_GEOM_CAPSULE && t2 == MJS_GEOM_CAPSULE) {
            // mjc_CapsuleCapsule (oracle capsule_capsule): closest points of the two axis segments; parallel axes give the
            // two end-point tests of segment 1
            real a1[3] = {m1[2], m1[5], m1[8]}, a2[3] = {m2[2], m2[5], m2[8]}, df[3] = {-dif[0], -dif[1], -dif[2]};
            real ma = dot3(a1, a1), mb = -dot3(a1, a2), mc = dot3(a2, a2), u = -dot3(a1, df), v = dot3(a2, df);
            real det = ma * mc - mb * mb;
            r1 = s1[0]; r2 = s2[0];
            if (fabs(det) >= (real)MJS_MINVAL) {
              real x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
              if (x1 > s1[1]) { x1 = s1[1]; x2 = (v - mb * s1[1]) / mc; }
              else if (x1 < -s1[1]) { x1 = -s1[1]; x2 = (v + mb * s1[1]) / mc; }
              if (x2 > s2[1]) { x2 = s2[1]; x1 = (u - mb * s2[1]) / ma; }
              else if (x2 < -s2[1]) { x2 = -s2[1]; x1 = (u + mb * s2[1]) / ma; }
              x1 = fmin(fmax(x1, -s1[1]), s1[1]);
              x2 = fmin(fmax(x2, -s2[1]), s2[1]);
              for (int k = 0; k < 3; k++) { q1[k] = p1[k] + a1[k] * x1; q2[k] = p2[k] + a2[k] * x2; }
            } else {
              mode = 0;
#pragma unroll
              for (int e2 = 0; e2 < 2; e2++) {
                real x1 = e2 == 0 ? s1[1] : -s1[1];
                real x2 = fmin(fmax((v - mb * x1) / mc, -s2[1]), s2[1]);
                real qa[3], qb[3], d3[3];
                for (int k = 0; k < 3; k++) { qa[k] = p1[k] + a1[k] * x1; qb[k] = p2[k] + a2[k] * x2; d3[k] = qb[k] - qa[k]; }
                real len = sqrt(dot3(d3, d3)), dist = len - r1 - r2;
                if (dist <= margin) {
                  if (len < (real)MJS_MINVAL) { d3[0] = 1; d3[1] = 0; d3[2] = 0; } else { d3[0] /= len; d3[1] /= len; d3[2] /= len; }
                  for (int k = 0; k < 3; k++) nrm[k] = d3[k];  // parallel axes: both end-point contacts share the normal
                  MJS_CPUT(n, dist, qa[0] + d3[0] * (r1 + (real)0.5 * dist), qa[1] + d3[1] * (r1 + (real)0.5 * dist), qa[2] + d3[2] * (r1 + (real)0.5 * dist));
                  n++;
                }
              }
            }
          } else if ((t1 == MJS_GEOM_CAPSULE || t1 == MJS_GEOM_CYLINDER) && (t2 == MJS_GEOM_CYLINDER || t2 == MJS_GEOM_BOX)) {
            needs_convex = true;
            mode = 0;
          } else {
            atomicOr(&W.status, MJS_STATUS_UNSUPPORTED_PAIR);
            mode = 0;
          }
          // The general convex routine is NOT called from here: a call inside this function costs the hot path registers
          // (measured: 2.33 -> 1.85 M env-steps/s with the call here, whatever the call rate).  The lane only reports the
          // pair; substep() runs convex_fixup() after collide, when the narrowphase's registers are dead.
          if (needs_convex) cvx = true;
          if (mode == 1) {
            real d3[3] = {q2[0] - q1[0], q2[1] - q1[1], q2[2] - q1[2]};
            real len = sqrt(dot3(d3, d3)), dist = len - r1 - r2;
            if (dist <= margin) {
              if (len < (real)MJS_MINVAL) { d3[0] = 1; d3[1] = 0; d3[2] = 0; } else { d3[0] /= len; d3[1] /= len; d3[2] /= len; }
              cd[0] = dist;
              for (int k = 0; k < 3; k++) { nrm[k] = d3[k]; cp[0][k] = q1[k] + d3[k] * (r1 + (real)0.5 * dist); }
              n = 1;
            }
          } else if (mode == 2) n = 1;
        }
        // mj_collision: contacts with dist >= margin - gap produce no rows
        real incl = GM().pair_incl[lane];
#pragma unroll
        for (int k = 0; k < 4; k++)
          if (k < n && cd[k] < incl) valid |= 1u << k;
        n = __popc(valid);
      }
    } else if (lane >= A.npair && lane < A.npair + ARM_NLINK) {
      // joint limits (mj_instantiateLimit)
      int j = lane - A.npair;
      if (GM().jnt_limited[j]) {
        real q = (real)W.qpos[j];
        real dlo = q - GM().jnt_lo[j], dhi = GM().jnt_hi[j] - q;
        if (dlo < GM().jnt_margin[j]) { n = 1; valid = 1; cd[0] = dlo; cp[0][0] = 1; }
        else if (dhi < GM().jnt_margin[j]) { n = 1; valid = 1; cd[0] = dhi; cp[0][0] = -1; }
        cp[0][1] = 0; cp[0][2] = 0;
        code = -1 - j;
      }
    }
    // warp exclusive scan of n
    int incl_sum = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(FULLMASK, incl_sum, o); if (lane >= o) incl_sum += t; }
    int base = incl_sum - n;
    int total = __shfl_sync(FULLMASK, incl_sum, 31);
    if (total > ARM_MAXC) { if (lane == 0) atomicOr(&W.status, MJS_STATUS_CONTACT_OVERFLOW); total = ARM_MAXC; }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int c = base + __popc(valid & ((1u << k) - 1u));
      if (((valid >> k) & 1u) && c < ARM_MAXC) {
        W.craw[0][c] = cd[k];
        W.craw[1][c] = cp[k][0]; W.craw[2][c] = cp[k][1]; W.craw[3][c] = cp[k][2];
        W.craw[4][c] = nrm[0]; W.craw[5][c] = nrm[1]; W.craw[6][c] = nrm[2];
        W.craw[7][c] = yh[0]; W.craw[8][c] = yh[1]; W.craw[9][c] = yh[2];
        W.cpair[c] = code;
      }
    }
    if (lane == 0) { W.ncon = total; W.coup = 0; W.objcoup = 0; }
    if (lane <= ARM_MAXOBJ) W.bcount[lane] = 0;
    __syncwarp();
  }

  // per-contact registers that live across the Newton iterations
  struct ConReg {
    real D, mu[3], r0, rk[3], e0, ek[3], de0, dek[3];
    int dim;  // 0: lane has no contact
  };

  // ----------------------------------------------------------- [P5] constraint rows
  // Lane c builds contact c: frame, contact-frame Jacobian blocks (side 0 = body1
  // with sign -1, side 1 = body2), impedance -> D, reference acceleration.
  __device__ static void make_rows(const AM& A, WM& W, int lane, ConReg& C) {
    C.dim = 0;
    C.D = 0; C.r0 = 0; C.e0 = 0; C.de0 = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) { C.mu[k] = 0; C.rk[k] = 0; C.ek[k] = 0; C.dek[k] = 0; }
    int ncon = W.ncon;
    if (lane < ncon) {
      int c = lane;
      int code = W.cpair[c];
      real dist = W.craw[0][c];
#pragma unroll 4
      for (int e = 0; e < 48; e++) W.cJ[e][c] = 0;
      real u[4] = {0, 0, 0, 0};
      real kk, bb, imp, incl, diag0;
      if (code < 0) {
        int j = -1 - code;
        real sgn = W.craw[1][c];
        W.cJ[j][c] = sgn;  // side 0, row 0, column j
        W.cblk[0][c] = 0; W.ccols[0][c] = ARM_NLINK; W.cblk[1][c] = -1; W.ccols[1][c] = 0;
        u[0] = sgn * W.qv[j];
        kk = GM().jnt_kb[j][0]; bb = GM().jnt_kb[j][1];
        incl = GM().jnt_margin[j];
        imp = mjs_impedance<real>(GM().jnt_solimp[j], dist, incl);
        diag0 = GM().jnt_diag[j];
        C.dim = 1;
      } else {
        int p = code;
        real fr[9] = {W.craw[4][c], W.craw[5][c], W.craw[6][c], W.craw[7][c], W.craw[8][c], W.craw[9][c], 0, 0, 0};
        real pos[3] = {W.craw[1][c], W.craw[2][c], W.craw[3][c]};
        // mju_makeFrame
        {
          real nn = rsqrt_r<real>(dot3(fr, fr));
          fr[0] *= nn; fr[1] *= nn; fr[2] *= nn;
          if (dot3(fr + 3, fr + 3) < (real)0.25) {
            fr[3] = 0; fr[4] = 0; fr[5] = 0;
            if (fr[1] < (real)0.5 && fr[1] > (real)-0.5) fr[4] = 1; else fr[5] = 1;
          }
          real t = dot3(fr, fr + 3);
          fr[3] -= t * fr[0]; fr[4] -= t * fr[1]; fr[5] -= t * fr[2];
          nn = rsqrt_r<real>(dot3(fr + 3, fr + 3));
          fr[3] *= nn; fr[4] *= nn; fr[5] *= nn;
          cross3(fr + 6, fr, fr + 3);
        }
        int dim = GM().pair_dim[p];
        C.dim = dim;
        int nrow = dim == 1 ? 1 : dim;  // contact-space rows: normal, 2 tangents, (torsion)
#pragma unroll ROLL
        for (int side = 0; side < 2; side++) {
          int g = side == 0 ? GM().pair_g1[p] : GM().pair_g2[p];
          int slot = GM().geom_slot[g];
          real sg = side == 0 ? (real)-1 : (real)1;
          int blk = -1, cols = 0;
          // Row loops are unrolled with `k < nrow` guards so that fr[] and u[] are indexed statically (registers, not the
          // thread's stack frame), and the lever-arm cross product of a column is formed once, not once per row.
          if (slot >= SLOT_LINK0 && slot < SLOT_LINK0 + ARM_NLINK) {
            blk = 0; cols = slot - SLOT_LINK0 + 1;
            for (int j = 0; j < cols; j++) {
              real Sa[3] = {W.S[j][0], W.S[j][1], W.S[j][2]}, jp[3];
              cross3(jp, Sa, pos);
              jp[0] += W.S[j][3]; jp[1] += W.S[j][4]; jp[2] += W.S[j][5];
              const real w = W.qv[j];
#pragma unroll
              for (int k = 0; k < 4; k++) {
                if (k < nrow) {
                  real val = k < 3 ? dot3(fr + 3 * k, jp) : dot3(fr, Sa);
                  W.cJ[(side * 4 + k) * 6 + j][c] = sg * val;
                  u[k] += sg * val * w;
                }
              }
            }
          } else if (slot >= SLOT_OBJ0 && slot < SLOT_OBJ0 + ARM_MAXOBJ) {
            int ob = slot - SLOT_OBJ0;
            blk = 1 + ob; cols = 6;
            const real* R = W.xmat[slot];
            real r[3] = {pos[0] - W.xpos[slot][0], pos[1] - W.xpos[slot][1], pos[2] - W.xpos[slot][2]};
            int b0 = ARM_NLINK + 6 * ob;
#pragma unroll
            for (int k = 0; k < 4; k++) {
              if (k < nrow) {
#pragma unroll
                for (int j = 0; j < 3; j++) {
                  real val = k < 3 ? fr[3 * k + j] : (real)0;
                  W.cJ[(side * 4 + k) * 6 + j][c] = sg * val;
                  u[k] += sg * val * W.qv[b0 + j];
                }
              }
            }
#pragma unroll ROLL
            for (int j = 0; j < 3; j++) {
              real Ra[3] = {R[j], R[3 + j], R[6 + j]}, jp[3];
              cross3(jp, Ra, r);
              const real w = W.qv[b0 + 3 + j];
#pragma unroll
              for (int k = 0; k < 4; k++) {
                if (k < nrow) {
                  real val = k < 3 ? dot3(fr + 3 * k, jp) : dot3(fr, Ra);
                  W.cJ[(side * 4 + k) * 6 + 3 + j][c] = sg * val;
                  u[k] += sg * val * w;
                }
              }
            }
          }
          W.cblk[side][c] = blk; W.ccols[side][c] = cols;
        }
        C.mu[0] = GM().pair_mu[p][0]; C.mu[1] = GM().pair_mu[p][0]; C.mu[2] = GM().pair_mu[p][1];
        kk = GM().pair_kb[p][0]; bb = GM().pair_kb[p][1];
        incl = GM().pair_incl[p];
        imp = mjs_impedance<real>(GM().pair_solimp[p], dist, incl);
        diag0 = dim == 1 ? GM().pair_tran[p] : GM().pair_tran[p] * (1 + C.mu[0] * C.mu[0]);
      }
      real R0 = fmax((real)MJS_MINVAL, (1 - imp) * diag0 / imp);
      if (C.dim == 1) C.D = 1 / R0;
      else {
        real mu = C.mu[0] * rsqrt_r<real>(A.impratio);
        C.D = 1 / (2 * mu * mu * R0);
      }
      // aref rows: r0 +- rk ; jar rows: e0 +- ek with e = (J a) - aref
      C.r0 = -bb * u[0] - kk * imp * (dist - incl);
#pragma unroll
      for (int k = 0; k < 3; k++) C.rk[k] = -bb * C.mu[k] * u[1 + k];
      // block coupling pattern of the Hessian
      {
        int ba = W.cblk[0][c], bb = W.cblk[1][c];
        if (ba >= 0 && bb >= 0 && ba != bb) {
          if (ba == 0) atomicOr(&W.coup, 1 << (bb - 1));
          else if (bb == 0) atomicOr(&W.coup, 1 << (ba - 1));
          else atomicOr(&W.objcoup, 1);
        }
      }
      // register this contact with the dof blocks it touches
#pragma unroll ROLL
      for (int side = 0; side < 2; side++) {
        int blk = W.cblk[side][c];
        if (blk >= 0) {
          int slotn = atomicAdd(&W.bcount[blk], 1);
          if (slotn < ARM_MAXCB) W.blist[blk][slotn] = c | (side << 8);
          else atomicOr(&W.status, MJS_STATUS_CONTACT_OVERFLOW);
        }
      }
    }
    __syncwarp();
    if (lane <= ARM_MAXOBJ && W.bcount[lane] > ARM_MAXCB) W.bcount[lane] = ARM_MAXCB;
    __syncwarp();
  }

  // -------------------------------------------------------------- [P9] Newton solve
  // contact lane: forces, arrow weights and cost of the current e = J a - aref
  __device__ static __forceinline__ real contact_eval(WM& W, int lane, const ConReg& C, bool write) {
    real cost = 0;
    if (C.dim > 0) {
      real f0 = 0, fk[3] = {0, 0, 0}, w00 = 0, w0k[3] = {0, 0, 0}, wkk[3] = {0, 0, 0};
      if (C.dim == 1) {
        real jar = C.e0;
        if (jar < 0) { f0 = -C.D * jar; w00 = C.D; cost = (real)0.5 * C.D * jar * jar; }
      } else {
#pragma unroll
        for (int k = 0; k < 3; k++) {
          if (k < C.dim - 1) {
            real jp = C.e0 + C.ek[k], jm = C.e0 - C.ek[k];
            real ap = jp < 0 ? (real)1 : (real)0, am = jm < 0 ? (real)1 : (real)0;
            real fp = -C.D * jp * ap, fm = -C.D * jm * am;
            f0 += fp + fm;
            fk[k] = C.mu[k] * (fp - fm);
            cost += (real)0.5 * C.D * (jp * jp * ap + jm * jm * am);
            w00 += C.D * (ap + am);
            w0k[k] = C.D * C.mu[k] * (ap - am);
            wkk[k] = C.D * C.mu[k] * C.mu[k] * (ap + am);
          }
        }
      }
      if (write) {
        W.cf[0][lane] = f0; W.cf[1][lane] = fk[0]; W.cf[2][lane] = fk[1]; W.cf[3][lane] = fk[2];
        W.cW[0][lane] = w00; W.cW[1][lane] = w0k[0]; W.cW[2][lane] = w0k[1]; W.cW[3][lane] = w0k[2];
        W.cW[4][lane] = wkk[0]; W.cW[5][lane] = wkk[1]; W.cW[6][lane] = wkk[2];
      }
    }
    return cost;
  }

  // contact lane: u = Jc * x for a dof vector x in shared memory (contact space: normal, t1, t2, torsion)
  __device__ static __forceinline__ void contact_apply(const WM& W, int lane, const ConReg& C, const real* x, real* u) {
    u[0] = u[1] = u[2] = u[3] = 0;
    if (C.dim > 0) {
#pragma unroll ROLL
      for (int side = 0; side < 2; side++) {
        int blk = W.cblk[side][lane];
        if (blk < 0) continue;
        int b0 = blk_base(blk), cols = W.ccols[side][lane];
        for (int j = 0; j < cols; j++) {
          real xj = x[b0 + j];
#pragma unroll
          for (int k = 0; k < 4; k++) u[k] += W.cJ[(side * 4 + k) * 6 + j][lane] * xj;
        }
      }
    }
  }

  // dof lane: (J^T f)_i over the contacts of this dof's block
  __device__ static __forceinline__ real jt_force(const WM& W, int myblk, int d) {
    real s = 0;
    int n = W.bcount[myblk];
    for (int r = 0; r < n; r++) {
      int e = W.blist[myblk][r], c = e & 0xff, side = e >> 8;
#pragma unroll
      for (int k = 0; k < 4; k++) s += W.cJ[(side * 4 + k) * 6 + d][c] * W.cf[k][c];
    }
    return s;
  }

  // ---- 6x6 systems in registers (lower triangle, (i, j) -> i (i + 1) / 2 + j).  Every lane of a dof block factorises the whole
  // block redundantly: ~90 dependent FP operations without a shuffle, a shared-memory round trip or a __syncwarp, instead of
  // six cooperative column steps of ~150 cycles each.  The idle lanes pay nothing; the block's latency drops ~3x.
  __device__ static __forceinline__ void chol6(real* h, real* rd) {
#pragma unroll
    for (int cc = 0; cc < 6; cc++) {
#pragma unroll
      for (int i = cc; i < 6; i++) {
        real s = h[i * (i + 1) / 2 + cc];
#pragma unroll
        for (int j = 0; j < cc; j++) s -= h[i * (i + 1) / 2 + j] * h[cc * (cc + 1) / 2 + j];
        if (i == cc) { real r = rsqrt_r<real>(s); rd[cc] = r; h[i * (i + 1) / 2 + cc] = s * r; }
        else h[i * (i + 1) / 2 + cc] = s * rd[cc];
      }
    }
  }
  __device__ static __forceinline__ void solve6(const real* h, const real* rd, real* y) {
#pragma unroll
    for (int i = 0; i < 6; i++) {
      real s = y[i];
#pragma unroll
      for (int j = 0; j < i; j++) s -= h[i * (i + 1) / 2 + j] * y[j];
      y[i] = s * rd[i];
    }
#pragma unroll
    for (int i = 5; i >= 0; i--) {
      real s = y[i];
#pragma unroll
      for (int j = i + 1; j < 6; j++) s -= h[j * (j + 1) / 2 + i] * y[j];
      y[i] = s * rd[i];
    }
  }

  // ---- Newton direction, block-sparse path.  Solver order: props first, arm last.
  // Without prop-prop contacts the Hessian is block-arrow: the prop blocks are
  // factorised in parallel (6 column steps for all props at once), then the arm
  // rows against the coupled props, then the 6x6 arm Schur complement.  When the
  // arm shares no contact with a prop either (coup == 0, the common case) the
  // Hessian is block-diagonal and the arm block rides in the same 6 column steps
  // and the same 6 + 6 substitution steps as the props: half the serial chain.
  __device__ static __forceinline__ real block_direction(WM& W, int lane, real b) {
    const bool isarm = lane < ARM_NLINK, isobj = lane >= ARM_NLINK && lane < NV;
    const int myc = isarm ? lane : (lane - ARM_NLINK) % 6;
    const int base = isarm ? 6 * NOBJ : ((lane - ARM_NLINK) / 6) * 6;
    const int si = base + myc;
    const int coup = W.coup;
    const bool indep = coup == 0;                   // warp-uniform
    const bool isblk = isobj || (isarm && indep);   // lanes of the parallel diagonal-block pass
    const int lane0 = isarm ? 0 : ARM_NLINK + base; // lane that owns column 0 of this lane's block
    if (indep) {
      // block-diagonal Hessian (the common case): every lane solves its own 6x6 block in registers
      real rhs[6];
#pragma unroll
      for (int j = 0; j < 6; j++) rhs[j] = shfl(b, (isarm || isobj) ? lane0 + j : lane);
      real x = 0;
      if (isarm || isobj) {
        real h[21], rd[6];
#pragma unroll
        for (int i = 0; i < 6; i++)
#pragma unroll
          for (int j = 0; j <= i; j++) h[i * (i + 1) / 2 + j] = W.H[base + i][base + j];
        chol6(h, rd);
        solve6(h, rd, rhs);
#pragma unroll
        for (int j = 0; j < 6; j++) if (j == myc) x = rhs[j];
      }
      return x;
    }
    real rdiag = 1;
    // diagonal blocks (props, and the arm when it is uncoupled)
#pragma unroll CHOLU
    for (int c = 0; c < 6; c++) {
      real s = 0;
      if (isblk && myc >= c) {
        s = W.H[si][base + c];
        for (int j = 0; j < c; j++) s -= W.H[si][base + j] * W.H[base + c][base + j];
      }
      real dk = shfl(s, isblk ? lane0 + c : lane);
      real rd = rsqrt_r<real>(dk);
      if (isblk) {
        if (myc == c) { W.H[si][si] = dk * rd; rdiag = rd; }
        else if (myc > c) W.H[si][base + c] = s * rd;
      }
      __syncwarp();
    }
    if (!indep) {
      // arm rows x coupled prop columns
      for (int ob = 0; ob < NOBJ; ob++) {
        if (!((coup >> ob) & 1)) continue;
#pragma unroll CHOLU
        for (int c = 0; c < 6; c++) {
          int k = 6 * ob + c;
          real rdk = shfl(rdiag, ARM_NLINK + k);
          if (isarm) {
            real s = W.H[si][k];
            for (int j = 0; j < c; j++) s -= W.H[si][6 * ob + j] * W.H[k][6 * ob + j];
            W.H[si][k] = s * rdk;
          }
        }
      }
      __syncwarp();
      // arm Schur complement block
#pragma unroll CHOLU
      for (int c = 0; c < 6; c++) {
        int k = 6 * NOBJ + c;
        real s = 0;
        if (isarm && myc >= c) {
          s = W.H[si][k];
          for (int ob = 0; ob < NOBJ; ob++) {
            if (!((coup >> ob) & 1)) continue;
#pragma unroll
            for (int j = 0; j < 6; j++) s -= W.H[si][6 * ob + j] * W.H[k][6 * ob + j];
          }
          for (int j = 0; j < c; j++) s -= W.H[si][6 * NOBJ + j] * W.H[k][6 * NOBJ + j];
        }
        real dk = shfl(s, c);
        real rd = rsqrt_r<real>(dk);
        if (isarm) {
          if (myc == c) { W.H[si][si] = dk * rd; rdiag = rd; }
          else if (myc > c) W.H[si][k] = s * rd;
        }
        __syncwarp();
      }
    }
    // forward substitution: diagonal blocks in parallel, then the coupled arm
#pragma unroll CHOLU
    for (int c = 0; c < 6; c++) {
      real yk = shfl(b * rdiag, isblk ? lane0 + c : lane);
      if (isblk) { if (myc == c) b = yk; else if (myc > c) b -= W.H[si][base + c] * yk; }
    }
    if (!indep) {
      if (isobj) W.dir[lane] = b;
      __syncwarp();
      if (isarm) {
        for (int ob = 0; ob < NOBJ; ob++) {
          if (!((coup >> ob) & 1)) continue;
#pragma unroll
          for (int j = 0; j < 6; j++) b -= W.H[si][6 * ob + j] * W.dir[ARM_NLINK + 6 * ob + j];
        }
      }
#pragma unroll CHOLU
      for (int c = 0; c < 6; c++) {
        real yk = shfl(b * rdiag, c);
        if (isarm) { if (myc == c) b = yk; else if (myc > c) b -= W.H[si][6 * NOBJ + c] * yk; }
      }
      // backward substitution: arm, then props
#pragma unroll CHOLU
      for (int c = 5; c >= 0; c--) {
        real xk = shfl(b * rdiag, c);
        if (isarm) { if (myc == c) b = xk; else if (myc < c) b -= W.H[6 * NOBJ + c][si] * xk; }
      }
      __syncwarp();
      if (isarm) W.dir[lane] = b;
      __syncwarp();
      if (isobj && ((coup >> (base / 6)) & 1)) {
#pragma unroll
        for (int i = 0; i < 6; i++) b -= W.H[6 * NOBJ + i][si] * W.dir[i];
      }
    }
    // backward substitution of the parallel blocks
#pragma unroll CHOLU
    for (int c = 5; c >= 0; c--) {
      real xk = shfl(b * rdiag, isblk ? lane0 + c : lane);
      if (isblk) { if (myc == c) b = xk; else if (myc < c) b -= W.H[base + c][si] * xk; }
    }
    return (isarm || isobj) ? b : (real)0;
  }

  // ---- Newton direction, dense path (prop-prop contacts present): lane = row,
  // left-looking column Cholesky over the whole NV x NV Hessian.
  __device__ static __noinline__ real dense_direction(WM& W, int lane, real b) {
    const bool isdof = lane < NV;
    const int si = isdof ? sidx(lane) : 0;
    real rdiag = 1;
    for (int k = 0; k < NV; k++) {
      real s = 0;
      if (isdof && si >= k) {
        s = W.H[si][k];
        for (int j = 0; j < k; j++) s -= W.H[si][j] * W.H[k][j];
      }
      int owner = dof_of_sidx(k);
      real dk = shfl(s, owner);
      real rd = rsqrt_r<real>(dk);
      if (isdof) {
        if (si == k) { W.H[si][k] = dk * rd; rdiag = rd; }
        else if (si > k) W.H[si][k] = s * rd;
      }
      __syncwarp();
    }
    for (int k = 0; k < NV; k++) {
      int owner = dof_of_sidx(k);
      real yk = shfl(b * rdiag, owner);
      if (isdof) { if (si == k) b = yk; else if (si > k) b -= W.H[si][k] * yk; }
    }
    for (int k = NV - 1; k >= 0; k--) {
      int owner = dof_of_sidx(k);
      real xk = shfl(b * rdiag, owner);
      if (isdof) { if (si == k) b = xk; else if (si < k) b -= W.H[k][si] * xk; }
    }
    return isdof ? b : (real)0;
  }

  // Primal Newton with a safeguarded-Newton exact line search (cost is piecewise
  // quadratic).  On exit W.a = qacc, W.fcon = qfrc_constraint.  Returns iterations.
  __device__ static int solve(const AM& A, WM& W, int lane, ConReg& C, real& a_i) {
    const int myblk = lane < ARM_NLINK ? 0 : 1 + (lane - ARM_NLINK) / 6;
    const int d = lane < ARM_NLINK ? lane : (lane - ARM_NLINK) % 6;
    const bool isdof = lane < NV;
    const int si = isdof ? sidx(lane) : 0;
    const int ncon = W.ncon;
    real mdiag = 0;
    if (isdof && lane >= ARM_NLINK) { int k = (lane - ARM_NLINK) / 6; mdiag = d < 3 ? GM().obj_mass[k] : GM().obj_I[k][d - 3]; }
    if (isdof) W.a[lane] = a_i;
    __syncwarp();
    // e = J a - aref
    {
      real u[4];
      contact_apply(W, lane, C, W.a, u);
      C.e0 = u[0] - C.r0;
#pragma unroll
      for (int k = 0; k < 3; k++) C.ek[k] = C.mu[k] * u[1 + k] - C.rk[k];
    }
    real prev_cost = 0;
    int it = 0;
    long long tclk = clock64(); (void)tclk;
    const real fsm_i = isdof ? W.fsm[lane] : (real)0;
    bool done = false;
    int nit = 0;
    // lock-step Newton loop: the group iterates (and leaves) together; barriers inside the loop are legal only then
    const bool vote = W.sync >= 2 && !W.novote;
    for (; it < A.max_iter; it++) {
      real Ma = 0, grad = 0, ccost = 0, cost = 0, gn = 0, dir_i = 0;
      if (!done) {
        ccost = contact_eval(W, lane, C, true);
        __syncwarp();
        // gradient and cost
        if (isdof) {
          if (lane < ARM_NLINK) {
  #pragma unroll
            for (int j = 0; j < ARM_NLINK; j++) Ma += W.Marm[lane][j] * W.a[j];
          } else Ma = mdiag * a_i;
          grad = Ma - fsm_i - jt_force(W, myblk, d);
        }
        cost = warp_sum(ccost + (isdof ? a_i * ((real)0.5 * Ma - fsm_i) : (real)0));
        gn = sqrt(warp_sum(grad * grad));
        if (it > 0 && (prev_cost - cost) * A.scale < A.tol) done = true;
        if (gn * A.scale < A.tol) done = true;
        prev_cost = cost;
      }
      STAGE_CLK(W, 4, tclk);
      // lock-step mode: the block leaves the loop together (shared instruction fetch)
      if (vote) { if (!group_or(!done)) break; }
      else if (done) break;
      STAGE_CLK(W, 10, tclk);
      if (!done) {
        nit++;
        // ---- Hessian rows H = M + sum_c Jc^T W_c Jc (solver ordering)
        // The row's own diagonal block accumulates in registers and is stored once; entries outside the block exist only
        // when a contact couples two blocks (coup / objcoup), and only then is the row cleared and updated in shared memory
        // (block_direction never reads an off-block entry of an uncoupled Hessian).
        if (isdof) {
          const int ownbase = lane < ARM_NLINK ? 6 * NOBJ : (myblk - 1) * 6;
          real hown[6];
#pragma unroll
          for (int j = 0; j < 6; j++) hown[j] = lane < ARM_NLINK ? W.Marm[lane][j] : (j == d ? mdiag : (real)0);
          if ((W.coup | W.objcoup) != 0)
            for (int j = 0; j < NV; j++) W.H[si][j] = 0;
          int n = W.bcount[myblk];
          for (int r = 0; r < n; r++) {
            int e = W.blist[myblk][r], c = e & 0xff, side = e >> 8;
            real jm[4], z[4];
#pragma unroll
            for (int k = 0; k < 4; k++) jm[k] = W.cJ[(side * 4 + k) * 6 + d][c];
            z[0] = W.cW[0][c] * jm[0] + W.cW[1][c] * jm[1] + W.cW[2][c] * jm[2] + W.cW[3][c] * jm[3];
#pragma unroll
            for (int k = 0; k < 3; k++) z[1 + k] = W.cW[1 + k][c] * jm[0] + W.cW[4 + k][c] * jm[1 + k];
#pragma unroll ROLL
            for (int s2 = 0; s2 < 2; s2++) {
              int blk2 = W.cblk[s2][c];
              if (blk2 < 0) continue;
              int cols2 = W.ccols[s2][c];
              if (blk2 == myblk) {
#pragma unroll
                for (int j = 0; j < 6; j++) {
                  if (j < cols2) {
                    real t = 0;
#pragma unroll
                    for (int k = 0; k < 4; k++) t += z[k] * W.cJ[(s2 * 4 + k) * 6 + j][c];
                    hown[j] += t;
                  }
                }
              } else {
                int b2 = blk_base(blk2);
                for (int j = 0; j < cols2; j++) {
                  real t = 0;
#pragma unroll
                  for (int k = 0; k < 4; k++) t += z[k] * W.cJ[(s2 * 4 + k) * 6 + j][c];
                  W.H[si][sidx(b2 + j)] += t;
                }
              }
            }
          }
#pragma unroll
          for (int j = 0; j < 6; j++) W.H[si][ownbase + j] = hown[j];
        }
        __syncwarp();
      }
      STAGE_CLK(W, 5, tclk);
      if (vote) stage_sync(W, 3);
      if (!done) {
        real b = -grad;
        if (W.objcoup == 0) b = block_direction(W, lane, b);
        else b = dense_direction(W, lane, b);
        dir_i = isdof ? b : (real)0;
        if (isdof) W.dir[lane] = dir_i;
        __syncwarp();
      }
      STAGE_CLK(W, 6, tclk);
      if (vote) stage_sync(W, 3);
      if (!done) {
        // ---- exact line search along dir
        real Md = 0;
        if (isdof) {
          if (lane < ARM_NLINK) {
  #pragma unroll
            for (int j = 0; j < ARM_NLINK; j++) Md += W.Marm[lane][j] * W.dir[j];
          } else Md = mdiag * dir_i;
        }
        real q1 = warp_sum(isdof ? dir_i * (Ma - fsm_i) : (real)0);
        real q2 = warp_sum(isdof ? dir_i * Md : (real)0);
        {
          real u[4];
          contact_apply(W, lane, C, W.dir, u);
          C.de0 = u[0];
  #pragma unroll
          for (int k = 0; k < 3; k++) C.dek[k] = C.mu[k] * u[1 + k];
        }
        real alpha = 0, lo = 0, hi = (real)-1, s_lo = 0;
        real s0abs = 0;
        const int LSMAX = sizeof(real) == 8 ? 40 : 12;
        for (int ls = 0; ls < LSMAX; ls++) {
          // derivative and curvature of the cost at alpha
          real sl = 0, cu = 0;
          if (C.dim == 1) {
            real jar = C.e0 + alpha * C.de0;
            if (jar < 0 || (jar == 0 && C.de0 < 0)) { sl = C.D * jar * C.de0; cu = C.D * C.de0 * C.de0; }
          } else if (C.dim > 1) {
  #pragma unroll
            for (int k = 0; k < 3; k++) {
              if (k < C.dim - 1) {
                real jp = C.e0 + C.ek[k] + alpha * (C.de0 + C.dek[k]), dp = C.de0 + C.dek[k];
                real jm = C.e0 - C.ek[k] + alpha * (C.de0 - C.dek[k]), dm = C.de0 - C.dek[k];
                if (jp < 0 || (jp == 0 && dp < 0)) { sl += C.D * jp * dp; cu += C.D * dp * dp; }
                if (jm < 0 || (jm == 0 && dm < 0)) { sl += C.D * jm * dm; cu += C.D * dm * dm; }
              }
            }
          }
          sl = warp_sum(sl) + q1 + alpha * q2;
          cu = warp_sum(cu) + q2;
          if (ls == 0) s0abs = fabs(sl);
          const real lstol = sizeof(real) == 8 ? (real)1e-14 : (real)1e-6;
          if (fabs(sl) <= lstol * s0abs || cu <= 0) break;
          if (sl < 0) { lo = alpha; s_lo = sl; } else hi = alpha;
          real an = alpha - sl / cu;
          if (hi >= 0 && !(an > lo && an < hi)) an = (real)0.5 * (lo + hi);
          if (an == alpha) break;
          alpha = an;
        }
        (void)s_lo;
        a_i += alpha * dir_i;
        if (isdof) W.a[lane] = a_i;
        C.e0 += alpha * C.de0;
  #pragma unroll
        for (int k = 0; k < 3; k++) C.ek[k] += alpha * C.dek[k];
        __syncwarp();
      }
      STAGE_CLK(W, 7, tclk);
    }
    // final forces -> qfrc_constraint.  A converged warp evaluated them at its last
    // iterate already (cf is current); only an exit at max_iter leaves a newer iterate.
    if (!done) contact_eval(W, lane, C, true);
    __syncwarp();
    if (isdof) W.fcon[lane] = jt_force(W, myblk, d);
    __syncwarp();
    STAGE_CLK(W, 8, tclk);
    return nit;
  }

  // ----------------------------------------------------- [P10] touch sensor (mj_sensorAcc)
  __device__ static real touch_force(const AM& A, const WM& W, int lane, const ConReg& C) {
    real f = 0;
    if (A.touch_slot >= 0 && C.dim > 0 && W.cpair[lane] >= 0 && ((A.touch_pair_mask >> W.cpair[lane]) & 1)) {
      real fn = W.cf[0][lane];
      if (fn > 0) {
        const real* R = W.xmat[A.touch_slot];
        real sp[3], sm[9], v[3];
        real tp[3] = {A.touch_pos[0], A.touch_pos[1], A.touch_pos[2]};
        mat_vec(v, R, tp);
        for (int k = 0; k < 3; k++) sp[k] = W.xpos[A.touch_slot][k] + v[k];
        for (int r = 0; r < 3; r++)
          for (int c = 0; c < 3; c++) sm[3 * r + c] = R[3 * r] * A.touch_mat[c] + R[3 * r + 1] * A.touch_mat[3 + c] + R[3 * r + 2] * A.touch_mat[6 + c];
        real dif[3] = {W.craw[1][lane] - sp[0], W.craw[2][lane] - sp[1], W.craw[3][lane] - sp[2]}, loc[3];
        matT_vec(loc, sm, dif);
        if (loc[0] * loc[0] + loc[1] * loc[1] <= A.touch_size[0] * A.touch_size[0] && fabs(loc[2]) <= A.touch_size[1]) f = fn;
      }
    }
    return warp_sum(f);
  }

  // ------------------------------------------------- [P11] implicitfast / Euler + advance
  // (M - h dF/dv) qacc' = qfrc_smooth + qfrc_constraint; dF/dv = -kv of unclamped
  // actuators (mjd_actuator_vel).  Props have no velocity-dependent forces.
  __device__ static void integrate(const AM& A, WM& W, int lane, bool clamped, double& time) {
    const real h = A.timestep;
    real qacc = 0;
    if (lane < ARM_NLINK) {
      real dv = (A.integrator == MJS_INT_IMPLICITFAST && !clamped) ? (real)GM().kv[lane] : (real)0;
      // 6x6 system in the (now free) Hessian scratch, lanes 0..5 = rows
#pragma unroll
      for (int j = 0; j < ARM_NLINK; j++) W.H[lane][j] = W.Marm[lane][j] + (j == lane ? h * dv : (real)0);
    }
    __syncwarp();
    {
      real rhs[6];
      const real bi = lane < NV ? W.fsm[lane] + W.fcon[lane] : (real)0;
#pragma unroll
      for (int j = 0; j < 6; j++) rhs[j] = shfl(bi, j);
      if (lane < ARM_NLINK) {
        real h6[21], rd[6];
#pragma unroll
        for (int i = 0; i < 6; i++)
#pragma unroll
          for (int j = 0; j <= i; j++) h6[i * (i + 1) / 2 + j] = W.H[i][j];
        chol6(h6, rd);
        solve6(h6, rd, rhs);
#pragma unroll
        for (int j = 0; j < 6; j++) if (j == lane) qacc = rhs[j];
      } else if (lane < NV) {
        int k = (lane - ARM_NLINK) / 6, d = (lane - ARM_NLINK) % 6;
        qacc = bi / (d < 3 ? GM().obj_mass[k] : GM().obj_I[k][d - 3]);
      }
    }
    // mj_advance in double
    if (lane < NV) W.qvel[lane] += (double)h * (double)qacc;
    __syncwarp();
    if (lane < ARM_NLINK) W.qpos[lane] += (double)h * W.qvel[lane];
    else if (lane < NV) {
      int k = (lane - ARM_NLINK) / 6, d = (lane - ARM_NLINK) % 6;
      int qa = ARM_NLINK + 7 * k, da = ARM_NLINK + 6 * k;
      if (d < 3) W.qpos[qa + d] += (double)h * W.qvel[da + d];
      else if (d == 3) {
        // quaternion exponential map: q <- q * exp(h * omega_body / 2)
        double w0 = W.qvel[da + 3], w1 = W.qvel[da + 4], w2 = W.qvel[da + 5];
        double nrm = sqrt(w0 * w0 + w1 * w1 + w2 * w2);
        double ax0 = 1, ax1 = 0, ax2 = 0;
        if (nrm >= MJS_MINVAL) { ax0 = w0 / nrm; ax1 = w1 / nrm; ax2 = w2 / nrm; }
        double ang = nrm * (double)h, sn, cs;
        if (sizeof(real) == 8) sincos_r<double>(0.5 * ang, &sn, &cs);
        else { float fs, fc; sincosf((float)(0.5 * ang), &fs, &fc); sn = fs; cs = fc; }
        double dq[4] = {cs, ax0 * sn, ax1 * sn, ax2 * sn};
        double q[4] = {W.qpos[qa + 3], W.qpos[qa + 4], W.qpos[qa + 5], W.qpos[qa + 6]}, r[4];
        quat_mul<double>(r, q, dq);
        double n2 = 1.0 / sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3]);
        W.qpos[qa + 3] = r[0] * n2; W.qpos[qa + 4] = r[1] * n2; W.qpos[qa + 5] = r[2] * n2; W.qpos[qa + 6] = r[3] * n2;
      }
    }
    time += (double)h;
    __syncwarp();
  }

  // Pairs that collide() handed to the general convex routine (lane = pair): MPR penetration, the contact appended to the
  // warp's contact list, the separating axis of a miss returned for the next substep.  Runs between collide and make_rows.
  __device__ static __noinline__ SepAxis<real> convex_fixup(WM* Wp, int lane, bool cvx, SepAxis<real> sep) {
    WM& W = *Wp;
    if (cvx) {
      const int g1 = GM().pair_g1[lane], g2 = GM().pair_g2[lane];
#if defined(MJS_CONVEX_MODE) && MJS_CONVEX_MODE == 0   // A/B build: the pre-MPR behaviour (status bit instead of a contact)
      atomicOr(&W.status, MJS_STATUS_UNSUPPORTED_PAIR);
#else
#ifdef MJS_CONVEX_COUNT
      atomicAdd(&W.ncvx, 1);
#endif
      const ConvexHit<real> ch = convex_pair_dev<real>(g1, g2, W.gpos[g1], W.gmat[g1], W.gpos[g2], W.gmat[g2]);
      MJS_SEP_SET(sep, ch.hit, ch.n);
      if (ch.hit && ch.dist < GM().pair_incl[lane]) {  // mj_collision: contacts with dist >= margin - gap produce no rows
#ifdef MJS_CONVEX_COUNT
        atomicAdd(&W.ncvx_hit, 1);
#endif
        const int c = atomicAdd(&W.ncon, 1);
        if (c < ARM_MAXC) {
          W.craw[0][c] = ch.dist;
          W.craw[1][c] = ch.p[0]; W.craw[2][c] = ch.p[1]; W.craw[3][c] = ch.p[2];
          W.craw[4][c] = ch.n[0]; W.craw[5][c] = ch.n[1]; W.craw[6][c] = ch.n[2];
          W.craw[7][c] = 0; W.craw[8][c] = 0; W.craw[9][c] = 0;
          W.cpair[c] = lane;
        } else atomicOr(&W.status, MJS_STATUS_CONTACT_OVERFLOW);
      }
#endif
    }
    __syncwarp();
    if (lane == 0 && W.ncon > ARM_MAXC) W.ncon = ARM_MAXC;
    __syncwarp();
    return sep;
  }

  // ------------------------------------------------ one physics substep (legacy_step)
  struct SubRes { real a_i; double time; real touch; int niter; SepAxis<real> sep; };

  // One physics step for the current (qpos, qvel, ctrl): mj_step1 (kinematics ..
  // constraint rows) followed by mj_step2 (solve, touch sensor, integrate).  The
  // composer loop's "step2 then step1" order is the same sequence seen from the
  // state's point of view (ArmTask::control_step).  Deliberately NOT inlined:
  // one copy of the heavy code keeps the instruction footprint inside the i-cache.
  // The substep is two non-inlined halves -- mj_step1 up to the narrowphase, then rows / solve / integrate -- with the rare
  // general-convex post-pass called between them from this small inlined wrapper: a call from inside either half costs the
  // half its registers (profiles/r2/convex_pairs.txt).
  __device__ static __noinline__ int substep_front(WM* Wp, int lane, int do_integrate, SepAxis<real> sep) {
    const AM& A = cmodel<real>();
    WM& W = *Wp;
    if (!(do_integrate & 2)) stage_sync(W, 1);
    long long tclk = clock64(); (void)tclk;
#ifdef MJS_STAGE_CLOCKS_REPEAT  // experiment: the same (non-inlined) stage twice; the second run is hot in the instruction cache
    kinematics_only(&W, lane);
    STAGE_CLK(W, 0, tclk);
    kinematics_only(&W, lane);
    STAGE_CLK(W, 11, tclk);
#else
    kinematics(A, W, lane);
    STAGE_CLK(W, 0, tclk);
#endif
    stage_sync(W, 3);
#ifdef MJS_STAGE_CLOCKS_REPEAT
    bool clamped = dynamics_only(&W, lane);
    STAGE_CLK(W, 1, tclk);
    clamped = dynamics_only(&W, lane);
    STAGE_CLK(W, 12, tclk);
    collide_only(&W, lane);
    STAGE_CLK(W, 2, tclk);
    collide_only(&W, lane);
    STAGE_CLK(W, 13, tclk);
    bool cvx = false;
#else
    bool clamped = dynamics(A, W, lane);
    STAGE_CLK(W, 1, tclk);
    stage_sync(W, 3);
    bool cvx;
    collide(A, W, lane, sep, cvx);
    STAGE_CLK(W, 2, tclk);
#endif
    return (clamped ? 1 : 0) | (cvx ? 2 : 0);
  }
  __device__ static __noinline__ SubRes substep_back(WM* Wp, int lane, real a_i, double time, int do_integrate, bool clamped) {
    const AM& A = cmodel<real>();
    WM& W = *Wp;
    ConReg C;
    long long tclk = clock64(); (void)tclk;
    stage_sync(W, 3);
    make_rows(A, W, lane, C);
    STAGE_CLK(W, 3, tclk);
    stage_sync(W, 2);
    SubRes r;
    r.niter = solve(A, W, lane, C, a_i);
    tclk = clock64();
    r.touch = touch_force(A, W, lane, C);
    stage_sync(W, 3);
    if (do_integrate & 1) integrate(A, W, lane, clamped, time);
    STAGE_CLK(W, 9, tclk);
    r.a_i = a_i;
    r.time = time;
    return r;
  }
  __device__ static __forceinline__ SubRes substep(WM* Wp, int lane, real a_i, double time, int do_integrate, SepAxis<real> sep = SepAxis<real>MJS_SEP_ZERO) {
    const int f = substep_front(Wp, lane, do_integrate, sep);
    if (__any_sync(FULLMASK, (f & 2) != 0)) sep = convex_fixup(Wp, lane, (f & 2) != 0, sep);  // rare: pairs without an analytic routine (cap contacts)
    SubRes r = substep_back(Wp, lane, a_i, time, do_integrate, (f & 1) != 0);
    r.sep = sep;
    return r;
  }
  __device__ static __noinline__ void kinematics_only(WM* Wp, int lane) { kinematics(cmodel<real>(), *Wp, lane); }
  __device__ static __noinline__ bool dynamics_only(WM* Wp, int lane) { return dynamics(cmodel<real>(), *Wp, lane); }
  __device__ static __noinline__ void collide_only(WM* Wp, int lane) { SepAxis<real> sep = MJS_SEP_ZERO; bool cvx; collide(cmodel<real>(), *Wp, lane, sep, cvx); }
};

// ===================================================================== analytic UR IK
// Device restatement of oracle/mjs_oracle.cpp ur_ik_all / ur_ik_closest (which
// restates ur_analytic_ik.ur5e.inverse_kinematics_closest, robot.py:31-38).
// Lane s (0..7) evaluates branch (s1, s5, s3) = bits (4, 2, 1); double precision.
// DH transform with the link twist given by its sine / cosine (the twists of the UR chain are 0 and +-pi/2: constants at every call
// site, where the library sincos(alpha) cost ~160 instructions each time); the joint angle uses the kernel's short sincos.
__device__ __forceinline__ void dh_T(double* T, double th, double dd, double a, double sa, double ca) {
  double st, ct;
  sincos_r<double>(th, &st, &ct);
  T[0] = ct; T[1] = -st * ca; T[2] = st * sa; T[3] = a * ct;
  T[4] = st; T[5] = ct * ca; T[6] = -ct * sa; T[7] = a * st;
  T[8] = 0; T[9] = sa; T[10] = ca; T[11] = dd;
}
// (loops fully unrolled: the 3x4 arrays must be indexed statically to stay in registers -- the once-per-step code was the
// kernel's only local-memory traffic, 3.2 kB per env-step of DRAM writes in round 1)
__device__ __forceinline__ void mul34(double* C, const double* A, const double* B) {  // rigid 3x4 products
  double R[12];
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) R[4 * i + j] = A[4 * i] * B[j] + A[4 * i + 1] * B[4 + j] + A[4 * i + 2] * B[8 + j];
    R[4 * i + 3] = A[4 * i] * B[3] + A[4 * i + 1] * B[7] + A[4 * i + 2] * B[11] + A[4 * i + 3];
  }
#pragma unroll
  for (int i = 0; i < 12; i++) C[i] = R[i];
}
__device__ __forceinline__ void inv34(double* B, const double* A) {
  double R[12] = {A[0], A[4], A[8], 0, A[1], A[5], A[9], 0, A[2], A[6], A[10], 0};
#pragma unroll
  for (int i = 0; i < 3; i++) R[4 * i + 3] = -(R[4 * i] * A[3] + R[4 * i + 1] * A[7] + R[4 * i + 2] * A[11]);
#pragma unroll
  for (int i = 0; i < 12; i++) B[i] = R[i];
}
#define MJS_PI 3.14159265358979323846

// all 32 lanes call; returns true (warp-uniform) and q[6] (identical in all lanes) on success
__device__ __forceinline__ bool ur_ik_closest_warp(const double* dh, const double* T06, const double* seed, double* qout, int lane) {
  double d1 = dh[0], a2 = dh[1], a3 = dh[2], d4 = dh[3], d5 = dh[4], d6 = dh[5];
  double q[6] = {0, 0, 0, 0, 0, 0};
  double dist = 1e300;
  bool valid = lane < 8;
  int s1 = (lane >> 2) & 1, s5 = (lane >> 1) & 1, s3 = lane & 1;
  double p05[3] = {T06[3] - d6 * T06[2], T06[7] - d6 * T06[6], T06[11] - d6 * T06[10]};
  double rxy = hypot(p05[0], p05[1]);
  if (rxy < fabs(d4)) valid = false;
  if (valid) {
    double psi = atan2(p05[1], p05[0]), phi = acos(d4 / rxy);
    double th1 = psi + (s1 == 0 ? phi : -phi) + MJS_PI / 2;
    double T01[12], T10[12], T16[12];
    dh_T(T01, th1, d1, 0, 1.0, 0.0);
    inv34(T10, T01);
    mul34(T16, T10, T06);
    double c5 = T16[10];
    if (fabs(c5) > 1) { if (fabs(c5) < 1 + 1e-9) c5 = c5 > 0 ? 1 : -1; else valid = false; }
    if (valid) {
      double th5 = (s5 == 0 ? 1 : -1) * acos(c5);
      double sn5 = sin(th5), th6, th234;
      if (fabs(sn5) < 1e-9) {
        th6 = 0;
        th234 = c5 < 0 ? atan2(-T16[4], -T16[0]) : atan2(T16[4], T16[0]);
      } else {
        th6 = atan2(-T16[9] / sn5, T16[8] / sn5);
        th234 = atan2(T16[6] / -sn5, T16[2] / -sn5);
      }
      double T45[12], T56[12], T46[12], T64[12], T14[12];
      dh_T(T45, th5, d5, 0, -1.0, 0.0);
      dh_T(T56, th6, d6, 0, 0.0, 1.0);
      mul34(T46, T45, T56);
      inv34(T64, T46);
      mul34(T14, T16, T64);
      double x = T14[3], y = T14[7];
      double c3 = (x * x + y * y - a2 * a2 - a3 * a3) / (2 * a2 * a3);
      if (fabs(c3) > 1) { if (fabs(c3) < 1 + 1e-9) c3 = c3 > 0 ? 1 : -1; else valid = false; }
      if (valid) {
        double th3 = (s3 == 0 ? 1 : -1) * acos(c3);
        double th2 = atan2(y, x) - atan2(a3 * sin(th3), a2 + a3 * cos(th3));
        double th4 = th234 - th2 - th3;
        q[0] = th1; q[1] = th2; q[2] = th3; q[3] = th4; q[4] = th5; q[5] = th6;
        dist = 0;
#pragma unroll
        for (int i = 0; i < 6; i++) {
          double v = fmod(q[i] + MJS_PI, 2 * MJS_PI);
          if (v < 0) v += 2 * MJS_PI;
          v -= MJS_PI;
          double bq = v, bd = fabs(v - seed[i]);
#pragma unroll
          for (int w = -1; w <= 1; w += 2) {
            double c = v + w * 2 * MJS_PI;
            if (c >= -2 * MJS_PI && c <= 2 * MJS_PI && fabs(c - seed[i]) < bd) { bd = fabs(c - seed[i]); bq = c; }
          }
          q[i] = bq;
          dist += bd * bd;
        }
      }
    }
  }
  if (!valid) dist = 1e300;
  // argmin with lowest-lane tie break (the oracle keeps the first strict minimum)
  double best = dist;
  int bl = lane;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double od = __shfl_xor_sync(FULLMASK, best, o);
    int ol = __shfl_xor_sync(FULLMASK, bl, o);
    if (od < best || (od == best && ol < bl)) { best = od; bl = ol; }
  }
#pragma unroll
  for (int i = 0; i < 6; i++) qout[i] = __shfl_sync(FULLMASK, q[i], bl);
  return best < 1e299;
}

// ===================================================================== state layout
// Per-env AoS record in HBM: SD doubles + SI ints.  A warp streams its record
// with lane-strided (coalesced) loads once per control step.
template <int NOBJ>
struct ArmLayout {
  static constexpr int NV = ARM_NLINK + 6 * NOBJ, NQ = ARM_NLINK + 7 * NOBJ;
  static constexpr int QPOS = 0, QVEL = QPOS + NQ, WARM = QVEL + NV, CTRL = WARM + NV, TIME = CTRL + 6, TARGET = TIME + 1;
  static constexpr int Q0 = TARGET + 3, Q1 = Q0 + 6, T0 = Q1 + 6, OVR = T0 + 1, RET = OVR + 3, SD = RET + 1;
  static constexpr int I_STEP = 0, I_TRAJ = 1, I_SWA = 2, I_SWP = 3, I_SWC = 4, I_STATUS = 5, I_EPISODE = 6, I_LEN = 7, SI = 8;
};

#define MJS_TIMES_WORDS (4 + MJS_NCLK)  // 16; per-env words of the mjs_debug_times buffer (4 used; 4..15 = stage cycles of a -DMJS_STAGE_CLOCKS build)
struct ArmIO {
  double* sd;
  int* si;
  const float* actions;
  float* obs;
  float* reward;
  unsigned char* flags;
  float* final_obs;
  double *obs64, *reward64, *final_obs64;  // optional float64 twins of obs / reward / final_obs (mjs_set_output64; the reference emits float64, dmc2gym.py:66-75)
  double* stats;
  const unsigned char* mask;
  int N, autoreset;
  unsigned long long seed;
  unsigned int env_offset, step_index;
  int sync_level;
  // pre-settled next-episode states (tasks with settle steps): see arm_reset_kernel
  double* slot_sd;
  int* slot_si;
  double* slot_obs;
  int* slot_ready;
  // settle jobs folded into the step kernel's grid (arm_jobs_kernel + the job blocks of arm_step_kernel)
  int* slot_prog;   // settle substeps already done for the slot under preparation (0: not started)
  int* jobs;        // env ids whose slot is empty or in preparation, rebuilt before every step launch
  int* njobs;       // number of entries wanted this launch (may exceed job_cap: the rest waits a step)
  int* njobs_next;  // the other counter of the ping-pong pair, cleared for the next launch
  int job_cap, job_blocks;
  int* queue;       // non-null: persistent step grid, warps pull work items from this counter (zeroed before the launch)
  float* gpose;      // mjs_render: world pose of every model geom [N][MJS_MAXGEOM][12] (pos 3, mat 9)
  float* aux;        // mjs_render: per-env anchors [N][8]: target xyz, switch xyz, switch active
  long long* times;  // test hook (mjs_debug_times): per env {start, end of control step, end, nit_sum<<32 | nit_max<<16 | smid} in ns
};

template <typename real, int NOBJ>
struct ArmTask {
  using E = ArmEngine<real, NOBJ>;
  using WM = WarpMem<real, NOBJ>;
  using AM = ArmModel<real>;
  using L = ArmLayout<NOBJ>;
  static constexpr int NV = L::NV, NQ = L::NQ;

  // task-level registers (identical in all lanes unless noted)
  struct TS {
    double time, target[3], t0, ovr[3], ret, tq0, tq1;  // tq0/tq1: lane j < 6 holds waypoint j
    int step, traj, swa, swp, swc, status, episode, len;
  };

  __device__ static double uniform(const ArmIO& io, unsigned genv, unsigned episode, unsigned purpose, unsigned idx, double lo, double hi) {
    Philox4 r = philox4x32_10(genv, episode, purpose, idx >> 1, (uint32_t)io.seed, (uint32_t)(io.seed >> 32));
    double u = (idx & 1) ? u53(r.v[2], r.v[3]) : u53(r.v[0], r.v[1]);
    return lo + (hi - lo) * u;
  }

  __device__ static void load(const AM& A, WM& W, const ArmIO& io, int e, int lane, TS& T, real& a_i, int sync = 0) {
    const double* s = io.sd + (size_t)e * L::SD;
    for (int k = lane; k < NQ; k += 32) W.qpos[k] = s[L::QPOS + k];
    if (lane < NV) { W.qvel[lane] = s[L::QVEL + lane]; a_i = (real)s[L::WARM + lane]; } else a_i = 0;
    if (lane < ARM_NLINK) { W.ctrl[lane] = s[L::CTRL + lane]; T.tq0 = s[L::Q0 + lane]; T.tq1 = s[L::Q1 + lane]; } else { T.tq0 = 0; T.tq1 = 0; }
    T.time = s[L::TIME]; T.t0 = s[L::T0]; T.ret = s[L::RET];
    for (int k = 0; k < 3; k++) { T.target[k] = s[L::TARGET + k]; T.ovr[k] = s[L::OVR + k]; }
    const int* si = io.si + (size_t)e * L::SI;
    T.step = si[L::I_STEP]; T.traj = si[L::I_TRAJ]; T.swa = si[L::I_SWA]; T.swp = si[L::I_SWP]; T.swc = si[L::I_SWC];
    T.status = si[L::I_STATUS]; T.episode = si[L::I_EPISODE]; T.len = si[L::I_LEN];
    if (lane == 0) { W.status = 0; W.sync = sync & 7; W.novote = (sync >> 3) & 1; W.period = 1 + ((sync >> 4) & 15); }
#ifdef MJS_CONVEX_COUNT
    if (lane == 0) { W.ncvx = 0; W.ncvx_hit = 0; }
#endif
#ifdef MJS_STAGE_CLOCKS
    if (lane < MJS_NCLK) W.clk[lane] = 0;
#endif
    init_static(A, W, lane, T);
    __syncwarp();
  }

  __device__ static void init_static(const AM& A, WM& W, int lane, const TS& T) {
    if (lane < 9) { W.xmat[SLOT_STATIC][lane] = (lane % 4 == 0) ? (real)1 : (real)0; W.xmat[SLOT_OVR][lane] = A.has_ovr ? A.ovr_mat[lane] : ((lane % 4 == 0) ? (real)1 : (real)0); }
    if (lane < 3) { W.xpos[SLOT_STATIC][lane] = 0; W.xpos[SLOT_OVR][lane] = (real)T.ovr[lane]; }
  }

  __device__ static void store(WM& W, const ArmIO& io, int e, int lane, const TS& T, real a_i) {
    double* s = io.sd + (size_t)e * L::SD;
    __syncwarp();
    for (int k = lane; k < NQ; k += 32) s[L::QPOS + k] = W.qpos[k];
    if (lane < NV) { s[L::QVEL + lane] = W.qvel[lane]; s[L::WARM + lane] = (double)a_i; }
    if (lane < ARM_NLINK) { s[L::CTRL + lane] = W.ctrl[lane]; s[L::Q0 + lane] = T.tq0; s[L::Q1 + lane] = T.tq1; }
    if (lane == 0) {
      s[L::TIME] = T.time; s[L::T0] = T.t0; s[L::RET] = T.ret;
      for (int k = 0; k < 3; k++) { s[L::TARGET + k] = T.target[k]; s[L::OVR + k] = T.ovr[k]; }
      int* si = io.si + (size_t)e * L::SI;
      si[L::I_STEP] = T.step; si[L::I_TRAJ] = T.traj; si[L::I_SWA] = T.swa; si[L::I_SWP] = T.swp; si[L::I_SWC] = T.swc;
      si[L::I_STATUS] = T.status | W.status; si[L::I_EPISODE] = T.episode; si[L::I_LEN] = T.len;
    }
  }

  // robot.py:218-259 servoL / servoJ.  tcp == nullptr: joint target in qt (lane j holds qt_j in all lanes' array)
  __device__ static __forceinline__ void servoJ(WM& W, int lane, TS& T, const double* qt) {
    if (lane < ARM_NLINK) {
      T.tq0 = W.qpos[lane];
      double v = 0;
#pragma unroll
      for (int k = 0; k < ARM_NLINK; k++) if (k == lane) v = qt[k];  // static indexing keeps the caller's array in registers
      T.tq1 = v;
    }
    T.t0 = T.time;
    T.traj = 1;
  }
  __device__ static __forceinline__ bool tcp_to_joints(const AM& A, const double* tcp, const double* seed, double* q, int lane) {
    // top-down flange pose (TOP_DOWN_QUATERNION = 180 deg about x), robot.py:138-151
    double T06[12] = {1, 0, 0, tcp[0] - (double)A.tcp_offset[0], 0, -1, 0, tcp[1] + (double)A.tcp_offset[1], 0, 0, -1, tcp[2] + (double)A.tcp_offset[2]};
    return ur_ik_closest_warp(A.ik_dh, T06, seed, q, lane);
  }
  __device__ static __forceinline__ void servoL(const AM& A, WM& W, int lane, TS& T, const double* tcp) {
    double seed[6], q[6];
#pragma unroll
    for (int i = 0; i < 6; i++) seed[i] = W.qpos[i];
    if (!tcp_to_joints(A, tcp, seed, q, lane)) { T.status |= MJS_STATUS_IK_FAIL; return; }
    servoJ(W, lane, T, q);
  }
  // joint_trajectory.py:33-47 + robot.py:261-272
  __device__ static void before_substep(const AM& A, WM& W, int lane, const TS& T) {
    if (T.traj && lane < ARM_NLINK) {
      double t0 = T.t0, t1 = T.t0 + A.control_dt;
      double tt = fmin(fmax(T.time, t0), t1);
      W.ctrl[lane] = T.tq0 + (T.tq1 - T.tq0) * (tt - t0) / (t1 - t0);
    }
    __syncwarp();
  }
  __device__ static void switch_update(const AM& A, TS& T, real force) {
    int was = T.swp;  // switch.py:51-60
    T.swp = ((double)force >= A.touch_lo && (double)force <= A.touch_hi) ? 1 : 0;
    if (T.swp && !was) T.swa = !T.swa;
    T.swc += T.swp;
  }

  __device__ static void tcp_position(const AM& A, const WM& W, double* tcp) {
    const real* R = W.xmat[A.tcp_slot];
    real v[3], sm[9], o[3];
    real tp[3] = {A.tcp_pos[0], A.tcp_pos[1], A.tcp_pos[2]}, off[3] = {A.tcp_offset[0], A.tcp_offset[1], A.tcp_offset[2]};
    mat_vec(v, R, tp);
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) sm[3 * r + c] = R[3 * r] * A.tcp_mat[c] + R[3 * r + 1] * A.tcp_mat[3 + c] + R[3 * r + 2] * A.tcp_mat[6 + c];
    mat_vec(o, sm, off);
    for (int k = 0; k < 3; k++) tcp[k] = (double)W.xpos[A.tcp_slot][k] + (double)v[k] + (double)o[k];
  }

  struct Out { double obs[16]; double reward; int terminated, truncated, success, done, nit_sum, nit_max, h4, h8, h16, h32; };

  // observation + reward + termination of the CURRENT state (kinematics must be fresh)
  __device__ static void observe(const AM& A, const WM& W, const TS& T, Out& o) {
    double tcp[3];
    tcp_position(A, W, tcp);
    bool accomplished = false, has_success = false;
    double reward = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) o.obs[k] = 0;
    if (A.kind == MJS_TASK_REACH) {
#pragma unroll
      for (int k = 0; k < 3; k++) { o.obs[k] = tcp[k]; o.obs[3 + k] = T.target[k]; }
      double dist = sqrt((tcp[0] - T.target[0]) * (tcp[0] - T.target[0]) + (tcp[1] - T.target[1]) * (tcp[1] - T.target[1]) + (tcp[2] - T.target[2]) * (tcp[2] - T.target[2]));
      reward = A.reward_type == 0 ? (dist < A.goal_threshold ? 1.0 : 0.0) : -dist;
    } else if (A.kind == MJS_TASK_PUSH) {
#pragma unroll
      for (int k = 0; k < 3; k++) o.obs[k] = tcp[k];
      o.obs[3] = T.target[0]; o.obs[4] = T.target[1];
      double sum = 0, nearest = 1e300;
      int within = 0;
#pragma unroll
      for (int b = 0; b < NOBJ; b++) {
        double x = W.qpos[ARM_NLINK + 7 * b], y = W.qpos[ARM_NLINK + 7 * b + 1];
        o.obs[5 + 2 * b] = x; o.obs[6 + 2 * b] = y;
        double dd = hypot(x - T.target[0], y - T.target[1]);
        sum += dd;
        if (dd < A.target_radius) within++;
        nearest = fmin(nearest, hypot(tcp[0] - x, tcp[1] - y));
      }
      if (A.reward_type == 0) reward = within;
      else reward = (-sum / NOBJ - A.nearest_coef * nearest) * 0.1;
      accomplished = within == NOBJ;
    } else {  // BUTTON
      // robot_push_button.py:109-113: joint_configuration with ABS_JOINT_ACTION (7 actions), tcp_position with ABS_EEF_ACTION (4)
      const bool joint_action = A.action_dim == 7;
      const real* R = W.xmat[A.touch_slot];
      real tp[3] = {A.touch_pos[0], A.touch_pos[1], A.touch_pos[2]}, v[3];
      mat_vec(v, R, tp);
      double quirk = 0.5 * ((double)A.touch_size[0] / 1.01);  // switch.py:86-87
      double swp[3];
#pragma unroll
      for (int k = 0; k < 3; k++) swp[k] = (double)W.xpos[A.touch_slot][k] + (double)v[k] + quirk;
#pragma unroll
      for (int k = 0; k < 6; k++) o.obs[k] = joint_action ? W.qpos[k] : (k < 3 ? tcp[k] : swp[k - 3]);
#pragma unroll
      for (int k = 0; k < 3; k++) o.obs[6 + k] = joint_action ? swp[k] : (k == 0 ? (double)T.swa : 0.0);
      o.obs[9] = joint_action ? (double)T.swa : 0.0;
      double dg = sqrt((tcp[0] - A.button_goal[0]) * (tcp[0] - A.button_goal[0]) + (tcp[1] - A.button_goal[1]) * (tcp[1] - A.button_goal[1]) + (tcp[2] - A.button_goal[2]) * (tcp[2] - A.button_goal[2]));
      bool goal = T.swa && dg < A.button_goal_radius;
      reward = goal ? 1.0 : 0.0;
      accomplished = goal; has_success = true;
    }
    bool terminate = accomplished;
    if (A.kind == MJS_TASK_PUSH && A.max_steps > 0 && T.step >= A.max_steps) terminate = true;
    bool timeup = A.time_limit > 0 && T.time >= A.time_limit;
    double discount = (accomplished && A.kind != MJS_TASK_REACH) ? 0.0 : 1.0;
    bool last = terminate || timeup;
    o.reward = reward;
    o.done = last; o.truncated = last && discount > 0; o.terminated = last && discount == 0;
    o.success = has_success ? (accomplished ? 1 : 0) : 0;
  }

  struct SubLoop { real a_i; double time, t0, tq0, tq1; int traj, swa, swp, swc, nit_sum, nit_max; };
  __device__ static __noinline__ SubLoop run_substeps(WM* Wp, int lane, SubLoop s) {
    const AM& A = cmodel<real>();
    WM& W = *Wp;
    SepAxis<real> sep = MJS_SEP_ZERO;  // per pair lane: separating axis cached across the substeps of this control step
    for (int k = 0; k < A.n_sub; k++) {
      if (s.traj && lane < ARM_NLINK) {  // before_substep: joint_trajectory.py:33-47 + robot.py:261-272
        double t0 = s.t0, t1 = s.t0 + A.control_dt;
        double tt = fmin(fmax(s.time, t0), t1);
        W.ctrl[lane] = s.tq0 + (s.tq1 - s.tq0) * (tt - t0) / (t1 - t0);
      }
      __syncwarp();
      typename E::SubRes so = E::substep(&W, lane, s.a_i, s.time, (k % W.period) ? 3 : 1, sep);  // bit 1: skip the substep-start barrier
      s.a_i = so.a_i; s.time = so.time; sep = so.sep;
      s.nit_sum += so.niter; s.nit_max = so.niter > s.nit_max ? so.niter : s.nit_max;
      if (A.kind == MJS_TASK_BUTTON) {  // switch.py:51-60, evaluated after every substep
        int was = s.swp;
        s.swp = ((double)so.touch >= A.touch_lo && (double)so.touch <= A.touch_hi) ? 1 : 0;
        if (s.swp && !was) s.swa = !s.swa;
        s.swc += s.swp;
      }
    }
    return s;
  }

  // One control step for the env owned by this warp (composer loop)
  __device__ static void control_step(const AM& A, WM& W, int lane, TS& T, real& a_i, const double* action, Out& o, const ArmIO& io, int e) {
    if (A.kind == MJS_TASK_REACH) servoL(A, W, lane, T, action);
    else if (A.kind == MJS_TASK_PUSH) {
      T.step += 1;
      double tcp[3] = {action[0], action[1], A.tcp_z};
      servoL(A, W, lane, T, tcp);
    } else {
      if (A.action_dim == 7) servoJ(W, lane, T, action);
      else servoL(A, W, lane, T, action);
    }
    o.h4 = o.h8 = o.h16 = o.h32 = 0;
    // the n_sub substeps run inside ONE non-inlined call that carries only what the loop needs: the two calls per substep
    // (substep_front / substep_back) then save and restore a dozen registers instead of this function's whole task state
    // (264 local-memory instructions per substep before, on the dependent path)
    SubLoop sl;
    sl.a_i = a_i; sl.time = T.time; sl.t0 = T.t0; sl.tq0 = T.tq0; sl.tq1 = T.tq1; sl.traj = T.traj;
    sl.swa = T.swa; sl.swp = T.swp; sl.swc = T.swc; sl.nit_sum = 0; sl.nit_max = 0;
    sl = run_substeps(&W, lane, sl);
    a_i = sl.a_i; T.time = sl.time; T.swa = sl.swa; T.swp = sl.swp; T.swc = sl.swc;
    const int nit_sum = sl.nit_sum, nit_max = sl.nit_max;
    // after_step (robot_push_button.py:159-165): an active, unpressed switch is deactivated with probability disturb_prob;
    // the draw is keyed by (seed, global env, episode, control step of the episode) on the DISTURB stream
    if (A.kind == MJS_TASK_BUTTON && A.button_disturbances && T.swa && !T.swp) {
      if (uniform(io, io.env_offset + e, (unsigned)T.episode, MJS_RNG_DISTURB, (unsigned)T.len, 0.0, 1.0) < A.disturb_prob) T.swa = 0;
    }
    E::kinematics_only(&W, lane);
    observe(A, W, T, o);
    o.nit_sum = nit_sum; o.nit_max = nit_max;
    T.ret += o.reward;
    T.len += 1;
    bool bad = false;
    if (lane < NV) bad = !(isfinite(W.qvel[lane]) && isfinite(W.qpos[lane]));
    if (__any_sync(FULLMASK, bad)) { T.status |= MJS_STATUS_NAN; o.done = 1; o.truncated = 1; o.terminated = 0; }
    // A geom pair the narrowphase has no routine for (or more contacts than the solver holds) came within margin during this
    // step: the bodies may interpenetrate from here on, so the episode is cut (truncated, status bit kept) instead of
    // returning rewards of a scene that is no longer simulated faithfully.
    __syncwarp();
    if (W.status & (MJS_STATUS_UNSUPPORTED_PAIR | MJS_STATUS_CONTACT_OVERFLOW)) { o.done = 1; o.truncated = 1; o.terminated = 0; }
  }

  // Scripted demonstrators: robot_reach.py:222-253, robot_push_button.py:231-300.
  // Kinematics must be fresh.  Returns false when the task has no scripted demonstrator (Planar Push's is a
  // keyboard tele-operation, robot_planar_push.py:288-308) or the IK of Button Push's joint action fails.
  __device__ static bool demo_policy(const AM& A, const WM& W, int lane, const TS& T, const ArmIO& io, int e, double noise, double* action) {
    double tcp[3];
    tcp_position(A, W, tcp);
    for (int k = 0; k < 8; k++) action[k] = 0;
    if (A.kind == MJS_TASK_REACH) {
      double d[3], m = 0;
      for (int k = 0; k < 3; k++) {
        d[k] = T.target[k] - tcp[k];
        if (noise > 0) {  // np.random.normal(0, noise_level, 3): Box-Muller on the env's Philox stream
          double u1 = uniform(io, io.env_offset + e, io.step_index, MJS_RNG_POLICY, 2 * k, 0.0, 1.0), u2 = uniform(io, io.env_offset + e, io.step_index, MJS_RNG_POLICY, 2 * k + 1, 0.0, 1.0);
          d[k] += noise * sqrt(-2.0 * log(fmax(u1, 1e-300))) * cos(6.283185307179586 * u2);
        }
        m = fmax(m, fabs(d[k]));
      }
      // "move at most 0.5 m/s": the reference only rescales differences above 0.5 m (and then by the control period)
      if (m > 0.5) for (int k = 0; k < 3; k++) d[k] = d[k] * 0.5 / m * A.control_dt;
      for (int k = 0; k < 3; k++) action[k] = tcp[k] + d[k];
      return true;
    }
    if (A.kind != MJS_TASK_BUTTON) return false;
    // switch.get_position (switch.py:86-87, incl. its scalar-offset quirk) == the task's position observable
    const real* R = W.xmat[A.touch_slot];
    real tp[3] = {A.touch_pos[0], A.touch_pos[1], A.touch_pos[2]}, v[3];
    mat_vec(v, R, tp);
    double quirk = 0.5 * ((double)A.touch_size[0] / 1.01), sw[3], goal[3];
    for (int k = 0; k < 3; k++) sw[k] = (double)W.xpos[A.touch_slot][k] + (double)v[k] + quirk;
    double dxy = hypot(tcp[0] - sw[0], tcp[1] - sw[1]);
    int phase = T.swa ? 3 : ((tcp[2] > sw[2] && dxy < 0.01) ? 2 : 1);
    if (phase == 1) {  // first up (clear the button), then above it
      if (tcp[2] < sw[2] + 0.02) { goal[0] = tcp[0]; goal[1] = tcp[1]; } else { goal[0] = sw[0]; goal[1] = sw[1]; }
      goal[2] = sw[2] + 0.05;
    } else if (phase == 2) {
      for (int k = 0; k < 3; k++) goal[k] = sw[k];
    } else {
      for (int k = 0; k < 3; k++) goal[k] = A.button_goal[k];
      if (dxy < 0.05) goal[2] = sw[2] + 0.1;  // do not touch the button again while leaving
    }
    double d[3], m = 0, lim = 0.5 * A.control_dt;  // MAX_SPEED 0.5 m/s
    for (int k = 0; k < 3; k++) { d[k] = goal[k] - tcp[k]; m = fmax(m, fabs(d[k])); }
    if (m > lim) for (int k = 0; k < 3; k++) d[k] = d[k] * 0.5 / m * A.control_dt;
    double tgt[3] = {tcp[0] + d[0], tcp[1] + d[1], tcp[2] + d[2]}, seed[6], q[6];
    if (A.action_dim != 7) {  // ABS_EEF_ACTION: the Cartesian target itself + the closed gripper (robot_push_button.py:289-297)
      for (int k = 0; k < 3; k++) action[k] = tgt[k];
      action[3] = 0.0;
      return true;
    }
    for (int i = 0; i < 6; i++) seed[i] = W.qpos[i];
    bool ok = tcp_to_joints(A, tgt, seed, q, lane);
    for (int i = 0; i < 6; i++) action[i] = ok ? q[i] : seed[i];
    action[6] = 0.0;  // "gripper, which is always closed"
    return ok;
  }

  // Episode reset: sample boxes, IK, mj_forward, settle (Planar Push)
  __device__ static void reset_env(const AM& A, WM& W, int lane, const ArmIO& io, int e, TS& T, real& a_i, Out& o) {
    reset_begin(A, W, lane, io, e, T, a_i);
    // mj_forward at the reset state (actuator force is zero: ctrl == qpos, qvel == 0),
    // then the settle steps (robot_planar_push.py:160-161).  substep = step1 + step2, so
    // N settle steps are N substeps; the reference's extra re-solve of the forward pass
    // restarts from a converged warm start and changes nothing.
    typename E::SubRes so;
    if (A.settle_steps > 0) {
      for (int k = 0; k < A.settle_steps; k++) {
        so = E::substep(&W, lane, a_i, T.time, 1);
        a_i = so.a_i; T.time = so.time;
      }
      reset_finish(A, W, lane, T, o);
    } else {
      so = E::substep(&W, lane, a_i, T.time, 0);
      a_i = so.a_i;
      if (A.kind == MJS_TASK_BUTTON) switch_update(A, T, so.touch);
      observe(A, W, T, o);
    }
  }
  // the observation of a settled reset state (settle_steps > 0)
  __device__ static void reset_finish(const AM& A, WM& W, int lane, TS& T, Out& o) {
    E::kinematics_only(&W, lane);
    if (A.kind == MJS_TASK_BUTTON) switch_update(A, T, (real)0);
    observe(A, W, T, o);
  }
  // sampling half of a reset: spawn pose -> IK, target, props; leaves qpos/qvel/ctrl in W and T zeroed
  __device__ static void reset_begin(const AM& A, WM& W, int lane, const ArmIO& io, int e, TS& T, real& a_i) {
    unsigned genv = io.env_offset + e;
    T.episode += 1;
    unsigned ep = (unsigned)T.episode;
    unsigned idx = 0;
    double spawn[3], q[6], seed[6] = {0, 0, 0, 0, 0, 0};
    for (int k = 0; k < 3; k++) spawn[k] = uniform(io, genv, ep, MJS_RNG_RESET, idx++, A.spawn_lo[k], A.spawn_hi[k]);
    bool ok = tcp_to_joints(A, spawn, seed, q, lane);
    if (!ok) for (int k = 0; k < 6; k++) q[k] = 0;  // robot.py:179-183: unreachable reset pose is ignored
    for (int k = 0; k < 3; k++) T.target[k] = uniform(io, genv, ep, MJS_RNG_RESET, idx++, A.target_lo[k], A.target_hi[k]);
    if (A.kind == MJS_TASK_BUTTON) { for (int k = 0; k < 3; k++) { T.ovr[k] = T.target[k]; T.target[k] = 0; } }
    if (lane < ARM_NLINK) { W.qpos[lane] = q[lane]; W.ctrl[lane] = q[lane]; }
    if (lane < NV) W.qvel[lane] = 0;
    if (NOBJ > 0) {
      // robot_planar_push.py:169-176 with a conservative overlap test: resample ALL props
      // until no two props and no prop/pusher overlap in the xy-plane (superset of ncon > 0)
      double px[NOBJ > 0 ? NOBJ : 1], py[NOBJ > 0 ? NOBJ : 1], pz[NOBJ > 0 ? NOBJ : 1];
      for (int attempt = 0; attempt < 64; attempt++) {
        for (int b = 0; b < NOBJ; b++) {
          px[b] = uniform(io, genv, ep, MJS_RNG_RESET, idx++, A.obj_lo[0], A.obj_hi[0]);
          py[b] = uniform(io, genv, ep, MJS_RNG_RESET, idx++, A.obj_lo[1], A.obj_hi[1]);
          pz[b] = uniform(io, genv, ep, MJS_RNG_RESET, idx++, A.obj_lo[2], A.obj_hi[2]);
        }
        bool coll = false;
        double dmin = 2 * A.obj_radius;
        for (int b = 0; b < NOBJ; b++) {
          if (hypot(px[b] - spawn[0], py[b] - spawn[1]) < dmin) coll = true;
          for (int c = 0; c < b; c++) if (hypot(px[b] - px[c], py[b] - py[c]) < dmin) coll = true;
        }
        if (!coll) break;
      }
      if (lane < NOBJ) {
        int qa = ARM_NLINK + 7 * lane;
        W.qpos[qa] = px[lane]; W.qpos[qa + 1] = py[lane]; W.qpos[qa + 2] = pz[lane];
        W.qpos[qa + 3] = 1; W.qpos[qa + 4] = 0; W.qpos[qa + 5] = 0; W.qpos[qa + 6] = 0;
      }
    }
    T.time = 0; T.t0 = 0; T.traj = 0; T.step = 0; T.ret = 0; T.len = 0; T.tq0 = 0; T.tq1 = 0;
    T.swa = 0; T.swc = 0;  // Switch.initialize_episode keeps _is_pressed (switch.py:62-65)
    a_i = 0;
    init_static(A, W, lane, T);
    __syncwarp();
  }
};

// ========================================================================= kernels
template <typename real, int NOBJ>
__device__ __forceinline__ void arm_write_out(const ArmModel<real>& A, const ArmIO& io, int e, int lane,
                                              const typename ArmTask<real, NOBJ>::Out& o, bool write_flags) {
  int O = A.obs_dim;
  if ((io.obs || io.obs64) && lane < O) {
    double v = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) if (k == lane) v = o.obs[k];
    if (io.obs) io.obs[(size_t)e * O + lane] = (float)v;
    if (io.obs64) io.obs64[(size_t)e * O + lane] = v;
  }
  if (write_flags && lane == 0) {
    if (io.reward) io.reward[e] = (float)o.reward;
    if (io.flags) reinterpret_cast<uchar4*>(io.flags)[e] = make_uchar4(o.terminated, o.truncated, o.success, o.done);
  }
}

// Settle jobs.  An episode reset of Planar Push costs settle_steps (150) physics substeps -- six control
// steps' worth -- so the NEXT episode of every env is prepared ahead of time into a slot, n_sub substeps
// per step launch, by warps of the step kernel's own grid (full blocks that share an SM like any other
// step block; a co-running kernel of small blocks kept whole SMs away from the step blocks instead).
// arm_jobs_kernel lists the envs whose slot is empty or in preparation; job warp j takes jobs[j].
static __global__ void arm_jobs_kernel(ArmIO io) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e == 0) *io.njobs_next = 0;
  if (e >= io.N) return;
  if (*((volatile int*)(io.slot_ready + e)) != 0) return;
  int k = atomicAdd(io.njobs, 1);
  if (k < io.job_cap) io.jobs[k] = e;
}

template <typename real, int NOBJ>
__device__ __noinline__ void arm_settle_job(const ArmModel<real>& A, WarpMem<real, NOBJ>& W, const ArmIO& io, int j, int lane) {
  using TK = ArmTask<real, NOBJ>;
  using E = ArmEngine<real, NOBJ>;
  using L = ArmLayout<NOBJ>;
  int nj = *io.njobs;
  if (j >= (nj < io.job_cap ? nj : io.job_cap)) return;
  int e = io.jobs[j];
  ArmIO ios = io;
  ios.sd = io.slot_sd; ios.si = io.slot_si;
  int prog = io.slot_prog[e];
  // the env's current episode id; its own warp may be rewriting the record right now, but the id only
  // changes there on an inline reset, and a slot prepared for the wrong episode is discarded at adoption
  int episode = *((volatile int*)(io.si + (size_t)e * L::SI + L::I_EPISODE));
  typename TK::TS T;
  real a_i;
  if (prog > 0) {
    TK::load(A, W, ios, e, lane, T, a_i);
    if (T.episode != episode + 1) prog = 0;  // overtaken: start again for the right episode
  }
  if (prog == 0) {
    if (lane == 0) { W.status = 0; W.sync = 0; W.novote = 0; W.period = 1; }
    T.episode = episode; T.swp = 0; T.status = 0;
    for (int k = 0; k < 3; k++) T.ovr[k] = io.sd[(size_t)e * L::SD + L::OVR + k];  // constant per env outside Button Push
    TK::reset_begin(A, W, lane, io, e, T, a_i);
  }
  int todo = A.settle_steps - prog;
  if (todo > A.n_sub) todo = A.n_sub;
  for (int k = 0; k < todo; k++) {
    typename E::SubRes so = E::substep(&W, lane, a_i, T.time, 1);
    a_i = so.a_i; T.time = so.time;
  }
  prog += todo;
  bool finished = prog >= A.settle_steps;
  if (finished) {
    typename TK::Out o;
    TK::reset_finish(A, W, lane, T, o);
    if (lane < 16) {
      double v = 0;
#pragma unroll
      for (int k = 0; k < 16; k++) if (k == lane) v = o.obs[k];
      io.slot_obs[(size_t)e * 16 + lane] = v;
    }
  }
  T.status = 0;
  TK::store(W, ios, e, lane, T, a_i);
  __syncwarp();
  if (lane == 0) {
    io.slot_prog[e] = finished ? 0 : prog;
    if (finished) { __threadfence(); *((volatile int*)(io.slot_ready + e)) = 1; }
  }
}

// One control step of env e by the calling warp (valid == false: a surplus warp of a lock-step block
// shadowing the last env so that barrier counts match; it writes nothing).
template <typename real, int NOBJ, bool SYNC>
__device__ __noinline__ void arm_step_env(const ArmModel<real>& A, WarpMem<real, NOBJ>& W, const ArmIO& io, int e, bool valid, int lane, double* bstat) {
  using TK = ArmTask<real, NOBJ>;
  typename TK::TS T;
  real a_i;
  long long t_start = 0;
  if (io.times) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
  TK::load(A, W, io, e, lane, T, a_i, SYNC ? io.sync_level : 0);
  // action (all lanes hold the full vector)
  double action[8];
  int AD = A.action_dim;
  if (io.actions) {
#pragma unroll
    for (int k = 0; k < 8; k++) action[k] = k < AD ? (double)io.actions[(size_t)e * AD + k] : 0.0;
  } else {
    unsigned genv = io.env_offset + e;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      action[k] = 0;
      if (k < AD) {
        Philox4 r = philox4x32_10(genv, io.step_index, MJS_RNG_ACTION, k >> 1, (uint32_t)io.seed, (uint32_t)(io.seed >> 32));
        double u = (k & 1) ? u53(r.v[2], r.v[3]) : u53(r.v[0], r.v[1]);
        action[k] = A.action_lo[k] + (A.action_hi[k] - A.action_lo[k]) * u;
        if (A.kind == MJS_TASK_BUTTON && A.action_dim == 7 && k < 6) action[k] = fmin(fmax(W.qpos[k] + action[k], -3.14), 3.14);
      }
    }
  }
  typename TK::Out o;
  TK::control_step(A, W, lane, T, a_i, action, o, io, e);
  if (!valid) return;  // all barriers of this launch are behind us
  if (io.times && lane == 0) {
    long long t1; unsigned smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
#ifdef MJS_STAGE_CLOCKS
    for (int k = 0; k < MJS_NCLK; k++) io.times[MJS_TIMES_WORDS * (size_t)e + 4 + k] = W.clk[k];
#endif
    io.times[MJS_TIMES_WORDS * (size_t)e] = t_start; io.times[MJS_TIMES_WORDS * (size_t)e + 1] = t1;
    io.times[MJS_TIMES_WORDS * (size_t)e + 3] = ((long long)o.nit_sum << 32) | ((long long)(o.nit_max & 0xffff) << 16) | (smid & 0xffff);
  }
  __syncwarp();
  if (lane == 0) W.sync = 0;  // an inline reset below runs on its own: no further block barriers
  __syncwarp();
  if ((io.final_obs || io.final_obs64) && lane < A.obs_dim) {
    double v = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) if (k == lane) v = o.obs[k];
    if (io.final_obs) io.final_obs[(size_t)e * A.obs_dim + lane] = (float)v;
    if (io.final_obs64) io.final_obs64[(size_t)e * A.obs_dim + lane] = v;
  }
  if (lane == 0) {
    if (io.reward) io.reward[e] = (float)o.reward;
    if (io.reward64) io.reward64[e] = o.reward;
    if (io.flags) reinterpret_cast<uchar4*>(io.flags)[e] = make_uchar4(o.terminated, o.truncated, o.success, o.done);
    // statistics go to the block's shared accumulator (one global atomic per counter and block at the end of the kernel,
    // SURVEY.md 8e) instead of two same-address global atomics per env and launch
    if (io.stats) {  // solver effort: env-substeps and Newton iterations (mjs_stats slots 6 and 7)
      atomicAdd(bstat + 6, (double)A.n_sub); atomicAdd(bstat + 7, (double)o.nit_sum);
#ifdef MJS_CONVEX_COUNT
      atomicAdd(bstat + 3, (double)W.ncvx); atomicAdd(bstat + 5, (double)W.ncvx_hit);
#endif
#ifdef MJS_NIT_HIST
      atomicAdd(bstat + 2, (double)o.h4); atomicAdd(bstat + 3, (double)o.h8); atomicAdd(bstat + 4, (double)o.h16); atomicAdd(bstat + 5, (double)o.h32);
#endif
    }
    if (o.done && io.stats) {
      atomicAdd(bstat + 0, T.ret); atomicAdd(bstat + 1, (double)T.len); atomicAdd(bstat + 2, 1.0);
      atomicAdd(bstat + 3, (double)o.success); atomicAdd(bstat + 4, (double)o.truncated);
      atomicAdd(bstat + 5, (T.status | W.status) ? 1.0 : 0.0);
    }
  }
  if (o.done && io.autoreset) {
    using L = ArmLayout<NOBJ>;
    // Fast path: adopt the pre-settled next episode prepared by the refill kernel on
    // the side stream (Planar Push resets cost 150 physics steps; doing them inline
    // would stretch this launch 8x for the whole GPU).
    bool took = false;
    if (io.slot_ready) {
      int rdy = 0;
      if (lane == 0) {
        rdy = *((volatile int*)(io.slot_ready + e));
        if (rdy) {
          __threadfence();
          rdy = (io.slot_si[(size_t)e * L::SI + L::I_EPISODE] == T.episode + 1) ? 1 : 0;
          if (!rdy) *((volatile int*)(io.slot_ready + e)) = 0;  // stale (an inline reset overtook it): prepare it again
        }
      }
      rdy = __shfl_sync(FULLMASK, rdy, 0);
      if (rdy) {
        const double* src = io.slot_sd + (size_t)e * L::SD;
        double* dst = io.sd + (size_t)e * L::SD;
        for (int k = lane; k < L::SD; k += 32) dst[k] = src[k];
        if (lane < L::SI) io.si[(size_t)e * L::SI + lane] = io.slot_si[(size_t)e * L::SI + lane];
        if (lane < A.obs_dim) {
          const double v = io.slot_obs[(size_t)e * 16 + lane];
          if (io.obs) io.obs[(size_t)e * A.obs_dim + lane] = (float)v;
          if (io.obs64) io.obs64[(size_t)e * A.obs_dim + lane] = v;
        }
        // the status word is sticky across episodes on the inline-reset path (T.status survives reset_begin): same here
        if (lane == 0 && (T.status | W.status)) io.si[(size_t)e * L::SI + L::I_STATUS] |= T.status | W.status;
        __syncwarp();
        if (lane == 0) { __threadfence(); *((volatile int*)(io.slot_ready + e)) = 0; }
        took = true;
      }
    }
    if (took && io.times && lane == 0) { long long t2; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2)); io.times[MJS_TIMES_WORDS * (size_t)e + 2] = t2; }
    if (took) return;
    typename TK::Out o2;
    T.status |= W.status;
    TK::reset_env(A, W, lane, io, e, T, a_i, o2);
    arm_write_out<real, NOBJ>(A, io, e, lane, o2, false);
  } else {
    arm_write_out<real, NOBJ>(A, io, e, lane, o, false);
  }
  TK::store(W, io, e, lane, T, a_i);
  if (io.times && lane == 0) { long long t2; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2)); io.times[MJS_TIMES_WORDS * (size_t)e + 2] = t2; }
}

// Work items of one launch: job_blocks * WPB settle jobs first, then the N envs.
//  * static mode (io.queue == nullptr): block b / warp w owns item b * WPB + w; the warps of a block may run the
//    substep stages in lock-step (io.sync_level > 0).
//  * queue mode (io.queue != nullptr, sync_level 0): a persistent grid of one block per SM whose warps each pull the
//    next item from an atomic counter until the items run out.  A warp never waits for the slowest env of its
//    block, and the launch has no wave quantisation: its tail is one env-step instead of one block-wave.
template <typename real, int NOBJ, int WPB, bool SYNC>
__global__ void __launch_bounds__(WPB * 32) arm_step_kernel(ArmIO io) {
  const ArmModel<real>& A = cmodel<real>();
  using WM = WarpMem<real, NOBJ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WM* wm = reinterpret_cast<WM*>(smem_raw);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WM& W = wm[warp];
  // block-level statistics accumulator: every warp reaches both block barriers (no early exit from this kernel)
  __shared__ double bstat[8];
  if (threadIdx.x < 8) bstat[threadIdx.x] = 0;
  __syncthreads();
  if (io.queue) {
    const int njob = io.job_blocks * WPB, total = njob + io.N;
    for (;;) {
      int item = 0;
      if (lane == 0) item = atomicAdd(io.queue, 1);
      item = __shfl_sync(FULLMASK, item, 0);
      if (item >= total) break;
      if (item < njob) arm_settle_job<real, NOBJ>(A, W, io, item, lane);
      else arm_step_env<real, NOBJ, SYNC>(A, W, io, item - njob, true, lane, bstat);  // the host passes sync_level 0 with a queue
      __syncwarp();
    }
  } else if ((int)blockIdx.x < io.job_blocks) {  // the first blocks of the grid prepare next-episode slots (no lock-step)
    arm_settle_job<real, NOBJ>(A, W, io, blockIdx.x * WPB + warp, lane);
  } else {
    int e = ((int)blockIdx.x - io.job_blocks) * WPB + warp;
    bool valid = e < io.N;
    // lock-step mode: surplus warps shadow the last env (barrier counts must match), write nothing
    if (valid || SYNC) arm_step_env<real, NOBJ, SYNC>(A, W, io, valid ? e : io.N - 1, valid, lane, bstat);
  }
  __syncthreads();
  if (io.stats && threadIdx.x < 8 && bstat[threadIdx.x] != 0) atomicAdd(io.stats + threadIdx.x, bstat[threadIdx.x]);
}

// refill == 0: reset the masked envs in place (mjs_reset).
// refill == 1: for every env whose slot is empty, generate its NEXT episode (sampling,
//              IK, mj_forward, settle steps) into the slot arrays and mark it ready.
template <typename real, int NOBJ, int WPB>
__global__ void __launch_bounds__(WPB * 32) arm_reset_kernel(ArmIO io, int refill) {
  const ArmModel<real>& A = cmodel<real>();
  using TK = ArmTask<real, NOBJ>;
  using WM = WarpMem<real, NOBJ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WM* wm = reinterpret_cast<WM*>(smem_raw);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e = blockIdx.x * WPB + warp;
  if (e >= io.N) return;
  if (refill) {
    if (*((volatile int*)(io.slot_ready + e)) != 0) return;
  } else if (io.mask && !io.mask[e]) return;
  WM& W = wm[warp];
  typename TK::TS T;
  real a_i;
  TK::load(A, W, io, e, lane, T, a_i);
  typename TK::Out o;
  TK::reset_env(A, W, lane, io, e, T, a_i, o);
  if (refill) {
    ArmIO io2 = io;
    io2.sd = io.slot_sd; io2.si = io.slot_si;
    T.status = 0;
    TK::store(W, io2, e, lane, T, a_i);
    if (lane < 16) {
      double v = 0;
#pragma unroll
      for (int k = 0; k < 16; k++) if (k == lane) v = o.obs[k];
      io.slot_obs[(size_t)e * 16 + lane] = v;
    }
    __syncwarp();
    if (lane == 0) { io.slot_prog[e] = 0; __threadfence(); *((volatile int*)(io.slot_ready + e)) = 1; }
  } else {
    arm_write_out<real, NOBJ>(A, io, e, lane, o, false);
    TK::store(W, io, e, lane, T, a_i);
    if (io.slot_ready && lane == 0) { io.slot_prog[e] = 0; *((volatile int*)(io.slot_ready + e)) = 0; }  // a stale slot would repeat this episode
  }
}

// Test hook (mjs_debug_substeps): continue the current control step by k physics substeps without
// the before_step hook, so a control step can be split at a substep boundary.
template <typename real, int NOBJ, int WPB>
__global__ void __launch_bounds__(WPB * 32) arm_substeps_kernel(ArmIO io, int k) {
  const ArmModel<real>& A = cmodel<real>();
  using TK = ArmTask<real, NOBJ>;
  using E = ArmEngine<real, NOBJ>;
  using WM = WarpMem<real, NOBJ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WM* wm = reinterpret_cast<WM*>(smem_raw);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e = blockIdx.x * WPB + warp;
  if (e >= io.N) return;
  WM& W = wm[warp];
  typename TK::TS T;
  real a_i;
  TK::load(A, W, io, e, lane, T, a_i);
  for (int i = 0; i < k; i++) {
    TK::before_substep(A, W, lane, T);
    typename E::SubRes so = E::substep(&W, lane, a_i, T.time, 1);
    a_i = so.a_i; T.time = so.time;
    if (A.kind == MJS_TASK_BUTTON) TK::switch_update(A, T, so.touch);
  }
  TK::store(W, io, e, lane, T, a_i);
}

// mjs_demo_policy: the task's scripted demonstrator evaluated on the current state of every env
template <typename real, int NOBJ, int WPB>
__global__ void __launch_bounds__(WPB * 32) arm_policy_kernel(ArmIO io, float* actions, int* ok_out, double noise) {
  const ArmModel<real>& A = cmodel<real>();
  using TK = ArmTask<real, NOBJ>;
  using E = ArmEngine<real, NOBJ>;
  using WM = WarpMem<real, NOBJ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WM* wm = reinterpret_cast<WM*>(smem_raw);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e = blockIdx.x * WPB + warp;
  if (e >= io.N) return;
  WM& W = wm[warp];
  typename TK::TS T;
  real a_i;
  TK::load(A, W, io, e, lane, T, a_i);
  E::kinematics_only(&W, lane);
  double action[8];
  bool ok = TK::demo_policy(A, W, lane, T, io, e, noise, action);
  if (lane < A.action_dim) {
    double v = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) if (k == lane) v = action[k];
    actions[(size_t)e * A.action_dim + lane] = (float)v;
  }
  if (ok_out && lane == 0) ok_out[e] = ok ? 1 : 0;
}

// mjs_render, first half: kinematics of the current state -> world pose of every model geom + the per-env anchors
template <typename real, int NOBJ, int WPB>
__global__ void __launch_bounds__(WPB * 32) arm_pose_kernel(ArmIO io) {
  const ArmModel<real>& A = cmodel<real>();
  using TK = ArmTask<real, NOBJ>;
  using E = ArmEngine<real, NOBJ>;
  using WM = WarpMem<real, NOBJ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WM* wm = reinterpret_cast<WM*>(smem_raw);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e = blockIdx.x * WPB + warp;
  if (e >= io.N) return;
  WM& W = wm[warp];
  typename TK::TS T;
  real a_i;
  TK::load(A, W, io, e, lane, T, a_i);
  E::kinematics_only(&W, lane);
  float* gp = io.gpose + (size_t)e * MJS_MAXGEOM * 12;
  if (lane < A.ngeom) {
    for (int k = 0; k < 3; k++) gp[lane * 12 + k] = (float)W.gpos[lane][k];
    for (int k = 0; k < 9; k++) gp[lane * 12 + 3 + k] = (float)W.gmat[lane][k];
  }
  if (lane == 0) {
    float* a = io.aux + (size_t)e * 8;
    for (int k = 0; k < 3; k++) { a[k] = (float)T.target[k]; a[3 + k] = (float)T.ovr[k]; }
    a[6] = (float)T.swa; a[7] = 0;
  }
}

// mj_forward on injected states (mjs_set_state with forward != 0): refresh the warm start
template <typename real, int NOBJ, int WPB>
__global__ void __launch_bounds__(WPB * 32) arm_forward_kernel(ArmIO io, int first, int count) {
  const ArmModel<real>& A = cmodel<real>();
  using TK = ArmTask<real, NOBJ>;
  using E = ArmEngine<real, NOBJ>;
  using WM = WarpMem<real, NOBJ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WM* wm = reinterpret_cast<WM*>(smem_raw);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e = first + blockIdx.x * WPB + warp;
  if (e >= first + count || e >= io.N) return;
  WM& W = wm[warp];
  typename TK::TS T;
  real a_i;
  TK::load(A, W, io, e, lane, T, a_i);
  typename E::SubRes so = E::substep(&W, lane, a_i, T.time, 0);
  a_i = so.a_i;
  TK::store(W, io, e, lane, T, a_i);
}

// debug: run mj_forward for one env and dump intermediate stages (tests only)
//   out: [0..35] Marm, [36..36+NV) fsm, [..+NV) qacc, [..+NV) fcon, then ncon, then per contact (dist,pos3,normal3) x ncon, then link xpos (18), S (36)
template <typename real, int NOBJ, int WPB>
__global__ void __launch_bounds__(WPB * 32) arm_debug_kernel(ArmIO io, int env, double* out, int pre) {
  const ArmModel<real>& A = cmodel<real>();
  using TK = ArmTask<real, NOBJ>;
  using E = ArmEngine<real, NOBJ>;
  using WM = WarpMem<real, NOBJ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WM* wm = reinterpret_cast<WM*>(smem_raw);
  int lane = threadIdx.x & 31;
  if (threadIdx.x >= 32) return;
  WM& W = wm[0];
  typename TK::TS T;
  real a_i;
  TK::load(A, W, io, env, lane, T, a_i);
  for (int i = 0; i < pre; i++) {  // continue the control step in-kernel first (no store)
    TK::before_substep(A, W, lane, T);
    typename E::SubRes s1 = E::substep(&W, lane, a_i, T.time, 1);
    a_i = s1.a_i; T.time = s1.time;
  }
  if (pre) TK::before_substep(A, W, lane, T);
  typename E::SubRes so = E::substep(&W, lane, a_i, T.time, 0);
  __syncwarp();
  constexpr int NV = WM::NV;
  if (lane == 0) {
    int p = 0;
    for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) out[p++] = W.Marm[i][j];
    for (int i = 0; i < NV; i++) out[p++] = W.fsm[i];
    for (int i = 0; i < NV; i++) out[p++] = W.a[i];
    for (int i = 0; i < NV; i++) out[p++] = W.fcon[i];
    out[p++] = W.ncon;
    out[p++] = so.niter;
    out[p++] = so.touch;
    for (int c = 0; c < ARM_MAXC; c++) for (int k = 0; k < 7; k++) out[p++] = c < W.ncon ? (double)W.craw[k][c] : 0.0;
    for (int i = 0; i < 6; i++) for (int k = 0; k < 3; k++) out[p++] = W.xpos[SLOT_LINK0 + i][k];
    for (int i = 0; i < 6; i++) for (int k = 0; k < 6; k++) out[p++] = W.S[i][k];
    for (int c = 0; c < ARM_MAXC; c++) for (int k = 0; k < 4; k++) out[p++] = c < W.ncon ? (double)W.cf[k][c] : 0.0;
  }
}
