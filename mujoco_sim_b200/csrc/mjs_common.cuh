// mjs_common.cuh -- shared device helpers: Philox counter RNG, small math,
// constraint impedance/reference formulas.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mjs_model.h"

#define MJS_MINVAL 1e-15
// plane x cylinder: tilt (rad) below which the cylinder lies flat on the plane.  MuJoCo uses
// mjMINVAL (1e-15), which rounding noise of a settled prop crosses at random; see DESIGN.md.
#define MJS_CYL_FLAT_TILT 1e-9
#define MJS_MINIMP 0.0001
#define MJS_MAXIMP 0.9999
#define MJS_MINMU 1e-5

// ---------------------------------------------------------------- Philox4x32-10
// Counter-based RNG: key = (seed lo, seed hi), counter = (global env id,
// episode id, purpose, index).  Results are independent of the GPU count
//  and are restated in numpy by tests/philox_ref.py.
struct Philox4 {
  uint32_t v[4];
};
__host__ __device__ inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
    uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0, hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  Philox4 o;
  o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
  return o;
}
// 53-bit uniform in [0,1) from two words (same construction as numpy's random_sample)
__host__ __device__ inline double u53(uint32_t a, uint32_t b) {
  return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}
#define MJS_RNG_RESET 1u
#define MJS_RNG_ACTION 2u
#define MJS_RNG_DISTURB 3u
#define MJS_RNG_POLICY 4u

// ------------------------------------------------------------------ impedance
template <typename real>
__host__ __device__ inline real mjs_impedance(const real* si, real pos, real margin) {
  real d0 = fmin(fmax(si[0], (real)MJS_MINIMP), (real)MJS_MAXIMP), dw = fmin(fmax(si[1], (real)MJS_MINIMP), (real)MJS_MAXIMP);
  real width = fmax(si[2], (real)MJS_MINVAL), mid = fmin(fmax(si[3], (real)MJS_MINIMP), (real)MJS_MAXIMP), power = fmax(si[4], (real)1);
  if (d0 == dw || width <= (real)MJS_MINVAL) return (real)0.5 * (d0 + dw);
  real x = fabs(pos - margin) / width;
  if (x >= 1) return dw;
  if (x <= 0) return d0;
  real y;
  if (power == 1) y = x;
  else if (power == 2) {  // the default solimp: avoid pow()
    if (x <= mid) y = x * x / mid;
    else y = 1 - (1 - x) * (1 - x) / (1 - mid);
  } else if (x <= mid) y = pow(x, power) / pow(mid, power - 1);
  else y = 1 - pow(1 - x, power) / pow(1 - mid, power - 1);
  return d0 + y * (dw - d0);
}

// stiffness k and damping b of the reference acceleration (refsafe applied)
template <typename real>
__host__ __device__ inline void mjs_kb(const real* solref, const real* solimp, real timestep, real* k, real* b) {
  real dmax = fmin(fmax(solimp[1], (real)MJS_MINIMP), (real)MJS_MAXIMP);
  if (solref[0] > 0) {
    real tc = fmax(solref[0], 2 * timestep);
    real dr = fmax((real)MJS_MINVAL, solref[1]);
    *k = 1 / fmax((real)MJS_MINVAL, dmax * dmax * tc * tc * dr * dr);
    *b = 2 / fmax((real)MJS_MINVAL, dmax * tc);
  } else {
    *k = -solref[0] / fmax((real)MJS_MINVAL, dmax * dmax);
    *b = -solref[1] / fmax((real)MJS_MINVAL, dmax);
  }
}

#define MJS_CUDA_CHECK(expr)                                                          \
  do {                                                                                \
    cudaError_t _e = (expr);                                                          \
    if (_e != cudaSuccess) {                                                          \
      mjs_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return -1;                                                                      \
    }                                                                                 \
  } while (0)

void mjs_set_error(const char* fmt, ...);
