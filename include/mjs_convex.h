/* mjs_convex.h -- penetration of two convex primitives by Minkowski Portal Refinement (MPR / XenoCollide).
 *
 * MuJoCo sends every pair it has no analytic routine for (cylinder x cylinder, cylinder x box, capsule x cylinder, and all
 * mesh pairs) through a general convex routine (mjc_Convex over libccd's ccdMPRPenetration, tolerance 1e-6, 50 iterations in
 * mjOption) [UPSTREAM-RECALL: neither MuJoCo nor libccd is in /root/reference or in this image].  This header restates that
 * algorithm -- portal discovery, portal refinement, penetration depth / direction / position from the final portal -- for
 * support-mapped primitives, ONE source compiled both into the CPU oracle (g++) and into the CUDA step kernel (nvcc), so
 * the two sides run the same arithmetic.
 *
 * Conventions: shapes are given by (type, size[3], pos[3], mat[9] row-major world rotation); the support point of the
 * Minkowski difference D = shape1 - shape2 in direction d is s1(d) - s2(-d).  Result: depth >= 0, dir = unit vector from
 * shape 1 towards shape 2 (moving shape 2 by depth * dir separates them), pos = midpoint of the two witness points.
 */
#ifndef MJS_CONVEX_H_
#define MJS_CONVEX_H_

#include <math.h>

#include "mjs_model.h"

#ifdef __CUDACC__
#define MJS_HD __host__ __device__
#define MJS_UNROLL _Pragma("unroll")   /* every array below must be indexed with compile-time constants to stay in registers */
#else
#define MJS_HD
#define MJS_UNROLL
#endif

#define MJS_CCD_TOLERANCE 1e-6
#define MJS_CCD_ITERATIONS 50   /* cap of each of the three loops (discovery, refinement, penetration) */

template <typename T>
struct MjsShape {
  int type;
  T size[3], pos[3], mat[9];
};

template <typename T>
struct MjsSupport {  /* a vertex of the Minkowski difference with its two witnesses */
  T v[3], v1[3], v2[3];
};

namespace mjsccd {
template <typename T> MJS_HD inline T dot(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename T> MJS_HD inline void cross(T* r, const T* a, const T* b) {
  T x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename T> MJS_HD inline void sub(T* r, const T* a, const T* b) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
template <typename T> MJS_HD inline bool normalize(T* a) {
  T n = sqrt(dot(a, a));
  if (n < (T)1e-30) return false;
  a[0] /= n; a[1] /= n; a[2] /= n;
  return true;
}

/* support point of one primitive in world direction d (need not be unit length) */
template <typename T>
MJS_HD inline void shape_support(const MjsShape<T>& s, const T* d, T* out) {
  const T* m = s.mat;
  T l[3] = {m[0] * d[0] + m[3] * d[1] + m[6] * d[2], m[1] * d[0] + m[4] * d[1] + m[7] * d[2], m[2] * d[0] + m[5] * d[1] + m[8] * d[2]};  /* R^T d */
  T p[3] = {0, 0, 0};
  if (s.type == MJS_GEOM_SPHERE || s.type == MJS_GEOM_CAPSULE) {
    T n = sqrt(dot(l, l));
    if (n > (T)1e-30) { p[0] = s.size[0] * l[0] / n; p[1] = s.size[0] * l[1] / n; p[2] = s.size[0] * l[2] / n; }
    if (s.type == MJS_GEOM_CAPSULE) p[2] += l[2] >= 0 ? s.size[1] : -s.size[1];
  } else if (s.type == MJS_GEOM_CYLINDER) {
    T n = sqrt(l[0] * l[0] + l[1] * l[1]);
    if (n > (T)1e-30) { p[0] = s.size[0] * l[0] / n; p[1] = s.size[0] * l[1] / n; }
    p[2] = l[2] >= 0 ? s.size[1] : -s.size[1];
  } else {  /* MJS_GEOM_BOX */
    p[0] = l[0] >= 0 ? s.size[0] : -s.size[0];
    p[1] = l[1] >= 0 ? s.size[1] : -s.size[1];
    p[2] = l[2] >= 0 ? s.size[2] : -s.size[2];
  }
  out[0] = s.pos[0] + m[0] * p[0] + m[1] * p[1] + m[2] * p[2];
  out[1] = s.pos[1] + m[3] * p[0] + m[4] * p[1] + m[5] * p[2];
  out[2] = s.pos[2] + m[6] * p[0] + m[7] * p[1] + m[8] * p[2];
}

#ifdef MJS_CCD_STATS  /* host-side experiments: number of support evaluations */
static long mjs_ccd_nsupport = 0;
#define MJS_CCD_COUNT() (++mjs_ccd_nsupport)
#else
#define MJS_CCD_COUNT() ((void)0)
#endif

template <typename T>
MJS_HD inline void support(const MjsShape<T>& a, const MjsShape<T>& b, const T* d, MjsSupport<T>& s) {
  MJS_CCD_COUNT();
  T nd[3] = {-d[0], -d[1], -d[2]};
  shape_support(a, d, s.v1);
  shape_support(b, nd, s.v2);
  sub(s.v, s.v1, s.v2);
}

template <typename T>
MJS_HD inline void portal_dir(const MjsSupport<T>* P, T* dir) {  /* normal of (v1, v2, v3), pointing away from v0 */
  T a[3], b[3];
  sub(a, P[2].v, P[1].v);
  sub(b, P[3].v, P[1].v);
  cross(dir, a, b);
  normalize(dir);
}

template <typename T>
MJS_HD inline void expand_portal(MjsSupport<T>* P, const MjsSupport<T>& v4) {
  T v4v0[3];
  cross(v4v0, v4.v, P[0].v);
  if (dot(P[1].v, v4v0) > 0) {
    if (dot(P[2].v, v4v0) > 0) P[1] = v4; else P[3] = v4;
  } else {
    if (dot(P[3].v, v4v0) > 0) P[2] = v4; else P[1] = v4;
  }
}

template <typename T>
MJS_HD inline bool reach_tolerance(const MjsSupport<T>* P, const MjsSupport<T>& v4, const T* dir) {
  T dv4 = dot(v4.v, dir);
  T d1 = dv4 - dot(P[1].v, dir), d2 = dv4 - dot(P[2].v, dir), d3 = dv4 - dot(P[3].v, dir);
  T mn = d1 < d2 ? d1 : d2;
  mn = mn < d3 ? mn : d3;
  return mn <= (T)MJS_CCD_TOLERANCE;
}

/* squared distance of the origin to triangle (a, b, c) and the closest point (the witness) */
template <typename T>
MJS_HD inline T origin_tri_dist2(const T* a, const T* b, const T* c, T* witness) {
  T ab[3], ac[3], ap[3] = {-a[0], -a[1], -a[2]};
  sub(ab, b, a); sub(ac, c, a);
  T d1 = dot(ab, ap), d2 = dot(ac, ap);
  T u = 0, v = 0;  /* closest = a + u ab + v ac */
  if (d1 <= 0 && d2 <= 0) { u = 0; v = 0; }
  else {
    T bp[3] = {-b[0], -b[1], -b[2]};
    T d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0 && d4 <= d3) { u = 1; v = 0; }
    else {
      T vc = d1 * d4 - d3 * d2;
      if (vc <= 0 && d1 >= 0 && d3 <= 0) { u = d1 / (d1 - d3); v = 0; }
      else {
        T cp[3] = {-c[0], -c[1], -c[2]};
        T d5 = dot(ab, cp), d6 = dot(ac, cp);
        if (d6 >= 0 && d5 <= d6) { u = 0; v = 1; }
        else {
          T vb = d5 * d2 - d1 * d6;
          if (vb <= 0 && d2 >= 0 && d6 <= 0) { u = 0; v = d2 / (d2 - d6); }
          else {
            T va = d3 * d6 - d5 * d4;
            if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) { T w = (d4 - d3) / ((d4 - d3) + (d5 - d6)); u = 1 - w; v = w; }
            else { T den = 1 / (va + vb + vc); u = vb * den; v = vc * den; }
          }
        }
      }
    }
  }
  witness[0] = a[0] + u * ab[0] + v * ac[0];
  witness[1] = a[1] + u * ab[1] + v * ac[1];
  witness[2] = a[2] + u * ab[2] + v * ac[2];
  return dot(witness, witness);
}

template <typename T>
MJS_HD inline void find_pos(const MjsSupport<T>* P, T* pos) {
  T dir[3], t[3];
  portal_dir(P, dir);
  T b0, b1, b2, b3;
  cross(t, P[1].v, P[2].v); b0 = dot(t, P[3].v);
  cross(t, P[3].v, P[2].v); b1 = dot(t, P[0].v);
  cross(t, P[0].v, P[1].v); b2 = dot(t, P[3].v);
  cross(t, P[2].v, P[1].v); b3 = dot(t, P[0].v);
  T sum = b0 + b1 + b2 + b3;
  if (sum <= 0) {
    b0 = 0;
    cross(t, P[2].v, P[3].v); b1 = dot(t, dir);
    cross(t, P[3].v, P[1].v); b2 = dot(t, dir);
    cross(t, P[1].v, P[2].v); b3 = dot(t, dir);
    sum = b1 + b2 + b3;
  }
  T inv = 1 / sum;
  MJS_UNROLL for (int k = 0; k < 3; k++) {
    T p1 = b0 * P[0].v1[k] + b1 * P[1].v1[k] + b2 * P[2].v1[k] + b3 * P[3].v1[k];
    T p2 = b0 * P[0].v2[k] + b1 * P[1].v2[k] + b2 * P[2].v2[k] + b3 * P[3].v2[k];
    pos[k] = (T)0.5 * (p1 + p2) * inv;
  }
}
}  // namespace mjsccd

/* Returns 1 and (depth, dir, pos) when the shapes intersect, 0 otherwise.  On a miss decided by a support point that does not
 * reach the origin, dir_out holds that direction: a separating axis of the two shapes (zero when the miss was decided by the
 * tolerance or the iteration cap).  Callers may test it first next time (sound: a separating axis proves a miss). */
template <typename T>
MJS_HD inline int mjs_mpr_penetration(const MjsShape<T>& A, const MjsShape<T>& B, T* depth, T* dir_out, T* pos) {
  using namespace mjsccd;
  MjsSupport<T> P[4], v4;
  T dir[3], va[3], vb[3];
  /* ---- portal discovery */
  MJS_UNROLL for (int k = 0; k < 3; k++) { P[0].v1[k] = A.pos[k]; P[0].v2[k] = B.pos[k]; P[0].v[k] = A.pos[k] - B.pos[k]; }
  if (dot(P[0].v, P[0].v) < (T)1e-20) P[0].v[0] += (T)1e-5;  /* centres coincide: nudge */
  dir[0] = -P[0].v[0]; dir[1] = -P[0].v[1]; dir[2] = -P[0].v[2];
  normalize(dir);
  dir_out[0] = dir_out[1] = dir_out[2] = 0;
  support(A, B, dir, P[1]);
  if (dot(P[1].v, dir) <= 0) { dir_out[0] = dir[0]; dir_out[1] = dir[1]; dir_out[2] = dir[2]; return 0; }
  cross(dir, P[0].v, P[1].v);
  if (dot(dir, dir) < (T)1e-20) {
    /* the origin lies on the ray v0 -> v1: the penetration is along that ray */
    T n = sqrt(dot(P[1].v, P[1].v));
    *depth = n;
    if (n > (T)1e-30) { dir_out[0] = P[1].v[0] / n; dir_out[1] = P[1].v[1] / n; dir_out[2] = P[1].v[2] / n; }
    else { dir_out[0] = -dir[0]; dir_out[1] = -dir[1]; dir_out[2] = -dir[2]; }
    MJS_UNROLL for (int k = 0; k < 3; k++) pos[k] = (T)0.5 * (P[1].v1[k] + P[1].v2[k]);
    return 1;
  }
  normalize(dir);
  support(A, B, dir, P[2]);
  if (dot(P[2].v, dir) <= 0) { dir_out[0] = dir[0]; dir_out[1] = dir[1]; dir_out[2] = dir[2]; return 0; }
  sub(va, P[1].v, P[0].v); sub(vb, P[2].v, P[0].v);
  cross(dir, va, vb);
  normalize(dir);
  if (dot(dir, P[0].v) > 0) {  /* orient the portal away from the interior point */
    MjsSupport<T> t = P[1]; P[1] = P[2]; P[2] = t;
    dir[0] = -dir[0]; dir[1] = -dir[1]; dir[2] = -dir[2];
  }
  for (int it = 0; it < MJS_CCD_ITERATIONS; it++) {
    support(A, B, dir, P[3]);
    if (dot(P[3].v, dir) <= 0) { dir_out[0] = dir[0]; dir_out[1] = dir[1]; dir_out[2] = dir[2]; return 0; }
    bool cont = false;
    cross(va, P[1].v, P[3].v);
    if (dot(va, P[0].v) < 0) { P[2] = P[3]; cont = true; }
    if (!cont) {
      cross(va, P[3].v, P[2].v);
      if (dot(va, P[0].v) < 0) { P[1] = P[3]; cont = true; }
    }
    if (!cont) break;
    sub(va, P[1].v, P[0].v); sub(vb, P[2].v, P[0].v);
    cross(dir, va, vb);
    normalize(dir);
  }
  /* ---- portal refinement: is the origin inside? */
  for (int it = 0;; it++) {
    portal_dir(P, dir);
    if (dot(dir, P[1].v) >= 0) break;  /* the portal encapsulates the origin */
    support(A, B, dir, v4);
    if (dot(v4.v, dir) < 0) { dir_out[0] = dir[0]; dir_out[1] = dir[1]; dir_out[2] = dir[2]; return 0; }
    if (reach_tolerance(P, v4, dir) || it > MJS_CCD_ITERATIONS) return 0;
    expand_portal(P, v4);
  }
  /* ---- penetration: push the portal to the surface */
  for (int it = 0;; it++) {
    portal_dir(P, dir);
    support(A, B, dir, v4);
    if (reach_tolerance(P, v4, dir) || it > MJS_CCD_ITERATIONS) {
      T w[3];
      T d2 = origin_tri_dist2(P[1].v, P[2].v, P[3].v, w);
      *depth = sqrt(d2);
      if (d2 < (T)1e-30) { dir_out[0] = dir[0]; dir_out[1] = dir[1]; dir_out[2] = dir[2]; }
      else { dir_out[0] = w[0]; dir_out[1] = w[1]; dir_out[2] = w[2]; normalize(dir_out); }
      find_pos(P, pos);
      return 1;
    }
    expand_portal(P, v4);
  }
}

#endif /* MJS_CONVEX_H_ */
