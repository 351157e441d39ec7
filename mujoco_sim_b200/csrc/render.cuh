// render.cuh -- camera observations as CUDA ray casting of the scene's primitives.
//
// Replaces entities/camera.py:94-103 (`physics.render(height, width, camera_id)`, channel-last uint8 RGB) for the fixed
// cameras of the tasks (point_reach.py:22,95-98; robot_reach.py:52; robot_planar_push.py:42-43; robot_push_button.py:87-96).
// MuJoCo's OpenGL renderer is unavailable on this box and is not bit-matchable; this is an analytic ray caster with its own,
// stated shading model, checked against a numpy restatement of the same math (oracle/render_ref.py):
//   pinhole camera looking along its -z axis, +y up, vertical field of view fovy, square pixels, ray through pixel centres;
//   nearest hit among plane (finite when its half-sizes are positive) / sphere / capsule / cylinder / box;
//   colour = rgb * (0.3 + 0.7 * |n . d|) (ambient + head light, two-sided normals), background black,
//   alpha ignored, 8-bit = floor(255 * min(c, 1) + 0.5).
// Two kernels: a pose kernel per engine writes the world pose of every model geom of every env (float [N][MJS_MAXGEOM][12]
// + per-env anchors [N][8]: target xyz, switch xyz, switch active), the render kernel below shades pixels from them.
#pragma once
#include "mjs_common.cuh"

#define RENDER_AMBIENT 0.3f
#define RENDER_DIFFUSE 0.7f

struct RenderVis {  // device copy of MjsVisual in float
  int type, geom, anchor, use_active;
  float pos[3], mat[9], size[3], rgb[3], rgb_active[3];
};
struct RenderScene {
  int nvis, width, height, cam_anchor, cam_geom;
  float cam_pos[3], cam_mat[9], tan_half_fovy;
  RenderVis vis[MJS_MAXVIS];
};

__device__ __forceinline__ void r_matvec(float* r, const float* m, const float* v) {
  float x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2], z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
__device__ __forceinline__ void r_matTvec(float* r, const float* m, const float* v) {
  float x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2], y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2], z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
__device__ __forceinline__ void r_matmul(float* r, const float* a, const float* b) {
  float t[9];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) t[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
#pragma unroll
  for (int i = 0; i < 9; i++) r[i] = t[i];
}

// nearest intersection t > 0 of the ray o + t d (local frame) with one primitive; writes the local outward normal
__device__ __forceinline__ bool r_hit(int type, const float* size, const float* o, const float* d, float& tbest, float* nrm) {
  const float EPS = 1e-6f;
  bool hit = false;
  if (type == MJS_GEOM_PLANE) {
    if (fabsf(d[2]) > EPS) {
      float t = -o[2] / d[2];
      if (t > EPS && t < tbest) {
        float x = o[0] + t * d[0], y = o[1] + t * d[1];
        if ((size[0] <= 0 || fabsf(x) <= size[0]) && (size[1] <= 0 || fabsf(y) <= size[1])) { tbest = t; nrm[0] = 0; nrm[1] = 0; nrm[2] = 1; hit = true; }
      }
    }
  } else if (type == MJS_GEOM_SPHERE) {
    float b = o[0] * d[0] + o[1] * d[1] + o[2] * d[2], c = o[0] * o[0] + o[1] * o[1] + o[2] * o[2] - size[0] * size[0];
    float disc = b * b - c;
    if (disc >= 0) {
      float t = -b - sqrtf(disc);
      if (t > EPS && t < tbest) { tbest = t; for (int k = 0; k < 3; k++) nrm[k] = (o[k] + t * d[k]) / size[0]; hit = true; }
    }
  } else if (type == MJS_GEOM_CYLINDER || type == MJS_GEOM_CAPSULE) {
    const float r = size[0], h = size[1];
    float a = d[0] * d[0] + d[1] * d[1], b = o[0] * d[0] + o[1] * d[1], c = o[0] * o[0] + o[1] * o[1] - r * r;
    if (a > EPS * EPS) {
      float disc = b * b - a * c;
      if (disc >= 0) {
        float t = (-b - sqrtf(disc)) / a, z = o[2] + t * d[2];
        if (t > EPS && t < tbest && fabsf(z) <= h) { tbest = t; nrm[0] = (o[0] + t * d[0]) / r; nrm[1] = (o[1] + t * d[1]) / r; nrm[2] = 0; hit = true; }
      }
    }
    for (int s = -1; s <= 1; s += 2) {
      if (type == MJS_GEOM_CYLINDER) {  // flat caps
        if (fabsf(d[2]) > EPS) {
          float t = (s * h - o[2]) / d[2], x = o[0] + t * d[0], y = o[1] + t * d[1];
          if (t > EPS && t < tbest && x * x + y * y <= r * r) { tbest = t; nrm[0] = 0; nrm[1] = 0; nrm[2] = (float)s; hit = true; }
        }
      } else {                         // hemispherical caps
        float oc[3] = {o[0], o[1], o[2] - s * h};
        float bb = oc[0] * d[0] + oc[1] * d[1] + oc[2] * d[2], cc = oc[0] * oc[0] + oc[1] * oc[1] + oc[2] * oc[2] - r * r;
        float disc = bb * bb - cc;
        if (disc >= 0) {
          float t = -bb - sqrtf(disc), z = oc[2] + t * d[2];
          if (t > EPS && t < tbest && z * s >= 0) { tbest = t; for (int k = 0; k < 3; k++) nrm[k] = (oc[k] + t * d[k]) / r; hit = true; }
        }
      }
    }
  } else if (type == MJS_GEOM_BOX) {
    float tn = -1e30f, tf = 1e30f;
    int ax = 0; float sg = 1;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      if (fabsf(d[k]) < EPS) { if (fabsf(o[k]) > size[k]) { tn = 1e30f; } continue; }
      float t1 = (-size[k] - o[k]) / d[k], t2 = (size[k] - o[k]) / d[k];
      float lo = fminf(t1, t2), hi = fmaxf(t1, t2);
      if (lo > tn) { tn = lo; ax = k; sg = d[k] > 0 ? -1.f : 1.f; }
      tf = fminf(tf, hi);
    }
    if (tn <= tf && tn > EPS && tn < tbest) { tbest = tn; nrm[0] = nrm[1] = nrm[2] = 0; nrm[ax] = sg; hit = true; }
  }
  return hit;
}

// grid (N, ceil(W*H / 128)), block 128: one thread per pixel; the env's resolved visuals sit in shared memory
static __global__ void __launch_bounds__(128) render_kernel(RenderScene S, const float* gpose, const float* aux, unsigned char* rgb) {
  __shared__ float v_pos[MJS_MAXVIS][3], v_mat[MJS_MAXVIS][9], v_rgb[MJS_MAXVIS][3], c_pos[3], c_mat[9];
  const int e = blockIdx.x;
  const float* gp = gpose + (size_t)e * MJS_MAXGEOM * 12;
  const float* ax = aux + (size_t)e * 8;
  for (int i = threadIdx.x; i < S.nvis; i += blockDim.x) {
    const RenderVis& V = S.vis[i];
    float p[3] = {V.pos[0], V.pos[1], V.pos[2]}, m[9];
    for (int k = 0; k < 9; k++) m[k] = V.mat[k];
    if (V.anchor == MJS_ANCHOR_GEOM) {
      for (int k = 0; k < 3; k++) p[k] = gp[V.geom * 12 + k];
      for (int k = 0; k < 9; k++) m[k] = gp[V.geom * 12 + 3 + k];
    } else if (V.anchor == MJS_ANCHOR_TARGET) {
      p[0] += ax[0]; p[1] += ax[1]; p[2] += ax[2];
    } else if (V.anchor == MJS_ANCHOR_OVR) {
      p[0] += ax[3]; p[1] += ax[4]; p[2] += ax[5];
    } else if (V.anchor == MJS_ANCHOR_GEOMFRAME) {
      float w[3];
      r_matvec(w, gp + V.geom * 12 + 3, p);
      for (int k = 0; k < 3; k++) p[k] = gp[V.geom * 12 + k] + w[k];
      r_matmul(m, gp + V.geom * 12 + 3, m);
    }
    const bool act = V.use_active && ax[6] != 0.f;
    for (int k = 0; k < 3; k++) { v_pos[i][k] = p[k]; v_rgb[i][k] = act ? V.rgb_active[k] : V.rgb[k]; }
    for (int k = 0; k < 9; k++) v_mat[i][k] = m[k];
  }
  if (threadIdx.x == 0) {
    float p[3] = {S.cam_pos[0], S.cam_pos[1], S.cam_pos[2]}, m[9];
    for (int k = 0; k < 9; k++) m[k] = S.cam_mat[k];
    if (S.cam_anchor == MJS_ANCHOR_GEOMFRAME) {
      float w[3];
      r_matvec(w, gp + S.cam_geom * 12 + 3, p);
      for (int k = 0; k < 3; k++) p[k] = gp[S.cam_geom * 12 + k] + w[k];
      r_matmul(m, gp + S.cam_geom * 12 + 3, m);
    }
    for (int k = 0; k < 3; k++) c_pos[k] = p[k];
    for (int k = 0; k < 9; k++) c_mat[k] = m[k];
  }
  __syncthreads();
  const int pix = blockIdx.y * blockDim.x + threadIdx.x;
  if (pix >= S.width * S.height) return;
  const int row = pix / S.width, col = pix % S.width;
  const float aspect = (float)S.width / (float)S.height;
  float dc[3] = {(2.f * (col + 0.5f) / S.width - 1.f) * S.tan_half_fovy * aspect, (1.f - 2.f * (row + 0.5f) / S.height) * S.tan_half_fovy, -1.f};
  float inv = rsqrtf(dc[0] * dc[0] + dc[1] * dc[1] + dc[2] * dc[2]);
  dc[0] *= inv; dc[1] *= inv; dc[2] *= inv;
  float d[3];
  r_matvec(d, c_mat, dc);
  float tbest = 1e30f, nbest[3] = {0, 0, 1};
  int vbest = -1;
  for (int i = 0; i < S.nvis; i++) {
    float rel[3] = {c_pos[0] - v_pos[i][0], c_pos[1] - v_pos[i][1], c_pos[2] - v_pos[i][2]}, ol[3], dl[3], nl[3];
    r_matTvec(ol, v_mat[i], rel);
    r_matTvec(dl, v_mat[i], d);
    if (r_hit(S.vis[i].type, S.vis[i].size, ol, dl, tbest, nl)) { vbest = i; r_matvec(nbest, v_mat[i], nl); }
  }
  float c[3] = {0, 0, 0};
  if (vbest >= 0) {
    float ndl = -(nbest[0] * d[0] + nbest[1] * d[1] + nbest[2] * d[2]);
    float shade = RENDER_AMBIENT + RENDER_DIFFUSE * fabsf(ndl);  // two-sided
    for (int k = 0; k < 3; k++) c[k] = fminf(v_rgb[vbest][k], 1.f) * shade;
  }
  unsigned char* out = rgb + ((size_t)e * S.width * S.height + pix) * 3;
  for (int k = 0; k < 3; k++) out[k] = (unsigned char)(255.f * fminf(c[k], 1.f) + 0.5f);
}

// PointMass pose kernel: one thread per env.  base = world pose of every model geom at q = 0; geoms of the point-mass body
// translate with (qx, qy).
struct PmPoseTable { int ngeom; int moves[MJS_MAXGEOM]; float base[MJS_MAXGEOM][12]; };
static __global__ void pm_pose_kernel(PmPoseTable T, const double* sd, int N, float* gpose, float* aux) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= N) return;
  const float qx = (float)sd[0 * (size_t)N + e], qy = (float)sd[1 * (size_t)N + e];  // PM_QX, PM_QY
  float* gp = gpose + (size_t)e * MJS_MAXGEOM * 12;
  for (int g = 0; g < T.ngeom; g++) {
    for (int k = 0; k < 12; k++) gp[g * 12 + k] = T.base[g][k];
    if (T.moves[g]) { gp[g * 12] += qx; gp[g * 12 + 1] += qy; }
  }
  float* a = aux + (size_t)e * 8;
  a[0] = (float)sd[9 * (size_t)N + e]; a[1] = (float)sd[10 * (size_t)N + e]; a[2] = 0;  // PM_TX, PM_TY
  a[3] = a[4] = a[5] = a[6] = a[7] = 0;
}
