// mjs_oracle.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Double-precision, single-env, readable CPU restatement of the computation the
// reference performs per control step behind
//   /root/reference/mujoco_sim/environments/dmc2gym.py:133-155
// i.e. dm_control.composer's hook loop around MuJoCo's mj_step2/mj_step1
//  plus the task hooks in
//   /root/reference/mujoco_sim/environments/tasks/{point_reach,robot_reach,
//   robot_planar_push,robot_push_button}.py.
//
// PARITY UNPINNED: the arithmetic of the reference lives in the third-party
// `mujoco` C library (unpinned: /root/reference/setup.py:10-19), which is not
// installed here or on the GPU box (probed), and the reference's tests hold no
// golden vectors for this path.  The engine stages below follow
// MuJoCo's documented computation pipeline from recall [UPSTREAM-RECALL]; each
// function names the MuJoCo stage it restates.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../include/mjs_model.h"
#include "../include/mjs_convex.h"

namespace {

constexpr double MINVAL = 1e-15;
constexpr double MINIMP = 0.0001, MAXIMP = 0.9999;
constexpr double MINMU = 1e-5;
constexpr double CYL_FLAT_TILT = 1e-9;

// ------------------------------------------------------------------ small math
inline double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline void cross3(double* r, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
inline void set3(double* r, const double* a) { r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; }
inline void add3(double* r, const double* a, const double* b) { r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; }
inline void sub3(double* r, const double* a, const double* b) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
inline void addscl3(double* r, const double* a, double s) { r[0] += a[0] * s; r[1] += a[1] * s; r[2] += a[2] * s; }
inline double norm3(const double* a) { return std::sqrt(dot3(a, a)); }
inline double normalize3(double* a) {
  double n = norm3(a);
  if (n < MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; return n; }
  a[0] /= n; a[1] /= n; a[2] /= n;
  return n;
}
inline void mulquat(double* r, const double* a, const double* b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
inline void normquat(double* q) {
  double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  for (int i = 0; i < 4; i++) q[i] /= n;
}
inline void quat2mat(double* m, const double* q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
inline void mulmatvec3(double* r, const double* m, const double* v) {
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  double y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  double z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
inline void mulmatTvec3(double* r, const double* m, const double* v) {
  double x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2];
  double y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2];
  double z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
inline void axisangle2quat(double* q, const double* axis, double ang) {
  double s = std::sin(ang * 0.5);
  q[0] = std::cos(ang * 0.5); q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}

// Spatial inertia about the world origin: mass, mass*com, rotational inertia
// about the origin (xx,yy,zz,xy,xz,yz).  Additive, like MuJoCo's cinert/crb.
struct SpI {
  double m, mh[3], I[6];
  void zero() { std::memset(this, 0, sizeof(*this)); }
  void add(const SpI& o) {
    m += o.m;
    for (int i = 0; i < 3; i++) mh[i] += o.mh[i];
    for (int i = 0; i < 6; i++) I[i] += o.I[i];
  }
  // f = I * v ; v = [ang; lin] motion vector, f = [torque; force]
  void mul(double* f, const double* v) const {
    const double* w = v; const double* l = v + 3;
    double c[3];
    cross3(c, w, mh);                       // w x (m h)
    f[3] = m * l[0] + c[0]; f[4] = m * l[1] + c[1]; f[5] = m * l[2] + c[2];
    double t[3];
    cross3(t, mh, l);                       // (m h) x v
    f[0] = I[0] * w[0] + I[3] * w[1] + I[4] * w[2] + t[0];
    f[1] = I[3] * w[0] + I[1] * w[1] + I[5] * w[2] + t[1];
    f[2] = I[4] * w[0] + I[5] * w[1] + I[2] * w[2] + t[2];
  }
};
inline void crossMotion(double* r, const double* vel, const double* v) {
  double a[3], b[3], c[3];
  cross3(a, vel, v);
  cross3(b, vel, v + 3);
  cross3(c, vel + 3, v);
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2];
  r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
inline void crossForce(double* r, const double* vel, const double* f) {
  double a[3], b[3], c[3];
  cross3(a, vel, f);
  cross3(b, vel + 3, f + 3);
  cross3(c, vel, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2];
  r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}

struct Contact {
  double dist, pos[3], frame[9], friction[5], solref[2], solimp[5], includemargin;
  int dim, g1, g2, exclude, efc_address;
};

enum { ROW_EQ = 0, ROW_LIMIT = 1, ROW_CONTACT = 2 };

struct Data {
  // kinematics
  double xpos[MJS_MAXBODY * 3], xquat[MJS_MAXBODY * 4], xmat[MJS_MAXBODY * 9];
  double xipos[MJS_MAXBODY * 3], ximat[MJS_MAXBODY * 9];
  double xanchor[MJS_MAXJNT * 3], xaxis[MJS_MAXJNT * 3];
  double geom_xpos[MJS_MAXGEOM * 3], geom_xmat[MJS_MAXGEOM * 9];
  double site_xpos[MJS_MAXSITE * 3], site_xmat[MJS_MAXSITE * 9];
  // dynamics
  double cdof[MJS_MAXV * 6];
  double M[MJS_MAXV * MJS_MAXV];
  double qfrc_bias[MJS_MAXV], qfrc_passive[MJS_MAXV], qfrc_actuator[MJS_MAXV];
  double qfrc_smooth[MJS_MAXV], qacc_smooth[MJS_MAXV], qacc[MJS_MAXV], qfrc_constraint[MJS_MAXV];
  double actuator_force[MJS_MAXU];
  // collision + constraints
  int ncon;
  Contact con[MJS_MAXCON];
  int nefc;
  int efc_type[MJS_MAXROW], efc_id[MJS_MAXROW];
  double efc_J[MJS_MAXROW * MJS_MAXV];
  double efc_pos[MJS_MAXROW], efc_margin[MJS_MAXROW], efc_diagApprox[MJS_MAXROW];
  double efc_R[MJS_MAXROW], efc_D[MJS_MAXROW], efc_aref[MJS_MAXROW], efc_force[MJS_MAXROW];
  double efc_solref[MJS_MAXROW * 2], efc_solimp[MJS_MAXROW * 5];
  double touch;
  int solver_niter;
  int status;
  long flops;  // rough multiply-add counter for the roofline figure (x2 = flops)
};

// ------------------------------------------------------------------ kinematics
// Restates mj_kinematics (+ mocap bodies) [UPSTREAM-RECALL].
void kinematics(const MjsModel* m, const MjsEnvState* s, Data* d) {
  d->xpos[0] = d->xpos[1] = d->xpos[2] = 0;
  d->xquat[0] = 1; d->xquat[1] = d->xquat[2] = d->xquat[3] = 0;
  quat2mat(d->xmat, d->xquat);
  set3(d->xipos, d->xpos);
  quat2mat(d->ximat, d->xquat);
  for (int b = 1; b < m->nbody; b++) {
    double* xp = d->xpos + 3 * b; double* xq = d->xquat + 4 * b;
    int p = m->body_parent[b];
    if (m->body_mocapid[b] >= 0) {
      set3(xp, s->mocap_pos + 3 * m->body_mocapid[b]);
      xq[0] = 1; xq[1] = xq[2] = xq[3] = 0;  // mocap_quat is never written by the reference
    } else {
      const double* bpos = (b == m->ovr_body) ? s->ovr_pos : m->body_pos + 3 * b;
      double v[3];
      mulmatvec3(v, d->xmat + 9 * p, bpos);
      add3(xp, d->xpos + 3 * p, v);
      mulquat(xq, d->xquat + 4 * p, m->body_quat + 4 * b);
      for (int k = 0; k < m->body_jntnum[b]; k++) {
        int j = m->body_jntadr[b] + k;
        int qa = m->jnt_qposadr[j];
        if (m->jnt_type[j] == MJS_JNT_FREE) {
          set3(xp, s->qpos + qa);
          for (int i = 0; i < 4; i++) xq[i] = s->qpos[qa + 3 + i];
          normquat(xq);
          set3(d->xanchor + 3 * j, xp);
          d->xaxis[3 * j] = 0; d->xaxis[3 * j + 1] = 0; d->xaxis[3 * j + 2] = 1;
          continue;
        }
        double rot[9], anchor[3], axis[3];
        quat2mat(rot, xq);
        mulmatvec3(anchor, rot, m->jnt_pos + 3 * j);
        add3(anchor, anchor, xp);
        mulmatvec3(axis, rot, m->jnt_axis + 3 * j);
        set3(d->xanchor + 3 * j, anchor);
        set3(d->xaxis + 3 * j, axis);
        double q = s->qpos[qa] - m->qpos0[qa];
        if (m->jnt_type[j] == MJS_JNT_SLIDE) {
          addscl3(xp, axis, q);
        } else {  // hinge
          double ql[4], qn[4], vec[3];
          axisangle2quat(ql, m->jnt_axis + 3 * j, q);
          mulquat(qn, xq, ql);
          for (int i = 0; i < 4; i++) xq[i] = qn[i];
          quat2mat(rot, xq);
          mulmatvec3(vec, rot, m->jnt_pos + 3 * j);
          sub3(xp, anchor, vec);
        }
      }
    }
    normquat(xq);
    quat2mat(d->xmat + 9 * b, xq);
    double v[3], q2[4];
    mulmatvec3(v, d->xmat + 9 * b, m->body_ipos + 3 * b);
    add3(d->xipos + 3 * b, xp, v);
    mulquat(q2, xq, m->body_iquat + 4 * b);
    quat2mat(d->ximat + 9 * b, q2);
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_body[g];
    double v[3], q2[4];
    mulmatvec3(v, d->xmat + 9 * b, m->geom_pos + 3 * g);
    add3(d->geom_xpos + 3 * g, d->xpos + 3 * b, v);
    mulquat(q2, d->xquat + 4 * b, m->geom_quat + 4 * g);
    quat2mat(d->geom_xmat + 9 * g, q2);
  }
  for (int i = 0; i < m->nsite; i++) {
    int b = m->site_body[i];
    double v[3], q2[4];
    mulmatvec3(v, d->xmat + 9 * b, m->site_pos + 3 * i);
    add3(d->site_xpos + 3 * i, d->xpos + 3 * b, v);
    mulquat(q2, d->xquat + 4 * b, m->site_quat + 4 * i);
    quat2mat(d->site_xmat + 9 * i, q2);
  }
}

// Motion axes of every dof about the world origin (mj_comPos's cdof, with the
// reference point moved from the subtree COM to the origin) [UPSTREAM-RECALL].
void com_pos(const MjsModel* m, Data* d) {
  for (int j = 0; j < m->njnt; j++) {
    int b = m->jnt_body[j], da = m->jnt_dofadr[j];
    if (m->jnt_type[j] == MJS_JNT_FREE) {
      for (int k = 0; k < 3; k++) {
        double* c = d->cdof + 6 * (da + k);
        c[0] = c[1] = c[2] = 0; c[3] = c[4] = c[5] = 0; c[3 + k] = 1;
        double* r = d->cdof + 6 * (da + 3 + k);
        const double* R = d->xmat + 9 * b;
        double ax[3] = {R[k], R[3 + k], R[6 + k]};
        set3(r, ax);
        cross3(r + 3, d->xpos + 3 * b, ax);
      }
    } else if (m->jnt_type[j] == MJS_JNT_SLIDE) {
      double* c = d->cdof + 6 * da;
      c[0] = c[1] = c[2] = 0;
      set3(c + 3, d->xaxis + 3 * j);
    } else {
      double* c = d->cdof + 6 * da;
      set3(c, d->xaxis + 3 * j);
      cross3(c + 3, d->xanchor + 3 * j, d->xaxis + 3 * j);
    }
  }
}

void body_spatial_inertia(const MjsModel* m, const Data* d, int b, SpI* out) {
  out->zero();
  double mass = m->body_mass[b];
  if (mass <= 0) return;
  const double* R = d->ximat + 9 * b; const double* I = m->body_inertia + 3 * b; const double* h = d->xipos + 3 * b;
  double Ic[9];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) Ic[3 * r + c] = R[3 * r] * I[0] * R[3 * c] + R[3 * r + 1] * I[1] * R[3 * c + 1] + R[3 * r + 2] * I[2] * R[3 * c + 2];
  double hh = dot3(h, h);
  out->m = mass;
  for (int i = 0; i < 3; i++) out->mh[i] = mass * h[i];
  out->I[0] = Ic[0] + mass * (hh - h[0] * h[0]);
  out->I[1] = Ic[4] + mass * (hh - h[1] * h[1]);
  out->I[2] = Ic[8] + mass * (hh - h[2] * h[2]);
  out->I[3] = Ic[1] - mass * h[0] * h[1];
  out->I[4] = Ic[2] - mass * h[0] * h[2];
  out->I[5] = Ic[5] - mass * h[1] * h[2];
}

// Restates mj_crb: composite rigid body inertias -> dense joint-space M.
void crb(const MjsModel* m, Data* d) {
  int nv = m->nv;
  SpI crbI[MJS_MAXBODY];
  for (int b = 0; b < m->nbody; b++) body_spatial_inertia(m, d, b, &crbI[b]);
  for (int b = m->nbody - 1; b > 0; b--) crbI[m->body_parent[b]].add(crbI[b]);
  std::memset(d->M, 0, sizeof(double) * nv * nv);
  for (int i = 0; i < nv; i++) {
    double f[6];
    crbI[m->dof_body[i]].mul(f, d->cdof + 6 * i);
    d->M[i * nv + i] = m->dof_armature[i];
    for (int j = i; j >= 0; j = m->dof_parent[j]) {
      const double* c = d->cdof + 6 * j;
      double v = c[0] * f[0] + c[1] * f[1] + c[2] * f[2] + c[3] * f[3] + c[4] * f[4] + c[5] * f[5];
      d->M[i * nv + j] += v;
      if (j != i) d->M[j * nv + i] += v;
      d->flops += 6;
    }
  }
}

// translational / rotational Jacobian of a world point attached to body b
void jac(const MjsModel* m, const Data* d, int b, const double* point, double* jacp, double* jacr) {
  int nv = m->nv;
  if (jacp) std::memset(jacp, 0, sizeof(double) * 3 * nv);
  if (jacr) std::memset(jacr, 0, sizeof(double) * 3 * nv);
  while (b > 0 && m->body_dofnum[b] == 0) b = m->body_parent[b];
  if (b <= 0) return;
  int i = m->body_dofadr[b] + m->body_dofnum[b] - 1;
  for (; i >= 0; i = m->dof_parent[i]) {
    const double* c = d->cdof + 6 * i;
    if (jacr) { jacr[i] = c[0]; jacr[nv + i] = c[1]; jacr[2 * nv + i] = c[2]; }
    if (jacp) {
      double t[3];
      cross3(t, c, point);
      jacp[i] = c[3] + t[0]; jacp[nv + i] = c[4] + t[1]; jacp[2 * nv + i] = c[5] + t[2];
    }
  }
}

// Restates mj_comVel + mj_rne(flg_acc=0): Coriolis/centrifugal + gravity.
void rne_bias(const MjsModel* m, const MjsEnvState* s, Data* d) {
  int nb = m->nbody, nv = m->nv;
  double cvel[MJS_MAXBODY * 6], cacc[MJS_MAXBODY * 6], cfrc[MJS_MAXBODY * 6], cdof_dot[MJS_MAXV * 6];
  std::memset(cvel, 0, sizeof(cvel));
  std::memset(cacc, 0, sizeof(cacc));
  cacc[3] = -m->gravity[0]; cacc[4] = -m->gravity[1]; cacc[5] = -m->gravity[2];
  for (int b = 1; b < nb; b++) {
    int p = m->body_parent[b];
    double v[6];
    std::memcpy(v, cvel + 6 * p, sizeof(v));
    for (int k = 0; k < m->body_jntnum[b]; k++) {
      int j = m->body_jntadr[b] + k, da = m->jnt_dofadr[j];
      if (m->jnt_type[j] == MJS_JNT_FREE) {
        for (int i = 0; i < 3; i++) {
          std::memset(cdof_dot + 6 * (da + i), 0, 6 * sizeof(double));
          for (int c = 0; c < 6; c++) v[c] += d->cdof[6 * (da + i) + c] * s->qvel[da + i];
        }
        for (int i = 3; i < 6; i++) crossMotion(cdof_dot + 6 * (da + i), v, d->cdof + 6 * (da + i));
        for (int i = 3; i < 6; i++)
          for (int c = 0; c < 6; c++) v[c] += d->cdof[6 * (da + i) + c] * s->qvel[da + i];
      } else {
        crossMotion(cdof_dot + 6 * da, v, d->cdof + 6 * da);
        for (int c = 0; c < 6; c++) v[c] += d->cdof[6 * da + c] * s->qvel[da];
      }
    }
    std::memcpy(cvel + 6 * b, v, sizeof(v));
    // acceleration with qacc = 0
    double a[6];
    std::memcpy(a, cacc + 6 * p, sizeof(a));
    for (int i = 0; i < m->body_dofnum[b]; i++) {
      int da = m->body_dofadr[b] + i;
      for (int c = 0; c < 6; c++) a[c] += cdof_dot[6 * da + c] * s->qvel[da];
    }
    std::memcpy(cacc + 6 * b, a, sizeof(a));
    SpI I;
    body_spatial_inertia(m, d, b, &I);
    double f[6], iv[6], g[6];
    I.mul(f, a);
    I.mul(iv, v);
    crossForce(g, v, iv);
    for (int c = 0; c < 6; c++) cfrc[6 * b + c] = f[c] + g[c];
    d->flops += 120;
  }
  std::memset(cfrc, 0, 6 * sizeof(double));
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parent[b];
    for (int c = 0; c < 6; c++) cfrc[6 * p + c] += cfrc[6 * b + c];
  }
  for (int i = 0; i < nv; i++) {
    const double* c = d->cdof + 6 * i; const double* f = cfrc + 6 * m->dof_body[i];
    d->qfrc_bias[i] = c[0] * f[0] + c[1] * f[1] + c[2] * f[2] + c[3] * f[3] + c[4] * f[4] + c[5] * f[5];
  }
}

// Restates mj_passive: joint damping + gravity compensation (robot.py:81-82).
void passive(const MjsModel* m, const MjsEnvState* s, Data* d) {
  int nv = m->nv;
  for (int i = 0; i < nv; i++) d->qfrc_passive[i] = -m->dof_damping[i] * s->qvel[i];
  double jp[3 * MJS_MAXV];
  for (int b = 1; b < m->nbody; b++) {
    double gc = m->body_gravcomp[b];
    if (gc == 0 || m->body_mass[b] <= 0) continue;
    jac(m, d, b, d->xipos + 3 * b, jp, nullptr);
    double f[3] = {-m->gravity[0] * m->body_mass[b] * gc, -m->gravity[1] * m->body_mass[b] * gc, -m->gravity[2] * m->body_mass[b] * gc};
    for (int i = 0; i < nv; i++) d->qfrc_passive[i] += jp[i] * f[0] + jp[nv + i] * f[1] + jp[2 * nv + i] * f[2];
    d->flops += 3 * nv;
  }
}

// Restates mj_fwdActuation for affine joint actuators.
void actuation(const MjsModel* m, const MjsEnvState* s, Data* d, bool disabled) {
  for (int i = 0; i < m->nv; i++) d->qfrc_actuator[i] = 0;
  if (disabled) { for (int u = 0; u < m->nu; u++) d->actuator_force[u] = 0; return; }
  for (int u = 0; u < m->nu; u++) {
    int dof = m->act_dof[u];
    int qa = m->jnt_qposadr[m->dof_jnt[dof]];
    double ctrl = s->ctrl[u];
    if (m->act_ctrllimited[u]) ctrl = std::min(std::max(ctrl, m->act_ctrlrange[2 * u]), m->act_ctrlrange[2 * u + 1]);
    double len = m->act_gear[u] * s->qpos[qa], vel = m->act_gear[u] * s->qvel[dof];
    double f = m->act_gain[u] * ctrl + m->act_bias[3 * u] + m->act_bias[3 * u + 1] * len + m->act_bias[3 * u + 2] * vel;
    if (m->act_forcelimited[u]) f = std::min(std::max(f, m->act_forcerange[2 * u]), m->act_forcerange[2 * u + 1]);
    d->actuator_force[u] = f;
    d->qfrc_actuator[dof] += m->act_gear[u] * f;
  }
}

// dense Cholesky solve  A x = b  (A spd, n<=32)
bool chol_factor(double* A, int n) {
  for (int j = 0; j < n; j++) {
    double s = A[j * n + j];
    for (int k = 0; k < j; k++) s -= A[j * n + k] * A[j * n + k];
    if (s <= 0) return false;
    double l = std::sqrt(s);
    A[j * n + j] = l;
    for (int i = j + 1; i < n; i++) {
      double t = A[i * n + j];
      for (int k = 0; k < j; k++) t -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = t / l;
    }
  }
  return true;
}
void chol_solve(const double* L, int n, double* x) {
  for (int i = 0; i < n; i++) {
    double t = x[i];
    for (int k = 0; k < i; k++) t -= L[i * n + k] * x[k];
    x[i] = t / L[i * n + i];
  }
  for (int i = n - 1; i >= 0; i--) {
    double t = x[i];
    for (int k = i + 1; k < n; k++) t -= L[k * n + i] * x[k];
    x[i] = t / L[i * n + i];
  }
}

// =================================================================== collision
// Narrowphase routines restate MuJoCo's analytic primitives
// (engine_collision_primitive.c) [UPSTREAM-RECALL]; normal points geom1->geom2,
// pos is the midpoint between the surfaces.  Each returns the number of
// contacts written (dist, pos, frame[0:3] = normal, frame[3:6] = optional y hint).
struct RawCon { double dist, pos[3], normal[3], yhint[3]; };

int plane_sphere(const double* p1, const double* m1, const double* p2, double r, double margin, RawCon* c) {
  double n[3] = {m1[2], m1[5], m1[8]}, v[3];
  sub3(v, p2, p1);
  double dist = dot3(v, n) - r;
  if (dist > margin) return 0;
  c->dist = dist;
  set3(c->normal, n);
  set3(c->pos, p2);
  addscl3(c->pos, n, -(r + dist * 0.5));
  c->yhint[0] = c->yhint[1] = c->yhint[2] = 0;
  return 1;
}

int plane_capsule(const double* p1, const double* m1, const double* p2, const double* m2, const double* size, double margin, RawCon* c) {
  double axis[3] = {m2[2], m2[5], m2[8]};
  int n = 0;
  for (int e = 0; e < 2; e++) {
    double end[3];
    set3(end, p2);
    addscl3(end, axis, (e == 0 ? 1.0 : -1.0) * size[1]);
    int k = plane_sphere(p1, m1, end, size[0], margin, c + n);
    if (k) { set3(c[n].yhint, axis); n++; }
  }
  return n;
}

int plane_cylinder(const double* p1, const double* m1, const double* p2, const double* m2, const double* size, double margin, RawCon* c) {
  double normal[3] = {m1[2], m1[5], m1[8]};
  double axis[3] = {m2[2], m2[5], m2[8]};
  double prjaxis = dot3(normal, axis);
  if (prjaxis > 0) { axis[0] = -axis[0]; axis[1] = -axis[1]; axis[2] = -axis[2]; prjaxis = -prjaxis; }
  double vec[3];
  sub3(vec, p2, p1);
  double dist0 = dot3(vec, normal);
  // remove component of -normal along axis
  vec[0] = axis[0] * prjaxis - normal[0]; vec[1] = axis[1] * prjaxis - normal[1]; vec[2] = axis[2] * prjaxis - normal[2];
  double len_sqr = dot3(vec, vec);
  // MuJoCo tests len_sqr against mjMINVAL^2 = 1e-30, i.e. a cylinder counts as "flat on the plane"
  // only below ~1e-15 rad of tilt.  A settled prop sits at 1e-16..1e-15 rad of rounding noise, so that
  // test is decided by noise and rotates the 3-point support triangle by an arbitrary angle, which
  // changes the response to the next impact at O(1).  Both this oracle and the CUDA path use
  // 1e-9 rad instead (DESIGN.md "deviations"): identical to MuJoCo for any tilt MuJoCo resolves.
  if (len_sqr >= CYL_FLAT_TILT * CYL_FLAT_TILT) {
    double s = size[0] / std::sqrt(len_sqr);
    vec[0] *= s; vec[1] *= s; vec[2] *= s;
  } else {
    vec[0] = m2[0] * size[0]; vec[1] = m2[3] * size[0]; vec[2] = m2[6] * size[0];
  }
  double prjvec = dot3(vec, normal);
  axis[0] *= size[1]; axis[1] *= size[1]; axis[2] *= size[1];
  prjaxis *= size[1];
  int cnt = 0;
  auto emit = [&](double dist, const double* a, double sa, const double* b, double sb, const double* e, double se) {
    RawCon* r = c + cnt++;
    r->dist = dist;
    set3(r->pos, p2);
    addscl3(r->pos, a, sa);
    addscl3(r->pos, b, sb);
    if (e) addscl3(r->pos, e, se);
    addscl3(r->pos, normal, -dist * 0.5);
    set3(r->normal, normal);
    r->yhint[0] = r->yhint[1] = r->yhint[2] = 0;
  };
  if (dist0 + prjaxis + prjvec <= margin) emit(dist0 + prjaxis + prjvec, vec, 1.0, axis, 1.0, nullptr, 0);
  else return 0;
  if (dist0 - prjaxis + prjvec <= margin) emit(dist0 - prjaxis + prjvec, vec, 1.0, axis, -1.0, nullptr, 0);
  double prjvec1 = -prjvec * 0.5;
  if (dist0 + prjaxis + prjvec1 <= margin) {
    double vec1[3];
    cross3(vec1, vec, axis);
    normalize3(vec1);
    double s = size[0] * std::sqrt(3.0) / 2;
    vec1[0] *= s; vec1[1] *= s; vec1[2] *= s;
    emit(dist0 + prjaxis + prjvec1, vec1, 1.0, axis, 1.0, vec, -0.5);
    emit(dist0 + prjaxis + prjvec1, vec1, -1.0, axis, 1.0, vec, -0.5);
  }
  return cnt;
}

int plane_box(const double* p1, const double* m1, const double* p2, const double* m2, const double* size, double margin, RawCon* c) {
  double normal[3] = {m1[2], m1[5], m1[8]}, dif[3];
  sub3(dif, p2, p1);
  double dist = dot3(dif, normal);
  int cnt = 0;
  for (int i = 0; i < 8; i++) {
    double vec[3] = {(i & 1 ? size[0] : -size[0]), (i & 2 ? size[1] : -size[1]), (i & 4 ? size[2] : -size[2])};
    double corner[3];
    mulmatvec3(corner, m2, vec);
    double ldist = dot3(normal, corner);
    if (dist + ldist > margin || ldist > 0) continue;
    RawCon* r = c + cnt++;
    r->dist = dist + ldist;
    add3(r->pos, corner, p2);
    addscl3(r->pos, normal, -r->dist * 0.5);
    set3(r->normal, normal);
    r->yhint[0] = r->yhint[1] = r->yhint[2] = 0;
    if (cnt >= 4) break;
  }
  return cnt;
}

int sphere_sphere_raw(const double* p1, double r1, const double* p2, double r2, double margin, RawCon* c) {
  double dif[3];
  sub3(dif, p2, p1);
  double len = norm3(dif);
  double dist = len - r1 - r2;
  if (dist > margin) return 0;
  if (len < MINVAL) { dif[0] = 1; dif[1] = 0; dif[2] = 0; }
  else { dif[0] /= len; dif[1] /= len; dif[2] /= len; }
  c->dist = dist;
  set3(c->normal, dif);
  set3(c->pos, p1);
  addscl3(c->pos, dif, r1 + 0.5 * dist);
  c->yhint[0] = c->yhint[1] = c->yhint[2] = 0;
  return 1;
}

int sphere_capsule(const double* p1, double r1, const double* p2, const double* m2, const double* size2, double margin, RawCon* c) {
  double axis[3] = {m2[2], m2[5], m2[8]}, vec[3];
  sub3(vec, p1, p2);
  double x = std::min(std::max(dot3(axis, vec), -size2[1]), size2[1]);
  double q[3];
  set3(q, p2);
  addscl3(q, axis, x);
  return sphere_sphere_raw(p1, r1, q, size2[0], margin, c);
}

int capsule_capsule(const double* p1, const double* m1, const double* s1, const double* p2, const double* m2, const double* s2, double margin, RawCon* c) {
  double a1[3] = {m1[2], m1[5], m1[8]}, a2[3] = {m2[2], m2[5], m2[8]}, dif[3];
  sub3(dif, p1, p2);
  double ma = dot3(a1, a1), mb = -dot3(a1, a2), mc = dot3(a2, a2);
  double u = -dot3(a1, dif), v = dot3(a2, dif);
  double det = ma * mc - mb * mb;
  if (std::fabs(det) >= MINVAL) {
    double x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
    if (x1 > s1[1]) { x1 = s1[1]; x2 = (v - mb * s1[1]) / mc; }
    else if (x1 < -s1[1]) { x1 = -s1[1]; x2 = (v + mb * s1[1]) / mc; }
    if (x2 > s2[1]) { x2 = s2[1]; x1 = (u - mb * s2[1]) / ma; }
    else if (x2 < -s2[1]) { x2 = -s2[1]; x1 = (u + mb * s2[1]) / ma; }
    x1 = std::min(std::max(x1, -s1[1]), s1[1]);
    x2 = std::min(std::max(x2, -s2[1]), s2[1]);
    double q1[3], q2[3];
    set3(q1, p1); addscl3(q1, a1, x1);
    set3(q2, p2); addscl3(q2, a2, x2);
    return sphere_sphere_raw(q1, s1[0], q2, s2[0], margin, c);
  }
  // parallel axes: two end-point tests (deduplicated by the caller's capacity)
  int n = 0;
  double q1[3], q2[3];
  double x1 = s1[1];
  double x2 = std::min(std::max((v - mb * x1) / mc, -s2[1]), s2[1]);
  set3(q1, p1); addscl3(q1, a1, x1); set3(q2, p2); addscl3(q2, a2, x2);
  n += sphere_sphere_raw(q1, s1[0], q2, s2[0], margin, c + n);
  x1 = -s1[1];
  x2 = std::min(std::max((v - mb * x1) / mc, -s2[1]), s2[1]);
  set3(q1, p1); addscl3(q1, a1, x1); set3(q2, p2); addscl3(q2, a2, x2);
  n += sphere_sphere_raw(q1, s1[0], q2, s2[0], margin, c + n);
  return n;
}

// Sound rejection before the general convex routine (same test as cyl_separated in arm_kernel.cuh): separating-axis test of
// two cylinders on both axes, their cross product and `extra`.
bool cyl_separated(const double* dif, const double* a1, double r1, double h1, const double* a2, double r2, double h2, const double* extra) {
  double cx[3];
  cross3(cx, a1, a2);
  bool sep = false;
  for (int k = 0; k < 4; k++) {
    const double* d = k == 0 ? a1 : k == 1 ? a2 : k == 2 ? cx : extra;
    double n2 = dot3(d, d);
    if (n2 > 1e-20) {
      double inv = 1.0 / std::sqrt(n2);
      double c1 = dot3(a1, d) * inv, c2 = dot3(a2, d) * inv;
      double e1 = h1 * std::fabs(c1) + r1 * std::sqrt(std::max(0.0, 1 - c1 * c1)), e2 = h2 * std::fabs(c2) + r2 * std::sqrt(std::max(0.0, 1 - c2 * c2));
      if (std::fabs(dot3(dif, d)) * inv > e1 + e2) sep = true;
    }
  }
  return sep;
}

// Cylinder x cylinder, SIDE-SIDE contact only (own analytic routine; MuJoCo
// routes this pair through its iterative convex solver).  One
// contact at the midpoint of the axial overlap; continuous in the tilt angle.
// Returns -1 when the cylinders overlap only through an end cap (unsupported).
int cylinder_cylinder(const double* p1, const double* m1, const double* s1, const double* p2, const double* m2, const double* s2, double margin, RawCon* c) {
  double a1[3] = {m1[2], m1[5], m1[8]}, a2[3] = {m2[2], m2[5], m2[8]}, dif[3];
  sub3(dif, p2, p1);
  double tc = dot3(a1, dif), ta = dot3(a1, a2) * s2[1];
  double lo = std::max(-s1[1], tc - std::fabs(ta)), hi = std::min(s1[1], tc + std::fabs(ta));
  if (lo > hi) {
    // no axial overlap: only a cap-cap / cap-rim configuration could touch
    double gap = lo - hi;
    double radial[3];
    set3(radial, dif);
    addscl3(radial, a1, -tc);
    (void)gap; (void)radial;
    // only a cap configuration can touch: the caller asks the general convex routine unless an axis separates the cylinders
    { bool sp = cyl_separated(dif, a1, s1[0], s1[1], a2, s2[0], s2[1], dif);
      return sp ? 0 : -1; }
  }
  double sm = 0.5 * (lo + hi);
  double q1[3], q2[3], w[3];
  set3(q1, p1); addscl3(q1, a1, sm);
  sub3(w, q1, p2);
  double u = std::min(std::max(dot3(a2, w), -s2[1]), s2[1]);
  set3(q2, p2); addscl3(q2, a2, u);
  // The side-side contact is only the right answer while the radial direction is the axis of least penetration.  When the
  // axial overlap is the smaller one (a prop resting on another prop's cap) the true contact is cap-cap / cap-rim: unsupported.
  double dq[3];
  sub3(dq, q2, q1);
  double radial_pen = s1[0] + s2[0] - norm3(dq);
  // extent of cylinder 2 along axis 1 (axis segment + rim), so that a prop lying on its side still overlaps axially
  double ca = dot3(a1, a2), ext = std::fabs(ca) * s2[1] + s2[0] * std::sqrt(std::max(0.0, 1.0 - ca * ca));
  double axial_pen = std::min(s1[1], tc + ext) - std::max(-s1[1], tc - ext);
  if (radial_pen > 0 && axial_pen < radial_pen) {
    bool sp = cyl_separated(dif, a1, s1[0], s1[1], a2, s2[0], s2[1], dq);
    return sp ? 0 : -1;
  }
  return sphere_sphere_raw(q1, s1[0], q2, s2[0], margin, c);
}

// General convex pair (mjc_Convex): MPR penetration of the two primitives, one contact.  Used for the pairs MuJoCo has no
// analytic routine for (cylinder x box, capsule x cylinder, capsule x box here) and for the cap configurations of
// cylinder x cylinder that the side-side routine above declines.  margin > 0 is not supported by this restatement (the
// scenes use margin 0): shapes that do not intersect give no contact.
int convex_pair(int t1, const double* p1, const double* m1, const double* s1, int t2, const double* p2, const double* m2, const double* s2, RawCon* c) {
  MjsShape<double> A, B;
  A.type = t1; B.type = t2;
  // coordinates relative to shape 1's centre (the CUDA kernel does the same: identical arithmetic on both sides)
  for (int k = 0; k < 3; k++) { A.size[k] = s1[k]; A.pos[k] = 0; B.size[k] = s2[k]; B.pos[k] = p2[k] - p1[k]; }
  for (int k = 0; k < 9; k++) { A.mat[k] = m1[k]; B.mat[k] = m2[k]; }
  double depth, dir[3], pos[3];
  if (!mjs_mpr_penetration<double>(A, B, &depth, dir, pos)) return 0;
  c->dist = -depth;
  set3(c->normal, dir);
  for (int k = 0; k < 3; k++) c->pos[k] = pos[k] + p1[k];
  c->yhint[0] = c->yhint[1] = c->yhint[2] = 0;
  return 1;
}

// Sphere x cylinder (restates mjc_SphereCylinder's three cases: side, cap, rim).
int sphere_cylinder(const double* p1, double r1, const double* p2, const double* m2, const double* s2, double margin, RawCon* c) {
  double axis[3] = {m2[2], m2[5], m2[8]}, vec[3];
  sub3(vec, p1, p2);
  double x = dot3(vec, axis);
  double perp[3];
  set3(perp, vec);
  addscl3(perp, axis, -x);
  double plen = norm3(perp);
  double R = s2[0], H = s2[1];
  bool inside_h = std::fabs(x) <= H, inside_r = plen <= R;
  if (inside_h && !inside_r) {  // side: like sphere vs capsule-side
    double q[3];
    set3(q, p2); addscl3(q, axis, x);
    return sphere_sphere_raw(p1, r1, q, R, margin, c);
  }
  double sgn = x >= 0 ? 1.0 : -1.0;
  if (inside_r) {  // cap (or centre inside the cylinder: push out through the nearer cap)
    double dist = std::fabs(x) - H - r1;
    if (dist > margin) return 0;
    c->dist = dist;
    c->normal[0] = -sgn * axis[0]; c->normal[1] = -sgn * axis[1]; c->normal[2] = -sgn * axis[2];
    set3(c->pos, p1);
    addscl3(c->pos, c->normal, r1 + 0.5 * dist);
    c->yhint[0] = c->yhint[1] = c->yhint[2] = 0;
    return 1;
  }
  // rim: closest point on the cap circle
  double q[3];
  set3(q, p2);
  addscl3(q, axis, sgn * H);
  addscl3(q, perp, R / plen);
  return sphere_sphere_raw(p1, r1, q, 0.0, margin, c);
}

// Sphere x box (restates mjc_SphereBox: clamp centre to the box).
int sphere_box(const double* p1, double r1, const double* p2, const double* m2, const double* s2, double margin, RawCon* c) {
  double dif[3], loc[3], clamped[3];
  sub3(dif, p1, p2);
  mulmatTvec3(loc, m2, dif);
  bool inside = true;
  for (int i = 0; i < 3; i++) {
    clamped[i] = std::min(std::max(loc[i], -s2[i]), s2[i]);
    if (clamped[i] != loc[i]) inside = false;
  }
  if (!inside) {
    double q[3];
    mulmatvec3(q, m2, clamped);
    add3(q, q, p2);
    return sphere_sphere_raw(p1, r1, q, 0.0, margin, c);
  }
  // centre inside the box: exit through the nearest face
  int best = 0; double bd = 1e30;
  for (int i = 0; i < 3; i++) { double dd = s2[i] - std::fabs(loc[i]); if (dd < bd) { bd = dd; best = i; } }
  double nl[3] = {0, 0, 0};
  nl[best] = loc[best] >= 0 ? -1.0 : 1.0;  // from sphere (geom1) towards box interior
  mulmatvec3(c->normal, m2, nl);
  c->dist = -bd - r1;
  set3(c->pos, p1);
  addscl3(c->pos, c->normal, r1 + 0.5 * c->dist);
  c->yhint[0] = c->yhint[1] = c->yhint[2] = 0;
  return 1;
}

// mju_makeFrame
void make_frame(double* f) {
  normalize3(f);
  if (norm3(f + 3) < 0.5) {
    f[3] = f[4] = f[5] = 0;
    if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1;
  }
  double t = dot3(f, f + 3);
  addscl3(f + 3, f, -t);
  normalize3(f + 3);
  cross3(f + 6, f, f + 3);
}

// Restates mj_collision for the static pair list + mj_contactParam mixing.
void collision(const MjsModel* m, Data* d) {
  d->ncon = 0;
  for (int p = 0; p < m->npair; p++) {
    int g1 = m->pair_geom1[p], g2 = m->pair_geom2[p];
    int t1 = m->geom_type[g1], t2 = m->geom_type[g2];
    double margin = std::max(m->geom_margin[g1], m->geom_margin[g2]);
    double gap = std::max(m->geom_gap[g1], m->geom_gap[g2]);
    const double *p1 = d->geom_xpos + 3 * g1, *p2 = d->geom_xpos + 3 * g2;
    const double *m1 = d->geom_xmat + 9 * g1, *m2 = d->geom_xmat + 9 * g2;
    const double *s1 = m->geom_size + 3 * g1, *s2 = m->geom_size + 3 * g2;
    // broadphase: bounding spheres (plane: signed distance)
    if (t1 == MJS_GEOM_PLANE) {
      double n[3] = {m1[2], m1[5], m1[8]}, v[3];
      sub3(v, p2, p1);
      if (dot3(v, n) - m->geom_rbound[g2] > margin) continue;
    } else {
      double v[3];
      sub3(v, p2, p1);
      if (norm3(v) > m->geom_rbound[g1] + m->geom_rbound[g2] + margin) continue;
    }
    RawCon rc[4];
    int n = 0;
    if (t1 == MJS_GEOM_PLANE && t2 == MJS_GEOM_SPHERE) n = plane_sphere(p1, m1, p2, s2[0], margin, rc);
    else if (t1 == MJS_GEOM_PLANE && t2 == MJS_GEOM_CAPSULE) n = plane_capsule(p1, m1, p2, m2, s2, margin, rc);
    else if (t1 == MJS_GEOM_PLANE && t2 == MJS_GEOM_CYLINDER) n = plane_cylinder(p1, m1, p2, m2, s2, margin, rc);
    else if (t1 == MJS_GEOM_PLANE && t2 == MJS_GEOM_BOX) n = plane_box(p1, m1, p2, m2, s2, margin, rc);
    else if (t1 == MJS_GEOM_SPHERE && t2 == MJS_GEOM_SPHERE) n = sphere_sphere_raw(p1, s1[0], p2, s2[0], margin, rc);
    else if (t1 == MJS_GEOM_SPHERE && t2 == MJS_GEOM_CAPSULE) n = sphere_capsule(p1, s1[0], p2, m2, s2, margin, rc);
    else if (t1 == MJS_GEOM_SPHERE && t2 == MJS_GEOM_CYLINDER) n = sphere_cylinder(p1, s1[0], p2, m2, s2, margin, rc);
    else if (t1 == MJS_GEOM_SPHERE && t2 == MJS_GEOM_BOX) n = sphere_box(p1, s1[0], p2, m2, s2, margin, rc);
    else if (t1 == MJS_GEOM_CAPSULE && t2 == MJS_GEOM_CAPSULE) n = capsule_capsule(p1, m1, s1, p2, m2, s2, margin, rc);
    else if (t1 == MJS_GEOM_CYLINDER && t2 == MJS_GEOM_CYLINDER) {
      n = cylinder_cylinder(p1, m1, s1, p2, m2, s2, margin, rc);
      if (n < 0) n = convex_pair(t1, p1, m1, s1, t2, p2, m2, s2, rc);  // cap-cap / cap-rim / cap-side
    }
    else if ((t1 == MJS_GEOM_CAPSULE || t1 == MJS_GEOM_CYLINDER) && (t2 == MJS_GEOM_CYLINDER || t2 == MJS_GEOM_BOX)) n = convex_pair(t1, p1, m1, s1, t2, p2, m2, s2, rc);
    else { d->status |= MJS_STATUS_UNSUPPORTED_PAIR; continue; }
    d->flops += 60;
    if (n < 0) { d->status |= MJS_STATUS_UNSUPPORTED_PAIR; continue; }
    for (int k = 0; k < n; k++) {
      if (d->ncon >= MJS_MAXCON) { d->status |= MJS_STATUS_CONTACT_OVERFLOW; break; }
      Contact* c = d->con + d->ncon++;
      c->dist = rc[k].dist;
      set3(c->pos, rc[k].pos);
      set3(c->frame, rc[k].normal);
      set3(c->frame + 3, rc[k].yhint);
      make_frame(c->frame);
      c->g1 = g1; c->g2 = g2;
      c->includemargin = margin - gap;
      c->exclude = (c->dist >= c->includemargin);
      // mj_contactParam
      double fr[3];
      if (m->geom_priority[g1] != m->geom_priority[g2]) {
        int g = m->geom_priority[g1] > m->geom_priority[g2] ? g1 : g2;
        c->dim = m->geom_condim[g];
        for (int i = 0; i < 2; i++) c->solref[i] = m->geom_solref[2 * g + i];
        for (int i = 0; i < 5; i++) c->solimp[i] = m->geom_solimp[5 * g + i];
        for (int i = 0; i < 3; i++) fr[i] = m->geom_friction[3 * g + i];
      } else {
        c->dim = std::max(m->geom_condim[g1], m->geom_condim[g2]);
        double x1 = m->geom_solmix[g1], x2 = m->geom_solmix[g2], mix;
        if (x1 >= MINVAL && x2 >= MINVAL) mix = x1 / (x1 + x2);
        else if (x1 < MINVAL && x2 < MINVAL) mix = 0.5;
        else mix = x1 < MINVAL ? 0.0 : 1.0;
        const double *r1 = m->geom_solref + 2 * g1, *r2 = m->geom_solref + 2 * g2;
        if (r1[0] > 0 && r2[0] > 0) for (int i = 0; i < 2; i++) c->solref[i] = mix * r1[i] + (1 - mix) * r2[i];
        else for (int i = 0; i < 2; i++) c->solref[i] = std::min(r1[i], r2[i]);
        for (int i = 0; i < 5; i++) c->solimp[i] = mix * m->geom_solimp[5 * g1 + i] + (1 - mix) * m->geom_solimp[5 * g2 + i];
        for (int i = 0; i < 3; i++) fr[i] = std::max(m->geom_friction[3 * g1 + i], m->geom_friction[3 * g2 + i]);
      }
      c->friction[0] = std::max(MINMU, fr[0]); c->friction[1] = std::max(MINMU, fr[0]);
      c->friction[2] = std::max(MINMU, fr[1]);
      c->friction[3] = std::max(MINMU, fr[2]); c->friction[4] = std::max(MINMU, fr[2]);
    }
  }
}

// ================================================================== constraints
// getsolparam / getimpedance / KBI of engine_core_constraint.c [UPSTREAM-RECALL]
double impedance(const double* si, double pos, double margin) {
  double d0 = std::min(std::max(si[0], MINIMP), MAXIMP), dw = std::min(std::max(si[1], MINIMP), MAXIMP);
  double width = std::max(si[2], MINVAL), mid = std::min(std::max(si[3], MINIMP), MAXIMP), power = std::max(si[4], 1.0);
  if (d0 == dw || width <= MINVAL) return 0.5 * (d0 + dw);
  double x = std::fabs(pos - margin) / width;
  if (x >= 1) return dw;
  if (x <= 0) return d0;
  double y;
  if (power == 1) y = x;
  else if (x <= mid) y = std::pow(x, power) / std::pow(mid, power - 1);
  else y = 1 - std::pow(1 - x, power) / std::pow(1 - mid, power - 1);
  return d0 + y * (dw - d0);
}

int add_row(Data* d, int nv, int type, int id, const double* J, double pos, double margin, double diagApprox, const double* solref, const double* solimp) {
  if (d->nefc >= MJS_MAXROW) { d->status |= MJS_STATUS_CONTACT_OVERFLOW; return -1; }
  int r = d->nefc++;
  d->efc_type[r] = type; d->efc_id[r] = id;
  std::memcpy(d->efc_J + r * nv, J, sizeof(double) * nv);
  d->efc_pos[r] = pos; d->efc_margin[r] = margin; d->efc_diagApprox[r] = diagApprox;
  d->efc_solref[2 * r] = solref[0]; d->efc_solref[2 * r + 1] = solref[1];
  for (int i = 0; i < 5; i++) d->efc_solimp[5 * r + i] = solimp[i];
  return r;
}

// Restates mj_makeConstraint (equality weld, joint limits, pyramidal contacts),
// mj_diagApprox and mj_makeImpedance [UPSTREAM-RECALL].
void make_constraint(const MjsModel* m, const MjsEnvState* s, Data* d) {
  int nv = m->nv;
  d->nefc = 0;
  double jp1[3 * MJS_MAXV], jp2[3 * MJS_MAXV], jr1[3 * MJS_MAXV], jr2[3 * MJS_MAXV], J[MJS_MAXV];
  // ---- equality: weld (body semantics; residual = body1 anchor - body2 anchor)
  for (int e = 0; e < m->neq; e++) {
    int b1 = m->eq_body1[e], b2 = m->eq_body2[e];
    const double* data = m->eq_data + 11 * e;
    double pos1[3], pos2[3], cpos[3];
    mulmatvec3(pos1, d->xmat + 9 * b1, data + 3); add3(pos1, pos1, d->xpos + 3 * b1);
    mulmatvec3(pos2, d->xmat + 9 * b2, data + 0); add3(pos2, pos2, d->xpos + 3 * b2);
    sub3(cpos, pos1, pos2);
    jac(m, d, b1, pos1, jp1, jr1);
    jac(m, d, b2, pos2, jp2, jr2);
    double tran = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
    double rot = m->body_invweight0[2 * b1 + 1] + m->body_invweight0[2 * b2 + 1];
    for (int k = 0; k < 3; k++) {
      for (int i = 0; i < nv; i++) J[i] = jp1[k * nv + i] - jp2[k * nv + i];
      add_row(d, nv, ROW_EQ, e, J, cpos[k], 0, tran, m->eq_solref + 2 * e, m->eq_solimp + 5 * e);
    }
    if (m->eq_rot_rows[e]) {
      // orientation residual: torquescale * vec( neg(q2) * (q1 * relquat) )
      double qa[4], qb[4], q2n[4] = {d->xquat[4 * b2], -d->xquat[4 * b2 + 1], -d->xquat[4 * b2 + 2], -d->xquat[4 * b2 + 3]};
      mulquat(qa, d->xquat + 4 * b1, data + 6);
      mulquat(qb, q2n, qa);
      double ts = data[10];
      // MuJoCo's quaternion-difference Jacobian: 0.5 * vec( neg(q2) * (0,w) * (q1*relquat) )
      double Jr[3 * MJS_MAXV];
      for (int i = 0; i < nv; i++) {
        double w4[4] = {0, jr1[i] - jr2[i], jr1[nv + i] - jr2[nv + i], jr1[2 * nv + i] - jr2[2 * nv + i]}, t4[4], r4[4];
        mulquat(t4, w4, qa);
        mulquat(r4, q2n, t4);
        for (int k = 0; k < 3; k++) Jr[k * nv + i] = 0.5 * ts * r4[1 + k];
      }
      for (int k = 0; k < 3; k++)
        add_row(d, nv, ROW_EQ, e, Jr + k * nv, ts * qb[1 + k], 0, rot, m->eq_solref + 2 * e, m->eq_solimp + 5 * e);
    }
  }
  // ---- joint limits
  for (int j = 0; j < m->njnt; j++) {
    if (!m->jnt_limited[j]) continue;
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    for (int side = -1; side <= 1; side += 2) {
      double dist = side * (m->jnt_range[2 * j + (side + 1) / 2] - s->qpos[qa]);
      if (dist < m->jnt_margin[j]) {
        std::memset(J, 0, sizeof(double) * nv);
        J[da] = -side;
        add_row(d, nv, ROW_LIMIT, j, J, dist, m->jnt_margin[j], m->dof_invweight0[da], m->jnt_solref + 2 * j, m->jnt_solimp + 5 * j);
      }
    }
  }
  // ---- contacts (pyramidal cone)
  for (int ci = 0; ci < d->ncon; ci++) {
    Contact* c = d->con + ci;
    c->efc_address = -1;
    if (c->exclude) continue;
    int b1 = m->geom_body[c->g1], b2 = m->geom_body[c->g2];
    jac(m, d, b1, c->pos, jp1, jr1);
    jac(m, d, b2, c->pos, jp2, jr2);
    // contact-frame Jacobian of the relative motion (body2 - body1)
    double Jc[6 * MJS_MAXV];
    for (int k = 0; k < 3; k++)
      for (int i = 0; i < nv; i++) {
        double tp = 0, tr = 0;
        for (int a = 0; a < 3; a++) {
          tp += c->frame[3 * k + a] * (jp2[a * nv + i] - jp1[a * nv + i]);
          tr += c->frame[3 * k + a] * (jr2[a * nv + i] - jr1[a * nv + i]);
        }
        Jc[k * nv + i] = tp; Jc[(3 + k) * nv + i] = tr;
      }
    d->flops += 18 * nv;
    double tran = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
    double rot = m->body_invweight0[2 * b1 + 1] + m->body_invweight0[2 * b2 + 1];
    c->efc_address = d->nefc;
    if (c->dim == 1) {
      add_row(d, nv, ROW_CONTACT, ci, Jc, c->dist, c->includemargin, tran, c->solref, c->solimp);
      continue;
    }
    for (int k = 1; k < c->dim; k++) {
      double fri = c->friction[k - 1];
      double da = tran + fri * fri * (k - 1 < 2 ? tran : rot);
      for (int sgn = 1; sgn >= -1; sgn -= 2) {
        for (int i = 0; i < nv; i++) J[i] = Jc[i] + sgn * fri * Jc[k * nv + i];
        add_row(d, nv, ROW_CONTACT, ci, J, c->dist, c->includemargin, da, c->solref, c->solimp);
      }
    }
  }
  // ---- mj_makeImpedance: R, D
  double imp[MJS_MAXROW];
  for (int r = 0; r < d->nefc; r++) {
    imp[r] = impedance(d->efc_solimp + 5 * r, d->efc_pos[r], d->efc_margin[r]);
    d->efc_R[r] = std::max(MINVAL, (1 - imp[r]) * d->efc_diagApprox[r] / imp[r]);
  }
  for (int ci = 0; ci < d->ncon; ci++) {
    Contact* c = d->con + ci;
    if (c->efc_address < 0 || c->dim == 1) continue;
    int r0 = c->efc_address;
    double mu = c->friction[0] / std::sqrt(m->impratio);
    double Rpy = 2 * mu * mu * d->efc_R[r0];
    for (int k = 0; k < 2 * (c->dim - 1); k++) d->efc_R[r0 + k] = Rpy;
  }
  for (int r = 0; r < d->nefc; r++) d->efc_D[r] = 1.0 / d->efc_R[r];
  // stash impedance in efc_force temporarily for reference_constraint
  for (int r = 0; r < d->nefc; r++) d->efc_force[r] = imp[r];
}

// Restates mj_referenceConstraint: aref = -b*vel - k*imp*(pos-margin).
void reference_constraint(const MjsModel* m, const MjsEnvState* s, Data* d) {
  int nv = m->nv;
  for (int r = 0; r < d->nefc; r++) {
    double imp = d->efc_force[r];
    double vel = 0;
    for (int i = 0; i < nv; i++) vel += d->efc_J[r * nv + i] * s->qvel[i];
    double sr0 = d->efc_solref[2 * r], sr1 = d->efc_solref[2 * r + 1];
    double dmax = std::min(std::max(d->efc_solimp[5 * r + 1], MINIMP), MAXIMP);
    double k, b;
    if (sr0 > 0) {
      double tc = std::max(sr0, 2 * m->timestep);  // refsafe
      double dr = std::max(MINVAL, sr1);
      k = 1 / std::max(MINVAL, dmax * dmax * tc * tc * dr * dr);
      b = 2 / std::max(MINVAL, dmax * tc);
    } else {
      k = -sr0 / std::max(MINVAL, dmax * dmax);
      b = -sr1 / std::max(MINVAL, dmax);
    }
    d->efc_aref[r] = -b * vel - k * imp * (d->efc_pos[r] - d->efc_margin[r]);
    d->flops += nv;
  }
}

// ---------------------------------------------------------------- Newton solver
// Primal Newton on  1/2 (a-a0)'M(a-a0) + sum_i s_i(J_i a - aref_i)  with an EXACT
// line search (the cost is piecewise quadratic for equality/limit/pyramidal rows).
// MuJoCo's mj_solNewton minimises the same strictly convex cost; iteration paths
// differ, the minimiser does not.  The oracle iterates to a
// much tighter tolerance than MuJoCo's 1e-8 so it is the exact minimiser.
double solver_cost(const MjsModel* m, const Data* d, const double* a, double* jar_out) {
  int nv = m->nv;
  double cost = 0;
  for (int i = 0; i < nv; i++) {
    double Ma = 0;
    for (int j = 0; j < nv; j++) Ma += d->M[i * nv + j] * (a[j] - d->qacc_smooth[j]);
    cost += 0.5 * Ma * (a[i] - d->qacc_smooth[i]);
  }
  for (int r = 0; r < d->nefc; r++) {
    double jar = -d->efc_aref[r];
    for (int i = 0; i < nv; i++) jar += d->efc_J[r * nv + i] * a[i];
    if (jar_out) jar_out[r] = jar;
    if (d->efc_type[r] == ROW_EQ || jar < 0) cost += 0.5 * d->efc_D[r] * jar * jar;
  }
  return cost;
}

void solve_constraints(const MjsModel* m, const MjsEnvState* s, Data* d, double tol_scale) {
  int nv = m->nv, nefc = d->nefc;
  d->solver_niter = 0;
  if (nefc == 0) {
    for (int i = 0; i < nv; i++) { d->qacc[i] = d->qacc_smooth[i]; d->qfrc_constraint[i] = 0; }
    return;
  }
  double a[MJS_MAXV], jar[MJS_MAXROW], grad[MJS_MAXV], dir[MJS_MAXV], H[MJS_MAXV * MJS_MAXV], jd[MJS_MAXROW], Md[MJS_MAXV];
  // warm start: the cheaper of qacc_warmstart and qacc_smooth
  double cw = solver_cost(m, d, s->qacc_warm, nullptr), cs = solver_cost(m, d, d->qacc_smooth, nullptr);
  for (int i = 0; i < nv; i++) a[i] = (cw < cs) ? s->qacc_warm[i] : d->qacc_smooth[i];
  double scale = 1.0 / (m->meaninertia * std::max(1, nv));
  double tol = m->tolerance * tol_scale;
  double cost = solver_cost(m, d, a, jar);
  int maxiter = m->solver_iterations;
  for (int it = 0; it < maxiter; it++) {
    // gradient and Hessian
    for (int i = 0; i < nv; i++) {
      double g = 0;
      for (int j = 0; j < nv; j++) g += d->M[i * nv + j] * (a[j] - d->qacc_smooth[j]);
      grad[i] = g;
    }
    std::memcpy(H, d->M, sizeof(double) * nv * nv);
    for (int r = 0; r < nefc; r++) {
      bool active = d->efc_type[r] == ROW_EQ || jar[r] < 0;
      if (!active) continue;
      const double* Jr = d->efc_J + r * nv;
      double Dj = d->efc_D[r];
      for (int i = 0; i < nv; i++) {
        if (Jr[i] == 0) continue;
        grad[i] += Jr[i] * Dj * jar[r];
        for (int j = 0; j < nv; j++) H[i * nv + j] += Dj * Jr[i] * Jr[j];
      }
      d->flops += (long)nv * nv / 2;
    }
    double gn = 0;
    for (int i = 0; i < nv; i++) gn += grad[i] * grad[i];
    if (std::sqrt(gn) * scale < tol) break;
    if (!chol_factor(H, nv)) { d->status |= MJS_STATUS_NAN; break; }
    for (int i = 0; i < nv; i++) dir[i] = -grad[i];
    chol_solve(H, nv, dir);
    d->flops += (long)nv * nv * nv / 6 + (long)nv * nv;
    // exact line search along dir
    double s0 = 0, c0 = 0;  // phi'(0), phi''(0+)
    for (int i = 0; i < nv; i++) {
      double t = 0;
      for (int j = 0; j < nv; j++) t += d->M[i * nv + j] * dir[j];
      Md[i] = t;
      c0 += dir[i] * t;
      double am = 0;
      for (int j = 0; j < nv; j++) am += d->M[i * nv + j] * (a[j] - d->qacc_smooth[j]);
      s0 += dir[i] * am;
    }
    std::vector<std::pair<double, int>> bp;
    for (int r = 0; r < nefc; r++) {
      double t = 0;
      for (int i = 0; i < nv; i++) t += d->efc_J[r * nv + i] * dir[i];
      jd[r] = t;
      bool eq = d->efc_type[r] == ROW_EQ;
      bool act0 = eq || jar[r] < 0 || (jar[r] == 0 && t < 0);
      if (act0) { s0 += d->efc_D[r] * jar[r] * t; c0 += d->efc_D[r] * t * t; }
      if (!eq && t != 0) {
        double al = -jar[r] / t;
        if (al > 0) bp.push_back({al, r});
      }
    }
    std::sort(bp.begin(), bp.end());
    double alpha = 0, sl = s0, cu = c0, best = -1;
    for (size_t k = 0; k <= bp.size(); k++) {
      double next = k < bp.size() ? bp[k].first : 1e300;
      double astar = cu > 0 ? alpha - sl / cu : 1e300;
      if (sl >= 0) { best = alpha; break; }
      if (astar <= next) { best = astar; break; }
      sl += cu * (next - alpha);
      alpha = next;
      int r = bp[k].second;
      double q = d->efc_D[r] * jd[r] * jd[r];
      // row switches: active->inactive if jd>0 (jar rising through 0), else inactive->active
      if (jd[r] > 0) cu -= q; else cu += q;
    }
    if (best < 0) best = alpha;
    for (int i = 0; i < nv; i++) a[i] += best * dir[i];
    double newcost = solver_cost(m, d, a, jar);
    d->solver_niter = it + 1;
    d->flops += (long)nefc * nv * 2;
    double improvement = cost - newcost;
    cost = newcost;
    if (improvement * scale < tol) break;
  }
  for (int i = 0; i < nv; i++) { d->qacc[i] = a[i]; d->qfrc_constraint[i] = 0; }
  for (int r = 0; r < nefc; r++) {
    bool active = d->efc_type[r] == ROW_EQ || jar[r] < 0;
    d->efc_force[r] = active ? -d->efc_D[r] * jar[r] : 0;
    for (int i = 0; i < nv; i++) d->qfrc_constraint[i] += d->efc_J[r * nv + i] * d->efc_force[r];
  }
}

// Restates mj_sensorAcc for the touch sensor: sum of the normal
// forces (pyramidal: sum of the edge forces) of contacts that involve the
// site's body and whose point lies inside the cylindrical site volume.
void touch_sensor(const MjsModel* m, Data* d) {
  d->touch = 0;
  if (m->touch_site < 0) return;
  int sb = m->site_body[m->touch_site];
  const double* sp = d->site_xpos + 3 * m->touch_site; const double* sm = d->site_xmat + 9 * m->touch_site;
  for (int ci = 0; ci < d->ncon; ci++) {
    const Contact* c = d->con + ci;
    if (c->efc_address < 0) continue;
    if (m->geom_body[c->g1] != sb && m->geom_body[c->g2] != sb) continue;
    double f = 0;
    int nr = c->dim == 1 ? 1 : 2 * (c->dim - 1);
    for (int k = 0; k < nr; k++) f += d->efc_force[c->efc_address + k];
    if (f <= 0) continue;
    double dif[3], loc[3];
    sub3(dif, c->pos, sp);
    mulmatTvec3(loc, sm, dif);
    if (loc[0] * loc[0] + loc[1] * loc[1] <= m->touch_size[0] * m->touch_size[0] && std::fabs(loc[2]) <= m->touch_size[1]) d->touch += f;
  }
}

// ------------------------------------------------------------------ step halves
// mj_step1 = mj_fwdPosition + mj_fwdVelocity (SURVEY.md 3.3 "legacy_step").
void step1(const MjsModel* m, const MjsEnvState* s, Data* d) {
  kinematics(m, s, d);
  com_pos(m, d);
  crb(m, d);
  collision(m, d);
  make_constraint(m, s, d);
  passive(m, s, d);
  reference_constraint(m, s, d);
  rne_bias(m, s, d);
}

// mj_step2 = fwdActuation + fwdAcceleration + fwdConstraint + sensorAcc + integrate.
void step2(const MjsModel* m, MjsEnvState* s, Data* d, bool actuation_disabled, bool integrate, double tol_scale) {
  int nv = m->nv;
  actuation(m, s, d, actuation_disabled);
  double L[MJS_MAXV * MJS_MAXV];
  std::memcpy(L, d->M, sizeof(double) * nv * nv);
  if (nv && !chol_factor(L, nv)) d->status |= MJS_STATUS_NAN;
  for (int i = 0; i < nv; i++) {
    d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i] + d->qfrc_actuator[i];
    d->qacc_smooth[i] = d->qfrc_smooth[i];
  }
  if (nv) chol_solve(L, nv, d->qacc_smooth);
  d->flops += (long)nv * nv * nv / 6 + (long)nv * nv;
  solve_constraints(m, s, d, tol_scale);
  for (int i = 0; i < nv; i++) s->qacc_warm[i] = d->qacc[i];  // mj_fwdConstraint: save warmstart
  touch_sensor(m, d);
  if (!integrate) return;
  double h = m->timestep;
  double qacc[MJS_MAXV];
  std::memcpy(qacc, d->qacc, sizeof(double) * nv);
  // velocity-derivative damping terms handled implicitly
  double dv[MJS_MAXV];
  bool any = false;
  for (int i = 0; i < nv; i++) { dv[i] = m->dof_damping[i]; if (dv[i] > 0) any = true; }
  if (m->integrator == MJS_INT_IMPLICITFAST) {
    // mjd_actuator_vel: affine bias velocity gain; skipped when the force is clamped
    for (int u = 0; u < m->nu; u++) {
      if (m->act_forcelimited[u] && (d->actuator_force[u] <= m->act_forcerange[2 * u] || d->actuator_force[u] >= m->act_forcerange[2 * u + 1])) continue;
      dv[m->act_dof[u]] += -m->act_bias[3 * u + 2] * m->act_gear[u] * m->act_gear[u];
      any = true;
    }
  }
  if (any) {
    // (M + h*diag(dv)) qacc' = qfrc_smooth + qfrc_constraint   (mj_EulerSkip / mj_implicit)
    std::memcpy(L, d->M, sizeof(double) * nv * nv);
    for (int i = 0; i < nv; i++) { L[i * nv + i] += h * dv[i]; qacc[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i]; }
    if (!chol_factor(L, nv)) d->status |= MJS_STATUS_NAN;
    chol_solve(L, nv, qacc);
    d->flops += (long)nv * nv * nv / 6 + (long)nv * nv;
  }
  // mj_advance: qvel += h*qacc; qpos integrated with the NEW velocity
  for (int i = 0; i < nv; i++) s->qvel[i] += h * qacc[i];
  for (int j = 0; j < m->njnt; j++) {
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    if (m->jnt_type[j] == MJS_JNT_FREE) {
      for (int k = 0; k < 3; k++) s->qpos[qa + k] += h * s->qvel[da + k];
      double w[3] = {s->qvel[da + 3], s->qvel[da + 4], s->qvel[da + 5]};
      double ang = norm3(w) * h;
      double ax[3] = {w[0], w[1], w[2]};
      normalize3(ax);
      double dq[4], qn[4];
      axisangle2quat(dq, ax, ang);
      mulquat(qn, s->qpos + qa + 3, dq);
      normquat(qn);
      for (int k = 0; k < 4; k++) s->qpos[qa + 3 + k] = qn[k];
    } else {
      s->qpos[qa] += h * s->qvel[da];
    }
  }
  s->time += h;
  for (int i = 0; i < nv; i++) if (!std::isfinite(s->qpos[i]) || !std::isfinite(s->qvel[i])) d->status |= MJS_STATUS_NAN;
}

// ================================================================ analytic UR IK
// Closed-form 6R inverse kinematics for the UR DH convention, restating what the
// reference obtains from `ur_analytic_ik.ur5e.inverse_kinematics_closest`
// (/root/reference/mujoco_sim/entities/robots/robot.py:31-38) [UPSTREAM-RECALL:
// package absent; "closest" = minimum squared joint distance to the seed after
// moving every joint to its nearest 2*pi-equivalent inside [-2pi, 2pi]].
void dh_T(double* T, double th, double dd, double a, double al) {
  double ct = std::cos(th), st = std::sin(th), ca = std::cos(al), sa = std::sin(al);
  double R[16] = {ct, -st * ca, st * sa, a * ct, st, ct * ca, -ct * sa, a * st, 0, sa, ca, dd, 0, 0, 0, 1};
  std::memcpy(T, R, sizeof(R));
}
void mul44(double* C, const double* A, const double* B) {
  double R[16];
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      double t = 0;
      for (int k = 0; k < 4; k++) t += A[4 * i + k] * B[4 * k + j];
      R[4 * i + j] = t;
    }
  std::memcpy(C, R, sizeof(R));
}
void inv44(double* B, const double* A) {  // rigid transform inverse
  double R[16] = {A[0], A[4], A[8], 0, A[1], A[5], A[9], 0, A[2], A[6], A[10], 0, 0, 0, 0, 1};
  for (int i = 0; i < 3; i++) R[4 * i + 3] = -(R[4 * i] * A[3] + R[4 * i + 1] * A[7] + R[4 * i + 2] * A[11]);
  std::memcpy(B, R, sizeof(R));
}
const double UR_ALPHA[6] = {M_PI / 2, 0, 0, M_PI / 2, -M_PI / 2, 0};

void ur_fk(const double* dh, const double* q, double* T06) {
  double dd[6] = {dh[0], 0, 0, dh[3], dh[4], dh[5]}, aa[6] = {0, dh[1], dh[2], 0, 0, 0};
  double T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, A[16];
  for (int i = 0; i < 6; i++) { dh_T(A, q[i], dd[i], aa[i], UR_ALPHA[i]); mul44(T, T, A); }
  std::memcpy(T06, T, sizeof(T));
}

// returns number of solutions (<=8) written to sols[8][6]
int ur_ik_all(const double* dh, const double* T06, double sols[8][6]) {
  double d1 = dh[0], a2 = dh[1], a3 = dh[2], d4 = dh[3], d6 = dh[5];
  int n = 0;
  double p05[3] = {T06[3] - d6 * T06[2], T06[7] - d6 * T06[6], T06[11] - d6 * T06[10]};
  double rxy = std::hypot(p05[0], p05[1]);
  if (rxy < std::fabs(d4)) return 0;
  double psi = std::atan2(p05[1], p05[0]), phi = std::acos(d4 / rxy);
  for (int s1 = 0; s1 < 2; s1++) {
    double th1 = psi + (s1 == 0 ? phi : -phi) + M_PI / 2;
    double T01[16], T10[16], T16[16];
    dh_T(T01, th1, d1, 0, M_PI / 2);
    inv44(T10, T01);
    mul44(T16, T10, T06);
    double c5 = T16[10];  // R16[2][2]
    if (std::fabs(c5) > 1) { if (std::fabs(c5) < 1 + 1e-9) c5 = c5 > 0 ? 1 : -1; else continue; }
    for (int s5 = 0; s5 < 2; s5++) {
      double th5 = (s5 == 0 ? 1 : -1) * std::acos(c5);
      double sn5 = std::sin(th5);
      double th6, th234;
      if (std::fabs(sn5) < 1e-9) {
        th6 = 0;
        // R16 = Rz(th234 -/+ th6): with th6 = 0
        th234 = std::atan2(T16[4], T16[0]) * (c5 > 0 ? 1 : 1);
        if (c5 < 0) th234 = std::atan2(-T16[4], -T16[0]);
      } else {
        th6 = std::atan2(-T16[9] / sn5, T16[8] / sn5);
        th234 = std::atan2(T16[6] / -sn5, T16[2] / -sn5);
      }
      double T45[16], T56[16], T46[16], T64[16], T14[16];
      dh_T(T45, th5, dh[4], 0, -M_PI / 2);
      dh_T(T56, th6, d6, 0, 0);
      mul44(T46, T45, T56);
      inv44(T64, T46);
      mul44(T14, T16, T64);
      double x = T14[3], y = T14[7];
      double c3 = (x * x + y * y - a2 * a2 - a3 * a3) / (2 * a2 * a3);
      if (std::fabs(c3) > 1) { if (std::fabs(c3) < 1 + 1e-9) c3 = c3 > 0 ? 1 : -1; else continue; }
      for (int s3 = 0; s3 < 2; s3++) {
        double th3 = (s3 == 0 ? 1 : -1) * std::acos(c3);
        double th2 = std::atan2(y, x) - std::atan2(a3 * std::sin(th3), a2 + a3 * std::cos(th3));
        double th4 = th234 - th2 - th3;
        double* q = sols[n++];
        q[0] = th1; q[1] = th2; q[2] = th3; q[3] = th4; q[4] = th5; q[5] = th6;
        for (int i = 0; i < 6; i++) {  // wrap to (-pi, pi]
          q[i] = std::fmod(q[i] + M_PI, 2 * M_PI);
          if (q[i] < 0) q[i] += 2 * M_PI;
          q[i] -= M_PI;
        }
      }
    }
  }
  return n;
}

bool ur_ik_closest(const double* dh, const double* T06, const double* seed, double* qout) {
  double sols[8][6];
  int n = ur_ik_all(dh, T06, sols);
  double best = 1e300;
  int bi = -1;
  for (int k = 0; k < n; k++) {
    double dist = 0;
    for (int i = 0; i < 6; i++) {
      double q = sols[k][i], bq = q, bd = std::fabs(q - seed[i]);
      for (int w = -1; w <= 1; w += 2) {
        double c = q + w * 2 * M_PI;
        if (c >= -2 * M_PI && c <= 2 * M_PI && std::fabs(c - seed[i]) < bd) { bd = std::fabs(c - seed[i]); bq = c; }
      }
      sols[k][i] = bq;
      dist += bd * bd;
    }
    if (dist < best) { best = dist; bi = k; }
  }
  if (bi < 0) return false;
  for (int i = 0; i < 6; i++) qout[i] = sols[bi][i];
  return true;
}

// top-down TCP position -> flange pose -> joints (robot.py:113-121,138-151,218-225)
bool tcp_to_joints(const MjsTask* t, const double* tcp, const double* seed, double* q) {
  // TOP_DOWN_QUATERNION (scalar-last [1,0,0,0]) = 180 deg about x: R = diag(1,-1,-1)
  double T[16] = {1, 0, 0, 0, 0, -1, 0, 0, 0, 0, -1, 0, 0, 0, 0, 1};
  T[3] = tcp[0] - (t->tcp_offset[0]);
  T[7] = tcp[1] + (t->tcp_offset[1]);
  T[11] = tcp[2] + (t->tcp_offset[2]);
  return ur_ik_closest(t->ik_dh, T, seed, q);
}

// ==================================================================== task layer
// The composer hook loop of SURVEY.md 3.3 for one env, one control step.
void tcp_position(const MjsModel* m, const MjsTask* t, const Data* d, double* tcp) {
  // robot.py:153-168 + :126-136: tcp = flange_pos + flange_R * tcp_in_flange
  const double* R = d->site_xmat + 9 * t->tcp_site;
  mulmatvec3(tcp, R, t->tcp_offset);
  add3(tcp, tcp, d->site_xpos + 3 * t->tcp_site);
}

void servoJ(const MjsTask* t, MjsEnvState* s, const double* qtarget) {
  // robot.py:227-259: two-waypoint trajectory (q_now, t) -> (q*, t + T)
  for (int i = 0; i < 6; i++) { s->traj_q0[i] = s->qpos[t->arm_qadr + i]; s->traj_q1[i] = qtarget[i]; }
  s->traj_t0 = s->time;
  s->traj_active = 1;
}

void servoL(const MjsTask* t, MjsEnvState* s, const double* tcp) {
  double q[6];
  if (!tcp_to_joints(t, tcp, s->qpos + t->arm_qadr, q)) {
    // the reference raises ValueError("IK failed") (robot.py:221-224); the batched
    // env flags the env and keeps the previous trajectory
    s->status |= MJS_STATUS_IK_FAIL;
    return;
  }
  servoJ(t, s, q);
}

void before_substep(const MjsTask* t, MjsEnvState* s) {
  if (t->kind == MJS_TASK_POINTMASS || !s->traj_active) return;
  // joint_trajectory.py:33-47 (time clipped to the waypoints), robot.py:261-272
  double t0 = s->traj_t0, t1 = s->traj_t0 + t->control_dt;
  double tt = std::min(std::max(s->time, t0), t1);
  for (int i = 0; i < 6; i++) s->ctrl[t->arm_qadr + i] = s->traj_q0[i] + (s->traj_q1[i] - s->traj_q0[i]) * (tt - t0) / (t1 - t0);
}

void switch_update(const MjsTask* t, MjsEnvState* s, double force) {
  // switch.py:51-60
  int was = s->sw_pressed;
  s->sw_pressed = (force >= t->touch_lo && force <= t->touch_hi) ? 1 : 0;
  if (s->sw_pressed && !was) s->sw_active = !s->sw_active;
  s->sw_count += s->sw_pressed;
}

void write_obs_reward(const MjsModel* m, const MjsTask* t, MjsEnvState* s, const Data* d, MjsStepOut* out) {
  std::memset(out, 0, sizeof(*out));
  bool accomplished = false, has_success = false, success = false;
  double reward = 0;
  if (t->kind == MJS_TASK_POINTMASS) {
    const double* xp = d->xpos + 3 * t->pm_body;
    out->obs[0] = xp[0]; out->obs[1] = xp[1]; out->obs[2] = s->target[0]; out->obs[3] = s->target[1];
    // point_reach.py:172-202
    switch (t->reward_type) {
      case 0: reward = s->dist < t->goal_threshold; break;
      case 1: reward = -s->dist; break;
      case 2: reward = -s->dist + 0.5; break;
      default: reward = s->prev_dist - s->dist; break;
    }
    accomplished = s->dist < t->goal_threshold;
    has_success = true; success = accomplished;
  } else {
    double tcp[3];
    tcp_position(m, t, d, tcp);
    if (t->kind == MJS_TASK_REACH) {
      for (int i = 0; i < 3; i++) { out->obs[i] = tcp[i]; out->obs[3 + i] = s->target[i]; }
      double df[3];
      sub3(df, tcp, s->target);
      double dist = norm3(df);
      reward = t->reward_type == 0 ? (dist < t->goal_threshold) : -dist;  // robot_reach.py:174-181
      accomplished = false;  // composer.Task default: never terminates
    } else if (t->kind == MJS_TASK_PUSH) {
      for (int i = 0; i < 3; i++) out->obs[i] = tcp[i];
      out->obs[3] = s->target[0]; out->obs[4] = s->target[1];
      double sum = 0, nearest = 1e300;
      int within = 0;
      for (int o = 0; o < t->n_obj; o++) {
        const double* xp = d->xpos + 3 * t->obj_body[o];
        out->obs[5 + 2 * o] = xp[0]; out->obs[6 + 2 * o] = xp[1];
        double dd = std::hypot(xp[0] - s->target[0], xp[1] - s->target[1]);
        sum += dd;
        if (dd < t->target_radius) within++;
        nearest = std::min(nearest, std::hypot(tcp[0] - xp[0], tcp[1] - xp[1]));
      }
      // robot_planar_push.py:203-220
      if (t->reward_type == 0) reward = within;
      else reward = (-sum / t->n_obj - t->nearest_coef * nearest) * 0.1;
      accomplished = within == t->n_obj;
    } else {  // BUTTON
      // robot_push_button.py:109-113: joint_configuration with ABS_JOINT_ACTION (7 actions), tcp_position with ABS_EEF_ACTION (4)
      int base = 6;
      if (t->action_dim == 7) for (int i = 0; i < 6; i++) out->obs[i] = s->qpos[t->arm_qadr + i];
      else { base = 3; for (int i = 0; i < 3; i++) out->obs[i] = tcp[i]; }
      // switch.py:86-87: button geom xpos + 0.5*size[1] added to all three coordinates
      // (the button geom is geom `touch geom` = the site's own position)
      const double* bp = d->site_xpos + 3 * m->touch_site;
      double quirk = 0.5 * (m->touch_size[0] / 1.01);
      for (int i = 0; i < 3; i++) out->obs[base + i] = bp[i] + quirk;
      out->obs[base + 3] = s->sw_active;
      double df[3];
      sub3(df, tcp, t->button_goal);
      bool goal = s->sw_active && norm3(df) < t->button_goal_radius;  // robot_push_button.py:205-210
      reward = goal ? 1.0 : 0.0;
      accomplished = goal;
      has_success = true; success = goal;
    }
  }
  bool terminate = accomplished;
  if ((t->kind == MJS_TASK_PUSH) && t->max_steps > 0 && s->episode_step >= t->max_steps) terminate = true;  // base.py:47-51
  bool timeup = t->time_limit > 0 && s->time >= t->time_limit;  // composer time_limit
  double discount = accomplished ? 0.0 : 1.0;
  if (t->kind == MJS_TASK_REACH) discount = 1.0;
  bool last = terminate || timeup;
  out->reward = reward;
  out->discount = discount;
  out->done = last;
  out->truncated = last && discount > 0;   // dmc2gym.py:144
  out->terminated = last && discount == 0; // dmc2gym.py:145
  out->success = has_success ? success : 0;
  s->ep_return += reward;
}

// touch-sensor force after every substep of the last control_step on this thread (test hook: margin of the 5 N / 200 N
// thresholds of switch.py:51-60, so flag comparisons can exclude envs that sit on a threshold)
static thread_local double g_touch_trace[64];
// uniform draw of the button-disturbance test of the current control step (robot_push_button.py:159-165 `random_state.rand()`),
// supplied by the caller (oracle_control_step_disturb); 1.0 = never
static thread_local double g_disturb_u = 1.0;

void control_step(const MjsModel* m, const MjsTask* t, MjsEnvState* s, const double* action, MjsStepOut* out, double tol_scale, long* flops) {
  static thread_local Data d;
  d.status = 0; d.flops = 0;
  // derived quantities of the CURRENT state (what the previous mj_step1 left behind;
  // for PointMass this still sees the old mocap_pos -- SURVEY.md 3.4-1)
  step1(m, s, &d);
  // ---- before_step
  if (t->kind == MJS_TASK_POINTMASS) {
    const double* xp = d.xpos + 3 * t->pm_body;  // point_reach.py:150-163
    for (int k = 0; k < 2; k++) s->mocap_pos[3 * t->pm_mocap + k] = std::min(std::max(xp[k] + action[k], t->arena_range[0]), t->arena_range[1]);
  } else if (t->kind == MJS_TASK_REACH) {
    servoL(t, s, action);
  } else if (t->kind == MJS_TASK_PUSH) {
    s->episode_step += 1;
    double tcp[3] = {action[0], action[1], t->tcp_z};
    servoL(t, s, tcp);
  } else if (t->kind == MJS_TASK_BUTTON) {
    if (t->action_dim == 7) servoJ(t, s, action);
    else servoL(t, s, action);
  }
  // ---- substeps (legacy_step: mj_step2 then mj_step1)
  for (int k = 0; k < t->n_sub; k++) {
    before_substep(t, s);
    step2(m, s, &d, false, true, tol_scale);
    step1(m, s, &d);
    if (k < 64) g_touch_trace[k] = d.touch;
    if (t->kind == MJS_TASK_BUTTON) switch_update(t, s, d.touch);
  }
  // ---- after_step
  if (t->kind == MJS_TASK_BUTTON && t->button_disturbances && s->sw_active && !s->sw_pressed && g_disturb_u < t->disturb_prob) s->sw_active = 0;
  if (t->kind == MJS_TASK_POINTMASS) {
    const double* xp = d.xpos + 3 * t->pm_body;  // point_reach.py:165-170
    s->prev_dist = s->dist;
    s->dist = std::hypot(xp[0] - s->target[0], xp[1] - s->target[1]);
  }
  s->status |= d.status;
  write_obs_reward(m, t, s, &d, out);
  // an unsupported geom pair or a contact overflow during this step cuts the episode (same rule as the CUDA step kernel)
  if (d.status & (MJS_STATUS_UNSUPPORTED_PAIR | MJS_STATUS_CONTACT_OVERFLOW)) { out->done = 1; out->truncated = 1; out->terminated = 0; }
  if (flops) *flops += d.flops;
}

// physics.reset_context exit: mj_forward with actuation disabled
void forward_after_reset(const MjsModel* m, MjsEnvState* s, Data* d) {
  for (int i = 0; i < m->nv; i++) s->qacc_warm[i] = 0;
  step1(m, s, d);
  step2(m, s, d, true, false, 1e-4);
}

}  // namespace

// ======================================================================== C API
extern "C" {

// one control step for n envs (AoS states), OpenMP-free; callers thread by range
void oracle_control_step(const MjsModel* m, const MjsTask* t, MjsEnvState* states, const double* actions, MjsStepOut* outs, int n, double tol_scale, long* flops) {
  for (int e = 0; e < n; e++) control_step(m, t, states + e, actions + (size_t)e * t->action_dim, outs + e, tol_scale, flops);
}

// MPR penetration of two primitives (include/mjs_convex.h) for the CPU tests: out = depth, dir[3], pos[3]; returns 1 on contact
int oracle_convex_pair(int t1, const double* size1, const double* pos1, const double* mat1, int t2, const double* size2, const double* pos2, const double* mat2, double* out) {
  MjsShape<double> A, B;
  A.type = t1; B.type = t2;
  for (int k = 0; k < 3; k++) { A.size[k] = size1[k]; A.pos[k] = pos1[k]; B.size[k] = size2[k]; B.pos[k] = pos2[k]; }
  for (int k = 0; k < 9; k++) { A.mat[k] = mat1[k]; B.mat[k] = mat2[k]; }
  return mjs_mpr_penetration<double>(A, B, out, out + 1, out + 4);
}

// world poses of every model geom at the given state (mj_kinematics): xpos[ngeom][3], xmat[ngeom][9].  Feeds the numpy
// restatement of the ray caster (oracle/render_ref.py).
void oracle_geom_poses(const MjsModel* m, const MjsEnvState* state, double* xpos, double* xmat) {
  static thread_local Data d;
  MjsEnvState s = *state;
  step1(m, &s, &d);
  std::memcpy(xpos, d.geom_xpos, sizeof(double) * 3 * m->ngeom);
  std::memcpy(xmat, d.geom_xmat, sizeof(double) * 9 * m->ngeom);
}

// oracle_control_step with the button-disturbance draw of every env supplied by the caller: u[n] in [0, 1)
void oracle_control_step_disturb(const MjsModel* m, const MjsTask* t, MjsEnvState* states, const double* actions, MjsStepOut* outs, int n, double tol_scale, const double* u) {
  for (int e = 0; e < n; e++) {
    g_disturb_u = u[e];
    control_step(m, t, states + e, actions + (size_t)e * t->action_dim, outs + e, tol_scale, nullptr);
  }
  g_disturb_u = 1.0;
}

// oracle_control_step that also returns the touch-sensor force after every substep: touch[n][n_sub]
void oracle_control_step_touch(const MjsModel* m, const MjsTask* t, MjsEnvState* states, const double* actions, MjsStepOut* outs, int n, double tol_scale, double* touch) {
  for (int e = 0; e < n; e++) {
    control_step(m, t, states + e, actions + (size_t)e * t->action_dim, outs + e, tol_scale, nullptr);
    for (int k = 0; k < t->n_sub && k < 64; k++) touch[(size_t)e * t->n_sub + k] = g_touch_trace[k];
  }
}

// finish a reset whose (qpos, qvel, mocap, ctrl, target...) were written by the caller:
// mj_forward; for Planar Push additionally the 150 settle steps (robot_planar_push.py:160-161)
void oracle_finish_reset(const MjsModel* m, const MjsTask* t, MjsEnvState* states, MjsStepOut* outs, int n, int settle_steps) {
  static thread_local Data d;
  for (int e = 0; e < n; e++) {
    MjsEnvState* s = states + e;
    d.status = 0; d.flops = 0;
    forward_after_reset(m, s, &d);
    for (int k = 0; k < settle_steps; k++) { step2(m, s, &d, false, true, 1e-4); step1(m, s, &d); }
    if (t->kind == MJS_TASK_BUTTON) switch_update(t, s, d.touch);  // Switch.initialize_episode
    s->status |= d.status;
    double keep = s->ep_return;
    write_obs_reward(m, t, s, &d, outs + e);
    s->ep_return = keep;
  }
}

// number of contacts at the given state (Planar Push rejection sampling: physics.data.ncon)
// continue the current control step by k more physics substeps (no before_step hook): lets tests
// split one control step at a substep boundary
void oracle_substeps(const MjsModel* m, const MjsTask* t, MjsEnvState* states, int n, int k, double tol_scale) {
  for (int e = 0; e < n; e++) {
    Data d;
    d.status = 0; d.flops = 0;
    MjsEnvState* s = states + e;
    step1(m, s, &d);
    for (int i = 0; i < k; i++) {
      before_substep(t, s);
      step2(m, s, &d, false, true, tol_scale);
      step1(m, s, &d);
      if (t->kind == MJS_TASK_BUTTON) switch_update(t, s, d.touch);
    }
    s->status |= d.status;
  }
}

// Noise-free scripted demonstrators on the current states: point_reach.py:227-240, robot_reach.py:222-253,
// robot_push_button.py:231-300.  ok[e] = 0 where the task has none or the IK of the joint action fails.
void oracle_demo_policy(const MjsModel* m, const MjsTask* t, const MjsEnvState* states, int n, double* actions, int* ok) {
  static thread_local Data d;
  for (int e = 0; e < n; e++) {
    const MjsEnvState* s = states + e;
    double* a = actions + (size_t)e * t->action_dim;
    for (int k = 0; k < t->action_dim; k++) a[k] = 0;
    ok[e] = 1;
    d.status = 0;
    kinematics(m, s, &d);
    if (t->kind == MJS_TASK_POINTMASS) {
      const double* xp = d.xpos + 3 * t->pm_body;
      double v[2] = {s->target[0] - xp[0], s->target[1] - xp[1]};
      double mx = std::max(std::fabs(v[0]), std::fabs(v[1]));
      double k = mx > 0 ? t->action_hi[0] / mx : 0.0;  // MAX_STEP_SIZE
      a[0] = v[0] * k; a[1] = v[1] * k;
      continue;
    }
    double tcp[3];
    tcp_position(m, t, &d, tcp);
    if (t->kind == MJS_TASK_REACH) {
      double df[3], mx = 0;
      for (int k = 0; k < 3; k++) { df[k] = s->target[k] - tcp[k]; mx = std::max(mx, std::fabs(df[k])); }
      if (mx > 0.5) for (int k = 0; k < 3; k++) df[k] = df[k] * 0.5 / mx * t->control_dt;
      for (int k = 0; k < 3; k++) a[k] = tcp[k] + df[k];
    } else if (t->kind == MJS_TASK_BUTTON) {
      const double* bp = d.site_xpos + 3 * m->touch_site;
      double quirk = 0.5 * (m->touch_size[0] / 1.01), sw[3], goal[3];
      for (int k = 0; k < 3; k++) sw[k] = bp[k] + quirk;
      double dxy = std::hypot(tcp[0] - sw[0], tcp[1] - sw[1]);
      int phase = s->sw_active ? 3 : ((tcp[2] > sw[2] && dxy < 0.01) ? 2 : 1);
      if (phase == 1) {
        if (tcp[2] < sw[2] + 0.02) { goal[0] = tcp[0]; goal[1] = tcp[1]; } else { goal[0] = sw[0]; goal[1] = sw[1]; }
        goal[2] = sw[2] + 0.05;
      } else if (phase == 2) {
        for (int k = 0; k < 3; k++) goal[k] = sw[k];
      } else {
        for (int k = 0; k < 3; k++) goal[k] = t->button_goal[k];
        if (dxy < 0.05) goal[2] = sw[2] + 0.1;
      }
      double df[3], mx = 0, lim = 0.5 * t->control_dt;
      for (int k = 0; k < 3; k++) { df[k] = goal[k] - tcp[k]; mx = std::max(mx, std::fabs(df[k])); }
      if (mx > lim) for (int k = 0; k < 3; k++) df[k] = df[k] * 0.5 / mx * t->control_dt;
      double tgt[3] = {tcp[0] + df[0], tcp[1] + df[1], tcp[2] + df[2]}, q[6];
      if (t->action_dim != 7) {  // ABS_EEF_ACTION: the Cartesian target + the closed gripper (robot_push_button.py:289-297)
        for (int k = 0; k < 3; k++) a[k] = tgt[k];
        a[3] = 0.0; ok[e] = 1;
        continue;
      }
      bool good = tcp_to_joints(t, tgt, s->qpos + t->arm_qadr, q);
      for (int i = 0; i < 6; i++) a[i] = good ? q[i] : s->qpos[t->arm_qadr + i];
      a[6] = 0.0;
      ok[e] = good ? 1 : 0;
    } else {
      ok[e] = 0;  // Planar Push: keyboard tele-operation in the reference
    }
  }
}

int oracle_ncon(const MjsModel* m, const MjsEnvState* s) {
  static thread_local Data d;
  d.status = 0;
  kinematics(m, s, &d);
  collision(m, &d);
  return d.ncon;
}

// stage-wise introspection for tests: runs step1 (+ optional smooth/solve) and dumps
void oracle_stage_dump(const MjsModel* m, MjsEnvState* s, double* xpos, double* xquat, double* site_xpos, double* site_xmat,
                       double* M, double* qfrc_bias, double* qfrc_passive, double* qacc_smooth, double* qacc,
                       int* ncon, double* con_dist, double* con_pos, double* con_frame, int* nefc, double* efc_J,
                       double* efc_aref, double* efc_D, double* efc_force, double* touch, int* niter) {
  static thread_local Data d;
  d.status = 0; d.flops = 0;
  MjsEnvState tmp = *s;
  step1(m, &tmp, &d);
  step2(m, &tmp, &d, false, false, 1e-4);
  std::memcpy(xpos, d.xpos, sizeof(double) * 3 * m->nbody);
  std::memcpy(xquat, d.xquat, sizeof(double) * 4 * m->nbody);
  std::memcpy(site_xpos, d.site_xpos, sizeof(double) * 3 * m->nsite);
  std::memcpy(site_xmat, d.site_xmat, sizeof(double) * 9 * m->nsite);
  std::memcpy(M, d.M, sizeof(double) * m->nv * m->nv);
  std::memcpy(qfrc_bias, d.qfrc_bias, sizeof(double) * m->nv);
  std::memcpy(qfrc_passive, d.qfrc_passive, sizeof(double) * m->nv);
  std::memcpy(qacc_smooth, d.qacc_smooth, sizeof(double) * m->nv);
  std::memcpy(qacc, d.qacc, sizeof(double) * m->nv);
  *ncon = d.ncon;
  for (int i = 0; i < d.ncon; i++) {
    con_dist[i] = d.con[i].dist;
    std::memcpy(con_pos + 3 * i, d.con[i].pos, 3 * sizeof(double));
    std::memcpy(con_frame + 9 * i, d.con[i].frame, 9 * sizeof(double));
  }
  *nefc = d.nefc;
  std::memcpy(efc_J, d.efc_J, sizeof(double) * d.nefc * m->nv);
  std::memcpy(efc_aref, d.efc_aref, sizeof(double) * d.nefc);
  std::memcpy(efc_D, d.efc_D, sizeof(double) * d.nefc);
  std::memcpy(efc_force, d.efc_force, sizeof(double) * d.nefc);
  *touch = d.touch;
  *niter = d.solver_niter;
}

void oracle_ur_fk(const double* dh, const double* q, double* T06) { ur_fk(dh, q, T06); }
int oracle_ur_ik_all(const double* dh, const double* T06, double* sols) { return ur_ik_all(dh, T06, (double(*)[6])sols); }
int oracle_ur_ik_closest(const double* dh, const double* T06, const double* seed, double* q) { return ur_ik_closest(dh, T06, seed, q) ? 1 : 0; }
int oracle_tcp_to_joints(const MjsTask* t, const double* tcp, const double* seed, double* q) { return tcp_to_joints(t, tcp, seed, q) ? 1 : 0; }
int oracle_sizeof_model() { return (int)sizeof(MjsModel); }
int oracle_sizeof_task() { return (int)sizeof(MjsTask); }
int oracle_sizeof_state() { return (int)sizeof(MjsEnvState); }
int oracle_sizeof_out() { return (int)sizeof(MjsStepOut); }

}  // extern "C"
