/* mjs_model.h -- flat, fixed-capacity model + task description shared by the
 * CUDA product path (mujoco_sim_b200/csrc) and the CPU oracle (oracle/).
 *
 * This is DATA only: it is what the reference gets from
 * `mjcf.Physics.from_mjcf_model(task.root_entity.mjcf_model)` (the compiled
 * mjModel behind /root/reference/mujoco_sim/__init__.py:19-23) restricted to
 * the fields the four composer tasks exercise, plus the task constants of
 * /root/reference/mujoco_sim/environments/tasks/ (all four task files).  The host-side model
 * flattener (mujoco_sim_b200/model/compiler.py) fills it.
 *
 * The layout is deliberately "one declaration per line, plain arrays" so the
 * Python side can mirror it into a ctypes.Structure by parsing this file
 * (mujoco_sim_b200/model/cstruct.py).  Keep that grammar:
 *     <int|double> name;            or
 *     <int|double> name[EXPR];      EXPR = product of MJS_* macros / integers
 */
#ifndef MJS_MODEL_H_
#define MJS_MODEL_H_

#define MJS_MAXBODY 16
#define MJS_MAXJNT 16
#define MJS_MAXQ 40
#define MJS_MAXV 32
#define MJS_MAXGEOM 24
#define MJS_MAXSITE 8
#define MJS_MAXEQ 2
#define MJS_MAXU 8
#define MJS_MAXPAIR 64
#define MJS_MAXMOCAP 2
#define MJS_MAXOBJ 5
#define MJS_MAXCON 24   /* contacts kept per env per substep            */
#define MJS_MAXROW 160  /* constraint rows per env per substep          */

/* joint types (values follow mjtJoint) */
#define MJS_JNT_FREE 0
#define MJS_JNT_SLIDE 2
#define MJS_JNT_HINGE 3
/* geom types (values follow mjtGeom) */
#define MJS_GEOM_PLANE 0
#define MJS_GEOM_SPHERE 2
#define MJS_GEOM_CAPSULE 3
#define MJS_GEOM_CYLINDER 5
#define MJS_GEOM_BOX 6
/* integrators */
#define MJS_INT_EULER 0
#define MJS_INT_IMPLICITFAST 1
/* task kinds */
#define MJS_TASK_NONE 0
#define MJS_TASK_POINTMASS 1
#define MJS_TASK_REACH 2
#define MJS_TASK_PUSH 3
#define MJS_TASK_BUTTON 4
/* per-env status bits written by the step kernels */
#define MJS_STATUS_NAN 1
#define MJS_STATUS_IK_FAIL 2
#define MJS_STATUS_CONTACT_OVERFLOW 4
#define MJS_STATUS_UNSUPPORTED_PAIR 8
/* flag bytes, out_flags[env*4 + k] */
#define MJS_FLAG_TERMINATED 0
#define MJS_FLAG_TRUNCATED 1
#define MJS_FLAG_SUCCESS 2
#define MJS_FLAG_DONE 3

typedef struct MjsModel {
  /* ---- sizes ---- */
  int nbody;
  int njnt;
  int nq;
  int nv;
  int ngeom;
  int nsite;
  int neq;
  int nu;
  int npair;
  int nmocap;
  /* ---- options (mjOption) ---- */
  int integrator;
  int solver_iterations;
  double timestep;
  double gravity[3];
  double impratio;
  double tolerance;
  double meaninertia;
  /* ---- bodies ---- */
  int body_parent[MJS_MAXBODY];
  int body_jntadr[MJS_MAXBODY];
  int body_jntnum[MJS_MAXBODY];
  int body_dofadr[MJS_MAXBODY];
  int body_dofnum[MJS_MAXBODY];
  int body_mocapid[MJS_MAXBODY];
  int body_weldid[MJS_MAXBODY];
  double body_pos[MJS_MAXBODY * 3];
  double body_quat[MJS_MAXBODY * 4];
  double body_ipos[MJS_MAXBODY * 3];
  double body_iquat[MJS_MAXBODY * 4];
  double body_mass[MJS_MAXBODY];
  double body_inertia[MJS_MAXBODY * 3];
  double body_gravcomp[MJS_MAXBODY];
  double body_invweight0[MJS_MAXBODY * 2];
  /* ---- joints ---- */
  int jnt_type[MJS_MAXJNT];
  int jnt_body[MJS_MAXJNT];
  int jnt_qposadr[MJS_MAXJNT];
  int jnt_dofadr[MJS_MAXJNT];
  int jnt_limited[MJS_MAXJNT];
  double jnt_pos[MJS_MAXJNT * 3];
  double jnt_axis[MJS_MAXJNT * 3];
  double jnt_range[MJS_MAXJNT * 2];
  double jnt_margin[MJS_MAXJNT];
  double jnt_solref[MJS_MAXJNT * 2];
  double jnt_solimp[MJS_MAXJNT * 5];
  double qpos0[MJS_MAXQ];
  /* ---- dofs ---- */
  int dof_body[MJS_MAXV];
  int dof_parent[MJS_MAXV];
  int dof_jnt[MJS_MAXV];
  double dof_armature[MJS_MAXV];
  double dof_damping[MJS_MAXV];
  double dof_invweight0[MJS_MAXV];
  /* ---- geoms ---- */
  int geom_type[MJS_MAXGEOM];
  int geom_body[MJS_MAXGEOM];
  int geom_condim[MJS_MAXGEOM];
  int geom_priority[MJS_MAXGEOM];
  double geom_pos[MJS_MAXGEOM * 3];
  double geom_quat[MJS_MAXGEOM * 4];
  double geom_size[MJS_MAXGEOM * 3];
  double geom_friction[MJS_MAXGEOM * 3];
  double geom_solref[MJS_MAXGEOM * 2];
  double geom_solimp[MJS_MAXGEOM * 5];
  double geom_solmix[MJS_MAXGEOM];
  double geom_margin[MJS_MAXGEOM];
  double geom_gap[MJS_MAXGEOM];
  double geom_rbound[MJS_MAXGEOM];
  /* ---- sites ---- */
  int site_body[MJS_MAXSITE];
  double site_pos[MJS_MAXSITE * 3];
  double site_quat[MJS_MAXSITE * 4];
  /* ---- equality constraints (weld only) ---- */
  int eq_body1[MJS_MAXEQ];
  int eq_body2[MJS_MAXEQ];
  int eq_rot_rows[MJS_MAXEQ]; /* 1: emit the 3 rotational weld rows too */
  double eq_data[MJS_MAXEQ * 11];
  double eq_solref[MJS_MAXEQ * 2];
  double eq_solimp[MJS_MAXEQ * 5];
  /* ---- actuators (general, joint transmission, affine bias) ---- */
  int act_dof[MJS_MAXU];
  int act_ctrllimited[MJS_MAXU];
  int act_forcelimited[MJS_MAXU];
  double act_gear[MJS_MAXU];
  double act_gain[MJS_MAXU];
  double act_bias[MJS_MAXU * 3];
  double act_ctrlrange[MJS_MAXU * 2];
  double act_forcerange[MJS_MAXU * 2];
  /* ---- static candidate collision pairs (after contype/parent/weld filter) ---- */
  int pair_geom1[MJS_MAXPAIR];
  int pair_geom2[MJS_MAXPAIR];
  /* ---- per-env overridable body position (Button Push: switch frame) ---- */
  int ovr_body; /* -1: none.  state.ovr_pos[3] replaces body_pos of this body */
  /* ---- touch sensor (Button Push) ---- */
  int touch_site;          /* -1: none */
  double touch_size[2];    /* cylinder site radius, half-height */
} MjsModel;

/* Task constants: the numbers in /root/reference/mujoco_sim/environments/tasks/ (all four task files) */
typedef struct MjsTask {
  int kind;
  int n_sub;               /* round(control_timestep / physics_timestep) */
  int reward_type;
  int max_steps;           /* task-side step limit (RobotTask, base.py:47-51); 0 = none */
  int n_obj;
  int action_dim;              /* PointMass 2, Reach 3, Planar Push 2; Button Push 7 = ABS_JOINT_ACTION (servoJ) or 4 = ABS_EEF_ACTION (servoL), robot_push_button.py:143-157 */
  int obs_dim;
  int tcp_site;            /* flange site id (robot tasks) */
  int arm_qadr;            /* first arm qpos/dof/ctrl index */
  int obj_body[MJS_MAXOBJ];
  int obj_qadr[MJS_MAXOBJ];
  int obj_dadr[MJS_MAXOBJ];
  int pm_body;             /* PointMass: body whose xpos is observed */
  int pm_mocap;
  double control_dt;
  double time_limit;       /* composer time_limit (seconds); <=0 = infinity */
  double goal_threshold;
  double target_radius;
  double nearest_coef;
  double tcp_offset[3];    /* tcp_in_flange translation */
  double tcp_z;            /* Planar Push commanded z (robot_planar_push.py:199) */
  double arena_range[2];   /* PointMass X_RANGE */
  double spawn_lo[3];
  double spawn_hi[3];
  double target_lo[3];
  double target_hi[3];
  double obj_lo[3];
  double obj_hi[3];
  double action_lo[8];     /* box the synthetic-rollout actions are drawn from */
  double action_hi[8];
  double obj_radius;
  double ik_dh[6];         /* d1 a2 a3 d4 d5 d6 of the analytic UR IK */
  double button_goal[3];
  double button_goal_radius;
  double touch_lo;
  double touch_hi;
  int settle_steps;        /* Planar Push: 150 physics steps after placing objects */
  int button_disturbances; /* Button Push: after_step may deactivate the switch (robot_push_button.py:159-165) */
  double disturb_prob;     /* ... with this probability per control step (0.01 in the reference) */
} MjsTask;

/* Logical per-env state (AoS, double).  The oracle steps this struct directly;
 * the CUDA library keeps the same fields as env-major SoA tensors in HBM and
 * converts in mjs_get_state / mjs_set_state. */
typedef struct MjsEnvState {
  double qpos[MJS_MAXQ];
  double qvel[MJS_MAXV];
  double qacc_warm[MJS_MAXV];
  double ctrl[MJS_MAXU];
  double mocap_pos[MJS_MAXMOCAP * 3];
  double time;
  double target[3];
  double dist;
  double prev_dist;
  double traj_q0[6];
  double traj_q1[6];
  double traj_t0;
  double ovr_pos[3];
  double ep_return;
  int episode_step;
  int traj_active;
  int sw_active;
  int sw_pressed;
  int sw_count;
  int status;
  int episode_id;
  int pad0;
} MjsEnvState;

#define MJS_MAXOBS 24
/* Result of one control step for one env (what dmc2gym.py:133-155 returns). */
typedef struct MjsStepOut {
  double obs[MJS_MAXOBS];
  double reward;
  double discount;
  int terminated;
  int truncated;
  int success;
  int done;
} MjsStepOut;

/* ---- camera observations as ray casting (replaces entities/camera.py:76-103 `physics.render`) ---- */
#define MJS_MAXVIS 32
#define MJS_ANCHOR_WORLD 0   /* pose given in the world frame */
#define MJS_ANCHOR_GEOM 1    /* pose of model geom `geom` (its own pos/quat/size; `pos`/`quat` unused) */
#define MJS_ANCHOR_TARGET 2  /* translated to the env's task target (state.target; z from `pos[2]` when the target is planar) */
#define MJS_ANCHOR_OVR 3     /* translated by the env's overridable body position (Button Push switch) */
#define MJS_ANCHOR_GEOMFRAME 4 /* `pos`/`quat` are local to the frame of model geom `geom` (wrist camera) */
/* One visible primitive: a model geom, or a marker (target site, switch parts) that follows per-env state. */
typedef struct MjsVisual {
  int type;              /* MJS_GEOM_* */
  int geom;              /* model geom index for MJS_ANCHOR_GEOM / GEOMFRAME, else -1 */
  int anchor;
  int use_active;        /* 1: colour is rgba_active while the env's switch is active (props/switch.py:59) */
  double pos[3];
  double quat[4];
  double size[3];
  double rgba[4];
  double rgba_active[4];
} MjsVisual;

/* Fixed camera (entities/camera.py:86-88: mode="fixed", looks along its -z axis, +y up), optionally riding on a geom frame. */
typedef struct MjsCamera {
  int anchor;            /* MJS_ANCHOR_WORLD or MJS_ANCHOR_GEOMFRAME */
  int geom;
  double pos[3];
  double quat[4];
  double fovy;           /* vertical field of view, degrees */
} MjsCamera;

#endif /* MJS_MODEL_H_ */
