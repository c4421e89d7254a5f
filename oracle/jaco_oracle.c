/* jaco_oracle.c -- CPU ORACLE (test infrastructure, NOT the product).  See jaco_oracle.h for scope and the
 * "parity unpinned" statement.  Plain C, one env per object, dense generic algebra, written for clarity.
 * Build: see oracle/Makefile (double build libjaco_oracle.so, float build libjaco_oracle_f32.so). */
#include "jaco_oracle.h"
#include "../include/jaco_model_layout.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifdef JO_FLOAT
typedef float real;
#else
typedef double real;
#endif
#define R_(x) ((real)(x))
#define rsqrt_(x) ((real)sqrt((double)(x)))
#define rsin(x) ((real)sin((double)(x)))
#define rcos(x) ((real)cos((double)(x)))
#define ratan2(y, x) ((real)atan2((double)(y), (double)(x)))
#define racos(x) ((real)acos((double)(x)))
#define rasin(x) ((real)asin((double)(x)))
#define rexp(x) ((real)exp((double)(x)))
#define rpow(x, y) ((real)pow((double)(x), (double)(y)))
#define rabs(x) ((real)fabs((double)(x)))
#define MJ_MINVAL R_(1e-15)
#define PI_ R_(3.14159265358979323846)

enum { NB = 12, NCH = 9, NV = JO_NV, NQ = JO_NQ };
enum { G_PLANE = 0, G_BOX = 1, G_CAPSULE = 2 };
static const int CH_PARENT[NCH] = {-1, 0, 1, 2, 3, 4, 5, 5, 5};

/* ------------------------------------------------------------------------------------------- small math */
static void v3set(real *r, real a, real b, real c) { r[0] = a; r[1] = b; r[2] = c; }
static void v3copy(real *r, const real *a) { r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; }
static void v3add(real *r, const real *a, const real *b) { r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; }
static void v3sub(real *r, const real *a, const real *b) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
static void v3addscl(real *r, const real *a, const real *b, real s) { r[0] = a[0] + s * b[0]; r[1] = a[1] + s * b[1]; r[2] = a[2] + s * b[2]; }
static real v3dot(const real *a, const real *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static void v3cross(real *r, const real *a, const real *b) {
  real x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static real v3norm(const real *a) { return rsqrt_(v3dot(a, a)); }
/* row-major 3x3: r = M v, r = M^T v */
static void m3mulv(real *r, const real *M, const real *v) {
  real x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2], y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2], z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void m3tmulv(real *r, const real *M, const real *v) {
  real x = M[0] * v[0] + M[3] * v[1] + M[6] * v[2], y = M[1] * v[0] + M[4] * v[1] + M[7] * v[2], z = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void m3mul(real *r, const real *A, const real *B) {
  real t[9];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  memcpy(r, t, sizeof t);
}
static void m3mult(real *r, const real *A, const real *B) { /* A * B^T */
  real t[9];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) t[3 * i + j] = A[3 * i] * B[3 * j] + A[3 * i + 1] * B[3 * j + 1] + A[3 * i + 2] * B[3 * j + 2];
  memcpy(r, t, sizeof t);
}
static void quat_mul(real *r, const real *a, const real *b) {
  real w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  real x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  real y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  real z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
static void quat_normalize(real *q) { /* mju_normalize4: zero quaternion -> identity */
  real n = rsqrt_(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < MJ_MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
  else { for (int i = 0; i < 4; i++) q[i] /= n; }
}
static void quat2mat(real *M, const real *q) {
  real w = q[0], x = q[1], y = q[2], z = q[3];
  M[0] = w * w + x * x - y * y - z * z; M[1] = 2 * (x * y - w * z); M[2] = 2 * (x * z + w * y);
  M[3] = 2 * (x * y + w * z); M[4] = w * w - x * x + y * y - z * z; M[5] = 2 * (y * z - w * x);
  M[6] = 2 * (x * z - w * y); M[7] = 2 * (y * z + w * x); M[8] = w * w - x * x - y * y + z * z;
}
static void axisangle2quat(real *q, const real *axis, real ang) {
  real s = rsin(ang * R_(0.5));
  q[0] = rcos(ang * R_(0.5)); q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
/* Gohlke transformations, axes='rxyz' (SURVEY App. A.1): R = Rx(a) Ry(b) Rz(g) */
static void euler_rxyz_to_quat(real *q, real a, real b, real g) {
  real qx[4] = {rcos(a / 2), rsin(a / 2), 0, 0}, qy[4] = {rcos(b / 2), 0, rsin(b / 2), 0}, qz[4] = {rcos(g / 2), 0, 0, rsin(g / 2)}, t[4];
  quat_mul(t, qx, qy); quat_mul(q, t, qz);
}
static void mat_to_euler_rxyz(real *e, const real *M) {
  real cy = rsqrt_(M[8] * M[8] + M[5] * M[5]);
  if (cy > R_(8.881784197001252e-16)) { e[0] = ratan2(-M[5], M[8]); e[1] = ratan2(M[2], cy); e[2] = ratan2(-M[1], M[0]); }
  else { e[0] = 0; e[1] = ratan2(M[2], cy); e[2] = ratan2(M[3], M[4]); }
}
static void quat_to_euler_rxyz(real *e, const real *q) { /* euler_from_quaternion = euler_from_matrix(quaternion_matrix(q)) */
  real n = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3], M[9];
  if (n < R_(8.881784197001252e-16)) { real I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; mat_to_euler_rxyz(e, I); return; }
  real s = rsqrt_(1 / n), qq[4] = {q[0] * s, q[1] * s, q[2] * s, q[3] * s};
  quat2mat(M, qq);
  mat_to_euler_rxyz(e, M);
}
/* util:448-468 _get_rotation */
static void get_rotation(real *out, real roll, real pitch, real yaw, const real *vec, int inv) {
  real Rx[9] = {1, 0, 0, 0, rcos(roll), -rsin(roll), 0, rsin(roll), rcos(roll)};
  real Ry[9] = {rcos(pitch), 0, rsin(pitch), 0, 1, 0, -rsin(pitch), 0, rcos(pitch)};
  real Rz[9] = {rcos(yaw), -rsin(yaw), 0, rsin(yaw), rcos(yaw), 0, 0, 0, 1};
  real T[9], Rm[9];
  m3mul(T, Ry, Rz); m3mul(Rm, Rx, T);
  if (inv) m3tmulv(out, Rm, vec); else m3mulv(out, Rm, vec);
}

/* dense Cholesky of the leading n x n of a row-major (ld) SPD matrix; returns 0 on a non-positive pivot */
static int chol_factor(real *A, int n, int ld) {
  for (int j = 0; j < n; j++) {
    real d = A[j * ld + j];
    for (int k = 0; k < j; k++) d -= A[j * ld + k] * A[j * ld + k];
    if (!(d > 0)) return 0;
    d = rsqrt_(d); A[j * ld + j] = d;
    for (int i = j + 1; i < n; i++) {
      real s = A[i * ld + j];
      for (int k = 0; k < j; k++) s -= A[i * ld + k] * A[j * ld + k];
      A[i * ld + j] = s / d;
    }
  }
  return 1;
}
static void chol_solve(const real *L, int n, int ld, real *x) {
  for (int i = 0; i < n; i++) { real s = x[i]; for (int k = 0; k < i; k++) s -= L[i * ld + k] * x[k]; x[i] = s / L[i * ld + i]; }
  for (int i = n - 1; i >= 0; i--) { real s = x[i]; for (int k = i + 1; k < n; k++) s -= L[k * ld + i] * x[k]; x[i] = s / L[i * ld + i]; }
}

/* ------------------------------------------------------------------------------------------- Philox4x32-10 */
void jo_philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t out[4]) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32), c[4] = {c0, c1, c2, c3};
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  memcpy(out, c, sizeof(uint32_t) * 4);
}

/* ------------------------------------------------------------------------------------------- model + data */
typedef struct {
  int type, body;
  real pos[3], R[9], size[3], rbound;
} geom_t;
typedef struct {
  int g1, g2, group, condim;
  real friction[3], solref[2], solimp[5], margin, invw_tran, invw_rot;
} pair_t;
typedef struct { int type, body, geom; real pos[3], R[9], size[3]; } site_t;

typedef struct {
  real pos[3], frame[9], dist;
  int g1, g2, pair, dim, efc;
  real mu[5];
  real nforce;
} contact_t;

struct jo_env {
  /* model */
  real ch_pos[NCH][3], ch_quat[NCH][4], ch_axis[NCH][3], ch_ref[NCH], ch_range[NCH][2], ch_damping[NCH];
  int ch_limited[NCH];
  real mass[NB], com[NB][3], inertia[NB][9]; /* body-frame, index by body id (0 world unused) */
  real free_qpos0[2][7];
  real motor_fr[6][2], finger_act[3][5];
  real ee_off[3], ee_obj_off[3], base_pos[3], base_link_pos[3], base_link_eul[3];
  real dof_invw[NV], meaninertia, hand_sphere[4];
  real dt, gravz, tol; int maxiter;
  geom_t geoms[JM_NGEOM]; pair_t pairs[JM_NPAIR]; site_t sites[JM_NSITE];
  /* state */
  real qpos[NQ], qvel[NV], qacc_warm[NV], ctrl[JO_NU];
  /* mjData of the last forward */
  real dq[NQ], dv[NV]; /* state the derived data below were evaluated at */
  real xpos[NB][3], xmat[NB][9], axis_w[NCH][3], com_w[NB][3], iw[NB][9];
  real M[NV][NV], bias[NV], qfrc_smooth[NV], qacc_smooth[NV], qacc[NV], qfrc_constraint[NV];
  real ee_pos[3], ee_quat[4], Jee[6][6], obj_xpos[3];
  real sensordata[20];
  int touch_knife_edge;   /* test probe: see touch_sensors */
  int probe, contact_knife_edge; real margin_shift;   /* test probe (jo_set_probe): contact count of the last forward under margins +-2 um */
  int ncon, nefc, solver_iters; real solver_grad;
  contact_t con[JO_MAXCON];
  real w_jar[JO_MAXEFC], w_jp[JO_MAXEFC], w_bp[JO_MAXEFC];
  real efc_J[JO_MAXEFC][NV], efc_pos[JO_MAXEFC], efc_D[JO_MAXEFC], efc_aref[JO_MAXEFC], efc_force[JO_MAXEFC];
  /* env logic (env_mujoco_util.py) */
  int task, variant, obs_mode, binary, env_id; uint64_t seed; uint32_t rng_counter;
  real target[6], gripper_angle, ramp[50];
  int gripper_iter, action_in, picked, touch_index, num_episodes, current_steps, task_max_steps;
  real obj_goal[3], dest_goal[3], reach_goal[6];
  real gripper_pose[6];
};

static void load_model(jo_env *e, const double *b) {
  for (int k = 0; k < NCH; k++) {
    for (int i = 0; i < 3; i++) { e->ch_pos[k][i] = R_(b[JM_CHAIN_POS + 3 * k + i]); e->ch_axis[k][i] = R_(b[JM_CHAIN_AXIS + 3 * k + i]); e->com[k + 1][i] = R_(b[JM_CHAIN_COM + 3 * k + i]); }
    for (int i = 0; i < 4; i++) e->ch_quat[k][i] = R_(b[JM_CHAIN_QUAT + 4 * k + i]);
    e->ch_ref[k] = R_(b[JM_CHAIN_REF + k]); e->ch_range[k][0] = R_(b[JM_CHAIN_RANGE + 2 * k]); e->ch_range[k][1] = R_(b[JM_CHAIN_RANGE + 2 * k + 1]);
    e->ch_limited[k] = b[JM_CHAIN_LIMITED + k] != 0; e->ch_damping[k] = R_(b[JM_CHAIN_DAMPING + k]);
    e->mass[k + 1] = R_(b[JM_CHAIN_MASS + k]);
    const double *I = b + JM_CHAIN_INERTIA + 6 * k;
    real T[9] = {R_(I[0]), R_(I[3]), R_(I[4]), R_(I[3]), R_(I[1]), R_(I[5]), R_(I[4]), R_(I[5]), R_(I[2])};
    memcpy(e->inertia[k + 1], T, sizeof T);
  }
  for (int f = 0; f < 2; f++) {
    e->mass[10 + f] = R_(b[JM_FREE_MASS + f]);
    for (int i = 0; i < 3; i++) e->com[10 + f][i] = R_(b[JM_FREE_COM + 3 * f + i]);
    const double *I = b + JM_FREE_INERTIA + 6 * f;
    real T[9] = {R_(I[0]), R_(I[3]), R_(I[4]), R_(I[3]), R_(I[1]), R_(I[5]), R_(I[4]), R_(I[5]), R_(I[2])};
    memcpy(e->inertia[10 + f], T, sizeof T);
    for (int i = 0; i < 7; i++) e->free_qpos0[f][i] = R_(b[JM_FREE_QPOS0 + 7 * f + i]);
  }
  for (int i = 0; i < 6; i++) { e->motor_fr[i][0] = R_(b[JM_MOTOR_FORCERANGE + 2 * i]); e->motor_fr[i][1] = R_(b[JM_MOTOR_FORCERANGE + 2 * i + 1]); }
  for (int i = 0; i < 3; i++) for (int j = 0; j < 5; j++) e->finger_act[i][j] = R_(b[JM_FINGER_ACT + 5 * i + j]);
  for (int i = 0; i < 3; i++) { e->ee_off[i] = R_(b[JM_EE_OFFSET + i]); e->ee_obj_off[i] = R_(b[JM_EE_OBJ_OFFSET + i]); e->base_pos[i] = R_(b[JM_BASE_POSITION + i]); e->base_link_pos[i] = R_(b[JM_BASE_LINK_POSITION + i]); e->base_link_eul[i] = R_(b[JM_BASE_LINK_EULER + i]); }
  for (int i = 0; i < NV; i++) e->dof_invw[i] = R_(b[JM_DOF_INVWEIGHT0 + i]);
  e->meaninertia = R_(b[JM_MEANINERTIA]);
  for (int i = 0; i < 4; i++) e->hand_sphere[i] = R_(b[JM_HAND_SPHERE + i]);
  e->dt = R_(b[JM_OPT]); e->gravz = R_(b[JM_OPT + 1]); e->maxiter = (int)b[JM_OPT + 2]; e->tol = R_(b[JM_OPT + 3]);
  for (int g = 0; g < JM_NGEOM; g++) {
    const double *r = b + JM_GEOMS + JM_NGEOM_FIELDS * g; geom_t *G = &e->geoms[g];
    G->type = (int)r[0]; G->body = (int)r[1];
    for (int i = 0; i < 3; i++) { G->pos[i] = R_(r[2 + i]); G->size[i] = R_(r[14 + i]); }
    for (int i = 0; i < 9; i++) G->R[i] = R_(r[5 + i]);
    G->rbound = R_(r[17]);
  }
  for (int p = 0; p < JM_NPAIR; p++) {
    const double *r = b + JM_PAIRS + JM_NPAIR_FIELDS * p; pair_t *P = &e->pairs[p];
    P->g1 = (int)r[0]; P->g2 = (int)r[1]; P->group = (int)r[2]; P->condim = (int)r[3];
    for (int i = 0; i < 3; i++) P->friction[i] = R_(r[4 + i]);
    P->solref[0] = R_(r[7]); P->solref[1] = R_(r[8]);
    for (int i = 0; i < 5; i++) P->solimp[i] = R_(r[9 + i]);
    P->margin = R_(r[14]); P->invw_tran = R_(r[15]); P->invw_rot = R_(r[16]);
  }
  for (int s = 0; s < JM_NSITE; s++) {
    const double *r = b + JM_SITES + JM_NSITE_FIELDS * s; site_t *S = &e->sites[s];
    S->type = (int)r[0]; S->body = (int)r[1]; S->geom = (int)r[17];
    for (int i = 0; i < 3; i++) { S->pos[i] = R_(r[2 + i]); S->size[i] = R_(r[14 + i]); }
    for (int i = 0; i < 9; i++) S->R[i] = R_(r[5 + i]);
  }
}

/* ------------------------------------------------------------------------------------------- kinematics */
/* mj_kinematics + mj_comPos for the merged tree; normalises free-joint quaternions in qpos in place */
static void kinematics(jo_env *e) {
  memset(e->xpos[0], 0, sizeof e->xpos[0]);
  real I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  memcpy(e->xmat[0], I3, sizeof I3);
  for (int k = 0; k < NCH; k++) {
    int pb = CH_PARENT[k] + 1, b = k + 1;
    real t[3], Rq[9], Rj[9], qj[4], T[9];
    m3mulv(t, e->xmat[pb], e->ch_pos[k]); v3add(e->xpos[b], e->xpos[pb], t);
    quat2mat(Rq, e->ch_quat[k]);
    axisangle2quat(qj, e->ch_axis[k], e->qpos[k] - e->ch_ref[k]); quat2mat(Rj, qj);
    m3mul(T, e->xmat[pb], Rq); m3mul(e->xmat[b], T, Rj);
    m3mulv(e->axis_w[k], e->xmat[b], e->ch_axis[k]);
  }
  for (int f = 0; f < 2; f++) {
    real *q = e->qpos + 9 + 7 * f;
    quat_normalize(q + 3);
    v3copy(e->xpos[10 + f], q); quat2mat(e->xmat[10 + f], q + 3);
  }
  for (int b = 1; b < NB; b++) {
    real t[3], T[9];
    m3mulv(t, e->xmat[b], e->com[b]); v3add(e->com_w[b], e->xpos[b], t);
    m3mul(T, e->xmat[b], e->inertia[b]); m3mult(e->iw[b], T, e->xmat[b]);
  }
}
static int chain_has_dof(int k, int j) { /* is dof j an ancestor joint of chain body k */
  if (k < 6) return j <= k;
  return j < 6 || j == k;
}
/* mj_jac: translational / rotational Jacobian (3 x NV, row-major) of world point p fixed to body b */
static void jac_point(const jo_env *e, int b, const real *p, real Jp[3][NV], real Jr[3][NV]) {
  memset(Jp, 0, sizeof(real) * 3 * NV); memset(Jr, 0, sizeof(real) * 3 * NV);
  if (b >= 1 && b <= 9) {
    int k = b - 1;
    for (int j = 0; j < NCH; j++) if (chain_has_dof(k, j)) {
      real r[3], c[3];
      v3sub(r, p, e->xpos[j + 1]); v3cross(c, e->axis_w[j], r);
      for (int i = 0; i < 3; i++) { Jp[i][j] = c[i]; Jr[i][j] = e->axis_w[j][i]; }
    }
  } else if (b >= 10) {
    int s = 9 + 6 * (b - 10);
    real r[3]; v3sub(r, p, e->xpos[b]);
    for (int a = 0; a < 3; a++) {
      Jp[a][s + a] = 1;
      real ax[3] = {e->xmat[b][a], e->xmat[b][3 + a], e->xmat[b][6 + a]}, c[3];
      v3cross(c, ax, r);
      for (int i = 0; i < 3; i++) { Jr[i][s + 3 + a] = ax[i]; Jp[i][s + 3 + a] = c[i]; }
    }
  }
}
/* joint-space inertia as sum over bodies of J^T diag(m, I) J (equals mj_crb + mj_fullM) */
static void mass_matrix(jo_env *e) {
  memset(e->M, 0, sizeof e->M);
  for (int b = 1; b < NB; b++) {
    real Jp[3][NV], Jr[3][NV];
    jac_point(e, b, e->com_w[b], Jp, Jr);
    for (int i = 0; i < NV; i++) for (int j = 0; j < NV; j++) {
      real s = 0;
      for (int a = 0; a < 3; a++) {
        s += e->mass[b] * Jp[a][i] * Jp[a][j];
        for (int c = 0; c < 3; c++) s += Jr[a][i] * e->iw[b][3 * a + c] * Jr[c][j];
      }
      e->M[i][j] += s;
    }
  }
}
/* mj_rne(flg_acc=0): Coriolis/centrifugal + gravity */
static void rne_bias(jo_env *e) {
  real w[NB][3], al[NB][3], ao[NB][3], f[NB][3], n[NB][3];
  memset(w, 0, sizeof w); memset(al, 0, sizeof al); memset(ao, 0, sizeof ao);
  ao[0][2] = -e->gravz;
  for (int k = 0; k < NCH; k++) {
    int b = k + 1, p = CH_PARENT[k] + 1;
    real wj[3], r[3], t[3], t2[3], c[3], ac[3], Iw[3];
    v3addscl(wj, (real[3]){0, 0, 0}, e->axis_w[k], e->qvel[k]);
    v3add(w[b], w[p], wj);
    v3cross(t, w[p], wj); v3add(al[b], al[p], t);
    v3sub(r, e->xpos[b], e->xpos[p]);
    v3cross(t, al[p], r); v3cross(t2, w[p], r); v3cross(t2, w[p], t2);
    for (int i = 0; i < 3; i++) ao[b][i] = ao[p][i] + t[i] + t2[i];
    v3sub(c, e->com_w[b], e->xpos[b]);
    v3cross(t, al[b], c); v3cross(t2, w[b], c); v3cross(t2, w[b], t2);
    for (int i = 0; i < 3; i++) ac[i] = ao[b][i] + t[i] + t2[i];
    for (int i = 0; i < 3; i++) f[b][i] = e->mass[b] * ac[i];
    m3mulv(Iw, e->iw[b], w[b]); v3cross(t, w[b], Iw); m3mulv(t2, e->iw[b], al[b]);
    v3cross(n[b], c, f[b]);
    for (int i = 0; i < 3; i++) n[b][i] += t[i] + t2[i];
  }
  for (int k = NCH - 1; k >= 0; k--) {
    int b = k + 1, p = CH_PARENT[k] + 1;
    e->bias[k] = v3dot(e->axis_w[k], n[b]);
    if (p > 0) {
      real r[3], t[3];
      v3sub(r, e->xpos[b], e->xpos[p]); v3cross(t, r, f[b]);
      for (int i = 0; i < 3; i++) { f[p][i] += f[b][i]; n[p][i] += n[b][i] + t[i]; }
    }
  }
  for (int fb = 0; fb < 2; fb++) {
    int b = 10 + fb, s = 9 + 6 * fb;
    real ww[3], c[3], t[3], ac[3], F[3], Iw[3], N[3], no[3];
    m3mulv(ww, e->xmat[b], e->qvel + s + 3);
    v3sub(c, e->com_w[b], e->xpos[b]);
    v3cross(t, ww, c); v3cross(t, ww, t);
    v3set(ac, t[0], t[1], t[2] - e->gravz);
    for (int i = 0; i < 3; i++) F[i] = e->mass[b] * ac[i];
    m3mulv(Iw, e->iw[b], ww); v3cross(N, ww, Iw);
    v3cross(no, c, F); v3add(no, no, N);
    for (int i = 0; i < 3; i++) e->bias[s + i] = F[i];
    m3tmulv(e->bias + s + 3, e->xmat[b], no);
  }
}

/* ------------------------------------------------------------------------------------------- collision */
static void geom_world(const jo_env *e, int g, real *pos, real *R) {
  const geom_t *G = &e->geoms[g];
  real t[3]; m3mulv(t, e->xmat[G->body], G->pos); v3add(pos, e->xpos[G->body], t);
  m3mul(R, e->xmat[G->body], G->R);
}
static void make_frame(real *fr) { /* mju_makeFrame: fr[0:3] given */
  real *x = fr, *y = fr + 3, *z = fr + 6;
  if (rabs(x[1]) < R_(0.5)) v3set(y, 0, 1, 0); else v3set(y, 0, 0, 1);
  real d = v3dot(x, y); v3addscl(y, y, x, -d);
  real n = v3norm(y); for (int i = 0; i < 3; i++) y[i] /= n;
  v3cross(z, x, y);
}
static void add_contact(jo_env *e, int pair, const real *pos, const real *normal, real dist) {
  if (e->ncon >= JO_MAXCON) return;
  const pair_t *P = &e->pairs[pair];
  contact_t *c = &e->con[e->ncon++];
  v3copy(c->pos, pos); v3copy(c->frame, normal); make_frame(c->frame);
  c->dist = dist; c->g1 = P->g1; c->g2 = P->g2; c->pair = pair; c->dim = P->condim; c->efc = -1; c->nforce = 0;
  c->mu[0] = c->mu[1] = P->friction[0]; c->mu[2] = P->friction[1]; c->mu[3] = c->mu[4] = P->friction[2];
}
/* mjc_PlaneBox: corners below the box centre along the plane normal, at most 4 */
static void collide_plane_box(jo_env *e, int pair, const real *pp, const real *pR, const real *bp, const real *bR, const real *size, real margin) {
  real n[3] = {pR[2], pR[5], pR[8]}, d[3];
  v3sub(d, bp, pp);
  real dist = v3dot(d, n);
  int cnt = 0;
  for (int i = 0; i < 8; i++) {
    real v[3] = {(i & 1 ? size[0] : -size[0]), (i & 2 ? size[1] : -size[1]), (i & 4 ? size[2] : -size[2])}, corner[3], pos[3];
    m3mulv(corner, bR, v);
    real ld = v3dot(n, corner);
    if (dist + ld > margin || ld > 0) continue;
    real cd = dist + ld;
    v3add(pos, bp, corner); v3addscl(pos, pos, n, -cd * R_(0.5));
    add_contact(e, pair, pos, n, cd);
    if (++cnt >= 4) return;
  }
}
/* sphere (centre c, radius r) against box: returns 1 and fills normal (box->sphere), dist, pos */
static int sphere_box(const real *c, real r, const real *bp, const real *bR, const real *h, real margin, real *normal, real *dist, real *pos) {
  real d[3], l[3], q[3];
  v3sub(d, c, bp); m3tmulv(l, bR, d);
  int inside = 1;
  for (int i = 0; i < 3; i++) { q[i] = l[i]; if (q[i] > h[i]) { q[i] = h[i]; inside = 0; } else if (q[i] < -h[i]) { q[i] = -h[i]; inside = 0; } }
  real nl[3], dd;
  if (!inside) {
    real df[3]; v3sub(df, l, q); dd = v3norm(df);
    if (dd - r > margin) return 0;
    for (int i = 0; i < 3; i++) nl[i] = df[i] / dd;
    *dist = dd - r;
  } else { /* centre inside the box: push out through the nearest face */
    int best = 0; real bd = h[0] - rabs(l[0]);
    for (int i = 1; i < 3; i++) { real t = h[i] - rabs(l[i]); if (t < bd) { bd = t; best = i; } }
    v3set(nl, 0, 0, 0); nl[best] = l[best] >= 0 ? R_(1) : R_(-1);
    q[best] = nl[best] * h[best];
    *dist = -bd - r;
  }
  real qw[3]; m3mulv(normal, bR, nl); m3mulv(qw, bR, q); v3add(qw, qw, bp);
  v3addscl(pos, qw, normal, *dist * R_(0.5));
  return 1;
}
/* derivative of squared distance from segment point p0 + t*d (box frame) to the box, w.r.t. t (up to factor 2) */
static real seg_box_dfdt(const real *p0, const real *d, const real *h, real t) {
  real s = 0;
  for (int i = 0; i < 3; i++) { real x = p0[i] + t * d[i]; if (x > h[i]) s += (x - h[i]) * d[i]; else if (x < -h[i]) s += (x + h[i]) * d[i]; }
  return s;
}
/* capsule vs box.  Contacts: each end-cap sphere that penetrates; plus, unless both ends penetrate, the sphere at
 * the exact closest point of the axis to the box when that point is interior to the axis. */
static void collide_capsule_box(jo_env *e, int pair, int box_first, const real *cp, const real *cR, const real *csize, const real *bp, const real *bR, const real *bsize, real margin) {
  real ax[3] = {cR[2], cR[5], cR[8]}, p0[3], p1[3], r = csize[0];
  v3addscl(p0, cp, ax, -csize[1]); v3addscl(p1, cp, ax, csize[1]);
  real nrm[3], dist, pos[3];
  int hit0, hit1;
  real sgn = box_first ? R_(1) : R_(-1);
  hit0 = sphere_box(p0, r, bp, bR, bsize, margin, nrm, &dist, pos);
  if (hit0) { real n2[3] = {sgn * nrm[0], sgn * nrm[1], sgn * nrm[2]}; add_contact(e, pair, pos, n2, dist); }
  hit1 = sphere_box(p1, r, bp, bR, bsize, margin, nrm, &dist, pos);
  if (hit1) { real n2[3] = {sgn * nrm[0], sgn * nrm[1], sgn * nrm[2]}; add_contact(e, pair, pos, n2, dist); }
  if (hit0 && hit1) return;
  /* exact minimiser of the convex piecewise-quadratic distance along the axis */
  real d[3], l0[3], dl[3], t[3];
  v3sub(d, p1, p0); v3sub(t, p0, bp); m3tmulv(l0, bR, t); m3tmulv(dl, bR, d);
  real lo = 0, hi = 1, flo = seg_box_dfdt(l0, dl, bsize, 0), fhi = seg_box_dfdt(l0, dl, bsize, 1);
  if (flo >= 0 || fhi <= 0) return; /* minimum at an end cap: already handled */
  for (int i = 0; i < 3; i++) for (int s = -1; s <= 1; s += 2) {
    if (rabs(dl[i]) < MJ_MINVAL) continue;
    real tb = (s * bsize[i] - l0[i]) / dl[i];
    if (tb <= lo || tb >= hi) continue;
    real fb = seg_box_dfdt(l0, dl, bsize, tb);
    if (fb <= 0) { lo = tb; flo = fb; } else { hi = tb; fhi = fb; }
  }
  real ts = (fhi - flo > MJ_MINVAL) ? lo - flo * (hi - lo) / (fhi - flo) : R_(0.5) * (lo + hi);
  if (ts <= R_(0.02) || ts >= R_(0.98)) return;
  real c[3]; v3addscl(c, p0, d, ts);
  if (sphere_box(c, r, bp, bR, bsize, margin, nrm, &dist, pos)) { real n2[3] = {sgn * nrm[0], sgn * nrm[1], sgn * nrm[2]}; add_contact(e, pair, pos, n2, dist); }
}
/* mjc_PlaneCapsule: the two end-cap spheres */
static void collide_plane_capsule(jo_env *e, int pair, const real *pp, const real *pR, const real *cp, const real *cR, const real *csize, real margin) {
  real n[3] = {pR[2], pR[5], pR[8]}, ax[3] = {cR[2], cR[5], cR[8]};
  for (int s = -1; s <= 1; s += 2) {
    real c[3], d[3], pos[3];
    v3addscl(c, cp, ax, s * csize[1]); v3sub(d, c, pp);
    real dist = v3dot(d, n) - csize[0];
    if (dist > margin) continue;
    v3addscl(pos, c, n, -(csize[0] + dist * R_(0.5)));
    add_contact(e, pair, pos, n, dist);
  }
}
/* box-box: separating-axis test over 15 axes, then face clipping (<= 8 points) or one edge-edge point.
 * Own formulation, not MuJoCo's mjc_BoxBox. Normal points from box A (g1) to box B (g2). */
static void collide_box_box(jo_env *e, int pair, const real *pa, const real *Ra, const real *ha, const real *pb, const real *Rb, const real *hb, real margin) {
  real t[3], A[3][3], B[3][3], C[3][3], AC[3][3];
  v3sub(t, pb, pa);
  for (int i = 0; i < 3; i++) for (int k = 0; k < 3; k++) { A[i][k] = Ra[3 * k + i]; B[i][k] = Rb[3 * k + i]; }
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { C[i][j] = v3dot(A[i], B[j]); AC[i][j] = rabs(C[i][j]); }
  real ta[3] = {v3dot(t, A[0]), v3dot(t, A[1]), v3dot(t, A[2])}, tb[3] = {v3dot(t, B[0]), v3dot(t, B[1]), v3dot(t, B[2])};
  real best = R_(-1e30); int code = -1;
  for (int i = 0; i < 3; i++) { /* faces of A */
    real s = rabs(ta[i]) - (ha[i] + hb[0] * AC[i][0] + hb[1] * AC[i][1] + hb[2] * AC[i][2]);
    if (s > margin) return;
    if (s > best) { best = s; code = i; }
  }
  for (int j = 0; j < 3; j++) { /* faces of B */
    real s = rabs(tb[j]) - (hb[j] + ha[0] * AC[0][j] + ha[1] * AC[1][j] + ha[2] * AC[2][j]);
    if (s > margin) return;
    if (s > best) { best = s; code = 3 + j; }
  }
  real ebest = R_(-1e30); int ecode = -1; real eaxis[3] = {0, 0, 0};
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { /* edge x edge */
    real L[3]; v3cross(L, A[i], B[j]);
    real l = v3norm(L);
    if (l < R_(1e-6)) continue;
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    real ra = ha[i1] * AC[i2][j] + ha[i2] * AC[i1][j], rb = hb[j1] * AC[i][j2] + hb[j2] * AC[i][j1];
    real s = (rabs(v3dot(t, L)) - ra - rb) / l;
    if (s > margin) return;
    if (s > ebest) { ebest = s; ecode = 3 * i + j; for (int k = 0; k < 3; k++) eaxis[k] = L[k] / l; }
  }
  if (ecode >= 0 && ebest * R_(1.05) > best) { /* prefer faces unless the edge axis is 5% shallower */
    {
      int i = ecode / 3, j = ecode % 3;
      real n[3]; v3copy(n, eaxis);
      if (v3dot(n, t) < 0) for (int k = 0; k < 3; k++) n[k] = -n[k];
      /* supporting edge of A along +n, of B along -n */
      real pA[3], pB[3];
      v3copy(pA, pa); v3copy(pB, pb);
      for (int k = 0; k < 3; k++) if (k != i) { real s = v3dot(n, A[k]) > 0 ? R_(1) : R_(-1); v3addscl(pA, pA, A[k], s * ha[k]); }
      for (int k = 0; k < 3; k++) if (k != j) { real s = v3dot(n, B[k]) > 0 ? R_(-1) : R_(1); v3addscl(pB, pB, B[k], s * hb[k]); }
      /* closest points of the two lines pA + u A[i], pB + v B[j] */
      real w[3]; v3sub(w, pA, pB);
      real b_ = C[i][j], d_ = v3dot(A[i], w), e_ = v3dot(B[j], w), den = 1 - b_ * b_;
      real u = (b_ * e_ - d_) / den, v = (e_ - b_ * d_) / den;
      real qa[3], qb[3], pos[3];
      v3addscl(qa, pA, A[i], u); v3addscl(qb, pB, B[j], v);
      for (int k = 0; k < 3; k++) pos[k] = R_(0.5) * (qa[k] + qb[k]);
      add_contact(e, pair, pos, n, ebest);
      return;
    }
  }
  /* face contact: reference box owns the axis, incident face of the other box is clipped against it */
  int refA = code < 3, ax = refA ? code : code - 3;
  const real *pr = refA ? pa : pb, *hr = refA ? ha : hb, *pi = refA ? pb : pa, *hi = refA ? hb : ha;
  real (*Rr)[3] = refA ? A : B, (*Ri)[3] = refA ? B : A;
  real nr[3]; /* reference face normal, pointing towards the incident box */
  real sg = (refA ? ta[ax] : -tb[ax]) >= 0 ? R_(1) : R_(-1);
  for (int k = 0; k < 3; k++) nr[k] = sg * Rr[ax][k];
  /* incident face: most anti-parallel to nr */
  int ia = 0; real bd = -1;
  for (int k = 0; k < 3; k++) { real dd = rabs(v3dot(Ri[k], nr)); if (dd > bd) { bd = dd; ia = k; } }
  real si = v3dot(Ri[ia], nr) > 0 ? R_(-1) : R_(1);
  int i1 = (ia + 1) % 3, i2 = (ia + 2) % 3, r1 = (ax + 1) % 3, r2 = (ax + 2) % 3;
  real fc[3]; v3addscl(fc, pi, Ri[ia], si * hi[ia]);
  real poly[16][2], depth[16], tmp[16][2], tmpd[16]; int np = 4;
  static const int sx[4] = {1, -1, -1, 1}, sy[4] = {1, 1, -1, -1};
  for (int k = 0; k < 4; k++) {
    real v[3], dlt[3];
    v3addscl(v, fc, Ri[i1], sx[k] * hi[i1]); v3addscl(v, v, Ri[i2], sy[k] * hi[i2]);
    v3sub(dlt, v, pr);
    poly[k][0] = v3dot(dlt, Rr[r1]); poly[k][1] = v3dot(dlt, Rr[r2]); depth[k] = v3dot(dlt, nr) - hr[ax];
  }
  /* Sutherland-Hodgman against |x| <= hr[r1], |y| <= hr[r2]; depth is interpolated linearly */
  for (int c = 0; c < 4; c++) {
    int dim = c >> 1; real sgn = (c & 1) ? R_(-1) : R_(1), lim = dim ? hr[r2] : hr[r1];
    int nn = 0;
    for (int k = 0; k < np; k++) {
      int k2 = (k + 1) % np;
      real a = sgn * poly[k][dim] - lim, b2 = sgn * poly[k2][dim] - lim;
      if (a <= 0) { tmp[nn][0] = poly[k][0]; tmp[nn][1] = poly[k][1]; tmpd[nn] = depth[k]; nn++; }
      if ((a <= 0) != (b2 <= 0)) {
        real s = a / (a - b2);
        tmp[nn][0] = poly[k][0] + s * (poly[k2][0] - poly[k][0]); tmp[nn][1] = poly[k][1] + s * (poly[k2][1] - poly[k][1]);
        tmpd[nn] = depth[k] + s * (depth[k2] - depth[k]); nn++;
      }
    }
    np = nn; if (np == 0) return;
    memcpy(poly, tmp, sizeof(real) * 2 * np); memcpy(depth, tmpd, sizeof(real) * np);
  }
  real n[3]; for (int k = 0; k < 3; k++) n[k] = refA ? nr[k] : -nr[k];
  int cnt = 0;
  for (int k = 0; k < np && cnt < 8; k++) {
    if (depth[k] > margin) continue;
    real pos[3];
    v3addscl(pos, pr, Rr[r1], poly[k][0]); v3addscl(pos, pos, Rr[r2], poly[k][1]);
    v3addscl(pos, pos, nr, hr[ax] + depth[k] * R_(0.5));
    add_contact(e, pair, pos, n, depth[k]);
    cnt++;
  }
}
static void collision(jo_env *e) {
  e->ncon = 0;
  /* group broad phase: hand-cluster sphere (link6 frame) against the other body's bounding sphere */
  real hs[3], t[3];
  m3mulv(t, e->xmat[6], e->hand_sphere); v3add(hs, e->xpos[6], t);
  for (int p = 0; p < JM_NPAIR; p++) {
    const pair_t *P = &e->pairs[p];
    const geom_t *G1 = &e->geoms[P->g1], *G2 = &e->geoms[P->g2];
    real p1[3], R1[9], p2[3], R2[9];
    geom_world(e, P->g1, p1, R1); geom_world(e, P->g2, p2, R2);
    if (P->group >= 1 && P->group <= 4) { /* cull the whole hand cluster at once */
      const geom_t *O = (P->group == 1) ? G2 : G1; const real *po = (P->group == 1) ? p2 : p1;
      if (O->type != G_PLANE) { real d[3]; v3sub(d, hs, po); if (v3norm(d) > e->hand_sphere[3] + O->rbound + (P->margin + e->margin_shift)) continue; }
      else { real d[3]; v3sub(d, hs, po); if (d[2] > e->hand_sphere[3] + (P->margin + e->margin_shift)) continue; }
    }
    if (G1->type != G_PLANE && G2->type != G_PLANE) { real d[3]; v3sub(d, p1, p2); if (v3norm(d) > G1->rbound + G2->rbound + (P->margin + e->margin_shift)) continue; }
    if (G1->type == G_PLANE && G2->type == G_BOX) collide_plane_box(e, p, p1, R1, p2, R2, G2->size, (P->margin + e->margin_shift));
    else if (G1->type == G_PLANE && G2->type == G_CAPSULE) collide_plane_capsule(e, p, p1, R1, p2, R2, G2->size, (P->margin + e->margin_shift));
    else if (G1->type == G_BOX && G2->type == G_BOX) collide_box_box(e, p, p1, R1, G1->size, p2, R2, G2->size, (P->margin + e->margin_shift));
    else if (G1->type == G_BOX && G2->type == G_CAPSULE) collide_capsule_box(e, p, 1, p2, R2, G2->size, p1, R1, G1->size, (P->margin + e->margin_shift));
    else if (G1->type == G_CAPSULE && G2->type == G_BOX) collide_capsule_box(e, p, 0, p1, R1, G1->size, p2, R2, G2->size, (P->margin + e->margin_shift));
  }
}

/* ------------------------------------------------------------------------------------------- constraints */
static real impedance(const real *solimp, real x) { /* getimpedance: x = pos - margin */
  real d0 = solimp[0], dw = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  if (d0 == dw || width <= MJ_MINVAL) return R_(0.5) * (d0 + dw);
  real xx = rabs(x) / width;
  if (xx >= 1) return dw;
  if (xx == 0) return d0;
  real y;
  if (power == 1) y = xx;
  else if (xx <= mid) y = rpow(xx / mid, power) * mid;           /* a = 1/mid^(p-1) */
  else y = 1 - rpow((1 - xx) / (1 - mid), power) * (1 - mid);   /* b = 1/(1-mid)^(p-1) */
  return d0 + y * (dw - d0);
}
static void add_row(jo_env *e, const real *J, real pos, real margin, const real *solref, const real *solimp, real diagApprox, real *Rfirst) {
  if (e->nefc >= JO_MAXEFC) return;
  int r = e->nefc++;
  memcpy(e->efc_J[r], J, sizeof(real) * NV);
  real tc = solref[0] > 2 * e->dt ? solref[0] : 2 * e->dt, dr = solref[1], dmax = solimp[1];
  real k = 1 / fmax((double)MJ_MINVAL, (double)(dmax * dmax * tc * tc * dr * dr)), b = 2 / fmax((double)MJ_MINVAL, (double)(dmax * tc));
  real imp = impedance(solimp, pos - margin);
  real Rr = (1 - imp) * diagApprox / imp; if (Rr < MJ_MINVAL) Rr = MJ_MINVAL;
  real vel = 0; for (int i = 0; i < NV; i++) vel += J[i] * e->qvel[i];
  e->efc_pos[r] = pos; e->efc_D[r] = 1 / Rr;
  e->efc_aref[r] = -b * vel - k * imp * (pos - margin);
  if (Rfirst) *Rfirst = Rr;
}
static void make_constraints(jo_env *e) {
  static const real def_solref[2] = {R_(0.02), R_(1)}, def_solimp[5] = {R_(0.9), R_(0.95), R_(0.001), R_(0.5), R_(2)};
  e->nefc = 0;
  for (int k = 0; k < NCH; k++) if (e->ch_limited[k]) { /* mj_instantiateLimit */
    real J[NV] = {0}, d;
    d = e->qpos[k] - e->ch_range[k][0];
    if (d < 0) { J[k] = 1; add_row(e, J, d, 0, def_solref, def_solimp, e->dof_invw[k], 0); }
    d = e->ch_range[k][1] - e->qpos[k];
    if (d < 0) { J[k] = -1; add_row(e, J, d, 0, def_solref, def_solimp, e->dof_invw[k], 0); }
  }
  for (int c = 0; c < e->ncon; c++) { /* mj_instantiateContact, pyramidal cone */
    contact_t *con = &e->con[c];
    const pair_t *P = &e->pairs[con->pair];
    real Jp1[3][NV], Jr1[3][NV], Jp2[3][NV], Jr2[3][NV], Jc[6][NV];
    jac_point(e, e->geoms[con->g1].body, con->pos, Jp1, Jr1);
    jac_point(e, e->geoms[con->g2].body, con->pos, Jp2, Jr2);
    for (int a = 0; a < 3; a++) for (int i = 0; i < NV; i++) {
      real sp = 0, sr = 0;
      for (int x = 0; x < 3; x++) { sp += con->frame[3 * a + x] * (Jp2[x][i] - Jp1[x][i]); sr += con->frame[3 * a + x] * (Jr2[x][i] - Jr1[x][i]); }
      Jc[a][i] = sp; Jc[3 + a][i] = sr;
    }
    con->efc = e->nefc;
    if (con->dim == 1) { add_row(e, Jc[0], con->dist, P->margin, P->solref, P->solimp, P->invw_tran, 0); continue; }
    int first = e->nefc; real R0 = 0;
    for (int k = 1; k < con->dim; k++) for (int s = 1; s >= -1; s -= 2) {
      real J[NV], mu = con->mu[k - 1];
      for (int i = 0; i < NV; i++) J[i] = Jc[0][i] + s * mu * Jc[k][i];
      real dA = P->invw_tran + mu * mu * (k - 1 < 2 ? P->invw_tran : P->invw_rot);
      add_row(e, J, con->dist, P->margin, P->solref, P->solimp, dA, (k == 1 && s == 1) ? &R0 : 0);
    }
    real Rpy = 2 * con->mu[0] * con->mu[0] * R0; if (Rpy < MJ_MINVAL) Rpy = MJ_MINVAL;
    for (int r = first; r < e->nefc; r++) e->efc_D[r] = 1 / Rpy;
  }
}
/* primal cost, gradient pieces for qacc = a */
static real cost_at(const jo_env *e, const real *a, real *jar_out) {
  real c = 0, Ma[NV];
  for (int i = 0; i < NV; i++) { real s = 0; for (int j = 0; j < NV; j++) s += e->M[i][j] * (a[j] - e->qacc_smooth[j]); Ma[i] = s; }
  for (int i = 0; i < NV; i++) c += R_(0.5) * Ma[i] * (a[i] - e->qacc_smooth[i]);
  for (int r = 0; r < e->nefc; r++) {
    real jar = -e->efc_aref[r];
    for (int i = 0; i < NV; i++) jar += e->efc_J[r][i] * a[i];
    if (jar_out) jar_out[r] = jar;
    if (jar < 0) c += R_(0.5) * e->efc_D[r] * jar * jar;
  }
  return c;
}
static int cmp_real(const void *a, const void *b) { real x = *(const real *)a, y = *(const real *)b; return (x > y) - (x < y); }
/* Newton on the primal problem (mj_solNewton semantics; exact line search by break-point walk) */
static void solve_constraints(jo_env *e) {
  real a[NV];
  e->solver_iters = 0; e->solver_grad = 0;
  if (e->nefc == 0) { memcpy(e->qacc, e->qacc_smooth, sizeof a); memset(e->qfrc_constraint, 0, sizeof a); return; }
  /* warm start: whichever of qacc_warmstart / qacc_smooth has the lower cost */
  real cw = cost_at(e, e->qacc_warm, 0), cs = cost_at(e, e->qacc_smooth, 0);
  memcpy(a, cw < cs ? e->qacc_warm : e->qacc_smooth, sizeof a);
  real scale = 1 / (e->meaninertia * NV);
  real *jar = e->w_jar, *jp = e->w_jp, *bp = e->w_bp;
  real cost = cost_at(e, a, jar);
  for (int it = 0; it < e->maxiter; it++) {
    real g[NV], H[NV][NV], p[NV];
    for (int i = 0; i < NV; i++) { real s = 0; for (int j = 0; j < NV; j++) s += e->M[i][j] * (a[j] - e->qacc_smooth[j]); g[i] = s; }
    memcpy(H, e->M, sizeof H);
    for (int r = 0; r < e->nefc; r++) if (jar[r] < 0) {
      real D = e->efc_D[r], f = -D * jar[r];
      for (int i = 0; i < NV; i++) { if (e->efc_J[r][i] == 0) continue; g[i] -= e->efc_J[r][i] * f; for (int j = 0; j < NV; j++) H[i][j] += D * e->efc_J[r][i] * e->efc_J[r][j]; }
    }
    real gn = 0; for (int i = 0; i < NV; i++) gn += g[i] * g[i];
    gn = rsqrt_(gn); e->solver_grad = scale * gn;
    if (scale * gn < e->tol) break;
    if (!chol_factor(&H[0][0], NV, NV)) break;
    for (int i = 0; i < NV; i++) p[i] = -g[i];
    chol_solve(&H[0][0], NV, NV, p);
    /* phi(al) = cost(a + al p): derivative is piecewise linear and increasing */
    real q1 = 0, q2 = 0;
    for (int i = 0; i < NV; i++) { real s = 0; for (int j = 0; j < NV; j++) s += e->M[i][j] * p[j]; q2 += p[i] * s; }
    for (int i = 0; i < NV; i++) { real s = 0; for (int j = 0; j < NV; j++) s += e->M[i][j] * (a[j] - e->qacc_smooth[j]); q1 += p[i] * s; }
    int nb = 0;
    for (int r = 0; r < e->nefc; r++) {
      real s = 0; for (int i = 0; i < NV; i++) s += e->efc_J[r][i] * p[i];
      jp[r] = s;
      if (s != 0) { real al = -jar[r] / s; if (al > 0) bp[nb++] = al; }
    }
    qsort(bp, nb, sizeof(real), cmp_real);
    real lo = 0, alpha = 0; int done = 0;
    for (int k = 0; k <= nb && !done; k++) {
      real hi = k < nb ? bp[k] : R_(1e30), mid = k < nb ? R_(0.5) * (lo + hi) : lo + 1;
      real d1 = q1, d2 = q2; /* phi'(al) = d1 + d2*al on this segment, using the active set at mid */
      for (int r = 0; r < e->nefc; r++) if (jar[r] + mid * jp[r] < 0) { d1 += e->efc_D[r] * jar[r] * jp[r]; d2 += e->efc_D[r] * jp[r] * jp[r]; }
      if (d2 > 0) { real al = -d1 / d2; if (al <= hi) { alpha = al < lo ? lo : al; done = 1; } }
      lo = hi;
    }
    if (!done || !(alpha > 0)) break;
    for (int i = 0; i < NV; i++) a[i] += alpha * p[i];
    real newcost = cost_at(e, a, jar);
    e->solver_iters = it + 1;
    real improvement = scale * (cost - newcost);
    cost = newcost;
    if (improvement < e->tol) break;
  }
  memcpy(e->qacc, a, sizeof a);
  memset(e->qfrc_constraint, 0, sizeof a);
  for (int r = 0; r < e->nefc; r++) {
    real f = jar[r] < 0 ? -e->efc_D[r] * jar[r] : 0;
    e->efc_force[r] = f;
    for (int i = 0; i < NV; i++) e->qfrc_constraint[i] += e->efc_J[r][i] * f;
  }
}

/* ------------------------------------------------------------------------------------------- sensors */
static int ray_hits_site(const site_t *S, const real *spos, const real *sR, const real *pnt, const real *vec) {
  real d[3], p[3], v[3];
  v3sub(d, pnt, spos); m3tmulv(p, sR, d); m3tmulv(v, sR, vec);
  real t0 = R_(-1e30), t1 = R_(1e30);
  int naxes = S->type == 0 ? 3 : 1;
  for (int k = 0; k < naxes; k++) {
    int i = S->type == 0 ? k : 2; real h = S->type == 0 ? S->size[i] : S->size[1];
    if (rabs(v[i]) < MJ_MINVAL) { if (rabs(p[i]) > h) return 0; continue; }
    real ta = (-h - p[i]) / v[i], tb = (h - p[i]) / v[i];
    if (ta > tb) { real s = ta; ta = tb; tb = s; }
    if (ta > t0) t0 = ta; if (tb < t1) t1 = tb;
  }
  if (S->type == 1) { /* cylinder side: |p_xy + t v_xy|^2 <= r^2 */
    real a = v[0] * v[0] + v[1] * v[1], b = p[0] * v[0] + p[1] * v[1], c = p[0] * p[0] + p[1] * p[1] - S->size[0] * S->size[0];
    if (a < MJ_MINVAL) { if (c > 0) return 0; }
    else { real disc = b * b - a * c; if (disc < 0) return 0; real sq = rsqrt_(disc), ta = (-b - sq) / a, tb = (-b + sq) / a; if (ta > t0) t0 = ta; if (tb < t1) t1 = tb; }
  }
  return t1 >= (t0 > 0 ? t0 : 0);
}
/* mjSENS_TOUCH: sum of normal forces of contacts on the site's body whose force ray meets the site volume */
static void touch_sensors(jo_env *e) {
  e->touch_knife_edge = 0;
  for (int c = 0; c < e->ncon; c++) {
    contact_t *con = &e->con[c]; real f = 0;
    int nr = con->dim == 1 ? 1 : 2 * (con->dim - 1);
    if (con->efc >= 0) for (int r = 0; r < nr && con->efc + r < e->nefc; r++) f += e->efc_force[con->efc + r];
    con->nforce = f;
  }
  for (int s = 0; s < JM_NSITE; s++) {
    const site_t *S = &e->sites[s];
    real spos[3], sR[9], t[3], sum = 0;
    m3mulv(t, e->xmat[S->body], S->pos); v3add(spos, e->xpos[S->body], t); m3mul(sR, e->xmat[S->body], S->R);
    for (int c = 0; c < e->ncon; c++) {
      const contact_t *con = &e->con[c];
      if (con->g1 != S->geom && con->g2 != S->geom) continue;
      if (con->nforce <= 0) continue;
      real ray[3]; real sg = con->g1 == S->geom ? R_(1) : R_(-1);
      for (int i = 0; i < 3; i++) ray[i] = sg * con->frame[i];
      if (ray_hits_site(S, spos, sR, con->pos, ray)) sum += con->nforce;
      { /* knife-edge probe for the parity tests: does the decision change when the site grows / shrinks by 2 um?  (pad geoms and
           their sensor sites have the same lateral size, so a contact at a pad edge sits exactly on the site boundary) */
        site_t big = *S, small = *S;
        for (int i = 0; i < 3; i++) { big.size[i] += R_(2e-6); small.size[i] -= R_(2e-6); }
        if (ray_hits_site(&big, spos, sR, con->pos, ray) != ray_hits_site(&small, spos, sR, con->pos, ray)) e->touch_knife_edge = 1;
      }
    }
    e->sensordata[s] = sum;
  }
}

/* ------------------------------------------------------------------------------------------- forward / step */
static void fwd_actuation(const jo_env *e, real *qfrc) {
  memset(qfrc, 0, sizeof(real) * NV);
  for (int i = 0; i < 6; i++) { real f = e->ctrl[i]; if (f < e->motor_fr[i][0]) f = e->motor_fr[i][0]; if (f > e->motor_fr[i][1]) f = e->motor_fr[i][1]; qfrc[i] = f; }
  for (int i = 0; i < 3; i++) {
    const real *A = e->finger_act[i]; real c = e->ctrl[6 + i];
    if (c < A[1]) c = A[1]; if (c > A[2]) c = A[2];
    real f = A[0] * c - A[0] * e->qpos[6 + i];
    if (f < A[3]) f = A[3]; if (f > A[4]) f = A[4];
    qfrc[6 + i] = f;
  }
}
/* body quaternion of link6 (= xquat of body EE): accumulate like mj_kinematics */
static void ee_quaternion(const jo_env *e, const real *qpos, real *q) {
  real acc[4] = {1, 0, 0, 0}, t[4], qj[4];
  for (int k = 0; k < 6; k++) { quat_mul(t, acc, e->ch_quat[k]); axisangle2quat(qj, e->ch_axis[k], qpos[k] - e->ch_ref[k]); quat_mul(acc, t, qj); }
  memcpy(q, acc, sizeof acc);
}
static void mj_forward(jo_env *e) {
  kinematics(e);
  mass_matrix(e);
  if (e->probe) { /* parity-test probe: is a geom pair within 2 um of making / breaking contact at this state? */
    e->margin_shift = R_(2e-6); collision(e); int np_ = e->ncon;
    e->margin_shift = R_(-2e-6); collision(e); int nm_ = e->ncon;
    e->margin_shift = 0; e->contact_knife_edge = np_ != nm_;
  }
  collision(e);
  rne_bias(e);
  real act[NV];
  fwd_actuation(e, act);
  for (int i = 0; i < NV; i++) e->qfrc_smooth[i] = act[i] - e->bias[i] - (i < NCH ? e->ch_damping[i] * e->qvel[i] : 0);
  real L[NV][NV];
  memcpy(L, e->M, sizeof L); chol_factor(&L[0][0], NV, NV);
  memcpy(e->qacc_smooth, e->qfrc_smooth, sizeof e->qacc_smooth); chol_solve(&L[0][0], NV, NV, e->qacc_smooth);
  make_constraints(e);
  solve_constraints(e);
  touch_sensors(e);
  /* quantities the controller / env read from mjData */
  memcpy(e->dq, e->qpos, sizeof e->dq); memcpy(e->dv, e->qvel, sizeof e->dv);
  real t[3], Jp[3][NV], Jr[3][NV];
  m3mulv(t, e->xmat[6], e->ee_off); v3add(e->ee_pos, e->xpos[6], t);
  jac_point(e, 6, e->ee_pos, Jp, Jr);
  for (int a = 0; a < 3; a++) for (int j = 0; j < 6; j++) { e->Jee[a][j] = Jp[a][j]; e->Jee[3 + a][j] = Jr[a][j]; }
  v3copy(e->obj_xpos, e->xpos[10]);
  ee_quaternion(e, e->qpos, e->ee_quat);
}
static void mj_euler(jo_env *e) {
  real a[NV], h = e->dt;
  int damp = 0; for (int i = 0; i < NCH; i++) if (e->ch_damping[i] > 0) damp = 1;
  if (damp) { /* (M + h D) qacc = qfrc_smooth + qfrc_constraint */
    real H[NV][NV];
    memcpy(H, e->M, sizeof H);
    for (int i = 0; i < NCH; i++) H[i][i] += h * e->ch_damping[i];
    for (int i = 0; i < NV; i++) a[i] = e->qfrc_smooth[i] + e->qfrc_constraint[i];
    chol_factor(&H[0][0], NV, NV); chol_solve(&H[0][0], NV, NV, a);
  } else memcpy(a, e->qacc, sizeof a);
  for (int i = 0; i < NV; i++) e->qvel[i] += h * a[i];
  for (int i = 0; i < NCH; i++) e->qpos[i] += h * e->qvel[i];
  for (int f = 0; f < 2; f++) { /* mju_quatIntegrate with the local angular velocity */
    real *q = e->qpos + 9 + 7 * f, *v = e->qvel + 9 + 6 * f;
    for (int i = 0; i < 3; i++) q[i] += h * v[i];
    real w[3] = {v[3], v[4], v[5]}, n = v3norm(w);
    if (n > MJ_MINVAL) { real ax[3] = {w[0] / n, w[1] / n, w[2] / n}, dq[4], r[4]; axisangle2quat(dq, ax, n * h); quat_mul(r, q + 3, dq); memcpy(q + 3, r, sizeof r); quat_normalize(q + 3); }
  }
  memcpy(e->qacc_warm, e->qacc, sizeof e->qacc_warm); /* mj_advance saves the warm start */
}
static void mj_step_(jo_env *e) { mj_forward(e); mj_euler(e); }

/* ------------------------------------------------------------------------------------------- OSC (App. A.2) */
/* parity-test probe (jo_osc_knife_edge): set when an OSC call sits on a decision boundary that float32 rounding may take either way:
   |det(J M^-1 J^T)| within 2 % of the 1e-3 branch threshold, an eigenvalue of the pinv branch within 30 % of the rcond cut-off, or the
   orientation error within 1e-4 of 180 degrees (sign flip of the error vector) */
static int g_osc_knife_edge = 0;
/* symmetric pseudo-inverse via cyclic Jacobi (numpy.linalg.pinv(X, rcond) for symmetric PSD X) */
static void sym_pinv6(const real *X, real rcond, real *out) {
  real A[6][6], V[6][6];
  for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) { A[i][j] = X[6 * i + j]; V[i][j] = i == j; }
  for (int sweep = 0; sweep < 30; sweep++) {
    real off = 0; for (int i = 0; i < 6; i++) for (int j = i + 1; j < 6; j++) off += A[i][j] * A[i][j];
    if (off < R_(1e-30)) break;
    for (int p = 0; p < 6; p++) for (int q = p + 1; q < 6; q++) {
      if (rabs(A[p][q]) < R_(1e-300)) continue;
      real th = (A[q][q] - A[p][p]) / (2 * A[p][q]);
      real t = (th >= 0 ? R_(1) : R_(-1)) / (rabs(th) + rsqrt_(th * th + 1)), c = 1 / rsqrt_(t * t + 1), s = t * c;
      for (int k = 0; k < 6; k++) { real akp = A[k][p], akq = A[k][q]; A[k][p] = c * akp - s * akq; A[k][q] = s * akp + c * akq; }
      for (int k = 0; k < 6; k++) { real apk = A[p][k], aqk = A[q][k]; A[p][k] = c * apk - s * aqk; A[q][k] = s * apk + c * aqk; }
      for (int k = 0; k < 6; k++) { real vkp = V[k][p], vkq = V[k][q]; V[k][p] = c * vkp - s * vkq; V[k][q] = s * vkp + c * vkq; }
    }
  }
  real lmax = 0; for (int i = 0; i < 6; i++) if (rabs(A[i][i]) > lmax) lmax = rabs(A[i][i]);
  for (int k = 0; k < 6; k++) { real ratio = rabs(A[k][k]) / (lmax > 0 ? lmax : 1); if (ratio > R_(0.7) * rcond && ratio < R_(1.3) * rcond) g_osc_knife_edge = 1; }
  for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) {
    real s = 0;
    for (int k = 0; k < 6; k++) if (rabs(A[k][k]) > rcond * lmax) s += V[i][k] * V[j][k] / A[k][k];
    out[6 * i + j] = s;
  }
}
static void osc_generate(const real J[6][6], const real M[6][6], const real *bias, const real *ee_pos, const real *ee_quat, const real *dq, const real *target, real *u) {
  const real kp = 50, ko = 180, kv = 20, vmax0 = R_(0.4), vmax1 = R_(1.0472);
  real L[6][6], Minv[6][6], X[6][6], Mx[6][6];
  memcpy(L, M, sizeof L); chol_factor(&L[0][0], 6, 6);
  for (int c = 0; c < 6; c++) { real col[6] = {0}; col[c] = 1; chol_solve(&L[0][0], 6, 6, col); for (int r = 0; r < 6; r++) Minv[r][c] = col[r]; }
  for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) { real s = 0; for (int a = 0; a < 6; a++) for (int b = 0; b < 6; b++) s += J[i][a] * Minv[a][b] * J[j][b]; X[i][j] = s; }
  real Lx[6][6]; memcpy(Lx, X, sizeof Lx);
  int ok = chol_factor(&Lx[0][0], 6, 6); real det = 1;
  if (ok) for (int i = 0; i < 6; i++) det *= Lx[i][i] * Lx[i][i];
  if (ok && rabs(det) > R_(0.98e-3) && rabs(det) < R_(1.02e-3)) g_osc_knife_edge = 1;
  if (ok && rabs(det) >= R_(1e-3)) {
    for (int c = 0; c < 6; c++) { real col[6] = {0}; col[c] = 1; chol_solve(&Lx[0][0], 6, 6, col); for (int r = 0; r < 6; r++) Mx[r][c] = col[r]; }
  } else sym_pinv6(&X[0][0], R_(1e-4), &Mx[0][0]);
  real ut[6];
  for (int i = 0; i < 3; i++) ut[i] = ee_pos[i] - target[i];
  real qd[4], qc[4] = {ee_quat[0], -ee_quat[1], -ee_quat[2], -ee_quat[3]}, qr[4];
  euler_rxyz_to_quat(qd, target[3], target[4], target[5]);
  { real n = rsqrt_(qd[0] * qd[0] + qd[1] * qd[1] + qd[2] * qd[2] + qd[3] * qd[3]); for (int i = 0; i < 4; i++) qd[i] /= n; }
  quat_mul(qr, qd, qc);
  real sg = qr[0] > 0 ? R_(1) : (qr[0] < 0 ? R_(-1) : R_(0));
  if (rabs(qr[0]) < R_(1e-4)) g_osc_knife_edge = 1;   /* orientation error at 180 deg: the sign of the error vector is a coin flip */
  for (int i = 0; i < 3; i++) ut[3 + i] = -qr[1 + i] * sg;
  /* velocity limiting */
  real sat_xyz = vmax0 / kp * kv, sat_abg = vmax1 / ko * kv, nx = v3norm(ut), na = v3norm(ut + 3), sc[6] = {1, 1, 1, 1, 1, 1};
  if (nx > sat_xyz) for (int i = 0; i < 3; i++) sc[i] *= sat_xyz / nx;
  if (na > sat_abg) for (int i = 3; i < 6; i++) sc[i] *= sat_abg / na;
  for (int i = 0; i < 6; i++) ut[i] = kv * sc[i] * ((i < 3 ? kp : ko) / kv) * ut[i];
  real f[6];
  for (int i = 0; i < 6; i++) { real s = 0; for (int j = 0; j < 6; j++) s += Mx[i][j] * ut[j]; f[i] = s; }
  for (int i = 0; i < 6; i++) {
    real s = 0, t = 0;
    for (int j = 0; j < 6; j++) { s += M[i][j] * dq[j]; t += J[j][i] * f[j]; }
    u[i] = -kv * s - t + bias[i];
  }
}

/* ------------------------------------------------------------------------------------------- env logic */
static real uni(const jo_env *e, int stream, int k) {
  uint32_t o[4];
  jo_philox(e->seed, (uint32_t)(k >> 2), (uint32_t)stream, (uint32_t)e->env_id, e->rng_counter, o);
  return R_((double)(o[k & 3] >> 8) * (1.0 / 16777216.0));
}
static real uniab(const jo_env *e, int stream, int k, double a, double b) { return R_(a) + R_(b - a) * uni(e, stream, k); }

static void get_gripper_pose(jo_env *e) { /* util:309 / 667-673 */
  v3copy(e->gripper_pose, e->ee_pos);
  quat_to_euler_rxyz(e->gripper_pose + 3, e->ee_quat);
}
static int get_touch(const jo_env *e) { /* util:470-490 */
  real ta[20];
  for (int i = 0; i < 19; i++) ta[i] = e->sensordata[i];
  ta[19] = e->sensordata[19];
  const real th = R_(0.001);
  int thumb = 0, index = 0, pinky = 0, inner = 0, outer = 0;
  for (int i = 1; i < 5; i++) if (ta[i] > th) thumb = 1;
  for (int i = 5; i < 9; i++) if (ta[i] > th) index = 1;
  for (int i = 9; i < 13; i++) if (ta[i] > th) pinky = 1;
  for (int i = 0; i < 13; i++) if (ta[i] > th) inner = 1;
  for (int i = 13; i < 20; i++) if (ta[i] > th) outer = 1;
  if ((thumb && index) || (thumb && pinky)) return 3;
  if (inner) return 1;
  if (outer) return 2;
  return 0;
}
static void rulebased_subgoal(jo_env *e, real *pos, real *ori) { /* util:273-300 */
  if (e->task == JO_TASK_PLACING) { v3copy(pos, e->dest_goal); ori[0] = 0; ori[1] = PI_ / 2; ori[2] = 0; return; }
  real D[3], n;
  v3sub(D, e->gripper_pose, e->obj_goal); n = v3norm(D);
  for (int i = 0; i < 3; i++) pos[i] = D[i] / n * R_(0.12) + e->obj_goal[i] + (uni(e, 1, i) - R_(0.5)) / 25;
  if (pos[2] < R_(0.1898) + R_(0.1)) pos[2] = R_(0.1898) + R_(0.1);
  if (pos[1] > e->obj_xpos[1] - R_(0.15)) pos[1] = e->obj_xpos[1] - R_(0.15);
  real x = -D[0] / n, y = -D[1] / n, z = -D[2] / n;
  real qp = rsqrt_(R_(2)) / 2, qv[4] = {qp, qp * x, qp * y, qp * z}, qd[4] = {rsqrt_(R_(3)) / 2, R_(0.5) * x, R_(0.5) * y, R_(0.5) * z}, rv[3], rd[3], rr[3];
  quat_to_euler_rxyz(rv, qv); quat_to_euler_rxyz(rd, qd);
  v3cross(rr, rv, rd);
  rr[0] -= PI_ / 2;
  for (int i = 0; i < 3; i++) ori[i] = rr[i] + (uni(e, 1, 3 + i) - R_(0.5)) / 10;
}
/* env_mujoco_util_trans.py:848-854: global_to_local(prop) = (Rx Ry Rz)^T (prop - base_position), applied AS WRITTEN to positions and to
   Euler triples alike (the 'orientation' branch is the same expression: an Euler triple minus the base translation, rotated) */
static void global_to_local(const jo_env *e, const real *prop, real *out) {
  real d[3]; v3sub(d, prop, e->base_link_pos);
  get_rotation(out, e->base_link_eul[0], e->base_link_eul[1], e->base_link_eul[2], d, 1);
}
/* env_mujoco_util_trans.py:856-862: local_to_global(prop) = (Rx Ry Rz) prop + base_position */
static void local_to_global(const jo_env *e, const real *prop, real *out) {
  real r[3]; get_rotation(r, e->base_link_eul[0], e->base_link_eul[1], e->base_link_eul[2], prop, 0);
  v3add(out, r, e->base_link_pos);
}
static void get_observation_trans(jo_env *e, float *obs) { /* env_mujoco_util_trans.py:295-313 (rulebased_subgoal branch) */
  real o[JO_OBS], sp[3], so[3], t[3];
  const real zero[3] = {0, 0, 0}, dori[3] = {0, PI_ / 2, 0};
  o[0] = (real)e->touch_index;
  global_to_local(e, e->gripper_pose, o + 1);
  global_to_local(e, e->gripper_pose + 3, t); for (int i = 0; i < 3; i++) o[4 + i] = t[i] / PI_;
  o[7] = (e->gripper_angle - R_(0.8)) / R_(0.2);
  global_to_local(e, e->obj_xpos, o + 8);
  global_to_local(e, zero, o + 11);
  global_to_local(e, e->dest_goal, o + 14);
  rulebased_subgoal(e, sp, so);
  global_to_local(e, sp, o + 17);
  global_to_local(e, so, t); for (int i = 0; i < 3; i++) o[20 + i] = t[i] / PI_;
  global_to_local(e, dori, o + 23);
  for (int i = 0; i < JO_OBS; i++) obs[i] = (float)o[i];
}
static void get_observation(jo_env *e, float *obs) { /* util:223-271 (+ _cis :267-328, _trans :274-) */
  get_gripper_pose(e);
  e->touch_index = get_touch(e);
  /* env_mujoco_util_cis.py:568-578: the variants' _get_touch uses `if` (not `elif`) for the inner-touch test, so the grasp class 3 is
     always overwritten by 1.  Reproduced as written. */
  if (e->variant != JO_VAR_BASE && e->touch_index == 3) e->touch_index = 1;
  if (e->variant == JO_VAR_TRANS && e->obs_mode == 1) { get_observation_trans(e, obs); return; }
  real o[JO_OBS], sp[3], so[3], base[3] = {0, 0, 0};
  if (e->variant != JO_VAR_BASE) v3copy(base, e->base_link_pos);
  o[0] = (real)e->touch_index;
  for (int i = 0; i < 3; i++) { o[1 + i] = e->gripper_pose[i] - base[i]; o[4 + i] = e->gripper_pose[3 + i] / PI_; }
  o[7] = (e->gripper_angle - R_(0.8)) / R_(0.2);
  for (int i = 0; i < 3; i++) { o[8 + i] = e->obj_xpos[i] - base[i]; o[11 + i] = 0; o[14 + i] = e->dest_goal[i] - ((e->variant != JO_VAR_BASE && e->obs_mode != 0) ? base[i] : 0); }
  if (e->obs_mode == 1) { rulebased_subgoal(e, sp, so); for (int i = 0; i < 3; i++) { o[17 + i] = sp[i] - base[i]; o[20 + i] = so[i] / PI_; } }
  else if (e->obs_mode == 2) { for (int i = 0; i < 3; i++) { o[17 + i] = e->gripper_pose[i]; o[20 + i] = e->gripper_pose[3 + i] / PI_; } }
  else { for (int i = 0; i < 3; i++) { o[17 + i] = e->reach_goal[i] - base[i]; o[20 + i] = e->reach_goal[3 + i] / PI_; } }
  o[23] = 0; o[24] = PI_ / 2; o[25] = 0;
  for (int i = 0; i < JO_OBS; i++) obs[i] = (float)o[i];
}
static real reach_ang_diff(const jo_env *e) { /* util:323-333 */
  real g[4], t[4], ge[3], te[3], d[3];
  euler_rxyz_to_quat(g, e->gripper_pose[3], e->gripper_pose[4], e->gripper_pose[5]);
  euler_rxyz_to_quat(t, e->reach_goal[3], e->reach_goal[4], e->reach_goal[5]);
  quat_to_euler_rxyz(ge, g); quat_to_euler_rxyz(te, t);
  v3sub(d, ge, te);
  real a = v3norm(d);
  if (a > PI_) a = 2 * PI_ - a;
  return a;
}
static real get_reward(jo_env *e) { /* util:312-446 */
  if (e->task == JO_TASK_REACHING) {
    real d[3]; v3sub(d, e->gripper_pose, e->reach_goal);
    real dist = v3norm(d), ang = reach_ang_diff(e);
    real r = 5 * rexp(-dist) / 2 + 2 * rexp(-ang / (PI_ / 6)) / (2 * (dist * 15 + 1));
    real w[3]; v3sub(w, e->ee_pos, e->variant == JO_VAR_BASE ? e->base_pos : e->base_link_pos); /* cis:42,409 */
    real wb = v3norm(w);
    if (wb < R_(0.15)) r -= R_(0.15) - wb;
    if (e->gripper_pose[2] < R_(0.1)) r -= R_(0.1) - e->gripper_pose[2];
    return R_(0.05) * r;
  }
  if (e->task == JO_TASK_PICKING && e->variant != JO_VAR_BASE && e->binary) return 0; /* cis:459,500 / trans:465 */
  if (e->task == JO_TASK_GRASPING || e->task == JO_TASK_PICKING) {
    real scale = e->task == JO_TASK_GRASPING ? R_(0.05) : R_(0.01);
    real ev[3], down[3] = {0, 0, -1}, up[3] = {0, 0, 1}, xyz[3], tv[3], dv[3];
    get_rotation(ev, e->gripper_pose[3], e->gripper_pose[4], e->gripper_pose[5], down, 0);
    v3sub(xyz, e->obj_xpos, e->gripper_pose);
    real od = v3norm(xyz), x = xyz[0] / od, z = xyz[2] / od;
    real cz = z < -1 ? -1 : (z > 1 ? 1 : z), cx = -x / rsqrt_(1 - z * z > R_(1e-30) ? 1 - z * z : R_(1e-30));
    if (cx < -1) cx = -1; if (cx > 1) cx = 1; /* rounding guard: the reference would return NaN here (measure-zero set) */
    real pitch = -racos(cz), yaw = -racos(cx);
    get_rotation(tv, 0, pitch, yaw, up, 1);
    tv[2] *= -1;
    v3sub(dv, ev, tv);
    real ad = v3norm(dv);
    real r = 5 * rexp(-od / R_(0.2)) / 2 + 2 * rexp(-ad / (PI_ / 6)) / 2 / (od * 15 + 1);
    if (e->touch_index == 1) r += 5 * R_(0.5) * R_(0.3);
    else if (e->touch_index == 2) r -= 5 * R_(0.5) * R_(0.3);
    else if (e->touch_index == 3) r += 5 * R_(0.5);
    r += 100 * (e->obj_xpos[2] - R_(0.1898));
    return r * scale;
  }
  return 0;
}
/* util:492-600; returns done, fills bonus / wb / succ */
static int terminal_inspection(jo_env *e, real *bonus, real *wb, int *succ) {
  e->num_episodes += 1;
  real d[3]; *succ = 0; *bonus = 0;
  v3sub(d, e->ee_pos, e->variant == JO_VAR_BASE ? e->base_pos : e->base_link_pos); *wb = v3norm(d); /* util:35,499; cis:42,590 */
  const real *op = e->obj_xpos;
  real dd[2] = {e->dest_goal[0] - op[0], e->dest_goal[1] - op[1]}, dest_diff = rsqrt_(dd[0] * dd[0] + dd[1] * dd[1]);
  real q2 = e->qpos[2];
  if (PI_ - R_(0.1) < q2 && q2 < PI_ + R_(0.1)) { *bonus = -1; return 1; }
  real oz = R_(0.1898);
  switch (e->task) {
  case JO_TASK_REACHING: {
    v3sub(d, e->gripper_pose, e->reach_goal);
    if (v3norm(d) < R_(0.025) && reach_ang_diff(e) < PI_ / 6) { *bonus = 200 - e->num_episodes * R_(0.1); *succ = 1; return 1; }
    return 0; }
  case JO_TASK_GRASPING: {
    v3sub(d, e->gripper_pose, op);
    if (v3norm(d) > R_(0.2)) { *bonus = -20; return 1; }
    if (op[2] > oz + R_(0.07) && (e->touch_index == 1 || e->touch_index == 3)) { *bonus = 200 - e->num_episodes * R_(0.1); *succ = 1; return 1; }
    if (op[2] < R_(0.1)) { *bonus = -20; return 1; }
    return 0; }
  case JO_TASK_PICKING: {
    const int bin = e->variant != JO_VAR_BASE && e->binary; /* cis:637-660 / trans:643-: sparse bonuses 10 / -1 */
    if (op[2] > oz + R_(0.07) && (e->touch_index == 1 || e->touch_index == 3)) { *bonus = bin ? R_(10) : 200 - e->num_episodes * R_(0.1); *succ = 1; return 1; }
    if (op[2] < R_(0.1)) { *bonus = bin ? R_(-1) : R_(-20); return 1; }
    return 0; }
  case JO_TASK_PLACING: {
    real th = e->variant == JO_VAR_BASE ? R_(0.02) : R_(0.04);
    if (op[2] < R_(0.1)) { *bonus = -20; return 1; }
    if (dest_diff < th && e->touch_index == 0 && op[2] < R_(0.35)) { *bonus = 200 - e->num_episodes * R_(0.1); *succ = 1; return 1; }
    if (dest_diff > th && e->touch_index == 0 && op[2] < R_(0.20)) { *bonus = -20; return 1; }
    return 0; }
  case JO_TASK_CARRYING: return 1; /* util:549-550: (True, 0, wb) */
  case JO_TASK_RELEASING: { /* util:551-566; obj_velocity = |qvel[9:12]| (mujoco.py:212-215) */
    real ov = rsqrt_(e->qvel[9] * e->qvel[9] + e->qvel[10] * e->qvel[10] + e->qvel[11] * e->qvel[11]);
    if (op[2] < R_(0.1)) { *bonus = -20; return 1; }
    if (dest_diff < R_(0.04) && e->touch_index == 0 && op[2] < R_(0.35) && ov < R_(0.01)) { *bonus = 200 - e->num_episodes * R_(0.1); *succ = 1; return 1; }
    if (dest_diff > R_(0.04) && e->touch_index == 0 && op[2] < R_(0.20)) { *bonus = -20; return 1; }
    return 0; }
  case JO_TASK_PICKANDPLACE:
    if (op[2] > oz + R_(0.07) && (e->touch_index == 1 || e->touch_index == 3) && !e->picked) { e->picked = 1; *bonus = 20; return 0; }
    if (dest_diff < R_(0.04) && e->touch_index == 0 && op[2] < R_(0.35)) { *bonus = 180; *succ = 1; return 1; }
    if (op[2] < R_(0.1)) { *bonus = -20; return 1; }
    return 0;
  }
  return 0;
}
static void step_simulation(jo_env *e) { /* util:73-83 */
  real M6[6][6], u[6];
  for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) M6[i][j] = e->M[i][j];
  osc_generate(e->Jee, M6, e->bias, e->ee_pos, e->ee_quat, e->qvel, e->target, u);
  real g = e->action_in ? e->ramp[e->gripper_iter < 50 ? e->gripper_iter : 49] : e->gripper_angle;
  for (int i = 0; i < 6; i++) e->ctrl[i] = u[i];
  e->ctrl[6] = e->ctrl[7] = e->ctrl[8] = g;
  mj_step_(e);
  e->gripper_iter += 1;
}
static void set_obj(jo_env *e, const real *xyz, const real *quat) { /* mujoco.py:217-227 */
  for (int i = 0; i < 3; i++) e->qpos[9 + i] = xyz[i];
  for (int i = 0; i < 4; i++) e->qpos[12 + i] = quat[i];
  for (int i = 9; i < NV; i++) e->qvel[i] = 0;
  mj_forward(e);
}
static void take_action(jo_env *e, const real *a) { /* util:602-646 */
  e->action_in = 1; e->gripper_iter = 0;
  get_gripper_pose(e);
  /* base: a[:3]/25, a[3:6]/5 (util:621); _cis / _trans single arm: a[:3]/100, a[3:6]/20
   * (env_mujoco_util_cis.py:776, env_mujoco_util_trans.py:780) */
  const real sp = e->variant == JO_VAR_BASE ? R_(25) : R_(100), sa = e->variant == JO_VAR_BASE ? R_(5) : R_(20);
  for (int i = 0; i < 3; i++) { e->target[i] = e->gripper_pose[i] + a[i] / sp; e->target[3 + i] = e->gripper_pose[3 + i] + a[3 + i] / sa; }
  if (rabs(e->target[5]) > PI_) e->target[5] += -(e->target[5] > 0 ? R_(1) : R_(-1)) * 2 * PI_;
  real prev = e->gripper_angle, g = prev + a[6] / 10;
  if (g < R_(0.6)) g = R_(0.6); if (g > 1) g = 1;
  e->gripper_angle = g;
  real stepv = (g - prev) / 49;
  for (int i = 0; i < 50; i++) e->ramp[i] = prev + i * stepv;
  e->ramp[49] = g;
}

void jo_reset(jo_env *e, float *obs) { /* env_mujoco.py:99-102 -> util:92-174 */
  e->current_steps = 0;
  /* sim.reset() */
  for (int k = 0; k < NCH; k++) e->qpos[k] = e->ch_ref[k];
  for (int f = 0; f < 2; f++) for (int i = 0; i < 7; i++) e->qpos[9 + 7 * f + i] = e->free_qpos0[f][i];
  memset(e->qvel, 0, sizeof e->qvel); memset(e->qacc_warm, 0, sizeof e->qacc_warm); memset(e->ctrl, 0, sizeof e->ctrl);
  e->action_in = 0; e->picked = 0; e->gripper_angle = R_(0.6); e->gripper_iter = 0; e->touch_index = 0; e->num_episodes = 0;
  /* _create_init_angle, util:176-190 */
  real ia[6];
  if (e->task == JO_TASK_REACHING || e->task == JO_TASK_PICKING || e->task == JO_TASK_PICKANDPLACE) {
    ia[0] = uniab(e, 0, 0, 0.7, 2.5); ia[1] = uniab(e, 0, 1, 3.8, 4); ia[2] = uniab(e, 0, 2, 1, 1.7);
    ia[3] = uniab(e, 0, 3, 1.8, 2.5); ia[4] = uniab(e, 0, 4, 1, 2.5); ia[5] = uniab(e, 0, 5, 0.8, 2.3);
  } else if (e->task == JO_TASK_RELEASING) { /* util:186-189: nine values, the fingers start at 0.6 (mujoco.py:342-346) */
    ia[0] = uniab(e, 0, 0, 1.9, 2); ia[1] = uniab(e, 0, 1, 3.3, 3.6); ia[2] = uniab(e, 0, 2, 0.5, 0.8);
    ia[3] = uniab(e, 0, 3, 1.8, 2.5); ia[4] = uniab(e, 0, 4, 1.3, 2); ia[5] = uniab(e, 0, 5, -0.4, -0.9);
    e->qpos[6] = e->qpos[7] = e->qpos[8] = R_(0.6);
  } else {
    real c0 = uniab(e, 0, 0, 3 * 3.14159265358979323846 / 8, 3.14159265358979323846 / 2), c1 = uniab(e, 0, 1, 3.14159265358979323846 / 2, 5 * 3.14159265358979323846 / 8);
    ia[0] = uni(e, 0, 6) < R_(0.5) ? c0 : c1; ia[1] = R_(3.85); ia[2] = uniab(e, 0, 2, 1, 1.1);
    ia[3] = uniab(e, 0, 3, 2, 2.1); ia[4] = uniab(e, 0, 4, 0.8, 2.3); ia[5] = uniab(e, 0, 5, -1.2, -1.1);
  }
  for (int i = 0; i < 6; i++) { e->qpos[i] = ia[i]; e->qvel[i] = 0; }
  /* __sample_goal, util:192-221 (goal_buffer None) */
  real rg[3], xyz[3];
  rg[0] = uniab(e, 0, 7, 0.3, 0.42) * (uni(e, 0, 8) < R_(0.5) ? -1 : 1);
  rg[1] = uniab(e, 0, 9, 0.3, 0.42) * (uni(e, 0, 10) < R_(0.5) ? -1 : 1);
  rg[2] = uniab(e, 0, 11, 0.3, 0.5);
  v3sub(xyz, rg, e->variant == JO_VAR_BASE ? e->base_pos : e->base_link_pos);
  { real n = v3norm(xyz); for (int i = 0; i < 3; i++) xyz[i] /= n; }
  real sx = xyz[0] > 0 ? R_(1) : (xyz[0] < 0 ? R_(-1) : 0);
  real alpha = -(real)asin((double)(xyz[1] / rsqrt_(xyz[1] * xyz[1] + xyz[2] * xyz[2]))) * sx;
  real beta = racos(xyz[0] / v3norm(xyz)) * sx, gamma = uniab(e, 0, 12, -0.1, 0.1);
  v3copy(e->reach_goal, rg);
  /* np.float16 rounding of the orientation (util:207) */
  { real o3[3] = {alpha, beta, gamma};
    for (int i = 0; i < 3; i++) { /* round to 11 significant bits like IEEE half (values here are normal halfs) */
      double v = (double)o3[i]; if (v != 0) { int ex; double m = frexp(v, &ex); m = nearbyint(m * 2048.0) / 2048.0; v = ldexp(m, ex); }
      e->reach_goal[3 + i] = R_(v); } }
  e->obj_goal[0] = uniab(e, 0, 13, -0.1, 0.1); e->obj_goal[1] = R_(0.65) + uniab(e, 0, 14, -0.08, 0.02); e->obj_goal[2] = R_(0.1898);
  e->dest_goal[0] = R_(0.4) + uniab(e, 0, 15, -0.05, 0.05); e->dest_goal[1] = R_(0.3) + uniab(e, 0, 16, -0.05, 0.05); e->dest_goal[2] = R_(0.3468);
  if (e->variant == JO_VAR_TRANS) { real g[3]; local_to_global(e, e->obj_goal, g); v3copy(e->obj_goal, g); local_to_global(e, e->dest_goal, g); v3copy(e->dest_goal, g); } /* trans:265-270 */
  else if (e->variant != JO_VAR_BASE) for (int i = 0; i < 3; i++) { e->obj_goal[i] += e->base_link_pos[i]; e->dest_goal[i] += e->base_link_pos[i]; }
  if (e->variant == JO_VAR_BASE) { /* set_dest_xyz, mujoco.py:229-237 (qvel[16:] quirk kept) */
    e->qpos[16] = e->dest_goal[0]; e->qpos[17] = e->dest_goal[1];
    for (int i = 16; i < NV; i++) e->qvel[i] = 0;
  }
  mj_forward(e);
  const int held = e->task == JO_TASK_PLACING || e->task == JO_TASK_RELEASING || e->task == JO_TASK_CARRYING; /* util:106 */
  if (held) { /* util:106-117: object held in the hand for 150 settle substeps */
    real R6[9], p[3], t[3], eul[3], q[4], dp[3], off[3] = {R_(-0.04), 0, 0};
    memcpy(R6, e->xmat[6], sizeof R6);
    m3mulv(t, R6, e->ee_obj_off); v3add(p, e->xpos[6], t);
    quat_to_euler_rxyz(eul, e->ee_quat);
    euler_rxyz_to_quat(q, eul[0], eul[1], eul[2]);
    get_rotation(dp, eul[0], eul[1], eul[2], off, 0);
    v3add(p, p, dp);
    set_obj(e, p, q);
    get_gripper_pose(e); memcpy(e->target, e->gripper_pose, sizeof e->target);
    for (int i = 0; i < 150; i++) { step_simulation(e); set_obj(e, p, q); }
  } else if (e->task != JO_TASK_REACHING) {
    real zq[4] = {0, 0, 0, 0};
    set_obj(e, e->obj_goal, zq);
  }
  if (e->task == JO_TASK_REACHING) memcpy(e->target, e->reach_goal, sizeof e->target);
  else if (e->task == JO_TASK_GRASPING || e->task == JO_TASK_CARRYING) { /* util:123-167: two approach servo loops (unbounded in the reference; bounded here, JO_GRASP_RESET_MAX substeps each) */
    get_gripper_pose(e);
    real d[3]; v3sub(d, e->obj_goal, e->gripper_pose);
    real sx = d[0] > 0 ? R_(1) : (d[0] < 0 ? R_(-1) : R_(0));
    real ori[3] = {-rasin(d[1] / rsqrt_(d[1] * d[1] + d[2] * d[2])) * sx, racos(d[0] / v3norm(d)) * sx, uniab(e, 0, 17, -0.1, 0.1)};
    for (int i = 0; i < 3; i++) { double v = (double)ori[i]; if (v != 0) { int ex; double m = frexp(v, &ex); m = nearbyint(m * 2048.0) / 2048.0; v = ldexp(m, ex); } ori[i] = R_(v); } /* dtype=np.float16 */
    for (int i = 0; i < 3; i++) { e->target[i] = e->obj_goal[i]; e->target[3 + i] = ori[i]; }
    real qt[4]; euler_rxyz_to_quat(qt, e->reach_goal[3], e->reach_goal[4], e->reach_goal[5]);
    { real n = rsqrt_(qt[0] * qt[0] + qt[1] * qt[1] + qt[2] * qt[2] + qt[3] * qt[3]); for (int i = 0; i < 4; i++) qt[i] /= n; }
    for (int it = 0; it < JO_GRASP_RESET_MAX; it++) {
      for (int i = 9; i < NV; i++) e->qvel[i] = 0; /* stop_obj (mujoco.py:239-246) */
      mj_forward(e);
      step_simulation(e);
      get_gripper_pose(e);
      v3sub(d, e->gripper_pose, e->obj_goal);
      real qg[4]; euler_rxyz_to_quat(qg, e->gripper_pose[3], e->gripper_pose[4], e->gripper_pose[5]);
      { real n = rsqrt_(qg[0] * qg[0] + qg[1] * qg[1] + qg[2] * qg[2] + qg[3] * qg[3]); for (int i = 0; i < 4; i++) qg[i] /= n; }
      real ad = 0; for (int i = 0; i < 4; i++) ad += (qg[i] - qt[i]) * (qg[i] - qt[i]);
      if (v3norm(d) < R_(0.2)) { for (int i = 0; i < 3; i++) e->target[i] = e->gripper_pose[i]; break; }
      if (rsqrt_(ad) < PI_ / 6) break;
    }
    for (int it = 0; it < JO_GRASP_RESET_MAX; it++) {
      step_simulation(e);
      get_gripper_pose(e);
      for (int i = 0; i < 3; i++) e->target[i] = e->obj_goal[i];
      v3sub(d, e->gripper_pose, e->obj_goal);
      if (v3norm(d) < R_(0.15)) break;
    }
  }
  else if (!held) { get_gripper_pose(e); memcpy(e->target, e->gripper_pose, sizeof e->target); } /* util:65 (ctor) */
  get_observation(e, obs);
  e->rng_counter += 1;
}

void jo_step(jo_env *e, const double *action, float *obs, double *reward, int *done, int *succ, double *wb) { /* env_mujoco.py:116-139 */
  real a[7];
  for (int i = 0; i < 7; i++) { double v = action[i]; if (v < -1) v = -1; if (v > 1) v = 1; a[i] = R_(v); }
  if (e->task == JO_TASK_REACHING) a[6] = 0;
  take_action(e, a);
  /* len(a) == 6 (util:625-626): gripper_angle = 0.6 for the observation, but gripper_angle_array is never refilled and
   * stays np.zeros (util:37), so once action_in is set the finger servos are commanded 0.0 (util:76-79) */
  if (e->task == JO_TASK_REACHING) { e->gripper_angle = R_(0.6); for (int i = 0; i < 50; i++) e->ramp[i] = 0; }
  for (int i = 0; i < 50; i++) { step_simulation(e); }
  get_observation(e, obs);
  real r = get_reward(e), bonus, w; int s, d;
  e->current_steps += 1; /* env_mujoco.py:144-150 */
  if (e->current_steps < e->task_max_steps) d = terminal_inspection(e, &bonus, &w, &s);
  else { d = 1; bonus = -10; w = 0; s = 0; }
  *reward = (double)(r + bonus); *done = d; *succ = s; *wb = (double)w;
  e->rng_counter += 1;
}

/* ------------------------------------------------------------------------------------------- API plumbing */
jo_env *jo_create(const double *blob, int nblob, int task, int variant, uint64_t seed, int env_id) {
  if (nblob != JM_TOTAL) return 0;
  jo_env *e = (jo_env *)calloc(1, sizeof(jo_env));
  load_model(e, blob);
  e->task = task; e->variant = variant & 15; e->obs_mode = (variant >> 4) & 15; e->binary = (variant >> 8) & 1; e->seed = seed; e->env_id = env_id;
  e->task_max_steps = (task == JO_TASK_REACHING || task == JO_TASK_GRASPING || task == JO_TASK_RELEASING || task == JO_TASK_CARRYING) ? 500 : ((task == JO_TASK_PICKING || task == JO_TASK_PLACING) ? 700 : 1200);
  float obs[JO_OBS];
  jo_reset(e, obs);
  e->rng_counter = 0;
  return e;
}
void jo_destroy(jo_env *e) { free(e); }
int jo_sizeof_real(void) { return (int)sizeof(real); }

void jo_get_state(const jo_env *e, double *qpos, double *qvel, double *aux) {
  for (int i = 0; i < NQ; i++) qpos[i] = e->qpos[i];
  for (int i = 0; i < NV; i++) qvel[i] = e->qvel[i];
  if (!aux) return;
  memset(aux, 0, sizeof(double) * JO_AUX);
  for (int i = 0; i < 6; i++) aux[i] = e->target[i];
  aux[6] = e->gripper_angle; aux[7] = e->action_in; aux[8] = e->picked; aux[9] = e->current_steps; aux[10] = e->num_episodes;
  for (int i = 0; i < 3; i++) { aux[11 + i] = e->obj_goal[i]; aux[14 + i] = e->dest_goal[i]; }
  for (int i = 0; i < 6; i++) aux[17 + i] = e->reach_goal[i];
  aux[23] = e->rng_counter; aux[24] = e->touch_index;
  for (int i = 0; i < NV; i++) aux[25 + i] = e->qacc_warm[i];
  for (int i = 0; i < 9; i++) { aux[46 + i] = e->dq[i]; aux[55 + i] = e->dv[i]; }
}
/* recompute the mjData quantities the controller reads, at the lagged chain state */
static void refresh_lagged(jo_env *e, const real *lq, const real *lv) {
  real sq[NQ], sv[NV];
  memcpy(sq, e->qpos, sizeof sq); memcpy(sv, e->qvel, sizeof sv);
  for (int i = 0; i < 9; i++) { e->qpos[i] = lq[i]; e->qvel[i] = lv[i]; }
  real sw[NV]; memcpy(sw, e->qacc_warm, sizeof sw);
  mj_forward(e);
  memcpy(e->qpos, sq, sizeof sq); memcpy(e->qvel, sv, sizeof sv); memcpy(e->qacc_warm, sw, sizeof sw);
  for (int i = 9; i < NQ; i++) e->dq[i] = sq[i];
}
void jo_set_state(jo_env *e, const double *qpos, const double *qvel, const double *aux) {
  for (int i = 0; i < NQ; i++) e->qpos[i] = R_(qpos[i]);
  for (int i = 0; i < NV; i++) e->qvel[i] = R_(qvel[i]);
  real lq[9], lv[9];
  for (int i = 0; i < 9; i++) { lq[i] = e->qpos[i]; lv[i] = e->qvel[i]; }
  if (aux) {
    for (int i = 0; i < 6; i++) e->target[i] = R_(aux[i]);
    e->gripper_angle = R_(aux[6]); e->action_in = (int)aux[7]; e->picked = (int)aux[8]; e->current_steps = (int)aux[9]; e->num_episodes = (int)aux[10];
    for (int i = 0; i < 3; i++) { e->obj_goal[i] = R_(aux[11 + i]); e->dest_goal[i] = R_(aux[14 + i]); }
    for (int i = 0; i < 6; i++) e->reach_goal[i] = R_(aux[17 + i]);
    e->rng_counter = (uint32_t)aux[23]; e->touch_index = (int)aux[24];
    for (int i = 0; i < NV; i++) e->qacc_warm[i] = R_(aux[25 + i]);
    for (int i = 0; i < 9; i++) { lq[i] = R_(aux[46 + i]); lv[i] = R_(aux[55 + i]); }
  }
  refresh_lagged(e, lq, lv);
}
void jo_mj_step(jo_env *e, const double *ctrl, int *ncon, double *sensordata) {
  for (int i = 0; i < JO_NU; i++) e->ctrl[i] = R_(ctrl[i]);
  mj_step_(e);
  if (ncon) *ncon = e->ncon;
  if (sensordata) for (int i = 0; i < 20; i++) sensordata[i] = e->sensordata[i];
}
void jo_forward(jo_env *e, const double *ctrl, double *M, double *bias, double *J, double *ee_pos, double *ee_quat, double *qacc, int *ncon, double *sensordata) {
  if (ctrl) for (int i = 0; i < JO_NU; i++) e->ctrl[i] = R_(ctrl[i]);
  mj_forward(e);
  if (M) for (int i = 0; i < NV; i++) for (int j = 0; j < NV; j++) M[NV * i + j] = e->M[i][j];
  if (bias) for (int i = 0; i < NV; i++) bias[i] = e->bias[i];
  if (J) for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) J[6 * i + j] = e->Jee[i][j];
  if (ee_pos) for (int i = 0; i < 3; i++) ee_pos[i] = e->ee_pos[i];
  if (ee_quat) for (int i = 0; i < 4; i++) ee_quat[i] = e->ee_quat[i];
  if (qacc) for (int i = 0; i < NV; i++) qacc[i] = e->qacc[i];
  if (ncon) *ncon = e->ncon;
  if (sensordata) for (int i = 0; i < 20; i++) sensordata[i] = e->sensordata[i];
}
/* world primitives the headless camera draws (restates mujoco_jaco_b200/csrc/jaco_render.cuh render_prims; util:302-307 is an OpenGL
 * call in the reference): per primitive pos[3], R[9] row-major, size[3], type; the model's geoms, then two capsules (radius 0.04)
 * between the origins of bodies 1-2 and 2-3 for the upper-arm links that carry no collision geom.  Evaluated at the current qpos. */
int jo_render_prims(jo_env *e, double *out, int maxprim) {
  kinematics(e);
  int n = 0;
  for (int g = 0; g < JM_NGEOM && n < maxprim; g++, n++) {
    real p[3], R[9]; geom_world(e, g, p, R);
    double *o = out + 16 * n;
    for (int i = 0; i < 3; i++) { o[i] = p[i]; o[12 + i] = e->geoms[g].size[i]; }
    for (int i = 0; i < 9; i++) o[3 + i] = R[i];
    o[15] = e->geoms[g].type;
  }
  for (int b = 0; b < 2 && n < maxprim; b++, n++) {
    double *o = out + 16 * n, d[3], len = 0;
    for (int i = 0; i < 3; i++) { d[i] = (double)e->xpos[b + 2][i] - (double)e->xpos[b + 1][i]; len += d[i] * d[i]; o[i] = 0.5 * ((double)e->xpos[b + 1][i] + (double)e->xpos[b + 2][i]); }
    len = sqrt(len);
    if (len < 1e-6) { d[0] = 0; d[1] = 0; d[2] = 1; len = 0; } else for (int i = 0; i < 3; i++) d[i] /= len;
    double a[3] = {1, 0, 0}; if (fabs(d[0]) > 0.9) { a[0] = 0; a[1] = 1; }
    double x[3] = {a[1] * d[2] - a[2] * d[1], a[2] * d[0] - a[0] * d[2], a[0] * d[1] - a[1] * d[0]}, nx = sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    for (int i = 0; i < 3; i++) x[i] /= nx;
    double y[3] = {d[1] * x[2] - d[2] * x[1], d[2] * x[0] - d[0] * x[2], d[0] * x[1] - d[1] * x[0]};
    for (int i = 0; i < 3; i++) { o[3 + 3 * i] = x[i]; o[4 + 3 * i] = y[i]; o[5 + 3 * i] = d[i]; }
    o[12] = 0.04; o[13] = 0.5 * len; o[14] = 0; o[15] = 2;
  }
  return n;
}
/* the constraint problem of the last forward, for independent optimality checks (tests/test_oracle_independent.py):
 * J [nefc][NV], D, aref, qacc_smooth; returns nefc */
int jo_get_efc(const jo_env *e, double *J, double *D, double *aref, double *qacc_smooth, int maxrows) {
  int n = e->nefc < maxrows ? e->nefc : maxrows;
  for (int r = 0; r < n; r++) { for (int i = 0; i < NV; i++) J[r * NV + i] = e->efc_J[r][i]; D[r] = e->efc_D[r]; aref[r] = e->efc_aref[r]; }
  for (int i = 0; i < NV; i++) qacc_smooth[i] = e->qacc_smooth[i];
  return e->nefc;
}
int jo_get_contacts(const jo_env *e, double *out, int maxcon) {
  int n = e->ncon < maxcon ? e->ncon : maxcon;
  for (int c = 0; c < n; c++) {
    const contact_t *k = &e->con[c]; double *o = out + 10 * c;
    for (int i = 0; i < 3; i++) { o[i] = k->pos[i]; o[3 + i] = k->frame[i]; }
    o[6] = k->dist; o[7] = k->g1; o[8] = k->g2; o[9] = k->nforce;
  }
  return e->ncon;
}
void jo_osc(const jo_env *e, const double *J, const double *M, const double *bias, const double *ee_pos, const double *ee_quat, const double *dq, const double *target, double *u) {
  (void)e;
  real J_[6][6], M_[6][6], b_[6], p_[3], q_[4], d_[6], t_[6], u_[6];
  for (int i = 0; i < 6; i++) { for (int j = 0; j < 6; j++) { J_[i][j] = R_(J[6 * i + j]); M_[i][j] = R_(M[6 * i + j]); } b_[i] = R_(bias[i]); d_[i] = R_(dq[i]); t_[i] = R_(target[i]); }
  for (int i = 0; i < 3; i++) p_[i] = R_(ee_pos[i]);
  for (int i = 0; i < 4; i++) q_[i] = R_(ee_quat[i]);
  osc_generate(J_, M_, b_, p_, q_, d_, t_, u_);
  for (int i = 0; i < 6; i++) u[i] = u_[i];
}
/* 1 when, in the last forward, some contact's touch-sensor attribution flips if the sensor site is resized by +-2 um */
/* 1 when an OSC call since the last jo_osc_knife_edge_clear sat on the pinv branch threshold / rcond cut-off (process-wide flag) */
int jo_osc_knife_edge(void) { return g_osc_knife_edge; }
void jo_osc_knife_edge_clear(void) { g_osc_knife_edge = 0; }
int jo_touch_knife_edge(const jo_env *e) { return e->touch_knife_edge | (e->probe ? e->contact_knife_edge : 0); }
/* enables the (slower) contact-margin probe evaluated by every forward */
void jo_set_probe(jo_env *e, int on) { e->probe = on; e->margin_shift = 0; e->contact_knife_edge = 0; }
void jo_solver_stats(const jo_env *e, int *iters, double *gradnorm, int *nefc) {
  if (iters) *iters = e->solver_iters; if (gradnorm) *gradnorm = e->solver_grad; if (nefc) *nefc = e->nefc;
}
