// jaco_fast.cuh -- the register-resident fast path of the env step (one thread = one environment).
//
// Same arithmetic contract as jaco_sim.cuh (reference env_script/env_mujoco.py:116-150 and the MuJoCo-2.0 /
// abr_control semantics restated in SURVEY.md App. A), specialised to the regime a picking roll-out is in for
// ~99 % of its physics substeps: no arm contact, no active joint limit, the object and the destination box each
// touching at most one static scene geom through <= 4 coplanar contacts.  In that regime
//   * the 9-DoF chain is unconstrained (qacc = qacc_smooth) and needs one pass of spatial-vector dynamics,
//   * each free box is an independent 6-DoF convex problem solved by Newton with a 6x6 Hessian in registers.
// A thread that leaves the regime returns FAST_BAIL *without touching the state*; the host side then advances
// that environment with the general path (jaco_sim.cuh) from the same pre-step state, so results never depend
// on which path ran beyond fp32 rounding.  Everything here is __host__ __device__ so tests/harness can run the
// identical source on the CPU in the parity tests.
#pragma once
#include "jaco_sim.cuh"
#include "jaco_coop.cuh"

#define FAST_OK 0
#define FAST_BAIL 1
#define FAST_CLEAR_EPS 1e-4f   // slack (m) of the conservative "no arm contact possible" tests

// bail reasons (bit mask, statistics only)
enum { BAIL_LIMIT = 1, BAIL_HAND = 2, BAIL_FOREARM = 4, BAIL_OBJ = 8, BAIL_DEST = 16, BAIL_OSC = 32, BAIL_NAN = 64 };

#if defined(JACO_COLLISION_STATS) && !defined(__CUDACC__)
static long long g_fast_wstats[16];   // CPU harness only: workload mix of the hybrid step (scripts/workload_mix.py)
#define FAST_WSTAT(k) g_fast_wstats[k]++
#define FAST_WSTAT_ADD(k, v) g_fast_wstats[k] += (v)
#else
#define FAST_WSTAT(k)
#define FAST_WSTAT_ADD(k, v)
#endif

struct FastSim {
  float qpos[JACO_NQ], qvel[JACO_NV], warm[JACO_NV];
  float u[6];                 // OSC torque for the coming substep (built from the previous forward: SURVEY 3.1 lag)
  float eepos[3], eeq[4];     // EE pose of the last forward
  float objp[3];              // object xpos of the last forward (what obs / reward / terminal read)
  float tq[4];                // target orientation as a unit quaternion (constant during one env step)
  // env logic: same member names as JacoSim so the templated logic of jaco_sim.cuh applies unchanged
  float target[6], grip, ramp0, ramp1;
  int action_in, picked, touch_index, num_episodes, current_steps;
  uint32_t rng_counter;
  float obj_goal[3], dest_goal[3], reach_goal[6], gripper_pose[6];
  float lagq[9], lagv[9], accum_rew;
  int touch_now;              // touch class of the last forward (set by the cooperative path; 0 without arm contact)
};
JHD const float* obj_pos(const FastSim& s) { return s.objp; }
JHD int get_touch(const FastSim& s) { return s.touch_now; }  // 0 unless the cooperative path found pad contacts

// ------------------------------------------------------------------------------------------- chain pass
// Outputs of one pass of the chain dynamics at (q, v): inertia blocks, bias, EE Jacobian / pose.
struct ChainOut {
  float A[21];        // arm 6x6 block of M, packed lower
  float B[3][6];      // finger-arm coupling rows M[6+f][0..5]
  float Dg[3];        // finger diagonals M[6+f][6+f]
  float bias[9];      // qfrc_bias (gravity + Coriolis), mj_rne(flg_acc = 0)
  float J[6][6];      // EE Jacobian (rows: lin xyz, ang xyz; columns: arm joints)
  float eepos[3], eeq[4];
  int near;           // BAIL_* bits of the proximity tests that could not prove "no arm contact"
  float bp[7][3], bRm[7][9];  // world pose of bodies 3..9 (the bodies that carry collision geoms); consumed only by the cooperative path
};
struct SceneBoxes {   // world pose of the three scene boxes the arm can touch (geoms 1 holder, 2 object, 3 dest)
  float hp[3], hR[9], hh[3];
  float op[3], oR[9], oh[3];
  float dp[3], dR[9], dh[3];
};

JHD bool sphere_clear_of_box(const float* c, float r, const float* bp, const float* bR, const float* h) {
  float rr = r + FAST_CLEAR_EPS;
  return point_box_dist2(c, bp, bR, h) > rr * rr;
}

// spatial inertia (m, h = m*(com - O), I about O: xx yy zz xy xz yz) applied to a motion vector (wa ; wl)
JHD void sinertia_mul(float m, const float* h, const float* I, const float* wa, const float* wl, float* outA, float* outL) {
  float t[3], c[3];
  s3mulv(t, I, wa); v3cross(c, h, wl); v3add(outA, t, c);
  v3cross(c, wa, h); outL[0] = m * wl[0] + c[0]; outL[1] = m * wl[1] + c[1]; outL[2] = m * wl[2] + c[2];
}

// One body of the forward sweep.  In: body pose (R, p), world joint axis a, joint rate qd, parent spatial velocity
// (w, vl) / acceleration (aa, al) about the fixed reference point O -- updated in place to this body's.
// Out: motion axis lin part s, spatial inertia about O (m from the model; h, IO), net spatial force (fa ; fl).
JHD void chain_body(int b, const float* R, const float* p, const float* a, float qd, const float* O,
                    float* w, float* vl, float* aa, float* al, float* s, float* h, float* IO, float* fa, float* fl) {
  const JacoModel& m = MDL;
  float t[3], d[3], r[3];
  const float ms = m.mass[b];
  m3mulv(t, R, m.com[b]); v3add(d, p, t); v3sub(d, d, O);          // d = com - O
  {  // world inertia about the COM (R I R^T), then shifted to O
    const float* I = m.inertia[b];
    float S[9] = {I[0], I[3], I[4], I[3], I[1], I[5], I[4], I[5], I[2]}, T[9];
    m3mul(T, R, S);
    float dd = v3dot(d, d);
    IO[0] = T[0] * R[0] + T[1] * R[1] + T[2] * R[2] + ms * (dd - d[0] * d[0]);
    IO[1] = T[3] * R[3] + T[4] * R[4] + T[5] * R[5] + ms * (dd - d[1] * d[1]);
    IO[2] = T[6] * R[6] + T[7] * R[7] + T[8] * R[8] + ms * (dd - d[2] * d[2]);
    IO[3] = T[0] * R[3] + T[1] * R[4] + T[2] * R[5] - ms * d[0] * d[1];
    IO[4] = T[0] * R[6] + T[1] * R[7] + T[2] * R[8] - ms * d[0] * d[2];
    IO[5] = T[3] * R[6] + T[4] * R[7] + T[5] * R[8] - ms * d[1] * d[2];
  }
  h[0] = ms * d[0]; h[1] = ms * d[1]; h[2] = ms * d[2];
  v3sub(r, p, O); v3cross(s, r, a);                                // s = (p - O) x a
  // velocity-product acceleration (parent velocity x S) * qd, then this body's velocity
  float ca[3], cl[3];
  v3cross(ca, w, a); v3cross(cl, w, s); v3cross(t, vl, a); v3add(cl, cl, t);
  v3addscl(w, w, a, qd); v3addscl(vl, vl, s, qd);
  v3addscl(aa, aa, ca, qd); v3addscl(al, al, cl, qd);
  // f = I a + v x* (I v)
  float L[3], P[3], fa0[3], fl0[3];
  sinertia_mul(ms, h, IO, w, vl, L, P);
  sinertia_mul(ms, h, IO, aa, al, fa0, fl0);
  v3cross(t, w, L); v3add(fa, fa0, t); v3cross(t, vl, P); v3add(fa, fa, t);
  v3cross(t, w, P); v3add(fl, fl0, t);
}

// joint transform of chain body k: parent quaternion Qp -> body quaternion q and rotation matrix R (mj_kinematics)
JHD void chain_joint(int k, const float* Qp, float ang, float* q, float* R) {
  const JacoModel& m = MDL;
  float qb[4], qj[4], sh, chh;
  quat_mul(qb, Qp, m.ch_quat[k]);
  sincosf(0.5f * (ang - m.ch_ref[k]), &sh, &chh);
  qj[0] = chh; qj[1] = m.ch_axis[k][0] * sh; qj[2] = m.ch_axis[k][1] * sh; qj[3] = m.ch_axis[k][2] * sh;
  quat_mul(q, qb, qj);
  quat2mat(R, q);
}

// capsule geom g (on a body at pose R, p) against floor / holder / dest: true when contact is provably impossible
JHD bool capsule_clear(int g, const float* R, const float* p, const SceneBoxes& sb) {
  const JacoGeom& G = MDL.geoms[g];
  float c[3], t[3], ax[3], gax[3] = {G.R[2], G.R[5], G.R[8]};
  m3mulv(t, R, G.pos); v3add(c, p, t); m3mulv(ax, R, gax);
  float zmin = c[2] - fabsf(ax[2]) * G.size[1] - G.size[0];
  bool ok = zmin > FAST_CLEAR_EPS;
  ok = ok && sphere_clear_of_box(c, G.rbound, sb.hp, sb.hR, sb.hh);
  ok = ok && sphere_clear_of_box(c, G.rbound, sb.dp, sb.dR, sb.dh);
  return ok;
}

// bounding sphere of the collision geoms on hand-cluster body 6 + j against floor / holder / object / dest
JHD bool body_sphere_clear(int j, const float* R, const float* p, const SceneBoxes& sb) {
  const float* S = MDL.body_sphere[j];
  float c[3], t[3];
  m3mulv(t, R, S); v3add(c, p, t);
  bool ok = c[2] > S[3] + FAST_CLEAR_EPS;
  ok = ok && sphere_clear_of_box(c, S[3], sb.hp, sb.hR, sb.hh);
  ok = ok && sphere_clear_of_box(c, S[3], sb.op, sb.oR, sb.oh);
  ok = ok && sphere_clear_of_box(c, S[3], sb.dp, sb.dR, sb.dh);
  return ok;
}

// DYN = false: kinematics only (eepos, eeq).  O: fixed world reference point of the spatial quantities (any point;
// a point near the wrist keeps the fp32 cancellation in the distal inertias small).
template <bool DYN>
JHD void fast_chain(const float* q, const float* v, const float* O, const SceneBoxes* sb, ChainOut& o) {
  const JacoModel& m = MDL;
  float Qp[4] = {1, 0, 0, 0}, Rp[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, pp[3] = {0, 0, 0};
  float w[3] = {0, 0, 0}, vl[3] = {0, 0, 0}, aa[3] = {0, 0, 0}, al[3] = {0, 0, -m.gravz};
  float Sa[6][3], Sl[6][3], Ih[6][3], II[6][6], Im[6], Fa[6][3], Fl[6][3];
  o.near = 0;
#pragma unroll
  for (int k = 0; k < 6; k++) {
    float p[3], t[3], qn[4], R[9];
    m3mulv(t, Rp, m.ch_pos[k]); v3add(p, pp, t);
    chain_joint(k, Qp, q[k], qn, R);
    if (DYN) {
      m3mulv(Sa[k], R, m.ch_axis[k]);
      chain_body(k + 1, R, p, Sa[k], v[k], O, w, vl, aa, al, Sl[k], Ih[k], II[k], Fa[k], Fl[k]);
      Im[k] = m.mass[k + 1];
      if (sb && k >= 2 && k <= 4) { if (!capsule_clear(28 + k, R, p, *sb)) o.near |= BAIL_FOREARM; }  // geoms 30,31,32 on bodies 3,4,5
      if (sb && k >= 2) {
        v3copy(o.bp[k - 2], p);
#pragma unroll
        for (int i = 0; i < 9; i++) o.bRm[k - 2][i] = R[i];
      }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) Qp[i] = qn[i];
#pragma unroll
    for (int i = 0; i < 9; i++) Rp[i] = R[i];
    v3copy(pp, p);
  }
  // body 6 (hand): EE pose, hand bounding sphere
  { float t[3]; m3mulv(t, Rp, m.ee_off); v3add(o.eepos, pp, t); }
#pragma unroll
  for (int i = 0; i < 4; i++) o.eeq[i] = Qp[i];
  if (!DYN) return;
  bool fine = false;   // coarse hand sphere could not prove clearance: test the four per-body spheres
  if (sb) {
    float hs[3], t[3]; const float hr = m.hand_sphere[3];
    m3mulv(t, Rp, m.hand_sphere); v3add(hs, pp, t);
    bool ok = hs[2] > hr + FAST_CLEAR_EPS;
    ok = ok && sphere_clear_of_box(hs, hr, sb->hp, sb->hR, sb->hh);
    ok = ok && sphere_clear_of_box(hs, hr, sb->op, sb->oR, sb->oh);
    ok = ok && sphere_clear_of_box(hs, hr, sb->dp, sb->dR, sb->dh);
    fine = !ok;
    if (fine && !body_sphere_clear(0, Rp, pp, *sb)) o.near |= BAIL_HAND;
  }
  // fingers: leaves hanging off body 6; folded into the hand's composite right away
#pragma unroll
  for (int f = 0; f < 3; f++) {
    const int k = 6 + f;
    float p[3], t[3], qn[4], R[9], a[3], s[3], h[3], IO[6], fa[3], fl[3];
    float wf[3], vf[3], af[3], lf[3];
    v3copy(wf, w); v3copy(vf, vl); v3copy(af, aa); v3copy(lf, al);
    m3mulv(t, Rp, m.ch_pos[k]); v3add(p, pp, t);
    chain_joint(k, Qp, q[k], qn, R);
    if (sb) {
      v3copy(o.bp[4 + f], p);
#pragma unroll
      for (int i = 0; i < 9; i++) o.bRm[4 + f][i] = R[i];
    }
    if (fine && !body_sphere_clear(1 + f, R, p, *sb)) o.near |= BAIL_HAND;
    m3mulv(a, R, m.ch_axis[k]);
    chain_body(k + 1, R, p, a, v[k], O, wf, vf, af, lf, s, h, IO, fa, fl);
    o.bias[k] = v3dot(a, fa) + v3dot(s, fl);
    float FA[3], FL[3];
    sinertia_mul(m.mass[k + 1], h, IO, a, s, FA, FL);
    o.Dg[f] = v3dot(a, FA) + v3dot(s, FL);
#pragma unroll
    for (int j = 0; j < 6; j++) o.B[f][j] = v3dot(Sa[j], FA) + v3dot(Sl[j], FL);
    v3add(Fa[5], Fa[5], fa); v3add(Fl[5], Fl[5], fl);
    Im[5] += m.mass[k + 1]; v3add(Ih[5], Ih[5], h);
#pragma unroll
    for (int i = 0; i < 6; i++) II[5][i] += IO[i];
  }
  // backward sweep over the arm: bias, composite inertia, inertia columns
#pragma unroll
  for (int k = 5; k >= 0; k--) {
    o.bias[k] = v3dot(Sa[k], Fa[k]) + v3dot(Sl[k], Fl[k]);
    float FA[3], FL[3];
    sinertia_mul(Im[k], Ih[k], II[k], Sa[k], Sl[k], FA, FL);
#pragma unroll
    for (int j = 0; j <= k; j++) o.A[PK(k, j)] = v3dot(Sa[j], FA) + v3dot(Sl[j], FL);
    if (k > 0) {
      v3add(Fa[k - 1], Fa[k - 1], Fa[k]); v3add(Fl[k - 1], Fl[k - 1], Fl[k]);
      Im[k - 1] += Im[k]; v3add(Ih[k - 1], Ih[k - 1], Ih[k]);
#pragma unroll
      for (int i = 0; i < 6; i++) II[k - 1][i] += II[k][i];
    }
  }
  // EE Jacobian: lin col = a_j x (ee - p_j) = Sl_j + a_j x (ee - O)
  float de[3]; v3sub(de, o.eepos, O);
#pragma unroll
  for (int j = 0; j < 6; j++) {
    float c[3]; v3cross(c, Sa[j], de);
    o.J[0][j] = Sl[j][0] + c[0]; o.J[1][j] = Sl[j][1] + c[1]; o.J[2][j] = Sl[j][2] + c[2];
    o.J[3][j] = Sa[j][0]; o.J[4][j] = Sa[j][1]; o.J[5][j] = Sa[j][2];
  }
}

// ------------------------------------------------------------------------------------------- chain algebra
// Solve M x = r for the 9-DoF chain with the fingers eliminated first (they are leaves that only couple to the
// arm): S = A - sum_f B_f B_f^T / d_f.  damp_h = h * damping is added to the diagonal (mj_Euler's implicit
// damping); damp_h = 0 gives qacc_smooth.  Returns 0 on a bad pivot.
JHD int chain_solve(const ChainOut& c, float hdt, const float* r, float* x) {
  const JacoModel& m = MDL;
  float S[21], ra[6], id[3];
#pragma unroll
  for (int i = 0; i < 21; i++) S[i] = c.A[i];
#pragma unroll
  for (int i = 0; i < 6; i++) { S[PK(i, i)] += hdt * m.ch_damping[i]; ra[i] = r[i]; }
#pragma unroll
  for (int f = 0; f < 3; f++) {
    id[f] = 1.0f / (c.Dg[f] + hdt * m.ch_damping[6 + f]);
    float rf = r[6 + f] * id[f];
#pragma unroll
    for (int i = 0; i < 6; i++) {
      float bi = c.B[f][i] * id[f];
      ra[i] -= c.B[f][i] * rf;
#pragma unroll
      for (int j = 0; j <= i; j++) S[PK(i, j)] -= bi * c.B[f][j];
    }
  }
  int ok = chol_packed<6>(S);
  chol_solve_packed<6>(S, ra);
#pragma unroll
  for (int i = 0; i < 6; i++) x[i] = ra[i];
#pragma unroll
  for (int f = 0; f < 3; f++) {
    float sm = r[6 + f];
#pragma unroll
    for (int j = 0; j < 6; j++) sm -= c.B[f][j] * ra[j];
    x[6 + f] = sm * id[f];
  }
  return ok;
}

// abr_control OSC.generate on the quantities of one forward, without explicit inverses:
// A = L L^T, Y = L^-1 J^T, Mx_inv = Y^T Y, Mx u_task by a second Cholesky.
// When the reference takes the pinv branch (|det Mx_inv| < 1e-3, numpy.linalg.pinv(Mx_inv, rcond = 1e-4)) or a pivot
// fails, `mode` decides: OSC_BAIL returns 0 (caller leaves the fast path), OSC_INLINE runs the Jacobi
// eigen-decomposition here, OSC_DEFER fills `od` with what osc_finish_pinv needs and returns 2 (the pipeline hands
// these rare, slow cases to its dirty-env stream so the main kernel never executes the eigen-solver).
enum { OSC_BAIL = 0, OSC_INLINE = 1, OSC_DEFER = 2 };
struct OscDefer { float X[21], ut[6], u0[6], J[6][6]; };   // Mx_inv (packed lower), scaled task error, -kv M dq + bias, EE Jacobian
JHD void osc_finish_pinv(const OscDefer& od, float* u) {
  float Xf[36], Mx[36], f[6];
  for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) Xf[6 * i + j] = od.X[i >= j ? PK(i, j) : PK(j, i)];
  FAST_WSTAT(15);
  sym_pinv6(Xf, 1e-4f, Mx);
  for (int i = 0; i < 6; i++) { float sm = 0; for (int j = 0; j < 6; j++) sm += Mx[6 * i + j] * od.ut[j]; f[i] = sm; }
  for (int i = 0; i < 6; i++) { float tt = 0; for (int j = 0; j < 6; j++) tt += od.J[j][i] * f[j]; u[i] = od.u0[i] - tt; }
}
JHD int fast_osc(const ChainOut& c, const float* tq, const float* target, const float* dq, float* u, int mode = OSC_BAIL, OscDefer* od = nullptr) {
  const float kp = 50, ko = 180, kv = 20, vmax0 = 0.4f, vmax1 = 1.0472f;
  float L[21], Y[6][6], X[21];
#pragma unroll
  for (int i = 0; i < 21; i++) L[i] = c.A[i];
  int ok = chol_packed<6>(L);
  // Y[:, i] = L^-1 (row i of J)^T  (forward substitution only)
#pragma unroll
  for (int i = 0; i < 6; i++) {
#pragma unroll
    for (int r = 0; r < 6; r++) {
      float sm = c.J[i][r];
#pragma unroll
      for (int k = 0; k < r; k++) sm -= L[PK(r, k)] * Y[k][i];
      Y[r][i] = sm / L[PK(r, r)];
    }
  }
  float X0[21];
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j <= i; j++) {
      float sm = 0;
#pragma unroll
      for (int k = 0; k < 6; k++) sm += Y[k][i] * Y[k][j];
      X[PK(i, j)] = sm; X0[PK(i, j)] = sm;
    }
  ok &= chol_packed<6>(X);
  float det = 1;
#pragma unroll
  for (int i = 0; i < 6; i++) det *= X[PK(i, i)] * X[PK(i, i)];
  const bool singular = !ok || !(fabsf(det) >= 1e-3f);
  if (singular && mode == OSC_BAIL) return 0;
  float ut[6];
#pragma unroll
  for (int i = 0; i < 3; i++) ut[i] = c.eepos[i] - target[i];
  float qc[4] = {c.eeq[0], -c.eeq[1], -c.eeq[2], -c.eeq[3]}, qr[4];
  quat_mul(qr, tq, qc);
  float sgq = qr[0] > 0 ? 1.0f : (qr[0] < 0 ? -1.0f : 0.0f);
  ut[3] = -qr[1] * sgq; ut[4] = -qr[2] * sgq; ut[5] = -qr[3] * sgq;
  float sat_xyz = vmax0 / kp * kv, sat_abg = vmax1 / ko * kv, nx = v3norm(ut), na = v3norm(ut + 3), sc[6] = {1, 1, 1, 1, 1, 1};
  if (nx > sat_xyz) { float r = sat_xyz / nx; sc[0] *= r; sc[1] *= r; sc[2] *= r; }
  if (na > sat_abg) { float r = sat_abg / na; sc[3] *= r; sc[4] *= r; sc[5] *= r; }
#pragma unroll
  for (int i = 0; i < 6; i++) ut[i] = kv * sc[i] * ((i < 3 ? kp : ko) / kv) * ut[i];
  float u0[6];
#pragma unroll
  for (int i = 0; i < 6; i++) {
    float sm = 0;
#pragma unroll
    for (int j = 0; j < 6; j++) sm += c.A[i >= j ? PK(i, j) : PK(j, i)] * dq[j];
    u0[i] = -kv * sm + c.bias[i];
  }
  if (singular) {  // numpy.linalg.pinv(Mx_inv, rcond = 1e-4) branch of the reference controller (rare)
    OscDefer loc; OscDefer& d = (mode == OSC_DEFER) ? *od : loc;
    for (int i = 0; i < 21; i++) d.X[i] = X0[i];
    for (int i = 0; i < 6; i++) { d.ut[i] = ut[i]; d.u0[i] = u0[i]; for (int j = 0; j < 6; j++) d.J[i][j] = c.J[i][j]; }
    if (mode == OSC_DEFER) return 2;
    osc_finish_pinv(d, u);
    return 3;   // done, through the pinv branch (callers count it: STAT_PINV)
  }
  chol_solve_packed<6>(X, ut);  // ut <- Mx u_task
#pragma unroll
  for (int i = 0; i < 6; i++) {
    float tt = 0;
#pragma unroll
    for (int j = 0; j < 6; j++) tt += c.J[j][i] * ut[j];
    u[i] = u0[i] - tt;
  }
  return 1;
}

// ------------------------------------------------------------------------------------------- free bodies
struct FreeBody {       // one free box at the current state
  float x[3], R[9];     // body origin, rotation (world-from-body)
  float M[21], L[21];   // 6x6 inertia in (world-linear, body-angular) dofs, packed lower; its Cholesky factor
  float bias[6];
  float gp[3], gR[9];   // world pose of its box geom
};
JHD void free_body_setup(int b, int geom, float* q7, const float* v6, FreeBody& fb) {
  const JacoModel& m = MDL;
  quat_normalize(q7 + 3);
  v3copy(fb.x, q7); quat2mat(fb.R, q7 + 3);
  const float* R = fb.R; const float ms = m.mass[b];
  float d[3], Iw[6];
  m3mulv(d, R, m.com[b]);  // com - origin, world
  { const float* I = m.inertia[b];
    float S[9] = {I[0], I[3], I[4], I[3], I[1], I[5], I[4], I[5], I[2]}, T[9];
    m3mul(T, R, S);
    Iw[0] = T[0] * R[0] + T[1] * R[1] + T[2] * R[2]; Iw[1] = T[3] * R[3] + T[4] * R[4] + T[5] * R[5]; Iw[2] = T[6] * R[6] + T[7] * R[7] + T[8] * R[8];
    Iw[3] = T[0] * R[3] + T[1] * R[4] + T[2] * R[5]; Iw[4] = T[0] * R[6] + T[1] * R[7] + T[2] * R[8]; Iw[5] = T[3] * R[6] + T[4] * R[7] + T[5] * R[8]; }
  // crba (jaco_sim.cuh crba_free)
  float* M = fb.M;
#pragma unroll
  for (int i = 0; i < 21; i++) M[i] = 0;
  M[PK(0, 0)] = M[PK(1, 1)] = M[PK(2, 2)] = ms;
  float r[3][3], f[3][3], h[3][3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    v3set(r[i], R[i], R[3 + i], R[6 + i]);
    v3cross(f[i], r[i], d); f[i][0] *= ms; f[i][1] *= ms; f[i][2] *= ms;
    float t[3]; s3mulv(h[i], Iw, r[i]); v3cross(t, d, f[i]); v3add(h[i], h[i], t);
  }
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int k = 0; k < 3; k++) M[PK(3 + i, k)] = f[i][k];
#pragma unroll
    for (int j = 0; j <= i; j++) M[PK(3 + i, 3 + j)] = v3dot(r[j], h[i]);
  }
#pragma unroll
  for (int i = 0; i < 21; i++) fb.L[i] = M[i];
  chol_packed<6>(fb.L);
  // bias (jaco_sim.cuh rne_free)
  { float bw[3], t[3], F[3], Iwv[3], N[3], no[3];
    m3mulv(bw, R, v6 + 3);
    v3cross(t, bw, d); v3cross(t, bw, t);
    F[0] = ms * t[0]; F[1] = ms * t[1]; F[2] = ms * (t[2] - m.gravz);
    s3mulv(Iwv, Iw, bw); v3cross(N, bw, Iwv);
    v3cross(no, d, F); v3add(no, no, N);
    v3copy(fb.bias, F); m3tmulv(fb.bias + 3, R, no); }
  const JacoGeom& G = m.geoms[geom];
  { float t[3]; m3mulv(t, R, G.pos); v3add(fb.gp, fb.x, t); m3mul(fb.gR, R, G.R); }
}
JHD void mul_M6(const float* M, const float* x, float* y) {
#pragma unroll
  for (int i = 0; i < 6; i++) {
    float a = 0;
#pragma unroll
    for (int j = 0; j < 6; j++) a += M[i >= j ? PK(i, j) : PK(j, i)] * x[j];
    y[i] = a;
  }
}

// ------------------------------------------------------------------------------------------- fast collision
// Up to 4 contacts of ONE geom pair (static geom -> free box), all sharing one normal.
struct Con4 { float n[3]; float pos[4][3], dist[4]; int valid[4]; int pair; };
JHD void con4_clear(Con4& c) {
#pragma unroll
  for (int k = 0; k < 4; k++) { c.valid[k] = 0; c.dist[k] = 0; v3set(c.pos[k], 0, 0, 0); }
  v3set(c.n, 0, 0, 1); c.pair = 0;
}
JHD int con4_count(const Con4& c) { return c.valid[0] + c.valid[1] + c.valid[2] + c.valid[3]; }

// plane (g1) vs box: same contacts, same order as collide_plane_box (first 4 qualifying corners)
JHD void fast_plane_box(int pair, const float* pp, const float* pR, const float* bp, const float* bR, const float* size, float margin, Con4& out) {
  con4_clear(out); out.pair = pair;
  float n[3] = {pR[2], pR[5], pR[8]}, d[3];
  v3copy(out.n, n);
  v3sub(d, bp, pp);
  float dist = v3dot(d, n);
  int cnt = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    float v[3] = {(i & 1) ? size[0] : -size[0], (i & 2) ? size[1] : -size[1], (i & 4) ? size[2] : -size[2]}, corner[3], pos[3];
    m3mulv(corner, bR, v);
    float ld = v3dot(n, corner);
    bool q = !(dist + ld > margin || ld > 0) && cnt < 4;
    float cd = dist + ld;
    v3add(pos, bp, corner); v3addscl(pos, pos, n, -cd * 0.5f);
#pragma unroll
    for (int s = 0; s < 4; s++) if (q && cnt == s) { v3copy(out.pos[s], pos); out.dist[s] = cd; out.valid[s] = 1; }
    cnt += q ? 1 : 0;
  }
}

// box (A, static) vs box (B): the face-contact case of collide_box_box whose incident face lies wholly inside
// the reference face (what a box resting on a larger box produces).  Returns 0 = no contact (separated),
// 1 = contacts in `out` (identical to collide_box_box), -1 = another case (edge contact, clipped polygon): bail.
JHD int fast_box_box_face(int pair, const float* pa, const float* Ra, const float* ha, const float* pb, const float* Rb, const float* hb, float margin, Con4& out) {
  con4_clear(out); out.pair = pair;
  float t[3], A[3][3], B[3][3], C[3][3], AC[3][3];
  v3sub(t, pb, pa);
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int k = 0; k < 3; k++) { A[i][k] = Ra[3 * k + i]; B[i][k] = Rb[3 * k + i]; }
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) { C[i][j] = v3dot(A[i], B[j]); AC[i][j] = fabsf(C[i][j]); }
  float ta[3] = {v3dot(t, A[0]), v3dot(t, A[1]), v3dot(t, A[2])}, tb[3] = {v3dot(t, B[0]), v3dot(t, B[1]), v3dot(t, B[2])};
  float best = -1e30f; int code = -1; bool sep = false;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    float sp = fabsf(ta[i]) - (ha[i] + hb[0] * AC[i][0] + hb[1] * AC[i][1] + hb[2] * AC[i][2]);
    sep = sep || sp > margin;
    if (sp > best) { best = sp; code = i; }
  }
#pragma unroll
  for (int j = 0; j < 3; j++) {
    float sp = fabsf(tb[j]) - (hb[j] + ha[0] * AC[0][j] + ha[1] * AC[1][j] + ha[2] * AC[2][j]);
    sep = sep || sp > margin;
    if (sp > best) { best = sp; code = 3 + j; }
  }
  float ebest = -1e30f; int ecode = -1;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      float Lx[3]; v3cross(Lx, A[i], B[j]);
      float l = v3norm(Lx);
      if (l < 1e-6f) continue;
      const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      float ra = ha[i1] * AC[i2][j] + ha[i2] * AC[i1][j], rb = hb[j1] * AC[i][j2] + hb[j2] * AC[i][j1];
      float sp = (fabsf(v3dot(t, Lx)) - ra - rb) / l;
      sep = sep || sp > margin;
      if (sp > ebest) { ebest = sp; ecode = 3 * i + j; }
    }
  if (sep) return 0;
  if (ecode >= 0 && ebest * 1.05f > best) return -1;  // edge-edge contact
  // face contact: reference box / axis, incident face
  const bool refA = code < 3; const int axn = refA ? code : code - 3;
  float pr[3], hr[3], pi[3], hi[3], Rr[3][3], Ri[3][3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    pr[i] = refA ? pa[i] : pb[i]; hr[i] = refA ? ha[i] : hb[i]; pi[i] = refA ? pb[i] : pa[i]; hi[i] = refA ? hb[i] : ha[i];
#pragma unroll
    for (int k = 0; k < 3; k++) { Rr[i][k] = refA ? A[i][k] : B[i][k]; Ri[i][k] = refA ? B[i][k] : A[i][k]; }
  }
  float tsel = refA ? (axn == 0 ? ta[0] : (axn == 1 ? ta[1] : ta[2])) : -(axn == 0 ? tb[0] : (axn == 1 ? tb[1] : tb[2]));
  float sg = tsel >= 0 ? 1.0f : -1.0f;
  float Rn[3], R1[3], R2[3], hn, h1, h2;  // reference axis and the two in-face axes (cyclic order)
#pragma unroll
  for (int k = 0; k < 3; k++) {
    Rn[k] = axn == 0 ? Rr[0][k] : (axn == 1 ? Rr[1][k] : Rr[2][k]);
    R1[k] = axn == 0 ? Rr[1][k] : (axn == 1 ? Rr[2][k] : Rr[0][k]);
    R2[k] = axn == 0 ? Rr[2][k] : (axn == 1 ? Rr[0][k] : Rr[1][k]);
  }
  hn = axn == 0 ? hr[0] : (axn == 1 ? hr[1] : hr[2]); h1 = axn == 0 ? hr[1] : (axn == 1 ? hr[2] : hr[0]); h2 = axn == 0 ? hr[2] : (axn == 1 ? hr[0] : hr[1]);
  float nr[3] = {sg * Rn[0], sg * Rn[1], sg * Rn[2]};
  int ia = 0; float bd = -1;
#pragma unroll
  for (int k = 0; k < 3; k++) { float dd = fabsf(v3dot(Ri[k], nr)); if (dd > bd) { bd = dd; ia = k; } }
  float Ia[3], I1[3], I2[3], ga, g1, g2;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    Ia[k] = ia == 0 ? Ri[0][k] : (ia == 1 ? Ri[1][k] : Ri[2][k]);
    I1[k] = ia == 0 ? Ri[1][k] : (ia == 1 ? Ri[2][k] : Ri[0][k]);
    I2[k] = ia == 0 ? Ri[2][k] : (ia == 1 ? Ri[0][k] : Ri[1][k]);
  }
  ga = ia == 0 ? hi[0] : (ia == 1 ? hi[1] : hi[2]); g1 = ia == 0 ? hi[1] : (ia == 1 ? hi[2] : hi[0]); g2 = ia == 0 ? hi[2] : (ia == 1 ? hi[0] : hi[1]);
  float si = v3dot(Ia, nr) > 0 ? -1.0f : 1.0f;
  float fc[3]; v3addscl(fc, pi, Ia, si * ga);
  const float sx[4] = {1, -1, -1, 1}, sy[4] = {1, 1, -1, -1};
  bool inside = true;
  v3set(out.n, refA ? nr[0] : -nr[0], refA ? nr[1] : -nr[1], refA ? nr[2] : -nr[2]);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    float v[3], dlt[3];
    v3addscl(v, fc, I1, sx[k] * g1); v3addscl(v, v, I2, sy[k] * g2);
    v3sub(dlt, v, pr);
    float px = v3dot(dlt, R1), py = v3dot(dlt, R2), depth = v3dot(dlt, nr) - hn;
    inside = inside && (px - h1 <= 0) && (-px - h1 <= 0) && (py - h2 <= 0) && (-py - h2 <= 0);
    float pos[3];
    v3addscl(pos, pr, R1, px); v3addscl(pos, pos, R2, py); v3addscl(pos, pos, nr, hn + depth * 0.5f);
    v3copy(out.pos[k], pos); out.dist[k] = depth; out.valid[k] = depth > margin ? 0 : 1;
  }
  if (!inside) return -1;
  return 1;
}

// ------------------------------------------------------------------------------------------- 6-DoF contact solve
// Constraint rows of one free box touching a static geom through <= 4 contacts with a common frame (n, t1, t2),
// condim 3, pyramidal cone: rows (n +- mu t1), (n +- mu t2) per contact, all with the same D (make_constraints).
struct Rows4 {
  float n[3], t1[3], t2[3];   // shared contact frame
  float jr[4][3][3];          // jr[c][d] = R^T (r_c x dir_d): angular part of the Jacobian of component d
  float A[4][3], D[4];        // aref components, 1/R (0 for an empty slot)
  float e[4][3], p[4][3];     // (J a - aref) components at the iterate, J p components
  float mu;
};
JHD void rows4_build(const Con4& c, const FreeBody& fb, const float* v6, Rows4& r) {
  const JacoPair& P = MDL.pairs[c.pair];
  v3copy(r.n, c.n); make_frame(r.n, r.t1, r.t2);
  r.mu = P.friction[0];
  const float mu0 = P.friction[0];
  const float dA = P.invw_tran + mu0 * mu0 * P.invw_tran;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    float rr[3], x[3], comp[3];
    v3sub(rr, c.pos[k], fb.x);
    v3cross(x, rr, r.n); m3tmulv(r.jr[k][0], fb.R, x);
    v3cross(x, rr, r.t1); m3tmulv(r.jr[k][1], fb.R, x);
    v3cross(x, rr, r.t2); m3tmulv(r.jr[k][2], fb.R, x);
    comp[0] = v3dot(r.n, v6) + v3dot(r.jr[k][0], v6 + 3);
    comp[1] = v3dot(r.t1, v6) + v3dot(r.jr[k][1], v6 + 3);
    comp[2] = v3dot(r.t2, v6) + v3dot(r.jr[k][2], v6 + 3);
    float kk, b, imp, R0;
    row_params(P.solref, P.solimp, c.dist[k], P.margin, dA, &kk, &b, &imp, &R0);
    float Rr = fmaxf(JMINVAL, 2 * mu0 * mu0 * R0);
    r.D[k] = c.valid[k] ? 1.0f / Rr : 0.0f;
    r.A[k][0] = -b * comp[0] - kk * imp * (c.dist[k] - P.margin);
    r.A[k][1] = -b * comp[1]; r.A[k][2] = -b * comp[2];
  }
}
// components of J x for all slots; out = J x - sub
JHD void rows4_eval(const Rows4& r, const float* x, bool minus_aref, float out[4][3]) {
  float l0 = v3dot(r.n, x), l1 = v3dot(r.t1, x), l2 = v3dot(r.t2, x);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    out[k][0] = l0 + v3dot(r.jr[k][0], x + 3) - (minus_aref ? r.A[k][0] : 0.0f);
    out[k][1] = l1 + v3dot(r.jr[k][1], x + 3) - (minus_aref ? r.A[k][1] : 0.0f);
    out[k][2] = l2 + v3dot(r.jr[k][2], x + 3) - (minus_aref ? r.A[k][2] : 0.0f);
  }
}
// cost 0.5 D jar^2 over active rows at e + al p, with first / second derivative in al
JHD void rows4_cost(const Rows4& r, const float e[4][3], float al, float* cost, float* d1, float* d2) {
  float c0 = 0, g1 = 0, g2 = 0; const float mu = r.mu;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const float D = r.D[k];
#pragma unroll
    for (int d = 1; d <= 2; d++) {
#pragma unroll
      for (int sg = 1; sg >= -1; sg -= 2) {
        float jp = r.p[k][0] + sg * mu * r.p[k][d];
        float j = e[k][0] + sg * mu * e[k][d] + al * jp;
        if (j < 0) { c0 += 0.5f * D * j * j; g1 += D * j * jp; g2 += D * jp * jp; }
      }
    }
  }
  *cost = c0; *d1 = g1; *d2 = g2;
}
// J^T f at the iterate (6-vector)
JHD void rows4_force(const Rows4& r, const float e[4][3], float* out) {
  float s0 = 0, s1 = 0, s2 = 0, oa[3] = {0, 0, 0}; const float mu = r.mu;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const float D = r.D[k];
    float sc0 = 0, sc[3] = {0, 0, 0};
#pragma unroll
    for (int d = 1; d <= 2; d++) {
      float jp_ = e[k][0] + mu * e[k][d], jm_ = e[k][0] - mu * e[k][d];
      float fp = jp_ < 0 ? -D * jp_ : 0.0f, fm = jm_ < 0 ? -D * jm_ : 0.0f;
      sc0 += fp + fm; sc[d] = mu * (fp - fm);
    }
    s0 += sc0; s1 += sc[1]; s2 += sc[2];
#pragma unroll
    for (int i = 0; i < 3; i++) oa[i] += sc0 * r.jr[k][0][i] + sc[1] * r.jr[k][1][i] + sc[2] * r.jr[k][2][i];
  }
#pragma unroll
  for (int i = 0; i < 3; i++) { out[i] = s0 * r.n[i] + s1 * r.t1[i] + s2 * r.t2[i]; out[3 + i] = oa[i]; }
}
// H = M + J^T diag(D active) J, packed lower 6x6
JHD void rows4_hessian(const Rows4& r, const float e[4][3], const float* M, float* H) {
  const float mu = r.mu;
  float Ws[5] = {0, 0, 0, 0, 0};                       // sum_c W_c: w00 w01 w02 w11 w22 (w12 = 0)
  float Z[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};   // sum_c W_c Jr_c
  float Haa[6] = {0, 0, 0, 0, 0, 0};                   // sum_c Jr_c^T W_c Jr_c : 00 10 11 20 21 22
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const float D = r.D[k];
    float a1p = (e[k][0] + mu * e[k][1] < 0) ? D : 0.0f, a1m = (e[k][0] - mu * e[k][1] < 0) ? D : 0.0f;
    float a2p = (e[k][0] + mu * e[k][2] < 0) ? D : 0.0f, a2m = (e[k][0] - mu * e[k][2] < 0) ? D : 0.0f;
    float w00 = a1p + a1m + a2p + a2m, w01 = mu * (a1p - a1m), w02 = mu * (a2p - a2m), w11 = mu * mu * (a1p + a1m), w22 = mu * mu * (a2p + a2m);
    Ws[0] += w00; Ws[1] += w01; Ws[2] += w02; Ws[3] += w11; Ws[4] += w22;
    float WJ[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
      WJ[0][i] = w00 * r.jr[k][0][i] + w01 * r.jr[k][1][i] + w02 * r.jr[k][2][i];
      WJ[1][i] = w01 * r.jr[k][0][i] + w11 * r.jr[k][1][i];
      WJ[2][i] = w02 * r.jr[k][0][i] + w22 * r.jr[k][2][i];
      Z[0][i] += WJ[0][i]; Z[1][i] += WJ[1][i]; Z[2][i] += WJ[2][i];
    }
    int q = 0;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) Haa[q++] += r.jr[k][0][i] * WJ[0][j] + r.jr[k][1][i] * WJ[1][j] + r.jr[k][2][i] * WJ[2][j];
  }
  // dirs^T Ws dirs and dirs^T Z
  float Y[3][3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    Y[0][i] = Ws[0] * r.n[i] + Ws[1] * r.t1[i] + Ws[2] * r.t2[i];
    Y[1][i] = Ws[1] * r.n[i] + Ws[3] * r.t1[i];
    Y[2][i] = Ws[2] * r.n[i] + Ws[4] * r.t2[i];
  }
#pragma unroll
  for (int i = 0; i < 21; i++) H[i] = M[i];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j <= i; j++) H[PK(i, j)] += r.n[i] * Y[0][j] + r.t1[i] * Y[1][j] + r.t2[i] * Y[2][j];
#pragma unroll
  for (int i = 0; i < 3; i++)   // angular row i, linear column j
#pragma unroll
    for (int j = 0; j < 3; j++) H[PK(3 + i, j)] += r.n[j] * Z[0][i] + r.t1[j] * Z[1][i] + r.t2[j] * Z[2][i];
  { int q = 0;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) H[PK(3 + i, 3 + j)] += Haa[q++]; }
}

// Newton on min 0.5 (a - as)' M (a - as) + sum 0.5 D min(0, J a - aref)^2 for one free box (solve_constraints
// semantics restricted to its 6 dofs).  In: as = qacc_smooth, aw = warm start.  Out: a, fcon = J^T f.
JHD int fast_solve6(Rows4& r, const FreeBody& fb, const float* as_, const float* aw, float* a, float* fcon) {
  float e[4][3], ew[4][3], cw_, cs_, t1_, t2_;
  const float zero4[4][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
#pragma unroll
  for (int k = 0; k < 4; k++) { r.p[k][0] = r.p[k][1] = r.p[k][2] = 0; }
  rows4_eval(r, aw, true, ew); rows4_cost(r, ew, 0, &cw_, &t1_, &t2_);
  { float d[6], Md[6];
#pragma unroll
    for (int i = 0; i < 6; i++) d[i] = aw[i] - as_[i];
    mul_M6(fb.M, d, Md);
#pragma unroll
    for (int i = 0; i < 6; i++) cw_ += 0.5f * Md[i] * d[i]; }
  rows4_eval(r, as_, true, e); rows4_cost(r, e, 0, &cs_, &t1_, &t2_);
  const bool usew = cw_ < cs_;
#pragma unroll
  for (int i = 0; i < 6; i++) a[i] = usew ? aw[i] : as_[i];
#pragma unroll
  for (int k = 0; k < 4; k++) { e[k][0] = usew ? ew[k][0] : e[k][0]; e[k][1] = usew ? ew[k][1] : e[k][1]; e[k][2] = usew ? ew[k][2] : e[k][2]; }
  (void)zero4;
  int it = 0;
  for (; it < JACO_SOLVER_MAXIT; it++) {
    float d[6], g[6], jf[6], H[21], p[6];
#pragma unroll
    for (int i = 0; i < 6; i++) d[i] = a[i] - as_[i];
    mul_M6(fb.M, d, g);
    rows4_force(r, e, jf);
    float gn = 0, fn = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) { fn += jf[i] * jf[i] + g[i] * g[i]; g[i] -= jf[i]; gn += g[i] * g[i]; }
    if (gn <= 1e-12f * fn || gn < 1e-20f) break;
    rows4_hessian(r, e, fb.M, H);
    chol_packed<6>(H);
#pragma unroll
    for (int i = 0; i < 6; i++) p[i] = -g[i];
    chol_solve_packed<6>(H, p);
    rows4_eval(r, p, false, r.p);
    float Mp[6], q1 = 0, q2 = 0;
    mul_M6(fb.M, p, Mp);
#pragma unroll
    for (int i = 0; i < 6; i++) { q2 += p[i] * Mp[i]; q1 += Mp[i] * d[i]; }
    float lo = 0, hi = -1, al = 0, d1, d2, c0;
    rows4_cost(r, e, 0, &c0, &d1, &d2);
    float dlo = q1 + d1, slope0 = dlo;
    if (!(dlo < 0)) break;
    al = -dlo / (q2 + d2);
    for (int ls = 0; ls < 12; ls++) {
      rows4_cost(r, e, al, &c0, &d1, &d2);
      float dd = q1 + q2 * al + d1, hh = q2 + d2;
      if (fabsf(dd) <= 1e-6f * fabsf(slope0)) break;
      if (dd < 0) { lo = al; dlo = dd; } else { hi = al; }
      float nx = al - dd / hh;
      if (hi > 0 && (nx <= lo || nx >= hi)) nx = 0.5f * (lo + hi);
      if (nx == al) break;
      al = nx;
    }
    float amax = 0, smax = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) { a[i] += al * p[i]; amax = fmaxf(amax, fabsf(a[i])); smax = fmaxf(smax, fabsf(al * p[i])); }
#pragma unroll
    for (int k = 0; k < 4; k++) { e[k][0] += al * r.p[k][0]; e[k][1] += al * r.p[k][1]; e[k][2] += al * r.p[k][2]; }
    if (smax <= 2e-7f * fmaxf(1.0f, amax)) break;
  }
  rows4_force(r, e, fcon);
  return it;
}

// forward + Euler of one free box whose contacts (if any) are in `c`: updates q7 / v6 / warm6
JHD void fast_free_step(FreeBody& fb, const Con4& c, float* q7, float* v6, float* warm6) {
  const float h = MDL.dt;
#if defined(JACO_COLLISION_STATS) && !defined(__CUDACC__)
  float in_[19]; for (int i = 0; i < 7; i++) in_[i] = q7[i]; for (int i = 0; i < 6; i++) { in_[7 + i] = v6[i]; in_[13 + i] = warm6[i]; }
  struct Probe { float* q; float* v; float* w; float* in; ~Probe() { int same = 1; for (int i = 0; i < 7; i++) same &= (q[i] == in[i]); for (int i = 0; i < 6; i++) same &= (v[i] == in[7 + i]) && (w[i] == in[13 + i]); FAST_WSTAT(11); if (same) FAST_WSTAT(12); } } probe_{q7, v6, warm6, in_};
#endif
  float fsm[6], as_[6], a[6], fcon[6];
#pragma unroll
  for (int i = 0; i < 6; i++) { fsm[i] = -fb.bias[i]; as_[i] = fsm[i]; }
  chol_solve_packed<6>(fb.L, as_);
  if (con4_count(c) > 0) {
    Rows4 r;
    rows4_build(c, fb, v6, r);
    fast_solve6(r, fb, as_, warm6, a, fcon);
  } else {
#pragma unroll
    for (int i = 0; i < 6; i++) { a[i] = as_[i]; fcon[i] = 0; }
  }
  float y[6];
#pragma unroll
  for (int i = 0; i < 6; i++) { warm6[i] = a[i]; y[i] = fsm[i] + fcon[i]; }
  chol_solve_packed<6>(fb.L, y);
#pragma unroll
  for (int i = 0; i < 6; i++) v6[i] += h * y[i];
  q7[0] += h * v6[0]; q7[1] += h * v6[1]; q7[2] += h * v6[2];
  float n = sqrtf(v6[3] * v6[3] + v6[4] * v6[4] + v6[5] * v6[5]);
  if (n > JMINVAL) {
    float sh, chh; sincosf(0.5f * n * h, &sh, &chh);
    float inv = sh / n, dq[4] = {chh, v6[3] * inv, v6[4] * inv, v6[5] * inv}, rq[4];
    quat_mul(rq, q7 + 3, dq);
    q7[3] = rq[0]; q7[4] = rq[1]; q7[5] = rq[2]; q7[6] = rq[3];
    quat_normalize(q7 + 3);
  }
}

// ------------------------------------------------------------------------------------------- substep / env step
// scene pair indices the fast path relies on (checked against the model by fast_model_ok)
#define FP_FLOOR_OBJ 0
#define FP_HOLDER_OBJ 1
#define FP_DEST_OBJ 2
#define FP_FLOOR_DEST 3
#define FG_FLOOR 0
#define FG_HOLDER 1
#define FG_OBJ 2
#define FG_DEST 3

// host-side check that the compiled model has the structure hard-wired above
static inline bool fast_model_ok(const JacoModel& m) {
  const int want[4][2] = {{FG_FLOOR, FG_OBJ}, {FG_HOLDER, FG_OBJ}, {FG_DEST, FG_OBJ}, {FG_FLOOR, FG_DEST}};
  for (int p = 0; p < 4; p++) if (m.pairs[p].g1 != want[p][0] || m.pairs[p].g2 != want[p][1] || m.pairs[p].condim != 3 || m.pairs[p].group != 0) return false;
  for (int p = 4; p < JM_NPAIR; p++) if (m.pairs[p].group == 0) return false;
  if (m.geoms[FG_FLOOR].type != JACO_GEOM_PLANE || m.geoms[FG_FLOOR].body != 0) return false;
  if (m.geoms[FG_HOLDER].type != JACO_GEOM_BOX || m.geoms[FG_HOLDER].body != 0) return false;
  if (m.geoms[FG_OBJ].type != JACO_GEOM_BOX || m.geoms[FG_OBJ].body != 10) return false;
  if (m.geoms[FG_DEST].type != JACO_GEOM_BOX || m.geoms[FG_DEST].body != 11) return false;
  for (int g = 4; g < JM_NGEOM; g++) {
    int b = m.geoms[g].body;
    bool fore = (g >= 30 && g <= 32);
    if (fore ? (m.geoms[g].type != JACO_GEOM_CAPSULE || b != g - 27) : (b < 6 || b > 9)) return false;  // hand cluster on bodies 6..9
  }
  // every non-scene pair must involve only hand-cluster / fore-arm geoms on one side and a scene geom on the other
  for (int p = 4; p < JM_NPAIR; p++) {
    int a = m.pairs[p].g1, b = m.pairs[p].g2;
    if (!((a <= 3 && b >= 4) || (b <= 3 && a >= 4))) return false;
  }
  return true;
}

// Owner thread -> warp workspace, stage A: poses of the geom-carrying bodies only (cheap; no kinematics redone)
JHD void coop_publish_poses(CoopPub& w, const ChainOut& c, const FreeBody& fo, const FreeBody& fd, int flags) {
  v3set(w.bx[0], 0, 0, 0);
  { const float I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; for (int i = 0; i < 9; i++) w.bR[0][i] = I[i]; }
  for (int b = 3; b <= 9; b++) { v3copy(w.bx[b], c.bp[b - 3]); for (int i = 0; i < 9; i++) w.bR[b][i] = c.bRm[b - 3][i]; }
  v3copy(w.bx[10], fo.x); v3copy(w.bx[11], fd.x);
  for (int i = 0; i < 9; i++) { w.bR[10][i] = fo.R[i]; w.bR[11][i] = fd.R[i]; }
  w.flags = flags;
}
// Owner thread -> warp workspace: the mj_forward quantities of this substep the cooperative path consumes.
JHD void coop_publish(CoopPub& w, const FastSim& s, const ChainOut& c, const FreeBody& fo, const FreeBody& fd, const float* fsm9, int want_touch) {
  const JacoModel& m = MDL;
  v3set(w.bx[0], 0, 0, 0); v3set(w.bw[0], 0, 0, 0); v3set(w.bv[0], 0, 0, 0);
  { const float I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; for (int i = 0; i < 9; i++) w.bR[0][i] = I[i]; }
  float Q6[4] = {1, 0, 0, 0};
  {  // chain kinematics + body velocities (mj_kinematics, mj_comVel restricted to what the rows need)
    float Qp[4] = {1, 0, 0, 0};
    for (int k = 0; k < 9; k++) {
      const int pb = ch_parent(k) + 1, b = k + 1;
      float t[3], qn[4], r[3];
      m3mulv(t, w.bR[pb], m.ch_pos[k]); v3add(w.bx[b], w.bx[pb], t);
      chain_joint(k, k <= 5 ? Qp : Q6, s.qpos[k], qn, w.bR[b]);
      m3mulv(w.ax[k], w.bR[b], m.ch_axis[k]);
      v3sub(r, w.bx[b], w.bx[pb]); v3cross(t, w.bw[pb], r); v3add(w.bv[b], w.bv[pb], t);
      v3addscl(w.bw[b], w.bw[pb], w.ax[k], s.qvel[k]);
      if (k <= 5) for (int i = 0; i < 4; i++) Qp[i] = qn[i];
      if (k == 5) for (int i = 0; i < 4; i++) Q6[i] = qn[i];
    }
  }
  for (int f = 0; f < 2; f++) {
    const FreeBody& fb = f ? fd : fo; const int b = 10 + f, d = 9 + 6 * f;
    v3copy(w.bx[b], fb.x);
    for (int i = 0; i < 9; i++) w.bR[b][i] = fb.R[i];
    v3copy(w.bv[b], s.qvel + d); m3mulv(w.bw[b], fb.R, s.qvel + d + 3);
    for (int i = 0; i < 21; i++) w.Mf[f][i] = fb.M[i];
    float y[6];
    for (int i = 0; i < 6; i++) y[i] = -fb.bias[i];
    chol_solve_packed<6>(fb.L, y);
    for (int i = 0; i < 6; i++) w.asm_[d + i] = y[i];
  }
  for (int i = 0; i < 9; i++) w.qpos9[i] = s.qpos[i];
  for (int i = 0; i < JACO_NV; i++) { w.qvel[i] = s.qvel[i]; w.warm[i] = s.warm[i]; }
  for (int i = 0; i < 45; i++) w.Mc[i] = 0;
  for (int i = 0; i < 21; i++) w.Mc[i] = c.A[i];
  for (int f = 0; f < 3; f++) { for (int j = 0; j < 6; j++) w.Mc[PK(6 + f, j)] = c.B[f][j]; w.Mc[PK(6 + f, 6 + f)] = c.Dg[f]; }
  { float a9[9]; chain_solve(c, 0.0f, fsm9, a9); for (int i = 0; i < 9; i++) w.asm_[i] = a9[i]; }
  w.want_touch = want_touch;
}

// ------------------------------------------------------------------------------------------- thread-local arm collision
// The owner thread's own broad + narrow phase over the arm pairs (model pairs 4..59: an arm geom against the
// holder / floor / object / dest).  Every lane that is near the scene runs it for its own environment at the same
// time, so the warp pays for the slowest lane instead of the sum over lanes; the warp-cooperative stage B is then
// entered only by lanes that actually found contacts (or have an active joint limit / non-simple scene contacts).
// Semantics: collide_pair of jaco_sim.cuh in pair order, i.e. exactly what coop_collision would have produced for
// the candidate pairs; the sink holds at most MAXC arm contacts (the scene contacts of pairs 0..3 come first in
// the workspace, the total is capped at JACO_MAXCON there).
// Block-wide re-alignment points of the substep loop (hybrid kernel only).  The substep body is ~24 k straight-line
// instructions (~380 KB of SASS), far beyond the 32 KB L1.5 instruction cache: warps that run it at different
// program counters each stream it from L2 on their own (ncu: stall_no_instruction dominates).  Warps of a block that
// pass these points together share one instruction stream.  JACO_SUBSTEP_SYNC: 0 off, 1 per substep, 2 per phase.
#ifndef JACO_SUBSTEP_SYNC
#define JACO_SUBSTEP_SYNC 2
#endif
#if defined(__CUDA_ARCH__)
#define JSYNC1() do { if (COOP && JACO_SUBSTEP_SYNC >= 1) __syncthreads(); } while (0)
#define JSYNC2() do { if (COOP && JACO_SUBSTEP_SYNC >= 2) __syncthreads(); } while (0)
#define JSYNC3() do { if (COOP && JACO_SUBSTEP_SYNC >= 3) __syncthreads(); } while (0)
#else
#define JSYNC1()
#define JSYNC2()
#define JSYNC3()
#endif
#ifndef JACO_NOCOOP_OSC_MODE
#define JACO_NOCOOP_OSC_MODE OSC_INLINE   // bail-only fast path (pass 1 of the env step): the OSC's pinv branch runs in place, it is not a reason to bail
#endif
#ifndef JACO_COOP_NWS
#define JACO_COOP_NWS 8   // cooperative workspaces (= serving warps) per block of the fused kernel: 8 x 22 KB of dynamic shared memory
#endif
#ifndef JACO_TIMING_SKIP   /* timing experiments only (scripts/build_variant.sh): bit 0 no stage B, bit 1 no arm collision, bit 2 arm stage A only */
#define JACO_TIMING_SKIP 0
#endif
struct LocalSink {
  static const int MAXC = 24;
  int ncon, cpair[MAXC];
  float cpos[MAXC][3], cn[MAXC][3], cdist[MAXC];
};
struct LocalPoses {
  static JHD const JacoModel& model() { return MDLG; }
  const float* bx[JACO_NBODY];
  const float* bR[JACO_NBODY];
};
// cand_out != nullptr: when more than FAST_DEFER_PAIRS candidate pairs survive stage A, stop there, return their mask and k.ncon = -1:
// the caller hands the narrow phase of such an env to the cooperative path (lanes over pairs, same collide_pair, same contact order)
// instead of running 20+ box-box tests one after the other on the owner thread -- one arm lying across the scene (28 contacts) made
// its whole block 7 ms late.  nullptr: always run the narrow phase here (CPU harness, in-place reset).
#ifndef FAST_DEFER_PAIRS
#define FAST_DEFER_PAIRS 6
#endif
JHD JNOINLINE void fast_arm_collide(const ChainOut& c, const FreeBody& fo, const FreeBody& fd, LocalSink& k, unsigned long long* cand_out = nullptr) {
  const JacoModel& m = MDL;
  const float Z3[3] = {0, 0, 0}, I9[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  k.ncon = 0;
  // world pose of the four scene geoms (0 floor, 1 holder, 2 object, 3 dest)
  float sp[4][3], sR[4][9];
#pragma unroll
  for (int g = 0; g < 2; g++) { v3copy(sp[g], m.geoms[g].pos);
#pragma unroll
    for (int i = 0; i < 9; i++) sR[g][i] = m.geoms[g].R[i]; }
  v3copy(sp[2], fo.gp); v3copy(sp[3], fd.gp);
#pragma unroll
  for (int i = 0; i < 9; i++) { sR[2][i] = fo.gR[i]; sR[3][i] = fd.gR[i]; }
  // ---- stage A (uniform loop over the arm pairs: constant-bank reads).  Level 1: bounding sphere of the arm geom
  // against the exact scene box / plane.  Level 2, conservative and cheap: separating face axes (box-box), the
  // capsule's bounding box in the scene-box frame (capsule-box), the exact lowest point (plane-capsule).
  // Level 0: (arm body, scene geom) combinations whose body bounding sphere reaches the scene geom; bit 4 * (body - 3)
  // + scene.  The fore-arm bodies 3..5 carry one capsule each and are always kept.
  unsigned nearbits = 0xFFFu;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const float* BS = m.body_sphere[j];
    float cc[3], t[3];
    m3mulv(t, c.bRm[3 + j], BS); v3add(cc, c.bp[3 + j], t);
    const float rr = BS[3] + 1e-5f;
#pragma unroll
    for (int g = 0; g < 4; g++) {
      float d[3], l[3];
      v3sub(d, cc, sp[g]); m3tmulv(l, sR[g], d);
      bool hit;
      if (g == 0) hit = !(l[2] > rr);
      else {
        const float* hsz = m.geoms[g].size;
        const float e0 = fmaxf(fabsf(l[0]) - hsz[0], 0.0f), e1 = fmaxf(fabsf(l[1]) - hsz[1], 0.0f), e2 = fmaxf(fabsf(l[2]) - hsz[2], 0.0f);
        hit = !(e0 * e0 + e1 * e1 + e2 * e2 > rr * rr);
      }
      if (hit) nearbits |= 1u << (4 * (3 + j) + g);
    }
  }
#ifdef __CUDA_ARCH__
  unsigned warpbits = nearbits;   // union over the lanes that are in here together: keeps the pair loop uniform
  { const unsigned am = __activemask();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warpbits |= __shfl_xor_sync(am, warpbits, o) ; }
#else
  const unsigned warpbits = nearbits;
#endif
  unsigned long long cand = 0;
  for (int p = 4; p < JM_NPAIR; p++) {
    const JacoPair& P = m.pairs[p];
    const bool arm1 = P.g1 > 3;
    const JacoGeom& A = arm1 ? m.geoms[P.g1] : m.geoms[P.g2];
    const int sg = arm1 ? P.g2 : P.g1;
    if (!((warpbits >> (4 * (A.body - 3) + sg)) & 1u)) continue;   // uniform skip
    if (!((nearbits >> (4 * (A.body - 3) + sg)) & 1u)) continue;
    const JacoGeom& S = m.geoms[sg];
    const float* bR = c.bRm[A.body - 3];
    float cc[3], t[3], d[3], l[3];
    m3mulv(t, bR, A.pos); v3add(cc, c.bp[A.body - 3], t);
    const float rr = A.rbound + P.margin + 1e-5f;
    v3sub(d, cc, sp[sg]); m3tmulv(l, sR[sg], d);             // arm geom centre in the scene geom's frame
    bool hit;
    if (S.type == JACO_GEOM_PLANE) {
      hit = !(l[2] > rr);
      if (hit && A.type == JACO_GEOM_CAPSULE) {
        float ax[3], ga[3] = {A.R[2], A.R[5], A.R[8]}; m3mulv(ax, bR, ga);
        const float az = sR[sg][2] * ax[0] + sR[sg][5] * ax[1] + sR[sg][8] * ax[2];
        hit = !(l[2] - fabsf(az) * A.size[1] - A.size[0] > P.margin + 1e-5f);
      }
    } else {
      const float e0 = fmaxf(fabsf(l[0]) - S.size[0], 0.0f), e1 = fmaxf(fabsf(l[1]) - S.size[1], 0.0f), e2 = fmaxf(fabsf(l[2]) - S.size[2], 0.0f);
      hit = !(e0 * e0 + e1 * e1 + e2 * e2 > rr * rr);
      if (hit && A.type == JACO_GEOM_CAPSULE) {
        float ax[3], al[3], ga[3] = {A.R[2], A.R[5], A.R[8]}; m3mulv(ax, bR, ga); m3tmulv(al, sR[sg], ax);
        const float hl = A.size[1], r2 = A.size[0] + P.margin + 1e-5f;
        const float g0 = fmaxf(fabsf(l[0]) - hl * fabsf(al[0]) - S.size[0], 0.0f), g1 = fmaxf(fabsf(l[1]) - hl * fabsf(al[1]) - S.size[1], 0.0f),
                    g2 = fmaxf(fabsf(l[2]) - hl * fabsf(al[2]) - S.size[2], 0.0f);
        hit = !(g0 * g0 + g1 * g1 + g2 * g2 > r2 * r2);
      } else if (hit && A.type == JACO_GEOM_BOX) {
        float Ra[9], C[9];                                     // C = Rs^T Ra: arm box axes in the scene box frame
        m3mul(Ra, bR, A.R);
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) C[3 * i + j] = sR[sg][i] * Ra[j] + sR[sg][3 + i] * Ra[3 + j] + sR[sg][6 + i] * Ra[6 + j];
        const float mg = P.margin + 1e-5f;
        bool sep = false;
#pragma unroll
        for (int i = 0; i < 3; i++) {   // scene box face axes
          const float ra = A.size[0] * fabsf(C[3 * i]) + A.size[1] * fabsf(C[3 * i + 1]) + A.size[2] * fabsf(C[3 * i + 2]);
          sep = sep || (fabsf(l[i]) - S.size[i] - ra > mg);
        }
#pragma unroll
        for (int j = 0; j < 3; j++) {   // arm box face axes
          const float la = l[0] * C[j] + l[1] * C[3 + j] + l[2] * C[6 + j];
          const float rb = S.size[0] * fabsf(C[j]) + S.size[1] * fabsf(C[3 + j]) + S.size[2] * fabsf(C[6 + j]);
          sep = sep || (fabsf(la) - A.size[j] - rb > mg);
        }
        hit = !sep;
      }
    }
    if (hit) cand |= 1ull << p;
  }
  FAST_WSTAT(11);
  if (cand) FAST_WSTAT(12);
#ifdef __CUDA_ARCH__
  if (cand_out && __popcll(cand) > FAST_DEFER_PAIRS) { *cand_out = cand; k.ncon = -1; return; }
#endif
  if (!cand) return;
  if (JACO_TIMING_SKIP & 4) { k.ncon = (int)(cand >> 63); return; }
  // ---- narrow phase, one routine at a time so that lanes working on different pairs stay converged
  LocalPoses ps;
  ps.bx[0] = Z3; ps.bR[0] = I9;
  ps.bx[1] = ps.bx[2] = Z3; ps.bR[1] = ps.bR[2] = I9;   // bodies 1, 2 carry no collision geoms
#pragma unroll
  for (int b = 3; b <= 9; b++) { ps.bx[b] = c.bp[b - 3]; ps.bR[b] = c.bRm[b - 3]; }
  ps.bx[10] = fo.x; ps.bR[10] = fo.R; ps.bx[11] = fd.x; ps.bR[11] = fd.R;
  float hs[3];
  { float t[3]; m3mulv(t, ps.bR[6], m.hand_sphere); v3add(hs, ps.bx[6], t); }
  // contacts must come out in pair order: collect per routine, then merge by pair index
  LocalSink tmp; tmp.ncon = 0;
  const unsigned long long masks[3] = {m.pm_boxbox, m.pm_capbox, m.pm_plane};
  for (int r = 0; r < 3; r++) {
    unsigned long long w = cand & masks[r];
    while (w) {
#ifdef __CUDA_ARCH__
      const int p = __ffsll((long long)w) - 1;
#else
      const int p = __builtin_ffsll((long long)w) - 1;
#endif
      w &= w - 1;
      collide_pair(ps, p, hs, tmp);
    }
  }
  // stable merge by pair index (few contacts; contacts of one pair are contiguous and keep their order)
  for (int p = 4; p < JM_NPAIR && k.ncon < tmp.ncon; p++) {
    if (!((cand >> p) & 1ull)) continue;
    for (int j = 0; j < tmp.ncon; j++) if (tmp.cpair[j] == p) {
      const int d = k.ncon++;
      v3copy(k.cpos[d], tmp.cpos[j]); v3copy(k.cn[d], tmp.cn[j]); k.cdist[d] = tmp.cdist[j]; k.cpair[d] = p;
    }
  }
  if (k.ncon) FAST_WSTAT(13);
}

// Owner thread -> warp workspace: its pre-collided arm contacts; stage B then runs its narrow phase on the scene
// pairs 0..3 only.
JHD void coop_publish_contacts(CoopPub& w, const LocalSink& k, int flags) {
  w.flags = flags; w.cand[0] = 0xFu; w.cand[1] = 0u;
  w.pre_ncon = k.ncon;
  for (int j = 0; j < k.ncon; j++) { v3copy(w.pre_pos[j], k.cpos[j]); v3copy(w.pre_n[j], k.cn[j]); w.pre_dist[j] = k.cdist[j]; w.pre_pair[j] = k.cpair[j]; }
}

JHD int touch_class(const float* touch) {  // env_mujoco_util.py:470-490 on the 20 sensor readings
  const float th = 0.001f;
  int thumb = 0, index = 0, pinky = 0, inner = 0, outer = 0;
  for (int i = 1; i < 5; i++) if (touch[i] > th) thumb = 1;
  for (int i = 5; i < 9; i++) if (touch[i] > th) index = 1;
  for (int i = 9; i < 13; i++) if (touch[i] > th) pinky = 1;
  for (int i = 0; i < 13; i++) if (touch[i] > th) inner = 1;
  for (int i = 13; i < 20; i++) if (touch[i] > th) outer = 1;
  if ((thumb && index) || (thumb && pinky)) return 3;
  if (inner) return 1;
  if (outer) return 2;
  return 0;
}
// Euler step of one free box given the solved acceleration (warm start) and constraint force
JHD void free_integrate(const FreeBody& fb, const float* a6, const float* fcon6, float* q7, float* v6, float* warm6) {
  const float h = MDL.dt;
  float y[6];
#pragma unroll
  for (int i = 0; i < 6; i++) { warm6[i] = a6[i]; y[i] = -fb.bias[i] + fcon6[i]; }
  chol_solve_packed<6>(fb.L, y);
#pragma unroll
  for (int i = 0; i < 6; i++) v6[i] += h * y[i];
  q7[0] += h * v6[0]; q7[1] += h * v6[1]; q7[2] += h * v6[2];
  float n = sqrtf(v6[3] * v6[3] + v6[4] * v6[4] + v6[5] * v6[5]);
  if (n > JMINVAL) {
    float sh, chh; sincosf(0.5f * n * h, &sh, &chh);
    float inv = sh / n, dq[4] = {chh, v6[3] * inv, v6[4] * inv, v6[5] * inv}, rq[4];
    quat_mul(rq, q7 + 3, dq);
    q7[3] = rq[0]; q7[4] = rq[1]; q7[5] = rq[2]; q7[6] = rq[3];
    quat_normalize(q7 + 3);
  }
}

// ------------------------------------------------------------------------------------------- step pieces
// The env step in pieces (fast_env_step below strings them together):
//   step_begin    lagged controller data, _take_action, first OSC torque
//   sub_forward   mj_forward of the fast regime at the current state: limits, free boxes, scene contacts, chain pass
//   sub_finish    mj_Euler for the three blocks (with the cooperative path's qacc / fcon where it solved a block),
//                 bookkeeping of the last substep, OSC torque for the next substep
//   step_end      observation, reward, terminal logic (env_mujoco.py:116-150)
struct SubFwd {
  ChainOut c; FreeBody fo, fd; Con4 co, cd; float fsm[9]; int bail;
};
JHD void scene_static(SceneBoxes& sb) {
  const JacoModel& m = MDL;
  const JacoGeom& G = m.geoms[FG_HOLDER];
  v3copy(sb.hp, G.pos); v3copy(sb.hh, G.size);
#pragma unroll
  for (int i = 0; i < 9; i++) sb.hR[i] = G.R[i];
  v3copy(sb.oh, m.geoms[FG_OBJ].size); v3copy(sb.dh, m.geoms[FG_DEST].size);
}
// returns 0 when the first OSC torque needs the pinv branch and osc_mode == OSC_BAIL; 2 when deferred into *od; 3 = pinv ran inline
JHD int step_begin(const JacoCfg& cfg, FastSim& s, const float* action, int osc_mode, OscDefer* od) {
  ChainOut c;
  {  // rebuild the controller-visible mjData at the lagged chain state
    const float O0[3] = {0, 0, 0.5f};
    fast_chain<false>(s.lagq, s.lagv, O0, nullptr, c);
    float O[3]; v3copy(O, c.eepos);
    fast_chain<true>(s.lagq, s.lagv, O, nullptr, c);
  }
  v3copy(s.eepos, c.eepos);
#pragma unroll
  for (int i = 0; i < 4; i++) s.eeq[i] = c.eeq[i];
  float a[7];
#pragma unroll
  for (int i = 0; i < 7; i++) a[i] = fminf(fmaxf(action[i], -1.0f), 1.0f);
  if (cfg.task == JACO_TASK_REACHING) a[6] = 0;
  take_action(s, a, cfg.variant);
  if (cfg.task == JACO_TASK_REACHING) { s.grip = 0.6f; s.ramp0 = s.ramp1 = 0.0f; }   // util:37,76-79,625-626
  euler_rxyz_to_quat(s.tq, s.target[3], s.target[4], s.target[5]);
  quat_normalize(s.tq);
  s.touch_now = 0;
  return fast_osc(c, s.tq, s.target, s.qvel, s.u, osc_mode, od);
}
template <bool COOP>
JHD void sub_forward(FastSim& s, SceneBoxes& sb, int it, SubFwd& f) {
  const JacoModel& m = MDL;
  const JacoGeom& GF = m.geoms[FG_FLOOR];
  // ---- joint limits (mj_instantiateLimit): any active limit row leaves the fast regime
  int bail = 0;
#pragma unroll
  for (int k = 0; k < 9; k++)
    if (m.ch_limited[k] && (s.qpos[k] - m.ch_range[k][0] < 0 || m.ch_range[k][1] - s.qpos[k] < 0)) bail |= BAIL_LIMIT;
  // ---- free boxes at the current state
  FreeBody& fo = f.fo; FreeBody& fd = f.fd;
  free_body_setup(10, FG_OBJ, s.qpos + 9, s.qvel + 9, fo);
  JSYNC3();
  free_body_setup(11, FG_DEST, s.qpos + 16, s.qvel + 15, fd);
  JSYNC3();
  v3copy(s.objp, fo.x);
  v3copy(sb.op, fo.gp); v3copy(sb.dp, fd.gp);
#pragma unroll
  for (int i = 0; i < 9; i++) { sb.oR[i] = fo.gR[i]; sb.dR[i] = fd.gR[i]; }
  // ---- scene contacts
  {  // object: holder (box-box face) or floor (plane-box); both at once or anything else -> general path
    Con4 cf; con4_clear(cf);
    float dz[3]; v3sub(dz, fo.gp, GF.pos);
    float pn[3] = {GF.R[2], GF.R[5], GF.R[8]};
    if (v3dot(dz, pn) - m.geoms[FG_OBJ].rbound <= m.pairs[FP_FLOOR_OBJ].margin)
      fast_plane_box(FP_FLOOR_OBJ, GF.pos, GF.R, fo.gp, fo.gR, sb.oh, m.pairs[FP_FLOOR_OBJ].margin, cf);
    int rc = fast_box_box_face(FP_HOLDER_OBJ, sb.hp, sb.hR, sb.hh, fo.gp, fo.gR, sb.oh, m.pairs[FP_HOLDER_OBJ].margin, f.co);
    if (rc < 0) bail |= BAIL_OBJ;
    int nf = con4_count(cf), nh = rc > 0 ? con4_count(f.co) : 0;
    if (nf > 0 && nh > 0) bail |= BAIL_OBJ;
    if (nh == 0) f.co = cf;
    float dd[3]; v3sub(dd, fd.gp, fo.gp);
    if (!(v3norm(dd) > m.geoms[FG_DEST].rbound + m.geoms[FG_OBJ].rbound + m.pairs[FP_DEST_OBJ].margin)) bail |= BAIL_OBJ;
  }
  JSYNC3();
  fast_plane_box(FP_FLOOR_DEST, GF.pos, GF.R, fd.gp, fd.gR, sb.dh, m.pairs[FP_FLOOR_DEST].margin, f.cd);
  JSYNC2();
  // ---- chain dynamics at (q, v), reference point = EE position of the previous forward
  float O[3]; v3copy(O, s.eepos);
  fast_chain<true>(s.qpos, s.qvel, O, &sb, f.c);
  bail |= f.c.near;
  f.bail = bail;
  JSYNC2();
  // ---- actuation + smooth forces of the chain (fwd_actuation, mj_passive)
  const float gcmd = ramp_value(s, it);
#pragma unroll
  for (int i = 0; i < 6; i++) f.fsm[i] = fminf(fmaxf(s.u[i], m.motor_fr[i][0]), m.motor_fr[i][1]) - f.c.bias[i] - m.ch_damping[i] * s.qvel[i];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const float* A = m.finger_act[i];
    float cc = fminf(fmaxf(gcmd, A[1]), A[2]);
    float fr = A[0] * cc - A[0] * s.qpos[6 + i];
    f.fsm[6 + i] = fminf(fmaxf(fr, A[3]), A[4]) - f.c.bias[6 + i] - m.ch_damping[6 + i] * s.qvel[6 + i];
  }
}
// solved: bit mask of the blocks (1 chain, 2 object, 4 dest) whose qacc_g / fcon_g come from the cooperative path.
// Returns BAIL_* bits (0 = fine); osc_ret: result of the OSC for the next substep (1 done, 3 done via pinv, 0 bail, 2 deferred into *od).
template <bool COOP>
JHD int sub_finish(FastSim& s, SubFwd& f, int it, int solved, const float* qacc_g, const float* fcon_g, int osc_mode, OscDefer* od, int* osc_ret) {
  const float h = MDL.dt;
  int dead = 0;
  float acc[9];
  *osc_ret = 1;
  if (it == 49) {  // what the next env step needs from this forward
    if (solved & 1) { for (int i = 0; i < 9; i++) s.warm[i] = qacc_g[i]; }
    else { chain_solve(f.c, 0.0f, f.fsm, acc); for (int i = 0; i < 9; i++) s.warm[i] = acc[i]; }
#pragma unroll
    for (int i = 0; i < 9; i++) { s.lagq[i] = s.qpos[i]; s.lagv[i] = s.qvel[i]; }
  }
  // ---- mj_Euler, chain (implicit joint damping)
  if (solved & 1) { for (int i = 0; i < 9; i++) f.fsm[i] += fcon_g[i]; }
  if (!chain_solve(f.c, h, f.fsm, acc)) dead |= BAIL_NAN;
  JSYNC3();
#pragma unroll
  for (int i = 0; i < 9; i++) { s.qvel[i] += h * acc[i]; s.qpos[i] += h * s.qvel[i]; }
  // ---- free boxes: contact solve + Euler
  if (solved & 2) free_integrate(f.fo, qacc_g + 9, fcon_g + 9, s.qpos + 9, s.qvel + 9, s.warm + 9);
  else fast_free_step(f.fo, f.co, s.qpos + 9, s.qvel + 9, s.warm + 9);
  JSYNC3();
  if (solved & 4) free_integrate(f.fd, qacc_g + 15, fcon_g + 15, s.qpos + 16, s.qvel + 15, s.warm + 15);
  else fast_free_step(f.fd, f.cd, s.qpos + 16, s.qvel + 15, s.warm + 15);
  JSYNC2();
  // ---- controller output for the next substep from this forward's mjData and the fresh joint velocity
  v3copy(s.eepos, f.c.eepos);
#pragma unroll
  for (int i = 0; i < 4; i++) s.eeq[i] = f.c.eeq[i];
  if (it < 49) { *osc_ret = fast_osc(f.c, s.tq, s.target, s.qvel, s.u, osc_mode, od); if (*osc_ret == 0) dead |= BAIL_OSC; }
  return dead;
}
// env_mujoco.py:128-150 after the 50 substeps.  cause: 0 running, 1 success, 2 time-out, 3 singular exit, 4 drop / -20 exit.
// Returns 0 when the result is not finite (caller: NaN guard).
JHD int step_end(const JacoCfg& cfg, FastSim& s, int env_id, float* obs, float* reward, int* done, int* succ, float* wb, int* cause) {
  get_observation(cfg, s, env_id, obs);
  float r = get_reward(cfg, s), bonus, w; int sc, d;
  s.current_steps += 1;
  if (s.current_steps < cfg.task_max_steps) d = terminal_inspection(cfg, s, &bonus, &w, &sc);
  else { d = 1; bonus = -10; w = 0; sc = 0; }
  *cause = !d ? 0 : (sc ? 1 : (s.current_steps >= cfg.task_max_steps ? 2 : (JPI - 0.1f < s.qpos[2] && s.qpos[2] < JPI + 0.1f ? 3 : (bonus < 0 ? 4 : 5))));
  r += bonus;
  if (!(r == r)) return 0;
  *reward = r; *done = d; *succ = sc; *wb = w;
  s.rng_counter += 1;
  return 1;
}

// ------------------------------------------------------------------------------------------- in-kernel reset
// reset() of util:92-174 for the tasks whose reset needs no settle loop, without the general mj_forward: valid when no
// arm geom can touch anything at the sampled pose (then every touch sensor reads 0, which is all the first observation
// takes from that forward).  Returns 0 when the proximity tests cannot prove it: the caller hands the env to the
// general reset kernel.
JHD JNOINLINE int fast_reset(const JacoCfg& cfg, FastSim& s, int env_id, float* obs) {
  if (cfg.task == JACO_TASK_PLACING || cfg.task == JACO_TASK_GRASPING || cfg.task == JACO_TASK_RELEASING || cfg.task == JACO_TASK_CARRYING) return 0;
  reset_sample(cfg, s, env_id);
  if (cfg.task != JACO_TASK_REACHING) {   // set_obj_xyz(obj_goal, [0,0,0,0]) (util:119-121, mujoco.py:217-227)
    for (int i = 0; i < 3; i++) s.qpos[9 + i] = s.obj_goal[i];
    for (int i = 0; i < 4; i++) s.qpos[12 + i] = 0;
    for (int i = 9; i < JACO_NV; i++) s.qvel[i] = 0;
  }
  SceneBoxes sb; scene_static(sb);
  FreeBody fo, fd;
  free_body_setup(10, FG_OBJ, s.qpos + 9, s.qvel + 9, fo);    // also normalises the quaternions like mj_kinematics
  free_body_setup(11, FG_DEST, s.qpos + 16, s.qvel + 15, fd);
  v3copy(s.objp, fo.x);
  v3copy(sb.op, fo.gp); v3copy(sb.dp, fd.gp);
  for (int i = 0; i < 9; i++) { sb.oR[i] = fo.gR[i]; sb.dR[i] = fd.gR[i]; }
  ChainOut c;
  const float O0[3] = {0, 0, 0.5f};
  fast_chain<true>(s.qpos, s.qvel, O0, &sb, c);
  if (c.near) {   // bounding spheres overlap: the owner-thread narrow phase decides; only a real arm contact needs the general reset
    LocalSink ks; ks.ncon = 0;
    fast_arm_collide(c, fo, fd, ks);
    if (ks.ncon > 0) return 0;
  }
  for (int k = 0; k < 9; k++) if (MDL.ch_limited[k] && (s.qpos[k] - MDL.ch_range[k][0] < 0 || MDL.ch_range[k][1] - s.qpos[k] < 0)) return 0;
  v3copy(s.eepos, c.eepos);
  for (int i = 0; i < 4; i++) s.eeq[i] = c.eeq[i];
  for (int i = 0; i < 9; i++) { s.lagq[i] = s.qpos[i]; s.lagv[i] = s.qvel[i]; }
  s.touch_now = 0;
  if (cfg.task == JACO_TASK_REACHING) { for (int i = 0; i < 6; i++) s.target[i] = s.reach_goal[i]; }
  else { get_gripper_pose(s); for (int i = 0; i < 6; i++) s.target[i] = s.gripper_pose[i]; }
  get_observation(cfg, s, env_id, obs);
  s.rng_counter += 1;
  s.accum_rew = 0;
  return 1;
}

// One env step on the fast path, all 50 substeps in one call.  Returns FAST_OK with all outputs written into s / obs / ..., or
// FAST_BAIL (reason bits in *why) with s in an unspecified state -- the caller must not store it.
// COOP = true: nothing bails for contact reasons; environments that need the general contact handling get it from
// the warp-cooperative path (jaco_coop.cuh) through the per-warp workspace ws, substep by substep.  All lanes of a
// warp must then call this function together (lanes without an environment pass active = false).
template <bool COOP>
JHD int fast_env_step(const JacoCfg& cfg, FastSim& s, int env_id, const float* action, float* obs, float* reward, int* done, int* succ, float* wb, int* why,
                      CoopWS* ws = nullptr, bool active = true, int* cause = nullptr, float* lst = nullptr, int* busy = nullptr) {
  *why = 0;
  int anyB = 0;
  int dead = 0;  // COOP: bail reason latched, the thread keeps pace with its warp until the step ends
  { const int rb = step_begin(cfg, s, action, COOP ? OSC_INLINE : JACO_NOCOOP_OSC_MODE, nullptr);
    if (!rb) { *why = BAIL_OSC; if (!COOP) return FAST_BAIL; dead |= BAIL_OSC; }
    if (rb == 3 && lst && active) lst[STAT_PINV] += 1; }
  SceneBoxes sb;
  scene_static(sb);
  for (int it = 0; it < 50; it++) {
    JSYNC1();
    SubFwd f;
    sub_forward<COOP>(s, sb, it, f);
    const int bail = f.bail;
    if (!COOP && bail) {
      // pass 1: "near" is only a bounding-sphere statement.  Run the owner-thread narrow phase over the arm pairs and leave the
      // fast path only when there is something to solve (an arm contact, an active limit, non-simple scene contacts).
      bool leave = (bail & (BAIL_LIMIT | BAIL_OBJ | BAIL_DEST)) != 0;
      if (!leave) { LocalSink ks; ks.ncon = 0; fast_arm_collide(f.c, f.fo, f.fd, ks); leave = ks.ncon > 0; }
      if (leave) { *why = bail; return FAST_BAIL; }
    }
    // ---- general contact handling on demand (warp-cooperative)
    int solved = 0;   // bit mask of the blocks (1 chain, 2 object, 4 dest) the cooperative path solved for this env
    float qacc_g[JACO_NV], fcon_g[JACO_NV];
    if (COOP) {
      const bool need = active && !dead && bail != 0;
      if (need && busy) *busy += 1;   // substeps that left the plain fast path (placement key of the step kernel)
      const int flags = ((bail & (BAIL_OBJ | BAIL_DEST)) ? 0 : COOP_SIMPLE_OBJ) | ((bail & (BAIL_LIMIT | BAIL_OBJ | BAIL_DEST)) ? COOP_FORCE_B : 0);
      // own broad + narrow phase over the arm pairs; stage B only for lanes with something to solve
      LocalSink ks; ks.ncon = 0;
      unsigned long long cand_def = 0;   // != 0: the narrow phase of this env's arm pairs is deferred to the cooperative path
      if (!(JACO_TIMING_SKIP & 2) && need && (bail & (BAIL_HAND | BAIL_FOREARM))) {
        fast_arm_collide(f.c, f.fo, f.fd, ks, &cand_def);
        if (ks.ncon < 0) ks.ncon = 0; else cand_def = 0;
      }
      const bool needB = !(JACO_TIMING_SKIP & 1) && need && (ks.ncon > 0 || cand_def != 0 || (flags & COOP_FORCE_B));
      FAST_WSTAT(0);
      if (need) { FAST_WSTAT(1); if (bail & BAIL_LIMIT) FAST_WSTAT(8); if (bail & (BAIL_HAND | BAIL_FOREARM)) FAST_WSTAT(9); if (bail & (BAIL_OBJ | BAIL_DEST)) FAST_WSTAT(10); }
#ifdef __CUDA_ARCH__
      // Block-wide task queue: every lane that needs stage B takes a slot; the block's warps then serve the slots one
      // task per warp and round, so the block waits for ceil(tasks / warps) cooperative solves instead of for the
      // warp that happens to own the most.  Owners publish into the workspace of the serving warp (all owners of a
      // round at the same time) and read their results back from it.
      __shared__ int q_n;
      const int mywarp = threadIdx.x >> 5;
      const int nwarps = min((int)(blockDim.x >> 5), JACO_COOP_NWS);   // serving warps = workspaces of the block
      CoopWS* const wsbase = ws - mywarp;
      if (threadIdx.x == 0) q_n = 0;
      __syncthreads();
      int slot = -1;
      if (needB) slot = atomicAdd(&q_n, 1);
      __syncthreads();
      const int ntask = q_n;
      for (int base = 0; base < ntask; base += nwarps) {
        const bool mine = slot >= base && slot < base + nwarps;
        JPROF_T(t1);
        if (mine) {
          CoopWS& wt = wsbase[slot - base]; coop_publish(wt, s, f.c, f.fo, f.fd, f.fsm, it == 49); coop_publish_contacts(wt, ks, flags);
          if (cand_def) { wt.cand[0] |= (unsigned)(cand_def & 0xFFFFFFFFull); wt.cand[1] = (unsigned)(cand_def >> 32); }   // lanes collide these pairs
        }
        __syncthreads();
        JPROF_T(t2);
        if (mywarp < nwarps && base + mywarp < ntask) {
          const int sv = coop_stageB(*ws);
          if ((threadIdx.x & 31) == 0) ws->solved = sv;
          JPROF_T(t3);
          JPROF_ADD(8, 1); JPROF_ADD(9, t2 - t1); JPROF_ADD(2, t3 - t2); if (sv) JPROF_ADD(3, 1);
        }
        __syncthreads();
        if (mine) {
          const CoopWS& wt = wsbase[slot - base];
          solved = wt.solved;
          if (solved) { for (int i = 0; i < JACO_NV; i++) { qacc_g[i] = wt.qacc[i]; fcon_g[i] = wt.fcon[i]; } if (it == 49) s.touch_now = touch_class(wt.touch); }
#if defined(JACO_DIAG) && defined(__CUDA_ARCH__)
          if (busy) { busy[1] += wt.iters; busy[2] = max(busy[2], wt.ncon); busy[3] += (wt.iters >= JACO_SOLVER_MAXIT); }   // per-env solve cost (diagnostics)
#endif
          if (lst) {   // workload / cap counters (jaco_stats)
            lst[STAT_STAGEB] += 1;
            if (wt.ncon >= JACO_MAXCON || ks.ncon >= LocalSink::MAXC) lst[STAT_CONCAP] += 1;   // incl. the owner thread's 24 arm contacts
            if (wt.iters >= JACO_SOLVER_MAXIT) lst[STAT_ITERCAP] += 1;
            if (wt.fingerobj) lst[STAT_FINGEROBJ] += 1;
          }
        }
        if (base + nwarps < ntask) __syncthreads();   // the next round's owners overwrite the workspaces
      }
#else
      if (needB) {
        FAST_WSTAT(2); anyB = 1;
        coop_publish(*ws, s, f.c, f.fo, f.fd, f.fsm, it == 49); coop_publish_contacts(*ws, ks, flags);
        solved = coop_stageB(*ws);
        if (lst) { lst[STAT_STAGEB] += 1; if (ws->ncon >= JACO_MAXCON) lst[STAT_CONCAP] += 1; if (ws->iters >= JACO_SOLVER_MAXIT) lst[STAT_ITERCAP] += 1; if (ws->fingerobj) lst[STAT_FINGEROBJ] += 1; }
        if (solved) { FAST_WSTAT(3); FAST_WSTAT_ADD(4, ws->iters); FAST_WSTAT_ADD(5, ws->ncon); if (solved & 1) FAST_WSTAT(6); if (ws->ncon > 0) FAST_WSTAT(7);
          for (int i = 0; i < JACO_NV; i++) { qacc_g[i] = ws->qacc[i]; fcon_g[i] = ws->fcon[i]; } if (it == 49) s.touch_now = touch_class(ws->touch); }
      }
#endif
    }
    JSYNC2();
    int osc_ret;
    const int dd = sub_finish<COOP>(s, f, it, solved, qacc_g, fcon_g, COOP ? OSC_INLINE : JACO_NOCOOP_OSC_MODE, nullptr, &osc_ret);
    if (dd) { *why = dd; if (!COOP) return FAST_BAIL; dead |= dd; }
    if (osc_ret == 3 && lst && active && !dead) lst[STAT_PINV] += 1;
  }
  if (anyB) FAST_WSTAT(14);
  (void)anyB;
  if (dead) { *why = dead; return FAST_BAIL; }
  bool bad = false;
#pragma unroll
  for (int k = 0; k < JACO_NQ; k++) bad = bad || !(fabsf(s.qpos[k]) < 1e6f);
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) bad = bad || !(fabsf(s.qvel[k]) < 1e6f);
  if (bad) { *why = BAIL_NAN; return FAST_BAIL; }
  int cz = 0;
  if (!step_end(cfg, s, env_id, obs, reward, done, succ, wb, &cz)) { *why = BAIL_NAN; return FAST_BAIL; }
  if (cause) *cause = cz;
  return FAST_OK;
}

// ------------------------------------------------------------------------------------------- state I/O
#define ST(r) state[(size_t)(r) * N + i]
JHD void fast_load(const float* state, int N, int i, FastSim& s) {
#pragma unroll
  for (int k = 0; k < JACO_NQ; k++) s.qpos[k] = ST(k);
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) s.qvel[k] = ST(JACO_NQ + k);
  const int A = JACO_NQ + JACO_NV;
#pragma unroll
  for (int k = 0; k < 6; k++) s.target[k] = ST(A + AUX_TARGET + k);
  s.grip = ST(A + AUX_GRIP); s.ramp0 = s.ramp1 = s.grip;
  s.action_in = (int)ST(A + AUX_ACTION_IN); s.picked = (int)ST(A + AUX_PICKED);
  s.current_steps = (int)ST(A + AUX_CUR_STEPS); s.num_episodes = (int)ST(A + AUX_NUM_EPISODES);
#pragma unroll
  for (int k = 0; k < 3; k++) { s.obj_goal[k] = ST(A + AUX_OBJ_GOAL + k); s.dest_goal[k] = ST(A + AUX_DEST_GOAL + k); }
#pragma unroll
  for (int k = 0; k < 6; k++) s.reach_goal[k] = ST(A + AUX_REACH_GOAL + k);
  s.rng_counter = (uint32_t)ST(A + AUX_RNG) + ((uint32_t)ST(A + AUX_RNG_HI) << 24); s.touch_index = (int)ST(A + AUX_TOUCH);
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) s.warm[k] = ST(A + AUX_WARM + k);
#pragma unroll
  for (int k = 0; k < 9; k++) { s.lagq[k] = ST(A + AUX_LAGQ + k); s.lagv[k] = ST(A + AUX_LAGV + k); }
  s.accum_rew = ST(A + AUX_ACCUM_REW);
}
JHD void fast_store(float* state, int N, int i, const FastSim& s) {
#pragma unroll
  for (int k = 0; k < JACO_NQ; k++) ST(k) = s.qpos[k];
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) ST(JACO_NQ + k) = s.qvel[k];
  const int A = JACO_NQ + JACO_NV;
#pragma unroll
  for (int k = 0; k < 6; k++) ST(A + AUX_TARGET + k) = s.target[k];
  ST(A + AUX_GRIP) = s.grip; ST(A + AUX_ACTION_IN) = (float)s.action_in; ST(A + AUX_PICKED) = (float)s.picked;
  ST(A + AUX_CUR_STEPS) = (float)s.current_steps; ST(A + AUX_NUM_EPISODES) = (float)s.num_episodes;
  ST(A + AUX_RNG) = (float)(s.rng_counter & 0xFFFFFFu); ST(A + AUX_RNG_HI) = (float)(s.rng_counter >> 24); ST(A + AUX_TOUCH) = (float)s.touch_index;
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) ST(A + AUX_WARM + k) = s.warm[k];
#pragma unroll
  for (int k = 0; k < 9; k++) { ST(A + AUX_LAGQ + k) = s.lagq[k]; ST(A + AUX_LAGV + k) = s.lagv[k]; }
  ST(A + AUX_ACCUM_REW) = s.accum_rew;
}
#undef ST

// Fast-path entry for env i.  Returns FAST_BAIL (state untouched, nothing written) or FAST_OK with the state
// advanced and every output row written; *needs_reset = 1 when the episode ended and auto_reset is on (the
// caller re-initialises the env with the general reset and overwrites obs_row with the new episode's first obs).
// COOP = true: the warp-cooperative general path serves this env's contacts (all 32 lanes call together; lanes
// with active = false shadow env i without writing anything).
template <bool COOP>
JHD int entry_step_fast(const JacoCfg& cfg, float* state, int N, int i, int env_id, const float* action_row, float* obs_row,
                        float* reward, unsigned char* done, unsigned char* succ, float* wb, float* term_obs_row, int auto_reset,
                        float* local_stats, int* needs_reset, int* why, CoopWS* ws = nullptr, bool active = true, int* busy = nullptr) {
  FastSim s;
  fast_load(state, N, i, s);
  float obs[JACO_NOBS], r, w; int d, sc;
  *needs_reset = 0;
  int cause = 0;
  int rc = fast_env_step<COOP>(cfg, s, env_id, action_row, obs, &r, &d, &sc, &w, why, ws, active, &cause, local_stats, busy);
  if (!active) return FAST_OK;
  if (rc != FAST_OK) return FAST_BAIL;
  s.accum_rew += r;
  local_stats[STAT_STEPS] += 1;
  if (d) {
    local_stats[STAT_EPISODES] += 1; local_stats[STAT_SUCCESS] += sc; local_stats[STAT_RETURN] += s.accum_rew; local_stats[STAT_LENGTH] += s.current_steps;
    if (cause == 2) local_stats[STAT_TIMEOUT] += 1;
    else if (cause == 3) local_stats[STAT_SINGULAR] += 1;
    else if (cause == 4) local_stats[STAT_DROP] += 1;
  }
  if (term_obs_row) for (int k = 0; k < JACO_NOBS; k++) term_obs_row[k] = obs[k];
  if (d && auto_reset) {
    // reset() in place when the proximity tests prove the general mj_forward is not needed at the sampled pose
    // (util:92-174); otherwise the caller queues the env for the general reset kernel
    fast_store(state, N, i, s);
    if (!fast_reset(cfg, s, env_id, obs)) { *needs_reset = 1; fast_load(state, N, i, s); }
    else {   // the new episode's goals (rows fast_store leaves alone: they are constant within an episode)
      const int A = JACO_NQ + JACO_NV;
#pragma unroll
      for (int k = 0; k < 3; k++) { state[(size_t)(A + AUX_OBJ_GOAL + k) * N + i] = s.obj_goal[k]; state[(size_t)(A + AUX_DEST_GOAL + k) * N + i] = s.dest_goal[k]; }
#pragma unroll
      for (int k = 0; k < 6; k++) state[(size_t)(A + AUX_REACH_GOAL + k) * N + i] = s.reach_goal[k];
    }
  }
  fast_store(state, N, i, s);
  for (int k = 0; k < JACO_NOBS; k++) obs_row[k] = obs[k];
  *reward = r; *done = (unsigned char)d; *succ = (unsigned char)sc; *wb = w;
  return FAST_OK;
}
