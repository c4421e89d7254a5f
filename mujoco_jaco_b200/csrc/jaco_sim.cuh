// jaco_sim.cuh -- one Jaco environment advanced by one thread (fp32).
//
// This is the product's arithmetic: everything JacoMujocoEnv.step does between the action and the
// (obs, reward, done) triple (reference env_script/env_mujoco.py:116-150, env_mujoco_util.py:73-90,223-271,
// 312-431,470-646) including the MuJoCo-2.0-semantics physics substep and the abr_control OSC it calls
// (SURVEY.md App. A).  Specialised to the compiled Jaco scene: a 9-DoF tree (6 arm joints + 3 finger hinges)
// plus two free boxes, collision primitives from the model blob.  All functions are __host__ __device__ so the
// same source also compiles into tests/host_harness.cpp, which lets the CPU test-suite exercise the kernel
// logic against the oracle where no GPU exists; the shipped library only ever launches the CUDA kernels.
#pragma once
#include <math.h>
#include <stdint.h>
#include "jaco_model.h"

#ifdef __CUDACC__
#define JHD __host__ __device__ inline
#define JNOINLINE __noinline__   // cold / shared helpers: one copy of the code (instruction-cache footprint), ABI call
static __constant__ JacoModel c_jaco_model;   // uniform (same address in every lane) accesses: the per-thread paths
static __device__ JacoModel g_jaco_model_dev; // lane-varying accesses (pairs / geoms indexed by lane): the cooperative path
#else
#define JHD inline
#define JNOINLINE __attribute__((noinline))
#endif
static JacoModel g_jaco_model_host;
#ifdef __CUDA_ARCH__
#define MDL c_jaco_model
#define MDLG g_jaco_model_dev
#else
#define MDL g_jaco_model_host
#define MDLG g_jaco_model_host
#endif

#define JPI 3.14159265358979323846f
#define JMINVAL 1e-15f
#define JACO_GRASP_RESET_MAX 4000     // bound (substeps) of each of the reference's two unbounded grasping-reset servo loops (util:135-167)
#ifndef JACO_SOLVER_MAXIT
#define JACO_SOLVER_MAXIT 16           // Newton iterations per component (fast path / cooperative path).  Round 1 used 8: on the in-contact
                                       // snapshots 23 % of the cooperative solves then stopped on the cap and 3 of 4096 env steps flipped touch / success
                                       // against the oracle; from 12 upwards the results are identical to cap 40 (the solves still running there
                                       // stagnate at float32 resolution: 4.6 % at 16, 1.2 % at 40).  jaco_stats counts the cap hits.
#endif
#ifndef JACO_SOLVER_RTOL
#define JACO_SOLVER_RTOL 2e-8f         // cooperative path: mj_solNewton's improvement test (stop when scale * (cost - newcost) < tolerance) in float32
                                       // terms -- stop once the exact line search lowers the cost by less than RTOL x the cost (a third of a float32
                                       // ulp): such iterates only move within rounding.  Without it 3.8 % of the cooperative solves on the in-contact
                                       // snapshots (1 % on the roll-out) ran into the iteration cap, each a 13-iteration tail that its whole block waits
                                       // for: 18.1 -> 16.4 ms per step; 2e-8 / 5e-8 keep the GPU parity suite green, 1e-7 puts one carrying env 7e-5 off,
                                       // 1e-6 flips one touch class of the 4096 snapshots (profiles/r02_solver_rtol_ab.log).
#endif
#define JACO_SOLVER_MAXIT_GENERAL 40   // general path: all coupled blocks are solved as one system, which needs more
#define PK(i, j) ((i) * ((i) + 1) / 2 + (j))  // packed lower-triangular index, j <= i

// ------------------------------------------------------------------------------------------- small math
JHD void v3set(float* r, float a, float b, float c) { r[0] = a; r[1] = b; r[2] = c; }
JHD void v3copy(float* r, const float* a) { r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; }
JHD void v3add(float* r, const float* a, const float* b) { r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; }
JHD void v3sub(float* r, const float* a, const float* b) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
JHD void v3addscl(float* r, const float* a, const float* b, float s) { r[0] = a[0] + s * b[0]; r[1] = a[1] + s * b[1]; r[2] = a[2] + s * b[2]; }
JHD float v3dot(const float* a, const float* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
JHD void v3cross(float* r, const float* a, const float* b) {
  float x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
JHD float v3norm(const float* a) { return sqrtf(v3dot(a, a)); }
JHD void m3mulv(float* r, const float* M, const float* v) {
  float x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2], y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2], z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
JHD void m3tmulv(float* r, const float* M, const float* v) {
  float x = M[0] * v[0] + M[3] * v[1] + M[6] * v[2], y = M[1] * v[0] + M[4] * v[1] + M[7] * v[2], z = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
JHD void m3mul(float* r, const float* A, const float* B) {
  float t[9];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
#pragma unroll
  for (int i = 0; i < 9; i++) r[i] = t[i];
}
// symmetric 3x3 stored xx yy zz xy xz yz times vector
JHD void s3mulv(float* r, const float* S, const float* v) {
  float x = S[0] * v[0] + S[3] * v[1] + S[4] * v[2], y = S[3] * v[0] + S[1] * v[1] + S[5] * v[2], z = S[4] * v[0] + S[5] * v[1] + S[2] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
JHD void quat_mul(float* r, const float* a, const float* b) {
  float w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  float x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  float y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  float z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
JHD void quat_normalize(float* q) {
  float n = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < JMINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
  else { float s = 1.0f / n; q[0] *= s; q[1] *= s; q[2] *= s; q[3] *= s; }
}
JHD void quat2mat(float* M, const float* q) {
  float w = q[0], x = q[1], y = q[2], z = q[3];
  M[0] = w * w + x * x - y * y - z * z; M[1] = 2 * (x * y - w * z); M[2] = 2 * (x * z + w * y);
  M[3] = 2 * (x * y + w * z); M[4] = w * w - x * x + y * y - z * z; M[5] = 2 * (y * z - w * x);
  M[6] = 2 * (x * z - w * y); M[7] = 2 * (y * z + w * x); M[8] = w * w - x * x - y * y + z * z;
}
// Gohlke 'rxyz' conventions
JHD void euler_rxyz_to_quat(float* q, float a, float b, float g) {
  float sa, ca, sb, cb, sg, cg;
  sincosf(0.5f * a, &sa, &ca); sincosf(0.5f * b, &sb, &cb); sincosf(0.5f * g, &sg, &cg);
  float qx[4] = {ca, sa, 0, 0}, qy[4] = {cb, 0, sb, 0}, qz[4] = {cg, 0, 0, sg}, t[4];
  quat_mul(t, qx, qy); quat_mul(q, t, qz);
}
JHD void mat_to_euler_rxyz(float* e, const float* M) {
  float cy = sqrtf(M[8] * M[8] + M[5] * M[5]);
  if (cy > 8.881784197001252e-16f) { e[0] = atan2f(-M[5], M[8]); e[1] = atan2f(M[2], cy); e[2] = atan2f(-M[1], M[0]); }
  else { e[0] = 0; e[1] = atan2f(M[2], cy); e[2] = atan2f(M[3], M[4]); }
}
JHD void quat_to_euler_rxyz(float* e, const float* q) {
  float n = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3], M[9];
  if (n < 8.881784197001252e-16f) { float I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; mat_to_euler_rxyz(e, I); return; }
  float s = sqrtf(1.0f / n), qq[4] = {q[0] * s, q[1] * s, q[2] * s, q[3] * s};
  quat2mat(M, qq);
  mat_to_euler_rxyz(e, M);
}
JHD void get_rotation(float* out, float roll, float pitch, float yaw, const float* vec, int inv) {  // util:448-468
  float sr, cr, sp, cp, sy, cy;
  sincosf(roll, &sr, &cr); sincosf(pitch, &sp, &cp); sincosf(yaw, &sy, &cy);
  float Rx[9] = {1, 0, 0, 0, cr, -sr, 0, sr, cr}, Ry[9] = {cp, 0, sp, 0, 1, 0, -sp, 0, cp}, Rz[9] = {cy, -sy, 0, sy, cy, 0, 0, 0, 1}, T[9], Rm[9];
  m3mul(T, Ry, Rz); m3mul(Rm, Rx, T);
  if (inv) m3tmulv(out, Rm, vec); else m3mulv(out, Rm, vec);
}

// packed-lower Cholesky of an N x N SPD matrix held in registers (fully unrolled); returns 0 on a bad pivot
template <int N>
JHD int chol_packed(float* A) {
  int ok = 1;
#pragma unroll
  for (int j = 0; j < N; j++) {
    float d = A[PK(j, j)];
#pragma unroll
    for (int k = 0; k < j; k++) d -= A[PK(j, k)] * A[PK(j, k)];
    if (!(d > 0)) { ok = 0; d = 1; }
    float r = sqrtf(d), ir = 1.0f / r;
    A[PK(j, j)] = r;
#pragma unroll
    for (int i = j + 1; i < N; i++) {
      float s = A[PK(i, j)];
#pragma unroll
      for (int k = 0; k < j; k++) s -= A[PK(i, k)] * A[PK(j, k)];
      A[PK(i, j)] = s * ir;
    }
  }
  return ok;
}
template <int N>
JHD void chol_solve_packed(const float* L, float* x) {
#pragma unroll
  for (int i = 0; i < N; i++) {
    float s = x[i];
#pragma unroll
    for (int k = 0; k < i; k++) s -= L[PK(i, k)] * x[k];
    x[i] = s / L[PK(i, i)];
  }
#pragma unroll
  for (int i = N - 1; i >= 0; i--) {
    float s = x[i];
#pragma unroll
    for (int k = i + 1; k < N; k++) s -= L[PK(k, i)] * x[k];
    x[i] = s / L[PK(i, i)];
  }
}
// runtime-size variants (solver Hessian lives in local memory)
JHD int chol_packed_rt(float* A, int n) {
  int ok = 1;
  for (int j = 0; j < n; j++) {
    float d = A[PK(j, j)];
    for (int k = 0; k < j; k++) d -= A[PK(j, k)] * A[PK(j, k)];
    if (!(d > 0)) { ok = 0; d = 1; }
    float r = sqrtf(d), ir = 1.0f / r;
    A[PK(j, j)] = r;
    for (int i = j + 1; i < n; i++) {
      float s = A[PK(i, j)];
      for (int k = 0; k < j; k++) s -= A[PK(i, k)] * A[PK(j, k)];
      A[PK(i, j)] = s * ir;
    }
  }
  return ok;
}
JHD void chol_solve_packed_rt(const float* L, float* x, int n) {
  for (int i = 0; i < n; i++) { float s = x[i]; for (int k = 0; k < i; k++) s -= L[PK(i, k)] * x[k]; x[i] = s / L[PK(i, i)]; }
  for (int i = n - 1; i >= 0; i--) { float s = x[i]; for (int k = i + 1; k < n; k++) s -= L[PK(k, i)] * x[k]; x[i] = s / L[PK(i, i)]; }
}

// ------------------------------------------------------------------------------------------- Philox4x32-10
JHD void philox4x32(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t* out) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// ------------------------------------------------------------------------------------------- per-env working set
struct JacoCfg { int task, variant, obs_mode, task_max_steps; uint64_t seed; int binary; };   // binary: the variants' sparse picking reward (cis:56,459,637)

struct JacoSim {
  static const int MAXC = JACO_MAXCON_GENERAL;
  static JHD const JacoModel& model() { return MDL; }
  // state
  float qpos[JACO_NQ], qvel[JACO_NV], warm[JACO_NV], ctrl[JACO_NU];
  // kinematics (index = body id: 0 world, 1..9 chain, 10 object, 11 dest)
  float bx[JACO_NBODY][3], bR[JACO_NBODY][9], cw[JACO_NBODY][3], Iw[JACO_NBODY][6];
  float ax[9][3];                                    // joint axes, world
  float bw[JACO_NBODY][3], bv[JACO_NBODY][3];        // body angular velocity / linear velocity of the body origin
  float eeq[4], eepos[3], Jee[6][6];
  // dynamics
  float Mc[45], Mf[2][21], Lf[2][21];                // chain / free-body inertia (packed), free-body Cholesky factors
  float bias[JACO_NV], fsm[JACO_NV], asm_[JACO_NV], qacc[JACO_NV], fcon[JACO_NV];
  // contacts
  int ncon, cpair[JACO_MAXCON_GENERAL];
  float cpos[JACO_MAXCON_GENERAL][3], cn[JACO_MAXCON_GENERAL][3], cdist[JACO_MAXCON_GENERAL];
  float cD[JACO_MAXCON_GENERAL], cA[JACO_MAXCON_GENERAL][6];         // 1/R of the contact's rows; aref components (A0, B1..B5)
  float ce[JACO_MAXCON_GENERAL][6], cp[JACO_MAXCON_GENERAL][6];      // (J a - aref) components at the iterate; J p components
  float cf0[JACO_MAXCON_GENERAL];                            // total normal force (touch sensors)
  // joint limits
  int nlim, ldof[5];
  float lsign[5], lD[5], laref[5], le[5], lp[5];
  int solver_iters;
  // env logic
  float target[6], grip, ramp0, ramp1;               // gripper command ramps linearly ramp0 -> ramp1 over the 50 substeps
  int action_in, picked, touch_index, num_episodes, current_steps;
  uint32_t rng_counter;
  float obj_goal[3], dest_goal[3], reach_goal[6], gripper_pose[6];
  float touch[20];
  float lagq[9], lagv[9];                             // chain state the working set (mjData) was evaluated at
  float accum_rew;
};

JHD const float* obj_pos(const JacoSim& s) { return s.bx[10]; }  // lagged xpos of the object body

// chain topology helpers
JHD int ch_parent(int k) { return k < 6 ? k - 1 : 5; }
JHD bool ch_has_dof(int k, int j) { return k < 6 ? (j <= k) : (j < 6 || j == k); }

// ------------------------------------------------------------------------------------------- kinematics
// mj_kinematics + mj_comPos for the chain bodies (quaternion composition like MuJoCo's xquat)
JHD void fk_chain(JacoSim& s) {
  const JacoModel& m = MDL;
  v3set(s.bx[0], 0, 0, 0);
  { float I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
#pragma unroll
    for (int i = 0; i < 9; i++) s.bR[0][i] = I[i]; }
  float Q5[4] = {1, 0, 0, 0}, Qp[4] = {1, 0, 0, 0};
#pragma unroll
  for (int k = 0; k < 9; k++) {
    int pb = ch_parent(k) + 1, b = k + 1;
    float t[3], qb[4], qj[4], q[4], sh, chh;
    m3mulv(t, s.bR[pb], m.ch_pos[k]); v3add(s.bx[b], s.bx[pb], t);
    const float* qpar = (k <= 5) ? Qp : Q5;
    quat_mul(qb, qpar, m.ch_quat[k]);
    sincosf(0.5f * (s.qpos[k] - m.ch_ref[k]), &sh, &chh);
    qj[0] = chh; qj[1] = m.ch_axis[k][0] * sh; qj[2] = m.ch_axis[k][1] * sh; qj[3] = m.ch_axis[k][2] * sh;
    quat_mul(q, qb, qj);
    quat2mat(s.bR[b], q);
    m3mulv(s.ax[k], s.bR[b], m.ch_axis[k]);
    if (k <= 5) { Qp[0] = q[0]; Qp[1] = q[1]; Qp[2] = q[2]; Qp[3] = q[3]; }
    if (k == 5) { Q5[0] = q[0]; Q5[1] = q[1]; Q5[2] = q[2]; Q5[3] = q[3]; }
  }
  s.eeq[0] = Q5[0]; s.eeq[1] = Q5[1]; s.eeq[2] = Q5[2]; s.eeq[3] = Q5[3];
  { float t[3]; m3mulv(t, s.bR[6], m.ee_off); v3add(s.eepos, s.bx[6], t); }
}
JHD void fk_free(JacoSim& s) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    float* q = s.qpos + 9 + 7 * f;
    quat_normalize(q + 3);
    v3copy(s.bx[10 + f], q); quat2mat(s.bR[10 + f], q + 3);
  }
}
// world COM and world inertia (about the COM) of bodies b0..b1-1
JHD void com_inertia(JacoSim& s, int b0, int b1) {
  const JacoModel& m = MDL;
  for (int b = b0; b < b1; b++) {
    float t[3]; const float* R = s.bR[b]; const float* I = m.inertia[b];
    m3mulv(t, R, m.com[b]); v3add(s.cw[b], s.bx[b], t);
    float S[9] = {I[0], I[3], I[4], I[3], I[1], I[5], I[4], I[5], I[2]}, T[9];
    m3mul(T, R, S);  // T = R I
    s.Iw[b][0] = T[0] * R[0] + T[1] * R[1] + T[2] * R[2];
    s.Iw[b][1] = T[3] * R[3] + T[4] * R[4] + T[5] * R[5];
    s.Iw[b][2] = T[6] * R[6] + T[7] * R[7] + T[8] * R[8];
    s.Iw[b][3] = T[0] * R[3] + T[1] * R[4] + T[2] * R[5];
    s.Iw[b][4] = T[0] * R[6] + T[1] * R[7] + T[2] * R[8];
    s.Iw[b][5] = T[3] * R[6] + T[4] * R[7] + T[5] * R[8];
  }
}
// body "twists" (angular part, linear part at the body origin) generated by a generalised vector x[21];
// x = qvel gives velocities, x = qacc gives the J*qacc part of accelerations.
JHD void body_twists(const JacoSim& s, const float* x, float W[JACO_NBODY][3], float V[JACO_NBODY][3]) {
  v3set(W[0], 0, 0, 0); v3set(V[0], 0, 0, 0);
#pragma unroll
  for (int k = 0; k < 9; k++) {
    int pb = ch_parent(k) + 1, b = k + 1;
    float r[3], t[3];
    v3sub(r, s.bx[b], s.bx[pb]); v3cross(t, W[pb], r);
    v3add(V[b], V[pb], t);
    v3addscl(W[b], W[pb], s.ax[k], x[k]);
  }
#pragma unroll
  for (int f = 0; f < 2; f++) {
    int b = 10 + f, d = 9 + 6 * f;
    v3copy(V[b], x + d); m3mulv(W[b], s.bR[b], x + d + 3);
  }
}
// transpose of the above: generalised forces from per-body (force F at the origin, torque N about the origin)
JHD void wrench_to_joint(const JacoSim& s, float F[JACO_NBODY][3], float N[JACO_NBODY][3], float* out) {
#pragma unroll
  for (int k = 8; k >= 0; k--) {
    int pb = ch_parent(k) + 1, b = k + 1;
    out[k] = v3dot(s.ax[k], N[b]);
    if (pb > 0) {
      float r[3], t[3];
      v3sub(r, s.bx[b], s.bx[pb]); v3cross(t, r, F[b]);
      v3add(F[pb], F[pb], F[b]); v3add(N[pb], N[pb], N[b]); v3add(N[pb], N[pb], t);
    }
  }
#pragma unroll
  for (int f = 0; f < 2; f++) {
    int b = 10 + f, d = 9 + 6 * f;
    v3copy(out + d, F[b]); m3tmulv(out + d + 3, s.bR[b], N[b]);
  }
}

// ------------------------------------------------------------------------------------------- inertia + bias
// composite-rigid-body inertia of the 9-DoF tree (mj_crb), packed lower
JHD void crba_chain(JacoSim& s) {
  const JacoModel& m = MDL;
  float cm[9], cc[9][3], cI[9][6];
#pragma unroll
  for (int k = 0; k < 9; k++) {
    cm[k] = m.mass[k + 1]; v3copy(cc[k], s.cw[k + 1]);
#pragma unroll
    for (int i = 0; i < 6; i++) cI[k][i] = s.Iw[k + 1][i];
  }
#pragma unroll
  for (int k = 8; k >= 1; k--) {
    int p = ch_parent(k);
    float mt = cm[p] + cm[k], c[3], da[3], db[3];
#pragma unroll
    for (int i = 0; i < 3; i++) c[i] = (cm[p] * cc[p][i] + cm[k] * cc[k][i]) / mt;
    v3sub(da, cc[p], c); v3sub(db, cc[k], c);
    float na = v3dot(da, da), nb = v3dot(db, db);
    cI[p][0] += cI[k][0] + cm[p] * (na - da[0] * da[0]) + cm[k] * (nb - db[0] * db[0]);
    cI[p][1] += cI[k][1] + cm[p] * (na - da[1] * da[1]) + cm[k] * (nb - db[1] * db[1]);
    cI[p][2] += cI[k][2] + cm[p] * (na - da[2] * da[2]) + cm[k] * (nb - db[2] * db[2]);
    cI[p][3] += cI[k][3] - cm[p] * da[0] * da[1] - cm[k] * db[0] * db[1];
    cI[p][4] += cI[k][4] - cm[p] * da[0] * da[2] - cm[k] * db[0] * db[2];
    cI[p][5] += cI[k][5] - cm[p] * da[1] * da[2] - cm[k] * db[1] * db[2];
    cm[p] = mt; v3copy(cc[p], c);
  }
#pragma unroll
  for (int i = 0; i < 45; i++) s.Mc[i] = 0;
#pragma unroll
  for (int i = 0; i < 9; i++) {
    float d[3], f[3], h[3];
    v3sub(d, cc[i], s.bx[i + 1]); v3cross(f, s.ax[i], d);
    f[0] *= cm[i]; f[1] *= cm[i]; f[2] *= cm[i];
    s3mulv(h, cI[i], s.ax[i]);
#pragma unroll
    for (int j = 0; j < 9; j++) {
      if (j > i || !ch_has_dof(i, j)) continue;
      float r[3], t[3];
      v3sub(r, cc[i], s.bx[j + 1]); v3cross(t, r, f);
      s.Mc[PK(i, j)] = s.ax[j][0] * (h[0] + t[0]) + s.ax[j][1] * (h[1] + t[1]) + s.ax[j][2] * (h[2] + t[2]);
    }
  }
}
JHD void crba_free(JacoSim& s) {
  const JacoModel& m = MDL;
#pragma unroll
  for (int fb = 0; fb < 2; fb++) {
    int b = 10 + fb; float* M = s.Mf[fb]; const float* R = s.bR[b];
    float d[3]; v3sub(d, s.cw[b], s.bx[b]);
#pragma unroll
    for (int i = 0; i < 21; i++) M[i] = 0;
    M[PK(0, 0)] = M[PK(1, 1)] = M[PK(2, 2)] = m.mass[b];
    float r[3][3], f[3][3], h[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
      v3set(r[i], R[i], R[3 + i], R[6 + i]);
      v3cross(f[i], r[i], d); f[i][0] *= m.mass[b]; f[i][1] *= m.mass[b]; f[i][2] *= m.mass[b];
      float t[3]; s3mulv(h[i], s.Iw[b], r[i]); v3cross(t, d, f[i]); v3add(h[i], h[i], t);
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
#pragma unroll
      for (int k = 0; k < 3; k++) M[PK(3 + i, k)] = f[i][k];
#pragma unroll
      for (int j = 0; j <= i; j++) M[PK(3 + i, 3 + j)] = v3dot(r[j], h[i]);
    }
  }
}
// mj_rne(flg_acc = 0) on the chain; also leaves the body velocities in bw / bv
JHD void rne_chain(JacoSim& s) {
  const JacoModel& m = MDL;
  float al[10][3], ao[10][3], f[10][3], n[10][3];
  v3set(s.bw[0], 0, 0, 0); v3set(s.bv[0], 0, 0, 0); v3set(al[0], 0, 0, 0); v3set(ao[0], 0, 0, -m.gravz);
#pragma unroll
  for (int k = 0; k < 9; k++) {
    int b = k + 1, p = ch_parent(k) + 1;
    float wj[3], r[3], t[3], t2[3], c[3], ac[3], Iw[3];
    v3set(wj, s.ax[k][0] * s.qvel[k], s.ax[k][1] * s.qvel[k], s.ax[k][2] * s.qvel[k]);
    v3add(s.bw[b], s.bw[p], wj);
    v3cross(t, s.bw[p], wj); v3add(al[b], al[p], t);
    v3sub(r, s.bx[b], s.bx[p]);
    v3cross(t2, s.bw[p], r); v3add(s.bv[b], s.bv[p], t2);
    v3cross(t, al[p], r); v3cross(t2, s.bw[p], t2);
    ao[b][0] = ao[p][0] + t[0] + t2[0]; ao[b][1] = ao[p][1] + t[1] + t2[1]; ao[b][2] = ao[p][2] + t[2] + t2[2];
    v3sub(c, s.cw[b], s.bx[b]);
    v3cross(t, al[b], c); v3cross(t2, s.bw[b], c); v3cross(t2, s.bw[b], t2);
    ac[0] = ao[b][0] + t[0] + t2[0]; ac[1] = ao[b][1] + t[1] + t2[1]; ac[2] = ao[b][2] + t[2] + t2[2];
    f[b][0] = m.mass[b] * ac[0]; f[b][1] = m.mass[b] * ac[1]; f[b][2] = m.mass[b] * ac[2];
    s3mulv(Iw, s.Iw[b], s.bw[b]); v3cross(t, s.bw[b], Iw); s3mulv(t2, s.Iw[b], al[b]);
    v3cross(n[b], c, f[b]);
    n[b][0] += t[0] + t2[0]; n[b][1] += t[1] + t2[1]; n[b][2] += t[2] + t2[2];
  }
#pragma unroll
  for (int k = 8; k >= 0; k--) {
    int b = k + 1, p = ch_parent(k) + 1;
    s.bias[k] = v3dot(s.ax[k], n[b]);
    if (p > 0) {
      float r[3], t[3];
      v3sub(r, s.bx[b], s.bx[p]); v3cross(t, r, f[b]);
      v3add(f[p], f[p], f[b]); v3add(n[p], n[p], n[b]); v3add(n[p], n[p], t);
    }
  }
}
JHD void rne_free(JacoSim& s) {
  const JacoModel& m = MDL;
#pragma unroll
  for (int fb = 0; fb < 2; fb++) {
    int b = 10 + fb, d = 9 + 6 * fb;
    float c[3], t[3], F[3], Iw[3], N[3], no[3];
    m3mulv(s.bw[b], s.bR[b], s.qvel + d + 3); v3copy(s.bv[b], s.qvel + d);
    v3sub(c, s.cw[b], s.bx[b]);
    v3cross(t, s.bw[b], c); v3cross(t, s.bw[b], t);
    F[0] = m.mass[b] * t[0]; F[1] = m.mass[b] * t[1]; F[2] = m.mass[b] * (t[2] - m.gravz);
    s3mulv(Iw, s.Iw[b], s.bw[b]); v3cross(N, s.bw[b], Iw);
    v3cross(no, c, F); v3add(no, no, N);
    v3copy(s.bias + d, F); m3tmulv(s.bias + d + 3, s.bR[b], no);
  }
}
JHD void ee_jacobian(JacoSim& s) {
#pragma unroll
  for (int j = 0; j < 6; j++) {
    float r[3], c[3];
    v3sub(r, s.eepos, s.bx[j + 1]); v3cross(c, s.ax[j], r);
    s.Jee[0][j] = c[0]; s.Jee[1][j] = c[1]; s.Jee[2][j] = c[2];
    s.Jee[3][j] = s.ax[j][0]; s.Jee[4][j] = s.ax[j][1]; s.Jee[5][j] = s.ax[j][2];
  }
}
// the part of mj_forward the controller reads (mujoco_config.py J/M/g/R/Tx): chain only
JHD void forward_chain_smooth(JacoSim& s) {
  fk_chain(s);
  com_inertia(s, 1, 10);
  crba_chain(s);
  rne_chain(s);
  ee_jacobian(s);
}

// ------------------------------------------------------------------------------------------- collision
// squared distance from point c to an oriented box (centre bp, rows of bR = world-from-box, half sizes h)
JHD float point_box_dist2(const float* c, const float* bp, const float* bR, const float* h) {
  float d[3], l[3];
  v3sub(d, c, bp); m3tmulv(l, bR, d);
  float e0 = fmaxf(fabsf(l[0]) - h[0], 0.0f), e1 = fmaxf(fabsf(l[1]) - h[1], 0.0f), e2 = fmaxf(fabsf(l[2]) - h[2], 0.0f);
  return e0 * e0 + e1 * e1 + e2 * e2;
}
JHD void make_frame(const float* n, float* t1, float* t2) {  // mju_makeFrame
  if (fabsf(n[1]) < 0.5f) v3set(t1, 0, 1, 0); else v3set(t1, 0, 0, 1);
  float d = v3dot(n, t1); v3addscl(t1, t1, n, -d);
  float inv = 1.0f / v3norm(t1); t1[0] *= inv; t1[1] *= inv; t1[2] *= inv;
  v3cross(t2, n, t1);
}
template <class K>
JHD void add_contact(K& s, int pair, const float* pos, const float* n, float dist) {
  if (s.ncon >= K::MAXC) return;
  int c = s.ncon++;
  v3copy(s.cpos[c], pos); v3copy(s.cn[c], n); s.cdist[c] = dist; s.cpair[c] = pair;
}
template <class K>
JHD void collide_plane_box(K& s, int pair, const float* pp, const float* pR, const float* bp, const float* bR, const float* size, float margin) {
  float n[3] = {pR[2], pR[5], pR[8]}, d[3];
  v3sub(d, bp, pp);
  float dist = v3dot(d, n);
  int cnt = 0;
  for (int i = 0; i < 8; i++) {
    float v[3] = {(i & 1) ? size[0] : -size[0], (i & 2) ? size[1] : -size[1], (i & 4) ? size[2] : -size[2]}, corner[3], pos[3];
    m3mulv(corner, bR, v);
    float ld = v3dot(n, corner);
    if (dist + ld > margin || ld > 0) continue;
    float cd = dist + ld;
    v3add(pos, bp, corner); v3addscl(pos, pos, n, -cd * 0.5f);
    add_contact(s, pair, pos, n, cd);
    if (++cnt >= 4) return;
  }
}
JHD int sphere_box(const float* c, float r, const float* bp, const float* bR, const float* h, float margin, float* normal, float* dist, float* pos) {
  float d[3], l[3], q[3];
  v3sub(d, c, bp); m3tmulv(l, bR, d);
  int inside = 1;
#pragma unroll
  for (int i = 0; i < 3; i++) { q[i] = l[i]; if (q[i] > h[i]) { q[i] = h[i]; inside = 0; } else if (q[i] < -h[i]) { q[i] = -h[i]; inside = 0; } }
  float nl[3];
  if (!inside) {
    float df[3]; v3sub(df, l, q);
    float dd = v3norm(df);
    if (dd - r > margin) return 0;
    float inv = 1.0f / dd;
    nl[0] = df[0] * inv; nl[1] = df[1] * inv; nl[2] = df[2] * inv;
    *dist = dd - r;
  } else {
    int best = 0; float bd = h[0] - fabsf(l[0]);
#pragma unroll
    for (int i = 1; i < 3; i++) { float t = h[i] - fabsf(l[i]); if (t < bd) { bd = t; best = i; } }
    v3set(nl, 0, 0, 0);
    float sg = l[best] >= 0 ? 1.0f : -1.0f;
    if (best == 0) { nl[0] = sg; q[0] = sg * h[0]; } else if (best == 1) { nl[1] = sg; q[1] = sg * h[1]; } else { nl[2] = sg; q[2] = sg * h[2]; }
    *dist = -bd - r;
  }
  float qw[3];
  m3mulv(normal, bR, nl); m3mulv(qw, bR, q); v3add(qw, qw, bp);
  v3addscl(pos, qw, normal, *dist * 0.5f);
  return 1;
}
JHD float seg_box_dfdt(const float* p0, const float* d, const float* h, float t) {
  float s = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) { float x = p0[i] + t * d[i]; if (x > h[i]) s += (x - h[i]) * d[i]; else if (x < -h[i]) s += (x + h[i]) * d[i]; }
  return s;
}
template <class K>
JHD void collide_capsule_box(K& s, int pair, int box_first, const float* cpn, const float* cR, const float* csize, const float* bp, const float* bR, const float* bsize, float margin) {
  float ax[3] = {cR[2], cR[5], cR[8]}, p0[3], p1[3], r = csize[0];
  v3addscl(p0, cpn, ax, -csize[1]); v3addscl(p1, cpn, ax, csize[1]);
  float nrm[3], dist, pos[3], sgn = box_first ? 1.0f : -1.0f;
  int hit0 = sphere_box(p0, r, bp, bR, bsize, margin, nrm, &dist, pos);
  if (hit0) { float n2[3] = {sgn * nrm[0], sgn * nrm[1], sgn * nrm[2]}; add_contact(s, pair, pos, n2, dist); }
  int hit1 = sphere_box(p1, r, bp, bR, bsize, margin, nrm, &dist, pos);
  if (hit1) { float n2[3] = {sgn * nrm[0], sgn * nrm[1], sgn * nrm[2]}; add_contact(s, pair, pos, n2, dist); }
  if (hit0 && hit1) return;
  float d[3], l0[3], dl[3], t[3];
  v3sub(d, p1, p0); v3sub(t, p0, bp); m3tmulv(l0, bR, t); m3tmulv(dl, bR, d);
  float lo = 0, hi = 1, flo = seg_box_dfdt(l0, dl, bsize, 0), fhi = seg_box_dfdt(l0, dl, bsize, 1);
  if (flo >= 0 || fhi <= 0) return;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int sg = -1; sg <= 1; sg += 2) {
      if (fabsf(dl[i]) < JMINVAL) continue;
      float tb = (sg * bsize[i] - l0[i]) / dl[i];
      if (tb <= lo || tb >= hi) continue;
      float fb = seg_box_dfdt(l0, dl, bsize, tb);
      if (fb <= 0) { lo = tb; flo = fb; } else { hi = tb; fhi = fb; }
    }
  float ts = (fhi - flo > JMINVAL) ? lo - flo * (hi - lo) / (fhi - flo) : 0.5f * (lo + hi);
  if (ts <= 0.02f || ts >= 0.98f) return;
  float c[3]; v3addscl(c, p0, d, ts);
  if (sphere_box(c, r, bp, bR, bsize, margin, nrm, &dist, pos)) { float n2[3] = {sgn * nrm[0], sgn * nrm[1], sgn * nrm[2]}; add_contact(s, pair, pos, n2, dist); }
}
template <class K>
JHD void collide_plane_capsule(K& s, int pair, const float* pp, const float* pR, const float* cpn, const float* cR, const float* csize, float margin) {
  float n[3] = {pR[2], pR[5], pR[8]}, ax[3] = {cR[2], cR[5], cR[8]};
#pragma unroll
  for (int sg = -1; sg <= 1; sg += 2) {
    float c[3], d[3], pos[3];
    v3addscl(c, cpn, ax, sg * csize[1]); v3sub(d, c, pp);
    float dist = v3dot(d, n) - csize[0];
    if (dist > margin) continue;
    v3addscl(pos, c, n, -(csize[0] + dist * 0.5f));
    add_contact(s, pair, pos, n, dist);
  }
}
// box-box: SAT over 15 axes, then face clipping (<= 8 points) or one edge-edge point (DESIGN.md "collision")
template <class K>
JHD void collide_box_box(K& s, int pair, const float* pa, const float* Ra, const float* ha, const float* pb, const float* Rb, const float* hb, float margin) {
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
  float best = -1e30f; int code = -1;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    float sp = fabsf(ta[i]) - (ha[i] + hb[0] * AC[i][0] + hb[1] * AC[i][1] + hb[2] * AC[i][2]);
    if (sp > margin) return;
    if (sp > best) { best = sp; code = i; }
  }
#pragma unroll
  for (int j = 0; j < 3; j++) {
    float sp = fabsf(tb[j]) - (hb[j] + ha[0] * AC[0][j] + ha[1] * AC[1][j] + ha[2] * AC[2][j]);
    if (sp > margin) return;
    if (sp > best) { best = sp; code = 3 + j; }
  }
  float ebest = -1e30f; int ecode = -1; float eaxis[3] = {0, 0, 0};
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      float L[3]; v3cross(L, A[i], B[j]);
      float l = v3norm(L);
      if (l < 1e-6f) continue;
      const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      float ra = ha[i1] * AC[i2][j] + ha[i2] * AC[i1][j], rb = hb[j1] * AC[i][j2] + hb[j2] * AC[i][j1];
      float sp = (fabsf(v3dot(t, L)) - ra - rb) / l;
      if (sp > margin) return;
      if (sp > ebest) { ebest = sp; ecode = 3 * i + j; eaxis[0] = L[0] / l; eaxis[1] = L[1] / l; eaxis[2] = L[2] / l; }
    }
  if (ecode >= 0 && ebest * 1.05f > best) {
    int i = ecode / 3, j = ecode % 3;
    float n[3]; v3copy(n, eaxis);
    if (v3dot(n, t) < 0) { n[0] = -n[0]; n[1] = -n[1]; n[2] = -n[2]; }
    float pA[3], pB[3];
    v3copy(pA, pa); v3copy(pB, pb);
    for (int k = 0; k < 3; k++) if (k != i) { float sg = v3dot(n, A[k]) > 0 ? 1.0f : -1.0f; v3addscl(pA, pA, A[k], sg * ha[k]); }
    for (int k = 0; k < 3; k++) if (k != j) { float sg = v3dot(n, B[k]) > 0 ? -1.0f : 1.0f; v3addscl(pB, pB, B[k], sg * hb[k]); }
    float w[3]; v3sub(w, pA, pB);
    float b_ = C[i][j], d_ = v3dot(A[i], w), e_ = v3dot(B[j], w), den = 1 - b_ * b_;
    float u = (b_ * e_ - d_) / den, v = (e_ - b_ * d_) / den;
    float qa[3], qb[3], pos[3];
    v3addscl(qa, pA, A[i], u); v3addscl(qb, pB, B[j], v);
    pos[0] = 0.5f * (qa[0] + qb[0]); pos[1] = 0.5f * (qa[1] + qb[1]); pos[2] = 0.5f * (qa[2] + qb[2]);
    add_contact(s, pair, pos, n, ebest);
    return;
  }
  int refA = code < 3, axn = refA ? code : code - 3;
  const float *pr = refA ? pa : pb, *hr = refA ? ha : hb, *pi = refA ? pb : pa, *hi = refA ? hb : ha;
  float(*Rr)[3] = refA ? A : B; float(*Ri)[3] = refA ? B : A;
  float nr[3], sg = (refA ? ta[axn] : -tb[axn]) >= 0 ? 1.0f : -1.0f;
  nr[0] = sg * Rr[axn][0]; nr[1] = sg * Rr[axn][1]; nr[2] = sg * Rr[axn][2];
  int ia = 0; float bd = -1;
  for (int k = 0; k < 3; k++) { float dd = fabsf(v3dot(Ri[k], nr)); if (dd > bd) { bd = dd; ia = k; } }
  float si = v3dot(Ri[ia], nr) > 0 ? -1.0f : 1.0f;
  int i1 = (ia + 1) % 3, i2 = (ia + 2) % 3, r1 = (axn + 1) % 3, r2 = (axn + 2) % 3;
  float fc[3]; v3addscl(fc, pi, Ri[ia], si * hi[ia]);
  float poly[16][2], depth[16], tmp[16][2], tmpd[16]; int np = 4;
  const int sx[4] = {1, -1, -1, 1}, sy[4] = {1, 1, -1, -1};
  for (int k = 0; k < 4; k++) {
    float v[3], dlt[3];
    v3addscl(v, fc, Ri[i1], sx[k] * hi[i1]); v3addscl(v, v, Ri[i2], sy[k] * hi[i2]);
    v3sub(dlt, v, pr);
    poly[k][0] = v3dot(dlt, Rr[r1]); poly[k][1] = v3dot(dlt, Rr[r2]); depth[k] = v3dot(dlt, nr) - hr[axn];
  }
  for (int c = 0; c < 4; c++) {
    int dim = c >> 1; float sgn = (c & 1) ? -1.0f : 1.0f, lim = dim ? hr[r2] : hr[r1];
    int nn = 0;
    for (int k = 0; k < np; k++) {
      int k2 = (k + 1 == np) ? 0 : k + 1;
      float a = sgn * poly[k][dim] - lim, b2 = sgn * poly[k2][dim] - lim;
      if (a <= 0) { tmp[nn][0] = poly[k][0]; tmp[nn][1] = poly[k][1]; tmpd[nn] = depth[k]; nn++; }
      if ((a <= 0) != (b2 <= 0)) {
        float sp = a / (a - b2);
        tmp[nn][0] = poly[k][0] + sp * (poly[k2][0] - poly[k][0]); tmp[nn][1] = poly[k][1] + sp * (poly[k2][1] - poly[k][1]);
        tmpd[nn] = depth[k] + sp * (depth[k2] - depth[k]); nn++;
      }
    }
    np = nn; if (np == 0) return;
    for (int k = 0; k < np; k++) { poly[k][0] = tmp[k][0]; poly[k][1] = tmp[k][1]; depth[k] = tmpd[k]; }
  }
  float n[3] = {refA ? nr[0] : -nr[0], refA ? nr[1] : -nr[1], refA ? nr[2] : -nr[2]};
  int cnt = 0;
  for (int k = 0; k < np && cnt < 8; k++) {
    if (depth[k] > margin) continue;
    float pos[3];
    v3addscl(pos, pr, Rr[r1], poly[k][0]); v3addscl(pos, pos, Rr[r2], poly[k][1]);
    v3addscl(pos, pos, nr, hr[axn] + depth[k] * 0.5f);
    add_contact(s, pair, pos, n, depth[k]);
    cnt++;
  }
}
JHD void geom_world(const JacoSim& s, const JacoGeom& G, float* pos, float* R) {
  float t[3]; m3mulv(t, s.bR[G.body], G.pos); v3add(pos, s.bx[G.body], t);
  m3mul(R, s.bR[G.body], G.R);
}
#if defined(JACO_COLLISION_STATS) && !defined(__CUDACC__)
static long long g_cstats[16];
#define CSTAT(i) g_cstats[i]++
#else
#define CSTAT(i)
#endif
// one geom pair: group culling by the hand sphere hs, bounding-sphere cull, narrow phase into sink k.
// PS supplies the body poses (bx, bR); K receives the contacts (add_contact).
template <class PS, class K>
JHD void collide_pair(const PS& s, int p, const float* hs, K& k) {
  const JacoModel& m = PS::model();
  const JacoPair& P = m.pairs[p];
  const JacoGeom& G1 = m.geoms[P.g1]; const JacoGeom& G2 = m.geoms[P.g2];
  float p1[3], p2[3];
  { float u[3]; m3mulv(u, s.bR[G1.body], G1.pos); v3add(p1, s.bx[G1.body], u); m3mulv(u, s.bR[G2.body], G2.pos); v3add(p2, s.bx[G2.body], u); }
  if (P.group >= 1 && P.group <= 4) {
    const JacoGeom& O = (P.group == 1) ? G2 : G1; const float* po = (P.group == 1) ? p2 : p1;
    float d[3]; v3sub(d, hs, po);
    if (O.type != JACO_GEOM_PLANE) { if (v3norm(d) > m.hand_sphere[3] + O.rbound + P.margin) return; }
    else { if (d[2] > m.hand_sphere[3] + P.margin) return; }
    CSTAT(P.group);
  }
  if (G1.type != JACO_GEOM_PLANE && G2.type != JACO_GEOM_PLANE) { float d[3]; v3sub(d, p1, p2); if (v3norm(d) > G1.rbound + G2.rbound + P.margin) return; }
  CSTAT(5 + P.group);
  float R1[9], R2[9];
  m3mul(R1, s.bR[G1.body], G1.R); m3mul(R2, s.bR[G2.body], G2.R);
  if (G1.type == JACO_GEOM_PLANE && G2.type == JACO_GEOM_BOX) collide_plane_box(k, p, p1, R1, p2, R2, G2.size, P.margin);
  else if (G1.type == JACO_GEOM_PLANE && G2.type == JACO_GEOM_CAPSULE) collide_plane_capsule(k, p, p1, R1, p2, R2, G2.size, P.margin);
  else if (G1.type == JACO_GEOM_BOX && G2.type == JACO_GEOM_BOX) collide_box_box(k, p, p1, R1, G1.size, p2, R2, G2.size, P.margin);
  else if (G1.type == JACO_GEOM_BOX && G2.type == JACO_GEOM_CAPSULE) collide_capsule_box(k, p, 1, p2, R2, G2.size, p1, R1, G1.size, P.margin);
  else if (G1.type == JACO_GEOM_CAPSULE && G2.type == JACO_GEOM_BOX) collide_capsule_box(k, p, 0, p1, R1, G1.size, p2, R2, G2.size, P.margin);
}
JHD void collision(JacoSim& s) {
  const JacoModel& m = MDL;
  s.ncon = 0;
  CSTAT(0);
  float hs[3], t[3];
  m3mulv(t, s.bR[6], m.hand_sphere); v3add(hs, s.bx[6], t);
  for (int p = 0; p < JM_NPAIR; p++) collide_pair(s, p, hs, s);
}

// ------------------------------------------------------------------------------------------- constraints
JHD float impedance(const float* solimp, float x) {  // getimpedance, x = pos - margin
  float d0 = solimp[0], dw = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  if (d0 == dw || width <= JMINVAL) return 0.5f * (d0 + dw);
  float xx = fabsf(x) / width;
  if (xx >= 1) return dw;
  if (xx == 0) return d0;
  float y;
  if (power == 1) y = xx;
  else if (power == 2) {   // MuJoCo's default exponent: a plain square instead of powf (same value to the last bits that matter, ~60 fewer instructions per row)
    if (xx <= mid) { const float a = xx / mid; y = a * a * mid; }
    else { const float a = (1 - xx) / (1 - mid); y = 1 - a * a * (1 - mid); }
  }
  else if (xx <= mid) y = powf(xx / mid, power) * mid;
  else y = 1 - powf((1 - xx) / (1 - mid), power) * (1 - mid);
  return d0 + y * (dw - d0);
}
// k, b, imp, R for one scalar constraint (mj_makeImpedance)
JHD void row_params(const float* solref, const float* solimp, float pos, float margin, float diagApprox, float* k, float* b, float* imp, float* R) {
  float dt2 = 2 * MDL.dt, tc = solref[0] > dt2 ? solref[0] : dt2, dr = solref[1], dmax = solimp[1];
  *k = 1.0f / fmaxf(JMINVAL, dmax * dmax * tc * tc * dr * dr);
  *b = 2.0f / fmaxf(JMINVAL, dmax * tc);
  *imp = impedance(solimp, pos - margin);
  *R = fmaxf(JMINVAL, (1 - *imp) * diagApprox / *imp);
}
template <class S>
JHD int contact_dim(int c, const S& s) { return S::model().pairs[s.cpair[c]].condim; }
JHD float contact_mu(const JacoPair& P, int k) { return k <= 2 ? P.friction[0] : (k == 3 ? P.friction[1] : P.friction[2]); }  // k = 1..5

// relative point motion at a contact for body twists (W, V): u = linear (b2 - b1) at pos, w = angular (b2 - b1)
template <class S>
JHD void contact_rel(const S& s, int c, const float W[JACO_NBODY][3], const float V[JACO_NBODY][3], float* u, float* w) {
  const JacoPair& P = S::model().pairs[s.cpair[c]];
  int b1 = S::model().geoms[P.g1].body, b2 = S::model().geoms[P.g2].body;
  float r[3], t[3], u1[3], u2[3];
  v3sub(r, s.cpos[c], s.bx[b1]); v3cross(t, W[b1], r); v3add(u1, V[b1], t);
  v3sub(r, s.cpos[c], s.bx[b2]); v3cross(t, W[b2], r); v3add(u2, V[b2], t);
  v3sub(u, u2, u1); v3sub(w, W[b2], W[b1]);
}
// six frame components (n,t1,t2 of u ; n,t1,t2 of w)
template <class S>
JHD void contact_comp(const S& s, int c, const float* u, const float* w, float* out) {
  float t1[3], t2[3];
  make_frame(s.cn[c], t1, t2);
  out[0] = v3dot(s.cn[c], u); out[1] = v3dot(t1, u); out[2] = v3dot(t2, u);
  out[3] = v3dot(s.cn[c], w); out[4] = v3dot(t1, w); out[5] = v3dot(t2, w);
}
// mj_instantiateLimit + mj_instantiateContact + mj_makeImpedance + mj_referenceConstraint (pyramidal cones)
JHD void make_constraints(JacoSim& s) {
  const JacoModel& m = MDL;
  const float def_solref[2] = {0.02f, 1.0f}, def_solimp[5] = {0.9f, 0.95f, 0.001f, 0.5f, 2.0f};
  s.nlim = 0;
#pragma unroll
  for (int k = 0; k < 9; k++) {
    if (!m.ch_limited[k]) continue;
    float dlo = s.qpos[k] - m.ch_range[k][0], dhi = m.ch_range[k][1] - s.qpos[k];
    for (int side = 0; side < 2; side++) {
      float d = side ? dhi : dlo, sg = side ? -1.0f : 1.0f;
      if (d >= 0 || s.nlim >= 5) continue;
      float kk, b, imp, R;
      row_params(def_solref, def_solimp, d, 0, m.dof_invw[k], &kk, &b, &imp, &R);
      int r = s.nlim++;
      s.ldof[r] = k; s.lsign[r] = sg; s.lD[r] = 1.0f / R; s.laref[r] = -b * (sg * s.qvel[k]) - kk * imp * d;
    }
  }
  for (int c = 0; c < s.ncon; c++) {
    const JacoPair& P = m.pairs[s.cpair[c]];
    float u[3], w[3], comp[6], kk, b, imp, R0;
    contact_rel(s, c, s.bw, s.bv, u, w);
    contact_comp(s, c, u, w, comp);
    float mu0 = P.friction[0];
    float dA = (P.condim == 1) ? P.invw_tran : P.invw_tran + mu0 * mu0 * P.invw_tran;
    row_params(P.solref, P.solimp, s.cdist[c], P.margin, dA, &kk, &b, &imp, &R0);
    float Rr = (P.condim == 1) ? R0 : fmaxf(JMINVAL, 2 * mu0 * mu0 * R0);
    s.cD[c] = 1.0f / Rr;
    s.cA[c][0] = -b * comp[0] - kk * imp * (s.cdist[c] - P.margin);
#pragma unroll
    for (int k = 1; k < 6; k++) s.cA[c][k] = -b * comp[k];
  }
}
// evaluate residual components e = (J x) - aref pieces for all rows at generalised acceleration x.
// For a contact: e[0] = n-part minus A0, e[k] = k-th friction part minus B_k; row(k,+-) = e[0] +- mu_k e[k].
JHD void eval_rows(JacoSim& s, const float* x, bool is_dir) {
  float W[JACO_NBODY][3], V[JACO_NBODY][3];
  body_twists(s, x, W, V);
  for (int r = 0; r < s.nlim; r++) {
    float v = s.lsign[r] * x[s.ldof[r]];
    if (is_dir) s.lp[r] = v; else s.le[r] = v - s.laref[r];
  }
  for (int c = 0; c < s.ncon; c++) {
    float u[3], w[3], comp[6];
    contact_rel(s, c, W, V, u, w);
    contact_comp(s, c, u, w, comp);
    if (is_dir) { for (int k = 0; k < 6; k++) s.cp[c][k] = comp[k]; }
    else { for (int k = 0; k < 6; k++) s.ce[c][k] = comp[k] - s.cA[c][k]; }
  }
}
// constraint cost 0.5*D*jar^2 over active rows at step al along the stored direction (al = 0: at the iterate);
// also first and second derivative w.r.t. al
JHD void rows_cost(const JacoSim& s, float al, float* cost, float* d1, float* d2) {
  float c0 = 0, g1 = 0, g2 = 0;
  for (int r = 0; r < s.nlim; r++) {
    float j = s.le[r] + al * s.lp[r];
    if (j < 0) { c0 += 0.5f * s.lD[r] * j * j; g1 += s.lD[r] * j * s.lp[r]; g2 += s.lD[r] * s.lp[r] * s.lp[r]; }
  }
  for (int c = 0; c < s.ncon; c++) {
    const JacoPair& P = MDL.pairs[s.cpair[c]];
    float D = s.cD[c];
    if (P.condim == 1) {
      float j = s.ce[c][0] + al * s.cp[c][0], jp = s.cp[c][0];
      if (j < 0) { c0 += 0.5f * D * j * j; g1 += D * j * jp; g2 += D * jp * jp; }
      continue;
    }
    for (int k = 1; k < P.condim; k++) {
      float mu = contact_mu(P, k);
#pragma unroll
      for (int sg = 1; sg >= -1; sg -= 2) {
        float jp = s.cp[c][0] + sg * mu * s.cp[c][k];
        float j = s.ce[c][0] + sg * mu * s.ce[c][k] + al * jp;
        if (j < 0) { c0 += 0.5f * D * j * j; g1 += D * j * jp; g2 += D * jp * jp; }
      }
    }
  }
  *cost = c0; if (d1) *d1 = g1; if (d2) *d2 = g2;
}
// M * x for the block-diagonal inertia
JHD void mul_M(const JacoSim& s, const float* x, float* y) {
#pragma unroll
  for (int i = 0; i < 9; i++) {
    float a = 0;
#pragma unroll
    for (int j = 0; j < 9; j++) a += s.Mc[i >= j ? PK(i, j) : PK(j, i)] * x[j];
    y[i] = a;
  }
#pragma unroll
  for (int f = 0; f < 2; f++)
#pragma unroll
    for (int i = 0; i < 6; i++) {
      float a = 0;
#pragma unroll
      for (int j = 0; j < 6; j++) a += s.Mf[f][i >= j ? PK(i, j) : PK(j, i)] * x[9 + 6 * f + j];
      y[9 + 6 * f + i] = a;
    }
}
JHD float gauss_cost(const JacoSim& s, const float* a) {
  float d[JACO_NV], Md[JACO_NV], c = 0;
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) d[i] = a[i] - s.asm_[i];
  mul_M(s, d, Md);
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) c += 0.5f * Md[i] * d[i];
  return c;
}
// J^T f of the active rows at the iterate (uses ce / le); also fills cf0 (total normal force per contact)
JHD void rows_force(JacoSim& s, float* out) {
  float F[JACO_NBODY][3], N[JACO_NBODY][3];
  for (int b = 0; b < JACO_NBODY; b++) { v3set(F[b], 0, 0, 0); v3set(N[b], 0, 0, 0); }
  for (int c = 0; c < s.ncon; c++) {
    const JacoPair& P = MDL.pairs[s.cpair[c]];
    float D = s.cD[c], sc[6] = {0, 0, 0, 0, 0, 0};
    if (P.condim == 1) { float j = s.ce[c][0]; if (j < 0) sc[0] = -D * j; }
    else for (int k = 1; k < P.condim; k++) {
      float mu = contact_mu(P, k);
      float jp_ = s.ce[c][0] + mu * s.ce[c][k], jm_ = s.ce[c][0] - mu * s.ce[c][k];
      float fp = jp_ < 0 ? -D * jp_ : 0.0f, fm = jm_ < 0 ? -D * jm_ : 0.0f;
      sc[0] += fp + fm; sc[k] = mu * (fp - fm);
    }
    s.cf0[c] = sc[0];
    float t1[3], t2[3], Fc[3], Tc[3];
    make_frame(s.cn[c], t1, t2);
#pragma unroll
    for (int i = 0; i < 3; i++) { Fc[i] = sc[0] * s.cn[c][i] + sc[1] * t1[i] + sc[2] * t2[i]; Tc[i] = sc[3] * s.cn[c][i] + sc[4] * t1[i] + sc[5] * t2[i]; }
    int b1 = MDL.geoms[P.g1].body, b2 = MDL.geoms[P.g2].body;
    float r[3], t[3];
    v3sub(r, s.cpos[c], s.bx[b2]); v3cross(t, r, Fc);
    v3add(F[b2], F[b2], Fc); v3add(N[b2], N[b2], t); v3add(N[b2], N[b2], Tc);
    v3sub(r, s.cpos[c], s.bx[b1]); v3cross(t, r, Fc);
    v3sub(F[b1], F[b1], Fc); v3sub(N[b1], N[b1], t); v3sub(N[b1], N[b1], Tc);
  }
  wrench_to_joint(s, F, N, out);
  for (int r = 0; r < s.nlim; r++) if (s.le[r] < 0) out[s.ldof[r]] += s.lsign[r] * (-s.lD[r] * s.le[r]);
}
// dense Jacobian row pieces of one contact direction (dl linear, dr angular) -> j[21]
template <class S>
JHD void contact_row(const S& s, int c, const float* dl, const float* dr, float* j) {
  const JacoPair& P = S::model().pairs[s.cpair[c]];
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) j[i] = 0;
#pragma unroll
  for (int side = 0; side < 2; side++) {
    int b = side ? S::model().geoms[P.g2].body : S::model().geoms[P.g1].body; float sg = side ? 1.0f : -1.0f;
    if (b == 0) continue;
    if (b <= 9) {
      int k = b - 1;
#pragma unroll
      for (int d = 0; d < 9; d++) {
        if (!ch_has_dof(k, d)) continue;
        float r[3], t[3];
        v3sub(r, s.cpos[c], s.bx[d + 1]); v3cross(t, r, dl);
        j[d] += sg * (s.ax[d][0] * (t[0] + dr[0]) + s.ax[d][1] * (t[1] + dr[1]) + s.ax[d][2] * (t[2] + dr[2]));
      }
    } else {
      int d0 = 9 + 6 * (b - 10);
      float r[3], t[3], l[3];
      v3sub(r, s.cpos[c], s.bx[b]); v3cross(t, r, dl); v3add(t, t, dr);
      m3tmulv(l, s.bR[b], t);
      j[d0] += sg * dl[0]; j[d0 + 1] += sg * dl[1]; j[d0 + 2] += sg * dl[2];
      j[d0 + 3] += sg * l[0]; j[d0 + 4] += sg * l[1]; j[d0 + 5] += sg * l[2];
    }
  }
}
JHD int body_block(int b) { return b == 0 ? 0 : (b <= 9 ? 1 : (b == 10 ? 2 : 4)); }  // bit mask: chain, object, dest
// Newton on the primal problem min 0.5 (a-a_s)' M (a-a_s) + sum 0.5 D min(0, J a - aref)^2  (mj_solNewton
// semantics), matrix-free in the rows: J is never stored, rows are re-derived from contact
// geometry and body twists.  Fixed iteration cap with an fp32 convergence test.
JHD void solve_constraints(JacoSim& s) {
  const JacoModel& m = MDL;
  s.solver_iters = 0;
  if (s.nlim == 0 && s.ncon == 0) {
#pragma unroll
    for (int i = 0; i < JACO_NV; i++) { s.qacc[i] = s.asm_[i]; s.fcon[i] = 0; }
    return;
  }
  // the direction components enter rows_cost as al * cp with al = 0 before the first line search: they must be
  // finite (a stale Inf / NaN in this uninitialised working set would deactivate rows: found as a launched object)
  for (int c = 0; c < s.ncon; c++)
#pragma unroll
    for (int k = 0; k < 6; k++) s.cp[c][k] = 0;
#pragma unroll
  for (int r = 0; r < 5; r++) s.lp[r] = 0;
  int mask = s.nlim ? 1 : 0;
  for (int c = 0; c < s.ncon; c++) { const JacoPair& P = m.pairs[s.cpair[c]]; mask |= body_block(m.geoms[P.g1].body) | body_block(m.geoms[P.g2].body); }
  // compact index of every dof inside the active system
  int off[3], n = 0;
  off[0] = n; if (mask & 1) n += 9;
  off[1] = n; if (mask & 2) n += 6;
  off[2] = n; if (mask & 4) n += 6;
  float a[JACO_NV], aw[JACO_NV];
  // warm start (blocks without rows sit at qacc_smooth, their exact optimum)
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) { int blk = i < 9 ? 1 : (i < 15 ? 2 : 4); aw[i] = (mask & blk) ? s.warm[i] : s.asm_[i]; }
  float cw_, cs_, tmp;
  eval_rows(s, aw, false); rows_cost(s, 0, &cw_, 0, 0); cw_ += gauss_cost(s, aw);
  eval_rows(s, s.asm_, false); rows_cost(s, 0, &cs_, 0, 0);
  if (cw_ < cs_) {
#pragma unroll
    for (int i = 0; i < JACO_NV; i++) a[i] = aw[i];
    eval_rows(s, a, false);
  } else {
#pragma unroll
    for (int i = 0; i < JACO_NV; i++) a[i] = s.asm_[i];
  }
  float H[231], g[JACO_NV], p[JACO_NV], jf[JACO_NV];
  for (int it = 0; it < JACO_SOLVER_MAXIT_GENERAL; it++) {
    // gradient
    float d[JACO_NV];
#pragma unroll
    for (int i = 0; i < JACO_NV; i++) d[i] = a[i] - s.asm_[i];
    mul_M(s, d, g);
    rows_force(s, jf);
    // block-wise relative gradient test: each block (chain, object, dest) against its own force scale -- one global
    // test lets the 1280 kg destination box hide an unconverged 17 g object (same rule as jaco_coop.cuh)
    bool conv = true;
    {
      float gb[3] = {0, 0, 0}, fb[3] = {0, 0, 0};
#pragma unroll
      for (int i = 0; i < JACO_NV; i++) { const int b = i < 9 ? 0 : (i < 15 ? 1 : 2); fb[b] += jf[i] * jf[i] + g[i] * g[i]; g[i] -= jf[i]; gb[b] += g[i] * g[i]; }
#pragma unroll
      for (int b = 0; b < 3; b++) if ((mask >> b) & 1) { if (!(gb[b] <= 1e-12f * fb[b] || gb[b] < 1e-20f)) conv = false; }
    }
    if (conv) break;
    // Hessian over the active dofs
    for (int i = 0; i < n * (n + 1) / 2; i++) H[i] = 0;
    if (mask & 1) for (int i = 0; i < 9; i++) for (int j = 0; j <= i; j++) H[PK(off[0] + i, off[0] + j)] = s.Mc[PK(i, j)];
    if (mask & 2) for (int i = 0; i < 6; i++) for (int j = 0; j <= i; j++) H[PK(off[1] + i, off[1] + j)] = s.Mf[0][PK(i, j)];
    if (mask & 4) for (int i = 0; i < 6; i++) for (int j = 0; j <= i; j++) H[PK(off[2] + i, off[2] + j)] = s.Mf[1][PK(i, j)];
    for (int r = 0; r < s.nlim; r++) if (s.le[r] < 0) { int i = off[0] + s.ldof[r]; H[PK(i, i)] += s.lD[r]; }
    for (int c = 0; c < s.ncon; c++) {
      const JacoPair& P = m.pairs[s.cpair[c]];
      float D = s.cD[c], t1[3], t2[3], zero[3] = {0, 0, 0}, jn[JACO_NV], jk[JACO_NV], row[JACO_NV];
      int bm = body_block(m.geoms[P.g1].body) | body_block(m.geoms[P.g2].body);
      bool any = false;
      if (P.condim == 1) any = s.ce[c][0] < 0;
      else for (int k = 1; k < P.condim; k++) { float mu = contact_mu(P, k); if (s.ce[c][0] + mu * s.ce[c][k] < 0 || s.ce[c][0] - mu * s.ce[c][k] < 0) any = true; }
      if (!any) continue;
      make_frame(s.cn[c], t1, t2);
      contact_row(s, c, s.cn[c], zero, jn);
      int nk = P.condim == 1 ? 1 : P.condim - 1;
      for (int k = 1; k <= nk; k++) {
        float mu = 0;
        if (P.condim > 1) {
          mu = contact_mu(P, k);
          const float* dirv = (k == 1 || k == 4) ? t1 : ((k == 2 || k == 5) ? t2 : s.cn[c]);
          if (k <= 2) contact_row(s, c, dirv, zero, jk); else contact_row(s, c, zero, dirv, jk);
        }
        for (int sg = 1; sg >= -1; sg -= 2) {
          if (P.condim == 1 && sg < 0) break;
          float jar = P.condim == 1 ? s.ce[c][0] : s.ce[c][0] + sg * mu * s.ce[c][k];
          if (!(jar < 0)) continue;
#pragma unroll
          for (int i = 0; i < JACO_NV; i++) row[i] = P.condim == 1 ? jn[i] : jn[i] + sg * mu * jk[i];
          // rank-1 update restricted to the blocks this contact touches
          for (int bi = 0; bi < 3; bi++) {
            if (!(bm & (1 << bi))) continue;
            int i0 = bi == 0 ? 0 : (bi == 1 ? 9 : 15), ni = bi == 0 ? 9 : 6;
            for (int bj = 0; bj <= bi; bj++) {
              if (!(bm & (1 << bj))) continue;
              int j0 = bj == 0 ? 0 : (bj == 1 ? 9 : 15), nj = bj == 0 ? 9 : 6;
              for (int i = 0; i < ni; i++) {
                float di = D * row[i0 + i];
                int jmax = (bi == bj) ? i + 1 : nj;
                for (int j = 0; j < jmax; j++) H[PK(off[bi] + i, off[bj] + j)] += di * row[j0 + j];
              }
            }
          }
        }
      }
    }
    chol_packed_rt(H, n);
    // p = -H^-1 g on the active dofs
    float pc[JACO_NV];
    for (int bi = 0; bi < 3; bi++) if (mask & (1 << bi)) { int i0 = bi == 0 ? 0 : (bi == 1 ? 9 : 15), ni = bi == 0 ? 9 : 6; for (int i = 0; i < ni; i++) pc[off[bi] + i] = -g[i0 + i]; }
    chol_solve_packed_rt(H, pc, n);
#pragma unroll
    for (int i = 0; i < JACO_NV; i++) p[i] = 0;
    for (int bi = 0; bi < 3; bi++) if (mask & (1 << bi)) { int i0 = bi == 0 ? 0 : (bi == 1 ? 9 : 15), ni = bi == 0 ? 9 : 6; for (int i = 0; i < ni; i++) p[i0 + i] = pc[off[bi] + i]; }
    // exact line search on the piecewise-quadratic phi(al): safeguarded 1-D Newton
    eval_rows(s, p, true);
    float Mp[JACO_NV], q1 = 0, q2 = 0;
    mul_M(s, p, Mp);
#pragma unroll
    for (int i = 0; i < JACO_NV; i++) { q2 += p[i] * Mp[i]; q1 += Mp[i] * d[i]; }
    float lo = 0, hi = -1, al = 0, d1, d2, c0;
    rows_cost(s, 0, &c0, &d1, &d2);
    float dlo = q1 + d1, slope0 = dlo;
    if (!(dlo < 0)) break;
    al = -dlo / (q2 + d2);
    for (int ls = 0; ls < 12; ls++) {
      rows_cost(s, al, &c0, &d1, &d2);
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
    for (int i = 0; i < JACO_NV; i++) { a[i] += al * p[i]; amax = fmaxf(amax, fabsf(a[i])); smax = fmaxf(smax, fabsf(al * p[i])); }
    // residuals at the new iterate: e += al * p-components
    for (int r = 0; r < s.nlim; r++) s.le[r] += al * s.lp[r];
    for (int c = 0; c < s.ncon; c++) for (int k = 0; k < 6; k++) s.ce[c][k] += al * s.cp[c][k];
    s.solver_iters = it + 1;
    if (smax <= 2e-7f * fmaxf(1.0f, amax)) break;
  }
  (void)tmp;
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) s.qacc[i] = a[i];
  rows_force(s, s.fcon);
}

// ------------------------------------------------------------------------------------------- sensors
JHD int ray_hits_site(const JacoSite& S, const float* spos, const float* sR, const float* pnt, const float* vec) {
  float d[3], p[3], v[3];
  v3sub(d, pnt, spos); m3tmulv(p, sR, d); m3tmulv(v, sR, vec);
  float t0 = -1e30f, t1 = 1e30f;
  int naxes = S.type == 0 ? 3 : 1;
  for (int k = 0; k < naxes; k++) {
    int i = S.type == 0 ? k : 2; float h = S.type == 0 ? S.size[i] : S.size[1];
    if (fabsf(v[i]) < JMINVAL) { if (fabsf(p[i]) > h) return 0; continue; }
    float ta = (-h - p[i]) / v[i], tb = (h - p[i]) / v[i];
    if (ta > tb) { float sw = ta; ta = tb; tb = sw; }
    if (ta > t0) t0 = ta;
    if (tb < t1) t1 = tb;
  }
  if (S.type == 1) {
    float a = v[0] * v[0] + v[1] * v[1], b = p[0] * v[0] + p[1] * v[1], c = p[0] * p[0] + p[1] * p[1] - S.size[0] * S.size[0];
    if (a < JMINVAL) { if (c > 0) return 0; }
    else {
      float disc = b * b - a * c;
      if (disc < 0) return 0;
      float sq = sqrtf(disc), ta = (-b - sq) / a, tb = (-b + sq) / a;
      if (ta > t0) t0 = ta;
      if (tb < t1) t1 = tb;
    }
  }
  return t1 >= (t0 > 0 ? t0 : 0);
}
JHD void touch_sensors(JacoSim& s) {  // mjSENS_TOUCH over the 20 sites
  const JacoModel& m = MDL;
  for (int k = 0; k < JM_NSITE; k++) s.touch[k] = 0;
  for (int c = 0; c < s.ncon; c++) {
    if (!(s.cf0[c] > 0)) continue;
    const JacoPair& P = m.pairs[s.cpair[c]];
    for (int k = 0; k < JM_NSITE; k++) {
      const JacoSite& S = m.sites[k];
      if (P.g1 != S.geom && P.g2 != S.geom) continue;
      float spos[3], sR[9], t[3], ray[3], sg = (P.g1 == S.geom) ? 1.0f : -1.0f;
      m3mulv(t, s.bR[S.body], S.pos); v3add(spos, s.bx[S.body], t); m3mul(sR, s.bR[S.body], S.R);
      ray[0] = sg * s.cn[c][0]; ray[1] = sg * s.cn[c][1]; ray[2] = sg * s.cn[c][2];
      if (ray_hits_site(S, spos, sR, s.cpos[c], ray)) s.touch[k] += s.cf0[c];
    }
  }
}

// ------------------------------------------------------------------------------------------- forward / step
JHD void fwd_actuation(const JacoSim& s, float* qfrc) {
  const JacoModel& m = MDL;
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) qfrc[i] = 0;
#pragma unroll
  for (int i = 0; i < 6; i++) qfrc[i] = fminf(fmaxf(s.ctrl[i], m.motor_fr[i][0]), m.motor_fr[i][1]);
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const float* A = m.finger_act[i];
    float c = fminf(fmaxf(s.ctrl[6 + i], A[1]), A[2]);
    float f = A[0] * c - A[0] * s.qpos[6 + i];
    qfrc[6 + i] = fminf(fmaxf(f, A[3]), A[4]);
  }
}
// mj_forward at (qpos, qvel, ctrl); sensors only when asked (they are read once per env step)
JHD void mj_forward(JacoSim& s, bool want_touch) {
  const JacoModel& m = MDL;
#pragma unroll
  for (int i = 0; i < 9; i++) { s.lagq[i] = s.qpos[i]; s.lagv[i] = s.qvel[i]; }
  fk_chain(s); fk_free(s);
  com_inertia(s, 1, JACO_NBODY);
  crba_chain(s); crba_free(s);
  rne_chain(s); rne_free(s);
  ee_jacobian(s);
  collision(s);
  float act[JACO_NV];
  fwd_actuation(s, act);
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) s.fsm[i] = act[i] - s.bias[i] - (i < 9 ? m.ch_damping[i] * s.qvel[i] : 0.0f);
  {  // qacc_smooth = M^-1 qfrc_smooth, block by block
    float L[45];
#pragma unroll
    for (int i = 0; i < 45; i++) L[i] = s.Mc[i];
    chol_packed<9>(L);
    float x[9];
#pragma unroll
    for (int i = 0; i < 9; i++) x[i] = s.fsm[i];
    chol_solve_packed<9>(L, x);
#pragma unroll
    for (int i = 0; i < 9; i++) s.asm_[i] = x[i];
#pragma unroll
    for (int f = 0; f < 2; f++) {
#pragma unroll
      for (int i = 0; i < 21; i++) s.Lf[f][i] = s.Mf[f][i];
      chol_packed<6>(s.Lf[f]);
      float y[6];
#pragma unroll
      for (int i = 0; i < 6; i++) y[i] = s.fsm[9 + 6 * f + i];
      chol_solve_packed<6>(s.Lf[f], y);
#pragma unroll
      for (int i = 0; i < 6; i++) s.asm_[9 + 6 * f + i] = y[i];
    }
  }
  make_constraints(s);
  solve_constraints(s);
  if (want_touch) touch_sensors(s);
}
// mj_Euler: implicit in joint damping, semi-implicit in position; then save the warm start (mj_advance)
JHD void mj_euler(JacoSim& s) {
  const JacoModel& m = MDL;
  float h = m.dt, a[JACO_NV];
  {
    float L[45];
#pragma unroll
    for (int i = 0; i < 45; i++) L[i] = s.Mc[i];
#pragma unroll
    for (int i = 0; i < 9; i++) L[PK(i, i)] += h * m.ch_damping[i];
    chol_packed<9>(L);
    float x[9];
#pragma unroll
    for (int i = 0; i < 9; i++) x[i] = s.fsm[i] + s.fcon[i];
    chol_solve_packed<9>(L, x);
#pragma unroll
    for (int i = 0; i < 9; i++) a[i] = x[i];
#pragma unroll
    for (int f = 0; f < 2; f++) {
      float y[6];
#pragma unroll
      for (int i = 0; i < 6; i++) y[i] = s.fsm[9 + 6 * f + i] + s.fcon[9 + 6 * f + i];
      chol_solve_packed<6>(s.Lf[f], y);
#pragma unroll
      for (int i = 0; i < 6; i++) a[9 + 6 * f + i] = y[i];
    }
  }
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) s.qvel[i] += h * a[i];
#pragma unroll
  for (int i = 0; i < 9; i++) s.qpos[i] += h * s.qvel[i];
#pragma unroll
  for (int f = 0; f < 2; f++) {
    float* q = s.qpos + 9 + 7 * f; const float* v = s.qvel + 9 + 6 * f;
    q[0] += h * v[0]; q[1] += h * v[1]; q[2] += h * v[2];
    float n = sqrtf(v[3] * v[3] + v[4] * v[4] + v[5] * v[5]);
    if (n > JMINVAL) {
      float sh, chh; sincosf(0.5f * n * h, &sh, &chh);
      float inv = sh / n, dq[4] = {chh, v[3] * inv, v[4] * inv, v[5] * inv}, r[4];
      quat_mul(r, q + 3, dq);
      q[3] = r[0]; q[4] = r[1]; q[5] = r[2]; q[6] = r[3];
      quat_normalize(q + 3);
    }
  }
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) s.warm[i] = s.qacc[i];
}

// ------------------------------------------------------------------------------------------- OSC (App. A.2)
JHD void sym_pinv6(const float* X, float rcond, float* out) {  // numpy.linalg.pinv for a symmetric PSD 6x6 (cyclic Jacobi)
  float A[6][6], V[6][6];
  for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) { A[i][j] = X[6 * i + j]; V[i][j] = i == j ? 1.0f : 0.0f; }
  for (int sweep = 0; sweep < 12; sweep++) {
    float off = 0, dg = 0;
    for (int i = 0; i < 6; i++) { dg += A[i][i] * A[i][i]; for (int j = i + 1; j < 6; j++) off += A[i][j] * A[i][j]; }
    if (off <= 1e-14f * dg) break;
    for (int p = 0; p < 6; p++) for (int q = p + 1; q < 6; q++) {
      const float apq = A[p][q];
      if (fabsf(apq) < 1e-30f) continue;
      float th = (A[q][q] - A[p][p]) / (2 * apq);
      float t = (th >= 0 ? 1.0f : -1.0f) / (fabsf(th) + sqrtf(th * th + 1)), c = 1.0f / sqrtf(t * t + 1), sn = t * c;
      // symmetric two-sided rotation: only rows / columns p, q change, the (p, q) entry is annihilated exactly
      // (half the work of rotating columns and rows separately; this branch is 16 % of the aged step's instructions)
      for (int k = 0; k < 6; k++) {
        if (k == p || k == q) continue;
        const float akp = A[k][p], akq = A[k][q];
        const float np_ = c * akp - sn * akq, nq_ = sn * akp + c * akq;
        A[k][p] = np_; A[p][k] = np_; A[k][q] = nq_; A[q][k] = nq_;
      }
      A[p][p] -= t * apq; A[q][q] += t * apq; A[p][q] = 0; A[q][p] = 0;
      for (int k = 0; k < 6; k++) { float vkp = V[k][p], vkq = V[k][q]; V[k][p] = c * vkp - sn * vkq; V[k][q] = sn * vkp + c * vkq; }
    }
  }
  float lmax = 0;
  for (int i = 0; i < 6; i++) lmax = fmaxf(lmax, fabsf(A[i][i]));
  float inv[6];   // 1 / lambda_k of the kept eigenvalues, 0 for the truncated ones (6 divisions instead of 216)
  for (int k = 0; k < 6; k++) inv[k] = fabsf(A[k][k]) > rcond * lmax ? 1.0f / A[k][k] : 0.0f;
  for (int i = 0; i < 6; i++) for (int j = 0; j <= i; j++) {
    float sm = 0;
    for (int k = 0; k < 6; k++) sm += V[i][k] * inv[k] * V[j][k];
    out[6 * i + j] = sm; out[6 * j + i] = sm;
  }
}
// abr_control OSC.generate with kp=50 ko=180 kv=20 vmax=[0.4,1.0472] (env_mujoco_util.py:59-63), reading the
// lagged mjData quantities (Jee, Mc, bias, eepos, eeq) and the fresh joint velocity dq
JHD void osc_generate(const JacoSim& s, const float* dq, float* u) {
  const float kp = 50, ko = 180, kv = 20, vmax0 = 0.4f, vmax1 = 1.0472f;
  float L[21], Minv[6][6], X[21], Mx[6][6];
#pragma unroll
  for (int i = 0; i < 21; i++) L[i] = s.Mc[i];  // leading 6x6 of the packed chain inertia
  chol_packed<6>(L);
#pragma unroll
  for (int c = 0; c < 6; c++) {
    float col[6] = {0, 0, 0, 0, 0, 0}; col[c] = 1;
    chol_solve_packed<6>(L, col);
#pragma unroll
    for (int r = 0; r < 6; r++) Minv[r][c] = col[r];
  }
  float JM[6][6];
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int b = 0; b < 6; b++) { float sm = 0;
#pragma unroll
      for (int a = 0; a < 6; a++) sm += s.Jee[i][a] * Minv[a][b];
      JM[i][b] = sm; }
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j <= i; j++) { float sm = 0;
#pragma unroll
      for (int b = 0; b < 6; b++) sm += JM[i][b] * s.Jee[j][b];
      X[PK(i, j)] = sm; }
  float Lx[21];
#pragma unroll
  for (int i = 0; i < 21; i++) Lx[i] = X[i];
  int ok = chol_packed<6>(Lx);
  float det = 1;
#pragma unroll
  for (int i = 0; i < 6; i++) det *= Lx[PK(i, i)] * Lx[PK(i, i)];
  if (ok && fabsf(det) >= 1e-3f) {
#pragma unroll
    for (int c = 0; c < 6; c++) {
      float col[6] = {0, 0, 0, 0, 0, 0}; col[c] = 1;
      chol_solve_packed<6>(Lx, col);
#pragma unroll
      for (int r = 0; r < 6; r++) Mx[r][c] = col[r];
    }
  } else {
    float Xf[36];
    for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) Xf[6 * i + j] = X[i >= j ? PK(i, j) : PK(j, i)];
    sym_pinv6(Xf, 1e-4f, &Mx[0][0]);
  }
  float ut[6];
#pragma unroll
  for (int i = 0; i < 3; i++) ut[i] = s.eepos[i] - s.target[i];
  float qd[4], qc[4] = {s.eeq[0], -s.eeq[1], -s.eeq[2], -s.eeq[3]}, qr[4];
  euler_rxyz_to_quat(qd, s.target[3], s.target[4], s.target[5]);
  quat_normalize(qd);
  quat_mul(qr, qd, qc);
  float sgq = qr[0] > 0 ? 1.0f : (qr[0] < 0 ? -1.0f : 0.0f);
  ut[3] = -qr[1] * sgq; ut[4] = -qr[2] * sgq; ut[5] = -qr[3] * sgq;
  float sat_xyz = vmax0 / kp * kv, sat_abg = vmax1 / ko * kv, nx = v3norm(ut), na = v3norm(ut + 3), sc[6] = {1, 1, 1, 1, 1, 1};
  if (nx > sat_xyz) { float r = sat_xyz / nx; sc[0] *= r; sc[1] *= r; sc[2] *= r; }
  if (na > sat_abg) { float r = sat_abg / na; sc[3] *= r; sc[4] *= r; sc[5] *= r; }
#pragma unroll
  for (int i = 0; i < 6; i++) ut[i] = kv * sc[i] * ((i < 3 ? kp : ko) / kv) * ut[i];
  float f[6];
#pragma unroll
  for (int i = 0; i < 6; i++) { float sm = 0;
#pragma unroll
    for (int j = 0; j < 6; j++) sm += Mx[i][j] * ut[j];
    f[i] = sm; }
#pragma unroll
  for (int i = 0; i < 6; i++) {
    float sm = 0, tt = 0;
#pragma unroll
    for (int j = 0; j < 6; j++) { sm += s.Mc[i >= j ? PK(i, j) : PK(j, i)] * dq[j]; tt += s.Jee[j][i] * f[j]; }
    u[i] = -kv * sm - tt + s.bias[i];
  }
}

// ------------------------------------------------------------------------------------------- env logic
template <class S>
JHD float uni(const JacoCfg& cfg, const S& s, int env_id, int stream, int k) {
  uint32_t o[4];
  philox4x32(cfg.seed, (uint32_t)(k >> 2), (uint32_t)stream, (uint32_t)env_id, s.rng_counter, o);
  return (float)(o[k & 3] >> 8) * (1.0f / 16777216.0f);
}
template <class S>
JHD float uniab(const JacoCfg& cfg, const S& s, int env_id, int stream, int k, float a, float b) { return a + (b - a) * uni(cfg, s, env_id, stream, k); }

template <class S>
JHD void get_gripper_pose(S& s) {  // util:309, 667-673
  v3copy(s.gripper_pose, s.eepos);
  quat_to_euler_rxyz(s.gripper_pose + 3, s.eeq);
}
JHD int get_touch(const JacoSim& s) {  // util:470-490
  const float th = 0.001f;
  int thumb = 0, index = 0, pinky = 0, inner = 0, outer = 0;
  for (int i = 1; i < 5; i++) if (s.touch[i] > th) thumb = 1;
  for (int i = 5; i < 9; i++) if (s.touch[i] > th) index = 1;
  for (int i = 9; i < 13; i++) if (s.touch[i] > th) pinky = 1;
  for (int i = 0; i < 13; i++) if (s.touch[i] > th) inner = 1;
  for (int i = 13; i < 20; i++) if (s.touch[i] > th) outer = 1;
  if ((thumb && index) || (thumb && pinky)) return 3;
  if (inner) return 1;
  if (outer) return 2;
  return 0;
}
template <class S>
JHD void rulebased_subgoal(const JacoCfg& cfg, S& s, int env_id, float* pos, float* ori) {  // util:273-300
  if (cfg.task == JACO_TASK_PLACING) { v3copy(pos, s.dest_goal); ori[0] = 0; ori[1] = JPI / 2; ori[2] = 0; return; }
  float D[3], n;
  v3sub(D, s.gripper_pose, s.obj_goal); n = v3norm(D);
  for (int i = 0; i < 3; i++) pos[i] = D[i] / n * 0.12f + s.obj_goal[i] + (uni(cfg, s, env_id, 1, i) - 0.5f) / 25;
  if (pos[2] < 0.1898f + 0.1f) pos[2] = 0.1898f + 0.1f;
  if (pos[1] > obj_pos(s)[1] - 0.15f) pos[1] = obj_pos(s)[1] - 0.15f;
  float x = -D[0] / n, y = -D[1] / n, z = -D[2] / n;
  float qp = sqrtf(2.0f) / 2, qv[4] = {qp, qp * x, qp * y, qp * z}, qd[4] = {sqrtf(3.0f) / 2, 0.5f * x, 0.5f * y, 0.5f * z}, rv[3], rd[3], rr[3];
  quat_to_euler_rxyz(rv, qv); quat_to_euler_rxyz(rd, qd);
  v3cross(rr, rv, rd);
  rr[0] -= JPI / 2;
  for (int i = 0; i < 3; i++) ori[i] = rr[i] + (uni(cfg, s, env_id, 1, 3 + i) - 0.5f) / 10;
}
// env_mujoco_util_trans.py:848-862: global_to_local(prop) = (Rx Ry Rz)^T (prop - base_position), applied as written to positions and
// to Euler triples alike; local_to_global(prop) = (Rx Ry Rz) prop + base_position
JHD void global_to_local(const float* prop, float* out) {
  float d[3]; v3sub(d, prop, MDL.base_link_pos);
  get_rotation(out, MDL.base_link_eul[0], MDL.base_link_eul[1], MDL.base_link_eul[2], d, 1);
}
JHD void local_to_global(const float* prop, float* out) {
  float r[3]; get_rotation(r, MDL.base_link_eul[0], MDL.base_link_eul[1], MDL.base_link_eul[2], prop, 0);
  v3add(out, r, MDL.base_link_pos);
}
template <class S>
JHD void get_observation(const JacoCfg& cfg, S& s, int env_id, float* o) {  // util:223-271
  get_gripper_pose(s);
  s.touch_index = get_touch(s);
  // _cis / _trans _get_touch (env_mujoco_util_cis.py:568-578) tests "inner touch" with a second `if`, not `elif`: the grasp class 3 is
  // always overwritten by 1 (thumb / index / pinky pads are inner pads).  Reproduced, not normalised: observation, reward (+0.75 for
  // class 1 instead of +2.5 for class 3) and the success test (touch in {1, 3}) then match what the reference computes.
  if (cfg.variant != 0 && s.touch_index == 3) s.touch_index = 1;
  if (cfg.variant == 2 && cfg.obs_mode == 1) {   // env_mujoco_util_trans.py:295-313
    float sp[3], so[3], t[3];
    const float zero[3] = {0, 0, 0}, dori[3] = {0, JPI / 2, 0};
    o[0] = (float)s.touch_index;
    global_to_local(s.gripper_pose, o + 1);
    global_to_local(s.gripper_pose + 3, t); for (int i = 0; i < 3; i++) o[4 + i] = t[i] / JPI;
    o[7] = (s.grip - 0.8f) / 0.2f;
    global_to_local(obj_pos(s), o + 8);
    global_to_local(zero, o + 11);
    global_to_local(s.dest_goal, o + 14);
    rulebased_subgoal(cfg, s, env_id, sp, so);
    global_to_local(sp, o + 17);
    global_to_local(so, t); for (int i = 0; i < 3; i++) o[20 + i] = t[i] / JPI;
    global_to_local(dori, o + 23);
    return;
  }
  float sp[3], so[3], base[3] = {0, 0, 0};
  if (cfg.variant != 0) v3copy(base, MDL.base_link_pos);
  o[0] = (float)s.touch_index;
  for (int i = 0; i < 3; i++) { o[1 + i] = s.gripper_pose[i] - base[i]; o[4 + i] = s.gripper_pose[3 + i] / JPI; }
  o[7] = (s.grip - 0.8f) / 0.2f;
  for (int i = 0; i < 3; i++) { o[8 + i] = obj_pos(s)[i] - base[i]; o[11 + i] = 0; o[14 + i] = s.dest_goal[i] - ((cfg.variant != 0 && cfg.obs_mode != 0) ? base[i] : 0.0f); }
  if (cfg.obs_mode == 1) { rulebased_subgoal(cfg, s, env_id, sp, so); for (int i = 0; i < 3; i++) { o[17 + i] = sp[i] - base[i]; o[20 + i] = so[i] / JPI; } }
  else if (cfg.obs_mode == 2) { for (int i = 0; i < 3; i++) { o[17 + i] = s.gripper_pose[i]; o[20 + i] = s.gripper_pose[3 + i] / JPI; } }
  else { for (int i = 0; i < 3; i++) { o[17 + i] = s.reach_goal[i] - base[i]; o[20 + i] = s.reach_goal[3 + i] / JPI; } }
  o[23] = 0; o[24] = JPI / 2; o[25] = 0;
}
template <class S>
JHD float reach_ang_diff(const S& s) {  // util:323-333
  float g[4], t[4], ge[3], te[3], d[3];
  euler_rxyz_to_quat(g, s.gripper_pose[3], s.gripper_pose[4], s.gripper_pose[5]);
  euler_rxyz_to_quat(t, s.reach_goal[3], s.reach_goal[4], s.reach_goal[5]);
  quat_to_euler_rxyz(ge, g); quat_to_euler_rxyz(te, t);
  v3sub(d, ge, te);
  float a = v3norm(d);
  if (a > JPI) a = 2 * JPI - a;
  return a;
}
template <class S>
JHD float get_reward(const JacoCfg& cfg, const S& s) {  // util:312-446
  if (cfg.task == JACO_TASK_REACHING) {
    float d[3]; v3sub(d, s.gripper_pose, s.reach_goal);
    float dist = v3norm(d), ang = reach_ang_diff(s);
    float r = 5 * expf(-dist) / 2 + 2 * expf(-ang / (JPI / 6)) / (2 * (dist * 15 + 1));
    float w[3]; v3sub(w, s.eepos, cfg.variant == 0 ? MDL.base_pos : MDL.base_link_pos);   // cis:42,409
    float wb = v3norm(w);
    if (wb < 0.15f) r -= 0.15f - wb;
    if (s.gripper_pose[2] < 0.1f) r -= 0.1f - s.gripper_pose[2];
    return 0.05f * r;
  }
  if (cfg.task == JACO_TASK_PICKING && cfg.variant != 0 && cfg.binary) return 0;   // cis:459,500 / trans:465: `if not self.binary ... else: return 0`
  if (cfg.task == JACO_TASK_GRASPING || cfg.task == JACO_TASK_PICKING) {
    float scale = cfg.task == JACO_TASK_GRASPING ? 0.05f : 0.01f;
    float ev[3], down[3] = {0, 0, -1}, up[3] = {0, 0, 1}, xyz[3], tv[3], dv[3];
    get_rotation(ev, s.gripper_pose[3], s.gripper_pose[4], s.gripper_pose[5], down, 0);
    v3sub(xyz, obj_pos(s), s.gripper_pose);
    float od = v3norm(xyz), x = xyz[0] / od, z = xyz[2] / od;
    // fp32 guard: |x| / sqrt(1 - z^2) can round past 1 when the object sits in the gripper's x-z plane (DESIGN.md "deviations")
    float pitch = -acosf(fminf(fmaxf(z, -1.0f), 1.0f)), yaw = -acosf(fminf(fmaxf(-x / sqrtf(fmaxf(1 - z * z, 1e-30f)), -1.0f), 1.0f));
    get_rotation(tv, 0, pitch, yaw, up, 1);
    tv[2] *= -1;
    v3sub(dv, ev, tv);
    float ad = v3norm(dv);
    float r = 5 * expf(-od / 0.2f) / 2 + 2 * expf(-ad / (JPI / 6)) / 2 / (od * 15 + 1);
    if (s.touch_index == 1) r += 5 * 0.5f * 0.3f;
    else if (s.touch_index == 2) r -= 5 * 0.5f * 0.3f;
    else if (s.touch_index == 3) r += 5 * 0.5f;
    r += 100 * (obj_pos(s)[2] - 0.1898f);
    return r * scale;
  }
  return 0;
}
template <class S>
JHD int terminal_inspection(const JacoCfg& cfg, S& s, float* bonus, float* wb, int* succ) {  // util:492-600
  s.num_episodes += 1;
  float d[3]; *succ = 0; *bonus = 0;
  v3sub(d, s.eepos, cfg.variant == 0 ? MDL.base_pos : MDL.base_link_pos); *wb = v3norm(d);   // util:35,499; cis:42,590
  const float* op = obj_pos(s);
  float dx = s.dest_goal[0] - op[0], dy = s.dest_goal[1] - op[1], dest_diff = sqrtf(dx * dx + dy * dy);
  float q2 = s.qpos[2];
  if (JPI - 0.1f < q2 && q2 < JPI + 0.1f) { *bonus = -1; return 1; }
  const float oz = 0.1898f;
  switch (cfg.task) {
  case JACO_TASK_REACHING:
    v3sub(d, s.gripper_pose, s.reach_goal);
    if (v3norm(d) < 0.025f && reach_ang_diff(s) < JPI / 6) { *bonus = 200 - s.num_episodes * 0.1f; *succ = 1; return 1; }
    return 0;
  case JACO_TASK_GRASPING:
    v3sub(d, s.gripper_pose, op);
    if (v3norm(d) > 0.2f) { *bonus = -20; return 1; }
    if (op[2] > oz + 0.07f && (s.touch_index == 1 || s.touch_index == 3)) { *bonus = 200 - s.num_episodes * 0.1f; *succ = 1; return 1; }
    if (op[2] < 0.1f) { *bonus = -20; return 1; }
    return 0;
  case JACO_TASK_PICKING: {
    const bool bin = cfg.variant != 0 && cfg.binary;   // cis:637-660 / trans:643-: sparse bonuses 10 / -1 instead of 200 - 0.1 n / -20
    if (op[2] > oz + 0.07f && (s.touch_index == 1 || s.touch_index == 3)) { *bonus = bin ? 10.0f : 200 - s.num_episodes * 0.1f; *succ = 1; return 1; }
    if (op[2] < 0.1f) { *bonus = bin ? -1.0f : -20.0f; return 1; }
    return 0; }
  case JACO_TASK_PLACING: {
    float th = cfg.variant == 0 ? 0.02f : 0.04f;
    if (op[2] < 0.1f) { *bonus = -20; return 1; }
    if (dest_diff < th && s.touch_index == 0 && op[2] < 0.35f) { *bonus = 200 - s.num_episodes * 0.1f; *succ = 1; return 1; }
    if (dest_diff > th && s.touch_index == 0 && op[2] < 0.20f) { *bonus = -20; return 1; }
    return 0; }
  case JACO_TASK_CARRYING: return 1;   // util:549-550: (True, 0, wb)
  case JACO_TASK_RELEASING: {          // util:551-566; obj_velocity = |qvel[9:12]| (mujoco.py:212-215)
    float ov = sqrtf(s.qvel[9] * s.qvel[9] + s.qvel[10] * s.qvel[10] + s.qvel[11] * s.qvel[11]);
    if (op[2] < 0.1f) { *bonus = -20; return 1; }
    if (dest_diff < 0.04f && s.touch_index == 0 && op[2] < 0.35f && ov < 0.01f) { *bonus = 200 - s.num_episodes * 0.1f; *succ = 1; return 1; }
    if (dest_diff > 0.04f && s.touch_index == 0 && op[2] < 0.20f) { *bonus = -20; return 1; }
    return 0; }
  case JACO_TASK_PICKANDPLACE:
    if (op[2] > oz + 0.07f && (s.touch_index == 1 || s.touch_index == 3) && !s.picked) { s.picked = 1; *bonus = 20; return 0; }
    if (dest_diff < 0.04f && s.touch_index == 0 && op[2] < 0.35f) { *bonus = 180; *succ = 1; return 1; }
    if (op[2] < 0.1f) { *bonus = -20; return 1; }
    return 0;
  }
  return 0;
}
// gripper command of substep i (numpy.linspace(prev, new, 50), util:631)
template <class S>
JHD float ramp_value(const S& s, int i) {
  if (!s.action_in) return s.grip;
  if (i >= 49) return s.ramp1;
  return s.ramp0 + i * ((s.ramp1 - s.ramp0) / 49);
}
// one physics substep driven by the controller output u already computed from the lagged data (util:73-83)
JHD void substep(JacoSim& s, const float* u, int i, bool want_touch) {
#pragma unroll
  for (int k = 0; k < 6; k++) s.ctrl[k] = u[k];
  s.ctrl[6] = s.ctrl[7] = s.ctrl[8] = ramp_value(s, i);
  mj_forward(s, want_touch);
  mj_euler(s);
}
template <class S>
JHD void take_action(S& s, const float* a, int variant = 0) {  // util:602-646
  s.action_in = 1;
  get_gripper_pose(s);
  // base: a[:3]/25, a[3:6]/5 (util:621); _cis / _trans, single arm: a[:3]/100, a[3:6]/20 (env_mujoco_util_cis.py:776)
  const float sp = variant == 0 ? 25.0f : 100.0f, sa = variant == 0 ? 5.0f : 20.0f;
  for (int i = 0; i < 3; i++) { s.target[i] = s.gripper_pose[i] + a[i] / sp; s.target[3 + i] = s.gripper_pose[3 + i] + a[3 + i] / sa; }
  if (fabsf(s.target[5]) > JPI) s.target[5] += -(s.target[5] > 0 ? 1.0f : -1.0f) * 2 * JPI;
  float prev = s.grip, g = fminf(fmaxf(prev + a[6] / 10, 0.6f), 1.0f);
  s.grip = g; s.ramp0 = prev; s.ramp1 = g;
}
JHD void set_obj(JacoSim& s, const float* xyz, const float* quat) {  // mujoco.py:217-227
  for (int i = 0; i < 3; i++) s.qpos[9 + i] = xyz[i];
  for (int i = 0; i < 4; i++) s.qpos[12 + i] = quat[i];
  for (int i = 9; i < JACO_NV; i++) s.qvel[i] = 0;
}
JHD float round_half(float v) {  // value after a round trip through np.float16 (normal range)
  if (v == 0) return v;
  int ex; float mnt = frexpf(v, &ex);
  mnt = rintf(mnt * 2048.0f) / 2048.0f;
  return ldexpf(mnt, ex);
}
// sim.reset() + _create_init_angle + __sample_goal + set_dest_xyz (util:92-103,176-221; mujoco.py:229-237): everything of
// reset() that precedes the first mj_forward.  Shared by the general reset (env_reset) and the in-place reset of the fast path
// (jaco_fast.cuh fast_reset); S = JacoSim or FastSim.
template <class S>
JHD void reset_sample(const JacoCfg& cfg, S& s, int env_id) {
  const JacoModel& m = MDL;
  s.current_steps = 0;
  for (int k = 0; k < 9; k++) s.qpos[k] = m.ch_ref[k];
  for (int f = 0; f < 2; f++) for (int i = 0; i < 7; i++) s.qpos[9 + 7 * f + i] = m.free_qpos0[f][i];
  for (int i = 0; i < JACO_NV; i++) { s.qvel[i] = 0; s.warm[i] = 0; }
  s.action_in = 0; s.picked = 0; s.grip = 0.6f; s.ramp0 = s.ramp1 = 0.6f; s.touch_index = 0; s.num_episodes = 0;
  float ia[6];
  if (cfg.task == JACO_TASK_REACHING || cfg.task == JACO_TASK_PICKING || cfg.task == JACO_TASK_PICKANDPLACE) {
    ia[0] = uniab(cfg, s, env_id, 0, 0, 0.7f, 2.5f); ia[1] = uniab(cfg, s, env_id, 0, 1, 3.8f, 4.0f); ia[2] = uniab(cfg, s, env_id, 0, 2, 1.0f, 1.7f);
    ia[3] = uniab(cfg, s, env_id, 0, 3, 1.8f, 2.5f); ia[4] = uniab(cfg, s, env_id, 0, 4, 1.0f, 2.5f); ia[5] = uniab(cfg, s, env_id, 0, 5, 0.8f, 2.3f);
  } else if (cfg.task == JACO_TASK_RELEASING) {   // util:186-189: nine values, the fingers start at 0.6 (mujoco.py:342-346)
    ia[0] = uniab(cfg, s, env_id, 0, 0, 1.9f, 2.0f); ia[1] = uniab(cfg, s, env_id, 0, 1, 3.3f, 3.6f); ia[2] = uniab(cfg, s, env_id, 0, 2, 0.5f, 0.8f);
    ia[3] = uniab(cfg, s, env_id, 0, 3, 1.8f, 2.5f); ia[4] = uniab(cfg, s, env_id, 0, 4, 1.3f, 2.0f); ia[5] = uniab(cfg, s, env_id, 0, 5, -0.4f, -0.9f);
    s.qpos[6] = s.qpos[7] = s.qpos[8] = 0.6f;
  } else {
    float c0 = uniab(cfg, s, env_id, 0, 0, 3 * JPI / 8, JPI / 2), c1 = uniab(cfg, s, env_id, 0, 1, JPI / 2, 5 * JPI / 8);
    ia[0] = uni(cfg, s, env_id, 0, 6) < 0.5f ? c0 : c1; ia[1] = 3.85f; ia[2] = uniab(cfg, s, env_id, 0, 2, 1.0f, 1.1f);
    ia[3] = uniab(cfg, s, env_id, 0, 3, 2.0f, 2.1f); ia[4] = uniab(cfg, s, env_id, 0, 4, 0.8f, 2.3f); ia[5] = uniab(cfg, s, env_id, 0, 5, -1.2f, -1.1f);
  }
  for (int i = 0; i < 6; i++) s.qpos[i] = ia[i];
  float rg[3], xyz[3];
  rg[0] = uniab(cfg, s, env_id, 0, 7, 0.3f, 0.42f) * (uni(cfg, s, env_id, 0, 8) < 0.5f ? -1.0f : 1.0f);
  rg[1] = uniab(cfg, s, env_id, 0, 9, 0.3f, 0.42f) * (uni(cfg, s, env_id, 0, 10) < 0.5f ? -1.0f : 1.0f);
  rg[2] = uniab(cfg, s, env_id, 0, 11, 0.3f, 0.5f);
  v3sub(xyz, rg, cfg.variant == 0 ? m.base_pos : m.base_link_pos);
  { float n = v3norm(xyz); xyz[0] /= n; xyz[1] /= n; xyz[2] /= n; }
  float sx = xyz[0] > 0 ? 1.0f : (xyz[0] < 0 ? -1.0f : 0.0f);
  float alpha = -asinf(xyz[1] / sqrtf(xyz[1] * xyz[1] + xyz[2] * xyz[2])) * sx;
  float beta = acosf(xyz[0] / v3norm(xyz)) * sx, gamma = uniab(cfg, s, env_id, 0, 12, -0.1f, 0.1f);
  v3copy(s.reach_goal, rg);
  s.reach_goal[3] = round_half(alpha); s.reach_goal[4] = round_half(beta); s.reach_goal[5] = round_half(gamma);
  s.obj_goal[0] = uniab(cfg, s, env_id, 0, 13, -0.1f, 0.1f); s.obj_goal[1] = 0.65f + uniab(cfg, s, env_id, 0, 14, -0.08f, 0.02f); s.obj_goal[2] = 0.1898f;
  s.dest_goal[0] = 0.4f + uniab(cfg, s, env_id, 0, 15, -0.05f, 0.05f); s.dest_goal[1] = 0.3f + uniab(cfg, s, env_id, 0, 16, -0.05f, 0.05f); s.dest_goal[2] = 0.3468f;
  if (cfg.variant == 2) { float g[3]; local_to_global(s.obj_goal, g); v3copy(s.obj_goal, g); local_to_global(s.dest_goal, g); v3copy(s.dest_goal, g); }   // trans:265-270
  else if (cfg.variant != 0) for (int i = 0; i < 3; i++) { s.obj_goal[i] += m.base_link_pos[i]; s.dest_goal[i] += m.base_link_pos[i]; }
  if (cfg.variant == 0) { s.qpos[16] = s.dest_goal[0]; s.qpos[17] = s.dest_goal[1]; for (int i = 16; i < JACO_NV; i++) s.qvel[i] = 0; }
}
// env_mujoco.py:99-102 -> util:92-174.  Leaves mjData (the JacoSim working set) evaluated at the reset state.
JHD void env_reset(const JacoCfg& cfg, JacoSim& s, int env_id, float* obs) {
  const JacoModel& m = MDL;
  reset_sample(cfg, s, env_id);
  for (int i = 0; i < JACO_NU; i++) s.ctrl[i] = 0;
  const bool held = cfg.task == JACO_TASK_PLACING || cfg.task == JACO_TASK_RELEASING || cfg.task == JACO_TASK_CARRYING;   // util:106
  if (held) {  // util:106-117: hold the object in the hand for 150 settle substeps
    mj_forward(s, false);
    float p[3], t[3], eul[3], q[4], dp[3], off[3] = {-0.04f, 0, 0}, u[6];
    m3mulv(t, s.bR[6], m.ee_obj_off); v3add(p, s.bx[6], t);
    quat_to_euler_rxyz(eul, s.eeq);
    euler_rxyz_to_quat(q, eul[0], eul[1], eul[2]);
    get_rotation(dp, eul[0], eul[1], eul[2], off, 0);
    v3add(p, p, dp);
    set_obj(s, p, q);
    mj_forward(s, false);
    get_gripper_pose(s);
    for (int i = 0; i < 6; i++) s.target[i] = s.gripper_pose[i];
    for (int i = 0; i < 150; i++) {
      osc_generate(s, s.qvel, u);
      substep(s, u, i, false);
      set_obj(s, p, q);
      mj_forward(s, false);
    }
    mj_forward(s, true);
  } else {
    if (cfg.task != JACO_TASK_REACHING) { float zq[4] = {0, 0, 0, 0}; set_obj(s, s.obj_goal, zq); }
    mj_forward(s, true);
  }
  if (cfg.task == JACO_TASK_REACHING) { for (int i = 0; i < 6; i++) s.target[i] = s.reach_goal[i]; }
  else if (cfg.task == JACO_TASK_GRASPING || cfg.task == JACO_TASK_CARRYING) {
    // util:123-167: two approach servo loops.  Unbounded `while True` in the reference; bounded here (JACO_GRASP_RESET_MAX substeps
    // each: 4 s of simulated time, the approach takes ~0.5-1.5 s at the controller's 0.4 m/s limit).
    get_gripper_pose(s);
    float d[3], u[6]; v3sub(d, s.obj_goal, s.gripper_pose);
    const float sx = d[0] > 0 ? 1.0f : (d[0] < 0 ? -1.0f : 0.0f);
    float ori[3] = {round_half(-asinf(d[1] / sqrtf(d[1] * d[1] + d[2] * d[2])) * sx), round_half(acosf(d[0] / v3norm(d)) * sx),
                    round_half(uniab(cfg, s, env_id, 0, 17, -0.1f, 0.1f))};
    for (int i = 0; i < 3; i++) { s.target[i] = s.obj_goal[i]; s.target[3 + i] = ori[i]; }
    float qt[4]; euler_rxyz_to_quat(qt, s.reach_goal[3], s.reach_goal[4], s.reach_goal[5]); quat_normalize(qt);
    for (int it = 0; it < JACO_GRASP_RESET_MAX; it++) {
      for (int i = 9; i < JACO_NV; i++) s.qvel[i] = 0;   // stop_obj (mujoco.py:239-246), then sim.forward()
      mj_forward(s, false);
      osc_generate(s, s.qvel, u);
      substep(s, u, 0, true);
      get_gripper_pose(s);
      v3sub(d, s.gripper_pose, s.obj_goal);
      float qg[4]; euler_rxyz_to_quat(qg, s.gripper_pose[3], s.gripper_pose[4], s.gripper_pose[5]); quat_normalize(qg);
      float ad = 0; for (int i = 0; i < 4; i++) ad += (qg[i] - qt[i]) * (qg[i] - qt[i]);
      if (v3norm(d) < 0.2f) { for (int i = 0; i < 3; i++) s.target[i] = s.gripper_pose[i]; break; }
      if (sqrtf(ad) < JPI / 6) break;
    }
    for (int it = 0; it < JACO_GRASP_RESET_MAX; it++) {
      osc_generate(s, s.qvel, u);
      substep(s, u, 0, true);
      get_gripper_pose(s);
      for (int i = 0; i < 3; i++) s.target[i] = s.obj_goal[i];
      v3sub(d, s.gripper_pose, s.obj_goal);
      if (v3norm(d) < 0.15f) break;
    }
  }
  else if (!held) { get_gripper_pose(s); for (int i = 0; i < 6; i++) s.target[i] = s.gripper_pose[i]; }
  get_observation(cfg, s, env_id, obs);
  s.rng_counter += 1;
}
// env_mujoco.py:116-139.  On entry the working set holds the chain quantities at the lagged state
// (see load path in jaco_kernels.cu); on exit it holds mjData of the last substep's forward.
JHD void env_step(const JacoCfg& cfg, JacoSim& s, int env_id, const float* action, float* obs, float* reward, int* done, int* succ, float* wb, int* cause = nullptr) {
  float a[7], u[6];
  for (int i = 0; i < 7; i++) a[i] = fminf(fmaxf(action[i], -1.0f), 1.0f);
  if (cfg.task == JACO_TASK_REACHING) a[6] = 0;
  take_action(s, a, cfg.variant);
  // 6-D action (util:625-626): grip 0.6 for the observation; the finger servos get gripper_angle_array = zeros (util:37,76-79)
  if (cfg.task == JACO_TASK_REACHING) { s.grip = 0.6f; s.ramp0 = s.ramp1 = 0.0f; }
  for (int i = 0; i < 50; i++) {
    osc_generate(s, s.qvel, u);
    substep(s, u, i, i == 49);
  }
  get_observation(cfg, s, env_id, obs);
  float r = get_reward(cfg, s), bonus, w; int sc, d;
  s.current_steps += 1;
  if (s.current_steps < cfg.task_max_steps) d = terminal_inspection(cfg, s, &bonus, &w, &sc);
  else { d = 1; bonus = -10; w = 0; sc = 0; }
  // cause: 0 running, 1 success, 2 time-out, 3 singular exit, 4 drop / -20 exit, 5 other
  if (cause) *cause = !d ? 0 : (sc ? 1 : (s.current_steps >= cfg.task_max_steps ? 2 : (JPI - 0.1f < s.qpos[2] && s.qpos[2] < JPI + 0.1f ? 3 : (bonus < 0 ? 4 : 5))));
  *reward = r + bonus; *done = d; *succ = sc; *wb = w;
  s.rng_counter += 1;
}

// ------------------------------------------------------------------------------------------- state I/O
// SoA state buffer: row r of env i lives at state[r * N + i]; rows = qpos(23) | qvel(21) | aux(72)
#define ST(r) state[(size_t)(r) * N + i]
JHD void load_env(const float* state, int N, int i, JacoSim& s) {
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
#pragma unroll
  for (int k = 0; k < JACO_NU; k++) s.ctrl[k] = 0;
}
JHD void store_env(float* state, int N, int i, const JacoSim& s) {
#pragma unroll
  for (int k = 0; k < JACO_NQ; k++) ST(k) = s.qpos[k];
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) ST(JACO_NQ + k) = s.qvel[k];
  const int A = JACO_NQ + JACO_NV;
#pragma unroll
  for (int k = 0; k < 6; k++) ST(A + AUX_TARGET + k) = s.target[k];
  ST(A + AUX_GRIP) = s.grip; ST(A + AUX_ACTION_IN) = (float)s.action_in; ST(A + AUX_PICKED) = (float)s.picked;
  ST(A + AUX_CUR_STEPS) = (float)s.current_steps; ST(A + AUX_NUM_EPISODES) = (float)s.num_episodes;
#pragma unroll
  for (int k = 0; k < 3; k++) { ST(A + AUX_OBJ_GOAL + k) = s.obj_goal[k]; ST(A + AUX_DEST_GOAL + k) = s.dest_goal[k]; }
#pragma unroll
  for (int k = 0; k < 6; k++) ST(A + AUX_REACH_GOAL + k) = s.reach_goal[k];
  ST(A + AUX_RNG) = (float)(s.rng_counter & 0xFFFFFFu); ST(A + AUX_RNG_HI) = (float)(s.rng_counter >> 24); ST(A + AUX_TOUCH) = (float)s.touch_index;
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) ST(A + AUX_WARM + k) = s.warm[k];
#pragma unroll
  for (int k = 0; k < 9; k++) { ST(A + AUX_LAGQ + k) = s.lagq[k]; ST(A + AUX_LAGV + k) = s.lagv[k]; }
  ST(A + AUX_ACCUM_REW) = s.accum_rew;
}
#undef ST
// rebuild the controller-visible mjData (chain kinematics/dynamics) at the lagged state
JHD void refresh_lagged(JacoSim& s) {
  float q[9], v[9];
#pragma unroll
  for (int k = 0; k < 9; k++) { q[k] = s.qpos[k]; v[k] = s.qvel[k]; s.qpos[k] = s.lagq[k]; s.qvel[k] = s.lagv[k]; }
  forward_chain_smooth(s);
#pragma unroll
  for (int k = 0; k < 9; k++) { s.qpos[k] = q[k]; s.qvel[k] = v[k]; }
}

// episode / workload counters (jaco_stats): 0-8 episode statistics, 9-15 workload and cap counters of the step kernels
enum { STAT_EPISODES = 0, STAT_SUCCESS = 1, STAT_RETURN = 2, STAT_TIMEOUT = 3, STAT_DROP = 4, STAT_SINGULAR = 5, STAT_NAN = 6, STAT_STEPS = 7, STAT_LENGTH = 8,
       STAT_CONCAP = 9,      // env-substeps whose contact list hit JACO_MAXCON / the row pool (contacts were dropped)
       STAT_ITERCAP = 10,    // Newton solves that stopped on the iteration cap instead of the convergence test
       STAT_STAGEB = 11,     // env-substeps served by the general (cooperative) contact path
       STAT_PINV = 12,       // env-substeps whose OSC took the numpy.linalg.pinv branch
       STAT_FINGEROBJ = 13,  // env-substeps with >= 1 contact between a hand-cluster geom and the object
       STAT_FALLBACK = 14,   // env steps re-run by the general thread-per-env path (NaN guard / bad pivot)
       STAT_GENRESET = 15 }; // resets that went through the general reset kernel (hand near the scene at the reset pose)

// per-env entry points shared by the CUDA kernels (jaco_kernels.cu) and the CPU logic harness (tests/)
JHD void entry_reset(const JacoCfg& cfg, float* state, int N, int i, int env_id, float* obs_row) {
  JacoSim s;
  load_env(state, N, i, s);
  float obs[JACO_NOBS];
  env_reset(cfg, s, env_id, obs);
  s.accum_rew = 0;
  store_env(state, N, i, s);
  if (obs_row) for (int k = 0; k < JACO_NOBS; k++) obs_row[k] = obs[k];
}
// local_stats: 16 floats accumulated by the caller (block-reduced on the GPU)
JHD void entry_step(const JacoCfg& cfg, float* state, int N, int i, int env_id, const float* action_row, float* obs_row,
                    float* reward, unsigned char* done, unsigned char* succ, float* wb, float* term_obs_row, int auto_reset, float* local_stats) {
  JacoSim s;
#ifdef JACO_POISON   /* CPU harness debugging: fill the working set with a byte pattern to expose reads of unset members */
  memset((void*)&s, JACO_POISON, sizeof s);
#endif
  load_env(state, N, i, s);
  refresh_lagged(s);
  float obs[JACO_NOBS], r, w; int d, sc, cause = 0;
  env_step(cfg, s, env_id, action_row, obs, &r, &d, &sc, &w, &cause);
  bool bad = !(r == r);
  for (int k = 0; k < JACO_NQ; k++) bad = bad || !(fabsf(s.qpos[k]) < 1e6f);
  for (int k = 0; k < JACO_NV; k++) bad = bad || !(fabsf(s.qvel[k]) < 1e6f);
  if (bad) { d = 1; sc = 0; r = 0; local_stats[STAT_NAN] += 1; }  // NaN/Inf guard: forced termination, counted, never silent
  s.accum_rew += r;
  local_stats[STAT_STEPS] += 1;
  if (d) {
    local_stats[STAT_EPISODES] += 1; local_stats[STAT_SUCCESS] += sc; local_stats[STAT_RETURN] += s.accum_rew; local_stats[STAT_LENGTH] += s.current_steps;
    if (!bad && cause == 2) local_stats[STAT_TIMEOUT] += 1;
    else if (!bad && cause == 3) local_stats[STAT_SINGULAR] += 1;
    else if (!bad && cause == 4) local_stats[STAT_DROP] += 1;
  }
  if (term_obs_row) for (int k = 0; k < JACO_NOBS; k++) term_obs_row[k] = obs[k];
  if (d && (auto_reset || bad)) { env_reset(cfg, s, env_id, obs); s.accum_rew = 0; }
  store_env(state, N, i, s);
  for (int k = 0; k < JACO_NOBS; k++) obs_row[k] = obs[k];
  *reward = r; *done = (unsigned char)d; *succ = (unsigned char)sc; *wb = w;
}
// one raw mj_step with explicit ctrl (mujoco.py:258-278)
JHD void entry_substep(float* state, int N, int i, const float* ctrl_row, int* ncon, float* sensordata_row, float* qacc_row) {
  JacoSim s;
  load_env(state, N, i, s);
  for (int k = 0; k < JACO_NU; k++) s.ctrl[k] = ctrl_row[k];
  mj_forward(s, true);
  if (ncon) *ncon = s.ncon;
  if (sensordata_row) for (int k = 0; k < 20; k++) sensordata_row[k] = s.touch[k];
  if (qacc_row) for (int k = 0; k < JACO_NV; k++) qacc_row[k] = s.qacc[k];
  mj_euler(s);
  store_env(state, N, i, s);
}
