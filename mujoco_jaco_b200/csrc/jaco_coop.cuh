// jaco_coop.cuh -- warp-cooperative general contact handling for ONE environment.
//
// The fast path (jaco_fast.cuh) advances one environment per thread.  When an environment needs what the fast
// path cannot do in registers -- narrow phase over all 60 geom pairs, joint-limit rows, contacts that couple the
// arm, the object and the destination box -- its thread publishes the mj_forward quantities of that substep
// into a per-warp shared-memory workspace and all 32 lanes of the warp cooperate on:
//   collision (lanes over geom pairs) -> constraint rows (lanes over contacts / Jacobian rows) ->
//   Newton on the primal problem per connected component of {chain, object, dest} (lanes over rows, Hessian
//   entries, Cholesky rows) -> touch sensors (lanes over sites).
// Semantics are those of jaco_sim.cuh (mj_collision, mj_makeConstraint, mj_solNewton with pyramidal cones:
// SURVEY.md App. A.3); only the schedule differs.  The code is written as bulk-synchronous phases over `lane`
// with all inter-lane traffic through the workspace, so the CPU harness runs the identical source by looping
// lane = 0..31 inside each phase.
#pragma once
#include "jaco_sim.cuh"

#if defined(JACO_PROF) && defined(__CUDACC__)
static __device__ unsigned long long g_jprof[16];
#endif
#if defined(JACO_PROF) && defined(__CUDA_ARCH__)
#define JPROF_T(v) long long v = clock64()
#define JPROF_ADD(k_, v_) do { if ((threadIdx.x & 31) == 0) atomicAdd(&g_jprof[k_], (unsigned long long)(v_)); } while (0)
#else
#define JPROF_T(v)
#define JPROF_ADD(k, x)
#endif
#ifdef __CUDA_ARCH__
#define COOP_LANES for (int lane = (int)(threadIdx.x & 31), once_ = 1; once_; once_ = 0)
#define COOP_SYNC() __syncwarp()
#define COOP_LANE0 if ((threadIdx.x & 31) == 0)
#else
#define COOP_LANES for (int lane = 0; lane < 32; lane++)
#define COOP_SYNC() ((void)0)
#define COOP_LANE0
#endif
#ifdef __CUDA_ARCH__
__device__ __forceinline__ float coop_wsum(float v) {   // butterfly sum: every lane gets the same total
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
#endif
#define COOP_BEGIN COOP_SYNC(); COOP_LANES
#define COOP_END COOP_SYNC();

#define COOP_SIMPLE_OBJ 1   // the owner's fast scene collision of the object is valid (it can solve the object alone)
#define COOP_FORCE_B 2      // active joint limit or non-simple scene contacts: stage B regardless of the arm proximity
#ifndef COOP_NROW
#define COOP_NROW 128   // pool of direction-Jacobian rows (3 per condim-3 contact, 6 per condim-6 contact)
#endif

#if defined(JACO_COLLISION_STATS) && !defined(__CUDACC__)
static long long g_coop_pairA[64], g_coop_pairC[64];   // CPU harness only: per-pair stage-A survivors / pairs with contacts
#endif
struct CoopSink {  // contacts of one geom pair (<= 8: box-box clipping cap)
  static const int MAXC = 8;
  int ncon, cpair[8];
  float cpos[8][3], cn[8][3], cdist[8];
};

// What the owner thread publishes: the mj_forward quantities of this substep the cooperative path consumes.
struct CoopPub {
  float bx[JACO_NBODY][3], bR[JACO_NBODY][9], ax[9][3], bw[JACO_NBODY][3], bv[JACO_NBODY][3];
  float qpos9[9], qvel[JACO_NV], Mc[45], Mf[2][21], asm_[JACO_NV], warm[JACO_NV];
  int want_touch, flags;      // flags: COOP_SIMPLE_OBJ | COOP_FORCE_B
  unsigned cand[2];           // stage A: candidate bit per geom pair (pairs >= 4; scene pairs are always tested in stage B)
  // ---- arm contacts the owner thread found with its own narrow phase (fast_arm_collide), in pair order
  int pre_ncon, pre_pair[24];
  float pre_pos[24][3], pre_n[24][3], pre_dist[24];
};
struct CoopWS : CoopPub {
  static JHD const JacoModel& model() { return MDLG; }
  // ---- contacts and rows
  int ncon, nlim, arm_rows;
  int cpair[JACO_MAXCON], cmask[JACO_MAXCON], ccond[JACO_MAXCON], croff[JACO_MAXCON];
  float cpos[JACO_MAXCON][3], cn[JACO_MAXCON][3], cdist[JACO_MAXCON];
  float cD[JACO_MAXCON], cA[JACO_MAXCON][6], ce[JACO_MAXCON][6], cp[JACO_MAXCON][6], cmu[JACO_MAXCON][6];
  float csc[JACO_MAXCON][6], cW[JACO_MAXCON][11], cf0[JACO_MAXCON];
  int ldof[5];
  float lsign[5], lD[5], laref[5], le[5], lp[5];
  union {
    float Jd[COOP_NROW][JACO_NV];   // direction Jacobians (n, t1, t2 linear; n, t1, t2 angular) of each contact
    CoopSink sink[32];              // per-lane narrow-phase output (collision phase only)
  };
  // ---- solver
  float H[231], a[JACO_NV], g[JACO_NV], p[JACO_NV], Mp[JACO_NV], jf[JACO_NV], red[32][4];
  // ---- results
  float qacc[JACO_NV], fcon[JACO_NV], touch[20];
  int iters;
  int fingerobj;   // >= 1 contact between a hand-cluster geom (bodies 6..9) and the object (body 10): workload counter
  int solved;   // stage-B result (block mask) for the owner thread, which may sit in another warp (block task queue)
};

JHD int coop_ndir(int condim) { return condim <= 3 ? (condim == 1 ? 1 : 3) : 6; }
JHD int coop_block(int dof) { return dof < 9 ? 1 : (dof < 15 ? 2 : 4); }

// ------------------------------------------------------------------------------------------- stage A: proximity
// Tight, conservative broad phase of every arm-geom pair: bounding sphere of the arm geom against the exact scene
// box / plane.  Needs only the poses of bodies 0, 3..11.  Returns (uniformly) whether any pair survives.
JHD int coop_stageA(CoopWS& w) {
  const JacoModel& m = MDLG;
  COOP_BEGIN {
    for (int round = 0; round < 2; round++) {
      const int p = round * 32 + lane;
      int hit = 0;
      if (p >= 4 && p < JM_NPAIR) {
        const JacoPair& P = m.pairs[p];
        const bool arm1 = m.geoms[P.g1].body >= 1 && m.geoms[P.g1].body <= 9;
        const JacoGeom& A = arm1 ? m.geoms[P.g1] : m.geoms[P.g2];
        const JacoGeom& S = arm1 ? m.geoms[P.g2] : m.geoms[P.g1];
        float c[3], t[3];
        m3mulv(t, w.bR[A.body], A.pos); v3add(c, w.bx[A.body], t);
        const float rr = A.rbound + P.margin + 1e-5f;
        if (S.type == JACO_GEOM_PLANE) {
          float sp[3], d[3], R[9];
          m3mulv(t, w.bR[S.body], S.pos); v3add(sp, w.bx[S.body], t); m3mul(R, w.bR[S.body], S.R);
          v3sub(d, c, sp);
          hit = !(d[0] * R[2] + d[1] * R[5] + d[2] * R[8] > rr);
        } else if (S.type == JACO_GEOM_BOX) {
          float sp[3], R[9];
          m3mulv(t, w.bR[S.body], S.pos); v3add(sp, w.bx[S.body], t); m3mul(R, w.bR[S.body], S.R);
          hit = !(point_box_dist2(c, sp, R, S.size) > rr * rr);
        } else hit = 1;
      }
      w.red[lane][round] = hit ? 1.0f : 0.0f;
    }
  } COOP_END
  unsigned c0 = 0, c1 = 0;
  for (int l = 0; l < 32; l++) { if (w.red[l][0] != 0.0f) c0 |= 1u << l; if (w.red[l][1] != 0.0f) c1 |= 1u << l; }
  COOP_SYNC();
#if defined(JACO_COLLISION_STATS) && !defined(__CUDACC__)
  for (int l = 0; l < 32; l++) { if ((c0 >> l) & 1) g_coop_pairA[l]++; if ((c1 >> l) & 1) g_coop_pairA[32 + l]++; }
#endif
  COOP_LANE0 { w.cand[0] = c0 | 0xFu; w.cand[1] = c1; }   // scene pairs 0..3 always go through the narrow phase
  COOP_SYNC();
  return (c0 | c1) != 0 || (w.flags & COOP_FORCE_B) != 0;
}

// ------------------------------------------------------------------------------------------- collision
JHD JNOINLINE void coop_collision(CoopWS& w) {
  const JacoModel& m = MDLG;
  float hs[3];
  { float t[3]; m3mulv(t, w.bR[6], m.hand_sphere); v3add(hs, w.bx[6], t); }
  int base = 0;
  for (int round = 0; round < (JM_NPAIR + 31) / 32; round++) {
    if (w.cand[round] == 0u) continue;   // uniform: nothing to test in this round
    COOP_BEGIN {
      int p = round * 32 + lane;
      CoopSink& k = w.sink[lane];
      k.ncon = 0;
      if (p < JM_NPAIR && ((w.cand[round] >> lane) & 1u)) collide_pair(w, p, hs, k);
#if defined(JACO_COLLISION_STATS) && !defined(__CUDACC__)
      if (k.ncon > 0) g_coop_pairC[p]++;
#endif
    } COOP_END
    COOP_BEGIN {
      int off = base;
      for (int l = 0; l < lane; l++) off += w.sink[l].ncon;
      const CoopSink& k = w.sink[lane];
      for (int j = 0; j < k.ncon; j++) {
        int c = off + j;
        if (c >= JACO_MAXCON) break;
        v3copy(w.cpos[c], k.cpos[j]); v3copy(w.cn[c], k.cn[j]); w.cdist[c] = k.cdist[j]; w.cpair[c] = k.cpair[j];
      }
    } COOP_END
    int tot = 0;
    for (int l = 0; l < 32; l++) tot += w.sink[l].ncon;
    base = base + tot > JACO_MAXCON ? JACO_MAXCON : base + tot;
  }
  COOP_BEGIN {  // append the owner's pre-collided arm contacts (pair indices >= 4, already in pair order)
    const int j = lane;
    if (j < w.pre_ncon && base + j < JACO_MAXCON) {
      const int c = base + j;
      v3copy(w.cpos[c], w.pre_pos[j]); v3copy(w.cn[c], w.pre_n[j]); w.cdist[c] = w.pre_dist[j]; w.cpair[c] = w.pre_pair[j];
    }
  } COOP_END
  base = base + w.pre_ncon > JACO_MAXCON ? JACO_MAXCON : base + w.pre_ncon;
  COOP_SYNC();
  COOP_LANE0 { w.ncon = base; }
  COOP_SYNC();
}

// ------------------------------------------------------------------------------------------- constraint rows
JHD JNOINLINE void coop_constraints(CoopWS& w) {
  const JacoModel& m = MDLG;
  COOP_SYNC();
  COOP_LANE0 {  // joint limits (mj_instantiateLimit) + row-pool offsets
    const float def_solref[2] = {0.02f, 1.0f}, def_solimp[5] = {0.9f, 0.95f, 0.001f, 0.5f, 2.0f};
    int nl = 0;
    for (int k = 0; k < 9; k++) {
      if (!m.ch_limited[k]) continue;
      float dlo = w.qpos9[k] - m.ch_range[k][0], dhi = m.ch_range[k][1] - w.qpos9[k];
      for (int side = 0; side < 2; side++) {
        float d = side ? dhi : dlo, sg = side ? -1.0f : 1.0f;
        if (d >= 0 || nl >= 5) continue;
        float kk, b, imp, R;
        row_params(def_solref, def_solimp, d, 0, m.dof_invw[k], &kk, &b, &imp, &R);
        w.ldof[nl] = k; w.lsign[nl] = sg; w.lD[nl] = 1.0f / R; w.laref[nl] = -b * (sg * w.qvel[k]) - kk * imp * d;
        nl++;
      }
    }
    w.nlim = nl;
    int off = 0, nc = w.ncon, arm = nl > 0;
    for (int c = 0; c < nc; c++) {
      const JacoPair& P = m.pairs[w.cpair[c]];
      int nd = coop_ndir(P.condim);
      if (off + nd > COOP_NROW) { nc = c; break; }   // row pool exhausted: later contacts are dropped (documented cap)
      w.croff[c] = off; off += nd;
      int mk = body_block(m.geoms[P.g1].body) | body_block(m.geoms[P.g2].body);
      w.cmask[c] = mk; w.ccond[c] = P.condim;
      if (mk & 1) arm = 1;
    }
    w.ncon = nc; w.arm_rows = arm;
  }
  COOP_SYNC();
  COOP_BEGIN {  // per contact: frame-independent parameters, aref components (mj_makeImpedance / mj_referenceConstraint)
    int c = lane;
    if (c < w.ncon) {
      const JacoPair& P = m.pairs[w.cpair[c]];
      float u[3], ww[3], comp[6], kk, b, imp, R0;
      contact_rel(w, c, w.bw, w.bv, u, ww);
      contact_comp(w, c, u, ww, comp);
      float mu0 = P.friction[0];
      float dA = (P.condim == 1) ? P.invw_tran : P.invw_tran + mu0 * mu0 * P.invw_tran;
      row_params(P.solref, P.solimp, w.cdist[c], P.margin, dA, &kk, &b, &imp, &R0);
      float Rr = (P.condim == 1) ? R0 : fmaxf(JMINVAL, 2 * mu0 * mu0 * R0);
      w.cD[c] = 1.0f / Rr;
      w.cA[c][0] = -b * comp[0] - kk * imp * (w.cdist[c] - P.margin);
      for (int k = 1; k < 6; k++) w.cA[c][k] = -b * comp[k];
      w.cmu[c][0] = 0;
      for (int k = 1; k < 6; k++) w.cmu[c][k] = k < P.condim ? contact_mu(P, k) : 0.0f;
    }
  } COOP_END
  COOP_BEGIN {  // direction Jacobians, one (contact, direction) task at a time
    for (int t = lane; t < w.ncon * 6; t += 32) {
      int c = t / 6, d = t - 6 * c;
      if (d >= coop_ndir(w.ccond[c])) continue;
      float t1[3], t2[3], zero[3] = {0, 0, 0};
      make_frame(w.cn[c], t1, t2);
      const float* dir = (d % 3 == 0) ? w.cn[c] : (d % 3 == 1 ? t1 : t2);
      if (d < 3) contact_row(w, c, dir, zero, w.Jd[w.croff[c] + d]);
      else contact_row(w, c, zero, dir, w.Jd[w.croff[c] + d]);
    }
  } COOP_END
}

// ------------------------------------------------------------------------------------------- solver pieces
// per-contact cost of the active pyramid rows at e + al p (rows_cost of jaco_sim.cuh for one contact)
JHD void coop_con_cost(const CoopWS& w, int c, float al, float* c0, float* g1, float* g2) {
  const int condim = w.ccond[c]; const float D = w.cD[c];
  if (condim == 1) {
    float j = w.ce[c][0] + al * w.cp[c][0], jp = w.cp[c][0];
    if (j < 0) { *c0 += 0.5f * D * j * j; *g1 += D * j * jp; *g2 += D * jp * jp; }
    return;
  }
  for (int k = 1; k < condim; k++) {
    float mu = w.cmu[c][k];
    for (int sg = 1; sg >= -1; sg -= 2) {
      float jp = w.cp[c][0] + sg * mu * w.cp[c][k];
      float j = w.ce[c][0] + sg * mu * w.ce[c][k] + al * jp;
      if (j < 0) { *c0 += 0.5f * D * j * j; *g1 += D * j * jp; *g2 += D * jp * jp; }
    }
  }
}
// reduce (cost, d1, d2) of the rows of component `mask` at step al; every lane returns the same totals
JHD JNOINLINE void coop_rows_cost(CoopWS& w, int mask, float al, float* cost, float* d1, float* d2) {
  COOP_BEGIN {
    float c0 = 0, g1 = 0, g2 = 0;
    int c = lane;
    if (c < w.ncon && (w.cmask[c] & mask)) coop_con_cost(w, c, al, &c0, &g1, &g2);
    if (lane == 31 && (mask & 1)) {
      for (int r = 0; r < w.nlim; r++) {
        float j = w.le[r] + al * w.lp[r];
        if (j < 0) { c0 += 0.5f * w.lD[r] * j * j; g1 += w.lD[r] * j * w.lp[r]; g2 += w.lD[r] * w.lp[r] * w.lp[r]; }
      }
    }
#ifdef __CUDA_ARCH__
    *cost = coop_wsum(c0); *d1 = coop_wsum(g1); *d2 = coop_wsum(g2);
#else
    w.red[lane][0] = c0; w.red[lane][1] = g1; w.red[lane][2] = g2;
#endif
  } COOP_END
#ifndef __CUDA_ARCH__
  float c0 = 0, g1 = 0, g2 = 0;
  for (int l = 0; l < 32; l++) { c0 += w.red[l][0]; g1 += w.red[l][1]; g2 += w.red[l][2]; }
  *cost = c0; *d1 = g1; *d2 = g2;
#endif
}
// component Jacobian times a workspace vector x: into ce (minus aref) or cp
JHD JNOINLINE void coop_eval(CoopWS& w, int mask, const float* x, bool is_dir) {
  COOP_BEGIN {
    for (int t = lane; t < w.ncon * 6; t += 32) {
      int c = t / 6, d = t - 6 * c;
      if (!(w.cmask[c] & mask) || d >= coop_ndir(w.ccond[c])) continue;
      const float* J = w.Jd[w.croff[c] + d];
      float sm = 0;
      for (int i = 0; i < JACO_NV; i++) sm += J[i] * x[i];
      if (is_dir) w.cp[c][d] = sm; else w.ce[c][d] = sm - w.cA[c][d];
    }
    if (lane == 31 && (mask & 1))
      for (int r = 0; r < w.nlim; r++) {
        float v = w.lsign[r] * x[w.ldof[r]];
        if (is_dir) w.lp[r] = v; else w.le[r] = v - w.laref[r];
      }
  } COOP_END
}
// y = M x restricted to the blocks in mask (row i by lane i)
JHD float coop_mul_M_row(const CoopWS& w, int i, const float* x) {
  float a = 0;
  if (i < 9) { for (int j = 0; j < 9; j++) a += w.Mc[i >= j ? PK(i, j) : PK(j, i)] * x[j]; }
  else { int f = (i - 9) / 6, r = i - 9 - 6 * f; for (int j = 0; j < 6; j++) a += w.Mf[f][r >= j ? PK(r, j) : PK(j, r)] * x[9 + 6 * f + j]; }
  return a;
}
// comps-space forces of every contact of the component (rows_force), then jf = J^T f by dof
JHD JNOINLINE void coop_force(CoopWS& w, int mask) {
  COOP_BEGIN {
    int c = lane;
    if (c < w.ncon && (w.cmask[c] & mask)) {
      const int condim = w.ccond[c]; const float D = w.cD[c];
      float sc[6] = {0, 0, 0, 0, 0, 0};
      if (condim == 1) { float j = w.ce[c][0]; if (j < 0) sc[0] = -D * j; }
      else for (int k = 1; k < condim; k++) {
        float mu = w.cmu[c][k];
        float jp_ = w.ce[c][0] + mu * w.ce[c][k], jm_ = w.ce[c][0] - mu * w.ce[c][k];
        float fp = jp_ < 0 ? -D * jp_ : 0.0f, fm = jm_ < 0 ? -D * jm_ : 0.0f;
        sc[0] += fp + fm; sc[k] = mu * (fp - fm);
      }
      for (int k = 0; k < 6; k++) w.csc[c][k] = sc[k];
      w.cf0[c] = sc[0];
    }
  } COOP_END
  COOP_BEGIN {
    int i = lane;
    if (i < JACO_NV) {
      float sm = 0;
      if (coop_block(i) & mask) {
        for (int c = 0; c < w.ncon; c++) {
          if (!(w.cmask[c] & mask) || !(w.cmask[c] & coop_block(i))) continue;
          const int nd = coop_ndir(w.ccond[c]);
          for (int d = 0; d < nd; d++) sm += w.Jd[w.croff[c] + d][i] * w.csc[c][d];
        }
        if (i < 9) for (int r = 0; r < w.nlim; r++) if (w.ldof[r] == i && w.le[r] < 0) sm += w.lsign[r] * (-w.lD[r] * w.le[r]);
      }
      w.jf[i] = sm;
    }
  } COOP_END
}
// comps-space Hessian weights of each contact: cW = {w00, w0k (k=1..5), wkk (k=1..5)}
JHD JNOINLINE void coop_weights(CoopWS& w, int mask) {
  COOP_BEGIN {
    int c = lane;
    if (c < w.ncon && (w.cmask[c] & mask)) {
      const int condim = w.ccond[c]; const float D = w.cD[c];
      float W[11];
      for (int k = 0; k < 11; k++) W[k] = 0;
      if (condim == 1) { if (w.ce[c][0] < 0) W[0] = D; }
      else for (int k = 1; k < condim; k++) {
        float mu = w.cmu[c][k];
        float ap = (w.ce[c][0] + mu * w.ce[c][k] < 0) ? D : 0.0f, am = (w.ce[c][0] - mu * w.ce[c][k] < 0) ? D : 0.0f;
        W[0] += ap + am; W[k] = mu * (ap - am); W[5 + k] = mu * mu * (ap + am);
      }
      for (int k = 0; k < 11; k++) w.cW[c][k] = W[k];
    }
  } COOP_END
}

// ------------------------------------------------------------------------------------------- device-only warp primitives
#ifdef __CUDA_ARCH__
// Newton direction p = -H^-1 g of component `mask`, H = M + J^T W J held in registers: lane l owns column l of the
// (symmetric) 21 x 21 Hessian, dofs outside the component are identity rows.  Assembly reads the direction
// Jacobians from the workspace (broadcast loads), the Cholesky factorisation is right-looking with one shuffle per
// rank-1 term, the two triangular solves exchange the factor through w.H.  Same arithmetic as the portable path
// below (coop_solve_component, CPU harness) up to summation order.
__device__ __noinline__ void coop_newton_dir_dev(CoopWS& w, int mask) {
  const int lane = threadIdx.x & 31;
  const int j = lane < JACO_NV ? lane : JACO_NV - 1;          // lanes >= 21 shadow dof 20 (never stored)
  const int bj = coop_block(j);
  const bool inb = (bj & mask) != 0;
  float h[JACO_NV];
#pragma unroll
  for (int i = 0; i < JACO_NV; i++) {
    float v = 0;
    const int bi = coop_block(i);
    if (inb && bi == bj) {
      if (i < 9) v = w.Mc[i >= j ? PK(i, j) : PK(j, i)];
      else { const int f = (i - 9) / 6, a = i - 9 - 6 * f, b = j - 9 - 6 * f; v = w.Mf[f][a >= b ? PK(a, b) : PK(b, a)]; }
    }
    if (!inb && i == j) v = 1.0f;
    h[i] = v;
  }
  if (mask & 1) {
    float add = 0;
    for (int r = 0; r < w.nlim; r++) if (w.ldof[r] == j && w.le[r] < 0) add += w.lD[r];
#pragma unroll
    for (int i = 0; i < 9; i++) if (i == j) h[i] += add;
  }
  for (int c = 0; c < w.ncon; c++) {
    if (!(w.cmask[c] & mask)) continue;
    const float* W = w.cW[c];
    const int r0 = w.croff[c], condim = w.ccond[c];
    const float J0 = w.Jd[r0][j];
    float T0 = W[0] * J0, Tk[5];
#pragma unroll
    for (int k = 1; k < 6; k++) {
      float t = 0;
      if (k < condim) { const float Jk = w.Jd[r0 + k][j]; T0 += W[k] * Jk; t = W[k] * J0 + W[5 + k] * Jk; }
      Tk[k - 1] = t;
    }
    if (condim <= 3) {
#pragma unroll
      for (int i = 0; i < JACO_NV; i++) {
        float a = w.Jd[r0][i] * T0;
        if (condim == 3) { a += w.Jd[r0 + 1][i] * Tk[0]; a += w.Jd[r0 + 2][i] * Tk[1]; }
        h[i] += a;
      }
    } else {
#pragma unroll
      for (int i = 0; i < JACO_NV; i++) {
        float a = w.Jd[r0][i] * T0;
#pragma unroll
        for (int k = 1; k < 6; k++) a += w.Jd[r0 + k][i] * Tk[k - 1];
        h[i] += a;
      }
    }
  }
  if (!inb) {   // keep excluded dofs exactly decoupled (their Jacobian entries are zero, this guards rounding paths)
#pragma unroll
    for (int i = 0; i < JACO_NV; i++) h[i] = (i == j) ? 1.0f : 0.0f;
  }
  // ---- right-looking Cholesky: after step c, h[c] of lane l >= c holds L[l][c]
#pragma unroll
  for (int c = 0; c < JACO_NV; c++) {
    float d = __shfl_sync(0xffffffffu, h[c], c);
    if (!(d > 0)) d = 1;
    const float ir = rsqrtf(d);
    const float lc = (lane == c) ? d * ir : h[c] * ir;
    h[c] = lc;
#pragma unroll
    for (int k = c + 1; k < JACO_NV; k++) {
      const float lk = __shfl_sync(0xffffffffu, lc, k);   // L[k][c]
      h[k] -= lc * lk;                                   // lane l: H[k][l] -= L[l][c] L[k][c]  (used for l >= k)
    }
  }
  // ---- L y = -g (column oriented), then L^T x = y with the factor's rows read back from the workspace
  float b = (lane < JACO_NV && inb) ? -w.g[lane] : 0.0f;
  float dinv = 1.0f;
#pragma unroll
  for (int c = 0; c < JACO_NV; c++) if (c == j) dinv = 1.0f / h[c];
#pragma unroll
  for (int c = 0; c < JACO_NV; c++) {
    const float yc = __shfl_sync(0xffffffffu, b * dinv, c);
    if (lane == c) b = yc; else if (lane > c) b -= h[c] * yc;
  }
  if (lane < JACO_NV) {
#pragma unroll
    for (int c = 0; c < JACO_NV; c++) if (c <= lane) w.H[PK(lane, c)] = h[c];
  }
  __syncwarp();
#pragma unroll
  for (int c = JACO_NV - 1; c >= 0; c--) {
    const float xc = __shfl_sync(0xffffffffu, b * dinv, c);
    if (lane == c) b = xc; else if (lane < c) b -= w.H[PK(c, j)] * xc;   // L[c][lane]
  }
  __syncwarp();
  if (lane < JACO_NV) w.p[lane] = inb ? b : 0.0f;
  __syncwarp();
}
#endif

// ------------------------------------------------------------------------------------------- Newton per component
// Solves the rows of the blocks in `mask` (bit 1 chain, 2 object, 4 dest) jointly; w.a holds the iterate (all 21
// entries; blocks outside `mask` are not touched).
JHD JNOINLINE void coop_solve_component(CoopWS& w, int mask) {
  // compact dof list of the component
  int n = 0, dofs[JACO_NV];
  for (int i = 0; i < JACO_NV; i++) if (coop_block(i) & mask) dofs[n++] = i;
  const int nent = n * (n + 1) / 2;
  JPROF_T(s0);
  // ---- start point: warm start if it beats qacc_smooth (mj_solNewton warm-start rule)
  float cw_, cs_, t1_, t2_;
  COOP_BEGIN {
    int i = lane;
    if (i < JACO_NV) { const bool in = (coop_block(i) & mask) != 0; if (in) w.a[i] = w.warm[i]; w.p[i] = in ? w.warm[i] - w.asm_[i] : 0.0f; }
    int c = lane;
    if (c < w.ncon) for (int k = 0; k < 6; k++) w.cp[c][k] = 0;   // direction components are unused until the first line search
    if (lane == 31) for (int r = 0; r < 5; r++) w.lp[r] = 0;
  } COOP_END
  coop_eval(w, mask, w.a, false);
  COOP_BEGIN { int i = lane; if (i < JACO_NV) w.Mp[i] = (coop_block(i) & mask) ? coop_mul_M_row(w, i, w.p) : 0.0f; } COOP_END
  coop_rows_cost(w, mask, 0.0f, &cw_, &t1_, &t2_);
  { float gs = 0; for (int i = 0; i < JACO_NV; i++) if (coop_block(i) & mask) gs += 0.5f * w.Mp[i] * w.p[i]; cw_ += gs; }
  // park the warm-start residuals (csc / g are free here), evaluate at qacc_smooth
  COOP_BEGIN { int c = lane; if (c < w.ncon) for (int k = 0; k < 6; k++) w.csc[c][k] = w.ce[c][k]; if (lane == 31) for (int r = 0; r < 5; r++) w.g[r] = w.le[r]; } COOP_END
  coop_eval(w, mask, w.asm_, false);
  coop_rows_cost(w, mask, 0.0f, &cs_, &t1_, &t2_);
  const bool usew = cw_ < cs_;
  COOP_BEGIN {
    int i = lane;
    if (i < JACO_NV && (coop_block(i) & mask)) w.a[i] = usew ? w.warm[i] : w.asm_[i];
    int c = lane;
    if (usew) { if (c < w.ncon && (w.cmask[c] & mask)) for (int k = 0; k < 6; k++) w.ce[c][k] = w.csc[c][k]; if (lane == 31 && (mask & 1)) for (int r = 0; r < w.nlim; r++) w.le[r] = w.g[r]; }
  } COOP_END
  int iters = 0;
  JPROF_T(s1); JPROF_ADD(10, s1 - s0);
  for (int it = 0; it < JACO_SOLVER_MAXIT; it++) {
    JPROF_T(i0);
    // ---- gradient g = M (a - asm) - J^T f
    coop_force(w, mask);
    COOP_BEGIN { int i = lane; if (i < JACO_NV) w.p[i] = (coop_block(i) & mask) ? w.a[i] - w.asm_[i] : 0.0f; } COOP_END
    COOP_BEGIN { int i = lane; if (i < JACO_NV) { float md = (coop_block(i) & mask) ? coop_mul_M_row(w, i, w.p) : 0.0f; w.Mp[i] = md; w.g[i] = md - w.jf[i]; } } COOP_END
    bool conv = true;  // block-wise relative gradient test (each block against its own force scale)
#ifdef __CUDA_ARCH__
    {
      const int ln = threadIdx.x & 31;
      const bool v = ln < JACO_NV;
      const int bl = v ? coop_block(ln) : 0;
      const float gi = v ? w.g[ln] * w.g[ln] : 0.0f, fi = v ? w.jf[ln] * w.jf[ln] + w.Mp[ln] * w.Mp[ln] : 0.0f;
      for (int b = 1; b <= 4; b <<= 1) {
        if (!(mask & b)) continue;
        const float gn = coop_wsum(bl == b ? gi : 0.0f), fn = coop_wsum(bl == b ? fi : 0.0f);
        if (!(gn <= 1e-12f * fn || gn < 1e-20f)) conv = false;
      }
    }
#else
    for (int b = 1; b <= 4; b <<= 1) {
      if (!(mask & b)) continue;
      float gn = 0, fn = 0;
      for (int i = 0; i < JACO_NV; i++) if (coop_block(i) == b) { fn += w.jf[i] * w.jf[i] + w.Mp[i] * w.Mp[i]; gn += w.g[i] * w.g[i]; }
      if (!(gn <= 1e-12f * fn || gn < 1e-20f)) conv = false;
    }
#endif
    if (conv) break;
    JPROF_T(i1); JPROF_ADD(11, i1 - i0);
    // ---- Hessian over the component's dofs
    coop_weights(w, mask);
#ifdef __CUDA_ARCH__
    coop_newton_dir_dev(w, mask);
    JPROF_T(i2); JPROF_ADD(12, i2 - i1);
#else
    COOP_BEGIN {
      for (int e = lane; e < nent; e += 32) {
        int ci = 0; while ((ci + 1) * (ci + 2) / 2 <= e) ci++;
        int cj = e - ci * (ci + 1) / 2;
        const int i = dofs[ci], j = dofs[cj];
        const int bi = coop_block(i), bj = coop_block(j);
        float hv = 0;
        if (bi == bj) {
          if (i < 9) hv = w.Mc[PK(i, j)];
          else { int f = (i - 9) / 6; hv = w.Mf[f][PK(i - 9 - 6 * f, j - 9 - 6 * f)]; }
        }
        if (i == j && i < 9) for (int r = 0; r < w.nlim; r++) if (w.ldof[r] == i && w.le[r] < 0) hv += w.lD[r];
        for (int c = 0; c < w.ncon; c++) {
          const int cm = w.cmask[c];
          if (!(cm & mask) || !(cm & bi) || !(cm & bj)) continue;
          const float* W = w.cW[c];
          const int r0 = w.croff[c], condim = w.ccond[c];
          const float J0i = w.Jd[r0][i], J0j = w.Jd[r0][j];
          float acc = W[0] * J0j, sm = 0;
          for (int k = 1; k < condim; k++) {
            const float Jki = w.Jd[r0 + k][i], Jkj = w.Jd[r0 + k][j];
            acc += W[k] * Jkj;
            sm += Jki * (W[k] * J0j + W[5 + k] * Jkj);
          }
          hv += J0i * acc + sm;
        }
        w.H[e] = hv;
      }
    } COOP_END
    // ---- Cholesky (left-looking, lanes over rows), then p = -H^-1 g
    for (int j = 0; j < n; j++) {
      COOP_BEGIN {
        int i = j + lane;
        if (i < n) { float s = w.H[PK(i, j)]; for (int k = 0; k < j; k++) s -= w.H[PK(i, k)] * w.H[PK(j, k)]; w.red[lane][0] = s; }
      } COOP_END
      float d = w.red[0][0];
      if (!(d > 0)) d = 1;
      const float ir = 1.0f / sqrtf(d);
      COOP_BEGIN { int i = j + lane; if (i < n) w.H[PK(i, j)] = (lane == 0) ? sqrtf(d) : w.red[lane][0] * ir; } COOP_END
    }
    COOP_BEGIN { int k = lane; if (k < n) w.red[k][1] = -w.g[dofs[k]]; } COOP_END
    for (int j = 0; j < n; j++) {  // forward substitution, column oriented
      COOP_SYNC();
      const float xj = w.red[j][1] / w.H[PK(j, j)];
      COOP_BEGIN { int i = j + lane; if (i == j) w.red[j][1] = xj; else if (i < n) w.red[i][1] -= w.H[PK(i, j)] * xj; } COOP_END
    }
    for (int j = n - 1; j >= 0; j--) {  // back substitution with L^T
      COOP_SYNC();
      const float xj = w.red[j][1] / w.H[PK(j, j)];
      COOP_BEGIN { int i = lane; if (i == j) w.red[j][1] = xj; else if (i < j) w.red[i][1] -= w.H[PK(j, i)] * xj; } COOP_END
    }
    COOP_BEGIN { int i = lane; if (i < JACO_NV) w.p[i] = 0; } COOP_END
    COOP_BEGIN { int k = lane; if (k < n) w.p[dofs[k]] = w.red[k][1]; } COOP_END
#endif
    JPROF_T(i3); JPROF_ADD(13, i3 - i2);
    // ---- exact line search on the piecewise-quadratic phi(al): safeguarded 1-D Newton
    coop_eval(w, mask, w.p, true);
    float q1 = 0, q2 = 0;
    {  // Mp currently holds M (a - asm); q1 = p . M d, q2 = p . M p
#ifdef __CUDA_ARCH__
      const int ln = threadIdx.x & 31;
      float mp = 0, pi = 0, mdi = 0;
      if (ln < JACO_NV && (coop_block(ln) & mask)) { mp = coop_mul_M_row(w, ln, w.p); pi = w.p[ln]; mdi = w.Mp[ln]; }
      q2 = coop_wsum(pi * mp); q1 = coop_wsum(pi * mdi);
#else
      COOP_BEGIN { int i = lane; if (i < JACO_NV) { w.red[i][2] = (coop_block(i) & mask) ? coop_mul_M_row(w, i, w.p) : 0.0f; } } COOP_END
      for (int i = 0; i < JACO_NV; i++) { q2 += w.p[i] * w.red[i][2]; q1 += w.p[i] * w.Mp[i]; }
#endif
    }
    float lo = 0, hi = -1, al = 0, d1, d2, c0;
    coop_rows_cost(w, mask, 0.0f, &c0, &d1, &d2);
    const float crows0 = c0;   // constraint part of the cost at the current iterate
    float dlo = q1 + d1; const float slope0 = dlo;
    if (!(dlo < 0)) break;
    al = -dlo / (q2 + d2);
    for (int ls = 0; ls < 12; ls++) {
      coop_rows_cost(w, mask, al, &c0, &d1, &d2);
      float dd = q1 + q2 * al + d1, hh = q2 + d2;
      if (fabsf(dd) <= 1e-6f * fabsf(slope0)) break;
      if (dd < 0) { lo = al; dlo = dd; } else { hi = al; }
      float nx = al - dd / hh;
      if (hi > 0 && (nx <= lo || nx >= hi)) nx = 0.5f * (lo + hi);
      if (nx == al) break;
      al = nx;
    }
    JPROF_T(i4); JPROF_ADD(14, i4 - i3);
    // ---- update iterate and residuals; per-dof step test
    bool small = true;
    bool stagnant = false;
#ifdef JACO_SOLVER_RTOL
    // mj_solNewton's improvement test (stop when scale * (cost - newcost) < tolerance) in float32 terms: the exact line search lowers
    // the cost by about -slope0 * al / 2; once that is below JACO_SOLVER_RTOL x the cost itself the iterates only move within float32
    // rounding of the cost (the float64 tolerance of 1e-8 / scale is far below that resolution).  The step is still taken.
    {
      float gauss = 0;
#ifdef __CUDA_ARCH__
      const int ln = threadIdx.x & 31;
      gauss = coop_wsum((ln < JACO_NV && (coop_block(ln) & mask)) ? (w.a[ln] - w.asm_[ln]) * w.Mp[ln] : 0.0f);
#else
      for (int i = 0; i < JACO_NV; i++) if (coop_block(i) & mask) gauss += (w.a[i] - w.asm_[i]) * w.Mp[i];
#endif
      stagnant = -0.5f * slope0 * al <= JACO_SOLVER_RTOL * (0.5f * gauss + crows0);
    }
#endif
#ifdef __CUDA_ARCH__
    {
      const int ln = threadIdx.x & 31;
      bool ok = true;
      if (ln < JACO_NV && (coop_block(ln) & mask)) { float st = fabsf(al * w.p[ln]); ok = st <= 2e-7f * fmaxf(1.0f, fabsf(w.a[ln] + al * w.p[ln])); }
      small = __all_sync(0xffffffffu, ok);
    }
#else
    for (int i = 0; i < JACO_NV; i++) if (coop_block(i) & mask) { float st = fabsf(al * w.p[i]); if (!(st <= 2e-7f * fmaxf(1.0f, fabsf(w.a[i] + al * w.p[i])))) small = false; }
#endif
    COOP_BEGIN {
      int i = lane;
      if (i < JACO_NV && (coop_block(i) & mask)) w.a[i] += al * w.p[i];
      int c = lane;
      if (c < w.ncon && (w.cmask[c] & mask)) for (int k = 0; k < 6; k++) w.ce[c][k] += al * w.cp[c][k];
      if (lane == 31 && (mask & 1)) for (int r = 0; r < w.nlim; r++) w.le[r] += al * w.lp[r];
    } COOP_END
    iters = it + 1;
    JPROF_T(i5); JPROF_ADD(15, i5 - i4);
    if (small || stagnant) break;
  }
  coop_force(w, mask);
  COOP_BEGIN { int i = lane; if (i < JACO_NV && (coop_block(i) & mask)) { w.qacc[i] = w.a[i]; w.fcon[i] = w.jf[i]; } if (lane == 0 && iters > w.iters) w.iters = iters; } COOP_END
}

// ------------------------------------------------------------------------------------------- touch sensors
JHD JNOINLINE void coop_touch(CoopWS& w) {
  const JacoModel& m = MDLG;
  COOP_BEGIN {
    int k = lane;
    if (k < JM_NSITE) {
      const JacoSite& S = m.sites[k];
      float acc = 0, spos[3], sR[9], t[3];
      m3mulv(t, w.bR[S.body], S.pos); v3add(spos, w.bx[S.body], t); m3mul(sR, w.bR[S.body], S.R);
      for (int c = 0; c < w.ncon; c++) {
        if (!(w.cf0[c] > 0)) continue;
        const JacoPair& P = m.pairs[w.cpair[c]];
        if (P.g1 != S.geom && P.g2 != S.geom) continue;
        float ray[3], sg = (P.g1 == S.geom) ? 1.0f : -1.0f;
        ray[0] = sg * w.cn[c][0]; ray[1] = sg * w.cn[c][1]; ray[2] = sg * w.cn[c][2];
        if (ray_hits_site(S, spos, sR, w.cpos[c], ray)) acc += w.cf0[c];
      }
      w.touch[k] = acc;
    }
  } COOP_END
}

// ------------------------------------------------------------------------------------------- driver
// Stage B.  In: the fully published forward quantities + cand[] of stage A.  Out: bit mask of the blocks
// (1 chain, 2 object, 4 dest) whose qacc / fcon were solved here -- the owner thread keeps its own fast solve for
// the others -- and (want_touch) touch[20].
JHD int coop_stageB(CoopWS& w) {
  COOP_SYNC();
  JPROF_T(c0);
  coop_collision(w);
  JPROF_T(c1);
  coop_constraints(w);
  JPROF_T(c2);
  JPROF_ADD(4, c1 - c0); JPROF_ADD(5, c2 - c1);
  const int nc = w.ncon, nl = w.nlim, flags = w.flags;
  COOP_BEGIN { int i = lane; if (i < JACO_NV) { w.qacc[i] = w.asm_[i]; w.fcon[i] = 0; w.a[i] = w.asm_[i]; } if (lane < JACO_MAXCON) w.cf0[lane] = 0; if (lane == 0) w.iters = 0; if (lane < 20) w.touch[lane] = 0; } COOP_END
  // connected components of the blocks under the contacts
  int comp[3] = {1, 2, 4}, rows[3] = {nl > 0 ? 1 : 0, 0, 0};
  for (int c = 0; c < nc; c++) {
    const int mk = w.cmask[c];
    int first = -1;
    for (int b = 0; b < 3; b++) if (mk & (1 << b)) {
      int root = 0;
      for (int r = 0; r < 3; r++) if (comp[r] & (1 << b)) root = r;
      if (first < 0) first = root;
      else if (root != first) { comp[first] |= comp[root]; comp[root] = 0; rows[first] |= rows[root]; rows[root] = 0; }
    }
    if (first >= 0) rows[first] = 1;
  }
  COOP_LANE0 {
    int fo = 0;
    for (int c = 0; c < nc; c++) {
      const JacoPair& P = MDLG.pairs[w.cpair[c]];
      const int b1 = MDLG.geoms[P.g1].body, b2 = MDLG.geoms[P.g2].body;
      if ((b1 == 10 && b2 >= 6 && b2 <= 9) || (b2 == 10 && b1 >= 6 && b1 <= 9)) fo = 1;
    }
    w.fingerobj = fo;
  }
  JPROF_T(c3);
  int solved = 0;
  for (int b = 0; b < 3; b++) {
    const int cm = comp[b];
    if (!cm || !rows[b]) continue;
    // a free box on its own with simple scene contacts stays with the owner's register solve
    const bool needs = (cm & 1) || ((cm & 2) && !(flags & COOP_SIMPLE_OBJ)) || ((cm & 6) == 6);
    if (!needs) continue;
    coop_solve_component(w, cm);
    solved |= cm;
  }
  JPROF_T(c4);
  JPROF_ADD(6, c4 - c3); JPROF_ADD(7, w.iters);
  if (w.want_touch && (solved & 1)) coop_touch(w);
  return solved;
}
