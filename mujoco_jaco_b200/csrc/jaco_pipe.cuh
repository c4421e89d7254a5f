// jaco_pipe.cuh -- the production env step as a per-substep pipeline with stream compaction of the rare paths.
//
// One JacoMujocoEnv.step (reference env_script/env_mujoco.py:116-150) = 50 physics substeps.  ~98 % of the environments
// of a picking roll-out stay in the "fast regime" (jaco_fast.cuh) for a whole env step; the rest need, on some or most
// substeps, work that is one to two orders of magnitude slower per environment and that diverges badly inside a warp:
// the narrow phase over the arm geoms, the general Newton contact solve (jaco_coop.cuh), the eigen-decomposition
// behind the OSC's numpy.linalg.pinv branch.  Round 1 served those inside one fused kernel; every block then waited
// for its slowest lane on every substep (ncu: 3.8 stalled warps per issue on the barrier, 16 active threads per
// instruction).  Here the env step is a pipeline over two CUDA streams per env range instead:
//
//   main stream   K_S(k), k = 0..49   one thread per env, substep k of the fast regime for every env that has not left
//                                     it ("clean").  An env that needs the general path at substep k publishes its
//                                     mj_forward quantities into its PipeRec, joins the dirty list and is skipped by
//                                     K_S(k+1..49).
//   dirty stream  iteration j = 0..49 (waits for K_S(j)):
//                   K_B   one WARP per pending task: cooperative collision / constraint rows / Newton solve / touch
//                         sensors (coop_stageB), then the Euler step and the next OSC torque (lanes 0..2), or the
//                         deferred pinv of the OSC;
//                   K_S'  one thread per dirty env: its next substep (same code as K_S), which may publish again.
//   join          K_E     one thread per env: observation, reward, terminal logic, statistics, in-kernel reset.
//
// State lives in the SoA buffer of the handle and makes one L2 round trip per substep (30 MB working set, 126 MB L2);
// HBM sees it once per env step.  Everything here is __host__ __device__: tests/harness runs the identical pipeline on
// the CPU (sequentially) for the CPU-side parity tests.
#pragma once
#include "jaco_fast.cuh"

// scratch rows [PIPE_NSCR][N]: per-env data that lives only inside one env step
enum { SC_U = 0, SC_TQ = 6, SC_EEPOS = 10, SC_EEQ = 13, SC_OBJP = 17, SC_RAMP0 = 20, SC_RAMP1 = 21, SC_TOUCH = 22, PIPE_NSCR = 23 };
enum { PEND_NONE = 0, PEND_B = 1, PEND_PINV = 2 };

// per-env record of a pending task (global memory, written by the owner thread, read by the serving warp)
struct PipeRec {
  CoopPub pub;                       // what coop_stageB consumes (copied word by word into the warp's workspace)
  float fsm[9], bias9[9], J[6][6];   // smooth chain forces (before constraint forces), chain bias, EE Jacobian
  float eepos[3], eeq[4];
  float fb_bias[2][6], fb_L[2][21];  // free boxes: bias force, Cholesky factor of the 6x6 inertia
  Con4 co, cd;                       // the owner's scene contacts (used when the cooperative path leaves a box to the 6-DoF solve)
  OscDefer od;                       // deferred pinv branch of the OSC
};

struct PipeBufs {
  float* state;      // [JACO_NSTATE][N]
  float* scratch;    // [PIPE_NSCR][N]
  PipeRec* rec;      // [N]
  int* it;           // [N] next substep of the env, -1 = step not begun
  int* pend;         // [N] PEND_*
  int* dirty;        // [N] env has left the main stream for this env step
  int* dead;         // [N] BAIL_* bits: the pipeline gave up on the env, the general path re-runs the whole step
  int N;
};
#define PST(r) B.state[(size_t)(r) * B.N + i]
#define PSC(r) B.scratch[(size_t)(r) * B.N + i]

#if defined(__CUDA_ARCH__)
#define PIPE_STAT(gs, k) atomicAdd((gs) + (k), 1.0)
#else
#define PIPE_STAT(gs, k) ((gs)[k] += 1.0)
#endif

// ------------------------------------------------------------------------------------------- K_S: one substep
// Substep B.it[i] of env i (the env step is begun first when it has not been).  Returns 1 when the env now has a pending
// task for the dirty stream (general contact pass or deferred pinv), else 0.  dirty_stream: the call comes from K_S'
// (the pinv branch then runs inline instead of being deferred once more).
JHD int pipe_env_substep(const JacoCfg& cfg, const PipeBufs& B, int i, const float* action_row, bool dirty_stream, double* gstats) {
  const int A = JACO_NQ + JACO_NV;
  int it = B.it[i];
  if (it >= 50) return 0;
  FastSim s;
  PipeRec& r = B.rec[i];
#pragma unroll
  for (int k = 0; k < JACO_NQ; k++) s.qpos[k] = PST(k);
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) s.qvel[k] = PST(JACO_NQ + k);
#pragma unroll
  for (int k = 9; k < JACO_NV; k++) s.warm[k] = PST(A + AUX_WARM + k);
  if (it < 0) {   // ---- begin the env step: controller data at the lagged state, _take_action, first OSC torque
#pragma unroll
    for (int k = 0; k < 9; k++) { s.lagq[k] = PST(A + AUX_LAGQ + k); s.lagv[k] = PST(A + AUX_LAGV + k); }
    s.grip = PST(A + AUX_GRIP);
    int rb = step_begin(cfg, s, action_row, OSC_DEFER, &r.od);
    if (rb == 2 && dirty_stream) { osc_finish_pinv(r.od, s.u); PIPE_STAT(gstats, STAT_PINV); rb = 1; }
#pragma unroll
    for (int k = 0; k < 6; k++) { PST(A + AUX_TARGET + k) = s.target[k]; PSC(SC_U + k) = s.u[k]; }
    PST(A + AUX_GRIP) = s.grip; PST(A + AUX_ACTION_IN) = 1.0f;
#pragma unroll
    for (int k = 0; k < 4; k++) { PSC(SC_TQ + k) = s.tq[k]; PSC(SC_EEQ + k) = s.eeq[k]; }
#pragma unroll
    for (int k = 0; k < 3; k++) PSC(SC_EEPOS + k) = s.eepos[k];
    PSC(SC_RAMP0) = s.ramp0; PSC(SC_RAMP1) = s.ramp1; PSC(SC_TOUCH) = 0.0f;
    it = 0;
    if (rb == 2) { B.it[i] = 0; B.pend[i] = PEND_PINV; return 1; }
    if (rb == 0) { B.dead[i] = BAIL_OSC; B.it[i] = 50; return 0; }
  } else {
#pragma unroll
    for (int k = 0; k < 6; k++) { s.u[k] = PSC(SC_U + k); s.target[k] = PST(A + AUX_TARGET + k); }
#pragma unroll
    for (int k = 0; k < 4; k++) s.tq[k] = PSC(SC_TQ + k);
#pragma unroll
    for (int k = 0; k < 3; k++) s.eepos[k] = PSC(SC_EEPOS + k);
    s.grip = PST(A + AUX_GRIP); s.ramp0 = PSC(SC_RAMP0); s.ramp1 = PSC(SC_RAMP1); s.action_in = 1;
  }
  SceneBoxes sb; scene_static(sb);
  SubFwd f;
  sub_forward<false>(s, sb, it, f);
  if (it == 49) {
#pragma unroll
    for (int k = 0; k < 3; k++) PSC(SC_OBJP + k) = s.objp[k];
  }
  if (f.bail) {
    // stage A of the arm collision by the owner: candidate pairs only; the narrow phase belongs to the serving warp
    unsigned long long cand = 0;
    if (f.bail & (BAIL_HAND | BAIL_FOREARM)) { LocalSink ks; fast_arm_collide(f.c, f.fo, f.fd, ks, &cand); }
    const int flags = ((f.bail & (BAIL_OBJ | BAIL_DEST)) ? 0 : COOP_SIMPLE_OBJ) | ((f.bail & (BAIL_LIMIT | BAIL_OBJ | BAIL_DEST)) ? COOP_FORCE_B : 0);
    if (cand != 0 || (flags & COOP_FORCE_B)) {
      coop_publish(r.pub, s, f.c, f.fo, f.fd, f.fsm, it == 49);
      r.pub.flags = flags; r.pub.cand[0] = (unsigned)(cand & 0xFFFFFFFFull) | 0xFu; r.pub.cand[1] = (unsigned)(cand >> 32); r.pub.pre_ncon = 0;
      for (int k = 0; k < 9; k++) { r.fsm[k] = f.fsm[k]; r.bias9[k] = f.c.bias[k]; }
      for (int a = 0; a < 6; a++) for (int b = 0; b < 6; b++) r.J[a][b] = f.c.J[a][b];
      v3copy(r.eepos, f.c.eepos);
      for (int k = 0; k < 4; k++) r.eeq[k] = f.c.eeq[k];
      for (int k = 0; k < 6; k++) { r.fb_bias[0][k] = f.fo.bias[k]; r.fb_bias[1][k] = f.fd.bias[k]; }
      for (int k = 0; k < 21; k++) { r.fb_L[0][k] = f.fo.L[k]; r.fb_L[1][k] = f.fd.L[k]; }
      r.co = f.co; r.cd = f.cd;
      // the free-box quaternions were normalised in registers by the forward: the serving warp integrates from the state rows
#pragma unroll
      for (int k = 12; k < JACO_NQ; k++) if (k != 16 && k != 17 && k != 18) PST(k) = s.qpos[k];
      B.it[i] = it; B.pend[i] = PEND_B;
      PIPE_STAT(gstats, STAT_STAGEB);
      return 1;
    }
  }
  int osc_ret;
  const int dd = sub_finish<false>(s, f, it, 0, nullptr, nullptr, OSC_DEFER, &r.od, &osc_ret);
  if (osc_ret == 2 && dirty_stream) { osc_finish_pinv(r.od, s.u); PIPE_STAT(gstats, STAT_PINV); osc_ret = 1; }
#pragma unroll
  for (int k = 0; k < JACO_NQ; k++) PST(k) = s.qpos[k];
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) PST(JACO_NQ + k) = s.qvel[k];
#pragma unroll
  for (int k = 9; k < JACO_NV; k++) PST(A + AUX_WARM + k) = s.warm[k];
  if (it == 49) {
#pragma unroll
    for (int k = 0; k < 9; k++) { PST(A + AUX_WARM + k) = s.warm[k]; PST(A + AUX_LAGQ + k) = s.lagq[k]; PST(A + AUX_LAGV + k) = s.lagv[k]; }
  }
#pragma unroll
  for (int k = 0; k < 3; k++) PSC(SC_EEPOS + k) = s.eepos[k];
#pragma unroll
  for (int k = 0; k < 4; k++) PSC(SC_EEQ + k) = s.eeq[k];
  if (dd & ~BAIL_OSC) { B.dead[i] = dd; B.it[i] = 50; return 0; }
  B.it[i] = it + 1;
  if (osc_ret == 2) { B.pend[i] = PEND_PINV; return 1; }
  if (osc_ret == 0) { B.dead[i] = BAIL_OSC; B.it[i] = 50; return 0; }
  if (it < 49) {
#pragma unroll
    for (int k = 0; k < 6; k++) PSC(SC_U + k) = s.u[k];
  }
  return 0;
}

// ------------------------------------------------------------------------------------------- K_B: one pending task
// Lanes 0 (chain), 1 (object), 2 (dest) finish the substep after the cooperative solve: mj_Euler of their block and,
// lane 0, the bookkeeping of the last substep and the OSC torque for the next one.
JHD void pipe_tail_chain(const JacoCfg& cfg, const PipeBufs& B, int i, const PipeRec& r, const CoopWS& w, int sv, int it, double* gstats) {
  (void)cfg;
  const int A = JACO_NQ + JACO_NV;
  const float h = MDL.dt;
  ChainOut c;
  for (int k = 0; k < 21; k++) c.A[k] = w.Mc[k];
  for (int f = 0; f < 3; f++) { for (int j = 0; j < 6; j++) c.B[f][j] = w.Mc[PK(6 + f, j)]; c.Dg[f] = w.Mc[PK(6 + f, 6 + f)]; }
  for (int k = 0; k < 9; k++) c.bias[k] = r.bias9[k];
  for (int a = 0; a < 6; a++) for (int b = 0; b < 6; b++) c.J[a][b] = r.J[a][b];
  v3copy(c.eepos, r.eepos);
  for (int k = 0; k < 4; k++) c.eeq[k] = r.eeq[k];
  float q[9], v[9], fsm[9], acc[9];
  for (int k = 0; k < 9; k++) { q[k] = w.qpos9[k]; v[k] = w.qvel[k]; fsm[k] = r.fsm[k]; }
  if (it == 49) {
    if (sv & 1) { for (int k = 0; k < 9; k++) acc[k] = w.qacc[k]; }
    else chain_solve(c, 0.0f, fsm, acc);
    for (int k = 0; k < 9; k++) { PST(A + AUX_WARM + k) = acc[k]; PST(A + AUX_LAGQ + k) = q[k]; PST(A + AUX_LAGV + k) = v[k]; }
  }
  if (sv & 1) { for (int k = 0; k < 9; k++) fsm[k] += w.fcon[k]; }
  int dead = 0;
  if (!chain_solve(c, h, fsm, acc)) dead |= BAIL_NAN;
  for (int k = 0; k < 9; k++) { v[k] += h * acc[k]; q[k] += h * v[k]; PST(k) = q[k]; PST(JACO_NQ + k) = v[k]; }
  for (int k = 0; k < 3; k++) PSC(SC_EEPOS + k) = r.eepos[k];
  for (int k = 0; k < 4; k++) PSC(SC_EEQ + k) = r.eeq[k];
  if (it < 49) {
    float tq[4], target[6], u[6];
    for (int k = 0; k < 4; k++) tq[k] = PSC(SC_TQ + k);
    for (int k = 0; k < 6; k++) target[k] = PST(A + AUX_TARGET + k);
    OscDefer od;
    const int rc = fast_osc(c, tq, target, v, u, OSC_DEFER, &od);
    if (rc == 2) { osc_finish_pinv(od, u); PIPE_STAT(gstats, STAT_PINV); }
    for (int k = 0; k < 6; k++) PSC(SC_U + k) = u[k];
  }
  if (dead) { B.dead[i] = dead; B.it[i] = 50; } else B.it[i] = it + 1;
}
JHD void pipe_tail_free(const PipeBufs& B, int i, const PipeRec& r, const CoopWS& w, int sv, int fbi) {
  const int A = JACO_NQ + JACO_NV;
  const int qo = 9 + 7 * fbi, vo = 9 + 6 * fbi, b = 10 + fbi;
  FreeBody fb;
  v3copy(fb.x, w.bx[b]);
  for (int k = 0; k < 9; k++) fb.R[k] = w.bR[b][k];
  for (int k = 0; k < 21; k++) { fb.M[k] = w.Mf[fbi][k]; fb.L[k] = r.fb_L[fbi][k]; }
  for (int k = 0; k < 6; k++) fb.bias[k] = r.fb_bias[fbi][k];
  float q7[7], v6[6], warm6[6];
  for (int k = 0; k < 7; k++) q7[k] = PST(qo + k);
  for (int k = 0; k < 6; k++) { v6[k] = PST(JACO_NQ + vo + k); warm6[k] = PST(A + AUX_WARM + vo + k); }
  if (sv & (2 << fbi)) free_integrate(fb, w.qacc + vo, w.fcon + vo, q7, v6, warm6);
  else fast_free_step(fb, fbi ? r.cd : r.co, q7, v6, warm6);
  for (int k = 0; k < 7; k++) PST(qo + k) = q7[k];
  for (int k = 0; k < 6; k++) { PST(JACO_NQ + vo + k) = v6[k]; PST(A + AUX_WARM + vo + k) = warm6[k]; }
}
// All 32 lanes of the serving warp (CPU harness: the COOP_LANES loops emulate them).
JHD void pipe_task(const JacoCfg& cfg, const PipeBufs& B, int i, CoopWS& w, double* gstats) {
  const int pend = B.pend[i];
  if (pend == PEND_NONE) return;
  PipeRec& r = B.rec[i];
  if (pend == PEND_PINV) {
    COOP_BEGIN {
      if (lane == 0) {
        float u[6];
        osc_finish_pinv(r.od, u);
        for (int k = 0; k < 6; k++) PSC(SC_U + k) = u[k];
        B.pend[i] = PEND_NONE;
        PIPE_STAT(gstats, STAT_PINV);
      }
    } COOP_END
    return;
  }
  const int it = B.it[i];
  COOP_BEGIN {   // record -> workspace, word by word (coalesced)
    const unsigned* src = reinterpret_cast<const unsigned*>(&r.pub);
    unsigned* dst = reinterpret_cast<unsigned*>(static_cast<CoopPub*>(&w));
    for (int k = lane; k < (int)(sizeof(CoopPub) / 4); k += 32) dst[k] = src[k];
  } COOP_END
  const int sv = coop_stageB(w);
  COOP_BEGIN {
    if (lane == 1 || lane == 2) pipe_tail_free(B, i, r, w, sv, lane - 1);
    if (lane == 0) {
      if (it == 49 && (sv & 1)) PSC(SC_TOUCH) = (float)touch_class(w.touch);
      if (w.ncon >= JACO_MAXCON) PIPE_STAT(gstats, STAT_CONCAP);
      if (w.iters >= JACO_SOLVER_MAXIT) PIPE_STAT(gstats, STAT_ITERCAP);
      if (w.fingerobj) PIPE_STAT(gstats, STAT_FINGEROBJ);
      pipe_tail_chain(cfg, B, i, r, w, sv, it, gstats);
      B.pend[i] = PEND_NONE;
    }
  } COOP_END
}

// ------------------------------------------------------------------------------------------- K_E: end of the env step
// Returns 0 done in place, 1 the env needs the general path for the whole step (state untouched by this call), 2 the
// episode ended and needs the general reset (outputs of the step are written; the reset kernel overwrites obs_row).
JHD int pipe_env_end(const JacoCfg& cfg, const PipeBufs& B, int i, int env_id, float* obs_row, float* reward, unsigned char* done, unsigned char* succ,
                     float* wb, float* term_obs_row, int auto_reset, float* local_stats) {
  const int A = JACO_NQ + JACO_NV;
  const int was_dead = B.dead[i];
  B.it[i] = -1; B.pend[i] = PEND_NONE; B.dirty[i] = 0; B.dead[i] = 0;
  if (was_dead) return 1;
  FastSim s;
  bool bad = false;
#pragma unroll
  for (int k = 0; k < JACO_NQ; k++) { s.qpos[k] = PST(k); bad = bad || !(fabsf(s.qpos[k]) < 1e6f); }
#pragma unroll
  for (int k = 0; k < JACO_NV; k++) { s.qvel[k] = PST(JACO_NQ + k); bad = bad || !(fabsf(s.qvel[k]) < 1e6f); }
  if (bad) return 1;
#pragma unroll
  for (int k = 0; k < 3; k++) { s.eepos[k] = PSC(SC_EEPOS + k); s.objp[k] = PSC(SC_OBJP + k); s.obj_goal[k] = PST(A + AUX_OBJ_GOAL + k); s.dest_goal[k] = PST(A + AUX_DEST_GOAL + k); }
#pragma unroll
  for (int k = 0; k < 4; k++) s.eeq[k] = PSC(SC_EEQ + k);
#pragma unroll
  for (int k = 0; k < 6; k++) { s.reach_goal[k] = PST(A + AUX_REACH_GOAL + k); s.target[k] = PST(A + AUX_TARGET + k); }
  s.touch_now = (int)PSC(SC_TOUCH);
  s.grip = PST(A + AUX_GRIP); s.action_in = 1; s.picked = (int)PST(A + AUX_PICKED);
  s.current_steps = (int)PST(A + AUX_CUR_STEPS); s.num_episodes = (int)PST(A + AUX_NUM_EPISODES);
  s.rng_counter = (uint32_t)PST(A + AUX_RNG) + ((uint32_t)PST(A + AUX_RNG_HI) << 24);
  s.accum_rew = PST(A + AUX_ACCUM_REW);
  float obs[JACO_NOBS], r, w; int d, sc, cause;
  if (!step_end(cfg, s, env_id, obs, &r, &d, &sc, &w, &cause)) return 1;
  s.accum_rew += r;
  local_stats[STAT_STEPS] += 1;
  int rc = 0;
  if (d) {
    local_stats[STAT_EPISODES] += 1; local_stats[STAT_SUCCESS] += sc; local_stats[STAT_RETURN] += s.accum_rew; local_stats[STAT_LENGTH] += s.current_steps;
    if (cause == 2) local_stats[STAT_TIMEOUT] += 1;
    else if (cause == 3) local_stats[STAT_SINGULAR] += 1;
    else if (cause == 4) local_stats[STAT_DROP] += 1;
  }
  if (term_obs_row) for (int k = 0; k < JACO_NOBS; k++) term_obs_row[k] = obs[k];
  *reward = r; *done = (unsigned char)d; *succ = (unsigned char)sc; *wb = w;
  PST(A + AUX_TOUCH) = (float)s.touch_index; PST(A + AUX_PICKED) = (float)s.picked;
  PST(A + AUX_CUR_STEPS) = (float)s.current_steps; PST(A + AUX_NUM_EPISODES) = (float)s.num_episodes;
  PST(A + AUX_RNG) = (float)(s.rng_counter & 0xFFFFFFu); PST(A + AUX_RNG_HI) = (float)(s.rng_counter >> 24);
  PST(A + AUX_ACCUM_REW) = s.accum_rew;
  if (d && auto_reset) {
    if (fast_reset(cfg, s, env_id, obs)) {
#pragma unroll
      for (int k = 0; k < JACO_NQ; k++) PST(k) = s.qpos[k];
#pragma unroll
      for (int k = 0; k < JACO_NV; k++) { PST(JACO_NQ + k) = s.qvel[k]; PST(A + AUX_WARM + k) = s.warm[k]; }
#pragma unroll
      for (int k = 0; k < 6; k++) { PST(A + AUX_TARGET + k) = s.target[k]; PST(A + AUX_REACH_GOAL + k) = s.reach_goal[k]; }
#pragma unroll
      for (int k = 0; k < 3; k++) { PST(A + AUX_OBJ_GOAL + k) = s.obj_goal[k]; PST(A + AUX_DEST_GOAL + k) = s.dest_goal[k]; }
#pragma unroll
      for (int k = 0; k < 9; k++) { PST(A + AUX_LAGQ + k) = s.lagq[k]; PST(A + AUX_LAGV + k) = s.lagv[k]; }
      PST(A + AUX_GRIP) = s.grip; PST(A + AUX_ACTION_IN) = 0.0f; PST(A + AUX_PICKED) = 0.0f;
      PST(A + AUX_CUR_STEPS) = 0.0f; PST(A + AUX_NUM_EPISODES) = 0.0f; PST(A + AUX_TOUCH) = (float)s.touch_index;
      PST(A + AUX_RNG) = (float)(s.rng_counter & 0xFFFFFFu); PST(A + AUX_RNG_HI) = (float)(s.rng_counter >> 24);
      PST(A + AUX_ACCUM_REW) = 0.0f;
    } else rc = 2;
  }
  for (int k = 0; k < JACO_NOBS; k++) obs_row[k] = obs[k];
  return rc;
}
#undef PST
#undef PSC
