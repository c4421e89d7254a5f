// host_harness.cpp -- TEST INFRASTRUCTURE.  Compiles the per-env device code (mujoco_jaco_b200/csrc/jaco_sim.cuh)
// for the CPU so that the CPU-only test-suite can check the kernel logic against the oracle on machines
// without a GPU.  Never loaded by the product; the shipped library only launches the CUDA kernels.
#include <string.h>
#define JACO_COLLISION_STATS 1
#include <vector>
#include "../../mujoco_jaco_b200/csrc/jaco_fast.cuh"
#include "../../mujoco_jaco_b200/csrc/jaco_render.cuh"

extern "C" {
void hh_init(const double* blob) { jaco_model_from_blob(blob, &g_jaco_model_host); }
static JacoCfg mkcfg(int task, int variant, unsigned long long seed) {
  JacoCfg c; c.task = task; c.variant = variant & 15; c.obs_mode = (variant >> 4) & 15; c.binary = (variant >> 8) & 1; c.seed = seed;
  c.task_max_steps = (task == 0 || task == 1 || task == 5 || task == 6) ? 500 : ((task == 2 || task == 3) ? 700 : 1200);
  return c;
}
// the headless camera's per-env and per-pixel code (jaco_render.cuh) without the tile culling: every primitive is tested per pixel
void hh_render(const float* state, int N, int i, const float* camera19, const float* prim_rgb, int W, int H, unsigned char* rgb, float* depth, float* prims_out) {
  JacoSim s;
  for (int k = 0; k < JACO_NQ; k++) s.qpos[k] = state[(size_t)k * N + i];
  std::vector<float> P(JACO_RENDER_NPRIM * 16);
  render_prims(s, P.data());
  if (prims_out) memcpy(prims_out, P.data(), sizeof(float) * P.size());
  JacoCamera cam;
  for (int k = 0; k < 3; k++) cam.pos[k] = camera19[k];
  for (int k = 0; k < 9; k++) cam.R[k] = camera19[3 + k];
  cam.tan_half_fovy = tanf(0.5f * camera19[12] * 3.14159265358979323846f / 180.0f);
  cam.znear = camera19[13]; cam.zfar = camera19[14]; cam.ambient = camera19[15];
  for (int k = 0; k < JACO_RENDER_NPRIM; k++) for (int c = 0; c < 3; c++) cam.rgb[k][c] = prim_rgb[3 * k + c];
  std::vector<float> Q(JACO_RENDER_NPRIM * 16);
  for (int g = 0; g < JACO_RENDER_NPRIM; g++) render_to_camera(cam, P.data() + 16 * g, Q.data() + 16 * g);
  for (int r = 0; r < H; r++) for (int c = 0; c < W; c++) {
    float x, y; render_pixel_xy(cam, W, H, c + 0.5f, r + 0.5f, &x, &y);
    render_shade(cam, Q.data(), nullptr, JACO_RENDER_NPRIM, x, y, rgb + 3 * ((size_t)r * W + c), depth + (size_t)r * W + c);
  }
}
void hh_reset(int task, int variant, unsigned long long seed, int env_offset, float* state, int N, float* obs) {
  JacoCfg c = mkcfg(task, variant, seed);
  for (int i = 0; i < N; i++) entry_reset(c, state, N, i, env_offset + i, obs ? obs + (size_t)i * JACO_NOBS : 0);
}
void hh_step(int task, int variant, unsigned long long seed, int env_offset, float* state, int N, const float* action, float* obs,
             float* reward, unsigned char* done, unsigned char* succ, float* wb, int auto_reset, float* stats16) {
  JacoCfg c = mkcfg(task, variant, seed);
  for (int i = 0; i < N; i++)
    entry_step(c, state, N, i, env_offset + i, action + (size_t)i * 7, obs + (size_t)i * JACO_NOBS, reward + i, done + i, succ + i, wb + i, 0, auto_reset, stats16);
}
void hh_substep(float* state, int N, const float* ctrl, int* ncon, float* sensordata, float* qacc) {
  for (int i = 0; i < N; i++) entry_substep(state, N, i, ctrl + (size_t)i * 9, ncon ? ncon + i : 0, sensordata ? sensordata + (size_t)i * 20 : 0, qacc ? qacc + (size_t)i * JACO_NV : 0);
}
// forward-only probe: M (chain packed 45), bias(21), Jee(36), eepos(3), eeq(4), qacc(21), ncon, iters
void hh_forward(float* state, int N, int i, const float* ctrl, float* Mc, float* bias, float* Jee, float* eepos, float* eeq, float* qacc, int* ncon, int* iters, float* contacts8) {
  JacoSim s; load_env(state, N, i, s);
  for (int k = 0; k < 9; k++) s.ctrl[k] = ctrl[k];
  mj_forward(s, true);
  memcpy(Mc, s.Mc, sizeof s.Mc); memcpy(bias, s.bias, sizeof s.bias); memcpy(Jee, s.Jee, sizeof s.Jee);
  memcpy(eepos, s.eepos, sizeof s.eepos); memcpy(eeq, s.eeq, sizeof s.eeq); memcpy(qacc, s.qacc, sizeof s.qacc);
  *ncon = s.ncon; *iters = s.solver_iters;
  if (contacts8) for (int c = 0; c < s.ncon; c++) { float* o = contacts8 + 8 * c; for (int k = 0; k < 3; k++) { o[k] = s.cpos[c][k]; o[3 + k] = s.cn[c][k]; } o[6] = s.cdist[c]; o[7] = s.cf0[c]; }
}
void hh_osc(const float* J, const float* Mc45, const float* bias6, const float* eepos, const float* eeq, const float* dq, const float* target, float* u) {
  JacoSim s; memset(&s, 0, sizeof s);
  memcpy(s.Jee, J, sizeof s.Jee); memcpy(s.Mc, Mc45, sizeof s.Mc); memcpy(s.bias, bias6, 6 * sizeof(float));
  memcpy(s.eepos, eepos, 12); memcpy(s.eeq, eeq, 16); memcpy(s.target, target, 24);
  osc_generate(s, dq, u);
}
}

// per-substep trace of one env step (debugging aid for the parity tests): out[k] = {qpos 23, qvel 21, ncon, iters, nlim}
extern "C" void hh_step_trace(int task, int variant, unsigned long long seed, int env_id, float* state, int N, int i, const float* action, float* out47x50) {
  JacoCfg c = mkcfg(task, variant, seed);
  JacoSim s; load_env(state, N, i, s); refresh_lagged(s);
  float a[7], u[6];
  for (int k = 0; k < 7; k++) a[k] = fminf(fmaxf(action[k], -1.0f), 1.0f);
  take_action(s, a, c.variant);
  for (int k = 0; k < 50; k++) {
    osc_generate(s, s.qvel, u);
    substep(s, u, k, k == 49);
    float* o = out47x50 + 47 * k;
    for (int j = 0; j < JACO_NQ; j++) o[j] = s.qpos[j];
    for (int j = 0; j < JACO_NV; j++) o[23 + j] = s.qvel[j];
    o[44] = (float)s.ncon; o[45] = (float)s.solver_iters;
    int mask = 0; for (int c = 0; c < s.ncon; c++) { const JacoPair& P = MDL.pairs[s.cpair[c]]; int m1 = body_block(MDL.geoms[P.g1].body), m2 = body_block(MDL.geoms[P.g2].body); if (m1 && m2 && m1 != m2) mask |= (m1 | m2); }
    o[46] = (float)(s.nlim + 100 * mask);
  }
}

extern "C" void hh_cstats(long long* out, int clear) { for (int i = 0; i < 16; i++) { out[i] = g_cstats[i]; if (clear) g_cstats[i] = 0; } }

// fast path (jaco_fast.cuh) with the same fall-back protocol as the CUDA launcher: bail -> general path from the
// untouched pre-step state; finished episodes are re-initialised by the general reset.  why[i] = bail reason bits.
extern "C" void hh_step_fast(int task, int variant, unsigned long long seed, int env_offset, float* state, int N, const float* action, float* obs,
             float* reward, unsigned char* done, unsigned char* succ, float* wb, int auto_reset, float* stats16, int* why, int allow_fallback) {
  JacoCfg c = mkcfg(task, variant, seed);
  for (int i = 0; i < N; i++) {
    int nr = 0; why[i] = 0;
    static CoopWS ws;
    int rc = allow_fallback == 2 ? entry_step_fast<true>(c, state, N, i, env_offset + i, action + (size_t)i * 7, obs + (size_t)i * JACO_NOBS, reward + i, done + i, succ + i, wb + i, 0, auto_reset, stats16, &nr, why + i, &ws, true)
           : entry_step_fast<false>(c, state, N, i, env_offset + i, action + (size_t)i * 7, obs + (size_t)i * JACO_NOBS, reward + i, done + i, succ + i, wb + i, 0, auto_reset, stats16, &nr, why + i);
    if (rc == FAST_OK) { if (nr) entry_reset(c, state, N, i, env_offset + i, obs + (size_t)i * JACO_NOBS); }
    else if (allow_fallback) entry_step(c, state, N, i, env_offset + i, action + (size_t)i * 7, obs + (size_t)i * JACO_NOBS, reward + i, done + i, succ + i, wb + i, 0, auto_reset, stats16);
  }
}
extern "C" void hh_wstats(long long* out, int clear) { for (int i = 0; i < 16; i++) { out[i] = g_fast_wstats[i]; if (clear) g_fast_wstats[i] = 0; } }
extern "C" void hh_pairstats(long long* a, long long* c) { for (int i = 0; i < 64; i++) { a[i] = g_coop_pairA[i]; c[i] = g_coop_pairC[i]; } }
extern "C" int hh_fast_model_ok() { return fast_model_ok(g_jaco_model_host) ? 1 : 0; }
// chain pass probe: M (9x9 dense), bias9, J 6x6, eepos, eeq at (q9, v9) with reference point O
extern "C" void hh_fast_chain(const float* q9, const float* v9, const float* O, float* M81, float* bias9, float* J36, float* eepos, float* eeq) {
  ChainOut c; fast_chain<true>(q9, v9, O, nullptr, c);
  for (int i = 0; i < 81; i++) M81[i] = 0;
  for (int i = 0; i < 6; i++) for (int j = 0; j <= i; j++) M81[9 * i + j] = M81[9 * j + i] = c.A[PK(i, j)];
  for (int f = 0; f < 3; f++) { M81[9 * (6 + f) + 6 + f] = c.Dg[f]; for (int j = 0; j < 6; j++) M81[9 * (6 + f) + j] = M81[9 * j + 6 + f] = c.B[f][j]; }
  memcpy(bias9, c.bias, 36); memcpy(J36, c.J, 144); memcpy(eepos, c.eepos, 12); memcpy(eeq, c.eeq, 16);
}
extern "C" void hh_body_spheres(float* out16) { memcpy(out16, g_jaco_model_host.body_sphere, 64); }
extern "C" void hh_scale_spheres(float f) { for (int b = 0; b < 4; b++) g_jaco_model_host.body_sphere[b][3] *= f; }
extern "C" int hh_sizeof_coopws() { return (int)sizeof(CoopWS); }

// The production env step (jaco_kernel_coop.cu) with the device's protocol: the hybrid path (fast path + cooperative general
// contact handling) for every env; the general path for what it hands back (NaN guard / bad pivot); the general reset where
// the in-place reset declined.  info[i]: bit 0 the cooperative path served the env, bit 1 general-path fall-back, bit 2 general reset.
extern "C" void hh_step_hybrid(int task, int variant, unsigned long long seed, int env_offset, float* state, int N, const float* action, float* obs,
             float* reward, unsigned char* done, unsigned char* succ, float* wb, int auto_reset, float* stats16, int* info) {
  JacoCfg c = mkcfg(task, variant, seed);
  static CoopWS ws;
  for (int i = 0; i < N; i++) {
    int nr = 0, why = 0; info[i] = 0;
    const float b0 = stats16[STAT_STAGEB];
    int rc = entry_step_fast<true>(c, state, N, i, env_offset + i, action + (size_t)i * 7, obs + (size_t)i * JACO_NOBS, reward + i, done + i, succ + i, wb + i, 0, auto_reset, stats16, &nr, &why, &ws, true);
    if (stats16[STAT_STAGEB] != b0) info[i] |= 1;
    if (rc == FAST_OK) { if (nr) { entry_reset(c, state, N, i, env_offset + i, obs + (size_t)i * JACO_NOBS); stats16[STAT_GENRESET] += 1; info[i] |= 4; } }
    else { entry_step(c, state, N, i, env_offset + i, action + (size_t)i * 7, obs + (size_t)i * JACO_NOBS, reward + i, done + i, succ + i, wb + i, 0, auto_reset, stats16); stats16[STAT_FALLBACK] += 1; info[i] |= 2; }
  }
}
