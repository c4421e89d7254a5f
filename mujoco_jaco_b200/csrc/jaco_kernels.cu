// jaco_kernels.cu -- the C ABI of libjaco_b200.so (include/jaco_b200.h) and the general-path kernels.
// The production env step is the fused hybrid kernel of jaco_kernel_coop.cu; the kernels here are the general
// thread-per-env path (reset, raw mj_step for parity injection, and the work-list kernel that re-runs the few envs the
// step kernel hands back).  State is SoA in HBM so every load/store of a warp is one coalesced 128-byte line per state
// row; model constants sit in __constant__ memory (uniform broadcast reads).  See DESIGN.md.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <new>
#include "../../include/jaco_b200.h"
#include "jaco_launch.h"
#include "jaco_render.cuh"

#define JACO_BLOCK 128

struct jaco_sim_handle {
  int n, device, env_offset, sms;
  JacoCfg cfg;
  float* state;       // [JACO_NSTATE][n]
  double* stats;      // [16] (double on the device: a float counter stops counting at 2^24)
  float *s_action, *s_obs, *s_reward, *s_wb; unsigned char *s_done, *s_succ;  // staging for jaco_step_host
  int use_fast;        // 2 fused hybrid kernel (production), 0 general kernel
  int* worklist;       // [2][n] env indices: row 0 = general-path fall-backs, row 1 = finished episodes that need the general reset
  int* counters;       // [16]: 0 n_fallback, 1 n_reset, 2..7 bail-reason histogram (cumulative, diagnostics)
  JacoModel* model;    // host copy: the model symbols are per library, a handle re-uploads them when another handle used them last
  long long launches;
  int* perm; unsigned char* key; int wave;   // env placement of the step kernel (jaco_launch.h JacoStepArgs)
  float* render_prims; int render_cap;   // jaco_render scratch: world primitives of the rendered envs (grown on demand)
  char err[256];
};
static jaco_sim_handle* g_model_owner = nullptr;   // handle whose model the __constant__ / __device__ symbols hold

// ---- scratch of jaco_step_noalloc (include/jaco_b200.h): the step kernels' work lists and counters in a file-scope device array
// instead of cudaMalloc'ed buffers.  Every GPU that loads the library holds these 128.25 KB until exit.  Layout (ints):
// [0, 2 * MAX) work lists (row 0 general-path fall-backs, row 1 general resets; row stride = the handle's n_envs), then 64 counters.
/* JACO_NOALLOC_MAX_ENVS (16384) comes from include/jaco_b200.h */
#define JACO_NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) int g_noalloc_scratch[2 * JACO_NOALLOC_MAX_ENVS + 64];
// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address a kernel may use is resolved once per
// device, with that device current, by cudaGetSymbolAddress (at set-up: jaco_create) and cached here.
static int* g_noalloc_ptr[JACO_NOALLOC_MAX_DEVICES];
static cudaError_t noalloc_resolve(int device) {
  if (device < 0 || device >= JACO_NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
  if (g_noalloc_ptr[device]) return cudaSuccess;
  void* p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);   // `device` is current (jaco_create has just set it)
  if (e == cudaSuccess) g_noalloc_ptr[device] = static_cast<int*>(p);
  return e;
}

__global__ void __launch_bounds__(JACO_BLOCK) jaco_reset_kernel(JacoCfg cfg, float* state, int N, int env_offset, const unsigned char* mask, float* obs) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  if (mask && !mask[i]) return;
  entry_reset(cfg, state, N, i, env_offset + i, obs ? obs + (size_t)i * JACO_NOBS : nullptr);
}

__global__ void __launch_bounds__(JACO_BLOCK) jaco_step_kernel(JacoCfg cfg, float* state, int N, int env_offset, const float* action, float* obs, float* reward,
                                                               unsigned char* done, unsigned char* succ, float* wb, float* term_obs, int auto_reset, double* gstats) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  float local[16];
#pragma unroll
  for (int k = 0; k < 16; k++) local[k] = 0;
  if (i < N) {
    float a[JACO_NACT];
#pragma unroll
    for (int k = 0; k < JACO_NACT; k++) a[k] = action[(size_t)i * JACO_NACT + k];
    float r, w; unsigned char d, sc;
    entry_step(cfg, state, N, i, env_offset + i, a, obs + (size_t)i * JACO_NOBS, &r, &d, &sc, &w,
               term_obs ? term_obs + (size_t)i * JACO_NOBS : nullptr, auto_reset, local);
    reward[i] = r; done[i] = d; succ[i] = sc; wb[i] = w;
  }
  block_stats_flush(local, gstats);
}

#define JACO_LIST_BLOCK 32

// general path over the hand-back list (from the pre-step snapshot), then the general reset over the finished-episode list
__global__ void __launch_bounds__(JACO_LIST_BLOCK) jaco_step_list_kernel(JacoCfg cfg, float* state, const float* state_prev, int N, int env_offset, const float* action, float* obs, float* reward,
                                                                         unsigned char* done, unsigned char* succ, float* wb, float* term_obs, int auto_reset, double* gstats,
                                                                         const int* worklist, const int* counters) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int nb = counters[0], nr = counters[1];
  if ((int)(blockIdx.x * blockDim.x) >= nb + nr) return;
  float local[16];
#pragma unroll
  for (int k = 0; k < 16; k++) local[k] = 0;
  if (t < nb) {
    int i = worklist[t];
    if (state_prev) for (int r = 0; r < JACO_NSTATE; r++) state[(size_t)r * N + i] = state_prev[(size_t)r * N + i];
    local[STAT_FALLBACK] += 1;
    float a[JACO_NACT];
#pragma unroll
    for (int k = 0; k < JACO_NACT; k++) a[k] = action[(size_t)i * JACO_NACT + k];
    float r, w; unsigned char d, sc;
    entry_step(cfg, state, N, i, env_offset + i, a, obs + (size_t)i * JACO_NOBS, &r, &d, &sc, &w,
               term_obs ? term_obs + (size_t)i * JACO_NOBS : nullptr, auto_reset, local);
    reward[i] = r; done[i] = d; succ[i] = sc; wb[i] = w;
  } else if (t < nb + nr) {
    int i = worklist[N + (t - nb)];
    local[STAT_GENRESET] += 1;
    entry_reset(cfg, state, N, i, env_offset + i, obs + (size_t)i * JACO_NOBS);
  }
  block_stats_flush(local, gstats);
}

__global__ void __launch_bounds__(JACO_BLOCK) jaco_substep_kernel(float* state, int N, const float* ctrl, int* ncon, float* sensordata, float* qacc) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float c[JACO_NU];
  for (int k = 0; k < JACO_NU; k++) c[k] = ctrl[(size_t)i * JACO_NU + k];
  entry_substep(state, N, i, c, ncon ? ncon + i : nullptr, sensordata ? sensordata + (size_t)i * 20 : nullptr, qacc ? qacc + (size_t)i * JACO_NV : nullptr);
}

// row-major API tensors <-> SoA state rows
__global__ void jaco_gather_kernel(const float* state, int N, int row0, int nrows, float* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  for (int r = 0; r < nrows; r++) out[(size_t)i * nrows + r] = state[(size_t)(row0 + r) * N + i];
}
__global__ void jaco_scatter_kernel(float* state, int N, int row0, int nrows, const float* in) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  for (int r = 0; r < nrows; r++) state[(size_t)(row0 + r) * N + i] = in[(size_t)i * nrows + r];
}

__global__ void jaco_stats_read_kernel(double* stats, float* out, int clear) {
  int k = threadIdx.x;
  if (k < JACO_NSTATS) { out[k] = (float)stats[k]; if (clear) stats[k] = 0.0; }
}

// FP32 FMA peak micro-benchmark: the denominator of the step kernel's compute roofline
__global__ void __launch_bounds__(256) jaco_fma_peak_kernel(float* out, int iters) {
  float x = 1.0f + 1e-7f * threadIdx.x, y = 1e-9f * blockIdx.x;
  float a0 = 0.1f, a1 = 0.2f, a2 = 0.3f, a3 = 0.4f, a4 = 0.5f, a5 = 0.6f, a6 = 0.7f, a7 = 0.8f;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 16; u++) {
      a0 = fmaf(a0, x, y); a1 = fmaf(a1, x, y); a2 = fmaf(a2, x, y); a3 = fmaf(a3, x, y);
      a4 = fmaf(a4, x, y); a5 = fmaf(a5, x, y); a6 = fmaf(a6, x, y); a7 = fmaf(a7, x, y);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

#define CK(call)                                                                                     \
  do {                                                                                               \
    cudaError_t e_ = (call);                                                                         \
    if (e_ != cudaSuccess) { snprintf(h->err, sizeof h->err, "%s: %s", #call, cudaGetErrorString(e_)); return JACO_ECUDA; } \
  } while (0)

// The model symbols (__constant__ / __device__, one copy per translation unit) belong to the library, not to a handle:
// a handle whose model is not the one currently uploaded re-uploads it, stream-ordered, before it launches anything.
static cudaError_t upload_model(jaco_sim_handle* h) {
  if (g_model_owner == h) return cudaSuccess;
  cudaError_t e = cudaMemcpyToSymbol(c_jaco_model, h->model, sizeof(JacoModel));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_jaco_model_dev, h->model, sizeof(JacoModel));
  if (e == cudaSuccess) e = jaco_coop_upload_model(h->model);
  if (e == cudaSuccess) g_model_owner = h;
  return e;
}

extern "C" {

int jaco_create(const void* model_blob, size_t nbytes, int n_envs, int task_id, int variant, int device, uint64_t seed, int env_offset, jaco_handle* out) {
  if (!model_blob || !out || n_envs <= 0 || nbytes != sizeof(double) * JM_TOTAL || task_id < 0 || task_id > JACO_CARRYING) return JACO_EINVAL;
  jaco_sim_handle* h = new (std::nothrow) jaco_sim_handle();
  if (!h) return JACO_ENOMEM;
  memset(h, 0, sizeof *h);
  *out = h;
  h->n = n_envs; h->device = device; h->env_offset = env_offset;
  h->cfg.task = task_id; h->cfg.variant = variant & 15; h->cfg.obs_mode = (variant >> 4) & 15; h->cfg.binary = (variant >> 8) & 1; h->cfg.seed = seed;
  h->cfg.task_max_steps = (task_id == JACO_REACHING || task_id == JACO_GRASPING || task_id == JACO_RELEASING || task_id == JACO_CARRYING) ? 500 : ((task_id == JACO_PICKING || task_id == JACO_PLACING) ? 700 : 1200);
  CK(cudaSetDevice(device));
  CK(noalloc_resolve(device));   /* set-up for jaco_step_noalloc: device address of its static scratch, once per device */
  h->model = new JacoModel();
  jaco_model_from_blob((const double*)model_blob, h->model);
  CK(upload_model(h));
  CK(cudaMalloc(&h->state, sizeof(float) * JACO_NSTATE * (size_t)n_envs));
  CK(cudaMemset(h->state, 0, sizeof(float) * JACO_NSTATE * (size_t)n_envs));
  h->use_fast = fast_model_ok(*h->model) ? 2 : 0;        /* production: the fused hybrid kernel; a model without the hard-wired scene structure runs the general kernel */
#ifdef JACO_DIAG   /* diagnostic builds only: the general thread-per-env kernel for cross-checks */
  if (getenv("JACO_NO_FAST")) h->use_fast = 0;
#endif
  CK(cudaMalloc(&h->worklist, sizeof(int) * 2 * (size_t)n_envs));
  h->sms = 148; cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, device);
  { /* env placement: on for every handle with at least two blocks; wave = one block per SM */
    const int grid = (n_envs + JACO_FAST_BLOCK - 1) / JACO_FAST_BLOCK;
    h->wave = grid < 2 ? 0 : (grid < h->sms ? grid : h->sms);
#ifdef JACO_DIAG
    if (getenv("JACO_SORT_WAVE")) { int w = atoi(getenv("JACO_SORT_WAVE")); if (w >= 0) h->wave = w > grid ? grid : w; }
#endif
    CK(cudaMalloc(&h->perm, sizeof(int) * (size_t)n_envs));
    CK(cudaMalloc(&h->key, (size_t)n_envs));
    CK(cudaMemset(h->key, 0, (size_t)n_envs));
    int* iota = (int*)malloc(sizeof(int) * (size_t)n_envs);
    if (!iota) return JACO_ENOMEM;
    for (int i = 0; i < n_envs; i++) iota[i] = i;
    cudaError_t ce = cudaMemcpy(h->perm, iota, sizeof(int) * (size_t)n_envs, cudaMemcpyHostToDevice);
    free(iota);
    CK(ce);
  }
  CK(cudaMalloc(&h->counters, sizeof(int) * 16));
  CK(cudaMemset(h->counters, 0, sizeof(int) * 16));
  CK(cudaMalloc(&h->stats, sizeof(double) * JACO_NSTATS));
  CK(cudaMemset(h->stats, 0, sizeof(double) * JACO_NSTATS));
  CK(cudaFuncSetAttribute(jaco_step_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 0));
  CK(cudaFuncSetAttribute(jaco_reset_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 0));
  return JACO_OK;
}

void jaco_destroy(jaco_handle h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  cudaFree(h->state); cudaFree(h->stats); cudaFree(h->worklist); cudaFree(h->counters);
  cudaFree(h->render_prims); cudaFree(h->perm); cudaFree(h->key);
  cudaFree(h->s_action); cudaFree(h->s_obs); cudaFree(h->s_reward); cudaFree(h->s_wb); cudaFree(h->s_done); cudaFree(h->s_succ);
  if (g_model_owner == h) g_model_owner = nullptr;
  delete h->model;
  delete h;
}
const char* jaco_last_error(jaco_handle h) { return h ? h->err : "null handle"; }

int jaco_render(jaco_handle h, const float* camera19, const float* prim_rgb, int width, int height, int env_begin, int env_count,
                uint8_t* rgb, float* depth, void* stream) {
  if (!h || !camera19 || !prim_rgb || !rgb || !depth || width <= 0 || height <= 0 || env_begin < 0 || env_count <= 0 || env_begin + env_count > h->n ||
      env_count > 65535) {
    if (h) snprintf(h->err, sizeof h->err, "jaco_render: bad arguments (env range [%d, %d) of %d, %d x %d; at most 65535 envs per call)", env_begin, env_begin + env_count, h->n, width, height);
    return JACO_EINVAL;
  }
  cudaStream_t st = (cudaStream_t)stream;
  CK(upload_model(h));
  if (env_count > h->render_cap) {
    CK(cudaStreamSynchronize(st));
    cudaFree(h->render_prims); h->render_prims = nullptr; h->render_cap = 0;
    CK(cudaMalloc(&h->render_prims, sizeof(float) * 16 * JACO_RENDER_NPRIM * (size_t)env_count));
    h->render_cap = env_count;
  }
  JacoCamera cam;
  for (int i = 0; i < 3; i++) cam.pos[i] = camera19[i];
  for (int i = 0; i < 9; i++) cam.R[i] = camera19[3 + i];
  cam.tan_half_fovy = tanf(0.5f * camera19[12] * 3.14159265358979323846f / 180.0f);
  cam.znear = camera19[13]; cam.zfar = camera19[14]; cam.ambient = camera19[15];
  if (!(cam.znear > 0) || !(cam.zfar > cam.znear)) { snprintf(h->err, sizeof h->err, "jaco_render: need 0 < znear < zfar"); return JACO_EINVAL; }
  for (int k = 0; k < JACO_RENDER_NPRIM; k++) for (int c = 0; c < 3; c++) cam.rgb[k][c] = prim_rgb[3 * k + c];
  jaco_render_prims_kernel<<<(env_count + 127) / 128, 128, 0, st>>>(cam, h->state, h->n, env_begin, env_count, h->render_prims);
  dim3 grid((width + JACO_RENDER_TX - 1) / JACO_RENDER_TX, (height + JACO_RENDER_TY - 1) / JACO_RENDER_TY, env_count);
  jaco_render_kernel<<<grid, dim3(JACO_RENDER_TX, JACO_RENDER_TY), 0, st>>>(cam, h->render_prims, width, height, rgb, depth);
  h->launches += 2;
  CK(cudaGetLastError());
  return JACO_OK;
}
int jaco_debug_counters(jaco_handle h, int* out16) {  /* host copy of the work-list counters (synchronises) */
  if (!h || !out16) return JACO_EINVAL;
  CK(cudaMemcpy(out16, h->counters, 16 * sizeof(int), cudaMemcpyDeviceToHost));
  return JACO_OK;
}
int jaco_n_envs(jaco_handle h) { return h ? h->n : 0; }
long long jaco_launch_count(jaco_handle h) { return h ? h->launches : 0; }

static inline int nblocks(int n) { return (n + JACO_BLOCK - 1) / JACO_BLOCK; }

int jaco_reset(jaco_handle h, const uint8_t* mask, float* obs, void* stream) {
  if (!h) return JACO_EINVAL;
  CK(upload_model(h));
  jaco_reset_kernel<<<nblocks(h->n), JACO_BLOCK, 0, (cudaStream_t)stream>>>(h->cfg, h->state, h->n, h->env_offset, mask, obs);
  h->launches++;
  CK(cudaGetLastError());
  return JACO_OK;
}

int jaco_step(jaco_handle h, const float* action, float* obs, float* reward, uint8_t* done, uint8_t* succ, float* wb, float* term_obs, int auto_reset, void* stream) {
  if (!h || !action || !obs || !reward || !done || !succ || !wb) return JACO_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  CK(upload_model(h));
  JacoStepArgs A = {h->cfg, h->state, h->n, h->env_offset, action, obs, reward, done, succ, wb, term_obs, auto_reset, h->stats, h->worklist, h->counters,
                    h->wave > 0 ? h->perm : nullptr, h->key, h->wave};
  const int nlist = (h->n + JACO_LIST_BLOCK - 1) / JACO_LIST_BLOCK;
  if (h->use_fast == 2) {
    CK(cudaMemsetAsync(h->counters, 0, 2 * sizeof(int), st));
    CK(jaco_coop_launch(A, st));
    jaco_step_list_kernel<<<nlist, JACO_LIST_BLOCK, 0, st>>>(h->cfg, h->state, nullptr, h->n, h->env_offset, action, obs, reward, done, succ, wb,
                                                           term_obs, auto_reset, h->stats, h->worklist, h->counters);
    h->launches += 2;
    if (h->wave > 0) { CK(jaco_place_launch(h->key, h->n, h->perm, h->counters + 8, st)); h->launches++; }   /* next step's placement */
  } else {
    jaco_step_kernel<<<nblocks(h->n), JACO_BLOCK, 0, st>>>(h->cfg, h->state, h->n, h->env_offset, action, obs, reward, done, succ, wb, term_obs, auto_reset, h->stats);
    h->launches++;
  }
  CK(cudaGetLastError());
  return JACO_OK;
}

// jaco_step with its scratch (work lists, counters) in the library's static device array: allocates nothing.  Same kernels, same
// arguments otherwise, so the outputs are bit-identical to jaco_step's.  All calls (all handles of a device) share the array: calls
// that overlap on the GPU race.
int jaco_step_noalloc(jaco_handle h, const float* action, float* obs, float* reward, uint8_t* done, uint8_t* succ, float* wb, float* term_obs, int auto_reset, void* stream) {
  if (!h || !action || !obs || !reward || !done || !succ || !wb) return JACO_EINVAL;
  if (h->n > JACO_NOALLOC_MAX_ENVS) {
    snprintf(h->err, sizeof h->err, "jaco_step_noalloc: %d envs exceed the static scratch (JACO_NOALLOC_MAX_ENVS = %d); use jaco_step", h->n, JACO_NOALLOC_MAX_ENVS);
    return JACO_EINVAL;
  }
  if (h->use_fast != 2) { snprintf(h->err, sizeof h->err, "jaco_step_noalloc: this model runs the general kernel, which has no scratch; use jaco_step"); return JACO_EINVAL; }
  int* scratch = (h->device >= 0 && h->device < JACO_NOALLOC_MAX_DEVICES) ? g_noalloc_ptr[h->device] : nullptr;
  if (!scratch) { snprintf(h->err, sizeof h->err, "jaco_step_noalloc: static scratch not resolved on device %d", h->device); return JACO_ECUDA; }
  cudaStream_t st = (cudaStream_t)stream;
  CK(upload_model(h));
  int* worklist = scratch;                                  // [2][h->n]
  int* counters = scratch + 2 * JACO_NOALLOC_MAX_ENVS;      // [16] of 64
  JacoStepArgs A = {h->cfg, h->state, h->n, h->env_offset, action, obs, reward, done, succ, wb, term_obs, auto_reset, h->stats, worklist, counters,
                    nullptr, h->key, 0};   /* identity placement: the entry point launches the two step kernels and nothing else */
  const int nlist = (h->n + JACO_LIST_BLOCK - 1) / JACO_LIST_BLOCK;
  CK(cudaMemsetAsync(counters, 0, 16 * sizeof(int), st));   // the array keeps the last call's contents: the list lengths must start at zero
  CK(jaco_coop_launch(A, st));
  jaco_step_list_kernel<<<nlist, JACO_LIST_BLOCK, 0, st>>>(h->cfg, h->state, nullptr, h->n, h->env_offset, action, obs, reward, done, succ, wb,
                                                         term_obs, auto_reset, h->stats, worklist, counters);
  h->launches += 2;
  CK(cudaGetLastError());
  return JACO_OK;
}

int jaco_step_host(jaco_handle h, const float* action, float* obs, float* reward, uint8_t* done, uint8_t* succ, float* wb, int auto_reset, void* stream) {
  if (!h || !action || !obs || !reward || !done || !succ || !wb) return JACO_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  size_t n = h->n;
  if (!h->s_action) {
    CK(cudaMalloc(&h->s_action, n * JACO_NACT * sizeof(float))); CK(cudaMalloc(&h->s_obs, n * JACO_NOBS * sizeof(float)));
    CK(cudaMalloc(&h->s_reward, n * sizeof(float))); CK(cudaMalloc(&h->s_wb, n * sizeof(float)));
    CK(cudaMalloc(&h->s_done, n)); CK(cudaMalloc(&h->s_succ, n));
  }
  CK(cudaMemcpyAsync(h->s_action, action, n * JACO_NACT * sizeof(float), cudaMemcpyHostToDevice, st));
  int rc = jaco_step(h, h->s_action, h->s_obs, h->s_reward, h->s_done, h->s_succ, h->s_wb, nullptr, auto_reset, stream);
  if (rc) return rc;
  CK(cudaMemcpyAsync(obs, h->s_obs, n * JACO_NOBS * sizeof(float), cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(reward, h->s_reward, n * sizeof(float), cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(wb, h->s_wb, n * sizeof(float), cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(done, h->s_done, n, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(succ, h->s_succ, n, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return JACO_OK;
}

int jaco_get_state(jaco_handle h, float* qpos, float* qvel, float* aux, void* stream) {
  if (!h) return JACO_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  int nb = (h->n + 255) / 256;
  if (qpos) { jaco_gather_kernel<<<nb, 256, 0, st>>>(h->state, h->n, 0, JACO_NQ, qpos); h->launches++; }
  if (qvel) { jaco_gather_kernel<<<nb, 256, 0, st>>>(h->state, h->n, JACO_NQ, JACO_NV, qvel); h->launches++; }
  if (aux) { jaco_gather_kernel<<<nb, 256, 0, st>>>(h->state, h->n, JACO_NQ + JACO_NV, JACO_NAUX, aux); h->launches++; }
  CK(cudaGetLastError());
  return JACO_OK;
}
int jaco_set_state(jaco_handle h, const float* qpos, const float* qvel, const float* aux, void* stream) {
  if (!h) return JACO_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  int nb = (h->n + 255) / 256;
  if (qpos) { jaco_scatter_kernel<<<nb, 256, 0, st>>>(h->state, h->n, 0, JACO_NQ, qpos); h->launches++; }
  if (qvel) { jaco_scatter_kernel<<<nb, 256, 0, st>>>(h->state, h->n, JACO_NQ, JACO_NV, qvel); h->launches++; }
  if (aux) { jaco_scatter_kernel<<<nb, 256, 0, st>>>(h->state, h->n, JACO_NQ + JACO_NV, JACO_NAUX, aux); h->launches++; }
  CK(cudaGetLastError());
  return JACO_OK;
}
int jaco_debug_substep(jaco_handle h, const float* ctrl, int32_t* ncon, float* sensordata, float* qacc, void* stream) {
  if (!h || !ctrl) return JACO_EINVAL;
  CK(upload_model(h));
  jaco_substep_kernel<<<nblocks(h->n), JACO_BLOCK, 0, (cudaStream_t)stream>>>(h->state, h->n, ctrl, ncon, sensordata, qacc);
  h->launches++;
  CK(cudaGetLastError());
  return JACO_OK;
}
int jaco_stats(jaco_handle h, float* out16, int clear, void* stream) {
  if (!h || !out16) return JACO_EINVAL;
  jaco_stats_read_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(h->stats, out16, clear);
  h->launches++;
  CK(cudaGetLastError());
  return JACO_OK;
}

// Measured FP32 FMA throughput of `device` in TFLOP/s (FMA = 2 flops); best of `reps` timed launches.
double jaco_fp32_peak_tflops(int device, int reps) {
  if (cudaSetDevice(device) != cudaSuccess) return -1.0;
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  int blocks = sms * 8, iters = 4096; float* out = nullptr;
  if (cudaMalloc(&out, sizeof(float) * blocks * 256) != cudaSuccess) return -1.0;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0;
  for (int r = 0; r < reps + 2; r++) {
    cudaEventRecord(e0);
    jaco_fma_peak_kernel<<<blocks, 256>>>(out, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    double tf = 2.0 * 128.0 * iters * (double)blocks * 256 / (ms * 1e-3) / 1e12;
    if (r >= 2 && tf > best) best = tf;
  }
  cudaFree(out); cudaEventDestroy(e0); cudaEventDestroy(e1);
  return best;
}

}  // extern "C"
