// jaco_kernel_coop.cu -- the production step kernel: fast path per thread + warp-cooperative general contact
// handling on demand (jaco_fast.cuh, jaco_coop.cuh), all 50 substeps of one env step in one launch.
//
// Alternatives measured and rejected this round (profiles/r02_*): a per-substep kernel pipeline with the rare
// paths on a second stream (90-114 ms per 65 536-env step against 44), and a two-pass split (bail-only fast kernel, then
// this kernel over the dirty list: 84-98 ms) -- the general contact pass is throughput-bound on instruction delivery and
// needs every SM's warps, which this single kernel gives it substep by substep.
#include "jaco_launch.h"
#define JACO_COOP_WSN ((JACO_FAST_BLOCK / 32) < JACO_COOP_NWS ? (JACO_FAST_BLOCK / 32) : JACO_COOP_NWS)

#ifdef JACO_DIAG   /* diagnostic builds: start / end time (globaltimer, ns) and SM of every block of the last step kernel launch */
__device__ unsigned long long g_block_t[3][2048];
__device__ int g_env_dbg[4][65536];   // per env of the last launch (first 65536): block, solve iterations (sum), max contacts, cap hits
__device__ __forceinline__ unsigned long long jaco_gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ unsigned jaco_smid() { unsigned s; asm volatile("mov.u32 %0, %smid;" : "=r"(s)); return s; }
#endif

// hybrid kernel: fast path per thread + warp-cooperative general contact handling on demand (jaco_coop.cuh);
// one CoopWS per warp in dynamic shared memory
__global__ void __launch_bounds__(JACO_FAST_BLOCK, JACO_FAST_MINBLOCKS) jaco_step_coop_kernel(const JacoStepArgs A) {
  const JacoCfg& cfg = A.cfg; float* state = A.state; const int N = A.N, env_offset = A.env_offset; const float* action = A.action;
  float* obs = A.obs; float* reward = A.reward; unsigned char* done = A.done; unsigned char* succ = A.succ; float* wb = A.wb;
  float* term_obs = A.term_obs; const int auto_reset = A.auto_reset; double* gstats = A.gstats; int* worklist = A.worklist; int* counters = A.counters;

#ifdef JACO_DIAG
  if (threadIdx.x == 0 && blockIdx.x < 2048) { g_block_t[0][blockIdx.x] = jaco_gtime(); g_block_t[2][blockIdx.x] = jaco_smid(); }
#endif
  extern __shared__ __align__(16) unsigned char coop_smem[];
  CoopWS* ws = reinterpret_cast<CoopWS*>(coop_smem) + (threadIdx.x >> 5);
  // Placement: a block's cost per substep is set by the union of the code paths its lanes take, hardly by how many lanes
  // take them.  The envs are served in the order of last step's workload key (perm: solving envs first, then the other busy ones, then
  // the quiet majority); the first `wave` blocks -- one per SM, they start together -- are dealt that order in snake fashion, so every
  // one of them gets the same share of every workload level and the busy envs sit in the same warps; the remaining blocks take the tail
  // of the order contiguously: quiet envs only, close to the uniform-flow rate, in the shadow of the busy blocks' tail.
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int wave = A.wave;
  if ((int)blockIdx.x < wave) p = threadIdx.x * wave + ((threadIdx.x & 1) ? wave - 1 - (int)blockIdx.x : (int)blockIdx.x);
  const bool active = p < N;
  const int i = A.perm ? A.perm[active ? p : N - 1] : (active ? p : N - 1);
  float local[16];
#pragma unroll
  for (int k = 0; k < 16; k++) local[k] = 0;
  float a[JACO_NACT];
#pragma unroll
  for (int k = 0; k < JACO_NACT; k++) a[k] = action[(size_t)i * JACO_NACT + k];
  float r, w; unsigned char d, sc; int needs_reset, why, busy[4] = {0, 0, 0, 0};
  int rc = entry_step_fast<true>(cfg, state, N, i, env_offset + i, a, obs + (size_t)i * JACO_NOBS, &r, &d, &sc, &w,
                                 term_obs ? term_obs + (size_t)i * JACO_NOBS : nullptr, auto_reset, local, &needs_reset, &why, ws, active, busy);
  if (active) {
    if (rc == FAST_OK) {
      reward[i] = r; done[i] = d; succ[i] = sc; wb[i] = w;
      if (needs_reset) worklist[N + atomicAdd(counters + 1, 1)] = i;
    } else {
      worklist[atomicAdd(counters + 0, 1)] = i;
      int bit = 31 - __clz(why & -why);
      atomicAdd(counters + 2 + (bit < 0 ? 0 : bit), 1);
    }
  }
  // placement key: substeps of this step that left the plain fast path (arm near the scene / joint limit / non-simple scene contacts,
  // whether or not a cooperative solve followed) + substeps on the pinv branch
  // 32..63: the env needed cooperative solves (ordered by how many); 1..31: busy without a solve (near / pinv substeps, 4 per level)
  if (active && A.key) {
    const int nb = (int)local[STAT_STAGEB], nq = busy[0] + (int)local[STAT_PINV];
    A.key[i] = (unsigned char)(nb > 0 ? 32 + min(31, (nb * 31 + 49) / 50) : min(31, (nq + 3) >> 2));
  }
  block_stats_flush(local, gstats);
#ifdef JACO_DIAG
  if (active && i < 65536) { g_env_dbg[0][i] = blockIdx.x; g_env_dbg[1][i] = busy[1]; g_env_dbg[2][i] = busy[2]; g_env_dbg[3][i] = busy[3] + 1000 * (int)local[STAT_STAGEB] + 100000 * (int)local[STAT_PINV]; }
  __syncthreads();
  if (threadIdx.x == 0 && blockIdx.x < 2048) g_block_t[1][blockIdx.x] = jaco_gtime();
#endif
}
#ifdef JACO_DIAG
extern "C" int jaco_diag_env_dbg(int* out4x65536) {
  return cudaMemcpyFromSymbol(out4x65536, g_env_dbg, sizeof(int) * 4 * 65536) == cudaSuccess ? 0 : -2;
}
extern "C" int jaco_diag_block_times(unsigned long long* out3x2048) {
  return cudaMemcpyFromSymbol(out3x2048, g_block_t, sizeof(unsigned long long) * 3 * 2048) == cudaSuccess ? 0 : -2;
}
#endif

// perm = env indices sorted by key (descending; 64 buckets), one block: warp-aggregated shared-memory counting sort.
// The order inside a bucket is arbitrary: an env's arithmetic does not depend on where it is placed.
__global__ void __launch_bounds__(1024) jaco_place_kernel(const unsigned char* key, int n, int* perm, int* dbg) {
  __shared__ int hist[64], off[64];
  const int tid = threadIdx.x, lane = tid & 31;
  if (tid < 64) hist[tid] = 0;
  __syncthreads();
  for (int e0 = 0; e0 < n; e0 += 1024) {
    const int e = e0 + tid;
    const int k = e < n ? key[e] & 63 : 64;
    const unsigned m = __match_any_sync(0xffffffffu, k);
    if (k < 64 && lane == __ffs(m) - 1) atomicAdd(&hist[k], __popc(m));
  }
  __syncthreads();
  if (tid == 0) {
    int run = 0; for (int k = 63; k >= 0; k--) { off[k] = run; run += hist[k]; }
    if (dbg) {   // diagnostics (jaco_debug_counters 8..11): quiet envs, busy envs without / with cooperative solves, envs at the top key
      int nA = 0, nB = 0;
      for (int k = 32; k < 64; k++) nA += hist[k];
      for (int k = 1; k < 32; k++) nB += hist[k];
      dbg[0] = hist[0]; dbg[1] = nB; dbg[2] = nA; dbg[3] = hist[63];
    }
  }
  __syncthreads();
  for (int e0 = 0; e0 < n; e0 += 1024) {
    const int e = e0 + tid;
    const int k = e < n ? key[e] & 63 : 64;
    const unsigned m = __match_any_sync(0xffffffffu, k);
    const int leader = __ffs(m) - 1;
    int base = 0;
    if (k < 64 && lane == leader) base = atomicAdd(&off[k], __popc(m));
    base = __shfl_sync(m, base, leader);
    if (k < 64) perm[base + __popc(m & ((1u << lane) - 1))] = e;
  }
}
cudaError_t jaco_place_launch(const unsigned char* key, int n, int* perm, int* dbg, cudaStream_t st) {
  jaco_place_kernel<<<1, 1024, 0, st>>>(key, n, perm, dbg);
  return cudaGetLastError();
}


cudaError_t jaco_coop_upload_model(const JacoModel* m) {
  cudaError_t e = cudaMemcpyToSymbol(c_jaco_model, m, sizeof(JacoModel));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_jaco_model_dev, m, sizeof(JacoModel));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(jaco_step_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(CoopWS) * JACO_COOP_WSN));
  return e;
}
cudaError_t jaco_coop_launch(const JacoStepArgs& a, cudaStream_t st) {
  const size_t smem = sizeof(CoopWS) * JACO_COOP_WSN;
  jaco_step_coop_kernel<<<(a.N + JACO_FAST_BLOCK - 1) / JACO_FAST_BLOCK, JACO_FAST_BLOCK, smem, st>>>(a);
  return cudaGetLastError();
}
#ifdef JACO_PROF
extern "C" int jaco_debug_prof(unsigned long long* out16);   /* profiling builds: what scripts/prof_coop.py reads */
int jaco_debug_prof(unsigned long long* out16) { return jaco_coop_prof(out16); }
#endif
int jaco_coop_prof(unsigned long long* out16) {
#ifdef JACO_PROF
  return cudaMemcpyFromSymbol(out16, g_jprof, 16 * sizeof(unsigned long long)) == cudaSuccess ? 0 : -2;
#else
  (void)out16; return -1;
#endif
}
