// jaco_kernel_coop.cu -- the production step kernel: fast path per thread + warp-cooperative general contact
// handling on demand (jaco_fast.cuh, jaco_coop.cuh), all 50 substeps of one env step in one launch.
//
// Alternatives measured and rejected this round (profiles/r02_*): a per-substep kernel pipeline with the rare
// paths on a second stream (90-114 ms per 65 536-env step against 44), and a two-pass split (bail-only fast kernel, then
// this kernel over the dirty list: 84-98 ms) -- the general contact pass is throughput-bound on instruction delivery and
// needs every SM's warps, which this single kernel gives it substep by substep.
#include "jaco_launch.h"
#define JACO_COOP_WSN ((JACO_FAST_BLOCK / 32) < JACO_COOP_NWS ? (JACO_FAST_BLOCK / 32) : JACO_COOP_NWS)

// hybrid kernel: fast path per thread + warp-cooperative general contact handling on demand (jaco_coop.cuh);
// one CoopWS per warp in dynamic shared memory
__global__ void __launch_bounds__(JACO_FAST_BLOCK, JACO_FAST_MINBLOCKS) jaco_step_coop_kernel(const JacoStepArgs A) {
  const JacoCfg& cfg = A.cfg; float* state = A.state; const int N = A.N, env_offset = A.env_offset; const float* action = A.action;
  float* obs = A.obs; float* reward = A.reward; unsigned char* done = A.done; unsigned char* succ = A.succ; float* wb = A.wb;
  float* term_obs = A.term_obs; const int auto_reset = A.auto_reset; double* gstats = A.gstats; int* worklist = A.worklist; int* counters = A.counters;

  extern __shared__ __align__(16) unsigned char coop_smem[];
  CoopWS* ws = reinterpret_cast<CoopWS*>(coop_smem) + (threadIdx.x >> 5);
  const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = i0 < N;
  const int i = active ? i0 : N - 1;
  float local[16];
#pragma unroll
  for (int k = 0; k < 16; k++) local[k] = 0;
  float a[JACO_NACT];
#pragma unroll
  for (int k = 0; k < JACO_NACT; k++) a[k] = action[(size_t)i * JACO_NACT + k];
  float r, w; unsigned char d, sc; int needs_reset, why;
  int rc = entry_step_fast<true>(cfg, state, N, i, env_offset + i, a, obs + (size_t)i * JACO_NOBS, &r, &d, &sc, &w,
                                 term_obs ? term_obs + (size_t)i * JACO_NOBS : nullptr, auto_reset, local, &needs_reset, &why, ws, active);
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
  block_stats_flush(local, gstats);
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
int jaco_coop_prof(unsigned long long* out16) {
#ifdef JACO_PROF
  return cudaMemcpyFromSymbol(out16, g_jprof, 16 * sizeof(unsigned long long)) == cudaSuccess ? 0 : -2;
#else
  (void)out16; return -1;
#endif
}
