// jaco_kernel_fast.cu -- the bail-out variant of the fast kernel (diagnostics / comparison: JACO_NO_COOP=1).
#include "jaco_launch.h"

__global__ void __launch_bounds__(JACO_FAST_BLOCK, JACO_FAST_MINBLOCKS) jaco_step_fast_kernel(const JacoStepArgs A) {
  const JacoCfg& cfg = A.cfg; float* state = A.state; const int N = A.N, env_offset = A.env_offset; const float* action = A.action;
  float* obs = A.obs; float* reward = A.reward; unsigned char* done = A.done; unsigned char* succ = A.succ; float* wb = A.wb;
  float* term_obs = A.term_obs; const int auto_reset = A.auto_reset; double* gstats = A.gstats; int* worklist = A.worklist; int* counters = A.counters;

  int i = blockIdx.x * blockDim.x + threadIdx.x;
  float local[16];
#pragma unroll
  for (int k = 0; k < 16; k++) local[k] = 0;
  if (i < N) {
    float a[JACO_NACT];
#pragma unroll
    for (int k = 0; k < JACO_NACT; k++) a[k] = action[(size_t)i * JACO_NACT + k];
    float r, w; unsigned char d, sc; int needs_reset, why;
    int rc = entry_step_fast<false>(cfg, state, N, i, env_offset + i, a, obs + (size_t)i * JACO_NOBS, &r, &d, &sc, &w,
                             term_obs ? term_obs + (size_t)i * JACO_NOBS : nullptr, auto_reset, local, &needs_reset, &why);
    if (rc == FAST_OK) {
      reward[i] = r; done[i] = d; succ[i] = sc; wb[i] = w;
      if (needs_reset) worklist[N + atomicAdd(counters + 1, 1)] = i;
    } else {
      worklist[atomicAdd(counters + 0, 1)] = i;
      int bit = 31 - __clz(why & -why);  // lowest reason bit
      atomicAdd(counters + 2 + (bit < 0 ? 0 : bit), 1);
    }
  }
  block_stats_flush(local, gstats);
}


cudaError_t jaco_fast_upload_model(const JacoModel* m) {
  cudaError_t e = cudaMemcpyToSymbol(c_jaco_model, m, sizeof(JacoModel));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_jaco_model_dev, m, sizeof(JacoModel));
  return e;
}
cudaError_t jaco_fast_launch(const JacoStepArgs& a, cudaStream_t st) {
  jaco_step_fast_kernel<<<(a.N + JACO_FAST_BLOCK - 1) / JACO_FAST_BLOCK, JACO_FAST_BLOCK, 0, st>>>(a);
  return cudaGetLastError();
}
