// jaco_kernel_pipe.cu -- the production env step: kernels and launch schedule of the per-substep pipeline (jaco_pipe.cuh).
//
// Per env range ("pipe") two CUDA streams:
//   main   K_S(0) .. K_S(49)                                  one thread per env of the range
//   dirty  for j = 0..49: wait K_S(j); K_B(j); K_S'(j)   then K_B(50)      persistent grids over the dirty list
// and on the caller's stream, after both joined: K_E over all envs, then the general-path list kernel (jaco_kernels.cu)
// for the few envs the pipeline handed back (NaN guard) and for resets that need the general mj_forward.
#include "jaco_launch.h"

#ifndef PIPE_SBLOCK
#define PIPE_SBLOCK 128      // K_S: threads per block (one env per thread)
#endif
#ifndef PIPE_SMINB
#define PIPE_SMINB 2
#endif
#define PIPE_BWARPS 4        // K_B: warps (= concurrent tasks) per block, one CoopWS each in dynamic shared memory
#define PIPE_DBLOCK 64       // K_S': threads per block

// K_S: substep k of every clean env of [e0, e0 + ne).  cnt[0] = length of the dirty list, cnt[1] = blocks finished (the
// last block to finish publishes snap[k] = dirty count as of "K_S(k) complete": the dirty stream's loop bound).
__global__ void __launch_bounds__(PIPE_SBLOCK, PIPE_SMINB) jaco_pipe_substep_kernel(const JacoCfg cfg, const PipeBufs B, int e0, int ne, const float* __restrict__ action,
                                                                                   double* gstats, int* dlist, int* cnt, int* snap, int k) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < ne) {
    const int i = e0 + t;
    if (!B.dirty[i]) {
      if (pipe_env_substep(cfg, B, i, action + (size_t)i * JACO_NACT, false, gstats)) {
        B.dirty[i] = 1;
        dlist[atomicAdd(cnt, 1)] = i;
      }
    }
  }
  __shared__ int last;
  __syncthreads();
  if (threadIdx.x == 0) { __threadfence(); last = (atomicAdd(cnt + 1, 1) == (int)gridDim.x - 1); }
  __syncthreads();
  if (last && threadIdx.x == 0) { __threadfence(); snap[k] = *(volatile int*)cnt; cnt[1] = 0; }
}

// K_S': next substep of every dirty env that has no pending task (thread per dirty env, grid-stride over the list)
__global__ void __launch_bounds__(PIPE_DBLOCK) jaco_pipe_dirty_substep_kernel(const JacoCfg cfg, const PipeBufs B, const float* __restrict__ action, double* gstats,
                                                                               const int* dlist, const int* snap, int k) {
  const int n = snap[k];
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const int i = dlist[t];
    if (B.pend[i] == PEND_NONE && B.it[i] < 50) pipe_env_substep(cfg, B, i, action + (size_t)i * JACO_NACT, true, gstats);
  }
}

// K_B: one warp per pending task (general contact pass or deferred pinv), grid-stride over the dirty list
__global__ void __launch_bounds__(PIPE_BWARPS * 32) jaco_pipe_task_kernel(const JacoCfg cfg, const PipeBufs B, double* gstats, const int* dlist, const int* snap, int k) {
  extern __shared__ __align__(16) unsigned char pipe_smem[];
  CoopWS& ws = reinterpret_cast<CoopWS*>(pipe_smem)[threadIdx.x >> 5];
  const int n = snap[k];
  const int nw = gridDim.x * PIPE_BWARPS;
  for (int t = blockIdx.x * PIPE_BWARPS + (threadIdx.x >> 5); t < n; t += nw) {
    const int i = dlist[t];
    pipe_task(cfg, B, i, ws, gstats);
    __syncwarp();
  }
}

// K_E: end of the env step for every env; envs handed to the general path go onto the work lists of jaco_kernels.cu
__global__ void __launch_bounds__(128) jaco_pipe_end_kernel(const JacoCfg cfg, const PipeBufs B, int env_offset, float* obs, float* reward, unsigned char* done,
                                                            unsigned char* succ, float* wb, float* term_obs, int auto_reset, double* gstats, int* worklist, int* counters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float local[16];
#pragma unroll
  for (int k = 0; k < 16; k++) local[k] = 0;
  if (i < B.N) {
    const int rc = pipe_env_end(cfg, B, i, env_offset + i, obs + (size_t)i * JACO_NOBS, reward + i, done + i, succ + i, wb + i,
                                term_obs ? term_obs + (size_t)i * JACO_NOBS : nullptr, auto_reset, local);
    if (rc == 1) worklist[atomicAdd(counters + 0, 1)] = i;
    else if (rc == 2) worklist[B.N + atomicAdd(counters + 1, 1)] = i;
  }
  block_stats_flush(local, gstats);
}

cudaError_t jaco_pipe_upload_model(const JacoModel* m) {
  cudaError_t e = cudaMemcpyToSymbol(c_jaco_model, m, sizeof(JacoModel));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_jaco_model_dev, m, sizeof(JacoModel));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(jaco_pipe_task_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(CoopWS) * PIPE_BWARPS));
  return e;
}

cudaError_t jaco_pipe_create(JacoPipe* P, int n_envs, int device) {
  memset(P, 0, sizeof *P);
  int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  P->sms = sms;
  P->nranges = n_envs >= 16384 ? 2 : 1;
  cudaError_t e;
  const size_t n = (size_t)n_envs;
  if ((e = cudaMalloc(&P->scratch, sizeof(float) * PIPE_NSCR * n)) != cudaSuccess) return e;
  if ((e = cudaMalloc(&P->rec, sizeof(PipeRec) * n)) != cudaSuccess) return e;
  if ((e = cudaMalloc(&P->flags, sizeof(int) * 4 * n)) != cudaSuccess) return e;
  if ((e = cudaMalloc(&P->dlist, sizeof(int) * n)) != cudaSuccess) return e;
  if ((e = cudaMalloc(&P->cnt, sizeof(int) * JACO_PIPE_MAXRANGES * 64)) != cudaSuccess) return e;
  if ((e = cudaMalloc(&P->state_prev, sizeof(float) * JACO_NSTATE * n)) != cudaSuccess) return e;
  cudaMemset(P->scratch, 0, sizeof(float) * PIPE_NSCR * n);
  cudaMemset(P->cnt, 0, sizeof(int) * JACO_PIPE_MAXRANGES * 64);
  int lo = 0, hi = 0; cudaDeviceGetStreamPriorityRange(&lo, &hi);   // hi = numerically lowest = highest priority
  for (int p = 0; p < P->nranges; p++) {
    if ((e = cudaStreamCreateWithPriority(&P->s_main[p], cudaStreamNonBlocking, lo)) != cudaSuccess) return e;
    if ((e = cudaStreamCreateWithPriority(&P->s_dirty[p], cudaStreamNonBlocking, hi)) != cudaSuccess) return e;
    for (int k = 0; k < 50; k++) if ((e = cudaEventCreateWithFlags(&P->ev_sub[p][k], cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&P->ev_join[p][0], cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&P->ev_join[p][1], cudaEventDisableTiming)) != cudaSuccess) return e;
  }
  if ((e = cudaEventCreateWithFlags(&P->ev_fork, cudaEventDisableTiming)) != cudaSuccess) return e;
  return jaco_pipe_clear(P, n_envs, 0);
}
void jaco_pipe_destroy(JacoPipe* P) {
  for (int p = 0; p < P->nranges; p++) {
    if (P->s_main[p]) cudaStreamDestroy(P->s_main[p]);
    if (P->s_dirty[p]) cudaStreamDestroy(P->s_dirty[p]);
    for (int k = 0; k < 50; k++) if (P->ev_sub[p][k]) cudaEventDestroy(P->ev_sub[p][k]);
    for (int k = 0; k < 2; k++) if (P->ev_join[p][k]) cudaEventDestroy(P->ev_join[p][k]);
  }
  if (P->ev_fork) cudaEventDestroy(P->ev_fork);
  cudaFree(P->scratch); cudaFree(P->rec); cudaFree(P->flags); cudaFree(P->dlist); cudaFree(P->cnt); cudaFree(P->state_prev);
}
// it = -1 (step not begun), pend = dirty = dead = 0 for every env
cudaError_t jaco_pipe_clear(JacoPipe* P, int n_envs, cudaStream_t st) {
  const size_t n = (size_t)n_envs;
  cudaError_t e = cudaMemsetAsync(P->flags, 0xFF, sizeof(int) * n, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(P->flags + n, 0, sizeof(int) * 3 * n, st);
  return e;
}

cudaError_t jaco_pipe_step(JacoPipe* P, const JacoStepArgs& a, double* gstats, cudaStream_t st) {
  const int N = a.N;
  const size_t n = (size_t)N;
  PipeBufs B = {a.state, P->scratch, P->rec, P->flags, P->flags + n, P->flags + 2 * n, P->flags + 3 * n, N};
  cudaError_t e;
  // pre-step snapshot: what the general path restarts from for an env the pipeline hands back
  if ((e = cudaMemcpyAsync(P->state_prev, a.state, sizeof(float) * JACO_NSTATE * n, cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(P->cnt, 0, sizeof(int) * JACO_PIPE_MAXRANGES * 64, st)) != cudaSuccess) return e;
  if ((e = cudaEventRecord(P->ev_fork, st)) != cudaSuccess) return e;
  const int R = P->nranges;
  int e0[JACO_PIPE_MAXRANGES], ne[JACO_PIPE_MAXRANGES];
  for (int p = 0; p < R; p++) {
    e0[p] = (int)((long long)N * p / R); ne[p] = (int)((long long)N * (p + 1) / R) - e0[p];
    cudaStreamWaitEvent(P->s_main[p], P->ev_fork, 0); cudaStreamWaitEvent(P->s_dirty[p], P->ev_fork, 0);
  }
  const size_t smem = sizeof(CoopWS) * PIPE_BWARPS;
  const int bgrid = P->sms * 2, dgrid = P->sms;
  for (int k = 0; k < 50; k++) {
    for (int p = 0; p < R; p++) {
      int* cnt = P->cnt + 64 * p; int* snap = cnt + 2; int* dl = P->dlist + e0[p];
      jaco_pipe_substep_kernel<<<(ne[p] + PIPE_SBLOCK - 1) / PIPE_SBLOCK, PIPE_SBLOCK, 0, P->s_main[p]>>>(a.cfg, B, e0[p], ne[p], a.action, gstats, dl, cnt, snap, k);
      cudaEventRecord(P->ev_sub[p][k], P->s_main[p]);
    }
    for (int p = 0; p < R; p++) {
      int* cnt = P->cnt + 64 * p; int* snap = cnt + 2; int* dl = P->dlist + e0[p];
      cudaStreamWaitEvent(P->s_dirty[p], P->ev_sub[p][k], 0);
      jaco_pipe_task_kernel<<<bgrid, PIPE_BWARPS * 32, smem, P->s_dirty[p]>>>(a.cfg, B, gstats, dl, snap, k);
      jaco_pipe_dirty_substep_kernel<<<dgrid, PIPE_DBLOCK, 0, P->s_dirty[p]>>>(a.cfg, B, a.action, gstats, dl, snap, k);
    }
  }
  for (int p = 0; p < R; p++) {
    int* cnt = P->cnt + 64 * p; int* snap = cnt + 2; int* dl = P->dlist + e0[p];
    jaco_pipe_task_kernel<<<bgrid, PIPE_BWARPS * 32, smem, P->s_dirty[p]>>>(a.cfg, B, gstats, dl, snap, 49);   // tasks published by K_S'(49)
    cudaEventRecord(P->ev_join[p][0], P->s_main[p]); cudaEventRecord(P->ev_join[p][1], P->s_dirty[p]);
    cudaStreamWaitEvent(st, P->ev_join[p][0], 0); cudaStreamWaitEvent(st, P->ev_join[p][1], 0);
  }
  jaco_pipe_end_kernel<<<(N + 127) / 128, 128, 0, st>>>(a.cfg, B, a.env_offset, a.obs, a.reward, a.done, a.succ, a.wb, a.term_obs, a.auto_reset, gstats, a.worklist, a.counters);
  P->launches += R * (50 + 50 * 2 + 1) + 1;
  return cudaGetLastError();
}
