// jaco_launch.h -- internal glue between the translation units of libjaco_b200.so (each .cu is compiled on its
// own so the big kernels build in parallel; every unit keeps its own copy of the model symbols).
#pragma once
#include <cuda_runtime.h>
#include <string.h>
#include "../../include/jaco_b200.h"
#include "jaco_pipe.cuh"

#ifndef JACO_FAST_BLOCK
#define JACO_FAST_BLOCK 256   // fused hybrid kernel: one 8-warp block per SM at 255 registers
#endif
#ifndef JACO_FAST_MINBLOCKS
#define JACO_FAST_MINBLOCKS 1
#endif

// episode statistics accumulate in double on the device (a float counter stops counting at 2^24)
__device__ __forceinline__ void block_stats_flush(float* local, double* gstats) {
  // warp-shuffle reduction of the per-thread counters, one atomic per warp per non-zero counter
#pragma unroll
  for (int k = 0; k < 16; k++) {
    float v = local[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v != 0.0f) atomicAdd(gstats + k, (double)v);
  }
}

struct JacoStepArgs {
  JacoCfg cfg; float* state; int N, env_offset; const float* action; float* obs; float* reward;
  unsigned char* done; unsigned char* succ; float* wb; float* term_obs; int auto_reset; double* gstats; int* worklist; int* counters;
};

// ---- the per-substep pipeline (jaco_kernel_pipe.cu; diagnostic builds only: measured 3x slower than the fused kernel)
#define JACO_PIPE_MAXRANGES 4
struct JacoPipe {
  int nranges, sms;                 // env ranges ("pipes"), each with a main and a dirty stream
  float* scratch; PipeRec* rec; int* flags /* it | pend | dirty | dead, [4][N] */; int* dlist; int* cnt; float* state_prev;
  cudaStream_t s_main[JACO_PIPE_MAXRANGES], s_dirty[JACO_PIPE_MAXRANGES];
  cudaEvent_t ev_fork, ev_sub[JACO_PIPE_MAXRANGES][50], ev_join[JACO_PIPE_MAXRANGES][2];
  long long launches;
};
cudaError_t jaco_pipe_upload_model(const JacoModel* m);
cudaError_t jaco_pipe_create(JacoPipe* P, int n_envs, int device);
void jaco_pipe_destroy(JacoPipe* P);
cudaError_t jaco_pipe_clear(JacoPipe* P, int n_envs, cudaStream_t st);
cudaError_t jaco_pipe_step(JacoPipe* P, const JacoStepArgs& a, double* gstats, cudaStream_t st);

// ---- the production fused hybrid kernel (jaco_kernel_coop.cu)
cudaError_t jaco_coop_upload_model(const JacoModel* m);
cudaError_t jaco_coop_launch(const JacoStepArgs& a, cudaStream_t st);
int jaco_coop_prof(unsigned long long* out16);
#ifdef JACO_DIAG   // bail-only fast kernel, kept for A/B timing in diagnostic builds only
cudaError_t jaco_fast_upload_model(const JacoModel* m);
cudaError_t jaco_fast_launch(const JacoStepArgs& a, cudaStream_t st);
#endif
