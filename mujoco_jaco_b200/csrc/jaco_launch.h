// jaco_launch.h -- internal glue between the translation units of libjaco_b200.so (each .cu is compiled on its
// own so the big kernels build in parallel; every unit keeps its own copy of the model symbols).
#pragma once
#include <cuda_runtime.h>
#include <string.h>
#include "../../include/jaco_b200.h"
#include "jaco_fast.cuh"

#ifndef JACO_FAST_BLOCK
#define JACO_FAST_BLOCK 256   // the step kernel: one 8-warp block per SM at 255 registers
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
  // env placement: perm[p] = env served by placement slot p (sorted by key, busiest first; nullptr: identity);
  // key[env] = workload of the env's last step; wave = number of leading blocks that are dealt the sorted order in snake fashion
  const int* perm; unsigned char* key; int wave;
};

// ---- the production step kernel (jaco_kernel_coop.cu)
cudaError_t jaco_coop_upload_model(const JacoModel* m);
cudaError_t jaco_coop_launch(const JacoStepArgs& a, cudaStream_t st);
// perm = envs sorted by workload key, descending (one block, shared-memory counting sort); dbg (nullable): 4 ints of diagnostics
cudaError_t jaco_place_launch(const unsigned char* key, int n, int* perm, int* dbg, cudaStream_t st);
int jaco_coop_prof(unsigned long long* out16);

