"""Instruction-fetch micro-benchmark 2: how tightly must warps be aligned to share one instruction stream?
8 warps per SM run the same 16384-instruction (256 KB) straight-line body in a loop; warp w first executes a private
prologue of w * SKEW extra instructions, so the warps stay SKEW instructions apart.  A second variant re-aligns the warps
with __syncthreads() every SEG instructions."""
N = 16384
print("#include <cstdio>\n#include <cuda_runtime.h>")
print("template <int SYNC> __global__ void body(float* out, int iters, float x, float y, int skew) {")
print("  float a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, a4 = 4, a5 = 5, a6 = 6, a7 = 7;")
print("  const int w = threadIdx.x >> 5;")
print("  for (int k = 0; k < w * skew; k++) { a0 = fmaf(a0, x, 0.5f); a1 = fmaf(a1, y, 0.25f); a2 = fmaf(a2, x, 0.75f); a3 = fmaf(a3, y, 0.125f); a4 = fmaf(a4, x, 0.375f); a5 = fmaf(a5, y, 0.625f); a6 = fmaf(a6, x, 0.875f); a7 = fmaf(a7, y, 0.0625f); }")
print("  for (int it = 0; it < iters; it++) {")
for i in range(N // 8):
    if i % 64 == 0:
        print("    if (SYNC) __syncthreads();")
    print(f"    a0 = fmaf(a0, x, {i}.25f); a1 = fmaf(a1, y, {i}.5f); a2 = fmaf(a2, x, {i}.75f); a3 = fmaf(a3, y, {i}.125f); a4 = fmaf(a4, x, {i}.375f); a5 = fmaf(a5, y, {i}.625f); a6 = fmaf(a6, x, {i}.875f); a7 = fmaf(a7, y, {i}.0625f);")
print("  }")
print("  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;\n}")
print("""template <int SYNC> void run(float* out, int sms, int skew8) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 100;
  body<SYNC><<<sms, 256>>>(out, 2, 1.0001f, 0.9999f, skew8); cudaDeviceSynchronize();
  cudaEventRecord(e0); body<SYNC><<<sms, 256>>>(out, iters, 1.0001f, 0.9999f, skew8); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double winst = (double)%d * iters * 8;
  printf("sync-every-512 %%d  skew %%5d instr/warp  ms %%8.3f  IPC/SM %%.3f\\n", SYNC, skew8 * 8, ms, winst / (ms * 1e-3 * 1.965e9));
}
int main() {
  float* out; cudaMalloc(&out, 148 * 256 * 4);
  int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int skews[] = {0, 1, 4, 16, 64, 256, 1024};
  for (int s : skews) run<0>(out, sms, s);
  for (int s : skews) run<1>(out, sms, s);
  return 0;
}""" % N)
