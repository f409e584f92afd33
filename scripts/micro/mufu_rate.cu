// Micro-benchmark: ex2.approx throughput per SM as a function of resident warps per SM sub-partition, with the
// instruction mix of the attention softmax inner loop (per pair: 1 FFMA2, 2 MUFU.EX2, 1 FADD2, 1 F2FP).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_rate mufu_rate.cu && ./mufu_rate
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
template <int MODE>
__global__ void k(float* out, int iters, float a, float b) {
  float s[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) s[i] = threadIdx.x * 1e-3f + i;
  float acc0 = 0.f, acc1 = 0.f;
  unsigned pk = 0;
  for (int it = 0; it < iters; ++it) {
    b += 1e-6f;  // loop-variant inputs: nothing can be hoisted
#pragma unroll 1
    for (int rep = 0; rep < 1; ++rep)
#pragma unroll
    for (int i = 0; i < 64; i += 2) {
      float x0 = fmaf(s[i], a, b), x1 = fmaf(s[i + 1], a, b);
      float p0 = ex2(x0), p1 = ex2(x1);
      if (MODE == 1) {
        acc0 += p0; acc1 += p1;
        __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
        pk ^= *reinterpret_cast<unsigned*>(&t);
      } else {
        acc0 += p0 + p1;
      }
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 + acc1 + __uint_as_float(pk);
}
int main() {
  float* d; cudaMalloc(&d, 1 << 24);
  int sms = 148;
  for (int mode = 0; mode < 2; ++mode)
    for (int warps = 4; warps <= 16; warps *= 2) {
      const int iters = 2000;
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        if (mode) k<1><<<sms, warps * 32>>>(d, iters, 1e-3f, -1.f); else k<0><<<sms, warps * 32>>>(d, iters, 1e-3f, -1.f);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
      }
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
      double ex = (double)warps * 32 * iters * 64;  // per SM
      printf("mode %d warps/SM %2d (%d per SMSP): %.3f ms  %.2f ex2/ns/SM  (%.1f per clk at %.2f GHz nominal)\n", mode, warps, warps / 4, ms,
             ex / (ms * 1e6), ex / (ms * 1e6) / (clk * 1e-6), clk * 1e-6);
    }
  return 0;
}
