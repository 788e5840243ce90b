// Microbenchmark: issue rate of FFMA vs FHFMA.BF16 (fma.rn.f32.bf16) per SM, independent accumulators.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fma_rate fma_rate.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
template <int MODE>
__global__ void k(float* out, uint32_t x, uint32_t w, int iters, long long* cyc) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
  uint32_t xv = x + threadIdx.x, wv = w;
  float xf = __uint_as_float(xv), wf = __uint_as_float(wv);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) {
        asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(xf), "f"(wf));
      } else {
        asm volatile("{\n\t.reg .b16 xl, xh, wl, wh;\n\t mov.b32 {xl, xh}, %1;\n\t mov.b32 {wl, wh}, %2;\n\t"
                     "fma.rn.f32.bf16 %0, xl, wl, %0;\n\t}" : "+f"(a[i]) : "r"(xv), "r"(wv));
      }
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  const int iters = 20000;
  for (int warps = 4; warps <= 32; warps *= 2) {
    for (int mode = 0; mode < 2; ++mode) {
      if (mode == 0) k<0><<<148, warps * 32>>>(out, 0x3f803f80u, 0x3f803f80u, iters, cyc);
      else k<1><<<148, warps * 32>>>(out, 0x3f803f80u, 0x3f803f80u, iters, cyc);
      cudaDeviceSynchronize();
      long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      double inst = (double)iters * 16 * warps;
      printf("%s warps/SM %2d: %.3f warp-FMA/clk/SM (%.1f lanes/clk/SM)\n", mode ? "FHFMA.BF16" : "FFMA      ", warps, inst / h, inst / h * 32);
    }
  }
  return 0;
}
