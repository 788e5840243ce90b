// Microbenchmark + semantics check of TMA tile::gather4 (4 arbitrary rows of a 2-D tensor per op).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_gather4 tma_gather4.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint32_t b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c)); }
__device__ __forceinline__ void mb_expect(uint32_t b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(n) : "memory"); }
__device__ __forceinline__ void mb_wait(uint32_t b, uint32_t ph) {
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(b), "r"(ph) : "memory");
}
__device__ __forceinline__ void g4(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int r0, int r1, int r2, int r3) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
               ::"r"(dst), "l"((uint64_t)m), "r"(bar), "r"(c0), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
}
// each stage = OPS gather4 ops (4 rows x box_cols each); rows pseudo-random
__global__ void bench(const __grid_constant__ CUtensorMap tm, int iters, int stages, int ops, int box_cols, int rows_total, long long* cyc, uint16_t* dump) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[16];
  const int op_bytes = 4 * box_cols * 2, stage_bytes = ops * op_bytes;
  const int NT = 4;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) mb_init(s32(&bars[s]), NT);
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  if (threadIdx.x < NT * 32 && (threadIdx.x & 31) == 0) {
    const int tid = threadIdx.x >> 5;
    uint32_t rng = 12345u + blockIdx.x * 977u + tid * 31337u;
    long long t0 = clock64();
    int is = 0, issued = 0;
    auto issue = [&]() {
      mb_expect(s32(&bars[is]), stage_bytes / NT);
      for (int o = tid; o < ops; o += NT) {
        rng = rng * 1664525u + 1013904223u;
        int r = (rng >> 8) & 0x7FFF;
        g4(s32(smem + (size_t)is * stage_bytes + o * op_bytes), &tm, s32(&bars[is]), 0, r, r + 77, r + 1031, r + 3001);
      }
      if (++is == stages) is = 0;
    };
    for (; issued < stages && issued < iters; ++issued) issue();
    int ds = 0; uint32_t ph = 0;
    for (int done = 0; done < iters; ++done) {
      mb_wait(s32(&bars[ds]), ph);
      if (issued < iters) { issue(); ++issued; }
      if (++ds == stages) { ds = 0; ph ^= 1; }
    }
    if (tid == 0) cyc[blockIdx.x] = clock64() - t0;
  }
  __syncthreads();
  if (blockIdx.x == 0 && dump) {   // semantics check: one op with known rows
    if (threadIdx.x == 0) {
      mb_init(s32(&bars[15]), 1);
      asm volatile("fence.mbarrier_init.release.cluster;");
      mb_expect(s32(&bars[15]), op_bytes);
      g4(s32(smem), &tm, s32(&bars[15]), 0, 5, 100, 7, 60000);
      mb_wait(s32(&bars[15]), 0);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 4 * box_cols; i += blockDim.x) dump[i] = ((uint16_t*)smem)[i];
  }
}
int main(int argc, char** argv) {
  int box_cols = argc > 1 ? atoi(argv[1]) : 256, ops = argc > 2 ? atoi(argv[2]) : 8, stages = argc > 3 ? atoi(argv[3]) : 4;
  int grid = argc > 4 ? atoi(argv[4]) : 148;
  const int D = 512, rows = 65536;
  uint16_t* h = (uint16_t*)malloc((size_t)rows * D * 2);
  for (int r = 0; r < rows; ++r) for (int c = 0; c < D; ++c) h[(size_t)r * D + c] = (uint16_t)(r * 7 + c);
  void* d; cudaMalloc(&d, (size_t)rows * D * 2); cudaMemcpy(d, h, (size_t)rows * D * 2, cudaMemcpyHostToDevice);
  CUtensorMap tm; cuuint64_t gd[2] = {D, rows}, gs[1] = {D * 2}; cuuint32_t bx[2] = {(cuuint32_t)box_cols, 1}, es[2] = {1, 1};
  void* fn; cudaDriverEntryPointQueryResult q; cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  typedef CUresult (*F)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUresult r = ((F)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r) { printf("encode failed %d\n", r); return 1; }
  int iters = 4000, stage_bytes = ops * 4 * box_cols * 2;
  long long* cyc; cudaMalloc(&cyc, grid * 8);
  uint16_t* dump; cudaMalloc(&dump, 4 * box_cols * 2);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, stages * stage_bytes);
  for (int it = 0; it < 2; ++it) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    bench<<<grid, 128, stages * stage_bytes>>>(tm, iters, stages, ops, box_cols, rows, cyc, dump);
    cudaEventRecord(e1); cudaError_t e = cudaDeviceSynchronize();
    if (e) { printf("err %s\n", cudaGetErrorString(e)); return 1; }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long hc; cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
    double nops = (double)iters * ops;
    printf("gather4 box_cols %d ops/stage %d stages %d grid %d: %.3f ms, %.1f clk/op, %.1f B/clk/SM, total %.2f TB/s\n", box_cols, ops, stages, grid, ms,
           hc / nops, nops * 4 * box_cols * 2 / hc, nops * 4 * box_cols * 2 * grid / ms / 1e9);
  }
  uint16_t* hd = (uint16_t*)malloc(4 * box_cols * 2); cudaMemcpy(hd, dump, 4 * box_cols * 2, cudaMemcpyDeviceToHost);
  int rr[4] = {5, 100, 7, 60000}; int bad = 0;
  for (int j = 0; j < 4; ++j) for (int c = 0; c < box_cols; ++c) if (hd[j * box_cols + c] != (uint16_t)(rr[j] * 7 + c)) ++bad;
  printf("semantics: smem[j*box_cols + c] == T[row_j][c] mismatches: %d (first words %u %u | %u)\n", bad, hd[0], hd[1], hd[box_cols]);
  return 0;
}
