// Microbenchmark: how many bytes/clk can one SM ingest through TMA (2-D tiled loads from L2-resident data)?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_ingest tma_ingest.cu -lcuda
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
__device__ __forceinline__ void tma2d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"((uint64_t)m), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// one thread issues TMA loads into a ring of `stages` buffers and immediately recycles each buffer when it lands
__global__ void ingest(const __grid_constant__ CUtensorMap tm, int tiles, int stages, int tile_bytes, int box_rows, int rows_total, int ncol_blocks, int box_cols, long long* cyc, int start_stride) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[16];
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) mb_init(s32(&bars[s]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;");
    long long t0 = clock64();
    int issued = 0, done = 0;
    uint32_t phase[16] = {0};
    const int row_blocks = rows_total / box_rows;
    int is = 0, rb = (int)(((long long)blockIdx.x * start_stride) % row_blocks), cb = 0;
    auto issue = [&]() {
      mb_expect(s32(&bars[is]), tile_bytes);
      tma2d(s32(smem + (size_t)is * tile_bytes), &tm, s32(&bars[is]), cb * box_cols, rb * box_rows);
      if (++is == stages) is = 0;
      if (++cb == ncol_blocks) { cb = 0; if (++rb == row_blocks) rb = 0; }
    };
    for (; issued < stages && issued < tiles; ++issued) issue();
    int ds = 0; uint32_t ph = 0;
    for (; done < tiles; ++done) {
      mb_wait(s32(&bars[ds]), ph);
      if (issued < tiles) { issue(); ++issued; }
      if (++ds == stages) { ds = 0; ph ^= 1; }
    }
    cyc[blockIdx.x] = clock64() - t0;
  }
}
int main(int argc, char** argv) {
  int box_cols = argc > 1 ? atoi(argv[1]) : 64, box_rows = argc > 2 ? atoi(argv[2]) : 128, stages = argc > 3 ? atoi(argv[3]) : 4;
  int swz = argc > 4 ? atoi(argv[4]) : 0, grid = argc > 5 ? atoi(argv[5]) : 148;
  const int D = 512;
  const int rows = argc > 6 ? atoi(argv[6]) : 65536;   // default 64 MB bf16 matrix: L2 resident
  const int start_stride = argc > 7 ? atoi(argv[7]) : 7;   // 0: every CTA streams the same tiles in lockstep
  const int ncb = argc > 8 ? atoi(argv[8]) : D / box_cols;  // column blocks cycled (1: one D-slice only)
  void* d; cudaMalloc(&d, (size_t)rows * D * 2); cudaMemset(d, 1, (size_t)rows * D * 2);
  CUtensorMap tm; cuuint64_t gd[2] = {D, rows}, gs[1] = {D * 2}; cuuint32_t bx[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows}, es[2] = {1, 1};
  void* fn; cudaDriverEntryPointQueryResult q; cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  typedef CUresult (*F)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUresult r = ((F)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r) { printf("encode failed %d\n", r); return 1; }
  int tile_bytes = box_cols * box_rows * 2, tiles = argc > 9 ? atoi(argv[9]) : 20000;
  long long* cyc; cudaMalloc(&cyc, grid * 8);
  cudaFuncSetAttribute(ingest, cudaFuncAttributeMaxDynamicSharedMemorySize, stages * tile_bytes);
  for (int it = 0; it < 2; ++it) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    ingest<<<grid, 32, stages * tile_bytes>>>(tm, tiles, stages, tile_bytes, box_rows, rows, ncb, box_cols, cyc, start_stride);
    cudaEventRecord(e1); cudaError_t e = cudaDeviceSynchronize();
    if (e) { printf("err %s\n", cudaGetErrorString(e)); return 1; }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[1]; cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
    double tot = (double)tiles * tile_bytes * grid;
    printf("rows %d stride %d ncb %d box %dx%d stages %d swz %d grid %d: %.3f ms, %.2f TB/s total, %.1f B/clk/SM (clock64), %.1f B/clk/SM @1.9GHz\n", rows, start_stride, ncb, box_cols, box_rows, stages, swz, grid, ms,
           tot / ms / 1e9, (double)tiles * tile_bytes / h[0], tot / ms / 1e-3 / grid / 1.9e9);
  }
  return 0;
}
