// Helpers shared by the K2 streaming kernels (maxsim_bwd_v2.cu: dTxt + dImg v3, maxsim_bwd_dimg4.cu: dImg v4).
#pragma once
#include "common.cuh"

namespace cxl {

// pair weight as the 16-bit operand of the mixed-precision FMA: bf16, or fp16 when the operands are fp16
__device__ __forceinline__ uint32_t cvt_w16(float w, int f16) {
  return f16 ? static_cast<uint32_t>(__half_as_ushort(__float2half_rn(w)))
             : static_cast<uint32_t>(__bfloat16_as_ushort(__float2bfloat16_rn(w)));
}
__device__ __forceinline__ uint32_t pack_entry_lo(float w, uint32_t hi16, int f16 = 0) {
  return (hi16 << 16) | cvt_w16(w, f16);
}

// acc[0..7] += x8(v) * w(low half of e): bf16 x bf16 (SASS FHFMA.BF16) or fp16 x fp16 (FHFMA) products added into
// fp32 with one rounding
template <bool F16 = false>
__device__ __forceinline__ void fhfma8_lo(float (&a)[8], const uint4& v, uint32_t e) {
  if (F16) {
    asm("{\n\t.reg .b16 x0, x1, x2, x3, x4, x5, x6, x7, wl, wh;\n\t"
        "mov.b32 {x0, x1}, %8;\n\t mov.b32 {x2, x3}, %9;\n\t mov.b32 {x4, x5}, %10;\n\t mov.b32 {x6, x7}, %11;\n\t"
        "mov.b32 {wl, wh}, %12;\n\t"
        "fma.rn.f32.f16 %0, x0, wl, %0;\n\t fma.rn.f32.f16 %1, x1, wl, %1;\n\t"
        "fma.rn.f32.f16 %2, x2, wl, %2;\n\t fma.rn.f32.f16 %3, x3, wl, %3;\n\t"
        "fma.rn.f32.f16 %4, x4, wl, %4;\n\t fma.rn.f32.f16 %5, x5, wl, %5;\n\t"
        "fma.rn.f32.f16 %6, x6, wl, %6;\n\t fma.rn.f32.f16 %7, x7, wl, %7;\n\t}"
        : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7])
        : "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(e));
  } else {
    asm("{\n\t.reg .b16 x0, x1, x2, x3, x4, x5, x6, x7, wl, wh;\n\t"
        "mov.b32 {x0, x1}, %8;\n\t mov.b32 {x2, x3}, %9;\n\t mov.b32 {x4, x5}, %10;\n\t mov.b32 {x6, x7}, %11;\n\t"
        "mov.b32 {wl, wh}, %12;\n\t"
        "fma.rn.f32.bf16 %0, x0, wl, %0;\n\t fma.rn.f32.bf16 %1, x1, wl, %1;\n\t"
        "fma.rn.f32.bf16 %2, x2, wl, %2;\n\t fma.rn.f32.bf16 %3, x3, wl, %3;\n\t"
        "fma.rn.f32.bf16 %4, x4, wl, %4;\n\t fma.rn.f32.bf16 %5, x5, wl, %5;\n\t"
        "fma.rn.f32.bf16 %6, x6, wl, %6;\n\t fma.rn.f32.bf16 %7, x7, wl, %7;\n\t}"
        : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7])
        : "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(e));
  }
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// host helpers defined in maxsim_bwd_v2.cu
// [rows, D] bf16 row-major, box = box_cols x box_rows, no swizzle (rows land densely, box_cols*2 bytes apart)
int make_tmap_plain(CUtensorMap* out, const void* base, long long rows, int D, int box_cols, int box_rows, int f16 = 0);
int round_up_i(int x, int m);
int env_int(const char* name, int dflt);

// ---- dImg v4 (maxsim_bwd_dimg4.cu) ----
struct Dimg4Geom {
  int nw, npo, nimg;       // consumer warps, patches per quarter warp, images per CTA
  int ring, nbuf;          // sub-tile slots in the ring (nimg + depth), step-table buffers (depth + 1)
  int sub_rows, sub_bytes; // text rows per sub-tile (multiple of 16), sub_rows * 128
  int box_rows, n_box;     // TMA boxes per sub-tile
  int H;                   // sub-tiles = ceil(rows / sub_rows)
  int tau_max;             // last walk = H + nimg - 2
  int walks;               // walks per image: walk m of the image at position c of its CTA is tau = m * nimg + c
  int hs;                  // u16 header slots per consumer warp: npo step counts, then the first step of the warp
  int hdr_bytes;           // header bytes of a block (multiple of 16)
  int blk_stride;          // bytes reserved per block (worst case: every row of a walk in one quarter's list)
};
bool dimg4_covers(int P);
Dimg4Geom dimg4_geom(long long rows, int P, int max_smem);
long long dimg4_bytes_per_image(const Dimg4Geom& g);
// f16: fp16 operands; w_scale: device [s, 1/s] (power of two applied to the 16-bit pair weights, undone at the store) or null
int launch_dimg_v4(const void* txt, const void* argmax, long long ld_arg, const float* w, int Bt, int Bi, int T, int P,
                   int D, int ld_pairs, int k_off, float* dimg, int accumulate_img, const Dimg4Geom& g, void* scratch,
                   long long cap, int max_smem, int f16, const float* w_scale, cudaStream_t stream);

}  // namespace cxl
