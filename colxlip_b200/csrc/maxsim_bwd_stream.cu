// K2 (fast path): MaxSim backward from the saved argmax, bf16 operands, fp32 register accumulators.
//
// Replaces autograd of loss.py:34-44 in the reference (max -> scatter into zeros, then two bmm
// against a one-hot dS):
//   w[m,k]      = dS[m,k] * g / (cnt[m,k] + 1e-8)          (masked mean)   or  dS[m,k] * g / T
//   dTxt[m,n,:] = sum_k w[m,k] * img[k, q*(m,k,n), :]
//   dImg[k,q,:] = sum_{m,n : q*(m,k,n) = q} w[m,k] * txt[m,n,:]
// Both are sparse products with exactly Bt*Bi*T non-zero "items" of D fused multiply-adds each.
// There is no operand reuse (every FMA needs its own operand element), so the kernels are built
// around data movement: accumulators never leave registers, the other operand is staged ONCE per
// CTA in shared memory by TMA and read with conflict-free vector loads, the per-item argmax /
// weight pairs are prepared by a producer warp, and the FMAs are packed (FFMA2).
//
//   dTxt kernel: CTA = (tile of text-token rows, D-slice).  Every row has exactly one item per
//                image, so the loop is regular: for each image k the [P, DS] slab of img[k] is
//                staged and warp w gathers row q*(r,k) for each of its R rows.
//   dImg kernel: CTA = (image k, D-slice).  Warp g owns patches [g*G, g*G+G) (accumulators
//                acc[G][.] in registers); text-row tiles stream through shared memory and each
//                warp picks the rows whose argmax falls into its patch group (ballot scan).
#include "common.cuh"

namespace cxl {

struct BwdStreamParams {
  const float* w;       // [Bt, ld_pairs]  per-pair weight dL/dS / cnt  (cxl_pair_weights)
  const void* argmax;   // [k_off + Bi][ld_arg]  image-major
  long long ld_arg;
  int Bt, Bi, T, P, D, p_stride;
  int rows;             // Bt*T
  int ld_pairs, k_off, argmax_u16;
  float* dtxt;          // [rows, D]
  float* dimg;          // [Bi*P, D]
  int accumulate_img;
  int stages;
  int stage_bytes;      // bytes of one ring stage (operand tile + entries)
  int tile_bytes;       // bytes of the operand tile inside a stage
};

struct __align__(8) Entry {   // one per (row, image): weight and argmax-derived offset / patch id
  float w;
  uint32_t q;
};

__device__ __forceinline__ uint32_t load_q(const BwdStreamParams& p, size_t off) {
  return p.argmax_u16 ? static_cast<uint32_t>(__ldg(reinterpret_cast<const uint16_t*>(p.argmax) + off))
                      : static_cast<uint32_t>(__ldg(reinterpret_cast<const uint8_t*>(p.argmax) + off));
}

// acc(2 x f32) += x(2 x f32) * w, one FFMA2
__device__ __forceinline__ void ffma2(float& a0, float& a1, float x0, float x1, float w) {
  asm("{\n\t.reg .b64 a, b, c;\n\t"
      "mov.b64 a, {%2, %3};\n\t"
      "mov.b64 b, {%4, %4};\n\t"
      "mov.b64 c, {%0, %1};\n\t"
      "fma.rn.f32x2 c, a, b, c;\n\t"
      "mov.b64 {%0, %1}, c;\n\t}"
      : "+f"(a0), "+f"(a1)
      : "f"(x0), "f"(x1), "f"(w));
}

template <int VEC> struct PackedLoad;   // VEC bf16 values per lane from shared memory -> VEC/2 words
template <> struct PackedLoad<2> {
  __device__ static void ld(uint32_t addr, uint32_t (&v)[1]) { asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v[0]) : "r"(addr)); }
};
template <> struct PackedLoad<4> {
  __device__ static void ld(uint32_t addr, uint32_t (&v)[2]) {
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v[0]), "=r"(v[1]) : "r"(addr));
  }
};
template <> struct PackedLoad<8> {
  __device__ static void ld(uint32_t addr, uint32_t (&v)[4]) {
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(addr));
  }
};

template <int VEC>
__device__ __forceinline__ void fma_row(float (&acc)[VEC], const uint32_t (&v)[VEC / 2], float w) {
#pragma unroll
  for (int i = 0; i < VEC / 2; ++i)
    ffma2(acc[2 * i], acc[2 * i + 1], __uint_as_float(v[i] << 16), __uint_as_float(v[i] & 0xffff0000u), w);
}

struct __align__(8) StreamBarriers {
  uint64_t full[8];
  uint64_t empty[8];
};

// ------------------------------------------------------------------------------------------------
// dTxt
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void sts_v2(uint32_t addr, float a, uint32_t b) {
  asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "r"(b) : "memory");
}

// Warp layout: warps 0..NWC-1 consume (R rows each, lanes span the DS-wide slice), warp NWC issues TMA.
// Each consumer warp prepares the (weight, slab offset) entries of ITS OWN rows (lane <-> row, loads
// issued two images ahead) and exchanges them through a private double-buffered smem strip, so the only
// cross-warp synchronisation is the slab ring.
template <int DS, int R, int NWC>
__global__ void __launch_bounds__((NWC + 1) * 32, 1)
maxsim_bwd_dtxt_stream_kernel(const __grid_constant__ CUtensorMap tmImg, const BwdStreamParams p) {
  constexpr int VEC = DS / 32;
  constexpr int ROWS = NWC * R;
  static_assert(R == 32 || R == 16, "lane <-> row entry production");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  StreamBarriers* bars = reinterpret_cast<StreamBarriers*>(smem + p.stages * p.stage_bytes);
  uint8_t* ent_area = reinterpret_cast<uint8_t*>(bars + 1);          // [NWC][2][R] entries

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int r0 = blockIdx.x * ROWS;
  const int d0 = blockIdx.y * DS;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(smem_u32(&bars->full[s]), 1);
      mbar_init(smem_u32(&bars->empty[s]), NWC);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmImg);
  }
  __syncthreads();

  if (warp == NWC) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int k = 0; k < p.Bi; ++k) {
        mbar_wait(smem_u32(&bars->empty[stage]), phase ^ 1);
        const uint32_t st = smem_u32(smem + stage * p.stage_bytes);
        const uint32_t full = smem_u32(&bars->full[stage]);
        mbar_arrive_expect_tx(full, p.tile_bytes);
        for (int q0 = 0; q0 < p.P; q0 += 256) tma_load_2d(st + q0 * DS * 2, &tmImg, full, d0, k * p.p_stride + q0);
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    float acc[R][VEC];
#pragma unroll
    for (int j = 0; j < R; ++j)
#pragma unroll
      for (int i = 0; i < VEC; ++i) acc[j][i] = 0.f;

    // entry production state (lanes >= R idle when R == 16)
    const int my_row = min(r0 + warp * R + (lane % R), p.rows - 1);       // clamped: tail rows are never stored
    const uint8_t* qptr8 = reinterpret_cast<const uint8_t*>(p.argmax) + static_cast<size_t>(p.k_off) * p.ld_arg + my_row;
    const uint16_t* qptr16 = reinterpret_cast<const uint16_t*>(p.argmax) + static_cast<size_t>(p.k_off) * p.ld_arg + my_row;
    const float* wptr = p.w + static_cast<size_t>(my_row / p.T) * p.ld_pairs + p.k_off;
    const bool u16 = p.argmax_u16 != 0;
    auto ldq = [&](int k) -> uint32_t {
      const size_t o = static_cast<size_t>(k) * p.ld_arg;
      return u16 ? static_cast<uint32_t>(__ldg(qptr16 + o)) : static_cast<uint32_t>(__ldg(qptr8 + o));
    };
    uint32_t qn[2];
    float wn[2];
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int kk = min(b, p.Bi - 1);
      qn[b] = ldq(kk);
      wn[b] = __ldg(wptr + kk);
    }
    const uint32_t ent_base = smem_u32(ent_area) + warp * (2 * R * 8);
    const uint32_t lane_off = lane * VEC * 2;
    int stage = 0;
    uint32_t phase = 0;
    for (int k0 = 0; k0 < p.Bi; k0 += 2) {
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int k = k0 + b;
        if (k < p.Bi) {
          const uint32_t ebuf = ent_base + b * (R * 8);
          if (R == 32 || lane < R) sts_v2(ebuf + (lane % R) * 8, wn[b], qn[b] * (DS * 2));
          __syncwarp();
          {   // prefetch the entry of image k + 2 (clamped; an unused extra load at the very end is harmless)
            const int kk = min(k + 2, p.Bi - 1);
            qn[b] = ldq(kk);
            wn[b] = __ldg(wptr + kk);
          }
          mbar_wait(smem_u32(&bars->full[stage]), phase);
          const uint32_t stl = smem_u32(smem + stage * p.stage_bytes) + lane_off;
#pragma unroll
          for (int j = 0; j < R; j += 2) {
            float w0, w1;
            uint32_t o0, o1;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=f"(w0), "=r"(o0), "=f"(w1), "=r"(o1) : "r"(ebuf + j * 8));
            uint32_t v0[VEC / 2], v1[VEC / 2];
            PackedLoad<VEC>::ld(stl + o0, v0);
            PackedLoad<VEC>::ld(stl + o1, v1);
            fma_row<VEC>(acc[j], v0, w0);
            fma_row<VEC>(acc[j + 1], v1, w1);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&bars->empty[stage]));
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const int r = r0 + warp * R + j;
      if (r < p.rows) {
        float* o = p.dtxt + static_cast<size_t>(r) * p.D + d0 + lane * VEC;
        if constexpr (VEC == 2) {
          *reinterpret_cast<float2*>(o) = make_float2(acc[j][0], acc[j][1]);
        } else {
#pragma unroll
          for (int i = 0; i < VEC; i += 4)
            *reinterpret_cast<float4*>(o + i) = make_float4(acc[j][i], acc[j][i + 1], acc[j][i + 2], acc[j][i + 3]);
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// dImg
// ------------------------------------------------------------------------------------------------
template <int VEC, int G>
__device__ __forceinline__ void fma_select(float (&acc)[G][VEC], const uint32_t (&v)[VEC / 2], float w, int ql) {
  // warp-uniform switch: the item (and therefore ql) is the same for all lanes
  switch (ql) {
#define CXL_CASE(i) \
  case i:           \
    if constexpr (i < G) fma_row<VEC>(acc[i < G ? i : 0], v, w); \
    break;
    CXL_CASE(0) CXL_CASE(1) CXL_CASE(2) CXL_CASE(3) CXL_CASE(4) CXL_CASE(5) CXL_CASE(6) CXL_CASE(7)
    CXL_CASE(8) CXL_CASE(9) CXL_CASE(10) CXL_CASE(11) CXL_CASE(12) CXL_CASE(13) CXL_CASE(14) CXL_CASE(15)
    CXL_CASE(16) CXL_CASE(17) CXL_CASE(18) CXL_CASE(19) CXL_CASE(20) CXL_CASE(21) CXL_CASE(22) CXL_CASE(23)
#undef CXL_CASE
    default: break;
  }
}

// Warp layout: warps 0..nwc-1 consume (warp g owns patches [g*G, g*G+G)), warps nwc..nwc+RT/32-1 produce:
// producer warp pw prepares the (weight, patch) entry of row pw*32 + lane of every tile (argmax bytes are
// prefetched two tiles ahead; the weights of 32 consecutive texts sit in one register per lane, fetched a
// whole batch ahead) and producer warp 0 also issues the TMA load of the text-row tile.
// The host guarantees T >= 8: a tile's RT = 128 rows span at most 17 texts, i.e. at most two batches.
template <int DS, int G, int RT, int MAXW>
__global__ void __launch_bounds__(MAXW * 32, 1)
maxsim_bwd_dimg_stream_kernel(const __grid_constant__ CUtensorMap tmTxt, const BwdStreamParams p, int nwc) {
  constexpr int VEC = DS / 32;
  constexpr int NPW = RT / 32;
  static_assert(G <= 24, "patch group too large");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  StreamBarriers* bars = reinterpret_cast<StreamBarriers*>(smem + p.stages * p.stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int k = blockIdx.x;
  const int d0 = blockIdx.y * DS;
  const int num_tiles = (p.rows + RT - 1) / RT;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(smem_u32(&bars->full[s]), NPW + 1);     // one arrive per producer warp + the expect_tx arrive
      mbar_init(smem_u32(&bars->empty[s]), nwc);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmTxt);
  }
  __syncthreads();

  if (warp >= nwc) {
    const int pw = warp - nwc;
    const int slot = pw * 32 + lane;                     // row of the tile this lane prepares
    const bool u16 = p.argmax_u16 != 0;
    const uint8_t* q8 = reinterpret_cast<const uint8_t*>(p.argmax) + static_cast<size_t>(p.k_off + k) * p.ld_arg;
    const uint16_t* q16 = reinterpret_cast<const uint16_t*>(p.argmax) + static_cast<size_t>(p.k_off + k) * p.ld_arg;
    auto ldq = [&](int r) -> uint32_t {
      if (r >= p.rows) return 0xFFFFu;                   // owned by nobody
      return u16 ? static_cast<uint32_t>(__ldg(q16 + r)) : static_cast<uint32_t>(__ldg(q8 + r));
    };
    const float* wcol = p.w + p.k_off + k;
    auto load_w = [&](int m) { return m < p.Bt ? __ldg(wcol + static_cast<size_t>(m) * p.ld_pairs) : 0.f; };
    int m_base = 0;
    float w_cur = load_w(lane), w_nxt = load_w(32 + lane);
    int m_row = slot / p.T;                              // text of this lane's row in the current tile
    int n_row = slot - m_row * p.T;
    int m_first = 0, n_first = 0;                        // text / token of the tile's first row (warp-uniform)
    uint32_t qn[2] = {ldq(slot), ldq(RT + slot)};
    int stage = 0;
    uint32_t phase = 0;
    for (int t0 = 0; t0 < num_tiles; t0 += 2) {
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int t = t0 + b;
        if (t < num_tiles) {
          while (m_first >= m_base + 32) {               // keep the tile's first text inside w_cur
            m_base += 32;
            w_cur = w_nxt;
            w_nxt = load_w(m_base + 32 + lane);
          }
          const int mi = min(m_row, p.Bt - 1) - m_base;  // 0 .. 63
          const float wa = __shfl_sync(0xffffffffu, w_cur, mi & 31);
          const float wb = __shfl_sync(0xffffffffu, w_nxt, mi & 31);
          const float wv = mi < 32 ? wa : wb;
          mbar_wait(smem_u32(&bars->empty[stage]), phase ^ 1);
          const uint32_t st = smem_u32(smem + stage * p.stage_bytes);
          const uint32_t full = smem_u32(&bars->full[stage]);
          if (pw == 0 && lane == 0) {
            mbar_arrive_expect_tx(full, p.tile_bytes);
            tma_load_2d(st, &tmTxt, full, d0, t * RT);   // rows past the end are zero-filled
          }
          sts_v2(st + p.tile_bytes + slot * 8, wv, qn[b]);
          __syncwarp();
          if (lane == 0) mbar_arrive(full);
          qn[b] = ldq((t + 2) * RT + slot);
          n_row += RT;
          while (n_row >= p.T) { n_row -= p.T; ++m_row; }
          n_first += RT;
          while (n_first >= p.T) { n_first -= p.T; ++m_first; }
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ============ consumer warp `warp` owns patches [warp*G, warp*G + G) ============
    float acc[G][VEC];
#pragma unroll
    for (int j = 0; j < G; ++j)
#pragma unroll
      for (int i = 0; i < VEC; ++i) acc[j][i] = 0.f;
    const uint32_t q_lo = warp * G;
    const uint32_t lane_off = lane * VEC * 2;
    int stage = 0;
    uint32_t phase = 0;
    for (int t = 0; t < num_tiles; ++t) {
      mbar_wait(smem_u32(&bars->full[stage]), phase);
      const uint32_t st = smem_u32(smem + stage * p.stage_bytes);
      const uint32_t ent = st + p.tile_bytes + lane * 8;
      const uint32_t stl = st + lane_off;
#pragma unroll
      for (int c = 0; c < RT / 32; ++c) {
        float w;
        uint32_t q;
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=f"(w), "=r"(q) : "r"(ent + c * 256));
        const uint32_t ql = q - q_lo;
        uint32_t mask = __ballot_sync(0xffffffffu, ql < static_cast<uint32_t>(G));
        while (mask) {
          const int j = __ffs(mask) - 1;
          mask &= mask - 1;
          const float wj = __shfl_sync(0xffffffffu, w, j);
          const int qj = __shfl_sync(0xffffffffu, static_cast<int>(ql), j);
          uint32_t v[VEC / 2];
          PackedLoad<VEC>::ld(stl + (c * 32 + j) * (DS * 2), v);
          fma_select<VEC, G>(acc, v, wj, qj);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars->empty[stage]));
      if (++stage == p.stages) { stage = 0; phase ^= 1; }
    }
#pragma unroll
    for (int j = 0; j < G; ++j) {
      const int q = q_lo + j;
      if (q < p.P) {
        float* o = p.dimg + (static_cast<size_t>(k) * p.P + q) * p.D + d0 + lane * VEC;
#pragma unroll
        for (int i = 0; i < VEC; i += 2) {
          float2 val = make_float2(acc[j][i], acc[j][i + 1]);
          if (p.accumulate_img) {
            const float2 old = *reinterpret_cast<const float2*>(o + i);
            val.x += old.x;
            val.y += old.y;
          }
          *reinterpret_cast<float2*>(o + i) = val;
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn2 encode_fn() {
  static EncodeTiledFn2 fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn2>(ptr);
  }
  return fn;
}

// [rows, D] bf16 row-major, box = box_cols x box_rows, no swizzle (rows land densely, box_cols*2 bytes apart)
static int make_tmap_plain(CUtensorMap* out, const void* base, long long rows, int D, int box_cols, int box_rows) {
  EncodeTiledFn2 fn = encode_fn();
  if (fn == nullptr) return CXL_ERR_CUDA;
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(D), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(D) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_cuda_error(cudaErrorInvalidValue, "cuTensorMapEncodeTiled (bwd)");
    return CXL_ERR_CUDA;
  }
  return CXL_OK;
}

static int round_up_i(int x, int m) { return (x + m - 1) / m * m; }

template <int DS, int R, int NWC>
static int launch_dtxt(const void* img, BwdStreamParams p, int max_smem, cudaStream_t stream) {
  constexpr int ROWS = NWC * R;
  // slab rows: the TMA box covers min(P,256) rows; P > 256 takes several boxes (last one may read into the
  // next image or past the end -> zero fill; those rows are never gathered)
  const int box_rows = p.P < 256 ? p.P : 256;
  const int slab_rows = round_up_i(p.P, box_rows);
  p.tile_bytes = slab_rows * DS * 2;
  p.stage_bytes = round_up_i(p.tile_bytes, 128);
  const int overhead = 128 + static_cast<int>(sizeof(StreamBarriers)) + NWC * 2 * R * 8;
  int stages = (max_smem - overhead) / p.stage_bytes;
  if (stages < 2) return CXL_ERR_UNSUPPORTED;
  p.stages = stages > 4 ? 4 : stages;
  const int smem_bytes = overhead + p.stages * p.stage_bytes;
  CUtensorMap tm;
  int rc = make_tmap_plain(&tm, img, static_cast<long long>(p.Bi) * p.p_stride, p.D, DS, box_rows);
  if (rc != CXL_OK) return rc;
  auto kern = maxsim_bwd_dtxt_stream_kernel<DS, R, NWC>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  dim3 grid((p.rows + ROWS - 1) / ROWS, p.D / DS);
  kern<<<grid, (NWC + 1) * 32, smem_bytes, stream>>>(tm, p);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

template <int DS, int G, int RT, int MAXW>
static int launch_dimg(const void* txt, BwdStreamParams p, int max_smem, cudaStream_t stream) {
  const int nwc = (p.P + G - 1) / G;
  if (nwc + RT / 32 > MAXW) return CXL_ERR_UNSUPPORTED;
  p.tile_bytes = RT * DS * 2;
  p.stage_bytes = round_up_i(p.tile_bytes + RT * 8, 128);
  const int overhead = 128 + static_cast<int>(sizeof(StreamBarriers));
  int stages = (max_smem - overhead) / p.stage_bytes;
  if (stages < 2) return CXL_ERR_UNSUPPORTED;
  p.stages = stages > 4 ? 4 : stages;
  const int smem_bytes = overhead + p.stages * p.stage_bytes;
  CUtensorMap tm;
  int rc = make_tmap_plain(&tm, txt, p.rows, p.D, DS, RT);
  if (rc != CXL_OK) return rc;
  auto kern = maxsim_bwd_dimg_stream_kernel<DS, G, RT, MAXW>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  dim3 grid(p.Bi, p.D / DS);
  kern<<<grid, (nwc + RT / 32) * 32, smem_bytes, stream>>>(tm, p, nwc);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

// returns CXL_ERR_UNSUPPORTED when the shape is outside the fast path (caller falls back to the generic kernels)
int maxsim_bwd_stream(const void* txt, const void* img, const void* argmax, long long ld_arg, const float* w, int Bt,
                      int Bi, int T, int P, int D, int p_stride, int ld_pairs, int k_off, float* dtxt, float* dimg,
                      int accumulate_img, cudaStream_t stream) {
  if (D % 128 != 0 || T < 8) return CXL_ERR_UNSUPPORTED;
  if (static_cast<long long>(Bt) * T >= (1LL << 31)) return CXL_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(txt) & 15) || (reinterpret_cast<uintptr_t>(img) & 15)) return CXL_ERR_UNSUPPORTED;
  BwdStreamParams p;
  memset(&p, 0, sizeof(p));
  p.w = w; p.argmax = argmax; p.ld_arg = ld_arg;
  p.Bt = Bt; p.Bi = Bi; p.T = T; p.P = P; p.D = D; p.p_stride = p_stride;
  p.rows = Bt * T;
  p.ld_pairs = ld_pairs; p.k_off = k_off; p.argmax_u16 = P > 256;
  p.dtxt = dtxt; p.dimg = dimg; p.accumulate_img = accumulate_img;
  int dev = 0, max_smem = 0;
  CXL_CUDA_CHECK(cudaGetDevice(&dev));
  CXL_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));

  int rc = CXL_OK;
  if (dtxt != nullptr) {
    if (P <= 1024) rc = launch_dtxt<64, 32, 19>(img, p, max_smem, stream);
    else rc = CXL_ERR_UNSUPPORTED;
    if (rc != CXL_OK) return rc;
  }
  if (dimg != nullptr) {
    if (P <= 28 * 8) rc = launch_dimg<128, 8, 128, 32>(txt, p, max_smem, stream);
    else if (P <= 24 * 24) rc = launch_dimg<64, 24, 128, 28>(txt, p, max_smem, stream);
    else rc = CXL_ERR_UNSUPPORTED;
    if (rc != CXL_OK) return rc;
  }
  return CXL_OK;
}

}  // namespace cxl
