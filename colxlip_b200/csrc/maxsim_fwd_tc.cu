// K1: fused MaxSim forward on 5th-gen tensor cores (tcgen05 + TMEM, operands fed by TMA).
//
// Replaces loss.py:34-44 / model.py:626-628 of the reference: the [Bt*T, D] x [D, Bi*P]
// contraction, the max over each image's patches and the per-text token sum, without ever
// writing the [Bt,Bi,T,P] similarity tensor.  Per (text token, image) only the argmax index
// is stored (u8/u16) for the backward.
//
// Tiling: M-tile = 128 consecutive rows of the flat [Bt*T, D] text-token matrix (one TMEM lane
// per row), N-tile = the patches of one image (or `ipt` small images, or one `part` of a large
// image), K = D in 64-wide blocks (128 B rows, SWIZZLE_128B).  The text tile stays resident in
// shared memory for a whole work unit while image tiles stream through a TMA ring; two 256-
// column TMEM accumulators let the max/argmax epilogue of tile i overlap the MMAs of tile i+1.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM allocator,
// warps 4..7 = epilogue (warp w reads TMEM lanes 32*(w%4)..+31).
#include "common.cuh"

namespace cxl {

static constexpr int kBlockM = 128;
static constexpr int kBlockK = 64;     // bf16 elements per 128-byte swizzled row
static constexpr int kUmmaK = 16;
static constexpr int kThreads = 256;
static constexpr int kThreads2 = 384;   // pair kernel: 8 epilogue warps (two per TMEM lane quadrant)
static constexpr int kABlockBytes = kBlockM * kBlockK * 2;   // 16 KB
static constexpr int kMaxStages = 8;

struct FwdTcParams {
  int Bt, Bi, T, P, D;
  int rows_a;            // Bt*T
  int num_kb;            // D / 64
  int n_tile;            // UMMA N (multiple of 16, <= 256)
  int ipt;               // images per N-tile (>= 1)
  int p_stride;          // rows per image in the img matrix == column stride between images in a tile
  int parts;             // N-tiles per image (1 unless P > 256)
  int part_cols;         // patch columns covered by one part
  int total_n_tiles;     // ceil(Bi / ipt) * parts
  int tiles_per_unit;    // multiple of parts
  int num_m_tiles;
  int num_units;
  int a_resident;        // 1: text tile resident in smem for the whole unit, 0: streamed with B
  int num_stages;
  int stage_bytes;       // bytes of one ring stage (B block, plus A block when streaming)
  int b_block_bytes;     // n_tile * 128
  int mean_mode;
  int argmax_u16;
  int f16;               // operands are fp16 instead of bf16 (same tiles, same rate; instruction descriptor formats differ)
  float* S;              // [Bt,Bi] accumulator of row maxima sums: pre-zeroed and added to atomically, or (S_extra != null)
                         // slot 0 of the fixed-order partial sums
  float* S_extra;        // deterministic sums: slots 1.. of the per-32-row-block partial sums of a text, [nslots-1][Bt*Bi];
                         // the block that holds rows R..R+31 of text m STORES its partial into slot (R>>5) - ((m*T)>>5)
                         // (exactly one writer per cell), the finalize kernel adds a text's slots in ascending order
  long long slot_stride; // Bt*Bi
  float* cnt;            // [Bt,Bi] pre-zeroed count of non-zero maxima (or null)
  void* argmax;          // [Bi][ld_arg] image-major (row = flat text-token index) or null
  long long ld_arg;
  float* dbg;            // debug: raw accumulators of (unit 0, first tile) as [128, n_tile], or null
  float* raw_c;          // plain GEMM mode (cxl_gemm_nt): C[rows_a, raw_n] = A * B^T, every accumulator stored; or null
  long long raw_ldc;     // row pitch of C in elements
  int raw_n;             // columns of C (rows of B)
#ifdef CXL_EXPERIMENTS
  int exp_mode;          // timing experiments, compiled only with -DCXL_EXPERIMENTS (env CXL_K1_EXP): bit0 = skip the
                         // epilogue scan (WRONG results); release builds carry no such switch
#endif
  // Pair kernel, packed tiles (env CXL_K1_PACK=1; P % 16 != 0): instead of padding every image to round_up(P,16)
  // columns, `pack_g` consecutive images (pack_g * P a multiple of 16) share pack_g tiles whose widths add up to
  // pack_g * P exactly.  Tile t of a group starts at flat patch row pack_start[t] of the group and is pack_n[t]
  // wide: image t from its column pack_start[t] - t*P on, followed by the first (< 16) columns of image t + 1,
  // whose running max is carried into the next tile.  0 = padded tiles.
  int round16;           // CXL_MEAN_ROUND16: row maxima are rounded to the operand dtype before they are summed (the
                         // reference's autocast einsum hands torch.max 16-bit similarities, loss.py:34-35)
  int l2_hint;           // pair kernel: bit0 = TMA loads carry L2 hints (image tiles evict_last, text tiles evict_first),
                         // bit1 = argmax bytes are stored with the streaming (.cs) qualifier
  int pack_g;
  uint16_t pack_start[16];
  uint16_t pack_n[16];
};

// RN to bf16 / fp16 and back (NaN stays NaN): what a 16-bit einsum output holds
__device__ __forceinline__ float round_to_16(float x, int f16) {
  return f16 ? __half2float(__float2half_rn(x)) : __bfloat162float(__float2bfloat16_rn(x));
}

struct __align__(8) FwdTcBarriers {
  uint64_t b_full[kMaxStages];
  uint64_t b_empty[kMaxStages];
  uint64_t a_full;
  uint64_t a_empty;
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  uint64_t xfull[2];     // pair kernel: xchg[buf] written by the upper-column epilogue warps (count 4)
  float2 xchg[2][128];   // pair kernel: (max, argmax) handed from the upper-column epilogue warp to the lower one
};

// Running (max, FIRST argmax -- torch.max semantics) over 16 more accumulator columns; pad columns never win.
__device__ __forceinline__ void scan16(const uint32_t (&v)[16], int c0, int ncols, int qoff, float& best, int& bidx) {
  float lb = -INFINITY;
  int lj = 0;
  if (c0 + 16 <= ncols) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float x = __uint_as_float(v[j]);
      if (x > lb) { lb = x; lj = j; }
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float x = __uint_as_float(v[j]);
      if (c0 + j < ncols && x > lb) { lb = x; lj = j; }
    }
  }
  if (lb > best) { best = lb; bidx = qoff + c0 + lj; }
}

// Packed tiles: the last `h` (< 16) columns of a tile's last chunk are patches 0..h-1 of the NEXT image.
__device__ __forceinline__ void scan_head(const uint32_t (&v)[16], int h, float& best, int& bidx) {
  best = -INFINITY;
  bidx = 0;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float x = __uint_as_float(v[j]);
    if (j >= 16 - h && x > best) { best = x; bidx = j - (16 - h); }
  }
}

__global__ void __launch_bounds__(kThreads, 1)
maxsim_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const FwdTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operand tiles need 1024-byte alignment
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int a_bytes = p.a_resident ? p.num_kb * kABlockBytes : 0;
  uint8_t* smem_a = smem;
  uint8_t* smem_ring = smem + a_bytes;
  FwdTcBarriers* bars = reinterpret_cast<FwdTcBarriers*>(smem_ring + p.num_stages * p.stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(smem_u32(&bars->b_full[s]), 1);
      mbar_init(smem_u32(&bars->b_empty[s]), 1);
    }
    mbar_init(smem_u32(&bars->a_full), 1);
    mbar_init(smem_u32(&bars->a_empty), 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(smem_u32(&bars->tmem_full[b]), 1);
      mbar_init(smem_u32(&bars->tmem_empty[b]), 4);   // one arrive per epilogue warp
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 2) tmem_alloc_512(smem_u32(&bars->tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer =====================
    int stage = 0;
    uint32_t phase = 0;
    uint32_t a_phase = 0;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      const int chunk = unit / p.num_m_tiles;
      const int m_tile = unit - chunk * p.num_m_tiles;
      const int j0 = chunk * p.tiles_per_unit;
      const int j1 = min(j0 + p.tiles_per_unit, p.total_n_tiles);
      if (p.a_resident) {
        mbar_wait(smem_u32(&bars->a_empty), a_phase ^ 1);      // previous unit's MMAs are done with A
        mbar_arrive_expect_tx(smem_u32(&bars->a_full), p.num_kb * kABlockBytes);
        for (int kb = 0; kb < p.num_kb; ++kb)
          tma_load_2d(smem_u32(smem_a + kb * kABlockBytes), &tmA, smem_u32(&bars->a_full), kb * kBlockK,
                      m_tile * kBlockM);
        a_phase ^= 1;
      }
      for (int j = j0; j < j1; ++j) {
        const int g = j / p.parts;
        const int part = j - g * p.parts;
        const int row0 = g * p.ipt * p.p_stride + part * p.part_cols;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&bars->b_empty[stage]), phase ^ 1);
          uint8_t* st = smem_ring + stage * p.stage_bytes;
          const uint32_t full = smem_u32(&bars->b_full[stage]);
          if (p.a_resident) {
            mbar_arrive_expect_tx(full, p.b_block_bytes);
            tma_load_2d(smem_u32(st), &tmB, full, kb * kBlockK, row0);
          } else {
            mbar_arrive_expect_tx(full, p.b_block_bytes + kABlockBytes);
            tma_load_2d(smem_u32(st), &tmA, full, kb * kBlockK, m_tile * kBlockM);
            tma_load_2d(smem_u32(st + kABlockBytes), &tmB, full, kb * kBlockK, row0);
          }
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // whole warp runs the warp-uniform control flow (descriptors stay in uniform registers); one elected lane issues
    const uint32_t idesc = umma_idesc_bf16(kBlockM, p.n_tile, p.f16);
    const uint64_t desc_a0 = umma_smem_desc_sw128(smem_u32(smem_a));
    const uint64_t desc_r0 = umma_smem_desc_sw128(smem_u32(smem_ring));
    const uint32_t a_step = kABlockBytes >> 4;            // descriptor address units are 16 bytes
    const uint32_t s_step = static_cast<uint32_t>(p.stage_bytes) >> 4;
    int stage = 0;
    uint32_t phase = 0;
    uint32_t a_phase = 0;
    int buf = 0;
    uint32_t acc_phase = 0;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      const int chunk = unit / p.num_m_tiles;
      const int j0 = chunk * p.tiles_per_unit;
      const int j1 = min(j0 + p.tiles_per_unit, p.total_n_tiles);
      if (p.a_resident) {
        mbar_wait(smem_u32(&bars->a_full), a_phase);
        a_phase ^= 1;
      }
      for (int j = j0; j < j1; ++j) {
        mbar_wait(smem_u32(&bars->tmem_empty[buf]), acc_phase ^ 1);   // epilogue drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * 256;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&bars->b_full[stage]), phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t ds = desc_r0 + static_cast<uint64_t>(stage * s_step);
            const uint64_t da = p.a_resident ? desc_a0 + static_cast<uint64_t>(kb * a_step) : ds;
            const uint64_t db = p.a_resident ? ds : ds + a_step;
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k)
              umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);       // +32 bytes per UMMA_K
            umma_commit(smem_u32(&bars->b_empty[stage]));              // frees the ring slot when MMAs retire
          }
          __syncwarp();
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit(smem_u32(&bars->tmem_full[buf]));  // accumulator ready for the epilogue
        __syncwarp();
        buf ^= 1;
        if (buf == 0) acc_phase ^= 1;
      }
      if (p.a_resident) {
        if (elect_one()) umma_commit(smem_u32(&bars->a_empty));
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: max/argmax over patches, per-text sums =====================
    const int quad = warp & 3;
    const int row_in_tile = quad * 32 + lane;
    int buf = 0;
    uint32_t acc_phase = 0;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      const int chunk = unit / p.num_m_tiles;
      const int m_tile = unit - chunk * p.num_m_tiles;
      const int j0 = chunk * p.tiles_per_unit;
      const int j1 = min(j0 + p.tiles_per_unit, p.total_n_tiles);
      const int R = m_tile * kBlockM + row_in_tile;
      const bool valid = R < p.rows_a;
      const int m = R / p.T;
      // segmented-reduction predicates: lanes of one text are contiguous
      uint32_t same = 0;
#pragma unroll
      for (int o = 0; o < 5; ++o) {
        const int off = 1 << o;
        const int mo = __shfl_down_sync(0xffffffffu, m, off);
        if (lane + off < 32 && mo == m) same |= 1u << o;
      }
      const int m_prev = __shfl_up_sync(0xffffffffu, m, 1);
      const bool head = valid && (lane == 0 || m_prev != m);
      const size_t pair_base = static_cast<size_t>(m) * p.Bi;

      float best = -INFINITY;
      int bidx = 0;
      for (int j = j0; j < j1; ++j) {
        const int g = j / p.parts;
        const int part = j - g * p.parts;
        mbar_wait(smem_u32(&bars->tmem_full[buf]), acc_phase);
        tc_fence_after();
        const uint32_t tbase = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + buf * 256;

        if (p.dbg != nullptr && unit == 0 && j == j0) {
          for (int c0 = 0; c0 < p.n_tile; c0 += 16) {
            uint32_t v[16];
            tmem_ld16(tbase + c0, v);
            tmem_wait_ld16(v);
#pragma unroll
            for (int jj = 0; jj < 16; ++jj)
              p.dbg[static_cast<size_t>(row_in_tile) * p.n_tile + c0 + jj] = __uint_as_float(v[jj]);
          }
        }

        if (p.raw_c != nullptr) {
          // plain GEMM: tile j covers C columns [j * p_stride, j * p_stride + n_tile); lane = row, 16 columns per tcgen05.ld
          const int col0 = j * p.p_stride;
          float* crow = p.raw_c + static_cast<size_t>(R) * p.raw_ldc;
          const bool vec_ok = (p.raw_ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(p.raw_c) & 15) == 0;
          uint32_t va[16], vb[16];
          tmem_ld16(tbase, va);
          for (int c0 = 0; c0 < p.n_tile; c0 += 16) {
            tmem_wait_ld16(va);
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) vb[jj] = va[jj];
            if (c0 + 16 < p.n_tile) tmem_ld16(tbase + c0 + 16, va);     // next chunk in flight while this one is stored
            const int cc = col0 + c0;
            if (valid && cc < p.raw_n) {
              if (vec_ok && cc + 16 <= p.raw_n) {
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4)
                  *reinterpret_cast<float4*>(crow + cc + 4 * q4) =
                      make_float4(__uint_as_float(vb[4 * q4]), __uint_as_float(vb[4 * q4 + 1]),
                                  __uint_as_float(vb[4 * q4 + 2]), __uint_as_float(vb[4 * q4 + 3]));
              } else {
#pragma unroll
                for (int jj = 0; jj < 16; ++jj)
                  if (cc + jj < p.raw_n) crow[cc + jj] = __uint_as_float(vb[jj]);
              }
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&bars->tmem_empty[buf]));
          buf ^= 1;
          if (buf == 0) acc_phase ^= 1;
          continue;
        }
        const int qoff = part * p.part_cols;
        const int ncols = (p.parts == 1) ? p.P : min(p.part_cols, p.P - qoff);
        const int nchunks = (ncols + 15) >> 4;
        for (int s = 0; s < p.ipt; ++s) {
          const int k = g * p.ipt + s;
          if (k >= p.Bi) break;
          if (part == 0) { best = -INFINITY; bidx = 0; }
          const uint32_t cbase = tbase + s * p.p_stride;
          uint32_t va[16], vb[16];
          tmem_ld16(cbase, va);
          for (int c = 0; c < nchunks; c += 2) {
            tmem_wait_ld16(va);
            if (c + 1 < nchunks) tmem_ld16(cbase + (c + 1) * 16, vb);
            scan16(va, c * 16, ncols, qoff, best, bidx);
            if (c + 1 < nchunks) {
              tmem_wait_ld16(vb);
              if (c + 2 < nchunks) tmem_ld16(cbase + (c + 2) * 16, va);
              scan16(vb, (c + 1) * 16, ncols, qoff, best, bidx);
            }
          }
          if (part == p.parts - 1) {
            // ---- emit (row, image k): argmax byte, then per-text segmented sum of the maxima ----
            if (valid && p.argmax != nullptr) {
              const size_t off = static_cast<size_t>(k) * p.ld_arg + R;
              if (p.argmax_u16) reinterpret_cast<uint16_t*>(p.argmax)[off] = static_cast<uint16_t>(bidx);
              else reinterpret_cast<uint8_t*>(p.argmax)[off] = static_cast<uint8_t>(bidx);
            }
            if (p.round16) best = round_to_16(best, p.f16);
            float v = valid ? best : 0.f;
            float nz = (valid && best != 0.f) ? 1.f : 0.f;
#pragma unroll
            for (int o = 0; o < 5; ++o) {
              const float tv = __shfl_down_sync(0xffffffffu, v, 1 << o);
              const float tn = __shfl_down_sync(0xffffffffu, nz, 1 << o);
              if (same & (1u << o)) { v += tv; nz += tn; }
            }
            if (head) {
              if (p.S_extra != nullptr) {
                const int slot = (R >> 5) - ((m * p.T) >> 5);
                float* dst = slot == 0 ? p.S : p.S_extra + static_cast<size_t>(slot - 1) * p.slot_stride;
                dst[pair_base + k] = v;
              } else {
                atomicAdd(p.S + pair_base + k, v);
              }
              if (p.cnt != nullptr) atomicAdd(p.cnt + pair_base + k, nz);
            }
          }
        }
        // all tcgen05.ld of this warp have completed (wait::ld above): hand the accumulator back
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars->tmem_empty[buf]));
        buf ^= 1;
        if (buf == 0) acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_512(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------
// K1, CTA-pair variant (cta_group::2): one UMMA covers 256 text-token rows x one image.  Each CTA of the pair
// keeps ITS 128-row text tile resident and streams only HALF of every image tile (n_tile/2 patch rows), so the
// per-SM L2->smem traffic and the shared-memory operand reads of the B operand are halved against the single-
// CTA kernel, which is bound by the 128 B/clk shared-memory port (TMA writes + UMMA operand reads).
// Barriers: the even CTA owns a_full / b_full (both CTAs' TMA loads credit them) and tmem_empty (8 epilogue
// warps); b_empty / a_empty / tmem_full exist in both CTAs and are signalled by multicast tcgen05.commit.
// SA = false: 128 < P <= 256 (one image per N tile) with a text tile that fits next to the ring.
// SA = true ("streamed A"): the text tile does not stay resident (D = 768: 192 KB) or an image spans several N tiles
// (P > 256: `parts` tiles of part_cols columns, running max carried from part to part).  A ring stage then holds this
// CTA's [128 x 64] text block AND its half of the image block; per k-block and CTA the shared-memory port moves
// 16 + 12 KB of TMA writes and 16 + 12 KB of operand reads (N = 192) against 16 + 24 and 16 + 24 KB in the single-CTA
// kernel, which is what bounds that kernel at the cfg-5 geometry.
template <bool SA>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, 1)
maxsim_fwd_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const FwdTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_ring = SA ? smem : smem + p.num_kb * kABlockBytes;
  FwdTcBarriers* bars = reinterpret_cast<FwdTcBarriers*>(smem_ring + p.num_stages * p.stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();              // 0 = even CTA (issues the MMAs)
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int half_rows = p.n_tile >> 1;                  // image-tile rows loaded by each CTA

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(smem_u32(&bars->b_full[s]), 1);
      mbar_init(smem_u32(&bars->b_empty[s]), 1);
    }
    mbar_init(smem_u32(&bars->a_full), 1);
    mbar_init(smem_u32(&bars->a_empty), 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(smem_u32(&bars->tmem_full[b]), 1);
      mbar_init(smem_u32(&bars->xfull[b]), 4);
      mbar_init(smem_u32(&bars->tmem_empty[b]), 16);   // 8 epilogue warps of each CTA (even CTA's copy is used)
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 2) tmem_alloc_512_2cta(smem_u32(&bars->tmem_base));
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                    // peer barriers initialised, peer TMEM allocated
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer (both CTAs) =====================
    int stage = 0;
    uint32_t phase = 0;
    uint32_t a_phase = 0;
    const uint32_t a_full = smem_u32(&bars->a_full);
    // L2 residency: the image block every cluster sweeps is re-read by all waves of units (keep it), a text tile is
    // read once per block by one CTA (let it pass through)
    const bool hinted = (p.l2_hint & 1) != 0;
    const uint64_t pol_img = l2_policy_evict_last(), pol_txt = l2_policy_evict_first();
    for (int unit = cluster_id; unit < p.num_units; unit += num_clusters) {
      const int chunk = unit / p.num_m_tiles;            // num_m_tiles counts 256-row PAIRS here
      const int m_tile = (unit - chunk * p.num_m_tiles) * 2 + static_cast<int>(rank);
      const int j0 = chunk * p.tiles_per_unit;
      const int j1 = min(j0 + p.tiles_per_unit, p.total_n_tiles);
      if constexpr (!SA) {
        mbar_wait(smem_u32(&bars->a_empty), a_phase ^ 1);        // previous unit's MMAs are done with A
        if (rank == 0) mbar_arrive_expect_tx(a_full, 2 * p.num_kb * kABlockBytes);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          if (hinted) tma_load_2d_2cta_hint(smem_u32(smem_a + kb * kABlockBytes), &tmA, a_full, kb * kBlockK, m_tile * kBlockM, pol_txt);
          else tma_load_2d_2cta(smem_u32(smem_a + kb * kABlockBytes), &tmA, a_full, kb * kBlockK, m_tile * kBlockM);
        }
        a_phase ^= 1;
      }
      for (int j = j0; j < j1; ++j) {
        int row0 = j * p.p_stride + static_cast<int>(rank) * half_rows;
        if constexpr (SA) {                              // tile j = part (j % parts) of image j / parts
          const int g = j / p.parts;
          row0 = g * p.p_stride + (j - g * p.parts) * p.part_cols + static_cast<int>(rank) * half_rows;
        } else if (p.pack_g) {                           // packed tiles: this CTA's half of tile t of group grp
          const int grp = j / p.pack_g, t = j - grp * p.pack_g;
          row0 = grp * p.pack_g * p.P + p.pack_start[t] + static_cast<int>(rank) * (p.pack_n[t] >> 1);
        }
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(smem_u32(&bars->b_empty[stage]), phase ^ 1);
          const uint32_t full = smem_u32(&bars->b_full[stage]);
          if constexpr (SA) {
            if (rank == 0) mbar_arrive_expect_tx(full, 2 * (kABlockBytes + p.b_block_bytes));
            const uint32_t st = smem_u32(smem_ring + stage * p.stage_bytes);
            tma_load_2d_2cta(st, &tmA, full, kb * kBlockK, m_tile * kBlockM);
            tma_load_2d_2cta(st + kABlockBytes, &tmB, full, kb * kBlockK, row0);
          } else {
            if (rank == 0) mbar_arrive_expect_tx(full, 2 * p.b_block_bytes);
            if (hinted) tma_load_2d_2cta_hint(smem_u32(smem_ring + stage * p.stage_bytes), &tmB, full, kb * kBlockK, row0, pol_img);
            else tma_load_2d_2cta(smem_u32(smem_ring + stage * p.stage_bytes), &tmB, full, kb * kBlockK, row0);
          }
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ===================== MMA issuer (even CTA only) =====================
    // The whole warp runs the (warp-uniform) control flow so that addresses and descriptors stay in the uniform
    // datapath; one elected lane issues the tcgen05 instructions.
    const uint32_t idesc = umma_idesc_bf16(2 * kBlockM, p.n_tile, p.f16);
    const uint64_t desc_a0 = umma_smem_desc_sw128(smem_u32(smem_a));
    const uint64_t desc_b0 = umma_smem_desc_sw128(smem_u32(smem_ring));
    const uint32_t a_step = kABlockBytes >> 4;            // descriptor address units are 16 bytes
    const uint32_t b_step = static_cast<uint32_t>(p.stage_bytes) >> 4;
    const int num_kb = p.num_kb;
    const int num_stages = p.num_stages;
    int stage = 0;
    uint32_t phase = 0;
    uint32_t a_phase = 0;
    int buf = 0;
    uint32_t acc_phase = 0;
    for (int unit = cluster_id; unit < p.num_units; unit += num_clusters) {
      const int chunk = unit / p.num_m_tiles;
      const int j0 = chunk * p.tiles_per_unit;
      const int j1 = min(j0 + p.tiles_per_unit, p.total_n_tiles);
      if constexpr (!SA) {
        mbar_wait(smem_u32(&bars->a_full), a_phase);
        a_phase ^= 1;
      }
      for (int j = j0; j < j1; ++j) {
        mbar_wait(smem_u32(&bars->tmem_empty[buf]), acc_phase ^ 1);   // both CTAs' epilogues drained this buffer
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * 256;
        const uint32_t idesc_j = p.pack_g ? umma_idesc_bf16(2 * kBlockM, p.pack_n[j % p.pack_g], p.f16) : idesc;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(smem_u32(&bars->b_full[stage]), phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t da = SA ? desc_b0 + static_cast<uint64_t>(stage * b_step)       // text block at the head of the stage
                                   : desc_a0 + static_cast<uint64_t>(kb * a_step);
            const uint64_t db = desc_b0 + static_cast<uint64_t>(stage * b_step) + (SA ? (kABlockBytes >> 4) : 0);
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k)
              umma_bf16_2cta(tmem_d, da + 2 * k, db + 2 * k, idesc_j, (kb | k) != 0);   // +32 bytes per UMMA_K
            umma_commit_2cta(smem_u32(&bars->b_empty[stage]), 3);      // frees the ring slot in both CTAs
          }
          __syncwarp();
          if (++stage == num_stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit_2cta(smem_u32(&bars->tmem_full[buf]), 3);   // accumulator ready in both CTAs
        __syncwarp();
        buf ^= 1;
        if (buf == 0) acc_phase ^= 1;
      }
      if constexpr (!SA) {
        if (elect_one()) umma_commit_2cta(smem_u32(&bars->a_empty), 3);
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (both CTAs): max/argmax over patches, per-text sums =====================
    // Two warps per TMEM lane quadrant: warp 4+q scans the lower column chunks of rows 32q..32q+31, warp 8+q the
    // upper ones and hands its (max, argmax) over through shared memory; the lower warp merges and emits.
    const int quad = warp & 3;
    const bool upper = warp >= 8;
    const int row_in_tile = quad * 32 + lane;
    int buf = 0;
    uint32_t acc_phase = 0;
    int nchunks = (p.P + 15) >> 4;
    int c_split = (nchunks + 1) >> 1;
    int c_lo = upper ? c_split : 0;
    int c_hi = upper ? nchunks : c_split;
    int col_off = 0;                                      // packed tiles: image column of the tile's first column
    int head_n = 0;                                       // packed tiles: columns of the NEXT image at the tile's end
    float carry_best = -INFINITY;                         // lower warp: max over the next image's head columns
    int carry_idx = 0;
    float2* xhead = reinterpret_cast<float2*>(bars + 1);  // [2][128], allocated only for packed tiles
    for (int unit = cluster_id; unit < p.num_units; unit += num_clusters) {
      const int chunk = unit / p.num_m_tiles;
      const int m_tile = (unit - chunk * p.num_m_tiles) * 2 + static_cast<int>(rank);
      const int j0 = chunk * p.tiles_per_unit;
      const int j1 = min(j0 + p.tiles_per_unit, p.total_n_tiles);
      const int R = m_tile * kBlockM + row_in_tile;
      const bool valid = R < p.rows_a;
      const int m = R / p.T;
      uint32_t same = 0;                                  // segmented-reduction predicates: lanes of one text are contiguous
#pragma unroll
      for (int o = 0; o < 5; ++o) {
        const int off = 1 << o;
        const int mo = __shfl_down_sync(0xffffffffu, m, off);
        if (lane + off < 32 && mo == m) same |= 1u << o;
      }
      const int m_prev = __shfl_up_sync(0xffffffffu, m, 1);
      const bool head = valid && (lane == 0 || m_prev != m);
      float* s_row = p.S + static_cast<size_t>(m) * p.Bi;
      float* c_row = p.cnt != nullptr ? p.cnt + static_cast<size_t>(m) * p.Bi : nullptr;
      float* s_slot = nullptr;
      if (p.S_extra != nullptr) {
        const int slot = (R >> 5) - ((m * p.T) >> 5);
        s_slot = slot == 0 ? s_row : p.S_extra + static_cast<size_t>(slot - 1) * p.slot_stride + static_cast<size_t>(m) * p.Bi;
      }

      float run_best = -INFINITY;                         // SA: lower warp's running (max, argmax) over the parts of an image
      int run_idx = 0;
      for (int k = j0; k < j1; ++k) {                     // one image per tile (SA: tile k = part k % parts of image k / parts)
        mbar_wait(smem_u32(&bars->tmem_full[buf]), acc_phase);
        tc_fence_after();
        const uint32_t cbase = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + buf * 256;
        float best = -INFINITY;
        int bidx = 0;
        int img = k, qoff = 0, ncols_t = p.P;
        bool last_part = true;
        if constexpr (SA) {
          img = k / p.parts;
          const int part = k - img * p.parts;
          qoff = part * p.part_cols;
          ncols_t = min(p.part_cols, p.P - qoff);
          last_part = part == p.parts - 1;
          nchunks = (ncols_t + 15) >> 4;
          c_split = (nchunks + 1) >> 1;
          c_lo = upper ? c_split : 0;
          c_hi = upper ? nchunks : c_split;
          if (!upper && part > 0) { best = run_best; bidx = run_idx; }     // earlier parts hold the lower indices
        } else if (p.pack_g) {                            // geometry of tile t = k % pack_g (units start at t = 0)
          const int t = k % p.pack_g;
          col_off = static_cast<int>(p.pack_start[t]) - t * p.P;
          head_n = static_cast<int>(p.pack_start[t]) + static_cast<int>(p.pack_n[t]) - (t + 1) * p.P;
          nchunks = p.pack_n[t] >> 4;
          c_split = (nchunks + 1) >> 1;
          c_lo = upper ? c_split : 0;
          c_hi = upper ? nchunks : c_split;
          if (!upper && col_off > 0) { best = carry_best; bidx = carry_idx; }   // columns 0..col_off-1 came first
        }
        float hbest = -INFINITY;
        int hidx = 0;
        if (c_lo < c_hi) {
          uint32_t va[16], vb[16];
          tmem_ld16(cbase + c_lo * 16, va);
#ifdef CXL_EXPERIMENTS
          const int c_end = (p.exp_mode & 1) ? c_lo + 1 : c_hi;
#else
          const int c_end = c_hi;
#endif
          for (int c = c_lo; c < c_end; c += 2) {
            tmem_wait_ld16(va);
            if (c + 1 < c_hi) tmem_ld16(cbase + (c + 1) * 16, vb);
            scan16(va, c * 16 + col_off, ncols_t, qoff, best, bidx);
            if (c + 1 < c_hi) {
              tmem_wait_ld16(vb);
              if (c + 2 < c_hi) tmem_ld16(cbase + (c + 2) * 16, va);
              scan16(vb, (c + 1) * 16 + col_off, ncols_t, qoff, best, bidx);
            }
          }
          if (upper && head_n > 0) {                      // the tile's last chunk again: head of the next image
            tmem_ld16(cbase + (nchunks - 1) * 16, va);
            tmem_wait_ld16(va);
            scan_head(va, head_n, hbest, hidx);
          }
        }
        tc_fence_before();
        if (upper) {
          // all tcgen05.ld of this warp have completed (wait::ld above): hand the accumulator back, pass the result on
          if (p.pack_g) xhead[buf * 128 + row_in_tile] = make_float2(hbest, __int_as_float(hidx));
          bars->xchg[buf][row_in_tile] = make_float2(best, __int_as_float(bidx));
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(smem_u32(&bars->xfull[buf]));
            mbar_arrive_even_cta(smem_u32(&bars->tmem_empty[buf]));
          }
        } else {
          mbar_wait(smem_u32(&bars->xfull[buf]), acc_phase);
          const float2 o = bars->xchg[buf][row_in_tile];  // upper columns: wins only if strictly greater
          if (p.pack_g) {
            const float2 hv = xhead[buf * 128 + row_in_tile];
            carry_best = hv.x;
            carry_idx = __float_as_int(hv.y);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive_even_cta(smem_u32(&bars->tmem_empty[buf]));   // also orders xchg reuse after this read
          if (o.x > best) { best = o.x; bidx = __float_as_int(o.y); }
          if (SA && !last_part) {                         // more parts of this image to come: carry, emit nothing yet
            run_best = best;
            run_idx = bidx;
            buf ^= 1;
            if (buf == 0) acc_phase ^= 1;
            continue;
          }
          if (valid && p.argmax != nullptr) {
            const size_t off = static_cast<size_t>(img) * p.ld_arg + R;
            if (p.argmax_u16) reinterpret_cast<uint16_t*>(p.argmax)[off] = static_cast<uint16_t>(bidx);
            else if (p.l2_hint & 2) __stcs(reinterpret_cast<unsigned char*>(p.argmax) + off, static_cast<unsigned char>(bidx));
            else reinterpret_cast<uint8_t*>(p.argmax)[off] = static_cast<uint8_t>(bidx);
          }
          if (p.round16) best = round_to_16(best, p.f16);
          float v = valid ? best : 0.f;
          float nz = (valid && best != 0.f) ? 1.f : 0.f;
#pragma unroll
          for (int o2 = 0; o2 < 5; ++o2) {
            const float tv = __shfl_down_sync(0xffffffffu, v, 1 << o2);
            const float tn = __shfl_down_sync(0xffffffffu, nz, 1 << o2);
            if (same & (1u << o2)) { v += tv; nz += tn; }
          }
          if (head) {
            if (s_slot != nullptr) s_slot[img] = v;        // fixed-order sums: this block's own slot, one writer
            else atomicAdd(s_row + img, v);
            if (c_row != nullptr) atomicAdd(c_row + img, nz);
          }
        }
        buf ^= 1;
        if (buf == 0) acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                    // no CTA leaves while its peer may still signal it
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_512_2cta(tmem_base);
  }
}

// S <- final score: masked mean  sum / (cnt + 1e-8)   (loss.py:40-44)  or plain mean sum / T (model.py:628)
__global__ void maxsim_finalize_kernel(float* __restrict__ S, const float* __restrict__ cnt, size_t n, float inv_T,
                                       int mean_mode) {
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (; i < n; i += stride) {
    const float s = S[i];
    S[i] = (mean_mode == CXL_MEAN_MASKED) ? s / (cnt[i] + 1e-8f) : s * inv_T;
  }
}

// the same with the fixed-order slot sums in front: S[m,k] = slot0 + slot1 + ... over the 32-row (slot_rows-row) blocks
// text m touches, ascending -- every run adds the same numbers in the same order
__global__ void maxsim_finalize_slots_kernel(float* __restrict__ S, const float* __restrict__ S_extra,
                                             const float* __restrict__ cnt, int Bt, int Bi, int T, int slot_rows,
                                             float inv_T, int mean_mode) {
  const size_t n = static_cast<size_t>(Bt) * Bi;
  for (int m = blockIdx.y; m < Bt; m += gridDim.y) {
    const int last = (m * T + T - 1) / slot_rows - (m * T) / slot_rows;      // highest slot text m wrote
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < Bi; k += gridDim.x * blockDim.x) {
      const size_t i = static_cast<size_t>(m) * Bi + k;
      float s = S[i];
      for (int sl = 1; sl <= last; ++sl) s += S_extra[static_cast<size_t>(sl - 1) * n + i];
      S[i] = (mean_mode == CXL_MEAN_MASKED) ? s / (cnt[i] + 1e-8f) : s * inv_T;
    }
  }
}

int maxsim_fwd_slots(int T, int slot_rows) { return (T + slot_rows - 2) / slot_rows + 1; }

int launch_maxsim_finalize_slots(float* S, const float* S_extra, const float* cnt, int Bt, int Bi, int T, int slot_rows,
                                 int mean_mode, cudaStream_t stream) {
  if (Bt <= 0 || Bi <= 0) return CXL_OK;
  int gx = (Bi + 255) / 256;
  if (gx > 64) gx = 64;
  dim3 g(gx, Bt > 65535 ? 65535 : Bt);
  maxsim_finalize_slots_kernel<<<g, 256, 0, CXL_COUNTED(stream)>>>(S, S_extra, cnt, Bt, Bi, T, slot_rows,
                                                                  1.0f / static_cast<float>(T), mean_mode);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

int launch_maxsim_finalize(float* S, const float* cnt, size_t npairs, int T, int mean_mode, cudaStream_t stream) {
  const int fb = 256;
  const int sms = num_sms();
  size_t fg = (npairs + fb - 1) / fb;
  if (fg > static_cast<size_t>(sms) * 16) fg = static_cast<size_t>(sms) * 16;
  if (fg == 0) return CXL_OK;
  maxsim_finalize_kernel<<<static_cast<unsigned>(fg), fb, 0, CXL_COUNTED(stream)>>>(S, cnt, npairs, 1.0f / static_cast<float>(T),
                                                                      mean_mode);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// [rows, D] bf16 row-major matrix, box = 64 columns x box_rows rows, 128-byte swizzle
static int make_tmap(CUtensorMap* out, const void* base, long long rows, int D, int box_rows, int f16) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) return CXL_ERR_CUDA;
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(D), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(D) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kBlockK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_cuda_error(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    return CXL_ERR_CUDA;
  }
  return CXL_OK;
}

static int round_up(int x, int m) { return (x + m - 1) / m * m; }

int img_row_stride(int P) { return P <= 128 ? round_up(P, 8) : P; }

int maxsim_fwd_tc(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int mean_mode, float* S,
                  float* cnt, void* argmax, long long ld_arg, float* dbg, int force_stream_a, int force_grid, int f16,
                  cudaStream_t stream, float* S_extra) {
  if (D % kBlockK != 0 || D < kBlockK) return CXL_ERR_UNSUPPORTED;
  if (P > 65535) return CXL_ERR_UNSUPPORTED;
  const int round16 = (mean_mode & CXL_MEAN_ROUND16) ? 1 : 0;
  mean_mode &= ~CXL_MEAN_ROUND16;
  if (mean_mode == CXL_MEAN_MASKED && cnt == nullptr) return CXL_ERR_INVALID_ARG;
  FwdTcParams p;
  memset(&p, 0, sizeof(p));
  p.round16 = round16;
  p.S_extra = S_extra;                                   // null: atomic sums; else the extra slots of the fixed-order sums
  p.slot_stride = static_cast<long long>(Bt) * Bi;
  p.Bt = Bt; p.Bi = Bi; p.T = T; p.P = P; p.D = D;
  p.f16 = f16 ? 1 : 0;
  p.rows_a = Bt * T;
  p.num_kb = D / kBlockK;
  p.p_stride = img_row_stride(P);
  if (P <= 128) {
    p.ipt = (256 - round_up(P, 16)) / p.p_stride + 1;
    if (p.ipt > Bi) p.ipt = Bi;
    p.parts = 1;
    p.part_cols = P;
    p.n_tile = round_up((p.ipt - 1) * p.p_stride + P, 16);
  } else if (P <= 256) {
    p.ipt = 1; p.parts = 1; p.part_cols = P;
    p.n_tile = round_up(P, 16);
  } else {
    p.ipt = 1;
    p.parts = (P + 255) / 256;
    p.part_cols = round_up((P + p.parts - 1) / p.parts, 16);
    p.n_tile = p.part_cols;
  }
  p.total_n_tiles = ((Bi + p.ipt - 1) / p.ipt) * p.parts;
  p.num_m_tiles = (p.rows_a + kBlockM - 1) / kBlockM;
  p.b_block_bytes = p.n_tile * kBlockK * 2;

  int dev = 0, max_smem = 0;
  CXL_CUDA_CHECK(cudaGetDevice(&dev));
  CXL_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const int overhead = 1024 /*alignment*/ + static_cast<int>(sizeof(FwdTcBarriers)) + 64;

  // ---- CTA-pair kernel: one image per N tile, both operands' tiles fit, no scheduling overrides ----
  static const int pair_env = getenv("CXL_K1_PAIR") ? atoi(getenv("CXL_K1_PAIR")) : 1;
  if (pair_env && p.ipt == 1 && p.parts == 1 && dbg == nullptr && !force_stream_a && force_grid <= 0 &&
      p.n_tile % 16 == 0) {
    const int half_bytes = (p.n_tile / 2) * kBlockK * 2;
    const int half_stage = round_up(half_bytes, 1024);
    int st2 = (max_smem - overhead - p.num_kb * kABlockBytes) / half_stage;
    if ((p.n_tile / 2) % 8 == 0 && st2 >= 4) {
      FwdTcParams q = p;
      q.a_resident = 1;
      q.b_block_bytes = half_bytes;
      q.stage_bytes = half_stage;
      q.num_stages = st2 > kMaxStages ? kMaxStages : st2;
      q.num_m_tiles = (p.rows_a + 2 * kBlockM - 1) / (2 * kBlockM);      // 256-row pairs
      const int sms2 = num_sms();
      const int clusters = sms2 / 2;
      // units = (256-row pair, chunk of images).  Whole image sweeps per unit keep the text tile resident longest;
      // split into more chunks only while the static round-robin over clusters would leave > 3 % of them idle.
      int chunks2 = 1;
      for (; chunks2 < 64 && chunks2 < q.total_n_tiles; ++chunks2) {
        const long long u = static_cast<long long>(chunks2) * q.num_m_tiles;
        const long long waves = (u + clusters - 1) / clusters;
        if (static_cast<double>(u) >= 0.97 * static_cast<double>(waves * clusters)) break;
      }
      q.tiles_per_unit = (q.total_n_tiles + chunks2 - 1) / chunks2;
      // Image blocks: units are ordered chunk-major, so all clusters work on the SAME block of images at a time.  A block
      // that fits L2 (CXL_K1_IMG_BLOCK images, default 192 = 38 MB of bf16 patches at P 196, D 512; the two dies each
      // cache their own copy, so 384 images = 77 MB already thrash the 126 MB L2) is pulled from DRAM once and re-read
      // from L2 by every wave of row-pair units; whole sweeps per unit re-stream the 6.6 GB operand from DRAM for every
      // wave.  Costs one extra text-tile load (128 KB) per unit.  K1 is power-capped, so DRAM energy is clock:
      // one B200, [16384 x 32768] launches interleaved: block 0 5871 / 5949 ms, 96 5810, 192 5801 / 5807, 384 6161 ms;
      // [4096 x 32768]: 1473 -> 1450 ms, DRAM reads 534 -> 150 GB; one [4096 x 4096] shard of the 8-GPU pipeline
      // 180.8 -> 176.1 ms (profiles/r4j_k1_img_block_ab_1gpu.txt, r4j_k1_img_block_ncu.txt).
      {
        static const int blk_env = getenv("CXL_K1_IMG_BLOCK") ? atoi(getenv("CXL_K1_IMG_BLOCK")) : 192;
        if (blk_env > 0 && q.tiles_per_unit > blk_env && q.num_m_tiles >= 4 * clusters) {
          const int nblk = (q.total_n_tiles + blk_env - 1) / blk_env;
          q.tiles_per_unit = (q.total_n_tiles + nblk - 1) / nblk;
        }
      }
      // packed tiles (opt-in, CXL_K1_PACK=1): groups of g images with g*P % 16 == 0 share g tiles without pad columns
      int extra_smem = 0;
      static const int pack_env = getenv("CXL_K1_PACK") ? atoi(getenv("CXL_K1_PACK")) : 0;
      if (pack_env && P % 16 != 0) {
        int g = 16;
        for (int d = 1; d <= 16; d <<= 1)
          if ((P * d) % 16 == 0) { g = d; break; }
        bool ok = true;
        int start = 0;
        for (int t = 0; t < g; ++t) {
          const int end = round_up((t + 1) * P, 16);
          const int n = end - start;
          ok = ok && n <= q.n_tile && n >= 32 && start - t * P >= 0 && start - t * P < 16;
          q.pack_start[t] = static_cast<uint16_t>(start);
          q.pack_n[t] = static_cast<uint16_t>(n);
          start = end;
        }
        if (ok) {
          q.pack_g = g;
          q.tiles_per_unit = round_up(q.tiles_per_unit, g);     // every unit starts on a group boundary
          extra_smem = 2 * 128 * static_cast<int>(sizeof(float2));
        }
      }
      chunks2 = (q.total_n_tiles + q.tiles_per_unit - 1) / q.tiles_per_unit;
      q.num_units = chunks2 * q.num_m_tiles;
      {
        static const int hint_env = getenv("CXL_K1_L2HINT") ? atoi(getenv("CXL_K1_L2HINT")) : 0;
        q.l2_hint = hint_env;
      }
      q.mean_mode = mean_mode;
      q.argmax_u16 = P > 256;
      q.S = S; q.cnt = cnt; q.argmax = argmax; q.ld_arg = ld_arg; q.dbg = nullptr;
#ifdef CXL_EXPERIMENTS
      q.exp_mode = getenv("CXL_K1_EXP") ? atoi(getenv("CXL_K1_EXP")) : 0;
#endif
      CUtensorMap tmA2, tmB2;
      int rc2 = make_tmap(&tmA2, txt, q.rows_a, D, kBlockM, p.f16);
      if (rc2 != CXL_OK) return rc2;
      rc2 = make_tmap(&tmB2, img, static_cast<long long>(Bi) * q.p_stride, D, q.n_tile / 2, p.f16);
      if (rc2 != CXL_OK) return rc2;
      const size_t npairs2 = static_cast<size_t>(Bt) * Bi;
      if (S_extra == nullptr) CXL_CUDA_CHECK(cudaMemsetAsync(S, 0, npairs2 * sizeof(float), stream));
      if (cnt != nullptr) CXL_CUDA_CHECK(cudaMemsetAsync(cnt, 0, npairs2 * sizeof(float), stream));
      if (overhead + q.num_kb * kABlockBytes + q.num_stages * q.stage_bytes + extra_smem > max_smem) q.num_stages -= 1;
      const int smem2 = overhead + q.num_kb * kABlockBytes + q.num_stages * q.stage_bytes + extra_smem;
      CXL_CUDA_CHECK(cudaFuncSetAttribute(maxsim_fwd_tc2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
      int ncl = q.num_units < clusters ? q.num_units : clusters;
      maxsim_fwd_tc2_kernel<false><<<2 * ncl, kThreads2, smem2, CXL_COUNTED(stream)>>>(tmA2, tmB2, q);
      CXL_CUDA_CHECK(cudaGetLastError());
      if (S_extra != nullptr) return launch_maxsim_finalize_slots(S, S_extra, cnt, Bt, Bi, T, 32, mean_mode, stream);
      return launch_maxsim_finalize(S, cnt, npairs2, T, mean_mode, stream);
    }
  }
  // ---- CTA-pair kernel with the text block streamed through the ring (SA): large D (text tile not resident) and / or
  //      images of several N tiles (P > 256) -- the cfg-5 geometry (P 576 = 3 x 192 columns, D 768) ----
  static const int pair_sa_env = getenv("CXL_K1_PAIR_STREAM") ? atoi(getenv("CXL_K1_PAIR_STREAM")) : 1;
  if (pair_env && pair_sa_env && p.ipt == 1 && dbg == nullptr && !force_stream_a && force_grid <= 0 &&
      p.n_tile % 16 == 0 && P > 128) {
    const int half_bytes = (p.n_tile / 2) * kBlockK * 2;
    const int stage_sa = kABlockBytes + round_up(half_bytes, 1024);
    const int st = (max_smem - overhead) / stage_sa;
    if (st >= 4) {
      FwdTcParams q = p;
      q.a_resident = 0;
      q.b_block_bytes = half_bytes;
      q.stage_bytes = stage_sa;
      q.num_stages = st > kMaxStages ? kMaxStages : st;
      q.num_m_tiles = (p.rows_a + 2 * kBlockM - 1) / (2 * kBlockM);      // 256-row pairs
      const int clusters = num_sms() / 2;
      const int n_img = (Bi + p.ipt - 1) / p.ipt;
      // units = (256-row pair, chunk of whole images), chunk-major: all clusters stream the same images at a time.
      // As few chunks as keep the static round-robin over clusters balanced ...
      int chunks = 1;
      for (; chunks < 64 && chunks < n_img; ++chunks) {
        const long long u = static_cast<long long>(chunks) * q.num_m_tiles;
        const long long waves = (u + clusters - 1) / clusters;
        if (static_cast<double>(u) >= 0.97 * static_cast<double>(waves * clusters)) break;
      }
      int img_per_unit = (n_img + chunks - 1) / chunks;
      // ... but no more images than stay in L2 while every wave of row pairs re-reads them, when there are several waves
      // at all.  The streamed text tiles of a wave (74 x 393 KB = 29 MB at D = 768) live in L2 next to the block and
      // each die keeps its own copy, so the block is small: 14 MB = 16 images at cfg-5.  One B200, [8192 x 8192] launch,
      // P 576 / D 768: no blocks 1020 TF/s, 4 images 1298, 8 1311, 16 1308-1313, 24 1306, 47 1286-1291, 96 1246
      // (profiles/r4z_k1_sa_block.txt)
      {
        const long long img_bytes = static_cast<long long>(p.p_stride) * D * 2;
        int blk = static_cast<int>((14LL << 20) / img_bytes);
        static const int blk_env = getenv("CXL_K1_IMG_BLOCK") ? atoi(getenv("CXL_K1_IMG_BLOCK")) : -1;
        if (blk_env >= 0) blk = blk_env;
        if (blk > 0 && img_per_unit > blk && q.num_m_tiles >= 4 * clusters) {
          const int nblk = (n_img + blk - 1) / blk;
          img_per_unit = (n_img + nblk - 1) / nblk;
        }
      }
      q.tiles_per_unit = img_per_unit * p.parts;
      chunks = (q.total_n_tiles + q.tiles_per_unit - 1) / q.tiles_per_unit;
      q.num_units = chunks * q.num_m_tiles;
      q.mean_mode = mean_mode;
      q.argmax_u16 = P > 256;
      q.S = S; q.cnt = cnt; q.argmax = argmax; q.ld_arg = ld_arg; q.dbg = nullptr;
      q.pack_g = 0;
      q.l2_hint = 0;
      CUtensorMap tmA2, tmB2;
      int rc2 = make_tmap(&tmA2, txt, q.rows_a, D, kBlockM, p.f16);
      if (rc2 != CXL_OK) return rc2;
      rc2 = make_tmap(&tmB2, img, static_cast<long long>(Bi) * q.p_stride, D, q.n_tile / 2, p.f16);
      if (rc2 != CXL_OK) return rc2;
      const size_t npairs2 = static_cast<size_t>(Bt) * Bi;
      if (S_extra == nullptr) CXL_CUDA_CHECK(cudaMemsetAsync(S, 0, npairs2 * sizeof(float), stream));
      if (cnt != nullptr) CXL_CUDA_CHECK(cudaMemsetAsync(cnt, 0, npairs2 * sizeof(float), stream));
      const int smem2 = overhead + q.num_stages * q.stage_bytes;
      CXL_CUDA_CHECK(cudaFuncSetAttribute(maxsim_fwd_tc2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
      const int ncl = q.num_units < clusters ? q.num_units : clusters;
      maxsim_fwd_tc2_kernel<true><<<2 * ncl, kThreads2, smem2, CXL_COUNTED(stream)>>>(tmA2, tmB2, q);
      CXL_CUDA_CHECK(cudaGetLastError());
      if (S_extra != nullptr) return launch_maxsim_finalize_slots(S, S_extra, cnt, Bt, Bi, T, 32, mean_mode, stream);
      return launch_maxsim_finalize(S, cnt, npairs2, T, mean_mode, stream);
    }
  }
  const int b_stage = round_up(p.b_block_bytes, 1024);
  // text tile resident if it leaves room for >= 3 ring stages; otherwise stream A with B
  p.a_resident = 1;
  int stages = (max_smem - overhead - p.num_kb * kABlockBytes) / b_stage;
  if (stages < 3 || force_stream_a) {
    p.a_resident = 0;
    p.stage_bytes = b_stage + kABlockBytes;
    stages = (max_smem - overhead) / p.stage_bytes;
  } else {
    p.stage_bytes = b_stage;
  }
  if (stages < 2) return CXL_ERR_UNSUPPORTED;
  p.num_stages = stages > kMaxStages ? kMaxStages : stages;
  const int smem_bytes = overhead + (p.a_resident ? p.num_kb * kABlockBytes : 0) + p.num_stages * p.stage_bytes;

  const int sms = num_sms();
  long long total = static_cast<long long>(p.total_n_tiles) * p.num_m_tiles;
  long long tpu = (total + sms * 8LL - 1) / (sms * 8LL);
  tpu = (tpu + p.num_m_tiles - 1) / p.num_m_tiles;       // n-tiles per unit so that units ~ 8 x SMs
  if (tpu > 256) tpu = 256;
  if (tpu < 1) tpu = 1;
  p.tiles_per_unit = round_up(static_cast<int>(tpu), p.parts);
  if (p.tiles_per_unit > p.total_n_tiles) p.tiles_per_unit = p.total_n_tiles;
  const int chunks = (p.total_n_tiles + p.tiles_per_unit - 1) / p.tiles_per_unit;
  p.num_units = chunks * p.num_m_tiles;
  p.mean_mode = mean_mode;
  p.argmax_u16 = P > 256;
  p.S = S; p.cnt = cnt; p.argmax = argmax; p.ld_arg = ld_arg; p.dbg = dbg;

  CUtensorMap tmA, tmB;
  int rc = make_tmap(&tmA, txt, p.rows_a, D, kBlockM, p.f16);
  if (rc != CXL_OK) return rc;
  rc = make_tmap(&tmB, img, static_cast<long long>(Bi) * p.p_stride, D, p.n_tile, p.f16);
  if (rc != CXL_OK) return rc;

  const size_t npairs = static_cast<size_t>(Bt) * Bi;
  if (S_extra == nullptr) CXL_CUDA_CHECK(cudaMemsetAsync(S, 0, npairs * sizeof(float), stream));
  if (cnt != nullptr) CXL_CUDA_CHECK(cudaMemsetAsync(cnt, 0, npairs * sizeof(float), stream));

  CXL_CUDA_CHECK(cudaFuncSetAttribute(maxsim_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  int grid = p.num_units < sms ? p.num_units : sms;
  if (force_grid > 0 && force_grid < grid) grid = force_grid;
  maxsim_fwd_tc_kernel<<<grid, kThreads, smem_bytes, CXL_COUNTED(stream)>>>(tmA, tmB, p);
  CXL_CUDA_CHECK(cudaGetLastError());

  if (S_extra != nullptr) return launch_maxsim_finalize_slots(S, S_extra, cnt, Bt, Bi, T, 32, mean_mode, stream);
  return launch_maxsim_finalize(S, cnt, npairs, T, mean_mode, stream);
}

// ------------------------------------------------------------------------------------------------
// Plain NT GEMM on the same pipeline: C[M,N] (f32) = A[M,K] * B[N,K]^T, 16-bit operands (bf16 / fp16), fp32 accumulate.
// Serves the global-embedding half of the losses (loss.py:252-253, 132-160): logits [b,B] = txt [b,D] x img [B,D]^T and
// the two gradient GEMMs of their backward.  M tile 128 rows, N tile 256 rows of B, K in 64-wide blocks; the A tile
// stays resident when K is small (the logits GEMM), is streamed with B otherwise (the gradient GEMMs, K = batch).
int gemm_nt_tc(const void* A, const void* B, int M, int N, int K, float* C, long long ldc, int f16, cudaStream_t stream) {
  if (K % kBlockK != 0 || K < kBlockK || M <= 0 || N <= 0 || ldc < N) return CXL_ERR_UNSUPPORTED;
  FwdTcParams p;
  memset(&p, 0, sizeof(p));
  p.Bt = M; p.T = 1; p.rows_a = M; p.D = K; p.num_kb = K / kBlockK;
  p.P = 256; p.p_stride = 256; p.ipt = 1; p.parts = 1; p.part_cols = 256; p.n_tile = 256;
  p.Bi = (N + 255) / 256;
  p.f16 = f16 ? 1 : 0;
  p.total_n_tiles = p.Bi;
  p.num_m_tiles = (M + kBlockM - 1) / kBlockM;
  p.b_block_bytes = p.n_tile * kBlockK * 2;
  p.raw_c = C; p.raw_ldc = ldc; p.raw_n = N;
  int dev = 0, max_smem = 0;
  CXL_CUDA_CHECK(cudaGetDevice(&dev));
  CXL_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const int overhead = 1024 + static_cast<int>(sizeof(FwdTcBarriers)) + 64;
  const int b_stage = p.b_block_bytes;                    // 32 KB, already a multiple of 1024
  p.a_resident = 1;
  int stages = (max_smem - overhead - p.num_kb * kABlockBytes) / b_stage;
  if (stages < 3) {
    p.a_resident = 0;
    p.stage_bytes = b_stage + kABlockBytes;
    stages = (max_smem - overhead) / p.stage_bytes;
  } else {
    p.stage_bytes = b_stage;
  }
  p.num_stages = stages > kMaxStages ? kMaxStages : stages;
  const int smem_bytes = overhead + (p.a_resident ? p.num_kb * kABlockBytes : 0) + p.num_stages * p.stage_bytes;
  const int sms = num_sms();
  // units = (m tile, chunk of n tiles); whole sweeps per unit when there are enough m tiles, else split the sweep
  long long tpu = p.total_n_tiles;
  while (tpu > 1 && static_cast<long long>(p.num_m_tiles) * ((p.total_n_tiles + tpu - 1) / tpu) < 2LL * sms) tpu = (tpu + 1) / 2;
  p.tiles_per_unit = static_cast<int>(tpu);
  const int chunks = (p.total_n_tiles + p.tiles_per_unit - 1) / p.tiles_per_unit;
  p.num_units = chunks * p.num_m_tiles;
  CUtensorMap tmA, tmB;
  int rc = make_tmap(&tmA, A, M, K, kBlockM, p.f16);
  if (rc != CXL_OK) return rc;
  rc = make_tmap(&tmB, B, N, K, p.n_tile, p.f16);
  if (rc != CXL_OK) return rc;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(maxsim_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  const int grid = p.num_units < sms ? p.num_units : sms;
  maxsim_fwd_tc_kernel<<<grid, kThreads, smem_bytes, CXL_COUNTED(stream)>>>(tmA, tmB, p);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

}  // namespace cxl
