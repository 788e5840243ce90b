// K2 (fast path, second generation): MaxSim backward from the saved argmax, bf16 operands, fp32 register
// accumulators, mixed-precision FMAs (fma.rn.f32.bf16 -> SASS FHFMA.BF16: bf16 x bf16 products added into fp32
// with one rounding, no unpack instructions).
//
// Replaces autograd of loss.py:34-44 in the reference (max -> scatter into zeros, then two bmm against a one-hot
// dS); under autocast the reference runs those bmm with bf16 dS, so rounding the pair weight w to bf16 here is
// the reference's own precision:
//   dTxt[m,n,:] = sum_k w[m,k] * img[k, q*(m,k,n), :]
//   dImg[k,q,:] = sum_{m,n : q*(m,k,n) = q} w[m,k] * txt[m,n,:]
// Both are sparse products with Bt*Bi*T "items" of D multiply-adds each and no operand reuse, so the floor is
// the shared-memory read of one D-wide bf16 row per item (D*2 bytes at 128 B/clk/SM).
//
//   dTxt: CTA = (tile of 32*NWC text-token rows, 64-wide D slice).  The [P,64] slab of image k is staged once
//         per CTA by TMA; a QUARTER warp handles one row (8 lanes x 8 channels = one 128 B slab row per
//         LDS.128 phase, conflict-free), so one warp instruction serves four items.  Per-item entries
//         (bf16 weight | slab row offset) are prepared by the consuming warp itself two images ahead.
//   dImg: CTA = (image k, DS-wide D slice); consumer warp g owns patches [g*G, g*G+G) in registers.  Text-row
//         tiles stream through a TMA ring (multicast to the C images of a cluster); producer warps bucket the
//         tile's items into per-consumer queues (match.any ranks, no atomics), and the consumer dispatches on
//         the patch with one brx.idx into straight-line FMA groups.
#include "common.cuh"

namespace cxl {

struct BwdV2Params {
  const float* w;       // [Bt, ld_pairs]  per-pair weight dL/dS / cnt
  const void* argmax;   // [k_off + Bi][ld_arg]  image-major
  long long ld_arg;
  int Bt, Bi, T, P, D, p_stride;
  int rows;             // Bt*T
  int ld_pairs, k_off, argmax_u16;
  float* dtxt;          // [rows, D]
  float* dimg;          // [Bi*P, D]
  int accumulate_img;
  int stages;
  int stage_bytes;      // bytes of one ring stage
  int tile_bytes;       // bytes TMA delivers into a stage
  int nwc;              // dImg: consumer warps
};

struct __align__(8) V2Barriers {
  uint64_t full[8];
  uint64_t empty[8];
  uint64_t ready[8];    // cluster leader only: every CTA of the cluster has freed + armed the stage
  uint64_t qfull[8];    // dImg: the producers have bucketed the stage's items
};

__device__ __forceinline__ uint32_t pack_entry(float w, uint32_t lo16) {
  return (static_cast<uint32_t>(__bfloat16_as_ushort(__float2bfloat16_rn(w))) << 16) | lo16;
}

// acc[0..7] += bf16x8(v) * bf16(high half of e)
__device__ __forceinline__ void fhfma8(float (&a)[8], const uint4& v, uint32_t e) {
  asm("{\n\t.reg .b16 x0, x1, x2, x3, x4, x5, x6, x7, wl, wh;\n\t"
      "mov.b32 {x0, x1}, %8;\n\t mov.b32 {x2, x3}, %9;\n\t mov.b32 {x4, x5}, %10;\n\t mov.b32 {x6, x7}, %11;\n\t"
      "mov.b32 {wl, wh}, %12;\n\t"
      "fma.rn.f32.bf16 %0, x0, wh, %0;\n\t fma.rn.f32.bf16 %1, x1, wh, %1;\n\t"
      "fma.rn.f32.bf16 %2, x2, wh, %2;\n\t fma.rn.f32.bf16 %3, x3, wh, %3;\n\t"
      "fma.rn.f32.bf16 %4, x4, wh, %4;\n\t fma.rn.f32.bf16 %5, x5, wh, %5;\n\t"
      "fma.rn.f32.bf16 %6, x6, wh, %6;\n\t fma.rn.f32.bf16 %7, x7, wh, %7;\n\t}"
      : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7])
      : "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(e));
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// ------------------------------------------------------------------------------------------------
// dTxt
// ------------------------------------------------------------------------------------------------
// Warps 0..NWC-1 consume (32 rows each), warp NWC issues the TMA loads.  A stage holds the [P,64] slab of one
// image AND the CTA's slice of that image's argmax row (ROWS contiguous elements, one bulk copy), so no
// consumer ever waits on a global load for its indices.  Lane = 8*qw + l8: quarter qw handles rows 4*jj + qw
// (jj = 0..7) of its warp, l8 selects the 8-channel chunk of the 64-wide slice.
template <int NWC, bool U16>
__global__ void __launch_bounds__((NWC + 1) * 32, 1)
maxsim_bwd_dtxt_v2_kernel(const __grid_constant__ CUtensorMap tmImg, const BwdV2Params p) {
  constexpr int DS = 64;
  constexpr int ROWS = NWC * 32;
  constexpr int ES = U16 ? 2 : 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  V2Barriers* bars = reinterpret_cast<V2Barriers*>(smem + p.stages * p.stage_bytes);
  uint8_t* ent_area = reinterpret_cast<uint8_t*>(bars + 1);          // [NWC][2][32] packed entries

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int r0 = blockIdx.x * ROWS;
  const int d0 = blockIdx.y * DS;
  const int Bi = p.Bi;
  const int stages = p.stages;
  const uint32_t stage_bytes = p.stage_bytes;
  const uint32_t smem0 = smem_u32(smem);
  const uint32_t full0 = smem_u32(&bars->full[0]);
  const uint32_t empty0 = smem_u32(&bars->empty[0]);

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, NWC);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmImg);
  }
  __syncthreads();

  if (warp == NWC) {
    if (lane == 0) {
      // argmax elements of this CTA's rows (clamped to the padded row length; both multiples of 16 bytes)
      const long long avail = p.ld_arg - r0;
      const uint32_t qbytes = static_cast<uint32_t>((avail < ROWS ? avail : ROWS) * ES);
      const uint8_t* qsrc = reinterpret_cast<const uint8_t*>(p.argmax) +
                            (static_cast<size_t>(p.k_off) * p.ld_arg + r0) * ES;
      const size_t qstep = static_cast<size_t>(p.ld_arg) * ES;
      int stage = 0;
      uint32_t phase = 0;
      for (int k = 0; k < Bi; ++k) {
        mbar_wait(empty0 + 8 * stage, phase ^ 1);
        const uint32_t st = smem0 + stage * stage_bytes;
        const uint32_t full = full0 + 8 * stage;
        mbar_arrive_expect_tx(full, p.tile_bytes + qbytes);
        for (int q0 = 0; q0 < p.P; q0 += 256) tma_load_2d(st + q0 * DS * 2, &tmImg, full, d0, k * p.p_stride + q0);
        bulk_load_1d(st + p.tile_bytes, qsrc, qbytes, full);
        qsrc += qstep;
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
    return;
  }

  float acc[8][8];
#pragma unroll
  for (int j = 0; j < 8; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;

  // entry production: lane <-> row (warp*32 + lane); tail rows are clamped and never stored
  const int my_row = min(r0 + warp * 32 + lane, p.rows - 1);
  const float* wptr = p.w + static_cast<size_t>(my_row / p.T) * p.ld_pairs + p.k_off;
  const uint32_t q_off = p.tile_bytes + (warp * 32 + lane) * ES;          // this lane's index inside a stage
  const uint32_t q_max = static_cast<uint32_t>(p.P - 1);
  const int qw = lane >> 3;
  // entry strip of this warp: [2][4 quarters][8 rows]; row j = 4*jj + qw lives at word qw*8 + jj
  const uint32_t ent_base = smem_u32(ent_area) + warp * 256;
  const uint32_t ent_st = ent_base + ((lane & 3) * 8 + (lane >> 2)) * 4;   // where this lane's produced entry goes
  const uint32_t ent_ld = ent_base + qw * 32;                              // the 8 entries this lane consumes
  const uint32_t stl0 = smem0 + (lane & 7) * 16;

  auto ldq_smem = [&](uint32_t addr) -> uint32_t {
    uint32_t q;
    if (U16) asm volatile("ld.shared.u16 %0, [%1];" : "=r"(q) : "r"(addr));
    else asm volatile("ld.shared.u8 %0, [%1];" : "=r"(q) : "r"(addr));
    return min(q, q_max);                                  // stale bytes of tail rows must stay inside the slab
  };

  // pair weights of this lane's row for images k .. k+3 (prefetch ring; consecutive k are contiguous)
  float wr[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) wr[i] = __ldg(wptr + min(i, Bi - 1));

  // entries of image 0
  mbar_wait(full0, 0);
  sts32(ent_st, pack_entry(wr[0], ldq_smem(smem0 + q_off) * 8));   // slab row offset in 16-byte units (q*128/16)
  __syncwarp();

  int stage = 0;                 // stage of image k
  uint32_t phase = 0;
  for (int k0 = 0; k0 < Bi; k0 += 4) {
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int k = k0 + b;
      if (k < Bi) {
        // entries of image k (buffer b & 1) were stored one iteration ago
        const uint4 ea = lds128(ent_ld + (b & 1) * 128);
        const uint4 eb = lds128(ent_ld + (b & 1) * 128 + 16);
        // stage the entries of image k+1 into the other buffer (its stage has normally landed already)
        int nstage = stage + 1;
        uint32_t nphase = phase;
        if (nstage == stages) { nstage = 0; nphase ^= 1; }
        if (k + 1 < Bi) {
          mbar_wait(full0 + 8 * nstage, nphase);
          sts32(ent_st + ((b & 1) ^ 1) * 128,
                pack_entry(wr[(b + 1) & 3], ldq_smem(smem0 + nstage * stage_bytes + q_off) * 8));
        }
        wr[b] = __ldg(wptr + min(k + 4, Bi - 1));
        const uint32_t stl = stl0 + stage * stage_bytes;
        const uint32_t e[8] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y, eb.z, eb.w};
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
          const uint4 v = lds128(stl + ((e[jj] & 0xFFFFu) << 4));
          fhfma8(acc[jj], v, e[jj]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * stage);
        stage = nstage;
        phase = nphase;
      }
    }
  }
#pragma unroll
  for (int jj = 0; jj < 8; ++jj) {
    const int r = r0 + warp * 32 + 4 * jj + qw;
    if (r < p.rows) {
      float* o = p.dtxt + static_cast<size_t>(r) * p.D + d0 + (lane & 7) * 8;
      *reinterpret_cast<float4*>(o) = make_float4(acc[jj][0], acc[jj][1], acc[jj][2], acc[jj][3]);
      *reinterpret_cast<float4*>(o + 4) = make_float4(acc[jj][4], acc[jj][5], acc[jj][6], acc[jj][7]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// dImg
// ------------------------------------------------------------------------------------------------
// Patch dispatch: one brx.idx into straight-line FHFMA groups (registers cannot be indexed dynamically).
template <int G, int VEC> struct PatchFma;

#define CXL_F4(L, a, b, c, d)                                                                              \
  "C" #L ": fma.rn.f32.bf16 %" #a ", x0, wh, %" #a "; fma.rn.f32.bf16 %" #b ", x1, wh, %" #b               \
  "; fma.rn.f32.bf16 %" #c ", x2, wh, %" #c "; fma.rn.f32.bf16 %" #d ", x3, wh, %" #d "; bra.uni DONE;\n\t"
#define CXL_F2(L, a, b) \
  "C" #L ": fma.rn.f32.bf16 %" #a ", x0, wh, %" #a "; fma.rn.f32.bf16 %" #b ", x1, wh, %" #b "; bra.uni DONE;\n\t"
#define CXL_A4(j) "+f"(acc[j][0]), "+f"(acc[j][1]), "+f"(acc[j][2]), "+f"(acc[j][3])
#define CXL_A2(j) "+f"(acc[j][0]), "+f"(acc[j][1])

template <> struct PatchFma<8, 4> {
  // v[0..1]: 4 bf16 channels of the text row, e: entry (bf16 weight in the high half), ql: patch within the group
  __device__ static __forceinline__ void run(float (&acc)[8][4], const uint32_t (&v)[2], uint32_t e, uint32_t ql) {
    asm volatile(
        "{\n\t.reg .b16 x0, x1, x2, x3, wl, wh;\n\t"
        "mov.b32 {x0, x1}, %33;\n\t mov.b32 {x2, x3}, %34;\n\t mov.b32 {wl, wh}, %35;\n\t"
        "DT: .branchtargets C0, C1, C2, C3, C4, C5, C6, C7, DONE;\n\t"
        "brx.idx.uni %32, DT;\n\t"
        CXL_F4(0, 0, 1, 2, 3) CXL_F4(1, 4, 5, 6, 7) CXL_F4(2, 8, 9, 10, 11) CXL_F4(3, 12, 13, 14, 15)
        CXL_F4(4, 16, 17, 18, 19) CXL_F4(5, 20, 21, 22, 23) CXL_F4(6, 24, 25, 26, 27) CXL_F4(7, 28, 29, 30, 31)
        "DONE:\n\t}"
        : CXL_A4(0), CXL_A4(1), CXL_A4(2), CXL_A4(3), CXL_A4(4), CXL_A4(5), CXL_A4(6), CXL_A4(7)
        : "r"(ql), "r"(v[0]), "r"(v[1]), "r"(e));
  }
};

template <> struct PatchFma<24, 2> {
  __device__ static __forceinline__ void run(float (&acc)[24][2], const uint32_t (&v)[1], uint32_t e, uint32_t ql) {
    asm volatile(
        "{\n\t.reg .b16 x0, x1, wl, wh;\n\t"
        "mov.b32 {x0, x1}, %49;\n\t mov.b32 {wl, wh}, %50;\n\t"
        "DT: .branchtargets C0, C1, C2, C3, C4, C5, C6, C7, C8, C9, C10, C11, C12, C13, C14, C15, C16, C17, C18, C19, "
        "C20, C21, C22, C23, DONE;\n\t"
        "brx.idx.uni %48, DT;\n\t"
        CXL_F2(0, 0, 1) CXL_F2(1, 2, 3) CXL_F2(2, 4, 5) CXL_F2(3, 6, 7) CXL_F2(4, 8, 9) CXL_F2(5, 10, 11)
        CXL_F2(6, 12, 13) CXL_F2(7, 14, 15) CXL_F2(8, 16, 17) CXL_F2(9, 18, 19) CXL_F2(10, 20, 21) CXL_F2(11, 22, 23)
        CXL_F2(12, 24, 25) CXL_F2(13, 26, 27) CXL_F2(14, 28, 29) CXL_F2(15, 30, 31) CXL_F2(16, 32, 33)
        CXL_F2(17, 34, 35) CXL_F2(18, 36, 37) CXL_F2(19, 38, 39) CXL_F2(20, 40, 41) CXL_F2(21, 42, 43)
        CXL_F2(22, 44, 45) CXL_F2(23, 46, 47)
        "DONE:\n\t}"
        : CXL_A2(0), CXL_A2(1), CXL_A2(2), CXL_A2(3), CXL_A2(4), CXL_A2(5), CXL_A2(6), CXL_A2(7), CXL_A2(8), CXL_A2(9),
          CXL_A2(10), CXL_A2(11), CXL_A2(12), CXL_A2(13), CXL_A2(14), CXL_A2(15), CXL_A2(16), CXL_A2(17), CXL_A2(18),
          CXL_A2(19), CXL_A2(20), CXL_A2(21), CXL_A2(22), CXL_A2(23)
        : "r"(ql), "r"(v[0]), "r"(e));
  }
};
#undef CXL_F4
#undef CXL_F2
#undef CXL_A4
#undef CXL_A2

template <int VEC> struct RowLoad;
template <> struct RowLoad<4> {
  __device__ static __forceinline__ void ld(uint32_t addr, uint32_t (&v)[2]) {
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v[0]), "=r"(v[1]) : "r"(addr));
  }
};
template <> struct RowLoad<2> {
  __device__ static __forceinline__ void ld(uint32_t addr, uint32_t (&v)[1]) {
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v[0]) : "r"(addr));
  }
};

// Warps [0, nwc): consumers; [nwc, nwc+4): producers (one tile row per lane); nwc+4: TMA issue.
// Stage = [tile RT x DS bf16][argmax of the tile's rows RT x ES][queues nwc x RT x u32][counts nwc x 4 x u32].
// Entry = bf16(w) << 16 | row-in-tile << 5 | patch-in-group  (patch-in-group == G: padding, skipped by the dispatch).
// The host guarantees T >= 8: a tile's RT = 128 rows then span at most 17 texts, i.e. at most two 32-text batches
// of the per-lane pair-weight cache.
template <int DS, int G, int C, int MAXW, bool U16>
__global__ void __launch_bounds__(MAXW * 32, 1)
maxsim_bwd_dimg_v2_kernel(const __grid_constant__ CUtensorMap tmTxt, const BwdV2Params p) {
  constexpr int VEC = DS / 32;
  constexpr int RT = 128;
  constexpr int NPW = 4;
  constexpr int ES = U16 ? 2 : 1;
  constexpr int OFF_SHIFT = (DS == 128) ? 3 : 2;       // (row << 5) -> row * DS * 2 bytes
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  V2Barriers* bars = reinterpret_cast<V2Barriers*>(smem + p.stages * p.stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nwc = p.nwc;
  const int k = blockIdx.x;
  const bool active = k < p.Bi;                          // cluster padding CTAs run the pipeline but own nothing
  const int kk = active ? k : p.Bi - 1;
  const int d0 = blockIdx.y * DS;
  const int rows = p.rows;
  const int num_tiles = (rows + RT - 1) / RT;
  const int stages = p.stages;
  const uint32_t stage_bytes = p.stage_bytes;
  const uint32_t tile_bytes = p.tile_bytes;
  const uint32_t a_area = tile_bytes;                    // offsets inside a stage
  const uint32_t q_area = a_area + RT * ES;
  const uint32_t c_area = q_area + nwc * RT * 4;
  const uint32_t smem0 = smem_u32(smem);
  const uint32_t full0 = smem_u32(&bars->full[0]);
  const uint32_t empty0 = smem_u32(&bars->empty[0]);
  const uint32_t qfull0 = smem_u32(&bars->qfull[0]);
  const uint32_t crank = (C > 1) ? cluster_ctarank() : 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, nwc);
      mbar_init(qfull0 + 8 * s, NPW);
      mbar_init(smem_u32(&bars->ready[s]), C);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmTxt);
  }
  __syncthreads();
  if (C > 1) cluster_sync_all();

  if (warp == nwc + NPW) {
    // ===================== TMA issue: text-row tile (multicast) + this image's argmax bytes =====================
    if (lane == 0) {
      const uint8_t* qsrc = reinterpret_cast<const uint8_t*>(p.argmax) + static_cast<size_t>(p.k_off + kk) * p.ld_arg * ES;
      int stage = 0;
      uint32_t phase = 0;
      for (int t = 0; t < num_tiles; ++t) {
        mbar_wait(empty0 + 8 * stage, phase ^ 1);
        const uint32_t st = smem0 + stage * stage_bytes;
        const uint32_t full = full0 + 8 * stage;
        const long long avail = p.ld_arg - static_cast<long long>(t) * RT;
        const uint32_t qbytes = static_cast<uint32_t>((avail < RT ? avail : RT) * ES);
        mbar_arrive_expect_tx(full, tile_bytes + qbytes);
        bulk_load_1d(st + a_area, qsrc + static_cast<size_t>(t) * RT * ES, qbytes, full);
        if (C == 1) {
          tma_load_2d(st, &tmTxt, full, d0, t * RT);     // rows past the end are zero-filled
        } else {
          const uint32_t ready = smem_u32(&bars->ready[stage]);
          mbar_arrive_cluster(mapa_shared(ready, 0));
          if (crank == 0) {
            mbar_wait_cluster(ready, phase);             // every CTA of the cluster freed and armed this stage
            tma_load_2d_mcast(st, &tmTxt, full, d0, t * RT, static_cast<uint16_t>((1u << C) - 1));
          }
        }
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= nwc) {
    // ===================== producers: bucket the tile's items by consumer warp =====================
    const int pw = warp - nwc;
    const int slot = pw * 32 + lane;
    const int T = p.T;
    const float* wcol = p.w + p.k_off + kk;
    auto load_w = [&](int m) { return m < p.Bt ? __ldg(wcol + static_cast<size_t>(m) * p.ld_pairs) : 0.f; };
    int m_base = 0;
    float w_cur = load_w(lane), w_nxt = load_w(32 + lane);
    int m_row = slot / T;                                // text of this lane's row in the current tile
    int n_row = slot - m_row * T;
    int m_first = 0, n_first = 0;                        // text / token of the tile's first row (warp-uniform)
    const uint32_t lt_mask = (1u << lane) - 1;
    const uint32_t q_max = static_cast<uint32_t>(p.P - 1);
    int stage = 0;
    uint32_t phase = 0;
    for (int t = 0; t < num_tiles; ++t) {
      while (m_first >= m_base + 32) {                   // keep the tile's first text inside w_cur
        m_base += 32;
        w_cur = w_nxt;
        w_nxt = load_w(m_base + 32 + lane);
      }
      const int mi = min(m_row, p.Bt - 1) - m_base;      // 0 .. 63
      const float wa = __shfl_sync(0xffffffffu, w_cur, mi & 31);
      const float wb = __shfl_sync(0xffffffffu, w_nxt, mi & 31);
      const float wv = mi < 32 ? wa : wb;
      const bool valid = t * RT + slot < rows;
      const uint32_t st = smem0 + stage * stage_bytes;
      mbar_wait(full0 + 8 * stage, phase);               // tile + argmax bytes have landed
      uint32_t q;
      if (U16) asm volatile("ld.shared.u16 %0, [%1];" : "=r"(q) : "r"(st + a_area + slot * ES));
      else asm volatile("ld.shared.u8 %0, [%1];" : "=r"(q) : "r"(st + a_area + slot * ES));
      q = min(q, q_max);
      const uint32_t g = (G == 8) ? (q >> 3) : ((q * 2731u) >> 16);
      const uint32_t ql = q - g * G;
      const uint32_t mask = __match_any_sync(0xffffffffu, valid ? g : 0xFFFFu);
      const uint32_t rank = __popc(mask & lt_mask);
      if (lane < nwc) sts32(st + c_area + (lane * 4 + pw) * 4, 0u);
      __syncwarp();
      if (valid && rank == 0) sts32(st + c_area + (g * 4 + pw) * 4, __popc(mask));
      named_bar_sync(1, NPW * 32);
      if (valid) {
        uint32_t base = rank;
        for (int pp = 0; pp < pw; ++pp) {
          uint32_t c;
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(c) : "r"(st + c_area + (g * 4 + pp) * 4));
          base += c;
        }
        sts32(st + q_area + (g * RT + base) * 4, pack_entry(wv, (static_cast<uint32_t>(slot) << 5) | ql));
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(qfull0 + 8 * stage);
      n_row += RT;
      while (n_row >= T) { n_row -= T; ++m_row; }
      n_first += RT;
      while (n_first >= T) { n_first -= T; ++m_first; }
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  } else {
    // ===================== consumer warp `warp` owns patches [warp*G, warp*G + G) =====================
    float acc[G][VEC];
#pragma unroll
    for (int j = 0; j < G; ++j)
#pragma unroll
      for (int i = 0; i < VEC; ++i) acc[j][i] = 0.f;
    const uint32_t lane_off = lane * VEC * 2;
    const uint32_t my_q = q_area + warp * (RT * 4);
    const uint32_t my_c = c_area + warp * 16;
    int stage = 0;
    uint32_t phase = 0;
    for (int t = 0; t < num_tiles; ++t) {
      const uint32_t st = smem0 + stage * stage_bytes;
      mbar_wait(qfull0 + 8 * stage, phase);              // queues ready (the producers saw the data land)
      mbar_wait(full0 + 8 * stage, phase);               // observe the TMA completion ourselves (already done)
      const uint4 c4 = lds128(st + my_c);
      const int n = static_cast<int>(c4.x + c4.y + c4.z + c4.w);
      const uint32_t qa = st + my_q;
      const uint32_t stl = st + lane_off;
      for (int i = 0; i < n; i += 4) {
        const uint4 e4 = lds128(qa + i * 4);
        const uint32_t e[4] = {e4.x, e4.y, e4.z, e4.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t eu = (u == 0 || i + u < n) ? e[u] : static_cast<uint32_t>(G);   // row 0, weight +0, skipped
          uint32_t v[VEC / 2];
          RowLoad<VEC>::ld(stl + ((eu & 0xFFE0u) << OFF_SHIFT), v);
          PatchFma<G, VEC>::run(acc, v, eu, eu & 31u);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * stage);
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
    if (active) {
#pragma unroll
      for (int j = 0; j < G; ++j) {
        const int q = warp * G + j;
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
  if (C > 1) cluster_sync_all();
}

// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFnV2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFnV2 encode_fn_v2() {
  static EncodeTiledFnV2 fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFnV2>(ptr);
  }
  return fn;
}

// [rows, D] bf16 row-major, box = box_cols x box_rows, no swizzle (rows land densely, box_cols*2 bytes apart)
static int make_tmap_plain(CUtensorMap* out, const void* base, long long rows, int D, int box_cols, int box_rows) {
  EncodeTiledFnV2 fn = encode_fn_v2();
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

template <int NWC, bool U16>
static int launch_dtxt_v2(const void* img, BwdV2Params p, int max_smem, cudaStream_t stream) {
  constexpr int DS = 64;
  constexpr int ROWS = NWC * 32;
  // the TMA box covers min(P,256) slab rows; P > 256 takes several boxes (the last one may read into the next
  // image or past the end -> zero fill; those rows are never gathered)
  const int box_rows = p.P < 256 ? p.P : 256;
  const int slab_rows = round_up_i(p.P, box_rows);
  p.tile_bytes = slab_rows * DS * 2;
  p.stage_bytes = round_up_i(p.tile_bytes + ROWS * (U16 ? 2 : 1), 128);
  const int overhead = 128 + static_cast<int>(sizeof(V2Barriers)) + NWC * 256;
  int stages = (max_smem - overhead) / p.stage_bytes;
  if (stages < 2) return CXL_ERR_UNSUPPORTED;
  p.stages = stages > 6 ? 6 : stages;
  const int smem_bytes = overhead + p.stages * p.stage_bytes;
  CUtensorMap tm;
  int rc = make_tmap_plain(&tm, img, static_cast<long long>(p.Bi) * p.p_stride, p.D, DS, box_rows);
  if (rc != CXL_OK) return rc;
  auto kern = maxsim_bwd_dtxt_v2_kernel<NWC, U16>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  dim3 grid((p.rows + ROWS - 1) / ROWS, p.D / DS);
  kern<<<grid, (NWC + 1) * 32, smem_bytes, stream>>>(tm, p);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

template <int DS, int G, int C, int MAXW, bool U16>
static int launch_dimg_v2(const void* txt, BwdV2Params p, int max_smem, cudaStream_t stream) {
  constexpr int RT = 128;
  p.nwc = (p.P + G - 1) / G;
  if (p.nwc + 5 > MAXW) return CXL_ERR_UNSUPPORTED;
  p.tile_bytes = RT * DS * 2;
  p.stage_bytes = round_up_i(p.tile_bytes + RT * (U16 ? 2 : 1) + p.nwc * RT * 4 + p.nwc * 16, 128);
  const int overhead = 128 + static_cast<int>(sizeof(V2Barriers));
  int stages = (max_smem - overhead) / p.stage_bytes;
  if (stages < 2) return CXL_ERR_UNSUPPORTED;
  p.stages = stages > 6 ? 6 : stages;
  const int smem_bytes = overhead + p.stages * p.stage_bytes;
  CUtensorMap tm;
  int rc = make_tmap_plain(&tm, txt, p.rows, p.D, DS, RT);
  if (rc != CXL_OK) return rc;
  auto kern = maxsim_bwd_dimg_v2_kernel<DS, G, C, MAXW, U16>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(round_up_i(p.Bi, C), p.D / DS);
  cfg.blockDim = dim3((p.nwc + 5) * 32);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CXL_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tm, p));
  return CXL_OK;
}

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return (s != nullptr && *s != 0) ? atoi(s) : dflt;
}

// returns CXL_ERR_UNSUPPORTED when the shape is outside the fast path (caller falls back to the generic kernels)
int maxsim_bwd_stream(const void* txt, const void* img, const void* argmax, long long ld_arg, const float* w, int Bt,
                      int Bi, int T, int P, int D, int p_stride, int ld_pairs, int k_off, float* dtxt, float* dimg,
                      int accumulate_img, cudaStream_t stream) {
  if (D % 128 != 0 || P > 576 || T < 8) return CXL_ERR_UNSUPPORTED;
  if (static_cast<long long>(Bt) * T >= (1LL << 31)) return CXL_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(txt) & 15) || (reinterpret_cast<uintptr_t>(img) & 15)) return CXL_ERR_UNSUPPORTED;
  BwdV2Params p;
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
    rc = P > 256 ? launch_dtxt_v2<15, true>(img, p, max_smem, stream) : launch_dtxt_v2<15, false>(img, p, max_smem, stream);
    if (rc != CXL_OK) return rc;
  }
  if (dimg != nullptr) {
    static const int cluster = env_int("CXL_DIMG_CLUSTER", 2);
    if (P <= 25 * 8) {
      if (cluster >= 4) rc = launch_dimg_v2<128, 8, 4, 30, false>(txt, p, max_smem, stream);
      else if (cluster >= 2) rc = launch_dimg_v2<128, 8, 2, 30, false>(txt, p, max_smem, stream);
      else rc = launch_dimg_v2<128, 8, 1, 30, false>(txt, p, max_smem, stream);
    } else if (P <= 256) {
      if (cluster >= 2) rc = launch_dimg_v2<64, 24, 2, 29, false>(txt, p, max_smem, stream);
      else rc = launch_dimg_v2<64, 24, 1, 29, false>(txt, p, max_smem, stream);
    } else {
      if (cluster >= 2) rc = launch_dimg_v2<64, 24, 2, 29, true>(txt, p, max_smem, stream);
      else rc = launch_dimg_v2<64, 24, 1, 29, true>(txt, p, max_smem, stream);
    }
    if (rc != CXL_OK) return rc;
  }
  return CXL_OK;
}

}  // namespace cxl
