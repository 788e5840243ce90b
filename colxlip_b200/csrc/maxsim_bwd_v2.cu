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
//         (slab row offset | bf16 weight) are prepared by the consuming warp itself one image ahead.
//   dImg: also a gather with static accumulators: a pre-pass sorts every 512-row super-tile of text rows by
//         argmax patch (per image), then CTA = (2 images, 64-wide D slice) streams the super-tiles + the list
//         blocks through a TMA ring and each quarter warp walks the row lists of the patches it owns.
#include "bwd_v2_common.cuh"

namespace cxl {

struct BwdV2Params {
  const float* w;       // [Bt, ld_pairs]  per-pair weight dL/dS / cnt
  const void* argmax;   // [k_off + Bi][ld_arg]  image-major (already advanced to this call's first text row)
  long long ld_arg;
  long long arg_avail;  // elements of an argmax row from this call's first text row to the end of the pitch
  int Bt, Bi, T, P, D, p_stride;
  int rows;             // Bt*T
  int ld_pairs, k_off, argmax_u16;
  float* dtxt;          // [rows, D]
  float* dimg;          // [Bi*P, D]
  int accumulate_img;
  int accumulate_txt;   // dTxt: start from what is already there (image blocks after the first)
  int img_k0;           // dTxt: first image of this launch inside the image operand (TMA row coordinate)
  int stages;
  int stage_bytes;      // bytes of one ring stage
  int tile_bytes;       // bytes TMA delivers into a stage
  int slab_box_rows;    // dTxt: slab rows per TMA box (P when P <= 256, else P split evenly into <= 256-row boxes)
  int slab_n_box;       // dTxt: TMA boxes per slab (even when row-tile pairs share them by multicast)
  // dImg (one chunk of images at a time)
  const uint8_t* lists; // [chunk_n][n_st][block_bytes] row lists of the chunk (maxsim_bwd_dimg_lists_kernel)
  int n_st;             // 512-row super-tiles
  int hdr_words, block_bytes;
  int chunk_k0;         // first image of the chunk (index into dimg)
  int chunk_n;          // images in the chunk
  int f16;              // operands (and the 16-bit pair weights) are fp16 instead of bf16
  const float* w_scale; // device [s, 1/s] or null: power of two applied to the pair weights before the 16-bit rounding
                        // (keeps fp16 weights in range whatever the loss scale), undone exactly when results are stored
};

struct __align__(8) V2Barriers {
  uint64_t full[8];
  uint64_t empty[8];
};


// ------------------------------------------------------------------------------------------------
// dTxt
// ------------------------------------------------------------------------------------------------
// Warps 0..NWC-1 consume (32 rows each), warp NWC issues the TMA loads.  A stage holds the [P,64] slab of one
// image AND the CTA's slice of that image's argmax row (ROWS contiguous elements, one bulk copy), so no
// consumer ever waits on a global load for its indices.  Lane = 8*qw + l8: quarter qw handles rows 4*jj + qw
// (jj = 0..7) of its warp, l8 selects the 8-channel chunk of the 64-wide slice.
// MC = true: clusters of two row-tile CTAs of the same D slice.  Both walk the same images in the same order, so each
// CTA loads HALF of every slab's TMA boxes and multicasts them into both CTAs' stages (L2 -> SM traffic of the slabs
// halves; at P = 576 the kernel is bound by exactly that).  A stage may be overwritten only when the consumers of BOTH
// CTAs have released it: every consumer warp arrives on both CTAs' empty barriers (count 2 * NWC).  The argmax slice
// of a stage stays private.  A CTA past the last row tile (odd number of tiles) takes part in the protocol and stores
// nothing.
template <int NWC, bool U16, bool F16, bool MC>
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

  const uint32_t crank = MC ? cluster_ctarank() : 0u;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, MC ? 2 * NWC : NWC);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmImg);
  }
  __syncthreads();
  if constexpr (MC) cluster_sync_all();                   // the peer's barriers exist before anything is sent to them

  if (warp == NWC) {
    if (lane == 0) {
      // argmax elements of this CTA's rows (clamped to the padded row length; both multiples of 16 bytes)
      long long avail = p.arg_avail - r0;
      if (avail < 0) avail = 0;                            // MC: the CTA past the last row tile has no rows
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
        if constexpr (MC) {                                // boxes crank, crank + 2, ...: into both CTAs of the pair
          for (int b = static_cast<int>(crank); b < p.slab_n_box; b += 2) {
            const int q0 = b * p.slab_box_rows;
            tma_load_2d_mc(st + q0 * DS * 2, &tmImg, full, d0, (p.img_k0 + k) * p.p_stride + q0, 3);
          }
        } else {
          for (int b = 0; b < p.slab_n_box; ++b) {
            const int q0 = b * p.slab_box_rows;
            tma_load_2d(st + q0 * DS * 2, &tmImg, full, d0, (p.img_k0 + k) * p.p_stride + q0);
          }
        }
        if (qbytes > 0) {
          bulk_load_1d(st + p.tile_bytes, qsrc, qbytes, full);
          if (k + 12 < Bi) bulk_prefetch_l2(qsrc + 12 * qstep, qbytes);   // argmax bytes come from DRAM: stay 12 images ahead
        }
        qsrc += qstep;
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
    if constexpr (MC) cluster_sync_all();                 // pairs with the consumers' exit barrier below
    return;
  }

  // image blocks after the first continue from the partial sums in dTxt, so the additions happen in the same
  // order whatever the block size
  const float w_s = p.w_scale != nullptr ? __ldg(p.w_scale) : 1.f;       // powers of two: scaling is exact
  const float w_inv = p.w_scale != nullptr ? __ldg(p.w_scale + 1) : 1.f;
  float acc[8][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int r = r0 + warp * 32 + 4 * j + (lane >> 3);
    if (p.accumulate_txt && r < p.rows) {
      const float* o = p.dtxt + static_cast<size_t>(r) * p.D + d0 + (lane & 7) * 8;
      const float4 a = *reinterpret_cast<const float4*>(o);
      const float4 b = *reinterpret_cast<const float4*>(o + 4);
      acc[j][0] = a.x * w_s; acc[j][1] = a.y * w_s; acc[j][2] = a.z * w_s; acc[j][3] = a.w * w_s;
      acc[j][4] = b.x * w_s; acc[j][5] = b.y * w_s; acc[j][6] = b.z * w_s; acc[j][7] = b.w * w_s;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
    }
  }

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
  // entry = slab row offset << 16 | bf16(w): bytes (q*128) when P <= 256, 16-byte units (q*8) otherwise
  sts32(ent_st, pack_entry_lo(wr[0] * w_s, ldq_smem(smem0 + q_off) * (U16 ? 8u : 128u), F16));
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
                pack_entry_lo(wr[(b + 1) & 3] * w_s, ldq_smem(smem0 + nstage * stage_bytes + q_off) * (U16 ? 8u : 128u), F16));
        }
        wr[b] = __ldg(wptr + min(k + 4, Bi - 1));
        const uint32_t stl = stl0 + stage * stage_bytes;
        const uint32_t e[8] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y, eb.z, eb.w};
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
          const uint4 v = lds128(stl + (U16 ? ((e[jj] >> 16) << 4) : (e[jj] >> 16)));   // u8: one LEA.HI
          fhfma8_lo<F16>(acc[jj], v, e[jj]);
        }
        __syncwarp();
        if constexpr (MC) {
          if (lane < 2) mbar_arrive_cluster(empty0 + 8 * stage, lane);     // this warp is done with the slot in BOTH CTAs' eyes
        } else {
          if (lane == 0) mbar_arrive(empty0 + 8 * stage);
        }
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
      *reinterpret_cast<float4*>(o) = make_float4(acc[jj][0] * w_inv, acc[jj][1] * w_inv, acc[jj][2] * w_inv, acc[jj][3] * w_inv);
      *reinterpret_cast<float4*>(o + 4) = make_float4(acc[jj][4] * w_inv, acc[jj][5] * w_inv, acc[jj][6] * w_inv, acc[jj][7] * w_inv);
    }
  }
  if constexpr (MC) cluster_sync_all();                   // the peer may still arrive on this CTA's barriers
}

// ------------------------------------------------------------------------------------------------
// dImg
// ------------------------------------------------------------------------------------------------
// Gather formulation with STATIC accumulators (registers cannot be indexed dynamically, and a per-item indirect
// branch costs far more than the item's FMAs): a quarter warp owns NPO patches of NIMG images for one 64-wide D
// slice and walks the list of text rows whose argmax is that patch.  The lists come from a cheap pre-pass
// (one warp per (image, 512-row super-tile)): a deterministic counting sort of the super-tile's rows by patch.
// Lock-step padding of the four quarter warps uses a dummy entry: row 0 with weight +0.
//
// List block of one (image, super-tile), `block_bytes` long:
//   u32 hdr[hdr_words]   hdr[q] = first entry | number of entries << 16   (zero for q >= P)
//   u32 ent[512]         entries grouped by patch, ascending row inside a patch;
//                        entry = row-in-super-tile * 128 << 16 | bf16(w)   (high half: byte offset of the row in the tile)
constexpr int kStRows = 512;

struct DimgListParams {
  const void* argmax;   // [k_off + Bi][ld_arg] image-major
  long long ld_arg;
  const float* w;       // [Bt, ld_pairs]
  int ld_pairs;
  int k_first;          // first image of the chunk, as a column of argmax / w
  int n_img;            // images in the chunk
  int Bt, T, P, rows, n_st;
  int hdr_words, block_bytes;
  uint8_t* lists;       // [n_img][n_st][block_bytes]
  int f16;
  const float* w_scale;
};

template <bool U16>
__global__ void __launch_bounds__(256) maxsim_bwd_dimg_lists_kernel(const DimgListParams p) {
  extern __shared__ uint32_t lsm[];
  constexpr int kArgWords = U16 ? 256 : 128;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int per_warp = 2 * p.hdr_words + kStRows + kArgWords + 256;      // words
  uint32_t* cur = lsm + warp * per_warp;                 // per-patch count, then fill cursor
  uint32_t* hdr = cur + p.hdr_words;
  uint32_t* ent = hdr + p.hdr_words;
  uint32_t* arg_w = ent + kStRows;                       // the super-tile's argmax values, row order
  uint16_t* wrow = reinterpret_cast<uint16_t*>(arg_w + kArgWords);          // bf16 pair weight of each row
  const long long unit = static_cast<long long>(blockIdx.x) * 8 + warp;
  if (unit >= static_cast<long long>(p.n_img) * p.n_st) return;
  const int kl = static_cast<int>(unit / p.n_st);
  const int s = static_cast<int>(unit % p.n_st);
  const int kcol = p.k_first + kl;
  const int rows0 = s * kStRows;
  const int nrows = min(kStRows, p.rows - rows0);
  const uint32_t q_max = static_cast<uint32_t>(p.P - 1);
  const float w_s = p.w_scale != nullptr ? __ldg(p.w_scale) : 1.f;

  for (int i = lane; i < p.hdr_words; i += 32) cur[i] = 0u;
  __syncwarp();

  // ---- pass A: lane l takes rows [16 l, 16 l + 16): count per patch, stage argmax + pair weights in row order ----
  {
    const int rl = 16 * lane;
    const long long g0 = static_cast<long long>(rows0) + rl;                // first global row of this lane
    uint32_t a[U16 ? 8 : 4];
#pragma unroll
    for (int i = 0; i < (U16 ? 8 : 4); ++i) a[i] = 0u;
    if (g0 < p.ld_arg) {                                                     // ld_arg is a multiple of 16
      const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(p.argmax) +
                                                        (static_cast<size_t>(kcol) * p.ld_arg + g0) * (U16 ? 2 : 1));
      const uint4 v0 = __ldg(src);
      a[0] = v0.x; a[1] = v0.y; a[2] = v0.z; a[3] = v0.w;
      if (U16) {
        const uint4 v1 = __ldg(src + 1);
        a[4] = v1.x; a[5] = v1.y; a[6] = v1.z; a[7] = v1.w;
      }
    }
    uint4* adst = reinterpret_cast<uint4*>(arg_w) + lane * (U16 ? 2 : 1);
    adst[0] = make_uint4(a[0], a[1], a[2], a[3]);
    if (U16) adst[1] = make_uint4(a[4], a[5], a[6], a[7]);
    int m = static_cast<int>(g0 / p.T);
    int rem = static_cast<int>(g0 - static_cast<long long>(m) * p.T);
    uint32_t wpk[8];
    if (p.T >= 16 && nrows == kStRows) {
      // common case (full super-tile, a lane's 16 rows span at most two texts): straight-line
      const int nb = p.T - rem;                            // rows of this lane that still belong to text m
      const float w0 = __ldg(p.w + static_cast<size_t>(m) * p.ld_pairs + kcol);
      const float w1 = (nb < 16) ? __ldg(p.w + static_cast<size_t>(m + 1) * p.ld_pairs + kcol) : 0.f;
      const uint32_t b0 = cvt_w16(w0 * w_s, p.f16);
      const uint32_t b1 = cvt_w16(w1 * w_s, p.f16);
#pragma unroll
      for (int i = 0; i < 16; i += 2) wpk[i >> 1] = (i < nb ? b0 : b1) | ((i + 1 < nb ? b0 : b1) << 16);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint32_t q = U16 ? ((a[i >> 1] >> (16 * (i & 1))) & 0xFFFFu) : ((a[i >> 2] >> (8 * (i & 3))) & 0xFFu);
        atomicAdd(&cur[min(q, q_max)], 1u);
      }
    } else {
      float wv = (g0 < p.rows) ? __ldg(p.w + static_cast<size_t>(m) * p.ld_pairs + kcol) : 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        uint32_t wb = 0u;
        if (rl + i < nrows) {
          if (rem == p.T) {
            rem = 0;
            ++m;
            wv = __ldg(p.w + static_cast<size_t>(m) * p.ld_pairs + kcol);
          }
          ++rem;
          const uint32_t q = U16 ? ((a[i >> 1] >> (16 * (i & 1))) & 0xFFFFu) : ((a[i >> 2] >> (8 * (i & 3))) & 0xFFu);
          atomicAdd(&cur[min(q, q_max)], 1u);
          wb = cvt_w16(wv * w_s, p.f16);
        }
        if (i & 1) wpk[i >> 1] |= wb << 16;
        else wpk[i >> 1] = wb;
      }
    }
    uint4* wdst = reinterpret_cast<uint4*>(wrow) + lane * 2;
    wdst[0] = make_uint4(wpk[0], wpk[1], wpk[2], wpk[3]);
    wdst[1] = make_uint4(wpk[4], wpk[5], wpk[6], wpk[7]);
  }
  __syncwarp();

  // ---- pass B: exclusive prefix of the counts -> header, fill cursors ----
  uint32_t carry = 0;
  for (int b0 = 0; b0 < p.hdr_words; b0 += 32) {
    const int b = b0 + lane;
    const uint32_t cnt = cur[b];
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    const uint32_t start = carry + incl - cnt;
    hdr[b] = start | (cnt << 16);
    cur[b] = start;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  __syncwarp();

  // ---- pass C: place the entries, 32 consecutive rows per round; lanes with the same patch are ranked by
  //      lane (= row) order, so every list comes out in ascending row order whatever the scheduling ----
  const uint32_t lt_mask = (1u << lane) - 1;
  const uint8_t* arg8 = reinterpret_cast<const uint8_t*>(arg_w);
  const uint16_t* arg16 = reinterpret_cast<const uint16_t*>(arg_w);
  for (int r = lane; r < ((nrows + 31) & ~31); r += 32) {
    const bool valid = r < nrows;
    const uint32_t q = min(static_cast<uint32_t>(U16 ? arg16[r] : arg8[r]), q_max);
    const uint32_t mask = __match_any_sync(0xffffffffu, valid ? q : 0xFFFFFFFFu);
    const uint32_t base = valid ? cur[q] : 0u;           // every lane of a group reads the cursor ...
    __syncwarp();
    const uint32_t rank = __popc(mask & lt_mask);
    if (valid) {
      if (rank == 0) cur[q] = base + __popc(mask);       // ... before its first lane moves it on
      ent[base + rank] = (static_cast<uint32_t>(r * 128) << 16) | static_cast<uint32_t>(wrow[r]);
    }
    __syncwarp();
  }
  uint4* dst = reinterpret_cast<uint4*>(p.lists + static_cast<size_t>(unit) * p.block_bytes);
  const uint4* h4 = reinterpret_cast<const uint4*>(hdr);
  const uint4* e4 = reinterpret_cast<const uint4*>(ent);
  const int hq = p.hdr_words / 4;
  for (int i = lane; i < hq; i += 32) dst[i] = h4[i];
  for (int i = lane; i < kStRows / 4; i += 32) dst[hq + i] = e4[i];
}

struct __align__(8) DimgBarriers {
  uint64_t full[4];
  uint64_t empty[4];
};

// Warps [0, NW): consumers (quarter warp o = 4 warp + lane / 8 owns patches [o NPO, o NPO + NPO) of the CTA's NIMG
// images); warp NW issues the loads.  Stage = [512 x 64 bf16 super-tile of text rows | NIMG list blocks].
template <int NW, int NPO, int NIMG, bool F16>
__global__ void __launch_bounds__((NW + 1) * 32, 1)
maxsim_bwd_dimg_v3_kernel(const __grid_constant__ CUtensorMap tmTxt, const BwdV2Params p) {
  constexpr int DS = 64;
  constexpr uint32_t kTileBytes = kStRows * DS * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  DimgBarriers* bars = reinterpret_cast<DimgBarriers*>(smem + p.stages * p.stage_bytes);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int d0 = blockIdx.x * DS;
  const int kl0 = blockIdx.y * NIMG;                     // first image of this CTA inside the chunk
  const int n_st = p.n_st;
  const int stages = p.stages;
  const uint32_t stage_bytes = p.stage_bytes;
  const uint32_t block_bytes = p.block_bytes;
  const uint32_t smem0 = smem_u32(smem);
  const uint32_t full0 = smem_u32(&bars->full[0]);
  const uint32_t empty0 = smem_u32(&bars->empty[0]);

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, NW);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmTxt);
  }
  __syncthreads();

  if (warp == NW) {
    if (lane == 0) {
      const uint8_t* lsrc[NIMG];
#pragma unroll
      for (int i = 0; i < NIMG; ++i)
        lsrc[i] = p.lists + static_cast<size_t>(min(kl0 + i, p.chunk_n - 1)) * n_st * block_bytes;
      int stage = 0;
      uint32_t phase = 0;
      for (int s = 0; s < n_st; ++s) {
        mbar_wait(empty0 + 8 * stage, phase ^ 1);
        const uint32_t st = smem0 + stage * stage_bytes;
        const uint32_t full = full0 + 8 * stage;
        mbar_arrive_expect_tx(full, kTileBytes + NIMG * block_bytes);
        tma_load_2d(st, &tmTxt, full, d0, s * kStRows);                          // rows past the end: zero-filled
        tma_load_2d(st + kTileBytes / 2, &tmTxt, full, d0, s * kStRows + kStRows / 2);
#pragma unroll
        for (int i = 0; i < NIMG; ++i)
          bulk_load_1d(st + kTileBytes + i * block_bytes, lsrc[i] + static_cast<size_t>(s) * block_bytes, block_bytes, full);
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
    return;
  }

  float acc[NIMG][NPO][8];
#pragma unroll
  for (int i = 0; i < NIMG; ++i)
#pragma unroll
    for (int j = 0; j < NPO; ++j)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[i][j][c] = 0.f;
  const int owner = warp * 4 + (lane >> 3);
  const uint32_t lane_off = (lane & 7) * 16;
  const uint32_t hdr_bytes = p.hdr_words * 4;
  int stage = 0;
  uint32_t phase = 0;
  for (int s = 0; s < n_st; ++s) {
    const uint32_t st = smem0 + stage * stage_bytes;
    mbar_wait(full0 + 8 * stage, phase);
    const uint32_t row_base = st + lane_off;
    // headers of this quarter warp's NIMG x NPO lists first (their latency overlaps), then the lists
    uint32_t h[NIMG][NPO];
#pragma unroll
    for (int i = 0; i < NIMG; ++i)
#pragma unroll
      for (int j = 0; j < NPO; ++j)
        asm volatile("ld.shared.b32 %0, [%1];"
                     : "=r"(h[i][j]) : "r"(st + kTileBytes + i * block_bytes + (owner * NPO + j) * 4));
#pragma unroll
    for (int i = 0; i < NIMG; ++i) {
      const uint32_t ent = st + kTileBytes + i * block_bytes + hdr_bytes;
#pragma unroll
      for (int j = 0; j < NPO; ++j) {
        const int len = static_cast<int>(h[i][j] >> 16);
        uint32_t ep = ent + (h[i][j] & 0xFFFFu) * 4;
        const int maxlen = __reduce_max_sync(0xffffffffu, len);
        // two items per trip, then at most one single; a quarter past the end of its own list takes a dummy
        // entry (row 0 with weight +0)
        int it = 0;
#pragma unroll 1
        for (; it + 2 <= maxlen; it += 2, ep += 8) {
          uint32_t ea, eb;
          asm volatile("{\n\t.reg .pred pa, pb;\n\t"
                       "setp.lt.s32 pa, %3, %4;\n\t setp.lt.s32 pb, %5, %4;\n\t"
                       "mov.b32 %0, 0;\n\t mov.b32 %1, 0;\n\t"
                       "@pa ld.shared.b32 %0, [%2];\n\t @pb ld.shared.b32 %1, [%2+4];\n\t}"
                       : "=&r"(ea), "=&r"(eb) : "r"(ep), "r"(it), "r"(len), "r"(it + 1));
          const uint4 va = lds128(row_base + (ea >> 16));
          const uint4 vb = lds128(row_base + (eb >> 16));
          fhfma8_lo<F16>(acc[i][j], va, ea);
          fhfma8_lo<F16>(acc[i][j], vb, eb);
        }
        if (it < maxlen) {
          uint32_t ea;
          asm volatile("{\n\t.reg .pred pa;\n\t"
                       "setp.lt.s32 pa, %2, %3;\n\t mov.b32 %0, 0;\n\t @pa ld.shared.b32 %0, [%1];\n\t}"
                       : "=&r"(ea) : "r"(ep), "r"(it), "r"(len));
          const uint4 va = lds128(row_base + (ea >> 16));
          fhfma8_lo<F16>(acc[i][j], va, ea);
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty0 + 8 * stage);
    if (++stage == stages) { stage = 0; phase ^= 1; }
  }
#pragma unroll
  for (int i = 0; i < NIMG; ++i) {
    const int kl = kl0 + i;
    if (kl < p.chunk_n) {
#pragma unroll
      for (int j = 0; j < NPO; ++j) {
        const int q = owner * NPO + j;
        if (q < p.P) {
          float* o = p.dimg + (static_cast<size_t>(p.chunk_k0 + kl) * p.P + q) * p.D + d0 + (lane & 7) * 8;
          const float w_inv = p.w_scale != nullptr ? __ldg(p.w_scale + 1) : 1.f;
          float4 lo = make_float4(acc[i][j][0] * w_inv, acc[i][j][1] * w_inv, acc[i][j][2] * w_inv, acc[i][j][3] * w_inv);
          float4 hi = make_float4(acc[i][j][4] * w_inv, acc[i][j][5] * w_inv, acc[i][j][6] * w_inv, acc[i][j][7] * w_inv);
          if (p.accumulate_img) {
            const float4 a = *reinterpret_cast<const float4*>(o);
            const float4 b = *reinterpret_cast<const float4*>(o + 4);
            lo.x += a.x; lo.y += a.y; lo.z += a.z; lo.w += a.w;
            hi.x += b.x; hi.y += b.y; hi.z += b.z; hi.w += b.w;
          }
          *reinterpret_cast<float4*>(o) = lo;
          *reinterpret_cast<float4*>(o + 4) = hi;
        }
      }
    }
  }
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
int make_tmap_plain(CUtensorMap* out, const void* base, long long rows, int D, int box_cols, int box_rows, int f16) {
  EncodeTiledFnV2 fn = encode_fn_v2();
  if (fn == nullptr) return CXL_ERR_CUDA;
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(D), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(D) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_cuda_error(cudaErrorInvalidValue, "cuTensorMapEncodeTiled (bwd)");
    return CXL_ERR_CUDA;
  }
  return CXL_OK;
}

int round_up_i(int x, int m) { return (x + m - 1) / m * m; }

int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return (s != nullptr && *s != 0) ? atoi(s) : dflt;
}

template <int NWC, bool U16, bool F16, bool MC>
static int launch_dtxt_v2_k(const void* img, BwdV2Params p, int max_smem, cudaStream_t stream) {
  constexpr int DS = 64;
  constexpr int ROWS = NWC * 32;
  // the TMA box covers P slab rows when P <= 256; a larger P is split EVENLY into ceil(P/256) boxes (576 = 3 x 192:
  // the stage holds exactly the image's 72 KB and three stages fit; with 256-row boxes it held 768 rows = 96 KB, a
  // third of them the next image's, and only two stages fit).  A box that is rounded up may read into the next image
  // or past the end (zero fill); those rows are never gathered.  MC: an even number of boxes, half per CTA of the pair.
  const int n_box = MC ? 2 * ((p.P + 511) / 512) : (p.P + 255) / 256;
  const int box_rows = (!MC && p.P <= 256) ? p.P : round_up_i((p.P + n_box - 1) / n_box, 8);
  const int slab_rows = n_box * box_rows;
  p.slab_box_rows = box_rows;
  p.slab_n_box = n_box;
  p.tile_bytes = slab_rows * DS * 2;
  p.stage_bytes = round_up_i(p.tile_bytes + ROWS * (U16 ? 2 : 1), 128);
  const int overhead = 128 + static_cast<int>(sizeof(V2Barriers)) + NWC * 256;
  int stages = (max_smem - overhead) / p.stage_bytes;
  if (stages < 2) return CXL_ERR_UNSUPPORTED;
  p.stages = stages > 6 ? 6 : stages;
  const int smem_bytes = overhead + p.stages * p.stage_bytes;
  CUtensorMap tm;
  int rc = make_tmap_plain(&tm, img, static_cast<long long>(p.Bi) * p.p_stride, p.D, DS, box_rows, F16);
  if (rc != CXL_OK) return rc;
  auto kern = maxsim_bwd_dtxt_v2_kernel<NWC, U16, F16, MC>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  int gx = (p.rows + ROWS - 1) / ROWS;
  if (MC) gx = (gx + 1) & ~1;                             // whole pairs; the extra CTA only takes part in the protocol
  dim3 grid(gx, p.D / DS);
  // Image blocks: every CTA of a D slice streams the same [P,64] slabs, but the CTAs of later waves start
  // while earlier ones are mid-stream, so without blocking each CTA ends up pulling its slabs from DRAM
  // (measured at cfg-2: 25.8 TB of DRAM reads, 88 % of HBM bandwidth, L2 hit 31 %).  One launch per block of
  // images whose slabs for one slice (~32 MB) stay in L2 across all waves; partial dTxt is accumulated in
  // place (rows*D*8 B of traffic per extra block, negligible).
  const long long slab_bytes = static_cast<long long>(p.p_stride) * DS * 2;
  int kb = static_cast<int>((static_cast<long long>(env_int("CXL_DTXT_BLOCK_MB", 32)) << 20) / slab_bytes);
  if (kb < 8) kb = 8;
  const int Bi = p.Bi, k_off = p.k_off;
  const int acc_first = p.accumulate_txt;          // shard calls continue from the previous shards' partial sums
  for (int k0 = 0; k0 < Bi; k0 += kb) {
    p.Bi = (Bi - k0 < kb) ? Bi - k0 : kb;
    p.k_off = k_off + k0;
    p.img_k0 = k0;
    p.accumulate_txt = (k0 > 0) || acc_first;
    if (MC) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = grid;
      cfg.blockDim = dim3((NWC + 1) * 32);
      cfg.dynamicSmemBytes = smem_bytes;
      cfg.stream = stream;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      note_launch();
      CXL_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tm, p));
    } else {
      kern<<<grid, (NWC + 1) * 32, smem_bytes, CXL_COUNTED(stream)>>>(tm, p);
    }
    CXL_CUDA_CHECK(cudaGetLastError());
  }
  return CXL_OK;
}

// CXL_DTXT_MC: 0 (default) = independent CTAs, 1 = row-tile pairs with multicast slabs for P > 256, 2 = always.
// NEGATIVE RESULT (one B200, interleaved, profiles/r4x_dtxt_multicast_ab.txt): bit-identical output, but 61.9 vs 39.2 ms
// at P 576 / D 768 (B 2048) and 133.1 vs 55.5 ms at P 196 / D 512 (B 4096).  Halving the L2 -> SM slab traffic does not
// pay for coupling the two CTAs: a stage is refilled only when BOTH CTAs' 15 consumer warps have released it (remote
// mbarrier arrives), so the pair runs at the pace of its slower half with the latency of a cluster round trip per
// stage, and the 3-6 stage ring is too shallow to absorb that.  Kept as an option for A/B.
template <int NWC, bool U16, bool F16>
static int launch_dtxt_v2(const void* img, BwdV2Params p, int max_smem, cudaStream_t stream) {
  const int mc = env_int("CXL_DTXT_MC", 0);
  const bool use_mc = (mc >= 2 || (mc == 1 && p.P > 256)) && p.rows > NWC * 32;
  if (use_mc) return launch_dtxt_v2_k<NWC, U16, F16, true>(img, p, max_smem, stream);
  return launch_dtxt_v2_k<NWC, U16, F16, false>(img, p, max_smem, stream);
}

// One chunk of images: lists pre-pass, then the gather kernel.  `p.lists` is scratch for chunk_cap images.
template <int NW, int NPO, int NIMG, bool F16>
static int launch_dimg_v3(const void* txt, BwdV2Params p, int max_smem, int chunk_cap, cudaStream_t stream) {
  constexpr int DS = 64;
  p.stage_bytes = kStRows * DS * 2 + NIMG * p.block_bytes;           // tile | list blocks
  const int overhead = 128 + static_cast<int>(sizeof(DimgBarriers));
  int stages = (max_smem - overhead) / p.stage_bytes;
  if (stages < 2) return CXL_ERR_UNSUPPORTED;
  p.stages = stages > 4 ? 4 : stages;
  const int smem_bytes = overhead + p.stages * p.stage_bytes;
  CUtensorMap tm;
  int rc = make_tmap_plain(&tm, txt, p.rows, p.D, DS, kStRows / 2, F16);
  if (rc != CXL_OK) return rc;
  auto kern = maxsim_bwd_dimg_v3_kernel<NW, NPO, NIMG, F16>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));

  DimgListParams lp;
  lp.argmax = p.argmax; lp.ld_arg = p.ld_arg; lp.w = p.w; lp.ld_pairs = p.ld_pairs;
  lp.Bt = p.Bt; lp.T = p.T; lp.P = p.P; lp.rows = p.rows; lp.n_st = p.n_st;
  lp.hdr_words = p.hdr_words; lp.block_bytes = p.block_bytes;
  lp.lists = const_cast<uint8_t*>(p.lists);
  lp.f16 = p.f16; lp.w_scale = p.w_scale;
  const int lists_smem = 8 * (2 * lp.hdr_words + kStRows + (p.argmax_u16 ? 256 : 128) + 256) * 4;
  auto lkern = p.argmax_u16 ? maxsim_bwd_dimg_lists_kernel<true> : maxsim_bwd_dimg_lists_kernel<false>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(lkern, cudaFuncAttributeMaxDynamicSharedMemorySize, lists_smem));

  for (int k0 = 0; k0 < p.Bi; k0 += chunk_cap) {
    const int n = (p.Bi - k0 < chunk_cap) ? p.Bi - k0 : chunk_cap;
    lp.k_first = p.k_off + k0;
    lp.n_img = n;
    const long long units = static_cast<long long>(n) * p.n_st;
    lkern<<<static_cast<unsigned>((units + 7) / 8), 256, lists_smem, CXL_COUNTED(stream)>>>(lp);
    CXL_CUDA_CHECK(cudaGetLastError());
    p.chunk_k0 = k0;
    p.chunk_n = n;
    dim3 grid(p.D / DS, (n + NIMG - 1) / NIMG);
    kern<<<grid, (NW + 1) * 32, smem_bytes, CXL_COUNTED(stream)>>>(tm, p);
    CXL_CUDA_CHECK(cudaGetLastError());
  }
  return CXL_OK;
}


static bool stream_path_covers(const void* txt, const void* img, int Bt, int T, int P, int D) {
  if (D % 128 != 0 || P > 576 || T < 8) return false;
  if (static_cast<long long>(Bt) * T >= (1LL << 31)) return false;
  if ((reinterpret_cast<uintptr_t>(txt) & 15) || (reinterpret_cast<uintptr_t>(img) & 15)) return false;
  return true;
}


// ---- dImg v4 (maxsim_bwd_dimg4.cu) ----
static int max_optin_smem() {
  int dev = 0, v = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 232448;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess || v <= 0) return 232448;
  return v;
}
// dImg generation: 3 = 512-row super-tiles + per-patch row lists; 4 = staggered long walks + pre-padded step tables
// (maxsim_bwd_dimg4.cu).  Bit-identical results.  Measured on B200: P = 196 within +-2 % of each other
// (profiles/r4a_dimg_v4_ab.txt) -> v3; P = 576, D = 768: v4 22 % faster (65.4 vs 83.9 ms at B = 2048, lists of 0.9
// rows make v3's per-list overhead dominate) -> v4.  CXL_DIMG_V=3|4 forces one.
static bool use_dimg4(int P) {
  const int v = env_int("CXL_DIMG_V", 0);
  if (!dimg4_covers(P)) return false;
  return v == 4 || (v != 3 && P > 304);
}
// images per chunk for `budget_bytes` of scratch: whole waves of the gather kernel (1 CTA per SM, D/64 CTAs per image group)
static long long dimg4_cap(const Dimg4Geom& g, int Bi, int D, long long budget_bytes) {
  long long cap = (budget_bytes - 256) / dimg4_bytes_per_image(g);
  if (cap >= Bi) return Bi;
  int a = num_sms(), b = D / 64;
  while (b != 0) { const int t = a % b; a = b; b = t; }      // a = gcd(SMs, D/64)
  const long long groups_per_wave = num_sms() / a;
  const long long per_wave = groups_per_wave * g.nimg;
  if (cap >= per_wave) cap = cap / per_wave * per_wave;
  else cap = cap / g.nimg * g.nimg;
  return cap;
}

// dImg plan: kernel variant (consumer warps x patches per quarter warp x images per CTA), list-block geometry and
// how many images' row lists fit `budget_bytes` of scratch (whole waves of the gather kernel when chunking)
struct DimgPlan {
  int npo, nw, nimg, n_st, hdr_words, block_bytes;
  long long per_img;   // scratch bytes per image
  long long cap;       // images per chunk
};
static DimgPlan dimg_plan(int Bt, int Bi, int T, int P, long long budget_bytes) {
  DimgPlan d;
  d.npo = P <= 200 ? 2 : P <= 300 ? 3 : 8;
  d.nw = P <= 300 ? 25 : 19;
  d.nimg = P <= 200 ? 2 : 1;
  d.n_st = (Bt * T + kStRows - 1) / kStRows;
  d.hdr_words = round_up_i(d.nw * 4 * d.npo, 32);
  d.block_bytes = d.hdr_words * 4 + kStRows * 4;
  d.per_img = static_cast<long long>(d.n_st) * d.block_bytes;
  long long cap = budget_bytes / d.per_img;
  if (cap >= Bi) {
    cap = Bi;
  } else {
    const long long per_wave = static_cast<long long>(num_sms()) * d.nimg;   // images per D/64 waves (1 CTA/SM)
    if (cap >= per_wave) cap = cap / per_wave * per_wave;
    cap = cap / d.nimg * d.nimg;
  }
  d.cap = cap;
  return d;
}
static long long dimg_budget_bytes() {
  const int mb = env_int("CXL_DIMG_SCRATCH_MB", 4608);       // read per call: tests shrink it
  return static_cast<long long>(mb < 1 ? 1 : mb) << 20;
}

// preferred workspace of maxsim_bwd_stream for this shape (0: none needed / shape not on the fast path)
size_t maxsim_bwd_workspace_bytes(int Bt, int Bi, int T, int P, int D) {
  if (D % 128 != 0 || P > 576 || T < 8 || static_cast<long long>(Bt) * T >= (1LL << 31)) return 0;
  if (use_dimg4(P)) {
    const Dimg4Geom g = dimg4_geom(static_cast<long long>(Bt) * T, P, max_optin_smem());
    long long cap = dimg4_cap(g, Bi, D, dimg_budget_bytes());
    if (cap < g.nimg) cap = g.nimg < Bi ? g.nimg : Bi;
    return static_cast<size_t>(cap * dimg4_bytes_per_image(g) + 256);
  }
  DimgPlan d = dimg_plan(Bt, Bi, T, P, dimg_budget_bytes());
  if (d.cap < d.nimg) d.cap = d.nimg;
  return static_cast<size_t>(d.cap * d.per_img);
}

// returns CXL_ERR_UNSUPPORTED when the shape is outside the fast path (caller falls back to the generic kernels).
// ws/ws_bytes: caller-provided scratch for the dImg row lists (16-byte aligned; see maxsim_bwd_workspace_bytes);
// ws == nullptr: stream-ordered allocation from the device's default memory pool.
// arg_row_off: first text-token row of this call inside the argmax rows (row-slice calls, dTxt only).
int maxsim_bwd_stream(const void* txt, const void* img, const void* argmax, long long ld_arg, long long arg_row_off,
                      const float* w, int Bt, int Bi, int T, int P, int D, int p_stride, int ld_pairs, int k_off,
                      float* dtxt, int accumulate_txt, float* dimg, int accumulate_img, void* ws, size_t ws_bytes,
                      int f16, const float* w_scale, cudaStream_t stream) {
  if (!stream_path_covers(txt, img, Bt, T, P, D)) return CXL_ERR_UNSUPPORTED;
  const int es = P > 256 ? 2 : 1;
  if ((arg_row_off * es) % 16 != 0) return CXL_ERR_UNSUPPORTED;     // bulk copies of argmax runs need 16 B alignment
  argmax = static_cast<const uint8_t*>(argmax) + arg_row_off * es;
  if (ws != nullptr && (reinterpret_cast<uintptr_t>(ws) & 15)) return CXL_ERR_INVALID_ARG;
  BwdV2Params p;
  memset(&p, 0, sizeof(p));
  p.w = w; p.argmax = argmax; p.ld_arg = ld_arg; p.arg_avail = ld_arg - arg_row_off;
  p.Bt = Bt; p.Bi = Bi; p.T = T; p.P = P; p.D = D; p.p_stride = p_stride;
  p.rows = Bt * T;
  p.ld_pairs = ld_pairs; p.k_off = k_off; p.argmax_u16 = P > 256;
  p.dtxt = dtxt; p.dimg = dimg; p.accumulate_img = accumulate_img; p.accumulate_txt = accumulate_txt;
  p.f16 = f16 ? 1 : 0; p.w_scale = w_scale;
  int dev = 0, max_smem = 0;
  CXL_CUDA_CHECK(cudaGetDevice(&dev));
  CXL_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));

  int rc = CXL_OK;
  if (dtxt != nullptr) {
    if (f16) rc = P > 256 ? launch_dtxt_v2<15, true, true>(img, p, max_smem, stream) : launch_dtxt_v2<15, false, true>(img, p, max_smem, stream);
    else rc = P > 256 ? launch_dtxt_v2<15, true, false>(img, p, max_smem, stream) : launch_dtxt_v2<15, false, false>(img, p, max_smem, stream);
    if (rc != CXL_OK) return rc;
  }
  if (dimg != nullptr && use_dimg4(P)) {
    const Dimg4Geom g = dimg4_geom(p.rows, P, max_smem);
    const long long budget = ws != nullptr ? static_cast<long long>(ws_bytes) : dimg_budget_bytes();
    long long cap = dimg4_cap(g, Bi, D, budget);
    const long long min_cap = g.nimg < Bi ? g.nimg : Bi;
    void* scratch = ws;
    if (ws != nullptr) {
      if (cap < min_cap) return CXL_ERR_WORKSPACE;
    } else {
      if (cap < min_cap) cap = min_cap;
      for (;;) {
        const cudaError_t ae = cudaMallocAsync(&scratch, static_cast<size_t>(cap * dimg4_bytes_per_image(g) + 256), stream);
        if (ae == cudaSuccess) break;
        (void)cudaGetLastError();
        if (ae != cudaErrorMemoryAllocation || cap <= min_cap) {
          set_cuda_error(ae, "cudaMallocAsync (dImg step tables)");
          return CXL_ERR_CUDA;
        }
        cap = (cap / 2 + g.nimg - 1) / g.nimg * g.nimg;
      }
    }
    rc = launch_dimg_v4(txt, argmax, ld_arg, w, Bt, Bi, T, P, D, ld_pairs, k_off, dimg, accumulate_img, g, scratch, cap,
                        max_smem, f16, w_scale, stream);
    if (ws == nullptr) {
      const cudaError_t fe = cudaFreeAsync(scratch, stream);
      if (rc == CXL_OK && fe != cudaSuccess) {
        set_cuda_error(fe, "cudaFreeAsync (dImg step tables)");
        rc = CXL_ERR_CUDA;
      }
    }
    if (rc != CXL_OK) return rc;
  } else if (dimg != nullptr) {
    DimgPlan d = dimg_plan(Bt, Bi, T, P, ws != nullptr ? static_cast<long long>(ws_bytes) : dimg_budget_bytes());
    p.n_st = d.n_st;
    p.hdr_words = d.hdr_words;
    p.block_bytes = d.block_bytes;
    long long cap = d.cap;
    void* scratch = ws;
    if (ws != nullptr) {
      if (cap < d.nimg && cap < Bi) return CXL_ERR_WORKSPACE;
    } else {
      if (cap < d.nimg) cap = d.nimg;
      static bool pool_set[64] = {false};
      if (dev < 64 && !pool_set[dev]) {           // keep freed scratch cached in the device's pool between calls
        cudaMemPool_t pool;
        CXL_CUDA_CHECK(cudaDeviceGetDefaultMemPool(&pool, dev));
        unsigned long long keep = ~0ull;
        CXL_CUDA_CHECK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
        pool_set[dev] = true;
      }
      // when the device is nearly full (the framework's own allocator holds the rest) settle for smaller chunks
      for (;;) {
        const cudaError_t ae = cudaMallocAsync(&scratch, static_cast<size_t>(cap * d.per_img), stream);
        if (ae == cudaSuccess) break;
        (void)cudaGetLastError();
        if (ae != cudaErrorMemoryAllocation || cap <= d.nimg) {
          set_cuda_error(ae, "cudaMallocAsync (dImg lists)");
          return CXL_ERR_CUDA;
        }
        cap = (cap / 2 + d.nimg - 1) / d.nimg * d.nimg;
      }
    }
    p.lists = static_cast<const uint8_t*>(scratch);
    const int cap_i = static_cast<int>(cap);
    if (f16) {
      if (d.npo == 2) rc = launch_dimg_v3<25, 2, 2, true>(txt, p, max_smem, cap_i, stream);
      else if (d.npo == 3) rc = launch_dimg_v3<25, 3, 1, true>(txt, p, max_smem, cap_i, stream);
      else rc = launch_dimg_v3<19, 8, 1, true>(txt, p, max_smem, cap_i, stream);
    } else {
      if (d.npo == 2) rc = launch_dimg_v3<25, 2, 2, false>(txt, p, max_smem, cap_i, stream);
      else if (d.npo == 3) rc = launch_dimg_v3<25, 3, 1, false>(txt, p, max_smem, cap_i, stream);
      else rc = launch_dimg_v3<19, 8, 1, false>(txt, p, max_smem, cap_i, stream);
    }
    if (ws == nullptr) {
      const cudaError_t fe = cudaFreeAsync(scratch, stream);
      if (rc == CXL_OK && fe != cudaSuccess) {
        set_cuda_error(fe, "cudaFreeAsync (dImg lists)");
        rc = CXL_ERR_CUDA;
      }
    }
    if (rc != CXL_OK) return rc;
  }
  return CXL_OK;
}

}  // namespace cxl
