// K2 dImg, fourth generation: gather with static register accumulators over LONG row lists.
//
//   dImg[k,q,:] = sum_{r : q*(r,k) = q} w[t(r),k] * txt[r,:]        (autograd of loss.py:34-35 w.r.t. the image tokens)
//
// The third generation (maxsim_bwd_v2.cu) sorted every 512-row super-tile by patch and let quarter warps walk
// lists of 512/P = 2.6 rows in 4-way lock step: 45 % of its slots were padding, and every list paid a header
// decode, a warp-wide max and predicated entry loads.  This version changes three things:
//
//  * STAGGERED super-tiles.  A CTA owns NIMG images of one 64-wide D slice.  Text rows stream through a ring of
//    NIMG+DEPTH sub-tiles; image c of the CTA walks the NIMG sub-tiles ending at sub-tile tau whenever
//    tau % NIMG == c.  Every sub-tile is loaded once, used by all NIMG images, has DEPTH walks of lead time, and
//    yet each image walks lists over NIMG sub-tiles (longer lists: less lock-step padding per item).
//  * PRE-PADDED step tables.  The pre-pass (one warp per (image, walk)) emits, per consumer warp and per
//    accumulator set ("phase"), a table of steps: 4 words per step, one per quarter warp, already padded to the
//    warp's longest list (rounded up to an even number of steps) with zero words.  A warp's phases are contiguous,
//    so its whole walk is ONE stream of step pairs: LDS.64 (two steps' entries), address arithmetic, two LDS.128
//    and 16 FHFMA per pair; no per-quarter headers, no warp reductions, no predicate setup, no odd tails
//    (optionally one compare per step that switches a padding slot's row read off).
//  * entry = (byte offset of the row inside the RING) >> 2 << 16 | bf16(w): the pre-pass knows which ring slot
//    every sub-tile will occupy (slot = sub-tile index % (NIMG+1)), so the consumer never computes a slot.
//
// Every (image, patch) accumulator still sees its rows in ascending order whatever the walk partition is, so
// results are bit-identical across chunkings, image shards and positions of an image inside its CTA.
#include "bwd_v2_common.cuh"

namespace cxl {

struct Dimg4StepsParams {
  const void* argmax;      // [k_first + n_img][ld_arg] image-major
  long long ld_arg;
  const float* w;          // [Bt, ld_pairs]
  int ld_pairs;
  int k_first;             // column of argmax / w of the chunk's first image
  int n_img;               // images in the chunk
  int T, P, rows;
  Dimg4Geom g;
  uint8_t* blocks;         // [n_img][walks][blk_stride]
  uint16_t* blk_units;     // [n_img][walks] bytes / 16 of each block that are in use (0: empty walk)
  int cap_words;           // table words that fit the pre-pass's shared-memory staging (larger tables go straight to global)
  int per_warp_words;
  int f16;                 // 16-bit pair weights are fp16 (fp16 operands) instead of bf16
  const float* w_scale;    // device [s, 1/s] or null
};

// Step table of one (image, walk):
//   u16 hdr[nw][hs]      hdr[w][j] = steps of phase j (j < npo), hdr[w][npo] = index of warp w's first step
//   u32 step[..][4]      phase after phase, warp after warp; every phase holds an EVEN number of steps, pairs of
//                        steps are interleaved ([pair][quarter][2]) so a quarter reads two steps with one LDS.64;
//                        padding slots are zero words (row offset 0, weight +0)
template <bool U16, int NPO>
__global__ void __launch_bounds__(128) dimg4_steps_kernel(const Dimg4StepsParams p) {
  extern __shared__ __align__(16) uint32_t sm4[];
  constexpr int ES = U16 ? 2 : 1;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const Dimg4Geom& g = p.g;
  const int rows_max = g.nimg * g.sub_rows;
  const int PQ = g.nw * 4 * NPO;
  const int NPH = g.nw * NPO;
  uint32_t* cnt = sm4 + static_cast<size_t>(warp) * p.per_warp_words;
  uint32_t* pst = cnt + PQ;
  uint32_t* argw = pst + ((NPH + 3) & ~3);
  const int arg_words = ((rows_max * ES + 15) / 16) * 4;
  uint16_t* wrow = reinterpret_cast<uint16_t*>(argw + arg_words);
  uint16_t* srow = wrow + rows_max;
  uint32_t* hdr = reinterpret_cast<uint32_t*>(srow + rows_max);
  uint32_t* tab_s = hdr + g.hdr_bytes / 4;

  const long long unit = static_cast<long long>(blockIdx.x) * 4 + warp;
  if (unit >= static_cast<long long>(p.n_img) * g.walks) return;
  const int kl = static_cast<int>(unit / g.walks);
  const int m = static_cast<int>(unit - static_cast<long long>(kl) * g.walks);
  const int tau = m * g.nimg + kl % g.nimg;
  if (tau > g.tau_max) {
    if (lane == 0) p.blk_units[unit] = 0;
    return;
  }
  uint8_t* blk = p.blocks + static_cast<size_t>(unit) * g.blk_stride;
  const int lo_sub = max(0, tau - g.nimg + 1);
  const int hi_sub = min(g.H - 1, tau);
  const int r_lo = lo_sub * g.sub_rows;
  const int nrows = min(p.rows, (hi_sub + 1) * g.sub_rows) - r_lo;
  const int kcol = p.k_first + kl;
  const uint32_t q_max = static_cast<uint32_t>(p.P - 1);

  for (int i = lane; i < PQ; i += 32) cnt[i] = 0u;
  {  // the walk's argmax elements, row order (r_lo and the row pitch are multiples of 16 elements)
    const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(p.argmax) +
                                                      (static_cast<size_t>(kcol) * p.ld_arg + r_lo) * ES);
    const int nvec = (nrows * ES + 15) / 16;
    for (int i = lane; i < nvec; i += 32) reinterpret_cast<uint4*>(argw)[i] = __ldg(src + i);
  }
  {  // bf16 pair weight of every text the walk touches (one round of global-load latency), then one per row
    const int m_lo = r_lo / p.T;
    const int rem0 = r_lo - m_lo * p.T;
    const int ntext = (rem0 + nrows - 1) / p.T + 1;
    const float* wcol = p.w + static_cast<size_t>(m_lo) * p.ld_pairs + kcol;
    const float w_s = p.w_scale != nullptr ? __ldg(p.w_scale) : 1.f;
    if (ntext <= rows_max) {                                 // always, unless T == 1 and the walk starts mid-text
      for (int i = lane; i < ntext; i += 32)
        srow[i] = static_cast<uint16_t>(cvt_w16(__ldg(wcol + static_cast<size_t>(i) * p.ld_pairs) * w_s, p.f16));
      __syncwarp();
      const float invT = 1.0f / static_cast<float>(p.T);
      for (int r = lane; r < nrows; r += 32) {
        // floor((rem0 + r) / T): the operand is < 2^13 and at least 0.5/T away from an integer, so the float product is exact enough
        const int tl = p.T <= 2048 ? __float2int_rz((static_cast<float>(rem0 + r) + 0.5f) * invT) : (rem0 + r) / p.T;
        wrow[r] = srow[tl];
      }
    } else {
      for (int r = lane; r < nrows; r += 32)
        wrow[r] = static_cast<uint16_t>(cvt_w16(__ldg(wcol + static_cast<size_t>((rem0 + r) / p.T) * p.ld_pairs) * w_s, p.f16));
    }
  }
  __syncwarp();
  const uint8_t* arg8 = reinterpret_cast<const uint8_t*>(argw);
  const uint16_t* arg16 = reinterpret_cast<const uint16_t*>(argw);

  // ---- pass 1: position of every row inside its patch list (ascending rows), list lengths ----
  const uint32_t lt_mask = (1u << lane) - 1;
  for (int r0 = 0; r0 < nrows; r0 += 32) {
    const int r = r0 + lane;
    const bool valid = r < nrows;
    const uint32_t q = valid ? min(static_cast<uint32_t>(U16 ? arg16[r] : arg8[r]), q_max) : 0xFFFFFFFFu;
    const uint32_t mask = __match_any_sync(0xffffffffu, q);
    const uint32_t base = valid ? cnt[q] : 0u;             // every lane of a group reads the count ...
    __syncwarp();
    const uint32_t rank = __popc(mask & lt_mask);
    if (valid) {
      if (rank == 0) cnt[q] = base + __popc(mask);         // ... before its first lane moves it on
      srow[r] = static_cast<uint16_t>(base + rank);
    }
    __syncwarp();
  }

  // ---- pass 2: steps of every (warp, phase) = longest of its four quarters' lists; exclusive prefix ----
  uint32_t carry = 0;
  for (int p0 = 0; p0 < NPH; p0 += 32) {
    const int ph = p0 + lane;
    uint32_t nst = 0;
    if (ph < NPH) {
      const int w_ = ph / NPO, j = ph - w_ * NPO;
#pragma unroll
      for (int o = 0; o < 4; ++o) nst = max(nst, cnt[(w_ * 4 + o) * NPO + j]);
      nst = (nst + 1u) & ~1u;                               // whole pairs: the consumer has no odd tail
    }
    uint32_t incl = nst;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    if (ph < NPH) pst[ph] = (carry + incl - nst) | (nst << 16);
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  const uint32_t total_steps = carry;
  __syncwarp();
  {  // header
    uint16_t* h16 = reinterpret_cast<uint16_t*>(hdr);
    for (int i = lane; i < g.hdr_bytes / 2; i += 32) {
      const int w_ = i / g.hs, j = i - w_ * g.hs;
      uint32_t v = 0;
      if (w_ < g.nw) {
        if (j < NPO) v = pst[w_ * NPO + j] >> 16;
        else if (j == NPO) v = pst[w_ * NPO] & 0xFFFFu;
      }
      h16[i] = static_cast<uint16_t>(v);
    }
  }
  const bool in_smem = static_cast<int>(total_steps + 2) * 4 <= p.cap_words;
  uint32_t* tab = in_smem ? tab_s : reinterpret_cast<uint32_t*>(blk + g.hdr_bytes);
  for (uint32_t i = lane; i < total_steps + 2; i += 32) reinterpret_cast<uint4*>(tab)[i] = make_uint4(0u, 0u, 0u, 0u);   // + end sentinel
  __syncwarp();

  // ---- pass 3: place the entries ----
  uint32_t slot_off[3];                                     // ring byte offset of the walk's (<= nimg <= 3) sub-tiles
#pragma unroll
  for (int i = 0; i < 3; ++i) slot_off[i] = static_cast<uint32_t>((lo_sub + i) % g.ring) * g.sub_bytes;
  for (int r = lane; r < nrows; r += 32) {
    const uint32_t q = min(static_cast<uint32_t>(U16 ? arg16[r] : arg8[r]), q_max);
    const uint32_t s = srow[r];
    const uint32_t O = q / NPO, j = q - O * NPO;
    const uint32_t ps = pst[(O >> 2) * NPO + j];
    const uint32_t idx = (ps & 0xFFFFu) * 4 + (s >> 1) * 8 + (O & 3u) * 2 + (s & 1u);
    const bool s1 = r >= g.sub_rows, s2 = r >= 2 * g.sub_rows;
    const uint32_t phys = (s2 ? slot_off[2] : s1 ? slot_off[1] : slot_off[0]) +
                          static_cast<uint32_t>(r - (s2 ? 2 * g.sub_rows : s1 ? g.sub_rows : 0)) * 128u;
    tab[idx] = ((phys >> 2) << 16) | wrow[r];
  }
  __syncwarp();
  uint4* dst = reinterpret_cast<uint4*>(blk);
  const int hq = g.hdr_bytes / 16;
  for (int i = lane; i < hq; i += 32) dst[i] = reinterpret_cast<const uint4*>(hdr)[i];
  if (in_smem)
    for (uint32_t i = lane; i < total_steps + 2; i += 32) dst[hq + i] = reinterpret_cast<const uint4*>(tab_s)[i];
  if (lane == 0) p.blk_units[unit] = static_cast<uint16_t>(hq + total_steps + 2);
}

struct Dimg4Params {
  const uint8_t* blocks;
  const uint16_t* blk_units;
  Dimg4Geom g;
  int chunk_k0, chunk_n;
  int P, D;
  float* dimg;
  int accumulate_img;
  const float* w_scale;
};

struct __align__(8) Dimg4Barriers {
  uint64_t sub_full[6];
  uint64_t sub_empty[6];
  uint64_t lst_full[4];
  uint64_t lst_empty[4];
};

__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
// row read of one step; PRED: a padding slot (zero entry) issues no shared-memory request and keeps the stale
// (finite) row registers, whose product with the +0 weight leaves the accumulators untouched
template <bool PRED>
__device__ __forceinline__ void row_load(uint4& v, uint32_t ring_lane, uint32_t e) {
  const uint32_t addr = ring_lane + __byte_perm(e, 0u, 0x4432) * 4u;       // PRMT (high half) + LEA
  if (PRED) {
    asm volatile("{\n\t.reg .pred p;\n\t setp.ne.b32 p, %4, 0;\n\t @p ld.shared.v4.b32 {%0, %1, %2, %3}, [%5];\n\t}"
                 : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w) : "r"(e), "r"(addr));
  } else {
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  }
}
// fhfma8_lo as a volatile asm: keeps both row loads of a pair ahead of the first FMA group
#define CXL_FHFMA8_V(TY)                                                                                               \
  asm volatile("{\n\t.reg .b16 x0, x1, x2, x3, x4, x5, x6, x7, wl, wh;\n\t"                                           \
      "mov.b32 {x0, x1}, %8;\n\t mov.b32 {x2, x3}, %9;\n\t mov.b32 {x4, x5}, %10;\n\t mov.b32 {x6, x7}, %11;\n\t"   \
      "mov.b32 {wl, wh}, %12;\n\t"                                                                                   \
      "fma.rn.f32." TY " %0, x0, wl, %0;\n\t fma.rn.f32." TY " %1, x1, wl, %1;\n\t"                                  \
      "fma.rn.f32." TY " %2, x2, wl, %2;\n\t fma.rn.f32." TY " %3, x3, wl, %3;\n\t"                                  \
      "fma.rn.f32." TY " %4, x4, wl, %4;\n\t fma.rn.f32." TY " %5, x5, wl, %5;\n\t"                                  \
      "fma.rn.f32." TY " %6, x6, wl, %6;\n\t fma.rn.f32." TY " %7, x7, wl, %7;\n\t}"                                 \
      : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7])               \
      : "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(e))
template <bool F16>
__device__ __forceinline__ void fhfma8_lo_v(float (&a)[8], const uint4& v, uint32_t e) {
  if (F16) CXL_FHFMA8_V("f16");
  else CXL_FHFMA8_V("bf16");
}

// Warps [0, NW): consumers (quarter warp o = 4 warp + lane / 8 owns patches [o NPO, o NPO + NPO) of the CTA's NIMG
// images); warp NW issues the loads.  PF: the row reads of pair i+1 are issued before the FMAs of pair i.
template <int NW, int NPO, int NIMG, bool PRED, bool PF, bool F16>
__global__ void __launch_bounds__((NW + 1) * 32, 1)
dimg4_gather_kernel(const __grid_constant__ CUtensorMap tmTxt, const Dimg4Params p) {
  constexpr int DS = 64;
  constexpr int HS = (NPO + 1 + 3) / 4 * 4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  const Dimg4Geom& g = p.g;
  const int RING = g.ring, NB = g.nbuf;
  const uint32_t sub_bytes = g.sub_bytes;
  const uint32_t blk_stride = g.blk_stride;
  Dimg4Barriers* bars = reinterpret_cast<Dimg4Barriers*>(smem + RING * sub_bytes + NB * blk_stride + 64);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int d0 = blockIdx.x * DS;
  const int kl0 = blockIdx.y * NIMG;                     // first image of this CTA inside the chunk
  const int H = g.H, tau_max = g.tau_max;
  const uint32_t ring0 = smem_u32(smem);
  const uint32_t lst0 = ring0 + RING * sub_bytes;
  const uint32_t sub_full0 = smem_u32(&bars->sub_full[0]);
  const uint32_t sub_empty0 = smem_u32(&bars->sub_empty[0]);
  const uint32_t lst_full0 = smem_u32(&bars->lst_full[0]);
  const uint32_t lst_empty0 = smem_u32(&bars->lst_empty[0]);

  if (threadIdx.x == 0) {
    for (int s = 0; s < RING; ++s) {
      mbar_init(sub_full0 + 8 * s, 1);
      mbar_init(sub_empty0 + 8 * s, NW);
    }
    for (int s = 0; s < NB; ++s) {
      mbar_init(lst_full0 + 8 * s, 1);
      mbar_init(lst_empty0 + 8 * s, NW);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmTxt);
  }
  __syncthreads();

  if (warp == NW) {
    if (lane == 0) {
      auto units_of = [&](int tau) -> uint32_t {
        const int kl = kl0 + tau % NIMG;
        return (tau <= tau_max && kl < p.chunk_n) ? p.blk_units[static_cast<size_t>(kl) * g.walks + tau / NIMG] : 0u;
      };
      uint32_t units = units_of(0);
      int slot = 0, lb = 0;                                 // tau % RING, tau % NB
      uint32_t sph = 0, lph = 0;                            // (tau / RING) & 1, (tau / NB) & 1
      for (int tau = 0; tau <= tau_max; ++tau) {
        const uint32_t units_next = units_of(tau + 1);    // its latency hides behind this walk's waits
        mbar_wait(lst_empty0 + 8 * lb, lph ^ 1);
        if (units != 0) {
          const int kl = kl0 + tau % NIMG;
          mbar_arrive_expect_tx(lst_full0 + 8 * lb, units * 16);
          bulk_load_1d(lst0 + lb * blk_stride,
                       p.blocks + (static_cast<size_t>(kl) * g.walks + tau / NIMG) * blk_stride, units * 16,
                       lst_full0 + 8 * lb);
        } else {
          mbar_arrive(lst_full0 + 8 * lb);
        }
        if (tau < H) {
          mbar_wait(sub_empty0 + 8 * slot, sph ^ 1);
          mbar_arrive_expect_tx(sub_full0 + 8 * slot, sub_bytes);
          for (int b = 0; b < g.n_box; ++b)                 // rows past the end: zero-filled, never referenced
            tma_load_2d(ring0 + slot * sub_bytes + b * g.box_rows * 128, &tmTxt, sub_full0 + 8 * slot, d0,
                        tau * g.sub_rows + b * g.box_rows);
        }
        units = units_next;
        if (++slot == RING) { slot = 0; sph ^= 1; }
        if (++lb == NB) { lb = 0; lph ^= 1; }
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
  uint4 va = make_uint4(0u, 0u, 0u, 0u), vb = make_uint4(0u, 0u, 0u, 0u);
  const uint32_t q4 = lane >> 3;
  const uint32_t ring_lane = ring0 + (lane & 7) * 16;
  const uint32_t hdr_off = warp * HS * 2;
  int slot = 0, lb = 0, rel = 0;                            // tau % RING, tau % NB, (tau - NIMG + 1) % RING
  uint32_t sph = 0, lph = 0;

  for (int tau0 = 0; tau0 <= tau_max; tau0 += NIMG) {
#pragma unroll
    for (int c = 0; c < NIMG; ++c) {
      const int tau = tau0 + c;
      if (tau <= tau_max) {
        if (tau < H) mbar_wait(sub_full0 + 8 * slot, sph);
        mbar_wait(lst_full0 + 8 * lb, lph);
        if (kl0 + c < p.chunk_n) {
          const uint32_t lstb = lst0 + lb * blk_stride;
          uint32_t h[HS / 2];
#pragma unroll
          for (int i = 0; i < HS / 4; ++i) {
            const uint2 t = lds64(lstb + hdr_off + i * 8);
            h[2 * i] = t.x;
            h[2 * i + 1] = t.y;
          }
          const uint32_t first = (h[NPO >> 1] >> (16 * (NPO & 1))) & 0xFFFFu;
          uint32_t ep = lstb + g.hdr_bytes + first * 16 + q4 * 8;   // this quarter's slot of the current pair
          uint2 cur = lds64(ep);                                     // (a pair past the warp's last region is read, never used)
          if (PF) {
            row_load<PRED>(va, ring_lane, cur.x);
            row_load<PRED>(vb, ring_lane, cur.y);
          }
#pragma unroll
          for (int j = 0; j < NPO; ++j) {
            const uint32_t nst = (h[j >> 1] >> (16 * (j & 1))) & 0xFFFFu;
            const uint32_t end = ep + nst * 16;                       // nst is even: whole pairs of 32 bytes
#pragma unroll 1
            while (ep != end) {
              ep += 32;
              const uint2 nxt = lds64(ep);
              if (PF) {
                uint4 na, nb;
                row_load<false>(na, ring_lane, nxt.x);
                row_load<false>(nb, ring_lane, nxt.y);
                fhfma8_lo_v<F16>(acc[c][j], va, cur.x);
                fhfma8_lo_v<F16>(acc[c][j], vb, cur.y);
                va = na;
                vb = nb;
              } else {
                row_load<PRED>(va, ring_lane, cur.x);
                row_load<PRED>(vb, ring_lane, cur.y);
                fhfma8_lo_v<F16>(acc[c][j], va, cur.x);
                fhfma8_lo_v<F16>(acc[c][j], vb, cur.y);
              }
              cur = nxt;
            }
          }
        }
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(lst_empty0 + 8 * lb);
          if (tau >= NIMG - 1 && tau - NIMG + 1 < H) mbar_arrive(sub_empty0 + 8 * rel);   // that sub-tile's last walk
        }
        if (tau >= NIMG - 1 && ++rel == RING) rel = 0;
        if (++slot == RING) { slot = 0; sph ^= 1; }
        if (++lb == NB) { lb = 0; lph ^= 1; }
      }
    }
  }
  const int owner = warp * 4 + static_cast<int>(q4);
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
static constexpr int kGatherSlack = 128 /*alignment*/ + 64 /*read-ahead past the last region*/ +
                                    static_cast<int>(sizeof(Dimg4Barriers));

// (consumer warps, patches per quarter warp, images per CTA).  P <= 200/204/208 has several candidates
// (CXL_DIMG_CFG selects; warps are allocated in fours: 18 -> 20 warps = 96 registers per thread, 14 -> 16 = 128,
// 26 -> 28 = 72).
static void dimg4_variant(int P, int* nw, int* npo, int* nimg) {
  const int cfg = env_int("CXL_DIMG_CFG", 0);
  if (P <= 204 && cfg == 1733) { *nw = 17; *npo = 3; *nimg = 3; }
  else if (P <= 204 && (cfg == 0 || cfg == 1732)) { *nw = 17; *npo = 3; *nimg = 2; }
  else if (P <= 200 && cfg == 2522) { *nw = 25; *npo = 2; *nimg = 2; }
  else if (P <= 208 && cfg == 1343) { *nw = 13; *npo = 4; *nimg = 3; }
  else if (P <= 208 && cfg == 1342) { *nw = 13; *npo = 4; *nimg = 2; }
  else if (P <= 304) { *nw = 19; *npo = 4; *nimg = 2; }
  else { *nw = 19; *npo = 8; *nimg = 1; }
}

bool dimg4_covers(int P) { return P <= 608; }

Dimg4Geom dimg4_geom(long long rows, int P, int max_smem) {
  Dimg4Geom g;
  memset(&g, 0, sizeof(g));
  dimg4_variant(P, &g.nw, &g.npo, &g.nimg);
  int depth = env_int("CXL_DIMG_DEPTH", 1);                 // walks of lead time of the sub-tile / step-table loads
  if (depth < 1) depth = 1;
  if (depth > 3) depth = 3;
  g.ring = g.nimg + depth;
  g.nbuf = depth + 1;
  g.hs = (g.npo + 1 + 3) / 4 * 4;
  g.hdr_bytes = round_up_i(g.nw * g.hs * 2, 16);
  // ring of sub-tiles (128 B per row) + nbuf worst-case step tables (4 words per row of a walk, + the end sentinel)
  const int per_row = 128 * g.ring + g.nbuf * 16 * g.nimg;
  int sub = (max_smem - kGatherSlack - g.nbuf * (g.hdr_bytes + 32 + 16 * 2 * g.nw * g.npo)) / per_row;
  const int lim = 65535 / (32 * g.ring);                    // ring offsets >> 2 must fit 16 bits
  if (sub > lim) sub = lim;
  if (static_cast<long long>(sub) > rows) sub = static_cast<int>(rows);
  int n_box = (sub + 255) / 256;
  sub = sub / (16 * n_box) * (16 * n_box);                  // multiple of 16 rows and of the box count
  if (sub < 16 * n_box) { sub = 16; n_box = 1; }
  g.sub_rows = sub;
  g.sub_bytes = sub * 128;
  g.n_box = n_box;
  g.box_rows = sub / n_box;
  g.H = static_cast<int>((rows + sub - 1) / sub);
  g.tau_max = g.H + g.nimg - 2;
  g.walks = g.tau_max / g.nimg + 1;
  // worst case: 4 words per row + one padding step per phase (even step counts) + the two-step end sentinel
  g.blk_stride = g.hdr_bytes + 16 * g.nimg * g.sub_rows + 16 * g.nw * g.npo + 32;
  return g;
}

long long dimg4_bytes_per_image(const Dimg4Geom& g) {
  return static_cast<long long>(g.walks) * (g.blk_stride + 2);
}

template <int NW, int NPO, int NIMG, bool PRED, bool F16>
static int launch_gather4_k(const CUtensorMap& tm, const Dimg4Params& gp, dim3 grid, int smem_bytes, cudaStream_t stream) {
  auto kern = dimg4_gather_kernel<NW, NPO, NIMG, PRED, false, F16>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  kern<<<grid, (NW + 1) * 32, smem_bytes, CXL_COUNTED(stream)>>>(tm, gp);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}
// fp16 operands are built for the default variants only (the CXL_DIMG_CFG experiments are bf16)
template <int NW, int NPO, int NIMG, bool WITH_F16>
static int launch_gather4(const CUtensorMap& tm, const Dimg4Params& gp, dim3 grid, int smem_bytes, bool pred, int f16,
                          cudaStream_t stream) {
  if (f16) {
    if (!WITH_F16) return CXL_ERR_UNSUPPORTED;
    return launch_gather4_k<NW, NPO, NIMG, false, WITH_F16>(tm, gp, grid, smem_bytes, stream);
  }
  if (pred) return launch_gather4_k<NW, NPO, NIMG, true, false>(tm, gp, grid, smem_bytes, stream);
  return launch_gather4_k<NW, NPO, NIMG, false, false>(tm, gp, grid, smem_bytes, stream);
}

// dImg of images [k_off, k_off + Bi) in chunks of `cap` images (scratch = cap * dimg4_bytes_per_image)
int launch_dimg_v4(const void* txt, const void* argmax, long long ld_arg, const float* w, int Bt, int Bi, int T, int P,
                   int D, int ld_pairs, int k_off, float* dimg, int accumulate_img, const Dimg4Geom& g, void* scratch,
                   long long cap, int max_smem, int f16, const float* w_scale, cudaStream_t stream) {
  const int rows = Bt * T;
  const bool u16 = P > 256;
  CUtensorMap tm;
  int rc = make_tmap_plain(&tm, txt, rows, D, 64, g.box_rows, f16);
  if (rc != CXL_OK) return rc;

  Dimg4StepsParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.argmax = argmax; sp.ld_arg = ld_arg; sp.w = w; sp.ld_pairs = ld_pairs;
  sp.T = T; sp.P = P; sp.rows = rows; sp.g = g;
  sp.f16 = f16 ? 1 : 0; sp.w_scale = w_scale;
  const long long units_bytes = (cap * g.walks * 2 + 255) / 256 * 256;
  sp.blk_units = static_cast<uint16_t*>(scratch);
  sp.blocks = static_cast<uint8_t*>(scratch) + units_bytes;
  const int rows_max = g.nimg * g.sub_rows;
  const int fixed_words = g.nw * 4 * g.npo + ((g.nw * g.npo + 3) & ~3) + ((rows_max * (u16 ? 2 : 1) + 15) / 16) * 4 +
                          rows_max /* wrow + srow */ + g.hdr_bytes / 4;
  const int worst_words = (g.blk_stride - g.hdr_bytes) / 4;
  sp.cap_words = worst_words < 2304 ? worst_words : 2304;   // typical tables are ~1.6 words per row; larger ones go straight to global
  sp.per_warp_words = (fixed_words + sp.cap_words + 3) & ~3;
  const int steps_smem = 4 * sp.per_warp_words * 4;

  Dimg4Params gp;
  memset(&gp, 0, sizeof(gp));
  gp.blocks = sp.blocks; gp.blk_units = sp.blk_units; gp.g = g;
  gp.P = P; gp.D = D; gp.dimg = dimg; gp.accumulate_img = accumulate_img; gp.w_scale = w_scale;
  const int smem_bytes = kGatherSlack + g.ring * g.sub_bytes + g.nbuf * g.blk_stride;
  if (smem_bytes > max_smem) return CXL_ERR_UNSUPPORTED;
  const bool pred = env_int("CXL_DIMG_PRED", 0) != 0;

  void (*skern)(const Dimg4StepsParams) = nullptr;
  if (g.npo == 2) skern = u16 ? dimg4_steps_kernel<true, 2> : dimg4_steps_kernel<false, 2>;
  else if (g.npo == 3) skern = u16 ? dimg4_steps_kernel<true, 3> : dimg4_steps_kernel<false, 3>;
  else if (g.npo == 4) skern = u16 ? dimg4_steps_kernel<true, 4> : dimg4_steps_kernel<false, 4>;
  else skern = u16 ? dimg4_steps_kernel<true, 8> : dimg4_steps_kernel<false, 8>;
  CXL_CUDA_CHECK(cudaFuncSetAttribute(skern, cudaFuncAttributeMaxDynamicSharedMemorySize, steps_smem));

  for (int k0 = 0; k0 < Bi; k0 += static_cast<int>(cap)) {
    const int n = (Bi - k0 < cap) ? Bi - k0 : static_cast<int>(cap);
    sp.k_first = k_off + k0;
    sp.n_img = n;
    const long long units = static_cast<long long>(n) * g.walks;
    skern<<<static_cast<unsigned>((units + 3) / 4), 128, steps_smem, CXL_COUNTED(stream)>>>(sp);
    CXL_CUDA_CHECK(cudaGetLastError());
    gp.chunk_k0 = k0;
    gp.chunk_n = n;
    dim3 grid(D / 64, (n + g.nimg - 1) / g.nimg);
    if (g.nw == 17 && g.nimg == 3) rc = launch_gather4<17, 3, 3, false>(tm, gp, grid, smem_bytes, pred, f16, stream);
    else if (g.nw == 17) rc = launch_gather4<17, 3, 2, false>(tm, gp, grid, smem_bytes, pred, f16, stream);
    else if (g.nw == 25) rc = launch_gather4<25, 2, 2, false>(tm, gp, grid, smem_bytes, pred, f16, stream);
    else if (g.nw == 13 && g.nimg == 3) rc = launch_gather4<13, 4, 3, false>(tm, gp, grid, smem_bytes, pred, f16, stream);
    else if (g.nw == 13) rc = launch_gather4<13, 4, 2, false>(tm, gp, grid, smem_bytes, pred, f16, stream);
    else if (g.npo == 4) rc = launch_gather4<19, 4, 2, false>(tm, gp, grid, smem_bytes, pred, f16, stream);
    else rc = launch_gather4<19, 8, 1, true>(tm, gp, grid, smem_bytes, pred, f16, stream);
    if (rc != CXL_OK) return rc;
  }
  return CXL_OK;
}

}  // namespace cxl
