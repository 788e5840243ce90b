// fp32 parity-mode MaxSim forward (SIMT FFMA).  Same contract as the tensor-core kernel but
// with fp32 multiplicands end to end, so results agree with the reference's fp32 CPU path
// to ~1e-6 (north_star "fp32 mode <= 1e-5").  Used for BASELINE config 1 (fp32) and as an
// on-device cross-check of the tcgen05 kernel.  Reference: loss.py:34-44, model.py:626-628.
#include "common.cuh"

namespace cxl {

static constexpr int FM = 64;   // text-token rows per CTA tile
static constexpr int FN = 64;   // patch columns per step
static constexpr int FK = 16;   // k-chunk

struct FwdF32Params {
  const float* txt;   // [Bt*T, D]
  const float* img;   // [Bi*P, D]
  int Bt, Bi, T, P, D;
  int rows_a;
  int imgs_per_cta;
  float* S;
  float* cnt;
  float* S_extra;      // fixed-order sums (see maxsim_fwd_tc.cu): slots 1.. of the per-FM-row-block partials, or null (atomics)
  long long slot_stride;
  void* argmax;
  long long ld_arg;
  int argmax_u16;
};

__global__ void __launch_bounds__(256)
maxsim_fwd_f32_kernel(const FwdF32Params p) {
  __shared__ __align__(16) float As[FK][FM + 4];
  __shared__ __align__(16) float Bs[FK][FN + 4];
  __shared__ float row_best[FM];

  const int tid = threadIdx.x;
  const int tx = tid & 15;        // column group
  const int ty = tid >> 4;        // row group
  const int row0 = blockIdx.x * FM;
  const int k_begin = blockIdx.y * p.imgs_per_cta;
  const int k_end = min(k_begin + p.imgs_per_cta, p.Bi);

  // loader mapping: 64 rows x 16 k -> each thread one float4 along k
  const int lrow = tid >> 2;
  const int lk = (tid & 3) * 4;

  for (int k = k_begin; k < k_end; ++k) {
    float best[4];
    int bidx[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { best[i] = -INFINITY; bidx[i] = 0; }

    for (int c0 = 0; c0 < p.P; c0 += FN) {
      float acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

      for (int d0 = 0; d0 < p.D; d0 += FK) {
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        const int ar = row0 + lrow;
        if (ar < p.rows_a && d0 + lk < p.D)
          a = *reinterpret_cast<const float4*>(p.txt + static_cast<size_t>(ar) * p.D + d0 + lk);
        const int bc = c0 + lrow;
        if (bc < p.P && d0 + lk < p.D)
          b = *reinterpret_cast<const float4*>(p.img + (static_cast<size_t>(k) * p.P + bc) * p.D + d0 + lk);
        __syncthreads();
        As[lk + 0][lrow] = a.x; As[lk + 1][lrow] = a.y; As[lk + 2][lrow] = a.z; As[lk + 3][lrow] = a.w;
        Bs[lk + 0][lrow] = b.x; Bs[lk + 1][lrow] = b.y; Bs[lk + 2][lrow] = b.z; Bs[lk + 3][lrow] = b.w;
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < FK; ++kk) {
          const float4 av = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
          const float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
          const float ar4[4] = {av.x, av.y, av.z, av.w};
          const float br4[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar4[i], br4[j], acc[i][j]);
        }
      }
      // running max / first-index argmax over this 64-column step
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float lb = -INFINITY;
        int li = 0x7fffffff;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int q = c0 + tx * 4 + j;
          if (q < p.P && acc[i][j] > lb) { lb = acc[i][j]; li = q; }
        }
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
          const float ob = __shfl_xor_sync(0xffffffffu, lb, o);
          const int oi = __shfl_xor_sync(0xffffffffu, li, o);
          if (ob > lb || (ob == lb && oi < li)) { lb = ob; li = oi; }
        }
        if (lb > best[i]) { best[i] = lb; bidx[i] = li; }
      }
    }

    __syncthreads();   // previous image's row_best fully consumed
    if (tx == 0) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = ty * 4 + i;
        const int R = row0 + r;
        row_best[r] = best[i];
        if (R < p.rows_a && p.argmax != nullptr) {
          const size_t off = static_cast<size_t>(k) * p.ld_arg + R;
          if (p.argmax_u16) reinterpret_cast<uint16_t*>(p.argmax)[off] = static_cast<uint16_t>(bidx[i]);
          else reinterpret_cast<uint8_t*>(p.argmax)[off] = static_cast<uint8_t>(bidx[i]);
        }
      }
    }
    __syncthreads();
    if (tid < FM) {
      const int R = row0 + tid;
      if (R < p.rows_a) {
        const int m = R / p.T;
        const bool head = (tid == 0) || ((R - 1) / p.T != m);
        if (head) {
          float s = 0.f, nz = 0.f;
          for (int r = tid; r < FM && row0 + r < p.rows_a && (row0 + r) / p.T == m; ++r) {
            const float v = row_best[r];
            s += v;
            nz += (v != 0.f) ? 1.f : 0.f;
          }
          const size_t pi = static_cast<size_t>(m) * p.Bi + k;
          if (p.S_extra != nullptr) {
            const int slot = row0 / FM - (m * p.T) / FM;               // one writer per (slot, text, image)
            (slot == 0 ? p.S : p.S_extra + static_cast<size_t>(slot - 1) * p.slot_stride)[pi] = s;
          } else {
            atomicAdd(p.S + pi, s);
          }
          if (p.cnt != nullptr) atomicAdd(p.cnt + pi, nz);
        }
      }
    }
  }
}

int launch_maxsim_finalize(float* S, const float* cnt, size_t npairs, int T, int mean_mode, cudaStream_t stream);
int launch_maxsim_finalize_slots(float* S, const float* S_extra, const float* cnt, int Bt, int Bi, int T, int slot_rows,
                                 int mean_mode, cudaStream_t stream);
int maxsim_fwd_f32_slot_rows() { return FM; }

int maxsim_fwd_f32(const float* txt, const float* img, int Bt, int Bi, int T, int P, int D, int mean_mode, float* S,
                   float* cnt, void* argmax, long long ld_arg, cudaStream_t stream, float* S_extra) {
  if (D % 4 != 0) return CXL_ERR_UNSUPPORTED;
  if (P > 65535) return CXL_ERR_UNSUPPORTED;
  if (mean_mode == CXL_MEAN_MASKED && cnt == nullptr) return CXL_ERR_INVALID_ARG;
  FwdF32Params p;
  p.txt = txt; p.img = img; p.Bt = Bt; p.Bi = Bi; p.T = T; p.P = P; p.D = D;
  p.rows_a = Bt * T;
  p.S = S; p.cnt = cnt; p.argmax = argmax; p.ld_arg = ld_arg; p.argmax_u16 = P > 256;
  p.S_extra = S_extra; p.slot_stride = static_cast<long long>(Bt) * Bi;
  const int m_tiles = (p.rows_a + FM - 1) / FM;
  const int sms = num_sms();
  int chunks = (4 * sms + m_tiles - 1) / m_tiles;
  if (chunks > Bi) chunks = Bi;
  if (chunks < 1) chunks = 1;
  if (chunks > 65535) chunks = 65535;
  p.imgs_per_cta = (Bi + chunks - 1) / chunks;
  chunks = (Bi + p.imgs_per_cta - 1) / p.imgs_per_cta;
  const size_t npairs = static_cast<size_t>(Bt) * Bi;
  if (S_extra == nullptr) CXL_CUDA_CHECK(cudaMemsetAsync(S, 0, npairs * sizeof(float), stream));
  if (cnt != nullptr) CXL_CUDA_CHECK(cudaMemsetAsync(cnt, 0, npairs * sizeof(float), stream));
  dim3 grid(m_tiles, chunks);
  maxsim_fwd_f32_kernel<<<grid, 256, 0, CXL_COUNTED(stream)>>>(p);
  CXL_CUDA_CHECK(cudaGetLastError());
  if (S_extra != nullptr) return launch_maxsim_finalize_slots(S, S_extra, cnt, Bt, Bi, T, FM, mean_mode, stream);
  return launch_maxsim_finalize(S, cnt, npairs, T, mean_mode, stream);
}

}  // namespace cxl
