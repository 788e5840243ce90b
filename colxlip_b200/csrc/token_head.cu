// Producer side of the MaxSim operands: the reference's token heads
//     nn.Sequential(LayerNorm(width), Linear(width, D), GELU(), LayerNorm(D))        model.py:518-530
// preceded, for text, by zeroing the tokens at and after the EOT position (model.py:580-593), followed by
// F.normalize(dim=-1) (model.py:551-554, 605-607) and the autocast copy einsum makes of the result (loss.py:34).
// Three launches instead of ~12 elementwise / reduction kernels around a library GEMM:
//   1. head_ln_in_kernel   x [rows, W] -> LayerNorm (masked rows see a zero input) -> 16-bit GEMM operand      (HBM)
//   2. cxl_gemm_nt         [rows, W] x [D, W]^T -> fp32 [rows, D] on tcgen05                                    (tensor)
//   3. head_out_kernel     + bias -> exact GELU -> LayerNorm -> L2 normalise -> cast: K1's operand layout       (HBM)
// One warp per row, 16-byte accesses; statistics in fp32 (two-pass variance on registers).
#include "common.cuh"

namespace cxl {

namespace {

template <typename T> struct Ld8;
template <> struct Ld8<float> {
  __device__ static void load(const float* p, float (&o)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p));
    const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
  }
  __device__ static void store(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *(reinterpret_cast<float4*>(p) + 1) = make_float4(v[4], v[5], v[6], v[7]);
  }
};
template <> struct Ld8<__nv_bfloat16> {
  __device__ static void load(const __nv_bfloat16* p, float (&o)[8]) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      o[2 * i] = __uint_as_float(w[i] << 16);
      o[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <> struct Ld8<__half> {
  __device__ static void load(const __half* p, float (&o)[8]) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      o[2 * i] = f.x;
      o[2 * i + 1] = f.y;
    }
  }
  __device__ static void store(__half* p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __half2 h = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int kMaxVec = 4;     // 32 lanes x 8 channels x 4 = rows of up to 1024 channels stay in registers

// y = LayerNorm(x) * gamma + beta over `width` channels (biased variance, eps inside the sqrt, as nn.LayerNorm);
// rows with (row % T) >= valid_len[row / T] are computed from a ZERO input (model.py:589-591), i.e. y = beta.
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) head_ln_in_kernel(const TI* __restrict__ x, long long rows, int width,
                                                         const float* __restrict__ gamma, const float* __restrict__ beta,
                                                         float eps, const int* __restrict__ valid_len, int T,
                                                         TO* __restrict__ y) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  const long long nwarps = static_cast<long long>(gridDim.x) * 8;
  const int nvec = width >> 8;                 // 256 channels per warp pass
  for (long long r = warp0; r < rows; r += nwarps) {
    const bool masked = valid_len != nullptr && static_cast<int>(r % T) >= __ldg(valid_len + r / T);
    float v[kMaxVec][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i) {
      if (i < nvec) {
        if (masked) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[i][j] = 0.f;
        } else {
          Ld8<TI>::load(x + r * width + i * 256 + lane * 8, v[i]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[i][j];
      }
    }
    const float mean = wsum(s) / static_cast<float>(width);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i)
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[i][j] - mean;
          q += d * d;
        }
      }
    const float rstd = rsqrtf(wsum(q) / static_cast<float>(width) + eps);
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i)
      if (i < nvec) {
        float g[8], b[8], o[8];
        Ld8<float>::load(gamma + i * 256 + lane * 8, g);
        Ld8<float>::load(beta + i * 256 + lane * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[j] + b[j];
        Ld8<TO>::store(y + r * width + i * 256 + lane * 8, o);
      }
  }
}

// y = [L2 normalise] LayerNorm(GELU(h + bias)) * gamma + beta, h [rows, D] fp32 (the GEMM's output)
template <typename TO>
__global__ void __launch_bounds__(256) head_out_kernel(const float* __restrict__ h, long long rows, int D,
                                                       const float* __restrict__ bias, const float* __restrict__ gamma,
                                                       const float* __restrict__ beta, float eps, int l2_normalize,
                                                       TO* __restrict__ y) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  const long long nwarps = static_cast<long long>(gridDim.x) * 8;
  const int nvec = D >> 8;
  for (long long r = warp0; r < rows; r += nwarps) {
    float v[kMaxVec][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i)
      if (i < nvec) {
        float b[8];
        Ld8<float>::load(h + r * D + i * 256 + lane * 8, v[i]);
        Ld8<float>::load(bias + i * 256 + lane * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float t = v[i][j] + b[j];
          v[i][j] = 0.5f * t * (1.f + erff(t * 0.70710678118654752f));      // nn.GELU(), approximate='none'
          s += v[i][j];
        }
      }
    const float mean = wsum(s) / static_cast<float>(D);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i)
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[i][j] - mean;
          q += d * d;
        }
      }
    const float rstd = rsqrtf(wsum(q) / static_cast<float>(D) + eps);
    float n2 = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i)
      if (i < nvec) {
        float g[8], b[8];
        Ld8<float>::load(gamma + i * 256 + lane * 8, g);
        Ld8<float>::load(beta + i * 256 + lane * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[i][j] = (v[i][j] - mean) * rstd * g[j] + b[j];
          n2 += v[i][j] * v[i][j];
        }
      }
    const float inv = l2_normalize ? 1.f / fmaxf(sqrtf(wsum(n2)), 1e-12f) : 1.f;     // F.normalize eps
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i)
      if (i < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = v[i][j] * inv;
        Ld8<TO>::store(y + r * D + i * 256 + lane * 8, o);
      }
  }
}

unsigned head_grid(long long rows) {
  long long blocks = (rows + 7) / 8;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  return static_cast<unsigned>(blocks < 1 ? 1 : blocks);
}

bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

int head_ln_in(const void* x, int x_dtype, long long rows, int width, const float* gamma, const float* beta, float eps,
               const int* valid_len, int T, void* y, int y_dtype, cudaStream_t stream) {
  if (rows < 0 || width <= 0 || width % 256 != 0 || width > 256 * kMaxVec) return CXL_ERR_UNSUPPORTED;
  if (!al16(x) || !al16(y) || !al16(gamma) || !al16(beta)) return CXL_ERR_INVALID_ARG;
  if (valid_len != nullptr && T <= 0) return CXL_ERR_INVALID_ARG;
  if (rows == 0) return CXL_OK;
  const unsigned grid = head_grid(rows);
  using bf = __nv_bfloat16;
#define CXL_HLN(TI, TO)                                                                                              \
  head_ln_in_kernel<TI, TO><<<grid, 256, 0, CXL_COUNTED(stream)>>>(static_cast<const TI*>(x), rows, width, gamma, beta, \
                                                                   eps, valid_len, T, static_cast<TO*>(y))
  if (x_dtype == CXL_DTYPE_F32 && y_dtype == CXL_DTYPE_BF16) CXL_HLN(float, bf);
  else if (x_dtype == CXL_DTYPE_BF16 && y_dtype == CXL_DTYPE_BF16) CXL_HLN(bf, bf);
  else if (x_dtype == CXL_DTYPE_F32 && y_dtype == CXL_DTYPE_F16) CXL_HLN(float, __half);
  else if (x_dtype == CXL_DTYPE_F16 && y_dtype == CXL_DTYPE_F16) CXL_HLN(__half, __half);
  else return CXL_ERR_INVALID_ARG;
#undef CXL_HLN
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

int head_out(const float* h, long long rows, int D, const float* bias, const float* gamma, const float* beta, float eps,
             int l2_normalize, void* y, int y_dtype, cudaStream_t stream) {
  if (rows < 0 || D <= 0 || D % 256 != 0 || D > 256 * kMaxVec) return CXL_ERR_UNSUPPORTED;
  if (!al16(h) || !al16(y) || !al16(bias) || !al16(gamma) || !al16(beta)) return CXL_ERR_INVALID_ARG;
  if (rows == 0) return CXL_OK;
  const unsigned grid = head_grid(rows);
#define CXL_HOUT(TO)                                                                                                  \
  head_out_kernel<TO><<<grid, 256, 0, CXL_COUNTED(stream)>>>(h, rows, D, bias, gamma, beta, eps, l2_normalize,          \
                                                             static_cast<TO*>(y))
  if (y_dtype == CXL_DTYPE_BF16) CXL_HOUT(__nv_bfloat16);
  else if (y_dtype == CXL_DTYPE_F16) CXL_HOUT(__half);
  else if (y_dtype == CXL_DTYPE_F32) CXL_HOUT(float);
  else return CXL_ERR_INVALID_ARG;
#undef CXL_HOUT
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

}  // namespace cxl
