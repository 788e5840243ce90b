// L2 normalisation of token embeddings fused with the cast to the tensor-core operand dtype.
//
// Replaces F.normalize(tokens, dim=-1) at the end of ColXLIP.encode_image / encode_text (model.py:551-554,
// 605-607) plus the autocast copy torch.einsum makes of its result (loss.py:34): one pass reads the head's
// output row, writes y = x / max(||x||, 1e-12) directly as bf16 (or f32) and keeps 1/max(||x||, eps) for the
// backward  dx = (g - y * <g, y>) / max(||x||, eps)   (rows with ||x|| < eps: dx = g / eps, as torch's clamp_min).
// HBM-bound: rows * D * (in_bytes + out_bytes); one warp per row, 16-byte accesses.
#include "common.cuh"

namespace cxl {

template <typename T> struct Row8;
template <> struct Row8<float> {
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
template <> struct Row8<__nv_bfloat16> {
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

template <> struct Row8<__half> {
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

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr float kNormEps = 1e-12f;   // F.normalize default

template <typename TI, typename TO>
__global__ void __launch_bounds__(256) l2norm_fwd_kernel(const TI* __restrict__ x, long long rows, int D,
                                                         TO* __restrict__ y, float* __restrict__ inv_norm) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  const long long nwarps = static_cast<long long>(gridDim.x) * 8;
  for (long long r = warp0; r < rows; r += nwarps) {
    const TI* xr = x + r * D;
    float ss = 0.f;
    for (int c = lane * 8; c < D; c += 256) {
      float v[8];
      Row8<TI>::load(xr + c, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) ss = fmaf(v[i], v[i], ss);
    }
    ss = warp_sum(ss);
    const float inv = 1.f / fmaxf(sqrtf(ss), kNormEps);
    TO* yr = y + r * D;
    for (int c = lane * 8; c < D; c += 256) {        // second read of the row comes from L1
      float v[8];
      Row8<TI>::load(xr + c, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] *= inv;
      Row8<TO>::store(yr + c, v);
    }
    if (lane == 0 && inv_norm != nullptr) inv_norm[r] = inv;
  }
}

template <typename TY, typename TG, typename TO>
__global__ void __launch_bounds__(256) l2norm_bwd_kernel(const TY* __restrict__ y, const TG* __restrict__ g,
                                                         const float* __restrict__ inv_norm, long long rows, int D,
                                                         TO* __restrict__ dx) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  const long long nwarps = static_cast<long long>(gridDim.x) * 8;
  for (long long r = warp0; r < rows; r += nwarps) {
    const TY* yr = y + r * D;
    const TG* gr = g + r * D;
    float dot = 0.f;
    for (int c = lane * 8; c < D; c += 256) {
      float a[8], b[8];
      Row8<TY>::load(yr + c, a);
      Row8<TG>::load(gr + c, b);
#pragma unroll
      for (int i = 0; i < 8; ++i) dot = fmaf(a[i], b[i], dot);
    }
    dot = warp_sum(dot);
    const float inv = inv_norm[r];
    TO* o = dx + r * D;
    for (int c = lane * 8; c < D; c += 256) {
      float a[8], b[8];
      Row8<TY>::load(yr + c, a);
      Row8<TG>::load(gr + c, b);
#pragma unroll
      for (int i = 0; i < 8; ++i) b[i] = (b[i] - a[i] * dot) * inv;
      Row8<TO>::store(o + c, b);
    }
  }
}

static unsigned l2norm_grid(long long rows) {
  long long blocks = (rows + 7) / 8;
  const long long cap = static_cast<long long>(num_sms()) * 16;   // grid-stride beyond 16 CTAs/SM worth of rows
  if (blocks > cap) blocks = cap;
  return static_cast<unsigned>(blocks < 1 ? 1 : blocks);
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int l2norm_fwd(const void* x, int x_dtype, long long rows, int D, void* y, int y_dtype, float* inv_norm,
               cudaStream_t stream) {
  if (rows < 0 || D <= 0 || D % 8 != 0) return CXL_ERR_INVALID_ARG;
  if (!aligned16(x) || !aligned16(y)) return CXL_ERR_INVALID_ARG;
  if (rows == 0) return CXL_OK;
  const unsigned grid = l2norm_grid(rows);
  using bf = __nv_bfloat16;
#define CXL_L2F(TI, TO)                                                                                         \
  l2norm_fwd_kernel<TI, TO><<<grid, 256, 0, CXL_COUNTED(stream)>>>(static_cast<const TI*>(x), rows, D,           \
                                                                   static_cast<TO*>(y), inv_norm)
  if (x_dtype == CXL_DTYPE_F32 && y_dtype == CXL_DTYPE_BF16) CXL_L2F(float, bf);
  else if (x_dtype == CXL_DTYPE_F32 && y_dtype == CXL_DTYPE_F32) CXL_L2F(float, float);
  else if (x_dtype == CXL_DTYPE_BF16 && y_dtype == CXL_DTYPE_BF16) CXL_L2F(bf, bf);
  else if (x_dtype == CXL_DTYPE_BF16 && y_dtype == CXL_DTYPE_F32) CXL_L2F(bf, float);
  else if (x_dtype == CXL_DTYPE_F32 && y_dtype == CXL_DTYPE_F16) CXL_L2F(float, __half);      // fp16 autocast (--precision amp)
  else if (x_dtype == CXL_DTYPE_F16 && y_dtype == CXL_DTYPE_F16) CXL_L2F(__half, __half);
  else if (x_dtype == CXL_DTYPE_F16 && y_dtype == CXL_DTYPE_F32) CXL_L2F(__half, float);
  else return CXL_ERR_INVALID_ARG;
#undef CXL_L2F
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

template <typename TY, typename TG>
static int l2norm_bwd_out(const void* y, const void* g, const float* inv_norm, long long rows, int D, void* dx,
                          int dx_dtype, unsigned grid, cudaStream_t stream) {
  if (dx_dtype == CXL_DTYPE_F32)
    l2norm_bwd_kernel<TY, TG, float><<<grid, 256, 0, CXL_COUNTED(stream)>>>(
        static_cast<const TY*>(y), static_cast<const TG*>(g), inv_norm, rows, D, static_cast<float*>(dx));
  else if (dx_dtype == CXL_DTYPE_BF16)
    l2norm_bwd_kernel<TY, TG, __nv_bfloat16><<<grid, 256, 0, CXL_COUNTED(stream)>>>(
        static_cast<const TY*>(y), static_cast<const TG*>(g), inv_norm, rows, D, static_cast<__nv_bfloat16*>(dx));
  else if (dx_dtype == CXL_DTYPE_F16)
    l2norm_bwd_kernel<TY, TG, __half><<<grid, 256, 0, CXL_COUNTED(stream)>>>(
        static_cast<const TY*>(y), static_cast<const TG*>(g), inv_norm, rows, D, static_cast<__half*>(dx));
  else
    return CXL_ERR_INVALID_ARG;
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

int l2norm_bwd(const void* y, int y_dtype, const void* g, int g_dtype, const float* inv_norm, long long rows, int D,
               void* dx, int dx_dtype, cudaStream_t stream) {
  if (rows < 0 || D <= 0 || D % 8 != 0 || inv_norm == nullptr) return CXL_ERR_INVALID_ARG;
  if (!aligned16(y) || !aligned16(g) || !aligned16(dx)) return CXL_ERR_INVALID_ARG;
  if (rows == 0) return CXL_OK;
  const unsigned grid = l2norm_grid(rows);
  using bf = __nv_bfloat16;
  if (y_dtype == CXL_DTYPE_BF16 && g_dtype == CXL_DTYPE_BF16)
    return l2norm_bwd_out<bf, bf>(y, g, inv_norm, rows, D, dx, dx_dtype, grid, stream);
  if (y_dtype == CXL_DTYPE_BF16 && g_dtype == CXL_DTYPE_F32)
    return l2norm_bwd_out<bf, float>(y, g, inv_norm, rows, D, dx, dx_dtype, grid, stream);
  if (y_dtype == CXL_DTYPE_F32 && g_dtype == CXL_DTYPE_F32)
    return l2norm_bwd_out<float, float>(y, g, inv_norm, rows, D, dx, dx_dtype, grid, stream);
  if (y_dtype == CXL_DTYPE_F32 && g_dtype == CXL_DTYPE_BF16)
    return l2norm_bwd_out<float, bf>(y, g, inv_norm, rows, D, dx, dx_dtype, grid, stream);
  if (y_dtype == CXL_DTYPE_F16 && g_dtype == CXL_DTYPE_F16)
    return l2norm_bwd_out<__half, __half>(y, g, inv_norm, rows, D, dx, dx_dtype, grid, stream);
  if (y_dtype == CXL_DTYPE_F16 && g_dtype == CXL_DTYPE_F32)
    return l2norm_bwd_out<__half, float>(y, g, inv_norm, rows, D, dx, dx_dtype, grid, stream);
  return CXL_ERR_INVALID_ARG;
}

}  // namespace cxl
