// K2: MaxSim backward driven by the saved argmax (no dS[Bt,Bi,T,P] one-hot tensor, no GEMMs).
//
// Replaces autograd of loss.py:34-44 in the reference (max -> scatter into zeros, then two bmm):
//   w[m,k]      = dS[m,k] * g / (cnt[m,k] + 1e-8)          (masked mean)   or  dS[m,k] * g / T
//   dTxt[m,n,:] = sum_k w[m,k] * img[k, q*(m,k,n), :]
//   dImg[k,q,:] = sum_{m,n : q*(m,k,n) = q} w[m,k] * txt[m,n,:]
// Both are sparse products with exactly Bt*Bi*T non-zeros; the work is gather/scatter bound.
#include "common.cuh"

namespace cxl {

template <typename T> struct Vec;
template <> struct Vec<float> {
  static constexpr int N = 4;
  __device__ static void load(const float* p, float (&o)[4]) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
  }
};
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const __nv_bfloat16* p, float (&o)[8]) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      o[2 * i] = __uint_as_float(w[i] << 16);
      o[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};
template <> struct Vec<__half> {
  static constexpr int N = 8;
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
};
__device__ __forceinline__ float to_f32(__half x) { return __half2float(x); }
__device__ __forceinline__ float to_f32(float x) { return x; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }

struct BwdParams {
  const void* txt;      // [Bt*T, D]
  const void* img;      // [Bi*p_stride, D]
  const void* argmax;   // [k_off + Bi][ld_arg] image-major
  long long ld_arg;
  const float* w;       // [Bt, ld_pairs] per-pair weight dL/dS / cnt (cxl_pair_weights)
  int Bt, Bi, T, P, D, p_stride;
  int ld_pairs;         // row stride (in pairs) of dS / cnt / argmax: total images of the forward
  int k_off;            // first image of this call inside that row
  int argmax_u16;
  float* dtxt;          // [Bt*T, D]
  float* dimg;          // [Bi*P, D]
  int accumulate_img;
  int accumulate_txt;
};

__device__ __forceinline__ float pair_weight(const BwdParams& p, size_t pair) { return p.w[pair]; }
__device__ __forceinline__ int load_arg(const BwdParams& p, size_t off) {
  return p.argmax_u16 ? static_cast<int>(reinterpret_cast<const uint16_t*>(p.argmax)[off])
                      : static_cast<int>(reinterpret_cast<const uint8_t*>(p.argmax)[off]);
}

// ---- dTxt: one warp per text-token row and D-slice; lanes own Vec::N consecutive channels ----
template <typename T>
__global__ void __launch_bounds__(256) maxsim_bwd_dtxt_kernel(const BwdParams p) {
  constexpr int V = Vec<T>::N;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const long long R = static_cast<long long>(blockIdx.x) * 8 + warp;
  if (R >= static_cast<long long>(p.Bt) * p.T) return;
  const int d = (blockIdx.y * 32 + lane) * V;
  const bool d_ok = d < p.D;
  const int m = static_cast<int>(R / p.T);
  const int n = static_cast<int>(R - static_cast<long long>(m) * p.T);
  const T* img = reinterpret_cast<const T*>(p.img);
  const size_t pair0 = static_cast<size_t>(m) * p.ld_pairs + p.k_off;

  float acc[V];
#pragma unroll
  for (int i = 0; i < V; ++i) acc[i] = 0.f;

  for (int k0 = 0; k0 < p.Bi; k0 += 32) {
    const int kk = k0 + lane;
    int qv = 0;
    float wv = 0.f;
    if (kk < p.Bi) {
      qv = load_arg(p, static_cast<size_t>(p.k_off + kk) * p.ld_arg + R);
      wv = pair_weight(p, pair0 + kk);
    }
    const int cnt_k = min(32, p.Bi - k0);
#pragma unroll 4
    for (int j = 0; j < cnt_k; ++j) {
      const int q = __shfl_sync(0xffffffffu, qv, j);
      const float w = __shfl_sync(0xffffffffu, wv, j);
      if (d_ok) {
        float x[V];
        Vec<T>::load(img + (static_cast<size_t>(k0 + j) * p.p_stride + q) * p.D + d, x);
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] = fmaf(w, x[i], acc[i]);
      }
    }
  }
  if (d_ok) {
    float* out = p.dtxt + static_cast<size_t>(R) * p.D + d;
#pragma unroll
    for (int i = 0; i < V; i += 4) {
      float4 v = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
      if (p.accumulate_txt) {
        const float4 o = *reinterpret_cast<const float4*>(out + i);
        v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
      }
      *reinterpret_cast<float4*>(out + i) = v;
    }
  }
}

// ---- dImg: one CTA per (image, D-slice); fp32 accumulators [P, DS] in shared memory.  Warp w owns
// channels [32w, 32w+32) of the slice for every patch, so no two threads ever touch the same
// accumulator and plain read-modify-write is race free. ----
template <typename T>
__global__ void __launch_bounds__(128) maxsim_bwd_dimg_kernel(const BwdParams p, int DS) {
  extern __shared__ float acc[];   // [P][DS]
  const int k = blockIdx.x;
  const int lane = threadIdx.x & 31;
  const int c = threadIdx.x;                 // channel within slice (blockDim.x == DS)
  const int d = blockIdx.y * DS + c;
  const bool d_ok = d < p.D;
  const T* txt = reinterpret_cast<const T*>(p.txt);

  for (int i = threadIdx.x; i < p.P * DS; i += blockDim.x) acc[i] = 0.f;
  __syncthreads();

  for (int m = 0; m < p.Bt; ++m) {
    const size_t pair = static_cast<size_t>(m) * p.ld_pairs + p.k_off + k;
    const float w = pair_weight(p, pair);
    const size_t abase = static_cast<size_t>(p.k_off + k) * p.ld_arg + static_cast<size_t>(m) * p.T;
    const T* trow = txt + static_cast<size_t>(m) * p.T * p.D + d;
    for (int n0 = 0; n0 < p.T; n0 += 32) {
      const int nn = n0 + lane;
      const int qv = (nn < p.T) ? load_arg(p, abase + nn) : 0;
      const int cnt_n = min(32, p.T - n0);
#pragma unroll 4
      for (int j = 0; j < cnt_n; ++j) {
        const int q = __shfl_sync(0xffffffffu, qv, j);
        if (d_ok) {
          const float x = to_f32(trow[static_cast<size_t>(n0 + j) * p.D]);
          acc[q * DS + c] = fmaf(w, x, acc[q * DS + c]);
        }
      }
    }
  }
  __syncthreads();
  if (d_ok) {
    for (int q = 0; q < p.P; ++q) {
      float* o = p.dimg + (static_cast<size_t>(k) * p.P + q) * p.D + d;
      const float v = acc[q * DS + c];
      *o = p.accumulate_img ? (*o + v) : v;
    }
  }
}

int maxsim_bwd_stream(const void* txt, const void* img, const void* argmax, long long ld_arg, long long arg_row_off,
                      const float* w, int Bt, int Bi, int T, int P, int D, int p_stride, int ld_pairs, int k_off,
                      float* dtxt, int accumulate_txt, float* dimg, int accumulate_img, void* ws, size_t ws_bytes,
                      int f16, const float* w_scale, cudaStream_t stream);

// arg_row_off != 0: txt / w / dtxt point at a slice of Bt texts whose first flat text-token row is `arg_row_off`
// inside the forward's argmax rows (pitch ld_arg); only dTxt can be computed from a row slice.
int maxsim_bwd(const void* txt, const void* img, const void* argmax, long long ld_arg, long long arg_row_off,
               const float* w, int Bt, int Bi, int T, int P, int D, int dtype, int p_stride, int ld_pairs, int k_off,
               float* dtxt, int accumulate_txt, float* dimg, int accumulate_img, int force_generic, void* ws,
               size_t ws_bytes, const float* w_scale, cudaStream_t stream) {
  if (ld_pairs < Bi + k_off || k_off < 0) return CXL_ERR_INVALID_ARG;
  if (arg_row_off < 0 || ld_arg < arg_row_off + static_cast<long long>(Bt) * T) return CXL_ERR_INVALID_ARG;
  if (arg_row_off != 0 && dimg != nullptr) return CXL_ERR_INVALID_ARG;
  const bool half16 = dtype == CXL_DTYPE_BF16 || dtype == CXL_DTYPE_F16;
  if (half16 && !force_generic) {
    // fast path: TMA-staged streaming kernels; shapes it does not cover fall through to the generic kernels
    const int rc = maxsim_bwd_stream(txt, img, argmax, ld_arg, arg_row_off, w, Bt, Bi, T, P, D, p_stride, ld_pairs,
                                     k_off, dtxt, accumulate_txt, dimg, accumulate_img, ws, ws_bytes,
                                     dtype == CXL_DTYPE_F16, w_scale, stream);
    if (rc != CXL_ERR_UNSUPPORTED) return rc;
  }
  // (the generic kernels multiply with the fp32 pair weight: w_scale does not apply)
  const int V = half16 ? 8 : 4;
  if (D % V != 0) return CXL_ERR_UNSUPPORTED;
  BwdParams p;
  p.txt = txt; p.img = img; p.ld_arg = ld_arg; p.w = w;
  p.argmax = static_cast<const uint8_t*>(argmax) + arg_row_off * (P > 256 ? 2 : 1);
  p.Bt = Bt; p.Bi = Bi; p.T = T; p.P = P; p.D = D; p.p_stride = p_stride;
  p.ld_pairs = ld_pairs; p.k_off = k_off;
  p.argmax_u16 = P > 256;
  p.dtxt = dtxt; p.dimg = dimg; p.accumulate_img = accumulate_img; p.accumulate_txt = accumulate_txt;

  if (dtxt != nullptr) {
    const long long rows = static_cast<long long>(Bt) * T;
    dim3 grid(static_cast<unsigned>((rows + 7) / 8), (D + 32 * V - 1) / (32 * V));
    if (dtype == CXL_DTYPE_BF16) maxsim_bwd_dtxt_kernel<__nv_bfloat16><<<grid, 256, 0, CXL_COUNTED(stream)>>>(p);
    else if (dtype == CXL_DTYPE_F16) maxsim_bwd_dtxt_kernel<__half><<<grid, 256, 0, CXL_COUNTED(stream)>>>(p);
    else maxsim_bwd_dtxt_kernel<float><<<grid, 256, 0, CXL_COUNTED(stream)>>>(p);
    CXL_CUDA_CHECK(cudaGetLastError());
  }
  if (dimg != nullptr) {
    int DS = 128;
    while (DS > 32 && static_cast<size_t>(P) * DS * 4 > 200 * 1024) DS >>= 1;
    if (D < DS) DS = D >= 64 ? 64 : 32;
    const size_t smem = static_cast<size_t>(P) * DS * 4;
    if (smem > 220 * 1024) return CXL_ERR_UNSUPPORTED;
    dim3 grid(Bi, (D + DS - 1) / DS);
    if (dtype == CXL_DTYPE_F16) {
      CXL_CUDA_CHECK(cudaFuncSetAttribute(maxsim_bwd_dimg_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          static_cast<int>(smem)));
      maxsim_bwd_dimg_kernel<__half><<<grid, DS, smem, CXL_COUNTED(stream)>>>(p, DS);
    } else if (dtype == CXL_DTYPE_BF16) {
      CXL_CUDA_CHECK(cudaFuncSetAttribute(maxsim_bwd_dimg_kernel<__nv_bfloat16>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
      maxsim_bwd_dimg_kernel<__nv_bfloat16><<<grid, DS, smem, CXL_COUNTED(stream)>>>(p, DS);
    } else {
      CXL_CUDA_CHECK(cudaFuncSetAttribute(maxsim_bwd_dimg_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          static_cast<int>(smem)));
      maxsim_bwd_dimg_kernel<float><<<grid, DS, smem, CXL_COUNTED(stream)>>>(p, DS);
    }
    CXL_CUDA_CHECK(cudaGetLastError());
  }
  return CXL_OK;
}

}  // namespace cxl
