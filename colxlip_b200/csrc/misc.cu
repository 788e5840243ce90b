// Small supporting kernels: bf16 pack/cast of token embeddings and retrieval rank search.
#include "common.cuh"

namespace cxl {

// src [items, n, D] (f32, bf16 or f16, arbitrary item/row strides, unit channel stride)
//   -> dst bf16 or f16 [items, n_pad, D] contiguous; rows n..n_pad-1 are zero.
__device__ __forceinline__ float pk_to_f32(float x) { return x; }
__device__ __forceinline__ float pk_to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
__device__ __forceinline__ float pk_to_f32(__half x) { return __half2float(x); }
__device__ __forceinline__ uint32_t pk_pair(float a, float b, __nv_bfloat16*) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t pk_pair(float a, float b, __half*) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

template <typename T, typename O>
__global__ void pack_operand_kernel(const T* __restrict__ src, long long item_stride, long long row_stride, int items,
                                    int n, int n_pad, int D, O* __restrict__ dst) {
  const int dq = D >> 2;   // groups of 4 channels
  const size_t total = static_cast<size_t>(items) * n_pad * dq;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c4 = static_cast<int>(i % dq);
    const size_t rr = i / dq;
    const int r = static_cast<int>(rr % n_pad);
    const size_t it = rr / n_pad;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (r < n) {
      const T* s = src + it * item_stride + static_cast<long long>(r) * row_stride + c4 * 4;
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = pk_to_f32(s[j]);
    }
    uint2 o;
    o.x = pk_pair(v[0], v[1], static_cast<O*>(nullptr));
    o.y = pk_pair(v[2], v[3], static_cast<O*>(nullptr));
    *reinterpret_cast<uint2*>(dst + (rr * D + c4 * 4)) = o;
  }
}

template <typename O>
static int pack_to(const void* src, int src_dtype, long long item_stride, long long row_stride, int items, int n,
                   int n_pad, int D, O* dst, unsigned blocks, cudaStream_t stream) {
  if (src_dtype == CXL_DTYPE_F32)
    pack_operand_kernel<float, O><<<blocks, 256, 0, CXL_COUNTED(stream)>>>(
        reinterpret_cast<const float*>(src), item_stride, row_stride, items, n, n_pad, D, dst);
  else if (src_dtype == CXL_DTYPE_BF16)
    pack_operand_kernel<__nv_bfloat16, O><<<blocks, 256, 0, CXL_COUNTED(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(src), item_stride, row_stride, items, n, n_pad, D, dst);
  else if (src_dtype == CXL_DTYPE_F16)
    pack_operand_kernel<__half, O><<<blocks, 256, 0, CXL_COUNTED(stream)>>>(
        reinterpret_cast<const __half*>(src), item_stride, row_stride, items, n, n_pad, D, dst);
  else
    return CXL_ERR_INVALID_ARG;
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

int pack_operand(const void* src, int src_dtype, long long item_stride, long long row_stride, int items, int n,
                 int n_pad, int D, void* dst, int dst_dtype, cudaStream_t stream) {
  if (D % 4 != 0 || n_pad < n || items < 0) return CXL_ERR_INVALID_ARG;
  const size_t total = static_cast<size_t>(items) * n_pad * (D >> 2);
  if (total == 0) return CXL_OK;
  size_t blocks = (total + 255) / 256;
  const size_t cap = static_cast<size_t>(num_sms()) * 32;
  if (blocks > cap) blocks = cap;
  if (dst_dtype == CXL_DTYPE_BF16)
    return pack_to(src, src_dtype, item_stride, row_stride, items, n, n_pad, D, reinterpret_cast<__nv_bfloat16*>(dst),
                   static_cast<unsigned>(blocks), stream);
  if (dst_dtype == CXL_DTYPE_F16)
    return pack_to(src, src_dtype, item_stride, row_stride, items, n, n_pad, D, reinterpret_cast<__half*>(dst),
                   static_cast<unsigned>(blocks), stream);
  return CXL_ERR_INVALID_ARG;
}

int pack_bf16(const void* src, int src_dtype, long long item_stride, long long row_stride, int items, int n, int n_pad,
              int D, void* dst, cudaStream_t stream) {
  return pack_operand(src, src_dtype, item_stride, row_stride, items, n, n_pad, D, dst, CXL_DTYPE_BF16, stream);
}

// w = dS * g / (cnt + 1e-8)   (masked mean, loss.py:40-44)   or   dS * g / T   (plain mean, model.py:628)
__global__ void pair_weights_kernel(const float* __restrict__ dS, const float* __restrict__ cnt,
                                    const float* __restrict__ gscale, int mean_mode, float inv_T, size_t n,
                                    float* __restrict__ w) {
  const float g = gscale ? *gscale : 1.f;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float d = dS[i] * g;
    w[i] = (mean_mode == CXL_MEAN_MASKED) ? d / (cnt[i] + 1e-8f) : d * inv_T;
  }
}

int pair_weights(const float* dS, const float* cnt, const float* gscale, int mean_mode, int T, long long n, float* w,
                 cudaStream_t stream) {
  if (n <= 0) return CXL_OK;
  if (mean_mode == CXL_MEAN_MASKED && cnt == nullptr) return CXL_ERR_INVALID_ARG;
  size_t blocks = (static_cast<size_t>(n) + 255) / 256;
  const size_t cap = static_cast<size_t>(num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  pair_weights_kernel<<<static_cast<unsigned>(blocks), 256, 0, CXL_COUNTED(stream)>>>(dS, cnt, gscale, mean_mode,
                                                                        1.0f / static_cast<float>(T),
                                                                        static_cast<size_t>(n), w);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

// rank of the target column in a stable descending ordering of each row (train.py:467-469):
// #entries strictly greater + #equal entries at a lower index.
__global__ void __launch_bounds__(256) rank_of_target_kernel(const float* __restrict__ scores, long long row_stride,
                                                             long long col_stride, int cols,
                                                             const int* __restrict__ target, int* __restrict__ rank_out) {
  const int r = blockIdx.x;
  const float* row = scores + static_cast<long long>(r) * row_stride;
  const int t = target[r];
  if (t < 0 || t >= cols) {                                // stale / sentinel index: no out-of-bounds read, rank -1
    if (threadIdx.x == 0) rank_out[r] = -1;
    return;
  }
  const float tv = row[static_cast<long long>(t) * col_stride];
  int cnt = 0;
  if (col_stride == 1 && (cols & 3) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    for (int i = threadIdx.x; i < (cols >> 2); i += blockDim.x) {          // contiguous rows: 16-byte loads
      const float4 v = __ldg(reinterpret_cast<const float4*>(row) + i);
      const int c = 4 * i;
      cnt += (v.x > tv) || (v.x == tv && c < t);
      cnt += (v.y > tv) || (v.y == tv && c + 1 < t);
      cnt += (v.z > tv) || (v.z == tv && c + 2 < t);
      cnt += (v.w > tv) || (v.w == tv && c + 3 < t);
    }
  } else {
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      const float v = row[static_cast<long long>(c) * col_stride];
      cnt += (v > tv) || (v == tv && c < t);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  __shared__ int part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
    for (int w = 0; w < 8; ++w) s += part[w];
    rank_out[r] = s;
  }
}

// Same ranking for a TRANSPOSED view (row_stride == 1: consecutive rows are adjacent in memory, e.g. the t2i
// direction scores.t() of compute_retrieval): lane <-> row so that every warp load is 32 consecutive floats;
// the 8 warps of a CTA split the column range of the CTA's 32 rows.
__global__ void __launch_bounds__(256) rank_of_target_tr_kernel(const float* __restrict__ scores, long long col_stride,
                                                                int rows, int cols, const int* __restrict__ target,
                                                                int* __restrict__ rank_out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int r = blockIdx.x * 32 + lane;
  const bool ok = r < rows;
  const int t_raw = ok ? target[r] : 0;
  const bool t_ok = t_raw >= 0 && t_raw < cols;            // stale / sentinel index: no out-of-bounds read, rank -1
  const int t = t_ok ? t_raw : 0;
  const float* base = scores + (ok ? r : 0);
  const float tv = base[static_cast<long long>(t) * col_stride];
  int cnt = 0;
  int c = warp;
  for (; c + 24 < cols; c += 32) {                      // four independent loads in flight per lane
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = __ldg(base + static_cast<long long>(c + 8 * j) * col_stride);
#pragma unroll
    for (int j = 0; j < 4; ++j) cnt += (v[j] > tv) || (v[j] == tv && c + 8 * j < t);
  }
  for (; c < cols; c += 8) {
    const float v = base[static_cast<long long>(c) * col_stride];
    cnt += (v > tv) || (v == tv && c < t);
  }
  __shared__ int part[8][32];
  part[warp][lane] = cnt;
  __syncthreads();
  if (warp == 0 && ok) {
    int s = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += part[w][lane];
    rank_out[r] = t_ok ? s : -1;
  }
}

int rank_of_target(const float* scores, long long row_stride, long long col_stride, int rows, int cols,
                   const int* target, int* rank_out, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return CXL_ERR_INVALID_ARG;
  if (row_stride == 1 && col_stride != 1) {
    rank_of_target_tr_kernel<<<(rows + 31) / 32, 256, 0, CXL_COUNTED(stream)>>>(scores, col_stride, rows, cols, target,
                                                                               rank_out);
    CXL_CUDA_CHECK(cudaGetLastError());
    return CXL_OK;
  }
  rank_of_target_kernel<<<rows, 256, 0, CXL_COUNTED(stream)>>>(scores, row_stride, col_stride, cols, target, rank_out);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

// X [R,C] f32 -> Y [R,C] and Yt [C,R] in bf16 / fp16 (either output may be null): the 16-bit copies of dL/dlogits the two
// gradient GEMMs of the global-embedding loss consume (the reference's autocast backward multiplies a 16-bit dlogits,
// loss.py:252-253).  32 x 32 tiles through shared memory: both the row-major and the transposed store are coalesced.
template <typename O>
__global__ void __launch_bounds__(256) cast_transpose_kernel(const float* __restrict__ X, int R, int C,
                                                             O* __restrict__ Y, O* __restrict__ Yt,
                                                             const float* __restrict__ scale) {
  __shared__ float tile[32][33];
  const float sc = scale != nullptr ? __ldg(scale) : 1.f;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8 threads
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = r0 + ty + 8 * i, c = c0 + tx;
    float v = 0.f;
    if (r < R && c < C) v = X[static_cast<size_t>(r) * C + c] * sc;
    tile[ty + 8 * i][tx] = v;
    if (Y != nullptr && r < R && c < C) Y[static_cast<size_t>(r) * C + c] = static_cast<O>(v);
  }
  __syncthreads();
  if (Yt != nullptr) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = c0 + ty + 8 * i, r = r0 + tx;
      if (r < R && c < C) Yt[static_cast<size_t>(c) * R + r] = static_cast<O>(tile[tx][ty + 8 * i]);
    }
  }
}

int cast_transpose(const float* X, int R, int C, void* Y, void* Yt, int dtype, const float* scale, cudaStream_t stream) {
  if (R <= 0 || C <= 0) return CXL_ERR_INVALID_ARG;
  dim3 grid((C + 31) / 32, (R + 31) / 32);
  if (grid.y > 65535) return CXL_ERR_UNSUPPORTED;
  if (dtype == CXL_DTYPE_BF16)
    cast_transpose_kernel<__nv_bfloat16><<<grid, 256, 0, CXL_COUNTED(stream)>>>(X, R, C, static_cast<__nv_bfloat16*>(Y),
                                                                                static_cast<__nv_bfloat16*>(Yt), scale);
  else if (dtype == CXL_DTYPE_F16)
    cast_transpose_kernel<__half><<<grid, 256, 0, CXL_COUNTED(stream)>>>(X, R, C, static_cast<__half*>(Y),
                                                                         static_cast<__half*>(Yt), scale);
  else
    return CXL_ERR_INVALID_ARG;
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

// NaN propagation of torch.max (loss.py:35): the K1 epilogues keep a running maximum with `x > best`, which drops NaN
// similarities instead of propagating them.  A NaN / Inf similarity can only come from a NaN / Inf operand element, so
// the operands are scanned once (HBM speed) and every score of a text / an image that holds a non-finite element is
// set to NaN: the loss and the gradients of that step become non-finite, which is what AMP's GradScaler looks for.
template <typename T> __device__ __forceinline__ bool nonfinite16(uint4 v);
template <> __device__ __forceinline__ bool nonfinite16<__nv_bfloat16>(uint4 v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  bool bad = false;
#pragma unroll
  for (int i = 0; i < 4; ++i) bad |= ((w[i] & 0x7F80u) == 0x7F80u) || ((w[i] & 0x7F800000u) == 0x7F800000u);
  return bad;
}
template <> __device__ __forceinline__ bool nonfinite16<__half>(uint4 v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  bool bad = false;
#pragma unroll
  for (int i = 0; i < 4; ++i) bad |= ((w[i] & 0x7C00u) == 0x7C00u) || ((w[i] & 0x7C000000u) == 0x7C000000u);
  return bad;
}
template <> __device__ __forceinline__ bool nonfinite16<float>(uint4 v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  bool bad = false;
#pragma unroll
  for (int i = 0; i < 4; ++i) bad |= (w[i] & 0x7F800000u) == 0x7F800000u;
  return bad;
}

// one CTA per group of `vecs` 16-byte vectors (a text's tokens / an image's patch rows); a flagged group writes NaN to
// S[group * s_group + e * s_elem], e < n
template <typename T>
__global__ void __launch_bounds__(256) poison_nonfinite_kernel(const uint4* __restrict__ op, long long vecs, float* S,
                                                               long long s_group, long long s_elem, int n) {
  const uint4* g = op + static_cast<long long>(blockIdx.x) * vecs;
  bool bad = false;
  for (long long i = threadIdx.x; i < vecs; i += blockDim.x) bad |= nonfinite16<T>(__ldg(g + i));
  if (__syncthreads_or(bad ? 1 : 0)) {
    float* row = S + static_cast<long long>(blockIdx.x) * s_group;
    for (int e = threadIdx.x; e < n; e += blockDim.x) row[static_cast<long long>(e) * s_elem] = __int_as_float(0x7FC00000);
  }
}

// dtype: CXL_DTYPE_* of the operands; txt [Bt, T*D], img [Bi, p_stride*D] contiguous, 16-byte aligned rows
int poison_nonfinite(const void* txt, const void* img, int Bt, int Bi, int T, int p_stride, int D, int dtype, float* S,
                     cudaStream_t stream) {
  const int es = dtype == CXL_DTYPE_F32 ? 4 : 2;
  const long long tv = static_cast<long long>(T) * D * es, iv = static_cast<long long>(p_stride) * D * es;
  if (tv % 16 != 0 || iv % 16 != 0 || ((reinterpret_cast<uintptr_t>(txt) | reinterpret_cast<uintptr_t>(img)) & 15))
    return CXL_OK;                                          // unaligned operands (odd T*D): no vector scan, no propagation
#define CXL_POISON(TY)                                                                                                  \
  poison_nonfinite_kernel<TY><<<Bt, 256, 0, CXL_COUNTED(stream)>>>(static_cast<const uint4*>(txt), tv / 16, S, Bi, 1, Bi); \
  poison_nonfinite_kernel<TY><<<Bi, 256, 0, CXL_COUNTED(stream)>>>(static_cast<const uint4*>(img), iv / 16, S, 1, Bi, Bt)
  if (dtype == CXL_DTYPE_BF16) { CXL_POISON(__nv_bfloat16); }
  else if (dtype == CXL_DTYPE_F16) { CXL_POISON(__half); }
  else { CXL_POISON(float); }
#undef CXL_POISON
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

}  // namespace cxl
