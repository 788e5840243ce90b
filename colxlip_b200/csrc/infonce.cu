// K3: logit scale + symmetric InfoNCE (row- and column-softmax cross-entropy against the
// diagonal) over a [B_loc, B] block of unscaled logits, plus its gradient.  HBM-bound.
//
// Replaces `logit_scale * S`, `.T` and the four F.cross_entropy calls of the reference
// (loss.py:252-263, 285-293) and their autograd (log_softmax / nll_loss backward).
// Rows are this rank's texts (global row index row0 + m), columns are all B images.
#include <mutex>

#include "common.cuh"

namespace cxl {

// online (max, sum exp) pair
struct MS {
  float m, s;
};
__device__ __forceinline__ MS ms_combine(MS a, MS b) {
  const float mx = fmaxf(a.m, b.m);
  if (mx == -INFINITY) return MS{-INFINITY, 0.f};
  return MS{mx, a.s * expf(a.m - mx) + b.s * expf(b.m - mx)};
}
__device__ __forceinline__ void ms_push(MS& a, float v) {
  if (v > a.m) {
    a.s = a.s * expf(a.m - v) + 1.f;
    a.m = v;
  } else {
    a.s += expf(v - a.m);
  }
}

// four values at once, branch-free: one rescale of the running sum + four exponentials (values are finite)
__device__ __forceinline__ void ms_push4(MS& a, float v0, float v1, float v2, float v3) {
  const float mn = fmaxf(fmaxf(a.m, fmaxf(v0, v1)), fmaxf(v2, v3));
  a.s = a.s * expf(a.m - mn) + ((expf(v0 - mn) + expf(v1 - mn)) + (expf(v2 - mn) + expf(v3 - mn)));
  a.m = mn;
}

// ---- row log-sum-exp: one CTA per row ----
__global__ void __launch_bounds__(256) infonce_row_kernel(const float* __restrict__ X, const float* __restrict__ scale_p,
                                                          int B, int row0, float* __restrict__ row_lse,
                                                          float* __restrict__ diag) {
  const int m = blockIdx.x;
  const float scale = *scale_p;
  const float* x = X + static_cast<size_t>(m) * B;
  MS acc{-INFINITY, 0.f}, acc2{-INFINITY, 0.f};
  const int nvec = ((reinterpret_cast<uintptr_t>(x) & 15) == 0) ? (B >> 2) : 0;
  int i = threadIdx.x;
  for (; i + static_cast<int>(blockDim.x) < nvec; i += 2 * blockDim.x) {     // two independent 16-byte loads in flight
    const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    const float4 u = __ldg(reinterpret_cast<const float4*>(x) + i + blockDim.x);
    ms_push4(acc, scale * v.x, scale * v.y, scale * v.z, scale * v.w);
    ms_push4(acc2, scale * u.x, scale * u.y, scale * u.z, scale * u.w);
  }
  for (; i < nvec; i += blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    ms_push4(acc, scale * v.x, scale * v.y, scale * v.z, scale * v.w);
  }
  for (int j = nvec * 4 + threadIdx.x; j < B; j += blockDim.x) ms_push(acc, scale * x[j]);
  acc = ms_combine(acc, acc2);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MS other{__shfl_xor_sync(0xffffffffu, acc.m, o), __shfl_xor_sync(0xffffffffu, acc.s, o)};
    acc = ms_combine(acc, other);
  }
  __shared__ MS part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    MS t = part[0];
    for (int w = 1; w < (blockDim.x >> 5); ++w) t = ms_combine(t, part[w]);
    row_lse[m] = t.m + logf(t.s);
    diag[m] = scale * x[row0 + m];
  }
}

// ---- column (max, sumexp) partials over a block of rows: thread owns 4 adjacent columns ----
static constexpr int kColRowsPerBlock = 128;
__global__ void __launch_bounds__(128) infonce_col_partial_kernel(const float* __restrict__ X,
                                                                  const float* __restrict__ scale_p, int B_loc, int B,
                                                                  float* __restrict__ pmax, float* __restrict__ psum) {
  const int c0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (c0 >= B) return;
  const float scale = *scale_p;
  const int r0 = blockIdx.y * kColRowsPerBlock;
  const int r1 = min(r0 + kColRowsPerBlock, B_loc);
  MS a[4] = {{-INFINITY, 0.f}, {-INFINITY, 0.f}, {-INFINITY, 0.f}, {-INFINITY, 0.f}};
  const bool vec = (c0 + 3 < B) && ((B & 3) == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
  int r = r0;
  if (vec) {
    for (; r + 4 <= r1; r += 4) {                       // four independent 16-byte loads in flight
      float4 t[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) t[j] = __ldg(reinterpret_cast<const float4*>(X + static_cast<size_t>(r + j) * B + c0));
      ms_push4(a[0], scale * t[0].x, scale * t[1].x, scale * t[2].x, scale * t[3].x);
      ms_push4(a[1], scale * t[0].y, scale * t[1].y, scale * t[2].y, scale * t[3].y);
      ms_push4(a[2], scale * t[0].z, scale * t[1].z, scale * t[2].z, scale * t[3].z);
      ms_push4(a[3], scale * t[0].w, scale * t[1].w, scale * t[2].w, scale * t[3].w);
    }
  }
  for (; r < r1; ++r) {
    const float* x = X + static_cast<size_t>(r) * B + c0;
    float v[4];
    if (vec) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(x));
      v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) v[i] = (c0 + i < B) ? x[i] : -INFINITY;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (c0 + i < B) ms_push(a[i], scale * v[i]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (c0 + i < B) {
      pmax[static_cast<size_t>(blockIdx.y) * B + c0 + i] = a[i].m;
      psum[static_cast<size_t>(blockIdx.y) * B + c0 + i] = a[i].s;
    }
}
__global__ void infonce_col_combine_kernel(const float* __restrict__ pmax, const float* __restrict__ psum, int nblk,
                                           int B, float* __restrict__ col_max, float* __restrict__ col_sum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= B) return;
  MS t{-INFINITY, 0.f};
  for (int b = 0; b < nblk; ++b) t = ms_combine(t, MS{pmax[static_cast<size_t>(b) * B + c], psum[static_cast<size_t>(b) * B + c]});
  col_max[c] = t.m;
  col_sum[c] = t.s;
}

// ---- gradient: dX = coef*scale*(softmax_row + softmax_col - 2*delta); dscale += sum(coef*(...)*X) ----
// (rows-only when col_lse is null: dX = coef*scale*(softmax_row - delta))
__global__ void __launch_bounds__(256) infonce_grad_kernel(const float* X, const float* __restrict__ scale_p,
                                                           const float* __restrict__ coef_p, int B_loc, int B, int row0,
                                                           const float* __restrict__ row_lse,
                                                           const float* __restrict__ col_lse,
                                                           const float* __restrict__ cnt, float* dX,
                                                           float* __restrict__ dscale_accum,
                                                           float* __restrict__ partials) {
  const int m = blockIdx.y;
  const float scale = *scale_p;
  const float coef = *coef_p / (2.f * static_cast<float>(B));
  const float rl = row_lse[m];
  const float* x = X + static_cast<size_t>(m) * B;
  float* dx = dX + static_cast<size_t>(m) * B;
  const float* cn = cnt ? cnt + static_cast<size_t>(m) * B : nullptr;
  const int diag_c = row0 + m;
  float ds = 0.f;
  auto one = [&](float xv, int c, float cl, float cnv) -> float {
    const float sx = scale * xv;
    // col_lse == nullptr: rows-only cross-entropy (ClipLoss local_loss, loss.py:144-146): softmax_row - delta
    float t = expf(sx - rl);
    if (col_lse) t += expf(sx - cl);
    if (c == diag_c) t -= col_lse ? 2.f : 1.f;
    t *= coef;
    ds = fmaf(t, xv, ds);
    const float dxv = t * scale;
    return cn ? dxv / (cnv + 1e-8f) : dxv;            // with cnt: the MaxSim pair weight  dL/dS / (cnt + 1e-8)
  };
  // rows of X / dX / cnt are 16-byte aligned when the bases are and B % 4 == 0: four columns per thread and access
  const bool vec = ((B & 3) == 0) && (((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(dX) |
                                        reinterpret_cast<uintptr_t>(cnt) | reinterpret_cast<uintptr_t>(col_lse)) & 15) == 0);
  if (vec) {
    const int nv = B >> 2;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += gridDim.x * blockDim.x) {
      const float4 xv = *(reinterpret_cast<const float4*>(x) + i);
      const float4 cl = col_lse ? __ldg(reinterpret_cast<const float4*>(col_lse) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 cv = cn ? __ldg(reinterpret_cast<const float4*>(cn) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 o;
      o.x = one(xv.x, 4 * i, cl.x, cv.x);
      o.y = one(xv.y, 4 * i + 1, cl.y, cv.y);
      o.z = one(xv.z, 4 * i + 2, cl.z, cv.z);
      o.w = one(xv.w, 4 * i + 3, cl.w, cv.w);
      *(reinterpret_cast<float4*>(dx) + i) = o;
    }
  } else {
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < B; c += gridDim.x * blockDim.x)
      dx[c] = one(x[c], c, col_lse ? col_lse[c] : 0.f, cn ? cn[c] : 0.f);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ds += __shfl_xor_sync(0xffffffffu, ds, o);
  __shared__ float part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = ds;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < (blockDim.x >> 5); ++w) t += part[w];
    // partials != null: this block's own cell, summed afterwards in a fixed order (reproducible d logit_scale)
    if (partials != nullptr) partials[static_cast<size_t>(blockIdx.y) * gridDim.x + blockIdx.x] = t;
    else atomicAdd(dscale_accum, t);
  }
}

// *accum += sum(partials[0..n)) in an order that depends only on n: thread t adds elements t, t + 1024, ... in
// ascending order, then a fixed shared-memory tree
__global__ void __launch_bounds__(1024) sum_fixed_order_kernel(const float* __restrict__ partials, size_t n,
                                                               float* __restrict__ accum) {
  __shared__ float sh[1024];
  float t = 0.f;
  for (size_t i = threadIdx.x; i < n; i += 1024) t += partials[i];
  sh[threadIdx.x] = t;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (static_cast<int>(threadIdx.x) < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *accum += sh[0];
}

size_t infonce_workspace_bytes(int B_loc, int B) {
  const size_t nblk = (static_cast<size_t>(B_loc) + kColRowsPerBlock - 1) / kColRowsPerBlock;
  return 2 * nblk * static_cast<size_t>(B) * sizeof(float);
}

int infonce_stats(const float* X, const float* scale_dev, int B_loc, int B, int row0, float* row_lse, float* col_max,
                  float* col_sum, float* diag, void* ws, size_t ws_bytes, cudaStream_t stream) {
  if (B_loc <= 0 || B <= 0 || row0 < 0 || row0 + B_loc > B) return CXL_ERR_INVALID_ARG;
  if ((col_max == nullptr) != (col_sum == nullptr)) return CXL_ERR_INVALID_ARG;
  if (col_max != nullptr && (ws == nullptr || ws_bytes < infonce_workspace_bytes(B_loc, B))) return CXL_ERR_WORKSPACE;
  const int nblk = (B_loc + kColRowsPerBlock - 1) / kColRowsPerBlock;
  float* pmax = reinterpret_cast<float*>(ws);
  float* psum = pmax + static_cast<size_t>(nblk) * B;
  infonce_row_kernel<<<B_loc, 256, 0, CXL_COUNTED(stream)>>>(X, scale_dev, B, row0, row_lse, diag);
  CXL_CUDA_CHECK(cudaGetLastError());
  if (col_max == nullptr && col_sum == nullptr) return CXL_OK;   // rows-only statistics (ClipLoss local_loss)
  dim3 g((B + 511) / 512, nblk);
  infonce_col_partial_kernel<<<g, 128, 0, CXL_COUNTED(stream)>>>(X, scale_dev, B_loc, B, pmax, psum);
  CXL_CUDA_CHECK(cudaGetLastError());
  infonce_col_combine_kernel<<<(B + 255) / 256, 256, 0, CXL_COUNTED(stream)>>>(pmax, psum, nblk, B, col_max, col_sum);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

static int infonce_grad_grid_x(int B) {
  int gx = (B + 256 * 8 - 1) / (256 * 8);
  return gx < 1 ? 1 : gx;
}

size_t infonce_grad_workspace_bytes(int B_loc, int B) {
  if (B_loc <= 0 || B <= 0) return 0;
  return static_cast<size_t>(infonce_grad_grid_x(B)) * static_cast<size_t>(B_loc) * sizeof(float);
}

// partials: null = d logit_scale accumulated with atomicAdd; else one float per block (infonce_grad_workspace_bytes),
// summed in a fixed order after the last slab
static int infonce_grad_slab(const float* X, const float* scale_dev, const float* coef_dev, int B_loc, int B, int row0,
                             const float* row_lse, const float* col_lse, const float* cnt, float* dX, float* dscale_accum,
                             float* partials, cudaStream_t stream) {
  dim3 g(infonce_grad_grid_x(B), B_loc);
  infonce_grad_kernel<<<g, 256, 0, CXL_COUNTED(stream)>>>(X, scale_dev, coef_dev, B_loc, B, row0, row_lse, col_lse, cnt, dX,
                                             dscale_accum, partials);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}

int infonce_grad(const float* X, const float* scale_dev, const float* coef_dev, int B_loc, int B, int row0,
                 const float* row_lse, const float* col_lse, const float* cnt, float* dX, float* dscale_accum,
                 cudaStream_t stream, float* partials) {
  if (B_loc <= 0 || B <= 0 || row0 < 0 || row0 + B_loc > B) return CXL_ERR_INVALID_ARG;
  const size_t gx = static_cast<size_t>(infonce_grad_grid_x(B));
  for (int r = 0; r < B_loc; r += 65535) {               // grid.y limit: slabs of rows
    const int nr = min(65535, B_loc - r);
    int rc = infonce_grad_slab(X + static_cast<size_t>(r) * B, scale_dev, coef_dev, nr, B, row0 + r, row_lse + r, col_lse,
                               cnt ? cnt + static_cast<size_t>(r) * B : nullptr, dX + static_cast<size_t>(r) * B,
                               dscale_accum, partials ? partials + static_cast<size_t>(r) * gx : nullptr, stream);
    if (rc != CXL_OK) return rc;
  }
  if (partials != nullptr) {
    sum_fixed_order_kernel<<<1, 1024, 0, CXL_COUNTED(stream)>>>(partials, gx * static_cast<size_t>(B_loc), dscale_accum);
    CXL_CUDA_CHECK(cudaGetLastError());
  }
  return CXL_OK;
}

// ---- allocation-free statistics pass: column partials in a file-scope device array ----------------------------
// (no counterpart in the reference; added for callers that may not call
// cudaMalloc: CUDA-graph capture, tight serving loops.)  Every device that loads the library holds
// kInfonceStaticScratchBytes until exit.  All calls on one device share the array: calls that overlap on the GPU
// (different streams) race.  Nothing in it has to start at zero: the partial kernel writes every (row block, column)
// cell the combine kernel reads, so stale contents of an earlier call are never consumed.
static constexpr size_t kInfonceStaticScratchBytes = size_t(CXL_INFONCE_STATIC_SCRATCH_BYTES);
static __device__ __align__(256) unsigned char g_infonce_static_scratch[kInfonceStaticScratchBytes];

static constexpr int kMaxStaticDevices = 64;
static std::mutex g_static_mu;
static void* g_static_ptr[kMaxStaticDevices];        // device address of the array, per device ordinal (0 = not yet asked)

// The array's name is a host-side stand-in here, not a device address: ask the runtime once per device, with that
// device current (the caller's current device is where its stream and buffers live).
static int infonce_static_scratch(void** out) {
  int dev = -1;
  CXL_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxStaticDevices) return CXL_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    CXL_CUDA_CHECK(cudaGetSymbolAddress(&p, g_infonce_static_scratch));
    if (p == nullptr) return CXL_ERR_CUDA;
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return CXL_OK;
}

int infonce_stats_static(const float* X, const float* scale_dev, int B_loc, int B, int row0, float* row_lse,
                         float* col_max, float* col_sum, float* diag, cudaStream_t stream) {
  if (B_loc <= 0 || B <= 0 || row0 < 0 || row0 + B_loc > B) return CXL_ERR_INVALID_ARG;
  if (col_max == nullptr || col_sum == nullptr) return CXL_ERR_INVALID_ARG;   // rows-only needs no scratch: cxl_infonce_stats
  if (infonce_workspace_bytes(B_loc, B) > kInfonceStaticScratchBytes) return CXL_ERR_STATIC_SCRATCH;
  void* scratch = nullptr;
  const int rc = infonce_static_scratch(&scratch);
  if (rc != CXL_OK) return rc;
  const int nblk = (B_loc + kColRowsPerBlock - 1) / kColRowsPerBlock;
  float* pmax = reinterpret_cast<float*>(scratch);
  float* psum = pmax + static_cast<size_t>(nblk) * B;
  infonce_row_kernel<<<B_loc, 256, 0, CXL_COUNTED(stream)>>>(X, scale_dev, B, row0, row_lse, diag);
  CXL_CUDA_CHECK(cudaGetLastError());
  dim3 g((B + 511) / 512, nblk);
  infonce_col_partial_kernel<<<g, 128, 0, CXL_COUNTED(stream)>>>(X, scale_dev, B_loc, B, pmax, psum);
  CXL_CUDA_CHECK(cudaGetLastError());
  infonce_col_combine_kernel<<<(B + 255) / 256, 256, 0, CXL_COUNTED(stream)>>>(pmax, psum, nblk, B, col_max, col_sum);
  CXL_CUDA_CHECK(cudaGetLastError());
  return CXL_OK;
}


}  // namespace cxl
