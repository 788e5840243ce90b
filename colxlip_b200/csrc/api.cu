// extern "C" boundary of libcolxlip_b200.so (see include/colxlip_b200.h).
#include <stdio.h>
#include <string.h>

#include "common.cuh"

namespace cxl {

static thread_local char g_cuda_err[512] = "";

void set_cuda_error(cudaError_t e, const char* where) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
}

static unsigned long long g_launches = 0;
void note_launch() { __atomic_fetch_add(&g_launches, 1ull, __ATOMIC_RELAXED); }
unsigned long long launches_so_far() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int img_row_stride(int P);
int maxsim_fwd_tc(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int mean_mode, float* S,
                  float* cnt, void* argmax, long long ld_arg, float* dbg, int force_stream_a, int force_grid, int f16,
                  cudaStream_t stream, float* S_extra = nullptr);
int maxsim_fwd_f32(const float* txt, const float* img, int Bt, int Bi, int T, int P, int D, int mean_mode, float* S,
                   float* cnt, void* argmax, long long ld_arg, cudaStream_t stream, float* S_extra = nullptr);
int maxsim_fwd_slots(int T, int slot_rows);
int maxsim_fwd_f32_slot_rows();
int maxsim_bwd(const void* txt, const void* img, const void* argmax, long long ld_arg, long long arg_row_off,
               const float* w, int Bt, int Bi, int T, int P, int D, int dtype, int p_stride, int ld_pairs, int k_off,
               float* dtxt, int accumulate_txt, float* dimg, int accumulate_img, int force_generic, void* ws,
               size_t ws_bytes, const float* w_scale, cudaStream_t stream);
size_t maxsim_bwd_workspace_bytes(int Bt, int Bi, int T, int P, int D);
int pair_weights(const float* dS, const float* cnt, const float* gscale, int mean_mode, int T, long long n, float* w,
                 cudaStream_t stream);
size_t infonce_workspace_bytes(int B_loc, int B);
int infonce_stats(const float* X, const float* scale_dev, int B_loc, int B, int row0, float* row_lse, float* col_max,
                  float* col_sum, float* diag, void* ws, size_t ws_bytes, cudaStream_t stream);
int infonce_stats_static(const float* X, const float* scale_dev, int B_loc, int B, int row0, float* row_lse,
                         float* col_max, float* col_sum, float* diag, cudaStream_t stream);
int infonce_grad(const float* X, const float* scale_dev, const float* coef_dev, int B_loc, int B, int row0,
                 const float* row_lse, const float* col_lse, const float* cnt, float* dX, float* dscale_accum,
                 cudaStream_t stream, float* partials = nullptr);
size_t infonce_grad_workspace_bytes(int B_loc, int B);
int pack_bf16(const void* src, int src_dtype, long long item_stride, long long row_stride, int items, int n, int n_pad,
              int D, void* dst, cudaStream_t stream);
int pack_operand(const void* src, int src_dtype, long long item_stride, long long row_stride, int items, int n,
                 int n_pad, int D, void* dst, int dst_dtype, cudaStream_t stream);
int gemm_nt_tc(const void* A, const void* B, int M, int N, int K, float* C, long long ldc, int f16, cudaStream_t stream);
int cast_transpose(const float* X, int R, int C, void* Y, void* Yt, int dtype, const float* scale, cudaStream_t stream);
int head_ln_in(const void* x, int x_dtype, long long rows, int width, const float* gamma, const float* beta, float eps,
               const int* valid_len, int T, void* y, int y_dtype, cudaStream_t stream);
int head_out(const float* h, long long rows, int D, const float* bias, const float* gamma, const float* beta, float eps,
             int l2_normalize, void* y, int y_dtype, cudaStream_t stream);
int poison_nonfinite(const void* txt, const void* img, int Bt, int Bi, int T, int p_stride, int D, int dtype, float* S,
                     cudaStream_t stream);
int rank_of_target(const float* scores, long long row_stride, long long col_stride, int rows, int cols,
                   const int* target, int* rank_out, cudaStream_t stream);

int l2norm_fwd(const void* x, int x_dtype, long long rows, int D, void* y, int y_dtype, float* inv_norm,
               cudaStream_t stream);
int l2norm_bwd(const void* y, int y_dtype, const void* g, int g_dtype, const float* inv_norm, long long rows, int D,
               void* dx, int dx_dtype, cudaStream_t stream);

static int device_ok() {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return CXL_ERR_NO_DEVICE;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return CXL_ERR_NO_DEVICE;
  return major == 10 ? CXL_OK : CXL_ERR_NO_DEVICE;
}

}  // namespace cxl

using namespace cxl;

extern "C" {

unsigned long long cxl_kernel_launches(void) { return cxl::launches_so_far(); }

const char* cxl_strerror(int code) {
  switch (code) {
    case CXL_OK: return "ok";
    case CXL_ERR_INVALID_ARG: return "invalid argument";
    case CXL_ERR_UNSUPPORTED: return "unsupported shape or dtype for this build";
    case CXL_ERR_CUDA: return "CUDA call failed (see cxl_last_cuda_error)";
    case CXL_ERR_NO_DEVICE: return "no sm_100 (Blackwell B200) device; there is no CPU fallback";
    case CXL_ERR_WORKSPACE: return "workspace too small";
    case CXL_ERR_STATIC_SCRATCH:
      return "problem exceeds the built-in static scratch of this *_static entry point (see its size limit in "
             "colxlip_b200.h); use the workspace-taking entry point";
    default: return "unknown error";
  }
}

const char* cxl_last_cuda_error(void) { return g_cuda_err; }

int cxl_version(void) { return 100; }

int cxl_device_check(void) { return device_ok(); }

int cxl_argmax_elem_bytes(int P) { return P > 256 ? 2 : 1; }

long long cxl_argmax_ld(int Bt, int T) { return (static_cast<long long>(Bt) * T + 15) / 16 * 16; }

int cxl_img_row_stride(int P) { return img_row_stride(P); }

int cxl_pack_bf16(const void* src, int src_dtype, long long src_item_stride, long long src_row_stride, int items,
                  int n, int n_pad, int D, void* dst, void* stream) {
  if (src == nullptr || dst == nullptr) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return pack_bf16(src, src_dtype, src_item_stride, src_row_stride, items, n, n_pad, D, dst,
                   static_cast<cudaStream_t>(stream));
}

int cxl_pack_operand(const void* src, int src_dtype, long long src_item_stride, long long src_row_stride, int items,
                     int n, int n_pad, int D, void* dst, int dst_dtype, void* stream) {
  if (src == nullptr || dst == nullptr) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return pack_operand(src, src_dtype, src_item_stride, src_row_stride, items, n, n_pad, D, dst, dst_dtype,
                      static_cast<cudaStream_t>(stream));
}

static int maxsim_fwd_entry(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int dtype, int mean_mode,
                            float* S, float* cnt, void* argmax, void* ws, size_t ws_bytes, void* stream) {
  if (txt == nullptr || img == nullptr || S == nullptr) return CXL_ERR_INVALID_ARG;
  if (Bt <= 0 || Bi <= 0 || T <= 0 || P <= 0 || D <= 0) return CXL_ERR_INVALID_ARG;
  if (static_cast<long long>(Bt) * T >= (1LL << 31) || static_cast<long long>(Bi) * P >= (1LL << 31))
    return CXL_ERR_UNSUPPORTED;
  const int base_mode = mean_mode & ~CXL_MEAN_ROUND16;
  if (base_mode != CXL_MEAN_MASKED && base_mode != CXL_MEAN_PLAIN) return CXL_ERR_INVALID_ARG;
  if ((mean_mode & CXL_MEAN_ROUND16) && dtype == CXL_DTYPE_F32) return CXL_ERR_INVALID_ARG;   // fp32 einsum rounds nothing
  if (dtype != CXL_DTYPE_BF16 && dtype != CXL_DTYPE_F16 && dtype != CXL_DTYPE_F32) return CXL_ERR_INVALID_ARG;
  float* S_extra = nullptr;                              // ws given: fixed-order sums, S is slot 0 and ws holds the others
  if (ws != nullptr) {
    if (ws_bytes < cxl_maxsim_fwd_workspace_bytes(Bt, Bi, T, dtype)) return CXL_ERR_WORKSPACE;
    if ((reinterpret_cast<uintptr_t>(ws) & 3) != 0) return CXL_ERR_INVALID_ARG;
    S_extra = static_cast<float*>(ws);
  }
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long ld_arg = cxl_argmax_ld(Bt, T);
  if (dtype == CXL_DTYPE_BF16 || dtype == CXL_DTYPE_F16)
    rc = maxsim_fwd_tc(txt, img, Bt, Bi, T, P, D, mean_mode, S, cnt, argmax, ld_arg, nullptr, 0, 0,
                       dtype == CXL_DTYPE_F16, st, S_extra);
  else
    rc = maxsim_fwd_f32(static_cast<const float*>(txt), static_cast<const float*>(img), Bt, Bi, T, P, D, mean_mode, S,
                        cnt, argmax, ld_arg, st, S_extra);
  if (rc != CXL_OK) return rc;
  // torch.max propagates NaN, the epilogue's running max does not: scores of texts / images with a non-finite element -> NaN
  return poison_nonfinite(txt, img, Bt, Bi, T, dtype == CXL_DTYPE_F32 ? P : img_row_stride(P), D, dtype, S, st);
}

int cxl_maxsim_fwd(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int dtype, int mean_mode,
                   float* S, float* cnt, void* argmax, void* stream) {
  return maxsim_fwd_entry(txt, img, Bt, Bi, T, P, D, dtype, mean_mode, S, cnt, argmax, nullptr, 0, stream);
}

size_t cxl_maxsim_fwd_workspace_bytes(int Bt, int Bi, int T, int dtype) {
  if (Bt <= 0 || Bi <= 0 || T <= 0) return 0;
  const int slot_rows = dtype == CXL_DTYPE_F32 ? maxsim_fwd_f32_slot_rows() : 32;
  const size_t extra = static_cast<size_t>(maxsim_fwd_slots(T, slot_rows) - 1);
  const size_t bytes = extra * static_cast<size_t>(Bt) * Bi * sizeof(float);
  return bytes > 0 ? bytes : 4;                          // T = 1: a text never spans two blocks; keep ws != NULL meaningful
}

int cxl_maxsim_fwd_ws(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int dtype, int mean_mode,
                      float* S, float* cnt, void* argmax, void* ws, size_t ws_bytes, void* stream) {
  if (ws == nullptr) return CXL_ERR_WORKSPACE;
  return maxsim_fwd_entry(txt, img, Bt, Bi, T, P, D, dtype, mean_mode, S, cnt, argmax, ws, ws_bytes, stream);
}

int cxl_pair_weights(const float* dS, const float* cnt, const float* gscale_dev, int mean_mode, int T, long long n,
                     float* w, void* stream) {
  if (dS == nullptr || w == nullptr || T <= 0) return CXL_ERR_INVALID_ARG;
  if (mean_mode != CXL_MEAN_MASKED && mean_mode != CXL_MEAN_PLAIN) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return pair_weights(dS, cnt, gscale_dev, mean_mode, T, n, w, static_cast<cudaStream_t>(stream));
}

static int bwd_entry(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T, int P,
                     int D, int dtype, int ld_pairs, int k_off, float* dtxt, float* dimg, int accumulate_img,
                     int force_generic, void* ws, size_t ws_bytes, void* stream) {
  if (txt == nullptr || img == nullptr || argmax == nullptr || w == nullptr) return CXL_ERR_INVALID_ARG;
  if (Bt <= 0 || Bi <= 0 || T <= 0 || P <= 0 || D <= 0) return CXL_ERR_INVALID_ARG;
  if (dtype != CXL_DTYPE_BF16 && dtype != CXL_DTYPE_F32 && dtype != CXL_DTYPE_F16) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  const int p_stride = dtype != CXL_DTYPE_F32 ? img_row_stride(P) : P;
  return maxsim_bwd(txt, img, argmax, cxl_argmax_ld(Bt, T), 0, w, Bt, Bi, T, P, D, dtype, p_stride, ld_pairs, k_off,
                    dtxt, 0, dimg, accumulate_img, force_generic, ws, ws_bytes, nullptr, static_cast<cudaStream_t>(stream));
}

int cxl_maxsim_bwd_dtxt_rows(const void* img, const void* argmax, const float* w, int Bt_total, int m0, int m_cnt,
                             int Bi, int T, int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt,
                             int force_generic, void* stream) {
  if (img == nullptr || argmax == nullptr || w == nullptr || dtxt == nullptr) return CXL_ERR_INVALID_ARG;
  if (Bt_total <= 0 || m0 < 0 || m_cnt <= 0 || m0 + m_cnt > Bt_total) return CXL_ERR_INVALID_ARG;
  if (Bi <= 0 || T <= 0 || P <= 0 || D <= 0) return CXL_ERR_INVALID_ARG;
  if (dtype != CXL_DTYPE_BF16 && dtype != CXL_DTYPE_F32 && dtype != CXL_DTYPE_F16) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  const int p_stride = dtype != CXL_DTYPE_F32 ? img_row_stride(P) : P;
  // dTxt never reads the text operand; `img` stands in for the (unused) txt pointer of the shared dispatcher
  return maxsim_bwd(img, img, argmax, cxl_argmax_ld(Bt_total, T), static_cast<long long>(m0) * T,
                    w + static_cast<size_t>(m0) * ld_pairs, m_cnt, Bi, T, P, D, dtype, p_stride, ld_pairs, k_off, dtxt,
                    0, nullptr, 0, force_generic, nullptr, 0, nullptr, static_cast<cudaStream_t>(stream));
}

// Shard-local operands: `argmax` is the shard's own buffer; the kernels index argmax rows and w columns with the same
// k_off, so the argmax base is moved back by k_off rows (never dereferenced below the shard's first row).
static const void* shard_argmax_base(const void* argmax, long long ld_arg, int P, int k_off) {
  const uintptr_t back = static_cast<uintptr_t>(k_off) * static_cast<uintptr_t>(ld_arg) * (P > 256 ? 2u : 1u);
  return reinterpret_cast<const void*>(reinterpret_cast<uintptr_t>(argmax) - back);
}

int cxl_maxsim_bwd_shard(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T,
                         int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt, int accumulate_txt,
                         float* dimg, int accumulate_img, const float* w_scale_dev, void* ws, size_t ws_bytes,
                         void* stream) {
  if (txt == nullptr || img == nullptr || argmax == nullptr || w == nullptr) return CXL_ERR_INVALID_ARG;
  if (Bt <= 0 || Bi <= 0 || T <= 0 || P <= 0 || D <= 0 || k_off < 0) return CXL_ERR_INVALID_ARG;
  if (dtype != CXL_DTYPE_BF16 && dtype != CXL_DTYPE_F32 && dtype != CXL_DTYPE_F16) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  const int p_stride = dtype != CXL_DTYPE_F32 ? img_row_stride(P) : P;
  const long long ld_arg = cxl_argmax_ld(Bt, T);
  return maxsim_bwd(txt, img, shard_argmax_base(argmax, ld_arg, P, k_off), ld_arg, 0, w, Bt, Bi, T, P, D, dtype,
                    p_stride, ld_pairs, k_off, dtxt, accumulate_txt, dimg, accumulate_img, 0, ws, ws_bytes, w_scale_dev,
                    static_cast<cudaStream_t>(stream));
}

int cxl_maxsim_bwd_dtxt_rows_shard(const void* img, const void* argmax, const float* w, int Bt_total, int m0, int m_cnt,
                                   int Bi, int T, int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt,
                                   int accumulate_txt, const float* w_scale_dev, void* stream) {
  if (img == nullptr || argmax == nullptr || w == nullptr || dtxt == nullptr) return CXL_ERR_INVALID_ARG;
  if (Bt_total <= 0 || m0 < 0 || m_cnt <= 0 || m0 + m_cnt > Bt_total || k_off < 0) return CXL_ERR_INVALID_ARG;
  if (Bi <= 0 || T <= 0 || P <= 0 || D <= 0) return CXL_ERR_INVALID_ARG;
  if (dtype != CXL_DTYPE_BF16 && dtype != CXL_DTYPE_F32 && dtype != CXL_DTYPE_F16) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  const int p_stride = dtype != CXL_DTYPE_F32 ? img_row_stride(P) : P;
  const long long ld_arg = cxl_argmax_ld(Bt_total, T);
  return maxsim_bwd(img, img, shard_argmax_base(argmax, ld_arg, P, k_off), ld_arg, static_cast<long long>(m0) * T,
                    w + static_cast<size_t>(m0) * ld_pairs, m_cnt, Bi, T, P, D, dtype, p_stride, ld_pairs, k_off, dtxt,
                    accumulate_txt, nullptr, 0, 0, nullptr, 0, w_scale_dev, static_cast<cudaStream_t>(stream));
}

int cxl_maxsim_bwd(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T, int P,
                   int D, int dtype, int ld_pairs, int k_off, float* dtxt, float* dimg, int accumulate_img,
                   void* stream) {
  return bwd_entry(txt, img, argmax, w, Bt, Bi, T, P, D, dtype, ld_pairs, k_off, dtxt, dimg, accumulate_img, 0, nullptr,
                   0, stream);
}

size_t cxl_maxsim_bwd_workspace_bytes(int Bt, int Bi, int T, int P, int D, int dtype) {
  if ((dtype != CXL_DTYPE_BF16 && dtype != CXL_DTYPE_F16) || Bt <= 0 || Bi <= 0 || T <= 0 || P <= 0 || D <= 0) return 0;
  return maxsim_bwd_workspace_bytes(Bt, Bi, T, P, D);
}

int cxl_maxsim_bwd_ws(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T,
                      int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt, float* dimg, int accumulate_img,
                      void* ws, size_t ws_bytes, void* stream) {
  return bwd_entry(txt, img, argmax, w, Bt, Bi, T, P, D, dtype, ld_pairs, k_off, dtxt, dimg, accumulate_img, 0, ws,
                   ws_bytes, stream);
}

size_t cxl_infonce_workspace_bytes(int B_loc, int B) { return infonce_workspace_bytes(B_loc, B); }

int cxl_infonce_stats(const float* X, const float* scale_dev, int B_loc, int B, int row0, float* row_lse,
                      float* col_max, float* col_sum, float* diag, void* ws, size_t ws_bytes, void* stream) {
  if (!X || !scale_dev || !row_lse || !diag) return CXL_ERR_INVALID_ARG;   // col_max/col_sum/ws: all null = rows only
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return infonce_stats(X, scale_dev, B_loc, B, row0, row_lse, col_max, col_sum, diag, ws, ws_bytes,
                       static_cast<cudaStream_t>(stream));
}

int cxl_infonce_stats_static(const float* X, const float* scale_dev, int B_loc, int B, int row0, float* row_lse,
                             float* col_max, float* col_sum, float* diag, void* stream) {
  if (!X || !scale_dev || !row_lse || !col_max || !col_sum || !diag) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return infonce_stats_static(X, scale_dev, B_loc, B, row0, row_lse, col_max, col_sum, diag,
                              static_cast<cudaStream_t>(stream));
}

int cxl_infonce_grad(const float* X, const float* scale_dev, const float* coef_dev, int B_loc, int B, int row0,
                     const float* row_lse, const float* col_lse, const float* cnt, float* dX, float* dscale_accum,
                     void* stream) {
  if (!X || !scale_dev || !coef_dev || !row_lse || !dX || !dscale_accum) return CXL_ERR_INVALID_ARG;  // col_lse may be null
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return infonce_grad(X, scale_dev, coef_dev, B_loc, B, row0, row_lse, col_lse, cnt, dX, dscale_accum,
                      static_cast<cudaStream_t>(stream));
}

size_t cxl_infonce_grad_workspace_bytes(int B_loc, int B) { return infonce_grad_workspace_bytes(B_loc, B); }

int cxl_infonce_grad_ws(const float* X, const float* scale_dev, const float* coef_dev, int B_loc, int B, int row0,
                        const float* row_lse, const float* col_lse, const float* cnt, float* dX, float* dscale_accum,
                        void* ws, size_t ws_bytes, void* stream) {
  if (!X || !scale_dev || !coef_dev || !row_lse || !dX || !dscale_accum) return CXL_ERR_INVALID_ARG;  // col_lse may be null
  if (ws == nullptr || ws_bytes < infonce_grad_workspace_bytes(B_loc, B)) return CXL_ERR_WORKSPACE;
  if ((reinterpret_cast<uintptr_t>(ws) & 3) != 0) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return infonce_grad(X, scale_dev, coef_dev, B_loc, B, row0, row_lse, col_lse, cnt, dX, dscale_accum,
                      static_cast<cudaStream_t>(stream), static_cast<float*>(ws));
}

int cxl_l2norm_fwd(const void* x, int x_dtype, long long rows, int D, void* y, int y_dtype, float* inv_norm,
                   void* stream) {
  if (!x || !y) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return l2norm_fwd(x, x_dtype, rows, D, y, y_dtype, inv_norm, static_cast<cudaStream_t>(stream));
}

int cxl_l2norm_bwd(const void* y, int y_dtype, const void* g, int g_dtype, const float* inv_norm, long long rows,
                   int D, void* dx, int dx_dtype, void* stream) {
  if (!y || !g || !inv_norm || !dx) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return l2norm_bwd(y, y_dtype, g, g_dtype, inv_norm, rows, D, dx, dx_dtype, static_cast<cudaStream_t>(stream));
}

int cxl_gemm_nt(const void* A, const void* B, int M, int N, int K, int dtype, float* C, long long ldc, void* stream) {
  if (!A || !B || !C) return CXL_ERR_INVALID_ARG;
  if (dtype != CXL_DTYPE_BF16 && dtype != CXL_DTYPE_F16) return CXL_ERR_INVALID_ARG;
  if ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return gemm_nt_tc(A, B, M, N, K, C, ldc, dtype == CXL_DTYPE_F16, static_cast<cudaStream_t>(stream));
}

int cxl_cast_transpose(const float* X, int R, int C, void* Y, void* Yt, int dtype, const float* scale_dev, void* stream) {
  if (!X || (!Y && !Yt)) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return cast_transpose(X, R, C, Y, Yt, dtype, scale_dev, static_cast<cudaStream_t>(stream));
}

int cxl_head_ln_in(const void* x, int x_dtype, long long rows, int width, const float* gamma, const float* beta, float eps,
                   const int* valid_len, int T, void* y, int y_dtype, void* stream) {
  if (!x || !y || !gamma || !beta) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return head_ln_in(x, x_dtype, rows, width, gamma, beta, eps, valid_len, T, y, y_dtype, static_cast<cudaStream_t>(stream));
}

int cxl_head_out(const float* h, long long rows, int D, const float* bias, const float* gamma, const float* beta,
                 float eps, int l2_normalize, void* y, int y_dtype, void* stream) {
  if (!h || !y || !bias || !gamma || !beta) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return head_out(h, rows, D, bias, gamma, beta, eps, l2_normalize, y, y_dtype, static_cast<cudaStream_t>(stream));
}

int cxl_rank_of_target(const float* scores, long long row_stride, long long col_stride, int rows, int cols,
                       const int* target, int* rank_out, void* stream) {
  if (!scores || !target || !rank_out) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return rank_of_target(scores, row_stride, col_stride, rows, cols, target, rank_out, static_cast<cudaStream_t>(stream));
}

// Debug aid: run the tensor-core kernel on ONE text tile x ONE image tile (P = n_tile columns,
// T = 128 rows) and dump the raw fp32 accumulators.  S/cnt scratch are allocated by the caller
// through `out` being followed by nothing: we use stream-ordered temporaries here instead.
int cxl_debug_tile(const void* txt_bf16, const void* img_bf16, int rowsA, int rowsB, int D, int n_tile, float* out,
                   void* stream) {
  if (!txt_bf16 || !img_bf16 || !out) return CXL_ERR_INVALID_ARG;
  if (rowsA < 1 || rowsA > 128 || rowsB < 129 || rowsB > 256) return CXL_ERR_INVALID_ARG;  // single-image tile path
  (void)n_tile;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* scratch = nullptr;
  CXL_CUDA_CHECK(cudaMallocAsync(&scratch, 2 * sizeof(float), st));
  // one "text" of rowsA tokens against one "image" of rowsB patches
  rc = maxsim_fwd_tc(txt_bf16, img_bf16, 1, 1, rowsA, rowsB, D, CXL_MEAN_PLAIN, scratch, scratch + 1, nullptr, 0, out, 0,
                     0, 0, st);
  cudaFreeAsync(scratch, st);
  return rc;
}

// Test hook: tensor-core forward with explicit scheduling overrides (stream-A mode, grid size).
int cxl_maxsim_fwd_tc_ex(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int mean_mode, float* S,
                         float* cnt, void* argmax, int force_stream_a, int force_grid, void* stream) {
  if (txt == nullptr || img == nullptr || S == nullptr) return CXL_ERR_INVALID_ARG;
  int rc = device_ok();
  if (rc != CXL_OK) return rc;
  return maxsim_fwd_tc(txt, img, Bt, Bi, T, P, D, mean_mode, S, cnt, argmax, cxl_argmax_ld(Bt, T), nullptr,
                       force_stream_a, force_grid, 0, static_cast<cudaStream_t>(stream));
}

// Test hook: backward through the generic (non-streaming) kernels regardless of shape.
int cxl_maxsim_bwd_generic(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T,
                           int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt, float* dimg,
                           int accumulate_img, void* stream) {
  return bwd_entry(txt, img, argmax, w, Bt, Bi, T, P, D, dtype, ld_pairs, k_off, dtxt, dimg, accumulate_img, 1, nullptr, 0, stream);
}

}  // extern "C"
