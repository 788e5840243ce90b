/*
 * colxlip_b200 -- C-ABI of the B200-native MaxSim contrastive scoring + loss path.
 *
 * The reference (lezhang7/colxlip) has no FFI: its boundary for this path is the Python
 * API of src/colxlip/loss.py.  The entry points below are what a binding of that API
 * calls; each one cites the reference lines it replaces.  Conventions:
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - no allocation inside: the caller passes outputs and workspaces;
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*), nothing synchronises;
 *   - return value 0 = success, negative = CXL_ERR_* (see cxl_strerror); no exceptions;
 *   - thread-compatible; two calls may not share a workspace concurrently.
 * Layouts are row-major contiguous:  txt [Bt,T,D], img [Bi,P,D], S [Bt,Bi], grads shaped like
 * their inputs.  The saved argmax is IMAGE-major: argmax[k][r] = best patch of image k for the
 * flat text-token row r = m*T + n, row pitch cxl_argmax_ld(Bt,T) elements (Bt*T rounded up to
 * 16), element uint8 if P <= 256 else uint16 -- so one image's items are one contiguous stream
 * for the backward and a row tile's items are a contiguous run per image.
 */
#ifndef COLXLIP_B200_H_
#define COLXLIP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CXL_OK 0
#define CXL_ERR_INVALID_ARG (-1)   /* bad shape / null pointer / unsupported combination   */
#define CXL_ERR_UNSUPPORTED (-2)   /* valid request this build cannot serve (e.g. D % 64)   */
#define CXL_ERR_CUDA (-3)          /* a CUDA runtime/driver call failed (see cxl_last_cuda_error) */
#define CXL_ERR_NO_DEVICE (-4)     /* no sm_100 device: there is NO CPU fallback             */
#define CXL_ERR_WORKSPACE (-5)     /* workspace too small                                   */
#define CXL_ERR_STATIC_SCRATCH (-6) /* problem needs more than the built-in scratch of a *_static entry point */

#define CXL_DTYPE_F32 0
#define CXL_DTYPE_BF16 1
#define CXL_DTYPE_F16 2 /* fp16 operands (the reference's `--precision amp`): same kernels, kind::f16 with F16 formats */

#define CXL_MEAN_MASKED 0 /* loss.py:38-44  sum / (count(max != 0) + 1e-8) */
#define CXL_MEAN_PLAIN 1  /* model.py:628   mean over all T tokens          */
/* OR into mean_mode of cxl_maxsim_fwd (16-bit dtypes only): every row maximum is rounded to the operand dtype (RN)
 * before it enters the sum and the != 0 test -- the arithmetic of the reference under torch.autocast, whose einsum
 * (loss.py:34) returns bf16 / fp16 similarities to torch.max while `sum` and the division run in fp32.  The argmax is
 * unchanged: the fp32 maximum is always one of the patches that tie at 16-bit resolution. */
#define CXL_MEAN_ROUND16 0x10

const char* cxl_strerror(int code);

/* Number of CUDA kernels this library has launched in the process so far (all entry points; the dImg backward
 * launches two kernels per chunk of images, so the count per call depends on the shape). */
unsigned long long cxl_kernel_launches(void);
const char* cxl_last_cuda_error(void);
int cxl_version(void);
/* 0 if the current device can run the kernels (compute capability 10.x), else CXL_ERR_NO_DEVICE. */
int cxl_device_check(void);
/* size in bytes of one argmax element for P patches (1 or 2). */
int cxl_argmax_elem_bytes(int P);
/* row pitch (elements) of the image-major argmax buffer: total elements = Bi * cxl_argmax_ld(Bt, T). */
long long cxl_argmax_ld(int Bt, int T);

/*
 * Pack/cast token embeddings for the tensor-core path: src (f32 or bf16, row stride
 * `src_row_stride` elements between consecutive tokens, `src_item_stride` between items)
 * -> dst bf16 [items, n_pad, D] contiguous, rows >= n zero-filled.
 * Replaces the implicit autocast copy torch.einsum performs at loss.py:34.
 */
int cxl_pack_bf16(const void* src, int src_dtype, long long src_item_stride, long long src_row_stride,
                  int items, int n, int n_pad, int D, void* dst, void* stream);
/* Same with the operand dtype chosen by the caller: dst_dtype CXL_DTYPE_BF16 or CXL_DTYPE_F16 (fp16 autocast,
 * train.py:131-134 with --precision amp); src f32 / bf16 / f16. */
int cxl_pack_operand(const void* src, int src_dtype, long long src_item_stride, long long src_row_stride,
                     int items, int n, int n_pad, int D, void* dst, int dst_dtype, void* stream);

/*
 * MaxSim forward.  Replaces compute_colbert_similarity (loss.py:20-46) when
 * mean_mode == CXL_MEAN_MASKED and ColXLIP.compute_colbert_similarity (model.py:611-629)
 * when mean_mode == CXL_MEAN_PLAIN.
 *   dtype BF16: tcgen05/TMEM tensor-core kernel fed by TMA (fp32 accumulate, max over fp32);
 *               requires D % 64 == 0.  `img` must be laid out [Bi, p_stride, D] where
 *               p_stride = cxl_img_row_stride(P) (== P unless P < 128, see cxl_pack_bf16).
 *   dtype F32 : SIMT fp32 kernel (parity mode, rel <= 1e-5), img is [Bi, P, D], D % 4 == 0.
 *   S      [Bt,Bi] f32 out (final scores).
 *   cnt    [Bt,Bi] f32 out, count of non-zero maxima (no epsilon); may be NULL for MEAN_PLAIN.
 *   argmax [Bi][cxl_argmax_ld(Bt,T)] out or NULL (inference: nothing saved).
 * Non-finite inputs: torch.max propagates a NaN similarity; the kernels' running max does not, so the operands are
 * scanned and EVERY score of a text / an image that holds a NaN or Inf element is returned as NaN (the reference can
 * return +Inf instead of NaN for an Inf operand; either way the loss and its gradients are non-finite, which is what
 * AMP's GradScaler tests for).
 */
int cxl_img_row_stride(int P);
int cxl_maxsim_fwd(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int dtype,
                   int mean_mode, float* S, float* cnt, void* argmax, void* stream);
/*
 * The same with run-to-run reproducible scores (loss.py:38-44 sums a text's T maxima in one fixed order; torch's
 * reduction is deterministic).  cxl_maxsim_fwd adds the partial sums of the 3-5 warps that hold one text's rows into
 * S[m,k] with atomicAdd, so S can differ in its last bit between runs (cnt and the argmax cannot).  With a workspace
 * every 32-row block (64-row for F32) STORES its partial sum of text m into its own slot -- slot = block index relative to the block
 * that holds the text's first row, slot 0 is S itself, slots 1.. live in ws -- and the finalize kernel adds a text's
 * slots in ascending order: same inputs, same bits, on every run and for every split of the images into calls.
 * ws: cxl_maxsim_fwd_workspace_bytes(Bt, Bi, T, dtype) = ((T + 30) / 32) * Bt * Bi * 4 bytes for the 16-bit dtypes
 * (12.9 GB at cfg-2 on one GPU, 0.8 GB per shard call at cfg-5), contents undefined on entry and on return.
 */
size_t cxl_maxsim_fwd_workspace_bytes(int Bt, int Bi, int T, int dtype);
int cxl_maxsim_fwd_ws(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int dtype,
                      int mean_mode, float* S, float* cnt, void* argmax, void* ws, size_t ws_bytes, void* stream);

/*
 * Per-pair backward weights  w[m,k] = dS[m,k] * g / (cnt[m,k] + 1e-8)   (CXL_MEAN_MASKED, loss.py:40-44)
 *                                    = dS[m,k] * g / T                   (CXL_MEAN_PLAIN,  model.py:628)
 * over n = Bt*Bi contiguous pairs; gscale_dev is an optional device scalar g; w may alias dS.
 * (cxl_infonce_grad writes w directly when given cnt, so the fused loss never calls this.)
 */
int cxl_pair_weights(const float* dS, const float* cnt, const float* gscale_dev, int mean_mode, int T, long long n,
                     float* w, void* stream);

/*
 * MaxSim backward from the saved argmax.  Replaces autograd of loss.py:34-44
 * (value_selecting_reduction_backward + two bmm on a one-hot dS):
 *   dtxt[m,n,:] = sum_k w[m,k] * img[k, argmax(m,k,n), :]      [Bt,T,D] f32 out (overwritten) or NULL
 *   dimg[k,q,:] = sum_{m,n: argmax(m,k,n)=q} w[m,k] * txt[m,n,:]  [Bi,P,D] f32 out (overwritten unless
 *                                                                accumulate_img != 0) or NULL
 *   w [Bt, ld_pairs] f32 pair weights (cxl_pair_weights / cxl_infonce_grad).
 * txt/img: dtype F32 -> [.,.,D] contiguous; BF16 -> img uses row stride cxl_img_row_stride(P).
 * Image-shard calls (multi-GPU reduce-scatter pipeline): `img`/`dimg` point at the first of Bi
 * images of a shard, while w / argmax are the forward's full buffers over `ld_pairs` images and
 * the shard starts at image `k_off` of them (single call: ld_pairs = Bi, k_off = 0).
 * bf16 shapes with D % 128 == 0, T >= 8 and P <= 576 run the TMA-staged kernels (dTxt gather; dImg = row-list
 * pre-pass + gather, which takes up to CXL_DIMG_SCRATCH_MB (default 4608) MB of stream-ordered scratch from the
 * device's default memory pool); the rest a generic kernel.
 */
int cxl_maxsim_bwd(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T, int P,
                   int D, int dtype, int ld_pairs, int k_off, float* dtxt, float* dimg, int accumulate_img,
                   void* stream);

/* Same computation (autograd of loss.py:34-44), with caller-provided scratch for the dImg row lists (no
 * allocation inside the library): this is the entry the Python mirror binds.
 * cxl_maxsim_bwd_workspace_bytes() is the preferred size for a call with these dimensions (0 when the shape
 * needs none); a smaller `ws` makes the library work through the images in smaller chunks, CXL_ERR_WORKSPACE
 * if not even one CTA's images fit.  `ws` must be 16-byte aligned and is only used until the call's kernels
 * finish on `stream`.  cxl_maxsim_bwd() itself takes the same scratch from the device's default memory pool. */
size_t cxl_maxsim_bwd_workspace_bytes(int Bt, int Bi, int T, int P, int D, int dtype);
int cxl_maxsim_bwd_ws(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T,
                      int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt, float* dimg, int accumulate_img,
                      void* ws, size_t ws_bytes, void* stream);

/*
 * dTxt of a SLICE of the texts: rows m0 .. m0+m_cnt-1 of the forward's Bt_total texts, against images
 * k_off .. k_off+Bi-1.  `argmax` and `w` are the forward's full buffers (pitches cxl_argmax_ld(Bt_total,T) and
 * ld_pairs), `img` points at image k_off, dtxt [m_cnt,T,D] f32 is overwritten.  Serves the reference's
 * gradient-accumulation loop (train.py:166-183), where only one sub-batch of the concatenated features
 * carries gradient: together with an image-shard call of cxl_maxsim_bwd_ws (dimg only, k_off = the sub-batch)
 * the backward does 1/accum_freq of the full work.  The TMA-staged kernel needs m0*T*elem_bytes % 16 == 0;
 * other offsets run the generic kernel.
 */
int cxl_maxsim_bwd_dtxt_rows(const void* img, const void* argmax, const float* w, int Bt_total, int m0, int m_cnt,
                             int Bi, int T, int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt,
                             int force_generic, void* stream);

/*
 * The same backward for ONE SHARD of the images with shard-local operands -- the multi-GPU pipeline (reference
 * loss.py:48-92 + autograd of loss.py:34-44): each rank scores its texts against one home rank's images at a time,
 * keeps that shard's operand and argmax as separate buffers (so they can be freed shard by shard: at ViT-L/14@336,
 * B = 65536 on 8 GPUs the gathered operand + argmax are 131 GiB per rank) and sends the dImg partial home.
 *   img    [Bi, p_stride, D]   the shard's image operand          argmax [Bi][cxl_argmax_ld(Bt,T)]  the shard's own
 *   w      [Bt, ld_pairs]      pair weights over ALL images; the shard is columns [k_off, k_off + Bi)
 *   dtxt   [Bt,T,D] or NULL; accumulate_txt != 0: dtxt += this shard's sum.  Shards taken in ascending image order
 *          continue the same chain of additions, so the result has the bits of one call over all images.
 *   dimg   [Bi,P,D] or NULL (the shard's partial), accumulate_img as above.
 * A sub-range [lo, lo+n) of a shard is the same call with img / argmax / dimg advanced by lo images and k_off + lo.
 *   w_scale_dev  NULL, or a device pair {s, 1/s} with s a power of two: the streaming kernels multiply with a 16-bit
 *          copy of w (bf16, or fp16 for CXL_DTYPE_F16 operands); s is applied before that rounding and undone exactly
 *          at the store, which keeps fp16 weights inside their range whatever the loss scale (GradScaler) is.
 */
int cxl_maxsim_bwd_shard(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T,
                         int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt, int accumulate_txt,
                         float* dimg, int accumulate_img, const float* w_scale_dev, void* ws, size_t ws_bytes,
                         void* stream);
/* cxl_maxsim_bwd_dtxt_rows for one shard with shard-local img / argmax (see above), optionally accumulating. */
int cxl_maxsim_bwd_dtxt_rows_shard(const void* img, const void* argmax, const float* w, int Bt_total, int m0, int m_cnt,
                                   int Bi, int T, int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt,
                                   int accumulate_txt, const float* w_scale_dev, void* stream);

/*
 * Symmetric InfoNCE over a [B_loc, B] block of UNSCALED logits (rows = texts row0..row0+B_loc,
 * columns = all B images; labels are the diagonal).  Replaces `logit_scale *` and the
 * F.cross_entropy pairs of loss.py:252-263, 285-293.
 * Stats pass: row log-sum-exp [B_loc] and per-rank column (max, sumexp) partials [2,B] of scale*X.
 * col_max, col_sum and ws may all be NULL: only the row statistics are computed (the one-directional CEs of
 * ClipLoss with local_loss=True, loss.py:144-146).
 */
size_t cxl_infonce_workspace_bytes(int B_loc, int B);
int cxl_infonce_stats(const float* X, const float* scale_dev, int B_loc, int B, int row0,
                      float* row_lse /*[B_loc]*/, float* col_max /*[B]*/, float* col_sum /*[B]*/,
                      float* diag /*[B_loc] scale*X[m,row0+m]*/, void* ws, size_t ws_bytes, void* stream);
/*
 * Allocation-free variant of cxl_infonce_stats (both directions): the column partials live in a file-scope
 * __device__ array inside the library instead of a caller workspace, so the call needs no cudaMalloc on either side
 * (CUDA-graph capture, tight serving loops, harnesses that forbid allocation).  NO COUNTERPART IN THE REFERENCE.
 * Same kernels, same results as cxl_infonce_stats.
 *   - limit: cxl_infonce_workspace_bytes(B_loc, B) <= CXL_INFONCE_STATIC_SCRATCH_BYTES (16 MiB: e.g. B_loc = B <= 16384,
 *     or one rank of 8 at B = 65536 with B_loc = 4096); larger problems return CXL_ERR_STATIC_SCRATCH and launch nothing;
 *   - col_max and col_sum are required (the rows-only form of cxl_infonce_stats needs no scratch at all);
 *   - every GPU that loads the library holds the 16 MiB until the process exits;
 *   - all calls on one device share the array: calls that may overlap on the GPU (different streams) race -- order
 *     them on one stream or with events.  Nothing in the array has to be cleared between calls.
 */
#define CXL_INFONCE_STATIC_SCRATCH_BYTES (16u << 20)
int cxl_infonce_stats_static(const float* X, const float* scale_dev, int B_loc, int B, int row0,
                             float* row_lse /*[B_loc]*/, float* col_max /*[B]*/, float* col_sum /*[B]*/,
                             float* diag /*[B_loc]*/, void* stream);
/*
 * Gradient pass: dX[m,k] = coef * scale * (exp(scale*X - row_lse[m]) + exp(scale*X - col_lse[k]) - 2*delta)
 * with coef = *coef_dev / (2B); dscale_partial += sum(dX/scale * X) (one float, atomically added).
 * If cnt != NULL (token logits) the output is the MaxSim pair weight dX / (cnt + 1e-8) instead of dX.
 * If col_lse == NULL only the row cross-entropy is differentiated: dX = coef * scale * (exp(scale*X - row_lse[m]) - delta)
 * -- the two one-directional CEs of ClipLoss with local_loss=True (loss.py:144-146,173-176); the caller folds
 * its own normalisation into *coef_dev (the kernel still divides by 2B).
 * dX may alias X.
 */
int cxl_infonce_grad(const float* X, const float* scale_dev, const float* coef_dev, int B_loc, int B, int row0,
                     const float* row_lse, const float* col_lse /*[B] global, or NULL*/, const float* cnt /*[B_loc,B] or NULL*/,
                     float* dX, float* dscale_accum, void* stream);
/* The same with a reproducible d logit_scale: every block stores its partial of sum(dX/scale * X) into ws
 * (cxl_infonce_grad_workspace_bytes: one float per block, 2 MB at B_loc = B = 32768) and one fixed-order pass adds them
 * into *dscale_accum, instead of one atomicAdd per block.  dX is identical to cxl_infonce_grad's. */
size_t cxl_infonce_grad_workspace_bytes(int B_loc, int B);
int cxl_infonce_grad_ws(const float* X, const float* scale_dev, const float* coef_dev, int B_loc, int B, int row0,
                        const float* row_lse, const float* col_lse, const float* cnt, float* dX, float* dscale_accum,
                        void* ws, size_t ws_bytes, void* stream);

/*
 * Global-embedding logits and their gradient GEMMs on the tcgen05 pipeline of cxl_maxsim_fwd (plain-GEMM epilogue).
 * Replaces `image_features @ text_features.T` and its autograd (loss.py:252-253 in ColClipLoss, 132-160 in ClipLoss):
 *   C[M,N] (f32, row pitch ldc) = A[M,K] * B[N,K]^T,  A / B row-major 16-bit (dtype CXL_DTYPE_BF16 or CXL_DTYPE_F16,
 *   what the reference's autocast matmul multiplies), fp32 accumulate.  K % 64 == 0, A / B 16-byte aligned;
 *   other shapes: CXL_ERR_UNSUPPORTED (the caller keeps its library GEMM).
 * cxl_cast_transpose: X [R,C] f32 -> Y [R,C] and / or Yt [C,R] in the 16-bit dtype (either may be NULL): the operand
 * copies of dL/dlogits for  d_txt = dG * img  and  d_img = dG^T * txt.  scale_dev: NULL or a device scalar multiplied
 * in before the cast (a power of two keeps fp16 copies of small gradients out of the subnormal range).
 */
int cxl_gemm_nt(const void* A, const void* B, int M, int N, int K, int dtype, float* C, long long ldc, void* stream);
int cxl_cast_transpose(const float* X, int R, int C, void* Y, void* Yt, int dtype, const float* scale_dev, void* stream);

/*
 * L2 normalisation of embedding rows fused with the cast to the operand dtype.  Replaces
 * F.normalize(x, dim=-1) at the end of ColXLIP.encode_image / encode_text (model.py:551-554, 605-607) and the
 * autocast copy einsum makes of its result (loss.py:34):
 *   y[r,:] = x[r,:] / max(||x[r,:]||_2, 1e-12)       x [rows,D] f32|bf16 -> y [rows,D] bf16|f32, contiguous
 *   inv_norm[r] = 1 / max(||x[r,:]||, 1e-12)         [rows] f32, kept for the backward (may be NULL: inference)
 * and its backward  dx[r,:] = (g[r,:] - y[r,:] * <g[r,:], y[r,:]>) * inv_norm[r].
 * D % 8 == 0; all row pointers 16-byte aligned.
 */
int cxl_l2norm_fwd(const void* x, int x_dtype, long long rows, int D, void* y, int y_dtype, float* inv_norm,
                   void* stream);
int cxl_l2norm_bwd(const void* y, int y_dtype, const void* g, int g_dtype, const float* inv_norm, long long rows,
                   int D, void* dx, int dx_dtype, void* stream);

/*
 * Token heads of ColXLIP (model.py:518-530: LayerNorm -> Linear -> GELU -> LayerNorm), the EOT masking of the text
 * tokens in front of them (model.py:580-593) and the L2 normalise + operand cast behind them (model.py:551-554,
 * 605-607; loss.py:34), as three launches: cxl_head_ln_in -> cxl_gemm_nt -> cxl_head_out.
 *   cxl_head_ln_in: y[r,:] = LayerNorm(x[r,:]) * gamma + beta over `width` channels, written in the 16-bit GEMM dtype;
 *       valid_len (device int [rows / T], may be NULL): rows with (r % T) >= valid_len[r / T] are computed from a ZERO
 *       input, as the reference zeroes the tokens at and after the EOT position before the head (so y = beta there).
 *   cxl_head_out:   y[r,:] = [L2 normalise] LayerNorm(GELU(h[r,:] + bias)) * gamma + beta, h [rows, D] fp32 = the GEMM
 *       output; exact (erf) GELU; y in bf16 / fp16 (K1's operand layout for P >= 128) or f32.
 * width and D: multiples of 256, at most 1024; gamma / beta / bias fp32, 16-byte aligned.
 */
int cxl_head_ln_in(const void* x, int x_dtype, long long rows, int width, const float* gamma, const float* beta, float eps,
                   const int* valid_len, int T, void* y, int y_dtype, void* stream);
int cxl_head_out(const float* h, long long rows, int D, const float* bias, const float* gamma, const float* beta,
                 float eps, int l2_normalize, void* y, int y_dtype, void* stream);

/*
 * Retrieval ranking: rank of the ground-truth column in every row of a score matrix
 * (number of entries strictly greater, ties broken by lower index first, as a stable
 * descending argsort does).  Replaces the per-row argsort loop of train.py:467-469.
 * A target outside [0, cols) is not dereferenced; its rank is -1 (the reference raises IndexError there).
 */
int cxl_rank_of_target(const float* scores, long long row_stride, long long col_stride, int rows, int cols,
                       const int* target /*[rows]*/, int* rank_out /*[rows]*/, void* stream);

/* Test hook: tensor-core forward with scheduling overrides (force the streamed-A pipeline, cap the grid). */
int cxl_maxsim_fwd_tc_ex(const void* txt, const void* img, int Bt, int Bi, int T, int P, int D, int mean_mode,
                         float* S, float* cnt, void* argmax, int force_stream_a, int force_grid, void* stream);

/* Test hook: cxl_maxsim_bwd through the generic (non-streaming) kernels regardless of shape. */
int cxl_maxsim_bwd_generic(const void* txt, const void* img, const void* argmax, const float* w, int Bt, int Bi, int T,
                           int P, int D, int dtype, int ld_pairs, int k_off, float* dtxt, float* dimg,
                           int accumulate_img, void* stream);

/* Debug aid: raw fp32 accumulators of the first 128 x n_tile tile of the tensor-core kernel. */
int cxl_debug_tile(const void* txt_bf16, const void* img_bf16, int rowsA, int rowsB, int D, int n_tile,
                   float* out /*[128, n_tile]*/, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* COLXLIP_B200_H_ */
