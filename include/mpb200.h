/* mpb200.h -- C ABI of libmpb200.so (mpi_pytorch_b200/lib/libmpb200.so), sm_100a only.
 *
 * Every function returns 0 on success, a cudaError_t value (> 0) when a CUDA call or launch failed,
 * or a small negative number for an argument the kernel cannot handle (shape / alignment).
 * Pointers are device pointers unless stated otherwise; `stream` is a cudaStream_t passed as void*.
 * dtype: 0 = fp32, 1 = bf16.  4-D activations are NHWC (channels-last), C must be a multiple of 8.
 * Citations point into the reference (/root/reference) at the call each entry point replaces.
 */
#ifndef MPB200_H_
#define MPB200_H_

#ifdef __cplusplus
extern "C" {
#endif

/* ---- K10 softmax cross-entropy, fused forward+backward (main.py:134,150,152) ------------------------------
 * loss_sum (1 float, zeroed inside) receives the MEAN loss; dlogits = (softmax - onehot)/B in logits' dtype. */
int mpb_softmax_ce(const void* logits, const long long* target, void* dlogits, float* loss_sum, float* row_loss,
                   int B, int C, int dtype, void* stream);
/* argmax + compare + count (main.py:182-183, evaluation_pipeline.py:157-158); count is one int64, zeroed inside. */
int mpb_count_correct(const void* logits, const long long* target, long long* count, int B, int C, int dtype,
                      void* stream);

/* ---- flat optimizers over an arena slice (main.py:155) ---------------------------------------------------- */
int mpb_adam_flat(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n, int step, float lr,
                  float b1, float b2, float eps, float wd, float grad_scale, void* stream);
int mpb_sgd_flat(float* p, const float* g, float* buf, void* shadow_bf16, long long n, int step, float lr,
                 float momentum, float wd, float grad_scale, void* stream);
int mpb_cast_f32_bf16(const float* src, void* dst, long long n, void* stream);

/* ---- K7 BatchNorm (+residual) (+ReLU), training and eval (torchvision BN layers built at models.py:33..87) -
 * x is [M = N*H*W, C].  `sums` / `dsum` are 2*C floats of caller-provided scratch, zeroed inside.           */
int mpb_bn_stats(const void* x, float* sums, long long M, int C, int dtype, void* stream);
int mpb_bn_fwd_apply(const void* x, const void* res, void* out, const float* sums, const float* gamma,
                     const float* beta, float* run_mean, float* run_var, float* save_mean, float* save_invstd,
                     long long M, int C, float eps, float momentum, int relu, int training, int dtype, void* stream);
int mpb_bn_bwd(const void* dz, const void* z, const void* x, const float* gamma, const float* save_mean,
               const float* save_invstd, float* dsum, void* dx, void* dres, float* dgamma, float* dbeta, long long M,
               int C, int relu, int dtype, void* stream);

/* BN -> ReLU backward (no residual): the ReLU mask is recomputed from x, so the forward output is never read. */
int mpb_bn_relu_bwd_recompute(const void* dz, const void* x, const float* gamma, const float* beta,
                              const float* save_mean, const float* save_invstd, float* dsum, void* dx, float* dgamma,
                              float* dbeta, long long M, int C, int dtype, void* stream);

/* Same kernels with the training-engine extras: `num_batches` (nullable) is the module's int64 step counter, bumped
 * inside the forward kernel; `flags` bit 0 adds dgamma/dbeta to what the buffers hold (slots of the flat gradient
 * arena) instead of overwriting them, bit 1 says `dsum` is already zero (skip the memset).                      */
int mpb_bn_fwd_apply_ex(const void* x, const void* res, void* out, const float* sums, const float* gamma,
                        const float* beta, float* run_mean, float* run_var, float* save_mean, float* save_invstd,
                        long long* num_batches, long long M, int C, float eps, float momentum, int relu, int training,
                        int dtype, void* stream);
int mpb_bn_bwd_ex(const void* dz, const void* z, const void* x, const float* gamma, const float* save_mean,
                  const float* save_invstd, float* dsum, void* dx, void* dres, float* dgamma, float* dbeta,
                  long long M, int C, int relu, int flags, int dtype, void* stream);
int mpb_bn_relu_bwd_recompute_ex(const void* dz, const void* x, const float* gamma, const float* beta,
                                 const float* save_mean, const float* save_invstd, float* dsum, void* dx,
                                 float* dgamma, float* dbeta, long long M, int C, int flags, int dtype,
                                 void* stream);

/* BatchNorm backward without caller scratch.  This entry point has no counterpart in the reference: it exists
 * for hosts that cannot allocate device memory themselves (all of the GPU already belongs to their own pool,
 * or the implicit synchronisation of cudaMalloc/cudaFree is not acceptable where they call us).  Same maths
 * and argument meaning as mpb_bn_bwd minus `dsum`; the 2*C floats live in a fixed 16 KB device array inside
 * the library, so C <= 2048 (larger C returns -4).  The first call on a device performs a one-time address
 * look-up and should happen before any stream capture; calls that can overlap on the GPU share the array and
 * would race -- keep them ordered.  Calls no cudaMalloc/cudaFree.                                            */
int mpb_bn_bwd_noalloc(const void* dz, const void* z, const void* x, const float* gamma, const float* save_mean,
                       const float* save_invstd, void* dx, void* dres, float* dgamma, float* dbeta, long long M,
                       int C, int relu, int dtype, void* stream);

/* ---- K8 pooling, K11 elementwise (ATen pooling / relu / dropout reached from models.py) ------------------- */
int mpb_maxpool_fwd(const void* x, void* y, unsigned char* idx, int N, int H, int W, int C, int Ho, int Wo, int kh,
                    int kw, int sh, int sw, int ph, int pw, int dtype, void* stream);
int mpb_maxpool_bwd(const void* dy, const unsigned char* idx, void* dx, int N, int H, int W, int C, int Ho, int Wo,
                    int kh, int kw, int sh, int sw, int ph, int pw, int dtype, void* stream);
int mpb_avgpool(const void* in, void* out, int N, int H, int W, int C, int Ho, int Wo, int kh, int kw, int sh, int sw,
                int ph, int pw, int count_include_pad, int backward, int dtype, void* stream);
int mpb_adaptive_avgpool(const void* in, void* out, int N, int H, int W, int C, int Ho, int Wo, int backward,
                         int dtype, void* stream);
int mpb_relu_fwd(const void* x, void* y, long long n, int dtype, void* stream);
int mpb_relu_bwd(const void* dy, const void* y, void* dx, long long n, int dtype, void* stream);
/* dx = dy * (y > 0) and sums[0:C] += column sums of dx in one pass (sums: 2*C zeroed floats): the backward of a
 * Conv2d(+bias) -> ReLU pair, where the column sums are the bias gradient (models.py:47-72 stacks).           */
int mpb_relu_bwd_colsum(const void* dy, const void* y, void* dx, float* sums, long long M, int C, int dtype,
                        void* stream);
int mpb_dropout(const void* x, void* y, long long n, unsigned int seed, float p, int dtype, void* stream);
/* M rows of row_bytes (multiple of 16) between row-pitched buffers: channel concat into an NHWC slice and its inverse
 * (torch.cat of the DenseNet / Inception / Fire blocks, models.py:56-87).                                       */
int mpb_copy_rows(const void* src, long long src_pitch, void* dst, long long dst_pitch, long long M, long long row_bytes,
                  void* stream);
/* dst[r][j] (+)= src[r][idx[j]] (idx[j] < 0 -> 0): the stem filter's relayout between KRSC and the space-to-depth
 * form of the tensor-core stem kernels (torchvision stems reached from models.py:33,41,77).  dtype 0 fp32, 1 bf16. */
int mpb_gather_cols(const void* src, const long long* idx, void* dst, int rows, int src_cols, int dst_cols,
                    int src_dtype, int dst_dtype, int accumulate, void* stream);
/* `epoch` (nullable device pointer): a step counter mixed into the seed on the device, so a captured launch draws a
 * new mask on every CUDA-graph replay; the backward call of the same step passes the same pointer and seed.      */
int mpb_dropout_ex(const void* x, void* y, long long n, unsigned int seed, const long long* epoch, float p, int dtype,
                   void* stream);

/* ---- K9 resize + normalise, K14 synthetic data (main.py:62-65, evaluation_pipeline.py:89,116-122) ---------
 * mode 0: u8 -> u8 NHWC; 1: u8 -> normalised float via a u8 rounding (PIL semantics); 2: u8 -> float directly. */
int mpb_resize(const unsigned char* in, void* out, int N, int Hi, int Wi, int Ho, int Wo, int bicubic, int mode,
               int out_nhwc, int out_dtype, void* stream);
int mpb_normalize_u8(const unsigned char* in, void* out, int N, int H, int W, int out_nhwc, int out_dtype,
                     void* stream);
int mpb_synth_images(void* out, const long long* index, int n, int C, int H, int W, long long seed, int out_nhwc,
                     int out_dtype, void* stream);

/* ---- K3-K6 tcgen05 / TMEM / TMA GEMM and implicit-GEMM convolution (every nn.Conv2d / nn.Linear) ----------
 * bf16 operands, fp32 accumulation.  GEMM: D[M,N] = A[M,K] * B[N,K]^T with either operand K-major
 * ([rows][K]) or MN-major ([K][rows]).  out_mode 0: bf16, 1: fp32, 2: fp32 atomic accumulate (split-K).     */
int mpb_tc_gemm(const void* A, const void* B, void* D, const float* bias, float* stats, int M, int N, int K,
                long long lda, long long ldb, long long ldd, int a_mn, int b_mn, int out_mode, int splitk, void* stream);
/* explicit im2col (rows = output pixels, columns = (r,s,c) zero-padded to Kpad) for inputs whose channel count is
 * not a multiple of 8 (3-channel stems); feeds mpb_tc_gemm for fprop and wgrad.                               */
int mpb_im2col_nhwc(const void* x, void* out, int N, int H, int W, int C, int R, int S, int P, int Q, int stride,
                    int pad_h, int pad_w, int Kpad, void* stream);
/* y[N,P,Q,K] = conv(x[N,H,W,C], w[K,R,S,C]); C % 64 == 0, K % 16 == 0.  stats: optional 2*K floats, zeroed by the
 * caller, receive per-channel sum / sum of squares of the fp32 accumulator (BatchNorm statistics).            */
int mpb_tc_conv_fprop(const void* x, const void* w, void* y, const float* bias, float* stats, int N, int H, int W,
                      int C, int K, int R, int S, int P, int Q, int stride, int pad_h, int pad_w, int dil,
                      void* stream);
/* dx for stride-1 convolutions; the filter is read in place as an MN-major operand (no transposed copy).      */
int mpb_tc_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int K, int R, int S, int P,
                      int Q, int pad_h, int pad_w, void* stream);
/* Halo-tile variants for stride-1 filters larger than 1x1: the activation tile (with halo) is fetched once per
 * 64-channel block and every filter tap is a row-shifted UMMA view of it.  mpb_tc_halo_efficiency = fraction of the
 * 128 GEMM rows that are real output pixels for a geometry.                                                     */
float mpb_tc_halo_efficiency(int Ho, int Wo, int R, int S);
int mpb_tc_halo_persistent_ok(int Ci, int Co, int R, int S, int Ho, int Wo);
int mpb_tc_conv_fprop_halo(const void* x, const void* w, void* y, const float* bias, float* stats, int N, int H, int W,
                           int C, int K, int R, int S, int P, int Q, int pad_h, int pad_w, void* stream);
/* fprop with an activation epilogue: y = relu?(conv(x, w) + bias).  With BatchNorm folded into w / bias (eval mode:
 * w' = w * gamma / sqrt(var + eps), bias' = beta - mean * gamma / sqrt(var + eps)) one kernel is the Conv2d + BatchNorm2d
 * (+ ReLU) triple of a torchvision block (models.py:30-95 build them; evaluation_pipeline.py:155-156 runs them).   */
int mpb_tc_conv_fprop_act(const void* x, const void* w, void* y, const float* bias, float* stats, int N, int H, int W,
                          int C, int K, int R, int S, int P, int Q, int stride, int pad_h, int pad_w, int dil, int relu,
                          void* stream);
int mpb_tc_conv_fprop_halo_act(const void* x, const void* w, void* y, const float* bias, float* stats, int N, int H,
                               int W, int C, int K, int R, int S, int P, int Q, int pad_h, int pad_w, int relu,
                               void* stream);
int mpb_tc_conv_dgrad_halo(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int K, int R, int S,
                           int P, int Q, int pad_h, int pad_w, void* stream);
/* dx for stride >= 2: decomposed into stride^2 output-parity classes, each an implicit GEMM over the taps that reach it. */
int mpb_tc_conv_dgrad_strided(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int K, int R, int S,
                              int P, int Q, int stride, int pad_h, int pad_w, void* stream);
/* dw[K,R,S,C] (fp32) += split-K partial sums; dw must hold zeros or a partial gradient to accumulate into.     */
int mpb_tc_conv_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int K, int R, int S,
                      int P, int Q, int stride, int pad_h, int pad_w, int dil, int splitk, void* stream);
/* Persistent halo-tile wgrad for stride-1, dilation-1 filters larger than 1x1 (tc_wgrad_halo.cu): same contract
 * as mpb_tc_conv_wgrad.  mpb_tc_wgrad_halo_ok says whether a layer fits it (C == 64 or C % 128 == 0, K % 64 == 0,
 * accumulators within 512 TMEM columns, a >= 2-stage ring within shared memory); -2 otherwise.                   */
/* Variants taking explicit element strides of x between pixels / rows / images; a pixel stride smaller than C
 * describes overlapping super-pixels.  With mpb_s2d_stem (x NHWC, 4*C <= 16, H and W even -> out [N][H/2][Wp][16],
 * lpad zero pixels on the left) a 7x7 stride-2 stem becomes a 4x1 stride-1 convolution with 64 channels:
 * no im2col matrix (models.py:33,68,81 stems).                                                                  */
int mpb_s2d_stem(const void* x, void* out, int N, int H, int W, int C, int Wp, int lpad, void* stream);
int mpb_tc_conv_fprop_halo_strided(const void* x, const void* w, void* y, const float* bias, float* stats, int N, int H,
                                   int W, int C, int K, int R, int S, int P, int Q, int pad_h, int pad_w,
                                   long long xs_w, long long xs_h, long long xs_n, void* stream);
int mpb_tc_conv_wgrad_halo_strided(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int K, int R,
                                   int S, int P, int Q, int pad_h, int pad_w, long long xs_w, long long xs_h,
                                   long long xs_n, void* stream);
int mpb_tc_wgrad_halo_ok(int C, int K, int R, int S, int Ho, int Wo);
int mpb_tc_conv_wgrad_halo(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int K, int R, int S,
                           int P, int Q, int pad_h, int pad_w, void* stream);

/* ---- K1 / K2 / K13 peer-memory collectives over NVLink (mpi_tools.py:30-37,47-53; evaluation_pipeline.py:196)
 * Symmetric buffers: cudaMalloc on every rank, handles exchanged by the caller, mapped with the ipc helpers.
 * `grads/params/shadows/pads/bufs/vals` are HOST arrays of W device pointers indexed by rank, already offset
 * to the range being reduced.  `epochs` is this rank's private device array of mpb_epoch_bytes() bytes.       */
int mpb_sym_alloc(void** ptr, long long bytes);
int mpb_sym_free(void* ptr);
int mpb_ipc_get_handle(void* ptr, void* handle_out64);
int mpb_ipc_open_handle(const void* handle64, void** ptr);
int mpb_ipc_close_handle(void* ptr);
int mpb_ipc_handle_bytes(void);
int mpb_signal_pad_bytes(void);
int mpb_epoch_bytes(void);
long long mpb_oneshot_max_elems(int blocks);
/* algo 0 one-shot, 1 two-shot, 2 two-shot through NVLS multicast addresses (mc_* = multicast views of the ranges).  opt_kind 0: grads averaged in place; 1 Adam / 2 SGD-momentum: fp32 params and
 * optional bf16 shadows updated on every rank from the averaged gradient (two-shot: owner slice + all-gather).
 * step_dev: optional device int holding the optimizer step (used instead of `step` when non-null).            */
int mpb_allreduce_fused(void* const* grads, void* const* params, void* const* shadows, void* const* pads,
                        void* epochs, float* m, float* v, long long n, int rank, int W, int algo, int blocks,
                        int opt_kind, int step, float lr, float b1, float b2, float eps, float wd, float momentum,
                        const int* step_dev, void* mc_grads, void* mc_params, void* mc_shadow, void* stream);
/* VMM / NVLS multicast symmetric memory (file descriptors are exchanged by the caller, e.g. over SCM_RIGHTS). */
int mpb_vmm_multicast_supported(void);
long long mpb_vmm_round_size(long long bytes, int num_devices);
int mpb_vmm_alloc(long long bytes, void** ptr, int* fd_out, unsigned long long* handle_out);
int mpb_vmm_import(int fd, long long bytes, void** ptr);
int mpb_mc_create(long long bytes, int num_devices, int* fd_out, unsigned long long* mc_out);
int mpb_mc_import(int fd, unsigned long long* mc_out);
int mpb_mc_add_device(unsigned long long mc);
int mpb_mc_bind_map(unsigned long long mc, unsigned long long mem_handle, long long bytes, void** mc_ptr);
/* GRAD_DTYPE=bf16 form of mpb_allreduce_fused (same reference sites: mpi_tools.py:24-26,30-37 forces an fp32 wire,
 * main.py:155): `grads` is this rank's local fp32 bucket, packed by the kernel into the symmetric bf16 `wire`
 * buffers the peers read (2 bytes per element in the reduce-scatter), fp32 accumulation, Adam/SGD on the fp32
 * masters, all-gather of params + bf16 shadow.  opt_kind 1|2, algo 1 (peer) | 2 (NVLS), n % (8*W) == 0.        */
int mpb_allreduce_fused_bf16wire(const float* grads, void* const* wire, void* const* params, void* const* shadows,
                                 void* const* pads, void* epochs, float* m, float* v, long long n, int rank, int W,
                                 int algo, int blocks, int opt_kind, int step, float lr, float b1, float b2,
                                 float eps, float wd, float momentum, const int* step_dev, void* mc_wire,
                                 void* mc_params, void* mc_shadow, void* stream);
/* in-place average of W symmetric bf16 buffers (n % (8*W) == 0): fp32 accumulation in rank order, one rounding */
int mpb_allreduce_bf16(void* const* bufs, void* const* pads, void* epochs, long long n, int rank, int W, int blocks,
                       void* stream);
int mpb_broadcast(void* const* bufs, void* const* pads, void* epochs, long long n_floats, int rank, int W, int root,
                  int blocks, void* stream);
int mpb_allgather_slices(void* const* bufs, void* const* pads, void* epochs, long long n_floats, int rank, int W,
                         int blocks, void* stream);
int mpb_count_allreduce(void* const* vals, void* const* pads, void* epochs, long long* out, int rank, int W,
                        void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MPB200_H_ */
