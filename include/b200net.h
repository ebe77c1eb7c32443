/* b200net C-ABI: the drop-in boundary for the ResNet-50 data-parallel hot path on B200 (sm_100a).
 *
 * The reference (VaibhavKamat/Performance-Comparison-of-TensorFlow-PyTorch-and-their-Distributed-Counterparts)
 * has no FFI of its own: its scripts program against the PyTorch object protocol and the arithmetic lives in
 * ATen/cuDNN/oneDNN.  Each entry point below therefore cites the reference statement whose work it carries
 * (file:line in /root/reference; .ipynb lines are raw JSON lines) and the ATen operator it stands in for.
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer is a DEVICE pointer unless it says "host";
 *  - activations are NHWC ("[rows = N*H*W][C]"), dtype B200_F32 or B200_BF16; statistics, gradients of
 *    parameters and optimizer state are always fp32; parameters cross the boundary as OIHW fp32;
 *  - every call returns a status (0 = OK), never throws, never exits; b200_last_error() gives the message of
 *    the calling thread's last failure;
 *  - `stream` is a cudaStream_t passed as void*; work is enqueued, not synchronised;
 *  - the library never allocates device memory: scratch is caller-provided (`ws`, `ws_bytes`; the matching
 *    *_workspace_bytes query gives the size) and nothing is retained after the call returns;
 *  - there is no CPU / cuDNN / cuBLAS fallback: an unsupported shape is a hard error.
 */
#ifndef B200NET_H_
#define B200NET_H_
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { B200_OK = 0, B200_INVALID_ARG = 1, B200_UNSUPPORTED = 2, B200_CUDA_ERROR = 3, B200_NCCL_ERROR = 4,
       B200_WORKSPACE_TOO_SMALL = 5 };
enum { B200_F32 = 0, B200_BF16 = 1 };

int32_t b200_version(void);
const char* b200_last_error(void);
/* measurement aids (bench.py): kernels launched by this library so far; per-launch CUDA-event timing of the
 * tcgen05 GEMM kernel (enable -> run -> read sums the launches recorded since enable). */
int64_t b200_launch_count(void);
int32_t b200_gemm_timing_enable(int32_t on);
int32_t b200_gemm_timing_read(double* total_ms, int64_t* launches, double* flops);

/* ------------------------------------------------------------------ layout / packing
 * inputs.to(device) hands the model NCHW fp32 (pytorch_training_inference_on_image.ipynb:667); the first kernel
 * converts to the internal NHWC layout (and to bf16 in bf16 mode). */
/* Cpad >= C: channels of the NHWC output (extra channels are written as zero; used to widen the 3-channel
 * image to 8-byte pixels for the stem convolution). */
int32_t b200_nchw_to_nhwc(const float* x_nchw, void* y_nhwc, int32_t N, int32_t C, int32_t H, int32_t W,
                          int32_t Cpad, int32_t out_dtype, void* stream);
int32_t b200_nhwc_to_nchw(const void* x_nhwc, float* y_nchw, int32_t N, int32_t C, int32_t H, int32_t W,
                          int32_t in_dtype, void* stream);
/* OIHW fp32 parameter -> [K][ld] (dtype) with column (r*S+s)*C+c: the K-major operand of the implicit GEMM.
 * ld >= R*S*C is the row stride, zero-filled past R*S*C (b200_conv_weight_ld gives the value the conv entry
 * points expect: R*S*C rounded up to a multiple of 8). */
int32_t b200_conv_weight_ld(int32_t R, int32_t S, int32_t C);
/* Cw = channels of the parameter, C >= Cw = channels of the packed operand (zero-filled beyond Cw). */
int32_t b200_pack_weight(const float* w_oihw, void* w_krsc, int32_t K, int32_t Cw, int32_t R, int32_t S,
                         int32_t C, int32_t ld, int32_t out_dtype, void* stream);
/* b200_pack_weight for `count` parameters in one launch per 48 (host arrays of device pointers and of the
 * per-tensor K, Cw, C, R*S, ld): the training step re-packs every convolution weight after each optimizer update. */
int32_t b200_pack_weight_multi(int32_t count, const void* const* w_oihw, void* const* w_krsc, const int32_t* K,
                               const int32_t* Cw, const int32_t* C, const int32_t* RS, const int32_t* ld,
                               int32_t out_dtype, void* stream);
/* Explicit im2col staging: col[m][(r*S+s)*C + c] for output pixel m = (n*Ho + p)*Wo + q, row stride ld >= R*S*C
 * (tail zero-filled).  Used for the 3-channel stem: the image is widened to 4 channels, staged once per step with
 * ld = 256, and the stem then runs as a 1x1 convolution over 256 "channels" on the implicit tcgen05 path (forward
 * and weight gradient share the staged matrix). */
int32_t b200_im2col(const void* x, void* col, int32_t N, int32_t H, int32_t W, int32_t C, int32_t R, int32_t S,
                    int32_t stride, int32_t pad, int32_t ld, int32_t dtype, void* stream);
/* Row-pair window staging for the stem (Conv2d(3, 64, kernel_size=7, stride=2, padding=3): the first line of the
 * stored repr, pytorch_training_inference_on_image.ipynb:455), bf16 mode.  x is the channel-padded image
 * [N][H][W][4] bf16 (b200_nchw_to_nhwc with Cpad = 4); out is [N][Ho + R' - 1][Wo][64] bf16 with R' = (R+1)/2:
 * out[n][j][q][r2*32 + s*4 + c] = x[n][2j + r2 - pad][2q + s - pad][c] (zero outside the image), s = 0..7.
 * The stem then runs as an R' x 1 stride-1 convolution over 64 channels on the implicit tcgen05 path (desc
 * {N, Ho+R'-1, Wo, 64, K, R', 1, 1, 0}) with the weights packed by b200_stem_pack_weight into [K][R'*64]
 * (column j*64 + r2*32 + s*4 + c = w[k][c][2j+r2][s]); forward and weight gradient share the staged tensor,
 * which is a quarter of a full im2col matrix.  Stride 2, S <= 8, at most 4 channels. */
int32_t b200_stem_stage(const void* x_nhwc4, void* out, int32_t N, int32_t H, int32_t W, int32_t R, int32_t S,
                        int32_t pad, void* stream);
/* The same staged tensor straight from the loaders' fp32 NCHW batch (Cw <= 4 channels): layout conversion and staging
 * in one pass (pytorch_training_inference_on_image.ipynb:667 hands the model NCHW fp32). */
int32_t b200_stem_stage_nchw(const float* x_nchw, void* out, int32_t N, int32_t Cw, int32_t H, int32_t W, int32_t R,
                             int32_t S, int32_t pad, void* stream);
int32_t b200_stem_pack_weight(const float* w_oihw, void* w_packed, int32_t K, int32_t Cw, int32_t R, int32_t S,
                              void* stream);
/* Input pipeline: the reference's per-image transform
 * transforms.Compose([transforms.Resize((224,224)), transforms.ToTensor()])
 * (pytorch_training_inference_on_image.ipynb:325-326, :847-848; another_neural_net.py:38-46) for a batch of DECODED
 * images.  image_ptrs / heights / widths / row_strides are DEVICE arrays of N entries: uint8 HWC (RGB) pixel
 * pointers, sizes and row strides in bytes -- images of one batch may differ in size.  Bilinear resize to Ho x Wo
 * with PIL's antialiasing (antialias != 0; 0 = plain bilinear, torch's tensor default), rounded to a byte as PIL
 * does, then x/255 into out_layout 0: [N][Ho][Wo][4] bf16 (the stem's 8-byte pixels; the model accepts this tensor
 * directly) or 1: [N][3][Ho][Wo] fp32 (what ToTensor() returns).  JPEG entropy decoding stays on the host. */
int32_t b200_preprocess_u8(const void* const* image_ptrs, const int32_t* heights, const int32_t* widths,
                           const int64_t* row_strides, int32_t N, int32_t Ho, int32_t Wo, void* out,
                           int32_t out_layout, int32_t antialias, void* stream);
/* dtype cast of a flat buffer (fp32 <-> bf16) */
int32_t b200_cast(const void* x, void* y, int64_t n, int32_t in_dtype, int32_t out_dtype, void* stream);

/* ------------------------------------------------------------------ convolution  (Conv2d.forward / backward;
 * 53 instances, repr at pytorch_training_inference_on_image.ipynb:455-626; called through model.forward(inputs)
 * :669 and loss.backward() :671; replaces aten::cudnn_convolution / convolution_backward)
 * Implicit GEMM: fprop M=N*Ho*Wo, N=K(out channels), K=R*S*C. */
typedef struct {
  int32_t N, H, W, C;   /* input  NHWC */
  int32_t K, R, S;      /* filter: K output channels, RxS */
  int32_t stride, pad;
  int32_t dtype;        /* B200_F32 (exact FFMA path) or B200_BF16 (tcgen05 path, fp32 accumulate) */
} b200_conv_desc;

enum { B200_CONV_FPROP = 0, B200_CONV_DGRAD = 1, B200_CONV_WGRAD = 2 };
int64_t b200_conv_workspace_bytes(const b200_conv_desc* d, int32_t which);
/* y[N,Ho,Wo,K] = conv(x[N,H,W,C], w[K,R,S,C]) */
int32_t b200_conv_fprop(const b200_conv_desc* d, const void* x, const void* w_krsc, void* y, void* ws,
                        int64_t ws_bytes, void* stream);
/* Forward convolution that also emits the batch statistics the following BatchNorm2d needs (every conv of the
 * model is followed by one: repr at pytorch_training_inference_on_image.ipynb:455-626): `partials` receives
 * [rows][2][K] fp32 = per pixel-tile column sums and sums of squares of the bf16-rounded output, one row per pixel
 * tile; b200_conv_fprop_bnstats_rows gives the row count (0 = this shape is not on the fused path: use
 * b200_conv_fprop + b200_bn_train_stats).  b200_bn_finalize_stats turns the partials into mean / invstd /
 * scale / shift and updates the running statistics. */
int64_t b200_conv_fprop_bnstats_rows(const b200_conv_desc* d);
int32_t b200_conv_fprop_bnstats(const b200_conv_desc* d, const void* x, const void* w_krsc, void* y,
                                float* partials, int64_t partial_rows, void* stream);
/* Forward convolution with an eval-mode BatchNorm2d (+ReLU) folded into the epilogue (model.eval() paths:
 * pytorch_training_inference_on_image.ipynb:678-690, :853-861): y = act(conv(x,w) * scale[k] + shift[k] (+ residual)),
 * scale/shift from b200_bn_eval_coeffs, residual (optional, may be NULL) = the block's identity / downsample branch,
 * same shape as y.  Same shape support as b200_conv_fprop_bnstats (rows query > 0). */
int32_t b200_conv_fprop_affine(const b200_conv_desc* d, const void* x, const void* w_krsc, void* y,
                               const float* scale, const float* shift, const void* residual, int32_t relu,
                               void* stream);
/* dx[N,H,W,C] (+)= conv_transpose(dy[N,Ho,Wo,K], w[K,R,S,C]);  accumulate!=0 adds into dx */
int32_t b200_conv_dgrad(const b200_conv_desc* d, const void* dy, const void* w_krsc, void* dx, int32_t accumulate,
                        void* ws, int64_t ws_bytes, void* stream);
/* dw_oihw[K,C,R,S] fp32 = sum_{n,p,q} dy * x   (deterministic split-K; written in the parameter's own layout) */
int32_t b200_conv_wgrad(const b200_conv_desc* d, const void* x, const void* dy, float* dw_oihw, void* ws,
                        int64_t ws_bytes, void* stream);
/* Deferred split-K reductions of the weight gradient (no reference counterpart: launch-count optimisation of the
 * training step, which otherwise runs 53 tiny reduce kernels).  After b200_wgrad_defer(1), on the same host thread,
 * b200_conv_wgrad leaves its slabs in the workspace it was given -- each call needs its OWN workspace, alive until the
 * flush -- and b200_wgrad_flush(stream) writes all the queued dw tensors in one launch per 48 layers.  Layers whose
 * kernel writes dw directly are unaffected.  b200_wgrad_defer(0) leaves the mode (and drops what is still queued). */
int32_t b200_wgrad_defer(int32_t on);
int32_t b200_wgrad_flush(void* stream);

/* ------------------------------------------------------------------ GEMM (Linear layers of the head,
 * pytorch_training_inference_on_image.ipynb:439-443; replaces aten::addmm / mm)
 * C[M,N] (+)= op(A) * op(B) (+ bias[N]) (ReLU).  a_mn==0: A is [M][lda] (K contiguous); a_mn!=0: A is [K][lda]
 * (M contiguous).  b_mn==0: B is [N][ldb] (K contiguous); b_mn!=0: B is [K][ldb] (N contiguous). */
int32_t b200_gemm_f32(const float* A, const float* B, float* C, int32_t M, int32_t N, int32_t K, int64_t lda,
                      int64_t ldb, int64_t ldc, int32_t a_mn, int32_t b_mn, const float* bias, int32_t relu,
                      int32_t accumulate, void* stream);
/* bf16 operands on tcgen05 tensor cores, fp32 accumulate; out_dtype selects C's type.  splits>1 needs
 * ws >= splits*M*ldc*4 bytes and writes the reduced result to C. */
int32_t b200_gemm_bf16(const void* A, const void* B, void* C, int32_t M, int32_t N, int32_t K, int64_t lda,
                       int64_t ldb, int64_t ldc, int32_t a_mn, int32_t b_mn, int32_t out_dtype, int32_t accumulate,
                       int32_t splits, void* ws, int64_t ws_bytes, void* stream);

/* ------------------------------------------------------------------ BatchNorm2d (+ReLU, +residual)
 * (53 instances, eps 1e-5, momentum 0.1: pytorch_training_inference_on_image.ipynb:456; train mode in the
 * training loop :664-673 and in Standalone_Inference_Imagenette_trial.ipynb:180, eval mode :678,853;
 * replaces aten::cudnn_batch_norm / native_batch_norm(_backward), relu_, add_)
 * x is [rows][C].  All per-channel vectors are fp32 [C]. */
int64_t b200_bn_workspace_bytes(int64_t rows, int32_t C);
/* batch statistics -> mean, invstd; scale = gamma*invstd, shift = beta - mean*scale;
 * running_mean/var EMA update (unbiased variance) when running_mean != NULL. */
int32_t b200_bn_train_stats(const void* x, int64_t rows, int32_t C, int32_t dtype, const float* gamma,
                            const float* beta, float* running_mean, float* running_var, float momentum, float eps,
                            float* save_mean, float* save_invstd, float* scale, float* shift, void* ws,
                            int64_t ws_bytes, void* stream);
/* second half of b200_bn_train_stats for partial sums produced elsewhere (b200_conv_fprop_bnstats):
 * partials is [partial_rows][2][C]; rows = N*H*W of the normalised tensor. */
int32_t b200_bn_finalize_stats(const float* partials, int64_t partial_rows, int64_t rows, int32_t C,
                               const float* gamma, const float* beta, float* running_mean, float* running_var,
                               float momentum, float eps, float* save_mean, float* save_invstd, float* scale,
                               float* shift, void* stream);
/* Same computation and outputs as b200_bn_train_stats, for callers that have no workspace to give and must not
 * allocate one.  (No counterpart in the reference: it exists for hosts of this library that cannot call
 * cudaMalloc -- a serving process whose framework claimed all device memory for its own pool at start-up, or a
 * latency-bound loop that cannot afford cudaMalloc's device-wide synchronisation.)  The per-block partial sums
 * live in a 2 MiB array that is part of the library image itself (every GPU that loads the library holds those
 * 2 MiB until the process exits); the call makes no cudaMalloc / cudaFree / allocating library call.
 *   limit  : C <= B200_BN_NOALLOC_MAX_C; anything larger is rejected with B200_UNSUPPORTED and a message.
 *   set-up : the first call on each device looks up the array's device address (cudaGetSymbolAddress), with that
 *            device current; make that first call before starting any stream capture.
 *   sharing: every call on a device uses the same array, which keeps whatever the previous call left in it
 *            (nothing in it has to start at zero: each launch overwrites every slot it later reads).  Two calls
 *            that overlap on the GPU (different streams, no ordering between them) race on it: serialise them. */
#define B200_BN_NOALLOC_MAX_C 4096
int32_t b200_bn_train_stats_noalloc(const void* x, int64_t rows, int32_t C, int32_t dtype, const float* gamma,
                                    const float* beta, float* running_mean, float* running_var, float momentum,
                                    float eps, float* save_mean, float* save_invstd, float* scale, float* shift,
                                    void* stream);
/* eval mode: scale/shift from running statistics */
int32_t b200_bn_eval_coeffs(const float* gamma, const float* beta, const float* running_mean,
                            const float* running_var, float eps, float* scale, float* shift, int32_t C, void* stream);
/* y = act(x*scale + shift (+ residual)) ; act = ReLU when relu != 0.
 * mask (optional, may be NULL): receives the ReLU mask of the STORED output, one byte per 16-byte vector of y
 * (bit e = element e of the vector is > 0; b200_bn_mask_bytes gives the size, 0 = not available for this C). */
int64_t b200_bn_mask_bytes(int64_t rows, int32_t C, int32_t dtype);
/* reverse != 0: the kernel walks the tensor from its last element to its first.  A bandwidth kernel that starts
 * where its producer finished finds that end of the tensor still in the 126 MB L2 (the result does not depend on
 * the direction); callers alternate the direction along a producer -> consumer chain. */
int32_t b200_bn_apply(const void* x, const void* residual, void* y, uint8_t* mask, int64_t rows, int32_t C,
                      int32_t dtype, const float* scale, const float* shift, int32_t relu, int32_t reverse,
                      void* stream);
/* Backward of y = act(bn(x) (+res)).  g = dz * [y > 0] when relu, else g = dz.  The sign of y comes from `mask`
 * (as written by b200_bn_apply) when it is given, otherwise from z = the saved output.
 *   dgamma = sum g*xhat, dbeta = sum g, dx = gamma*invstd*(g - dbeta/n - xhat*dgamma/n);
 *   dres (optional, may be NULL) receives g (gradient of the residual branch).
 *   walk: bit 0 = the reduction pass, bit 1 = the dx pass walks the tensors from the last row to the first (see
 *   b200_bn_apply; results are independent of it). */
int32_t b200_bn_backward(const void* dz, const void* z, const uint8_t* mask, const void* x, void* dx, void* dres,
                         int64_t rows, int32_t C, int32_t dtype, const float* gamma, const float* save_mean,
                         const float* save_invstd, int32_t relu, float* dgamma, float* dbeta, void* ws,
                         int64_t ws_bytes, int32_t walk, void* stream);
/* standalone ReLU backward / elementwise add used between blocks: y = a + b ; y = dz*[z>0] */
int32_t b200_add(const void* a, const void* b, void* y, int64_t n, int32_t dtype, void* stream);
int32_t b200_relu_bwd(const void* dz, const void* z, void* dx, int64_t n, int32_t dtype, void* stream);

/* ------------------------------------------------------------------ pooling
 * MaxPool2d(3, stride 2, pad 1) (:458) and AdaptiveAvgPool2d((1,1)) (:627). */
int32_t b200_maxpool3x3s2_fwd(const void* x, void* y, uint8_t* argmax, int32_t N, int32_t H, int32_t W, int32_t C,
                              int32_t dtype, void* stream);
int32_t b200_maxpool3x3s2_bwd(const void* dy, const uint8_t* argmax, void* dx, int32_t N, int32_t H, int32_t W,
                              int32_t C, int32_t dtype, void* stream);
/* The stem's BatchNorm -> ReLU -> MaxPool2d(3, 2, 1) in one pass each way (torchvision ResNet.forward bn1 / relu /
 * maxpool of models.resnet50, pytorch_training_inference_on_image.ipynb:389, repr :455-458): the ReLU output and the pooling
 * input gradient -- the largest tensors of the step -- are never written.
 *   fwd: pooled[N,Ho,Wo,C] = maxpool(relu(y * scale + shift)); argmax (one byte per pooled element) may be NULL;
 *        scale == shift == NULL pools y as it is.
 *   bwd: dy[N,H,W,C] = BatchNorm-backward(relu-backward(maxpool-backward(dpool))) with batch statistics mean / invstd;
 *        dgamma / dbeta may be NULL; ws as for b200_bn_backward (b200_bn_workspace_bytes(N*H*W, C)).
 * C % 8 == 0 (bf16) / C % 4 == 0 (fp32), 16-byte aligned tensors; B200_UNSUPPORTED otherwise.
 * b200_maxpool3x3s2_bwd_rows: the pooling backward alone with the same indexing (b200_maxpool3x3s2_bwd uses it). */
int32_t b200_bn_relu_maxpool_fwd(const void* y, const float* scale, const float* shift, void* pooled, uint8_t* argmax,
                                 int32_t N, int32_t H, int32_t W, int32_t C, int32_t dtype, void* stream);
int32_t b200_maxpool_bn_bwd(const void* dpool, const uint8_t* argmax, const void* pooled, const void* y, void* dy,
                            int32_t N, int32_t H, int32_t W, int32_t C, int32_t dtype, const float* gamma,
                            const float* mean, const float* invstd, float* dgamma, float* dbeta, void* ws,
                            int64_t ws_bytes, void* stream);
int32_t b200_maxpool3x3s2_bwd_rows(const void* dy, const uint8_t* argmax, void* dx, int32_t N, int32_t H, int32_t W,
                                   int32_t C, int32_t dtype, void* stream);
/* y[N,C] fp32 = mean over HW */
int32_t b200_avgpool_fwd(const void* x, float* y, int32_t N, int32_t HW, int32_t C, int32_t dtype, void* stream);
int32_t b200_avgpool_bwd(const float* dy, void* dx, int32_t N, int32_t HW, int32_t C, int32_t dtype, void* stream);

/* ------------------------------------------------------------------ head: ReLU / Dropout / LogSoftmax / NLLLoss
 * nn.Sequential(Linear, ReLU, Dropout(0.2), Linear, LogSoftmax(dim=1)) + nn.NLLLoss()
 * (pytorch_training_inference_on_image.ipynb:439-444). fp32 [rows][cols]. */
int32_t b200_relu_fwd_f32(const float* x, float* y, int64_t n, void* stream);
/* Philox4x32-10 keyed by (seed, offset); mask[i] in {0,1}; y = x*mask/(1-p) */
int32_t b200_dropout_fwd(const float* x, float* y, uint8_t* mask, int64_t n, float p, uint64_t seed,
                         uint64_t offset, void* stream);
int32_t b200_dropout_bwd(const float* dy, const uint8_t* mask, float* dx, int64_t n, float p, void* stream);
int32_t b200_log_softmax_fwd(const float* x, float* y, int32_t rows, int32_t cols, void* stream);
int32_t b200_log_softmax_bwd(const float* dy, const float* y, float* dx, int32_t rows, int32_t cols, void* stream);
/* loss[0] = -mean_n logp[n, label[n]];  dlogp = -dloss/N at the label */
int32_t b200_nll_loss_fwd(const float* logp, const int64_t* labels, float* loss, int32_t rows, int32_t cols,
                          void* stream);
int32_t b200_nll_loss_bwd(const float* dloss, const int64_t* labels, float* dlogp, int32_t rows, int32_t cols,
                          void* stream);
/* bias gradient: db[N] = sum over rows of dy[rows][N] */
int32_t b200_colsum_f32(const float* dy, float* db, int32_t rows, int32_t cols, void* stream);

/* ------------------------------------------------------------------ optimizers (multi-tensor, fp32)
 * optim.Adam(model.fc.parameters(), lr=0.003) (:445), SGD(lr=0.001) (resnet.py:24), AdamW(lr=2e-5, eps=1e-8)
 * + clip_grad_norm_(.., 1.0) (pytorch_on_language_distr.py:167-170,273).  Pointer tables are HOST arrays of
 * device pointers; they are passed to the kernels by value (no device allocation). */
int32_t b200_multi_tensor_sgd(int32_t n, float* const* params, const float* const* grads, float* const* momentum_bufs,
                              const int64_t* sizes, float lr, float momentum, float dampening, float weight_decay,
                              int32_t nesterov, int32_t first_step, float grad_scale, void* stream);
/* decoupled_wd!=0 -> AdamW.  step is the 1-based step count used for bias correction. */
int32_t b200_multi_tensor_adam(int32_t n, float* const* params, const float* const* grads, float* const* exp_avg,
                               float* const* exp_avg_sq, const int64_t* sizes, float lr, float beta1, float beta2,
                               float eps, float weight_decay, int32_t decoupled_wd, int64_t step, float grad_scale,
                               void* stream);
/* out[0] = sum of squares over all tensors (deterministic two-stage); ws >= b200_l2norm_workspace_bytes() */
int64_t b200_l2norm_workspace_bytes(void);
int32_t b200_multi_tensor_l2norm_sq(int32_t n, const float* const* tensors, const int64_t* sizes, float* out,
                                    void* ws, int64_t ws_bytes, void* stream);
/* t *= min(1, max_norm / (sqrt(norm_sq[0]) + 1e-6))   (clip_grad_norm_ semantics, scale read on device) */
int32_t b200_multi_tensor_clip_scale(int32_t n, float* const* tensors, const int64_t* sizes, const float* norm_sq,
                                     float max_norm, void* stream);
int32_t b200_multi_tensor_zero(int32_t n, float* const* tensors, const int64_t* sizes, void* stream);
/* gradient bucket pack/unpack for DistributedDataParallel (commented constructor at
 * pytorch_on_language_distr.py:220-221): flat[off_i ..] = src_i * scale ; dst_i = flat[off_i ..] * scale */
int32_t b200_bucket_pack(int32_t n, const float* const* srcs, const int64_t* sizes, float* flat, float scale,
                         void* stream);
int32_t b200_bucket_unpack(int32_t n, float* const* dsts, const int64_t* sizes, const float* flat, float scale,
                           void* stream);
/* dst_i = src_i * scale for n independent tensor pairs, 48 pairs per launch: how the captured backward graphs move
 * a whole segment's gradients into their (non-adjacent) bucket slices with a handful of launches */
int32_t b200_multi_tensor_copy(int32_t n, float* const* dsts, const float* const* srcs, const int64_t* sizes,
                               float scale, void* stream);

/* ------------------------------------------------------------------ BERT-base encoder (IMDB fine-tuning step)
 * model = BertForSequenceClassification.from_pretrained("bert-base-uncased", num_labels=2, ...)
 * (pytorch_on_language_distr.py:155-161); one step = model(b_input_ids, token_type_ids=None,
 * attention_mask=b_input_mask, labels=b_labels) (:258-261), loss.backward() (:268).  The Linear layers run on
 * b200_conv_fprop_affine / b200_conv_dgrad / b200_conv_wgrad as flat 1x1 problems (bias in the epilogue); these
 * entry points are everything else.  Activations are [tokens][hidden] in `dtype`; parameters and statistics fp32.
 * Dropout masks are never stored: forward and backward derive them from (seed, offset, flat element index) with
 * Philox4x32-10, so the same (seed, offset) must be passed to the matching backward call. */

/* Dropout under CUDA-graph replay: offsets passed by value are frozen into a captured graph.  While a device
 * pointer is registered here (per calling thread; NULL clears it), every dropout site launched from this thread
 * adds (*device_counter) << 40 to its offset when the kernel EXECUTES, so bumping the counter before a replay gives
 * the replayed step fresh masks (and forward / backward graphs of one step agree with each other). */
int32_t b200_dropout_counter(const uint64_t* device_counter);

/* BertEmbeddings: s = (word[ids] + type[type_ids]) + pos[0..S) ; y = dropout(LayerNorm(s)).  type_ids may be NULL
 * (token_type_ids=None at :259 = all zeros).  sum_out receives s (saved for the backward), *status is set to 1 if an
 * id is outside [0, vocab) (torch raises a device-side assert there). */
int32_t b200_embed_ln_fwd(const int64_t* ids, const int64_t* type_ids, const float* word, const float* pos,
                          const float* type, const float* gamma, const float* beta, float eps, int32_t N, int32_t S,
                          int32_t H, int32_t vocab, int32_t max_pos, float p_drop, uint64_t seed, uint64_t offset,
                          void* sum_out, void* y, float* mean, float* rstd, int32_t* status, int32_t dtype,
                          void* stream);
/* gradient of the three tables from ds = dL/ds: dword (fp32 atomics, table zeroed here), dpos, dtype_tab.
 * ws: b200_colsum_workspace_bytes(N*S, H). */
int32_t b200_embed_bwd(const int64_t* ids, const int64_t* type_ids, const void* dsum, int32_t N, int32_t S, int32_t H,
                       int32_t vocab, int32_t max_pos, float* dword, float* dpos, float* dtype_tab, void* ws,
                       int64_t ws_bytes, int32_t dtype, void* stream);
/* BertSelfOutput / BertOutput: s = dropout(x) + residual ; y = LayerNorm(s) (eps 1e-12).  sum_out receives s (may
 * alias x); residual may be NULL. */
int32_t b200_ln_fwd(const void* x, const void* residual, void* sum_out, void* y, float* mean, float* rstd,
                    const float* gamma, const float* beta, float eps, int64_t rows, int32_t H, float p_drop,
                    uint64_t seed, uint64_t offset, int32_t dtype, void* stream);
/* dy = dy_a (+ dy_b, may be NULL: the two consumers of a hidden state).  drop_after == 0: dres = dL/ds,
 * dx = dres * dropout mask (dx may be NULL when p_drop == 0: it then equals dres), dbias = column sums of dx (bias
 * gradient of the Linear that produced x; may be NULL).  drop_after != 0 (embeddings): dy is first multiplied by the
 * dropout mask, dres = dL/ds, dx / dbias unused. */
int64_t b200_ln_bwd_workspace_bytes(int64_t rows, int32_t H);
int32_t b200_ln_bwd(const void* dy_a, const void* dy_b, const void* sum, const float* mean, const float* rstd,
                    const float* gamma, int64_t rows, int32_t H, float p_drop, uint64_t seed, uint64_t offset,
                    int32_t drop_after, void* dres, void* dx, float* dgamma, float* dbeta, float* dbias, void* ws,
                    int64_t ws_bytes, int32_t dtype, void* stream);
/* BertIntermediate activation: erf-GELU (hidden_act = "gelu") and its derivative; n % (16 bytes / element) == 0 */
int32_t b200_gelu_fwd(const void* x, void* y, int64_t n, int32_t dtype, void* stream);
int32_t b200_gelu_bwd(const void* dy, const void* x, void* dx, int64_t n, int32_t dtype, void* stream);
/* gelu_bwd fused with the bias gradient of the Linear in front of it: dx = dy * gelu'(x) on [rows][cols] and
 * dbias[cols] = column sums of dx, one pass.  ws: b200_colsum_workspace_bytes(rows, cols); dx may alias dy. */
int32_t b200_gelu_bwd_bias(const void* dy, const void* x, void* dx, float* dbias, int64_t rows, int32_t cols,
                           int32_t dtype, void* ws, int64_t ws_bytes, void* stream);
/* out[cols] fp32 = column sums of x[rows][cols] (bias gradients; deterministic two-stage) */
int64_t b200_colsum_workspace_bytes(int64_t rows, int32_t cols);
int32_t b200_colsum(const void* x, float* out, int64_t rows, int32_t cols, int32_t dtype, void* ws, int64_t ws_bytes,
                    void* stream);
/* BertSelfAttention: qkv [N*S][3*heads*D] = (q | k | v) from the fused QKV Linear; mask [N][S] (attention_mask at
 * :260, 0 = padded key, may be NULL); probs [N][heads][S][S] = softmax(q k^T / sqrt(D)) before dropout (saved);
 * ctx [N*S][heads*D] = dropout(probs) v.  D must be 64, S <= 512 and a multiple of 4. */
int32_t b200_attention_fwd(const void* qkv, const int64_t* mask, void* probs, void* ctx, int32_t N, int32_t S,
                           int32_t heads, int32_t D, float p_drop, uint64_t seed, uint64_t offset, int32_t dtype,
                           void* stream);
/* dqkv from dctx; dscores is scratch of the size of probs */
int32_t b200_attention_bwd(const void* qkv, const void* probs, const void* dctx, void* dscores, void* dqkv, int32_t N,
                           int32_t S, int32_t heads, int32_t D, float p_drop, uint64_t seed, uint64_t offset,
                           int32_t dtype, void* stream);
/* Fused tensor-core attention (bf16 only; S in {128, 256}, D = 64: b200_attention_flash_supported says which
 * shapes qualify, others use the two entry points above).  Nothing of size S*S*2 bytes reaches HBM: forward saves
 * lse [N][heads][S] (log2 units) and, when p_drop > 0, the dropout keep mask as bits keep_bits [N][heads][S][S/32]
 * (same Philox indices as above; may be NULL when p_drop == 0); backward recomputes the probabilities.
 * dvec: scratch [N][heads][S]. */
int32_t b200_attention_flash_supported(int32_t S, int32_t D);
int32_t b200_attention_flash_fwd(const void* qkv, const int64_t* mask, float* lse, uint32_t* keep_bits, void* ctx,
                                 int32_t N, int32_t S, int32_t heads, int32_t D, float p_drop, uint64_t seed,
                                 uint64_t offset, void* stream);
int32_t b200_attention_flash_bwd(const void* qkv, const int64_t* mask, const float* lse, const uint32_t* keep_bits,
                                 const void* ctx, const void* dctx, float* dvec, void* dqkv, int32_t N, int32_t S,
                                 int32_t heads, int32_t D, float p_drop, void* stream);
/* BertPooler: tanh on fp32 [rows][H]; first-token gather (out[n] = x[n*row_stride ..], to fp32) and its scatter */
int32_t b200_tanh_fwd_f32(const float* x, float* y, int64_t n, void* stream);
int32_t b200_tanh_bwd_f32(const float* dy, const float* y, float* dx, int64_t n, void* stream);
int32_t b200_gather_rows(const void* x, float* out, int32_t N, int64_t row_stride, int32_t H, int32_t dtype,
                         void* stream);
int32_t b200_scatter_rows(const float* d, void* out, int32_t N, int64_t row_stride, int32_t H, int32_t dtype,
                          void* stream);

/* ------------------------------------------------------------------ VGG16 path
 * models.vgg16(pretrained=True) + classifier[6] head (pytorch_training_inference_on_image.ipynb:1968-2041,
 * another_neural_net.py:244-258).  The 13 conv3x3+bias+ReLU layers are b200_conv_fprop_affine (scale = 1,
 * shift = bias, relu = 1) / b200_conv_dgrad / b200_conv_wgrad, the classifier's Linears the same kernel as flat
 * 1x1 problems; these are the remaining layers.  NHWC, `dtype` storage. */
/* MaxPool2d(kernel_size=2, stride=2) (floor mode): y [N,H/2,W/2,C]; argmax (optional) = window position 0..3 */
int32_t b200_maxpool2x2_fwd(const void* x, void* y, uint8_t* argmax, int32_t N, int32_t H, int32_t W, int32_t C,
                            int32_t dtype, void* stream);
int32_t b200_maxpool2x2_bwd(const void* dy, const uint8_t* argmax, void* dx, int32_t N, int32_t H, int32_t W,
                            int32_t C, int32_t dtype, void* stream);
/* AdaptiveAvgPool2d((OH, OW)): windows [floor(i*H/OH), ceil((i+1)*H/OH)) as in ATen */
int32_t b200_adaptive_avgpool_fwd(const void* x, void* y, int32_t N, int32_t H, int32_t W, int32_t C, int32_t OH,
                                  int32_t OW, int32_t dtype, void* stream);
int32_t b200_adaptive_avgpool_bwd(const void* dy, void* dx, int32_t N, int32_t H, int32_t W, int32_t C, int32_t OH,
                                  int32_t OW, int32_t dtype, void* stream);
/* Dropout on an activation buffer: y = x * keep / (1 - p), mask from Philox (seed, offset, index); applying the
 * same call to the output gradient is the backward.  n % 8 == 0. */
int32_t b200_dropout_act(const void* x, void* y, int64_t n, float p_drop, uint64_t seed, uint64_t offset,
                         int32_t dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200NET_H_ */
