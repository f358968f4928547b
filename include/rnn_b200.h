/*
 * rnn_b200.h -- C ABI of librnn_b200.so, the sm_100a (B200) kernels behind replicann_b200.
 *
 * The reference (aaakulchyk/replicann) has no FFI of its own: every FLOP it performs is a call into
 * torch (`nn.Linear`, `@`, `F.softmax`, `nn.LayerNorm`, `nn.Dropout`, `F.relu`, autograd).  Each entry
 * point below therefore cites the reference call site (file:line under /root/reference/src/replicann)
 * whose library call it replaces, or -- for the CNN training path BASELINE.json's north_star names but the
 * reference does not contain -- the torch.nn operator whose semantics it follows.
 *
 * Conventions
 *   - All pointers are caller-owned DEVICE memory unless stated; nothing here allocates or frees device
 *     memory and nothing synchronises the device.  Work is enqueued on `stream`.
 *   - Return value: RNN_OK (0) or a negative rnn_status.  rnn_last_error() returns a static,
 *     thread-local description of the most recent failure.
 *   - Activations are NHWC (torch channels_last), filters are KRSC (channels_last [K,C,R,S]).
 *   - "bf16" tensors are __nv_bfloat16; statistics, optimizer state and reductions are fp32.
 *   - Kernels that need scratch take (ws, ws_bytes) and have a matching *_workspace() size query.
 */
#ifndef RNN_B200_H_
#define RNN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

typedef enum {
  RNN_OK = 0,
  RNN_ERR_SHAPE = -1,     /* unsupported or inconsistent shape                              */
  RNN_ERR_ALIGN = -2,     /* pointer not 16-byte aligned                                    */
  RNN_ERR_WORKSPACE = -3, /* workspace missing or too small                                 */
  RNN_ERR_CUDA = -4,      /* CUDA runtime error (see rnn_last_error)                        */
  RNN_ERR_TENSORMAP = -5, /* cuTensorMapEncode* rejected the descriptor                     */
  RNN_ERR_DTYPE = -6      /* unsupported dtype / mask type (reference: TypeError, attention.py:43-44) */
} rnn_status;

/* ---- runtime ------------------------------------------------------------------------------ */
const char* rnn_last_error(void);
int rnn_version(void);
int rnn_sm_count(void); /* SMs of the current device (148 on B200) */
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
unsigned long long rnn_launch_count(void);
/* Measurement aid (bench.py): with timing enabled, the tensor-core entry points record a CUDA-event pair around each
 * of their kernel launches, on the stream of the launch, tagged with a family name ("tc_kernel", "splitk_reduce",
 * "bn_finalize").  rnn_timing_enable(on) clears what was recorded and switches recording on / off (the one place
 * in the library that creates CUDA events; never during stream capture); rnn_timing_read() waits for the recorded
 * events of one family and returns their summed device time and count.  No counterpart in the reference. */
int rnn_timing_enable(int on);
int rnn_timing_read(const char* family, float* ms, int* launches);
/* Dropout under CUDA graphs.  Every Philox-drawing kernel (rnn_dropout_fwd, the attention families) adds a
 * device-side epoch to its `offset` argument; the epoch is 0 until this call, which enqueues kernels that advance
 * it by 2^32.  Put one call at the head of a captured training step: each replay then draws fresh masks although
 * the seed / offset arguments are frozen in the graph; forward and backward of one step see the same epoch.
 * No counterpart in the reference (torch keeps its Philox offset in device memory under graphs for the same reason). */
int rnn_rng_advance(cudaStream_t stream);

/* ---- tensor-core contractions (tcgen05 + TMEM + TMA) -------------------------------------- */

/* D[M,N] = A.B^T (+bias[N]) (ReLU).  a_mn=0: A is [M,K] row-major; a_mn=1: A is stored [K,M].
 * b_mn=0: B is [N,K] row-major (nn.Linear.weight); b_mn=1: B is stored [K,N].
 * Replaces nn.Linear forward at nn/attention.py:25,85,131-132 and arch/transformer.py:24-26,
 * and autograd's dgrad (a_mn=0,b_mn=1) / wgrad (a_mn=1,b_mn=1) of the same.
 * D is bf16, or fp32 when out_f32 != 0.  Small tile grids are split along K into `ws`. */
size_t rnn_gemm_bf16_workspace(int M, int N, int K, int out_f32, int has_epilogue);
int rnn_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int a_mn, int b_mn,
                  const float* bias, int relu, int out_f32, void* ws, size_t ws_bytes, cudaStream_t stream);

typedef struct {
  int N, H, W, C;       /* input  x[N,H,W,C]            */
  int K, R, S;          /* filter w[K,R,S,C]            */
  int stride_h, stride_w, pad_h, pad_w;
  int P, Q;             /* output y[N,P,Q,K]            */
} rnn_conv_desc;

/* torch.nn.Conv2d forward as an implicit GEMM (im2col-mode TMA): y = conv(x, w), then optionally
 * y = y*scale[k] + shift[k] (bias / folded BatchNorm), y += residual, y = relu(y).   C % 64 == 0. */
int rnn_conv2d_fprop_bf16(const rnn_conv_desc* d, const void* x, const void* w, void* y, const float* scale,
                          const float* shift, const void* residual, int relu, cudaStream_t stream);
/* Conv2d forward + torch.nn.BatchNorm2d (training) batch statistics of its output in one pass: the epilogue
 * accumulates per-channel sums of the bf16 values it stores; mean / invstd / scale / shift and the running
 * statistics come out as from rnn_bn_stats_bf16.  K must be a multiple of its N tile (64 / 128 / 256). */
size_t rnn_conv2d_fprop_bn_stats_workspace(const rnn_conv_desc* d);
int rnn_conv2d_fprop_bn_stats_bf16(const rnn_conv_desc* d, const void* x, const void* w, void* y, const float* gamma,
                                   const float* beta, float eps, float momentum, float* running_mean,
                                   float* running_var, float* mean, float* invstd, float* scale, float* shift,
                                   void* ws, size_t ws_bytes, cudaStream_t stream);
/* autograd of Conv2d w.r.t. its input.  accumulate != 0: dx += result. */
int rnn_conv2d_dgrad_bf16(const rnn_conv_desc* d, const void* dy, const void* w, void* dx, int accumulate,
                          cudaStream_t stream);
/* F.relu(Conv2d(x)) with the ReLU's backward moved into the CONSUMER's dgrad (a chain conv -> relu -> conv, as in
 * VGG / AlexNet; torch.nn.Conv2d + F.relu semantics, no counterpart kernel in the reference):
 *   fprop_relu_mask: y = relu(conv(x, w) + shift) and mask[N,P,Q,K/8] (bit k&7 of byte k>>3 = y > 0); K % 64 == 0.
 *   dgrad_masked   : dx = conv_transpose(dy, w) where the mask bit of that element is set, else 0 (mask: [N,H,W,C/8]
 *                    of the tensor this convolution consumed); unit stride, C % 64 == 0.
 * The producer's backward then needs no pass over its output (its incoming gradient is already masked). */
int rnn_conv2d_fprop_relu_mask_bf16(const rnn_conv_desc* d, const void* x, const void* w, void* y, const float* shift,
                                    unsigned char* mask_out, cudaStream_t stream);
int rnn_conv2d_dgrad_masked_bf16(const rnn_conv_desc* d, const void* dy, const void* w, const unsigned char* mask,
                                 void* dx, cudaStream_t stream);
/* autograd of Conv2d w.r.t. its weight; dw is bf16 [K,R,S,C] (fp32 when dw_f32 != 0). */
size_t rnn_conv2d_wgrad_workspace(const rnn_conv_desc* d);
int rnn_conv2d_wgrad_bf16(const rnn_conv_desc* d, const void* x, const void* dy, void* dw, int dw_f32,
                          int accumulate, void* ws, size_t ws_bytes, cudaStream_t stream);

/* First-layer convolution for inputs with C <= 8 (RGB stems): eight horizontally adjacent pixels x 8 padded
 * channels form one 64-wide K step ("packed pixel" view through an overlapping-stride im2col tensor map), so a
 * 7x7 stem costs 7 K steps instead of 49.  `d` describes the REAL problem (C = 3 ...).  x_packed is
 * [N, Hp, Wp, 8] bf16, spatially zero-padded (rnn_nchw_f32_to_nhwc_bf16_padded writes it; Hp/Wp/SG from
 * rnn_conv2d_first_geometry); w_packed / dw_packed are [K, R, SG, 8 taps, 8 ch] with zeros beyond S and C.
 * Same torch.nn.Conv2d semantics as above; there is no dgrad (the network input needs no gradient). */
int rnn_conv2d_first_geometry(const rnn_conv_desc* d, int* Hp, int* Wp, int* SG);
int rnn_conv2d_first_fprop_bf16(const rnn_conv_desc* d, const void* x_packed, const void* w_packed, void* y,
                                const float* scale, const float* shift, int relu, cudaStream_t stream);
/* first layer + 1-bit ReLU mask, as rnn_conv2d_fprop_relu_mask_bf16 (K % 64 == 0) */
int rnn_conv2d_first_fprop_relu_mask_bf16(const rnn_conv_desc* d, const void* x_packed, const void* w_packed, void* y,
                                          const float* shift, unsigned char* mask_out, cudaStream_t stream);
size_t rnn_conv2d_first_wgrad_workspace(const rnn_conv_desc* d);
int rnn_conv2d_first_wgrad_bf16(const rnn_conv_desc* d, const void* x_packed, const void* dy, void* dw_packed,
                                int dw_f32, void* ws, size_t ws_bytes, cudaStream_t stream);

/* ---- fp32 CUDA-core path (LeNet-5 class networks; BASELINE.json configs[0]) -------------------- */
/* torch.nn.Conv2d (fp32, NHWC / KRSC) forward incl. bias, and autograd dgrad / wgrad. */
int rnn_conv2d_fprop_f32(const rnn_conv_desc* d, const float* x, const float* w, const float* bias, float* y,
                         cudaStream_t stream);
int rnn_conv2d_dgrad_f32(const rnn_conv_desc* d, const float* dy, const float* w, float* dx, cudaStream_t stream);
int rnn_conv2d_wgrad_f32(const rnn_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate,
                         cudaStream_t stream);
/* fp32 counterpart of rnn_gemm_bf16 (same a_mn / b_mn meaning); nn.Linear in fp32 models. */
int rnn_gemm_f32(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn,
                 const float* bias, int relu, int accumulate, cudaStream_t stream);
/* ReLU backward fused with the bias gradient that follows it (autograd of relu + bias add: transformer.py:24-25,
 * CNN conv+bias+ReLU layers): dx = dy * (y > 0), dbias[c] (+)= sum_r dx[r, c].  C % 8 == 0 (bf16) / % 4 (fp32);
 * ws: rnn_colsum_workspace(M, C). */
int rnn_relu_bwd_colsum(const void* dy, const void* y, void* dx, long long M, int C, void* dbias, int accumulate,
                        int is_f32, void* ws, size_t ws_bytes, cudaStream_t stream);
/* The same GEMM at fp32 accuracy on the bf16 tensor cores: both operands are split into three bf16 pieces
 * (24 mantissa bits) and the six significant piece products run as one tcgen05 GEMM over a 6x longer reduction.
 * No accumulate mode; N % 8 == 0, K % 4 == 0 (and M % 8 == 0 when a_mn) -- rnn_gemm_f32_tc_supported().
 * ws: rnn_gemm_f32_tc_workspace() bytes (the split operands and split-K partials). */
int rnn_gemm_f32_tc_supported(int M, int N, int K, int a_mn, int b_mn);
size_t rnn_gemm_f32_tc_workspace(int M, int N, int K, int has_epilogue);
int rnn_gemm_f32_tc(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn,
                    const float* bias, int relu, void* ws, size_t ws_bytes, cudaStream_t stream);
/* out[c] (+)= sum_r x[r,c] -- bias gradient of nn.Linear / nn.Conv2d (autograd of attention.py:85,
 * transformer.py:24-25). */
size_t rnn_colsum_workspace(long long M, int C);
int rnn_colsum(const void* x, long long M, int C, void* out, int accumulate, int is_f32, void* ws, size_t ws_bytes,
               cudaStream_t stream);

/* ---- elementwise (HBM-bound) -------------------------------------------------------------------- */
/* Input pipeline: NCHW fp32 -> NHWC bf16 with the channel dimension zero-padded to Cpad. */
int rnn_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int H, int W, int Cpad, cudaStream_t stream);
/* Same, into a spatially zero-padded [N, Hp, Wp, Cpad] buffer (image at rows pad_h.., columns pad_w..). */
int rnn_nchw_f32_to_nhwc_bf16_padded(const float* x, void* y, int N, int C, int H, int W, int Cpad, int pad_h,
                                     int pad_w, int Hp, int Wp, cudaStream_t stream);
/* The same two conversions with the image pipeline's per-channel normalisation folded in (SURVEY 8f.2:
 * "fp32 NCHW -> bf16 NHWC conversion + normalisation"; torchvision's Normalize(mean, std), which the reference's
 * notebooks would apply on the host): v = (x - mean[c]) * inv_std[c] before the bf16 rounding.  mean / inv_std are
 * device arrays of C floats, both NULL = no normalisation.  Padding (channels, borders) stays exactly zero. */
int rnn_nchw_f32_to_nhwc_bf16_norm(const float* x, void* y, int N, int C, int H, int W, int Cpad, const float* mean,
                                   const float* inv_std, cudaStream_t stream);
int rnn_nchw_f32_to_nhwc_bf16_padded_norm(const float* x, void* y, int N, int C, int H, int W, int Cpad, int pad_h,
                                          int pad_w, int Hp, int Wp, const float* mean, const float* inv_std,
                                          cudaStream_t stream);
int rnn_nhwc_bf16_to_nchw_f32(const void* x, float* y, int N, int C, int H, int W, int Cpad, cudaStream_t stream);
/* F.relu (arch/transformer.py:30) and its autograd: dx = dy * (y > 0). */
int rnn_relu_fwd(const void* x, void* y, long long n, int is_f32, cudaStream_t stream);
int rnn_relu_bwd(const void* dy, const void* y, void* dx, long long n, int is_f32, cudaStream_t stream);
/* tanh activation of the classic LeNet-5 (fp32 only). */
int rnn_tanh_fwd_f32(const float* x, float* y, long long n, cudaStream_t stream);
int rnn_tanh_bwd_f32(const float* dy, const float* y, float* dx, long long n, cudaStream_t stream);
/* residual add x + sublayer(x) (arch/transformer.py:111-112) when it is not fused into a norm kernel. */
int rnn_add_bf16(const void* a, const void* b, void* y, long long n, cudaStream_t stream);
int rnn_cast_f32_to_bf16(const float* x, void* y, long long n, cudaStream_t stream);
int rnn_cast_bf16_to_f32(const void* x, float* y, long long n, cudaStream_t stream);
/* nn.Dropout (nn/attention.py:26,46,86 ; arch/transformer.py:26,32): Philox counter RNG keyed by
 * (seed, offset); mask holds one keep-byte per element for the backward pass. */
int rnn_dropout_fwd(const void* x, void* y, uint8_t* mask, long long n, float p, unsigned long long seed,
                    unsigned long long offset, int is_f32, cudaStream_t stream);
int rnn_dropout_bwd(const void* dy, const uint8_t* mask, void* dx, long long n, float p, int is_f32,
                    cudaStream_t stream);

/* ---- normalisation ------------------------------------------------------------------------------- */
/* torch.nn.BatchNorm2d, training mode, over x[M = N*H*W, C] bf16: batch statistics (biased variance for
 * normalisation, unbiased for running_var), scale = gamma*invstd, shift = beta - mean*scale. */
size_t rnn_bn_workspace(long long M, int C);
int rnn_bn_stats_bf16(const void* x, long long M, int C, const float* gamma, const float* beta, float eps,
                      float momentum, float* running_mean, float* running_var, float* mean, float* invstd,
                      float* scale, float* shift, void* ws, size_t ws_bytes, cudaStream_t stream);
/* Variant of rnn_bn_stats_bf16 that takes no workspace and allocates nothing (no counterpart in the
 * reference, which has neither this operator nor any allocation contract).  It exists for hosts that can
 * hand over input and output buffers but have no device allocator to lend: a process whose HBM was carved
 * up in full by another framework's pool at start-up, or a thin foreign-function binding that can marshal
 * only the tensors themselves.  The partial sums live in an array compiled into the library
 * (RNN_BN_STATIC_SCRATCH_BYTES, resident on every GPU that loads the library until exit).
 * Limit: C % 8 == 0 and C <= RNN_BN_STATIC_MAX_C, else RNN_ERR_WORKSPACE.  The first call on a device also
 * resolves the array's address -- make it before any stream capture.  Calls share the array: calls whose
 * kernels may overlap on the GPU race with each other and must be ordered by the caller. */
#define RNN_BN_STATIC_SCRATCH_BYTES (1u << 20)
#define RNN_BN_STATIC_MAX_C 4096
int rnn_bn_stats_bf16_static(const void* x, long long M, int C, const float* gamma, const float* beta, float eps,
                             float momentum, float* running_mean, float* running_var, float* mean, float* invstd,
                             float* scale, float* shift, cudaStream_t stream);
int rnn_bn_eval_prepare(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                        float eps, int C, float* scale, float* shift, cudaStream_t stream);
/* y = relu(x*scale + shift + residual): fused BatchNorm-apply (+ residual add) (+ ReLU).  relu_mask (nullable,
 * M*C/8 bytes) receives one bit per element (y > 0) -- the compact ReLU mask the backward kernels can read
 * instead of y. */
int rnn_bn_apply_bf16(const void* x, const float* scale, const float* shift, const void* residual, int relu, void* y,
                      unsigned char* relu_mask, long long M, int C, cudaStream_t stream);
/* Backward of the fused op above.  The ReLU mask comes from y_out (the forward output), from the forward's
 * relu_mask bits, or -- when the forward had no residual -- is recomputed from x with the forward's
 * (fwd_scale, fwd_shift): pass at most one of the three.
 * dres (nullable) receives the gradient of the residual branch.  dgamma/dbeta are fp32 or bf16. */
int rnn_bn_bwd_bf16(const void* dy, const void* x, const void* y_out, const float* mean, const float* invstd,
                    const float* gamma, long long M, int C, void* dx, void* dres, void* dgamma, void* dbeta,
                    int grads_f32, int accumulate_grads, float* sums /* 5*C floats of scratch */,
                    const float* fwd_scale, const float* fwd_shift, const unsigned char* relu_mask, void* ws,
                    size_t ws_bytes, cudaStream_t stream);
/* nn.LayerNorm(E) over y = LN(x + residual) (arch/transformer.py:65,72,111-112,188); xsum (nullable)
 * receives x + residual for the backward pass; mean / rstd are per-row fp32. */
int rnn_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, float eps, void* y,
                      void* xsum, float* mean, float* rstd, long long M, int E, int is_f32, cudaStream_t stream);
size_t rnn_layernorm_bwd_workspace(long long M, int E);
int rnn_layernorm_bwd(const void* dy, const void* xin, const void* gamma, const float* mean, const float* rstd,
                      void* dx, void* dgamma, void* dbeta, int accumulate_grads, long long M, int E, int is_f32,
                      void* ws, size_t ws_bytes, cudaStream_t stream);

/* ---- pooling (NHWC) ------------------------------------------------------------------------------ */
/* torch.nn.MaxPool2d(return_indices=True): indices (nullable) are h*W+w inside the input plane (int32), first
 * maximum in window order, NaN propagates -- bit-exact with ATen.  window_idx (nullable) is the same argmax as
 * a 1-byte position inside the window (r*S+s), the compact form the backward pass reads; pass either to bwd. */
int rnn_maxpool2d_fwd(const void* x, void* y, int* indices, unsigned char* window_idx, int N, int H, int W, int C,
                      int R, int S, int sh, int sw, int ph, int pw, int is_f32, cudaStream_t stream);
/* relu(BatchNorm-apply(x)) followed by the max pool, in one pass: pools relu(x * scale + shift) (rounded to the
 * storage type first, so values and argmax ties equal rnn_bn_apply_bf16 + rnn_maxpool2d_fwd) without writing the
 * normalised tensor -- the stem of the ResNets the north star names (torch: F.max_pool2d(F.relu(bn(x)))). */
int rnn_bn_relu_maxpool2d_fwd(const void* x, const float* scale, const float* shift, void* y, int* indices,
                              unsigned char* window_idx, int N, int H, int W, int C, int R, int S, int sh, int sw, int ph,
                              int pw, int is_f32, cudaStream_t stream);
int rnn_maxpool2d_bwd(const void* dy, const int* indices, const unsigned char* window_idx, void* dx, int N, int H,
                      int W, int C, int R, int S, int sh, int sw, int ph, int pw, int is_f32, cudaStream_t stream);
/* max_pool2d backward through a ReLU in front of the pool (autograd of max_pool2d(relu(.)) w.r.t. the ReLU's input,
 * given the POOLED output y): a window's gradient lands on its argmax element, whose value is the pooled value, so
 * the ReLU mask is (y > 0) per window -- dx = scatter(dy * (y > 0)).  Same index arguments as rnn_maxpool2d_bwd. */
int rnn_maxpool2d_relu_bwd(const void* dy, const void* y, const int* indices, const unsigned char* window_idx, void* dx,
                           int N, int H, int W, int C, int R, int S, int sh, int sw, int ph, int pw, int is_f32,
                           cudaStream_t stream);
/* torch.nn.AvgPool2d (count_include_pad=True, ceil_mode=False). */
int rnn_avgpool2d_fwd(const void* x, void* y, int N, int H, int W, int C, int R, int S, int sh, int sw, int ph, int pw,
                      int is_f32, cudaStream_t stream);
int rnn_avgpool2d_bwd(const void* dy, void* dx, int N, int H, int W, int C, int R, int S, int sh, int sw, int ph, int pw,
                      int is_f32, cudaStream_t stream);
/* torch.nn.AdaptiveAvgPool2d((1,1)). */
int rnn_global_avgpool_fwd(const void* x, void* y, int N, int HW, int C, int is_f32, cudaStream_t stream);
int rnn_global_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, int is_f32, cudaStream_t stream);

/* ---- softmax / loss ------------------------------------------------------------------------------ */
/* F.softmax(w, dim=-1) (nn/attention.py:45) and its autograd. */
int rnn_softmax_fwd(const void* x, void* y, long long rows, int cols, int is_f32, cudaStream_t stream);
int rnn_softmax_bwd(const void* dy, const void* y, void* dx, long long rows, int cols, int is_f32,
                    cudaStream_t stream);
/* F.cross_entropy(logits, target) (mean) + dlogits = (softmax - onehot)*grad_scale in one pass. */
int rnn_softmax_ce(const void* logits, const long long* target, long long rows, int cols, float grad_scale,
                   float* row_loss, float* loss, void* dlogits, int is_f32, cudaStream_t stream);
/* torch.argmax(x, dim=-1): first maximum wins. */
int rnn_argmax(const void* x, long long* out, long long rows, int cols, int is_f32, cudaStream_t stream);

/* ---- embeddings ------------------------------------------------------------------------------------------ */
/* torch.nn.Embedding lookup for a token table tok[V, E] fused with the learned positional table pos[T, E]
 * (nullable): out[n, :] = tok[ids[n], :] + pos[n % T, :] for n < n_tokens (= B * T).  Not in the reference: the
 * embedding stage of the encoder / decoder stacks SURVEY.md 8(f).4 names as the next replication; it feeds
 * TransformerEncoder / TransformerDecoder (arch/transformer.py:87-151).  An index outside [0, V) reads as zeros. */
int rnn_embedding_fwd(const long long* ids, const void* tok, const void* pos, void* out, long long n_tokens, int T,
                      int V, int E, int is_f32, cudaStream_t stream);
/* Its autograd: dtok[v, :] (+)= sum of dy rows whose index is v (ascending token order, no floating-point atomics:
 * bitwise reproducible), dpos[t, :] (+)= sum over the batch.  Either output may be NULL.  ws: V ints. */
size_t rnn_embedding_bwd_workspace(int V);
int rnn_embedding_bwd(const long long* ids, const void* dy, void* dtok, void* dpos, long long n_tokens, int T, int V,
                      int E, int is_f32, int accumulate_tok, int accumulate_pos, void* ws, size_t ws_bytes,
                      cudaStream_t stream);

/* ---- batched GEMM (attention contractions on the tensor cores) -------------------------------------- */
/* D_z = A_z . B_z^T, bf16 in, fp32 accumulate, for z = b * heads + h.  No counterpart in the reference: it is the
 * two torch.bmm-shaped `@` products of _AttentionHead._attention (nn/attention.py:37,47) and their autograd, for
 * all heads of a multi-head module at once.  Operand X in {a, b, d} is a 3-D tensor of row-major slabs
 * (x_ld elements between rows, x_bstride between slabs).  x_per_head = 0: problem z owns slab z.
 * x_per_head = 1: the heads of batch b share slab b and head h's columns start at h * x_hoff -- the layout of a
 * packed projection [B][T][H*hs] (x_hoff = hs; hs and the extents along a shared row must be multiples of 64).
 * a_mn / b_mn: the operand is stored [K][M] / [K][N] (transposed use of the same buffer, no copy).
 * out_f32: D is fp32, one dense [M][N] slab per problem. */
typedef struct {
  int M, N, K, batches, heads;
  int a_mn, b_mn, out_f32;
  long long a_ld, a_bstride; int a_per_head, a_hoff;
  long long b_ld, b_bstride; int b_per_head, b_hoff;
  long long d_ld, d_bstride; int d_per_head, d_hoff;
} rnn_bgemm_desc;
int rnn_gemm_batched_bf16(const rnn_bgemm_desc* d, const void* A, const void* B, void* D, cudaStream_t stream);

/* ---- fused attention ------------------------------------------------------------------------------- */
/* All heads of one multi-head module in one launch.  q/k/v/z/dz are addressed as
 * base + b*bs + row*rs + head*hs + d (elements): the packed projection output and the reference's
 * torch.cat(heads, -1) layout (nn/attention.py:157-160,180).  mask: [mask_b][T][Tk] through (mask_bs, mask_rs);
 * mask_kind 1 = bool/uint8, nonzero masks OUT (-inf, :39-40); 2 = float32 additive (:41-42). */
typedef struct {
  int B, H, T, Tk, hs;
  long long q_bs, q_rs, k_bs, k_rs, v_bs, v_rs, z_bs, z_rs, dz_bs, dz_rs;
  float scale;          /* 1/sqrt(in_dim): the EMBEDDING size (nn/attention.py:24) */
  int mask_kind;
  long long mask_bs, mask_rs;
  float p_drop;         /* dropout on the attention weights (:46); 0 in eval mode */
  unsigned long long seed, offset;
} rnn_attn_desc;
/* _AttentionHead._attention (nn/attention.py:28-50) for H heads: z = dropout(softmax(q k^T*scale (+mask))) v.
 * lse [B,H,T] (nullable in inference) is saved for the backward pass. */
int rnn_attention_fwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask, void* z,
                      float* lse, int is_f32, cudaStream_t stream);
/* autograd of the same: dq, dk, dv (same strides as q, k, v); delta is [B,H,T] fp32 scratch. */
int rnn_attention_bwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask,
                      const void* z, const void* dz, const float* lse, float* delta, void* dq, void* dk, void* dv,
                      int is_f32, cudaStream_t stream);

/* Tensor-core form of the same attention (bf16; head size a multiple of 64, T and Tk multiples of 8, strides
 * multiples of 8 elements -- rnn_attention_tc_supported() says whether a descriptor qualifies).  The contractions
 * (nn/attention.py:37,47 and their autograd) run as batched tcgen05 GEMMs over all heads, the softmax / mask /
 * dropout step (:38-46) as one row-wise kernel in between.  `saved` (rnn_attention_tc_saved_bytes) holds the
 * probabilities from forward to backward; `ws` (rnn_attention_tc_workspace) is scratch for either direction.
 * dq / dk / dv use the strides of q / k / v. */
int rnn_attention_tc_supported(const rnn_attn_desc* d);
size_t rnn_attention_tc_saved_bytes(const rnn_attn_desc* d);
size_t rnn_attention_tc_workspace(const rnn_attn_desc* d);
int rnn_attention_tc_fwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask,
                         void* z, void* saved, void* ws, size_t ws_bytes, cudaStream_t stream);
int rnn_attention_tc_bwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* saved,
                         const void* dz, void* dq, void* dk, void* dv, void* ws, size_t ws_bytes,
                         cudaStream_t stream);

/* Fused form of the same attention (bf16, head size 64 or 128, T and Tk multiples of 8, strides multiples of 8
 * elements -- rnn_attention_fused_supported()): ONE kernel per direction pair keeps scores and probabilities on the
 * SM -- q k^T into TMEM by tcgen05.mma, scale / mask / softmax / Philox dropout on the accumulator rows read back with
 * tcgen05.ld, probabilities handed to the second tcgen05.mma through shared memory -- so the [B,H,T,Tk] tensors of
 * nn/attention.py:37-47 (and of rnn_attention_tc_*) never touch HBM.  Saved for backward: z and lse [B,H,T] only.
 * Backward recomputes the probabilities from lse (same Philox counters): dq, then dk / dv; `delta` is [B,H,T] fp32
 * scratch.  Same masks (bool: nonzero = masked out; float32: additive), scale and dropout stream as rnn_attention_fwd. */
int rnn_attention_fused_supported(const rnn_attn_desc* d);
int rnn_attention_fused_fwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask,
                            void* z, float* lse, cudaStream_t stream);
int rnn_attention_fused_bwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask,
                            const void* z, const void* dz, const float* lse, float* delta, void* dq, void* dk, void* dv,
                            cudaStream_t stream);

/* fp32 form of the same pipeline at fp32 accuracy (the reference's default dtype): every operand of the six
 * contractions is split into three bf16 pieces and each contraction is one batched tcgen05 GEMM over a reduction
 * six times as long, fp32 out; head size, T and Tk multiples of 8.  `saved` keeps the fp32 probabilities. */
int rnn_attention_f32_tc_supported(const rnn_attn_desc* d);
size_t rnn_attention_f32_tc_saved_bytes(const rnn_attn_desc* d);
size_t rnn_attention_f32_tc_workspace(const rnn_attn_desc* d);
int rnn_attention_f32_tc_fwd(const rnn_attn_desc* d, const float* q, const float* k, const float* v, const void* mask,
                             float* z, void* saved, void* ws, size_t ws_bytes, cudaStream_t stream);
int rnn_attention_f32_tc_bwd(const rnn_attn_desc* d, const float* q, const float* k, const float* v,
                             const void* saved, const float* dz, float* dq, float* dk, float* dv, void* ws,
                             size_t ws_bytes, cudaStream_t stream);

/* ---- fused optimizers over flat buckets ---------------------------------------------------------- */
/* torch.optim.SGD (momentum, dampening, nesterov, weight_decay) on an fp32 master copy. */
int rnn_sgd_step(float* w, float* buf, const void* grad, int grad_bf16, void* wcopy, int wcopy_bf16, long long n,
                 float lr, float momentum, float dampening, float weight_decay, int nesterov, int first_step,
                 float grad_scale, cudaStream_t stream);
/* torch.optim.AdamW (amsgrad=False). */
int rnn_adamw_step(float* w, float* m, float* v, const void* grad, int grad_bf16, void* wcopy, int wcopy_bf16,
                   long long n, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                   float grad_scale, cudaStream_t stream);

/* AdamW with the step count in device memory, and the kernel that moves such a counter: a step built from these
 * has no host-side value that changes between steps, so it can be captured once and replayed (CUDA graph). */
int rnn_adamw_step_dev(float* w, float* m, float* v, const void* grad, int grad_bf16, void* wcopy, int wcopy_bf16,
                       long long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                       const long long* step_dev, float grad_scale, cudaStream_t stream);
int rnn_counter_add(long long* counter, long long inc, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* RNN_B200_H_ */
