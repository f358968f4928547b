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
/* autograd of Conv2d w.r.t. its input.  accumulate != 0: dx += result. */
int rnn_conv2d_dgrad_bf16(const rnn_conv_desc* d, const void* dy, const void* w, void* dx, int accumulate,
                          cudaStream_t stream);
/* autograd of Conv2d w.r.t. its weight; dw is bf16 [K,R,S,C] (fp32 when dw_f32 != 0). */
size_t rnn_conv2d_wgrad_workspace(const rnn_conv_desc* d);
int rnn_conv2d_wgrad_bf16(const rnn_conv_desc* d, const void* x, const void* dy, void* dw, int dw_f32,
                          int accumulate, void* ws, size_t ws_bytes, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* RNN_B200_H_ */
