// Row softmax and fused softmax + cross-entropy (forward and gradient in one pass).
//   softmax      : reference F.softmax(w, dim=-1) at nn/attention.py:45 (standalone form; the fused
//                  attention kernel carries its own copy)
//   softmax-CE   : torch.nn.functional.cross_entropy(logits, target) with mean reduction, plus
//                  dlogits = (softmax - onehot) * grad_scale in the same pass over the logits.
#include "common.cuh"

namespace {

template <typename T> __device__ __forceinline__ float ldf(const T* p, long long i) { return (float)p[i]; }

// one block per row; the row is read twice from L1/L2 (max, then exp-sum) and written once.
template <typename T>
__global__ void __launch_bounds__(256)
softmax_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int cols) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const T* xr = x + row * cols;
  float m = -INFINITY;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) m = fmaxf(m, (float)xr[c]);
  m = block_max(m, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) s += __expf((float)xr[c] - m);
  s = block_sum(s, red);
  const float inv = 1.f / s;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) y[row * cols + c] = (T)(__expf((float)xr[c] - m) * inv);
}

// dx = y * (dy - sum(dy*y))
template <typename T>
__global__ void __launch_bounds__(256)
softmax_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y, T* __restrict__ dx, int cols) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  float s = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) s += (float)dy[row * cols + c] * (float)y[row * cols + c];
  s = block_sum(s, red);
  for (int c = threadIdx.x; c < cols; c += blockDim.x)
    dx[row * cols + c] = (T)((float)y[row * cols + c] * ((float)dy[row * cols + c] - s));
}

template <typename T>
__global__ void __launch_bounds__(256)
softmax_ce_kernel(const T* __restrict__ logits, const long long* __restrict__ target, int cols, float grad_scale,
                  float* __restrict__ row_loss, T* __restrict__ dlogits) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const T* xr = logits + row * cols;
  float m = -INFINITY;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) m = fmaxf(m, (float)xr[c]);
  m = block_max(m, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) s += expf((float)xr[c] - m);
  s = block_sum(s, red);
  const float lse = m + logf(s);
  const int t = (int)target[row];
  if (threadIdx.x == 0) row_loss[row] = (t >= 0 && t < cols) ? lse - (float)xr[t] : 0.f;
  if (dlogits) {
    const float inv = 1.f / s;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      const float p = expf((float)xr[c] - m) * inv;
      dlogits[row * cols + c] = (T)((p - (c == t ? 1.f : 0.f)) * grad_scale);
    }
  }
}

__global__ void mean_kernel(const float* __restrict__ v, int n, float* __restrict__ out) {
  __shared__ float red[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += v[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) *out = s / (float)n;
}

// argmax over the last dimension (first maximum wins, as torch.argmax) -- bit-exact index check
template <typename T>
__global__ void __launch_bounds__(256)
argmax_kernel(const T* __restrict__ x, long long* __restrict__ out, int cols) {
  __shared__ float bv[256];
  __shared__ int bi[256];
  const long long row = blockIdx.x;
  float best = -INFINITY;
  int idx = 0x7fffffff;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    const float v = (float)x[row * cols + c];
    if (v > best || (v == best && c < idx)) { best = v; idx = c; }
  }
  bv[threadIdx.x] = best;
  bi[threadIdx.x] = idx;
  __syncthreads();
  for (int s = blockDim.x / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      const float v = bv[threadIdx.x + s];
      const int i = bi[threadIdx.x + s];
      if (v > bv[threadIdx.x] || (v == bv[threadIdx.x] && i < bi[threadIdx.x])) { bv[threadIdx.x] = v; bi[threadIdx.x] = i; }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) out[row] = bi[0] == 0x7fffffff ? 0 : bi[0];
}

}  // namespace

extern "C" {

int rnn_softmax_fwd(const void* x, void* y, long long rows, int cols, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(rows > 0 && cols > 0 && rows < (1ll << 31), RNN_ERR_SHAPE);
  if (is_f32) softmax_fwd_kernel<float><<<(int)rows, 256, 0, stream>>>((const float*)x, (float*)y, cols);
  else softmax_fwd_kernel<bf16><<<(int)rows, 256, 0, stream>>>((const bf16*)x, (bf16*)y, cols);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_softmax_bwd(const void* dy, const void* y, void* dx, long long rows, int cols, int is_f32,
                    cudaStream_t stream) {
  RNN_REQUIRE(rows > 0 && cols > 0 && rows < (1ll << 31), RNN_ERR_SHAPE);
  if (is_f32) softmax_bwd_kernel<float><<<(int)rows, 256, 0, stream>>>((const float*)dy, (const float*)y, (float*)dx, cols);
  else softmax_bwd_kernel<bf16><<<(int)rows, 256, 0, stream>>>((const bf16*)dy, (const bf16*)y, (bf16*)dx, cols);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
/* loss = mean_rows(logsumexp(logits) - logits[target]); dlogits (nullable) = (softmax - onehot)*grad_scale.
 * row_loss is a caller buffer of `rows` floats. */
int rnn_softmax_ce(const void* logits, const long long* target, long long rows, int cols, float grad_scale,
                   float* row_loss, float* loss, void* dlogits, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(rows > 0 && cols > 0 && rows < (1ll << 31), RNN_ERR_SHAPE);
  if (is_f32)
    softmax_ce_kernel<float><<<(int)rows, 256, 0, stream>>>((const float*)logits, target, cols, grad_scale, row_loss,
                                                            (float*)dlogits);
  else
    softmax_ce_kernel<bf16><<<(int)rows, 256, 0, stream>>>((const bf16*)logits, target, cols, grad_scale, row_loss,
                                                           (bf16*)dlogits);
  RNN_CHECK_LAUNCH();
  if (loss) {
    mean_kernel<<<1, 256, 0, stream>>>(row_loss, (int)rows, loss);
    RNN_CHECK_LAUNCH();
  }
  return RNN_OK;
}
int rnn_argmax(const void* x, long long* out, long long rows, int cols, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(rows > 0 && cols > 0 && rows < (1ll << 31), RNN_ERR_SHAPE);
  if (is_f32) argmax_kernel<float><<<(int)rows, 256, 0, stream>>>((const float*)x, out, cols);
  else argmax_kernel<bf16><<<(int)rows, 256, 0, stream>>>((const bf16*)x, out, cols);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

}  // extern "C"
