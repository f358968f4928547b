// Fused optimizer updates over flat parameter buckets (torch.optim.SGD / AdamW semantics).
// One pass: read fp32 master weight + state + gradient, write master weight + state + the bf16 (or
// fp32) working copy the forward kernels read.  18 B/param for SGD-momentum with bf16 grads.
#include "common.cuh"

namespace {

template <typename G> __device__ __forceinline__ void load4(const G* g, long long i, float* f);
template <> __device__ __forceinline__ void load4<float>(const float* g, long long i, float* f) {
  const float4 v = *reinterpret_cast<const float4*>(g + i);
  f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}
template <> __device__ __forceinline__ void load4<bf16>(const bf16* g, long long i, float* f) {
  const uint2 u = *reinterpret_cast<const uint2*>(g + i);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}
__device__ __forceinline__ void store4(float* p, long long i, const float* f) {
  *reinterpret_cast<float4*>(p + i) = make_float4(f[0], f[1], f[2], f[3]);
}
__device__ __forceinline__ void store4(bf16* p, long long i, const float* f) {
  uint2 u;
  u.x = pack2(f[0], f[1]);
  u.y = pack2(f[2], f[3]);
  *reinterpret_cast<uint2*>(p + i) = u;
}

// torch.optim.SGD: g += wd*w; buf = first ? g : mu*buf + (1-damp)*g; g = nesterov ? g + mu*buf : buf; w -= lr*g
template <typename G, typename Wc>
__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ w, float* __restrict__ buf, const G* __restrict__ grad, Wc* __restrict__ wcopy,
           long long n, float lr, float mu, float damp, float wd, int nesterov, int first, float gscale) {
  for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n;
       i += (long long)gridDim.x * blockDim.x * 4) {
    float p[4], g[4], b[4] = {0.f, 0.f, 0.f, 0.f};
    load4<float>(w, i, p);
    load4<G>(grad, i, g);
    if (buf && !first) load4<float>(buf, i, b);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float d = g[e] * gscale + wd * p[e];
      if (buf) {
        b[e] = first ? d : mu * b[e] + (1.f - damp) * d;
        d = nesterov ? d + mu * b[e] : b[e];
      }
      p[e] -= lr * d;
    }
    store4(w, i, p);
    if (buf) store4(buf, i, b);
    if (wcopy) store4(wcopy, i, p);
  }
}

// torch.optim.AdamW (amsgrad=False): w *= 1-lr*wd; m,v EMA; w -= lr/bc1 * m / (sqrt(v)/sqrt(bc2) + eps)
template <typename G, typename Wc>
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ w, float* __restrict__ m, float* __restrict__ v, const G* __restrict__ grad,
             Wc* __restrict__ wcopy, long long n, float lr, float b1, float b2, float eps, float wd, float bc1,
             float bc2_sqrt, float gscale, const long long* __restrict__ step_dev) {
  if (step_dev != nullptr) {   // step count kept on the device (CUDA-graph replays): bias corrections from it
    const float st = (float)(*step_dev);
    bc1 = 1.f - powf(b1, st);
    bc2_sqrt = sqrtf(1.f - powf(b2, st));
  }
  for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n;
       i += (long long)gridDim.x * blockDim.x * 4) {
    float p[4], g[4], mm[4], vv[4];
    load4<float>(w, i, p);
    load4<G>(grad, i, g);
    load4<float>(m, i, mm);
    load4<float>(v, i, vv);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float d = g[e] * gscale;
      p[e] *= 1.f - lr * wd;
      mm[e] = b1 * mm[e] + (1.f - b1) * d;
      vv[e] = b2 * vv[e] + (1.f - b2) * d * d;
      const float denom = sqrtf(vv[e]) / bc2_sqrt + eps;
      p[e] -= (lr / bc1) * (mm[e] / denom);
    }
    store4(w, i, p);
    store4(m, i, mm);
    store4(v, i, vv);
    if (wcopy) store4(wcopy, i, p);
  }
}

int ogrid(long long n) {
  long long b = (n / 4 + 255) / 256;
  const long long cap = (long long)rnn_num_sms() * 8;
  return (int)(b > cap ? cap : (b < 1 ? 1 : b));
}

__global__ void counter_add_kernel(long long* c, long long inc) { *c += inc; }

}  // namespace

extern "C" {

/* n % 4 == 0 (buckets are padded).  grad_bf16: gradient dtype; wcopy (nullable): working copy, bf16 when
 * wcopy_bf16 != 0 else fp32.  buf == NULL disables momentum.  grad_scale multiplies the gradient first
 * (1/world_size after an all-reduce sum). */
int rnn_sgd_step(float* w, float* buf, const void* grad, int grad_bf16, void* wcopy, int wcopy_bf16, long long n,
                 float lr, float momentum, float dampening, float weight_decay, int nesterov, int first_step,
                 float grad_scale, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && n % 4 == 0, RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  const int g = ogrid(n);
#define SGD_LAUNCH(G, W) \
  sgd_kernel<G, W><<<g, 256, 0, stream>>>(w, buf, (const G*)grad, (W*)wcopy, n, lr, momentum, dampening, weight_decay, \
                                          nesterov, first_step, grad_scale)
  if (grad_bf16) { if (wcopy_bf16) SGD_LAUNCH(bf16, bf16); else SGD_LAUNCH(bf16, float); }
  else { if (wcopy_bf16) SGD_LAUNCH(float, bf16); else SGD_LAUNCH(float, float); }
#undef SGD_LAUNCH
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

static int adamw_launch(float* w, float* m, float* v, const void* grad, int grad_bf16, void* wcopy, int wcopy_bf16,
                        long long n, float lr, float beta1, float beta2, float eps, float weight_decay, float bc1,
                        float bc2s, const long long* step_dev, float grad_scale, cudaStream_t stream) {
  const int g = ogrid(n);
#define ADAM_LAUNCH(G, W) \
  adamw_kernel<G, W><<<g, 256, 0, stream>>>(w, m, v, (const G*)grad, (W*)wcopy, n, lr, beta1, beta2, eps, weight_decay, \
                                            bc1, bc2s, grad_scale, step_dev)
  if (grad_bf16) { if (wcopy_bf16) ADAM_LAUNCH(bf16, bf16); else ADAM_LAUNCH(bf16, float); }
  else { if (wcopy_bf16) ADAM_LAUNCH(float, bf16); else ADAM_LAUNCH(float, float); }
#undef ADAM_LAUNCH
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int rnn_adamw_step(float* w, float* m, float* v, const void* grad, int grad_bf16, void* wcopy, int wcopy_bf16,
                   long long n, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                   float grad_scale, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && n % 4 == 0 && step >= 1, RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float bc2s = sqrtf(1.f - powf(beta2, (float)step));
  return adamw_launch(w, m, v, grad, grad_bf16, wcopy, wcopy_bf16, n, lr, beta1, beta2, eps, weight_decay, bc1, bc2s,
                      nullptr, grad_scale, stream);
}

/* Same update with the step count read from device memory (*step_dev >= 1 when the kernel runs): nothing about the
 * call changes from step to step, so it can be replayed from a CUDA graph.  rnn_counter_add moves the counter. */
int rnn_adamw_step_dev(float* w, float* m, float* v, const void* grad, int grad_bf16, void* wcopy, int wcopy_bf16,
                       long long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                       const long long* step_dev, float grad_scale, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && n % 4 == 0 && step_dev != nullptr, RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  return adamw_launch(w, m, v, grad, grad_bf16, wcopy, wcopy_bf16, n, lr, beta1, beta2, eps, weight_decay, 1.f, 1.f,
                      step_dev, grad_scale, stream);
}
int rnn_counter_add(long long* counter, long long inc, cudaStream_t stream) {
  RNN_REQUIRE(counter != nullptr, RNN_ERR_SHAPE);
  counter_add_kernel<<<1, 1, 0, stream>>>(counter, inc);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

}  // extern "C"
