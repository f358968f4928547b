// HBM-bound elementwise kernels: 128-bit vectorised, grid-stride, one pass over each tensor.
//   - input pipeline: NCHW fp32 -> NHWC bf16 (channel padding) -- SURVEY.md section 8(f).2
//   - ReLU forward / backward (reference: F.relu, arch/transformer.py:30)
//   - dropout forward / backward (reference: nn.Dropout, nn/attention.py:26,46 ; arch/transformer.py:26,32)
//   - add, casts
#include "common.cuh"

namespace {

constexpr int EW_THREADS = 256;

inline int ew_grid(long long nvec) {
  long long b = (nvec + EW_THREADS - 1) / EW_THREADS;
  const long long cap = (long long)rnn_num_sms() * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// ---------------------------------------------------------------- layout conversion
// One block handles one (n, h) row: reads W*C floats channel-planar (coalesced along w), transposes
// through shared memory, writes W*Cp bf16 pixel-major (coalesced along c).
//
// Normalisation (the torchvision `Normalize(mean, std)` step of an image pipeline) is folded into the same pass:
// `mean` / `inv_std` are per-channel device arrays (nullable = none), v = (x - mean[c]) * inv_std[c] before the
// bf16 rounding.  Spatial and channel padding stay exact zeros (zero padding happens after normalisation).
__global__ void nchw_f32_to_nhwc_bf16_kernel(const float* __restrict__ x, bf16* __restrict__ y, int C, int H,
                                             int W, int Cp, const float* __restrict__ mean,
                                             const float* __restrict__ inv_std) {
  extern __shared__ float tile[];  // [C][W+1]
  const int n = blockIdx.y, h = blockIdx.x;
  const float* src = x + ((long long)n * C * H + h) * W;
  for (int i = threadIdx.x; i < C * W; i += blockDim.x) {
    const int c = i / W, w = i - c * W;
    const float v = src[(long long)c * H * W + w];
    tile[c * (W + 1) + w] = mean ? (v - mean[c]) * inv_std[c] : v;
  }
  __syncthreads();
  bf16* dst = y + ((long long)n * H + h) * W * Cp;
  for (int i = threadIdx.x; i < W * Cp; i += blockDim.x) {
    const int w = i / Cp, c = i - w * Cp;
    dst[i] = __float2bfloat16(c < C ? tile[c * (W + 1) + w] : 0.f);
  }
}

// Same conversion into a spatially zero-padded buffer [N, Hp, Wp, Cp]: block (hp group, n) writes ROWS_PER_BLOCK
// padded rows; border rows/columns and padded channels are written as zeros, so no separate memset is needed.
// Cp == 8 (the packed-pixel first layer): one thread builds one whole pixel and stores it as 16 bytes.
constexpr int PAD_ROWS_PER_BLOCK = 4;
__global__ void __launch_bounds__(256)
nchw_f32_to_nhwc_bf16_padded_kernel(const float* __restrict__ x, bf16* __restrict__ y, int C, int H, int W, int Cp,
                                    int pad_h, int pad_w, int Hp, int Wp, const float* __restrict__ mean,
                                    const float* __restrict__ inv_std) {
  extern __shared__ float tile[];  // [C][W+1]
  const int n = blockIdx.y;
  for (int hp = blockIdx.x * PAD_ROWS_PER_BLOCK; hp < min((int)(blockIdx.x + 1) * PAD_ROWS_PER_BLOCK, Hp); ++hp) {
    const int h = hp - pad_h;
    const bool inside = h >= 0 && h < H;
    __syncthreads();   // the previous row's readers are done with the tile
    if (inside) {
      const float* src = x + ((long long)n * C * H + h) * W;
      if ((W & 3) == 0) {   // rows are 16-byte aligned: 128-bit loads
        const int w4 = W >> 2;
        for (int i = threadIdx.x; i < C * w4; i += blockDim.x) {
          const int c = i / w4, j = i - c * w4;
          float4 v = *reinterpret_cast<const float4*>(src + (long long)c * H * W + 4 * j);
          if (mean) {
            const float m = mean[c], is = inv_std[c];
            v.x = (v.x - m) * is; v.y = (v.y - m) * is; v.z = (v.z - m) * is; v.w = (v.w - m) * is;
          }
          float* t = tile + c * (W + 1) + 4 * j;
          t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
        }
      } else {
        for (int i = threadIdx.x; i < C * W; i += blockDim.x) {
          const int c = i / W, w = i - c * W;
          const float v = src[(long long)c * H * W + w];
          tile[c * (W + 1) + w] = mean ? (v - mean[c]) * inv_std[c] : v;
        }
      }
    }
    __syncthreads();
    bf16* dst = y + ((long long)n * Hp + hp) * Wp * Cp;
    if (Cp == 8) {
      for (int wp = threadIdx.x; wp < Wp; wp += blockDim.x) {
        const int w = wp - pad_w;
        float v[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) v[c] = (inside && w >= 0 && w < W && c < C) ? tile[c * (W + 1) + w] : 0.f;
        *reinterpret_cast<uint4*>(dst + wp * 8) = pack8(v);
      }
    } else {
      for (int i = threadIdx.x; i < Wp * Cp; i += blockDim.x) {
        const int wp = i / Cp, c = i - wp * Cp;
        const int w = wp - pad_w;
        const bool ok = inside && w >= 0 && w < W && c < C;
        dst[i] = __float2bfloat16(ok ? tile[c * (W + 1) + w] : 0.f);
      }
    }
  }
}

// The image path proper (Cp == 8, W % 4 == 0: every ImageNet-shaped input): a block owns IMG_ROWS padded rows, issues
// ALL of its 128-bit loads first (IMG_ROWS x C x W/4 float4: ~5 in flight per thread, against less than one in the
// row-at-a-time kernel above, which was latency-bound at 3.2 TB/s), keeps the rows in shared memory as dense
// float4s (no bank conflicts either way: threads walk consecutive addresses), then writes whole 16-byte pixels.
constexpr int IMG_ROWS = 8;
__global__ void __launch_bounds__(256)
image_to_packed_pixels_kernel(const float* __restrict__ x, bf16* __restrict__ y, int C, int H, int W, int pad_h,
                              int pad_w, int Hp, int Wp, const float* __restrict__ mean,
                              const float* __restrict__ inv_std) {
  extern __shared__ __align__(16) float rows[];  // [IMG_ROWS][C][W]
  const int n = blockIdx.y;
  const int hp0 = blockIdx.x * IMG_ROWS;
  const int w4 = W >> 2;
  const int per_row = C * w4;
  for (int i = threadIdx.x; i < IMG_ROWS * per_row; i += blockDim.x) {
    const int r = i / per_row, rem = i - r * per_row;
    const int c = rem / w4, j = rem - c * w4;
    const int h = hp0 + r - pad_h;
    if (h < 0 || h >= H || hp0 + r >= Hp) continue;
    float4 v = __ldcs(reinterpret_cast<const float4*>(x + (((long long)n * C + c) * H + h) * W + 4 * j));
    if (mean) {
      const float m = mean[c], is = inv_std[c];
      v.x = (v.x - m) * is; v.y = (v.y - m) * is; v.z = (v.z - m) * is; v.w = (v.w - m) * is;
    }
    reinterpret_cast<float4*>(rows)[i] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < IMG_ROWS * Wp; i += blockDim.x) {
    const int r = i / Wp, wp = i - r * Wp;
    const int hp = hp0 + r;
    if (hp >= Hp) break;
    const int h = hp - pad_h, w = wp - pad_w;
    const bool ok = h >= 0 && h < H && w >= 0 && w < W;
    float v[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) v[c] = (ok && c < C) ? rows[(r * C + c) * W + w] : 0.f;
    st_stream(y + (((long long)n * Hp + hp) * Wp + wp) * 8, pack8(v));
  }
}

__global__ void nhwc_bf16_to_nchw_f32_kernel(const bf16* __restrict__ x, float* __restrict__ y, int C, int H,
                                             int W, int Cp) {
  extern __shared__ float tile[];  // [C][W+1]
  const int n = blockIdx.y, h = blockIdx.x;
  const bf16* src = x + ((long long)n * H + h) * W * Cp;
  for (int i = threadIdx.x; i < W * Cp; i += blockDim.x) {
    const int w = i / Cp, c = i - w * Cp;
    if (c < C) tile[c * (W + 1) + w] = __bfloat162float(src[i]);
  }
  __syncthreads();
  float* dst = y + ((long long)n * C * H + h) * W;
  for (int i = threadIdx.x; i < C * W; i += blockDim.x) {
    const int c = i / W, w = i - c * W;
    dst[(long long)c * H * W + w] = tile[c * (W + 1) + w];
  }
}

// ---------------------------------------------------------------- relu / add
__global__ void relu_fwd_bf16_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float f[8];
    unpack8(ld_stream(x + i), f);
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
    st_stream(y + i, pack8(f));
  }
}
// dx = dy * (y > 0)   (y = the ReLU output)
__global__ void relu_bwd_bf16_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y, uint4* __restrict__ dx,
                                     long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float g[8], v[8];
    unpack8(ld_stream(dy + i), g);
    unpack8(ld_stream(y + i), v);
#pragma unroll
    for (int e = 0; e < 8; ++e) g[e] = v[e] > 0.f ? g[e] : 0.f;
    st_stream(dx + i, pack8(g));
  }
}
__global__ void add_bf16_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ y,
                                long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float f[8], g[8];
    unpack8(ld_stream(a + i), f);
    unpack8(ld_stream(b + i), g);
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] += g[e];
    st_stream(y + i, pack8(f));
  }
}
__global__ void relu_fwd_f32_kernel(const float4* __restrict__ x, float4* __restrict__ y, long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float4 v = x[i];
    v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
    y[i] = v;
  }
}
__global__ void relu_bwd_f32_kernel(const float4* __restrict__ dy, const float4* __restrict__ y, float4* __restrict__ dx,
                                    long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float4 g = dy[i];
    const float4 v = y[i];
    g.x = v.x > 0.f ? g.x : 0.f; g.y = v.y > 0.f ? g.y : 0.f; g.z = v.z > 0.f ? g.z : 0.f; g.w = v.w > 0.f ? g.w : 0.f;
    dx[i] = g;
  }
}
__global__ void relu_fwd_f32_scalar_kernel(const float* __restrict__ x, float* __restrict__ y, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = fmaxf(x[i], 0.f);
}
__global__ void relu_bwd_f32_scalar_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                           float* __restrict__ dx, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dx[i] = y[i] > 0.f ? dy[i] : 0.f;
}
__global__ void tanh_fwd_f32_kernel(const float* __restrict__ x, float* __restrict__ y, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = tanhf(x[i]);
}
__global__ void tanh_bwd_f32_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ dx,
                                    long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dx[i] = dy[i] * (1.f - y[i] * y[i]);
}

// ---------------------------------------------------------------- casts
__global__ void cast_f32_bf16_kernel(const float4* __restrict__ x, uint2* __restrict__ y, long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = x[i];
    uint2 u;
    u.x = pack2(v.x, v.y);
    u.y = pack2(v.z, v.w);
    y[i] = u;
  }
}
__global__ void cast_bf16_f32_kernel(const uint2* __restrict__ x, float4* __restrict__ y, long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    const uint2 u = x[i];
    const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
    y[i] = make_float4(a.x, a.y, b.x, b.y);
  }
}

// ---------------------------------------------------------------- dropout (counter-based RNG)
// Philox-4x32-10: each call yields four 32-bit words for (key=seed, counter=(idx, offset)).
__device__ __forceinline__ uint4 philox4x32(uint2 key, uint4 ctr) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// Device-side RNG epoch: added to every Philox `offset` of this file.  It stays 0 unless rnn_rng_advance() is
// used; a training step captured in a CUDA graph puts one rnn_rng_advance() at its head, so every replay draws new
// dropout masks although the (seed, offset) kernel arguments are frozen in the graph.
static __device__ unsigned long long g_rng_epoch_ew = 0ull;
__global__ void rng_advance_ew_kernel(unsigned long long inc) { g_rng_epoch_ew += inc; }

// y = x * keep / (1 - p); the keep mask is written as one byte per element for the backward pass.
template <typename T>
__global__ void dropout_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ mask, long long n,
                                   float p, unsigned long long seed, unsigned long long offset) {
  const float inv_keep = 1.f / (1.f - p);
  const uint32_t thresh = (uint32_t)(p * 4294967296.0);
  const long long n4 = (n + 3) / 4;
  offset += g_rng_epoch_ew;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const uint4 r = philox4x32(make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)),
                               make_uint4((uint32_t)i, (uint32_t)(i >> 32), (uint32_t)offset, (uint32_t)(offset >> 32)));
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const long long j = i * 4 + e;
      if (j < n) {
        const bool keep = rr[e] >= thresh;
        mask[j] = keep ? 1 : 0;
        y[j] = (T)(keep ? (float)x[j] * inv_keep : 0.f);
      }
    }
  }
}
template <typename T>
__global__ void dropout_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ mask, T* __restrict__ dx,
                                   long long n, float p) {
  const float inv_keep = 1.f / (1.f - p);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dx[i] = (T)(mask[i] ? (float)dy[i] * inv_keep : 0.f);
}

}  // namespace

int rnn_internal_rng_advance_ew(unsigned long long inc, cudaStream_t stream) {
  rng_advance_ew_kernel<<<1, 1, 0, stream>>>(inc);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

extern "C" {

int rnn_nchw_f32_to_nhwc_bf16_norm(const float* x, void* y, int N, int C, int H, int W, int Cpad, const float* mean,
                                   const float* inv_std, cudaStream_t stream) {
  RNN_REQUIRE(N > 0 && C > 0 && H > 0 && W > 0 && Cpad >= C, RNN_ERR_SHAPE);
  RNN_REQUIRE((mean == nullptr) == (inv_std == nullptr), RNN_ERR_SHAPE);
  const size_t smem = (size_t)C * (W + 1) * sizeof(float);
  RNN_REQUIRE(smem <= 48 * 1024, RNN_ERR_SHAPE);
  nchw_f32_to_nhwc_bf16_kernel<<<dim3(H, N), 256, smem, stream>>>(x, reinterpret_cast<bf16*>(y), C, H, W, Cpad, mean,
                                                                  inv_std);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int H, int W, int Cpad, cudaStream_t stream) {
  return rnn_nchw_f32_to_nhwc_bf16_norm(x, y, N, C, H, W, Cpad, nullptr, nullptr, stream);
}
int rnn_nchw_f32_to_nhwc_bf16_padded_norm(const float* x, void* y, int N, int C, int H, int W, int Cpad, int pad_h,
                                          int pad_w, int Hp, int Wp, const float* mean, const float* inv_std,
                                          cudaStream_t stream) {
  RNN_REQUIRE(N > 0 && C > 0 && H > 0 && W > 0 && Cpad >= C, RNN_ERR_SHAPE);
  RNN_REQUIRE(pad_h >= 0 && pad_w >= 0 && Hp >= H + pad_h && Wp >= W + pad_w, RNN_ERR_SHAPE);
  RNN_REQUIRE((mean == nullptr) == (inv_std == nullptr), RNN_ERR_SHAPE);
  const size_t img_smem = (size_t)IMG_ROWS * C * W * sizeof(float);
  if (Cpad == 8 && (W & 3) == 0 && img_smem <= 48 * 1024 && (reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
    image_to_packed_pixels_kernel<<<dim3(ceil_div_i(Hp, IMG_ROWS), N), 256, img_smem, stream>>>(
        x, reinterpret_cast<bf16*>(y), C, H, W, pad_h, pad_w, Hp, Wp, mean, inv_std);
    RNN_CHECK_LAUNCH();
    return RNN_OK;
  }
  const size_t smem = (size_t)C * (W + 1) * sizeof(float);
  RNN_REQUIRE(smem <= 48 * 1024, RNN_ERR_SHAPE);
  nchw_f32_to_nhwc_bf16_padded_kernel<<<dim3(ceil_div_i(Hp, PAD_ROWS_PER_BLOCK), N), 256, smem, stream>>>(
      x, reinterpret_cast<bf16*>(y), C, H, W, Cpad, pad_h, pad_w, Hp, Wp, mean, inv_std);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_nchw_f32_to_nhwc_bf16_padded(const float* x, void* y, int N, int C, int H, int W, int Cpad, int pad_h,
                                     int pad_w, int Hp, int Wp, cudaStream_t stream) {
  return rnn_nchw_f32_to_nhwc_bf16_padded_norm(x, y, N, C, H, W, Cpad, pad_h, pad_w, Hp, Wp, nullptr, nullptr, stream);
}
int rnn_nhwc_bf16_to_nchw_f32(const void* x, float* y, int N, int C, int H, int W, int Cpad, cudaStream_t stream) {
  RNN_REQUIRE(N > 0 && C > 0 && H > 0 && W > 0 && Cpad >= C, RNN_ERR_SHAPE);
  const size_t smem = (size_t)C * (W + 1) * sizeof(float);
  RNN_REQUIRE(smem <= 48 * 1024, RNN_ERR_SHAPE);
  nhwc_bf16_to_nchw_f32_kernel<<<dim3(H, N), 256, smem, stream>>>(reinterpret_cast<const bf16*>(x), y, C, H, W, Cpad);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int rnn_relu_fwd(const void* x, void* y, long long n, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && (is_f32 || n % 8 == 0), RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  if (is_f32 && n % 4 != 0) {
    relu_fwd_f32_scalar_kernel<<<ew_grid(n), EW_THREADS, 0, stream>>>((const float*)x, (float*)y, n);
    RNN_CHECK_LAUNCH();
    return RNN_OK;
  }
  const long long nvec = n / (is_f32 ? 4 : 8);
  if (is_f32) relu_fwd_f32_kernel<<<ew_grid(nvec), EW_THREADS, 0, stream>>>((const float4*)x, (float4*)y, nvec);
  else relu_fwd_bf16_kernel<<<ew_grid(nvec), EW_THREADS, 0, stream>>>((const uint4*)x, (uint4*)y, nvec);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_relu_bwd(const void* dy, const void* y, void* dx, long long n, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && (is_f32 || n % 8 == 0), RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  if (is_f32 && n % 4 != 0) {
    relu_bwd_f32_scalar_kernel<<<ew_grid(n), EW_THREADS, 0, stream>>>((const float*)dy, (const float*)y, (float*)dx, n);
    RNN_CHECK_LAUNCH();
    return RNN_OK;
  }
  const long long nvec = n / (is_f32 ? 4 : 8);
  if (is_f32)
    relu_bwd_f32_kernel<<<ew_grid(nvec), EW_THREADS, 0, stream>>>((const float4*)dy, (const float4*)y, (float4*)dx, nvec);
  else
    relu_bwd_bf16_kernel<<<ew_grid(nvec), EW_THREADS, 0, stream>>>((const uint4*)dy, (const uint4*)y, (uint4*)dx, nvec);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_tanh_fwd_f32(const float* x, float* y, long long n, cudaStream_t stream) {
  if (n == 0) return RNN_OK;
  tanh_fwd_f32_kernel<<<ew_grid(n), EW_THREADS, 0, stream>>>(x, y, n);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_tanh_bwd_f32(const float* dy, const float* y, float* dx, long long n, cudaStream_t stream) {
  if (n == 0) return RNN_OK;
  tanh_bwd_f32_kernel<<<ew_grid(n), EW_THREADS, 0, stream>>>(dy, y, dx, n);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_add_bf16(const void* a, const void* b, void* y, long long n, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && n % 8 == 0, RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  add_bf16_kernel<<<ew_grid(n / 8), EW_THREADS, 0, stream>>>((const uint4*)a, (const uint4*)b, (uint4*)y, n / 8);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_cast_f32_to_bf16(const float* x, void* y, long long n, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && n % 4 == 0, RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  cast_f32_bf16_kernel<<<ew_grid(n / 4), EW_THREADS, 0, stream>>>((const float4*)x, (uint2*)y, n / 4);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_cast_bf16_to_f32(const void* x, float* y, long long n, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && n % 4 == 0, RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  cast_bf16_f32_kernel<<<ew_grid(n / 4), EW_THREADS, 0, stream>>>((const uint2*)x, (float4*)y, n / 4);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_dropout_fwd(const void* x, void* y, uint8_t* mask, long long n, float p, unsigned long long seed,
                    unsigned long long offset, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && p >= 0.f && p < 1.f, RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  const int grid = ew_grid((n + 3) / 4);
  if (is_f32) dropout_fwd_kernel<float><<<grid, EW_THREADS, 0, stream>>>((const float*)x, (float*)y, mask, n, p, seed, offset);
  else dropout_fwd_kernel<bf16><<<grid, EW_THREADS, 0, stream>>>((const bf16*)x, (bf16*)y, mask, n, p, seed, offset);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_dropout_bwd(const void* dy, const uint8_t* mask, void* dx, long long n, float p, int is_f32,
                    cudaStream_t stream) {
  RNN_REQUIRE(n >= 0 && p >= 0.f && p < 1.f, RNN_ERR_SHAPE);
  if (n == 0) return RNN_OK;
  if (is_f32) dropout_bwd_kernel<float><<<ew_grid(n), EW_THREADS, 0, stream>>>((const float*)dy, mask, (float*)dx, n, p);
  else dropout_bwd_kernel<bf16><<<ew_grid(n), EW_THREADS, 0, stream>>>((const bf16*)dy, mask, (bf16*)dx, n, p);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

}  // extern "C"
