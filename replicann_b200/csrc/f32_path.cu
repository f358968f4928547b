// fp32 CUDA-core path for the small classic networks (LeNet-5, BASELINE.json configs[0]) whose channel
// counts are far below a tensor-core tile: direct NHWC convolution (forward, dgrad, wgrad) and a tiled
// SGEMM with arbitrary operand strides.  torch.nn.Conv2d / nn.Linear semantics; fp32 FMA accumulation so
// the result stays within 1e-4 relative of the CPU oracle.  Also the column-sum used for bias gradients.
#include "common.cuh"

namespace {

struct CG {  // conv geometry
  int N, H, W, C, K, R, S, sh, sw, ph, pw, P, Q;
};

__global__ void __launch_bounds__(256)
conv_fwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                    float* __restrict__ y, CG g) {
  const long long total = (long long)g.N * g.P * g.Q * g.K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % g.K);
    long long t = i / g.K;
    const int q = (int)(t % g.Q); t /= g.Q;
    const int p = (int)(t % g.P);
    const int n = (int)(t / g.P);
    float acc = bias ? bias[k] : 0.f;
    for (int r = 0; r < g.R; ++r) {
      const int h = p * g.sh - g.ph + r;
      if (h < 0 || h >= g.H) continue;
      for (int s = 0; s < g.S; ++s) {
        const int ww = q * g.sw - g.pw + s;
        if (ww < 0 || ww >= g.W) continue;
        const float* xp = x + (((long long)n * g.H + h) * g.W + ww) * g.C;
        const float* wp = w + (((long long)k * g.R + r) * g.S + s) * g.C;
        for (int c = 0; c < g.C; ++c) acc = fmaf(xp[c], wp[c], acc);
      }
    }
    y[i] = acc;
  }
}

__global__ void __launch_bounds__(256)
conv_dgrad_f32_kernel(const float* __restrict__ dy, const float* __restrict__ w, float* __restrict__ dx, CG g) {
  const long long total = (long long)g.N * g.H * g.W * g.C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % g.C);
    long long t = i / g.C;
    const int ww = (int)(t % g.W); t /= g.W;
    const int h = (int)(t % g.H);
    const int n = (int)(t / g.H);
    float acc = 0.f;
    for (int r = 0; r < g.R; ++r) {
      const int hp = h + g.ph - r;
      if (hp < 0 || hp % g.sh != 0) continue;
      const int p = hp / g.sh;
      if (p >= g.P) continue;
      for (int s = 0; s < g.S; ++s) {
        const int wq = ww + g.pw - s;
        if (wq < 0 || wq % g.sw != 0) continue;
        const int q = wq / g.sw;
        if (q >= g.Q) continue;
        const float* dp = dy + (((long long)n * g.P + p) * g.Q + q) * g.K;
        const float* wp = w + ((long long)r * g.S + s) * g.C + c;
        const long long wk = (long long)g.R * g.S * g.C;
        for (int k = 0; k < g.K; ++k) acc = fmaf(dp[k], wp[k * wk], acc);
      }
    }
    dx[i] = acc;
  }
}

// one block per filter element (k,r,s,c); threads stride over the N*P*Q output pixels
__global__ void __launch_bounds__(512)
conv_wgrad_f32_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dw, CG g,
                      int accumulate) {
  __shared__ float red[32];
  const int e = blockIdx.x;
  const int c = e % g.C;
  int t = e / g.C;
  const int s = t % g.S; t /= g.S;
  const int r = t % g.R;
  const int k = t / g.R;
  // 512 threads stride over the N*P*Q output pixels with 32-bit index arithmetic (the 128-thread form with 64-bit
  // divisions was 47 % of a LeNet-5 training step); four independent pixels in flight per thread
  const unsigned npq = (unsigned)g.N * g.P * g.Q, PQ = (unsigned)g.P * g.Q;
  float acc = 0.f;
#pragma unroll 4
  for (unsigned i = threadIdx.x; i < npq; i += blockDim.x) {
    const unsigned n = i / PQ, rem = i - n * PQ;
    const unsigned p = rem / g.Q, q = rem - p * g.Q;
    const int h = (int)p * g.sh - g.ph + r, ww = (int)q * g.sw - g.pw + s;
    const bool ok = h >= 0 && h < g.H && ww >= 0 && ww < g.W;
    const float xv = ok ? x[(((long long)n * g.H + h) * g.W + ww) * g.C + c] : 0.f;
    acc = fmaf(xv, dy[(long long)i * g.K + k], acc);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) dw[e] = (accumulate ? dw[e] : 0.f) + acc;
}

// D[m,n] = sum_k A(m,k)*B(n,k) (+bias[n]) ; A(m,k) = A[m*sam + k*sak], B(n,k) = B[n*sbn + k*sbk]
constexpr int TS = 64, TK = 16;
__global__ void __launch_bounds__(256)
sgemm_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int M, int N, int K,
             long long sam, long long sak, long long sbn, long long sbk, const float* __restrict__ bias, int relu,
             int accumulate) {
  __shared__ float sa[TK][TS + 1], sb[TK][TS + 1];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * TS, n0 = blockIdx.x * TS;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += TK) {
    for (int i = threadIdx.x; i < TS * TK; i += 256) {
      const int kk = i % TK, mm = i / TK;
      const int m = m0 + mm, n = n0 + mm, k = k0 + kk;
      sa[kk][mm] = (m < M && k < K) ? A[m * sam + k * sak] : 0.f;
      sb[kk][mm] = (n < N && k < K) ? B[n * sbn + k * sbk] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = sa[kk][ty * 4 + i]; b[i] = sb[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (m < M && n < N) {
        float v = acc[i][j] + (bias ? bias[n] : 0.f);
        if (accumulate) v += D[(long long)m * N + n];
        if (relu) v = fmaxf(v, 0.f);
        D[(long long)m * N + n] = v;
      }
    }
}

// column sums of an [M, C] matrix: partials[gridDim.y][C] then a finalize pass (deterministic order)
template <typename T>
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const T* __restrict__ x, long long M, int C, float* __restrict__ partials) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  float s = 0.f;
  for (long long r = blockIdx.y; r < M; r += gridDim.y) s += (float)x[r * C + c];
  partials[(long long)blockIdx.y * C + c] = s;
}
// Vector form for C a multiple of the 16-byte vector width (every bias gradient of the CNNs): a thread owns one
// vector column and walks its rows four at a time with 16-byte streaming loads; the row-lanes of a block are
// combined through shared memory.  (The scalar kernel above used 64 of 256 threads and 2-byte loads at C = 64 and
// ran at a tenth of HBM speed: 24 % of a VGG-16 training step.)
template <typename T> struct CsVec;
template <> struct CsVec<bf16> {
  static constexpr int W = 8;
  __device__ static void load(const bf16* p, float* f) { unpack8(ld_stream(p), f); }
};
template <> struct CsVec<float> {
  static constexpr int W = 4;
  __device__ static void load(const float* p, float* f) {
    const uint4 u = ld_stream(p);
    f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
  }
};
template <typename T>
__global__ void __launch_bounds__(256)
colsum_vec_kernel(const T* __restrict__ x, long long M, int C, float* __restrict__ partials) {
  constexpr int W = CsVec<T>::W, U = 4;
  const int cgs = C / W;
  const int tpr = cgs < 256 ? cgs : 256;   // threads per row
  const int rpp = 256 / tpr;               // rows per pass
  const int cg = blockIdx.x * tpr + (threadIdx.x % tpr);
  const int rlocal = threadIdx.x / tpr;
  float acc[W];
#pragma unroll
  for (int e = 0; e < W; ++e) acc[e] = 0.f;
  const bool active = cg < cgs && rlocal < rpp;
  if (active) {
    const long long stride = (long long)gridDim.y * rpp;
    long long r = (long long)blockIdx.y * rpp + rlocal;
    for (; r + (U - 1) * stride < M; r += U * stride) {
      float v[U][W];
#pragma unroll
      for (int u = 0; u < U; ++u) CsVec<T>::load(x + (r + u * stride) * C + cg * W, v[u]);
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int e = 0; e < W; ++e) acc[e] += v[u][e];
    }
    for (; r < M; r += stride) {
      float v[W];
      CsVec<T>::load(x + r * C + cg * W, v);
#pragma unroll
      for (int e = 0; e < W; ++e) acc[e] += v[e];
    }
  }
  __shared__ float sm[256 * W];
#pragma unroll
  for (int e = 0; e < W; ++e) sm[threadIdx.x * W + e] = active ? acc[e] : 0.f;
  __syncthreads();
  if (rlocal == 0 && cg < cgs) {
    for (int rr = 1; rr < rpp; ++rr) {
      const int t = rr * tpr + (threadIdx.x % tpr);
#pragma unroll
      for (int e = 0; e < W; ++e) acc[e] += sm[t * W + e];
    }
    float* p = partials + (long long)blockIdx.y * C + cg * W;
#pragma unroll
    for (int e = 0; e < W; ++e) p[e] = acc[e];
  }
}
// ReLU backward fused with the bias gradient that follows it in Conv2d / Linear backward: dx = dy * (y > 0) is
// written and its column sums are accumulated in the same pass (the separate column-sum kernel re-read dx).
template <typename T> struct CsStore;
template <> struct CsStore<bf16> { __device__ static void store(bf16* p, const float* f) { st_stream(p, pack8(f)); } };
template <> struct CsStore<float> {
  __device__ static void store(float* p, const float* f) {
    st_stream(p, make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3])));
  }
};
template <typename T>
__global__ void __launch_bounds__(256)
relu_bwd_colsum_kernel(const T* __restrict__ dy, const T* __restrict__ y, T* __restrict__ dx, long long M, int C,
                       float* __restrict__ partials) {
  constexpr int W = CsVec<T>::W, U = 2;
  const int cgs = C / W;
  const int tpr = cgs < 256 ? cgs : 256;
  const int rpp = 256 / tpr;
  const int cg = blockIdx.x * tpr + (threadIdx.x % tpr);
  const int rlocal = threadIdx.x / tpr;
  float acc[W];
#pragma unroll
  for (int e = 0; e < W; ++e) acc[e] = 0.f;
  const bool active = cg < cgs && rlocal < rpp;
  if (active) {
    const long long stride = (long long)gridDim.y * rpp;
    long long r = (long long)blockIdx.y * rpp + rlocal;
    for (; r + (U - 1) * stride < M; r += U * stride) {
      float g[U][W], o[U][W];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        CsVec<T>::load(dy + (r + u * stride) * C + cg * W, g[u]);
        CsVec<T>::load(y + (r + u * stride) * C + cg * W, o[u]);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int e = 0; e < W; ++e) { g[u][e] = o[u][e] > 0.f ? g[u][e] : 0.f; acc[e] += g[u][e]; }
        CsStore<T>::store(dx + (r + u * stride) * C + cg * W, g[u]);
      }
    }
    for (; r < M; r += stride) {
      float g[W], o[W];
      CsVec<T>::load(dy + r * C + cg * W, g);
      CsVec<T>::load(y + r * C + cg * W, o);
#pragma unroll
      for (int e = 0; e < W; ++e) { g[e] = o[e] > 0.f ? g[e] : 0.f; acc[e] += g[e]; }
      CsStore<T>::store(dx + r * C + cg * W, g);
    }
  }
  __shared__ float sm[256 * W];
#pragma unroll
  for (int e = 0; e < W; ++e) sm[threadIdx.x * W + e] = active ? acc[e] : 0.f;
  __syncthreads();
  if (rlocal == 0 && cg < cgs) {
    for (int rr = 1; rr < rpp; ++rr) {
      const int t = rr * tpr + (threadIdx.x % tpr);
#pragma unroll
      for (int e = 0; e < W; ++e) acc[e] += sm[t * W + e];
    }
    float* p = partials + (long long)blockIdx.y * C + cg * W;
#pragma unroll
    for (int e = 0; e < W; ++e) p[e] = acc[e];
  }
}

// row blocks of the vector form: every thread sees >= 16 rows, at most about 4 resident blocks per SM in total
int colsum_vec_rows(long long M, int C, int W) {
  const int cgs = C / W, tpr = cgs < 256 ? cgs : 256, rpp = 256 / tpr, gx = (cgs + tpr - 1) / tpr;
  long long gy = (M + (long long)rpp * 16 - 1) / ((long long)rpp * 16);   // >= 16 rows per thread: few partial rows
  long long cap = (long long)rnn_num_sms() * 4 / gx;
  if (cap < 1) cap = 1;
  if (gy > cap) gy = cap;
  return (int)(gy < 1 ? 1 : gy);
}

// 128 threads = 32 columns x 4 row lanes: coalesced partial rows, four independent sums per column, fixed order
template <typename T>
__global__ void colsum_final_kernel(const float* __restrict__ partials, int nparts, int C, T* out, int accumulate) {
  __shared__ float sm[4][33];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (c < C)
    for (int p = w; p < nparts; p += 4) s += partials[(long long)p * C + c];
  sm[w][lane] = s;
  __syncthreads();
  if (w != 0 || c >= C) return;
  s = (sm[0][lane] + sm[1][lane]) + (sm[2][lane] + sm[3][lane]);
  if (accumulate) s += (float)out[c];
  out[c] = (T)s;
}

int fgrid(long long total) {
  long long b = (total + 255) / 256;
  const long long cap = (long long)rnn_num_sms() * 16;
  return (int)(b > cap ? cap : (b < 1 ? 1 : b));
}
CG to_cg(const rnn_conv_desc* d) {
  CG g{d->N, d->H, d->W, d->C, d->K, d->R, d->S, d->stride_h, d->stride_w, d->pad_h, d->pad_w, d->P, d->Q};
  return g;
}
int check_desc_f32(const rnn_conv_desc* d) {
  RNN_REQUIRE(d != nullptr, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->N > 0 && d->H > 0 && d->W > 0 && d->C > 0 && d->K > 0 && d->R > 0 && d->S > 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->stride_h > 0 && d->stride_w > 0 && d->pad_h >= 0 && d->pad_w >= 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->P == (d->H + 2 * d->pad_h - d->R) / d->stride_h + 1, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->Q == (d->W + 2 * d->pad_w - d->S) / d->stride_w + 1, RNN_ERR_SHAPE);
  return RNN_OK;
}
int colsum_rows(long long M) {
  long long r = (M + 63) / 64;
  return (int)(r > 512 ? 512 : (r < 1 ? 1 : r));
}

}  // namespace

extern "C" {

int rnn_conv2d_fprop_f32(const rnn_conv_desc* d, const float* x, const float* w, const float* bias, float* y,
                         cudaStream_t stream) {
  int rc = check_desc_f32(d);
  if (rc != RNN_OK) return rc;
  conv_fwd_f32_kernel<<<fgrid((long long)d->N * d->P * d->Q * d->K), 256, 0, stream>>>(x, w, bias, y, to_cg(d));
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_conv2d_dgrad_f32(const rnn_conv_desc* d, const float* dy, const float* w, float* dx, cudaStream_t stream) {
  int rc = check_desc_f32(d);
  if (rc != RNN_OK) return rc;
  conv_dgrad_f32_kernel<<<fgrid((long long)d->N * d->H * d->W * d->C), 256, 0, stream>>>(dy, w, dx, to_cg(d));
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_conv2d_wgrad_f32(const rnn_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate,
                         cudaStream_t stream) {
  int rc = check_desc_f32(d);
  if (rc != RNN_OK) return rc;
  conv_wgrad_f32_kernel<<<d->K * d->R * d->S * d->C, 512, 0, stream>>>(x, dy, dw, to_cg(d), accumulate);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
/* D[M,N] = op(A).op(B)^T (+bias)(relu); a_mn / b_mn as in rnn_gemm_bf16. */
int rnn_gemm_f32(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn,
                 const float* bias, int relu, int accumulate, cudaStream_t stream) {
  RNN_REQUIRE(M > 0 && N > 0 && K > 0, RNN_ERR_SHAPE);
  const long long sam = a_mn ? 1 : K, sak = a_mn ? M : 1, sbn = b_mn ? 1 : K, sbk = b_mn ? N : 1;
  sgemm_kernel<<<dim3(ceil_div_i(N, TS), ceil_div_i(M, TS)), 256, 0, stream>>>(A, B, D, M, N, K, sam, sak, sbn, sbk,
                                                                                bias, relu, accumulate);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
size_t rnn_colsum_workspace(long long M, int C) {
  int rows = colsum_rows(M);
  if (C > 0 && C % 4 == 0) { const int r4 = colsum_vec_rows(M, C, 4); rows = r4 > rows ? r4 : rows; }
  if (C > 0 && C % 8 == 0) { const int r8 = colsum_vec_rows(M, C, 8); rows = r8 > rows ? r8 : rows; }
  return (size_t)rows * C * sizeof(float);
}
/* out[c] (+)= sum_r x[r,c]; x and out share a dtype (bias gradients of Conv2d / Linear). */
int rnn_colsum(const void* x, long long M, int C, void* out, int accumulate, int is_f32, void* ws, size_t ws_bytes,
               cudaStream_t stream) {
  RNN_REQUIRE(M > 0 && C > 0, RNN_ERR_SHAPE);
  const int W = is_f32 ? 4 : 8;
  if (C % W == 0 && (reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
    const int rows = colsum_vec_rows(M, C, W);
    RNN_REQUIRE(ws != nullptr && ws_bytes >= (size_t)rows * C * sizeof(float), RNN_ERR_WORKSPACE);
    const int cgs = C / W, tpr = cgs < 256 ? cgs : 256;
    const dim3 grid((cgs + tpr - 1) / tpr, rows);
    if (is_f32) {
      colsum_vec_kernel<float><<<grid, 256, 0, stream>>>((const float*)x, M, C, (float*)ws);
      RNN_CHECK_LAUNCH();
      colsum_final_kernel<float><<<ceil_div_i(C, 32), 128, 0, stream>>>((const float*)ws, rows, C, (float*)out, accumulate);
    } else {
      colsum_vec_kernel<bf16><<<grid, 256, 0, stream>>>((const bf16*)x, M, C, (float*)ws);
      RNN_CHECK_LAUNCH();
      colsum_final_kernel<bf16><<<ceil_div_i(C, 32), 128, 0, stream>>>((const float*)ws, rows, C, (bf16*)out, accumulate);
    }
    RNN_CHECK_LAUNCH();
    return RNN_OK;
  }
  const int rows = colsum_rows(M);
  RNN_REQUIRE(ws != nullptr && ws_bytes >= (size_t)rows * C * sizeof(float), RNN_ERR_WORKSPACE);
  const dim3 grid(ceil_div_i(C, 256), rows);
  if (is_f32) {
    colsum_partial_kernel<float><<<grid, 256, 0, stream>>>((const float*)x, M, C, (float*)ws);
    RNN_CHECK_LAUNCH();
    colsum_final_kernel<float><<<ceil_div_i(C, 32), 128, 0, stream>>>((const float*)ws, rows, C, (float*)out, accumulate);
  } else {
    colsum_partial_kernel<bf16><<<grid, 256, 0, stream>>>((const bf16*)x, M, C, (float*)ws);
    RNN_CHECK_LAUNCH();
    colsum_final_kernel<bf16><<<ceil_div_i(C, 32), 128, 0, stream>>>((const float*)ws, rows, C, (bf16*)out, accumulate);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}


/* dx = dy * (y > 0) and dbias[c] (+)= sum_r dx[r, c] in one pass (Conv2d / Linear backward with a fused ReLU).
 * C must be a multiple of the 16-byte vector (8 bf16 / 4 fp32); ws: rnn_colsum_workspace(M, C). */
int rnn_relu_bwd_colsum(const void* dy, const void* y, void* dx, long long M, int C, void* dbias, int accumulate,
                        int is_f32, void* ws, size_t ws_bytes, cudaStream_t stream) {
  const int W = is_f32 ? 4 : 8;
  RNN_REQUIRE(M > 0 && C > 0 && C % W == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(dx)) & 15u) == 0,
              RNN_ERR_ALIGN);
  const int rows = colsum_vec_rows(M, C, W);
  RNN_REQUIRE(ws != nullptr && ws_bytes >= (size_t)rows * C * sizeof(float), RNN_ERR_WORKSPACE);
  const int cgs = C / W, tpr = cgs < 256 ? cgs : 256;
  const dim3 grid((cgs + tpr - 1) / tpr, rows);
  if (is_f32) {
    relu_bwd_colsum_kernel<float><<<grid, 256, 0, stream>>>((const float*)dy, (const float*)y, (float*)dx, M, C, (float*)ws);
    RNN_CHECK_LAUNCH();
    colsum_final_kernel<float><<<ceil_div_i(C, 32), 128, 0, stream>>>((const float*)ws, rows, C, (float*)dbias, accumulate);
  } else {
    relu_bwd_colsum_kernel<bf16><<<grid, 256, 0, stream>>>((const bf16*)dy, (const bf16*)y, (bf16*)dx, M, C, (float*)ws);
    RNN_CHECK_LAUNCH();
    colsum_final_kernel<bf16><<<ceil_div_i(C, 32), 128, 0, stream>>>((const float*)ws, rows, C, (bf16*)dbias, accumulate);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
}  // extern "C"
