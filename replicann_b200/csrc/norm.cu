// Normalisation kernels (HBM-bound; 128-bit accesses, warp-shuffle / shared-memory reductions).
//
// BatchNorm2d over NHWC bf16 (torch.nn.BatchNorm2d semantics, training mode):
//   stats   : per-channel sum / sum of squares  -> partials in workspace -> finalize (mean, invstd,
//             scale = gamma*invstd, shift = beta - mean*scale, running stats with unbiased variance)
//   apply   : y = relu(x*scale + shift + residual)            (fused BN-apply + add + ReLU)
//   backward: reduce  (sum dy', sum dy'*xhat with dy' = dy * [y > 0])  -> partials -> finalize
//             apply   dx = scale * (dy' - sum_dy/M - xhat * sum_dy_xhat/M) ; dres = dy'
//
// LayerNorm over the last dimension (reference: nn.LayerNorm at arch/transformer.py:65,72,188 with
// the residual add of :111-112,115-116,149-150,198-200 fused in): one warp per row.
#include "common.cuh"
#include <stdlib.h>

#include <atomic>

namespace {

// --------------------------------------------------------------------------------- BatchNorm
// Thread layout for an [M, C] bf16 matrix: each thread owns one 8-channel vector column (cg) and
// walks rows r0, r0 + rows_per_pass, ...   blockDim.x = 256; cgs = C/8 column groups.
template <bool BWD>
__device__ __forceinline__ void bn_reduce_body(const bf16* __restrict__ x, const bf16* __restrict__ dy,
                                               const bf16* __restrict__ yout, const float* __restrict__ mean,
                                               const float* __restrict__ invstd, long long M, int C,
                                               float* __restrict__ partials /* [nby][2][C] */,
                                               const float* __restrict__ fscale, const float* __restrict__ fshift,
                                               const unsigned char* __restrict__ bitmask, const int bx, const int by,
                                               const int nby) {
  // grid.x tiles the column groups (when C/8 > 256), grid.y tiles rows.  Each thread owns one 8-channel
  // vector column and walks its rows four at a time so that 4 (x streams) independent 16-byte loads are
  // in flight per thread.  In the backward flavour the ReLU mask comes from the saved output `yout`, or is
  // recomputed from x (fscale/fshift = the forward's scale/shift) when the forward had no residual.
  constexpr int U = 4;
  const int cgs = C >> 3;
  const int tpr = cgs < 256 ? cgs : 256;          // threads per row
  const int rpp = 256 / tpr;                      // rows per pass
  const int cg = bx * tpr + (threadIdx.x % tpr);
  const int rlocal = threadIdx.x / tpr;
  float s0[8], s1[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) s0[e] = s1[e] = 0.f;
  float mu[8], is[8], fs[8], fb[8];
  const bool active = cg < cgs && rlocal < rpp;
  if (BWD && active) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      mu[e] = mean[cg * 8 + e];
      is[e] = invstd[cg * 8 + e];
      fs[e] = fscale ? fscale[cg * 8 + e] : 0.f;
      fb[e] = fscale ? fshift[cg * 8 + e] : 0.f;
    }
  }
  auto accum = [&](const uint4& ux, const uint4& ug, const uint4& uo, unsigned mbits) {
    float v[8];
    unpack8(ux, v);
    if (!BWD) {
#pragma unroll
      for (int e = 0; e < 8; ++e) { s0[e] += v[e]; s1[e] = fmaf(v[e], v[e], s1[e]); }
    } else {
      float g[8];
      unpack8(ug, g);
      if (yout) {
        float o[8];
        unpack8(uo, o);
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = o[e] > 0.f ? g[e] : 0.f;
      } else if (fscale) {
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = fmaf(v[e], fs[e], fb[e]) > 0.f ? g[e] : 0.f;
      } else if (bitmask) {
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = ((mbits >> e) & 1u) ? g[e] : 0.f;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) { s0[e] += g[e]; s1[e] = fmaf(g[e], (v[e] - mu[e]) * is[e], s1[e]); }
    }
  };
  if (active) {
    const long long stride = (long long)nby * rpp;
    long long r = (long long)by * rpp + rlocal;
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (; r + (U - 1) * stride < M; r += U * stride) {
      uint4 ux[U], ug[U], uo[U];
      unsigned mb[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long off = (r + u * stride) * C + cg * 8;
        ux[u] = ld_stream(x + off);
        ug[u] = BWD ? ld_stream(dy + off) : z;
        uo[u] = (BWD && yout) ? ld_stream(yout + off) : z;
        mb[u] = (BWD && bitmask) ? bitmask[(r + u * stride) * cgs + cg] : 0u;
      }
#pragma unroll
      for (int u = 0; u < U; ++u) accum(ux[u], ug[u], uo[u], mb[u]);
    }
    for (; r < M; r += stride) {
      const long long off = r * C + cg * 8;
      accum(ld_stream(x + off), BWD ? ld_stream(dy + off) : z, (BWD && yout) ? ld_stream(yout + off) : z,
            (BWD && bitmask) ? bitmask[r * cgs + cg] : 0u);
    }
  }
  // combine the rpp row-lanes of each column group through shared memory
  __shared__ float sm[256 * 16];
#pragma unroll
  for (int e = 0; e < 8; ++e) { sm[threadIdx.x * 16 + e] = s0[e]; sm[threadIdx.x * 16 + 8 + e] = s1[e]; }
  __syncthreads();
  if (rlocal == 0 && cg < cgs) {
    for (int rr = 1; rr < rpp; ++rr) {
      const int t = rr * tpr + (threadIdx.x % tpr);
#pragma unroll
      for (int e = 0; e < 8; ++e) { s0[e] += sm[t * 16 + e]; s1[e] += sm[t * 16 + 8 + e]; }
    }
    float* p = partials + (long long)by * 2 * C;
#pragma unroll
    for (int e = 0; e < 8; ++e) { p[cg * 8 + e] = s0[e]; p[C + cg * 8 + e] = s1[e]; }
  }
}

template <bool BWD>
__global__ void __launch_bounds__(256)
bn_reduce_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, const bf16* __restrict__ yout,
                 const float* __restrict__ mean, const float* __restrict__ invstd, long long M, int C,
                 float* __restrict__ partials /* [gridDim.y][2][C] */, const float* __restrict__ fscale = nullptr,
                 const float* __restrict__ fshift = nullptr, const unsigned char* __restrict__ bitmask = nullptr) {
  bn_reduce_body<BWD>(x, dy, yout, mean, invstd, M, C, partials, fscale, fshift, bitmask, blockIdx.x, blockIdx.y,
                      gridDim.y);
}

// Sum the per-block partials.  A 1024-thread block owns 32 channels: warp w walks partial rows w, w+32, ... with
// its 32 lanes on 32 consecutive channels (one coalesced 128-byte load per row and quantity), then lane-columns
// are combined through shared memory in a fixed order (deterministic).  With ~600 partial rows every thread has
// all of its ~19 loads in flight at once, so the kernel costs about two L2 round trips instead of a chain of them.
// Partial layout: [C / blk][nparts][2][blk] -- blk == C for the row-block partials of bn_reduce_kernel, blk = the
// N tile of the convolution when the statistics come out of its epilogue (one partial row per epilogue warp).
// Returns true in the threads of warp 0 that hold a finished channel (c < C); s / ss are valid only there.
constexpr int FIN_THREADS = 1024, FIN_CH = 32;
__device__ __forceinline__ bool sum_partials_block(const float* __restrict__ partials, int nparts, int C, int c,
                                                   float& s, float& ss, int blk) {
  __shared__ float sm[2][32][FIN_CH + 1];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float a = 0.f, b = 0.f;
  if (c < C) {
    const float* base = partials + (long long)(c / blk) * nparts * 2 * blk + (c % blk);
    const long long rs = 2ll * blk;
    int p = w;
    for (; p + 96 < nparts; p += 128) {
      const float a0 = base[p * rs], b0 = base[p * rs + blk];
      const float a1 = base[(p + 32) * rs], b1 = base[(p + 32) * rs + blk];
      const float a2 = base[(p + 64) * rs], b2 = base[(p + 64) * rs + blk];
      const float a3 = base[(p + 96) * rs], b3 = base[(p + 96) * rs + blk];
      a += (a0 + a1) + (a2 + a3);
      b += (b0 + b1) + (b2 + b3);
    }
    for (; p < nparts; p += 32) {
      a += base[p * rs];
      b += base[p * rs + blk];
    }
  }
  sm[0][w][lane] = a;
  sm[1][w][lane] = b;
  __syncthreads();
  if (w != 0) return false;
  a = 0.f; b = 0.f;
#pragma unroll
  for (int i = 0; i < 32; ++i) { a += sm[0][i][lane]; b += sm[1][i][lane]; }
  s = a; ss = b;
  return c < C;
}

__global__ void __launch_bounds__(FIN_THREADS)
bn_stats_finalize_kernel(const float* __restrict__ partials, int nparts, int C, float count,
                         const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                         float momentum, float* running_mean, float* running_var,
                         float* __restrict__ mean, float* __restrict__ invstd,
                         float* __restrict__ scale, float* __restrict__ shift, int blk) {
  const int c = blockIdx.x * FIN_CH + (threadIdx.x & 31);
  float s, ss;
  if (!sum_partials_block(partials, nparts, C, c, s, ss, blk)) return;
  const float m = s / count;
  float var = ss / count - m * m;
  var = var < 0.f ? 0.f : var;
  const float is = rsqrtf(var + eps);
  mean[c] = m;
  invstd[c] = is;
  const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  scale[c] = g * is;
  shift[c] = b - m * g * is;
  if (running_mean) {
    const float unb = count > 1.f ? var * count / (count - 1.f) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unb;
  }
}

// eval mode: scale/shift from running statistics
__global__ void bn_eval_prepare_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                                       const float* __restrict__ rm, const float* __restrict__ rv, float eps, int C,
                                       float* __restrict__ scale, float* __restrict__ shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float is = rsqrtf(rv[c] + eps);
  const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  scale[c] = g * is;
  shift[c] = b - rm[c] * g * is;
}

__global__ void __launch_bounds__(FIN_THREADS)
bn_bwd_finalize_kernel(const float* __restrict__ partials, int nparts, int C,
                       float* __restrict__ sum_dy, float* __restrict__ sum_dy_xhat,
                       void* dgamma, void* dbeta, int grads_f32, int accumulate,
                       const float* __restrict__ mean, const float* __restrict__ invstd,
                       const float* __restrict__ gamma, float inv_count, float* __restrict__ coef /* [3][C] */) {
  const int c = blockIdx.x * FIN_CH + (threadIdx.x & 31);
  float s, ss;
  if (!sum_partials_block(partials, nparts, C, c, s, ss, C)) return;
  sum_dy[c] = s;
  sum_dy_xhat[c] = ss;
  {
    // dx = gamma*invstd*(dy' - sum_dy/M - xhat*sum_dy_xhat/M), xhat = (x-mean)*invstd
    //    = A*dy' + B*x + D  with per-channel A, B, D: the apply kernel needs three vector loads per 8 channels
    const float is = invstd[c], gm = gamma ? gamma[c] : 1.f;
    const float A = gm * is;
    const float B = -gm * is * is * ss * inv_count;
    coef[c] = A;
    coef[C + c] = B;
    coef[2 * C + c] = -A * s * inv_count - B * mean[c];
  }
  if (dgamma) {
    if (grads_f32) {
      float* dg = (float*)dgamma; float* db = (float*)dbeta;
      dg[c] = (accumulate ? dg[c] : 0.f) + ss;
      db[c] = (accumulate ? db[c] : 0.f) + s;
    } else {
      bf16* dg = (bf16*)dgamma; bf16* db = (bf16*)dbeta;
      dg[c] = __float2bfloat16((accumulate ? __bfloat162float(dg[c]) : 0.f) + ss);
      db[c] = __float2bfloat16((accumulate ? __bfloat162float(db[c]) : 0.f) + s);
    }
  }
}

// The streaming apply kernels run grid-stride with APPLY_UNROLL 16-byte vectors per thread per trip, all loads
// issued before any arithmetic (measured: 1 -> 2 vectors is worth 1.3 % of a ResNet-50 step).
#ifndef RNN_APPLY_UNROLL
#define RNN_APPLY_UNROLL 2
#endif
constexpr int APPLY_UNROLL = RNN_APPLY_UNROLL;
__global__ void __launch_bounds__(256)
bn_apply_kernel(const bf16* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                const bf16* __restrict__ residual, int relu, bf16* __restrict__ y, long long nvec, int cgs,
                unsigned char* __restrict__ relu_mask) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const uint4 z = make_uint4(0, 0, 0, 0);
  auto finish = [&](long long i, int cg, const uint4& ux, const uint4& ur) {
    float v[8];
    unpack8(ux, v);
    const float4 a0 = __ldg(reinterpret_cast<const float4*>(scale + cg * 8));
    const float4 a1 = __ldg(reinterpret_cast<const float4*>(scale + cg * 8 + 4));
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(shift + cg * 8));
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(shift + cg * 8 + 4));
    v[0] = fmaf(v[0], a0.x, b0.x); v[1] = fmaf(v[1], a0.y, b0.y); v[2] = fmaf(v[2], a0.z, b0.z); v[3] = fmaf(v[3], a0.w, b0.w);
    v[4] = fmaf(v[4], a1.x, b1.x); v[5] = fmaf(v[5], a1.y, b1.y); v[6] = fmaf(v[6], a1.z, b1.z); v[7] = fmaf(v[7], a1.w, b1.w);
    if (residual) {
      float r[8];
      unpack8(ur, r);
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] += r[e];
    }
    if (relu) {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
    }
    const uint4 packed = pack8(v);
    st_stream(y + i * 8, packed);
    if (relu_mask) {  // one bit per element: (stored bf16 output > 0), for the backward kernels
      float o[8];
      unpack8(packed, o);
      unsigned m = 0;
#pragma unroll
      for (int e = 0; e < 8; ++e) m |= (o[e] > 0.f ? 1u : 0u) << e;
      relu_mask[i] = (unsigned char)m;
    }
  };
  constexpr int U = APPLY_UNROLL;
  const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  // channel group of each vector without a 64-bit modulo per trip: it advances by (stride mod cgs), wrapped
  int cg = (int)(i0 % cgs);
  const int dcg = (int)(stride % cgs);
  for (long long i = i0; i < nvec; i += U * stride) {
    uint4 xs[U], rs[U];
    int cgu[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      xs[u] = j < nvec ? ld_stream(x + j * 8) : z;
      rs[u] = (j < nvec && residual) ? ld_stream(residual + j * 8) : z;
      cgu[u] = cg;
      cg += dcg; cg -= cg >= cgs ? cgs : 0;
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < nvec) finish(i + u * stride, cgu[u], xs[u], rs[u]);
  }
}

template <int U>
__device__ __forceinline__ void bn_bwd_apply_body(const bf16* __restrict__ dy, const bf16* __restrict__ x,
                                                  const bf16* __restrict__ yout,
                                                  const float* __restrict__ coef /* [3][C]: A, B, D */,
                                                  bf16* __restrict__ dx, bf16* __restrict__ dres, long long nvec, int cgs,
                                                  const float* __restrict__ fscale, const float* __restrict__ fshift,
                                                  const unsigned char* __restrict__ bitmask, const int bx, const int nbx) {
  const int C = cgs * 8;
  const long long stride = (long long)nbx * blockDim.x;
  const uint4 z = make_uint4(0, 0, 0, 0);
  auto finish = [&](long long i, int cg, const uint4& ug, const uint4& ux, const uint4& uo, unsigned mbits) {
    const int c0 = cg * 8;
    float g[8], v[8];
    unpack8(ug, g);
    unpack8(ux, v);
    if (yout) {
      float o[8];
      unpack8(uo, o);
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] = o[e] > 0.f ? g[e] : 0.f;
    } else if (fscale) {
      float fs[8], fb[8];
      *reinterpret_cast<float4*>(fs) = __ldg(reinterpret_cast<const float4*>(fscale + c0));
      *reinterpret_cast<float4*>(fs + 4) = __ldg(reinterpret_cast<const float4*>(fscale + c0 + 4));
      *reinterpret_cast<float4*>(fb) = __ldg(reinterpret_cast<const float4*>(fshift + c0));
      *reinterpret_cast<float4*>(fb + 4) = __ldg(reinterpret_cast<const float4*>(fshift + c0 + 4));
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] = fmaf(v[e], fs[e], fb[e]) > 0.f ? g[e] : 0.f;
    } else if (bitmask) {
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] = ((mbits >> e) & 1u) ? g[e] : 0.f;
    }
    if (dres) st_stream(dres + i * 8, pack8(g));
    float a[8], b[8], d[8], r[8];
    *reinterpret_cast<float4*>(a) = __ldg(reinterpret_cast<const float4*>(coef + c0));
    *reinterpret_cast<float4*>(a + 4) = __ldg(reinterpret_cast<const float4*>(coef + c0 + 4));
    *reinterpret_cast<float4*>(b) = __ldg(reinterpret_cast<const float4*>(coef + C + c0));
    *reinterpret_cast<float4*>(b + 4) = __ldg(reinterpret_cast<const float4*>(coef + C + c0 + 4));
    *reinterpret_cast<float4*>(d) = __ldg(reinterpret_cast<const float4*>(coef + 2 * C + c0));
    *reinterpret_cast<float4*>(d + 4) = __ldg(reinterpret_cast<const float4*>(coef + 2 * C + c0 + 4));
#pragma unroll
    for (int e = 0; e < 8; ++e) r[e] = fmaf(a[e], g[e], fmaf(b[e], v[e], d[e]));
    st_stream(dx + i * 8, pack8(r));
  };
  const long long i0 = (long long)bx * blockDim.x + threadIdx.x;
  int cg = (int)(i0 % cgs);
  const int dcg = (int)(stride % cgs);
  for (long long i = i0; i < nvec; i += U * stride) {
    uint4 gs[U], xs[U], os[U];
    unsigned ms[U];
    int cgu[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      const bool ok = j < nvec;
      gs[u] = ok ? ld_stream(dy + j * 8) : z;
      xs[u] = ok ? ld_stream(x + j * 8) : z;
      os[u] = (ok && yout) ? ld_stream(yout + j * 8) : z;
      ms[u] = (ok && bitmask) ? bitmask[j] : 0u;
      cgu[u] = cg;
      cg += dcg; cg -= cg >= cgs ? cgs : 0;
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < nvec) finish(i + u * stride, cgu[u], gs[u], xs[u], os[u], ms[u]);
  }
}

__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, const bf16* __restrict__ yout,
                    const float* __restrict__ coef, bf16* __restrict__ dx, bf16* __restrict__ dres, long long nvec,
                    int cgs, const float* __restrict__ fscale, const float* __restrict__ fshift,
                    const unsigned char* __restrict__ bitmask) {
  bn_bwd_apply_body<APPLY_UNROLL>(dy, x, yout, coef, dx, dres, nvec, cgs, fscale, fshift, bitmask, blockIdx.x,
                                  gridDim.x);
}

// ---- apply kernels with the per-channel constants in registers ------------------------------------------------------
// ncu on the kernels above (round 1 / round 2) showed L1, not DRAM, as their busiest unit (l1tex 62 %, DRAM 24-60 %):
// every 16-byte vector of payload came with 4 (forward) or 10 (backward) 16-byte loads of per-channel constants, and
// the asymptotic rate stayed at 4.0-4.8 TB/s.  When the grid stride is a multiple of the number of channel groups
// (it is for every power-of-two channel count) a thread sees ONE channel group for its whole life, so the constants
// are loaded once into registers and the loop touches only the tensors.  One wave of blocks (grid = SMs x resident
// blocks), four vectors per stream in flight per thread.
constexpr int HOIST_U = 4;

__global__ void __launch_bounds__(256, 2)
bn_apply_hoisted_kernel(const bf16* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                        const bf16* __restrict__ residual, int relu, bf16* __restrict__ y, long long nvec, int cgs,
                        unsigned char* __restrict__ relu_mask) {
  constexpr int U = HOIST_U;
  const long long stride = (long long)gridDim.x * blockDim.x;      // multiple of cgs (checked by the launcher)
  const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cg = (int)(i0 % cgs);
  float sc[8], sh[8];
  *reinterpret_cast<float4*>(sc) = __ldg(reinterpret_cast<const float4*>(scale + cg * 8));
  *reinterpret_cast<float4*>(sc + 4) = __ldg(reinterpret_cast<const float4*>(scale + cg * 8 + 4));
  *reinterpret_cast<float4*>(sh) = __ldg(reinterpret_cast<const float4*>(shift + cg * 8));
  *reinterpret_cast<float4*>(sh + 4) = __ldg(reinterpret_cast<const float4*>(shift + cg * 8 + 4));
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (long long i = i0; i < nvec; i += U * stride) {
    uint4 xs[U], rs[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      xs[u] = j < nvec ? ld_stream(x + j * 8) : z;
      rs[u] = (j < nvec && residual) ? ld_stream(residual + j * 8) : z;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      if (j >= nvec) break;
      float v[8];
      unpack8(xs[u], v);
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = fmaf(v[e], sc[e], sh[e]);
      if (residual) {
        float r[8];
        unpack8(rs[u], r);
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] += r[e];
      }
      if (relu) {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
      }
      const uint4 packed = pack8(v);
      st_stream(y + j * 8, packed);
      if (relu_mask) {
        float o[8];
        unpack8(packed, o);
        unsigned m = 0;
#pragma unroll
        for (int e = 0; e < 8; ++e) m |= (o[e] > 0.f ? 1u : 0u) << e;
        relu_mask[j] = (unsigned char)m;
      }
    }
  }
}

// MASK: 0 none, 1 recompute from x with the forward's scale / shift, 2 bit mask, 3 saved output
template <int MASK>
__global__ void __launch_bounds__(256, 2)
bn_bwd_apply_hoisted_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, const bf16* __restrict__ yout,
                            const float* __restrict__ coef /* [3][C]: A, B, D */, bf16* __restrict__ dx,
                            bf16* __restrict__ dres, long long nvec, int cgs, const float* __restrict__ fscale,
                            const float* __restrict__ fshift, const unsigned char* __restrict__ bitmask) {
  constexpr int U = HOIST_U;
  const int C = cgs * 8;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = (int)(i0 % cgs) * 8;
  float a[8], b[8], d[8], fs[8], fb[8];
  *reinterpret_cast<float4*>(a) = __ldg(reinterpret_cast<const float4*>(coef + c0));
  *reinterpret_cast<float4*>(a + 4) = __ldg(reinterpret_cast<const float4*>(coef + c0 + 4));
  *reinterpret_cast<float4*>(b) = __ldg(reinterpret_cast<const float4*>(coef + C + c0));
  *reinterpret_cast<float4*>(b + 4) = __ldg(reinterpret_cast<const float4*>(coef + C + c0 + 4));
  *reinterpret_cast<float4*>(d) = __ldg(reinterpret_cast<const float4*>(coef + 2 * C + c0));
  *reinterpret_cast<float4*>(d + 4) = __ldg(reinterpret_cast<const float4*>(coef + 2 * C + c0 + 4));
  if (MASK == 1) {
    *reinterpret_cast<float4*>(fs) = __ldg(reinterpret_cast<const float4*>(fscale + c0));
    *reinterpret_cast<float4*>(fs + 4) = __ldg(reinterpret_cast<const float4*>(fscale + c0 + 4));
    *reinterpret_cast<float4*>(fb) = __ldg(reinterpret_cast<const float4*>(fshift + c0));
    *reinterpret_cast<float4*>(fb + 4) = __ldg(reinterpret_cast<const float4*>(fshift + c0 + 4));
  }
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (long long i = i0; i < nvec; i += U * stride) {
    uint4 gs[U], xs[U], os[MASK == 3 ? U : 1];
    unsigned ms[MASK == 2 ? U : 1];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      const bool ok = j < nvec;
      gs[u] = ok ? ld_stream(dy + j * 8) : z;
      xs[u] = ok ? ld_stream(x + j * 8) : z;
      if (MASK == 3) os[u] = ok ? ld_stream(yout + j * 8) : z;
      if (MASK == 2) ms[u] = ok ? bitmask[j] : 0u;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      if (j >= nvec) break;
      float g[8], v[8], r[8];
      unpack8(gs[u], g);
      unpack8(xs[u], v);
      if (MASK == 3) {
        float o[8];
        unpack8(os[u], o);
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = o[e] > 0.f ? g[e] : 0.f;
      } else if (MASK == 1) {
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = fmaf(v[e], fs[e], fb[e]) > 0.f ? g[e] : 0.f;
      } else if (MASK == 2) {
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = ((ms[u] >> e) & 1u) ? g[e] : 0.f;
      }
      if (dres) st_stream(dres + j * 8, pack8(g));
#pragma unroll
      for (int e = 0; e < 8; ++e) r[e] = fmaf(a[e], g[e], fmaf(b[e], v[e], d[e]));
      st_stream(dx + j * 8, pack8(r));
    }
  }
}

// grid of the hoisted kernels: one wave, and a stride that is a multiple of the channel groups (0 = not applicable)
template <typename K>
int hoisted_grid(K kernel, long long nvec, int cgs) {
  if (256 % cgs != 0 && cgs % 256 != 0) return 0;
  static int per_sm_cache = 0;
  if (per_sm_cache == 0) {
    int v = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, kernel, 256, 0) != cudaSuccess || v < 1) v = 2;
    per_sm_cache = v;
  }
  long long blocks = (nvec + 256ll * HOIST_U - 1) / (256ll * HOIST_U);
  const long long cap = (long long)rnn_num_sms() * per_sm_cache;
  if (blocks > cap) blocks = cap;
  if (cgs > 256) {   // the stride (blocks * 256) must be a multiple of cgs
    const int m = cgs / 256;
    blocks = blocks / m * m;
  }
  return blocks < 1 ? 0 : (int)blocks;
}

// CTAs per SM of the streaming apply kernels (grid-stride loops).  Experiments: RNN_BN_APPLY_CTAS.
int apply_ctas_per_sm() {
  static const int v = [] {
    const char* e = getenv("RNN_BN_APPLY_CTAS");
    const int n = e ? atoi(e) : 0;
    return n > 0 ? n : 16;
  }();
  return v;
}

// RNN_BN_HOIST=0: the apply kernels that reload their constants per vector (A/B timing; any channel count)
bool apply_hoist_enabled() {
  static const bool v = [] {
    const char* e = getenv("RNN_BN_HOIST");
    return e == nullptr || atoi(e) != 0;
  }();
  return v;
}

int bn_row_blocks(long long M, int C) {
  const int cgs = C >> 3;
  const int tpr = cgs < 256 ? cgs : 256;
  const int rpp = 256 / tpr;
  const int gx = (cgs + tpr - 1) / tpr;
  // every thread should see >= 4 rows (one unrolled batch)
  long long gy = (M + (long long)rpp * 4 - 1) / ((long long)rpp * 4);
  static const int per_sm = [] {
    const char* e = getenv("RNN_BN_REDUCE_CTAS");
    const int n = e ? atoi(e) : 0;
    return n > 0 ? n : 2;
  }();
  // 124 registers per thread: two blocks are resident per SM, so 2 per SM is ONE wave.  A second wave cost 4 us on
  // every tensor size (bench_bn_bwd.py, round 2): the kernel is a single latency-bound sweep per thread.
  long long cap = (long long)rnn_num_sms() * per_sm / gx;
  if (cap < 1) cap = 1;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  return (int)gy;
}

// --------------------------------------------------------------------------------- LayerNorm
// One warp per row; E % 8 == 0 (bf16) / E % 4 == 0 (fp32).  y = LN(x + res) * gamma + beta.
template <typename T>
struct Vec;
template <>
struct Vec<bf16> {
  static constexpr int W = 8;
  __device__ static void load(const bf16* p, float* f) { unpack8(*reinterpret_cast<const uint4*>(p), f); }
  __device__ static void store(bf16* p, const float* f) { *reinterpret_cast<uint4*>(p) = pack8(f); }
};
template <>
struct Vec<float> {
  static constexpr int W = 4;
  __device__ static void load(const float* p, float* f) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
  __device__ static void store(float* p, const float* f) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  }
};

template <typename T>
__global__ void __launch_bounds__(256)
layernorm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res, const T* __restrict__ gamma,
                     const T* __restrict__ beta, float eps, T* __restrict__ y, T* __restrict__ xsum,
                     float* __restrict__ mean, float* __restrict__ rstd, long long M, int E) {
  constexpr int W = Vec<T>::W;
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  const T* xr = x + row * E;
  const T* rr = res ? res + row * E : nullptr;
  // pass 1: sum (and keep x+res if requested so the backward pass can reread it)
  float s = 0.f;
  for (int c = lane * W; c < E; c += 32 * W) {
    float v[W];
    Vec<T>::load(xr + c, v);
    if (rr) {
      float r[W];
      Vec<T>::load(rr + c, r);
#pragma unroll
      for (int e = 0; e < W; ++e) v[e] += r[e];
      if (xsum) {
        Vec<T>::store(xsum + row * E + c, v);
      }
    }
#pragma unroll
    for (int e = 0; e < W; ++e) s += v[e];
  }
  s = warp_sum(s);
  const float mu = s / E;
  // pass 2: variance about the mean (two-pass, as torch does)
  const T* src = (rr && xsum) ? xsum + row * E : xr;
  const bool readd = rr && !xsum;
  float q = 0.f;
  for (int c = lane * W; c < E; c += 32 * W) {
    float v[W];
    Vec<T>::load(src + c, v);
    if (readd) {
      float r[W];
      Vec<T>::load(rr + c, r);
#pragma unroll
      for (int e = 0; e < W; ++e) v[e] += r[e];
    }
#pragma unroll
    for (int e = 0; e < W; ++e) { const float d = v[e] - mu; q += d * d; }
  }
  q = warp_sum(q);
  const float rs = rsqrtf(q / E + eps);
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
  for (int c = lane * W; c < E; c += 32 * W) {
    float v[W], g[W], b[W];
    Vec<T>::load(src + c, v);
    if (readd) {
      float r[W];
      Vec<T>::load(rr + c, r);
#pragma unroll
      for (int e = 0; e < W; ++e) v[e] += r[e];
    }
    Vec<T>::load(gamma + c, g);
    Vec<T>::load(beta + c, b);
#pragma unroll
    for (int e = 0; e < W; ++e) v[e] = (v[e] - mu) * rs * g[e] + b[e];
    Vec<T>::store(y + row * E + c, v);
  }
}

// dx = rstd * (g - mean(g) - xhat*mean(g*xhat)), g = dy*gamma.  Per-block partial dgamma/dbeta rows are
// written to `partials` [gridDim.x][2][E] and summed by layernorm_bwd_finalize_kernel.
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ xin, const T* __restrict__ gamma,
                     const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                     float* __restrict__ partials, long long M, int E) {
  constexpr int W = Vec<T>::W;
  extern __shared__ float acc[];  // [2][E] block accumulators
  for (int i = threadIdx.x; i < 2 * E; i += blockDim.x) acc[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (long long row = (long long)blockIdx.x * nw + wid; row < M; row += (long long)gridDim.x * nw) {
    const float mu = mean[row], rs = rstd[row];
    float sg = 0.f, sgx = 0.f;
    for (int c = lane * W; c < E; c += 32 * W) {
      float d[W], v[W], g[W];
      Vec<T>::load(dy + row * E + c, d);
      Vec<T>::load(xin + row * E + c, v);
      Vec<T>::load(gamma + c, g);
#pragma unroll
      for (int e = 0; e < W; ++e) {
        const float xh = (v[e] - mu) * rs;
        const float gg = d[e] * g[e];
        sg += gg;
        sgx += gg * xh;
        atomicAdd(&acc[c + e], d[e] * xh);
        atomicAdd(&acc[E + c + e], d[e]);
      }
    }
    sg = warp_sum(sg) / E;
    sgx = warp_sum(sgx) / E;
    for (int c = lane * W; c < E; c += 32 * W) {
      float d[W], v[W], g[W], o[W];
      Vec<T>::load(dy + row * E + c, d);
      Vec<T>::load(xin + row * E + c, v);
      Vec<T>::load(gamma + c, g);
#pragma unroll
      for (int e = 0; e < W; ++e) {
        const float xh = (v[e] - mu) * rs;
        o[e] = rs * (d[e] * g[e] - sg - xh * sgx);
      }
      Vec<T>::store(dx + row * E + c, o);
    }
  }
  __syncthreads();
  float* p = partials + (long long)blockIdx.x * 2 * E;
  for (int i = threadIdx.x; i < 2 * E; i += blockDim.x) p[i] = acc[i];
}

// Same computation for rows of at most 32 * W * NV elements: each lane owns NV fixed W-wide column groups, keeps
// the row it is working on in registers (one global read of dy and x instead of two) and accumulates its share of
// dgamma / dbeta in registers across all of its rows; shared memory sees one add per column per WARP at the end
// instead of two atomics per element per row.
template <typename T, int NV>
__global__ void __launch_bounds__(256)
layernorm_bwd_reg_kernel(const T* __restrict__ dy, const T* __restrict__ xin, const T* __restrict__ gamma,
                         const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                         float* __restrict__ partials, long long M, int E) {
  constexpr int W = Vec<T>::W;
  extern __shared__ float acc[];  // [2][E] block accumulators
  for (int i = threadIdx.x; i < 2 * E; i += blockDim.x) acc[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float gam[NV][W], dgam[NV][W], dbet[NV][W];
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int c = (k * 32 + lane) * W;
#pragma unroll
    for (int e = 0; e < W; ++e) { gam[k][e] = 0.f; dgam[k][e] = 0.f; dbet[k][e] = 0.f; }
    if (c < E) Vec<T>::load(gamma + c, gam[k]);
  }
  for (long long row = (long long)blockIdx.x * nw + wid; row < M; row += (long long)gridDim.x * nw) {
    const float mu = mean[row], rs = rstd[row];
    float d[NV][W], xh[NV][W];
    float sg = 0.f, sgx = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * W;
      if (c < E) {
        float v[W];
        Vec<T>::load(dy + row * E + c, d[k]);
        Vec<T>::load(xin + row * E + c, v);
#pragma unroll
        for (int e = 0; e < W; ++e) {
          xh[k][e] = (v[e] - mu) * rs;
          const float gg = d[k][e] * gam[k][e];
          sg += gg;
          sgx += gg * xh[k][e];
          dgam[k][e] += d[k][e] * xh[k][e];
          dbet[k][e] += d[k][e];
        }
      }
    }
    sg = warp_sum(sg) / E;
    sgx = warp_sum(sgx) / E;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * W;
      if (c < E) {
        float o[W];
#pragma unroll
        for (int e = 0; e < W; ++e) o[e] = rs * (d[k][e] * gam[k][e] - sg - xh[k][e] * sgx);
        Vec<T>::store(dx + row * E + c, o);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int c = (k * 32 + lane) * W;
    if (c < E) {
#pragma unroll
      for (int e = 0; e < W; ++e) { atomicAdd(&acc[c + e], dgam[k][e]); atomicAdd(&acc[E + c + e], dbet[k][e]); }
    }
  }
  __syncthreads();
  float* p = partials + (long long)blockIdx.x * 2 * E;
  for (int i = threadIdx.x; i < 2 * E; i += blockDim.x) p[i] = acc[i];
}

template <typename T>
__global__ void __launch_bounds__(FIN_THREADS)
layernorm_bwd_finalize_kernel(const float* __restrict__ partials, int nparts, int E, T* dgamma, T* dbeta,
                              int accumulate) {
  const int c = blockIdx.x * FIN_CH + (threadIdx.x & 31);
  float a, b;
  if (!sum_partials_block(partials, nparts, E, c, a, b, E)) return;   // partials: [nparts][2][E]
  if (accumulate) { a += (float)dgamma[c]; b += (float)dbeta[c]; }
  dgamma[c] = (T)a;
  dbeta[c] = (T)b;
}

int ln_bwd_blocks(long long M) {
  long long b = (M + 31) / 32;   // 8 warps per block, >= 4 rows per warp: the per-block partial rows are the overhead
  const long long cap = (long long)rnn_num_sms() * 2;
  return (int)(b > cap ? cap : (b < 1 ? 1 : b));
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Built-in scratch for rnn_bn_stats_bf16_static (see include/rnn_b200.h).
//
// The array below is part of the library image: every GPU that loads librnn_b200.so carries its
// RNN_BN_STATIC_SCRATCH_BYTES until the process exits, whether or not the entry point is ever used.
// On the host side the identifier `g_bn_static_scratch` is only a handle for the runtime -- it is NOT a
// device pointer and must never be passed to a kernel as one.  The real address is looked up once per
// device with cudaGetSymbolAddress (with that device current) and cached here.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_bn_static_scratch[RNN_BN_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
// One slot per device ordinal.  Host threads may make their first call on a device at the same time: each then
// looks the address up for itself and stores the same value, so an atomic slot (no lock) is all that is needed.
// A cudaDeviceReset() destroys the context the address belongs to; the cached value is then stale, so do not
// reset a device between calls of rnn_bn_stats_bf16_static on it.
std::atomic<float*> g_bn_static_addr[kMaxDevices];

// Resolve (first use on this device) or fetch (later uses) the device address of the built-in scratch.
int bn_static_scratch_address(float** out) {
  int dev = -1;
  RNN_CUDA(cudaGetDevice(&dev));
  RNN_REQUIRE(dev >= 0 && dev < kMaxDevices, RNN_ERR_CUDA);
  float* addr = g_bn_static_addr[dev].load(std::memory_order_acquire);
  if (addr == nullptr) {
    void* p = nullptr;
    RNN_CUDA(cudaGetSymbolAddress(&p, g_bn_static_scratch));
    RNN_REQUIRE(p != nullptr, RNN_ERR_CUDA);
    addr = reinterpret_cast<float*>(p);
    g_bn_static_addr[dev].store(addr, std::memory_order_release);
  }
  *out = addr;
  return RNN_OK;
}
}  // namespace

// Finalize BatchNorm statistics from partial sums produced elsewhere (the convolution epilogue): used by
// tc_gemm.cu.  `blk` = columns per partial block (see sum_partials_block).
int rnn_internal_bn_finalize(const float* partials, int nparts, int C, int blk, float count, const float* gamma,
                             const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                             float* mean, float* invstd, float* scale, float* shift, cudaStream_t stream) {
  RnnTimedScope timed("bn_finalize", stream);
  bn_stats_finalize_kernel<<<ceil_div_i(C, FIN_CH), FIN_THREADS, 0, stream>>>(partials, nparts, C, count, gamma, beta, eps, momentum,
                                                                running_mean, running_var, mean, invstd, scale, shift, blk);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

extern "C" {

size_t rnn_bn_workspace(long long M, int C) { return (size_t)bn_row_blocks(M, C) * 2 * C * sizeof(float); }

int rnn_bn_stats_bf16(const void* x, long long M, int C, const float* gamma, const float* beta, float eps,
                      float momentum, float* running_mean, float* running_var, float* mean, float* invstd,
                      float* scale, float* shift, void* ws, size_t ws_bytes, cudaStream_t stream) {
  RNN_REQUIRE(M > 0 && C > 0 && C % 8 == 0, RNN_ERR_SHAPE);
  const int gy = bn_row_blocks(M, C);
  RNN_REQUIRE(ws != nullptr && ws_bytes >= (size_t)gy * 2 * C * sizeof(float), RNN_ERR_WORKSPACE);
  const int cgs = C >> 3, tpr = cgs < 256 ? cgs : 256, gx = (cgs + tpr - 1) / tpr;
  bn_reduce_kernel<false><<<dim3(gx, gy), 256, 0, stream>>>((const bf16*)x, nullptr, nullptr, nullptr, nullptr, M, C,
                                                           (float*)ws);
  RNN_CHECK_LAUNCH();
  bn_stats_finalize_kernel<<<ceil_div_i(C, FIN_CH), FIN_THREADS, 0, stream>>>((const float*)ws, gy, C, (float)M, gamma, beta, eps,
                                                                   momentum, running_mean, running_var, mean, invstd,
                                                                   scale, shift, C);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

// Same result as rnn_bn_stats_bf16, for callers that cannot provide (or allocate) a workspace: the
// per-block partial sums go to the library's built-in scratch array instead of a caller buffer.
//
//  * Nothing is allocated or freed here, and no allocating routine is called.
//  * The first call on a given device also looks up the scratch array's device address
//    (cudaGetSymbolAddress, device must be current).  Make that first call before starting any stream
//    capture on that device; later calls only launch kernels.
//  * Limit: C % 8 == 0 and C <= RNN_BN_STATIC_MAX_C.  Wider inputs are refused with RNN_ERR_WORKSPACE and
//    rnn_last_error() says why -- use rnn_bn_stats_bf16 with your own workspace for those.  M is
//    unlimited: the row range is simply cut into as many blocks as the scratch can hold partials for.
//  * All calls on a device share the one array and it keeps whatever the previous call left in it.  That
//    is harmless here because every partial-sum slot that the second kernel reads has just been
//    overwritten by the first, so nothing needs clearing -- but it also means two calls whose kernels can
//    overlap on the GPU (different streams, no ordering between them) corrupt each other.  Serialise them.
int rnn_bn_stats_bf16_static(const void* x, long long M, int C, const float* gamma, const float* beta, float eps,
                             float momentum, float* running_mean, float* running_var, float* mean, float* invstd,
                             float* scale, float* shift, cudaStream_t stream) {
  RNN_REQUIRE(M > 0 && C > 0 && C % 8 == 0, RNN_ERR_SHAPE);
  if (C > RNN_BN_STATIC_MAX_C) {
    rnn_set_last_message(__FILE__, __LINE__,
                         "rnn_bn_stats_bf16_static: C exceeds RNN_BN_STATIC_MAX_C (built-in scratch too small); "
                         "call rnn_bn_stats_bf16 with a caller workspace instead");
    return RNN_ERR_WORKSPACE;
  }
  float* scratch = nullptr;
  int rc = bn_static_scratch_address(&scratch);
  if (rc != RNN_OK) return rc;
  // as many row blocks as the regular entry point would use, capped by what fits in the built-in scratch
  const size_t per_block = (size_t)2 * C * sizeof(float);
  int gy = bn_row_blocks(M, C);
  const int fit = (int)(RNN_BN_STATIC_SCRATCH_BYTES / per_block);
  if (gy > fit) gy = fit;
  RNN_REQUIRE(gy >= 1, RNN_ERR_WORKSPACE);
  const int cgs = C >> 3, tpr = cgs < 256 ? cgs : 256, gx = (cgs + tpr - 1) / tpr;
  bn_reduce_kernel<false><<<dim3(gx, gy), 256, 0, stream>>>((const bf16*)x, nullptr, nullptr, nullptr, nullptr, M, C,
                                                           scratch);
  RNN_CHECK_LAUNCH();
  bn_stats_finalize_kernel<<<ceil_div_i(C, FIN_CH), FIN_THREADS, 0, stream>>>(scratch, gy, C, (float)M, gamma, beta, eps,
                                                                   momentum, running_mean, running_var, mean, invstd,
                                                                   scale, shift, C);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int rnn_bn_eval_prepare(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                        float eps, int C, float* scale, float* shift, cudaStream_t stream) {
  RNN_REQUIRE(C > 0, RNN_ERR_SHAPE);
  bn_eval_prepare_kernel<<<ceil_div_i(C, 128), 128, 0, stream>>>(gamma, beta, running_mean, running_var, eps, C, scale,
                                                                 shift);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int rnn_bn_apply_bf16(const void* x, const float* scale, const float* shift, const void* residual, int relu, void* y,
                      unsigned char* relu_mask, long long M, int C, cudaStream_t stream) {
  RNN_REQUIRE(M > 0 && C > 0 && C % 8 == 0, RNN_ERR_SHAPE);
  const long long nvec = M * (C / 8);
  if (const int hg = apply_hoist_enabled() ? hoisted_grid(bn_apply_hoisted_kernel, nvec, C / 8) : 0) {
    bn_apply_hoisted_kernel<<<hg, 256, 0, stream>>>((const bf16*)x, scale, shift, (const bf16*)residual, relu, (bf16*)y,
                                                    nvec, C / 8, relu ? relu_mask : nullptr);
    RNN_CHECK_LAUNCH();
    return RNN_OK;
  }
  long long blocks = (nvec + 255) / 256;
  const long long cap = (long long)rnn_num_sms() * apply_ctas_per_sm();
  if (blocks > cap) blocks = cap;
  bn_apply_kernel<<<(int)blocks, 256, 0, stream>>>((const bf16*)x, scale, shift, (const bf16*)residual, relu, (bf16*)y,
                                                   nvec, C / 8, relu ? relu_mask : nullptr);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int rnn_bn_bwd_bf16(const void* dy, const void* x, const void* y_out /* nullable: ReLU output */, const float* mean,
                    const float* invstd, const float* gamma, long long M, int C, void* dx,
                    void* dres /* nullable */, void* dgamma, void* dbeta, int grads_f32, int accumulate_grads,
                    float* sums /* [5][C] scratch: sum_dy, sum_dy_xhat, A, B, D */, const float* fwd_scale,
                    const float* fwd_shift, const unsigned char* relu_mask, void* ws, size_t ws_bytes,
                    cudaStream_t stream) {
  float* sum_dy = sums;
  float* sum_dy_xhat = sums + C;
  float* coef = sums + 2 * (long long)C;
  RNN_REQUIRE((fwd_scale == nullptr) == (fwd_shift == nullptr), RNN_ERR_SHAPE);
  RNN_REQUIRE((y_out != nullptr) + (fwd_scale != nullptr) + (relu_mask != nullptr) <= 1, RNN_ERR_SHAPE);  // one mask source
  RNN_REQUIRE(M > 0 && C > 0 && C % 8 == 0, RNN_ERR_SHAPE);
  const int gy = bn_row_blocks(M, C);
  RNN_REQUIRE(ws != nullptr && ws_bytes >= (size_t)gy * 2 * C * sizeof(float), RNN_ERR_WORKSPACE);
  const int cgs = C >> 3, tpr = cgs < 256 ? cgs : 256, gx = (cgs + tpr - 1) / tpr;
  {
    RnnTimedScope timed("bn_bwd_reduce", stream);
    bn_reduce_kernel<true><<<dim3(gx, gy), 256, 0, stream>>>((const bf16*)x, (const bf16*)dy, (const bf16*)y_out, mean,
                                                            invstd, M, C, (float*)ws, fwd_scale, fwd_shift, relu_mask);
    RNN_CHECK_LAUNCH();
  }
  {
    RnnTimedScope timed("bn_bwd_finalize", stream);
    bn_bwd_finalize_kernel<<<ceil_div_i(C, FIN_CH), FIN_THREADS, 0, stream>>>((const float*)ws, gy, C, sum_dy, sum_dy_xhat, dgamma,
                                                                  dbeta, grads_f32, accumulate_grads, mean, invstd, gamma,
                                                                  1.f / (float)M, coef);
    RNN_CHECK_LAUNCH();
  }
  RnnTimedScope timed_apply("bn_bwd_apply", stream);
  const long long nvec = M * (C / 8);
  if (apply_hoist_enabled()) {
    const bf16 *a_dy = (const bf16*)dy, *a_x = (const bf16*)x, *a_y = (const bf16*)y_out;
    int hg = 0;
#define RNN_BWD_HOISTED(MASK)                                                                                        \
  if ((hg = hoisted_grid(bn_bwd_apply_hoisted_kernel<MASK>, nvec, C / 8)) > 0)                                        \
    bn_bwd_apply_hoisted_kernel<MASK><<<hg, 256, 0, stream>>>(a_dy, a_x, a_y, coef, (bf16*)dx, (bf16*)dres, nvec, C / 8, \
                                                              fwd_scale, fwd_shift, relu_mask)
    if (y_out != nullptr) { RNN_BWD_HOISTED(3); }
    else if (fwd_scale != nullptr) { RNN_BWD_HOISTED(1); }
    else if (relu_mask != nullptr) { RNN_BWD_HOISTED(2); }
    else { RNN_BWD_HOISTED(0); }
#undef RNN_BWD_HOISTED
    if (hg > 0) {
      RNN_CHECK_LAUNCH();
      return RNN_OK;
    }
  }
  long long blocks = (nvec + 255) / 256;
  const long long cap = (long long)rnn_num_sms() * apply_ctas_per_sm();
  if (blocks > cap) blocks = cap;
  bn_bwd_apply_kernel<<<(int)blocks, 256, 0, stream>>>((const bf16*)dy, (const bf16*)x, (const bf16*)y_out, coef,
                                                       (bf16*)dx, (bf16*)dres, nvec, C / 8, fwd_scale, fwd_shift, relu_mask);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int rnn_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, float eps, void* y,
                      void* xsum, float* mean, float* rstd, long long M, int E, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(M > 0 && E > 0 && E % (is_f32 ? 4 : 8) == 0, RNN_ERR_SHAPE);
  const int blocks = ceil_div_i(M, 8);
  if (is_f32)
    layernorm_fwd_kernel<float><<<blocks, 256, 0, stream>>>((const float*)x, (const float*)residual, (const float*)gamma,
                                                            (const float*)beta, eps, (float*)y, (float*)xsum, mean, rstd,
                                                            M, E);
  else
    layernorm_fwd_kernel<bf16><<<blocks, 256, 0, stream>>>((const bf16*)x, (const bf16*)residual, (const bf16*)gamma,
                                                           (const bf16*)beta, eps, (bf16*)y, (bf16*)xsum, mean, rstd, M,
                                                           E);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

size_t rnn_layernorm_bwd_workspace(long long M, int E) { return (size_t)ln_bwd_blocks(M) * 2 * E * sizeof(float); }

int rnn_layernorm_bwd(const void* dy, const void* xin, const void* gamma, const float* mean, const float* rstd,
                      void* dx, void* dgamma, void* dbeta, int accumulate_grads, long long M, int E, int is_f32,
                      void* ws, size_t ws_bytes, cudaStream_t stream) {
  RNN_REQUIRE(M > 0 && E > 0 && E % (is_f32 ? 4 : 8) == 0, RNN_ERR_SHAPE);
  const int blocks = ln_bwd_blocks(M);
  RNN_REQUIRE(ws != nullptr && ws_bytes >= (size_t)blocks * 2 * E * sizeof(float), RNN_ERR_WORKSPACE);
  const size_t smem = (size_t)2 * E * sizeof(float);
  RNN_REQUIRE(smem <= 48 * 1024, RNN_ERR_SHAPE);
  if (is_f32) {
    const int nv = ceil_div_i(E, 32 * 4);   // 4-wide column groups per lane
    if (nv <= 2)
      layernorm_bwd_reg_kernel<float, 2><<<blocks, 256, smem, stream>>>((const float*)dy, (const float*)xin, (const float*)gamma,
                                                                        mean, rstd, (float*)dx, (float*)ws, M, E);
    else if (nv <= 4)
      layernorm_bwd_reg_kernel<float, 4><<<blocks, 256, smem, stream>>>((const float*)dy, (const float*)xin, (const float*)gamma,
                                                                        mean, rstd, (float*)dx, (float*)ws, M, E);
    else if (nv <= 8)
      layernorm_bwd_reg_kernel<float, 8><<<blocks, 256, smem, stream>>>((const float*)dy, (const float*)xin, (const float*)gamma,
                                                                        mean, rstd, (float*)dx, (float*)ws, M, E);
    else
      layernorm_bwd_kernel<float><<<blocks, 256, smem, stream>>>((const float*)dy, (const float*)xin, (const float*)gamma,
                                                                 mean, rstd, (float*)dx, (float*)ws, M, E);
    RNN_CHECK_LAUNCH();
    layernorm_bwd_finalize_kernel<float><<<ceil_div_i(E, FIN_CH), FIN_THREADS, 0, stream>>>((const float*)ws, blocks, E,
                                                                                (float*)dgamma, (float*)dbeta,
                                                                                accumulate_grads);
  } else {
    const int nv = ceil_div_i(E, 32 * 8);   // 8-wide column groups per lane
    if (nv <= 2)
      layernorm_bwd_reg_kernel<bf16, 2><<<blocks, 256, smem, stream>>>((const bf16*)dy, (const bf16*)xin, (const bf16*)gamma,
                                                                       mean, rstd, (bf16*)dx, (float*)ws, M, E);
    else if (nv <= 4)
      layernorm_bwd_reg_kernel<bf16, 4><<<blocks, 256, smem, stream>>>((const bf16*)dy, (const bf16*)xin, (const bf16*)gamma,
                                                                       mean, rstd, (bf16*)dx, (float*)ws, M, E);
    else
      layernorm_bwd_kernel<bf16><<<blocks, 256, smem, stream>>>((const bf16*)dy, (const bf16*)xin, (const bf16*)gamma,
                                                                mean, rstd, (bf16*)dx, (float*)ws, M, E);
    RNN_CHECK_LAUNCH();
    layernorm_bwd_finalize_kernel<bf16><<<ceil_div_i(E, FIN_CH), FIN_THREADS, 0, stream>>>((const float*)ws, blocks, E,
                                                                               (bf16*)dgamma, (bf16*)dbeta,
                                                                               accumulate_grads);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

}  // extern "C"
