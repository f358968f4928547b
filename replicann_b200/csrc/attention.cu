// Fused scaled-dot-product attention for the reference's heads (nn/attention.py:28-50):
//     w = q k^T * scale ; mask (bool: True -> -inf | float32: additive) ; softmax ; dropout ; z = w v
// in one kernel per direction -- the [B,H,T,Tk] score tensor the reference materialises (:37,45,47) never
// touches HBM (flash-style online softmax; backward recomputes the probabilities from the saved
// log-sum-exp).  All heads of a multi-head module run in one launch: q/k/v/z are addressed as
// [batch][row][head*hs + d] with explicit strides, which is exactly the layout of the packed projection
// output and of the reference's `torch.cat(heads, dim=-1)` (:157-160,180) -- the concat disappears.
//
// The contraction sizes here (head size <= 128, T a few hundred) are tiny next to the projection / FFN
// GEMMs (24*M*E^2 vs 4*B*T^2*E FLOPs), so this kernel uses fp32 CUDA-core FMAs with
// shared-memory K/V tiles; the GEMMs around it are the tcgen05 kernels in tc_gemm.cu.
#include <string.h>

#include "tc_ptx.cuh"

namespace {

struct AttnParams {
  const void *q, *k, *v;
  void* z;
  long long q_bs, q_rs, k_bs, k_rs, v_bs, v_rs, z_bs, z_rs;  // batch / row strides in elements
  int B, H, T, Tk, hs;
  float scale;
  const void* mask;        // nullable; [mask_b][T][Tk] with strides
  int mask_kind;           // 0 none, 1 bool (uint8, nonzero = masked out), 2 float32 additive
  long long mask_bs, mask_rs;
  float p_drop;
  unsigned long long seed, offset;
  float* lse;              // [B,H,T]
  // backward
  const void* dz;
  long long dz_bs, dz_rs;
  void *dq, *dk, *dv;      // same strides as q / k / v
  float* delta;            // [B,H,T]
};

constexpr int KC = 64;     // keys (or queries, in the dk/dv kernel) per shared-memory tile
constexpr int ROWS = 16;   // rows handled per CTA (4 warps x 4)
constexpr int MAXD = 4;    // head size <= 32 * MAXD = 128

// Device-side RNG epoch of this file (see elementwise.cu): added to every Philox offset.
static __device__ unsigned long long g_rng_epoch_attn = 0ull;
__global__ void rng_advance_attn_kernel(unsigned long long inc) { g_rng_epoch_attn += inc; }

// Philox-4x32-10: one call yields four 32-bit words = EIGHT 16-bit keep decisions.  Attention-weight element `idx`
// uses half (idx & 1) of word ((idx >> 1) & 3) of the block with counter (idx >> 3), so a kernel that walks eight
// consecutive elements pays for one call (the generator, not the exponential, is what the dropout path of the
// fused kernels spends its ALU time on).  A 16-bit decision resolves p to 1.5e-5.
__device__ __forceinline__ uint4 philox4(unsigned long long seed, unsigned long long offset, unsigned long long ctr) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  offset += g_rng_epoch_attn;
  uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = (uint32_t)offset, c3 = (uint32_t)(offset >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0, hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  return make_uint4(c0, c1, c2, c3);
}
// 16-bit random value of attention-weight element `idx`
__device__ __forceinline__ uint32_t philox_half(unsigned long long seed, unsigned long long offset,
                                                unsigned long long idx) {
  const uint4 r = philox4(seed, offset, idx >> 3);
  const int w = (int)((idx >> 1) & 3);
  const uint32_t word = w == 0 ? r.x : w == 1 ? r.y : w == 2 ? r.z : r.w;
  return (idx & 1) ? (word >> 16) : (word & 0xFFFFu);
}
// keep threshold on the 16-bit scale: an element is kept when its value is >= the threshold
__device__ __forceinline__ uint32_t keep_thresh16(float p_drop) { return (uint32_t)(p_drop * 65536.0f); }

__device__ __forceinline__ float keep_scale(const AttnParams& p, int bh, int i, int j) {
  if (p.p_drop <= 0.f) return 1.f;
  const unsigned long long idx = ((unsigned long long)bh * p.T + i) * p.Tk + j;
  const uint32_t r = philox_half(p.seed, p.offset, idx);
  return r >= keep_thresh16(p.p_drop) ? 1.f / (1.f - p.p_drop) : 0.f;
}

__device__ __forceinline__ float masked_score(const AttnParams& p, float s, int b, int i, int j) {
  if (p.mask_kind == 1) {
    const uint8_t m = reinterpret_cast<const uint8_t*>(p.mask)[b * p.mask_bs + i * p.mask_rs + j];
    return m ? -INFINITY : s;
  }
  if (p.mask_kind == 2) return s + reinterpret_cast<const float*>(p.mask)[b * p.mask_bs + i * p.mask_rs + j];
  return s;
}

template <typename T>
__device__ __forceinline__ void load_tile(float* dst /*[KC][hs+1]*/, const T* src, long long rs, int row0, int nrows,
                                          int total_rows, int hs) {
  for (int e = threadIdx.x; e < KC * hs; e += blockDim.x) {
    const int r = e / hs, d = e - r * hs;
    const int row = row0 + r;
    dst[r * (hs + 1) + d] = (r < nrows && row < total_rows) ? (float)src[(long long)row * rs + d] : 0.f;
  }
}

// ------------------------------------------------------------------------------------------ forward
template <typename T>
__global__ void __launch_bounds__(128)
attn_fwd_kernel(const AttnParams p) {
  extern __shared__ float sm[];
  const int hs = p.hs;
  float* ks = sm;                         // [KC][hs+1]
  float* vs = ks + KC * (hs + 1);         // [KC][hs+1]
  float* qs = vs + KC * (hs + 1);         // [4 warps][hs]
  float* ps = qs + 4 * hs;                // [4 warps][KC]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const T* qb = reinterpret_cast<const T*>(p.q) + b * p.q_bs + h * hs;
  const T* kb = reinterpret_cast<const T*>(p.k) + b * p.k_bs + h * hs;
  const T* vb = reinterpret_cast<const T*>(p.v) + b * p.v_bs + h * hs;
  T* zb = reinterpret_cast<T*>(p.z) + b * p.z_bs + h * hs;
  const int row_base = blockIdx.x * ROWS + warp * 4;

  float m[4], l[4], acc[4][MAXD];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    m[r] = -INFINITY; l[r] = 0.f;
#pragma unroll
    for (int t = 0; t < MAXD; ++t) acc[r][t] = 0.f;
  }
  for (int k0 = 0; k0 < p.Tk; k0 += KC) {
    __syncthreads();
    load_tile<T>(ks, kb, p.k_rs, k0, KC, p.Tk, hs);
    load_tile<T>(vs, vb, p.v_rs, k0, KC, p.Tk, hs);
    __syncthreads();
    const int nk = min(KC, p.Tk - k0);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int i = row_base + r;
      if (i >= p.T) continue;  // warp-uniform
      for (int d = lane; d < hs; d += 32) qs[warp * hs + d] = (float)qb[(long long)i * p.q_rs + d];
      __syncwarp();
      // scores for keys lane and lane+32 of this tile
      float s[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int j = lane + 32 * u;
        float a = 0.f;
        if (j < nk) {
          const float* kr = ks + j * (hs + 1);
          for (int d = 0; d < hs; ++d) a = fmaf(qs[warp * hs + d], kr[d], a);
          a = masked_score(p, a * p.scale, b, i, k0 + j);
        } else {
          a = -INFINITY;
        }
        s[u] = a;
      }
      const float mt = warp_max(fmaxf(s[0], s[1]));
      const float mn = fmaxf(m[r], mt);
      const float corr = (m[r] == -INFINITY) ? 0.f : __expf(m[r] - mn);
      float psum = 0.f;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int j = lane + 32 * u;
        // a row masked out so far contributes nothing; if it stays fully masked the final 0/0 gives NaN, as torch
        const float pe = (j < nk && mn != -INFINITY) ? __expf(s[u] - mn) : 0.f;
        psum += pe;
        ps[warp * KC + j] = pe * ((j < nk) ? keep_scale(p, bh, i, k0 + j) : 0.f);
      }
      psum = warp_sum(psum);
      l[r] = l[r] * corr + psum;
      m[r] = mn;
      __syncwarp();
#pragma unroll
      for (int t = 0; t < MAXD; ++t) {
        const int d = lane + 32 * t;
        if (d < hs) {
          float a = acc[r][t] * corr;
          for (int j = 0; j < nk; ++j) a = fmaf(ps[warp * KC + j], vs[j * (hs + 1) + d], a);
          acc[r][t] = a;
        }
      }
      __syncwarp();
    }
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int i = row_base + r;
    if (i >= p.T) continue;
    const float inv = 1.f / l[r];
#pragma unroll
    for (int t = 0; t < MAXD; ++t) {
      const int d = lane + 32 * t;
      if (d < hs) zb[(long long)i * p.z_rs + d] = (T)(acc[r][t] * inv);
    }
    if (lane == 0 && p.lse) p.lse[(long long)bh * p.T + i] = m[r] + __logf(l[r]);
  }
}

// ------------------------------------------------------------------------------------------ backward: dq (+delta)
template <typename T>
__global__ void __launch_bounds__(128)
attn_bwd_dq_kernel(const AttnParams p) {
  extern __shared__ float sm[];
  const int hs = p.hs;
  float* ks = sm;
  float* vs = ks + KC * (hs + 1);
  float* qs = vs + KC * (hs + 1);         // [4][hs]
  float* gs = qs + 4 * hs;                // [4][hs]   dz row
  float* ds = gs + 4 * hs;                // [4][KC]   dS
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const T* qb = reinterpret_cast<const T*>(p.q) + b * p.q_bs + h * hs;
  const T* kb = reinterpret_cast<const T*>(p.k) + b * p.k_bs + h * hs;
  const T* vb = reinterpret_cast<const T*>(p.v) + b * p.v_bs + h * hs;
  const T* zb = reinterpret_cast<const T*>(p.z) + b * p.z_bs + h * hs;
  const T* gb = reinterpret_cast<const T*>(p.dz) + b * p.dz_bs + h * hs;
  T* dqb = reinterpret_cast<T*>(p.dq) + b * p.q_bs + h * hs;
  const int row_base = blockIdx.x * ROWS + warp * 4;

  float acc[4][MAXD], delta[4], lse[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int i = row_base + r;
    float dl = 0.f;
    if (i < p.T)
      for (int d = lane; d < hs; d += 32) dl += (float)gb[(long long)i * p.dz_rs + d] * (float)zb[(long long)i * p.z_rs + d];
    delta[r] = warp_sum(dl);
    lse[r] = (i < p.T) ? p.lse[(long long)bh * p.T + i] : 0.f;
    if (i < p.T && lane == 0) p.delta[(long long)bh * p.T + i] = delta[r];
#pragma unroll
    for (int t = 0; t < MAXD; ++t) acc[r][t] = 0.f;
  }
  for (int k0 = 0; k0 < p.Tk; k0 += KC) {
    __syncthreads();
    load_tile<T>(ks, kb, p.k_rs, k0, KC, p.Tk, hs);
    load_tile<T>(vs, vb, p.v_rs, k0, KC, p.Tk, hs);
    __syncthreads();
    const int nk = min(KC, p.Tk - k0);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int i = row_base + r;
      if (i >= p.T) continue;
      for (int d = lane; d < hs; d += 32) {
        qs[warp * hs + d] = (float)qb[(long long)i * p.q_rs + d];
        gs[warp * hs + d] = (float)gb[(long long)i * p.dz_rs + d];
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int j = lane + 32 * u;
        float dsv = 0.f;
        if (j < nk) {
          const float* kr = ks + j * (hs + 1);
          const float* vr = vs + j * (hs + 1);
          float a = 0.f, dp = 0.f;
          for (int d = 0; d < hs; ++d) { a = fmaf(qs[warp * hs + d], kr[d], a); dp = fmaf(gs[warp * hs + d], vr[d], dp); }
          a = masked_score(p, a * p.scale, b, i, k0 + j);
          const float pe = __expf(a - lse[r]);
          dsv = pe * (dp * keep_scale(p, bh, i, k0 + j) - delta[r]);
        }
        ds[warp * KC + j] = dsv;
      }
      __syncwarp();
#pragma unroll
      for (int t = 0; t < MAXD; ++t) {
        const int d = lane + 32 * t;
        if (d < hs) {
          float a = acc[r][t];
          for (int j = 0; j < nk; ++j) a = fmaf(ds[warp * KC + j], ks[j * (hs + 1) + d], a);
          acc[r][t] = a;
        }
      }
      __syncwarp();
    }
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int i = row_base + r;
    if (i >= p.T) continue;
#pragma unroll
    for (int t = 0; t < MAXD; ++t) {
      const int d = lane + 32 * t;
      if (d < hs) dqb[(long long)i * p.q_rs + d] = (T)(acc[r][t] * p.scale);
    }
  }
}

// ------------------------------------------------------------------------------------------ backward: dk, dv
// One warp-row per KEY; queries stream through shared memory in tiles of KC.
template <typename T>
__global__ void __launch_bounds__(128)
attn_bwd_dkv_kernel(const AttnParams p) {
  extern __shared__ float sm[];
  const int hs = p.hs;
  float* qs = sm;                          // [KC][hs+1] queries
  float* gs = qs + KC * (hs + 1);          // [KC][hs+1] dz
  float* kr_s = gs + KC * (hs + 1);        // [4][hs]    this warp's key row
  float* vr_s = kr_s + 4 * hs;             // [4][hs]    this warp's value row
  float* pw = vr_s + 4 * hs;               // [4][KC]    P*keep   (for dv)
  float* dw = pw + 4 * KC;                 // [4][KC]    dS       (for dk)
  float* ls = dw + 4 * KC;                 // [KC] lse, [KC] delta
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const T* qb = reinterpret_cast<const T*>(p.q) + b * p.q_bs + h * hs;
  const T* kb = reinterpret_cast<const T*>(p.k) + b * p.k_bs + h * hs;
  const T* vb = reinterpret_cast<const T*>(p.v) + b * p.v_bs + h * hs;
  const T* gb = reinterpret_cast<const T*>(p.dz) + b * p.dz_bs + h * hs;
  T* dkb = reinterpret_cast<T*>(p.dk) + b * p.k_bs + h * hs;
  T* dvb = reinterpret_cast<T*>(p.dv) + b * p.v_bs + h * hs;
  const int key_base = blockIdx.x * ROWS + warp * 4;

  float dk[4][MAXD], dv[4][MAXD];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int t = 0; t < MAXD; ++t) dk[r][t] = dv[r][t] = 0.f;

  for (int q0 = 0; q0 < p.T; q0 += KC) {
    __syncthreads();
    load_tile<T>(qs, qb, p.q_rs, q0, KC, p.T, hs);
    load_tile<T>(gs, gb, p.dz_rs, q0, KC, p.T, hs);
    for (int e = threadIdx.x; e < KC; e += blockDim.x) {
      const int i = q0 + e;
      ls[e] = i < p.T ? p.lse[(long long)bh * p.T + i] : 0.f;
      ls[KC + e] = i < p.T ? p.delta[(long long)bh * p.T + i] : 0.f;
    }
    __syncthreads();
    const int nq = min(KC, p.T - q0);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int j = key_base + r;
      if (j >= p.Tk) continue;
      for (int d = lane; d < hs; d += 32) {
        kr_s[warp * hs + d] = (float)kb[(long long)j * p.k_rs + d];
        vr_s[warp * hs + d] = (float)vb[(long long)j * p.v_rs + d];
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int e = lane + 32 * u;
        float pk = 0.f, dsv = 0.f;
        if (e < nq) {
          const int i = q0 + e;
          const float* qr = qs + e * (hs + 1);
          const float* gr = gs + e * (hs + 1);
          float a = 0.f, dp = 0.f;
          for (int d = 0; d < hs; ++d) { a = fmaf(qr[d], kr_s[warp * hs + d], a); dp = fmaf(gr[d], vr_s[warp * hs + d], dp); }
          a = masked_score(p, a * p.scale, b, i, j);
          const float pe = __expf(a - ls[e]);
          const float ksc = keep_scale(p, bh, i, j);
          pk = pe * ksc;
          dsv = pe * (dp * ksc - ls[KC + e]);
        }
        pw[warp * KC + e] = pk;
        dw[warp * KC + e] = dsv;
      }
      __syncwarp();
#pragma unroll
      for (int t = 0; t < MAXD; ++t) {
        const int d = lane + 32 * t;
        if (d < hs) {
          float a = dv[r][t], c = dk[r][t];
          for (int e = 0; e < nq; ++e) {
            a = fmaf(pw[warp * KC + e], gs[e * (hs + 1) + d], a);
            c = fmaf(dw[warp * KC + e], qs[e * (hs + 1) + d], c);
          }
          dv[r][t] = a; dk[r][t] = c;
        }
      }
      __syncwarp();
    }
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int j = key_base + r;
    if (j >= p.Tk) continue;
#pragma unroll
    for (int t = 0; t < MAXD; ++t) {
      const int d = lane + 32 * t;
      if (d < hs) {
        dkb[(long long)j * p.k_rs + d] = (T)(dk[r][t] * p.scale);
        dvb[(long long)j * p.v_rs + d] = (T)dv[r][t];
      }
    }
  }
}

size_t fwd_smem(int hs) { return (size_t)(2 * KC * (hs + 1) + 4 * hs + 4 * KC) * sizeof(float); }
size_t dq_smem(int hs) { return (size_t)(2 * KC * (hs + 1) + 8 * hs + 4 * KC) * sizeof(float); }
size_t dkv_smem(int hs) { return (size_t)(2 * KC * (hs + 1) + 8 * hs + 8 * KC + 2 * KC) * sizeof(float); }

int check(const rnn_attn_desc* d) {
  RNN_REQUIRE(d != nullptr, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->B > 0 && d->H > 0 && d->T > 0 && d->Tk > 0 && d->hs > 0 && d->hs <= 32 * MAXD, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->mask_kind >= 0 && d->mask_kind <= 2, RNN_ERR_DTYPE);
  RNN_REQUIRE(d->p_drop >= 0.f && d->p_drop < 1.f, RNN_ERR_SHAPE);
  RNN_REQUIRE((long long)d->B * d->H < 65536, RNN_ERR_SHAPE);
  return RNN_OK;
}

AttnParams to_params(const rnn_attn_desc* d) {
  AttnParams p;
  memset(&p, 0, sizeof(p));
  p.B = d->B; p.H = d->H; p.T = d->T; p.Tk = d->Tk; p.hs = d->hs;
  p.q_bs = d->q_bs; p.q_rs = d->q_rs; p.k_bs = d->k_bs; p.k_rs = d->k_rs; p.v_bs = d->v_bs; p.v_rs = d->v_rs;
  p.z_bs = d->z_bs; p.z_rs = d->z_rs; p.dz_bs = d->dz_bs; p.dz_rs = d->dz_rs;
  p.scale = d->scale; p.mask_kind = d->mask_kind; p.mask_bs = d->mask_bs; p.mask_rs = d->mask_rs;
  p.p_drop = d->p_drop; p.seed = d->seed; p.offset = d->offset;
  return p;
}

template <typename K>
int set_smem(K kern, size_t bytes) {
  if (bytes > 48 * 1024) RNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return RNN_OK;
}


// ---------------------------------------------------------------------------------------------------------------
// Tensor-core form (bf16, head size a multiple of 64): the two contractions of the forward pass and the four of
// the backward pass are batched tcgen05 GEMMs over all (batch, head) problems (rnn_gemm_batched_bf16, operands read
// in place from the packed projections); the row-wise softmax / mask / dropout step between them runs here.  The
// scores are kept in fp32 between the QK^T GEMM and the softmax; the probabilities are stored in bf16 because
// they are tensor-core operands next.
// ---------------------------------------------------------------------------------------------------------------
struct SmaxParams {
  int H, T, Tk;
  float scale;
  const void* mask;
  int mask_kind;
  long long mask_bs, mask_rs;
  float p_drop;
  unsigned long long seed, offset;
};

__device__ __forceinline__ float smax_score(const SmaxParams& p, float raw, int b, int i, int j) {
  const float s = raw * p.scale;
  if (p.mask_kind == 1) {
    const uint8_t m = reinterpret_cast<const uint8_t*>(p.mask)[b * p.mask_bs + i * p.mask_rs + j];
    return m ? -INFINITY : s;
  }
  if (p.mask_kind == 2) return s + reinterpret_cast<const float*>(p.mask)[b * p.mask_bs + i * p.mask_rs + j];
  return s;
}

// four consecutive probabilities in their storage type: bf16 (operands of the bf16 pipeline) or fp32 (the fp32
// pipeline keeps them exact and splits them into bf16 pieces afterwards)
__device__ __forceinline__ void store4(bf16* p, const float* f) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(f[0], f[1]), hi = __floats2bfloat162_rn(f[2], f[3]);
  uint2 out;
  out.x = *reinterpret_cast<uint32_t*>(&lo); out.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(p) = out;
}
__device__ __forceinline__ void store4(float* p, const float* f) {
  *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
}
__device__ __forceinline__ void load4(const bf16* p, float* f) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}
__device__ __forceinline__ void load4(const float* p, float* f) {
  const float4 v = *reinterpret_cast<const float4*>(p);
  f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}

// One warp per score row: P = softmax(S * scale (+ mask)), Pd = dropout(P).  Lanes walk the row four elements at a
// time (Tk is a multiple of 8 here): 128-bit loads of the scores, one Philox call per four probabilities, 64-bit
// stores.  Three passes over a row of a few KB (served by L1 after the first); same Philox word per element as
// the fused kernel (keep_scale above).
template <typename TP>
__global__ void __launch_bounds__(128)
attn_softmax_fwd_kernel(const float* __restrict__ S, TP* __restrict__ P, TP* __restrict__ Pd, long long rows,
                        SmaxParams p) {
  const int lane = threadIdx.x & 31;
  const long long r = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int bh = (int)(r / p.T), i = (int)(r - (long long)bh * p.T), b = bh / p.H;
  const float4* srow = reinterpret_cast<const float4*>(S + r * p.Tk);
  const int n4 = p.Tk >> 2;
  auto score4 = [&](int j4, float* s) {
    const float4 raw = srow[j4];
    s[0] = smax_score(p, raw.x, b, i, 4 * j4 + 0); s[1] = smax_score(p, raw.y, b, i, 4 * j4 + 1);
    s[2] = smax_score(p, raw.z, b, i, 4 * j4 + 2); s[3] = smax_score(p, raw.w, b, i, 4 * j4 + 3);
  };
  float m = -INFINITY;
  for (int j4 = lane; j4 < n4; j4 += 32) {
    float s[4];
    score4(j4, s);
    m = fmaxf(fmaxf(m, fmaxf(s[0], s[1])), fmaxf(s[2], s[3]));
  }
  m = warp_max(m);
  float l = 0.f;
  for (int j4 = lane; j4 < n4; j4 += 32) {
    float s[4];
    score4(j4, s);
    l += (__expf(s[0] - m) + __expf(s[1] - m)) + (__expf(s[2] - m) + __expf(s[3] - m));
  }
  l = warp_sum(l);
  const float inv = 1.f / l;
  const float keep = 1.f / (1.f - p.p_drop);
  const uint32_t thresh = keep_thresh16(p.p_drop);
  const unsigned long long row_idx = ((unsigned long long)bh * p.T + i) * p.Tk;   // multiple of 8
  for (int j4 = lane; j4 < n4; j4 += 32) {
    float s[4], pv[4];
    score4(j4, s);
#pragma unroll
    for (int e = 0; e < 4; ++e) pv[e] = __expf(s[e] - m) * inv;
    store4(P + r * p.Tk + 4 * j4, pv);
    if (Pd != P) {
      // elements 4*j4 .. 4*j4+3 of the row: words (x, y) or (z, w) of the 8-element block they fall into
      const uint4 rnd = philox4(p.seed, p.offset, (row_idx >> 3) + (j4 >> 1));
      const uint32_t w0 = (j4 & 1) ? rnd.z : rnd.x, w1 = (j4 & 1) ? rnd.w : rnd.y;
      const float pk[4] = {(w0 & 0xFFFFu) >= thresh ? pv[0] * keep : 0.f, (w0 >> 16) >= thresh ? pv[1] * keep : 0.f,
                           (w1 & 0xFFFFu) >= thresh ? pv[2] * keep : 0.f, (w1 >> 16) >= thresh ? pv[3] * keep : 0.f};
      store4(Pd + r * p.Tk + 4 * j4, pk);
    }
  }
}

// dS = scale * P * (dPd * keep - delta), delta_i = sum_j dPd_ij * Pd_ij (= dz_i . z_i).  `keep` is read off the
// stored pair: an element with P != 0 and Pd == 0 was dropped (P == 0 contributes nothing either way).
template <typename TP>
__global__ void __launch_bounds__(128)
attn_softmax_bwd_kernel(const TP* __restrict__ P, const TP* __restrict__ Pd, const float* dPd, TP* dS,
                        long long rows, int Tk, float scale, float p_drop) {   // dS may alias dPd (fp32 pipeline)
  const int lane = threadIdx.x & 31;
  const long long r = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const TP* prow = P + r * Tk;      // four elements per step (Tk % 8 == 0)
  const TP* pdrow = Pd + r * Tk;
  const float4* grow = reinterpret_cast<const float4*>(dPd + r * Tk);
  const int n4 = Tk >> 2;
  float delta = 0.f;
  for (int j4 = lane; j4 < n4; j4 += 32) {
    float pd[4];
    load4(pdrow + 4 * j4, pd);
    const float4 g = grow[j4];
    delta += (g.x * pd[0] + g.y * pd[1]) + (g.z * pd[2] + g.w * pd[3]);
  }
  delta = warp_sum(delta);
  const float keep = 1.f / (1.f - p_drop);
  for (int j4 = lane; j4 < n4; j4 += 32) {
    float pv[4], pd[4];
    load4(prow + 4 * j4, pv);
    const float4 g4 = grow[j4];
    float g[4] = {g4.x, g4.y, g4.z, g4.w};
    if (Pd != P) {
      load4(pdrow + 4 * j4, pd);
#pragma unroll
      for (int e = 0; e < 4; ++e) g[e] = pd[e] != 0.f ? g[e] * keep : 0.f;
    }
    const float ds[4] = {scale * pv[0] * (g[0] - delta), scale * pv[1] * (g[1] - delta),
                         scale * pv[2] * (g[2] - delta), scale * pv[3] * (g[3] - delta)};
    store4(dS + r * Tk + 4 * j4, ds);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// fp32 pipeline on the tensor cores at fp32 accuracy.  Every operand of the six contractions is split into three
// bf16 pieces (x = x1 + x2 + x3, see split3_kernel in tc_gemm.cu) and repacked into dense per-(batch, head) slabs
// with the pieces laid out along the reduction dimension, so that each contraction is one batched bf16 GEMM over
// a reduction six times as long, with fp32 output.  Piece orders: operand A {1,1,2,1,2,3}, operand B {1,2,1,3,2,1}.
//   side layout    [z][rows][6 * cols]  (K-major operand:  K = cols, pieces side by side)
//   stacked layout [z][6 * rows][cols]  (MN-major operand: K = rows, pieces stacked)
// Source element (b, r, h, c) = x[b * bs + r * rs + h * cols + c]; H = 1 describes a dense [B][rows][cols] tensor.
// ---------------------------------------------------------------------------------------------------------------
__constant__ int kAttnPieceA[6] = {0, 0, 1, 0, 1, 2};
__constant__ int kAttnPieceB[6] = {0, 1, 0, 2, 1, 0};

__global__ void __launch_bounds__(256)
split3_heads_kernel(const float* __restrict__ x, long long bs, long long rs, int B, int H, int rows, int cols,
                    bf16* __restrict__ out, int stacked, int is_b) {
  const int c4n = cols >> 2;
  const long long total = (long long)B * H * rows * c4n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long t = i;
    const int c = (int)(t % c4n) * 4; t /= c4n;
    const int r = (int)(t % rows); t /= rows;
    const int h = (int)(t % H);
    const int b = (int)(t / H);
    const long long z = (long long)b * H + h;
    float f[4];
    load4(x + b * bs + r * rs + (long long)h * cols + c, f);
    float piece[3][4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      piece[0][e] = __bfloat162float(__float2bfloat16(f[e]));
      const float r1 = f[e] - piece[0][e];
      piece[1][e] = __bfloat162float(__float2bfloat16(r1));
      piece[2][e] = r1 - piece[1][e];
    }
#pragma unroll
    for (int sidx = 0; sidx < 6; ++sidx) {
      const int k = is_b ? kAttnPieceB[sidx] : kAttnPieceA[sidx];
      bf16* dst = stacked ? out + ((z * 6 + sidx) * rows + r) * cols + c
                          : out + (z * rows + r) * (6ll * cols) + (long long)sidx * cols + c;
      store4(dst, k == 0 ? piece[0] : k == 1 ? piece[1] : piece[2]);
    }
  }
}

// dense fp32 slabs [B*H][rows][cols] -> the concatenated-heads layout out[b*bs + r*rs + h*cols + c]
__global__ void __launch_bounds__(256)
gather_heads_kernel(const float* __restrict__ slab, float* __restrict__ out, long long bs, long long rs, int B, int H,
                    int rows, int cols) {
  const int c4n = cols >> 2;
  const long long total = (long long)B * H * rows * c4n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long t = i;
    const int c = (int)(t % c4n) * 4; t /= c4n;
    const int r = (int)(t % rows); t /= rows;
    const int h = (int)(t % H);
    const int b = (int)(t / H);
    const float4 v = *reinterpret_cast<const float4*>(slab + i * 4);
    *reinterpret_cast<float4*>(out + b * bs + r * rs + (long long)h * cols + c) = v;
  }
}

int ew_blocks(long long total) {
  const long long b = (total + 255) / 256, cap = (long long)rnn_num_sms() * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
int launch_split(const float* x, long long bs, long long rs, int B, int H, int rows, int cols, bf16* out, int stacked,
                 int is_b, cudaStream_t stream) {
  split3_heads_kernel<<<ew_blocks((long long)B * H * rows * (cols / 4)), 256, 0, stream>>>(x, bs, rs, B, H, rows, cols,
                                                                                         out, stacked, is_b);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int launch_gather(const float* slab, float* out, long long bs, long long rs, int B, int H, int rows, int cols,
                  cudaStream_t stream) {
  gather_heads_kernel<<<ew_blocks((long long)B * H * rows * (cols / 4)), 256, 0, stream>>>(slab, out, bs, rs, B, H, rows,
                                                                                         cols);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
// one batched GEMM over dense slabs: D[z] (fp32 [M][N]) = A[z] . B[z]^T with a reduction of 6 * K
int slab_gemm(int nz, int M, int N, int K, int a_mn, int b_mn, const bf16* A, const bf16* Bm, float* D,
              cudaStream_t stream) {
  rnn_bgemm_desc g;
  memset(&g, 0, sizeof(g));
  g.batches = nz; g.heads = 1;
  g.M = M; g.N = N; g.K = 6 * K; g.a_mn = a_mn; g.b_mn = b_mn; g.out_f32 = 1;
  if (a_mn) { g.a_ld = M; g.a_bstride = 6ll * K * M; } else { g.a_ld = 6ll * K; g.a_bstride = 6ll * K * M; }
  if (b_mn) { g.b_ld = N; g.b_bstride = 6ll * K * N; } else { g.b_ld = 6ll * K; g.b_bstride = 6ll * K * N; }
  g.d_ld = N; g.d_bstride = (long long)M * N;
  return rnn_gemm_batched_bf16(&g, A, Bm, D, stream);
}

#include "attention_fused.cuh"

// ---- host side of the fused tcgen05 attention ------------------------------------------------------------------
int fused_ok(const rnn_attn_desc* d) {
  if (d == nullptr) return 0;
  if (!(d->hs == 64 || d->hs == 128) || d->T % 8 != 0 || d->Tk % 8 != 0) return 0;
  if (d->q_rs % 8 || d->k_rs % 8 || d->v_rs % 8 || d->z_rs % 8 || d->q_bs % 8 || d->k_bs % 8 || d->v_bs % 8 || d->z_bs % 8)
    return 0;
  return (long long)d->B * d->H < 65536 && d->B <= 65535 && d->H <= 65535;
}

fa::Params fused_params(const rnn_attn_desc* d, const void* mask) {
  fa::Params p;
  memset(&p, 0, sizeof(p));
  p.B = d->B; p.H = d->H; p.T = d->T; p.Tk = d->Tk;
  p.scale = d->scale;
  p.mask = mask; p.mask_kind = d->mask_kind; p.mask_bs = d->mask_bs; p.mask_rs = d->mask_rs;
  const long long esz = d->mask_kind == 2 ? 4 : 1;
  p.mask_vec = mask != nullptr && (reinterpret_cast<uintptr_t>(mask) & 15u) == 0 && (d->mask_rs * esz) % 16 == 0 &&
               (d->mask_bs * esz) % 16 == 0;
  p.p_drop = d->p_drop; p.seed = d->seed; p.offset = d->offset;
  return p;
}

// operand [B][rows][H*hs] with explicit row / batch strides -> 3-D tensor map, box {64, box_rows, 1}
int fused_map(CUtensorMap* m, const void* ptr, const rnn_attn_desc* d, int rows, long long rs, long long bs, int box_rows) {
  RNN_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15u) == 0, RNN_ERR_ALIGN);
  const long long bstride = d->B == 1 ? (long long)rows * rs : bs;
  return rnn_internal_tmap_3d(m, ptr, d->B, rows, (long long)d->H * d->hs, rs, bstride, box_rows);
}

template <int HS>
int fused_fwd_launch(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask, void* z,
                     float* lse, cudaStream_t stream) {
  CUtensorMap tq, tk, tv, tz;
  int rc = fused_map(&tq, q, d, d->T, d->q_rs, d->q_bs, 128);
  if (rc == RNN_OK) rc = fused_map(&tk, k, d, d->Tk, d->k_rs, d->k_bs, 128);
  if (rc == RNN_OK) rc = fused_map(&tv, v, d, d->Tk, d->v_rs, d->v_bs, 128);
  if (rc == RNN_OK) rc = fused_map(&tz, z, d, d->T, d->z_rs, d->z_bs, 32);
  if (rc != RNN_OK) return rc;
  fa::Params p = fused_params(d, mask);
  p.lse = lse;
  auto kern = fa::fwd_kernel<HS>;
  static bool attr_set[64] = {false};
  int dev = 0;
  RNN_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    RNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, fa::FwdCfg<HS>::SMEM));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  kern<<<dim3(ceil_div_i(d->T, fa::BQ), d->H, d->B), fa::NTHREADS, fa::FwdCfg<HS>::SMEM, stream>>>(tq, tk, tv, tz, p);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

template <int HS>
int fused_bwd_launch(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask, const void* z,
                     const void* dz, const float* lse, float* delta, void* dq, void* dk, void* dv, cudaStream_t stream) {
  CUtensorMap tq, tk, tv, tdo, tdq, tdk, tdv;
  int rc = fused_map(&tq, q, d, d->T, d->q_rs, d->q_bs, 128);
  if (rc == RNN_OK) rc = fused_map(&tk, k, d, d->Tk, d->k_rs, d->k_bs, 128);
  if (rc == RNN_OK) rc = fused_map(&tv, v, d, d->Tk, d->v_rs, d->v_bs, 128);
  if (rc == RNN_OK) rc = fused_map(&tdo, dz, d, d->T, d->dz_rs, d->dz_bs, 128);
  if (rc == RNN_OK) rc = fused_map(&tdq, dq, d, d->T, d->q_rs, d->q_bs, 32);
  if (rc == RNN_OK) rc = fused_map(&tdk, dk, d, d->Tk, d->k_rs, d->k_bs, 32);
  if (rc == RNN_OK) rc = fused_map(&tdv, dv, d, d->Tk, d->v_rs, d->v_bs, 32);
  if (rc != RNN_OK) return rc;
  fa::Params p = fused_params(d, mask);
  p.lse = const_cast<float*>(lse);
  p.delta = delta;
  p.z = reinterpret_cast<const bf16*>(z); p.dz = reinterpret_cast<const bf16*>(dz);
  p.z_bs = d->z_bs; p.z_rs = d->z_rs; p.dz_bs = d->dz_bs; p.dz_rs = d->dz_rs;
  auto k1 = fa::dq_kernel<HS>;
  auto k2 = fa::dkdv_kernel<HS>;
  static bool attr_set[64] = {false};
  int dev = 0;
  RNN_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    RNN_CUDA(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, fa::DqCfg<HS>::SMEM));
    RNN_CUDA(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, fa::DkvCfg<HS>::SMEM));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  k1<<<dim3(ceil_div_i(d->T, fa::BQ), d->H, d->B), fa::NTHREADS, fa::DqCfg<HS>::SMEM, stream>>>(tq, tk, tv, tdo, tdq, p);
  RNN_CHECK_LAUNCH();
  k2<<<dim3(ceil_div_i(d->Tk, fa::BKV), d->H, d->B), fa::NTHREADS, fa::DkvCfg<HS>::SMEM, stream>>>(tq, tk, tv, tdo, tdk, tdv, p);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int tc_check(const rnn_attn_desc* d) {
  int rc = check(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(d->hs % 64 == 0 && d->T % 8 == 0 && d->Tk % 8 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->q_rs % 8 == 0 && d->k_rs % 8 == 0 && d->v_rs % 8 == 0 && d->z_rs % 8 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->q_bs % 8 == 0 && d->k_bs % 8 == 0 && d->v_bs % 8 == 0 && d->z_bs % 8 == 0, RNN_ERR_SHAPE);
  return RNN_OK;
}

// operand specs of the batched GEMMs
void spec_heads(long long& ld, long long& bstride, int& per_head, int& hoff, long long rs, long long bs, int hs) {
  ld = rs; bstride = bs; per_head = 1; hoff = hs;
}
void spec_slab(long long& ld, long long& bstride, int& per_head, int& hoff, int rows, int cols) {
  ld = cols; bstride = (long long)rows * cols; per_head = 0; hoff = 0;
}

}  // namespace

int rnn_internal_rng_advance_attn(unsigned long long inc, cudaStream_t stream) {
  rng_advance_attn_kernel<<<1, 1, 0, stream>>>(inc);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

extern "C" {

int rnn_attention_fwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask, void* z,
                      float* lse, int is_f32, cudaStream_t stream) {
  int rc = check(d);
  if (rc != RNN_OK) return rc;
  AttnParams p = to_params(d);
  p.q = q; p.k = k; p.v = v; p.z = z; p.mask = mask; p.lse = lse;
  RNN_REQUIRE((d->mask_kind == 0) == (mask == nullptr), RNN_ERR_DTYPE);
  const dim3 grid(ceil_div_i(d->T, ROWS), d->B * d->H);
  const size_t smem = fwd_smem(d->hs);
  if (is_f32) {
    rc = set_smem(attn_fwd_kernel<float>, smem);
    if (rc != RNN_OK) return rc;
    attn_fwd_kernel<float><<<grid, 128, smem, stream>>>(p);
  } else {
    rc = set_smem(attn_fwd_kernel<bf16>, smem);
    if (rc != RNN_OK) return rc;
    attn_fwd_kernel<bf16><<<grid, 128, smem, stream>>>(p);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int rnn_attention_bwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask,
                      const void* z, const void* dz, const float* lse, float* delta, void* dq, void* dk, void* dv,
                      int is_f32, cudaStream_t stream) {
  int rc = check(d);
  if (rc != RNN_OK) return rc;
  AttnParams p = to_params(d);
  p.q = q; p.k = k; p.v = v; p.z = const_cast<void*>(z); p.mask = mask; p.lse = const_cast<float*>(lse);
  p.dz = dz; p.dq = dq; p.dk = dk; p.dv = dv; p.delta = delta;
  RNN_REQUIRE((d->mask_kind == 0) == (mask == nullptr), RNN_ERR_DTYPE);
  const dim3 gq(ceil_div_i(d->T, ROWS), d->B * d->H), gk(ceil_div_i(d->Tk, ROWS), d->B * d->H);
  const size_t s1 = dq_smem(d->hs), s2 = dkv_smem(d->hs);
  if (is_f32) {
    rc = set_smem(attn_bwd_dq_kernel<float>, s1); if (rc != RNN_OK) return rc;
    rc = set_smem(attn_bwd_dkv_kernel<float>, s2); if (rc != RNN_OK) return rc;
    attn_bwd_dq_kernel<float><<<gq, 128, s1, stream>>>(p);
    RNN_CHECK_LAUNCH();
    attn_bwd_dkv_kernel<float><<<gk, 128, s2, stream>>>(p);
  } else {
    rc = set_smem(attn_bwd_dq_kernel<bf16>, s1); if (rc != RNN_OK) return rc;
    rc = set_smem(attn_bwd_dkv_kernel<bf16>, s2); if (rc != RNN_OK) return rc;
    attn_bwd_dq_kernel<bf16><<<gq, 128, s1, stream>>>(p);
    RNN_CHECK_LAUNCH();
    attn_bwd_dkv_kernel<bf16><<<gk, 128, s2, stream>>>(p);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}


int rnn_attention_tc_supported(const rnn_attn_desc* d) {
  if (d == nullptr || d->hs <= 0) return 0;
  return d->hs % 64 == 0 && d->T % 8 == 0 && d->Tk % 8 == 0 && d->q_rs % 8 == 0 && d->k_rs % 8 == 0 &&
         d->v_rs % 8 == 0 && d->z_rs % 8 == 0 && d->q_bs % 8 == 0 && d->k_bs % 8 == 0 && d->v_bs % 8 == 0 &&
         d->z_bs % 8 == 0;
}
// probabilities kept from forward to backward: P, and dropout(P) when dropout is active (both bf16 [B*H][T][Tk])
size_t rnn_attention_tc_saved_bytes(const rnn_attn_desc* d) {
  if (d == nullptr) return 0;
  const size_t one = (size_t)d->B * d->H * d->T * d->Tk * sizeof(bf16);
  return d->p_drop > 0.f ? 2 * one : one;
}
// scratch of either direction: fp32 scores (forward), fp32 d(probabilities) + bf16 d(scores) (backward)
size_t rnn_attention_tc_workspace(const rnn_attn_desc* d) {
  if (d == nullptr) return 0;
  const size_t n = (size_t)d->B * d->H * d->T * d->Tk;
  return n * sizeof(float) + n * sizeof(bf16);
}

int rnn_attention_tc_fwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask,
                         void* z, void* saved, void* ws, size_t ws_bytes, cudaStream_t stream) {
  int rc = tc_check(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE((d->mask_kind == 0) == (mask == nullptr), RNN_ERR_DTYPE);
  RNN_REQUIRE(saved != nullptr && ws != nullptr && ws_bytes >= rnn_attention_tc_workspace(d), RNN_ERR_WORKSPACE);
  const long long n = (long long)d->B * d->H * d->T * d->Tk;
  float* S = reinterpret_cast<float*>(ws);
  bf16* P = reinterpret_cast<bf16*>(saved);
  bf16* Pd = d->p_drop > 0.f ? P + n : P;
  rnn_bgemm_desc g;
  memset(&g, 0, sizeof(g));
  g.batches = d->B; g.heads = d->H;
  // S = q k^T  (nn/attention.py:37)
  g.M = d->T; g.N = d->Tk; g.K = d->hs; g.out_f32 = 1;
  spec_heads(g.a_ld, g.a_bstride, g.a_per_head, g.a_hoff, d->q_rs, d->q_bs, d->hs);
  spec_heads(g.b_ld, g.b_bstride, g.b_per_head, g.b_hoff, d->k_rs, d->k_bs, d->hs);
  spec_slab(g.d_ld, g.d_bstride, g.d_per_head, g.d_hoff, d->T, d->Tk);
  rc = rnn_gemm_batched_bf16(&g, q, k, S, stream);
  if (rc != RNN_OK) return rc;
  // scale, mask, softmax, dropout (:38-46)
  SmaxParams sp;
  sp.H = d->H; sp.T = d->T; sp.Tk = d->Tk; sp.scale = d->scale; sp.mask = mask; sp.mask_kind = d->mask_kind;
  sp.mask_bs = d->mask_bs; sp.mask_rs = d->mask_rs; sp.p_drop = d->p_drop; sp.seed = d->seed; sp.offset = d->offset;
  const long long rows = (long long)d->B * d->H * d->T;
  attn_softmax_fwd_kernel<bf16><<<ceil_div_i(rows, 4), 128, 0, stream>>>(S, P, Pd, rows, sp);
  RNN_CHECK_LAUNCH();
  // z = w v  (:47), written straight into the concatenated-heads layout
  memset(&g, 0, sizeof(g));
  g.batches = d->B; g.heads = d->H;
  g.M = d->T; g.N = d->hs; g.K = d->Tk; g.b_mn = 1;
  spec_slab(g.a_ld, g.a_bstride, g.a_per_head, g.a_hoff, d->T, d->Tk);
  spec_heads(g.b_ld, g.b_bstride, g.b_per_head, g.b_hoff, d->v_rs, d->v_bs, d->hs);
  spec_heads(g.d_ld, g.d_bstride, g.d_per_head, g.d_hoff, d->z_rs, d->z_bs, d->hs);
  return rnn_gemm_batched_bf16(&g, Pd, v, z, stream);
}

int rnn_attention_tc_bwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* saved,
                         const void* dz, void* dq, void* dk, void* dv, void* ws, size_t ws_bytes,
                         cudaStream_t stream) {
  int rc = tc_check(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(d->dz_rs % 8 == 0 && d->dz_bs % 8 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(saved != nullptr && ws != nullptr && ws_bytes >= rnn_attention_tc_workspace(d), RNN_ERR_WORKSPACE);
  const long long n = (long long)d->B * d->H * d->T * d->Tk;
  float* dPd = reinterpret_cast<float*>(ws);
  bf16* dS = reinterpret_cast<bf16*>(dPd + n);
  const bf16* P = reinterpret_cast<const bf16*>(saved);
  const bf16* Pd = d->p_drop > 0.f ? P + n : P;
  rnn_bgemm_desc g;
  // d(dropout(w)) = dz v^T
  memset(&g, 0, sizeof(g));
  g.batches = d->B; g.heads = d->H;
  g.M = d->T; g.N = d->Tk; g.K = d->hs; g.out_f32 = 1;
  spec_heads(g.a_ld, g.a_bstride, g.a_per_head, g.a_hoff, d->dz_rs, d->dz_bs, d->hs);
  spec_heads(g.b_ld, g.b_bstride, g.b_per_head, g.b_hoff, d->v_rs, d->v_bs, d->hs);
  spec_slab(g.d_ld, g.d_bstride, g.d_per_head, g.d_hoff, d->T, d->Tk);
  rc = rnn_gemm_batched_bf16(&g, dz, v, dPd, stream);
  if (rc != RNN_OK) return rc;
  // dv = dropout(w)^T dz
  memset(&g, 0, sizeof(g));
  g.batches = d->B; g.heads = d->H;
  g.M = d->Tk; g.N = d->hs; g.K = d->T; g.a_mn = 1; g.b_mn = 1;
  spec_slab(g.a_ld, g.a_bstride, g.a_per_head, g.a_hoff, d->T, d->Tk);
  spec_heads(g.b_ld, g.b_bstride, g.b_per_head, g.b_hoff, d->dz_rs, d->dz_bs, d->hs);
  spec_heads(g.d_ld, g.d_bstride, g.d_per_head, g.d_hoff, d->v_rs, d->v_bs, d->hs);
  rc = rnn_gemm_batched_bf16(&g, Pd, dz, dv, stream);
  if (rc != RNN_OK) return rc;
  // softmax / dropout / scale backward
  const long long rows = (long long)d->B * d->H * d->T;
  attn_softmax_bwd_kernel<bf16><<<ceil_div_i(rows, 4), 128, 0, stream>>>(P, Pd, dPd, dS, rows, d->Tk, d->scale, d->p_drop);
  RNN_CHECK_LAUNCH();
  // dq = ds k
  memset(&g, 0, sizeof(g));
  g.batches = d->B; g.heads = d->H;
  g.M = d->T; g.N = d->hs; g.K = d->Tk; g.b_mn = 1;
  spec_slab(g.a_ld, g.a_bstride, g.a_per_head, g.a_hoff, d->T, d->Tk);
  spec_heads(g.b_ld, g.b_bstride, g.b_per_head, g.b_hoff, d->k_rs, d->k_bs, d->hs);
  spec_heads(g.d_ld, g.d_bstride, g.d_per_head, g.d_hoff, d->q_rs, d->q_bs, d->hs);
  rc = rnn_gemm_batched_bf16(&g, dS, k, dq, stream);
  if (rc != RNN_OK) return rc;
  // dk = ds^T q
  memset(&g, 0, sizeof(g));
  g.batches = d->B; g.heads = d->H;
  g.M = d->Tk; g.N = d->hs; g.K = d->T; g.a_mn = 1; g.b_mn = 1;
  spec_slab(g.a_ld, g.a_bstride, g.a_per_head, g.a_hoff, d->T, d->Tk);
  spec_heads(g.b_ld, g.b_bstride, g.b_per_head, g.b_hoff, d->q_rs, d->q_bs, d->hs);
  spec_heads(g.d_ld, g.d_bstride, g.d_per_head, g.d_hoff, d->k_rs, d->k_bs, d->hs);
  return rnn_gemm_batched_bf16(&g, dS, q, dk, stream);
}

// ---- fused tcgen05 attention (scores stay in TMEM / shared memory; see attention_fused.cuh) ----------------------
int rnn_attention_fused_supported(const rnn_attn_desc* d) { return fused_ok(d); }

int rnn_attention_fused_fwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask,
                            void* z, float* lse, cudaStream_t stream) {
  int rc = check(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(fused_ok(d), RNN_ERR_SHAPE);
  RNN_REQUIRE((d->mask_kind == 0) == (mask == nullptr), RNN_ERR_DTYPE);
  if (d->hs == 64) return fused_fwd_launch<64>(d, q, k, v, mask, z, lse, stream);
  return fused_fwd_launch<128>(d, q, k, v, mask, z, lse, stream);
}

int rnn_attention_fused_bwd(const rnn_attn_desc* d, const void* q, const void* k, const void* v, const void* mask,
                            const void* z, const void* dz, const float* lse, float* delta, void* dq, void* dk, void* dv,
                            cudaStream_t stream) {
  int rc = check(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(fused_ok(d) && d->dz_rs % 8 == 0 && d->dz_bs % 8 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE((d->mask_kind == 0) == (mask == nullptr), RNN_ERR_DTYPE);
  RNN_REQUIRE(lse != nullptr && delta != nullptr, RNN_ERR_WORKSPACE);
  if (d->hs == 64) return fused_bwd_launch<64>(d, q, k, v, mask, z, dz, lse, delta, dq, dk, dv, stream);
  return fused_bwd_launch<128>(d, q, k, v, mask, z, dz, lse, delta, dq, dk, dv, stream);
}

// ---- fp32 attention on the tensor cores at fp32 accuracy (three-piece bf16 splits, see above) -----------------
namespace {
struct F32Plan {
  long long n;          // B*H*T*Tk
  size_t x, big, f, o;  // byte sizes of: a small split operand, a probability-sized split operand, an fp32 score
                        // tensor, an fp32 output slab set
  size_t total;
};
size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }
F32Plan f32_plan(const rnn_attn_desc* d) {
  F32Plan p;
  const long long nz = (long long)d->B * d->H, tm = d->T > d->Tk ? d->T : d->Tk;
  p.n = nz * d->T * d->Tk;
  p.x = up256((size_t)nz * 6 * tm * d->hs * sizeof(bf16));
  p.big = up256((size_t)p.n * 6 * sizeof(bf16));
  p.f = up256((size_t)p.n * sizeof(float));
  p.o = up256((size_t)nz * tm * d->hs * sizeof(float));
  p.total = 2 * p.x + p.big + p.f + p.o;
  return p;
}
int f32_tc_check(const rnn_attn_desc* d) {
  int rc = check(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(d->hs % 8 == 0 && d->T % 8 == 0 && d->Tk % 8 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->q_rs % 4 == 0 && d->k_rs % 4 == 0 && d->v_rs % 4 == 0 && d->z_rs % 4 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->q_bs % 4 == 0 && d->k_bs % 4 == 0 && d->v_bs % 4 == 0 && d->z_bs % 4 == 0, RNN_ERR_SHAPE);
  return RNN_OK;
}
}  // namespace

int rnn_attention_f32_tc_supported(const rnn_attn_desc* d) {
  if (d == nullptr || d->hs <= 0) return 0;
  return d->hs % 8 == 0 && d->hs <= 32 * MAXD && d->T % 8 == 0 && d->Tk % 8 == 0 && d->q_rs % 4 == 0 &&
         d->k_rs % 4 == 0 && d->v_rs % 4 == 0 && d->z_rs % 4 == 0 && d->q_bs % 4 == 0 && d->k_bs % 4 == 0 &&
         d->v_bs % 4 == 0 && d->z_bs % 4 == 0;
}
// fp32 probabilities kept from forward to backward: P, and dropout(P) when dropout is active
size_t rnn_attention_f32_tc_saved_bytes(const rnn_attn_desc* d) {
  if (d == nullptr) return 0;
  const size_t one = (size_t)d->B * d->H * d->T * d->Tk * sizeof(float);
  return d->p_drop > 0.f ? 2 * one : one;
}
size_t rnn_attention_f32_tc_workspace(const rnn_attn_desc* d) { return d == nullptr ? 0 : f32_plan(d).total; }

int rnn_attention_f32_tc_fwd(const rnn_attn_desc* d, const float* q, const float* k, const float* v, const void* mask,
                             float* z, void* saved, void* ws, size_t ws_bytes, cudaStream_t stream) {
  int rc = f32_tc_check(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE((d->mask_kind == 0) == (mask == nullptr), RNN_ERR_DTYPE);
  const F32Plan pl = f32_plan(d);
  RNN_REQUIRE(saved != nullptr && ws != nullptr && ws_bytes >= pl.total, RNN_ERR_WORKSPACE);
  unsigned char* w = reinterpret_cast<unsigned char*>(ws);
  bf16* X1 = reinterpret_cast<bf16*>(w);
  bf16* X2 = reinterpret_cast<bf16*>(w + pl.x);
  bf16* BIG = reinterpret_cast<bf16*>(w + 2 * pl.x);
  float* S = reinterpret_cast<float*>(w + 2 * pl.x + pl.big);
  float* O = reinterpret_cast<float*>(w + 2 * pl.x + pl.big + pl.f);
  float* P = reinterpret_cast<float*>(saved);
  float* Pd = d->p_drop > 0.f ? P + pl.n : P;
  const int nz = d->B * d->H;
  // S = q k^T  (nn/attention.py:37)
  rc = launch_split(q, d->q_bs, d->q_rs, d->B, d->H, d->T, d->hs, X1, 0, 0, stream);
  if (rc != RNN_OK) return rc;
  rc = launch_split(k, d->k_bs, d->k_rs, d->B, d->H, d->Tk, d->hs, X2, 0, 1, stream);
  if (rc != RNN_OK) return rc;
  rc = slab_gemm(nz, d->T, d->Tk, d->hs, 0, 0, X1, X2, S, stream);
  if (rc != RNN_OK) return rc;
  // scale, mask, softmax, dropout (:38-46)
  SmaxParams sp;
  sp.H = d->H; sp.T = d->T; sp.Tk = d->Tk; sp.scale = d->scale; sp.mask = mask; sp.mask_kind = d->mask_kind;
  sp.mask_bs = d->mask_bs; sp.mask_rs = d->mask_rs; sp.p_drop = d->p_drop; sp.seed = d->seed; sp.offset = d->offset;
  const long long rows = (long long)nz * d->T;
  attn_softmax_fwd_kernel<float><<<ceil_div_i(rows, 4), 128, 0, stream>>>(S, P, Pd, rows, sp);
  RNN_CHECK_LAUNCH();
  // z = w v  (:47)
  rc = launch_split(Pd, (long long)d->T * d->Tk, d->Tk, nz, 1, d->T, d->Tk, BIG, 0, 0, stream);
  if (rc != RNN_OK) return rc;
  rc = launch_split(v, d->v_bs, d->v_rs, d->B, d->H, d->Tk, d->hs, X1, 1, 1, stream);
  if (rc != RNN_OK) return rc;
  rc = slab_gemm(nz, d->T, d->hs, d->Tk, 0, 1, BIG, X1, O, stream);
  if (rc != RNN_OK) return rc;
  return launch_gather(O, z, d->z_bs, d->z_rs, d->B, d->H, d->T, d->hs, stream);
}

int rnn_attention_f32_tc_bwd(const rnn_attn_desc* d, const float* q, const float* k, const float* v,
                             const void* saved, const float* dz, float* dq, float* dk, float* dv, void* ws,
                             size_t ws_bytes, cudaStream_t stream) {
  int rc = f32_tc_check(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(d->dz_rs % 4 == 0 && d->dz_bs % 4 == 0, RNN_ERR_SHAPE);
  const F32Plan pl = f32_plan(d);
  RNN_REQUIRE(saved != nullptr && ws != nullptr && ws_bytes >= pl.total, RNN_ERR_WORKSPACE);
  unsigned char* w = reinterpret_cast<unsigned char*>(ws);
  bf16* X1 = reinterpret_cast<bf16*>(w);
  bf16* X2 = reinterpret_cast<bf16*>(w + pl.x);
  bf16* BIG = reinterpret_cast<bf16*>(w + 2 * pl.x);
  float* G = reinterpret_cast<float*>(w + 2 * pl.x + pl.big);     // d(dropout(w)), then ds in place
  float* O = reinterpret_cast<float*>(w + 2 * pl.x + pl.big + pl.f);
  const float* P = reinterpret_cast<const float*>(saved);
  const float* Pd = d->p_drop > 0.f ? P + pl.n : P;
  const int nz = d->B * d->H;
  const long long pslab = (long long)d->T * d->Tk;
  // d(dropout(w)) = dz v^T
  rc = launch_split(dz, d->dz_bs, d->dz_rs, d->B, d->H, d->T, d->hs, X1, 0, 0, stream);
  if (rc != RNN_OK) return rc;
  rc = launch_split(v, d->v_bs, d->v_rs, d->B, d->H, d->Tk, d->hs, X2, 0, 1, stream);
  if (rc != RNN_OK) return rc;
  rc = slab_gemm(nz, d->T, d->Tk, d->hs, 0, 0, X1, X2, G, stream);
  if (rc != RNN_OK) return rc;
  // dv = dropout(w)^T dz
  rc = launch_split(Pd, pslab, d->Tk, nz, 1, d->T, d->Tk, BIG, 1, 0, stream);
  if (rc != RNN_OK) return rc;
  rc = launch_split(dz, d->dz_bs, d->dz_rs, d->B, d->H, d->T, d->hs, X1, 1, 1, stream);
  if (rc != RNN_OK) return rc;
  rc = slab_gemm(nz, d->Tk, d->hs, d->T, 1, 1, BIG, X1, O, stream);
  if (rc != RNN_OK) return rc;
  rc = launch_gather(O, dv, d->v_bs, d->v_rs, d->B, d->H, d->Tk, d->hs, stream);
  if (rc != RNN_OK) return rc;
  // softmax / dropout / scale backward, in place over G (every element is read by its own thread before it is
  // written; the row sum is complete before the first write)
  const long long rows = (long long)nz * d->T;
  attn_softmax_bwd_kernel<float><<<ceil_div_i(rows, 4), 128, 0, stream>>>(P, Pd, G, G, rows, d->Tk, d->scale, d->p_drop);
  RNN_CHECK_LAUNCH();
  // dq = ds k
  rc = launch_split(G, pslab, d->Tk, nz, 1, d->T, d->Tk, BIG, 0, 0, stream);
  if (rc != RNN_OK) return rc;
  rc = launch_split(k, d->k_bs, d->k_rs, d->B, d->H, d->Tk, d->hs, X1, 1, 1, stream);
  if (rc != RNN_OK) return rc;
  rc = slab_gemm(nz, d->T, d->hs, d->Tk, 0, 1, BIG, X1, O, stream);
  if (rc != RNN_OK) return rc;
  rc = launch_gather(O, dq, d->q_bs, d->q_rs, d->B, d->H, d->T, d->hs, stream);
  if (rc != RNN_OK) return rc;
  // dk = ds^T q
  rc = launch_split(G, pslab, d->Tk, nz, 1, d->T, d->Tk, BIG, 1, 0, stream);
  if (rc != RNN_OK) return rc;
  rc = launch_split(q, d->q_bs, d->q_rs, d->B, d->H, d->T, d->hs, X1, 1, 1, stream);
  if (rc != RNN_OK) return rc;
  rc = slab_gemm(nz, d->Tk, d->hs, d->T, 1, 1, BIG, X1, O, stream);
  if (rc != RNN_OK) return rc;
  return launch_gather(O, dk, d->k_bs, d->k_rs, d->B, d->H, d->Tk, d->hs, stream);
}
}  // extern "C"
