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

#include "common.cuh"

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

__device__ __forceinline__ uint32_t philox_first(unsigned long long seed, unsigned long long offset,
                                                 unsigned long long idx) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  uint32_t c0 = (uint32_t)idx, c1 = (uint32_t)(idx >> 32), c2 = (uint32_t)offset, c3 = (uint32_t)(offset >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0, hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  return c0;
}

__device__ __forceinline__ float keep_scale(const AttnParams& p, int bh, int i, int j) {
  if (p.p_drop <= 0.f) return 1.f;
  const unsigned long long idx = ((unsigned long long)bh * p.T + i) * p.Tk + j;
  const uint32_t r = philox_first(p.seed, p.offset, idx);
  return r >= (uint32_t)(p.p_drop * 4294967296.0) ? 1.f / (1.f - p.p_drop) : 0.f;
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

}  // namespace

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

}  // extern "C"
