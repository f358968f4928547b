// Pooling over NHWC tensors (torch.nn.MaxPool2d / AvgPool2d / AdaptiveAvgPool2d((1,1)) semantics).
// Max pooling returns the argmax as torch does: the flat index h*W + w inside the input plane,
// first maximum in row-major window order, NaN wins.  Indices are int32 here (torch: int64).
#include "common.cuh"

namespace {

template <typename T, bool VEC> struct PV;
template <> struct PV<bf16, false> {
  static constexpr int W = 1;
  __device__ static void load(const bf16* p, float* f) { f[0] = __bfloat162float(*p); }
  __device__ static void store(bf16* p, const float* f) { *p = __float2bfloat16(f[0]); }
};
template <> struct PV<float, false> {
  static constexpr int W = 1;
  __device__ static void load(const float* p, float* f) { f[0] = *p; }
  __device__ static void store(float* p, const float* f) { *p = f[0]; }
};
template <> struct PV<bf16, true> {
  static constexpr int W = 8;
  __device__ static void load(const bf16* p, float* f) { unpack8(*reinterpret_cast<const uint4*>(p), f); }
  __device__ static void store(bf16* p, const float* f) { *reinterpret_cast<uint4*>(p) = pack8(f); }
};
template <> struct PV<float, true> {
  static constexpr int W = 4;
  __device__ static void load(const float* p, float* f) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
  __device__ static void store(float* p, const float* f) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  }
};

struct PoolGeom {
  int N, H, W, C, R, S, sh, sw, ph, pw, P, Q;
};

// Element-group index -> (n, a, b, cg) for a [N, A, B, cgs] iteration space.  32-bit arithmetic whenever the
// space fits (64-bit integer division costs ~10x more and these kernels decode one index per 16 bytes moved).
template <typename I>
__device__ __forceinline__ void decode4(long long i, int cgs, int B, int A, int& n, int& a, int& b, int& cg) {
  I t = (I)i;
  cg = (int)(t % (I)cgs); t /= (I)cgs;
  b = (int)(t % (I)B); t /= (I)B;
  a = (int)(t % (I)A);
  n = (int)(t / (I)A);
}

// window position (1 byte per element) packed VW to a word
template <int VW> struct WIdx;
template <> struct WIdx<1> {
  __device__ static void store(unsigned char* p, const int* l) { p[0] = (unsigned char)l[0]; }
  __device__ static void load(const unsigned char* p, int* l) { l[0] = p[0]; }
};
template <> struct WIdx<4> {
  __device__ static void store(unsigned char* p, const int* l) {
    *reinterpret_cast<unsigned*>(p) = (unsigned)l[0] | ((unsigned)l[1] << 8) | ((unsigned)l[2] << 16) | ((unsigned)l[3] << 24);
  }
  __device__ static void load(const unsigned char* p, int* l) {
    const unsigned v = *reinterpret_cast<const unsigned*>(p);
#pragma unroll
    for (int e = 0; e < 4; ++e) l[e] = (v >> (8 * e)) & 0xffu;
  }
};
template <> struct WIdx<8> {
  __device__ static void store(unsigned char* p, const int* l) {
    uint2 v;
    v.x = (unsigned)l[0] | ((unsigned)l[1] << 8) | ((unsigned)l[2] << 16) | ((unsigned)l[3] << 24);
    v.y = (unsigned)l[4] | ((unsigned)l[5] << 8) | ((unsigned)l[6] << 16) | ((unsigned)l[7] << 24);
    *reinterpret_cast<uint2*>(p) = v;
  }
  __device__ static void load(const unsigned char* p, int* l) {
    const uint2 v = *reinterpret_cast<const uint2*>(p);
#pragma unroll
    for (int e = 0; e < 4; ++e) { l[e] = (v.x >> (8 * e)) & 0xffu; l[4 + e] = (v.y >> (8 * e)) & 0xffu; }
  }
};

// Max pooling runs one grid row per output (fwd) / input (bwd) image row: blockIdx.y = row, blockIdx.z = image,
// so a thread decodes only (column, channel group) -- the flat-index version spent ~250 instructions per 16 bytes
// on integer divisions and was instruction-bound at 1 TB/s.
template <typename T, bool VEC, int FX>
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int* __restrict__ idx,
                   unsigned char* __restrict__ widx, PoolGeom g, int cgs_shift,
                   const float* __restrict__ scale = nullptr, const float* __restrict__ shift = nullptr) {
  constexpr int VW = PV<T, VEC>::W;
  // FX != 0: the window geometry is a compile-time constant (1: 3x3 stride 2 pad 1, 2: 2x2 stride 2, 3: 3x3 stride 2
  // pad 0 -- the pools of ResNet, VGG and AlexNet), which turns the divisions below into shifts
  const int gR = FX == 2 ? 2 : (FX ? 3 : g.R), gS = FX == 2 ? 2 : (FX ? 3 : g.S);
  const int gsh = FX ? 2 : g.sh, gsw = FX ? 2 : g.sw;
  const int gph = FX == 1 ? 1 : (FX ? 0 : g.ph), gpw = FX == 1 ? 1 : (FX ? 0 : g.pw);
  const int cgs = g.C / VW;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= g.Q * cgs) return;
  const int q = cgs_shift >= 0 ? (e >> cgs_shift) : (int)((unsigned)e / (unsigned)cgs);
  const int cg = e - q * cgs;
  const int p = blockIdx.y, n = blockIdx.z;
  float best[VW];
  int bl[VW];   // argmax as the position inside the (unclipped) R x S window
  const int h0 = p * gsh - gph, w0 = q * gsw - gpw;
  const int hs = h0 < 0 ? 0 : h0, ws = w0 < 0 ? 0 : w0;
  const int he = min(h0 + gR, g.H), we = min(w0 + gS, g.W);
#pragma unroll
  for (int e2 = 0; e2 < VW; ++e2) { best[e2] = -INFINITY; bl[e2] = (hs - h0) * gS + (ws - w0); }
  // scale != nullptr: the input is a pre-BatchNorm activation and the pooled quantity is relu(x * scale + shift)
  // rounded to the storage type first -- exactly what a separate BatchNorm-apply kernel would have stored, so values
  // AND argmax ties are those of the unfused pair (the ResNet stem: saves writing and re-reading that tensor)
  float sc[VW], sf[VW];
  if (scale != nullptr) {
#pragma unroll
    for (int e2 = 0; e2 < VW; ++e2) { sc[e2] = scale[cg * VW + e2]; sf[e2] = shift[cg * VW + e2]; }
  }
  const T* xn = x + (long long)n * g.H * g.W * g.C + cg * VW;
  for (int h = hs; h < he; ++h)
    for (int w = ws; w < we; ++w) {
      float v[VW];
      PV<T, VEC>::load(xn + ((long long)h * g.W + w) * g.C, v);
      if (scale != nullptr) {
        float t[VW];
#pragma unroll
        for (int e2 = 0; e2 < VW; ++e2) t[e2] = fmaxf(fmaf(v[e2], sc[e2], sf[e2]), 0.f);
        __align__(16) T tmp[VW];
        PV<T, VEC>::store(tmp, t);      // round to the storage type
        PV<T, VEC>::load(tmp, v);
      }
      const int local = (h - h0) * gS + (w - w0);
#pragma unroll
      for (int e2 = 0; e2 < VW; ++e2)
        if (v[e2] > best[e2] || v[e2] != v[e2]) { best[e2] = v[e2]; bl[e2] = local; }
    }
  const long long o = (((long long)n * g.P + p) * g.Q + q) * g.C + cg * VW;
  PV<T, VEC>::store(y + o, best);
  if (idx) {  // torch's flat index h * W + w inside the input plane
#pragma unroll
    for (int e2 = 0; e2 < VW; ++e2) {
      const int lh = bl[e2] / gS, lw = bl[e2] - lh * gS;
      idx[o + e2] = (h0 + lh) * g.W + (w0 + lw);
    }
  }
  if (widx) WIdx<VW>::store(widx + o, bl);  // 1 byte instead of 4 for the backward pass
}

// Gather form of the backward pass: every input element sums the dy of the windows whose argmax it is.
template <typename T, bool VEC, int FX>
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const T* __restrict__ dy, const int* __restrict__ idx, const unsigned char* __restrict__ widx,
                   T* __restrict__ dx, PoolGeom g, int cgs_shift, const T* __restrict__ ypool = nullptr) {
  constexpr int VW = PV<T, VEC>::W;
  // FX != 0: the window geometry is a compile-time constant (1: 3x3 stride 2 pad 1, 2: 2x2 stride 2, 3: 3x3 stride 2
  // pad 0 -- the pools of ResNet, VGG and AlexNet), which turns the divisions below into shifts
  const int gR = FX == 2 ? 2 : (FX ? 3 : g.R), gS = FX == 2 ? 2 : (FX ? 3 : g.S);
  const int gsh = FX ? 2 : g.sh, gsw = FX ? 2 : g.sw;
  const int gph = FX == 1 ? 1 : (FX ? 0 : g.ph), gpw = FX == 1 ? 1 : (FX ? 0 : g.pw);
  const int cgs = g.C / VW;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= g.W * cgs) return;
  const int w = cgs_shift >= 0 ? (e >> cgs_shift) : (int)((unsigned)e / (unsigned)cgs);
  const int cg = e - w * cgs;
  const int h = blockIdx.y, n = blockIdx.z;
  const int me = h * g.W + w;
  float acc[VW];
#pragma unroll
  for (int e2 = 0; e2 < VW; ++e2) acc[e2] = 0.f;
  // windows p with p*sh - ph <= h < p*sh - ph + R (the row bounds are the same for the whole block)
  int p_lo = (h + gph - gR + gsh) / gsh; if (h + gph - gR + 1 <= 0) p_lo = 0;
  int p_hi = (h + gph) / gsh; if (p_hi > g.P - 1) p_hi = g.P - 1;
  int q_lo = (w + gpw - gS + gsw) / gsw; if (w + gpw - gS + 1 <= 0) q_lo = 0;
  int q_hi = (w + gpw) / gsw; if (q_hi > g.Q - 1) q_hi = g.Q - 1;
  for (int p = p_lo; p <= p_hi; ++p)
    for (int q = q_lo; q <= q_hi; ++q) {
      const long long o = (((long long)n * g.P + p) * g.Q + q) * g.C + cg * VW;
      float d[VW];
      PV<T, VEC>::load(dy + o, d);
      if (ypool) {   // ReLU backward of the pooled tensor's producer: the gradient passes where the pooled value > 0
        float yv[VW];
        PV<T, VEC>::load(ypool + o, yv);
#pragma unroll
        for (int e2 = 0; e2 < VW; ++e2) d[e2] = yv[e2] > 0.f ? d[e2] : 0.f;
      }
      if (widx) {
        const int local = (h - (p * gsh - gph)) * gS + (w - (q * gsw - gpw));
        int l[VW];
        WIdx<VW>::load(widx + o, l);
#pragma unroll
        for (int e2 = 0; e2 < VW; ++e2)
          if (l[e2] == local) acc[e2] += d[e2];
      } else {
#pragma unroll
        for (int e2 = 0; e2 < VW; ++e2)
          if (idx[o + e2] == me) acc[e2] += d[e2];
      }
    }
  PV<T, VEC>::store(dx + (((long long)n * g.H + h) * g.W + w) * g.C + cg * VW, acc);
}

// Backward of NON-OVERLAPPING windows (kernel == stride, no padding; VGG's 2x2/2 pools): every input element belongs
// to exactly one window, so a thread owns one window, reads its dy / argmax once and writes the R*S input elements
// (dy at the argmax, zero elsewhere).  Rows / columns beyond the last full window get zeros from the tail threads.
template <typename T, int K>
__global__ void __launch_bounds__(256)
maxpool_bwd_tiled_kernel(const T* __restrict__ dy, const unsigned char* __restrict__ widx, T* __restrict__ dx,
                         PoolGeom g, int cgs_shift, const T* __restrict__ ypool = nullptr) {
  constexpr int VW = PV<T, true>::W;
  const int cgs = g.C / VW;
  const int Qx = (g.W + K - 1) / K;                 // windows per row including a partial last one
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= Qx * cgs) return;
  const int q = cgs_shift >= 0 ? (e >> cgs_shift) : (int)((unsigned)e / (unsigned)cgs);
  const int cg = e - q * cgs;
  const int p = blockIdx.y, n = blockIdx.z;         // p over ceil(H / K)
  float d[VW];
  int l[VW];
  const bool full = p < g.P && q < g.Q;
  if (full) {
    const long long o = (((long long)n * g.P + p) * g.Q + q) * g.C + cg * VW;
    PV<T, true>::load(dy + o, d);
    WIdx<VW>::load(widx + o, l);
    if (ypool) {
      float yv[VW];
      PV<T, true>::load(ypool + o, yv);
#pragma unroll
      for (int e2 = 0; e2 < VW; ++e2) d[e2] = yv[e2] > 0.f ? d[e2] : 0.f;
    }
  }
#pragma unroll
  for (int r = 0; r < K; ++r)
#pragma unroll
    for (int sx = 0; sx < K; ++sx) {
      const int h = p * K + r, w = q * K + sx;
      if (h >= g.H || w >= g.W) continue;
      float o[VW];
#pragma unroll
      for (int e2 = 0; e2 < VW; ++e2) o[e2] = (full && l[e2] == r * K + sx) ? d[e2] : 0.f;
      PV<T, true>::store(dx + (((long long)n * g.H + h) * g.W + w) * g.C + cg * VW, o);
    }
}

// Backward of the 3x3 / stride 2 / pad 1 pool (ResNet stem): a thread owns the 2x2 input quad (2p+a, 2q+b) and the
// four windows (p..p+1, q..q+1) that can cover it -- 4 window loads per quad instead of 9 (2.25 per element), and
// the positions inside the windows are constants.  Sums run in the same window order as the generic kernel.
template <typename T>
__global__ void __launch_bounds__(256)
maxpool_bwd_3x3s2_kernel(const T* __restrict__ dy, const unsigned char* __restrict__ widx, T* __restrict__ dx,
                         PoolGeom g, int cgs_shift, const T* __restrict__ ypool = nullptr) {
  constexpr int VW = PV<T, true>::W;
  const int cgs = g.C / VW;
  const int Qx = (g.W + 1) / 2;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= Qx * cgs) return;
  const int q = cgs_shift >= 0 ? (e >> cgs_shift) : (int)((unsigned)e / (unsigned)cgs);
  const int cg = e - q * cgs;
  const int p = blockIdx.y, n = blockIdx.z;
  float d[2][2][VW];
  int l[2][2][VW];
  bool ok[2][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      ok[a][b] = (p + a) < g.P && (q + b) < g.Q;
      if (ok[a][b]) {
        const long long o = (((long long)n * g.P + (p + a)) * g.Q + (q + b)) * g.C + cg * VW;
        PV<T, true>::load(dy + o, d[a][b]);
        WIdx<VW>::load(widx + o, l[a][b]);
        if (ypool) {
          float yv[VW];
          PV<T, true>::load(ypool + o, yv);
#pragma unroll
          for (int e2 = 0; e2 < VW; ++e2) d[a][b][e2] = yv[e2] > 0.f ? d[a][b][e2] : 0.f;
        }
      }
    }
  // position of input (2p+a, 2q+b) inside window (p+wa, q+wb): row a+1-2*wa, column b+1-2*wb (only 0..2 is inside)
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int h = 2 * p + a, w = 2 * q + b;
      if (h >= g.H || w >= g.W) continue;
      float acc[VW];
#pragma unroll
      for (int e2 = 0; e2 < VW; ++e2) acc[e2] = 0.f;
#pragma unroll
      for (int wa = 0; wa <= a; ++wa)
#pragma unroll
        for (int wb = 0; wb <= b; ++wb) {
          if (!ok[wa][wb]) continue;
          const int local = (a + 1 - 2 * wa) * 3 + (b + 1 - 2 * wb);
#pragma unroll
          for (int e2 = 0; e2 < VW; ++e2)
            if (l[wa][wb][e2] == local) acc[e2] += d[wa][wb][e2];
        }
      PV<T, true>::store(dx + (((long long)n * g.H + h) * g.W + w) * g.C + cg * VW, acc);
    }
}

// ---- marching forms of the 3x3 / stride 2 / pad 1 pool (ResNet stem, bf16, C % 8 == 0) --------------------------
// A thread owns one (column, channel group) and walks MROWS rows downwards; what two neighbouring rows share stays in
// registers.  Results (values, argmax positions, summation order) are those of the kernels above, bit for bit.
//   forward : input row 2p+1 is the last row of window p and the first of window p+1 -- its 3-wide row maximum (value
//             and column; "first maximum, NaN wins" composes over rows) is carried, so a window costs 6 loads (and 6
//             BatchNorm+ReLU evaluations in the fused form) instead of 9;
//   backward: window row p+1 serves the input quads of rows p and p+1 -- 2 window loads per quad instead of 4.
// Measured on [256,112,112,64] (tools/bench_pool.py): backward 255 -> 134 us (4.2 TB/s), forward 211 -> 134 us,
// BatchNorm+ReLU-fused forward 285 -> 189 us.
constexpr int MROWS = 4;

// The forward kernel compares in packed bf16x2 (the values ARE bf16, so the packed compare is exact): one HSET2 pair
// + three LOP3 per two channels and window position instead of ~12 scalar instructions -- the scalar form ran at 123
// registers and 780 instructions per output vector (ncu), i.e. instruction-bound at 2 TB/s.
__device__ __forceinline__ unsigned pool_take_mask(unsigned v, unsigned best) {   // v > best || isnan(v), per half
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&v);
  const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&best);
  return __hgt2_mask(a, b) | __hneu2_mask(a, a);
}
__device__ __forceinline__ unsigned pool_sel(unsigned m, unsigned a, unsigned b) { return (a & m) | (b & ~m); }

// 3-wide maximum of one input row for 8 channels: rv[4] packed values, rc[4] the column (0..2) in each 16-bit half
template <bool BN>
__device__ __forceinline__ void pool_row_max(const bf16* __restrict__ xrow, int w0, int W, int C, const float* sc,
                                             const float* sf, unsigned* rv, unsigned* rc) {
  uint4 raw[3];
  bool in[3];
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    const int w = w0 + s;
    in[s] = w >= 0 && w < W;
    if (in[s]) raw[s] = *reinterpret_cast<const uint4*>(xrow + (long long)w * C);
  }
  const unsigned first = w0 < 0 ? 0x00010001u : 0u;
#pragma unroll
  for (int k = 0; k < 4; ++k) { rv[k] = 0xff80ff80u; rc[k] = first; }   // -inf, first column inside the image
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    if (!in[s]) continue;
    unsigned v[4] = {raw[s].x, raw[s].y, raw[s].z, raw[s].w};
    if (BN) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {   // what a separate BatchNorm-apply kernel would have stored (rounded to bf16)
        const float lo = fmaxf(fmaf(__uint_as_float(v[k] << 16), sc[2 * k], sf[2 * k]), 0.f);
        const float hi = fmaxf(fmaf(__uint_as_float(v[k] & 0xffff0000u), sc[2 * k + 1], sf[2 * k + 1]), 0.f);
        const __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
        v[k] = *reinterpret_cast<const unsigned*>(&t);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned m = pool_take_mask(v[k], rv[k]);
      rv[k] = pool_sel(m, v[k], rv[k]);
      rc[k] = pool_sel(m, (unsigned)s * 0x00010001u, rc[k]);
    }
  }
}

template <bool BN, int PAD>
__global__ void __launch_bounds__(256, BN ? 3 : 4)   // measured: fused 189 us at 3 (235 at 4: spills), plain 134 us at 4
maxpool_fwd_3x3s2_march_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int* __restrict__ idx,
                               unsigned char* __restrict__ widx, PoolGeom g, int cgs_shift,
                               const float* __restrict__ scale, const float* __restrict__ shift) {
  const int cgs = g.C >> 3;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= g.Q * cgs) return;
  const int q = cgs_shift >= 0 ? (e >> cgs_shift) : (int)((unsigned)e / (unsigned)cgs);
  const int cg = e - q * cgs;
  const int n = blockIdx.z, p0 = blockIdx.y * MROWS;
  const int p1 = min(p0 + MROWS, g.P);
  const int w0 = 2 * q - PAD;
  float sc[8], sf[8];
  if (BN) {
#pragma unroll
    for (int e2 = 0; e2 < 8; ++e2) { sc[e2] = scale[cg * 8 + e2]; sf[e2] = shift[cg * 8 + e2]; }
  }
  const bf16* xn = x + (long long)n * g.H * g.W * g.C + cg * 8;
  const long long rowb = (long long)g.W * g.C;
  unsigned cv[4], cc[4];
  bool have = PAD == 0 || p0 > 0;           // window p0's top row 2*p0 - PAD exists
  if (have) pool_row_max<BN>(xn + (long long)(2 * p0 - PAD) * rowb, w0, g.W, g.C, sc, sf, cv, cc);
  for (int p = p0; p < p1; ++p) {
    const int h0 = 2 * p - PAD;
    unsigned best[4], bl[4];                // bl: position inside the 3x3 window, per 16-bit half
    const unsigned first = (unsigned)((h0 < 0 ? 3 : 0) + (w0 < 0 ? 1 : 0)) * 0x00010001u;
#pragma unroll
    for (int k = 0; k < 4; ++k) { best[k] = 0xff80ff80u; bl[k] = first; }
    if (have) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const unsigned m = pool_take_mask(cv[k], best[k]);
        best[k] = pool_sel(m, cv[k], best[k]);
        bl[k] = pool_sel(m, cc[k], bl[k]);
      }
    }
    unsigned rv[4], rc[4];
    pool_row_max<BN>(xn + (long long)(h0 + 1) * rowb, w0, g.W, g.C, sc, sf, rv, rc);   // middle row: always inside
    have = h0 + 2 < g.H;
    if (have) pool_row_max<BN>(xn + (long long)(h0 + 2) * rowb, w0, g.W, g.C, sc, sf, cv, cc);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned m = pool_take_mask(rv[k], best[k]);
      best[k] = pool_sel(m, rv[k], best[k]);
      bl[k] = pool_sel(m, rc[k] + 0x00030003u, bl[k]);
    }
    if (have) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const unsigned m = pool_take_mask(cv[k], best[k]);
        best[k] = pool_sel(m, cv[k], best[k]);
        bl[k] = pool_sel(m, cc[k] + 0x00060006u, bl[k]);
      }
    }
    const long long o = (((long long)n * g.P + p) * g.Q + q) * g.C + cg * 8;
    *reinterpret_cast<uint4*>(y + o) = make_uint4(best[0], best[1], best[2], best[3]);
    if (idx) {
#pragma unroll
      for (int e2 = 0; e2 < 8; ++e2) {
        const int l = (int)((bl[e2 >> 1] >> (16 * (e2 & 1))) & 0xffu);
        const int lh = l / 3, lw = l - lh * 3;
        idx[o + e2] = (h0 + lh) * g.W + (w0 + lw);
      }
    }
    if (widx)   // bytes 0 and 2 of each position word -> one byte per channel
      *reinterpret_cast<uint2*>(widx + o) = make_uint2(__byte_perm(bl[0], bl[1], 0x6420), __byte_perm(bl[2], bl[3], 0x6420));
  }
}

// Quads for padding PAD: quad (p, q) holds the inputs (2p+a+OFF, 2q+b+OFF), OFF = 1 - PAD, a, b in {0, 1}; it is
// covered by the windows (p..p+1, q..q+1) at row a+1-2*wa, column b+1-2*wb.  With PAD = 0 the quads start at -1
// (row / column 0 is the second element of quad -1).
__host__ __device__ inline int pool_quads(int extent, int pad) { return (extent - 1 - (1 - pad)) / 2 + (1 - pad) + 1; }

template <int PAD>
__global__ void __launch_bounds__(256)
maxpool_bwd_3x3s2_march_kernel(const bf16* __restrict__ dy, const unsigned char* __restrict__ widx,
                               bf16* __restrict__ dx, PoolGeom g, int cgs_shift,
                               const bf16* __restrict__ ypool = nullptr) {
  constexpr int OFF = 1 - PAD;
  const int cgs = g.C >> 3;
  const int Qx = pool_quads(g.W, PAD), Px = pool_quads(g.H, PAD);
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= Qx * cgs) return;
  const int qi = cgs_shift >= 0 ? (e >> cgs_shift) : (int)((unsigned)e / (unsigned)cgs);
  const int cg = e - qi * cgs;
  const int q = qi - OFF;
  const int n = blockIdx.z, p0 = blockIdx.y * MROWS - OFF;
  const int p1 = min(p0 + MROWS, Px - OFF);
  // window row r, columns q and q+1: output gradients (packed bf16) and argmax positions (8 bytes)
  uint4 d0[2], d1[2];
  uint2 l0[2], l1[2];
  bool ok0[2], ok1[2];
  auto load_row = [&](int r, uint4* d, uint2* l, bool* ok) {
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      ok[b] = r >= 0 && r < g.P && (q + b) >= 0 && (q + b) < g.Q;
      if (ok[b]) {
        const long long o = (((long long)n * g.P + r) * g.Q + (q + b)) * g.C + cg * 8;
        d[b] = *reinterpret_cast<const uint4*>(dy + o);
        l[b] = *reinterpret_cast<const uint2*>(widx + o);
        if (ypool) {   // ReLU backward of the pooled tensor's producer, packed: keep the halves whose pooled value > 0
          const uint4 yv = *reinterpret_cast<const uint4*>(ypool + o);
          const __nv_bfloat162 z = __floats2bfloat162_rn(0.f, 0.f);
          d[b].x &= __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&yv.x), z);
          d[b].y &= __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&yv.y), z);
          d[b].z &= __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&yv.z), z);
          d[b].w &= __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&yv.w), z);
        }
      }
    }
  };
  load_row(p0, d0, l0, ok0);
  for (int p = p0; p < p1; ++p) {
    load_row(p + 1, d1, l1, ok1);
    // sums in (wa, wb) order: ascending window index, as the generic kernel
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int h = 2 * p + a + OFF, w = 2 * q + b + OFF;
        if (h < 0 || w < 0 || h >= g.H || w >= g.W) continue;
        float acc[8];
#pragma unroll
        for (int e2 = 0; e2 < 8; ++e2) acc[e2] = 0.f;
#pragma unroll
        for (int wa = 0; wa <= a; ++wa)
#pragma unroll
          for (int wb = 0; wb <= b; ++wb) {
            const bool ok = wa ? ok1[wb] : ok0[wb];
            if (!ok) continue;
            const unsigned local = (unsigned)((a + 1 - 2 * wa) * 3 + (b + 1 - 2 * wb));
            const uint4 dr = wa ? d1[wb] : d0[wb];
            const uint2 lr = wa ? l1[wb] : l0[wb];
            float d[8];
            unpack8(dr, d);
#pragma unroll
            for (int e2 = 0; e2 < 8; ++e2) {
              const unsigned byte = ((e2 < 4 ? lr.x : lr.y) >> (8 * (e2 & 3))) & 0xffu;
              if (byte == local) acc[e2] += d[e2];
            }
          }
        *reinterpret_cast<uint4*>(dx + (((long long)n * g.H + h) * g.W + w) * g.C + cg * 8) = pack8(acc);
      }
#pragma unroll
    for (int b = 0; b < 2; ++b) { d0[b] = d1[b]; l0[b] = l1[b]; ok0[b] = ok1[b]; }
  }
}

// count_include_pad = True, ceil_mode = False (torch defaults): divisor is always R*S.
template <typename T, bool VEC>
__global__ void __launch_bounds__(256)
avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, PoolGeom g) {
  constexpr int VW = PV<T, VEC>::W;
  const int cgs = g.C / VW;
  const long long total = (long long)g.N * g.P * g.Q * cgs;
  const float inv = 1.f / (float)(g.R * g.S);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % cgs);
    long long t = i / cgs;
    const int q = (int)(t % g.Q); t /= g.Q;
    const int p = (int)(t % g.P);
    const int n = (int)(t / g.P);
    float acc[VW];
#pragma unroll
    for (int e = 0; e < VW; ++e) acc[e] = 0.f;
    const int h0 = p * g.sh - g.ph, w0 = q * g.sw - g.pw;
    for (int h = max(h0, 0); h < min(h0 + g.R, g.H); ++h)
      for (int w = max(w0, 0); w < min(w0 + g.S, g.W); ++w) {
        float v[VW];
        PV<T, VEC>::load(x + (((long long)n * g.H + h) * g.W + w) * g.C + cg * VW, v);
#pragma unroll
        for (int e = 0; e < VW; ++e) acc[e] += v[e];
      }
#pragma unroll
    for (int e = 0; e < VW; ++e) acc[e] *= inv;
    PV<T, VEC>::store(y + (((long long)n * g.P + p) * g.Q + q) * g.C + cg * VW, acc);
  }
}

template <typename T, bool VEC>
__global__ void __launch_bounds__(256)
avgpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, PoolGeom g) {
  constexpr int VW = PV<T, VEC>::W;
  const int cgs = g.C / VW;
  const long long total = (long long)g.N * g.H * g.W * cgs;
  const float inv = 1.f / (float)(g.R * g.S);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % cgs);
    long long t = i / cgs;
    const int w = (int)(t % g.W); t /= g.W;
    const int h = (int)(t % g.H);
    const int n = (int)(t / g.H);
    float acc[VW];
#pragma unroll
    for (int e = 0; e < VW; ++e) acc[e] = 0.f;
    int p_lo = (h + g.ph - g.R + g.sh) / g.sh; if (h + g.ph - g.R + 1 <= 0) p_lo = 0;
    int p_hi = (h + g.ph) / g.sh; if (p_hi > g.P - 1) p_hi = g.P - 1;
    int q_lo = (w + g.pw - g.S + g.sw) / g.sw; if (w + g.pw - g.S + 1 <= 0) q_lo = 0;
    int q_hi = (w + g.pw) / g.sw; if (q_hi > g.Q - 1) q_hi = g.Q - 1;
    for (int p = p_lo; p <= p_hi; ++p)
      for (int q = q_lo; q <= q_hi; ++q) {
        float d[VW];
        PV<T, VEC>::load(dy + (((long long)n * g.P + p) * g.Q + q) * g.C + cg * VW, d);
#pragma unroll
        for (int e = 0; e < VW; ++e) acc[e] += d[e];
      }
#pragma unroll
    for (int e = 0; e < VW; ++e) acc[e] *= inv;
    PV<T, VEC>::store(dx + (((long long)n * g.H + h) * g.W + w) * g.C + cg * VW, acc);
  }
}

// Global average pool: x[N, HW, C] -> y[N, C].  Block = (image, 256-thread slab of column groups x rows).
template <typename T, bool VEC>
__global__ void __launch_bounds__(256)
global_avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int HW, int C) {
  constexpr int VW = PV<T, VEC>::W;
  const int cgs = C / VW;
  const int tpr = cgs < 256 ? cgs : 256;
  const int rpp = 256 / tpr;
  const int cg = blockIdx.x * tpr + threadIdx.x % tpr;
  const int rl = threadIdx.x / tpr;
  const int n = blockIdx.y;
  float acc[VW];
#pragma unroll
  for (int e = 0; e < VW; ++e) acc[e] = 0.f;
  if (cg < cgs && rl < rpp)
    for (int r = rl; r < HW; r += rpp) {
      float v[VW];
      PV<T, VEC>::load(x + ((long long)n * HW + r) * C + cg * VW, v);
#pragma unroll
      for (int e = 0; e < VW; ++e) acc[e] += v[e];
    }
  __shared__ float sm[256 * 8];
#pragma unroll
  for (int e = 0; e < VW; ++e) sm[threadIdx.x * 8 + e] = acc[e];
  __syncthreads();
  if (rl == 0 && cg < cgs) {
    for (int rr = 1; rr < rpp; ++rr) {
      const int t = rr * tpr + threadIdx.x % tpr;
#pragma unroll
      for (int e = 0; e < VW; ++e) acc[e] += sm[t * 8 + e];
    }
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int e = 0; e < VW; ++e) acc[e] *= inv;
    PV<T, VEC>::store(y + (long long)n * C + cg * VW, acc);
  }
}

template <typename T, bool VEC>
__global__ void __launch_bounds__(256)
global_avgpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int HW, int C, long long total_vec) {
  constexpr int VW = PV<T, VEC>::W;
  const int cgs = C / VW;
  const float inv = 1.f / (float)HW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec; i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % cgs);
    const long long n = i / ((long long)cgs * HW);
    float d[VW];
    PV<T, VEC>::load(dy + n * C + cg * VW, d);
#pragma unroll
    for (int e = 0; e < VW; ++e) d[e] *= inv;
    PV<T, VEC>::store(dx + i * VW, d);
  }
}

// log2(v) when v is a power of two, else -1 (the kernels then divide)
int pow2_shift(int v) {
  if (v <= 0 || (v & (v - 1)) != 0) return -1;
  int s = 0;
  while ((1 << s) < v) ++s;
  return s;
}

// 1: 3x3 stride 2 pad 1, 2: 2x2 stride 2 pad 0, 3: 3x3 stride 2 pad 0, else 0 (see the FX template parameter)
int fixed_geometry(const PoolGeom& g) {
  if (g.sh != 2 || g.sw != 2 || g.R != g.S || g.ph != g.pw) return 0;
  if (g.R == 3 && g.ph == 1) return 1;
  if (g.R == 2 && g.ph == 0) return 2;
  if (g.R == 3 && g.ph == 0) return 3;
  return 0;
}

// RNN_POOL_MARCH=0 selects the one-row-per-CTA kernels for the 3x3 / stride 2 / pad 1 pool (tests compare the two)
bool pool_march_enabled() {
  const char* e = getenv("RNN_POOL_MARCH");
  return !(e != nullptr && e[0] == '0');
}

int pgrid(long long total) {
  long long b = (total + 255) / 256;
  const long long cap = (long long)rnn_num_sms() * 16;
  return (int)(b > cap ? cap : (b < 1 ? 1 : b));
}

int make_geom(PoolGeom* g, int N, int H, int W, int C, int R, int S, int sh, int sw, int ph, int pw, int vw) {
  (void)vw;
  RNN_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(R > 0 && S > 0 && sh > 0 && sw > 0 && ph >= 0 && pw >= 0 && ph * 2 <= R && pw * 2 <= S, RNN_ERR_SHAPE);
  g->N = N; g->H = H; g->W = W; g->C = C; g->R = R; g->S = S; g->sh = sh; g->sw = sw; g->ph = ph; g->pw = pw;
  g->P = (H + 2 * ph - R) / sh + 1;
  g->Q = (W + 2 * pw - S) / sw + 1;
  RNN_REQUIRE(g->P > 0 && g->Q > 0, RNN_ERR_SHAPE);
  return RNN_OK;
}

}  // namespace

extern "C" {

static int maxpool_fwd_impl(const void* x, void* y, int* indices, unsigned char* window_idx, int N, int H, int W, int C,
                            int R, int S, int sh, int sw, int ph, int pw, int is_f32, const float* scale,
                            const float* shift, cudaStream_t stream) {
  RNN_REQUIRE(R * S <= 255, RNN_ERR_SHAPE);
  PoolGeom g;
  const int vw = is_f32 ? 4 : 8;
  const bool vec = C % vw == 0;
  int rc = make_geom(&g, N, H, W, C, R, S, sh, sw, ph, pw, vw);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(g.P <= 65535 && N <= 65535, RNN_ERR_SHAPE);
  const int cgs = vec ? C / vw : C;
  const int sh_ = pow2_shift(cgs);
  const dim3 grid(ceil_div_i((long long)g.Q * cgs, 256), g.P, N);
  const int fx = fixed_geometry(g);
  if (is_f32) {
    if (vec) maxpool_fwd_kernel<float, true, 0><<<grid, 256, 0, stream>>>((const float*)x, (float*)y, indices, window_idx, g, sh_, scale, shift);
    else maxpool_fwd_kernel<float, false, 0><<<grid, 256, 0, stream>>>((const float*)x, (float*)y, indices, window_idx, g, sh_, scale, shift);
  } else if (!vec) {
    maxpool_fwd_kernel<bf16, false, 0><<<grid, 256, 0, stream>>>((const bf16*)x, (bf16*)y, indices, window_idx, g, sh_, scale, shift);
  } else {
    if ((fx == 1 || fx == 3) && pool_march_enabled()) {   // 3x3 stride 2, pad 1 (ResNet stem) or 0 (AlexNet)
      const dim3 mg(grid.x, ceil_div_i(g.P, MROWS), N);
      const bf16* xi = (const bf16*)x;
      bf16* yo = (bf16*)y;
      if (fx == 1 && scale != nullptr)
        maxpool_fwd_3x3s2_march_kernel<true, 1><<<mg, 256, 0, stream>>>(xi, yo, indices, window_idx, g, sh_, scale, shift);
      else if (fx == 1)
        maxpool_fwd_3x3s2_march_kernel<false, 1><<<mg, 256, 0, stream>>>(xi, yo, indices, window_idx, g, sh_, scale, shift);
      else if (scale != nullptr)
        maxpool_fwd_3x3s2_march_kernel<true, 0><<<mg, 256, 0, stream>>>(xi, yo, indices, window_idx, g, sh_, scale, shift);
      else
        maxpool_fwd_3x3s2_march_kernel<false, 0><<<mg, 256, 0, stream>>>(xi, yo, indices, window_idx, g, sh_, scale, shift);
      RNN_CHECK_LAUNCH();
      return RNN_OK;
    }
    switch (fx) {
      case 1: maxpool_fwd_kernel<bf16, true, 1><<<grid, 256, 0, stream>>>((const bf16*)x, (bf16*)y, indices, window_idx, g, sh_, scale, shift); break;
      case 2: maxpool_fwd_kernel<bf16, true, 2><<<grid, 256, 0, stream>>>((const bf16*)x, (bf16*)y, indices, window_idx, g, sh_, scale, shift); break;
      case 3: maxpool_fwd_kernel<bf16, true, 3><<<grid, 256, 0, stream>>>((const bf16*)x, (bf16*)y, indices, window_idx, g, sh_, scale, shift); break;
      default: maxpool_fwd_kernel<bf16, true, 0><<<grid, 256, 0, stream>>>((const bf16*)x, (bf16*)y, indices, window_idx, g, sh_, scale, shift);
    }
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_maxpool2d_fwd(const void* x, void* y, int* indices, unsigned char* window_idx, int N, int H, int W, int C,
                      int R, int S, int sh, int sw, int ph, int pw, int is_f32, cudaStream_t stream) {
  return maxpool_fwd_impl(x, y, indices, window_idx, N, H, W, C, R, S, sh, sw, ph, pw, is_f32, nullptr, nullptr, stream);
}
int rnn_bn_relu_maxpool2d_fwd(const void* x, const float* scale, const float* shift, void* y, int* indices,
                              unsigned char* window_idx, int N, int H, int W, int C, int R, int S, int sh, int sw, int ph,
                              int pw, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(scale != nullptr && shift != nullptr, RNN_ERR_SHAPE);
  return maxpool_fwd_impl(x, y, indices, window_idx, N, H, W, C, R, S, sh, sw, ph, pw, is_f32, scale, shift, stream);
}
static int maxpool_bwd_impl(const void* dy, const void* ypool, const int* indices, const unsigned char* window_idx,
                            void* dx, int N, int H, int W, int C, int R, int S, int sh, int sw, int ph, int pw,
                            int is_f32, cudaStream_t stream) {
  RNN_REQUIRE((indices != nullptr) != (window_idx != nullptr), RNN_ERR_SHAPE);  // exactly one index source
  PoolGeom g;
  const int vw = is_f32 ? 4 : 8;
  const bool vec = C % vw == 0;
  int rc = make_geom(&g, N, H, W, C, R, S, sh, sw, ph, pw, vw);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(H <= 65535 && N <= 65535, RNN_ERR_SHAPE);
  const int cgs = vec ? C / vw : C;
  const int sh_ = pow2_shift(cgs);
  const dim3 grid(ceil_div_i((long long)W * cgs, 256), H, N);
  const int fx = fixed_geometry(g);
  if (!is_f32 && vec && window_idx != nullptr && fx == 2) {   // 2x2 stride 2, no padding: non-overlapping windows
    const dim3 tg(ceil_div_i((long long)((W + 1) / 2) * cgs, 256), (H + 1) / 2, N);
    maxpool_bwd_tiled_kernel<bf16, 2><<<tg, 256, 0, stream>>>((const bf16*)dy, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool);
    RNN_CHECK_LAUNCH();
    return RNN_OK;
  }
  if (!is_f32 && vec && window_idx != nullptr && (fx == 1 || fx == 3) && pool_march_enabled()) {
    // 3x3 stride 2, pad 1 (ResNet stem) or pad 0 (AlexNet): marching 2x2 input quads
    const int pad = fx == 1 ? 1 : 0;
    const dim3 mg(ceil_div_i((long long)pool_quads(W, pad) * cgs, 256), ceil_div_i(pool_quads(H, pad), MROWS), N);
    if (pad) maxpool_bwd_3x3s2_march_kernel<1><<<mg, 256, 0, stream>>>((const bf16*)dy, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool);
    else maxpool_bwd_3x3s2_march_kernel<0><<<mg, 256, 0, stream>>>((const bf16*)dy, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool);
    RNN_CHECK_LAUNCH();
    return RNN_OK;
  }
  if (!is_f32 && vec && window_idx != nullptr && fx == 1) {   // 3x3 stride 2 pad 1: one thread per 2x2 input quad
    const dim3 tg(ceil_div_i((long long)((W + 1) / 2) * cgs, 256), (H + 1) / 2, N);
    maxpool_bwd_3x3s2_kernel<bf16><<<tg, 256, 0, stream>>>((const bf16*)dy, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool);
    RNN_CHECK_LAUNCH();
    return RNN_OK;
  }
  if (is_f32) {
    if (vec) maxpool_bwd_kernel<float, true, 0><<<grid, 256, 0, stream>>>((const float*)dy, indices, window_idx, (float*)dx, g, sh_, (const float*)ypool);
    else maxpool_bwd_kernel<float, false, 0><<<grid, 256, 0, stream>>>((const float*)dy, indices, window_idx, (float*)dx, g, sh_, (const float*)ypool);
  } else if (!vec) {
    maxpool_bwd_kernel<bf16, false, 0><<<grid, 256, 0, stream>>>((const bf16*)dy, indices, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool);
  } else {
    switch (fx) {
      case 1: maxpool_bwd_kernel<bf16, true, 1><<<grid, 256, 0, stream>>>((const bf16*)dy, indices, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool); break;
      case 2: maxpool_bwd_kernel<bf16, true, 2><<<grid, 256, 0, stream>>>((const bf16*)dy, indices, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool); break;
      case 3: maxpool_bwd_kernel<bf16, true, 3><<<grid, 256, 0, stream>>>((const bf16*)dy, indices, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool); break;
      default: maxpool_bwd_kernel<bf16, true, 0><<<grid, 256, 0, stream>>>((const bf16*)dy, indices, window_idx, (bf16*)dx, g, sh_, (const bf16*)ypool);
    }
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_maxpool2d_bwd(const void* dy, const int* indices, const unsigned char* window_idx, void* dx, int N, int H,
                      int W, int C, int R, int S, int sh, int sw, int ph, int pw, int is_f32, cudaStream_t stream) {
  return maxpool_bwd_impl(dy, nullptr, indices, window_idx, dx, N, H, W, C, R, S, sh, sw, ph, pw, is_f32, stream);
}
int rnn_maxpool2d_relu_bwd(const void* dy, const void* y, const int* indices, const unsigned char* window_idx, void* dx,
                           int N, int H, int W, int C, int R, int S, int sh, int sw, int ph, int pw, int is_f32,
                           cudaStream_t stream) {
  RNN_REQUIRE(y != nullptr, RNN_ERR_SHAPE);
  return maxpool_bwd_impl(dy, y, indices, window_idx, dx, N, H, W, C, R, S, sh, sw, ph, pw, is_f32, stream);
}
int rnn_avgpool2d_fwd(const void* x, void* y, int N, int H, int W, int C, int R, int S, int sh, int sw, int ph, int pw,
                      int is_f32, cudaStream_t stream) {
  PoolGeom g;
  const int vw = is_f32 ? 4 : 8;
  const bool vec = C % vw == 0;
  int rc = make_geom(&g, N, H, W, C, R, S, sh, sw, ph, pw, vw);
  if (rc != RNN_OK) return rc;
  const long long total = (long long)N * g.P * g.Q * (vec ? C / vw : C);
  if (is_f32) {
    if (vec) avgpool_fwd_kernel<float, true><<<pgrid(total), 256, 0, stream>>>((const float*)x, (float*)y, g);
    else avgpool_fwd_kernel<float, false><<<pgrid(total), 256, 0, stream>>>((const float*)x, (float*)y, g);
  } else {
    if (vec) avgpool_fwd_kernel<bf16, true><<<pgrid(total), 256, 0, stream>>>((const bf16*)x, (bf16*)y, g);
    else avgpool_fwd_kernel<bf16, false><<<pgrid(total), 256, 0, stream>>>((const bf16*)x, (bf16*)y, g);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_avgpool2d_bwd(const void* dy, void* dx, int N, int H, int W, int C, int R, int S, int sh, int sw, int ph, int pw,
                      int is_f32, cudaStream_t stream) {
  PoolGeom g;
  const int vw = is_f32 ? 4 : 8;
  const bool vec = C % vw == 0;
  int rc = make_geom(&g, N, H, W, C, R, S, sh, sw, ph, pw, vw);
  if (rc != RNN_OK) return rc;
  const long long total = (long long)N * H * W * (vec ? C / vw : C);
  if (is_f32) {
    if (vec) avgpool_bwd_kernel<float, true><<<pgrid(total), 256, 0, stream>>>((const float*)dy, (float*)dx, g);
    else avgpool_bwd_kernel<float, false><<<pgrid(total), 256, 0, stream>>>((const float*)dy, (float*)dx, g);
  } else {
    if (vec) avgpool_bwd_kernel<bf16, true><<<pgrid(total), 256, 0, stream>>>((const bf16*)dy, (bf16*)dx, g);
    else avgpool_bwd_kernel<bf16, false><<<pgrid(total), 256, 0, stream>>>((const bf16*)dy, (bf16*)dx, g);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_global_avgpool_fwd(const void* x, void* y, int N, int HW, int C, int is_f32, cudaStream_t stream) {
  const int vw = is_f32 ? 4 : 8;
  RNN_REQUIRE(N > 0 && HW > 0 && C > 0, RNN_ERR_SHAPE);
  const bool vec = C % vw == 0;
  const int cgs = vec ? C / vw : C, tpr = cgs < 256 ? cgs : 256, gx = (cgs + tpr - 1) / tpr;
  const dim3 grid(gx, N);
  if (is_f32) {
    if (vec) global_avgpool_fwd_kernel<float, true><<<grid, 256, 0, stream>>>((const float*)x, (float*)y, HW, C);
    else global_avgpool_fwd_kernel<float, false><<<grid, 256, 0, stream>>>((const float*)x, (float*)y, HW, C);
  } else {
    if (vec) global_avgpool_fwd_kernel<bf16, true><<<grid, 256, 0, stream>>>((const bf16*)x, (bf16*)y, HW, C);
    else global_avgpool_fwd_kernel<bf16, false><<<grid, 256, 0, stream>>>((const bf16*)x, (bf16*)y, HW, C);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}
int rnn_global_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, int is_f32, cudaStream_t stream) {
  const int vw = is_f32 ? 4 : 8;
  RNN_REQUIRE(N > 0 && HW > 0 && C > 0, RNN_ERR_SHAPE);
  const bool vec = C % vw == 0;
  const long long total = (long long)N * HW * (vec ? C / vw : C);
  if (is_f32) {
    if (vec) global_avgpool_bwd_kernel<float, true><<<pgrid(total), 256, 0, stream>>>((const float*)dy, (float*)dx, HW, C, total);
    else global_avgpool_bwd_kernel<float, false><<<pgrid(total), 256, 0, stream>>>((const float*)dy, (float*)dx, HW, C, total);
  } else {
    if (vec) global_avgpool_bwd_kernel<bf16, true><<<pgrid(total), 256, 0, stream>>>((const bf16*)dy, (bf16*)dx, HW, C, total);
    else global_avgpool_bwd_kernel<bf16, false><<<pgrid(total), 256, 0, stream>>>((const bf16*)dy, (bf16*)dx, HW, C, total);
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

}  // extern "C"
