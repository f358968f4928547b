// Fused tcgen05 attention (bf16, head size 64 or 128): the scores never leave the SM.
//
//     w = q k^T * scale ; mask ; softmax ; dropout ; z = w v          (reference nn/attention.py:37-47)
//
// Forward, one CTA per (batch, head, 128-query tile), two sweeps over the keys in 128-key tiles:
//   sweep 1:  S_j = Q K_j^T into TMEM (tcgen05.mma), the softmax warps read it back (tcgen05.ld) and keep the running
//             row maximum of s*scale + mask -- no exponentials, nothing stored;
//   sweep 2:  S_j again (the QK^T product is ~a third of the kernel's tensor work; repeating it is cheaper than
//             rescaling an accumulator that lives in TMEM), P_j = exp(s - max) [* dropout keep], written as bf16
//             into shared memory in the K-major SWIZZLE_128B layout of an MMA A operand, then O += P_j V_j
//             accumulates in a second TMEM region; the row sums are carried in registers.
//   epilogue: O / rowsum -> bf16 -> TMA store into the concatenated-heads output; log-sum-exp saved (fp32 [B,H,T]).
// The [B,H,T,Tk] score / probability tensors of the reference (:37,45) and of the unfused tensor-core pipeline
// (rnn_attention_tc_*) never reach HBM; the backward kernels below recompute the probabilities from the saved
// log-sum-exp and regenerate the dropout mask from the same Philox counters.
//
// Roles (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected thread), warps 2..9 = softmax /
// epilogue: a query row (= TMEM lane) is shared by two threads, one per 64-key half of every tile (the softmax,
// mask and Philox arithmetic of these warps, not the tensor pipe, is what bounds the kernel; partial row maxima and
// sums are combined through shared memory once per sweep).  Two S buffers in TMEM, two P buffers in shared memory:
// the S product of tile j+1 overlaps the softmax of tile j, which overlaps the P V product of tile j-1.
//
// Included by attention.cu inside its anonymous namespace (shares philox4 / the dropout epoch with the other
// attention kernels, so all paths draw the same keep mask for the same seed / offset).
#pragma once

namespace fa {

constexpr int BQ = 128;    // query rows per CTA = TMEM lanes
constexpr int BKV = 128;   // keys per tile
constexpr int TILE64 = 16384;   // a [128 rows x 64 bf16] SWIZZLE_128B block
constexpr int NTHREADS = 320;   // producer warp, MMA warp, 8 softmax / epilogue warps (two threads per query row)
constexpr float kLog2e = 1.4426950408889634f;
constexpr int kMaskScratch = 8 * 4096;   // 4 KB per softmax warp (see scores32)
constexpr int kMaxSmem = 232448;         // 227 KB of dynamic shared memory per CTA

struct Params {
  int B, H, T, Tk;
  float scale;
  const void* mask;
  int mask_kind;            // 0 none, 1 bool (nonzero = masked out), 2 float32 additive
  long long mask_bs, mask_rs;
  int mask_vec;             // rows are 16-byte aligned: vector loads
  float p_drop;
  unsigned long long seed, offset;
  float* lse;               // [B,H,T]
  float* delta;             // [B,H,T] (backward)
  const bf16 *z, *dz;       // backward: forward output and its gradient, for delta_i = dz_i . z_i
  long long z_bs, z_rs, dz_bs, dz_rs;
};

template <int HS>
struct Cfg {
  static constexpr int NKB = HS / 64;              // 64-wide blocks of the head dimension
  static constexpr int TILE = NKB * TILE64;        // a [128 x HS] operand tile (Q, K, V, dO)
  static constexpr int PBYTES = 2 * TILE64;        // a [128 x 128] probability / dS tile
};

// ---- score of one element: s * scale + mask, in the log2 domain ---------------------------------------------------
// 32 consecutive keys of one query row.  `ok` = the row exists (i < T).  Keys at or beyond Tk get -inf.
// `mscr`: 4 KB of shared memory owned by the calling warp, or nullptr.  With it, an additive float mask is fetched by
// the warp as a whole -- four fully used 128-byte lines per load instruction -- and transposed through the scratch
// (XOR-swizzled, conflict-free both ways); without it every thread walks its own mask row and each load instruction
// touches 32 different lines, which made the masked kernels L1-bound (the reference's decoder blocks always pass
// such a mask, arch/transformer.py:148).  Must be called by all 32 lanes of the warp (rows i - lane .. i - lane + 31).
__device__ __forceinline__ void scores32(const Params& p, const uint32_t* raw, float* s, int b, int i, int j0, bool ok,
                                         float* mscr = nullptr, int lane = 0) {
  const float c = p.scale * kLog2e;
#pragma unroll
  for (int e = 0; e < 32; ++e) s[e] = __uint_as_float(raw[e]) * c;
  if (p.mask_kind == 2 && mscr != nullptr && p.mask_vec && j0 + 32 <= p.Tk) {
    const float* mb = reinterpret_cast<const float*>(p.mask) + b * p.mask_bs + j0;
    const int i_base = i - lane;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int row_l = it * 4 + (lane >> 3), piece = lane & 7;
      const int row = i_base + row_l;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row < p.T) v = __ldg(reinterpret_cast<const float4*>(mb + (long long)row * p.mask_rs) + piece);
      *reinterpret_cast<float4*>(mscr + row_l * 32 + ((piece ^ (row_l & 7)) << 2)) = v;
    }
    __syncwarp();
#pragma unroll
    for (int e4 = 0; e4 < 8; ++e4) {
      const float4 m = *reinterpret_cast<const float4*>(mscr + lane * 32 + ((e4 ^ (lane & 7)) << 2));
      s[4 * e4 + 0] = fmaf(m.x, kLog2e, s[4 * e4 + 0]); s[4 * e4 + 1] = fmaf(m.y, kLog2e, s[4 * e4 + 1]);
      s[4 * e4 + 2] = fmaf(m.z, kLog2e, s[4 * e4 + 2]); s[4 * e4 + 3] = fmaf(m.w, kLog2e, s[4 * e4 + 3]);
    }
    __syncwarp();
    return;   // j0 + 32 <= Tk: no key of this chunk is out of range
  }
  if (p.mask_kind == 1 && ok) {
    const uint8_t* mrow = reinterpret_cast<const uint8_t*>(p.mask) + b * p.mask_bs + i * p.mask_rs + j0;
    if (p.mask_vec && j0 + 32 <= p.Tk) {
      uint32_t w[8];
      *reinterpret_cast<uint4*>(w) = __ldg(reinterpret_cast<const uint4*>(mrow));
      *reinterpret_cast<uint4*>(w + 4) = __ldg(reinterpret_cast<const uint4*>(mrow + 16));
#pragma unroll
      for (int e = 0; e < 32; ++e)
        if ((w[e >> 2] >> ((e & 3) * 8)) & 0xFFu) s[e] = -INFINITY;
    } else {
#pragma unroll
      for (int e = 0; e < 32; ++e)
        if (j0 + e < p.Tk && __ldg(mrow + e)) s[e] = -INFINITY;
    }
  } else if (p.mask_kind == 2 && ok) {
    const float* mrow = reinterpret_cast<const float*>(p.mask) + b * p.mask_bs + i * p.mask_rs + j0;
    if (p.mask_vec && j0 + 32 <= p.Tk) {
#pragma unroll
      for (int e4 = 0; e4 < 8; ++e4) {
        const float4 m = __ldg(reinterpret_cast<const float4*>(mrow) + e4);
        s[4 * e4 + 0] = fmaf(m.x, kLog2e, s[4 * e4 + 0]); s[4 * e4 + 1] = fmaf(m.y, kLog2e, s[4 * e4 + 1]);
        s[4 * e4 + 2] = fmaf(m.z, kLog2e, s[4 * e4 + 2]); s[4 * e4 + 3] = fmaf(m.w, kLog2e, s[4 * e4 + 3]);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 32; ++e)
        if (j0 + e < p.Tk) s[e] = fmaf(__ldg(mrow + e), kLog2e, s[e]);
    }
  }
  if (j0 + 32 > p.Tk) {
#pragma unroll
    for (int e = 0; e < 32; ++e)
      if (j0 + e >= p.Tk) s[e] = -INFINITY;
  }
}

// dropout keep factors (0 or 1/(1-p)) of 32 consecutive keys of row (bh, i): the Philox words of keep_scale()
__device__ __forceinline__ void keep32(const Params& p, float* kf, int bh, int i, int j0) {
  const float keep = 1.f / (1.f - p.p_drop);
  const uint32_t thresh = keep_thresh16(p.p_drop);
  const unsigned long long idx0 = ((unsigned long long)bh * p.T + i) * p.Tk + j0;   // multiple of 8 (Tk % 8 == 0)
#pragma unroll
  for (int e8 = 0; e8 < 4; ++e8) {   // one Philox call = eight 16-bit decisions (see philox_half)
    const uint4 r = philox4(p.seed, p.offset, (idx0 >> 3) + e8);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      kf[8 * e8 + 2 * k] = (w[k] & 0xFFFFu) >= thresh ? keep : 0.f;
      kf[8 * e8 + 2 * k + 1] = (w[k] >> 16) >= thresh ? keep : 0.f;
    }
  }
}

// Write 32 bf16 values of row `r` (columns c0 .. c0+31 of a [128 x 128] tile) into the K-major SWIZZLE_128B layout:
// 64-column blocks of TILE64 bytes, 128-byte rows, 16-byte pieces XOR-swizzled with the row index.
__device__ __forceinline__ void store_tile_row32(uint8_t* tile, int r, int c0, const float* v) {
  uint8_t* rowp = tile + (c0 >> 6) * TILE64 + r * 128;
  const int piece0 = (c0 & 63) >> 3;
#pragma unroll
  for (int g = 0; g < 4; ++g)
    *reinterpret_cast<uint4*>(rowp + (((piece0 + g) ^ (r & 7)) << 4)) = pack8(v + 8 * g);
}

// TMEM accumulator rows -> bf16 -> per-warp swizzled staging -> TMA store (clips rows / columns outside the tensor).
// `stage_w`: 4 KB of 1024-byte-aligned shared memory owned by this warp.  NCH 64-column chunks.
template <int NCH>
__device__ __forceinline__ void store_acc_rows(uint32_t t_acc, float row_scale, uint8_t* stage_w, const CUtensorMap* tm,
                                               int col0, int row0, int slab, int lane) {
#pragma unroll 1
  for (int chunk = 0; chunk < NCH; ++chunk) {
    uint32_t r[64];
    tc::tmem_ld32(t_acc + (uint32_t)(chunk * 64), r);
    tc::tmem_ld32(t_acc + (uint32_t)(chunk * 64 + 32), r + 32);
    tc::tmem_ld_wait();
    uint4 out[8];
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[8 * g + e]) * row_scale;
      out[g] = pack8(v);
    }
    if (lane == 0) tc::tma_store_wait_read<0>();
    __syncwarp();
    uint4* srow = reinterpret_cast<uint4*>(stage_w + lane * 128);
#pragma unroll
    for (int j = 0; j < 8; ++j) srow[j ^ (lane & 7)] = out[j];
    tc::fence_proxy_async();
    __syncwarp();
    if (lane == 0) {
      tc::tma_store_3d(tm, stage_w, col0 + chunk * 64, row0, slab);
      tc::tma_store_commit();
    }
  }
  if (lane == 0) tc::tma_store_wait<0>();
  __syncwarp();
}

// one [128 x HS] operand tile = HS/64 boxes of {64 columns, 128 rows}
template <int HS>
__device__ __forceinline__ void load_tile(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int col0, int row0, int slab) {
#pragma unroll
  for (int kb = 0; kb < HS / 64; ++kb) tc::tma_ld3(dst + kb * TILE64, tm, bar, col0 + kb * 64, row0, slab);
}

// D[128 x 128] = A[128 x HS] . B[128 x HS]^T, both K-major tiles (Q K^T, dO V^T)
template <int HS>
__device__ __forceinline__ void mma_nt(uint32_t d_tmem, uint32_t sA, uint32_t sB) {
  constexpr uint32_t idesc = tc::make_idesc_bf16(128, 128, false, false);
  uint32_t acc = 0;
#pragma unroll
  for (int kb = 0; kb < HS / 64; ++kb) {
    const uint64_t ad = tc::make_smem_desc(sA + kb * TILE64, 16, 1024);
    const uint64_t bd = tc::make_smem_desc(sB + kb * TILE64, 16, 1024);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tc::umma_bf16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, acc);
      acc = 1;
    }
  }
}
// D[128 x HS] (+)= A[128 x 128] . B[128 x HS]: A = a K-major probability-like tile, B = an operand tile read MN-major
// (rows = the reduction index).  (P V, dS K)
template <int HS>
__device__ __forceinline__ void mma_nn(uint32_t d_tmem, uint32_t sA, uint32_t sB, uint32_t accumulate) {
  constexpr uint32_t idesc = tc::make_idesc_bf16(128, HS, false, true);
  const uint64_t bd = tc::make_smem_desc(sB, TILE64, 1024);
  uint32_t acc = accumulate;
#pragma unroll
  for (int kb = 0; kb < 2; ++kb) {
    const uint64_t ad = tc::make_smem_desc(sA + kb * TILE64, 16, 1024);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tc::umma_bf16(d_tmem, ad + 2 * k, bd + (uint64_t)((kb * 4 + k) * (2048 >> 4)), idesc, acc);
      acc = 1;
    }
  }
}
// D[128 x HS] (+)= A^T . B with A = a [128 x 128] tile stored K-major for its untransposed use (so M-major here) and
// B = a [128 x HS] operand tile read MN-major; the reduction runs over the 128 rows of both.  (P^T dO, dS^T Q)
template <int HS>
__device__ __forceinline__ void mma_tn(uint32_t d_tmem, uint32_t sA, uint32_t sB, uint32_t accumulate) {
  constexpr uint32_t idesc = tc::make_idesc_bf16(128, HS, true, true);
  const uint64_t ad = tc::make_smem_desc(sA, TILE64, 1024);
  const uint64_t bd = tc::make_smem_desc(sB, TILE64, 1024);
  uint32_t acc = accumulate;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    tc::umma_bf16(d_tmem, ad + (uint64_t)(k * (2048 >> 4)), bd + (uint64_t)(k * (2048 >> 4)), idesc, acc);
    acc = 1;
  }
}


// =================================================================================================== forward
template <int HS>
struct FwdCfg : Cfg<HS> {
  static constexpr int STAGES = HS == 64 ? 3 : 2;                      // (K_j, V_j) ring
  static constexpr int BASE = Cfg<HS>::TILE * (1 + 2 * STAGES) + 2 * Cfg<HS>::PBYTES + 1024 + 256 + 1024;
  static constexpr bool MSCR = BASE + kMaskScratch <= kMaxSmem;    // room for the per-warp mask scratch?
  static constexpr int SMEM = BASE + (MSCR ? kMaskScratch : 0);
};

template <int HS>
__global__ void __launch_bounds__(NTHREADS, 1)
fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
           const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmZ,
           const __grid_constant__ Params p) {
  using C = FwdCfg<HS>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = tc::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + C::TILE;                          // stage s: K at s * 2 * TILE, V after it
  uint8_t* sP = sKV + C::STAGES * 2 * C::TILE;          // two [128 x 128] bf16 tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * C::PBYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;
  uint64_t* kv_empty = kv_full + C::STAGES;
  uint64_t* s_full = kv_empty + C::STAGES;              // [2]
  uint64_t* s_empty = s_full + 2;                       // [2]
  uint64_t* p_full = s_empty + 2;                       // [2]
  uint64_t* p_empty = p_full + 2;                       // [2]
  uint64_t* o_full = p_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);
  float* xchg = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [2 halves][128 rows]
  float* mscr_base = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256 + 1024);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, h = blockIdx.y, b = blockIdx.z;
  const int nkv = (p.Tk + BKV - 1) / BKV;

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmQ); tc::prefetch_tmap(&tmK); tc::prefetch_tmap(&tmV); tc::prefetch_tmap(&tmZ);
    tc::mbar_init(q_full, 1);
    for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&kv_full[s], 1); tc::mbar_init(&kv_empty[s], 1); }
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&s_full[i], 1); tc::mbar_init(&s_empty[i], 8);
      tc::mbar_init(&p_full[i], 8); tc::mbar_init(&p_empty[i], 1);
    }
    tc::mbar_init(o_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) { tc::tmem_alloc(tmem_slot, 512); tc::tmem_relinquish(); }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + 256;   // S buffers at columns 0 / 128, O at 256

  if (warp == 0) {
    // ------------------------------------------------------------------------------------------ TMA producer
    if (tc::elect_one()) {
      tc::mbar_expect_tx(tc::smem_u32(q_full), C::TILE);
      load_tile<HS>(tc::smem_u32(sQ), &tmQ, tc::smem_u32(q_full), h * HS, q0, b);
      int stage = 0;
      uint32_t phase = 0;
      for (int sweep = 0; sweep < 2; ++sweep)
        for (int j = 0; j < nkv; ++j) {
          tc::mbar_wait(tc::smem_u32(&kv_empty[stage]), phase ^ 1u);
          const uint32_t fb = tc::smem_u32(&kv_full[stage]);
          const uint32_t dK = tc::smem_u32(sKV + stage * 2 * C::TILE);
          tc::mbar_expect_tx(fb, sweep == 0 ? C::TILE : 2 * C::TILE);
          load_tile<HS>(dK, &tmK, fb, h * HS, j * BKV, b);
          if (sweep == 1) load_tile<HS>(dK + C::TILE, &tmV, fb, h * HS, j * BKV, b);
          if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
        }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------------------------------ MMA issuer
    if (tc::elect_one()) {
      const uint32_t aQ = tc::smem_u32(sQ);
      tc::mbar_wait(tc::smem_u32(q_full), 0);
      tc::tc_fence_after();
      int stage = 0;
      uint32_t phase = 0;
      int t = 0;    // S tiles issued so far (both sweeps): buffer t & 1, barrier phase (t >> 1) & 1
      // sweep 1: S tiles only; the K stage is free as soon as its product has retired
      for (int j = 0; j < nkv; ++j, ++t) {
        tc::mbar_wait(tc::smem_u32(&kv_full[stage]), phase);
        tc::mbar_wait(tc::smem_u32(&s_empty[t & 1]), (((uint32_t)t >> 1) & 1u) ^ 1u);
        tc::tc_fence_after();
        mma_nt<HS>(tS + (uint32_t)((t & 1) * 128), aQ, tc::smem_u32(sKV + stage * 2 * C::TILE));
        tc::umma_commit(&kv_empty[stage]);
        tc::umma_commit(&s_full[t & 1]);
        if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
      }
      // sweep 2: S_j, then O += P_{j-1} V_{j-1} one tile behind, so that the softmax of tile j-1 overlaps S_j
      int pstage = stage;           // stage of tile j-1
      for (int j = 0; j <= nkv; ++j) {
        if (j < nkv) {
          tc::mbar_wait(tc::smem_u32(&kv_full[stage]), phase);
          tc::mbar_wait(tc::smem_u32(&s_empty[t & 1]), (((uint32_t)t >> 1) & 1u) ^ 1u);
          tc::tc_fence_after();
          mma_nt<HS>(tS + (uint32_t)((t & 1) * 128), aQ, tc::smem_u32(sKV + stage * 2 * C::TILE));
          tc::umma_commit(&s_full[t & 1]);
          ++t;
        }
        if (j >= 1) {
          const int u = j - 1;      // P tile index: buffer u & 1, phase (u >> 1) & 1
          tc::mbar_wait(tc::smem_u32(&p_full[u & 1]), ((uint32_t)u >> 1) & 1u);
          tc::tc_fence_after();
          mma_nn<HS>(tO, tc::smem_u32(sP + (u & 1) * C::PBYTES), tc::smem_u32(sKV + pstage * 2 * C::TILE + C::TILE),
                     u > 0 ? 1u : 0u);
          tc::umma_commit(&kv_empty[pstage]);
          tc::umma_commit(&p_empty[u & 1]);
        }
        pstage = stage;
        if (j < nkv) { if (++stage == C::STAGES) { stage = 0; phase ^= 1u; } }
      }
      tc::umma_commit(o_full);
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------------------------------ softmax / epilogue
    const int qd = warp & 3;                     // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;            // which 64 keys of every 128-key tile this thread owns
    const int r = qd * 32 + lane;                // row inside the tile
    const int i = q0 + r;
    const bool row_ok = i < p.T;
    const int bh = b * p.H + h;
    const uint32_t lane_off = (uint32_t)(qd * 32) << 16;
    float* mscr = C::MSCR ? mscr_base + (warp - 2) * 1024 : nullptr;
    float m = -INFINITY;
    int t = 0;
    for (int j = 0; j < nkv; ++j, ++t) {
      tc::mbar_wait(&s_full[t & 1], ((uint32_t)t >> 1) & 1u);
      tc::tc_fence_after();
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const int ch = half * 2 + cc;
        uint32_t raw[32];
        float s[32];
        tc::tmem_ld32(tS + lane_off + (uint32_t)((t & 1) * 128 + ch * 32), raw);
        tc::tmem_ld_wait();
        scores32(p, raw, s, b, i, j * BKV + ch * 32, row_ok, mscr, lane);
#pragma unroll
        for (int e = 0; e < 32; ++e) m = fmaxf(m, s[e]);
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&s_empty[t & 1]);
    }
    // row maximum over both halves
    xchg[half * 128 + r] = m;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    m = fmaxf(m, xchg[(half ^ 1) * 128 + r]);
    asm volatile("bar.sync 1, 256;" ::: "memory");   // xchg is reused for the row sums below
    float l = 0.f;
    for (int j = 0; j < nkv; ++j, ++t) {
      tc::mbar_wait(&s_full[t & 1], ((uint32_t)t >> 1) & 1u);
      tc::mbar_wait(&p_empty[j & 1], (((uint32_t)j >> 1) & 1u) ^ 1u);
      tc::tc_fence_after();
      uint8_t* pt = sP + (j & 1) * C::PBYTES;
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const int ch = half * 2 + cc;
        uint32_t raw[32];
        float s[32];
        tc::tmem_ld32(tS + lane_off + (uint32_t)((t & 1) * 128 + ch * 32), raw);
        tc::tmem_ld_wait();
        scores32(p, raw, s, b, i, j * BKV + ch * 32, row_ok, mscr, lane);
#pragma unroll
        for (int e = 0; e < 32; ++e) { s[e] = exp2f(s[e] - m); l += s[e]; }
        if (p.p_drop > 0.f) {
          float kf[32];
          keep32(p, kf, bh, i, j * BKV + ch * 32);
#pragma unroll
          for (int e = 0; e < 32; ++e) s[e] *= kf[e];
        }
        store_tile_row32(pt, r, ch * 32, s);
      }
      tc::tc_fence_before();
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) { tc::mbar_arrive(&s_empty[t & 1]); tc::mbar_arrive(&p_full[j & 1]); }
    }
    // row sum over both halves, in a fixed order (half 0 + half 1) so that both threads hold the same value
    xchg[half * 128 + r] = l;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    l = xchg[r] + xchg[128 + r];
    // epilogue: O / l -> z, log-sum-exp (natural log) for the backward pass
    tc::mbar_wait(o_full, 0);
    tc::tc_fence_after();
    if (half == 0 && row_ok && p.lse != nullptr) p.lse[(long long)bh * p.T + i] = (m + log2f(l)) * 0.6931471805599453f;
    // every P V product has retired: the first P buffer is free and serves as the store staging area.  The 64-column
    // chunks of O are dealt out to the two halves (head size 64: one chunk, half 0 stores it).
    if (half < HS / 64)
      store_acc_rows<1>(tO + lane_off + (uint32_t)(half * 64), 1.f / l, sP + (warp - 2) * 4096, &tmZ, h * HS + half * 64,
                        q0 + qd * 32, b, lane);
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, 512);
  }
}

// =================================================================================================== backward
// Probabilities are recomputed from the saved log-sum-exp:  P = exp(s*scale + mask - lse),  Pd = P * keep,
//   dPd = dO V^T,  delta_i = dO_i . z_i (= sum_j Pd_ij dPd_ij),  dS = scale * P * (dPd * keep - delta)
//   dQ = dS K,  dK = dS^T Q,  dV = Pd^T dO.
// Two kernels, no atomics, fixed summation order (bitwise reproducible):
//   dq_kernel   : CTA per (b, h, query tile), loops over key tiles; also writes delta for the second kernel
//   dkdv_kernel : CTA per (b, h, key tile), loops over query tiles
// Each recomputes S and dPd for its (query tile, key tile) pairs -- 7 tile products in total instead of 5; what
// they save is every [B,H,T,Tk] tensor of the unfused pipeline (P, Pd, dPd, dS: 4 writes + 5 reads of T*Tk each).

// p (normalised probability) and dS of 32 keys of one row; `raw_s`, `raw_dp` = S and dPd accumulator words
__device__ __forceinline__ void bwd_row32(const Params& p, const uint32_t* raw_s, const uint32_t* raw_dp, float* pd, float* ds,
                                          int b, int bh, int i, int j0, bool ok, float lse2, float delta,
                                          float* mscr = nullptr, int lane = 0) {
  float s[32];
  scores32(p, raw_s, s, b, i, j0, ok, mscr, lane);
  float kf[32];
  if (p.p_drop > 0.f) keep32(p, kf, bh, i, j0);
#pragma unroll
  for (int e = 0; e < 32; ++e) {
    const float pe = ok ? exp2f(s[e] - lse2) : 0.f;
    const float k = p.p_drop > 0.f ? kf[e] : 1.f;
    pd[e] = pe * k;
    ds[e] = p.scale * pe * (__uint_as_float(raw_dp[e]) * k - delta);
  }
}

template <int HS>
struct DqCfg : Cfg<HS> {
  static constexpr int STAGES = 2;                                     // (K_j, V_j) ring
  static constexpr int BASE = Cfg<HS>::TILE * (2 + 2 * STAGES) + Cfg<HS>::PBYTES + 1024 + 256;
  static constexpr bool MSCR = BASE + kMaskScratch <= kMaxSmem;
  static constexpr int SMEM = BASE + (MSCR ? kMaskScratch : 0);
};

template <int HS>
__global__ void __launch_bounds__(NTHREADS, 1)
dq_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
          const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
          const __grid_constant__ CUtensorMap tmDQ, const __grid_constant__ Params p) {
  using C = DqCfg<HS>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = tc::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint8_t* sQ = smem;
  uint8_t* sDO = sQ + C::TILE;
  uint8_t* sKV = sDO + C::TILE;
  uint8_t* sDS = sKV + C::STAGES * 2 * C::TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + C::PBYTES);
  uint64_t* qdo_full = bars;
  uint64_t* kv_full = bars + 1;
  uint64_t* kv_empty = kv_full + C::STAGES;
  uint64_t* sd_full = kv_empty + C::STAGES;
  uint64_t* sd_empty = sd_full + 1;
  uint64_t* ds_full = sd_empty + 1;
  uint64_t* ds_empty = ds_full + 1;
  uint64_t* dq_full = ds_empty + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dq_full + 1);
  float* mscr_base = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, h = blockIdx.y, b = blockIdx.z;
  const int nkv = (p.Tk + BKV - 1) / BKV;

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmQ); tc::prefetch_tmap(&tmK); tc::prefetch_tmap(&tmV); tc::prefetch_tmap(&tmDO);
    tc::prefetch_tmap(&tmDQ);
    tc::mbar_init(qdo_full, 1);
    for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&kv_full[s], 1); tc::mbar_init(&kv_empty[s], 1); }
    tc::mbar_init(sd_full, 1); tc::mbar_init(sd_empty, 8);
    tc::mbar_init(ds_full, 8); tc::mbar_init(ds_empty, 1);
    tc::mbar_init(dq_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) { tc::tmem_alloc(tmem_slot, 512); tc::tmem_relinquish(); }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDQ = tmem_base + 256;

  if (warp == 0) {
    if (tc::elect_one()) {
      const uint32_t fb0 = tc::smem_u32(qdo_full);
      tc::mbar_expect_tx(fb0, 2 * C::TILE);
      load_tile<HS>(tc::smem_u32(sQ), &tmQ, fb0, h * HS, q0, b);
      load_tile<HS>(tc::smem_u32(sDO), &tmDO, fb0, h * HS, q0, b);
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < nkv; ++j) {
        tc::mbar_wait(tc::smem_u32(&kv_empty[stage]), phase ^ 1u);
        const uint32_t fb = tc::smem_u32(&kv_full[stage]);
        const uint32_t dK = tc::smem_u32(sKV + stage * 2 * C::TILE);
        tc::mbar_expect_tx(fb, 2 * C::TILE);
        load_tile<HS>(dK, &tmK, fb, h * HS, j * BKV, b);
        load_tile<HS>(dK + C::TILE, &tmV, fb, h * HS, j * BKV, b);
        if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (tc::elect_one()) {
      tc::mbar_wait(tc::smem_u32(qdo_full), 0);
      tc::tc_fence_after();
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < nkv; ++j) {
        const uint32_t aK = tc::smem_u32(sKV + stage * 2 * C::TILE);
        tc::mbar_wait(tc::smem_u32(&kv_full[stage]), phase);
        tc::mbar_wait(tc::smem_u32(sd_empty), ((uint32_t)j & 1u) ^ 1u);
        tc::tc_fence_after();
        mma_nt<HS>(tS, tc::smem_u32(sQ), aK);                 // S   = Q  K_j^T
        mma_nt<HS>(tDP, tc::smem_u32(sDO), aK + C::TILE);     // dPd = dO V_j^T
        tc::umma_commit(sd_full);
        tc::mbar_wait(tc::smem_u32(ds_full), (uint32_t)j & 1u);
        tc::tc_fence_after();
        mma_nn<HS>(tDQ, tc::smem_u32(sDS), aK, j > 0 ? 1u : 0u);   // dQ += dS K_j
        tc::umma_commit(&kv_empty[stage]);
        tc::umma_commit(ds_empty);
        if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
      }
      tc::umma_commit(dq_full);
    }
    __syncwarp();
  } else {
    const int qd = warp & 3;
    const int half = (warp - 2) >> 2;            // which 64 keys of every tile this thread owns
    const int r = qd * 32 + lane;
    const int i = q0 + r;
    const bool row_ok = i < p.T;
    const int bh = b * p.H + h;
    const uint32_t lane_off = (uint32_t)(qd * 32) << 16;
    // delta_i = dO_i . z_i over this head's columns (fp32; both threads of a row compute it, one keeps it for the
    // dK / dV kernel)
    float delta = 0.f, lse2 = 0.f;
    if (row_ok) {
      const bf16* zr = p.z + (long long)b * p.z_bs + (long long)i * p.z_rs + h * HS;
      const bf16* gr = p.dz + (long long)b * p.dz_bs + (long long)i * p.dz_rs + h * HS;
#pragma unroll
      for (int c = 0; c < HS; c += 8) {
        float a[8], g[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(zr + c)), a);
        unpack8(__ldg(reinterpret_cast<const uint4*>(gr + c)), g);
#pragma unroll
        for (int e = 0; e < 8; ++e) delta = fmaf(a[e], g[e], delta);
      }
      if (half == 0) p.delta[(long long)bh * p.T + i] = delta;
      lse2 = p.lse[(long long)bh * p.T + i] * kLog2e;
    }
    for (int j = 0; j < nkv; ++j) {
      tc::mbar_wait(sd_full, (uint32_t)j & 1u);
      tc::mbar_wait(ds_empty, ((uint32_t)j & 1u) ^ 1u);
      tc::tc_fence_after();
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const int ch = half * 2 + cc;
        uint32_t rs[32], rp[32];
        float pd[32], ds[32];
        tc::tmem_ld32(tS + lane_off + (uint32_t)(ch * 32), rs);
        tc::tmem_ld32(tDP + lane_off + (uint32_t)(ch * 32), rp);
        tc::tmem_ld_wait();
        bwd_row32(p, rs, rp, pd, ds, b, bh, i, j * BKV + ch * 32, row_ok, lse2, delta,
                  C::MSCR ? mscr_base + (warp - 2) * 1024 : nullptr, lane);
        store_tile_row32(sDS, r, ch * 32, ds);
      }
      tc::tc_fence_before();
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) { tc::mbar_arrive(sd_empty); tc::mbar_arrive(ds_full); }
    }
    tc::mbar_wait(dq_full, 0);
    tc::tc_fence_after();
    if (half < HS / 64)
      store_acc_rows<1>(tDQ + lane_off + (uint32_t)(half * 64), 1.f, sDS + (warp - 2) * 4096, &tmDQ, h * HS + half * 64,
                        q0 + qd * 32, b, lane);
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, 512);
  }
}

template <int HS>
struct DkvCfg : Cfg<HS> {
  static constexpr int STAGES = HS == 64 ? 2 : 1;                      // (Q_i, dO_i) ring
  static constexpr int BASE = Cfg<HS>::TILE * (2 + 2 * STAGES) + 2 * Cfg<HS>::PBYTES + 1024 + 256;
  static constexpr bool MSCR = BASE + kMaskScratch <= kMaxSmem;
  static constexpr int SMEM = BASE + (MSCR ? kMaskScratch : 0);
};

template <int HS>
__global__ void __launch_bounds__(NTHREADS, 1)
dkdv_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
            const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
            const __grid_constant__ CUtensorMap tmDK, const __grid_constant__ CUtensorMap tmDV,
            const __grid_constant__ Params p) {
  using C = DkvCfg<HS>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = tc::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint8_t* sK = smem;
  uint8_t* sV = sK + C::TILE;
  uint8_t* sQD = sV + C::TILE;                          // stage s: Q_i at s * 2 * TILE, dO_i after it
  uint8_t* sPD = sQD + C::STAGES * 2 * C::TILE;         // Pd tile, then dS tile
  uint8_t* sDS = sPD + C::PBYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + C::PBYTES);
  uint64_t* kv_full = bars;
  uint64_t* qd_full = bars + 1;
  uint64_t* qd_empty = qd_full + C::STAGES;
  uint64_t* sd_full = qd_empty + C::STAGES;
  uint64_t* sd_empty = sd_full + 1;
  uint64_t* pds_full = sd_empty + 1;
  uint64_t* pds_empty = pds_full + 1;
  uint64_t* out_full = pds_empty + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(out_full + 1);
  float* mscr_base = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * BKV, h = blockIdx.y, b = blockIdx.z;
  const int nq = (p.T + BQ - 1) / BQ;

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmQ); tc::prefetch_tmap(&tmK); tc::prefetch_tmap(&tmV); tc::prefetch_tmap(&tmDO);
    tc::prefetch_tmap(&tmDK); tc::prefetch_tmap(&tmDV);
    tc::mbar_init(kv_full, 1);
    for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&qd_full[s], 1); tc::mbar_init(&qd_empty[s], 1); }
    tc::mbar_init(sd_full, 1); tc::mbar_init(sd_empty, 8);
    tc::mbar_init(pds_full, 8); tc::mbar_init(pds_empty, 1);
    tc::mbar_init(out_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) { tc::tmem_alloc(tmem_slot, 512); tc::tmem_relinquish(); }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDV = tmem_base + 256, tDK = tmem_base + 384;

  if (warp == 0) {
    if (tc::elect_one()) {
      const uint32_t fb0 = tc::smem_u32(kv_full);
      tc::mbar_expect_tx(fb0, 2 * C::TILE);
      load_tile<HS>(tc::smem_u32(sK), &tmK, fb0, h * HS, k0, b);
      load_tile<HS>(tc::smem_u32(sV), &tmV, fb0, h * HS, k0, b);
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0; it < nq; ++it) {
        tc::mbar_wait(tc::smem_u32(&qd_empty[stage]), phase ^ 1u);
        const uint32_t fb = tc::smem_u32(&qd_full[stage]);
        const uint32_t dQ = tc::smem_u32(sQD + stage * 2 * C::TILE);
        tc::mbar_expect_tx(fb, 2 * C::TILE);
        load_tile<HS>(dQ, &tmQ, fb, h * HS, it * BQ, b);
        load_tile<HS>(dQ + C::TILE, &tmDO, fb, h * HS, it * BQ, b);
        if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (tc::elect_one()) {
      tc::mbar_wait(tc::smem_u32(kv_full), 0);
      tc::tc_fence_after();
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0; it < nq; ++it) {
        const uint32_t aQ = tc::smem_u32(sQD + stage * 2 * C::TILE);
        tc::mbar_wait(tc::smem_u32(&qd_full[stage]), phase);
        tc::mbar_wait(tc::smem_u32(sd_empty), ((uint32_t)it & 1u) ^ 1u);
        tc::tc_fence_after();
        mma_nt<HS>(tS, aQ, tc::smem_u32(sK));                  // S   = Q_i  K^T   (rows = queries)
        mma_nt<HS>(tDP, aQ + C::TILE, tc::smem_u32(sV));       // dPd = dO_i V^T
        tc::umma_commit(sd_full);
        tc::mbar_wait(tc::smem_u32(pds_full), (uint32_t)it & 1u);
        tc::tc_fence_after();
        mma_tn<HS>(tDV, tc::smem_u32(sPD), aQ + C::TILE, it > 0 ? 1u : 0u);   // dV += Pd^T dO_i
        mma_tn<HS>(tDK, tc::smem_u32(sDS), aQ, it > 0 ? 1u : 0u);             // dK += dS^T Q_i
        tc::umma_commit(&qd_empty[stage]);
        tc::umma_commit(pds_empty);
        if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
      }
      tc::umma_commit(out_full);
    }
    __syncwarp();
  } else {
    const int qd = warp & 3;
    const int half = (warp - 2) >> 2;            // which 64 keys of the tile this thread owns
    const int r = qd * 32 + lane;
    const int bh = b * p.H + h;
    const uint32_t lane_off = (uint32_t)(qd * 32) << 16;
    for (int it = 0; it < nq; ++it) {
      const int i = it * BQ + r;
      const bool row_ok = i < p.T;
      const float lse2 = row_ok ? p.lse[(long long)bh * p.T + i] * kLog2e : 0.f;
      const float delta = row_ok ? p.delta[(long long)bh * p.T + i] : 0.f;
      tc::mbar_wait(sd_full, (uint32_t)it & 1u);
      tc::mbar_wait(pds_empty, ((uint32_t)it & 1u) ^ 1u);
      tc::tc_fence_after();
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const int ch = half * 2 + cc;
        uint32_t rs[32], rp[32];
        float pd[32], ds[32];
        tc::tmem_ld32(tS + lane_off + (uint32_t)(ch * 32), rs);
        tc::tmem_ld32(tDP + lane_off + (uint32_t)(ch * 32), rp);
        tc::tmem_ld_wait();
        bwd_row32(p, rs, rp, pd, ds, b, bh, i, k0 + ch * 32, row_ok, lse2, delta,
                  C::MSCR ? mscr_base + (warp - 2) * 1024 : nullptr, lane);
        if (!row_ok) {
#pragma unroll
          for (int e = 0; e < 32; ++e) ds[e] = 0.f;     // rows past T must not reach dK / dV
        }
        store_tile_row32(sPD, r, ch * 32, pd);
        store_tile_row32(sDS, r, ch * 32, ds);
      }
      tc::tc_fence_before();
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) { tc::mbar_arrive(sd_empty); tc::mbar_arrive(pds_full); }
    }
    tc::mbar_wait(out_full, 0);
    tc::tc_fence_after();
    // rows of the accumulators are keys here; one half of the warps stores dV, the other dK
    if (half == 0)
      store_acc_rows<HS / 64>(tDV + lane_off, 1.f, sPD + (warp - 2) * 4096, &tmDV, h * HS, k0 + qd * 32, b, lane);
    else
      store_acc_rows<HS / 64>(tDK + lane_off, 1.f, sPD + (warp - 2) * 4096, &tmDK, h * HS, k0 + qd * 32, b, lane);
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace fa
