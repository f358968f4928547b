// tcgen05 / TMEM / TMA tensor-core kernels for sm_100a.
//
// One persistent, warp-specialised kernel template covers the three contraction shapes of the
// training hot path:
//   KIND_GEMM  : D[M,N] = A . B^T with A and B each either K-major or MN-major in memory
//                (Linear forward / dgrad / wgrad -- the nn.Linear call sites of the reference,
//                nn/attention.py:25,85 and arch/transformer.py:24-26).
//   KIND_CONV  : implicit-GEMM convolution, A gathered by TMA in im2col mode from an NHWC tensor,
//                B = filter matrix (K-major for fprop, MN-major for dgrad).
//   KIND_WGRAD : filter gradient; the reduction runs over output pixels, so both operands are
//                MN-major: A = dY tiles, B = im2col tiles of X.
//
// Roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer,
// warps 2..5 = epilogue (TMEM -> registers -> global).  Three mbarrier rings: smem full/empty
// between producer and MMA, TMEM full/empty between MMA and epilogue (two accumulator buffers so
// the epilogue of tile i overlaps the main loop of tile i+1).
#include <stdlib.h>
#include <string.h>

#include "tc_ptx.cuh"

namespace {

constexpr int BM = 128;       // UMMA M (cta_group::1)
constexpr int BK = 64;        // 64 bf16 = one 128-byte swizzle row
constexpr int A_BYTES = BM * BK * 2;
constexpr int NUM_THREADS = 320;        // TMA producer warp, MMA warp, 2 groups of 4 epilogue warps
constexpr int NUM_THREADS_NARROW = 192; // ... or 1 group: the launch decides (see launch_inst2)
constexpr int MAX_TAPS = 128;

enum { KIND_GEMM = 0, KIND_CONV = 1, KIND_WGRAD = 2, KIND_BGEMM = 3 };

struct TapDesc {
  int16_t ow, oh;  // filter-tap offsets handed to the im2col TMA
  int32_t bcol;    // column (or row-block) of this tap inside the filter matrix
};

struct TcParams {
  int M, N;             // extents of D (WGRAD: N = channels per tap)
  int k_iters;          // 64-deep iterations of the full reduction
  int splits;           // split-K factor (partials go to ws)
  int n_mblk, n_nblk;   // tile grid
  // im2col base-pixel mapping: linear index -> (img, i, j) -> (w, h) = (j*ts_w+lo_w, i*ts_h+lo_h)
  int Wo, HoWo, ts_w, ts_h, lo_w, lo_h;
  int cchunks;          // CONV: 64-channel chunks per tap
  int ntaps;
  int ncblk;            // WGRAD: 64-channel sub-blocks per tap
  // output mapping
  long long ldd;        // row stride of D (elements)
  int pix_map;          // CONV: rows of D are scattered pixels (strided dgrad)
  int OW, OHOW, os_h, os_w, oa, ob;
  void* D;              // bf16 output (when ws == nullptr)
  float* ws;            // fp32 output / split-K partials [splits][M][ldw]
  long long ldw;
  const float* scale;   // per-column scale (nullable)
  const float* shift;   // per-column shift / bias (nullable)
  const bf16* residual; // added before ReLU (nullable), same indexing as D
  unsigned char* mask_out;       // FUSED: bit c&7 of byte [row][c>>3] = (stored value > 0)  (1-bit ReLU mask, nullable)
  const unsigned char* mask_in;  // FUSED: D = bit ? acc : 0 -- the backward of that ReLU (nullable); N % 64 == 0
  int relu;
  int tma_store;        // epilogue writes through the third tensor map (dense row mapping only)
  int reduce_add;       // TMA stores add into D instead of overwriting it (gradient accumulation)
  float* stat_partials; // STATS: per-CTA column sums / sums of squares [n_nblk][stat_rows][2][BN]
  int stat_rows;        // = gridDim.x / n_nblk
  int force_grid;       // launch exactly this many CTAs (STATS needs gridDim.x % n_nblk == 0)
  int cg;               // host only: 2 = launch as CTA pairs (cta_group::2), else single CTAs
  // CONV, single CTAs: the whole B operand (filter) of this CTA's n-block stays in shared memory for the life of the
  // CTA and the ring carries A tiles only (ring_stages stages of A_BYTES).  Needs a constant n-block per CTA
  // (gridDim.x % n_nblk == 0) and splits == 1.
  int b_resident, ring_stages;
  // Two CTAs per SM (TcParams::dual): narrow tiles are paced by the two single-thread role loops of a CTA, not by
  // memory or the tensor pipe (profiles/r02_tc_role_elimination.txt), so a second CTA with its own producer and MMA
  // threads on the same SM doubles the step rate.  Each CTA then gets `nst` ring stages in `ring_area` bytes, one
  // epilogue group (192 threads) and half of the SM's TMEM.
  int dual, nst, ring_area;
  int no_b_resident;    // host only: the packed-pixel first layer -- never the resident form (measured slower with it)
  // BGEMM: `splits` independent problems z = b * bg_heads + h, each with the full K range.  Operand X lives in a
  // 3-D tensor; its slab index is b (x_zdiv = 1: heads of a batch share rows, head h starts x_hoff * h elements
  // into the inner dimension) or z (x_zdiv = 0).
  int bg_heads;
  int a_zdiv, b_zdiv, d_zdiv;
  int a_hoff, b_hoff, d_hoff;
  long long d_bstride;  // elements between slabs of a bf16 D
  int d_slabs, d_inner; // extents of the bf16 D tensor map (slabs, inner elements)
  TapDesc taps[MAX_TAPS];
};

// CG = CTAs per MMA (cta_group): 1, or 2 = a CTA pair computing a 256 x BN tile with each CTA staging its own
// 128 rows of A and HALF of B (the pair's tensor cores read both halves), which halves the operand bytes every
// SM pulls through L2 per FLOP and deepens the pipeline.
// Shared memory given to the operand ring.  The weight-gradient kernels take less: they run on a side stream
// beside the BatchNorm backward kernels of the main stream, whose blocks need 16 KB each to be co-resident.
constexpr int kStageBudget = 196608;
constexpr int kStageBudgetWgrad = 163840;      // 3 stages of 48 KB (BN = 256)
constexpr int kStageBudgetWgradPair = 131072;  // 4 stages of 32 KB
template <int BN, int CG = 1, int KIND = KIND_GEMM>
struct Cfg {
  static constexpr int BUDGET = KIND == KIND_WGRAD ? (CG == 2 ? kStageBudgetWgradPair : kStageBudgetWgrad) : kStageBudget;
  static constexpr int B_BYTES = BN * BK * 2 / CG;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BUDGET / STAGE_BYTES) > 8 ? 8 : (BUDGET / STAGE_BYTES);
  static constexpr int ACC_STRIDE = BN == 192 ? 256 : BN;  // TMEM columns between the two accumulators
  static constexpr int TMEM_COLS = 2 * ACC_STRIDE;         // 128 / 256 / 512: powers of two
  static constexpr int STAGING_BYTES = 8 * 4096;           // per epilogue warp: 32 rows x 128 B
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + STAGING_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

// FUSED = the bf16 epilogue applies scale / shift / residual / ReLU; the plain variant (most convolutions of
// a BatchNorm network, every dgrad) compiles to convert + store only.
// STATS = the epilogue also accumulates, per output channel, the sum and the sum of squares of the bf16 values it
// stores (BatchNorm batch statistics without a second pass over the activation).
template <int BN, int KIND, bool A_MN, bool B_MN, bool FUSED, bool STATS, int CG>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
          const __grid_constant__ CUtensorMap tmD, const __grid_constant__ TcParams p) {
  using C = Cfg<BN, CG, KIND>;
  static_assert(CG == 1 || (BN % 128 == 0), "a CTA pair splits B in halves of whole 64-wide blocks");
  const int cta_rank = CG == 2 ? (int)tc::cluster_ctarank() : 0;   // 0 = the CTA that issues the MMAs
  const int cluster_id = blockIdx.x / CG, num_clusters = gridDim.x / CG;
  const int n_pm = (p.n_mblk + CG - 1) / CG;                       // M tiles of 128*CG rows
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = tc::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

  const uint32_t ring_area = p.dual ? (uint32_t)p.ring_area : (uint32_t)(C::STAGES * C::STAGE_BYTES);
  uint8_t* staging = smem + ring_area;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + C::STAGING_BYTES);
  constexpr int MAXS = 8;   // barrier slots: the resident-filter form may run more ring stages than C::STAGES
  uint64_t* empty_bar = full_bar + MAXS;
  uint64_t* tfull_bar = empty_bar + MAXS;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* bres_bar = tempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bres_bar + 1);
  // resident-B form (see TcParams::b_resident): B sits at the end of the ring area, the ring strides by A_BYTES
  const bool bres = KIND == KIND_CONV && CG == 1 && p.b_resident != 0;
  const int nstages = bres ? p.ring_stages : (p.dual ? p.nst : C::STAGES);
  const uint32_t stage_bytes = bres ? (uint32_t)A_BYTES : (uint32_t)C::STAGE_BYTES;
  const uint32_t bres_off = ring_area - (uint32_t)(p.k_iters * C::B_BYTES);

  // warp index through a shuffle: the compiler then knows it is warp-uniform and keeps the role loops' state
  // (stage counters, barrier addresses, TMA coordinates, MMA descriptors) in uniform registers
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int num_tiles = n_pm * p.n_nblk * p.splits;

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmA);
    tc::prefetch_tmap(&tmB);
    if (p.tma_store) tc::prefetch_tmap(&tmD);
    for (int s = 0; s < MAXS; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(&tfull_bar[a], 1);
      tc::mbar_init(&tempty_bar[a], 4 * CG);  // one arrive per epilogue warp (of every CTA of the pair)
    }
    tc::mbar_init(bres_bar, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) {
    if (CG == 2) { tc::tmem_alloc_pair(tmem_slot, C::TMEM_COLS); tc::tmem_relinquish_pair(); }
    else { tc::tmem_alloc(tmem_slot, C::TMEM_COLS); tc::tmem_relinquish(); }
  }
  tc::tc_fence_before();
  if (CG == 2) tc::cluster_sync_all();  // barriers of both CTAs initialised before any remote arrive / TMA
  else __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const uint32_t smem_a0 = tc::smem_u32(smem);              // stage s: A at smem_a0 + s * STAGE_BYTES, B after it
  const uint32_t full0 = tc::smem_u32(full_bar), empty0 = tc::smem_u32(empty_bar);

  if (warp == 0) {
    // ===================================================================== TMA producer
    // One elected thread.  Everything that does not change from one K step to the next is computed per tile; a
    // K step costs a barrier wait, the expect_tx arrive and the TMA instructions (the step rate of this thread,
    // not the memory system, bounded the narrow-N tiles before: ~800 cycles per step).
    if (tc::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      constexpr int NB_LOCAL = BN / 64 / CG;   // 64-wide B blocks staged by this CTA (MN-major B)
      const int jb0 = cta_rank * NB_LOCAL;     // first 64-wide B block this CTA stages
      if (KIND == KIND_CONV && bres) {
        // every tile of this CTA has n-block (cluster_id % n_nblk): fetch its whole filter slice once
        const int n0r = (cluster_id % p.n_nblk) * BN;
        const uint32_t rb = tc::smem_u32(bres_bar);
        tc::mbar_expect_tx(rb, (uint32_t)(p.k_iters * C::B_BYTES));
        int tap = 0, cc = 0;
        for (int it = 0; it < p.k_iters; ++it) {
          const TapDesc td = p.taps[tap];
          const uint32_t dB = smem_a0 + bres_off + (uint32_t)(it * C::B_BYTES);
          if (!B_MN) {
            tc::tma_ld2<1>(dB, &tmB, rb, td.bcol + cc * BK, n0r);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tc::tma_ld2<1>(dB + j * 8192, &tmB, rb, td.bcol + n0r + j * 64, cc * BK);
          }
          if (++cc == p.cchunks) { cc = 0; ++tap; }
        }
      }
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        int t = tile;
        const int nblk = t % p.n_nblk; t /= p.n_nblk;
        const int mblk = (t % n_pm) * CG + cta_rank;
        const int split = t / n_pm;
        const int k_lo = KIND == KIND_BGEMM ? 0 : (int)(((long long)p.k_iters * split) / p.splits);
        const int k_hi = KIND == KIND_BGEMM ? p.k_iters : (int)(((long long)p.k_iters * (split + 1)) / p.splits);
        const int m0 = mblk * BM;
        const int n0 = nblk * BN;
        // BGEMM: slab and inner offset of this problem's operands
        int za = 0, zb = 0, a_off = 0, b_off = 0;
        if (KIND == KIND_BGEMM) {
          const int bb = split / p.bg_heads, hh = split - bb * p.bg_heads;
          za = p.a_zdiv ? bb : split; zb = p.b_zdiv ? bb : split;
          a_off = hh * p.a_hoff; b_off = hh * p.b_hoff;
        }
        // CONV: base pixel of this M tile
        int img0 = 0, w0 = 0, h0 = 0;
        if (KIND == KIND_CONV) {
          img0 = m0 / p.HoWo;
          const int rem = m0 - img0 * p.HoWo;
          const int i = rem / p.Wo, j = rem - i * p.Wo;
          w0 = j * p.ts_w + p.lo_w;
          h0 = i * p.ts_h + p.lo_h;
        }
        // WGRAD: the N dimension is the flattened (tap, channel) space in 64-wide sub-blocks; a tile covers
        // BN/64 consecutive sub-blocks, which may straddle taps (e.g. 3 taps x 64 channels for BN = 192).
        int sb0 = 0, nsb_valid = BN / 64;
        if (KIND == KIND_WGRAD) {
          sb0 = nblk * (BN / 64);
          const int total_sb = p.ntaps * p.ncblk;  // ncblk = 64-channel blocks per tap
          if (sb0 + nsb_valid > total_sb) nsb_valid = total_sb - sb0;
        }
        int sb_c0[BN / 64];
        uint16_t sb_ow[BN / 64], sb_oh[BN / 64];
        if (KIND == KIND_WGRAD) {
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) {
            const int sb = sb0 + (j < nsb_valid ? j : 0);
            const int tap = sb / p.ncblk;
            sb_c0[j] = (sb - tap * p.ncblk) * 64;
            sb_ow[j] = (uint16_t)p.taps[tap].ow;
            sb_oh[j] = (uint16_t)p.taps[tap].oh;
          }
        }
        // loads of a pair credit the even CTA's barrier; it alone announces the bytes of BOTH CTAs
        const uint32_t tx_bytes = KIND == KIND_WGRAD ? (uint32_t)(CG * A_BYTES + nsb_valid * 8192)
                                                     : (bres ? (uint32_t)A_BYTES : (uint32_t)(CG * C::STAGE_BYTES));
        // CONV: (tap, channel chunk) of the K step, advanced without divisions
        int tap = 0, cc = 0;
        TapDesc td = p.taps[0];
        if (KIND == KIND_CONV) {
          tap = k_lo / p.cchunks;
          cc = k_lo - tap * p.cchunks;
          td = p.taps[tap];
        }
        // WGRAD: pixel coordinates of the K step, advanced by 64 pixels per step
        int wg_img = 0, wg_i = 0, wg_j = 0;
        const int wg_ho = KIND == KIND_WGRAD ? p.HoWo / p.Wo : 1;
        if (KIND == KIND_WGRAD) {
          const int pix0 = k_lo * BK;
          wg_img = pix0 / p.HoWo;
          const int rem = pix0 - wg_img * p.HoWo;
          wg_i = rem / p.Wo;
          wg_j = rem - wg_i * p.Wo;
        }
        for (int it = k_lo; it < k_hi; ++it) {
          const uint32_t fb = full0 + stage * 8;
          const uint32_t sA = smem_a0 + stage * stage_bytes;
          const uint32_t sB = sA + A_BYTES;
          tc::mbar_wait(empty0 + stage * 8, phase ^ 1u);
          if (cta_rank == 0) tc::mbar_expect_tx(fb, tx_bytes);
          if (KIND == KIND_GEMM) {
            const int kc = it * BK;
            if (!A_MN) {
              tc::tma_ld2<CG>(sA, &tmA, fb, kc, m0);
            } else {
              tc::tma_ld2<CG>(sA, &tmA, fb, m0, kc);
              tc::tma_ld2<CG>(sA + 8192, &tmA, fb, m0 + 64, kc);
            }
            if (!B_MN) {
              tc::tma_ld2<CG>(sB, &tmB, fb, kc, n0 + cta_rank * (BN / CG));   // box rows = BN / CG
            } else {
#pragma unroll
              for (int j = 0; j < NB_LOCAL; ++j) tc::tma_ld2<CG>(sB + j * 8192, &tmB, fb, n0 + (jb0 + j) * 64, kc);
            }
          } else if (KIND == KIND_BGEMM) {
            const int kc = it * BK;
            if (!A_MN) {
              tc::tma_ld3(sA, &tmA, fb, kc + a_off, m0, za);
            } else {
              tc::tma_ld3(sA, &tmA, fb, m0 + a_off, kc, za);
              tc::tma_ld3(sA + 8192, &tmA, fb, m0 + 64 + a_off, kc, za);
            }
            if (!B_MN) {
              tc::tma_ld3(sB, &tmB, fb, kc + b_off, n0, zb);
            } else {
#pragma unroll
              for (int j = 0; j < NB_LOCAL; ++j) tc::tma_ld3(sB + j * 8192, &tmB, fb, n0 + j * 64 + b_off, kc, zb);
            }
          } else if (KIND == KIND_CONV) {
            tc::tma_ldi<CG>(sA, &tmA, fb, cc * BK, w0, h0, img0, (uint16_t)td.ow, (uint16_t)td.oh);
            if (bres) {
              // B is resident (loaded once below, before the first tile)
            } else if (!B_MN) {
              tc::tma_ld2<CG>(sB, &tmB, fb, td.bcol + cc * BK, n0 + cta_rank * (BN / CG));
            } else {
#pragma unroll
              for (int j = 0; j < NB_LOCAL; ++j)
                tc::tma_ld2<CG>(sB + j * 8192, &tmB, fb, td.bcol + n0 + (jb0 + j) * 64, cc * BK);
            }
            if (++cc == p.cchunks) {
              cc = 0;
              ++tap;
              td = p.taps[tap < MAX_TAPS ? tap : 0];
            }
          } else {  // KIND_WGRAD: reduction over pixels
            const int pix0 = it * BK;
            tc::tma_ld2<CG>(sA, &tmA, fb, m0, pix0);
            tc::tma_ld2<CG>(sA + 8192, &tmA, fb, m0 + 64, pix0);
            const int cw = wg_j * p.ts_w + p.lo_w, ch = wg_i * p.ts_h + p.lo_h;
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) {
              if (j >= jb0 && j < jb0 + NB_LOCAL && j < nsb_valid)
                tc::tma_ldi<CG>(sB + (j - jb0) * 8192, &tmB, fb, sb_c0[j], cw, ch, wg_img, sb_ow[j], sb_oh[j]);
            }
            // next step starts 64 pixels further: carry into rows and images without dividing
            wg_j += BK;
            while (wg_j >= p.Wo) { wg_j -= p.Wo; if (++wg_i == wg_ho) { wg_i = 0; ++wg_img; } }
          }
          if (++stage == nstages) { stage = 0; phase ^= 1u; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    if (cta_rank == 0 && tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_bf16(BM * CG, BN, A_MN, B_MN);
      // descriptors of stage 0; a stage adds STAGE_BYTES >> 4 and a 16-wide K slice adds KADV to the address field
      const uint64_t adesc0 = A_MN ? tc::make_smem_desc(smem_a0, 8192, 1024) : tc::make_smem_desc(smem_a0, 16, 1024);
      const uint64_t bdesc0 = B_MN ? tc::make_smem_desc(smem_a0 + A_BYTES, 8192, 1024)
                                   : tc::make_smem_desc(smem_a0 + A_BYTES, 16, 1024);
      constexpr uint32_t A_KADV = (A_MN ? 2048 : 32) >> 4, B_KADV = (B_MN ? 2048 : 32) >> 4;
      const uint32_t tfull0 = tc::smem_u32(tfull_bar), tempty0 = tc::smem_u32(tempty_bar);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint64_t bdesc_res = B_MN ? tc::make_smem_desc(smem_a0 + bres_off, 8192, 1024)
                                      : tc::make_smem_desc(smem_a0 + bres_off, 16, 1024);
      if (bres) {
        tc::mbar_wait(tc::smem_u32(bres_bar), 0);
        tc::tc_fence_after();
      }
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int split = tile / (p.n_nblk * n_pm);
        const int k_lo = KIND == KIND_BGEMM ? 0 : (int)(((long long)p.k_iters * split) / p.splits);
        const int k_hi = KIND == KIND_BGEMM ? p.k_iters : (int)(((long long)p.k_iters * (split + 1)) / p.splits);
        tc::mbar_wait(tempty0 + acc * 8, acc_phase ^ 1u);
        tc::tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * C::ACC_STRIDE);
        uint32_t accumulate = 0;
        for (int it = k_lo; it < k_hi; ++it) {
          tc::mbar_wait(full0 + stage * 8, phase);
          tc::tc_fence_after();
          const uint64_t ad = adesc0 + (uint64_t)(stage * (stage_bytes >> 4));
          const uint64_t bd = bres ? bdesc_res + (uint64_t)(it * (C::B_BYTES >> 4))
                                   : bdesc0 + (uint64_t)(stage * (C::STAGE_BYTES >> 4));
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            if (CG == 2) tc::umma_bf16_pair(d_tmem, ad + k * A_KADV, bd + k * B_KADV, idesc, accumulate);
            else tc::umma_bf16(d_tmem, ad + k * A_KADV, bd + k * B_KADV, idesc, accumulate);
            accumulate = 1;
          }
          // frees the smem slot (in both CTAs of a pair) once these MMAs retire
          tc::umma_commit_u32<CG>(empty0 + stage * 8);
          if (++stage == nstages) { stage = 0; phase ^= 1u; }
        }
        // accumulator complete -> epilogue warps (of both CTAs)
        tc::umma_commit_u32<CG>(tfull0 + acc * 8);
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
    __syncwarp();
  } else {
    // ===================================================================== epilogue (warps 2..9)
    // One or two groups of four warps (blockDim.x = 192 or 320).  With two, group g drains every other tile of
    // this CTA, so the epilogues of two consecutive tiles overlap (memory-bound layers spend longer here than in
    // their MMAs); the side-stream weight-gradient launches use one group to stay small next to other kernels.
    // TMEM -> registers (one row per thread) -> per-warp smem staging (XOR-swizzled 16-byte pieces) ->
    // TMA store, or coalesced 128-bit global stores: each store instruction writes 4 rows x 128 contiguous bytes.
    const int q = warp & 3;          // TMEM lane quarter this warp may read
    const int grp = (warp - 2) >> 2; // accumulator buffer this warp's group owns
    uint8_t* stage_w = staging + (grp * 4 + q) * 4096;
    float st[STATS ? BN / 64 : 1][4];       // STATS: {sum, sum} {sumsq, sumsq} of columns 2*lane, 2*lane+1 per chunk
#pragma unroll
    for (int c = 0; c < (STATS ? BN / 64 : 1); ++c) st[c][0] = st[c][1] = st[c][2] = st[c][3] = 0.f;
    const int ngrp = blockDim.x >= NUM_THREADS ? 2 : 1;
    int ti = grp;   // index of the tile in this CTA's sequence: accumulator buffer ti & 1, barrier phase (ti >> 1) & 1
    for (int tile = cluster_id + grp * num_clusters; tile < num_tiles; tile += ngrp * num_clusters, ti += ngrp) {
      const int acc = ti & 1;
      const uint32_t acc_phase = (uint32_t)(ti >> 1) & 1u;
      int t = tile;
      const int nblk = t % p.n_nblk; t /= p.n_nblk;
      const int mblk = (t % n_pm) * CG + cta_rank;
      const int split = t / n_pm;
      const int row = mblk * BM + q * 32 + lane;
      const bool row_ok = row < p.M;
      const int col_base = nblk * BN;
      const int col_lim_local = p.N - col_base;  // valid columns inside the tile
      long long out_row = row;
      if (KIND == KIND_CONV && p.pix_map && row_ok) {
        const int img = row / p.HoWo;
        const int rem = row - img * p.HoWo;
        const int i = rem / p.Wo, j = rem - i * p.Wo;
        out_row = (long long)img * p.OHOW + (long long)(i * p.os_h + p.oa) * p.OW + (j * p.os_w + p.ob);
      }
      // byte address of this thread's row at the tile's first column (0 = row does not exist)
      unsigned long long row_ptr = 0;
      if (row_ok) {
        row_ptr = p.ws != nullptr
                      ? reinterpret_cast<unsigned long long>(p.ws + ((long long)split * p.M + row) * p.ldw + col_base)
                      : reinterpret_cast<unsigned long long>(reinterpret_cast<bf16*>(p.D) + out_row * p.ldd + col_base);
      }
      tc::mbar_wait(&tfull_bar[acc], acc_phase);
      tc::tc_fence_after();
      const uint32_t t_acc = tmem_base + (uint32_t)(acc * C::ACC_STRIDE) + ((uint32_t)(q * 32) << 16);
      // staged write of one 128-byte-per-row chunk: v = this thread's 8 x 16 B, chunk_bytes0 = byte offset of
      // the chunk inside the row, valid16 = number of 16-byte pieces of the chunk that are inside D
      // TMA-store variant: the swizzled staging buffer is exactly the SWIZZLE_128B box the output tensor map
      // expects; one lane hands it to the TMA engine, which also clips rows/columns outside D.
      auto flush_tma = [&](const uint4* v, int col0, int row0) {
        if (lane == 0) tc::tma_store_wait_read<0>();  // this warp's previous store has drained the buffer
        __syncwarp();
        uint8_t* buf = stage_w;
        uint4* srow = reinterpret_cast<uint4*>(buf + lane * 128);
#pragma unroll
        for (int j = 0; j < 8; ++j) srow[j ^ (lane & 7)] = v[j];
        tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (p.ws != nullptr) tc::tma_store_3d(&tmD, buf, col0, row0, split);
          else if (KIND == KIND_BGEMM) {
            const int bb = split / p.bg_heads, hh = split - bb * p.bg_heads;
            tc::tma_store_3d(&tmD, buf, col0 + hh * p.d_hoff, row0, p.d_zdiv ? bb : split);
          }
          else if (p.reduce_add) tc::tma_reduce_add_2d(&tmD, buf, col0, row0);
          else tc::tma_store_2d(&tmD, buf, col0, row0);
          tc::tma_store_commit();
        }
      };
      auto flush = [&](const uint4* v, int chunk_byte0, int valid16) {
        uint4* srow = reinterpret_cast<uint4*>(stage_w + lane * 128);
#pragma unroll
        for (int j = 0; j < 8; ++j) srow[j ^ (lane & 7)] = v[j];
        __syncwarp();
        const int j = lane & 7;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rr = (lane >> 3) + 4 * i;
          const uint4 val = reinterpret_cast<const uint4*>(stage_w + rr * 128)[j ^ (rr & 7)];
          const unsigned long long pr = __shfl_sync(0xffffffffu, row_ptr, rr);
          if (pr != 0 && j < valid16) *reinterpret_cast<uint4*>(pr + chunk_byte0 + j * 16) = val;
        }
        __syncwarp();
      };
      if (p.ws != nullptr) {
        // fp32 partials / fp32 output: 32 columns = 128 bytes per row per chunk
#pragma unroll 1
        for (int chunk = 0; chunk < BN / 32; ++chunk) {
          uint32_t r[32];
          __syncwarp();
          tc::tmem_ld32(t_acc + (uint32_t)(chunk * 32), r);
          tc::tmem_ld_wait();
          int valid16 = (col_lim_local - chunk * 32) / 4;
          valid16 = valid16 < 0 ? 0 : (valid16 > 8 ? 8 : valid16);
          if (p.tma_store) flush_tma(reinterpret_cast<const uint4*>(r), col_base + chunk * 32, mblk * BM + q * 32);
          else flush(reinterpret_cast<const uint4*>(r), chunk * 128, valid16);
        }
      } else {
        // bf16 output with the fused epilogue: 64 columns = 128 bytes per row per chunk
        auto do_chunk = [&](const int chunk) {
          uint32_t r[64];
          __syncwarp();
          tc::tmem_ld32(t_acc + (uint32_t)(chunk * 64), r);
          tc::tmem_ld32(t_acc + (uint32_t)(chunk * 64 + 32), r + 32);
          tc::tmem_ld_wait();
          uint4 out[8];
          const int c_first = col_base + chunk * 64;
          const bf16* res = (FUSED && p.residual && row_ok) ? p.residual + out_row * p.ldd + c_first : nullptr;
          // 1-bit ReLU masks: 8 bytes per row and 64-column chunk
          uint2 mbits = make_uint2(0xffffffffu, 0xffffffffu), obits = make_uint2(0u, 0u);
          const bool chunk_in = chunk * 64 < col_lim_local;
          if (FUSED && p.mask_in && row_ok && chunk_in)
            mbits = *reinterpret_cast<const uint2*>(p.mask_in + out_row * (p.ldd >> 3) + (c_first >> 3));
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[8 * g + e]);
            if (FUSED && chunk * 64 + g * 8 < col_lim_local) {
              const int c = c_first + g * 8;
              if (p.scale) {
                const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.scale + c));
                const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.scale + c + 4));
                v[0] *= s0.x; v[1] *= s0.y; v[2] *= s0.z; v[3] *= s0.w;
                v[4] *= s1.x; v[5] *= s1.y; v[6] *= s1.z; v[7] *= s1.w;
              }
              if (p.shift) {
                const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.shift + c));
                const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.shift + c + 4));
                v[0] += s0.x; v[1] += s0.y; v[2] += s0.z; v[3] += s0.w;
                v[4] += s1.x; v[5] += s1.y; v[6] += s1.z; v[7] += s1.w;
              }
              if (res) {
                float rr[8];
                unpack8(*reinterpret_cast<const uint4*>(res + g * 8), rr);
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] += rr[e];
              }
              if (p.relu) {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
              }
              if (p.mask_in) {
                const unsigned byte = ((g < 4 ? mbits.x : mbits.y) >> (8 * (g & 3))) & 0xffu;
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = ((byte >> e) & 1u) ? v[e] : 0.f;
              }
              if (p.mask_out) {
                unsigned byte = 0;
#pragma unroll
                for (int e = 0; e < 8; ++e) byte |= (v[e] > 0.f ? 1u : 0u) << e;
                if (g < 4) obits.x |= byte << (8 * g); else obits.y |= byte << (8 * (g - 4));
              }
            }
            out[g] = pack8(v);
          }
          if (FUSED && p.mask_out && row_ok && chunk_in)
            *reinterpret_cast<uint2*>(p.mask_out + out_row * (p.ldd >> 3) + (c_first >> 3)) = obits;
          int valid16 = (col_lim_local - chunk * 64) / 8;
          valid16 = valid16 < 0 ? 0 : (valid16 > 8 ? 8 : valid16);
          if (p.tma_store) flush_tma(out, c_first, mblk * BM + q * 32);
          else flush(out, chunk * 128, valid16);
          if constexpr (STATS) {
            // column sums over this warp's 32 rows, read back from the swizzled staging tile: lane owns the
            // 4-byte word holding columns 2*lane, 2*lane+1 (conflict-free: 32 distinct words per row)
            float a0 = 0.f, a1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 8
            for (int rr = 0; rr < 32; ++rr) {
              const uint32_t wv = *reinterpret_cast<const uint32_t*>(
                  stage_w + rr * 128 + ((((lane >> 2) ^ (rr & 7)) << 4) | ((lane & 3) << 2)));
              const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&wv));
              a0 += f.x; a1 += f.y;
              q0 = fmaf(f.x, f.x, q0); q1 = fmaf(f.y, f.y, q1);
            }
            st[STATS ? chunk : 0][0] += a0; st[STATS ? chunk : 0][1] += a1;
            st[STATS ? chunk : 0][2] += q0; st[STATS ? chunk : 0][3] += q1;
          }
        };
        if constexpr (STATS) {
#pragma unroll
          for (int chunk = 0; chunk < BN / 64; ++chunk) do_chunk(chunk);
        } else {
#pragma unroll 1
          for (int chunk = 0; chunk < BN / 64; ++chunk) do_chunk(chunk);
        }
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) {  // the accumulator buffer may be overwritten once every epilogue warp of the pair has drained it
        if (CG == 2 && cta_rank != 0) tc::mbar_arrive_remote(&tempty_bar[acc], 0);
        else tc::mbar_arrive(&tempty_bar[acc]);
      }
    }
    if (p.tma_store && lane == 0) tc::tma_store_wait<0>();
    if constexpr (STATS) {
      // The number of clusters is a multiple of n_nblk: every tile of this CTA has the same n-block.  The eight
      // epilogue warps combine their column sums through shared memory (their staging buffers are idle now) and
      // the CTA writes ONE partial row, so the finalize kernel reads 8x fewer rows (it was 0.6 ms of a ResNet-50
      // step for O(C) work).  Fixed order of the eight terms: bitwise reproducible.
      __syncwarp();   // lane 0 has waited for this warp's last TMA store to finish reading the staging buffer
      float* mine = reinterpret_cast<float*>(staging + (grp * 4 + q) * 4096);
#pragma unroll
      for (int c = 0; c < BN / 64; ++c) {
        *reinterpret_cast<float2*>(mine + c * 64 + 2 * lane) = make_float2(st[c][0], st[c][1]);
        *reinterpret_cast<float2*>(mine + BN + c * 64 + 2 * lane) = make_float2(st[c][2], st[c][3]);
      }
      const int nepi = (int)(blockDim.x >> 5) - 2;      // 8 epilogue warps, or 4 when two CTAs share the SM
      asm volatile("bar.sync 1, %0;" ::"r"(nepi * 32) : "memory");
      const int nb = cluster_id % p.n_nblk;
      const int prow = (cluster_id / p.n_nblk) * CG + cta_rank;
      float* base = p.stat_partials + ((long long)(nb * p.stat_rows + prow) * 2) * BN;
      for (int e = (warp - 2) * 32 + lane; e < 2 * BN; e += nepi * 32) {
        float a = 0.f;
        for (int w = 0; w < nepi; ++w) a += reinterpret_cast<const float*>(staging + w * 4096)[e];
        base[e] = a;
      }
    }
  }

  tc::tc_fence_before();
  if (CG == 2) tc::cluster_sync_all();  // the peer may still be reading this CTA's smem / signalling its barriers
  else __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    if (CG == 2) tc::tmem_dealloc_pair(tmem_base, C::TMEM_COLS);
    else tc::tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// fp32 GEMM on the bf16 tensor cores: x = x1 + x2 + x3 with x1 = bf16(x), x2 = bf16(x - x1), x3 = bf16(x - x1 - x2)
// (24 mantissa bits in three pieces), and a.b ~= a1b1 + a1b2 + a2b1 + a1b3 + a2b2 + a3b1: every product of two bf16
// numbers is exact in the fp32 accumulator and the dropped terms are below 2^-24 relative.  The six products are
// ONE bf16 GEMM over a reduction six times as long: the pieces are laid out along K in the orders below.
// layout 0: [rows][6 * cols] (pieces side by side; K-major operand), layout 1: [6 * rows][cols] (pieces stacked;
// MN-major operand, rows = K).
__constant__ int kPieceA[6] = {0, 0, 1, 0, 1, 2};
__constant__ int kPieceB[6] = {0, 1, 0, 2, 1, 0};
__global__ void __launch_bounds__(256)
split3_kernel(const float* __restrict__ x, bf16* __restrict__ out, long long rows, int cols, int stacked, int is_b) {
  const long long nvec = rows * (cols >> 2);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / (cols >> 2);
    const int c = (int)(i - r * (cols >> 2)) * 4;
    const float4 v = *reinterpret_cast<const float4*>(x + r * cols + c);
    const float f[4] = {v.x, v.y, v.z, v.w};
    uint2 piece[3];
    {
      float p1[4], p2[4], p3[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        p1[e] = __bfloat162float(__float2bfloat16(f[e]));
        const float r1 = f[e] - p1[e];
        p2[e] = __bfloat162float(__float2bfloat16(r1));
        p3[e] = r1 - p2[e];
      }
      const float* ps[3] = {p1, p2, p3};
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(ps[k][0], ps[k][1]), hi = __floats2bfloat162_rn(ps[k][2], ps[k][3]);
        piece[k].x = *reinterpret_cast<uint32_t*>(&lo);
        piece[k].y = *reinterpret_cast<uint32_t*>(&hi);
      }
    }
#pragma unroll
    for (int s = 0; s < 6; ++s) {
      const int k = is_b ? kPieceB[s] : kPieceA[s];
      bf16* dst = stacked ? out + ((long long)s * rows + r) * cols + c : out + r * (6ll * cols) + (long long)s * cols + c;
      *reinterpret_cast<uint2*>(dst) = piece[k];
    }
  }
}

// Sum split-K partials (fp32) and finish the epilogue: out = [out +] sum_s ws[s] (+shift) (relu).
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, int splits, long long slab, int ncols,
                                     const float* __restrict__ shift, int relu, int accumulate, void* out,
                                     int out_f32) {
  const long long i4 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= slab) return;
  float4 a = *reinterpret_cast<const float4*>(ws + i4);
  for (int s = 1; s < splits; ++s) {
    const float4 b = *reinterpret_cast<const float4*>(ws + (long long)s * slab + i4);
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  }
  if (shift) {
    const int c = (int)(i4 % ncols);
    a.x += shift[c]; a.y += shift[c + 1]; a.z += shift[c + 2]; a.w += shift[c + 3];
  }
  if (out_f32) {
    float* o = reinterpret_cast<float*>(out) + i4;
    if (accumulate) {
      const float4 b = *reinterpret_cast<const float4*>(o);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    if (relu) { a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); a.z = fmaxf(a.z, 0.f); a.w = fmaxf(a.w, 0.f); }
    *reinterpret_cast<float4*>(o) = a;
  } else {
    bf16* o = reinterpret_cast<bf16*>(out) + i4;
    if (accumulate) {
      const uint2 b = *reinterpret_cast<const uint2*>(o);
      const float2 b0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&b.x));
      const float2 b1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&b.y));
      a.x += b0.x; a.y += b0.y; a.z += b1.x; a.w += b1.y;
    }
    if (relu) { a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); a.z = fmaxf(a.z, 0.f); a.w = fmaxf(a.w, 0.f); }
    uint2 u;
    u.x = pack2(a.x, a.y);
    u.y = pack2(a.z, a.w);
    *reinterpret_cast<uint2*>(o) = u;
  }
}

// ------------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn g_encode_tiled = nullptr;
EncodeIm2colFn g_encode_im2col = nullptr;

int load_driver_entry_points() {
  // cuTensorMapEncode* are driver-API calls and need a CUDA context that is current on THIS host thread.  A thread
  // that has not yet made a runtime call of its own (torch runs backward passes on autograd worker threads) has
  // none bound, and the encode then fails with CUDA_ERROR_INVALID_CONTEXT (seen when a weight-gradient launch was
  // the first CUDA work of a backward thread).  cudaSetDevice binds the primary context; it is a thread-local
  // no-op when already bound.
  {
    int dev = 0;
    RNN_CUDA(cudaGetDevice(&dev));
    RNN_CUDA(cudaSetDevice(dev));
  }
  if (g_encode_tiled && g_encode_im2col) return RNN_OK;
  void* f = nullptr;
  cudaDriverEntryPointQueryResult qr;
  RNN_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr));
  RNN_REQUIRE(f != nullptr && qr == cudaDriverEntryPointSuccess, RNN_ERR_CUDA);
  g_encode_tiled = reinterpret_cast<EncodeTiledFn>(f);
  f = nullptr;
  RNN_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f, cudaEnableDefault, &qr));
  RNN_REQUIRE(f != nullptr && qr == cudaDriverEntryPointSuccess, RNN_ERR_CUDA);
  g_encode_im2col = reinterpret_cast<EncodeIm2colFn>(f);
  return RNN_OK;
}

// Row-major bf16 matrix [rows, cols] (leading dimension ld elements); box = {64 cols, box_rows}.
int make_tmap_2d(CUtensorMap* m, const void* ptr, long long rows, long long cols, long long ld, int box_rows) {
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  RNN_REQUIRE((ld * 2) % 16 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15u) == 0, RNN_ERR_ALIGN);
  CUresult r = g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstr, box,
                              estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char msg[256];
    snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled(2d) -> CUresult %d: ptr %p rows %lld cols %lld ld %lld box_rows %d",
             (int)r, ptr, rows, cols, ld, box_rows);
    rnn_set_last_message(__FILE__, __LINE__, msg);
    return RNN_ERR_TENSORMAP;
  }
  return RNN_OK;
}

// bf16 tensor [slabs][rows][inner] (row pitch ld, slab pitch bstride, elements); box = {64 inner, box_rows, 1}.
int make_tmap_3d(CUtensorMap* m, const void* ptr, long long slabs, long long rows, long long inner, long long ld,
                 long long bstride, int box_rows) {
  cuuint64_t gdim[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)slabs};
  cuuint64_t gstr[2] = {(cuuint64_t)ld * 2, (cuuint64_t)bstride * 2};
  cuuint32_t box[3] = {64u, (cuuint32_t)box_rows, 1u};
  cuuint32_t estr[3] = {1u, 1u, 1u};
  RNN_REQUIRE((ld * 2) % 16 == 0 && (bstride * 2) % 16 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15u) == 0, RNN_ERR_ALIGN);
  CUresult r = g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box,
                              estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RNN_REQUIRE(r == CUDA_SUCCESS, RNN_ERR_TENSORMAP);
  return RNN_OK;
}

// NHWC bf16 activation tensor seen as (C, W, H, N) for im2col-mode TMA.  Strides are explicit so
// that callers may describe virtual (overlapping) views.
int make_tmap_im2col(CUtensorMap* m, const void* ptr, int C, int W, int H, int N, long long str_w,
                     long long str_h, long long str_n, int lo_w, int lo_h, int up_w, int up_h, int ts_w, int ts_h,
                     int pixels) {
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t gstr[3] = {(cuuint64_t)str_w * 2, (cuuint64_t)str_h * 2, (cuuint64_t)str_n * 2};
  int lower[2] = {lo_w, lo_h};
  int upper[2] = {up_w, up_h};
  cuuint32_t estr[4] = {1u, (cuuint32_t)ts_w, (cuuint32_t)ts_h, 1u};
  RNN_REQUIRE((str_w * 2) % 16 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15u) == 0, RNN_ERR_ALIGN);
  RNN_REQUIRE(lo_w >= -128 && lo_w <= 127 && lo_h >= -128 && lo_h <= 127, RNN_ERR_SHAPE);
  RNN_REQUIRE(up_w >= -128 && up_w <= 127 && up_h >= -128 && up_h <= 127, RNN_ERR_SHAPE);
  CUresult r = g_encode_im2col(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), gdim, gstr, lower,
                               upper, 64u, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RNN_REQUIRE(r == CUDA_SUCCESS, RNN_ERR_TENSORMAP);
  return RNN_OK;
}

// Output tensor map for the epilogue's TMA stores: bf16 D[M, ldd] (box 64 x 32) or fp32 partials
// ws[splits][M][ldw] (box 32 x 32 x 1), SWIZZLE_128B.  Scattered-row outputs (strided dgrad) keep the
// register->smem->global store path and get a dummy map.
int make_tmap_out(CUtensorMap* m, TcParams* p) {
  p->tma_store = 0;
  memset(m, 0, sizeof(*m));
  if (p->pix_map) return RNN_OK;
  cuuint32_t estr[3] = {1u, 1u, 1u};
  CUresult r;
  if (p->ws != nullptr) {
    if ((p->ldw * 4) % 16 != 0 || (reinterpret_cast<uintptr_t>(p->ws) & 15u)) return RNN_OK;
    cuuint64_t gdim[3] = {(cuuint64_t)p->ldw, (cuuint64_t)p->M, (cuuint64_t)p->splits};
    cuuint64_t gstr[2] = {(cuuint64_t)p->ldw * 4, (cuuint64_t)p->ldw * 4 * (cuuint64_t)p->M};
    cuuint32_t box[3] = {32u, 32u, 1u};
    r = g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, p->ws, gdim, gstr, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else if (p->bg_heads > 0) {
    if ((p->ldd * 2) % 16 != 0 || (p->d_bstride * 2) % 16 != 0 || (reinterpret_cast<uintptr_t>(p->D) & 15u)) return RNN_OK;
    cuuint64_t gdim[3] = {(cuuint64_t)p->d_inner, (cuuint64_t)p->M, (cuuint64_t)p->d_slabs};
    cuuint64_t gstr[2] = {(cuuint64_t)p->ldd * 2, (cuuint64_t)p->d_bstride * 2};
    cuuint32_t box[3] = {64u, 32u, 1u};
    r = g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, p->D, gdim, gstr, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    if (p->residual != nullptr) return RNN_OK;  // residual path reads rows per thread; keep it simple
    if ((p->ldd * 2) % 16 != 0 || (reinterpret_cast<uintptr_t>(p->D) & 15u)) return RNN_OK;
    cuuint64_t gdim[2] = {(cuuint64_t)p->N, (cuuint64_t)p->M};
    cuuint64_t gstr[1] = {(cuuint64_t)p->ldd * 2};
    cuuint32_t box[2] = {64u, 32u};
    r = g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p->D, gdim, gstr, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  RNN_REQUIRE(r == CUDA_SUCCESS, RNN_ERR_TENSORMAP);
  p->tma_store = 1;
  return RNN_OK;
}

// Whether a launch runs as CTA pairs.  A pair owns a 256-row M tile, so the M tiles must come in twos (or be
// many enough that one idle half tile is noise) and B must split into whole 64-wide halves.  Pairs cut the
// operand bytes each SM pulls through L2 by a third (BN = 256), which is what bounds the long-K tiles
// (measured: 3x3 and C >= 512 layers +5..20 %); tiles with few K steps are bound by their epilogue and lose
// 20-30 % to the cross-CTA hand-offs, so they stay single.  RNN_TC_PAIR=0 turns pairs off, =2 forces them
// wherever the shape allows (A/B timing and tests of both launch forms).
int pick_cg(int BN, int n_mblk, int k_iters) {
  static const int mode = [] {
    const char* e = getenv("RNN_TC_PAIR");
    return e == nullptr ? 1 : atoi(e);
  }();
  if (mode == 0 || BN % 128 != 0) return 1;
  if (!(n_mblk % 2 == 0 || n_mblk >= 16)) return 1;
  return (mode == 2 || k_iters >= 8) ? 2 : 1;
}

// RNN_TC_BRES=0 switches the resident-filter form off, =2 forces it wherever the shape allows, however few tiles a
// CTA walks (A/B timing; tests of both forms on small cases)
int b_resident_mode() {
  static const int v = [] {
    const char* e = getenv("RNN_TC_BRES");
    return e == nullptr ? 1 : atoi(e);
  }();
  return v;
}

// STATS launches: every CTA (pair) keeps one n-block for all of its tiles, so the number of CTAs (pairs) is a
// multiple of n_nblk.  Partial-sum rows: one per CTA.
// RNN_TC_DUAL: two CTAs per SM for the N = 64 convolution tiles (see TcParams::dual).
//   0 never   1 every eligible launch   2 the packed-pixel first layer only   3 (default) every eligible launch except
//   those with the statistics epilogue.
// Timed alone the C64 layers gain 18-21 % (C64 3x3 111 -> 90 us, VGG 224^2 C64 800 -> 650 us, 1x1 C64->64 -11 %).  Inside
// a step (with 128-byte-aligned filters, see optim.py -- before that fix the filter-streaming dual form lost badly):
// mode 3 vs mode 2: VGG-16 13.66 -> 13.25 ms, AlexNet 2.32 -> 2.23 ms, ResNet-50 18.60 -> 18.59 ms.  The statistics
// variant stays on one CTA: with ONE epilogue group per CTA its heavier epilogue does not overlap across tiles
// (ncu: 162-218 us against 135 us); a second group per CTA would need <= 96 registers per thread.
static int dual_mode() {
  static const int v = [] { const char* e = getenv("RNN_TC_DUAL"); return e != nullptr ? atoi(e) : 3; }();
  return v;
}
// the N = 64 convolution tiles on single CTAs with enough tiles to keep 2 x SMs CTAs busy (BN = 128 would get only two
// ring stages per CTA and loses even alone)
static bool want_dual(int BN, int cg, long long tiles, bool first_layer = false, bool stats = false) {
  const int mode = dual_mode();   // 3: every N = 64 convolution except the statistics-epilogue launches
  return (mode == 1 || (mode == 2 && first_layer) || (mode == 3 && !stats)) && cg == 1 && BN == 64 &&
         tiles >= 4ll * rnn_num_sms();
}

void plan_stats_grid(TcParams* p, bool dual) {
  const int cg = p->cg == 2 ? 2 : 1;
  const int units = rnn_num_sms() / cg * (dual ? 2 : 1);
  p->force_grid = (units / p->n_nblk) * p->n_nblk * cg;
  p->stat_rows = p->force_grid / p->n_nblk;
}

template <int BN, int KIND, bool A_MN, bool B_MN, bool FUSED, bool STATS = false, int CG = 1>
int launch_inst2(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p_in, cudaStream_t stream) {
  TcParams p = p_in;
  CUtensorMap td;
  {
    const int rc = make_tmap_out(&td, &p);
    if (rc != RNN_OK) return rc;
  }
  static bool attr_set[64] = {false};   // the attribute is per device
  auto kern = tc_kernel<BN, KIND, A_MN, B_MN, FUSED, STATS, CG>;
  constexpr int smem = Cfg<BN, CG, KIND>::SMEM_BYTES;
  int dev = 0;
  RNN_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    RNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    RNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  const int units = rnn_num_sms() / CG;   // CTAs, or CTA pairs
  const int tiles = ceil_div_i(p.n_mblk, CG) * p.n_nblk * p.splits;
  int grid = (tiles < units ? tiles : units) * CG;
  if (p.force_grid > 0) grid = p.force_grid;
  // Resident filter (TcParams::b_resident): convolutions whose whole B slice fits beside an A-only ring of at least
  // four stages, on single CTAs, when every CTA walks several tiles (otherwise there is nothing to re-use).  These
  // are the narrow layers that run at the L2 -> SM throughput cap: a 128 x 64 tile of the C64 3x3 layers pulled
  // 144 KB of A and 72 KB of filter through L2; the filter share disappears.
  p.b_resident = 0;
  if constexpr (KIND == KIND_CONV && CG == 1) {
    using C = Cfg<BN, CG, KIND>;
    const int need = p.k_iters * C::B_BYTES;
    const int ring = (C::STAGES * C::STAGE_BYTES - need) / A_BYTES;
    const int g2 = p.force_grid > 0 ? grid : (units / p.n_nblk) * p.n_nblk;
    if (b_resident_mode() != 0 && !p.no_b_resident && p.splits == 1 && need > 0 && ring >= 4 && g2 > 0 && g2 % p.n_nblk == 0 &&
        (tiles >= 3 * g2 || b_resident_mode() == 2)) {
      p.b_resident = 1;
      p.ring_stages = ring > 8 ? 8 : ring;
      grid = g2;
    }
  }
  // weight-gradient kernels run on a side stream next to other kernels: one epilogue group keeps them small
  int threads = (KIND == KIND_WGRAD && !STATS) ? NUM_THREADS_NARROW : NUM_THREADS;
  int smem_launch = smem;
  p.dual = 0;
  if constexpr (KIND == KIND_CONV && CG == 1 && BN == 64) {
    using C = Cfg<BN, CG, KIND>;
    // per CTA: ring + staging + barriers must fit twice into the SM's 227 KB (1 KB per block is the driver's)
    constexpr int kPerCta = (232448 - 2 * 1024) / 2;
    constexpr int fixed = C::STAGING_BYTES + 1024 + 256;
    constexpr int nst = (kPerCta - fixed) / C::STAGE_BYTES;
    static_assert(nst >= 3, "two CTAs per SM need a ring of at least three stages each");
    // statistics launches planned their (doubled) grid with the same predicate (conv_fprop_impl)
    if (want_dual(BN, CG, tiles, p.no_b_resident != 0, STATS) && (p.force_grid == 0 || STATS)) {
      p.dual = 1;
      p.b_resident = 0;
      p.nst = nst > 8 ? 8 : nst;
      p.ring_area = p.nst * C::STAGE_BYTES;
      smem_launch = p.ring_area + fixed;
      threads = NUM_THREADS_NARROW;
      if (p.force_grid == 0) grid = tiles < 2 * units ? tiles : 2 * units;
    }
  }
  RnnTimedScope timed("tc_kernel", stream);
  if (CG == 1) {
    kern<<<grid, threads, smem_launch, stream>>>(ta, tb, td, p);
  } else {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = CG; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    RNN_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, td, p));
  }
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

template <int BN, int KIND, bool A_MN, bool B_MN, int CG = 1>
int launch_inst(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t stream) {
  if constexpr (KIND == KIND_CONV && !A_MN && !B_MN && BN != 192) {
    if (p.stat_partials != nullptr) return launch_inst2<BN, KIND, A_MN, B_MN, false, true, CG>(ta, tb, p, stream);
  }
  // wgrad and the MN-major-A GEMM (Linear wgrad) never carry a fused epilogue
  if constexpr (KIND != KIND_WGRAD && !A_MN) {
    if (p.ws == nullptr && (p.scale || p.shift || p.residual || p.relu || p.mask_in || p.mask_out))
      return launch_inst2<BN, KIND, A_MN, B_MN, true, false, CG>(ta, tb, p, stream);
  }
  return launch_inst2<BN, KIND, A_MN, B_MN, false, false, CG>(ta, tb, p, stream);
}

template <int KIND, bool A_MN, bool B_MN>
int launch_bn(int BN, const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t s) {
  if (p.cg == 2) {
    switch (BN) {
      case 128: return launch_inst<128, KIND, A_MN, B_MN, 2>(ta, tb, p, s);
      case 256: return launch_inst<256, KIND, A_MN, B_MN, 2>(ta, tb, p, s);
    }
    return RNN_ERR_SHAPE;
  }
  switch (BN) {
    case 64: return launch_inst<64, KIND, A_MN, B_MN>(ta, tb, p, s);
    case 128: return launch_inst<128, KIND, A_MN, B_MN>(ta, tb, p, s);
    case 192: return launch_inst<192, KIND, A_MN, B_MN>(ta, tb, p, s);
    case 256: return launch_inst<256, KIND, A_MN, B_MN>(ta, tb, p, s);
  }
  return RNN_ERR_SHAPE;
}

// batched GEMMs always launch single CTAs (their tiles have 1-8 K steps)
template <bool A_MN, bool B_MN>
int launch_bgemm(int BN, const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t s) {
  switch (BN) {
    case 64: return launch_inst2<64, KIND_BGEMM, A_MN, B_MN, false>(ta, tb, p, s);
    case 128: return launch_inst2<128, KIND_BGEMM, A_MN, B_MN, false>(ta, tb, p, s);
    case 256: return launch_inst2<256, KIND_BGEMM, A_MN, B_MN, false>(ta, tb, p, s);
  }
  return RNN_ERR_SHAPE;
}

int launch_reduce(const float* ws, int splits, long long slab, int ncols, const float* shift, int relu,
                  int accumulate, void* out, int out_f32, cudaStream_t stream) {
  RNN_REQUIRE(slab % 4 == 0 && ncols % 4 == 0, RNN_ERR_SHAPE);
  const long long n4 = slab / 4;
  const int threads = 256;
  const int blocks = ceil_div_i(n4, threads);
  RnnTimedScope timed("splitk_reduce", stream);
  splitk_reduce_kernel<<<blocks, threads, 0, stream>>>(ws, splits, slab, ncols, shift, relu, accumulate, out,
                                                       out_f32);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

int pick_bn(int n) {
  if (n % 256 == 0) return 256;
  if (n % 128 == 0) return 128;
  if (n > 128 && n % 64 != 0) return 128;  // ragged: fewer, wider tiles
  return 64;
}

// Split-K policy for plain GEMMs: split only when the tile grid leaves most SMs idle.
int pick_splits(int base, int k_iters);

int gemm_splits(int M, int N, int K, int BN) {
  const int tiles = ceil_div_i(M, BM) * ceil_div_i(N, BN);
  const int k_iters = ceil_div_i(K, BK);
  if (tiles * 2 > rnn_num_sms() || k_iters < 8) return 1;
  return pick_splits(tiles, k_iters);
}

struct WgradPlan {
  int BN, n_mblk, ncblk, k_iters, splits;
  size_t ws_bytes;
};

// Split factor for a reduction of `k_iters` steps over `base` output tiles on a persistent grid of one CTA per
// SM: maximise wave efficiency tiles / (ceil(tiles / SMs) * SMs) -- a 149th tile costs a whole extra wave --
// keep at least 4 K-steps per split, and among near-equal choices take the smallest (fewer partials).
int pick_splits(int base, int k_iters) {
  const int sms = rnn_num_sms();
  int max_s = k_iters / 4;
  if (max_s > 64) max_s = 64;
  if (max_s < 1) max_s = 1;
  int best = 1;
  double best_eff = 0.0;
  for (int s = 1; s <= max_s; ++s) {
    const long long tiles = (long long)base * s;
    const long long waves = (tiles + sms - 1) / sms;
    const double eff = (double)tiles / (double)(waves * sms);
    if (eff > best_eff + 0.03) { best_eff = eff; best = s; }
    if (waves > 4) break;  // enough parallelism already; more splits only add partial traffic
  }
  return best;
}

// Tile width (in columns of the [K, taps*C] gradient matrix) for `nsb` 64-wide sub-blocks.
int wgrad_bn(int nsb) {
  if (const char* e = getenv("RNN_WGRAD_BN")) {  // tuning experiments only
    const int v = atoi(e);
    if (v == 64 || v == 128 || v == 192 || v == 256) return v;
  }
  if (nsb % 4 == 0) return 256;
  if (nsb % 3 == 0) return 192;
  if (nsb % 2 == 0) return nsb > 8 ? 256 : 128;
  return nsb >= 5 ? 256 : 64;
}

WgradPlan plan_wgrad_cols(int K, int ntaps, int C, long long mpix) {
  WgradPlan w;
  const int nsb = ntaps * (C / 64);
  w.BN = wgrad_bn(nsb);
  w.n_mblk = ceil_div_i(K, BM);
  w.ncblk = ceil_div_i(nsb, w.BN / 64);  // number of N tiles
  w.k_iters = ceil_div_i(mpix, BK);
  const int base = w.n_mblk * w.ncblk;
  const int s = pick_splits(base, w.k_iters);
  w.splits = s;
  w.ws_bytes = (size_t)s * K * ntaps * C * sizeof(float);
  return w;
}

WgradPlan plan_wgrad(const rnn_conv_desc* d) {
  return plan_wgrad_cols(d->K, d->R * d->S, d->C, (long long)d->N * d->P * d->Q);
}

}  // namespace

// Tensor-map builders for the other translation units (the fused attention kernels in attention.cu): a bf16 tensor
// [slabs][rows][inner] with a {64 inner, box_rows, 1} box in SWIZZLE_128B, for TMA loads and stores alike.
int rnn_internal_tmap_3d(CUtensorMap* m, const void* ptr, long long slabs, long long rows, long long inner, long long ld,
                         long long bstride, int box_rows) {
  int rc = load_driver_entry_points();
  if (rc != RNN_OK) return rc;
  return make_tmap_3d(m, ptr, slabs, rows, inner, ld, bstride, box_rows);
}

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

// Batched D_z = A_z . B_z^T for z = b * heads + h (see rnn_bgemm_desc in the header): the attention contractions
// of all heads of a multi-head module in one launch, operands addressed in place inside the packed projections.
int rnn_gemm_batched_bf16(const rnn_bgemm_desc* d, const void* A, const void* B, void* D, cudaStream_t stream) {
  RNN_REQUIRE(d != nullptr && d->M > 0 && d->N > 0 && d->K > 0 && d->batches > 0 && d->heads > 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(!(d->a_mn && !d->b_mn), RNN_ERR_SHAPE);  // (MN,K) combination is not instantiated
  // a head's slice of a shared row must be whole 64-wide blocks, or a tile would read / write its neighbour
  if (d->a_per_head) RNN_REQUIRE(d->a_hoff % 64 == 0 && (d->a_mn ? d->M : d->K) % 64 == 0, RNN_ERR_SHAPE);
  if (d->b_per_head) RNN_REQUIRE(d->b_hoff % 64 == 0 && (d->b_mn ? d->N : d->K) % 64 == 0, RNN_ERR_SHAPE);
  if (d->d_per_head) RNN_REQUIRE(d->d_hoff % 64 == 0 && d->N % 64 == 0, RNN_ERR_SHAPE);
  int rc = load_driver_entry_points();
  if (rc != RNN_OK) return rc;
  const long long nz = (long long)d->batches * d->heads;
  RNN_REQUIRE(nz < (1 << 24), RNN_ERR_SHAPE);
  const int BN = pick_bn(d->N);
  CUtensorMap ta, tb;
  {
    const long long slabs = d->a_per_head ? d->batches : nz;
    const long long rows = d->a_mn ? d->K : d->M;
    const long long inner = d->a_per_head ? (long long)d->heads * d->a_hoff : (d->a_mn ? d->M : d->K);
    rc = make_tmap_3d(&ta, A, slabs, rows, inner, d->a_ld, d->a_bstride, d->a_mn ? 64 : BM);
    if (rc != RNN_OK) return rc;
  }
  {
    const long long slabs = d->b_per_head ? d->batches : nz;
    const long long rows = d->b_mn ? d->K : d->N;
    const long long inner = d->b_per_head ? (long long)d->heads * d->b_hoff : (d->b_mn ? d->N : d->K);
    rc = make_tmap_3d(&tb, B, slabs, rows, inner, d->b_ld, d->b_bstride, d->b_mn ? 64 : BN);
    if (rc != RNN_OK) return rc;
  }
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = d->M; p.N = d->N;
  p.k_iters = ceil_div_i(d->K, BK);
  p.splits = (int)nz;
  p.n_mblk = ceil_div_i(d->M, BM);
  p.n_nblk = ceil_div_i(d->N, BN);
  p.cg = 1;
  p.bg_heads = d->heads;
  p.a_zdiv = d->a_per_head; p.b_zdiv = d->b_per_head; p.d_zdiv = d->d_per_head;
  p.a_hoff = d->a_per_head ? d->a_hoff : 0; p.b_hoff = d->b_per_head ? d->b_hoff : 0;
  p.d_hoff = d->d_per_head ? d->d_hoff : 0;
  if (d->out_f32) {
    // fp32 D is always one dense slab per problem: [z][M][N]
    RNN_REQUIRE(!d->d_per_head && d->d_ld == d->N && d->d_bstride == (long long)d->M * d->N, RNN_ERR_SHAPE);
    RNN_REQUIRE(d->N % 4 == 0 && (reinterpret_cast<uintptr_t>(D) & 15u) == 0, RNN_ERR_ALIGN);
    p.ws = reinterpret_cast<float*>(D);
    p.ldw = d->N;
  } else {
    p.D = D;
    p.ldd = d->d_ld;
    p.d_bstride = d->d_bstride;
    p.d_slabs = (int)(d->d_per_head ? d->batches : nz);
    p.d_inner = (int)(d->d_per_head ? (long long)d->heads * d->d_hoff : d->N);
    RNN_REQUIRE((d->d_ld * 2) % 16 == 0 && (d->d_bstride * 2) % 16 == 0, RNN_ERR_SHAPE);
    RNN_REQUIRE((reinterpret_cast<uintptr_t>(D) & 15u) == 0, RNN_ERR_ALIGN);
  }
  if (!d->a_mn && !d->b_mn) return launch_bgemm<false, false>(BN, ta, tb, p, stream);
  if (!d->a_mn && d->b_mn) return launch_bgemm<false, true>(BN, ta, tb, p, stream);
  return launch_bgemm<true, true>(BN, ta, tb, p, stream);
}

size_t rnn_gemm_bf16_workspace(int M, int N, int K, int out_f32, int has_epilogue) {
  const int BN = pick_bn(N);
  const int s = gemm_splits(M, N, K, BN);
  if (s == 1 && !(out_f32 && has_epilogue)) return 0;
  return (size_t)s * M * N * sizeof(float);
}

int rnn_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int a_mn, int b_mn,
                  const float* bias, int relu, int out_f32, void* ws, size_t ws_bytes, cudaStream_t stream) {
  RNN_REQUIRE(M > 0 && N > 0 && K > 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(N % 8 == 0, RNN_ERR_SHAPE);                       // 16-byte stores of D rows
  RNN_REQUIRE(a_mn ? (M % 8 == 0) : (K % 8 == 0), RNN_ERR_SHAPE);  // 16-byte row pitch of A (TMA)
  RNN_REQUIRE(b_mn ? true : (K % 8 == 0), RNN_ERR_SHAPE);          // 16-byte row pitch of B (TMA)
  RNN_REQUIRE(!(a_mn && !b_mn), RNN_ERR_SHAPE);  // (MN,K) combination is not instantiated
  int rc = load_driver_entry_points();
  if (rc != RNN_OK) return rc;

  const int BN = pick_bn(N);
  const int cg = gemm_splits(M, N, K, BN) > 1 ? 1 : pick_cg(BN, ceil_div_i(M, BM), ceil_div_i(K, BK));
  CUtensorMap ta, tb;
  if (!a_mn) rc = make_tmap_2d(&ta, A, M, K, K, BM);   // A[M,K], box {64 k, 128 m}
  else       rc = make_tmap_2d(&ta, A, K, M, M, 64);   // A stored [K,M], box {64 m, 64 k}
  if (rc != RNN_OK) return rc;
  if (!b_mn) rc = make_tmap_2d(&tb, B, N, K, K, BN / cg);   // B[N,K], box {64 k, BN n} (half per CTA of a pair)
  else       rc = make_tmap_2d(&tb, B, K, N, N, 64);   // B stored [K,N], box {64 n, 64 k}
  if (rc != RNN_OK) return rc;

  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.N = N;
  p.cg = cg;
  p.k_iters = ceil_div_i(K, BK);
  p.splits = gemm_splits(M, N, K, BN);
  p.n_mblk = ceil_div_i(M, BM);
  p.n_nblk = ceil_div_i(N, BN);
  p.ldd = N;
  p.ldw = N;
  const bool has_epi = (bias != nullptr) || relu;
  const bool via_ws = p.splits > 1 || (out_f32 && has_epi);
  if (via_ws) {
    const size_t need = (size_t)p.splits * M * N * sizeof(float);
    RNN_REQUIRE(ws != nullptr && ws_bytes >= need, RNN_ERR_WORKSPACE);
    p.ws = reinterpret_cast<float*>(ws);
  } else if (out_f32) {
    p.ws = reinterpret_cast<float*>(D);
  } else {
    p.D = D;
    p.shift = bias;
    p.relu = relu;
  }
  if (!a_mn && !b_mn) rc = launch_bn<KIND_GEMM, false, false>(BN, ta, tb, p, stream);
  else if (!a_mn && b_mn) rc = launch_bn<KIND_GEMM, false, true>(BN, ta, tb, p, stream);
  else rc = launch_bn<KIND_GEMM, true, true>(BN, ta, tb, p, stream);
  if (rc != RNN_OK) return rc;
  if (via_ws) {
    rc = launch_reduce(p.ws, p.splits, (long long)M * N, N, bias, relu, 0, D, out_f32, stream);
  }
  return rc;
}

// fp32 in / fp32 out GEMM at fp32 accuracy on the tensor cores (three-piece bf16 split, see split3_kernel).
static bool f32_tc_shape_ok(int M, int N, int K, int a_mn, int b_mn) {
  if (M <= 0 || N <= 0 || K <= 0 || (a_mn && !b_mn)) return false;
  if (N % 8 != 0 || K % 4 != 0) return false;       // 16-byte rows of D and of the split operands
  if (a_mn && M % 8 != 0) return false;
  return (long long)6 * K < (1ll << 30);
}
int rnn_gemm_f32_tc_supported(int M, int N, int K, int a_mn, int b_mn) { return f32_tc_shape_ok(M, N, K, a_mn, b_mn); }
size_t rnn_gemm_f32_tc_workspace(int M, int N, int K, int has_epilogue) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  const size_t a = ((size_t)6 * M * K * sizeof(bf16) + 255) & ~(size_t)255;
  const size_t b = ((size_t)6 * N * K * sizeof(bf16) + 255) & ~(size_t)255;
  return a + b + rnn_gemm_bf16_workspace(M, N, 6 * K, 1, has_epilogue);
}
int rnn_gemm_f32_tc(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn,
                    const float* bias, int relu, void* ws, size_t ws_bytes, cudaStream_t stream) {
  RNN_REQUIRE(f32_tc_shape_ok(M, N, K, a_mn, b_mn), RNN_ERR_SHAPE);
  const int has_epi = (bias != nullptr) || relu;
  RNN_REQUIRE(ws != nullptr && ws_bytes >= rnn_gemm_f32_tc_workspace(M, N, K, has_epi), RNN_ERR_WORKSPACE);
  RNN_REQUIRE(((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15u) == 0, RNN_ERR_ALIGN);
  const size_t a_bytes = ((size_t)6 * M * K * sizeof(bf16) + 255) & ~(size_t)255;
  const size_t b_bytes = ((size_t)6 * N * K * sizeof(bf16) + 255) & ~(size_t)255;
  bf16* A3 = reinterpret_cast<bf16*>(ws);
  bf16* B3 = reinterpret_cast<bf16*>(reinterpret_cast<unsigned char*>(ws) + a_bytes);
  void* gws = reinterpret_cast<unsigned char*>(ws) + a_bytes + b_bytes;
  // A is [M][K] (K-major: pieces side by side) or [K][M] (MN-major: pieces stacked along K); same for B
  const long long a_rows = a_mn ? K : M, b_rows = b_mn ? K : N;
  const int a_cols = a_mn ? M : K, b_cols = b_mn ? N : K;
  RNN_REQUIRE(a_cols % 4 == 0 && b_cols % 4 == 0, RNN_ERR_SHAPE);
  const int ga = ceil_div_i(a_rows * (a_cols / 4), 256), gb = ceil_div_i(b_rows * (b_cols / 4), 256);
  const int cap = rnn_num_sms() * 16;
  split3_kernel<<<ga < cap ? ga : cap, 256, 0, stream>>>(A, A3, a_rows, a_cols, a_mn, 0);
  RNN_CHECK_LAUNCH();
  split3_kernel<<<gb < cap ? gb : cap, 256, 0, stream>>>(B, B3, b_rows, b_cols, b_mn, 1);
  RNN_CHECK_LAUNCH();
  return rnn_gemm_bf16(A3, B3, D, M, N, 6 * K, a_mn, b_mn, bias, relu, 1, gws,
                       ws_bytes - a_bytes - b_bytes, stream);
}

static int check_conv_desc(const rnn_conv_desc* d) {
  RNN_REQUIRE(d != nullptr, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->N > 0 && d->H > 0 && d->W > 0 && d->C > 0 && d->K > 0 && d->R > 0 && d->S > 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->stride_h > 0 && d->stride_w > 0 && d->pad_h >= 0 && d->pad_w >= 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->P == (d->H + 2 * d->pad_h - d->R) / d->stride_h + 1, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->Q == (d->W + 2 * d->pad_w - d->S) / d->stride_w + 1, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->C % 64 == 0, RNN_ERR_SHAPE);  // one 128-byte swizzle row of channels per K step
  RNN_REQUIRE(d->K % 8 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->R * d->S <= MAX_TAPS, RNN_ERR_SHAPE);
  return RNN_OK;
}

static int conv_fprop_impl(const rnn_conv_desc* d, const void* x, const void* w, void* y, const float* scale,
                           const float* shift, const void* residual, int relu, float* stat_partials, int* stat_rows,
                           cudaStream_t stream, unsigned char* mask_out = nullptr) {
  int rc = check_conv_desc(d);
  if (rc != RNN_OK) return rc;
  rc = load_driver_entry_points();
  if (rc != RNN_OK) return rc;
  const int BN = pick_bn(d->K);
  CUtensorMap ta, tb;
  rc = make_tmap_im2col(&ta, x, d->C, d->W, d->H, d->N, d->C, (long long)d->W * d->C,
                        (long long)d->H * d->W * d->C, -d->pad_w, -d->pad_h, d->pad_w - (d->S - 1),
                        d->pad_h - (d->R - 1), d->stride_w, d->stride_h, BM);
  if (rc != RNN_OK) return rc;
  const long long rsc = (long long)d->R * d->S * d->C;
  const long long M = (long long)d->N * d->P * d->Q;
  RNN_REQUIRE(M < (1ll << 31), RNN_ERR_SHAPE);
  const int cg = pick_cg(BN, ceil_div_i(M, BM), d->R * d->S * (d->C / 64));
  rc = make_tmap_2d(&tb, w, d->K, rsc, rsc, BN / cg);
  if (rc != RNN_OK) return rc;

  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = (int)M; p.N = d->K;
  p.cg = cg;
  p.cchunks = d->C / 64;
  p.ntaps = d->R * d->S;
  p.k_iters = p.ntaps * p.cchunks;
  p.splits = 1;
  p.n_mblk = ceil_div_i(M, BM);
  p.n_nblk = ceil_div_i(d->K, BN);
  p.Wo = d->Q; p.HoWo = d->P * d->Q;
  p.ts_w = d->stride_w; p.ts_h = d->stride_h; p.lo_w = -d->pad_w; p.lo_h = -d->pad_h;
  p.ldd = d->K;
  p.D = y; p.scale = scale; p.shift = shift; p.residual = reinterpret_cast<const bf16*>(residual); p.relu = relu;
  p.mask_out = mask_out;
  if (stat_partials != nullptr) {
    RNN_REQUIRE(d->K % BN == 0, RNN_ERR_SHAPE);  // statistics need whole N tiles
    RNN_REQUIRE(p.n_nblk <= rnn_num_sms() / 2, RNN_ERR_SHAPE);
    plan_stats_grid(&p, want_dual(BN, cg, (long long)p.n_mblk * p.n_nblk, false, true));
    p.stat_partials = stat_partials;
    *stat_rows = p.stat_rows;
  }
  for (int r = 0; r < d->R; ++r)
    for (int s = 0; s < d->S; ++s) {
      TapDesc& t = p.taps[r * d->S + s];
      t.ow = (int16_t)s; t.oh = (int16_t)r; t.bcol = (r * d->S + s) * d->C;
    }
  return launch_bn<KIND_CONV, false, false>(BN, ta, tb, p, stream);
}

// y = relu(conv(x, w) + shift) and the 1-bit mask (y > 0) the backward pass needs: [N,P,Q,K/8] bytes, bit k&7 of byte
// k>>3.  The consumer's dgrad applies it (rnn_conv2d_dgrad_masked_bf16) in place of a ReLU backward pass over y.
int rnn_conv2d_fprop_relu_mask_bf16(const rnn_conv_desc* d, const void* x, const void* w, void* y, const float* shift,
                                    unsigned char* mask_out, cudaStream_t stream) {
  RNN_REQUIRE(d != nullptr && mask_out != nullptr && d->K % 64 == 0, RNN_ERR_SHAPE);
  return conv_fprop_impl(d, x, w, y, nullptr, shift, nullptr, 1, nullptr, nullptr, stream, mask_out);
}

// y[N,P,Q,K] = conv(x[N,H,W,C], w[K,R,S,C]); optional per-channel scale/shift, residual add, ReLU.
int rnn_conv2d_fprop_bf16(const rnn_conv_desc* d, const void* x, const void* w, void* y, const float* scale,
                          const float* shift, const void* residual, int relu, cudaStream_t stream) {
  return conv_fprop_impl(d, x, w, y, scale, shift, residual, relu, nullptr, nullptr, stream);
}

// Convolution + BatchNorm batch statistics of its output in one pass: the epilogue accumulates per-channel sums
// of the bf16 values it stores, a finalize kernel turns them into mean / invstd / scale / shift and updates the
// running statistics (torch.nn.BatchNorm2d training semantics).  ws: rnn_conv2d_fprop_bn_stats_workspace() bytes.
size_t rnn_conv2d_fprop_bn_stats_workspace(const rnn_conv_desc* d) {
  if (d == nullptr || d->K <= 0) return 0;
  const int BN = pick_bn(d->K);
  const int n_nblk = ceil_div_i(d->K, BN);
  const int sms = rnn_num_sms();
  if (n_nblk > sms) return 0;
  const int grid = (sms * (BN == 64 ? 2 : 1) / n_nblk) * n_nblk;   // N = 64 tiles may run two CTAs per SM
  return (size_t)n_nblk * (grid / n_nblk) * 2 * BN * sizeof(float);
}

int rnn_conv2d_fprop_bn_stats_bf16(const rnn_conv_desc* d, const void* x, const void* w, void* y, const float* gamma,
                                   const float* beta, float eps, float momentum, float* running_mean,
                                   float* running_var, float* mean, float* invstd, float* scale, float* shift,
                                   void* ws, size_t ws_bytes, cudaStream_t stream) {
  RNN_REQUIRE(d != nullptr, RNN_ERR_SHAPE);
  const size_t need = rnn_conv2d_fprop_bn_stats_workspace(d);
  RNN_REQUIRE(need > 0 && ws != nullptr && ws_bytes >= need, RNN_ERR_WORKSPACE);
  int rows = 0;
  int rc = conv_fprop_impl(d, x, w, y, nullptr, nullptr, nullptr, 0, reinterpret_cast<float*>(ws), &rows, stream);
  if (rc != RNN_OK) return rc;
  const long long M = (long long)d->N * d->P * d->Q;
  return rnn_internal_bn_finalize(reinterpret_cast<const float*>(ws), rows, d->K, pick_bn(d->K), (float)M, gamma, beta,
                                  eps, momentum, running_mean, running_var, mean, invstd, scale, shift, stream);
}

// dx[N,H,W,C] = conv_transpose(dy[N,P,Q,K], w[K,R,S,C]).  Strided convolutions are decomposed
// into stride_h*stride_w parity classes, each a unit-stride gather over dy with a subset of taps.
// `accumulate` != 0 adds into the existing dx (gradient of a tensor consumed by two branches).
static int conv_dgrad_impl(const rnn_conv_desc* d, const void* dy, const void* w, void* dx, int accumulate,
                           const void* relu_mask, cudaStream_t stream);
int rnn_conv2d_dgrad_bf16(const rnn_conv_desc* d, const void* dy, const void* w, void* dx, int accumulate,
                          cudaStream_t stream) {
  return conv_dgrad_impl(d, dy, w, dx, accumulate, nullptr, stream);
}
// dx = conv_transpose(dy, w) masked by the 1-bit ReLU mask of the tensor x this convolution consumed ([N,H,W,C/8]
// bytes, from rnn_conv2d_fprop_relu_mask_bf16 of the producing layer): the producer's ReLU backward in this epilogue
// instead of a separate pass that reads dx and x and writes dx.  Unit stride, C % 64 == 0.
int rnn_conv2d_dgrad_masked_bf16(const rnn_conv_desc* d, const void* dy, const void* w, const unsigned char* mask,
                                 void* dx, cudaStream_t stream) {
  RNN_REQUIRE(d != nullptr && mask != nullptr && d->stride_h == 1 && d->stride_w == 1 && d->C % 64 == 0, RNN_ERR_SHAPE);
  return conv_dgrad_impl(d, dy, w, dx, 0, mask, stream);
}
static int conv_dgrad_impl(const rnn_conv_desc* d, const void* dy, const void* w, void* dx, int accumulate,
                           const void* relu_mask, cudaStream_t stream) {
  int rc = check_conv_desc(d);
  if (rc != RNN_OK) return rc;
  RNN_REQUIRE(d->K % 64 == 0 && d->C % 8 == 0, RNN_ERR_SHAPE);
  rc = load_driver_entry_points();
  if (rc != RNN_OK) return rc;
  const int BN = pick_bn(d->C);
  const long long rsc = (long long)d->R * d->S * d->C;
  CUtensorMap tb;
  rc = make_tmap_2d(&tb, w, d->K, rsc, rsc, 64);  // MN-major B: box {64 c, 64 k}
  if (rc != RNN_OK) return rc;

  // Classes with no contributing tap leave their dx pixels untouched: zero them first.
  const bool has_empty = (d->R < d->stride_h) || (d->S < d->stride_w);
  if (has_empty && !accumulate) {
    RNN_CUDA(cudaMemsetAsync(dx, 0, (size_t)d->N * d->H * d->W * d->C * 2, stream));
  }

  for (int a = 0; a < d->stride_h; ++a) {
    for (int b = 0; b < d->stride_w; ++b) {
      const int Hc = (d->H - a + d->stride_h - 1) / d->stride_h;
      const int Wc = (d->W - b + d->stride_w - 1) / d->stride_w;
      if (Hc <= 0 || Wc <= 0) continue;
      // taps: r with (a + pad - r) % stride == 0; dy row = i + (a + pad - r) / stride
      int offh[MAX_TAPS], rr[MAX_TAPS], nh = 0, offw[MAX_TAPS], ss[MAX_TAPS], nw = 0;
      for (int r = 0; r < d->R; ++r) {
        const int t = a + d->pad_h - r;
        if (((t % d->stride_h) + d->stride_h) % d->stride_h != 0) continue;
        offh[nh] = (t >= 0) ? t / d->stride_h : -((-t) / d->stride_h);
        rr[nh++] = r;
      }
      for (int s = 0; s < d->S; ++s) {
        const int t = b + d->pad_w - s;
        if (((t % d->stride_w) + d->stride_w) % d->stride_w != 0) continue;
        offw[nw] = (t >= 0) ? t / d->stride_w : -((-t) / d->stride_w);
        ss[nw++] = s;
      }
      if (nh == 0 || nw == 0) continue;
      int lo_h = offh[0], lo_w = offw[0];
      for (int i = 1; i < nh; ++i) lo_h = offh[i] < lo_h ? offh[i] : lo_h;
      for (int i = 1; i < nw; ++i) lo_w = offw[i] < lo_w ? offw[i] : lo_w;
      const int up_h = Hc - d->P + lo_h, up_w = Wc - d->Q + lo_w;

      CUtensorMap ta;
      rc = make_tmap_im2col(&ta, dy, d->K, d->Q, d->P, d->N, d->K, (long long)d->Q * d->K,
                            (long long)d->P * d->Q * d->K, lo_w, lo_h, up_w, up_h, 1, 1, BM);
      if (rc != RNN_OK) return rc;

      TcParams p;
      memset(&p, 0, sizeof(p));
      const long long M = (long long)d->N * Hc * Wc;
      p.M = (int)M; p.N = d->C;
      p.cchunks = d->K / 64;
      p.ntaps = nh * nw;
      p.k_iters = p.ntaps * p.cchunks;
      p.splits = 1;
      p.n_mblk = ceil_div_i(M, BM);
      p.n_nblk = ceil_div_i(d->C, BN);
      p.cg = pick_cg(BN, p.n_mblk, p.k_iters);
      p.Wo = Wc; p.HoWo = Hc * Wc;
      p.ts_w = 1; p.ts_h = 1; p.lo_w = lo_w; p.lo_h = lo_h;
      p.ldd = d->C;
      p.pix_map = (d->stride_h != 1 || d->stride_w != 1) ? 1 : 0;
      p.OW = d->W; p.OHOW = d->H * d->W; p.os_h = d->stride_h; p.os_w = d->stride_w; p.oa = a; p.ob = b;
      p.D = dx;
      // accumulation: dense outputs add through the TMA reduce-store (bf16 add at the destination, the same
      // rounding as adding two bf16 gradient tensors); scattered (strided) outputs read-modify-write per thread
      if (relu_mask != nullptr) p.mask_in = reinterpret_cast<const unsigned char*>(relu_mask);
      else if (accumulate && !p.pix_map) p.reduce_add = 1;
      else p.residual = accumulate ? reinterpret_cast<const bf16*>(dx) : nullptr;
      for (int i = 0; i < nh; ++i)
        for (int j = 0; j < nw; ++j) {
          TapDesc& t = p.taps[i * nw + j];
          t.oh = (int16_t)(offh[i] - lo_h); t.ow = (int16_t)(offw[j] - lo_w);
          t.bcol = (rr[i] * d->S + ss[j]) * d->C;
        }
      rc = launch_bn<KIND_CONV, false, true>(BN, ta, tb, p, stream);
      if (rc != RNN_OK) return rc;
    }
  }
  return RNN_OK;
}

size_t rnn_conv2d_wgrad_workspace(const rnn_conv_desc* d) {
  if (check_conv_desc(d) != RNN_OK) return 0;
  return plan_wgrad(d).ws_bytes;
}

// dw[K,R,S,C] = sum over pixels of dy[.,k] * x_im2col[., (r,s,c)].  Split over the pixel range;
// fp32 partials live in `ws` and are summed in a fixed order (bitwise run-to-run reproducible).
int rnn_conv2d_wgrad_bf16(const rnn_conv_desc* d, const void* x, const void* dy, void* dw, int dw_f32,
                          int accumulate, void* ws, size_t ws_bytes, cudaStream_t stream) {
  int rc = check_conv_desc(d);
  if (rc != RNN_OK) return rc;
  rc = load_driver_entry_points();
  if (rc != RNN_OK) return rc;
  const WgradPlan w = plan_wgrad(d);
  RNN_REQUIRE(ws != nullptr && ws_bytes >= w.ws_bytes, RNN_ERR_WORKSPACE);
  const long long mpix = (long long)d->N * d->P * d->Q;
  RNN_REQUIRE(mpix < (1ll << 31), RNN_ERR_SHAPE);
  CUtensorMap ta, tb;
  rc = make_tmap_2d(&ta, dy, mpix, d->K, d->K, 64);  // MN-major A: box {64 k_out, 64 pixels}
  if (rc != RNN_OK) return rc;
  rc = make_tmap_im2col(&tb, x, d->C, d->W, d->H, d->N, d->C, (long long)d->W * d->C,
                        (long long)d->H * d->W * d->C, -d->pad_w, -d->pad_h, d->pad_w - (d->S - 1),
                        d->pad_h - (d->R - 1), d->stride_w, d->stride_h, 64);
  if (rc != RNN_OK) return rc;

  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = d->K; p.N = d->R * d->S * d->C;   // D = [K, taps*C]
  p.k_iters = w.k_iters;
  p.splits = w.splits;
  p.ntaps = d->R * d->S;
  p.ncblk = d->C / 64;                    // 64-channel sub-blocks per tap
  p.n_mblk = w.n_mblk;
  p.n_nblk = w.ncblk;
  // CTA pairs when K_out fills a 256-row tile (measured with the per-K-step issue cost out of the way: VGG-16
  // weight gradients +30 %, ResNet-50 stages 3-4 +3..18 %)
  p.cg = pick_cg(w.BN, w.n_mblk, w.k_iters / w.splits);
  p.Wo = d->Q; p.HoWo = d->P * d->Q;
  p.ts_w = d->stride_w; p.ts_h = d->stride_h; p.lo_w = -d->pad_w; p.lo_h = -d->pad_h;
  p.ws = reinterpret_cast<float*>(ws);
  p.ldw = (long long)d->R * d->S * d->C;
  for (int r = 0; r < d->R; ++r)
    for (int s = 0; s < d->S; ++s) {
      TapDesc& t = p.taps[r * d->S + s];
      t.ow = (int16_t)s; t.oh = (int16_t)r; t.bcol = (r * d->S + s) * d->C;
    }
  rc = launch_bn<KIND_WGRAD, true, true>(w.BN, ta, tb, p, stream);
  if (rc != RNN_OK) return rc;
  return launch_reduce(p.ws, p.splits, (long long)d->K * p.ldw, (int)p.ldw, nullptr, 0, accumulate, dw, dw_f32,
                       stream);
}


// =============================================================================================
// First-layer convolution (few input channels): "packed pixel" implicit GEMM.
//
// A 3-channel image wastes 61/64 of every 64-channel K step.  Instead the input is stored NHWC with the
// channels padded to 8 and physically zero-padded in space; eight horizontally adjacent pixels (8 x 8
// channels = 64 bf16 = 128 B, contiguous in memory) are then presented to the im2col TMA as ONE pixel of a
// virtual tensor whose W stride is a single real pixel (overlapping strides are legal for TMA).  One K step
// so covers 8 horizontal filter taps; the filter is laid out to match, [K][R][SG][8 taps][8 ch] with
// SG = ceil(S/8) and zeros where s >= S or c >= C.  ResNet's 7x7 stem: 7 K steps instead of 49.
// =============================================================================================
struct FirstGeom {
  int Hp, Wp, SG, Wv, up_w, up_h, ntaps;
};

static int first_geom(const rnn_conv_desc* d, FirstGeom* g) {
  RNN_REQUIRE(d != nullptr, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->N > 0 && d->H > 0 && d->W > 0 && d->C > 0 && d->C <= 8 && d->K > 0 && d->K % 8 == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->R > 0 && d->S > 0 && d->stride_h > 0 && d->stride_w > 0 && d->pad_h >= 0 && d->pad_w >= 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->P == (d->H + 2 * d->pad_h - d->R) / d->stride_h + 1, RNN_ERR_SHAPE);
  RNN_REQUIRE(d->Q == (d->W + 2 * d->pad_w - d->S) / d->stride_w + 1, RNN_ERR_SHAPE);
  g->SG = (d->S + 7) / 8;
  g->Hp = d->H + 2 * d->pad_h;
  const int need = d->stride_w * (d->Q - 1) + 8 * g->SG;
  g->Wp = d->W + 2 * d->pad_w > need ? d->W + 2 * d->pad_w : need;
  g->Wv = g->Wp - 7;
  g->up_w = d->stride_w * (d->Q - 1) - (g->Wv - 1);
  g->up_h = d->stride_h * (d->P - 1) - (g->Hp - 1);
  g->ntaps = d->R * g->SG;
  RNN_REQUIRE(g->ntaps <= MAX_TAPS && 8 * (g->SG - 1) <= 255 && d->R - 1 <= 255, RNN_ERR_SHAPE);
  return RNN_OK;
}

static void first_taps(const rnn_conv_desc* d, const FirstGeom& g, TcParams* p) {
  for (int r = 0; r < d->R; ++r)
    for (int sg = 0; sg < g.SG; ++sg) {
      TapDesc& t = p->taps[r * g.SG + sg];
      t.ow = (int16_t)(8 * sg); t.oh = (int16_t)r; t.bcol = (r * g.SG + sg) * 64;
    }
}

/* Padded-input geometry a caller must allocate for: x_packed is [N, Hp, Wp, 8] bf16 with the image at
 * rows pad_h.., columns pad_w.. and zeros elsewhere; the filter is [K, R, SG, 64] bf16. */
int rnn_conv2d_first_geometry(const rnn_conv_desc* d, int* Hp, int* Wp, int* SG) {
  FirstGeom g;
  int rc = first_geom(d, &g);
  if (rc != RNN_OK) return rc;
  if (Hp) *Hp = g.Hp;
  if (Wp) *Wp = g.Wp;
  if (SG) *SG = g.SG;
  return RNN_OK;
}

static int first_fprop_impl(const rnn_conv_desc* d, const void* x_packed, const void* w_packed, void* y,
                            const float* scale, const float* shift, int relu, unsigned char* mask_out,
                            cudaStream_t stream);
int rnn_conv2d_first_fprop_bf16(const rnn_conv_desc* d, const void* x_packed, const void* w_packed, void* y,
                                const float* scale, const float* shift, int relu, cudaStream_t stream) {
  return first_fprop_impl(d, x_packed, w_packed, y, scale, shift, relu, nullptr, stream);
}
// first layer with the 1-bit ReLU mask as in rnn_conv2d_fprop_relu_mask_bf16
int rnn_conv2d_first_fprop_relu_mask_bf16(const rnn_conv_desc* d, const void* x_packed, const void* w_packed, void* y,
                                          const float* shift, unsigned char* mask_out, cudaStream_t stream) {
  RNN_REQUIRE(d != nullptr && mask_out != nullptr && d->K % 64 == 0, RNN_ERR_SHAPE);
  return first_fprop_impl(d, x_packed, w_packed, y, nullptr, shift, 1, mask_out, stream);
}
static int first_fprop_impl(const rnn_conv_desc* d, const void* x_packed, const void* w_packed, void* y,
                            const float* scale, const float* shift, int relu, unsigned char* mask_out,
                            cudaStream_t stream) {
  FirstGeom g;
  int rc = first_geom(d, &g);
  if (rc != RNN_OK) return rc;
  rc = load_driver_entry_points();
  if (rc != RNN_OK) return rc;
  const int BN = pick_bn(d->K);
  CUtensorMap ta, tb;
  rc = make_tmap_im2col(&ta, x_packed, 64, g.Wv, g.Hp, d->N, 8, (long long)g.Wp * 8, (long long)g.Hp * g.Wp * 8, 0, 0,
                        g.up_w, g.up_h, d->stride_w, d->stride_h, BM);
  if (rc != RNN_OK) return rc;
  const long long kcols = (long long)g.ntaps * 64;
  rc = make_tmap_2d(&tb, w_packed, d->K, kcols, kcols, BN);
  if (rc != RNN_OK) return rc;
  TcParams p;
  memset(&p, 0, sizeof(p));
  const long long M = (long long)d->N * d->P * d->Q;
  RNN_REQUIRE(M < (1ll << 31), RNN_ERR_SHAPE);
  p.M = (int)M; p.N = d->K;
  p.cchunks = 1; p.ntaps = g.ntaps; p.k_iters = g.ntaps; p.splits = 1;
  p.n_mblk = ceil_div_i(M, BM); p.n_nblk = ceil_div_i(d->K, BN);
  p.Wo = d->Q; p.HoWo = d->P * d->Q;
  p.ts_w = d->stride_w; p.ts_h = d->stride_h; p.lo_w = 0; p.lo_h = 0;
  p.ldd = d->K;
  p.D = y; p.scale = scale; p.shift = shift; p.relu = relu; p.mask_out = mask_out;
  p.no_b_resident = 1;   // 338 -> 362 us in-step with the filter resident: the 7 K steps of a tile are epilogue-bound
  first_taps(d, g, &p);
  return launch_bn<KIND_CONV, false, false>(BN, ta, tb, p, stream);
}

static WgradPlan plan_first_wgrad(const rnn_conv_desc* d, const FirstGeom& g) {
  return plan_wgrad_cols(d->K, g.ntaps, 64, (long long)d->N * d->P * d->Q);
}

size_t rnn_conv2d_first_wgrad_workspace(const rnn_conv_desc* d) {
  FirstGeom g;
  if (first_geom(d, &g) != RNN_OK) return 0;
  return plan_first_wgrad(d, g).ws_bytes;
}

/* dw_packed[K, R, SG, 64] (bf16, or fp32 when dw_f32) = gradient w.r.t. the packed filter. */
int rnn_conv2d_first_wgrad_bf16(const rnn_conv_desc* d, const void* x_packed, const void* dy, void* dw_packed,
                                int dw_f32, void* ws, size_t ws_bytes, cudaStream_t stream) {
  FirstGeom g;
  int rc = first_geom(d, &g);
  if (rc != RNN_OK) return rc;
  rc = load_driver_entry_points();
  if (rc != RNN_OK) return rc;
  const WgradPlan w = plan_first_wgrad(d, g);
  RNN_REQUIRE(ws != nullptr && ws_bytes >= w.ws_bytes, RNN_ERR_WORKSPACE);
  const long long mpix = (long long)d->N * d->P * d->Q;
  RNN_REQUIRE(mpix < (1ll << 31), RNN_ERR_SHAPE);
  CUtensorMap ta, tb;
  rc = make_tmap_2d(&ta, dy, mpix, d->K, d->K, 64);
  if (rc != RNN_OK) return rc;
  rc = make_tmap_im2col(&tb, x_packed, 64, g.Wv, g.Hp, d->N, 8, (long long)g.Wp * 8, (long long)g.Hp * g.Wp * 8, 0, 0,
                        g.up_w, g.up_h, d->stride_w, d->stride_h, 64);
  if (rc != RNN_OK) return rc;
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = d->K; p.N = g.ntaps * 64;
  p.k_iters = w.k_iters; p.splits = w.splits;
  p.ntaps = g.ntaps; p.ncblk = 1;
  p.n_mblk = w.n_mblk; p.n_nblk = w.ncblk;
  p.Wo = d->Q; p.HoWo = d->P * d->Q;
  p.ts_w = d->stride_w; p.ts_h = d->stride_h; p.lo_w = 0; p.lo_h = 0;
  p.ws = reinterpret_cast<float*>(ws);
  p.ldw = (long long)g.ntaps * 64;
  first_taps(d, g, &p);
  rc = launch_bn<KIND_WGRAD, true, true>(w.BN, ta, tb, p, stream);
  if (rc != RNN_OK) return rc;
  return launch_reduce(p.ws, p.splits, (long long)d->K * p.ldw, (int)p.ldw, nullptr, 0, 0, dw_packed, dw_f32, stream);
}

}  // extern "C"
