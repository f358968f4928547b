// Token / positional embedding kernels (HBM-bound gather and a deterministic scatter-add).
//
// Not in the reference: SURVEY.md section 8(f).4 names "encoder/decoder stacks + embeddings + LM loss" as the next
// network replication upstream would add; semantics are torch.nn.Embedding's (weight [V, E], int64 indices).
//
//   forward : out[n, :] = tok[ids[n], :] (+ pos[n % T, :])                 one warp per token row, 128-bit accesses
//   backward: dtok[v, :] (+)= sum over tokens n with ids[n] == v of dy[n, :], summed in ASCENDING n
//             dpos[t, :] (+)= sum over the batch b of dy[b * T + t, :], summed in ascending b
//
// The backward pass uses no floating-point atomics: the summation order is fixed, so gradients are bitwise
// reproducible from run to run (the same property the split-K and BatchNorm reductions of this library have).
// A histogram of the indices (integer atomics: order-independent) tells every vocabulary row how many tokens it
// has to find; rows without tokens only clear (or keep) their gradient, rows with tokens scan the index list 32 at
// a time (ballot) and stop as soon as they have found all of theirs.  Cost: (rows that occur) x n_tokens / 32 warp
// steps -- 4 M for 4096 tokens, microseconds; the index list (<= a few hundred KB) stays in L1/L2.
#include "common.cuh"

namespace {

template <typename T>
struct EVec;
template <>
struct EVec<bf16> {
  static constexpr int W = 8;
  __device__ static void load(const bf16* p, float* f) { unpack8(*reinterpret_cast<const uint4*>(p), f); }
  __device__ static void store(bf16* p, const float* f) { *reinterpret_cast<uint4*>(p) = pack8(f); }
};
template <>
struct EVec<float> {
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
embedding_fwd_kernel(const long long* __restrict__ ids, const T* __restrict__ tok, const T* __restrict__ pos,
                     T* __restrict__ out, long long n_tokens, int Tctx, int V, int E) {
  constexpr int W = EVec<T>::W;
  const int lane = threadIdx.x & 31;
  const long long n = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (n >= n_tokens) return;
  const long long id = ids[n];
  const bool ok = id >= 0 && id < V;          // an index outside the table reads as a zero row
  const T* src = tok + (ok ? id : 0) * E;
  const T* pr = pos ? pos + (n % Tctx) * (long long)E : nullptr;
  for (int c = lane * W; c < E; c += 32 * W) {
    float v[W];
    if (ok) EVec<T>::load(src + c, v);
    else {
#pragma unroll
      for (int e = 0; e < W; ++e) v[e] = 0.f;
    }
    if (pr) {
      float p[W];
      EVec<T>::load(pr + c, p);
#pragma unroll
      for (int e = 0; e < W; ++e) v[e] += p[e];
    }
    EVec<T>::store(out + n * E + c, v);
  }
}

__global__ void embedding_hist_kernel(const long long* __restrict__ ids, long long n_tokens, int V,
                                      int* __restrict__ counts) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_tokens;
       i += (long long)gridDim.x * blockDim.x) {
    const long long id = ids[i];
    if (id >= 0 && id < V) atomicAdd(&counts[id], 1);
  }
}

// grid.x: vocabulary rows, 8 per block (one warp each); grid.y: column tiles of 32 lanes x W elements
template <typename T>
__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const long long* __restrict__ ids, const int* __restrict__ counts, const T* __restrict__ dy,
                     T* __restrict__ dtok, long long n_tokens, int V, int E, int accumulate) {
  constexpr int W = EVec<T>::W;
  const int lane = threadIdx.x & 31;
  const int v = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (v >= V) return;
  const int c = (blockIdx.y * 32 + lane) * W;
  const bool col_ok = c < E;
  int remaining = counts[v];
  float acc[W];
#pragma unroll
  for (int e = 0; e < W; ++e) acc[e] = 0.f;
  if (remaining == 0) {
    if (!accumulate && col_ok) EVec<T>::store(dtok + (long long)v * E + c, acc);
    return;
  }
  for (long long base = 0; base < n_tokens && remaining > 0; base += 32) {
    const long long i = base + lane;
    const bool hit = i < n_tokens && ids[i] == (long long)v;
    unsigned m = __ballot_sync(0xffffffffu, hit);
    remaining -= __popc(m);
    while (m) {                                  // ascending token order: lowest set bit first
      const int b = __ffs(m) - 1;
      m &= m - 1;
      if (col_ok) {
        float g[W];
        EVec<T>::load(dy + (base + b) * E + c, g);
#pragma unroll
        for (int e = 0; e < W; ++e) acc[e] += g[e];
      }
    }
  }
  if (!col_ok) return;
  T* dst = dtok + (long long)v * E + c;
  if (accumulate) {
    float old[W];
    EVec<T>::load(dst, old);
#pragma unroll
    for (int e = 0; e < W; ++e) acc[e] += old[e];
  }
  EVec<T>::store(dst, acc);
}

// dpos[t, :] (+)= sum_b dy[b*T + t, :]; one thread per (t, W-wide column group)
template <typename T>
__global__ void __launch_bounds__(256)
pos_embedding_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dpos, long long n_tokens, int Tctx, int E,
                         int accumulate) {
  constexpr int W = EVec<T>::W;
  const int cgs = E / W;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)Tctx * cgs) return;
  const int t = (int)(idx / cgs), c = (int)(idx - (long long)t * cgs) * W;
  float acc[W];
#pragma unroll
  for (int e = 0; e < W; ++e) acc[e] = 0.f;
  for (long long n = t; n < n_tokens; n += Tctx) {
    float g[W];
    EVec<T>::load(dy + n * E + c, g);
#pragma unroll
    for (int e = 0; e < W; ++e) acc[e] += g[e];
  }
  T* dst = dpos + (long long)t * E + c;
  if (accumulate) {
    float old[W];
    EVec<T>::load(dst, old);
#pragma unroll
    for (int e = 0; e < W; ++e) acc[e] += old[e];
  }
  EVec<T>::store(dst, acc);
}

}  // namespace

extern "C" {

int rnn_embedding_fwd(const long long* ids, const void* tok, const void* pos, void* out, long long n_tokens, int T,
                      int V, int E, int is_f32, cudaStream_t stream) {
  RNN_REQUIRE(n_tokens > 0 && T > 0 && V > 0 && E > 0 && E % (is_f32 ? 4 : 8) == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(pos == nullptr || n_tokens % T == 0, RNN_ERR_SHAPE);
  const int blocks = ceil_div_i(n_tokens, 8);
  if (is_f32)
    embedding_fwd_kernel<float><<<blocks, 256, 0, stream>>>(ids, (const float*)tok, (const float*)pos, (float*)out,
                                                            n_tokens, T, V, E);
  else
    embedding_fwd_kernel<bf16><<<blocks, 256, 0, stream>>>(ids, (const bf16*)tok, (const bf16*)pos, (bf16*)out,
                                                           n_tokens, T, V, E);
  RNN_CHECK_LAUNCH();
  return RNN_OK;
}

size_t rnn_embedding_bwd_workspace(int V) { return V > 0 ? (size_t)V * sizeof(int) : 0; }

int rnn_embedding_bwd(const long long* ids, const void* dy, void* dtok, void* dpos, long long n_tokens, int T, int V,
                      int E, int is_f32, int accumulate_tok, int accumulate_pos, void* ws, size_t ws_bytes,
                      cudaStream_t stream) {
  RNN_REQUIRE(n_tokens > 0 && T > 0 && V > 0 && E > 0 && E % (is_f32 ? 4 : 8) == 0, RNN_ERR_SHAPE);
  RNN_REQUIRE(dpos == nullptr || n_tokens % T == 0, RNN_ERR_SHAPE);
  const int W = is_f32 ? 4 : 8;
  if (dtok != nullptr) {
    RNN_REQUIRE(ws != nullptr && ws_bytes >= rnn_embedding_bwd_workspace(V), RNN_ERR_WORKSPACE);
    int* counts = reinterpret_cast<int*>(ws);
    RNN_CUDA(cudaMemsetAsync(counts, 0, (size_t)V * sizeof(int), stream));
    int hb = ceil_div_i(n_tokens, 256);
    if (hb > rnn_num_sms() * 8) hb = rnn_num_sms() * 8;
    embedding_hist_kernel<<<hb, 256, 0, stream>>>(ids, n_tokens, V, counts);
    RNN_CHECK_LAUNCH();
    const dim3 grid(ceil_div_i(V, 8), ceil_div_i(E, 32 * W));
    if (is_f32)
      embedding_bwd_kernel<float><<<grid, 256, 0, stream>>>(ids, counts, (const float*)dy, (float*)dtok, n_tokens, V, E,
                                                            accumulate_tok);
    else
      embedding_bwd_kernel<bf16><<<grid, 256, 0, stream>>>(ids, counts, (const bf16*)dy, (bf16*)dtok, n_tokens, V, E,
                                                           accumulate_tok);
    RNN_CHECK_LAUNCH();
  }
  if (dpos != nullptr) {
    const int blocks = ceil_div_i((long long)T * (E / W), 256);
    if (is_f32)
      pos_embedding_bwd_kernel<float><<<blocks, 256, 0, stream>>>((const float*)dy, (float*)dpos, n_tokens, T, E,
                                                                  accumulate_pos);
    else
      pos_embedding_bwd_kernel<bf16><<<blocks, 256, 0, stream>>>((const bf16*)dy, (bf16*)dpos, n_tokens, T, E,
                                                                 accumulate_pos);
    RNN_CHECK_LAUNCH();
  }
  return RNN_OK;
}

}  // extern "C"
