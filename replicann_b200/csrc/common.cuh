// Shared device/host helpers for the replicann_b200 sm_100a kernels.
// Everything here is header-only; each .cu translation unit includes it.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/rnn_b200.h"

#define RNN_CHECK_LAUNCH()                                   \
  do {                                                       \
    rnn_count_launch();                                      \
    cudaError_t e__ = cudaGetLastError();                    \
    if (e__ != cudaSuccess) {                                \
      rnn_set_last_error(__FILE__, __LINE__, e__);           \
      return RNN_ERR_CUDA;                                   \
    }                                                        \
  } while (0)

#define RNN_CUDA(call)                                       \
  do {                                                       \
    cudaError_t e__ = (call);                                \
    if (e__ != cudaSuccess) {                                \
      rnn_set_last_error(__FILE__, __LINE__, e__);           \
      return RNN_ERR_CUDA;                                   \
    }                                                        \
  } while (0)

#define RNN_REQUIRE(cond, code)                              \
  do {                                                       \
    if (!(cond)) {                                           \
      rnn_set_last_message(__FILE__, __LINE__, #cond);       \
      return (code);                                         \
    }                                                        \
  } while (0)

void rnn_set_last_error(const char* file, int line, cudaError_t e);
void rnn_set_last_message(const char* file, int line, const char* msg);
int rnn_num_sms();
void rnn_count_launch();
int rnn_internal_bn_finalize(const float* partials, int nparts, int C, int blk, float count, const float* gamma,
                             const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                             float* mean, float* invstd, float* scale, float* shift, cudaStream_t stream);

// per-family launch timing (runtime.cu); a scope records an event pair around the launches it encloses
bool rnn_timing_begin(const char* family, cudaStream_t stream, void** token);
void rnn_timing_end(void* token, cudaStream_t stream);
struct RnnTimedScope {
  void* token;
  cudaStream_t stream;
  RnnTimedScope(const char* family, cudaStream_t s) : token(nullptr), stream(s) { rnn_timing_begin(family, s, &token); }
  ~RnnTimedScope() { rnn_timing_end(token, stream); }
};

int rnn_internal_tmap_3d(CUtensorMap* m, const void* ptr, long long slabs, long long rows, long long inner, long long ld,
                         long long bstride, int box_rows);
int rnn_internal_rng_advance_ew(unsigned long long inc, cudaStream_t stream);
int rnn_internal_rng_advance_attn(unsigned long long inc, cudaStream_t stream);

typedef __nv_bfloat16 bf16;

static inline int ceil_div_i(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum; `red` is shared scratch of >= 32 floats. Result valid in every thread.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  r = warp_sum(r);
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -INFINITY;
  r = warp_max(r);
  return r;
}

// 128-bit streaming load / store (read-once data; keep it out of L1)
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

#endif  // __CUDACC__
