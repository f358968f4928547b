// Error reporting and device queries shared by every kernel file.
#include <string.h>

#include <atomic>

#include "common.cuh"

namespace {
thread_local char g_err[512] = "no error";
std::atomic<unsigned long long> g_launches{0};
}

void rnn_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

void rnn_set_last_error(const char* file, int line, cudaError_t e) {
  const char* base = strrchr(file, '/');
  snprintf(g_err, sizeof(g_err), "%s:%d: CUDA error %d (%s)", base ? base + 1 : file, line, (int)e,
           cudaGetErrorString(e));
}
void rnn_set_last_message(const char* file, int line, const char* msg) {
  const char* base = strrchr(file, '/');
  snprintf(g_err, sizeof(g_err), "%s:%d: requirement failed: %s", base ? base + 1 : file, line, msg);
}

int rnn_num_sms() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}

extern "C" {
const char* rnn_last_error(void) { return g_err; }
int rnn_version(void) { return 100; }
int rnn_sm_count(void) { return rnn_num_sms(); }
unsigned long long rnn_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
/* Advance the device-side dropout epoch (one tiny kernel per Philox-using translation unit). */
int rnn_rng_advance(cudaStream_t stream) {
  const unsigned long long inc = 1ull << 32;   // host-side offsets are small call counters: epochs never collide
  int rc = rnn_internal_rng_advance_ew(inc, stream);
  if (rc != RNN_OK) return rc;
  return rnn_internal_rng_advance_attn(inc, stream);
}
}
