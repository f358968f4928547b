// Error reporting and device queries shared by every kernel file.
#include <string.h>

#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace {
thread_local char g_err[512] = "no error";
std::atomic<unsigned long long> g_launches{0};
}

void rnn_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- optional per-family device timing (bench.py): CUDA-event pairs recorded around individual launches ON THE
// STREAM THE KERNEL IS LAUNCHED ON, so a family that one entry point launches together with another (the split-K
// reduction after a weight-gradient kernel) gets its own time.  Off by default: no events, no overhead.
namespace {
struct TimedLaunch {
  std::string family;
  cudaEvent_t e0, e1;
};
std::atomic<int> g_timing_on{0};
std::mutex g_timing_mu;
std::vector<TimedLaunch> g_timed;
}  // namespace

bool rnn_timing_begin(const char* family, cudaStream_t stream, void** token) {
  *token = nullptr;
  if (!g_timing_on.load(std::memory_order_relaxed)) return false;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return false;
  TimedLaunch* t = new TimedLaunch{family, nullptr, nullptr};
  if (cudaEventCreate(&t->e0) != cudaSuccess || cudaEventCreate(&t->e1) != cudaSuccess) { delete t; return false; }
  cudaEventRecord(t->e0, stream);
  *token = t;
  return true;
}
void rnn_timing_end(void* token, cudaStream_t stream) {
  if (token == nullptr) return;
  TimedLaunch* t = reinterpret_cast<TimedLaunch*>(token);
  cudaEventRecord(t->e1, stream);
  std::lock_guard<std::mutex> lk(g_timing_mu);
  g_timed.push_back(*t);
  delete t;
}

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
int rnn_timing_enable(int on) {
  std::lock_guard<std::mutex> lk(g_timing_mu);
  for (auto& t : g_timed) { cudaEventDestroy(t.e0); cudaEventDestroy(t.e1); }
  g_timed.clear();
  g_timing_on.store(on ? 1 : 0, std::memory_order_relaxed);
  return RNN_OK;
}
int rnn_timing_read(const char* family, float* ms, int* launches) {
  RNN_REQUIRE(family != nullptr && ms != nullptr && launches != nullptr, RNN_ERR_SHAPE);
  std::lock_guard<std::mutex> lk(g_timing_mu);
  float total = 0.f;
  int n = 0;
  for (auto& t : g_timed) {
    if (t.family != family) continue;
    RNN_CUDA(cudaEventSynchronize(t.e1));
    float el = 0.f;
    RNN_CUDA(cudaEventElapsedTime(&el, t.e0, t.e1));
    total += el;
    ++n;
  }
  *ms = total;
  *launches = n;
  return RNN_OK;
}
/* Advance the device-side dropout epoch (one tiny kernel per Philox-using translation unit). */
int rnn_rng_advance(cudaStream_t stream) {
  const unsigned long long inc = 1ull << 32;   // host-side offsets are small call counters: epochs never collide
  int rc = rnn_internal_rng_advance_ew(inc, stream);
  if (rc != RNN_OK) return rc;
  return rnn_internal_rng_advance_attn(inc, stream);
}
}
