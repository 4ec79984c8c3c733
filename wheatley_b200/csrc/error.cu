// Error string, launch counter and the optional per-kernel-class event timing.
#include <stdlib.h>
#include <stdarg.h>
#include <atomic>
#include <mutex>
#include <vector>
#include "common.cuh"

namespace b200gat {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static std::atomic<int64_t> g_launches{0};
void note_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static std::atomic<int> g_timing{0};
static std::mutex g_tmu;
struct EvPair { cudaEvent_t a, b; };
static std::vector<EvPair> g_pending[B200GAT_TIMING_CLASSES];
// Event pairs recorded as EXTERNAL event nodes of a CUDA graph under capture: they are re-recorded by every
// replay and read (not destroyed) by b200gat_timing_collect_graph().
static std::vector<EvPair> g_graph_pairs[B200GAT_TIMING_CLASSES];
static double g_ms[B200GAT_TIMING_CLASSES];
static int64_t g_cnt[B200GAT_TIMING_CLASSES];

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200GAT_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

bool timing_enabled() { return g_timing.load(std::memory_order_relaxed); }

ScopedTimer::ScopedTimer(int cls, cudaStream_t st) : cls_(cls), st_(st), on_(false) {
  if (!g_timing.load(std::memory_order_relaxed) || cls < 0 || cls >= B200GAT_TIMING_CLASSES) return;
  if (cudaEventCreate(&e0_) != cudaSuccess || cudaEventCreate(&e1_) != cudaSuccess) return;
  on_ = true;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  capturing_ = cudaStreamIsCapturing(st_, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusActive;
  if (capturing_) cudaEventRecordWithFlags(e0_, st_, cudaEventRecordExternal);
  else cudaEventRecord(e0_, st_);
}
ScopedTimer::~ScopedTimer() {
  if (!on_) return;
  if (capturing_) cudaEventRecordWithFlags(e1_, st_, cudaEventRecordExternal);
  else cudaEventRecord(e1_, st_);
  std::lock_guard<std::mutex> lk(g_tmu);
  (capturing_ ? g_graph_pairs : g_pending)[cls_].push_back({e0_, e1_});
}

static void drain() {
  for (int c = 0; c < B200GAT_TIMING_CLASSES; ++c) {
    for (auto& p : g_pending[c]) {
      float ms = 0.f;
      if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
        g_ms[c] += ms;
        g_cnt[c] += 1;
      }
      cudaEventDestroy(p.a);
      cudaEventDestroy(p.b);
    }
    g_pending[c].clear();
  }
}
}  // namespace b200gat

using namespace b200gat;

extern "C" const char* b200gat_last_error(void) { return g_err; }
extern "C" int b200gat_version(void) { return 210; }   // 2.1: graph flags, no-alloc colsum, side stream, dependent launches
extern "C" int64_t b200gat_launch_count(void) { return g_launches.load(); }
extern "C" void b200gat_timing_enable(int on) {
  std::lock_guard<std::mutex> lk(g_tmu);
  drain();
  for (int c = 0; c < B200GAT_TIMING_CLASSES; ++c) { g_ms[c] = 0; g_cnt[c] = 0; }
  g_timing.store(on ? 1 : 0);
}
extern "C" int b200gat_timing_collect_graph(void) {
  std::lock_guard<std::mutex> lk(g_tmu);
  int n = 0;
  for (int c = 0; c < B200GAT_TIMING_CLASSES; ++c)
    for (auto& p : g_graph_pairs[c]) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
        g_ms[c] += ms;
        g_cnt[c] += 1;
        ++n;
      }
    }
  cudaGetLastError();   // an unrecorded pair is not an error of the caller
  return n;
}
extern "C" void b200gat_timing_clear_graph(void) {
  std::lock_guard<std::mutex> lk(g_tmu);
  for (int c = 0; c < B200GAT_TIMING_CLASSES; ++c) {
    for (auto& p : g_graph_pairs[c]) {
      cudaEventDestroy(p.a);
      cudaEventDestroy(p.b);
    }
    g_graph_pairs[c].clear();
  }
}
extern "C" int b200gat_timing_read(int cls, double* total_ms, int64_t* count) {
  B200_CHECK_ARG(cls >= 0 && cls < B200GAT_TIMING_CLASSES && total_ms && count, "bad timing class");
  std::lock_guard<std::mutex> lk(g_tmu);
  drain();
  *total_ms = g_ms[cls];
  *count = g_cnt[cls];
  return B200GAT_OK;
}
