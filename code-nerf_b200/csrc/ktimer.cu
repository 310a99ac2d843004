// Kernel timers for bench.py's roofline leg and the tuning scripts: CUDA events around one of the three tensor-core
// kernels on its launching stream (off by default; no host synchronisation until the query).
#include "common.cuh"
#include "internal.h"

namespace {
constexpr int EV_MAX = 512;                 // launches per timing window
struct KTimer {
  int64_t min_rows = 0;
  cudaEvent_t ev[EV_MAX][2];
  int made = 0, used = 0;
  double ms_sum = 0.0;
  long long samples = 0, last_rows = 0;
  void collect() {
    for (int i = 0; i < used; ++i) {
      float ms = 0.f;
      if (cudaEventSynchronize(ev[i][1]) == cudaSuccess && cudaEventElapsedTime(&ms, ev[i][0], ev[i][1]) == cudaSuccess) {
        ms_sum += ms;
        ++samples;
      }
    }
    used = 0;
  }
};
KTimer g_kt[CNRF_KT_COUNT];
}  // namespace

bool cnrf_ktimer_begin(int which, int64_t rows, cudaStream_t st) {
  KTimer& t = g_kt[which];
  if (!t.min_rows || rows < t.min_rows || t.used >= EV_MAX) return false;
  if (t.used == t.made) { cudaEventCreate(&t.ev[t.made][0]); cudaEventCreate(&t.ev[t.made][1]); ++t.made; }
  cudaEventRecord(t.ev[t.used][0], st);
  t.last_rows = rows;
  return true;
}
void cnrf_ktimer_end(int which, cudaStream_t st) {
  KTimer& t = g_kt[which];
  cudaEventRecord(t.ev[t.used][1], st);
  ++t.used;
}

// which: 0 forward chain, 1 backward-data chain, 2 weight gradients.  min_rows > 0: time every launch of that kernel over
// at least min_rows sample rows from now on (resets the statistics); min_rows = 0: stop; min_rows < 0: query only.
// Returns the mean kernel duration in ms of the launches timed so far (synchronises on the last one).
extern "C" double cnrf_debug_kernel_ms(int which, int64_t min_rows, int64_t* launches, int64_t* rows) {
  if (which < 0 || which >= CNRF_KT_COUNT) return -1.0;
  KTimer& t = g_kt[which];
  t.collect();
  const double mean = t.samples ? t.ms_sum / (double)t.samples : 0.0;
  if (launches) *launches = t.samples;
  if (rows) *rows = t.last_rows;
  if (min_rows >= 0) { t.min_rows = min_rows; t.ms_sum = 0.0; t.samples = 0; }
  return mean;
}
extern "C" double cnrf_debug_fwd_kernel_ms(int64_t min_rows, int64_t* launches, int64_t* rows) {
  return cnrf_debug_kernel_ms(CNRF_KT_FWD, min_rows, launches, rows);
}
