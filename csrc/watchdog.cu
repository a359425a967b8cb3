// Exchange watchdog: the host side of the data plane's bounded waits, in native code.
//
// reference site replaced: the aggregation timeout of the learning thread
// (p2pfl/learning/aggregators/aggregator.py:163-186 - a lock acquired with a timeout) for rounds
// whose exchange is ONE kernel that waits on its peers' arrival flags (SURVEY.md section 5.8,
// "watchdog").  Such a kernel ends by itself after its cycle budget; somebody still has to end it EARLY
// when the control plane learns that a contributor is gone, and to notice promptly when it is done.
// That somebody was a Python loop (event query + 30 us sleep): it took the interpreter lock a few
// thousand times per second per node, from the thread that launches the next round - while seven
// reader threads of the same process need that lock to answer their peers.
//
// p2pfl_rt_watch blocks the calling thread WITHOUT the interpreter (ctypes drops the lock around
// foreign calls): it polls the completion event, a host cancel word that any thread may set, and
// the clock, and flips the device abort word itself - through a copy on a stream of its own, so it
// overtakes the kernel it is about to stop - before it waits for the kernel to leave.
#include <time.h>

#include <atomic>
#include <mutex>

#include "common.cuh"

using namespace p2pfl;

namespace {

constexpr int kMaxDevices = 64;

struct AbortPath {
  cudaStream_t stream = nullptr;
  uint32_t* one = nullptr;  // pinned host word holding 1: the source of the abort copy
};

std::mutex g_mu;
AbortPath g_path[kMaxDevices];

int abort_path(int device, AbortPath* out) {
  P2PFL_REQUIRE(device >= 0 && device < kMaxDevices, "device index out of range");
  std::lock_guard<std::mutex> lock(g_mu);
  AbortPath& p = g_path[device];
  if (p.stream == nullptr) {
    cudaStream_t s;
    uint32_t* one;
    P2PFL_CUDA_OK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    cudaError_t e = cudaHostAlloc((void**)&one, sizeof(uint32_t), cudaHostAllocPortable);
    if (e != cudaSuccess) {
      cudaStreamDestroy(s);
      P2PFL_CUDA_OK(e);
    }
    *one = 1u;
    p.stream = s;
    p.one = one;
  }
  *out = p;
  return 0;
}

inline double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

// 0: completed, 1: still running, <0 / >1: error (already recorded)
inline int query(cudaEvent_t ev) {
  cudaError_t e = cudaEventQuery(ev);
  if (e == cudaSuccess) return 0;
  if (e == cudaErrorNotReady) {
    (void)cudaGetLastError();  // "not ready" is an answer, not a fault: do not leave it as the thread's last error
    return 1;
  }
  set_error("cudaEventQuery failed: %s", cudaGetErrorString(e));
  return -(int)e;
}

inline void pause_for(double waited) {
  // the first 100 us of a wait are polled back to back (an exchange among synchronised peers takes
  // tens of microseconds); after that the thread sleeps 20 us between polls
  if (waited < 100e-6) {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
    return;
  }
  timespec ts{0, 20000};
  nanosleep(&ts, nullptr);
}

}  // namespace

extern "C" {

int p2pfl_rt_watch(int device, void* done_event, uint32_t* abort_dev, const volatile int32_t* cancel, double timeout_s,
                   double slice_s, double grace_s, double* waited_s) {
  P2PFL_REQUIRE(done_event != nullptr, "null event");
  cudaEvent_t ev = (cudaEvent_t)done_event;
  const double t0 = now_s();
  int verdict = P2PFL_WATCH_DONE;
  for (;;) {
    int q = query(ev);
    const double waited = now_s() - t0;
    if (q == 0) {
      if (waited_s) *waited_s = waited;
      return P2PFL_WATCH_DONE;
    }
    if (q != 1) return q;
    if (cancel != nullptr && *cancel != 0) {
      verdict = P2PFL_WATCH_CANCELLED;
      break;
    }
    if (timeout_s >= 0 && waited >= timeout_s) {
      verdict = P2PFL_WATCH_TIMEOUT;
      break;
    }
    if (slice_s > 0 && waited >= slice_s) {
      if (waited_s) *waited_s = waited;
      return P2PFL_WATCH_SLICE;
    }
    pause_for(waited);
  }
  // stop the kernel: it reads the abort word in every wait loop and finalises what it holds
  if (abort_dev != nullptr) {
    AbortPath p;
    P2PFL_CUDA_OK(cudaSetDevice(device));
    int rc = abort_path(device, &p);
    if (rc != 0) return rc;
    P2PFL_CUDA_OK(cudaMemcpyAsync(abort_dev, p.one, sizeof(uint32_t), cudaMemcpyHostToDevice, p.stream));
  }
  const double t1 = now_s();
  for (;;) {
    int q = query(ev);
    if (q == 0) break;
    if (q != 1) return q;
    if (grace_s >= 0 && now_s() - t1 > grace_s) {
      set_error("the kernel did not leave within %.1f s of the abort word being set", grace_s);
      return -2;
    }
    pause_for(1.0);
  }
  if (waited_s) *waited_s = now_s() - t0;
  return verdict;
}

}  // extern "C"
