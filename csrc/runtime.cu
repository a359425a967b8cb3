// Host runtime: peer-mappable arenas, CUDA IPC, peer access, error plumbing.
// (No counterpart in the reference, which keeps models in host state_dicts.)
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace p2pfl {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static int g_pdl = 1;
}  // namespace p2pfl
namespace p2pfl {
bool pdl_enabled() { return g_pdl != 0; }

int sm_count(int device) {
  static int cache[64] = {0};
  if (device < 0 || device >= 64) return 148;
  if (cache[device] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
    cache[device] = n;
  }
  return cache[device];
}

}  // namespace p2pfl

using namespace p2pfl;

extern "C" {

const char* p2pfl_last_error(void) { return g_err; }
int p2pfl_version(void) { return 100; }

int p2pfl_set_pdl(int enabled) {
  int old = g_pdl;
  g_pdl = enabled ? 1 : 0;
  return old;
}

int p2pfl_rt_alloc(int device, uint64_t bytes, void** out_ptr) {
  P2PFL_REQUIRE(out_ptr != nullptr && bytes > 0, "bad arguments to p2pfl_rt_alloc");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  P2PFL_CUDA_OK(cudaMalloc(out_ptr, bytes));
  P2PFL_CUDA_OK(cudaMemset(*out_ptr, 0, bytes));
  return 0;
}

int p2pfl_rt_free(int device, void* ptr) {
  P2PFL_CUDA_OK(cudaSetDevice(device));
  P2PFL_CUDA_OK(cudaFree(ptr));
  return 0;
}

int p2pfl_rt_ipc_export(int device, void* ptr, uint8_t out_handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  P2PFL_CUDA_OK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(out_handle, &h, 64);
  return 0;
}

int p2pfl_rt_ipc_open(int device, const uint8_t handle[64], void** out_ptr) {
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  P2PFL_CUDA_OK(cudaIpcOpenMemHandle(out_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int p2pfl_rt_ipc_close(int device, void* ptr) {
  P2PFL_CUDA_OK(cudaSetDevice(device));
  P2PFL_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int p2pfl_rt_can_access_peer(int device, int peer_device, int* out_yes) {
  P2PFL_CUDA_OK(cudaDeviceCanAccessPeer(out_yes, device, peer_device));
  return 0;
}

int p2pfl_rt_enable_peer(int device, int peer_device) {
  if (device == peer_device) return 0;
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  P2PFL_CUDA_OK(e);
  return 0;
}

int p2pfl_rt_memset32(int device, void* stream, void* ptr, uint32_t value, uint64_t count) {
  P2PFL_CUDA_OK(cudaSetDevice(device));
  if (value == 0 || ((value & 0xff) == ((value >> 8) & 0xff) && (value & 0xffff) == (value >> 16))) {
    P2PFL_CUDA_OK(cudaMemsetAsync(ptr, (int)(value & 0xff), count * 4, as_stream(stream)));
    return 0;
  }
  set_error("p2pfl_rt_memset32 supports byte-replicated values only");
  return -1;
}

/* Threads of this framework call CUDA while ANOTHER thread may be capturing a graph (several
 * federated nodes per process, each building its epoch graph on first use).  A thread in CUDA's
 * default "global" capture mode is refused every potentially unsafe call (cudaMalloc, a device
 * synchronize, ...) while any capture is running - and the refusal also invalidates that capture.
 * In thread-local mode a thread is only restricted by captures of its own. */
int p2pfl_rt_thread_capture_mode_local(void) {
  cudaStreamCaptureMode mode = cudaStreamCaptureModeThreadLocal;
  P2PFL_CUDA_OK(cudaThreadExchangeStreamCaptureMode(&mode));
  return 0;
}

/* A stream of its own for every learner.  (Streams handed out by a framework's pool are shared
 * round-robin: with a few dozen federated nodes in one process two learners ended up on ONE stream,
 * and one's work landed in the other's graph capture.) */
int p2pfl_rt_stream_create(int device, void** out_stream) {
  P2PFL_REQUIRE(out_stream != nullptr, "null argument");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cudaStream_t s;
  P2PFL_CUDA_OK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *out_stream = s;
  return 0;
}

int p2pfl_rt_stream_destroy(int device, void* stream) {
  P2PFL_CUDA_OK(cudaSetDevice(device));
  P2PFL_CUDA_OK(cudaStreamDestroy(as_stream(stream)));
  return 0;
}

int p2pfl_rt_sm_clock_khz(int device, int* out_khz) {
  P2PFL_REQUIRE(out_khz != nullptr, "null argument");
  P2PFL_CUDA_OK(cudaDeviceGetAttribute(out_khz, cudaDevAttrClockRate, device));
  return 0;
}

int p2pfl_rt_device_info(int device, int* sms, int* cc_major, int* cc_minor, uint64_t* total_mem) {
  cudaDeviceProp prop;
  P2PFL_CUDA_OK(cudaGetDeviceProperties(&prop, device));
  if (sms) *sms = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (total_mem) *total_mem = (uint64_t)prop.totalGlobalMem;
  return 0;
}

}  // extern "C"
