// Error reporting, device queries and the cudaIpc peer-memory fabric.
//
// The reference has no shared-memory transport at all: every model crosses process boundaries as
// protobuf bytes over gRPC (controller.cc:696-759, learner.py:197-206).  Here each GPU process
// allocates one arena with cudaMalloc, exports a cudaIpc handle, and maps every peer's arena; NVLink
// P2P makes those mappings directly loadable/storable from kernels (aggregate.cu).
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace mfl {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
  return (int)e;
}

}  // namespace mfl

using namespace mfl;

extern "C" const char* mfl_last_error(void) { return g_err; }

// Number of in-kernel waits on this device that gave up after their time budget (common.cuh: SpinGuard).  Reads
// device symbols, so it synchronises with the device: for tests and error paths, not for the hot loop.
extern "C" int mfl_spin_timeouts_aggregate(void);
extern "C" int mfl_spin_timeouts_async(void);
extern "C" int mfl_spin_timeouts(void) {
  const int a = mfl_spin_timeouts_aggregate(), b = mfl_spin_timeouts_async();
  return (a < 0 || b < 0) ? -1 : a + b;
}

extern "C" int mfl_sm_count(int device) {
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return -1;
  return n;
}

extern "C" int mfl_fabric_alloc(void** ptr, size_t bytes, void* ipc_handle_out) {
  MFL_ARG(ptr && bytes > 0, "fabric_alloc: bad arguments");
  MFL_CUDA(cudaMalloc(ptr, bytes));
  MFL_CUDA(cudaMemset(*ptr, 0, bytes));
  MFL_CUDA(cudaDeviceSynchronize());
  if (ipc_handle_out) {
    cudaIpcMemHandle_t h;
    MFL_CUDA(cudaIpcGetMemHandle(&h, *ptr));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(ipc_handle_out, &h, sizeof(h));
  }
  return 0;
}

extern "C" int mfl_fabric_open(void** ptr, const void* ipc_handle) {
  MFL_ARG(ptr && ipc_handle, "fabric_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle, sizeof(h));
  MFL_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int mfl_fabric_close(void* ptr) {
  MFL_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int mfl_fabric_free(void* ptr) {
  MFL_CUDA(cudaFree(ptr));
  return 0;
}

extern "C" int mfl_fabric_enable_peers(int device, int world) {
  for (int p = 0; p < world; ++p) {
    if (p == device) continue;
    int can = 0;
    MFL_CUDA(cudaDeviceCanAccessPeer(&can, device, p));
    MFL_ARG(can, "fabric_enable_peers: GPU %d cannot access GPU %d over P2P", device, p);
    cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); continue; }
    MFL_CUDA(e);
  }
  return 0;
}
