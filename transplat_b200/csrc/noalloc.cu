// noalloc.cu — tsplat_forward_noalloc: the forward pass for callers that cannot allocate device memory at all (code
// under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// Every kernel of the forward chain works in a caller-owned workspace (tsplat_workspace_layout): that workspace is the
// scratch memory of the path.  Here it is a file-scope device array of a fixed, modest size instead, so the caller
// passes inputs and outputs only.  Nothing on this path calls cudaMalloc, cudaFree or an allocating library routine.
//
//   * The array exists once per GPU that loads this library and is held until the process exits
//     (TSPLAT_NOALLOC_WORKSPACE_BYTES = 32 MiB).
//   * Problems whose workspace does not fit are rejected with TSPLAT_ENOSPC and a message that names both sizes.
//   * Every call on a device shares the array.  It keeps the previous call's contents: the forward chain clears what
//     must start at zero itself (control block, look-back state, tile ranges, per-tile counters, gradient
//     accumulators: DESIGN.md 3), everything else is written before it is read.  Two calls that OVERLAP on the GPU
//     (different streams, or graphs replayed concurrently) race on it: serialise them.
//   * In host code the array's name is a host-side stand-in, not a device address; the device address is resolved here,
//     once per device and with that device current, by cudaGetSymbolAddress.
#include <mutex>

#include "common.cuh"

static __device__ __align__(256) unsigned char g_noalloc_workspace[TSPLAT_NOALLOC_WORKSPACE_BYTES];

namespace tsplat {
namespace {

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
void* g_noalloc_ptr[kMaxDevices] = {};

// Device address of the static workspace on the CURRENT device (resolved once per device).
cudaError_t noalloc_workspace(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_workspace);
    if (e != cudaSuccess) return e;
    if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 255u) != 0) return cudaErrorInvalidSymbol;
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return cudaSuccess;
}

}  // namespace
}  // namespace tsplat

extern "C" int tsplat_forward_noalloc(const TsplatProblem* p, uint64_t d_cap, float* out_color, float* out_depth,
                                      int32_t* out_radii, void* cuda_stream) {
  using namespace tsplat;
  if (p == nullptr) {
    set_error("tsplat_forward_noalloc: null problem");
    return TSPLAT_EINVAL;
  }
  TsplatLayout lay;
  const int rc = tsplat_workspace_layout(p, d_cap, &lay);
  if (rc != TSPLAT_OK) return rc;  // message set by tsplat_workspace_layout
  if (lay.total_bytes > (uint64_t)TSPLAT_NOALLOC_WORKSPACE_BYTES) {
    set_error("tsplat_forward_noalloc: this problem needs a workspace of %llu bytes, the static workspace holds %llu "
              "(TSPLAT_NOALLOC_WORKSPACE_BYTES); use tsplat_forward with a caller-owned workspace",
              (unsigned long long)lay.total_bytes, (unsigned long long)TSPLAT_NOALLOC_WORKSPACE_BYTES);
    return TSPLAT_ENOSPC;
  }
  void* ws = nullptr;
  const cudaError_t e = noalloc_workspace(&ws);
  if (e != cudaSuccess) {
    set_error("tsplat_forward_noalloc: cudaGetSymbolAddress(static workspace): %s", cudaGetErrorString(e));
    return TSPLAT_ECUDA;
  }
  return tsplat_forward(p, ws, lay.total_bytes, d_cap, out_color, out_depth, out_radii, cuda_stream);
}
