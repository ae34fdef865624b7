// pn_mlp_forward_tc_static: the tensor-core ResnetFC chain for callers that cannot allocate.
//
// No counterpart in the reference.  Some callers cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  pn_mlp_forward_tc
// needs a caller-provided workspace (pn_mlp_tc_workspace_bytes: a 256-byte header whose first word is the watchdog
// error word, plus -- when n_views > 1 -- the fp16 hand-off of h between the per-view and the merged segment,
// 1 KB per (view, point) row).  This variant takes that scratch from a file-scope device array instead, so the
// call allocates nothing: no cudaMalloc / cudaFree, no allocating library routine, no torch allocation.
//
// Limits and sharing (also in include/pixelnerf_b200.h):
//   * the array is PN_STATIC_SCRATCH_BYTES = 32 MiB, held by every GPU that loads the library until exit;
//     a call whose workspace would exceed it (n_views > 1 and more than 32 767 rows) is rejected with PN_ERR_INVALID;
//   * EVERY call on a device shares the one array, which keeps the previous call's contents.  What must start at
//     zero (the error word) is cleared by every call on its stream; the hand-off region is fully written by the
//     first launch of a call before the second launch reads it, so stale contents are never consumed;
//   * calls that overlap on the GPU (two streams, two host threads on one device) RACE on the array: serialise them.
#include <atomic>

#include "pn_common.cuh"

// In host code this name is a host-side stand-in for the symbol, NOT a device address: passing it to a kernel
// faults (or, where the GPU can reach host memory, silently uses the wrong memory).  Host code below resolves the
// device address once per device with cudaGetSymbolAddress and passes that.
static __device__ __align__(256) unsigned char pn_static_scratch[PN_STATIC_SCRATCH_BYTES];

namespace pn {

// Device address of pn_static_scratch on the CURRENT device, resolved once per device (with that device current).
static int static_scratch_ptr(void** out) {
  static std::atomic<void*> cached[64];
  int dev = 0;
  PN_CUDA_OK(cudaGetDevice(&dev));
  PN_CHECK_ARG(dev >= 0 && dev < 64, "pn_mlp_forward_tc_static: device ordinal %d out of range", dev);
  void* p = cached[dev].load(std::memory_order_acquire);
  if (!p) {
    PN_CUDA_OK(cudaGetSymbolAddress((void**)&p, pn_static_scratch));
    PN_CHECK_ARG(p != nullptr && ((uintptr_t)p & 255) == 0, "pn_mlp_forward_tc_static: bad scratch address");
    cached[dev].store(p, std::memory_order_release);   // a racing thread resolves the same address: harmless
  }
  *out = p;
  return PN_OK;
}

}  // namespace pn

extern "C" size_t pn_static_scratch_bytes(void) { return (size_t)PN_STATIC_SCRATCH_BYTES; }

extern "C" int pn_mlp_forward_tc_static(const void* packed, const void* tables, int64_t n_texels, const void* feat_x,
                                        int n_objects, int n_views, int64_t points_per_object, float* out,
                                        float* pre_act, pn_stream_t stream) {
  PN_CHECK_ARG(n_objects > 0 && n_views > 0 && points_per_object >= 0, "pn_mlp_forward_tc_static: bad sizes");
  const int64_t n_rows = (int64_t)n_objects * n_views * points_per_object;
  const size_t need = pn_mlp_tc_workspace_bytes(n_rows, n_views);
  PN_CHECK_ARG(need <= (size_t)PN_STATIC_SCRATCH_BYTES,
               "pn_mlp_forward_tc_static: %lld rows x %d views need %zu bytes of scratch, the static scratch array "
               "holds %zu (use pn_mlp_forward_tc with a caller-allocated workspace, or split the call)",
               (long long)n_rows, n_views, need, (size_t)PN_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  int rc = pn::static_scratch_ptr(&scratch);
  if (rc != PN_OK) return rc;
  // pn_mlp_forward_tc clears the error word (first 4 bytes) on `stream` before its launches; nothing else of the
  // scratch is read before it is written
  return pn_mlp_forward_tc(packed, tables, n_texels, feat_x, n_objects, n_views, points_per_object, out, pre_act,
                           scratch, (size_t)PN_STATIC_SCRATCH_BYTES, nullptr, stream);
}
