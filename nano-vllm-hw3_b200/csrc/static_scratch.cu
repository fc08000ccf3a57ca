// ql_linear_static: ql_linear for callers that have no way to hand the library a workspace.
//
// ql_linear needs a scratch area (the quantised activations qx [M,K] and their scales sx [M])
// and leaves its allocation to the caller -- inside the engine that is a torch tensor.  Two
// kinds of caller have nothing to allocate it with:
//   * a plain C / ctypes / FFI client that binds this ABI without torch, so it has no caching
//     allocator to draw from, only the raw CUDA runtime;
//   * anything that runs after the serving engine has sized its KV cache to whatever HBM was
//     left (the reference does that at start-up, nanovllm/engine/model_runner.py allocate_kv_cache):
//     a cudaMalloc at that point fails with out-of-memory, and one that succeeds synchronises
//     the whole device in the middle of a forward.
// For them the scratch lives in the library image itself: one file-scope __device__ array,
// present on every GPU that loads the library from the first CUDA call until the process ends.
//
// The array name below is only a handle on the host; the device address is fetched once per
// device with cudaGetSymbolAddress (scratch_for_current_device) and that address is what the
// kernels get.
#include "common.cuh"
#include "kernels.h"

namespace ql {

// qx [M,K] + sx [M], laid out exactly like a ql_linear workspace (ql_workspace_bytes(M, K)).
// 8 MiB holds e.g. 4096 tokens x K = 1024 (+ 16 KiB of scales) or 2048 tokens x K = 3072.
static __device__ __align__(256) unsigned char g_static_scratch[QL_STATIC_SCRATCH_BYTES];

// Device address of g_static_scratch on the CURRENT device, resolved on that device's first
// call and cached per ordinal.  No allocation: the array is part of the loaded module.
static int scratch_for_current_device(void** out) {
  static void* addr[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return QL_ERR_CUDA; }
  const bool track = dev >= 0 && dev < 64;
  void* p = track ? __atomic_load_n(&addr[dev], __ATOMIC_ACQUIRE) : nullptr;
  if (p == nullptr) {
    if (cudaGetSymbolAddress(&p, g_static_scratch) != cudaSuccess || p == nullptr) {
      (void)cudaGetLastError();
      return QL_ERR_CUDA;
    }
    if (track) __atomic_store_n(&addr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return QL_OK;
}

}  // namespace ql

using namespace ql;

extern "C" {

size_t ql_static_scratch_bytes(void) { return (size_t)QL_STATIC_SCRATCH_BYTES; }

int ql_linear_static(const void* x, int64_t ldx, const void* qw, const float* sw, const void* bias,
                     void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int fmt, int64_t group,
                     ql_stream_t stream) {
  if (M > 0 && K > 0 && ql_workspace_bytes(M, K) > (size_t)QL_STATIC_SCRATCH_BYTES) {
    // Too large for the built-in scratch.  M <= 16 only gets here with an absurd K; otherwise let
    // ql_linear validate the rest first (it launches nothing without a workspace when M > 16), so
    // a bad format / shape / pointer / device keeps its own error code.
    if (M <= 16) return QL_ERR_STATIC_SCRATCH;
    const int r = ql_linear(x, ldx, qw, sw, bias, y, ldy, M, N, K, fmt, group, nullptr, 0, stream);
    return r == QL_ERR_WORKSPACE ? QL_ERR_STATIC_SCRATCH : r;
  }
  void* scratch = nullptr;
  if (M > 0 && N > 0 && ql_device_check() == QL_OK) {
    if (int e = scratch_for_current_device(&scratch)) return e;
  }
  // Nothing in the scratch has to start at zero: the quantiser rewrites every byte of qx and sx
  // that the GEMM of the same call reads (rows past M are out of bounds of the TMA box and come
  // back as zeros from the hardware), so the previous call's contents are never observed.
  // Argument validation and dispatch are ql_linear's own; with scratch == NULL (an empty
  // problem or the wrong device) it returns before any launch.
  return ql_linear(x, ldx, qw, sw, bias, y, ldy, M, N, K, fmt, group, scratch,
                   scratch ? (size_t)QL_STATIC_SCRATCH_BYTES : 0, stream);
}

}  // extern "C"
