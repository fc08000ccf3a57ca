// C ABI of libqlinear_b200.so (see include/qlinear_b200.h): argument validation, device
// check, dispatch by M.  No CPU fallback: anything but a compute-capability-10.x device gets
// QL_ERR_UNSUPPORTED_ARCH.
#include "common.cuh"
#include "kernels.h"

namespace ql {

struct DevInfo {
  int ok;   // 1 = sm_100-class device, 0 = something else, -1 = query failed
  int sms;
};

static DevInfo query_device() {
  // cached per device ordinal; the engine runs one process per GPU (model_runner.py:29)
  static DevInfo cache[64];
  static bool have[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return {-1, 0};
  if (!have[dev]) {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return {-1, 0};
    cache[dev].ok = (prop.major == 10) ? 1 : 0;
    cache[dev].sms = prop.multiProcessorCount;
    have[dev] = true;
  }
  return cache[dev];
}

int num_sms() {
  const DevInfo d = query_device();
  return d.sms > 0 ? d.sms : 148;
}

int ensure_max_smem(const void* kern, int bytes, SmemAttrOnce& once) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return QL_ERR_CUDA;
  const bool track = dev >= 0 && dev < 64;
  if (track && __atomic_load_n(&once.done[dev], __ATOMIC_ACQUIRE)) return QL_OK;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) {
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  if (track) __atomic_store_n(&once.done[dev], (unsigned char)1, __ATOMIC_RELEASE);
  return QL_OK;
}

static unsigned long long g_launches = 0;
void count_launch(int n) { __atomic_fetch_add(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }

static int check_arch() {
  const DevInfo d = query_device();
  if (d.ok < 0) return QL_ERR_CUDA;
  return d.ok ? QL_OK : QL_ERR_UNSUPPORTED_ARCH;
}

static bool fmt_ok(int fmt) { return fmt == QL_INT8 || fmt == QL_FP8_E4M3; }
static bool aligned(const void* p, uintptr_t a) { return ((uintptr_t)p % a) == 0; }

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// M <= 16: one launch, quantiser fused.  The cluster split-K tcgen05 kernel when a split plan
// fits, the mma.sync GEMV otherwise.
static int decode_linear(const void* x, int64_t ldx, const void* qw, const float* sw,
                         const void* bias, void* y, int64_t ldy, int64_t M, int64_t N, int64_t K,
                         int fmt, int64_t group, cudaStream_t st) {
  if (skinny_fused_ok(M, N, K, group)) {
    const int r = launch_skinny_fused(x, ldx, qw, sw, bias, y, ldy, M, N, K, fmt, group, st);
    if (r != QL_ERR_BAD_SHAPE) return r;
  }
  return launch_gemv(x, ldx, nullptr, nullptr, qw, sw, bias, y, ldy, M, N, K, fmt, group, st);
}

}  // namespace ql

using namespace ql;

extern "C" {

int ql_version(void) { return QL_ABI_VERSION; }

const char* ql_status_string(int status) {
  switch (status) {
    case QL_OK: return "ok";
    case QL_ERR_BAD_FORMAT: return "unsupported element format (expected int8 or float8_e4m3fn)";
    case QL_ERR_BAD_SHAPE: return "unsupported shape (K % 128, K % group, group % 128 or M limit violated)";
    case QL_ERR_MISALIGNED: return "pointer or stride not 16-byte aligned";
    case QL_ERR_UNSUPPORTED_ARCH: return "device is not compute capability 10.x (B200, sm_100a); no fallback exists";
    case QL_ERR_WORKSPACE: return "workspace missing or too small";
    case QL_ERR_CUDA: return "CUDA runtime/driver call failed";
    case QL_ERR_NULL: return "required pointer is NULL";
    case QL_ERR_STATIC_SCRATCH:
      return "problem too large for the library's built-in scratch (ql_workspace_bytes(M, K) > ql_static_scratch_bytes()); "
             "split the tokens or call ql_linear with a workspace";
    case QL_ERR_GUARDED:
      return "two-channel tensor-parallel mode is disabled: it has a known cross-GPU ordering hazard (can hang at TP=8); "
             "set QL_TP_ALLOW_CONCURRENT=1 to opt in";
    default: return "unknown status";
  }
}

int ql_device_check(void) { return check_arch(); }

void ql_debug_set_trace(void* device_buf16_i64) { set_skinny_trace(device_buf16_i64); }

uint64_t ql_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

int ql_quant_rows(const void* x, int64_t M, int64_t K, int64_t ldx, int fmt, void* q, float* s,
                  ql_stream_t stream) {
  if (!fmt_ok(fmt)) return QL_ERR_BAD_FORMAT;
  if (M < 0 || K <= 0 || ldx < K) return QL_ERR_BAD_SHAPE;
  if (M == 0) return QL_OK;
  if (!x || !q || !s) return QL_ERR_NULL;
  if (int e = check_arch()) return e;
  return launch_quant_rows(x, M, K, ldx, fmt, q, s, (cudaStream_t)stream);
}

int ql_quant_groups(const void* w, int64_t N, int64_t K, int64_t group, int fmt, void* q,
                    float* s, ql_stream_t stream) {
  if (!fmt_ok(fmt)) return QL_ERR_BAD_FORMAT;
  if (N < 0 || K <= 0 || group <= 0 || K % group != 0) return QL_ERR_BAD_SHAPE;
  if (N == 0) return QL_OK;
  if (!w || !q || !s) return QL_ERR_NULL;
  if (int e = check_arch()) return e;
  return launch_quant_groups(w, N, K, group, fmt, q, s, (cudaStream_t)stream);
}

int ql_quant_tensor(const void* x, int64_t numel, int fmt, void* q, void* scale_bf16,
                    void* workspace, size_t workspace_bytes, ql_stream_t stream) {
  if (!fmt_ok(fmt)) return QL_ERR_BAD_FORMAT;
  if (numel < 0) return QL_ERR_BAD_SHAPE;
  if (numel == 0) return QL_OK;
  if (!x || !q || !scale_bf16) return QL_ERR_NULL;
  if (!workspace || workspace_bytes < 4) return QL_ERR_WORKSPACE;
  if (!aligned(workspace, 4)) return QL_ERR_MISALIGNED;
  if (int e = check_arch()) return e;
  return launch_quant_tensor(x, numel, fmt, q, scale_bf16, workspace, (cudaStream_t)stream);
}

static int check_gemm_common(int64_t M, int64_t N, int64_t K, int fmt, int64_t group) {
  if (!fmt_ok(fmt)) return QL_ERR_BAD_FORMAT;
  if (M < 0 || N < 0 || K <= 0 || K % 128 != 0) return QL_ERR_BAD_SHAPE;  // reference :328
  if (group < 0 || (group > 0 && (group % 128 != 0 || K % group != 0))) return QL_ERR_BAD_SHAPE;
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) return QL_ERR_BAD_SHAPE;
  return QL_OK;
}

int ql_gemm_w8a8(const void* qx, const float* sx, const void* qw, const float* sw,
                 const void* bias, void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int fmt,
                 int64_t group, ql_stream_t stream) {
  if (int e = check_gemm_common(M, N, K, fmt, group)) return e;
  if (ldy < N) return QL_ERR_BAD_SHAPE;
  if (M == 0 || N == 0) return QL_OK;
  if (!qx || !sx || !qw || !sw || !y) return QL_ERR_NULL;
  if (!aligned(qx, 16) || !aligned(qw, 16)) return QL_ERR_MISALIGNED;  // TMA global address
  if (int e = check_arch()) return e;
  if (skinny_ok(M, N, K, group)) {
    const int r = launch_skinny(qx, sx, qw, sw, bias, y, ldy, nullptr, M, N, K, fmt, group,
                                (cudaStream_t)stream);
    if (r != QL_ERR_BAD_SHAPE) return r;   // BAD_SHAPE: no split plan fits, use the tiled GEMM
  }
  return launch_gemm(qx, sx, qw, sw, bias, y, ldy, nullptr, M, N, K, fmt, group,
                     (cudaStream_t)stream);
}

int ql_gemm_i8_acc(const void* qx, const void* qw, int32_t* acc, int64_t M, int64_t N, int64_t K,
                   ql_stream_t stream) {
  if (int e = check_gemm_common(M, N, K, QL_INT8, 0)) return e;
  if (M == 0 || N == 0) return QL_OK;
  if (!qx || !qw || !acc) return QL_ERR_NULL;
  if (!aligned(qx, 16) || !aligned(qw, 16)) return QL_ERR_MISALIGNED;
  if (int e = check_arch()) return e;
  if (skinny_ok(M, N, K, 0)) {
    const int r = launch_skinny(qx, nullptr, qw, nullptr, nullptr, nullptr, N, acc, M, N, K, QL_INT8,
                                0, (cudaStream_t)stream);
    if (r != QL_ERR_BAD_SHAPE) return r;
  }
  return launch_gemm(qx, nullptr, qw, nullptr, nullptr, nullptr, N, acc, M, N, K, QL_INT8, 0,
                     (cudaStream_t)stream);
}

int ql_gemv_w8a8(const void* x, int64_t ldx, const void* qw, const float* sw, const void* bias,
                 void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int fmt, int64_t group,
                 ql_stream_t stream) {
  if (int e = check_gemm_common(M, N, K, fmt, group)) return e;
  if (M > 16 || ldx < K || ldy < N) return QL_ERR_BAD_SHAPE;
  if (M == 0 || N == 0) return QL_OK;
  if (!x || !qw || !sw || !y) return QL_ERR_NULL;
  if (!aligned(x, 16) || ldx % 8 != 0 || !aligned(qw, 16)) return QL_ERR_MISALIGNED;
  if (int e = check_arch()) return e;
  return decode_linear(x, ldx, qw, sw, bias, y, ldy, M, N, K, fmt, group, (cudaStream_t)stream);
}

int ql_rmsnorm_quant(const void* x, int64_t ldx, const void* residual_in, void* residual_out,
                      const void* weight, float eps, int64_t M, int64_t K, int fmt, void* q,
                      float* s, void* h_out, ql_stream_t stream) {
  if (q != nullptr && !fmt_ok(fmt)) return QL_ERR_BAD_FORMAT;
  if (M < 0 || K <= 0 || K % 8 != 0 || ldx < K) return QL_ERR_BAD_SHAPE;
  if (M == 0) return QL_OK;
  if (!x || !weight || (q == nullptr && h_out == nullptr) || (q != nullptr && s == nullptr))
    return QL_ERR_NULL;
  if ((residual_in == nullptr) != (residual_out == nullptr)) return QL_ERR_NULL;
  if (!aligned(x, 16) || ldx % 8 != 0 || !aligned(weight, 16) || !aligned(residual_in, 16) ||
      !aligned(residual_out, 16) || !aligned(q, 8) || !aligned(h_out, 16))
    return QL_ERR_MISALIGNED;
  if ((size_t)K * 4 > 200 * 1024) return QL_ERR_BAD_SHAPE;   // the row lives in shared memory
  if (int e = check_arch()) return e;
  return launch_rmsnorm_quant(x, ldx, residual_in, residual_out, weight, eps, M, K, fmt, q, s,
                              h_out, (cudaStream_t)stream);
}

int ql_silu_mul_quant(const void* gate_up, int64_t ld, int64_t M, int64_t I, int fmt, void* q,
                      float* s, void* h_out, ql_stream_t stream) {
  if (q != nullptr && !fmt_ok(fmt)) return QL_ERR_BAD_FORMAT;
  if (M < 0 || I <= 0 || I % 8 != 0 || ld < 2 * I) return QL_ERR_BAD_SHAPE;
  if (M == 0) return QL_OK;
  if (!gate_up || (q == nullptr && h_out == nullptr) || (q != nullptr && s == nullptr))
    return QL_ERR_NULL;
  if (!aligned(gate_up, 16) || ld % 8 != 0 || !aligned(q, 8) || !aligned(h_out, 16))
    return QL_ERR_MISALIGNED;
  if ((size_t)I * 2 > 200 * 1024) return QL_ERR_BAD_SHAPE;   // the row lives in shared memory
  if (int e = check_arch()) return e;
  return launch_silu_mul_quant(gate_up, ld, M, I, fmt, q, s, h_out, (cudaStream_t)stream);
}

size_t ql_workspace_bytes(int64_t M, int64_t K) {
  if (M <= 0 || K <= 0) return 0;
  // qx [M,K] (16-byte aligned) followed by sx [M].  M <= 16 can run without it (fused quantiser)
  // but long K slices are faster through the two-launch path, which needs it.
  return align_up((size_t)M * (size_t)K, 256) + align_up((size_t)M * sizeof(float), 256);
}

const char* ql_linear_path(int64_t M, int64_t N, int64_t K, int64_t group) {
  if (M <= 16) {
    if (skinny_fused_preferred(M, N, K, group)) return "skinny_tcgen05_fused";
    if (skinny_ok(M, N, K, group)) return "quant+skinny_tcgen05";
    return skinny_fused_ok(M, N, K, group) ? "skinny_tcgen05_fused" : "gemv_mma";
  }
  if (skinny_ok(M, N, K, group)) return "skinny_tcgen05";
  const int choice = gemm_kernel_choice(M, N, K, group);
  return choice == 2 ? "gemm_tcgen05_2cta" : (choice == 3 ? "gemm_tcgen05_2cta_grouped" : "gemm_tcgen05");
}

int ql_linear(const void* x, int64_t ldx, const void* qw, const float* sw, const void* bias,
              void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int fmt, int64_t group,
              void* workspace, size_t workspace_bytes, ql_stream_t stream) {
  if (int e = check_gemm_common(M, N, K, fmt, group)) return e;
  if (ldx < K || ldy < N) return QL_ERR_BAD_SHAPE;
  if (M == 0 || N == 0) return QL_OK;
  if (!x || !qw || !sw || !y) return QL_ERR_NULL;
  if (!aligned(x, 16) || ldx % 8 != 0 || !aligned(qw, 16)) return QL_ERR_MISALIGNED;
  if (int e = check_arch()) return e;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t need = ql_workspace_bytes(M, K);
  if (M <= 16) {
    // one fused launch while the K slice is short, else quantiser + TMA-fed decode kernel;
    // without a workspace the fused / GEMV kernels still serve every M <= 16
    const bool two_launch = !skinny_fused_preferred(M, N, K, group) && skinny_ok(M, N, K, group) &&
                            workspace != nullptr && workspace_bytes >= need && aligned(workspace, 16);
    if (!two_launch) return decode_linear(x, ldx, qw, sw, bias, y, ldy, M, N, K, fmt, group, st);
  }
  if (workspace == nullptr || workspace_bytes < need) return QL_ERR_WORKSPACE;
  if (!aligned(workspace, 16)) return QL_ERR_MISALIGNED;
  uint8_t* qx = static_cast<uint8_t*>(workspace);
  float* sx = reinterpret_cast<float*>(qx + align_up((size_t)M * (size_t)K, 256));
  if (int e = launch_quant_rows(x, M, K, ldx, fmt, qx, sx, st)) return e;
  if (skinny_ok(M, N, K, group)) {
    const int r = launch_skinny(qx, sx, qw, sw, bias, y, ldy, nullptr, M, N, K, fmt, group, st);
    if (r != QL_ERR_BAD_SHAPE) return r;
  }
  return launch_gemm(qx, sx, qw, sw, bias, y, ldy, nullptr, M, N, K, fmt, group, st);
}

}  // extern "C"
