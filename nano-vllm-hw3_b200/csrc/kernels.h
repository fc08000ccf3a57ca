// Internal launcher declarations shared by the translation units of libqlinear_b200.so.
// The public surface is include/qlinear_b200.h; these take validated arguments.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ql {

int launch_quant_rows(const void* x, int64_t M, int64_t K, int64_t ldx, int fmt, void* q,
                      float* s, cudaStream_t st);
int launch_quant_groups(const void* w, int64_t N, int64_t K, int64_t g, int fmt, void* q,
                        float* s, cudaStream_t st);
int launch_quant_tensor(const void* x, int64_t numel, int fmt, void* q, void* scale_bf16,
                        void* ws, cudaStream_t st);

// Fused quantise + GEMV for M <= 16.  Exactly one of (x) or (qx_pre, sx_pre) is used:
// x != nullptr quantises the activations inside the kernel.
int launch_gemv(const void* x, int64_t ldx, const void* qx_pre, const float* sx_pre,
                const void* qw, const float* sw, const void* bias, void* y, int64_t ldy,
                int64_t M, int64_t N, int64_t K, int fmt, int64_t group, cudaStream_t st);

// tcgen05 GEMM.  out_acc != nullptr stores raw int32 accumulators (INT8, per-row only).
int launch_gemm(const void* qx, const float* sx, const void* qw, const float* sw,
                const void* bias, void* y, int64_t ldy, int32_t* out_acc, int64_t M, int64_t N,
                int64_t K, int fmt, int64_t group, cudaStream_t st);

// Cluster split-K weight-streaming tcgen05 kernel for decode-sized M (gemm_skinny.cu).
int launch_skinny(const void* qx, const float* sx, const void* qw, const float* sw,
                  const void* bias, void* y, int64_t ldy, int32_t* out_acc, int64_t M, int64_t N,
                  int64_t K, int fmt, int64_t group, cudaStream_t st);
int launch_skinny_fused(const void* x, int64_t ldx, const void* qw, const float* sw,
                        const void* bias, void* y, int64_t ldy, int64_t M, int64_t N, int64_t K,
                        int fmt, int64_t group, cudaStream_t st);
bool skinny_ok(int64_t M, int64_t N, int64_t K, int64_t group);        // pre-quantised path usable
bool skinny_fused_ok(int64_t M, int64_t N, int64_t K, int64_t group);  // fused path usable (M <= 16)
bool skinny_fused_preferred(int64_t M, int64_t N, int64_t K, int64_t group);  // ... and the faster choice
long long* skinny_trace_ptr();
void set_skinny_trace(void* buf);  // debug: 16 x int64 device buffer, or nullptr
bool pdl_enabled();  // QL_PDL=0 disables programmatic dependent launch

// Producers fused with the per-token quantiser (producers.cu).  q == nullptr: bf16 output only.
int launch_rmsnorm_quant(const void* x, int64_t ldx, const void* res_in, void* res_out,
                         const void* w, float eps, int64_t M, int64_t K, int fmt, void* q,
                         float* s, void* h_out, cudaStream_t st);
int launch_silu_mul_quant(const void* gate_up, int64_t ld, int64_t M, int64_t I, int fmt, void* q,
                          float* s, void* h_out, cudaStream_t st);

// Peer-memory communicator state (allreduce.cu), for kernels that fuse the all-reduce.
struct ArView {
  uint8_t* area[8];     // every rank's receive area as mapped in this process
  uint32_t* ctr;        // local {epoch, done}
  size_t region;        // bytes per (parity, source rank) region
  size_t max_bytes;     // largest payload of one call
  uint64_t timeout_ns;  // flag-wait bound (0 = forever)
  int world, rank;
};
bool ar_view(ArView* v);   // false until ql_ar_open succeeded
uint64_t ar_timeout_ns();
int launch_ar_rmsnorm_quant(const ArView& ar, const void* partial, const void* res_in, void* res_out,
                            const void* w, float eps, int64_t M, int64_t K, int fmt, void* q,
                            float* s, cudaStream_t st);

// The "big" section of the symmetric area: the reduce-scatter / all-gather path behind prefill-
// sized row-parallel linears (tp_collective.cu).  Offsets are relative to a rank's big section.
struct TpLayout {
  size_t flags;   // 4 KB: ready[8] (128 B apart) | +1024: done[8] | +2048: {epoch, rows_done, -, copy_done}
  size_t part;    // bf16 partial sums: [rows_cap, N] of this rank (pull) or W receive regions (push)
  size_t q;       // gathered codes  [rows_cap, N] 1 B
  size_t s;       // gathered scales [rows_cap] fp32
  size_t y;       // gathered bf16   [rows_cap, N]
  size_t total;
};
struct TpView {
  uint8_t* big[8];      // every rank's big section as mapped in this process
  TpLayout lay;
  size_t rows, rows_cap, hidden;
  uint64_t timeout_ns;
  int world, rank;
  bool emulated;        // all ranks live in this process: one launch runs them all (tests)
};
// false unless ql_ar_init was given a large-path capacity and ql_ar_open succeeded.  `channel`
// selects one of kTpChannels independent sets of buffers, flags and call counters: two token
// micro-batches on two streams keep one collective in flight each, so the wire time of one hides
// under the GEMMs of the other.
constexpr int kTpChannels = 2;
bool tp_view(TpView* v, int channel = 0);
// big sections in the current area (1, or kTpChannels when QL_TP_ALLOW_CONCURRENT=1 was set at init);
// before init: what an init would allocate now
int tp_channels_allocated();
constexpr size_t kTpReadyOff = 0, kTpDoneOff = 1024, kTpCtrOff = 2048, kTpFlagStride = 128;
// rows per token owner: whole 256-token pair tiles, so that a GEMM tile never straddles two owners
inline int64_t tp_rows_per_owner(int64_t M, int world) {
  const int64_t tiles = (M + 255) / 256;
  return ((tiles + world - 1) / world) * 256;
}

// CTA-pair (cta_group::2) tcgen05 GEMM for prefill-sized M, per-row scales only.
int launch_gemm2(const void* qx, const float* sx, const void* qw, const float* sw,
                 const void* bias, void* y, int64_t ldy, int32_t* out_acc, int64_t M, int64_t N,
                 int64_t K, int fmt, cudaStream_t st);
bool gemm2_eligible(const void* y, int64_t ldy, const void* sw, const void* bias, bool out_acc);
// The same kernel as a fused GEMM + reduce-scatter over NVLink: every 256-token output tile is
// TMA-stored into the receive region of the rank that owns those tokens (peer memory), and the last
// reduce kernel that follows in the stream raises this rank's `ready` flag on every peer once the
// grid (and with it every bulk store) has completed.  M rows in total, owner slices of
// tp_rows_per_owner(M, world) rows.
int launch_gemm2_push(const TpView& tp, const void* qx, const float* sx, const void* qw, const float* sw,
                      const void* bias, int64_t M, int64_t N, int64_t K, int fmt, cudaStream_t st);
// Reduce (fp32, rank order) the W partials of this rank's token slice, round to bf16, then either
// quantise per token and all-gather codes + scales (fmt = QL_INT8 / QL_FP8_E4M3) or all-gather the
// bf16 rows and copy the full result to `out` (fmt < 0).  `pushed`: partials were delivered by
// launch_gemm2_push; otherwise every rank's partial sits in its own `part` buffer and is pulled.
int launch_tp_reduce(const TpView& tp, int64_t M, int64_t N, int fmt, int pushed, void* out, int64_t ldo,
                     cudaStream_t st);
// Which GEMM kernel launch_gemm picks: 1 = single-CTA weight-stationary, 2 = CTA pair.
int gemm_kernel_choice(int64_t M, int64_t N, int64_t K, int64_t group);
// Per-group scales on the CTA pair (gemm2g_kernel): 16 epilogue warps, running sums in registers.
bool gemm2_grouped_eligible(const void* y, int64_t ldy, const void* sw, int64_t K, int64_t group);
int launch_gemm2_grouped(const void* qx, const float* sx, const void* qw, const float* sw, const void* bias,
                         void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int fmt, int64_t group,
                         cudaStream_t st);

// 2-D tiled TMA descriptor: tensor [outer, inner] with `row_stride_bytes` between rows, box
// {box_inner, box_outer} elements, 128-byte swizzle, out-of-bounds reads as zero / writes clipped.
bool make_tmap_2d(CUtensorMap* tm, const void* base, CUtensorMapDataType dt, int64_t inner,
                  int64_t outer, int64_t row_stride_bytes, int box_inner, int box_outer);

bool make_tmap_2d_plain(CUtensorMap* tm, const void* base, CUtensorMapDataType dt, int64_t inner,
                        int64_t outer, int64_t row_stride_bytes, int box_inner, int box_outer);

int num_sms();

// cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE attribute and the library serves
// several devices from one process (api.cu caches per ordinal, the binding switches devices per
// call), so "set once" is tracked per device ordinal.  Setting it twice is harmless, so the
// flags need no lock.
struct SmemAttrOnce {
  unsigned char done[64] = {};
};
int ensure_max_smem(const void* kern, int bytes, SmemAttrOnce& once);

// Every kernel launch of the library bumps this (read through ql_launch_count()).
void count_launch(int n = 1);

}  // namespace ql
