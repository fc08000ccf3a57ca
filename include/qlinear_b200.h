/*
 * qlinear_b200.h -- C ABI of the B200-native (sm_100a) quantized-linear engine.
 *
 * This is the drop-in boundary for ONE hot path of songxxzp/nano-vllm-hw3: the INT8 / FP8
 * (e4m3) per-row and per-group quantized linear layers.  Every entry point names the
 * reference interface it replaces (paths relative to the reference repository root).  The
 * reference is pure Python, so the binding a maintainer adds is a ctypes stub; it is shown
 * in INTEGRATION.md and shipped as nano-vllm-hw3_b200/nanovllm/_ext.py.
 *
 * Conventions
 *   - plain pointers and sizes only; all pointers are DEVICE pointers unless stated;
 *   - every call is stream-ordered on `stream` (a cudaStream_t passed as void*), never
 *     synchronises, never allocates device memory, keeps no per-call global state and is
 *     safe inside CUDA-graph capture (the reference captures its decode step,
 *     nanovllm/engine/model_runner.py:316-356);
 *   - returns QL_OK (0) or a QL_ERR_* code; never throws, never aborts;
 *   - there is NO CPU fallback: on a device that is not compute capability 10.0 every
 *     compute entry point returns QL_ERR_UNSUPPORTED_ARCH.
 *
 * Layouts (identical to the reference's in-memory tensors)
 *   x    bf16  [M, K], row stride ldx elements, K contiguous (16-byte aligned rows)
 *   qx   1 B   [M, K] dense             int8 or e4m3 bit patterns
 *   sx   fp32  [M]                      per-token scales
 *   qw   1 B   [N, K] dense, K contiguous
 *   sw   fp32  [N] (per-row, group == 0) or [N, K/group] row-major (per-group)
 *   bias bf16  [N] or NULL
 *   y    bf16  [M, N], row stride ldy elements
 */
#ifndef QLINEAR_B200_H_
#define QLINEAR_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QL_ABI_VERSION 9

/* status codes */
#define QL_OK 0
#define QL_ERR_BAD_FORMAT 1       /* fmt is not QL_INT8 / QL_FP8_E4M3 (reference: assert dtype in [...], quantization.py:151,203,353) */
#define QL_ERR_BAD_SHAPE 2        /* K % 128 != 0 (quantization.py:328), K % group != 0 (:205), M > 16 for the GEMV, negative sizes */
#define QL_ERR_MISALIGNED 3       /* pointer / stride alignment the kernels need (16 B) */
#define QL_ERR_UNSUPPORTED_ARCH 4 /* current device is not sm_100 */
#define QL_ERR_WORKSPACE 5        /* caller-provided workspace too small or NULL */
#define QL_ERR_CUDA 6             /* a CUDA runtime/driver call failed (launch error, TMA descriptor encode) */
#define QL_ERR_NULL 7             /* required pointer is NULL */
#define QL_ERR_GUARDED 9          /* channel > 0 / ql_tp_set_concurrent(1) without QL_TP_ALLOW_CONCURRENT=1 (known hazard, see ql_tp_channels) */
#define QL_ERR_STATIC_SCRATCH 8   /* ql_linear_static: the problem needs more than QL_STATIC_SCRATCH_BYTES of scratch */

/* element formats of the quantized tensors */
#define QL_INT8 0     /* torch.int8,          qmax 127 */
#define QL_FP8_E4M3 1 /* torch.float8_e4m3fn, qmax 448 */

typedef void* ql_stream_t; /* cudaStream_t */

#if defined(__GNUC__)
#define QL_API __attribute__((visibility("default")))
#else
#define QL_API
#endif

/* ABI version of the loaded library (== QL_ABI_VERSION it was built with). */
QL_API int ql_version(void);

/* Static, NUL-terminated description of a status code. */
QL_API const char* ql_status_string(int status);

/* QL_OK when the current CUDA device can run the kernels (compute capability 10.x). */
QL_API int ql_device_check(void);

/* Number of CUDA kernels this library has launched in this process (monotonic; for bench.py's
 * gpu_launches and for tests that assert the CUDA path -- not a fallback -- ran). */
QL_API uint64_t ql_launch_count(void);

/* Debug aid (tools/trace_skinny.py): while a device buffer of 16 int64 PER CTA is registered, every CTA of
 * the decode kernel stores clock64() stamps of its pipeline phases into it.  NULL switches it
 * off (the default).  Not part of the reference-facing surface. */
QL_API void ql_debug_set_trace(void* device_buf16_i64);

/*
 * Per-row (per-token) symmetric quantiser.
 * Replaces per_row_quant / _per_row_quant_kernel, nanovllm/utils/quantization.py:13-51,147-163:
 *   amax = max(max_k |x[m,k]|, 1e-8); s[m] = amax/qmax; q = round_half_away(x * (qmax/amax))
 *   (int8) or e4m3_rn(x * (qmax/amax)) (fp8).  Used for activations every forward and for
 * weights once at load (QLinear.__init__, :355).
 */
QL_API int ql_quant_rows(const void* x, int64_t M, int64_t K, int64_t ldx, int fmt,
                  void* q, float* s, ql_stream_t stream);

/*
 * Per-group weight quantiser (load time).
 * Replaces per_group_quant, quantization.py:194-234 (torch semantics: s = amax/qmax,
 * q = clamp(round_half_even(w / s), -128, 127) or e4m3(w / s)).  w is bf16 [N, K] dense,
 * s is [N, K/group].  group == K gives the per-row layout of the reference's CPU path.
 */
QL_API int ql_quant_groups(const void* w, int64_t N, int64_t K, int64_t group, int fmt,
                    void* q, float* s, ql_stream_t stream);

/*
 * Per-tensor quantiser (load time).
 * Replaces per_tensor_quant, quantization.py:166-191: one scale for the whole tensor, kept in
 * the input dtype (a bf16 scalar) and applied with a bf16 division, as torch does there:
 *   scale = bf16(max(amax, 1e-8) / qmax); q = clamp(round_half_even(bf16(x / scale))) or e4m3(.)
 * x is bf16, `numel` contiguous elements; scale_bf16 points at ONE bf16 value; workspace needs
 * >= 4 bytes (the running amax).  Two stream-ordered kernels + one 4-byte memset.
 */
QL_API int ql_quant_tensor(const void* x, int64_t numel, int fmt, void* q, void* scale_bf16,
                    void* workspace, size_t workspace_bytes, ql_stream_t stream);

/*
 * W8A8 GEMM on tcgen05 tensor cores with the scale/bias epilogue fused.
 * group == 0: replaces _per_row_post_process_mm, quantization.py:54-144 (launched :330-345):
 *   y[m,n] = bf16( float(sum_k qx[m,k]*qw[n,k]) * sx[m] * sw[n] + bias[n] )
 * group > 0 (multiple of 128): the per-group W8A8 form that replaces per_group_matmul,
 * quantization.py:265-307, on the GPU:
 *   y[m,n] = bf16( sx[m] * sum_j sw[n,j] * float(sum_{k in block j} qx*qw) + bias[n] )
 * Requires K % 128 == 0 (as the reference, :328).
 */
QL_API int ql_gemm_w8a8(const void* qx, const float* sx, const void* qw, const float* sw,
                 const void* bias, void* y, int64_t ldy, int64_t M, int64_t N, int64_t K,
                 int fmt, int64_t group, ql_stream_t stream);

/*
 * Parity entry point: raw int32 accumulators acc[m,n] = sum_k qx[m,k]*qw[n,k] of the same
 * tcgen05 INT8 kernel (dense [M,N] int32).  Mirrors the `acc` tile of the reference kernel,
 * quantization.py:96-106, which a bf16 output cannot prove equal.
 */
QL_API int ql_gemm_i8_acc(const void* qx, const void* qw, int32_t* acc, int64_t M, int64_t N,
                   int64_t K, ql_stream_t stream);

/*
 * Decode-regime fused quantise + GEMV (M <= 16), one launch per linear.
 * Replaces the per_row_quant + _per_row_post_process_mm pair of per_row_matmul,
 * quantization.py:310-346, for small M (and the per-group form, as ql_gemm_w8a8).
 * group == 0 or a multiple of 32 that divides K.  Same arithmetic as ql_quant_rows followed
 * by ql_gemm_w8a8 (bit-identical for INT8 per-row).
 */
QL_API int ql_gemv_w8a8(const void* x, int64_t ldx, const void* qw, const float* sw, const void* bias,
                 void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int fmt, int64_t group,
                 ql_stream_t stream);

/*
 * One quantized linear, dispatched by M: the whole of per_row_matmul (quantization.py:310-346)
 * / the W8A8 replacement of per_group_matmul (:265-307) behind QLinear.forward (:363-364) and
 * QGroupLinear.forward (:400-401).  M <= 16 -> ql_gemv_w8a8; otherwise ql_quant_rows into
 * `workspace` followed by ql_gemm_w8a8.  workspace must hold ql_workspace_bytes(M, K) bytes
 * (may be NULL when that is 0).
 */
QL_API int ql_linear(const void* x, int64_t ldx, const void* qw, const float* sw, const void* bias,
              void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int fmt, int64_t group,
              void* workspace, size_t workspace_bytes, ql_stream_t stream);

/*
 * RMSNorm (optionally with the residual add) fused with the per-token quantiser: the producer of
 * qkv_proj's and gate_up_proj's input.
 * Replaces RMSNorm.rms_forward / add_rms_forward, nanovllm/layers/layernorm.py:16-40 (called at
 * nanovllm/models/qwen3.py:156-165) followed by per_row_quant (quantization.py:147-163) inside
 * per_row_matmul (:325):
 *   xf = float(x) [+ float(residual_in)];  residual_out = bf16(xf)
 *   h  = bf16( float(bf16(xf * rsqrt(mean(xf^2) + eps))) * float(weight) )
 *   (q, s) = ql_quant_rows(h)
 * x bf16 [M, K] with row stride ldx; residual_in / residual_out bf16 [M, K] dense, both NULL or
 * both set (they may alias: the engine updates the residual stream in place); weight bf16 [K];
 * q [M, K] codes + s [M] (q == NULL: no quantised output); h_out bf16 [M, K] dense or NULL (the
 * bf16 result is only materialised on request).  K % 8 == 0, K <= 51200.  The pair (q, s) feeds
 * ql_gemm_w8a8 directly.
 */
QL_API int ql_rmsnorm_quant(const void* x, int64_t ldx, const void* residual_in, void* residual_out,
                     const void* weight, float eps, int64_t M, int64_t K, int fmt, void* q,
                     float* s, void* h_out, ql_stream_t stream);

/*
 * SiLU(gate) * up fused with the per-token quantiser: the producer of down_proj's input.
 * Replaces SiluAndMul.forward, nanovllm/layers/activation.py:11-14 (called at
 * nanovllm/models/qwen3.py:117) followed by per_row_quant inside per_row_matmul:
 *   h = bf16( float(bf16(silu(float(gate)))) * float(up) );  (q, s) = ql_quant_rows(h)
 * gate_up bf16 [M, 2*I] with row stride ld (gate = columns [0, I), up = [I, 2I), the layout
 * MergedColumnParallelLinear produces, linear.py:76-93); q [M, I], s [M], h_out bf16 [M, I] dense
 * or NULL; q == NULL: bf16 output only.  I % 8 == 0, I <= 102400.
 */
QL_API int ql_silu_mul_quant(const void* gate_up, int64_t ld, int64_t M, int64_t I, int fmt, void* q,
                      float* s, void* h_out, ql_stream_t stream);

/*
 * One-shot all-reduce over NVLink peer memory for RowParallelLinear's decode-sized outputs.
 * Replaces `dist.all_reduce(y)` of RowParallelLinear.forward, nanovllm/layers/linear.py:150-152,
 * when y ([M, hidden] bf16) is at most max_bytes; larger outputs stay on NCCL.  One communicator
 * per process (the engine runs one process per GPU, model_runner.py:26-29).
 *   ql_ar_can_access_peers(world): QL_OK when the current device can map the memory of devices
 *     0..world-1 (one process per GPU, rank == device ordinal); call before ql_ar_init and fall back
 *     to NCCL on every rank if any rank says no;
 *   ql_ar_init(max_bytes, rs_tokens, rs_hidden, world, rank, handle_out): cudaMalloc of this rank's
 *     symmetric area -- 2 parities x world regions x 2*max_bytes for the one-shot all-reduce, plus,
 *     when rs_tokens > 0, the buffers of the large path (ql_tp_*) for [rs_tokens, rs_hidden]
 *     outputs -- and writes its CUDA IPC handle (ql_ar_handle_bytes() bytes) to HOST memory;
 *   the caller all-gathers the handles (torch.distributed) and passes the concatenation, in rank
 *     order, to ql_ar_open (HOST pointer), which maps every peer's area;
 *   ql_ar_destroy(): unmaps the peers and frees the area (failed set-up, shutdown); idempotent;
 *   ql_ar_allreduce_bf16(in, out, numel, stream): out[i] = bf16(sum over ranks r = 0..W-1 of
 *     in_r[i], fp32, rank order); in/out are ordinary local device tensors (may alias); numel % 8
 *     == 0, numel * 2 <= ql_ar_max_bytes(); stream-ordered, no host synchronisation, CUDA-graph
 *     safe; every rank must issue the same sequence of calls.
 * world in {2, 4, 8}.  Flag waits are bounded by QL_AR_TIMEOUT_S seconds (default 10, 0 = wait
 * forever): past it the kernel traps (CUDA error) instead of hanging on a dead peer.  These kernels
 * cannot be profiled with ncu kernel replay (a replayed launch never sees matching epochs).
 */
QL_API size_t ql_ar_handle_bytes(void);
QL_API int ql_ar_can_access_peers(int world);
QL_API int ql_ar_init(size_t max_bytes, size_t rs_tokens, size_t rs_hidden, int world, int rank,
               void* handle_out);
QL_API int ql_ar_open(const void* all_handles);
QL_API int ql_ar_destroy(void);
QL_API size_t ql_ar_max_bytes(void);
QL_API int ql_ar_allreduce_bf16(const void* in, void* out, int64_t numel, ql_stream_t stream);

/*
 * All-reduce fused with the residual add, RMSNorm and the per-token quantiser: everything between
 * a row-parallel linear and the next column-parallel one in TP decode, in one kernel.
 * Replaces dist.all_reduce(y) (nanovllm/layers/linear.py:150-152) + RMSNorm.add_rms_forward
 * (nanovllm/layers/layernorm.py:27-40, called at models/qwen3.py:156-165) + per_row_quant
 * (utils/quantization.py:147-163):
 *   y = bf16(sum over ranks of partial_r, fp32, rank order)      == ql_ar_allreduce_bf16
 *   (residual_out, q, s) = ql_rmsnorm_quant(y, residual_in, weight, eps)   (bit-identical)
 * partial: this rank's un-reduced row-parallel output, bf16 [M, K] dense (K = hidden size, K % 8
 * == 0, K <= 8192); M * K * 2 <= ql_ar_max_bytes().  Same calling rules as ql_ar_allreduce_bf16.
 */
QL_API int ql_ar_rmsnorm_quant(const void* partial, const void* residual_in, void* residual_out,
                        const void* weight, float eps, int64_t M, int64_t K, int fmt, void* q,
                        float* s, ql_stream_t stream);

/*
 * Prefill-sized row-parallel outputs under tensor parallelism: reduce-scatter + per-token quantiser
 * + all-gather over NVLink peer memory (csrc/tp_collective.cu, csrc/gemm_2cta.cu).
 * Replaces `dist.all_reduce(y)` of RowParallelLinear.forward (nanovllm/layers/linear.py:150-152)
 * and the `per_row_quant(x)` of the NEXT quantized linear (utils/quantization.py:147-163, :323):
 * tokens are owned by ranks in slices of whole 256-token tiles; the owner adds the W bf16 partials
 * in fp32 in rank order, rounds once to bf16, quantises its tokens and sends codes + scales to
 * every rank, whose next column-parallel GEMM reads them through ql_gemm_w8a8.
 *   ql_tp_max_tokens / ql_tp_max_hidden: capacity given to ql_ar_init (0 = path unavailable);
 *   ql_tp_buffer(which, channel): this rank's buffers inside the symmetric area (device pointers that stay
 *     valid until ql_ar_destroy): 0 = bf16 partial [M, N] dense (write the row-parallel GEMM's
 *     output here for pushed = 0), 1 = gathered codes [M, N], 2 = gathered scales fp32 [M];
 *   ql_tp_gemm_push: the row-parallel W8A8 GEMM (per-row scales, as ql_gemm_w8a8) whose epilogue
 *     TMA-stores every 256-token tile into the receive region of the rank that owns those tokens --
 *     GEMM and reduce-scatter in ONE kernel, the transfer overlapping the remaining tiles;
 *   ql_tp_reduce_quant_gather(M, N, fmt, pushed): after it (stream-ordered) buffers 1 and 2 hold
 *     ql_quant_rows(bf16(sum of the ranks' partials)) for all M tokens on every rank, bit-identical
 *     to all-reduce (fp32, rank order) followed by ql_quant_rows.  pushed = 1: partials came from
 *     ql_tp_gemm_push; pushed = 0: every rank left its partial in buffer 0 and the owners pull;
 *   ql_tp_allreduce_bf16(M, N, pushed, out, ldo): same reduction, bf16 rows gathered instead; the
 *     complete [M, N] result is written to the local tensor `out` (dist.all_reduce semantics).
 * M <= ql_tp_max_tokens(), N <= ql_tp_max_hidden(), N % 8 == 0.  Same calling rules as
 * ql_ar_allreduce_bf16 (every rank, same order, same shapes; CUDA-graph safe) -- PER CHANNEL:
 * `channel` (0 .. ql_tp_channels()-1) selects one of the independent sets of buffers, flags and
 * call counters.  A caller that splits a prefill batch into two token micro-batches and runs them
 * on two streams, one channel each, hides the wire time, flag latency and write acknowledgements
 * of one micro-batch's collective under the other's GEMMs: the reduce kernel is one 128-thread CTA
 * per SM and stays co-resident with a CTA-pair GEMM (per-row / per-tensor scales; do not use two
 * channels concurrently with per-group linears, whose GEMM does not leave room for it).
 */
/*
 * Test hook: a W-rank communicator emulated on ONE GPU (the driver's GPU test tier has a single
 * GPU).  ql_ar_init_local allocates all W symmetric areas in this process; ql_tp_emu_set_rank(r)
 * makes the following ql_tp_gemm_push / ql_tp_buffer calls act as rank r; ql_tp_reduce_quant_gather /
 * ql_tp_allreduce_bf16 then run ALL ranks in one co-resident launch (for the bf16 form `out` holds W
 * stacked [M, ldo] tensors).  Same kernels, same flags, same tensor maps as the multi-GPU path --
 * only the addresses are local.  Not part of the reference-facing surface.
 */
QL_API int ql_ar_init_local(size_t max_bytes, size_t rs_tokens, size_t rs_hidden, int world);
QL_API int ql_tp_emu_set_rank(int rank);

QL_API size_t ql_tp_max_tokens(void);
QL_API size_t ql_tp_max_hidden(void);
/*
 *   ql_tp_set_concurrent(on): declare that two channels will be in flight at the same time (also
 *     implied by the first call on a channel > 0).  The reduce kernels then release their stream's
 *     next kernel only after their last cross-GPU wait: a GEMM parked behind a reduce kernel that
 *     waits for a peer would otherwise hold the shared memory the other channel's GEMM -- the one
 *     the peer is waiting for -- needs.  Call it before capturing CUDA graphs of channel 0.
 *
 * KNOWN HAZARD -- two channels in flight at once can hang.  One TP=8 run ended in the reduce
 * kernel's bounded-wait trap (profiles/r2n_n8_ag_push_micro2_trap.err).  Only the reduce kernel's
 * GRID is capped at one CTA per SM; the scheduler may still stack its spinning CTAs on one SM, the
 * other channel's persistent GEMM CTA then has no SM to run its statically assigned tiles on, and
 * the peer waits for those tiles: a cross-GPU cycle.  Until this is fixed, every call with
 * channel > 0 and ql_tp_set_concurrent(1) return QL_ERR_GUARDED unless the environment variable
 * QL_TP_ALLOW_CONCURRENT=1 is set (read at each call).  Channel 0 alone is unaffected.
 *
 *   ql_tp_channels(): channels the symmetric area holds -- 1 by default; 2 only when
 *     QL_TP_ALLOW_CONCURRENT=1 was set when ql_ar_init / ql_ar_init_local ran (channel 1's buffers
 *     are as large as channel 0's, ~1.4 GB at 65536 tokens x 4096, and would otherwise come out of
 *     the KV-cache budget for nothing).  A channel that was not allocated gets QL_ERR_WORKSPACE.
 *     Every rank must run with the same setting.
 */
QL_API int ql_tp_channels(void);
QL_API int ql_tp_set_concurrent(int on);
QL_API void* ql_tp_buffer(int which, int channel);
QL_API int ql_tp_gemm_push(const void* qx, const float* sx, const void* qw, const float* sw,
                    const void* bias, int64_t M, int64_t N, int64_t K, int fmt, int channel,
                    ql_stream_t stream);
QL_API int ql_tp_reduce_quant_gather(int64_t M, int64_t N, int fmt, int pushed, int channel,
                              ql_stream_t stream);
QL_API int ql_tp_allreduce_bf16(int64_t M, int64_t N, int pushed, void* out, int64_t ldo, int channel,
                         ql_stream_t stream);

/* Bytes of caller-provided scratch ql_linear needs for an [M, K] activation (0 for M <= 16). */
QL_API size_t ql_workspace_bytes(int64_t M, int64_t K);

/*
 * ql_linear without a workspace argument, for callers that cannot allocate one.
 * (No counterpart in the reference: there every forward allocates through torch,
 * nanovllm/utils/quantization.py:320-329.  This entry point exists for users of the C ABI.)
 *
 * Who needs it: a client that binds this library from C, ctypes or another FFI without torch has
 * no caching allocator, only cudaMalloc -- and once the serving engine has given all remaining
 * HBM to its KV cache, cudaMalloc fails, or at best stalls every stream of the device while it
 * runs.  ql_linear_static therefore takes the same inputs and outputs as ql_linear and keeps the
 * quantised activations (qx [M,K], sx [M]) in an array that is part of the library image:
 * QL_STATIC_SCRATCH_BYTES (8 MiB) of device memory on EVERY GPU that loads the library, held
 * until the process exits.  It calls no cudaMalloc / cudaFree and no allocating library routine.
 *
 * Limit: ql_workspace_bytes(M, K) <= QL_STATIC_SCRATCH_BYTES (also returned by
 * ql_static_scratch_bytes()), e.g. 4096 tokens at K = 1024 or 2048 tokens at K = 3072.  Anything
 * larger is refused with QL_ERR_STATIC_SCRATCH and nothing is launched -- split the tokens into
 * several calls, or use ql_linear with your own workspace.
 *
 * Results are those of ql_linear on the same arguments, bit for bit (same kernels, same dispatch).
 *
 * Rules that come with a shared, never-freed scratch:
 *   - the first call on each device looks the array's device address up (cudaGetSymbolAddress,
 *     with that device current) and caches it; make that first call before any stream capture;
 *   - all calls on a device use the same array.  Calls issued to ONE stream are safe back to
 *     back; calls that can overlap on the GPU (different streams, or several host threads
 *     feeding different streams) race on it and corrupt each other's activations -- order
 *     them with events or use ql_linear with separate workspaces;
 *   - the array keeps the previous call's bytes.  Nothing in it has to be zero: every call
 *     rewrites all of qx and sx it later reads, so no clearing happens and none is needed.
 */
#define QL_STATIC_SCRATCH_BYTES (8u << 20)
QL_API size_t ql_static_scratch_bytes(void);
QL_API int ql_linear_static(const void* x, int64_t ldx, const void* qw, const float* sw,
                     const void* bias, void* y, int64_t ldy, int64_t M, int64_t N, int64_t K,
                     int fmt, int64_t group, ql_stream_t stream);

/*
 * Decode-regime CHAIN of quantized linears in one persistent launch (csrc/decode_chain.cu).
 * Replaces, for M <= 16 tokens, the per-linear per_row_quant + per_row_matmul launches of the
 * reference's decode step (nanovllm/utils/quantization.py:310-346, called from the four linears of
 * every decoder layer, nanovllm/models/qwen3.py:80,90,116,118) wherever one quantized linear feeds
 * the next: op i with x == NULL consumes the first K_i columns of op i-1's output y (row pitch
 * ldy of that op); an op with x != NULL reads its own bf16 input [M, ldx] (the first op must).
 * Every op is y = bf16((acc * sx[m]) * sw[n] + bias[n]) on per-token-quantised x, per-row weight
 * scales, exactly as ql_linear (INT8: bit-identical; FP8: fp32 summation order differs).
 *   ql_chain_workspace_bytes: device scratch the chain needs (0 = the chain is not supported:
 *     M > 16, K % 128, N % 8, K > 12288 ...).  Two kernels: a tcgen05 consumer when every
 *     K % 512 == 0, an mma.sync consumer otherwise (QL_CHAIN_TC=0 forces the latter);
 *   ql_chain_prepare: plans the chain for the current device (clusters of 4 CTAs, as many as
 *     cudaOccupancyMaxActiveClusters reports co-resident), encodes one TMA descriptor per op,
 *     writes descriptors and plan tables into `workspace` (256-byte aligned, stays owned by the
 *     caller, must outlive every run) and fills `plan`; synchronises `stream`; NOT capturable.
 *     The op array, M, every pointer and pitch are frozen into the workspace: prepare again when
 *     any of them changes;
 *   ql_chain_run: ONE kernel launch (cooperative + cluster launch, one CTA per SM, all
 *     co-resident; every CTA spins on a bounded grid barrier between ops); capturable into a CUDA
 *     graph.  Runs of one workspace must not overlap each other.  If the driver refuses the
 *     cooperative launch the call fails with QL_ERR_CUDA and launches nothing: without the
 *     residency guarantee another stream holding SMs would make the grid barrier spin into its
 *     bounded-wait trap.  QL_CHAIN_ALLOW_NONCOOP=1 falls back to a plain cluster launch instead
 *     (only for a caller that keeps the device otherwise idle).
 * Measured on B200: correct but SLOWER than one ql_linear per op (a
 * dependent op costs ~5 us of release / barrier / activation round trips either way); kept as an
 * experiment that bench.py reports next to the per-linear launches.
 * group must be 0 (per-group scales: use ql_linear).
 */
typedef struct ql_chain_op {
  const void* qw;     /* [N, K] codes, row-major, 16-byte aligned */
  const float* sw;    /* [N] weight scales */
  const void* bias;   /* bf16 [N] or NULL */
  const void* x;      /* bf16 [M, ldx], or NULL = previous op's y */
  void* y;            /* bf16 [M, ldy] */
  int64_t ldx, ldy;   /* row pitches in elements (ldx ignored when x == NULL) */
  int64_t N, K;
} ql_chain_op;
typedef struct ql_chain_plan {
  unsigned char opaque[128];
} ql_chain_plan;
QL_API size_t ql_chain_workspace_bytes(const ql_chain_op* ops, int n_ops, int64_t M);
QL_API int ql_chain_prepare(const ql_chain_op* ops, int n_ops, int64_t M, int fmt, int64_t group,
                            void* workspace, size_t workspace_bytes, ql_chain_plan* plan, ql_stream_t stream);
QL_API int ql_chain_run(const ql_chain_plan* plan, void* workspace, ql_stream_t stream);

/* Name of the kernel family ql_linear would pick ("gemv_mma", "gemm_tcgen05"); for logs/tests. */
QL_API const char* ql_linear_path(int64_t M, int64_t N, int64_t K, int64_t group);

#ifdef __cplusplus
}
#endif
#endif /* QLINEAR_B200_H_ */
