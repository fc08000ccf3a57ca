// Reduce-scatter + per-token quantiser + all-gather over NVLink peer memory: what follows a
// PREFILL-sized row-parallel linear under tensor parallelism.
//
// Replaces `dist.all_reduce(y)` of RowParallelLinear.forward (nanovllm/layers/linear.py:150-152)
// together with the next quantized linear's `per_row_quant(x)` (utils/quantization.py:147-163,
// called from per_row_matmul :323).  NCCL's ring all-reduce moves 2 * 2 B * (W-1)/W per element
// and every rank then quantises all W*T tokens again; here
//   1. every rank's bf16 partial sums reach the rank that OWNS a token (reduce-scatter): either
//      pushed tile by tile from the GEMM epilogue (gemm2_push_kernel, TMA stores into the owner's
//      receive region while the tensor core works on the next tile) or pulled by the owner;
//   2. the owner adds the W partials in fp32 in rank order, rounds once to bf16 -- identical on
//      every rank, deterministic, more accurate than NCCL's bf16 ring -- and quantises ITS tokens
//      only (the Triton quantiser's arithmetic, bit for bit, as quant_rows_kernel);
//   3. codes (1 B) + scales travel to every peer (all-gather by posted writes), so the following
//      column-parallel GEMM consumes (qx, sx) directly: 3 B * (W-1)/W per element on the wire.
// The bf16 form (fmt < 0) all-gathers the rounded bf16 rows instead and leaves the complete
// [M, N] tensor in `out`: `dist.all_reduce` semantics for callers that need the tensor itself.
//
// Synchronisation: two flag arrays per rank in its symmetric area, written with st.release.sys by
// the peers and polled locally with ld.acquire.sys: ready[src] = "src's partials for call e are
// in place", done[src] = "src has delivered everything it owns for call e".  The call number e
// is a device-side counter advanced by the last block, so the kernels replay inside CUDA graphs.
// Buffers are single per CHANNEL: a rank can start call e+1's pushes only after its own call-e
// kernel saw every done flag, i.e. after every peer finished reading call e's partials.  Every
// rank must issue the same sequence of calls with the same shapes on a channel.  The kTpChannels
// channels are independent (own buffers, flags, counters): two token micro-batches on two streams
// keep one collective each in flight, and because this kernel is launched with at most one small
// CTA per SM it stays co-resident with the other channel's GEMM -- flag latencies, write
// acknowledgements and wire time of one micro-batch hide under the tensor-core work of the other.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace ql {

constexpr int QL_FMT_BF16 = -1;

struct TpArgs {
  uint8_t* big[8];
  TpLayout lay;
  int64_t M, Mo;         // tokens in total / per owner
  int N;
  int rank, pushed;      // rank < 0: single-GPU emulation, blockIdx.y is the rank (see ql_ar_init_local)
  uint16_t* out;         // bf16 form: complete result, row stride ldo
  int64_t ldo;
  int64_t out_rank_stride;   // emulation: elements between the ranks' `out` tensors
  uint64_t timeout_ns;
  int late_pdl;          // channels run concurrently: release the stream's next kernel only after the cross-GPU waits
  int ag_pull;           // all-gather of codes + scales by remote LOADS after the done flags (else posted stores)
};

template <int FMT> struct TpQ { static constexpr float qmax = 1.0f; };
template <> struct TpQ<QL_INT8> { static constexpr float qmax = 127.0f; };
template <> struct TpQ<QL_FP8_E4M3> { static constexpr float qmax = 448.0f; };

template <int WPR> struct TpRowCfg {
  static constexpr int threads = WPR * 32 > 128 ? WPR * 32 : 128;
  static constexpr int rows = threads / 32 / WPR;
};

__device__ __forceinline__ uint32_t tp_pack_bf(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void tp_unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}

__device__ __forceinline__ uint32_t* tp_flag(uint8_t* big, size_t base, int src) {
  return reinterpret_cast<uint32_t*>(big + base + (size_t)src * kTpFlagStride);
}

// wait until the flags of all W ranks (except `self`) in MY area have reached call e
template <int W>
__device__ __forceinline__ void tp_wait_flags(uint8_t* my_big, size_t base, int self, uint32_t e,
                                              uint64_t timeout_ns) {
  if (threadIdx.x < W && (int)threadIdx.x != self) {
    const uint32_t* f = tp_flag(my_big, base, (int)threadIdx.x);
    SpinGuard guard(timeout_ns);
    while ((int32_t)(ld_acquire_sys(f) - e) < 0) guard.tick();
  }
  __syncthreads();
}

template <int FMT, int W, int NV, int WPR>
// <= 128 registers x 128 threads: fits next to a 384-thread, 128-register GEMM CTA on one SM
__global__ void __launch_bounds__(TpRowCfg<WPR>::threads, 4)
tp_reduce_kernel(const TpArgs a) {
  __shared__ float red[8];
  // With one channel the next kernel of the stream may start (and park at its griddepcontrol.wait)
  // right away.  With two channels on two streams it must not: a parked GEMM holds its SM's shared
  // memory while this kernel waits for a PEER's GEMM -- if the peer is in the mirrored state on the
  // other channel, neither rank can place the GEMM the other one waits for.  Then the dependents
  // are released after the last cross-GPU wait instead.
  if (!a.late_pdl) pdl_launch_dependents();
  pdl_wait();                       // this rank's GEMM (its partial sums / its pushes) is complete
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int ROWS = TpRowCfg<WPR>::rows;
  const int rank = a.rank >= 0 ? a.rank : (int)blockIdx.y;
  uint16_t* const out = a.out + (a.rank >= 0 ? 0 : (int64_t)blockIdx.y * a.out_rank_stride);
  uint8_t* mine = a.big[rank];
  uint32_t* ctr = reinterpret_cast<uint32_t*>(mine + a.lay.flags + kTpCtrOff);
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(ctr) + 1;
  const int nvec = a.N >> 3;

  // ---- 1. the partials of call e are in place on every rank ---------------------------------
  // (the GEMM grid in front of this kernel has completed, and with it its stores -- the bulk
  //  stores into the owners' receive regions in the pushed form)
  if (blockIdx.x == 0 && threadIdx.x < W && (int)threadIdx.x != rank) {
    __threadfence_system();
    st_release_sys(tp_flag(a.big[threadIdx.x], a.lay.flags + kTpReadyOff, rank), e);
  }
  tp_wait_flags<W>(mine, a.lay.flags + kTpReadyOff, rank, e, a.timeout_ns);

  // ---- 2. my token slice: sum, round, quantise, deliver --------------------------------------
  const int64_t row0 = (int64_t)rank * a.Mo;
  const int64_t mine_rows = a.M - row0 < 0 ? 0 : (a.M - row0 < a.Mo ? a.M - row0 : a.Mo);
  for (int64_t rb = (int64_t)blockIdx.x * ROWS; rb < mine_rows; rb += (int64_t)gridDim.x * ROWS) {
    const int64_t lr = rb + warp / WPR;
    const int part = warp % WPR;
    const bool live = lr < mine_rows;
    const int64_t lrc = live ? lr : 0;
    const int64_t grow = row0 + lrc;             // global token index
    float f[NV][8];
#pragma unroll
    for (int r = 0; r < W; ++r) {                // fp32 sum in rank order (as ar_oneshot_kernel)
      const uint8_t* src = a.pushed
          ? mine + a.lay.part + (size_t)((int64_t)r * a.Mo + lrc) * (size_t)a.N * 2
          : a.big[r] + a.lay.part + (size_t)grow * (size_t)a.N * 2;
      uint4 v[NV];
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int idx = (i * WPR + part) * 32 + lane;
        v[i] = (live && idx < nvec) ? ld_sys128(src + (size_t)idx * 16) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        float g[8];
        tp_unpack8(v[i], g);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[i][k] = r == 0 ? g[k] : __fadd_rn(f[i][k], g[k]);
      }
    }
    uint4 hv[NV];                                // y = bf16(sum): what dist.all_reduce leaves in y
    float amax = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      hv[i] = make_uint4(tp_pack_bf(f[i][0], f[i][1]), tp_pack_bf(f[i][2], f[i][3]),
                         tp_pack_bf(f[i][4], f[i][5]), tp_pack_bf(f[i][6], f[i][7]));
      amax = fmaxf(amax, amax8(hv[i]));
    }
    if constexpr (FMT == QL_FMT_BF16) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int idx = (i * WPR + part) * 32 + lane;
        if (!(live && idx < nvec)) continue;
        *reinterpret_cast<uint4*>(out + grow * a.ldo + (int64_t)idx * 8) = hv[i];
#pragma unroll
        for (int d = 1; d < W; ++d) {
          uint8_t* dst = a.big[(rank + d) % W] + a.lay.y + (size_t)grow * (size_t)a.N * 2;
          *reinterpret_cast<uint4*>(dst + (size_t)idx * 16) = hv[i];
        }
      }
    } else {
      constexpr float qmax = TpQ<FMT>::qmax;
      amax = warp_max(amax);
      if constexpr (WPR > 1) {
        __syncthreads();
        if (lane == 0) red[warp] = amax;
        __syncthreads();
        const int w0 = (warp / WPR) * WPR;
        amax = red[w0];
#pragma unroll
        for (int i = 1; i < WPR; ++i) amax = fmaxf(amax, red[w0 + i]);
      }
      amax = fmaxf(amax, 1e-8f);
      const float inv = __fdiv_rn(qmax, amax);
      const float sc = __fdiv_rn(amax, qmax);
      uint2 qv[NV];
#pragma unroll
      for (int i = 0; i < NV; ++i) qv[i] = quant8_mul<FMT>(hv[i], inv);
      // all-gather: posted stores to every peer as the codes are produced, or (ag_pull) the own copy
      // only -- the peers fetch it with loads after the done flags
      const int ndst = a.ag_pull ? 1 : W;
#pragma unroll
      for (int d = 0; d < W; ++d) {              // own copy first, then the peers, staggered by rank
        if (d >= ndst) break;
        uint8_t* dbig = a.big[(rank + d) % W];
        uint2* qr = reinterpret_cast<uint2*>(dbig + a.lay.q + (size_t)grow * (size_t)a.N);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int idx = (i * WPR + part) * 32 + lane;
          if (live && idx < nvec) qr[idx] = qv[i];
        }
        if (live && part == 0 && lane == 0)
          reinterpret_cast<float*>(dbig + a.lay.s)[grow] = sc;
      }
    }
  }

  // ---- 3. everything I own has been delivered: tell the peers, wait for theirs --------------
  __syncthreads();
  __shared__ uint32_t last_s;
  if (threadIdx.x == 0) {
    __threadfence_system();
    last_s = atomicAdd(ctr + 1, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last_s) {
    if (threadIdx.x < W && (int)threadIdx.x != rank) {
      __threadfence_system();
      st_release_sys(tp_flag(a.big[threadIdx.x], a.lay.flags + kTpDoneOff, rank), e);
    }
  }
  tp_wait_flags<W>(mine, a.lay.flags + kTpDoneOff, rank, e, a.timeout_ns);
  if (a.late_pdl) pdl_launch_dependents();

  if constexpr (FMT != QL_FMT_BF16) {
    if (a.ag_pull) {
      // ---- 3b. fetch the other owners' codes and scales: 8 x 16-byte peer loads in flight per thread.
      // (An owner rewrites its slice only in its next call, behind that call's ready flags, i.e.
      //  after every peer has finished this kernel.)
      constexpr int U = 8;
      const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
      const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
      for (int d = 1; d < W; ++d) {
        const int o = (rank + d) % W;
        const int64_t r0 = (int64_t)o * a.Mo;
        const int64_t rows_o = a.M - r0 < 0 ? 0 : (a.M - r0 < a.Mo ? a.M - r0 : a.Mo);
        const int64_t vecs = rows_o * (int64_t)a.N / 16;            // N % 16 == 0 (checked by the host)
        const uint8_t* src = a.big[o] + a.lay.q + (size_t)r0 * (size_t)a.N;
        uint8_t* dst = mine + a.lay.q + (size_t)r0 * (size_t)a.N;
        for (int64_t i = tid; i < vecs; i += nthr * U) {
          uint4 v[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int64_t j = i + (int64_t)u * nthr;
            v[u] = j < vecs ? ld_sys128(src + (size_t)j * 16) : make_uint4(0, 0, 0, 0);
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int64_t j = i + (int64_t)u * nthr;
            if (j < vecs) *reinterpret_cast<uint4*>(dst + (size_t)j * 16) = v[u];
          }
        }
        const float* ssrc = reinterpret_cast<const float*>(a.big[o] + a.lay.s) + r0;
        float* sdst = reinterpret_cast<float*>(mine + a.lay.s) + r0;
        for (int64_t i = tid; i < rows_o; i += nthr)
          sdst[i] = __uint_as_float(ld_acquire_sys(reinterpret_cast<const uint32_t*>(ssrc + i)));
      }
    }
  }

  if constexpr (FMT == QL_FMT_BF16) {
    // the peers' rows sit in my gather buffer: copy them next to my own rows in `out`
    const int64_t total_vec = a.M * (int64_t)nvec;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec;
         i += (int64_t)gridDim.x * blockDim.x) {
      const int64_t row = i / nvec;
      const int idx = (int)(i - row * nvec);
      if (row >= row0 && row < row0 + a.Mo) continue;
      const uint4 v = ld_sys128(mine + a.lay.y + (size_t)row * (size_t)a.N * 2 + (size_t)idx * 16);
      *reinterpret_cast<uint4*>(out + row * a.ldo + (int64_t)idx * 8) = v;
    }
  }

  // ---- 4. the last block to leave advances the call counter ------------------------------------
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(ctr + 3, 1u) == gridDim.x - 1) {
      ctr[1] = 0;
      ctr[3] = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(ctr) = e;
    }
  }
}

// ---------------------------------------------------------------------------------- host side

static bool tp_pick_cfg(int nvec, int& nv, int& wpr) {
  static const int cfgs[][2] = {{4, 1}, {4, 2}, {4, 4}, {8, 4}};
  for (const auto& c : cfgs)
    if (nvec <= c[0] * c[1] * 32) {
      nv = c[0]; wpr = c[1];
      return true;
    }
  return false;
}

template <int FMT, int W, int NV, int WPR>
static int launch_tp_cfg(const TpArgs& a, cudaStream_t st) {
  auto kern = tp_reduce_kernel<FMT, W, NV, WPR>;
  constexpr int threads = TpRowCfg<WPR>::threads;
  // All blocks wait on cross-GPU flags, so the grid must be co-resident; and it must stay small
  // enough to share the SMs with a GEMM of the other channel (one 384-thread CTA per SM holding
  // ~200 KB of shared memory and most of the register file): QL_TP_REDUCE_CTAS CTAs per SM,
  // default 1.
  static int per_sm = 0;
  if (per_sm == 0) {
    const char* e = getenv("QL_TP_REDUCE_CTAS");
    int n = e ? atoi(e) : 1;
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, 0) != cudaSuccess || occ < 1) {
      (void)cudaGetLastError();
      occ = 1;
    }
    per_sm = n < 1 ? 1 : (n > occ ? occ : n);
  }
  const int64_t want = (a.Mo + TpRowCfg<WPR>::rows - 1) / TpRowCfg<WPR>::rows;
  int64_t blocks = (int64_t)num_sms() * per_sm / (a.rank < 0 ? W : 1);   // emulation: W ranks share the GPU
  if (blocks > want) blocks = want;
  if (blocks < 1) blocks = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)blocks, a.rank < 0 ? W : 1);
  cfg.blockDim = dim3(threads);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, kern, a);
  count_launch();
  if (err != cudaSuccess) {
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  return QL_OK;
}

template <int FMT, int W>
static int launch_tp_w(const TpArgs& a, cudaStream_t st) {
  int nv = 0, wpr = 0;
  if (!tp_pick_cfg(a.N / 8, nv, wpr)) return QL_ERR_BAD_SHAPE;
  if (nv == 4 && wpr == 1) return launch_tp_cfg<FMT, W, 4, 1>(a, st);
  if (nv == 4 && wpr == 2) return launch_tp_cfg<FMT, W, 4, 2>(a, st);
  if (nv == 4 && wpr == 4) return launch_tp_cfg<FMT, W, 4, 4>(a, st);
  return launch_tp_cfg<FMT, W, 8, 4>(a, st);
}

template <int FMT>
static int launch_tp_fmt(const TpArgs& a, int world, cudaStream_t st) {
  switch (world) {
    case 2: return launch_tp_w<FMT, 2>(a, st);
    case 4: return launch_tp_w<FMT, 4>(a, st);
    case 8: return launch_tp_w<FMT, 8>(a, st);
    default: return QL_ERR_BAD_SHAPE;
  }
}

// set once two channels are in use (ql_tp_set_concurrent, or any call on a channel > 0)
static int g_tp_concurrent = 0;

int launch_tp_reduce(const TpView& tp, int64_t M, int64_t N, int fmt, int pushed, void* out, int64_t ldo,
                     cudaStream_t st) {
  TpArgs a;
  a.late_pdl = __atomic_load_n(&g_tp_concurrent, __ATOMIC_RELAXED);
  for (int r = 0; r < 8; ++r) a.big[r] = tp.big[r];
  a.lay = tp.lay;
  a.M = M;
  a.Mo = tp_rows_per_owner(M, tp.world);
  a.N = (int)N;
  a.rank = tp.emulated ? -1 : tp.rank;
  a.pushed = pushed;
  a.out = static_cast<uint16_t*>(out);
  a.ldo = ldo;
  a.out_rank_stride = M * ldo;
  a.timeout_ns = tp.timeout_ns;
  {
    // QL_TP_AG_PULL=1 (read per launch: tests flip it): all-gather by peer loads.  Off by default:
    // on 8 x B200 it measured 11.8 ms / step against 10.6 ms with posted stores (profiles/r2n_*);
    // N % 16 required
    const char* e = getenv("QL_TP_AG_PULL");
    a.ag_pull = (N % 16 == 0) && e != nullptr && e[0] == '1';
  }
  if (fmt == QL_INT8) return launch_tp_fmt<QL_INT8>(a, tp.world, st);
  if (fmt == QL_FP8_E4M3) return launch_tp_fmt<QL_FP8_E4M3>(a, tp.world, st);
  return launch_tp_fmt<QL_FMT_BF16>(a, tp.world, st);
}

}  // namespace ql

using namespace ql;

extern "C" {

static int tp_check_shape(const TpView& tp, int64_t M, int64_t N) {
  if (M < 0 || N <= 0 || N % 8 != 0 || N > 8192 * 4) return QL_ERR_BAD_SHAPE;
  if ((size_t)M > tp.rows || (size_t)N > tp.hidden) return QL_ERR_WORKSPACE;
  // owner slices are rounded up to whole tiles: the rounded total must fit the buffers
  if ((size_t)(tp_rows_per_owner(M, tp.world) * tp.world) > tp.rows_cap) return QL_ERR_WORKSPACE;
  return QL_OK;
}

size_t ql_tp_max_tokens(void) {
  TpView tp;
  return tp_view(&tp) ? tp.rows : 0;
}

size_t ql_tp_max_hidden(void) {
  TpView tp;
  return tp_view(&tp) ? tp.hidden : 0;
}

int ql_tp_channels(void) { return tp_channels_allocated(); }

// Two channels in flight at once have a KNOWN cross-GPU ordering hazard (bounded-wait trap at TP=8,
// profiles/r2n_n8_ag_push_micro2_trap.err): only the GRID of tp_reduce_kernel is capped at one CTA
// per SM, the hardware scheduler may still stack several of its spinning CTAs on one SM, where a
// persistent, statically strided GEMM CTA of the other channel then cannot be placed -- and the peer
// spins on the rows that GEMM would have pushed.  Until that is fixed the mode is refused unless the
// caller opts in explicitly.
static bool tp_concurrent_opt_in() {
  const char* e = getenv("QL_TP_ALLOW_CONCURRENT");
  return e != nullptr && e[0] == '1';
}

int ql_tp_set_concurrent(int on) {
  if (on && !tp_concurrent_opt_in()) return QL_ERR_GUARDED;
  __atomic_store_n(&g_tp_concurrent, on ? 1 : 0, __ATOMIC_RELAXED);
  return QL_OK;
}

void* ql_tp_buffer(int which, int channel) {
  TpView tp;
  if (!tp_view(&tp, channel)) return nullptr;
  uint8_t* b = tp.big[tp.rank];
  switch (which) {
    case 0: return b + tp.lay.part;
    case 1: return b + tp.lay.q;
    case 2: return b + tp.lay.s;
    default: return nullptr;
  }
}

int ql_tp_reduce_quant_gather(int64_t M, int64_t N, int fmt, int pushed, int channel, ql_stream_t stream) {
  if (fmt != QL_INT8 && fmt != QL_FP8_E4M3) return QL_ERR_BAD_FORMAT;
  if (channel < 0 || channel >= kTpChannels) return QL_ERR_BAD_SHAPE;
  if (channel > 0 && !tp_concurrent_opt_in()) return QL_ERR_GUARDED;
  if (channel > 0) __atomic_store_n(&g_tp_concurrent, 1, __ATOMIC_RELAXED);
  TpView tp;
  if (!tp_view(&tp, channel)) return QL_ERR_WORKSPACE;
  if (int e = tp_check_shape(tp, M, N)) return e;
  if (M == 0) return QL_OK;
  return launch_tp_reduce(tp, M, N, fmt, pushed ? 1 : 0, nullptr, 0, (cudaStream_t)stream);
}

int ql_tp_allreduce_bf16(int64_t M, int64_t N, int pushed, void* out, int64_t ldo, int channel,
                         ql_stream_t stream) {
  if (channel < 0 || channel >= kTpChannels) return QL_ERR_BAD_SHAPE;
  if (channel > 0 && !tp_concurrent_opt_in()) return QL_ERR_GUARDED;
  if (channel > 0) __atomic_store_n(&g_tp_concurrent, 1, __ATOMIC_RELAXED);
  TpView tp;
  if (!tp_view(&tp, channel)) return QL_ERR_WORKSPACE;
  if (int e = tp_check_shape(tp, M, N)) return e;
  if (ldo < N) return QL_ERR_BAD_SHAPE;
  if (M == 0) return QL_OK;
  if (!out) return QL_ERR_NULL;
  if ((uintptr_t)out % 16 != 0 || ldo % 8 != 0) return QL_ERR_MISALIGNED;
  return launch_tp_reduce(tp, M, N, -1, pushed ? 1 : 0, out, ldo, (cudaStream_t)stream);
}

int ql_tp_gemm_push(const void* qx, const float* sx, const void* qw, const float* sw, const void* bias,
                    int64_t M, int64_t N, int64_t K, int fmt, int channel, ql_stream_t stream) {
  if (fmt != QL_INT8 && fmt != QL_FP8_E4M3) return QL_ERR_BAD_FORMAT;
  if (M < 0 || N <= 0 || K <= 0 || K % 128 != 0) return QL_ERR_BAD_SHAPE;
  if (channel < 0 || channel >= kTpChannels) return QL_ERR_BAD_SHAPE;
  if (channel > 0 && !tp_concurrent_opt_in()) return QL_ERR_GUARDED;
  if (channel > 0) __atomic_store_n(&g_tp_concurrent, 1, __ATOMIC_RELAXED);
  TpView tp;
  if (!tp_view(&tp, channel)) return QL_ERR_WORKSPACE;
  if (int e = tp_check_shape(tp, M, N)) return e;
  if (M == 0) return QL_OK;
  if (!qx || !sx || !qw || !sw) return QL_ERR_NULL;
  if ((uintptr_t)qx % 16 != 0 || (uintptr_t)qw % 16 != 0) return QL_ERR_MISALIGNED;
  return launch_gemm2_push(tp, qx, sx, qw, sw, bias, M, N, K, fmt, (cudaStream_t)stream);
}

}  // extern "C"
