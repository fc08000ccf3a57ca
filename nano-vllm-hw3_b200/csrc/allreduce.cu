// One-shot all-reduce over NVLink peer memory for the row-parallel linears' decode-sized outputs
// (SURVEY.md section 8e/8f-2).  Replaces the NCCL call of RowParallelLinear.forward
// (nanovllm/layers/linear.py:150-152: `dist.all_reduce(y)`) when [M, hidden] bf16 is small: there
// latency, not bandwidth, is the cost.
//
// Push protocol, no barriers and no fences: every rank owns a receive area (cudaMalloc + CUDA
// IPC, mapped by all peers) with one region per (parity, source rank).  A thread loads 16 bytes
// of its rank's partial sums, and for every peer stores them as two 16-byte packets
// {data, epoch, data, epoch} straight into the peer's region (posted NVLink writes; data and flag
// travel in the same store, so a packet whose two flags equal the call's epoch is complete).  It
// then spins on the matching packets in its OWN region (local memory), adds the W partials in
// fp32 in rank order -- identical on all ranks, deterministic, one rounding -- and stores bf16.
// A first version (profiles/r1p_*) pulled the peers' data behind start/end flag barriers with
// st.release.sys / ld.acquire.sys: 3 NVLink round trips + system fences per call made it slower
// than NCCL.
// Epochs live in device memory (one counter, advanced by the last block to finish), regions
// alternate with the epoch's parity (a rank can be at most one call ahead of a peer: to start call
// i+2 it needs the peer's packets of call i+1, which the peer sends after it has consumed call i),
// so the kernel replays inside a CUDA graph.  No host synchronisation, no allocation per call.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "kernels.h"

namespace ql {

constexpr int kArMaxWorld = 8;
constexpr int kArMaxBlocks = 64;
constexpr int kArThreads = 256;

struct ArState {
  bool ready = false;
  bool emulated = false;                    // all W areas live in THIS process (ql_ar_init_local)
  int world = 0, rank = 0;
  size_t bytes = 0;                         // largest payload (bf16 bytes) of one one-shot call
  size_t big_rows = 0, big_hidden = 0;      // capacity of the large (reduce-scatter) path, 0 = off
  int channels = 0;                         // big sections allocated: 1, or 2 with QL_TP_ALLOW_CONCURRENT=1 at init
  uint8_t* local = nullptr;                 // this rank's receive area + counters
  uint8_t* peer[kArMaxWorld] = {};          // everybody's area as mapped here (peer[rank] == local)
};
static ArState g_ar;

// area layout: [parity 0: W regions of 2*bytes][parity 1: same][epoch u32][done u32] (256 B)
//              [big section x kTpChannels, 4 KB aligned: flags 4 KB | partial | codes | scales | bf16 gather]
static size_t ar_region_bytes(size_t bytes) { return 2 * bytes; }          // 8 B packet per 4 B payload
static size_t ar_ctr_offset(size_t bytes, int world) { return 2 * (size_t)world * ar_region_bytes(bytes); }
static size_t ar_big_offset(size_t bytes, int world) { return (ar_ctr_offset(bytes, world) + 256 + 4095) & ~(size_t)4095; }
static size_t rup(size_t v, size_t a) { return (v + a - 1) / a * a; }
// rows of the partial / receive buffer: every owner's slice is rounded up to whole 256-token tiles
static size_t big_rows_cap(size_t rows, int world) { return rup(rows, 256) + 256 * (size_t)world; }
static void big_layout(size_t rows, size_t hidden, int world, TpLayout* L) {
  const size_t cap = big_rows_cap(rows, world);
  L->flags = 0;
  L->part = 4096;
  L->q = L->part + rup(cap * hidden * 2, 1024);
  L->s = L->q + rup(cap * hidden, 1024);
  L->y = L->s + rup(cap * 4, 1024);
  L->total = L->y + rup(cap * hidden * 2, 1024);
}
// Channel 1 only serves the opt-in two-channel pipeline (tp_collective.cu): its section -- as large
// as channel 0's, ~1.4 GB at the default 65536 x 4096 capacity -- is allocated only when the opt-in
// is set at init time.  Channel 0's offsets do not depend on the count.
static int ar_want_channels() {
  const char* e = getenv("QL_TP_ALLOW_CONCURRENT");
  return (e != nullptr && e[0] == '1') ? kTpChannels : 1;
}
static size_t ar_alloc_bytes(size_t bytes, int world, size_t rows, size_t hidden, int channels) {
  size_t total = ar_big_offset(bytes, world);
  if (rows > 0 && hidden > 0) {
    TpLayout L;
    big_layout(rows, hidden, world, &L);
    total += (size_t)channels * rup(L.total, 4096);
  }
  return total;
}

// flags and call counters of every channel start at epoch 0
static bool big_clear_flags(uint8_t* area, size_t bytes, int world, size_t rows, size_t hidden, int channels) {
  if (rows == 0 || hidden == 0) return true;
  TpLayout L;
  big_layout(rows, hidden, world, &L);
  for (int c = 0; c < channels; ++c)
    if (cudaMemset(area + ar_big_offset(bytes, world) + (size_t)c * rup(L.total, 4096) + L.flags, 0, 4096) !=
        cudaSuccess)
      return false;
  return true;
}

struct ArParams {
  uint8_t* area[kArMaxWorld];         // all ranks' receive areas
  const uint4* in;                    // local partial sums (bf16)
  uint4* out;                         // local bf16 output
  uint32_t* ctr;                      // local {epoch, done}
  int64_t nvec;                       // 16-byte payload vectors
  size_t region;                      // bytes per (parity, source) region
  uint64_t timeout_ns;                // flag-wait bound (QL_AR_TIMEOUT_S; 0 = wait forever)
  int world, rank;
};

__device__ __forceinline__ void st_packet(void* p, uint32_t d0, uint32_t d1, uint32_t e) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(d0), "r"(e), "r"(d1),
               "r"(e)
               : "memory");
}
__device__ __forceinline__ uint4 ld_packet(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

template <int W>
__global__ void __launch_bounds__(kArThreads)
ar_oneshot_kernel(const ArParams p) {
  pdl_launch_dependents();
  pdl_wait();                               // the partial sums come from the previous kernel (the GEMM)
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(p.ctr) + 1;
  const size_t par = (size_t)(e & 1) * W * p.region;
  const int64_t stride = (int64_t)gridDim.x * kArThreads;
  const int64_t i0 = (int64_t)blockIdx.x * kArThreads + threadIdx.x;

  // push: my vector i -> region (parity, my rank) of every peer, packets 2i and 2i+1
  for (int64_t i = i0; i < p.nvec; i += stride) {
    const uint4 v = p.in[i];
#pragma unroll
    for (int d = 1; d < W; ++d) {
      const int r = (p.rank + d) % W;
      uint8_t* dst = p.area[r] + par + (size_t)p.rank * p.region + (size_t)i * 32;
      st_packet(dst, v.x, v.y, e);
      st_packet(dst + 16, v.z, v.w, e);
    }
  }
  // pull from my own area and reduce in rank order
  for (int64_t i = i0; i < p.nvec; i += stride) {
    const uint4 mine = p.in[i];
    float acc[8];
#pragma unroll
    for (int r = 0; r < W; ++r) {
      uint4 v;
      if (r == p.rank) {
        v = mine;
      } else {
        const uint8_t* src = p.area[p.rank] + par + (size_t)r * p.region + (size_t)i * 32;
        uint4 a, b;
        SpinGuard guard(p.timeout_ns);
        while (true) { a = ld_packet(src); if (a.y == e && a.w == e) break; guard.tick(); }
        while (true) { b = ld_packet(src + 16); if (b.y == e && b.w == e) break; guard.tick(); }
        v = make_uint4(a.x, a.z, b.x, b.z);
      }
      const float f[8] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y),
                          bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w)};
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] = r == 0 ? f[k] : __fadd_rn(acc[k], f[k]);
    }
    uint4 o;
    o.x = (uint32_t)f2bf(acc[0]) | ((uint32_t)f2bf(acc[1]) << 16);
    o.y = (uint32_t)f2bf(acc[2]) | ((uint32_t)f2bf(acc[3]) << 16);
    o.z = (uint32_t)f2bf(acc[4]) | ((uint32_t)f2bf(acc[5]) << 16);
    o.w = (uint32_t)f2bf(acc[6]) | ((uint32_t)f2bf(acc[7]) << 16);
    p.out[i] = o;
  }
  // the last block to finish publishes the epoch for the next call
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(p.ctr + 1, 1u) == gridDim.x - 1) {
      p.ctr[1] = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(p.ctr) = e;
    }
  }
}

uint64_t ar_timeout_ns() {
  static long long v = -1;
  if (v < 0) {
    const char* e = getenv("QL_AR_TIMEOUT_S");
    const double sec = (e != nullptr && e[0] != 0) ? atof(e) : 10.0;
    v = sec <= 0.0 ? 0 : (long long)(sec * 1e9);
  }
  return (uint64_t)v;
}

bool ar_view(ArView* v) {
  if (!g_ar.ready) return false;
  for (int r = 0; r < kArMaxWorld; ++r) v->area[r] = r < g_ar.world ? g_ar.peer[r] : nullptr;
  v->ctr = reinterpret_cast<uint32_t*>(g_ar.local + ar_ctr_offset(g_ar.bytes, g_ar.world));
  v->region = ar_region_bytes(g_ar.bytes);
  v->max_bytes = g_ar.bytes;
  v->world = g_ar.world;
  v->rank = g_ar.rank;
  v->timeout_ns = ar_timeout_ns();
  return true;
}

bool tp_view(TpView* v, int channel) {
  if (!g_ar.ready || g_ar.big_rows == 0 || channel < 0 || channel >= g_ar.channels) return false;
  big_layout(g_ar.big_rows, g_ar.big_hidden, g_ar.world, &v->lay);
  const size_t off = ar_big_offset(g_ar.bytes, g_ar.world) + (size_t)channel * rup(v->lay.total, 4096);
  for (int r = 0; r < kArMaxWorld; ++r) v->big[r] = r < g_ar.world ? g_ar.peer[r] + off : nullptr;
  v->rows = g_ar.big_rows;
  v->rows_cap = big_rows_cap(g_ar.big_rows, g_ar.world);
  v->hidden = g_ar.big_hidden;
  v->world = g_ar.world;
  v->rank = g_ar.rank;
  v->emulated = g_ar.emulated;
  v->timeout_ns = ar_timeout_ns();
  return true;
}

int tp_channels_allocated() { return (g_ar.ready && g_ar.big_rows != 0) ? g_ar.channels : ar_want_channels(); }

}  // namespace ql

using namespace ql;

extern "C" {

size_t ql_ar_handle_bytes(void) { return sizeof(cudaIpcMemHandle_t); }

int ql_ar_can_access_peers(int world) {
  if (world < 1 || world > kArMaxWorld) return QL_ERR_BAD_SHAPE;
  if (int e = ql_device_check()) return e;
  int dev = 0, count = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceCount(&count) != cudaSuccess) {
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  if (count < world) return QL_ERR_UNSUPPORTED_ARCH;
  // one process per GPU, rank == device ordinal on the node (model_runner.py:29)
  for (int d = 0; d < world; ++d) {
    if (d == dev) continue;
    int ok = 0;
    if (cudaDeviceCanAccessPeer(&ok, dev, d) != cudaSuccess) { (void)cudaGetLastError(); return QL_ERR_CUDA; }
    if (!ok) return QL_ERR_UNSUPPORTED_ARCH;
  }
  return QL_OK;
}

int ql_ar_init(size_t max_bytes, size_t rs_tokens, size_t rs_hidden, int world, int rank, void* handle_out) {
  if ((world != 2 && world != 4 && world != 8) || rank < 0 || rank >= world || max_bytes == 0 ||
      max_bytes % 16 != 0 || rs_hidden % 8 != 0 || (rs_tokens == 0) != (rs_hidden == 0))
    return QL_ERR_BAD_SHAPE;
  if (!handle_out) return QL_ERR_NULL;
  if (g_ar.local != nullptr) return QL_ERR_BAD_SHAPE;   // one communicator per process (one process per GPU)
  if (int e = ql_device_check()) return e;
  const int channels = ar_want_channels();
  const size_t total = ar_alloc_bytes(max_bytes, world, rs_tokens, rs_hidden, channels);
  uint8_t* ptr = nullptr;
  if (cudaMalloc(&ptr, total) != cudaSuccess) { (void)cudaGetLastError(); return QL_ERR_CUDA; }
  // only the packet regions, counters and flags need a defined initial state (epoch 0)
  const size_t head = ar_big_offset(max_bytes, world);
  if (cudaMemset(ptr, 0, head) != cudaSuccess || !big_clear_flags(ptr, max_bytes, world, rs_tokens, rs_hidden, channels) ||
      cudaDeviceSynchronize() != cudaSuccess) {
    (void)cudaGetLastError();
    cudaFree(ptr);
    return QL_ERR_CUDA;
  }
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, ptr) != cudaSuccess) {
    (void)cudaGetLastError();
    cudaFree(ptr);
    return QL_ERR_CUDA;
  }
  memcpy(handle_out, &h, sizeof(h));
  g_ar.local = ptr;
  g_ar.bytes = max_bytes;
  g_ar.big_rows = rs_tokens;
  g_ar.big_hidden = rs_hidden;
  g_ar.channels = channels;
  g_ar.world = world;
  g_ar.rank = rank;
  return QL_OK;
}

int ql_ar_open(const void* all_handles) {
  if (g_ar.local == nullptr) return QL_ERR_BAD_SHAPE;
  if (!all_handles) return QL_ERR_NULL;
  for (int r = 0; r < g_ar.world; ++r) {
    if (r == g_ar.rank) { g_ar.peer[r] = g_ar.local; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const uint8_t*>(all_handles) + (size_t)r * sizeof(h), sizeof(h));
    void* p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      (void)cudaGetLastError();
      return QL_ERR_CUDA;
    }
    g_ar.peer[r] = static_cast<uint8_t*>(p);
  }
  g_ar.ready = true;
  return QL_OK;
}

// Single-GPU emulation of a W-rank communicator for tests (the GPU test tier has one GPU): all W
// symmetric areas are allocated in this process; ql_tp_emu_set_rank selects which rank the next
// ql_tp_gemm_push / ql_tp_buffer call acts as, and the reduce kernels run ALL ranks in one launch
// (blockIdx.y = rank, co-resident grid) -- kernels that wait on one another must never be separate
// launches on one GPU.  Only the prefill-sized path (ql_tp_*) is emulated.
int ql_ar_init_local(size_t max_bytes, size_t rs_tokens, size_t rs_hidden, int world) {
  if ((world != 2 && world != 4 && world != 8) || max_bytes == 0 || max_bytes % 16 != 0 || rs_tokens == 0 ||
      rs_hidden == 0 || rs_hidden % 8 != 0)
    return QL_ERR_BAD_SHAPE;
  if (g_ar.local != nullptr) return QL_ERR_BAD_SHAPE;
  if (int e = ql_device_check()) return e;
  const int channels = ar_want_channels();
  const size_t total = ar_alloc_bytes(max_bytes, world, rs_tokens, rs_hidden, channels);
  const size_t head = ar_big_offset(max_bytes, world);
  for (int r = 0; r < world; ++r) {
    uint8_t* ptr = nullptr;
    if (cudaMalloc(&ptr, total) != cudaSuccess || cudaMemset(ptr, 0, head) != cudaSuccess ||
        !big_clear_flags(ptr, max_bytes, world, rs_tokens, rs_hidden, channels)) {
      (void)cudaGetLastError();
      if (ptr != nullptr) cudaFree(ptr);
      for (int k = 0; k < r; ++k) { cudaFree(g_ar.peer[k]); g_ar.peer[k] = nullptr; }
      return QL_ERR_CUDA;
    }
    g_ar.peer[r] = ptr;
  }
  if (cudaDeviceSynchronize() != cudaSuccess) { (void)cudaGetLastError(); return QL_ERR_CUDA; }
  g_ar.local = g_ar.peer[0];
  g_ar.bytes = max_bytes;
  g_ar.big_rows = rs_tokens;
  g_ar.big_hidden = rs_hidden;
  g_ar.channels = channels;
  g_ar.world = world;
  g_ar.rank = 0;
  g_ar.emulated = true;
  g_ar.ready = true;
  return QL_OK;
}

int ql_tp_emu_set_rank(int rank) {
  if (!g_ar.ready || !g_ar.emulated) return QL_ERR_WORKSPACE;
  if (rank < 0 || rank >= g_ar.world) return QL_ERR_BAD_SHAPE;
  g_ar.rank = rank;
  g_ar.local = g_ar.peer[rank];
  return QL_OK;
}

// Unmap the peers' areas and free this rank's (after a failed set-up, or at shutdown once every
// rank has stopped issuing calls).  Idempotent.
int ql_ar_destroy(void) {
  if (g_ar.local == nullptr) return QL_OK;
  (void)cudaDeviceSynchronize();
  if (g_ar.emulated) {
    for (int r = 0; r < g_ar.world; ++r)
      if (g_ar.peer[r] != nullptr) (void)cudaFree(g_ar.peer[r]);
    (void)cudaGetLastError();
    g_ar = ArState();
    return QL_OK;
  }
  for (int r = 0; r < g_ar.world; ++r) {
    if (r != g_ar.rank && g_ar.peer[r] != nullptr) (void)cudaIpcCloseMemHandle(g_ar.peer[r]);
    g_ar.peer[r] = nullptr;
  }
  (void)cudaFree(g_ar.local);
  (void)cudaGetLastError();
  g_ar = ArState();
  return QL_OK;
}

size_t ql_ar_max_bytes(void) { return g_ar.ready ? g_ar.bytes : 0; }

int ql_ar_allreduce_bf16(const void* in, void* out, int64_t numel, ql_stream_t stream) {
  if (!g_ar.ready || g_ar.emulated) return QL_ERR_WORKSPACE;
  if (numel < 0 || numel % 8 != 0 || (size_t)numel * 2 > g_ar.bytes) return QL_ERR_BAD_SHAPE;
  if (numel == 0) return QL_OK;
  if (!in || !out) return QL_ERR_NULL;
  if ((uintptr_t)in % 16 != 0 || (uintptr_t)out % 16 != 0) return QL_ERR_MISALIGNED;
  ArParams p;
  for (int r = 0; r < kArMaxWorld; ++r) p.area[r] = r < g_ar.world ? g_ar.peer[r] : nullptr;
  p.in = static_cast<const uint4*>(in);
  p.out = static_cast<uint4*>(out);
  p.ctr = reinterpret_cast<uint32_t*>(g_ar.local + ar_ctr_offset(g_ar.bytes, g_ar.world));
  p.nvec = numel / 8;
  p.region = ar_region_bytes(g_ar.bytes);
  p.timeout_ns = ar_timeout_ns();
  p.world = g_ar.world;
  p.rank = g_ar.rank;
  int64_t blocks = (p.nvec + kArThreads - 1) / kArThreads;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)blocks);
  cfg.blockDim = dim3(kArThreads);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaError_t err;
  switch (g_ar.world) {
    case 2: err = cudaLaunchKernelEx(&cfg, ar_oneshot_kernel<2>, p); break;
    case 4: err = cudaLaunchKernelEx(&cfg, ar_oneshot_kernel<4>, p); break;
    default: err = cudaLaunchKernelEx(&cfg, ar_oneshot_kernel<8>, p); break;
  }
  count_launch();
  if (err != cudaSuccess) { (void)cudaGetLastError(); return QL_ERR_CUDA; }
  return QL_OK;
}

int ql_ar_rmsnorm_quant(const void* partial, const void* residual_in, void* residual_out,
                        const void* weight, float eps, int64_t M, int64_t K, int fmt, void* q,
                        float* s, ql_stream_t stream) {
  if (fmt != QL_INT8 && fmt != QL_FP8_E4M3) return QL_ERR_BAD_FORMAT;
  ArView ar;
  if (!ar_view(&ar) || g_ar.emulated) return QL_ERR_WORKSPACE;
  if (M < 0 || K <= 0 || K % 8 != 0 || (size_t)M * (size_t)K * 2 > ar.max_bytes) return QL_ERR_BAD_SHAPE;
  if (M == 0) return QL_OK;
  if (!partial || !residual_in || !residual_out || !weight || !q || !s) return QL_ERR_NULL;
  if ((uintptr_t)partial % 16 || (uintptr_t)residual_in % 16 || (uintptr_t)residual_out % 16 ||
      (uintptr_t)weight % 16 || (uintptr_t)q % 8)
    return QL_ERR_MISALIGNED;
  return launch_ar_rmsnorm_quant(ar, partial, residual_in, residual_out, weight, eps, M, K, fmt, q, s,
                                 (cudaStream_t)stream);
}

}  // extern "C"
