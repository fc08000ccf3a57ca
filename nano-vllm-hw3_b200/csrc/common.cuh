// Shared device helpers for the sm_100a quantized-linear kernels: number formats, cache-hinted
// vector loads, mbarrier / TMA / tcgen05 PTX wrappers.  sm_100a only -- no other arch is built.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/qlinear_b200.h"

namespace ql {

constexpr int kWarp = 32;

// ------------------------------------------------------------------ number formats

template <int FMT> struct Fmt;
template <> struct Fmt<QL_INT8> { static constexpr float qmax = 127.0f; };
template <> struct Fmt<QL_FP8_E4M3> { static constexpr float qmax = 448.0f; };

__device__ __forceinline__ float bf16_lo(uint32_t p) { return __uint_as_float(p << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t p) { return __uint_as_float(p & 0xffff0000u); }

// float -> bf16 bits, round to nearest even (cvt.rn.bf16.f32).
__device__ __forceinline__ uint16_t f2bf(float f) {
  return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}

// roundf(v) (half away from zero) as an integer, for |v| < 2^22, without the quarter-rate F2I /
// FRND conversions: n = floor(|v| + 0.5) where both additions round toward zero -- RZ never
// carries a sum across an integer, so floor(RZ(|v| + 0.5)) == floor(|v| + 0.5) exactly, and
// RZ(t + 2^23) leaves floor(t) in the low mantissa bits.
__device__ __forceinline__ int round_away_i8(float v) {
  const float t = __fadd_rz(fabsf(v), 0.5f);
  const int n = __float_as_int(__fadd_rz(t, 8388608.0f)) & 0x7fffff;
  return v < 0.0f ? -n : n;
}

// Quantise two scaled values to one 16-bit pair of codes.
//   int8: round half away from zero (libdevice.round, reference quantization.py:41)
//   fp8 : cvt.rn.satfinite.e4m3x2.f32 (reference :39)
template <int FMT>
__device__ __forceinline__ uint32_t quant_pair_away(float a, float b) {
  if constexpr (FMT == QL_INT8) {
    return (uint32_t)(round_away_i8(a) & 0xff) | ((uint32_t)(round_away_i8(b) & 0xff) << 8);
  } else {
    return (uint32_t)__nv_cvt_float2_to_fp8x2(make_float2(a, b), __NV_SATFINITE, __NV_E4M3);
  }
}

// torch semantics (per_group_quant, reference :224-228): int8 = clamp(rint(v)), fp8 = cast.
template <int FMT>
__device__ __forceinline__ uint32_t quant_pair_even(float a, float b) {
  if constexpr (FMT == QL_INT8) {
    float ra = fminf(fmaxf(rintf(a), -128.0f), 127.0f);
    float rb = fminf(fmaxf(rintf(b), -128.0f), 127.0f);
    return (uint32_t)((int)ra & 0xff) | ((uint32_t)((int)rb & 0xff) << 8);
  } else {
    return (uint32_t)__nv_cvt_float2_to_fp8x2(make_float2(a, b), __NV_SATFINITE, __NV_E4M3);
  }
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// max |.| over the 8 bf16 values of a 16-byte vector
__device__ __forceinline__ float amax8(const uint4& v) {
  float m = fmaxf(fabsf(bf16_lo(v.x)), fabsf(bf16_hi(v.x)));
  m = fmaxf(m, fmaxf(fabsf(bf16_lo(v.y)), fabsf(bf16_hi(v.y))));
  m = fmaxf(m, fmaxf(fabsf(bf16_lo(v.z)), fabsf(bf16_hi(v.z))));
  m = fmaxf(m, fmaxf(fabsf(bf16_lo(v.w)), fabsf(bf16_hi(v.w))));
  return m;
}

// 8 bf16 -> 8 codes (uint2), v * inv then quantise (Triton per-row semantics).
template <int FMT>
__device__ __forceinline__ uint2 quant8_mul(const uint4& v, float inv) {
  uint32_t p0 = quant_pair_away<FMT>(__fmul_rn(bf16_lo(v.x), inv), __fmul_rn(bf16_hi(v.x), inv));
  uint32_t p1 = quant_pair_away<FMT>(__fmul_rn(bf16_lo(v.y), inv), __fmul_rn(bf16_hi(v.y), inv));
  uint32_t p2 = quant_pair_away<FMT>(__fmul_rn(bf16_lo(v.z), inv), __fmul_rn(bf16_hi(v.z), inv));
  uint32_t p3 = quant_pair_away<FMT>(__fmul_rn(bf16_lo(v.w), inv), __fmul_rn(bf16_hi(v.w), inv));
  return make_uint2(p0 | (p1 << 16), p2 | (p3 << 16));
}

// ------------------------------------------------------------------ global memory access

// streaming 16-byte load that does not allocate in L1 (weights are read exactly once)
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// ------------------------------------------------------------------ mbarrier

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Bounded wait: a pipeline bug (bad tensor map, wrong byte count) must surface as a CUDA error
// (trap -> cudaErrorLaunchFailure), never as a kernel that spins forever on a shared GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = globaltimer_ns();
  while (!mbar_try_wait(bar, parity)) {
    if (globaltimer_ns() - t0 > 4000000000ull) __trap();  // 4 s
  }
}

// ------------------------------------------------------------------ clusters / DSMEM / PDL

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of `local` (a shared::cta address) inside CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
// split cluster barrier, usable from divergent code (no .aligned)
__device__ __forceinline__ void cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ void st_cluster_u32(uint32_t cluster_addr, uint32_t v) {
  asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(cluster_addr), "r"(v) : "memory");
}
// remote arrive that publishes this thread's earlier DSMEM stores to the waiting CTA
__device__ __forceinline__ void mbar_arrive_release_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_acq_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_acq_cluster(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait_acq_cluster(bar, parity)) return;
  const uint64_t t0 = globaltimer_ns();
  while (!mbar_try_wait_acq_cluster(bar, parity)) {
    if (globaltimer_ns() - t0 > 4000000000ull) __trap();
  }
}
// generic-proxy shared-memory writes -> visible to the async proxy (TMA / tcgen05 operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// Bounded spin for cross-GPU flag waits: a peer that died must not hang this GPU forever.  After
// `limit_ns` of %globaltimer (QL_AR_TIMEOUT_S, default 10 s; 0 = wait forever, as NCCL would) the
// kernel traps, which surfaces as a CUDA error on the host -- loud, not silent.  Legitimate skew
// (a peer in first-call module load, under a debugger) needs a larger bound or 0.  These kernels
// cannot run under ncu kernel replay: a replayed launch sees stale epochs and can never match.
struct SpinGuard {
  uint32_t n = 0;
  uint64_t t0 = 0, limit;
  __device__ __forceinline__ explicit SpinGuard(uint64_t limit_ns) : limit(limit_ns) {}
  __device__ __forceinline__ void tick() {
    if ((++n & 0xfffu) == 0 && limit != 0) {
      uint64_t t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (t0 == 0) t0 = t;
      else if (t - t0 > limit) __trap();
    }
  }
};

// system-scope flag accesses for cross-GPU hand-offs (peer-mapped or local memory)
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// 16-byte load that is never served from a stale L1 line (peer-written or peer-resident data)
__device__ __forceinline__ uint4 ld_sys128(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

// Programmatic dependent launch: a kernel launched with the programmatic-stream-serialization
// attribute starts while its predecessor drains; everything that reads the predecessor's output
// (or overwrites its input) must sit behind pdl_wait().  Both are no-ops for a plain launch.
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ------------------------------------------------------------------ TMA (cp.async.bulk.tensor)

__device__ __forceinline__ void tma_prefetch_desc(const void* desc) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(desc) : "memory");
}

// 2-D tiled load: coordinates (c0 = innermost/K byte offset, c1 = row)
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* desc, uint64_t* bar,
                                            int32_t c0, int32_t c1, uint64_t cache_hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(desc), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(cache_hint)
      : "memory");
}

// L2 cache-policy constants (same encodings CUTLASS uses for TMA::CacheHintSm90)
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;

// ------------------------------------------------------------------ tcgen05 / TMEM

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_out, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_out)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// commit all prior tcgen05.mma of this thread; arrive on `bar` when they complete
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]^T ; int8 x int8 -> int32   (SASS: UTCIMMA)
__device__ __forceinline__ void mma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                       uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// e4m3 x e4m3 -> fp32   (SASS: UTCQMMA)
__device__ __forceinline__ void mma_f8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                       uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// TMEM -> registers: this warp's 32 lanes x 32 consecutive 32-bit columns (SASS: LDTM)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
        "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// wait + make the loaded registers depend on it, so that consumers of `r` cannot be scheduled
// above the wait when other tcgen05.ld's are kept in flight around it
__device__ __forceinline__ void tmem_ld_wait_regs(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) asm volatile("" : "+r"(r[i]));
}

// K-major, 128-byte-swizzled shared-memory operand descriptor (sm_100 format):
//   [0,14) start address >> 4 | [16,30) LBO >> 4 (unused for swizzled K-major) |
//   [32,46) SBO >> 4 = 1024 B (8 rows x 128 B) | [46,48) version = 1 | [61,64) layout = SWIZZLE_128B (2)
__device__ __forceinline__ uint64_t make_sw128_kmajor_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffffu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Instruction descriptor for kind::i8 / kind::f8f6f4, both operands K-major, dense:
//   [4,6) D format (1 = F32, 2 = S32) | [7,10) A format | [10,13) B format |
//   [17,23) N >> 3 | [24,29) M >> 4
template <int FMT>
__host__ __device__ constexpr uint32_t make_idesc(uint32_t umma_m, uint32_t umma_n) {
  uint32_t d = 0;
  if (FMT == QL_INT8) {
    d |= 2u << 4;   // S32 accumulate
    d |= 1u << 7;   // A = signed int8
    d |= 1u << 10;  // B = signed int8
  } else {
    d |= 1u << 4;   // F32 accumulate
    d |= 0u << 7;   // A = E4M3
    d |= 0u << 10;  // B = E4M3
  }
  d |= (umma_n >> 3) << 17;
  d |= (umma_m >> 4) << 24;
  return d;
}

// ------------------------------------------------------------------ persistent tile order

// Tile order: token tiles vary fastest inside super-groups of `group_m` token tiles, so the
// tiles in flight at any moment share a few weight tiles and a set of activation tiles that
// stays L2-resident, and every weight tile is fetched from HBM once per super-group.
__device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int group_m,
                                            int& tm, int& tn) {
  const int per_group = group_m * tiles_n;
  const int g = tile / per_group, r = tile - g * per_group;
  const int gm = min(group_m, tiles_m - g * group_m);
  tn = r / gm;
  tm = g * group_m + (r - tn * gm);
}

}  // namespace ql
