// Prefill-regime W8A8 GEMM: CTA-pair (cta_group::2) tcgen05 kernel with per-row scales.
// Replaces the reference's Triton kernel _per_row_post_process_mm
// (nanovllm/utils/quantization.py:54-144) for token counts that fill the machine.
//
// Why a second GEMM kernel: at Qwen3 shapes K is short (1024..4096), so a tile's operands are
// loaded once per output tile and the L2 -> shared-memory stream, not the tensor pipe, is the
// first limit.  A CTA pair computes a 256-token x 256-feature tile with ONE
// tcgen05.mma.cta_group::2 (UMMA 256x256x32): each CTA stages only its own 128 token rows (A)
// and HALF of the 256 weight rows (B); the tensor core reads the other half from the peer's
// shared memory.  Per CTA that is 32 KB of operands per 128-byte K block for 128x256 outputs
// (256 FLOP per L2 byte, against 128 for the single-CTA 128x128 tile).
//
// Orientation: tokens on the TMEM lanes, features on the TMEM columns.  An epilogue thread owns
// one token row; after tcgen05.ld it holds 32 consecutive features of that token, applies
// (float(acc) * sx[m]) * sw[n] (+ bias[n]) -- the reference's product order, :111-121 -- packs
// bf16 pairs and writes 16-byte pieces into a 128B-swizzled staging tile that one lane hands
// to the TMA store engine (cp.async.bulk.tensor ... global.shared::cta).  No per-element
// address arithmetic, no transposition, stores fully coalesced by the TMA unit.
//
// Pipeline per CTA (384 threads; both CTAs of a pair run the same roles except the MMA issue):
//   warp 0 / lane 0  TMA producer: own A rows + own half of B per K block -> STAGES-deep ring;
//                    completion bytes are counted on the LEADER CTA's `full` barrier
//   warp 1 / lane 0  (leader CTA only) MMA issuer; tcgen05.commit multicast to both CTAs frees
//                    the ring slot (`empty`) and publishes the accumulator (`tfull`)
//   warp 2           TMEM allocator (512 columns = two 256-column accumulators)
//   warps 4-11       epilogue (TMEM -> registers -> scales -> bf16 -> smem -> TMA store), two
//                    warps per SM sub-partition so their tcgen05.ld / fence / store waits overlap;
//                    arrives on the leader's `tempty` when an accumulator has been drained
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace ql {

struct Gemm2Params {
  const float* sx;       // [M]
  const float* sw;       // [N]
  const uint16_t* bias;  // bf16 [N] or nullptr
  int32_t* acc_out;      // int32 [M, N] (OUT_ACC only)
  int M, N, K;
  int tiles_m, tiles_n;  // pair tiles (256 tokens x 256 features)
  int group_m;           // token tiles per rasterisation super-group
  long long* trace;      // debug (ql_debug_set_trace): globaltimer stamps [launch slot][cta][8], else nullptr
};
#define G2_TRACE(i)                                                                                   \
  do {                                                                                                \
    if (p.trace != nullptr) p.trace[(size_t)blockIdx.x * 8 + (i)] = (long long)globaltimer_ns();       \
  } while (0)

// Fused reduce-scatter (PUSH): where the tiles go and whom to tell.
struct Gemm2Push {
  CUtensorMap ymap[8];        // per token owner: its receive region for THIS rank's partials, [rows_o, N] bf16
  int tiles_per_owner;        // 256-token tiles per owner slice
  int rot;                    // token-tile rotation: rank r starts with owner r+1's tiles and ends with its own,
                              // so at any moment the W senders address W different receivers and the local
                              // (wire-free) tiles are the tail
};

constexpr int kG2Threads = 384;   // 4 control warps + 8 epilogue warps
constexpr int kG2EpiWarps = 8;
constexpr int kG2RowsA = 128;    // token rows per CTA (UMMA M = 256 per pair)
constexpr int kG2TileN = 256;    // features per pair tile (UMMA N)
constexpr int kG2RowsB = 128;    // weight rows staged per CTA (half of the tile)
constexpr int kG2BlockK = 128;   // bytes of K per stage (one 128B swizzle atom)
constexpr int kG2Stages = 5;
constexpr int kG2StageA = kG2RowsA * kG2BlockK;
constexpr int kG2StageB = kG2RowsB * kG2BlockK;
constexpr int kG2Stage = kG2StageA + kG2StageB;                 // 32 KB
constexpr int kG2StoreCols = 64;                                // bf16 columns per TMA store box (128 B)
constexpr int kG2StageOut = 32 * kG2StoreCols * 2;              // 4 KB: [32 tokens][64 features]
constexpr int kG2EpiSmem = kG2EpiWarps * kG2StageOut;            // one staging tile per epilogue warp
constexpr int kG2VecSmem = 2 * 2 * kG2TileN * 4;                // sw / bias tiles (fp32), double buffered
constexpr int kG2Smem = kG2Stages * kG2Stage + kG2EpiSmem + kG2VecSmem + 256 + 1024;

// ---- cluster / peer helpers -----------------------------------------------------------------

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::
                   : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load issued by either CTA of a pair; completion bytes land on the barrier at cluster
// address `bar_cluster` (the leader's)
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const void* desc,
                                                 uint32_t bar_cluster, int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(desc), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void* desc, uint32_t smem_src, int32_t c0,
                                             int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   desc),
               "r"(smem_src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* smem_out, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_out)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols)
               : "memory");
}
// commit all prior MMAs of this thread; arrive on the barrier at the same offset in both CTAs
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
      "[%0], %1;" ::"r"(smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void mma2_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                        uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma2_f8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                        uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c,
                                       uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c),
               "r"(d)
               : "memory");
}
__device__ __forceinline__ float4 lds128f(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts32f(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

template <int FMT, bool OUT_ACC, bool HAS_BIAS, bool PUSH>
__device__ __forceinline__ void gemm2_body(const CUtensorMap& tm_x, const CUtensorMap& tm_w,
                                           const CUtensorMap* tm_y, const Gemm2Params& p,
                                           const Gemm2Push* push) {
  static_assert(!(OUT_ACC && FMT != QL_INT8), "raw accumulators: int8 only");
  static_assert(!(PUSH && OUT_ACC), "the fused reduce-scatter stores bf16 tiles");
  extern __shared__ uint8_t smem_raw[];
  // all offsets below are identical in both CTAs of the pair (the MMA and the multicast
  // commits address the peer's shared memory by offset)
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi0 = smem0 + kG2Stages * kG2Stage;   // TMA-store staging, 1024-aligned
  const uint32_t vec0 = epi0 + kG2EpiSmem;              // sw / bias tiles
  uint8_t* smem_gen = smem_raw + (smem0 - smem_u32(smem_raw));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_gen + kG2Stages * kG2Stage + kG2EpiSmem +
                                                   kG2VecSmem);
  uint64_t* empty_bar = full_bar + kG2Stages;
  uint64_t* tfull_bar = empty_bar + kG2Stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();   // 0 = leader (issues the MMAs)
  if (threadIdx.x == 0) G2_TRACE(0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    tma_prefetch_desc(&tm_w);
    if constexpr (!PUSH) tma_prefetch_desc(tm_y);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kG2Stages; ++i) {
      mbar_init(&full_bar[i], 1);    // the leader's arrive.expect_tx (bytes of BOTH CTAs)
      mbar_init(&empty_bar[i], 1);   // one multicast tcgen05.commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);   // one multicast tcgen05.commit
      mbar_init(&tempty_bar[i], 2 * kG2EpiWarps);  // epilogue warps of both CTAs
    }
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc2(tmem_slot, 512);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) G2_TRACE(1);
  pdl_launch_dependents();   // the next kernel (a quantiser) may become resident and wait

  const int num_kb = p.K / kG2BlockK;
  const int total_tiles = p.tiles_m * p.tiles_n;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      // Programmatic dependent launch: the weights never depend on the previous kernel, so the
      // first ring fill's B halves are requested before griddepcontrol.wait (the quantiser that
      // writes qx / sx may still be running); A loads and everything after sit behind the wait.
      int pre = 0;
      {
        int stage = 0;
        for (int tile = pair; tile < total_tiles && pre < kG2Stages; tile += num_pairs) {
          int tm, tn;
          tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
          if constexpr (PUSH) { tm += push->rot; if (tm >= p.tiles_m) tm -= p.tiles_m; }
          const int row_b = tn * kG2TileN + (int)cta * kG2RowsB;
          for (int kb = 0; kb < num_kb && pre < kG2Stages; ++kb, ++pre, ++stage) {
            const uint32_t full_leader = mapa_u32(smem_u32(&full_bar[stage]), 0);
            if (cta == 0) mbar_expect_tx(&full_bar[stage], 2 * kG2Stage);
            tma_load_2d_pair(smem_gen + stage * kG2Stage + kG2StageA, &tm_w, full_leader,
                             kb * kG2BlockK, row_b);
          }
        }
      }
      pdl_wait();
      G2_TRACE(2);
      int stage = 0, it = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs) {
        int tm, tn;
        tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
          if constexpr (PUSH) { tm += push->rot; if (tm >= p.tiles_m) tm -= p.tiles_m; }
        const int row_a = tm * 2 * kG2RowsA + (int)cta * kG2RowsA;
        const int row_b = tn * kG2TileN + (int)cta * kG2RowsB;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const uint32_t full_leader = mapa_u32(smem_u32(&full_bar[stage]), 0);
          uint8_t* sa = smem_gen + stage * kG2Stage;
          if (it >= pre) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            // Both CTAs' bytes are counted on the leader's barrier; only the leader arrives
            // (a local, fence-free operation).  The peer's bytes may land before the leader's
            // expect_tx of the same phase: the transaction count goes negative transiently and
            // the phase still cannot complete before the leader's arrival.
            if (cta == 0) mbar_expect_tx(&full_bar[stage], 2 * kG2Stage);
            tma_load_2d_pair(sa + kG2StageA, &tm_w, full_leader, kb * kG2BlockK, row_b);
          }
          tma_load_2d_pair(sa, &tm_x, full_leader, kb * kG2BlockK, row_a);
          if (++stage == kG2Stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (cta == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc<FMT>(2 * kG2RowsA, kG2TileN);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kG2TileN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (tile == pair && kb == 0) G2_TRACE(3);
          const uint32_t sa = smem0 + stage * kG2Stage;
          const uint64_t a_desc = make_sw128_kmajor_desc(sa);
          const uint64_t b_desc = make_sw128_kmajor_desc(sa + kG2StageA);
#pragma unroll
          for (int k = 0; k < kG2BlockK / 32; ++k) {
            const uint32_t accumulate = (kb > 0 || k > 0);
            if constexpr (FMT == QL_INT8)
              mma2_i8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, accumulate);
            else
              mma2_f8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, accumulate);
          }
          tc_commit_pair(&empty_bar[stage]);
          if (++stage == kG2Stages) { stage = 0; phase ^= 1; }
        }
        tc_commit_pair(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      G2_TRACE(4);
    }
  } else if (warp >= 4) {
    // ===================== epilogue (both CTAs, 8 warps) =====================
    // Warp ew reads TMEM lane quarter q = ew % 4 (hardware: a warp may only touch lanes
    // 32*(warp%4)..+31) and column half ew / 4, i.e. a [32 tokens][128 features] block per tile
    // in two 64-column steps.  Two warps per SM sub-partition: while one waits on tcgen05.ld,
    // the proxy fence or the TMA store engine, the other issues.
    const int ew = warp - 4;
    const int q = ew & 3, half = ew >> 2;
    const int et = threadIdx.x - 128;             // 0..255 among the epilogue threads
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16) + half * 128;
    const uint32_t stage_out = epi0 + ew * kG2StageOut;
    const uint32_t tempty_leader0 = mapa_u32(smem_u32(&tempty_bar[0]), 0);
    const uint32_t swz = (uint32_t)(lane & 7) << 4;   // 128B swizzle: 16-byte chunk ^= row & 7
    const uint32_t out = stage_out + lane * 128;
    int acc = 0;
    uint32_t acc_phase = 0;
    pdl_wait();   // sx and the y buffer belong to the previous kernels

    for (int tile = pair; tile < total_tiles; tile += num_pairs) {
      int tm, tn;
      tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
          if constexpr (PUSH) { tm += push->rot; if (tm >= p.tiles_m) tm -= p.tiles_m; }
      const int n0 = tn * kG2TileN + half * 128;                          // first feature of this warp
      const int m0 = tm * 2 * kG2RowsA + (int)cta * kG2RowsA + q * 32;   // first token of this warp
      const int m = m0 + lane;
      // PUSH: the tile belongs to the rank that owns these 256 tokens; its rows are addressed
      // relative to the owner's slice, in the owner's receive region for this rank (peer memory)
      const CUtensorMap* ymap = tm_y;
      int m_store = m0;
      if constexpr (PUSH) {
        const int owner = tm / push->tiles_per_owner;
        ymap = &push->ymap[owner];
        m_store = m0 - owner * push->tiles_per_owner * 2 * kG2RowsA;
      }
      float sxm = 0.0f;
      const uint32_t sw_s = vec0 + acc * (2 * kG2TileN * 4) + half * 128 * 4;
      const uint32_t bias_s = sw_s + kG2TileN * 4;
      if constexpr (!OUT_ACC) {
        // stage this tile's weight scales / bias (fp32) for broadcast LDS.128 reads: one
        // column per epilogue thread, double-buffered by accumulator index
        const int n = tn * kG2TileN + et;
        sts32f(vec0 + acc * (2 * kG2TileN * 4) + et * 4, n < p.N ? __ldg(p.sw + n) : 0.0f);
        if constexpr (HAS_BIAS)
          sts32f(vec0 + acc * (2 * kG2TileN * 4) + kG2TileN * 4 + et * 4,
                 n < p.N ? bf16_lo(p.bias[n]) : 0.0f);
        sxm = m < p.M ? __ldg(p.sx + m) : 0.0f;
        named_bar_sync(1, 32 * kG2EpiWarps);
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      if (threadIdx.x == 128 && tile + num_pairs >= total_tiles) G2_TRACE(5);   // last tile's accumulator is ready

      if constexpr (OUT_ACC) {
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          uint32_t r[32];
          tmem_ld32(lane_addr + acc * kG2TileN + c * 32, r);
          tmem_ld_wait();
          if (m < p.M) {
            const int nb = n0 + c * 32;
            int32_t* dst = p.acc_out + (int64_t)m * p.N + nb;
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (nb + i < p.N) dst[i] = (int32_t)r[i];
          }
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {                 // 64 features per step
          const int nc = n0 + c * 64;                 // first feature of this step
          uint32_t r0[32], r1[32];
          tmem_ld32(lane_addr + acc * kG2TileN + c * 64, r0);
          tmem_ld32(lane_addr + acc * kG2TileN + c * 64 + 32, r1);
          tmem_ld_wait();
          uint32_t pk[32];                            // 64 bf16 results of this token
#pragma unroll
          for (int h = 0; h < 2; ++h) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {             // 8 features -> one 16-byte chunk
              const int col = c * 64 + h * 32 + j * 8;   // column inside this warp's 128
              const float4 s0 = lds128f(sw_s + col * 4), s1 = lds128f(sw_s + col * 4 + 16);
              const float sv[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
              float bv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
              if constexpr (HAS_BIAS) {
                const float4 b0 = lds128f(bias_s + col * 4), b1 = lds128f(bias_s + col * 4 + 16);
                bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
                bv[4] = b1.x; bv[5] = b1.y; bv[6] = b1.z; bv[7] = b1.w;
              }
              float v[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const uint32_t raw = h == 0 ? r0[j * 8 + e] : r1[j * 8 + e];
                const float a = FMT == QL_INT8 ? __int2float_rn((int)raw) : __uint_as_float(raw);
                float t = __fmul_rn(__fmul_rn(a, sxm), sv[e]);   // (acc * sa) * sb, reference :121
                if constexpr (HAS_BIAS) t = __fadd_rn(t, bv[e]);
                v[e] = t;
              }
#pragma unroll
              for (int e = 0; e < 4; ++e) pk[(h * 4 + j) * 4 + e] = pack_bf16x2(v[2 * e], v[2 * e + 1]);
            }
          }
          // the staging tile is rewritten: the TMA store issued from it one step ago (a whole
          // step of arithmetic earlier) must have finished reading shared memory
          if (lane == 0) tma_store_wait_read<0>();
          __syncwarp();
#pragma unroll
          for (int ch = 0; ch < 8; ++ch)
            sts128(out + (((uint32_t)ch << 4) ^ swz), pk[ch * 4], pk[ch * 4 + 1], pk[ch * 4 + 2],
                   pk[ch * 4 + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            // rows >= M and columns >= N are clipped by the TMA unit; a box entirely outside
            // the tensor is not issued
            if (m0 < p.M && nc < p.N) tma_store_2d(ymap, stage_out, nc, m_store);
            tma_store_commit();
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(tempty_leader0 + acc * 8);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    // Every bulk store of this CTA must be COMPLETE (written, not merely read out of the staging
    // tiles) before the CTA exits.  For PUSH the destinations are peer memory and the reduce kernel
    // behind this one in the stream tells the peers "your rows have landed" with an st.release.sys
    // flag: that is only true if the async-proxy stores were performed first.  Grid completion is
    // widely assumed to imply it, but nothing we can cite says so for cp.async.bulk groups that
    // were never waited on, and a violation would show up as rare stale rows -- so the last thing
    // the epilogue does is wait for the whole group (one NVLink round trip per CTA, paid while the
    // dependent reduce kernel is still parked at its griddepcontrol.wait anyway).
    if (lane == 0) tma_store_wait_all();
    if (threadIdx.x == 128) G2_TRACE(6);
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc2(tmem_base, 512);
  if (threadIdx.x == 0) G2_TRACE(7);
}

// 128 registers x 384 threads leave a quarter of the register file free: the tensor-parallel
// reduce kernel of the OTHER channel (tp_collective.cu, 128 threads) stays co-resident with a GEMM
// CTA, which the two-channel pipeline relies on (no spills at this bound).
template <int FMT, bool OUT_ACC, bool HAS_BIAS>
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(128)
gemm2_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
             const __grid_constant__ CUtensorMap tm_y, const Gemm2Params p) {
  gemm2_body<FMT, OUT_ACC, HAS_BIAS, false>(tm_x, tm_w, &tm_y, p, nullptr);
}

// GEMM + reduce-scatter over NVLink in one kernel: UTC*MMA tiles whose epilogue TMA-stores into
// peer memory (see Gemm2Push).
template <int FMT, bool HAS_BIAS>
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(128)
gemm2_push_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                  const __grid_constant__ Gemm2Push push, const Gemm2Params p) {
  gemm2_body<FMT, false, HAS_BIAS, true>(tm_x, tm_w, nullptr, p, &push);
}

// ------------------------------------------------------------------ per-group scales on the CTA pair
//
// per_group_matmul (reference quantization.py:265-307; W8A8 form): weight scales
// sw[n, j] per `group` values of K.  One TMEM accumulator per 128-byte K block, ping-ponged (2 x 256
// columns = all of TMEM): while the tensor core fills P_{b+1} the epilogue warps promote P_b into
// fp32 running sums in REGISTERS, t[n] = fmaf(float(P_b[n]), sw[n, b / blocks_per_group], t[n]), in
// K order -- the arithmetic of gemm_kernel<GROUPED> and of the oracle's gemm_groups, bit for bit.
// What bounds it is the TMEM READ path: every K block the whole 128 x 256 x 4 B accumulator of a
// CTA goes through tcgen05.ld (64 B/clk/SM nominal) -- 64 tensor FLOP per TMEM byte, a ceiling of
// ~1.2-1.6 PFLOP/s whatever the tile shape (measured: 89 B/clk here, 71 B/clk in gemm_kernel).
//
// The 128 x 256 fp32 running sums of a CTA are half of its register file, so this kernel runs
// SIXTEEN epilogue warps (4 per TMEM lane quarter, 64 feature columns each: t[64] per thread) next
// to two control warps (TMA producer; TMEM allocator + MMA issuer).  The scales of 8 groups x 256
// features are staged in shared memory transposed ([group][feature], double buffered, the next
// panel in flight as registers while the current one is used), so a thread reads the 4 scales of
// 4 consecutive columns with one broadcast LDS.128.  INT8 partial sums are converted with the
// exact magic-number add (|P_b| <= 128 * 127^2 < 2^22: IADD + FADD).  Results leave as 16-byte global stores (a thread owns 128 contiguous bytes of
// its token's row).

constexpr int kG2gEpiWarps = 16;
constexpr int kG2gThreads = 64 + 32 * kG2gEpiWarps;             // 576
constexpr int kG2gStages = 5;
constexpr int kG2gPanel = 8;                                    // scale groups per staged panel
constexpr int kG2gSwSmem = 2 * kG2gPanel * kG2TileN * 4;        // [2][8][256] fp32
constexpr int kG2gBiasSmem = 2 * kG2TileN * 4;                  // [2][256] fp32
constexpr int kG2gSmem = kG2gStages * kG2Stage + kG2gSwSmem + kG2gBiasSmem + 256 + 1024;

struct Gemm2gParams {
  const float* sx;       // [M]
  const float* sw;       // [N, G]
  const uint16_t* bias;  // bf16 [N] or nullptr
  uint16_t* y;           // bf16 [M, ldy]
  int64_t ldy;
  int M, N, K;
  int G, kb_per_g;       // scale groups along K; 128-byte K blocks per group
  int tiles_m, tiles_n, group_m;
};

__device__ __forceinline__ void tmem_ld_wait_regs16(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) asm volatile("" : "+r"(r[i]));
}
__device__ __forceinline__ void stg128(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

template <int FMT, bool HAS_BIAS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kG2gThreads, 1)
gemm2g_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
              const Gemm2gParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem0 - smem_u32(smem_raw));
  const uint32_t sw0 = smem0 + kG2gStages * kG2Stage;     // [2][8][256] fp32
  const uint32_t bias0 = sw0 + kG2gSwSmem;                // [2][256] fp32
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_gen + kG2gStages * kG2Stage + kG2gSwSmem + kG2gBiasSmem);
  uint64_t* empty_bar = full_bar + kG2gStages;
  uint64_t* tfull_bar = empty_bar + kG2gStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();   // 0 = leader (issues the MMAs)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    tma_prefetch_desc(&tm_w);
    for (int i = 0; i < kG2gStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], 2 * kG2gEpiWarps);   // epilogue warps of both CTAs
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc2(tmem_slot, 512);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();

  const int num_kb = p.K / kG2BlockK;
  const int total_tiles = p.tiles_m * p.tiles_n;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs): as gemm2_body =====================
    if (lane == 0) {
      int pre = 0;
      {
        int stage = 0;
        for (int tile = pair; tile < total_tiles && pre < kG2gStages; tile += num_pairs) {
          int tm, tn;
          tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
          const int row_b = tn * kG2TileN + (int)cta * kG2RowsB;
          for (int kb = 0; kb < num_kb && pre < kG2gStages; ++kb, ++pre, ++stage) {
            const uint32_t full_leader = mapa_u32(smem_u32(&full_bar[stage]), 0);
            if (cta == 0) mbar_expect_tx(&full_bar[stage], 2 * kG2Stage);
            tma_load_2d_pair(smem_gen + stage * kG2Stage + kG2StageA, &tm_w, full_leader, kb * kG2BlockK, row_b);
          }
        }
      }
      pdl_wait();
      int stage = 0, it = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs) {
        int tm, tn;
        tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
        const int row_a = tm * 2 * kG2RowsA + (int)cta * kG2RowsA;
        const int row_b = tn * kG2TileN + (int)cta * kG2RowsB;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const uint32_t full_leader = mapa_u32(smem_u32(&full_bar[stage]), 0);
          uint8_t* sa = smem_gen + stage * kG2Stage;
          if (it >= pre) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (cta == 0) mbar_expect_tx(&full_bar[stage], 2 * kG2Stage);
            tma_load_2d_pair(sa + kG2StageA, &tm_w, full_leader, kb * kG2BlockK, row_b);
          }
          tma_load_2d_pair(sa, &tm_x, full_leader, kb * kG2BlockK, row_a);
          if (++stage == kG2gStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only): one accumulator per scale group ==========
    if (cta == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc<FMT>(2 * kG2RowsA, kG2TileN);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs) {
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * kG2TileN;
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem0 + stage * kG2Stage;
          const uint64_t a_desc = make_sw128_kmajor_desc(sa);
          const uint64_t b_desc = make_sw128_kmajor_desc(sa + kG2StageA);
#pragma unroll
          for (int k = 0; k < kG2BlockK / 32; ++k) {
            if constexpr (FMT == QL_INT8)
              mma2_i8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, k > 0);
            else
              mma2_f8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, k > 0);
          }
          tc_commit_pair(&empty_bar[stage]);
          if (++stage == kG2gStages) { stage = 0; phase ^= 1; }
          tc_commit_pair(&tfull_bar[acc]);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue: 16 warps, lane quarter q = warp % 4, column quarter cq ========
    const int ew = warp - 2;
    const int q = warp & 3, cq = ew >> 2;
    const int et = threadIdx.x - 64;                       // 0..511
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(cq * 64);
    const uint32_t tempty_leader0 = mapa_u32(smem_u32(&tempty_bar[0]), 0);
    const int panels = (p.G + kG2gPanel - 1) / kG2gPanel;
    const int pcol = et >> 1, phf = et & 1;                // panel loader: feature column, which 4 groups
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t pbuf = 0, bbuf = 0;                           // scale panel / bias buffers in use
    pdl_wait();   // sx and the y buffer belong to the previous kernels

    // scales of panel `pn` of the tile with feature origin `n0`: 4 consecutive groups of one column
    auto load_panel = [&](int n0, int pn) -> float4 {
      const int n = n0 + pcol, gi = pn * kG2gPanel + phf * 4;
      if (n < p.N && gi < p.G) return __ldg(reinterpret_cast<const float4*>(p.sw + (size_t)n * p.G + gi));
      return make_float4(0.f, 0.f, 0.f, 0.f);
    };
    auto store_panel = [&](uint32_t buf, const float4& v) {
      const uint32_t base = sw0 + buf * (kG2gPanel * kG2TileN * 4) + (uint32_t)(phf * 4 * kG2TileN + pcol) * 4u;
      sts32f(base, v.x);
      sts32f(base + kG2TileN * 4, v.y);
      sts32f(base + 2 * kG2TileN * 4, v.z);
      sts32f(base + 3 * kG2TileN * 4, v.w);
    };
    auto load_bias = [&](int n0) -> float {
      if constexpr (HAS_BIAS) {
        const int n = n0 + et;
        return (et < kG2TileN && n < p.N) ? bf16_lo(p.bias[n]) : 0.0f;
      } else {
        return 0.0f;
      }
    };

    // first tile's first panel (+ bias)
    if (pair < total_tiles) {
      int tm, tn;
      tile_coords(pair, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
      store_panel(0, load_panel(tn * kG2TileN, 0));
      if constexpr (HAS_BIAS) { if (et < kG2TileN) sts32f(bias0 + et * 4, load_bias(tn * kG2TileN)); }
    }
    named_bar_sync(1, 32 * kG2gEpiWarps);

    for (int tile = pair; tile < total_tiles; tile += num_pairs) {
      int tm, tn;
      tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
      const int n0 = tn * kG2TileN;
      const int m = tm * 2 * kG2RowsA + (int)cta * kG2RowsA + q * 32 + lane;
      // where the NEXT panel comes from once this tile's panels are used up
      int n0_next = -1;
      if (tile + num_pairs < total_tiles) {
        int tm2, tn2;
        tile_coords(tile + num_pairs, p.tiles_m, p.tiles_n, p.group_m, tm2, tn2);
        n0_next = tn2 * kG2TileN;
      }
      const float sxm = m < p.M ? __ldg(p.sx + m) : 0.0f;
      float t[64];
#pragma unroll
      for (int i = 0; i < 64; ++i) t[i] = 0.0f;

      for (int pn = 0; pn < panels; ++pn) {
        // the panel after this one (of this tile, or panel 0 of the pair's next tile): in flight
        // as registers while this panel is used
        const bool last_panel = pn + 1 == panels;
        const bool have_next = !last_panel || n0_next >= 0;
        float4 nxt = make_float4(0.f, 0.f, 0.f, 0.f);
        float nbias = 0.0f;
        if (have_next) nxt = last_panel ? load_panel(n0_next, 0) : load_panel(n0, pn + 1);
        if (last_panel && n0_next >= 0) nbias = load_bias(n0_next);
        const int g_end = min(p.G, (pn + 1) * kG2gPanel);
        for (int gk = pn * kG2gPanel * p.kb_per_g; gk < g_end * p.kb_per_g; ++gk) {
          // K block gk of the tile belongs to scale group gk / kb_per_g; every 128-byte block is
          // promoted on its own (the engine's per-group semantics: oracle gemm_groups, gp = 128)
          const int g = gk / p.kb_per_g;
          const uint32_t srow = sw0 + pbuf * (kG2gPanel * kG2TileN * 4) +
                                (uint32_t)((g - pn * kG2gPanel) * kG2TileN + cq * 64) * 4u;
          mbar_wait(&tfull_bar[acc], acc_phase);
          tc_fence_after();
          uint32_t r[2][16];
          const uint32_t tadr = lane_addr + acc * kG2TileN;
          tmem_ld16(tadr, r[0]);
          tmem_ld_wait_regs16(r[0]);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            if (c + 1 < 4) tmem_ld16(tadr + (c + 1) * 16, r[(c + 1) & 1]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float4 s4 = lds128f(srow + (uint32_t)(c * 16 + j * 4) * 4u);
              const float sv[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const uint32_t raw = r[c & 1][j * 4 + e];
                float a;
                if constexpr (FMT == QL_INT8) {
                  // float(P) by integer add + FADD: |P| <= 128 * 127^2 < 2^22, so the integer
                  // add into the mantissa of 1.5 * 2^23 and the subtraction are both exact
                  a = __fsub_rn(__uint_as_float(raw + 0x4B400000u), 12582912.0f);
                } else {
                  a = __uint_as_float(raw);
                }
                t[c * 16 + j * 4 + e] = fmaf(a, sv[e], t[c * 16 + j * 4 + e]);
              }
            }
            if (c + 1 < 4) tmem_ld_wait_regs16(r[(c + 1) & 1]);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(tempty_leader0 + acc * 8);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (have_next) {
          store_panel(pbuf ^ 1, nxt);
          if constexpr (HAS_BIAS) {
            if (last_panel && et < kG2TileN) sts32f(bias0 + (bbuf ^ 1) * (kG2TileN * 4) + et * 4, nbias);
          }
        }
        if (!last_panel) {
          named_bar_sync(1, 32 * kG2gEpiWarps);   // everyone is done with this panel, the next one is staged
          pbuf ^= 1;
        }
      }

      // ---- the tile is complete: (t * sx[m]) (+ bias[n]) -> bf16 -> 128 contiguous bytes of the token row
      const uint32_t brow = bias0 + bbuf * (kG2TileN * 4) + (uint32_t)(cq * 64) * 4u;
      const int nb0 = n0 + cq * 64;
      uint16_t* yrow = p.y + (int64_t)m * p.ldy + nb0;
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __fmul_rn(t[c8 * 8 + e], sxm);
        if constexpr (HAS_BIAS) {
          const float4 b0 = lds128f(brow + (uint32_t)(c8 * 8) * 4u), b1 = lds128f(brow + (uint32_t)(c8 * 8 + 4) * 4u);
          v[0] = __fadd_rn(v[0], b0.x); v[1] = __fadd_rn(v[1], b0.y); v[2] = __fadd_rn(v[2], b0.z); v[3] = __fadd_rn(v[3], b0.w);
          v[4] = __fadd_rn(v[4], b1.x); v[5] = __fadd_rn(v[5], b1.y); v[6] = __fadd_rn(v[6], b1.z); v[7] = __fadd_rn(v[7], b1.w);
        }
        const int nb = nb0 + c8 * 8;
        if (m < p.M) {
          if (nb + 8 <= p.N) {
            stg128(yrow + c8 * 8, pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                   pack_bf16x2(v[6], v[7]));
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e)
              if (nb + e < p.N) yrow[c8 * 8 + e] = f2bf(v[e]);
          }
        }
      }
      // the next tile's first panel (and bias) was staged above: publish it
      named_bar_sync(1, 32 * kG2gEpiWarps);
      pbuf ^= 1;
      bbuf ^= 1;
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc2(tmem_base, 512);
}

// ------------------------------------------------------------------ host side

template <int FMT, bool OUT_ACC, bool HAS_BIAS>
static int launch_g2(const void* qx, const void* qw, void* y, int64_t ldy, Gemm2Params p,
                     cudaStream_t st) {
  auto kern = gemm2_kernel<FMT, OUT_ACC, HAS_BIAS>;
  static SmemAttrOnce smem_once;   // per kernel instantiation, per device ordinal
  if (int e = ensure_max_smem(reinterpret_cast<const void*>(kern), kG2Smem, smem_once)) return e;
  CUtensorMap tm_x, tm_w, tm_y;
  if (!make_tmap_2d(&tm_x, qx, CU_TENSOR_MAP_DATA_TYPE_UINT8, p.K, p.M, p.K, kG2BlockK, kG2RowsA) ||
      !make_tmap_2d(&tm_w, qw, CU_TENSOR_MAP_DATA_TYPE_UINT8, p.K, p.N, p.K, kG2BlockK, kG2RowsB))
    return QL_ERR_CUDA;
  if constexpr (OUT_ACC) {
    tm_y = tm_x;  // unused by the kernel
  } else {
    if (!make_tmap_2d(&tm_y, y, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p.N, p.M, ldy * 2, kG2StoreCols, 32))
      return QL_ERR_CUDA;
  }
  p.tiles_m = (p.M + 2 * kG2RowsA - 1) / (2 * kG2RowsA);
  p.tiles_n = (p.N + kG2TileN - 1) / kG2TileN;
  // keep a super-group's activations (group_m x 256 x K bytes) within ~48 MB of the 126 MB L2
  {
    const int64_t g = (48ll << 20) / (256ll * p.K);
    p.group_m = (int)(g < 1 ? 1 : (g > p.tiles_m ? p.tiles_m : g));
  }
  const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
  // A persistent kernel must not launch more clusters than can be co-resident: a GPC with an
  // odd number of free SMs strands one, so ask the driver instead of assuming num_sms / 2.
  static int pairs_max = 0;
  if (pairs_max == 0) {
    cudaLaunchConfig_t qc = {};
    qc.gridDim = dim3(num_sms() & ~1);
    qc.blockDim = dim3(kG2Threads);
    qc.dynamicSmemBytes = kG2Smem;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &qc) != cudaSuccess || n < 1) {
      (void)cudaGetLastError();
      n = num_sms() / 2;
    }
    pairs_max = n;
    if (getenv("QL_DEBUG") != nullptr)
      fprintf(stderr, "[qlinear_b200] gemm2: %d co-resident CTA pairs on %d SMs\n", n, num_sms());
  }
  const int pairs = (int)(tiles < pairs_max ? tiles : pairs_max);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * pairs));
  cfg.blockDim = dim3(kG2Threads);
  cfg.dynamicSmemBytes = kG2Smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, kern, tm_x, tm_w, tm_y, p);
  count_launch();
  if (err != cudaSuccess) {
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  return QL_OK;
}

static int g2_pairs_max(const void* kern) {
  static int pairs_max = 0;
  if (pairs_max == 0) {
    cudaLaunchConfig_t qc = {};
    qc.gridDim = dim3(num_sms() & ~1);
    qc.blockDim = dim3(kG2Threads);
    qc.dynamicSmemBytes = kG2Smem;
    int n = 0;   // the kernel carries __cluster_dims__(2, 1, 1)
    if (cudaOccupancyMaxActiveClusters(&n, kern, &qc) != cudaSuccess || n < 1) {
      (void)cudaGetLastError();
      n = num_sms() / 2;
    }
    pairs_max = n;
  }
  return pairs_max;
}

template <int FMT, bool HAS_BIAS>
static int launch_g2_push(const TpView& tp, const void* qx, const void* qw, Gemm2Params p, cudaStream_t st) {
  auto kern = gemm2_push_kernel<FMT, HAS_BIAS>;
  static SmemAttrOnce smem_once;
  if (int e = ensure_max_smem(reinterpret_cast<const void*>(kern), kG2Smem, smem_once)) return e;
  CUtensorMap tm_x, tm_w;
  if (!make_tmap_2d(&tm_x, qx, CU_TENSOR_MAP_DATA_TYPE_UINT8, p.K, p.M, p.K, kG2BlockK, kG2RowsA) ||
      !make_tmap_2d(&tm_w, qw, CU_TENSOR_MAP_DATA_TYPE_UINT8, p.K, p.N, p.K, kG2BlockK, kG2RowsB))
    return QL_ERR_CUDA;
  Gemm2Push push;
  const int64_t Mo = tp_rows_per_owner(p.M, tp.world);
  for (int o = 0; o < 8; ++o) {
    if (o >= tp.world) { push.ymap[o] = tm_x; continue; }
    int64_t rows_o = (int64_t)p.M - o * Mo;
    rows_o = rows_o < 1 ? 1 : (rows_o > Mo ? Mo : rows_o);       // an owner without rows is never addressed
    uint8_t* base = tp.big[o] + tp.lay.part + (size_t)((int64_t)tp.rank * Mo) * (size_t)p.N * 2;
    if (!make_tmap_2d(&push.ymap[o], base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p.N, rows_o, (int64_t)p.N * 2,
                      kG2StoreCols, 32))
      return QL_ERR_CUDA;
  }
  push.tiles_per_owner = (int)(Mo / (2 * kG2RowsA));
  push.rot = 0;   // set below, once tiles_m is known
  p.tiles_m = (p.M + 2 * kG2RowsA - 1) / (2 * kG2RowsA);
  p.tiles_n = (p.N + kG2TileN - 1) / kG2TileN;
  // one owner's token tiles form a rasterisation group: the grid works through the owners one
  // after the other (all feature tiles of owner r+1, then r+2, ..., its own last)
  p.group_m = push.tiles_per_owner < p.tiles_m ? push.tiles_per_owner : p.tiles_m;
  {
    const int64_t first = (int64_t)((tp.rank + 1) % tp.world) * push.tiles_per_owner;
    push.rot = first < p.tiles_m ? (int)first : 0;       // an owner without tokens: start at owner 0
  }
  const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
  const int pairs_max = g2_pairs_max(reinterpret_cast<const void*>(kern));
  const int pairs = (int)(tiles < pairs_max ? tiles : pairs_max);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * pairs));
  cfg.blockDim = dim3(kG2Threads);
  cfg.dynamicSmemBytes = kG2Smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, kern, tm_x, tm_w, push, p);
  count_launch();
  if (err != cudaSuccess) {
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  return QL_OK;
}

int launch_gemm2_push(const TpView& tp, const void* qx, const float* sx, const void* qw, const float* sw,
                      const void* bias, int64_t M, int64_t N, int64_t K, int fmt, cudaStream_t st) {
  if (M == 0 || N == 0) return QL_OK;
  Gemm2Params p;
  p.sx = sx; p.sw = sw;
  p.bias = static_cast<const uint16_t*>(bias);
  p.acc_out = nullptr;
  p.trace = nullptr;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.tiles_m = p.tiles_n = 0;
  p.group_m = 1;
  if (fmt == QL_INT8)
    return bias ? launch_g2_push<QL_INT8, true>(tp, qx, qw, p, st) : launch_g2_push<QL_INT8, false>(tp, qx, qw, p, st);
  return bias ? launch_g2_push<QL_FP8_E4M3, true>(tp, qx, qw, p, st)
              : launch_g2_push<QL_FP8_E4M3, false>(tp, qx, qw, p, st);
}

bool gemm2_eligible(const void* y, int64_t ldy, const void* sw, const void* bias, bool out_acc) {
  (void)sw; (void)bias;
  if (out_acc) return true;
  return (reinterpret_cast<uintptr_t>(y) & 15) == 0 && ldy % 8 == 0;   // TMA store: 16-byte pitch
}

int launch_gemm2(const void* qx, const float* sx, const void* qw, const float* sw,
                 const void* bias, void* y, int64_t ldy, int32_t* out_acc, int64_t M, int64_t N,
                 int64_t K, int fmt, cudaStream_t st) {
  if (M == 0 || N == 0) return QL_OK;
  Gemm2Params p;
  p.sx = sx; p.sw = sw;
  p.bias = static_cast<const uint16_t*>(bias);
  p.acc_out = out_acc;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.tiles_m = p.tiles_n = 0;
  p.group_m = 1;
  {
    // debug timeline: launch i of a traced sequence stamps slot i % 8 of the registered buffer
    static unsigned launch_no = 0;
    long long* tr = skinny_trace_ptr();
    p.trace = tr != nullptr ? tr + (size_t)(launch_no++ % 8) * 148 * 8 : nullptr;
  }
  if (out_acc != nullptr) return launch_g2<QL_INT8, true, false>(qx, qw, y, ldy, p, st);
  if (fmt == QL_INT8)
    return bias ? launch_g2<QL_INT8, false, true>(qx, qw, y, ldy, p, st)
                : launch_g2<QL_INT8, false, false>(qx, qw, y, ldy, p, st);
  return bias ? launch_g2<QL_FP8_E4M3, false, true>(qx, qw, y, ldy, p, st)
              : launch_g2<QL_FP8_E4M3, false, false>(qx, qw, y, ldy, p, st);
}

// ---- per-group scales on the CTA pair -------------------------------------------------------

template <int FMT, bool HAS_BIAS>
static int launch_g2g(const void* qx, const void* qw, Gemm2gParams p, cudaStream_t st) {
  auto kern = gemm2g_kernel<FMT, HAS_BIAS>;
  static SmemAttrOnce smem_once;
  if (int e = ensure_max_smem(reinterpret_cast<const void*>(kern), kG2gSmem, smem_once)) return e;
  CUtensorMap tm_x, tm_w;
  if (!make_tmap_2d(&tm_x, qx, CU_TENSOR_MAP_DATA_TYPE_UINT8, p.K, p.M, p.K, kG2BlockK, kG2RowsA) ||
      !make_tmap_2d(&tm_w, qw, CU_TENSOR_MAP_DATA_TYPE_UINT8, p.K, p.N, p.K, kG2BlockK, kG2RowsB))
    return QL_ERR_CUDA;
  p.tiles_m = (p.M + 2 * kG2RowsA - 1) / (2 * kG2RowsA);
  p.tiles_n = (p.N + kG2TileN - 1) / kG2TileN;
  {
    const int64_t g = (48ll << 20) / (256ll * p.K);
    p.group_m = (int)(g < 1 ? 1 : (g > p.tiles_m ? p.tiles_m : g));
  }
  const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
  static int pairs_max = 0;
  if (pairs_max == 0) {
    cudaLaunchConfig_t qc = {};
    qc.gridDim = dim3(num_sms() & ~1);
    qc.blockDim = dim3(kG2gThreads);
    qc.dynamicSmemBytes = kG2gSmem;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &qc) != cudaSuccess || n < 1) {
      (void)cudaGetLastError();
      n = num_sms() / 2;
    }
    pairs_max = n;
  }
  const int pairs = (int)(tiles < pairs_max ? tiles : pairs_max);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * pairs));
  cfg.blockDim = dim3(kG2gThreads);
  cfg.dynamicSmemBytes = kG2gSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, kern, tm_x, tm_w, p);
  count_launch();
  if (err != cudaSuccess) {
    if (getenv("QL_DEBUG") != nullptr)
      fprintf(stderr, "[qlinear_b200] gemm2g launch failed: %s\n", cudaGetErrorString(err));
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  return QL_OK;
}

// float4 scale-panel loads: 4 groups at a time, 16-byte aligned rows of sw
bool gemm2_grouped_eligible(const void* y, int64_t ldy, const void* sw, int64_t K, int64_t group) {
  if (group <= 0 || group % kG2BlockK != 0 || K % group != 0) return false;
  const int64_t G = K / group;
  return G % 4 == 0 && (reinterpret_cast<uintptr_t>(sw) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0 &&
         ldy % 8 == 0;
}

int launch_gemm2_grouped(const void* qx, const float* sx, const void* qw, const float* sw, const void* bias,
                         void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int fmt, int64_t group,
                         cudaStream_t st) {
  if (M == 0 || N == 0) return QL_OK;
  Gemm2gParams p;
  p.sx = sx; p.sw = sw;
  p.bias = static_cast<const uint16_t*>(bias);
  p.y = static_cast<uint16_t*>(y);
  p.ldy = ldy;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.G = (int)(K / group);
  p.kb_per_g = (int)(group / kG2BlockK);
  p.tiles_m = p.tiles_n = 0;
  p.group_m = 1;
  if (fmt == QL_INT8)
    return bias ? launch_g2g<QL_INT8, true>(qx, qw, p, st) : launch_g2g<QL_INT8, false>(qx, qw, p, st);
  return bias ? launch_g2g<QL_FP8_E4M3, true>(qx, qw, p, st) : launch_g2g<QL_FP8_E4M3, false>(qx, qw, p, st);
}

}  // namespace ql
