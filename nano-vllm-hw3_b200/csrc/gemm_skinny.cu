// Decode-regime W8A8 linear (M <= 256 tokens): a weight-streaming tcgen05 kernel with the
// K dimension split over a thread-block CLUSTER and the partial accumulators reduced through
// distributed shared memory.  Replaces per_row_quant + _per_row_post_process_mm of the
// reference's per_row_matmul (nanovllm/utils/quantization.py:310-346; fixed 128^3 Triton tile on
// N/128 CTAs -> 8..48 CTAs for a decode batch) and per_group_matmul (:265-307).
//
// This regime is HBM-bound: algorithmic bytes = N*K weights (+ scales, + M*(2K + 2N)
// activations); the tensor core is idle most of the time.  What matters is (1) every SM
// streaming weights from the first cycle, (2) enough bytes in flight, (3) one launch.
//
//   * Orientation as gemm_tcgen05.cu (swap-AB): UMMA A = 128 weight rows, UMMA B = BT tokens
//     (BT = 16..256 = UMMA N), accumulator D[128 x BT] in TMEM, weight row on the TMEM lane.
//   * grid = ceil(N/128) weight tiles x S cluster ranks; rank r streams the K blocks
//     [r*U/S, (r+1)*U/S) of its weight tile through a TMA ring (16 KB per 128-byte K block),
//     so N/128 = 8..48 tiles still occupy 148 SMs (two CTAs per SM are co-resident: <= 110 KB of
//     shared memory and <= 256 TMEM columns each).
//   * FUSED (M <= 16): the activations are quantised inside the kernel, by the epilogue warps,
//     while the weight ring fills: each rank reads only ITS K slice of x (bf16), the row amax is
//     combined across the cluster through DSMEM (remote st.shared::cluster + remote mbarrier
//     arrive), and the codes are written straight into the UMMA's 128B-swizzled K-major B layout.
//     Arithmetic is the Triton quantiser's (:26-41), bit for bit.  Otherwise (16 < M) qx comes
//     from quant_rows_kernel through TMA alongside the weights.
//   * split-K reduction = reduce-scatter over DSMEM: token column c of every rank's partial
//     accumulator goes to rank c / (BT/S); that rank adds the S partials in rank order (int32:
//     exact, so INT8 per-row results stay bit-identical to the unsplit GEMM; fp32: fixed order,
//     deterministic), applies (float(acc) * sx[m]) * sw[n] (+ bias[n]) -- reference product order
//     :111-121 -- and stores 16 bytes (8 features) of one token per thread.
//   * per-group (g multiple of 128): one TMEM accumulator per scale group, promoted by the
//     epilogue warps with t = fmaf(float(P_j), sw[n,j], t) while the tensor core fills the next.
//   * programmatic dependent launch: weights do not depend on the previous kernel, so the
//     producer starts streaming them before griddepcontrol.wait; only the activation reads and
//     the output stores sit behind it.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace ql {

struct SkinnyParams {
  const uint16_t* x;     // FUSED: bf16 [M, ldx]
  int64_t ldx;
  const float* sx;       // !FUSED: [M]
  const float* sw;       // [N] or [N, G]
  const uint16_t* bias;  // bf16 [N] or nullptr
  uint16_t* y;           // bf16 [M, ldy]
  int64_t ldy;
  int32_t* acc_out;      // OUT_ACC: int32 [M, N]
  int M, N, K;
  int G;                 // scale columns (1 for per-row)
  int kb_per_g;          // 128-byte K blocks per split unit (= per scale group)
  int units;             // split units along K
  int stages;            // ring depth (<= kSkMaxStages)
  int ring_bytes;        // bytes reserved for the ring (>= reduction buffer when aliased)
  int bres_bytes;        // FUSED: resident B operand bytes
  long long* trace;      // debug: clock64() stamps of CTA 0 (ql_debug_set_trace), else nullptr
};

// Debug timeline (ql_debug_set_trace): every CTA stamps clock64() into trace[blockIdx.x * 16 + i];
// slot 15 holds globaltimer at entry so CTAs on different SMs can be aligned.  The buffer must
// hold 16 int64 per CTA of the largest grid launched while tracing is on.
#define SK_TRACE(i)                                                              \
  do {                                                                           \
    if (p.trace != nullptr) {                                                    \
      p.trace[(size_t)blockIdx.x * 16 + (i)] = clock64();                        \
      if ((i) == 0) p.trace[(size_t)blockIdx.x * 16 + 15] = (long long)globaltimer_ns(); \
    }                                                                            \
  } while (0)

constexpr int kSkThreads = 256;
constexpr int kSkTileN = 128;
constexpr int kSkBlockK = 128;
constexpr int kSkStageA = kSkTileN * kSkBlockK;  // 16 KB
constexpr int kSkMaxStages = 8;
constexpr int kSkSmall = 1024;                   // barriers, tmem slot, scales, amax exchange

template <bool GROUPED, int BT> struct SkCfg {
  static constexpr int nacc_raw = GROUPED ? 256 / BT : 1;
  static constexpr int nacc = nacc_raw > 4 ? 4 : (nacc_raw < 1 ? 1 : nacc_raw);
  static constexpr int cols_raw = nacc * BT;
  static constexpr int tmem_cols = cols_raw <= 32 ? 32 : cols_raw <= 64 ? 64 : cols_raw <= 128 ? 128
                                   : cols_raw <= 256 ? 256 : 512;
  static constexpr bool two_per_sm = BT <= 64 || (BT == 128 && !GROUPED);
  static constexpr int min_blocks = two_per_sm ? 2 : 1;
  static constexpr int recv_bytes = kSkTileN * BT * 4;
  static constexpr bool alias_recv = BT > 32;      // reduction buffer reuses the (drained) ring
  static constexpr int CH = BT < 32 ? BT : 32;     // TMEM columns per tcgen05.ld
};

template <int CH>
__device__ __forceinline__ void tmem_ld_ch(uint32_t taddr, uint32_t (&r)[CH]) {
  if constexpr (CH == 16) tmem_ld16(taddr, r);
  else tmem_ld32(taddr, r);
}

template <int FMT, bool GROUPED, int BT, bool FUSED, bool OUT_ACC>
__global__ void __launch_bounds__(kSkThreads, SkCfg<GROUPED, BT>::min_blocks)
skinny_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
              const SkinnyParams p) {
  using Cfg = SkCfg<GROUPED, BT>;
  constexpr int NACC = Cfg::nacc;
  constexpr int CH = Cfg::CH;
  static_assert(!FUSED || BT == 16, "fused quantisation: one 16-token tile");
  static_assert(!(OUT_ACC && (GROUPED || FMT != QL_INT8)), "raw accumulators: int8 per-row only");
  constexpr int STAGE = kSkStageA + (FUSED ? 0 : BT * kSkBlockK);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* ring = smem;
  uint8_t* bres = ring + p.ring_bytes;                                   // FUSED only
  uint8_t* recv_ded = bres + p.bres_bytes;                               // !alias_recv only
  uint8_t* small = recv_ded + (Cfg::alias_recv ? 0 : Cfg::recv_bytes);
  uint32_t* recv = reinterpret_cast<uint32_t*>(Cfg::alias_recv ? ring : recv_ded);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(small);
  uint64_t* empty_bar = full_bar + kSkMaxStages;
  uint64_t* tfull_bar = empty_bar + kSkMaxStages;
  uint64_t* tempty_bar = tfull_bar + 4;
  uint64_t* bready_bar = tempty_bar + 4;
  uint64_t* amax_bar = bready_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(amax_bar + 1);
  float* sx_s = reinterpret_cast<float*>(small + 256);                   // [16]  FUSED
  float* pmax_s = sx_s + 16;                                             // [16]
  float* amax_part = pmax_s + 16;                                        // [8][16]
  float* inv_s = amax_part + 128;                                        // [16]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) SK_TRACE(0);
  const uint32_t S = cluster_nctarank();
  const uint32_t rank = cluster_ctarank();
  const int tile_n = blockIdx.x / S;
  const int STAGES = p.stages;

  // this rank's K range, in split units and in 128-byte blocks
  const int u_begin = (int)((int64_t)p.units * rank / S), u_end = (int)((int64_t)p.units * (rank + 1) / S);
  const int kb_begin = u_begin * p.kb_per_g, kb_end = u_end * p.kb_per_g;
  const int nkb = kb_end - kb_begin;

  // Warp 0 owns the ring: it initialises the ring barriers itself and starts streaming weights
  // (which never depend on the previous kernel) before the rest of the CTA has finished its setup.
  const int pre = nkb < STAGES ? nkb : STAGES;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_w);
    if constexpr (!FUSED) tma_prefetch_desc(&tm_x);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_fence_init();
    for (int i = 0; i < pre; ++i) {
      mbar_expect_tx(&full_bar[i], STAGE);
      tma_load_2d(ring + i * STAGE, &tm_w, &full_bar[i], (kb_begin + i) * kSkBlockK,
                  tile_n * kSkTileN, kEvictFirst);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < NACC; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], 4);
    }
    mbar_init(bready_bar, 4);
    mbar_init(amax_bar, S);
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, Cfg::tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) SK_TRACE(1);
  if (S > 1) cluster_arrive();           // #1: "my barriers and buffers exist" (waited before any DSMEM op)
  pdl_launch_dependents();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      if constexpr (!FUSED) {
        pdl_wait();
        for (int i = 0; i < pre; ++i)
          tma_load_2d(ring + i * STAGE + kSkStageA, &tm_x, &full_bar[i], (kb_begin + i) * kSkBlockK, 0,
                      kEvictLast);
      }
      SK_TRACE(2);
      int stage = pre == STAGES ? 0 : pre;
      uint32_t phase = pre == STAGES ? 1 : 0;
      for (int i = pre; i < nkb; ++i) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = ring + stage * STAGE;
        mbar_expect_tx(&full_bar[stage], STAGE);
        tma_load_2d(sa, &tm_w, &full_bar[stage], (kb_begin + i) * kSkBlockK, tile_n * kSkTileN,
                    kEvictFirst);
        if constexpr (!FUSED)
          tma_load_2d(sa + kSkStageA, &tm_x, &full_bar[stage], (kb_begin + i) * kSkBlockK, 0, kEvictLast);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc<FMT>(kSkTileN, BT);
      if constexpr (FUSED) {
        mbar_wait(bready_bar, 0);
        tc_fence_after();
      }
      SK_TRACE(7);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int i = 0; i < nkb; ++i) {
        const int in_unit = i % p.kb_per_g;
        if constexpr (GROUPED) {
          if (in_unit == 0) {
            mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
            tc_fence_after();
          }
        }
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (i == 0) SK_TRACE(8);
        const uint32_t sa = smem_u32(ring + stage * STAGE);
        const uint64_t a_desc = make_sw128_kmajor_desc(sa);
        const uint64_t b_desc = make_sw128_kmajor_desc(FUSED ? smem_u32(bres) + i * (BT * kSkBlockK)
                                                             : sa + kSkStageA);
        const uint32_t d_tmem = tmem_base + acc * BT;
#pragma unroll
        for (int k = 0; k < kSkBlockK / 32; ++k) {
          const uint32_t accumulate = GROUPED ? (in_unit > 0 || k > 0) : (i > 0 || k > 0);
          if constexpr (FMT == QL_INT8)
            mma_i8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, accumulate);
          else
            mma_f8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, accumulate);
        }
        tc_commit(&empty_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
        if constexpr (GROUPED) {
          if (in_unit == p.kb_per_g - 1) {
            tc_commit(&tfull_bar[acc]);
            if (++acc == NACC) { acc = 0; acc_phase ^= 1; }
          }
        }
      }
      if constexpr (!GROUPED) tc_commit(&tfull_bar[0]);
      SK_TRACE(9);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;            // TMEM lane quarter / quantiser warp index
    const int e = threadIdx.x - 128;   // 0..127
    bool waited1 = false;              // cluster barrier #1 already waited by this thread

    // ===================== fused activation quantiser (M <= 16) =====================
    if constexpr (FUSED) {
      // Work item = 32 bytes of one activation row (16 bf16 -> one 16-byte chunk of a swizzled
      // 128-byte B row).  Items are dealt round-robin to the 128 threads, kQU items (2*kQU
      // 16-byte loads) in flight per thread, so the cost is a few L2 round trips whatever M is.
      // Every aligned group of 8 lanes works on one row and one K block.
      constexpr float qmax = Fmt<FMT>::qmax;
      constexpr int kQU = 4;
      int* pmax_i = reinterpret_cast<int*>(pmax_s);
      if (e < 16) pmax_i[e] = 0;         // |x| >= 0: float order == int order of the bit patterns
      asm volatile("bar.sync 1, 128;" ::: "memory");
      pdl_wait();
      if (e == 0) SK_TRACE(3);
      const int nit = nkb * 8;           // items per row in this rank's K slice
      const int items = p.M * nit;
      const uint16_t* xk = p.x + (int64_t)kb_begin * kSkBlockK;
      // pass 1: amax of this rank's slice of every row
#pragma unroll 1
      for (int base = 0; base < items; base += 128 * kQU) {
        uint4 v[kQU][2];
        int mm[kQU];
#pragma unroll
        for (int j = 0; j < kQU; ++j) {
          const int idx = base + j * 128 + e;
          const bool ok = idx < items;
          const int m = ok ? idx / nit : 0;
          const int it = ok ? idx - m * nit : 0;
          const uint4* src = reinterpret_cast<const uint4*>(xk + (int64_t)m * p.ldx + it * 16);
          mm[j] = ok ? m : -1;
          v[j][0] = ok ? __ldg(src) : make_uint4(0, 0, 0, 0);
          v[j][1] = ok ? __ldg(src + 1) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int j = 0; j < kQU; ++j) {
          float a = fmaxf(amax8(v[j][0]), amax8(v[j][1]));
          a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, 1));
          a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, 2));
          a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, 4));
          if ((lane & 7) == 0 && mm[j] >= 0) atomicMax(&pmax_i[mm[j]], __float_as_int(a));
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (e == 0) SK_TRACE(4);
      if (S > 1) {
        cluster_wait();                // #1: every peer's amax_bar / amax_part exist
        waited1 = true;
        if (e < (int)S) {
          const uint32_t dst = mapa_u32(smem_u32(amax_part + rank * 16), (uint32_t)e);
#pragma unroll
          for (int m = 0; m < 16; ++m) st_cluster_u32(dst + m * 4, __float_as_uint(pmax_s[m]));
          mbar_arrive_release_cluster(mapa_u32(smem_u32(amax_bar), (uint32_t)e));
        }
        mbar_wait_acq_cluster(amax_bar, 0);
      }
      if (e == 0) SK_TRACE(5);
      if (e < 16) {
        float amax = pmax_s[e];
        if (S > 1) {
          amax = amax_part[e];
          for (uint32_t s = 1; s < S; ++s) amax = fmaxf(amax, amax_part[s * 16 + e]);
        }
        amax = fmaxf(amax, 1e-8f);
        inv_s[e] = __fdiv_rn(qmax, amax);
        sx_s[e] = e < p.M ? __fdiv_rn(amax, qmax) : 0.0f;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // pass 2: quantise into the UMMA B layout (K-major, 128B swizzle, 8-row groups of 1 KB).
      // Rows >= M of the 16-token operand stay unwritten: a B row only feeds its own
      // accumulator column, and columns >= M are never read back.
#pragma unroll 1
      for (int base = 0; base < items; base += 128 * kQU) {
        uint4 v[kQU][2];
        int mm[kQU], ii[kQU];
#pragma unroll
        for (int j = 0; j < kQU; ++j) {
          const int idx = base + j * 128 + e;
          const bool ok = idx < items;
          const int m = ok ? idx / nit : 0;
          const int it = ok ? idx - m * nit : 0;
          const uint4* src = reinterpret_cast<const uint4*>(xk + (int64_t)m * p.ldx + it * 16);
          mm[j] = ok ? m : -1;
          ii[j] = it;
          v[j][0] = ok ? __ldg(src) : make_uint4(0, 0, 0, 0);
          v[j][1] = ok ? __ldg(src + 1) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int j = 0; j < kQU; ++j) {
          if (mm[j] < 0) continue;
          const int m = mm[j];
          const float inv = inv_s[m];
          const uint2 q0 = quant8_mul<FMT>(v[j][0], inv), q1 = quant8_mul<FMT>(v[j][1], inv);
          uint8_t* dst = bres + (m >> 3) * 1024 + (m & 7) * 128 + (ii[j] >> 3) * (BT * kSkBlockK) +
                         (((ii[j] & 7) ^ (m & 7)) << 4);
          *reinterpret_cast<uint4*>(dst) = make_uint4(q0.x, q0.y, q1.x, q1.y);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bready_bar);
      if (e == 0) SK_TRACE(6);
    }

    // ===================== epilogue =====================
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int n_lane = tile_n * kSkTileN + q * 32 + lane;   // the weight row on this thread's TMEM lane
    const int lgS = 31 - __clz((int)S);                     // S is a power of two
    const int CPB = BT >> lgS;                              // token columns reduced by each rank
    const int lgC = 31 - __clz(CPB);
    const uint32_t recv_local = smem_u32(recv);
    // final-store coordinates; the weight scales / bias are static, so their loads are issued
    // here, long before the accumulator is ready, instead of after the reduce barrier
    const int fg = e & 15, tp = e >> 4;
    const int nb = tile_n * kSkTileN + fg * 8;
    const bool vec_ok = !OUT_ACC && (p.ldy % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.y) & 15) == 0);
    float swv[8], bv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const bool ok = nb + j < p.N;
      swv[j] = (!GROUPED && !OUT_ACC && ok) ? __ldg(p.sw + nb + j) : 0.0f;
      bv[j] = (p.bias != nullptr && ok) ? bf16_lo(p.bias[nb + j]) : 0.0f;
    }
    const bool has_bias = p.bias != nullptr;

    // column `col` of this rank's partial -> slot [rank][col % CPB][n] of rank col / CPB
    auto scatter = [&](int col, uint32_t v) {
      const int dest = col >> lgC, cl = col & (CPB - 1);
      const uint32_t off = (uint32_t)(((int)rank * CPB + cl) * kSkTileN + q * 32 + lane) * 4u;
      st_cluster_u32(mapa_u32(recv_local + off, (uint32_t)dest), v);
    };
    auto before_scatter = [&]() {
      if (S > 1) {
        if (!waited1) { cluster_wait(); waited1 = true; }
        if constexpr (Cfg::alias_recv) {   // #2: every rank's ring is drained before it is overwritten
          cluster_arrive();
          cluster_wait();
        }
      } else if constexpr (Cfg::alias_recv) {
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
    };

    if constexpr (!GROUPED) {
      mbar_wait(&tfull_bar[0], 0);
      tc_fence_after();
      if (e == 0) SK_TRACE(10);
      before_scatter();
#pragma unroll 1
      for (int c0 = 0; c0 < BT; c0 += CH) {
        if (c0 >= p.M) break;
        uint32_t r[CH];
        tmem_ld_ch<CH>(lane_addr + c0, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < CH; ++i)
          if (c0 + i < p.M) scatter(c0 + i, r[i]);
      }
    } else {
      float t[BT];
#pragma unroll
      for (int i = 0; i < BT; ++i) t[i] = 0.0f;
      const float* swrow = p.sw + (int64_t)(n_lane < p.N ? n_lane : 0) * p.G;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int u = u_begin; u < u_end; ++u) {
        const float sc = __ldg(swrow + u);
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < BT; c0 += CH) {
          uint32_t r[CH];
          tmem_ld_ch<CH>(lane_addr + acc * BT + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < CH; ++i) {
            const float a = FMT == QL_INT8 ? __int2float_rn((int)r[i]) : __uint_as_float(r[i]);
            t[c0 + i] = fmaf(a, sc, t[c0 + i]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        if (++acc == NACC) { acc = 0; acc_phase ^= 1; }
      }
      before_scatter();
#pragma unroll
      for (int i = 0; i < BT; ++i)
        if (i < p.M) scatter(i, __float_as_uint(t[i]));
    }

    // reduce-scatter complete once every rank has stored its partials
    if (e == 0) SK_TRACE(11);
    if (S > 1) {
      cluster_arrive();                // #3
      cluster_wait();
    } else {
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }

    if (e == 0) SK_TRACE(12);
    // ---- final: sum the S partials of my CPB token columns, scale, store ----------------------
    if constexpr (!FUSED) pdl_wait();  // sx (and the y buffer) belong to the previous kernels
    for (int c = tp; c < CPB; c += 8) {
      const int m = (int)rank * CPB + c;
      if (m >= p.M) break;
      uint32_t sum[8];
      {
        const uint4* src = reinterpret_cast<const uint4*>(recv + (size_t)c * kSkTileN + fg * 8);
        const uint4 a = src[0], b = src[1];
        sum[0] = a.x; sum[1] = a.y; sum[2] = a.z; sum[3] = a.w;
        sum[4] = b.x; sum[5] = b.y; sum[6] = b.z; sum[7] = b.w;
      }
      for (uint32_t s = 1; s < S; ++s) {
        const uint4* src =
            reinterpret_cast<const uint4*>(recv + ((size_t)s * CPB + c) * kSkTileN + fg * 8);
        const uint4 a = src[0], b = src[1];
        const uint32_t o[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if constexpr (FMT == QL_INT8 && !GROUPED) sum[j] = (uint32_t)((int)sum[j] + (int)o[j]);
          else sum[j] = __float_as_uint(__fadd_rn(__uint_as_float(sum[j]), __uint_as_float(o[j])));
        }
      }
      if constexpr (OUT_ACC) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (nb + j < p.N) p.acc_out[(int64_t)m * p.N + nb + j] = (int32_t)sum[j];
      } else {
        const float sxm = FUSED ? sx_s[m] : __ldg(p.sx + m);
        uint16_t o16[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float v;
          if constexpr (GROUPED) {
            v = __fmul_rn(__uint_as_float(sum[j]), sxm);
          } else {
            const float a = FMT == QL_INT8 ? __int2float_rn((int)sum[j]) : __uint_as_float(sum[j]);
            v = __fmul_rn(__fmul_rn(a, sxm), swv[j]);   // (acc * sa) * sb, reference :121
          }
          if (has_bias) v = __fadd_rn(v, bv[j]);
          o16[j] = f2bf(v);
        }
        uint16_t* dst = p.y + (int64_t)m * p.ldy + nb;
        if (vec_ok && nb + 8 <= p.N) {
          uint4 pk;
          pk.x = o16[0] | ((uint32_t)o16[1] << 16);
          pk.y = o16[2] | ((uint32_t)o16[3] << 16);
          pk.z = o16[4] | ((uint32_t)o16[5] << 16);
          pk.w = o16[6] | ((uint32_t)o16[7] << 16);
          *reinterpret_cast<uint4*>(dst) = pk;
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (nb + j < p.N) dst[j] = o16[j];
        }
      }
    }
  }

  if (threadIdx.x == 128) SK_TRACE(13);
  // warps 0-3 (and idle lanes) take part in the cluster barriers the epilogue warps use
  if (warp < 4 && S > 1) {
    cluster_wait();                    // #1
    if constexpr (Cfg::alias_recv) {
      cluster_arrive();                // #2
      cluster_wait();
    }
    cluster_arrive();                  // #3
    cluster_wait();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, Cfg::tmem_cols);
  if (threadIdx.x == 0) SK_TRACE(14);
}

// ------------------------------------------------------------------ host side

static long long* g_trace = nullptr;
void set_skinny_trace(void* buf) { g_trace = static_cast<long long*>(buf); }
long long* skinny_trace_ptr() { return g_trace; }

static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return (e != nullptr && e[0] != 0) ? atoi(e) : dflt;
}

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) v = env_int("QL_PDL", 1) != 0 ? 1 : 0;
  return v == 1;
}

struct SkPlan {
  int S, stages, ring_bytes, bres_bytes, smem_bytes;
};

// Split factor: double S while the grid still fits co-resident on the machine, every rank keeps
// at least one split unit and at least one token column.
template <bool GROUPED, int BT, bool FUSED>
static bool plan_skinny(const SkinnyParams& p, SkPlan& pl) {
  using Cfg = SkCfg<GROUPED, BT>;
  const int tiles_n = (p.N + kSkTileN - 1) / kSkTileN;
  const int resident = num_sms() * Cfg::min_blocks;
  int S = 1;
  const int forced = env_int("QL_SKINNY_S", 0);
  if (forced > 0) {
    S = forced;
    if (S > 8 || (S & (S - 1)) != 0 || S > p.units || BT % S != 0) return false;
  } else {
    while (S < 8 && S * 2 <= p.units && BT % (S * 2) == 0 && (int64_t)tiles_n * S * 2 <= resident) S *= 2;
  }
  const int stage = kSkStageA + (FUSED ? 0 : BT * kSkBlockK);
  const int nkb_max = ((p.units + S - 1) / S) * p.kb_per_g;
  const int bres = FUSED ? nkb_max * BT * kSkBlockK : 0;
  const int budget = (Cfg::two_per_sm ? 110 : 220) * 1024;
  const int fixed = 1024 + kSkSmall + bres + (Cfg::alias_recv ? 0 : Cfg::recv_bytes);
  int stages = (budget - fixed) / stage;
  if (stages > kSkMaxStages) stages = kSkMaxStages;
  if (stages > nkb_max) stages = nkb_max;
  if (stages < 1 || (stages < 2 && nkb_max > 1)) return false;
  int ring = stages * stage;
  if (Cfg::alias_recv && ring < Cfg::recv_bytes) ring = Cfg::recv_bytes;
  if (fixed + ring > 227 * 1024 - 1024) return false;
  pl.S = S; pl.stages = stages; pl.ring_bytes = ring; pl.bres_bytes = bres;
  pl.smem_bytes = fixed + ring;
  return true;
}

template <int FMT, bool GROUPED, int BT, bool FUSED, bool OUT_ACC>
static int launch_sk(const void* qx, const void* qw, SkinnyParams p, cudaStream_t st) {
  auto kern = skinny_kernel<FMT, GROUPED, BT, FUSED, OUT_ACC>;
  SkPlan pl;
  if (!plan_skinny<GROUPED, BT, FUSED>(p, pl)) return QL_ERR_BAD_SHAPE;
  static SmemAttrOnce smem_once;   // per kernel instantiation, per device ordinal
  if (int e = ensure_max_smem(reinterpret_cast<const void*>(kern), 226 * 1024, smem_once)) return e;
  CUtensorMap tm_w, tm_x;
  if (!make_tmap_2d(&tm_w, qw, CU_TENSOR_MAP_DATA_TYPE_UINT8, p.K, p.N, p.K, kSkBlockK, kSkTileN))
    return QL_ERR_CUDA;
  if constexpr (FUSED) {
    tm_x = tm_w;  // unused by the kernel
  } else {
    if (!make_tmap_2d(&tm_x, qx, CU_TENSOR_MAP_DATA_TYPE_UINT8, p.K, p.M, p.K, kSkBlockK, BT))
      return QL_ERR_CUDA;
  }
  p.trace = g_trace;
  p.stages = pl.stages;
  p.ring_bytes = pl.ring_bytes;
  p.bres_bytes = pl.bres_bytes;
  const int tiles_n = (p.N + kSkTileN - 1) / kSkTileN;

  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(tiles_n * pl.S));
  cfg.blockDim = dim3(kSkThreads);
  cfg.dynamicSmemBytes = (size_t)pl.smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (pl.S > 1) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = (unsigned)pl.S;
    attrs[na].val.clusterDim.y = 1;
    attrs[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_enabled()) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = (unsigned)na;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, kern, tm_w, tm_x, p);
  count_launch();
  if (err != cudaSuccess) {
    if (getenv("QL_DEBUG") != nullptr)
      fprintf(stderr, "[qlinear_b200] skinny launch failed: %s (S=%d smem=%d)\n",
              cudaGetErrorString(err), pl.S, pl.smem_bytes);
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  return QL_OK;
}

template <int FMT, bool GROUPED, bool OUT_ACC>
static int launch_sk_bt(const void* qx, const void* qw, const SkinnyParams& p, cudaStream_t st) {
  if (p.M <= 16) return launch_sk<FMT, GROUPED, 16, false, OUT_ACC>(qx, qw, p, st);
  if (p.M <= 32) return launch_sk<FMT, GROUPED, 32, false, OUT_ACC>(qx, qw, p, st);
  if (p.M <= 64) return launch_sk<FMT, GROUPED, 64, false, OUT_ACC>(qx, qw, p, st);
  if (p.M <= 128) return launch_sk<FMT, GROUPED, 128, false, OUT_ACC>(qx, qw, p, st);
  if constexpr (GROUPED) return QL_ERR_BAD_SHAPE;
  else return launch_sk<FMT, GROUPED, 256, false, OUT_ACC>(qx, qw, p, st);
}

int skinny_max_m(int64_t group) { return group > 0 ? 128 : 256; }

static SkinnyParams base_params(const float* sw, const void* bias, void* y, int64_t ldy, int64_t M,
                                int64_t N, int64_t K, int64_t group) {
  SkinnyParams p = {};
  p.sw = sw;
  p.bias = static_cast<const uint16_t*>(bias);
  p.y = static_cast<uint16_t*>(y);
  p.ldy = ldy;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.G = group > 0 ? (int)(K / group) : 1;
  p.kb_per_g = group > 0 ? (int)(group / kSkBlockK) : 1;
  p.units = (int)(K / kSkBlockK) / p.kb_per_g;
  return p;
}

// Pre-quantised activations (16 < M <= skinny_max_m, or the raw-accumulator test entry point).
int launch_skinny(const void* qx, const float* sx, const void* qw, const float* sw,
                  const void* bias, void* y, int64_t ldy, int32_t* out_acc, int64_t M, int64_t N,
                  int64_t K, int fmt, int64_t group, cudaStream_t st) {
  if (M == 0 || N == 0) return QL_OK;
  if (M > skinny_max_m(group)) return QL_ERR_BAD_SHAPE;
  SkinnyParams p = base_params(sw, bias, y, ldy, M, N, K, group);
  p.sx = sx;
  p.acc_out = out_acc;
  if (out_acc != nullptr) return launch_sk_bt<QL_INT8, false, true>(qx, qw, p, st);
  if (fmt == QL_INT8)
    return group > 0 ? launch_sk_bt<QL_INT8, true, false>(qx, qw, p, st)
                     : launch_sk_bt<QL_INT8, false, false>(qx, qw, p, st);
  return group > 0 ? launch_sk_bt<QL_FP8_E4M3, true, false>(qx, qw, p, st)
                   : launch_sk_bt<QL_FP8_E4M3, false, false>(qx, qw, p, st);
}

// Fused quantise + linear for M <= 16.  QL_ERR_BAD_SHAPE when the K slice of a rank does not
// fit the resident B operand (the caller falls back to gemv_kernel).
int launch_skinny_fused(const void* x, int64_t ldx, const void* qw, const float* sw,
                        const void* bias, void* y, int64_t ldy, int64_t M, int64_t N, int64_t K,
                        int fmt, int64_t group, cudaStream_t st) {
  if (M == 0 || N == 0) return QL_OK;
  if (M > 16) return QL_ERR_BAD_SHAPE;
  SkinnyParams p = base_params(sw, bias, y, ldy, M, N, K, group);
  p.x = static_cast<const uint16_t*>(x);
  p.ldx = ldx;
  if (fmt == QL_INT8)
    return group > 0 ? launch_sk<QL_INT8, true, 16, true, false>(nullptr, qw, p, st)
                     : launch_sk<QL_INT8, false, 16, true, false>(nullptr, qw, p, st);
  return group > 0 ? launch_sk<QL_FP8_E4M3, true, 16, true, false>(nullptr, qw, p, st)
                   : launch_sk<QL_FP8_E4M3, false, 16, true, false>(nullptr, qw, p, st);
}

static bool fused_plan(int64_t M, int64_t N, int64_t K, int64_t group, SkPlan& pl, int& nkb_max) {
  SkinnyParams p = base_params(nullptr, nullptr, nullptr, 0, M, N, K, group);
  const bool ok = group > 0 ? plan_skinny<true, 16, true>(p, pl) : plan_skinny<false, 16, true>(p, pl);
  if (ok) nkb_max = ((p.units + pl.S - 1) / pl.S) * p.kb_per_g;
  return ok;
}

bool skinny_fused_ok(int64_t M, int64_t N, int64_t K, int64_t group) {
  if (M > 16 || env_int("QL_NO_SKINNY", 0)) return false;
  SkPlan pl;
  int nkb_max = 0;
  return fused_plan(M, N, K, group, pl, nkb_max);
}

// Measured on B200 (profiles/r1j_fused_vs_two_launch.log): quantiser launch + TMA-fed decode
// kernel, chained by programmatic dependent launch, beats the fused kernel on every Qwen3 shape
// (0.6B qkv M=1: 4.9 vs 6.9 us; 8B gate_up M=16: 20.8 vs 44.9 us).  The fused kernel pays two
// dependent L2 round trips (~1.2 us each while the weight stream saturates the memory system)
// before its first MMA and its resident 16-row operand squeezes the weight ring.  It therefore
// only serves callers without a workspace (ql_gemv_w8a8); QL_FUSED=1 forces it for experiments.
bool skinny_fused_preferred(int64_t M, int64_t N, int64_t K, int64_t group) {
  if (env_int("QL_FUSED", 0) <= 0) return false;
  return skinny_fused_ok(M, N, K, group);
}

// Where the split-K decode kernel beats the tiled GEMMs, measured on B200 over the Qwen3 shapes
// (profiles/r1n_dispatch_crossover.log; whole ql_linear, graph replay): always up to 32 tokens;
// beyond that only while the K split buys parallelism the N x M tiling cannot -- long K relative
// to N -- because above 64 tokens a 256-wide CTA-pair tile already amortises the weight stream.
bool skinny_ok(int64_t M, int64_t N, int64_t K, int64_t group) {
  if (M > skinny_max_m(group) || env_int("QL_NO_SKINNY", 0)) return false;
  if (group > 0 || M <= 32 || env_int("QL_SKINNY_ALWAYS", 0)) return true;
  if (M <= 64) return K >= 2048;
  if (M <= 128) return K >= 2 * N;
  return K >= 2 * N && K >= 8192;
}

}  // namespace ql
