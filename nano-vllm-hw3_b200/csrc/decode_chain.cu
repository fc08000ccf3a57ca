// Decode-regime CHAIN of W8A8 linears (M <= 16 tokens) in ONE persistent launch.
//
// The reference runs every decode-step linear as per_row_quant + _per_row_post_process_mm
// (nanovllm/utils/quantization.py:310-346), i.e. two launches per linear and 4 linears per decoder
// layer (nanovllm/models/qwen3.py:80,90,116,118).  At M <= 16 a linear moves N*K weight bytes and
// almost nothing else: the regime is HBM-bound, and with one launch (or two) per linear the fixed
// launch / barrier-init / drain cost (~2.8 us per launch measured, profiles/r1n_bench_T16.json)
// is 5-10x the time the bytes need.  This kernel walks a whole chain of linears -- op i consumes
// the first K_i columns of op i-1's output, as the decoder stack's qkv -> o -> gate_up -> down
// order does in bench.py's LinearStack -- with
//
//   * one CTA per SM, resident for the whole chain; a producer warp streams the CTA's share of
//     EVERY op's weights through one shared-memory ring with 1-D bulk copies
//     (cp.async.bulk.shared.global, 8 weight rows x <= 2 KB of K per stage, ~200 KB in flight per
//     SM).  Weights depend on nothing, so the ring runs ahead across op boundaries: the weights
//     of op i+1 are already in shared memory while op i drains, and HBM never idles at a boundary;
//   * per op a 2-D split (row blocks x K slices of <= 2 KB) sized so that every SM gets one unit;
//     8 consumer warps split the unit's K slice 8 ways, quantise THEIR piece of the activation
//     row straight from L2 into mma fragments (no shared-memory staging, arithmetic of the Triton
//     quantiser :26-41 bit for bit) and run m16n8k32 warp MMAs (int8 -> int32 exact, e4m3 ->
//     fp32) -- the tensor-core flavour is irrelevant at 16 tokens, bytes in flight are what count;
//   * K-slice partials meet in L2: every slice's CTA stores its partial tile and bumps the row
//     block's counter; the LAST arriver adds the slices in slice order (int32: exact; fp32: fixed
//     order, deterministic), applies (acc * sx[m]) * sw[n] (+ bias[n]) -- reference product order
//     :111-121 -- stores bf16 and folds |y| into the NEXT op's per-token amax (atomicMax on the
//     float bit pattern), so the next op's quantiser needs no second pass over the row;
//   * ONE grid-wide barrier per op (red.release.gpu + ld.acquire.gpu spin), nothing else.
//
// Results are bit-identical to ql_linear op by op for INT8 per-row (int32 accumulation is exact in
// any order); FP8 per-row differs from the tcgen05 kernels only in fp32 summation order.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "kernels.h"

namespace ql {

constexpr int kDcConsumers = 256;              // 8 MMA warps
constexpr int kDcThreads = kDcConsumers + 32;  // + 1 producer warp
constexpr int kDcMaxSlice = 4096;              // bytes of K per unit: 32 pieces of 128 B, 4 per warp
constexpr int kDcPieces = kDcMaxSlice / 128 / 8;   // pieces per warp
constexpr int kDcSlotK = 1024;                 // bytes of K per ring slot (8 weight rows x <= 1 KB)
constexpr int kDcSlotPitch = kDcSlotK + 64;    // rows 64 B apart modulo 128: conflict-free LDS.128
constexpr int kDcSlotBytes = 8 * kDcSlotPitch;
constexpr int kDcSlots = 22;                   // ~187 KB of weights in flight per SM
constexpr int kDcR = 4;                        // row groups accumulated between cross-warp reductions
constexpr int kDcRedBytes = 2 * 8 * kDcR * 128 * 4;   // double-buffered [8 warps][R][16 x 8] partial tiles
constexpr int kDcSmem = kDcSlots * kDcSlotBytes + kDcRedBytes + 1024 + 128;
constexpr uint32_t kDcMagic = 0x51444332u;     // "QDC2"
static_assert(kDcSmem <= 227 * 1024, "decode chain shared memory");

struct ChainOp {            // device-side descriptor, 128 bytes
  const uint8_t* qw;        // [N, K] codes
  const float* sw;          // [N]
  const uint16_t* bias;     // bf16 [N] or nullptr
  const uint16_t* x;        // bf16 [M, ldx]
  uint16_t* y;              // bf16 [M, ldy]
  int64_t ldx, ldy;
  int N, K;
  int KS, RB;               // K slices x row blocks = units (<= grid)
  int ksz, rot;             // bytes of K per slice; unit -> CTA rotation
  int RG;                   // groups of 8 weight rows
  int amax_cols;            // leading columns of y folded into the next op's amax (0: none)
  int x_amax;               // 1: amax[op] was produced by the previous op's epilogue
  uint32_t ctr_off;         // first row-block counter of this op
  uint32_t pad[8];
};
static_assert(sizeof(ChainOp) == 128, "ChainOp layout");

struct ChainParams {
  const ChainOp* ops;
  int n_ops, M;
  uint32_t* sync;           // [0] grid barrier counter, [1] exit counter
  uint32_t* amax;           // [n_ops][16] float bit patterns
  uint32_t* blk_ctr;
  uint32_t* part;           // K-slice partials [KS][16][N] (int32 / fp32 bits)
  long long* trace;         // debug (ql_debug_set_trace): clock64 stamps [cta][16 ops][16], else nullptr
  int flags;                // debug (QL_CHAIN_DEBUG): skip parts of the kernel for attribution -- WRONG RESULTS
};
#define DC_TRACE(i)                                                                          \
  do {                                                                                       \
    if (p.trace != nullptr && tid == 0 && l < 16)                                            \
      p.trace[((size_t)cta * 16 + l) * 16 + (i)] = clock64();                                \
  } while (0)

// mbarrier wait without the globaltimer watchdog of mbar_wait (flag 64: attribution experiment)
__device__ __forceinline__ void dc_wait(uint64_t* bar, uint32_t parity, int flags) {
  if (flags & 64) {
    uint32_t n = 0;
    while (!mbar_try_wait(bar, parity))
      if (++n > (1u << 28)) __trap();
  } else {
    mbar_wait(bar, parity);
  }
}

__device__ __forceinline__ void mma_i8_sync(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                            uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_f8_sync(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                            uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* src, uint32_t bytes, uint64_t* bar,
                                          uint64_t hint) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(smem_dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(hint)
      : "memory");
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// 16 bf16 (two 16-byte vectors) -> 16 codes (4 words)
template <int FMT>
__device__ __forceinline__ void quant16(const uint4& v0, const uint4& v1, float inv, uint32_t (&w)[4]) {
  const uint2 q0 = quant8_mul<FMT>(v0, inv), q1 = quant8_mul<FMT>(v1, inv);
  w[0] = q0.x; w[1] = q0.y; w[2] = q1.x; w[3] = q1.y;
}

// unit of CTA `cta` in an op: row block rb, K slice ks (false: the CTA idles through this op)
__device__ __forceinline__ bool dc_unit(int cta, int G, int rot, int KS, int RB, int& rb, int& ks) {
  int v = cta + rot;
  if (v >= G) v -= G;
  rb = ks = 0;
  if (v >= KS * RB) return false;
  rb = v / KS;
  ks = v - rb * KS;
  return true;
}

template <int FMT>
__global__ void __launch_bounds__(kDcThreads, 1) decode_chain_kernel(const ChainParams p) {
  using acc_t = typename std::conditional<FMT == QL_INT8, int, float>::type;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  uint8_t* ring = smem;
  uint32_t* red = reinterpret_cast<uint32_t*>(ring + kDcSlots * kDcSlotBytes);   // [2][8][R][128]
  uint8_t* small = reinterpret_cast<uint8_t*>(red) + kDcRedBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(small);                       // [kDcSlots]
  uint64_t* empty_bar = full_bar + kDcSlots;                                     // [kDcSlots] (352 B so far)
  float* sx_s = reinterpret_cast<float*>(small + 384);     // [16]
  int* amax_s = reinterpret_cast<int*>(small + 448);       // [16]
  uint32_t* flag_s = reinterpret_cast<uint32_t*>(small + 512);
  ChainOp* desc_s = reinterpret_cast<ChainOp*>(small + 640);   // [2]: this op's / the next op's descriptor

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = gridDim.x, cta = blockIdx.x;
  constexpr float qmax = Fmt<FMT>::qmax;

  if (tid == 0) {
    for (int i = 0; i < kDcSlots; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 8);
    }
    mbar_fence_init();
  }
  if (tid < 8) reinterpret_cast<uint4*>(desc_s)[tid] = __ldg(reinterpret_cast<const uint4*>(p.ops) + tid);
  __syncthreads();

  if (warp == 8) {
    // ===================== producer: every op's weights, as far ahead as the ring allows ==========
    int slot = 0;
    uint32_t phase = 0;
    for (int l = 0; l < p.n_ops; ++l) {
      const ChainOp& op = p.ops[l];
      const int KS = __ldg(&op.KS), RB = __ldg(&op.RB);
      int rb, ks;
      if (!dc_unit(cta, G, __ldg(&op.rot), KS, RB, rb, ks)) continue;
      const int RG = __ldg(&op.RG), ksz = __ldg(&op.ksz), K = __ldg(&op.K);
      const int rg0 = (int)((int64_t)rb * RG / RB), rg1 = (int)((int64_t)(rb + 1) * RG / RB);
      const uint8_t* src0 = op.qw + (size_t)ks * ksz;
      for (int rg = rg0; rg < rg1; ++rg) {
        const uint8_t* srow = src0 + (size_t)(rg * 8 + (lane & 7)) * K;
        for (int off = 0; off < ksz; off += kDcSlotK) {
          const uint32_t bytes = (uint32_t)(ksz - off < kDcSlotK ? ksz - off : kDcSlotK);
          dc_wait(&empty_bar[slot], phase ^ 1, p.flags);
          if (p.flags & 8) {
            if (lane == 0) mbar_arrive(&full_bar[slot]);
            if (++slot == kDcSlots) { slot = 0; phase ^= 1; }
            continue;
          }
          if (lane == 0) mbar_expect_tx(&full_bar[slot], 8u * bytes);
          __syncwarp();
          if (lane < 8)
            bulk_load(ring + slot * kDcSlotBytes + lane * kDcSlotPitch, srow + off, bytes, &full_bar[slot],
                      kEvictFirst);
          if (++slot == kDcSlots) { slot = 0; phase ^= 1; }
        }
      }
    }
    return;
  }

  // ===================== consumers =====================
  const int g = lane >> 2, t = lane & 3;     // mma fragment coordinates
  const int e_m = (tid & 127) >> 3, e_j = tid & 7, e_r = tid >> 7;   // finalizer: token, row in group, group parity
  int slot = 0;
  uint32_t phase = 0;
  uint32_t red_buf = 0;

  for (int l = 0; l < p.n_ops; ++l) {
    const ChainOp& op = desc_s[l & 1];
    const int KS = op.KS, RB = op.RB;
    int rb, ks;
    const bool active = dc_unit(cta, G, op.rot, KS, RB, rb, ks);
    DC_TRACE(0);
    if (active) {
      const int RG = op.RG, ksz = op.ksz, N = op.N;
      const int rg0 = (int)((int64_t)rb * RG / RB), rg1 = (int)((int64_t)(rb + 1) * RG / RB);
      const int nsc = ksz >> 7;                          // 128-byte pieces of the K slice
      const int nslot = (ksz + kDcSlotK - 1) / kDcSlotK;
      const int64_t ldx = op.ldx, ldy = op.ldy;
      const uint16_t* xk = op.x + (size_t)ks * ksz;      // this slice's first column
      const float* sw = op.sw;
      const uint16_t* bias = op.bias;
      uint16_t* y = op.y;
      const int amax_cols = op.amax_cols;

      // ---- this warp's raw activation pieces (L2): in flight together with the amax read --------
      uint4 xv[kDcPieces][2][2][2];                      // [piece][half][token g / g+8][vector]
#pragma unroll
      for (int i = 0; i < kDcPieces; ++i) {
        const int sc = warp + 8 * i;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int kb = sc * 128 + h * 64 + t * 16;
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int m = g + 8 * u;
            xv[i][h][u][0] = xv[i][h][u][1] = make_uint4(0, 0, 0, 0);
            if (sc < nsc && m < p.M && !(p.flags & 16)) {
              const uint4* s = reinterpret_cast<const uint4*>(xk + (size_t)m * ldx + kb);
              xv[i][h][u][0] = __ldcg(s);
              xv[i][h][u][1] = __ldcg(s + 1);
            }
          }
        }
      }

      // ---- per-token scale of the input rows ------------------------------------------------
      if (op.x_amax) {
        if (tid < 16) amax_s[tid] = (int)__ldcg(p.amax + l * 16 + tid);
      } else {
        // no producer in the chain: every CTA scans the (small) activation once
        if (tid < 16) amax_s[tid] = 0;
        consumer_sync();
        const int nvec = op.K >> 3;
        for (int m = warp; m < p.M; m += 8) {
          float a = 0.0f;
          const uint4* row = reinterpret_cast<const uint4*>(op.x + (size_t)m * ldx);
          for (int i = lane; i < nvec; i += 32) a = fmaxf(a, amax8(__ldcg(row + i)));
          a = warp_max(a);
          if (lane == 0) amax_s[m] = __float_as_int(a);
        }
      }
      consumer_sync();
      DC_TRACE(1);
      const float am_lo = fmaxf(__int_as_float(amax_s[g]), 1e-8f);
      const float am_hi = fmaxf(__int_as_float(amax_s[g + 8]), 1e-8f);
      const float inv_lo = __fdiv_rn(qmax, am_lo), inv_hi = __fdiv_rn(qmax, am_hi);
      if (tid < 16) sx_s[tid] = __fdiv_rn(fmaxf(__int_as_float(amax_s[tid]), 1e-8f), qmax);

      // ---- quantise into mma A fragments (Triton quantiser arithmetic :26-41) ------------------
      uint32_t a_lo[kDcPieces][2][4], a_hi[kDcPieces][2][4];
#pragma unroll
      for (int i = 0; i < kDcPieces; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          quant16<FMT>(xv[i][h][0][0], xv[i][h][0][1], inv_lo, a_lo[i][h]);
          quant16<FMT>(xv[i][h][1][0], xv[i][h][1][1], inv_hi, a_hi[i][h]);
        }

      // ---- stream the unit's weight row groups, kDcR at a time ----------------------------------
      float run_max = 0.0f;
      if (p.trace != nullptr && tid == 0 && l < 16 && (a_lo[0][0][0] ^ a_hi[0][0][0]) != 0x12345678u) DC_TRACE(2);
      for (int rgb = rg0; rgb < rg1; rgb += kDcR) {
        const int nr = rg1 - rgb < kDcR ? rg1 - rgb : kDcR;
        // scales / bias of the rows this thread finalises (static data: the loads overlap the MMAs)
        float swv[kDcR / 2], bv[kDcR / 2];
#pragma unroll
        for (int j = 0; j < kDcR / 2; ++j) {
          const int r = e_r + 2 * j;
          const int n = (rgb + r) * 8 + e_j;
          const bool ok = r < nr && KS == 1;
          swv[j] = ok ? __ldg(sw + n) : 0.0f;
          bv[j] = (ok && bias != nullptr) ? bf16_lo(__ldg(bias + n)) : 0.0f;
        }
        acc_t c[kDcR][4];
#pragma unroll
        for (int r = 0; r < kDcR; ++r) c[r][0] = c[r][1] = c[r][2] = c[r][3] = 0;
#pragma unroll
        for (int r = 0; r < kDcR; ++r) {
          if (r < nr) {
#pragma unroll
            for (int i = 0; i < kDcPieces; ++i) {
              if (i < nslot) {
                dc_wait(&full_bar[slot], phase, p.flags);
                if (r == 0 && i == 0 && rgb == rg0) DC_TRACE(3);
                if (warp + 8 * i < nsc && !(p.flags & 1)) {
                  const uint8_t* wrow = ring + slot * kDcSlotBytes + g * kDcSlotPitch + warp * 128 + t * 16;
#pragma unroll
                  for (int h = 0; h < 2; ++h) {
                    const uint4 w = *reinterpret_cast<const uint4*>(wrow + h * 64);
                    if constexpr (FMT == QL_INT8) {
                      mma_i8_sync(c[r], a_lo[i][h][0], a_hi[i][h][0], a_lo[i][h][1], a_hi[i][h][1], w.x, w.y);
                      mma_i8_sync(c[r], a_lo[i][h][2], a_hi[i][h][2], a_lo[i][h][3], a_hi[i][h][3], w.z, w.w);
                    } else {
                      mma_f8_sync(c[r], a_lo[i][h][0], a_hi[i][h][0], a_lo[i][h][1], a_hi[i][h][1], w.x, w.y);
                      mma_f8_sync(c[r], a_lo[i][h][2], a_hi[i][h][2], a_lo[i][h][3], a_hi[i][h][3], w.z, w.w);
                    }
                  }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty_bar[slot]);
                if (r == 0 && i == 0 && rgb == rg0) DC_TRACE(4);
                if (++slot == kDcSlots) { slot = 0; phase ^= 1; }
              }
            }
            if (r == 0 && rgb == rg0) DC_TRACE(5);
          }
        }

        if (rgb == rg0) DC_TRACE(6);
        // cross-warp reduction of the nr 16 x 8 tiles (every warp holds a K piece of each)
        uint32_t* rt = red + red_buf * (8 * kDcR * 128) + warp * (kDcR * 128);
#pragma unroll
        for (int r = 0; r < kDcR; ++r) {
          if (r < nr) {
            uint32_t cu[4];
            if constexpr (FMT == QL_INT8) { cu[0] = c[r][0]; cu[1] = c[r][1]; cu[2] = c[r][2]; cu[3] = c[r][3]; }
            else {
              cu[0] = __float_as_uint(c[r][0]); cu[1] = __float_as_uint(c[r][1]);
              cu[2] = __float_as_uint(c[r][2]); cu[3] = __float_as_uint(c[r][3]);
            }
            *reinterpret_cast<uint2*>(rt + r * 128 + g * 8 + 2 * t) = make_uint2(cu[0], cu[1]);
            *reinterpret_cast<uint2*>(rt + r * 128 + (g + 8) * 8 + 2 * t) = make_uint2(cu[2], cu[3]);
          }
        }
        consumer_sync();
        if (rgb == rg0) DC_TRACE(7);
#pragma unroll
        for (int j = 0; j < kDcR / 2; ++j) {
          const int r = e_r + 2 * j;
          if (r < nr) {
            const uint32_t* rs = red + red_buf * (8 * kDcR * 128) + r * 128 + (tid & 127);
            uint32_t sum = rs[0];
#pragma unroll
            for (int w = 1; w < 8; ++w) {
              if constexpr (FMT == QL_INT8) sum = (uint32_t)((int)sum + (int)rs[w * (kDcR * 128)]);
              else sum = __float_as_uint(__fadd_rn(__uint_as_float(sum), __uint_as_float(rs[w * (kDcR * 128)])));
            }
            const int n = (rgb + r) * 8 + e_j;
            if (e_m < p.M) {
              if (KS == 1) {
                const float a = FMT == QL_INT8 ? __int2float_rn((int)sum) : __uint_as_float(sum);
                float v = __fmul_rn(__fmul_rn(a, sx_s[e_m]), swv[j]);     // (acc * sa) * sb, reference :121
                if (bias != nullptr) v = __fadd_rn(v, bv[j]);
                const uint16_t o = f2bf(v);
                y[(size_t)e_m * ldy + n] = o;
                if (n < amax_cols) run_max = fmaxf(run_max, fabsf(bf16_lo(o)));
              } else {
                p.part[((size_t)ks * 16 + e_m) * N + n] = sum;
              }
            }
          }
        }
        if (rgb == rg0) DC_TRACE(8);
        red_buf ^= 1;
      }

      DC_TRACE(9);
      // ---- K-slice partials: the last slice to arrive finishes the row block --------------------
      if (KS > 1) {
        __threadfence();
        consumer_sync();
        if (tid == 0) {
          uint32_t* ctr = p.blk_ctr + op.ctr_off + rb;
          const uint32_t old = atomicAdd(ctr, 1u);
          const uint32_t last = old == (uint32_t)(KS - 1);
          if (last) *ctr = 0;                              // ready for the next launch
          *flag_s = last;
        }
        consumer_sync();
        if (*flag_s) {
          __threadfence();
          if (e_m < p.M) {
            const float sxm = sx_s[e_m];
            for (int rg = rg0 + e_r; rg < rg1; rg += 2) {
              const int n = rg * 8 + e_j;
              const uint32_t* src = p.part + (size_t)e_m * N + n;
              uint32_t sum = __ldcg(src);
              for (int s = 1; s < KS; ++s) {
                const uint32_t o = __ldcg(src + (size_t)s * 16 * N);
                if constexpr (FMT == QL_INT8) sum = (uint32_t)((int)sum + (int)o);
                else sum = __float_as_uint(__fadd_rn(__uint_as_float(sum), __uint_as_float(o)));
              }
              const float a = FMT == QL_INT8 ? __int2float_rn((int)sum) : __uint_as_float(sum);
              float v = __fmul_rn(__fmul_rn(a, sxm), __ldg(sw + n));
              if (bias != nullptr) v = __fadd_rn(v, bf16_lo(__ldg(bias + n)));
              const uint16_t o = f2bf(v);
              y[(size_t)e_m * ldy + n] = o;
              if (n < amax_cols) run_max = fmaxf(run_max, fabsf(bf16_lo(o)));
            }
          }
        }
      }

      // ---- fold |y| into the next op's per-token amax --------------------------------------------
      if (amax_cols > 0) {
        run_max = fmaxf(run_max, __shfl_xor_sync(0xffffffffu, run_max, 1));
        run_max = fmaxf(run_max, __shfl_xor_sync(0xffffffffu, run_max, 2));
        run_max = fmaxf(run_max, __shfl_xor_sync(0xffffffffu, run_max, 4));
        if (e_j == 0 && e_m < p.M && run_max > 0.0f && !(p.flags & 4))
          atomicMax(reinterpret_cast<int*>(p.amax) + (l + 1) * 16 + e_m, __float_as_int(run_max));
      }
      if (!(p.flags & 32)) __threadfence();
      DC_TRACE(10);
    }

    // ---- grid barrier: op l is complete everywhere (its output and the next amax are visible) ----
    if (l + 1 < p.n_ops) {
      // the next op's descriptor is static: fetch it while the barrier is in flight
      if (tid >= 32 && tid < 40)
        reinterpret_cast<uint4*>(desc_s + ((l + 1) & 1))[tid - 32] =
            __ldg(reinterpret_cast<const uint4*>(p.ops + l + 1) + (tid - 32));
      consumer_sync();
      DC_TRACE(11);
      if (tid == 0) {
        red_release_gpu_add(p.sync, 1u);
        DC_TRACE(12);
        const uint32_t target = (p.flags & 2) ? 0u : (uint32_t)(l + 1) * (uint32_t)G;
        if (ld_acquire_gpu(p.sync) < target) {
          if (p.flags & 64) {
            uint32_t n = 0;
            while (ld_acquire_gpu(p.sync) < target)
              if (++n > (1u << 26)) __trap();
          } else {
            const uint64_t t0 = globaltimer_ns();
            while (ld_acquire_gpu(p.sync) < target)
              if (globaltimer_ns() - t0 > 4000000000ull) __trap();   // 4 s: never hang the GPU
          }
        }
        DC_TRACE(13);
      }
      consumer_sync();
    }
  }

  // ---- the last CTA out re-arms the counters for the next launch --------------------------------
  consumer_sync();
  if (tid == 0) {
    __threadfence();
    *flag_s = atomicAdd(p.sync + 1, 1u) == (uint32_t)(G - 1);
  }
  consumer_sync();
  if (*flag_s) {
    for (int i = tid; i < p.n_ops * 16; i += kDcConsumers) p.amax[i] = 0u;
    if (tid == 0) {
      p.sync[0] = 0u;
      p.sync[1] = 0u;
    }
  }
}

// ------------------------------------------------------------------ host side

struct ChainPlanHost {      // what ql_chain_run needs; lives in the caller's ql_chain_plan blob
  uint32_t magic;
  int n_ops, M, fmt, grid, smem;
  size_t off_amax, off_ctr, off_ops, off_part, total;
};
static_assert(sizeof(ChainPlanHost) <= sizeof(ql_chain_plan), "ql_chain_plan too small");

static size_t rup(size_t v, size_t a) { return (v + a - 1) / a * a; }

// K slices: the fewest that keep a slice <= 2 KB with K % (KS * 128) == 0
static int pick_ks(int64_t K) {
  for (int ks = (int)((K + kDcMaxSlice - 1) / kDcMaxSlice); ks <= 64; ++ks)
    if (K % ((int64_t)ks * 128) == 0) return ks;
  return 0;
}

static int plan_chain(const ql_chain_op* ops, int n_ops, int64_t M, int grid, ChainPlanHost* pl, ChainOp* dev_ops) {
  if (ops == nullptr || n_ops <= 0 || n_ops > 4096 || M <= 0 || M > 16) return QL_ERR_BAD_SHAPE;
  size_t n_ctr = 0, part = 0;
  for (int l = 0; l < n_ops; ++l) {
    const ql_chain_op& o = ops[l];
    if (o.N <= 0 || o.K <= 0 || o.K % 128 != 0 || o.N % 8 != 0 || o.ldy < o.N) return QL_ERR_BAD_SHAPE;
    if (!o.qw || !o.sw || !o.y) return QL_ERR_NULL;
    if (l == 0 && o.x == nullptr) return QL_ERR_NULL;
    const bool chained = o.x == nullptr;
    const void* x = chained ? ops[l - 1].y : o.x;
    const int64_t ldx = chained ? ops[l - 1].ldy : o.ldx;
    if (chained && ops[l - 1].N < o.K) return QL_ERR_BAD_SHAPE;
    if (ldx < o.K) return QL_ERR_BAD_SHAPE;
    if ((reinterpret_cast<uintptr_t>(x) & 15) || ldx % 8 != 0 || (reinterpret_cast<uintptr_t>(o.qw) & 15))
      return QL_ERR_MISALIGNED;
    const int KS = pick_ks(o.K);
    if (KS == 0 || KS > grid) return QL_ERR_BAD_SHAPE;
    const int RG = (int)(o.N / 8);
    int RB = grid / KS;
    if (RB > RG) RB = RG;
    if (dev_ops != nullptr) {
      ChainOp& d = dev_ops[l];
      memset(&d, 0, sizeof(d));
      d.qw = static_cast<const uint8_t*>(o.qw);
      d.sw = o.sw;
      d.bias = static_cast<const uint16_t*>(o.bias);
      d.x = static_cast<const uint16_t*>(x);
      d.y = static_cast<uint16_t*>(o.y);
      d.ldx = ldx;
      d.ldy = o.ldy;
      d.N = (int)o.N;
      d.K = (int)o.K;
      d.KS = KS;
      d.RB = RB;
      d.ksz = (int)(o.K / KS);
      d.rot = (int)(((int64_t)l * 41) % grid);   // spread the idle / one-group-longer units over the SMs
      d.RG = RG;
      d.x_amax = chained ? 1 : 0;
      d.amax_cols = (l + 1 < n_ops && ops[l + 1].x == nullptr) ? (int)ops[l + 1].K : 0;
      d.ctr_off = (uint32_t)n_ctr;
    }
    if (KS > 1) {
      n_ctr += (size_t)RB;
      const size_t need = (size_t)KS * 16 * (size_t)o.N * 4;
      if (need > part) part = need;
    }
  }
  pl->magic = kDcMagic;
  pl->n_ops = n_ops;
  pl->M = (int)M;
  pl->grid = grid;
  pl->smem = kDcSmem;
  pl->off_amax = 256;
  pl->off_ctr = pl->off_amax + rup((size_t)n_ops * 64, 256);
  pl->off_ops = pl->off_ctr + rup((n_ctr + 1) * 4, 256);
  pl->off_part = pl->off_ops + rup((size_t)n_ops * sizeof(ChainOp), 256);
  pl->total = pl->off_part + rup(part, 256);
  return QL_OK;
}

}  // namespace ql

using namespace ql;

extern "C" {

size_t ql_chain_workspace_bytes(const ql_chain_op* ops, int n_ops, int64_t M) {
  ChainPlanHost pl;
  if (plan_chain(ops, n_ops, M, num_sms(), &pl, nullptr) != QL_OK) return 0;
  return pl.total;
}

int ql_chain_prepare(const ql_chain_op* ops, int n_ops, int64_t M, int fmt, int64_t group, void* workspace,
                     size_t workspace_bytes, ql_chain_plan* plan, ql_stream_t stream) {
  if (fmt != QL_INT8 && fmt != QL_FP8_E4M3) return QL_ERR_BAD_FORMAT;
  if (group != 0) return QL_ERR_BAD_SHAPE;          // per-row scales only (per-group: ql_linear)
  if (plan == nullptr || workspace == nullptr) return QL_ERR_NULL;
  if (reinterpret_cast<uintptr_t>(workspace) & 255) return QL_ERR_MISALIGNED;
  if (int e = ql_device_check()) return e;
  ChainPlanHost pl;
  ChainOp* host_ops = static_cast<ChainOp*>(malloc(sizeof(ChainOp) * (size_t)(n_ops > 0 ? n_ops : 1)));
  if (host_ops == nullptr) return QL_ERR_CUDA;
  int e = plan_chain(ops, n_ops, M, num_sms(), &pl, host_ops);
  if (e == QL_OK && workspace_bytes < pl.total) e = QL_ERR_WORKSPACE;
  if (e == QL_OK) {
    pl.fmt = fmt;
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    cudaStream_t st = (cudaStream_t)stream;
    // counters, amax arrays and row-block counters start at zero; the kernel re-arms them itself
    if (cudaMemsetAsync(ws, 0, pl.off_ops, st) != cudaSuccess ||
        cudaMemcpyAsync(ws + pl.off_ops, host_ops, sizeof(ChainOp) * (size_t)n_ops, cudaMemcpyHostToDevice, st) !=
            cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) {
      (void)cudaGetLastError();
      e = QL_ERR_CUDA;
    }
  }
  free(host_ops);
  if (e != QL_OK) return e;
  memset(plan, 0, sizeof(*plan));
  memcpy(plan, &pl, sizeof(pl));
  return QL_OK;
}

int ql_chain_run(const ql_chain_plan* plan, void* workspace, ql_stream_t stream) {
  if (plan == nullptr || workspace == nullptr) return QL_ERR_NULL;
  ChainPlanHost pl;
  memcpy(&pl, plan, sizeof(pl));
  if (pl.magic != kDcMagic) return QL_ERR_BAD_SHAPE;
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  ChainParams p;
  p.ops = reinterpret_cast<const ChainOp*>(ws + pl.off_ops);
  p.n_ops = pl.n_ops;
  p.M = pl.M;
  p.sync = reinterpret_cast<uint32_t*>(ws);
  p.amax = reinterpret_cast<uint32_t*>(ws + pl.off_amax);
  p.blk_ctr = reinterpret_cast<uint32_t*>(ws + pl.off_ctr);
  p.part = reinterpret_cast<uint32_t*>(ws + pl.off_part);
  p.trace = skinny_trace_ptr();
  {
    const char* e = getenv("QL_CHAIN_DEBUG");
    p.flags = (e != nullptr && e[0] != 0) ? atoi(e) : 0;
  }
  const void* kern = pl.fmt == QL_INT8 ? reinterpret_cast<const void*>(decode_chain_kernel<QL_INT8>)
                                       : reinterpret_cast<const void*>(decode_chain_kernel<QL_FP8_E4M3>);
  static SmemAttrOnce once_i8, once_f8;
  if (int e = ensure_max_smem(kern, pl.smem, pl.fmt == QL_INT8 ? once_i8 : once_f8)) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)pl.grid);
  cfg.blockDim = dim3(kDcThreads);
  cfg.dynamicSmemBytes = (size_t)pl.smem;
  cfg.stream = (cudaStream_t)stream;
  // every CTA spins on the grid barrier: the grid must be co-resident (one CTA per SM by shared memory)
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  void* args[] = {&p};
  const cudaError_t err = cudaLaunchKernelExC(&cfg, kern, args);
  count_launch();
  if (err != cudaSuccess) {
    if (getenv("QL_DEBUG") != nullptr)
      fprintf(stderr, "[qlinear_b200] decode chain launch failed: %s\n", cudaGetErrorString(err));
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  return QL_OK;
}

}  // extern "C"
