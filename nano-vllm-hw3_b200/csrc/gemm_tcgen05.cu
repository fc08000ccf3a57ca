// W8A8 GEMM on the 5th-generation tensor cores (tcgen05) with the scale/bias epilogue fused.
// Replaces the reference's Triton kernel _per_row_post_process_mm
// (nanovllm/utils/quantization.py:54-144) and, for group > 0, the fp32 dequant + SGEMM of
// per_group_matmul (:265-307).
//
// Orientation ("weight-stationary", swap-AB): the UMMA A operand is a 128-row tile of the
// WEIGHT matrix qw[N,K]; the B operand is a BT-token tile of the quantised activations
// qx[M,K] (BT = 32..256 = UMMA N).  The accumulator D[128 x BT] lives in TMEM with the weight
// row n on the TMEM lane and the token m on the column, so in the epilogue each thread owns one
// output feature n: its weight scale(s) sw[n] / sw[n,j] and bias[n] are per-thread scalars and
// the bf16 stores of a warp cover 32 consecutive n of one token (64 contiguous bytes).
//
// Pipeline per CTA (persistent over tiles, 256 threads):
//   warp 0 (1 lane)  TMA producer: cp.async.bulk.tensor 128-byte-swizzled tiles of qw and qx
//                    into a STAGES-deep shared-memory ring, completion on `full` mbarriers
//   warp 1 (1 lane)  MMA issuer: 4 x tcgen05.mma (K = 32 bytes each) per 128-byte K block,
//                    tcgen05.commit -> `empty` (smem slot free) and `tfull` (accumulator ready)
//   warp 2           TMEM allocator (512 columns = 2..4 accumulator buffers)
//   warps 4-7        epilogue: tcgen05.ld 32 lanes x 32 columns -> registers -> scales -> bf16
// Per-row mode: one accumulator per tile (all of K), double-buffered so the epilogue of tile i
// overlaps the MMAs of tile i+1.  Per-group mode: one accumulator per 128-wide K block; the
// epilogue warps promote each block with its weight scale into fp32 registers
// (t = fmaf(float(P_j), sw[n,j], t)) while the tensor core fills the next buffer.
//
// INT8 accumulators are exact int32 (kind::i8, SASS UTCIMMA); FP8 accumulates in fp32
// (kind::f8f6f4, SASS UTCQMMA).  Epilogue order is the reference's: (float(acc) * sx[m]) * sw[n].
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace ql {

struct GemmParams {
  const float* sx;       // [M]
  const float* sw;       // [N] or [N, G]
  const uint16_t* bias;  // bf16 [N] or nullptr
  uint16_t* y;           // bf16 [M, ldy]
  int64_t ldy;
  int32_t* acc_out;      // int32 [M, N] (OUT_ACC only)
  int M, N, K;
  int G;                 // scale columns (per-group) or 1
  int kb_per_g;          // 128-byte K blocks per scale group
  int tiles_n, tiles_m;
  int group_m;           // token tiles per rasterisation super-group (common.cuh tile_coords)
};

constexpr int kGemmThreads = 256;
constexpr int kTileN = 128;   // weight rows per tile (UMMA M)
constexpr int kBlockK = 128;  // bytes of K per pipeline stage (one 128B swizzle atom)
constexpr int kSmemBudget = 227 * 1024;
// epilogue staging per epilogue warp: a [32 tokens][32 features] bf16 tile (64-byte rows) used to
// transpose the accumulator (feature on the TMEM lane) into token-major 16-byte global stores,
// plus the 32 per-token scales of the chunk
constexpr int kEpiWarpSmem = 32 * 32 * 2 + 32 * 4;
constexpr int kEpiSmem = 8 * kEpiWarpSmem;   // up to 8 epilogue warps (per-group mode)
// Per-group mode runs 8 epilogue warps (two per TMEM lane quarter, each promoting half of the token
// columns): the per-K-block promotion, not the tensor core, bounds that mode.
template <bool GROUPED, int BT> struct EpiCfg {
  static constexpr int halves = (GROUPED && BT >= 64) ? 2 : 1;
  static constexpr int threads = GROUPED ? 384 : 256;
};

template <int BT> struct GemmCfg {
  static constexpr int stage_a = kTileN * kBlockK;
  static constexpr int stage_b = BT * kBlockK;
  static constexpr int stage = stage_a + stage_b;
  static constexpr int stages_raw = (kSmemBudget - 2048 - kEpiSmem) / stage;
  static constexpr int stages = stages_raw > 8 ? 8 : stages_raw;
  static constexpr int nacc_raw = 512 / BT;
  static constexpr int nacc = nacc_raw > 4 ? 4 : nacc_raw;
  static constexpr int tmem_cols_raw = nacc * BT;
  static constexpr int tmem_cols = tmem_cols_raw <= 32 ? 32 : tmem_cols_raw <= 64 ? 64
                                   : tmem_cols_raw <= 128 ? 128 : tmem_cols_raw <= 256 ? 256 : 512;
  static constexpr int smem_bytes = stages * stage + 1024 /*align slack*/ + 256 /*barriers*/ + kEpiSmem;
};

template <int FMT, bool GROUPED, int BT, bool OUT_ACC>
__global__ void __launch_bounds__(EpiCfg<GROUPED, BT>::threads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
            const GemmParams p) {
  using Cfg = GemmCfg<BT>;
  constexpr int STAGES = Cfg::stages;
  constexpr int NACC = Cfg::nacc;
  static_assert(!(GROUPED && BT > 128), "per-group epilogue keeps BT fp32 sums in registers");
  static_assert(!(OUT_ACC && (GROUPED || FMT != QL_INT8)), "raw accumulators: int8 per-row only");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::stage);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + NACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + NACC);
  uint8_t* epi_smem = smem + STAGES * Cfg::stage + 256;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_w);
    tma_prefetch_desc(&tm_x);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < NACC; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], 4 * EpiCfg<GROUPED, BT>::halves);  // one arrive per active epilogue warp
    }
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
  pdl_launch_dependents();

  const int num_kb = p.K / kBlockK;
  const int total_tiles = p.tiles_n * p.tiles_m;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      // programmatic dependent launch: weight tiles of the first ring fill are requested before
      // griddepcontrol.wait; the activation loads (written by the quantiser) come after it
      int pre = 0;
      {
        int stage = 0;
        for (int tile = blockIdx.x; tile < total_tiles && pre < STAGES; tile += gridDim.x) {
          int tm, tn;
          tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
          for (int kb = 0; kb < num_kb && pre < STAGES; ++kb, ++pre, ++stage) {
            mbar_expect_tx(&full_bar[stage], Cfg::stage);
            tma_load_2d(smem + stage * Cfg::stage, &tm_w, &full_bar[stage], kb * kBlockK,
                        tn * kTileN, kEvictNormal);
          }
        }
      }
      pdl_wait();
      int stage = 0, it = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int tm, tn;
        tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          uint8_t* sa = smem + stage * Cfg::stage;
          if (it >= pre) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_expect_tx(&full_bar[stage], Cfg::stage);
            tma_load_2d(sa, &tm_w, &full_bar[stage], kb * kBlockK, tn * kTileN, kEvictNormal);
          }
          tma_load_2d(sa + Cfg::stage_a, &tm_x, &full_bar[stage], kb * kBlockK, tm * BT,
                      kEvictNormal);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc<FMT>(kTileN, BT);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        if constexpr (!GROUPED) {
          mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
          tc_fence_after();
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          if constexpr (GROUPED) {
            mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
            tc_fence_after();
          }
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::stage);
          const uint64_t a_desc = make_sw128_kmajor_desc(sa);
          const uint64_t b_desc = make_sw128_kmajor_desc(sa + Cfg::stage_a);
          const uint32_t d_tmem = tmem_base + acc * BT;
#pragma unroll
          for (int k = 0; k < kBlockK / 32; ++k) {
            const uint32_t accumulate = GROUPED ? (k > 0) : (kb > 0 || k > 0);
            // +32 bytes of K inside the swizzle atom = +2 in the (addr >> 4) field
            if constexpr (FMT == QL_INT8)
              mma_i8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, accumulate);
            else
              mma_f8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, accumulate);
          }
          tc_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
          if constexpr (GROUPED) {
            tc_commit(&tfull_bar[acc]);
            if (++acc == NACC) { acc = 0; acc_phase ^= 1; }
          }
        }
        if constexpr (!GROUPED) {
          tc_commit(&tfull_bar[acc]);
          if (++acc == NACC) { acc = 0; acc_phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (TMEM -> registers -> smem transpose -> global) ==========
    const int q = (warp - 4) & 3;   // TMEM lane quarter: this warp reads lanes [32q, 32q+32) (= warp % 4)
    const int half = (warp - 4) >> 2;   // per-group mode: which half of the token columns
    constexpr int HALVES = EpiCfg<GROUPED, BT>::halves;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    uint16_t* stg = reinterpret_cast<uint16_t*>(epi_smem + (warp - 4) * kEpiWarpSmem);  // [32 m][32 n]
    float* sxs = reinterpret_cast<float*>(stg + 32 * 32);                      // [32 m]
    const bool vec_ok = (p.ldy % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.y) & 15) == 0);
    int acc = 0;
    uint32_t acc_phase = 0;
    pdl_wait();   // sx and the y buffer belong to the previous kernels

    // v[i] = value of (token mc + i, this thread's feature n), already scaled; -> bf16 -> y
    auto store_chunk = [&](const float (&v)[32], int mc, int nq0) {
#pragma unroll
      for (int i = 0; i < 32; ++i) stg[i * 32 + lane] = f2bf(v[i]);
      __syncwarp();
      if (vec_ok) {
        const int nb = nq0 + (lane & 3) * 8;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int row = j * 8 + (lane >> 2);
          const int m = mc + row;
          const uint4 val = *reinterpret_cast<const uint4*>(stg + row * 32 + (lane & 3) * 8);
          if (m < p.M) {
            uint16_t* dst = p.y + (int64_t)m * p.ldy + nb;
            if (nb + 8 <= p.N) {
              *reinterpret_cast<uint4*>(dst) = val;
            } else if (nb < p.N) {
              const uint16_t* sv = stg + row * 32 + (lane & 3) * 8;
              for (int e = 0; e < p.N - nb; ++e) dst[e] = sv[e];
            }
          }
        }
      } else {
        const int n = nq0 + lane;
        for (int row = 0; row < 32; ++row) {
          const int m = mc + row;
          if (m < p.M && n < p.N) p.y[(int64_t)m * p.ldy + n] = stg[row * 32 + lane];
        }
      }
      __syncwarp();
    };

    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int tm, tn;
      tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, tm, tn);
      const int nq0 = tn * kTileN + q * 32;
      const int n = nq0 + lane;
      const bool n_ok = n < p.N;
      const int m0 = tm * BT;
      const bool has_bias = p.bias != nullptr;
      const float biasv = (has_bias && n_ok) ? bf16_lo(p.bias[n]) : 0.0f;

      if constexpr (OUT_ACC) {
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < BT / 32; ++c) {
          uint32_t r[32];
          tmem_ld32(lane_addr + acc * BT + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int m = m0 + c * 32 + i;
            if (m < p.M && n_ok) p.acc_out[(int64_t)m * p.N + n] = (int32_t)r[i];
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        if (++acc == NACC) { acc = 0; acc_phase ^= 1; }
      } else if constexpr (!GROUPED) {
        const float swn = n_ok ? __ldg(p.sw + n) : 0.0f;
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < BT / 32; ++c) {
          const int mc = m0 + c * 32;
          if (mc >= p.M) break;  // whole chunk past the last token (warp-uniform)
          uint32_t r[32];
          tmem_ld32(lane_addr + acc * BT + c * 32, r);
          sxs[lane] = (mc + lane < p.M) ? __ldg(p.sx + mc + lane) : 0.0f;
          __syncwarp();
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 s4 = *reinterpret_cast<const float4*>(sxs + i);
            const float sxv[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float a = FMT == QL_INT8 ? __int2float_rn((int)r[i + e]) : __uint_as_float(r[i + e]);
              float t = __fmul_rn(__fmul_rn(a, sxv[e]), swn);  // (acc * sa) * sb, reference :121
              if (has_bias) t = __fadd_rn(t, biasv);
              v[i + e] = t;
            }
          }
          store_chunk(v, mc, nq0);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        if (++acc == NACC) { acc = 0; acc_phase ^= 1; }
      } else if (half < HALVES) {
        // per-group: this warp promotes token columns [half * CW, (half + 1) * CW) of every K block.
        // The tcgen05.ld of chunk c+1 is in flight while chunk c is promoted.
        constexpr int CW = BT / HALVES;
        constexpr int NCH = CW / 32;
        const uint32_t col0 = (uint32_t)(half * CW);
        float t[CW];
#pragma unroll
        for (int i = 0; i < CW; ++i) t[i] = 0.0f;
        const float* swrow = p.sw + (int64_t)(n_ok ? n : 0) * p.G;
        float sc = __ldg(swrow);
        for (int kb = 0; kb < num_kb; ++kb) {
          const int nxt = (kb + 1 < num_kb) ? (kb + 1) / p.kb_per_g : 0;
          const float sc_next = __ldg(swrow + nxt);
          mbar_wait(&tfull_bar[acc], acc_phase);
          tc_fence_after();
          uint32_t r[2][32];
          tmem_ld32(lane_addr + acc * BT + col0, r[0]);
          tmem_ld_wait_regs(r[0]);
#pragma unroll
          for (int c = 0; c < NCH; ++c) {
            if (c + 1 < NCH) tmem_ld32(lane_addr + acc * BT + col0 + (c + 1) * 32, r[(c + 1) & 1]);
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float a = FMT == QL_INT8 ? __int2float_rn((int)r[c & 1][i]) : __uint_as_float(r[c & 1][i]);
              t[c * 32 + i] = fmaf(a, sc, t[c * 32 + i]);
            }
            if (c + 1 < NCH) tmem_ld_wait_regs(r[(c + 1) & 1]);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[acc]);
          if (++acc == NACC) { acc = 0; acc_phase ^= 1; }
          sc = sc_next;
        }
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          const int mc = m0 + (int)col0 + c * 32;
          if (mc < p.M) {
            sxs[lane] = (mc + lane < p.M) ? __ldg(p.sx + mc + lane) : 0.0f;
            __syncwarp();
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              float u = __fmul_rn(t[c * 32 + i], sxs[i]);
              if (has_bias) u = __fadd_rn(u, biasv);
              v[i] = u;
            }
            store_chunk(v, mc, nq0);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, Cfg::tmem_cols);
}

// ------------------------------------------------------------------ host side

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// A tensor map is a pure function of its arguments (the driver call touches no device state), so
// the descriptors of the eager path -- the same weights, activations and outputs every step -- are
// kept in a small per-thread direct-mapped cache instead of being re-encoded on every launch.
struct TmapKey {
  const void* base;
  int64_t inner, outer, stride;
  int dt, box_inner, box_outer;
  bool operator==(const TmapKey& o) const {
    return base == o.base && inner == o.inner && outer == o.outer && stride == o.stride && dt == o.dt &&
           box_inner == o.box_inner && box_outer == o.box_outer;
  }
};
struct TmapSlot {
  TmapKey key;
  bool valid;
  CUtensorMap map;
};
constexpr int kTmapSlots = 1024;   // power of two; 28 layers x 4 linears x 3 maps fit several times over

static inline uint32_t tmap_hash(const TmapKey& k) {
  uint64_t h = reinterpret_cast<uint64_t>(k.base) * 0x9E3779B97F4A7C15ull;
  h ^= (uint64_t)k.outer * 0xC2B2AE3D27D4EB4Full + (uint64_t)k.inner * 0x165667B19E3779F9ull;
  h ^= ((uint64_t)k.stride << 20) ^ ((uint64_t)k.box_outer << 8) ^ (uint64_t)k.box_inner ^ ((uint64_t)k.dt << 40);
  return (uint32_t)(h >> 40) & (kTmapSlots - 1);
}

bool make_tmap_2d(CUtensorMap* tm, const void* base, CUtensorMapDataType dt, int64_t inner,
                  int64_t outer, int64_t row_stride_bytes, int box_inner, int box_outer) {
  static thread_local TmapSlot* cache = nullptr;
  if (cache == nullptr) cache = new TmapSlot[kTmapSlots]();   // lives as long as the thread's launches
  const TmapKey key{base, inner, outer, row_stride_bytes, (int)dt, box_inner, box_outer};
  TmapSlot& slot = cache[tmap_hash(key)];
  if (slot.valid && slot.key == key) {
    *tm = slot.map;
    return true;
  }
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t gstride[1] = {(cuuint64_t)row_stride_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  if (fn(tm, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  slot.key = key;
  slot.map = *tm;
  slot.valid = true;
  return true;
}

// The same tensor without swizzle (dense box rows in shared memory, boxes wider than 128 bytes);
// not cached: the decode chain encodes its maps once, in ql_chain_prepare.
bool make_tmap_2d_plain(CUtensorMap* tm, const void* base, CUtensorMapDataType dt, int64_t inner, int64_t outer,
                        int64_t row_stride_bytes, int box_inner, int box_outer) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t gstride[1] = {(cuuint64_t)row_stride_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  return fn(tm, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// 2-D uint8 tensor [rows, K] (K contiguous), box = {128 K-bytes, box_rows}
static bool make_tmap(CUtensorMap* tm, const void* base, int64_t rows, int64_t K, int box_rows) {
  return make_tmap_2d(tm, base, CU_TENSOR_MAP_DATA_TYPE_UINT8, K, rows, K, kBlockK, box_rows);
}

template <int FMT, bool GROUPED, int BT, bool OUT_ACC>
static int launch_cfg(const void* qx, const void* qw, GemmParams p, cudaStream_t st) {
  using Cfg = GemmCfg<BT>;
  auto kern = gemm_kernel<FMT, GROUPED, BT, OUT_ACC>;
  static SmemAttrOnce smem_once;   // per kernel instantiation, per device ordinal
  if (int e = ensure_max_smem(reinterpret_cast<const void*>(kern), Cfg::smem_bytes, smem_once)) return e;
  CUtensorMap tm_w, tm_x;
  if (!make_tmap(&tm_w, qw, p.N, p.K, kTileN) || !make_tmap(&tm_x, qx, p.M, p.K, BT))
    return QL_ERR_CUDA;
  p.tiles_n = (p.N + kTileN - 1) / kTileN;
  p.tiles_m = (p.M + BT - 1) / BT;
  {
    const int64_t g = (48ll << 20) / ((int64_t)BT * p.K);   // activations of a super-group stay in L2
    p.group_m = (int)(g < 1 ? 1 : (g > p.tiles_m ? p.tiles_m : g));
  }
  const int64_t tiles = (int64_t)p.tiles_n * p.tiles_m;
  const int grid = (int)(tiles < num_sms() ? tiles : num_sms());
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(EpiCfg<GROUPED, BT>::threads);
  cfg.dynamicSmemBytes = Cfg::smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, kern, tm_w, tm_x, p);
  count_launch();
  if (err != cudaSuccess) {
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  return QL_OK;
}

// Token-tile width: smallest tile covering M for M <= 128; above that the width with the
// fewest tensor-core cycles over whole waves (tile time ~ BT per K block).
static int pick_bt(int64_t M, int64_t N, bool grouped) {
  if (M <= 32) return 32;
  if (M <= 64) return 64;
  if (M <= 128 || grouped) return 128;
  const int64_t tn = (N + kTileN - 1) / kTileN;
  const int sms = num_sms();
  const int64_t c128 = ((tn * ((M + 127) / 128) + sms - 1) / sms) * 128;
  const int64_t c256 = ((tn * ((M + 255) / 256) + sms - 1) / sms) * 256;
  return c128 < c256 ? 128 : 256;
}

template <int FMT, bool GROUPED, bool OUT_ACC>
static int launch_bt(const void* qx, const void* qw, const GemmParams& p, cudaStream_t st) {
  switch (pick_bt(p.M, p.N, GROUPED)) {
    case 32: return launch_cfg<FMT, GROUPED, 32, OUT_ACC>(qx, qw, p, st);
    case 64: return launch_cfg<FMT, GROUPED, 64, OUT_ACC>(qx, qw, p, st);
    case 128: return launch_cfg<FMT, GROUPED, 128, OUT_ACC>(qx, qw, p, st);
    default:
      if constexpr (GROUPED) return QL_ERR_BAD_SHAPE;
      else return launch_cfg<FMT, GROUPED, 256, OUT_ACC>(qx, qw, p, st);
  }
}

// Kernel choice between the CTA-pair kernel (256x256 tiles) and the single-CTA kernel
// (128 x {32..256} tiles, the only one with per-group scales).
// QL_FORCE_GEMM=1|2 (read once) pins the choice for sweeps and tests.
int gemm_kernel_choice(int64_t M, int64_t N, int64_t K, int64_t group) {
  (void)K;
  if (group > 0) {
    // per-group: the CTA-pair kernel (3) from 256 tokens on; QL_GROUPED_PAIR=0 keeps the single-CTA kernel
    static int pair_ok = -1;
    if (pair_ok < 0) {
      const char* e = getenv("QL_GROUPED_PAIR");
      pair_ok = (e != nullptr && e[0] == '0') ? 0 : ((e != nullptr && e[0] == '2') ? 2 : 1);   // 2: always
    }
    // measured (profiles/r2k_sweep_*): the pair kernel wins once its 256 x 256 tiles fill the 74 pairs
    // and a tile holds >= 16 K blocks; with K = 1024 (8 blocks) the per-tile costs (panel staging,
    // direct stores) outweigh its faster promotion
    const int64_t tiles = ((M + 255) / 256) * ((N + 255) / 256);
    return (pair_ok && M >= 256 && (pair_ok == 2 || (K >= 2048 && tiles >= 64))) ? 3 : 1;
  }
  static int forced = -1;
  if (forced < 0) {
    const char* e = getenv("QL_FORCE_GEMM");
    forced = (e != nullptr && (e[0] == '1' || e[0] == '2')) ? e[0] - '0' : 0;
  }
  if (forced) return forced;
  // measured (profiles/r1n_dispatch_crossover.log): from 128 tokens on the pair kernel wins on
  // every Qwen3 shape, even with as few as 4 tiles (its TMA-store epilogue is much shorter)
  (void)N;
  return M >= 128 ? 2 : 1;
}

int launch_gemm(const void* qx, const float* sx, const void* qw, const float* sw,
                const void* bias, void* y, int64_t ldy, int32_t* out_acc, int64_t M, int64_t N,
                int64_t K, int fmt, int64_t group, cudaStream_t st) {
  if (M == 0 || N == 0) return QL_OK;
  const int choice = gemm_kernel_choice(M, N, K, group);
  if (choice == 2 && gemm2_eligible(y, ldy, sw, bias, out_acc != nullptr))
    return launch_gemm2(qx, sx, qw, sw, bias, y, ldy, out_acc, M, N, K, fmt, st);
  if (choice == 3 && out_acc == nullptr && gemm2_grouped_eligible(y, ldy, sw, K, group))
    return launch_gemm2_grouped(qx, sx, qw, sw, bias, y, ldy, M, N, K, fmt, group, st);
  GemmParams p;
  p.sx = sx; p.sw = sw;
  p.bias = static_cast<const uint16_t*>(bias);
  p.y = static_cast<uint16_t*>(y);
  p.ldy = ldy;
  p.acc_out = out_acc;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.G = group > 0 ? (int)(K / group) : 1;
  p.kb_per_g = group > 0 ? (int)(group / kBlockK) : 1;
  p.tiles_n = p.tiles_m = 0;
  p.group_m = 1;
  if (out_acc != nullptr) return launch_bt<QL_INT8, false, true>(qx, qw, p, st);
  if (fmt == QL_INT8)
    return group > 0 ? launch_bt<QL_INT8, true, false>(qx, qw, p, st)
                     : launch_bt<QL_INT8, false, false>(qx, qw, p, st);
  return group > 0 ? launch_bt<QL_FP8_E4M3, true, false>(qx, qw, p, st)
                   : launch_bt<QL_FP8_E4M3, false, false>(qx, qw, p, st);
}

}  // namespace ql
