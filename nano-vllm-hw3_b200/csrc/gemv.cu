// Decode-regime kernel (M <= 16): per-token activation quantisation fused with a
// weight-streaming GEMV -- ONE launch per quantized linear.  Replaces the
// per_row_quant + _per_row_post_process_mm pair of the reference's per_row_matmul
// (nanovllm/utils/quantization.py:310-346) where the reference launches a 128x128 tile kernel
// on 8..48 CTAs for a single token.
//
// This kernel is HBM-bound: algorithmic bytes = N*K (weights) + S_w + 2*M*K + 2*M*N.
//   * every CTA quantises the M activation rows into shared memory (qx_s, K+16 byte pitch ->
//     conflict-free 16-byte fragment loads); x is L2-resident and tiny next to the weights;
//   * weights stream straight from HBM into registers with 16-byte L1-bypassing loads, 8
//     weight rows x 128 K-bytes per warp step, 4 steps in flight per lane;
//   * the dot products run on the legacy warp-level tensor path (mma.sync m16n8k32, int8 ->
//     int32 exact, e4m3 -> fp32): at M <= 16 the math is <2% of the byte time, and CUDA-core
//     dp4a would already be issue-bound at M = 16.  The K order inside an MMA is permuted
//     (lane j owns bytes [32j, 32j+32) of each 128-byte block) identically for both operands,
//     which leaves every dot product unchanged and keeps the weight loads 16-byte coalesced;
//   * K is split over KSPLIT warps of a CTA when N alone cannot fill 148 SMs; partial sums
//     are combined through shared memory (int32: exact; fp32: order differs from the GEMM).
#include "common.cuh"
#include "kernels.h"

namespace ql {

struct GemvParams {
  const uint16_t* x;      // bf16 [M, ldx] or nullptr
  int64_t ldx;
  const uint8_t* qx_pre;  // [M, K] or nullptr
  const float* sx_pre;    // [M]
  const uint8_t* qw;      // [N, K]
  const float* sw;        // [N] or [N, K/g]
  const uint16_t* bias;   // bf16 [N] or nullptr
  uint16_t* y;            // bf16 [M, ldy]
  int64_t ldy;
  int M, N, K, g;
};

__device__ __forceinline__ void mma_m16n8k32_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_m16n8k32_e4m3(float (&c)[4], uint32_t a0, uint32_t a1,
                                                  uint32_t a2, uint32_t a3, uint32_t b0,
                                                  uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int FMT>
struct AccT { using type = int; };
template <>
struct AccT<QL_FP8_E4M3> { using type = float; };

constexpr int kGemvThreads = 256;
constexpr int kGemvUnroll = 4;  // 128-byte K blocks in flight per lane (8 x LDG.128)

template <int FMT, bool GROUPED, int MROWS, int KSPLIT>
__global__ void __launch_bounds__(kGemvThreads)
gemv_kernel(const GemvParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int K = p.K, M = p.M, N = p.N;
  const int P = K + 16;                                   // qx_s row pitch (bytes)
  uint8_t* qx_s = smem;                                   // [MROWS][P]
  float* sx_s = reinterpret_cast<float*>(smem + (size_t)MROWS * P);  // [16]
  uint32_t* red = reinterpret_cast<uint32_t*>(sx_s + 16);            // [8][32][4] if KSPLIT > 1

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr float qmax = Fmt<FMT>::qmax;

  // ---- prologue: activations -> qx_s / sx_s (warp w owns rows w and w + 8) ------------------
  for (int m = warp; m < MROWS; m += 8) {
    uint8_t* dst = qx_s + (size_t)m * P;
    const int nvec = K >> 3;
    if (m >= M) {
      for (int i = lane; i < nvec; i += 32) *reinterpret_cast<uint2*>(dst + i * 8) = make_uint2(0, 0);
      if (lane == 0) sx_s[m] = 0.0f;
    } else if (p.x != nullptr) {
      const uint4* xr = reinterpret_cast<const uint4*>(p.x + (int64_t)m * p.ldx);
      float amax = 0.0f;
      for (int i = lane; i < nvec; i += 32) amax = fmaxf(amax, amax8(__ldg(xr + i)));
      amax = fmaxf(warp_max(amax), 1e-8f);
      const float inv = __fdiv_rn(qmax, amax);
      if (lane == 0) sx_s[m] = __fdiv_rn(amax, qmax);
      for (int i = lane; i < nvec; i += 32)
        *reinterpret_cast<uint2*>(dst + i * 8) = quant8_mul<FMT>(__ldg(xr + i), inv);
    } else {
      const uint4* qr = reinterpret_cast<const uint4*>(p.qx_pre + (int64_t)m * K);
      for (int i = lane; i < (K >> 4); i += 32) *reinterpret_cast<uint4*>(dst + i * 16) = __ldg(qr + i);
      if (lane == 0) sx_s[m] = p.sx_pre[m];
    }
  }
  __syncthreads();

  // ---- main loop -----------------------------------------------------------------------------
  using acc_t = typename AccT<FMT>::type;
  constexpr int RPC = 8 / KSPLIT;       // 8-row groups per CTA iteration
  const int wk = warp % KSPLIT;         // this warp's K slice
  const int wr = warp / KSPLIT;         // this warp's row group inside the iteration
  const int r = lane >> 2, j = lane & 3;
  const int KB = K >> 7;                // 128-byte K blocks
  const int G = GROUPED ? K / p.g : 1;
  const int kb_per_g = GROUPED ? p.g >> 7 : 1;
  const int64_t RG = ((int64_t)N + 7) >> 3;
  const int64_t iters = (RG + RPC - 1) / RPC;
  const uint8_t* xa_base = qx_s + (size_t)r * P + j * 32;
  const uint8_t* xb_base = xa_base + (size_t)8 * P;

  for (int64_t it = blockIdx.x; it < iters; it += gridDim.x) {
    const int64_t rg = it * RPC + wr;
    const bool active = rg < RG;
    const int64_t n0 = rg * 8;
    const int64_t nload = min(n0 + r, (int64_t)N - 1);  // clamp: rows past N are loaded, not stored
    const uint8_t* wp = p.qw + nload * K + j * 32;
    const int64_t nc0 = min(n0 + 2 * j, (int64_t)N - 1), nc1 = min(n0 + 2 * j + 1, (int64_t)N - 1);

    acc_t acc[4] = {0, 0, 0, 0};      // per-row: running accumulators (int32 exact / fp32)
    float facc[4] = {0.f, 0.f, 0.f, 0.f};  // grouped: scaled fp32 sums

    if (active) {
      for (int kb0 = wk; kb0 < KB; kb0 += KSPLIT * kGemvUnroll) {
        uint4 w0[kGemvUnroll], w1[kGemvUnroll];
        float sc0[kGemvUnroll], sc1[kGemvUnroll];
#pragma unroll
        for (int u = 0; u < kGemvUnroll; ++u) {
          const int kb = kb0 + u * KSPLIT;
          if (kb < KB) {
            w0[u] = ldg_stream(wp + (size_t)kb * 128);
            w1[u] = ldg_stream(wp + (size_t)kb * 128 + 16);
            if constexpr (GROUPED) {
              sc0[u] = __ldg(p.sw + nc0 * G + kb / kb_per_g);
              sc1[u] = __ldg(p.sw + nc1 * G + kb / kb_per_g);
            }
          }
        }
#pragma unroll
        for (int u = 0; u < kGemvUnroll; ++u) {
          const int kb = kb0 + u * KSPLIT;
          if (kb < KB) {
            const uint4 xa0 = *reinterpret_cast<const uint4*>(xa_base + (size_t)kb * 128);
            const uint4 xa1 = *reinterpret_cast<const uint4*>(xa_base + (size_t)kb * 128 + 16);
            uint4 xb0 = make_uint4(0, 0, 0, 0), xb1 = make_uint4(0, 0, 0, 0);
            if constexpr (MROWS == 16) {
              xb0 = *reinterpret_cast<const uint4*>(xb_base + (size_t)kb * 128);
              xb1 = *reinterpret_cast<const uint4*>(xb_base + (size_t)kb * 128 + 16);
            }
            acc_t part[4];
            if constexpr (GROUPED) {
              part[0] = part[1] = part[2] = part[3] = 0;
            } else {
              part[0] = acc[0]; part[1] = acc[1]; part[2] = acc[2]; part[3] = acc[3];
            }
            if constexpr (FMT == QL_INT8) {
              mma_m16n8k32_s8(part, xa0.x, xb0.x, xa0.y, xb0.y, w0[u].x, w0[u].y);
              mma_m16n8k32_s8(part, xa0.z, xb0.z, xa0.w, xb0.w, w0[u].z, w0[u].w);
              mma_m16n8k32_s8(part, xa1.x, xb1.x, xa1.y, xb1.y, w1[u].x, w1[u].y);
              mma_m16n8k32_s8(part, xa1.z, xb1.z, xa1.w, xb1.w, w1[u].z, w1[u].w);
            } else {
              mma_m16n8k32_e4m3(part, xa0.x, xb0.x, xa0.y, xb0.y, w0[u].x, w0[u].y);
              mma_m16n8k32_e4m3(part, xa0.z, xb0.z, xa0.w, xb0.w, w0[u].z, w0[u].w);
              mma_m16n8k32_e4m3(part, xa1.x, xb1.x, xa1.y, xb1.y, w1[u].x, w1[u].y);
              mma_m16n8k32_e4m3(part, xa1.z, xb1.z, xa1.w, xb1.w, w1[u].z, w1[u].w);
            }
            if constexpr (GROUPED) {
              facc[0] = fmaf((float)part[0], sc0[u], facc[0]);
              facc[1] = fmaf((float)part[1], sc1[u], facc[1]);
              facc[2] = fmaf((float)part[2], sc0[u], facc[2]);
              facc[3] = fmaf((float)part[3], sc1[u], facc[3]);
            } else {
              acc[0] = part[0]; acc[1] = part[1]; acc[2] = part[2]; acc[3] = part[3];
            }
          }
        }
      }
    }

    // ---- combine the KSPLIT partial sums ------------------------------------------------------
    if constexpr (KSPLIT > 1) {
      uint32_t* mine = red + (warp * 32 + lane) * 4;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if constexpr (GROUPED) mine[i] = __float_as_uint(facc[i]);
        else if constexpr (FMT == QL_INT8) mine[i] = (uint32_t)acc[i];
        else mine[i] = __float_as_uint(acc[i]);
      }
      __syncthreads();
      if (wk == 0) {
#pragma unroll
        for (int k = 1; k < KSPLIT; ++k) {
          const uint32_t* other = red + ((warp + k) * 32 + lane) * 4;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if constexpr (GROUPED) facc[i] += __uint_as_float(other[i]);
            else if constexpr (FMT == QL_INT8) acc[i] += (int)other[i];
            else acc[i] += __uint_as_float(other[i]);
          }
        }
      }
      __syncthreads();
    }

    // ---- epilogue: scales, bias, bf16 ----------------------------------------------------------
    if (active && wk == 0) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = r + (i >> 1) * 8;
        const int64_t n = n0 + 2 * j + (i & 1);
        if (m < M && n < N) {
          float v;
          if constexpr (GROUPED) {
            v = __fmul_rn(facc[i], sx_s[m]);
          } else {
            // (acc * sa) * sb, reference quantization.py:111,121
            v = __fmul_rn(__fmul_rn((float)acc[i], sx_s[m]), __ldg(p.sw + n));
          }
          if (p.bias != nullptr) v = __fadd_rn(v, bf16_lo(p.bias[n]));
          p.y[(int64_t)m * p.ldy + n] = f2bf(v);
        }
      }
    }
  }
}

// ------------------------------------------------------------------ launcher

template <int FMT, bool GROUPED, int MROWS, int KSPLIT>
static int launch_one(const GemvParams& p, cudaStream_t st) {
  auto kern = gemv_kernel<FMT, GROUPED, MROWS, KSPLIT>;
  const size_t smem = (size_t)MROWS * (p.K + 16) + 64 + (KSPLIT > 1 ? 8 * 32 * 4 * 4 : 0);
  if (smem > 227 * 1024) return QL_ERR_BAD_SHAPE;
  static SmemAttrOnce smem_once;   // per kernel instantiation, per device ordinal
  if (int e = ensure_max_smem(reinterpret_cast<const void*>(kern), 227 * 1024, smem_once)) return e;
  const int64_t RG = ((int64_t)p.N + 7) / 8;
  const int64_t iters = (RG * KSPLIT + 7) / 8;
  int occ = (int)((220 * 1024) / smem);
  occ = occ < 1 ? 1 : (occ > 4 ? 4 : occ);
  int64_t grid = (int64_t)num_sms() * occ;
  if (grid > iters) grid = iters;
  kern<<<(unsigned)grid, kGemvThreads, smem, st>>>(p);
  count_launch();
  return cudaGetLastError() == cudaSuccess ? QL_OK : QL_ERR_CUDA;
}

template <int FMT, bool GROUPED, int MROWS>
static int launch_ksplit(const GemvParams& p, cudaStream_t st) {
  const int64_t RG = ((int64_t)p.N + 7) / 8;
  const int KB = p.K / 128;
  int ks = 1;
  while (ks < 8 && ks * 2 <= KB && (RG * ks + 7) / 8 < num_sms()) ks *= 2;
  switch (ks) {
    case 1: return launch_one<FMT, GROUPED, MROWS, 1>(p, st);
    case 2: return launch_one<FMT, GROUPED, MROWS, 2>(p, st);
    case 4: return launch_one<FMT, GROUPED, MROWS, 4>(p, st);
    default: return launch_one<FMT, GROUPED, MROWS, 8>(p, st);
  }
}

template <int FMT, bool GROUPED>
static int launch_mrows(const GemvParams& p, cudaStream_t st) {
  return p.M <= 8 ? launch_ksplit<FMT, GROUPED, 8>(p, st) : launch_ksplit<FMT, GROUPED, 16>(p, st);
}

int launch_gemv(const void* x, int64_t ldx, const void* qx_pre, const float* sx_pre,
                const void* qw, const float* sw, const void* bias, void* y, int64_t ldy,
                int64_t M, int64_t N, int64_t K, int fmt, int64_t group, cudaStream_t st) {
  if (M == 0 || N == 0) return QL_OK;
  GemvParams p;
  p.x = static_cast<const uint16_t*>(x);
  p.ldx = ldx;
  p.qx_pre = static_cast<const uint8_t*>(qx_pre);
  p.sx_pre = sx_pre;
  p.qw = static_cast<const uint8_t*>(qw);
  p.sw = sw;
  p.bias = static_cast<const uint16_t*>(bias);
  p.y = static_cast<uint16_t*>(y);
  p.ldy = ldy;
  p.M = (int)M; p.N = (int)N; p.K = (int)K; p.g = (int)group;
  const bool grouped = group > 0;
  if (fmt == QL_INT8)
    return grouped ? launch_mrows<QL_INT8, true>(p, st) : launch_mrows<QL_INT8, false>(p, st);
  return grouped ? launch_mrows<QL_FP8_E4M3, true>(p, st)
                 : launch_mrows<QL_FP8_E4M3, false>(p, st);
}

}  // namespace ql
