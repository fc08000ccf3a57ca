// Activation producers fused with the per-token quantiser.
//
// In the reference every quantized linear re-reads its bf16 input to quantise it
// (per_row_matmul -> per_row_quant, nanovllm/utils/quantization.py:310-346).  Three of the four
// linears of a Qwen3 layer are fed by an elementwise/row-wise op whose output nothing else reads:
//   input_layernorm          -> qkv_proj      (models/qwen3.py:156-161, layers/layernorm.py:16-40)
//   post_attention_layernorm -> gate_up_proj  (models/qwen3.py:164-165)
//   SiluAndMul               -> down_proj     (models/qwen3.py:116-118, layers/activation.py:11-14)
// These kernels produce (q [M,K], s [M]) directly, so the bf16 intermediate never goes to HBM
// (2K bytes written + 2K bytes re-read per row saved) and each linear needs one launch less.
// Both are HBM-bound row kernels: one CTA per token row, 16-byte coalesced loads, the row kept
// on chip (shared memory) between the reduction and the quantise pass.
//
// Arithmetic follows the reference's Python source in eager semantics (every torch op rounds
// to its output dtype):
//   RMSNorm   xf = float(x) [+ float(residual)]; residual' = bf16(xf); var = mean(xf^2);
//             t = bf16(xf * rsqrt(var + eps)); h = bf16(float(t) * float(w))     layernorm.py:21-25,33-39
//   SiluAndMul  g = bf16(a / (1 + exp(-a))) (fp32 inside); h = bf16(float(g) * float(b))  activation.py:13-14
//   quantiser   exactly quant_rows_kernel / the Triton kernel (:26-41) applied to h
// The fp32 reductions (sum of squares) use a fixed tree order, so results are deterministic;
// against the reference they are subject to its (unspecified) reduction order: parity is
// tolerance-based for h (<= 1 bf16 ulp) and for the codes (<= 1 step), see tests.
#include <cuda.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace ql {

constexpr int QL_FMT_NONE = -1;   // no quantised output requested (plain bf16 producer)

template <int FMT> struct FmtQ { static constexpr float qmax = 1.0f; };
template <> struct FmtQ<QL_INT8> { static constexpr float qmax = 127.0f; };
template <> struct FmtQ<QL_FP8_E4M3> { static constexpr float qmax = 448.0f; };

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide reductions in a fixed order (warp butterfly, then warps in index order).
template <bool IS_MAX>
__device__ __forceinline__ float block_reduce(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = IS_MAX ? warp_max(v) : warp_sum(v);
  __syncthreads();                       // red[] may still be read from the previous reduction
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = red[0];
  for (int i = 1; i < nw; ++i) r = IS_MAX ? fmaxf(r, red[i]) : r + red[i];
  return r;
}

__device__ __forceinline__ uint32_t pack_bf(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// F.silu on a bf16 tensor computes x / (1 + exp(-x)) in fp32 and rounds to bf16, the product with
// `up` rounds again (activation.py:14).  ex2.approx + rcp.approx are ~2 fp32 ulp off the IEEE
// result, 2^-15 of a bf16 ulp: the rounded bf16 differs from the exact one for ~1e-4 of the
// inputs (and the reference's own inductor-compiled kernel uses approximate instructions too).
// History of this kernel (profiles/r1m_*): exact expf + IEEE division + scalar cvt.rn.bf16.f32
// (an F2F on the quarter-rate XU pipe) ran at 0.40 of HBM peak, instruction-bound; the packed
// conversions / multiply and the approximate sigmoid below bring it to the bandwidth regime.
// For a < -87.3 the clamped sigmoid gives a * 2^-126 (flushed to -0) instead of a value of
// magnitude < 1.1e-36: the tests compare such values absolutely (2^-100), not in ulp.
// sigmoid(a) = 1 / (1 + 2^min(-a log2 e, 126)): the clamp keeps the divisor finite (rcp.approx of
// 2^126 is the smallest normal), so no sign select and no range checks are needed.
__device__ __forceinline__ float silu_f32(float a) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fminf(-1.4426950408889634f * a, 126.0f)));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return a * r;
}
// two elements: bf16x2( bf16(silu(a)) * up ).  The second product runs on the packed bf16x2
// multiplier (exact fp32 product of two bf16 values, rounded once -- the same result as the
// scalar fp32 multiply + cvt.rn), so `up` is never unpacked.
__device__ __forceinline__ uint32_t silu_mul_pair(uint32_t a, uint32_t up) {
  const uint32_t g = pack_bf(silu_f32(bf16_lo(a)), silu_f32(bf16_hi(a)));   // bf16(silu(a))
  uint32_t h;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(h) : "r"(g), "r"(up));
  return h;
}
__device__ __forceinline__ uint4 silu_mul8(const uint4& a, const uint4& b) {
  return make_uint4(silu_mul_pair(a.x, b.x), silu_mul_pair(a.y, b.y), silu_mul_pair(a.z, b.z),
                    silu_mul_pair(a.w, b.w));
}

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf(f[0], f[1]), pack_bf(f[2], f[3]), pack_bf(f[4], f[5]),
                    pack_bf(f[6], f[7]));
}

// Quantise the row held in shared memory as packed bf16 vectors and write the codes.
template <int FMT>
__device__ __forceinline__ void quant_row_from_smem(const uint4* hrow, int hstride_vec, int nvec,
                                                    float amax, uint8_t* qrow, float* s_out) {
  constexpr float qmax = FmtQ<FMT>::qmax;
  amax = fmaxf(amax, 1e-8f);
  const float inv = __fdiv_rn(qmax, amax);
  if (threadIdx.x == 0) *s_out = __fdiv_rn(amax, qmax);
  uint2* qr = reinterpret_cast<uint2*>(qrow);
  for (int idx = threadIdx.x; idx < nvec; idx += blockDim.x)
    qr[idx] = quant8_mul<FMT>(hrow[idx * hstride_vec], inv);
}

// ---------------------------------------------------------------------------------- RMSNorm
// Long-row fallback (K > 16384): one CTA per row, the row staged in shared memory.
// Shared memory: 32 bytes per 8 elements (the fp32 sums xf); after normalisation the packed bf16
// result overwrites the first 16 bytes of the same slot (same thread, no cross-thread hazard).
template <int FMT, bool HAS_RES>
__global__ void __launch_bounds__(512)
rmsnorm_quant_long_kernel(const uint16_t* __restrict__ x, int64_t ldx, const uint16_t* res_in,
                     uint16_t* res_out, const uint16_t* __restrict__ w, float eps, int K,
                     uint8_t* __restrict__ q, float* __restrict__ s, uint16_t* __restrict__ h_out) {
  extern __shared__ uint4 row_smem[];      // [K/8][2] uint4
  __shared__ float red[16];
  pdl_launch_dependents();
  pdl_wait();                              // x / residual belong to the previous kernels
  const int64_t row = blockIdx.x;
  const int nvec = K >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * ldx);
  const uint4* rr = HAS_RES ? reinterpret_cast<const uint4*>(res_in + row * (int64_t)K) : nullptr;
  uint4* ro = HAS_RES ? reinterpret_cast<uint4*>(res_out + row * (int64_t)K) : nullptr;
  const uint4* wv = reinterpret_cast<const uint4*>(w);

  float ss = 0.0f;
  for (int idx = threadIdx.x; idx < nvec; idx += blockDim.x) {
    const uint4 a = ldg_stream(xr + idx);
    float f[8] = {bf16_lo(a.x), bf16_hi(a.x), bf16_lo(a.y), bf16_hi(a.y),
                  bf16_lo(a.z), bf16_hi(a.z), bf16_lo(a.w), bf16_hi(a.w)};
    if constexpr (HAS_RES) {
      const uint4 b = rr[idx];             // plain load: res_out may alias res_in
      const float g[8] = {bf16_lo(b.x), bf16_hi(b.x), bf16_lo(b.y), bf16_hi(b.y),
                          bf16_lo(b.z), bf16_hi(b.z), bf16_lo(b.w), bf16_hi(b.w)};
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = __fadd_rn(f[e], g[e]);
      ro[idx] = make_uint4(pack_bf(f[0], f[1]), pack_bf(f[2], f[3]), pack_bf(f[4], f[5]),
                           pack_bf(f[6], f[7]));
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) ss = __fadd_rn(ss, __fmul_rn(f[e], f[e]));   // x.pow(2), then sum
    float4* dst = reinterpret_cast<float4*>(row_smem + 2 * idx);
    dst[0] = make_float4(f[0], f[1], f[2], f[3]);
    dst[1] = make_float4(f[4], f[5], f[6], f[7]);
  }
  ss = block_reduce<false>(ss, red);
  const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)K), eps));

  float amax = 0.0f;
  for (int idx = threadIdx.x; idx < nvec; idx += blockDim.x) {
    const float4* src = reinterpret_cast<const float4*>(row_smem + 2 * idx);
    const float4 f0 = src[0], f1 = src[1];
    const float f[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
    const uint4 wq = __ldg(wv + idx);
    const float g[8] = {bf16_lo(wq.x), bf16_hi(wq.x), bf16_lo(wq.y), bf16_hi(wq.y),
                        bf16_lo(wq.z), bf16_hi(wq.z), bf16_lo(wq.w), bf16_hi(wq.w)};
    float h[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float t = bf16_lo((uint32_t)f2bf(__fmul_rn(f[e], rs)));   // x.to(bf16)
      h[e] = bf16_lo((uint32_t)f2bf(__fmul_rn(t, g[e])));             // .mul_(weight) in bf16
      amax = fmaxf(amax, fabsf(h[e]));
    }
    const uint4 hv = make_uint4(pack_bf(h[0], h[1]), pack_bf(h[2], h[3]), pack_bf(h[4], h[5]),
                                pack_bf(h[6], h[7]));
    row_smem[2 * idx] = hv;
    if (h_out != nullptr) reinterpret_cast<uint4*>(h_out + row * (int64_t)K)[idx] = hv;
  }
  if constexpr (FMT != QL_FMT_NONE) {
    amax = block_reduce<true>(amax, red);
    quant_row_from_smem<FMT>(row_smem, 2, nvec, amax, q + row * (int64_t)K, s + row);
  }
}

// ---------------------------------------------------------------------------------- SiluAndMul
// Long-row fallback (I > 16384): one CTA per row.
// gate_up row = [gate (I) | up (I)], ld elements between rows.  Shared memory: the packed bf16
// products, 16 bytes per 8 elements.
template <int FMT>
__global__ void __launch_bounds__(512)
silu_mul_quant_long_kernel(const uint16_t* __restrict__ gu, int64_t ld, int I, uint8_t* __restrict__ q,
                      float* __restrict__ s, uint16_t* __restrict__ h_out) {
  extern __shared__ uint4 row_smem[];      // [I/8] uint4
  __shared__ float red[16];
  pdl_launch_dependents();
  pdl_wait();
  const int64_t row = blockIdx.x;
  const int nvec = I >> 3;
  const uint4* ar = reinterpret_cast<const uint4*>(gu + row * ld);
  const uint4* br = reinterpret_cast<const uint4*>(gu + row * ld + I);

  float amax = 0.0f;
  for (int idx = threadIdx.x; idx < nvec; idx += blockDim.x) {
    const uint4 a = ldg_stream(ar + idx), b = ldg_stream(br + idx);
    const uint4 hv = silu_mul8(a, b);
    amax = fmaxf(amax, amax8(hv));
    row_smem[idx] = hv;
    if (h_out != nullptr) reinterpret_cast<uint4*>(h_out + row * (int64_t)I)[idx] = hv;
  }
  if constexpr (FMT != QL_FMT_NONE) {
    amax = block_reduce<true>(amax, red);
    quant_row_from_smem<FMT>(row_smem, 1, nvec, amax, q + row * (int64_t)I, s + row);
  }
}

// ---------------------------------------------------------------------------------- register-resident kernels
// The production shapes (rows up to 16384 elements).  As quant_rows_kernel: WPR warps own one row,
// every lane keeps NV 16-byte vectors of the row in registers between the reduction and the
// quantise pass, so all of a lane's loads are issued back to back (NV, or 2 NV with a residual /
// the up half, independent 16-byte requests per lane -- the bytes in flight, not the arithmetic,
// decide the speed of these kernels) and nothing but the cross-warp amax / sum goes through
// shared memory.  Blocks are 4 warps = 4 / WPR rows (8 warps = one row for WPR = 8).
template <int WPR> struct RowCfg {
  static constexpr int threads = WPR * 32 > 128 ? WPR * 32 : 128;
  static constexpr int rows = threads / 32 / WPR;
};

template <int WPR, bool IS_MAX>
__device__ __forceinline__ float row_reduce(float v, float* red, int warp) {
  v = IS_MAX ? warp_max(v) : warp_sum(v);
  if constexpr (WPR > 1) {
    const int lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    const int w0 = (warp / WPR) * WPR;
    v = red[w0];
#pragma unroll
    for (int i = 1; i < WPR; ++i) v = IS_MAX ? fmaxf(v, red[w0 + i]) : v + red[w0 + i];
  }
  return v;
}

template <int FMT, bool HAS_RES, int NV, int WPR>
__global__ void __launch_bounds__(RowCfg<WPR>::threads)
rmsnorm_quant_kernel(const uint16_t* __restrict__ x, int64_t ldx, const uint16_t* res_in,
                     uint16_t* res_out, const uint16_t* __restrict__ w, float eps, int64_t M, int K,
                     uint8_t* __restrict__ q, float* __restrict__ s, uint16_t* __restrict__ h_out) {
  __shared__ float red[8];
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int ROWS = RowCfg<WPR>::rows;
  const int64_t row = (int64_t)blockIdx.x * ROWS + warp / WPR;
  const int part = warp % WPR;
  const bool live = row < M;
  const int64_t rowc = live ? row : 0;
  const int nvec = K >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + rowc * ldx);
  const uint4* rr = reinterpret_cast<const uint4*>(res_in + rowc * (int64_t)K);
  uint4* ro = reinterpret_cast<uint4*>(res_out + rowc * (int64_t)K);
  const uint4* wv = reinterpret_cast<const uint4*>(w);

  uint4 xv[NV], rv[HAS_RES ? NV : 1];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = (i * WPR + part) * 32 + lane;
    const bool ok = live && idx < nvec;
    xv[i] = ok ? ldg_stream(xr + idx) : make_uint4(0, 0, 0, 0);
    if constexpr (HAS_RES) rv[i] = ok ? rr[idx] : make_uint4(0, 0, 0, 0);   // may alias res_out
  }
  float f[NV][8];
  float ss = 0.0f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    unpack8(xv[i], f[i]);
    if constexpr (HAS_RES) {
      float g[8];
      unpack8(rv[i], g);
#pragma unroll
      for (int e = 0; e < 8; ++e) f[i][e] = __fadd_rn(f[i][e], g[e]);
      const int idx = (i * WPR + part) * 32 + lane;
      if (live && idx < nvec) ro[idx] = pack8(f[i]);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) ss = __fadd_rn(ss, __fmul_rn(f[i][e], f[i][e]));   // x.pow(2), then sum
  }
  ss = row_reduce<WPR, false>(ss, red, warp);
  const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)K), eps));

  uint4 hv[NV];
  float amax = 0.0f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = (i * WPR + part) * 32 + lane;
    const bool ok = live && idx < nvec;
    float g[8], h[8];
    unpack8(ok ? __ldg(wv + idx) : make_uint4(0, 0, 0, 0), g);
#pragma unroll
    for (int e = 0; e < 8; e += 2) {
      const uint32_t t = pack_bf(__fmul_rn(f[i][e], rs), __fmul_rn(f[i][e + 1], rs));   // x.to(bf16)
      h[e] = __fmul_rn(bf16_lo(t), g[e]);                                              // .mul_(weight)
      h[e + 1] = __fmul_rn(bf16_hi(t), g[e + 1]);
    }
    hv[i] = pack8(h);
    amax = fmaxf(amax, amax8(hv[i]));
    if (h_out != nullptr && ok) reinterpret_cast<uint4*>(h_out + rowc * (int64_t)K)[idx] = hv[i];
  }
  if constexpr (FMT != QL_FMT_NONE) {
    constexpr float qmax = FmtQ<FMT>::qmax;
    amax = fmaxf(row_reduce<WPR, true>(amax, red, warp), 1e-8f);
    const float inv = __fdiv_rn(qmax, amax);
    if (live && part == 0 && lane == 0) s[row] = __fdiv_rn(amax, qmax);
    uint2* qr = reinterpret_cast<uint2*>(q + rowc * (int64_t)K);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = (i * WPR + part) * 32 + lane;
      if (live && idx < nvec) qr[idx] = quant8_mul<FMT>(hv[i], inv);
    }
  }
}

template <int FMT, int NV, int WPR>
__global__ void __launch_bounds__(RowCfg<WPR>::threads)
silu_mul_quant_kernel(const uint16_t* __restrict__ gu, int64_t ld, int64_t M, int I,
                      uint8_t* __restrict__ q, float* __restrict__ s, uint16_t* __restrict__ h_out) {
  __shared__ float red[8];
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int ROWS = RowCfg<WPR>::rows;
  const int64_t row = (int64_t)blockIdx.x * ROWS + warp / WPR;
  const int part = warp % WPR;
  const bool live = row < M;
  const int64_t rowc = live ? row : 0;
  const int nvec = I >> 3;
  const uint4* ar = reinterpret_cast<const uint4*>(gu + rowc * ld);
  const uint4* br = reinterpret_cast<const uint4*>(gu + rowc * ld + I);

  uint4 av[NV], bv[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = (i * WPR + part) * 32 + lane;
    const bool ok = live && idx < nvec;
    av[i] = ok ? ldg_stream(ar + idx) : make_uint4(0, 0, 0, 0);
    bv[i] = ok ? ldg_stream(br + idx) : make_uint4(0, 0, 0, 0);
  }
  float amax = 0.0f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    av[i] = silu_mul8(av[i], bv[i]);      // the row's products, packed bf16, stay in registers
    amax = fmaxf(amax, amax8(av[i]));
    const int idx = (i * WPR + part) * 32 + lane;
    if (h_out != nullptr && live && idx < nvec)
      reinterpret_cast<uint4*>(h_out + rowc * (int64_t)I)[idx] = av[i];
  }
  if constexpr (FMT != QL_FMT_NONE) {
    constexpr float qmax = FmtQ<FMT>::qmax;
    amax = fmaxf(row_reduce<WPR, true>(amax, red, warp), 1e-8f);
    const float inv = __fdiv_rn(qmax, amax);
    if (live && part == 0 && lane == 0) s[row] = __fdiv_rn(amax, qmax);
    uint2* qr = reinterpret_cast<uint2*>(q + rowc * (int64_t)I);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = (i * WPR + part) * 32 + lane;
      if (live && idx < nvec) qr[idx] = quant8_mul<FMT>(av[i], inv);
    }
  }
}

// ---------------------------------------------------------------------------------- all-reduce + RMSNorm + quant
// TP decode: o_proj / down_proj (row-parallel) -> all-reduce -> add residual -> RMSNorm -> quant for
// the next qkv / gate_up (reference: linear.py:150-152, then models/qwen3.py:156-165).  One kernel:
// the row layout of rmsnorm_quant_kernel with the one-shot push all-reduce of allreduce.cu as its
// load stage -- every lane pushes its NV vectors of this rank's partial sums to the peers' receive
// areas ({data, epoch} packets over NVLink), picks up the peers' packets for the same vectors from
// its own area, adds them in fp32 in rank order and rounds to bf16 (bit-identical to the separate
// all-reduce kernel), then continues as the RMSNorm producer.  Saves a launch and the bf16 round
// trip of the reduced tensor per row-parallel linear.
struct ArArgs {
  uint8_t* area[8];
  uint32_t* ctr;
  size_t region;
  uint64_t timeout_ns;
  int rank;
};

__device__ __forceinline__ void ar_st_packet(void* p, uint32_t d0, uint32_t d1, uint32_t e) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(d0), "r"(e), "r"(d1),
               "r"(e)
               : "memory");
}
__device__ __forceinline__ uint4 ar_ld_packet(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

template <int FMT, int W, int NV, int WPR>
__global__ void __launch_bounds__(RowCfg<WPR>::threads)
ar_rmsnorm_quant_kernel(const ArArgs ar, const uint16_t* __restrict__ partial, const uint16_t* res_in,
                        uint16_t* res_out, const uint16_t* __restrict__ w, float eps, int64_t M, int K,
                        uint8_t* __restrict__ q, float* __restrict__ s) {
  __shared__ float red[8];
  pdl_launch_dependents();
  pdl_wait();                               // the partial sums come from the row-parallel GEMM
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int ROWS = RowCfg<WPR>::rows;
  const int64_t row = (int64_t)blockIdx.x * ROWS + warp / WPR;
  const int part = warp % WPR;
  const bool live = row < M;
  const int64_t rowc = live ? row : 0;
  const int nvec = K >> 3;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(ar.ctr) + 1;
  const size_t par = (size_t)(e & 1) * W * ar.region;
  const uint4* pr = reinterpret_cast<const uint4*>(partial + rowc * (int64_t)K);
  const uint4* rr = reinterpret_cast<const uint4*>(res_in + rowc * (int64_t)K);
  uint4* ro = reinterpret_cast<uint4*>(res_out + rowc * (int64_t)K);
  const uint4* wv = reinterpret_cast<const uint4*>(w);

  uint4 xv[NV], rv[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = (i * WPR + part) * 32 + lane;
    const bool ok = live && idx < nvec;
    xv[i] = ok ? pr[idx] : make_uint4(0, 0, 0, 0);
    rv[i] = ok ? rr[idx] : make_uint4(0, 0, 0, 0);
    if (ok) {                               // push this rank's vector to every peer
      const size_t off = par + (size_t)ar.rank * ar.region + (size_t)(rowc * nvec + idx) * 32;
#pragma unroll
      for (int d = 1; d < W; ++d) {
        uint8_t* dst = ar.area[(ar.rank + d) % W] + off;
        ar_st_packet(dst, xv[i].x, xv[i].y, e);
        ar_st_packet(dst + 16, xv[i].z, xv[i].w, e);
      }
    }
  }
  float f[NV][8];
  float ss = 0.0f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = (i * WPR + part) * 32 + lane;
    const bool ok = live && idx < nvec;
    float acc[8];
#pragma unroll
    for (int r = 0; r < W; ++r) {           // fp32 sum in rank order, as ar_oneshot_kernel
      uint4 v = xv[i];
      if (r != ar.rank && ok) {
        const uint8_t* src = ar.area[ar.rank] + par + (size_t)r * ar.region +
                             (size_t)(rowc * nvec + idx) * 32;
        uint4 a, b;
        SpinGuard guard(ar.timeout_ns);
        while (true) { a = ar_ld_packet(src); if (a.y == e && a.w == e) break; guard.tick(); }
        while (true) { b = ar_ld_packet(src + 16); if (b.y == e && b.w == e) break; guard.tick(); }
        v = make_uint4(a.x, a.z, b.x, b.z);
      }
      float g[8];
      unpack8(v, g);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] = r == 0 ? g[k] : __fadd_rn(acc[k], g[k]);
    }
    unpack8(pack8(acc), f[i]);              // y = bf16(sum): what dist.all_reduce leaves in y
    float g[8];
    unpack8(rv[i], g);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[i][k] = __fadd_rn(f[i][k], g[k]);
    if (ok) ro[idx] = pack8(f[i]);
#pragma unroll
    for (int k = 0; k < 8; ++k) ss = __fadd_rn(ss, __fmul_rn(f[i][k], f[i][k]));
  }
  ss = row_reduce<WPR, false>(ss, red, warp);
  const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)K), eps));

  uint4 hv[NV];
  float amax = 0.0f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = (i * WPR + part) * 32 + lane;
    const bool ok = live && idx < nvec;
    float g[8], h[8];
    unpack8(ok ? __ldg(wv + idx) : make_uint4(0, 0, 0, 0), g);
#pragma unroll
    for (int k = 0; k < 8; k += 2) {
      const uint32_t t = pack_bf(__fmul_rn(f[i][k], rs), __fmul_rn(f[i][k + 1], rs));
      h[k] = __fmul_rn(bf16_lo(t), g[k]);
      h[k + 1] = __fmul_rn(bf16_hi(t), g[k + 1]);
    }
    hv[i] = pack8(h);
    amax = fmaxf(amax, amax8(hv[i]));
  }
  constexpr float qmax = FmtQ<FMT>::qmax;
  amax = fmaxf(row_reduce<WPR, true>(amax, red, warp), 1e-8f);
  const float inv = __fdiv_rn(qmax, amax);
  if (live && part == 0 && lane == 0) s[row] = __fdiv_rn(amax, qmax);
  uint2* qr = reinterpret_cast<uint2*>(q + rowc * (int64_t)K);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = (i * WPR + part) * 32 + lane;
    if (live && idx < nvec) qr[idx] = quant8_mul<FMT>(hv[i], inv);
  }
  // the last block to finish publishes the epoch for the next call (as ar_oneshot_kernel)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(ar.ctr + 1, 1u) == gridDim.x - 1) {
      ar.ctr[1] = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(ar.ctr) = e;
    }
  }
}

// ---------------------------------------------------------------------------------- host side

static int block_threads(int nvec) { return nvec <= 128 ? 128 : nvec <= 1024 ? 256 : 512; }

template <typename Kern, typename... Args>
static int launch_rows(Kern kern, int64_t rows, int threads, size_t smem, cudaStream_t st,
                       Args... args) {
  if (smem > 48 * 1024) {
    // opt-in above the default limit; cheap enough to repeat per launch for these long rows
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
        cudaSuccess) {
      (void)cudaGetLastError();
      return QL_ERR_BAD_SHAPE;
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)rows);
  cfg.blockDim = dim3((unsigned)threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, kern, args...);
  count_launch();
  if (err != cudaSuccess) {
    (void)cudaGetLastError();
    return QL_ERR_CUDA;
  }
  return QL_OK;
}

// (NV, WPR) for a row of nvec 16-byte vectors: the fewest warps per row that keep NV <= max_nv.
static bool pick_cfg(int nvec, int max_nv, int& nv, int& wpr) {
  // measured on B200 (tools/silu_cfg_probe.py): one row per 4-warp block beats two rows per block
  static const int cfgs[][2] = {{4, 1}, {4, 2}, {4, 4}, {6, 4}, {8, 4}, {6, 8}, {8, 8}};
  if (const char* e = getenv("QL_PROD_CFG")) {   // tuning aid: "NV,WPR" (must cover the row)
    int a = 0, b = 0;
    if (sscanf(e, "%d,%d", &a, &b) == 2 && nvec <= a * b * 32) {
      nv = a; wpr = b;
      return true;
    }
  }
  for (const auto& c : cfgs)
    if (c[0] <= max_nv && nvec <= c[0] * c[1] * 32) {
      nv = c[0]; wpr = c[1];
      return true;
    }
  return false;
}

template <int FMT, bool HAS_RES>
static int launch_rms_fmt(const uint16_t* xp, int64_t ldx, const uint16_t* rp, uint16_t* rop,
                          const uint16_t* wp, float eps, int64_t M, int K, uint8_t* qp, float* s,
                          uint16_t* hp, cudaStream_t st) {
  const int nvec = K / 8;
  int nv = 0, wpr = 0;
  if (!pick_cfg(nvec, 8, nv, wpr))
    return launch_rows(rmsnorm_quant_long_kernel<FMT, HAS_RES>, M, block_threads(nvec),
                       (size_t)nvec * 32, st, xp, ldx, rp, rop, wp, eps, K, qp, s, hp);
  const int rows_pb = wpr >= 4 ? 1 : 4 / wpr;
  const int64_t blocks = (M + rows_pb - 1) / rows_pb;
#define QL_RMS(NV, WPR)                                                                        \
  if (nv == NV && wpr == WPR)                                                                  \
    return launch_rows(rmsnorm_quant_kernel<FMT, HAS_RES, NV, WPR>, blocks, RowCfg<WPR>::threads, 0, st, xp, ldx, \
                       rp, rop, wp, eps, M, K, qp, s, hp)
  QL_RMS(4, 1); QL_RMS(4, 2); QL_RMS(4, 4); QL_RMS(6, 4); QL_RMS(8, 4);
  QL_RMS(6, 8); QL_RMS(8, 8);
#undef QL_RMS
  return QL_ERR_BAD_SHAPE;
}

int launch_rmsnorm_quant(const void* x, int64_t ldx, const void* res_in, void* res_out,
                         const void* w, float eps, int64_t M, int64_t K, int fmt, void* q,
                         float* s, void* h_out, cudaStream_t st) {
  const auto* xp = static_cast<const uint16_t*>(x);
  const auto* rp = static_cast<const uint16_t*>(res_in);
  auto* rop = static_cast<uint16_t*>(res_out);
  const auto* wp = static_cast<const uint16_t*>(w);
  auto* qp = static_cast<uint8_t*>(q);
  auto* hp = static_cast<uint16_t*>(h_out);
  const int K_ = (int)K;
#define QL_RMS(F, R) launch_rms_fmt<F, R>(xp, ldx, rp, rop, wp, eps, M, K_, qp, s, hp, st)
  const bool has_res = res_in != nullptr;
  if (q == nullptr) return has_res ? QL_RMS(QL_FMT_NONE, true) : QL_RMS(QL_FMT_NONE, false);
  if (fmt == QL_INT8) return has_res ? QL_RMS(QL_INT8, true) : QL_RMS(QL_INT8, false);
  return has_res ? QL_RMS(QL_FP8_E4M3, true) : QL_RMS(QL_FP8_E4M3, false);
#undef QL_RMS
}

template <int FMT>
static int launch_silu_fmt(const uint16_t* gp, int64_t ld, int64_t M, int I, uint8_t* qp, float* s,
                           uint16_t* hp, cudaStream_t st) {
  const int nvec = I / 8;
  int nv = 0, wpr = 0;
  if (!pick_cfg(nvec, 8, nv, wpr))
    return launch_rows(silu_mul_quant_long_kernel<FMT>, M, block_threads(nvec), (size_t)nvec * 16,
                       st, gp, ld, I, qp, s, hp);
  const int rows_pb = wpr >= 4 ? 1 : 4 / wpr;
  const int64_t blocks = (M + rows_pb - 1) / rows_pb;
#define QL_SILU(NV, WPR)                                                                       \
  if (nv == NV && wpr == WPR)                                                                  \
    return launch_rows(silu_mul_quant_kernel<FMT, NV, WPR>, blocks, RowCfg<WPR>::threads, 0, st, gp, ld, M, I, qp, \
                       s, hp)
  QL_SILU(4, 1); QL_SILU(4, 2); QL_SILU(4, 4); QL_SILU(6, 4); QL_SILU(8, 4);
  QL_SILU(6, 8); QL_SILU(8, 8);
#undef QL_SILU
  return QL_ERR_BAD_SHAPE;
}

int launch_silu_mul_quant(const void* gate_up, int64_t ld, int64_t M, int64_t I, int fmt, void* q,
                          float* s, void* h_out, cudaStream_t st) {
  const auto* gp = static_cast<const uint16_t*>(gate_up);
  auto* qp = static_cast<uint8_t*>(q);
  auto* hp = static_cast<uint16_t*>(h_out);
  if (q == nullptr) return launch_silu_fmt<QL_FMT_NONE>(gp, ld, M, (int)I, qp, s, hp, st);
  if (fmt == QL_INT8) return launch_silu_fmt<QL_INT8>(gp, ld, M, (int)I, qp, s, hp, st);
  return launch_silu_fmt<QL_FP8_E4M3>(gp, ld, M, (int)I, qp, s, hp, st);
}

template <int FMT, int W>
static int launch_arn_fw(const ArArgs& a, const uint16_t* pp, const uint16_t* rp, uint16_t* rop,
                         const uint16_t* wp, float eps, int64_t M, int K, uint8_t* qp, float* s,
                         cudaStream_t st) {
  int nv = 0, wpr = 0;
  if (!pick_cfg(K / 8, 8, nv, wpr)) return QL_ERR_BAD_SHAPE;
  const int rows_pb = wpr >= 4 ? 1 : 4 / wpr;
  const int64_t blocks = (M + rows_pb - 1) / rows_pb;
#define QL_ARN(NV, WPR)                                                                         \
  if (nv == NV && wpr == WPR)                                                                   \
    return launch_rows(ar_rmsnorm_quant_kernel<FMT, W, NV, WPR>, blocks, RowCfg<WPR>::threads, 0, \
                       st, a, pp, rp, rop, wp, eps, M, K, qp, s)
  QL_ARN(4, 1); QL_ARN(4, 2); QL_ARN(4, 4); QL_ARN(6, 4); QL_ARN(8, 4);
#undef QL_ARN
  return QL_ERR_BAD_SHAPE;               // rows longer than 8192 elements: use the unfused pair
}

int launch_ar_rmsnorm_quant(const ArView& ar, const void* partial, const void* res_in, void* res_out,
                            const void* w, float eps, int64_t M, int64_t K, int fmt, void* q,
                            float* s, cudaStream_t st) {
  ArArgs a;
  for (int r = 0; r < 8; ++r) a.area[r] = ar.area[r];
  a.ctr = ar.ctr;
  a.region = ar.region;
  a.timeout_ns = ar.timeout_ns;
  a.rank = ar.rank;
  const auto* pp = static_cast<const uint16_t*>(partial);
  const auto* rp = static_cast<const uint16_t*>(res_in);
  auto* rop = static_cast<uint16_t*>(res_out);
  const auto* wp = static_cast<const uint16_t*>(w);
  auto* qp = static_cast<uint8_t*>(q);
#define QL_ARN_W(F)                                                                              \
  switch (ar.world) {                                                                            \
    case 2: return launch_arn_fw<F, 2>(a, pp, rp, rop, wp, eps, M, (int)K, qp, s, st);           \
    case 4: return launch_arn_fw<F, 4>(a, pp, rp, rop, wp, eps, M, (int)K, qp, s, st);           \
    case 8: return launch_arn_fw<F, 8>(a, pp, rp, rop, wp, eps, M, (int)K, qp, s, st);           \
    default: return QL_ERR_BAD_SHAPE;                                                            \
  }
  if (fmt == QL_INT8) { QL_ARN_W(QL_INT8) }
  QL_ARN_W(QL_FP8_E4M3)
#undef QL_ARN_W
}

}  // namespace ql
