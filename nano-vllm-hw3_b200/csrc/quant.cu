// Quantiser kernels (memory-bound, CUDA cores): per-row/per-token (Triton semantics of the
// reference, nanovllm/utils/quantization.py:13-51) and per-group (torch semantics, :194-234).
//
// HBM layout: x bf16 [M,K] (row stride ldx), q 1 B [M,K] dense, s fp32 [M] / [N,K/g].
// Algorithmic bytes per row: 2K read + K written + 4.  One warp owns one row: 16-byte
// coalesced loads (512 B per warp-instruction), amax by warp shuffle, 8-byte stores.
#include "common.cuh"
#include "kernels.h"

namespace ql {

// WPR warps own one row (128-thread blocks, 4 / WPR rows per block).  A lane keeps NV 16-byte
// vectors (8 bf16 each) in registers between the amax pass and the quantise pass, so a row of
// up to WPR * NV * 256 elements is read exactly once.  NV is kept <= 8 (about 64 registers) so
// that 6+ blocks are resident per SM and the grid covers the machine in about one wave; the
// cross-warp amax (WPR > 1) goes through shared memory.  NV = 0: rows longer than 16 K elements
// are read twice (the second read comes from L2).
template <int FMT, int NV, int WPR>
__global__ void __launch_bounds__(128)
quant_rows_kernel(const uint16_t* __restrict__ x, int64_t M, int K, int64_t ldx,
                  uint8_t* __restrict__ q, float* __restrict__ s) {
  __shared__ float red[4];
  // Programmatic dependent launch: let the consumer (the decode kernel streams its weights while
  // it waits for qx) start right away; x itself belongs to the previous kernel.
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int ROWS = 4 / WPR;
  const int64_t row = (int64_t)blockIdx.x * ROWS + warp / WPR;
  const int part = warp % WPR;                      // this warp's slice of the row
  const bool live = row < M;
  const int64_t rowc = live ? row : 0;
  const uint4* xr = reinterpret_cast<const uint4*>(x + rowc * ldx);
  uint2* qr = reinterpret_cast<uint2*>(q + rowc * (int64_t)K);
  const int nvec = K >> 3;
  constexpr float qmax = Fmt<FMT>::qmax;

  float amax = 0.0f;
  if constexpr (NV > 0) {
    uint4 v[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = (i * WPR + part) * 32 + lane;
      v[i] = (live && idx < nvec) ? ldg_stream(xr + idx) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) amax = fmaxf(amax, amax8(v[i]));
    amax = warp_max(amax);
    if constexpr (WPR > 1) {
      if (lane == 0) red[warp] = amax;
      __syncthreads();
      const int w0 = (warp / WPR) * WPR;
#pragma unroll
      for (int i = 0; i < WPR; ++i) amax = fmaxf(amax, red[w0 + i]);
    }
    amax = fmaxf(amax, 1e-8f);
    const float inv = __fdiv_rn(qmax, amax);
    if (live && part == 0 && lane == 0) s[row] = __fdiv_rn(amax, qmax);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = (i * WPR + part) * 32 + lane;
      if (live && idx < nvec) qr[idx] = quant8_mul<FMT>(v[i], inv);
    }
  } else {
    static_assert(NV > 0 || WPR == 4, "long rows use the whole block");
    if (live)
      for (int idx = threadIdx.x; idx < nvec; idx += 128) amax = fmaxf(amax, amax8(__ldg(xr + idx)));
    amax = warp_max(amax);
    if (lane == 0) red[warp] = amax;
    __syncthreads();
    amax = fmaxf(fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3])), 1e-8f);
    const float inv = __fdiv_rn(qmax, amax);
    if (live && threadIdx.x == 0) s[row] = __fdiv_rn(amax, qmax);
    if (live)
      for (int idx = threadIdx.x; idx < nvec; idx += 128) qr[idx] = quant8_mul<FMT>(__ldg(xr + idx), inv);
  }
}

// Unaligned / K % 8 != 0 rows: scalar loads, same arithmetic.  Slow path, kept so that the
// entry point accepts everything the reference's per_row_quant accepts.
template <int FMT>
__global__ void __launch_bounds__(256)
quant_rows_scalar_kernel(const uint16_t* __restrict__ x, int64_t M, int K, int64_t ldx,
                         uint8_t* __restrict__ q, float* __restrict__ s) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  const uint16_t* xr = x + row * ldx;
  uint8_t* qr = q + row * (int64_t)K;
  constexpr float qmax = Fmt<FMT>::qmax;
  float amax = 0.0f;
  for (int k = lane; k < K; k += 32) amax = fmaxf(amax, fabsf(bf16_lo(xr[k])));
  amax = fmaxf(warp_max(amax), 1e-8f);
  const float inv = __fdiv_rn(qmax, amax);
  if (lane == 0) s[row] = __fdiv_rn(amax, qmax);
  for (int k = lane; k < K; k += 32)
    qr[k] = (uint8_t)(quant_pair_away<FMT>(__fmul_rn(bf16_lo(xr[k]), inv), 0.0f) & 0xff);
}

// Per-group weight quantiser (load time).  One warp per (row, group): amax over the group,
// s = amax/qmax, q = quant(w / s) with torch rounding.  Any group size that divides K.
template <int FMT>
__global__ void __launch_bounds__(256)
quant_groups_kernel(const uint16_t* __restrict__ w, int64_t N, int K, int g,
                    uint8_t* __restrict__ q, float* __restrict__ s) {
  const int lane = threadIdx.x & 31;
  const int G = K / g;
  const int64_t gid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (gid >= N * (int64_t)G) return;
  const uint16_t* wr = w + gid * g;  // [N,K] dense => group gid starts at gid*g
  uint8_t* qr = q + gid * g;
  constexpr float qmax = Fmt<FMT>::qmax;
  float amax = 0.0f;
  for (int k = lane; k < g; k += 32) amax = fmaxf(amax, fabsf(bf16_lo(wr[k])));
  amax = fmaxf(warp_max(amax), 1e-8f);
  const float sc = __fdiv_rn(amax, qmax);
  if (lane == 0) s[gid] = sc;
  for (int k = lane; k < g; k += 32)
    qr[k] = (uint8_t)(quant_pair_even<FMT>(__fdiv_rn(bf16_lo(wr[k]), sc), 0.0f) & 0xff);
}

// Per-tensor quantiser (load time), torch semantics of per_tensor_quant (reference :166-191):
// the scale stays in the input dtype (bf16) and x / scale is a bf16 division.
//   pass 1: amax over the tensor -> ws[0] (uint32 bit pattern of a non-negative float; atomicMax)
//   pass 2: scale = bf16(max(amax, bf16(1e-8)) / qmax); q = quant(bf16(x / scale))
__global__ void __launch_bounds__(256)
tensor_amax_kernel(const uint16_t* __restrict__ x, int64_t numel, uint32_t* __restrict__ ws) {
  float amax = 0.0f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
       i += (int64_t)gridDim.x * blockDim.x)
    amax = fmaxf(amax, fabsf(bf16_lo(x[i])));
  amax = warp_max(amax);
  if ((threadIdx.x & 31) == 0) atomicMax(ws, __float_as_uint(amax));
}

__device__ __forceinline__ float round_bf16(float f) { return bf16_lo(f2bf(f)); }

template <int FMT>
__global__ void __launch_bounds__(256)
tensor_quant_kernel(const uint16_t* __restrict__ x, int64_t numel, const uint32_t* __restrict__ ws,
                    uint8_t* __restrict__ q, uint16_t* __restrict__ scale_out) {
  constexpr float qmax = Fmt<FMT>::qmax;
  const float amax = fmaxf(__uint_as_float(*ws), round_bf16(1e-8f));
  const float sc = round_bf16(__fdiv_rn(amax, qmax));
  if (blockIdx.x == 0 && threadIdx.x == 0) *scale_out = f2bf(sc);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float v = round_bf16(__fdiv_rn(bf16_lo(x[i]), sc));
    q[i] = (uint8_t)(quant_pair_even<FMT>(v, 0.0f) & 0xff);
  }
}

// ------------------------------------------------------------------ launchers

int launch_quant_tensor(const void* x, int64_t numel, int fmt, void* q, void* scale_bf16,
                        void* ws, cudaStream_t st) {
  if (numel == 0) return QL_OK;
  if (cudaMemsetAsync(ws, 0, 4, st) != cudaSuccess) return QL_ERR_CUDA;
  int64_t blocks = (numel + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const auto* xp = static_cast<const uint16_t*>(x);
  tensor_amax_kernel<<<(unsigned)blocks, 256, 0, st>>>(xp, numel, static_cast<uint32_t*>(ws));
  if (fmt == QL_INT8)
    tensor_quant_kernel<QL_INT8><<<(unsigned)blocks, 256, 0, st>>>(
        xp, numel, static_cast<const uint32_t*>(ws), static_cast<uint8_t*>(q),
        static_cast<uint16_t*>(scale_bf16));
  else
    tensor_quant_kernel<QL_FP8_E4M3><<<(unsigned)blocks, 256, 0, st>>>(
        xp, numel, static_cast<const uint32_t*>(ws), static_cast<uint8_t*>(q),
        static_cast<uint16_t*>(scale_bf16));
  count_launch(2);
  return cudaGetLastError() == cudaSuccess ? QL_OK : QL_ERR_CUDA;
}

template <int FMT>
static int launch_quant_rows_fmt(const void* x, int64_t M, int64_t K, int64_t ldx, void* q,
                                 float* s, cudaStream_t st) {
  const auto* xp = static_cast<const uint16_t*>(x);
  auto* qp = static_cast<uint8_t*>(q);
  const bool vec_ok = (K % 8 == 0) && (ldx % 8 == 0) && ((uintptr_t)x % 16 == 0) &&
                      ((uintptr_t)q % 8 == 0);
  const int K_ = (int)K;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(128);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
#define QL_QROWS(NV, WPR)                                                     \
  do {                                                                        \
    cfg.gridDim = dim3((unsigned)((M + (4 / WPR) - 1) / (4 / WPR)));          \
    (void)cudaLaunchKernelEx(&cfg, quant_rows_kernel<FMT, NV, WPR>, xp, M, K_, ldx, qp, s); \
  } while (0)
  if (!vec_ok)
    quant_rows_scalar_kernel<FMT><<<(unsigned)((M + 7) / 8), 256, 0, st>>>(xp, M, K_, ldx, qp, s);
  else if (K <= 1024) QL_QROWS(4, 1);
  else if (K <= 2048) QL_QROWS(8, 1);
  else if (K <= 3072) QL_QROWS(6, 2);
  else if (K <= 4096) QL_QROWS(8, 2);
  else if (K <= 8192) QL_QROWS(8, 4);
  else if (K <= 12288) QL_QROWS(12, 4);
  else if (K <= 16384) QL_QROWS(16, 4);
  else QL_QROWS(0, 4);
#undef QL_QROWS
  count_launch();
  return cudaGetLastError() == cudaSuccess ? QL_OK : QL_ERR_CUDA;
}

int launch_quant_rows(const void* x, int64_t M, int64_t K, int64_t ldx, int fmt, void* q,
                      float* s, cudaStream_t st) {
  if (M == 0) return QL_OK;
  return fmt == QL_INT8 ? launch_quant_rows_fmt<QL_INT8>(x, M, K, ldx, q, s, st)
                        : launch_quant_rows_fmt<QL_FP8_E4M3>(x, M, K, ldx, q, s, st);
}

int launch_quant_groups(const void* w, int64_t N, int64_t K, int64_t g, int fmt, void* q,
                        float* s, cudaStream_t st) {
  if (N == 0) return QL_OK;
  const int64_t groups = N * (K / g);
  const unsigned blocks = (unsigned)((groups + 7) / 8);
  const auto* wp = static_cast<const uint16_t*>(w);
  auto* qp = static_cast<uint8_t*>(q);
  if (fmt == QL_INT8)
    quant_groups_kernel<QL_INT8><<<blocks, 256, 0, st>>>(wp, N, (int)K, (int)g, qp, s);
  else
    quant_groups_kernel<QL_FP8_E4M3><<<blocks, 256, 0, st>>>(wp, N, (int)K, (int)g, qp, s);
  count_launch();
  return cudaGetLastError() == cudaSuccess ? QL_OK : QL_ERR_CUDA;
}

}  // namespace ql
