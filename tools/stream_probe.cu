// HBM streaming probe for the decode kernel's weight fetch pattern (measurement tool, not product):
// how fast can 148 SMs pull an [N, K] byte matrix through shared memory with
//   mode 0: 2-D tiled TMA, box = 128 K-bytes x 128 rows (the UMMA SW128 operand tile), KPS boxes per stage
//   mode 1: 1-D bulk copies of contiguous KPS*16 KB chunks (what a pre-tiled private weight copy allows)
//   mode 2: plain 16-byte LDG by every thread (no shared memory)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/stream_probe tools/stream_probe.cu -lcuda
// run:   stream_probe N K
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_2d(void* dst, const void* desc, uint64_t* bar, int c0, int c1, uint64_t hint) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
               ::"r"(smem_u32(dst)), "l"(desc), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(hint) : "memory");
}
__device__ __forceinline__ void bulk_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t hint) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(hint) : "memory");
}
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;

struct P {
  const uint8_t* w;
  int N, K, mode, kps, stages, hint;
  unsigned long long* sink;
};

__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tm, const P p) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full[16], empty[16];
  const int stage_bytes = p.kps * 16384;
  const int nkb = p.K / 128, tiles = (p.N + 127) / 128;
  const int gpt = nkb / p.kps;                      // stage-groups per tile
  const long long units = (long long)tiles * gpt;
  const long long u0 = units * blockIdx.x / gridDim.x, u1 = units * (blockIdx.x + 1) / gridDim.x;
  if (p.mode == 2) {
    const uint4* src = (const uint4*)p.w;
    const long long nvec = (long long)p.N * p.K / 16;
    uint32_t acc = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x * 4) {
      uint4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        long long idx = i + (long long)j * gridDim.x * blockDim.x;
        v[j] = idx < nvec ? __ldcs(src + idx) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) acc ^= v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
    }
    if (acc == 0x12345678u) p.sink[0] = acc;
    return;
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint64_t hint = p.hint ? kEvictFirst : kEvictNormal;
  if (threadIdx.x == 0) {
    int stage = 0; uint32_t phase = 0;
    for (long long u = u0; u < u1; ++u) {
      mbar_wait(&empty[stage], phase ^ 1);
      const int tile = (int)(u / gpt), kg = (int)(u % gpt);
      uint8_t* dst = smem + stage * stage_bytes;
      mbar_expect_tx(&full[stage], stage_bytes);
      if (p.mode == 0) {
        for (int j = 0; j < p.kps; ++j)
          tma_2d(dst + j * 16384, &tm, &full[stage], (kg * p.kps + j) * 128, tile * 128, hint);
      } else {
        bulk_1d(dst, p.w + ((long long)tile * nkb + kg * p.kps) * 16384, stage_bytes, &full[stage], hint);
      }
      if (++stage == p.stages) { stage = 0; phase ^= 1; }
    }
  } else if (threadIdx.x == 32) {
    int stage = 0; uint32_t phase = 0;
    for (long long u = u0; u < u1; ++u) {
      mbar_wait(&full[stage], phase);
      mbar_arrive(&empty[stage]);
      if (++stage == p.stages) { stage = 0; phase ^= 1; }
    }
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 24576, K = argc > 2 ? atoi(argv[2]) : 4096;
  const size_t bytes = (size_t)N * K;
  const int copies = (int)((400ull << 20) / bytes) + 2;
  uint8_t* w;
  CK(cudaMalloc(&w, bytes * copies));
  CK(cudaMemset(w, 1, bytes * copies));
  unsigned long long* sink;
  CK(cudaMalloc(&sink, 8));
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &qr));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  printf("N=%d K=%d  %.1f MB, %d copies\n", N, K, bytes / 1e6, copies);
  struct Cfg { int mode, kps, stages, cps, hint, promo; };
  const Cfg cfgs[] = {
      {2, 1, 1, 8, 0, 0},
      {0, 1, 6, 2, 1, 2}, {0, 1, 6, 2, 0, 2}, {0, 1, 6, 2, 1, 1}, {0, 1, 6, 2, 1, 0},
      {0, 1, 12, 1, 1, 2}, {0, 2, 6, 1, 1, 2}, {0, 4, 3, 1, 1, 2}, {0, 4, 3, 1, 1, 0}, {0, 2, 3, 2, 1, 2},
      {0, 8, 1, 1, 1, 2}, {0, 4, 1, 2, 1, 2},
      {1, 1, 6, 2, 1, 0}, {1, 1, 12, 1, 1, 0}, {1, 2, 6, 1, 1, 0}, {1, 4, 3, 1, 1, 0}, {1, 4, 3, 1, 0, 0},
  };
  for (const Cfg& c : cfgs) {
    if ((K / 128) % c.kps) continue;
    const CUtensorMapL2promotion promo = c.promo == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B
                                         : c.promo == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B
                                                        : CU_TENSOR_MAP_L2_PROMOTION_NONE;
    const int smem = c.mode == 2 ? 0 : c.kps * 16384 * c.stages + 1024;
    const int grid = 148 * c.cps;
    const int reps = 24;
    float ms_pdl[2];
    for (int pdl = 0; pdl < 2; ++pdl) {
      for (int pass = 0; pass < 2; ++pass) {   // pass 0 = warm-up
        CK(cudaEventRecord(e0));
        for (int r = 0; r < reps; ++r) {
          const uint8_t* base = w + (size_t)(r % copies) * bytes;
          CUtensorMap tm;
          cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)N};
          cuuint64_t gstr[1] = {(cuuint64_t)K};
          cuuint32_t box[2] = {128, 128}, es[2] = {1, 1};
          if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void*)base, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
          P p = {base, N, K, c.mode, c.kps, c.stages, c.hint, sink};
          cudaLaunchConfig_t lc = {};
          lc.gridDim = dim3(grid); lc.blockDim = dim3(128); lc.dynamicSmemBytes = smem;
          cudaLaunchAttribute at[1];
          at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
          at[0].val.programmaticStreamSerializationAllowed = 1;
          lc.attrs = at; lc.numAttrs = pdl ? 1 : 0;
          CK(cudaLaunchKernelEx(&lc, probe, tm, p));
        }
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
      }
      CK(cudaEventElapsedTime(&ms_pdl[pdl], e0, e1));
    }
    printf("mode=%d kps=%d stages=%2d cta/sm=%d hint=%d promo=%d : %.2f us/launch %.0f GB/s | PDL %.2f us %.0f GB/s\n", c.mode, c.kps,
           c.stages, c.cps, c.hint, c.promo, ms_pdl[0] / reps * 1e3, bytes * reps / ms_pdl[0] / 1e6, ms_pdl[1] / reps * 1e3,
           bytes * reps / ms_pdl[1] / 1e6);
  }
  return 0;
}
