// NVLink write/read bandwidth probe for the reduce-scatter design (tools only, not part of the
// library).  One process, two GPUs with peer access; each test runs on BOTH GPUs at once (the
// collective's traffic is symmetric) and reports the slower side.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o nvlink_probe nvlink_probe.cu -lcuda
//   ./nvlink_probe [MiB per launch, default 64]
//
// Variants: st.v4 (coalesced thread stores), ld.v4 (coalesced thread loads from the peer),
// bulk-K (cp.async.bulk shared->global, K-byte contiguous chunks), tma-2d (cp.async.bulk.tensor
// 2-D store of [32 rows][128 B] boxes into a row-major matrix with a 2 KB pitch: the GEMM
// epilogue's pattern), each to the local GPU and to the peer.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(256) st_v4_kernel(uint4* dst, size_t n16) {
  const uint4 v = make_uint4(threadIdx.x, blockIdx.x, 3, 4);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = v;
}

__global__ void __launch_bounds__(256) ld_v4_kernel(const uint4* src, size_t n16, uint4* sink) {
  uint4 a = make_uint4(0, 0, 0, 0);
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t step = (size_t)gridDim.x * blockDim.x;
  for (; i + 3 * step < n16; i += 4 * step) {
    uint4 v0, v1, v2, v3;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v0.x), "=r"(v0.y), "=r"(v0.z), "=r"(v0.w) : "l"(src + i));
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v1.x), "=r"(v1.y), "=r"(v1.z), "=r"(v1.w) : "l"(src + i + step));
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v2.x), "=r"(v2.y), "=r"(v2.z), "=r"(v2.w) : "l"(src + i + 2 * step));
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v3.x), "=r"(v3.y), "=r"(v3.z), "=r"(v3.w) : "l"(src + i + 3 * step));
    a.x ^= v0.x ^ v1.x ^ v2.x ^ v3.x; a.y ^= v0.y ^ v1.y ^ v2.y ^ v3.y;
  }
  if (a.x == 0x12345678 && a.y == 0x9abcdef0) sink[0] = a;
}

// cp.async.bulk shared::cta -> global, CHUNK contiguous bytes per copy, DEPTH copies in flight / CTA
template <int CHUNK, int DEPTH>
__global__ void __launch_bounds__(128) bulk_kernel(uint8_t* dst, size_t bytes) {
  extern __shared__ __align__(128) uint8_t sm[];
  for (int i = threadIdx.x; i < CHUNK * DEPTH / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const size_t nchunks = bytes / CHUNK;
    int slot = 0;
    for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(DEPTH - 1) : "memory");
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + c * CHUNK),
                   "r"(smem_u32(sm + slot * CHUNK)), "r"(CHUNK) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      slot = (slot + 1) % DEPTH;
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

// cp.async.bulk.tensor 2-D store: [32 rows][64 bf16] boxes (128B swizzle) into [rows][1024] bf16
template <int DEPTH>
__global__ void __launch_bounds__(128) tma2d_kernel(const __grid_constant__ CUtensorMap tm, int rows, int cols) {
  extern __shared__ __align__(128) uint8_t sm[];
  for (int i = threadIdx.x; i < 4096 * DEPTH / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const int bx = cols / 64, by = rows / 32;
    int slot = 0;
    for (int b = blockIdx.x; b < bx * by; b += gridDim.x) {
      // same order as the GEMM epilogue: a CTA walks 4 column boxes of one 32-row band
      const int c0 = (b % bx) * 64, r0 = (b / bx) * 32;
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(DEPTH - 1) : "memory");
      asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(&tm),
                   "r"(smem_u32(sm + slot * 4096)), "r"(c0), "r"(r0) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      slot = (slot + 1) % DEPTH;
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

// flag ping-pong: one-way latency of st.release.sys -> ld.acquire.sys across the link
__global__ void pingpong_kernel(volatile uint32_t* mine, uint32_t* theirs, int iters, int first) {
  for (int i = 1; i <= iters; ++i) {
    if (first) {
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(i) : "memory");
      while (*mine < (uint32_t)i) {}
    } else {
      while (*mine < (uint32_t)i) {}
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(i) : "memory");
    }
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(void* base, int rows, int cols) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    fn = (EncodeTiledFn)p;
  }
  CUtensorMap tm;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {64, 32}, es[2] = {1, 1};
  if (fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
    printf("tensor map encode failed\n");
    exit(1);
  }
  return tm;
}

struct Side {
  int dev;
  uint8_t* buf;
  cudaStream_t st;
  cudaEvent_t e0, e1;
};

template <typename F>
static void run(const char* name, const char* where, Side* s, size_t bytes, int reps, F launch) {
  for (int w = 0; w < 2; ++w)
    for (int g = 0; g < 2; ++g) { CK(cudaSetDevice(s[g].dev)); launch(g); }
  for (int g = 0; g < 2; ++g) { CK(cudaSetDevice(s[g].dev)); CK(cudaStreamSynchronize(s[g].st)); }
  for (int g = 0; g < 2; ++g) { CK(cudaSetDevice(s[g].dev)); CK(cudaEventRecord(s[g].e0, s[g].st)); }
  for (int r = 0; r < reps; ++r)
    for (int g = 0; g < 2; ++g) { CK(cudaSetDevice(s[g].dev)); launch(g); }
  float worst = 0;
  for (int g = 0; g < 2; ++g) {
    CK(cudaSetDevice(s[g].dev));
    CK(cudaEventRecord(s[g].e1, s[g].st));
    CK(cudaEventSynchronize(s[g].e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, s[g].e0, s[g].e1));
    if (ms > worst) worst = ms;
  }
  printf("%-14s %-6s %8.1f us/launch  %8.1f GB/s per direction\n", name, where, worst * 1e3 / reps,
         bytes / (worst * 1e-3 / reps) / 1e9);
  fflush(stdout);
}

int main(int argc, char** argv) {
  const size_t mib = argc > 1 ? atoi(argv[1]) : 64;
  const size_t bytes = mib << 20;
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (ndev < 2) { printf("needs 2 GPUs\n"); return 0; }
  Side s[2];
  for (int g = 0; g < 2; ++g) {
    s[g].dev = g;
    CK(cudaSetDevice(g));
    int can = 0;
    CK(cudaDeviceCanAccessPeer(&can, g, 1 - g));
    if (!can) { printf("no peer access\n"); return 0; }
    CK(cudaDeviceEnablePeerAccess(1 - g, 0));
    CK(cudaMalloc(&s[g].buf, bytes));
    CK(cudaMemset(s[g].buf, 0, bytes));
    CK(cudaStreamCreate(&s[g].st));
    CK(cudaEventCreate(&s[g].e0));
    CK(cudaEventCreate(&s[g].e1));
  }
  const int reps = 20, sms = 148;
  for (int g = 0; g < 2; ++g) {
    CK(cudaSetDevice(g));
    CK(cudaFuncSetAttribute(bulk_kernel<16384, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 4));
  }
  printf("# %zu MiB per launch per direction, both GPUs active, %d launches timed\n", mib, reps);
  for (int peer = 0; peer < 2; ++peer) {
    const char* where = peer ? "peer" : "local";
    auto dst = [&](int g) { return s[peer ? 1 - g : g].buf; };
    for (int ctas : {sms, 2 * sms, 4 * sms}) {
      char nm[64];
      snprintf(nm, sizeof nm, "st.v4 x%d", ctas);
      run(nm, where, s, bytes, reps, [&](int g) { st_v4_kernel<<<ctas, 256, 0, s[g].st>>>((uint4*)dst(g), bytes / 16); });
    }
    run("ld.v4 x592", where, s, bytes, reps, [&](int g) { ld_v4_kernel<<<4 * sms, 256, 0, s[g].st>>>((const uint4*)dst(g), bytes / 16, (uint4*)s[g].buf); });
    run("ld.v4 x1184", where, s, bytes, reps, [&](int g) { ld_v4_kernel<<<8 * sms, 256, 0, s[g].st>>>((const uint4*)dst(g), bytes / 16, (uint4*)s[g].buf); });
    run("bulk-4K d4", where, s, bytes, reps, [&](int g) { bulk_kernel<4096, 4><<<sms, 128, 4096 * 4, s[g].st>>>(dst(g), bytes); });
    run("bulk-4K d8 x2", where, s, bytes, reps, [&](int g) { bulk_kernel<4096, 8><<<2 * sms, 128, 4096 * 8, s[g].st>>>(dst(g), bytes); });
    run("bulk-16K d4", where, s, bytes, reps, [&](int g) { bulk_kernel<16384, 4><<<sms, 128, 16384 * 4, s[g].st>>>(dst(g), bytes); });
    run("bulk-1K d8", where, s, bytes, reps, [&](int g) { bulk_kernel<1024, 8><<<sms, 128, 1024 * 8, s[g].st>>>(dst(g), bytes); });
    const int rows = (int)(bytes / 2048);
    CUtensorMap tm[2];
    for (int g = 0; g < 2; ++g) tm[g] = make_map(dst(g), rows, 1024);
    run("tma-2d d4", where, s, bytes, reps, [&](int g) { tma2d_kernel<4><<<sms, 128, 4096 * 4 + 1024, s[g].st>>>(tm[g], rows, 1024); });
    run("tma-2d d8 x2", where, s, bytes, reps, [&](int g) { tma2d_kernel<8><<<2 * sms, 128, 4096 * 8 + 1024, s[g].st>>>(tm[g], rows, 1024); });
    run("tma-2d d8 x4", where, s, bytes, reps, [&](int g) { tma2d_kernel<8><<<4 * sms, 128, 4096 * 8 + 1024, s[g].st>>>(tm[g], rows, 1024); });
  }
  // one-way flag latency
  {
    uint32_t* f[2];
    for (int g = 0; g < 2; ++g) { CK(cudaSetDevice(g)); CK(cudaMalloc(&f[g], 256)); CK(cudaMemset(f[g], 0, 256)); }
    for (int g = 0; g < 2; ++g) { CK(cudaSetDevice(g)); CK(cudaDeviceSynchronize()); }
    const int iters = 2000;
    CK(cudaSetDevice(0));
    CK(cudaEventRecord(s[0].e0, s[0].st));
    pingpong_kernel<<<1, 1, 0, s[0].st>>>(f[0], f[1], iters, 1);
    CK(cudaEventRecord(s[0].e1, s[0].st));
    CK(cudaSetDevice(1));
    pingpong_kernel<<<1, 1, 0, s[1].st>>>(f[1], f[0], iters, 0);
    CK(cudaSetDevice(0));
    CK(cudaEventSynchronize(s[0].e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, s[0].e0, s[0].e1));
    printf("flag ping-pong: %.2f us round trip, %.2f us one way\n", ms * 1e3 / iters, ms * 1e3 / iters / 2);
  }
  return 0;
}
