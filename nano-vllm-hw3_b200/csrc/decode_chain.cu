// Decode-regime CHAIN of W8A8 linears (M <= 16 tokens) in ONE persistent launch.
//
// The reference runs every decode-step linear as per_row_quant + _per_row_post_process_mm
// (nanovllm/utils/quantization.py:310-346), i.e. two launches per linear and 4 linears per decoder
// layer (nanovllm/models/qwen3.py:80,90,116,118).  At M <= 16 a linear moves N*K weight bytes and
// almost nothing else: the regime is HBM-bound, and with one launch (or two) per linear the fixed
// launch / barrier-init / drain cost is 5-10x the time the bytes need.  This kernel walks a whole
// chain of linears -- op i consumes the first K_i columns of op i-1's output, as the decoder
// stack's qkv -> o -> gate_up -> down order does in bench.py's LinearStack.
//
// Measured on B200 (tools/chain_probe.py, profiles/r2f_*): what bounds a chained op is not the
// weight stream but (a) the work every CTA repeats -- reading and quantising the op's whole
// activation (16 x K values: 9 ALU instructions each, ~5500 cycles at K = 4096 when every CTA does
// all of it) -- and (b) the global round trips between dependent ops (release ~1000 cycles, grid
// barrier ~1500, L2 read ~600).  Hence:
//
//   * thread-block CLUSTERS of 4 CTAs split K: rank r of a cluster reads and quantises only ITS
//     K/4 slice of the activation (Triton quantiser arithmetic :26-41, bit for bit) and streams
//     only that slice of its weight rows; the four partial accumulator tiles meet through
//     DISTRIBUTED SHARED MEMORY (reduce-scatter: 8-row group i goes to rank i % 4 by st.async, the
//     bytes counted on the receiver's mbarrier -- no fences; summed in rank order -- int32: exact,
//     fp32: fixed order), no L2 round trip, no second grid barrier;
//   * rows are dealt to the clusters in groups of 8 (the mma.sync n = 8), so every SM gets the
//     same bytes +- one group for any N; weights arrive by TMA (one 8-row x <= 1 KB box per
//     instruction -- whole DRAM pages per row: with 512-byte rows the stream reached 5 TB/s, with
//     1 KB rows the HBM peak --, two TMA warps) through a 17-slot ring that runs ahead ACROSS op boundaries (weights depend on
//     nothing): HBM keeps streaming while an op's epilogue and the grid barrier are in flight;
//   * each consumer warp owns whole row groups (its own ring slots, its own accumulators in
//     registers: m16n8k32 warp MMAs, int8 -> int32 exact / e4m3 -> fp32), activations as
//     register-resident A fragments loaded once per K chunk: no block-wide sync in the stream;
//   * the finishing rank applies (acc * sx[m]) * sw[n] (+ bias[n]) -- reference product order
//     :111-121 --, stores bf16 and folds |y| into the NEXT op's per-token amax (atomicMax on the
//     float bit pattern), so the next op's quantiser needs no pass over the row;
//   * ONE grid barrier per op (red.release.gpu by one thread per CTA + ld.acquire.gpu spin);
//     static data is fetched ahead of its use: op descriptors two ops ahead into shared memory,
//     the TMA warps read the next op's plan and prefetch its tensor map while issuing this op's
//     boxes, the scales / bias of a pass are loaded before its weight stream and staged in shared
//     memory.
//
// Results are bit-identical to ql_linear op by op for INT8 per-row (int32 accumulation is exact in
// any order); FP8 per-row differs from the tcgen05 kernels only in fp32 summation order.
//
// OUTCOME: correct, and slower than one ql_linear per op on every Qwen3
// stack measured (per op: 0.6B 5.8 us with the tcgen05 consumer below / 7.2 us with this mma.sync
// one vs 5.4 us; 8B 13.4 / 15.9 vs 12 us): the dependent chain release ->
// grid barrier -> activation read -> quantise -> first MMAs -> exchange -> epilogue costs ~10 000
// cycles per op, more than a kernel boundary with programmatic dependent launch.  bench.py times
// it next to the per-linear launches; nothing else in the engine depends on it.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "kernels.h"

namespace ql {

constexpr int kDcC = 4;                        // CTAs per cluster
constexpr int kDcConsumers = 256;              // 8 MMA warps
constexpr int kDcProducers = 2;                // TMA warps (alternate ring slots)
constexpr int kDcThreads = kDcConsumers + 32 * kDcProducers;
constexpr int kDcJ = 8;                        // row groups per warp per pass
constexpr int kDcPassRG = 8 * kDcJ;            // row groups per CTA per pass
constexpr int kDcMaxBlk = 48;                  // 64-byte K blocks per rank slice (3072 B: K <= 12288)
constexpr int kDcChunkBlk = 8;                 // 64-byte K blocks whose A fragments a warp holds at a time
constexpr int kDcSlotBlk = 16;                 // 64-byte K blocks per ring slot row (<= 1 KB: whole DRAM pages per row)
constexpr int kDcSlotBytes = 8 * kDcSlotBlk * 64;    // 8 KB: 8 weight rows x <= 1 KB
constexpr int kDcSlots = 17;
constexpr int kDcABytes = kDcMaxBlk * 1024;    // quantised activation slice in mma-fragment order
constexpr int kDcRecvBytes = kDcPassRG * 512;  // partial tiles received from the cluster
constexpr int kDcSmall = 2048 + 2 * kDcPassRG * 8 * 4;   // barriers, descriptors + staged scales / bias of a pass
constexpr int kDcSmem = kDcSlots * kDcSlotBytes + kDcABytes + kDcRecvBytes + kDcSmall + 128;
constexpr uint32_t kDcMagic = 0x51444333u;     // "QDC3"
static_assert(kDcSmem <= 227 * 1024, "decode chain shared memory");

struct ChainOp {            // device-side descriptor, 128 bytes
  const uint8_t* qw;        // [N, K] codes (streamed through the op's tensor map)
  const float* sw;          // [N]
  const uint16_t* bias;     // bf16 [N] or nullptr
  const uint16_t* x;        // bf16 [M, ldx]
  uint16_t* y;              // bf16 [M, ldy]
  int64_t ldx, ldy;
  int N, K;
  int kse, nblk;            // K split inside the cluster (2 or 4); 64-byte blocks per rank slice
  int cb, nchunk;           // blocks per chunk, chunks per slice
  int RG, rot;              // groups of 8 weight rows; cluster rotation
  int amax_cols;            // leading columns of y folded into the next op's amax (0: none)
  int x_amax;               // 1: amax[op] was produced by the previous op's epilogue
  uint32_t pad[8];
};
static_assert(sizeof(ChainOp) == 128, "ChainOp layout");

struct ChainParams {
  const ChainOp* ops;
  const CUtensorMap* tmaps; // one per op: uint32 [N, K/4], box {cb*16, 8}, no swizzle
  int n_ops, M, NC;         // NC clusters
  uint32_t* sync;           // [0] grid barrier counter, [1] exit counter
  uint32_t* amax;           // [n_ops][16] float bit patterns
  long long* trace;         // debug (ql_debug_set_trace): clock64 stamps [cta][16 ops][16], else nullptr
};
#define DC_TRACE(i)                                                                          \
  do {                                                                                       \
    if (p.trace != nullptr && tid == 0 && l < 16)                                            \
      p.trace[((size_t)cta * 16 + l) * 16 + (i)] = clock64();                                \
  } while (0)

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

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// 8-byte store into a peer's shared memory through the async proxy: the bytes are counted on the
// peer's mbarrier (complete_tx), so the receiver needs no fence -- it waits for the byte count.
__device__ __forceinline__ void st_async_v2(uint32_t cluster_addr, uint32_t a, uint32_t b, uint32_t cluster_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];" ::"r"(
                   cluster_addr),
               "r"(a), "r"(b), "r"(cluster_bar)
               : "memory");
}
// "my reads of the buffer are done" (all of them completed before the bar.sync that precedes this)
__device__ __forceinline__ void mbar_arrive_relaxed_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

// What CTA (cluster q, rank r) does in an op: K slice ksl of row groups [rg0, rg1); nrg_max = the
// largest row-group count among the cluster's sub-blocks (the same for all 4 ranks: it fixes the
// number of passes and therefore of DSMEM exchanges every rank of the cluster takes part in).
struct DcUnit {
  int ksl, sbi, rg0, rg1, nrg_max;
};
__device__ __forceinline__ DcUnit dc_unit(int cluster, int rank, int NC, int rot, int kse, int RG) {
  DcUnit u;
  int q = cluster + rot;
  if (q >= NC) q -= NC;
  const int nsub = kDcC / kse, NSB = NC * nsub;
  u.sbi = rank / kse;
  u.ksl = rank - u.sbi * kse;
  u.nrg_max = 0;
  u.rg0 = u.rg1 = 0;
  for (int s = 0; s < nsub; ++s) {
    const int sb = q * nsub + s;
    // sb <= 148 and RG <= 2^22 (checked by the planner): the products fit 32 bits
    const int r0 = (int)((uint32_t)sb * (uint32_t)RG / (uint32_t)NSB),
              r1 = (int)((uint32_t)(sb + 1) * (uint32_t)RG / (uint32_t)NSB);
    u.nrg_max = r1 - r0 > u.nrg_max ? r1 - r0 : u.nrg_max;
    if (s == u.sbi) { u.rg0 = r0; u.rg1 = r1; }
  }
  return u;
}

template <int FMT>
__global__ void __launch_bounds__(kDcThreads, 1) decode_chain_kernel(const ChainParams p) {
  using acc_t = typename std::conditional<FMT == QL_INT8, int, float>::type;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  uint8_t* ring = smem;
  uint4* A_s = reinterpret_cast<uint4*>(ring + kDcSlots * kDcSlotBytes);          // [blk][hi][g][t]
  uint32_t* recv = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(A_s) + kDcABytes);  // [src][tile][128]
  uint8_t* small = reinterpret_cast<uint8_t*>(recv) + kDcRecvBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(small);                       // [kDcSlots]
  uint64_t* empty_bar = full_bar + kDcSlots;                                     // [kDcSlots]
  uint64_t* recv_bar = empty_bar + kDcSlots;                                     // partial tiles of a pass have landed
  uint64_t* free_bar = recv_bar + 1;                                             // every rank has read its tiles
  float* sx_s = reinterpret_cast<float*>(small + 640);     // [16]
  int* amax_s = reinterpret_cast<int*>(small + 704);       // [16]
  uint32_t* flag_s = reinterpret_cast<uint32_t*>(small + 768);
  ChainOp* desc_s = reinterpret_cast<ChainOp*>(small + 1024);   // [3]: descriptors of ops l, l+1, l+2
  float* sw_s = reinterpret_cast<float*>(small + 2048);         // [512] weight scales of this pass's rows
  float* b_s = sw_s + kDcPassRG * 8;                            // [512] bias of this pass's rows

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = gridDim.x, cta = blockIdx.x;
  const int rank = (int)cluster_ctarank(), cluster = cta / kDcC;
  constexpr float qmax = Fmt<FMT>::qmax;

  if (tid == 0) {
    for (int i = 0; i < kDcSlots; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(recv_bar, 1);
    mbar_init(free_bar, kDcC);
    mbar_fence_init();
  }
  if (tid < 16 && (tid >> 3) < p.n_ops)
    reinterpret_cast<uint4*>(desc_s)[tid] = __ldg(reinterpret_cast<const uint4*>(p.ops) + tid);
  __syncthreads();
  cluster_arrive();          // every rank's barriers exist before any remote arrive / DSMEM store
  cluster_wait();

  if (warp >= 8) {
    // ===================== producers: every op's weights, as far ahead as the ring allows =========
    // Two TMA warps take alternate ring slots; the next op's plan fields and tensor map are fetched
    // while this op's boxes are being issued (both are static data that would miss to DRAM).
    if (lane == 0) {
      const uint32_t pw = (uint32_t)(warp - 8);
      uint32_t seq = 0;
      int kse = __ldg(&p.ops[0].kse), nblk = __ldg(&p.ops[0].nblk), cb = __ldg(&p.ops[0].cb),
          nchunk = __ldg(&p.ops[0].nchunk), rot = __ldg(&p.ops[0].rot), RG = __ldg(&p.ops[0].RG);
      tma_prefetch_desc(p.tmaps);
      for (int l = 0; l < p.n_ops; ++l) {
        const DcUnit u = dc_unit(cluster, rank, p.NC, rot, kse, RG);
        const CUtensorMap* tm = p.tmaps + l;
        const uint32_t bytes = 8u * (uint32_t)cb * 64u;
        const int ksl_blk = u.ksl * nblk, cbc = cb, nch = nchunk;
        if (l + 1 < p.n_ops) {
          const ChainOp& nx = p.ops[l + 1];
          kse = __ldg(&nx.kse); nblk = __ldg(&nx.nblk); cb = __ldg(&nx.cb);
          nchunk = __ldg(&nx.nchunk); rot = __ldg(&nx.rot); RG = __ldg(&nx.RG);
          tma_prefetch_desc(p.tmaps + l + 1);
        }
        for (int r0p = u.rg0; r0p < u.rg1; r0p += kDcPassRG) {
          const int n = u.rg1 - r0p < kDcPassRG ? u.rg1 - r0p : kDcPassRG;
          // Blocks of 8 row groups (one per consumer warp), chunk-major inside a block: the chunks of
          // a row group follow each other within 8 slots (32 KB), so its row segments are fetched
          // close together (DRAM page reuse), and a warp's consecutive slots are 8 apart -- never
          // more than one ring lap ahead of the producer (mbarrier parity waits tolerate only one).
          for (int i0 = 0; i0 < n; i0 += 8) {
            const int m = n - i0 < 8 ? n - i0 : 8;
            for (int c = 0; c < nch; ++c) {
              const int c0 = (ksl_blk + c * cbc) * 16;   // uint32 elements along K
              for (int w = 0; w < m; ++w, ++seq) {
                if ((seq & 1u) != pw) continue;
                const uint32_t slot = seq % kDcSlots, par = (seq / kDcSlots) & 1u;
                mbar_wait(&empty_bar[slot], par ^ 1u);
                mbar_expect_tx(&full_bar[slot], bytes);
                tma_load_2d(ring + slot * kDcSlotBytes, tm, &full_bar[slot], c0, (r0p + i0 + w) * 8, kEvictFirst);
              }
            }
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== consumers =====================
    const int g = lane >> 2, t = lane & 3;                 // mma fragment coordinates
    uint32_t seq_base = 0;                                  // ring sequence number of this pass's first slot
    uint32_t n_exch = 0;                                    // DSMEM exchanges this cluster has done
    const uint32_t recv_local = smem_u32(recv), recv_bar_local = smem_u32(recv_bar),
                   free_bar_local = smem_u32(free_bar);

    for (int l = 0; l < p.n_ops; ++l) {
      const ChainOp& op = desc_s[l % 3];
      const int kse = op.kse, nblk = op.nblk, cb = op.cb, nchunk = op.nchunk;
      const DcUnit u = dc_unit(cluster, rank, p.NC, op.rot, kse, op.RG);
      DC_TRACE(0);
      // descriptor of op l+2: in flight during this whole op, parked in shared memory at its end
      uint4 dpre = make_uint4(0, 0, 0, 0);
      if (tid >= 32 && tid < 40 && l + 2 < p.n_ops)
        dpre = __ldg(reinterpret_cast<const uint4*>(p.ops + l + 2) + (tid - 32));

      if (u.nrg_max > 0) {
        const int64_t ldx = op.ldx, ldy = op.ldy;
        const float* sw = op.sw;
        const uint16_t* bias = op.bias;
        uint16_t* y = op.y;
        const int amax_cols = op.amax_cols;
        const int tps = kDcPassRG / kse;                   // tile slots per source rank in recv
        const bool have_rows = u.rg1 > u.rg0;

        // ---- this rank's K slice of the activation: read (L2), quantise, park as A fragments ----
        // item = 16 consecutive K values of one token = one uint4 of codes at A_s[item]
        // (item = ((blk * 2 + token / 8) * 8 + token % 8) * 4 + t: a warp's fragment load is 512 B contiguous)
        const int nitems = have_rows ? nblk * 64 : 0;
        const uint16_t* xk = op.x + (size_t)u.ksl * nblk * 64;
        uint4 xv[4][2];
        auto load_items = [&](int base) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int it = base + i * kDcConsumers + tid;
            const int m = ((it >> 5) & 1) * 8 + ((it >> 2) & 7);
            xv[i][0] = xv[i][1] = make_uint4(0, 0, 0, 0);
            if (it < nitems && m < p.M) {
              const uint4* s = reinterpret_cast<const uint4*>(xk + (size_t)m * ldx + (it >> 6) * 64 + (it & 3) * 16);
              xv[i][0] = __ldcg(s);
              xv[i][1] = __ldcg(s + 1);
            }
          }
        };
        load_items(0);
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
        if (tid < 16) sx_s[tid] = __fdiv_rn(fmaxf(__int_as_float(amax_s[tid]), 1e-8f), qmax);
        for (int base = 0; base < nitems; base += 4 * kDcConsumers) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int it = base + i * kDcConsumers + tid;
            if (it < nitems) {
              const int m = ((it >> 5) & 1) * 8 + ((it >> 2) & 7);
              const float inv = __fdiv_rn(qmax, fmaxf(__int_as_float(amax_s[m]), 1e-8f));
              const uint2 q0 = quant8_mul<FMT>(xv[i][0], inv), q1 = quant8_mul<FMT>(xv[i][1], inv);
              A_s[it] = make_uint4(q0.x, q0.y, q1.x, q1.y);
            }
          }
          if (base + 4 * kDcConsumers < nitems) load_items(base + 4 * kDcConsumers);
        }
        consumer_sync();
        if (tid < 16) amax_s[tid] = 0;        // from here on: running max of |y| for the NEXT op (finalize)
        DC_TRACE(2);

        // ---- passes of <= 64 row groups: stream, exchange, finish -------------------------------
        int pass = 0;
        for (int r0c = 0; r0c < u.nrg_max; r0c += kDcPassRG, ++pass) {
          const int r0p = u.rg0 + r0c;
          int n = u.rg1 - r0p;
          n = n < 0 ? 0 : (n > kDcPassRG ? kDcPassRG : n);
          const int nj = n > warp ? (n - warp + 7) >> 3 : 0;
          // this pass expects kse partial tiles for every row group it finishes
          const int my_tiles = n > u.ksl ? (n - u.ksl + kse - 1) / kse : 0;
          if (tid == 0) mbar_expect_tx(recv_bar, (uint32_t)(my_tiles * kse * 512));
          // scales / bias of the pass's rows (static data, DRAM latency): in flight during the stream
          float swr[2] = {0.0f, 0.0f}, br[2] = {0.0f, 0.0f};
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int r = tid + i * kDcConsumers;
            if (r < n * 8) {
              swr[i] = __ldg(sw + r0p * 8 + r);
              if (bias != nullptr) br[i] = bf16_lo(__ldg(bias + r0p * 8 + r));
            }
          }
          acc_t acc[kDcJ][4];
#pragma unroll
          for (int j = 0; j < kDcJ; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0;
#pragma unroll
          for (int j = 0; j < kDcJ; ++j) {
            if (j < nj) {
              for (int c = 0; c < nchunk; ++c) {
                const int mj = n - 8 * j < 8 ? n - 8 * j : 8;      // row groups in this block of 8
                const uint32_t s = seq_base + (uint32_t)(8 * j * nchunk + c * mj + warp);
                const uint32_t slot = s % kDcSlots, par = (s / kDcSlots) & 1u;
                const uint8_t* wrow = ring + slot * kDcSlotBytes + g * (cb * 64) + t * 16;
                // a slot row holds cb <= 16 blocks; the warp keeps the A fragments of 8 at a time
                for (int h0 = 0; h0 < cb; h0 += kDcChunkBlk) {
                  const int hb = cb - h0 < kDcChunkBlk ? cb - h0 : kDcChunkBlk;
                  uint4 a_lo[kDcChunkBlk], a_hi[kDcChunkBlk];
#pragma unroll
                  for (int b = 0; b < kDcChunkBlk; ++b) {
                    if (b < hb) {
                      a_lo[b] = A_s[((c * cb + h0 + b) * 2 + 0) * 32 + lane];
                      a_hi[b] = A_s[((c * cb + h0 + b) * 2 + 1) * 32 + lane];
                    }
                  }
                  if (h0 == 0) {
                    mbar_wait(&full_bar[slot], par);
                    if (j == 0 && c == 0 && pass == 0) DC_TRACE(3);
                  }
#pragma unroll
                  for (int b = 0; b < kDcChunkBlk; ++b) {
                    if (b < hb) {
                      const uint4 w = *reinterpret_cast<const uint4*>(wrow + (h0 + b) * 64);
                      if constexpr (FMT == QL_INT8) {
                        mma_i8_sync(acc[j], a_lo[b].x, a_hi[b].x, a_lo[b].y, a_hi[b].y, w.x, w.y);
                        mma_i8_sync(acc[j], a_lo[b].z, a_hi[b].z, a_lo[b].w, a_hi[b].w, w.z, w.w);
                      } else {
                        mma_f8_sync(acc[j], a_lo[b].x, a_hi[b].x, a_lo[b].y, a_hi[b].y, w.x, w.y);
                        mma_f8_sync(acc[j], a_lo[b].z, a_hi[b].z, a_lo[b].w, a_hi[b].w, w.z, w.w);
                      }
                    }
                  }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty_bar[slot]);
              }
            }
          }
          seq_base += (uint32_t)(nchunk * n);
          if (pass == 0) DC_TRACE(4);

          // ---- reduce-scatter of the partial tiles over DSMEM: row group idx -> rank idx % kse ------
          // (st.async: the bytes are counted on the receiver's mbarrier, no fences on either side)
          if (n_exch > 0) mbar_wait(free_bar, (n_exch - 1) & 1u);   // receivers are done with the last pass
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            sw_s[tid + i * kDcConsumers] = swr[i];
            b_s[tid + i * kDcConsumers] = br[i];
          }
#pragma unroll
          for (int j = 0; j < kDcJ; ++j) {
            if (j < nj) {
              const int idx = warp + 8 * j;
              const int d = idx % kse, ts = idx / kse;
              const uint32_t peer = (uint32_t)(u.sbi * kse + d);
              const uint32_t dst = mapa_u32(recv_local + (uint32_t)((u.ksl * tps + ts) * 512), peer);
              const uint32_t bar = mapa_u32(recv_bar_local, peer);
              uint32_t cu[4];
              if constexpr (FMT == QL_INT8) {
                cu[0] = acc[j][0]; cu[1] = acc[j][1]; cu[2] = acc[j][2]; cu[3] = acc[j][3];
              } else {
                cu[0] = __float_as_uint(acc[j][0]); cu[1] = __float_as_uint(acc[j][1]);
                cu[2] = __float_as_uint(acc[j][2]); cu[3] = __float_as_uint(acc[j][3]);
              }
              st_async_v2(dst + (uint32_t)(g * 8 + 2 * t) * 4u, cu[0], cu[1], bar);
              st_async_v2(dst + (uint32_t)((g + 8) * 8 + 2 * t) * 4u, cu[2], cu[3], bar);
            }
          }
          consumer_sync();                                  // sw_s / b_s are staged
          mbar_wait(recv_bar, n_exch & 1u);
          if (pass == 0) DC_TRACE(5);

          // ---- finish the tiles this rank owns (idx = ts * kse + ksl): one (tile, token) per thread =
          // 8 consecutive features: 2 x LDS.128 per source rank, one 16-byte store
          for (int it = tid; it < my_tiles * 16; it += kDcConsumers) {
            const int ts = it >> 4, m = it & 15;
            const int idx = ts * kse + u.ksl;
            const uint4* rs = reinterpret_cast<const uint4*>(recv + ts * 128 + m * 8);
            uint4 s0 = rs[0], s1 = rs[1];
            for (int s = 1; s < kse; ++s) {
              const uint4 o0 = rs[s * tps * 32], o1 = rs[s * tps * 32 + 1];
              if constexpr (FMT == QL_INT8) {
                s0.x += o0.x; s0.y += o0.y; s0.z += o0.z; s0.w += o0.w;
                s1.x += o1.x; s1.y += o1.y; s1.z += o1.z; s1.w += o1.w;
              } else {
                s0.x = __float_as_uint(__fadd_rn(__uint_as_float(s0.x), __uint_as_float(o0.x)));
                s0.y = __float_as_uint(__fadd_rn(__uint_as_float(s0.y), __uint_as_float(o0.y)));
                s0.z = __float_as_uint(__fadd_rn(__uint_as_float(s0.z), __uint_as_float(o0.z)));
                s0.w = __float_as_uint(__fadd_rn(__uint_as_float(s0.w), __uint_as_float(o0.w)));
                s1.x = __float_as_uint(__fadd_rn(__uint_as_float(s1.x), __uint_as_float(o1.x)));
                s1.y = __float_as_uint(__fadd_rn(__uint_as_float(s1.y), __uint_as_float(o1.y)));
                s1.z = __float_as_uint(__fadd_rn(__uint_as_float(s1.z), __uint_as_float(o1.z)));
                s1.w = __float_as_uint(__fadd_rn(__uint_as_float(s1.w), __uint_as_float(o1.w)));
              }
            }
            if (m < p.M) {
              const uint32_t raw[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
              const float sxm = sx_s[m];
              const float4 w0 = *reinterpret_cast<const float4*>(sw_s + idx * 8),
                           w1 = *reinterpret_cast<const float4*>(sw_s + idx * 8 + 4);
              const float4 c0 = *reinterpret_cast<const float4*>(b_s + idx * 8),
                           c1 = *reinterpret_cast<const float4*>(b_s + idx * 8 + 4);
              const float swv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
              const float bv[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
              const int nn = (r0p + idx) * 8;
              uint16_t o16[8];
              float mx = 0.0f;
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float a = FMT == QL_INT8 ? __int2float_rn((int)raw[e]) : __uint_as_float(raw[e]);
                float v = __fmul_rn(__fmul_rn(a, sxm), swv[e]);            // (acc * sa) * sb, reference :121
                if (bias != nullptr) v = __fadd_rn(v, bv[e]);
                o16[e] = f2bf(v);
                if (nn + e < amax_cols) mx = fmaxf(mx, fabsf(bf16_lo(o16[e])));
              }
              uint4 pk;
              pk.x = o16[0] | ((uint32_t)o16[1] << 16);
              pk.y = o16[2] | ((uint32_t)o16[3] << 16);
              pk.z = o16[4] | ((uint32_t)o16[5] << 16);
              pk.w = o16[6] | ((uint32_t)o16[7] << 16);
              *reinterpret_cast<uint4*>(y + (size_t)m * ldy + nn) = pk;
              if (mx > 0.0f) atomicMax(amax_s + m, __float_as_int(mx));   // shared: this op's input amax is consumed
            }
          }
          consumer_sync();
          if (tid < kDcC) mbar_arrive_relaxed_cluster(mapa_u32(free_bar_local, (uint32_t)tid));
          ++n_exch;
          if (pass == 0) DC_TRACE(6);
        }

        // ---- fold |y| into the next op's per-token amax (collected per CTA in amax_s) ---------------
        if (amax_cols > 0 && tid < 16 && tid < p.M && amax_s[tid] > 0)
          atomicMax(reinterpret_cast<int*>(p.amax) + (l + 1) * 16 + tid, amax_s[tid]);
        DC_TRACE(7);
      }

      // ---- grid barrier: op l is complete everywhere (its output and the next amax are visible) ----
      if (l + 1 < p.n_ops) {
        if (tid >= 32 && tid < 40 && l + 2 < p.n_ops)
          reinterpret_cast<uint4*>(desc_s + (l + 2) % 3)[tid - 32] = dpre;
        consumer_sync();             // orders every thread's stores / atomics before thread 0's release
        DC_TRACE(8);
        if (tid == 0) {
          red_release_gpu_add(p.sync, 1u);
          DC_TRACE(9);
          const uint32_t target = (uint32_t)(l + 1) * (uint32_t)G;
          if (ld_acquire_gpu(p.sync) < target) {
            const uint64_t t0 = globaltimer_ns();
            while (ld_acquire_gpu(p.sync) < target)
              if (globaltimer_ns() - t0 > 4000000000ull) __trap();   // 4 s: never hang the GPU
          }
          DC_TRACE(10);
        }
        consumer_sync();
      }
    }

    // ---- the last CTA out re-arms the counters for the next launch ------------------------------
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
  // no rank leaves while a peer may still store into / arrive on its shared memory
  cluster_arrive();
  cluster_wait();
}

// ------------------------------------------------------------------ tcgen05 form of the chain
//
// Same chain, same protocol (clusters of 4 split K, amax carried by the epilogue, one grid barrier
// per op), but the consumer is the tensor core of gemm_skinny.cu: UMMA A = a 128-row weight tile
// streamed by TMA in 128-byte K blocks (consecutive ring stages = consecutive K blocks of the same
// rows: whole DRAM pages), UMMA B = this rank's K/4 slice of the 16 quantised tokens, written by
// the quantiser threads straight into the 128B-swizzled K-major operand layout, accumulator
// D[128 x 16] in TMEM (two buffers: the epilogue of tile i overlaps the MMAs of tile i+1).
// Nothing of the weight stream passes through registers, so the consumer runs far above the HBM
// rate and what the ring prefetched during the previous op's epilogue and the grid barrier is
// consumed at once.  Split-K reduce-scatter over DSMEM: token columns 4r..4r+3 of every rank's
// accumulator go to rank r (st.async, bytes counted on the receiver's mbarrier), which adds the
// four partials in rank order, scales, stores bf16 and folds |y| into the next op's amax.
// Requires K % 512 == 0 and K <= 12288 for every op (else the mma.sync kernel above is used).

constexpr int kTcThreads = 384;                     // producer, MMA, 2 spare, 8 quantiser warps (4 of them the epilogue)
constexpr int kTcAcc = 4;                           // accumulators per tile: the 4 MMAs of a K block go to 4 TMEM
                                                    // tiles (a dependent M128 x N16 MMA costs ~140 cycles)
constexpr int kTcStages = 10;
constexpr int kTcStage = 128 * 128;                 // 16 KB: 128 weight rows x 128 B of K
constexpr int kTcBres = 24 * 16 * 128;              // 48 KB: 16 tokens x 3072 B of K (K <= 12288)
constexpr int kTcRecv = 4 * 128 * 16;               // 8 KB: [src rank][row][4 token columns] words
constexpr int kTcBGroups = 6;                       // B-operand readiness is signalled per 4 K blocks (24 / 4)
constexpr int kTcSmall = 1024;
constexpr int kTcSmem = kTcStages * kTcStage + kTcBres + kTcRecv + kTcSmall + 1024;
static_assert(kTcSmem <= 227 * 1024, "tc chain shared memory");

__device__ __forceinline__ void st_async_v4(uint32_t cluster_addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                            uint32_t cluster_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(
                   cluster_addr),
               "r"(a), "r"(b), "r"(c), "r"(d), "r"(cluster_bar)
               : "memory");
}
__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 2, 128;" ::: "memory"); }
__device__ __forceinline__ void quant_sync() { asm volatile("bar.sync 3, 256;" ::: "memory"); }

// tiles of 128 weight rows dealt to the clusters round robin, rotated per op
__device__ __forceinline__ int tc_first_tile(int cluster, int NC, int rot) {
  int q = cluster - rot;
  if (q < 0) q += NC;
  return q;
}

#define TC_TRACE(i)                                                                          \
  do {                                                                                       \
    if (p.trace != nullptr && e == 0 && l < 16)                                              \
      p.trace[((size_t)cta * 16 + l) * 16 + (i)] = clock64();                                \
  } while (0)

template <int FMT>
__global__ void __launch_bounds__(kTcThreads, 1) decode_chain_tc_kernel(const ChainParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* ring = smem;
  uint8_t* bres = ring + kTcStages * kTcStage;
  uint32_t* recv = reinterpret_cast<uint32_t*>(bres + kTcBres);
  uint8_t* small = reinterpret_cast<uint8_t*>(recv) + kTcRecv;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(small);      // [10]
  uint64_t* empty_bar = full_bar + kTcStages;                   // [10]
  uint64_t* tfull_bar = empty_bar + kTcStages;                  // [2]
  uint64_t* tempty_bar = tfull_bar + 2;                         // [2]
  uint64_t* bready_bar = tempty_bar + 2;                        // [6]: groups of 4 K blocks of the B operand
  uint64_t* recv_bar = bready_bar + kTcBGroups;
  uint64_t* free_bar = recv_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(free_bar + 1);
  float* sx_s = reinterpret_cast<float*>(small + 320);          // [16]
  int* amax_s = reinterpret_cast<int*>(small + 384);            // [16]
  uint32_t* flag_s = reinterpret_cast<uint32_t*>(small + 448);
  ChainOp* desc_s = reinterpret_cast<ChainOp*>(small + 512);    // [3]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = gridDim.x, cta = blockIdx.x;
  const int rank = (int)cluster_ctarank(), cluster = cta / kDcC;
  constexpr float qmax = Fmt<FMT>::qmax;

  if (tid == 0) {
    for (int i = 0; i < kTcStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], 4);
    }
    for (int i = 0; i < kTcBGroups; ++i) mbar_init(&bready_bar[i], 8);
    mbar_init(recv_bar, 1);
    mbar_init(free_bar, kDcC);
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 2 * kTcAcc * 16);
    tmem_relinquish();
  }
  if (tid >= 64 && tid < 80 && ((tid - 64) >> 3) < p.n_ops)
    reinterpret_cast<uint4*>(desc_s)[tid - 64] = __ldg(reinterpret_cast<const uint4*>(p.ops) + (tid - 64));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  cluster_arrive();
  cluster_wait();

  if (warp == 0) {
    // ===================== TMA producer: every op's weight tiles, as far ahead as the ring allows ==
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int N = __ldg(&p.ops[0].N), K = __ldg(&p.ops[0].K), rot = __ldg(&p.ops[0].rot);
      tma_prefetch_desc(p.tmaps);
      for (int l = 0; l < p.n_ops; ++l) {
        const int tiles = (N + 127) >> 7, nkb = K >> 9;          // K / 4 ranks / 128 B
        const int t0 = tc_first_tile(cluster, p.NC, rot);
        const CUtensorMap* tm = p.tmaps + l;
        if (l + 1 < p.n_ops) {
          N = __ldg(&p.ops[l + 1].N); K = __ldg(&p.ops[l + 1].K); rot = __ldg(&p.ops[l + 1].rot);
          tma_prefetch_desc(p.tmaps + l + 1);
        }
        for (int tile = t0; tile < tiles; tile += p.NC) {
          for (int kb = 0; kb < nkb; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_expect_tx(&full_bar[stage], kTcStage);
            tma_load_2d(ring + stage * kTcStage, tm, &full_bar[stage], (rank * nkb + kb) * 128, tile * 128, kEvictFirst);
            if (p.trace != nullptr && l < 16 && tile == t0 && (kb == 0 || kb == nkb - 1))
              p.trace[((size_t)cta * 16 + l) * 16 + (kb == 0 ? 14 : 15)] = clock64();
            if (++stage == kTcStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer ==========================================================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc<FMT>(128, 16);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0, b_phases = 0;   // b_phases: one parity bit per B-operand group
      int N = __ldg(&p.ops[0].N), K = __ldg(&p.ops[0].K), rot = __ldg(&p.ops[0].rot);
      for (int l = 0; l < p.n_ops; ++l) {
        const int tiles = (N + 127) >> 7, nkb = K >> 9;
        const int t0 = tc_first_tile(cluster, p.NC, rot);
        if (l + 1 < p.n_ops) {
          N = __ldg(&p.ops[l + 1].N); K = __ldg(&p.ops[l + 1].K); rot = __ldg(&p.ops[l + 1].rot);
        }
        if (t0 >= tiles) continue;
        // (the B operand becomes ready group by group: see the quantiser)
        for (int tile = t0; tile < tiles; tile += p.NC) {
          mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * (kTcAcc * 16);
          for (int kb = 0; kb < nkb; ++kb) {
            if (tile == t0 && (kb & 3) == 0) {
              // K blocks kb..kb+3 of this op's activations are quantised into bres (the quantiser
              // works through the slice in K order while the first blocks are already multiplied)
              const int gi = kb >> 2;
              mbar_wait(&bready_bar[gi], (b_phases >> gi) & 1u);
              b_phases ^= 1u << gi;
              tc_fence_after();
              if (kb == 0 && p.trace != nullptr && l < 16) p.trace[((size_t)cta * 16 + l) * 16 + 11] = clock64();
            }
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (p.trace != nullptr && l < 16 && tile == t0 && (kb == 0 || kb == nkb - 1))
              p.trace[((size_t)cta * 16 + l) * 16 + (kb == 0 ? 12 : 13)] = clock64();
            const uint64_t a_desc = make_sw128_kmajor_desc(smem_u32(ring + stage * kTcStage));
            const uint64_t b_desc = make_sw128_kmajor_desc(smem_u32(bres) + kb * (16 * 128));
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if constexpr (FMT == QL_INT8) mma_i8(d_tmem + k * 16, a_desc + 2 * k, b_desc + 2 * k, idesc, kb > 0);
              else mma_f8(d_tmem + k * 16, a_desc + 2 * k, b_desc + 2 * k, idesc, kb > 0);
            }
            tc_commit(&empty_bar[stage]);
            if (++stage == kTcStages) { stage = 0; phase ^= 1; }
          }
          tc_commit(&tfull_bar[acc]);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== quantiser (256 threads: warps 4-11) + epilogue (warps 4-7) ============
    const int e = tid - 128;                       // 0..255
    const int q = (warp - 4) & 3;                  // TMEM lane quarter of an epilogue warp (= warp % 4)
    const bool epi = warp < 8;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    const uint32_t recv_local = smem_u32(recv), recv_bar_local = smem_u32(recv_bar),
                   free_bar_local = smem_u32(free_bar);
    int acc = 0;
    uint32_t acc_phase = 0, n_exch = 0;

    for (int l = 0; l < p.n_ops; ++l) {
      const ChainOp& op = desc_s[l % 3];
      const int N = op.N, K = op.K;
      const int tiles = (N + 127) >> 7, nkb = K >> 9;
      const int t0 = tc_first_tile(cluster, p.NC, op.rot);
      uint4 dpre = make_uint4(0, 0, 0, 0);
      if (e < 8 && l + 2 < p.n_ops) dpre = __ldg(reinterpret_cast<const uint4*>(p.ops + l + 2) + e);
      TC_TRACE(0);

      if (t0 < tiles) {
        const int64_t ldx = op.ldx, ldy = op.ldy;
        const float* sw = op.sw;
        const uint16_t* bias = op.bias;
        uint16_t* y = op.y;
        const int amax_cols = op.amax_cols;

        // ---- this rank's K/4 slice of the activation: read, quantise into the UMMA B layout --------
        // item = 16 consecutive K values of one token (one 16-byte chunk of a swizzled 128-byte B
        // row), ordered K-block-major: it = (kb * 16 + token) * 8 + chunk.  The slice is processed
        // in groups of 4 K blocks (512 items, 2 per thread) and every group is published to the MMA
        // warp on its own barrier, so the first K blocks are being multiplied (their weights are
        // already in the ring) while the rest of the slice is still being quantised.
        const int items = nkb * 128;
        const int ngroups = (nkb + 3) >> 2;
        const uint16_t* xk = op.x + (size_t)rank * nkb * 128;
        uint4 xa[2][2], xb[2][2];                       // two groups in flight: [item][vector]
        auto load_group = [&](int gi, uint4 (&xv)[2][2]) {
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int it = gi * 512 + i * 256 + e;
            const int m = (it >> 3) & 15;
            xv[i][0] = xv[i][1] = make_uint4(0, 0, 0, 0);
            if (it < items && m < p.M) {
              const uint4* s = reinterpret_cast<const uint4*>(xk + (size_t)m * ldx + ((it >> 7) * 8 + (it & 7)) * 16);
              xv[i][0] = __ldcg(s);
              xv[i][1] = __ldcg(s + 1);
            }
          }
        };
        load_group(0, xa);
        if (op.x_amax) {
          if (e < 16) amax_s[e] = (int)__ldcg(p.amax + l * 16 + e);
        } else {
          if (e < 16) amax_s[e] = 0;
          quant_sync();
          const int nvec = K >> 3;
          for (int m = warp - 4; m < p.M; m += 8) {
            float a = 0.0f;
            const uint4* row = reinterpret_cast<const uint4*>(op.x + (size_t)m * ldx);
            for (int i = lane; i < nvec; i += 32) a = fmaxf(a, amax8(__ldcg(row + i)));
            a = warp_max(a);
            if (lane == 0) amax_s[m] = __float_as_int(a);
          }
        }
        quant_sync();
        TC_TRACE(1);
        if (e < 16) sx_s[e] = __fdiv_rn(fmaxf(__int_as_float(amax_s[e]), 1e-8f), qmax);
        // quantise group gi from `cur` while group gi + 1 is being loaded into `nxt`, then publish it
        auto do_group = [&](int gi, uint4 (&cur)[2][2], uint4 (&nxt)[2][2]) {
          if (gi + 1 < ngroups) load_group(gi + 1, nxt);
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int it = gi * 512 + i * 256 + e;
            if (it < items) {
              const int m = (it >> 3) & 15, c = it & 7, kb = it >> 7;
              const float inv = __fdiv_rn(qmax, fmaxf(__int_as_float(amax_s[m]), 1e-8f));
              const uint2 q0 = quant8_mul<FMT>(cur[i][0], inv), q1 = quant8_mul<FMT>(cur[i][1], inv);
              uint8_t* dst = bres + kb * (16 * 128) + (m >> 3) * 1024 + (m & 7) * 128 + ((c ^ (m & 7)) << 4);
              *reinterpret_cast<uint4*>(dst) = make_uint4(q0.x, q0.y, q1.x, q1.y);
            }
          }
          if (gi + 1 == ngroups) {
            quant_sync();                                // every thread is done with this op's input amax
            if (e < 16) amax_s[e] = 0;                   // from here on: running max of |y| for the next op
          }
          fence_proxy_async_smem();
          quant_sync();
          if (lane == 0) mbar_arrive(&bready_bar[gi]);
        };
#pragma unroll 1
        for (int gi = 0; gi < ngroups; gi += 2) {
          do_group(gi, xa, xb);
          if (gi + 1 < ngroups) do_group(gi + 1, xb, xa);
        }
        TC_TRACE(2);

        if (epi) {
          for (int tile = t0; tile < tiles; tile += p.NC) {
            const int n = tile * 128 + q * 32 + lane;    // the weight row on this thread's TMEM lane
            const bool n_ok = n < N;
            const float swn = n_ok ? __ldg(sw + n) : 0.0f;
            const float bn = (n_ok && bias != nullptr) ? bf16_lo(__ldg(bias + n)) : 0.0f;
            if (e == 0) mbar_expect_tx(recv_bar, kTcRecv);
            mbar_wait(&tfull_bar[acc], acc_phase);
            tc_fence_after();
            if (tile == t0) TC_TRACE(3);
            // the tile's kTcAcc partial accumulators (one per 32-byte K step of a block): add them
            uint32_t r[16];
            {
              uint32_t r1[16], r2[16], r3[16];
              const uint32_t ta = lane_addr + acc * (kTcAcc * 16);
              tmem_ld16(ta, r);
              tmem_ld16(ta + 16, r1);
              tmem_ld16(ta + 32, r2);
              tmem_ld16(ta + 48, r3);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                if constexpr (FMT == QL_INT8) {
                  r[i] = (uint32_t)((int)r[i] + (int)r1[i] + (int)r2[i] + (int)r3[i]);
                } else {
                  const float s01 = __fadd_rn(__uint_as_float(r[i]), __uint_as_float(r1[i]));
                  const float s23 = __fadd_rn(__uint_as_float(r2[i]), __uint_as_float(r3[i]));
                  r[i] = __float_as_uint(__fadd_rn(s01, s23));
                }
              }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[acc]);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            // reduce-scatter: token columns 4d..4d+3 of this row go to rank d
            if (n_exch > 0) mbar_wait(free_bar, (n_exch - 1) & 1u);
            const uint32_t off = (uint32_t)((rank * 128 + q * 32 + lane) * 16);
#pragma unroll
            for (int d = 0; d < kDcC; ++d)
              st_async_v4(mapa_u32(recv_local + off, (uint32_t)d), r[4 * d], r[4 * d + 1], r[4 * d + 2], r[4 * d + 3],
                          mapa_u32(recv_bar_local, (uint32_t)d));
            mbar_wait(recv_bar, n_exch & 1u);
            if (tile == t0) TC_TRACE(4);
            uint4 sum = *reinterpret_cast<const uint4*>(recv + (q * 32 + lane) * 4);
#pragma unroll
            for (int s = 1; s < kDcC; ++s) {
              const uint4 o = *reinterpret_cast<const uint4*>(recv + (s * 128 + q * 32 + lane) * 4);
              if constexpr (FMT == QL_INT8) {
                sum.x += o.x; sum.y += o.y; sum.z += o.z; sum.w += o.w;
              } else {
                sum.x = __float_as_uint(__fadd_rn(__uint_as_float(sum.x), __uint_as_float(o.x)));
                sum.y = __float_as_uint(__fadd_rn(__uint_as_float(sum.y), __uint_as_float(o.y)));
                sum.z = __float_as_uint(__fadd_rn(__uint_as_float(sum.z), __uint_as_float(o.z)));
                sum.w = __float_as_uint(__fadd_rn(__uint_as_float(sum.w), __uint_as_float(o.w)));
              }
            }
            const uint32_t raw[4] = {sum.x, sum.y, sum.z, sum.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const int m = rank * 4 + c;
              float mx = 0.0f;
              if (m < p.M && n_ok) {
                const float a = FMT == QL_INT8 ? __int2float_rn((int)raw[c]) : __uint_as_float(raw[c]);
                float v = __fmul_rn(__fmul_rn(a, sx_s[m]), swn);              // (acc * sa) * sb, reference :121
                if (bias != nullptr) v = __fadd_rn(v, bn);
                const uint16_t o = f2bf(v);
                y[(size_t)m * ldy + n] = o;
                if (n < amax_cols) mx = fabsf(bf16_lo(o));
              }
              if (amax_cols > 0 && m < p.M) {
                mx = warp_max(mx);
                if (lane == 0 && mx > 0.0f) atomicMax(amax_s + m, __float_as_int(mx));
              }
            }
            epi_sync();
            if (e < kDcC) mbar_arrive_relaxed_cluster(mapa_u32(free_bar_local, (uint32_t)e));
            ++n_exch;
          }
          if (amax_cols > 0 && e < 16 && e < p.M && amax_s[e] > 0)
            atomicMax(reinterpret_cast<int*>(p.amax) + (l + 1) * 16 + e, amax_s[e]);
          TC_TRACE(7);
        }
      }

      // ---- grid barrier (the epilogue warps; the helper quantiser warps wait at the 256-thread sync) --
      if (l + 1 < p.n_ops) {
        if (e < 8 && l + 2 < p.n_ops) reinterpret_cast<uint4*>(desc_s + (l + 2) % 3)[e] = dpre;
        if (epi) {
          epi_sync();                // orders the epilogue threads' stores / atomics before the release
          TC_TRACE(8);
          if (e == 0) {
            red_release_gpu_add(p.sync, 1u);
            TC_TRACE(9);
            const uint32_t target = (uint32_t)(l + 1) * (uint32_t)G;
            if (ld_acquire_gpu(p.sync) < target) {
              const uint64_t t0g = globaltimer_ns();
              while (ld_acquire_gpu(p.sync) < target)
                if (globaltimer_ns() - t0g > 4000000000ull) __trap();
            }
            TC_TRACE(10);
          }
        }
        quant_sync();
      }
    }

    quant_sync();
    if (e == 0) {
      __threadfence();
      *flag_s = atomicAdd(p.sync + 1, 1u) == (uint32_t)(G - 1);
    }
    quant_sync();
    if (*flag_s) {
      for (int i = e; i < p.n_ops * 16; i += 256) p.amax[i] = 0u;
      if (e == 0) {
        p.sync[0] = 0u;
        p.sync[1] = 0u;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_arrive();
  cluster_wait();
  if (warp == 2) tmem_dealloc(tmem_base, 2 * kTcAcc * 16);
}

// ------------------------------------------------------------------ host side

struct ChainPlanHost {      // what ql_chain_run needs; lives in the caller's ql_chain_plan blob
  uint32_t magic;
  int n_ops, M, fmt, clusters, smem, cooperative;
  int tc;                   // 1: tcgen05 kernel (every K % 512 == 0, K <= 12288), 0: mma.sync kernel
  size_t off_amax, off_ops, off_tmaps, total;
};
static_assert(sizeof(ChainPlanHost) <= sizeof(ql_chain_plan), "ql_chain_plan too small");

static size_t rup(size_t v, size_t a) { return (v + a - 1) / a * a; }

static const void* chain_kernel(int fmt, int tc) {
  if (tc)
    return fmt == QL_INT8 ? reinterpret_cast<const void*>(decode_chain_tc_kernel<QL_INT8>)
                          : reinterpret_cast<const void*>(decode_chain_tc_kernel<QL_FP8_E4M3>);
  return fmt == QL_INT8 ? reinterpret_cast<const void*>(decode_chain_kernel<QL_INT8>)
                        : reinterpret_cast<const void*>(decode_chain_kernel<QL_FP8_E4M3>);
}

static void chain_launch_cfg(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, int clusters, int cooperative,
                             cudaStream_t st, int tc) {
  memset(cfg, 0, sizeof(*cfg));
  cfg->gridDim = dim3((unsigned)(clusters * kDcC));
  cfg->blockDim = dim3(tc ? kTcThreads : kDcThreads);
  cfg->dynamicSmemBytes = (size_t)(tc ? kTcSmem : kDcSmem);
  cfg->stream = st;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kDcC;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeCooperative;
  attr[1].val.cooperative = 1;
  cfg->attrs = attr;
  cfg->numAttrs = cooperative ? 2u : 1u;
}

// Clusters of kDcC CTAs (one CTA per SM: ~220 KB of shared memory each) the device can hold at once.
// Every CTA spins on the grid barrier, so the grid must never exceed this.
static int chain_max_clusters(int fmt, int tc) {
  static int cached[64][4];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  const int f = (fmt == QL_INT8 ? 0 : 1) + 2 * (tc ? 1 : 0);
  if (cached[dev][f] > 0) return cached[dev][f];
  static SmemAttrOnce once[4];
  if (ensure_max_smem(chain_kernel(fmt, tc), tc ? kTcSmem : kDcSmem, once[f])) return 0;
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[2];
  chain_launch_cfg(&cfg, attr, num_sms() / kDcC, 0, nullptr, tc);
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, chain_kernel(fmt, tc), &cfg) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  if (n > num_sms() / kDcC) n = num_sms() / kDcC;
  if (const char* e = getenv("QL_CHAIN_CLUSTERS")) {
    const int v = atoi(e);
    if (v > 0 && v < n) n = v;
  }
  cached[dev][f] = n;
  return n;
}

// every op fits the tcgen05 kernel?  (QL_CHAIN_TC=0 forces the mma.sync kernel)
static int chain_wants_tc(const ql_chain_op* ops, int n_ops) {
  const char* e = getenv("QL_CHAIN_TC");
  if (e != nullptr && e[0] == '0') return 0;
  for (int l = 0; l < n_ops; ++l)
    if (ops[l].K % 512 != 0 || ops[l].K > 12288) return 0;
  return 1;
}

static int plan_chain(const ql_chain_op* ops, int n_ops, int64_t M, int clusters, ChainPlanHost* pl, ChainOp* dev_ops,
                      CUtensorMap* tmaps, int tc) {
  if (ops == nullptr || n_ops <= 0 || n_ops > 4096 || M <= 0 || M > 16) return QL_ERR_BAD_SHAPE;
  for (int l = 0; l < n_ops; ++l) {
    const ql_chain_op& o = ops[l];
    if (o.N <= 0 || o.K <= 0 || o.K % 128 != 0 || o.N % 8 != 0 || o.ldy < o.N) return QL_ERR_BAD_SHAPE;
    if (o.N / 8 > (1 << 22)) return QL_ERR_BAD_SHAPE;
    if (!o.qw || !o.sw || !o.y) return QL_ERR_NULL;
    if (l == 0 && o.x == nullptr) return QL_ERR_NULL;
    const bool chained = o.x == nullptr;
    const void* x = chained ? ops[l - 1].y : o.x;
    const int64_t ldx = chained ? ops[l - 1].ldy : o.ldx;
    if (chained && ops[l - 1].N < o.K) return QL_ERR_BAD_SHAPE;
    if (ldx < o.K) return QL_ERR_BAD_SHAPE;
    if ((reinterpret_cast<uintptr_t>(x) & 15) || ldx % 8 != 0 || (reinterpret_cast<uintptr_t>(o.qw) & 15) ||
        (reinterpret_cast<uintptr_t>(o.y) & 15) || o.ldy % 8 != 0)
      return QL_ERR_MISALIGNED;
    const int nb64 = (int)(o.K / 64);
    const int kse = nb64 % 4 == 0 ? 4 : 2;                // K % 128 == 0: at least 2 blocks
    const int nblk = nb64 / kse;
    if (nblk > kDcMaxBlk) return QL_ERR_BAD_SHAPE;
    int cb = 1;
    for (int d = kDcSlotBlk; d >= 1; --d)
      if (nblk % d == 0) { cb = d; break; }
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
      d.kse = kse;
      d.nblk = nblk;
      d.cb = cb;
      d.nchunk = nblk / cb;
      d.RG = (int)(o.N / 8);
      d.rot = clusters > 0 ? (int)(((int64_t)l * 11) % clusters) : 0;   // spread the one-group-longer units
      d.x_amax = chained ? 1 : 0;
      d.amax_cols = (l + 1 < n_ops && ops[l + 1].x == nullptr) ? (int)ops[l + 1].K : 0;
      if (tc) {
        if (!make_tmap_2d(&tmaps[l], o.qw, CU_TENSOR_MAP_DATA_TYPE_UINT8, o.K, o.N, o.K, 128, 128)) return QL_ERR_CUDA;
      } else if (!make_tmap_2d_plain(&tmaps[l], o.qw, CU_TENSOR_MAP_DATA_TYPE_UINT32, o.K / 4, o.N, o.K, cb * 16, 8)) {
        return QL_ERR_CUDA;
      }
    }
  }
  pl->magic = kDcMagic;
  pl->n_ops = n_ops;
  pl->M = (int)M;
  pl->clusters = clusters;
  pl->smem = tc ? kTcSmem : kDcSmem;
  pl->tc = tc;
  pl->off_amax = 256;
  pl->off_ops = pl->off_amax + rup((size_t)n_ops * 64, 256);
  pl->off_tmaps = pl->off_ops + rup((size_t)n_ops * sizeof(ChainOp), 256);
  pl->total = pl->off_tmaps + rup((size_t)n_ops * sizeof(CUtensorMap), 256);
  return QL_OK;
}

}  // namespace ql

using namespace ql;

extern "C" {

size_t ql_chain_workspace_bytes(const ql_chain_op* ops, int n_ops, int64_t M) {
  ChainPlanHost pl;
  if (plan_chain(ops, n_ops, M, 0, &pl, nullptr, nullptr, 0) != QL_OK) return 0;
  return pl.total;
}

int ql_chain_prepare(const ql_chain_op* ops, int n_ops, int64_t M, int fmt, int64_t group, void* workspace,
                     size_t workspace_bytes, ql_chain_plan* plan, ql_stream_t stream) {
  if (fmt != QL_INT8 && fmt != QL_FP8_E4M3) return QL_ERR_BAD_FORMAT;
  if (group != 0) return QL_ERR_BAD_SHAPE;          // per-row scales only (per-group: ql_linear)
  if (plan == nullptr || workspace == nullptr) return QL_ERR_NULL;
  if (reinterpret_cast<uintptr_t>(workspace) & 255) return QL_ERR_MISALIGNED;
  if (int e = ql_device_check()) return e;
  if (ops == nullptr || n_ops <= 0) return QL_ERR_BAD_SHAPE;
  const int tc = chain_wants_tc(ops, n_ops);
  const int clusters = chain_max_clusters(fmt, tc);
  if (clusters <= 0) return QL_ERR_CUDA;
  ChainPlanHost pl;
  const size_t nn = (size_t)(n_ops > 0 ? n_ops : 1);
  ChainOp* host_ops = static_cast<ChainOp*>(malloc(sizeof(ChainOp) * nn));
  CUtensorMap* host_tm = static_cast<CUtensorMap*>(aligned_alloc(64, sizeof(CUtensorMap) * nn));
  if (host_ops == nullptr || host_tm == nullptr) {
    free(host_ops);
    free(host_tm);
    return QL_ERR_CUDA;
  }
  int e = plan_chain(ops, n_ops, M, clusters, &pl, host_ops, host_tm, tc);
  if (e == QL_OK && workspace_bytes < pl.total) e = QL_ERR_WORKSPACE;
  if (e == QL_OK) {
    pl.fmt = fmt;
    pl.cooperative = 1;
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    cudaStream_t st = (cudaStream_t)stream;
    // counters and amax arrays start at zero; the kernel re-arms them itself
    if (cudaMemsetAsync(ws, 0, pl.off_ops, st) != cudaSuccess ||
        cudaMemcpyAsync(ws + pl.off_ops, host_ops, sizeof(ChainOp) * (size_t)n_ops, cudaMemcpyHostToDevice, st) !=
            cudaSuccess ||
        cudaMemcpyAsync(ws + pl.off_tmaps, host_tm, sizeof(CUtensorMap) * (size_t)n_ops, cudaMemcpyHostToDevice,
                        st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) {
      (void)cudaGetLastError();
      e = QL_ERR_CUDA;
    }
  }
  free(host_ops);
  free(host_tm);
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
  p.tmaps = reinterpret_cast<const CUtensorMap*>(ws + pl.off_tmaps);
  p.n_ops = pl.n_ops;
  p.M = pl.M;
  p.NC = pl.clusters;
  p.sync = reinterpret_cast<uint32_t*>(ws);
  p.amax = reinterpret_cast<uint32_t*>(ws + pl.off_amax);
  p.trace = skinny_trace_ptr();
  const void* kern = chain_kernel(pl.fmt, pl.tc);
  // every CTA spins on the grid barrier: the grid must be co-resident.  Cooperative launch makes the
  // driver check that; where it refuses the cluster + cooperative combination the grid is still
  // sized by cudaOccupancyMaxActiveClusters (chain_max_clusters) and the barrier spin is bounded.
  // A refused cooperative launch is an ERROR by default: a plain cluster launch has no residency
  // guarantee, and if another stream holds SMs the grid barrier would spin until its bounded wait
  // traps and poisons the context.  QL_CHAIN_ALLOW_NONCOOP=1 accepts that risk (a caller that
  // knows the device is otherwise idle).
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[2];
  void* args[] = {&p};
  chain_launch_cfg(&cfg, attr, pl.clusters, 1, (cudaStream_t)stream, pl.tc);
  cudaError_t err = cudaLaunchKernelExC(&cfg, kern, args);
  if (err != cudaSuccess) {
    const char* allow = getenv("QL_CHAIN_ALLOW_NONCOOP");
    if (getenv("QL_DEBUG") != nullptr)
      fprintf(stderr, "[qlinear_b200] decode chain: cooperative cluster launch refused (%s)%s\n",
              cudaGetErrorString(err), (allow && allow[0] == '1') ? ", plain cluster launch" : "");
    (void)cudaGetLastError();
    if (allow == nullptr || allow[0] != '1') return QL_ERR_CUDA;
    chain_launch_cfg(&cfg, attr, pl.clusters, 0, (cudaStream_t)stream, pl.tc);
    err = cudaLaunchKernelExC(&cfg, kern, args);
  }
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
