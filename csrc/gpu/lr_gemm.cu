// K12 - logistic-regression gradient as a tcgen05 GEMM whose output tiles are reduce-scattered to
// the owner GPUs as they retire (consumer: ytk-learn's distributed L-BFGS / LR,
// R/README.md:268-278, which the reference serves with reduceScatterArray, PCS:1174-1209).
//
//   G[F x C] = sum over ranks of  Xt[F x K] * Rt[C x K]^T        (K = local rows, bf16 in, fp32 out)
//   rank q ends up owning rows [q*F/p, (q+1)*F/p) of G           (the reduce-scatter)
//
// Both operands are K-major (the design matrix is kept feature-major, Xt, next to the row-major X
// of the forward pass; the residuals are produced transposed), which is the canonical tcgen05
// "TN" case: TMA (cp.async.bulk.tensor, 128-byte swizzle) stages [128 x 64] A tiles and [C x 64]
// B tiles in shared memory through a 4-stage mbarrier ring; ONE elected thread issues
// tcgen05.mma.cta_group::1.kind::f16 (M = 128, N = C, K = 16) with the fp32 accumulator in TMEM;
// four epilogue warps read the accumulator back with tcgen05.ld (32 lanes x 32 columns per warp)
// and add every row straight into the OWNER rank's shard in the symmetric heap with
// red.global.add.f32 over NVLink - no local G is ever materialised, no separate collective runs.
// Split-K across CTAs fills the 148 SMs when F/128 tiles are too few; the same atomics merge the
// splits. Roles per CTA (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA
// issuer, warps 2-5 = epilogue.
#include <cuda.h>
#include <cooperative_groups.h>
#include <cuda_bf16.h>

#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "cuda_check.h"
#include "gpu_plane.h"
#include "launch.h"

namespace cg = cooperative_groups;

namespace mp4j {

namespace {

constexpr int kTileM = 128;    // features per CTA tile (UMMA M)
constexpr int kBlockK = 64;    // bf16 elements per stage row = 128 bytes = one swizzle atom
constexpr int kUmmaK = 16;
constexpr int kMaxStages = 8;  // ring depth is a launch parameter: as deep as 200 KB of smem allows
constexpr int kGemmThreads = 192;
constexpr int kMaxC = 256;

struct LrArgs {
  CUtensorMap tma_a[kMaxRanks];  // Xt of each local rank: dims {K, F}
  CUtensorMap tma_b[kMaxRanks];  // Rt of each local rank: dims {K, C}
  PeerTable tab;
  unsigned long long out_off;    // heap offset of this rank's G shard [feats_per_rank x C] fp32
  int F, C, K;
  int feats_per_rank;            // ceil(F / p)
  int k_blocks_per_split;        // 64-element K blocks handled by one CTA
  int tmem_cols;                 // power of two >= max(32, C)
  int stages;                    // smem ring depth (2..kMaxStages)
  int cluster;                   // CTAs per cluster along split-K (1 = none): partials meet in DSMEM
  int mt;                        // 128-row M tiles per CTA (1 or 2) sharing every staged B tile
  // world > 1: rows owned by OTHER ranks are first merged over the split-K CTAs in a local fp32
  // scratch [F x C] (L2-resident atomics); the last CTA of a tile forwards the merged rows to their
  // owners, so a remote shard receives one contribution per rank instead of one per split
  float* scratch[kMaxRanks];     // per local rank; nullptr = add straight into the owner's shard
  int* tile_cnt[kMaxRanks];      // per local rank: arrivals per M tile (reset by the last CTA)
  int splits;
  // the forwarded rows are STORED (coalesced 128-byte NVLink writes, ~5x the throughput of 16-byte
  // remote atomics) into the owner's receive area [source rank][feats_per_rank x C]; the owner adds
  // the p-1 areas to its shard after the closing barrier (lr_sum_recv_kernel)
  unsigned long long recv_off;
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
// bounded wait: a protocol bug must surface as a trap, never as a hung GPU
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const unsigned long long t0 = globaltimer_ns();
  for (;;) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
    if (done) return;
    if (globaltimer_ns() - t0 > 4000000000ull) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_addr(smem_dst)),
      "l"(map), "r"(c0), "r"(c1), "r"(smem_addr(bar))
      : "memory");
}
// K-major, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart (SBO), descriptor version 1
__device__ __forceinline__ uint64_t umma_desc_sw128(const void* smem_tile, int k_byte_offset) {
  const uint32_t addr = smem_addr(smem_tile) + (uint32_t)k_byte_offset;
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3ffffu) >> 4);       // start address      [0,14)
  d |= (uint64_t)1 << 16;                          // leading byte off.  [16,30) (unused for SW128 K-major)
  d |= (uint64_t)(1024 >> 4) << 32;                // stride byte offset [32,46)
  d |= (uint64_t)1 << 46;                          // version = 1 (Blackwell)
  d |= (uint64_t)2 << 61;                          // layout type SWIZZLE_128B
  return d;
}
// kind::f16 instruction descriptor: D = fp32, A = B = bf16, both K-major, M = 128, N = n
__device__ __forceinline__ uint32_t umma_idesc_bf16(int n) {
  uint32_t d = 0;
  d |= 1u << 4;                    // c_format = F32
  d |= 1u << 7;                    // a_format = BF16
  d |= 1u << 10;                   // b_format = BF16
  d |= (uint32_t)(n >> 3) << 17;   // n_dim
  d |= (uint32_t)(kTileM >> 4) << 24;  // m_dim
  return d;
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(bar))
               : "memory");
}

__global__ void __launch_bounds__(kGemmThreads, 2) lr_xtr_rs_kernel(const __grid_constant__ LrArgs a) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages], empty_bar[kMaxStages], tmem_full_bar;
  __shared__ uint32_t tmem_base_s;
  const PeerTable& tab = a.tab;
  const int rank = tab.emulate ? (int)blockIdx.z : tab.rank;
  const int row = tab.emulate ? rank : 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_m = blockIdx.x;
  const int kb_total = (a.K + kBlockK - 1) / kBlockK;
  const int kb_begin = blockIdx.y * a.k_blocks_per_split;
  const int kb_end = min(kb_total, kb_begin + a.k_blocks_per_split);
  const int nkb = kb_end - kb_begin;  // >= 1 by construction of the grid
  // mt M tiles share one staged B tile: B is re-read from L2 by every CTA, so at mt = 1 the
  // L2 -> SM bytes are twice the DRAM bytes (profiles/README.md §4); mt = 2 cuts that to 1.5x
  constexpr uint32_t a_tile_bytes = kTileM * kBlockK * 2;
  const uint32_t a_bytes = (uint32_t)a.mt * a_tile_bytes, b_bytes = (uint32_t)a.C * kBlockK * 2;
  // 1024-byte aligned stage buffers (128B swizzle atoms repeat every 1024 B)
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  unsigned char* smem_a = smem;
  unsigned char* smem_b = smem + (size_t)a.stages * a_bytes;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tma_a[row]) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tma_b[row]) : "memory");
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {  // one warp allocates TMEM and later frees it
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&tmem_base_s)),
                 "r"((uint32_t)a.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_d = tmem_base_s;

  if (warp == 0) {
    // ===== TMA producer
    if (lane == 0) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % a.stages;
        if (i >= a.stages) mbar_wait(&empty_bar[s], ((i / a.stages) - 1) & 1);
        mbar_expect_tx(&full_bar[s], a_bytes + b_bytes);
        const int k0 = (kb_begin + i) * kBlockK;
        for (int t = 0; t < a.mt; ++t)
          tma_load_2d(smem_a + (size_t)s * a_bytes + (size_t)t * a_tile_bytes, &a.tma_a[row], k0,
                      (tile_m * a.mt + t) * kTileM, &full_bar[s]);
        tma_load_2d(smem_b + (size_t)s * b_bytes, &a.tma_b[row], k0, 0, &full_bar[s]);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: a single thread drives the tensor core
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(a.C);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % a.stages;
        mbar_wait(&full_bar[s], (i / a.stages) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        for (int t = 0; t < a.mt; ++t) {
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            const uint64_t da =
                umma_desc_sw128(smem_a + (size_t)s * a_bytes + (size_t)t * a_tile_bytes, k * kUmmaK * 2);
            const uint64_t db = umma_desc_sw128(smem_b + (size_t)s * b_bytes, k * kUmmaK * 2);
            umma_bf16(tmem_d + (uint32_t)(t * a.C), da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[s]);  // frees the smem stage once these MMAs have read it
      }
      umma_commit(&tmem_full_bar);   // accumulator complete
    }
  } else {
    // ===== epilogue warps 2..5: TMEM -> registers -> red.add into the owner's shard
    const int quarter = warp & 3;  // a warp may only touch TMEM lanes [32*(warp%4), +32)
    mbar_wait(&tmem_full_bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
   for (int t = 0; t < a.mt; ++t) {
    const int f = (tile_m * a.mt + t) * kTileM + quarter * 32 + lane;  // global feature row of this thread
    const int owner = min(f / a.feats_per_rank, tab.world - 1);
    const int local_row = f - owner * a.feats_per_rank;
    float* dst_row = (float*)(tab.heap[owner] + a.out_off) + (size_t)local_row * a.C;
    if (a.scratch[row] != nullptr && owner != rank) dst_row = a.scratch[row] + (size_t)f * a.C;
    for (int c0 = 0; c0 < a.C; c0 += 32) {
      uint32_t v[32];
      const uint32_t taddr = tmem_d + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(t * a.C + c0);
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
          "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
            "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
            "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
          : "r"(taddr)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (a.cluster > 1) {
        // stash this CTA's partial tile in (now idle) stage memory; the cluster sums it below
        float* prow = (float*)smem + (size_t)(quarter * 32 + lane) * (a.C + 4) + c0;
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          if (c0 + j < a.C)
            *(float4*)(prow + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                               __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
      } else if (f < a.F) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          if (c0 + j < a.C)
            asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst_row + c0 + j),
                         "f"(__uint_as_float(v[j])), "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])),
                         "f"(__uint_as_float(v[j + 3]))
                         : "memory");
        }
      }
    }
   }
   if (a.scratch[row] != nullptr) {
     // ---- split-K fix-up: the last CTA of this tile forwards the locally merged foreign rows
     __shared__ int last_s;
     const int et = threadIdx.x - 64;  // 0..127 within the epilogue warps
     __threadfence();                  // my adds are performed before the arrival is counted
     asm volatile("bar.sync 1, 128;" ::: "memory");
     if (et == 0) {
       const int old = atomicAdd(&a.tile_cnt[row][tile_m], 1);
       last_s = (old == a.splits - 1);
       if (last_s) a.tile_cnt[row][tile_m] = 0;  // ready for the next call
     }
     asm volatile("bar.sync 1, 128;" ::: "memory");
     if (last_s) {
       __threadfence();
       const int c4n = a.C / 4;
       const int rows = a.mt * kTileM;
       for (int idx = et; idx < rows * c4n; idx += 128) {
         const int f2 = tile_m * rows + idx / c4n;
         if (f2 >= a.F) break;
         const int owner2 = min(f2 / a.feats_per_rank, tab.world - 1);
         if (owner2 == rank) continue;
         const int c = (idx % c4n) * 4;
         const float* src = a.scratch[row] + (size_t)f2 * a.C + c;
         float4 x;
         asm volatile("ld.relaxed.gpu.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                      : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w)
                      : "l"(src)
                      : "memory");
         float* dst = (float*)(tab.heap[owner2] + a.recv_off) +
                      ((size_t)rank * a.feats_per_rank + (size_t)(f2 - owner2 * a.feats_per_rank)) * a.C + c;
         asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(x.x), "f"(x.y),
                      "f"(x.z), "f"(x.w)
                      : "memory");
       }
     }
   }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)a.tmem_cols)
                 : "memory");
  }
  if (a.cluster > 1) {
    // Split-K partials of one output tile live in the shared memory of the CTAs of one cluster:
    // CTA j sums row slice j over all of them through DSMEM (fixed order) and sends ONE contribution
    // per rank to the owner, instead of `splits` remote atomics per element.
    cg::cluster_group cluster = cg::this_cluster();
    cluster.sync();
    const int cs = a.cluster;
    const int cr = (int)cluster.block_rank();
    const int rows_per = kTileM / cs;
    const int c4n = a.C / 4;
    const int pitch = a.C + 4;
    const uint32_t base = smem_addr(smem);
    for (int idx = threadIdx.x; idx < rows_per * c4n; idx += blockDim.x) {
      const int r = cr * rows_per + idx / c4n;
      const int c = (idx % c4n) * 4;
      const uint32_t local = base + (uint32_t)((r * pitch + c) * 4);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j = 0; j < cs; ++j) {
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(j));
        float4 x;
        asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w)
                     : "r"(remote)
                     : "memory");
        acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
      }
      const int f = tile_m * kTileM + r;
      if (f < a.F) {
        const int owner = min(f / a.feats_per_rank, tab.world - 1);
        float* dst = (float*)(tab.heap[owner] + a.out_off) + (size_t)(f - owner * a.feats_per_rank) * a.C + c;
        asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(acc.x), "f"(acc.y),
                     "f"(acc.z), "f"(acc.w)
                     : "memory");
      }
    }
    cluster.sync();  // nobody leaves while a peer may still read its shared memory
  }
}

// owner side of the store-forwarded contributions: shard += sum over source ranks of recv[src]
__global__ void lr_sum_recv_kernel(PeerTable tab, unsigned long long out_off, unsigned long long recv_off,
                                   int feats_per_rank, int F, int C) {
  const int rank = tab.emulate ? (int)blockIdx.y : tab.rank;
  const int rows = max(0, min(feats_per_rank, F - rank * feats_per_rank));
  const long long n4 = (long long)rows * C / 4;
  float4* shard = (float4*)(tab.heap[rank] + out_off);
  const float4* recv = (const float4*)(tab.heap[rank] + recv_off);
  const long long per_src4 = (long long)feats_per_rank * C / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 acc = shard[i];
    for (int q = 0; q < tab.world; ++q) {
      if (q == rank) continue;
      const float4 x = recv[(long long)q * per_src4 + i];
      acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
    }
    shard[i] = acc;
  }
}

// one-block cross-rank barrier between the phases (zero shard | accumulate | consume)
constexpr int kLrCtrlRow = kMaxBlocks - 2;
__global__ void lr_peer_barrier_kernel(PeerTable tab) {
  const int rank = tab.emulate ? (int)blockIdx.y : tab.rank;
  uint32_t* eslot = (uint32_t*)((char*)tab.sig[rank] + kSigEpochOff) + kLrCtrlRow;
  __shared__ uint32_t e_s;
  if (threadIdx.x == 0) e_s = *(volatile uint32_t*)eslot + 1;
  __syncthreads();
  const uint32_t e = e_s;
  if (threadIdx.x < (unsigned)tab.world) {
    st_release_sys(tab.sig[threadIdx.x] + kLrCtrlRow * kMaxRanks + rank, e);
    wait_flag(tab, tab.sig[rank] + kLrCtrlRow * kMaxRanks + threadIdx.x, e);
  }
  __syncthreads();
  if (threadIdx.x == 0) *(volatile uint32_t*)eslot = e;
}

}  // namespace

// shared by the fused consumers (K12, K13): one-block cross-rank barrier on the stream
void launch_peer_barrier(const PeerTable& tab, cudaStream_t s) {
  lr_peer_barrier_kernel<<<dim3(1, tab.emulate ? tab.world : 1), 32, 0, s>>>(tab);
  MP4J_CUDA(cudaGetLastError());
}

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult st;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &st);
    if (e != cudaSuccess || st != cudaDriverEntryPointSuccess || !p) throw Error("cuTensorMapEncodeTiled not available");
    fn = (EncodeTiledFn)p;
  }
  return fn;
}

// row-major [rows x K] bf16, K contiguous; box = [box_rows x 64]
void make_map(CUtensorMap* m, const void* base, long long rows, long long K, int box_rows, bool promo256 = false) {
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  cuuint32_t box[2] = {(cuuint32_t)kBlockK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = encode_tiled()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)base, dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              promo256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) throw Error("cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
}

}  // namespace

// Host entry: G shards are plane-owned symmetric buffers; returns their address for every local rank.
void GpuPlane::lr_xtr_reduce_scatter(const void* const* xt, const void* const* rt, int F, int C, long long K,
                                     float** shard_out, int* feats_per_rank_out, cudaStream_t s) {
  entry_check();
  const int p = world();
  const int L = local_ranks();
  if (F <= 0 || K <= 0) throw Error("lr_xtr_reduce_scatter: bad shape");
  if (C < 16 || C > kMaxC || C % 16 != 0) throw Error("lr_xtr_reduce_scatter: C must be a multiple of 16 in [16, 256]");
  if (K % 8 != 0) throw Error("lr_xtr_reduce_scatter: K (rows) must be a multiple of 8 (16-byte TMA row pitch)");
  const int fpr = (F + p - 1) / p;
  const size_t shard_bytes = (size_t)fpr * C * 4;
  if (lr_shard_bytes_ < shard_bytes) {
    MP4J_CUDA(cudaStreamSynchronize(s));
    if (lr_shard_bytes_) free(lr_shard_off_);
    lr_shard_off_ = alloc(shard_bytes);
    lr_shard_bytes_ = shard_bytes;
  }
  LrArgs a{};
  a.tab = table();
  a.out_off = lr_shard_off_;
  a.F = F;
  a.C = C;
  a.K = (int)K;
  a.feats_per_rank = fpr;
  // two M tiles per CTA when the problem is large enough to still fill the GPU (MP4J_LR_MT=1|2 overrides)
  const char* mt_env = getenv("MP4J_LR_MT");
  const char* mode = getenv("MP4J_LR_SPLIT");
  const bool want_cluster = mode && strcmp(mode, "cluster") == 0;
  a.mt = mt_env ? atoi(mt_env) : ((F >= 2 * kTileM * 16 && !want_cluster) ? 2 : 1);
  if (a.mt != 2 || want_cluster) a.mt = 1;
  int cols = 32;
  while (cols < a.mt * C) cols <<= 1;
  a.tmem_cols = cols;
  for (int i = 0; i < L; ++i) {
    // 256-byte L2 promotion on the streamed operand: the neighbouring 128 B of every row is the next
    // k-block's data, so DRAM sees 256-byte bursts per row (MP4J_LR_L2PROMO=128 restores the default)
    static const bool promo256 = !(getenv("MP4J_LR_L2PROMO") && atoi(getenv("MP4J_LR_L2PROMO")) == 128);
    make_map(&a.tma_a[i], xt[i], F, K, kTileM, promo256);
    make_map(&a.tma_b[i], rt[i], C, K, C);
  }
  const int tiles_m = (F + a.mt * kTileM - 1) / (a.mt * kTileM);
  const int kb_total = (int)((K + kBlockK - 1) / kBlockK);
  // Ring depth: as deep as ~200 KB of shared memory allows (one CTA per SM): HBM latency under load is
  // 2-3 us, so a 4-stage ring (64 KB of A in flight) left DRAM at 53 % (profiles/README.md §4).
  const size_t stage_bytes = (size_t)a.mt * kTileM * kBlockK * 2 + (size_t)C * kBlockK * 2;
  // Two CTAs per SM with half the ring each (MP4J_LR_CTAS_PER_SM=1 disables): one CTA's prologue /
  // epilogue (TMEM round trip + atomics, no loads in flight) overlaps the other's main loop.
  const char* cps_env = getenv("MP4J_LR_CTAS_PER_SM");
  const int cps = (!(cps_env && atoi(cps_env) == 1) && (size_t)a.tmem_cols * 2 <= 512 && a.cluster <= 1 && !want_cluster) ? 2 : 1;
  a.stages = (int)std::max<size_t>(2, std::min<size_t>(kMaxStages, ((cps == 2 ? 108 : 200) * 1024) / stage_bytes));
  const int wave = sms_ * cps;
  // Split K so the grid is close to a whole number of waves of 148 CTAs (one CTA per SM): the first
  // version launched 320 CTAs = 2.16 waves, i.e. a third wave at 16 % occupancy.
  int best_splits = 1;
  double best_eff = 0.0;
  const int max_splits = std::max(1, std::min(kb_total / 4, 64));
  for (int sp = 1; sp <= max_splits; ++sp) {
    const int kps = (kb_total + sp - 1) / sp;
    const int real = (kb_total + kps - 1) / kps;
    const long long ctas = (long long)tiles_m * real;
    const long long waves = (ctas + wave - 1) / wave;
    const double eff = (double)ctas / (double)(waves * wave);
    if (eff > best_eff + 0.02 || (ctas >= 2 * wave && best_eff < 0.9 && eff > best_eff)) {
      best_eff = eff;
      best_splits = real;
    }
    if (best_eff >= 0.95 && ctas >= wave) break;
  }
  int splits = best_splits;
  a.k_blocks_per_split = (kb_total + splits - 1) / splits;
  splits = (kb_total + a.k_blocks_per_split - 1) / a.k_blocks_per_split;
  // Optional (MP4J_LR_SPLIT=cluster): put the CTAs of one output tile into one thread-block cluster
  // (2, 4 or 8 along K) and let them sum their partial tiles through distributed shared memory, so a
  // tile costs one remote contribution per rank instead of one per split. Measured on B200 it loses
  // to plain atomics (0.294 vs 0.209 ms at F=8192, C=128, K=65536): gang-scheduling 8-CTA clusters at
  // one CTA per SM strands SMs and serialises tile completion, so atomics stay the default.
  a.cluster = 1;
  if (splits > 1 && want_cluster) {
    for (int cs : {8, 4, 2}) {
      if (cs > splits + 1 && cs != 2) continue;  // do not split far more than the grid needs
      const int kps = (kb_total + cs - 1) / cs;
      if (kps < 1 || (kb_total + kps - 1) / kps != cs) continue;
      if ((size_t)a.stages * stage_bytes < (size_t)kTileM * (C + 4) * 4) continue;
      a.cluster = cs;
      a.k_blocks_per_split = kps;
      splits = cs;
      break;
    }
  }
  a.splits = splits;
  // local-first merge of the split-K partials of foreign rows (MP4J_LR_REMOTE=direct disables)
  const char* rm = getenv("MP4J_LR_REMOTE");
  const bool local_first = p > 1 && splits > 1 && a.cluster == 1 && !(rm && strcmp(rm, "direct") == 0);
  // Two receive areas used alternately: call n may write area n%2 of a peer without asking, because
  // the peer finished reading it (sum of call n-2) before it entered the closing barrier of call n-1,
  // which this rank has passed. That removes the opening cross-GPU barrier from this path.
  if (local_first && lr_recv_bytes_ < shard_bytes * p) {
    MP4J_CUDA(cudaStreamSynchronize(s));
    host_barrier();  // nobody may still be writing into the old area of a peer
    if (lr_recv_bytes_) free(lr_recv_off_);
    lr_recv_off_ = alloc(2 * shard_bytes * p);
    lr_recv_bytes_ = shard_bytes * p;
    host_barrier();
  }
  if (local_first) lr_recv_next_ ^= 1;
  a.recv_off = lr_recv_off_ + (size_t)lr_recv_next_ * lr_recv_bytes_;
  for (int i = 0; i < L; ++i) {
    a.scratch[i] = nullptr;
    a.tile_cnt[i] = nullptr;
    if (!local_first) continue;
    const size_t need = (size_t)F * C * 4;
    if (lr_scratch_bytes_[i] < need) {
      MP4J_CUDA(cudaStreamSynchronize(s));
      if (lr_scratch_[i]) MP4J_CUDA(cudaFree(lr_scratch_[i]));
      MP4J_CUDA(cudaMalloc(&lr_scratch_[i], need));
      lr_scratch_bytes_[i] = need;
    }
    if (lr_tiles_[i] < tiles_m) {
      MP4J_CUDA(cudaStreamSynchronize(s));
      if (lr_tile_cnt_[i]) MP4J_CUDA(cudaFree(lr_tile_cnt_[i]));
      MP4J_CUDA(cudaMalloc(&lr_tile_cnt_[i], (size_t)tiles_m * sizeof(int)));
      MP4J_CUDA(cudaMemset(lr_tile_cnt_[i], 0, (size_t)tiles_m * sizeof(int)));
      lr_tiles_[i] = tiles_m;
    }
    MP4J_CUDA(cudaMemsetAsync(lr_scratch_[i], 0, need, s));
    a.scratch[i] = lr_scratch_[i];
    a.tile_cnt[i] = lr_tile_cnt_[i];
  }
  const size_t smem = (size_t)a.stages * stage_bytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    MP4J_CUDA(cudaFuncSetAttribute(lr_xtr_rs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    attr_set = true;
  }
  // zero my shard | everyone zeroed | accumulate into owners | everyone done
  for (int i = 0; i < L; ++i) {
    const int r = emulated() ? i : rank();
    MP4J_CUDA(cudaMemsetAsync(heap_.base(r) + lr_shard_off_, 0, shard_bytes, s));
    shard_out[i] = (float*)(heap_.base(r) + lr_shard_off_);
  }
  *feats_per_rank_out = fpr;
  PeerTable tab = a.tab;
  // direct mode adds into remote shards: everyone must have zeroed first. The store-forwarding mode
  // never touches a remote shard and double-buffers its receive areas, so it needs no opening barrier.
  // (Under graph capture the alternation is frozen into the graph, so the barrier stays.)
  if (p > 1 && (!local_first || capturing(s)))
    lr_peer_barrier_kernel<<<dim3(1, emulated() ? p : 1), 32, 0, s>>>(tab);
  {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(tiles_m, splits, emulated() ? p : 1);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = a.cluster;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    MP4J_CUDA(cudaLaunchKernelEx(&cfg, lr_xtr_rs_kernel, a));
  }
  MP4J_CUDA(cudaGetLastError());
  if (p > 1) lr_peer_barrier_kernel<<<dim3(1, emulated() ? p : 1), 32, 0, s>>>(tab);
  if (local_first) {
    const long long n4 = (long long)fpr * C / 4;
    const int blocks = (int)std::max<long long>(1, std::min<long long>(2 * sms_, (n4 + 255) / 256));
    lr_sum_recv_kernel<<<dim3(blocks, emulated() ? p : 1), 256, 0, s>>>(tab, a.out_off, a.recv_off, fpr, F, C);
  }
  MP4J_CUDA(cudaGetLastError());
  last_ = CollStat{"lr_xtr_reduce_scatter", a.cluster > 1 ? "tcgen05+dsmem-splitk+red.add" : (local_first ? "tcgen05+local-splitk+red.add" : (a.mt == 2 ? "tcgen05-2mt+red.add" : "tcgen05+red.add")), (size_t)F * K * 2, tiles_m * splits, false};
}

}  // namespace mp4j
