// K19 + C3: k-means assign + accumulate on the 5th-generation tensor cores.
//
// Not in the reference source: BASELINE.json config 5 builds k-means from DoWhile / Apply +
// GroupBy-aggregate (LinqToDryad/DryadLinqQueryable.cs:1280-1306 is the client loop).  One
// iteration = for every point the nearest centroid (argmin_j |c_j|^2 - 2 x.c_j), then per-centroid
// sums and counts, then an all-reduce of those partials.
//
// km_assign is a persistent, warp-specialised kernel (one CTA per SM):
//   warp 0      TMA producer: point tiles [128 x 128] fp32 (double buffered) and centroid k-blocks
//               [128 x 32] fp32 through a 5-stage ring, all SWIZZLE_128B, mbarrier complete_tx.
//   warp 1      tcgen05.mma issuer (one elected lane): kind::tf32, M=128 N=128 K=8, accumulating
//               x.c_j for 128 points x 128 centroids into one of two TMEM accumulator buffers;
//               tcgen05.commit releases ring slots and publishes accumulators.
//   warps 2-9   epilogue (two warps per TMEM lane quarter, each scanning half of the accumulator
//               columns with two independent running minima so the argmin chain is not
//               latency-bound): tcgen05.ld the accumulator rows (one point per thread pair), fused
//               argmin with the fp32 centroid norms, then the point's exact fp32 coordinates are
//               read back from the swizzled shared tile and added to the winner's sum with
//               16-byte vector reductions (red.global.add.v4.f32) — sums stay fp32-exact, the
//               TF32 product only decides near-ties.
// km_update all-reduces sums/counts by loading every peer's partials over NVLink (peer-mapped
// arena, fixed rank order => deterministic) and writes the new centroids and norms: no NCCL.
#include "common.cuh"
#include "../../include/dryad_b200.h"
#include <cuda.h>
#include <cudaTypedefs.h>
#include <stdlib.h>

namespace dryad { namespace km {

constexpr int BM = 128, BN = 128, D = 128, KB = 32, NKB = D / KB;
constexpr int B_STAGES = 5;
constexpr int A_KB_BYTES = BM * KB * 4;          // 16 KiB
constexpr int A_BYTES = A_KB_BYTES * NKB;        // 64 KiB
constexpr int B_STAGE_BYTES = BN * KB * 4;       // 16 KiB
constexpr int MAX_K = 2048;
constexpr int THREADS = 320;                     // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue
constexpr int TMEM_COLS = 256;                   // two accumulator buffers of 128 fp32 columns

constexpr size_t OFF_A = 0;
constexpr size_t OFF_B = OFF_A + 2 * A_BYTES;
constexpr size_t OFF_CN = OFF_B + (size_t)B_STAGES * B_STAGE_BYTES;
constexpr size_t OFF_BAR = OFF_CN + MAX_K * 4;
constexpr size_t SMEM_BYTES = OFF_BAR + 192 + 128 * 8 + 1024;   // barriers, pair-exchange, alignment slack   // + slack for 1024-byte alignment

// ---- PTX wrappers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return done != 0;
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y) : "memory");
}
// TMA reduce-add of one 128-byte shared-memory row segment into a global fp32 tensor (the
// per-centroid sums): one L2 reduction transaction of 32 floats instead of eight 16-byte REDs.
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* map, int x, int y, const void* smem_src) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2}], [%3];"
               ::"l"(map), "r"(x), "r"(y), "r"(smem_u32(smem_src)) : "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_ld_32x32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void red_add_v4(float* addr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor: 8-row x 128-byte atoms stacked every
// 1024 bytes (stride byte offset), descriptor version 1 (sm_100), layout type 2.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);             // start address
  d |= (uint64_t)1 << 16;                              // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;                    // stride byte offset
  d |= (uint64_t)1 << 46;                              // version
  d |= (uint64_t)2 << 61;                              // SWIZZLE_128B
  return d;
}

// c=F32 (1<<4), a=b=TF32 (2<<7, 2<<10), K-major both, N>>3 at bit 17, M>>4 at bit 24.
constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

struct Bars {
  uint64_t a_full[2], a_empty[2], b_full[B_STAGES], b_empty[B_STAGES], t_full[2], t_empty[2];
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(THREADS, 1)
km_assign(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_c,
          const __grid_constant__ CUtensorMap map_s,
          const float* __restrict__ cnorm, int64_t n, int k, int32_t* __restrict__ assign,
          float* __restrict__ sums, unsigned long long* __restrict__ counts, float* __restrict__ inertia, int dbg) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* sA = smem + OFF_A;
  unsigned char* sB = smem + OFF_B;
  float* s_cn = reinterpret_cast<float*>(smem + OFF_CN);
  Bars* bars = reinterpret_cast<Bars*>(smem + OFF_BAR);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunks = k / BN;
  const int64_t ntiles_real = (n + BM - 1) / BM;
  const int64_t ntiles = ntiles_real;

  for (int i = threadIdx.x; i < k; i += THREADS) s_cn[i] = cnorm[i];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&bars->a_full[i], 1); mbar_init(&bars->a_empty[i], 256);
                                  mbar_init(&bars->t_full[i], 1); mbar_init(&bars->t_empty[i], 256); }
    for (int i = 0; i < B_STAGES; ++i) { mbar_init(&bars->b_full[i], 1); mbar_init(&bars->b_empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t it = 0, bs = 0, bphase = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const uint32_t ab = it & 1;
        mbar_wait(&bars->a_empty[ab], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&bars->a_full[ab], A_BYTES);
        for (int kb = 0; kb < NKB; ++kb)
          tma_load_2d(sA + ab * A_BYTES + kb * A_KB_BYTES, &map_x, kb * KB, (int)(tile * BM), &bars->a_full[ab]);
        for (int c = 0; c < nchunks; ++c)
          for (int kb = 0; kb < NKB; ++kb) {
            mbar_wait(&bars->b_empty[bs], bphase ^ 1);
            mbar_expect_tx(&bars->b_full[bs], B_STAGE_BYTES);
            tma_load_2d(sB + bs * B_STAGE_BYTES, &map_c, kb * KB, c * BN, &bars->b_full[bs]);
            if (++bs == B_STAGES) { bs = 0; bphase ^= 1; }
          }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    uint32_t it = 0, bs = 0, bphase = 0, cc = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ab = it & 1;
      mbar_wait(&bars->a_full[ab], (it >> 1) & 1);
      for (int c = 0; c < nchunks; ++c, ++cc) {
        const uint32_t tb = cc & 1;
        mbar_wait(&bars->t_empty[tb], ((cc >> 1) & 1) ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < NKB; ++kb) {
          mbar_wait(&bars->b_full[bs], bphase);
          tc_fence_after();
          if (lane == 0) {
            const uint64_t ad = make_desc(smem_u32(sA + ab * A_BYTES + kb * A_KB_BYTES));
            const uint64_t bd = make_desc(smem_u32(sB + bs * B_STAGE_BYTES));
#pragma unroll
            for (int kk = 0; kk < KB / 8; ++kk)   // K = 8 tf32 = 32 bytes per instruction
              tc_mma_tf32(tmem_base + tb * BN, ad + (uint64_t)(kk * 2), bd + (uint64_t)(kk * 2), IDESC, (kb | kk) != 0);
            tc_commit(&bars->b_empty[bs]);
            if (kb == NKB - 1) tc_commit(&bars->t_full[tb]);
          }
          __syncwarp();
          if (++bs == B_STAGES) { bs = 0; bphase ^= 1; }
        }
      }
    }
  } else {
    // ===== epilogue: one point per thread pair (column halves) =====
    const int ew = warp - 2;                        // 0..7
    const int q = warp & 3;                         // TMEM lane quarter this warp may read
    const int half = ew >> 2;                       // which 64 of the 128 accumulator columns
    const int r = q * 32 + lane;                    // row inside the tile
    float* s_best = reinterpret_cast<float*>(smem + OFF_BAR + 192);        // [128] partner exchange
    int* s_bi = reinterpret_cast<int*>(s_best + 128);                      // [128]
    uint32_t it = 0, cc = 0;
    float local_inertia = 0.f;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ab = it & 1;
      float best0 = 3.4e38f, best1 = 3.4e38f;
      int bi0 = 0, bi1 = 0;
      for (int c = 0; c < nchunks; ++c, ++cc) {
        const uint32_t tb = cc & 1;
        mbar_wait(&bars->t_full[tb], (cc >> 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int sub = 0; sub < 2; ++sub) {
          float v[32];
          const int col = half * 64 + sub * 32;
          tc_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + tb * BN + col, v);
          const int j0 = c * BN + col;
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float v0 = fmaf(-2.f, v[j], s_cn[j0 + j]);
            float v1 = fmaf(-2.f, v[j + 1], s_cn[j0 + j + 1]);
            if (v0 < best0) { best0 = v0; bi0 = j0 + j; }
            if (v1 < best1) { best1 = v1; bi1 = j0 + j + 1; }
          }
        }
        tc_fence_before();
        mbar_arrive(&bars->t_empty[tb]);
      }
      float best = best0; int bi = bi0;
      if (best1 < best || (best1 == best && bi1 < bi)) { best = best1; bi = bi1; }
      // merge the two column halves of each row through shared memory (named barrier: epilogue only)
      if (half == 1) { s_best[r] = best; s_bi[r] = bi; }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (half == 0) {
        const float ob = s_best[r]; const int oi = s_bi[r];
        if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        s_best[r] = best; s_bi[r] = bi;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      best = s_best[r]; bi = s_bi[r];
      mbar_wait(&bars->a_full[ab], (it >> 1) & 1);  // TMA-written tile visible to this thread
      const int64_t row = tile * BM + r;
      if (row < n) {
        if (half == 0) { assign[row] = bi; atomicAdd(&counts[bi], 1ull); }
        float xn = 0.f;
        float* dst = sums + (size_t)bi * D;
        const unsigned char* base = sA + ab * A_BYTES + (r >> 3) * 1024 + (r & 7) * 128;
        // each thread of the pair adds one half of the point's 128 coordinates: the exact fp32 row
        // segments go from the swizzled shared tile straight into sums[bi] by TMA reduce-add
#pragma unroll
        for (int kk = 0; kk < NKB / 2; ++kk) {
          const int kb = half * (NKB / 2) + kk;
          if (!(dbg & 4)) {     // default: vector REDs (measured 5 % faster than the TMA reduce-add form)
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
              float4 x = *reinterpret_cast<const float4*>(base + kb * A_KB_BYTES + ((ch ^ (r & 7)) << 4));
              xn += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
              red_add_v4(dst + kb * KB + ch * 4, x);
            }
          } else {
            tma_reduce_add_2d(&map_s, kb * KB, bi, base + kb * A_KB_BYTES);
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
              float4 x = *reinterpret_cast<const float4*>(base + kb * A_KB_BYTES + (ch << 4));
              xn += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;     // order inside the row is irrelevant for the norm
            }
          }
        }
        local_inertia += xn + (half == 0 ? best : 0.f);
      }
      // the TMA engine must have finished READING this thread's row segments before the tile
      // buffer is handed back to the producer
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      asm volatile("bar.sync 1, 256;" ::: "memory");   // s_best/s_bi reused by the next tile
      mbar_arrive(&bars->a_empty[ab]);
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // all reductions performed
    local_inertia = warp_sum(local_inertia);
    if (lane == 0 && inertia) atomicAdd(inertia, local_inertia);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
  }
}

// ---- CTA-pair variant (tcgen05 cta_group::2) -----------------------------------------------------
// Two CTAs on the two SMs of a TPC work on 256 points x 256 centroids per instruction: every CTA
// stages its own 128 points (A) and HALF of the centroid block (128 rows of B); the leader CTA's
// single MMA thread issues M=256 N=256 K=8 instructions that read both halves, so each SM's shared
// memory feeds 8 KB per 128 tensor-clocks instead of 8 KB per 64 — the SS-mode operand-bandwidth
// bound of the one-CTA kernel is halved.  Accumulators: 128 lanes x 256 columns per CTA, double
// buffered = the whole 512-column TMEM.
//   both CTAs   warp 0: TMA producer.  A tile -> local barrier; B half-blocks -> complete_tx on the
//               LEADER's barrier (cp.async.bulk.tensor ... .cta_group::2, barrier address with the
//               CTA-rank bit cleared), freed by the leader's multicast tcgen05.commit.
//   leader      warp 1: MMA issuer.   peer warp 1: relays "my A tile has landed" to the leader.
//   both CTAs   warps 2-9: epilogue on their own 128 accumulator rows (same fused argmin + exact
//               fp32 sums as above); accumulator buffers are released to the leader per warp.
constexpr int BN2 = 256;                         // centroids per instruction (128 staged per CTA)
constexpr int TMEM_COLS2 = 512;
constexpr uint32_t IDESC2 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN2 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
constexpr uint32_t PEER_BIT_MASK = 0xFEFFFFFFu;  // shared::cluster address of the same offset in CTA rank 0

// 64 accumulator columns per instruction, NOT waited for: the caller issues several loads back to
// back and calls tc_ld_wait() once, so the round trips overlap.
__device__ __forceinline__ void tc_ld_32x32_x64_nowait(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float4 lds_f4(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}
// (d & mask) | idx as ONE three-input LOP3 (and + or with two immediates compiled to two
// instructions: ncu showed 2.1 LOP3 per accumulator column, the largest stall bucket of the scan).
// `mask` must be a value the assembler cannot fold (the kernel derives it from a run-time argument):
// then it lives in one register and every column index becomes the instruction's immediate.
__device__ __forceinline__ float pack_idx(float d, int idx, uint32_t mask) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(__float_as_uint(d)), "r"(mask), "r"((uint32_t)idx));
  return __uint_as_float(r);
}

struct Bars2 {
  uint64_t a_full[2], a_empty[2], a_ready[2], b_full[B_STAGES], b_empty[B_STAGES], t_full[2], t_empty[2];
  uint32_t tmem_base;
};
static_assert(sizeof(Bars2) <= 192, "barrier block overlaps the pair-exchange area");

__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar) & PEER_BIT_MASK), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
// arrive on the barrier at the same offset in the leader CTA (rank 0) of the pair
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(bar)), "r"(0));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(ra) : "memory");
}
// same, without the release fence: the caller only orders tcgen05.ld completions (already waited
// for and fenced with tcgen05.fence::before_thread_sync), no global-memory writes — a release at
// cluster scope would drain the outstanding reductions (ncu: 28 % of the peer CTA's epilogue time).
__device__ __forceinline__ void mbar_arrive_leader_relaxed(uint64_t* bar) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(bar)), "r"(0));
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(ra) : "memory");
}

constexpr size_t SMEM_BYTES2 = OFF_BAR + 192 + 2 * 4 * 128 * 8 + 1024;   // + [parity][part][row] exchange
template <int EW>      // epilogue warps per CTA: 8 or 16
__global__ void __launch_bounds__((2 + EW) * 32, 1)
km_assign_pair(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_c,
               const float* __restrict__ cnorm, int64_t n, int k, int32_t* __restrict__ assign,
               float* __restrict__ sums, unsigned long long* __restrict__ counts, float* __restrict__ inertia, int dbg,
               const float* __restrict__ cent, const float* __restrict__ sep2, uint32_t* __restrict__ ulist,
               unsigned int* __restrict__ ucount) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* sA = smem + OFF_A;
  unsigned char* sB = smem + OFF_B;
  float* s_cn = reinterpret_cast<float*>(smem + OFF_CN);
  Bars2* bars = reinterpret_cast<Bars2*>(smem + OFF_BAR);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunks = k / BN2;
  const uint32_t crank = cluster_rank();
  const bool leader = crank == 0;
  const int64_t npairs = gridDim.x >> 1, pair = blockIdx.x >> 1;
  const int64_t ntiles = (n + 2 * BM - 1) / (2 * BM);          // pair tiles of 256 points

  for (int i = threadIdx.x; i < k; i += (2 + EW) * 32) s_cn[i] = cnorm[i];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->a_full[i], 1); mbar_init(&bars->a_empty[i], EW * 32); mbar_init(&bars->a_ready[i], 1);
      mbar_init(&bars->t_full[i], 1); mbar_init(&bars->t_empty[i], 2 * EW);     // epilogue warps x 2 CTAs
    }
    for (int i = 0; i < B_STAGES; ++i) { mbar_init(&bars->b_full[i], 1); mbar_init(&bars->b_empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {      // the same warp of BOTH CTAs allocates collectively
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)), "n"(TMEM_COLS2) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                 // the peer's barriers exist before anything is signalled remotely
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ===== TMA producer (both CTAs) =====
    // The point tile of the NEXT pair-tile is requested as soon as its buffer is free (polled
    // between centroid stages), not after the current tile's last centroid stage: a 64 KiB tile
    // straight from HBM takes longer than the five stages of look-ahead (ncu: the MMA issuer spent
    // 76 % of its time waiting for A before this).
    if (lane == 0) {
      uint32_t it = 0, bs = 0, bphase = 0;
      auto issue_a = [&](int64_t tile, uint32_t itx) {
        const uint32_t ab = itx & 1;
        mbar_expect_tx(&bars->a_full[ab], A_BYTES);
        for (int kb = 0; kb < NKB; ++kb)
          tma_load_2d(sA + ab * A_BYTES + kb * A_KB_BYTES, &map_x, kb * KB, (int)(tile * 2 * BM + crank * BM), &bars->a_full[ab]);
      };
      if (pair < ntiles) issue_a(pair, 0);                       // buffers start empty
      for (int64_t tile = pair; tile < ntiles; tile += npairs, ++it) {
        const int64_t next = tile + npairs;
        bool next_done = next >= ntiles;
        const uint32_t nab = (it + 1) & 1, nparity = (((it + 1) >> 1) & 1) ^ 1;
        for (int c = 0; c < nchunks; ++c)
          for (int kb = 0; kb < NKB; ++kb) {
            if (!next_done && mbar_try(&bars->a_empty[nab], nparity)) { issue_a(next, it + 1); next_done = true; }
            mbar_wait(&bars->b_empty[bs], bphase ^ 1);            // leader's multicast commit freed the slot in both CTAs
            if (leader) mbar_expect_tx(&bars->b_full[bs], 2 * B_STAGE_BYTES);   // both halves report to the leader
            tma_load_2d_pair(sB + bs * B_STAGE_BYTES, &map_c, kb * KB, c * BN2 + (int)crank * BN, &bars->b_full[bs]);
            if (++bs == B_STAGES) { bs = 0; bphase ^= 1; }
          }
        if (!next_done) { mbar_wait(&bars->a_empty[nab], nparity); issue_a(next, it + 1); }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      // ===== MMA issuer =====
      uint32_t it = 0, bs = 0, bphase = 0, cc = 0;
      for (int64_t tile = pair; tile < ntiles; tile += npairs, ++it) {
        const uint32_t ab = it & 1;
        mbar_wait(&bars->a_full[ab], (it >> 1) & 1);
        mbar_wait(&bars->a_ready[ab], (it >> 1) & 1);            // the peer's half of A
        for (int c = 0; c < nchunks; ++c, ++cc) {
          const uint32_t tb = cc & 1;
          mbar_wait(&bars->t_empty[tb], ((cc >> 1) & 1) ^ 1);
          tc_fence_after();
          for (int kb = 0; kb < NKB; ++kb) {
            mbar_wait(&bars->b_full[bs], bphase);
            tc_fence_after();
            if (lane == 0) {
              const uint64_t ad = make_desc(smem_u32(sA + ab * A_BYTES + kb * A_KB_BYTES));
              const uint64_t bd = make_desc(smem_u32(sB + bs * B_STAGE_BYTES));
#pragma unroll
              for (int kk = 0; kk < KB / 8; ++kk)
                tc_mma_tf32_pair(tmem_base + tb * BN2, ad + (uint64_t)(kk * 2), bd + (uint64_t)(kk * 2), IDESC2, (kb | kk) != 0);
              tc_commit_pair(&bars->b_empty[bs]);
              if (kb == NKB - 1) tc_commit_pair(&bars->t_full[tb]);
            }
            __syncwarp();
            if (++bs == B_STAGES) { bs = 0; bphase ^= 1; }
          }
        }
      }
    } else {
      // ===== relay: tell the leader that this CTA's A tile has landed =====
      uint32_t it = 0;
      for (int64_t tile = pair; tile < ntiles; tile += npairs, ++it) {
        const uint32_t ab = it & 1;
        mbar_wait(&bars->a_full[ab], (it >> 1) & 1);
        if (lane == 0) mbar_arrive_leader(&bars->a_ready[ab]);
        __syncwarp();
      }
    }
  } else {
    // ===== epilogue (both CTAs): PARTS threads per point, each scanning 256/PARTS accumulator columns =====
    // argmin without a compare/select chain: the column index (7 bits) replaces the low mantissa
    // bits of the distance, so a running fminf carries value and index together (relative error
    // 2^-16, far below the TF32 product noise); four independent minima per thread.
    constexpr int PARTS = EW / 4, COLS = BN2 / PARTS, SUBS = COLS / 32;
    const uint32_t pmask = 0xFFFFFF80u | ((uint32_t)k >> 31);      // = 0xFFFFFF80, opaque to the assembler
    const int ew = warp - 2;
    const int q = warp & 3;                         // TMEM lane quarter this warp may read
    const int part = ew >> 2;
    const int r = q * 32 + lane;
    float2* s_x = reinterpret_cast<float2*>(smem + OFF_BAR + 192);          // [2][PARTS][128] (best, index)
    uint32_t it = 0, cc = 0;
    float local_inertia = 0.f;
    for (int64_t tile = pair; tile < ntiles; tile += npairs, ++it) {
      const uint32_t ab = it & 1;
      float best = 3.4e38f;
      int bchunk = 0;
      // Accumulator columns travel TMEM -> registers in 64-column granules, two per chunk, always
      // one load in flight while the previous granule is reduced; the buffer is handed back to the
      // MMA issuer the moment its last granule has landed.
      constexpr int G = 64, NG = COLS / G;
      static_assert(NG == 2, "the granule pipeline below is written for two granules per chunk");
      uint32_t va[G], vb[G];
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + part * COLS;
      mbar_wait(&bars->t_full[cc & 1], (cc >> 1) & 1);
      tc_fence_after();
      tc_ld_32x32_x64_nowait(trow + (cc & 1) * BN2, va);
      for (int c = 0; c < nchunks; ++c, ++cc) {
        const uint32_t tb = cc & 1;
        const uint32_t cn_s = smem_u32(s_cn + c * BN2 + part * COLS);
        float m0 = 3.4e38f, m1 = 3.4e38f, m2 = 3.4e38f, m3 = 3.4e38f;
        tc_ld_wait();                                             // granule A of chunk c is in registers
        tc_ld_32x32_x64_nowait(trow + tb * BN2 + G, vb);          // granule B in flight ...
#pragma unroll
        for (int j = 0; j < G; j += 4) {                          // ... while A is reduced
          const float4 n4 = lds_f4(cn_s + j * 4);
          m0 = fminf(m0, pack_idx(fmaf(-2.f, __uint_as_float(va[j]), n4.x), j, pmask));
          m1 = fminf(m1, pack_idx(fmaf(-2.f, __uint_as_float(va[j + 1]), n4.y), j + 1, pmask));
          m2 = fminf(m2, pack_idx(fmaf(-2.f, __uint_as_float(va[j + 2]), n4.z), j + 2, pmask));
          m3 = fminf(m3, pack_idx(fmaf(-2.f, __uint_as_float(va[j + 3]), n4.w), j + 3, pmask));
        }
        tc_ld_wait();                                             // the whole chunk has left TMEM
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (leader) mbar_arrive(&bars->t_empty[tb]); else mbar_arrive_leader_relaxed(&bars->t_empty[tb]); }
        if (c + 1 < nchunks) {                                    // first granule of the next chunk in flight
          mbar_wait(&bars->t_full[tb ^ 1], ((cc + 1) >> 1) & 1);
          tc_fence_after();
          tc_ld_32x32_x64_nowait(trow + (tb ^ 1) * BN2, va);
        }
#pragma unroll
        for (int j = 0; j < G; j += 4) {
          const float4 n4 = lds_f4(cn_s + (G + j) * 4);
          m0 = fminf(m0, pack_idx(fmaf(-2.f, __uint_as_float(vb[j]), n4.x), G + j, pmask));
          m1 = fminf(m1, pack_idx(fmaf(-2.f, __uint_as_float(vb[j + 1]), n4.y), G + j + 1, pmask));
          m2 = fminf(m2, pack_idx(fmaf(-2.f, __uint_as_float(vb[j + 2]), n4.z), G + j + 2, pmask));
          m3 = fminf(m3, pack_idx(fmaf(-2.f, __uint_as_float(vb[j + 3]), n4.w), G + j + 3, pmask));
        }
        const float cm = fminf(fminf(m0, m1), fminf(m2, m3));
        if (cm < best) { best = cm; bchunk = c; }
      }
      int bi = bchunk * BN2 + part * COLS + (int)(__float_as_uint(best) & 0x7Fu);
      best = __uint_as_float(__float_as_uint(best) & 0xFFFFFF80u);
      // every part publishes its candidate, one named barrier, every reader merges all candidates
      float2* xrow = s_x + (size_t)ab * PARTS * 128;
      xrow[part * 128 + r] = make_float2(best, __int_as_float(bi));
      asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory");
      mbar_wait(&bars->a_full[ab], (it >> 1) & 1);  // TMA-written tile visible to this thread
      const int64_t row0 = tile * 2 * BM + crank * BM;
      if (part == 0) {                              // one thread per point: assignment and count
#pragma unroll
        for (int p = 1; p < PARTS; ++p) {
          const float2 o = xrow[p * 128 + r];
          const int oi = __float_as_int(o.y);
          if (o.x < best || (o.x == best && oi < bi)) { best = o.x; bi = oi; }
        }
        if (row0 + r < n) { assign[row0 + r] = bi; atomicAdd(&counts[bi], 1ull); local_inertia += best; }
      }
      // sums: one WARP per point — 32 lanes x 16 bytes = the point's whole fp32 row, read from the
      // swizzled tile and added to its centroid's row with one coalesced 512-byte reduction
      // (the lane-per-point form issued 32 half-used sectors per instruction; ncu: lg_throttle).
      {
        constexpr int ROWS_PER_WARP = BM / EW;
        const int kb = lane >> 3, ch = lane & 7;
        float xn = 0.f;
#pragma unroll 4
        for (int i = 0; i < ROWS_PER_WARP; ++i) {
          const int rr = ew * ROWS_PER_WARP + i;
          float2 o = xrow[rr];
          int wi = __float_as_int(o.y);
#pragma unroll
          for (int p = 1; p < PARTS; ++p) {
            const float2 o2 = xrow[p * 128 + rr];
            const int oi = __float_as_int(o2.y);
            if (o2.x < o.x || (o2.x == o.x && oi < wi)) { o = o2; wi = oi; }
          }
          if (row0 + rr < n) {
            const float4 x = lds_f4(smem_u32(sA + ab * A_BYTES + kb * A_KB_BYTES + (rr >> 3) * 1024 + (rr & 7) * 128 + ((ch ^ (rr & 7)) << 4)));
            const float x2 = x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
            xn += x2;
            if (!(dbg & 1)) red_add_v4(sums + (size_t)wi * D + kb * KB + ch * 4, x);
            if (sep2) {
              // Certificate (Elkan's first lemma): the TF32 products picked centroid wi; if the EXACT
              // fp32 distance to it is at most half the distance to wi's nearest other centroid, no
              // other centroid can be closer, so the choice equals the fp32 argmin.  Everything else
              // is queued for the fp32 re-check kernel.
              const float4 c4 = __ldg(reinterpret_cast<const float4*>(cent + (size_t)wi * D + kb * KB + ch * 4));
              float dp = x2 - 2.f * (x.x * c4.x + x.y * c4.y + x.z * c4.z + x.w * c4.w);
              dp = warp_sum(dp);
              if (lane == 0 && !(dp + s_cn[wi] <= sep2[wi])) ulist[atomicAdd(ucount, 1u)] = (uint32_t)(row0 + rr);
            }
          }
        }
        local_inertia += xn;
      }
      mbar_arrive(&bars->a_empty[ab]);
    }
    local_inertia = warp_sum(local_inertia);
    if (lane == 0 && inertia) atomicAdd(inertia, local_inertia);
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                 // nobody exits (or frees TMEM) while the pair still uses it
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS2) : "memory");
  }
}

// sep2[b] = (1/4) * min over j != b of |c_b - c_j|^2, shaved by a relative 1e-4 so that fp32 rounding in
// either side of the certificate can only make it MORE conservative.  One block per centroid.
__global__ void __launch_bounds__(256) km_sep(const float* __restrict__ c, int k, float* __restrict__ sep2) {
  __shared__ float s_c[D];
  __shared__ float s_min[8];
  const int b = blockIdx.x;
  if (threadIdx.x < D) s_c[threadIdx.x] = c[(size_t)b * D + threadIdx.x];
  __syncthreads();
  float best = 3.4e38f;
  for (int j = threadIdx.x; j < k; j += 256) {
    if (j == b) continue;
    float d2 = 0.f;
    const float4* cj = reinterpret_cast<const float4*>(c + (size_t)j * D);
#pragma unroll 8
    for (int q = 0; q < D / 4; ++q) {
      const float4 v = __ldg(cj + q);
      const float a0 = v.x - s_c[4 * q], a1 = v.y - s_c[4 * q + 1], a2 = v.z - s_c[4 * q + 2], a3 = v.w - s_c[4 * q + 3];
      d2 += a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3;
    }
    best = fminf(best, d2);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = fminf(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane_id() == 0) s_min[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = s_min[0];
    for (int w = 1; w < 8; ++w) m = fminf(m, s_min[w]);
    sep2[b] = (k > 1) ? 0.25f * m * (1.f - 1e-4f) : 3.4e38f;
  }
}

// Certificate pass: one warp per point (four points in flight per warp), exact fp32 distance to the
// assigned centroid against sep2.  Measured on 8 M x 128 x 1024: 0.8 ms as its own streaming pass
// over the points, against +2.8 ms when the same arithmetic sits in the tcgen05 kernel's epilogue
// (where it lengthens the critical path of every tile).
__global__ void __launch_bounds__(256) km_certify(const float* __restrict__ pts, int64_t n, const float* __restrict__ cent,
                                                  const float* __restrict__ cnorm, const float* __restrict__ sep2,
                                                  const int32_t* __restrict__ assign, uint32_t* __restrict__ ulist,
                                                  unsigned int* __restrict__ ucount) {
  const int lane = threadIdx.x & 31;
  const int64_t w0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  constexpr int U = 4;
  for (int64_t base = w0 * U; base < n; base += nw * U) {
    int b[U];
    float4 x[U], c4[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t r = base + u;
      b[u] = (r < n) ? assign[r] : 0;
      x[u] = (r < n) ? __ldg(reinterpret_cast<const float4*>(pts + (size_t)r * D) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) c4[u] = __ldg(reinterpret_cast<const float4*>(cent + (size_t)b[u] * D) + lane);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float dp = x[u].x * x[u].x + x[u].y * x[u].y + x[u].z * x[u].z + x[u].w * x[u].w -
                 2.f * (x[u].x * c4[u].x + x[u].y * c4[u].y + x[u].z * c4[u].z + x[u].w * c4[u].w);
      dp = warp_sum(dp);
      if (lane == 0 && base + u < n && !(dp + cnorm[b[u]] <= sep2[b[u]])) ulist[atomicAdd(ucount, 1u)] = (uint32_t)(base + u);
    }
  }
}

// fp32 re-check of the queued points: exact argmin of cnorm[j] - 2 x.c_j over ALL centroids in fp32
// FMA arithmetic (ties -> lowest index).  When it differs from the tensor-core choice the point is
// moved: assignment, counts and the centroid sums (exact fp32 rows) are corrected in place.
__global__ void __launch_bounds__(256) km_recheck(const float* __restrict__ pts, const float* __restrict__ cent,
                                                  const float* __restrict__ cnorm, int k, int32_t* __restrict__ assign,
                                                  float* __restrict__ sums, unsigned long long* __restrict__ counts,
                                                  const uint32_t* __restrict__ ulist, const unsigned int* __restrict__ ucount,
                                                  unsigned int* __restrict__ changed) {
  __shared__ __align__(16) float s_x[8][D];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const unsigned int total = *ucount;
  unsigned int w = blockIdx.x * 8 + wib;
  const unsigned int nw = gridDim.x * 8;
  for (; w < total; w += nw) {
    const uint32_t row = ulist[w];
    const float4 xv = __ldg(reinterpret_cast<const float4*>(pts + (size_t)row * D) + lane);
    reinterpret_cast<float4*>(s_x[wib])[lane] = xv;
    __syncwarp();
    float best = 3.4e38f;
    int bi = 0x7fffffff;
    for (int j = lane; j < k; j += 32) {
      const float4* cj = reinterpret_cast<const float4*>(cent + (size_t)j * D);
      float dot = 0.f;
#pragma unroll 8
      for (int q = 0; q < D / 4; ++q) {
        const float4 c4 = __ldg(cj + q);
        const float4 x4 = reinterpret_cast<const float4*>(s_x[wib])[q];
        dot = fmaf(x4.x, c4.x, dot); dot = fmaf(x4.y, c4.y, dot); dot = fmaf(x4.z, c4.z, dot); dot = fmaf(x4.w, c4.w, dot);
      }
      const float d = fmaf(-2.f, dot, cnorm[j]);
      if (d < best || (d == best && j < bi)) { best = d; bi = j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
    }
    const int old = assign[row];
    if (bi != old) {
      red_add_v4(sums + (size_t)bi * D + lane * 4, xv);
      red_add_v4(sums + (size_t)old * D + lane * 4, make_float4(-xv.x, -xv.y, -xv.z, -xv.w));
      if (lane == 0) {
        assign[row] = bi;
        atomicAdd(&counts[bi], 1ull);
        atomicAdd(&counts[old], ~0ull);              // -1 in two's complement
        if (changed) atomicAdd(changed, 1u);
      }
    }
    __syncwarp();
  }
}

__global__ void km_norms(const float* __restrict__ c, int k, float* __restrict__ cnorm) {
  int j = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  if (j >= k) return;
  float s = 0.f;
  for (int d = lane_id(); d < D; d += 32) { float v = c[(size_t)j * D + d]; s += v * v; }
  s = warp_sum(s);
  if (lane_id() == 0) cnorm[j] = s;
}

// All-reduce of (sums, counts) by peer loads + centroid update.  One thread per (centroid, dim).
__global__ void km_update(float* const* __restrict__ peer_sums, unsigned long long* const* __restrict__ peer_counts,
                          int world, int k, float* __restrict__ centroids, float* __restrict__ cnorm,
                          unsigned long long* __restrict__ total_counts) {
  int j = blockIdx.x;
  int d = threadIdx.x;
  unsigned long long cnt = 0;
  float s = 0.f;
  for (int r = 0; r < world; ++r) {            // fixed rank order: every rank computes identical bits
    s += peer_sums[r][(size_t)j * D + d];
    cnt += peer_counts[r][j];
  }
  float cval = cnt ? s / (float)cnt : centroids[(size_t)j * D + d];
  centroids[(size_t)j * D + d] = cval;
  float sq = warp_sum(cval * cval);
  __shared__ float part[D / 32];
  if ((d & 31) == 0) part[d >> 5] = sq;
  __syncthreads();
  if (d == 0) {
    float t = 0.f;
    for (int w = 0; w < D / 32; ++w) t += part[w];
    cnorm[j] = t;
    if (total_counts) total_counts[j] = cnt;
  }
}

static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess) return nullptr;
    fn = (PFN_cuTensorMapEncodeTiled_v12000)p;
  }
  return fn;
}

static int make_map(CUtensorMap* m, const float* base, uint64_t rows, uint32_t box_rows = BM) {
  auto enc = get_encode();
  if (!enc) return DRYAD_ERR_BAD_ARG;
  cuuint64_t dims[2] = {(cuuint64_t)D, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)D * 4};
  cuuint32_t box[2] = {(cuuint32_t)KB, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? DRYAD_OK : (int)(1000 + r);
}

}}  // namespace dryad::km

extern "C" {
using namespace dryad::km;

int dryad_kmeans_norms(const float* centroids, int k, float* cnorm, void* stream) {
  if (k <= 0) return DRYAD_ERR_BAD_ARG;
  km_norms<<<(k + 7) / 8, 256, 0, (cudaStream_t)stream>>>(centroids, k, cnorm);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

// points [n,128] fp32, centroids [k,128] fp32 (k multiple of 128, <= 2048), cnorm [k].
// Outputs: assign[n] int32; sums[k,128] and counts[k] are ACCUMULATED into (zero them per
// iteration); inertia (optional, 1 float) accumulates sum of squared distances.
static int kmeans_assign_impl(const float* points, int64_t n, const float* centroids, const float* cnorm, int k,
                              int32_t* assign, float* sums, unsigned long long* counts, float* inertia,
                              int max_ctas, const float* sep2, uint32_t* ulist, unsigned int* ucount, void* stream) {
  if (n < 0 || k <= 0 || k % BN != 0 || k > MAX_K) return DRYAD_ERR_BAD_ARG;
  if (n == 0) return DRYAD_OK;
  if (((uintptr_t)points & 15) || ((uintptr_t)centroids & 15) || ((uintptr_t)sums & 15)) return DRYAD_ERR_BAD_ARG;
  const char* penv = getenv("DRYAD_KM_PAIR");
  // default: the CTA-pair kernel (measured 4.07 ms vs 5.39 ms per iteration for 8 M x 128 x 1024)
  const bool pair_mode = (penv ? atoi(penv) != 0 : true) && k % BN2 == 0 && n >= 2 * BM;
  if (pair_mode) {
    CUtensorMap mx, mc;
    int rc = make_map(&mx, points, (uint64_t)n); if (rc) return rc;
    rc = make_map(&mc, centroids, (uint64_t)k, (uint32_t)BN); if (rc) return rc;
    int dev = 0, sms = 148;
    DRYAD_CUDA_TRY(cudaGetDevice(&dev));
    DRYAD_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int64_t ntiles2 = (n + 2 * BM - 1) / (2 * BM);
    int64_t pairs = sms / 2;
    if (max_ctas > 0 && pairs > max_ctas / 2) pairs = max_ctas / 2;
    if (pairs > ntiles2) pairs = ntiles2;
    if (pairs < 1) pairs = 1;
    const char* dbg = getenv("DRYAD_KM_DEBUG");
    const int dbgv = dbg ? atoi(dbg) : 0;
    cudaLaunchConfig_t cfg = {};
    // 8 epilogue warps; a 16-warp variant (64 columns per thread) measured slower (4.3-4.6 ms)
    cfg.gridDim = dim3((unsigned)(2 * pairs)); cfg.blockDim = dim3((unsigned)((2 + 8) * 32));
    cfg.dynamicSmemBytes = SMEM_BYTES2; cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    DRYAD_CUDA_TRY(cudaFuncSetAttribute(km_assign_pair<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES2));
    const char* fenv = getenv("DRYAD_KM_FUSED_CERT");
    const bool fused_cert = sep2 && fenv && atoi(fenv) != 0;      // certificate inside the epilogue (slower, see km_certify)
    DRYAD_CUDA_TRY(cudaLaunchKernelEx(&cfg, km_assign_pair<8>, mx, mc, cnorm, n, k, assign, sums, counts, inertia, dbgv,
                                      centroids, fused_cert ? sep2 : (const float*)nullptr, ulist, ucount));
    DRYAD_LAUNCH_CHECK();
    if (sep2 && !fused_cert) {
      int64_t cg = (n + 31) / 32;
      if (cg > 148 * 16) cg = 148 * 16;
      km_certify<<<(int)cg, 256, 0, (cudaStream_t)stream>>>(points, n, centroids, cnorm, sep2, assign, ulist, ucount);
      DRYAD_LAUNCH_CHECK();
    }
    return DRYAD_OK;
  }
  CUtensorMap mx, mc;
  int rc = make_map(&mx, points, (uint64_t)n); if (rc) return rc;
  rc = make_map(&mc, centroids, (uint64_t)k, (uint32_t)BN); if (rc) return rc;
  CUtensorMap ms;
  rc = make_map(&ms, sums, (uint64_t)k, 1u); if (rc) return rc;
  int dev = 0, sms = 148;
  DRYAD_CUDA_TRY(cudaGetDevice(&dev));
  DRYAD_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  int64_t ntiles = (n + BM - 1) / BM;
  int grid = (int)(ntiles < sms ? ntiles : sms);
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  if (grid == 0) grid = 1;
  const char* dbg = getenv("DRYAD_KM_DEBUG");
  const int dbgv = dbg ? atoi(dbg) : 0;
  DRYAD_CUDA_TRY(cudaFuncSetAttribute(km_assign, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
  km_assign<<<grid, THREADS, SMEM_BYTES, (cudaStream_t)stream>>>(mx, mc, ms, cnorm, n, k, assign, sums, counts, inertia, dbgv);
  DRYAD_LAUNCH_CHECK();
  if (sep2) {                          // this kernel does not certify in its epilogue: one extra pass over the points
    int64_t cg = (n + 31) / 32;
    if (cg > 148 * 16) cg = 148 * 16;
    km_certify<<<(int)cg, 256, 0, (cudaStream_t)stream>>>(points, n, centroids, cnorm, sep2, assign, ulist, ucount);
    DRYAD_LAUNCH_CHECK();
  }
  return DRYAD_OK;
}

int dryad_kmeans_assign(const float* points, int64_t n, const float* centroids, const float* cnorm, int k,
                        int32_t* assign, float* sums, unsigned long long* counts, float* inertia,
                        int max_ctas, void* stream) {
  return kmeans_assign_impl(points, n, centroids, cnorm, k, assign, sums, counts, inertia, max_ctas, nullptr, nullptr, nullptr, stream);
}

int dryad_kmeans_sep(const float* centroids, int k, float* sep2, void* stream) {
  if (k <= 0) return DRYAD_ERR_BAD_ARG;
  km_sep<<<k, 256, 0, (cudaStream_t)stream>>>(centroids, k, sep2);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_kmeans_assign_exact(const float* points, int64_t n, const float* centroids, const float* cnorm, const float* sep2,
                              int k, int32_t* assign, float* sums, unsigned long long* counts, float* inertia,
                              uint32_t* uncertain, unsigned int* uncertain_count, unsigned int* changed, int max_ctas,
                              void* stream) {
  if (!sep2 || !uncertain || !uncertain_count || n >= (1ll << 32)) return DRYAD_ERR_BAD_ARG;
  int rc = kmeans_assign_impl(points, n, centroids, cnorm, k, assign, sums, counts, inertia, max_ctas, sep2, uncertain,
                              uncertain_count, stream);
  if (rc) return rc;
  if (n == 0) return DRYAD_OK;
  km_recheck<<<148 * 4, 256, 0, (cudaStream_t)stream>>>(points, centroids, cnorm, k, assign, sums, counts, uncertain,
                                                       uncertain_count, changed);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_kmeans_update(float* const* peer_sums, unsigned long long* const* peer_counts, int world, int k,
                        float* centroids, float* cnorm, unsigned long long* total_counts, void* stream) {
  if (k <= 0 || world < 1 || world > DRYAD_MAX_PEERS) return DRYAD_ERR_BAD_ARG;
  km_update<<<k, D, 0, (cudaStream_t)stream>>>(peer_sums, peer_counts, world, k, centroids, cnorm, total_counts);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"
