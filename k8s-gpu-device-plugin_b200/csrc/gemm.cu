// gemm.cu — dense bf16 GEMM probe on the 5th-generation tensor cores (K3 of SURVEY.md §8).
//
//   C[i][j] = SUM_k A[i][k] * B[j][k]     A: M x K, B: N x K, both K-major bf16; C fp32 row-major
//
// No reference counterpart (the reference never touches the GPU: device/devices.go:74).
// Algorithmic work per launch: 2 M N K flop, bound by the tensor pipe.
//
// Kernel shape (sm_100a only):
//   * persistent, one CTA per SM; with kCtaGroup == 2 the CTAs of a 2-CTA cluster (one TPC)
//     share each 256 x 256 output tile: tcgen05.mma.cta_group::2 with UMMA M = 256, each CTA
//     staging its own 128 rows of A and its own 128 rows of B (half the operand traffic per
//     SM), accumulators in each CTA's tensor memory;
//   * warp 0: TMA producer (cp.async.bulk.tensor.2d, 128-byte swizzle) into a ring of
//     kStages shared-memory stages, completion counted on mbarriers;
//   * warp 1: allocates the 512 TMEM columns, and its elected lane issues tcgen05.mma
//     (kind::f16, bf16 x bf16 -> fp32), releasing stages with tcgen05.commit;
//   * warps 2-5: epilogue, tcgen05.ld 32 lanes x 32 columns at a time -> registers ->
//     256-bit st.global (one full L2 sector per lane); two accumulator stages (2 x 256 TMEM columns) so the epilogue of tile t
//     overlaps the main loop of tile t+1.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "common.cuh"
#include "engine.h"

namespace gp {

// ---------------------------------------------------------------- tiling ---
constexpr int kBlockM = 128;  // rows of A (and of the accumulator) per CTA
constexpr int kBlockK = 64;   // 64 bf16 = 128 bytes = one swizzle span
constexpr int kUmmaK = 16;    // K per tcgen05.mma for 16-bit inputs
constexpr int kTileN = 256;   // accumulator columns per tile (UMMA N)
constexpr int kGemmThreads = 192;

// CG = CTAs per MMA (tcgen05 cta_group), NT = 256-wide N sub-tiles accumulated per tile.
// NT = 2 keeps BOTH 256-column TMEM accumulators busy with one 256 x 512 tile that shares every
// A stage between two MMAs, so a stage moves A + 2B for twice the flops: 25 % fewer operand
// bytes per flop from L2 (the measured bottleneck) at the price of an
// epilogue that no longer overlaps the main loop.
template <int CG, int NT>
struct GemmCfg {
  static_assert(NT == 1 || (NT == 2 && CG == 2), "two N sub-tiles need the 2-CTA stage budget");
  static constexpr int kLoadN = kTileN / CG;  // rows of each B sub-tile staged by each CTA
  static constexpr uint32_t kABytes = kBlockM * kBlockK * 2;
  static constexpr uint32_t kBBytes = kLoadN * kBlockK * 2;
  static constexpr uint32_t kStageBytes = kABytes + NT * kBBytes;
  static constexpr int kStages = (192 * 1024) / kStageBytes;  // 6 x 32 KiB, 4 x 48 KiB
  static constexpr uint32_t kTxBytes = kStageBytes * CG;  // bytes landing per stage on the leader's barrier
  static constexpr uint32_t kEpiBytes = 2 * 4 * 4096;  // 2 buffers x 4 epilogue warps x (32 rows x 128 B)
  static constexpr size_t kSmemBytes = (size_t)kStages * kStageBytes + kEpiBytes + 1024 /*align*/ + 256 /*barriers*/;
  static constexpr int kTileM = kBlockM * CG;
  static constexpr int kTileNAll = kTileN * NT;   // output columns per tile
  static constexpr int kAccStages = 2 / NT;       // TMEM: 512 columns = kAccStages x kTileNAll
};

// ------------------------------------------------------------ PTX helpers ---
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred;
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* p) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(p) : "memory");
}
template <int CG>
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const void* map, uint32_t mbar,
                                            int c0, int c1, uint64_t policy) {
  if constexpr (CG == 1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;" ::
            "r"(smem_dst), "l"(map), "r"(mbar), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
  } else {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;" ::
            "r"(smem_dst), "l"(map), "r"(mbar), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
  }
}
__device__ __forceinline__ uint64_t l2_policy(int kind) {  // 0 normal, 1 evict_last, 2 evict_first
  uint64_t p;
  if (kind == 1) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  else if (kind == 2) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// 32 bytes (one full L2 sector) per lane per store: STG.E.256
__device__ __forceinline__ void st_f8_hint(float* p, const float* v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8}, %9;" ::"l"(p),
               "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <int CG>
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
  if constexpr (CG == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
}
template <int CG>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  if constexpr (CG == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
  else
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread for the CTA (pair)
template <int CG>
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                         uint32_t idesc, uint32_t accumulate) {
  if constexpr (CG == 1) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed
template <int CG>
__device__ __forceinline__ void umma_commit(uint32_t mbar) {
  if constexpr (CG == 1) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar) : "memory");
  } else {  // same offset in both CTAs of the pair
    const uint16_t mask = 0x3;
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(mbar),
        "h"(mask)
        : "memory");
  }
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (lane t owns row t)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor for a K-major tile whose rows are 128 bytes (64 bf16) stored
// with the 128-byte swizzle TMA wrote: 8-row groups are 1024 bytes apart (SBO), the leading
// dimension offset is unused for swizzled K-major, descriptor version 1 (sm_100), layout
// type 2 = SWIZZLE_128B.  Advancing K by 16 elements inside the span adds 32 bytes to the
// start address.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);  // start address, 16-byte units
  d |= (uint64_t)1 << 16;                      // LBO (ignored for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;            // SBO = 1024 bytes between 8-row groups
  d |= (uint64_t)1 << 46;                      // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                      // SWIZZLE_128B
  return d;
}
// Instruction descriptor, kind::f16: D = fp32 (1 @ bit 4), A = B = bf16 (1 @ bits 7, 10),
// both K-major (bits 15, 16 = 0), N >> 3 @ bits 17-22, M >> 4 @ bits 24-28.
__host__ __device__ constexpr uint32_t make_idesc(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// ---------------------------------------------------------------- kernel ---
// OUT = 0: C is fp32; OUT = 1: C is bf16 (round-to-nearest-even of the fp32 accumulator)
template <int CG, int NT, int OUT>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_c, float* __restrict__ C, int M, int N, int K,
                 int group_m, int hints) {
  using Cfg = GemmCfg<CG, NT>;
  constexpr int kAccStages = Cfg::kAccStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // swizzle-128B needs 1024-byte alignment
  const uint32_t epi_base = smem_base + Cfg::kStages * Cfg::kStageBytes;  // 1024-aligned: stage bytes are multiples of 1024
  const uint32_t bar_base = epi_base + Cfg::kEpiBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * Cfg::kStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * Cfg::kStages + kAccStages + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * Cfg::kStages + 2 * kAccStages);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = CG == 2 ? cluster_ctarank() : 0;
  const bool leader = cta_rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_c);
    for (int s = 0; s < Cfg::kStages; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(full_bar(s)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(empty_bar(s)));
    }
    for (int a = 0; a < kAccStages; ++a) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tfull_bar(a)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tempty_bar(a)), "r"(4 * CG));
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<CG>(tmem_slot, 512);
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int tiles_m = M / Cfg::kTileM, tiles_n = N / Cfg::kTileNAll;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = K / kBlockK;
  const int unit = blockIdx.x / CG, num_units = gridDim.x / CG;
  // raster: groups of kGroupM tile rows, columns fastest inside a group, so that the ~148 tiles
  // in flight share A and B panels that fit in L2 together
  auto tile_coords = [&](int t, int& tm, int& tn) {
    const int per_group = group_m * tiles_n;
    const int g = t / per_group, r = t % per_group;
    const int rows = min(group_m, tiles_m - g * group_m);
    tm = g * group_m + r % rows;
    tn = r / rows;
  };

  if (warp == 0) {
    // ============================ TMA producer ============================
    if (elect_one()) {
      int stage = 0; uint32_t phase = 0;
      const uint32_t full_leader = CG == 2 ? map_to_cta(full_bar(0), 0) : full_bar(0);
      const uint64_t pol_a = l2_policy(hints & 3), pol_b = l2_policy((hints >> 2) & 3);
      for (int t = unit; t < num_tiles; t += num_units) {
        int tm, tn;
        tile_coords(t, tm, tn);
        const int row_a = tm * Cfg::kTileM + (int)cta_rank * kBlockM;
        const int row_b = tn * Cfg::kTileNAll + (int)cta_rank * Cfg::kLoadN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait_addr(empty_bar(stage), phase ^ 1);
          const uint32_t sa = smem_base + stage * Cfg::kStageBytes;
          const uint32_t sb = sa + Cfg::kABytes;
          const uint32_t fb = full_leader + 8u * stage;
          if (leader)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full_bar(stage)),
                         "r"(Cfg::kTxBytes) : "memory");
          tma_load_2d<CG>(sa, &map_a, fb, kb * kBlockK, row_a, pol_a);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt)
            tma_load_2d<CG>(sb + nt * Cfg::kBBytes, &map_b, fb, kb * kBlockK, row_b + nt * kTileN, pol_b);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ============================= MMA issuer =============================
    if (leader && elect_one()) {
      constexpr uint32_t idesc = make_idesc(Cfg::kTileM, kTileN);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int t = unit; t < num_tiles; t += num_units) {
        mbar_wait_addr(tempty_bar(acc), acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * Cfg::kTileNAll);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait_addr(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * Cfg::kStageBytes;
          const uint64_t adesc = make_smem_desc(sa);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
              // +32 bytes per K step of 16 bf16 -> +2 in the 16-byte-unit address field
              const uint64_t bdesc = make_smem_desc(sa + Cfg::kABytes + nt * Cfg::kBBytes);
              umma_f16<CG>(d_tmem + (uint32_t)(nt * kTileN), adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k),
                           idesc, (uint32_t)((kb | k) != 0));
            }
          }
          umma_commit<CG>(empty_bar(stage));  // frees the stage once these MMAs have read it
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit<CG>(tfull_bar(acc));      // accumulator complete -> epilogue
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ============================== epilogue ==============================
    const int quarter = warp & 3;  // a warp may only touch TMEM lanes [32 q, 32 q + 32)
    int acc = 0; uint32_t acc_phase = 0;
    const uint32_t tempty_leader = CG == 2 ? map_to_cta(tempty_bar(0), 0) : tempty_bar(0);
    const uint64_t pol_c = l2_policy((hints >> 4) & 3);
    uint32_t epi_count = 0;
    for (int t = unit; t < num_tiles; t += num_units) {
      int tm, tn;
      tile_coords(t, tm, tn);
      mbar_wait_addr(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int row = tm * Cfg::kTileM + (int)cta_rank * kBlockM + quarter * 32 + lane;
      float* crow = C + (size_t)row * N + (size_t)tn * Cfg::kTileNAll;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * Cfg::kTileNAll);
      if constexpr (OUT == 1) {
        // bf16 output: 64 accumulator columns -> one 128-byte staging row per lane -> TMA store of a
        // 32 x 64 bf16 box; same swizzle and double buffering as the fp32 path below
        const int ewarp = warp - 2;
        const int tile_row0 = tm * Cfg::kTileM + (int)cta_rank * kBlockM + quarter * 32;
#pragma unroll 1
        for (int c = 0; c < Cfg::kTileNAll / 64; ++c) {
          float v0[32], v1[32];
          tmem_ld_32x32(taddr + (uint32_t)(c * 64), v0);
          tmem_ld_32x32(taddr + (uint32_t)(c * 64 + 32), v1);
          const uint32_t buf = epi_base + (uint32_t)(((epi_count & 1) * 4 + ewarp) * 4096);
          if (lane == 0) bulk_wait_read<1>();
          __syncwarp();
          tmem_ld_wait();
          uint32_t w[32];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w[i]) : "f"(v0[2 * i + 1]), "f"(v0[2 * i]));
            asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w[16 + i]) : "f"(v1[2 * i + 1]), "f"(v1[2 * i]));
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t addr = buf + (uint32_t)(lane * 128 + ((j ^ (lane & 7)) << 4));
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(w[4 * j]), "r"(w[4 * j + 1]),
                         "r"(w[4 * j + 2]), "r"(w[4 * j + 3]) : "memory");
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            asm volatile(
                "cp.async.bulk.tensor.2d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%2, %3}], [%1], %4;" ::"l"(&map_c),
                "r"(buf), "r"(tn * Cfg::kTileNAll + c * 64), "r"(tile_row0), "l"(pol_c)
                : "memory");
            bulk_commit();
          }
          ++epi_count;
        }
      } else if (hints & 128) {  // development path: direct 256-bit stores from registers
#pragma unroll 1
        for (int c = 0; c < Cfg::kTileNAll / 32; ++c) {
          float v[32];
          tmem_ld_32x32(taddr + (uint32_t)(c * 32), v);
          tmem_ld_wait();
          if (!(hints & 64) || v[0] == 123.456f)
#pragma unroll
            for (int q = 0; q < 4; ++q) st_f8_hint(crow + c * 32 + q * 8, v + 8 * q, pol_c);
        }
      } else {
        // registers -> this warp's 32 x 32 fp32 staging tile (128-byte rows, 16-byte chunks XOR-swizzled
        // by row & 7 exactly as the tensor map's SWIZZLE_128B expects, which also makes the st.shared.v4
        // conflict-free) -> one TMA store per chunk; two staging buffers per warp
        const int ewarp = warp - 2;
        const int tile_row0 = tm * Cfg::kTileM + (int)cta_rank * kBlockM + quarter * 32;
#pragma unroll 1
        for (int c = 0; c < Cfg::kTileNAll / 32; ++c) {
          float v[32];
          tmem_ld_32x32(taddr + (uint32_t)(c * 32), v);
          const uint32_t buf = epi_base + (uint32_t)(((epi_count & 1) * 4 + ewarp) * 4096);
          if (lane == 0) bulk_wait_read<1>();  // the store that last used this buffer has read it
          __syncwarp();
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t addr = buf + (uint32_t)(lane * 128 + ((j ^ (lane & 7)) << 4));
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v[4 * j]), "f"(v[4 * j + 1]),
                         "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            asm volatile(
                "cp.async.bulk.tensor.2d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%2, %3}], [%1], %4;" ::"l"(&map_c),
                "r"(buf), "r"(tn * Cfg::kTileNAll + c * 32), "r"(tile_row0), "l"(pol_c)
                : "memory");
            bulk_commit();
          }
          ++epi_count;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 2) mbar_arrive_cluster(tempty_leader + 8u * acc);
        else asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty_bar(acc)) : "memory");
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) bulk_wait_all<0>();  // this warp's TMA stores have completed
  }

  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) tmem_dealloc<CG>(tmem_base, 512);
}

// --------------------------------------------------- check / helper kernels --
__device__ __forceinline__ float c_load(const float* C, size_t i) { return C[i]; }
__device__ __forceinline__ float c_load(const uint16_t* C, size_t i) { return bf16_bits_to_f32(C[i]); }
// gather sampled entries of C
template <class T>
__global__ void gather_kernel(const T* __restrict__ C, int N, const uint32_t* __restrict__ ij,
                              int n, float* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) out[t] = c_load(C, (size_t)ij[2 * t] * N + ij[2 * t + 1]);
}
// Freivalds in fp64 on the CUDA cores: x_j = +-1 from the generator.
__global__ void fv_signs_kernel(float* __restrict__ x, int N, uint64_t seed) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < N) x[j] = (mix64(seed ^ 0x7777, (uint64_t)j) & 1) ? 1.0f : -1.0f;
}
// u[k] += SUM_j B[j][k] x[j] ; ub[k] += SUM_j |B[j][k]|   (thread per k, blockIdx.y splits j; u, ub zeroed)
__global__ void fv_bt_x_kernel(const uint16_t* __restrict__ B, int N, int K, const float* __restrict__ x,
                               double* __restrict__ u, double* __restrict__ ub) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const int per = (N + gridDim.y - 1) / gridDim.y;
  const int j0 = blockIdx.y * per, j1 = min(N, j0 + per);
  double s = 0, a = 0;
  for (int j = j0; j < j1; ++j) {
    const double b = (double)bf16_bits_to_f32(B[(size_t)j * K + k]);
    s += b * (double)x[j];
    a += fabs(b);
  }
  atomicAdd(&u[k], s);
  atomicAdd(&ub[k], a);
}
// v[i] = SUM_k A[i][k] u[k] ; sc[i] = SUM_k |A[i][k]| ub[k] ; w[i] = SUM_j C[i][j] x[j] ;
// cabs[i] = SUM_j |C[i][j]|   (one warp per row)
template <class T>
__global__ void fv_rows_kernel(const uint16_t* __restrict__ A, const T* __restrict__ C, int M, int N,
                               int K, const float* __restrict__ x, const double* __restrict__ u,
                               const double* __restrict__ ub, double* __restrict__ v,
                               double* __restrict__ sc, double* __restrict__ w, double* __restrict__ cabs) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= M) return;
  double sv = 0, ss = 0, sw = 0, sa = 0;
  for (int k = lane; k < K; k += 32) {
    const double a = (double)bf16_bits_to_f32(A[(size_t)row * K + k]);
    sv += a * u[k];
    ss += fabs(a) * ub[k];
  }
  for (int j = lane; j < N; j += 32) {
    const double cv = (double)c_load(C, (size_t)row * N + j);
    sw += cv * (double)x[j];
    sa += fabs(cv);
  }
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) {
    sv += __shfl_xor_sync(0xffffffffu, sv, m);
    ss += __shfl_xor_sync(0xffffffffu, ss, m);
    sw += __shfl_xor_sync(0xffffffffu, sw, m);
    sa += __shfl_xor_sync(0xffffffffu, sa, m);
  }
  if (lane == 0) { v[row] = sv; sc[row] = ss; w[row] = sw; cabs[row] = sa; }
}

// ------------------------------------------------------------------ host ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// K-major bf16 matrix [rows][K] -> 2-D map with a (64 x box_rows) box and 128-byte swizzle
static int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t K, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled not available from the driver"); return GP_ERR_UNSUPPORTED; }
  const cuuint64_t dims[2] = {K, rows};
  const cuuint64_t strides[1] = {K * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kBlockK, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box,
                         estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed: %d", (int)r); return GP_ERR_CUDA; }
  return GP_OK;
}

static int gemm_cta_group() {
  const char* e = getenv("GPUPROBE_GEMM_CTA_GROUP");  // development switch; default is the 2-CTA kernel
  if (e && e[0] == '1') return 1;
  return 2;
}

template <int CG, int NT, int OUT>
static cudaError_t launch_gemm_cg(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, float* C,
                                  int M, int N, int K, int sm_count, cudaStream_t stream) {
  using Cfg = GemmCfg<CG, NT>;
  const int tiles = (M / Cfg::kTileM) * (N / Cfg::kTileNAll);
  int units = std::min(tiles, sm_count / CG);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(units * CG));
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  // raster group height in tile rows (development override: GPUPROBE_GEMM_GROUP_M)
  static const int env_group = getenv("GPUPROBE_GEMM_GROUP_M") ? atoi(getenv("GPUPROBE_GEMM_GROUP_M")) : 0;
  // measured at 8192^3: 256x256 tiles 4: 1430, 8: 1475, 16: 1458; 256x512 tiles 4: 1493, 8: 1529, 16: 1540, 32: 1519
  const int group_m = env_group > 0 ? env_group : (NT == 2 ? 16 : 16 / CG);
  // L2 policies, 2 bits each for A | B << 2 | C << 4 (0 normal, 1 evict_last, 2 evict_first)
  static const int env_hints = getenv("GPUPROBE_GEMM_HINTS") ? atoi(getenv("GPUPROBE_GEMM_HINTS")) : -1;
  // measured at 8192^3, 20 launches each: operands evict_last + C evict_first 1462 TFLOP/s median,
  // no hints 1438; operand panels are re-read by the next wave, C is written once and never read
  const int hints = env_hints >= 0 ? env_hints : (1 | (1 << 2) | (2 << 4));
  return cudaLaunchKernelEx(&cfg, gemm_bf16_kernel<CG, NT, OUT>, ma, mb, mc, C, M, N, K, group_m, hints);
}

// function attributes are per device: called once from device_prepare with the device current
template <int CG, int NT, int OUT>
static cudaError_t configure_one() {
  return cudaFuncSetAttribute(gemm_bf16_kernel<CG, NT, OUT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)GemmCfg<CG, NT>::kSmemBytes);
}
cudaError_t gemm_kernels_configure() {
  cudaError_t e;
  if ((e = configure_one<1, 1, 0>()) != cudaSuccess) return e;
  if ((e = configure_one<2, 1, 0>()) != cudaSuccess) return e;
  if ((e = configure_one<2, 2, 0>()) != cudaSuccess) return e;
  if ((e = configure_one<1, 1, 1>()) != cudaSuccess) return e;
  if ((e = configure_one<2, 1, 1>()) != cudaSuccess) return e;
  return configure_one<2, 2, 1>();
}

struct GemmPlan {
  int cg, nt, out;
  CUtensorMap ma, mb, mc;
};

// C (row-major M x N) as a 2-D map with 128-byte box rows (32 fp32 or 64 bf16) x 32 rows, 128-byte swizzle
static int make_map_c(CUtensorMap* map, const void* base, uint64_t M, uint64_t N, bool out_bf16) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled not available from the driver"); return GP_ERR_UNSUPPORTED; }
  const cuuint64_t dims[2] = {N, M};
  const cuuint64_t strides[1] = {N * (out_bf16 ? 2u : 4u)};
  const cuuint32_t box[2] = {out_bf16 ? 64u : 32u, 32};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, out_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                         const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (C) failed: %d", (int)r); return GP_ERR_CUDA; }
  return GP_OK;
}

static int make_plan(GemmPlan* p, const void* A, const void* B, const void* C, uint32_t m, uint32_t n, uint32_t k,
                     bool out_bf16) {
  p->out = out_bf16 ? 1 : 0;
  p->cg = gemm_cta_group();
  // 256 x 512 tiles (two N sub-tiles per CTA pair) whenever the shape allows; GPUPROBE_GEMM_NT=1 forces 256 x 256
  static const int env_nt = getenv("GPUPROBE_GEMM_NT") ? atoi(getenv("GPUPROBE_GEMM_NT")) : 0;
  p->nt = (p->cg == 2 && n % 512 == 0 && env_nt != 1) ? 2 : 1;
  if (m % 256 || n % 256 || k % 64 || m == 0 || n == 0 || k == 0) {
    set_error("gemm needs m, n multiples of 256 and k a multiple of 64");
    return GP_ERR_INVALID;
  }
  int rc = make_map(&p->ma, A, m, k, kBlockM);
  if (rc) return rc;
  rc = make_map_c(&p->mc, C, m, n, out_bf16);
  if (rc) return rc;
  return make_map(&p->mb, B, n, k, (uint32_t)(kTileN / p->cg));
}

static cudaError_t launch_gemm(const GemmPlan& p, float* C, int M, int N, int K, int sm_count,
                               cudaStream_t stream) {
  if (p.out) {
    if (p.cg == 2 && p.nt == 2) return launch_gemm_cg<2, 2, 1>(p.ma, p.mb, p.mc, C, M, N, K, sm_count, stream);
    return p.cg == 2 ? launch_gemm_cg<2, 1, 1>(p.ma, p.mb, p.mc, C, M, N, K, sm_count, stream)
                     : launch_gemm_cg<1, 1, 1>(p.ma, p.mb, p.mc, C, M, N, K, sm_count, stream);
  }
  if (p.cg == 2 && p.nt == 2) return launch_gemm_cg<2, 2, 0>(p.ma, p.mb, p.mc, C, M, N, K, sm_count, stream);
  return p.cg == 2 ? launch_gemm_cg<2, 1, 0>(p.ma, p.mb, p.mc, C, M, N, K, sm_count, stream)
                   : launch_gemm_cg<1, 1, 0>(p.ma, p.mb, p.mc, C, M, N, K, sm_count, stream);
}

#define CU_OR_RETURN(call)                                               \
  do {                                                                   \
    cudaError_t e__ = (call);                                            \
    if (e__ != cudaSuccess) return ::gp::cuda_fail(e__, #call, nullptr); \
  } while (0)

static int check_device(Device* d) {
  if (d->cc_major != 10) {
    set_error("gemm probe needs an sm_100 device (tcgen05/TMEM); this one is sm_%d%d", d->cc_major, d->cc_minor);
    return GP_ERR_UNSUPPORTED;
  }
  return GP_OK;
}

int gemm_host_locked(Ctx* ctx, Device* d, const uint16_t* a, const uint16_t* b, float* c, uint32_t m,
                     uint32_t n, uint32_t k) {
  int rc = check_device(d);
  if (rc) return rc;
  const size_t a_bytes = (size_t)m * k * 2, b_bytes = (size_t)n * k * 2, c_bytes = (size_t)m * n * 4;
  auto up = [](size_t x) { return (x + 1023) & ~(size_t)1023; };
  rc = ensure_arena(ctx, d, up(a_bytes) + up(b_bytes) + up(c_bytes));
  if (rc) return rc;
  d->triad_operands_bytes = 0;
  uint8_t* A = d->arena;
  uint8_t* B = A + up(a_bytes);
  float* C = reinterpret_cast<float*>(B + up(b_bytes));
  GemmPlan plan;
  rc = make_plan(&plan, A, B, C, m, n, k, /*out_bf16=*/false);
  if (rc) return rc;
  CU_OR_RETURN(cudaMemcpyAsync(A, a, a_bytes, cudaMemcpyHostToDevice, d->stream));
  CU_OR_RETURN(cudaMemcpyAsync(B, b, b_bytes, cudaMemcpyHostToDevice, d->stream));
  CU_OR_RETURN(cudaMemsetAsync(C, 0xFF, c_bytes, d->stream));
  CU_OR_RETURN(launch_gemm(plan, C, (int)m, (int)n, (int)k, d->sm_count, d->stream));
  CU_OR_RETURN(cudaMemcpyAsync(c, C, c_bytes, cudaMemcpyDeviceToHost, d->stream));
  CU_OR_RETURN(cudaStreamSynchronize(d->stream));
  d->stats[GP_K_GEMM].launches += 1;
  return GP_OK;
}

int gemm_probe_locked(Ctx* ctx, Device* d, uint32_t mnk, uint64_t seed, uint32_t warmup, uint32_t iters,
                      uint32_t samples, gp_gemm_result* out) {
  memset(out, 0, sizeof(*out));
  out->struct_size = (uint32_t)sizeof(*out);
  int rc = check_device(d);
  if (rc) return rc;
  const uint32_t m = mnk, n = mnk, k = mnk;
  out->m = m; out->n = n; out->k = k;
  if (iters + 1 > (uint32_t)kEventPool) return GP_ERR_TOO_LARGE;
  if (samples > 65536) samples = 65536;
  const bool out_bf16 = ctx->cfg.gemm_output == 1;
  out->output_bf16 = out_bf16 ? 1 : 0;
  const size_t a_bytes = (size_t)m * k * 2, b_bytes = (size_t)n * k * 2, c_bytes = (size_t)m * n * (out_bf16 ? 2 : 4);
  auto up = [](size_t x) { return (x + 1023) & ~(size_t)1023; };
  const size_t vec_bytes = up((size_t)std::max(std::max(m, n), k) * sizeof(double));
  const size_t aux = 7 * vec_bytes + up((size_t)samples * 8) + up((size_t)samples * 4);
  rc = ensure_arena(ctx, d, up(a_bytes) + up(b_bytes) + up(c_bytes) + aux);
  if (rc) return rc;
  d->triad_operands_bytes = 0;
  uint8_t* p = d->arena;
  uint16_t* A = reinterpret_cast<uint16_t*>(p); p += up(a_bytes);
  uint16_t* B = reinterpret_cast<uint16_t*>(p); p += up(b_bytes);
  float* C = reinterpret_cast<float*>(p); p += up(c_bytes);
  double* u = reinterpret_cast<double*>(p); p += vec_bytes;
  double* ub = reinterpret_cast<double*>(p); p += vec_bytes;
  double* v = reinterpret_cast<double*>(p); p += vec_bytes;
  double* sc = reinterpret_cast<double*>(p); p += vec_bytes;
  double* w = reinterpret_cast<double*>(p); p += vec_bytes;
  double* cabs = reinterpret_cast<double*>(p); p += vec_bytes;
  float* xs = reinterpret_cast<float*>(p); p += vec_bytes;
  uint32_t* d_ij = reinterpret_cast<uint32_t*>(p); p += up((size_t)samples * 8);
  float* d_sampled = reinterpret_cast<float*>(p);

  GemmPlan plan;
  rc = make_plan(&plan, A, B, C, m, n, k, out_bf16);
  if (rc) return rc;
  cudaStream_t s = d->stream;
  CU_OR_RETURN(launch_gen_bf16(A, (uint64_t)m * k, seed, d->sm_count, s));
  CU_OR_RETURN(launch_gen_bf16(B, (uint64_t)n * k, seed ^ kSeedBXor, d->sm_count, s));
  CU_OR_RETURN(cudaMemsetAsync(C, 0xFF, c_bytes, s));
  d->stats[GP_K_INIT].launches += 2;
  for (uint32_t i = 0; i < warmup; ++i) CU_OR_RETURN(launch_gemm(plan, C, m, n, k, d->sm_count, s));
  for (uint32_t i = 0; i < iters; ++i) {
    CU_OR_RETURN(cudaEventRecord(d->events[i], s));
    CU_OR_RETURN(launch_gemm(plan, C, m, n, k, d->sm_count, s));
  }
  CU_OR_RETURN(cudaEventRecord(d->events[iters], s));

  // ---- checks: stratified samples (every 256x256 tile gets its share) + Freivalds ----
  std::vector<uint32_t> ij(2 * (size_t)samples);
  const uint32_t tm = m / 256, tn = n / 256, ntile = tm * tn;
  for (uint32_t t = 0; t < samples; ++t) {
    const uint64_t r = mix64(seed ^ 0xC0FFEE, t);
    const uint32_t tile = t % ntile;
    ij[2 * t] = (tile / tn) * 256 + (uint32_t)(r & 255);
    ij[2 * t + 1] = (tile % tn) * 256 + (uint32_t)((r >> 8) & 255);
  }
  std::vector<float> got(samples);
  if (samples) {
    CU_OR_RETURN(cudaMemcpyAsync(d_ij, ij.data(), ij.size() * 4, cudaMemcpyHostToDevice, s));
    if (out_bf16)
      gather_kernel<uint16_t><<<(samples + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint16_t*>(C), (int)n, d_ij, (int)samples, d_sampled);
    else
      gather_kernel<float><<<(samples + 255) / 256, 256, 0, s>>>(C, (int)n, d_ij, (int)samples, d_sampled);
    CU_OR_RETURN(cudaGetLastError());
    CU_OR_RETURN(cudaMemcpyAsync(got.data(), d_sampled, (size_t)samples * 4, cudaMemcpyDeviceToHost, s));
  }
  fv_signs_kernel<<<(n + 255) / 256, 256, 0, s>>>(xs, (int)n, seed);
  CU_OR_RETURN(cudaMemsetAsync(u, 0, 2 * vec_bytes, s));  // u and ub are adjacent
  fv_bt_x_kernel<<<dim3((k + 127) / 128, 32), 128, 0, s>>>(B, (int)n, (int)k, xs, u, ub);
  CU_OR_RETURN(cudaGetLastError());
  if (out_bf16)
    fv_rows_kernel<uint16_t><<<(m * 32 + 255) / 256, 256, 0, s>>>(A, reinterpret_cast<const uint16_t*>(C), (int)m, (int)n, (int)k, xs, u, ub, v, sc, w, cabs);
  else
    fv_rows_kernel<float><<<(m * 32 + 255) / 256, 256, 0, s>>>(A, C, (int)m, (int)n, (int)k, xs, u, ub, v, sc, w, cabs);
  CU_OR_RETURN(cudaGetLastError());
  std::vector<double> hv(m), hs(m), hw(m), hc(m);
  CU_OR_RETURN(cudaMemcpyAsync(hc.data(), cabs, (size_t)m * 8, cudaMemcpyDeviceToHost, s));
  CU_OR_RETURN(cudaMemcpyAsync(hv.data(), v, (size_t)m * 8, cudaMemcpyDeviceToHost, s));
  CU_OR_RETURN(cudaMemcpyAsync(hs.data(), sc, (size_t)m * 8, cudaMemcpyDeviceToHost, s));
  CU_OR_RETURN(cudaMemcpyAsync(hw.data(), w, (size_t)m * 8, cudaMemcpyDeviceToHost, s));
  CU_OR_RETURN(launch_pattern_verify(reinterpret_cast<uint64_t*>(C), c_bytes / 8, kPatNone, 0, d->scratch,
                                     d->d_results, d->sm_count, s));
  CU_OR_RETURN(cudaMemcpyAsync(d->h_results, d->d_results, sizeof(gp_pattern_result), cudaMemcpyDeviceToHost, s));
  CU_OR_RETURN(cudaStreamSynchronize(s));

  std::vector<float> ms;
  for (uint32_t i = 0; i < iters; ++i) {
    float t = 0;
    CU_OR_RETURN(cudaEventElapsedTime(&t, d->events[i], d->events[i + 1]));
    ms.push_back(t);
  }
  summarize(ms, &out->t);
  const double flop = 2.0 * m * n * k;
  for (float t : ms) {
    d->stats[GP_K_GEMM].launches += 1;
    d->stats[GP_K_GEMM].total_ms += t;
    d->stats[GP_K_GEMM].algo_bytes += flop;
  }
  out->tflops_best = out->t.best_ms > 0 ? flop / (out->t.best_ms * 1e-3) / 1e12 : 0;
  out->tflops_median = out->t.median_ms > 0 ? flop / (out->t.median_ms * 1e-3) / 1e12 : 0;
  out->c_checksum = d->h_results[0].xor_;

  // host reference of the sampled entries: fp64 dot products of the bf16-rounded generator output
  const double tol = ctx->cfg.gemm_tol > 0 ? ctx->cfg.gemm_tol : 1e-3;
  const double out_round = out_bf16 ? 1.0 / 256.0 : 0.0;
  std::vector<double> err(samples, 0.0);
  const unsigned nthreads = std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
  std::vector<std::thread> th;
  for (unsigned w0 = 0; w0 < nthreads; ++w0) {
    th.emplace_back([&, w0] {
      for (uint32_t t = w0; t < samples; t += nthreads) {
        const uint64_t i = ij[2 * t], j = ij[2 * t + 1];
        double acc = 0, aacc = 0;
        for (uint32_t kk = 0; kk < k; ++kk) {
          const double x = (double)bf16_bits_to_f32(f32_to_bf16_rne(gen_f32(seed, i * k + kk)));
          const double y = (double)bf16_bits_to_f32(f32_to_bf16_rne(gen_f32(seed ^ kSeedBXor, j * k + kk)));
          acc += x * y;
          aacc += fabs(x * y);
        }
        // a bf16 result may differ from the exact value by its own rounding, 2^-8 |ref| at most
        const double e = std::max(0.0, fabs((double)got[t] - acc) - out_round * fabs(acc));
        err[t] = (std::isfinite((double)got[t]) && aacc > 0) ? e / aacc : (e == 0 ? 0.0 : 1e30);
      }
    });
  }
  for (auto& t : th) t.join();
  out->samples = samples;
  for (uint32_t t = 0; t < samples; ++t) {
    out->max_scaled_err = std::max(out->max_scaled_err, err[t]);
    if (!(err[t] <= tol)) out->samples_bad++;
  }
  for (uint32_t i = 0; i < m; ++i) {
    const double e = std::max(0.0, fabs(hw[i] - hv[i]) - out_round * hc[i]);
    const double r = (std::isfinite(hw[i]) && hs[i] > 0) ? e / hs[i] : (e == 0 ? 0.0 : 1e30);
    out->freivalds_err = std::max(out->freivalds_err, r);
  }
  if (!(out->freivalds_err <= tol)) out->samples_bad++;
  return GP_OK;
}

}  // namespace gp
