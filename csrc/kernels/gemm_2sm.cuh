// K10 on CTA PAIRS: the same bf16 GEMM as gemm_bf16_tcgen05_kernel, issued as tcgen05.mma.cta_group::2.
//
//   C[M,N] = act( op(A) . op(B)^T + bias[N] )
//
// Two CTAs of one cluster (two SMs of a TPC) own one 256 x BN output tile.  Each CTA loads ITS 128 rows of A and ITS
// BN/2 rows of B; one thread of the even CTA issues M = 256, N = BN instructions that read A from each CTA's own shared
// memory and the two halves of B from both, and accumulate 128 rows x BN columns into each CTA's own TMEM.  Against the
// single-CTA kernel this halves the B bytes every SM stages per flop (a 32 KB stage instead of 48 KB for BN = 256: six
// stages instead of four, and a third less shared-memory write traffic and power) and halves the number of MMA issues.
// No operand is multicast: every TMA load is a plain load into the issuing CTA's own shared memory, only its
// completion is signalled on the even CTA's mbarrier (.cta_group::2).
//
// Barriers (same offsets in both CTAs):
//   full[s]       even CTA only, count 1: the even CTA's producer arms it with the bytes of BOTH CTAs' stage
//   empty[s]      both CTAs, count 1: tcgen05.commit multicast to the pair once the MMAs have read stage s
//   tmem_full[b]  both CTAs, count 1: tcgen05.commit multicast when accumulator b is complete
//   tmem_empty[b] even CTA only, count 8: one lane of every epilogue warp of the pair arrives (remotely from the odd CTA)
// Tiles are walked in bands of kGroupM tile rows so that the clusters running at the same time share A and B panels
// in L2.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "tcgen05_common.cuh"

namespace mfl {
namespace gemm2 {
using namespace mfl::gemm;

constexpr int kThreads = 256;
constexpr uint32_t kABytes = kBlockM * kBlockK * 2;
constexpr int kGroupM = 8;

template <int BN> struct Cfg {
  static_assert(BN == 256 || BN == 192 || BN == 128, "pair tile widths");
  static constexpr int kStages = BN == 128 ? 8 : 6;
  static constexpr uint32_t kBBytes = (BN / 2) * kBlockK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr uint32_t kTmemCols = BN == 128 ? 256 : 512;  // two fp32 accumulators of BN columns, power of two
  static constexpr uint32_t kStoreBytes = 4 * 2 * 4096;
  static constexpr uint32_t kBiasBytes = 256 * 4;
  static constexpr size_t kSmemBytes = 1024 + (size_t)kStages * kStageBytes + kStoreBytes + kBiasBytes + 256;
  static_assert(kSmemBytes <= 232448, "shared memory budget of one CTA");
};

struct Params {
  void* c;
  const void* bias;
  int64_t ldc;
  int32_t M, N, K;
  int32_t bias_fp32, tma_store, a_mn, b_mn;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  // per-thread (non-.aligned) forms: the single-lane producer / issuer loops leave their warps diverged
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
// shared::cluster address of the same shared-memory location in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
// Plain (non-multicast) tile load into THIS CTA's shared memory; completion bytes go to `bar_cluster_addr`, which may be
// the partner CTA's mbarrier.
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, uint32_t bar_cluster_addr, int c0,
                                                 int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1) : "memory");
}
template <bool ACC>
__device__ __forceinline__ void umma2_f16_c(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc) {
  if constexpr (ACC) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.eq.u32 p, 1, 1;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc) : "memory");
  } else {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.eq.u32 p, 1, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc) : "memory");
  }
}
// arrive on the mbarrier at this offset in BOTH CTAs of the pair once all previously issued MMAs have completed
__device__ __forceinline__ void umma2_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr) {
  // default semantics (no .release.cluster): the arrival orders nothing but itself; the data it stands for
  // (tcgen05.ld results) was already waited for, and a cluster-scope release costs a MEMBAR.GPU per arriving thread
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}

// work item -> tile coordinates: bands of kGroupM tile rows, walked row-fastest inside a band
__device__ __forceinline__ void tile_coords(int w, int tiles_m, int tiles_n, int& tm, int& tn) {
  const int per_band = kGroupM * tiles_n;
  const int band = w / per_band, r = w - band * per_band;
  const int first = band * kGroupM;
  const int rows = min(kGroupM, tiles_m - first);
  tn = r / rows;
  tm = first + (r - tn * rows);
}

template <int BN, int ACT, bool C_FP32>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm2sm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ Params p) {
  using C = Cfg<BN>;
  constexpr int kStages = C::kStages;
  extern __shared__ uint8_t smem_raw[];
  // the dynamic shared-memory window starts at the same offset in both CTAs, so the aligned carve-up is identical
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* store_slabs = smem + (size_t)kStages * C::kStageBytes;
  float* sbias = reinterpret_cast<float*>(store_slabs + C::kStoreBytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(store_slabs + C::kStoreBytes + C::kBiasBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int tiles_m = (p.M + 2 * kBlockM - 1) / (2 * kBlockM), tiles_n = (p.N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (p.K + kBlockK - 1) / kBlockK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_b) : "memory");
    if (p.tma_store) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_c) : "memory");
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full_bar + b, 1);
      mbar_init(tmem_empty_bar + b, 8);  // one lane of each epilogue warp of both CTAs
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  } else if (warp == 2) {
    // the same warp of both CTAs allocates the same columns in both tensor memories
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(C::kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  cluster_sync_all();  // the partner's barriers exist before anything is signalled on them
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer (both CTAs): own half of A, own half of B =====
      uint32_t it = 0;
      for (int w = cluster_id; w < num_tiles; w += n_clusters) {
        int tm, tn;
        tile_coords(w, tiles_m, tiles_n, tm, tn);
        const int m0 = tm * 2 * kBlockM + (int)rank * kBlockM;
        const int n0 = tn * BN + (int)rank * (BN / 2);
        // A half-tile that lies entirely past M or N is not loaded at all (whatever the stage holds there only reaches
        // accumulator rows / columns that are never stored); the even CTA arms the barrier with what WILL arrive
        const bool a_ok = m0 < p.M, b_ok = n0 < p.N;
        const uint32_t pair_bytes = kABytes + (tm * 2 * kBlockM + kBlockM < p.M ? kABytes : 0u) + C::kBBytes +
                                    (tn * BN + BN / 2 < p.N ? C::kBBytes : 0u);
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t phase = (it / kStages) & 1;
          mbar_wait(empty_bar + s, phase ^ 1);
          if (rank == 0) mbar_expect_tx(full_bar + s, pair_bytes);
          const uint32_t bar = map_to_cta(smem_u32(full_bar + s), 0);
          uint8_t* a_dst = smem + (size_t)s * C::kStageBytes;
          if (!a_ok) {
          } else if (p.a_mn) {
#pragma unroll
            for (int j = 0; j < kBlockM / 64; ++j) tma_load_2d_pair(a_dst + j * 8192, &tmap_a, bar, m0 + 64 * j, kb * kBlockK);
          } else {
            tma_load_2d_pair(a_dst, &tmap_a, bar, kb * kBlockK, m0);
          }
          if (!b_ok) {
          } else if (p.b_mn) {
            if constexpr (BN % 128 == 0) {
#pragma unroll
              for (int j = 0; j < BN / 128; ++j)
                tma_load_2d_pair(a_dst + kABytes + j * 8192, &tmap_b, bar, n0 + 64 * j, kb * kBlockK);
            }
          } else {
            tma_load_2d_pair(a_dst + kABytes, &tmap_b, bar, kb * kBlockK, n0);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {  // ===== MMA issuer: one lane of the even CTA drives both tensor cores =====
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) |
                             ((uint32_t)((2 * kBlockM) >> 4) << 24) | (p.a_mn ? (1u << 15) : 0u) | (p.b_mn ? (1u << 16) : 0u);
      const uint64_t a_kstep = p.a_mn ? 128 : 2, b_kstep = p.b_mn ? 128 : 2;
      const uint32_t a_lbo = p.a_mn ? 512 : 1, b_lbo = p.b_mn ? 512 : 1;
      uint32_t it = 0, t = 0;
      for (int w = cluster_id; w < num_tiles; w += n_clusters, ++t) {
        const uint32_t buf = t & 1, use = t >> 1;
        mbar_wait(tmem_empty_bar + buf, (use & 1) ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + buf * BN;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t phase = (it / kStages) & 1;
          mbar_wait(full_bar + s, phase);
          tcgen05_fence_after();
          const uint32_t a_addr = smem_u32(smem + (size_t)s * C::kStageBytes);
          const uint64_t adesc = make_sw128_desc(a_addr, a_lbo), bdesc = make_sw128_desc(a_addr + kABytes, b_lbo);
          if (elect_one()) {
            if (kb == 0) umma2_f16_c<false>(tmem_d, adesc, bdesc, idesc);
            else umma2_f16_c<true>(tmem_d, adesc, bdesc, idesc);
#pragma unroll
            for (int k = 1; k < kBlockK / kUmmaK; ++k)
              umma2_f16_c<true>(tmem_d, adesc + (uint64_t)k * a_kstep, bdesc + (uint64_t)k * b_kstep, idesc);
            umma2_commit_pair(empty_bar + s);
          }
          __syncwarp();
        }
        if (elect_one()) umma2_commit_pair(tmem_full_bar + buf);
        __syncwarp();
      }
    }
  } else if (warp >= 4) {  // ===== epilogue (both CTAs): own 128 rows of the pair's tile =====
    const int wq = warp & 3;
    uint32_t t = 0;
    // staging-slab parity runs ACROSS tiles: BN = 192 has three bf16 chunks per tile, and a per-tile counter would
    // reuse the slab whose TMA store `wait_group.read 1` still allows to be in flight
    uint32_t chunk = 0;
    for (int w = cluster_id; w < num_tiles; w += n_clusters, ++t) {
      int tm, tn;
      tile_coords(w, tiles_m, tiles_n, tm, tn);
      const int m0 = tm * 2 * kBlockM + (int)rank * kBlockM, n0 = tn * BN;
      const uint32_t buf = t & 1, use = t >> 1;
      asm volatile("bar.sync 1, 128;" ::: "memory");  // the previous tile's readers of sbias are done
      for (int j = threadIdx.x - 128; j < BN; j += 128) {
        const int col = n0 + j;
        float b = 0.f;
        if (p.bias && col < p.N)
          b = p.bias_fp32 ? __ldg(reinterpret_cast<const float*>(p.bias) + col)
                          : __bfloat162float(__ldg(reinterpret_cast<const __nv_bfloat16*>(p.bias) + col));
        sbias[j] = b;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(tmem_full_bar + buf, use & 1);
      tcgen05_fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(wq * 32) << 16) + buf * BN;
      const int row = m0 + wq * 32 + lane;
      if (p.tma_store) {
        constexpr int cols = C_FP32 ? 32 : 64;
        uint8_t* slabs = store_slabs + wq * 8192;
#pragma unroll 1
        for (int c = 0; c < BN; c += cols) {
          uint32_t r[cols];
          tmem_ld_32x32(tacc + (uint32_t)c, r);
          if constexpr (!C_FP32) tmem_ld_32x32(tacc + (uint32_t)(c + 32), r + 32);
          if (n0 + c >= p.N || m0 + wq * 32 >= p.M) continue;  // warp-uniform
          uint8_t* slab = slabs + (chunk & 1) * 4096;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          uint8_t* my_row = slab + lane * 128;
          constexpr int per = C_FP32 ? 4 : 8;
#pragma unroll
          for (int j8 = 0; j8 < 8; ++j8) {
            float v[per];
#pragma unroll
            for (int j = 0; j < per; ++j) v[j] = apply_act<ACT>(__uint_as_float(r[j8 * per + j]) + sbias[c + j8 * per + j]);
            uint4 piece;
            if constexpr (C_FP32) piece = VecTraits<MFL_FLOAT32>::pack(v);
            else piece = VecTraits<MFL_BFLOAT16>::pack(v);
            *reinterpret_cast<uint4*>(my_row + ((j8 ^ (lane & 7)) << 4)) = piece;
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmap_c, slab, n0 + c, m0 + wq * 32);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          ++chunk;  // counts issued stores only (skipped chunks must not flip the slab parity)
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t r[32];
          tmem_ld_32x32(tacc + (uint32_t)c, r);
          if (n0 + c >= p.N || row >= p.M) continue;
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = apply_act<ACT>(__uint_as_float(r[j]) + sbias[c + j]);
          if constexpr (C_FP32) {
            float* out = reinterpret_cast<float*>(p.c) + (int64_t)row * p.ldc + n0 + c;
            for (int j = 0; j < 32 && n0 + c + j < p.N; ++j) out[j] = v[j];
          } else {
            __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.c) + (int64_t)row * p.ldc + n0 + c;
            for (int j = 0; j < 32 && n0 + c + j < p.N; ++j) out[j] = __float2bfloat16_rn(v[j]);
          }
        }
      }
      tcgen05_fence_before();
      __syncwarp();  // every lane's tcgen05.ld of this buffer has completed (tcgen05.wait::ld inside tmem_ld_32x32)
      if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(tmem_empty_bar + buf), 0));
    }
  }
  if (warp >= 4 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  tcgen05_fence_before();
  cluster_sync_all();  // neither CTA leaves (or frees tensor memory) while the pair's MMAs can still touch it
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(C::kTmemCols) : "memory");
  }
}

}  // namespace gemm2
}  // namespace mfl
