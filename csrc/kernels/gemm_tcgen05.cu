// K10: bf16 GEMM on 5th-generation tensor cores (tcgen05 + TMEM), operands fed by TMA.
//
//   C[M,N] = act( A[M,K] . B[N,K]^T + bias[N] )      A, B K-major bf16, fp32 accumulation in TMEM
//
// This is the matmul under the learners' Linear layers (y = x W^T + b): the reference delegates it
// to TensorFlow / user PyTorch code (keras_model_ops.py:156, examples/pytorch/models/mlp.py:49-68).
//
// Structure (persistent: one CTA per SM walks 128 x BN output tiles, 256 threads):
//   warp 0, one lane : TMA producer - cp.async.bulk.tensor 128x64 (A) and BNx64 (B) bf16 boxes with
//                      128-byte swizzle into a kStages-deep shared-memory ring, mbarrier expect_tx
//   warp 1, one lane : MMA issuer  - 4 x tcgen05.mma (M128, N=BN, K16) per stage, accumulator in TMEM;
//                      tcgen05.commit releases the stage / signals the epilogue
//   warp 2           : TMEM allocator (BN fp32 columns) and deallocator
//   warps 4-7        : epilogue - tcgen05.ld 32 lanes x 32 columns per warp, bias + activation in
//                      registers, bf16 / fp32 store
// The accumulator is double-buffered in TMEM (2 x BN columns, tmem_full / tmem_empty mbarriers), so the
// epilogue of tile i overlaps the TMA + MMA main loop of tile i+1; the smem ring runs continuously
// across tiles.
// Shared-memory operand layout = what TMA's SWIZZLE_128B writes = the canonical K-major SW128 UMMA
// layout (rows of 128 B, 8-row groups 1024 B apart), so the smem descriptors are just
// {start, LBO=1, SBO=1024 B, version 1, layout SWIZZLE_128B} and a K step of 16 elements advances the
// start address by 32 bytes.
#include <cuda.h>

#include <algorithm>

#include "common.cuh"
#include "tcgen05_common.cuh"
#include "gemm_2sm.cuh"

namespace mfl {
namespace gemm {

constexpr int kThreads = 256;
constexpr uint32_t kABytes = kBlockM * kBlockK * 2;

template <int BN> struct Cfg {
  static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
  static constexpr uint32_t kBBytes = BN * kBlockK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr uint32_t kTmemCols = 2 * BN;  // two fp32 accumulators of BN columns
  static constexpr uint32_t kStoreBytes = 4 * 2 * 4096;  // 4 epilogue warps x 2 staging slabs of 32 rows x 128 B
  static constexpr uint32_t kBiasBytes = 256 * 4;        // fp32 bias of the current tile's columns
  static constexpr size_t kSmemBytes = 1024 /*align slack*/ + (size_t)kStages * kStageBytes + kStoreBytes + kBiasBytes + 128 /*barriers*/;
};


struct Epilogue {
  void* c;
  const void* bias;   // bf16 or fp32 [N], may be null
  int64_t ldc;
  int32_t M, N, K;
  int32_t c_fp32;     // output dtype: 0 bf16, 1 fp32
  int32_t bias_fp32;
  int32_t act;        // 0 none, 1 relu, 2 gelu(tanh)
  int32_t tma_store;  // C goes through swizzled smem slabs + cp.async.bulk.tensor stores (needs 16 B row pitch)
  // implicit-GEMM convolution (K11): A rows are output pixels, A columns are (r, s, c) - loaded with im2col TMA
  int32_t conv, out_h, out_w, stride, pad, filt_s, c_blocks;
  // operand majors: 0 = K-major (reduction index contiguous), 1 = MN-major (tensor map is [reduction rows, M or N
  // contiguous]; the tile arrives as 64 x 64 boxes, 8 KB apart, and the instruction descriptor transposes it)
  int32_t a_mn, b_mn;
  // ---- convolution gradients on the same pipeline (K11) ------------------------------------------------------
  // mode 0: plain GEMM / forward convolution (the `conv` flag above)
  // mode 2: data gradient.  A = im2col of dY (rows = dX pixels of one stride-parity class, columns = (tap, k));
  //         B = the filters W[K, R*S*C] read MN-major (a [64 k, 64 c] box per tap); k-block kb -> tap kb / kb_per_tap
  // mode 3: weight gradient.  A = dY [P, K] read MN-major (M = output channels); B = im2col of X read MN-major:
  //         one [64 pixels, 64 channels] im2col box per 64 columns of the (r, s, c) axis; the reduction runs over
  //         the P = N*OH*OW output pixels
  int32_t mode;
  int32_t ntaps, kb_per_tap;
  int8_t tap_ow[16], tap_oh[16], tap_w[16];   // im2col offsets of each tap and the filter tap it multiplies with
  int32_t c_in;                               // mode 2: channels of dX = column pitch of one filter tap in W
  // mode 2 with stride 2: the (h mod 2, w mod 2) classes of dX are independent problems over the same dY; one launch
  // walks all of them (work item = class * tiles + tile); class q uses taps [cls_tap0[q], cls_tap0[q] + cls_ntaps[q])
  int32_t n_cls;
  int8_t cls_tap0[4], cls_ntaps[4], cls_ph[4], cls_pw[4];
  // split-K: the k-blocks of a tile are divided among `splits` CTAs, each adds its fp32 partial tile into a
  // zero-initialised workspace with TMA reduce-add stores (tmap_c is the workspace map); a finalize kernel
  // converts it and re-zeroes the workspace
  int32_t splits, kb_per_split;
  // strided store (mode 2, stride 2): dX is seen as {C, w mod 2, W/2, h mod 2, N*H/2}; the 32 rows of an epilogue
  // slab are a {64 c, 1, bw, 1, bh} box of it
  int32_t store4d, box_w, box_h, box_n, cls_w, cls_h;
};

__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2, int c3,
                                             int c4) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];"
               ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
}

template <int BN, int ACT, bool C_FP32>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ Epilogue ep) {
  using C = Cfg<BN>;
  constexpr int kStages = C::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* store_slabs = smem + (size_t)kStages * C::kStageBytes;  // 1024-byte aligned
  float* sbias = reinterpret_cast<float*>(store_slabs + C::kStoreBytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(store_slabs + C::kStoreBytes + C::kBiasBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;    // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_n = (ep.N + BN - 1) / BN, tiles_m = (ep.M + kBlockM - 1) / kBlockM;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (ep.K + kBlockK - 1) / kBlockK;
  // work item w = tile * splits + split; a split owns k-blocks [split * kb_per_split, +kb_per_split) (clipped)
  const int splits = ep.splits > 1 ? ep.splits : 1;
  const int n_cls = ep.n_cls > 1 ? ep.n_cls : 1;
  const int work_per_cls = num_tiles * splits;
  const int num_work = work_per_cls * n_cls;
  const int kbps = splits > 1 ? ep.kb_per_split : (1 << 30);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_b) : "memory");
    if (ep.tma_store) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_c) : "memory");
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full_bar + b, 1);
      mbar_init(tmem_empty_bar + b, 128);  // every epilogue thread arrives
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  } else if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(C::kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      uint32_t it = 0;  // k-block counter across all tiles of this CTA: the ring never drains
      const int ohw = ep.out_h * ep.out_w;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        const int cls = w / work_per_cls, wi = w - cls * work_per_cls;
        const int tile = wi / splits, split = wi - tile * splits;
        const int m0 = (tile / tiles_n) * kBlockM, n0 = (tile % tiles_n) * BN;
        const int nkb = ep.n_cls ? (int)ep.cls_ntaps[cls] * ep.kb_per_tap : num_kb;
        const int kb0 = split * kbps, kb1 = min(nkb, kb0 + kbps);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t phase = (it / kStages) & 1;
          mbar_wait(empty_bar + s, phase ^ 1);
          uint8_t* a_dst = smem + (size_t)s * C::kStageBytes;
          if (ep.mode == 3) {  // weight gradient
            int nvalid = 0;
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) nvalid += (n0 + 64 * j < ep.N) ? 1 : 0;
            mbar_expect_tx(full_bar + s, kABytes + (uint32_t)nvalid * 8192u);
#pragma unroll
            for (int j = 0; j < kBlockM / 64; ++j) tma_load_2d(a_dst + j * 8192, &tmap_a, full_bar + s, m0 + 64 * j, kb * kBlockK);
            const int p0 = kb * kBlockK, img = p0 / ohw, rem = p0 - img * ohw;
            const int oh = rem / ep.out_w, ow = rem - oh * ep.out_w;
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) {
              const int nb = (n0 >> 6) + j;
              if (nb * 64 < ep.N) {
                const int tap = nb / ep.c_blocks, cb = nb - tap * ep.c_blocks;
                tma_load_im2col_4d(a_dst + kABytes + j * 8192, &tmap_b, full_bar + s, cb * kBlockK,
                                   ow * ep.stride - ep.pad, oh * ep.stride - ep.pad, img,
                                   (uint16_t)(tap % ep.filt_s), (uint16_t)(tap / ep.filt_s));
              }
            }
            continue;
          }
          mbar_expect_tx(full_bar + s, C::kStageBytes);
          if (ep.mode == 2) {  // data gradient: A = im2col(dY) for tap kb / kb_per_tap, B = W tap, MN-major
            const int tl = kb / ep.kb_per_tap, kk = kb - tl * ep.kb_per_tap;
            const int tap = (ep.n_cls ? (int)ep.cls_tap0[cls] : 0) + tl;
            const int img = m0 / ohw, rem = m0 - img * ohw;
            const int oh = rem / ep.out_w, ow = rem - oh * ep.out_w;
            tma_load_im2col_4d(a_dst, &tmap_a, full_bar + s, kk * kBlockK, ow - ep.pad, oh - ep.pad, img,
                               (uint16_t)ep.tap_ow[tap], (uint16_t)ep.tap_oh[tap]);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(a_dst + kABytes + j * 8192, &tmap_b, full_bar + s, (int)ep.tap_w[tap] * ep.c_in + n0 + 64 * j,
                          kk * kBlockK);
            continue;
          }
          if (ep.conv) {
            const int img = m0 / ohw, rem = m0 % ohw;
            const int oh = rem / ep.out_w, ow = rem % ep.out_w;
            const int rs = kb / ep.c_blocks, cb = kb % ep.c_blocks;
            tma_load_im2col_4d(a_dst, &tmap_a, full_bar + s, cb * kBlockK, ow * ep.stride - ep.pad,
                               oh * ep.stride - ep.pad, img, (uint16_t)(rs % ep.filt_s), (uint16_t)(rs / ep.filt_s));
          } else if (ep.a_mn) {
#pragma unroll
            for (int j = 0; j < kBlockM / 64; ++j) tma_load_2d(a_dst + j * 8192, &tmap_a, full_bar + s, m0 + 64 * j, kb * kBlockK);
          } else {
            tma_load_2d(a_dst, &tmap_a, full_bar + s, kb * kBlockK, m0);
          }
          if (ep.b_mn) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(a_dst + kABytes + j * 8192, &tmap_b, full_bar + s, n0 + 64 * j, kb * kBlockK);
          } else {
            tma_load_2d(a_dst + kABytes, &tmap_b, full_bar + s, kb * kBlockK, n0);
          }
        }
      }
    }
  } else if (warp == 1) {  // ===== MMA issuer =====
    // The whole warp runs the loop (so the descriptor arithmetic is warp-uniform and stays in uniform registers:
    // with a lane-0-only loop every tcgen05.mma cost an ELECT + two R2UR moves and the issue interval was ~140
    // cycles, capping the tensor pipe at 88 % for N = 256 and far lower for narrow tiles); one elected lane issues.
    const uint32_t idesc = make_idesc<BN>() | (ep.a_mn ? (1u << 15) : 0u) | (ep.b_mn ? (1u << 16) : 0u);
    // descriptor units are 16 B: a K step of 16 elements is +32 B inside the swizzle row for a K-major operand and
    // two 1024-byte atoms (16 reduction rows of 128 B) for an MN-major one, whose 64-wide blocks lie 8 KB apart
    const uint64_t a_kstep = ep.a_mn ? 128 : 2, b_kstep = ep.b_mn ? 128 : 2;
    const uint32_t a_lbo = ep.a_mn ? 512 : 1, b_lbo = ep.b_mn ? 512 : 1;
    uint32_t it = 0, t = 0;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++t) {
      const int cls = w / work_per_cls, split = (w - cls * work_per_cls) % splits;
      const int nkb = ep.n_cls ? (int)ep.cls_ntaps[cls] * ep.kb_per_tap : num_kb;
      const int kb0 = split * kbps, kb1 = min(nkb, kb0 + kbps);
      const uint32_t buf = t & 1, use = t >> 1;
      mbar_wait(tmem_empty_bar + buf, (use & 1) ^ 1);  // epilogue has drained this accumulator
      tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + buf * BN;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % kStages;
        const uint32_t phase = (it / kStages) & 1;
        mbar_wait(full_bar + s, phase);
        tcgen05_fence_after();
        const uint32_t a_addr = smem_u32(smem + (size_t)s * C::kStageBytes);
        const uint64_t adesc = make_sw128_desc(a_addr, a_lbo), bdesc = make_sw128_desc(a_addr + kABytes, b_lbo);
        if (elect_one()) {
          // +32 bytes per K step inside the 128-byte swizzle atom (encoded >> 4)
          if (kb == kb0) umma_f16_c<false>(tmem_d, adesc, bdesc, idesc);
          else umma_f16_c<true>(tmem_d, adesc, bdesc, idesc);
#pragma unroll
          for (int k = 1; k < kBlockK / kUmmaK; ++k)
            umma_f16_c<true>(tmem_d, adesc + (uint64_t)k * a_kstep, bdesc + (uint64_t)k * b_kstep, idesc);
          umma_commit(empty_bar + s);  // frees the smem slot once these MMAs have read it
        }
        __syncwarp();
      }
      if (elect_one()) umma_commit(tmem_full_bar + buf);  // accumulator complete
      __syncwarp();
    }
  } else if (warp >= 4) {  // ===== epilogue: TMEM -> registers -> global =====
    const int wq = warp & 3;  // TMEM lane quarter this warp may access
    uint32_t t = 0;
    // staging-slab parity runs ACROSS tiles: with an odd number of chunks per tile (one for BN = 64 bf16) a per-tile
    // counter would reuse the slab whose TMA store `wait_group.read 1` still allows to be in flight
    uint32_t chunk = 0;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++t) {
      const int cls = w / work_per_cls;
      const int tile = (w - cls * work_per_cls) / splits;
      const int m0 = (tile / tiles_n) * kBlockM, n0 = (tile % tiles_n) * BN;
      const uint32_t buf = t & 1, use = t >> 1;
      mbar_wait(tmem_full_bar + buf, use & 1);
      tcgen05_fence_after();
      const int row = m0 + wq * 32 + lane;
      const bool row_ok = row < ep.M;
      // stage this tile's bias (zero where there is none / past N) so the inner loops are branch-free
      asm volatile("bar.sync 1, 128;" ::: "memory");  // previous tile's readers are done
      for (int j = threadIdx.x - 128; j < BN; j += 128) {
        const int col = n0 + j;
        float b = 0.f;
        if (ep.bias && col < ep.N)
          b = ep.bias_fp32 ? __ldg(reinterpret_cast<const float*>(ep.bias) + col)
                           : __bfloat162float(__ldg(reinterpret_cast<const __nv_bfloat16*>(ep.bias) + col));
        sbias[j] = b;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (ep.tma_store) {
        // Coalesced path: each warp packs 32 rows x 128 B into a swizzled slab and one lane hands it to
        // the TMA store engine (which also clips the ragged M / N edges).
        constexpr int cols = C_FP32 ? 32 : 64;  // columns per 128-byte slab row
        uint8_t* slabs = store_slabs + wq * 8192;
#pragma unroll 1
        for (int c = 0; c < BN; c += cols) {
          uint32_t r[cols];
          tmem_ld_32x32(tmem_base + ((uint32_t)(wq * 32) << 16) + buf * BN + (uint32_t)c, r);
          if constexpr (!C_FP32) tmem_ld_32x32(tmem_base + ((uint32_t)(wq * 32) << 16) + buf * BN + (uint32_t)(c + 32), r + 32);
          if (n0 + c >= ep.N) continue;  // warp-uniform
          uint8_t* slab = slabs + (chunk & 1) * 4096;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // slab used 2 chunks ago is free
          __syncwarp();
          uint8_t* my_row = slab + lane * 128;
          constexpr int per = C_FP32 ? 4 : 8;  // elements per 16-byte piece
#pragma unroll
          for (int j8 = 0; j8 < 8; ++j8) {  // eight 16-byte pieces per row, XOR-swizzled like TMA's SWIZZLE_128B
            float v[per];
#pragma unroll
            for (int j = 0; j < per; ++j) v[j] = apply_act<ACT>(__uint_as_float(r[j8 * per + j]) + sbias[c + j8 * per + j]);
            uint4 piece;
            if constexpr (C_FP32) piece = VecTraits<MFL_FLOAT32>::pack(v);
            else piece = VecTraits<MFL_BFLOAT16>::pack(v);
            *reinterpret_cast<uint4*>(my_row + ((j8 ^ (lane & 7)) << 4)) = piece;
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> async proxy
          __syncwarp();
          if (lane == 0) {
            if (splits > 1) {
              tma_reduce_add_2d(&tmap_c, slab, n0 + c, m0 + wq * 32);  // fp32 partial tile += into the workspace
            } else if (ep.store4d) {
              // rows m0 + wq*32 .. +31 of the class's pixel grid = a {bw, bh} box of the strided view of dX
              const int r0 = m0 + wq * 32;
              tma_store_5d(&tmap_c, slab, n0 + c, (int)ep.cls_pw[cls], r0 % ep.cls_w, (int)ep.cls_ph[cls], r0 / ep.cls_w);
            } else {
              tma_store_2d(&tmap_c, slab, n0 + c, m0 + wq * 32);
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          ++chunk;  // counts issued stores only (skipped chunks past N must not flip the slab parity)
        }
      } else {
#pragma unroll 1
      for (int c = 0; c < BN; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(wq * 32) << 16) + buf * BN + (uint32_t)c, r);
        if (n0 + c >= ep.N || !row_ok) continue;
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = apply_act<ACT>(__uint_as_float(r[j]) + sbias[c + j]);
        const bool full = (n0 + c + 32 <= ep.N);
        if constexpr (C_FP32) {
          float* out = reinterpret_cast<float*>(ep.c) + (int64_t)row * ep.ldc + n0 + c;
          if (full && (ep.ldc % 4 == 0)) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(out + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
            for (int j = 0; j < 32 && n0 + c + j < ep.N; ++j) out[j] = v[j];
          }
        } else {
          __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(ep.c) + (int64_t)row * ep.ldc + n0 + c;
          if (full && (ep.ldc % 8 == 0)) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) *reinterpret_cast<uint4*>(out + j) = VecTraits<MFL_BFLOAT16>::pack(v + j);
          } else {
            for (int j = 0; j < 32 && n0 + c + j < ep.N; ++j) out[j] = __float2bfloat16_rn(v[j]);
          }
        }
      }
      }  // direct-store path
      tcgen05_fence_before();
      mbar_arrive(tmem_empty_bar + buf);  // this thread's TMEM reads of the buffer are done
    }
  }
  if (warp >= 4 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // stores have left smem
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(C::kTmemCols) : "memory");
  }
}

// ---- host side: tensor maps through the driver entry point (no link-time dependency on libcuda) ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

static int make_map(CUtensorMap* map, const void* base, int rows, int k, int64_t ld, int box_rows, bool fp32 = false) {
  bind_context();  // driver-API call below: this thread (e.g. an autograd worker) may not have the context bound
  EncodeTiledFn fn = encode_fn();
  MFL_ARG(fn != nullptr, "gemm: cuTensorMapEncodeTiled is not available from this driver");
  const int esz = fp32 ? 4 : 2;
  cuuint64_t dims[2] = {(cuuint64_t)k, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)(128 / esz), (cuuint32_t)box_rows};  // 128-byte inner extent
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MFL_ARG(r == CUDA_SUCCESS, "gemm: cuTensorMapEncodeTiled failed with CUresult %d (rows=%d k=%d ld=%lld)", (int)r, rows, k,
          (long long)ld);
  return 0;
}

}  // namespace gemm
}  // namespace mfl

using namespace mfl;

namespace mfl { namespace gemm {
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const Epilogue& ep, int bn,
                       int act, bool c32, cudaStream_t st) {
  const int tiles = ((ep.N + bn - 1) / bn) * ((ep.M + kBlockM - 1) / kBlockM);
  const int grid = std::min(tiles * std::max(1, (int)ep.splits) * std::max(1, (int)ep.n_cls), sm_count());
  int rc = 0;
#define MFL_GEMM_LAUNCH(BN_, ACT_, F32_)                                                                          \
  do {                                                                                                            \
    auto kern = gemm_bf16_tcgen05_kernel<BN_, ACT_, F32_>;                                                        \
    rc = check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg<BN_>::kSmemBytes), \
               "cudaFuncSetAttribute(gemm smem)");                                                                \
    if (!rc) kern<<<grid, kThreads, Cfg<BN_>::kSmemBytes, st>>>(ta, tb, tc, ep);                                  \
  } while (0)
#define MFL_GEMM_ACT(BN_, F32_)                                         \
  do {                                                                  \
    if (act == 0) MFL_GEMM_LAUNCH(BN_, 0, F32_);                        \
    else if (act == 1) MFL_GEMM_LAUNCH(BN_, 1, F32_);                   \
    else MFL_GEMM_LAUNCH(BN_, 2, F32_);                                 \
  } while (0)
  if (bn == 256) { if (c32) MFL_GEMM_ACT(256, true); else MFL_GEMM_ACT(256, false); }
  else if (bn == 128) { if (c32) MFL_GEMM_ACT(128, true); else MFL_GEMM_ACT(128, false); }
  else { if (c32) MFL_GEMM_ACT(64, true); else MFL_GEMM_ACT(64, false); }
#undef MFL_GEMM_ACT
#undef MFL_GEMM_LAUNCH
  if (rc) return rc;
  return check(cudaGetLastError(), "gemm_bf16_tcgen05_kernel launch");
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeIm2colFn encode_im2col_fn() {
  static EncodeIm2colFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeIm2colFn>(p);
  }
  return fn;
}
}}  // namespace mfl::gemm

namespace mfl { namespace gemm {
// [reduction rows, mn contiguous] matrix -> 64 x 64 boxes (one 128-byte swizzle row per reduction index)
static int make_map_mn(CUtensorMap* map, const void* base, int mn, int kred, int64_t ld) {
  bind_context();
  EncodeTiledFn fn = encode_fn();
  MFL_ARG(fn != nullptr, "gemm: cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {(cuuint64_t)mn, (cuuint64_t)kred};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64, 64};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MFL_ARG(r == CUDA_SUCCESS, "gemm: cuTensorMapEncodeTiled (MN-major) failed with CUresult %d (mn=%d k=%d ld=%lld)", (int)r, mn,
          kred, (long long)ld);
  return 0;
}
}}  // namespace mfl::gemm

static int gemm_bf16_impl(void* c, int c_dtype, int64_t ldc, const void* a, int64_t lda, int a_mn, const void* b,
                          int64_t ldb, int b_mn, const void* bias, int bias_dtype, int m, int n, int k, int act,
                          void* stream, float* ws = nullptr, size_t ws_bytes = 0);

namespace mfl { namespace gemm {
// ---- CTA-pair variant (gemm_2sm.cuh) ----------------------------------------------------------------------------
// How many 2-CTA clusters of this kernel the device runs at once (GPCs with an odd SM count leave an SM unpaired).
template <typename K> static int max_pair_clusters(K kern, size_t smem) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * (unsigned)(sm_count() / 2));
  cfg.blockDim = dim3(gemm2::kThreads);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    n = sm_count() / 2;
  }
  return std::min(n, sm_count() / 2);
}

struct Epilogue;
static void plan_splits(Epilogue& ep, int bn, const float* ws, size_t ws_bytes);
static int finalize_splits(const Epilogue& ep, float* ws, void* out, int64_t ld, cudaStream_t st);
static int g_pair_mode = [] { const char* e = getenv("MFL_GEMM_2SM"); return e ? atoi(e) : -1; }();

static int pair_cluster_estimate() { return std::max(1, sm_count() / 2); }

// Tile width of the pair kernel for this problem, or 0 to stay on the single-CTA kernel.  MFL_GEMM_2SM=0 disables,
// =1 forces it wherever it is legal.  Cost model: waves over the resident clusters x (tile width + a fixed per-tile
// cost); BN = 192 exists for shapes like N = 3072 whose 256-wide tiling ends in a nearly empty wave.
static int pair_tile_width(int m, int n, int k, bool b_mn) {
  const int mode = g_pair_mode;
  if (mode == 0) return 0;
  if (m < 2 * kBlockM || n < 128 || k < kBlockK) return 0;
  const int clusters = pair_cluster_estimate();
  const int tiles_m = (m + 2 * kBlockM - 1) / (2 * kBlockM);
  if (mode != 1 && tiles_m * ((n + 127) / 128) < clusters / 2) return 0;  // too small to fill the machine in pairs
  int best = 0;
  long best_cost = 0;
  for (int bn : {256, 192, 128}) {
    if (bn == 192 && b_mn) continue;  // an MN-major B half is made of 64-column boxes
    if (n < bn && bn != 128) continue;
    const long tiles = (long)tiles_m * ((n + bn - 1) / bn);
    const long cost = ((tiles + clusters - 1) / clusters) * (bn + 32);
    if (!best || cost < best_cost) { best = bn; best_cost = cost; }
  }
  return best;
}

static int launch_gemm_pair(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const gemm2::Params& pp,
                            int bn, int act, bool c32, cudaStream_t st) {
  const int tiles = ((pp.N + bn - 1) / bn) * ((pp.M + 2 * kBlockM - 1) / (2 * kBlockM));
  int rc = 0;
#define MFL_GEMM2_LAUNCH(BN_, ACT_, F32_)                                                                           \
  do {                                                                                                              \
    auto kern = gemm2::gemm2sm_kernel<BN_, ACT_, F32_>;                                                             \
    static int clusters = 0;                                                                                        \
    rc = check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,                              \
                                    (int)gemm2::Cfg<BN_>::kSmemBytes), "cudaFuncSetAttribute(gemm2 smem)");         \
    if (!rc && !clusters) clusters = max_pair_clusters(kern, gemm2::Cfg<BN_>::kSmemBytes);                          \
    if (!rc) kern<<<2 * std::min(tiles, clusters), gemm2::kThreads, gemm2::Cfg<BN_>::kSmemBytes, st>>>(ta, tb, tc, pp); \
  } while (0)
#define MFL_GEMM2_ACT(BN_, F32_)                                        \
  do {                                                                  \
    if (act == 0) MFL_GEMM2_LAUNCH(BN_, 0, F32_);                       \
    else if (act == 1) MFL_GEMM2_LAUNCH(BN_, 1, F32_);                  \
    else MFL_GEMM2_LAUNCH(BN_, 2, F32_);                                \
  } while (0)
  if (bn == 256) { if (c32) MFL_GEMM2_ACT(256, true); else MFL_GEMM2_ACT(256, false); }
  else if (bn == 192) { if (c32) MFL_GEMM2_ACT(192, true); else MFL_GEMM2_ACT(192, false); }
  else { if (c32) MFL_GEMM2_ACT(128, true); else MFL_GEMM2_ACT(128, false); }
#undef MFL_GEMM2_ACT
#undef MFL_GEMM2_LAUNCH
  if (rc) return rc;
  return check(cudaGetLastError(), "gemm2sm_kernel launch");
}
}}  // namespace mfl::gemm

// -1 automatic (default, or MFL_GEMM_2SM from the environment), 0 never use the CTA-pair kernel, 1 wherever it is legal
extern "C" int mfl_gemm_set_pair_mode(int mode) {
  const int old = mfl::gemm::g_pair_mode;
  mfl::gemm::g_pair_mode = mode < 0 ? -1 : (mode ? 1 : 0);
  return old;
}

extern "C" int mfl_gemm_bf16(void* c, int c_dtype, int64_t ldc, const void* a, int64_t lda, const void* b, int64_t ldb,
                             const void* bias, int bias_dtype, int m, int n, int k, int act, void* stream) {
  return gemm_bf16_impl(c, c_dtype, ldc, a, lda, 0, b, ldb, 0, bias, bias_dtype, m, n, k, act, stream);
}

// C[m,n] = act(op(A) . op(B)^T + bias): a_mn = 0: A is [m, k] row-major; a_mn = 1: A is given as [k, m] row-major
// (its transpose: the m index contiguous).  Same for B with n.  This is what the backward GEMMs of a Linear layer
// need (dX = dY . W reads W [n_red, k_out] as an MN-major B; dW = dY^T . X reads both operands MN-major), so no
// transposed copies are ever made.
extern "C" int mfl_gemm_bf16_ex(void* c, int c_dtype, int64_t ldc, const void* a, int64_t lda, int a_mn, const void* b,
                                int64_t ldb, int b_mn, const void* bias, int bias_dtype, int m, int n, int k, int act,
                                void* stream) {
  return gemm_bf16_impl(c, c_dtype, ldc, a, lda, a_mn, b, ldb, b_mn, bias, bias_dtype, m, n, k, act, stream);
}

// Same product with an optional split of the reduction: ws is a ZERO-initialised fp32 scratch of at least m * n floats
// (handed back zeroed).  It is used when the output has too few tiles to occupy the machine (the weight gradient of a
// square Linear layer: 1024 x 1024 outputs, reduction over thousands of tokens), bf16 output, no bias / activation.
extern "C" int mfl_gemm_bf16_ws(void* c, int c_dtype, int64_t ldc, const void* a, int64_t lda, int a_mn, const void* b,
                                int64_t ldb, int b_mn, int m, int n, int k, float* ws, size_t ws_bytes, void* stream) {
  return gemm_bf16_impl(c, c_dtype, ldc, a, lda, a_mn, b, ldb, b_mn, nullptr, 0, m, n, k, 0, stream, ws, ws_bytes);
}

static int gemm_bf16_impl(void* c, int c_dtype, int64_t ldc, const void* a, int64_t lda, int a_mn, const void* b,
                          int64_t ldb, int b_mn, const void* bias, int bias_dtype, int m, int n, int k, int act,
                          void* stream, float* ws, size_t ws_bytes) {
  using namespace mfl::gemm;
  MFL_ARG(a && b && c && m > 0 && n > 0 && k > 0, "gemm: null operand or empty shape");
  MFL_ARG(c_dtype == MFL_BFLOAT16 || c_dtype == MFL_FLOAT32, "gemm: output must be BFLOAT16 or FLOAT32");
  MFL_ARG(!bias || bias_dtype == MFL_BFLOAT16 || bias_dtype == MFL_FLOAT32, "gemm: bias must be BFLOAT16 or FLOAT32");
  MFL_ARG(lda % 8 == 0 && ldb % 8 == 0 && ((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0,
          "gemm: A and B need 16-byte aligned bases and row pitches that are multiples of 8 elements (TMA)");
  MFL_ARG(act >= 0 && act <= 2, "gemm: unknown activation %d", act);
  CUtensorMap ta, tb, tc;
  if (const int pbn = pair_tile_width(m, n, k, b_mn != 0)) {  // CTA pairs (tcgen05.mma.cta_group::2)
    if (int rc = a_mn ? make_map_mn(&ta, a, m, k, lda) : make_map(&ta, a, m, k, lda, kBlockM)) return rc;
    if (int rc = b_mn ? make_map_mn(&tb, b, n, k, ldb) : make_map(&tb, b, n, k, ldb, pbn / 2)) return rc;
    const bool c32p = c_dtype == MFL_FLOAT32;
    const bool tma_store_p = ((uintptr_t)c & 15) == 0 && (ldc * (c32p ? 4 : 2)) % 16 == 0;
    if (tma_store_p) {
      if (int rc = make_map(&tc, c, m, n, ldc, 32, c32p)) return rc;
    } else {
      tc = ta;
    }
    gemm2::Params pp{c, bias, ldc, m, n, k, bias_dtype == MFL_FLOAT32, tma_store_p, a_mn ? 1 : 0, b_mn ? 1 : 0};
    return launch_gemm_pair(ta, tb, tc, pp, pbn, act, c32p, (cudaStream_t)stream);
  }
  const int bn = n > 128 ? 256 : (n > 64 ? 128 : 64);
  if (int rc = a_mn ? make_map_mn(&ta, a, m, k, lda) : make_map(&ta, a, m, k, lda, kBlockM)) return rc;
  if (int rc = b_mn ? make_map_mn(&tb, b, n, k, ldb) : make_map(&tb, b, n, k, ldb, bn)) return rc;
  const bool c32 = c_dtype == MFL_FLOAT32;
  const bool tma_store = ((uintptr_t)c & 15) == 0 && (ldc * (c32 ? 4 : 2)) % 16 == 0;
  if (tma_store) {
    if (int rc = make_map(&tc, c, m, n, ldc, 32, c32)) return rc;
  } else {
    tc = ta;  // unused by the kernel
  }
  Epilogue ep{c, bias, ldc, m, n, k, c32, bias_dtype == MFL_FLOAT32, act, tma_store, 0, 0, 0, 0, 0, 0, 0, a_mn ? 1 : 0, b_mn ? 1 : 0};
  if (ws && !bias && act == 0 && !c32 && tma_store) {
    plan_splits(ep, bn, ws, ws_bytes);
    if (ep.splits > 1) {
      ep.c = ws; ep.c_fp32 = 1; ep.ldc = n;
      if (int rc = make_map(&tc, ws, m, n, n, 32, true)) return rc;
      if (int rc = launch_gemm(ta, tb, tc, ep, bn, 0, true, (cudaStream_t)stream)) return rc;
      return finalize_splits(ep, ws, c, ldc, (cudaStream_t)stream);
    }
  }
  return launch_gemm(ta, tb, tc, ep, bn, act, c32, (cudaStream_t)stream);
}

// ---- K11: implicit-GEMM convolution forward (also used for stride-1 data gradients) ----------------
// y[N, OH, OW, K] = conv(x[N, H, W, C], w[K, R, S, C]), NHWC bf16, C % 64 == 0, K % 8 == 0.
// Replaces the framework convolutions under the learners' training loop (keras_model_ops.py:156).  The A
// operand is never materialised: the TMA unit gathers 128 output pixels x 64 channels per filter tap
// straight into the swizzled shared-memory tile the tensor core reads (im2col tensor map).
namespace mfl { namespace gemm {
__global__ void splitk_finalize_kernel(float* __restrict__ ws, __nv_bfloat16* __restrict__ out, int rows, int cols,
                                       int64_t ld, int accumulate);
// Split the reduction of a convolution whose output has too few tiles to fill the machine (the 8x8 and 4x4 layers
// of ResNet-18 give 32-64 tiles for 148 SMs): each (tile, split) work item adds its fp32 partial into `ws`.
// Output-tile width for a convolution: the widest N tile that still gives (almost) every SM a tile.  A 128 x 256 tile
// is the most efficient instruction shape, but the 8x8 / 4x4 layers of ResNet-18 have only 64 / 16 row tiles, and a CTA
// that owns a whole 256-wide tile then runs its 36-72 k-blocks alone while half the machine idles.
static int pick_bn(int m, int n) {
  const int tiles_m = (m + kBlockM - 1) / kBlockM;
  for (int bn : {256, 128}) {
    if (n < bn && !(bn == 128 && n > 64)) continue;
    if (tiles_m * ((n + bn - 1) / bn) * 5 >= sm_count() * 4) return bn;
  }
  return n > 64 ? 128 : 64;
}

static void plan_splits(Epilogue& ep, int bn, const float* ws, size_t ws_bytes) {
  const int tiles = ((ep.N + bn - 1) / bn) * ((ep.M + kBlockM - 1) / kBlockM);
  const int num_kb = (ep.K + kBlockK - 1) / kBlockK;
  ep.splits = 1;
  ep.kb_per_split = num_kb;
  // the fp32 reduce-add pass and the finalize launch cost ~10 us: only worth it when under half of the SMs have a tile
  if (!ws || (size_t)ep.M * ep.N * sizeof(float) > ws_bytes || tiles * 2 > sm_count() || ep.N % 8 != 0) return;
  int splits = std::min(num_kb / 4, (sm_count() + tiles - 1) / tiles);
  if (splits < 2) return;
  ep.kb_per_split = (num_kb + splits - 1) / splits;
  ep.splits = (num_kb + ep.kb_per_split - 1) / ep.kb_per_split;
  if (ep.splits < 2) { ep.splits = 1; ep.kb_per_split = num_kb; }
}
static int finalize_splits(const Epilogue& ep, float* ws, void* out, int64_t ld, cudaStream_t st) {
  const int64_t vecs = (int64_t)ep.M * ep.N / 8;
  const int blocks = (int)std::min<int64_t>((vecs + 255) / 256, (int64_t)sm_count() * 4);
  splitk_finalize_kernel<<<blocks, 256, 0, st>>>(ws, static_cast<__nv_bfloat16*>(out), ep.M, ep.N, ld, 0);
  return check(cudaGetLastError(), "splitk_finalize_kernel launch");
}
}}  // namespace mfl::gemm

// ws / ws_bytes: optional zero-initialised fp32 scratch (handed back zeroed) that lets small-output layers split
// their reduction over more SMs; pass NULL / 0 to run without it.
extern "C" int mfl_conv2d_fprop_bf16_ws(void* y, const void* x, const void* w, float* ws, size_t ws_bytes, int n, int h,
                                        int wd, int c, int k, int r, int s, int stride, int pad, void* stream) {
  using namespace mfl::gemm;
  MFL_ARG(y && x && w && n > 0 && h > 0 && wd > 0, "conv2d: null operand or empty shape");
  MFL_ARG(c % 64 == 0 && k % 8 == 0, "conv2d: needs C %% 64 == 0 and K %% 8 == 0 (got C=%d K=%d)", c, k);
  MFL_ARG(stride >= 1 && stride <= 8 && pad >= 0 && pad < 128 && r >= 1 && s >= 1 && r * s <= 49, "conv2d: bad geometry");
  const int oh = (h + 2 * pad - r) / stride + 1, ow = (wd + 2 * pad - s) / stride + 1;
  MFL_ARG(oh > 0 && ow > 0, "conv2d: empty output");
  EncodeIm2colFn fn = encode_im2col_fn();
  MFL_ARG(fn != nullptr, "conv2d: cuTensorMapEncodeIm2col is not available from this driver");
  CUtensorMap ta, tb, tc;
  {
    cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)wd, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)wd * c * 2, (cuuint64_t)h * wd * c * 2};
    int lower[2] = {-pad, -pad};
    int upper[2] = {pad - (s - 1), pad - (r - 1)};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult res = fn(&ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, lower, upper,
                      (cuuint32_t)kBlockK, (cuuint32_t)kBlockM, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MFL_ARG(res == CUDA_SUCCESS, "conv2d: cuTensorMapEncodeIm2col failed with CUresult %d", (int)res);
  }
  const int m = n * oh * ow, kk = r * s * c;
  const int bn = pick_bn(m, k);
  if (int rc = make_map(&tb, w, k, kk, kk, bn)) return rc;
  Epilogue ep{y, nullptr, k, m, k, kk, 0, 0, 0, 1, 1, oh, ow, stride, pad, s, c / 64};
  plan_splits(ep, bn, ws, ws_bytes);
  const bool split = ep.splits > 1;
  if (split) {
    ep.c = ws; ep.c_fp32 = 1;
    if (int rc = make_map(&tc, ws, m, k, k, 32, true)) return rc;
  } else {
    if (int rc = make_map(&tc, y, m, k, k, 32, false)) return rc;
  }
  if (int rc = launch_gemm(ta, tb, tc, ep, bn, 0, split, (cudaStream_t)stream)) return rc;
  return split ? finalize_splits(ep, ws, y, k, (cudaStream_t)stream) : 0;
}

extern "C" int mfl_conv2d_fprop_bf16(void* y, const void* x, const void* w, int n, int h, int wd, int c, int k, int r,
                                     int s, int stride, int pad, void* stream) {
  return mfl_conv2d_fprop_bf16_ws(y, x, w, nullptr, 0, n, h, wd, c, k, r, s, stride, pad, stream);
}

// =====================================================================================================================
// K11 gradients: data gradient (stride 1 and, by stride-parity classes, stride 2) and weight gradient (split-K over the
// output pixels) on the same persistent tcgen05 pipeline.  Reference counterpart: the framework convolution backward
// under model.fit (keras_model_ops.py:156).
// =====================================================================================================================
namespace mfl { namespace gemm {

// ws (fp32 [rows, cols], the split-K workspace) -> out (bf16, row pitch ld), optionally accumulating; re-zeroes ws
__global__ void splitk_finalize_kernel(float* __restrict__ ws, __nv_bfloat16* __restrict__ out, int rows, int cols,
                                       int64_t ld, int accumulate) {
  const int64_t vec_per_row = cols / 8;
  const int64_t total = (int64_t)rows * vec_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / vec_per_row, c = (i - r * vec_per_row) * 8;
    float4* p = reinterpret_cast<float4*>(ws + r * cols + c);
    const float4 a = p[0], b = p[1];
    float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint4* o = reinterpret_cast<uint4*>(out + r * ld + c);
    if (accumulate) {
      float f[8];
      VecTraits<MFL_BFLOAT16>::unpack(*o, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += f[j];
    }
    *o = VecTraits<MFL_BFLOAT16>::pack(v);
    p[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    p[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

static int make_im2col_map(CUtensorMap* map, const void* base, int n, int h, int wd, int c, int lower_w, int lower_h,
                           int upper_w, int upper_h, int trav_stride, int pixels) {
  bind_context();
  EncodeIm2colFn fn = encode_im2col_fn();
  MFL_ARG(fn != nullptr, "conv2d: cuTensorMapEncodeIm2col is not available from this driver");
  cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)wd, (cuuint64_t)h, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)wd * c * 2, (cuuint64_t)h * wd * c * 2};
  int lower[2] = {lower_w, lower_h};
  int upper[2] = {upper_w, upper_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)trav_stride, (cuuint32_t)trav_stride, 1};
  CUresult res = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper,
                    (cuuint32_t)kBlockK, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MFL_ARG(res == CUDA_SUCCESS, "conv2d: cuTensorMapEncodeIm2col failed with CUresult %d", (int)res);
  return 0;
}

// strided parity-class view of dX for the 4-D epilogue store: {C, W/s, H/s, N}, box {64, bw, bh, bn}
// strided view of dX for the stride-2 epilogue store: {C, w mod 2, W/2, h mod 2, N*H/2}, box {64, 1, bw, 1, bh}
static int make_store5d_map(CUtensorMap* map, void* base, int n, int h, int wd, int c, int bw, int bh) {
  bind_context();
  EncodeTiledFn fn = encode_fn();
  MFL_ARG(fn != nullptr, "conv2d: cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[5] = {(cuuint64_t)c, 2, (cuuint64_t)wd / 2, 2, (cuuint64_t)n * h / 2};
  cuuint64_t strides[4] = {(cuuint64_t)c * 2, (cuuint64_t)2 * c * 2, (cuuint64_t)wd * c * 2, (cuuint64_t)2 * wd * c * 2};
  cuuint32_t box[5] = {64, 1, (cuuint32_t)bw, 1, (cuuint32_t)bh};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MFL_ARG(r == CUDA_SUCCESS, "conv2d dgrad: strided output tensor map failed with CUresult %d", (int)r);
  return 0;
}

static bool pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

}}  // namespace mfl::gemm

// ws (fp32 [rows, cols]) -> out (bf16, row pitch ld; out += ws when accumulate), ws re-zeroed; cols % 8 == 0
extern "C" int mfl_splitk_finalize(float* ws, void* out_bf16, int rows, int cols, int64_t ld, int accumulate, void* stream) {
  using namespace mfl::gemm;
  MFL_ARG(ws && out_bf16 && rows > 0 && cols > 0 && cols % 8 == 0 && ld % 8 == 0, "splitk_finalize: bad arguments");
  const int64_t vecs = (int64_t)rows * cols / 8;
  const int blocks = (int)std::min<int64_t>((vecs + 255) / 256, (int64_t)sm_count() * 4);
  splitk_finalize_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(ws, static_cast<__nv_bfloat16*>(out_bf16), rows, cols, ld,
                                                                  accumulate);
  return check(cudaGetLastError(), "splitk_finalize_kernel launch");
}

extern "C" size_t mfl_conv2d_wgrad_workspace_bytes(int c, int k, int r, int s) { return (size_t)k * r * s * c * sizeof(float); }

// dW[K,R,S,C] (+)= sum over output pixels of dY[p, k] * X[p + (r, s), c].  ws: zero-initialised fp32 scratch of
// mfl_conv2d_wgrad_workspace_bytes() bytes; it is handed back zeroed, so one buffer can serve every layer of a model on
// one stream.  NHWC bf16, C % 64 == 0, K % 8 == 0.
extern "C" int mfl_conv2d_wgrad_bf16(void* dw, const void* dy, const void* x, float* ws, int n, int h, int wd, int c, int k,
                                     int r, int s, int stride, int pad, int accumulate, void* stream) {
  using namespace mfl::gemm;
  MFL_ARG(dw && dy && x && ws && n > 0 && h > 0 && wd > 0, "conv2d wgrad: null operand or empty shape");
  MFL_ARG(c % 64 == 0 && k % 8 == 0, "conv2d wgrad: needs C %% 64 == 0 and K %% 8 == 0 (got C=%d K=%d)", c, k);
  MFL_ARG(stride >= 1 && stride <= 8 && pad >= 0 && r >= 1 && s >= 1 && r * s <= 49, "conv2d wgrad: bad geometry");
  const int oh = (h + 2 * pad - r) / stride + 1, ow = (wd + 2 * pad - s) / stride + 1;
  const int P = n * oh * ow, rsc = r * s * c;
  CUtensorMap ta, tb, tc;
  if (int rc = make_map_mn(&ta, dy, k, P, k)) return rc;                         // A: dY [P, K], M = K contiguous
  if (int rc = make_im2col_map(&tb, x, n, h, wd, c, -pad, -pad, pad - (s - 1), pad - (r - 1), stride, 64)) return rc;
  if (int rc = make_map(&tc, ws, k, rsc, rsc, 32, true)) return rc;             // fp32 workspace [K, RSC]
  const int bn = rsc > 128 ? 256 : (rsc > 64 ? 128 : 64);
  Epilogue ep{};
  ep.c = ws; ep.bias = nullptr; ep.ldc = rsc; ep.M = k; ep.N = rsc; ep.K = P;
  ep.c_fp32 = 1; ep.tma_store = 1; ep.a_mn = 1; ep.b_mn = 1; ep.mode = 3;
  ep.out_h = oh; ep.out_w = ow; ep.stride = stride; ep.pad = pad; ep.filt_s = s; ep.c_blocks = c / 64;
  const int tiles = ((rsc + bn - 1) / bn) * ((k + kBlockM - 1) / kBlockM);
  const int num_kb = (P + kBlockK - 1) / kBlockK;
  // enough (tile, split) work items for about two per SM; every split owns at least one k-block
  int splits = std::max(1, std::min(num_kb, (2 * sm_count() + tiles - 1) / tiles));
  ep.kb_per_split = (num_kb + splits - 1) / splits;
  ep.splits = (num_kb + ep.kb_per_split - 1) / ep.kb_per_split;
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = launch_gemm(ta, tb, tc, ep, bn, 0, true, st)) return rc;
  const int64_t vecs = (int64_t)k * rsc / 8;
  const int blocks = (int)std::min<int64_t>((vecs + 255) / 256, (int64_t)sm_count() * 4);
  splitk_finalize_kernel<<<blocks, 256, 0, st>>>(ws, static_cast<__nv_bfloat16*>(dw), k, rsc, rsc, accumulate);
  return check(cudaGetLastError(), "splitk_finalize_kernel launch");
}

// dX[N,H,W,C] = sum_{r,s,k} dY[n, (h + pad - r) / stride, (w + pad - s) / stride, k] * W[k, r, s, c]  (exact divisions only).
// stride 1: one launch over all R*S taps.  stride 2: one launch per (h mod 2, w mod 2) class - each class is a
// stride-1 correlation of dY with the taps of matching parity (for 3x3 / pad 1: 1, 2, 2 and 4 taps: no wasted math) -
// written through a strided 4-D tensor map.  NHWC bf16; K % 64 == 0, C % 8 == 0; stride 2 needs H/2 and W/2 powers of
// two (<= 32) so that 32-row store slabs tile the class grid.
extern "C" int mfl_conv2d_dgrad_bf16_ws(void* dx, const void* dy, const void* w, float* ws, size_t ws_bytes, int n,
                                        int h, int wd, int c, int k, int r, int s, int stride, int pad, void* stream);
extern "C" int mfl_conv2d_dgrad_bf16(void* dx, const void* dy, const void* w, int n, int h, int wd, int c, int k, int r,
                                     int s, int stride, int pad, void* stream) {
  return mfl_conv2d_dgrad_bf16_ws(dx, dy, w, nullptr, 0, n, h, wd, c, k, r, s, stride, pad, stream);
}
extern "C" int mfl_conv2d_dgrad_bf16_ws(void* dx, const void* dy, const void* w, float* ws, size_t ws_bytes, int n,
                                        int h, int wd, int c, int k, int r, int s, int stride, int pad, void* stream) {
  using namespace mfl::gemm;
  MFL_ARG(dx && dy && w && n > 0 && h > 0 && wd > 0, "conv2d dgrad: null operand or empty shape");
  MFL_ARG(k % 64 == 0 && c % 8 == 0, "conv2d dgrad: needs K %% 64 == 0 and C %% 8 == 0 (got C=%d K=%d)", c, k);
  MFL_ARG((stride == 1 || stride == 2) && r * s <= 16 && pad >= 0 && pad < r && pad < s, "conv2d dgrad: bad geometry");
  const int oh = (h + 2 * pad - r) / stride + 1, ow = (wd + 2 * pad - s) / stride + 1;
  MFL_ARG(oh > 0 && ow > 0, "conv2d dgrad: empty output");
  cudaStream_t st = (cudaStream_t)stream;
  const int rsc = r * s * c;
  CUtensorMap ta, tb, tc;
  if (int rc = make_map_mn(&tb, w, rsc, k, rsc)) return rc;   // B: W [K, RSC]: rows = reduction k, columns contiguous
  const int bn = pick_bn(n * (h / stride) * (wd / stride), c);
  if (stride == 2) {
    MFL_ARG(h % 2 == 0 && wd % 2 == 0 && h / 2 == oh && wd / 2 == ow && pow2(oh) && pow2(ow) && c % 64 == 0,
            "conv2d dgrad: stride 2 needs even H, W with H/2, W/2 powers of two and C %% 64 == 0");
    if (r * s < 4)  // fewer taps than classes (1x1 filters): the untouched classes of dX are zero
      MFL_CUDA(cudaMemsetAsync(dx, 0, (size_t)n * h * wd * c * 2, st));
  }
  // tap tables of the stride-parity classes (one class when stride == 1), heaviest class first
  struct Cls { int ph, pw, nt, dmin, dmax; int th[16], tw[16], tf[16]; };
  Cls cls[4];
  int ncls = 0;
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      Cls q{}; q.ph = ph; q.pw = pw; q.dmin = 1 << 20; q.dmax = -(1 << 20);
      for (int rr = 0; rr < r; ++rr) {
        if ((ph + pad - rr) % stride != 0) continue;
        const int dh = (ph + pad - rr) / stride;   // exact division, may be negative
        for (int ss = 0; ss < s; ++ss) {
          if ((pw + pad - ss) % stride != 0) continue;
          const int dw_ = (pw + pad - ss) / stride;
          q.th[q.nt] = dh; q.tw[q.nt] = dw_; q.tf[q.nt] = rr * s + ss; ++q.nt;
          q.dmin = std::min(q.dmin, std::min(dh, dw_)); q.dmax = std::max(q.dmax, std::max(dh, dw_));
        }
      }
      if (q.nt) cls[ncls++] = q;
    }
  MFL_ARG(ncls >= 1, "conv2d dgrad: no filter tap reaches the input");
  std::sort(cls, cls + ncls, [](const Cls& a, const Cls& b) { return a.nt > b.nt; });
  int dmin = cls[0].dmin;
  for (int q = 1; q < ncls; ++q) dmin = std::min(dmin, cls[q].dmin);
  const int cls_h = h / stride, cls_w = wd / stride;   // stride 2: H, W even (checked above); stride 1: H, W
  Epilogue ep{};
  ep.mode = 2; ep.kb_per_tap = k / 64; ep.c_in = c;
  int nt_total = 0;
  for (int q = 0; q < ncls; ++q) {
    ep.cls_tap0[q] = (int8_t)nt_total; ep.cls_ntaps[q] = (int8_t)cls[q].nt;
    ep.cls_ph[q] = (int8_t)cls[q].ph; ep.cls_pw[q] = (int8_t)cls[q].pw;
    for (int t = 0; t < cls[q].nt; ++t, ++nt_total) {
      MFL_ARG(nt_total < 16, "conv2d dgrad: more than 16 filter taps");
      // base pixel of class position (h', w') is (h' + dmin, w' + dmin); tap t reads base + (d_t - dmin)
      ep.tap_oh[nt_total] = (int8_t)(cls[q].th[t] - dmin); ep.tap_ow[nt_total] = (int8_t)(cls[q].tw[t] - dmin);
      ep.tap_w[nt_total] = (int8_t)cls[q].tf[t];
    }
  }
  ep.n_cls = ncls; ep.ntaps = nt_total;
  // im2col box of base positions: [dmin, dmin + cls - 1] per axis -> #positions = dim + upper - lower = cls
  if (int rc = make_im2col_map(&ta, dy, n, oh, ow, k, dmin, dmin, cls_w - ow + dmin, cls_h - oh + dmin, 1, kBlockM)) return rc;
  ep.c = dx; ep.ldc = c; ep.M = n * cls_h * cls_w; ep.N = c; ep.K = cls[0].nt * k;  // K of the heaviest class (grid sizing)
  ep.c_fp32 = 0; ep.tma_store = 1; ep.a_mn = 0; ep.b_mn = 1;
  ep.out_h = cls_h; ep.out_w = cls_w; ep.stride = 1; ep.pad = -dmin;  // the kernel subtracts `pad` from both axes
  bool split = false;
  if (stride == 1) {
    plan_splits(ep, bn, ws, ws_bytes);
    split = ep.splits > 1;
    if (split) {
      ep.c = ws; ep.c_fp32 = 1;
      if (int rc = make_map(&tc, ws, ep.M, c, c, 32, true)) return rc;
    } else if (int rc = make_map(&tc, dx, ep.M, c, c, 32, false)) return rc;
  } else {
    const int bw = std::min(cls_w, 32), bh = 32 / bw;
    MFL_ARG((n * cls_h) % bh == 0, "conv2d dgrad: N * H/2 must be a multiple of %d", bh);
    ep.store4d = 1; ep.box_w = bw; ep.box_h = bh; ep.cls_w = cls_w; ep.cls_h = cls_h;
    if (int rc = make_store5d_map(&tc, dx, n, h, wd, c, bw, bh)) return rc;
  }
  if (int rc = launch_gemm(ta, tb, tc, ep, bn, 0, split, st)) return rc;
  if (split)
    if (int rc = finalize_splits(ep, ws, dx, c, st)) return rc;
  return 0;
}
