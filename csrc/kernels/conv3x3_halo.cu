// K15: weight-stationary 3x3 / stride-1 / pad-1 convolution (forward, and data gradient with transformed
// filters) on tcgen05 tensor cores, NHWC bf16, fp32 accumulation in TMEM.
//
// K11 (im2col TMA) re-fetches the A operand through L2 once per filter tap: 9 x 128 pixels x 128 B per
// 64-channel block and output tile, which is what bounds it on the L2-resident ResNet shapes
// (profiles/README.md).  Here each CTA
//   * keeps ALL filter taps of its 64 output channels in shared memory for its whole lifetime
//     (9 * C/64 tiles of 64 x 64 bf16, one TMA burst at start-up), and
//   * loads ONE halo tile per (output tile, 64-channel block): (TH + 2) rows x (W + 2) pixels x 64
//     channels by a tiled 4-D TMA whose out-of-bounds elements are zero-filled (that is the padding).
// The nine taps are nine tcgen05.mma sequences that read the SAME halo tile through shared-memory
// descriptors shifted by (r * (W + 2) + s) pixel rows of 128 B: with 128-byte swizzling the XOR pattern is a
// function of the absolute shared-memory address, so a descriptor may start at any 128-byte row.  The MMA
// M dimension runs over TH * (W + 2) consecutive halo pixels; the two junk columns per image row (and the
// rows past TH) are computed and never stored.
//
//   warp 0 / lane 0 : TMA producer (filters once, then a 2-deep ring of halo tiles)
//   warp 1 / lane 0 : MMA issuer, 36 x tcgen05.mma (M128 N64 K16) per 64-channel block
//   warp 2          : TMEM allocator (2 x 64 fp32 columns: the epilogue of tile i overlaps tile i + 1)
//   warps 4-7       : epilogue, TMEM -> registers -> bf16 -> global, plus optional per-channel sum / sum of
//                     squares of the stored outputs (the BatchNorm statistics of the next layer, K14)
//
// Reference counterpart: the framework convolutions under model.fit (keras_model_ops.py:156).
#include <cuda.h>

#include <algorithm>

#include "common.cuh"
#include "tcgen05_common.cuh"

namespace mfl {
namespace conv3 {

using namespace mfl::gemm;

constexpr int kThreads = 256;
constexpr int kBN = 64;                          // output channels per CTA
constexpr int kBTileBytes = kBN * kBlockK * 2;   // one tap x one 64-channel block of filters: 8 KB
constexpr int kMaxAStages = 6;               // halo ring depth is chosen at launch from the shared memory left over
constexpr int kMaxCBlocks = 2;                   // C <= 128: 18 filter tiles = 144 KB stay resident
constexpr int kTmemCols = 2 * kBN;

struct Params {
  __nv_bfloat16* y;
  float* stats;          // optional [2][K] fp32: atomically accumulated sum / sum of squares per output channel
  int N, H, W, C, K;
  int TH, Wp;            // output rows per tile, padded width W + 2
  int halo_bytes;        // shared-memory bytes reserved per halo stage (multiple of 1024)
  int box_bytes;         // bytes one halo TMA delivers: (TH + 2) * Wp * 128
  int tiles_per_img, num_m_tiles, n_tiles;
  int a_stages;          // depth of the halo ring (2..kMaxAStages)
  // data-gradient mode: the halo tile is dY (C = its channels, the reduction), K = channels of dX; the filter tensor
  // is the forward one, read MN-major ([64 k, 64 c] boxes: reduction rows, output channels contiguous) and with the
  // taps mirrored, so no transformed copy of the filters is ever made
  int dgrad;
  int w_cols_per_tap;    // column pitch of one tap in the [rows, 9 * cols] filter matrix
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

__global__ void __launch_bounds__(kThreads, 1)
conv3x3_halo_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                    const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int cblocks = p.C / kBlockK;
  const int kAStages = p.a_stages;
  uint8_t* halo = smem;                                             // a_stages x halo_bytes
  uint8_t* filt = smem + (size_t)kAStages * p.halo_bytes;           // 9 * cblocks tiles of 8 KB, 1024-aligned
  uint64_t* a_full = reinterpret_cast<uint64_t*>(filt + (size_t)9 * cblocks * kBTileBytes);
  uint64_t* a_empty = a_full + kMaxAStages;
  uint64_t* w_bar = a_empty + kMaxAStages;
  uint64_t* tmem_full = w_bar + 1;    // [2]
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x % p.n_tiles;
  const int m_first = blockIdx.x / p.n_tiles, m_step = gridDim.x / p.n_tiles;
  const int n0 = n_tile * kBN;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_w) : "memory");
    for (int s = 0; s < kAStages; ++s) {
      mbar_init(a_full + s, 1);
      mbar_init(a_empty + s, 1);
    }
    mbar_init(w_bar, 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full + b, 1);
      mbar_init(tmem_empty + b, 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  } else if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      mbar_expect_tx(w_bar, (uint32_t)(9 * cblocks * kBTileBytes));
      for (int cb = 0; cb < cblocks; ++cb)
        for (int tap = 0; tap < 9; ++tap)
          if (p.dgrad)  // tile for halo tap `tap` = forward tap 8 - tap, rows = reduction block cb, columns = n0..n0+63
            tma_load_2d(filt + (size_t)(cb * 9 + tap) * kBTileBytes, &tmap_w, w_bar, (8 - tap) * p.w_cols_per_tap + n0,
                        cb * kBlockK);
          else
            tma_load_2d(filt + (size_t)(cb * 9 + tap) * kBTileBytes, &tmap_w, w_bar, tap * p.C + cb * kBlockK, n0);
      uint32_t it = 0;
      for (int mt = m_first; mt < p.num_m_tiles; mt += m_step) {
        const int img = mt / p.tiles_per_img, h0 = (mt % p.tiles_per_img) * p.TH;
        for (int cb = 0; cb < cblocks; ++cb, ++it) {
          const int s = it % kAStages;
          mbar_wait(a_empty + s, ((it / kAStages) & 1) ^ 1);
          mbar_expect_tx(a_full + s, (uint32_t)p.box_bytes);
          tma_load_4d(halo + (size_t)s * p.halo_bytes, &tmap_x, a_full + s, cb * kBlockK, -1, h0 - 1, img);
        }
      }
    }
  } else if (warp == 1) {  // ===== MMA issuer: the whole warp runs the loop, one elected lane issues =====
    const uint32_t idesc = make_idesc<kBN>() | (p.dgrad ? (1u << 16) : 0u);  // B is MN-major in data-gradient mode
    const uint64_t b_kstep = p.dgrad ? 128 : 2;  // 16 reduction rows of 128 B vs 32 B inside the swizzle row
    mbar_wait(w_bar, 0);
    uint32_t it = 0, t = 0;
    for (int mt = m_first; mt < p.num_m_tiles; mt += m_step, ++t) {
      const uint32_t buf = t & 1, use = t >> 1;
      mbar_wait(tmem_empty + buf, (use & 1) ^ 1);
      tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + buf * kBN;
      for (int cb = 0; cb < cblocks; ++cb, ++it) {
        const int s = it % kAStages;
        mbar_wait(a_full + s, (it / kAStages) & 1);
        tcgen05_fence_after();
        const uint64_t a0 = make_sw128_desc(smem_u32(halo + (size_t)s * p.halo_bytes));
        const uint64_t b0 = make_sw128_desc(smem_u32(filt + (size_t)cb * 9 * kBTileBytes));
        const uint64_t row = (uint64_t)((p.Wp * 128) >> 4);  // one halo row of pixels, in descriptor units of 16 B
        const uint64_t px = 8;                                // one pixel (128 B)
        if (elect_one()) {
          if (cb == 0) umma_f16_c<false>(tmem_d, a0, b0, idesc);
          else umma_f16_c<true>(tmem_d, a0, b0, idesc);
#pragma unroll
          for (int q = 1; q < 36; ++q) {  // q = tap * 4 + k: 9 taps x 4 K-steps of 16 channels, straight-line
            const int tap = q >> 2, k = q & 3, r = tap / 3, sx = tap - 3 * r;
            umma_f16_c<true>(tmem_d, a0 + (uint64_t)r * row + (uint64_t)sx * px + (uint64_t)(k * 2),
                             b0 + (uint64_t)(tap * (kBTileBytes >> 4)) + (uint64_t)k * b_kstep, idesc);
          }
          umma_commit(a_empty + s);
        }
        __syncwarp();
      }
      if (elect_one()) umma_commit(tmem_full + buf);
      __syncwarp();
    }
  } else if (warp >= 4) {  // ===== epilogue =====
    const int wq = warp & 3;
    const int m = wq * 32 + lane;           // halo-pixel row of the MMA tile this thread owns
    const int t_row = m / p.Wp, col = m - t_row * p.Wp;
    uint32_t t = 0;
    for (int mt = m_first; mt < p.num_m_tiles; mt += m_step, ++t) {
      const int img = mt / p.tiles_per_img, h0 = (mt % p.tiles_per_img) * p.TH;
      const uint32_t buf = t & 1, use = t >> 1;
      mbar_wait(tmem_full + buf, use & 1);
      tcgen05_fence_after();
      const bool ok = t_row < p.TH && col < p.W && h0 + t_row < p.H;
      __nv_bfloat16* out = p.y + ((size_t)((size_t)img * p.H + h0 + t_row) * p.W + col) * p.K + n0;
#pragma unroll 1
      for (int c = 0; c < kBN; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(wq * 32) << 16) + buf * kBN + (uint32_t)c, r);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = ok ? __uint_as_float(r[j]) : 0.f;
        if (ok) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) *reinterpret_cast<uint4*>(out + c + j) = VecTraits<MFL_BFLOAT16>::pack(v + j);
        }
        if (p.stats) {  // per-channel sums over the 32 rows of this warp (bf16-rounded values, as stored)
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float x = __bfloat162float(__float2bfloat16_rn(v[j]));
            float s1 = x, s2 = x * x;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
              s1 += __shfl_xor_sync(0xffffffffu, s1, o);
              s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            }
            if (lane == j) {
              atomicAdd(p.stats + n0 + c + j, s1);
              atomicAdd(p.stats + p.K + n0 + c + j, s2);
            }
          }
        }
      }
      tcgen05_fence_before();
      mbar_arrive(tmem_empty + buf);
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

}  // namespace conv3
}  // namespace mfl

// ---- host side ---------------------------------------------------------------------------------------
namespace mfl { namespace conv3 {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(q);
  }
  return fn;
}

}}  // namespace mfl::conv3

extern "C" int mfl_conv3x3_supported(int n, int h, int wd, int c, int k) {
  using namespace mfl::conv3;
  if (n <= 0 || h <= 0 || wd <= 0) return 0;
  if (c % 64 != 0 || c / 64 > kMaxCBlocks || k % kBN != 0) return 0;
  if (wd + 2 > 128 || wd % 2 != 0) return 0;
  return 1;
}

// y[N,H,W,K] = conv3x3(x[N,H,W,C], w[K,3,3,C]), stride 1, zero padding 1, NHWC bf16.
// stats (optional): [2][K] fp32, ADDED to: per-channel sum and sum of squares of the stored outputs.
static int conv3x3_impl(void* y, const void* x, const void* w, float* stats, int n, int h, int wd, int c, int k,
                        int dgrad, void* stream);

extern "C" int mfl_conv3x3_bf16(void* y, const void* x, const void* w, float* stats, int n, int h, int wd, int c, int k,
                                void* stream) {
  return conv3x3_impl(y, x, w, stats, n, h, wd, c, k, 0, stream);
}

// dX[N,H,W,C] of y = conv3x3(x, w[K,3,3,C]) (stride 1, padding 1) from dY[N,H,W,K]: the same kernel with dY as the halo
// operand and the forward filters read MN-major with mirrored taps.  Shapes: mfl_conv3x3_supported(n, h, w, K, C).
extern "C" int mfl_conv3x3_dgrad_bf16(void* dx, const void* dy, const void* w, int n, int h, int wd, int c, int k,
                                      void* stream) {
  return conv3x3_impl(dx, dy, w, nullptr, n, h, wd, /*reduction channels=*/k, /*output channels=*/c, 1, stream);
}

static int conv3x3_impl(void* y, const void* x, const void* w, float* stats, int n, int h, int wd, int c, int k,
                        int dgrad, void* stream) {
  using namespace mfl;
  using namespace mfl::conv3;
  MFL_ARG(y && x && w, "conv3x3: null operand");
  MFL_ARG(mfl_conv3x3_supported(n, h, wd, c, k), "conv3x3: unsupported shape n=%d h=%d w=%d c=%d k=%d (need c in {64,128}, "
          "k %% 64 == 0, even w <= 126)", n, h, wd, c, k);
  bind_context();  // the driver-API tensor-map calls need the primary context bound to this thread
  EncodeTiledFn fn = encode_fn();
  MFL_ARG(fn != nullptr, "conv3x3: cuTensorMapEncodeTiled is not available from this driver");
  Params p;
  p.y = static_cast<__nv_bfloat16*>(y);
  p.stats = stats;
  p.N = n; p.H = h; p.W = wd; p.C = c; p.K = k;
  p.dgrad = dgrad;
  p.w_cols_per_tap = dgrad ? k : c;   // the filter matrix is always the forward one: [K_fwd, 9 * C_fwd]
  p.Wp = wd + 2;
  p.TH = std::max(1, std::min(h, 128 / p.Wp));
  p.box_bytes = (p.TH + 2) * p.Wp * 128;
  const int need = std::max(p.box_bytes, (2 * p.Wp + 2 + 128) * 128);
  p.halo_bytes = (need + 1023) / 1024 * 1024;
  p.tiles_per_img = (h + p.TH - 1) / p.TH;
  p.num_m_tiles = n * p.tiles_per_img;
  p.n_tiles = k / kBN;
  const int cblocks = c / 64;
  const size_t fixed = 1024 + (size_t)9 * cblocks * kBTileBytes + 256;
  p.a_stages = (int)std::min<size_t>(kMaxAStages, (227 * 1024 - fixed) / p.halo_bytes);
  MFL_ARG(p.a_stages >= 2, "conv3x3: not enough shared memory for two halo stages (%d bytes each)", p.halo_bytes);
  const size_t smem = fixed + (size_t)p.a_stages * p.halo_bytes;

  CUtensorMap tx, tw;
  {  // activations: 4-D {C, W, H, N}, box {64, W + 2, TH + 2, 1}; out-of-bounds elements read as zero = the padding
    cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)wd, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)wd * c * 2, (cuuint64_t)h * wd * c * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wp, (cuuint32_t)(p.TH + 2), 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = fn(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MFL_ARG(r == CUDA_SUCCESS, "conv3x3: activation tensor map failed with CUresult %d", (int)r);
  }
  {  // filters as a [K, 9 * C] row-major matrix (forward orientation), box {64, 64}
    const int wc = dgrad ? k : c, wk = dgrad ? c : k;
    cuuint64_t dims[2] = {(cuuint64_t)9 * wc, (cuuint64_t)wk};
    cuuint64_t strides[1] = {(cuuint64_t)9 * wc * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)kBN};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(&tw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MFL_ARG(r == CUDA_SUCCESS, "conv3x3: filter tensor map failed with CUresult %d", (int)r);
  }
  int grid = std::min(sm_count(), p.num_m_tiles * p.n_tiles);
  grid = std::max(p.n_tiles, grid / p.n_tiles * p.n_tiles);  // every CTA keeps one filter slice: multiple of n_tiles
  static bool attr_set = false;
  if (!attr_set) {
    MFL_CUDA(cudaFuncSetAttribute(conv3x3_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  conv3x3_halo_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(tx, tw, p);
  return check(cudaGetLastError(), "conv3x3_halo_kernel launch");
}
