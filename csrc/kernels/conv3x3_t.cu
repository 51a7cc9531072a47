// K15t: 3x3 / stride-1 / padding-1 convolution, forward and data gradient, "pixels as N" formulation on tcgen05
// tensor cores.  NHWC bf16, fp32 accumulation in TMEM.
//
//   forward : Y[n,h,w,k] = sum_{r,s,c} X[n, h+r-1, w+s-1, c] * W[k,r,s,c]
//   dgrad   : dX[n,h,w,c] = sum_{r,s,k} dY[n, h+1-r, w+1-s, k] * W[k,r,s,c]     (same kernel, taps mirrored,
//             the filter tensor read MN-major, so no flipped / transposed copy is made)
//
// Why a third formulation.  With the output pixels as the MMA's M dimension (K11 im2col, K15 halo) a 64- or 128-channel
// layer issues M128 x N64/128 instructions whose 4 KB A tile is re-read from shared memory for every instruction; the
// tensor pipe then idles behind the shared-memory port (38 % busy, profiles/README.md).  Here the operands swap roles:
//
//   D^T[out-ch (M = 128), pixels (N <= 256)] += Wtap[out-ch, 64 red-ch] (A, 16 KB per tap and 64-channel block)
//                                               * Halo[pixels + tap shift, 64 red-ch] (B)
//
//   * B is ONE zero-padded halo tile per (pixel tile, 64-channel block): (TH + 2) rows x (W + 2) pixels x 128 B,
//     loaded by a tiled 4-D TMA whose out-of-bounds elements are the padding; the nine taps read it through
//     descriptors shifted by (r * (W + 2) + s) pixel rows (the 128-byte swizzle is a function of the absolute
//     shared-memory address, so a descriptor may start at any 128-byte row);
//   * N = TH * (W + 2) padded pixels per instruction (up to 256): M128 x N x K16 is compute-bound (A 4 KB + B N * 32 B
//     per instruction against N / 2 tensor cycles), and the halo is read from L2 once instead of nine times;
//   * A streams through a ring (16 KB per tap and channel block), B through a second ring;
//   * the accumulator is double-buffered in TMEM (2 x N <= 512 columns): the epilogue of pixel tile i overlaps the main
//     loop of tile i + 1.  The epilogue reads D^T (lane = channel, column = pixel), swaps odd/even pixels between
//     neighbouring lanes with one shuffle and stores channel pairs, so every warp writes 64 contiguous bytes per
//     pixel; junk columns (w >= W) and rows past the image are simply not stored.
//
//   warp 0 / lane 0 : TMA producer (halo ring + weight ring)        warp 1 : MMA issuer (one elected lane)
//   warp 2          : TMEM allocator                                warps 4-11 : epilogue (two warps per TMEM lane
//                                                                   quarter, each takes every other 32-column chunk)
// Reference counterpart: the framework convolutions under model.fit (keras_model_ops.py:156).
#include <cuda.h>

#include <algorithm>

#include "common.cuh"
#include "tcgen05_common.cuh"

namespace mfl {
namespace conv3t {

using namespace mfl::gemm;

constexpr int kThreads = 384;                // warps 0-2: producer / MMA / TMEM; 3 idle; 4-11: epilogue (2 per lane quarter)
constexpr int kAStages = 6;                 // weight tiles in flight (16 KB each)
constexpr int kABytes16 = 128 * 64 * 2;     // one weight tile: 128 out-channels x 64 reduction channels
constexpr int kMaxBStages = 4;

struct Params {
  __nv_bfloat16* y;
  int N, H, W;
  int Cred, Cout;          // reduction channels (of the halo operand) / output channels
  int dgrad;
  int w_cols_per_tap;      // column pitch of one tap in the forward filter matrix [K_fwd, 9 * C_fwd] (= C_fwd)
  int TH, Wp, NN;          // output rows per tile, padded width, MMA N (multiple of 16, <= 256)
  int halo_bytes;          // shared memory per halo stage (multiple of 1024)
  int box_bytes;           // bytes one halo TMA delivers
  int b_stages;
  int tiles_per_img, num_ptiles, mtiles, cblocks;
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

// kind::f16, bf16 x bf16 -> fp32, M = 128, N = n; bit 15 = A is MN-major
__device__ __forceinline__ uint32_t idesc_n(int n, bool a_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn ? (1u << 15) : 0u) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

__global__ void __launch_bounds__(kThreads, 1)
conv3x3_t_kernel(const __grid_constant__ CUtensorMap tmap_act, const __grid_constant__ CUtensorMap tmap_w,
                 const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* a_ring = smem;                                            // kAStages x 16 KB
  uint8_t* b_ring = smem + (size_t)kAStages * kABytes16;             // b_stages x halo_bytes
  uint64_t* a_full = reinterpret_cast<uint64_t*>(b_ring + (size_t)p.b_stages * p.halo_bytes);
  uint64_t* a_empty = a_full + kAStages;
  uint64_t* b_full = a_empty + kAStages;
  uint64_t* b_empty = b_full + kMaxBStages;
  uint64_t* tmem_full = b_empty + kMaxBStages;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;          // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  // output offset (in pixels) of every padded pixel of a tile, -1 for the junk columns; rows are checked per tile
  int* pix_off = reinterpret_cast<int*>(tmem_ptr_smem + 4);          // [256]
  for (int q = threadIdx.x; q < 256; q += kThreads) {
    const int trow = q / p.Wp, wcol = q - trow * p.Wp;
    pix_off[q] = (q < p.NN && wcol < p.W && trow < p.TH) ? ((trow << 16) | wcol) : -1;
  }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_items = p.num_ptiles * p.mtiles;   // item = pixel tile * mtiles + out-channel tile

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_act) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_w) : "memory");
    for (int s = 0; s < kAStages; ++s) { mbar_init(a_full + s, 1); mbar_init(a_empty + s, 1); }
    for (int s = 0; s < p.b_stages; ++s) { mbar_init(b_full + s, 1); mbar_init(b_empty + s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(tmem_full + b, 1); mbar_init(tmem_empty + b, 256); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  } else if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer: per item and channel block one halo tile, then its nine weight tiles =====
      uint32_t na = 0, nb = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const int pt = item / p.mtiles, mt = item - pt * p.mtiles;
        const int img = pt / p.tiles_per_img, h0 = (pt - img * p.tiles_per_img) * p.TH;
        const int m0 = mt * 128;
        for (int cb = 0; cb < p.cblocks; ++cb, ++nb) {
          const int sb = nb % p.b_stages;
          mbar_wait(b_empty + sb, ((nb / p.b_stages) & 1) ^ 1);
          mbar_expect_tx(b_full + sb, (uint32_t)p.box_bytes);
          tma_load_4d(b_ring + (size_t)sb * p.halo_bytes, &tmap_act, b_full + sb, cb * 64, -1, h0 - 1, img);
          for (int tap = 0; tap < 9; ++tap, ++na) {
            const int sa = na % kAStages;
            mbar_wait(a_empty + sa, ((na / kAStages) & 1) ^ 1);
            mbar_expect_tx(a_full + sa, kABytes16);
            uint8_t* dst = a_ring + (size_t)sa * kABytes16;
            if (p.dgrad) {
              // halo tap `tap` multiplies with forward tap 8 - tap; rows = reduction (k) block cb, columns = out (c)
              tma_load_2d(dst, &tmap_w, a_full + sa, (8 - tap) * p.w_cols_per_tap + m0, cb * 64);
              tma_load_2d(dst + 8192, &tmap_w, a_full + sa, (8 - tap) * p.w_cols_per_tap + m0 + 64, cb * 64);
            } else {
              tma_load_2d(dst, &tmap_w, a_full + sa, tap * p.w_cols_per_tap + cb * 64, m0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {  // ===== MMA issuer =====
    const uint32_t idesc = idesc_n(p.NN, p.dgrad != 0);
    const uint64_t a_kstep = p.dgrad ? 128 : 2;   // MN-major: 16 reduction rows of 128 B; K-major: 32 B in the swizzle row
    const uint32_t a_lbo = p.dgrad ? 512 : 1;     // MN-major: the two 64-wide blocks of out-channels are 8 KB apart
    uint32_t na = 0, nb = 0, t = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++t) {
      const uint32_t buf = t & 1, use = t >> 1;
      mbar_wait(tmem_empty + buf, (use & 1) ^ 1);
      tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + buf * 256;
      for (int cb = 0; cb < p.cblocks; ++cb, ++nb) {
        const int sb = nb % p.b_stages;
        mbar_wait(b_full + sb, (nb / p.b_stages) & 1);
        tcgen05_fence_after();
        const uint64_t b0 = make_sw128_desc(smem_u32(b_ring + (size_t)sb * p.halo_bytes));
        for (int tap = 0; tap < 9; ++tap, ++na) {
          const int sa = na % kAStages;
          mbar_wait(a_full + sa, (na / kAStages) & 1);
          tcgen05_fence_after();
          const uint64_t a0 = make_sw128_desc(smem_u32(a_ring + (size_t)sa * kABytes16), a_lbo);
          const int r = tap / 3, sx = tap - 3 * r;
          const uint64_t bt = b0 + (uint64_t)((r * p.Wp + sx) * 8);   // shifted by whole pixel rows (128 B = 8 units)
          if (elect_one()) {
            if (cb == 0 && tap == 0) umma_f16_c<false>(tmem_d, a0, bt, idesc);
            else umma_f16_c<true>(tmem_d, a0, bt, idesc);
#pragma unroll
            for (int k = 1; k < 4; ++k) umma_f16_c<true>(tmem_d, a0 + (uint64_t)k * a_kstep, bt + (uint64_t)(k * 2), idesc);
            umma_commit(a_empty + sa);
            if (tap == 8) umma_commit(b_empty + sb);
          }
          __syncwarp();
        }
      }
      if (elect_one()) umma_commit(tmem_full + buf);
      __syncwarp();
    }
  } else if (warp >= 4) {  // ===== epilogue: D^T[channel lanes, pixel columns] -> NHWC =====
    const int wq = warp & 3, half = (warp - 4) >> 2;
    uint32_t t = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++t) {
      const int pt = item / p.mtiles, mt = item - pt * p.mtiles;
      const int img = pt / p.tiles_per_img, h0 = (pt - img * p.tiles_per_img) * p.TH;
      const uint32_t buf = t & 1, use = t >> 1;
      mbar_wait(tmem_full + buf, use & 1);
      tcgen05_fence_after();
      const int ch_warp = mt * 128 + wq * 32;          // first channel of this warp's lane quarter
      // after the pair swap an even lane stores (its channel, its neighbour's) for the even pixel, an odd lane the
      // same pair for the odd pixel
      const int ch_pair = ch_warp + (lane & ~1);
      const bool live = ch_pair < p.Cout;
      const int rows_ok = min(p.TH, p.H - h0);         // ragged last tile of an image
      __nv_bfloat16* out_tile = p.y + ((size_t)img * p.H + h0) * p.W * p.Cout + ch_pair;
      const int odd = lane & 1;
#pragma unroll 1
      for (int c0 = half * 32; c0 < p.NN; c0 += 64) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(wq * 32) << 16) + buf * 256 + (uint32_t)c0, v);
        if (ch_warp >= p.Cout) continue;               // warp-uniform: rows of zero-filled weights
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          // a lane keeps the pixel of its own parity and gets its neighbour's value for that pixel
          const uint32_t mine = odd ? v[j + 1] : v[j];
          const uint32_t got = __shfl_xor_sync(0xffffffffu, odd ? v[j] : v[j + 1], 1);
          const int po = pix_off[c0 + j + odd];
          if (po >= 0 && (po >> 16) < rows_ok && live) {
            const __nv_bfloat162 pk = __floats2bfloat162_rn(__uint_as_float(odd ? got : mine), __uint_as_float(odd ? mine : got));
            *reinterpret_cast<__nv_bfloat162*>(out_tile + ((size_t)(po >> 16) * p.W + (po & 0xffff)) * p.Cout) = pk;
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
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

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

static bool plan(Params& p, int n, int h, int wd, int cred, int cout) {
  p.N = n; p.H = h; p.W = wd; p.Cred = cred; p.Cout = cout;
  p.Wp = wd + 2;
  if (cred % 64 != 0 || cout % 8 != 0 || p.Wp > 128) return false;
  // rows per tile: minimise the padded pixels processed per image, ceil(H / TH) * ceil16(TH * Wp); ties -> taller tile
  int best_th = 0;
  long best = 1L << 60;
  for (int th = 1; th <= h && th * p.Wp <= 256; ++th) {
    const int nn = (th * p.Wp + 15) / 16 * 16;
    if (nn > 256) break;
    const long cost = (long)((h + th - 1) / th) * nn;
    if (cost <= best) { best = cost; best_th = th; }
  }
  if (!best_th) return false;
  p.TH = best_th;
  p.NN = (p.TH * p.Wp + 15) / 16 * 16;
  p.box_bytes = (p.TH + 2) * p.Wp * 128;
  // the last tap of the last (junk) columns reads up to NN + 2 * Wp + 2 rows
  const int need_rows = std::max((p.TH + 2) * p.Wp, p.NN + 2 * p.Wp + 2);
  p.halo_bytes = (need_rows * 128 + 1023) / 1024 * 1024;
  const size_t fixed = 1024 + (size_t)kAStages * kABytes16 + 512 + 1024;
  p.b_stages = (int)std::min<size_t>(kMaxBStages, (227 * 1024 - fixed) / p.halo_bytes);
  if (p.b_stages < 2) return false;
  p.tiles_per_img = (h + p.TH - 1) / p.TH;
  p.num_ptiles = n * p.tiles_per_img;
  p.mtiles = (cout + 127) / 128;
  p.cblocks = cred / 64;
  return true;
}

static int launch(void* y, const void* act, const void* w, int n, int h, int wd, int cred, int cout, int dgrad, void* stream) {
  MFL_ARG(y && act && w, "conv3x3 (pixels-as-N): null operand");
  Params p;
  MFL_ARG(n > 0 && h > 0 && wd > 0 && plan(p, n, h, wd, cred, cout),
          "conv3x3 (pixels-as-N): unsupported shape n=%d h=%d w=%d reduction channels=%d output channels=%d", n, h, wd, cred, cout);
  MFL_ARG(!dgrad || cout % 64 == 0, "conv3x3 dgrad: the output channels must be a multiple of 64 (got %d)", cout);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.dgrad = dgrad;
  // the filter matrix is always the forward one [K_fwd, 9 * C_fwd]: forward reduces over C_fwd, dgrad over K_fwd
  const int c_fwd = dgrad ? cout : cred, k_fwd = dgrad ? cred : cout;
  p.w_cols_per_tap = c_fwd;
  bind_context();
  EncodeTiledFn fn = encode_fn();
  MFL_ARG(fn != nullptr, "conv3x3: cuTensorMapEncodeTiled is not available from this driver");
  CUtensorMap ta, tw;
  {
    cuuint64_t dims[4] = {(cuuint64_t)cred, (cuuint64_t)wd, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)cred * 2, (cuuint64_t)wd * cred * 2, (cuuint64_t)h * wd * cred * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wp, (cuuint32_t)(p.TH + 2), 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = fn(&ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(act), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MFL_ARG(r == CUDA_SUCCESS, "conv3x3: activation tensor map failed with CUresult %d", (int)r);
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)9 * c_fwd, (cuuint64_t)k_fwd};
    cuuint64_t strides[1] = {(cuuint64_t)9 * c_fwd * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)(dgrad ? 64 : 128)};   // dgrad: [64 k rows, 64 c] MN-major blocks
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(&tw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MFL_ARG(r == CUDA_SUCCESS, "conv3x3: filter tensor map failed with CUresult %d", (int)r);
  }
  const size_t smem = 1024 + (size_t)kAStages * kABytes16 + (size_t)p.b_stages * p.halo_bytes + 512 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    MFL_CUDA(cudaFuncSetAttribute(conv3x3_t_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  const int grid = std::min(sm_count(), p.num_ptiles * p.mtiles);
  conv3x3_t_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(ta, tw, p);
  return check(cudaGetLastError(), "conv3x3_t_kernel launch");
}

}  // namespace conv3t
}  // namespace mfl

extern "C" int mfl_conv3x3_t_supported(int n, int h, int wd, int cred, int cout) {
  mfl::conv3t::Params p;
  return n > 0 && h > 0 && wd > 0 && mfl::conv3t::plan(p, n, h, wd, cred, cout) ? 1 : 0;
}

// y[N,H,W,K] = conv3x3(x[N,H,W,C], w[K,3,3,C]), stride 1, zero padding 1.  C % 64 == 0, K % 8 == 0.
extern "C" int mfl_conv3x3_t_fprop_bf16(void* y, const void* x, const void* w, int n, int h, int wd, int c, int k,
                                        void* stream) {
  return mfl::conv3t::launch(y, x, w, n, h, wd, c, k, 0, stream);
}

// dX[N,H,W,C] of the same convolution from dY[N,H,W,K]; w is the FORWARD filter.  K % 64 == 0, C % 64 == 0.
extern "C" int mfl_conv3x3_t_dgrad_bf16(void* dx, const void* dy, const void* w, int n, int h, int wd, int c, int k,
                                        void* stream) {
  return mfl::conv3t::launch(dx, dy, w, n, h, wd, /*reduction=*/k, /*output=*/c, 1, stream);
}
