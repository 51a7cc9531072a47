// K15w: weight gradient of a 3x3 / stride-1 / padding-1 convolution on tcgen05 tensor cores, NHWC bf16.
//
//   dW[k, r, s, c] (+)= sum over (n, oh, ow) of dY[n, oh, ow, k] * X[n, oh + r - 1, ow + s - 1, c]
//
// As a GEMM this is D[k, (r,s,c)] = A[k, p] * B[(r,s,c), p] with the reduction p over the N*H*W output pixels, both
// operands "MN-major" (in memory the pixel index is the slow one).  The im2col formulation (K11, mode 3) fetches the
// activations nine times - once per filter tap - and is bound by that L2 traffic.  Here:
//
//   * a work item owns one filter row r, one 64-channel block of X, one 128-row block of output channels, and a
//     contiguous range of "chunks" (TH image rows of NI images); its accumulator D[128 k, 3 taps x 64 c] = 192 fp32
//     columns stays in TMEM over the whole range;
//   * per chunk ONE tiled TMA brings dY ({64 k, W + 2, TH, NI}: the two extra columns are out of bounds and read as
//     zero) and ONE brings the rows of X the filter row needs ({64 c, W + 2, TH, NI} starting at column -1 and row
//     h0 + r - 1: out-of-bounds = the zero padding).  Both land as [pixel rows of 128 B], swizzle 128 B;
//   * the three taps s = 0, 1, 2 of the filter row are ONE tcgen05.mma with N = 192: the B descriptor's leading-dimension
//     byte offset - the distance between consecutive 64-column blocks - is set to 128 bytes, ONE pixel row, so block
//     s of the instruction reads the same tile shifted by s pixels.  With 128-byte swizzling the XOR pattern is a
//     function of the absolute shared-memory address, so overlapping, row-shifted blocks read consistently.
//     The reduction index runs over the padded pixel grid t * (W + 2) + w; the junk positions w >= W contribute
//     nothing because dY is zero there.
//   * M128 x N192 x K16 costs 96 tensor-core cycles and reads 4 KB + 6 KB of shared memory: compute-bound, unlike the
//     N = 64 instructions (one per tap) a per-tap formulation would issue.
//   * partial results of the work items that share an output tile are combined with TMA reduce-add stores into a
//     zero-initialised fp32 workspace; mfl_splitk_finalize adds it into the bf16 gradient and re-zeroes it.
//
//   warp 0 / lane 0 : TMA producer          warp 1 : MMA issuer (one elected lane)
//   warp 2          : TMEM allocator        warps 4-7 : epilogue (TMEM -> swizzled smem slabs -> TMA reduce-add)
// Reference counterpart: the framework convolution backward under model.fit (keras_model_ops.py:156).
#include <cuda.h>

#include <algorithm>

#include "common.cuh"
#include "tcgen05_common.cuh"

extern "C" int mfl_splitk_finalize(float* ws, void* out_bf16, int rows, int cols, int64_t ld, int accumulate, void* stream);

namespace mfl {
namespace wgrad3 {

using namespace mfl::gemm;

constexpr int kThreads = 256;
constexpr int kN = 192;            // 3 taps x 64 channels
constexpr int kTmemCols = 256;
constexpr int kTailRows = 8;       // zeroed rows after every tile: the s = 2 shift reads two rows past the reduction
constexpr int kMaxStages = 4;
constexpr int kSlabBytes = 4 * 2 * 4096;  // 4 epilogue warps x 2 fp32 slabs of 32 rows x 128 B; they reuse stage 0, which
                                          // is free once the last MMA has retired (the epilogue waits for exactly that)
constexpr int kMaxRows = 272;             // pixel rows per chunk: two 3-block stages of this size fit in 227 KB

struct Params {
  int N, H, W, C, K;
  int TH, NI, Wp;
  int rows, rows16;        // pixel rows one TMA delivers / rounded up to the MMA K step
  int blk_bytes;           // shared-memory bytes of one operand block (multiple of 1024)
  int stages;
  int chunks_h, chunks;    // chunks per image column of NI images / total
  int cblocks, mtiles, groups, chunks_per_group;
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
// kind::f16, bf16 x bf16 -> fp32, both operands MN-major, M = 128, N = 192
__device__ __forceinline__ uint32_t idesc_mn() {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(kN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

__global__ void __launch_bounds__(kThreads, 1)
conv3x3_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                     const __grid_constant__ CUtensorMap tmap_ws, const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = 3 * p.blk_bytes;           // dY block 0, dY block 1, X
  uint8_t* slabs = smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* tmem_full = empty_bar + kMaxStages;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // work item: blockIdx.x = ((grp * cblocks + cb) * 3 + r) * mtiles + mt
  int it = blockIdx.x;
  const int mt = it % p.mtiles; it /= p.mtiles;
  const int r = it % 3; it /= 3;
  const int cb = it % p.cblocks; it /= p.cblocks;
  const int grp = it;
  const int chunk0 = grp * p.chunks_per_group, chunk1 = min(p.chunks, chunk0 + p.chunks_per_group);

  // rows past the TMA box are never written again: zero them once (dY there must be zero, X there must not be NaN)
  for (int s = 0; s < p.stages; ++s)
    for (int b = 0; b < 3; ++b) {
      uint4* tail = reinterpret_cast<uint4*>(smem + (size_t)s * stage_bytes + (size_t)b * p.blk_bytes + (size_t)p.rows * 128);
      const int n16 = (p.blk_bytes - p.rows * 128) / 16;
      for (int i = threadIdx.x; i < n16; i += kThreads) tail[i] = make_uint4(0, 0, 0, 0);
    }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy zeros before the tensor core reads them

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_dy) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_ws) : "memory");
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, 1);
    }
    mbar_init(tmem_full, 1);
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
      uint32_t n = 0;
      for (int ch = chunk0; ch < chunk1; ++ch, ++n) {
        const int s = n % p.stages;
        mbar_wait(empty_bar + s, ((n / p.stages) & 1) ^ 1);
        mbar_expect_tx(full_bar + s, (uint32_t)(3 * p.rows * 128));
        const int img0 = (ch / p.chunks_h) * p.NI, h0 = (ch % p.chunks_h) * p.TH;
        uint8_t* base = smem + (size_t)s * stage_bytes;
        tma_load_4d(base, &tmap_dy, full_bar + s, mt * 128, 0, h0, img0);
        tma_load_4d(base + p.blk_bytes, &tmap_dy, full_bar + s, mt * 128 + 64, 0, h0, img0);
        tma_load_4d(base + 2 * p.blk_bytes, &tmap_x, full_bar + s, cb * 64, -1, h0 + r - 1, img0);
      }
    }
  } else if (warp == 1) {  // ===== MMA issuer =====
    const uint32_t idesc = idesc_mn();
    const int ksteps = p.rows16 / 16;
    uint32_t n = 0;
    for (int ch = chunk0; ch < chunk1; ++ch, ++n) {
      const int s = n % p.stages;
      mbar_wait(full_bar + s, (n / p.stages) & 1);
      tcgen05_fence_after();
      const uint32_t base = smem_u32(smem + (size_t)s * stage_bytes);
      // A: two 64-wide blocks of output channels, blk_bytes apart; B: three 64-wide blocks ONE pixel row (128 B) apart
      const uint64_t a0 = make_sw128_desc(base, (uint32_t)(p.blk_bytes >> 4));
      const uint64_t b0 = make_sw128_desc(base + 2 * p.blk_bytes, 128 >> 4);
      if (elect_one()) {
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint64_t off = (uint64_t)ks * 128;  // 16 pixel rows = 2048 B, in 16-byte units
          if (n == 0 && ks == 0) umma_f16_c<false>(tmem_base, a0, b0, idesc);
          else umma_f16_c<true>(tmem_base, a0 + off, b0 + off, idesc);
        }
        umma_commit(empty_bar + s);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(tmem_full);
    __syncwarp();
  } else if (warp >= 4) {  // ===== epilogue: 192 fp32 columns -> six 32-column slabs -> reduce-add into the workspace =====
    const int wq = warp & 3;
    mbar_wait(tmem_full, 0);
    tcgen05_fence_after();
    uint8_t* my_slabs = slabs + wq * 8192;
    const int row0 = mt * 128 + wq * 32;
#pragma unroll 1
    for (int g = 0; g < 6; ++g) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(g * 32), v);
      if (row0 >= p.K) continue;  // warp-uniform
      uint8_t* slab = my_slabs + (g & 1) * 4096;
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      __syncwarp();
      uint8_t* my_row = slab + lane * 128;
#pragma unroll
      for (int j8 = 0; j8 < 8; ++j8)
        *reinterpret_cast<uint4*>(my_row + ((j8 ^ (lane & 7)) << 4)) = make_uint4(v[4 * j8], v[4 * j8 + 1], v[4 * j8 + 2], v[4 * j8 + 3]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        const int tap = r * 3 + (g >> 1);
        tma_reduce_add_2d(&tmap_ws, slab, tap * p.C + cb * 64 + (g & 1) * 32, row0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
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

static bool plan(Params& p, int n, int h, int wd, int c, int k) {
  p.N = n; p.H = h; p.W = wd; p.C = c; p.K = k;
  p.Wp = wd + 2;
  if (p.Wp > 128 || c % 64 != 0 || k % 8 != 0) return false;
  p.TH = std::min(h, std::max(1, std::min(8, kMaxRows / p.Wp)));
  p.NI = std::max(1, std::min(n, kMaxRows / (p.TH * p.Wp)));
  if (p.TH < h) p.NI = 1;                      // several images per box only when a box holds whole images
  p.NI = std::min(p.NI, 255);
  p.rows = p.NI * p.TH * p.Wp;
  p.rows16 = (p.rows + 15) / 16 * 16;
  p.blk_bytes = ((p.rows16 + kTailRows) * 128 + 1023) / 1024 * 1024;
  const size_t fixed = 1024 + 256;
  p.stages = (int)std::min<size_t>(kMaxStages, (227 * 1024 - fixed) / ((size_t)3 * p.blk_bytes));
  if (p.stages < 2 || (size_t)3 * p.blk_bytes < (size_t)kSlabBytes) return false;
  p.chunks_h = (h + p.TH - 1) / p.TH;
  p.chunks = ((n + p.NI - 1) / p.NI) * p.chunks_h;
  p.cblocks = c / 64;
  p.mtiles = (k + 127) / 128;
  // split the chunks into groups so that the work items fill whole waves of SMs: cost = waves x (chunks per item +
  // about 1.5 chunk-times for the fill, the epilogue and its reduce-add traffic)
  const int base = p.cblocks * 3 * p.mtiles, sms = sm_count();
  double best = 1e30;
  for (int g = 1; g <= std::min(p.chunks, 128); ++g) {
    const int cpg = (p.chunks + g - 1) / g, G = (p.chunks + cpg - 1) / cpg;
    const double cost = (double)((G * base + sms - 1) / sms) * (cpg + 1.5);
    if (cost < best) { best = cost; p.chunks_per_group = cpg; p.groups = G; }
  }
  return true;
}

}  // namespace wgrad3
}  // namespace mfl

extern "C" int mfl_conv3x3_wgrad_supported(int n, int h, int wd, int c, int k) {
  mfl::wgrad3::Params p;
  return n > 0 && h > 0 && wd > 0 && mfl::wgrad3::plan(p, n, h, wd, c, k) ? 1 : 0;
}

// dW[K,3,3,C] (+)= weight gradient of y = conv3x3(x, w) (stride 1, padding 1) from dY[N,H,W,K] and X[N,H,W,C], NHWC
// bf16.  ws: zero-initialised fp32 scratch of K * 9 * C floats, handed back zeroed.
extern "C" int mfl_conv3x3_wgrad_bf16(void* dw, const void* dy, const void* x, float* ws, int n, int h, int wd, int c,
                                      int k, int accumulate, void* stream) {
  using namespace mfl;
  using namespace mfl::wgrad3;
  MFL_ARG(dw && dy && x && ws, "conv3x3 wgrad: null operand");
  Params p;
  MFL_ARG(n > 0 && h > 0 && wd > 0 && plan(p, n, h, wd, c, k),
          "conv3x3 wgrad: unsupported shape n=%d h=%d w=%d c=%d k=%d (need C %% 64 == 0, K %% 8 == 0, W <= 126)", n, h, wd, c, k);
  bind_context();
  EncodeTiledFn fn = encode_fn();
  MFL_ARG(fn != nullptr, "conv3x3 wgrad: cuTensorMapEncodeTiled is not available from this driver");
  CUtensorMap tdy, tx, tws;
  auto act_map = [&](CUtensorMap* m, const void* base, int ch) -> CUresult {
    cuuint64_t dims[4] = {(cuuint64_t)ch, (cuuint64_t)wd, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)ch * 2, (cuuint64_t)wd * ch * 2, (cuuint64_t)h * wd * ch * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wp, (cuuint32_t)p.TH, (cuuint32_t)p.NI};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  CUresult r1 = act_map(&tdy, dy, k), r2 = act_map(&tx, x, c);
  MFL_ARG(r1 == CUDA_SUCCESS && r2 == CUDA_SUCCESS, "conv3x3 wgrad: activation tensor maps failed (%d, %d)", (int)r1, (int)r2);
  {
    const int cols = 9 * c;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)k};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r3 = fn(&tws, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ws, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MFL_ARG(r3 == CUDA_SUCCESS, "conv3x3 wgrad: workspace tensor map failed (%d)", (int)r3);
  }
  const size_t smem = 1024 + (size_t)p.stages * 3 * p.blk_bytes + 256;
  static bool attr_set = false;
  if (!attr_set) {
    MFL_CUDA(cudaFuncSetAttribute(conv3x3_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  const int grid = p.groups * p.cblocks * 3 * p.mtiles;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  conv3x3_wgrad_kernel<<<grid, kThreads, smem, st>>>(tdy, tx, tws, p);
  if (int rc = check(cudaGetLastError(), "conv3x3_wgrad_kernel launch")) return rc;
  return mfl_splitk_finalize(ws, dw, k, 9 * c, 9 * c, accumulate, stream);
}
