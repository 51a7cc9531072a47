// K11s: the network stem - a 3x3 / stride-1 / padding-1 convolution over an image with 1..4 channels (3 -> 64 for the
// CIFAR ResNet-18), forward and weight gradient, NHWC bf16.
//
// With C = 3 the implicit-GEMM reduction is only 27 long: there is nothing for a 128 x N x 16 tensor-core tile to
// chew on (K11 needs C % 64 == 0), the layer is 0.5 % of the network's FLOPs and it is bound by writing (forward) or
// reading (weight gradient) the 64-channel activation.  So this is a plain FMA kernel shaped for those streams:
//   forward : one output pixel per thread, its 27 inputs in registers, the 64 x 27 filters in shared memory (read as
//             broadcast float4), 128-byte rows of bf16 outputs written 16 bytes at a time;
//   wgrad   : thread (pixel group, k) keeps the 27 partial sums of output channel k in registers while the block
//             streams its pixels (dY rows read coalesced, the 27-value input patches staged in shared memory);
//             blocks combine with fp32 atomics into a 64 x 27 scratch that a last step folds into the bf16 gradient.
// There is no data gradient: the stem's input is the image.
// Reference counterpart: the first framework convolution under model.fit (keras_model_ops.py:156).
#include <algorithm>

#include "common.cuh"

namespace mfl {
namespace stem {

constexpr int kMaxK = 128;    // output channels
constexpr int kTaps = 36;     // 9 taps x up to 4 channels

template <int C>
__global__ void __launch_bounds__(256) stem_fprop_kernel(const __nv_bfloat16* __restrict__ x,
                                                         const __nv_bfloat16* __restrict__ w,
                                                         __nv_bfloat16* __restrict__ y, int N, int H, int W, int K) {
  __shared__ __align__(16) float sw[kMaxK * kTaps];  // [k][tap*C + c], rows padded to kTaps
  const int taps = 9 * C;
  for (int i = threadIdx.x; i < K * kTaps; i += blockDim.x) {
    const int k = i / kTaps, j = i - k * kTaps;
    sw[i] = j < taps ? __bfloat162float(w[k * taps + j]) : 0.f;
  }
  __syncthreads();
  const long long P = (long long)N * H * W;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (long long)gridDim.x * blockDim.x) {
    const int wx = (int)(p % W), hy = (int)((p / W) % H);
    const long long img = p / ((long long)W * H);
    float in[kTaps];
#pragma unroll
    for (int j = 0; j < kTaps; ++j) in[j] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int yy = hy + r - 1;
      if (yy < 0 || yy >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int xx = wx + s - 1;
        if (xx < 0 || xx >= W) continue;
        const __nv_bfloat16* px = x + ((img * H + yy) * W + xx) * C;
#pragma unroll
        for (int c = 0; c < C; ++c) in[(r * 3 + s) * C + c] = __bfloat162float(px[c]);
      }
    }
    __nv_bfloat16* out = y + p * K;
    for (int k0 = 0; k0 < K; k0 += 8) {
      float acc[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4* wr = reinterpret_cast<const float4*>(sw + (k0 + q) * kTaps);
        float a = 0.f;
#pragma unroll
        for (int j4 = 0; j4 < kTaps / 4; ++j4) {
          const float4 wv = wr[j4];
          a = fmaf(in[4 * j4], wv.x, a); a = fmaf(in[4 * j4 + 1], wv.y, a);
          a = fmaf(in[4 * j4 + 2], wv.z, a); a = fmaf(in[4 * j4 + 3], wv.w, a);
        }
        acc[q] = a;
      }
      *reinterpret_cast<uint4*>(out + k0) = VecTraits<MFL_BFLOAT16>::pack(acc);
    }
  }
}

constexpr int kPixTile = 64;   // pixels staged per step
constexpr int kGroups = 4;     // pixel groups per block (blockDim = kGroups * K)

template <int C>
__global__ void __launch_bounds__(512) stem_wgrad_kernel(const __nv_bfloat16* __restrict__ x,
                                                         const __nv_bfloat16* __restrict__ dy, float* __restrict__ ws,
                                                         int N, int H, int W, int K, int pix_per_block) {
  __shared__ __align__(16) float patch[kPixTile * kTaps];
  const int taps = 9 * C;
  const int k = threadIdx.x % K, grp = threadIdx.x / K;
  const long long P = (long long)N * H * W;
  const long long p_begin = (long long)blockIdx.x * pix_per_block;
  const long long p_end = min(P, p_begin + pix_per_block);
  float acc[kTaps];
#pragma unroll
  for (int j = 0; j < kTaps; ++j) acc[j] = 0.f;
  for (long long p0 = p_begin; p0 < p_end; p0 += kPixTile) {
    __syncthreads();
    for (int i = threadIdx.x; i < kPixTile * kTaps; i += blockDim.x) {
      const int q = i / kTaps, j = i - q * kTaps;
      const long long p = p0 + q;
      float v = 0.f;
      if (p < p_end && j < taps) {
        const int tap = j / C, c = j - tap * C, r = tap / 3, s = tap - 3 * r;
        const int wx = (int)(p % W) + s - 1, hy = (int)((p / W) % H) + r - 1;
        const long long img = p / ((long long)W * H);
        if (wx >= 0 && wx < W && hy >= 0 && hy < H) v = __bfloat162float(x[((img * H + hy) * W + wx) * C + c]);
      }
      patch[i] = v;
    }
    __syncthreads();
    for (int q = grp; q < kPixTile; q += kGroups) {
      const long long p = p0 + q;
      if (p >= p_end) break;
      const float g = __bfloat162float(dy[p * K + k]);
      const float4* pr = reinterpret_cast<const float4*>(patch + q * kTaps);
#pragma unroll
      for (int j4 = 0; j4 < kTaps / 4; ++j4) {
        const float4 v = pr[j4];
        acc[4 * j4] = fmaf(g, v.x, acc[4 * j4]); acc[4 * j4 + 1] = fmaf(g, v.y, acc[4 * j4 + 1]);
        acc[4 * j4 + 2] = fmaf(g, v.z, acc[4 * j4 + 2]); acc[4 * j4 + 3] = fmaf(g, v.w, acc[4 * j4 + 3]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kTaps; ++j)
    if (j < taps) atomicAdd(ws + k * taps + j, acc[j]);
}

__global__ void stem_wgrad_finalize_kernel(float* __restrict__ ws, __nv_bfloat16* __restrict__ dw, int n, int accumulate) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float v = ws[i];
    if (accumulate) v += __bfloat162float(dw[i]);
    dw[i] = __float2bfloat16_rn(v);
    ws[i] = 0.f;
  }
}

// ---- the stem as a tensor-core problem --------------------------------------------------------------------------
// patches[p][(r*3+s)*C + c] = x[pixel p shifted by tap (r, s)][c], zero outside the image, rows padded to 32 bf16
// (64 bytes: the shortest row a TMA tensor map accepts cleanly).  With this 8 MB matrix the stem's forward is a plain
// K10 GEMM  Y[P, K] = patches[P, 32] . W32[K, 32]^T  and its weight gradient a split-k GEMM  dW32[K, 32] = dY^T . patches
// (both operands read MN-major), instead of an FMA loop per pixel: 16 us + 17 us against 43 + 62 us for the library's
// direct kernels on the CIFAR stem (bench/stem_bench.py).  One thread per pixel: nine 2C-byte reads, one 64-byte row out.
constexpr int kPatchCols = 32;

template <int C>
__global__ void __launch_bounds__(256) stem_im2col_kernel(const __nv_bfloat16* __restrict__ x,
                                                          __nv_bfloat16* __restrict__ patches, int N, int H, int W) {
  static_assert(9 * C <= kPatchCols, "patch row");
  const long long P = (long long)N * H * W;
  const __nv_bfloat16 zero = __float2bfloat16_rn(0.f);
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (long long)gridDim.x * blockDim.x) {
    const int wx = (int)(p % W), hy = (int)((p / W) % H);
    const long long img = p / ((long long)W * H);
    __align__(16) __nv_bfloat16 row[kPatchCols];
#pragma unroll
    for (int j = 0; j < kPatchCols; ++j) row[j] = zero;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int yy = hy + r - 1;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int xx = wx + s - 1;
        if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
          const __nv_bfloat16* px = x + ((img * H + yy) * W + xx) * C;
#pragma unroll
          for (int c = 0; c < C; ++c) row[(r * 3 + s) * C + c] = px[c];
        }
      }
    }
    uint4* out = reinterpret_cast<uint4*>(patches + p * kPatchCols);
    const uint4* src = reinterpret_cast<const uint4*>(row);
#pragma unroll
    for (int q = 0; q < kPatchCols * 2 / 16; ++q) out[q] = src[q];
  }
}

}  // namespace stem
}  // namespace mfl

using namespace mfl;

// patches[N*H*W, 32] (bf16, rows of 64 bytes) from x[N,H,W,C], C <= 3: column (r*3+s)*C + c, zero padded.
extern "C" int mfl_conv_stem_im2col_bf16(void* patches, const void* x, int n, int h, int wd, int c, void* stream) {
  MFL_ARG(patches && x && n > 0 && h > 0 && wd > 0, "conv stem im2col: null operand or empty shape");
  MFL_ARG(c >= 1 && c <= 3, "conv stem im2col: 1 <= C <= 3 (9 C values fit a 32-wide patch row), got %d", c);
  const long long P = (long long)n * h * wd;
  const int blocks = (int)std::min<long long>((P + 255) / 256, (long long)sm_count() * 8);
  auto xs = static_cast<const __nv_bfloat16*>(x);
  auto ps = static_cast<__nv_bfloat16*>(patches);
  cudaStream_t st = (cudaStream_t)stream;
  switch (c) {
    case 1: stem::stem_im2col_kernel<1><<<blocks, 256, 0, st>>>(xs, ps, n, h, wd); break;
    case 2: stem::stem_im2col_kernel<2><<<blocks, 256, 0, st>>>(xs, ps, n, h, wd); break;
    default: stem::stem_im2col_kernel<3><<<blocks, 256, 0, st>>>(xs, ps, n, h, wd); break;
  }
  return check(cudaGetLastError(), "stem_im2col_kernel launch");
}

extern "C" int mfl_conv_stem_supported(int c, int k) { return c >= 1 && c <= 4 && k % 8 == 0 && k <= stem::kMaxK && 512 % k == 0; }

// y[N,H,W,K] = conv3x3(x[N,H,W,C], w[K,3,3,C]), stride 1, zero padding 1, C <= 4, NHWC bf16.
extern "C" int mfl_conv_stem_fprop_bf16(void* y, const void* x, const void* w, int n, int h, int wd, int c, int k,
                                        void* stream) {
  MFL_ARG(y && x && w && n > 0 && h > 0 && wd > 0, "conv stem: null operand or empty shape");
  MFL_ARG(mfl_conv_stem_supported(c, k), "conv stem: needs 1 <= C <= 4 and K in {8,16,32,64,128} (got C=%d K=%d)", c, k);
  const long long P = (long long)n * h * wd;
  const int blocks = (int)std::min<long long>((P + 255) / 256, (long long)sm_count() * 8);
  auto xs = static_cast<const __nv_bfloat16*>(x), wsrc = static_cast<const __nv_bfloat16*>(w);
  auto ys = static_cast<__nv_bfloat16*>(y);
  cudaStream_t st = (cudaStream_t)stream;
  switch (c) {
    case 1: stem::stem_fprop_kernel<1><<<blocks, 256, 0, st>>>(xs, wsrc, ys, n, h, wd, k); break;
    case 2: stem::stem_fprop_kernel<2><<<blocks, 256, 0, st>>>(xs, wsrc, ys, n, h, wd, k); break;
    case 3: stem::stem_fprop_kernel<3><<<blocks, 256, 0, st>>>(xs, wsrc, ys, n, h, wd, k); break;
    default: stem::stem_fprop_kernel<4><<<blocks, 256, 0, st>>>(xs, wsrc, ys, n, h, wd, k); break;
  }
  return check(cudaGetLastError(), "stem_fprop_kernel launch");
}

// dW[K,3,3,C] (+)= the weight gradient of the stem; ws: zero-initialised fp32 scratch of K * 9 * C floats, handed
// back zeroed.
extern "C" int mfl_conv_stem_wgrad_bf16(void* dw, const void* dy, const void* x, float* ws, int n, int h, int wd, int c,
                                        int k, int accumulate, void* stream) {
  MFL_ARG(dw && dy && x && ws && n > 0 && h > 0 && wd > 0, "conv stem wgrad: null operand or empty shape");
  MFL_ARG(mfl_conv_stem_supported(c, k), "conv stem wgrad: needs 1 <= C <= 4 and K in {8,...,128} (got C=%d K=%d)", c, k);
  const long long P = (long long)n * h * wd;
  int blocks = (int)std::min<long long>((P + 511) / 512, (long long)sm_count() * 2);
  const int ppb = (int)((P + blocks - 1) / blocks);
  blocks = (int)((P + ppb - 1) / ppb);
  cudaStream_t st = (cudaStream_t)stream;
  auto xs = static_cast<const __nv_bfloat16*>(x), gs = static_cast<const __nv_bfloat16*>(dy);
  const int threads = stem::kGroups * k;
  switch (c) {
    case 1: stem::stem_wgrad_kernel<1><<<blocks, threads, 0, st>>>(xs, gs, ws, n, h, wd, k, ppb); break;
    case 2: stem::stem_wgrad_kernel<2><<<blocks, threads, 0, st>>>(xs, gs, ws, n, h, wd, k, ppb); break;
    case 3: stem::stem_wgrad_kernel<3><<<blocks, threads, 0, st>>>(xs, gs, ws, n, h, wd, k, ppb); break;
    default: stem::stem_wgrad_kernel<4><<<blocks, threads, 0, st>>>(xs, gs, ws, n, h, wd, k, ppb); break;
  }
  if (int rc = check(cudaGetLastError(), "stem_wgrad_kernel launch")) return rc;
  const int nw = k * 9 * c;
  stem::stem_wgrad_finalize_kernel<<<(nw + 255) / 256, 256, 0, st>>>(ws, static_cast<__nv_bfloat16*>(dw), nw, accumulate);
  return check(cudaGetLastError(), "stem_wgrad_finalize_kernel launch");
}
