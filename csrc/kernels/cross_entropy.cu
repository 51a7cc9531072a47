// K16: fused softmax cross-entropy over a bf16 logits matrix (language-model head): per-row loss AND the gradient
// with respect to the logits in one kernel, no fp32 copy of the logits.
//
// The framework sequence for GPT-2 (config #5) is logits.float() (an 823 MB copy for 4096 x 50304), log-softmax,
// NLL, and in backward the softmax gradient plus a conversion back to bf16: about 2.2 ms of a 23.6 ms step and the
// largest activation of the model.  Here one CTA owns one row: pass 1 is an online (max, sum-exp) reduction over
// 16-byte loads, pass 2 re-reads the row (it is still in L1 / L2: 100 KB) and writes
//   d logits = (softmax - onehot(target)) * scale        (bf16; rows whose target is ignore_index get zeros)
// and the row loss  logsumexp - logit[target].  The backward pass is then just dlogits * upstream gradient.
//
// Reference counterpart: the loss inside model.fit (keras_model_ops.py:156, SparseCategoricalCrossentropy in
// examples/keras/models/*.py).
#include "common.cuh"

namespace mfl {
namespace xent {

constexpr int kThreads = 256;
using V = VecTraits<MFL_BFLOAT16>;

__device__ __forceinline__ void online_merge(float& m, float& s, float m2, float s2) {
  if (m2 == -INFINITY) return;             // the other side saw no element (short rows): exp(-inf - -inf) is NaN
  if (m == -INFINITY) { m = m2; s = s2; return; }
  const float mn = fmaxf(m, m2);
  s = s * __expf(m - mn) + s2 * __expf(m2 - mn);
  m = mn;
}

__global__ void __launch_bounds__(kThreads) lm_xent_kernel(const __nv_bfloat16* __restrict__ logits,
                                                           const long long* __restrict__ targets,
                                                           __nv_bfloat16* __restrict__ dlogits,
                                                           float* __restrict__ row_loss, int vocab, int64_t ld,
                                                           int64_t ldd, float scale, long long ignore_index) {
  __shared__ float red_m[kThreads / 32], red_s[kThreads / 32];
  __shared__ float bc[2];
  const int row = blockIdx.x;
  const __nv_bfloat16* x = logits + (int64_t)row * ld;
  __nv_bfloat16* dx = dlogits + (int64_t)row * ldd;
  const long long tgt = targets[row];
  const int nvec = vocab / 8;
  if (tgt == ignore_index || tgt < 0 || tgt >= vocab) {  // no contribution: zero gradient row, zero loss
    for (int i = threadIdx.x; i < nvec; i += kThreads) reinterpret_cast<uint4*>(dx)[i] = make_uint4(0, 0, 0, 0);
    for (int i = nvec * 8 + threadIdx.x; i < vocab; i += kThreads) dx[i] = __float2bfloat16_rn(0.f);
    if (threadIdx.x == 0) row_loss[row] = 0.f;
    return;
  }
  // ---- pass 1: online max / sum of exponentials -------------------------------------------------------
  float m = -INFINITY, s = 0.f;
  for (int i = threadIdx.x; i < nvec; i += kThreads) {
    float f[8];
    V::unpack(reinterpret_cast<const uint4*>(x)[i], f);
    float vm = f[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) vm = fmaxf(vm, f[k]);
    const float mn = fmaxf(m, vm);
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc += __expf(f[k] - mn);
    s = s * __expf(m - mn) + acc;
    m = mn;
  }
  for (int i = nvec * 8 + threadIdx.x; i < vocab; i += kThreads) online_merge(m, s, __bfloat162float(x[i]), 1.f);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
    online_merge(m, s, m2, s2);
  }
  if ((threadIdx.x & 31) == 0) { red_m[threadIdx.x >> 5] = m; red_s[threadIdx.x >> 5] = s; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float mm = red_m[0], ss = red_s[0];
    for (int w = 1; w < kThreads / 32; ++w) online_merge(mm, ss, red_m[w], red_s[w]);
    const float lse = mm + __logf(ss);
    bc[0] = lse;
    row_loss[row] = lse - __bfloat162float(x[tgt]);
  }
  __syncthreads();
  const float lse = bc[0];
  // ---- pass 2: gradient -------------------------------------------------------------------------------
  const int tvec = (int)(tgt >> 3), tk = (int)(tgt & 7);
  for (int i = threadIdx.x; i < nvec; i += kThreads) {
    float f[8];
    V::unpack(reinterpret_cast<const uint4*>(x)[i], f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = __expf(f[k] - lse) * scale;
    if (i == tvec) f[tk] -= scale;
    reinterpret_cast<uint4*>(dx)[i] = V::pack(f);
  }
  for (int i = nvec * 8 + threadIdx.x; i < vocab; i += kThreads) {
    float g = __expf(__bfloat162float(x[i]) - lse) * scale;
    if (i == tgt) g -= scale;
    dx[i] = __float2bfloat16_rn(g);
  }
}

}  // namespace xent
}  // namespace mfl

using namespace mfl;

// logits [rows, vocab] bf16 (row pitch ld elements), targets [rows] int64, dlogits [rows, vocab] bf16 (pitch ldd),
// row_loss [rows] fp32.  scale multiplies the gradient (1 / number of contributing rows for a mean loss).
extern "C" int mfl_lm_cross_entropy(void* dlogits, int64_t ldd, float* row_loss, const void* logits, int64_t ld,
                                    const long long* targets, int rows, int vocab, float scale, long long ignore_index,
                                    void* stream) {
  MFL_ARG(dlogits && row_loss && logits && targets, "lm_cross_entropy: null tensor");
  MFL_ARG(rows > 0 && vocab > 0, "lm_cross_entropy: empty shape");
  MFL_ARG(ld % 8 == 0 && ldd % 8 == 0 && ((uintptr_t)logits & 15) == 0 && ((uintptr_t)dlogits & 15) == 0,
          "lm_cross_entropy: rows must start on 16-byte boundaries (pitches multiples of 8 elements)");
  xent::lm_xent_kernel<<<rows, xent::kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(logits), targets, static_cast<__nv_bfloat16*>(dlogits), row_loss, vocab, ld, ldd,
      scale, ignore_index);
  return check(cudaGetLastError(), "lm_xent_kernel launch");
}
