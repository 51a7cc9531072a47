// K14 - fused BatchNorm2d (+ residual add) (+ ReLU) for NHWC bf16 activations, forward and backward.
//
// The framework kernels this replaces cost 35 % of a ResNet-18 training step on a B200 (four ATen
// batch-norm kernels per layer plus separate ReLU, ReLU-backward, residual-add, semaphore-fill and
// running-statistics launches; profiles/r1_step_breakdown.md).  The work is purely bandwidth bound,
// so the design is: 16-byte vector accesses with every thread pinned to one group of 8 channels
// (per-channel constants live in registers), deterministic two-level reductions (per-CTA partials in
// a workspace, folded in double precision by a small finalize kernel - no floating-point atomics),
// and everything elementwise that follows the normalisation fused into the same pass:
//
//   forward :  partial sums -> finalize (mean / invstd / running stats / step counter / scale+shift)
//              -> y = relu(x * scale + shift + residual)
//   backward:  g = dy * (y > 0);  partial sums of g and g * xhat -> finalize (dgamma, dbeta accumulated
//              straight into the flat gradient arena) -> dx = gamma * invstd * (g - mean(g) - xhat * mean(g xhat)),
//              d_residual = g
//
// Reference counterpart: none (the reference delegates local training to TensorFlow/Keras,
// keras_model_ops.py:156); this is part of the "local training math" row of SURVEY section 2.4.
#include "common.cuh"

namespace mfl {
namespace bn {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;
constexpr int kMaxPartials = 256;  // finalize folds 32 lanes x 8 partials per channel
using V = VecTraits<MFL_BFLOAT16>;

__device__ __forceinline__ void load8(const float* p, float* f) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// Block-level fold of the per-thread accumulators: red[(ty * 2 + s) * C + c] -> ws[(blk * 2 + s) * C + c].
__device__ __forceinline__ void fold_block(float* red, const float* s0, const float* s1, float* ws, int C,
                                           int tpr, int rpi) {
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    red[(ty * 2 + 0) * C + tx * 8 + k] = s0[k];
    red[(ty * 2 + 1) * C + tx * 8 + k] = s1[k];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < 2 * C; o += kThreads) {
    const int s = o / C, c = o - s * C;
    float acc = 0.f;
    for (int r = 0; r < rpi; ++r) acc += red[(r * 2 + s) * C + c];
    ws[(size_t)blockIdx.x * 2 * C + o] = acc;
  }
}

// ---- forward statistics: per-CTA sum(x), sum(x^2) ----------------------------------------------
__global__ void __launch_bounds__(kThreads) stats_partial_kernel(const __nv_bfloat16* __restrict__ x,
                                                                 float* __restrict__ ws, int M, int C,
                                                                 int rows_per_block) {
  __shared__ float red[kThreads * 16];
  const int tpr = C >> 3, rpi = kThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int row0 = blockIdx.x * rows_per_block;
  const int row1 = min(M, row0 + rows_per_block);
  float s0[8], s1[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s0[k] = s1[k] = 0.f;
  const __nv_bfloat16* base = x + (size_t)tx * 8;
  int r = row0 + ty;
  for (; r + (kUnroll - 1) * rpi < row1; r += kUnroll * rpi) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) v[u] = ld_stream(base + (size_t)(r + u * rpi) * C);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      float f[8];
      V::unpack(v[u], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) { s0[k] += f[k]; s1[k] = fmaf(f[k], f[k], s1[k]); }
    }
  }
  for (; r < row1; r += rpi) {
    float f[8];
    V::unpack(ld_stream(base + (size_t)r * C), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) { s0[k] += f[k]; s1[k] = fmaf(f[k], f[k], s1[k]); }
  }
  fold_block(red, s0, s1, ws, C, tpr, rpi);
}

// ---- backward statistics: per-CTA sum(g), sum(g * xhat) with g = dy * (y > 0) -------------------
template <bool RELU>
__global__ void __launch_bounds__(kThreads) bwd_partial_kernel(const __nv_bfloat16* __restrict__ x,
                                                               const __nv_bfloat16* __restrict__ dy,
                                                               const __nv_bfloat16* __restrict__ y,
                                                               const float* __restrict__ mean,
                                                               const float* __restrict__ invstd,
                                                               float* __restrict__ ws, int M, int C,
                                                               int rows_per_block) {
  __shared__ float red[kThreads * 16];
  const int tpr = C >> 3, rpi = kThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int row0 = blockIdx.x * rows_per_block;
  const int row1 = min(M, row0 + rows_per_block);
  float mu[8], is[8], s0[8], s1[8];
  load8(mean + tx * 8, mu);
  load8(invstd + tx * 8, is);
#pragma unroll
  for (int k = 0; k < 8; ++k) s0[k] = s1[k] = 0.f;
  const size_t col = (size_t)tx * 8;
  int r = row0 + ty;
  auto body = [&](const uint4& vx, const uint4& vdy, const uint4& vy) {
    float fx[8], fg[8], fy[8];
    V::unpack(vx, fx);
    V::unpack(vdy, fg);
    if (RELU) V::unpack(vy, fy);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float g = (!RELU || fy[k] > 0.f) ? fg[k] : 0.f;
      s0[k] += g;
      s1[k] = fmaf(g, (fx[k] - mu[k]) * is[k], s1[k]);
    }
  };
  constexpr int U = 2;  // three streams per row: keep the register footprint moderate
  for (; r + (U - 1) * rpi < row1; r += U * rpi) {
    uint4 vx[U], vd[U], vy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t off = (size_t)(r + u * rpi) * C + col;
      vx[u] = ld_stream(x + off);
      vd[u] = ld_stream(dy + off);
      vy[u] = RELU ? ld_stream(y + off) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) body(vx[u], vd[u], vy[u]);
  }
  for (; r < row1; r += rpi) {
    const size_t off = (size_t)r * C + col;
    body(ld_stream(x + off), ld_stream(dy + off), RELU ? ld_stream(y + off) : make_uint4(0, 0, 0, 0));
  }
  fold_block(red, s0, s1, ws, C, tpr, rpi);
}

// ---- finalize: fold G partials per channel; 32 channels per CTA, 32 partial lanes ----------------
// fp32 pairwise folding (fp64 units are scarce on this part: a double-precision fold of 256 partials cost
// 4 us per launch); only the handful of per-channel operations that follow are done in double.
__device__ __forceinline__ bool fold_partials(const float* ws, int G, int C, double& o0, double& o1) {
  __shared__ float red[2][32][33];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c = blockIdx.x * 32 + tx;
  // all loads first (G <= kMaxPartials = 32 lanes x 8), then the adds: one memory latency instead of G / 32
  float v0[8], v1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = ty + 32 * j;
    const bool ok = c < C && g < G;
    v0[j] = ok ? ws[((size_t)g * 2 + 0) * C + c] : 0.f;
    v1[j] = ok ? ws[((size_t)g * 2 + 1) * C + c] : 0.f;
  }
  red[0][ty][tx] = ((v0[0] + v0[1]) + (v0[2] + v0[3])) + ((v0[4] + v0[5]) + (v0[6] + v0[7]));
  red[1][ty][tx] = ((v1[0] + v1[1]) + (v1[2] + v1[3])) + ((v1[4] + v1[5]) + (v1[6] + v1[7]));
  __syncthreads();
  for (int s = 16; s > 0; s >>= 1) {
    if (ty < s) {
      red[0][ty][tx] += red[0][ty + s][tx];
      red[1][ty][tx] += red[1][ty + s][tx];
    }
    __syncthreads();
  }
  if (ty != 0 || c >= C) return false;
  o0 = (double)red[0][0][tx];
  o1 = (double)red[1][0][tx];
  return true;
}

// coef layout (floats): [0,C) scale, [C,2C) shift
__global__ void __launch_bounds__(1024) finalize_fwd_kernel(const float* __restrict__ ws, int G, int C, int M,
                                                            const float* gamma, const float* beta,
                                                            float* running_mean, float* running_var,
                                                            long long* num_batches_tracked, float* save_mean,
                                                            float* save_invstd, float* coef, float eps,
                                                            float momentum) {
  double s0, s1;
  if (blockIdx.x == 0 && threadIdx.x == 0 && threadIdx.y == 0 && num_batches_tracked) *num_batches_tracked += 1;
  if (!fold_partials(ws, G, C, s0, s1)) return;
  const int c = blockIdx.x * 32 + threadIdx.x;
  const double mean = s0 / M;
  double var = s1 / M - mean * mean;
  if (var < 0.0) var = 0.0;
  const float invstd = (float)(1.0 / sqrt(var + (double)eps));
  const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  save_mean[c] = (float)mean;
  save_invstd[c] = invstd;
  coef[c] = g * invstd;
  coef[C + c] = b - (float)mean * g * invstd;
  if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
  if (running_var) {
    const double unbiased = M > 1 ? var * (double)M / (double)(M - 1) : var;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
  }
}

// inference: scale / shift from the running statistics
__global__ void eval_coef_kernel(int C, const float* gamma, const float* beta, const float* running_mean,
                                 const float* running_var, float* coef, float eps) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = rsqrtf(running_var[c] + eps);
  const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  coef[c] = g * invstd;
  coef[C + c] = b - running_mean[c] * g * invstd;
}

// coef layout: [0,C) a = gamma * invstd, [C,2C) b = mean(g), [2C,3C) cc = mean(g * xhat)
__global__ void __launch_bounds__(1024) finalize_bwd_kernel(const float* __restrict__ ws, int G, int C, int M,
                                                            const float* gamma, const float* save_invstd,
                                                            float* dgamma, float* dbeta, int accumulate,
                                                            float* coef) {
  double s0, s1;
  if (!fold_partials(ws, G, C, s0, s1)) return;
  const int c = blockIdx.x * 32 + threadIdx.x;
  if (dgamma) dgamma[c] = (accumulate ? dgamma[c] : 0.f) + (float)s1;
  if (dbeta) dbeta[c] = (accumulate ? dbeta[c] : 0.f) + (float)s0;
  coef[c] = (gamma ? gamma[c] : 1.f) * save_invstd[c];
  coef[C + c] = (float)(s0 / M);
  coef[2 * C + c] = (float)(s1 / M);
}

// ---- forward apply: y = relu(x * scale + shift + residual) ---------------------------------------
template <bool RELU, bool RES>
__global__ void __launch_bounds__(kThreads) apply_fwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                             const __nv_bfloat16* __restrict__ res,
                                                             __nv_bfloat16* __restrict__ y,
                                                             const float* __restrict__ coef, int M, int C) {
  const int tpr = C >> 3, rpi = kThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  float sc[8], sh[8];
  load8(coef + tx * 8, sc);
  load8(coef + C + tx * 8, sh);
  const size_t col = (size_t)tx * 8;
  const int stride = gridDim.x * rpi;
  auto body = [&](size_t off, const uint4& vx, const uint4& vr) {
    float f[8], fr[8];
    V::unpack(vx, f);
    if (RES) V::unpack(vr, fr);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float t = fmaf(f[k], sc[k], sh[k]);
      if (RES) t += fr[k];
      f[k] = RELU ? fmaxf(t, 0.f) : t;
    }
    *reinterpret_cast<uint4*>(y + off) = V::pack(f);
  };
  int r = blockIdx.x * rpi + ty;
  for (; r + (kUnroll - 1) * stride < M; r += kUnroll * stride) {
    uint4 vx[kUnroll], vr[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t off = (size_t)(r + u * stride) * C + col;
      vx[u] = ld_stream(x + off);
      vr[u] = RES ? ld_stream(res + off) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) body((size_t)(r + u * stride) * C + col, vx[u], vr[u]);
  }
  for (; r < M; r += stride) {
    const size_t off = (size_t)r * C + col;
    body(off, ld_stream(x + off), RES ? ld_stream(res + off) : make_uint4(0, 0, 0, 0));
  }
}

// ---- backward apply: dx = a * (g - b - xhat * cc); d_residual = g --------------------------------
template <bool RELU, bool RES>
__global__ void __launch_bounds__(kThreads) apply_bwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                             const __nv_bfloat16* __restrict__ dy,
                                                             const __nv_bfloat16* __restrict__ y,
                                                             __nv_bfloat16* __restrict__ dx,
                                                             __nv_bfloat16* __restrict__ dres,
                                                             const float* __restrict__ mean,
                                                             const float* __restrict__ invstd,
                                                             const float* __restrict__ coef, int M, int C) {
  const int tpr = C >> 3, rpi = kThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  float mu[8], is[8], a[8], b[8], cc[8];
  load8(mean + tx * 8, mu);
  load8(invstd + tx * 8, is);
  load8(coef + tx * 8, a);
  load8(coef + C + tx * 8, b);
  load8(coef + 2 * C + tx * 8, cc);
  const size_t col = (size_t)tx * 8;
  const int stride = gridDim.x * rpi;
  auto body = [&](size_t off, const uint4& vx, const uint4& vd, const uint4& vy) {
    float fx[8], fg[8], fy[8];
    V::unpack(vx, fx);
    V::unpack(vd, fg);
    if (RELU) V::unpack(vy, fy);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float g = (!RELU || fy[k] > 0.f) ? fg[k] : 0.f;
      fg[k] = g;
      fx[k] = a[k] * (g - b[k] - (fx[k] - mu[k]) * is[k] * cc[k]);
    }
    *reinterpret_cast<uint4*>(dx + off) = V::pack(fx);
    if (RES) *reinterpret_cast<uint4*>(dres + off) = V::pack(fg);
  };
  constexpr int U = 2;
  int r = blockIdx.x * rpi + ty;
  for (; r + (U - 1) * stride < M; r += U * stride) {
    uint4 vx[U], vd[U], vy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t off = (size_t)(r + u * stride) * C + col;
      vx[u] = ld_stream(x + off);
      vd[u] = ld_stream(dy + off);
      vy[u] = RELU ? ld_stream(y + off) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) body((size_t)(r + u * stride) * C + col, vx[u], vd[u], vy[u]);
  }
  for (; r < M; r += stride) {
    const size_t off = (size_t)r * C + col;
    body(off, ld_stream(x + off), ld_stream(dy + off), RELU ? ld_stream(y + off) : make_uint4(0, 0, 0, 0));
  }
}

// ---- host side -----------------------------------------------------------------------------------
struct Plan {
  int tpr, rpi, partial_blocks, rows_per_block, apply_blocks;
};

static bool supported_channels(int c) {  // a row is C/8 threads wide and must divide the 256-thread CTA
  return c >= 8 && c <= 2048 && (c & (c - 1)) == 0;
}

static Plan make_plan(int m, int c) {
  Plan p;
  p.tpr = c / 8;
  p.rpi = kThreads / p.tpr;
  const int iters = (m + p.rpi - 1) / p.rpi;                   // row-iterations in the whole tensor
  p.partial_blocks = max(1, min(min(2 * sm_count(), kMaxPartials), (iters + 7) / 8));
  const int ipb = (iters + p.partial_blocks - 1) / p.partial_blocks;
  p.rows_per_block = ipb * p.rpi;
  p.partial_blocks = (m + p.rows_per_block - 1) / p.rows_per_block;
  p.apply_blocks = max(1, min(8 * sm_count(), (iters + kUnroll - 1) / kUnroll));
  return p;
}

static size_t workspace_floats(int m, int c) {
  const Plan p = make_plan(m, c);
  return (size_t)p.partial_blocks * 2 * c + 3 * (size_t)c;
}

}  // namespace bn
}  // namespace mfl

using namespace mfl;

extern "C" size_t mfl_bn_workspace_bytes(int m, int c) {
  if (m <= 0 || !bn::supported_channels(c)) return 0;
  return bn::workspace_floats(m, c) * sizeof(float);
}

extern "C" int mfl_bn_fwd(void* y, const void* x, const void* residual, const float* gamma, const float* beta,
                          float* running_mean, float* running_var, long long* num_batches_tracked,
                          float* save_mean, float* save_invstd, float* workspace, size_t workspace_bytes, int m,
                          int c, float eps, float momentum, int training, int relu, void* stream) {
  MFL_ARG(y && x && workspace, "mfl_bn_fwd: null tensor");
  MFL_ARG(m > 0 && bn::supported_channels(c), "mfl_bn_fwd: channels must be a power of two in [8, 2048], got %d", c);
  MFL_ARG(workspace_bytes >= mfl_bn_workspace_bytes(m, c), "mfl_bn_fwd: workspace too small");
  MFL_ARG(training ? (save_mean && save_invstd) : (running_mean && running_var), "mfl_bn_fwd: missing statistics");
  auto st = static_cast<cudaStream_t>(stream);
  const bn::Plan p = bn::make_plan(m, c);
  auto* xb = static_cast<const __nv_bfloat16*>(x);
  auto* rb = static_cast<const __nv_bfloat16*>(residual);
  auto* yb = static_cast<__nv_bfloat16*>(y);
  float* coef = workspace + (size_t)p.partial_blocks * 2 * c;
  if (training) {
    bn::stats_partial_kernel<<<p.partial_blocks, bn::kThreads, 0, st>>>(xb, workspace, m, c, p.rows_per_block);
    bn::finalize_fwd_kernel<<<(c + 31) / 32, dim3(32, 32), 0, st>>>(workspace, p.partial_blocks, c, m, gamma, beta,
                                                                    running_mean, running_var, num_batches_tracked,
                                                                    save_mean, save_invstd, coef, eps, momentum);
  } else {
    bn::eval_coef_kernel<<<(c + 127) / 128, 128, 0, st>>>(c, gamma, beta, running_mean, running_var, coef, eps);
  }
  const int g = p.apply_blocks;
  if (relu && rb) bn::apply_fwd_kernel<true, true><<<g, bn::kThreads, 0, st>>>(xb, rb, yb, coef, m, c);
  else if (relu) bn::apply_fwd_kernel<true, false><<<g, bn::kThreads, 0, st>>>(xb, rb, yb, coef, m, c);
  else if (rb) bn::apply_fwd_kernel<false, true><<<g, bn::kThreads, 0, st>>>(xb, rb, yb, coef, m, c);
  else bn::apply_fwd_kernel<false, false><<<g, bn::kThreads, 0, st>>>(xb, rb, yb, coef, m, c);
  MFL_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int mfl_bn_bwd(void* dx, void* dres, float* dgamma, float* dbeta, int accumulate, const void* dy,
                          const void* y, const void* x, const float* gamma, const float* save_mean,
                          const float* save_invstd, float* workspace, size_t workspace_bytes, int m, int c, int relu,
                          void* stream) {
  MFL_ARG(dx && dy && x && save_mean && save_invstd && workspace, "mfl_bn_bwd: null tensor");
  MFL_ARG(!relu || y, "mfl_bn_bwd: the ReLU mask needs the forward output");
  MFL_ARG(m > 0 && bn::supported_channels(c), "mfl_bn_bwd: channels must be a power of two in [8, 2048], got %d", c);
  MFL_ARG(workspace_bytes >= mfl_bn_workspace_bytes(m, c), "mfl_bn_bwd: workspace too small");
  auto st = static_cast<cudaStream_t>(stream);
  const bn::Plan p = bn::make_plan(m, c);
  auto* xb = static_cast<const __nv_bfloat16*>(x);
  auto* db = static_cast<const __nv_bfloat16*>(dy);
  auto* yb = static_cast<const __nv_bfloat16*>(y);
  auto* dxb = static_cast<__nv_bfloat16*>(dx);
  auto* drb = static_cast<__nv_bfloat16*>(dres);
  float* coef = workspace + (size_t)p.partial_blocks * 2 * c;
  if (relu)
    bn::bwd_partial_kernel<true><<<p.partial_blocks, bn::kThreads, 0, st>>>(xb, db, yb, save_mean, save_invstd,
                                                                             workspace, m, c, p.rows_per_block);
  else
    bn::bwd_partial_kernel<false><<<p.partial_blocks, bn::kThreads, 0, st>>>(xb, db, yb, save_mean, save_invstd,
                                                                              workspace, m, c, p.rows_per_block);
  bn::finalize_bwd_kernel<<<(c + 31) / 32, dim3(32, 32), 0, st>>>(workspace, p.partial_blocks, c, m, gamma,
                                                                  save_invstd, dgamma, dbeta, accumulate, coef);
  const int g = p.apply_blocks;
  if (relu && drb)
    bn::apply_bwd_kernel<true, true><<<g, bn::kThreads, 0, st>>>(xb, db, yb, dxb, drb, save_mean, save_invstd, coef, m, c);
  else if (relu)
    bn::apply_bwd_kernel<true, false><<<g, bn::kThreads, 0, st>>>(xb, db, yb, dxb, drb, save_mean, save_invstd, coef, m, c);
  else if (drb)
    bn::apply_bwd_kernel<false, true><<<g, bn::kThreads, 0, st>>>(xb, db, yb, dxb, drb, save_mean, save_invstd, coef, m, c);
  else
    bn::apply_bwd_kernel<false, false><<<g, bn::kThreads, 0, st>>>(xb, db, yb, dxb, drb, save_mean, save_invstd, coef, m, c);
  MFL_CUDA(cudaGetLastError());
  return 0;
}
