// K14 - fused BatchNorm2d (+ residual add) (+ ReLU) for NHWC bf16 activations, forward and backward.
//
// The framework kernels this replaces cost 35 % of a ResNet-18 training step on a B200 (four ATen
// batch-norm kernels per layer plus separate ReLU, ReLU-backward, residual-add, semaphore-fill and
// running-statistics launches; profiles/README.md "training step").  The work is purely bandwidth
// bound, so the design is: 16-byte vector accesses with every thread pinned to one group of 8
// channels (per-channel constants live in registers), every row loop fully software-pipelined
// (kUnroll independent 16-byte loads in flight per thread, out-of-range rows predicated off rather
// than peeled into a serial tail), deterministic two-level reductions (per-CTA partials in a
// workspace, folded in a fixed order - no floating-point atomics), and everything elementwise that
// follows the normalisation fused into the same pass:
//
//   forward :  partial sums -> finalize (mean / invstd / running stats / step counter / scale+shift)
//              -> y = relu(x * scale + shift + residual)
//   backward:  g = dy * (y > 0);  partial sums of g and g * xhat -> finalize (dgamma, dbeta accumulated
//              straight into the flat gradient arena) -> dx = gamma * invstd * (g - mean(g) - xhat * mean(g xhat)),
//              d_residual = g
//
// Two back-ends share the row loops: three launches per direction, or ONE cooperative launch whose
// phases are separated by grid barriers (a dependent launch costs ~3.8 us here - more than a whole
// phase of the small layers).
//
// Reference counterpart: none (the reference delegates local training to TensorFlow/Keras,
// keras_model_ops.py:156); this is part of the "local training math" row of SURVEY section 2.4.
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"

namespace mfl {
namespace bn {

constexpr int kThreads = 256;        // three-launch kernels
constexpr int kCoopThreads = 512;    // cooperative kernels (2 CTAs per SM)
constexpr int kUnroll = 4;           // independent row loads in flight per thread (one input stream)
constexpr int kUnrollBwd = 2;        // three input streams per row in backward
constexpr int kMaxPartials = 256;    // finalize_*_kernel folds 32 lanes x 8 partials per channel
using V = VecTraits<MFL_BFLOAT16>;

__device__ __forceinline__ void load8(const float* p, float* f) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void load8_cg(const float* p, float* f) {  // written earlier in this launch: bypass L1
  const float4 a = __ldcg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldcg(reinterpret_cast<const float4*>(p + 4));
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
// Programmatic dependent launch: the three kernels of one direction are launched back to back with the
// programmatic-stream-serialization attribute; a dependent grid may start (and sit in pdl_wait) while its producer
// is still running, which hides most of the ~3.8 us dependent-launch latency.  pdl_wait() returns once every
// prerequisite grid has completed and its writes are visible, so the data flow is unchanged.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ uint4 ld_row(const __nv_bfloat16* p, bool ok) {
  return ok ? ld_stream(p) : make_uint4(0, 0, 0, 0);
}

// ---- row loops (shared by both back-ends): rows r_begin, r_begin + step, ... < r_end -------------------
// A zero row contributes nothing to any of the sums, so predicated-off loads need no further masking.
__device__ __forceinline__ void rows_stats(const __nv_bfloat16* __restrict__ x, int C, size_t col, int r_begin,
                                           int r_end, int step, float* s0, float* s1) {
  for (int r = r_begin; r < r_end; r += kUnroll * step) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) v[u] = ld_row(x + (size_t)(r + u * step) * C + col, r + u * step < r_end);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      float f[8];
      V::unpack(v[u], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) { s0[k] += f[k]; s1[k] = fmaf(f[k], f[k], s1[k]); }
    }
  }
}

template <bool RELU>
__device__ __forceinline__ void rows_bwd_stats(const __nv_bfloat16* __restrict__ x,
                                               const __nv_bfloat16* __restrict__ dy,
                                               const __nv_bfloat16* __restrict__ y, int C, size_t col, int r_begin,
                                               int r_end, int step, const float* mu, const float* is, float* s0,
                                               float* s1) {
  constexpr int U = kUnrollBwd;
  for (int r = r_begin; r < r_end; r += U * step) {
    uint4 vx[U], vd[U], vy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t off = (size_t)(r + u * step) * C + col;
      const bool ok = r + u * step < r_end;
      vx[u] = ld_row(x + off, ok);
      vd[u] = ld_row(dy + off, ok);
      vy[u] = RELU ? ld_row(y + off, ok) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float fx[8], fg[8], fy[8];
      V::unpack(vx[u], fx);
      V::unpack(vd[u], fg);
      if (RELU) V::unpack(vy[u], fy);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float g = (!RELU || fy[k] > 0.f) ? fg[k] : 0.f;
        s0[k] += g;
        s1[k] = fmaf(g, (fx[k] - mu[k]) * is[k], s1[k]);
      }
    }
  }
}

template <bool RELU, bool RES>
__device__ __forceinline__ void rows_apply_fwd(const __nv_bfloat16* __restrict__ x,
                                               const __nv_bfloat16* __restrict__ res, __nv_bfloat16* __restrict__ y,
                                               int C, size_t col, int r_begin, int r_end, int step, const float* sc,
                                               const float* sh) {
  for (int r = r_begin; r < r_end; r += kUnroll * step) {
    uint4 vx[kUnroll], vr[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t off = (size_t)(r + u * step) * C + col;
      const bool ok = r + u * step < r_end;
      vx[u] = ld_row(x + off, ok);
      vr[u] = RES ? ld_row(res + off, ok) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (r + u * step < r_end) {
        float f[8], fr[8];
        V::unpack(vx[u], f);
        if (RES) V::unpack(vr[u], fr);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          float t = fmaf(f[k], sc[k], sh[k]);
          if (RES) t += fr[k];
          f[k] = RELU ? fmaxf(t, 0.f) : t;
        }
        *reinterpret_cast<uint4*>(y + (size_t)(r + u * step) * C + col) = V::pack(f);
      }
    }
  }
}

template <bool RELU, bool RES>
__device__ __forceinline__ void rows_apply_bwd(const __nv_bfloat16* __restrict__ x,
                                               const __nv_bfloat16* __restrict__ dy,
                                               const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ dx,
                                               __nv_bfloat16* __restrict__ dres, int C, size_t col, int r_begin,
                                               int r_end, int step, const float* mu, const float* is, const float* a,
                                               const float* b, const float* cc) {
  constexpr int U = kUnrollBwd;
  for (int r = r_begin; r < r_end; r += U * step) {
    uint4 vx[U], vd[U], vy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t off = (size_t)(r + u * step) * C + col;
      const bool ok = r + u * step < r_end;
      vx[u] = ld_row(x + off, ok);
      vd[u] = ld_row(dy + off, ok);
      vy[u] = RELU ? ld_row(y + off, ok) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (r + u * step < r_end) {
        float fx[8], fg[8], fy[8];
        V::unpack(vx[u], fx);
        V::unpack(vd[u], fg);
        if (RELU) V::unpack(vy[u], fy);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float g = (!RELU || fy[k] > 0.f) ? fg[k] : 0.f;
          fg[k] = g;
          fx[k] = a[k] * (g - b[k] - (fx[k] - mu[k]) * is[k] * cc[k]);
        }
        const size_t off = (size_t)(r + u * step) * C + col;
        *reinterpret_cast<uint4*>(dx + off) = V::pack(fx);
        if (RES) *reinterpret_cast<uint4*>(dres + off) = V::pack(fg);
      }
    }
  }
}

// Block-level fold of the per-thread accumulators through shared memory (NT * 16 floats):
// red[(ty * 2 + s) * C + c] -> ws_row[s * C + c].
template <int NT>
__device__ __forceinline__ void fold_block(float* red, const float* s0, const float* s1, float* ws_row, int C, int tpr,
                                           int rpi) {
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    red[(ty * 2 + 0) * C + tx * 8 + k] = s0[k];
    red[(ty * 2 + 1) * C + tx * 8 + k] = s1[k];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < 2 * C; o += NT) {
    const int s = o / C, c = o - s * C;
    float acc = 0.f;
    for (int r = 0; r < rpi; ++r) acc += red[(r * 2 + s) * C + c];
    ws_row[o] = acc;
  }
}

// Per-channel epilogues of the two reductions.  coef layouts (floats):
//   forward : [0,C) scale = gamma * invstd, [C,2C) shift = beta - mean * scale
//   backward: [0,C) a = gamma * invstd, [C,2C) b = mean(g), [2C,3C) cc = mean(g * xhat)
struct FwdFinal {
  const float *gamma, *beta;
  float *running_mean, *running_var;
  long long* num_batches_tracked;
  float *save_mean, *save_invstd, *coef;
  float eps, momentum;
};
struct BwdFinal {
  const float *gamma, *save_invstd;
  float *dgamma, *dbeta;
  int accumulate;
  float* coef;
};

__device__ __forceinline__ void finalize_fwd_channel(int c, float s0, float s1, int M, int C, const FwdFinal& f) {
  // E[x^2] - mean^2 cancels in fp32 when |mean| >> std: that one subtraction is done in double, the rest in fp32
  const double mean_d = (double)s0 / M;
  const double var_d = (double)s1 / M - mean_d * mean_d;
  const float mean = (float)mean_d;
  const float var = var_d > 0.0 ? (float)var_d : 0.f;
  const float invstd = rsqrtf(var + f.eps);
  const float g = f.gamma ? f.gamma[c] : 1.f, b = f.beta ? f.beta[c] : 0.f;
  f.save_mean[c] = mean;
  f.save_invstd[c] = invstd;
  f.coef[c] = g * invstd;
  f.coef[C + c] = b - mean * g * invstd;
  if (f.running_mean) f.running_mean[c] = (1.f - f.momentum) * f.running_mean[c] + f.momentum * mean;
  if (f.running_var) {
    const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
    f.running_var[c] = (1.f - f.momentum) * f.running_var[c] + f.momentum * unbiased;
  }
}

__device__ __forceinline__ void finalize_bwd_channel(int c, float s0, float s1, int M, int C, const BwdFinal& f) {
  if (f.dgamma) f.dgamma[c] = (f.accumulate ? f.dgamma[c] : 0.f) + s1;
  if (f.dbeta) f.dbeta[c] = (f.accumulate ? f.dbeta[c] : 0.f) + s0;
  f.coef[c] = (f.gamma ? f.gamma[c] : 1.f) * f.save_invstd[c];
  f.coef[C + c] = s0 / (float)M;
  f.coef[2 * C + c] = s1 / (float)M;
}

// ==== back-end A: three launches per direction ==========================================================
__global__ void __launch_bounds__(kThreads) stats_partial_kernel(const __nv_bfloat16* __restrict__ x,
                                                                 float* __restrict__ ws, int M, int C,
                                                                 int rows_per_block) {
  __shared__ float red[kThreads * 16];
  pdl_launch_dependents();
  const int tpr = C >> 3, rpi = kThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int row0 = blockIdx.x * rows_per_block;
  float s0[8], s1[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s0[k] = s1[k] = 0.f;
  rows_stats(x, C, (size_t)tx * 8, row0 + ty, min(M, row0 + rows_per_block), rpi, s0, s1);
  fold_block<kThreads>(red, s0, s1, ws + (size_t)blockIdx.x * 2 * C, C, tpr, rpi);
}

template <bool RELU>
__global__ void __launch_bounds__(kThreads) bwd_partial_kernel(const __nv_bfloat16* __restrict__ x,
                                                               const __nv_bfloat16* __restrict__ dy,
                                                               const __nv_bfloat16* __restrict__ y,
                                                               const float* __restrict__ mean,
                                                               const float* __restrict__ invstd,
                                                               float* __restrict__ ws, int M, int C,
                                                               int rows_per_block) {
  __shared__ float red[kThreads * 16];
  pdl_launch_dependents();
  const int tpr = C >> 3, rpi = kThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int row0 = blockIdx.x * rows_per_block;
  float mu[8], is[8], s0[8], s1[8];
  load8(mean + tx * 8, mu);
  load8(invstd + tx * 8, is);
#pragma unroll
  for (int k = 0; k < 8; ++k) s0[k] = s1[k] = 0.f;
  rows_bwd_stats<RELU>(x, dy, y, C, (size_t)tx * 8, row0 + ty, min(M, row0 + rows_per_block), rpi, mu, is, s0, s1);
  fold_block<kThreads>(red, s0, s1, ws + (size_t)blockIdx.x * 2 * C, C, tpr, rpi);
}

// finalize: fold G <= kMaxPartials partial rows per channel; 32 channels per CTA, 32 partial lanes.
// fp32 pairwise folding with all loads issued before the first add (one memory latency, not G / 32).
__device__ __forceinline__ bool fold_partials(const float* ws, int G, int C, float& o0, float& o1) {
  __shared__ float red[2][32][33];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c = blockIdx.x * 32 + tx;
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
  o0 = red[0][0][tx];
  o1 = red[1][0][tx];
  return true;
}

__global__ void __launch_bounds__(1024) finalize_fwd_kernel(const float* __restrict__ ws, int G, int C, int M,
                                                            FwdFinal fin) {
  float s0, s1;
  pdl_wait();
  pdl_launch_dependents();
  if (blockIdx.x == 0 && threadIdx.x == 0 && threadIdx.y == 0 && fin.num_batches_tracked)
    *fin.num_batches_tracked += 1;
  if (!fold_partials(ws, G, C, s0, s1)) return;
  finalize_fwd_channel(blockIdx.x * 32 + threadIdx.x, s0, s1, M, C, fin);
}

__global__ void __launch_bounds__(1024) finalize_bwd_kernel(const float* __restrict__ ws, int G, int C, int M,
                                                            BwdFinal fin) {
  float s0, s1;
  pdl_wait();
  pdl_launch_dependents();
  if (!fold_partials(ws, G, C, s0, s1)) return;
  finalize_bwd_channel(blockIdx.x * 32 + threadIdx.x, s0, s1, M, C, fin);
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

template <bool RELU, bool RES>
__global__ void __launch_bounds__(kThreads) apply_fwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                             const __nv_bfloat16* __restrict__ res,
                                                             __nv_bfloat16* __restrict__ y,
                                                             const float* __restrict__ coef, int M, int C) {
  const int tpr = C >> 3, rpi = kThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  float sc[8], sh[8];
  pdl_wait();
  load8(coef + tx * 8, sc);
  load8(coef + C + tx * 8, sh);
  rows_apply_fwd<RELU, RES>(x, res, y, C, (size_t)tx * 8, blockIdx.x * rpi + ty, M, gridDim.x * rpi, sc, sh);
}

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
  pdl_wait();
  load8(coef + tx * 8, a);
  load8(coef + C + tx * 8, b);
  load8(coef + 2 * C + tx * 8, cc);
  rows_apply_bwd<RELU, RES>(x, dy, y, dx, dres, C, (size_t)tx * 8, blockIdx.x * rpi + ty, M, gridDim.x * rpi, mu, is, a,
                            b, cc);
}

// ==== back-end B: one cooperative launch per direction ================================================
//   phase 1  per-CTA partial sums over a contiguous chunk of rows              -> ws[G][2C]
//   barrier
//   phase 2  CTAs fold 8 channels each (8 columns x 64 partial lanes)          -> coef, saved stats, running stats
//   barrier
//   phase 3  every CTA re-reads its own chunk (still in L2) and applies
//
// Sense-reversing grid barrier on two persistent words, 128 B apart so that the pollers of the generation word do
// not queue on the L2 sector the arrival atomics serialise on: bar[0] = count, bar[32] = generation.  Needs all
// CTAs co-resident (cooperative launch).  Release on arrival / acquire on the poll at GPU scope instead of full
// fences: a __threadfence() is a MEMBAR.SC + L1 invalidate here, and three of them per barrier cost more than the
// phases they separate.
// Arrival is two-level (groups of 16 CTAs, each with its own counter line, then one top counter): same-address L2
// atomics serialise at ~27 cycles per warp-level op, so 296 flat arrivals cost ~4 us, 16 + 19 cost ~0.5 us.
// Word layout (uint32): [0] top count, [32] generation, [64 + 32 k] count of group k (k < 32).
constexpr int kBarGenOffset = 32, kBarGroupBase = 64, kBarGroupShift = 4;
constexpr int kBarWords = kBarGroupBase + 32 * 32;
__device__ __forceinline__ unsigned int bar_arrive(unsigned int* p) {
  unsigned int prev;
  asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(prev) : "l"(p) : "memory");
  return prev;
}
__device__ __forceinline__ void bar_reset(unsigned int* p) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(0u) : "memory");
}
__device__ __forceinline__ void grid_barrier(unsigned int* bar) {
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* gen_p = bar + kBarGenOffset;
    unsigned int gen, cur;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(gen) : "l"(gen_p) : "memory");
    const unsigned int group = blockIdx.x >> kBarGroupShift;
    const unsigned int ngroups = (gridDim.x + (1u << kBarGroupShift) - 1) >> kBarGroupShift;
    const unsigned int gsize = min(1u << kBarGroupShift, gridDim.x - (group << kBarGroupShift));
    unsigned int* gcount = bar + kBarGroupBase + 32 * group;
    bool release = false;
    if (bar_arrive(gcount) == gsize - 1) {
      bar_reset(gcount);
      if (bar_arrive(bar) == ngroups - 1) {
        bar_reset(bar);
        release = true;
      }
    }
    if (release) {
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(gen_p), "r"(gen + 1u) : "memory");
    } else {
      do {
        __nanosleep(20);
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(cur) : "l"(gen_p) : "memory");
      } while (cur == gen);
    }
  }
  __syncthreads();
}

// Phase 2: a CTA folds channels [8cb, 8cb+8): thread = (column j in 0..7, lane l in 0..63).
// Returns true for the 8 threads (lane 0) that own a channel, with its two totals.
__device__ __forceinline__ bool coop_fold_channels(float* red, const float* ws, int C, int cb, float& o0, float& o1) {
  const int G = gridDim.x;
  const int j = threadIdx.x & 7, l = threadIdx.x >> 3;
  const int c = cb * 8 + j;
  float a0 = 0.f, a1 = 0.f;
#pragma unroll 5
  for (int g = l; g < G; g += 64) {
    a0 += __ldcg(ws + (size_t)g * 2 * C + c);
    a1 += __ldcg(ws + (size_t)g * 2 * C + C + c);
  }
  red[(l * 2 + 0) * 8 + j] = a0;
  red[(l * 2 + 1) * 8 + j] = a1;
  __syncthreads();
  for (int s = 32; s > 0; s >>= 1) {
    if (l < s) {
      red[(l * 2 + 0) * 8 + j] += red[((l + s) * 2 + 0) * 8 + j];
      red[(l * 2 + 1) * 8 + j] += red[((l + s) * 2 + 1) * 8 + j];
    }
    __syncthreads();
  }
  o0 = red[j];
  o1 = red[8 + j];
  return l == 0;
}

template <bool RELU, bool RES>
__global__ void __launch_bounds__(kCoopThreads, 2) coop_fwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                                   const __nv_bfloat16* __restrict__ res,
                                                                   __nv_bfloat16* __restrict__ y, float* ws, int M, int C,
                                                                   int rows_per_block, unsigned int* bar, FwdFinal fin) {
  __shared__ float red[kCoopThreads * 16];
  const int tpr = C >> 3, rpi = kCoopThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int row0 = blockIdx.x * rows_per_block;
  const int row1 = min(M, row0 + rows_per_block);
  const size_t col = (size_t)tx * 8;
  {
    float s0[8], s1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s0[k] = s1[k] = 0.f;
    rows_stats(x, C, col, row0 + ty, row1, rpi, s0, s1);
    fold_block<kCoopThreads>(red, s0, s1, ws + (size_t)blockIdx.x * 2 * C, C, tpr, rpi);
  }
  grid_barrier(bar);
  if (blockIdx.x == 0 && threadIdx.x == 0 && fin.num_batches_tracked) *fin.num_batches_tracked += 1;
  for (int cb = blockIdx.x; cb * 8 < C; cb += gridDim.x) {
    float t0, t1;
    if (coop_fold_channels(red, ws, C, cb, t0, t1)) finalize_fwd_channel(cb * 8 + (threadIdx.x & 7), t0, t1, M, C, fin);
    __syncthreads();
  }
  grid_barrier(bar);
  float sc[8], sh[8];
  load8_cg(fin.coef + tx * 8, sc);
  load8_cg(fin.coef + C + tx * 8, sh);
  rows_apply_fwd<RELU, RES>(x, res, y, C, col, row0 + ty, row1, rpi, sc, sh);
}

template <bool RELU, bool RES>
__global__ void __launch_bounds__(kCoopThreads, 2) coop_bwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                                   const __nv_bfloat16* __restrict__ dy,
                                                                   const __nv_bfloat16* __restrict__ y,
                                                                   __nv_bfloat16* __restrict__ dx,
                                                                   __nv_bfloat16* __restrict__ dres,
                                                                   const float* __restrict__ mean,
                                                                   const float* __restrict__ invstd, float* ws, int M,
                                                                   int C, int rows_per_block, unsigned int* bar,
                                                                   BwdFinal fin) {
  __shared__ float red[kCoopThreads * 16];
  const int tpr = C >> 3, rpi = kCoopThreads / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int row0 = blockIdx.x * rows_per_block;
  const int row1 = min(M, row0 + rows_per_block);
  const size_t col = (size_t)tx * 8;
  float mu[8], is[8];
  load8(mean + tx * 8, mu);
  load8(invstd + tx * 8, is);
  {
    float s0[8], s1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s0[k] = s1[k] = 0.f;
    rows_bwd_stats<RELU>(x, dy, y, C, col, row0 + ty, row1, rpi, mu, is, s0, s1);
    fold_block<kCoopThreads>(red, s0, s1, ws + (size_t)blockIdx.x * 2 * C, C, tpr, rpi);
  }
  grid_barrier(bar);
  for (int cb = blockIdx.x; cb * 8 < C; cb += gridDim.x) {
    float t0, t1;
    if (coop_fold_channels(red, ws, C, cb, t0, t1)) finalize_bwd_channel(cb * 8 + (threadIdx.x & 7), t0, t1, M, C, fin);
    __syncthreads();
  }
  grid_barrier(bar);
  float a[8], b[8], cc[8];
  load8_cg(fin.coef + tx * 8, a);
  load8_cg(fin.coef + C + tx * 8, b);
  load8_cg(fin.coef + 2 * C + tx * 8, cc);
  rows_apply_bwd<RELU, RES>(x, dy, y, dx, dres, C, col, row0 + ty, row1, rpi, mu, is, a, b, cc);
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
  return (size_t)max(p.partial_blocks, 2 * sm_count()) * 2 * c + 3 * (size_t)c;  // either plan's partial table
}

// Cooperative (single-launch) plan: 512-thread CTAs, at most ``max_blocks`` co-resident.
struct CoopPlan {
  int blocks, rows_per_block;
};
static CoopPlan make_coop_plan(int m, int c, int max_blocks) {
  const int rpi = kCoopThreads / (c / 8);
  const int iters = (m + rpi - 1) / rpi;
  CoopPlan p;
  p.blocks = max(1, min(min(min(max_blocks, 2 * sm_count()), 32 << kBarGroupShift), (iters + 1) / 2));
  int ipb = (iters + p.blocks - 1) / p.blocks;
  if (ipb > kUnroll) ipb = (ipb + kUnroll - 1) / kUnroll * kUnroll;  // whole software-pipelined groups per CTA
  p.rows_per_block = ipb * rpi;
  p.blocks = (m + p.rows_per_block - 1) / p.rows_per_block;
  return p;
}

template <typename K>
static int coop_max_blocks(K kernel) {
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kCoopThreads, 0) != cudaSuccess) occ = 0;
  return occ * sm_count();
}

}  // namespace bn
}  // namespace mfl

#include "bn_cluster.cuh"

using namespace mfl;

// -1 automatic (default; MFL_BN_CLUSTER in the environment), 0 never use the cluster back-end, 1 wherever a plan exists
extern "C" int mfl_bn_set_cluster_mode(int mode) {
  const int old = bn::g_cluster_mode;
  bn::g_cluster_mode = mode < 0 ? -1 : (mode ? 1 : 0);
  return old;
}

extern "C" size_t mfl_bn_workspace_bytes(int m, int c) {
  if (m <= 0 || !bn::supported_channels(c)) return 0;
  return bn::workspace_floats(m, c) * sizeof(float);
}

namespace {

// Launch with the programmatic-stream-serialization attribute (see pdl_wait above).
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = getenv("MFL_BN_NO_PDL") ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

template <bool RELU, bool RES>
int launch_coop_fwd(const __nv_bfloat16* x, const __nv_bfloat16* res, __nv_bfloat16* y, float* ws, int m, int c,
                    unsigned int* bar, bn::FwdFinal fin, cudaStream_t st) {
  static const int max_blocks = bn::coop_max_blocks(bn::coop_fwd_kernel<RELU, RES>);
  MFL_ARG(max_blocks > 0, "mfl_bn_fwd: cooperative kernel does not fit on this device");
  bn::CoopPlan p = bn::make_coop_plan(m, c, max_blocks);
  void* args[] = {&x, &res, &y, &ws, &m, &c, &p.rows_per_block, &bar, &fin};
  if (getenv("MFL_BN_COOP_PLAIN")) {  // experiment: same kernel through a plain launch
    bn::coop_fwd_kernel<RELU, RES><<<p.blocks, bn::kCoopThreads, 0, st>>>(x, res, y, ws, m, c, p.rows_per_block, bar, fin);
    return 0;
  }
  MFL_CUDA(cudaLaunchCooperativeKernel((const void*)bn::coop_fwd_kernel<RELU, RES>, dim3(p.blocks),
                                       dim3(bn::kCoopThreads), args, 0, st));
  return 0;
}

template <bool RELU, bool RES>
int launch_coop_bwd(const __nv_bfloat16* x, const __nv_bfloat16* dy, const __nv_bfloat16* y, __nv_bfloat16* dx,
                    __nv_bfloat16* dres, const float* mean, const float* invstd, float* ws, int m, int c,
                    unsigned int* bar, bn::BwdFinal fin, cudaStream_t st) {
  static const int max_blocks = bn::coop_max_blocks(bn::coop_bwd_kernel<RELU, RES>);
  MFL_ARG(max_blocks > 0, "mfl_bn_bwd: cooperative kernel does not fit on this device");
  bn::CoopPlan p = bn::make_coop_plan(m, c, max_blocks);
  void* args[] = {&x, &dy, &y, &dx, &dres, &mean, &invstd, &ws, &m, &c, &p.rows_per_block, &bar, &fin};
  if (getenv("MFL_BN_COOP_PLAIN")) {
    bn::coop_bwd_kernel<RELU, RES><<<p.blocks, bn::kCoopThreads, 0, st>>>(x, dy, y, dx, dres, mean, invstd, ws, m, c,
                                                                          p.rows_per_block, bar, fin);
    return 0;
  }
  MFL_CUDA(cudaLaunchCooperativeKernel((const void*)bn::coop_bwd_kernel<RELU, RES>, dim3(p.blocks),
                                       dim3(bn::kCoopThreads), args, 0, st));
  return 0;
}

}  // namespace

extern "C" int mfl_bn_fwd(void* y, const void* x, const void* residual, const float* gamma, const float* beta,
                          float* running_mean, float* running_var, long long* num_batches_tracked,
                          float* save_mean, float* save_invstd, float* workspace, size_t workspace_bytes,
                          unsigned int* barrier, int m, int c, float eps, float momentum, int training, int relu,
                          void* stream) {
  MFL_ARG(y && x && workspace, "mfl_bn_fwd: null tensor");
  MFL_ARG(m > 0 && bn::supported_channels(c), "mfl_bn_fwd: channels must be a power of two in [8, 2048], got %d", c);
  MFL_ARG(workspace_bytes >= mfl_bn_workspace_bytes(m, c), "mfl_bn_fwd: workspace too small");
  MFL_ARG(training ? (save_mean && save_invstd) : (running_mean && running_var), "mfl_bn_fwd: missing statistics");
  auto st = static_cast<cudaStream_t>(stream);
  const bn::Plan p = bn::make_plan(m, c);
  auto* xb = static_cast<const __nv_bfloat16*>(x);
  auto* rb = static_cast<const __nv_bfloat16*>(residual);
  auto* yb = static_cast<__nv_bfloat16*>(y);
  float* coef = workspace + bn::workspace_floats(m, c) - 3 * (size_t)c;
  const bn::FwdFinal fin{gamma, beta, running_mean, running_var, num_batches_tracked, save_mean, save_invstd, coef,
                         eps, momentum};
  bn::ClPlan cp;
  if (training && !barrier && bn::make_cluster_plan(m, c, 1, &cp)) {  // one launch on thread-block clusters
    if (relu && rb) return bn::launch_cluster(bn::cluster_fwd_kernel<true, true>, cp, st, xb, rb, yb, m, c, cp.slice, cp.rows_per_cta, cp.keep, fin);
    if (relu) return bn::launch_cluster(bn::cluster_fwd_kernel<true, false>, cp, st, xb, rb, yb, m, c, cp.slice, cp.rows_per_cta, cp.keep, fin);
    if (rb) return bn::launch_cluster(bn::cluster_fwd_kernel<false, true>, cp, st, xb, rb, yb, m, c, cp.slice, cp.rows_per_cta, cp.keep, fin);
    return bn::launch_cluster(bn::cluster_fwd_kernel<false, false>, cp, st, xb, rb, yb, m, c, cp.slice, cp.rows_per_cta, cp.keep, fin);
  }
  if (training && barrier) {  // one cooperative launch
    if (relu && rb) return launch_coop_fwd<true, true>(xb, rb, yb, workspace, m, c, barrier, fin, st);
    if (relu) return launch_coop_fwd<true, false>(xb, rb, yb, workspace, m, c, barrier, fin, st);
    if (rb) return launch_coop_fwd<false, true>(xb, rb, yb, workspace, m, c, barrier, fin, st);
    return launch_coop_fwd<false, false>(xb, rb, yb, workspace, m, c, barrier, fin, st);
  }
  if (training) {
    bn::stats_partial_kernel<<<p.partial_blocks, bn::kThreads, 0, st>>>(xb, workspace, m, c, p.rows_per_block);
    MFL_CUDA(launch_pdl(bn::finalize_fwd_kernel, dim3((c + 31) / 32), dim3(32, 32), st, workspace, p.partial_blocks, c, m,
                        fin));
  } else {
    bn::eval_coef_kernel<<<(c + 127) / 128, 128, 0, st>>>(c, gamma, beta, running_mean, running_var, coef, eps);
  }
  const dim3 g(p.apply_blocks), b(bn::kThreads);
  const float* cf = coef;
  if (relu && rb) MFL_CUDA(launch_pdl(bn::apply_fwd_kernel<true, true>, g, b, st, xb, rb, yb, cf, m, c));
  else if (relu) MFL_CUDA(launch_pdl(bn::apply_fwd_kernel<true, false>, g, b, st, xb, rb, yb, cf, m, c));
  else if (rb) MFL_CUDA(launch_pdl(bn::apply_fwd_kernel<false, true>, g, b, st, xb, rb, yb, cf, m, c));
  else MFL_CUDA(launch_pdl(bn::apply_fwd_kernel<false, false>, g, b, st, xb, rb, yb, cf, m, c));
  MFL_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int mfl_bn_bwd(void* dx, void* dres, float* dgamma, float* dbeta, int accumulate, const void* dy,
                          const void* y, const void* x, const float* gamma, const float* save_mean,
                          const float* save_invstd, float* workspace, size_t workspace_bytes, unsigned int* barrier,
                          int m, int c, int relu, void* stream) {
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
  float* coef = workspace + bn::workspace_floats(m, c) - 3 * (size_t)c;
  const bn::BwdFinal fin{gamma, save_invstd, dgamma, dbeta, accumulate, coef};
  bn::ClPlan cp;
  if (!barrier && bn::make_cluster_plan(m, c, 2, &cp)) {  // one launch on thread-block clusters
    if (relu && drb)
      return bn::launch_cluster(bn::cluster_bwd_kernel<true, true>, cp, st, xb, db, yb, dxb, drb, save_mean, save_invstd, m, c, cp.slice, cp.rows_per_cta, cp.keep, fin);
    if (relu)
      return bn::launch_cluster(bn::cluster_bwd_kernel<true, false>, cp, st, xb, db, yb, dxb, drb, save_mean, save_invstd, m, c, cp.slice, cp.rows_per_cta, cp.keep, fin);
    if (drb)
      return bn::launch_cluster(bn::cluster_bwd_kernel<false, true>, cp, st, xb, db, yb, dxb, drb, save_mean, save_invstd, m, c, cp.slice, cp.rows_per_cta, cp.keep, fin);
    return bn::launch_cluster(bn::cluster_bwd_kernel<false, false>, cp, st, xb, db, yb, dxb, drb, save_mean, save_invstd, m, c, cp.slice, cp.rows_per_cta, cp.keep, fin);
  }
  if (barrier) {  // one cooperative launch
    if (relu && drb)
      return launch_coop_bwd<true, true>(xb, db, yb, dxb, drb, save_mean, save_invstd, workspace, m, c, barrier, fin, st);
    if (relu)
      return launch_coop_bwd<true, false>(xb, db, yb, dxb, drb, save_mean, save_invstd, workspace, m, c, barrier, fin, st);
    if (drb)
      return launch_coop_bwd<false, true>(xb, db, yb, dxb, drb, save_mean, save_invstd, workspace, m, c, barrier, fin, st);
    return launch_coop_bwd<false, false>(xb, db, yb, dxb, drb, save_mean, save_invstd, workspace, m, c, barrier, fin, st);
  }
  if (relu)
    bn::bwd_partial_kernel<true><<<p.partial_blocks, bn::kThreads, 0, st>>>(xb, db, yb, save_mean, save_invstd,
                                                                             workspace, m, c, p.rows_per_block);
  else
    bn::bwd_partial_kernel<false><<<p.partial_blocks, bn::kThreads, 0, st>>>(xb, db, yb, save_mean, save_invstd,
                                                                              workspace, m, c, p.rows_per_block);
  MFL_CUDA(launch_pdl(bn::finalize_bwd_kernel, dim3((c + 31) / 32), dim3(32, 32), st, workspace, p.partial_blocks, c, m,
                      fin));
  const dim3 g(p.apply_blocks), b(bn::kThreads);
  const float* cf = coef;
  if (relu && drb)
    MFL_CUDA(launch_pdl(bn::apply_bwd_kernel<true, true>, g, b, st, xb, db, yb, dxb, drb, save_mean, save_invstd, cf, m, c));
  else if (relu)
    MFL_CUDA(launch_pdl(bn::apply_bwd_kernel<true, false>, g, b, st, xb, db, yb, dxb, drb, save_mean, save_invstd, cf, m, c));
  else if (drb)
    MFL_CUDA(launch_pdl(bn::apply_bwd_kernel<false, true>, g, b, st, xb, db, yb, dxb, drb, save_mean, save_invstd, cf, m, c));
  else
    MFL_CUDA(launch_pdl(bn::apply_bwd_kernel<false, false>, g, b, st, xb, db, yb, dxb, drb, save_mean, save_invstd, cf, m, c));
  MFL_CUDA(cudaGetLastError());
  return 0;
}
