// K14, back-end C: ONE launch per direction built on thread-block clusters, no grid barrier and no second launch.
//
// BatchNorm statistics are per channel, so the tensor is cut along CHANNELS first: a cluster owns a slice of S
// channels (S * 2 bytes contiguous per NHWC row: whole 32-byte sectors for S >= 16), its R CTAs split the rows.
//
//   phase 1  every CTA streams its rows once: partial sums per channel, and (when it fits) the rows themselves are
//            parked in shared memory
//   reduce   warp shuffles -> per-CTA partials in shared memory -> hardware cluster barrier -> every CTA adds the R
//            partials through distributed shared memory in rank order (deterministic) and finalizes its S channels
//            redundantly; rank 0 writes the saved / running statistics (forward) or dgamma / dbeta (backward)
//   phase 2  normalise (+ residual, + ReLU) / input gradient from the parked rows (or a second, L2-resident read)
//
// Against the three-launch back-end this removes two dependent launches (~3.8 us each inside a CUDA graph) and the
// finalize kernel per direction, and the forward reads x from global memory once instead of twice.  The cluster
// barrier is a hardware barrier among <= 8 co-scheduled CTAs, not a polled counter in L2 like the grid barrier of the
// cooperative back-end.  Included by bn_relu.cu (shares its finalize math and vector helpers).
#pragma once

namespace mfl {
namespace bn {

constexpr int kClThreads = 512;         // largest CTA (the plan picks 256 or 512 threads by rows per CTA)
constexpr int kClUnrollFwd = 8;         // one input stream per row in forward
constexpr int kClUnrollBwd = 4;         // three input streams per row in backward
constexpr int kClWarps = kClThreads / 32;
constexpr int kClMaxSlice = 64;          // channels per cluster (tpr <= 8 threads per row)
constexpr size_t kClSmemLimit = 200 * 1024;

struct ClPlan {
  int slice, n_slices, ranks, rows_per_cta, keep, threads;
  size_t smem;
};

__device__ __forceinline__ uint32_t cl_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cl_size() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
// warp-level (.aligned) forms behind an explicit reconvergence: they follow loops that only part of a warp runs, and
// per-thread arrivals of 2-8 K threads per cluster cost microseconds
__device__ __forceinline__ void cl_arrive() {
  __syncwarp();
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cl_wait() {
  __syncwarp();
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float cl_ld_remote(const float* local_smem_ptr, uint32_t rank) {
  uint32_t addr;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;"
               : "=r"(addr) : "r"((uint32_t)__cvta_generic_to_shared(local_smem_ptr)), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

// Sum the 16 per-thread accumulators over the threads of the CTA that share a channel group (same tx): xor-shuffles
// over the lane bits above log2(tpr), then one partial per warp in shared memory, then kClWarps adds.
// part[s * S + c] receives this CTA's total of sum s for slice channel c.
__device__ __forceinline__ void cl_fold_cta(float* wred, float* part, const float* s0, const float* s1, int S, int tpr) {
  float v[16];
#pragma unroll
  for (int k = 0; k < 8; ++k) { v[k] = s0[k]; v[8 + k] = s1[k]; }
  for (int m = tpr; m < 32; m <<= 1) {
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] += __shfl_xor_sync(0xffffffffu, v[k], m);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane < tpr) {
#pragma unroll
    for (int k = 0; k < 16; ++k) wred[(warp * kClMaxSlice / 8 + lane) * 16 + k] = v[k];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < 2 * S; o += blockDim.x) {
    const int s = o / S, c = o - s * S;
    float acc = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) acc += wred[(w * kClMaxSlice / 8 + (c >> 3)) * 16 + s * 8 + (c & 7)];
    part[o] = acc;
  }
}

// tot[o] = sum over the cluster's CTAs (rank order) of part[o]; call between cl_arrive/cl_wait pairs as below
__device__ __forceinline__ void cl_total(float* tot, const float* part, int S, uint32_t R) {
  for (int o = threadIdx.x; o < 2 * S; o += blockDim.x) {
    float v[8];   // all remote loads in flight before the first add (one DSMEM latency, not R)
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j) v[j] = j < R ? cl_ld_remote(part + o, j) : 0.f;
    float acc = 0.f;
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j) acc += v[j];
    for (uint32_t j = 8; j < R; ++j) acc += cl_ld_remote(part + o, j);
    tot[o] = acc;
  }
}

struct ClSmem {
  float* wred;   // [kClWarps][kClMaxSlice / 8][16]
  float* part;   // [2 S]   this CTA's partial sums (read by the whole cluster)
  float* tot;    // [2 S]
  float* coef;   // [3 S]
  uint4* slab;   // parked rows
};
__device__ __forceinline__ ClSmem cl_carve(uint8_t* base) {
  ClSmem s;
  s.wred = reinterpret_cast<float*>(base);
  s.part = s.wred + kClWarps * (kClMaxSlice / 8) * 16;
  s.tot = s.part + 2 * kClMaxSlice;
  s.coef = s.tot + 2 * kClMaxSlice;
  s.slab = reinterpret_cast<uint4*>(s.coef + 4 * kClMaxSlice);
  return s;
}
constexpr size_t kClFixedSmem = (kClWarps * (kClMaxSlice / 8) * 16 + 8 * kClMaxSlice) * sizeof(float);

template <bool RELU, bool RES>
__global__ void __launch_bounds__(kClThreads) cluster_fwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                                 const __nv_bfloat16* __restrict__ res,
                                                                 __nv_bfloat16* __restrict__ y, int M, int C, int S,
                                                                 int rows_per_cta, int keep, FwdFinal fin) {
  extern __shared__ __align__(16) uint8_t cl_smem_raw[];
  const ClSmem sm = cl_carve(cl_smem_raw);
  const uint32_t rank = cl_rank(), R = cl_size();
  const int tpr = S >> 3, rpi = (int)blockDim.x / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int c0 = blockIdx.y * S;
  const size_t col = (size_t)c0 + tx * 8;
  const int r0 = (int)rank * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float s0[8], s1[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s0[k] = s1[k] = 0.f;
  for (int r = r0 + ty; r < r1; r += kClUnrollFwd * rpi) {
    uint4 v[kClUnrollFwd];
#pragma unroll
    for (int u = 0; u < kClUnrollFwd; ++u) v[u] = ld_row(x + (size_t)(r + u * rpi) * C + col, r + u * rpi < r1);
#pragma unroll
    for (int u = 0; u < kClUnrollFwd; ++u) {
      if (keep && r + u * rpi < r1) sm.slab[(size_t)(r + u * rpi - r0) * tpr + tx] = v[u];
      float f[8];
      V::unpack(v[u], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) { s0[k] += f[k]; s1[k] = fmaf(f[k], f[k], s1[k]); }
    }
  }
  cl_fold_cta(sm.wred, sm.part, s0, s1, S, tpr);
  cl_arrive();
  cl_wait();                                   // every CTA's partials are in its shared memory
  cl_total(sm.tot, sm.part, S, R);
  cl_arrive();                                 // this thread is done with the other CTAs' shared memory
  __syncthreads();
  if (threadIdx.x < S) {
    const int c = threadIdx.x;
    const double mean_d = (double)sm.tot[c] / M;
    const double var_d = (double)sm.tot[S + c] / M - mean_d * mean_d;
    const float mean = (float)mean_d;
    const float var = var_d > 0.0 ? (float)var_d : 0.f;
    const float invstd = rsqrtf(var + fin.eps);
    const float g = fin.gamma ? fin.gamma[c0 + c] : 1.f, b = fin.beta ? fin.beta[c0 + c] : 0.f;
    sm.coef[c] = g * invstd;
    sm.coef[S + c] = b - mean * g * invstd;
    if (rank == 0) {
      fin.save_mean[c0 + c] = mean;
      fin.save_invstd[c0 + c] = invstd;
      if (fin.running_mean) fin.running_mean[c0 + c] = (1.f - fin.momentum) * fin.running_mean[c0 + c] + fin.momentum * mean;
      if (fin.running_var) {
        const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
        fin.running_var[c0 + c] = (1.f - fin.momentum) * fin.running_var[c0 + c] + fin.momentum * unbiased;
      }
      if (c == 0 && blockIdx.y == 0 && fin.num_batches_tracked) *fin.num_batches_tracked += 1;
    }
  }
  __syncthreads();
  float sc[8], sh[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { sc[k] = sm.coef[tx * 8 + k]; sh[k] = sm.coef[S + tx * 8 + k]; }
  for (int r = r0 + ty; r < r1; r += kClUnrollFwd * rpi) {
    uint4 vx[kClUnrollFwd], vr[kClUnrollFwd];
#pragma unroll
    for (int u = 0; u < kClUnrollFwd; ++u) {
      const bool ok = r + u * rpi < r1;
      const size_t off = (size_t)(r + u * rpi) * C + col;
      vx[u] = !ok ? make_uint4(0, 0, 0, 0) : (keep ? sm.slab[(size_t)(r + u * rpi - r0) * tpr + tx] : ld_stream(x + off));
      vr[u] = RES ? ld_row(res + off, ok) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < kClUnrollFwd; ++u) {
      if (r + u * rpi < r1) {
        float f[8], fr[8];
        V::unpack(vx[u], f);
        if (RES) V::unpack(vr[u], fr);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          float t = fmaf(f[k], sc[k], sh[k]);
          if (RES) t += fr[k];
          f[k] = RELU ? fmaxf(t, 0.f) : t;
        }
        *reinterpret_cast<uint4*>(y + (size_t)(r + u * rpi) * C + col) = V::pack(f);
      }
    }
  }
  cl_wait();                                   // nobody is still reading this CTA's partials
}

template <bool RELU, bool RES>
__global__ void __launch_bounds__(kClThreads) cluster_bwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                                 const __nv_bfloat16* __restrict__ dy,
                                                                 const __nv_bfloat16* __restrict__ y,
                                                                 __nv_bfloat16* __restrict__ dx,
                                                                 __nv_bfloat16* __restrict__ dres,
                                                                 const float* __restrict__ mean,
                                                                 const float* __restrict__ invstd, int M, int C, int S,
                                                                 int rows_per_cta, int keep, BwdFinal fin) {
  extern __shared__ __align__(16) uint8_t cl_smem_raw[];
  const ClSmem sm = cl_carve(cl_smem_raw);
  const uint32_t rank = cl_rank(), R = cl_size();
  const int tpr = S >> 3, rpi = (int)blockDim.x / tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int c0 = blockIdx.y * S;
  const size_t col = (size_t)c0 + tx * 8;
  const int r0 = (int)rank * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  uint4* slab_x = sm.slab;
  uint4* slab_g = sm.slab + (size_t)rows_per_cta * tpr;
  constexpr int U = kClUnrollBwd;
  float mu[8], is[8], s0[8], s1[8];
  load8(mean + col, mu);
  load8(invstd + col, is);
#pragma unroll
  for (int k = 0; k < 8; ++k) s0[k] = s1[k] = 0.f;
  for (int r = r0 + ty; r < r1; r += U * rpi) {
    uint4 vx[U], vd[U], vy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t off = (size_t)(r + u * rpi) * C + col;
      const bool ok = r + u * rpi < r1;
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
        fg[k] = g;
        s0[k] += g;
        s1[k] = fmaf(g, (fx[k] - mu[k]) * is[k], s1[k]);
      }
      if (keep && r + u * rpi < r1) {
        const size_t i = (size_t)(r + u * rpi - r0) * tpr + tx;
        slab_x[i] = vx[u];
        slab_g[i] = V::pack(fg);   // dy or 0: exact in bf16
      }
    }
  }
  cl_fold_cta(sm.wred, sm.part, s0, s1, S, tpr);
  cl_arrive();
  cl_wait();
  cl_total(sm.tot, sm.part, S, R);
  cl_arrive();
  __syncthreads();
  if (threadIdx.x < S) {
    const int c = threadIdx.x;
    const float t0 = sm.tot[c], t1 = sm.tot[S + c];
    sm.coef[c] = (fin.gamma ? fin.gamma[c0 + c] : 1.f) * fin.save_invstd[c0 + c];
    sm.coef[S + c] = t0 / (float)M;
    sm.coef[2 * S + c] = t1 / (float)M;
    if (rank == 0) {
      if (fin.dgamma) fin.dgamma[c0 + c] = (fin.accumulate ? fin.dgamma[c0 + c] : 0.f) + t1;
      if (fin.dbeta) fin.dbeta[c0 + c] = (fin.accumulate ? fin.dbeta[c0 + c] : 0.f) + t0;
    }
  }
  __syncthreads();
  float a[8], b[8], cc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    a[k] = sm.coef[tx * 8 + k];
    b[k] = sm.coef[S + tx * 8 + k];
    cc[k] = sm.coef[2 * S + tx * 8 + k];
  }
  for (int r = r0 + ty; r < r1; r += U * rpi) {
    uint4 vx[U], vg[U], vy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const bool ok = r + u * rpi < r1;
      const size_t off = (size_t)(r + u * rpi) * C + col;
      if (keep) {
        const size_t i = (size_t)(r + u * rpi - r0) * tpr + tx;
        vx[u] = ok ? slab_x[i] : make_uint4(0, 0, 0, 0);
        vg[u] = ok ? slab_g[i] : make_uint4(0, 0, 0, 0);
        vy[u] = make_uint4(0, 0, 0, 0);
      } else {
        vx[u] = ld_row(x + off, ok);
        vg[u] = ld_row(dy + off, ok);
        vy[u] = RELU ? ld_row(y + off, ok) : make_uint4(0, 0, 0, 0);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (r + u * rpi < r1) {
        float fx[8], fg[8], fy[8];
        V::unpack(vx[u], fx);
        V::unpack(vg[u], fg);
        if (RELU && !keep) V::unpack(vy[u], fy);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float g = (keep || !RELU || fy[k] > 0.f) ? fg[k] : 0.f;
          fg[k] = g;
          fx[k] = a[k] * (g - b[k] - (fx[k] - mu[k]) * is[k] * cc[k]);
        }
        const size_t off = (size_t)(r + u * rpi) * C + col;
        *reinterpret_cast<uint4*>(dx + off) = V::pack(fx);
        if (RES) *reinterpret_cast<uint4*>(dres + off) = V::pack(fg);
      }
    }
  }
  cl_wait();
}

// ---- host side --------------------------------------------------------------------------------------------------
// mode: -1 automatic, 0 off, 1 wherever a plan exists (MFL_BN_CLUSTER in the environment; tests / bn_bench flip it)
static int g_cluster_mode = [] { const char* e = getenv("MFL_BN_CLUSTER"); return e ? atoi(e) : -1; }();
static size_t g_cluster_auto_max_elems = [] {
  const char* e = getenv("MFL_BN_CLUSTER_MAX_ELEMS");
  return e ? (size_t)atoll(e) : (size_t)2200000;   // measured crossover (bn_bench): 128x512x4x4 and 128x256x8x8 win
}();

// 16-CTA clusters are a non-portable size: they need an opt-in per kernel and a GPC with 16 free SMs
static int g_cluster_max_ranks = [] { const char* e = getenv("MFL_BN_CLUSTER_RANKS"); return e ? atoi(e) : 16; }();

// Can this device co-schedule a 16-CTA cluster of these kernels at all?  Asked once, with the largest shared-memory
// request and the backward kernel (the heavier of the two).
static bool cluster16_fits() {
  static int fits = -1;
  if (fits < 0) {
    auto kernel = cluster_bwd_kernel<true, true>;
    fits = 0;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kClSmemLimit) == cudaSuccess &&
        cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(16, 8);
      cfg.blockDim = dim3(kClThreads);
      cfg.dynamicSmemBytes = kClSmemLimit;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = 16;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) == cudaSuccess && n > 0) fits = 1;
    }
    cudaGetLastError();
  }
  return fits == 1;
}

static bool make_cluster_plan(int m, int c, int slabs, ClPlan* p) {
  if (g_cluster_mode == 0 || c < 16) return false;
  if (g_cluster_mode < 0 && (size_t)m * c > g_cluster_auto_max_elems) return false;
  const int target = std::max(1, sm_count() - 20);
  int S = 16;
  // widen the slice while there are still enough (slice, rank) CTAs to cover the machine with 8-CTA clusters
  while (S < kClMaxSlice && S * 2 <= c && (c / (S * 2)) * 8 >= target) S *= 2;
  int R = 1;
  const int max_r = (g_cluster_max_ranks > 8 && cluster16_fits()) ? 16 : std::min(g_cluster_max_ranks, 8);
  while (R < max_r && (c / S) * R < target && (m + 2 * R - 1) / (2 * R) >= 64) R *= 2;
  p->slice = S;
  p->n_slices = c / S;
  p->ranks = R;
  // 512 threads (more loads in flight) once a CTA has at least eight row vectors per thread at that size
  p->threads = ((m + R - 1) / R) * (S / 8) >= 8 * 512 ? 512 : 256;
  const int rpi = p->threads / (S / 8);
  int rows = (m + R - 1) / R;
  rows = (rows + rpi - 1) / rpi * rpi;
  p->rows_per_cta = rows;
  const size_t slab = (size_t)rows * S * 2 * slabs;
  p->keep = kClFixedSmem + slab <= kClSmemLimit;
  p->smem = kClFixedSmem + (p->keep ? slab : 0);
  return true;
}

template <typename K, typename... Args>
static int launch_cluster(K kernel, const ClPlan& p, cudaStream_t st, Args... args) {
  MFL_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kClSmemLimit));
  if (p.ranks > 8) MFL_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(p.ranks, p.n_slices);
  cfg.blockDim = dim3(p.threads);
  cfg.dynamicSmemBytes = p.smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = p.ranks;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  MFL_CUDA(cudaLaunchKernelEx(&cfg, kernel, args...));
  return 0;
}

}  // namespace bn
}  // namespace mfl
