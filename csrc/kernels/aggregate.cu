// K1 / K3: weighted community-model aggregation kernels.
//
// What they replace (behaviour, not code): the reference controller's CPU loop
//   out_v = sum_l T(double(x_{l,v}) * w_l)          federated_average.cc:15-59,101-145
//   S += w_new x_new - w_old x_old ; C = T(S / z)   federated_rolling_average_base.cc:175-293
// and, for the fused variant, the three gRPC fan-outs around it (controller.cc:201-259 gather,
// :696-759 RunTask broadcast, :571-647 EvaluateModel broadcast).
//
// B200 design: the op is pure bandwidth (L reads + 1 write per element), so the kernels are
// persistent, 16-byte vectorised, keep >= 8 independent loads in flight per thread to cover the
// ~2 us NVLink load latency, accumulate in fp32 registers and cast once in the epilogue.  The
// fused kernel does reduce-scatter + all-gather in a single launch: rank r pulls shard r from
// every learner slot over NVLink (peer ld.global), and pushes the reduced shard into every
// rank's community buffer (peer st.global); two in-kernel flag barriers replace host syncs.
#include <algorithm>
#include <type_traits>

#include "common.cuh"

namespace mfl {

struct AggArgs {
  const void* src[MFL_MAX_SRC];
  double wd[MFL_MAX_SRC];
  float w[MFL_MAX_SRC];
  int n;
};

constexpr int kUnit = 8;  // elements per work unit (16 B of 16-bit types, 32 B of fp32)

template <int DT> struct Unit;  // load/store one 8-element unit as float[8]
template <> struct Unit<MFL_FLOAT32> {
  template <bool PEER> __device__ static void load(const void* base, size_t u, float* f) {
    const uint4* p = reinterpret_cast<const uint4*>(base) + 2 * u;
    uint4 a = PEER ? ld_peer(p) : ld_stream(p);
    uint4 b = PEER ? ld_peer(p + 1) : ld_stream(p + 1);
    VecTraits<MFL_FLOAT32>::unpack(a, f);
    VecTraits<MFL_FLOAT32>::unpack(b, f + 4);
  }
  __device__ static void store(void* base, size_t u, const float* f) {
    uint4* p = reinterpret_cast<uint4*>(base) + 2 * u;
    st_stream(p, VecTraits<MFL_FLOAT32>::pack(f));
    st_stream(p + 1, VecTraits<MFL_FLOAT32>::pack(f + 4));
  }
};
template <int DT16> struct Unit16 {
  template <bool PEER> __device__ static void load(const void* base, size_t u, float* f) {
    const uint4* p = reinterpret_cast<const uint4*>(base) + u;
    VecTraits<DT16>::unpack(PEER ? ld_peer(p) : ld_stream(p), f);
  }
  __device__ static void store(void* base, size_t u, const float* f) {
    st_stream(reinterpret_cast<uint4*>(base) + u, VecTraits<DT16>::pack(f));
  }
};
template <> struct Unit<MFL_BFLOAT16> : Unit16<MFL_BFLOAT16> {};
template <> struct Unit<MFL_FLOAT16> : Unit16<MFL_FLOAT16> {};

// Raw (unconverted) staging so that all loads of a batch are issued before any use.
template <int DT> struct Raw { uint4 v[DT == MFL_FLOAT32 ? 2 : 1]; };
template <int DT, bool PEER> __device__ __forceinline__ Raw<DT> load_raw(const void* base, size_t u) {
  Raw<DT> r;
  if constexpr (DT == MFL_FLOAT32) {
    const uint4* p = reinterpret_cast<const uint4*>(base) + 2 * u;
    r.v[0] = PEER ? ld_peer(p) : ld_stream(p);
    r.v[1] = PEER ? ld_peer(p + 1) : ld_stream(p + 1);
  } else {
    const uint4* p = reinterpret_cast<const uint4*>(base) + u;
    r.v[0] = PEER ? ld_peer(p) : ld_stream(p);
  }
  return r;
}
template <int DT> __device__ __forceinline__ void fma_raw(const Raw<DT>& r, float w, float* acc) {
  float f[kUnit];
  if constexpr (DT == MFL_FLOAT32) {
    VecTraits<MFL_FLOAT32>::unpack(r.v[0], f);
    VecTraits<MFL_FLOAT32>::unpack(r.v[1], f + 4);
  } else {
    VecTraits<DT>::unpack(r.v[0], f);
  }
#pragma unroll
  for (int i = 0; i < kUnit; ++i) acc[i] = fmaf(w, f[i], acc[i]);
}

// acc[8] += sum_l w_l * src_l[unit u]; loads are batched kBatch deep for memory-level parallelism.
template <int DT, bool PEER>
__device__ __forceinline__ void reduce_unit(const AggArgs& a, size_t byte_off, size_t u, float* acc) {
  // the fused kernel runs 2 CTAs x 512 threads per SM under a 64-register cap: half the batch depth,
  // twice the resident warps, same bytes in flight
  constexpr int kBatch = (DT == MFL_FLOAT32) ? (PEER ? 2 : 4) : (PEER ? 4 : 8);
  int l = 0;
  for (; l + kBatch <= a.n; l += kBatch) {
    Raw<DT> r[kBatch];
#pragma unroll
    for (int j = 0; j < kBatch; ++j)
      r[j] = load_raw<DT, PEER>(reinterpret_cast<const char*>(a.src[l + j]) + byte_off, u);
#pragma unroll
    for (int j = 0; j < kBatch; ++j) fma_raw<DT>(r[j], a.w[l + j], acc);
  }
  if (l < a.n) {
    Raw<DT> r[kBatch];
#pragma unroll
    for (int j = 0; j < kBatch; ++j)
      if (l + j < a.n) r[j] = load_raw<DT, PEER>(reinterpret_cast<const char*>(a.src[l + j]) + byte_off, u);
#pragma unroll
    for (int j = 0; j < kBatch; ++j)
      if (l + j < a.n) fma_raw<DT>(r[j], a.w[l + j], acc);
  }
}

// ------------------------------------------------------------------------------------------
// Local (single-GPU) fast path: float-family src/dst, fp32 accumulate.
// ------------------------------------------------------------------------------------------
template <int SRC, int DST>
__global__ void __launch_bounds__(512) weighted_sum_kernel(const __grid_constant__ AggArgs a, void* dst,
                                                           float* acc_f32, int accumulate, float post_scale,
                                                           size_t numel) {
  const size_t units = numel / kUnit;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t u = (size_t)blockIdx.x * blockDim.x + threadIdx.x; u < units; u += stride) {
    float acc[kUnit];
    // the accumulator and dst are rewritten below, so they are read with plain loads (PEER = true), never .nc
    if (acc_f32 && accumulate) {
      Unit<MFL_FLOAT32>::load<true>(acc_f32, u, acc);
    } else if (!acc_f32 && accumulate) {
      Unit<DST>::template load<true>(dst, u, acc);
    } else {
#pragma unroll
      for (int i = 0; i < kUnit; ++i) acc[i] = 0.f;
    }
    reduce_unit<SRC, false>(a, 0, u, acc);
    if (acc_f32) Unit<MFL_FLOAT32>::store(acc_f32, u, acc);
    if (dst) {
#pragma unroll
      for (int i = 0; i < kUnit; ++i) acc[i] *= post_scale;
      Unit<DST>::store(dst, u, acc);
    }
  }
  // scalar tail (< 8 elements), handled by the first few threads of block 0
  if (blockIdx.x == 0 && threadIdx.x < numel % kUnit) {
    const size_t i = units * kUnit + threadIdx.x;
    float acc = 0.f;
    if (acc_f32 && accumulate) acc = acc_f32[i];
    else if (!acc_f32 && accumulate) acc = load_as_float(dst, DST, i);
    for (int l = 0; l < a.n; ++l) acc = fmaf(a.w[l], load_as_float(a.src[l], SRC, i), acc);
    if (acc_f32) acc_f32[i] = acc;
    if (dst) store_from_float(dst, DST, i, acc * post_scale);
  }
}

// Reference-exact path (any dtype): per-term truncation to T, accumulation in T.
template <typename T>
__global__ void __launch_bounds__(256) weighted_sum_exact_kernel(const __grid_constant__ AggArgs a, T* dst,
                                                                 int accumulate, size_t numel) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel; i += stride) {
    T acc = accumulate ? dst[i] : from_double<T>(0.0);
    for (int l = 0; l < a.n; ++l) {
      T term = from_double<T>(to_double(reinterpret_cast<const T*>(a.src[l])[i]) * a.wd[l]);
      if constexpr (sizeof(T) == 2 && !std::is_integral<T>::value)
        acc = from_double<T>(to_double(acc) + to_double(term));
      else
        acc = (T)(acc + term);
    }
    dst[i] = acc;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) rolling_update_kernel(T* community, float* S, const T* x_new, float w_new,
                                                            const T* x_old, float w_old, float inv_z, size_t numel) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel; i += stride) {
    float s = S[i];
    s = fmaf(w_new, (float)to_double(x_new[i]), s);
    if (x_old) s = fmaf(-w_old, (float)to_double(x_old[i]), s);
    S[i] = s;
    if (community) community[i] = from_double<T>((double)(s * inv_z));
  }
}

// ------------------------------------------------------------------------------------------
// Fused multi-GPU aggregation.
// ------------------------------------------------------------------------------------------
struct FusedArgs {
  AggArgs agg;
  void* out[MFL_MAX_PEERS];
  uint32_t* pads[MFL_MAX_PEERS];
  mfl_segment seg[MFL_MAX_SEGMENTS];
  int n_seg, rank, world;
  float* install_f32;
  // NVLS mode (n_mc > 0): mc_src[j] / mc_out are MULTICAST addresses of symmetric buffers.  A multimem.ld_reduce on
  // mc_src[j] + x returns the sum over all GPUs of the value at offset x of slot j, added inside the NVSwitch; a
  // multimem.st on mc_out + x lands in every GPU's community buffer.  Per GPU and per byte of model that is one
  // pull of 1/world and one push of 1/world over NVLink instead of (world-1)/world each way with peer loads/stores.
  // The switch only adds, so the slots must hold pre-scaled models (K2 publishes w_l * theta_l).
  const void* mc_src[MFL_MAX_MC_SRC];
  void* mc_out;
  int n_mc;
};

__device__ __forceinline__ uint4 multimem_ld_reduce_bf16(const void* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 multimem_ld_reduce_f16(const void* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 multimem_ld_reduce_f32(const void* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(void* p, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// NVLS variant of fused_segment_fast: my shard of the segment, reduced by the switch, multicast back.
template <int DT>
__device__ __forceinline__ void fused_segment_nvls(const FusedArgs& fa, const mfl_segment& sg) {
  constexpr int kVecPerUnit = (DT == MFL_FLOAT32) ? 2 : 1;  // a unit is 8 elements
  const size_t units = sg.numel / kUnit;
  const size_t upr = (units + fa.world - 1) / fa.world;
  const size_t begin = min(units, upr * fa.rank), end = min(units, begin + upr);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  char* out = reinterpret_cast<char*>(fa.mc_out) + sg.offset_bytes;
  auto ld = [](const char* p) {
    return DT == MFL_FLOAT32 ? multimem_ld_reduce_f32(p)
                             : (DT == MFL_BFLOAT16 ? multimem_ld_reduce_bf16(p) : multimem_ld_reduce_f16(p));
  };
  // 16-byte vectors [vbegin, vend) of my shard; kNvlsBatch switch reductions in flight per thread
  constexpr int kNvlsBatch = 8;
  const size_t vbegin = begin * kVecPerUnit, vend = end * kVecPerUnit;
  const size_t first = vbegin + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (fa.n_mc == 1) {
    const char* src = reinterpret_cast<const char*>(fa.mc_src[0]) + sg.offset_bytes;
    for (size_t v = first; v < vend; v += stride * kNvlsBatch) {
      uint4 r[kNvlsBatch];
#pragma unroll
      for (int b = 0; b < kNvlsBatch; ++b) {
        const size_t idx = v + b * stride;
        if (idx < vend) r[b] = ld(src + idx * sizeof(uint4));
      }
#pragma unroll
      for (int b = 0; b < kNvlsBatch; ++b) {
        const size_t idx = v + b * stride;
        if (idx < vend) multimem_st(out + idx * sizeof(uint4), r[b]);
      }
    }
  } else {  // several learners per GPU: one switch reduction per local slot index, folded in fp32 here
    constexpr int N = VecTraits<DT>::N;
    for (size_t v = first; v < vend; v += stride) {
      float acc[N], f[N];
#pragma unroll
      for (int i = 0; i < N; ++i) acc[i] = 0.f;
      for (int jj = 0; jj < fa.n_mc; ++jj) {
        VecTraits<DT>::unpack(ld(reinterpret_cast<const char*>(fa.mc_src[jj]) + sg.offset_bytes + v * sizeof(uint4)), f);
#pragma unroll
        for (int i = 0; i < N; ++i) acc[i] += f[i];
      }
      multimem_st(out + v * sizeof(uint4), VecTraits<DT>::pack(acc));
    }
  }
  // tail elements (numel % 8) go through the peer-pointer path on the last rank
  if (fa.rank == fa.world - 1 && blockIdx.x == 0 && threadIdx.x < sg.numel % kUnit) {
    const size_t i = units * kUnit + threadIdx.x;
    float acc = 0.f;
    for (int l = 0; l < fa.agg.n; ++l)
      acc += load_as_float(reinterpret_cast<const char*>(fa.agg.src[l]) + sg.offset_bytes, DT, i);
    for (int t = 0; t < fa.world; ++t)
      store_from_float(reinterpret_cast<char*>(fa.out[t]) + sg.offset_bytes, DT, i, acc);
  }
}

template <int DT>
__device__ __forceinline__ void fused_segment_fast(const FusedArgs& fa, const mfl_segment& sg) {
  const size_t units = sg.numel / kUnit;
  const size_t upr = (units + fa.world - 1) / fa.world;
  const size_t begin = min(units, upr * fa.rank), end = min(units, begin + upr);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t u = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; u < end; u += stride) {
    float acc[kUnit];
#pragma unroll
    for (int i = 0; i < kUnit; ++i) acc[i] = 0.f;
    reduce_unit<DT, true>(fa.agg, sg.offset_bytes, u, acc);
    // pack once, store to every rank (own buffer included): all-gather by peer stores
    Raw<DT> r;
    if constexpr (DT == MFL_FLOAT32) {
      r.v[0] = VecTraits<MFL_FLOAT32>::pack(acc);
      r.v[1] = VecTraits<MFL_FLOAT32>::pack(acc + 4);
    } else {
      r.v[0] = VecTraits<DT>::pack(acc);
    }
    for (int t = 0; t < fa.world; ++t) {
      const int peer = (fa.rank + t) % fa.world;  // stagger destinations across ranks
      char* o = reinterpret_cast<char*>(fa.out[peer]) + sg.offset_bytes;
      if constexpr (DT == MFL_FLOAT32) {
        st_stream(reinterpret_cast<uint4*>(o) + 2 * u, r.v[0]);
        st_stream(reinterpret_cast<uint4*>(o) + 2 * u + 1, r.v[1]);
      } else {
        st_stream(reinterpret_cast<uint4*>(o) + u, r.v[0]);
      }
    }
  }
  // tail elements belong to the last rank
  if (fa.rank == fa.world - 1 && blockIdx.x == 0 && threadIdx.x < sg.numel % kUnit) {
    const size_t i = units * kUnit + threadIdx.x;
    float acc = 0.f;
    for (int l = 0; l < fa.agg.n; ++l)
      acc = fmaf(fa.agg.w[l], load_as_float(reinterpret_cast<const char*>(fa.agg.src[l]) + sg.offset_bytes, DT, i), acc);
    for (int t = 0; t < fa.world; ++t)
      store_from_float(reinterpret_cast<char*>(fa.out[t]) + sg.offset_bytes, DT, i, acc);
  }
}

template <typename T>
__device__ __noinline__ void fused_segment_exact(const FusedArgs& fa, const mfl_segment& sg) {
  const size_t per = (sg.numel + fa.world - 1) / fa.world;
  const size_t begin = min((size_t)sg.numel, per * fa.rank), end = min((size_t)sg.numel, begin + per);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride) {
    T acc = from_double<T>(0.0);
    for (int l = 0; l < fa.agg.n; ++l) {
      const T x = reinterpret_cast<const T*>(reinterpret_cast<const char*>(fa.agg.src[l]) + sg.offset_bytes)[i];
      T term = from_double<T>(to_double(x) * fa.agg.wd[l]);
      if constexpr (sizeof(T) == 2 && !std::is_integral<T>::value)
        acc = from_double<T>(to_double(acc) + to_double(term));
      else
        acc = (T)(acc + term);
    }
    for (int t = 0; t < fa.world; ++t)
      reinterpret_cast<T*>(reinterpret_cast<char*>(fa.out[t]) + sg.offset_bytes)[i] = acc;
  }
}

__global__ void __launch_bounds__(512, 2) fused_aggregate_kernel(const __grid_constant__ FusedArgs fa) {
  const int slot0 = (0 * gridDim.x + blockIdx.x) * MFL_MAX_PEERS;
  const int slot1 = (1 * gridDim.x + blockIdx.x) * MFL_MAX_PEERS;
  // Phase 0: every rank's slots are final (they were written by earlier work on each rank's stream).
  peer_barrier(fa.pads, fa.rank, fa.world, slot0, /*have_prior_writes=*/true);
  // Phase 1: reduce my shard of every segment and scatter it to all ranks.
  for (int s = 0; s < fa.n_seg; ++s) {
    const mfl_segment& sg = fa.seg[s];
    if (fa.n_mc > 0 && !sg.exact && (sg.dtype == MFL_BFLOAT16 || sg.dtype == MFL_FLOAT32 || sg.dtype == MFL_FLOAT16)) {
      if (sg.dtype == MFL_BFLOAT16) fused_segment_nvls<MFL_BFLOAT16>(fa, sg);
      else if (sg.dtype == MFL_FLOAT32) fused_segment_nvls<MFL_FLOAT32>(fa, sg);
      else fused_segment_nvls<MFL_FLOAT16>(fa, sg);
      continue;
    }
    if (!sg.exact && sg.dtype == MFL_BFLOAT16) fused_segment_fast<MFL_BFLOAT16>(fa, sg);
    else if (!sg.exact && sg.dtype == MFL_FLOAT32) fused_segment_fast<MFL_FLOAT32>(fa, sg);
    else if (!sg.exact && sg.dtype == MFL_FLOAT16) fused_segment_fast<MFL_FLOAT16>(fa, sg);
    else {
      MFL_DISPATCH_ALL(sg.dtype, T, fused_segment_exact<T>(fa, sg));
    }
  }
  // Phase 2: all shards have landed everywhere.
  peer_barrier(fa.pads, fa.rank, fa.world, slot1, /*have_prior_writes=*/true);
  // Phase 3 (optional): install the new community parameters as this rank's fp32 master copy.
  // The flag barriers pair block b here with block b of every peer and with nothing else, so block b may only read
  // what block b of some rank wrote in phase 1.  The loop therefore walks the shards rank by rank with the writer's
  // own index mapping (shard begin + block * blockDim + thread + k * grid stride) instead of striding over the whole
  // segment, whose units belong to blocks this one never synchronised with.
  if (fa.install_f32) {
    const mfl_segment& sg = fa.seg[0];
    const char* c = reinterpret_cast<const char*>(fa.out[fa.rank]) + sg.offset_bytes;
    const size_t units = sg.numel / kUnit;
    const size_t upr = (units + fa.world - 1) / fa.world;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t lane = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool vec_mapped = fa.n_mc > 0 && sg.dtype == MFL_FLOAT32;  // the NVLS writer strides 16-byte vectors
    for (int r = 0; r < fa.world; ++r) {
      const size_t begin = min(units, upr * r), end = min(units, begin + upr);
      if (vec_mapped) {
        for (size_t v = 2 * begin + lane; v < 2 * end; v += stride)
          st_stream(reinterpret_cast<uint4*>(fa.install_f32) + v, ld_peer(reinterpret_cast<const uint4*>(c) + v));
        continue;
      }
      for (size_t u = begin + lane; u < end; u += stride) {
        float f[kUnit];
        if (sg.dtype == MFL_BFLOAT16) VecTraits<MFL_BFLOAT16>::unpack(ld_peer(reinterpret_cast<const uint4*>(c) + u), f);
        else if (sg.dtype == MFL_FLOAT16) VecTraits<MFL_FLOAT16>::unpack(ld_peer(reinterpret_cast<const uint4*>(c) + u), f);
        else {
          VecTraits<MFL_FLOAT32>::unpack(ld_peer(reinterpret_cast<const uint4*>(c) + 2 * u), f);
          VecTraits<MFL_FLOAT32>::unpack(ld_peer(reinterpret_cast<const uint4*>(c) + 2 * u + 1), f + 4);
        }
        Unit<MFL_FLOAT32>::store(fa.install_f32, u, f);
      }
    }
    if (blockIdx.x == 0 && threadIdx.x < sg.numel % kUnit) {
      const size_t i = units * kUnit + threadIdx.x;
      fa.install_f32[i] = load_as_float(c, sg.dtype, i);
    }
  }
}

__global__ void barrier_kernel(const __grid_constant__ FusedArgs fa, int channel) {
  peer_barrier(fa.pads, fa.rank, fa.world, (channel * gridDim.x + blockIdx.x) * MFL_MAX_PEERS, true);
}

static bool is_float_fast(int dt) { return dt == MFL_FLOAT32 || dt == MFL_BFLOAT16 || dt == MFL_FLOAT16; }
static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int SRC>
static int launch_ws(int dst_dtype, const AggArgs& a, void* dst, float* acc, int accumulate, float post,
                     size_t numel, cudaStream_t st) {
  const size_t units = numel / kUnit;
  int blocks = (int)std::min<size_t>((units + 511) / 512, (size_t)sm_count() * 2);
  if (blocks < 1) blocks = 1;
  switch (dst_dtype) {
    case MFL_FLOAT32: weighted_sum_kernel<SRC, MFL_FLOAT32><<<blocks, 512, 0, st>>>(a, dst, acc, accumulate, post, numel); break;
    case MFL_BFLOAT16: weighted_sum_kernel<SRC, MFL_BFLOAT16><<<blocks, 512, 0, st>>>(a, dst, acc, accumulate, post, numel); break;
    case MFL_FLOAT16: weighted_sum_kernel<SRC, MFL_FLOAT16><<<blocks, 512, 0, st>>>(a, dst, acc, accumulate, post, numel); break;
    default: set_error("weighted_sum: unsupported dst dtype %d on the fast path", dst_dtype); return -1;
  }
  return check(cudaGetLastError(), "weighted_sum_kernel launch");
}

}  // namespace mfl

using namespace mfl;

MFL_SPIN_TIMEOUT_READER(mfl_spin_timeouts_aggregate)

extern "C" int mfl_weighted_sum(void* dst, int dst_dtype, const void* const* srcs, const double* weights,
                                int n_src, int src_dtype, size_t numel, float* acc_f32, int accumulate,
                                double post_scale, int exact, void* stream) {
  MFL_ARG(n_src >= 0 && n_src <= MFL_MAX_SRC, "weighted_sum: n_src=%d outside [0,%d]", n_src, MFL_MAX_SRC);
  MFL_ARG(dst || acc_f32, "weighted_sum: need dst or acc_f32");
  MFL_ARG(dtype_size(src_dtype) && (!dst || dtype_size(dst_dtype)), "weighted_sum: bad dtype");
  if (numel == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  AggArgs a;
  a.n = n_src;
  bool al = aligned16(dst) && aligned16(acc_f32);
  for (int l = 0; l < n_src; ++l) {
    a.src[l] = srcs[l];
    a.wd[l] = weights[l];
    a.w[l] = (float)weights[l];
    al = al && aligned16(srcs[l]);
  }
  if (!exact && is_float_fast(src_dtype) && (!dst || is_float_fast(dst_dtype)) && al) {
    const int dd = dst ? dst_dtype : MFL_FLOAT32;
    switch (src_dtype) {
      case MFL_FLOAT32: return launch_ws<MFL_FLOAT32>(dd, a, dst, acc_f32, accumulate, (float)post_scale, numel, st);
      case MFL_BFLOAT16: return launch_ws<MFL_BFLOAT16>(dd, a, dst, acc_f32, accumulate, (float)post_scale, numel, st);
      default: return launch_ws<MFL_FLOAT16>(dd, a, dst, acc_f32, accumulate, (float)post_scale, numel, st);
    }
  }
  MFL_ARG(dst && dst_dtype == src_dtype && !acc_f32 && post_scale == 1.0,
          "weighted_sum: exact/generic path needs dst of the source dtype, no fp32 accumulator, post_scale 1 "
          "(src=%d dst=%d aligned=%d)", src_dtype, dst_dtype, (int)al);
  int blocks = (int)std::min<size_t>((numel + 255) / 256, (size_t)sm_count() * 8);
  MFL_DISPATCH_ALL(src_dtype, T,
                   (weighted_sum_exact_kernel<T><<<blocks, 256, 0, st>>>(a, (T*)dst, accumulate, numel)));
  return check(cudaGetLastError(), "weighted_sum_exact_kernel launch");
}

extern "C" int mfl_rolling_update(void* community, int dtype, float* scaled_sum, const void* x_new, double w_new,
                                  const void* x_old, double w_old, double z, size_t numel, void* stream) {
  MFL_ARG(scaled_sum && x_new, "rolling_update: scaled_sum and x_new are required");
  MFL_ARG(z != 0.0 || !community, "rolling_update: z must be non-zero when a community model is written");
  if (numel == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  // float-family models with a community output take the vectorised fold kernel (16-byte accesses, all loads of a
  // unit in flight together); everything else keeps the element-wise kernel below
  if (community && is_float_fast(dtype) && aligned16(community) && aligned16(scaled_sum) && aligned16(x_new) &&
      aligned16(x_old)) {
    mfl_segment sg{0, numel, dtype, 0};
    const uint64_t off = 0;
    void* outs[1] = {community};
    const int blocks = (int)std::min<size_t>((numel / 8 + 255) / 256 + 1, (size_t)sm_count() * 4);
    return mfl_async_fold(x_new, w_new, x_old, w_old, 1.0, 1.0 / z, scaled_sum, &off, &sg, 1, outs, nullptr, 1,
                          nullptr, 0, nullptr, 0, nullptr, blocks, stream);
  }
  int blocks = (int)std::min<size_t>((numel + 255) / 256, (size_t)sm_count() * 8);
  const float inv_z = community ? (float)(1.0 / z) : 0.f;
  MFL_DISPATCH_ALL(dtype, T,
                   (rolling_update_kernel<T><<<blocks, 256, 0, st>>>((T*)community, scaled_sum, (const T*)x_new,
                                                                     (float)w_new, (const T*)x_old, (float)w_old,
                                                                     inv_z, numel)));
  return check(cudaGetLastError(), "rolling_update_kernel launch");
}

extern "C" size_t mfl_signal_pad_bytes(int n_blocks) {
  return (size_t)2 * n_blocks * MFL_MAX_PEERS * sizeof(uint32_t);
}

static int fill_fused(FusedArgs& fa, uint32_t* const* pads, int rank, int world) {
  MFL_ARG(world >= 1 && world <= MFL_MAX_PEERS && rank >= 0 && rank < world, "fused: bad rank/world %d/%d", rank, world);
  fa.rank = rank;
  fa.world = world;
  for (int t = 0; t < world; ++t) fa.pads[t] = pads[t];
  return 0;
}

static int fused_aggregate_impl(const void* const* src_slots, const double* weights, int n_src,
                                void* const* out_bufs, uint32_t* const* signal_pads, int rank, int world,
                                const mfl_segment* segments, int n_segments, float* install_f32, int n_blocks,
                                void* stream, const void* const* mc_src, int n_mc, void* mc_out) {
  MFL_ARG(n_src >= 1 && n_src <= MFL_MAX_SRC, "fused_aggregate: n_src=%d outside [1,%d]", n_src, MFL_MAX_SRC);
  MFL_ARG(n_segments >= 1 && n_segments <= MFL_MAX_SEGMENTS, "fused_aggregate: n_segments=%d", n_segments);
  MFL_ARG(n_blocks >= 1 && n_blocks <= 2 * sm_count(), "fused_aggregate: n_blocks=%d must be in [1, 2 x #SMs] "
          "(blocks spin on their peers, so all of them must be co-resident: 2 CTAs of 512 threads per SM)", n_blocks);
  FusedArgs local;
  if (int rc = fill_fused(local, signal_pads, rank, world)) return rc;
  local.agg.n = n_src;
  for (int l = 0; l < n_src; ++l) {
    local.agg.src[l] = src_slots[l];
    local.agg.wd[l] = weights[l];
    local.agg.w[l] = (float)weights[l];
    MFL_ARG(aligned16(src_slots[l]), "fused_aggregate: slot %d is not 16-byte aligned", l);
  }
  for (int t = 0; t < world; ++t) {
    local.out[t] = out_bufs[t];
    MFL_ARG(aligned16(out_bufs[t]), "fused_aggregate: output %d is not 16-byte aligned", t);
  }
  local.n_seg = n_segments;
  for (int s = 0; s < n_segments; ++s) {
    local.seg[s] = segments[s];
    MFL_ARG(segments[s].offset_bytes % 16 == 0 && dtype_size(segments[s].dtype),
            "fused_aggregate: segment %d misaligned or bad dtype", s);
  }
  local.install_f32 = install_f32;
  MFL_ARG(!install_f32 || (!segments[0].exact && is_float_fast(segments[0].dtype) && aligned16(install_f32)),
          "fused_aggregate: the fp32 install needs segment 0 to be a float-family, non-exact segment and a 16-byte "
          "aligned destination");
  MFL_ARG(n_mc >= 0 && n_mc <= MFL_MAX_MC_SRC, "fused_aggregate: n_mc=%d outside [0,%d]", n_mc, MFL_MAX_MC_SRC);
  MFL_ARG(n_mc == 0 || (mc_out && world > 1), "fused_aggregate: NVLS mode needs a multicast output and world > 1");
  local.n_mc = n_mc;
  local.mc_out = mc_out;
  for (int j = 0; j < n_mc; ++j) {
    local.mc_src[j] = mc_src[j];
    MFL_ARG(mc_src[j] && aligned16(mc_src[j]), "fused_aggregate: multicast source %d null or misaligned", j);
  }
  fused_aggregate_kernel<<<n_blocks, 512, 0, (cudaStream_t)stream>>>(local);
  return check(cudaGetLastError(), "fused_aggregate_kernel launch");
}

extern "C" int mfl_fused_aggregate(const void* const* src_slots, const double* weights, int n_src,
                                   void* const* out_bufs, uint32_t* const* signal_pads, int rank, int world,
                                   const mfl_segment* segments, int n_segments, float* install_f32,
                                   int n_blocks, void* stream) {
  return fused_aggregate_impl(src_slots, weights, n_src, out_bufs, signal_pads, rank, world, segments, n_segments,
                              install_f32, n_blocks, stream, nullptr, 0, nullptr);
}

extern "C" int mfl_fused_aggregate_nvls(const void* const* src_slots, const double* weights, int n_src,
                                        void* const* out_bufs, uint32_t* const* signal_pads, int rank, int world,
                                        const mfl_segment* segments, int n_segments, float* install_f32,
                                        int n_blocks, void* stream, const void* const* mc_src_slots, int n_mc,
                                        void* mc_out) {
  MFL_ARG(n_mc >= 1, "fused_aggregate_nvls: need at least one multicast source");
  return fused_aggregate_impl(src_slots, weights, n_src, out_bufs, signal_pads, rank, world, segments, n_segments,
                              install_f32, n_blocks, stream, mc_src_slots, n_mc, mc_out);
}

extern "C" int mfl_fabric_barrier(uint32_t* const* signal_pads, int rank, int world, int channel, void* stream) {
  FusedArgs local;
  if (int rc = fill_fused(local, signal_pads, rank, world)) return rc;
  MFL_ARG(channel == 0 || channel == 1, "fabric_barrier: channel must be 0 or 1");
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(local, channel);
  return check(cudaGetLastError(), "barrier_kernel launch");
}
