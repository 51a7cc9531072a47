// Shared device/host helpers for the metisfl_b200 kernel library (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "metisfl_b200.h"

namespace mfl {

void set_error(const char* fmt, ...);
int check(cudaError_t e, const char* what);

#define MFL_CUDA(call)                                   \
  do {                                                   \
    int _rc = ::mfl::check((call), #call);               \
    if (_rc) return _rc;                                 \
  } while (0)
#define MFL_ARG(cond, ...)                               \
  do {                                                   \
    if (!(cond)) {                                       \
      ::mfl::set_error(__VA_ARGS__);                     \
      return -1;                                         \
    }                                                    \
  } while (0)

// Bind the current device's primary context to the calling thread (legal during stream capture, unlike cudaFree(0)).
inline void bind_context() {
  int dev = 0;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaSetDevice(dev);
}

constexpr int kSMs = 148;

inline int sm_count() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = kSMs;
  }
  return n;
}

__host__ __device__ inline int dtype_size(int dt) {
  switch (dt) {
    case MFL_INT8: case MFL_UINT8: return 1;
    case MFL_INT16: case MFL_UINT16: case MFL_FLOAT16: case MFL_BFLOAT16: return 2;
    case MFL_INT32: case MFL_UINT32: case MFL_FLOAT32: return 4;
    case MFL_INT64: case MFL_UINT64: case MFL_FLOAT64: return 8;
  }
  return 0;
}

// ---- 16-byte streaming accesses ------------------------------------------------------------
// Peer (NVLink) and once-read HBM data must not pollute L1: nc + no_allocate for loads of
// immutable sources, plain .cg-like volatile paths for buffers another GPU writes.
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
// Load of data that another GPU (or an earlier phase of this kernel on another GPU) wrote: a plain
// weak load, ordered by the acquire side of the flag barrier that precedes it, bypassing L1
// allocation.  (.nc would be wrong here - the data is not read-only for the kernel's lifetime - and a
// .relaxed.sys load measured ~25% slower for no additional guarantee we need.)
__device__ __forceinline__ uint4 ld_peer(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- element <-> fp32 conversion of one 16-byte vector -------------------------------------
template <int DT> struct VecTraits;  // elements per 16 B and unpack/pack to float[]

template <> struct VecTraits<MFL_FLOAT32> {
  static constexpr int N = 4;
  __device__ static void unpack(const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y);
    f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
  }
  __device__ static uint4 pack(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                      __float_as_uint(f[3]));
  }
};
template <> struct VecTraits<MFL_BFLOAT16> {
  static constexpr int N = 8;
  __device__ static void unpack(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {  // bf16 -> fp32 is a 16-bit shift
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <> struct VecTraits<MFL_FLOAT16> {
  static constexpr int N = 8;
  __device__ static void unpack(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
  }
  __device__ static uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// scalar element access used by tails and the exact (reference-semantics) paths
template <typename T> __device__ __forceinline__ double to_double(T x) { return (double)x; }
template <> __device__ __forceinline__ double to_double(__nv_bfloat16 x) { return (double)__bfloat162float(x); }
template <> __device__ __forceinline__ double to_double(__half x) { return (double)__half2float(x); }
template <typename T> __device__ __forceinline__ T from_double(double x) { return (T)x; }
template <> __device__ __forceinline__ __nv_bfloat16 from_double(double x) { return __float2bfloat16_rn((float)x); }
template <> __device__ __forceinline__ __half from_double(double x) { return __float2half_rn((float)x); }

__device__ __forceinline__ float load_as_float(const void* p, int dt, size_t i) {
  switch (dt) {
    case MFL_FLOAT32: return ((const float*)p)[i];
    case MFL_BFLOAT16: return __bfloat162float(((const __nv_bfloat16*)p)[i]);
    case MFL_FLOAT16: return __half2float(((const __half*)p)[i]);
    case MFL_FLOAT64: return (float)((const double*)p)[i];
  }
  return 0.f;
}
__device__ __forceinline__ void store_from_float(void* p, int dt, size_t i, float v) {
  switch (dt) {
    case MFL_FLOAT32: ((float*)p)[i] = v; break;
    case MFL_BFLOAT16: ((__nv_bfloat16*)p)[i] = __float2bfloat16_rn(v); break;
    case MFL_FLOAT16: ((__half*)p)[i] = __float2half_rn(v); break;
    case MFL_FLOAT64: ((double*)p)[i] = (double)v; break;
  }
}

// Dispatch a generic lambda on a C++ element type for the wire dtype code.
#define MFL_DISPATCH_ALL(dt, T, ...)                                         \
  switch (dt) {                                                              \
    case MFL_INT8: { using T = int8_t; __VA_ARGS__; break; }                 \
    case MFL_INT16: { using T = int16_t; __VA_ARGS__; break; }               \
    case MFL_INT32: { using T = int32_t; __VA_ARGS__; break; }               \
    case MFL_INT64: { using T = int64_t; __VA_ARGS__; break; }               \
    case MFL_UINT8: { using T = uint8_t; __VA_ARGS__; break; }               \
    case MFL_UINT16: { using T = uint16_t; __VA_ARGS__; break; }             \
    case MFL_UINT32: { using T = uint32_t; __VA_ARGS__; break; }             \
    case MFL_UINT64: { using T = uint64_t; __VA_ARGS__; break; }             \
    case MFL_FLOAT32: { using T = float; __VA_ARGS__; break; }               \
    case MFL_FLOAT64: { using T = double; __VA_ARGS__; break; }              \
    case MFL_FLOAT16: { using T = __half; __VA_ARGS__; break; }              \
    case MFL_BFLOAT16: { using T = __nv_bfloat16; __VA_ARGS__; break; }      \
    default: break;                                                          \
  }

// ---- system-scope flag primitives (K13) ----------------------------------------------------
// put: flip the peer's flag 0 -> 1 (release); wait: flip our own flag 1 -> 0 (acquire).
// Self-resetting, so no epoch bookkeeping and the pads can be reused by back-to-back launches.
//
// Every spin is bounded: a rank that never shows up (dead process, a launch the host never made) must not wedge the
// GPU.  After kSpinTimeoutNs the waiter gives up, counts the event in g_spin_timeouts (read it with
// mfl_spin_timeouts()) and carries on; the result of that launch is garbage and the fabric has to be rebuilt, but
// the device stays usable.
constexpr unsigned long long kSpinTimeoutNs = 20ull * 1000 * 1000 * 1000;
// one counter per translation unit (the library is built without relocatable device code); a unit that spins
// exports a reader with MFL_SPIN_TIMEOUT_READER and mfl_spin_timeouts() in runtime.cu adds them up
static __device__ unsigned int g_spin_timeouts = 0;
#define MFL_SPIN_TIMEOUT_READER(name)                                                          \
  extern "C" int name(void) {                                                                  \
    unsigned int n = 0;                                                                        \
    if (cudaMemcpyFromSymbol(&n, ::mfl::g_spin_timeouts, sizeof(n)) != cudaSuccess) return -1; \
    return (int)n;                                                                             \
  }

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
struct SpinGuard {
  unsigned long long t0 = 0;
  unsigned int n = 0;
  __device__ __forceinline__ bool expired() {  // true once the budget is used up
    if ((++n & 0x3ff) != 0) return false;
    const unsigned long long now = global_ns();
    if (t0 == 0) { t0 = now; return false; }
    if (now - t0 < kSpinTimeoutNs) return false;
    atomicAdd(&g_spin_timeouts, 1u);
    return true;
  }
};
__device__ __forceinline__ void flag_put(uint32_t* addr) {
  uint32_t old;
  SpinGuard g;
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
  } while (old != 0u && !g.expired());
}
__device__ __forceinline__ void flag_wait(uint32_t* addr) {
  uint32_t old;
  SpinGuard g;
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
  } while (old != 1u && !g.expired());
}
// Monotonic 64-bit counters (version numbers, per-learner "task n is published"): one writer, any readers.
__device__ __forceinline__ void counter_publish(unsigned long long* addr, unsigned long long v) {
  asm volatile("st.global.release.sys.u64 [%0], %1;" :: "l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long counter_read(const unsigned long long* addr) {
  unsigned long long v;
  asm volatile("ld.global.acquire.sys.u64 %0, [%1];" : "=l"(v) : "l"(addr) : "memory");
  return v;
}
// returns false on timeout
__device__ __forceinline__ bool counter_wait_ge(const unsigned long long* addr, unsigned long long v) {
  SpinGuard g;
  while (counter_read(addr) < v)
    if (g.expired()) return false;
  return true;
}

// Block-level barrier with the same block index on every peer.  pads[t] is rank t's pad
// (peer-mapped); slot layout: pad[(channel * n_blocks + block) * MFL_MAX_PEERS + writer_rank].
__device__ __forceinline__ void peer_barrier(uint32_t* const* pads, int rank, int world, int slot_base,
                                             bool have_prior_writes) {
  if (have_prior_writes) __syncthreads();
  if (threadIdx.x < (unsigned)world) {
    const int t = threadIdx.x;
    if (have_prior_writes) __threadfence_system();
    flag_put(pads[t] + slot_base + rank);
    flag_wait(pads[rank] + slot_base + t);
  }
  __syncthreads();
}

}  // namespace mfl
