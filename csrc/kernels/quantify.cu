// K4: tensor quantifier - zeros / non-zeros per model variable in one pass over the flat arena.
//
// Behavioural spec: reference QuantifyTensor (controller/common/proto_tensor_serde.h:35-50) counts
// elements that compare equal to 0 per variable; the controller records the result for every
// variable of each new community model (controller.cc:952-1004).
//
// B200 design: the arena is split into equal contiguous element ranges, one per CTA, independent
// of variable boundaries, so a 100-element BatchNorm vector and a 50 M-element embedding both keep
// every SM busy.  A CTA writes the counts of variables that lie entirely inside its range
// directly; a variable cut by a range boundary yields at most two partial counts per CTA, which go
// to caller-provided scratch and are folded by a short in-order scan (no float, no atomics, so the
// result is deterministic).
#include <algorithm>
#include <mutex>

#include "common.cuh"

namespace mfl {

constexpr int kQBlocks = 2 * kSMs;  // scratch is sized for this many CTAs
constexpr int kQThreads = 256;

struct QPartial {          // one per CTA
  long long var_head;      // variable that began before this CTA's range (or -1)
  long long var_tail;      // variable that continues past this CTA's range (or -1)
  unsigned long long nz_head, nz_tail;
};

// non-zero lanes of one 16-byte vector holding elements of type U (sign bit masked for floats)
template <typename U>
__device__ __forceinline__ unsigned nz_in_vec(const uint4& v, U mask) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  unsigned c = 0;
  if constexpr (sizeof(U) == 8) {
    c += (((((unsigned long long)w[1] << 32) | w[0]) & mask) != 0);
    c += (((((unsigned long long)w[3] << 32) | w[2]) & mask) != 0);
  } else if constexpr (sizeof(U) == 4) {
#pragma unroll
    for (int i = 0; i < 4; ++i) c += ((w[i] & mask) != 0);
  } else if constexpr (sizeof(U) == 2) {
#pragma unroll
    for (int i = 0; i < 4; ++i) c += ((w[i] & mask) != 0) + (((w[i] >> 16) & mask) != 0);
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int b = 0; b < 4; ++b) c += (((w[i] >> (8 * b)) & mask) != 0);
  }
  return c;
}

// CTA-wide count over [lo, hi): scalar head up to a 16-byte boundary, 16-byte vector body, scalar tail
template <typename U>
__device__ __forceinline__ unsigned count_nonzero(const U* p, size_t lo, size_t hi, U mask) {
  constexpr size_t kPer = 16 / sizeof(U);
  unsigned c = 0;
  const uintptr_t addr = reinterpret_cast<uintptr_t>(p + lo);
  size_t head = ((16 - (addr & 15)) & 15) / sizeof(U);
  if (head > hi - lo) head = hi - lo;
  const size_t body_lo = lo + head, n_vec = (hi - body_lo) / kPer, tail_lo = body_lo + n_vec * kPer;
  if (threadIdx.x < head) c += ((p[lo + threadIdx.x] & mask) != 0);
  const uint4* v = reinterpret_cast<const uint4*>(p + body_lo);
  for (size_t i = threadIdx.x; i < n_vec; i += kQThreads) c += nz_in_vec<U>(ld_stream(v + i), mask);
  if (threadIdx.x < hi - tail_lo) c += ((p[tail_lo + threadIdx.x] & mask) != 0);
  return c;
}

__device__ __forceinline__ unsigned long long block_sum(unsigned c, unsigned long long* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  __syncthreads();  // smem reuse across calls
  if ((threadIdx.x & 31) == 0) smem[threadIdx.x >> 5] = c;
  __syncthreads();
  unsigned long long t = 0;
  for (int w = 0; w < kQThreads / 32; ++w) t += smem[w];
  return t;
}

template <typename U>
__global__ void __launch_bounds__(kQThreads) quantify_kernel(const U* data, U mask, const uint64_t* offsets,
                                                             int n_vars, unsigned long long* counts,
                                                             QPartial* partials) {
  __shared__ unsigned long long red[kQThreads / 32];
  const size_t total = offsets[n_vars];
  const size_t per = (total + gridDim.x - 1) / gridDim.x;
  const size_t lo = min(total, per * blockIdx.x), hi = min(total, lo + per);
  QPartial part{-1, -1, 0ull, 0ull};
  if (lo < hi) {
    // first variable intersecting [lo, hi): largest v with offsets[v] <= lo
    int a = 0, b = n_vars;
    while (b - a > 1) {
      const int m = (a + b) >> 1;
      if (offsets[m] <= lo) a = m; else b = m;
    }
    for (int v = a; v < n_vars && offsets[v] < hi; ++v) {
      const size_t vs = offsets[v], ve = offsets[v + 1];
      if (ve == vs) continue;  // empty variables are reported by the fold kernel
      const size_t s = max(vs, lo), e = min(ve, hi);
      const unsigned long long nz = block_sum(count_nonzero(data, s, e, mask), red);
      if (vs >= lo && ve <= hi) {
        if (threadIdx.x == 0) counts[2 * v] = nz;  // zeros are derived by the fold kernel
      } else if (vs < lo) {
        part.var_head = v; part.nz_head = nz;
      } else {
        part.var_tail = v; part.nz_tail = nz;
      }
    }
  }
  if (threadIdx.x == 0) partials[blockIdx.x] = part;
}

// In-order fold of the per-CTA partials: partial pieces of one variable are adjacent in CTA order.
__global__ void quantify_fold_kernel(const QPartial* partials, int n_blocks, const uint64_t* offsets, int n_vars,
                                     unsigned long long* counts) {
  __shared__ QPartial sp[kQBlocks];  // parallel load: a serial walk over global memory costs ~0.3 us per record
  for (int b = threadIdx.x; b < n_blocks; b += blockDim.x) sp[b] = partials[b];
  __syncthreads();
  if (threadIdx.x == 0) {  // stores are fire-and-forget, so this scan never waits on global memory
    long long cur = -1;
    unsigned long long sum = 0;
    for (int b = 0; b < n_blocks; ++b) {
      const QPartial p = sp[b];
      if (p.var_head >= 0) {
        if (p.var_head != cur) { if (cur >= 0) counts[2 * cur] = sum; cur = p.var_head; sum = 0; }
        sum += p.nz_head;
      }
      if (p.var_tail >= 0) {
        if (p.var_tail != cur) { if (cur >= 0) counts[2 * cur] = sum; cur = p.var_tail; sum = 0; }
        sum += p.nz_tail;
      }
    }
    if (cur >= 0) counts[2 * cur] = sum;
  }
  __syncthreads();
  for (int v = threadIdx.x; v < n_vars; v += blockDim.x) {
    const unsigned long long len = offsets[v + 1] - offsets[v];
    const unsigned long long nz = len ? counts[2 * v] : 0ull;
    counts[2 * v] = nz;
    counts[2 * v + 1] = len - nz;
  }
}

// Shared by the allocating and the caller-scratch entry points.
int quantify_launch(const void* data, int dtype, const uint64_t* offsets, int n_vars,
                    unsigned long long* counts, void* scratch, cudaStream_t st) {
  QPartial* parts = reinterpret_cast<QPartial*>(scratch);
  const int blocks = std::min(kQBlocks, 2 * sm_count());
  switch (dtype_size(dtype)) {
    case 1: quantify_kernel<uint8_t><<<blocks, kQThreads, 0, st>>>((const uint8_t*)data, 0xffu, offsets, n_vars, counts, parts); break;
    case 2: {
      const uint16_t mask = (dtype == MFL_FLOAT16 || dtype == MFL_BFLOAT16) ? 0x7fffu : 0xffffu;
      quantify_kernel<uint16_t><<<blocks, kQThreads, 0, st>>>((const uint16_t*)data, mask, offsets, n_vars, counts, parts);
      break;
    }
    case 4: {
      const uint32_t mask = dtype == MFL_FLOAT32 ? 0x7fffffffu : 0xffffffffu;
      quantify_kernel<uint32_t><<<blocks, kQThreads, 0, st>>>((const uint32_t*)data, mask, offsets, n_vars, counts, parts);
      break;
    }
    case 8: {
      const unsigned long long mask = dtype == MFL_FLOAT64 ? 0x7fffffffffffffffull : ~0ull;
      quantify_kernel<unsigned long long><<<blocks, kQThreads, 0, st>>>((const unsigned long long*)data, mask, offsets, n_vars, counts, parts);
      break;
    }
    default: set_error("quantify: bad dtype %d", dtype); return -1;
  }
  if (int rc = check(cudaGetLastError(), "quantify_kernel launch")) return rc;
  quantify_fold_kernel<<<1, 128, 0, st>>>(parts, blocks, offsets, n_vars, counts);
  return check(cudaGetLastError(), "quantify_fold_kernel launch");
}

}  // namespace mfl

using namespace mfl;

extern "C" size_t mfl_quantify_scratch_bytes(int /*n_vars*/) { return sizeof(QPartial) * kQBlocks; }

extern "C" int mfl_quantify(const void* data, int dtype, const uint64_t* offsets, int n_vars,
                            unsigned long long* counts, void* scratch, void* stream) {
  MFL_ARG(data && offsets && counts && scratch, "quantify: data, offsets, counts and scratch are required");
  MFL_ARG(n_vars >= 1, "quantify: n_vars must be >= 1");
  MFL_ARG((reinterpret_cast<uintptr_t>(scratch) & 15) == 0, "quantify: scratch must be 16-byte aligned");
  return quantify_launch(data, dtype, offsets, n_vars, counts, scratch, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// Quantifier for callers that cannot allocate device memory.
//
// Some hosts of this library have no way to get scratch for mfl_quantify(): a controller plug-in
// written in plain C that only ever sees the buffers its application hands it; a learner process in
// which a framework's caching allocator has already claimed the GPU's memory, so that a raw
// cudaMalloc may fail or fragment that pool; and code on the round's critical path, where the
// device-wide synchronisation hidden in cudaMalloc/cudaFree would stall the training streams.
// For them the per-CTA partial records live in the fixed array below, which is part of the loaded
// module: every GPU that loads this library holds these MFL_QUANTIFY_NOALLOC_SCRATCH_BYTES bytes
// from load until the process exits, whether or not the entry point is ever called.
//
// `g_noalloc_scratch` is a device symbol.  In host code that name is only a handle for the runtime,
// not a device address, so the host asks the runtime for the real address once per device
// (cudaGetSymbolAddress, with that device current) and passes the result to the kernels.
static __device__ __align__(256) unsigned char g_noalloc_scratch[MFL_QUANTIFY_NOALLOC_SCRATCH_BYTES];

namespace mfl {
constexpr int kMaxDevices = 64;
static void* g_noalloc_ptr[kMaxDevices] = {};
static std::mutex g_noalloc_mu;

// Device address of the static scratch array on the current device (resolved on first use).
static int noalloc_scratch(void** out) {
  int dev = -1;
  MFL_CUDA(cudaGetDevice(&dev));
  MFL_ARG(dev >= 0 && dev < kMaxDevices, "quantify_noalloc: device index %d outside [0,%d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    MFL_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    MFL_ARG(p != nullptr, "quantify_noalloc: cudaGetSymbolAddress returned a null address on device %d", dev);
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace mfl

extern "C" int mfl_quantify_noalloc(const void* data, int dtype, const uint64_t* offsets, int n_vars,
                                    unsigned long long* counts, void* stream) {
  MFL_ARG(data && offsets && counts, "quantify_noalloc: data, offsets and counts are required");
  MFL_ARG(n_vars >= 1, "quantify_noalloc: n_vars must be >= 1");
  // Today the scratch need does not depend on the call (one fixed-size record per CTA of a fixed-size grid: 9472 of the
  // 16384 bytes), so this refusal cannot trigger yet; it is here so that a later change of the grid or the record
  // size turns into a clear error for oversized calls instead of a write past the array.
  static_assert(sizeof(QPartial) * kQBlocks <= MFL_QUANTIFY_NOALLOC_SCRATCH_BYTES,
                "the built-in scratch array no longer holds one record per CTA");
  const size_t need = mfl_quantify_scratch_bytes(n_vars);
  MFL_ARG(need <= (size_t)MFL_QUANTIFY_NOALLOC_SCRATCH_BYTES,
          "quantify_noalloc: this call needs %zu bytes of scratch but the built-in array holds %d; "
          "use mfl_quantify() with caller-provided scratch instead",
          need, MFL_QUANTIFY_NOALLOC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (int rc = noalloc_scratch(&scratch)) return rc;
  // Nothing in the array has to be cleared first: each CTA overwrites its own record in full before
  // the fold kernel (same stream, later launch) reads it, so contents left by an earlier call are
  // never observed.
  return quantify_launch(data, dtype, offsets, n_vars, counts, scratch, (cudaStream_t)stream);
}
