// K12: batched cast/copy between wire-format tensor bytes and typed device arenas.
//
// Behavioural spec: the reference packs/unpacks every tensor on the host
// (`arr.flatten().tobytes()` / `np.frombuffer().reshape`, proto_messages_factory.py:419-465,484-494;
// memcpy (de)serialisation in controller/common/proto_tensor_serde.h:14-32).
//
// B200 design: the host only parses varints to find each `TensorSpec.value` span; the payload of a
// whole Model message is copied to the device once and ONE launch scatters/casts all variables into
// the contiguous parameter arena (or gathers them back), including big-endian wire data and
// fp32 <-> bf16/fp16 conversion.  A CTA group walks a (tensor, chunk) work list derived from the
// device-resident offset tables, so tiny and huge tensors share the same launch.
#include <algorithm>

#include "common.cuh"

namespace mfl {

template <typename T> __device__ __forceinline__ T bswap(T v) {
  if constexpr (sizeof(T) == 1) return v;
  unsigned char* b = reinterpret_cast<unsigned char*>(&v);
#pragma unroll
  for (int i = 0; i < (int)sizeof(T) / 2; ++i) {
    unsigned char t = b[i]; b[i] = b[sizeof(T) - 1 - i]; b[sizeof(T) - 1 - i] = t;
  }
  return v;
}

template <typename D, typename S>
__global__ void __launch_bounds__(256) pack_cast_kernel(char* dst, const char* src, const uint64_t* dst_off,
                                                        const uint64_t* src_off, const uint64_t* numel,
                                                        int n_tensors, int byte_swap) {
  constexpr size_t kChunk = 256 * 16;  // elements per (CTA, step)
  for (int t = 0; t < n_tensors; ++t) {
    const size_t n = numel[t];
    D* d = reinterpret_cast<D*>(dst + dst_off[t]);
    const S* s = reinterpret_cast<const S*>(src + src_off[t]);
    for (size_t base = (size_t)blockIdx.x * kChunk; base < n; base += (size_t)gridDim.x * kChunk) {
      const size_t end = min(n, base + kChunk);
      for (size_t i = base + threadIdx.x; i < end; i += 256) {
        S x = s[i];
        if (byte_swap) x = bswap(x);
        if constexpr (std::is_same<D, S>::value) d[i] = x;
        else d[i] = from_double<D>(to_double(x));
      }
    }
  }
}

}  // namespace mfl

using namespace mfl;

extern "C" int mfl_pack_cast(void* dst, int dst_dtype, const void* src, int src_dtype,
                             const uint64_t* dst_off_bytes, const uint64_t* src_off_bytes,
                             const uint64_t* numel, int n_tensors, int byte_swap, void* stream) {
  MFL_ARG(dst && src && dst_off_bytes && src_off_bytes && numel, "pack_cast: null argument");
  MFL_ARG(dtype_size(dst_dtype) && dtype_size(src_dtype), "pack_cast: bad dtype");
  if (n_tensors <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int blocks = sm_count() * 4;
  bool launched = false;
#define MFL_PACK_TO(DT)                                                                                     \
  MFL_DISPATCH_ALL(src_dtype, S, (pack_cast_kernel<DT, S><<<blocks, 256, 0, st>>>(                          \
                                      (char*)dst, (const char*)src, dst_off_bytes, src_off_bytes, numel,    \
                                      n_tensors, byte_swap), launched = true))
  if (dst_dtype == src_dtype) {
    // same-type copy (possibly byte-swapped): dispatch on width only
    switch (dtype_size(src_dtype)) {
      case 1: pack_cast_kernel<uint8_t, uint8_t><<<blocks, 256, 0, st>>>((char*)dst, (const char*)src, dst_off_bytes, src_off_bytes, numel, n_tensors, byte_swap); break;
      case 2: pack_cast_kernel<uint16_t, uint16_t><<<blocks, 256, 0, st>>>((char*)dst, (const char*)src, dst_off_bytes, src_off_bytes, numel, n_tensors, byte_swap); break;
      case 4: pack_cast_kernel<uint32_t, uint32_t><<<blocks, 256, 0, st>>>((char*)dst, (const char*)src, dst_off_bytes, src_off_bytes, numel, n_tensors, byte_swap); break;
      default: pack_cast_kernel<uint64_t, uint64_t><<<blocks, 256, 0, st>>>((char*)dst, (const char*)src, dst_off_bytes, src_off_bytes, numel, n_tensors, byte_swap); break;
    }
    launched = true;
  } else if (dst_dtype == MFL_FLOAT32) { MFL_PACK_TO(float); }
  else if (dst_dtype == MFL_BFLOAT16) { MFL_PACK_TO(__nv_bfloat16); }
  else if (dst_dtype == MFL_FLOAT16) { MFL_PACK_TO(__half); }
  else if (dst_dtype == MFL_FLOAT64) { MFL_PACK_TO(double); }
#undef MFL_PACK_TO
  MFL_ARG(launched, "pack_cast: unsupported conversion %d -> %d (casts target FLOAT32/64/16/BFLOAT16)", src_dtype, dst_dtype);
  return check(cudaGetLastError(), "pack_cast_kernel launch");
}
