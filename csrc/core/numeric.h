// Element types of the wire dtypes on the host, unaligned access, and dtype dispatch.
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>

namespace mfl::core {

// bfloat16 carried as its 16 raw bits; arithmetic goes through float (round-to-nearest-even back).
struct bf16_t {
  uint16_t bits;
  bf16_t() : bits(0) {}
  bf16_t(double v) { *this = from_float((float)v); }
  bf16_t(float v) { *this = from_float(v); }
  bf16_t(int v) { *this = from_float((float)v); }
  static bf16_t from_float(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    bf16_t r;
    if ((u & 0x7fffffffu) > 0x7f800000u) { r.bits = (uint16_t)((u >> 16) | 0x40); return r; }  // NaN
    u += 0x7fffu + ((u >> 16) & 1u);
    r.bits = (uint16_t)(u >> 16);
    return r;
  }
  operator float() const {
    uint32_t u = (uint32_t)bits << 16;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
  }
  operator double() const { return (double)(float)*this; }
  bf16_t operator+(bf16_t o) const { return bf16_t((float)*this + (float)o); }
  bf16_t operator-(bf16_t o) const { return bf16_t((float)*this - (float)o); }
};

using half_t = _Float16;

template <typename T> inline T load_unaligned(const T* p) {
  T v;
  std::memcpy(&v, p, sizeof(T));
  return v;
}
template <typename T> inline void store_unaligned(T* p, T v) { std::memcpy(p, &v, sizeof(T)); }

#define MFL_HOST_DISPATCH(dt, T, ...)                                     \
  switch (dt) {                                                           \
    case 0: { using T = int8_t; __VA_ARGS__; break; }                     \
    case 1: { using T = int16_t; __VA_ARGS__; break; }                    \
    case 2: { using T = int32_t; __VA_ARGS__; break; }                    \
    case 3: { using T = int64_t; __VA_ARGS__; break; }                    \
    case 4: { using T = uint8_t; __VA_ARGS__; break; }                    \
    case 5: { using T = uint16_t; __VA_ARGS__; break; }                   \
    case 6: { using T = uint32_t; __VA_ARGS__; break; }                   \
    case 7: { using T = uint64_t; __VA_ARGS__; break; }                   \
    case 8: { using T = float; __VA_ARGS__; break; }                      \
    case 9: { using T = double; __VA_ARGS__; break; }                     \
    case 10: { using T = ::mfl::core::half_t; __VA_ARGS__; break; }       \
    case 11: { using T = ::mfl::core::bf16_t; __VA_ARGS__; break; }       \
    default: throw std::runtime_error("unsupported tensor data type");    \
  }

}  // namespace mfl::core
