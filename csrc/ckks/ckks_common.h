// RNS-CKKS parameters, blob formats and modular arithmetic shared by the host oracle
// (ckks_host.cc) and the sm_100a kernels (ckks_kernels.cu).
//
// Behavioural spec: the reference wraps PALISADE CKKS with multDepth = 2, scaling factor 2^52,
// batch 4096 (encryption/palisade/ckks_scheme.cc:13-69); its shipped crypto context has ring
// dimension N = 16384 and three RNS moduli (resources/fheparams/cryptoparams/cryptocontext.txt,
// SURVEY §9.4).  Nothing of PALISADE is available here, so the scheme is implemented from the
// textbook construction; the ciphertext byte layout is this library's own (the proto treats
// CiphertextTensor.value as opaque, model.proto:69-72).
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define MFL_HD __host__ __device__ __forceinline__
#else
#define MFL_HD inline
#endif

namespace mfl::ckks {

constexpr uint32_t kMaxLimbs = 4;
// q_i == 1 (mod 2N) for N = 16384: the moduli of the reference's shipped context
constexpr uint64_t kDefaultModuli[3] = {1152921504606748673ull, 4503599626682369ull, 4503599627763713ull};
constexpr uint32_t kDefaultLogN = 14;

// 64-byte header in front of every ciphertext blob (one blob = all batches of one tensor).
struct CtHeader {
  char magic[4];        // "MFCK"
  uint16_t version;     // 1
  uint8_t log_n;
  uint8_t n_limbs;
  uint32_t n_ct;        // ciphertexts (batches) in the blob
  uint32_t slots;       // packed values per ciphertext
  uint64_t n_values;    // real values encoded in total
  double scale;         // current scaling factor (Delta, Delta^2 after a weighted average, ...)
  uint64_t q[kMaxLimbs];
};
static_assert(sizeof(CtHeader) == 64, "CtHeader must be 64 bytes");

MFL_HD uint64_t ct_poly_words(const CtHeader& h) { return (uint64_t)h.n_limbs << h.log_n; }
MFL_HD uint64_t ct_words(const CtHeader& h) { return 2 * ct_poly_words(h); }  // per ciphertext

// ---- modular arithmetic (moduli < 2^62) ------------------------------------------------------
MFL_HD uint64_t add_mod(uint64_t a, uint64_t b, uint64_t q) {
  uint64_t s = a + b;
  return s >= q ? s - q : s;
}
MFL_HD uint64_t sub_mod(uint64_t a, uint64_t b, uint64_t q) { return a >= b ? a - b : a + q - b; }

MFL_HD uint64_t mul_hi(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
  return __umul64hi(a, b);
#else
  return (uint64_t)(((unsigned __int128)a * b) >> 64);
#endif
}

// Shoup multiplication: w is a constant with precomputed wp = floor(w * 2^64 / q).
MFL_HD uint64_t mul_shoup(uint64_t a, uint64_t w, uint64_t wp, uint64_t q) {
  const uint64_t hi = mul_hi(a, wp);
  uint64_t r = a * w - hi * q;
  return r >= q ? r - q : r;
}

// Barrett-style generic product via 128-bit: used where no precomputation pays off.
MFL_HD uint64_t mul_mod(uint64_t a, uint64_t b, uint64_t q) {
#if defined(__CUDA_ARCH__)
  // 128-bit product reduced with a 2-step 64-bit estimate (q < 2^62 so the quotient fits)
  const uint64_t lo = a * b, hi = __umul64hi(a, b);
  // r = (hi * 2^64 + lo) mod q using double-word long division by parts
  unsigned __int128 x = ((unsigned __int128)hi << 64) | lo;
  return (uint64_t)(x % q);
#else
  return (uint64_t)(((unsigned __int128)a * b) % q);
#endif
}

inline uint64_t pow_mod(uint64_t b, uint64_t e, uint64_t q) {
  uint64_t r = 1;
  b %= q;
  while (e) {
    if (e & 1) r = (uint64_t)(((unsigned __int128)r * b) % q);
    b = (uint64_t)(((unsigned __int128)b * b) % q);
    e >>= 1;
  }
  return r;
}
inline uint64_t inv_mod(uint64_t a, uint64_t q) { return pow_mod(a, q - 2, q); }  // q prime
inline uint64_t shoup_const(uint64_t w, uint64_t q) { return (uint64_t)(((unsigned __int128)w << 64) / q); }

MFL_HD uint32_t bit_reverse(uint32_t x, uint32_t bits) {
  uint32_t r = 0;
  for (uint32_t i = 0; i < bits; ++i) {
    r = (r << 1) | (x & 1);
    x >>= 1;
  }
  return r;
}

}  // namespace mfl::ckks
