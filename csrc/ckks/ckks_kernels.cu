// K5-K9: RNS-CKKS kernels for sm_100a.
//
// Behavioural spec: the reference's PALISADE-backed CKKS wrapper
// (encryption/palisade/ckks_scheme.cc:110-252: batch-wise packed Encrypt, ComputeWeightedAverage =
// sum EvalMult(ct, w) / EvalAdd, Decrypt) and the PWA rule that calls it
// (controller/aggregation/private_weighted_average.cc:24-82).  No PALISADE code exists here; the
// math is the textbook RNS-CKKS of csrc/ckks/ckks_host.cc, which is also the oracle for every
// kernel in this file.
//
// B200 design: a limb polynomial is N = 16384 x uint64 = 128 KB, which fits the 227 KB shared
// memory of one CTA, so every (inverse) NTT runs entirely in shared memory: one global read, 14
// butterfly stages with Shoup multiplication (no 128-bit division anywhere), one global write.
// Encrypt fuses sampling + three NTTs + the public-key multiply-adds per (ciphertext, limb) row;
// decrypt fuses c0 + c1*s with the inverse NTT.  The weighted reduction (K8) is bandwidth bound
// (48x plaintext traffic) and has a fused NVLink variant with the same in-kernel barrier protocol as
// the plaintext aggregation kernel.
#include <algorithm>

#include "ckks/ckks_common.h"
#include "kernels/common.cuh"
#include "kernels/tcgen05_common.cuh"   // mbarrier helpers

namespace mfl {
namespace ckksdev {

using namespace ::mfl::ckks;

constexpr int kNttThreads = 1024;

// Shared-memory placement of coefficient i of the polynomial a CTA works on: one 8-byte pad after every 16
// coefficients.  The last two radix-4 passes touch coefficients 4 and 1 apart with 16- and 4-coefficient strides
// between lanes; unpadded, the 32 lanes of a warp then fall on 4 bank pairs (8-way conflicts on all eight accesses of
// a butterfly: about a third of the transform's time); padded they spread over all 16 (q4 = 1) or 11 (q4 = 4).
__device__ __forceinline__ uint32_t px(uint32_t i) { return i + (i >> 4); }
static size_t poly_smem_bytes(int log_n) { return sizeof(uint64_t) * (((size_t)1 << log_n) + ((size_t)1 << log_n >> 4) + 1); }

struct Ctx {  // by-value copy of mfl_ckks_ctx for kernel parameters
  mfl_ckks_ctx c;
};

__device__ __forceinline__ uint64_t d_mul_shoup(uint64_t a, uint64_t w, uint64_t wp, uint64_t q) {
  const uint64_t hi = __umul64hi(a, wp);
  uint64_t r = a * w - hi * q;
  return r >= q ? r - q : r;
}
__device__ __forceinline__ uint64_t d_add(uint64_t a, uint64_t b, uint64_t q) {
  const uint64_t s = a + b;
  return s >= q ? s - q : s;
}
__device__ __forceinline__ uint64_t d_sub(uint64_t a, uint64_t b, uint64_t q) { return a >= b ? a - b : a + q - b; }

// Forward negacyclic NTT of one polynomial held in shared memory (Cooley-Tukey, natural -> bit-reversed).
// Radix 4: two consecutive radix-2 stages (m, 2m) are applied to four elements in registers, so a polynomial of 2^14
// coefficients takes 7 passes over shared memory and 7 block barriers instead of 14 (one leading radix-2 stage when
// log_n is odd).  Same butterflies, same twiddle table, same results bit for bit as the radix-2 loop.
__device__ void ntt_forward_smem(uint64_t* x, const uint64_t* __restrict__ w, const uint64_t* __restrict__ wp,
                                 uint64_t q, int log_n) {
  const int n = 1 << log_n;
  int t = n, m = 1;
  if (log_n & 1) {  // one radix-2 stage first
    t >>= 1;
    for (int b = threadIdx.x; b < (n >> 1); b += blockDim.x) {
      const uint64_t s = __ldg(w + 1), sp = __ldg(wp + 1);
      const uint64_t u = x[px(b)], v = d_mul_shoup(x[px(b + t)], s, sp, q);
      x[px(b)] = d_add(u, v, q);
      x[px(b + t)] = d_sub(u, v, q);
    }
    __syncthreads();
    m = 2;
  }
  for (; m < n; m <<= 2) {
    // stage A: (m, t/2) pairs at distance t/2 ... written with t = distance of stage A's pairs
    t >>= 1;                       // pair distance of the first of the two stages
    const int q4 = t >> 1;         // pair distance of the second stage = quarter of the 2t-wide block
    const int log_q4 = __ffs(q4) - 1;
    for (int b = threadIdx.x; b < (n >> 2); b += blockDim.x) {
      const int i = b >> log_q4, j = b & (q4 - 1);
      const int idx = (i << (log_q4 + 2)) + j;
      const uint64_t w1 = __ldg(w + m + i), w1p = __ldg(wp + m + i);
      const uint64_t w2 = __ldg(w + 2 * m + 2 * i), w2p = __ldg(wp + 2 * m + 2 * i);
      const uint64_t w3 = __ldg(w + 2 * m + 2 * i + 1), w3p = __ldg(wp + 2 * m + 2 * i + 1);
      const uint64_t a0 = x[px(idx)], a1 = x[px(idx + q4)], a2 = x[px(idx + 2 * q4)], a3 = x[px(idx + 3 * q4)];
      const uint64_t v2 = d_mul_shoup(a2, w1, w1p, q), v3 = d_mul_shoup(a3, w1, w1p, q);
      const uint64_t b0 = d_add(a0, v2, q), b2 = d_sub(a0, v2, q);
      const uint64_t b1 = d_add(a1, v3, q), b3 = d_sub(a1, v3, q);
      const uint64_t u1 = d_mul_shoup(b1, w2, w2p, q), u3 = d_mul_shoup(b3, w3, w3p, q);
      x[px(idx)] = d_add(b0, u1, q);
      x[px(idx + q4)] = d_sub(b0, u1, q);
      x[px(idx + 2 * q4)] = d_add(b2, u3, q);
      x[px(idx + 3 * q4)] = d_sub(b2, u3, q);
    }
    __syncthreads();
    t >>= 1;
  }
}

// Inverse (Gentleman-Sande, bit-reversed -> natural), including the 1/N scaling; radix 4 like the forward transform
// (the trailing radix-2 stage when log_n is odd), with the 1/N factor folded into the last pass.
__device__ void ntt_inverse_smem(uint64_t* x, const uint64_t* __restrict__ w, const uint64_t* __restrict__ wp,
                                 uint64_t q, uint64_t n_inv, uint64_t n_inv_shoup, int log_n) {
  const int n = 1 << log_n;
  int t = 1, m = n;
  for (; m > 2; m >>= 2) {
    const int h = m >> 1;           // twiddle base of the first stage (pairs at distance t)
    const int log_t = __ffs(t) - 1;
    for (int b = threadIdx.x; b < (n >> 2); b += blockDim.x) {
      const int i = b >> log_t, j = b & (t - 1);      // i indexes blocks of 4t elements
      const int idx = (i << (log_t + 2)) + j;
      const uint64_t s0 = __ldg(w + h + 2 * i), s0p = __ldg(wp + h + 2 * i);
      const uint64_t s1 = __ldg(w + h + 2 * i + 1), s1p = __ldg(wp + h + 2 * i + 1);
      const uint64_t s2 = __ldg(w + (h >> 1) + i), s2p = __ldg(wp + (h >> 1) + i);
      const uint64_t a0 = x[px(idx)], a1 = x[px(idx + t)], a2 = x[px(idx + 2 * t)], a3 = x[px(idx + 3 * t)];
      const uint64_t b0 = d_add(a0, a1, q), b1 = d_mul_shoup(d_sub(a0, a1, q), s0, s0p, q);
      const uint64_t b2 = d_add(a2, a3, q), b3 = d_mul_shoup(d_sub(a2, a3, q), s1, s1p, q);
      x[px(idx)] = d_add(b0, b2, q);
      x[px(idx + 2 * t)] = d_mul_shoup(d_sub(b0, b2, q), s2, s2p, q);
      x[px(idx + t)] = d_add(b1, b3, q);
      x[px(idx + 3 * t)] = d_mul_shoup(d_sub(b1, b3, q), s2, s2p, q);
    }
    __syncthreads();
    t <<= 2;
  }
  if (m == 2) {  // odd log_n: the last radix-2 stage (h = 1, pairs at distance n / 2)
    const uint64_t s = __ldg(w + 1), sp = __ldg(wp + 1);
    for (int b = threadIdx.x; b < (n >> 1); b += blockDim.x) {
      const uint64_t u = x[px(b)], v = x[px(b + t)];
      x[px(b)] = d_add(u, v, q);
      x[px(b + t)] = d_mul_shoup(d_sub(u, v, q), s, sp, q);
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) x[px(i)] = d_mul_shoup(x[px(i)], n_inv, n_inv_shoup, q);
  __syncthreads();
}

__global__ void __launch_bounds__(kNttThreads, 1) ntt_rows_kernel(const __grid_constant__ Ctx cx, uint64_t* rows,
                                                                  size_t n_rows, int inverse) {
  extern __shared__ __align__(16) uint64_t sm[];
  const int n = 1 << cx.c.log_n;
  for (size_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
    const int l = (int)(r % cx.c.n_limbs);
    uint64_t* g = rows + r * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) sm[px(i)] = g[i];
    __syncthreads();
    if (inverse) ntt_inverse_smem(sm, cx.c.ipsi[l], cx.c.ipsi_shoup[l], cx.c.q[l], cx.c.n_inv[l], cx.c.n_inv_shoup[l], cx.c.log_n);
    else ntt_forward_smem(sm, cx.c.psi[l], cx.c.psi_shoup[l], cx.c.q[l], cx.c.log_n);
    for (int i = threadIdx.x; i < n; i += blockDim.x) g[i] = sm[px(i)];
    __syncthreads();
  }
}

// ---- counter-based CSPRNG: ChaCha20 block function (RFC 8439 state layout, 64-bit block counter) -------------------
// One 256-bit key per encrypt call (fresh from the OS on the host side).  A block is addressed by
// (counter = ciphertext index << 16 | block-in-polynomial, stream id): stream 0 = ternary u, 1 = e0, 2 = e1, so no
// block is ever used for two purposes and every limb row of a ciphertext regenerates the SAME u / e0 / e1.
struct ChaChaKey { uint32_t k[8]; };
struct ChaCha20 {
  ChaChaKey key;
  __device__ __forceinline__ static void qr(uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
    a += b; d ^= a; d = __funnelshift_l(d, d, 16);
    c += d; b ^= c; b = __funnelshift_l(b, b, 12);
    a += b; d ^= a; d = __funnelshift_l(d, d, 8);
    c += d; b ^= c; b = __funnelshift_l(b, b, 7);
  }
  __device__ __forceinline__ void block(uint64_t counter, uint32_t stream_id, uint32_t* out) const {
    const uint32_t in[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u,
                             key.k[0], key.k[1], key.k[2], key.k[3], key.k[4], key.k[5], key.k[6], key.k[7],
                             (uint32_t)counter, (uint32_t)(counter >> 32), stream_id, 0u};
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = in[i];
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      qr(x[0], x[4], x[8], x[12]); qr(x[1], x[5], x[9], x[13]); qr(x[2], x[6], x[10], x[14]); qr(x[3], x[7], x[11], x[15]);
      qr(x[0], x[5], x[10], x[15]); qr(x[1], x[6], x[11], x[12]); qr(x[2], x[7], x[8], x[13]); qr(x[3], x[4], x[9], x[14]);
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) out[i] = x[i] + in[i];
  }
};

// uniform {-1, 0, 1} from one 32-bit word by rejection-free multiply-shift (bias 2^-32, below the scheme's noise budget)
__device__ __forceinline__ int ternary_from_bits(uint32_t a) { return (int)__umulhi(a, 3u) - 1; }

__device__ __forceinline__ int gaussian_from_bits(uint32_t a, uint32_t b, float sigma) {
  const float u1 = (a + 1.0f) * 2.3283064e-10f;  // (0, 1]
  const float u2 = b * 2.3283064e-10f;
  return __float2int_rn(sigma * sqrtf(-2.0f * __logf(u1)) * __cosf(6.2831853f * u2));
}
__device__ __forceinline__ uint64_t signed_to_mod(int v, uint64_t q) { return v >= 0 ? (uint64_t)v : q - (uint64_t)(-v); }

// rounded (integral) double, |x| < 2^103 -> residue mod q
__device__ __forceinline__ uint64_t double_to_mod(double x, uint64_t q, uint64_t two40_mod_q) {
  const bool neg = x < 0;
  const double ax = fabs(x);
  uint64_t r;
  if (ax < 9.0e18) {
    r = (uint64_t)ax % q;
  } else {
    const double hi = floor(ax * 9.094947017729282e-13);  // / 2^40
    const double lo = ax - hi * 1099511627776.0;
    const uint64_t h = (uint64_t)hi % q;
    // (h * 2^40 + lo) mod q with 128-bit product via two halves (q < 2^62, 2^40 mod q < 2^40)
    const uint64_t plo = h * two40_mod_q, phi = __umul64hi(h, two40_mod_q);
    unsigned __int128 z = ((unsigned __int128)phi << 64) | plo;
    r = (uint64_t)((z + (uint64_t)lo) % q);
  }
  return neg && r ? q - r : r;
}


// ---- K6: canonical-embedding encode / decode (special FFT over the 2N-th roots, fp64) ----------
struct cplx { double re, im; };
__device__ __forceinline__ cplx cmul(cplx a, cplx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }

__global__ void __launch_bounds__(512) encode_kernel(const double* __restrict__ values, size_t n_values, uint32_t slots,
                                                     double scale, const double* __restrict__ ksi,
                                                     const uint32_t* __restrict__ rot, double* coeffs, int log_n) {
  extern __shared__ __align__(16) unsigned char raw[];
  cplx* v = reinterpret_cast<cplx*>(raw);
  const uint32_t ct = blockIdx.x, M = 2u << log_n;
  const size_t base = (size_t)ct * slots;
  for (uint32_t i = threadIdx.x; i < slots; i += blockDim.x)
    v[i] = {base + i < n_values ? values[base + i] : 0.0, 0.0};
  __syncthreads();
  for (uint32_t len = slots; len >= 2; len >>= 1) {
    const uint32_t lenh = len >> 1, lenq = len << 2;
    for (uint32_t b = threadIdx.x; b < slots / 2; b += blockDim.x) {
      const uint32_t i = (b / lenh) * len, j = b % lenh;
      const uint32_t idx = (lenq - (rot[j] % lenq)) * (M / lenq);
      const cplx a = v[i + j], c = v[i + j + lenh];
      const cplx w = {ksi[2 * idx], ksi[2 * idx + 1]};
      v[i + j] = {a.re + c.re, a.im + c.im};
      v[i + j + lenh] = cmul({a.re - c.re, a.im - c.im}, w);
    }
    __syncthreads();
  }
  const uint32_t bits = 31 - __clz(slots);
  const double f = scale / (double)slots;
  double* out = coeffs + (size_t)ct * 2 * slots;
  for (uint32_t i = threadIdx.x; i < slots; i += blockDim.x) {
    const uint32_t r = __brev(i) >> (32 - bits);
    out[r] = rint(v[i].re * f);
    out[slots + r] = rint(v[i].im * f);
  }
}

struct GarnerInv { uint64_t inv[16]; };

__device__ __forceinline__ uint64_t d_mul_mod_slow(uint64_t a, uint64_t b, uint64_t q) {
  return (uint64_t)(((unsigned __int128)a * b) % q);
}

// residues of one coefficient -> centred double (mixed-radix reconstruction, see ckks_host.cc)
__device__ double garner_to_double(const uint64_t* res, const uint64_t* q, int L, const GarnerInv& gi) {
  uint64_t d[4];
  for (int i = 0; i < L; ++i) {
    uint64_t t = res[i];
    for (int j = 0; j < i; ++j) t = d_mul_mod_slow(d_sub(t, d[j] % q[i], q[i]), gi.inv[j * 4 + i], q[i]);
    d[i] = t;
  }
  const bool neg = d[L - 1] > (q[L - 1] - 1) / 2;
  if (neg) {
    bool borrow = true;
    for (int i = 0; i < L; ++i) {
      uint64_t x = q[i] - 1 - d[i];
      if (borrow) {
        x += 1;
        if (x == q[i]) x = 0; else borrow = false;
      }
      d[i] = x;
    }
  }
  double acc = 0.0;
  for (int i = L - 1; i >= 0; --i) acc = acc * (double)q[i] + (double)d[i];
  return neg ? -acc : acc;
}

__global__ void __launch_bounds__(512) decode_kernel(const __grid_constant__ Ctx cx, const uint64_t* __restrict__ residues,
                                                     uint32_t slots, double inv_scale, const __grid_constant__ GarnerInv gi,
                                                     const double* __restrict__ ksi, const uint32_t* __restrict__ rot,
                                                     double* values, size_t n_values) {
  extern __shared__ __align__(16) unsigned char raw[];
  cplx* v = reinterpret_cast<cplx*>(raw);
  const int L = cx.c.n_limbs, log_n = cx.c.log_n;
  const uint32_t N = 1u << log_n, M = 2 * N, gap = (N / 2) / slots, ct = blockIdx.x;
  const uint32_t bits = 31 - __clz(slots);
  const uint64_t* rbase = residues + (size_t)ct * L * N;
  for (uint32_t k = threadIdx.x; k < slots; k += blockDim.x) {
    uint64_t r0[4], r1[4];
    for (int l = 0; l < L; ++l) {
      r0[l] = rbase[(size_t)l * N + (size_t)k * gap];
      r1[l] = rbase[(size_t)l * N + (size_t)k * gap + N / 2];
    }
    const uint32_t p = __brev(k) >> (32 - bits);  // bit-reversal permutation folded into the load
    v[p] = {garner_to_double(r0, cx.c.q, L, gi) * inv_scale, garner_to_double(r1, cx.c.q, L, gi) * inv_scale};
  }
  __syncthreads();
  for (uint32_t len = 2; len <= slots; len <<= 1) {
    const uint32_t lenh = len >> 1, lenq = len << 2;
    for (uint32_t b = threadIdx.x; b < slots / 2; b += blockDim.x) {
      const uint32_t i = (b / lenh) * len, j = b % lenh;
      const uint32_t idx = (rot[j] % lenq) * (M / lenq);
      const cplx a = v[i + j];
      const cplx c = cmul(v[i + j + lenh], {ksi[2 * idx], ksi[2 * idx + 1]});
      v[i + j] = {a.re + c.re, a.im + c.im};
      v[i + j + lenh] = {a.re - c.re, a.im - c.im};
    }
    __syncthreads();
  }
  const size_t base = (size_t)ct * slots;
  for (uint32_t k = threadIdx.x; k < slots; k += blockDim.x)
    if (base + k < n_values) values[base + k] = v[k].re;
}

// ---- K7: encrypt / decrypt ---------------------------------------------------------------------
__global__ void __launch_bounds__(kNttThreads, 1) encrypt_kernel(const __grid_constant__ Ctx cx, const double* __restrict__ coeffs,
                                                                 uint32_t slots, const uint64_t* __restrict__ pk,
                                                                 const uint64_t* __restrict__ pks,
                                                                 const __grid_constant__ ChaChaKey key,
                                                                 uint64_t* ct_out, uint32_t n_ct) {
  extern __shared__ __align__(16) uint64_t sm[];
  const int L = cx.c.n_limbs, log_n = cx.c.log_n;
  const uint32_t N = 1u << log_n, gap = (N / 2) / slots;
  const ChaCha20 rng{key};
  constexpr float kSigma = 3.2f;
  for (uint32_t row = blockIdx.x; row < n_ct * (uint32_t)L; row += gridDim.x) {
    const uint32_t ct = row / L, l = row % L;
    const uint64_t q = cx.c.q[l];
    const uint64_t two40 = (1ull << 40) % q;
    const uint64_t *w = cx.c.psi[l], *wp = cx.c.psi_shoup[l];
    uint64_t* c0 = ct_out + ((size_t)ct * 2 * L + l) * N;
    uint64_t* c1 = c0 + (size_t)L * N;
    const uint64_t *pk0 = pk + (size_t)l * N, *pk1 = pk + (size_t)(L + l) * N;
    const uint64_t *pk0s = pks + (size_t)l * N, *pk1s = pks + (size_t)(L + l) * N;
    const double* m = coeffs + (size_t)ct * 2 * slots;
    uint32_t r[16];
    // Word j of block b lands on coefficient j * (N / words-per-block-used) + b: consecutive threads (consecutive b)
    // write consecutive shared-memory words, and the map (b, j) -> coefficient is a bijection.
    const uint32_t nb16 = N / 16, nb8 = N / 8;
    const uint64_t ctr0 = (uint64_t)ct << 16;
    // pass 1: u (ternary, the same polynomial for every limb of this ciphertext), 16 coefficients per block
    for (uint32_t b = threadIdx.x; b < nb16; b += blockDim.x) {
      rng.block(ctr0 | b, 0, r);
#pragma unroll
      for (int j = 0; j < 16; ++j) sm[px(j * nb16 + b)] = signed_to_mod(ternary_from_bits(r[j]), q);
    }
    __syncthreads();
    ntt_forward_smem(sm, w, wp, q, log_n);
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) {
      const uint64_t u = sm[px(i)];
      c0[i] = d_mul_shoup(u, pk0[i], pk0s[i], q);
      c1[i] = d_mul_shoup(u, pk1[i], pk1s[i], q);
    }
    __syncthreads();
    // pass 2: c0 += NTT(m + e0)
    for (uint32_t b = threadIdx.x; b < nb8; b += blockDim.x) {
      rng.block(ctr0 | b, 1, r);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t i = j * nb8 + b;
        uint64_t t = signed_to_mod(gaussian_from_bits(r[2 * j], r[2 * j + 1], kSigma), q);
        if (i % gap == 0) t = d_add(t, double_to_mod(m[i / gap], q, two40), q);
        sm[px(i)] = t;
      }
    }
    __syncthreads();
    ntt_forward_smem(sm, w, wp, q, log_n);
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) c0[i] = d_add(c0[i], sm[px(i)], q);
    __syncthreads();
    // pass 3: c1 += NTT(e1)
    for (uint32_t b = threadIdx.x; b < nb8; b += blockDim.x) {
      rng.block(ctr0 | b, 2, r);
#pragma unroll
      for (int j = 0; j < 8; ++j) sm[px(j * nb8 + b)] = signed_to_mod(gaussian_from_bits(r[2 * j], r[2 * j + 1], kSigma), q);
    }
    __syncthreads();
    ntt_forward_smem(sm, w, wp, q, log_n);
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) c1[i] = d_add(c1[i], sm[px(i)], q);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kNttThreads, 1) decrypt_kernel(const __grid_constant__ Ctx cx, const uint64_t* __restrict__ ct,
                                                                 uint32_t n_ct, const uint64_t* __restrict__ sk,
                                                                 const uint64_t* __restrict__ sks, uint64_t* residues) {
  extern __shared__ __align__(16) uint64_t sm[];
  const int L = cx.c.n_limbs, log_n = cx.c.log_n;
  const uint32_t N = 1u << log_n;
  for (uint32_t row = blockIdx.x; row < n_ct * (uint32_t)L; row += gridDim.x) {
    const uint32_t c = row / L, l = row % L;
    const uint64_t q = cx.c.q[l];
    const uint64_t* c0 = ct + ((size_t)c * 2 * L + l) * N;
    const uint64_t* c1 = c0 + (size_t)L * N;
    const uint64_t *s = sk + (size_t)l * N, *ss = sks + (size_t)l * N;
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) sm[px(i)] = d_add(c0[i], d_mul_shoup(c1[i], s[i], ss[i], q), q);
    __syncthreads();
    ntt_inverse_smem(sm, cx.c.ipsi[l], cx.c.ipsi_shoup[l], q, cx.c.n_inv[l], cx.c.n_inv_shoup[l], log_n);
    uint64_t* out = residues + ((size_t)c * L + l) * N;
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) out[i] = sm[px(i)];
    __syncthreads();
  }
}


// ---- K8: limb-wise weighted reduction of ciphertext blobs --------------------------------------
struct WsArgs {
  const uint64_t* src[MFL_MAX_SRC];
  uint64_t w[MFL_MAX_SRC][4], wp[MFL_MAX_SRC][4];
  uint64_t q[4];
  int n_src, n_limbs, log_n;
};

// One multicast store lands in every GPU's output buffer (NVSwitch replicates it): 1/world of the outbound NVLink
// traffic of world separate peer stores.  `mc` is the multicast address of the symmetric output buffers.
__device__ __forceinline__ void ws_store(uint64_t* const* outs, int n_out, uint64_t* mc, size_t i, const uint4& o) {
  if (mc) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc + i), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
  } else {
    for (int t = 0; t < n_out; ++t) st_stream(outs[t] + i, o);
  }
}

// One ciphertext model per GPU (2, 4 or 8 sources): all sixteen 16-byte loads of 16 / NS element pairs are issued before
// the first modular multiply, so a thread keeps 256 B in flight over NVLink instead of 64 B (the 4-load batches of the
// general loop spent most of a round trip with nothing outstanding: 0.40 of the link roofline at 8 GPUs).
template <bool PEER, int NS>
__device__ __forceinline__ void ws_range_batched(const WsArgs& a, size_t begin, size_t end, uint64_t* const* outs,
                                                 int n_out, uint64_t* mc) {
  const size_t stride = (size_t)gridDim.x * blockDim.x * 2;
  constexpr int P = 16 / NS;
  for (size_t base = begin + ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 2; base < end; base += stride * P) {
    uint4 v[P][NS];
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const size_t i = base + p * stride;
#pragma unroll
      for (int j = 0; j < NS; ++j)
        v[p][j] = i < end ? (PEER ? ld_peer(a.src[j] + i) : ld_stream(a.src[j] + i)) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const size_t i = base + p * stride;
      if (i >= end) break;
      const int limb = (int)((i >> a.log_n) % a.n_limbs);
      const uint64_t q = a.q[limb];
      uint64_t acc0 = 0, acc1 = 0;
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        const uint64_t x0 = ((uint64_t)v[p][j].y << 32) | v[p][j].x, x1 = ((uint64_t)v[p][j].w << 32) | v[p][j].z;
        acc0 = d_add(acc0, d_mul_shoup(x0, a.w[j][limb], a.wp[j][limb], q), q);
        acc1 = d_add(acc1, d_mul_shoup(x1, a.w[j][limb], a.wp[j][limb], q), q);
      }
      const uint4 o = make_uint4((uint32_t)acc0, (uint32_t)(acc0 >> 32), (uint32_t)acc1, (uint32_t)(acc1 >> 32));
      ws_store(outs, n_out, mc, i, o);
    }
  }
}

// out[i] = sum_l src_l[i] * w_l[limb(i)] mod q[limb(i)] for words i in [begin, end), 2 words / thread-step
template <bool PEER>
__device__ __forceinline__ void ws_range(const WsArgs& a, size_t begin, size_t end, uint64_t* const* outs, int n_out,
                                         uint64_t* mc = nullptr) {
  const size_t stride = (size_t)gridDim.x * blockDim.x * 2;
  if (a.n_src == 8) return ws_range_batched<PEER, 8>(a, begin, end, outs, n_out, mc);
  if (a.n_src == 4) return ws_range_batched<PEER, 4>(a, begin, end, outs, n_out, mc);
  if (a.n_src == 2) return ws_range_batched<PEER, 2>(a, begin, end, outs, n_out, mc);
  for (size_t i = begin + ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 2; i < end; i += stride) {
    const int limb = (int)((i >> a.log_n) % a.n_limbs);  // both words of the pair share a row (N is even)
    const uint64_t q = a.q[limb];
    uint64_t acc0 = 0, acc1 = 0;
    int l = 0;
    for (; l + 4 <= a.n_src; l += 4) {  // batch the loads for memory-level parallelism
      uint4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = PEER ? ld_peer(a.src[l + j] + i) : ld_stream(a.src[l + j] + i);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint64_t x0 = ((uint64_t)v[j].y << 32) | v[j].x, x1 = ((uint64_t)v[j].w << 32) | v[j].z;
        acc0 = d_add(acc0, d_mul_shoup(x0, a.w[l + j][limb], a.wp[l + j][limb], q), q);
        acc1 = d_add(acc1, d_mul_shoup(x1, a.w[l + j][limb], a.wp[l + j][limb], q), q);
      }
    }
    for (; l < a.n_src; ++l) {
      const uint4 v = PEER ? ld_peer(a.src[l] + i) : ld_stream(a.src[l] + i);
      const uint64_t x0 = ((uint64_t)v.y << 32) | v.x, x1 = ((uint64_t)v.w << 32) | v.z;
      acc0 = d_add(acc0, d_mul_shoup(x0, a.w[l][limb], a.wp[l][limb], q), q);
      acc1 = d_add(acc1, d_mul_shoup(x1, a.w[l][limb], a.wp[l][limb], q), q);
    }
    const uint4 o = make_uint4((uint32_t)acc0, (uint32_t)(acc0 >> 32), (uint32_t)acc1, (uint32_t)(acc1 >> 32));
    ws_store(outs, n_out, mc, i, o);
  }
}

__global__ void __launch_bounds__(256) ckks_weighted_sum_kernel(const __grid_constant__ WsArgs a, size_t words, uint64_t* out) {
  uint64_t* outs[1] = {out};
  ws_range<false>(a, 0, words, outs, 1);
}

struct FusedWsArgs {
  WsArgs ws;
  uint64_t* out[MFL_MAX_PEERS];
  uint32_t* pads[MFL_MAX_PEERS];
  uint64_t* mc_out;   // multicast address of the output buffers, or NULL (peer stores)
  int rank, world;
  size_t words;
};

// ---- K8 fed by bulk copies --------------------------------------------------------------------------------------
// 16-byte peer loads issued by threads top out near 300 GB/s per GPU over NVLink here (2 and 8 GPUs alike): the link
// wants few, large requests.  This variant moves the operands with cp.async.bulk (no multicast: each copy reads one
// GPU's memory into this CTA's shared memory): one producer lane streams 4 KB pieces of every source's shard into a
// ring of stages (up to 192 KB in flight per SM), 256 consumer threads take one 16-byte element pair of every source
// per stage from shared memory, multiply-accumulate modulo q and store the result once (multimem.st) or to every peer.
constexpr int kWsChunkWords = 512;                 // 4 KB of one source per stage: 256 threads x 16 B
constexpr int kWsConsumers = 256;
constexpr int kWsThreads = kWsConsumers + 32;      // + the producer warp
constexpr int kWsMaxSrc = 8;
constexpr size_t kWsSmemBudget = 196 * 1024;

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(gemm::smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(gemm::smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(kWsThreads, 1) ckks_fused_weighted_sum_bulk_kernel(const __grid_constant__ FusedWsArgs fa,
                                                                                     int stages) {
  extern __shared__ __align__(128) uint8_t ws_smem[];
  using namespace gemm;
  const int ns = fa.ws.n_src;
  const uint32_t stage_bytes = (uint32_t)ns * kWsChunkWords * 8;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ws_smem + (size_t)stages * stage_bytes);
  uint64_t* empty_bar = full_bar + stages;
  const int slot0 = (0 * gridDim.x + blockIdx.x) * MFL_MAX_PEERS, slot1 = (1 * gridDim.x + blockIdx.x) * MFL_MAX_PEERS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, kWsConsumers / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  peer_barrier(fa.pads, fa.rank, fa.world, slot0, true);   // every rank's ciphertexts are complete (also a CTA barrier)
  // shards in whole chunks (a chunk never straddles a limb row: rows are 2^log_n words, chunks 2^9)
  const size_t chunks = fa.words / kWsChunkWords, per = (chunks + fa.world - 1) / fa.world;
  const size_t c_begin = min(chunks, per * fa.rank), c_end = min(chunks, per * (fa.rank + 1));
  if (warp == kWsConsumers / 32) {
    if (lane == 0) {  // ===== producer =====
      uint32_t it = 0;
      for (size_t c = c_begin + blockIdx.x; c < c_end; c += gridDim.x, ++it) {
        const int s = it % stages;
        mbar_wait(empty_bar + s, ((it / stages) & 1) ^ 1);
        mbar_expect_tx(full_bar + s, stage_bytes);
        uint8_t* dst = ws_smem + (size_t)s * stage_bytes;
        for (int j = 0; j < ns; ++j)
          bulk_load(dst + (size_t)j * kWsChunkWords * 8, fa.ws.src[j] + c * kWsChunkWords, kWsChunkWords * 8, full_bar + s);
      }
    }
  } else {  // ===== consumers =====
    uint32_t it = 0;
    for (size_t c = c_begin + blockIdx.x; c < c_end; c += gridDim.x, ++it) {
      const int s = it % stages;
      const size_t i = c * kWsChunkWords + (size_t)threadIdx.x * 2;
      const int limb = (int)(((c * kWsChunkWords) >> fa.ws.log_n) % fa.ws.n_limbs);
      const uint64_t q = fa.ws.q[limb];
      mbar_wait(full_bar + s, (it / stages) & 1);
      const uint4* src = reinterpret_cast<const uint4*>(ws_smem + (size_t)s * stage_bytes) + threadIdx.x;
      uint64_t acc0 = 0, acc1 = 0;
#pragma unroll
      for (int j = 0; j < kWsMaxSrc; ++j) {
        if (j < ns) {
          const uint4 v = src[(size_t)j * (kWsChunkWords / 2)];
          const uint64_t x0 = ((uint64_t)v.y << 32) | v.x, x1 = ((uint64_t)v.w << 32) | v.z;
          acc0 = d_add(acc0, d_mul_shoup(x0, fa.ws.w[j][limb], fa.ws.wp[j][limb], q), q);
          acc1 = d_add(acc1, d_mul_shoup(x1, fa.ws.w[j][limb], fa.ws.wp[j][limb], q), q);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar + s);   // this warp has read its part of the stage
      const uint4 o = make_uint4((uint32_t)acc0, (uint32_t)(acc0 >> 32), (uint32_t)acc1, (uint32_t)(acc1 >> 32));
      ws_store(fa.out, fa.world, fa.mc_out, i, o);
    }
  }
  // words past the last whole chunk (none for ring dimensions >= 512) would be handled here
  peer_barrier(fa.pads, fa.rank, fa.world, slot1, true);   // every shard has landed everywhere
}

__global__ void __launch_bounds__(512, 1) ckks_fused_weighted_sum_kernel(const __grid_constant__ FusedWsArgs fa) {
  const int slot0 = (0 * gridDim.x + blockIdx.x) * MFL_MAX_PEERS, slot1 = (1 * gridDim.x + blockIdx.x) * MFL_MAX_PEERS;
  peer_barrier(fa.pads, fa.rank, fa.world, slot0, true);   // every rank's ciphertexts are complete
  const size_t pairs = fa.words / 2, per = (pairs + fa.world - 1) / fa.world;
  const size_t begin = min(pairs, per * fa.rank) * 2, end = min(pairs, per * (fa.rank + 1)) * 2;
  ws_range<true>(fa.ws, begin, end, fa.out, fa.world, fa.mc_out);
  peer_barrier(fa.pads, fa.rank, fa.world, slot1, true);   // every shard has landed everywhere
}

// ---- K9: rescale (drop the last limb) -----------------------------------------------------------
struct RescaleArgs { uint64_t qinv[4], qinv_shoup[4]; };

__global__ void __launch_bounds__(kNttThreads, 1) rescale_kernel(const __grid_constant__ Ctx cx, const uint64_t* __restrict__ ct_in,
                                                                 uint32_t n_polys, const __grid_constant__ RescaleArgs ra,
                                                                 uint64_t* ct_out) {
  extern __shared__ __align__(16) uint64_t sm[];
  const int L = cx.c.n_limbs, log_n = cx.c.log_n, Lo = L - 1;
  const uint32_t N = 1u << log_n;
  const uint64_t ql = cx.c.q[L - 1];
  // one CTA per (polynomial, remaining limb): the dropped limb is brought back to coefficient form,
  // re-reduced modulo q_l around zero, transformed again and subtracted
  for (uint32_t job = blockIdx.x; job < n_polys * (uint32_t)Lo; job += gridDim.x) {
    const uint32_t p = job / Lo, l = job % Lo;
    const uint64_t q = cx.c.q[l];
    const uint64_t* last = ct_in + ((size_t)p * L + (L - 1)) * N;
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) sm[px(i)] = last[i];
    __syncthreads();
    ntt_inverse_smem(sm, cx.c.ipsi[L - 1], cx.c.ipsi_shoup[L - 1], ql, cx.c.n_inv[L - 1], cx.c.n_inv_shoup[L - 1], log_n);
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) {
      const uint64_t c = sm[px(i)];
      sm[px(i)] = c > ql / 2 ? d_sub(c % q, ql % q, q) : c % q;
    }
    __syncthreads();
    ntt_forward_smem(sm, cx.c.psi[l], cx.c.psi_shoup[l], q, log_n);
    const uint64_t* cin = ct_in + ((size_t)p * L + l) * N;
    uint64_t* co = ct_out + ((size_t)p * Lo + l) * N;
    for (uint32_t i = threadIdx.x; i < N; i += blockDim.x)
      co[i] = d_mul_shoup(d_sub(cin[i], sm[px(i)], q), ra.qinv[l], ra.qinv_shoup[l], q);
    __syncthreads();
  }
}

template <typename K> static int allow_smem(K kernel, size_t bytes) {
  return check(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes), "cudaFuncSetAttribute(smem)");
}
static int check_ctx(const mfl_ckks_ctx* c) {
  MFL_ARG(c && c->log_n >= 10 && c->log_n <= 14 && c->n_limbs >= 1 && c->n_limbs <= 4,
          "ckks: context must have 2^10..2^14 coefficients (one limb must fit shared memory) and 1..4 limbs");
  return 0;
}
static int fill_ws(WsArgs& a, const mfl_ckks_ctx* ctx, const uint64_t* const* cts, const uint64_t* w,
                   const uint64_t* w_shoup, int n_src) {
  MFL_ARG(n_src >= 1 && n_src <= MFL_MAX_SRC, "ckks weighted sum: n_src=%d outside [1,%d]", n_src, MFL_MAX_SRC);
  a.n_src = n_src; a.n_limbs = ctx->n_limbs; a.log_n = ctx->log_n;
  for (int i = 0; i < 4; ++i) a.q[i] = ctx->q[i];
  for (int l = 0; l < n_src; ++l) {
    a.src[l] = cts[l];
    MFL_ARG((reinterpret_cast<uintptr_t>(cts[l]) & 15) == 0, "ckks weighted sum: source %d is not 16-byte aligned", l);
    for (int i = 0; i < 4; ++i) { a.w[l][i] = w[l * 4 + i]; a.wp[l][i] = w_shoup[l * 4 + i]; }
  }
  return 0;
}

}  // namespace ckksdev
}  // namespace mfl

using namespace mfl;
using namespace mfl::ckksdev;

extern "C" int mfl_ckks_ntt(const mfl_ckks_ctx* ctx, uint64_t* rows, size_t n_rows, int inverse, void* stream) {
  if (int rc = check_ctx(ctx)) return rc;
  if (n_rows == 0) return 0;
  const size_t smem = poly_smem_bytes(ctx->log_n);
  if (int rc = allow_smem(ntt_rows_kernel, smem)) return rc;
  Ctx cx{*ctx};
  const int blocks = (int)std::min<size_t>(n_rows, (size_t)sm_count());
  ntt_rows_kernel<<<blocks, kNttThreads, smem, (cudaStream_t)stream>>>(cx, rows, n_rows, inverse);
  return check(cudaGetLastError(), "ntt_rows_kernel launch");
}

extern "C" int mfl_ckks_encode(const mfl_ckks_ctx* ctx, const double* values, size_t n_values, uint32_t slots, double scale,
                               const double* ksi_pows, const uint32_t* rot_group, double* coeffs, void* stream) {
  if (int rc = check_ctx(ctx)) return rc;
  MFL_ARG(slots >= 2 && (slots & (slots - 1)) == 0 && slots <= (1u << ctx->log_n) / 2, "ckks_encode: bad slot count %u", slots);
  if (n_values == 0) return 0;
  const uint32_t n_ct = (uint32_t)((n_values + slots - 1) / slots);
  const size_t smem = (size_t)slots * 16;
  if (int rc = allow_smem(encode_kernel, smem)) return rc;
  encode_kernel<<<n_ct, 512, smem, (cudaStream_t)stream>>>(values, n_values, slots, scale, ksi_pows, rot_group, coeffs, ctx->log_n);
  return check(cudaGetLastError(), "encode_kernel launch");
}

extern "C" int mfl_ckks_encrypt(const mfl_ckks_ctx* ctx, const double* coeffs, uint32_t n_ct, uint32_t slots,
                                const uint64_t* pk, const uint64_t* pk_shoup, const uint32_t* key256, uint64_t* ct_out,
                                void* stream) {
  if (int rc = check_ctx(ctx)) return rc;
  MFL_ARG(key256 != nullptr, "ckks_encrypt: a 256-bit key (8 host words from the OS entropy source) is required");
  MFL_ARG(ctx->log_n >= 4 && ctx->log_n <= 19, "ckks_encrypt: ring dimension 2^%d outside the block addressing", ctx->log_n);
  if (n_ct == 0) return 0;
  ChaChaKey key;
  for (int i = 0; i < 8; ++i) key.k[i] = key256[i];
  const size_t smem = poly_smem_bytes(ctx->log_n);
  if (int rc = allow_smem(encrypt_kernel, smem)) return rc;
  Ctx cx{*ctx};
  const int blocks = (int)std::min<size_t>((size_t)n_ct * ctx->n_limbs, (size_t)sm_count());
  encrypt_kernel<<<blocks, kNttThreads, smem, (cudaStream_t)stream>>>(cx, coeffs, slots, pk, pk_shoup, key, ct_out, n_ct);
  return check(cudaGetLastError(), "encrypt_kernel launch");
}

extern "C" int mfl_ckks_decrypt(const mfl_ckks_ctx* ctx, const uint64_t* ct, uint32_t n_ct, const uint64_t* sk,
                                const uint64_t* sk_shoup, uint64_t* residues, void* stream) {
  if (int rc = check_ctx(ctx)) return rc;
  if (n_ct == 0) return 0;
  const size_t smem = poly_smem_bytes(ctx->log_n);
  if (int rc = allow_smem(decrypt_kernel, smem)) return rc;
  Ctx cx{*ctx};
  const int blocks = (int)std::min<size_t>((size_t)n_ct * ctx->n_limbs, (size_t)sm_count());
  decrypt_kernel<<<blocks, kNttThreads, smem, (cudaStream_t)stream>>>(cx, ct, n_ct, sk, sk_shoup, residues);
  return check(cudaGetLastError(), "decrypt_kernel launch");
}

extern "C" int mfl_ckks_decode(const mfl_ckks_ctx* ctx, const uint64_t* residues, uint32_t n_ct, uint32_t slots,
                               double inv_scale, const uint64_t* garner_inv, const double* ksi_pows,
                               const uint32_t* rot_group, double* values, size_t n_values, void* stream) {
  if (int rc = check_ctx(ctx)) return rc;
  if (n_ct == 0) return 0;
  const size_t smem = (size_t)slots * 16;
  if (int rc = allow_smem(decode_kernel, smem)) return rc;
  Ctx cx{*ctx};
  GarnerInv gi;
  for (int i = 0; i < 16; ++i) gi.inv[i] = garner_inv[i];
  decode_kernel<<<n_ct, 512, smem, (cudaStream_t)stream>>>(cx, residues, slots, inv_scale, gi, ksi_pows, rot_group, values, n_values);
  return check(cudaGetLastError(), "decode_kernel launch");
}

extern "C" int mfl_ckks_weighted_sum(const mfl_ckks_ctx* ctx, const uint64_t* const* cts, const uint64_t* w,
                                     const uint64_t* w_shoup, int n_src, size_t words, uint64_t* out, void* stream) {
  if (int rc = check_ctx(ctx)) return rc;
  WsArgs a;
  if (int rc = fill_ws(a, ctx, cts, w, w_shoup, n_src)) return rc;
  if (words == 0) return 0;
  const int blocks = (int)std::max<size_t>(1, std::min<size_t>((words / 2 + 255) / 256, (size_t)sm_count() * 8));
  ckks_weighted_sum_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a, words, out);
  return check(cudaGetLastError(), "ckks_weighted_sum_kernel launch");
}

extern "C" int mfl_ckks_fused_weighted_sum_mc(const mfl_ckks_ctx* ctx, const uint64_t* const* cts, const uint64_t* w,
                                              const uint64_t* w_shoup, int n_src, size_t words, uint64_t* const* outs,
                                              uint64_t* mc_out, uint32_t* const* signal_pads, int rank, int world,
                                              int n_blocks, void* stream) {
  if (int rc = check_ctx(ctx)) return rc;
  MFL_ARG(world >= 1 && world <= MFL_MAX_PEERS && rank >= 0 && rank < world, "ckks fused: bad rank/world %d/%d", rank, world);
  MFL_ARG(n_blocks >= 1 && n_blocks <= sm_count(), "ckks fused: n_blocks=%d must be in [1, #SMs]", n_blocks);
  MFL_ARG((reinterpret_cast<uintptr_t>(mc_out) & 15) == 0, "ckks fused: the multicast output address is not 16-byte aligned");
  FusedWsArgs fa;
  if (int rc = fill_ws(fa.ws, ctx, cts, w, w_shoup, n_src)) return rc;
  for (int t = 0; t < world; ++t) { fa.out[t] = outs[t]; fa.pads[t] = signal_pads[t]; }
  fa.mc_out = mc_out;
  fa.rank = rank; fa.world = world; fa.words = words;
  // MFL_CKKS_BULK=1 selects the cp.async.bulk-fed variant; both variants are bound by the inbound link and measured
  // the same (5.85-6.0 ms for ResNet-18 on 8 GPUs), so the thread-load kernel stays the default
  static const int bulk_mode = [] { const char* e = getenv("MFL_CKKS_BULK"); return e ? atoi(e) : 0; }();
  if (bulk_mode && world > 1 && n_src <= kWsMaxSrc && words % kWsChunkWords == 0 && ctx->log_n >= 9) {
    const size_t stage_bytes = (size_t)n_src * kWsChunkWords * 8;
    const int stages = (int)std::min<size_t>(16, (kWsSmemBudget - 512) / stage_bytes);
    const size_t smem = (size_t)stages * stage_bytes + 2 * (size_t)stages * sizeof(uint64_t);
    if (int rc = allow_smem(ckks_fused_weighted_sum_bulk_kernel, smem)) return rc;
    ckks_fused_weighted_sum_bulk_kernel<<<n_blocks, kWsThreads, smem, (cudaStream_t)stream>>>(fa, stages);
    return check(cudaGetLastError(), "ckks_fused_weighted_sum_bulk_kernel launch");
  }
  ckks_fused_weighted_sum_kernel<<<n_blocks, 512, 0, (cudaStream_t)stream>>>(fa);
  return check(cudaGetLastError(), "ckks_fused_weighted_sum_kernel launch");
}

extern "C" int mfl_ckks_fused_weighted_sum(const mfl_ckks_ctx* ctx, const uint64_t* const* cts, const uint64_t* w,
                                           const uint64_t* w_shoup, int n_src, size_t words, uint64_t* const* outs,
                                           uint32_t* const* signal_pads, int rank, int world, int n_blocks, void* stream) {
  return mfl_ckks_fused_weighted_sum_mc(ctx, cts, w, w_shoup, n_src, words, outs, nullptr, signal_pads, rank, world,
                                        n_blocks, stream);
}

extern "C" int mfl_ckks_rescale(const mfl_ckks_ctx* ctx, const uint64_t* ct_in, uint32_t n_ct, const uint64_t* qlast_inv,
                                const uint64_t* qlast_inv_shoup, uint64_t* ct_out, void* stream) {
  if (int rc = check_ctx(ctx)) return rc;
  MFL_ARG(ctx->n_limbs >= 2, "ckks_rescale: no limb left to drop");
  if (n_ct == 0) return 0;
  const size_t smem = poly_smem_bytes(ctx->log_n);
  if (int rc = allow_smem(rescale_kernel, smem)) return rc;
  Ctx cx{*ctx};
  RescaleArgs ra{};
  for (int i = 0; i < ctx->n_limbs - 1; ++i) { ra.qinv[i] = qlast_inv[i]; ra.qinv_shoup[i] = qlast_inv_shoup[i]; }
  const uint32_t n_polys = n_ct * 2;
  const int blocks = (int)std::min<size_t>((size_t)n_polys * (ctx->n_limbs - 1), (size_t)sm_count());
  rescale_kernel<<<blocks, kNttThreads, smem, (cudaStream_t)stream>>>(cx, ct_in, n_polys, ra, ct_out);
  return check(cudaGetLastError(), "rescale_kernel launch");
}
