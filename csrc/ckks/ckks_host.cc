#include "ckks/ckks_host.h"

#include <omp.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <random>
#include <sstream>
#include <sys/random.h>
#include <stdexcept>

namespace mfl::ckks {

namespace {

// ChaCha20 keystream generator (RFC 8439 block function).  All secret material of the scheme - the ternary secret,
// the Gaussian errors, the encryption randomness u, e0, e1 - comes from a stream keyed with 256 bits of operating
// system entropy (getrandom); the PUBLIC uniform polynomial a of a key pair is drawn from a second, independently
// keyed stream, so nothing an observer of the public key sees is an output of the stream that produced the secret.
// A caller-supplied 64-bit seed (tests, reproducible demos) expands to a key deterministically and is NOT secure.
struct Rng {
  uint32_t key[8], nonce[3], counter = 0;
  uint32_t block[16];
  int used = 16;
  static uint32_t rotl(uint32_t x, int k) { return (x << k) | (x >> (32 - k)); }
  static void quarter(uint32_t* x, int a, int b, int c, int d) {
    x[a] += x[b]; x[d] = rotl(x[d] ^ x[a], 16);
    x[c] += x[d]; x[b] = rotl(x[b] ^ x[c], 12);
    x[a] += x[b]; x[d] = rotl(x[d] ^ x[a], 8);
    x[c] += x[d]; x[b] = rotl(x[b] ^ x[c], 7);
  }
  void refill() {
    uint32_t st[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u};
    for (int i = 0; i < 8; ++i) st[4 + i] = key[i];
    st[12] = counter++;
    for (int i = 0; i < 3; ++i) st[13 + i] = nonce[i];
    uint32_t x[16];
    std::memcpy(x, st, sizeof(x));
    for (int r = 0; r < 10; ++r) {
      quarter(x, 0, 4, 8, 12); quarter(x, 1, 5, 9, 13); quarter(x, 2, 6, 10, 14); quarter(x, 3, 7, 11, 15);
      quarter(x, 0, 5, 10, 15); quarter(x, 1, 6, 11, 12); quarter(x, 2, 7, 8, 13); quarter(x, 3, 4, 9, 14);
    }
    for (int i = 0; i < 16; ++i) block[i] = x[i] + st[i];
    used = 0;
    if (counter == 0) ++nonce[0];  // 256 GiB per nonce value; roll over into the nonce
  }
  // seed == 0: fresh OS entropy.  `stream` separates independent generators derived from one seed.
  explicit Rng(uint64_t seed, uint64_t stream = 0) {
    if (seed == 0) {
      size_t got = 0;
      while (got < sizeof(key)) {
        const ssize_t n = getrandom(reinterpret_cast<char*>(key) + got, sizeof(key) - got, 0);
        if (n <= 0) throw std::runtime_error("CKKS: the operating system entropy source is unavailable");
        got += (size_t)n;
      }
    } else {  // deterministic, for tests only: splitmix64 expansion of the seed
      uint64_t z0 = seed;
      for (int i = 0; i < 4; ++i) {
        z0 += 0x9e3779b97f4a7c15ull;
        uint64_t z = z0;
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
        z ^= z >> 31;
        key[2 * i] = (uint32_t)z; key[2 * i + 1] = (uint32_t)(z >> 32);
      }
    }
    nonce[0] = 0; nonce[1] = (uint32_t)stream; nonce[2] = (uint32_t)(stream >> 32);
  }
  uint64_t next() {
    if (used > 14) refill();
    const uint64_t r = (uint64_t)block[used] | ((uint64_t)block[used + 1] << 32);
    used += 2;
    return r;
  }
  double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
  int ternary() {  // unbiased: reject the one 64-bit value class that does not divide evenly by 3
    uint64_t x;
    do x = next(); while (x >= UINT64_MAX - UINT64_MAX % 3);
    return (int)(x % 3) - 1;
  }
  int gaussian(double sigma) {
    double u1 = uniform(), u2 = uniform();
    if (u1 < 1e-300) u1 = 1e-300;
    return (int)std::lrint(sigma * std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2));
  }
  uint64_t below(uint64_t q) {  // rejection sampling, unbiased
    const uint64_t lim = UINT64_MAX - UINT64_MAX % q;
    uint64_t x;
    do x = next(); while (x >= lim);
    return x % q;
  }
};

// 64 bits of OS entropy, used to key per-ciphertext streams of one Encrypt call
uint64_t EntropySeed() {
  uint64_t v = 0;
  if (getrandom(&v, sizeof(v), 0) != (ssize_t)sizeof(v)) throw std::runtime_error("CKKS: getrandom failed");
  return v | 1;
}

inline uint64_t SignedToMod(long long v, uint64_t q) {
  if (v >= 0) return (uint64_t)v % q;
  const uint64_t r = (uint64_t)(-v) % q;
  return r ? q - r : 0;
}

inline uint64_t DoubleToMod(double c, uint64_t q) {
  // c is integral (already rounded) and may exceed 2^63: go through __int128
  const __int128 x = (__int128)c;
  __int128 r = x % (__int128)q;
  if (r < 0) r += q;
  return (uint64_t)r;
}

constexpr double kSigma = 3.2;
constexpr uint32_t kKeyMagicLen = 4;

struct KeyHeader {
  char magic[4];
  uint16_t version;
  uint8_t log_n, n_limbs;
  uint32_t n_polys, reserved;
  uint64_t q[kMaxLimbs];
};

std::string PackKey(const char* magic, const Params& p, uint32_t n_polys, const std::vector<uint64_t>& data) {
  KeyHeader h{};
  std::memcpy(h.magic, magic, 4);
  h.version = 1; h.log_n = (uint8_t)p.log_n; h.n_limbs = (uint8_t)p.q.size(); h.n_polys = n_polys;
  for (size_t i = 0; i < p.q.size(); ++i) h.q[i] = p.q[i];
  std::string out(sizeof(h) + data.size() * 8, '\0');
  std::memcpy(out.data(), &h, sizeof(h));
  std::memcpy(out.data() + sizeof(h), data.data(), data.size() * 8);
  return out;
}

bool UnpackKey(const char* magic, const Params& p, std::string_view blob, uint32_t n_polys, std::vector<uint64_t>& out) {
  KeyHeader h;
  if (blob.size() < sizeof(h)) return false;
  std::memcpy(&h, blob.data(), sizeof(h));
  if (std::memcmp(h.magic, magic, 4) != 0 || h.log_n != p.log_n || h.n_limbs != p.q.size() || h.n_polys != n_polys) return false;
  const size_t words = (size_t)n_polys * p.q.size() * p.n();
  if (blob.size() != sizeof(h) + words * 8) return false;
  out.resize(words);
  std::memcpy(out.data(), blob.data() + sizeof(h), words * 8);
  return true;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
bool SaveParamsFile(const std::string& path, const Params& p) {
  std::ofstream f(path, std::ios::binary);
  if (!f) return false;
  f << "MFCX 1\nlog_n " << p.log_n << "\nscale_bits " << p.scale_bits << "\nbatch_size " << p.batch_size << "\nmoduli";
  for (auto q : p.q) f << " " << q;
  f << "\n";
  return (bool)f;
}

bool LoadParamsFile(const std::string& path, Params& p) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::string magic, key;
  int ver;
  if (!(f >> magic >> ver) || magic != "MFCX") return false;
  while (f >> key) {
    if (key == "log_n") f >> p.log_n;
    else if (key == "scale_bits") f >> p.scale_bits;
    else if (key == "batch_size") f >> p.batch_size;
    else if (key == "moduli") {
      p.q.clear();
      std::string line;
      std::getline(f, line);
      std::istringstream ls(line);
      uint64_t q;
      while (ls >> q) p.q.push_back(q);
    }
  }
  return !p.q.empty() && p.q.size() <= kMaxLimbs;
}

bool ParseHeader(std::string_view blob, CtHeader& h) {
  if (blob.size() < sizeof(CtHeader)) return false;
  std::memcpy(&h, blob.data(), sizeof(h));
  if (std::memcmp(h.magic, "MFCK", 4) != 0 || h.version != 1 || h.n_limbs == 0 || h.n_limbs > kMaxLimbs) return false;
  // the blob comes off the wire (a learner's ciphertext): every field later code divides by, shifts by or indexes
  // with is bounded here - ring dimension, odd moduli below 2^62, a finite positive scale, slots within the ring
  if (h.log_n < 1 || h.log_n > 17) return false;
  for (uint32_t i = 0; i < h.n_limbs; ++i)
    if (h.q[i] < 3 || (h.q[i] & 1) == 0 || h.q[i] >= (1ull << 62)) return false;
  if (!(h.scale > 0.0) || !std::isfinite(h.scale)) return false;
  if (h.slots == 0 || h.slots > (1u << h.log_n) / 2) return false;
  return blob.size() == sizeof(CtHeader) + (size_t)h.n_ct * ct_words(h) * 8;
}

// ---------------------------------------------------------------------------------------------
NttTables MakeNttTables(uint64_t q, uint32_t log_n) {
  const uint64_t n = 1ull << log_n;
  if ((q - 1) % (2 * n) != 0) throw std::invalid_argument("modulus is not 1 mod 2N");
  NttTables t;
  t.q = q;
  for (uint64_t x = 2;; ++x) {
    const uint64_t c = pow_mod(x, (q - 1) / (2 * n), q);
    if (pow_mod(c, n, q) == q - 1) { t.psi = c; break; }
  }
  const uint64_t ipsi = inv_mod(t.psi, q);
  t.n_inv = inv_mod(n % q, q);
  t.n_inv_shoup = shoup_const(t.n_inv, q);
  t.psi_br.resize(n); t.ipsi_br.resize(n); t.psi_br_shoup.resize(n); t.ipsi_br_shoup.resize(n);
  uint64_t p = 1, ip = 1;
  for (uint64_t i = 0; i < n; ++i) {
    const uint32_t r = bit_reverse((uint32_t)i, log_n);
    t.psi_br[r] = p; t.ipsi_br[r] = ip;
    p = mul_mod(p, t.psi, q);
    ip = mul_mod(ip, ipsi, q);
  }
  for (uint64_t i = 0; i < n; ++i) {
    t.psi_br_shoup[i] = shoup_const(t.psi_br[i], q);
    t.ipsi_br_shoup[i] = shoup_const(t.ipsi_br[i], q);
  }
  return t;
}

void NttForward(uint64_t* a, const NttTables& tb, uint32_t log_n) {
  const uint64_t n = 1ull << log_n, q = tb.q;
  uint64_t t = n;
  for (uint64_t m = 1; m < n; m <<= 1) {
    t >>= 1;
    for (uint64_t i = 0; i < m; ++i) {
      const uint64_t w = tb.psi_br[m + i], wp = tb.psi_br_shoup[m + i];
      uint64_t* x = a + 2 * i * t;
      for (uint64_t j = 0; j < t; ++j) {
        const uint64_t u = x[j], v = mul_shoup(x[j + t], w, wp, q);
        x[j] = add_mod(u, v, q);
        x[j + t] = sub_mod(u, v, q);
      }
    }
  }
}

void NttInverse(uint64_t* a, const NttTables& tb, uint32_t log_n) {
  const uint64_t n = 1ull << log_n, q = tb.q;
  uint64_t t = 1;
  for (uint64_t m = n; m > 1; m >>= 1) {
    const uint64_t h = m >> 1;
    uint64_t j1 = 0;
    for (uint64_t i = 0; i < h; ++i) {
      const uint64_t w = tb.ipsi_br[h + i], wp = tb.ipsi_br_shoup[h + i];
      for (uint64_t j = j1; j < j1 + t; ++j) {
        const uint64_t u = a[j], v = a[j + t];
        a[j] = add_mod(u, v, q);
        a[j + t] = mul_shoup(sub_mod(u, v, q), w, wp, q);
      }
      j1 += 2 * t;
    }
    t <<= 1;
  }
  for (uint64_t j = 0; j < n; ++j) a[j] = mul_shoup(a[j], tb.n_inv, tb.n_inv_shoup, q);
}

// ---------------------------------------------------------------------------------------------
Scheme::Scheme(const Params& p) : p_(p) {
  if (p.q.empty() || p.q.size() > kMaxLimbs) throw std::invalid_argument("1..4 RNS moduli supported");
  if (p.batch_size == 0 || (p.batch_size & (p.batch_size - 1)) || p.batch_size > p.n() / 2)
    throw std::invalid_argument("batch_size must be a power of two <= N/2");
  for (auto q : p.q) tables_.push_back(MakeNttTables(q, p.log_n));
  const uint32_t n = p.n(), m = 2 * n;
  rot_group_.resize(n / 2);
  uint64_t five = 1;
  for (uint32_t j = 0; j < n / 2; ++j) {
    rot_group_[j] = (uint32_t)five;
    five = five * 5 % m;
  }
  ksi_pows_.resize(m + 1);
  for (uint32_t k = 0; k <= m; ++k) {
    const double ang = 2.0 * M_PI * (double)k / (double)m;
    ksi_pows_[k] = {std::cos(ang), std::sin(ang)};
  }
}

static void BitReverseArray(std::complex<double>* v, uint32_t size) {
  for (uint32_t i = 1, j = 0; i < size; ++i) {
    uint32_t bit = size >> 1;
    for (; j >= bit; bit >>= 1) j -= bit;
    j += bit;
    if (i < j) std::swap(v[i], v[j]);
  }
}

void Scheme::FftSpecial(std::complex<double>* v, uint32_t size) const {
  const uint32_t m = 2 * p_.n();
  BitReverseArray(v, size);
  for (uint32_t len = 2; len <= size; len <<= 1) {
    const uint32_t lenh = len >> 1, lenq = len << 2;
    for (uint32_t i = 0; i < size; i += len)
      for (uint32_t j = 0; j < lenh; ++j) {
        const uint32_t idx = (rot_group_[j] % lenq) * (m / lenq);
        const auto u = v[i + j], w = v[i + j + lenh] * ksi_pows_[idx];
        v[i + j] = u + w;
        v[i + j + lenh] = u - w;
      }
  }
}

void Scheme::FftSpecialInv(std::complex<double>* v, uint32_t size) const {
  const uint32_t m = 2 * p_.n();
  for (uint32_t len = size; len >= 2; len >>= 1) {
    const uint32_t lenh = len >> 1, lenq = len << 2;
    for (uint32_t i = 0; i < size; i += len)
      for (uint32_t j = 0; j < lenh; ++j) {
        const uint32_t idx = (lenq - (rot_group_[j] % lenq)) * (m / lenq);
        const auto u = v[i + j] + v[i + j + lenh];
        const auto w = (v[i + j] - v[i + j + lenh]) * ksi_pows_[idx];
        v[i + j] = u;
        v[i + j + lenh] = w;
      }
  }
  BitReverseArray(v, size);
  const double inv = 1.0 / (double)size;
  for (uint32_t i = 0; i < size; ++i) v[i] *= inv;
}

void Scheme::EncodeCoeffs(const double* vals, size_t n, double scale, std::vector<double>& coeffs) const {
  const uint32_t slots = p_.batch_size, N = p_.n(), gap = (N / 2) / slots;
  std::vector<std::complex<double>> u(slots);
  for (uint32_t i = 0; i < slots; ++i) u[i] = {i < n ? vals[i] : 0.0, 0.0};
  FftSpecialInv(u.data(), slots);
  coeffs.assign(N, 0.0);
  for (uint32_t i = 0; i < slots; ++i) {
    coeffs[(size_t)i * gap] = std::nearbyint(u[i].real() * scale);
    coeffs[(size_t)i * gap + N / 2] = std::nearbyint(u[i].imag() * scale);
  }
}

void Scheme::DecodeCoeffs(const std::vector<double>& coeffs, size_t n, double* vals) const {
  const uint32_t slots = p_.batch_size, N = p_.n(), gap = (N / 2) / slots;
  std::vector<std::complex<double>> u(slots);
  for (uint32_t i = 0; i < slots; ++i) u[i] = {coeffs[(size_t)i * gap], coeffs[(size_t)i * gap + N / 2]};
  FftSpecial(u.data(), slots);
  for (uint32_t i = 0; i < slots && i < n; ++i) vals[i] = u[i].real();
}

// ---------------------------------------------------------------------------------------------
void Scheme::KeyGen(uint64_t seed) {
  const uint32_t N = p_.n(), L = (uint32_t)p_.q.size();
  Rng rng(seed, /*stream=*/1);      // secret material: s and the errors
  Rng pub(seed, /*stream=*/2);      // the public uniform polynomials (independently keyed when seed == 0)
  std::vector<int> s(N), e(N);
  for (auto& x : s) x = rng.ternary();
  sk_.assign((size_t)L * N, 0);
  for (uint32_t l = 0; l < L; ++l) {
    for (uint32_t i = 0; i < N; ++i) sk_[(size_t)l * N + i] = SignedToMod(s[i], p_.q[l]);
    NttForward(&sk_[(size_t)l * N], tables_[l], p_.log_n);
  }
  auto make_pair = [&](uint64_t* b, uint64_t* a, int s2_limb) {
    for (auto& x : e) x = rng.gaussian(kSigma);
    for (uint32_t l = 0; l < L; ++l) {
      const uint64_t q = p_.q[l];
      uint64_t* al = a + (size_t)l * N;
      uint64_t* bl = b + (size_t)l * N;
      for (uint32_t i = 0; i < N; ++i) { al[i] = pub.below(q); bl[i] = SignedToMod(e[i], q); }
      NttForward(bl, tables_[l], p_.log_n);  // e in evaluation form; a is sampled directly in it
      const uint64_t* sl = &sk_[(size_t)l * N];
      for (uint32_t i = 0; i < N; ++i) {
        uint64_t v = sub_mod(bl[i], mul_mod(al[i], sl[i], q), q);  // -a*s + e
        if ((int)l == s2_limb) v = add_mod(v, mul_mod(sl[i], sl[i], q), q);  // + s^2 on its own digit
        bl[i] = v;
      }
    }
  };
  pk_.assign((size_t)2 * L * N, 0);
  make_pair(&pk_[0], &pk_[(size_t)L * N], -1);
  // relinearisation key (RNS digit decomposition); generated for file parity with the reference's
  // key-eval-mult.txt - weighted averaging never needs it
  evk_.assign((size_t)L * 2 * L * N, 0);
  for (uint32_t j = 0; j < L; ++j) make_pair(&evk_[(size_t)j * 2 * L * N], &evk_[(size_t)j * 2 * L * N + (size_t)L * N], (int)j);
}

std::string Scheme::SerializePublicKey() const { return PackKey("MFPK", p_, 2, pk_); }
std::string Scheme::SerializeSecretKey() const { return PackKey("MFSK", p_, 1, sk_); }
std::string Scheme::SerializeEvalMultKey() const { return PackKey("MFEK", p_, 2 * (uint32_t)p_.q.size(), evk_); }
bool Scheme::LoadPublicKey(std::string_view b) { return UnpackKey("MFPK", p_, b, 2, pk_); }
bool Scheme::LoadSecretKey(std::string_view b) { return UnpackKey("MFSK", p_, b, 1, sk_); }

// ---------------------------------------------------------------------------------------------
std::string Scheme::Encrypt(const double* vals, size_t n, uint64_t seed) const {
  if (pk_.empty()) throw std::runtime_error("Encrypt: public key not loaded");
  const uint32_t N = p_.n(), L = (uint32_t)p_.q.size(), slots = p_.batch_size;
  const uint32_t n_ct = (uint32_t)((n + slots - 1) / slots);
  CtHeader h{};
  std::memcpy(h.magic, "MFCK", 4);
  h.version = 1; h.log_n = (uint8_t)p_.log_n; h.n_limbs = (uint8_t)L; h.n_ct = n_ct; h.slots = slots;
  h.n_values = n; h.scale = std::ldexp(1.0, (int)p_.scale_bits);
  for (uint32_t l = 0; l < L; ++l) h.q[l] = p_.q[l];
  std::string out(sizeof(h) + (size_t)n_ct * ct_words(h) * 8, '\0');
  std::memcpy(out.data(), &h, sizeof(h));
  uint64_t* data = reinterpret_cast<uint64_t*>(out.data() + sizeof(h));
  // one 256-bit key per call (OS entropy unless a test seed is given), one ChaCha20 stream per ciphertext
  const Rng master(seed, 0);
#pragma omp parallel for schedule(dynamic)
  for (long c = 0; c < (long)n_ct; ++c) {
    Rng rng = master;
    rng.nonce[1] = (uint32_t)(c + 1); rng.nonce[2] = (uint32_t)((uint64_t)(c + 1) >> 32); rng.counter = 0; rng.used = 16;
    std::vector<double> m;
    const size_t off = (size_t)c * slots;
    EncodeCoeffs(vals + off, std::min<size_t>(slots, n - off), h.scale, m);
    std::vector<int> u(N), e0(N), e1(N);
    for (uint32_t i = 0; i < N; ++i) { u[i] = rng.ternary(); e0[i] = rng.gaussian(kSigma); e1[i] = rng.gaussian(kSigma); }
    uint64_t* c0 = data + (size_t)c * ct_words(h);
    uint64_t* c1 = c0 + ct_poly_words(h);
    std::vector<uint64_t> ul(N), t(N);
    for (uint32_t l = 0; l < L; ++l) {
      const uint64_t q = p_.q[l];
      const uint64_t* pk0 = &pk_[(size_t)l * N];
      const uint64_t* pk1 = &pk_[(size_t)(L + l) * N];
      for (uint32_t i = 0; i < N; ++i) ul[i] = SignedToMod(u[i], q);
      NttForward(ul.data(), tables_[l], p_.log_n);
      // c0 = pk0*u + NTT(e0 + m)
      for (uint32_t i = 0; i < N; ++i) t[i] = add_mod(SignedToMod(e0[i], q), DoubleToMod(m[i], q), q);
      NttForward(t.data(), tables_[l], p_.log_n);
      for (uint32_t i = 0; i < N; ++i) c0[(size_t)l * N + i] = add_mod(mul_mod(pk0[i], ul[i], q), t[i], q);
      // c1 = pk1*u + NTT(e1)
      for (uint32_t i = 0; i < N; ++i) t[i] = SignedToMod(e1[i], q);
      NttForward(t.data(), tables_[l], p_.log_n);
      for (uint32_t i = 0; i < N; ++i) c1[(size_t)l * N + i] = add_mod(mul_mod(pk1[i], ul[i], q), t[i], q);
    }
  }
  return out;
}

// Garner mixed-radix reconstruction of one coefficient from its residues -> centred double.
static double CrtToDouble(const uint64_t* res, const uint64_t* q, uint32_t L, const uint64_t inv[kMaxLimbs][kMaxLimbs]) {
  uint64_t v[kMaxLimbs];
  for (uint32_t i = 0; i < L; ++i) {
    uint64_t t = res[i] % q[i];
    for (uint32_t j = 0; j < i; ++j) t = mul_mod(sub_mod(t, v[j] % q[i], q[i]), inv[j][i], q[i]);
    v[i] = t;
  }
  // x = v0 + q0 (v1 + q1 (v2 + ...)); negative iff the top digit is in the upper half
  const bool neg = v[L - 1] > (q[L - 1] - 1) / 2;
  if (neg) {  // digits of Q - x
    bool borrow = true;  // Q - x = sum (q_i - 1 - v_i) * prod + 1
    for (uint32_t i = 0; i < L; ++i) {
      uint64_t d = q[i] - 1 - v[i];
      if (borrow) {
        d += 1;
        if (d == q[i]) d = 0; else borrow = false;
      }
      v[i] = d;
    }
  }
  long double acc = 0.0L;
  for (int i = (int)L - 1; i >= 0; --i) acc = acc * (long double)q[i] + (long double)v[i];
  return neg ? -(double)acc : (double)acc;
}

std::vector<double> Scheme::Decrypt(std::string_view blob, size_t n) const {
  if (sk_.empty()) throw std::runtime_error("Decrypt: private key not loaded");
  CtHeader h;
  if (!ParseHeader(blob, h)) throw std::runtime_error("Decrypt: malformed ciphertext blob");
  if (h.log_n != p_.log_n || h.slots != p_.batch_size) throw std::runtime_error("Decrypt: ciphertext/context mismatch");
  const uint32_t N = p_.n(), L = h.n_limbs;  // L may be smaller than the key's limb count after a rescale
  for (uint32_t l = 0; l < L; ++l)
    if (h.q[l] != p_.q[l]) throw std::runtime_error("Decrypt: modulus chain mismatch");
  uint64_t inv[kMaxLimbs][kMaxLimbs] = {};
  for (uint32_t j = 0; j < L; ++j)
    for (uint32_t i = j + 1; i < L; ++i) inv[j][i] = inv_mod(h.q[j] % h.q[i], h.q[i]);
  const uint64_t* data = reinterpret_cast<const uint64_t*>(blob.data() + sizeof(h));
  std::vector<double> out(n, 0.0);
  const size_t slots = h.slots;
#pragma omp parallel for schedule(dynamic)
  for (long c = 0; c < (long)h.n_ct; ++c) {
    const size_t off = (size_t)c * slots;
    if (off >= n) continue;
    const uint64_t* c0 = data + (size_t)c * ct_words(h);
    const uint64_t* c1 = c0 + ct_poly_words(h);
    std::vector<uint64_t> m((size_t)L * N);
    for (uint32_t l = 0; l < L; ++l) {
      const uint64_t q = h.q[l];
      const uint64_t* s = &sk_[(size_t)l * N];
      uint64_t* ml = &m[(size_t)l * N];
      for (uint32_t i = 0; i < N; ++i) ml[i] = add_mod(c0[(size_t)l * N + i], mul_mod(c1[(size_t)l * N + i], s[i], q), q);
      NttInverse(ml, tables_[l], p_.log_n);
    }
    std::vector<double> coeffs(N);
    const double inv_scale = 1.0 / h.scale;
    for (uint32_t i = 0; i < N; ++i) {
      uint64_t res[kMaxLimbs];
      for (uint32_t l = 0; l < L; ++l) res[l] = m[(size_t)l * N + i];
      coeffs[i] = CrtToDouble(res, h.q, L, inv) * inv_scale;
    }
    DecodeCoeffs(coeffs, std::min<size_t>(slots, n - off), &out[off]);
  }
  return out;
}

// ---------------------------------------------------------------------------------------------
std::string WeightedAverage(const std::vector<std::string_view>& cts, const std::vector<double>& weights) {
  if (cts.empty() || cts.size() != weights.size()) throw std::invalid_argument("WeightedAverage: bad arguments");
  CtHeader h0;
  if (!ParseHeader(cts[0], h0)) throw std::runtime_error("WeightedAverage: malformed ciphertext blob");
  const size_t words = (size_t)h0.n_ct * ct_words(h0);
  const size_t N = 1ull << h0.log_n;
  const double delta = h0.scale;  // weights are encoded at the ciphertexts' own scale
  std::vector<std::vector<uint64_t>> wq(cts.size(), std::vector<uint64_t>(h0.n_limbs));
  for (size_t l = 0; l < cts.size(); ++l) {
    CtHeader h;
    if (!ParseHeader(cts[l], h) || std::memcmp(&h, &h0, sizeof(h)) != 0)
      throw std::runtime_error("WeightedAverage: ciphertexts are not compatible");
    const long double scaled = std::nearbyint((long double)weights[l] * (long double)delta);
    const bool neg = scaled < 0;
    const unsigned __int128 mag = (unsigned __int128)(neg ? -scaled : scaled);
    for (uint32_t i = 0; i < h0.n_limbs; ++i) {
      uint64_t r = (uint64_t)(mag % h0.q[i]);
      wq[l][i] = neg && r ? h0.q[i] - r : r;
    }
  }
  CtHeader ho = h0;
  ho.scale = h0.scale * delta;
  std::string out(sizeof(ho) + words * 8, '\0');
  std::memcpy(out.data(), &ho, sizeof(ho));
  uint64_t* dst = reinterpret_cast<uint64_t*>(out.data() + sizeof(ho));
  const size_t rows = words / N;  // (ct, poly, limb) rows of N coefficients
#pragma omp parallel for schedule(static)
  for (long r = 0; r < (long)rows; ++r) {
    const uint32_t limb = (uint32_t)(r % h0.n_limbs);
    const uint64_t q = h0.q[limb];
    uint64_t* d = dst + (size_t)r * N;
    for (size_t l = 0; l < cts.size(); ++l) {
      const uint64_t* s = reinterpret_cast<const uint64_t*>(cts[l].data() + sizeof(CtHeader)) + (size_t)r * N;
      const uint64_t w = wq[l][limb], wp = shoup_const(w, q);
      if (l == 0) for (size_t i = 0; i < N; ++i) d[i] = mul_shoup(s[i], w, wp, q);
      else for (size_t i = 0; i < N; ++i) d[i] = add_mod(d[i], mul_shoup(s[i], w, wp, q), q);
    }
  }
  return out;
}

std::string Rescale(std::string_view blob) {
  CtHeader h;
  if (!ParseHeader(blob, h)) throw std::runtime_error("Rescale: malformed ciphertext blob");
  if (h.n_limbs < 2) throw std::runtime_error("Rescale: no limb left to drop");
  const uint32_t L = h.n_limbs, N = 1u << h.log_n;
  const uint64_t ql = h.q[L - 1];
  std::vector<NttTables> tb;
  for (uint32_t l = 0; l < L; ++l) tb.push_back(MakeNttTables(h.q[l], h.log_n));
  CtHeader ho = h;
  ho.n_limbs = (uint8_t)(L - 1);
  ho.q[L - 1] = 0;
  ho.scale = h.scale / (double)ql;
  std::string out(sizeof(ho) + (size_t)h.n_ct * ct_words(ho) * 8, '\0');
  std::memcpy(out.data(), &ho, sizeof(ho));
  const uint64_t* src = reinterpret_cast<const uint64_t*>(blob.data() + sizeof(h));
  uint64_t* dst = reinterpret_cast<uint64_t*>(out.data() + sizeof(ho));
#pragma omp parallel for schedule(dynamic)
  for (long pc = 0; pc < (long)h.n_ct * 2; ++pc) {  // every polynomial of every ciphertext
    const uint64_t* p = src + (size_t)pc * L * N;
    uint64_t* o = dst + (size_t)pc * (L - 1) * N;
    std::vector<uint64_t> last(p + (size_t)(L - 1) * N, p + (size_t)L * N), t(N);
    NttInverse(last.data(), tb[L - 1], h.log_n);
    for (uint32_t l = 0; l + 1 < L; ++l) {
      const uint64_t q = h.q[l], qinv = inv_mod(ql % q, q), qinv_s = shoup_const(qinv, q);
      for (uint32_t i = 0; i < N; ++i) {
        // centred representative of the dropped limb, moved to modulus q
        const uint64_t c = last[i];
        t[i] = c > ql / 2 ? sub_mod(c % q, ql % q, q) : c % q;
      }
      NttForward(t.data(), tb[l], h.log_n);
      for (uint32_t i = 0; i < N; ++i)
        o[(size_t)l * N + i] = mul_shoup(sub_mod(p[(size_t)l * N + i], t[i], q), qinv, qinv_s, q);
    }
  }
  return out;
}

}  // namespace mfl::ckks
