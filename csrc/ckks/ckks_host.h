// Host RNS-CKKS: the CPU oracle and timed baseline (B4 `cpu-ckks`) for the GPU kernels, and the
// implementation behind PWA on the controller's host path.
// API shape follows the reference's HEScheme/CKKS (encryption/palisade/he_scheme.h:20-43,
// ckks_scheme.h:29-65): key/context files, Encrypt(vector<double>) -> bytes,
// ComputeWeightedAverage(list<bytes>, weights) -> bytes, Decrypt(bytes, n) -> vector<double>.
#pragma once
#include <complex>
#include <memory>
#include <string>
#include <string_view>
#include <vector>

#include "ckks/ckks_common.h"

namespace mfl::ckks {

struct Params {
  uint32_t log_n = kDefaultLogN;
  std::vector<uint64_t> q{kDefaultModuli[0], kDefaultModuli[1], kDefaultModuli[2]};
  uint32_t scale_bits = 52;
  uint32_t batch_size = 4096;
  uint32_t n() const { return 1u << log_n; }
};

bool LoadParamsFile(const std::string& path, Params& p);
bool SaveParamsFile(const std::string& path, const Params& p);

// Per-modulus NTT tables (bit-reversed powers of a primitive 2N-th root psi, with Shoup constants).
struct NttTables {
  uint64_t q = 0, psi = 0, n_inv = 0, n_inv_shoup = 0;
  std::vector<uint64_t> psi_br, psi_br_shoup;        // forward (Cooley-Tukey, natural -> bit-reversed)
  std::vector<uint64_t> ipsi_br, ipsi_br_shoup;      // inverse (Gentleman-Sande)
};
NttTables MakeNttTables(uint64_t q, uint32_t log_n);
void NttForward(uint64_t* a, const NttTables& t, uint32_t log_n);
void NttInverse(uint64_t* a, const NttTables& t, uint32_t log_n);

class Scheme {
 public:
  explicit Scheme(const Params& p);
  const Params& params() const { return p_; }
  const std::vector<NttTables>& tables() const { return tables_; }

  // keys (evaluation-form polynomials, limb-major)
  void KeyGen(uint64_t seed = 0);
  bool has_public_key() const { return !pk_.empty(); }
  bool has_secret_key() const { return !sk_.empty(); }
  const std::vector<uint64_t>& public_key() const { return pk_; }   // [2][L][N]
  const std::vector<uint64_t>& secret_key() const { return sk_; }   // [L][N]
  std::string SerializePublicKey() const, SerializeSecretKey() const, SerializeEvalMultKey() const;
  bool LoadPublicKey(std::string_view blob), LoadSecretKey(std::string_view blob);

  // plaintext <-> polynomial (coefficient form, signed integers scaled by `scale`)
  void EncodeCoeffs(const double* vals, size_t n, double scale, std::vector<double>& coeffs) const;
  void DecodeCoeffs(const std::vector<double>& coeffs, size_t n, double* vals) const;

  std::string Encrypt(const double* vals, size_t n, uint64_t seed = 0) const;
  std::vector<double> Decrypt(std::string_view blob, size_t n) const;

 private:
  void FftSpecial(std::complex<double>* v, uint32_t size) const;
  void FftSpecialInv(std::complex<double>* v, uint32_t size) const;
  Params p_;
  std::vector<NttTables> tables_;
  std::vector<uint32_t> rot_group_;
  std::vector<std::complex<double>> ksi_pows_;
  std::vector<uint64_t> pk_, sk_, evk_;
};

// Context-free homomorphic ops (moduli come from the blob header).
std::string WeightedAverage(const std::vector<std::string_view>& cts, const std::vector<double>& weights);
std::string Rescale(std::string_view blob);  // drop the last limb, scale /= q_last
bool ParseHeader(std::string_view blob, CtHeader& h);

}  // namespace mfl::ckks
