#include "core/aggregation.h"

#include <omp.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <type_traits>

#include "ckks/ckks_host.h"
#include "core/numeric.h"

namespace mfl::core {

using wire::VariableView;

namespace {

// out[i] = (acc ? out[i] : 0) + sum_l T(double(x_l[i]) * w_l), every term truncated to T and accumulated in T, in
// source order (federated_average.cc:25-35 semantics).  Chunked: one chunk of the accumulator stays in L1 while the
// sources stream through it one after the other, so the inner loop is a plain contiguous loop the compiler vectorises
// (the element-major form with a pointer per source compiled to scalar code: 4 GB/s on 8 threads).  The wire bytes are
// not necessarily aligned for T, hence the memcpy staging of every chunk.
constexpr size_t kSumChunk = 2048;

// The two floating-point types carry almost every model: their inner loops are compiled for AVX-512 / AVX2 / baseline
// and picked at load time (the rest of the core stays baseline x86-64).  The conversion to T between the multiply and
// the add is what the reference computes, and it also keeps the compiler from contracting them into an FMA.
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
#define MFL_CLONES __attribute__((target_clones("avx512f", "avx2", "default")))
#else
#define MFL_CLONES
#endif
MFL_CLONES void AddTermF32(float* a, const float* x, double w, size_t m) {
  for (size_t i = 0; i < m; ++i) a[i] = a[i] + float(double(x[i]) * w);
}
MFL_CLONES void SubTermF32(float* a, const float* x, double w, size_t m) {
  for (size_t i = 0; i < m; ++i) a[i] = a[i] - float(double(x[i]) * w);
}
MFL_CLONES void AddTermF64(double* a, const double* x, double w, size_t m) {
  for (size_t i = 0; i < m; ++i) a[i] = a[i] + x[i] * w;
}
MFL_CLONES void SubTermF64(double* a, const double* x, double w, size_t m) {
  for (size_t i = 0; i < m; ++i) a[i] = a[i] - x[i] * w;
}

template <typename T>
void WeightedSumT(size_t n, const std::vector<const char*>& srcs, const std::vector<double>& w, char* out, bool acc) {
  const size_t L = srcs.size();
  alignas(64) T a[kSumChunk];
  alignas(64) T x[kSumChunk];
  for (size_t base = 0; base < n; base += kSumChunk) {
    const size_t m = std::min(kSumChunk, n - base);
    if (acc) std::memcpy(a, out + base * sizeof(T), m * sizeof(T));
    else std::fill(a, a + m, T(0));   // (T may be a small class type: bf16 / fp16 wrappers)
    for (size_t l = 0; l < L; ++l) {
      std::memcpy(x, srcs[l] + base * sizeof(T), m * sizeof(T));
      // a negative weight subtracts the positive term (what the reference's SUBTRACTION merge does), which stays
      // well defined for unsigned T
      if constexpr (std::is_same_v<T, float>) {
        if (w[l] >= 0) AddTermF32(a, x, w[l], m); else SubTermF32(a, x, -w[l], m);
      } else if constexpr (std::is_same_v<T, double>) {
        if (w[l] >= 0) AddTermF64(a, x, w[l], m); else SubTermF64(a, x, -w[l], m);
      } else if (w[l] >= 0) {
        const double wl = w[l];
#pragma omp simd
        for (size_t i = 0; i < m; ++i) a[i] = T(a[i] + T(double(x[i]) * wl));
      } else {
        const double wl = -w[l];
#pragma omp simd
        for (size_t i = 0; i < m; ++i) a[i] = T(a[i] - T(double(x[i]) * wl));
      }
    }
    std::memcpy(out + base * sizeof(T), a, m * sizeof(T));
  }
}

template <typename T>
void ScaleT(size_t n, const char* x, double s, bool divide, char* out) {
  const T* xi = reinterpret_cast<const T*>(x);
  T* o = reinterpret_cast<T*>(out);
  for (size_t i = 0; i < n; ++i) {
    const double v = double(load_unaligned<T>(xi + i));
    store_unaligned<T>(o + i, T(divide ? v / s : v * s));
  }
}

std::vector<VariableView> ParseOrThrow(const std::string& m) {
  std::vector<VariableView> v;
  if (!wire::ParseModel(m, v)) throw std::runtime_error("malformed Model bytes");
  return v;
}

void CheckCompatible(const VariableView& a, const VariableView& b) {
  if (a.t.length != b.t.length) throw std::runtime_error("tensors have different sizes");
  if (a.t.dtype != b.t.dtype) throw std::runtime_error("tensors have different data types");
}

// A learner-supplied plaintext variable must match the reference metadata and carry as many bytes as it declares;
// everything downstream indexes value.data() by the reference length, so this is checked before any loop runs.
void CheckPayload(const VariableView& ref, const VariableView& x) {
  CheckCompatible(ref, x);
  const size_t esz = wire::DTypeSize(ref.t.dtype);
  if (esz == 0) throw std::runtime_error("unsupported tensor data type");
  if (x.t.value.size() < (size_t)ref.t.length * esz)
    throw std::runtime_error("tensor payload shorter than its declared length");
}

}  // namespace

void WeightedSumBytes(int dtype, size_t numel, const std::vector<const char*>& srcs,
                      const std::vector<double>& weights, char* out, bool accumulate) {
  MFL_HOST_DISPATCH(dtype, T, WeightedSumT<T>(numel, srcs, weights, out, accumulate));
}

void ScaleBytes(int dtype, size_t numel, const char* x, double s, bool divide, char* out) {
  MFL_HOST_DISPATCH(dtype, T, ScaleT<T>(numel, x, s, divide, out));
}

// ---------------------------------------------------------------------------------------------
Aggregated FedAvg::Aggregate(const std::vector<LearnerModels>& block) {
  Aggregated res;
  if (block.empty()) return res;
  std::vector<std::vector<VariableView>> parsed;
  std::vector<double> w;
  for (const auto& lm : block) {
    if (lm.lineage.empty()) continue;
    parsed.push_back(ParseOrThrow(*lm.lineage.front()));  // FedAvg reads the first lineage entry
    w.push_back(lm.weight);
  }
  if (parsed.empty()) return res;
  const size_t V = parsed[0].size();
  std::vector<std::string> values(V);
  // validate and size everything first (cheap, serial), then sum in parallel over (variable, element range) tasks:
  // per-variable parallelism alone leaves most threads idle behind the one that drew the largest tensor
  for (auto& p : parsed)
    if (p.size() != V) throw std::runtime_error("models have different numbers of variables");
  struct Task { size_t v, begin, end; };
  std::vector<Task> tasks;
  constexpr size_t kTaskElems = 1 << 17;
  for (size_t v = 0; v < V; ++v) {
    const VariableView& ref = parsed[0][v];
    if (ref.ciphertext) throw std::runtime_error("FedAvg aggregates plaintext variables only");
    const size_t esz = wire::DTypeSize(ref.t.dtype);
    if (esz == 0) throw std::runtime_error("unsupported tensor data type");
    for (auto& p : parsed) {
      CheckCompatible(ref, p[v]);
      if (p[v].t.value.size() < (size_t)ref.t.length * esz)
        throw std::runtime_error("tensor payload shorter than its declared length");
    }
    values[v].resize((size_t)ref.t.length * esz);
    for (size_t b = 0; b < (size_t)ref.t.length; b += kTaskElems)
      tasks.push_back({v, b, std::min((size_t)ref.t.length, b + kTaskElems)});
  }
  std::string err;
#pragma omp parallel for schedule(dynamic)
  for (long t = 0; t < (long)tasks.size(); ++t) {
    try {
      const Task& tk = tasks[t];
      const VariableView& ref = parsed[0][tk.v];
      const size_t esz = wire::DTypeSize(ref.t.dtype);
      std::vector<const char*> srcs;
      srcs.reserve(parsed.size());
      for (auto& p : parsed) srcs.push_back(p[tk.v].t.value.data() + tk.begin * esz);
      WeightedSumBytes(ref.t.dtype, tk.end - tk.begin, srcs, w, values[tk.v].data() + tk.begin * esz, false);
    } catch (const std::exception& e) {
#pragma omp critical
      err = e.what();
    }
  }
  if (!err.empty()) throw std::runtime_error(err);
  size_t total = 0;
  for (auto& s : values) total += s.size() + 64;
  res.model.reserve(total);
  for (size_t v = 0; v < V; ++v) {
    std::string_view sv(values[v]);
    wire::AppendVariable(res.model, parsed[0][v], &sv);
  }
  res.num_contributors = (uint32_t)parsed.size();
  return res;
}

// ---------------------------------------------------------------------------------------------
void RollingAverageBase::Clear() {
  meta_.clear();
  meta_backing_.clear();
  scaled_.clear();
  z_ = 0.0;
  contributors_ = 0;
}

void RollingAverageBase::Fold(const std::string* old_model, double w_old, const std::string& new_model, double w_new) {
  auto nv = ParseOrThrow(new_model);
  std::vector<VariableView> ov;
  if (old_model) ov = ParseOrThrow(*old_model);
  if (scaled_.empty()) {  // first contribution: remember metadata, S = 0
    std::string meta;
    for (auto& v : nv) {
      std::string_view none;
      wire::AppendVariable(meta, v, &none);
    }
    meta_backing_ = std::move(meta);
    meta_.clear();
    wire::ParseModel(meta_backing_, meta_);
    scaled_.resize(nv.size());
    for (size_t v = 0; v < nv.size(); ++v)
      scaled_[v].assign((size_t)nv[v].t.length * wire::DTypeSize(nv[v].t.dtype), '\0');
  }
  if (nv.size() != scaled_.size()) throw std::runtime_error("model shape changed between contributions");
  if (old_model && ov.size() != scaled_.size()) throw std::runtime_error("previous model has a different shape");
  // validate everything before S is touched: a malformed contribution must leave the rolling state intact
  for (size_t v = 0; v < nv.size(); ++v) {
    if (nv[v].ciphertext) continue;
    CheckPayload(meta_[v], nv[v]);
    if (old_model) CheckPayload(meta_[v], ov[v]);
    if (scaled_[v].size() < (size_t)meta_[v].t.length * wire::DTypeSize(meta_[v].t.dtype))
      throw std::runtime_error("rolling state does not match the model metadata");
  }
  std::string err;
#pragma omp parallel for schedule(dynamic)
  for (long v = 0; v < (long)nv.size(); ++v) {
    if (nv[v].ciphertext) continue;
    try {
      const int dt = meta_[v].t.dtype;
      const size_t n = meta_[v].t.length;
      if (old_model)  // S -= T(w_old * x_old)
        WeightedSumBytes(dt, n, {ov[v].t.value.data()}, {-w_old}, scaled_[v].data(), true);
      WeightedSumBytes(dt, n, {nv[v].t.value.data()}, {w_new}, scaled_[v].data(), true);  // S += T(w_new * x_new)
    } catch (const std::exception& e) {
#pragma omp critical
      err = e.what();
    }
  }
  if (!err.empty()) throw std::runtime_error(err);
}

Aggregated RollingAverageBase::Community() const {
  Aggregated res;
  if (scaled_.empty() || z_ == 0.0) return res;
  std::vector<std::string> values(scaled_.size());
#pragma omp parallel for schedule(dynamic)
  for (long v = 0; v < (long)scaled_.size(); ++v) {
    values[v].resize(scaled_[v].size());
    ScaleBytes(meta_[v].t.dtype, meta_[v].t.length, scaled_[v].data(), z_, /*divide=*/true, values[v].data());
  }
  for (size_t v = 0; v < scaled_.size(); ++v) {
    std::string_view sv(values[v]);
    wire::AppendVariable(res.model, meta_[v], &sv);
  }
  res.num_contributors = contributors_;
  return res;
}

Aggregated FedStride::Aggregate(const std::vector<LearnerModels>& block) {
  for (const auto& lm : block) {
    if (lm.lineage.empty()) continue;
    Fold(nullptr, 0.0, *lm.lineage.back(), lm.weight);
    z_ += lm.weight;
    ++contributors_;
  }
  return Community();
}

Aggregated FedRec::Aggregate(const std::vector<LearnerModels>& block) {
  // One completing learner per call (the controller passes exactly that learner: quirk Q2 fixed).
  for (const auto& lm : block) {
    if (lm.lineage.empty()) continue;
    if (lm.lineage.size() > 2) return Aggregated{};  // reference: lineage > 2 -> empty model
    if (lm.lineage.size() == 1) {
      Fold(nullptr, 0.0, *lm.lineage[0], lm.WeightOf(0));
      z_ += lm.WeightOf(0);
      ++contributors_;
    } else {
      // swap the learner's previous contribution for its newest one (z moves only once the fold went through)
      Fold(lm.lineage[0].get(), lm.WeightOf(0), *lm.lineage[1], lm.WeightOf(1));
      z_ += lm.WeightOf(1) - lm.WeightOf(0);
    }
  }
  return Community();
}

// ---------------------------------------------------------------------------------------------
double FedAsync::Staleness(int fn, double d, double a, double b) {
  d = d < 0 ? 0 : d;
  switch (fn) {
    case 1: return std::pow(1.0 + d, -a);
    case 2: return d <= b ? 1.0 : 1.0 / (a * (d - b) + 1.0);
    default: return 1.0;
  }
}

Aggregated FedAsync::Aggregate(const std::vector<LearnerModels>& block) {
  Aggregated res;
  for (const auto& lm : block) {
    if (lm.lineage.empty()) continue;
    const ModelBytes& x = lm.lineage.back();
    if (!community_) {  // first model bootstraps the community model
      community_ = x;
      contributors_ = 1;
      continue;
    }
    const double at = std::min(1.0, std::max(0.0, alpha_ * Staleness(fn_, lm.staleness, a_, b_)));
    auto cv = ParseOrThrow(*community_);
    auto xv = ParseOrThrow(*x);
    if (cv.size() != xv.size()) throw std::runtime_error("FedAsync: model shape mismatch");
    for (size_t v = 0; v < cv.size(); ++v) {
      if (cv[v].ciphertext || xv[v].ciphertext) throw std::runtime_error("FedAsync aggregates plaintext variables only");
      CheckPayload(cv[v], cv[v]);
      CheckPayload(cv[v], xv[v]);
    }
    std::vector<std::string> values(cv.size());
    std::string err;
#pragma omp parallel for schedule(dynamic)
    for (long v = 0; v < (long)cv.size(); ++v) {
      try {
        const int dt = cv[v].t.dtype;
        const size_t n = cv[v].t.length;
        values[v].resize(n * wire::DTypeSize(dt));
        if (dt <= 7) {  // integer bookkeeping variables: adopt the newest value
          std::memcpy(values[v].data(), xv[v].t.value.data(), values[v].size());
        } else {
          WeightedSumBytes(dt, n, {cv[v].t.value.data(), xv[v].t.value.data()}, {1.0 - at, at}, values[v].data(), false);
        }
      } catch (const std::exception& e) {
#pragma omp critical
        err = e.what();
      }
    }
    if (!err.empty()) throw std::runtime_error(err);
    auto next = std::make_shared<std::string>();
    for (size_t v = 0; v < cv.size(); ++v) {
      std::string_view sv(values[v]);
      wire::AppendVariable(*next, cv[v], &sv);
    }
    community_ = next;
    ++contributors_;
  }
  if (community_) res.model = *community_;
  res.num_contributors = contributors_;
  return res;
}

// ---------------------------------------------------------------------------------------------
PWA::PWA(const std::string& crypto_context_file, uint32_t batch_size, uint32_t scaling_factor_bits)
    : batch_size_(batch_size), scaling_bits_(scaling_factor_bits) {
  // The controller holds the crypto context only - never a key (private_weighted_average.cc:16-17).
  if (!crypto_context_file.empty()) {
    ckks::Params p;
    if (ckks::LoadParamsFile(crypto_context_file, p)) {
      batch_size_ = p.batch_size;
      scaling_bits_ = p.scale_bits;
    }
  }
}

Aggregated PWA::Aggregate(const std::vector<LearnerModels>& block) {
  Aggregated res;
  std::vector<std::vector<VariableView>> parsed;
  std::vector<double> w;
  for (const auto& lm : block) {
    if (lm.lineage.empty()) continue;
    parsed.push_back(ParseOrThrow(*lm.lineage.front()));
    w.push_back((double)(float)lm.weight);  // reference casts weights to float
  }
  if (parsed.empty()) return res;
  if (parsed.size() == 1) w[0] = 1.0;  // single learner -> weight 1 (private_weighted_average.cc:32-40)
  const size_t V = parsed[0].size();
  std::vector<std::string> values(V);
  std::string err;
#pragma omp parallel for schedule(dynamic)
  for (long v = 0; v < (long)V; ++v) {
    try {
      if (!parsed[0][v].ciphertext) throw std::runtime_error("PWA aggregates ciphertext variables only");
      std::vector<std::string_view> cts;
      for (auto& p : parsed) cts.push_back(p[v].t.value);
      values[v] = ckks::WeightedAverage(cts, w);
    } catch (const std::exception& e) {
#pragma omp critical
      err = e.what();
    }
  }
  if (!err.empty()) throw std::runtime_error(err);
  for (size_t v = 0; v < V; ++v) {
    std::string_view sv(values[v]);
    wire::AppendVariable(res.model, parsed[0][v], &sv);  // length/dims/dtype of the plaintext preserved
  }
  res.num_contributors = (uint32_t)parsed.size();
  return res;
}

}  // namespace mfl::core
