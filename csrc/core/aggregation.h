// Host aggregation rules of the controller core (also the CPU oracle of the device kernels).
//
// Behavioural spec = reference controller/aggregation/*: FedAvg (federated_average.cc:15-151),
// rolling scaled-model rules FedStride / FedRec (federated_rolling_average_base.cc:19-293,
// federated_stride.cc:6-64, federated_recency.cc:8-100) and PWA over CKKS ciphertexts
// (private_weighted_average.cc:24-82).  FedAsync (staleness-weighted mixing) is an extension.
// Models are serialized `metisfl.Model` bytes; tensors are read in place through wire::TensorView.
#pragma once
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "core/wire.h"

namespace mfl::core {

using ModelBytes = std::shared_ptr<const std::string>;
// one learner: its lineage oldest -> newest and its scaling factor
struct LearnerModels {
  std::vector<ModelBytes> lineage;
  double weight = 0.0;
  std::vector<double> lineage_weights;  // optional per-entry weights (FedRec {old,new} pair); default = weight
  double staleness = 0.0;  // FedAsync: community iterations since the learner's base model
  double WeightOf(size_t i) const { return i < lineage_weights.size() ? lineage_weights[i] : weight; }
};

struct Aggregated {
  std::string model;            // serialized metisfl.Model
  uint32_t num_contributors = 0;
};

class AggregationFunction {
 public:
  virtual ~AggregationFunction() = default;
  virtual Aggregated Aggregate(const std::vector<LearnerModels>& block) = 0;
  virtual int RequiredLearnerLineageLength() const = 0;
  virtual const char* Name() const = 0;
  virtual void Reset() = 0;
};

class FedAvg : public AggregationFunction {
 public:
  Aggregated Aggregate(const std::vector<LearnerModels>& block) override;
  int RequiredLearnerLineageLength() const override { return 1; }
  const char* Name() const override { return "FedAvg"; }
  void Reset() override {}
};

// Scaled model S (element type T, per-term truncation) + normaliser z; C = T(S / z).
class RollingAverageBase : public AggregationFunction {
 protected:
  void Fold(const std::string* old_model, double w_old, const std::string& new_model, double w_new);
  Aggregated Community() const;
  void Clear();
  std::vector<wire::VariableView> meta_;      // variable metadata of the first folded model
  std::string meta_backing_;                  // owns the bytes meta_ points into
  std::vector<std::string> scaled_;           // S per variable, raw T bytes
  double z_ = 0.0;
  uint32_t contributors_ = 0;
};

class FedStride : public RollingAverageBase {
 public:
  Aggregated Aggregate(const std::vector<LearnerModels>& block) override;
  int RequiredLearnerLineageLength() const override { return 1; }
  const char* Name() const override { return "FedStride"; }
  void Reset() override { Clear(); }
};

class FedRec : public RollingAverageBase {
 public:
  Aggregated Aggregate(const std::vector<LearnerModels>& block) override;
  int RequiredLearnerLineageLength() const override { return 2; }
  const char* Name() const override { return "FedRec"; }
  void Reset() override {}  // state persists across rounds
};

// C <- (1 - a_t) C + a_t x,  a_t = alpha * s(staleness)
class FedAsync : public AggregationFunction {
 public:
  FedAsync(double alpha, int staleness_fn, double a, double b) : alpha_(alpha), fn_(staleness_fn), a_(a), b_(b) {}
  Aggregated Aggregate(const std::vector<LearnerModels>& block) override;
  int RequiredLearnerLineageLength() const override { return 1; }
  const char* Name() const override { return "FedAsync"; }
  void Reset() override {}
  void SetCommunity(ModelBytes m) { community_ = std::move(m); }
  static double Staleness(int fn, double d, double a, double b);
 private:
  double alpha_; int fn_; double a_, b_;
  ModelBytes community_;
  uint32_t contributors_ = 0;
};

// Encrypted weighted average over ciphertext variables (CKKS limb-wise mul/add).
class PWA : public AggregationFunction {
 public:
  PWA(const std::string& crypto_context_file, uint32_t batch_size, uint32_t scaling_factor_bits);
  Aggregated Aggregate(const std::vector<LearnerModels>& block) override;
  int RequiredLearnerLineageLength() const override { return 1; }
  const char* Name() const override { return "PWA"; }
  void Reset() override {}
 private:
  uint32_t batch_size_, scaling_bits_;
};

// Weighted sum of raw tensors with the reference's element semantics; exposed for tests/oracle.
// out = sum_l T(double(x_l) * w_l) accumulated in T (+ existing `out` contents if accumulate).
void WeightedSumBytes(int dtype, size_t numel, const std::vector<const char*>& srcs,
                      const std::vector<double>& weights, char* out, bool accumulate);
// out = T(double(x) * s) (divide=false) or T(double(x) / s) (divide=true)
void ScaleBytes(int dtype, size_t numel, const char* x, double s, bool divide, char* out);

}  // namespace mfl::core
