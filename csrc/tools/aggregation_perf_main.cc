// Standalone aggregation micro-benchmark of the native controller core.
//
// Counterpart of the reference's scenarios harness
// (controller/scenarios/sync_model_aggregation_performance_main.cc:17-84 with
// scenarios_common.cc:76-289): argv = <learners> <tensors per model> <values per tensor>
// [rule=FedAvg|FedStride] [stride]; inserts identical synthetic models into the in-memory store,
// runs the block-wise community-model computation and prints the timings as JSON.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "core/aggregation.h"
#include "core/store.h"
#include "core/wire.h"

using namespace mfl::core;
namespace wire = mfl::wire;

static std::string GenerateModel(int tensors, long values) {
  std::string model;
  std::vector<float> data((size_t)values);
  for (long i = 0; i < values; ++i) data[i] = (float)(i % 97) * 0.01f;
  const std::string_view payload(reinterpret_cast<const char*>(data.data()), data.size() * sizeof(float));
  for (int t = 0; t < tensors; ++t) {
    const std::string name = "arr_" + std::to_string(t);
    wire::VariableView v;
    v.name = name;
    v.trainable = true;
    v.t.length = (uint32_t)values;
    v.t.dims = {values};
    v.t.dtype = 8;  // FLOAT32
    v.t.byte_order = 2;
    wire::AppendVariable(model, v, &payload);
  }
  return model;
}

static double Ms(std::chrono::steady_clock::time_point a) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - a).count();
}

int main(int argc, char** argv) {
  if (argc < 4) {
    std::fprintf(stderr, "usage: %s <learners> <tensors> <values_per_tensor> [FedAvg|FedStride] [stride]\n", argv[0]);
    return 2;
  }
  const int learners = std::atoi(argv[1]), tensors = std::atoi(argv[2]);
  const long values = std::atol(argv[3]);
  const std::string rule = argc > 4 ? argv[4] : "FedAvg";
  const int stride = argc > 5 ? std::atoi(argv[5]) : learners;
  auto t0 = std::chrono::steady_clock::now();
  auto model = std::make_shared<std::string>(GenerateModel(tensors, values));
  const double gen_ms = Ms(t0);
  HashMapModelStore store(EvictionPolicy{true, 1});
  t0 = std::chrono::steady_clock::now();
  for (int l = 0; l < learners; ++l) store.InsertModel("learner" + std::to_string(l), std::make_shared<std::string>(*model));
  const double insert_ms = Ms(t0);
  std::unique_ptr<AggregationFunction> agg;
  if (rule == "FedStride") agg = std::make_unique<FedStride>();
  else agg = std::make_unique<FedAvg>();
  const int block_len = rule == "FedStride" ? std::max(1, stride) : learners;
  std::vector<double> block_ms;
  Aggregated result;
  t0 = std::chrono::steady_clock::now();
  for (int b = 0; b < learners; b += block_len) {
    auto tb = std::chrono::steady_clock::now();
    std::vector<std::pair<std::string, int>> req;
    for (int l = b; l < std::min(learners, b + block_len); ++l) req.emplace_back("learner" + std::to_string(l), 1);
    auto sel = store.SelectModels(req);
    std::vector<LearnerModels> block;
    for (auto& kv : sel) {
      LearnerModels lm;
      lm.lineage = kv.second;
      lm.weight = 1.0 / learners;
      block.push_back(std::move(lm));
    }
    result = agg->Aggregate(block);
    store.ResetState();
    block_ms.push_back(Ms(tb));
  }
  agg->Reset();
  const double total_ms = Ms(t0);
  const double params = (double)tensors * (double)values;
  std::printf("{\"rule\": \"%s\", \"learners\": %d, \"tensors\": %d, \"values_per_tensor\": %ld, \"model_bytes\": %zu, "
              "\"generate_ms\": %.3f, \"insert_ms\": %.3f, \"aggregation_total_ms\": %.3f, \"blocks\": %zu, "
              "\"aggregated_params_per_s\": %.4g, \"num_contributors\": %u}\n",
              rule.c_str(), learners, tensors, values, model->size(), gen_ms, insert_ms, total_ms, block_ms.size(),
              learners * params / (total_ms * 1e-3), result.num_contributors);
  return 0;
}
