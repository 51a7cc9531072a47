// Model stores: per-learner model lineage with an eviction policy.
//
// Behavioural spec: store/model_store.h:22-60, hash_map/hash_map_model_store.cc:36-118,
// redis/redis_model_store.cc:60-278.  InsertModel appends; LineageLengthEviction(k) drops the oldest
// entry when size >= k before appending (k = 0 -> 1); SelectModels(id, n): n > stored -> empty,
// n <= 0 -> all, else the last n, oldest -> newest.  The Redis variant keeps every model as a LIST
// of serialized Model.Variable entries and fetches with MULTI / LRANGE / EXEC; its RESP client is
// written here from scratch (no hiredis in the image).
#pragma once
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "core/aggregation.h"

namespace mfl::core {

struct EvictionPolicy {
  bool lineage_length_eviction = false;
  uint32_t lineage_length = 0;  // only with lineage_length_eviction
  uint32_t Effective() const { return lineage_length == 0 ? 1 : lineage_length; }
};

class ModelStore {
 public:
  explicit ModelStore(EvictionPolicy p) : policy_(p) {}
  virtual ~ModelStore() = default;
  virtual const char* Name() const = 0;
  virtual void InsertModel(const std::string& learner_id, ModelBytes model) = 0;
  // request: learner id -> number of models (<= 0: all)
  virtual std::map<std::string, std::vector<ModelBytes>> SelectModels(
      const std::vector<std::pair<std::string, int>>& request) = 0;
  virtual void EraseModels(const std::vector<std::string>& learner_ids) = 0;
  virtual void Expunge() = 0;
  virtual void ResetState() = 0;
  virtual int GetLearnerLineageLength(const std::string& learner_id) = 0;
  virtual void Shutdown() {}
  const EvictionPolicy& policy() const { return policy_; }
 protected:
  EvictionPolicy policy_;
};

class HashMapModelStore : public ModelStore {
 public:
  using ModelStore::ModelStore;
  const char* Name() const override { return "HashMapModelStore"; }
  void InsertModel(const std::string& learner_id, ModelBytes model) override;
  std::map<std::string, std::vector<ModelBytes>> SelectModels(
      const std::vector<std::pair<std::string, int>>& request) override;
  void EraseModels(const std::vector<std::string>& learner_ids) override;
  void Expunge() override;
  void ResetState() override {}
  int GetLearnerLineageLength(const std::string& learner_id) override;
 private:
  std::mutex mu_;
  std::map<std::string, std::deque<ModelBytes>> models_;
};

// Minimal blocking RESP2 client (one connection, caller serialises access).
class RespClient {
 public:
  struct Reply {
    char type = 0;  // '+', '-', ':', '$', '*', or 0 for nil
    std::string str;
    long long integer = 0;
    std::vector<Reply> elements;
    bool ok() const { return type != '-' && type != 0; }
  };
  ~RespClient();
  bool Connect(const std::string& host, int port, double timeout_s, std::string* err);
  bool Command(const std::vector<std::string_view>& args, Reply* reply, std::string* err);
  void Close();
  bool connected() const { return fd_ >= 0; }
 private:
  bool SendAll(const std::string& buf);
  bool ReadReply(Reply* r);
  bool ReadLine(std::string* line);
  bool ReadExact(size_t n, std::string* out);
  int fd_ = -1;
  std::string rbuf_;
  size_t rpos_ = 0;
};

class RedisModelStore : public ModelStore {
 public:
  RedisModelStore(const std::string& host, int port, EvictionPolicy p);
  const char* Name() const override { return "RedisModelStore"; }
  void InsertModel(const std::string& learner_id, ModelBytes model) override;
  std::map<std::string, std::vector<ModelBytes>> SelectModels(
      const std::vector<std::pair<std::string, int>>& request) override;
  void EraseModels(const std::vector<std::string>& learner_ids) override;
  void Expunge() override;
  void ResetState() override;
  int GetLearnerLineageLength(const std::string& learner_id) override;
  void Shutdown() override;
 private:
  std::string NextKey(const std::string& learner_id);
  std::mutex mu_;
  RespClient client_;
  std::map<std::string, std::deque<std::string>> lineage_;  // learner -> redis keys, oldest first
  std::map<std::string, ModelBytes> cache_;                 // fetched models, cleared by ResetState
  uint64_t counter_ = 0;
};

}  // namespace mfl::core
