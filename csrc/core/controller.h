// Controller core: the federation state machine.
//
// Behavioural spec = reference controller/core/controller.cc: learner
// registry with auth tokens (:98-199), task templates (:148-153), global iteration + runtime metadata
// lineage (:394-426, :428-518), model-store insertion on task completion (:201-259), block-wise
// community-model computation (:795-950), tensor quantification (:952-1004), semi-synchronous task
// re-sizing (:520-569), evaluation lineage (:571-694), lineage getters (:261-332).
//
// Differences by design: no gRPC or protobuf library in here - RPC transport lives in Python
// (metisfl_b200/controller/service.py) and talks to this class through a Dispatcher; models are
// kept as shared serialized bytes and aggregated in place; reference quirks Q2/Q5/Q6/Q8/Q9/Q14 are
// fixed (completing learner folded for async rules, "has variables" check, random tokens, one lock
// for all bookkeeping, task deadline + eviction, weights normalised over actual contributors).
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "core/aggregation.h"
#include "core/policies.h"
#include "core/store.h"
#include "core/thread_pool.h"

namespace mfl::core {

enum class Protocol { kUnknown = 0, kSynchronous = 1, kAsynchronous = 2, kSemiSynchronous = 3 };

enum class Status { kOk = 0, kInvalidArgument, kAlreadyExists, kNotFound, kUnauthenticated, kFailedPrecondition, kInternal };

struct StatusOr {
  Status code = Status::kOk;
  std::string message;
  bool ok() const { return code == Status::kOk; }
};

struct ControllerParams {
  std::string hostname = "localhost";
  int port = 50051;
  std::string rule = "FedAvg";  // FedAvg | FedStride | FedRec | FedAsync | PWA
  uint32_t stride_length = 0;
  double fedasync_alpha = 0.5;
  int staleness_fn = 1;
  double staleness_a = 0.5, staleness_b = 4.0;
  ScalingKind scaling = ScalingKind::kNumTrainingExamples;
  Protocol protocol = Protocol::kSynchronous;
  int semi_sync_lambda = 0;
  bool semi_sync_recompute = false;
  double task_deadline_secs = 0.0;  // 0 = never evict (reference behaviour)
  bool redis = false;
  std::string redis_host = "127.0.0.1";
  int redis_port = 6379;
  EvictionPolicy eviction;
  uint32_t batch_size = 100, epochs = 5;
  float percent_validation = 0.f;
  // synchronous / semi-synchronous protocols: fraction of the active learners that get a task each round (client
  // sampling, the "C" of FedAvg).  The reference plumbs the field end to end and never reads it (metis.proto:307,
  // controller.cc:730-731); here a value in (0, 1) draws a fresh cohort per round.
  float participation_ratio = 1.f;
  uint64_t sampling_seed = 0;             // 0 = seeded from std::random_device
  std::string he_crypto_context_file;
  uint32_t he_batch_size = 4096, he_scaling_bits = 52;
  uint32_t community_lineage_length = 4;  // community checkpoints kept for GetCommunityModelLineage
  bool evaluate_community_model = true;
};

struct LearnerEndpoint {
  std::string id, hostname;
  int port = 0;
  std::string public_certificate;  // PEM stream sent by the learner on join (may be empty)
};

struct TaskItem {
  LearnerEndpoint learner;
  uint32_t global_iteration = 0;
  uint32_t num_local_updates = 0;
  float percent_validation = 0.f;
};

struct TensorQuant { uint32_t non_zeros = 0, zeros = 0, bytes = 0; };

// ---- device-resident federation ----------------------------------------------------------
// A learner that hosts a rank of the on-box NVLink fabric does not upload its model: MarkTaskCompleted carries a
// Model whose TensorSpecs have no value bytes but a device_slot{rank, slot, offset, version}.  When the models of
// an aggregation block are such references, the controller does not run its CPU loop; it names the slots, the
// scaling factors it computed and the rule in a DeviceAggregateRequest and the dispatcher makes the GPU processes
// run the aggregation kernel (K1 for FedAvg / FedStride, K3 for FedRec / FedAsync).  Learners outside the fabric
// still upload bytes; their models travel inside the request and are unpacked into a staging slot (K12).
struct DeviceSource {
  std::string learner_id;
  uint32_t rank = 0, slot = 0;
  uint64_t version = 0;
  double weight = 0.0;
  bool has_previous = false;  // FedRec: the contribution this one replaces is still in the learner's slot ring
  uint32_t previous_slot = 0;
  double previous_weight = 0.0;
  ModelBytes staged_model;  // non-null for an off-box learner (serialized metisfl.Model with bytes)
};

struct DeviceAggregateRequest {
  std::string rule;
  uint32_t global_iteration = 0;
  std::vector<DeviceSource> sources;
  double mix = 0.0;  // FedAsync: alpha * s(staleness) of the single source
  double z = 0.0;    // FedRec: normaliser after this fold
  bool first_block = true, last_block = true;  // FedStride streams blocks into one running result
  bool quantify = true;
};

struct DeviceAggregateResult {
  std::string model_ref;    // serialized Model, control-only (device_slot references to the community buffer)
  std::string model_bytes;  // serialized Model with values, when the service layer asked a rank for them
  double device_ms = 0.0;   // CUDA-event duration, max over the ranks that ran the kernel
  std::vector<TensorQuant> quantifiers;  // K4 counts of the new community model
  std::string error;
};

// Outbound RPC fan-out, implemented by the Python service layer.
class Dispatcher {
 public:
  virtual ~Dispatcher() = default;
  // federated_model = serialized metisfl.FederatedModel shared by all items; federated_model_ref (may be null) is
  // its control-only twin for learners that sit on the fabric
  virtual void SendRunTasks(const std::vector<TaskItem>& items, const ModelBytes& federated_model,
                            const ModelBytes& federated_model_ref) = 0;
  // model = serialized metisfl.Model; eval_index addresses the community evaluation lineage entry
  virtual void SendEvaluationTasks(const std::vector<LearnerEndpoint>& learners, const ModelBytes& model,
                                   const ModelBytes& model_ref, uint32_t batch_size, uint32_t eval_index,
                                   uint32_t global_iteration) = 0;
  // Blocking; returns false when there is no device fabric behind this dispatcher (host aggregation is used then).
  // Called with the controller lock held: the implementation must not call back into the controller.
  virtual bool AggregateOnDevice(const DeviceAggregateRequest&, DeviceAggregateResult*) { return false; }
};

struct TaskMeta {  // the fields of metisfl.TaskExecutionMetadata the controller interprets
  uint32_t global_iteration = 0;
  float completed_epochs = 0.f;
  uint32_t completed_batches = 0, batch_size = 0;
  float ms_per_epoch = 0.f, ms_per_batch = 0.f;
  std::string serialized;  // the whole TaskExecutionMetadata, echoed by GetLocalTaskLineage
};

struct RuntimeMetadata {
  uint32_t global_iteration = 0;
  double started_at = 0, completed_at = 0;  // unix seconds
  std::vector<std::string> assigned, completed;
  std::map<std::string, double> train_submitted_at, train_received_at, eval_submitted_at, eval_received_at;
  std::map<std::string, double> insertion_ms, selection_ms;
  double aggregation_started_at = 0, aggregation_completed_at = 0, aggregation_total_ms = 0;
  std::vector<double> block_size, block_memory_kb, block_duration_ms;
  std::vector<TensorQuant> quantifiers;
  bool device_timed = false;  // block durations are CUDA-event times reported by the GPU processes
};

struct CommunityEvaluation {
  uint32_t global_iteration = 0;
  std::map<std::string, std::string> evaluations;  // learner id -> serialized metisfl.ModelEvaluations
};

struct LearnerRecord {
  LearnerEndpoint endpoint;
  std::string auth_token;
  std::string server_entity;  // serialized metisfl.ServerEntity (echoed back)
  std::string dataset_spec;   // serialized metisfl.DatasetSpec
  uint32_t n_train = 0, n_valid = 0, n_test = 0;
  uint32_t num_local_updates = 0;
  uint32_t task_issued_iteration = 0;
  double task_issued_at = 0;
  bool task_in_flight = false;
};

class Controller {
 public:
  Controller(ControllerParams params, std::shared_ptr<Dispatcher> dispatcher);
  ~Controller();

  const ControllerParams& params() const { return params_; }
  StatusOr ReplaceCommunityModel(std::string federated_model_bytes);
  StatusOr AddLearner(const std::string& hostname, int port, std::string server_entity, std::string dataset_spec,
                      uint32_t n_train, uint32_t n_valid, uint32_t n_test, std::string public_cert,
                      std::string* learner_id, std::string* auth_token);
  StatusOr RemoveLearner(const std::string& id, const std::string& token);
  StatusOr LearnerCompletedTask(const std::string& id, const std::string& token, std::string model_bytes, TaskMeta meta);
  void EvaluationCompleted(const std::string& id, uint32_t eval_index, std::string evaluations_bytes);
  void TaskFailed(const std::string& id, const std::string& what);
  // Failure detector hook (heartbeat): drop a learner that stopped answering; returns false if it was unknown.
  bool EvictLearner(const std::string& id, const std::string& why);

  std::vector<LearnerRecord> GetLearners();
  ModelBytes CommunityModel();  // serialized FederatedModel
  uint32_t GlobalIteration();
  uint32_t AggregationFailures();
  std::vector<RuntimeMetadata> GetRuntimeMetadataLineage(int num_backtracks);
  std::vector<CommunityEvaluation> GetCommunityEvaluationLineage(int num_backtracks);
  std::map<std::string, std::vector<std::string>> GetLocalTaskLineage(const std::vector<std::string>& ids, int num_backtracks);
  std::vector<ModelBytes> GetCommunityModelLineage(int num_backtracks);
  std::map<std::string, std::vector<ModelBytes>> GetLearnerLocalModelLineage(const std::vector<std::string>& ids, int num_backtracks);
  void WaitIdle();  // block until the scheduling pool has drained (tests)
  void Shutdown();

 private:
  void ScheduleInitialTask(const std::string& id);
  void ScheduleInitialTaskLocked(const std::string& id);
  void ScheduleTasks(const std::string& id);
  void AfterBarrier(const std::vector<std::string>& to_schedule, const std::string& completing);
  ModelBytes ComputeCommunityModel(const std::vector<std::string>& selected, const std::string& completing,
                                   RuntimeMetadata& md, uint32_t* contributors);
  void RecordModelSize(const std::string& model, RuntimeMetadata& md);
  // Device path of one aggregation block; returns false when the block has to run on the host.
  bool AggregateBlockOnDevice(const std::vector<std::pair<std::string, int>>& req,
                              const std::map<std::string, std::vector<ModelBytes>>& models,
                              const std::map<std::string, double>& weights, bool first_block, bool last_block,
                              RuntimeMetadata& md, Aggregated* result);
  void UpdateTaskTemplates(const std::vector<std::string>& ids);
  void SendRunTasksLocked(const std::vector<std::string>& ids);
  void SendEvaluationTasksLocked(const std::vector<std::string>& ids, uint32_t eval_index);
  std::vector<std::string> ActiveIdsLocked() const;
  // client sampling: the learners the synchronous barrier waits for / the next round's cohort
  bool Sampling() const;
  std::vector<std::string> BarrierSetLocked() const;
  std::vector<std::string> SampleCohortLocked(const std::vector<std::string>& active);
  std::set<std::string> cohort_;          // learners holding a task of the current round (sampling only)
  std::mt19937_64 sample_rng_;
  void WatchdogLoop();
  void EvictLocked(const std::string& id, const char* why);
  // Dispatcher calls are queued while mu_ is held and issued after it is released, so a dispatcher
  // may call straight back into the controller.
  void FlushOutbox();
  std::vector<std::function<void()>> outbox_;
  std::mutex outbox_mu_;

  ControllerParams params_;
  std::shared_ptr<Dispatcher> dispatcher_;
  std::unique_ptr<AggregationFunction> aggregator_;
  std::unique_ptr<Scheduler> scheduler_;
  std::unique_ptr<ModelStore> store_;
  std::mutex mu_;  // guards everything below (learners, metadata, lineages)
  std::map<std::string, LearnerRecord> learners_;
  std::map<std::string, std::deque<TaskMeta>> local_tasks_;  // newest first
  std::vector<std::string> pending_initial_;
  ModelBytes community_;           // serialized FederatedModel
  ModelBytes community_model_;     // serialized Model inside it
  ModelBytes community_ref_;       // control-only FederatedModel (device_slot references), null on the host path
  ModelBytes community_model_ref_; // control-only Model inside it
  std::string pending_ref_;        // model_ref of the aggregation in progress
  // rolling state of FedRec on the device path (the sums live in HBM, the scalars here)
  double device_z_ = 0.0;
  uint32_t device_contributors_ = 0;
  std::map<std::string, double> device_w_;
  std::deque<ModelBytes> community_lineage_;
  uint32_t global_iteration_ = 0;
  uint32_t aggregation_failures_ = 0;
  std::vector<RuntimeMetadata> metadata_;
  std::vector<CommunityEvaluation> evaluations_;
  ThreadPool pool_;
  std::thread watchdog_;
  std::atomic<bool> stop_{false};
  std::condition_variable wd_cv_;
  std::mutex wd_mu_;
};

double NowSeconds();
long MaxRssKb();
std::string RandomToken(size_t n_chars = 64);

}  // namespace mfl::core
