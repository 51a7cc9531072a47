#include "core/controller.h"

#include <sys/resource.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <random>
#include <stdexcept>

#include "core/numeric.h"

namespace mfl::core {

double NowSeconds() {
  using namespace std::chrono;
  return duration_cast<duration<double>>(system_clock::now().time_since_epoch()).count();
}

long MaxRssKb() {
  rusage u{};
  getrusage(RUSAGE_SELF, &u);
  return u.ru_maxrss;
}

std::string RandomToken(size_t n) {
  static const char kHex[] = "0123456789abcdef";
  std::random_device rd;
  std::string s(n, '0');
  for (size_t i = 0; i < n; i += 8) {
    uint32_t r = rd();
    for (size_t j = i; j < std::min(n, i + 8); ++j, r >>= 4) s[j] = kHex[r & 15];
  }
  return s;
}

static double ElapsedMs(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

static void Log(const char* fmt, const std::string& a = "", const std::string& b = "") {
  std::fprintf(stderr, "[metisfl_b200.controller] ");
  std::fprintf(stderr, fmt, a.c_str(), b.c_str());
  std::fprintf(stderr, "\n");
}

// ---------------------------------------------------------------------------------------------
Controller::Controller(ControllerParams params, std::shared_ptr<Dispatcher> dispatcher)
    : params_(std::move(params)), dispatcher_(std::move(dispatcher)), pool_(2) {
  if (params_.batch_size == 0) throw std::invalid_argument("batch_size must be > 0");
  if (params_.epochs == 0) throw std::invalid_argument("epochs must be > 0");
  const std::string& r = params_.rule;
  if (r == "FedAvg") aggregator_ = std::make_unique<FedAvg>();
  else if (r == "FedStride") aggregator_ = std::make_unique<FedStride>();
  else if (r == "FedRec") aggregator_ = std::make_unique<FedRec>();
  else if (r == "FedAsync")
    aggregator_ = std::make_unique<FedAsync>(params_.fedasync_alpha, params_.staleness_fn, params_.staleness_a,
                                             params_.staleness_b);
  else if (r == "PWA")
    aggregator_ = std::make_unique<PWA>(params_.he_crypto_context_file, params_.he_batch_size, params_.he_scaling_bits);
  else throw std::invalid_argument("unknown aggregation rule: " + r);
  if (params_.protocol == Protocol::kAsynchronous) scheduler_ = std::make_unique<AsynchronousScheduler>();
  else if (params_.protocol == Protocol::kSynchronous || params_.protocol == Protocol::kSemiSynchronous)
    scheduler_ = std::make_unique<SynchronousScheduler>();
  else throw std::invalid_argument("unknown communication protocol");
  if (params_.redis) store_ = std::make_unique<RedisModelStore>(params_.redis_host, params_.redis_port, params_.eviction);
  else store_ = std::make_unique<HashMapModelStore>(params_.eviction);
  if (params_.participation_ratio < 0.f || params_.participation_ratio > 1.f)
    throw std::invalid_argument("participation_ratio must be in [0, 1] (0 = everybody)");
  sample_rng_.seed(params_.sampling_seed ? params_.sampling_seed : std::random_device{}());
  if (params_.task_deadline_secs > 0) watchdog_ = std::thread([this] { WatchdogLoop(); });
}

Controller::~Controller() { Shutdown(); }

void Controller::Shutdown() {
  if (stop_.exchange(true)) return;
  wd_cv_.notify_all();
  if (watchdog_.joinable()) watchdog_.join();
  pool_.Stop();
  store_->Shutdown();
}

void Controller::WaitIdle() { pool_.WaitIdle(); }

std::vector<std::string> Controller::ActiveIdsLocked() const {
  std::vector<std::string> ids;
  for (auto& kv : learners_) ids.push_back(kv.first);
  return ids;
}

bool Controller::Sampling() const {
  return params_.protocol != Protocol::kAsynchronous && params_.participation_ratio > 0.f &&
         params_.participation_ratio < 1.f;
}

std::vector<std::string> Controller::BarrierSetLocked() const {
  auto active = ActiveIdsLocked();
  if (!Sampling() || cohort_.empty()) return active;
  std::vector<std::string> out;
  for (auto& a : active)
    if (cohort_.count(a)) out.push_back(a);
  return out;
}

std::vector<std::string> Controller::SampleCohortLocked(const std::vector<std::string>& active) {
  std::vector<std::string> pool(active.begin(), active.end());
  const size_t k = std::min(pool.size(), (size_t)std::max(1.0, std::ceil((double)params_.participation_ratio * pool.size())));
  // partial Fisher-Yates: the first k entries are a uniform sample without replacement
  for (size_t i = 0; i < k; ++i) {
    std::uniform_int_distribution<size_t> pick(i, pool.size() - 1);
    std::swap(pool[i], pool[pick(sample_rng_)]);
  }
  pool.resize(k);
  std::sort(pool.begin(), pool.end());
  return pool;
}

// ---------------------------------------------------------------------------------------------
StatusOr Controller::ReplaceCommunityModel(std::string bytes) {
  uint32_t contributors, iteration;
  std::string_view model;
  if (!wire::ParseFederatedModel(bytes, contributors, iteration, model))
    return {Status::kInvalidArgument, "malformed FederatedModel"};
  std::vector<wire::VariableView> vars;
  if (!wire::ParseModel(model, vars)) return {Status::kInvalidArgument, "malformed Model"};
  std::vector<std::string> pending;
  {
    std::lock_guard<std::mutex> l(mu_);
    community_model_ = std::make_shared<std::string>(model);
    community_ = std::make_shared<std::string>(std::move(bytes));
    community_ref_.reset();  // fabric learners install these bytes once; references start with the first aggregation
    community_model_ref_.reset();
    if (auto* fa = dynamic_cast<FedAsync*>(aggregator_.get())) fa->SetCommunity(community_model_);
    community_lineage_.push_back(community_);
    while (community_lineage_.size() > std::max<uint32_t>(1, params_.community_lineage_length)) community_lineage_.pop_front();
    pending.swap(pending_initial_);
  }
  Log("community model replaced (%s variables)", std::to_string(vars.size()));
  for (auto& id : pending) pool_.Push([this, id] { ScheduleInitialTask(id); });
  return {};
}

StatusOr Controller::AddLearner(const std::string& hostname, int port, std::string server_entity, std::string dataset_spec,
                                uint32_t n_train, uint32_t n_valid, uint32_t n_test, std::string public_cert,
                                std::string* learner_id, std::string* auth_token) {
  if (hostname.empty() || port <= 0) return {Status::kInvalidArgument, "Must provide a valid hostname and port."};
  if (n_train == 0) return {Status::kInvalidArgument, "Learner training examples <= 0."};
  const std::string id = hostname + ":" + std::to_string(port);
  {
    std::lock_guard<std::mutex> l(mu_);
    if (learners_.count(id)) return {Status::kAlreadyExists, "Learner has already joined."};
    LearnerRecord rec;
    rec.endpoint = {id, hostname, port, std::move(public_cert)};
    rec.auth_token = RandomToken(64);
    rec.server_entity = std::move(server_entity);
    rec.dataset_spec = std::move(dataset_spec);
    rec.n_train = n_train; rec.n_valid = n_valid; rec.n_test = n_test;
    rec.num_local_updates = params_.epochs * (uint32_t)std::ceil((double)n_train / (double)params_.batch_size);
    *learner_id = id;
    *auth_token = rec.auth_token;
    learners_.emplace(id, std::move(rec));
  }
  Log("learner %s joined the federation", id);
  pool_.Push([this, id] { ScheduleInitialTask(id); });
  return {};
}

StatusOr Controller::RemoveLearner(const std::string& id, const std::string& token) {
  std::vector<std::string> resume;
  {
    std::lock_guard<std::mutex> l(mu_);
    auto it = learners_.find(id);
    if (it == learners_.end()) return {Status::kNotFound, "No such learner."};
    if (it->second.auth_token != token) return {Status::kUnauthenticated, "Incorrect credentials."};
    EvictLocked(id, "left the federation");
  }
  return {};
}

void Controller::EvictLocked(const std::string& id, const char* why) {
  learners_.erase(id);
  local_tasks_.erase(id);
  store_->EraseModels({id});
  device_w_.erase(id);
  const bool was_in_cohort = cohort_.erase(id) > 0;
  Log("learner %s %s", id, why);
  // a shrinking federation may complete a synchronous barrier
  if (params_.protocol != Protocol::kAsynchronous && !learners_.empty() && !stop_) {
    pool_.Push([this, was_in_cohort] {
      {
        std::lock_guard<std::mutex> l(mu_);
        auto* sync = dynamic_cast<SynchronousScheduler*>(scheduler_.get());
        if (!sync) return;
        if (Sampling() && was_in_cohort && cohort_.empty() && !learners_.empty() && global_iteration_ > 0) {
          // the whole cohort of this round is gone: nobody is left to reach the barrier, so the round closes without
          // a contribution and a fresh cohort is drawn from the learners that remain
          AfterBarrier({}, "");
        }
        auto active = BarrierSetLocked();
        if (active.empty()) return;
        // re-check the barrier by replaying an already-completed learner (idempotent insert)
        for (auto& a : active) {
          auto lt = local_tasks_.find(a);
          if (lt != local_tasks_.end() && !lt->second.empty() &&
              lt->second.front().global_iteration == global_iteration_) {
            auto next = scheduler_->ScheduleNext(a, active);
            if (!next.empty()) AfterBarrier(next, a);
            break;
          }
        }
      }
      FlushOutbox();
    });
  }
}

StatusOr Controller::LearnerCompletedTask(const std::string& id, const std::string& token, std::string model_bytes,
                                          TaskMeta meta) {
  {
    std::lock_guard<std::mutex> l(mu_);
    auto it = learners_.find(id);
    if (it == learners_.end() || it->second.auth_token != token)
      return {Status::kUnauthenticated, "Learner does not exist or incorrect credentials."};
    // store first: a failing insert (Redis down, malformed bytes) must leave the round bookkeeping untouched so the
    // learner can retry, and must reach the learner as a status instead of unwinding through the RPC thread
    auto t0 = std::chrono::steady_clock::now();
    try {
      store_->InsertModel(id, std::make_shared<std::string>(std::move(model_bytes)));
    } catch (const std::exception& e) {
      return {Status::kInternal, std::string("model store insert failed: ") + e.what()};
    }
    const double insert_ms = ElapsedMs(t0);
    it->second.task_in_flight = false;
    // the task's own iteration stamps the metadata slot (falls back to the current round)
    uint32_t iter = meta.global_iteration ? meta.global_iteration : global_iteration_;
    if (iter >= 1 && iter <= metadata_.size()) {
      auto& md = metadata_[iter - 1];
      md.completed.push_back(id);
      md.train_received_at[id] = NowSeconds();
    }
    if (!metadata_.empty()) metadata_.back().insertion_ms[id] = insert_ms;
    if (!meta.global_iteration) meta.global_iteration = global_iteration_;
    local_tasks_[id].push_front(std::move(meta));
  }
  pool_.Push([this, id] { ScheduleTasks(id); });
  return {};
}

void Controller::EvaluationCompleted(const std::string& id, uint32_t eval_index, std::string bytes) {
  std::lock_guard<std::mutex> l(mu_);
  if (eval_index < evaluations_.size()) {
    evaluations_[eval_index].evaluations[id] = std::move(bytes);
    const uint32_t it = evaluations_[eval_index].global_iteration;
    if (it >= 1 && it <= metadata_.size()) metadata_[it - 1].eval_received_at[id] = NowSeconds();
  }
}

void Controller::TaskFailed(const std::string& id, const std::string& what) {
  Log("RPC to learner %s failed: %s", id, what);
}

bool Controller::EvictLearner(const std::string& id, const std::string& why) {
  {
    std::lock_guard<std::mutex> l(mu_);
    if (!learners_.count(id)) return false;
    EvictLocked(id, why.c_str());
  }
  FlushOutbox();
  return true;
}

// ---------------------------------------------------------------------------------------------
void Controller::ScheduleInitialTask(const std::string& id) {
  {
    std::lock_guard<std::mutex> l(mu_);
    ScheduleInitialTaskLocked(id);
  }
  FlushOutbox();
}

void Controller::ScheduleInitialTaskLocked(const std::string& id) {
  if (stop_ || !learners_.count(id)) return;
  if (!community_model_ || community_model_->empty()) {  // nothing to train on yet (quirk Q5)
    pending_initial_.push_back(id);
    return;
  }
  if (global_iteration_ == 0) {
    global_iteration_ = 1;
    RuntimeMetadata md;
    md.global_iteration = 1;
    md.started_at = NowSeconds();
    metadata_.push_back(std::move(md));
  }
  SendRunTasksLocked({id});
}

void Controller::SendRunTasksLocked(const std::vector<std::string>& ids) {
  std::vector<TaskItem> items;
  const double now = NowSeconds();
  for (auto& id : ids) {
    auto it = learners_.find(id);
    if (it == learners_.end()) continue;
    auto& rec = it->second;
    rec.task_issued_iteration = global_iteration_;
    rec.task_issued_at = now;
    rec.task_in_flight = true;
    if (Sampling()) cohort_.insert(id);
    items.push_back({rec.endpoint, global_iteration_, rec.num_local_updates, params_.percent_validation});
    if (!metadata_.empty()) {
      metadata_.back().assigned.push_back(id);
      metadata_.back().train_submitted_at[id] = now;
    }
  }
  if (!items.empty() && dispatcher_) {
    auto d = dispatcher_;
    auto fm = community_;
    auto ref = community_ref_;
    std::lock_guard<std::mutex> l(outbox_mu_);
    outbox_.push_back([d, items, fm, ref] { d->SendRunTasks(items, fm, ref); });
  }
}

void Controller::FlushOutbox() {
  std::vector<std::function<void()>> calls;
  {
    std::lock_guard<std::mutex> l(outbox_mu_);
    calls.swap(outbox_);
  }
  for (auto& c : calls) c();
}

void Controller::SendEvaluationTasksLocked(const std::vector<std::string>& ids, uint32_t eval_index) {
  std::vector<LearnerEndpoint> eps;
  const double now = NowSeconds();
  const uint32_t it = evaluations_[eval_index].global_iteration;
  for (auto& id : ids) {
    auto f = learners_.find(id);
    if (f == learners_.end()) continue;
    eps.push_back(f->second.endpoint);
    if (it >= 1 && it <= metadata_.size()) metadata_[it - 1].eval_submitted_at[id] = now;
  }
  if (!eps.empty() && dispatcher_) {
    auto d = dispatcher_;
    auto model = community_model_;
    auto ref = community_model_ref_;
    const uint32_t bs = params_.batch_size;
    std::lock_guard<std::mutex> l(outbox_mu_);
    outbox_.push_back([d, eps, model, ref, bs, eval_index, it] { d->SendEvaluationTasks(eps, model, ref, bs, eval_index, it); });
  }
}

void Controller::ScheduleTasks(const std::string& id) {
  {
    std::lock_guard<std::mutex> l(mu_);
    if (stop_ || !learners_.count(id)) return;
    const auto active = BarrierSetLocked();   // with client sampling: this round's cohort
    auto to_schedule = scheduler_->ScheduleNext(id, active);
    if (to_schedule.empty()) return;  // synchronous barrier not reached yet
    AfterBarrier(to_schedule, id);
  }
  FlushOutbox();
}

void Controller::AfterBarrier(const std::vector<std::string>& completed, const std::string& completing) {
  const auto active = ActiveIdsLocked();
  // without sampling the learners that reached the barrier are the ones that train next; with it, the round
  // aggregates its cohort (not the stale models of learners that sat it out) and a new cohort is drawn
  const bool sampling = Sampling();
  const auto selected = sampling ? completed : ScheduledCardinality::Select(completed, active);
  const auto to_schedule = sampling ? SampleCohortLocked(active) : completed;
  if (sampling) cohort_.clear();
  if (metadata_.empty()) return;
  RuntimeMetadata& md = metadata_.back();
  uint32_t contributors = 0;
  ModelBytes model;
  try {
    model = ComputeCommunityModel(selected, completing, md, &contributors);
  } catch (const std::exception& e) {
    // One bad upload must not stall the federation: the scheduler's barrier set is already consumed and no task is
    // in flight, so nothing would ever call back in.  Keep the previous community model, close the round and
    // re-issue tasks from it; the offending learner's next upload supersedes the bad contribution.
    Log("aggregation failed (%s); round %s keeps the previous community model", e.what(),
        std::to_string(global_iteration_));
    aggregator_->Reset();
    store_->ResetState();
    // "SKIP:" = the dispatcher declined on purpose (the device fabric is dissolving because learners are leaving)
    if (std::string(e.what()).find("SKIP:") == std::string::npos) ++aggregation_failures_;
    pending_ref_.clear();
    model.reset();
  }
  if (model && !model->empty()) {
    if (md.quantifiers.empty()) RecordModelSize(*model, md);  // the device path already counted on the GPU (K4)
    community_model_ = model;
    community_ = std::make_shared<std::string>(wire::BuildFederatedModel(contributors, global_iteration_, *model));
    if (!pending_ref_.empty()) {
      community_model_ref_ = std::make_shared<std::string>(std::move(pending_ref_));
      community_ref_ = std::make_shared<std::string>(
          wire::BuildFederatedModel(contributors, global_iteration_, *community_model_ref_));
    } else {
      community_model_ref_.reset();
      community_ref_.reset();
    }
    pending_ref_.clear();
    if (auto* fa = dynamic_cast<FedAsync*>(aggregator_.get())) fa->SetCommunity(community_model_);
    community_lineage_.push_back(community_);
    while (community_lineage_.size() > std::max<uint32_t>(1, params_.community_lineage_length)) community_lineage_.pop_front();
  }
  CommunityEvaluation ev;
  ev.global_iteration = global_iteration_;
  evaluations_.push_back(std::move(ev));
  md.completed_at = NowSeconds();
  if (params_.evaluate_community_model) SendEvaluationTasksLocked(to_schedule, (uint32_t)evaluations_.size() - 1);
  ++global_iteration_;
  if (params_.protocol == Protocol::kSemiSynchronous &&
      (global_iteration_ == 2 || params_.semi_sync_recompute))
    UpdateTaskTemplates(sampling ? active : to_schedule);
  RuntimeMetadata next;
  next.global_iteration = global_iteration_;
  next.started_at = NowSeconds();
  metadata_.push_back(std::move(next));
  SendRunTasksLocked(to_schedule);
}

ModelBytes Controller::ComputeCommunityModel(const std::vector<std::string>& selected, const std::string& completing,
                                             RuntimeMetadata& md, uint32_t* contributors) {
  const auto t_start = std::chrono::steady_clock::now();
  md.aggregation_started_at = NowSeconds();
  const bool single = params_.rule == "FedRec" || params_.rule == "FedAsync";
  // participants: the completing learner for per-contribution rules (Q2), otherwise every selected
  // learner that actually has a stored model (Q14)
  std::vector<std::string> parts;
  if (single) { if (!completing.empty()) parts.push_back(completing); }  // empty: a round closed without contributions
  else
    for (auto& id : selected)
      if (store_->GetLearnerLineageLength(id) > 0) parts.push_back(id);
  std::vector<LearnerStats> stats;
  for (auto& id : parts) {
    LearnerStats s;
    s.id = id;
    auto f = learners_.find(id);
    if (f != learners_.end()) s.num_training_examples = f->second.n_train;
    auto lt = local_tasks_.find(id);
    if (lt != local_tasks_.end() && !lt->second.empty()) s.completed_batches = lt->second.front().completed_batches;  // latest
    stats.push_back(s);
  }
  auto weights = ComputeScalingFactors(params_.scaling, learners_.size(), stats);
  if (!single) {
    // A lone participant of a larger federation gets its RAW count from the scalers (train_dataset_size_scaler.cc:
    // 7-49 - meant for the per-contribution rolling rules, which divide by the running sum z).  FedAvg / PWA - and
    // FedStride on the device path, where the blocks are plain weighted sums - multiply by the factor as it is, so a
    // round that closes with one contributor (a sampled cohort of one, a first round that a late joiner missed) would
    // publish count x model.  The factors of a round are normalised to sum 1 over its contributors (for the host
    // FedStride, which divides by z = sum of factors, this changes nothing).
    double total = 0;
    for (auto& kv : weights) total += kv.second;
    if (total > 0 && std::fabs(total - 1.0) > 1e-12)
      for (auto& kv : weights) kv.second /= total;
  }
  size_t stride = parts.size();
  if (params_.rule == "FedStride" && params_.stride_length > 0) stride = params_.stride_length;
  if (stride == 0) stride = 1;
  const int need = aggregator_->RequiredLearnerLineageLength();
  Aggregated result;
  for (size_t b = 0; b < parts.size(); b += stride) {
    const auto t_block = std::chrono::steady_clock::now();
    std::vector<std::pair<std::string, int>> req;
    for (size_t i = b; i < std::min(parts.size(), b + stride); ++i) {
      const int have = store_->GetLearnerLineageLength(parts[i]);
      req.emplace_back(parts[i], std::min(have, need));
    }
    const auto t_sel = std::chrono::steady_clock::now();
    auto models = store_->SelectModels(req);
    const double sel_ms = ElapsedMs(t_sel);
    std::vector<LearnerModels> block;
    for (auto& [id, n] : req) {
      md.selection_ms[id] = sel_ms / (double)req.size();
      auto f = models.find(id);
      if (f == models.end() || f->second.empty()) continue;
      LearnerModels lm;
      lm.lineage = f->second;
      lm.weight = weights[id];
      auto lr = learners_.find(id);
      if (lr != learners_.end())
        lm.staleness = (double)global_iteration_ - (double)lr->second.task_issued_iteration;
      block.push_back(std::move(lm));
    }
    md.block_size.push_back((double)block.size());
    const bool first_block = b == 0, last_block = b + stride >= parts.size();
    if (!AggregateBlockOnDevice(req, models, weights, first_block, last_block, md, &result)) {
      result = aggregator_->Aggregate(block);
      md.block_duration_ms.push_back(ElapsedMs(t_block));
    }
    md.block_memory_kb.push_back((double)MaxRssKb());
    store_->ResetState();
  }
  aggregator_->Reset();
  md.aggregation_completed_at = NowSeconds();
  md.aggregation_total_ms = ElapsedMs(t_start);
  if (md.device_timed) {  // report what the GPUs measured, not the RPC round trip around it
    md.aggregation_total_ms = 0;
    for (double v : md.block_duration_ms) md.aggregation_total_ms += v;
  }
  *contributors = result.num_contributors;
  return std::make_shared<std::string>(std::move(result.model));
}

static bool DeviceSlotOf(const std::string& model, wire::DeviceSlot* slot) {
  std::vector<wire::VariableView> vars;
  if (!wire::ParseModel(model, vars) || vars.empty() || !vars[0].t.slot.present) return false;
  *slot = vars[0].t.slot;
  return true;
}

bool Controller::AggregateBlockOnDevice(const std::vector<std::pair<std::string, int>>& req,
                                        const std::map<std::string, std::vector<ModelBytes>>& models,
                                        const std::map<std::string, double>& weights, bool first_block,
                                        bool last_block, RuntimeMetadata& md, Aggregated* result) {
  if (!dispatcher_) return false;
  DeviceAggregateRequest dr;
  dr.rule = params_.rule;
  dr.global_iteration = global_iteration_;
  dr.first_block = first_block;
  dr.last_block = last_block;
  bool any_device = false;
  double z = device_z_;
  std::map<std::string, double> new_w;
  uint32_t new_contributors = 0;
  for (auto& [id, n] : req) {
    auto f = models.find(id);
    if (f == models.end() || f->second.empty()) continue;
    const auto& lineage = f->second;
    DeviceSource src;
    src.learner_id = id;
    auto w = weights.find(id);
    src.weight = w == weights.end() ? 0.0 : w->second;
    wire::DeviceSlot slot;
    if (DeviceSlotOf(*lineage.back(), &slot)) {
      any_device = true;
      src.rank = slot.device_rank; src.slot = slot.slot; src.version = slot.version;
    } else {
      src.staged_model = lineage.back();
    }
    if (params_.rule == "FedRec") {
      // recency swap (federated_recency.cc:27-94): z += w_new - w_old when the learner contributed before
      auto prev = device_w_.find(id);
      wire::DeviceSlot old;
      if (prev != device_w_.end() && lineage.size() >= 2 && DeviceSlotOf(*lineage[lineage.size() - 2], &old)) {
        src.has_previous = true;
        src.previous_slot = old.slot;
        src.previous_weight = prev->second;
        z += src.weight - prev->second;
      } else if (prev != device_w_.end()) {
        return false;  // the previous contribution is not addressable on the device: host path owns this learner
      } else {
        z += src.weight;
        ++new_contributors;
      }
      new_w[id] = src.weight;
    } else if (params_.rule == "FedAsync") {
      auto lr = learners_.find(id);
      const double staleness = lr == learners_.end() ? 0.0
                                                     : (double)global_iteration_ - (double)lr->second.task_issued_iteration;
      dr.mix = std::min(1.0, std::max(0.0, params_.fedasync_alpha * FedAsync::Staleness(params_.staleness_fn, staleness,
                                                                                        params_.staleness_a,
                                                                                        params_.staleness_b)));
    }
    dr.sources.push_back(std::move(src));
  }
  if (!any_device && !dr.sources.empty() && device_contributors_ > 0 &&
      (params_.rule == "FedRec" || params_.rule == "FedAsync"))
    // the rolling state of this federation lives on the device fabric: folding a byte-carried model on the host
    // would start a second, diverging community model - refuse loudly instead (counted as an aggregation failure)
    throw std::runtime_error("asynchronous rule on a device fabric: a contribution without a device slot cannot be "
                             "folded (keep every learner of an asynchronous federation inside the fabric)");
  if (!any_device || dr.sources.empty()) return false;
  if (params_.rule == "PWA")
    // ciphertexts that live on the GPUs are reduced there (K8); one that arrived as bytes cannot join them
    for (auto& s : dr.sources)
      if (s.staged_model)
        throw std::runtime_error("PWA on a device fabric: every contribution has to be a device reference");
  dr.z = z;
  DeviceAggregateResult out;
  if (!dispatcher_->AggregateOnDevice(dr, &out)) return false;
  if (!out.error.empty()) throw std::runtime_error("device aggregation failed: " + out.error);
  if (out.model_ref.empty()) throw std::runtime_error("device aggregation returned no community model reference");
  md.device_timed = true;
  md.block_duration_ms.push_back(out.device_ms);
  if (last_block) {
    if (!out.quantifiers.empty()) md.quantifiers = out.quantifiers;
    pending_ref_ = out.model_ref;
  }
  if (params_.rule == "FedRec") {
    device_z_ = z;
    device_contributors_ += new_contributors;
    for (auto& kv : new_w) device_w_[kv.first] = kv.second;
    result->num_contributors = device_contributors_;
  } else if (params_.rule == "FedAsync") {
    result->num_contributors = ++device_contributors_;
  } else {
    result->num_contributors = (first_block ? 0 : result->num_contributors) + (uint32_t)dr.sources.size();
  }
  result->model = out.model_bytes.empty() ? out.model_ref : std::move(out.model_bytes);
  return true;
}

template <typename T> static uint32_t CountZeros(const char* p, size_t n) {
  uint32_t z = 0;
  for (size_t i = 0; i < n; ++i) z += (double(load_unaligned<T>(reinterpret_cast<const T*>(p) + i)) == 0.0);
  return z;
}

void Controller::RecordModelSize(const std::string& model, RuntimeMetadata& md) {
  std::vector<wire::VariableView> vars;
  if (!wire::ParseModel(model, vars)) return;
  md.quantifiers.assign(vars.size(), TensorQuant{});
#pragma omp parallel for schedule(dynamic)
  for (long v = 0; v < (long)vars.size(); ++v) {
    const auto& t = vars[v].t;
    TensorQuant q;
    q.bytes = (uint32_t)t.value.size();
    if (!vars[v].ciphertext && t.value.size() >= (size_t)t.length * wire::DTypeSize(t.dtype)) {
      uint32_t zeros = 0;
      try {
        MFL_HOST_DISPATCH(t.dtype, T, zeros = CountZeros<T>(t.value.data(), t.length));
      } catch (...) {}
      q.zeros = zeros;
      q.non_zeros = t.length - zeros;
    }
    md.quantifiers[v] = q;
  }
}

void Controller::UpdateTaskTemplates(const std::vector<std::string>& ids) {
  // semi-synchronous: every learner trains for t_max = lambda * (slowest learner's epoch time)
  double slowest_epoch_ms = 0;
  for (auto& id : ids) {
    auto lt = local_tasks_.find(id);
    if (lt != local_tasks_.end() && !lt->second.empty())
      slowest_epoch_ms = std::max(slowest_epoch_ms, (double)lt->second.front().ms_per_epoch);
  }
  const double t_max = (double)params_.semi_sync_lambda * slowest_epoch_ms;
  for (auto& id : ids) {
    auto lt = local_tasks_.find(id);
    auto lr = learners_.find(id);
    if (lt == local_tasks_.end() || lt->second.empty() || lr == learners_.end()) continue;
    double ms_per_batch = lt->second.front().ms_per_batch;
    if (ms_per_batch <= 0) ms_per_batch = 1.0;
    const double n = std::ceil(t_max / ms_per_batch);
    if (n >= 1) lr->second.num_local_updates = (uint32_t)n;
  }
}

// ---------------------------------------------------------------------------------------------
void Controller::WatchdogLoop() {
  while (!stop_) {
    {
      std::unique_lock<std::mutex> l(wd_mu_);
      wd_cv_.wait_for(l, std::chrono::milliseconds(200), [this] { return stop_.load(); });
    }
    if (stop_) break;
    std::lock_guard<std::mutex> l(mu_);
    const double now = NowSeconds();
    std::vector<std::string> overdue;
    for (auto& [id, rec] : learners_)
      if (rec.task_in_flight && now - rec.task_issued_at > params_.task_deadline_secs) overdue.push_back(id);
    for (auto& id : overdue) EvictLocked(id, "evicted: task deadline exceeded");
  }
}

// ---------------------------------------------------------------------------------------------
std::vector<LearnerRecord> Controller::GetLearners() {
  std::lock_guard<std::mutex> l(mu_);
  std::vector<LearnerRecord> out;
  for (auto& kv : learners_) out.push_back(kv.second);
  return out;
}

ModelBytes Controller::CommunityModel() {
  std::lock_guard<std::mutex> l(mu_);
  return community_;
}

uint32_t Controller::AggregationFailures() {
  std::lock_guard<std::mutex> l(mu_);
  return aggregation_failures_;
}

uint32_t Controller::GlobalIteration() {
  std::lock_guard<std::mutex> l(mu_);
  return global_iteration_;
}

template <typename C> static C Tail(const C& c, int n) {
  if (n <= 0 || (size_t)n >= c.size()) return c;
  return C(c.end() - n, c.end());
}

std::vector<RuntimeMetadata> Controller::GetRuntimeMetadataLineage(int n) {
  std::lock_guard<std::mutex> l(mu_);
  return Tail(metadata_, n);
}

std::vector<CommunityEvaluation> Controller::GetCommunityEvaluationLineage(int n) {
  std::lock_guard<std::mutex> l(mu_);
  return Tail(evaluations_, n);
}

std::map<std::string, std::vector<std::string>> Controller::GetLocalTaskLineage(const std::vector<std::string>& ids, int n) {
  std::lock_guard<std::mutex> l(mu_);
  std::map<std::string, std::vector<std::string>> out;
  auto add = [&](const std::string& id, const std::deque<TaskMeta>& q) {
    auto& v = out[id];
    size_t k = 0;
    for (auto& m : q) {  // newest first
      if (n > 0 && k++ >= (size_t)n) break;
      v.push_back(m.serialized);
    }
  };
  if (ids.empty()) for (auto& kv : local_tasks_) add(kv.first, kv.second);
  else for (auto& id : ids) { auto f = local_tasks_.find(id); if (f != local_tasks_.end()) add(id, f->second); }
  return out;
}

std::vector<ModelBytes> Controller::GetCommunityModelLineage(int n) {
  std::lock_guard<std::mutex> l(mu_);
  std::vector<ModelBytes> all(community_lineage_.begin(), community_lineage_.end());
  return Tail(all, n);
}

std::map<std::string, std::vector<ModelBytes>> Controller::GetLearnerLocalModelLineage(const std::vector<std::string>& ids, int n) {
  std::vector<std::pair<std::string, int>> req;
  {
    std::lock_guard<std::mutex> l(mu_);
    if (ids.empty()) for (auto& kv : learners_) req.emplace_back(kv.first, 0);
    else for (auto& id : ids) req.emplace_back(id, 0);
  }
  auto all = store_->SelectModels(req);
  store_->ResetState();
  for (auto& kv : all) kv.second = Tail(kv.second, n);
  return all;
}

}  // namespace mfl::core
