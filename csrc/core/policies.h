// Scaling functions, schedulers and the selector of the controller core.
//
// Behavioural spec: batches_scaler.cc:19-47, participants_scaler.cc:20-38,
// train_dataset_size_scaler.cc:19-47 (federation of exactly one learner -> 1; a single participant
// -> its raw count; otherwise normalised over the participating learners);
// synchronous_scheduler.h:13-35 (barrier over the currently active set; also used for
// semi-synchronous), asynchronous_scheduler.h:12-19 (reschedule the caller);
// scheduled_cardinality.h:15-30 (scheduled < 2 -> all active learners).
#pragma once
#include <map>
#include <set>
#include <string>
#include <vector>

namespace mfl::core {

enum class ScalingKind { kUnknown = 0, kNumCompletedBatches = 1, kNumParticipants = 2, kNumTrainingExamples = 3 };

struct LearnerStats {
  std::string id;
  double num_training_examples = 0;
  double completed_batches = 0;
};

inline std::map<std::string, double> ComputeScalingFactors(ScalingKind kind, size_t federation_size,
                                                           const std::vector<LearnerStats>& participants) {
  std::map<std::string, double> out;
  if (federation_size == 1) {
    for (auto& p : participants) out[p.id] = 1.0;
    return out;
  }
  if (kind == ScalingKind::kNumParticipants) {
    for (auto& p : participants) out[p.id] = 1.0 / (double)participants.size();
    return out;
  }
  auto raw = [&](const LearnerStats& p) {
    return kind == ScalingKind::kNumCompletedBatches ? p.completed_batches : p.num_training_examples;
  };
  if (participants.size() == 1) {
    out[participants[0].id] = raw(participants[0]);
    return out;
  }
  double total = 0;
  for (auto& p : participants) total += raw(p);
  for (auto& p : participants) out[p.id] = total > 0 ? raw(p) / total : 1.0 / (double)participants.size();
  return out;
}

class Scheduler {
 public:
  virtual ~Scheduler() = default;
  // learner `id` just completed its task; returns the learners to schedule next (may be empty)
  virtual std::vector<std::string> ScheduleNext(const std::string& id, const std::vector<std::string>& active) = 0;
  virtual const char* Name() const = 0;
};

class SynchronousScheduler : public Scheduler {
 public:
  std::vector<std::string> ScheduleNext(const std::string& id, const std::vector<std::string>& active) override {
    done_.insert(id);
    // the barrier compares against the CURRENT active set, so a clean leave shrinks it
    for (auto& a : active)
      if (!done_.count(a)) return {};
    std::vector<std::string> out(active.begin(), active.end());
    done_.clear();
    return out;
  }
  const char* Name() const override { return "SynchronousScheduler"; }
 private:
  std::set<std::string> done_;
};

class AsynchronousScheduler : public Scheduler {
 public:
  std::vector<std::string> ScheduleNext(const std::string& id, const std::vector<std::string>&) override { return {id}; }
  const char* Name() const override { return "AsynchronousScheduler"; }
};

class ScheduledCardinality {
 public:
  static std::vector<std::string> Select(const std::vector<std::string>& scheduled, const std::vector<std::string>& active) {
    return scheduled.size() < 2 ? active : scheduled;
  }
};

}  // namespace mfl::core
