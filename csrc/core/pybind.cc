// pybind11 module `metisfl_b200._core`: the native controller core, aggregation rules / CPU
// oracle, model stores, wire codec and host CKKS.  Counterparts in the reference:
// controller/pybind/controller_pybind.cc:51-67 (ControllerWrapper) and
// encryption/pybind/ckks_pybind.cc:71-85 (fhe.CKKS); the Python-facing classes with those exact
// names live in metisfl_b200/controller/controller.py and metisfl_b200/encryption/fhe.py.
#include <pybind11/functional.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <deque>

#include "ckks/ckks_host.h"
#include "core/aggregation.h"
#include "core/controller.h"
#include "core/policies.h"
#include "core/store.h"
#include "core/wire.h"

namespace py = pybind11;
using namespace mfl::core;
namespace wire = mfl::wire;
namespace ckks = mfl::ckks;

namespace {

ModelBytes ToModel(const py::bytes& b) { return std::make_shared<std::string>(std::string(b)); }

struct StatusError : std::runtime_error {
  Status code;
  StatusError(Status c, const std::string& m) : std::runtime_error(m), code(c) {}
};
void Raise(const StatusOr& s) {
  if (!s.ok()) throw StatusError(s.code, s.message);
}

// block = [ (lineage: list[bytes], weight: float | list[float], staleness: float = 0) ... ]
std::vector<LearnerModels> ToBlock(const py::list& block) {
  std::vector<LearnerModels> out;
  for (auto item : block) {
    py::tuple t = py::cast<py::tuple>(item);
    LearnerModels lm;
    for (auto m : py::cast<py::list>(t[0])) lm.lineage.push_back(ToModel(py::cast<py::bytes>(m)));
    if (py::isinstance<py::float_>(t[1]) || py::isinstance<py::int_>(t[1])) lm.weight = py::cast<double>(t[1]);
    else {
      lm.lineage_weights = py::cast<std::vector<double>>(t[1]);
      lm.weight = lm.lineage_weights.empty() ? 0.0 : lm.lineage_weights.back();
    }
    if (t.size() > 2) lm.staleness = py::cast<double>(t[2]);
    out.push_back(std::move(lm));
  }
  return out;
}

py::tuple DoAggregate(AggregationFunction& f, const py::list& block) {
  auto b = ToBlock(block);
  Aggregated r;
  {
    py::gil_scoped_release rel;
    r = f.Aggregate(b);
  }
  return py::make_tuple(py::bytes(r.model), r.num_contributors);
}

class PyDispatcher : public Dispatcher {
 public:
  PyDispatcher(py::object run, py::object eval, py::object agg)
      : run_(std::move(run)), eval_(std::move(eval)), agg_(std::move(agg)) {}
  ~PyDispatcher() override {
    py::gil_scoped_acquire g;
    run_ = py::object();
    eval_ = py::object();
    agg_ = py::object();
  }
  // request -> dict; the callback answers None (no fabric: aggregate on the host) or a dict with model_ref,
  // model_bytes, device_ms, quantifiers [(non_zeros, zeros, bytes)], error
  bool AggregateOnDevice(const DeviceAggregateRequest& r, DeviceAggregateResult* out) override {
    py::gil_scoped_acquire g;
    if (agg_.is_none()) return false;
    try {
      py::dict d;
      d["rule"] = r.rule; d["global_iteration"] = r.global_iteration; d["mix"] = r.mix; d["z"] = r.z;
      d["first_block"] = r.first_block; d["last_block"] = r.last_block; d["quantify"] = r.quantify;
      py::list srcs;
      for (auto& s : r.sources) {
        py::dict sd;
        sd["learner_id"] = s.learner_id; sd["device_rank"] = s.rank; sd["slot"] = s.slot; sd["version"] = s.version;
        sd["weight"] = s.weight; sd["has_previous"] = s.has_previous; sd["previous_slot"] = s.previous_slot;
        sd["previous_weight"] = s.previous_weight;
        if (s.staged_model) sd["staged_model"] = py::bytes(*s.staged_model);
        else sd["staged_model"] = py::none();
        srcs.append(sd);
      }
      d["sources"] = srcs;
      py::object res = agg_(d);
      if (res.is_none()) return false;
      py::dict rd = py::cast<py::dict>(res);
      if (rd.contains("error") && !rd["error"].is_none()) out->error = py::cast<std::string>(rd["error"]);
      if (rd.contains("model_ref")) out->model_ref = std::string(py::cast<py::bytes>(rd["model_ref"]));
      if (rd.contains("model_bytes") && !rd["model_bytes"].is_none())
        out->model_bytes = std::string(py::cast<py::bytes>(rd["model_bytes"]));
      if (rd.contains("device_ms")) out->device_ms = py::cast<double>(rd["device_ms"]);
      if (rd.contains("quantifiers"))
        for (auto q : py::cast<py::list>(rd["quantifiers"])) {
          py::tuple t = py::cast<py::tuple>(q);
          out->quantifiers.push_back({py::cast<uint32_t>(t[0]), py::cast<uint32_t>(t[1]), py::cast<uint32_t>(t[2])});
        }
      return true;
    } catch (py::error_already_set& e) {
      out->error = e.what();
      return true;
    } catch (const std::exception& e) {
      out->error = e.what();
      return true;
    }
  }
  void SendRunTasks(const std::vector<TaskItem>& items, const ModelBytes& fm, const ModelBytes& ref) override {
    py::gil_scoped_acquire g;
    try {
      py::list l;
      for (auto& it : items)
        l.append(py::make_tuple(it.learner.id, it.learner.hostname, it.learner.port, py::bytes(it.learner.public_certificate),
                                it.global_iteration, it.num_local_updates, it.percent_validation));
      if (ref) run_(l, py::bytes(*fm), py::bytes(*ref));  // the control-only twin only exists on the device path
      else run_(l, py::bytes(*fm));
    } catch (py::error_already_set& e) {
      e.discard_as_unraisable("metisfl_b200 dispatcher (run tasks)");
    }
  }
  void SendEvaluationTasks(const std::vector<LearnerEndpoint>& learners, const ModelBytes& model,
                           const ModelBytes& ref, uint32_t batch_size, uint32_t eval_index,
                           uint32_t global_iteration) override {
    py::gil_scoped_acquire g;
    try {
      py::list l;
      for (auto& ep : learners) l.append(py::make_tuple(ep.id, ep.hostname, ep.port, py::bytes(ep.public_certificate)));
      if (ref) eval_(l, py::bytes(*model), batch_size, eval_index, global_iteration, py::bytes(*ref));
      else eval_(l, py::bytes(*model), batch_size, eval_index, global_iteration);
    } catch (py::error_already_set& e) {
      e.discard_as_unraisable("metisfl_b200 dispatcher (evaluation tasks)");
    }
  }
 private:
  py::object run_, eval_, agg_;
};

py::dict MetaToDict(const RuntimeMetadata& m) {
  py::dict d;
  d["global_iteration"] = m.global_iteration;
  d["started_at"] = m.started_at; d["completed_at"] = m.completed_at;
  d["assigned_to_learner_id"] = m.assigned; d["completed_by_learner_id"] = m.completed;
  d["train_task_submitted_at"] = m.train_submitted_at; d["train_task_received_at"] = m.train_received_at;
  d["eval_task_submitted_at"] = m.eval_submitted_at; d["eval_task_received_at"] = m.eval_received_at;
  d["model_insertion_duration_ms"] = m.insertion_ms; d["model_selection_duration_ms"] = m.selection_ms;
  d["model_aggregation_started_at"] = m.aggregation_started_at;
  d["model_aggregation_completed_at"] = m.aggregation_completed_at;
  d["model_aggregation_total_duration_ms"] = m.aggregation_total_ms;
  d["model_aggregation_block_size"] = m.block_size;
  d["model_aggregation_block_memory_kb"] = m.block_memory_kb;
  d["model_aggregation_block_duration_ms"] = m.block_duration_ms;
  py::list q;
  for (auto& t : m.quantifiers) q.append(py::make_tuple(t.non_zeros, t.zeros, t.bytes));
  d["model_tensor_quantifiers"] = q;
  d["device_timed"] = m.device_timed;
  return d;
}

py::list ModelsToList(const std::vector<ModelBytes>& v) {
  py::list l;
  for (auto& m : v) l.append(py::bytes(*m));
  return l;
}

}  // namespace

PYBIND11_MODULE(_core, m) {
  m.doc() = "metisfl_b200 native controller core";
  static py::exception<StatusError> exc(m, "ControllerError");
  py::register_exception_translator([](std::exception_ptr p) {
    try {
      if (p) std::rethrow_exception(p);
    } catch (const StatusError& e) {
      py::object err = py::reinterpret_steal<py::object>(PyObject_CallFunction(exc.ptr(), "s", e.what()));
      err.attr("code") = (int)e.code;
      PyErr_SetObject(exc.ptr(), err.ptr());
    }
  });
  py::enum_<Status>(m, "Status")
      .value("OK", Status::kOk).value("INVALID_ARGUMENT", Status::kInvalidArgument)
      .value("ALREADY_EXISTS", Status::kAlreadyExists).value("NOT_FOUND", Status::kNotFound)
      .value("UNAUTHENTICATED", Status::kUnauthenticated).value("FAILED_PRECONDITION", Status::kFailedPrecondition)
      .value("INTERNAL", Status::kInternal);

  // ---- wire codec ---------------------------------------------------------------------------
  m.def("parse_model", [](const py::bytes& b) {
    std::string s(b);
    std::vector<wire::VariableView> vars;
    if (!wire::ParseModel(s, vars)) throw std::runtime_error("malformed Model bytes");
    py::list out;
    for (auto& v : vars) {
      py::dict d;
      // names come off the wire: bytes that are not UTF-8 (a corrupted or hostile message) must not turn into an
      // exception half-way through the list
      d["name"] = py::reinterpret_steal<py::object>(PyUnicode_DecodeUTF8(v.name.data(), (Py_ssize_t)v.name.size(), "replace"));
      d["trainable"] = v.trainable; d["ciphertext"] = v.ciphertext;
      d["length"] = v.t.length; d["dimensions"] = v.t.dims; d["dtype"] = v.t.dtype;
      d["byte_order"] = v.t.byte_order; d["fortran_order"] = v.t.fortran;
      d["value"] = py::bytes(v.t.value.data(), v.t.value.size());
      if (v.t.slot.present)
        d["device_slot"] = py::make_tuple(v.t.slot.device_rank, v.t.slot.slot, v.t.slot.offset_bytes, v.t.slot.version);
      out.append(d);
    }
    return out;
  });
  m.def("build_model", [](const py::list& vars) {
    // values are taken through the buffer protocol (bytes, numpy arrays, memoryviews): the tensor bytes are copied
    // exactly once, from the caller's buffer into the serialized model
    std::string out;
    std::deque<std::string> names;
    std::vector<py::buffer_info> bufs;
    std::vector<wire::VariableView> views;
    size_t total = 0;
    for (auto item : vars) {
      py::dict d = py::cast<py::dict>(item);
      wire::VariableView v;
      names.push_back(py::cast<std::string>(d["name"]));
      v.name = names.back();
      v.trainable = py::cast<bool>(d["trainable"]);
      v.ciphertext = d.contains("ciphertext") && py::cast<bool>(d["ciphertext"]);
      v.t.length = py::cast<uint32_t>(d["length"]);
      v.t.dims = py::cast<std::vector<int64_t>>(d["dimensions"]);
      v.t.dtype = py::cast<int>(d["dtype"]);
      v.t.byte_order = d.contains("byte_order") ? py::cast<int>(d["byte_order"]) : 2;
      v.t.fortran = d.contains("fortran_order") && py::cast<bool>(d["fortran_order"]);
      py::buffer_info info = py::cast<py::buffer>(d["value"]).request();
      if (info.ndim > 1) throw std::runtime_error("build_model: tensor values must be flat, contiguous buffers");
      if (info.ndim == 1 && info.shape[0] > 1 && info.strides[0] != info.itemsize)
        throw std::runtime_error("build_model: tensor values must be contiguous");
      total += (size_t)info.size * (size_t)info.itemsize + v.name.size() + 64;
      bufs.push_back(std::move(info));
      if (d.contains("device_slot")) {
        py::tuple t = py::cast<py::tuple>(d["device_slot"]);
        v.t.slot = {true, py::cast<uint32_t>(t[0]), py::cast<uint32_t>(t[1]), py::cast<uint64_t>(t[2]), py::cast<uint64_t>(t[3])};
      }
      views.push_back(std::move(v));
    }
    out.reserve(total);
    for (size_t i = 0; i < views.size(); ++i) {
      std::string_view val(static_cast<const char*>(bufs[i].ptr), (size_t)bufs[i].size * (size_t)bufs[i].itemsize);
      wire::AppendVariable(out, views[i], &val);
    }
    return py::bytes(out);
  });
  m.def("model_payload_bytes", [](const py::buffer& b) -> py::object {
    // tensor bytes inside a serialized Model (0 for a control-only model of device-slot references); None if malformed
    py::buffer_info info = b.request();
    if (info.ndim > 1 || (info.ndim == 1 && info.shape[0] > 1 && info.strides[0] != info.itemsize)) return py::none();
    std::string_view sv(static_cast<const char*>(info.ptr), (size_t)info.size * (size_t)info.itemsize);
    std::vector<wire::VariableView> vars;
    if (!wire::ParseModel(sv, vars)) return py::none();
    uint64_t n = 0;
    for (auto& v : vars) n += v.t.value.size();
    return py::int_(n);
  });
  m.def("build_federated_model", [](uint32_t contributors, uint32_t iteration, const py::bytes& model) {
    return py::bytes(wire::BuildFederatedModel(contributors, iteration, std::string(model)));
  });
  m.def("parse_federated_model", [](const py::bytes& b) {
    std::string s(b);
    uint32_t c, i; std::string_view model;
    if (!wire::ParseFederatedModel(s, c, i, model)) throw std::runtime_error("malformed FederatedModel bytes");
    return py::make_tuple(c, i, py::bytes(model.data(), model.size()));
  });

  // ---- tensor math oracle -------------------------------------------------------------------
  m.def("weighted_sum_bytes", [](int dtype, const std::vector<py::bytes>& srcs, const std::vector<double>& w) {
    std::vector<std::string> hold;
    for (auto& s : srcs) hold.emplace_back(std::string(s));
    if (hold.empty()) throw std::invalid_argument("no sources");
    const size_t n = hold[0].size() / wire::DTypeSize(dtype);
    std::vector<const char*> p;
    for (auto& s : hold) p.push_back(s.data());
    std::string out(hold[0].size(), '\0');
    WeightedSumBytes(dtype, n, p, w, out.data(), false);
    return py::bytes(out);
  });

  // ---- aggregation rules --------------------------------------------------------------------
  py::class_<AggregationFunction>(m, "AggregationFunction")
      .def("aggregate", &DoAggregate)
      .def("reset", &AggregationFunction::Reset)
      .def("required_learner_lineage_length", &AggregationFunction::RequiredLearnerLineageLength)
      .def("name", &AggregationFunction::Name);
  py::class_<FedAvg, AggregationFunction>(m, "FedAvg").def(py::init<>());
  py::class_<FedStride, AggregationFunction>(m, "FedStride").def(py::init<>());
  py::class_<FedRec, AggregationFunction>(m, "FedRec").def(py::init<>());
  py::class_<FedAsync, AggregationFunction>(m, "FedAsync")
      .def(py::init<double, int, double, double>(), py::arg("alpha") = 0.5, py::arg("staleness_fn") = 1,
           py::arg("a") = 0.5, py::arg("b") = 4.0)
      .def("set_community", [](FedAsync& f, const py::bytes& b) { f.SetCommunity(ToModel(b)); })
      .def_static("staleness", &FedAsync::Staleness);
  py::class_<PWA, AggregationFunction>(m, "PWA")
      .def(py::init<const std::string&, uint32_t, uint32_t>(), py::arg("crypto_context_file") = "",
           py::arg("batch_size") = 4096, py::arg("scaling_factor_bits") = 52);

  // ---- policies -----------------------------------------------------------------------------
  m.def("compute_scaling_factors", [](int kind, size_t federation_size, const std::vector<std::tuple<std::string, double, double>>& parts) {
    std::vector<LearnerStats> s;
    for (auto& [id, ex, b] : parts) s.push_back({id, ex, b});
    return ComputeScalingFactors((ScalingKind)kind, federation_size, s);
  });
  py::class_<Scheduler>(m, "Scheduler").def("schedule_next", &Scheduler::ScheduleNext).def("name", &Scheduler::Name);
  py::class_<SynchronousScheduler, Scheduler>(m, "SynchronousScheduler").def(py::init<>());
  py::class_<AsynchronousScheduler, Scheduler>(m, "AsynchronousScheduler").def(py::init<>());
  m.def("scheduled_cardinality_select", &ScheduledCardinality::Select);

  // ---- stores -------------------------------------------------------------------------------
  // store calls may block on a socket (Redis): never hold the GIL across them
  py::class_<ModelStore>(m, "ModelStore")
      .def("insert_model", [](ModelStore& s, const std::string& id, const py::bytes& b) {
        ModelBytes mb = ToModel(b);
        py::gil_scoped_release rel;
        s.InsertModel(id, std::move(mb));
      })
      .def("select_models", [](ModelStore& s, const std::vector<std::pair<std::string, int>>& req) {
        std::map<std::string, std::vector<ModelBytes>> r;
        { py::gil_scoped_release rel; r = s.SelectModels(req); }
        py::dict d;
        for (auto& kv : r) d[py::str(kv.first)] = ModelsToList(kv.second);
        return d;
      })
      .def("erase_models", &ModelStore::EraseModels, py::call_guard<py::gil_scoped_release>())
      .def("expunge", &ModelStore::Expunge, py::call_guard<py::gil_scoped_release>())
      .def("reset_state", &ModelStore::ResetState, py::call_guard<py::gil_scoped_release>())
      .def("get_learner_lineage_length", &ModelStore::GetLearnerLineageLength, py::call_guard<py::gil_scoped_release>())
      .def("shutdown", &ModelStore::Shutdown, py::call_guard<py::gil_scoped_release>())
      .def("name", &ModelStore::Name);
  auto mk_policy = [](bool evict, uint32_t len) { EvictionPolicy p; p.lineage_length_eviction = evict; p.lineage_length = len; return p; };
  py::class_<HashMapModelStore, ModelStore>(m, "HashMapModelStore")
      .def(py::init([mk_policy](bool evict, uint32_t len) { return new HashMapModelStore(mk_policy(evict, len)); }),
           py::arg("lineage_length_eviction") = false, py::arg("lineage_length") = 0);
  py::class_<RedisModelStore, ModelStore>(m, "RedisModelStore")
      .def(py::init([mk_policy](const std::string& host, int port, bool evict, uint32_t len) {
             py::gil_scoped_release rel;
             return new RedisModelStore(host, port, mk_policy(evict, len));
           }),
           py::arg("hostname") = "127.0.0.1", py::arg("port") = 6379, py::arg("lineage_length_eviction") = false,
           py::arg("lineage_length") = 0);

  // ---- controller ---------------------------------------------------------------------------
  py::class_<ControllerParams>(m, "ControllerParams")
      .def(py::init<>())
      .def_readwrite("hostname", &ControllerParams::hostname).def_readwrite("port", &ControllerParams::port)
      .def_readwrite("rule", &ControllerParams::rule).def_readwrite("stride_length", &ControllerParams::stride_length)
      .def_readwrite("fedasync_alpha", &ControllerParams::fedasync_alpha)
      .def_readwrite("staleness_fn", &ControllerParams::staleness_fn)
      .def_readwrite("staleness_a", &ControllerParams::staleness_a).def_readwrite("staleness_b", &ControllerParams::staleness_b)
      .def_property("scaling", [](ControllerParams& p) { return (int)p.scaling; }, [](ControllerParams& p, int v) { p.scaling = (ScalingKind)v; })
      .def_property("protocol", [](ControllerParams& p) { return (int)p.protocol; }, [](ControllerParams& p, int v) { p.protocol = (Protocol)v; })
      .def_readwrite("semi_sync_lambda", &ControllerParams::semi_sync_lambda)
      .def_readwrite("semi_sync_recompute", &ControllerParams::semi_sync_recompute)
      .def_readwrite("task_deadline_secs", &ControllerParams::task_deadline_secs)
      .def_readwrite("redis", &ControllerParams::redis).def_readwrite("redis_host", &ControllerParams::redis_host)
      .def_readwrite("redis_port", &ControllerParams::redis_port)
      .def_property("lineage_length_eviction", [](ControllerParams& p) { return p.eviction.lineage_length_eviction; },
                    [](ControllerParams& p, bool v) { p.eviction.lineage_length_eviction = v; })
      .def_property("lineage_length", [](ControllerParams& p) { return p.eviction.lineage_length; },
                    [](ControllerParams& p, uint32_t v) { p.eviction.lineage_length = v; })
      .def_readwrite("batch_size", &ControllerParams::batch_size).def_readwrite("epochs", &ControllerParams::epochs)
      .def_readwrite("percent_validation", &ControllerParams::percent_validation)
      .def_readwrite("participation_ratio", &ControllerParams::participation_ratio)
      .def_readwrite("sampling_seed", &ControllerParams::sampling_seed)
      .def_readwrite("he_crypto_context_file", &ControllerParams::he_crypto_context_file)
      .def_readwrite("he_batch_size", &ControllerParams::he_batch_size)
      .def_readwrite("he_scaling_bits", &ControllerParams::he_scaling_bits)
      .def_readwrite("community_lineage_length", &ControllerParams::community_lineage_length)
      .def_readwrite("evaluate_community_model", &ControllerParams::evaluate_community_model);

  py::class_<Controller>(m, "Controller")
      .def(py::init([](const ControllerParams& p, py::object run_cb, py::object eval_cb, py::object agg_cb) {
             std::shared_ptr<Dispatcher> d;
             if (!run_cb.is_none()) d = std::make_shared<PyDispatcher>(run_cb, eval_cb, agg_cb);
             return new Controller(p, d);
           }),
           py::arg("params"), py::arg("send_run_tasks") = py::none(), py::arg("send_evaluation_tasks") = py::none(),
           py::arg("aggregate_on_device") = py::none())
      .def("replace_community_model", [](Controller& c, const py::bytes& b) {
        std::string s(b);
        py::gil_scoped_release rel;
        Raise(c.ReplaceCommunityModel(std::move(s)));
      })
      .def("add_learner", [](Controller& c, const std::string& host, int port, const py::bytes& se, const py::bytes& ds,
                             uint32_t n_train, uint32_t n_valid, uint32_t n_test, const py::bytes& cert) {
        std::string id, tok, ses(se), dss(ds), certs(cert);
        {
          py::gil_scoped_release rel;
          Raise(c.AddLearner(host, port, std::move(ses), std::move(dss), n_train, n_valid, n_test, std::move(certs), &id, &tok));
        }
        return py::make_tuple(id, tok);
      }, py::arg("hostname"), py::arg("port"), py::arg("server_entity") = py::bytes(), py::arg("dataset_spec") = py::bytes(),
         py::arg("num_training_examples") = 0, py::arg("num_validation_examples") = 0, py::arg("num_test_examples") = 0,
         py::arg("public_certificate") = py::bytes())
      .def("remove_learner", [](Controller& c, const std::string& id, const std::string& tok) {
        py::gil_scoped_release rel;
        Raise(c.RemoveLearner(id, tok));
      })
      .def("learner_completed_task", [](Controller& c, const std::string& id, const std::string& tok, const py::buffer& model,
                                        uint32_t global_iteration, float completed_epochs, uint32_t completed_batches,
                                        uint32_t batch_size, float ms_per_epoch, float ms_per_batch, const py::bytes& meta) {
        TaskMeta tm;
        tm.global_iteration = global_iteration; tm.completed_epochs = completed_epochs; tm.completed_batches = completed_batches;
        tm.batch_size = batch_size; tm.ms_per_epoch = ms_per_epoch; tm.ms_per_batch = ms_per_batch;
        tm.serialized = std::string(meta);
        // any contiguous byte buffer (bytes, or the memoryview slice of the request it arrived in): one copy, into the
        // string the model store keeps
        py::buffer_info mi = model.request();
        if (mi.ndim > 1 || (mi.ndim == 1 && mi.shape[0] > 1 && mi.strides[0] != mi.itemsize))
          throw std::runtime_error("learner_completed_task: the model must be a contiguous byte buffer");
        std::string ms(static_cast<const char*>(mi.ptr), (size_t)mi.size * (size_t)mi.itemsize);
        py::gil_scoped_release rel;
        Raise(c.LearnerCompletedTask(id, tok, std::move(ms), std::move(tm)));
      }, py::arg("learner_id"), py::arg("auth_token"), py::arg("model"), py::arg("global_iteration") = 0,
         py::arg("completed_epochs") = 0.f, py::arg("completed_batches") = 0, py::arg("batch_size") = 0,
         py::arg("processing_ms_per_epoch") = 0.f, py::arg("processing_ms_per_batch") = 0.f,
         py::arg("task_execution_metadata") = py::bytes())
      .def("evaluation_completed", [](Controller& c, const std::string& id, uint32_t idx, const py::bytes& ev) {
        std::string s(ev);
        py::gil_scoped_release rel;
        c.EvaluationCompleted(id, idx, std::move(s));
      })
      .def("task_failed", &Controller::TaskFailed, py::call_guard<py::gil_scoped_release>())
      .def("evict_learner", &Controller::EvictLearner, py::call_guard<py::gil_scoped_release>())
      .def("get_learners", [](Controller& c) {
        std::vector<LearnerRecord> recs;
        { py::gil_scoped_release rel; recs = c.GetLearners(); }
        py::list l;
        for (auto& r : recs) {
          py::dict d;
          d["id"] = r.endpoint.id; d["hostname"] = r.endpoint.hostname; d["port"] = r.endpoint.port;
          d["auth_token"] = r.auth_token; d["server_entity"] = py::bytes(r.server_entity);
          d["dataset_spec"] = py::bytes(r.dataset_spec); d["num_training_examples"] = r.n_train;
          d["num_local_updates"] = r.num_local_updates;
          d["public_certificate"] = py::bytes(r.endpoint.public_certificate);
          l.append(d);
        }
        return l;
      })
      .def("community_model", [](Controller& c) -> py::object {
        ModelBytes m;
        { py::gil_scoped_release rel; m = c.CommunityModel(); }
        if (!m) return py::none();
        return py::bytes(*m);
      })
      .def("global_iteration", &Controller::GlobalIteration, py::call_guard<py::gil_scoped_release>())
      .def("aggregation_failures", &Controller::AggregationFailures, py::call_guard<py::gil_scoped_release>())
      .def("get_runtime_metadata_lineage", [](Controller& c, int n) {
        std::vector<RuntimeMetadata> v;
        { py::gil_scoped_release rel; v = c.GetRuntimeMetadataLineage(n); }
        py::list l;
        for (auto& md : v) l.append(MetaToDict(md));
        return l;
      })
      .def("get_community_evaluation_lineage", [](Controller& c, int n) {
        std::vector<CommunityEvaluation> v;
        { py::gil_scoped_release rel; v = c.GetCommunityEvaluationLineage(n); }
        py::list l;
        for (auto& e : v) {
          py::dict ev;
          for (auto& kv : e.evaluations) ev[py::str(kv.first)] = py::bytes(kv.second);
          l.append(py::make_tuple(e.global_iteration, ev));
        }
        return l;
      })
      .def("get_local_task_lineage", [](Controller& c, const std::vector<std::string>& ids, int n) {
        std::map<std::string, std::vector<std::string>> v;
        { py::gil_scoped_release rel; v = c.GetLocalTaskLineage(ids, n); }
        py::dict d;
        for (auto& kv : v) {
          py::list l;
          for (auto& s : kv.second) l.append(py::bytes(s));
          d[py::str(kv.first)] = l;
        }
        return d;
      })
      .def("get_community_model_lineage", [](Controller& c, int n) {
        std::vector<ModelBytes> v;
        { py::gil_scoped_release rel; v = c.GetCommunityModelLineage(n); }
        return ModelsToList(v);
      })
      .def("get_learner_local_model_lineage", [](Controller& c, const std::vector<std::string>& ids, int n) {
        std::map<std::string, std::vector<ModelBytes>> v;
        { py::gil_scoped_release rel; v = c.GetLearnerLocalModelLineage(ids, n); }
        py::dict d;
        for (auto& kv : v) d[py::str(kv.first)] = ModelsToList(kv.second);
        return d;
      })
      .def("wait_idle", &Controller::WaitIdle, py::call_guard<py::gil_scoped_release>())
      .def("shutdown", &Controller::Shutdown, py::call_guard<py::gil_scoped_release>());

  // ---- host CKKS ----------------------------------------------------------------------------
  py::class_<ckks::Params>(m, "CKKSParams")
      .def(py::init<>())
      .def_readwrite("log_n", &ckks::Params::log_n).def_readwrite("moduli", &ckks::Params::q)
      .def_readwrite("scale_bits", &ckks::Params::scale_bits).def_readwrite("batch_size", &ckks::Params::batch_size);
  m.def("ckks_save_params", &ckks::SaveParamsFile);
  m.def("ckks_load_params", [](const std::string& path) {
    ckks::Params p;
    if (!ckks::LoadParamsFile(path, p)) throw std::runtime_error("cannot read crypto context " + path);
    return p;
  });
  py::class_<ckks::Scheme>(m, "CKKSScheme")
      .def(py::init<const ckks::Params&>())
      .def("keygen", &ckks::Scheme::KeyGen, py::arg("seed") = 0, py::call_guard<py::gil_scoped_release>())
      .def("has_public_key", &ckks::Scheme::has_public_key).def("has_secret_key", &ckks::Scheme::has_secret_key)
      .def("public_key_bytes", [](ckks::Scheme& s) { return py::bytes(s.SerializePublicKey()); })
      .def("secret_key_bytes", [](ckks::Scheme& s) { return py::bytes(s.SerializeSecretKey()); })
      .def("eval_mult_key_bytes", [](ckks::Scheme& s) { return py::bytes(s.SerializeEvalMultKey()); })
      .def("load_public_key", [](ckks::Scheme& s, const py::bytes& b) { return s.LoadPublicKey(std::string(b)); })
      .def("load_secret_key", [](ckks::Scheme& s, const py::bytes& b) { return s.LoadSecretKey(std::string(b)); })
      .def("public_key_words", [](ckks::Scheme& s) { return py::array_t<uint64_t>(s.public_key().size(), s.public_key().data()); })
      .def("secret_key_words", [](ckks::Scheme& s) { return py::array_t<uint64_t>(s.secret_key().size(), s.secret_key().data()); })
      .def("ntt_tables", [](ckks::Scheme& s, int limb) {
        const auto& t = s.tables().at(limb);
        py::dict d;
        d["q"] = t.q; d["psi"] = t.psi; d["n_inv"] = t.n_inv; d["n_inv_shoup"] = t.n_inv_shoup;
        d["psi_br"] = py::array_t<uint64_t>(t.psi_br.size(), t.psi_br.data());
        d["psi_br_shoup"] = py::array_t<uint64_t>(t.psi_br_shoup.size(), t.psi_br_shoup.data());
        d["ipsi_br"] = py::array_t<uint64_t>(t.ipsi_br.size(), t.ipsi_br.data());
        d["ipsi_br_shoup"] = py::array_t<uint64_t>(t.ipsi_br_shoup.size(), t.ipsi_br_shoup.data());
        return d;
      })
      .def("encode_coeffs", [](ckks::Scheme& s, py::array_t<double, py::array::c_style | py::array::forcecast> v, double scale) {
        std::vector<double> c;
        s.EncodeCoeffs(v.data(), (size_t)v.size(), scale, c);
        return py::array_t<double>(c.size(), c.data());
      })
      .def("decode_coeffs", [](ckks::Scheme& s, py::array_t<double, py::array::c_style | py::array::forcecast> c, size_t n) {
        std::vector<double> cv(c.data(), c.data() + c.size()), out(n);
        s.DecodeCoeffs(cv, n, out.data());
        return py::array_t<double>(out.size(), out.data());
      })
      .def("encrypt", [](ckks::Scheme& s, py::array_t<double, py::array::c_style | py::array::forcecast> v, uint64_t seed) {
        std::string out;
        const double* p = v.data();
        const size_t n = (size_t)v.size();
        { py::gil_scoped_release rel; out = s.Encrypt(p, n, seed); }
        return py::bytes(out);
      }, py::arg("values"), py::arg("seed") = 0)
      .def("decrypt", [](ckks::Scheme& s, const py::bytes& b, size_t n) {
        std::string blob(b);
        std::vector<double> out;
        { py::gil_scoped_release rel; out = s.Decrypt(blob, n); }
        return py::array_t<double>(out.size(), out.data());
      });
  m.def("ckks_weighted_average", [](const std::vector<py::bytes>& cts, const std::vector<double>& w) {
    std::vector<std::string> hold;
    for (auto& c : cts) hold.emplace_back(std::string(c));
    std::vector<std::string_view> v(hold.begin(), hold.end());
    std::string out;
    { py::gil_scoped_release rel; out = ckks::WeightedAverage(v, w); }
    return py::bytes(out);
  });
  m.def("ckks_rescale", [](const py::bytes& b) {
    std::string blob(b), out;
    { py::gil_scoped_release rel; out = ckks::Rescale(blob); }
    return py::bytes(out);
  });
  m.def("ckks_header", [](const py::bytes& b) {
    std::string blob(b);
    ckks::CtHeader h;
    if (!ckks::ParseHeader(blob, h)) throw std::runtime_error("malformed ciphertext blob");
    py::dict d;
    d["log_n"] = h.log_n; d["n_limbs"] = h.n_limbs; d["n_ct"] = h.n_ct; d["slots"] = h.slots;
    d["n_values"] = h.n_values; d["scale"] = h.scale;
    d["moduli"] = std::vector<uint64_t>(h.q, h.q + h.n_limbs);
    return d;
  });
}
