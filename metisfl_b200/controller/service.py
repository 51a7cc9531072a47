"""gRPC face of the controller: ``ControllerService`` (12 RPCs) on top of the native core.

Counterpart of the reference's C++ servicer (``controller/core/controller_servicer.cc:29-390``) and
of the async RunTask / EvaluateModel fan-out of its core (``controller.cc:571-759``).  The native
core (``metisfl_b200._core.Controller``) owns all federation state; this layer only (de)serialises
protobuf, maps status codes, and moves bytes over gRPC.  Unlike the reference, the two lineage RPCs
that it declares but never implements (``GetCommunityModelLineage``,
``GetLearnerLocalModelLineage``; reference gives UNIMPLEMENTED) are served.
"""
from __future__ import annotations

import concurrent.futures
import json
import os
import threading

import grpc
from google.protobuf.timestamp_pb2 import Timestamp

from metisfl_b200 import _core
from metisfl_b200.proto import (controller_pb2, controller_pb2_grpc, learner_pb2, learner_pb2_grpc, metis_pb2,
                                model_pb2, service_common_pb2)
from metisfl_b200.proto import envelope
from metisfl_b200.utils.grpc_services import GRPCChannelMaxMsgLength, GRPCServerMaxMsgLength
from metisfl_b200.utils.metis_logger import MetisLogger

_STATUS_TO_GRPC = {
    int(_core.Status.INVALID_ARGUMENT): grpc.StatusCode.INVALID_ARGUMENT,
    int(_core.Status.ALREADY_EXISTS): grpc.StatusCode.ALREADY_EXISTS,
    int(_core.Status.NOT_FOUND): grpc.StatusCode.NOT_FOUND,
    int(_core.Status.UNAUTHENTICATED): grpc.StatusCode.UNAUTHENTICATED,
    int(_core.Status.FAILED_PRECONDITION): grpc.StatusCode.FAILED_PRECONDITION,
    int(_core.Status.INTERNAL): grpc.StatusCode.INTERNAL,
}


def _ts(seconds: float) -> Timestamp:
    t = Timestamp()
    if seconds:
        t.FromNanoseconds(int(seconds * 1e9))
    return t


def _tensor_payload_bytes(model_bytes) -> int:
    """Tensor bytes inside a serialized ``Model`` (a control-only model with device-slot references has none)."""
    return int(_core.model_payload_bytes(model_bytes) or 0)


def _ack(status=True, message="") -> service_common_pb2.Ack:
    t = Timestamp()
    t.GetCurrentTime()
    return service_common_pb2.Ack(status=status, timestamp=t, message=message)


def params_pb_to_core(pb: metis_pb2.ControllerParams) -> "_core.ControllerParams":
    """``metisfl.ControllerParams`` -> native parameter struct (factories of
    ``controller_utils.cc:13-70``)."""
    p = _core.ControllerParams()
    p.hostname = pb.server_entity.hostname or "localhost"
    p.port = pb.server_entity.port or 50051
    rule = pb.global_model_specs.aggregation_rule
    which = rule.WhichOneof("rule") or "fed_avg"
    p.rule = {"fed_avg": "FedAvg", "fed_stride": "FedStride", "fed_rec": "FedRec", "pwa": "PWA",
              "fed_async": "FedAsync"}[which]
    if which == "fed_stride":
        p.stride_length = rule.fed_stride.stride_length
    if which == "fed_async":
        p.fedasync_alpha = rule.fed_async.alpha or 0.5
        p.staleness_fn = int(rule.fed_async.staleness_fn)
        p.staleness_a = rule.fed_async.a or 0.5
        p.staleness_b = rule.fed_async.b or 4.0
    if which == "pwa":
        he = rule.pwa.he_scheme_config
        p.he_crypto_context_file = he.crypto_context_file
        if he.WhichOneof("config") == "ckks_scheme_config":
            p.he_batch_size = he.ckks_scheme_config.batch_size
            p.he_scaling_bits = he.ckks_scheme_config.scaling_factor_bits
    p.scaling = int(rule.aggregation_rule_specs.scaling_factor) or 3
    p.participation_ratio = pb.global_model_specs.learners_participation_ratio or 1.0
    p.protocol = int(pb.communication_specs.protocol) or 1
    p.semi_sync_lambda = pb.communication_specs.protocol_specs.semi_sync_lambda
    p.semi_sync_recompute = pb.communication_specs.protocol_specs.semi_sync_recompute_num_updates
    p.task_deadline_secs = pb.communication_specs.protocol_specs.task_deadline_secs
    store = pb.model_store_config
    kind = store.WhichOneof("config")
    specs = store.redis_db_store.model_store_specs if kind == "redis_db_store" else store.in_memory_store.model_store_specs
    if kind == "redis_db_store":
        p.redis = True
        p.redis_host = store.redis_db_store.server_entity.hostname or "127.0.0.1"
        p.redis_port = store.redis_db_store.server_entity.port or 6379
    if specs.WhichOneof("eviction_policy") == "lineage_length_eviction":
        p.lineage_length_eviction = True
        p.lineage_length = specs.lineage_length_eviction.lineage_length
    # rules that swap a learner's previous model need it to still be stored
    need = 2 if p.rule == "FedRec" else 1
    if p.lineage_length_eviction and max(1, p.lineage_length) < need:
        p.lineage_length = need
    p.batch_size = pb.model_hyperparams.batch_size
    p.epochs = pb.model_hyperparams.epochs
    p.percent_validation = pb.model_hyperparams.percent_validation
    return p


class ControllerServicer(controller_pb2_grpc.ControllerServiceServicer):

    def __init__(self, params_pb: metis_pb2.ControllerParams, learner_client_workers: int = 16):
        self.params_pb = params_pb
        self._hyper = metis_pb2.Hyperparameters(batch_size=params_pb.model_hyperparams.batch_size,
                                                optimizer=params_pb.model_hyperparams.optimizer)
        self._rpc_pool = concurrent.futures.ThreadPoolExecutor(max_workers=learner_client_workers,
                                                               thread_name_prefix="ctrl-fanout")
        self._channels = {}
        self._channels_lock = threading.Lock()
        self._agg_pool = concurrent.futures.ThreadPoolExecutor(max_workers=16, thread_name_prefix="ctrl-aggregate")
        self._joined = set()
        # on-box fabrics announced by joining learners: fabric_id -> {world, endpoints[rank], certs[rank], members}
        self._fabrics = {}
        self._fabric_of = {}     # learner id -> (fabric_id, rank)
        self._fabric_lock = threading.Lock()
        self.controller_rx_model_bytes = 0   # model payload bytes received through MarkTaskCompleted (diagnostics)
        self.core = _core.Controller(params_pb_to_core(params_pb), self._send_run_tasks, self._send_evaluation_tasks,
                                     self._aggregate_on_device)
        self._shutdown_event = threading.Event()
        self._server = None
        # checkpoint / resume: the durable format is the serialized FederatedModel itself
        self._snapshot_dir = os.environ.get("METISFL_SNAPSHOT_DIR") or None
        self._snapshot_every = max(1, int(os.environ.get("METISFL_SNAPSHOT_EVERY", "1")))
        self._snapshot_iteration = -1
        self._snapshot_thread = None
        if self._snapshot_dir:
            os.makedirs(self._snapshot_dir, exist_ok=True)
            self._resume_from_snapshot()
        self._evaluation_metrics = metis_pb2.EvaluationMetrics()
        # failure detector: poll every learner's health RPC, evict after N consecutive misses
        self._heartbeat_secs = float(os.environ.get("METISFL_HEARTBEAT_SECS", "0") or 0)
        self._heartbeat_misses = max(1, int(os.environ.get("METISFL_HEARTBEAT_MISSES", "3")))
        self._missed = {}
        self._heartbeat_thread = None

    # -- server lifecycle -----------------------------------------------------------------------
    def start(self, max_workers: int = 16):
        self._server_holder = GRPCServerMaxMsgLength(max_workers=max_workers, server_entity=self.params_pb.server_entity)
        self._server = self._server_holder.server
        controller_pb2_grpc.add_ControllerServiceServicer_to_server(self, self._server)
        # the reference's server also answers the standard health-checking and reflection services
        # (controller_servicer.cc:80-84): grpc_health_probe / grpcurl work against this one too
        from metisfl_b200.proto.grpc_std_services import add_standard_services
        self._health = add_standard_services(self._server, ["metisfl.ControllerService"])
        self._server.start()
        from metisfl_b200.controller.metrics import ControllerMetrics
        self._metrics = ControllerMetrics(self)      # METISFL_PROMETHEUS_PORT: optional scrape endpoint
        self._metrics.start()
        if self._snapshot_dir:
            self._snapshot_thread = threading.Thread(target=self._snapshot_loop, daemon=True)
            self._snapshot_thread.start()
        if self._heartbeat_secs > 0:
            self._heartbeat_thread = threading.Thread(target=self._heartbeat_loop, daemon=True)
            self._heartbeat_thread.start()
        MetisLogger.info("Controller listening on %s:%s", self.params_pb.server_entity.hostname,
                         self.params_pb.server_entity.port)

    # -- heartbeat --------------------------------------------------------------------------------------
    def heartbeat_once(self, timeout=2.0):
        """One sweep of the failure detector; returns the ids evicted by it."""
        evicted = []
        for rec in self.core.get_learners():
            lid = rec["id"]
            try:
                stub = self._stub_for(lid, rec["hostname"], rec["port"], rec.get("public_certificate", b""))
                stub.GetServicesHealthStatus(service_common_pb2.GetServicesHealthStatusRequest(), timeout=timeout)
                self._missed[lid] = 0
            except Exception as e:
                self._missed[lid] = self._missed.get(lid, 0) + 1
                MetisLogger.warning("heartbeat miss %d/%d for %s: %s", self._missed[lid], self._heartbeat_misses, lid,
                                    getattr(e, "code", lambda: e)())
                if self._missed[lid] >= self._heartbeat_misses:
                    self._mark_fabric_dissolving(lid)
                    if self.core.evict_learner(lid, "evicted: heartbeat lost"):
                        evicted.append(lid)
                    self._missed.pop(lid, None)
                    with self._channels_lock:
                        ch = self._channels.pop(lid, None)
                    if ch:
                        ch[0].close()
        return evicted

    def _heartbeat_loop(self):
        while not self._shutdown_event.wait(self._heartbeat_secs):
            try:
                self.heartbeat_once()
            except Exception as e:
                MetisLogger.warning("heartbeat sweep failed: %r", e)

    # -- snapshots ------------------------------------------------------------------------------------
    def _resume_from_snapshot(self):
        state_fp = os.path.join(self._snapshot_dir, "state.json")
        if not os.path.exists(state_fp):
            return False
        with open(state_fp) as f:
            state = json.load(f)
        model_fp = os.path.join(self._snapshot_dir, state["community_model_file"])
        with open(model_fp, "rb") as f:
            self.core.replace_community_model(f.read())
        MetisLogger.info("Resumed community model of iteration %s from %s", state.get("global_iteration"), model_fp)
        return True

    def write_snapshot(self, force=False):
        """Persist the current community model + lineage metadata (atomic rename)."""
        if not self._snapshot_dir:
            return None
        blob = self.core.community_model()
        if blob is None:
            return None
        iteration = self.core.global_iteration()
        if not force and (iteration == self._snapshot_iteration or iteration % self._snapshot_every):
            return None
        name = "community_model_{:06d}.pb".format(iteration)
        tmp = os.path.join(self._snapshot_dir, name + ".tmp")
        with open(tmp, "wb") as f:
            f.write(blob)
        os.replace(tmp, os.path.join(self._snapshot_dir, name))
        state = {"global_iteration": iteration, "community_model_file": name,
                 "learners": [l["id"] for l in self.core.get_learners()],
                 "runtime_metadata": self.core.get_runtime_metadata_lineage(8)}
        tmp = os.path.join(self._snapshot_dir, "state.json.tmp")
        with open(tmp, "w") as f:
            json.dump(state, f, default=str)
        os.replace(tmp, os.path.join(self._snapshot_dir, "state.json"))
        self._snapshot_iteration = iteration
        return name

    def _snapshot_loop(self):
        while not self._shutdown_event.wait(0.5):
            try:
                self.write_snapshot()
            except Exception as e:  # a failed snapshot must never take the controller down
                MetisLogger.warning("snapshot failed: %r", e)

    def stop(self, grace=1.0):
        if getattr(self, "_health", None) is not None:
            self._health.enter_graceful_shutdown()
        if getattr(self, "_metrics", None) is not None:
            self._metrics.stop()
        self._shutdown_event.set()
        if self._snapshot_dir:
            try:
                self.write_snapshot(force=True)
            except Exception:
                pass
        self.core.shutdown()
        if self._server is not None:
            self._server.stop(grace).wait()
            self._server = None
        self._rpc_pool.shutdown(wait=False, cancel_futures=True)
        self._agg_pool.shutdown(wait=False, cancel_futures=True)
        with self._channels_lock:
            for ch, _ in self._channels.values():
                ch.close()
            self._channels.clear()

    def shutdown_request_received(self):
        return self._shutdown_event.is_set()

    # -- outbound fan-out (called by the native core) -------------------------------------------
    def _stub_for(self, learner_id, hostname, port, cert):
        with self._channels_lock:
            if learner_id not in self._channels:
                ssl_pb = None
                if cert:
                    ssl_pb = metis_pb2.SSLConfig(enable_ssl=True, ssl_config_stream=metis_pb2.SSLConfigStream(
                        public_certificate_stream=cert))
                entity = metis_pb2.ServerEntity(hostname=hostname, port=port, ssl_config=ssl_pb)
                ch = GRPCChannelMaxMsgLength(entity).channel
                self._channels[learner_id] = (ch, learner_pb2_grpc.LearnerServiceStub(ch))
            return self._channels[learner_id][1]

    def _send_run_tasks(self, items, federated_model_bytes, federated_model_ref_bytes=None):
        # The community model arrives from the core already serialized: the request is assembled around those bytes
        # (proto/envelope.py) instead of parse -> copy into a RunTaskRequest -> serialize per learner.  Learners with
        # the same task template share one request.
        # Learners on the fabric get the control-only twin (device-slot references, a few KB) when there is one.
        built = {}
        for (lid, host, port, cert, iteration, updates, pct_valid) in items:
            ref = bool(federated_model_ref_bytes) and lid in self._fabric_of
            key = (ref, iteration, updates, pct_valid)
            req = built.get(key)
            if req is None:
                rest = learner_pb2.RunTaskRequest(
                    task=metis_pb2.LearningTask(global_iteration=iteration, num_local_updates=updates,
                                                training_dataset_percentage_for_stratified_validation=pct_valid,
                                                metrics=self._evaluation_metrics),
                    hyperparameters=self._hyper)
                req = built[key] = envelope.embed(learner_pb2.RunTaskRequest, "federated_model",
                                                  federated_model_ref_bytes if ref else federated_model_bytes, rest)
            self._rpc_pool.submit(self._do_run_task, lid, host, port, cert, req)

    def _do_run_task(self, lid, host, port, cert, req):
        try:
            self._stub_for(lid, host, port, cert).RunTask(req, timeout=600)
        except Exception as e:  # grpc.RpcError and friends: the core decides what to do (deadline)
            self.core.task_failed(lid, f"RunTask: {e}")

    def _send_evaluation_tasks(self, learners, model_bytes, batch_size, eval_index, iteration, model_ref_bytes=None):
        E = learner_pb2.EvaluateModelRequest
        rest = E(batch_size=batch_size, evaluation_dataset=[E.TRAINING, E.VALIDATION, E.TEST],
                 metrics=self._evaluation_metrics).SerializeToString()
        req = envelope.embed(E, "model", model_bytes, rest)   # one request, shared by every learner
        req_ref = envelope.embed(E, "model", model_ref_bytes, rest) if model_ref_bytes else None
        for (lid, host, port, cert) in learners:
            r = req_ref if (req_ref is not None and lid in self._fabric_of) else req
            self._rpc_pool.submit(self._do_evaluate, lid, host, port, cert, r, eval_index)

    # -- device aggregation (called by the native core with its lock held: RPCs to learners only) ---------------
    def _fabric_stub(self, fabric, rank):
        host, _, port = fabric["endpoints"][rank].rpartition(":")
        return self._stub_for("fabric:{}:{}".format(fabric["id"], rank), host, int(port), fabric["certs"][rank])

    def _aggregate_on_device(self, req):
        """``DeviceAggregateRequest`` (dict) -> dict for the core, or None when no fabric is known (host path)."""
        with self._fabric_lock:
            fabric = next(iter(self._fabrics.values()), None)
        if fabric is None:
            return None
        if fabric.get("dissolving"):
            # a member of the fabric has left (shutdown in progress): the fused kernel needs every rank, so there is
            # nothing to launch any more - the round keeps the previous community model
            return {"error": "SKIP: device fabric is dissolving (a member left the federation)"}
        try:
            world = fabric["world"]
            pb = learner_pb2.AggregateOnDeviceRequest(
                rule=req["rule"], global_iteration=req["global_iteration"], mix=req["mix"], z=req["z"],
                first_block=req["first_block"], last_block=req["last_block"], quantify=req["quantify"],
                executor_rank=fabric["executor"])
            for s in req["sources"]:
                src = pb.sources.add(learner_id=s["learner_id"], device_rank=s["device_rank"], slot=s["slot"],
                                     version=s["version"], weight=s["weight"], has_previous=s["has_previous"],
                                     previous_slot=s["previous_slot"], previous_weight=s["previous_weight"])
                if s["staged_model"] is not None:
                    src.staged_model.CopyFrom(model_pb2.Model.FromString(s["staged_model"]))
            # bytes of the new community model are only pulled off the device when somebody needs them: learners
            # outside the fabric, or the snapshot writer
            off_box = any(rec["id"] not in self._fabric_of for rec in self._learners_cache())
            want_bytes = bool(req["last_block"] and (off_box or self._snapshot_dir))
            asynchronous = req["rule"] in ("FedRec", "FedAsync")
            # one-sided: only the executor GPU folds (its kernel reads the slot and pushes the result to everybody);
            # without peer-mapped memory (CPU tier) the fold is a collective and every rank takes part
            ranks = [fabric["executor"]] if (asynchronous and fabric["peer_mapped"]) else list(range(world))
            if any(not fabric["endpoints"][r] for r in ranks):
                return {"error": "fabric rank without a service endpoint"}

            def call(rank):
                r = learner_pb2.AggregateOnDeviceRequest()
                r.CopyFrom(pb)
                r.return_model_bytes = want_bytes and rank == fabric["executor"]
                if rank != fabric["executor"]:   # the staging GPU alone unpacks off-box models
                    for src in r.sources:
                        if len(src.staged_model.variables):
                            keep = src.staged_model.variables[0]
                            src.ClearField("staged_model")
                            src.staged_model.variables.add(name=keep.name)  # marker: "this source is staged"
                return self._fabric_stub(fabric, rank).AggregateOnDevice(r, timeout=120)

            if len(ranks) == 1:
                resps = [call(ranks[0])]
            else:
                resps = list(self._agg_pool.map(call, ranks))
            bad = [r for r in resps if not r.ack.status]
            if bad:
                return {"error": bad[0].ack.message or "AggregateOnDevice refused"}
            lead = resps[ranks.index(fabric["executor"])]
            out = {"model_ref": lead.community_ref.SerializeToString(),
                   "device_ms": max(r.device_ms for r in resps),   # max over the ranks that ran the kernel
                   "quantifiers": [(q.tensor_non_zeros, q.tensor_zeros, q.tensor_size_bytes) for q in lead.quantifiers],
                   "model_bytes": lead.community_model.SerializeToString() if want_bytes else None, "error": None}
            return out
        except Exception as e:  # grpc.RpcError and friends
            return {"error": "{}: {}".format(type(e).__name__, e)}

    def _learners_cache(self):
        # the core's lock is held by our caller; membership is mirrored here at join / leave time
        with self._fabric_lock:
            return [{"id": lid} for lid in self._joined]

    def _do_evaluate(self, lid, host, port, cert, req, eval_index):
        try:
            resp = self._stub_for(lid, host, port, cert).EvaluateModel(req, timeout=600)
            self.core.evaluation_completed(lid, eval_index, resp.evaluations.SerializeToString())
        except Exception as e:
            self.core.task_failed(lid, f"EvaluateModel: {e}")

    # -- helpers ----------------------------------------------------------------------------------
    @staticmethod
    def _abort(context, err):
        code = _STATUS_TO_GRPC.get(getattr(err, "code", None), grpc.StatusCode.INTERNAL)
        context.abort(code, str(err))

    @staticmethod
    def _metadata_pb(d) -> metis_pb2.FederatedTaskRuntimeMetadata:
        md = metis_pb2.FederatedTaskRuntimeMetadata(
            global_iteration=d["global_iteration"],
            assigned_to_learner_id=d["assigned_to_learner_id"],
            completed_by_learner_id=d["completed_by_learner_id"],
            model_insertion_duration_ms=d["model_insertion_duration_ms"],
            model_selection_duration_ms=d["model_selection_duration_ms"],
            model_aggregation_total_duration_ms=d["model_aggregation_total_duration_ms"],
            model_aggregation_block_size=d["model_aggregation_block_size"],
            model_aggregation_block_memory_kb=d["model_aggregation_block_memory_kb"],
            model_aggregation_block_duration_ms=d["model_aggregation_block_duration_ms"])
        for name in ("started_at", "completed_at", "model_aggregation_started_at", "model_aggregation_completed_at"):
            if d[name]:
                getattr(md, name).CopyFrom(_ts(d[name]))
        for name in ("train_task_submitted_at", "train_task_received_at", "eval_task_submitted_at",
                     "eval_task_received_at"):
            for k, v in d[name].items():
                getattr(md, name)[k].CopyFrom(_ts(v))
        for nz, z, nbytes in d["model_tensor_quantifiers"]:
            md.model_tensor_quantifiers.add(tensor_non_zeros=nz, tensor_zeros=z, tensor_size_bytes=nbytes)
        return md

    # -- RPC handlers -----------------------------------------------------------------------------
    def GetServicesHealthStatus(self, request, context):
        return service_common_pb2.GetServicesHealthStatusResponse(services_status={"controller": True})

    def JoinFederation(self, request, context):
        se, ds = request.server_entity, request.local_dataset_spec
        cert = b""
        if se.ssl_config.enable_ssl and se.ssl_config.WhichOneof("config") == "ssl_config_stream":
            cert = se.ssl_config.ssl_config_stream.public_certificate_stream
        try:
            learner_id, token = self.core.add_learner(
                se.hostname, se.port, se.SerializeToString(), ds.SerializeToString(), ds.num_training_examples,
                ds.num_validation_examples, ds.num_test_examples, cert)
        except _core.ControllerError as e:
            self._abort(context, e)
        with self._fabric_lock:
            self._joined.add(learner_id)
            if request.HasField("fabric") and request.fabric.world > 0:
                fb = request.fabric
                entry = self._fabrics.setdefault(fb.fabric_id, {
                    "id": fb.fabric_id, "world": fb.world, "executor": 0, "peer_mapped": fb.peer_mapped,
                    "endpoints": list(fb.endpoints) or [""] * fb.world,
                    "certs": list(fb.public_certificates) or [b""] * fb.world, "members": {}})
                if not entry["endpoints"][fb.rank]:
                    entry["endpoints"][fb.rank] = "{}:{}".format(se.hostname, se.port)
                    entry["certs"][fb.rank] = cert
                entry["members"][fb.rank] = learner_id
                self._fabric_of[learner_id] = (fb.fabric_id, fb.rank)
        return controller_pb2.JoinFederationResponse(ack=_ack(), learner_id=learner_id, auth_token=token)

    def LeaveFederation(self, request, context):
        self._mark_fabric_dissolving(request.learner_id)  # before the core re-checks the barrier without this learner
        try:
            self.core.remove_learner(request.learner_id, request.auth_token)
        except _core.ControllerError as e:
            self._abort(context, e)
        with self._channels_lock:
            ch = self._channels.pop(request.learner_id, None)
        if ch:
            ch[0].close()
        with self._fabric_lock:
            self._joined.discard(request.learner_id)
        return controller_pb2.LeaveFederationResponse(ack=_ack())

    def _mark_fabric_dissolving(self, learner_id):
        with self._fabric_lock:
            member = self._fabric_of.get(learner_id)
            if member is not None and member[0] in self._fabrics:
                self._fabrics[member[0]]["dissolving"] = True

    # the request is handed over as the bytes that arrived: the model record goes to the core as it is (the core has
    # its own wire parser), only the few hundred bytes around it are parsed here
    RAW_REQUEST_METHODS = ("MarkTaskCompleted",)

    def MarkTaskCompleted(self, request, context):
        R, T = controller_pb2.MarkTaskCompletedRequest, metis_pb2.CompletedLearningTask
        try:
            if isinstance(request, (bytes, bytearray, memoryview)):
                outer = envelope.split(request)
                task = envelope.merged_bytes(outer, R, "task")
                inner = envelope.split(task) if task is not None else {}
                model = envelope.merged_bytes(inner, T, "model")   # a view into the request: copied once, by the core
                model = model if model is not None else b""
                meta = envelope.merged_bytes(inner, T, "execution_metadata")
                md = metis_pb2.TaskExecutionMetadata.FromString(bytes(meta)) if meta is not None \
                    else metis_pb2.TaskExecutionMetadata()
                learner_id, auth_token = (bytes(v).decode("utf-8") if v is not None else "" for v in (
                    envelope.last_bytes(outer, R, "learner_id"), envelope.last_bytes(outer, R, "auth_token")))
            else:   # called with a parsed message (in-process callers, older stubs)
                learner_id, auth_token, md = request.learner_id, request.auth_token, request.task.execution_metadata
                model = request.task.model.SerializeToString()
        except Exception as e:   # ValueError from the splitter, DecodeError from protobuf
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, "malformed MarkTaskCompletedRequest: {}".format(e))
        n_payload = _tensor_payload_bytes(model)
        with self._fabric_lock:   # several RPC threads report at once
            self.controller_rx_model_bytes += n_payload
        try:
            self.core.learner_completed_task(
                learner_id, auth_token, model, md.global_iteration, md.completed_epochs, md.completed_batches,
                md.batch_size, md.processing_ms_per_epoch, md.processing_ms_per_batch, md.SerializeToString())
        except _core.ControllerError as e:
            self._abort(context, e)
        return controller_pb2.MarkTaskCompletedResponse(ack=_ack())

    def ReplaceCommunityModel(self, request, context):
        try:
            self.core.replace_community_model(request.model.SerializeToString())
        except _core.ControllerError as e:
            self._abort(context, e)
        return controller_pb2.ReplaceCommunityModelResponse(ack=_ack())

    def GetParticipatingLearners(self, request, context):
        resp = controller_pb2.GetParticipatingLearnersResponse()
        for rec in self.core.get_learners():
            resp.learner.add(id=rec["id"], server_entity=metis_pb2.ServerEntity.FromString(rec["server_entity"]),
                             dataset_spec=metis_pb2.DatasetSpec.FromString(rec["dataset_spec"]))
        return resp

    def GetRuntimeMetadataLineage(self, request, context):
        resp = controller_pb2.GetRuntimeMetadataLineageResponse()
        timed = False
        for d in self.core.get_runtime_metadata_lineage(request.num_backtracks):
            resp.metadata.append(self._metadata_pb(d))
            timed = timed or bool(d.get("device_timed"))
        # controller-wide counters that have no field of their own in the reference schema
        resp.json_metadata = json.dumps({
            "model_bytes_received": self.controller_rx_model_bytes,
            "aggregation_failures": self.core.aggregation_failures(),
            "aggregation_durations_device_timed": timed,
            "fabric_learners": len(self._fabric_of)})
        return resp

    def GetCommunityModelEvaluationLineage(self, request, context):
        resp = controller_pb2.GetCommunityModelEvaluationLineageResponse()
        for iteration, evals in self.core.get_community_evaluation_lineage(request.num_backtracks):
            ce = resp.community_evaluation.add(global_iteration=iteration)
            for lid, blob in evals.items():
                ce.evaluations[lid].CopyFrom(metis_pb2.ModelEvaluations.FromString(blob))
        return resp

    def GetLocalTaskLineage(self, request, context):
        resp = controller_pb2.GetLocalTaskLineageResponse()
        for lid, metas in self.core.get_local_task_lineage(list(request.learner_ids), request.num_backtracks).items():
            entry = resp.learner_task[lid]
            for blob in metas:
                entry.task_metadata.append(metis_pb2.TaskExecutionMetadata.FromString(blob))
        return resp

    def GetCommunityModelLineage(self, request, context):
        resp = controller_pb2.GetCommunityModelLineageResponse()
        for blob in self.core.get_community_model_lineage(request.num_backtracks):
            resp.federated_models.append(model_pb2.FederatedModel.FromString(blob))
        return resp

    def GetLearnerLocalModelLineage(self, request, context):
        ids = ["{}:{}".format(se.hostname, se.port) for se in request.server_entity]
        by_id = {rec["id"]: rec for rec in self.core.get_learners()}
        resp = controller_pb2.GetLearnerLocalModelLineageResponse()
        for lid, models in self.core.get_learner_local_model_lineage(ids, request.num_backtracks).items():
            entry = resp.learner_local_model.add()
            if lid in by_id:
                entry.server_entity.CopyFrom(metis_pb2.ServerEntity.FromString(by_id[lid]["server_entity"]))
            for blob in models:
                entry.model.append(model_pb2.Model.FromString(blob))
        return resp

    def ShutDown(self, request, context):
        self._shutdown_event.set()
        return service_common_pb2.ShutDownResponse(ack=_ack())
