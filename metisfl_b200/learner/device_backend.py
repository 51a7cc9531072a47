"""Learner backend for federations whose models never leave the GPUs.

Reference flow: ``RunTask`` ships the whole community model to the learner, the learner trains,
``MarkTaskCompleted`` ships the whole local model back (``learner.py:197-206``), the controller averages protobuf
bytes on the CPU (``controller.cc:795-950``) and ``EvaluateModel`` ships the result out again.  Here the learner
process hosts one rank of the on-box fabric (``parallel/engine.py``): it trains on an arena-backed module, the last
optimizer step publishes into a peer-visible slot, and every message that used to carry the model carries a
*reference* to it instead (``TensorSpec.device_slot``) - a few KB of metadata.  The controller keeps all of its
policy (scheduling, selection, scaling factors, lineage, runtime metadata) and, at the barrier, tells the GPU
processes which slots to aggregate with which weights (``LearnerService.AggregateOnDevice``); the fused kernels
(K1 / K3) do the rest over NVLink.  Only the very first community model arrives as bytes (the driver ships it to
the controller, ``driver_session.py:334-342``) and is unpacked on the device (K12).
"""
from __future__ import annotations

import threading
import time
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist

from metisfl_b200.models.model_proto_factory import ModelProtoFactory
from metisfl_b200.models.torch_model_ops import _Batches, load_torch_model, optim_hyper_from_pb
from metisfl_b200.models.trainer import cross_entropy_loss
from metisfl_b200.parallel.arena import ModelArena
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
from metisfl_b200.proto import learner_pb2, model_pb2
from metisfl_b200.utils.metis_logger import MetisLogger
from metisfl_b200.utils.proto_messages_factory import MetisProtoMessages


@dataclass
class FabricSpec:
    """Where this learner sits in the on-box fabric and which aggregation it has to be able to run."""
    rank: int = 0
    world: int = 1
    init_method: str = ""            # torch.distributed rendezvous, e.g. tcp://127.0.0.1:29511 (unused when world == 1)
    fabric_id: str = "fabric0"
    rule: str = "FedAvg"             # FedAvg | FedStride | FedRec | FedAsync | PWA (needs the learner's HE key files)
    staging_slots: int = 2           # per rank, for models of learners outside the fabric
    executor_rank: int = 0
    model_dtype: Optional[str] = None

    @staticmethod
    def parse(text: str) -> "FabricSpec":
        """``rank=0,world=8,init=tcp://127.0.0.1:29511,rule=FedAvg,...`` (the learner CLI's ``--fabric``)."""
        kv = dict(p.split("=", 1) for p in text.split(",") if "=" in p)
        return FabricSpec(rank=int(kv.get("rank", 0)), world=int(kv.get("world", 1)), init_method=kv.get("init", ""),
                          fabric_id=kv.get("id", "fabric0"), rule=kv.get("rule", "FedAvg"),
                          staging_slots=int(kv.get("staging", 2)), executor_rank=int(kv.get("executor", 0)),
                          model_dtype=kv.get("dtype"))

    def to_text(self) -> str:
        t = "rank={},world={},init={},id={},rule={},staging={},executor={}".format(
            self.rank, self.world, self.init_method, self.fabric_id, self.rule, self.staging_slots, self.executor_rank)
        return t + (",dtype=" + self.model_dtype if self.model_dtype else "")


def _spec_of(var_pb):
    return var_pb.ciphertext_tensor.tensor_spec if var_pb.HasField("ciphertext_tensor") else \
        var_pb.plaintext_tensor.tensor_spec


def _has_bytes(model_pb) -> bool:
    return len(model_pb.variables) > 0 and len(_spec_of(model_pb.variables[0]).value) > 0


def _ref_of(model_pb):
    if not len(model_pb.variables):
        return None
    spec = _spec_of(model_pb.variables[0])
    return spec.device_slot if spec.HasField("device_slot") else None


def _as_ciphertext_ref(model_pb):
    """A reference model whose variables are marked as ciphertext: what a learner of a secure federation reports.  The
    controller keeps treating it as an opaque reference (``controller.cc: DeviceSlotOf`` reads the slot of either
    kind)."""
    out = model_pb2.Model()
    for v in model_pb.variables:
        nv = out.variables.add(name=v.name, trainable=v.trainable)
        nv.ciphertext_tensor.tensor_spec.CopyFrom(_spec_of(v))
    return out


def _on_device(fn):
    """The RPC / task threads that call into the backend start with CUDA device 0 current; the kernels launch on the
    CALLING thread's current device and stream, so every entry point first makes the backend's GPU current."""
    import functools

    @functools.wraps(fn)
    def wrapper(self, *args, **kwargs):
        if self.device.type == "cuda":
            with torch.cuda.device(self.device):
                return fn(self, *args, **kwargs)
        return fn(self, *args, **kwargs)
    return wrapper


class DeviceBackend:
    """One per learner process: a ``FederationEngine`` rank plus a private evaluation copy of the module."""

    def __init__(self, model_dir: str, device, fabric: FabricSpec, loss_fn=None, endpoint=None, certificate=b"",
                 he_scheme_config_pb=None):
        self.fabric = fabric
        self._endpoint, self._certificate = endpoint, certificate or b""
        self.device = torch.device(device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu"))
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        if fabric.world > 1 and not dist.is_initialized():
            if self.device.type == "cuda":
                dist.init_process_group("nccl", init_method=fabric.init_method, rank=fabric.rank,
                                        world_size=fabric.world, device_id=self.device)
            else:
                dist.init_process_group("gloo", init_method=fabric.init_method, rank=fabric.rank,
                                        world_size=fabric.world)
        dtype = {"bf16": torch.bfloat16, "fp32": torch.float32, "fp16": torch.float16, None: None}[fabric.model_dtype]
        if dtype is None:
            dtype = torch.bfloat16 if self.device.type == "cuda" else torch.float32
        module = load_torch_model(model_dir)
        if loss_fn is None:
            from metisfl_b200.models.torch_model_ops import _overrides
            loss_fn = module.loss_fn if _overrides(module, "loss_fn") else cross_entropy_loss
        self.loss_fn = loss_fn
        # the controller owns the scaling factors, so nothing is pre-scaled at publish time (no in-switch reduction);
        # the working copy is separate from the community buffer so that evaluation of version v and training from
        # version v can overlap
        # PWA: the engine keeps its synchronous buffers; encryption / reduction / decryption are driven from here
        cfg = EngineConfig(learners_per_gpu=1, model_dtype=dtype, rule="FedAvg" if fabric.rule == "PWA" else fabric.rule,
                           nvls=False,
                           staging_slots=fabric.staging_slots, executor_rank=fabric.executor_rank,
                           direct_install=False)
        self.engine = FederationEngine(lambda: module, cfg, loss_fn=loss_fn, device=self.device)
        eval_module = load_torch_model(model_dir)
        if self.device.type == "cuda":
            eval_module = eval_module.to(memory_format=torch.channels_last)
            if dtype == torch.bfloat16:
                from metisfl_b200.ops import gemm as _gemm
                _gemm.use_b200_linear(eval_module)
        self.eval_arena = ModelArena(eval_module, self.device, dtype)
        assert self.eval_arena.layout.total_bytes == self.engine.layout.total_bytes
        self.secure = None
        if fabric.rule == "PWA":
            he = he_scheme_config_pb
            if he is None or not he.enabled or he.WhichOneof("config") != "ckks_scheme_config":
                raise RuntimeError("rule PWA on the device fabric needs the learner's CKKS configuration (key files)")
            from metisfl_b200.parallel.secure import FabricSecure
            self.secure = FabricSecure(self.engine, he.crypto_context_file, he.public_key_file, he.private_key_file,
                                       he.ckks_scheme_config.batch_size, he.ckks_scheme_config.scaling_factor_bits)
        # every rank learns every rank's service endpoint, so the controller can command the whole fabric as soon as
        # ONE of its learners has joined (the fused kernel needs all ranks to launch)
        self._peers = self.engine.fabric.all_gather_object((self._endpoint or "", bytes(self._certificate)))
        self._comm_lock = threading.Lock()   # community buffer: aggregation writes, evaluation / install read
        self._installed_bytes_iteration = None
        self._known_version = 0
        self._ds_cache = {}       # pinned host copies + gather buffers of the array datasets (models/torch_model_ops._Batches)
        self._eval_cache = {}     # (id(dataset), batch size) -> device-resident evaluation batches
        self._eval_cache_bytes = 0
        self._eval_graphs = {}    # batch signature -> captured forward + metric accumulation
        self._eval_acc = torch.zeros(2, dtype=torch.float32, device=self.device)   # [sum of loss * n, correct]

    # -- membership -----------------------------------------------------------------------------------------------
    def membership_pb(self):
        from metisfl_b200.proto import controller_pb2
        return controller_pb2.FabricMembership(fabric_id=self.fabric.fabric_id, rank=self.fabric.rank,
                                               world=self.fabric.world, staging_slots=self.fabric.staging_slots,
                                               peer_mapped=bool(self.engine.fabric.peer_mapped),
                                               endpoints=[p[0] for p in self._peers],
                                               public_certificates=[p[1] for p in self._peers])

    # -- community model --------------------------------------------------------------------------------------------
    def _adopt(self, model_pb, iteration=None):
        """Make the engine's community buffer hold the model the message talks about."""
        eng = self.engine
        if _has_bytes(model_pb):
            key = (iteration, len(model_pb.variables))
            with self._comm_lock:
                live = eng.async_rule and eng.version > 0  # folds are already landing in the buffers: keep them
                if (self._installed_bytes_iteration != key or iteration is None) and not live:
                    self._install_bytes(model_pb)
                    self._installed_bytes_iteration = key
            return None
        ref = _ref_of(model_pb)
        if ref is not None and eng.async_rule:
            self._known_version = max(self._known_version, int(ref.version))
            if not eng.is_executor or eng.version < self._known_version:
                eng.version = max(eng.version, self._known_version)
        return ref

    def _install_bytes(self, model_pb):
        """A model that arrived as bytes -> the community buffer.  Encrypted models (the driver's initial ship of a
        secure federation, ``driver_session.py:334-342`` with ``he_scheme``; an encrypted snapshot) are decrypted
        with this learner's private key first, like ``model_ops.py:24-57``."""
        eng = self.engine
        if not model_pb.variables[0].HasField("ciphertext_tensor"):
            return eng.install_community_model(model_pb)
        if self.secure is None:
            raise RuntimeError("an encrypted model reached a learner that holds no keys")
        from metisfl_b200.parallel.secure import FabricSecure
        if len(model_pb.variables) == 1 and model_pb.variables[0].name == FabricSecure.SNAPSHOT_VARIABLE:
            return self.secure.install_blob(model_pb.variables[0].ciphertext_tensor.tensor_spec.value)
        from metisfl_b200.utils.proto_messages_factory import ModelProtoMessages
        codec = ModelProtoMessages.TensorSpecProto
        names, trainables, values = [], [], []
        for var in model_pb.variables:
            spec = var.ciphertext_tensor.tensor_spec
            flat = self.secure.decrypt_blob(spec.value, spec.length)
            names.append(var.name)
            trainables.append(var.trainable)
            values.append(codec.proto_tensor_spec_with_list_values_to_numpy_array_like(spec, flat))
        eng.install_community_model(ModelProtoMessages.construct_model_pb_from_np(values, names, trainables))

    # -- training ---------------------------------------------------------------------------------------------------
    @_on_device
    def train(self, train_dataset, learning_task_pb, hyperparameters_pb, model_pb, cancel_event=None):
        eng = self.engine
        ref = self._adopt(model_pb, learning_task_pb.global_iteration)
        eng.trainer.set_hyper(optim_hyper_from_pb(hyperparameters_pb.optimizer))
        batch_size = hyperparameters_pb.batch_size or 32
        total_steps = max(1, int(learning_task_pb.num_local_updates))
        src = _Batches(train_dataset, batch_size, shuffle=True, pin=self.device.type == "cuda",
                       seed=learning_task_pb.global_iteration, cache=self._ds_cache, **self._batch_format())
        spe = max(1, src.steps_per_epoch)
        batches = []
        while len(batches) < total_steps:
            for x, y in src:
                if x.shape[0] != src.batch_size and len(batches) > 0:
                    continue  # ragged tail batches would re-capture the CUDA graph: skipped, like drop_last
                batches.append((x, y))
                if len(batches) >= total_steps:
                    break
        cuda = self.device.type == "cuda"
        if cuda:
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
        wall0 = time.perf_counter()
        loss = eng.train_learner(0, batches, publish_scale=1.0,
                                 min_version=int(ref.version) if (ref is not None and eng.async_rule) else None,
                                 cancel=cancel_event)
        if loss is None and (cancel_event is not None and cancel_event.is_set()):
            if cuda:   # copies out of the pinned gather buffer may still be in flight; the next task rewrites it
                torch.cuda.current_stream(self.device).synchronize()
            return None
        if cuda:
            t1.record()
            t1.synchronize()
            total_ms = t0.elapsed_time(t1)   # device time of the whole task (CUDA events on the training stream)
        else:
            total_ms = (time.perf_counter() - wall0) * 1e3
        steps = len(batches)
        ms_per_batch = total_ms / max(1, steps)
        stats = {"loss": [float(loss)]} if loss is not None else {}
        slot = eng.publish_slot(0)
        local_ref = eng.reference_model_pb(slot, version=eng.task_counter[0])
        if self.secure is not None:   # the published model leaves this learner's hands only as a ciphertext
            self.secure.encrypt_slot(slot)
            if cuda:   # the completion message says "the ciphertext is there": peers will read it
                torch.cuda.current_stream(self.device).synchronize()
            local_ref = _as_ciphertext_ref(local_ref)
        meta = ModelProtoFactory.CompletedLearningTaskProtoMessage(
            weights_values=[], weights_trainable=[], weights_names=[], train_stats=stats,
            completed_epochs=steps / spe, global_iteration=learning_task_pb.global_iteration,
            completes_batches=steps, batch_size=batch_size, processing_ms_per_epoch=ms_per_batch * spe,
            processing_ms_per_batch=ms_per_batch).construct_task_execution_metadata_pb()
        return MetisProtoMessages.construct_completed_learning_task_pb(
            model_pb=local_ref, task_execution_metadata_pb=meta,
            aux_metadata='{{"device_timed": {}, "fabric_rank": {}}}'.format("true" if cuda else "false", eng.rank))

    # -- evaluation -------------------------------------------------------------------------------------------------
    @_on_device
    @torch.no_grad()
    def evaluate(self, model_pb, datasets, batch_size, metrics):
        """``datasets`` = {code: dataset or None}; returns {code: {metric: value}}."""
        eng = self.engine
        nb = eng.layout.total_bytes
        self._adopt(model_pb)
        with self._comm_lock:
            if self.device.type == "cuda" and eng.async_rule and not eng.is_executor:
                # a fold may be landing: take a consistent snapshot through the version counter
                from metisfl_b200.ops import aggregation as agg_ops
                comm = [eng.community.ptrs[eng.rank] + b * nb for b in range(eng.n_comm)]
                agg_ops.snapshot_install(comm, eng.ctrl.ptrs[eng.rank], eng._status, self.eval_arena.state, nb,
                                         None, None)
            else:
                self.eval_arena.state[:nb].copy_(eng.community_bytes()[:nb], non_blocking=True)
            if self.device.type == "cuda":
                torch.cuda.current_stream(self.device).synchronize()
        module = self.eval_arena.module
        module.eval()
        out = {}
        for code, ds in datasets.items():
            if ds is None or not ds.get_size():
                out[code] = {}
                continue
            self._eval_acc.zero_()
            seen, classification = 0, False
            for x, y in self._eval_batches(ds, batch_size or 100):
                classification = self._eval_step(module, x, y) or classification
                seen += x.shape[0]
            acc = self._eval_acc.tolist()   # the one host read-back per split
            res = {"loss": acc[0] / max(1, seen)}
            if classification:
                res["accuracy"] = acc[1] / max(1, seen)
            if metrics:
                res = {k: v for k, v in res.items() if k in metrics or k == "loss"} or res
            out[code] = res
        return out

    def _batch_format(self):
        """Cached host copies of a dataset are kept in what the device consumes: the model dtype, NHWC on CUDA."""
        if self.device.type != "cuda":
            return {}
        return {"x_dtype": self.eval_arena.model_dtype, "channels_last": True}

    EVAL_CACHE_LIMIT_BYTES = 8 << 30   # evaluation splits up to this total stay resident in HBM between rounds

    def _eval_batches(self, ds, batch_size):
        """Evaluation batches on the device, in the model's dtype / memory format.  The splits of a learner do not
        change between rounds (``learner.py`` loads them once), so they are uploaded once and kept."""
        key = (id(ds), batch_size)
        hit = self._eval_cache.get(key)
        if hit is not None:
            return hit[1]
        cuda = self.device.type == "cuda"
        batches, nbytes = [], 0
        for x, y in _Batches(ds, batch_size, shuffle=False, pin=cuda, cache=self._ds_cache, **self._batch_format()):
            x = x.to(self.device, non_blocking=True)
            y = y.to(self.device, non_blocking=True)
            if x.is_floating_point():
                x = x.to(self.eval_arena.model_dtype)
                if x.dim() == 4 and cuda:
                    x = x.contiguous(memory_format=torch.channels_last)
            batches.append((x, y))
            nbytes += x.numel() * x.element_size() + y.numel() * y.element_size()
        if cuda and self._eval_cache_bytes + nbytes <= self.EVAL_CACHE_LIMIT_BYTES:
            self._eval_cache[key] = (ds, batches)   # holding ``ds`` keeps id(ds) from being reused
            self._eval_cache_bytes += nbytes
        return batches

    def _eval_body(self, module, x, y) -> bool:
        logits = module(x)
        self._eval_acc[0] += self.loss_fn(logits, y).float() * x.shape[0]
        if logits.dim() == 2 and not y.is_floating_point():
            self._eval_acc[1] += (logits.argmax(dim=-1) == y).sum()
            return True
        return False

    def _eval_step(self, module, x, y) -> bool:
        """One evaluation batch: forward + loss / accuracy accumulation on the device.  On CUDA the step is captured
        once per batch signature and replayed (a ResNet-18 forward is ~60 launches; eagerly the host is the bottleneck
        and it competes with the training thread for the interpreter)."""
        if self.device.type != "cuda":
            return self._eval_body(module, x, y)
        key = (tuple(x.shape), x.dtype, tuple(y.shape), y.dtype)
        ent = self._eval_graphs.get(key)
        if ent is None:
            ent = {"graph": None}
            try:
                sx, sy = torch.empty_like(x), torch.empty_like(y)
                sx.copy_(x)
                sy.copy_(y)
                saved = self._eval_acc.clone()   # warm-up passes add to the accumulator: put it back afterwards
                cur = torch.cuda.current_stream(self.device)
                side = torch.cuda.Stream(device=self.device)
                side.wait_stream(cur)
                with torch.cuda.stream(side):
                    for _ in range(2):   # lazy initialisations, cuDNN algorithm selection
                        cls = self._eval_body(module, sx, sy)
                cur.wait_stream(side)
                g = torch.cuda.CUDAGraph()
                # thread_local: the training thread keeps launching and allocating while this thread captures
                with torch.cuda.graph(g, capture_error_mode="thread_local"):
                    self._eval_body(module, sx, sy)
                self._eval_acc.copy_(saved)
                ent = {"graph": g, "x": sx, "y": sy, "cls": cls}
            except Exception as e:   # capture is an optimisation: fall back to eager evaluation for this signature
                from metisfl_b200.utils.metis_logger import MetisLogger
                MetisLogger.warning("evaluation graph capture failed (%r); evaluating eagerly", e)
                torch.cuda.synchronize(self.device)
            self._eval_graphs[key] = ent
        if ent["graph"] is None:
            return self._eval_body(module, x, y)
        ent["x"].copy_(x, non_blocking=True)
        ent["y"].copy_(y, non_blocking=True)
        ent["graph"].replay()
        return ent["cls"]

    # -- aggregation (commanded by the controller) ------------------------------------------------------------------
    @_on_device
    def aggregate(self, req) -> "learner_pb2.AggregateOnDeviceResponse":
        eng = self.engine
        cuda = self.device.type == "cuda"
        resp = learner_pb2.AggregateOnDeviceResponse()
        with self._comm_lock:
            # device time of the launch(es), on the stream they run on
            tstream = None
            if cuda:
                tstream = eng.fold_stream if eng.async_rule else torch.cuda.current_stream(self.device)
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record(tstream)
            wall0 = time.perf_counter()
            if self.secure is not None:
                # encrypted FedAvg: the controller's scaling factors, by fabric rank; the reduction runs on ciphertexts
                # (K8 over NVLink) and every learner decrypts its own copy of the encrypted community model
                w = [0.0] * eng.world
                for src in req.sources:
                    if len(src.staged_model.variables) or not 0 <= int(src.device_rank) < eng.world:
                        raise ValueError("secure aggregation on the fabric takes device references of fabric ranks only")
                    w[int(src.device_rank)] = float(src.weight)
                self.secure.reduce(w)
                self.secure.decrypt_install()
                version = eng.global_iteration
            elif not eng.async_rule:
                sources, weights, staged = [], [], 0
                for src in req.sources:
                    if len(src.staged_model.variables):
                        if eng.rank == req.executor_rank:
                            slot = eng.stage_model(src.staged_model, staged)
                        else:
                            slot = eng.k * eng.ring + staged
                        sources.append((req.executor_rank, slot))
                        staged += 1
                    else:
                        sources.append((src.device_rank, src.slot))
                    weights.append(src.weight)
                eng.aggregate_sync(sources=sources, weights=weights, accumulate=not req.first_block)
                version = eng.global_iteration
            else:
                assert len(req.sources) == 1, "asynchronous rules fold one learner at a time"
                src = req.sources[0]
                learner = int(src.device_rank) * eng.k
                eng.version = max(eng.version, self._known_version)
                if req.rule == "FedAsync":
                    eng.fold_async(learner, mix=req.mix, task_no=int(src.version), slot=int(src.slot))
                else:
                    eng.fold_async(learner, weight=src.weight, z=req.z, task_no=int(src.version), slot=int(src.slot),
                                   previous_weight=src.previous_weight if src.has_previous else None,
                                   previous_slot=int(src.previous_slot) if src.has_previous else None)
                version = eng.version
                self._known_version = version
            if cuda:
                t1.record(tstream)
                t1.synchronize()
                resp.device_ms = t0.elapsed_time(t1)
            else:
                resp.device_ms = (time.perf_counter() - wall0) * 1e3
            if req.quantify and req.last_block and self.secure is None:   # plaintext statistics only
                eng.quantify_community()
                rep = eng.quantifier_report()
                for info in eng.layout.variables:
                    nz, z, nbytes = rep[info.name]
                    resp.quantifiers.add(tensor_non_zeros=int(nz), tensor_zeros=int(z), tensor_size_bytes=int(nbytes))
            resp.version = version
            resp.community_ref.CopyFrom(eng.reference_model_pb(eng.COMMUNITY_SLOT, version))
            if self.secure is not None:
                resp.community_ref.CopyFrom(_as_ciphertext_ref(resp.community_ref))
                if req.return_model_bytes:   # what the controller may keep (snapshots) stays encrypted
                    from metisfl_b200.parallel.secure import FabricSecure
                    blob = self.secure.community_blob()
                    var = resp.community_model.variables.add(name=FabricSecure.SNAPSHOT_VARIABLE, trainable=True)
                    spec = var.ciphertext_tensor.tensor_spec
                    spec.length = self.secure.n
                    spec.dimensions.append(self.secure.n)
                    spec.type.type = model_pb2.DType.FLOAT64
                    spec.type.byte_order = model_pb2.DType.LITTLE_ENDIAN_ORDER
                    spec.value = blob
            elif req.return_model_bytes:
                resp.community_model.CopyFrom(eng.community_model_pb())
        from metisfl_b200.learner.learner_servicer import _ack
        resp.ack.CopyFrom(_ack(True))
        return resp

    @_on_device
    def close(self):
        try:
            self.engine.close()
        except Exception as e:  # peers may already be gone at shutdown
            MetisLogger.warning("fabric close: %r", e)
