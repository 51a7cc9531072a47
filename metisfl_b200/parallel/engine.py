"""On-box federation engine: the device-resident fast path of a federation round.

Reference round: RunTask fan-out of the whole model over gRPC -> local training ->
MarkTaskCompleted upload -> CPU weighted average -> EvaluateModel fan-out
(``controller.cc:428-518,696-759,571-647``; ``learner.py:376-396``).  On one NVSwitch box the same
round is: every GPU process trains its learner(s) on an arena-backed module, the last optimizer
step publishes into a peer-visible slot (K2), and ONE fused kernel per GPU (K1) reads all slots over
NVLink, applies the per-learner scaling factors and writes the new community model - and its fp32
master copy - into every GPU.  No tensor crosses the host.

Protocols / rules supported here: synchronous FedAvg and FedStride (streaming blocks), semi-synchronous
task sizing, asynchronous FedRec (recency swap) and FedAsync (staleness-weighted mixing, extension).
Scaling factors follow ``batches_scaler.cc`` / ``participants_scaler.cc`` /
``train_dataset_size_scaler.cc`` semantics.
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.nn as nn

from metisfl_b200.models.trainer import LocalTrainer, cross_entropy_loss
from metisfl_b200.ops import aggregation as agg_ops
from metisfl_b200.ops.optim import OptimHyper
from metisfl_b200.parallel.arena import ArenaLayout, ModelArena
from metisfl_b200.parallel.fabric import Fabric
from metisfl_b200.utils.tracing import DeviceTimer, nvtx_range


def staleness_factor(kind: str, staleness: float, a: float = 0.5, b: float = 4.0) -> float:
    """FedAsync staleness functions s(t - tau): constant, polynomial (1+d)^-a, hinge."""
    d = max(0.0, float(staleness))
    if kind in ("constant", "CONSTANT", 0):
        return 1.0
    if kind in ("polynomial", "POLYNOMIAL", 1):
        return (1.0 + d) ** (-a)
    if kind in ("hinge", "HINGE", 2):
        return 1.0 if d <= b else 1.0 / (a * (d - b) + 1.0)
    raise ValueError(f"unknown staleness function {kind}")


def scaling_factors(kind: str, examples: Sequence[float], batches: Sequence[float], federation_size: int):
    """Normalised contribution weights of the participating learners (SURVEY §9.2 'Scalers')."""
    n = len(examples)
    if federation_size == 1:
        return [1.0] * n
    if kind in ("NumParticipants", "NUM_PARTICIPANTS"):
        return [1.0 / n] * n
    raw = list(examples) if kind in ("NumTrainingExamples", "NUM_TRAINING_EXAMPLES") else list(batches)
    if n == 1:
        return [float(raw[0])]
    tot = float(sum(raw))
    return [float(r) / tot for r in raw]


@dataclass
class EngineConfig:
    learners_per_gpu: int = 1
    model_dtype: torch.dtype = torch.bfloat16
    optimizer: OptimHyper = field(default_factory=lambda: OptimHyper("MomentumSGD", lr=0.05, momentum=0.9))
    rule: str = "FedAvg"                 # FedAvg | FedStride | FedRec | FedAsync | PWA (CKKS secure FedAvg, needs crypto_dir)
    crypto_dir: Optional[str] = None     # PWA: directory with cryptocontext.txt / key-public.txt / key-private.txt
    he_batch_size: int = 4096
    he_scaling_factor_bits: int = 52
    scaling: str = "NumTrainingExamples"
    stride_length: int = 0
    num_train_examples: int = 6400       # per learner (can be overridden per learner)
    persist_optimizer_state: bool = False
    quantify: bool = False
    fedasync_alpha: float = 0.6
    staleness_fn: str = "polynomial"
    staleness_a: float = 0.5
    staleness_b: float = 4.0
    channels_last: bool = True
    use_cuda_graph: Optional[bool] = None
    fused_blocks: int = 296
    nvls: Optional[bool] = None          # NVSwitch in-switch reduction + multicast stores; None: when world >= 4 (measured: 1.83 vs 2.10 ms
                                         # at 8 GPUs for 355 M bf16 parameters, but 1.83 vs 1.16 ms at 2 GPUs)
    overlap_wgrad: bool = True           # convolution weight gradients on a side stream (ops/conv.py), CUDA only
    own_conv: bool = False               # every convolution (forward, data and weight gradient) on the hand-written
                                         # kernels (K11 / K11s / K15) instead of cuDNN; bf16 on CUDA only
    use_b200_linear: bool = True         # nn.Linear forward/backward on the tcgen05 GEMM (K10) when bf16 on CUDA
    seed: int = 1234
    ring: Optional[int] = None           # publish slots per learner (the device lineage): FedRec keeps the previous
                                         # contribution, so 2; every other rule 1
    staging_slots: int = 0               # extra slots per rank for models of off-box learners (unpacked by K12)
    executor_rank: int = 0               # asynchronous rules: the GPU that owns the fp32 rolling state and folds
    trace: bool = False                  # CUDA-event phase timers (train / aggregate / fold) + NVTX ranges; read them
                                         # with ``engine.phase_times()``
    fold_blocks: int = 64                # CTAs of one asynchronous fold (it shares the executor's GPU with training)
    standalone: bool = False             # ignore an initialised torch.distributed job: this engine is its own world
    direct_install: Optional[bool] = None  # train in place in the community buffer (one learner per GPU, synchronous
                                           # rules); False keeps a separate working copy (service path: evaluation of
                                           # version v overlaps training from version v)


class FederationEngine:
    """One instance per GPU process (``torch.distributed`` initialised by the launcher, or world 1)."""

    def __init__(self, model_factory: Callable[[], nn.Module], cfg: EngineConfig = None,
                 loss_fn: Callable = cross_entropy_loss, device=None, group=None,
                 examples_per_learner: Optional[Sequence[int]] = None):
        self.cfg = cfg = cfg or EngineConfig()
        self.fabric = Fabric(device, group, standalone=cfg.standalone)
        self.device = self.fabric.device
        self.rank, self.world = self.fabric.rank, self.fabric.world
        self.k = cfg.learners_per_gpu
        self.num_learners = self.world * self.k
        torch.manual_seed(cfg.seed)  # identical initial community model on every rank
        module = model_factory()
        if cfg.channels_last and self.device.type == "cuda":
            module = module.to(memory_format=torch.channels_last)
        if cfg.own_conv and self.device.type == "cuda" and cfg.model_dtype == torch.bfloat16:
            from metisfl_b200.ops import conv as _conv
            _conv.use_b200_conv(module)
        elif cfg.overlap_wgrad and self.device.type == "cuda":
            from metisfl_b200.ops import conv as _conv
            _conv.use_overlapped_wgrad(module)
        if cfg.use_b200_linear and self.device.type == "cuda" and cfg.model_dtype == torch.bfloat16:
            from metisfl_b200.ops import gemm as _gemm
            _gemm.use_b200_linear(module)
        self.layout = ArenaLayout.from_module(module, cfg.model_dtype)
        nb = self.layout.total_bytes
        # symmetric buffers: community model + k publish slots per rank
        want_mc = (cfg.nvls is True or (cfg.nvls is None and self.world >= 4)) and self.world > 1 and self.k <= 8
        sync_rule = cfg.rule in ("FedAvg", "FedStride", "PWA")
        self.async_rule = not sync_rule
        self.ring = int(cfg.ring or (2 if cfg.rule == "FedRec" else 1))
        self.n_slots = self.k * self.ring + int(cfg.staging_slots)
        # asynchronous rules rotate the community model through three buffers (version v lives in buffer v % 3) so
        # that a learner copying version v is not overwritten by the folds producing v + 1 and v + 2
        self.n_comm = 3 if self.async_rule else 1
        if self.async_rule:
            want_mc = (cfg.nvls is True or (cfg.nvls is None and self.world >= 4)) and self.world > 1
        self.community = self.fabric.alloc(nb * self.n_comm, multicast=want_mc)
        self.slots = self.fabric.alloc(nb * self.n_slots, multicast=want_mc and sync_rule)
        self.pads = self.fabric.signal_pads(cfg.fused_blocks)
        # control words (uint64): [0] community version, [8 + g] "task n of learner g is published"
        self.ctrl = self.fabric.alloc(8 * (8 + self.world * self.k))
        self.task_counter = [0] * self.k                       # tasks started per local learner
        self.version = 0                                       # community updates folded so far
        # NVLS aggregation: the switch adds, so learners publish w_l * theta_l; that needs weights known before the
        # task ends (example counts / participant counts) and every learner in every round
        self.nvls = bool(want_mc and self.community.mc_ptr and self.slots.mc_ptr and sync_rule
                         and cfg.scaling.lower() in ("numtrainingexamples", "numparticipants"))
        if cfg.nvls and not self.nvls and self.world > 1:
            raise RuntimeError("EngineConfig.nvls=True but multicast memory / a static scaling rule is unavailable")
        self._slot_scale = [1.0] * self.k
        self.direct_install = sync_rule and self.k == 1 and cfg.direct_install is not False
        storage = self.community.local if self.direct_install else None
        self.arena = ModelArena(module, self.device, cfg.model_dtype, storage=storage)
        if not self.direct_install:
            self.community.local[:nb].copy_(self.arena.state[:nb])
        self.trainer = LocalTrainer(self.arena, cfg.optimizer, loss_fn, cfg.use_cuda_graph)
        self.slot_views = [self.layout.views(self.slots.local[j * nb:(j + 1) * nb]) for j in range(self.n_slots)]
        self._comm_views = [self.layout.views(self.community.local[b * nb:(b + 1) * nb]) for b in range(self.n_comm)]
        if self.n_comm > 1:
            for b in range(1, self.n_comm):
                self.community.local[b * nb:(b + 1) * nb].copy_(self.community.local[:nb])
        ex = list(examples_per_learner) if examples_per_learner is not None else \
            [cfg.num_train_examples] * self.num_learners
        assert len(ex) == self.num_learners
        self.examples = ex
        self.completed_batches = [0] * self.num_learners
        self.global_iteration = 0
        self.quantifier = None
        self.gpu_launches = 0
        # async state (FedRec / FedAsync)
        self.learner_version = [0] * self.num_learners       # community iteration a learner started from
        self.contributed = [False] * self.num_learners
        self._remote_tasks = {}
        self._copy_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._pinned_loss = None
        self.timer = DeviceTimer(self.device) if cfg.trace else None
        self._init_async_state()
        self.secure = None
        if cfg.rule == "PWA":
            assert self.k == 1 and cfg.crypto_dir, "PWA: one learner per GPU and EngineConfig.crypto_dir (key files)"
            assert self.device.type == "cuda", "the PWA engine rule runs the CKKS kernels: CUDA only (host path: fhe.CKKS)"
            from metisfl_b200.parallel.secure import SecureAggregator
            self.secure = SecureAggregator(self, cfg.crypto_dir, cfg.he_batch_size, cfg.he_scaling_factor_bits)
        self.fabric.host_barrier()

    # ------------------------------------------------------------------------------------------
    @property
    def community_views(self):
        """Typed segment views of the CURRENT community model (the buffer of ``self.version``)."""
        return self._comm_views[self.version % self.n_comm]

    def community_bytes(self) -> torch.Tensor:
        nb, b = self.layout.total_bytes, self.version % self.n_comm
        return self.community.local[b * nb:(b + 1) * nb]

    def publish_slot(self, j: int, task_no: Optional[int] = None) -> int:
        """Slot index (inside this rank's slot buffer) that task ``task_no`` of local learner ``j`` publishes into."""
        n = self.task_counter[j] if task_no is None else task_no
        return j * self.ring + (max(n, 1) - 1) % self.ring

    # ------------------------------------------------------------------------------------------
    def learner_id(self, local_index: int) -> int:
        return self.rank * self.k + local_index

    def _slot_ptr(self, rank: int, slot: int) -> int:
        return self.slots.ptrs[rank] + slot * self.layout.total_bytes

    def _publish_buffers(self, slot: int, scale: float = 1.0):
        """Non-trainable segments (BatchNorm statistics ...) are copied into the slot after the task
        (pre-scaled like the trainable ones when the aggregation runs in the switch)."""
        for i, s in enumerate(self.layout.segments):
            if not s.trainable:
                if scale != 1.0 and not s.exact:
                    torch.mul(self.arena.seg[i], scale, out=self.slot_views[slot][i])
                else:
                    self.slot_views[slot][i].copy_(self.arena.seg[i], non_blocking=True)
                self.gpu_launches += 1

    def _publish_all(self, slot: int, scale: float = 1.0):
        nb = self.layout.total_bytes
        if scale == 1.0:
            self.slots.local[slot * nb:(slot + 1) * nb].copy_(self.arena.state[:nb], non_blocking=True)
            self.gpu_launches += 1
            return
        for i, s in enumerate(self.layout.segments):
            if s.exact:
                self.slot_views[slot][i].copy_(self.arena.seg[i], non_blocking=True)
            else:
                torch.mul(self.arena.seg[i], scale, out=self.slot_views[slot][i])
            self.gpu_launches += 1

    def _static_weight(self, learner: int) -> float:
        """Aggregation weight of a learner when it depends only on static facts (NVLS pre-scaling)."""
        ids = list(range(self.num_learners))
        w = scaling_factors(self.cfg.scaling, [self.examples[i] for i in ids], [1] * len(ids), self.num_learners)
        return float(w[learner])

    # ------------------------------------------------------------------------------------------
    def _phase(self, name: str):
        """NVTX range + (with ``EngineConfig.trace``) a CUDA-event timer around one phase of the round."""
        return self.timer.phase(name) if self.timer is not None else nvtx_range(name)

    def phase_times(self, reset: bool = True) -> Dict[str, float]:
        """Device milliseconds per phase since the last call (``EngineConfig.trace``): the numbers that go into
        ``FederatedTaskRuntimeMetadata`` on the service path, taken with CUDA events instead of host clocks."""
        return self.timer.report(reset) if self.timer is not None else {}

    def train_learner(self, j: int, batches: Sequence, publish_scale: Optional[float] = None,
                      min_version: Optional[int] = None, cancel=None):
        with self._phase("train_learner"):
            return self._train_learner(j, batches, publish_scale, min_version, cancel)

    def _train_learner(self, j: int, batches: Sequence, publish_scale: Optional[float] = None,
                       min_version: Optional[int] = None, cancel=None):
        """Run one learning task for local learner ``j``; returns the device loss of the last step.
        ``publish_scale`` defaults to the learner's aggregation weight in NVLS mode and to 1 otherwise.

        ``batches`` is a sequence of ``(x, y)``; pinned host tensors are copied with a prefetching
        side stream into the CUDA graph's static inputs.  Asynchronous rules first install the newest
        community model (at least ``min_version``) and, when the task is done, release "task n of this learner is
        published" to the executor.  ``cancel`` (a ``threading.Event``) pre-empts between optimizer steps; a
        pre-empted task publishes nothing and returns ``None``.
        """
        tr, arena = self.trainer, self.arena
        nb = self.layout.total_bytes
        if publish_scale is None:
            publish_scale = self._static_weight(self.learner_id(j)) if self.nvls else 1.0
        self._slot_scale[j] = float(publish_scale)
        self.task_counter[j] += 1
        slot = self.publish_slot(j)
        if self.async_rule:
            self._install_community(j, min_version)
        elif not self.direct_install:
            arena.load_state_bytes(self.community.local[:nb])
            self.gpu_launches += 1 + sum(1 for s, m, v in zip(self.layout.segments, arena.master, arena.seg)
                                         if s.trainable and m is not v)
        if not self.cfg.persist_optimizer_state:
            tr.reset_optimizer_state()
        if tr.hyper.kind == 2:
            tr.snapshot_anchor()
        loss = None
        n = len(batches)
        if n == 0:
            self._publish_all(slot, publish_scale)
            self._signal_ready(j)
            return None
        on_cuda = self.device.type == "cuda"
        for s, (x, y) in enumerate(batches):
            if cancel is not None and cancel.is_set():
                self.task_counter[j] -= 1
                return None
            last = s == n - 1
            if on_cuda and not x.is_cuda:
                sx, sy = tr.input_buffers(x, y)
                if sx is not None:
                    sx.copy_(x, non_blocking=True)
                    sy.copy_(y, non_blocking=True)
                    x, y = sx, sy
                else:
                    x, y = x.to(self.device, non_blocking=True), y.to(self.device, non_blocking=True)
                if self.cfg.channels_last and x.dim() == 4 and sx is None:
                    x = x.contiguous(memory_format=torch.channels_last)
            loss = tr.forward_backward(x, y)
            self.gpu_launches += 1
            self.gpu_launches += tr.apply_optimizer(self.slot_views[slot] if last else None, publish_scale)
        self._publish_buffers(slot, publish_scale)
        self._signal_ready(j)
        self.completed_batches[self.learner_id(j)] += n
        # the CUDA graph writes its loss into one static tensor: detach this task's value from it
        return loss.clone() if self.k > 1 else loss

    # ------------------------------------------------------------------------------------------
    def aggregate_sync(self, participants: Optional[Sequence[int]] = None, weights: Optional[Sequence[float]] = None,
                       sources: Optional[Sequence] = None, accumulate: bool = False):
        with self._phase("aggregate"):
            return self._aggregate_sync(participants, weights, sources, accumulate)

    def _aggregate_sync(self, participants: Optional[Sequence[int]] = None, weights: Optional[Sequence[float]] = None,
                        sources: Optional[Sequence] = None, accumulate: bool = False):
        """FedAvg / FedStride over all (or the given) learners: ONE fused launch per GPU.

        ``weights`` overrides the scaling factors computed here (the controller computes them on the service
        path); ``sources`` = ``[(rank, slot index)]`` names the slots explicitly (staging slots of off-box learners
        included) instead of deriving them from ``participants``; ``accumulate`` folds the current community
        model in with weight 1 (one block of a FedStride stream commanded from outside)."""
        cfg = self.cfg
        ids = list(range(self.num_learners)) if participants is None else list(participants)
        if sources is not None:
            assert weights is not None and len(weights) == len(sources)
            return self._aggregate_sources(list(sources), list(weights), accumulate)
        w = list(weights) if weights is not None else scaling_factors(
            cfg.scaling, [self.examples[i] for i in ids],
            [max(1, self.completed_batches[i]) for i in ids], self.num_learners)
        if weights is None and len(ids) == 1:
            w = [1.0]   # a lone participant's factor from the scalers is its RAW count (for the rolling rules, which
            #             divide by z); the synchronous rules multiply by it as it is
        stride = cfg.stride_length if (cfg.rule == "FedStride" and cfg.stride_length > 0) else len(ids)
        stride = min(stride, 64)
        install = self.arena.master[0] if (self.direct_install and self.arena.master[0] is not self.arena.seg[0]) else None
        if self.device.type != "cuda" or not self.fabric.peer_mapped:
            return self._aggregate_host(ids, w)
        scaled = [abs(self._slot_scale[j] - w_full) < 1e-12 for j, w_full in
                  ((j, self._static_weight(self.learner_id(j)) if self.nvls else 1.0) for j in range(self.k))]
        use_nvls = self.nvls and participants is None and weights is None and stride >= len(ids) and all(scaled)
        if self.nvls and not use_nvls and any(sc != 1.0 for sc in self._slot_scale):
            raise RuntimeError("slots hold pre-scaled models (NVLS publish) but this aggregation needs the "
                               "peer-pointer path: pass publish_scale=1.0 to train_learner or set nvls=False")
        if stride >= len(ids):
            srcs = [self._slot_ptr(i // self.k, (i % self.k) * self.ring) for i in ids]
            mc = [self.slots.mc_ptr + j * self.layout.total_bytes for j in range(self.k)] if use_nvls else None
            agg_ops.fused_aggregate(srcs, w, self.community.ptrs, self.pads.ptrs, self.rank, self.world,
                                    self.layout.segment_table(), install_f32=install, n_blocks=cfg.fused_blocks,
                                    mc_src_ptrs=mc, mc_out_ptr=self.community.mc_ptr if use_nvls else 0)
            self.gpu_launches += 1
        else:
            # FedStride: memory-bounded blocks; every block is a fused launch whose partial result
            # is folded into the running community model by giving it weight 1 as an extra source.
            first = True
            for b in range(0, len(ids), stride):
                blk = ids[b:b + stride]
                srcs = [self._slot_ptr(i // self.k, (i % self.k) * self.ring) for i in blk]
                ws = w[b:b + stride]
                if not first:
                    srcs = [self.community.ptrs[self.rank]] + srcs
                    ws = [1.0] + list(ws)
                lastb = b + stride >= len(ids)
                agg_ops.fused_aggregate(srcs, ws, self.community.ptrs, self.pads.ptrs, self.rank, self.world,
                                        self.layout.segment_table(), install_f32=install if lastb else None,
                                        n_blocks=cfg.fused_blocks)
                self.gpu_launches += 1
                first = False
        self.global_iteration += 1
        if cfg.quantify:
            self.quantify_community()

    def _aggregate_sources(self, sources, weights, accumulate: bool):
        """One fused launch over explicitly named slots (the controller-commanded path)."""
        install = self.arena.master[0] if (self.direct_install and self.arena.master[0] is not self.arena.seg[0]) else None
        nb = self.layout.total_bytes
        # the (rank, slot) pairs come from messages (a learner's device_slot reference, relayed by the controller): they
        # are turned into device addresses below, so they are bounded here
        for r, sl in sources:
            if not (0 <= int(r) < self.world and 0 <= int(sl) < self.n_slots):
                raise ValueError("aggregation source (rank {}, slot {}) is outside this fabric ({} ranks, {} slots per "
                                 "rank)".format(r, sl, self.world, self.n_slots))
        if len(weights) != len(sources):
            raise ValueError("aggregation needs one weight per source")
        if self.device.type != "cuda" or not self.fabric.peer_mapped:
            return self._aggregate_host(None, weights, sources=sources, accumulate=accumulate)
        srcs = [self._slot_ptr(int(r), int(sl)) for r, sl in sources]
        ws = list(weights)
        if accumulate:
            srcs = [self.community.ptrs[self.rank]] + srcs
            ws = [1.0] + ws
        if any(sc != 1.0 for sc in self._slot_scale):
            raise RuntimeError("slots hold pre-scaled models (NVLS publish); controller-commanded aggregation needs "
                               "publish_scale=1.0")
        for b in range(0, len(srcs), 64):  # MFL_MAX_SRC sources per launch; later launches fold the running result in
            blk, wb = srcs[b:b + 64], ws[b:b + 64]
            if b > 0:
                blk, wb = [self.community.ptrs[self.rank]] + blk[:63], [1.0] + wb[:63]
            last = b + 64 >= len(srcs)
            agg_ops.fused_aggregate(blk, wb, self.community.ptrs, self.pads.ptrs, self.rank, self.world,
                                    self.layout.segment_table(), install_f32=install if last else None,
                                    n_blocks=self.cfg.fused_blocks)
            self.gpu_launches += 1
        self.global_iteration += 1

    def _aggregate_host(self, ids, w, sources=None, accumulate: bool = False):
        """CPU / non-peer-mapped path (gloo): gather slots, run the same math with the CPU ops."""
        import torch.distributed as dist
        nb = self.layout.total_bytes
        if self.world > 1:
            gathered = [torch.empty_like(self.slots.local) for _ in range(self.world)]
            dist.all_gather(gathered, self.slots.local, group=self.fabric.group)
        else:
            gathered = [self.slots.local]
        pairs = list(sources) if sources is not None else [(i // self.k, (i % self.k) * self.ring) for i in ids]
        for si, s in enumerate(self.layout.segments):
            srcs = []
            for r, sl in pairs:
                buf = gathered[int(r)][int(sl) * nb:(int(sl) + 1) * nb]
                srcs.append(buf[s.byte_offset:s.byte_offset + s.nbytes].view(s.dtype))
            out = self.community_views[si]
            agg_ops.weighted_sum(srcs, w, out=out, exact=s.exact, accumulate=accumulate)
            if self.device.type == "cuda":
                self.gpu_launches += 1
        self.global_iteration += 1
        if self.direct_install:
            self.arena.sync_master_from_model()
        if self.cfg.quantify:
            self.quantify_community()

    # ------------------------------------------------------------------------------------------
    # -- asynchronous rules (FedRec / FedAsync): one executor GPU folds, everybody else only publishes ---------
    def _init_async_state(self):
        """fp32 rolling state on the executor, completion counter, status words, streams and events."""
        self._state = self._state_off = None
        self.rec_z, self.rec_w = 0.0, {}
        if not self.async_rule:
            return
        off, offs = 0, []
        for sg in self.layout.segments:
            offs.append(off)
            if not sg.exact:
                off += (sg.numel + 3) // 4 * 4
        self._state_off = offs
        self.is_executor = self.rank == self.cfg.executor_rank or not self.fabric.peer_mapped
        if self.is_executor:
            self._state = torch.zeros(max(off, 4), dtype=torch.float32, device=self.device)
            if self.cfg.rule == "FedAsync":  # the state is the fp32 master of the community model itself
                for si, sg in enumerate(self.layout.segments):
                    if not sg.exact:
                        self._state[offs[si]:offs[si] + sg.numel].copy_(self._comm_views[0][si])
        cuda = self.device.type == "cuda"
        self._done = torch.zeros(4, dtype=torch.int32, device=self.device) if cuda else None
        self._status = torch.zeros(2, dtype=torch.int64, device=self.device) if cuda else None
        self.fold_stream = torch.cuda.Stream(self.device) if cuda else None
        self._train_events = [None] * self.k      # task of local learner j is published (executor-local ordering)
        self._fold_events = {}                    # global learner -> event of its last fold (executor only)
        self._need_version = [0] * self.k         # version a local learner's next task must at least see
        if cuda:
            self.ctrl.local.zero_()

    def _ready_flag(self, learner: int) -> int:
        return self.ctrl.ptrs[self.cfg.executor_rank if self.fabric.peer_mapped else 0] + 8 * (8 + learner)

    def _install_community(self, j: int, min_version: Optional[int] = None):
        """Newest community model -> arena (+ fp32 master), after this learner's previous fold."""
        nb = self.layout.total_bytes
        need = max(self._need_version[j], int(min_version or 0))
        if self.device.type != "cuda":
            self.arena.load_state_bytes(self.community_bytes())
            return
        g = self.learner_id(j)
        if self.is_executor:
            ev = self._fold_events.get(g)
            if ev is not None:
                torch.cuda.current_stream(self.device).wait_event(ev)
        elif need > 0:
            agg_ops.flag_wait(self.ctrl.ptrs[self.rank], need)  # bounded in-kernel wait on the local version word
            self.gpu_launches += 1
        master = self.arena.master[0] if self.arena.master[0] is not self.arena.seg[0] else None
        seg0 = self.layout.segment_table()[0] if master is not None else None
        comm = [self.community.ptrs[self.rank] + b * nb for b in range(self.n_comm)]
        agg_ops.snapshot_install(comm, self.ctrl.ptrs[self.rank], self._status, self.arena.state, nb, master, seg0)
        for si, (sg, m, v) in enumerate(zip(self.layout.segments, self.arena.master, self.arena.seg)):
            if si > 0 and sg.trainable and m is not None and m is not v:
                m.copy_(v)  # further low-precision trainable segments (none in the shipped models)
        self.gpu_launches += 2

    def _signal_ready(self, j: int):
        if not self.async_rule or self.device.type != "cuda":
            return
        if self.is_executor:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self._train_events[j] = ev
        else:
            agg_ops.flag_signal(self._ready_flag(self.learner_id(j)), self.task_counter[j])
            self.gpu_launches += 1

    def fold_async(self, learner: int, staleness: Optional[float] = None, weight: Optional[float] = None,
                   task_no: Optional[int] = None, mix: Optional[float] = None, z: Optional[float] = None,
                   previous_weight: Optional[float] = None, slot: Optional[int] = None,
                   previous_slot: Optional[int] = None):
        """Fold the newest contribution of global learner ``learner`` into the community model (K3).

        Every rank calls this for every fold, in the same order - that is how they agree on version numbers -
        but only the executor launches anything: its kernel waits for the owner's "published" counter, reads
        the slot (and for FedRec the previous one) over NVLink, updates the fp32 state and pushes the new
        community model and version to every GPU.  ``mix`` / ``z`` / the weights default to what the rule and this
        engine's own bookkeeping give (the service path passes the controller's numbers)."""
        cfg = self.cfg
        assert self.async_rule, f"fold_async: rule {cfg.rule} is synchronous"
        r, j = divmod(learner, self.k)
        nb = self.layout.total_bytes
        # learner index and slot numbers arrive in messages on the service path and become device addresses below
        if not (0 <= r < self.world):
            raise ValueError("fold_async: learner {} is outside this fabric ({} learners)".format(learner, self.world * self.k))
        for name, val in (("slot", slot), ("previous_slot", previous_slot)):
            if val is not None and not (0 <= int(val) < self.n_slots):
                raise ValueError("fold_async: {} {} is outside the {} slots of a rank".format(name, val, self.n_slots))
        if task_no is None:
            task_no = self.task_counter[j] if r == self.rank else self._remote_tasks.setdefault(learner, 0) + 1
        if r != self.rank:
            self._remote_tasks[learner] = task_no
        if slot is None:
            slot = j * self.ring + (task_no - 1) % self.ring
        if staleness is None:
            staleness = self.version - self.learner_version[learner]
        x_old_slot = None
        if cfg.rule == "FedAsync":
            a = mix if mix is not None else cfg.fedasync_alpha * staleness_factor(
                cfg.staleness_fn, staleness, cfg.staleness_a, cfg.staleness_b)
            decay, w_new, w_old, post = 1.0 - a, a, 0.0, 1.0
        else:
            w_new = weight if weight is not None else (
                float(self.examples[learner]) if cfg.scaling != "NumParticipants" else 1.0)
            had = learner in self.rec_w
            w_old = (previous_weight if previous_weight is not None else self.rec_w[learner]) if had else 0.0
            if had:
                x_old_slot = previous_slot if previous_slot is not None else j * self.ring + (task_no - 2) % self.ring
                assert self.ring >= 2, "FedRec needs the previous contribution: ring >= 2"
            self.rec_z = z if z is not None else self.rec_z + w_new - w_old
            self.rec_w[learner] = w_new
            decay, post = 1.0, 1.0 / self.rec_z
        new_version = self.version + 1
        out_buf = new_version % self.n_comm
        if self.device.type == "cuda" and self.fabric.peer_mapped:
            if self.is_executor:
                cur = torch.cuda.current_stream(self.device)
                wait_flag = 0
                if r == self.rank:
                    if self._train_events[j] is not None:
                        self.fold_stream.wait_event(self._train_events[j])
                    else:
                        self.fold_stream.wait_stream(cur)
                else:
                    wait_flag = self._ready_flag(learner)
                outs = [self.community.ptrs[t] + out_buf * nb for t in range(self.world)]
                mc = self.community.mc_ptr + out_buf * nb if self.community.mc_ptr else 0
                agg_ops.async_fold(
                    self._slot_ptr(r, slot), w_new, self._slot_ptr(r, x_old_slot) if x_old_slot is not None else 0,
                    w_old, decay, post, self._state, self._state_off, self.layout.segment_table(), outs, mc,
                    [self.ctrl.ptrs[t] for t in range(self.world)], new_version, wait_flag, task_no, self._done,
                    n_blocks=cfg.fold_blocks, stream=self.fold_stream)
                ev = torch.cuda.Event()
                ev.record(self.fold_stream)
                self._fold_events[learner] = ev
                self.gpu_launches += 1
        else:
            self._fold_host(r, slot, x_old_slot, w_new, w_old, decay, post, out_buf)
        if r == self.rank:
            self._need_version[j] = new_version
        self.contributed[learner] = True
        self.version = new_version
        self.global_iteration = new_version
        self.learner_version[learner] = new_version

    def _fold_host(self, r, slot, x_old_slot, w_new, w_old, decay, post, out_buf):
        """CPU tier (gloo): the owner broadcasts its slot(s), every rank does the same arithmetic."""
        import torch.distributed as dist
        nb = self.layout.total_bytes

        def fetch(sl):
            buf = self.slots.local[sl * nb:(sl + 1) * nb].clone() if r == self.rank else \
                torch.empty(nb, dtype=torch.uint8, device=self.device)
            if self.world > 1:
                dist.broadcast(buf, src=r, group=self.fabric.group)
            return self.layout.views(buf)
        xn = fetch(slot)
        xo = fetch(x_old_slot) if x_old_slot is not None else None
        for si, sg in enumerate(self.layout.segments):
            out = self._comm_views[out_buf][si]
            if sg.exact:
                out.copy_(xn[si])
                continue
            st = self._state[self._state_off[si]:self._state_off[si] + sg.numel]
            st.mul_(decay).add_(xn[si].float(), alpha=w_new)
            if xo is not None:
                st.add_(xo[si].float(), alpha=-w_old)
            out.copy_((st * post).to(out.dtype))

    def async_health(self) -> Dict:
        """Torn community reads seen by this rank's installs and in-kernel waits that timed out (both should be 0)."""
        if self.device.type != "cuda" or not self.async_rule:
            return {"torn_reads": 0, "spin_timeouts": 0}
        torch.cuda.synchronize(self.device)
        return {"torn_reads": int(self._status[1].item()), "spin_timeouts": agg_ops.spin_timeouts()}

    def sync_folds(self):
        """Host-side join of the fold stream (executor) - call before reading the community model on the host."""
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        self.fabric.host_barrier()

    # ------------------------------------------------------------------------------------------
    def quantify_community(self):
        if self.quantifier is None:
            self.quantifier = []
            for si, s in enumerate(self.layout.segments):
                vs = [v for v in self.layout.variables if v.segment == si]
                # offsets incl. padding gaps as pseudo-variables so counts map 1:1 to real variables
                offs, keep, cur = [], [], 0
                for v in vs:
                    if v.elem_offset > cur:
                        offs.append(cur); keep.append(False)
                    offs.append(v.elem_offset); keep.append(True)
                    cur = v.elem_offset + v.numel
                offs.append(cur)
                self.quantifier.append((agg_ops.Quantifier(offs, self.device), keep, vs))
        out = []
        for si, (q, keep, vs) in enumerate(self.quantifier):
            counts = q(self.community_views[si])
            if self.device.type == "cuda":
                self.gpu_launches += 2
            out.append((counts, keep, vs))
        self._last_quant = out
        return out

    def quantifier_report(self):
        """Host-side dict name -> (non_zeros, zeros, bytes) of the last quantified community model."""
        rep = {}
        for counts, keep, vs in getattr(self, "_last_quant", []):
            c = counts.cpu().tolist()
            it = iter(vs)
            for row, k in zip(c, keep):
                if k:
                    v = next(it)
                    rep[v.name] = (row[0], row[1], v.numel * v.dtype.itemsize)
        return rep

    # ------------------------------------------------------------------------------------------
    def run_round(self, batches_per_learner: Sequence[Sequence], evaluate_batch=None) -> Dict:
        """One synchronous federation round over the learners hosted by this rank.

        ``batches_per_learner[j]`` = list of ``(x, y)`` for local learner j (host-pinned or device).
        Returns the per-learner last-step losses (host floats) and, if requested, the community
        model's evaluation on ``evaluate_batch`` - both read back through pinned memory.
        """
        assert self.cfg.rule in ("FedAvg", "FedStride", "PWA"), "run_round drives the synchronous rules"
        if self.secure is not None:   # encrypted FedAvg: encrypt -> fused ciphertext reduction over NVLink -> decrypt
            self.secure.round(batches_per_learner[0])
            return {"loss": [], "eval": None, "phase_ms": self.secure.phase_ms(reset=False)}
        losses = []
        for j in range(self.k):
            losses.append(self.train_learner(j, batches_per_learner[j]))
        self.aggregate_sync()
        ev = None
        if evaluate_batch is not None:
            if not self.direct_install:
                self.arena.load_state_bytes(self.community.local[:self.layout.total_bytes])
            x, y = evaluate_batch
            ev = self.trainer.evaluate(x.to(self.device, non_blocking=True), y.to(self.device, non_blocking=True))
        return self._read_back(losses, ev)

    def _read_back(self, losses, ev):
        vals = [l for l in losses if l is not None]
        if ev is not None:
            vals += [ev[0]] + ([ev[1].float()] if ev[1] is not None else [])
        if not vals:
            return {"loss": [], "eval": None}
        dev = torch.stack([v.float().reshape(()) for v in vals])
        if self.device.type == "cuda":
            if self._pinned_loss is None or self._pinned_loss.numel() != dev.numel():
                self._pinned_loss = torch.empty(dev.numel(), dtype=torch.float32, pin_memory=True)
            self._pinned_loss.copy_(dev, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            host = self._pinned_loss.tolist()
            self.gpu_launches += 2
        else:
            host = dev.tolist()
        nl = len([l for l in losses if l is not None])
        out = {"loss": host[:nl], "eval": None}
        if ev is not None:
            out["eval"] = {"loss": host[nl], "correct": host[nl + 1] if ev[1] is not None else None}
        return out

    # ------------------------------------------------------------------------------------------
    def run_round_streaming(self, n_local_learners: int, batches_fn: Callable[[int], Sequence],
                            examples_fn: Optional[Callable[[int], float]] = None) -> Dict:
        """Many-learner round (config #5): ``n_local_learners`` logical learners are time-multiplexed on
        this GPU and their models are folded into an fp32 running sum as soon as they are trained -
        no per-learner model is ever kept (FedStride's memory-bounded idea taken to stride 1,
        ``federated_stride.cc:6-48``).  The per-rank partial sums are then combined by ONE fused
        cross-GPU launch.  ``batches_fn(g)`` yields the batches of global learner ``g``.
        """
        assert self.k >= 2, "streaming needs learners_per_gpu >= 2 slots (slot 0: partial sum, slot 1: publish target)"
        n_total = n_local_learners * self.world
        ex = examples_fn or (lambda g: 1.0)
        # every rank needs the global normaliser; weights depend only on the (deterministic) example counts
        z = float(sum(ex(g) for g in range(n_total)))
        accs = [torch.zeros(sg.numel, dtype=torch.float32, device=self.device) if not sg.exact else None
                for sg in self.layout.segments]
        losses = []
        first = True
        for j in range(n_local_learners):
            g = self.rank * n_local_learners + j
            loss = self._train_into_slot(1, batches_fn(g))
            losses.append(loss)
            w = float(ex(g)) / z
            last = j == n_local_learners - 1
            for si, sg in enumerate(self.layout.segments):
                x = self.slot_views[1][si]
                if sg.exact:
                    agg_ops.weighted_sum([x], [w], out=self.slot_views[0][si], exact=True, accumulate=not first)
                else:
                    agg_ops.weighted_sum([x], [w], out=self.slot_views[0][si] if last else None, acc=accs[si],
                                         accumulate=not first, write_out=last)
                self.gpu_launches += 1
            first = False
        srcs = [self._slot_ptr(r, 0) for r in range(self.world)]
        if self.device.type == "cuda" and self.fabric.peer_mapped:
            mc_ok = bool(self.slots.mc_ptr and self.community.mc_ptr)
            agg_ops.fused_aggregate(srcs, [1.0] * self.world, self.community.ptrs, self.pads.ptrs, self.rank,
                                    self.world, self.layout.segment_table(), n_blocks=self.cfg.fused_blocks,
                                    mc_src_ptrs=[self.slots.mc_ptr] if mc_ok else None,
                                    mc_out_ptr=self.community.mc_ptr if mc_ok else 0)
            self.gpu_launches += 1
            self.global_iteration += 1
        else:
            self._aggregate_host([r * self.k for r in range(self.world)], [1.0] * self.world)
        return self._read_back([l for l in losses[-1:]], None)

    def _train_into_slot(self, j: int, batches: Sequence):
        """``train_learner`` for multiplexed learners that share slot ``j``."""
        done = self.completed_batches[self.learner_id(j)]
        loss = self.train_learner(j, batches, publish_scale=1.0)  # the streaming fold applies the weights
        self.completed_batches[self.learner_id(j)] = done
        return loss

    # -- checkpoint / resume: the durable format is the wire FederatedModel -------------
    def save_checkpoint(self, path: str):
        from types import SimpleNamespace

        from metisfl_b200.ops import pack
        from metisfl_b200.proto import model_pb2
        model = pack.arena_to_model_pb(SimpleNamespace(state=self.community_bytes(), layout=self.layout))
        fm = model_pb2.FederatedModel(num_contributors=self.num_learners, global_iteration=self.global_iteration,
                                      model=model)
        tmp = path + ".tmp"
        with open(tmp, "wb") as f:
            f.write(fm.SerializeToString())
        import os
        os.replace(tmp, path)

    def load_checkpoint(self, path: str):
        from types import SimpleNamespace

        from metisfl_b200.ops import pack
        from metisfl_b200.proto import model_pb2
        with open(path, "rb") as f:
            fm = model_pb2.FederatedModel.FromString(f.read())
        self.install_community_model(fm.model)
        self.global_iteration = fm.global_iteration
        return fm.global_iteration

    def install_community_model(self, model_pb):
        """Wire model (with bytes) -> this rank's community buffer(s), unpacked on the device (K12).  Used for the
        initial ship of the community model and for resuming from a checkpoint."""
        from types import SimpleNamespace

        from metisfl_b200.ops import pack
        nb = self.layout.total_bytes
        shim = SimpleNamespace(state=self.community_bytes(), layout=self.layout, device=self.device,
                               sync_master_from_model=lambda: None)
        pack.unpack_model_into_arena(model_pb, shim)
        cur = self.version % self.n_comm
        for b in range(self.n_comm):
            if b != cur:
                self.community.local[b * nb:(b + 1) * nb].copy_(self.community_bytes())
        if self.async_rule and self._state is not None and self.cfg.rule == "FedAsync":
            for si, sg in enumerate(self.layout.segments):
                if not sg.exact:
                    self._state[self._state_off[si]:self._state_off[si] + sg.numel].copy_(self.community_views[si])
        if self.direct_install:
            self.arena.sync_master_from_model()

    def stage_model(self, model_pb, staging_index: int) -> int:
        """Unpack an off-box learner's wire model into staging slot ``staging_index`` of this rank (K12); returns the
        slot index to name in an aggregation."""
        from types import SimpleNamespace

        from metisfl_b200.ops import pack
        assert 0 <= staging_index < self.cfg.staging_slots, "not enough staging slots (EngineConfig.staging_slots)"
        slot = self.k * self.ring + staging_index
        nb = self.layout.total_bytes
        shim = SimpleNamespace(state=self.slots.local[slot * nb:(slot + 1) * nb], layout=self.layout,
                               device=self.device, sync_master_from_model=lambda: None)
        pack.unpack_model_into_arena(model_pb, shim)
        return slot

    # -- control-only model references (TensorSpec.device_slot) ---------------------------------------------------
    COMMUNITY_SLOT = 0xFFFFFFFF

    def reference_model_pb(self, slot: int, version: int, rank: Optional[int] = None):
        """``metisfl.Model`` that describes every variable (name, shape, dtype) but carries no bytes: each TensorSpec
        points at ``slot`` of ``rank``'s slot buffer (``COMMUNITY_SLOT`` = the community buffer).  This is what
        ``MarkTaskCompleted`` / ``RunTask`` / ``EvaluateModel`` carry between processes that share the fabric."""
        from metisfl_b200.proto import model_pb2
        from metisfl_b200.utils import dtypes
        m = model_pb2.Model()
        rank = self.rank if rank is None else rank
        for info in self.layout.variables:
            seg = self.layout.segments[info.segment]
            var = m.variables.add(name=info.name, trainable=info.trainable)
            spec = var.plaintext_tensor.tensor_spec
            spec.length = info.numel
            spec.dimensions.extend(info.shape)
            spec.type.type = dtypes.torch_to_wire(seg.dtype)
            spec.type.byte_order = model_pb2.DType.NA if seg.dtype.itemsize == 1 else model_pb2.DType.LITTLE_ENDIAN_ORDER
            spec.device_slot.device_rank = rank
            spec.device_slot.slot = slot
            spec.device_slot.offset_bytes = seg.byte_offset + info.elem_offset * seg.dtype.itemsize
            spec.device_slot.version = version
        return m

    def community_model_pb(self):
        """The current community model WITH bytes (one device->host copy), e.g. for off-box learners or snapshots."""
        from types import SimpleNamespace

        from metisfl_b200.ops import pack
        return pack.arena_to_model_pb(SimpleNamespace(state=self.community_bytes(), layout=self.layout))

    def community_state_dict(self) -> Dict[str, torch.Tensor]:
        views = self.community_views
        return {v.name: self.layout.var_view(views, v).clone() for v in self.layout.variables}

    def close(self):
        self.fabric.close()
