"""Secure (CKKS) aggregation as an engine rule: ``EngineConfig(rule="PWA")``.

Reference flow: the driver generates the crypto context and the keys and ships the files
(``driver_session.py:110-140``); learners hold context + public + private key and encrypt every tensor of their local
model (``model_proto_factory.py:108-114``); the controller holds the context only and computes
``sum_l EvalMult(ct_l, w_l)`` (``private_weighted_average.cc:24-82``, ``ckks_scheme.cc:165-207``); learners decrypt the
community model (``model_ops.py:32-43``).  Here the same roles run on the GPUs of one box:

* keys come from FILES (``crypto_dir``: ``cryptocontext.txt``, ``key-public.txt``, ``key-private.txt``), generated once
  by whoever plays the driver (``SecureAggregator.generate_keys``) - never derived from a shared seed;
* each learner encrypts its published slot on its GPU (K6 encode + K7 encrypt) into a peer-visible ciphertext buffer;
* ONE fused kernel per GPU (K8) reduces the ciphertexts limb-wise across peers over NVLink and stores the encrypted
  community model into every GPU - the "controller" work needs no key and touches no plaintext;
* each learner decrypts (K7 + K6) and installs the result.

Every phase is timed with CUDA events (``phase_ms``), which is what ``examples/resnet_secure.py`` reports.
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Sequence

import torch

from metisfl_b200.utils.tracing import DeviceTimer

FILES = ("cryptocontext.txt", "key-public.txt", "key-private.txt")


class SecureAggregator:
    def __init__(self, engine, crypto_dir: str, batch_size: int = 4096, scaling_factor_bits: int = 52,
                 multicast_output: bool = False):
        from metisfl_b200.encryption import fhe
        from metisfl_b200.ops import ckks as gpu_ckks
        self.e = engine
        self.crypto_dir = crypto_dir
        for f in FILES:
            if not os.path.exists(os.path.join(crypto_dir, f)):
                raise FileNotFoundError(f"{f} is missing from {crypto_dir}: generate the keys first "
                                        "(SecureAggregator.generate_keys, or DriverSession with HomomorphicEncryption)")
        host = fhe.CKKS(batch_size, scaling_factor_bits, backend="host")
        host.load_crypto_context_from_file(os.path.join(crypto_dir, FILES[0]))
        host.load_public_key_from_file(os.path.join(crypto_dir, FILES[1]))
        host.load_private_key_from_file(os.path.join(crypto_dir, FILES[2]))
        self.host = host
        self.g = gpu_ckks.GpuCKKS(host._scheme, host._params, device=engine.device)
        seg = engine.layout.segments[0]
        assert seg.trainable and not seg.exact, "segment 0 (the trainable parameters) travels encrypted"
        self.n = seg.numel
        self.n_ct = self.g.n_ct(self.n)
        self.words = self.g.ct_words(self.n_ct)
        self.ct_slot = engine.fabric.alloc(self.words * 8)   # this learner's ciphertext, peer-visible
        # encrypted community model, written by all peers with plain peer stores.  (K8 can also store through a
        # multicast mapping - ``multicast_output=True`` - but the kernel is bound by INBOUND link bytes, and the
        # switch's copy of a GPU's own shard adds 1/world of a model to them: 5.99 vs 5.85 ms at 8 GPUs.)
        self.ct_out = engine.fabric.alloc(self.words * 8, multicast=bool(multicast_output) and engine.world > 1)
        self.timer = DeviceTimer(engine.device)
        self.last_error: Optional[float] = None

    @staticmethod
    def generate_keys(crypto_dir: str, batch_size: int = 4096, scaling_factor_bits: int = 52):
        """The driver's part (``driver_session.py:115-124``): fresh keys from OS entropy, written as files."""
        from metisfl_b200.encryption import fhe
        os.makedirs(crypto_dir, exist_ok=True)
        scheme = fhe.CKKS(batch_size, scaling_factor_bits, backend="host")
        scheme.gen_crypto_context_and_keys(crypto_dir)
        return scheme.get_crypto_params_files()

    @property
    def ciphertext_bytes(self) -> int:
        return self.words * 8

    def round(self, batches: Sequence, weights: Optional[Sequence[float]] = None, check: bool = False,
              sync_phases: bool = False):
        """One secure federation round for the learner of this rank (one learner per GPU).  ``sync_phases`` lines the
        ranks up (host barrier, outside the timers) before the cross-GPU reduction, so that its timer measures the
        kernel and not the slowest rank's encryption."""
        import torch.distributed as dist
        e, g = self.e, self.g
        w = list(weights) if weights is not None else [1.0 / e.world] * e.world
        with self.timer.phase("train"):
            e.train_learner(0, batches, publish_scale=1.0)   # the slot feeds the encoder: keep it unscaled
        x = e.slot_views[0][0]
        with self.timer.phase("encode_encrypt"):
            g.encrypt_device(x, out=self.ct_slot.local.view(torch.int64)[:self.words])  # randomness: os.urandom per call
        if sync_phases and e.world > 1:
            torch.cuda.synchronize(e.device)
            dist.barrier()
        with self.timer.phase("reduce_over_nvlink"):
            g.fused_weighted_sum(self.ct_slot.ptrs, w, self.words, self.ct_out.ptrs, e.pads.ptrs, e.rank, e.world,
                                 n_blocks=148, mc_out=self.ct_out.mc_ptr)
        with self.timer.phase("decrypt_decode"):
            avg = g.decrypt_device(self.ct_out.local.view(torch.int64)[:self.words], self.n, scale=g.scale * g.scale)
        with self.timer.phase("install"):
            e.arena.seg[0].copy_(avg.to(e.arena.seg[0].dtype))
            e.arena.sync_master_from_model()   # the fp32 master weights restart from the decrypted average
            if not e.direct_install:
                e.community_views[0].copy_(e.arena.seg[0])
        if check:  # plaintext oracle (NCCL all-reduce of the same models)
            ref = x.float() * w[e.rank]
            if e.world > 1:
                dist.all_reduce(ref)
            self.last_error = float((avg.float() - ref).abs().max())
        e.global_iteration += 1

    def phase_ms(self, reset: bool = True) -> Dict[str, float]:
        return self.timer.report(reset)
