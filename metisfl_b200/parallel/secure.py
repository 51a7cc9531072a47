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

from metisfl_b200.utils.tracing import DeviceTimer, nvtx_range

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


class FabricSecure:
    """The same three steps as ``SecureAggregator.round``, cut where the SERVICE path cuts them: the learner encrypts
    its published model when its task ends (``encrypt_slot``), the controller - which holds no key and sees no model -
    commands the reduction with the scaling factors it computed (``reduce``), and every learner decrypts the encrypted
    community model into its community buffer (``decrypt_install``).  Reference roles: ``learner.py:214-246`` (scheme
    per learner, keys from files), ``private_weighted_average.cc:24-82`` (controller side), ``model_ops.py:32-43``
    (decrypt at the learner).

    What travels encrypted is every floating-point segment of the arena (trainable parameters and, for models that
    have them, the floating-point buffers), concatenated; integer buffers (step counters) stay local to the learner.

    On CUDA the ciphertexts live in peer-visible buffers and the reduction is ONE fused launch per GPU over NVLink (K8);
    on the CPU tier (``gloo``, tests) the byte-blob host implementation does the same math.
    """

    SNAPSHOT_VARIABLE = "fabric/encrypted_segments"

    def __init__(self, engine, crypto_context_file: str, public_key_file: str, private_key_file: str,
                 batch_size: int = 4096, scaling_factor_bits: int = 52):
        from metisfl_b200.encryption import fhe
        self.e = engine
        assert engine.k == 1, "secure aggregation on the fabric: one learner per rank"
        host = fhe.CKKS(batch_size, scaling_factor_bits, backend="host")
        host.load_context_and_keys_from_files(crypto_context_file, public_key_file, private_key_file)
        self.host = host
        self.segments = [i for i, s in enumerate(engine.layout.segments) if not s.exact and s.dtype.is_floating_point]
        assert self.segments, "no floating-point segment to aggregate"
        self.n = sum(engine.layout.segments[i].numel for i in self.segments)
        self.cuda = engine.device.type == "cuda"
        if self.cuda:
            from metisfl_b200.ops import ckks as gpu_ckks
            self.g = gpu_ckks.GpuCKKS(host._scheme, host._params, device=engine.device)
            self.n_ct = self.g.n_ct(self.n)
            self.words = self.g.ct_words(self.n_ct)
            self.ct_slot = engine.fabric.alloc(self.words * 8)   # this learner's ciphertext, peer-visible
            self.ct_out = engine.fabric.alloc(self.words * 8)    # encrypted community model, written by the peers
        else:
            self.g = None
            self._blob = None       # this learner's ciphertext (host byte layout)
            self._out_blob = None   # encrypted community model

    # -- plaintext <-> the flat float64 vector the encoder takes ----------------------------------------------
    def _gather(self, views):
        parts = [views[i].reshape(-1).to(torch.float64) for i in self.segments]
        return parts[0] if len(parts) == 1 else torch.cat(parts)

    def _scatter(self, flat, views):
        off = 0
        for i in self.segments:
            n = self.e.layout.segments[i].numel
            views[i].copy_(flat[off:off + n].to(views[i].dtype).view_as(views[i]))
            off += n

    # -- learner: after its task -------------------------------------------------------------------------------
    def encrypt_slot(self, slot: int):
        x = self._gather(self.e.slot_views[slot])
        with nvtx_range("encode_encrypt"):
            if self.cuda:
                self.g.encrypt_device(x, out=self.ct_slot.local.view(torch.int64)[:self.words])
            else:
                self._blob = self.host.encrypt(x.numpy())

    # -- every rank: commanded by the controller ---------------------------------------------------------------
    def reduce(self, weights_by_rank: Sequence[float]):
        """``weights_by_rank[r]`` = scaling factor of rank r's learner (0 for a rank that sits this round out)."""
        e = self.e
        w = [float(v) for v in weights_by_rank]
        if len(w) != e.world:
            raise ValueError("secure reduction needs one weight per fabric rank")
        with nvtx_range("reduce_over_nvlink"):
            if self.cuda:
                self.g.fused_weighted_sum(self.ct_slot.ptrs, w, self.words, self.ct_out.ptrs, e.pads.ptrs, e.rank,
                                          e.world, n_blocks=148)
            else:
                blobs = e.fabric.all_gather_object(self._blob) if e.world > 1 else [self._blob]
                pairs = [(b, wi) for b, wi in zip(blobs, w) if wi != 0.0]
                if any(b is None for b, _ in pairs):
                    raise RuntimeError("a rank named in the aggregation has not published a ciphertext")
                self._out_blob = self.host.compute_weighted_average([b for b, _ in pairs], [wi for _, wi in pairs])
        e.global_iteration += 1

    def decrypt_install(self):
        e = self.e
        with nvtx_range("decrypt_decode"):
            if self.cuda:
                avg = self.g.decrypt_device(self.ct_out.local.view(torch.int64)[:self.words], self.n,
                                            scale=self.g.scale * self.g.scale)
            else:
                avg = torch.from_numpy(self.host.decrypt(self._out_blob, self.n))
        with nvtx_range("install"):
            self._scatter(avg, e.community_views)

    # -- encrypted snapshots (what the controller may store: it never sees plaintext) --------------------------
    def community_blob(self) -> bytes:
        if not self.cuda:
            return self._out_blob
        from metisfl_b200.ops import ckks as gpu_ckks
        g = self.g
        head = gpu_ckks.pack_header(g.log_n, g.n_limbs, self.n_ct, g.slots, self.n, g.scale * g.scale, g.moduli)
        return head + self.ct_out.local.view(torch.int64)[:self.words].cpu().numpy().tobytes()

    def decrypt_blob(self, blob: bytes, n_values: int):
        """Byte-layout ciphertext (the driver's initial model, a snapshot) -> float64 numpy array."""
        if self.cuda:
            return self.g.decrypt(blob, int(n_values))
        return self.host.decrypt(blob, int(n_values))

    def install_blob(self, blob: bytes):
        """An encrypted snapshot back into the community buffer (resume)."""
        avg = torch.from_numpy(self.decrypt_blob(blob, self.n)).to(self.e.device)
        self._scatter(avg, self.e.community_views)
