#!/usr/bin/env python
"""BASELINE.json config #4: ResNet-18 secure aggregation with CKKS, one learner per GPU, encrypted
FedAvg over NVLink: every learner encrypts its trained model on its GPU (K6/K7), ONE fused kernel per
GPU reduces the ciphertexts across peers limb-wise (K8), every learner decrypts the encrypted community
model (K7/K6).  The decrypted result is checked against the plaintext weighted average.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/resnet_secure.py
"""
import argparse
import tempfile

from _common import device_time_ms, finish, init

import torch
import torch.distributed as dist

from metisfl_b200.encryption import fhe
from metisfl_b200.models.zoo import ResNet18
from metisfl_b200.ops import ckks as gpu_ckks
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine

ap = argparse.ArgumentParser()
ap.add_argument("--local-steps", type=int, default=2)
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--rounds", type=int, default=2)
a = ap.parse_args()
rank, world, dev = init()
eng = FederationEngine(ResNet18, EngineConfig(), device=dev)
# every rank derives the same keys from the same seed (stand-in for the driver shipping key files)
host = fhe.CKKS(4096, 52, backend="host")
host.gen_crypto_context_and_keys(tempfile.mkdtemp(), seed=2024)
g = gpu_ckks.GpuCKKS(host._scheme, host._params, device=dev)
n = eng.layout.segments[0].numel  # the bf16 trainable segment travels encrypted
n_ct = g.n_ct(n)
words = g.ct_words(n_ct)
ct_slot = eng.fabric.alloc(words * 8)        # this learner's ciphertext, peer-visible
ct_out = eng.fabric.alloc(words * 8)         # encrypted community model, written by all peers
gen = torch.Generator().manual_seed(5 + rank)
batches = [(torch.randn(a.batch, 3, 32, 32, generator=gen).to(torch.bfloat16).to(dev),
            torch.randint(0, 10, (a.batch,), generator=gen).to(dev)) for _ in range(a.local_steps)]
w = [1.0 / world] * world
err = [0.0]


def secure_round():
    eng.train_learner(0, batches, publish_scale=1.0)  # the slot feeds the encoder: keep it unscaled
    x = eng.slot_views[0][0]
    g.encrypt_device(x, seed=1000 + rank, out=ct_slot.local.view(torch.int64)[:words])
    g.fused_weighted_sum(ct_slot.ptrs, w, words, ct_out.ptrs, eng.pads.ptrs, rank, world, n_blocks=148)
    avg = g.decrypt_device(ct_out.local.view(torch.int64)[:words], n, scale=g.scale * g.scale)
    # plaintext oracle for the check (NCCL all-reduce of the same models)
    ref = x.float() / world
    if world > 1:
        dist.all_reduce(ref)
    err[0] = float((avg.float() - ref).abs().max())
    eng.arena.seg[0].copy_(avg.to(torch.bfloat16))
    eng.arena.sync_master_from_model()


secure_round()  # warm-up
ms = device_time_ms(lambda: [secure_round() for _ in range(a.rounds)], dev, world)
finish(rank, world, {
    "config": "#4 ResNet-18 CKKS secure aggregation", "gpus": world, "learners": world, "encrypted_params": n,
    "ciphertexts_per_model": n_ct, "ciphertext_bytes_per_model": words * 8,
    "round_time_s": ms * 1e-3 / a.rounds, "max_abs_error_vs_plaintext": err[0]}, engine=eng)
