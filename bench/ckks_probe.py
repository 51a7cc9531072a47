#!/usr/bin/env python
"""CKKS phases on one GPU for `ncu -k regex:...` captures and per-kernel event timings: encode + encrypt, local
weighted sum, decrypt + decode of `--values` numbers (default: one ResNet-18 model, 2726 ciphertexts)."""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from metisfl_b200.encryption import fhe  # noqa: E402
from metisfl_b200.ops import ckks as gpu_ckks  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--values", type=int, default=11164368)
ap.add_argument("--iters", type=int, default=3)
a = ap.parse_args()
host = fhe.CKKS(4096, 52, backend="host")
host.gen_crypto_context_and_keys(tempfile.mkdtemp())
g = gpu_ckks.GpuCKKS(host._scheme, host._params, device="cuda:0")
x = torch.randn(a.values, device="cuda", dtype=torch.float64)
n_ct = g.n_ct(a.values)
ct = torch.empty(g.ct_words(n_ct), dtype=torch.int64, device="cuda")
out = torch.empty_like(ct)


def timed(fn):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / a.iters


t_enc = timed(lambda: g.encrypt_device(x, out=ct))
t_sum = timed(lambda: g.weighted_sum_device([ct, ct], [0.5, 0.5], out=out))
t_dec = timed(lambda: g.decrypt_device(out, a.values, scale=g.scale * g.scale))
res = g.decrypt_device(out, a.values, scale=g.scale * g.scale)
err = float((res - x).abs().max())
gb = ct.numel() * 8 / 1e9
print({"values": a.values, "ciphertexts": n_ct, "ciphertext_GB": gb, "encode_encrypt_ms": t_enc, "weighted_sum_ms": t_sum,
       "decrypt_decode_ms": t_dec, "weighted_sum_fraction_of_hbm_copy": (3 * gb / (t_sum * 1e-3)) / 6553.9,
       "max_abs_error": err})
