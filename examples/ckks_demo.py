"""CKKS walk-through (reference ``metisfl/encryption/palisade/ckks_demo.cc:39-96``): three learners,
weights 0.5 / 0.3 / 0.5, encrypt -> private weighted average -> decrypt."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from metisfl_b200.encryption import fhe  # noqa: E402

if __name__ == "__main__":
    scheme = fhe.CKKS(batch_size=4096, scaling_factor_bits=52)
    scheme.gen_crypto_context_and_keys(tempfile.mkdtemp(prefix="cryptoparams_"))
    print("key files:", scheme.get_crypto_params_files())
    rng = np.random.default_rng(0)
    models = [rng.normal(size=10000) for _ in range(3)]
    weights = [0.5, 0.3, 0.5]
    cts = [scheme.encrypt(m) for m in models]
    print("ciphertext bytes per model:", len(cts[0]))
    avg = scheme.compute_weighted_average(cts, weights)
    out = scheme.decrypt(avg, 10000)
    ref = sum(w * m for w, m in zip(weights, models))
    print("max |decrypt(PWA) - plaintext|:", float(np.abs(out - ref).max()))
