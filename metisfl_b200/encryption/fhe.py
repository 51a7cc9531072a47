"""``fhe.CKKS`` - the secure-aggregation scheme object used by driver, learners and tests.

Same Python surface as the reference's pybind class (``encryption/pybind/ckks_pybind.cc:71-85``):
``CKKS(batch_size, scaling_factor_bits)``, ``gen_crypto_context_and_keys(dir)``,
``get_crypto_params_files()``, ``load_*_from_file``, ``encrypt(float64[]) -> bytes``,
``compute_weighted_average(list[bytes], list[float]) -> bytes``, ``decrypt(bytes, n) -> float64[]``.
Key roles follow the reference: the controller needs only the crypto context; learners and the
driver hold context + public + private key (``driver_session.py:129-140``).

Backends: ``host`` = the C++ RNS-CKKS in ``csrc/ckks/ckks_host.cc`` (CPU oracle, always available);
``cuda`` = the sm_100a kernels in ``csrc/ckks/ckks_kernels.cu`` (used automatically for tensors that
are worth a launch when a GPU is present).  Both produce the same ciphertext byte layout.
"""
from __future__ import annotations

import os

import numpy as np

from metisfl_b200 import _core

CRYPTO_CONTEXT_FILE = "cryptocontext.txt"
PUBLIC_KEY_FILE = "key-public.txt"
PRIVATE_KEY_FILE = "key-private.txt"
EVAL_MULT_KEY_FILE = "key-eval-mult.txt"


class CKKS(object):

    def __init__(self, batch_size: int = 4096, scaling_factor_bits: int = 52, backend: str = "auto"):
        self.batch_size = int(batch_size)
        self.scaling_factor_bits = int(scaling_factor_bits)
        self.backend = backend
        self._params = _core.CKKSParams()
        self._params.batch_size = self.batch_size
        self._params.scale_bits = self.scaling_factor_bits
        self._scheme = _core.CKKSScheme(self._params)
        self._files = {}
        self._gpu = None

    # -- key / context management -----------------------------------------------------------------
    def gen_crypto_context_and_keys(self, crypto_dir: str, seed: int = 0):
        os.makedirs(crypto_dir, exist_ok=True)
        self._scheme.keygen(seed)
        self._files = {
            "crypto_context_file": os.path.join(crypto_dir, CRYPTO_CONTEXT_FILE),
            "public_key_file": os.path.join(crypto_dir, PUBLIC_KEY_FILE),
            "private_key_file": os.path.join(crypto_dir, PRIVATE_KEY_FILE),
            "eval_mult_key_file": os.path.join(crypto_dir, EVAL_MULT_KEY_FILE),
        }
        _core.ckks_save_params(self._files["crypto_context_file"], self._params)
        for key, blob in (("public_key_file", self._scheme.public_key_bytes()),
                          ("private_key_file", self._scheme.secret_key_bytes()),
                          ("eval_mult_key_file", self._scheme.eval_mult_key_bytes())):
            with open(self._files[key], "wb") as f:
                f.write(blob)
        self._gpu = None

    def get_crypto_params_files(self) -> dict:
        return dict(self._files)

    def load_crypto_context_from_file(self, filepath: str):
        keep_pk = self._scheme.public_key_bytes() if self._scheme.has_public_key() else None
        keep_sk = self._scheme.secret_key_bytes() if self._scheme.has_secret_key() else None
        self._params = _core.ckks_load_params(filepath)
        self.batch_size, self.scaling_factor_bits = self._params.batch_size, self._params.scale_bits
        self._scheme = _core.CKKSScheme(self._params)
        if keep_pk:
            self._scheme.load_public_key(keep_pk)
        if keep_sk:
            self._scheme.load_secret_key(keep_sk)
        self._files["crypto_context_file"] = filepath
        self._gpu = None

    def load_public_key_from_file(self, filepath: str):
        with open(filepath, "rb") as f:
            if not self._scheme.load_public_key(f.read()):
                raise RuntimeError(f"{filepath} is not a public key of this crypto context")
        self._files["public_key_file"] = filepath
        self._gpu = None

    def load_private_key_from_file(self, filepath: str):
        with open(filepath, "rb") as f:
            if not self._scheme.load_secret_key(f.read()):
                raise RuntimeError(f"{filepath} is not a private key of this crypto context")
        self._files["private_key_file"] = filepath
        self._gpu = None

    def load_context_and_keys_from_files(self, crypto_context_file, public_key_file, private_key_file):
        self.load_crypto_context_from_file(crypto_context_file)
        self.load_public_key_from_file(public_key_file)
        self.load_private_key_from_file(private_key_file)

    # -- operations -------------------------------------------------------------------------------
    def _use_gpu(self, n_values: int) -> bool:
        if self.backend == "host":
            return False
        try:
            import torch
            from metisfl_b200.ops import ckks as gpu_ckks
        except Exception:
            return False
        if not (torch.cuda.is_available() and gpu_ckks.available()):
            if self.backend == "cuda":
                raise RuntimeError("CKKS backend 'cuda' requested but no GPU / kernel library is available")
            return False
        return self.backend == "cuda" or n_values >= 4 * self.batch_size

    def _gpu_scheme(self):
        if self._gpu is None:
            from metisfl_b200.ops import ckks as gpu_ckks
            self._gpu = gpu_ckks.GpuCKKS(self._scheme, self._params)
        return self._gpu

    def encrypt(self, values, seed: int = 0) -> bytes:
        v = np.ascontiguousarray(np.asarray(values, dtype=np.float64).ravel())
        if self._use_gpu(v.size):
            return self._gpu_scheme().encrypt(v, seed)
        return self._scheme.encrypt(v, seed)

    def decrypt(self, ciphertext: bytes, num_values: int) -> np.ndarray:
        if self._use_gpu(int(num_values)):
            return self._gpu_scheme().decrypt(ciphertext, int(num_values))
        return self._scheme.decrypt(ciphertext, int(num_values))

    def compute_weighted_average(self, ciphertexts, scaling_factors) -> bytes:
        ciphertexts = list(ciphertexts)
        w = [float(s) for s in scaling_factors]
        n = _core.ckks_header(ciphertexts[0])["n_values"] if ciphertexts else 0
        if ciphertexts and self._use_gpu(n):
            from metisfl_b200.ops import ckks as gpu_ckks
            return gpu_ckks.weighted_average(ciphertexts, w)
        return _core.ckks_weighted_average(ciphertexts, w)

    def rescale(self, ciphertext: bytes) -> bytes:
        return _core.ckks_rescale(ciphertext)


if __name__ == "__main__":  # python -m metisfl_b200.encryption.fhe --gen DIR
    import argparse
    ap = argparse.ArgumentParser(description="generate a CKKS context and key files")
    ap.add_argument("--gen", required=True, metavar="DIR")
    ap.add_argument("--batch_size", type=int, default=4096)
    ap.add_argument("--scaling_factor_bits", type=int, default=52)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    scheme = CKKS(a.batch_size, a.scaling_factor_bits)
    scheme.gen_crypto_context_and_keys(a.gen, seed=a.seed)
    for k, v in scheme.get_crypto_params_files().items():
        print(k, v)
