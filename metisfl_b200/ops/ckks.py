"""GPU backend of ``fhe.CKKS``: device tables + front-ends of the K5-K9 kernels
(``csrc/ckks/ckks_kernels.cu``).  Ciphertexts can stay on the device (``*_device`` methods, used by
the on-box engine so that encrypted models never cross the host) or be exchanged as the same byte
blobs the host implementation produces (``encrypt`` / ``decrypt`` / ``weighted_average``).
Reference behaviour: ``encryption/palisade/ckks_scheme.cc:110-252``.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import struct
from typing import List, Sequence

import numpy as np
import torch

from metisfl_b200 import _core
from metisfl_b200.ops import _lib

HEADER_BYTES = 64


class _Ctx(C.Structure):
    _fields_ = [("q", C.c_uint64 * 4), ("n_inv", C.c_uint64 * 4), ("n_inv_shoup", C.c_uint64 * 4),
                ("psi", C.c_void_p * 4), ("psi_shoup", C.c_void_p * 4), ("ipsi", C.c_void_p * 4),
                ("ipsi_shoup", C.c_void_p * 4), ("log_n", C.c_int32), ("n_limbs", C.c_int32)]


_BOUND = False


def _bind():
    global _BOUND
    L = _lib.lib()
    if _BOUND:
        return L
    vp, u32, u64, sz, d, i = C.c_void_p, C.c_uint32, C.c_uint64, C.c_size_t, C.c_double, C.c_int
    P = C.POINTER
    L.mfl_ckks_ntt.argtypes = [P(_Ctx), vp, sz, i, vp]
    L.mfl_ckks_encode.argtypes = [P(_Ctx), vp, sz, u32, d, vp, vp, vp, vp]
    L.mfl_ckks_encrypt.argtypes = [P(_Ctx), vp, u32, u32, vp, vp, P(u32), vp, vp]
    L.mfl_ckks_decrypt.argtypes = [P(_Ctx), vp, u32, vp, vp, vp, vp]
    L.mfl_ckks_decode.argtypes = [P(_Ctx), vp, u32, u32, d, P(u64), vp, vp, vp, sz, vp]
    L.mfl_ckks_weighted_sum.argtypes = [P(_Ctx), P(vp), P(u64), P(u64), i, sz, vp, vp]
    L.mfl_ckks_fused_weighted_sum.argtypes = [P(_Ctx), P(vp), P(u64), P(u64), i, sz, P(vp), P(vp), i, i, i, vp]
    L.mfl_ckks_fused_weighted_sum_mc.argtypes = [P(_Ctx), P(vp), P(u64), P(u64), i, sz, P(vp), vp, P(vp), i, i, i, vp]
    L.mfl_ckks_rescale.argtypes = [P(_Ctx), vp, u32, P(u64), P(u64), vp, vp]
    _BOUND = True
    return L


def available() -> bool:
    return _lib.available()


def _u64(arr, device):
    return torch.from_numpy(np.ascontiguousarray(arr).view(np.int64)).to(device)


def _shoup(w: int, q: int) -> int:
    return (w << 64) // q


def pack_header(log_n, n_limbs, n_ct, slots, n_values, scale, moduli) -> bytes:
    q = list(moduli) + [0] * (4 - len(moduli))
    return struct.pack("<4sHBBIIQd4Q", b"MFCK", 1, log_n, n_limbs, n_ct, slots, n_values, scale, *q)


def scalar_weights(weights: Sequence[float], scale: float, moduli: Sequence[int]):
    """Delta-scaled integer encodings of the weights per limb, with Shoup companions (host side)."""
    n = len(weights)
    w = (C.c_uint64 * (4 * n))()
    wp = (C.c_uint64 * (4 * n))()
    for l, wt in enumerate(weights):
        mag = int(round(abs(wt) * scale))
        for i, q in enumerate(moduli):
            r = mag % q
            if wt < 0 and r:
                r = q - r
            w[4 * l + i] = r
            wp[4 * l + i] = _shoup(r, q)
    return w, wp


class GpuCKKS:
    """Device-resident crypto context: NTT tables, special-FFT tables and (optionally) the keys."""

    def __init__(self, host_scheme, params, device=None):
        self.L = _bind()
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.params = params
        self.log_n = int(params.log_n)
        self.N = 1 << self.log_n
        self.moduli = [int(q) for q in params.moduli]
        self.n_limbs = len(self.moduli)
        self.slots = int(params.batch_size)
        self.scale = float(2.0 ** int(params.scale_bits))
        self._keep = []
        self.ctx = self._make_ctx(host_scheme, self.n_limbs)
        m = 2 * self.N
        k = np.arange(m + 1, dtype=np.float64)
        ang = 2.0 * math.pi * k / m
        self.ksi = torch.from_numpy(np.stack([np.cos(ang), np.sin(ang)], axis=1).copy()).to(self.device)
        rot = np.empty(self.N // 2, dtype=np.uint32)
        five = 1
        for j in range(self.N // 2):
            rot[j] = five
            five = five * 5 % m
        self.rot = torch.from_numpy(rot.view(np.int32).copy()).to(self.device)
        self.pk = self.pk_shoup = self.sk = self.sk_shoup = None
        if host_scheme.has_public_key():
            self.pk, self.pk_shoup = self._key(host_scheme.public_key_words(), 2)
        if host_scheme.has_secret_key():
            self.sk, self.sk_shoup = self._key(host_scheme.secret_key_words(), 1)
        self._host = host_scheme

    def _make_ctx(self, host_scheme, n_limbs) -> _Ctx:
        ctx = _Ctx()
        ctx.log_n, ctx.n_limbs = self.log_n, n_limbs
        for l in range(n_limbs):
            t = host_scheme.ntt_tables(l)
            ctx.q[l], ctx.n_inv[l], ctx.n_inv_shoup[l] = int(t["q"]), int(t["n_inv"]), int(t["n_inv_shoup"])
            for name in ("psi", "psi_shoup", "ipsi", "ipsi_shoup"):
                dev = _u64(t[{"psi": "psi_br", "psi_shoup": "psi_br_shoup", "ipsi": "ipsi_br",
                              "ipsi_shoup": "ipsi_br_shoup"}[name]], self.device)
                self._keep.append(dev)
                getattr(ctx, name)[l] = dev.data_ptr()
        return ctx

    def _ctx_for(self, n_limbs) -> _Ctx:
        if n_limbs == self.n_limbs:
            return self.ctx
        ctx = _Ctx()
        C.memmove(C.byref(ctx), C.byref(self.ctx), C.sizeof(_Ctx))
        ctx.n_limbs = n_limbs
        return ctx

    def _key(self, words: np.ndarray, n_polys: int):
        w = np.ascontiguousarray(words).reshape(n_polys, self.n_limbs, self.N)
        sh = np.empty_like(w)
        for l, q in enumerate(self.moduli):
            # Shoup companions floor(w * 2^64 / q) with Python big ints (one-off, at key upload)
            flat = w[:, l, :].reshape(-1)
            sh[:, l, :] = np.array([(int(x) << 64) // q for x in flat], dtype=np.uint64).reshape(n_polys, self.N)
        return _u64(w.reshape(-1), self.device), _u64(sh.reshape(-1), self.device)

    # -- device-resident API ----------------------------------------------------------------------
    def n_ct(self, n_values: int) -> int:
        return (int(n_values) + self.slots - 1) // self.slots

    def ct_words(self, n_ct: int, n_limbs=None) -> int:
        return n_ct * 2 * (n_limbs or self.n_limbs) * self.N

    def encrypt_device(self, values: torch.Tensor, seed: int = 0, out: torch.Tensor = None) -> torch.Tensor:
        """float tensor (any shape, on this device) -> int64 tensor of ciphertext words."""
        assert self.pk is not None, "public key not loaded"
        v = values.detach().reshape(-1).to(self.device, torch.float64).contiguous()
        n = v.numel()
        n_ct = self.n_ct(n)
        coeffs = torch.empty(n_ct * 2 * self.slots, dtype=torch.float64, device=self.device)
        if out is None:
            out = torch.empty(self.ct_words(n_ct), dtype=torch.int64, device=self.device)
        st = _lib.stream_ptr()
        _lib.check(self.L.mfl_ckks_encode(C.byref(self.ctx), v.data_ptr(), n, self.slots, self.scale,
                                          self.ksi.data_ptr(), self.rot.data_ptr(), coeffs.data_ptr(), st),
                   "mfl_ckks_encode")
        # 256-bit ChaCha20 key per call: OS entropy, or (tests only) a fixed expansion of ``seed`` for reproducibility
        raw = os.urandom(32) if not seed else int(seed).to_bytes(32, "little", signed=False)
        key = (C.c_uint32 * 8).from_buffer_copy(raw)
        _lib.check(self.L.mfl_ckks_encrypt(C.byref(self.ctx), coeffs.data_ptr(), n_ct, self.slots, self.pk.data_ptr(),
                                           self.pk_shoup.data_ptr(), key, out.data_ptr(), st), "mfl_ckks_encrypt")
        return out

    def decrypt_device(self, ct_words: torch.Tensor, n_values: int, scale: float = None, n_limbs: int = None):
        assert self.sk is not None, "private key not loaded"
        n_limbs = n_limbs or self.n_limbs
        ctx = self._ctx_for(n_limbs)
        n_ct = self.n_ct(n_values)
        assert ct_words.numel() == self.ct_words(n_ct, n_limbs)
        st = _lib.stream_ptr()
        residues = torch.empty(n_ct * n_limbs * self.N, dtype=torch.int64, device=self.device)
        # the key is stored with the full limb count; rows of the first n_limbs limbs are a prefix
        _lib.check(self.L.mfl_ckks_decrypt(C.byref(ctx), ct_words.data_ptr(), n_ct, self.sk.data_ptr(),
                                           self.sk_shoup.data_ptr(), residues.data_ptr(), st), "mfl_ckks_decrypt")
        inv = (C.c_uint64 * 16)()
        for j in range(n_limbs):
            for i in range(j + 1, n_limbs):
                inv[j * 4 + i] = pow(self.moduli[j] % self.moduli[i], self.moduli[i] - 2, self.moduli[i])
        out = torch.empty(int(n_values), dtype=torch.float64, device=self.device)
        _lib.check(self.L.mfl_ckks_decode(C.byref(ctx), residues.data_ptr(), n_ct, self.slots,
                                          1.0 / (scale or self.scale), inv, self.ksi.data_ptr(), self.rot.data_ptr(),
                                          out.data_ptr(), int(n_values), st), "mfl_ckks_decode", launches=1)
        return out

    def weighted_sum_device(self, cts: Sequence[torch.Tensor], weights: Sequence[float], out: torch.Tensor = None,
                            n_limbs: int = None, scale: float = None):
        """Limb-wise ``sum_l ct_l * round(w_l * scale)``; the result's scale is ``scale**2``."""
        n_limbs = n_limbs or self.n_limbs
        ctx = self._ctx_for(n_limbs)
        words = cts[0].numel()
        if out is None:
            out = torch.empty_like(cts[0])
        w, wp = scalar_weights(weights, scale or self.scale, self.moduli[:n_limbs])
        _lib.check(self.L.mfl_ckks_weighted_sum(C.byref(ctx), _lib.voidp_array([c.data_ptr() for c in cts]), w, wp,
                                                len(cts), words, out.data_ptr(), _lib.stream_ptr()),
                   "mfl_ckks_weighted_sum")
        return out

    def fused_weighted_sum(self, src_ptrs: Sequence[int], weights: Sequence[float], words: int,
                           out_ptrs: Sequence[int], pad_ptrs: Sequence[int], rank: int, world: int,
                           n_blocks: int = 148, scale: float = None, mc_out: int = 0):
        """Every rank calls this with the same arguments: encrypted FedAvg over NVLink peer memory.  ``mc_out`` = the
        multicast address of the output buffers (``SymmetricBuffer.mc_ptr``): results go out as one ``multimem.st``."""
        w, wp = scalar_weights(weights, scale or self.scale, self.moduli)
        _lib.check(self.L.mfl_ckks_fused_weighted_sum_mc(
            C.byref(self.ctx), _lib.voidp_array(src_ptrs), w, wp, len(src_ptrs), int(words),
            _lib.voidp_array(out_ptrs), int(mc_out) or None, _lib.voidp_array(pad_ptrs), rank, world, n_blocks,
            _lib.stream_ptr()), "mfl_ckks_fused_weighted_sum_mc")

    def rescale_device(self, ct_words: torch.Tensor, n_ct: int, n_limbs: int = None):
        n_limbs = n_limbs or self.n_limbs
        ctx = self._ctx_for(n_limbs)
        ql = self.moduli[n_limbs - 1]
        qi = (C.c_uint64 * 4)()
        qis = (C.c_uint64 * 4)()
        for l in range(n_limbs - 1):
            q = self.moduli[l]
            qi[l] = pow(ql % q, q - 2, q)
            qis[l] = _shoup(qi[l], q)
        out = torch.empty(self.ct_words(n_ct, n_limbs - 1), dtype=torch.int64, device=self.device)
        _lib.check(self.L.mfl_ckks_rescale(C.byref(ctx), ct_words.data_ptr(), n_ct, qi, qis, out.data_ptr(),
                                           _lib.stream_ptr()), "mfl_ckks_rescale")
        return out

    def ntt_device(self, rows: torch.Tensor, inverse: bool = False):
        n_rows = rows.numel() // self.N
        _lib.check(self.L.mfl_ckks_ntt(C.byref(self.ctx), rows.data_ptr(), n_rows, int(inverse), _lib.stream_ptr()),
                   "mfl_ckks_ntt")
        return rows

    # -- byte-blob API (wire compatible with the host implementation) -----------------------------
    def encrypt(self, values: np.ndarray, seed: int = 0) -> bytes:
        v = torch.from_numpy(np.ascontiguousarray(values, dtype=np.float64)).to(self.device)
        words = self.encrypt_device(v, seed)
        n = v.numel()
        head = pack_header(self.log_n, self.n_limbs, self.n_ct(n), self.slots, n, self.scale, self.moduli)
        return head + words.cpu().numpy().tobytes()

    def decrypt(self, blob: bytes, n_values: int) -> np.ndarray:
        h = _core.ckks_header(blob)
        words = torch.frombuffer(bytearray(blob[HEADER_BYTES:]), dtype=torch.int64).to(self.device)
        # blobs may hold more ciphertexts than the requested prefix needs
        n_ct = self.n_ct(n_values)
        words = words[:self.ct_words(n_ct, h["n_limbs"])]
        return self.decrypt_device(words, n_values, scale=h["scale"], n_limbs=h["n_limbs"]).cpu().numpy()


def weighted_average(blobs: List[bytes], weights: Sequence[float]) -> bytes:
    """Byte-blob PWA on the GPU (what ``fhe.CKKS.compute_weighted_average`` calls for large tensors)."""
    h = _core.ckks_header(blobs[0])
    p = _core.CKKSParams()
    p.log_n, p.moduli, p.batch_size = h["log_n"], h["moduli"], h["slots"]
    p.scale_bits = max(1, int(round(math.log2(h["scale"]))))
    g = GpuCKKS(_core.CKKSScheme(p), p)
    dev = [torch.frombuffer(bytearray(b[HEADER_BYTES:]), dtype=torch.int64).to(g.device) for b in blobs]
    out = g.weighted_sum_device(dev, weights, n_limbs=h["n_limbs"], scale=h["scale"])
    head = pack_header(h["log_n"], h["n_limbs"], h["n_ct"], h["slots"], h["n_values"], h["scale"] * h["scale"],
                       h["moduli"])
    return head + out.cpu().numpy().tobytes()
