"""sm_100a CKKS kernels (K5-K9) against the host RNS-CKKS oracle (``csrc/ckks/ckks_host.cc``):
cross-implementation encrypt/decrypt, bit-exact weighted reduction and rescale, NTT round trip."""
import numpy as np
import pytest
import torch

from metisfl_b200 import _core
from metisfl_b200.encryption import fhe

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(tmp_path_factory):
    from metisfl_b200.ops import ckks as gpu
    host = fhe.CKKS(4096, 52, backend="host")
    host.gen_crypto_context_and_keys(str(tmp_path_factory.mktemp("crypto")), seed=11)
    g = gpu.GpuCKKS(host._scheme, host._params, device="cuda:0")
    return host, g


def test_ntt_round_trip_and_linearity(ctx):
    host, g = ctx
    rng = np.random.default_rng(0)
    rows = np.stack([rng.integers(0, q, size=g.N, dtype=np.uint64) for q in g.moduli * 2])
    x = torch.from_numpy(rows.view(np.int64).copy()).cuda()
    y = g.ntt_device(x.clone())
    assert not torch.equal(x, y)
    back = g.ntt_device(y.clone(), inverse=True)
    assert torch.equal(back, x)


def test_gpu_encrypt_host_decrypt_and_back(ctx):
    host, g = ctx
    rng = np.random.default_rng(1)
    x = rng.normal(size=10000) * 5
    blob = g.encrypt(x, seed=123)
    h = _core.ckks_header(blob)
    assert (h["n_ct"], h["n_limbs"], h["slots"]) == (3, 3, 4096)
    np.testing.assert_allclose(host.decrypt(blob, x.size), x, atol=1e-7)       # GPU -> host
    np.testing.assert_allclose(g.decrypt(host.encrypt(x), x.size), x, atol=1e-7)  # host -> GPU
    np.testing.assert_allclose(g.decrypt(blob, x.size), x, atol=1e-7)          # GPU -> GPU
    assert g.encrypt(x, seed=123) == blob and g.encrypt(x, seed=124) != blob   # counter-based RNG


def test_weighted_sum_and_rescale_are_bit_exact(ctx):
    host, g = ctx
    rng = np.random.default_rng(2)
    xs = [rng.normal(size=6000) for _ in range(5)]
    w = [0.1, 0.2, 0.3, 0.15, 0.25]
    blobs = [host.encrypt(x, seed=50 + i) for i, x in enumerate(xs)]
    ref = _core.ckks_weighted_average(blobs, w)
    from metisfl_b200.ops import ckks as gpu
    got = gpu.weighted_average(blobs, w)
    assert got == ref
    np.testing.assert_allclose(g.decrypt(got, 6000), sum(wi * x for wi, x in zip(w, xs)), atol=1e-7)
    # rescale: same words as the host implementation, message preserved
    words = torch.frombuffer(bytearray(ref[64:]), dtype=torch.int64).cuda()
    r = g.rescale_device(words, n_ct=2)
    href = _core.ckks_rescale(ref)
    assert r.cpu().numpy().tobytes() == href[64:]
    hh = _core.ckks_header(href)
    out = g.decrypt_device(r, 6000, scale=hh["scale"], n_limbs=2).cpu().numpy()
    np.testing.assert_allclose(out, sum(wi * x for wi, x in zip(w, xs)), atol=1e-7)


def test_fused_weighted_sum_single_rank(ctx):
    """The NVLink kernel with world = 1 (barriers against itself) equals the local kernel."""
    host, g = ctx
    from metisfl_b200.ops import _lib
    rng = np.random.default_rng(3)
    cts = [g.encrypt_device(torch.from_numpy(rng.normal(size=9000)).cuda(), seed=7 + i) for i in range(3)]
    w = [0.5, 0.3, 0.2]
    local = g.weighted_sum_device(cts, w)
    out = torch.zeros_like(local)
    pads = torch.zeros(_lib.lib().mfl_signal_pad_bytes(148), dtype=torch.uint8, device="cuda")
    for _ in range(2):
        g.fused_weighted_sum([c.data_ptr() for c in cts], w, local.numel(), [out.data_ptr()], [pads.data_ptr()], 0, 1)
    torch.cuda.synchronize()
    assert torch.equal(out, local)


def test_eight_source_weighted_sum_matches_decrypted_average(ctx):
    """Eight ciphertext models (the 8-GPU shape: all sixteen loads of two element pairs in flight per thread): the local
    and the fused kernel agree bit for bit, and the decrypted result is the plaintext weighted average."""
    host, g = ctx
    from metisfl_b200.ops import _lib
    rng = np.random.default_rng(11)
    xs = [rng.normal(size=9000) for _ in range(8)]
    w = [0.05, 0.1, 0.15, 0.2, 0.2, 0.15, 0.1, 0.05]
    cts = [g.encrypt_device(torch.from_numpy(x).cuda()) for x in xs]   # fresh OS randomness per ciphertext
    local = g.weighted_sum_device(cts, w)
    out = torch.zeros_like(local)
    pads = torch.zeros(_lib.lib().mfl_signal_pad_bytes(148), dtype=torch.uint8, device="cuda")
    g.fused_weighted_sum([c.data_ptr() for c in cts], w, local.numel(), [out.data_ptr()], [pads.data_ptr()], 0, 1)
    torch.cuda.synchronize()
    assert torch.equal(out, local)
    dec = g.decrypt_device(out, 9000, scale=g.scale * g.scale).cpu().numpy()
    np.testing.assert_allclose(dec, sum(wi * x for wi, x in zip(w, xs)), atol=1e-7)


def test_fhe_class_uses_gpu_for_large_tensors(ctx):
    host, _ = ctx
    auto = fhe.CKKS(4096, 52, backend="cuda")
    f = host.get_crypto_params_files()
    auto.load_context_and_keys_from_files(f["crypto_context_file"], f["public_key_file"], f["private_key_file"])
    x = np.linspace(-2, 2, 50000)
    a, b = auto.encrypt(x), auto.encrypt(-x)
    avg = auto.compute_weighted_average([a, b], [0.75, 0.25])
    np.testing.assert_allclose(auto.decrypt(avg, x.size), 0.5 * x, atol=1e-7)


def test_engine_rule_pwa_round_with_key_files(tmp_path):
    """``EngineConfig(rule="PWA")``: keys from files, encrypt -> fused ciphertext reduction -> decrypt, device-timed
    phases; at world 1 the decrypted community model must equal the learner's own published model."""
    import torch
    from metisfl_b200.models.zoo import FashionMnistMLP
    from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
    from metisfl_b200.parallel.secure import SecureAggregator
    files = SecureAggregator.generate_keys(str(tmp_path / "keys"))
    assert set(files) >= {"crypto_context_file", "public_key_file", "private_key_file"}
    eng = FederationEngine(FashionMnistMLP, EngineConfig(rule="PWA", crypto_dir=str(tmp_path / "keys"),
                                                         use_cuda_graph=False), device="cuda:0")
    x = torch.randn(32, 1, 28, 28).to(torch.bfloat16).cuda()
    y = torch.randint(0, 10, (32,)).cuda()
    eng.secure.round([(x, y)] * 2, check=True)
    assert eng.secure.last_error < 1e-6, eng.secure.last_error
    out = eng.run_round([[(x, y)]])
    assert set(out["phase_ms"]) >= {"train", "encode_encrypt", "reduce_over_nvlink", "decrypt_decode"}
    assert all(v > 0 for v in out["phase_ms"].values())
    slot = eng.slot_views[0][0].float()
    torch.testing.assert_close(eng.arena.seg[0].float(), slot, atol=1e-2, rtol=1e-2)
