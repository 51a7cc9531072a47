"""``FabricSecure``: the three cuts of a secure round on the service path (encrypt at task end, reduction commanded by
the controller, decrypt into the community buffer) against the plaintext weighted average - CPU tier (host CKKS).
Reference expectation: ``private_weighted_average_test.cc:41-120`` (decrypt(PWA(encrypt)) equals the average)."""
import numpy as np
import pytest
import torch

from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
from metisfl_b200.parallel.secure import FabricSecure, SecureAggregator


class _BNNet(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.fc = torch.nn.Linear(12, 8)
        self.bn = torch.nn.BatchNorm1d(8)
        self.out = torch.nn.Linear(8, 3)

    def forward(self, x):
        return self.out(torch.relu(self.bn(self.fc(x))))


@pytest.fixture(scope="module")
def keys(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("keys"))
    return SecureAggregator.generate_keys(d, 4096, 52)


def _engine(dtype=torch.float32):
    cfg = EngineConfig(learners_per_gpu=1, model_dtype=dtype, use_cuda_graph=False, direct_install=False)
    return FederationEngine(_BNNet, cfg, device="cpu")


def test_encrypt_reduce_decrypt_equals_the_weighted_model(keys):
    eng = _engine()
    sec = FabricSecure(eng, keys["crypto_context_file"], keys["public_key_file"], keys["private_key_file"])
    floats = [i for i, s in enumerate(eng.layout.segments) if not s.exact and s.dtype.is_floating_point]
    ints = [i for i, s in enumerate(eng.layout.segments) if s.exact]
    assert sec.segments == floats and len(floats) >= 2 and ints   # parameters, BatchNorm statistics; the step counter
    g = torch.Generator().manual_seed(1)
    for i in floats:
        eng.slot_views[0][i].copy_(torch.randn(eng.slot_views[0][i].shape, generator=g))
    before_int = [eng.community_views[i].clone() for i in ints]
    sec.encrypt_slot(0)
    it = eng.global_iteration
    sec.reduce([0.25])
    assert eng.global_iteration == it + 1
    sec.decrypt_install()
    for i in floats:
        np.testing.assert_allclose(eng.community_views[i].numpy(), 0.25 * eng.slot_views[0][i].numpy(), atol=1e-6)
    for i, b in zip(ints, before_int):   # integer buffers are not part of the encrypted payload
        assert torch.equal(eng.community_views[i], b)
    # an encrypted snapshot restores the same community model
    blob = sec.community_blob()
    want = [eng.community_views[i].clone() for i in floats]
    for i in floats:
        eng.community_views[i].zero_()
    sec.install_blob(blob)
    for i, w in zip(floats, want):
        np.testing.assert_allclose(eng.community_views[i].numpy(), w.numpy(), atol=1e-6)


def test_reduce_checks_its_arguments(keys):
    eng = _engine()
    sec = FabricSecure(eng, keys["crypto_context_file"], keys["public_key_file"], keys["private_key_file"])
    with pytest.raises(ValueError):
        sec.reduce([0.5, 0.5])            # one weight per fabric rank
    with pytest.raises(RuntimeError):
        sec.reduce([1.0])                 # nothing was published yet


def test_bfloat16_arena_round_trips_through_the_ciphertext(keys):
    """The GPU learners keep their models in bf16: the encoder takes float64, the decrypted average is cast back."""
    eng = _engine(torch.bfloat16)
    sec = FabricSecure(eng, keys["crypto_context_file"], keys["public_key_file"], keys["private_key_file"])
    g = torch.Generator().manual_seed(2)
    for i in sec.segments:
        v = eng.slot_views[0][i]
        v.copy_(torch.randn(v.shape, generator=g).to(v.dtype))
    sec.encrypt_slot(0)
    sec.reduce([1.0])
    sec.decrypt_install()
    for i in sec.segments:   # weight 1: the same bf16 values come back exactly (error 1e-10 << half a bf16 ulp)
        assert eng.community_views[i].dtype == eng.slot_views[0][i].dtype
        assert torch.equal(eng.community_views[i], eng.slot_views[0][i])
