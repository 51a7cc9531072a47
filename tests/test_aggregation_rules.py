"""Golden expectations of the reference's aggregation gtests against the native core:
``federated_average_test.cc:95-208``, ``federated_stride_test.cc:86-439``,
``federated_recency_test.cc:80-307`` - expectations ported, not code."""
import numpy as np
import pytest

from metisfl_b200 import _core
from helpers import model_arrays, model_bytes, one_to_ten

TRUNC = [0, 2, 2, 4, 4, 6, 6, 8, 8, 10]


@pytest.mark.parametrize("dtype,expect", [(np.uint16, TRUNC), (np.int32, TRUNC),
                                          (np.float32, list(range(1, 11))), (np.float64, list(range(1, 11)))])
def test_fedavg_two_identical_models_half_weights(dtype, expect):
    m = model_bytes([one_to_ten(dtype)])
    out, contributors = _core.FedAvg().aggregate([([m], 0.5), ([m], 0.5)])
    assert contributors == 2
    got = model_arrays(out)[0]
    assert got.dtype == np.dtype(dtype) and got.tolist() == expect


def test_fedavg_keeps_variable_metadata_and_uses_first_lineage_entry():
    a = model_bytes([one_to_ten(np.float32)], names=["dense/kernel"])
    b = model_bytes([one_to_ten(np.float32) * 3], names=["dense/kernel"])
    out, _ = _core.FedAvg().aggregate([([a, b], 1.0)])  # lineage oldest->newest: FedAvg reads the front
    v = _core.parse_model(out)[0]
    assert v["name"] == "dense/kernel" and v["trainable"] and v["dimensions"] == [10]
    assert model_arrays(out)[0].tolist() == list(range(1, 11))


def _strided(models_weights, stride):
    agg = _core.FedStride()
    out = None
    for i in range(0, len(models_weights), stride):
        out = agg.aggregate([([m], w) for m, w in models_weights[i:i + stride]])
    agg.reset()
    return out


@pytest.mark.parametrize("stride", [1, 2, 3])
@pytest.mark.parametrize("n_learners", [1, 2, 4])
def test_fedstride_float32_equal_models(stride, n_learners):
    m = model_bytes([one_to_ten(np.float32)])
    out, c = _strided([(m, 1.0 / n_learners)] * n_learners, stride)
    assert c == n_learners
    np.testing.assert_allclose(model_arrays(out)[0], np.arange(1, 11), rtol=1e-6)


def test_fedstride_int32_truncation_matches_reference():
    m = model_bytes([one_to_ten(np.int32)])
    assert model_arrays(_strided([(m, 1.0)], 1)[0])[0].tolist() == list(range(1, 11))
    assert model_arrays(_strided([(m, 0.5), (m, 0.5)], 1)[0])[0].tolist() == TRUNC
    # four learners x 0.25: int(0.25*k) summed four times
    exp = [4 * int(0.25 * k) for k in range(1, 11)]
    assert model_arrays(_strided([(m, 0.25)] * 4, 1)[0])[0].tolist() == exp


@pytest.mark.parametrize("stride", [1, 3])
def test_fedstride_unequal_models(stride):
    base = one_to_ten(np.float32)
    ms = [model_bytes([base * k]) for k in (1, 2, 3, 4)]
    w = [0.1, 0.2, 0.3, 0.4]
    out, _ = _strided(list(zip(ms, w)), stride)
    np.testing.assert_allclose(model_arrays(out)[0], base * (0.1 + 0.4 + 0.9 + 1.6), rtol=1e-6)


def test_fedstride_reset_clears_state():
    m = model_bytes([one_to_ten(np.float32)])
    agg = _core.FedStride()
    agg.aggregate([([m], 1.0)])
    agg.reset()
    out, c = agg.aggregate([([model_bytes([one_to_ten(np.float32) * 2])], 1.0)])
    assert c == 1
    np.testing.assert_allclose(model_arrays(out)[0], 2 * np.arange(1, 11))


# ---- FedRec ---------------------------------------------------------------------------------
def test_fedrec_lineage_longer_than_two_gives_empty_model():
    m = model_bytes([one_to_ten(np.float32)])
    out, _ = _core.FedRec().aggregate([([m, m, m], 1.0)])
    assert out == b""


def test_fedrec_single_learner_one_then_two_models():
    m1 = model_bytes([one_to_ten(np.float32)])
    m2 = model_bytes([one_to_ten(np.float32) * 2])
    agg = _core.FedRec()
    out, c = agg.aggregate([([m1], [1.0])])
    np.testing.assert_allclose(model_arrays(out)[0], np.arange(1, 11))
    out, c = agg.aggregate([([m1, m2], [1.0, 1.0])])  # swap m1 for m2
    assert c == 1
    np.testing.assert_allclose(model_arrays(out)[0], 2 * np.arange(1, 11))


def test_fedrec_first_time_committers_weights_1_2_3():
    base = one_to_ten(np.float32)
    agg = _core.FedRec()
    out = None
    for k, w in ((1, 1.0), (2, 2.0), (3, 3.0)):
        out, c = agg.aggregate([([model_bytes([base * k])], [w])])
    assert c == 3
    np.testing.assert_allclose(model_arrays(out)[0], base * (1 * 1 + 2 * 2 + 3 * 3) / 6, rtol=1e-6)


def test_fedrec_interleaved_recommits_reference_sequence():
    """federated_recency_test.cc:238-307: final z = 6, community = {0.75, 1.5, ..., 7.5}."""
    base = one_to_ten(np.float32)
    m1, m2, m3 = model_bytes([base]), model_bytes([base]), model_bytes([base * 0.5])
    agg = _core.FedRec()
    seq = [([m1], [1.0]), ([m1], [2.0]), ([m1, m2], [1.0, 1.0]), ([m1, m2], [2.0, 2.0]), ([m1], [3.0]),
           ([m2, m3], [1.0, 1.0]), ([m2, m3], [2.0, 2.0])]
    out = None
    for lineage, w in seq:
        out, _ = agg.aggregate([(lineage, w)])
    np.testing.assert_allclose(model_arrays(out)[0], [0.75, 1.5, 2.25, 3, 3.75, 4.5, 5.25, 6, 6.75, 7.5], rtol=1e-6)


# ---- FedAsync (extension) -------------------------------------------------------------------
def test_fedasync_staleness_functions():
    S = _core.FedAsync.staleness
    assert S(0, 10, 0.5, 4) == 1.0
    assert abs(S(1, 3, 0.5, 4) - (1 + 3) ** -0.5) < 1e-12
    assert S(2, 3, 0.5, 4) == 1.0 and abs(S(2, 6, 0.5, 4) - 1 / (0.5 * 2 + 1)) < 1e-12


def test_fedasync_mixes_with_staleness_weight():
    base = one_to_ten(np.float32)
    agg = _core.FedAsync(alpha=0.5, staleness_fn=1, a=0.5, b=4)
    agg.set_community(model_bytes([base]))
    out, _ = agg.aggregate([([model_bytes([base * 3])], 1.0, 3.0)])  # staleness 3 -> s = 0.5 -> a_t = 0.25
    np.testing.assert_allclose(model_arrays(out)[0], base * (0.75 + 0.25 * 3), rtol=1e-6)


# ---- element semantics oracle --------------------------------------------------------------
@pytest.mark.parametrize("code,dtype", [(0, np.int8), (1, np.int16), (2, np.int32), (3, np.int64), (4, np.uint8),
                                        (5, np.uint16), (6, np.uint32), (7, np.uint64), (8, np.float32),
                                        (9, np.float64), (10, np.float16)])
def test_weighted_sum_bytes_all_dtypes(code, dtype):
    x = np.arange(1, 11).astype(dtype)
    out = np.frombuffer(_core.weighted_sum_bytes(code, [x.tobytes(), x.tobytes()], [0.5, 0.5]), dtype=dtype)
    exp = TRUNC if np.issubdtype(dtype, np.integer) else list(range(1, 11))
    assert out.tolist() == exp


def test_weighted_sum_bytes_bfloat16():
    import torch
    x = torch.arange(1, 11).to(torch.bfloat16)
    raw = x.view(torch.uint8).numpy().tobytes()
    out = torch.frombuffer(bytearray(_core.weighted_sum_bytes(11, [raw, raw], [0.5, 0.5])), dtype=torch.bfloat16)
    assert out.float().tolist() == list(range(1, 11))
