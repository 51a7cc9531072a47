"""Host-side batch iterator (`models/torch_model_ops._Batches`): the per-learner cache of pinned / device-ready dataset
copies used by the device service path, and small helpers of the service benchmark."""
import os
import sys

import numpy as np
import torch

from metisfl_b200.models.model_dataset import ModelDatasetClassification
from metisfl_b200.models.torch_model_ops import _Batches

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ds(n=10, shape=(3, 4, 5), seed=0):
    rng = np.random.default_rng(seed)
    x = rng.normal(size=(n,) + shape).astype(np.float32)
    y = np.arange(n)
    return ModelDatasetClassification(x=x, y=y, size=n, examples_per_class={}), x, y


def test_uncached_iteration_is_unchanged():
    ds, x, y = _ds()
    got = list(_Batches(ds, 4, shuffle=False, pin=False))
    assert [len(b[0]) for b in got] == [4, 4, 2]
    assert torch.equal(torch.cat([b[0] for b in got]), torch.from_numpy(x))
    assert torch.equal(torch.cat([b[1] for b in got]), torch.from_numpy(y))


def test_cache_keeps_one_converted_copy_in_device_format():
    ds, x, _ = _ds()
    cache = {}
    a = _Batches(ds, 4, shuffle=False, pin=False, cache=cache, x_dtype=torch.bfloat16, channels_last=True)
    b = _Batches(ds, 4, shuffle=False, pin=False, cache=cache, x_dtype=torch.bfloat16, channels_last=True)
    assert len(cache) == 1 and a.x is b.x                      # converted once, shared by later tasks
    xb, _ = next(iter(a))
    assert xb.shape == (4, 3, 4, 5) and xb.dtype == torch.bfloat16
    assert xb.is_contiguous(memory_format=torch.channels_last)  # an NCHW view of NHWC rows: no copy on the way to the GPU
    assert torch.equal(xb.float(), torch.from_numpy(x[:4]).to(torch.bfloat16).float())
    # another format is another entry; the fp32 NCHW path is untouched
    c = _Batches(ds, 4, shuffle=False, pin=False, cache=cache)
    assert len(cache) == 2 and c.x.dtype == torch.float32


def test_shuffled_batches_are_a_permutation_and_depend_on_the_seed():
    ds, x, _ = _ds(n=32)
    cache = {}
    it = _Batches(ds, 8, shuffle=True, pin=False, seed=3, cache=cache, x_dtype=torch.bfloat16, channels_last=True)
    pairs = list(it)
    xs, ys = torch.cat([p[0] for p in pairs]), torch.cat([p[1] for p in pairs])
    assert sorted(ys.tolist()) == list(range(32))
    assert torch.equal(xs.float(), torch.from_numpy(x).to(torch.bfloat16).float()[ys])   # rows travel with their labels
    it2 = _Batches(ds, 8, shuffle=True, pin=False, seed=4, cache=cache, x_dtype=torch.bfloat16, channels_last=True)
    ys2 = torch.cat([p[1] for p in it2])
    assert ys2.tolist() != ys.tolist()


def test_service_bench_timestamp_parser():
    sys.path.insert(0, os.path.join(ROOT, "bench"))
    import service_bench as sb
    assert sb._ts({"seconds": 10, "nanos": 500_000_000}) == 10.5
    assert sb._ts(12.25) == 12.25
    a = sb._ts("2026-11-07T06:49:20.554416896Z")
    b = sb._ts("2026-11-07T06:49:21.054416896Z")
    assert abs((b - a) - 0.5) < 1e-6
    assert sb._ts("1970-01-01T00:00:01Z") == 1.0
