"""Every model family of the zoo trains through the framework's own ops (TorchModelOps -> arena -> fused optimizer)
and round-trips through the protobuf model format.  Reference zoo: ``examples/keras/models/*.py``,
``examples/pytorch/models/mlp.py``."""
import numpy as np
import pytest
import torch

from metisfl_b200.models import zoo
from metisfl_b200.models.model_dataset import ModelDatasetClassification, ModelDatasetRegression
from metisfl_b200.models.torch_model_ops import TorchModelOps
from metisfl_b200.proto import metis_pb2, model_pb2


def _task(steps):
    return metis_pb2.LearningTask(global_iteration=1, num_local_updates=steps)


def _hyper(batch, lr=0.01, kind="vanilla_sgd"):
    opt = model_pb2.OptimizerConfig()
    if kind == "adam":
        opt.adam.learning_rate = lr
    else:
        opt.vanilla_sgd.learning_rate = lr
    return metis_pb2.Hyperparameters(batch_size=batch, optimizer=opt)


CASES = {
    "cifar_cnn": (lambda: zoo.CifarCNN(), lambda n: (np.random.randn(n, 3, 32, 32).astype("f4"), np.random.randint(0, 10, n)), "cls"),
    "mnist_fc": (lambda: zoo.MnistFC(), lambda n: (np.random.randn(n, 28, 28).astype("f4"), np.random.randint(0, 10, n)), "cls"),
    "fashion_fc": (lambda: zoo.FashionMnistFC(), lambda n: (np.random.randn(n, 28, 28).astype("f4"), np.random.randint(0, 10, n)), "cls"),
    "housing": (lambda: zoo.HousingMLP(13, 10, 2), lambda n: (np.random.randn(n, 13).astype("f4"), np.random.randn(n).astype("f4")), "reg"),
    "imdb_lstm": (lambda: zoo.ImdbLSTM(500), lambda n: (np.random.randint(0, 500, (n, 12)), np.random.randint(0, 2, n).astype("f4")), "cls"),
    "melanoma": (lambda: zoo.MelanomaFC((32, 32), 8), lambda n: (np.random.randn(n, 3, 32, 32).astype("f4"), np.random.randint(0, 2, n).astype("f4")), "cls"),
    "brainage2d": (lambda: zoo.BrainAge2DCNN((4, 8, 8, 8, 8)), lambda n: (np.random.randn(n, 32, 32, 3).astype("f4"), 60 + np.random.randn(n).astype("f4")), "reg"),
    "brainage3d": (lambda: zoo.BrainAge3DCNN((4, 4, 8, 8, 8)), lambda n: (np.random.randn(n, 32, 32, 32).astype("f4"), 60 + np.random.randn(n).astype("f4")), "reg"),
    "alz2d": (lambda: zoo.AlzheimersDisease2DCNN((4, 8, 8, 8, 8)), lambda n: (np.random.randn(n, 3, 32, 32).astype("f4"), np.random.randint(0, 2, n).astype("f4")), "cls"),
    "alz3d": (lambda: zoo.AlzheimersDisease3DCNN((4, 4, 8, 8)), lambda n: (np.random.randn(n, 46, 46, 46).astype("f4"), np.random.randint(0, 2, n).astype("f4")), "cls"),
}


def test_cifar_cnn_parameter_count_matches_reference():
    m = zoo.CifarCNN()
    trainable = sum(p.numel() for p in m.parameters())
    moving = sum(b.numel() for n, b in m.named_buffers() if "running" in n)
    assert trainable + moving == 1610314  # the "1.6M" model of controller.cc:602


def test_fashion_mlp_parameter_count_matches_reference():
    assert sum(p.numel() for p in zoo.FashionMnistFC().parameters()) == 118282


@pytest.mark.parametrize("name", sorted(CASES))
def test_family_trains_and_round_trips(name):
    build, data, kind = CASES[name]
    np.random.seed(0); torch.manual_seed(0)
    x, y = data(8)
    ds = (ModelDatasetRegression if kind == "reg" else ModelDatasetClassification)(x=x, y=y, size=len(x))
    ops = TorchModelOps(model=build(), device="cpu")
    before = ops.get_model_pb()
    done = ops.train_model(ds, _task(2), _hyper(4, 1e-3))
    assert done.execution_metadata.completed_batches == 2
    after = ops.get_model_pb()
    assert len(before.variables) == len(after.variables) > 0
    changed = sum(a.plaintext_tensor.tensor_spec.value != b.plaintext_tensor.tensor_spec.value
                  for a, b in zip(before.variables, after.variables))
    assert changed > 0
    # install into a fresh instance and compare every tensor
    other = TorchModelOps(model=build(), device="cpu")
    other.set_model_weights_from_pb(after)
    again = other.get_model_pb()
    for a, b in zip(after.variables, again.variables):
        assert a.name == b.name and a.plaintext_tensor.tensor_spec.value == b.plaintext_tensor.tensor_spec.value
    res = ops.evaluate_model(ds, batch_size=4)
    assert np.isfinite(res["loss"])


def test_callbacks_step_counter_stops_exactly_and_profiler_records():
    from metisfl_b200.models.callbacks import PerformanceProfiler, StepCounter
    np.random.seed(0)
    x, y = np.random.randn(40, 28, 28).astype("f4"), np.random.randint(0, 10, 40)
    ds = ModelDatasetClassification(x=x, y=y, size=40)
    ops = TorchModelOps(model=zoo.MnistFC(), device="cpu")
    counter, prof = StepCounter(3), PerformanceProfiler("cpu")
    done = ops.train_model(ds, _task(100), _hyper(4), callbacks=[counter, prof])
    assert counter.steps_counter == 3 and done.execution_metadata.completed_batches == 3  # not total_steps + 1 (Q11)
    assert len(prof.batches_wall_clock_time_sec) == 3 and len(prof.epochs_wall_clock_time_sec) == 1


def test_fedprox_optimizer_matches_closed_form():
    from metisfl_b200.optim import FedProx
    torch.manual_seed(0)
    w = torch.nn.Parameter(torch.randn(17))
    opt = FedProx([w], learning_rate=0.1, proximal_term=0.5)
    w0 = w.detach().clone()
    ref = w0.clone()
    for _ in range(3):
        opt.zero_grad()
        (w ** 2).sum().backward()
        opt.step()
        ref = ref - 0.1 * (2 * ref + 0.5 * (ref - w0))
    torch.testing.assert_close(w.detach(), ref)
