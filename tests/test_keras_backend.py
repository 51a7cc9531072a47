"""KerasModelOps against a Keras-shaped model (TensorFlow itself is not in the image): the backend only uses the
``fit / evaluate / predict / get_weights / set_weights / weights / optimizer`` surface, so a small numpy logistic
regression with that surface exercises weight exchange, the step budget, optimizer reconfiguration and the metadata
(reference ``models/keras/keras_model_ops.py:117-283``, ``callbacks/step_counter.py:40-44`` incl. quirk Q11)."""
import math
from types import SimpleNamespace

import numpy as np

from metisfl_b200.models.keras.keras_model_ops import KerasModelOps
from metisfl_b200.models.model_dataset import ModelDatasetClassification
from metisfl_b200.proto import metis_pb2, model_pb2


class _Var:
    def __init__(self, name, trainable=True):
        self.name, self.trainable = name, trainable


class FakeKerasModel:
    """Softmax regression trained by plain SGD, shaped like ``tf.keras.Model``."""

    def __init__(self, d=8, classes=3):
        rng = np.random.default_rng(0)
        self.w = rng.normal(size=(d, classes)).astype(np.float32) * 0.1
        self.b = np.zeros(classes, dtype=np.float32)
        self.seen = np.zeros(1, dtype=np.int64)   # a non-trainable counter, like BatchNorm statistics
        self.weights = [_Var("dense/kernel:0"), _Var("dense/bias:0"), _Var("seen:0", trainable=False)]
        self.optimizer = SimpleNamespace(learning_rate=0.5, momentum=0.0)
        self.stop_training = False
        self.batches = 0

    def get_weights(self):
        return [self.w.copy(), self.b.copy(), self.seen.copy()]

    def set_weights(self, ws):
        self.w, self.b, self.seen = [np.array(a) for a in ws]

    def _probs(self, x):
        z = x @ self.w + self.b
        z -= z.max(1, keepdims=True)
        e = np.exp(z)
        return e / e.sum(1, keepdims=True)

    def fit(self, x, y, epochs=1, batch_size=32, verbose=0, callbacks=(), validation_data=None):
        hist = {"loss": []}
        for cb in callbacks:
            cb.set_model(self)
            cb.on_train_begin()
        for ep in range(epochs):
            for cb in callbacks:
                cb.on_epoch_begin(ep)
            for i in range(0, len(x), batch_size):
                if self.stop_training:
                    break
                for cb in callbacks:
                    cb.on_train_batch_begin(i // batch_size)
                xb, yb = x[i:i + batch_size], y[i:i + batch_size]
                p = self._probs(xb)
                p[np.arange(len(yb)), yb] -= 1
                self.w -= self.optimizer.learning_rate * xb.T @ p / len(yb)
                self.b -= self.optimizer.learning_rate * p.mean(0)
                self.seen += len(yb)
                self.batches += 1
                for cb in callbacks:
                    cb.on_train_batch_end(i // batch_size)
            hist["loss"].append(self.evaluate(x, y)["loss"])
            for cb in callbacks:
                cb.on_epoch_end(ep)
            if self.stop_training:
                break
        for cb in callbacks:
            cb.on_train_end()
        return SimpleNamespace(history=hist)

    def evaluate(self, x, y, batch_size=100, verbose=0, return_dict=True):
        p = self._probs(x)
        return {"loss": float(-np.log(p[np.arange(len(y)), y] + 1e-9).mean()),
                "accuracy": float((p.argmax(1) == y).mean())}

    def predict(self, x, batch_size=100, verbose=0):
        return self._probs(x)


def _dataset(n=96, seed=1):
    rng = np.random.default_rng(seed)
    y = rng.integers(0, 3, size=n)
    x = rng.normal(size=(n, 8)).astype(np.float32) + y[:, None]
    return ModelDatasetClassification(x=x, y=y, size=n)


def test_keras_backend_trains_within_the_step_budget_and_reports_metadata():
    ops = KerasModelOps(model=FakeKerasModel())
    names, trainables, values = ops.get_model_weights()
    assert names == ["dense/kernel:0", "dense/bias:0", "seen:0"] and trainables == [True, True, False]
    task = metis_pb2.LearningTask(global_iteration=3, num_local_updates=7)
    hyper = metis_pb2.Hyperparameters(batch_size=16, optimizer=model_pb2.OptimizerConfig(
        momentum_sgd=model_pb2.MomentumSGD(learning_rate=0.25, momentum_factor=0.9)))
    done = ops.train_model(_dataset(), task, hyper, None, _dataset(32, 2))
    m = ops._model
    assert m.batches == 7                               # exactly the budget (the reference runs total_steps + 1)
    assert m.optimizer.learning_rate == 0.25 and abs(m.optimizer.momentum - 0.9) < 1e-6
    md = done.execution_metadata
    assert md.completed_batches == 7 and md.batch_size == 16 and md.global_iteration == 3
    assert math.isclose(md.completed_epochs, 7 / 6, rel_tol=1e-6) and md.processing_ms_per_batch > 0
    test_ev = md.task_evaluation.test_evaluation[0].model_evaluation.metric_values
    assert "accuracy" in test_ev and "loss" in test_ev
    assert [v.name for v in done.model.variables] == names and not done.model.variables[2].trainable
    # the trained weights travel back through the wire format and can be installed again
    n2, t2, v2 = ops.get_model_weights_from_variables_pb(done.model.variables)
    fresh = KerasModelOps(model=FakeKerasModel())
    fresh.set_model_weights(n2, t2, v2)
    np.testing.assert_allclose(fresh._model.w, m.w, rtol=1e-6)
    assert int(fresh._model.seen[0]) == int(m.seen[0]) == 7 * 16
    ev = fresh.evaluate_model(_dataset(), 32, ["accuracy"])
    assert ev["accuracy"] > 0.5 and fresh.infer_model(_dataset(8, 5)).shape == (8, 3)


def test_optimizer_configs_map_onto_the_compiled_optimizer():
    """``construct_optimizer`` overwrites only what the federation configures (keras_model_ops.py:245-283): proto3
    zero values of optional hyper-parameters keep what the model was compiled with; the learning rate is always set."""
    m = FakeKerasModel()
    m.optimizer = SimpleNamespace(learning_rate=0.5, momentum=0.3, mu=0.01, beta_1=0.8, beta_2=0.9, epsilon=1e-5,
                                  weight_decay=0.0)
    ops = KerasModelOps(model=m)
    ops.construct_optimizer(model_pb2.OptimizerConfig(vanilla_sgd=model_pb2.VanillaSGD(learning_rate=0.1)))
    close = lambda a, b: abs(a - b) < 1e-6 * max(1.0, abs(b))   # proto floats are fp32
    assert close(m.optimizer.learning_rate, 0.1) and m.optimizer.momentum == 0.3
    ops.construct_optimizer(model_pb2.OptimizerConfig(fed_prox=model_pb2.FedProx(learning_rate=0.2, proximal_term=0.5)))
    assert close(m.optimizer.learning_rate, 0.2) and m.optimizer.mu == 0.5
    ops.construct_optimizer(model_pb2.OptimizerConfig(adam=model_pb2.Adam(learning_rate=0.001, beta_1=0.95)))
    assert close(m.optimizer.learning_rate, 0.001) and abs(m.optimizer.beta_1 - 0.95) < 1e-6
    assert m.optimizer.beta_2 == 0.9 and m.optimizer.epsilon == 1e-5            # unset in the message: untouched
    ops.construct_optimizer(model_pb2.OptimizerConfig(
        adam_weight_decay=model_pb2.AdamWeightDecay(learning_rate=0.003, weight_decay=0.02)))
    assert close(m.optimizer.learning_rate, 0.003) and abs(m.optimizer.weight_decay - 0.02) < 1e-7
    assert ops.construct_optimizer(None) is m.optimizer                          # nothing configured: nothing changes


def test_set_weights_casts_to_the_models_dtypes_and_rejects_a_different_architecture():
    import pytest
    ops = KerasModelOps(model=FakeKerasModel())
    names, trainables, values = ops.get_model_weights()
    as64 = [v.astype(np.float64) if v.dtype.kind == "f" else v for v in values]   # e.g. an fp64 community model
    as64[0] = as64[0] + 1.0
    ops.set_model_weights(names, trainables, as64)
    assert ops._model.w.dtype == np.float32 and np.allclose(ops._model.w, values[0] + 1.0)
    with pytest.raises(ValueError):
        ops.set_model_weights(names[:2], trainables[:2], values[:2])


def test_empty_splits_and_missing_training_data():
    import pytest
    ops = KerasModelOps(model=FakeKerasModel())
    assert ops.evaluate_model(None, 32) == {}
    task = metis_pb2.LearningTask(global_iteration=1, num_local_updates=2)
    hyper = metis_pb2.Hyperparameters(batch_size=16, optimizer=model_pb2.OptimizerConfig(
        vanilla_sgd=model_pb2.VanillaSGD(learning_rate=0.1)))
    with pytest.raises(RuntimeError):
        ops.train_model(None, task, hyper)
    done = ops.train_model(_dataset(), task, hyper)          # no validation / test split: metadata still complete
    assert done.execution_metadata.completed_batches == 2
    with pytest.raises(RuntimeError):
        ops.infer_model(None)
