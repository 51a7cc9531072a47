"""A learner that does not train: it sleeps, then echoes the received model back, and evaluates to nothing.

Counterpart of the reference's fake backend for load-testing the controller without any training
(``test/learner_notrain_noeval.py:16-198``: ``run_learning_task`` sleeps 3 s and returns the model through
``MarkTaskCompleted``; evaluation returns an empty ``ModelEvaluations``).  It is a drop-in ``Learner`` - same gRPC
servicer, same join / leave / task life cycle - so controller throughput (model receive, store, aggregate,
broadcast) can be measured with N of these and an arbitrarily large model.  ``python -m metisfl_b200.learner``
starts one when ``--neural_engine echo`` is given.
"""
import time

from metisfl_b200.learner.learner import Learner
from metisfl_b200.models.model_dataset import ModelDatasetClassification
from metisfl_b200.proto import metis_pb2


class EchoLearner(Learner):

    def __init__(self, *args, sleep_secs: float = 3.0, num_training_examples: int = 100, **kwargs):
        kwargs.setdefault("train_dataset_recipe_pkl", lambda: ModelDatasetClassification(size=num_training_examples))
        kwargs.setdefault("nn_engine", "echo")
        kwargs.setdefault("model_dir", "")
        kwargs.setdefault("train_dataset_fp", "")
        super().__init__(*args, **kwargs)
        self.sleep_secs = float(sleep_secs)
        self.num_training_examples = int(num_training_examples)
        self.tasks_echoed = 0

    def _load_model_datasets_size_specs_type_def(self):
        empty = (0, {}, ModelDatasetClassification)
        return [(self.num_training_examples, {}, ModelDatasetClassification), empty, empty]

    def model_train(self, learning_task_pb, hyperparameters_pb, model_pb, verbose=False, cancel_event=None):
        deadline = time.time() + self.sleep_secs
        while time.time() < deadline:
            if cancel_event is not None and cancel_event.is_set():
                return None  # pre-empted by a newer task, like a real training task
            time.sleep(min(0.05, max(0.0, deadline - time.time())))
        steps = max(1, int(learning_task_pb.num_local_updates))
        batch = int(hyperparameters_pb.batch_size) or 1
        meta = metis_pb2.TaskExecutionMetadata(
            global_iteration=learning_task_pb.global_iteration, completed_epochs=1.0, completed_batches=steps,
            batch_size=batch, processing_ms_per_epoch=self.sleep_secs * 1e3,
            processing_ms_per_batch=self.sleep_secs * 1e3 / steps)
        self.tasks_echoed += 1
        from metisfl_b200.proto.envelope import WireCompletedTask
        return WireCompletedTask(model_pb.SerializeToString(), meta)

    def model_evaluate(self, model_pb, batch_size, evaluation_datasets_pb, metrics_pb, verbose=False):
        return metis_pb2.ModelEvaluations()

    def model_infer(self, *args, **kwargs):
        return {}
