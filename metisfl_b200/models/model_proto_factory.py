"""Builders for the task-result messages a learner returns (API parity with reference
``metisfl/models/model_proto_factory.py:9-123``)."""
from metisfl_b200.utils.formatting import DictionaryFormatter
from metisfl_b200.utils.proto_messages_factory import MetisProtoMessages, ModelProtoMessages


class ModelProtoFactory:

    class CompletedLearningTaskProtoMessage(object):

        def __init__(self, weights_values, weights_trainable, weights_names, train_stats=None, completed_epochs=0,
                     global_iteration=0, validation_stats=None, test_stats=None, completes_batches=0,
                     batch_size=0, processing_ms_per_epoch=0.0, processing_ms_per_batch=0.0):
            """``*_stats`` are ``{metric: [value per epoch]}`` dictionaries."""
            self.weights_values, self.weights_trainable, self.weights_names = \
                weights_values, weights_trainable, weights_names
            self.train_stats = train_stats or {}
            self.validation_stats = validation_stats or {}
            self.test_stats = test_stats or {}
            self.completed_epochs = completed_epochs
            self.global_iteration = global_iteration
            self.completes_batches = completes_batches
            self.batch_size = batch_size
            self.processing_ms_per_epoch = processing_ms_per_epoch
            self.processing_ms_per_batch = processing_ms_per_batch

        def _construct_task_evaluation_pb(self, collection):
            """``{metric: [v_epoch1, v_epoch2, ...]}`` -> ``[EpochEvaluation]``."""
            listified = DictionaryFormatter.listify_values(collection)
            n_epochs = max((len(v) for v in listified.values()), default=0)
            out = []
            for e in range(n_epochs):
                metrics = {k: v[e] for k, v in listified.items() if e < len(v)}
                out.append(MetisProtoMessages.construct_epoch_evaluation_pb(
                    epoch_id=e + 1,
                    model_evaluation_pb=MetisProtoMessages.construct_model_evaluation_pb(
                        DictionaryFormatter.stringify(metrics))))
            return out

        def construct_task_execution_metadata_pb(self):
            task_evaluation_pb = MetisProtoMessages.construct_task_evaluation_pb(
                epoch_training_evaluations_pbs=self._construct_task_evaluation_pb(self.train_stats),
                epoch_validation_evaluations_pbs=self._construct_task_evaluation_pb(self.validation_stats),
                epoch_test_evaluations_pbs=self._construct_task_evaluation_pb(self.test_stats))
            return MetisProtoMessages.construct_task_execution_metadata_pb(
                global_iteration=self.global_iteration, task_evaluation_pb=task_evaluation_pb,
                completed_epochs=self.completed_epochs, completed_batches=self.completes_batches,
                batch_size=self.batch_size, processing_ms_per_epoch=self.processing_ms_per_epoch,
                processing_ms_per_batch=self.processing_ms_per_batch)

        def construct_completed_learning_task_pb(self, aux_metadata="", he_scheme=None):
            model_pb = ModelProtoMessages.construct_model_pb_from_np(
                self.weights_values, self.weights_names, self.weights_trainable, he_scheme=he_scheme)
            return MetisProtoMessages.construct_completed_learning_task_pb(
                model_pb=model_pb, task_execution_metadata_pb=self.construct_task_execution_metadata_pb(),
                aux_metadata=aux_metadata)

        def construct_completed_learning_task_wire(self, aux_metadata="", he_scheme=None):
            """Same task, with the model serialized straight from the tensors (``proto/envelope.py``)."""
            from metisfl_b200.proto.envelope import WireCompletedTask
            model_bytes = ModelProtoMessages.construct_model_bytes_from_np(
                self.weights_values, self.weights_names, self.weights_trainable, he_scheme=he_scheme)
            return WireCompletedTask(model_bytes, self.construct_task_execution_metadata_pb(), aux_metadata)

    class ModelEvaluationProtoMessage(object):

        def __init__(self, metric_values):
            self.metric_values = metric_values

        def construct_model_evaluation_pb(self):
            return MetisProtoMessages.construct_model_evaluation_pb(
                DictionaryFormatter.stringify(self.metric_values or {}))
