"""Context manager that installs a received model once and then trains it (API parity with reference
``metisfl/learner/learner_trainer.py:4-31``)."""


class LearnerTrainer(object):

    def __init__(self, model_ops_fn, model_pb):
        self._model_ops = model_ops_fn() if callable(model_ops_fn) else model_ops_fn
        if model_pb is not None and len(model_pb.variables) > 0:
            if hasattr(self._model_ops, "set_model_weights_from_pb"):  # bulk device unpack (K12)
                self._model_ops.set_model_weights_from_pb(model_pb)
            else:
                names, trainables, values = self._model_ops.get_model_weights_from_variables_pb(model_pb.variables)
                self._model_ops.set_model_weights(names, trainables, values)

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc_val, exc_tb):
        self._model_ops.cleanup()

    def train_model(self, train_dataset, learning_task_pb, hyperparameters_pb, validation_dataset=None,
                    test_dataset=None, verbose=False, **kwargs):
        return self._model_ops.train_model(train_dataset, learning_task_pb, hyperparameters_pb, validation_dataset,
                                           test_dataset, verbose, **kwargs)
