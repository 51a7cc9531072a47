"""Context manager that installs a received model once and then evaluates / infers with it (API
parity with reference ``metisfl/learner/learner_evaluator.py:4-35``)."""


class LearnerEvaluator(object):

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

    def evaluate_model(self, dataset, batch_size, metrics_pb, verbose=False):
        metrics = list(metrics_pb.metric) if metrics_pb is not None else None
        return self._model_ops.evaluate_model(dataset, batch_size, metrics, verbose)

    def infer_model(self, dataset, batch_size, verbose=False):
        return self._model_ops.infer_model(dataset, batch_size)
