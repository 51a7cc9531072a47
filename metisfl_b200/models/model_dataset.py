"""Dataset wrappers handed to the learners by the user's recipe functions (API parity with
reference ``metisfl/models/model_dataset.py:4-71``).  ``x`` / ``y`` may be numpy arrays, torch
tensors, or ``x`` may be a ``torch.utils.data.Dataset`` / ``DataLoader`` yielding ``(x, y)``."""
import abc


class ModelDataset(object):

    def __init__(self, x=None, y=None, size=0):
        self._x = x
        self._y = y
        self._size = size

    def get_x(self, *args, **kwargs):
        return self._x

    def get_y(self, *args, **kwargs):
        return self._y

    def get_size(self, *args, **kwargs):
        return self._size

    @abc.abstractmethod
    def get_model_dataset_specifications(self, *args, **kwargs):
        return dict()


class ModelDatasetClassification(ModelDataset):

    def __init__(self, x=None, y=None, size=0, examples_per_class=None):
        super().__init__(x=x, y=y, size=size)
        if examples_per_class is not None and not isinstance(examples_per_class, dict):
            raise AssertionError("examples_per_class must be a dict {class_id: count}")
        self.examples_per_class = dict(examples_per_class or {})

    def get_model_dataset_specifications(self, *args, **kwargs):
        return self.examples_per_class


class ModelDatasetRegression(ModelDataset):

    def __init__(self, x=None, y=None, size=0, min_val=0.0, max_val=0.0, mean_val=0.0, median_val=0.0,
                 mode_val=0.0, stddev_val=0.0):
        super().__init__(x=x, y=y, size=size)
        self.min_val, self.max_val, self.mean_val = min_val, max_val, mean_val
        self.median_val, self.mode_val, self.stddev = median_val, mode_val, stddev_val

    def get_model_dataset_specifications(self, *args, **kwargs):
        return {"min": self.min_val, "max": self.max_val, "mean": self.mean_val, "median": self.median_val,
                "mode": self.mode_val, "stddev": self.stddev}
