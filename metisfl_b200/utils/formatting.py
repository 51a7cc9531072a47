"""Metric dictionary helpers (API parity with reference ``metisfl/utils/formatting.py:7-45``):
model metrics travel as ``map<string,string>`` (``metis.proto:162-169``), NaNs included."""
import math

import numpy as np


class DictionaryFormatter(object):

    @classmethod
    def normalize(cls, d):
        """Flatten ``{"loss": [0.1]}`` style values to plain scalars / lists."""
        out = {}
        for k, v in d.items():
            if isinstance(v, np.ndarray):
                v = v.tolist()
            if isinstance(v, (list, tuple)) and len(v) == 1:
                v = v[0]
            out[k] = v
        return out

    @classmethod
    def _stringify_val(cls, val, stringify_nan=True):
        if isinstance(val, (float, np.floating)) and (math.isnan(val) or math.isinf(val)):
            return "NaN" if stringify_nan else str(val)
        if isinstance(val, (np.generic,)):
            val = val.item()
        return str(val)

    @classmethod
    def stringify(cls, d, stringify_nan=True):
        return {str(k): cls._stringify_val(v, stringify_nan) for k, v in d.items()}

    @classmethod
    def listify_values(cls, d):
        return {k: (v if isinstance(v, list) else [v]) for k, v in d.items()}
