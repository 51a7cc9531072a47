"""Import-path compatibility: ``import metisfl.<x>`` resolves to ``metisfl_b200.<x>``.

Code written against MetisFL (``metisfl.driver.driver_session.DriverSession``,
``metisfl.utils.fedenv_parser.FederationEnvironment``, ``metisfl.models.model_dataset.*``,
``metisfl.proto.*_pb2``, ``metisfl.encryption.fhe.CKKS``, ``python -m metisfl.controller`` ...) keeps
working unchanged; both names refer to the SAME module objects.
"""
import importlib
import importlib.abc
import importlib.util
import sys

import metisfl_b200

_PREFIX, _TARGET = "metisfl", "metisfl_b200"


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, target):
        self._target = target

    def create_module(self, spec):
        return importlib.import_module(self._target)

    def exec_module(self, module):  # already executed under its real name
        pass

    def get_code(self, fullname):  # `python -m metisfl.controller` / `metisfl.learner`
        spec = importlib.util.find_spec(self._target)
        return spec.loader.get_code(self._target)

    def get_source(self, fullname):
        spec = importlib.util.find_spec(self._target)
        return spec.loader.get_source(self._target)

    def is_package(self, fullname):
        return importlib.util.find_spec(self._target).submodule_search_locations is not None


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(_PREFIX + "."):
            return None
        real = _TARGET + fullname[len(_PREFIX):]
        try:
            real_spec = importlib.util.find_spec(real)
        except (ImportError, ValueError):
            return None
        if real_spec is None:
            return None
        spec = importlib.util.spec_from_loader(fullname, _AliasLoader(real),
                                               is_package=real_spec.submodule_search_locations is not None)
        return spec


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())

__version__ = metisfl_b200.__version__
__path__ = list(getattr(metisfl_b200, "__path__", []))
