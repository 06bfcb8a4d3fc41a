"""``import fedml`` compatibility alias: every name resolves to :mod:`fedml_b200`, and
``fedml.x.y`` sub-module imports are redirected to ``fedml_b200.x.y`` (with the reference's
``fedml.model`` -> ``fedml_b200.models`` rename), so scripts written for the reference run as is."""
import importlib
import importlib.abc
import importlib.util
import sys

import fedml_b200 as _impl
from fedml_b200 import *  # noqa: F401,F403
from fedml_b200 import __version__  # noqa: F401
import logging  # noqa: F401  (examples call fedml.logging.info)

_RENAMES = {"fedml.model": "fedml_b200.models"}


def _target(fullname):
    for old, new in _RENAMES.items():
        if fullname == old or fullname.startswith(old + "."):
            return new + fullname[len(old):]
    return "fedml_b200" + fullname[len("fedml"):]


class _AliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith("fedml."):
            return None
        try:
            real = importlib.import_module(_target(fullname))
        except ImportError:
            return None
        spec = importlib.util.spec_from_loader(fullname, self, is_package=hasattr(real, "__path__"))
        spec._fedml_real = real
        return spec

    def create_module(self, spec):
        return spec._fedml_real

    def exec_module(self, module):
        pass


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())


def __getattr__(name):
    return getattr(_impl, name)
