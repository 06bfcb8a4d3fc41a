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


def _compat_entry(fullname):
    """Reference module paths without a same-named module here (``fedml.simulation.single_process.fedavg.fedavg_api``,
    ``fedml.model.cv.resnet``, ...): fedml_b200/compat_table.py says which of the module's names exist in this package and where."""
    try:
        from fedml_b200.compat_table import TABLE
    except ImportError:
        return None
    return TABLE.get(fullname[len("fedml."):])


def _shim_module(fullname, entry):
    import types
    mod = types.ModuleType(fullname, f"Compatibility view of the reference module {fullname}: names resolve into fedml_b200.")
    if entry["pkg"]:
        mod.__path__ = []
    names = entry["names"]

    def __getattr__(name):
        target = names.get(name)
        if target is None:
            raise AttributeError(f"{fullname}.{name} has no counterpart in fedml_b200 (see docs/COMPONENTS.md)")
        tmod, _, attr = target.partition(":")
        value = getattr(importlib.import_module(tmod), attr)
        setattr(mod, name, value)
        return value

    mod.__getattr__ = __getattr__
    mod.__all__ = sorted(names)
    return mod


def _install_fallback(real, fullname, names):
    """Same-named module whose reference counterpart has names this one spells differently (``fedml.cli.cli.mlops_login``)."""
    def __getattr__(name):
        target = names.get(name)
        if target is None or target.partition(":")[0] == real.__name__ and target.partition(":")[2] == name:
            raise AttributeError(f"module {fullname!r} has no attribute {name!r}")
        tmod, _, attr = target.partition(":")
        return getattr(importlib.import_module(tmod), attr)

    real.__getattr__ = __getattr__


class _AliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith("fedml."):
            return None
        entry = _compat_entry(fullname)
        try:
            real = importlib.import_module(_target(fullname))
            if entry is not None and entry["pkg"] and not hasattr(real, "__path__"):
                real.__path__ = []              # a package in the reference, one module here: sub-module imports go through the table
            if entry is not None and entry["names"] and "__getattr__" not in vars(real):
                _install_fallback(real, fullname, entry["names"])
        except ImportError:
            if entry is None:
                return None
            real = _shim_module(fullname, entry)
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
