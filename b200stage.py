"""Alias: ``import b200stage`` (and ``b200stage.<submodule>``) resolves to the package in
``llm-for-distributed-egde-devices_b200/``, whose directory name is not a valid Python identifier.
Every alias is the SAME module object as the real one (no second copy of any class)."""
import importlib
import importlib.abc
import importlib.util
import os
import sys

_REAL = "llm-for-distributed-egde-devices_b200"
_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)


class _AliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname.startswith(__name__ + "."):
            return importlib.util.spec_from_loader(fullname, self)
        return None

    def create_module(self, spec):
        return importlib.import_module(_REAL + spec.name[len(__name__):])

    def exec_module(self, module):
        pass


sys.meta_path.insert(0, _AliasFinder())
_pkg = importlib.import_module(_REAL)
sys.modules[__name__] = _pkg
