"""Importable alias of the package directory `distributeddataparallel-cifar10_b200/` (the hyphen in
the mandated directory name is not a legal Python identifier): `import ddp_cifar10_b200 as nrd`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "distributeddataparallel-cifar10_b200")
_spec = importlib.util.spec_from_file_location(__name__, os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
