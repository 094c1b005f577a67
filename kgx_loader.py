"""Import helper: the package directory `kubernetes-gpu-exporter_b200/` is not a
valid Python identifier, so it is loaded under the module name `kgx_b200`."""
import importlib.util
import os
import sys

_NAME = "kgx_b200"
_ROOT = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.join(_ROOT, "kubernetes-gpu-exporter_b200")


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    spec = importlib.util.spec_from_file_location(
        _NAME, os.path.join(_PKG, "__init__.py"), submodule_search_locations=[_PKG])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod


def load_oracle():
    """The CPU oracle binding (tests / smoke / bench cpu_baseline only)."""
    name = "kgo_oracle"
    if name in sys.modules:
        return sys.modules[name]
    load()
    spec = importlib.util.spec_from_file_location(name, os.path.join(_ROOT, "oracle", "kgo.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod
