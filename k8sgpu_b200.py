"""Import shim: the package directory is `k8s-gpu-_b200/` (not a Python identifier), so this
module loads it and installs it in sys.modules under the name `k8sgpu_b200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "k8s-gpu-_b200")
_spec = importlib.util.spec_from_file_location(
    "k8sgpu_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["k8sgpu_b200"] = _mod
_spec.loader.exec_module(_mod)
