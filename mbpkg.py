"""Import shim: the package directory is named `multibody.jl_b200` (not a valid
Python identifier), so it is loaded under the module name `multibody_jl_b200`."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.join(ROOT, "multibody.jl_b200")


def load():
    if "multibody_jl_b200" in sys.modules:
        return sys.modules["multibody_jl_b200"]
    spec = importlib.util.spec_from_file_location(
        "multibody_jl_b200", os.path.join(PKG_DIR, "__init__.py"),
        submodule_search_locations=[PKG_DIR])
    m = importlib.util.module_from_spec(spec)
    sys.modules["multibody_jl_b200"] = m
    spec.loader.exec_module(m)
    return m
