"""The package directory is named `hip_llama.cpp_b200` (not a Python identifier);
load it, or one of its modules, by path."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "hip_llama.cpp_b200")
PKG_NAME = "hip_llama_cpp_b200"


def load_pkg(sub=None):
    if PKG_NAME not in sys.modules:
        spec = importlib.util.spec_from_file_location(
            PKG_NAME, os.path.join(PKG_DIR, "__init__.py"), submodule_search_locations=[PKG_DIR])
        mod = importlib.util.module_from_spec(spec)
        sys.modules[PKG_NAME] = mod
        spec.loader.exec_module(mod)
    if sub is None:
        return sys.modules[PKG_NAME]
    return importlib.import_module(f"{PKG_NAME}.{sub}")
