"""Imports the package directory `duckdb-cudasp-extension_b200/` (whose name is not a Python identifier)
under the module name `cudasp_b200`."""
import importlib.util
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))


def load():
    if "cudasp_b200" in sys.modules:
        return sys.modules["cudasp_b200"]
    path = os.path.join(_ROOT, "duckdb-cudasp-extension_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location("cudasp_b200", path, submodule_search_locations=[os.path.dirname(path)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["cudasp_b200"] = mod
    spec.loader.exec_module(mod)
    return mod
