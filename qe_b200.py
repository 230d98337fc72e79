"""Import alias: the package directory is `q-e_b200/` (the reference's repo name
plus `_b200`), which is not a valid Python identifier. `import qe_b200` loads
that directory as the package `qe_b200`."""
import importlib.util
import os
import sys

_d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "q-e_b200")
_spec = importlib.util.spec_from_file_location(
    "qe_b200", os.path.join(_d, "__init__.py"), submodule_search_locations=[_d]
)
_m = importlib.util.module_from_spec(_spec)
sys.modules["qe_b200"] = _m
_spec.loader.exec_module(_m)
