"""Import shim: the package directory is named after the reference repository
(`duckdb-robust-predicate-transfer_b200/`), which is not a valid Python identifier, so it is loaded
here under the module name ``rpt_b200``.  `import rpt_b200` is the public Python entry point."""
import importlib.util
import os
import sys

_PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "duckdb-robust-predicate-transfer_b200")

if "rpt_b200_pkg" not in sys.modules:
    _spec = importlib.util.spec_from_file_location(
        "rpt_b200_pkg", os.path.join(_PKG_DIR, "__init__.py"), submodule_search_locations=[_PKG_DIR])
    _mod = importlib.util.module_from_spec(_spec)
    sys.modules["rpt_b200_pkg"] = _mod
    _spec.loader.exec_module(_mod)

_pkg = sys.modules["rpt_b200_pkg"]
globals().update({k: v for k, v in vars(_pkg).items() if not k.startswith("__")})
PACKAGE_DIR = _PKG_DIR
