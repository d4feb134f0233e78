"""Import shim: the package directory is named after the reference project and contains a hyphen
(`physically-based-renderer_b200/`), so it cannot be imported by name.  `import pbr_b200` loads it."""
import importlib.util
import sys
from pathlib import Path

_PKG = Path(__file__).resolve().parent / "physically-based-renderer_b200"
_spec = importlib.util.spec_from_file_location(
    "pbr_b200", _PKG / "__init__.py", submodule_search_locations=[str(_PKG)])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["pbr_b200"] = _mod
_spec.loader.exec_module(_mod)
