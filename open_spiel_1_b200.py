"""Import shim: makes the package directory `open_spiel-1_b200/` importable as
`open_spiel_1_b200` (a hyphen is not valid in a module name)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "open_spiel-1_b200")
_spec = importlib.util.spec_from_file_location(
    "open_spiel_1_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["open_spiel_1_b200"] = _mod
_spec.loader.exec_module(_mod)
