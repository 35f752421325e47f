"""Import alias: the package directory is named `stablediffusion-pytorch_b200` (not a Python identifier), so this
bootstrap loads it under the importable name `sdb200` (`import sdb200; sdb200.runtime ...`)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stablediffusion-pytorch_b200")
_spec = importlib.util.spec_from_file_location("sdb200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["sdb200"] = _mod
_spec.loader.exec_module(_mod)
