"""Import alias: ``import b200ad`` loads the package in video-diffusion-experiments_b200/ (not a valid identifier)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "video-diffusion-experiments_b200")
_spec = importlib.util.spec_from_file_location("b200ad", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["b200ad"] = _mod
_spec.loader.exec_module(_mod)
