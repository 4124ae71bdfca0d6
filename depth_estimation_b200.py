"""Import alias: ``import depth_estimation_b200`` -> the package in ./depth-estimation_b200/.

The package directory carries the reference repository's name (with a hyphen),
which Python cannot import directly; this module replaces itself in
``sys.modules`` with that package so submodule imports work as usual.
"""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "depth-estimation_b200")
_spec = importlib.util.spec_from_file_location(
    __name__, os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
