"""Import alias: the package lives in the directory ``rvv-lut-vq_b200/`` (the name the project layout
prescribes), which is not a valid Python identifier; this module loads it as ``rvv_lut_vq_b200``."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "rvv-lut-vq_b200")
_spec = importlib.util.spec_from_file_location(
    "rvv_lut_vq_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir]
)
_mod = importlib.util.module_from_spec(_spec)
sys.modules["rvv_lut_vq_b200"] = _mod
_spec.loader.exec_module(_mod)
