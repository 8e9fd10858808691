"""Import shim: the product package lives in the directory `stereo.vision_b200/` at the repo root.

A directory name with a dot cannot be imported by name, so this one-file package registers it as
the submodule `stereo.vision_b200`; `import stereo.vision_b200.functions as f` then works as usual.
"""
import importlib.util as _ilu
import os as _os
import sys as _sys

_root = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "stereo.vision_b200")
if "stereo.vision_b200" not in _sys.modules:
    _spec = _ilu.spec_from_file_location(
        "stereo.vision_b200", _os.path.join(_root, "__init__.py"), submodule_search_locations=[_root])
    _mod = _ilu.module_from_spec(_spec)
    _sys.modules["stereo.vision_b200"] = _mod
    _spec.loader.exec_module(_mod)
vision_b200 = _sys.modules["stereo.vision_b200"]
