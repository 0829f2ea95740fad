"""Import helper: the package directory is named `opengl-ray-tracer_b200` (not a valid
Python identifier), so it is loaded by path under the module name `opengl_ray_tracer_b200`."""
import importlib.util
import os
import sys

_NAME = "opengl_ray_tracer_b200"


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    root = os.path.dirname(os.path.abspath(__file__))
    pkg = os.path.join(root, "opengl-ray-tracer_b200")
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(pkg, "__init__.py"), submodule_search_locations=[pkg])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod
