"""Import shim: the package directory is `sexy-raytracer_b200/` (a hyphen is not importable), so this
module loads it under the name `sexy_raytracer_b200`."""
import importlib.util
import os
import sys

_d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sexy-raytracer_b200")
_spec = importlib.util.spec_from_file_location(__name__, os.path.join(_d, "__init__.py"),
                                               submodule_search_locations=[_d])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
