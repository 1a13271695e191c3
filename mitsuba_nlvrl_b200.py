"""Import shim: the package directory is `mitsuba-nlvrl_b200/` (not a valid Python identifier), so
`import mitsuba_nlvrl_b200` from the repo root lands here and is replaced by that package."""
import importlib.util as _u, os as _os, sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "mitsuba-nlvrl_b200")
_spec = _u.spec_from_file_location(__name__, _os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = _u.module_from_spec(_spec)
_sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
