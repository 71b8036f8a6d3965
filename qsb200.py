"""Import alias: `import qsb200` loads the package directory `quantumsavory.jl_b200/`
(its name is not a valid Python identifier, so it is bound to the module name `qsb200`)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "quantumsavory.jl_b200")
_spec = importlib.util.spec_from_file_location(
    "qsb200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["qsb200"] = _mod
_spec.loader.exec_module(_mod)
