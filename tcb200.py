"""Import shim: the package directory is named `trajopt-1_b200` (not a Python identifier), so it is
registered in sys.modules as `trajopt1_b200`.  `import tcb200; pkg = tcb200.load()`."""
import importlib.util
import os
import sys

_NAME = "trajopt1_b200"
_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "trajopt-1_b200")


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(_DIR, "__init__.py"),
                                                  submodule_search_locations=[_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod
