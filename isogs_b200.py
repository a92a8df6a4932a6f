"""Import helper: loads the package directory `isogs-slam-cpp_b200/` (not a valid identifier)
as the module `isogs_slam_cpp_b200`.  Usage: `import isogs_b200; pkg = isogs_b200.pkg`."""
import importlib.util
import os
import sys

_NAME = "isogs_slam_cpp_b200"
_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "isogs-slam-cpp_b200")


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(_DIR, "__init__.py"),
                                                  submodule_search_locations=[_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    try:
        spec.loader.exec_module(mod)
    except BaseException:
        del sys.modules[_NAME]
        raise
    return mod


pkg = load()
