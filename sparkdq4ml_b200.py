"""Import shim: makes the package directory `net.jgp.labs.sparkdq4ml_b200/` (the reference's dotted package name)
importable as `sparkdq4ml_b200`.  `import sparkdq4ml_b200` returns the real package."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "net.jgp.labs.sparkdq4ml_b200")
_spec = importlib.util.spec_from_file_location(__name__, os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
