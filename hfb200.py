"""Import shim: the package directory is named after the reference repository and contains hyphens,
so it cannot be imported with a plain `import` statement. `import hfb200` gives the same module."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module("mitsuba3-differentiable-heightfield-rendering_b200")
workloads = importlib.import_module("mitsuba3-differentiable-heightfield-rendering_b200.workloads")
sys.modules[__name__] = _pkg
_pkg.workloads = workloads
_pkg.formats = importlib.import_module("mitsuba3-differentiable-heightfield-rendering_b200.formats")
