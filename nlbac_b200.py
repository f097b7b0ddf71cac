"""Import shim: ``import nlbac_b200`` loads the package whose directory carries the reference's name."""
import importlib
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
if _here not in sys.path:
    sys.path.insert(0, _here)
_pkg = importlib.import_module(
    "neural-ordinary-differential-equations-based-lyapunov-barrier-actor-critic-nlbac_b200")
sys.modules[__name__] = _pkg
