"""Import shim: ``import nlbac_b200`` loads the package whose directory carries the reference's name."""
import importlib
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
if _here not in sys.path:
    sys.path.insert(0, _here)
_long = "neural-ordinary-differential-equations-based-lyapunov-barrier-actor-critic-nlbac_b200"
_pkg = importlib.import_module(_long)
for _sub in ("dopri5", "parallel"):          # submodules the package itself imports lazily
    importlib.import_module(_long + "." + _sub)
# Alias every submodule under the short name too: otherwise ``from nlbac_b200.odeint import describe`` would load a SECOND
# copy of the submodule (its own caches) and rebind ``nlbac_b200.odeint`` from the function to that module.
for _k, _v in list(sys.modules.items()):
    if _k.startswith(_long + "."):
        sys.modules[__name__ + _k[len(_long):]] = _v
sys.modules[__name__] = _pkg
