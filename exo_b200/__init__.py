"""Import alias: `import exo_b200` loads the package directory
`a-deep-reinforcement-learning-enabled-soft-exoskeleton-for-parkinson-s-patients_b200/` (hyphens are not valid in
an import statement, so it is loaded through importlib and re-exported here)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _root not in sys.path:
    sys.path.insert(0, _root)
_NAME = "a-deep-reinforcement-learning-enabled-soft-exoskeleton-for-parkinson-s-patients_b200"
_pkg = importlib.import_module(_NAME)
for _sub in ("_capi", "env", "lap", "policy", "td7", "train", "evaluate", "urdf", "motions", "build", "sharding", "reference_stream", "hostpin", "metrics"):
    sys.modules[__name__ + "." + _sub] = importlib.import_module(_NAME + "." + _sub)
    globals()[_sub] = sys.modules[__name__ + "." + _sub]
globals().update({k: getattr(_pkg, k) for k in _pkg.__all__})
__all__ = list(_pkg.__all__)
