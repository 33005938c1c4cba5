"""Import shim: the package directory is `k8s-nvidia-gpus_b200/` (not a valid identifier), so
`import b200sd` resolves to it via importlib.  Submodules are attributes: `b200sd.pipeline`, `b200sd.ops`, ..."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("k8s-nvidia-gpus_b200")
sys.modules[__name__] = _pkg
