"""TEST INFRASTRUCTURE ONLY -- makes the UNMODIFIED reference importable in the build
container (where /root/reference exists): puts /root/reference and the PyG shim on sys.path
and aliases the reference's `datasets/` directory (it has no __init__.py and would lose to
the installed HuggingFace `datasets` package).

Never imported on the GPU box: /root/reference does not exist there.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("SGNN_REFERENCE_ROOT", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "sgnn"))


def load():
    """Returns the reference's `sgnn` package (imports happen lazily by the caller)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    shim = os.path.join(_HERE, "pyg_shim")
    for p in (shim, REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    if "datasets" not in sys.modules or not getattr(sys.modules["datasets"], "_sgnn_alias", False):
        pkg = types.ModuleType("datasets")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "datasets")]
        pkg._sgnn_alias = True
        sys.modules["datasets"] = pkg
    import sgnn  # noqa: F401  (the reference's package)
    return sys.modules["sgnn"]
