"""TEST INFRASTRUCTURE ONLY -- see torch_geometric/nn/__init__.py in this shim."""
from . import nn  # noqa: F401
