"""Multi-scale learned simulator (mirrors sgnn/multi_scale/__init__.py:3-7)."""
from .multi_scale_gnn import MultiScaleGNN
from .multi_scale_graph import MultiScaleConfig, MultiScaleGraph
from .multi_scale_simulator import MultiScaleSimulator

__all__ = ["MultiScaleGNN", "MultiScaleGraph", "MultiScaleConfig", "MultiScaleSimulator"]
