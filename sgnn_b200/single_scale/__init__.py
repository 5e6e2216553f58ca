"""Single-scale learned simulator (mirrors sgnn/single_scale/__init__.py:3-6)."""
from .graph_network import EncodeProcessDecode
from .learned_simulator import LearnedSimulator

__all__ = ["EncodeProcessDecode", "LearnedSimulator"]
