"""sgnn_b200: B200-native (sm_100a) implementation of the sgnn Encode-Process-Decode learned
simulator hot path, behind the reference's Python module API.  See DESIGN.md."""
__version__ = "0.1.0"
