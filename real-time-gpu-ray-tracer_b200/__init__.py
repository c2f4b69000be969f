"""rtb200 — B200-native per-frame ray-tracing hot path (host-side Python mirror over the C-ABI).

The directory name contains hyphens; import it with
``importlib.import_module("real-time-gpu-ray-tracer_b200")`` (``__graft_entry__.package()`` does so).
"""
from . import abi, scenes, sharding  # noqa: F401
from .renderer import Renderer, RtError, cameraStep  # noqa: F401

__all__ = ["abi", "scenes", "sharding", "Renderer", "RtError", "cameraStep"]
