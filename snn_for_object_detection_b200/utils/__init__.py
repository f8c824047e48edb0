"""Detection utilities kept next to the hot path (anchors, box decode / NMS)."""

from .anchors import AnchorGenerator  # noqa: F401
