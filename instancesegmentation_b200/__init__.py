"""B200-native hot path of YanMiaoW/instanceSegmentation: Segment forward + mask decode."""
from .segment import Segment, topology  # noqa: F401

__all__ = ["Segment", "topology"]
