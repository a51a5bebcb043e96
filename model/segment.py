"""Import path kept from the reference: ``from model.segment import Segment``
(train_instance.py:294 builds ``Segment(3)`` from this module)."""
from instancesegmentation_b200.segment import Segment  # noqa: F401

__all__ = ["Segment"]
