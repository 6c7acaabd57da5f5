"""apex_b200 -- B200-native drop-in for apex's mixed-precision gradient step.

Mirrors the import surface of the reference package root
(apex/__init__.py:1-5: fp16_utils, parallel, amp) for the hot path only; see
DESIGN.md for what is in and out of scope.
"""
from . import fp16_utils
from . import parallel
from . import amp

__all__ = ["fp16_utils", "parallel", "amp"]
