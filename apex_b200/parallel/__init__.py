"""reference: apex/parallel/__init__.py:1"""
from .distributed import DistributedDataParallel, Reducer
from .LARC import LARC
