"""apex_b200.parallel: DistributedDataParallel and Reducer as in the reference
(apex/parallel/__init__.py:1), plus LARC (the reference keeps it in apex.parallel.LARC)."""
from . import distributed as _distributed
from . import LARC as _larc_module

DistributedDataParallel = _distributed.DistributedDataParallel
Reducer = _distributed.Reducer
LARC = _larc_module.LARC
__all__ = ["DistributedDataParallel", "Reducer", "LARC"]
