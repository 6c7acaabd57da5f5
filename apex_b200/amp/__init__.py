"""apex.amp, hot-path part only.

In scope: the loss scaler's fused check+unscale
(`scaler.py`) and its two callers (`handle.py`, `opt.py`).  The function-patching
cast machinery of the reference (apex/amp/amp.py, wrap.py, utils.py, lists/*) is
forward-pass code, cannot initialise on torch >= 2 (torch.nn.backends is gone)
and is out of scope.
"""
from . import scaler
from .scaler import LossScaler, scale_check_overflow, iter_params
from .handle import AmpHandle, NoOpHandle
from .opt import OptimWrapper
