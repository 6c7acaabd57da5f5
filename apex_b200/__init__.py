"""apex_b200 -- B200-native drop-in for apex's mixed-precision gradient step.

Mirrors the import surface of the reference package root
(apex/__init__.py:1-5: fp16_utils, parallel, amp) for the hot path only; see
DESIGN.md for what is in and out of scope.
"""
from . import fp16_utils
from . import parallel
from . import amp

__all__ = ["fp16_utils", "parallel", "amp"]


def install_as_apex():
    """Register this package under the name ``apex`` (``import apex``, ``apex.fp16_utils``,
    ``apex.parallel``, ``apex.amp.scaler`` ...), for scripts written against the reference and for
    checkpoints it wrote: FP16_Optimizer.state_dict() pickles the loss-scaler OBJECT
    (apex/fp16_utils/fp16_optimizer.py:279), so loading one needs
    ``apex.fp16_utils.loss_scaler.DynamicLossScaler`` to be importable.  Opt-in, so that a real apex
    installation is never shadowed by accident."""
    import sys
    from .fp16_utils import fp16_optimizer, fp16util, loss_scaler, fused_weight_norm
    from .parallel import distributed
    from .amp import scaler, handle, opt
    this = sys.modules[__name__]
    aliases = {
        "apex": this,
        "apex.fp16_utils": fp16_utils,
        "apex.fp16_utils.fp16_optimizer": fp16_optimizer,
        "apex.fp16_utils.fp16util": fp16util,
        "apex.fp16_utils.loss_scaler": loss_scaler,
        "apex.fp16_utils.fused_weight_norm": fused_weight_norm,
        "apex.parallel": parallel,
        "apex.parallel.distributed": distributed,
        "apex.parallel.LARC": sys.modules[__name__ + ".parallel.LARC"],
        "apex.amp": amp,
        "apex.amp.scaler": scaler,
        "apex.amp.handle": handle,
        "apex.amp.opt": opt,
    }
    for name, module in aliases.items():
        existing = sys.modules.get(name)
        if existing is not None and existing is not module:
            raise RuntimeError("install_as_apex: %s is already imported from %s"
                               % (name, getattr(existing, "__file__", "?")))
        sys.modules[name] = module
    return this
