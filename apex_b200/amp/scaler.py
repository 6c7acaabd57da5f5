"""amp's loss scaler (reference: apex/amp/scaler.py).

The reference checks and unscales gradient by gradient, with a blocking
``float(d_grads.float().sum())`` for each (scaler.py:6-18, :35-40).  Its visible
side effect on overflow is a *partial* unscale: gradients before the first bad
tensor have been multiplied by 1/scale, the bad one and everything after it are
untouched.  Here: one check launch that leaves "index of the first bad tensor"
in a device word, one unscale launch that applies the multiply only to tensors
before that index, and one host read at the end -- same result, same return
value, two launches instead of 3n.
"""
import torch

from .. import _C
from ..multi_tensor import TensorTable, CHUNK_ZERO, new_flag, is_dense


def _flags_for(grads):
    flags = []
    for g in grads:
        if g.dtype == torch.float32:
            flags.append(_C.TF_GRAD_F32)
        elif g.dtype == torch.float16:
            flags.append(0)
        else:
            raise TypeError("apex_b200.amp: gradients must be fp16 or fp32, got %s" % g.dtype)
        if not is_dense(g):
            raise RuntimeError("apex_b200.amp: gradients must be dense (contiguous or channels-last)")
    return flags


class _Unscaler(object):
    def __init__(self):
        self.table = None
        self.flag = None

    def run(self, grads, inv_scale):
        """check + partial unscale of ``grads`` in order; returns True on overflow."""
        if not grads:
            return False
        dev = grads[0].device
        if self.table is None or self.table.device != dev:
            self.table = TensorTable(dev, 1, CHUNK_ZERO)     # shared by the scan and the in-place unscale
            self.flag = new_flag(dev)
        self.table.update([grads], _flags_for(grads))
        _C.flag_clear(self.flag)
        _C.mt_check_overflow(self.table, 0, self.flag)
        _C.mt_unscale(self.table, 0, inv_scale, self.flag, 1)
        return int(self.flag.item()) != 0


def scale_check_overflow(d_grads, scale):
    """``d_grads *= scale`` unless it holds an inf/NaN; returns True in that case.

    This is the function the reference meant to import from a native ``apex_C``
    (scaler.py:3-5).
    """
    return _Unscaler().run([d_grads], scale)


class LossScaler(object):
    """reference: scaler.py:20-55 -- 2^16 start, x2 every 2000 clean steps up to 2^24, /2 on overflow."""

    def __init__(self):
        self._loss_scale = 2.**16
        self._max_loss_scale = 2.**24
        self._scale_seq_len = 2000
        self._unskipped = 0
        self._has_overflow = False
        self._unscaler = _Unscaler()

    def loss_scale(self):
        return self._loss_scale

    def unscale_and_update(self, param_groups, scale):
        grads = [p.grad.data for p in iter_params(param_groups) if p.grad is not None]
        self._has_overflow = self._unscaler.run(grads, 1. / scale)

        if self._has_overflow:
            should_skip = True
            self._loss_scale /= 2.
            self._unskipped = 0
        else:
            should_skip = False
            self._unskipped += 1

        if self._unskipped == self._scale_seq_len:
            self._loss_scale = min(self._max_loss_scale, self._loss_scale * 2.)
            self._unskipped = 0

        return should_skip


def iter_params(param_groups):
    for group in param_groups:
        for p in group['params']:
            yield p
