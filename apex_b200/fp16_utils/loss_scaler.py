"""Static and dynamic loss scalers (reference: apex/fp16_utils/loss_scaler.py).

Host arithmetic (scale trajectory) is the reference's, bit for bit.  What
changes is inf/NaN detection: the reference does ``float(x.float().sum())`` per
tensor -- two launches and a blocking device-to-host read each
(loss_scaler.py:92-110); here all gradients are scanned by ONE multi-tensor
launch that raises a single device word, and the host reads that word once.
"""
import torch

from .. import _C
from ..multi_tensor import TensorTable, CHUNK_CHECK, new_flag, is_dense


def to_python_float(t):
    """reference: loss_scaler.py:4-8"""
    return t.item() if hasattr(t, "item") else t[0]


def _grad_flags(tensors):
    flags = []
    for t in tensors:
        if t.dtype == torch.float32:
            flags.append(_C.TF_GRAD_F32)
        elif t.dtype == torch.float16:
            flags.append(0)
        else:
            raise TypeError("apex_b200 overflow check supports fp16 and fp32 gradients, got %s" % t.dtype)
    return flags


class _OverflowProbe(object):
    """Device flag + table for one group of gradient tensors (not pickled)."""

    def __init__(self):
        self.table = None
        self.flag = None

    def launch(self, grads):
        """Clear the flag and scan ``grads``; returns the device flag tensor (no sync)."""
        dev = grads[0].device
        if self.table is None or self.table.device != dev:
            self.table = TensorTable(dev, 1, CHUNK_CHECK)
            self.flag = new_flag(dev)
        for g in grads:
            if not is_dense(g):
                raise RuntimeError("apex_b200: gradients must be dense (contiguous or channels-last)")
        self.table.update([grads], _grad_flags(grads))
        _C.flag_clear(self.flag)
        _C.mt_check_overflow(self.table, 0, self.flag)
        return self.flag

    def any_bad(self, grads):
        if len(grads) == 0:
            return False
        return int(self.launch(grads).item()) != 0


_shared_probes = {}     # device -> probe used by the class-level DynamicLossScaler._has_inf_or_nan


class LossScaler:
    """Fixed loss scale (reference: loss_scaler.py:10-45).  Never reports overflow."""

    def __init__(self, scale=1):
        self.cur_scale = scale

    def has_overflow(self, params):
        return False

    def _has_inf_or_nan(x):
        return False

    def update_scale(self, overflow):
        pass

    @property
    def loss_scale(self):
        return self.cur_scale

    def scale_gradient(self, module, grad_in, grad_out):
        return tuple(self.loss_scale * g for g in grad_in)

    def backward(self, loss):
        (loss * self.loss_scale).backward()


class DynamicLossScaler:
    """Dynamic loss scale (reference: loss_scaler.py:47-132).

    Trajectory rules, unchanged: on overflow ``cur_scale = max(cur_scale /
    scale_factor, 1)`` and the iteration is remembered; otherwise the scale is
    multiplied by ``scale_factor`` whenever ``(cur_iter - last_overflow_iter) %
    scale_window == 0``; ``cur_iter`` always advances (loss_scaler.py:113-121).
    """

    def __init__(self, init_scale=2**32, scale_factor=2., scale_window=1000):
        self.cur_scale = init_scale
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor = scale_factor
        self.scale_window = scale_window

    # The probe holds device memory; the reference's state_dict pickles the
    # scaler object itself (fp16_optimizer.py:279), so keep it out of __dict__'s
    # pickled form and keep the attribute names identical to the reference's.
    def __getstate__(self):
        d = dict(self.__dict__)
        d.pop("_probe", None)
        return d

    def _get_probe(self):
        probe = self.__dict__.get("_probe")
        if probe is None:
            probe = self.__dict__["_probe"] = _OverflowProbe()
        return probe

    def has_overflow(self, params):
        """One launch over every ``p.grad`` instead of a sync per tensor (loss_scaler.py:84-89)."""
        grads = [p.grad.data for p in params if p.grad is not None]
        return self._get_probe().any_bad(grads)

    def _has_inf_or_nan(x):
        """Single tensor form (loss_scaler.py:92-110); no ``self``, as in the reference."""
        probe = _shared_probes.get(x.device)
        if probe is None:
            probe = _shared_probes[x.device] = _OverflowProbe()
        return probe.any_bad([x])

    def update_scale(self, overflow):
        if overflow:
            self.cur_scale = max(self.cur_scale / self.scale_factor, 1)
            self.last_overflow_iter = self.cur_iter
        elif (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0:
            self.cur_scale *= self.scale_factor
        self.cur_iter += 1

    @property
    def loss_scale(self):
        return self.cur_scale

    def scale_gradient(self, module, grad_in, grad_out):
        return tuple(self.loss_scale * g for g in grad_in)

    def backward(self, loss):
        (loss * self.loss_scale).backward()
