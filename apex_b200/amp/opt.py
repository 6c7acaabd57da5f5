"""amp's optimizer wrapper: one loss scaler per loss (reference: apex/amp/opt.py).

With several losses per iteration the gradients of loss k can only be unscaled
while they are alone in ``.grad``; so what losses 0..k-1 accumulated is moved
aside before backward k and added back after unscaling (opt.py:33-58).  ``step``
is dropped for the iteration if any loss overflowed (opt.py:64-82).
"""
import contextlib
import logging

from ._scope import ScaledBackward
from .scaler import LossScaler, iter_params


class OptimWrapper(object):
    def __init__(self, optimizer, amp_handle, num_loss):
        self._optimizer = optimizer
        self._amp_handle = amp_handle
        self._num_loss = num_loss
        self._loss_idx = 0
        self._skip_next = [False] * num_loss
        self._loss_scaler = [LossScaler() for _ in range(num_loss)]

    def _cur_loss_scaler(self):
        assert 0 <= self._loss_idx < self._num_loss
        return self._loss_scaler[self._loss_idx]

    def _set_aside_grads(self):
        """Clone what earlier losses accumulated and zero it (only needed from the 2nd loss on)."""
        if self._loss_idx == 0:
            return None
        saved = [None if p.grad is None else p.grad.data.detach().clone()
                 for p in iter_params(self._optimizer.param_groups)]
        self._optimizer.zero_grad()
        return saved

    def _add_back(self, saved):
        if not saved:
            return
        for p, old in zip(iter_params(self._optimizer.param_groups), saved):
            if old is not None:
                p.grad.data.add_(old)

    @contextlib.contextmanager
    def scale_loss(self, loss):
        if not self._amp_handle.is_active():
            yield loss
            return
        scope = ScaledBackward(loss, self._cur_loss_scaler(), self._optimizer.param_groups)
        with scope:
            saved = self._set_aside_grads()
            yield scope.scaled_loss
        self._skip_next[self._loss_idx] = scope.overflowed
        self._loss_idx += 1
        self._add_back(saved)

    def step(self, closure=None):
        if not self._amp_handle.is_active():
            return self._optimizer.step(closure=closure)

        self._loss_idx = 0
        if self._amp_handle.has_cache:
            for p in iter_params(self._optimizer.param_groups):
                self._amp_handle.remove_cache(p)

        if closure is not None:
            raise NotImplementedError(
                'The `closure` argument is unsupported by the amp ' +
                'optimizer wrapper.')
        if any(self._skip_next):
            logging.info('Gradient overflow, skipping update')
            self._skip_next = [False] * self._num_loss
            return None
        return self._optimizer.step(closure=closure)

    # anything else (param_groups, state, ...) is looked up on the wrapped optimizer
    def __getattr__(self, attr):
        return getattr(self._optimizer, attr)


def _delegate(name):
    def method(self, *args, **kwargs):
        return getattr(self._optimizer, name)(*args, **kwargs)
    method.__name__ = name
    method.__doc__ = "Forwarded to the wrapped optimizer (reference: opt.py:84-108)."
    return method


# methods python would otherwise resolve on OptimWrapper itself (dunder lookups
# bypass __getattr__) or that the reference spells out one by one
for _name in ("__getstate__", "__setstate__", "__repr__", "state_dict", "load_state_dict",
              "zero_grad", "add_param_group"):
    setattr(OptimWrapper, _name, _delegate(_name))
del _name
