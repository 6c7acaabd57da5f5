"""amp handles, gradient-step half only (reference: apex/amp/handle.py).

``AmpHandle.scale_loss`` yields the scaled loss; when the block ends the
gradients are checked and unscaled by ``scaler.LossScaler`` (two kernel
launches) and, on overflow, the optimizer's next ``step()`` is turned into a
one-shot no-op (handle.py:54-61).  The cast cache the reference handle also
carries belongs to the function-patching machinery, which is out of scope:
``has_cache`` is kept (False unless asked for) so OptimWrapper's calls stay valid.
"""
import contextlib
import logging

from ._scope import ScaledBackward
from .opt import OptimWrapper
from .scaler import LossScaler


def _skip_next_step(optimizer):
    """Replace ``optimizer.step`` by a function that logs, restores the real one and
    does nothing else."""
    real_step = optimizer.step

    def skip_step():
        logging.info('Gradient overflow, skipping update')
        optimizer.step = real_step
    optimizer.step = skip_step


class AmpHandle(object):
    def __init__(self, enable_caching=True, verbose=False):
        self._enable_caching = enable_caching
        self._verbose = verbose
        self._cache = {}
        self._default_scaler = LossScaler()
        self._is_active = True

    def is_active(self):
        return self._is_active

    @contextlib.contextmanager
    def _disable_casts(self):
        self._is_active = False
        yield
        self._is_active = True

    def wrap_optimizer(self, optimizer, num_loss=1):
        # from here on each wrapped optimizer owns its scalers (handle.py:28-30)
        self._default_scaler = None
        return OptimWrapper(optimizer, self, num_loss)

    @contextlib.contextmanager
    def scale_loss(self, loss, optimizer):
        if not self.is_active():
            yield loss
            return
        if self._default_scaler is None:
            raise RuntimeError(
                'After calling `handle.wrap_optimizer()`, you must explicitly ' +
                'use `optimizer.scale_loss(loss)`.')
        with ScaledBackward(loss, self._default_scaler, optimizer.param_groups) as scope:
            yield scope.scaled_loss
        if scope.overflowed:
            _skip_next_step(optimizer)
        self._cache.clear()

    has_cache = property(lambda self: self._enable_caching)
    cache = property(lambda self: self._cache)
    verbose = property(lambda self: self._verbose)

    def remove_cache(self, param):
        self._cache.pop(param, None)


class NoOpHandle(object):
    """Inactive handle: everything passes through (handle.py:92-114)."""
    has_cache = False
    verbose = False

    def is_active(self):
        return False

    @contextlib.contextmanager
    def _disable_casts(self):
        yield

    def wrap_optimizer(self, optimizer, num_loss=1):
        return OptimWrapper(optimizer, self, num_loss)

    @contextlib.contextmanager
    def scale_loss(self, loss, optimizer):
        yield loss
