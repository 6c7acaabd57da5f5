"""The one piece of logic AmpHandle.scale_loss and OptimWrapper.scale_loss share.

The reference carries two copies of it (apex/amp/handle.py:41-55 and
apex/amp/opt.py:24-50, with a "TODO: ... Unify" between them): hand the user
``loss * scale``, warn if they backpropagate the unscaled loss instead, and on
exit check + unscale the gradients that backward produced.
"""
import warnings


class ScaledBackward(object):
    """``with ScaledBackward(loss, scaler, param_groups) as scope: scope.scaled_loss.backward()``

    After the block ``scope.overflowed`` tells whether the gradients held an
    inf/NaN (in which case, as in the reference, gradients before the first bad
    tensor are already unscaled and the rest are untouched).
    """

    _MESSAGE = ("You called .backward() on the unscaled loss "
                "inside a scale_loss block. This is almost "
                "certainly an error.")

    def __init__(self, loss, scaler, param_groups):
        self._loss = loss
        self._scaler = scaler
        self._param_groups = param_groups
        self._plain_backward = None
        self._scale = None
        self.scaled_loss = None
        self.overflowed = False

    def __enter__(self):
        plain = self._plain_backward = self._loss.backward
        message = self._MESSAGE

        def backward_with_warning():
            warnings.warn(message, stacklevel=2)
            plain()
        self._loss.backward = backward_with_warning
        self._scale = self._scaler.loss_scale()
        self.scaled_loss = self._loss * self._scale
        return self

    def __exit__(self, exc_type, exc, tb):
        if exc_type is not None:
            return False          # like a generator context manager: no cleanup on error
        self._loss.backward = self._plain_backward
        self.overflowed = self._scaler.unscale_and_update(self._param_groups, self._scale)
        return False
