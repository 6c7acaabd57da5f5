"""numpy-facing wrapper of the C oracle.  TEST INFRASTRUCTURE (see oracle/__init__.py)."""
import ctypes

import numpy as np

from . import build as _build

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(_build.build())
        LL, D, I, P = ctypes.c_longlong, ctypes.c_double, ctypes.c_int, ctypes.c_void_p
        _lib.ora_first_nonfinite_f16.restype = LL
        _lib.ora_first_nonfinite_f16.argtypes = [P, LL]
        _lib.ora_first_nonfinite_f32.restype = LL
        _lib.ora_first_nonfinite_f32.argtypes = [P, LL]
        _lib.ora_f16_to_f32_scaled.argtypes = [P, P, LL, D, I]
        _lib.ora_scale_f32.argtypes = [P, LL, D]
        _lib.ora_f32_to_f16.argtypes = [P, P, LL]
        F = ctypes.c_float
        _lib.ora_sgd.argtypes = [P, P, P, P, P, LL, D, D, D, D, I, I, D, I, F, I]
        _lib.ora_clip_coef.restype = F
        _lib.ora_clip_coef.argtypes = [F, D]
        _lib.ora_f16_to_f32_scaled_clipped.argtypes = [P, P, LL, D, F]
        _lib.ora_adam.argtypes = [P, P, P, P, P, P, LL, D, D, D, D, D, D, D, I, D, I]
        _lib.ora_adam_flavour.argtypes = [P, P, P, P, P, P, LL, D, D, D, D, D, D, D, I, D, I, I, I, F, I]
        _lib.ora_average_f16.argtypes = [P, LL, D]
        _lib.ora_average_f32.argtypes = [P, LL, D]
        _lib.ora_sumsq_f16.restype = D
        _lib.ora_sumsq_f16.argtypes = [P, LL, D, I]
        _lib.ora_sumsq_f32.restype = D
        _lib.ora_sumsq_f32.argtypes = [P, LL, D, I]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data


def _c(a, dt):
    a = np.ascontiguousarray(a)
    assert a.dtype == dt, (a.dtype, dt)
    return a


def first_nonfinite(g):
    g = np.ascontiguousarray(g)
    if g.dtype == np.uint16:
        return int(lib().ora_first_nonfinite_f16(_p(g), g.size))
    return int(lib().ora_first_nonfinite_f32(_p(_c(g, np.float32)), g.size))


def f16_to_f32_scaled(g_bits, inv_scale=1.0, apply=False, clip_coef=None):
    g = _c(g_bits, np.uint16)
    out = np.empty(g.shape, np.float32)
    if clip_coef is None:
        lib().ora_f16_to_f32_scaled(_p(g), _p(out), g.size, float(inv_scale), int(apply))
    else:
        lib().ora_f16_to_f32_scaled_clipped(_p(g), _p(out), g.size, float(inv_scale) if apply else 1.0,
                                            float(clip_coef))
    return out


def clip_coef(total_norm, max_norm):
    """clamp(max_norm / (norm + 1e-6), max=1) in fp32, as torch's clip_grad_norm_ computes it."""
    return float(lib().ora_clip_coef(float(np.float32(total_norm)), float(max_norm)))


def f32_to_f16(p):
    p = _c(p, np.float32)
    out = np.empty(p.shape, np.uint16)
    lib().ora_f32_to_f16(_p(p), _p(out), p.size)
    return out


def sgd(g, p, buf, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, first=False,
        inv_scale=1.0, apply_scale=False, want_p16=True, clip_coef=None):
    """In place on p (fp32) and buf (fp32 or None); g is uint16 (fp16 bits) or float32.  Returns p16 bits."""
    g = np.ascontiguousarray(g)
    g16 = g if g.dtype == np.uint16 else None
    g32 = g if g.dtype == np.float32 else None
    assert p.flags.c_contiguous and p.dtype == np.float32
    p16 = np.empty(p.shape, np.uint16) if want_p16 else None
    lib().ora_sgd(_p(g16), _p(g32), _p(p), _p(buf), _p(p16), p.size, float(lr), float(momentum), float(dampening),
                  float(weight_decay), int(nesterov), int(first), float(inv_scale), int(apply_scale),
                  1.0 if clip_coef is None else float(clip_coef), int(clip_coef is not None))
    return p16


def adam(g, p, m, v, lr, beta1, beta2, eps, weight_decay, step, decoupled=False, inv_scale=1.0,
         apply_scale=False, want_p16=True, flavour="cuda", cpu_fma=True, clip_coef=None):
    """``step`` is the 1-based step count; bias corrections follow torch/optim/adam.py:764-775.
    flavour "cuda" = torch's CUDA foreach association (what the GPU kernel must equal),
    "cpu" = torch's CPU kernels' association (what tests/golden was generated with)."""
    g = np.ascontiguousarray(g)
    g16 = g if g.dtype == np.uint16 else None
    g32 = g if g.dtype == np.float32 else None
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    step_size = (lr / bc1) * -1
    bc2_sqrt = bc2 ** 0.5
    p16 = np.empty(p.shape, np.uint16) if want_p16 else None
    lib().ora_adam_flavour(_p(g16), _p(g32), _p(p), _p(m), _p(v), _p(p16), p.size, float(lr), float(beta1),
                           float(beta2), float(eps), float(weight_decay), float(step_size), float(bc2_sqrt),
                           int(decoupled), float(inv_scale), int(apply_scale),
                           0 if flavour == "cuda" else 1, int(cpu_fma),
                           1.0 if clip_coef is None else float(clip_coef), int(clip_coef is not None))
    return p16


def average(x, divisor):
    x = np.ascontiguousarray(x)
    if x.dtype == np.uint16:
        lib().ora_average_f16(_p(x), x.size, float(divisor))
    else:
        lib().ora_average_f32(_p(_c(x, np.float32)), x.size, float(divisor))
    return x


def l2norm(grads, inv_scale=1.0, apply=False):
    total = 0.0
    for g in grads:
        g = np.ascontiguousarray(g)
        if g.dtype == np.uint16:
            total += lib().ora_sumsq_f16(_p(g), g.size, float(inv_scale), int(apply))
        else:
            total += lib().ora_sumsq_f32(_p(_c(g, np.float32)), g.size, float(inv_scale), int(apply))
    return float(np.float32(np.sqrt(total)))
