"""CPU test double for apex_b200._C's python wrappers, built on the oracle.

TEST INFRASTRUCTURE.  It lets `-m "not gpu"` tests drive the real host code
(FP16_Optimizer, scalers, DistributedDataParallel, TensorTable) on CPU tensors:
each kernel wrapper is replaced by a function that walks the table's tensor
lists and applies oracle/mp_step_oracle.c to them.  Nothing here is reachable
from the product package.
"""
import contextlib
from unittest import mock

import numpy as np
import torch

from apex_b200 import _C
from apex_b200.multi_tensor import TensorTable
from oracle import c_oracle

INT32_MAX = 0x7FFFFFFF
calls = []


def _bits(t):
    """numpy view sharing memory with CPU tensor t (fp16 as uint16)."""
    a = t.detach().numpy()
    return a.view(np.uint16) if t.dtype == torch.float16 else a


def _flat(t):
    """1-D numpy view of the tensor's MEMORY (dense tensors only), like the kernels see it."""
    t = t.detach()
    return _bits(t.as_strided((t.numel(),), (1,), t.storage_offset()))


def _flag_set(flag):
    return flag is not None and int(flag[0]) != 0


def flag_clear(flag, stream=None):
    flag.zero_()


def mt_check_overflow(table, grad_list, flag, stream=None):
    calls.append("check")
    for i, g in enumerate(table.lists[grad_list]):
        if c_oracle.first_nonfinite(_flat(g)) >= 0:
            flag[0] = max(int(flag[0]), INT32_MAX - i)
            # keep scanning: the word must end as the FIRST bad tensor = the max value
    # (max over all bad tensors of INT32_MAX - i is reached at the smallest i)


def mt_check_overflow_publish(table, grad_list, flag, ticket, host_mirror, rearm, stream=None):
    mt_check_overflow(table, grad_list, flag)
    if host_mirror is not None:
        host_mirror[0] = int(flag[0])
    if rearm is not None:
        rearm.zero_()


def mt_unscale(table, grad_list, inv_scale, flag, mode, stream=None):
    calls.append("unscale")
    first_bad = INT32_MAX
    if _flag_set(flag):
        if mode == 0:
            return
        first_bad = INT32_MAX - int(flag[0])
    for i, g in enumerate(table.lists[grad_list]):
        if i >= first_bad:
            continue
        if g.dtype == torch.float32:
            c_oracle.lib().ora_scale_f32(_flat(g).ctypes.data, g.numel(), float(inv_scale))
        else:
            f = c_oracle.f16_to_f32_scaled(_flat(g), inv_scale, True)
            _flat(g)[:] = c_oracle.f32_to_f16(f)


def _coef(clip):
    return None if clip is None else c_oracle.clip_coef(float(clip[0][0]), clip[1])


def mt_grads_to_master(table, src_list, dst_list, inv_scale, apply_scale, flag, clip=None, stream=None):
    calls.append("grads_to_master")
    if _flag_set(flag):
        return
    for s, d in zip(table.lists[src_list], table.lists[dst_list]):
        _flat(d)[:] = c_oracle.f16_to_f32_scaled(_flat(s), inv_scale, bool(apply_scale), _coef(clip))


def mt_master_to_model(table, src_list, dst_list, flag, stream=None):
    calls.append("master_to_model")
    if _flag_set(flag):
        return
    for s, d in zip(table.lists[src_list], table.lists[dst_list]):
        _flat(d)[:] = c_oracle.f32_to_f16(_flat(s))


def mt_copy(table, src_list, dst_list, elem_size, flag, stream=None):
    calls.append("copy")
    if _flag_set(flag):
        return
    for s, d in zip(table.lists[src_list], table.lists[dst_list]):
        _flat(d)[:] = _flat(s)


def mt_zero(table, lst, elem_size, stream=None):
    calls.append("zero")
    for t in table.lists[lst]:
        _flat(t)[:] = 0


def mt_fused_sgd(table, lr, momentum, dampening, weight_decay, nesterov, inv_scale, apply_scale, flag,
                 clip=None, stream=None):
    calls.append("fused_sgd")
    if _flag_set(flag):
        return
    grads, params, bufs, models = table.lists
    for i in range(len(grads)):
        fl = int(table.flags[i])
        apply = bool(apply_scale) and not (fl & _C.TF_NO_SCALE)
        buf = None if bufs[i] is None else _flat(bufs[i])
        p16 = c_oracle.sgd(_flat(grads[i]), _flat(params[i]), buf, lr, momentum, dampening, weight_decay,
                           nesterov, first=bool(fl & _C.TF_FIRST_STEP), inv_scale=inv_scale, apply_scale=apply,
                           clip_coef=None if (fl & _C.TF_NO_SCALE) else _coef(clip))
        if not (fl & _C.TF_NO_MODEL):
            _flat(models[i])[:] = p16


def mt_fused_adam(table, lr, beta1, beta2, eps, weight_decay, step_size, bias_correction2_sqrt,
                  decoupled_weight_decay, inv_scale, apply_scale, flag, clip=None, stream=None):
    calls.append("fused_adam")
    if _flag_set(flag):
        return
    grads, params, ms, vs, models = table.lists
    lib = c_oracle.lib()
    for i in range(len(grads)):
        fl = int(table.flags[i])
        apply = bool(apply_scale) and not (fl & _C.TF_NO_SCALE)
        g = _flat(grads[i])
        p = _flat(params[i])
        p16 = np.empty(p.shape, np.uint16)
        coef = None if (fl & _C.TF_NO_SCALE) else _coef(clip)
        lib.ora_adam_flavour(g.ctypes.data if g.dtype == np.uint16 else None,
                             g.ctypes.data if g.dtype == np.float32 else None,
                             p.ctypes.data, _flat(ms[i]).ctypes.data, _flat(vs[i]).ctypes.data, p16.ctypes.data,
                             p.size, float(lr), float(beta1), float(beta2), float(eps), float(weight_decay),
                             float(step_size), float(bias_correction2_sqrt), int(bool(decoupled_weight_decay)),
                             float(inv_scale), int(apply), 0, 0, 1.0 if coef is None else coef, int(coef is not None))
        if not (fl & _C.TF_NO_MODEL):
            _flat(models[i])[:] = p16


def mt_flatten(table, src_list, flat, stream=None):
    calls.append("flatten")
    off = 0
    out = _flat(flat)
    for t in table.lists[src_list]:
        n = t.numel()
        out[off:off + n] = _flat(t)
        off += n


def _average(arr, divisor):
    if arr.dtype == np.float64:
        arr *= 1.0 / divisor
    else:
        c_oracle.average(arr, divisor)


def _raise(flag, arr, index):
    if flag is None:
        return
    bad = (not np.isfinite(arr.view(np.float16) if arr.dtype == np.uint16 else arr).all())
    if bad:
        flag[0] = max(int(flag[0]), INT32_MAX - index)


def mt_scale_unflatten(table, flat, dst_list, divisor, apply_mul, flag, stream=None):
    calls.append("scale_unflatten")
    src = _flat(flat)
    off = 0
    for i, t in enumerate(table.lists[dst_list]):
        n = t.numel()
        piece = np.ascontiguousarray(src[off:off + n]).copy()
        if apply_mul:
            _average(piece, divisor)
        _flat(t)[:] = piece
        _raise(flag, piece, i)
        off += n


def scale_flat(flat, divisor, flag, stream=None):
    calls.append("scale_flat")
    arr = _flat(flat)
    piece = np.ascontiguousarray(arr).copy()
    _average(piece, divisor)
    arr[:] = piece
    _raise(flag, piece, 0)


def mt_l2norm(table, grad_list, inv_scale, apply_scale, partials, norm_out, stream=None):
    calls.append("l2norm")
    total = 0.0
    for i, g in enumerate(table.lists[grad_list]):
        apply = bool(apply_scale) and not (int(table.flags[i]) & _C.TF_NO_SCALE)
        total += c_oracle.l2norm([_flat(g)], inv_scale, apply) ** 2
    norm_out[0] = float(np.sqrt(total))


def mt_clip_by_norm(table, grad_list, norm, max_norm, flag, stream=None):
    calls.append("clip")
    if _flag_set(flag):
        return
    coef = np.float32(c_oracle.clip_coef(float(norm[0]), max_norm))
    if coef >= 1.0:
        return
    for g in table.lists[grad_list]:
        _flat(g)[:] = _flat(g) * coef


_real_update = TensorTable.update


def _recording_update(self, lists, flags=None):
    """TensorTable.update that also keeps the tensors themselves, so that the doubles above can walk
    them (the real table only holds addresses)."""
    self.lists = [list(l) for l in lists]
    self.flags = np.zeros(len(lists[0]), dtype=np.int32) if flags is None else np.asarray(flags, dtype=np.int32)
    return _real_update(self, lists, flags)


class _Event(object):
    def record(self, *a, **k):
        pass

    def synchronize(self):
        pass


class _Stream(object):
    def wait_stream(self, other):
        pass

    def wait_event(self, event):
        pass


_real_type = torch.Tensor.type


def _cuda_named_type(self, *a, **k):
    if a or k:
        return _real_type(self, *a, **k)
    name = _real_type(self)
    return name if name.startswith("torch.cuda") else name.replace("torch.", "torch.cuda.")


def install(monkeypatch):
    """Swap the kernel wrappers for the doubles and make torch.cuda plumbing inert."""
    del calls[:]
    monkeypatch.setattr(TensorTable, "update", _recording_update)
    for name in ("flag_clear", "mt_check_overflow", "mt_check_overflow_publish", "mt_unscale", "mt_grads_to_master", "mt_master_to_model",
                 "mt_copy", "mt_zero", "mt_fused_sgd", "mt_fused_adam", "mt_flatten", "mt_scale_unflatten",
                 "scale_flat", "mt_l2norm", "mt_clip_by_norm", "mt_larc"):
        monkeypatch.setattr(_C, name, globals()[name])
    monkeypatch.setattr(torch.cuda, "Event", _Event)
    monkeypatch.setattr(torch.cuda, "Stream", _Stream)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: _Stream())
    monkeypatch.setattr(torch.cuda, "stream", lambda s: contextlib.nullcontext())
    # FP16_Optimizer and DistributedDataParallel gate on 'torch.cuda.*Tensor' strings, as the reference does
    monkeypatch.setattr(torch.Tensor, "type", _cuda_named_type)


def mt_larc(table, grad_list, param_list, trust_coefficient, eps, weight_decay, lr, clip, scratch, norms=None,
            stream=None):
    """fp32 arithmetic in the order of apex_b200/csrc/mt_larc.cu (norms accumulated in double)."""
    calls.append("larc")
    f32 = np.float32
    for g, p in zip(table.lists[grad_list], table.lists[param_list]):
        gf = g.detach().float().numpy().reshape(-1)
        pf = p.detach().float().numpy().reshape(-1)
        gn = f32(np.sqrt(np.sum(gf.astype(np.float64) ** 2)))
        pn = f32(np.sqrt(np.sum(pf.astype(np.float64) ** 2)))
        if pn == 0 or gn == 0:
            continue
        r = f32(f32(trust_coefficient) * pn) / f32(f32(gn + f32(pn * f32(weight_decay))) + f32(eps))
        if clip:
            r = min(f32(r / f32(lr)), f32(1.0))
        if g.dtype == torch.float32:
            out = f32(f32(gf + f32(f32(weight_decay) * pf)) * r)
            _flat(g)[:] = out
        else:
            wp = (f32(weight_decay) * pf).astype(np.float16)
            s = (gf + wp.astype(np.float32)).astype(np.float16)
            _flat(g)[:] = (s.astype(np.float32) * r).astype(np.float16).view(np.uint16)
