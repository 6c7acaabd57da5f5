"""Master-parameter helpers (reference: apex/fp16_utils/fp16util.py).

``model_grads_to_master_grads`` and ``master_params_to_model_params`` keep the
reference's signatures and semantics; instead of one ``copy_`` per tensor
(fp16util.py:106-112, :127-129) every pair is handled by one multi-tensor
launch per dtype combination.  Model conversion helpers are setup-time code and
are the same few lines of torch as in the reference.
"""
import torch
import torch.nn as nn
from torch._utils import _flatten_dense_tensors

from .. import _C
from ..multi_tensor import TensorTable, CHUNK_STREAM, same_layout


class tofp16(nn.Module):
    """Casts its input to half (reference: fp16util.py:6-18)."""

    def __init__(self):
        super(tofp16, self).__init__()

    def forward(self, input):
        return input.half()


def BN_convert_float(module):
    """Put batch-norm layers back in fp32 after ``.half()`` (reference: fp16util.py:21-34)."""
    if isinstance(module, torch.nn.modules.batchnorm._BatchNorm):
        module.float()
    for child in module.children():
        BN_convert_float(child)
    return module


def network_to_half(network):
    """fp16 network with fp32 batch norm, fed through ``tofp16`` (reference: fp16util.py:37-41)."""
    return nn.Sequential(tofp16(), BN_convert_float(network.half()))


def backwards_debug_hook(grad):
    raise RuntimeError("master_params recieved a gradient in the backward pass!")


def prep_param_lists(model, flat_master=False):
    """Return ``(model_params, master_params)`` (reference: fp16util.py:47-90).

    ``flat_master=True`` gives a single flat fp32 ``nn.Parameter`` holding every
    model parameter in order, with a pre-allocated ``.grad``.
    """
    model_params = [p for p in model.parameters() if p.requires_grad]

    if not flat_master:
        master_params = [p.clone().float().detach() for p in model_params]
        for p in master_params:
            p.requires_grad = True
        return model_params, master_params

    try:
        flat = _flatten_dense_tensors([p.data for p in model_params]).float()
    except:  # noqa: E722  (the reference reports and re-raises whatever happened)
        print("Error in prep_param_lists:  model may contain a mixture of parameters "
              "of different types.  Use flat_master=False, or use F16_Optimizer.")
        raise
    flat = torch.nn.Parameter(flat)
    flat.requires_grad = True
    if flat.grad is None:
        flat.grad = flat.new(*flat.size())
    return model_params, [flat]


# ---- multi-tensor copies ------------------------------------------------------

_tables = {}


def _table(kind, device):
    key = (kind, str(device))
    t = _tables.get(key)
    if t is None:
        t = _tables[key] = TensorTable(device, 2, CHUNK_STREAM)
    return t


def _flat_views(flat, like):
    """Views of 1-D ``flat`` with the sizes of ``like``, in order (no copy)."""
    views, off = [], 0
    for t in like:
        n = t.numel()
        views.append(flat.narrow(0, off, n))
        off += n
    return views


def _convert_pairs(srcs, dsts, flag=None, inv_scale=1.0, apply_scale=False):
    """dst[i] <- src[i] for every pair, grouped by (src dtype, dst dtype)."""
    groups = {}
    for s, d in zip(srcs, dsts):
        if not ((s.is_contiguous() and d.is_contiguous()) or same_layout((s, d))):
            raise RuntimeError("apex_b200: tensors must be contiguous, or dense with one common layout")
        if s.numel() != d.numel():
            raise RuntimeError("apex_b200: size mismatch between model and master tensor")
        groups.setdefault((s.dtype, d.dtype), ([], []))
        groups[(s.dtype, d.dtype)][0].append(s)
        groups[(s.dtype, d.dtype)][1].append(d)
    for (sd, dd), (ss, ds) in groups.items():
        table = _table((sd, dd), ss[0].device)
        table.update([ss, ds])
        if sd == torch.float16 and dd == torch.float32:
            _C.mt_grads_to_master(table, 0, 1, inv_scale, apply_scale, flag)
        elif sd == torch.float32 and dd == torch.float16:
            if apply_scale:
                raise RuntimeError("apex_b200: scaling is only fused into the fp16->fp32 direction")
            _C.mt_master_to_model(table, 0, 1, flag)
        elif sd == dd and sd in (torch.float16, torch.float32, torch.float64):
            if apply_scale:
                raise RuntimeError("apex_b200: scaling is only fused into the fp16->fp32 direction")
            _C.mt_copy(table, 0, 1, ss[0].element_size(), flag)
        else:
            raise TypeError("apex_b200: unsupported copy %s -> %s" % (sd, dd))


def model_grads_to_master_grads(model_params, master_params, flat_master=False):
    """Copy model gradients into the master gradients (reference: fp16util.py:93-112)."""
    if flat_master:
        grads = [p.grad.data for p in model_params]
        _convert_pairs(grads, _flat_views(master_params[0].grad.data, grads))
        return
    srcs, dsts = [], []
    for model, master in zip(model_params, master_params):
        if model.grad is not None:
            if master.grad is None:
                master.grad = master.data.new(*master.data.size())
            srcs.append(model.grad.data)
            dsts.append(master.grad.data)
        else:
            master.grad = None
    if srcs:
        _convert_pairs(srcs, dsts)


def master_params_to_model_params(model_params, master_params, flat_master=False):
    """Copy master parameters into the model parameters (reference: fp16util.py:115-129)."""
    models = [p.data for p in model_params]
    if flat_master:
        masters = _flat_views(master_params[0].data, models)
    else:
        masters = [p.data for p in master_params]
    if models:
        _convert_pairs(masters, models)


def to_python_float(t):
    """reference: fp16util.py:133-137"""
    return t.item() if hasattr(t, "item") else t[0]


# reference: fp16util.py:139-144 picks the in-place name on torch > 0.4
clip_grad_norm = torch.nn.utils.clip_grad_norm_
