"""LARC: layer-wise adaptive rate control (reference: apex/parallel/LARC.py).

Same wrapper class and arguments.  The reference computes two norms per parameter, reads both
back to the host to compare them with zero, and rewrites the gradient with two or three more
launches -- per tensor.  Here one C-ABI call per param group (apex_b200_mt_larc: four launches)
does it for every tensor without any host read, then the wrapped optimizer steps as before.
"""
import torch

from .. import _C
from ..multi_tensor import TensorTable, CHUNK_NORM, same_layout


class LARC(object):
    """Wraps ``optimizer``; before each step every gradient is rescaled by the local rate
    ``trust_coefficient * |p| / (|g| + |p| * weight_decay + eps)`` (``clip=True``: by
    ``min(local_rate / lr, 1)``), with the group's weight decay absorbed into the gradient
    (reference: LARC.py:6-97)."""

    def __init__(self, optimizer, trust_coefficient=0.02, clip=True, eps=1e-8):
        self.param_groups = optimizer.param_groups
        self.optim = optimizer
        self.trust_coefficient = trust_coefficient
        self.eps = eps
        self.clip = clip
        self._tables = {}
        self._scratch = None

    def __getstate__(self):
        return self.optim.__getstate__()

    def __setstate__(self, state):
        self.optim.__setstate__(state)

    def __repr__(self):
        return self.optim.__repr__()

    def state_dict(self):
        return self.optim.state_dict()

    def load_state_dict(self, state_dict):
        self.optim.load_state_dict(state_dict)

    def zero_grad(self):
        self.optim.zero_grad()

    def add_param_group(self, param_group):
        self.optim.add_param_group(param_group)

    def _rescale_group(self, gi, group, weight_decay):
        grads, params, flags = [], [], []
        for p in group['params']:
            g = p.grad
            if g is None:
                continue
            if g.dtype != p.dtype or p.dtype not in (torch.float16, torch.float32):
                raise TypeError("apex_b200 LARC: parameters and gradients must both be fp16 or both fp32")
            if not same_layout((g, p)):
                raise RuntimeError("apex_b200 LARC: parameter and gradient must be dense and share a layout")
            grads.append(g)
            params.append(p.data)
            flags.append(_C.TF_GRAD_F32 if p.dtype == torch.float32 else 0)
        if not grads:
            return
        device = grads[0].device
        table = self._tables.get(gi)
        if table is None or table.device != device:
            table = self._tables[gi] = TensorTable(device, 2, CHUNK_NORM)
        if not table.built_from(grads):
            table.update([grads, params], flags)
            table.remember(grads)
        need = 2 * table.struct.total_chunks + len(grads)
        if self._scratch is None or self._scratch.numel() < need or self._scratch.device != device:
            self._scratch = torch.empty(need, dtype=torch.float32, device=device)
        _C.mt_larc(table, 0, 1, self.trust_coefficient, self.eps, weight_decay, group['lr'], self.clip,
                   self._scratch)

    def step(self):
        with torch.no_grad():
            weight_decays = []
            for gi, group in enumerate(self.optim.param_groups):
                # weight decay is applied here, inside the rescaled gradient, not by the optimizer
                weight_decay = group['weight_decay'] if 'weight_decay' in group else 0
                weight_decays.append(weight_decay)
                group['weight_decay'] = 0
                self._rescale_group(gi, group, weight_decay)
        self.optim.step()
        for group, weight_decay in zip(self.optim.param_groups, weight_decays):
            group['weight_decay'] = weight_decay
