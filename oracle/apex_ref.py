"""Restatement of the reference's hot path as plain torch ops (any device).

TEST INFRASTRUCTURE (see oracle/__init__.py).  Every function follows the cited
lines of /root/reference (mbrukman/apex-1).
"""
import math
from collections import OrderedDict

import torch

INF = float("inf")


# ---- overflow detection ---------------------------------------------------------

def has_inf_or_nan(x):
    """apex/fp16_utils/loss_scaler.py:92-110: decide by the fp32 SUM of the tensor."""
    s = float(x.float().sum())
    return s == INF or s == -INF or s != s


def any_overflow(grads):
    """apex/fp16_utils/loss_scaler.py:84-89: first bad gradient ends the scan."""
    for g in grads:
        if g is not None and has_inf_or_nan(g):
            return True
    return False


# ---- loss scale trajectories ------------------------------------------------------

class DynamicScale(object):
    """apex/fp16_utils/loss_scaler.py:73-121"""

    def __init__(self, init_scale=2**32, scale_factor=2., scale_window=1000):
        self.cur_scale = init_scale
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor = scale_factor
        self.scale_window = scale_window

    def update(self, overflow):
        if overflow:
            self.cur_scale = max(self.cur_scale / self.scale_factor, 1)
            self.last_overflow_iter = self.cur_iter
        elif (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0:
            self.cur_scale *= self.scale_factor
        self.cur_iter += 1


class StaticScale(object):
    """apex/fp16_utils/loss_scaler.py:23-38: fixed value, never overflows."""

    def __init__(self, scale=1):
        self.cur_scale = scale

    def update(self, overflow):
        pass


class AmpScale(object):
    """apex/amp/scaler.py:20-55"""

    def __init__(self):
        self.loss_scale = 2.**16
        self.max_loss_scale = 2.**24
        self.scale_seq_len = 2000
        self.unskipped = 0

    def unscale_and_update(self, grads, scale):
        """``grads`` in iter_params order.  Returns should_skip; mutates grads like :35-40."""
        overflow = False
        inv = 1. / scale
        for g in grads:
            if g is None:
                continue
            if has_inf_or_nan(g):          # scale_check_overflow, :6-18
                overflow = True
                break
            g.mul_(inv)
        if overflow:
            self.loss_scale /= 2.
            self.unskipped = 0
        else:
            self.unskipped += 1
        if self.unskipped == self.scale_seq_len:
            self.loss_scale = min(self.max_loss_scale, self.loss_scale * 2.)
            self.unskipped = 0
        return overflow


# ---- the mixed-precision step --------------------------------------------------------

class MixedPrecisionStep(object):
    """update_master_grads() + step() of FP16_Optimizer, restated.

    apex/fp16_utils/fp16_optimizer.py:105-175 (master creation), :213-241, :330-384,
    :486-497; apex/fp16_utils/fp16util.py:93-129.

    ``params``: list of leaf tensors (fp16 or fp32) whose ``.grad`` the caller fills.
    ``make_optimizer(list_of_fp32_leaves) -> torch.optim.Optimizer``.
    """

    def __init__(self, params, make_optimizer, static_loss_scale=1.0, dynamic_loss_scale=False,
                 dynamic_loss_args=None):
        self.fp16_params, self.masters, self.fp32_params = [], [], []
        owned = []
        for p in params:
            if p.dtype == torch.float16:
                m = p.detach().clone().float()          # :135
                m.requires_grad = True
                self.fp16_params.append(p)
                self.masters.append(m)
                owned.append(m)
            elif p.dtype == torch.float32:
                self.fp32_params.append(p)
                owned.append(p)
            else:
                raise TypeError("fp16 or fp32 only")     # :148-151
        self.optimizer = make_optimizer(owned)
        self.dynamic = bool(dynamic_loss_scale)
        if self.dynamic:
            self.scaler = DynamicScale(**(dynamic_loss_args or {}))
        else:
            self.scaler = StaticScale(static_loss_scale)
        self.overflow = False
        self.messages = []

    @property
    def loss_scale(self):
        return self.scaler.cur_scale

    def update_master_grads(self):
        if self.dynamic:                                                    # :493-495
            grads = [p.grad for p in self.fp16_params] + [p.grad for p in self.fp32_params]   # :213-221
            self.overflow = any_overflow([None if g is None else g.data for g in grads])
            if self.overflow:
                return
        for model, master in zip(self.fp16_params, self.masters):          # fp16util.py:106-112
            if model.grad is not None:
                if master.grad is None:
                    master.grad = torch.empty_like(master.data)
                master.grad.data.copy_(model.grad.data)
            else:
                master.grad = None
        if self.loss_scale != 1.0:                                          # :236-241
            inv = 1. / self.loss_scale
            for group in self.optimizer.param_groups:
                for p in group['params']:
                    if p.grad is not None:
                        p.grad.data.mul_(inv)

    def clip_master_grads(self, max_norm, norm_type=2):
        """:243-265"""
        if self.overflow:
            return -1
        ps = [p for g in self.optimizer.param_groups for p in g['params']]
        return torch.nn.utils.clip_grad_norm_(ps, max_norm, norm_type)

    def step(self):
        scale = self.scaler.cur_scale                                       # :369-370
        self.scaler.update(self.overflow)
        if self.overflow:                                                   # :372-375
            self.messages.append("OVERFLOW! Skipping step. Attempted loss scale: {}, reducing to {}"
                                 .format(scale, self.loss_scale))
            return
        self.optimizer.step()                                               # :380
        for model, master in zip(self.fp16_params, self.masters):          # fp16util.py:127-129
            model.data.copy_(master.data)

    def zero_grad(self):
        """:187-211 (set_grads_to_None=False)"""
        for group in self.optimizer.param_groups:
            for p in group['params']:
                if p.grad is not None:
                    p.grad.detach_()
                    p.grad.zero_()
        for p in self.fp16_params:
            if p.grad is not None:
                p.grad.detach_()
                p.grad.zero_()


# ---- functional helpers -----------------------------------------------------------------

def grads_to_master(model_grads, flat=False):
    """apex/fp16_utils/fp16util.py:93-112 (values only): fp32 copies, or one flat fp32 tensor."""
    if flat:
        return torch.cat([g.reshape(-1) for g in model_grads]).float()
    return [g.float() for g in model_grads]


def master_to_model(masters, like, flat=False):
    """apex/fp16_utils/fp16util.py:115-129 (values only): masters cast to each model dtype."""
    if flat:
        out, off = [], 0
        for t in like:
            n = t.numel()
            out.append(masters[off:off + n].view_as(t).to(t.dtype))
            off += n
        return out
    return [m.to(t.dtype) for m, t in zip(masters, like)]


# ---- bucket path ---------------------------------------------------------------------------

def flat_call(bucket, call, world_size, is_allreduce, extra_args=()):
    """apex/parallel/distributed.py:11-23 on one same-dtype bucket (in place)."""
    flat = torch.cat([t.reshape(-1) for t in bucket])
    call(flat, *extra_args)
    if is_allreduce:
        flat /= world_size
    off = 0
    for t in bucket:
        n = t.numel()
        t.copy_(flat[off:off + n].view_as(t))
        off += n


def grouped_by_type(tensors):
    """apex/parallel/distributed.py:27-38: first-seen order of tensor.type()."""
    groups = OrderedDict()
    for t in tensors:
        groups.setdefault(t.type(), []).append(t)
    return list(groups.values())


_CLASS = {torch.float16: 0, torch.float32: 1, torch.float64: 2}


def bucket_layout(arrival, numels, dtypes, message_size):
    """Bucket structure learned during the first backward.

    apex/parallel/distributed.py:248-261 (running bucket per dtype class, cut when
    numel >= message_size) and :176-180 (leftovers appended half, float, double).
    ``arrival``: parameter indices in the order their gradient hooks fire.
    Returns (buckets, where) with where[param_index] = (bucket_idx, bucket_loc).
    """
    buckets, running, count = [], [[], [], []], [0, 0, 0]
    for i in arrival:
        c = _CLASS[dtypes[i]]
        running[c].append(i)
        count[c] += numels[i]
        if count[c] >= message_size:
            buckets.append(running[c])
            running[c], count[c] = [], 0
    for r in running:
        if r:
            buckets.append(r)
    where = {}
    for b, bucket in enumerate(buckets):
        for loc, i in enumerate(bucket):
            where[i] = (b, loc)
    return buckets, where


def dispatch_order(ready_sequence, sizes):
    """Order in which full buckets are reduced given the order slots fill up.

    apex/parallel/distributed.py:279-315: a bucket goes out only when it is full AND
    it is the next index; later full buckets wait in a set and drain in sorted order.
    ``ready_sequence``: bucket index of each arriving gradient.  Returns the list of
    bucket indices in issue order.
    """
    filled = [0] * len(sizes)
    nxt, parked, issued = 0, set(), []
    for b in ready_sequence:
        filled[b] += 1
        if filled[b] == sizes[b]:
            if b == nxt:
                issued.append(b)
                nxt += 1
                for i in sorted(parked):
                    if i > nxt:
                        break
                    issued.append(i)
                    parked.remove(i)
                    nxt += 1
            else:
                parked.add(b)
    return issued


# ---- LARC -------------------------------------------------------------------------------------

def larc_rescale(params, lr, weight_decay, trust_coefficient=0.02, clip=True, eps=1e-8):
    """The gradient rewrite of LARC.step for one param group (apex/parallel/LARC.py:75-92), in place.
    ``params``: tensors with ``.grad``."""
    with torch.no_grad():
        for p in params:
            if p.grad is None:
                continue
            param_norm = torch.norm(p.data)
            grad_norm = torch.norm(p.grad.data)
            if param_norm != 0 and grad_norm != 0:
                adaptive_lr = trust_coefficient * (param_norm) / (grad_norm + param_norm * weight_decay + eps)
                if clip:
                    adaptive_lr = min(adaptive_lr / lr, 1)
                p.grad.data += weight_decay * p.data
                p.grad.data *= adaptive_lr
