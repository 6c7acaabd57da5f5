"""FP16_Optimizer on B200 kernels (reference: apex/fp16_utils/fp16_optimizer.py).

Public surface, attribute names, messages and return values are the
reference's.  What is different is what runs on the GPU:

reference, per iteration (fp16_optimizer.py:486-497, :330-384)
    has_overflow      2 launches + 1 blocking D2H read PER TENSOR   (loss_scaler.py:84-110)
    grads -> masters  1 copy_ per tensor                             (fp16util.py:106-112)
    downscale         1 mul_ per tensor                              (:236-241)
    optimizer.step()  torch foreach, 3-5 passes over every array     (:380)
    masters -> model  1 copy_ per tensor                             (fp16util.py:127-129)

here
    update_master_grads   ONE check launch over all gradients -> one device word,
                          copied to pinned host memory asynchronously.  The fp16
                          gradients are NOT expanded to fp32 yet: the master
                          gradients stay "virtual" (= float(g16) * 1/scale).
    step                  ONE fused launch per param group: unscale -> update of
                          master + state -> fp16 model copy, predicated on the
                          device word (a skipped step touches nothing).  Only
                          then does the host wait for the word, to drive
                          update_scale / the OVERFLOW message exactly as the
                          reference does.

Anything that looks at a master gradient in between (``clip_master_grads``, an
inner optimizer other than SGD/Adam/AdamW, a closure, user code reading
``master.grad``) makes the virtual gradients real first -- masters are a Tensor
subclass whose ``grad`` property does that -- and from there on the code follows
the reference's unfused sequence on multi-tensor kernels.  DESIGN.md lists the
(few, deliberate) observable differences.
"""
import operator
import weakref

import torch

from .. import _C
from ..multi_tensor import (TensorTable, CHUNK_STEP, CHUNK_STREAM, CHUNK_CHECK, CHUNK_ZERO, CHUNK_NORM,
                            new_flag, same_objects, is_dense, same_layout)
from .loss_scaler import DynamicLossScaler, LossScaler
from .fp16util import master_params_to_model_params, clip_grad_norm

_tensor_grad = torch._C.TensorBase.grad


class _MasterParam(torch.Tensor):
    """fp32 master copy of an fp16 parameter.

    Behaves like the plain fp32 leaf tensor the reference creates
    (fp16_optimizer.py:135-136).  Reading ``.grad`` first asks the owning
    FP16_Optimizer to materialise pending master gradients.
    """
    __torch_function__ = torch._C._disabled_torch_function_impl

    @property
    def grad(self):
        owner = self.__dict__.get("_apex_owner")
        if owner is not None:
            owner = owner()
            if owner is not None and owner._pending is not None:
                owner._materialize_master_grads()
        return _tensor_grad.__get__(self)

    @grad.setter
    def grad(self, value):
        _tensor_grad.__set__(self, value)

    @grad.deleter
    def grad(self):
        _tensor_grad.__delete__(self)

    def __reduce_ex__(self, proto):
        # checkpoints hold plain fp32 tensors, exactly what the reference saves
        # under 'fp32_from_fp16' (fp16_optimizer.py:284)
        plain = self.detach().requires_grad_(self.requires_grad)
        return plain.__reduce_ex__(proto)


# torch.optim picks its multi-tensor ("foreach") implementation only for tensor types it knows; the
# list is open for registration (torch/optim/optimizer.py: _foreach_supported_types).  Without this
# an inner optimizer stepping our masters itself (no fused kernel for it) would silently take the
# slower single-tensor loop, whose roundings differ from the foreach kernels the reference gets with
# its plain fp32 masters.
try:
    from torch.optim.optimizer import _foreach_supported_types as _foreach_types
    if _MasterParam not in _foreach_types:
        _foreach_types.append(_MasterParam)
except ImportError:          # older torch: no foreach selection by type
    pass


def _raw_grad(t):
    """``t.grad`` without triggering materialisation."""
    return _tensor_grad.__get__(t)


class _StepPlan(object):
    """Everything one fused launch of one param group needs, remembered between steps.

    Rebuilding it costs a python loop over every tensor (addresses, dtypes, state lookups); checking
    that it is still valid costs identity comparisons only.  It is valid while the group still holds
    the same parameter objects, the gradients are the same tensor objects, and optimizer.state still
    holds the same state tensors.
    """

    def __init__(self):
        self.kind = None
        self.from_masters = False
        self.owners = None       # keys of optimizer.state: masters then fp32-native params
        self.group_params = None # the group's parameter list as it was when the plan was built
        self.grads = None        # only used while building
        self.table = None
        self.state_tensors = None   # per owner: tuple of state tensors the kernel writes
        self.state_keys = ()
        self.fresh = None        # state created by us, published on the first committed step
        self.steps = None        # adam: list of per-parameter step tensors
        self.step_buf = None     # adam: the buffer our own step tensors are views of
        self.momentum = 0.0
        self.state_map = None
        self.state_len = 0
        self.state_dicts = None
        self._held_columns = ()
        self._watch = ()
        self._watch_ptrs = ()
        self._cursor = 0

    @staticmethod
    def build(opt, kind, gi, group, from_masters, halves, masters, floats, grads):
        want16 = torch.float32 if from_masters else torch.float16
        flag16 = (_C.TF_GRAD_F32 | _C.TF_NO_SCALE) if from_masters else 0
        if len(halves) + len(floats) != len(group['params']):
            return False          # param_groups edited behind our back
        params, models, flags = [], [], []
        n16 = len(halves)
        for i, g in enumerate(grads):
            want = want16 if i < n16 else torch.float32
            if g is None or g.dtype != want or g.is_sparse or not is_dense(g):
                return False
        for model, master in zip(halves, masters):
            params.append(master.data)
            models.append(model.data)
            flags.append(flag16)
        for p in floats:
            params.append(p.data)
            models.append(None)
            # already unscaled in place by update_master_grads, as the reference does
            flags.append(_C.TF_GRAD_F32 | _C.TF_NO_MODEL | _C.TF_NO_SCALE)
        # elementwise over flat memory: gradient, master and model copy must share one layout
        for g, pm, md in zip(grads, params, models):
            if not same_layout((g, pm, md)):
                return False
        opt._ensure_device(grads[0])
        plan = _StepPlan()
        plan.kind, plan.from_masters = kind, from_masters
        plan.owners = list(masters) + list(floats)
        plan.group_params = list(group['params'])
        state = opt.optimizer.state
        if kind == "sgd":
            lists = plan._build_sgd(group, state, grads, params, models, flags)
        else:
            lists = plan._build_adam(state, grads, params, models, flags)
        if lists is False:
            return False
        table = opt._step_tables.get(gi)
        if table is None or table.n_lists != len(lists):
            table = opt._step_tables[gi] = TensorTable(opt._device, len(lists), CHUNK_STEP)
        table.update(lists, flags)
        table.remember(grads)
        plan.table = table
        # objects whose storage could be swapped under an unchanged identity (`p.data = ...`,
        # `module.half()`): masters/params, fp16 model params, state tensors.  still_valid() compares
        # the addresses of the first, the last and one rotating entry of each list per step.
        plan._watch = [plan.owners, list(halves) + [None] * len(floats)] + [list(l) for l in lists[2:-1]]
        plan._watch_ptrs = [[0 if o is None else o.data_ptr() for o in objs] for objs in plan._watch]
        plan._cursor = 0
        if not plan.fresh:
            plan._bind_state(state)
        return plan

    def _build_sgd(self, group, state, grads, params, models, flags):
        self.momentum = group['momentum']
        bufs, fresh, held = [], [], []
        if self.momentum != 0:
            self.state_keys = ('momentum_buffer',)
            for i, p in enumerate(self.owners):
                buf = state[p].get('momentum_buffer') if p in state else None
                if buf is None:
                    # torch: buf = clone(grad) on the first step (sgd.py:361-362, :452-455)
                    buf = torch.empty_like(params[i])
                    fresh.append((p, {'momentum_buffer': buf}))
                    flags[i] |= _C.TF_FIRST_STEP
                elif buf.dtype != torch.float32 or not same_layout((buf, params[i])):
                    return False
                bufs.append(buf)
                held.append((buf,))
        else:
            bufs = [None] * len(self.owners)
        self.fresh = fresh
        self.state_tensors = held
        return [grads, params, bufs, models]

    def _build_adam(self, state, grads, params, models, flags):
        self.state_keys = ('exp_avg', 'exp_avg_sq')
        ms, vs, fresh, steps, held = [], [], [], [], []
        missing = [p for p in self.owners if p not in state or len(state[p]) == 0]
        if missing:
            # torch's lazy state init (adam.py:_init_group).  Our step counters are views of one
            # buffer so that a single add advances all of them.
            step_dtype = torch.float64 if torch.get_default_dtype() == torch.float64 else torch.float32
            self.step_buf = torch.zeros(len(missing), dtype=step_dtype)
        k = 0
        for i, p in enumerate(self.owners):
            st = state[p] if p in state else {}
            if len(st) == 0:
                st = {'step': self.step_buf[k],
                      'exp_avg': torch.zeros_like(params[i], memory_format=torch.preserve_format),
                      'exp_avg_sq': torch.zeros_like(params[i], memory_format=torch.preserve_format)}
                k += 1
                fresh.append((p, st))
            if not isinstance(st['step'], torch.Tensor) or st['step'].is_cuda:
                return False
            m, v = st['exp_avg'], st['exp_avg_sq']
            if m.dtype != torch.float32 or v.dtype != torch.float32 or not same_layout((m, v, params[i])):
                return False
            ms.append(m)
            vs.append(v)
            steps.append(st['step'])
            held.append((m, v))
        if len(set(float(s) for s in steps)) != 1:
            return False          # per-tensor step counts differ: let torch handle it
        self.fresh = fresh
        self.steps = steps
        self.state_tensors = held
        return [grads, params, ms, vs, models]

    def step_count(self):
        return float(self.steps[0])

    def _bind_state(self, state):
        """Remember the state mapping and the per-parameter dicts (only once they all exist)."""
        self.state_map = state
        self.state_len = len(state)
        self.state_dicts = [state[p] for p in self.owners] if self.state_keys else []
        self._held_columns = [[held[k] for held in self.state_tensors] for k in range(len(self.state_keys))]

    def still_valid(self, kind, from_masters, group, grads, state):
        if kind != self.kind or from_masters != self.from_masters or self.fresh:
            return False          # (a plan that still carries unpublished state is rebuilt)
        if not same_objects(group['params'], self.group_params):
            return False
        if kind == "sgd" and group['momentum'] != self.momentum:
            return False
        if not self.table.built_from(grads):
            return False
        # optimizer state: torch replaces the whole mapping on load_state_dict; inside it, the
        # per-parameter dicts we saw at build time must still hold the very same tensors
        if state is not self.state_map or len(state) != self.state_len:
            return False
        for k, key in enumerate(self.state_keys):
            held = self._held_columns[k]
            try:
                now = list(map(operator.itemgetter(key), self.state_dicts))
            except KeyError:
                return False
            if not all(map(operator.is_, now, held)):
                return False
        if self.steps is not None:
            if self.state_dicts[0].get('step') is not self.steps[0] or \
                    self.state_dicts[-1].get('step') is not self.steps[-1]:
                return False
        n = len(self.owners)
        k = self._cursor = (self._cursor + 1) % n
        for objs, ptrs in zip(self._watch, self._watch_ptrs):
            for i in (0, n - 1, k):
                o = objs[i]
                if o is not None and o.data_ptr() != ptrs[i]:
                    return False      # same object, new storage: rebuild (update() compares every address)
        return True

    def commit(self, state):
        """The step was not skipped: publish state created for it and advance Adam's counters."""
        fresh, self.fresh = self.fresh, []
        for p, new in fresh:
            state[p].update(new)
        if self.state_keys and not fresh and self.state_dicts is None:
            self._bind_state(state)
        if self.steps is not None:
            if self.step_buf is not None and len(self.steps) == self.step_buf.numel():
                self.step_buf += 1
            else:
                torch._foreach_add_(self.steps, 1)
        if fresh:
            self.table.invalidate()       # the FIRST_STEP flags must go: rebuild next time
            self.kind = None


class FP16_Optimizer(object):
    """Wraps ``init_optimizer`` and manages fp32 master weights and loss scaling.

    Same constructor and methods as the reference class
    (fp16_optimizer.py:11-551); see the module docstring for what runs where.
    """

    def __init__(self,
                 init_optimizer,
                 static_loss_scale=1.0,
                 dynamic_loss_scale=False,
                 dynamic_loss_args=None,
                 verbose=True):
        self.verbose = verbose
        self.optimizer = init_optimizer

        self.fp16_groups = []
        self.fp32_from_fp16_groups = []
        self.fp32_from_fp32_groups = []
        self._pending = None          # (inv_scale, apply_scale): master grads are float(g16)*inv_scale, not stored
        self._grads_ready = False     # master grads are materialised and current
        self._clip = None             # (device norm, max_norm): clipping still to be applied to virtual grads
        self_ref = weakref.ref(self)
        for gi, param_group in enumerate(self.optimizer.param_groups):
            self.maybe_print("FP16_Optimizer processing param group {}:".format(gi))
            halves, masters, floats = [], [], []
            for pi, param in enumerate(param_group['params']):
                if not param.requires_grad:
                    continue
                kind = param.type()
                if kind == 'torch.cuda.HalfTensor':
                    self.maybe_print("FP16_Optimizer received torch.cuda.HalfTensor with {}"
                                     .format(param.size()))
                    halves.append(param)
                    master = param.detach().clone().float().as_subclass(_MasterParam)
                    master.requires_grad = True
                    master.__dict__["_apex_owner"] = self_ref
                    param_group['params'][pi] = master
                    masters.append(master)
                    # keep any state the wrapped optimizer already had for this param
                    if param in self.optimizer.state:
                        self.optimizer.state[master] = self.optimizer.state.pop(param)
                elif kind == 'torch.cuda.FloatTensor':
                    self.maybe_print("FP16_Optimizer received torch.cuda.FloatTensor with {}"
                                     .format(param.size()))
                    floats.append(param)
                    param_group['params'][pi] = param
                else:
                    raise TypeError("Wrapped parameters must be either "
                                    "torch.cuda.FloatTensor or torch.cuda.HalfTensor. "
                                    "Received {}".format(kind))
            self.fp16_groups.append(halves)
            self.fp32_from_fp16_groups.append(masters)
            self.fp32_from_fp32_groups.append(floats)

        # flat views of the three list-of-lists, for the per-step walks (one comprehension each)
        self._flat_halves = [p for g in self.fp16_groups for p in g]
        self._flat_floats = [p for g in self.fp32_from_fp32_groups for p in g]
        self._spans = []
        a = b = 0
        for halves, floats in zip(self.fp16_groups, self.fp32_from_fp32_groups):
            self._spans.append((a, a + len(halves), b, b + len(floats)))
            a += len(halves)
            b += len(floats)
        self._g16 = self._g32 = None      # this iteration's gradients, read once in update_master_grads

        # recast pre-existing per-param state to the masters' dtype (reference :158)
        self.optimizer.load_state_dict(self.optimizer.state_dict())

        if dynamic_loss_scale:
            self.dynamic_loss_scale = True
            if dynamic_loss_args is not None:
                self.loss_scaler = DynamicLossScaler(**dynamic_loss_args)
            else:
                self.loss_scaler = DynamicLossScaler()
        else:
            self.dynamic_loss_scale = False
            self.loss_scaler = LossScaler(static_loss_scale)

        self._overflow = False
        self._overflow_event = None      # set while the device word has not been read yet
        self.first_closure_call_this_step = True
        self.clip_grad_norm = clip_grad_norm
        # clip_master_grads with the 2-norm: True = norm straight from the fp16 gradients and the
        # factor applied inside the fused update (a TOLERANCE path: the norm is summed in another order
        # than torch's, so the factor may differ in its last bits); False = materialise the master
        # gradients and call torch's clip_grad_norm_ as the reference does (bit-exact, ~3x the traffic)
        self.fused_clip = True
        # how many step() calls went through the wrapped optimizer's own step() + a copy-back launch
        # because no fused kernel applies (inner optimizer other than SGD/Adam/AdamW, amsgrad, ...)
        self.unfused_steps = 0

        # device-side plumbing (created on first use)
        self._device = None
        self._flag = None                # the single overflow word
        self._flag_src = None            # the word the pending decision lives in (ours, or DDP's)
        self._flag_host = None           # its pinned host mirror
        self._check_table = None         # all gradients, for the overflow scan
        self._unscale_table = None       # fp32-native gradients, unscaled in place
        self._grad_tables = {}           # group index -> table for materialisation
        self._step_tables = {}           # group index -> table for the fused step
        self._plans = {}                 # group index -> cached _StepPlan
        self._zero_table = None
        self._norm_buf = None            # [norm, per-chunk partial sums...] scratch of the gradient norm
        self._norm_table = None
        self._clip_table = None

    # ---- small helpers -------------------------------------------------------

    def maybe_print(self, msg):
        if self.verbose:
            print(msg)

    def __getstate__(self):
        raise RuntimeError("FP16_Optimizer should be serialized using state_dict().")

    def __setstate__(self, state):
        raise RuntimeError("FP16_Optimizer should be deserialized using load_state_dict().")

    def _ensure_device(self, tensor):
        if self._device is None:
            self._device = tensor.device
            # two overflow words used alternately (each check re-arms the other one: no memset on the
            # stream) and the ticket counter of the publishing check kernel
            words = torch.zeros(4, dtype=torch.int32, device=self._device)
            self._flag_pair = (words[0:1], words[1:2])
            self._flag_ticket = words[2:3]
            self._flag_turn = 0
            self._flag = self._flag_pair[0]
            self._flag_host = torch.zeros(1, dtype=torch.int32,
                                          pin_memory=(self._device.type == "cuda"))

    # ``overflow`` is a plain bool in the reference.  Here the truth may still
    # be on the device; reading the attribute waits for it (once).
    def _get_overflow(self):
        ev = self._overflow_event
        if ev is not None:
            ev.synchronize()
            self._overflow = int(self._flag_host[0]) != 0
            self._overflow_event = None
        return self._overflow

    def _set_overflow(self, value):
        self._overflow_event = None
        self._overflow = value

    overflow = property(_get_overflow, _set_overflow)

    def _live_flag(self):
        """Device word that predicates kernels, or None when no check is pending/needed."""
        return self._flag_src if self._overflow_event is not None else None

    # ---- zero_grad -------------------------------------------------------------

    def zero_grad(self, set_grads_to_None=False):
        """Zero (or drop) fp32 and fp16 gradients (reference: :187-211), in one launch."""
        self._pending = None
        self._grads_ready = False
        self._clip = None
        self._g16 = self._g32 = None
        if set_grads_to_None:
            for group in self.optimizer.param_groups:
                for p in group['params']:
                    p.grad = None
            for fp16_group in self.fp16_groups:
                for param in fp16_group:
                    param.grad = None
            return
        grads = []
        for group in self.optimizer.param_groups:
            for p in group['params']:
                g = _raw_grad(p)
                if g is not None:
                    g.detach_()
                    grads.append(g)
        for fp16_group in self.fp16_groups:
            for param in fp16_group:
                if param.grad is not None:
                    param.grad.detach_()
                    grads.append(param.grad)
        if not grads:
            return
        self._ensure_device(grads[0])
        table = self._zero_table
        if table is None:
            table = self._zero_table = TensorTable(self._device, 1, CHUNK_ZERO)
        if not table.built_from(grads):
            flags = []
            for g in grads:
                if not is_dense(g):
                    raise RuntimeError("apex_b200: gradients must be dense (contiguous or channels-last)")
                if g.dtype == torch.float32:
                    flags.append(_C.TF_GRAD_F32)
                elif g.dtype == torch.float16:
                    flags.append(0)
                else:
                    raise TypeError("apex_b200: unexpected gradient dtype %s" % g.dtype)
            table.update([grads], flags)
            table.remember(grads)
        _C.mt_zero(table, 0, 2)

    # ---- overflow check ----------------------------------------------------------

    def _check_overflow(self):
        """Scan fp16 grads then fp32-native grads (reference order, :213-221); no host sync."""
        if not isinstance(self.loss_scaler, DynamicLossScaler):
            # a user-supplied scaler object: ask it, as the reference does
            params = [p for g in self.fp16_groups for p in g] + \
                     [p for g in self.fp32_from_fp32_groups for p in g]
            self.overflow = self.loss_scaler.has_overflow(params)
            return
        # fp16 gradients, then the fp32-native ones (the reference's order, :214-220)
        grads = [g for g in self._g16 + self._g32 if g is not None]
        if not grads:
            self.overflow = False
            return
        self._ensure_device(grads[0])
        # Multi-GPU: DistributedDataParallel's allreduce kernel has already looked at every averaged
        # value it wrote; if that covers all of our gradients and nothing touched them since, its
        # flag word IS the answer and no extra pass over the gradients is needed.
        from ..parallel.distributed import verified_flag
        ready = verified_flag(grads)
        if ready is not None and ready.device == self._device:
            self._flag_src = ready
            self._flag_host.copy_(ready, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            self._overflow_event = ev
            return
        table = self._check_table
        if table is None:
            table = self._check_table = TensorTable(self._device, 1, CHUNK_CHECK)
        if not table.built_from(grads):
            flags = []
            for g in grads:
                if not is_dense(g):
                    raise RuntimeError("apex_b200: gradients must be dense (contiguous or channels-last)")
                if g.dtype == torch.float16:
                    flags.append(0)
                elif g.dtype == torch.float32:
                    flags.append(_C.TF_GRAD_F32)
                else:
                    raise TypeError("apex_b200: unexpected gradient dtype %s" % g.dtype)
            table.update([grads], flags)
            table.remember(grads)
        # One launch and nothing else on the stream: the kernel's last CTA writes the verdict into
        # pinned host memory and zeroes the word the next check will use.  (The previous verdict has
        # been read by the host -- step() does -- or is being discarded, before this launch re-arms.)
        turn = self._flag_turn = self._flag_turn ^ 1
        self._flag = self._flag_src = self._flag_pair[turn]
        _C.mt_check_overflow_publish(table, 0, self._flag, self._flag_ticket, self._flag_host,
                                     self._flag_pair[turn ^ 1])
        ev = torch.cuda.Event()
        ev.record()
        self._overflow_event = ev

    def _update_scale(self, has_overflow=False):
        self.loss_scaler.update_scale(has_overflow)

    # ---- master <-> model ------------------------------------------------------------

    def _master_params_to_model_params(self):
        for fp16_group, fp32_from_fp16_group in zip(self.fp16_groups, self.fp32_from_fp16_groups):
            if fp16_group:
                master_params_to_model_params(fp16_group, fp32_from_fp16_group)

    def _materialize_master_grads(self):
        """Make the virtual master gradients real: the reference's
        ``_model_grads_to_master_grads`` + ``_downscale_master`` for the masters
        (:232-241), including its early return on overflow (:493-495)."""
        pending, self._pending = self._pending, None
        clip, self._clip = self._clip, None
        if pending is None:
            return
        inv_scale, apply_scale = pending
        if self.overflow:      # waits for the device word; master grads stay as they were
            return
        self._grads_ready = True
        for gi, (fp16_group, masters) in enumerate(zip(self.fp16_groups, self.fp32_from_fp16_groups)):
            srcs, dsts = [], []
            for model, master in zip(fp16_group, masters):
                if model.grad is not None:
                    mg = _raw_grad(master)
                    if mg is None or not same_layout((mg, model.grad)):
                        mg = torch.empty_like(model.grad, dtype=torch.float32)     # the gradient's own layout
                        master.grad = mg
                    srcs.append(model.grad.data)
                    dsts.append(mg.data)
                else:
                    master.grad = None
            if not srcs:
                continue
            self._ensure_device(srcs[0])
            table = self._grad_tables.get(gi)
            if table is None:
                table = self._grad_tables[gi] = TensorTable(self._device, 2, CHUNK_STREAM)
            table.update([srcs, dsts])
            _C.mt_grads_to_master(table, 0, 1, inv_scale, apply_scale, None, clip)

    def _model_grads_to_master_grads(self):
        # kept for API parity with the reference (:232-234); scale 1 = plain copy
        self._pending = (1.0, False)
        self._materialize_master_grads()

    def _downscale_master(self):
        """In-place unscale of every fp32 gradient the inner optimizer owns (reference :236-241)."""
        if self.loss_scale != 1.0:
            grads = []
            for group in self.optimizer.param_groups:
                for param in group['params']:
                    if param.grad is not None:
                        grads.append(param.grad)
            self._unscale_in_place(grads, 1. / self.loss_scale, None)

    def _unscale_in_place(self, grads, inv_scale, flag):
        if not grads:
            return
        self._ensure_device(grads[0])
        table = self._unscale_table
        if table is None:
            table = self._unscale_table = TensorTable(self._device, 1, CHUNK_STREAM)
        if not table.built_from(grads):
            for g in grads:
                if g.dtype != torch.float32 or not is_dense(g):
                    raise RuntimeError("apex_b200: fp32 gradients must be dense float tensors")
            table.update([grads], [_C.TF_GRAD_F32] * len(grads))
            table.remember(grads)
        _C.mt_unscale(table, 0, inv_scale, flag, 0)

    # ---- clip ---------------------------------------------------------------------------

    def clip_master_grads(self, max_norm, norm_type=2):
        """Clip the fp32 master gradients (reference: :243-265).

        Returns the total norm, or -1 if the last gradients overflowed.  For the 2-norm the
        norm is computed straight from the fp16 gradients (one read of 2 B/param) and the
        clipping factor is applied inside the fused update: the master gradients are still
        not written unless somebody looks at them.
        """
        if self.overflow:
            return -1
        if self.fused_clip and self._pending is not None and self._clip is None and norm_type in (2, 2.0) \
                and self.clip_grad_norm is clip_grad_norm and self._groups_intact():
            return self._clip_virtual(max_norm)
        fp32_params = []
        for param_group in self.optimizer.param_groups:
            for param in param_group['params']:
                fp32_params.append(param)
        return self.clip_grad_norm(fp32_params, max_norm, norm_type)

    def _groups_intact(self):
        """The inner optimizer's param groups still hold exactly the parameters seen at construction."""
        groups = self.optimizer.param_groups
        if len(groups) != len(self.fp16_groups) or len(groups) != len(self._spans):
            return False
        for g, h, f, (a, b, c, d) in zip(groups, self.fp16_groups, self.fp32_from_fp32_groups, self._spans):
            if len(g['params']) != len(h) + len(f) or b - a != len(h) or d - c != len(f):
                return False
        return True

    def _clip_virtual(self, max_norm):
        inv_scale, apply_scale = self._pending
        grads, flags, floats = [], [], []
        for group in self.fp16_groups:
            for param in group:
                g = param.grad
                if g is not None:
                    grads.append(g)
                    flags.append(0)
        for group in self.fp32_from_fp32_groups:
            for param in group:
                g = param.grad
                if g is not None:
                    grads.append(g)
                    floats.append(g)
                    flags.append(_C.TF_GRAD_F32 | _C.TF_NO_SCALE)     # unscaled in place already
        if not grads:
            return torch.zeros(())
        self._ensure_device(grads[0])
        table = self._norm_table
        if table is None:
            table = self._norm_table = TensorTable(self._device, 1, CHUNK_NORM)
        if not table.built_from(grads):
            table.update([grads], flags)
            table.remember(grads)
        chunks = table.struct.total_chunks
        if self._norm_buf is None or self._norm_buf.numel() < chunks + 1:
            self._norm_buf = torch.empty(chunks + 1, dtype=torch.float32, device=self._device)
        norm, partials = self._norm_buf[:1], self._norm_buf[1:]
        _C.mt_l2norm(table, 0, inv_scale, apply_scale, partials, norm)
        if floats:
            # the user's own fp32 gradients are clipped where they are, as torch does
            ftable = self._clip_table
            if ftable is None:
                ftable = self._clip_table = TensorTable(self._device, 1, CHUNK_STREAM)
            if not ftable.built_from(floats):
                ftable.update([floats], [_C.TF_GRAD_F32] * len(floats))
                ftable.remember(floats)
            _C.mt_clip_by_norm(ftable, 0, norm, max_norm, None)
        self._clip = (norm, max_norm)
        return norm[0].clone()

    # ---- checkpointing --------------------------------------------------------------------

    def state_dict(self):
        """Same keys as the reference (:267-286)."""
        state_dict = {}
        state_dict['loss_scaler'] = self.loss_scaler
        state_dict['dynamic_loss_scale'] = self.dynamic_loss_scale
        state_dict['overflow'] = self.overflow
        state_dict['first_closure_call_this_step'] = self.first_closure_call_this_step
        state_dict['optimizer_state_dict'] = self.optimizer.state_dict()
        state_dict['fp32_from_fp16'] = self.fp32_from_fp16_groups
        return state_dict

    def load_state_dict(self, state_dict):
        """Restore scaler, flags, inner optimizer and the saved fp32 masters (reference: :288-328)."""
        self._pending = None
        self._grads_ready = False
        self._clip = None
        self.loss_scaler = state_dict['loss_scaler']
        self.dynamic_loss_scale = state_dict['dynamic_loss_scale']
        self.overflow = state_dict['overflow']
        self.first_closure_call_this_step = state_dict['first_closure_call_this_step']
        self.optimizer.load_state_dict(state_dict['optimizer_state_dict'])
        # A checkpoint loaded with map_location -> cuda brings Adam's step counters to the device, where
        # torch reads each of them back every step (non-capturable mode) and our launch plan would
        # decline: keep them where torch itself creates them, on the host.
        if self._fused_kind() == "adam":
            for group in self.optimizer.param_groups:
                if group.get('capturable') or group.get('fused'):
                    continue
                for p in group['params']:
                    st = self.optimizer.state.get(p)
                    if st and isinstance(st.get('step'), torch.Tensor) and st['step'].is_cuda:
                        st['step'] = st['step'].cpu()
        for current_group, saved_group in zip(self.fp32_from_fp16_groups, state_dict['fp32_from_fp16']):
            for current, saved in zip(current_group, saved_group):
                current.data.copy_(saved.data)
        for t in self._step_tables.values():
            t.invalidate()
        self._plans = {}

    # ---- step -----------------------------------------------------------------------------------

    def step(self, closure=None):
        """Update masters with the wrapped optimizer's rule, refresh the fp16 params
        (reference: :330-384).  On overflow: adjust the scale, print, skip."""
        scale = self.loss_scaler.loss_scale

        # Launch the (device-predicated) fused update before the host looks at
        # the overflow word, so the GPU never waits for the host.
        commit = None
        if closure is None:
            commit = self._launch_fused_step()

        overflow = self.overflow
        self._update_scale(overflow)

        if overflow:
            print("OVERFLOW! Skipping step. Attempted loss scale: {}, reducing to {}"
                  .format(scale, self.loss_scale))
            return

        if closure is not None:
            retval = self._step_with_closure(closure)
            self._master_params_to_model_params()
        elif commit is not None:
            commit()
            retval = None
        else:
            if self.unfused_steps == 0:
                import warnings
                warnings.warn("apex_b200: FP16_Optimizer.step() is not taking the fused update for %s "
                              "(no fused kernel for this optimizer or these options); using the wrapped "
                              "optimizer's own step() plus one copy-back launch.  Counted in .unfused_steps."
                              % type(self.optimizer).__name__)
            self.unfused_steps += 1
            retval = self.optimizer.step()
            self._master_params_to_model_params()

        return retval

    def _step_with_closure(self, closure):
        """reference: :386-423"""
        def wrapped_closure():
            if self.first_closure_call_this_step:
                # fp16 params are fresh on entry to step()
                self.first_closure_call_this_step = False
            else:
                # the inner optimizer moved the masters since the last call and
                # knows nothing about the fp16 copies
                self._master_params_to_model_params()
            temp_loss = closure()
            while self.overflow:
                scale = self.loss_scaler.loss_scale
                self._update_scale(self.overflow)
                print("OVERFLOW within closure! Skipping step. Attempted loss scale: {}, "
                      "reducing to {}".format(scale, self.loss_scale))
                temp_loss = closure()
            return temp_loss

        retval = self.optimizer.step(wrapped_closure)
        self.first_closure_call_this_step = True
        return retval

    # ---- the fused path ---------------------------------------------------------------------------

    def _fused_kind(self):
        opt = self.optimizer
        if type(opt) is torch.optim.SGD:
            return "sgd"
        if type(opt) is torch.optim.Adam or type(opt) is torch.optim.AdamW:
            return "adam"
        return None

    def _launch_fused_step(self):
        """Enqueue one fused launch per param group if every precondition holds.

        Returns a ``commit`` callable (to run once the step is known not to be
        skipped: it publishes new optimizer state such as momentum buffers and
        Adam's step count), or None when the generic path must be taken.
        Nothing is launched in that case.
        """
        pending = self._pending
        kind = self._fused_kind()
        if kind is None or (pending is None and not self._grads_ready):
            return None
        if self._overflow_event is None and self._overflow:
            return None           # overflow already known on the host: step() skips
        from_masters = pending is None     # gradients were materialised (e.g. for clipping)
        inv_scale, apply_scale = (1.0, False) if from_masters else pending
        clip = None if from_masters else self._clip
        plans = []
        if not self._groups_intact():
            return None           # groups were edited after wrapping: the inner optimizer handles it
        for gi, group in enumerate(self.optimizer.param_groups):
            plan = self._plan_for(kind, gi, group, from_masters)
            if plan is False:
                return None
            if plan is not None:
                plans.append((group, plan))
        flag = self._live_flag()
        for group, plan in plans:
            if kind == "sgd":
                _C.mt_fused_sgd(plan.table, group['lr'], group['momentum'], group['dampening'],
                                group['weight_decay'], group['nesterov'], inv_scale, apply_scale, flag, clip)
            else:
                beta1, beta2 = group['betas']
                step = plan.step_count() + 1
                # torch/optim/adam.py:764-775, in python doubles
                bias_correction1 = 1 - beta1 ** step
                bias_correction2 = 1 - beta2 ** step
                step_size = (group['lr'] / bias_correction1) * -1
                _C.mt_fused_adam(plan.table, group['lr'], beta1, beta2, group['eps'], group['weight_decay'],
                                 step_size, bias_correction2 ** 0.5,
                                 type(self.optimizer) is torch.optim.AdamW or
                                 bool(group.get('decoupled_weight_decay', False)), inv_scale, apply_scale, flag,
                                 clip)

        def commit():
            for _, plan in plans:
                plan.commit(self.optimizer.state)
        return commit

    def _plan_for(self, kind, gi, group, from_masters):
        """The cached launch plan of one param group, rebuilt when anything it was built from has
        been replaced.  False = this group cannot be fused, None = nothing to do."""
        if group.get('maximize', False) or group.get('differentiable', False):
            return False
        if isinstance(group['lr'], torch.Tensor) or group.get('fused'):
            return False          # (a torch-fused optimizer keeps its step counters on the device)
        if kind == "adam":
            if group.get('amsgrad', False) or group.get('capturable', False):
                return False
            if isinstance(group['betas'][0], torch.Tensor) or isinstance(group['betas'][1], torch.Tensor):
                return False
        halves = self.fp16_groups[gi]
        masters = self.fp32_from_fp16_groups[gi]
        floats = self.fp32_from_fp32_groups[gi]
        a, b, c, d = self._spans[gi]
        if from_masters:
            grads = [_raw_grad(m) for m in masters] + [p.grad for p in floats]
        elif self._g16 is not None and len(self._g16) == len(self._flat_halves):
            grads = self._g16[a:b] + self._g32[c:d]
        else:
            grads = [h.grad for h in halves] + [p.grad for p in floats]
        if not grads:
            return None
        plan = self._plans.get(gi)
        if plan is not None and plan.still_valid(kind, from_masters, group, grads, self.optimizer.state):
            return plan
        plan = _StepPlan.build(self, kind, gi, group, from_masters, halves, masters, floats, grads)
        if plan is False:
            self._plans.pop(gi, None)
            return False
        self._plans[gi] = plan
        return plan

    # ---- backward ---------------------------------------------------------------------------------------

    def backward(self, loss, update_master_grads=True):
        """Scale the loss in fp32, backpropagate, then refresh the master gradients
        (reference: :425-484)."""
        self.loss_scaler.backward(loss.float())
        if update_master_grads:
            self.update_master_grads()

    def update_master_grads(self):
        """Check the fresh fp16 gradients and (virtually) copy/unscale them into the
        masters (reference: :486-497)."""
        self._pending = None
        self._grads_ready = False
        self._clip = None
        # every param.grad is read ONCE per iteration; the check, the in-place unscale and the launch
        # plan of step() all work on these lists (a gradient replaced between here and step() is not
        # seen, as in the reference, whose master gradients are a snapshot taken here)
        if self._groups_intact():
            self._g16 = [p.grad for p in self._flat_halves]
            self._g32 = [p.grad for p in self._flat_floats]
        else:
            self._g16 = [p.grad for g in self.fp16_groups for p in g]
            self._g32 = [p.grad for g in self.fp32_from_fp32_groups for p in g]
        if self.dynamic_loss_scale:
            self._check_overflow()
            # reference: `if self.overflow: return`.  The decision is still on the
            # device; everything enqueued below is predicated on it instead.
        loss_scale = self.loss_scale
        apply_scale = loss_scale != 1.0
        inv_scale = 1. / loss_scale if apply_scale else 1.0
        flag = self._live_flag()
        if flag is None and self.dynamic_loss_scale and self.overflow:
            return          # overflow already known on the host (custom scaler path)
        # fp32-native parameters are the user's own tensors: unscale in place now.
        if apply_scale:
            groups = self.optimizer.param_groups
            if self._groups_intact():
                grads = [g for g in self._g32 if g is not None]
            else:
                # param_groups were edited after wrapping: the reference unscales the gradient of
                # EVERY parameter the inner optimizer owns (:236-241); masters are handled virtually
                grads = [p.grad for g in groups for p in g['params']
                         if type(p) is not _MasterParam and p.grad is not None]
            self._unscale_in_place(grads, inv_scale, flag)
        self._pending = (inv_scale, apply_scale)

    def inspect_master_grad_data(self):
        """reference: :499-522 -- not implemented there either."""
        raise NotImplementedError("Currently not implemented, working on it...")

    # ---- promoted attributes (reference: :524-550) --------------------------------------------------------

    def _get_loss_scale(self):
        return self.loss_scaler.loss_scale

    def _set_loss_scale(self, value):
        self.loss_scaler.cur_scale = value

    loss_scale = property(_get_loss_scale, _set_loss_scale)

    def _get_state(self):
        return self.optimizer.state

    def _set_state(self, value):
        self.optimizer.state = value
        self._plans = {}

    state = property(_get_state, _set_state)

    def _get_param_groups(self):
        return self.optimizer.param_groups

    def _set_param_groups(self, value):
        self.optimizer.param_groups = value

    param_groups = property(_get_param_groups, _set_param_groups)
