"""Data-parallel gradient averaging (reference: apex/parallel/distributed.py).

Observable behaviour is the reference's: parameters broadcast from rank 0 at
construction; gradients bucketed per dtype in the order their hooks fire on
rank 0, a bucket being cut once it holds >= ``message_size`` ELEMENTS, leftovers
appended half -> float -> double; the structure broadcast as one int32 vector;
buckets reduced strictly in index order on a side stream while backward keeps
running; sum across ranks, then divide by the world size in the bucket's own
dtype; the same three sanity errors.  The code is organised differently:

* ``_BucketLayout`` -- the learned structure (what rank 0 dictates).
* ``_BucketRound``  -- per-backward progress (which slots are filled, what may go).
* ``_FlatSlot``     -- a persistent coalesced buffer + device tensor table per
  bucket, so that the data path of one bucket is
      1 gather launch -> NCCL collective -> 1 (x 1/world, scatter, inf/NaN check) launch
  instead of the reference's alloc + cat + allreduce + divide + n copy_ launches
  (distributed.py:11-23).  A bucket of a single tensor is reduced in place.
"""
import copy
from collections import OrderedDict

import torch
import torch.distributed as dist
from torch.autograd import Variable
from torch.nn.modules import Module

from .. import _C
from ..multi_tensor import TensorTable, CHUNK_STREAM, is_dense
from . import symm

# bucket class per dtype; leftovers are appended in this order (reference :154-156, :178-180)
_DTYPE_CLASS = {torch.float16: 0, torch.float32: 1, torch.float64: 2}


# ---------------------------------------------------------------------------
# flat collective on a list of tensors
# ---------------------------------------------------------------------------

class _FlatSlot(object):
    """Coalesced staging buffer and tensor table, reused across iterations."""

    __slots__ = ("table", "flat")

    def __init__(self):
        self.table = None
        self.flat = None

    def bind(self, tensors):
        head = tensors[0]
        if head.dtype not in _DTYPE_CLASS:
            raise TypeError("apex_b200: unsupported bucket dtype %s" % head.dtype)
        total = 0
        for t in tensors:
            if t.dtype != head.dtype or t.device != head.device:
                raise TypeError("apply_flat_dist_call needs tensors of one type, got %s and %s"
                                % (head.type(), t.type()))
            if not is_dense(t):      # the flat buffer holds each tensor in ITS memory order; the round trip restores it
                raise RuntimeError("apex_b200: bucket tensors must be dense (contiguous or channels-last)")
            total += t.numel()
        if self.table is None or self.table.device != head.device:
            self.table = TensorTable(head.device, 1, CHUNK_STREAM)
        if self.flat is None or self.flat.numel() != total or self.flat.dtype != head.dtype \
                or self.flat.device != head.device:
            self.flat = torch.empty(total, dtype=head.dtype, device=head.device)
        self.table.update([tensors])
        return self.flat


_recent_slots = OrderedDict()
_RECENT_MAX = 64


def _recent_slot(tensors):
    head = tensors[0]
    key = (str(head.device), head.dtype, len(tensors), head.data_ptr())
    slot = _recent_slots.pop(key, None)
    if slot is None:
        slot = _FlatSlot()
    _recent_slots[key] = slot
    while len(_recent_slots) > _RECENT_MAX:
        _recent_slots.popitem(last=False)
    return slot


def apply_flat_dist_call(bucket, call, extra_args=None, _slot=None, _flag=None):
    """Run collective ``call`` on the concatenation of ``bucket`` and write the result
    back into the bucket's tensors; for ``dist.all_reduce`` the sum is then divided
    by the world size (sum first, divide after, in the bucket dtype -- a sum that
    overflows fp16 has to stay inf).  reference: distributed.py:11-23.

    ``_slot`` (a persistent ``_FlatSlot``) and ``_flag`` (device word raised when
    an averaged value is inf/NaN) are internal.
    """
    if len(bucket) == 0:
        return
    args = () if extra_args is None else tuple(extra_args)
    averaging = call is dist.all_reduce

    if len(bucket) == 1 and bucket[0].is_contiguous():
        # nothing to coalesce: reduce the tensor where it lives
        only = bucket[0]
        call(only, *args)
        if averaging:
            _C.scale_flat(only.view(-1), dist.get_world_size(), _flag)
        return

    slot = _slot if _slot is not None else _recent_slot(bucket)
    flat = slot.bind(bucket)
    _C.mt_flatten(slot.table, 0, flat)
    call(flat, *args)
    if averaging:
        _C.mt_scale_unflatten(slot.table, flat, 0, dist.get_world_size(), True, _flag)
    else:
        _C.mt_scale_unflatten(slot.table, flat, 0, 1.0, False, None)


class _Verified(object):
    """What a gradient tensor carries after its bucket went through the scatter-with-check pass:
    which device word would have been raised had the averaged values held an inf/NaN, in which
    round, and the tensor's version counter at that moment (any later in-place torch op on the
    gradient changes the counter and voids the verification)."""

    __slots__ = ("flag", "round_id", "done", "count")

    def __init__(self, flag, round_id):
        self.flag = flag
        self.round_id = round_id
        self.done = False          # set once the round's streams have been joined
        self.count = 0             # gradient tensors checked into this round's flag


VERIFIED_ATTR = "_apex_b200_verified"


def verified_flag(grads):
    """If every gradient in ``grads`` was averaged AND checked by one completed pipeline round and
    has not been touched since, return that round's device flag word; else None.  Used by
    FP16_Optimizer so that the overflow check of the reference's update_master_grads
    (fp16_optimizer.py:493-495) costs no extra pass over the gradients on multi-GPU runs."""
    token = None
    for g in grads:
        mark = getattr(g, VERIFIED_ATTR, None)
        if mark is None:
            return None
        t, version = mark
        if g._version != version or not t.done or (token is not None and t is not token):
            return None
        token = t
    # the word answers "any inf/NaN among ALL gradients of the round": only an optimizer that owns
    # exactly those gradients may take it as its own answer (the reference checks its own only)
    if token is None or token.count != len(grads):
        return None
    return token.flag


class _BucketPipeline(object):
    """Averages buckets with the three stages of consecutive buckets overlapped:

        gather stream :  gather(b+1)                 (apex_b200_mt_flatten)
        NCCL stream   :        all_reduce(b)         (torch.distributed, async)
        scatter stream:              x 1/world, scatter, check (b-1)   (apex_b200_mt_scale_unflatten)

    The reference runs the three steps of a bucket back to back on one stream
    (distributed.py:11-23 under :294-296), so HBM-bound copies and the NVLink-bound collective
    never overlap.  The order of collectives is unchanged: buckets are issued in the order given.
    """

    def __init__(self, gather_stream):
        self.gather_stream = gather_stream
        self.scatter_stream = torch.cuda.Stream()
        self.symm_pipe = None          # created when the symmetric-memory path is first used
        self.slots = {}
        self._spare = []
        self.check_flag = None         # device word raised by the scatter pass of the current round
        self._token = None
        self._round_id = 0

    def begin_round(self):
        """A new backward starts: gradients verified from now on belong to a fresh round."""
        self._round_id += 1
        self._token = None

    def _mark(self, bucket):
        """Called when a bucket is submitted: clear the round's flag before its first bucket, and
        tag every gradient of the bucket."""
        dev = bucket[0].device
        if self.check_flag is None or self.check_flag.device != dev:
            self.check_flag = torch.zeros(1, dtype=torch.int32, device=dev)
        if self._token is None:
            self._token = _Verified(self.check_flag, self._round_id)
            _C.flag_clear(self.check_flag)          # on the gather stream, before any scatter of this round
            self._cleared = torch.cuda.Event()
            self._cleared.record()
            self.scatter_stream.wait_event(self._cleared)
        token = self._token
        for g in bucket:
            setattr(g, VERIFIED_ATTR, (token, g._version))
        token.count += len(bucket)
        return self.check_flag

    def reset(self):
        """The bucket structure changed.  Symmetric-memory slices cannot be given back, so the
        slots are kept aside and reused by new buckets of the same size and dtype."""
        self._spare.extend(s for s in self.slots.values() if isinstance(s, symm.SymmSlot))
        self.slots = {}

    def _symm_slot(self, bucket):
        total = sum(t.numel() for t in bucket)
        for i, spare in enumerate(self._spare):
            if spare.total == total and spare.flat is not None and spare.flat.dtype == bucket[0].dtype:
                return self._spare.pop(i)
        return symm.SymmSlot()

    def submit(self, key, bucket, flag=None, verify=False):
        """Enqueue the averaging of ``bucket``; returns at once.  Call from inside
        ``with torch.cuda.stream(gather_stream)``.  ``verify=True``: the scatter pass also checks the
        averaged values and the gradients are tagged so that FP16_Optimizer can skip its own scan."""
        if len(bucket) == 0:
            return
        if verify and flag is None and bucket[0].dtype in (torch.float16, torch.float32):
            flag = self._mark(bucket)
        world = dist.get_world_size()
        if len(bucket) == 1 and bucket[0].is_contiguous():
            only = bucket[0]
            work = dist.all_reduce(only, async_op=True)
            with torch.cuda.stream(self.scatter_stream):
                work.wait()
                _C.scale_flat(only.view(-1), world, flag)
            return
        if symm.enabled() and bucket[0].dtype in (torch.float16, torch.float32):
            # our own allreduce over peer-mapped memory instead of NCCL (parallel/symm.py)
            slot = self.slots.get(key)
            try:
                if not isinstance(slot, symm.SymmSlot):
                    slot = self._symm_slot(bucket)
                    slot.bind(bucket)             # allocation + rendezvous: may not be possible on this box
                    self.slots[key] = slot
            except RuntimeError as err:           # every rank fails alike: the rendezvous is collective
                symm.mark_broken(err)
                self.slots.pop(key, None)
            else:
                if self.symm_pipe is None:
                    self.symm_pipe = symm.SymmPipeline(self.gather_stream, self.scatter_stream)
                self.symm_pipe.submit(slot, bucket, flag)
                return
        slot = self.slots.get(key)
        if not isinstance(slot, _FlatSlot):
            slot = self.slots[key] = _FlatSlot()
        flat = slot.bind(bucket)
        _C.mt_flatten(slot.table, 0, flat)
        work = dist.all_reduce(flat, async_op=True)
        with torch.cuda.stream(self.scatter_stream):
            work.wait()
            _C.mt_scale_unflatten(slot.table, flat, 0, world, True, flag)

    def join(self, stream):
        """Make ``stream`` wait for everything submitted so far; the round's verification is then
        usable by work enqueued on ``stream``."""
        stream.wait_stream(self.gather_stream)
        if self.symm_pipe is not None:
            self.symm_pipe.join(stream)
        stream.wait_stream(self.scatter_stream)
        if self._token is not None:
            self._token.done = True


def flat_dist_call(tensors, call, extra_args=None, _keep_buffers=True):
    """One flat call per tensor type, types taken in first-seen order.
    reference: distributed.py:27-38.  ``_keep_buffers=False`` (internal) releases the coalesced
    buffers afterwards: for one-off calls such as the initial broadcast or the flat allreduce of the
    iteration that learns the bucket structure (one buffer as large as all gradients)."""
    flat_dist_call.warn_on_half = True
    per_type = OrderedDict()
    for t in tensors:
        per_type.setdefault(t.type(), []).append(t)
    for same_type in per_type.values():
        apply_flat_dist_call(same_type, call, extra_args, _slot=None if _keep_buffers else _FlatSlot())


def extract_tensors(maybe_tensor, tensor_list):
    """Append every tensor found in a (nested) iterable.  reference: distributed.py:41-49"""
    if torch.is_tensor(maybe_tensor):
        tensor_list.append(maybe_tensor)
        return
    try:
        children = iter(maybe_tensor)
    except TypeError:
        return
    for child in children:
        extract_tensors(child, tensor_list)


class Reducer(object):
    """Averaging on demand: nothing is communicated until ``reduce()``.

    Built from a module, its parameters are broadcast from rank 0 and ``reduce()``
    averages their gradients; built from an iterable of tensors, ``reduce()``
    averages those tensors.  reference: distributed.py:52-93
    """

    def __init__(self, module_or_grads_list):
        if isinstance(module_or_grads_list, Module):
            self.module = module_or_grads_list
            flat_dist_call([p.data for p in self.module.parameters()], dist.broadcast, (0,), _keep_buffers=False)
        else:
            self.module = None
            self.grads = []
            extract_tensors(module_or_grads_list, self.grads)

    def reduce(self):
        if self.module:
            tensors = [p.grad.data for p in self.module.parameters() if p.grad is not None]
        else:
            tensors = self.grads
        flat_dist_call(tensors, dist.all_reduce)


# ---------------------------------------------------------------------------
# bucket bookkeeping
# ---------------------------------------------------------------------------

class _BucketLayout(object):
    """Which parameter sits where: ``members[b][loc]`` is an index into the list of
    active parameters.  Built on every rank from its own hook order, then replaced
    by rank 0's through ``encode``/``decode`` (reference: distributed.py:176-203)."""

    def __init__(self, message_size):
        self.message_size = message_size
        self.members = []                 # closed buckets
        self.open = [[], [], []]          # running bucket per dtype class
        self.open_numel = [0, 0, 0]

    def arrive(self, dtype_class, param_index, numel):
        """A gradient became ready while the structure is being learned (reference :250-261)."""
        self.open[dtype_class].append(param_index)
        self.open_numel[dtype_class] += numel
        if self.open_numel[dtype_class] >= self.message_size:
            self.members.append(self.open[dtype_class])
            self.open[dtype_class] = []
            self.open_numel[dtype_class] = 0

    def close(self):
        for leftover in self.open:
            if leftover:
                self.members.append(leftover)
        self.open = [[], [], []]
        self.open_numel = [0, 0, 0]

    def encode(self):
        """[num_buckets] + sizes + members, the vector the reference broadcasts (:185-187)."""
        out = [len(self.members)] + [len(b) for b in self.members]
        for b in self.members:
            out.extend(b)
        return out

    def decode(self, info):
        n = info[0]
        sizes = info[1:n + 1]
        body = info[n + 1:]
        self.members, at = [], 0
        for size in sizes:
            self.members.append(list(body[at:at + size]))
            at += size

    @property
    def sizes(self):
        return [len(b) for b in self.members]

    def positions(self):
        """param index -> (bucket_idx, bucket_loc)"""
        where = {}
        for b, bucket in enumerate(self.members):
            for loc, param_index in enumerate(bucket):
                where[param_index] = (b, loc)
        return where


class _BucketRound(object):
    """Progress of one backward pass through a fixed layout."""

    def __init__(self, sizes):
        self.sizes = sizes
        self.grads = [[None] * n for n in sizes]
        self.filled = [0] * len(sizes)
        self.next_bucket = 0
        self.parked = set()               # full, but an earlier bucket is not

    def put(self, bucket_idx, bucket_loc, grad):
        """Store a ready gradient; returns the bucket indices that may be reduced now,
        in the order they must be issued (identical on every rank)."""
        row = self.grads[bucket_idx]
        if row[bucket_loc] is not None:
            raise RuntimeError("The backward pass is attempting to replace an already-filled "
                               "bucket slot.  This is almost certainly an error.")
        row[bucket_loc] = grad
        self.filled[bucket_idx] += 1
        if self.filled[bucket_idx] != self.sizes[bucket_idx]:
            return []
        if bucket_idx != self.next_bucket:
            self.parked.add(bucket_idx)
            return []
        go = [bucket_idx]
        self.next_bucket += 1
        for waiting in sorted(self.parked):
            if waiting > self.next_bucket:
                break
            if waiting < self.next_bucket:
                raise ValueError("i should always be >= next_bucket")
            go.append(waiting)
            self.parked.remove(waiting)
            self.next_bucket += 1
        return go

    def verify_complete(self):
        if self.next_bucket != len(self.sizes):
            raise RuntimeError("In epilogue, next_bucket != num_buckets.  "
                               "This probably indicates some buckets were not allreduced.")
        for have, want in zip(self.filled, self.sizes):
            if have != want:
                raise RuntimeError("Some param buckets were not allreduced.")


# ---------------------------------------------------------------------------
# the module wrapper
# ---------------------------------------------------------------------------

class DistributedDataParallel(Module):
    """Wraps ``module``; gradients are averaged over all processes during ``backward()``,
    overlapped with the rest of backward.

    Args (reference: distributed.py:96-124):
        module: the network.
        message_size: minimum number of elements per communication bucket (default 1e7).
        delay_allreduce: communicate only once backward has finished (no overlap).
        shared_param: retired option; any value but None raises ValueError.
    """

    def __init__(self, module, message_size=10000000, delay_allreduce=False, shared_param=None):
        super(DistributedDataParallel, self).__init__()

        # the reference looks for dist.DistBackend / dist.dist_backend (both removed
        # from torch since); all it needs is the backend name
        self._backend = dist.get_backend()
        self.backend_enum_holder = dist.Backend
        self.warn_on_half = self._backend == dist.Backend.GLOO

        if shared_param is not None:
            raise ValueError("shared_param is no longer supported as an option.  It was misleadingly named from the start.  It turns out overlapping communication with computation should work fine with shared parameters.  If you still wish to delay communication to the end of the backward pass, use delay_allreduce=True|False instead.")

        self.delay_allreduce = delay_allreduce
        self.message_size = message_size
        self.reduction_stream = torch.cuda.Stream()
        self.module = module

        if self._backend == dist.Backend.NCCL:
            for p in self.module.parameters():
                assert p.is_cuda, "NCCL backend only supports model parameters to be on GPU."

        self.active_params = []
        self.needs_refresh = True
        self.callback_queued = False
        self._layout = None
        self._round = None
        self._where = {}          # id(param) -> (bucket_idx, bucket_loc)
        self._index_of = {}       # id(param) -> index in active_params
        self._pipeline = _BucketPipeline(self.reduction_stream)

        self.create_hooks()
        flat_dist_call([p.data for p in self.module.parameters()], dist.broadcast, (0,), _keep_buffers=False)

    # ---- views of the structure under the reference's attribute names ---------------

    @property
    def num_buckets(self):
        return 0 if self._layout is None else len(self._layout.members)

    @property
    def bucket_sizes(self):
        return [] if self._layout is None else self._layout.sizes

    @property
    def buckets(self):
        if self._round is not None:
            return self._round.grads
        return [] if self._layout is None else self._layout.members

    @property
    def param_id_to_bucket(self):
        return self._where

    @property
    def next_bucket(self):
        return 0 if self._round is None else self._round.next_bucket

    # ---- pickling: a CUDA stream cannot be pickled (reference :163-172) -------------

    def __getstate__(self):
        state = copy.copy(self.__dict__)
        for transient in ("reduction_stream", "_pipeline", "_round"):
            state.pop(transient, None)
        return state

    def __setstate__(self, state):
        super(DistributedDataParallel, self).__setstate__(state)
        self.reduction_stream = torch.cuda.Stream()
        self._pipeline = _BucketPipeline(self.reduction_stream)
        self._round = None

    # ---- structure ---------------------------------------------------------------------

    def sync_bucket_structure(self):
        """Finish this rank's layout, then take rank 0's (reference: distributed.py:176-203)."""
        self._layout.close()
        mine = self._layout.encode()
        device = next(self.module.parameters()).device
        info = torch.tensor(mine, dtype=torch.int32, device=device)
        dist.broadcast(info, 0)
        self._layout.decode(info.tolist())      # one device-to-host copy for the whole vector
        self._where = {id(self.active_params[i]): pos for i, pos in self._layout.positions().items()}
        self._pipeline.reset()

    # ---- hooks ---------------------------------------------------------------------------

    def create_hooks(self):
        """Hook every trainable parameter's gradient accumulator (reference: :206-276)."""
        self.grad_accs = []
        for param in self.module.parameters():
            if param.requires_grad:
                # a leaf's AccumulateGrad node is the first input of any trivial view of it
                accumulator = param.expand_as(param).grad_fn.next_functions[0][0]
                accumulator.register_hook(self._make_hook(param))
                self.grad_accs.append(accumulator)

    def _make_hook(self, param):
        def on_grad_ready(*unused):
            self._grad_ready(param)
        return on_grad_ready

    def _queue_once(self, callback):
        if not self.callback_queued:
            Variable._execution_engine.queue_callback(callback)
            self.callback_queued = True

    def _grad_ready(self, param):
        learning = self.delay_allreduce or self.needs_refresh
        if learning:
            if not self.delay_allreduce:
                self._layout.arrive(_DTYPE_CLASS[param.dtype], self._index_of[id(param)], param.numel())
            self._queue_once(self._finish_backward_flat)
        else:
            self._queue_once(self._finish_backward_overlapped)
            self.comm_ready_buckets(param)

    def _finish_backward_flat(self):
        """End of a backward that did no overlapped communication: adopt the bucket
        structure if one was being learned, then one flat allreduce per dtype over
        every gradient (reference ``allreduce_params``: :211-221)."""
        if not self.delay_allreduce and self.needs_refresh:
            self.sync_bucket_structure()
            self.needs_refresh = False
        grads = [p.grad.data for p in self.module.parameters() if p.grad is not None]
        # with delay_allreduce this runs every iteration: keep the buffers then
        flat_dist_call(grads, dist.all_reduce, _keep_buffers=self.delay_allreduce)

    def _finish_backward_overlapped(self):
        """reference ``overlapping_backward_epilogue``: :223-233"""
        self._pipeline.join(torch.cuda.current_stream())
        self._round.verify_complete()

    def comm_ready_buckets(self, param):
        """``param``'s gradient is final: file it, and reduce whatever became eligible on
        the side stream (reference: :279-315)."""
        bucket_idx, bucket_loc = self._where[id(param)]
        # the gradient tensor itself (not a `.data` alias): the scatter pass tags the very object the
        # optimizer will later find in param.grad
        eligible = self._round.put(bucket_idx, bucket_loc, param.grad)
        if not eligible:
            return
        # the gradients were produced on the current stream
        self.reduction_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.reduction_stream):
            for b in eligible:
                self._pipeline.submit(b, self._round.grads[b], verify=True)

    # ---- forward -----------------------------------------------------------------------------

    def forward(self, *inputs, **kwargs):
        """Run the wrapped module and arm the bookkeeping for the coming backward
        (reference: :318-351)."""
        result = self.module(*inputs, **kwargs)

        if not self.delay_allreduce:
            trainable = [p for p in self.module.parameters() if p.requires_grad]
            # only forward may ask for a refresh; only the end of backward clears it
            changed = (not self.active_params) or len(trainable) != len(self.active_params) or \
                any(a is not b for a, b in zip(trainable, self.active_params))
            if changed:
                self.needs_refresh = True

            if self.needs_refresh:
                self._layout = _BucketLayout(self.message_size)
                self._round = None
                self._where = {}
                self._index_of = {id(p): i for i, p in enumerate(trainable)}
            else:
                self._round = _BucketRound(self._layout.sizes)
                self._pipeline.begin_round()

            self.active_params = trainable

        self.callback_queued = False
        return result
