"""Data-parallel gradient averaging (reference: apex/parallel/distributed.py).

Observable behaviour is the reference's: parameters broadcast from rank 0 at
construction; gradients bucketed per dtype in the order their hooks fire on
rank 0, a bucket being cut once it holds >= ``message_size`` ELEMENTS, leftovers
appended half -> float -> double; the structure broadcast as one int32 vector;
buckets reduced strictly in index order on a side stream while backward keeps
running; sum across ranks, then divide by the world size in the bucket's own
dtype; the same three sanity errors.  The code is organised differently:

* ``_BucketLayout``   -- the learned structure (what rank 0 dictates).
* ``_BucketRound``    -- per-backward progress (which slots are filled, what may go).
* ``_BucketPipeline`` -- the data path.  Every bucket has a persistent flat home
  (parallel/buckets.py) that the gradients LIVE in: ``param.grad`` is a view of it
  after the first overlapped backward, so a bucket is averaged in place by ONE launch
  (parallel/symm.py + csrc/symm_allreduce.cu: barrier, sum across ranks, x 1/world,
  inf/NaN check, barrier) -- or NCCL ``all_reduce`` + one in-place scale launch --
  instead of the reference's alloc + cat + allreduce + divide + n copy_ launches
  (distributed.py:11-23).
* ``_FlatSlot``       -- coalesced staging buffer + table for one-off flat calls on
  arbitrary tensor lists (``flat_dist_call``: initial broadcast, the iteration that
  learns the structure, ``Reducer``): 1 gather launch -> collective -> 1 (x 1/world,
  scatter, inf/NaN check) launch.
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
from .buckets import GradBucket

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
    """What a gradient tensor carries after its bucket went through the averaging-with-check kernel:
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


# what setting up symmetric memory can raise on a box or torch build that does not have it (the module
# is a private torch API: missing, renamed or unsupported are all "use NCCL")
_SYMM_SETUP_ERRORS = (RuntimeError, ImportError, AttributeError, NotImplementedError)


class _BucketPipeline(object):
    """Averages gradient buckets in their persistent flat homes (parallel/buckets.py).

    NVLink domain, NCCL backend (parallel/symm.py): the bucket lives in symmetric memory and ONE kernel
    of ours per bucket does barrier -> two-shot sum -> x 1/world -> inf/NaN check -> barrier, in place,
    on the stream ``submit`` is called on.  Gradients that already live in the bucket cost nothing
    else; strays are copied in first (one launch) and either become views of the bucket (``owners``
    given: DistributedDataParallel) or get the result copied back (plain tensor lists).

    Otherwise (gloo, fp64, one rank, APEX_B200_SYMM=0): same flat homes in ordinary memory,
    ``dist.all_reduce`` on the flat buffer, then one in-place launch for x 1/world and the check on a
    second stream, so that the collective of bucket b overlaps the scaling of bucket b-1.

    The reference runs cat + all_reduce + ``/=`` + n copy_ per bucket on one stream
    (distributed.py:11-23 under :294-296).  The order of collectives is the same: as submitted.
    """

    MIN_ARENA = 32 << 20

    def __init__(self, gather_stream):
        self.gather_stream = gather_stream
        self.scatter_stream = torch.cuda.Stream()
        # symmetric path: consecutive buckets rotate over a few collective streams (one signal
        # channel each), so that the barriers / ramp-up of bucket b+1 overlap the drain of bucket
        # b; the submitting stream only carries stray copy-ins and the events that release the
        # kernels.  The choice depends on the bucket's position in the round only: same on every
        # rank.  APEX_B200_SYMM_STREAMS=1: everything on the submitting stream.
        n_comm = symm.comm_stream_count()
        self.comm_streams = [torch.cuda.Stream() for _ in range(n_comm)] if n_comm > 1 else None
        self._symm_count = 0
        self.slots = {}                # key -> (GradBucket, Arena or None, byte offset)
        self.arenas = []               # symmetric allocations of the current plan
        self._arena_flags = {}         # id(arena) -> (int32[1] view, byte offset)
        self._used_arenas = []         # arenas touched in the current round, primary first
        self.check_flag = None         # device word raised when an averaged value of the round is inf/NaN
        self._armed = True
        self._token = None
        self._round_id = 0

    # ---- rounds and the overflow word -------------------------------------------------------------

    def begin_round(self):
        """A new backward starts: the word is cleared before its first bucket, gradients verified
        from now on belong to a fresh round."""
        self._round_id += 1
        self._token = None
        self._armed = True
        self._symm_count = 0

    def _flag_of(self, arena):
        got = self._arena_flags.get(id(arena))
        if got is None:
            got = self._arena_flags[id(arena)] = arena.new_flag()
        return got

    def _arm(self, device, arena):
        """Make sure the round's word exists and is zero before the first kernel that may raise it."""
        if self.check_flag is None or self.check_flag.device != device:
            self.check_flag = self._flag_of(arena)[0] if arena is not None else \
                torch.zeros(1, dtype=torch.int32, device=device)
        if self._armed:
            self._armed = False
            self._used_arenas = []
            _C.flag_clear(self.check_flag)
            for a in self.arenas:
                word = self._flag_of(a)[0]
                if word.data_ptr() != self.check_flag.data_ptr():
                    _C.flag_clear(word)
            cleared = torch.cuda.Event()
            cleared.record()
            self.scatter_stream.wait_event(cleared)
        if arena is not None and arena not in self._used_arenas:
            self._used_arenas.append(arena)

    def _tag(self, tensors):
        if self._token is None:
            self._token = _Verified(self.check_flag, self._round_id)
        token = self._token
        for g in tensors:
            setattr(g, VERIFIED_ATTR, (token, g._version))
        token.count += len(tensors)

    # ---- flat homes -------------------------------------------------------------------------------------

    def reset(self):
        """The bucket structure changed: new homes will be planned.  Old arenas are not reused (old
        gradient views may still be alive); they are released with their last view."""
        self.slots = {}
        self.arenas = []
        self._arena_flags = {}
        self._used_arenas = []
        self.check_flag = None
        self._armed = True

    @staticmethod
    def _wants_symm(like):
        return symm.enabled() and like[0].dtype in (torch.float16, torch.float32) and like[0].is_cuda

    def _symm_home(self, like, world):
        """(storage, arena, offset) inside a symmetric allocation, growing the set of arenas if needed."""
        es = like[0].element_size()
        nbytes = GradBucket.nbytes_for(like, world)
        for arena in self.arenas:
            off = arena.take(nbytes)
            if off is not None:
                break
        else:
            size = max(self.MIN_ARENA, nbytes + symm.ALIGN, 2 * (self.arenas[-1].nbytes if self.arenas else 0))
            arena = symm.Arena(size, like[0].device)        # collective: every rank gets here alike
            self.arenas.append(arena)
            off = arena.take(nbytes)
        return arena.buf[off:off + nbytes].view(like[0].dtype), arena, off

    def _new_slot(self, like):
        world = dist.get_world_size()
        if self._wants_symm(like):
            try:
                storage, arena, off = self._symm_home(like, world)
                return (GradBucket(like, world, storage), arena, off)
            except _SYMM_SETUP_ERRORS as err:     # every rank fails alike: allocation + rendezvous are collective
                if "bucket tensors must be dense" in str(err):
                    raise
                symm.mark_broken(err)
        return (GradBucket(like, world), None, 0)

    def plan(self, buckets):
        """Create the homes of ``[(key, tensors-like)]`` at once: ONE symmetric allocation (one
        rendezvous) holds every bucket that can live there, back to back."""
        self.reset()
        world = dist.get_world_size()
        want = [like for _, like in buckets if like and self._wants_symm(like)]
        if want:
            total = sum((GradBucket.nbytes_for(like, world) + symm.ALIGN - 1) // symm.ALIGN * symm.ALIGN
                        for like in want) + symm.ALIGN
            try:
                self.arenas.append(symm.Arena(total, want[0][0].device))
            except _SYMM_SETUP_ERRORS as err:
                symm.mark_broken(err)
        for key, like in buckets:
            if like:
                try:
                    self.slots[key] = self._new_slot(like)
                except (RuntimeError, TypeError):
                    pass                      # not a bucket we can house: submit() reports it in context

    # ---- the data path -------------------------------------------------------------------------------------

    def submit(self, key, bucket, verify=False, owners=None):
        """Enqueue the averaging of ``bucket`` on the current stream (DDP: its reduction stream);
        returns at once.  ``owners`` (parameters, one per tensor): a gradient that does not live in
        the bucket yet is replaced by its view (``owner.grad = view``) instead of being copied back.
        ``verify``: tag the gradients so that FP16_Optimizer can take the round's word as its own
        overflow check.  Returns the list of tensors that now hold the averaged gradients."""
        if len(bucket) == 0:
            return bucket
        slot = self.slots.get(key)
        if slot is None or not slot[0].fits(bucket):
            slot = self.slots[key] = self._new_slot(bucket)
        home, arena, offset = slot
        world = dist.get_world_size()
        checks = home.dtype in (torch.float16, torch.float32)
        self._arm(home.device, arena)
        strays = home.strays(bucket)
        home.copy_in(bucket, strays)
        if arena is not None:
            word, word_off = self._flag_of(arena)
            streams, ctas = arena.rotation() if self.comm_streams is not None else (1, None)
            if streams == 1:
                arena.allreduce(offset, home.padded, home.dtype, world, word_off, 0, ctas=ctas)
                if strays and owners is None:
                    home.copy_out(bucket, strays)
            else:
                channel = self._symm_count % streams
                self._symm_count += 1
                ready = torch.cuda.Event()
                ready.record()                       # gradients final, strays copied in, word cleared
                cs = self.comm_streams[channel]
                with torch.cuda.stream(cs):
                    cs.wait_event(ready)
                    arena.allreduce(offset, home.padded, home.dtype, world, word_off, channel, ctas=ctas)
                    if strays and owners is None:
                        home.copy_out(bucket, strays)
        else:
            work = dist.all_reduce(home.flat, async_op=True)
            with torch.cuda.stream(self.scatter_stream):
                work.wait()
                _C.scale_flat(home.flat, world, self.check_flag if checks else None)
                if strays and owners is None:
                    home.copy_out(bucket, strays)
        result = bucket
        if strays and owners is not None:
            result = list(bucket)
            for i in strays:
                owners[i].grad = result[i] = home.views[i]
        if verify and checks:
            self._tag(result)
        return result

    def join(self, stream):
        """Make ``stream`` wait for everything submitted so far; the round's verification is then
        usable by work enqueued on ``stream``."""
        stream.wait_stream(self.gather_stream)
        stream.wait_stream(self.scatter_stream)
        for cs in self.comm_streams or ():
            stream.wait_stream(cs)
        if len(self._used_arenas) > 1 or (self._used_arenas and
                                          self._flag_of(self._used_arenas[0])[0].data_ptr() != self.check_flag.data_ptr()):
            # buckets of several allocations raised several words: fold them into the round's
            with torch.cuda.stream(stream):
                for a in self._used_arenas:
                    word = self._flag_of(a)[0]
                    if word.data_ptr() != self.check_flag.data_ptr():
                        torch.maximum(self.check_flag, word, out=self.check_flag)
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
        self._bucket_params = []  # bucket_idx -> parameters, in bucket order
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
        # the flat homes of all buckets at once (one symmetric allocation); the gradients move in
        # during the first overlapped backward (comm_ready_buckets)
        self._bucket_params = [[self.active_params[i] for i in members] for members in self._layout.members]
        self._pipeline.plan([(b, [p.data for p in params]) for b, params in enumerate(self._bucket_params)])

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
        # the gradient tensor itself (not a `.data` alias): the pipeline tags the very object the
        # optimizer will later find in param.grad
        eligible = self._round.put(bucket_idx, bucket_loc, param.grad)
        if not eligible:
            return
        # the gradients were produced on the current stream
        self.reduction_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.reduction_stream):
            for b in eligible:
                # gradients not yet living in the bucket's flat home are re-homed: param.grad becomes
                # a view of the bucket, so the next backward accumulates straight into it
                self._round.grads[b] = self._pipeline.submit(b, self._round.grads[b], verify=True,
                                                             owners=self._bucket_params[b])

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
