"""Flat buckets in peer-mapped (symmetric) memory and their allreduce without NCCL.

torch.distributed's symmetric memory is used for the plumbing only: allocating the buffer so that
every rank of the NVLink domain can address every other rank's copy (and, on NVSwitch, the multicast
alias of all copies), and the cross-rank barrier between phases.  The data path is ours:

    gather   (apex_b200_mt_flatten)        gradients -> my copy of the flat bucket
    barrier                                 every rank has gathered
    allreduce (apex_b200_symm_allreduce)    two-shot SUM: switch-side (multimem) or peer loads/stores
    barrier                                 every rank has written its slice everywhere
    scatter  (apex_b200_mt_scale_unflatten) x 1/world, back into the gradients, inf/NaN check

Measured on B200 boxes (GPT-2-XL gradients, 3.1 GB fp16, 256 MB buckets, whole bucket path):
    2 GPUs:  NCCL 7.87 ms   peer-to-peer 6.56 ms   multicast 9.47 ms
    8 GPUs:  NCCL 9.89 ms   peer-to-peer 10.95 ms  multicast 8.79 ms
so the default ("auto") uses peer loads/stores for 2 ranks and the switch multicast beyond that.
APEX_B200_SYMM=0 keeps NCCL; if symmetric memory cannot be set up the NCCL path is used with a warning.
"""
import os

import torch
import torch.distributed as dist

from .. import _C
from ..multi_tensor import TensorTable, CHUNK_STREAM, is_dense

_enabled = os.environ.get("APEX_B200_SYMM", "1") != "0"
_ctas = int(os.environ.get("APEX_B200_SYMM_CTAS", "0"))
_force_p2p = os.environ.get("APEX_B200_SYMM_P2P", "") == "1"
_force_mc = os.environ.get("APEX_B200_SYMM_P2P", "") == "0"
_broken = False


def enable(flag=True):
    global _enabled
    _enabled = bool(flag)


def enabled():
    return _enabled and not _broken and dist.is_initialized() and dist.get_backend() == "nccl" \
        and dist.get_world_size() > 1


def mark_broken(err):
    """Symmetric memory could not be set up on this box: stay on NCCL from now on."""
    global _broken
    _broken = True
    import warnings
    warnings.warn("apex_b200: symmetric-memory allreduce unavailable (%s); using NCCL" % (err,))


class _Arena(object):
    """One symmetric-memory allocation that many buckets are carved out of: a single rendezvous
    (collective, milliseconds) serves hundreds of buckets.  Sizes double, so a model needs a handful."""

    MIN_BYTES = 64 << 20
    ALIGN = 256

    def __init__(self, nbytes, device):
        import torch.distributed._symmetric_memory as symm_mem
        self.nbytes = nbytes
        self.buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, dist.group.WORLD)      # same order on every rank
        self.used = 0

    def take(self, nbytes):
        start = (self.used + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        if start + nbytes > self.nbytes:
            return None
        self.used = start + nbytes
        return start


_arenas = {}      # device index -> list of arenas (buckets of every DDP instance of the process share them)


def _carve(nbytes, device):
    arenas = _arenas.setdefault(device.index, [])
    for arena in arenas:
        off = arena.take(nbytes)
        if off is not None:
            return arena, off
    size = max(_Arena.MIN_BYTES, nbytes, 2 * arenas[-1].nbytes if arenas else 0)
    arena = _Arena(size, device)
    arenas.append(arena)
    return arena, arena.take(nbytes)


class SymmSlot(object):
    """A flat bucket living in symmetric memory + the table of the tensors it stands for."""

    def __init__(self):
        self.table = None
        self.flat = None          # padded slice of an arena, typed
        self.view = None          # first `total` elements
        self.handle = None
        self.offset = 0           # byte offset of the slice inside the arena
        self.total = 0

    def bind(self, tensors):
        head = tensors[0]
        if head.dtype not in (torch.float16, torch.float32):
            raise TypeError("apex_b200: symmetric-memory buckets are fp16 or fp32")
        total = 0
        for t in tensors:
            if t.dtype != head.dtype or not is_dense(t):
                raise RuntimeError("apex_b200: bucket tensors must be dense and of one type")
            total += t.numel()
        world = dist.get_world_size()
        if self.flat is None or total != self.total or self.flat.dtype != head.dtype:
            unit = (16 // head.element_size()) * world
            padded = (total + unit - 1) // unit * unit
            arena, off = _carve(padded * head.element_size(), head.device)
            self.flat = arena.buf[off:off + padded * head.element_size()].view(head.dtype)
            self.handle = arena.handle
            self.offset = off
            self.view = self.flat[:total]
            self.total = total
        if self.table is None:
            self.table = TensorTable(head.device, 1, CHUNK_STREAM)
        self.table.update([tensors])
        return self.view

    def barrier(self, channel=0):
        self.handle.barrier(channel=channel, timeout_ms=20000)

    def launch_allreduce(self):
        """This rank's half of the two-shot SUM; the caller provides the barriers."""
        h = self.handle
        mc = int(h.multicast_ptr or 0)
        use_mc = mc != 0 and not _force_p2p and (_force_mc or h.world_size > 2)
        ctas = _ctas if _ctas > 0 else (32 if use_mc else 148)
        peers = [int(p) + self.offset for p in h.buffer_ptrs]
        _C.symm_allreduce(peers, mc + self.offset if use_mc else 0, h.rank, h.world_size, self.flat.numel(),
                          self.flat.dtype, ctas)

    def allreduce(self):
        self.barrier(0)
        self.launch_allreduce()
        self.barrier(1)


class SymmPipeline(object):
    """gather / allreduce / scatter of consecutive buckets overlapped.

    Buckets alternate between two collective streams.  On its stream a bucket runs
        wait(gathered) -> barrier "every rank has gathered b" -> allreduce kernel -> barrier "every rank
        has reduced b" -> event for the scatter
    so the barriers of bucket b+1 (tiny kernels that spin until all ranks arrive) execute while the
    allreduce kernel of bucket b is still moving data: the NVLink-bound kernels run back to back
    instead of being separated by a cross-rank round trip.  Barriers in flight at the same time use
    different signal-pad channels (2 per stream).  The allreduce kernels never wait for a peer, so
    there is no circular wait.  APEX_B200_SYMM_STREAMS=1 serialises everything on one stream."""

    def __init__(self, gather_stream, scatter_stream):
        self.gather_stream = gather_stream
        self.scatter_stream = scatter_stream
        n = 1 if os.environ.get("APEX_B200_SYMM_STREAMS", "2") == "1" else 2
        self.comm_streams = [torch.cuda.Stream() for _ in range(n)]
        self.count = 0

    def submit(self, slot, bucket, flag):
        view = slot.bind(bucket)
        _C.mt_flatten(slot.table, 0, view)            # on the gather stream (current)
        gathered = torch.cuda.Event()
        gathered.record()
        k = self.count % len(self.comm_streams)
        self.count += 1
        cs = self.comm_streams[k]
        with torch.cuda.stream(cs):
            cs.wait_event(gathered)
            slot.barrier(2 * k)
            slot.launch_allreduce()
            slot.barrier(2 * k + 1)
            reduced = torch.cuda.Event()
            reduced.record()
        with torch.cuda.stream(self.scatter_stream):
            self.scatter_stream.wait_event(reduced)
            _C.mt_scale_unflatten(slot.table, view, 0, dist.get_world_size(), True, flag)

    def join(self, stream):
        for cs in self.comm_streams:
            stream.wait_stream(cs)


def average_bucket(slot, bucket, flag=None, comm_stream=None, scatter_stream=None):
    """gather -> symmetric allreduce -> x 1/world + scatter + check.

    With ``comm_stream``/``scatter_stream`` the three stages run on three streams chained by events,
    so that consecutive buckets overlap (gather of b+1, allreduce of b, scatter of b-1); without
    them everything is enqueued on the current stream."""
    view = slot.bind(bucket)
    _C.mt_flatten(slot.table, 0, view)
    world = dist.get_world_size()
    if comm_stream is None:
        slot.allreduce()
        _C.mt_scale_unflatten(slot.table, view, 0, world, True, flag)
        return
    gathered = torch.cuda.Event()
    gathered.record()
    with torch.cuda.stream(comm_stream):
        comm_stream.wait_event(gathered)
        slot.allreduce()
        reduced = torch.cuda.Event()
        reduced.record()
    with torch.cuda.stream(scatter_stream):
        scatter_stream.wait_event(reduced)
        _C.mt_scale_unflatten(slot.table, view, 0, world, True, flag)
