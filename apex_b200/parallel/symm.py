"""Symmetric (peer-mapped) memory for the gradient buckets and their averaging without NCCL.

torch.distributed's symmetric memory is used for plumbing only: allocating a buffer that every rank
of the NVLink domain can address on every other rank (and, on NVSwitch, through a multicast alias of
all copies), and exchanging the addresses.  Everything that happens per bucket is ours
(``apex_b200_symm_allreduce``, csrc/symm_allreduce.cu): ONE kernel per rank that waits for its peers,
sums the bucket across ranks (switch-side ``multimem`` or peer loads/stores), multiplies by 1/world,
tests the result for inf/NaN and waits for its peers again -- where the reference runs
``all_reduce`` + ``/=`` (apex/parallel/distributed.py:15-20) and, later, a scan of the averaged
gradients (apex/fp16_utils/fp16_optimizer.py:213-221).

Layout of one allocation (``Arena``):
    [0, 64 KiB)        signal words of the in-kernel barriers  [channel][cta][rank] uint32
    [96 KiB, +256 B)   overflow words (int32), one per user of the arena
    [128 KiB, ...)     buckets, 256 B aligned

APEX_B200_SYMM=0 keeps NCCL; if symmetric memory cannot be set up the NCCL path is used with a warning.
"""
import os

import torch
import torch.distributed as dist

from .. import _C

_enabled = os.environ.get("APEX_B200_SYMM", "1") != "0"
_ctas = int(os.environ.get("APEX_B200_SYMM_CTAS", "0"))
_force_p2p = os.environ.get("APEX_B200_SYMM_P2P", "") == "1"
_force_mc = os.environ.get("APEX_B200_SYMM_P2P", "") == "0"
_broken = False
_single_host = None

HEADER_BYTES = 128 << 10
FLAG_BASE = 96 << 10
MAX_FLAGS = 64
ALIGN = 256


def enable(flag=True):
    global _enabled
    _enabled = bool(flag)


def _one_nvlink_domain():
    """Symmetric memory spans one host.  torchrun exports LOCAL_WORLD_SIZE; without it (processes
    spawned by hand) every rank must have its own GPU on this host."""
    global _single_host
    if _single_host is None:
        world = dist.get_world_size()
        local = os.environ.get("LOCAL_WORLD_SIZE")
        if local is not None:
            _single_host = int(local) == world
        else:
            _single_host = world <= torch.cuda.device_count()
    return _single_host


def enabled():
    return _enabled and not _broken and dist.is_initialized() and dist.get_backend() == "nccl" \
        and dist.get_world_size() > 1 and dist.get_world_size() <= _C.SYMM_MAX_WORLD and _one_nvlink_domain()


def mark_broken(err):
    """Symmetric memory could not be set up on this box: stay on NCCL from now on."""
    global _broken
    _broken = True
    import warnings
    warnings.warn("apex_b200: symmetric-memory allreduce unavailable (%s); using NCCL" % (err,))


def comm_stream_count():
    """How many collective streams / signal channels consecutive buckets rotate over
    (APEX_B200_SYMM_STREAMS, default 4, at most the ABI's channel count); 1 = the submitting stream."""
    return max(1, min(int(os.environ.get("APEX_B200_SYMM_STREAMS", "4")), _C.SYMM_CHANNELS))


def _timeout_ns():
    """How long a rank waits for a peer inside the kernel before it traps.  Default: never (a late
    rank -- checkpointing, a data-loader stall -- is waited for, as NCCL's kernels do);
    APEX_B200_SYMM_TIMEOUT_S sets a limit."""
    return int(float(os.environ.get("APEX_B200_SYMM_TIMEOUT_S", "0")) * 1e9)


class Arena(object):
    """One symmetric allocation: a single rendezvous (collective, milliseconds) serves every bucket
    carved out of it.  Must be created by all ranks in the same order."""

    def __init__(self, payload_bytes, device):
        import torch.distributed._symmetric_memory as symm_mem
        self.nbytes = HEADER_BYTES + int(payload_bytes)
        self.buf = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        # the zeros (signal words!) must be in memory before any peer can learn the address
        torch.cuda.current_stream().synchronize()
        self.handle = symm_mem.rendezvous(self.buf, dist.group.WORLD)      # same order on every rank
        h = self.handle
        d = _C.SymmDomain()
        d.rank, d.world = h.rank, h.world_size
        for r, p in enumerate(h.buffer_ptrs):
            d.peer_base[r] = int(p)
        d.multicast_base = int(h.multicast_ptr or 0)
        d.signal_offset = 0
        d.timeout_ns = _timeout_ns()
        self.domain = d
        self.world = h.world_size
        self.has_multicast = d.multicast_base != 0
        self.used = HEADER_BYTES
        self.seq = [0] * _C.SYMM_CHANNELS
        self._flags = self.buf[FLAG_BASE:FLAG_BASE + 4 * MAX_FLAGS].view(torch.int32)
        self._flags_used = 0
        dist.barrier()          # every rank's zeros are in place before anybody launches

    def take(self, nbytes):
        """Byte offset of a fresh 256 B-aligned span, or None when the arena is full."""
        start = (self.used + ALIGN - 1) // ALIGN * ALIGN
        if start + nbytes > self.nbytes:
            return None
        self.used = start + nbytes
        return start

    def room(self):
        return self.nbytes - (self.used + ALIGN - 1) // ALIGN * ALIGN

    def new_flag(self):
        """(int32 tensor of one element inside the arena, its byte offset).  The word has the same
        offset on every rank; our kernel raises it everywhere."""
        if self._flags_used >= MAX_FLAGS:
            raise RuntimeError("apex_b200: out of overflow words in this symmetric arena")
        i = self._flags_used
        self._flags_used += 1
        return self._flags[i:i + 1], FLAG_BASE + 4 * i

    def use_multicast(self):
        return self.has_multicast and not _force_p2p and (_force_mc or self.world > 2)

    def rotation(self):
        """(streams, CTAs per kernel) for buckets in flight together.  Kernels that wait for their
        peers may only overlap on a GPU if ALL their CTAs are resident at once (else rank A could fill
        its SMs with bucket b while rank B fills its SMs with bucket b+1: a circular wait).  The
        multicast kernel is small (16-32 CTAs x 512 threads x 38 registers: four of them fit many
        times over); the peer-to-peer kernel needs a whole SM per CTA (512 threads x ~120 registers), so two
        of them share the 148 SMs with 74 CTAs each."""
        n = comm_stream_count()
        if self.use_multicast():
            # measured on 8 GPUs with 4 streams (3.1 GB): 16 CTAs per kernel 6.49 ms at 256 MB buckets
            # and 7.06 ms at 20 MB; 24: 6.53 / 7.33; 32: 6.58 / 7.63
            return n, (_ctas if _ctas > 0 else (16 if n >= 4 else 32))
        if n >= 2 and _ctas <= 0:
            return 2, 74
        return 1, (_ctas if _ctas > 0 else 148)

    def allreduce(self, offset, numel, dtype, divisor, flag_offset, channel=0, apply_mul=True, ctas=None):
        """This rank's kernel of the averaging of [offset, offset + numel * itemsize) on the current
        stream.  Every rank must issue the same sequence of calls per channel."""
        self.seq[channel] += 1
        mc = self.use_multicast()
        if ctas is None:
            ctas = _ctas if _ctas > 0 else (32 if mc else 148)
        _C.symm_allreduce(self.domain, offset, numel, dtype, divisor, apply_mul,
                          -1 if flag_offset is None else flag_offset, channel, self.seq[channel], mc, ctas)


def padded_numel(numel, element_size, world):
    """Bucket length our kernel accepts: every rank's slice is a whole number of 16 B vectors."""
    unit = (16 // element_size) * world
    return (numel + unit - 1) // unit * unit
