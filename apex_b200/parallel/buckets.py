"""Persistent flat homes for gradient buckets: the gradients LIVE in the bucket.

The reference builds a fresh coalesced copy of every bucket, reduces it and copies the result back
tensor by tensor, every iteration (apex/parallel/distributed.py:11-23: ``_flatten_dense_tensors``,
``all_reduce``, ``/=``, n x ``copy_``).  Here a bucket owns one flat buffer for its whole life and
hands out a view per gradient; ``DistributedDataParallel`` makes those views the ``.grad`` of its
parameters, so autograd accumulates straight into the bucket, the collective runs in place and there
is nothing to gather or scatter.  A gradient that is *not* its view when the
bucket is due (the user set ``.grad`` to None, so autograd produced a fresh tensor) is copied in with
one multi-tensor launch and, for DDP, replaced by its view afterwards -- no copy back either.

Inside the flat buffer every tensor starts on a 32-byte boundary (the 256-bit vector path of the
kernels) and the total is padded so that every rank's slice starts on a 2 KiB boundary (a warp's
512-byte request over NVLink must not straddle 128-byte lines: 16-byte-aligned slices cost 10 % of the
8-GPU bucket path); padding stays zero.
The flat layout is internal (the reference's coalesced buffer is a temporary nobody sees).
"""
import torch

from .. import _C
from ..multi_tensor import TensorTable, CHUNK_STREAM, is_dense

TENSOR_ALIGN_BYTES = 32
SLICE_ALIGN_BYTES = 2048


def bucket_layout(tensors, world):
    """(element offsets, padded total) of ``tensors`` in a flat buffer of their dtype."""
    es = tensors[0].element_size()
    step = TENSOR_ALIGN_BYTES // es
    offsets, at = [], 0
    for t in tensors:
        offsets.append(at)
        at += (t.numel() + step - 1) // step * step
    # every rank's slice of the bucket starts on a SLICE_ALIGN_BYTES boundary: NVLink/NVSwitch requests
    # of a warp (512 B) then never straddle 128 B lines (and the C ABI's rule, a whole number of
    # 16 B vectors per rank, is implied)
    unit = (SLICE_ALIGN_BYTES // es) * max(world, 1)
    unit = unit * step // _gcd(unit, step)
    total = (at + unit - 1) // unit * unit
    return offsets, max(total, unit)


def _gcd(a, b):
    while b:
        a, b = b, a % b
    return a


class GradBucket(object):
    """One bucket: ``flat`` (padded, zero-initialised, caller-provided storage or torch.zeros) and
    ``views[i]``, a tensor of the shape/strides of member i that aliases its span of ``flat``."""

    def __init__(self, like, world, storage=None):
        head = like[0]
        for t in like:
            if t.dtype != head.dtype or t.device != head.device:
                raise TypeError("apply_flat_dist_call needs tensors of one type, got %s and %s"
                                % (head.type(), t.type()))
            if not is_dense(t):
                raise RuntimeError("apex_b200: bucket tensors must be dense (contiguous or channels-last)")
        self.dtype, self.device = head.dtype, head.device
        self.offsets, self.padded = bucket_layout(like, world)
        self.numel = sum(t.numel() for t in like)
        if storage is None:
            storage = torch.zeros(self.padded, dtype=self.dtype, device=self.device)
        assert storage.numel() == self.padded and storage.dtype == self.dtype
        self.flat = storage
        self.signature = [(tuple(t.shape), tuple(t.stride())) for t in like]
        # `.data`: a view with its OWN version counter -- autograd accumulating into one gradient
        # must not look like a modification of all the others (parallel/distributed.py: verified_flag)
        self.views = [storage[o:o + t.numel()].as_strided(t.shape, t.stride(), storage.storage_offset() + o).data
                      for o, t in zip(self.offsets, like)]
        self._view_ptrs = [v.data_ptr() for v in self.views]
        self._in = None        # (key, table) of the last copy-in
        self._out = None

    @staticmethod
    def nbytes_for(like, world):
        return bucket_layout(like, world)[1] * like[0].element_size()

    def fits(self, tensors):
        return len(tensors) == len(self.signature) and tensors[0].dtype == self.dtype and \
            all((tuple(t.shape), tuple(t.stride())) == s for t, s in zip(tensors, self.signature))

    def strays(self, tensors):
        """Indices of the tensors that do not live in this bucket."""
        views, ptrs = self.views, self._view_ptrs
        return [i for i, t in enumerate(tensors) if t is not views[i] and t.data_ptr() != ptrs[i]]

    def _table(self, slot, tensors, idx, src_first):
        key = tuple(idx)
        cached = getattr(self, slot)
        table = cached[1] if cached is not None else TensorTable(self.device, 2, CHUNK_STREAM)
        outside = [tensors[i] for i in idx]
        inside = [self.views[i] for i in idx]
        table.update([outside, inside] if src_first else [inside, outside])
        setattr(self, slot, (key, table))
        return table

    def copy_in(self, tensors, idx):
        """tensors[i] -> views[i] for i in idx: one launch on the current stream."""
        if idx:
            _C.mt_copy(self._table("_in", tensors, idx, True), 0, 1, self.flat.element_size(), None)

    def copy_out(self, tensors, idx):
        """views[i] -> tensors[i] for i in idx."""
        if idx:
            _C.mt_copy(self._table("_out", tensors, idx, False), 0, 1, self.flat.element_size(), None)
