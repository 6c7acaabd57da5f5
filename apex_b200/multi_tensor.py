"""Device-resident tensor tables for the multi-tensor-apply kernels.

The reference walks python lists and launches one stock torch op per tensor
(e.g. apex/fp16_utils/fp16util.py:106-112, apex/fp16_utils/fp16_optimizer.py:236-241).
Here a list-of-lists of tensors is described once by a small table in HBM
(layout in include/apex_b200.h) and one launch covers all of it.  The table is
rebuilt only when a pointer, size or flag changes: callers pass the current
tensors every time and ``update`` compares their addresses with the cached key.

Ownership: torch owns every tensor; the table borrows raw addresses for the
duration of the launches the caller enqueues right after ``update`` and never
frees or retains them (SURVEY.md 8b "Ownership").
"""
import operator
import weakref

import numpy as np
import torch

from . import _C

# Elements per chunk = per CTA (one CTA per chunk; a chunk never straddles a tensor).  Measured with
# tools/kernel_bench.py on B200, GPT-2-XL shapes (GB/s of algorithmic bytes, torch copy_ = 6547):
#   fused Adam   2048: 6850   4096: 6799   16384: 6647   65536: 6517     (persistent grid: 6000)
#   fused SGD    2048: 6780   4096: 6822   16384: 6650
#   check        2048: 2777   8192: 6887   16384: 7076   65536: 7126
#   fp16<->fp32  2048: 4886   8192: 6022   16384: 6048   65536: 5832
#   flatten      2048: 5687   8192: 6729   16384: 6629   65536: 6167
#   zero         4096: 7050   16384: 7332  65536: 7141
#   l2 norm      8192: 6234   16384: 6848  65536: 6980
CHUNK_STEP = 2048       # fused update kernels: one 8-element vector per thread per CTA
CHUNK_CHECK = 65536     # read-only overflow scan
CHUNK_STREAM = 8192     # convert / copy / flatten / scatter / unscale
CHUNK_ZERO = 16384
CHUNK_NORM = 65536      # norm kernels: one fp32 partial sum per chunk


class TensorTable(object):
    """One packed table: ``n_lists`` parallel lists of ``n`` tensors each."""

    def __init__(self, device, n_lists, chunk_elems):
        self.device = torch.device(device)
        self.n_lists = n_lists
        self.chunk_elems = chunk_elems
        self.struct = _C.MtTableStruct()
        self.ref = None
        self.n = 0
        self.total_numel = 0
        self._key = None
        self._host = None      # pinned staging buffer
        self._dev = None       # device copy
        self._copied = None    # event: staging buffer free for reuse
        self.rebuilds = 0
        self._refs = None      # weak references to the tensors the table was built from
        self._ids = None
        self._ptrs = None      # their addresses when the table was built (spot-checked by built_from)
        self._cursor = 0
        self._alive = False
        self._generation = None

    # Fast path for callers that run every training step: "are these still the very same tensor
    # objects the table was built from?"  The table remembers their id()s; a weak reference per tensor
    # (the table must not keep gradients alive that the user has dropped) carries a callback that
    # voids the table the moment one of them dies -- an id can only be reused after that -- so the
    # per-step test is one C-speed pass over the ids: microseconds for hundreds of tensors, where
    # update() fetches every address again.  Identity alone would miss a tensor that keeps its python
    # object but gets new storage (`p.data = ...`, `module.half()`, `grad.set_(...)`): the addresses of
    # the first, the last and one rotating tensor are therefore compared on every call -- wholesale
    # swaps are caught at once, a single swapped tensor within n calls.  (Reassigning `.data` of
    # individual tensors between steps is outside what the fast path promises; `invalidate()` or any
    # changed object identity forces the full comparison of `update`.)
    def built_from(self, objs):
        if not (self._alive and self._key is not None and self._ids == list(map(id, objs))):
            return False
        ptrs = self._ptrs
        if ptrs:
            n = len(ptrs)
            k = self._cursor = (self._cursor + 1) % n
            if objs[0].data_ptr() != ptrs[0] or objs[-1].data_ptr() != ptrs[-1] or objs[k].data_ptr() != ptrs[k]:
                self.invalidate()
                return False
        return True

    def remember(self, objs):
        self._ids = list(map(id, objs))
        self._ptrs = [o.data_ptr() for o in objs]
        self._cursor = 0
        self._alive = True
        token = self._generation = object()
        me = weakref.ref(self)

        def died(_ref, token=token, me=me):
            table = me()
            if table is not None and table._generation is token:
                table._alive = False
        self._refs = [weakref.ref(o, died) for o in objs]

    def update(self, lists, flags=None):
        """Point the table at ``lists`` (n_lists sequences of equal length).

        Returns True when the device copy had to be rebuilt.  ``flags`` is an
        optional sequence of per-tensor APEX_B200_TF_* ints.  A ``None`` tensor
        is allowed in lists a kernel will not dereference (address 0).
        """
        assert len(lists) == self.n_lists
        n = len(lists[0])
        ptrs = np.empty(self.n_lists * n, dtype=np.uint64)
        for l, lst in enumerate(lists):
            assert len(lst) == n
            ptrs[l * n:(l + 1) * n] = [0 if t is None else t.data_ptr() for t in lst]
        numels = np.fromiter((t.numel() for t in lists[0]), dtype=np.int64, count=n)
        fl = np.zeros(n, dtype=np.int32) if flags is None else np.asarray(flags, dtype=np.int32)
        key = (ptrs.tobytes(), numels.tobytes(), fl.tobytes())
        if key == self._key:
            return False
        self._key = key
        self.n = n
        self.total_numel = int(numels.sum())
        self._pack(n, ptrs, numels, fl)
        self.rebuilds += 1
        return True

    def _pack(self, n, ptrs, numels, fl):
        nbytes = _C.table_bytes(n, self.n_lists, _C.count_chunks(n, numels.ctypes.data, self.chunk_elems))
        on_gpu = self.device.type == "cuda"
        if self._host is None or self._host.numel() < nbytes:
            cap = max(nbytes, 4096)
            self._host = torch.empty(cap, dtype=torch.uint8, pin_memory=on_gpu)
            self._dev = torch.empty(cap, dtype=torch.uint8, device=self.device) if on_gpu else self._host
            self._copied = None
        if self._copied is not None:
            self._copied.synchronize()   # previous upload still reading the staging buffer
        self.struct = _C.table_pack(self._host.data_ptr(), self._host.numel(), n, self.n_lists,
                                    numels.ctypes.data, ptrs.ctypes.data, fl.ctypes.data, self.chunk_elems)
        if on_gpu:
            self._dev[:nbytes].copy_(self._host[:nbytes], non_blocking=True)
            self._copied = torch.cuda.Event()
            self._copied.record()
        self.struct.d_table = self._dev.data_ptr()
        import ctypes
        self.ref = ctypes.byref(self.struct)

    def invalidate(self):
        self._key = None
        self._refs = None
        self._ids = None
        self._ptrs = None
        self._alive = False


def is_dense(t):
    """True when the tensor occupies one gap-free block of memory: row-major contiguous, or one of
    torch's channels-last formats (conv weights and their gradients under channels_last training).
    Elementwise kernels may then treat it as a flat array -- provided every tensor they combine it
    with has the same strides (``same_layout``)."""
    if t.is_contiguous():
        return True
    if t.dim() == 4:
        return t.is_contiguous(memory_format=torch.channels_last)
    if t.dim() == 5:
        return t.is_contiguous(memory_format=torch.channels_last_3d)
    return False


def same_layout(tensors):
    """All tensors (None entries ignored) dense and laid out identically."""
    first = None
    for t in tensors:
        if t is None:
            continue
        if not is_dense(t):
            return False
        if first is None:
            first = t
        elif t.shape != first.shape or t.stride() != first.stride():
            return False
    return True


def same_objects(a, b):
    """True when two equally long sequences hold identical objects position by position."""
    return len(a) == len(b) and all(map(operator.is_, a, b))


def new_flag(device):
    """The single device word every overflow check of one optimizer writes to."""
    return torch.zeros(1, dtype=torch.int32, device=device)
