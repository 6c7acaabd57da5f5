"""apex_b200_mt_l2norm_noalloc: the gradient norm for callers that cannot allocate scratch.

Who needs it: a clip_master_grads issued when the framework's memory pool has no room left (a
cudaMalloc there can fail, or flush the pool and synchronise the device), and code that runs while a
bucket allreduce is in flight on another stream (cudaMalloc/cudaFree wait for the whole device).  The
entry point keeps its per-chunk partial sums in a fixed 256 KiB array inside the library instead of a
caller-provided buffer.  It must return exactly what apex_b200_mt_l2norm returns.

This file is named to be collected last, so that a `-x` run has already been through every other test.
"""
import ctypes

import pytest
import torch

from apex_b200 import _C
from tests.gpu_util import SIZES, dev, gen, randn16, randn32, table

pytestmark = pytest.mark.gpu


def _grads(seed, sizes):
    g = gen(seed)
    grads, flags = [], []
    for i, n in enumerate(sizes):
        if i % 3 == 1:
            grads.append(randn32(n, g, 3.0))
            flags.append(_C.TF_GRAD_F32)
        else:
            grads.append(randn16(n, g, 40.0))
            flags.append(0)
    return grads, flags


def _both(t, inv_scale, apply_scale):
    """(norm from the existing entry point with caller scratch, norm from the no-alloc entry point)."""
    partials = torch.zeros(max(t.struct.total_chunks, 1), device=dev())
    a = torch.full((1,), -1.0, device=dev())
    b = torch.full((1,), -2.0, device=dev())
    _C.mt_l2norm(t, 0, inv_scale, apply_scale, partials, a)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = _C.lib.apex_b200_mt_l2norm_noalloc(t.ref, 0, ctypes.c_double(inv_scale), int(apply_scale),
                                            ctypes.c_void_p(b.data_ptr()), stream)
    assert rc == 0, _C.error_string(rc)
    torch.cuda.synchronize()
    return float(a.item()), float(b.item())


def test_no_alloc_norm_matches_on_two_calls_in_a_row():
    # call 1: many chunks (small chunk size), mixed fp16/fp32, scaled
    grads, flags = _grads(41, SIZES)
    t1 = table([grads], flags, 2048)
    assert t1.struct.total_chunks > 700
    a1, b1 = _both(t1, 1. / 512, True)
    assert a1 > 0 and abs(a1 - b1) <= 2e-6 * a1          # the tolerance the norm tests use ...
    assert a1 == b1                                      # ... and in fact the same bits (same reduction order)
    # call 2 right after, fewer chunks and other data: slots left over from call 1 must not leak in
    grads2, flags2 = _grads(42, [5, 70001, 2048, 0, 333])
    t2 = table([grads2], flags2, 65536)
    assert t2.struct.total_chunks < t1.struct.total_chunks
    a2, b2 = _both(t2, 1.0, False)
    assert a2 > 0 and a2 != a1
    assert abs(a2 - b2) <= 2e-6 * a2 and a2 == b2
    # and the python wrapper goes to the same entry point
    out = torch.zeros(1, device=dev())
    _C.mt_l2norm_noalloc(t2, 0, 1.0, False, out)
    assert float(out.item()) == a2
    # empty table: norm 0
    t0 = table([[]])
    out.fill_(7.0)
    _C.mt_l2norm_noalloc(t0, 0, 1.0, False, out)
    assert float(out.item()) == 0.0


def test_no_alloc_norm_rejects_sizes_beyond_its_fixed_array():
    out = torch.zeros(1, device=dev())
    grads, flags = _grads(43, [4096])
    t = table([grads], flags, 2048)
    too_big = _C.MtTableStruct(t.struct.d_table, 1, 1, 2048, _C.NOALLOC_MAX_CHUNKS + 1)
    rc = _C.lib.apex_b200_mt_l2norm_noalloc(ctypes.byref(too_big), 0, ctypes.c_double(1.0), 0,
                                            ctypes.c_void_p(out.data_ptr()), None)
    assert rc == -2
    assert "beyond the supported maximum" in _C.error_string(rc)
    assert _C.lib.apex_b200_mt_l2norm_noalloc(t.ref, 0, ctypes.c_double(1.0), 0, None, None) == -1
    assert _C.lib.apex_b200_mt_l2norm_noalloc(t.ref, 5, ctypes.c_double(1.0), 0,
                                              ctypes.c_void_p(out.data_ptr()), None) == -1
