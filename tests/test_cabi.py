"""The C ABI: the library loads, exports exactly what include/apex_b200.h declares, and
its host-only entry points (no GPU needed) behave.  No compute call is made here."""
import ctypes
import os
import re
import struct
import subprocess

import numpy as np
import pytest
import torch

from apex_b200 import _C
from tests.conftest import ROOT

HEADER = os.path.join(ROOT, "include", "apex_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(apex_b200_[a-z0-9_]+)\s*\(", text)))


def test_header_and_library_agree():
    declared = declared_symbols()
    assert len(declared) >= 20
    out = subprocess.check_output(["nm", "-D", "--defined-only", _C.LIB_PATH]).decode()
    exported = sorted(set(re.findall(r"\bT (apex_b200_[a-z0-9_]+)", out)))
    assert exported == declared
    # and the python binding binds every one of them
    assert sorted(_C.SIGNATURES) == declared
    for name in declared:
        assert getattr(_C.lib, name) is not None


def test_every_declaration_cites_the_reference():
    text = open(HEADER).read()
    # each entry point that replaces reference behaviour names the file it replaces
    names = declared_symbols()
    starts = sorted((text.rindex(" " + n + "("), n) for n in names)
    for k, (at, name) in enumerate(starts):
        if name in ("apex_b200_version", "apex_b200_error_string", "apex_b200_device_info",
                    "apex_b200_mt_count_chunks", "apex_b200_mt_table_bytes", "apex_b200_mt_table_pack",
                    "apex_b200_flag_clear"):
            continue
        # everything between the previous declaration and this one (comment + any struct)
        comment = text[starts[k - 1][0]:at]
        assert "apex/" in comment or "No counterpart in the reference" in comment, name


def test_version_and_error_strings():
    assert _C.version() == 100
    assert _C.error_string(0) == "ok"
    assert "invalid argument" in _C.error_string(-1)
    assert "maximum" in _C.error_string(-2)
    assert "too small" in _C.error_string(-3)
    assert isinstance(_C.error_string(700), str)       # cudaErrorIllegalAddress


def _pack(numels, n_lists, ptrs, flags, chunk):
    n = len(numels)
    # keep the arrays alive for the duration of the calls (ctypes.data of a temporary dangles)
    a_numel = np.asarray(numels, np.int64)
    a_ptrs = np.asarray(ptrs, np.uint64)
    a_flags = None if flags is None else np.asarray(flags, np.int32)
    nbytes = _C.table_bytes(n, n_lists, _C.count_chunks(n, a_numel.ctypes.data if n else None, chunk))
    buf = np.zeros(nbytes, np.uint8)
    st = _C.table_pack(buf.ctypes.data, nbytes, n, n_lists,
                       a_numel.ctypes.data if n else None, a_ptrs.ctypes.data if n else None,
                       None if a_flags is None else a_flags.ctypes.data, chunk)
    return st, buf


def test_table_pack_layout_ragged_and_empty():
    numels = [5, 0, 2048 * 8 + 3, 16384, 0, 1]
    n, L, chunk = len(numels), 2, 16384
    ptrs = [4096 * (i + 1) for i in range(n)] + [4096 * 100 + 32 * i for i in range(n)]
    ptrs[2] += 16                       # not 32 B aligned -> TF_UNALIGNED on tensor 2
    flags = [0, 1, 2, 4, 0, 16]
    st, buf = _pack(numels, L, ptrs, flags, chunk)
    assert (st.n_tensors, st.n_lists, st.chunk_elems) == (n, L, chunk)
    header = np.frombuffer(buf, np.int32, 4, 0)
    off = 16
    t_numel = np.frombuffer(buf, np.int64, n, off); off += 8 * n
    t_start = np.frombuffer(buf, np.int64, n + 1, off); off += 8 * (n + 1)
    t_ptr = np.frombuffer(buf, np.int64, n * L, off); off += 8 * n * L
    t_flags = np.frombuffer(buf, np.int32, n, off); off += 4 * n
    off = (off + 7) & ~7
    t_chunk = np.frombuffer(buf, np.int32, n + 1, off)
    off = (off + 4 * (n + 1) + 7) & ~7
    t_owner = np.frombuffer(buf, np.int32, st.total_chunks, off)
    assert t_owner.tolist() == [0, 2, 2, 3, 5]                 # tensor of every chunk
    assert t_numel.tolist() == numels
    assert t_start.tolist() == [0, 5, 5, 5 + 16387, 5 + 16387 + 16384, 5 + 16387 + 16384, 5 + 16387 + 16384 + 1]
    assert t_ptr.tolist() == ptrs
    chunks = [1, 0, 2, 1, 0, 1]
    assert t_chunk.tolist() == [0, 1, 1, 3, 4, 4, 5]
    assert st.total_chunks == sum(chunks) == header[3]
    want_flags = list(flags)
    want_flags[2] |= _C.TF_UNALIGNED
    assert t_flags.tolist() == want_flags
    assert header.tolist() == [n, L, chunk, 5]


def test_table_pack_rejects_bad_arguments():
    buf = np.zeros(4096, np.uint8)
    one = np.array([8], np.int64)
    ptr = np.array([256], np.uint64)
    out = _C.MtTableStruct()
    call = _C.lib.apex_b200_mt_table_pack
    ok = call(buf.ctypes.data, 4096, 1, 1, one.ctypes.data, ptr.ctypes.data, None, 2048, ctypes.byref(out))
    assert ok == 0 and out.total_chunks == 1
    assert call(buf.ctypes.data, 4096, 1, 1, one.ctypes.data, ptr.ctypes.data, None, 1000, ctypes.byref(out)) == -1
    assert call(buf.ctypes.data, 8, 1, 1, one.ctypes.data, ptr.ctypes.data, None, 2048, ctypes.byref(out)) == -3
    assert call(None, 4096, 1, 1, one.ctypes.data, ptr.ctypes.data, None, 2048, ctypes.byref(out)) == -1
    neg = np.array([-1], np.int64)
    assert call(buf.ctypes.data, 4096, 1, 1, neg.ctypes.data, ptr.ctypes.data, None, 2048, ctypes.byref(out)) == -1
    # empty table is fine
    assert call(buf.ctypes.data, 4096, 0, 1, None, None, None, 2048, ctypes.byref(out)) == 0
    assert out.total_chunks == 0


def test_null_arguments_are_rejected_without_touching_cuda():
    lib = _C.lib
    assert lib.apex_b200_mt_check_overflow(None, 0, None, None) == -1
    assert lib.apex_b200_mt_fused_sgd(None, None, None, None) == -1
    assert lib.apex_b200_mt_fused_adam(None, None, None, None) == -1
    assert lib.apex_b200_flag_clear(None, None) == -1
    t = _C.MtTableStruct(None, 0, 1, 2048, 0)
    # an empty table is a successful no-op for every kernel entry point
    assert lib.apex_b200_mt_unscale(ctypes.byref(t), 0, 1.0, None, 0, None) == 0
    assert lib.apex_b200_mt_zero(ctypes.byref(t), 0, 2, None) == 0
    assert lib.apex_b200_mt_unscale(ctypes.byref(t), 3, 1.0, None, 0, None) == -1     # list out of range
    assert lib.apex_b200_scale_flat(None, 0, 0, 2.0, None, None) == 0
    assert lib.apex_b200_scale_flat(None, 10, 7, 2.0, None, None) == -1


def test_product_fails_loudly_without_the_library(tmp_path):
    """No CPU fallback: importing the binding without the .so is an ImportError."""
    code = ("import importlib.util, sys; sys.path.insert(0, %r);"
            "import apex_b200._C as c" % ROOT)
    # run in a subprocess with LIB_PATH pointing nowhere
    probe = ("import sys; sys.path.insert(0, %r)\n"
             "import os, importlib\n"
             "real = os.path.exists\n"
             "os.path.exists = lambda p: False if p.endswith('libapex_b200.so') else real(p)\n"
             "try:\n"
             "    import apex_b200._C\n"
             "    print('IMPORTED')\n"
             "except ImportError as e:\n"
             "    print('IMPORTERROR', 'no python or CPU' in str(e))\n" % ROOT)
    out = subprocess.check_output(["python", "-c", probe]).decode()
    assert "IMPORTERROR True" in out


def test_table_pack_properties_on_random_tensor_lists():
    """For any list of sizes: chunks tile every tensor exactly, in order, and never straddle two tensors."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(st.lists(st.integers(min_value=0, max_value=300000), min_size=0, max_size=40),
           st.sampled_from([2048, 4096, 16384, 65536]), st.integers(min_value=1, max_value=5))
    def check(numels, chunk, n_lists):
        n = len(numels)
        ptrs = [32 * (7 + i + l * 1000) for l in range(n_lists) for i in range(n)]
        st_, buf = _pack(numels, n_lists, ptrs, None, chunk)
        off = 16 + 8 * n + 8 * (n + 1) + 8 * n * n_lists + 4 * n
        off = (off + 7) & ~7
        t_chunk = np.frombuffer(buf, np.int32, n + 1, off)
        off = (off + 4 * (n + 1) + 7) & ~7
        owner = np.frombuffer(buf, np.int32, st_.total_chunks, off)
        want_chunks = [-(-x // chunk) for x in numels]
        assert st_.total_chunks == sum(want_chunks)
        assert t_chunk.tolist() == [sum(want_chunks[:i]) for i in range(n + 1)]
        assert owner.tolist() == [i for i, c in enumerate(want_chunks) for _ in range(c)]
        # every chunk lies inside its tensor and the chunks of a tensor cover it exactly once
        covered = [0] * n
        for c, t in enumerate(owner.tolist()):
            start = (c - t_chunk[t]) * chunk
            assert 0 <= start < numels[t]
            covered[t] += min(chunk, numels[t] - start)
        assert covered == list(numels)
        assert len(buf) == _C.table_bytes(n, n_lists, st_.total_chunks)
    check()


def test_table_identity_fast_path_is_voided_when_a_tensor_dies():
    """TensorTable.built_from: same objects -> True; a different object, or a remembered tensor that has
    been freed since (its id may be reused), -> False."""
    import gc
    from apex_b200.multi_tensor import TensorTable
    a, b, c = torch.zeros(8), torch.zeros(8), torch.zeros(8)
    t = TensorTable("cpu", 1, 2048)
    t.update([[a, b]])
    t.remember([a, b])
    assert t.built_from([a, b]) and not t.built_from([a, c]) and not t.built_from([a]) and not t.built_from([b, a])
    assert t.update([[a, b]]) is False                    # same addresses: no rebuild
    del b
    gc.collect()
    assert t._alive is False and not t.built_from([a, c])
    # a table rebuilt afterwards is not affected by the death of tensors of its previous generation
    d = torch.zeros(8)
    t.update([[a, d]])
    t.remember([a, d])
    del c
    gc.collect()
    assert t.built_from([a, d])
    t.invalidate()
    assert not t.built_from([a, d])


def test_symm_allreduce_validates_before_any_cuda_call():
    """apex_b200_symm_allreduce: arguments that cannot describe a collective are refused on the host."""
    import ctypes
    lib = _C.lib
    d = _C.SymmDomain()
    d.rank, d.world = 0, 2
    d.peer_base[0], d.peer_base[1] = 1 << 20, 2 << 20

    def call(dom, offset=0, numel=32, dtype=0, divisor=2.0, apply_mul=1, flag_off=-1, channel=0, seq=1, mc=0, ctas=0):
        return lib.apex_b200_symm_allreduce(ctypes.byref(dom) if dom is not None else None, offset, numel, dtype,
                                            divisor, apply_mul, flag_off, channel, seq, mc, ctas, None)
    assert call(None) == -1
    assert call(d, numel=24) == -1                      # not a multiple of 8 halfs x 2 ranks
    assert call(d, offset=8) == -1                      # bucket not 16 B aligned
    assert call(d, dtype=2) == -1                       # fp64 is not supported
    assert call(d, seq=0) == -1                         # seq starts at 1
    assert call(d, channel=_C.SYMM_CHANNELS) == -1
    assert call(d, divisor=0.0) == -1
    assert call(d, mc=1) == -1                          # no multicast alias in this domain
    assert call(d, flag_off=6) == -1
    assert call(d, ctas=_C.SYMM_MAX_CTAS + 1) == -2     # APEX_B200_ERANGE
    assert call(d, numel=0) == 0                        # nothing to do
    bad = _C.SymmDomain()
    bad.rank, bad.world = 2, 2
    assert call(bad) == -1
    # the python-side constants mirror the header
    header = open(HEADER).read()
    assert "#define APEX_B200_SYMM_MAX_WORLD %d" % _C.SYMM_MAX_WORLD in header
    assert "#define APEX_B200_SYMM_MAX_CTAS %d" % _C.SYMM_MAX_CTAS in header
    assert "#define APEX_B200_SYMM_CHANNELS %d" % _C.SYMM_CHANNELS in header


def test_struct_layouts_match_the_header(tmp_path):
    """The ctypes mirrors in apex_b200/_C.py against what a C compiler makes of include/apex_b200.h."""
    import ctypes
    src = os.path.join(str(tmp_path), "layout.c")
    exe = os.path.join(str(tmp_path), "layout")
    fields = {
        "apex_b200_mt_table": ["d_table", "n_tensors", "n_lists", "chunk_elems", "total_chunks"],
        "apex_b200_sgd_args": ["lr", "momentum", "dampening", "weight_decay", "nesterov", "apply_scale", "inv_scale",
                               "clip_norm", "clip_max_norm"],
        "apex_b200_adam_args": ["lr", "beta1", "beta2", "eps", "weight_decay", "step_size", "bias_correction2_sqrt",
                                "decoupled_weight_decay", "apply_scale", "inv_scale", "clip_norm", "clip_max_norm"],
        "apex_b200_symm_domain": ["rank", "world", "peer_base", "multicast_base", "signal_offset", "timeout_ns"],
    }
    mirrors = {"apex_b200_mt_table": _C.MtTableStruct, "apex_b200_sgd_args": _C.SgdArgs,
               "apex_b200_adam_args": _C.AdamArgs, "apex_b200_symm_domain": _C.SymmDomain}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "apex_b200.h"', 'int main(void) {']
    for name, fs in fields.items():
        lines.append('printf("%s %%zu", sizeof(%s));' % (name, name))
        for f in fs:
            lines.append('printf(" %%zu", offsetof(%s, %s));' % (name, f))
        lines.append('printf("\\n");')
    lines += ['printf("signal_bytes %d\\n", APEX_B200_SYMM_SIGNAL_BYTES);', 'return 0; }']
    with open(src, "w") as f:
        f.write("\n".join(lines))
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), src, "-o", exe])   # plain C: the header is C
    out = subprocess.check_output([exe]).decode().splitlines()
    for line in out[:-1]:
        parts = line.split()
        mirror = mirrors[parts[0]]
        assert ctypes.sizeof(mirror) == int(parts[1]), parts[0]
        assert [getattr(mirror, f).offset for f, _ in mirror._fields_] == [int(x) for x in parts[2:]], parts[0]
        assert [f for f, _ in mirror._fields_] == fields[parts[0]]
    assert int(out[-1].split()[1]) == _C.SYMM_SIGNAL_BYTES
