"""apex_b200._C -- ctypes binding of libapex_b200.so (the C ABI in include/apex_b200.h).

This module is the only place the python host code touches native code.  Every
function below forwards to exactly one ``extern "C"`` entry point, converts a
non-zero return code into ``RuntimeError`` and never falls back to anything
else: if the shared library is missing the import fails with instructions to
build it (``python -c "import __graft_entry__ as g; g.build()"``).

The reference has no native layer; its intended seam is the commented-out
``from apex_C import scale_check_overflow`` at apex/amp/scaler.py:3.
"""
import ctypes
import os

import torch  # noqa: F401  (loads libcudart.so.12 into the process before our library binds to it)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libapex_b200.so")

TF_GRAD_F32 = 1
TF_FIRST_STEP = 2
TF_NO_MODEL = 4
TF_UNALIGNED = 8
TF_NO_SCALE = 16

FLAG_CLEAN = 0
INT32_MAX = 0x7FFFFFFF

DTYPE_CODE = {torch.float16: 0, torch.float32: 1, torch.float64: 2}


class MtTableStruct(ctypes.Structure):
    _fields_ = [("d_table", ctypes.c_void_p),
                ("n_tensors", ctypes.c_int32),
                ("n_lists", ctypes.c_int32),
                ("chunk_elems", ctypes.c_int32),
                ("total_chunks", ctypes.c_int32)]


class SgdArgs(ctypes.Structure):
    _fields_ = [("lr", ctypes.c_double), ("momentum", ctypes.c_double),
                ("dampening", ctypes.c_double), ("weight_decay", ctypes.c_double),
                ("nesterov", ctypes.c_int32), ("apply_scale", ctypes.c_int32),
                ("inv_scale", ctypes.c_double), ("clip_norm", ctypes.c_void_p), ("clip_max_norm", ctypes.c_double)]


class AdamArgs(ctypes.Structure):
    _fields_ = [("lr", ctypes.c_double), ("beta1", ctypes.c_double), ("beta2", ctypes.c_double),
                ("eps", ctypes.c_double), ("weight_decay", ctypes.c_double),
                ("step_size", ctypes.c_double), ("bias_correction2_sqrt", ctypes.c_double),
                ("decoupled_weight_decay", ctypes.c_int32), ("apply_scale", ctypes.c_int32),
                ("inv_scale", ctypes.c_double), ("clip_norm", ctypes.c_void_p), ("clip_max_norm", ctypes.c_double)]


SYMM_MAX_WORLD = 16
SYMM_MAX_CTAS = 256
SYMM_CHANNELS = 4
SYMM_SIGNAL_BYTES = SYMM_CHANNELS * SYMM_MAX_CTAS * SYMM_MAX_WORLD * 4


class SymmDomain(ctypes.Structure):
    _fields_ = [("rank", ctypes.c_int32), ("world", ctypes.c_int32),
                ("peer_base", ctypes.c_uint64 * SYMM_MAX_WORLD), ("multicast_base", ctypes.c_uint64),
                ("signal_offset", ctypes.c_uint64), ("timeout_ns", ctypes.c_uint64)]


_vp = ctypes.c_void_p
_i = ctypes.c_int
_d = ctypes.c_double
_tp = ctypes.POINTER(MtTableStruct)

# name -> (restype, argtypes); mirrors include/apex_b200.h one to one.
SIGNATURES = {
    "apex_b200_version": (_i, []),
    "apex_b200_error_string": (ctypes.c_char_p, [_i]),
    "apex_b200_device_info": (_i, [ctypes.POINTER(_i)] * 3),
    "apex_b200_mt_count_chunks": (ctypes.c_int64, [_i, _vp, _i]),
    "apex_b200_mt_table_bytes": (ctypes.c_size_t, [_i, _i, ctypes.c_int64]),
    "apex_b200_mt_table_pack": (_i, [_vp, ctypes.c_size_t, _i, _i, _vp, _vp, _vp, _i, _tp]),
    "apex_b200_flag_clear": (_i, [_vp, _vp]),
    "apex_b200_mt_check_overflow": (_i, [_tp, _i, _vp, _vp]),
    "apex_b200_mt_check_overflow_publish": (_i, [_tp, _i, _vp, _vp, _vp, _vp, _vp]),
    "apex_b200_mt_unscale": (_i, [_tp, _i, _d, _vp, _i, _vp]),
    "apex_b200_mt_grads_to_master": (_i, [_tp, _i, _i, _d, _i, _vp, _vp, _d, _vp]),
    "apex_b200_mt_master_to_model": (_i, [_tp, _i, _i, _vp, _vp]),
    "apex_b200_mt_copy": (_i, [_tp, _i, _i, _i, _vp, _vp]),
    "apex_b200_mt_zero": (_i, [_tp, _i, _i, _vp]),
    "apex_b200_mt_fused_sgd": (_i, [_tp, ctypes.POINTER(SgdArgs), _vp, _vp]),
    "apex_b200_mt_fused_adam": (_i, [_tp, ctypes.POINTER(AdamArgs), _vp, _vp]),
    "apex_b200_mt_flatten": (_i, [_tp, _i, _vp, _i, _vp]),
    "apex_b200_mt_scale_unflatten": (_i, [_tp, _vp, _i, _i, _d, _i, _vp, _vp]),
    "apex_b200_scale_flat": (_i, [_vp, ctypes.c_int64, _i, _d, _vp, _vp]),
    "apex_b200_symm_allreduce": (_i, [ctypes.POINTER(SymmDomain), ctypes.c_uint64, ctypes.c_int64, _i, _d, _i,
                                      ctypes.c_int64, _i, ctypes.c_uint32, _i, _i, _vp]),
    "apex_b200_mt_l2norm": (_i, [_tp, _i, _d, _i, _vp, ctypes.c_int64, _vp, _vp]),
    "apex_b200_mt_l2norm_noalloc": (_i, [_tp, _i, _d, _i, _vp, _vp]),
    "apex_b200_mt_clip_by_norm": (_i, [_tp, _i, _vp, _d, _vp, _vp]),
    "apex_b200_mt_larc": (_i, [_tp, _i, _i, _d, _d, _d, _d, _i, _vp, ctypes.c_int64, _vp, _vp]),
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "apex_b200: native library not built (%s missing). There is no python or CPU "
            "fallback for this path; build it with `make -C apex_b200/csrc` or "
            "`python -c \"import __graft_entry__ as g; g.build()\"`." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError here = header and library out of sync
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()

# Launch counter: bench.py reports how many of OUR kernels ran in the timed region.
launches = 0


def error_string(code):
    return lib.apex_b200_error_string(int(code)).decode()


def _check(code, what):
    if code != 0:
        raise RuntimeError("%s failed: %s (code %d)" % (what, error_string(code), code))


# APEX_B200_NVTX=1: every kernel wrapper below is bracketed by an NVTX range named after the entry
# point, so that a timeline (nsys) or `ncu --nvtx-include` can attribute launches to API calls.
# The reference has no tracing hooks at all.
if os.environ.get("APEX_B200_NVTX", "0") == "1":
    import functools

    def _traced(fn):
        @functools.wraps(fn)
        def wrapper(*args, **kwargs):
            torch.cuda.nvtx.range_push("apex_b200." + fn.__name__)
            try:
                return fn(*args, **kwargs)
            finally:
                torch.cuda.nvtx.range_pop()
        return wrapper
else:
    def _traced(fn):
        return fn


def version():
    return lib.apex_b200_version()


def device_info():
    sm, ma, mi = _i(), _i(), _i()
    _check(lib.apex_b200_device_info(ctypes.byref(sm), ctypes.byref(ma), ctypes.byref(mi)), "device_info")
    return sm.value, ma.value, mi.value


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def current_stream():
    """cudaStream_t of torch's current stream on the current device, as an int."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def count_chunks(n_tensors, numels_ptr, chunk_elems):
    n = lib.apex_b200_mt_count_chunks(n_tensors, numels_ptr, chunk_elems)
    if n < 0:
        _check(int(n), "mt_count_chunks")
    return int(n)


def table_bytes(n_tensors, n_lists, total_chunks):
    return lib.apex_b200_mt_table_bytes(n_tensors, n_lists, total_chunks)


def table_pack(host_ptr, host_bytes, n_tensors, n_lists, numels_ptr, ptrs_ptr, flags_ptr, chunk_elems):
    out = MtTableStruct()
    _check(lib.apex_b200_mt_table_pack(host_ptr, host_bytes, n_tensors, n_lists, numels_ptr, ptrs_ptr,
                                       flags_ptr, chunk_elems, ctypes.byref(out)), "mt_table_pack")
    return out


def flag_clear(flag, stream=None):
    _check(lib.apex_b200_flag_clear(_ptr(flag), current_stream() if stream is None else stream), "flag_clear")


def _count(table):
    global launches
    if table.struct.total_chunks > 0:
        launches += 1


@_traced
def mt_check_overflow(table, grad_list, flag, stream=None):
    _count(table)
    _check(lib.apex_b200_mt_check_overflow(table.ref, grad_list, _ptr(flag),
                                           current_stream() if stream is None else stream), "mt_check_overflow")


@_traced
def mt_check_overflow_publish(table, grad_list, flag, ticket, host_mirror, rearm, stream=None):
    """mt_check_overflow whose last CTA also writes the final word to pinned host memory and zeroes
    the word the next check will use (nothing else is enqueued between check and update)."""
    global launches
    launches += 1
    _check(lib.apex_b200_mt_check_overflow_publish(table.ref, grad_list, _ptr(flag), _ptr(ticket), _ptr(host_mirror),
                                                   _ptr(rearm), current_stream() if stream is None else stream),
           "mt_check_overflow_publish")


@_traced
def mt_unscale(table, grad_list, inv_scale, flag, mode, stream=None):
    _count(table)
    _check(lib.apex_b200_mt_unscale(table.ref, grad_list, float(inv_scale), _ptr(flag), mode,
                                    current_stream() if stream is None else stream), "mt_unscale")


@_traced
def mt_grads_to_master(table, src_list, dst_list, inv_scale, apply_scale, flag, clip=None, stream=None):
    """clip: None or (device norm tensor, max_norm)."""
    _count(table)
    _check(lib.apex_b200_mt_grads_to_master(table.ref, src_list, dst_list, float(inv_scale), int(apply_scale),
                                            _ptr(flag), None if clip is None else _ptr(clip[0]),
                                            0.0 if clip is None else float(clip[1]),
                                            current_stream() if stream is None else stream),
           "mt_grads_to_master")


@_traced
def mt_master_to_model(table, src_list, dst_list, flag, stream=None):
    _count(table)
    _check(lib.apex_b200_mt_master_to_model(table.ref, src_list, dst_list, _ptr(flag),
                                            current_stream() if stream is None else stream), "mt_master_to_model")


@_traced
def mt_copy(table, src_list, dst_list, elem_size, flag, stream=None):
    _count(table)
    _check(lib.apex_b200_mt_copy(table.ref, src_list, dst_list, elem_size, _ptr(flag),
                                 current_stream() if stream is None else stream), "mt_copy")


@_traced
def mt_zero(table, lst, elem_size, stream=None):
    _count(table)
    _check(lib.apex_b200_mt_zero(table.ref, lst, elem_size, current_stream() if stream is None else stream),
           "mt_zero")


@_traced
def mt_fused_sgd(table, lr, momentum, dampening, weight_decay, nesterov, inv_scale, apply_scale, flag,
                 clip=None, stream=None):
    _count(table)
    a = SgdArgs(float(lr), float(momentum), float(dampening), float(weight_decay), int(bool(nesterov)),
                int(bool(apply_scale)), float(inv_scale), None if clip is None else _ptr(clip[0]),
                0.0 if clip is None else float(clip[1]))
    _check(lib.apex_b200_mt_fused_sgd(table.ref, ctypes.byref(a), _ptr(flag),
                                      current_stream() if stream is None else stream), "mt_fused_sgd")


@_traced
def mt_fused_adam(table, lr, beta1, beta2, eps, weight_decay, step_size, bias_correction2_sqrt,
                  decoupled_weight_decay, inv_scale, apply_scale, flag, clip=None, stream=None):
    _count(table)
    a = AdamArgs(float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), float(step_size),
                 float(bias_correction2_sqrt), int(bool(decoupled_weight_decay)), int(bool(apply_scale)),
                 float(inv_scale), None if clip is None else _ptr(clip[0]), 0.0 if clip is None else float(clip[1]))
    _check(lib.apex_b200_mt_fused_adam(table.ref, ctypes.byref(a), _ptr(flag),
                                       current_stream() if stream is None else stream), "mt_fused_adam")


@_traced
def mt_flatten(table, src_list, flat, stream=None):
    _count(table)
    _check(lib.apex_b200_mt_flatten(table.ref, src_list, _ptr(flat), flat.element_size(),
                                    current_stream() if stream is None else stream), "mt_flatten")


@_traced
def mt_scale_unflatten(table, flat, dst_list, divisor, apply_mul, flag, stream=None):
    _count(table)
    _check(lib.apex_b200_mt_scale_unflatten(table.ref, _ptr(flat), dst_list, DTYPE_CODE[flat.dtype],
                                            float(divisor), int(bool(apply_mul)), _ptr(flag),
                                            current_stream() if stream is None else stream), "mt_scale_unflatten")


@_traced
def scale_flat(flat, divisor, flag, stream=None):
    global launches
    if flat.numel() > 0:
        launches += 1
    _check(lib.apex_b200_scale_flat(_ptr(flat), flat.numel(), DTYPE_CODE[flat.dtype], float(divisor), _ptr(flag),
                                    current_stream() if stream is None else stream), "scale_flat")


@_traced
def symm_allreduce(domain, offset, numel, dtype, divisor, apply_mul, flag_offset, channel, seq, use_multicast,
                   ctas=0, stream=None):
    """Average (sum, x 1/divisor, inf/NaN check) a bucket of a symmetric allocation over the domain's
    GPUs: this rank's kernel, with the cross-rank barriers inside.  flag_offset < 0: no check."""
    global launches
    if numel > 0:
        launches += 1
    _check(lib.apex_b200_symm_allreduce(ctypes.byref(domain), int(offset), int(numel), DTYPE_CODE[dtype],
                                        float(divisor), int(bool(apply_mul)), int(flag_offset), int(channel),
                                        int(seq), int(bool(use_multicast)), int(ctas),
                                        current_stream() if stream is None else stream), "symm_allreduce")


@_traced
def mt_l2norm(table, grad_list, inv_scale, apply_scale, partials, norm_out, stream=None):
    global launches
    launches += 2 if table.struct.total_chunks > 0 else 1
    _check(lib.apex_b200_mt_l2norm(table.ref, grad_list, float(inv_scale), int(bool(apply_scale)), _ptr(partials),
                                   0 if partials is None else partials.numel(), _ptr(norm_out),
                                   current_stream() if stream is None else stream), "mt_l2norm")


NOALLOC_MAX_CHUNKS = 65536


@_traced
def mt_l2norm_noalloc(table, grad_list, inv_scale, apply_scale, norm_out, stream=None):
    """apex_b200_mt_l2norm for callers that cannot allocate scratch (memory pool exhausted, or a
    collective in flight that a cudaMalloc would stall): the partial sums use a fixed 256 KiB array
    inside the library.  Tables with more than NOALLOC_MAX_CHUNKS chunks raise; calls on one GPU
    must not overlap; make the first call on a device before any stream capture."""
    global launches
    launches += 2 if table.struct.total_chunks > 0 else 1
    _check(lib.apex_b200_mt_l2norm_noalloc(table.ref, grad_list, float(inv_scale), int(bool(apply_scale)),
                                           _ptr(norm_out), current_stream() if stream is None else stream),
           "mt_l2norm_noalloc")


@_traced
def mt_clip_by_norm(table, grad_list, norm, max_norm, flag, stream=None):
    _count(table)
    _check(lib.apex_b200_mt_clip_by_norm(table.ref, grad_list, _ptr(norm), float(max_norm), _ptr(flag),
                                         current_stream() if stream is None else stream), "mt_clip_by_norm")


@_traced
def mt_larc(table, grad_list, param_list, trust_coefficient, eps, weight_decay, lr, clip, scratch, norms=None,
            stream=None):
    global launches
    launches += 4 if table.struct.total_chunks > 0 else 1
    _check(lib.apex_b200_mt_larc(table.ref, grad_list, param_list, float(trust_coefficient), float(eps),
                                 float(weight_decay), float(lr), int(bool(clip)), _ptr(scratch), scratch.numel(),
                                 _ptr(norms), current_stream() if stream is None else stream), "mt_larc")
