"""Generate tests/golden/*.npz|json by running the UNMODIFIED reference on CPU.

Run here (the build container), where /root/reference exists:
    python tests/golden/make_golden.py
The reference (mbrukman/apex-1) is pure python on torch, so it imports under this
torch (2.11).  Two shims are applied from the outside, the reference files are
not touched:
  * FP16_Optimizer / DistributedDataParallel gate on the strings
    'torch.cuda.HalfTensor' / 'torch.cuda.FloatTensor'
    (fp16_optimizer.py:131,143; distributed.py:154-156): ``torch.Tensor.type`` is
    patched to report the cuda names for CPU tensors while they run.
  * DistributedDataParallel wants dist.DistBackend, torch.cuda.Stream and
    torch.cuda.IntTensor (distributed.py:129-134,144,185): aliased to
    dist.Backend, a dummy stream and torch.IntTensor.
The fixtures hold inputs AND outputs so the tests do not depend on RNG streams.
The files are small (a few hundred KB in total).
"""
import contextlib
import importlib.util
import io
import json
import os
import sys
from unittest import mock

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden_cases import build_layout_model  # noqa: E402
REF_ROOT = os.environ.get("APEX_REFERENCE", "/root/reference")


def load_reference():
    # under its own name, so that objects it pickles carry the real class paths
    spec = importlib.util.spec_from_file_location(
        "apex", os.path.join(REF_ROOT, "apex", "__init__.py"),
        submodule_search_locations=[os.path.join(REF_ROOT, "apex")])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["apex"] = mod
    spec.loader.exec_module(mod)
    return mod


_real_type = torch.Tensor.type


def _cuda_named_type(self, *a, **k):
    if a or k:
        return _real_type(self, *a, **k)
    name = _real_type(self)
    return name.replace("torch.", "torch.cuda.") if not name.startswith("torch.cuda") else name


@contextlib.contextmanager
def cuda_type_names():
    with mock.patch.object(torch.Tensor, "type", _cuda_named_type):
        yield


def f16(a):
    """numpy COPY of a tensor; fp16 is stored as its uint16 bit pattern."""
    a = a.detach().cpu().numpy().copy()
    return a.view(np.uint16) if a.dtype == np.float16 else a


# ---------------------------------------------------------------------------

def golden_scalers(apex, out):
    LS = apex.fp16_utils.loss_scaler
    # (1) small window trajectory with overflows at iterations 4 and 5
    s = LS.DynamicLossScaler(init_scale=16, scale_window=3)
    traj = []
    for it in range(10):
        s.update_scale(it in (4, 5))
        traj.append(s.cur_scale)
    out["dyn_traj_small"] = {"init_scale": 16, "scale_window": 3, "overflow_iters": [4, 5],
                             "cur_scale": traj, "cur_iter": s.cur_iter,
                             "last_overflow_iter": s.last_overflow_iter}
    # (2) defaults: growth points
    s = LS.DynamicLossScaler()
    growth = []
    prev = s.cur_scale
    for it in range(2100):
        s.update_scale(False)
        if s.cur_scale != prev:
            growth.append(it)
            prev = s.cur_scale
    out["dyn_default_growth_calls"] = growth
    out["dyn_default_final_scale_log2"] = float(np.log2(float(s.cur_scale)))
    # (3) floor
    s = LS.DynamicLossScaler(init_scale=2)
    for _ in range(4):
        s.update_scale(True)
    out["dyn_floor"] = {"cur_scale": s.cur_scale, "is_int": isinstance(s.cur_scale, int)}
    # a long mixed trajectory with a seeded overflow pattern
    rng = np.random.RandomState(7)
    pattern = (rng.rand(400) < 0.04).tolist()
    s = LS.DynamicLossScaler(init_scale=2**20, scale_factor=2., scale_window=25)
    rows = []
    for ov in pattern:
        s.update_scale(ov)
        rows.append([float(s.cur_scale), s.cur_iter, s.last_overflow_iter])
    out["dyn_traj_long"] = {"init_scale": 2**20, "scale_window": 25, "pattern": pattern, "rows": rows}
    # (4) _has_inf_or_nan known answers
    f = LS.DynamicLossScaler._has_inf_or_nan
    out["has_inf_or_nan"] = {
        "inf_and_neg_inf_f32": f(torch.tensor([float("inf"), -float("inf")])),
        "one_nan_f16": f(torch.tensor([1.0, float("nan")]).half()),
        "million_max_f16": f(torch.full((1000000,), 65504.0).half()),
        "two_3e38_f32": f(torch.tensor([3e38, 3e38])),
        "finite_f16": f(torch.randn(1000).half()),
        "neg_inf_f16": f(torch.tensor([0.5, -float("inf"), 2.0]).half()),
    }
    # static scaler
    st = LS.LossScaler(128.0)
    p = torch.nn.Parameter(torch.zeros(2))
    p.grad = torch.tensor([float("inf"), 1.0])
    out["static_has_overflow_on_inf"] = st.has_overflow([p])


def golden_amp(apex, out, arrays):
    S = apex.amp.scaler
    # partial mutation: three fp32 grads, the middle one overflows
    ps = [torch.nn.Parameter(torch.zeros(2)) for _ in range(3)]
    ps[0].grad = torch.tensor([4.0, 4.0])
    ps[1].grad = torch.tensor([float("inf"), 4.0])
    ps[2].grad = torch.tensor([4.0, 4.0])
    sc = S.LossScaler()
    skip = sc.unscale_and_update([{"params": ps}], 4.0)
    out["amp_partial"] = {"skip": skip, "grads": [p.grad.tolist() for p in ps],
                          "loss_scale_after": sc.loss_scale()}
    # trajectory: 2^16 -> 2^17 after exactly 2000 clean calls, cap at 2^24, halves w/o floor
    sc = S.LossScaler()
    p = torch.nn.Parameter(torch.zeros(4))
    scales = []
    for it in range(2001):
        p.grad = torch.ones(4)
        sc.unscale_and_update([{"params": [p]}], 1.0)
        if it in (1998, 1999, 2000):
            scales.append(sc.loss_scale())
    out["amp_growth"] = {"scale_after_calls_1999_2000_2001": scales}
    sc = S.LossScaler()
    sc._loss_scale = 2.**24
    sc._unskipped = 1999
    p.grad = torch.ones(4)
    sc.unscale_and_update([{"params": [p]}], 1.0)
    out["amp_cap"] = sc.loss_scale()
    sc = S.LossScaler()
    for _ in range(20):
        p.grad = torch.tensor([float("nan"), 1, 1, 1])
        sc.unscale_and_update([{"params": [p]}], 1.0)
    out["amp_no_floor"] = sc.loss_scale()
    # numeric unscale: non power of two scale, fp32
    g = torch.randn(1000, generator=torch.Generator().manual_seed(3)) * 1000
    arrays["amp_unscale_in"] = g.numpy().copy()
    overflow = S.scale_check_overflow(g, 1. / 3000.0)
    arrays["amp_unscale_out"] = g.numpy().copy()
    out["amp_unscale_scale"] = 1. / 3000.0
    out["amp_unscale_overflow"] = overflow


def make_mlp_case(seed):
    gen = torch.Generator().manual_seed(seed)
    shapes = [(64, 32), (64,), (16, 64), (16,), (5, 7), (1,)]   # last two: odd sizes, not multiples of 8
    params = [(torch.randn(s, generator=gen) * 0.05).half() for s in shapes]
    return shapes, params


def run_fp16_optimizer_case(apex, name, opt_name, opt_kwargs, fp16_kwargs, inject, iters, out, arrays,
                            with_fp32_param=False, clip=None, cap=2.0**20):
    """Drive the reference FP16_Optimizer with synthetic gradients written straight into .grad."""
    FP16_Optimizer = apex.fp16_utils.FP16_Optimizer
    shapes, init = make_mlp_case(11)
    params = [torch.nn.Parameter(p.clone()) for p in init]
    if with_fp32_param:
        params.append(torch.nn.Parameter(torch.randn(24, generator=torch.Generator().manual_seed(5))))
    arrays[name + "/n_params"] = np.array(len(params))
    for i, p in enumerate(params):
        arrays["%s/init/%d" % (name, i)] = f16(p.data)
    inner = getattr(torch.optim, opt_name)(params, **opt_kwargs)
    with cuda_type_names(), contextlib.redirect_stdout(io.StringIO()):
        opt = FP16_Optimizer(inner, verbose=False, **fp16_kwargs)
    gen = torch.Generator().manual_seed(1234)
    rows, messages, norms = [], [], []
    for it in range(iters):
        scale = opt.loss_scale
        for i, p in enumerate(params):
            g = torch.randn(p.shape, generator=gen) * 1e-3 * float(min(scale, cap))
            g = g.to(p.dtype)
            kind = inject.get(it)
            if kind is not None and i == kind[0]:
                flat = g.view(-1)
                for pos, val in kind[1]:
                    flat[pos % flat.numel()] = val
            p.grad = g
            arrays["%s/grad/%d/%d" % (name, it, i)] = f16(g)
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            opt.update_master_grads()
            if clip is not None:
                n = opt.clip_master_grads(clip)
                norms.append(float(n))
            opt.step()
        messages.append(buf.getvalue())
        ls = opt.loss_scaler
        rows.append({"overflow": bool(opt.overflow), "loss_scale": float(opt.loss_scale),
                     "cur_iter": getattr(ls, "cur_iter", None),
                     "last_overflow_iter": getattr(ls, "last_overflow_iter", None)})
        # per-iteration snapshot of the first fp16 param's master (cheap, catches step-level drift)
        arrays["%s/master0/%d" % (name, it)] = opt.fp32_from_fp16_groups[0][0].data.numpy().copy()
    for i, m in enumerate(opt.fp32_from_fp16_groups[0]):
        arrays["%s/final_master/%d" % (name, i)] = m.data.numpy().copy()
        arrays["%s/final_model/%d" % (name, i)] = f16(opt.fp16_groups[0][i].data)
        st = opt.optimizer.state[m]
        for k in ("momentum_buffer", "exp_avg", "exp_avg_sq"):
            if k in st and st[k] is not None:
                arrays["%s/final_%s/%d" % (name, k, i)] = st[k].numpy().copy()
        if "step" in st:
            arrays["%s/final_step/%d" % (name, i)] = np.array(float(st["step"]))
    for i, p in enumerate(opt.fp32_from_fp32_groups[0]):
        arrays["%s/final_fp32/%d" % (name, i)] = p.data.numpy().copy()
        arrays["%s/final_fp32_grad/%d" % (name, i)] = p.grad.data.numpy().copy()
    out[name] = {"optimizer": opt_name, "opt_kwargs": opt_kwargs, "fp16_kwargs": fp16_kwargs,
                 "inject": {str(k): [v[0], [[a, repr(b)] for a, b in v[1]]] for k, v in inject.items()},
                 "iters": iters, "rows": rows, "messages": messages, "with_fp32_param": with_fp32_param,
                 "clip": clip, "norms": norms}


def golden_fp16_optimizer(apex, out, arrays):
    inf, nan = float("inf"), float("nan")
    inject = {2: (0, [(5, inf)]), 3: (3, [(0, -inf)]), 7: (4, [(34, nan)]),
              9: (1, [(3, inf), (4, -inf)]), 10: (5, [(0, nan)])}
    dyn = {"dynamic_loss_scale": True, "dynamic_loss_args": {"init_scale": 2**16, "scale_window": 4}}
    run_fp16_optimizer_case(apex, "sgd_mom_wd_dynamic", "SGD",
                            {"lr": 0.1, "momentum": 0.9, "weight_decay": 1e-4}, dyn, inject, 16, out, arrays)
    run_fp16_optimizer_case(apex, "sgd_nesterov_static", "SGD",
                            {"lr": 0.05, "momentum": 0.8, "nesterov": True, "dampening": 0.0},
                            {"static_loss_scale": 128.0}, {}, 6, out, arrays)
    run_fp16_optimizer_case(apex, "sgd_plain_dampening", "SGD",
                            {"lr": 0.02, "momentum": 0.5, "dampening": 0.25, "weight_decay": 0.0},
                            {"static_loss_scale": 3.0}, {}, 5, out, arrays)
    run_fp16_optimizer_case(apex, "sgd_no_momentum", "SGD", {"lr": 0.3},
                            {"static_loss_scale": 1.0}, {}, 4, out, arrays)
    run_fp16_optimizer_case(apex, "adam_dynamic", "Adam",
                            {"lr": 1e-3, "betas": (0.9, 0.999), "eps": 1e-8, "weight_decay": 0.0},
                            dyn, inject, 16, out, arrays)
    run_fp16_optimizer_case(apex, "adam_wd_static", "Adam",
                            {"lr": 2e-3, "betas": (0.8, 0.99), "eps": 1e-6, "weight_decay": 0.01},
                            {"static_loss_scale": 512.0}, {}, 6, out, arrays)
    run_fp16_optimizer_case(apex, "adamw_static", "AdamW",
                            {"lr": 2e-3, "weight_decay": 0.05}, {"static_loss_scale": 64.0}, {}, 5, out, arrays)
    run_fp16_optimizer_case(apex, "sgd_mixed_fp32_dynamic", "SGD",
                            {"lr": 0.1, "momentum": 0.9}, dyn, {4: (6, [(2, inf)])}, 10, out, arrays,
                            with_fp32_param=True)
    run_fp16_optimizer_case(apex, "sgd_clip_dynamic", "SGD",
                            {"lr": 0.1, "momentum": 0.9}, dyn, {3: (0, [(1, nan)])}, 8, out, arrays, clip=0.05)
    # default dynamic scaler start: 2^32 walks down (probe saw 2^32 -> 2^25)
    run_fp16_optimizer_case(apex, "sgd_default_dynamic", "SGD", {"lr": 0.1, "momentum": 0.9},
                            {"dynamic_loss_scale": True}, {}, 14, out, arrays, cap=float("inf"))


def golden_fp16util(apex, out, arrays):
    U = apex.fp16_utils
    model = torch.nn.Sequential(torch.nn.Linear(8, 5), torch.nn.Linear(5, 3)).half()
    gen = torch.Generator().manual_seed(21)
    for p in model.parameters():
        p.data = (torch.randn(p.shape, generator=gen)).half()
    for flat in (False, True):
        model_params, master_params = U.prep_param_lists(model, flat_master=flat)
        for i, p in enumerate(model_params):
            p.grad = (torch.randn(p.shape, generator=gen) * 100).half()
            arrays["fp16util/%s/model/%d" % (flat, i)] = f16(p.data)
            arrays["fp16util/%s/grad/%d" % (flat, i)] = f16(p.grad)
        U.model_grads_to_master_grads(model_params, master_params, flat_master=flat)
        for i, m in enumerate(master_params):
            arrays["fp16util/%s/master_grad/%d" % (flat, i)] = m.grad.data.numpy().copy()
            m.data.add_(torch.randn(m.shape, generator=gen) * 1e-3 + 1e-4)
            arrays["fp16util/%s/master_new/%d" % (flat, i)] = m.data.numpy().copy()
        U.master_params_to_model_params(model_params, master_params, flat_master=flat)
        for i, p in enumerate(model_params):
            arrays["fp16util/%s/model_new/%d" % (flat, i)] = f16(p.data)
    out["fp16util"] = {"n_model": len(model_params)}


class _DummyStream(object):
    def wait_stream(self, other):
        pass


def golden_ddp_layout(apex, out):
    """Bucket structure the reference DDP learns on its first backward (world size 1, gloo)."""
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("gloo", rank=0, world_size=1)
    DDP = apex.parallel.DistributedDataParallel

    build = build_layout_model

    cases = []
    patches = [mock.patch.object(dist, "DistBackend", dist.Backend, create=True),
               mock.patch.object(torch.cuda, "Stream", _DummyStream),
               mock.patch.object(torch.cuda, "IntTensor", torch.IntTensor, create=True),
               mock.patch.object(torch.cuda, "current_stream", lambda *a, **k: _DummyStream()),
               mock.patch.object(torch.cuda, "stream", lambda s: contextlib.nullcontext())]
    with contextlib.ExitStack() as stack:
        for p in patches:
            stack.enter_context(p)
        stack.enter_context(cuda_type_names())
        for kind, message_size in [("mlp3", 1), ("mlp3", 50), ("mlp3", 100), ("mlp3", 10**7),
                                   ("mixed", 1), ("mixed", 30), ("mixed", 130)]:
            model = build(kind)
            ddp = DDP(model, message_size=message_size)
            x = torch.randn(5, 6)
            for it in range(3):       # iteration 0 learns, 1 and 2 run overlapped
                model.zero_grad()
                y = ddp(x)
                y = y.float() if y.dtype != torch.float32 else y
                y.sum().backward()
            plist = [p for p in model.parameters() if p.requires_grad]
            where = [list(ddp.param_id_to_bucket[id(p)]) for p in plist]
            cases.append({"model": kind, "message_size": message_size, "num_buckets": ddp.num_buckets,
                          "bucket_sizes": list(ddp.bucket_sizes), "param_to_bucket": where,
                          "numels": [p.numel() for p in plist],
                          "dtypes": [str(p.dtype) for p in plist]})
    out["ddp_layout"] = cases
    dist.destroy_process_group()


def golden_checkpoint(apex):
    """A checkpoint written by the reference: FP16_Optimizer.state_dict() through torch.save
    (examples/FP16_Optimizer_simple/save_load.py:27-37).  It pickles the DynamicLossScaler object."""
    shapes, init = make_mlp_case(11)
    params = [torch.nn.Parameter(p.clone()) for p in init[:4]]
    inner = torch.optim.SGD(params, lr=0.1, momentum=0.9)
    with cuda_type_names(), contextlib.redirect_stdout(io.StringIO()):
        opt = apex.fp16_utils.FP16_Optimizer(inner, dynamic_loss_scale=True,
                                              dynamic_loss_args={"init_scale": 2**10, "scale_window": 3}, verbose=False)
    gen = torch.Generator().manual_seed(77)
    grads = []
    for it in range(7):
        gs = [(torch.randn(p.shape, generator=gen) * 0.5).half() for p in params]
        if it == 2:
            gs[1][0] = float("inf")
        grads.append(gs)
    with contextlib.redirect_stdout(io.StringIO()):
        for it in range(5):
            for p, g in zip(params, grads[it]):
                p.grad = g.clone()
            opt.update_master_grads()
            opt.step()
        ckpt = {"optimizer": opt.state_dict(), "model": [p.data.clone() for p in params]}
        torch.save(ckpt, os.path.join(HERE, "reference_checkpoint.pt"))
        for it in range(5, 7):                       # what the reference does after the checkpoint
            for p, g in zip(params, grads[it]):
                p.grad = g.clone()
            opt.update_master_grads()
            opt.step()
    torch.save({"grads": grads[5:], "final_masters": [m.data.clone() for m in opt.fp32_from_fp16_groups[0]],
                "final_model": [p.data.clone() for p in params], "final_scale": float(opt.loss_scale),
                "final_iter": opt.loss_scaler.cur_iter}, os.path.join(HERE, "reference_checkpoint_after.pt"))


def golden_larc(apex, out, arrays):
    """apex.parallel.LARC (apex/parallel/LARC.py) around SGD, clip and scale modes, a zero-norm
    parameter and a parameter without gradient."""
    from importlib import import_module
    LARC = import_module("apex.parallel.LARC").LARC
    for clip in (True, False):
        name = "larc_clip" if clip else "larc_scale"
        gen = torch.Generator().manual_seed(9)
        shapes = [(40, 7), (13,), (5, 5, 3), (1,)]
        init = [torch.randn(s, generator=gen) * 0.3 for s in shapes]
        init[3].zero_()
        ps = [torch.nn.Parameter(t.clone()) for t in init]
        opt = LARC(torch.optim.SGD(ps, lr=0.1, momentum=0.9, weight_decay=5e-4), trust_coefficient=0.02, clip=clip)
        for i, t in enumerate(init):
            arrays["%s/init/%d" % (name, i)] = f16(t)
        for it in range(4):
            for i, p in enumerate(ps):
                g = torch.randn(p.shape, generator=gen) * (10.0 if it % 2 else 0.01)
                arrays["%s/grad/%d/%d" % (name, it, i)] = f16(g)
                p.grad = g.clone()
            if it == 2:
                ps[1].grad = None
            opt.step()
        for i, p in enumerate(ps):
            arrays["%s/final/%d" % (name, i)] = f16(p.data)
    out["larc"] = {"lr": 0.1, "momentum": 0.9, "weight_decay": 5e-4, "trust_coefficient": 0.02, "iters": 4,
                   "no_grad": {"iteration": 2, "param": 1}}


def golden_amp_handle(apex, out):
    """AmpHandle.scale_loss / OptimWrapper (apex/amp/handle.py:32-63, opt.py:19-82) driven directly;
    amp.init()'s function patching is not involved (and cannot run on this torch)."""
    from importlib import import_module
    AmpHandle = import_module("apex.amp.handle").AmpHandle
    log = []
    w = torch.nn.Parameter(torch.tensor([1.0, 2.0, -1.0]))
    opt = torch.optim.SGD([w], lr=0.5)
    handle = AmpHandle()
    coeffs = [[1.0, 3.0, 0.5], [float("inf"), 1.0, 1.0], [2.0, -1.0, 0.25], [float("nan"), 0.0, 0.0], [1.0, 1.0, 1.0]]
    for c in coeffs:
        opt.zero_grad()
        loss = (w * torch.tensor(c)).sum()
        with handle.scale_loss(loss, opt) as scaled:
            scaled.backward()
        opt.step()
        log.append({"w": w.data.tolist(), "scale": handle._default_scaler.loss_scale()})
    out["amp_handle"] = {"coeffs": [[repr(x) for x in c] for c in coeffs], "log": log}
    # two losses per iteration through the optimizer wrapper; the second loss overflows in iteration 1
    w2 = torch.nn.Parameter(torch.tensor([1.0, 1.0]))
    wrapped = AmpHandle().wrap_optimizer(torch.optim.SGD([w2], lr=0.25), num_loss=2)
    log2 = []
    for it in range(3):
        wrapped.zero_grad()
        with wrapped.scale_loss((w2 * 2).sum()) as s0:
            s0.backward()
        second = torch.tensor([float("inf"), 1.0]) if it == 1 else torch.tensor([3.0, -1.0])
        with wrapped.scale_loss((w2 * second).sum()) as s1:
            s1.backward()
        wrapped.step()
        log2.append({"w": w2.data.tolist(), "grad": [repr(x) for x in w2.grad.tolist()],
                     "scales": [sc.loss_scale() for sc in wrapped._loss_scaler]})
    out["amp_optim_wrapper"] = {"log": log2}


def main():
    apex = load_reference()
    golden_checkpoint(apex)
    out, arrays = {}, {}
    out["torch_version"] = torch.__version__
    golden_scalers(apex, out)
    golden_amp(apex, out, arrays)
    golden_fp16_optimizer(apex, out, arrays)
    golden_fp16util(apex, out, arrays)
    golden_ddp_layout(apex, out)
    golden_larc(apex, out, arrays)
    golden_amp_handle(apex, out)
    with open(os.path.join(HERE, "reference_golden.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), **arrays)
    print("wrote", len(out), "json entries and", len(arrays), "arrays")


if __name__ == "__main__":
    main()
