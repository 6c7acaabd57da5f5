"""Host-side logic of the product package on CPU tensors.

The kernels cannot run here; the `cpu_backend` fixture swaps the python kernel
wrappers for an oracle-backed double (tests/fake_backend.py) so that everything
ABOVE the C ABI -- tensor tables, lazy master gradients, the commit protocol for
optimizer state, scaler trajectories, messages, checkpoints -- runs for real and is
compared with what the unmodified reference produced (tests/golden).
"""
import contextlib
import io

import numpy as np
import pytest
import torch

from tests.golden_cases import CASES, bits, grads_for, init_params, make_inner, tensor_from, ulp_diff


def _drive(name, meta, arrays, backend, probe=None):
    from apex_b200.fp16_utils import FP16_Optimizer
    params = init_params(arrays, name)
    inner = make_inner(meta)(params)
    opt = FP16_Optimizer(inner, verbose=False, **meta["fp16_kwargs"])
    rows, msgs, norms, m0 = [], [], [], []
    for it in range(meta["iters"]):
        for p, g in zip(params, grads_for(arrays, name, it, len(params))):
            p.grad = g
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            opt.update_master_grads()
            if meta["clip"] is not None:
                norms.append(float(opt.clip_master_grads(meta["clip"])))
            if probe is not None:
                probe(opt, it)
            opt.step()
        msgs.append(buf.getvalue())
        ls = opt.loss_scaler
        rows.append({"overflow": bool(opt.overflow), "loss_scale": float(opt.loss_scale),
                     "cur_iter": getattr(ls, "cur_iter", None),
                     "last_overflow_iter": getattr(ls, "last_overflow_iter", None)})
        m0.append(opt.fp32_from_fp16_groups[0][0].data.numpy().copy())
    return opt, rows, msgs, norms, m0


@pytest.mark.parametrize("name", CASES)
def test_fp16_optimizer_replays_reference(name, golden_json, golden_arrays, cpu_backend):
    meta = golden_json[name]
    adam = meta["optimizer"] in ("Adam", "AdamW")
    opt, rows, msgs, norms, m0 = _drive(name, meta, golden_arrays, cpu_backend)
    # decisions, trajectory and messages: bit-exact
    assert rows == meta["rows"]
    assert msgs == meta["messages"]
    if meta["clip"] is not None:
        # the norm is reduced in another order than torch's: same value to fp32 rounding
        assert np.allclose(norms, meta["norms"], rtol=2e-6, atol=0)
    steps_taken = sum(1 for r in rows if not r["overflow"])
    for i, (m, p) in enumerate(zip(opt.fp32_from_fp16_groups[0], opt.fp16_groups[0])):
        want = golden_arrays["%s/final_master/%d" % (name, i)]
        if meta["clip"] is not None:
            assert np.allclose(m.data.numpy(), want, rtol=1e-5, atol=1e-8), (name, i)
        elif not adam:
            assert np.array_equal(m.data.numpy(), want), (name, i)
            assert np.array_equal(bits(p.data), golden_arrays["%s/final_model/%d" % (name, i)])
        else:
            # the double computes Adam with torch's CUDA association, the golden run used torch's
            # CPU kernels (other association, sqrt not correctly rounded): a few ulp at the
            # parameter's scale per step.  Exact Adam parity is asserted on the GPU.
            scale = np.maximum(np.abs(want), meta["opt_kwargs"]["lr"])
            assert (np.abs(m.data.numpy() - want) <= 4 * steps_taken * np.spacing(scale.astype(np.float32))).all()
        # the fp16 copy is always the rounding of the master next to it
        assert torch.equal(p.data, m.data.half())
        st = opt.optimizer.state[m]
        gk = "%s/final_momentum_buffer/%d" % (name, i)
        if gk in golden_arrays:
            if meta["clip"] is not None:
                assert np.allclose(st["momentum_buffer"].numpy(), golden_arrays[gk], rtol=1e-5, atol=1e-9)
            else:
                assert np.array_equal(st["momentum_buffer"].numpy(), golden_arrays[gk])
        gk = "%s/final_step/%d" % (name, i)
        if gk in golden_arrays:
            assert float(st["step"]) == float(golden_arrays[gk]) == steps_taken
    for i, p in enumerate(opt.fp32_from_fp32_groups[0]):
        assert np.array_equal(p.data.numpy(), golden_arrays["%s/final_fp32/%d" % (name, i)])
        assert np.array_equal(p.grad.data.numpy(), golden_arrays["%s/final_fp32_grad/%d" % (name, i)])
    # which path ran
    kinds = set(cpu_backend.calls)
    assert ("fused_adam" if adam else "fused_sgd") in kinds
    assert "grads_to_master" not in kinds            # master grads never had to be materialised,
    if meta["clip"] is not None:                     # not even for clipping: norm from the fp16 grads
        assert "l2norm" in kinds


def test_skipped_step_leaves_no_optimizer_state(golden_json, golden_arrays, cpu_backend):
    """First iterations overflow at scale 2^32: the reference's inner optimizer has seen no
    step, so its state must still be empty (momentum buffers are only published on commit)."""
    name = "sgd_default_dynamic"
    meta = golden_json[name]
    seen = []

    def probe(opt, it):
        seen.append(len(opt.optimizer.state))
    opt, rows, _, _, _ = _drive(name, meta, golden_arrays, cpu_backend, probe)
    first_clean = [r["overflow"] for r in rows].index(False)
    assert first_clean == 8
    assert seen[:first_clean + 1] == [0] * (first_clean + 1)
    assert len(opt.optimizer.state) == 6


def test_master_grads_are_lazy_but_observable(cpu_backend):
    from apex_b200.fp16_utils import FP16_Optimizer
    p = torch.nn.Parameter((torch.arange(20, dtype=torch.float32) / 7).half())
    opt = FP16_Optimizer(torch.optim.SGD([p], lr=0.5), static_loss_scale=3.0, verbose=False)
    master = opt.fp32_from_fp16_groups[0][0]
    assert master.dtype == torch.float32 and master.requires_grad and master.is_leaf
    assert master.grad is None
    p.grad = (torch.arange(20, dtype=torch.float32) - 4).half()
    opt.update_master_grads()
    assert "grads_to_master" not in cpu_backend.calls
    # looking at the gradient makes it real: float(g16) * float(1/3.)
    want = p.grad.float() * torch.tensor(1. / 3.0, dtype=torch.float32)
    assert torch.equal(master.grad, want)
    assert cpu_backend.calls.count("grads_to_master") == 1
    master.grad            # second look does not launch again
    assert cpu_backend.calls.count("grads_to_master") == 1
    # a materialised gradient may be edited by the user; step() must use the edited values
    master.grad.mul_(2.0)
    before = master.data.clone()
    opt.step()
    assert torch.equal(master.data, before - 0.5 * (want * 2.0))
    assert torch.equal(p.data, master.data.half())
    # zero_grad: both gradients zeroed in ONE launch
    n_zero = cpu_backend.calls.count("zero")
    opt.zero_grad()
    assert cpu_backend.calls.count("zero") == n_zero + 1
    assert float(p.grad.abs().sum()) == 0 and float(master.grad.abs().sum()) == 0
    opt.zero_grad(set_grads_to_None=True)
    assert p.grad is None and master.grad is None


def test_generic_inner_optimizer_takes_the_unfused_path(cpu_backend):
    """An inner optimizer we have no fused kernel for (Adagrad) still works: gradients are
    materialised, torch does the update, one launch refreshes the fp16 copies."""
    from apex_b200.fp16_utils import FP16_Optimizer
    from oracle import apex_ref
    torch.manual_seed(3)
    init = [(torch.randn(33) * 0.1).half(), (torch.randn(4, 5) * 0.1).half()]
    grads = [[(torch.randn(t.shape) * 8).half() for t in init] for _ in range(3)]
    ps = [torch.nn.Parameter(t.clone()) for t in init]
    opt = FP16_Optimizer(torch.optim.Adagrad(ps, lr=0.1), static_loss_scale=8.0, verbose=False)
    rs = [torch.nn.Parameter(t.clone()) for t in init]
    ref = apex_ref.MixedPrecisionStep(rs, lambda q: torch.optim.Adagrad(q, lr=0.1), static_loss_scale=8.0)
    for it in range(3):
        for p, r, g in zip(ps, rs, grads[it]):
            p.grad = g.clone()
            r.grad = g.clone()
        opt.update_master_grads()
        opt.step()
        ref.update_master_grads()
        ref.step()
    assert "fused_sgd" not in cpu_backend.calls and "fused_adam" not in cpu_backend.calls
    assert "master_to_model" in cpu_backend.calls
    for p, r, m, rm in zip(ps, rs, opt.fp32_from_fp16_groups[0], ref.masters):
        assert torch.equal(m.data, rm.data)
        assert torch.equal(p.data, r.data)


def test_closure_step(cpu_backend):
    from apex_b200.fp16_utils import FP16_Optimizer
    p = torch.nn.Parameter(torch.ones(8).half())
    opt = FP16_Optimizer(torch.optim.SGD([p], lr=0.25), static_loss_scale=4.0, verbose=False)
    calls = []

    def closure():
        calls.append(1)
        p.grad = torch.full((8,), 4.0).half()       # "scaled" gradient of 1.0
        opt.update_master_grads()
        return torch.tensor(1.0)
    out = opt.step(closure)
    assert len(calls) == 1 and float(out) == 1.0
    assert torch.equal(p.data, torch.full((8,), 0.75).half())
    assert opt.first_closure_call_this_step is True


def test_constructor_contract(cpu_backend, capsys):
    from apex_b200.fp16_utils import FP16_Optimizer, DynamicLossScaler, LossScaler
    h = torch.nn.Parameter(torch.zeros(3).half())
    f = torch.nn.Parameter(torch.zeros(2))
    frozen = torch.nn.Parameter(torch.zeros(2).half(), requires_grad=False)
    inner = torch.optim.SGD([h, f, frozen], lr=0.1, momentum=0.9)
    opt = FP16_Optimizer(inner, dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 8, "scale_window": 2})
    out = capsys.readouterr().out
    assert "FP16_Optimizer processing param group 0:" in out
    assert "FP16_Optimizer received torch.cuda.HalfTensor with torch.Size([3])" in out
    assert "FP16_Optimizer received torch.cuda.FloatTensor with torch.Size([2])" in out
    assert opt.fp16_groups == [[h]] or (len(opt.fp16_groups[0]) == 1 and opt.fp16_groups[0][0] is h)
    assert opt.fp32_from_fp32_groups[0][0] is f
    assert inner.param_groups[0]["params"][0] is opt.fp32_from_fp16_groups[0][0]
    assert inner.param_groups[0]["params"][1] is f
    assert isinstance(opt.loss_scaler, DynamicLossScaler) and opt.loss_scale == 8
    assert opt.overflow is False and opt.first_closure_call_this_step is True
    opt.loss_scale = 4.0
    assert opt.loss_scaler.cur_scale == 4.0
    assert opt.param_groups is inner.param_groups and opt.state is inner.state
    with pytest.raises(NotImplementedError):
        opt.inspect_master_grad_data()
    with pytest.raises(RuntimeError):
        opt.__getstate__()
    static = FP16_Optimizer(torch.optim.SGD([torch.nn.Parameter(torch.zeros(1).half())], lr=1.0),
                            static_loss_scale=128.0, verbose=False)
    assert isinstance(static.loss_scaler, LossScaler) and static.loss_scale == 128.0
    with pytest.raises(TypeError, match="Wrapped parameters must be either"):
        FP16_Optimizer(torch.optim.SGD([torch.nn.Parameter(torch.zeros(1).double())], lr=1.0), verbose=False)


def test_state_dict_round_trip_and_pickle(cpu_backend, tmp_path):
    from apex_b200.fp16_utils import FP16_Optimizer

    def build():
        torch.manual_seed(0)
        ps = [torch.nn.Parameter((torch.randn(9) * 0.1).half()), torch.nn.Parameter((torch.randn(2, 3) * 0.1).half())]
        return ps, FP16_Optimizer(torch.optim.SGD(ps, lr=0.1, momentum=0.9), dynamic_loss_scale=True,
                                  dynamic_loss_args={"init_scale": 16., "scale_window": 3}, verbose=False)
    ps, opt = build()
    g = torch.Generator().manual_seed(5)
    for it in range(4):
        for p in ps:
            p.grad = (torch.randn(p.shape, generator=g) * 16).half()
        opt.update_master_grads()
        opt.step()
    sd = opt.state_dict()
    assert sorted(sd) == ['dynamic_loss_scale', 'first_closure_call_this_step', 'fp32_from_fp16', 'loss_scaler',
                          'optimizer_state_dict', 'overflow']
    path = tmp_path / "ckpt.pt"
    torch.save({"optimizer": sd, "model": [p.data for p in ps]}, str(path))
    loaded = torch.load(str(path), weights_only=False)
    # masters are stored as plain fp32 tensors, like the reference's
    assert type(loaded["optimizer"]["fp32_from_fp16"][0][0]) is torch.Tensor
    assert set(loaded["optimizer"]["loss_scaler"].__dict__) == {"cur_scale", "cur_iter", "last_overflow_iter",
                                                                 "scale_factor", "scale_window"}
    ps2, opt2 = build()
    for p2, saved in zip(ps2, loaded["model"]):
        p2.data.copy_(saved)
    opt2.load_state_dict(loaded["optimizer"])
    assert opt2.loss_scale == opt.loss_scale and opt2.loss_scaler.cur_iter == opt.loss_scaler.cur_iter
    for a, b in zip(opt.fp32_from_fp16_groups[0], opt2.fp32_from_fp16_groups[0]):
        assert torch.equal(a.data, b.data)
    # both continue identically
    for it in range(3):
        gs = [(torch.randn(p.shape, generator=g) * 16).half() for p in ps]
        for o, plist in ((opt, ps), (opt2, ps2)):
            for p, gg in zip(plist, gs):
                p.grad = gg.clone()
            o.update_master_grads()
            o.step()
    for a, b in zip(opt.fp32_from_fp16_groups[0], opt2.fp32_from_fp16_groups[0]):
        assert torch.equal(a.data, b.data)
    for a, b in zip(ps, ps2):
        assert torch.equal(a.data, b.data)


def test_product_scalers_follow_reference_trajectories(golden_json):
    from apex_b200.fp16_utils import DynamicLossScaler, LossScaler
    g = golden_json["dyn_traj_small"]
    s = DynamicLossScaler(init_scale=g["init_scale"], scale_window=g["scale_window"])
    traj = []
    for it in range(10):
        s.update_scale(it in g["overflow_iters"])
        traj.append(s.cur_scale)
    assert traj == g["cur_scale"]
    assert [type(x) for x in traj[:2]] == [int, int]          # python int until the first change
    g = golden_json["dyn_traj_long"]
    s = DynamicLossScaler(init_scale=g["init_scale"], scale_factor=2., scale_window=g["scale_window"])
    for ov, row in zip(g["pattern"], g["rows"]):
        s.update_scale(ov)
        assert [float(s.cur_scale), s.cur_iter, s.last_overflow_iter] == row
    s = DynamicLossScaler()
    assert (s.cur_scale, s.scale_factor, s.scale_window, s.cur_iter, s.last_overflow_iter) == (2**32, 2., 1000, 0, -1)
    growth, prev = [], s.cur_scale
    for it in range(2100):
        s.update_scale(False)
        if s.cur_scale != prev:
            growth.append(it)
            prev = s.cur_scale
    assert growth == golden_json["dyn_default_growth_calls"]
    s = DynamicLossScaler(init_scale=2)
    for _ in range(4):
        s.update_scale(True)
    assert s.cur_scale == 1 and isinstance(s.cur_scale, int)
    st = LossScaler(128.0)
    p = torch.nn.Parameter(torch.zeros(2))
    p.grad = torch.tensor([float("inf"), 1.0])
    assert st.has_overflow([p]) is False and st.loss_scale == 128.0
    st.update_scale(True)
    assert st.loss_scale == 128.0


def test_product_has_overflow_known_answers(golden_json, cpu_backend):
    from apex_b200.fp16_utils import DynamicLossScaler
    g = golden_json["has_inf_or_nan"]
    f = DynamicLossScaler._has_inf_or_nan
    assert f(torch.tensor([float("inf"), -float("inf")])) == g["inf_and_neg_inf_f32"]
    assert f(torch.tensor([1.0, float("nan")]).half()) == g["one_nan_f16"]
    assert f(torch.full((1000000,), 65504.0).half()) == g["million_max_f16"]
    assert f(torch.tensor([0.5, -float("inf"), 2.0]).half()) == g["neg_inf_f16"]
    assert f(torch.randn(1000).half()) is False
    s = DynamicLossScaler()
    ps = [torch.nn.Parameter(torch.zeros(4).half()) for _ in range(3)]
    assert s.has_overflow(ps) is False            # no grads at all
    for p in ps:
        p.grad = torch.ones(4).half()
    assert s.has_overflow(ps) is False
    ps[2].grad[1] = float("nan")
    assert s.has_overflow(ps) is True


def test_product_amp_scaler(golden_json, golden_arrays, cpu_backend):
    from apex_b200.amp import LossScaler, scale_check_overflow
    g = golden_json["amp_partial"]
    ps = [torch.nn.Parameter(torch.zeros(2)) for _ in range(3)]
    ps[0].grad = torch.tensor([4.0, 4.0])
    ps[1].grad = torch.tensor([float("inf"), 4.0])
    ps[2].grad = torch.tensor([4.0, 4.0])
    sc = LossScaler()
    assert sc.unscale_and_update([{"params": ps}], 4.0) == g["skip"]
    assert [p.grad.tolist() for p in ps] == g["grads"]           # the reference's partial mutation
    assert sc.loss_scale() == g["loss_scale_after"]
    sc = LossScaler()
    p = torch.nn.Parameter(torch.zeros(4))
    seen = []
    for it in range(2001):
        p.grad = torch.ones(4)
        sc.unscale_and_update([{"params": [p]}], 1.0)
        if it in (1998, 1999, 2000):
            seen.append(sc.loss_scale())
    assert seen == golden_json["amp_growth"]["scale_after_calls_1999_2000_2001"]
    sc = LossScaler()
    sc._loss_scale, sc._unskipped = 2.**24, 1999
    p.grad = torch.ones(4)
    sc.unscale_and_update([{"params": [p]}], 1.0)
    assert sc.loss_scale() == golden_json["amp_cap"]
    sc = LossScaler()
    for _ in range(20):
        p.grad = torch.tensor([float("nan"), 1, 1, 1])
        assert sc.unscale_and_update([{"params": [p]}], 1.0) is True
    assert sc.loss_scale() == golden_json["amp_no_floor"]
    x = tensor_from(golden_arrays["amp_unscale_in"])
    assert scale_check_overflow(x, golden_json["amp_unscale_scale"]) is False
    assert np.array_equal(x.numpy(), golden_arrays["amp_unscale_out"])
    y = torch.tensor([1.0, float("nan")])
    assert scale_check_overflow(y, 0.5) is True and y[0] == 1.0


def test_amp_handle_and_optim_wrapper(cpu_backend):
    from apex_b200.amp import AmpHandle, NoOpHandle
    w = torch.nn.Parameter(torch.tensor([1.0, 2.0]))
    opt = torch.optim.SGD([w], lr=0.5)
    handle = AmpHandle()
    loss = (w * torch.tensor([1.0, 3.0])).sum()
    with handle.scale_loss(loss, opt) as scaled:
        assert float(scaled) == float(loss) * 2.**16
        scaled.backward()
    assert torch.equal(w.grad, torch.tensor([1.0, 3.0]))        # unscaled in place
    opt.step()
    assert torch.equal(w.data, torch.tensor([0.5, 0.5]))
    # overflow: the next step() is a one-shot no-op and the scale halves
    opt.zero_grad()
    loss = (w * torch.tensor([float("inf"), 1.0])).sum()
    with handle.scale_loss(loss, opt) as scaled:
        scaled.backward()
    before = w.data.clone()
    opt.step()
    assert torch.equal(w.data, before)
    assert handle._default_scaler.loss_scale() == 2.**15
    w.grad = torch.tensor([1.0, 1.0])
    opt.step()                                                   # the real step is back
    assert torch.equal(w.data, before - 0.5)
    # wrapped optimizer with two losses: gradients of loss 0 are set aside and added back
    w2 = torch.nn.Parameter(torch.tensor([1.0, 1.0]))
    wrapped = AmpHandle().wrap_optimizer(torch.optim.SGD([w2], lr=1.0), num_loss=2)
    with wrapped.scale_loss((w2 * 2).sum()) as s0:
        s0.backward()
    with wrapped.scale_loss((w2 * 3).sum()) as s1:
        s1.backward()
    assert torch.equal(w2.grad, torch.tensor([5.0, 5.0]))
    wrapped.step()
    assert torch.equal(w2.data, torch.tensor([-4.0, -4.0]))
    with pytest.raises(NotImplementedError):
        wrapped.step(closure=lambda: None)
    noop = NoOpHandle()
    with noop.scale_loss(loss, opt) as same:
        assert same is loss


def test_fp16util_functions(golden_json, golden_arrays, cpu_backend):
    from apex_b200 import fp16_utils as U
    n = golden_json["fp16util"]["n_model"]
    for flat in (False, True):
        model = torch.nn.Sequential(torch.nn.Linear(8, 5), torch.nn.Linear(5, 3)).half()
        for i, p in enumerate(model.parameters()):
            p.data = tensor_from(golden_arrays["fp16util/%s/model/%d" % (flat, i)])
        model_params, master_params = U.prep_param_lists(model, flat_master=flat)
        assert len(master_params) == (1 if flat else n)
        assert all(m.dtype == torch.float32 and m.requires_grad for m in master_params)
        for i, p in enumerate(model_params):
            p.grad = tensor_from(golden_arrays["fp16util/%s/grad/%d" % (flat, i)])
        U.model_grads_to_master_grads(model_params, master_params, flat_master=flat)
        for i, m in enumerate(master_params):
            assert np.array_equal(m.grad.data.numpy(), golden_arrays["fp16util/%s/master_grad/%d" % (flat, i)])
            m.data.copy_(tensor_from(golden_arrays["fp16util/%s/master_new/%d" % (flat, i)]))
        U.master_params_to_model_params(model_params, master_params, flat_master=flat)
        for i, p in enumerate(model_params):
            assert np.array_equal(bits(p.data), golden_arrays["fp16util/%s/model_new/%d" % (flat, i)])
    # a parameter without gradient clears the master's gradient (fp16util.py:111-112)
    model_params, master_params = U.prep_param_lists(model)
    master_params[0].grad = torch.zeros_like(master_params[0])
    model_params[0].grad = None
    U.model_grads_to_master_grads(model_params, master_params)
    assert master_params[0].grad is None


def test_network_to_half_keeps_batchnorm_fp32():
    from apex_b200.fp16_utils import network_to_half, tofp16, BN_convert_float
    net = torch.nn.Sequential(torch.nn.Linear(4, 4), torch.nn.BatchNorm1d(4))
    half = network_to_half(net)
    assert isinstance(half[0], tofp16)
    assert half[1][0].weight.dtype == torch.float16 and half[1][1].weight.dtype == torch.float32


def test_checkpoint_written_by_the_reference_loads(cpu_backend):
    """tests/golden/reference_checkpoint.pt was written by the unmodified reference (torch.save of
    FP16_Optimizer.state_dict(), which pickles apex.fp16_utils.loss_scaler.DynamicLossScaler)."""
    import os
    import sys
    import apex_b200
    from apex_b200.fp16_utils import FP16_Optimizer, DynamicLossScaler
    from tests.conftest import GOLDEN_DIR
    from tests.golden_cases import tensor_from
    added = [k for k in ("apex",) if k not in sys.modules]
    apex_b200.install_as_apex()
    try:
        import apex
        assert apex.fp16_utils.FP16_Optimizer is FP16_Optimizer
        ckpt = torch.load(os.path.join(GOLDEN_DIR, "reference_checkpoint.pt"), weights_only=False)
        after = torch.load(os.path.join(GOLDEN_DIR, "reference_checkpoint_after.pt"), weights_only=False)
    finally:
        for name in [k for k in sys.modules if k == "apex" or k.startswith("apex.")]:
            del sys.modules[name]
    assert isinstance(ckpt["optimizer"]["loss_scaler"], DynamicLossScaler)
    params = [torch.nn.Parameter(p.clone()) for p in ckpt["model"]]
    opt = FP16_Optimizer(torch.optim.SGD(params, lr=0.1, momentum=0.9), dynamic_loss_scale=True, verbose=False)
    opt.load_state_dict(ckpt["optimizer"])
    assert opt.loss_scaler.scale_window == 3
    for gs in after["grads"]:
        for p, g in zip(params, gs):
            p.grad = g.clone()
        opt.update_master_grads()
        opt.step()
    assert float(opt.loss_scale) == after["final_scale"] and opt.loss_scaler.cur_iter == after["final_iter"]
    for m, want in zip(opt.fp32_from_fp16_groups[0], after["final_masters"]):
        assert torch.equal(m.data, want)
    for p, want in zip(params, after["final_model"]):
        assert torch.equal(p.data, want)


def test_clip_is_virtual_until_somebody_looks(cpu_backend):
    from apex_b200.fp16_utils import FP16_Optimizer
    from oracle import apex_ref
    torch.manual_seed(5)
    init = [(torch.randn(300) * 0.1).half(), (torch.randn(7, 9) * 0.1).half(), torch.randn(12) * 0.1]
    ps = [torch.nn.Parameter(t.clone()) for t in init]
    rs = [torch.nn.Parameter(t.clone()) for t in init]
    make = lambda q: torch.optim.SGD(q, lr=0.5, momentum=0.9)      # noqa: E731
    opt = FP16_Optimizer(make(ps), static_loss_scale=64.0, verbose=False)
    ref = apex_ref.MixedPrecisionStep(rs, make, static_loss_scale=64.0)
    for it in range(3):
        gs = [(torch.randn(t.shape) * 64).to(t.dtype) for t in init]
        for p, r, g in zip(ps, rs, gs):
            p.grad, r.grad = g.clone(), g.clone()
        opt.update_master_grads()
        ref.update_master_grads()
        n1, n2 = opt.clip_master_grads(0.7), ref.clip_master_grads(0.7)
        assert float(n1) == pytest.approx(float(n2), rel=2e-6)
        if it == 1:
            # looking at a master gradient now shows the CLIPPED value
            want = ref.masters[0].grad
            assert torch.allclose(opt.fp32_from_fp16_groups[0][0].grad, want, rtol=1e-5, atol=0)
        # the user's own fp32 gradient was clipped in place, like torch does
        assert torch.allclose(ps[2].grad, rs[2].grad, rtol=1e-5, atol=0)
        opt.step()
        ref.step()
    assert cpu_backend.calls.count("grads_to_master") == 1           # only the look in iteration 1
    assert cpu_backend.calls.count("l2norm") == 3
    for m, rm in zip(opt.fp32_from_fp16_groups[0], ref.masters):
        assert torch.allclose(m.data, rm.data, rtol=1e-5, atol=1e-7)
    # clipping twice in one iteration: the second call takes the generic route on real gradients
    for p in ps:
        p.grad = torch.ones_like(p)
    opt.update_master_grads()
    a = float(opt.clip_master_grads(0.1))
    b = float(opt.clip_master_grads(0.1))
    assert a == pytest.approx((300 + 63 + 12) ** 0.5 / 64, rel=1e-5) and b == pytest.approx(0.1, rel=1e-4)
    # other norm types go to torch as in the reference
    opt.update_master_grads()
    assert float(opt.clip_master_grads(1e9, norm_type=float("inf"))) == pytest.approx(1 / 64.)


@pytest.mark.parametrize("clip", [True, False])
def test_larc_matches_reference_loop(clip, cpu_backend):
    from apex_b200.parallel import LARC
    from oracle import apex_ref
    torch.manual_seed(9)
    shapes = [(40, 7), (13,), (5, 5, 3), (1,)]
    init = [torch.randn(s) * 0.3 for s in shapes]
    init[3].zero_()                                     # zero parameter norm: the reference skips it
    ps = [torch.nn.Parameter(t.clone()) for t in init]
    rs = [torch.nn.Parameter(t.clone()) for t in init]
    opt = LARC(torch.optim.SGD(ps, lr=0.1, momentum=0.9, weight_decay=5e-4), trust_coefficient=0.02, clip=clip)
    ropt = torch.optim.SGD(rs, lr=0.1, momentum=0.9, weight_decay=5e-4)
    for it in range(4):
        for p, r in zip(ps, rs):
            g = torch.randn(p.shape) * (10.0 if it % 2 else 0.01)
            p.grad, r.grad = g.clone(), g.clone()
        if it == 2:
            ps[1].grad, rs[1].grad = None, None         # parameter without gradient
        opt.step()
        apex_ref.larc_rescale(rs, 0.1, 5e-4, 0.02, clip)     # LARC.py:75-92 ...
        saved = ropt.param_groups[0]['weight_decay']
        ropt.param_groups[0]['weight_decay'] = 0             # ... with the weight decay absorbed (:73-74, :95-97)
        ropt.step()
        ropt.param_groups[0]['weight_decay'] = saved
        assert opt.optim.param_groups[0]['weight_decay'] == 5e-4      # handed back after the step
        for p, r in zip(ps, rs):
            if r.grad is not None:
                assert torch.allclose(p.grad, r.grad, rtol=2e-6, atol=0)
            assert torch.allclose(p.data, r.data, rtol=2e-6, atol=1e-9)
    assert cpu_backend.calls.count("larc") == 4
    assert repr(opt) == repr(opt.optim) and opt.state_dict().keys() == ropt.state_dict().keys()


def test_two_param_groups_with_their_own_hyperparameters(cpu_backend):
    from apex_b200.fp16_utils import FP16_Optimizer
    from oracle import apex_ref
    torch.manual_seed(11)
    a = [(torch.randn(50) * 0.1).half(), (torch.randn(6, 6) * 0.1).half()]
    b = [(torch.randn(17) * 0.1).half(), torch.randn(9) * 0.1]

    def make(pa, pb):
        return torch.optim.SGD([{"params": pa, "lr": 0.5, "momentum": 0.9},
                                {"params": pb, "lr": 0.05, "momentum": 0.0, "weight_decay": 0.01}], lr=1.0)
    pa, pb = [torch.nn.Parameter(t.clone()) for t in a], [torch.nn.Parameter(t.clone()) for t in b]
    ra, rb = [torch.nn.Parameter(t.clone()) for t in a], [torch.nn.Parameter(t.clone()) for t in b]
    opt = FP16_Optimizer(make(pa, pb), static_loss_scale=16.0, verbose=False)
    assert len(opt.fp16_groups) == 2 and len(opt.fp32_from_fp32_groups[1]) == 1

    class Ref(apex_ref.MixedPrecisionStep):       # same composition, two groups
        pass
    # the restatement takes one flat list; build its optimizer with the same two groups
    ref = apex_ref.MixedPrecisionStep(ra + rb, lambda owned: make(owned[:2], owned[2:]), static_loss_scale=16.0)
    for it in range(4):
        for p, r in zip(pa + pb, ra + rb):
            g = (torch.randn(p.shape) * 16).to(p.dtype)
            p.grad, r.grad = g.clone(), g.clone()
        if it == 2:
            opt.param_groups[0]["lr"] = 0.25               # schedulers edit the group dicts in place
            ref.optimizer.param_groups[0]["lr"] = 0.25
        opt.update_master_grads()
        opt.step()
        ref.update_master_grads()
        ref.step()
    assert cpu_backend.calls.count("fused_sgd") == 8        # one launch per group per step
    for m, rm in zip(opt.fp32_from_fp16_groups[0] + opt.fp32_from_fp16_groups[1], ref.masters):
        assert torch.equal(m.data, rm.data)
    assert torch.equal(pb[1].data, rb[1].data)
    # a group added after wrapping is left to the inner optimizer (generic path), as in the reference
    extra = torch.nn.Parameter(torch.ones(4))
    opt.optimizer.add_param_group({"params": [extra]})
    extra.grad = torch.ones(4)
    for p in pa + pb:
        p.grad = torch.zeros_like(p)
    opt.update_master_grads()
    opt.step()
    assert torch.allclose(extra.data, torch.full((4,), 1.0 - 1.0 / 16))   # lr 1.0, its gradient was unscaled too


def test_channels_last_tensors_are_handled_in_place(cpu_backend):
    """Conv weights (and their gradients) in torch.channels_last are dense but not row-major; the
    kernels walk memory order, which is legal because gradient, master, state and fp16 copy all
    share the parameter's strides."""
    from apex_b200.fp16_utils import FP16_Optimizer
    from oracle import apex_ref
    torch.manual_seed(13)
    init = [(torch.randn(8, 4, 3, 3) * 0.1).half().contiguous(memory_format=torch.channels_last),
            (torch.randn(16) * 0.1).half(),
            (torch.randn(6, 8, 1, 1) * 0.1).half().contiguous(memory_format=torch.channels_last)]
    assert not init[0].is_contiguous()
    ps = [torch.nn.Parameter(t.clone(memory_format=torch.preserve_format)) for t in init]
    rs = [torch.nn.Parameter(t.clone(memory_format=torch.preserve_format)) for t in init]
    make = lambda q: torch.optim.SGD(q, lr=0.1, momentum=0.9, weight_decay=1e-3)     # noqa: E731
    opt = FP16_Optimizer(make(ps), dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 64.0}, verbose=False)
    ref = apex_ref.MixedPrecisionStep(rs, make, dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 64.0})
    assert opt.fp32_from_fp16_groups[0][0].stride() == ps[0].stride()
    for it in range(4):
        for p, r in zip(ps, rs):
            g = (torch.randn(p.shape) * 8).half()
            if p.dim() == 4:
                g = g.contiguous(memory_format=torch.channels_last)     # autograd keeps the parameter's layout
            if it == 1 and p is ps[2]:
                g[3, 5, 0, 0] = float("inf")
            p.grad, r.grad = g.clone(memory_format=torch.preserve_format), g.clone(memory_format=torch.preserve_format)
        opt.update_master_grads()
        if it == 2:
            assert torch.equal(opt.fp32_from_fp16_groups[0][0].grad, ps[0].grad.float() * (1. / opt.loss_scale))
        opt.step()
        ref.update_master_grads()
        ref.step()
        assert opt.overflow == ref.overflow
    assert "fused_sgd" in cpu_backend.calls
    for m, rm, p, r in zip(opt.fp32_from_fp16_groups[0], ref.masters, ps, rs):
        assert torch.equal(m.data, rm.data) and torch.equal(p.data, r.data)
        assert torch.equal(opt.optimizer.state[m]["momentum_buffer"], ref.optimizer.state[rm]["momentum_buffer"])


def test_larc_against_golden_runs_of_the_reference(golden_json, golden_arrays, cpu_backend):
    from apex_b200.parallel import LARC
    meta = golden_json["larc"]
    for name, clip in (("larc_clip", True), ("larc_scale", False)):
        ps = [torch.nn.Parameter(tensor_from(golden_arrays["%s/init/%d" % (name, i)])) for i in range(4)]
        opt = LARC(torch.optim.SGD(ps, lr=meta["lr"], momentum=meta["momentum"], weight_decay=meta["weight_decay"]),
                   trust_coefficient=meta["trust_coefficient"], clip=clip)
        for it in range(meta["iters"]):
            for i, p in enumerate(ps):
                p.grad = tensor_from(golden_arrays["%s/grad/%d/%d" % (name, it, i)])
            if it == meta["no_grad"]["iteration"]:
                ps[meta["no_grad"]["param"]].grad = None
            opt.step()
        for i, p in enumerate(ps):
            want = golden_arrays["%s/final/%d" % (name, i)]
            # norms are reduced in another order than torch.norm: a few fp32 ulps on the rate
            assert np.allclose(p.data.numpy(), want, rtol=3e-6, atol=1e-9), (name, i)


def test_oracle_larc_restatement_is_pinned_too(golden_json, golden_arrays):
    from oracle import apex_ref
    meta = golden_json["larc"]
    for name, clip in (("larc_clip", True), ("larc_scale", False)):
        ps = [torch.nn.Parameter(tensor_from(golden_arrays["%s/init/%d" % (name, i)])) for i in range(4)]
        inner = torch.optim.SGD(ps, lr=meta["lr"], momentum=meta["momentum"], weight_decay=meta["weight_decay"])
        for it in range(meta["iters"]):
            for i, p in enumerate(ps):
                p.grad = tensor_from(golden_arrays["%s/grad/%d/%d" % (name, it, i)])
            if it == meta["no_grad"]["iteration"]:
                ps[meta["no_grad"]["param"]].grad = None
            apex_ref.larc_rescale(ps, meta["lr"], meta["weight_decay"], meta["trust_coefficient"], clip)
            inner.param_groups[0]["weight_decay"] = 0
            inner.step()
            inner.param_groups[0]["weight_decay"] = meta["weight_decay"]
        for i, p in enumerate(ps):
            assert np.array_equal(p.data.numpy(), golden_arrays["%s/final/%d" % (name, i)]), (name, i)


def test_amp_handle_and_wrapper_against_golden_runs_of_the_reference(golden_json, cpu_backend):
    from apex_b200.amp import AmpHandle
    g = golden_json["amp_handle"]
    w = torch.nn.Parameter(torch.tensor([1.0, 2.0, -1.0]))
    opt = torch.optim.SGD([w], lr=0.5)
    handle = AmpHandle()
    for c, row in zip(g["coeffs"], g["log"]):
        opt.zero_grad()
        loss = (w * torch.tensor([float(x) for x in c])).sum()
        with handle.scale_loss(loss, opt) as scaled:
            scaled.backward()
        opt.step()
        assert w.data.tolist() == row["w"] and handle._default_scaler.loss_scale() == row["scale"]
    g2 = golden_json["amp_optim_wrapper"]
    w2 = torch.nn.Parameter(torch.tensor([1.0, 1.0]))
    wrapped = AmpHandle().wrap_optimizer(torch.optim.SGD([w2], lr=0.25), num_loss=2)
    for it, row in enumerate(g2["log"]):
        wrapped.zero_grad()
        with wrapped.scale_loss((w2 * 2).sum()) as s0:
            s0.backward()
        second = torch.tensor([float("inf"), 1.0]) if it == 1 else torch.tensor([3.0, -1.0])
        with wrapped.scale_loss((w2 * second).sum()) as s1:
            s1.backward()
        wrapped.step()
        assert w2.data.tolist() == row["w"]
        assert [repr(x) for x in w2.grad.tolist()] == row["grad"]      # incl. the un-unscaled gradient of the bad loss
        assert [sc.loss_scale() for sc in wrapped._loss_scaler] == row["scales"]


def test_unfused_steps_are_counted_and_announced_once(cpu_backend):
    """An inner optimizer without a fused kernel goes through its own step() + one copy-back launch, as
    the reference does with every optimizer (fp16_optimizer.py:380-382): legitimate, but not silent."""
    import warnings
    from apex_b200.fp16_utils import FP16_Optimizer
    ps = [torch.nn.Parameter((torch.randn(40) * 0.1).half()), torch.nn.Parameter(torch.randn(9) * 0.1)]
    opt = FP16_Optimizer(torch.optim.RMSprop(ps, lr=0.01), static_loss_scale=4.0, verbose=False)
    with warnings.catch_warnings(record=True) as seen:
        warnings.simplefilter("always")
        for it in range(3):
            for p in ps:
                p.grad = torch.ones_like(p)
            opt.update_master_grads()
            opt.step()
    assert opt.unfused_steps == 3
    assert len([w for w in seen if "not taking the fused update for RMSprop" in str(w.message)]) == 1
    assert "master_to_model" in cpu_backend.calls
    # a fused optimizer never counts
    qs = [torch.nn.Parameter((torch.randn(40) * 0.1).half())]
    opt2 = FP16_Optimizer(torch.optim.SGD(qs, lr=0.01), static_loss_scale=4.0, verbose=False)
    qs[0].grad = torch.ones_like(qs[0])
    opt2.update_master_grads()
    opt2.step()
    assert opt2.unfused_steps == 0 and "fused_sgd" in cpu_backend.calls


def test_storage_swapped_under_an_unchanged_identity_is_noticed(cpu_backend):
    """`p.data = ...` keeps the python object and changes the address (EMA swap, module.half()): the
    cached launch plan and tables must not keep the old addresses."""
    from apex_b200.fp16_utils import FP16_Optimizer
    from oracle import apex_ref
    init = [(torch.randn(300) * 0.1).half(), (torch.randn(7, 9) * 0.1).half(), (torch.randn(50) * 0.1).half()]
    ps = [torch.nn.Parameter(t.clone()) for t in init]
    rs = [torch.nn.Parameter(t.clone()) for t in init]
    make = lambda q: torch.optim.SGD(q, lr=0.5, momentum=0.9)      # noqa: E731
    opt = FP16_Optimizer(make(ps), static_loss_scale=8.0, verbose=False)
    ref = apex_ref.MixedPrecisionStep(rs, make, static_loss_scale=8.0)
    for it in range(6):
        gs = [(torch.randn(t.shape) * 8).half() for t in init]
        if it == 3:
            for p, r in zip(ps, rs):                          # new storage, same objects, same values
                p.data = p.data.clone()
                r.data = r.data.clone()
            gs_objs = [p.grad for p in ps]
            for g_old in gs_objs:                             # gradients too: same objects, new storage
                g_old.data = g_old.data.clone()
        for p, r, g in zip(ps, rs, gs):
            if p.grad is None:
                p.grad = g.clone()
            else:
                p.grad.data.copy_(g)
            r.grad = g.clone()
        opt.update_master_grads()
        ref.update_master_grads()
        opt.step()
        ref.step()
    for p, r in zip(ps, rs):
        assert torch.equal(p.data, r.data)
    for m, rm in zip(opt.fp32_from_fp16_groups[0], ref.masters):
        assert torch.equal(m.data, rm.data)


def test_bucket_layout_alignment_and_views():
    """parallel/buckets.py: members on 32 B boundaries, every rank's slice on a 2 KiB boundary, views alias
    the flat buffer with the member's shape AND strides (channels-last too) and own version counters."""
    from apex_b200.parallel.buckets import GradBucket, bucket_layout, SLICE_ALIGN_BYTES
    for world in (1, 2, 3, 8):
        for dtype in (torch.float16, torch.float32):
            like = [torch.zeros(n, dtype=dtype) for n in (5, 4099, 64, 1000, 1)]
            offs, padded = bucket_layout(like, world)
            es = like[0].element_size()
            assert all((o * es) % 32 == 0 for o in offs)
            assert all(b - a >= t.numel() for a, b, t in zip(offs, offs[1:] + [padded], like))
            assert (padded * es) % (SLICE_ALIGN_BYTES * world) == 0
            assert (padded // world * es) % 16 == 0                       # the C ABI's own rule
    cl = torch.zeros(4, 3, 5, 5).to(memory_format=torch.channels_last)
    like = [torch.zeros(7, 9), cl, torch.zeros(11)]
    home = GradBucket(like, 2)
    assert home.flat.numel() == home.padded and float(home.flat.abs().sum()) == 0.0
    for v, t in zip(home.views, like):
        assert v.shape == t.shape and v.stride() == t.stride()
        lo = home.flat.data_ptr()
        assert lo <= v.data_ptr() < lo + home.flat.numel() * 4
    home.views[0].add_(1.0)
    assert home.views[0]._version == 1 and home.views[1]._version == 0     # not shared
    assert float(home.flat.sum()) == 63.0                                   # ... but the memory is
    assert home.fits(like) and not home.fits(like[:2]) and not home.fits([like[0], like[1].contiguous(), like[2]])
    assert home.strays(list(home.views)) == []
    assert home.strays([home.views[0], torch.zeros_like(cl), home.views[2]]) == [1]
    with pytest.raises(TypeError):
        GradBucket([torch.zeros(3), torch.zeros(3).half()], 2)
    with pytest.raises(RuntimeError, match="dense"):
        GradBucket([torch.zeros(4, 4)[:, :2]], 2)
