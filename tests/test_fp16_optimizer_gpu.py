"""FP16_Optimizer / scalers / fp16util on the GPU, end to end through the C ABI.

Three kinds of evidence:
  * the golden runs of the unmodified reference (tests/golden) replayed on the GPU;
  * the reference's composition restated in torch (oracle/apex_ref.py) run on the SAME
    GPU on the same inputs -- that is exactly what the reference would execute there;
  * at BASELINE.json's full sizes, properties that need no second copy of the state.
"""
import contextlib
import io

import numpy as np
import pytest
import torch

from apex_b200 import _C
from apex_b200.fp16_utils import FP16_Optimizer, DynamicLossScaler
from oracle import apex_ref
from tests.golden_cases import CASES, bits, grads_for, init_params, make_inner, ulp_diff
from tests.gpu_util import dev, gen

pytestmark = pytest.mark.gpu


def _replay(name, meta, arrays):
    params = init_params(arrays, name, dev())
    opt = FP16_Optimizer(make_inner(meta)(params), verbose=False, **meta["fp16_kwargs"])
    rows, msgs, norms = [], [], []
    for it in range(meta["iters"]):
        for p, g in zip(params, grads_for(arrays, name, it, len(params), dev())):
            p.grad = g
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            opt.update_master_grads()
            if meta["clip"] is not None:
                norms.append(float(opt.clip_master_grads(meta["clip"])))
            opt.step()
        msgs.append(buf.getvalue())
        ls = opt.loss_scaler
        rows.append({"overflow": bool(opt.overflow), "loss_scale": float(opt.loss_scale),
                     "cur_iter": getattr(ls, "cur_iter", None),
                     "last_overflow_iter": getattr(ls, "last_overflow_iter", None)})
    return opt, rows, msgs, norms


@pytest.mark.parametrize("name", CASES)
def test_golden_replay_on_gpu(name, golden_json, golden_arrays):
    meta = golden_json[name]
    adam = meta["optimizer"] in ("Adam", "AdamW")
    before = _C.launches
    opt, rows, msgs, norms = _replay(name, meta, golden_arrays)
    assert _C.launches > before                     # our kernels ran, not a fallback
    assert rows == meta["rows"]                     # overflow decisions + scale trajectory: bit-exact
    assert msgs == meta["messages"]
    steps_taken = sum(1 for r in rows if not r["overflow"])
    for i, (m, p) in enumerate(zip(opt.fp32_from_fp16_groups[0], opt.fp16_groups[0])):
        want = golden_arrays["%s/final_master/%d" % (name, i)]
        got = m.data.cpu().numpy()
        if not adam and meta["clip"] is None:
            assert ulp_diff(got, want).max() <= 1
            assert np.array_equal(got, want), (name, i)
            assert np.array_equal(bits(p.data), golden_arrays["%s/final_model/%d" % (name, i)])
        else:
            # golden = torch CPU kernels (other association in Adam, norm reduction order for clip)
            scale = np.maximum(np.abs(want), meta["opt_kwargs"]["lr"])
            assert (np.abs(got - want) <= 4 * max(steps_taken, 1) * np.spacing(scale.astype(np.float32))).all()
        assert torch.equal(p.data, m.data.half())
    if meta["clip"] is not None:
        assert np.allclose(norms, meta["norms"], rtol=1e-5, atol=0)
    for i, p in enumerate(opt.fp32_from_fp32_groups[0]):
        assert np.array_equal(p.data.cpu().numpy(), golden_arrays["%s/final_fp32/%d" % (name, i)])


def _mlp_shapes():
    # BASELINE config 1: Linear(1024,4096) - ReLU - Linear(4096,1024)
    return [(4096, 1024), (4096,), (1024, 4096), (1024,)]


def _resnet_like_shapes():
    # a slice of config 3: fp16 conv weights of odd sizes + fp32 batch-norm vectors of 64..2048
    conv = [(64, 3, 7, 7), (64, 64, 1, 1), (64, 64, 3, 3), (256, 64, 1, 1), (512, 256, 1, 1), (1000, 2048), (1000,)]
    bn = [(64,), (64,), (256,), (256,), (2048,), (2048,)]
    return conv, bn


def _make_pair(shapes16, shapes32, make_inner, seed, **fp16_kwargs):
    g = gen(seed)
    init16 = [(torch.randn(s, generator=g, device=dev()) * 0.02).half() for s in shapes16]
    init32 = [torch.randn(s, generator=g, device=dev()) * 0.02 + 1.0 for s in shapes32]
    mine = [torch.nn.Parameter(t.clone()) for t in init16 + init32]
    theirs = [torch.nn.Parameter(t.clone()) for t in init16 + init32]
    opt = FP16_Optimizer(make_inner(mine), verbose=False, **fp16_kwargs)
    ref = apex_ref.MixedPrecisionStep(theirs, make_inner, **fp16_kwargs)
    return mine, theirs, opt, ref, g


def _same_run(mine, theirs, opt, ref, g, iters, inject, clip=None, exact=True, lr=None):
    """exact=False (fused clip_master_grads only): the clip factor comes from a norm summed in
    another order than torch's and may differ from the reference's in its last bits (asserted below:
    the returned norm within 4 fp32 ulp, hence the factor within ~4 ulp = 2^-21 relative).  With plain
    SGD a relative change eps of the factor moves element i by lr * |clipped gradient i| * eps, plus one
    rounding of the parameter: the masters must stay within the element-wise budget
        sum over committed steps of ( lr * |g_i| * 2^-21 + ulp(p_i) ),
    g_i being the reference's own clipped master gradient."""
    committed = 0
    budget = [torch.zeros_like(rm, dtype=torch.float64) for rm in ref.masters] if not exact else None
    for it in range(iters):
        scale = float(opt.loss_scale)
        assert scale == float(ref.loss_scale)
        for k, (p, q) in enumerate(zip(mine, theirs)):
            grad = torch.randn(p.shape, generator=g, device=dev()) * (1e-3 * min(scale, 2.0**18))
            grad = grad.to(p.dtype)
            if it in inject and inject[it][0] == k:
                grad.view(-1)[inject[it][1] % grad.numel()] = inject[it][2]
            p.grad = grad.clone()
            q.grad = grad.clone()
        with contextlib.redirect_stdout(io.StringIO()):
            opt.update_master_grads()
            ref.update_master_grads()
            if clip is not None:
                a, b = opt.clip_master_grads(clip), ref.clip_master_grads(clip)
                if exact:
                    assert float(a) == float(b)
                else:
                    assert abs(float(a) - float(b)) <= 4 * float(np.spacing(np.float32(abs(float(b)))))
            opt.step()
            ref.step()
        committed += int(not opt.overflow)
        if not exact and not ref.overflow:
            for b, rm in zip(budget, ref.masters):
                b += lr * rm.grad.abs().double() * 2.0**-21
                b += torch.ldexp(torch.ones_like(b), torch.frexp(rm.data)[1] - 24)        # ulp of the fp32 value
        assert opt.overflow == ref.overflow, it
        assert float(opt.loss_scale) == float(ref.loss_scale), it
        if opt.dynamic_loss_scale:
            assert opt.loss_scaler.cur_iter == ref.scaler.cur_iter
            assert opt.loss_scaler.last_overflow_iter == ref.scaler.last_overflow_iter
    seen_budget = []
    for m, rm in zip(opt.fp32_from_fp16_groups[0], ref.masters):
        d = ulp_diff(m.data.cpu().numpy(), rm.data.cpu().numpy())
        if exact:
            assert d.max() <= 1                          # north_star tolerance
            assert torch.equal(m.data, rm.data)          # measured: identical
        else:
            diff = (m.data.double() - rm.data.double()).abs()
            used = float((diff / budget[len(seen_budget)]).max())
            seen_budget.append(used)
            print("fused clip: %.3f of the element-wise error budget used after %d committed steps; "
                  "max distance %d ulp of the parameter itself, %.4f%% of elements bit-equal"
                  % (used, committed, int(d.max()), 100.0 * (d == 0).mean()))
            assert used <= 1.0
    for p, q in zip(mine, theirs):
        if exact:
            assert torch.equal(p.data, q.data)
    for (p, q) in zip(opt.optimizer.param_groups[0]["params"], ref.optimizer.param_groups[0]["params"]):
        sa, sb = opt.optimizer.state[p], ref.optimizer.state[q]
        assert sorted(sa) == sorted(sb)
        for key in sa:
            if exact or key == "step":
                assert torch.equal(torch.as_tensor(sa[key]).cpu(), torch.as_tensor(sb[key]).cpu()), key


INF, NAN = float("inf"), float("nan")


@pytest.mark.parametrize("opt_name,kw", [
    ("SGD", dict(lr=0.1, momentum=0.9, weight_decay=1e-4)),
    ("SGD", dict(lr=0.1)),
    ("Adam", dict(lr=1e-3)),
    ("AdamW", dict(lr=1e-3, weight_decay=0.01)),
])
def test_config1_mlp_against_reference_composition_on_gpu(opt_name, kw):
    cls = getattr(torch.optim, opt_name)
    mine, theirs, opt, ref, g = _make_pair(_mlp_shapes(), [], lambda ps: cls(ps, **kw), 21, dynamic_loss_scale=True,
                                           dynamic_loss_args={"init_scale": 2**17, "scale_window": 5})
    inject = {2: (0, 4096 * 1024 - 1, INF), 3: (3, 0, NAN), 9: (1, 77, -INF), 16: (2, 1234567, INF)}
    _same_run(mine, theirs, opt, ref, g, 24, inject)


def test_default_dynamic_scaler_walks_down_like_the_reference():
    """scale 2^32: gradients of 1e-3*scale overflow fp16 until the scale has halved enough."""
    mine, theirs, opt, ref, g = _make_pair([(257, 33), (1000,)], [], lambda ps: torch.optim.SGD(ps, lr=0.1, momentum=0.9),
                                           22, dynamic_loss_scale=True)
    seen = []
    for it in range(14):
        scale = float(opt.loss_scale)
        for p, q in zip(mine, theirs):
            grad = (torch.randn(p.shape, generator=g, device=dev()) * (1e-3 * scale)).half()
            p.grad, q.grad = grad.clone(), grad.clone()
        with contextlib.redirect_stdout(io.StringIO()) as out:
            opt.update_master_grads()
            opt.step()
        ref.update_master_grads()
        ref.step()
        seen.append((opt.overflow, float(opt.loss_scale)))
        assert (ref.overflow, float(ref.loss_scale)) == seen[-1]
        if opt.overflow:
            assert out.getvalue().startswith("OVERFLOW! Skipping step. Attempted loss scale: ")
    assert seen[0] == (True, 2.0**31) and any(not o for o, _ in seen)
    for m, rm in zip(opt.fp32_from_fp16_groups[0], ref.masters):
        assert torch.equal(m.data, rm.data)


def test_config3_slice_mixed_fp16_fp32_params():
    conv, bn = _resnet_like_shapes()
    mine, theirs, opt, ref, g = _make_pair(conv, bn, lambda ps: torch.optim.SGD(ps, lr=0.1, momentum=0.9, weight_decay=1e-4),
                                           23, dynamic_loss_scale=True,
                                           dynamic_loss_args={"init_scale": 2**16, "scale_window": 4})
    n16 = len(conv)
    inject = {1: (n16 + 2, 5, INF), 5: (2, 9, NAN), 6: (n16, 63, -INF)}      # fp32 BN grads overflow too
    _same_run(mine, theirs, opt, ref, g, 14, inject)
    # the fp32-native gradients were unscaled in place, as the reference leaves them
    for p, q in zip(mine[n16:], theirs[n16:]):
        assert torch.equal(p.grad, q.grad)


def test_config4_clip_master_grads():
    """word_language_model flow: backward -> clip_master_grads -> step (main_fp16_optimizer.py:188-204)."""
    shapes = [(3331, 150), (600, 150), (600, 150), (600,), (600,), (3331,)]
    mine, theirs, opt, ref, g = _make_pair(shapes, [], lambda ps: torch.optim.SGD(ps, lr=1.0), 24,
                                           dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 2**15, "scale_window": 6})
    _same_run(mine, theirs, opt, ref, g, 12, {3: (1, 17, INF), 4: (5, 0, NAN)}, clip=0.25, exact=False, lr=1.0)
    # overflow -> clip returns -1 without touching anything
    mine[0].grad[0, 0] = INF
    opt.update_master_grads()
    assert opt.clip_master_grads(0.25) == -1


def test_config4_clip_master_grads_exact_path():
    """fused_clip = False: master gradients are materialised and torch's own clip_grad_norm_ runs on
    them, as in the reference (fp16_optimizer.py:243-265) -- bit-exact, norm included."""
    shapes = [(3331, 150), (600, 150), (600, 150), (600,), (600,), (3331,)]
    mine, theirs, opt, ref, g = _make_pair(shapes, [], lambda ps: torch.optim.SGD(ps, lr=20.0), 24,
                                           dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 2**15, "scale_window": 6})
    opt.fused_clip = False
    _same_run(mine, theirs, opt, ref, g, 12, {3: (1, 17, INF), 4: (5, 0, NAN)}, clip=0.25, exact=True)


def test_documented_exclusion_finite_fp32_values_whose_sum_overflows():
    """SURVEY 8a A11 / DESIGN section 3: the reference decides by float(x.float().sum()), so fp32
    gradients [3e38, 3e38] -- finite, but their SUM is inf -- count as an overflow there (golden
    `two_3e38_f32`: True; torch's reduction order decides such cases).  This implementation tests
    the ELEMENTS: the same input is NOT an overflow here.  Pinned so that the documented answer
    cannot move unnoticed.  fp16 gradients can never be in this situation (65504 * numel << 3.4e38)."""
    from apex_b200.fp16_utils import DynamicLossScaler
    x = torch.tensor([3e38, 3e38], device=dev())
    assert bool(torch.isinf(x.sum()))                                   # what the reference would see
    assert DynamicLossScaler._has_inf_or_nan(x) is False                # the documented answer here
    assert DynamicLossScaler._has_inf_or_nan(torch.tensor([3e38, float("inf")], device=dev())) is True
    big16 = torch.full((1000000,), 65504.0, device=dev()).half()
    assert DynamicLossScaler._has_inf_or_nan(big16) is False            # agrees with the reference's probe


def test_static_scale_does_no_check_and_one_launch():
    mine, theirs, opt, ref, g = _make_pair(_mlp_shapes(), [], lambda ps: torch.optim.SGD(ps, lr=0.1, momentum=0.9), 25,
                                           static_loss_scale=128.0)
    _same_run(mine, theirs, opt, ref, g, 3, {})
    for p in mine:
        p.grad = torch.randn_like(p)
    before = _C.launches
    opt.update_master_grads()
    opt.step()
    assert _C.launches - before == 1                # the fused step and nothing else
    assert opt.overflow is False
    # dynamic: check + fused step = 2 launches, and no table rebuild once pointers are stable
    mine, theirs, opt, ref, g = _make_pair(_mlp_shapes(), [], lambda ps: torch.optim.Adam(ps, lr=1e-3), 26,
                                           dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 256.0})
    grads = [torch.randn_like(p) for p in mine]
    rebuilds = []
    for it in range(4):
        for p, gr in zip(mine, grads):
            p.grad = gr
        before = _C.launches
        opt.update_master_grads()
        opt.step()
        assert _C.launches - before == 2
        rebuilds.append(opt._check_table.rebuilds + opt._step_tables[0].rebuilds)
    assert rebuilds[1] == rebuilds[-1]


def test_lazy_master_grads_on_gpu():
    p = torch.nn.Parameter((torch.arange(5000, dtype=torch.float32, device=dev()) / 977).half())
    opt = FP16_Optimizer(torch.optim.SGD([p], lr=0.5), static_loss_scale=3.0, verbose=False)
    master = opt.fp32_from_fp16_groups[0][0]
    p.grad = (torch.arange(5000, dtype=torch.float32, device=dev()) / 13 - 4).half()
    opt.update_master_grads()
    want = p.grad.float().mul_(1. / 3.0)
    assert torch.equal(master.grad, want)
    master.grad.mul_(2.0)
    before = master.data.clone()
    opt.step()
    assert torch.equal(master.data, torch.add(before, want * 2.0, alpha=-0.5))
    assert torch.equal(p.data, master.data.half())
    opt.zero_grad()
    assert float(p.grad.abs().sum()) == 0 and float(master.grad.abs().sum()) == 0


def test_has_overflow_api_and_fp16util_on_gpu(golden_json):
    g = golden_json["has_inf_or_nan"]
    f = DynamicLossScaler._has_inf_or_nan
    assert f(torch.tensor([INF, -INF], device=dev())) == g["inf_and_neg_inf_f32"]
    assert f(torch.tensor([1.0, NAN], device=dev()).half()) == g["one_nan_f16"]
    assert f(torch.full((1000000,), 65504.0, device=dev()).half()) == g["million_max_f16"]
    assert f(torch.tensor([0.5, -INF, 2.0], device=dev()).half()) == g["neg_inf_f16"]
    from apex_b200 import fp16_utils as U
    model = torch.nn.Sequential(torch.nn.Linear(37, 19), torch.nn.BatchNorm1d(19), torch.nn.Linear(19, 3)).to(dev())
    model = U.network_to_half(model)
    for flat in (False,):
        model_params, master_params = U.prep_param_lists(model, flat_master=flat)
        for p in model_params:
            p.grad = torch.randn_like(p)
        U.model_grads_to_master_grads(model_params, master_params)
        for p, m in zip(model_params, master_params):
            assert m.grad.dtype == torch.float32 and torch.equal(m.grad, p.grad.float())
            m.data.add_(0.001)
        U.master_params_to_model_params(model_params, master_params)
        for p, m in zip(model_params, master_params):
            assert torch.equal(p.data, m.data.to(p.dtype))
    half_only = torch.nn.Sequential(torch.nn.Linear(37, 19), torch.nn.Linear(19, 1000)).to(dev()).half()
    model_params, master_params = U.prep_param_lists(half_only, flat_master=True)
    for p in model_params:
        p.grad = torch.randn_like(p)
    U.model_grads_to_master_grads(model_params, master_params, flat_master=True)
    assert torch.equal(master_params[0].grad, torch.cat([p.grad.view(-1) for p in model_params]).float())
    master_params[0].data.mul_(1.5)
    U.master_params_to_model_params(model_params, master_params, flat_master=True)
    assert torch.equal(torch.cat([p.data.view(-1) for p in model_params]), master_params[0].data.half())


# ---- BASELINE.json full sizes: properties ----------------------------------------------------

def gpt2_shapes(L, d, V=50257, T=1024):
    per_layer = [(d,), (d,), (d, 3 * d), (3 * d,), (d, d), (d,), (d,), (d,), (d, 4 * d), (4 * d,), (4 * d, d), (d,)]
    return [(V, d), (T, d)] + per_layer * L + [(d,), (d,)]


def test_config2_gpt2_medium_full_size_properties():
    """355M fp16 params, 292 tensors (BASELINE config 2), fused unscale + overflow + SGD-momentum."""
    shapes = gpt2_shapes(24, 1024)
    assert len(shapes) == 292 and sum(int(np.prod(s)) for s in shapes) == 354823168
    g = gen(31)
    params = [torch.nn.Parameter((torch.randn(s, generator=g, device=dev()) * 0.02).half()) for s in shapes]
    opt = FP16_Optimizer(torch.optim.SGD(params, lr=0.1, momentum=0.9, weight_decay=1e-4), dynamic_loss_scale=True,
                         dynamic_loss_args={"init_scale": 2.0**16, "scale_window": 1000}, verbose=False)
    masters = opt.fp32_from_fp16_groups[0]
    grads = [(torch.randn(p.shape, generator=g, device=dev()) * (1e-3 * 65536)).half() for p in params]

    def checksum():
        return [float(m.data.double().sum()) for m in masters[:4]] + [float(p.data.double().sum()) for p in params[:4]]
    # step 1 against the formula, tensor by tensor, using torch's own kernels on the same GPU
    old = [m.data.clone() for m in masters]
    for p, gr in zip(params, grads):
        p.grad = gr
    opt.update_master_grads()
    opt.step()
    assert opt.overflow is False
    inv = torch.tensor(1. / 65536, device=dev())
    for m, o, gr, p in zip(masters, old, grads, params):
        gg = torch.add(gr.float() * inv, o, alpha=1e-4)
        assert torch.equal(m.data, torch.add(o, gg, alpha=-0.1))
        assert torch.equal(opt.optimizer.state[m]["momentum_buffer"], gg)
        assert torch.equal(p.data, m.data.half())
    del old
    # an overflow anywhere (last element of the last tensor) skips the whole step: nothing changes
    before = checksum()
    grads[-1].view(-1)[-1] = INF
    opt.update_master_grads()
    with contextlib.redirect_stdout(io.StringIO()) as out:
        opt.step()
    assert opt.overflow is True and opt.loss_scale == 2.0**15
    assert "OVERFLOW! Skipping step. Attempted loss scale: 65536.0, reducing to 32768.0" in out.getvalue()
    assert checksum() == before
    grads[-1].view(-1)[-1] = 0.0
    # idempotence of the model copy + the copy invariant after another step
    opt.update_master_grads()
    opt.step()
    assert opt.overflow is False
    for p, m in zip(params, masters):
        assert torch.equal(p.data, m.data.half())


def test_amp_handle_on_gpu():
    """apex.amp's scale_loss context (apex/amp/handle.py:32-63) with the two-launch unscale."""
    from apex_b200.amp import AmpHandle
    w = torch.nn.Parameter(torch.tensor([1.0, 2.0, 3.0], device=dev()))
    opt = torch.optim.SGD([w], lr=0.5)
    handle = AmpHandle()
    coeff = torch.tensor([1.0, 3.0, -2.0], device=dev())
    with handle.scale_loss((w * coeff).sum(), opt) as scaled:
        scaled.backward()
    assert torch.equal(w.grad, coeff)                       # unscaled in place by 1/2^16: exact
    opt.step()
    assert torch.equal(w.data, torch.tensor([0.5, 0.5, 4.0], device=dev()))
    opt.zero_grad()
    bad = torch.tensor([float("inf"), 1.0, 1.0], device=dev())
    with handle.scale_loss((w * bad).sum(), opt) as scaled:
        scaled.backward()
    before = w.data.clone()
    opt.step()                                              # one-shot no-op
    assert torch.equal(w.data, before) and handle._default_scaler.loss_scale() == 2.**15
    w.grad = torch.ones(3, device=dev())
    opt.step()
    assert torch.equal(w.data, before - 0.5)


def test_config3_resnet50_full_shapes_against_reference_composition():
    """All 161 ResNet-50 tensors (55 fp16 + 106 fp32 BatchNorm vectors of 64..2048 elements), SGD with
    momentum and weight decay, dynamic scaling, overflow injected into fp16 and fp32 gradients."""
    import bench
    spec = bench.resnet50_shapes()
    conv = [s for s, f in spec if not f]
    bn = [s for s, f in spec if f]
    assert len(conv) == 55 and len(bn) == 106
    mine, theirs, opt, ref, g = _make_pair(conv, bn, lambda ps: torch.optim.SGD(ps, lr=0.1, momentum=0.9, weight_decay=1e-4),
                                           41, dynamic_loss_scale=True,
                                           dynamic_loss_args={"init_scale": 2**15, "scale_window": 3})
    inject = {1: (54, 999, INF), 4: (55 + 100, 63, NAN), 5: (0, 0, -INF)}
    _same_run(mine, theirs, opt, ref, g, 10, inject)


def test_config4_lstm_full_shapes_with_clipping():
    """word_language_model LSTM, 135,891,278 fp16 parameters in 11 tensors (vocab 33,278 is not a
    multiple of 8: ragged tails), backward -> clip_master_grads(0.25) -> step with injected inf/NaN."""
    V, H = 33278, 1500
    shapes = [(V, H), (4 * H, H), (4 * H, H), (4 * H,), (4 * H,), (4 * H, H), (4 * H, H), (4 * H,), (4 * H,), (V, H), (V,)]
    assert sum(int(np.prod(s)) for s in shapes) == 135891278
    mine, theirs, opt, ref, g = _make_pair(shapes, [], lambda ps: torch.optim.SGD(ps, lr=20.0), 42,
                                           dynamic_loss_scale=True,
                                           dynamic_loss_args={"init_scale": 2**14, "scale_window": 4})
    inject = {2: (10, V - 1, INF), 3: (0, V * H - 3, NAN), 7: (5, 12345, -INF)}
    _same_run(mine, theirs, opt, ref, g, 10, inject, clip=0.25, exact=False, lr=20.0)


def test_config5_gpt2_xl_full_size_adam_properties():
    """1,557,611,200 fp16 params in 580 tensors with Adam (BASELINE config 5 / the bench workload):
    first step against torch's own kernels tensor by tensor, then skip-on-overflow and the copy invariant."""
    shapes = gpt2_shapes(48, 1600)
    assert len(shapes) == 580 and sum(int(np.prod(s)) for s in shapes) == 1557611200
    g = gen(43)
    params = [torch.nn.Parameter((torch.randn(s, generator=g, device=dev()) * 0.02).half()) for s in shapes]
    opt = FP16_Optimizer(torch.optim.Adam(params, lr=1e-3), dynamic_loss_scale=True,
                         dynamic_loss_args={"init_scale": 2.0**16}, verbose=False)
    masters = opt.fp32_from_fp16_groups[0]
    grads = [(torch.randn(p.shape, generator=g, device=dev()) * (1e-3 * 65536)).half() for p in params]
    for p, gr in zip(params, grads):
        p.grad = gr
    # reference composition for a sample of tensors (first, last, big, small, middle), computed BEFORE the step
    sample = [0, 1, 2, 5, 301, 578, 579]
    want = {}
    for i in sample:
        q = torch.nn.Parameter(masters[i].data.clone())
        q.grad = grads[i].float().mul_(1. / 65536)
        topt = torch.optim.Adam([q], lr=1e-3, foreach=True)
        topt.step()
        want[i] = (q.data, topt.state[q]["exp_avg"], topt.state[q]["exp_avg_sq"])
    opt.update_master_grads()
    opt.step()
    assert opt.overflow is False
    for i in sample:
        st = opt.optimizer.state[masters[i]]
        assert torch.equal(masters[i].data, want[i][0]), i
        assert torch.equal(st["exp_avg"], want[i][1]) and torch.equal(st["exp_avg_sq"], want[i][2]), i
        assert float(st["step"]) == 1.0
    for p, m in zip(params, masters):
        assert torch.equal(p.data, m.data.half())
    del want
    # overflow in the very last element: nothing moves, the step counters stay
    sums = [float(masters[i].data.double().sum()) for i in sample]
    grads[-1].view(-1)[-1] = NAN
    opt.update_master_grads()
    with contextlib.redirect_stdout(io.StringIO()):
        opt.step()
    assert opt.overflow is True and opt.loss_scale == 2.0**15
    assert [float(masters[i].data.double().sum()) for i in sample] == sums
    assert float(opt.optimizer.state[masters[0]]["step"]) == 1.0
    grads[-1].view(-1)[-1] = 0.0
    opt.update_master_grads()
    opt.step()
    assert opt.overflow is False and float(opt.optimizer.state[masters[579]]["step"]) == 2.0


def test_closure_retries_while_the_gradients_overflow():
    """_step_with_closure (apex/fp16_utils/fp16_optimizer.py:386-423): the closure is run again, with
    a smaller scale each time, until its gradients are finite."""
    p = torch.nn.Parameter(torch.ones(1000, device=dev()).half())
    opt = FP16_Optimizer(torch.optim.SGD([p], lr=0.25), dynamic_loss_scale=True,
                         dynamic_loss_args={"init_scale": 2.0**18}, verbose=False)
    seen = []

    def closure():
        seen.append(float(opt.loss_scale))
        g = torch.full((1000,), 0.5 * float(opt.loss_scale), device=dev())      # "scaled" gradient of 0.5
        p.grad = g.half()                                                          # inf in fp16 above 2^17 * 0.5
        opt.update_master_grads()
        return torch.tensor(1.0)
    with contextlib.redirect_stdout(io.StringIO()) as out:
        opt.step(closure)
    assert seen == [2.0**18, 2.0**17, 2.0**16]
    assert out.getvalue().count("OVERFLOW within closure! Skipping step.") == 2
    assert torch.equal(p.data, torch.full((1000,), 0.875, device=dev()).half())
    assert opt.first_closure_call_this_step is True


def test_channels_last_conv_model_trains_like_the_reference():
    """Real autograd gradients of a channels_last fp16 conv net: optimizer.backward / step against the
    reference composition (apex_ref) driven by the same losses."""
    def build():
        torch.manual_seed(3)
        net = torch.nn.Sequential(torch.nn.Conv2d(3, 16, 3, padding=1), torch.nn.ReLU(),
                                  torch.nn.Conv2d(16, 8, 1), torch.nn.Flatten(), torch.nn.Linear(8 * 8 * 8, 10))
        return net.to(dev()).half().to(memory_format=torch.channels_last)
    a, b = build(), build()
    assert not a[0].weight.is_contiguous()
    make = lambda ps: torch.optim.SGD(ps, lr=0.05, momentum=0.9, weight_decay=1e-4)     # noqa: E731
    kw = dict(dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 2.0**20, "scale_window": 3})
    opt = FP16_Optimizer(make(list(a.parameters())), verbose=False, **kw)
    ref = apex_ref.MixedPrecisionStep(list(b.parameters()), make, **kw)
    g = gen(44)
    for it in range(10):
        x = torch.randn(4, 3, 8, 8, generator=g, device=dev()).half().contiguous(memory_format=torch.channels_last)
        y = torch.randn(4, 10, generator=g, device=dev())
        a.zero_grad(set_to_none=True)
        b.zero_grad(set_to_none=True)
        with contextlib.redirect_stdout(io.StringIO()):
            opt.backward(torch.nn.functional.mse_loss(a(x).float(), y))
            opt.step()
            (torch.nn.functional.mse_loss(b(x).float(), y) * ref.loss_scale).backward()
            ref.update_master_grads()
            ref.step()
        assert opt.overflow == ref.overflow and float(opt.loss_scale) == float(ref.loss_scale), it
    assert float(opt.loss_scale) < 2.0**20      # the scale did walk down: overflowing iterations were skipped
    for m, rm, p, q in zip(opt.fp32_from_fp16_groups[0], ref.masters, a.parameters(), b.parameters()):
        assert m.stride() == p.stride()
        assert torch.equal(m.data, rm.data) and torch.equal(p.data, q.data)
