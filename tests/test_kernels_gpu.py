"""Parity of every CUDA kernel, called through the C ABI, with the C oracle (bit-exact) and
with the torch ops the reference composes on a GPU."""
import numpy as np
import pytest
import torch

from apex_b200 import _C
from oracle import c_oracle
from tests.gpu_util import SIZES, bits, dev, flag, gen, randn16, randn32, table, unaligned_view
from tests.golden_cases import ulp_diff

pytestmark = pytest.mark.gpu
INT32_MAX = 0x7FFFFFFF


def _mixed_grads(seed, unaligned_every=5):
    g = gen(seed)
    grads, flags = [], []
    for i, n in enumerate(SIZES):
        if i % 3 == 2:
            t, f = randn32(n, g), _C.TF_GRAD_F32
        else:
            t, f = randn16(n, g), 0
        if unaligned_every and i % unaligned_every == 4 and n > 0:
            t = unaligned_view(t)
        grads.append(t)
        flags.append(f)
    return grads, flags


def test_device_is_blackwell():
    sm, major, minor = _C.device_info()
    assert major == 10, "these kernels are compiled for sm_100a only"
    assert sm >= 100


@pytest.mark.parametrize("chunk", [2048, 16384, 65536])
def test_check_overflow_clean_and_first_bad(chunk):
    grads, flags = _mixed_grads(1)
    t = table([grads], flags, chunk)
    f = flag()
    _C.mt_check_overflow(t, 0, f)
    assert int(f.item()) == 0
    # largest finite values are not an overflow (loss_scaler.py probe: 1e6 x 65504 -> False)
    grads[-1].fill_(65504.0)
    _C.mt_check_overflow(t, 0, f)
    assert int(f.item()) == 0
    vals = [float("inf"), -float("inf"), float("nan")]
    k = 0
    for i, gt in enumerate(grads):
        n = gt.numel()
        if n == 0:
            continue
        for pos in sorted({0, n - 1, n // 2, min(n - 1, 3), max(0, n - 9)}):
            old = gt[pos].clone()
            gt[pos] = vals[k % 3]
            k += 1
            _C.flag_clear(f)
            _C.mt_check_overflow(t, 0, f)
            assert int(f.item()) == INT32_MAX - i, (i, pos)
            assert c_oracle.first_nonfinite(bits(gt)) == pos
            gt[pos] = old
    # several bad tensors: the word ends as the FIRST one; +inf and -inf in one tensor
    grads[3][0] = float("inf")
    grads[3][1] = -float("inf")
    grads[9][5] = float("nan")
    grads[15][77] = float("inf")
    _C.flag_clear(f)
    _C.mt_check_overflow(t, 0, f)
    assert int(f.item()) == INT32_MAX - 3
    # the flag is sticky until cleared
    grads[3][0] = 0
    grads[3][1] = 0
    _C.mt_check_overflow(t, 0, f)
    assert int(f.item()) == INT32_MAX - 3
    _C.flag_clear(f)
    _C.mt_check_overflow(t, 0, f)
    assert int(f.item()) == INT32_MAX - 9


@pytest.mark.parametrize("chunk", [2048, 65536])
def test_check_overflow_publish_mirrors_to_pinned_host_and_rearms(chunk):
    """The check whose last CTA publishes the verdict to pinned host memory and zeroes the NEXT word:
    same verdicts as apex_b200_mt_check_overflow, no memset and no device-to-host copy needed."""
    grads, flags = _mixed_grads(3)
    t = table([grads], flags, chunk)
    words = torch.zeros(4, dtype=torch.int32, device=dev())
    pair, ticket = (words[0:1], words[1:2]), words[2:3]
    host = torch.full((1,), -7, dtype=torch.int32).pin_memory()
    plain = flag()
    turn = 0
    for it, bad in enumerate([None, 5, None, 1, 11, 11, None]):
        saved = None
        if bad is not None:
            saved = grads[bad][grads[bad].numel() // 2].clone()
            grads[bad][grads[bad].numel() // 2] = float("nan") if it % 2 else float("inf")
        turn ^= 1
        _C.mt_check_overflow_publish(t, 0, pair[turn], ticket, host, pair[turn ^ 1])
        _C.flag_clear(plain)
        _C.mt_check_overflow(t, 0, plain)
        torch.cuda.synchronize()
        want = 0 if bad is None else INT32_MAX - bad
        assert int(plain.item()) == want
        assert int(pair[turn].item()) == want and int(host[0]) == want, (it, bad)
        assert int(pair[turn ^ 1].item()) == 0 and int(ticket.item()) == 0
        if bad is not None:
            grads[bad][grads[bad].numel() // 2] = saved
    # nothing to scan: the mirror and the re-armed word are still produced
    empty = table([[]])
    pair[0].fill_(9)
    _C.mt_check_overflow_publish(empty, 0, pair[1], ticket, host, pair[0])
    torch.cuda.synchronize()
    assert int(pair[0].item()) == 0 and int(ticket.item()) == 0


def test_check_overflow_empty_inputs():
    f = flag()
    t = table([[]])
    _C.mt_check_overflow(t, 0, f)
    t = table([[torch.empty(0, dtype=torch.float16, device=dev())]])
    _C.mt_check_overflow(t, 0, f)
    assert int(f.item()) == 0


@pytest.mark.parametrize("inv_scale", [1. / 65536, 1. / 3000.0, 0.37])
def test_unscale_modes(inv_scale):
    grads, flags = _mixed_grads(2)
    before = [bits(g).copy() for g in grads]

    def expect(i):
        b = before[i]
        if b.dtype == np.uint16:
            return c_oracle.f32_to_f16(c_oracle.f16_to_f32_scaled(b, inv_scale, True))
        out = b.copy()
        c_oracle.lib().ora_scale_f32(out.ctypes.data, out.size, inv_scale)
        return out
    t = table([grads], flags, 65536)
    f = flag()
    _C.mt_unscale(t, 0, inv_scale, f, 0)                      # clean flag: everything scaled
    for i, g in enumerate(grads):
        assert np.array_equal(bits(g), expect(i)), i
        if g.dtype == torch.float32:                           # and equal to torch's mul_ (fp16_optimizer.py:241)
            ref = torch.from_numpy(before[i]).to(dev()).mul_(inv_scale)
            assert torch.equal(g, ref)
    for i, g in enumerate(grads):
        g.copy_(torch.from_numpy(before[i].view(np.float16) if before[i].dtype == np.uint16 else before[i]))
    f.fill_(INT32_MAX - 6)                                      # tensor 6 is the first bad one
    _C.mt_unscale(t, 0, inv_scale, f, 0)                      # mode 0: nothing is touched
    for i, g in enumerate(grads):
        assert np.array_equal(bits(g), before[i]), i
    _C.mt_unscale(t, 0, inv_scale, f, 1)                      # mode 1: amp's partial mutation
    for i, g in enumerate(grads):
        assert np.array_equal(bits(g), expect(i) if i < 6 else before[i]), i


def test_amp_partial_mutation_kat(golden_json):
    """apex/amp/scaler.py:35-40 on [[4,4],[inf,4],[4,4]] with scale 4."""
    from apex_b200.amp import LossScaler
    ps = [torch.nn.Parameter(torch.zeros(2, device=dev())) for _ in range(3)]
    ps[0].grad = torch.tensor([4.0, 4.0], device=dev())
    ps[1].grad = torch.tensor([float("inf"), 4.0], device=dev())
    ps[2].grad = torch.tensor([4.0, 4.0], device=dev())
    sc = LossScaler()
    g = golden_json["amp_partial"]
    assert sc.unscale_and_update([{"params": ps}], 4.0) == g["skip"]
    assert [p.grad.tolist() for p in ps] == g["grads"]
    assert sc.loss_scale() == g["loss_scale_after"]


@pytest.mark.parametrize("apply_scale,inv_scale", [(False, 1.0), (True, 1. / 128), (True, 1. / 3.0)])
def test_grads_to_master_and_back(apply_scale, inv_scale):
    g = gen(3)
    src = [randn16(n, g, 50.0) for n in SIZES]
    src[5] = unaligned_view(src[5])
    dst = [torch.full((n,), -7.0, device=dev()) for n in SIZES]
    dst[9] = unaligned_view(dst[9])
    t = table([src, dst], None, 65536)
    _C.mt_grads_to_master(t, 0, 1, inv_scale, apply_scale, None)
    for s, d in zip(src, dst):
        assert np.array_equal(bits(d), c_oracle.f16_to_f32_scaled(bits(s), inv_scale, apply_scale))
        ref = s.float()                                        # fp16util.py:110 copy_, then :241 mul_
        if apply_scale:
            ref.mul_(inv_scale)
        assert torch.equal(d, ref)
    # predicated: a set flag leaves the destination alone
    f = flag()
    f.fill_(5)
    keep = [d.clone() for d in dst]
    for s in src:
        s.mul_(2)
    _C.mt_grads_to_master(t, 0, 1, inv_scale, apply_scale, f)
    assert all(torch.equal(a, b) for a, b in zip(keep, dst))
    # fp32 -> fp16 (round to nearest even), fp16util.py:129
    p32 = [randn32(n, g) for n in SIZES]
    p32[6] = unaligned_view(p32[6])
    p16 = [torch.zeros(n, dtype=torch.float16, device=dev()) for n in SIZES]
    t2 = table([p32, p16], None, 65536)
    _C.mt_master_to_model(t2, 0, 1, None)
    for s, d in zip(p32, p16):
        assert np.array_equal(bits(d), c_oracle.f32_to_f16(bits(s)))
        assert torch.equal(d, s.half())
    # ties and specials round like torch
    special = torch.tensor([1.0 + 2**-11, 1.0 + 3 * 2**-11, 65520.0, 65519.9, 1e-8, -0.0, float("inf"), float("nan"),
                            6.0e-8, 2.98e-8], device=dev())
    out = torch.zeros(10, dtype=torch.float16, device=dev())
    _C.mt_master_to_model(table([[special], [out]]), 0, 1, None)
    assert np.array_equal(bits(out), bits(special.half()))


def test_copy_zero():
    g = gen(4)
    for dtype in (torch.float16, torch.float32, torch.float64):
        src = [torch.randn(n, generator=g, device=dev()).to(dtype) for n in SIZES]
        src[7] = unaligned_view(src[7])
        dst = [torch.zeros(n, dtype=dtype, device=dev()) for n in SIZES]
        _C.mt_copy(table([src, dst], None, 65536), 0, 1, src[0].element_size(), None)
        assert all(torch.equal(a, b) for a, b in zip(src, dst))
    grads, flags = _mixed_grads(5)
    _C.mt_zero(table([grads], flags, 65536), 0, 2)
    assert all(float(x.float().abs().sum()) == 0 for x in grads)


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32, torch.float64])
@pytest.mark.parametrize("world", [2, 3, 8])
def test_flatten_scale_unflatten(dtype, world):
    g = gen(6)
    bucket = [(torch.randn(n, generator=g, device=dev()) * 100).to(dtype) for n in SIZES]
    bucket[8] = unaligned_view(bucket[8])
    total = sum(SIZES)
    flat = torch.empty(total, dtype=dtype, device=dev())
    t = table([bucket], None, 65536)
    _C.mt_flatten(t, 0, flat)
    want = torch.cat([b.reshape(-1) for b in bucket])          # _flatten_dense_tensors, distributed.py:12
    assert torch.equal(flat, want)
    out = [torch.zeros_like(b) for b in bucket]
    out[3] = unaligned_view(out[3])
    t_out = table([out], None, 65536)
    f = flag()
    _C.mt_scale_unflatten(t_out, flat, 0, world, True, f)
    ref = want.clone()
    ref /= world                                                # distributed.py:20, torch's own CUDA kernel
    off = 0
    for b in out:
        assert torch.equal(b, ref[off:off + b.numel()])
        off += b.numel()
    if dtype != torch.float64:
        avg = c_oracle.average(bits(want).copy(), world)
        off = 0
        for b in out:
            assert np.array_equal(bits(b), avg[off:off + b.numel()])
            off += b.numel()
    assert int(f.item()) == 0
    # a reduced sum that overflowed stays inf and raises the word at the right tensor
    flat[sum(SIZES[:10]) + 2] = float("inf")
    _C.mt_scale_unflatten(t_out, flat, 0, world, True, f)
    assert int(f.item()) == INT32_MAX - 10
    assert torch.isinf(out[10][2])
    # plain scatter (broadcast path)
    _C.mt_scale_unflatten(t_out, want, 0, 1.0, False, None)
    off = 0
    for b in out:
        assert torch.equal(b, want[off:off + b.numel()])
        off += b.numel()
    # in-place variant on one buffer
    one = want.clone()
    f2 = flag()
    _C.scale_flat(one, world, f2)
    assert torch.equal(one, ref) and int(f2.item()) == 0


SGD_VARIANTS = [
    dict(lr=0.1, momentum=0.9, dampening=0.0, weight_decay=1e-4, nesterov=False),
    dict(lr=0.1, momentum=0.9, dampening=0.0, weight_decay=0.0, nesterov=False),
    dict(lr=0.05, momentum=0.8, dampening=0.0, weight_decay=5e-4, nesterov=True),
    dict(lr=0.02, momentum=0.5, dampening=0.25, weight_decay=0.0, nesterov=False),
    dict(lr=0.3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False),
    dict(lr=0.3, momentum=0.0, dampening=0.0, weight_decay=0.01, nesterov=False),
]


def _step_inputs(seed, with_fp32=True):
    g = gen(seed)
    grads, masters, models, flags = [], [], [], []
    for i, n in enumerate(SIZES):
        p16 = randn16(n, g, 0.05)
        if with_fp32 and i % 4 == 3:                      # fp32-native parameter (e.g. BatchNorm)
            grads.append(randn32(n, g, 1e-2))
            masters.append(randn32(n, g, 0.05))
            models.append(None)
            flags.append(_C.TF_GRAD_F32 | _C.TF_NO_MODEL | _C.TF_NO_SCALE)
        else:
            grads.append(randn16(n, g, 30.0))
            masters.append(p16.float())
            models.append(p16.clone())
            flags.append(0)
    for k in (5, 10):
        grads[k] = unaligned_view(grads[k])
        masters[k] = unaligned_view(masters[k])
    return grads, masters, models, flags


@pytest.mark.parametrize("hp", SGD_VARIANTS)
@pytest.mark.parametrize("scale", [1.0, 128.0, 3000.0])
def test_fused_sgd_matches_oracle_and_torch(hp, scale):
    grads, masters, models, flags = _step_inputs(7)
    apply_scale = scale != 1.0
    inv = 1. / scale if apply_scale else 1.0
    use_mom = hp["momentum"] != 0
    bufs = [torch.zeros_like(m) if use_mom else None for m in masters]
    # oracle / torch copies
    o_p = [bits(m).copy() for m in masters]
    o_b = [np.zeros_like(x) for x in o_p]
    t_p = [torch.nn.Parameter(m.clone()) for m in masters]
    topt = torch.optim.SGD(t_p, foreach=True, **hp)
    t = table([grads, masters, bufs, models], None, 16384)
    for step in range(3):
        fl = [f | (_C.TF_FIRST_STEP if (use_mom and step == 0) else 0) for f in flags]
        t.update([grads, masters, bufs, models], fl)
        _C.mt_fused_sgd(t, hp["lr"], hp["momentum"], hp["dampening"], hp["weight_decay"], hp["nesterov"],
                        inv, apply_scale, None)
        for i in range(len(grads)):
            no_scale = bool(flags[i] & _C.TF_NO_SCALE)
            p16 = c_oracle.sgd(bits(grads[i]), o_p[i], o_b[i] if use_mom else None, first=(step == 0),
                               inv_scale=inv, apply_scale=apply_scale and not no_scale, **hp)
            assert np.array_equal(bits(masters[i]), o_p[i]), (step, i)
            if use_mom:
                assert np.array_equal(bits(bufs[i]), o_b[i]), (step, i)
            if models[i] is not None:
                assert np.array_equal(bits(models[i]), p16), (step, i)
            # the reference's composition on the GPU: copy_, mul_(1/scale), torch.optim.SGD
            tg = grads[i].float()
            if apply_scale and not no_scale:
                tg.mul_(inv)
            t_p[i].grad = tg
        topt.step()
        for i in range(len(grads)):
            d = ulp_diff(bits(masters[i]), bits(t_p[i].data))
            assert d.size == 0 or d.max() <= 1, (step, i, int(d.max()))       # north_star: within 1 ulp
            assert torch.equal(masters[i], t_p[i].data), (step, i)           # and in fact identical
            if use_mom:
                assert torch.equal(bufs[i], topt.state[t_p[i]]["momentum_buffer"])
            if models[i] is not None:
                assert torch.equal(models[i], t_p[i].data.half())
        for gt in grads:                                                      # fresh gradients next step
            gt.mul_(-0.7)


ADAM_VARIANTS = [
    dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, decoupled=False),
    dict(lr=2e-3, betas=(0.8, 0.99), eps=1e-6, weight_decay=0.01, decoupled=False),
    dict(lr=2e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.05, decoupled=True),
    dict(lr=1e-2, betas=(0.3, 0.9), eps=1e-8, weight_decay=0.0, decoupled=False),      # lerp weight >= 0.5
]


@pytest.mark.parametrize("hp", ADAM_VARIANTS)
@pytest.mark.parametrize("scale", [1.0, 65536.0, 3000.0])
def test_fused_adam_matches_oracle_and_torch(hp, scale):
    grads, masters, models, flags = _step_inputs(8)
    if scale > 1000:
        for gt in grads:
            if gt.dtype == torch.float16:
                gt.mul_(30.0)
    apply_scale = scale != 1.0
    inv = 1. / scale if apply_scale else 1.0
    ms = [torch.zeros_like(m) for m in masters]
    vs = [torch.zeros_like(m) for m in masters]
    o_p = [bits(m).copy() for m in masters]
    o_m = [np.zeros_like(x) for x in o_p]
    o_v = [np.zeros_like(x) for x in o_p]
    t_p = [torch.nn.Parameter(m.clone()) for m in masters]
    cls = torch.optim.AdamW if hp["decoupled"] else torch.optim.Adam
    topt = cls(t_p, lr=hp["lr"], betas=hp["betas"], eps=hp["eps"], weight_decay=hp["weight_decay"], foreach=True)
    t = table([grads, masters, ms, vs, models], flags, 16384)
    b1, b2 = hp["betas"]
    for step in range(1, 4):
        bc1 = 1 - b1 ** step
        bc2 = 1 - b2 ** step
        _C.mt_fused_adam(t, hp["lr"], b1, b2, hp["eps"], hp["weight_decay"], (hp["lr"] / bc1) * -1, bc2 ** 0.5,
                         hp["decoupled"], inv, apply_scale, None)
        for i in range(len(grads)):
            no_scale = bool(flags[i] & _C.TF_NO_SCALE)
            p16 = c_oracle.adam(bits(grads[i]), o_p[i], o_m[i], o_v[i], hp["lr"], b1, b2, hp["eps"],
                                hp["weight_decay"], step, decoupled=hp["decoupled"], inv_scale=inv,
                                apply_scale=apply_scale and not no_scale)
            assert np.array_equal(bits(ms[i]), o_m[i]), (step, i)
            assert np.array_equal(bits(vs[i]), o_v[i]), (step, i)
            assert np.array_equal(bits(masters[i]), o_p[i]), (step, i)
            if models[i] is not None:
                assert np.array_equal(bits(models[i]), p16), (step, i)
            tg = grads[i].float()
            if apply_scale and not no_scale:
                tg.mul_(inv)
            t_p[i].grad = tg
        topt.step()
        for i in range(len(grads)):
            st = topt.state[t_p[i]]
            assert torch.equal(ms[i], st["exp_avg"]), (step, i)
            assert torch.equal(vs[i], st["exp_avg_sq"]), (step, i)
            d = ulp_diff(bits(masters[i]), bits(t_p[i].data))
            assert d.size == 0 or d.max() <= 1, (step, i, int(d.max()))
            assert torch.equal(masters[i], t_p[i].data), (step, i)
        for gt in grads:
            gt.mul_(-0.7)


def test_fused_step_is_a_noop_when_the_flag_is_set():
    grads, masters, models, flags = _step_inputs(9)
    bufs = [torch.randn_like(m) for m in masters]
    vs = [torch.rand_like(m) for m in masters]
    keep = [[x.clone() if x is not None else None for x in lst] for lst in (masters, bufs, vs, models)]
    f = flag()
    f.fill_(INT32_MAX - 2)
    t = table([grads, masters, bufs, models], flags, 16384)
    _C.mt_fused_sgd(t, 0.1, 0.9, 0.0, 1e-4, False, 1. / 128, True, f)
    t5 = table([grads, masters, bufs, vs, models], flags, 16384)
    _C.mt_fused_adam(t5, 1e-3, 0.9, 0.999, 1e-8, 0.0, -1e-3, 1.0, False, 1. / 128, True, f)
    for lst, kept in zip((masters, bufs, vs, models), keep):
        for a, b in zip(lst, kept):
            assert a is None or torch.equal(a, b)
    # and with a clean flag it does update
    _C.flag_clear(f)
    _C.mt_fused_sgd(t, 0.1, 0.9, 0.0, 1e-4, False, 1. / 128, True, f)
    assert not torch.equal(masters[-1], keep[0][-1])


def test_l2norm_and_clip():
    grads, flags = _mixed_grads(10)
    inv = 1. / 1024
    t = table([grads], flags, 65536)
    partials = torch.zeros(t.struct.total_chunks, device=dev())
    norm = torch.zeros(1, device=dev())
    _C.mt_l2norm(t, 0, inv, True, partials, norm)
    want = np.sqrt(sum(c_oracle.l2norm([bits(g)], inv, True) ** 2 for g in grads))
    assert abs(float(norm.item()) - want) <= 2e-6 * want
    # against torch (what clip_master_grads calls): norm of the unscaled fp32 gradients
    fp32 = [g.float() * torch.tensor(inv) for g in grads]
    ps = [torch.nn.Parameter(torch.zeros_like(x)) for x in fp32]
    for p, x in zip(ps, fp32):
        p.grad = x.clone()
    tn = torch.nn.utils.clip_grad_norm_(ps, max_norm=0.01)
    assert abs(float(tn) - float(norm.item())) <= 2e-6 * float(tn)
    # deterministic: same bits on a second run
    norm2 = torch.zeros(1, device=dev())
    _C.mt_l2norm(t, 0, inv, True, partials, norm2)
    assert torch.equal(norm, norm2)
    # clip in place: every gradient times max_norm / (norm + 1e-6)
    mine = [x.clone() for x in fp32]
    tc = table([mine], [_C.TF_GRAD_F32] * len(mine), 65536)
    n32 = torch.zeros(1, device=dev())
    _C.mt_l2norm(tc, 0, 1.0, False, partials, n32)
    _C.mt_clip_by_norm(tc, 0, n32, 0.01, None)
    for a, p in zip(mine, ps):
        assert torch.allclose(a, p.grad, rtol=3e-6, atol=0)
    # a norm below max_norm leaves the gradients alone
    keep = [x.clone() for x in mine]
    _C.mt_l2norm(tc, 0, 1.0, False, partials, n32)
    _C.mt_clip_by_norm(tc, 0, n32, 1e9, None)
    assert all(torch.equal(a, b) for a, b in zip(mine, keep))
    # too small a scratch buffer is refused
    with pytest.raises(RuntimeError, match="too small"):
        _C.mt_l2norm(t, 0, inv, True, partials[:1], norm)


def test_two_streams_two_threads():
    """Entry points are re-entrant: DDP calls them from the autograd thread on a side stream."""
    import threading
    errs = []

    def work(seed):
        try:
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                g = gen(seed)
                src = [randn16(n, g) for n in SIZES]
                dst = [torch.zeros(n, device=dev()) for n in SIZES]
                t = table([src, dst], None, 65536)
                for _ in range(20):
                    _C.mt_grads_to_master(t, 0, 1, 0.5, True, None)
                s.synchronize()
                for a, b in zip(src, dst):
                    assert torch.equal(b, a.float() * 0.5)
        except Exception as e:      # noqa: BLE001
            errs.append(e)
    th = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [x.start() for x in th]
    [x.join() for x in th]
    assert not errs, errs


@pytest.mark.parametrize("norm_value,max_norm", [(37.5, 1.0), (0.02, 1.0), (3.0, 0.25)])
def test_clip_factor_folded_into_fused_step_and_materialisation(norm_value, max_norm):
    """clip_master_grads as a factor inside the kernels: g * inv_scale * clamp(max_norm/(norm+1e-6), 1)."""
    grads, masters, models, flags = _step_inputs(12)
    norm = torch.tensor([norm_value], device=dev())
    coef = c_oracle.clip_coef(norm_value, max_norm)
    assert (coef < 1.0) == (max_norm / (norm_value + 1e-6) < 1.0)
    inv = 1. / 1024
    # materialisation
    src = [g for g, f in zip(grads, flags) if f == 0]
    dst = [torch.zeros(g.numel(), device=dev()) for g in src]
    _C.mt_grads_to_master(table([src, dst], None, 8192), 0, 1, inv, True, None, (norm, max_norm))
    for s, d in zip(src, dst):
        assert np.array_equal(bits(d), c_oracle.f16_to_f32_scaled(bits(s), inv, True, coef))
        ref = s.float().mul_(inv).mul_(torch.tensor(coef, device=dev()))       # unscale, then clip_grad_norm_'s mul_
        assert torch.equal(d, ref)
    # header contract: apply_scale == 0 means exact widening, also when a clip norm is given and the
    # caller passes some inv_scale != 1 along (inv_scale leaked into this case)
    dst0 = [torch.zeros(g.numel(), device=dev()) for g in src]
    _C.mt_grads_to_master(table([src, dst0], None, 8192), 0, 1, inv, False, None, (norm, max_norm))
    for s, d in zip(src, dst0):
        assert torch.equal(d, s.float().mul_(torch.tensor(coef, device=dev())))
    # SGD with momentum
    bufs = [torch.zeros_like(m) for m in masters]
    o_p = [bits(m).copy() for m in masters]
    o_b = [np.zeros_like(x) for x in o_p]
    t = table([grads, masters, bufs, models], None, 2048)
    for step in range(2):
        fl = [f | (_C.TF_FIRST_STEP if step == 0 else 0) for f in flags]
        t.update([grads, masters, bufs, models], fl)
        _C.mt_fused_sgd(t, 0.1, 0.9, 0.0, 1e-4, False, inv, True, None, (norm, max_norm))
        for i in range(len(grads)):
            ns = bool(flags[i] & _C.TF_NO_SCALE)
            p16 = c_oracle.sgd(bits(grads[i]), o_p[i], o_b[i], 0.1, 0.9, 0.0, 1e-4, False, first=(step == 0),
                               inv_scale=inv, apply_scale=not ns, clip_coef=None if ns else coef)
            assert np.array_equal(bits(masters[i]), o_p[i]), (step, i)
            assert np.array_equal(bits(bufs[i]), o_b[i]), (step, i)
            if models[i] is not None:
                assert np.array_equal(bits(models[i]), p16)
    # Adam
    grads, masters, models, flags = _step_inputs(13)
    ms = [torch.zeros_like(m) for m in masters]
    vs = [torch.zeros_like(m) for m in masters]
    o_p = [bits(m).copy() for m in masters]
    o_m = [np.zeros_like(x) for x in o_p]
    o_v = [np.zeros_like(x) for x in o_p]
    t5 = table([grads, masters, ms, vs, models], flags, 2048)
    bc1, bc2 = 1 - 0.9, 1 - 0.999
    _C.mt_fused_adam(t5, 1e-3, 0.9, 0.999, 1e-8, 0.0, (1e-3 / bc1) * -1, bc2 ** 0.5, False, inv, True, None,
                     (norm, max_norm))
    for i in range(len(grads)):
        ns = bool(flags[i] & _C.TF_NO_SCALE)
        c_oracle.adam(bits(grads[i]), o_p[i], o_m[i], o_v[i], 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, inv_scale=inv,
                      apply_scale=not ns, clip_coef=None if ns else coef)
        assert np.array_equal(bits(masters[i]), o_p[i]) and np.array_equal(bits(ms[i]), o_m[i])
        assert np.array_equal(bits(vs[i]), o_v[i])


@pytest.mark.parametrize("clip", [True, False])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_larc_kernels_against_the_reference_loop_on_gpu(clip, dtype):
    from apex_b200.multi_tensor import TensorTable, CHUNK_NORM
    from oracle import apex_ref
    g = gen(14)
    sizes = [1, 7, 2048, 70001, 300000, 64, 0, 33]
    params = [(torch.randn(n, generator=g, device=dev()) * 0.3).to(dtype) for n in sizes]
    params[5].zero_()                                         # |p| == 0: left alone
    grads = [(torch.randn(n, generator=g, device=dev()) * (5.0 if i % 2 else 0.02)).to(dtype)
             for i, n in enumerate(sizes)]
    grads[3].zero_()                                          # |g| == 0: left alone
    ref = [torch.nn.Parameter(p.clone()) for p in params]
    for r, gr in zip(ref, grads):
        r.grad = gr.clone()
    apex_ref.larc_rescale(ref, 0.1, 5e-4, 0.02, clip)
    t = TensorTable(dev(), 2, CHUNK_NORM)
    t.update([grads, params], [_C.TF_GRAD_F32 if dtype == torch.float32 else 0] * len(sizes))
    scratch = torch.empty(2 * t.struct.total_chunks + len(sizes), device=dev())
    norms = torch.zeros(len(sizes), 2, device=dev())
    _C.mt_larc(t, 0, 1, 0.02, 1e-8, 5e-4, 0.1, clip, scratch, norms)
    for i, (mine, r, p) in enumerate(zip(grads, ref, params)):
        if sizes[i]:
            assert float(norms[i, 0]) == pytest.approx(float(torch.norm(p.float())), rel=3e-6)
        tol = 3e-6 if dtype == torch.float32 else 2e-3         # fp16: one more rounding of the rate's effect
        assert torch.allclose(mine.float(), r.grad.float(), rtol=tol, atol=1e-8 if dtype == torch.float32 else 1e-6), i
    assert torch.equal(grads[5], ref[5].grad) and torch.equal(grads[3], ref[3].grad)
