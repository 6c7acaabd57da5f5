"""FP16_Optimizer checkpoints on the GPU (reference: apex/fp16_utils/fp16_optimizer.py:267-328, flow of
examples/FP16_Optimizer_simple/save_load.py:27-56: torch.save(state_dict()) -> torch.load(map_location=cuda)
-> load_state_dict -> keep training)."""
import contextlib
import io
import os
import sys

import pytest
import torch

from tests.conftest import GOLDEN_DIR
from tests.gpu_util import dev, gen

pytestmark = pytest.mark.gpu


def _run(opt, params, grads_per_step):
    with contextlib.redirect_stdout(io.StringIO()):
        for gs in grads_per_step:
            for p, g in zip(params, gs):
                p.grad = g.clone()
            opt.update_master_grads()
            opt.step()


@pytest.mark.parametrize("opt_name", ["SGD", "Adam"])
def test_save_load_on_gpu_continues_bit_exactly(tmp_path, opt_name):
    from apex_b200.fp16_utils import FP16_Optimizer
    make = {"SGD": lambda ps: torch.optim.SGD(ps, lr=0.1, momentum=0.9, weight_decay=1e-4),
            "Adam": lambda ps: torch.optim.Adam(ps, lr=1e-3)}[opt_name]
    g = gen(77)
    shapes16, shapes32 = [(513, 129), (4099,), (64, 3, 7, 7)], [(64,), (2048,)]
    init = [(torch.randn(s, generator=g, device=dev()) * 0.02).half() for s in shapes16] + \
           [torch.randn(s, generator=g, device=dev()) * 0.02 + 1.0 for s in shapes32]
    kw = dict(dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 2.0**14, "scale_window": 3}, verbose=False)

    def grads(n, poison=None):
        out = []
        for it in range(n):
            gs = [(torch.randn(t.shape, generator=g, device=dev()) * 8.0).to(t.dtype) for t in init]
            if poison == it:
                gs[1][7] = float("inf")
            out.append(gs)
        return out
    first, second = grads(4, poison=2), grads(5, poison=1)

    # uninterrupted run
    straight = [torch.nn.Parameter(t.clone()) for t in init]
    opt_a = FP16_Optimizer(make(straight), **kw)
    _run(opt_a, straight, first + second)

    # run, save, restore into fresh objects on the GPU, continue
    before = [torch.nn.Parameter(t.clone()) for t in init]
    opt_b = FP16_Optimizer(make(before), **kw)
    _run(opt_b, before, first)
    path = os.path.join(str(tmp_path), "ckpt.pt")
    torch.save({"model": [p.data for p in before], "optimizer": opt_b.state_dict()}, path)
    ckpt = torch.load(path, map_location=lambda storage, loc: storage.cuda(0), weights_only=False)
    resumed = [torch.nn.Parameter(t.clone()) for t in ckpt["model"]]
    opt_c = FP16_Optimizer(make(resumed), **kw)
    opt_c.load_state_dict(ckpt["optimizer"])
    assert all(m.is_cuda for m in opt_c.fp32_from_fp16_groups[0])
    assert float(opt_c.loss_scale) == float(opt_b.loss_scale) and opt_c.loss_scaler.cur_iter == opt_b.loss_scaler.cur_iter
    _run(opt_c, resumed, second)

    assert opt_c.unfused_steps == 0                      # the resumed run stays on the fused kernels
    assert float(opt_c.loss_scale) == float(opt_a.loss_scale)
    assert opt_c.loss_scaler.cur_iter == opt_a.loss_scaler.cur_iter
    assert opt_c.loss_scaler.last_overflow_iter == opt_a.loss_scaler.last_overflow_iter
    for a, c in zip(opt_a.fp32_from_fp16_groups[0], opt_c.fp32_from_fp16_groups[0]):
        assert torch.equal(a.data, c.data)
    for a, c in zip(straight, resumed):
        assert torch.equal(a.data, c.data)
    for pa, pc in zip(opt_a.optimizer.param_groups[0]["params"], opt_c.optimizer.param_groups[0]["params"]):
        sa, sc = opt_a.optimizer.state[pa], opt_c.optimizer.state[pc]
        assert sorted(sa) == sorted(sc)
        for key in sa:
            assert torch.equal(torch.as_tensor(sa[key]).cpu(), torch.as_tensor(sc[key]).cpu()), key
    # the checkpoint holds plain tensors under the reference's keys
    assert sorted(ckpt["optimizer"]) == sorted(["loss_scaler", "dynamic_loss_scale", "overflow",
                                                "first_closure_call_this_step", "optimizer_state_dict",
                                                "fp32_from_fp16"])
    assert all(type(t) is torch.Tensor for t in ckpt["optimizer"]["fp32_from_fp16"][0])


def test_checkpoint_written_by_the_reference_loads_onto_the_gpu():
    """tests/golden/reference_checkpoint.pt was written by the UNMODIFIED reference (its pickled scaler is
    apex.fp16_utils.loss_scaler.DynamicLossScaler); loaded with map_location -> cuda as save_load.py:33-37
    does, training continues to exactly the state the reference itself reached."""
    import apex_b200
    from apex_b200.fp16_utils import FP16_Optimizer, DynamicLossScaler
    apex_b200.install_as_apex()
    try:
        to_gpu = lambda storage, loc: storage.cuda(0)      # noqa: E731
        ckpt = torch.load(os.path.join(GOLDEN_DIR, "reference_checkpoint.pt"), map_location=to_gpu, weights_only=False)
        after = torch.load(os.path.join(GOLDEN_DIR, "reference_checkpoint_after.pt"), map_location=to_gpu,
                           weights_only=False)
    finally:
        for name in [k for k in sys.modules if k == "apex" or k.startswith("apex.")]:
            del sys.modules[name]
    assert isinstance(ckpt["optimizer"]["loss_scaler"], DynamicLossScaler)
    params = [torch.nn.Parameter(p.clone()) for p in ckpt["model"]]
    assert all(p.is_cuda for p in params)
    opt = FP16_Optimizer(torch.optim.SGD(params, lr=0.1, momentum=0.9), dynamic_loss_scale=True, verbose=False)
    opt.load_state_dict(ckpt["optimizer"])
    _run(opt, params, after["grads"])
    assert float(opt.loss_scale) == after["final_scale"] and opt.loss_scaler.cur_iter == after["final_iter"]
    for m, want in zip(opt.fp32_from_fp16_groups[0], after["final_masters"]):
        assert torch.equal(m.data, want)
    for p, want in zip(params, after["final_model"]):
        assert torch.equal(p.data, want)
