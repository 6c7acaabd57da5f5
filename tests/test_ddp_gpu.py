"""apex_b200.parallel on the GPU.  One-GPU cases always run; the 2-rank NCCL cases (including
the reference's own known-answer test) run when the box has >= 2 GPUs."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.conftest import ROOT

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.fixture
def nccl_single():
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(_free_port())
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", rank=0, world_size=1)
    yield
    dist.destroy_process_group()


def test_single_rank_ddp_keeps_gradients_and_learns_layout(nccl_single, golden_json):
    from apex_b200.parallel import DistributedDataParallel as DDP
    from apex_b200 import _C
    from tests.golden_cases import build_layout_model
    for case in golden_json["ddp_layout"]:
        if case["model"] != "mlp3":
            continue
        model = build_layout_model("mlp3").cuda()
        ref = build_layout_model("mlp3").cuda()
        ddp = DDP(model, message_size=case["message_size"])
        x = torch.randn(5, 6, device="cuda")
        for it in range(3):
            model.zero_grad()
            ref.zero_grad()
            before = _C.launches
            ddp(x).sum().backward()
            ref(x).sum().backward()
            torch.cuda.synchronize()
            assert _C.launches > before
            for p, q in zip(model.parameters(), ref.parameters()):
                assert torch.equal(p.grad, q.grad)              # world size 1: x (1/1) is exact
        plist = [p for p in model.parameters()]
        assert [list(ddp.param_id_to_bucket[id(p)]) for p in plist] == case["param_to_bucket"]
        assert list(ddp.bucket_sizes) == case["bucket_sizes"]


def test_half_model_with_fp16_optimizer_single_rank(nccl_single):
    from apex_b200.parallel import DistributedDataParallel as DDP
    from apex_b200.fp16_utils import FP16_Optimizer
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.ReLU(), torch.nn.Linear(128, 16)).cuda().half()
    ddp = DDP(model, message_size=4096)
    opt = FP16_Optimizer(torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9), dynamic_loss_scale=True,
                         dynamic_loss_args={"init_scale": 2**10}, verbose=False)
    x = torch.randn(32, 64, device="cuda").half()
    y = torch.randn(32, 16, device="cuda").half()
    losses = []
    for it in range(12):
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(ddp(x).float(), y.float())
        opt.backward(loss)
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world)
    from apex_b200.parallel import DistributedDataParallel as DDP, symm as _symm
    from apex_b200.parallel.distributed import flat_dist_call
    ok = {}
    ok["symm_default_on"] = _symm.enabled()

    # the reference's known-answer test, verbatim sizes (tests/distributed/ddp_race_condition_test.py)
    class Model(torch.nn.Module):
        def __init__(self):
            super(Model, self).__init__()
            self.a = torch.nn.Parameter(torch.full((4096 * 4096,), 1.0, device="cuda"))
            self.b = torch.nn.Parameter(torch.full((4096 * 4096,), 2.0, device="cuda"))

        def forward(self, input):
            return (input * self.a) * self.b
    model = DDP(Model(), message_size=1)
    x = torch.empty(4096 * 4096, device="cuda")
    passed = True
    for i in range(10):
        x.fill_(i + rank)
        model.zero_grad()
        model(x).sum().backward()
        for param, val in ((model.module.a, 2.), (model.module.b, 1.)):
            # mean over ranks of val * n * (i + rank); for world = 2 this is the reference's formula
            expected = val * 4096 * 4096 * (2. * i + (world - 1)) / 2.
            passed &= (param.grad.data.sum().item() == expected)
    ok["race"] = bool(passed)

    # fp16 buckets: averaged gradients within rtol 1e-3 of the fp32 mean (north_star), sum-then-divide
    g = torch.Generator(device="cuda").manual_seed(5 + rank)
    tensors = [(torch.randn(n, generator=g, device="cuda") * 100).half() for n in (5, 4099, 1 << 20, 33)]
    exact = [t.float() for t in tensors]
    for e in exact:
        dist.all_reduce(e)
    # what the reference itself computes: cat, NCCL all_reduce in fp16, /= world (distributed.py:12-20)
    ref = torch.cat([t.view(-1) for t in tensors])
    dist.all_reduce(ref)
    ref /= world
    flat_dist_call(tensors, dist.all_reduce)
    torch.cuda.synchronize()
    mine = torch.cat([t.view(-1) for t in tensors])
    close = torch.allclose(mine.float(), ref.float(), rtol=1e-3, atol=1e-3)     # north_star: rtol 1e-3 for fp16 sums
    ok["fp16_same_bits_as_reference_composition"] = bool(torch.equal(mine, ref))
    # and not far from the true mean (fp16 partial sums round at every hop of NCCL's ring)
    close &= all(torch.allclose(t.float(), e / world, rtol=5e-3, atol=0.3) for t, e in zip(tensors, exact))
    ok["fp16_average"] = bool(close)
    big = torch.full((1000,), 40000.0, device="cuda").half()
    flat_dist_call([big, big.clone()], dist.all_reduce)
    ok["fp16_overflow_stays_inf"] = bool(torch.isinf(big).all())
    # a real model through DistributedDataParallel: learning iteration (flat call per dtype), then
    # overlapped iterations whose multi-tensor buckets go through the symmetric-memory pipeline;
    # gradients must equal the mean of the ranks' local gradients, with fresh .grad tensors every time
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 256), torch.nn.ReLU(),
                              torch.nn.Linear(256, 10)).cuda()
    twin = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 256), torch.nn.ReLU(),
                               torch.nn.Linear(256, 10)).cuda()
    twin.load_state_dict(net.state_dict())
    if rank != 0:
        for p in net.parameters():
            p.data.add_(1.0)                   # construction must broadcast rank 0's values
    ddp = DDP(net, message_size=20000)
    good = all(torch.equal(p.data, q.data) for p, q in zip(net.parameters(), twin.parameters()))
    for it in range(5):
        gg = torch.Generator(device="cuda").manual_seed(1000 * it + rank)
        xb = torch.randn(32, 64, generator=gg, device="cuda")
        net.zero_grad(set_to_none=True)
        twin.zero_grad(set_to_none=True)
        ddp(xb).pow(2).sum().backward()
        twin(xb).pow(2).sum().backward()
        torch.cuda.synchronize()
        for p, q in zip(net.parameters(), twin.parameters()):
            want = q.grad.clone()
            dist.all_reduce(want)
            want /= world
            good &= bool(torch.allclose(p.grad, want, rtol=1e-5, atol=1e-6))
    ok["ddp_real_model"] = bool(good)
    # the overlapped iterations tagged the gradients: an FP16_Optimizer would not need its own scan
    from apex_b200.parallel.distributed import verified_flag
    ok["ddp_grads_verified"] = verified_flag([p.grad for p in net.parameters()]) is ddp._pipeline.check_flag
    ok["ddp_real_model_buckets"] = ddp.num_buckets
    ok["ddp_used_symm"] = any(arena is not None for _, arena, _ in ddp._pipeline.slots.values())
    # after the overlapped iterations every gradient lives in its bucket (no gather, no scatter)
    ok["ddp_grads_rehomed"] = all(p.grad is v for b, ps in enumerate(ddp._bucket_params)
                                  for p, v in zip(ps, ddp._pipeline.slots[b][0].views))
    # in-place iterations (zero_grad keeps the tensors, autograd accumulates into the bucket)
    from apex_b200 import _C as _c
    good = True
    for it in range(3):
        gg = torch.Generator(device="cuda").manual_seed(7000 * it + rank)
        xb = torch.randn(32, 64, generator=gg, device="cuda")
        net.zero_grad(set_to_none=False)
        twin.zero_grad(set_to_none=True)
        ddp(xb).pow(2).sum().backward()
        twin(xb).pow(2).sum().backward()
        torch.cuda.synchronize()
        for p, q in zip(net.parameters(), twin.parameters()):
            want = q.grad.clone()
            dist.all_reduce(want)
            want /= world
            good &= bool(torch.allclose(p.grad, want, rtol=1e-5, atol=1e-6))
    ok["ddp_in_place"] = bool(good)

    # Reducer on the GPU (reference: distributed.py:52-93): tensor list and module
    from apex_b200.parallel import Reducer
    t = [torch.full((3,), float(rank), device="cuda"), [torch.full((2,), 2.0 * rank, device="cuda").half()]]
    Reducer(t).reduce()
    mean_rank = (world - 1) / 2.0
    ok["reducer_list"] = bool((t[0] == mean_rank).all() and (t[1][0] == 2.0 * mean_rank).all())
    m2 = torch.nn.Linear(3, 2).cuda()
    red2 = Reducer(m2)
    w0 = m2.weight.data.clone()
    dist.broadcast(w0, 0)
    ok["reducer_broadcast"] = bool(torch.equal(w0, m2.weight.data))
    m2(torch.full((1, 3), float(rank + 1), device="cuda")).sum().backward()
    red2.reduce()
    torch.cuda.synchronize()
    ok["reducer_module"] = bool(torch.equal(m2.weight.grad, torch.full((2, 3), mean_rank + 1.0, device="cuda")))

    # our own allreduce over symmetric memory (switch multicast if the box has it, and peer-to-peer),
    # and the NCCL fallback (APEX_B200_SYMM=0) through the same pipeline: same answers, same flag
    from apex_b200.parallel import symm
    from apex_b200.parallel.distributed import _BucketPipeline
    for mode, on, force_p2p in (("symm_auto", True, False), ("symm_p2p", True, True), ("nccl_fallback", False, False)):
        symm.enable(on)
        symm._force_p2p = force_p2p
        symm._force_mc = on and not force_p2p
        pipe = _BucketPipeline(torch.cuda.current_stream())
        g = torch.Generator(device="cuda").manual_seed(50 + rank)
        t16 = [(torch.randn(n, generator=g, device="cuda") * 100).half() for n in (5, 4099, 1 << 20, 33, 70001)]
        t32 = [torch.randn(n, generator=g, device="cuda") * 100 for n in (7, 65536, 12345)]
        e16 = [t.float() for t in t16]
        e32 = [t.clone() for t in t32]
        for e in e16 + e32:
            dist.all_reduce(e)
        # the reference composition on the same inputs: cat, NCCL all_reduce in fp16, /= world
        ref16 = torch.cat([t.view(-1) for t in t16])
        dist.all_reduce(ref16)
        ref16 /= world
        for it in range(3):           # the flat homes are reused: run it more than once
            w16 = [t.clone() for t in t16]
            w32 = [t.clone() for t in t32]
            pipe.begin_round()
            pipe.submit(0, w16)
            pipe.submit(1, w32)
            pipe.join(torch.cuda.current_stream())
            torch.cuda.synchronize()
            # against the exact mean: our kernels round the fp32 sum once; NCCL (the fallback, and the
            # reference's own transport) rounds partial sums to fp16 at every hop of its ring
            r16, a16 = (1e-3, 1e-2) if on else (5e-3, 0.3)
            good = all(torch.allclose(t.float(), e / world, rtol=r16, atol=a16) for t, e in zip(w16, e16))
            good &= all(torch.allclose(t, e / world, rtol=1e-5, atol=1e-4) for t, e in zip(w32, e32))
            got16 = torch.cat([t.view(-1) for t in w16])
            # north_star: rtol 1e-3 for fp16 sums.  NCCL rounds partial sums to fp16 at every hop of a
            # ring/tree (small buffers), we round once: near-zero means have no relative scale, so the
            # absolute slack is 1e-3 of the largest addend
            slack = 1e-3 * max(float(t.abs().max()) for t in t16)
            good &= bool(torch.allclose(got16.float(), ref16.float(), rtol=1e-3, atol=slack))
            ok[mode + "_max_abs_diff_to_reference"] = float((got16.float() - ref16.float()).abs().max())
            ok[mode + "_bits_equal_reference_fraction"] = float((got16 == ref16).float().mean())
            # every rank must hold identical bits
            mine = torch.cat([t.view(-1).float() for t in w16 + w32])
            other = mine.clone()
            dist.broadcast(other, 0)
            good &= bool(torch.equal(mine, other))
            ok[mode] = ok.get(mode, True) and bool(good) and int(pipe.check_flag.item()) == 0
        # a sum that overflows fp16 stays inf after the divide and raises the word on EVERY rank,
        # also when only one rank's slice holds it
        big = [torch.full((1000,), 40000.0, device="cuda").half(), torch.full((8,), 1.0, device="cuda").half()]
        pipe.begin_round()
        pipe.submit(2, big)
        pipe.join(torch.cuda.current_stream())
        torch.cuda.synchronize()
        ok[mode + "_overflow"] = bool(torch.isinf(big[0]).all()) and bool((big[1] == 1.0).all()) \
            and int(pipe.check_flag.item()) != 0
        one = [torch.zeros(4096 * world, device="cuda").half()]
        if rank == world - 1:
            one[0][-1] = float("nan")                      # lands in the LAST rank's slice
        pipe.begin_round()
        pipe.submit(3, one)
        pipe.join(torch.cuda.current_stream())
        torch.cuda.synchronize()
        ok[mode + "_overflow"] &= int(pipe.check_flag.item()) != 0 and bool(torch.isnan(one[0][-1]))
        # and the next round starts clean again
        pipe.begin_round()
        pipe.submit(3, [torch.ones(4096 * world, device="cuda").half()])
        pipe.join(torch.cuda.current_stream())
        torch.cuda.synchronize()
        ok[mode + "_overflow"] &= int(pipe.check_flag.item()) == 0
        ok[mode + "_multicast"] = bool(pipe.arenas and pipe.arenas[0].use_multicast())
        ok[mode + "_in_symmetric_memory"] = bool(pipe.arenas)
    symm.enable(True)
    symm._force_p2p = symm._force_mc = False
    # buckets that end up in DIFFERENT symmetric allocations (homes created one by one, the second
    # does not fit the first arena): each arena has its own overflow word, join() folds them into
    # the round's word
    pipe = _BucketPipeline(torch.cuda.current_stream())
    pipe.MIN_ARENA = 1 << 20
    first = [torch.ones(300000, device="cuda").half()]
    second = [torch.ones(1500000, device="cuda").half()]
    if rank == 0:
        second[0][777] = float("inf")
    pipe.begin_round()
    pipe.submit(0, first)
    pipe.submit(1, second)
    pipe.join(torch.cuda.current_stream())
    torch.cuda.synchronize()
    ok["two_arenas"] = len(pipe.arenas) == 2 and int(pipe.check_flag.item()) != 0 and bool((first[0] == 1).all()) \
        and bool(torch.isinf(second[0][777]))
    pipe.begin_round()
    pipe.submit(0, [torch.ones(300000, device="cuda").half()])
    pipe.submit(1, [torch.ones(1500000, device="cuda").half()])
    pipe.join(torch.cuda.current_stream())
    torch.cuda.synchronize()
    ok["two_arenas"] &= int(pipe.check_flag.item()) == 0
    torch.save(ok, os.path.join(out_dir, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


# every GPU of the box takes part (the driver's box has one: skipped there; logs of 2-, 4- and 8-GPU
# runs are kept under profiles/); APEX_B200_TEST_WORLD overrides
WORLD = int(os.environ.get("APEX_B200_TEST_WORLD", str(min(torch.cuda.device_count(), 8))))


@pytest.mark.skipif(torch.cuda.device_count() < max(2, WORLD), reason="needs %d GPUs" % max(2, WORLD))
def test_two_ranks_nccl(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(WORLD, port, str(tmp_path)), nprocs=WORLD, join=True)
    for r in range(WORLD):
        res = torch.load(os.path.join(str(tmp_path), "rank%d.pt" % r), weights_only=False)
        print("rank", r, res)
        assert res["ddp_real_model_buckets"] >= 2
        for key in ("race", "fp16_average", "fp16_overflow_stays_inf", "ddp_real_model", "ddp_used_symm",
                    "ddp_grads_verified", "ddp_grads_rehomed", "ddp_in_place",
                    "reducer_list", "reducer_broadcast", "reducer_module",
                    "symm_auto", "symm_p2p", "nccl_fallback",
                    "symm_auto_overflow", "symm_p2p_overflow", "nccl_fallback_overflow",
                    "symm_auto_in_symmetric_memory", "symm_p2p_in_symmetric_memory", "two_arenas"):
            assert res[key] is True, (key, res)
        assert res["nccl_fallback_in_symmetric_memory"] is False
        print("world", WORLD, "rank", r, "| multicast used:", res["symm_auto_multicast"],
              "| fraction of fp16 elements bit-equal to the reference composition (cat, NCCL all_reduce, /=):",
              "multicast %.6f" % res["symm_auto_bits_equal_reference_fraction"],
              "p2p %.6f" % res["symm_p2p_bits_equal_reference_fraction"],
              "nccl path %.6f" % res["nccl_fallback_bits_equal_reference_fraction"],
              "| symm default:", res["symm_default_on"])


def test_ddp_plus_fp16_optimizer_skips_its_own_scan_single_rank(nccl_single):
    """DDP (overlapped iterations) + FP16_Optimizer with dynamic scaling: from the second iteration on
    the step consists of the fused update only -- the check was done by DDP's scatter pass -- and an
    overflow produced by backward is still caught."""
    from apex_b200 import _C
    from apex_b200.parallel import DistributedDataParallel as DDP
    from apex_b200.fp16_utils import FP16_Optimizer
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.ReLU(), torch.nn.Linear(128, 16)).cuda().half()
    ddp = DDP(model, message_size=2000)
    opt = FP16_Optimizer(torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9), dynamic_loss_scale=True,
                         dynamic_loss_args={"init_scale": 2.0**8}, verbose=False)
    x = torch.randn(32, 64, device="cuda").half()
    y = torch.randn(32, 16, device="cuda")
    real_check = _C.mt_check_overflow_publish
    scans = []
    _C.mt_check_overflow_publish = lambda *a, **k: (scans.append(1), real_check(*a, **k))[1]
    try:
        for it in range(6):
            opt.zero_grad()
            scale_up = 1e6 if it == 4 else 1.0               # iteration 4 overflows in backward
            loss = torch.nn.functional.mse_loss(ddp(x).float() * scale_up, y)
            import contextlib, io
            with contextlib.redirect_stdout(io.StringIO()):
                opt.backward(loss)
                opt.step()
            assert opt.overflow == (it == 4), it
    finally:
        _C.mt_check_overflow_publish = real_check
    # zero_grad() zeroes the gradients in place through our own kernel (no version change), autograd then
    # accumulates into them (version change) BEFORE DDP averages and tags them -> still covered.
    assert len(scans) <= 2, scans        # the learning iteration (and at most one more) scanned; the rest did not
