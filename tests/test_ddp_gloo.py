"""apex_b200.parallel on CPU under gloo: bucket structure, ordering, averaging, N > 1.

Host-side logic only (the kernel wrappers are the oracle-backed double).  The
world_size-2 cases run as two spawned processes rendezvousing on 127.0.0.1.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.conftest import ROOT
from tests.golden_cases import build_layout_model


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.fixture
def gloo_single(cpu_backend):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(_free_port())
    dist.init_process_group("gloo", rank=0, world_size=1)
    yield cpu_backend
    dist.destroy_process_group()


def test_bucket_layout_matches_reference(golden_json, gloo_single):
    from apex_b200.parallel import DistributedDataParallel as DDP
    for case in golden_json["ddp_layout"]:
        model = build_layout_model(case["model"])
        ddp = DDP(model, message_size=case["message_size"])
        x = torch.randn(5, 6)
        for it in range(3):
            model.zero_grad()
            y = ddp(x)
            y.float().sum().backward()
            assert ddp.needs_refresh is False
        plist = [p for p in model.parameters() if p.requires_grad]
        where = [list(ddp.param_id_to_bucket[id(p)]) for p in plist]
        assert ddp.num_buckets == case["num_buckets"], case
        assert list(ddp.bucket_sizes) == case["bucket_sizes"], case
        assert where == case["param_to_bucket"], case
        # world size 1: averaging is the identity, gradients must be untouched and finite
        for p in plist:
            assert torch.isfinite(p.grad.float()).all()


def test_guards_and_options(gloo_single):
    from apex_b200.parallel import DistributedDataParallel as DDP
    from apex_b200.parallel.distributed import _BucketRound
    model = build_layout_model("mlp3")
    with pytest.raises(ValueError, match="shared_param is no longer supported"):
        DDP(model, shared_param=True)
    ddp = DDP(model)
    assert ddp.message_size == 10000000 and ddp.delay_allreduce is False
    assert ddp.module is model and ddp.reduction_stream is not None
    r = _BucketRound([2, 1])
    g = torch.zeros(1)
    assert r.put(1, 0, g) == []              # full but not next: parked
    assert r.put(0, 0, g) == []
    with pytest.raises(RuntimeError, match="already-filled"):
        r.put(0, 0, g)
    with pytest.raises(RuntimeError, match="next_bucket != num_buckets"):
        r.verify_complete()
    assert r.put(0, 1, g) == [0, 1]          # 0 goes, then parked 1 drains, in index order
    r.verify_complete()
    r2 = _BucketRound([1, 2])
    r2.put(0, 0, g)
    r2.next_bucket = 2
    r2.filled[1] = 1
    with pytest.raises(RuntimeError, match="Some param buckets were not allreduced"):
        r2.verify_complete()


def test_delay_allreduce_and_refresh_on_param_change(gloo_single):
    from apex_b200.parallel import DistributedDataParallel as DDP
    model = build_layout_model("mlp3")
    ddp = DDP(model, delay_allreduce=True)
    ddp(torch.randn(2, 6)).sum().backward()
    assert "flatten" in gloo_single.calls and ddp.num_buckets == 0      # no structure is ever learned
    ddp2 = DDP(build_layout_model("mlp3"), message_size=50)
    for _ in range(2):
        ddp2(torch.randn(2, 6)).sum().backward()
    assert ddp2.needs_refresh is False and ddp2.num_buckets == 3
    # freezing a parameter changes the active list -> structure is learned again
    ddp2.module[0].bias.requires_grad = False
    ddp2.module.zero_grad()
    ddp2(torch.randn(2, 6)).sum().backward()
    assert ddp2.needs_refresh is False
    assert sum(ddp2.bucket_sizes) == 5


def test_gradients_move_into_their_buckets(gloo_single):
    """After the first overlapped backward every param.grad is a view of its bucket's flat home
    (SURVEY 8f N2): later iterations neither gather nor scatter, autograd accumulates in place; a
    gradient the user dropped (set_to_none) is copied in once and becomes a view again."""
    from apex_b200.parallel import DistributedDataParallel as DDP
    from apex_b200.parallel.distributed import verified_flag
    calls = gloo_single.calls
    torch.manual_seed(1)
    net = build_layout_model("mlp3")
    ref = build_layout_model("mlp3")
    ref.load_state_dict(net.state_dict())
    ddp = DDP(net, message_size=50)

    def run(x, set_to_none):
        net.zero_grad(set_to_none=set_to_none)
        ref.zero_grad(set_to_none=True)
        del calls[:]
        ddp(x).pow(2).sum().backward()
        ref(x).pow(2).sum().backward()
        for p, q in zip(net.parameters(), ref.parameters()):
            assert torch.equal(p.grad, q.grad)                       # world 1: x (1/1) is exact

    def homes():
        out = []
        for b, params in enumerate(ddp._bucket_params):
            home = ddp._pipeline.slots[b][0]
            for p, v in zip(params, home.views):
                out.append(p.grad is v)
                lo = home.flat.data_ptr()
                assert lo <= v.data_ptr() < lo + home.flat.numel() * home.flat.element_size()
                assert v.data_ptr() % 32 == 0
        return out

    x = torch.randn(5, 6)
    run(x, True)                                                     # learns the structure (flat call)
    assert "flatten" in calls and ddp.num_buckets == 3
    run(x, True)                                                     # first overlapped: strays are re-homed
    assert calls.count("copy") == 3 and "flatten" not in calls and "scale_unflatten" not in calls
    assert all(homes())
    run(x, False)                                                    # zeroed in place: nothing to move
    assert "copy" not in calls and "flatten" not in calls and calls.count("scale_flat") == 3
    assert all(homes())
    assert verified_flag([p.grad for p in net.parameters()]) is ddp._pipeline.check_flag
    run(x, True)                                                     # dropped gradients come back as views
    assert calls.count("copy") == 3 and all(homes())
    # one gradient replaced by hand: only that bucket copies
    first = next(net.parameters())
    net.zero_grad(set_to_none=False)
    first.grad = torch.zeros_like(first)
    ref.zero_grad(set_to_none=True)
    del calls[:]
    ddp(x).pow(2).sum().backward()
    ref(x).pow(2).sum().backward()
    assert calls.count("copy") == 1 and all(homes())
    assert all(torch.equal(p.grad, q.grad) for p, q in zip(net.parameters(), ref.parameters()))


def test_symmetric_path_host_logic_with_a_fake_arena(gloo_single, monkeypatch):
    """The host side of the symmetric-memory bucket path without NVLink: homes carved out of ONE
    allocation by plan(), gradients re-homed, one allreduce call per bucket with per-channel sequence
    numbers that only depend on the bucket's position in the round, the round's overflow word living in
    the allocation.  (The kernel itself is covered by the multi-GPU tests.)"""
    from apex_b200.parallel import symm
    from apex_b200.parallel import distributed as D
    calls = []

    class FakeArena(object):
        def __init__(self, payload, device):
            self.nbytes = symm.HEADER_BYTES + int(payload)
            self.buf = torch.zeros(self.nbytes, dtype=torch.uint8)
            self.used = symm.HEADER_BYTES
            self.seq = [0, 0, 0, 0]
            self.flags = self.buf[symm.FLAG_BASE:symm.FLAG_BASE + 256].view(torch.int32)
            self.n_flags = 0

        def take(self, nbytes):
            start = (self.used + 255) // 256 * 256
            if start + nbytes > self.nbytes:
                return None
            self.used = start + nbytes
            return start

        def new_flag(self):
            self.n_flags += 1
            return self.flags[self.n_flags - 1:self.n_flags], symm.FLAG_BASE + 4 * (self.n_flags - 1)

        def use_multicast(self):
            return True

        def rotation(self):
            return 4, 32

        def allreduce(self, offset, numel, dtype, divisor, flag_offset, channel=0, apply_mul=True, ctas=None):
            self.seq[channel] += 1
            calls.append((offset, numel, channel, self.seq[channel], flag_offset, ctas))
            flat = self.buf[offset:offset + numel * dtype.itemsize].view(dtype)
            if not torch.isfinite(flat.float()).all():                    # world 1: average = identity
                self.buf[flag_offset:flag_offset + 4].view(torch.int32)[0] = 0x7fffffff
    monkeypatch.setattr(symm, "enabled", lambda: True)
    monkeypatch.setattr(symm, "Arena", FakeArena)
    monkeypatch.setattr(D._BucketPipeline, "_wants_symm",
                        staticmethod(lambda like: like[0].dtype in (torch.float16, torch.float32)))
    pipe = D._BucketPipeline(gloo_single._Stream())
    params = [torch.nn.Parameter(torch.zeros(n).half()) for n in (40, 7, 3000, 64, 5)]
    buckets = [params[:2], params[2:4], params[4:]]
    pipe.plan([(k, [p.data for p in b]) for k, b in enumerate(buckets)])
    assert len(pipe.arenas) == 1                                          # one allocation for all buckets
    offsets = [pipe.slots[k][2] for k in range(3)]
    assert offsets == sorted(offsets) and offsets[0] >= symm.HEADER_BYTES and all(o % 256 == 0 for o in offsets)
    for rnd in range(3):
        for p in params:
            if rnd != 1:                                                  # round 1: gradients kept (in place)
                p.grad = torch.full_like(p, float(rnd + 1))
        if rnd == 2:
            params[2].grad[5] = float("nan")
        del calls[:], gloo_single.calls[:]
        pipe.begin_round()
        for k, b in enumerate(buckets):
            pipe.submit(k, [p.grad for p in b], verify=True, owners=b)
        pipe.join(gloo_single._Stream())
        assert [c[2] for c in calls] == [0, 1, 2]                         # channel = position in the round
        assert [c[3] for c in calls] == [rnd + 1] * 3                     # one more per channel and round
        assert [c[0] for c in calls] == offsets and all(c[5] == 32 for c in calls)
        assert gloo_single.calls.count("copy") == (0 if rnd == 1 else 3)  # strays only
        for k, b in enumerate(buckets):
            assert all(p.grad is v for p, v in zip(b, pipe.slots[k][0].views))
        assert int(pipe.check_flag[0] != 0) == int(rnd == 2)
        assert pipe.check_flag.data_ptr() == pipe.arenas[0].flags.data_ptr()
        assert D.verified_flag([p.grad for p in params]) is pipe.check_flag


# ---- world size 2 -------------------------------------------------------------------------------

def _install_double():
    """fake_backend.install without pytest's monkeypatch (child process)."""
    import contextlib as _ctx
    from apex_b200 import _C
    from tests import fake_backend as fb
    from apex_b200.multi_tensor import TensorTable
    TensorTable.update = fb._recording_update
    for name in ("flag_clear", "mt_check_overflow", "mt_check_overflow_publish", "mt_unscale", "mt_grads_to_master", "mt_master_to_model",
                 "mt_copy", "mt_zero", "mt_fused_sgd", "mt_fused_adam", "mt_flatten", "mt_scale_unflatten",
                 "scale_flat", "mt_l2norm", "mt_clip_by_norm", "mt_larc"):
        setattr(_C, name, getattr(fb, name))
    torch.cuda.Event = fb._Event
    torch.cuda.Stream = fb._Stream
    torch.cuda.current_stream = lambda *a, **k: fb._Stream()
    torch.cuda.stream = lambda s: _ctx.nullcontext()
    torch.Tensor.type = fb._cuda_named_type
    return fb


class _RaceModel(torch.nn.Module):
    """The model of the reference's own known-answer test
    (tests/distributed/ddp_race_condition_test.py:28-34), smaller."""

    def __init__(self, n):
        super(_RaceModel, self).__init__()
        self.a = torch.nn.Parameter(torch.full((n,), 1.0))
        self.b = torch.nn.Parameter(torch.full((n,), 2.0))

    def forward(self, input):
        return (input * self.a) * self.b


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    fb = _install_double()
    from apex_b200.parallel import DistributedDataParallel as DDP, Reducer
    from apex_b200.parallel.distributed import flat_dist_call
    ok = {}

    # (7) of SURVEY 8c: flat_dist_call, fp16 [1,1,1] + [2,2,2] -> [1.5]*3 ; fp32 10 / 20 -> 15
    h = torch.full((3,), float(rank + 1)).half()
    f = torch.full((4,), 10.0 * (rank + 1))
    h2 = torch.full((5,), float(rank + 1)).half()
    flat_dist_call([h, f, h2], dist.all_reduce)
    ok["flat_half"] = bool((h == 1.5).all() and (h2 == 1.5).all() and h.dtype == torch.float16)
    ok["flat_float"] = bool((f == 15.0).all())

    # fp16 sum that overflows must stay inf after the divide (sum first, divide after)
    big = torch.full((4,), 40000.0).half()
    flat_dist_call([big, big.clone()], dist.all_reduce)
    ok["fp16_sum_overflow_stays_inf"] = bool(torch.isinf(big).all())

    # the reference's known answer, message_size=1 so every parameter is its own bucket
    n = 64
    model = _RaceModel(n)
    ddp = DDP(model, message_size=1)
    x = torch.empty(n)
    race = True
    for i in range(6):
        x.fill_(i + rank)
        model.zero_grad()
        ddp(x).sum().backward()
        race &= model.a.grad.sum().item() == 2. * n * (2. * i + 1) / 2.
        race &= model.b.grad.sum().item() == 1. * n * (2. * i + 1) / 2.
    ok["race_known_answer"] = bool(race)
    ok["race_buckets"] = ddp.num_buckets == 2

    # a real model with rank-dependent data: averaged gradients == mean of the local gradients,
    # in the learning iteration (one flat call per dtype) and in overlapped iterations alike
    torch.manual_seed(0)
    net = build_layout_model("mlp3")
    ref = build_layout_model("mlp3")
    if rank == 1:                      # rank 0's values must win at construction
        for p in net.parameters():
            p.data.add_(1.0)
    ddp = DDP(net, message_size=50)
    ok["broadcast_from_rank0"] = all(torch.equal(p.data, q.data) for p, q in zip(net.parameters(), ref.parameters()))
    same = True
    for it in range(4):
        g = torch.Generator().manual_seed(100 * it + rank)
        xb = torch.randn(7, 6, generator=g)
        net.zero_grad()
        ref.zero_grad()
        ddp(xb).pow(2).sum().backward()
        ref(xb).pow(2).sum().backward()
        for p, q in zip(net.parameters(), ref.parameters()):
            local = q.grad.clone()
            dist.all_reduce(local)
            want = local * torch.tensor(0.5)
            same &= torch.equal(p.grad, want)
    ok["ddp_average"] = bool(same)
    ok["ddp_calls"] = sorted(set(fb.calls))
    # gradients kept across iterations (zeroed in place): they live in the buckets, autograd accumulates
    # into the views and the averaging runs in place -- no copy in, no copy out
    del fb.calls[:]
    inplace = True
    for it in range(3):
        g = torch.Generator().manual_seed(900 * it + rank)
        xb = torch.randn(7, 6, generator=g)
        net.zero_grad(set_to_none=False)
        ref.zero_grad()
        ddp(xb).pow(2).sum().backward()
        ref(xb).pow(2).sum().backward()
        for p, q in zip(net.parameters(), ref.parameters()):
            local = q.grad.clone()
            dist.all_reduce(local)
            inplace &= torch.equal(p.grad, local * torch.tensor(0.5))
    inplace &= all(p.grad is v for b, ps in enumerate(ddp._bucket_params)
                   for p, v in zip(ps, ddp._pipeline.slots[b][0].views))
    ok["ddp_in_place"] = bool(inplace) and "copy" not in fb.calls and "flatten" not in fb.calls
    ok["ddp_layout"] = (ddp.num_buckets, list(ddp.bucket_sizes))

    # Reducer on a list of tensors and on a module
    t = [torch.full((3,), float(rank)), [torch.full((2,), 2.0 * rank).half()]]
    red = Reducer(t)
    red.reduce()
    ok["reducer_list"] = bool((t[0] == 0.5).all() and (t[1][0] == 1.0).all())
    m2 = torch.nn.Linear(3, 2)
    red2 = Reducer(m2)
    m2(torch.full((1, 3), float(rank + 1))).sum().backward()
    red2.reduce()
    ok["reducer_module"] = bool(torch.equal(m2.weight.grad, torch.full((2, 3), 1.5)))

    torch.save(ok, os.path.join(out_dir, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_two(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    results = [torch.load(os.path.join(str(tmp_path), "rank%d.pt" % r), weights_only=False) for r in range(2)]
    for r in results:
        for key in ("flat_half", "flat_float", "fp16_sum_overflow_stays_inf", "race_known_answer", "race_buckets",
                    "broadcast_from_rank0", "ddp_average", "ddp_in_place", "reducer_list", "reducer_module"):
            assert r[key] is True, (key, r)
        assert {"flatten", "scale_unflatten", "scale_flat", "copy"} <= set(r["ddp_calls"])
    # rank 0 dictates the structure: identical on both ranks
    assert results[0]["ddp_layout"] == results[1]["ddp_layout"] == (3, [3, 1, 2])


def test_scatter_pass_hands_its_overflow_check_to_the_optimizer(gloo_single):
    """On multi-GPU runs the averaged gradients are inspected by the scatter pass that writes them; the
    optimizer then needs no scan of its own -- unless something touched the gradients in between."""
    from apex_b200.fp16_utils import FP16_Optimizer
    from apex_b200.parallel.distributed import _BucketPipeline, verified_flag
    from tests import fake_backend as fb
    calls = gloo_single.calls
    ps = [torch.nn.Parameter((torch.randn(n) * 0.1).half()) for n in (40, 7, 300)]
    ps.append(torch.nn.Parameter(torch.randn(12) * 0.1))                     # an fp32 parameter too
    opt = FP16_Optimizer(torch.optim.SGD(ps, lr=0.1), dynamic_loss_scale=True,
                         dynamic_loss_args={"init_scale": 4.0}, verbose=False)
    pipe = _BucketPipeline(fb._Stream())

    def backward_with_ddp(bad=None, grads_seed=0):
        g = torch.Generator().manual_seed(grads_seed)
        for i, p in enumerate(ps):
            p.grad = (torch.randn(p.shape, generator=g) * 4).to(p.dtype)
            if bad == i:
                p.grad[3] = float("inf")
        pipe.begin_round()
        pipe.submit(0, [p.grad for p in ps[:3]], verify=True)               # the fp16 bucket
        pipe.submit(1, [ps[3].grad], verify=True)                            # the fp32 bucket
        pipe.join(fb._Stream())

    backward_with_ddp()
    assert verified_flag([p.grad for p in ps]) is pipe.check_flag
    n = calls.count("check")
    opt.update_master_grads()
    opt.step()
    assert calls.count("check") == n and opt.overflow is False               # no scan of its own
    backward_with_ddp(bad=1, grads_seed=1)
    before = [p.data.clone() for p in ps]
    opt.update_master_grads()
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()) as out:
        opt.step()
    assert calls.count("check") == n and opt.overflow is True and "OVERFLOW!" in out.getvalue()
    assert all(torch.equal(a, p.data) for a, p in zip(before, ps)) and opt.loss_scale == 2.0
    # a gradient edited after the averaging is no longer covered: the optimizer scans again
    backward_with_ddp(grads_seed=2)
    ps[0].grad.mul_(1.0)
    assert verified_flag([p.grad for p in ps]) is None
    opt.update_master_grads()
    opt.step()
    assert calls.count("check") == n + 1 and opt.overflow is False
    # an inf written after the averaging is caught by that scan
    backward_with_ddp(grads_seed=3)
    ps[2].grad[0] = float("nan")
    opt.update_master_grads()
    with contextlib.redirect_stdout(io.StringIO()):
        opt.step()
    assert opt.overflow is True and calls.count("check") == n + 2
    # gradients that were never averaged (no DDP) are scanned as before
    for p in ps:
        p.grad = torch.zeros_like(p)
    opt.update_master_grads()
    assert calls.count("check") == n + 3
    # an optimizer that owns only part of the round's gradients must not adopt the round's answer
    backward_with_ddp(grads_seed=4)
    assert verified_flag([p.grad for p in ps[:3]]) is None and verified_flag([p.grad for p in ps]) is not None
    # a round that has not been joined yet does not count
    for i, p in enumerate(ps):
        p.grad = torch.ones_like(p)
    pipe.begin_round()
    pipe.submit(0, [p.grad for p in ps[:3]], verify=True)
    assert verified_flag([p.grad for p in ps[:3]]) is None
    pipe.join(fb._Stream())
    assert verified_flag([p.grad for p in ps[:3]]) is pipe.check_flag


def test_bucket_bookkeeping_against_the_oracle_on_random_cases():
    """_BucketLayout / _BucketRound (host logic, no process group needed) against the restated
    reference algorithms (oracle/apex_ref.py: distributed.py:248-261, :176-180, :279-315)."""
    from hypothesis import given, settings, strategies as st
    from apex_b200.parallel.distributed import _BucketLayout, _BucketRound, _DTYPE_CLASS
    from oracle import apex_ref
    dts = [torch.float16, torch.float32, torch.float64]

    @settings(max_examples=150, deadline=None)
    @given(st.lists(st.tuples(st.integers(0, 2), st.integers(1, 5000)), min_size=1, max_size=30),
           st.integers(1, 12000), st.randoms(use_true_random=False))
    def check(spec, message_size, rnd):
        n = len(spec)
        dtypes = [dts[c] for c, _ in spec]
        numels = [k for _, k in spec]
        arrival = list(range(n))
        rnd.shuffle(arrival)
        layout = _BucketLayout(message_size)
        for i in arrival:
            layout.arrive(_DTYPE_CLASS[dtypes[i]], i, numels[i])
        layout.close()
        buckets, where = apex_ref.bucket_layout(arrival, numels, dtypes, message_size)
        assert layout.members == buckets
        assert layout.positions() == where
        # the vector that is broadcast decodes to the same structure
        copy = _BucketLayout(message_size)
        copy.decode(layout.encode())
        assert copy.members == buckets and copy.sizes == [len(b) for b in buckets]
        # a later backward in ANOTHER arrival order: issue order must still be 0, 1, 2, ...
        order = list(range(n))
        rnd.shuffle(order)
        rounds = _BucketRound(layout.sizes)
        issued = []
        for i in order:
            b, loc = where[i]
            issued += rounds.put(b, loc, object())
        assert issued == apex_ref.dispatch_order([where[i][0] for i in order], layout.sizes) == list(range(len(buckets)))
        rounds.verify_complete()
    check()
