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
    from apex_b200.parallel import DistributedDataParallel as DDP
    from apex_b200.parallel.distributed import flat_dist_call
    ok = {}

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
            expected = val * 4096 * 4096 * (2. * i + 1) / 2.
            passed &= (param.grad.data.sum().item() == expected)
    ok["race"] = bool(passed)

    # fp16 buckets: averaged gradients within rtol 1e-3 of the fp32 mean (north_star), sum-then-divide
    g = torch.Generator(device="cuda").manual_seed(5 + rank)
    tensors = [(torch.randn(n, generator=g, device="cuda") * 100).half() for n in (5, 4099, 1 << 20, 33)]
    exact = [t.float() for t in tensors]
    for e in exact:
        dist.all_reduce(e)
    flat_dist_call(tensors, dist.all_reduce)
    torch.cuda.synchronize()
    close = all(torch.allclose(t.float(), e / world, rtol=1e-3, atol=1e-2) for t, e in zip(tensors, exact))
    ok["fp16_average"] = bool(close)
    big = torch.full((1000,), 40000.0, device="cuda").half()
    flat_dist_call([big, big.clone()], dist.all_reduce)
    ok["fp16_overflow_stays_inf"] = bool(torch.isinf(big).all())
    torch.save(ok, os.path.join(out_dir, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_ranks_nccl(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        res = torch.load(os.path.join(str(tmp_path), "rank%d.pt" % r), weights_only=False)
        assert res == {"race": True, "fp16_average": True, "fp16_overflow_stays_inf": True}, res
