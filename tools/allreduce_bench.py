#!/usr/bin/env python
"""tools/allreduce_bench.py -- isolated rate of the bucket-averaging kernel at N GPUs (torchrun), next to
NCCL's all_reduce on the same buffer size.  CUDA events around back-to-back launches on one stream,
max over ranks.  Development tool.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/allreduce_bench.py
    SIZES_MB=20,64,256,1024  MODES=mc,p2p  CTAS=16,32,64,148
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from apex_b200.parallel import symm    # noqa: E402

rank = int(os.environ["RANK"])
world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def report(name, ms, nbytes):
    if rank == 0:
        alg = nbytes / ms / 1e6
        print("%-44s %8.3f ms  algbw %7.1f GB/s  busbw %7.1f GB/s" % (name, ms, alg, alg * 2 * (world - 1) / world),
              flush=True)


sizes = [int(x) for x in os.environ.get("SIZES_MB", "20,64,256,1024").split(",")]
modes = os.environ.get("MODES", "mc,p2p").split(",")
cta_list = [int(x) for x in os.environ.get("CTAS", "0").split(",")]
arena = symm.Arena(max(sizes) * (1 << 20) + 4096, dev)
flag, flag_off = arena.new_flag()
for mb in sizes:
    n = symm.padded_numel(mb * (1 << 20) // 2, 2, world)
    nbytes = n * 2
    if rank == 0:
        print("---- %d MB fp16, %d ranks" % (mb, world), flush=True)
    plain = torch.full((n,), 0.001, dtype=torch.float16, device=dev)
    report("nccl all_reduce", timed(lambda: dist.all_reduce(plain)), nbytes)
    report("nccl all_reduce ; div_ (reference)", timed(lambda: (dist.all_reduce(plain), plain.div_(world))), nbytes)
    del plain
    off = symm.HEADER_BYTES
    view = arena.buf[off:off + nbytes].view(torch.float16)
    for mode in modes:
        if mode == "mc" and not arena.has_multicast:
            continue
        symm._force_p2p, symm._force_mc = mode == "p2p", mode == "mc"
        for ctas in cta_list:
            symm._ctas = ctas
            view.fill_(0.001)
            torch.cuda.synchronize()
            dist.barrier()
            report("ours %s ctas=%s (sum, x1/N, check, barriers)" % (mode, ctas or "default"),
                   timed(lambda: arena.allreduce(off, n, torch.float16, world, flag_off)), nbytes)
            assert int(flag.item()) == 0 and abs(float(view[0]) - 0.001) < 1e-4, (float(view[0]), int(flag.item()))
dist.destroy_process_group()
