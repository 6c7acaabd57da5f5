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

# ---- the bench's situation: one round = the whole GPT-2-XL gradient set (3.1 GB) as K contiguous buckets,
# launched back to back; optionally an HBM-bound kernel of the length of the fused Adam between rounds
if os.environ.get("SPAN", "0") == "1":
    del arena
    torch.cuda.empty_cache()
    total = symm.padded_numel(1557611200, 2, world * 16)
    big = symm.Arena(total * 2 + (1 << 20), dev)
    bflag, bflag_off = big.new_flag()
    base = symm.HEADER_BYTES
    big.buf[base:base + total * 2].view(torch.float16).fill_(0.001)
    filler_a = torch.empty(1 << 30, dtype=torch.float32, device=dev)
    filler_b = torch.empty(1 << 30, dtype=torch.float32, device=dev)
    symm._force_p2p, symm._force_mc = False, False
    for ctas in (0, 16):
        symm._ctas = ctas
        for k in (1, 3, 6, 11, 44, 145):
            per = symm.padded_numel(total // k, 2, world * 16)
            spans = []
            at = 0
            while at < total:
                n = min(per, total - at)
                spans.append((base + at * 2, n))
                at += n

            def round_(with_filler=False):
                if with_filler:
                    filler_b.copy_(filler_a)           # ~8.6 GB of HBM traffic, about 1.3 ms
                for off, n in spans:
                    big.allreduce(off, n, torch.float16, world, bflag_off)
            report("3.1 GB as %d buckets ctas=%s" % (len(spans), ctas or "default"), timed(round_, reps=5), total * 2)
            if k == 11:
                ms = timed(lambda: round_(True), reps=5)
                ms_f = timed(lambda: filler_b.copy_(filler_a), reps=5)
                report("  same, an HBM copy before each round (copy alone %.3f ms)" % ms_f, ms - ms_f, total * 2)
dist.destroy_process_group()
