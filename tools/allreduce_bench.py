#!/usr/bin/env python
"""tools/allreduce_bench.py -- pure allreduce rates at N GPUs (torchrun): NCCL, torch's symmetric-memory
ops, and variants of our multicast kernel.  Development tool."""
import ctypes
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from apex_b200 import _C    # noqa: E402

rank = int(os.environ["RANK"])
world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
tune = ctypes.CDLL(os.path.join(ROOT, "tools", "libsymm_tune.so"))
tune.symm_tune_launch.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_int,
                                  ctypes.c_int64, ctypes.c_int, ctypes.c_void_p]


def timed(fn, reps=8):
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
        print("%-34s %8.3f ms  algbw %7.1f GB/s  busbw %7.1f GB/s" % (name, ms, alg, alg * 2 * (world - 1) / world), flush=True)


for mb in [int(x) for x in os.environ.get("SIZES_MB", "256,1024").split(",")]:
    n = mb * (1 << 20) // 2
    n = n // (8 * world) * (8 * world)
    nbytes = n * 2
    if rank == 0:
        print("---- %d MB fp16, %d ranks" % (mb, world), flush=True)
    plain = torch.ones(n, dtype=torch.float16, device=dev)
    report("nccl all_reduce", timed(lambda: dist.all_reduce(plain)), nbytes)
    buf = symm_mem.empty(n, dtype=torch.float16, device=dev)
    buf.fill_(1.0)
    h = symm_mem.rendezvous(buf, dist.group.WORLD)
    gname = dist.group.WORLD.group_name
    for opname in ("multimem_all_reduce_", "two_shot_all_reduce_"):
        try:
            op = getattr(torch.ops.symm_mem, opname)
            buf.fill_(0.001)
            report("torch symm_mem." + opname, timed(lambda: op(buf, "sum", gname)), nbytes)
        except Exception as e:      # noqa: BLE001
            if rank == 0:
                print("torch symm_mem.%s unavailable: %s" % (opname, str(e)[:80]))
    mc = int(h.multicast_ptr)
    mine = int(h.buffer_ptrs[h.rank])
    stream = torch.cuda.current_stream().cuda_stream

    def ours(variant, ctas):
        def fn():
            h.barrier(channel=0)
            rc = tune.symm_tune_launch(variant, mc, mine, rank, world, n, ctas, stream)
            assert rc == 0, rc
            h.barrier(channel=1)
        return fn
    full = os.environ.get("FULL", "0") == "1"
    for variant, label in ((0, "fused T512 U4"), (1, "fused T1024 U4"), (2, "fused T512 U8"), (3, "fused T256 U8"),
                           (4, "fused weak"), (5, "fused no acc::f32"), (6, "fused U1"), (7, "fused U16"),
                           (10, "reduce only"), (11, "bcast only"), (12, "reduce ; bcast U8")):
        if not full and variant not in (0, 10, 11):
            continue
        for ctas in (16, 32, 64, 148):
            if (variant in (1, 3, 6, 7, 10, 11) and ctas != 64) or (not full and variant == 0 and ctas > 32):
                continue
            buf.fill_(0.001)
            report("ours v%d %s ctas=%d" % (variant, label, ctas), timed(ours(variant, ctas)), nbytes)
    # reduce and broadcast as separate kernels on two streams, chunked so that bcast(c) overlaps reduce(c+1)
    tune.symm_tune_launch_range.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64,
                                            ctypes.c_int64, ctypes.c_int, ctypes.c_void_p]
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    count = n // 8 // world
    first = count * rank

    def split(chunks, ctas):
        def fn():
            cur = torch.cuda.current_stream()
            h.barrier(channel=0)
            start = torch.cuda.Event()
            start.record()
            sa.wait_event(start)
            sb.wait_event(start)
            per = (count + chunks - 1) // chunks
            for c in range(chunks):
                f0 = first + c * per
                cnt = min(per, count - c * per)
                if cnt <= 0:
                    break
                with torch.cuda.stream(sa):
                    tune.symm_tune_launch_range(10, mc, mine, f0, cnt, ctas, sa.cuda_stream)
                    ev = torch.cuda.Event()
                    ev.record()
                with torch.cuda.stream(sb):
                    sb.wait_event(ev)
                    tune.symm_tune_launch_range(11, mc, mine, f0, cnt, ctas, sb.cuda_stream)
            cur.wait_stream(sa)
            cur.wait_stream(sb)
            h.barrier(channel=1)
        return fn
    for chunks in (2, 4, 8):
        for ctas in (16, 32):
            buf.fill_(0.001)
            report("split reduce|bcast chunks=%d ctas=%d" % (chunks, ctas), timed(split(chunks, ctas)), nbytes)
    # two fused kernels on two streams, each half of the slice
    def two_fused(ctas):
        def fn():
            cur = torch.cuda.current_stream()
            h.barrier(channel=0)
            start = torch.cuda.Event()
            start.record()
            half = count // 2
            for st, f0, cnt in ((sa, first, half), (sb, first + half, count - half)):
                st.wait_event(start)
                tune.symm_tune_launch_range(0, mc, mine, f0, cnt, ctas, st.cuda_stream)
                cur.wait_stream(st)
            h.barrier(channel=1)
        return fn
    buf.fill_(0.001)
    report("two fused halves on two streams ctas=16", timed(two_fused(16)), nbytes)
    # the library entry point, both modes
    for label, mcp, ctas in (("lib multicast", mc, 64), ("lib p2p", 0, 148)):
        def fn():
            h.barrier(channel=0)
            _C.symm_allreduce(list(h.buffer_ptrs), mcp, rank, world, n, torch.float16, ctas)
            h.barrier(channel=1)
        buf.fill_(0.001)
        report(label, timed(fn), nbytes)
    del buf, h, plain
dist.destroy_process_group()
