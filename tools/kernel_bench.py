#!/usr/bin/env python
"""tools/kernel_bench.py -- per-kernel HBM throughput through the C ABI on a GPT-2-shaped tensor list.

Development tool: sweeps the chunk size (elements per CTA) of every kernel family and prints
algorithmic GB/s, so that the constants in apex_b200/multi_tensor.py are measured, not guessed.
    python tools/kernel_bench.py [--workload gpt2-xl|gpt2-medium] [--chunks 2048,4096,...]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from apex_b200 import _C                     # noqa: E402
from apex_b200.multi_tensor import TensorTable   # noqa: E402
from bench import WORKLOADS, numel           # noqa: E402


def timed(fn, reps=5):
    if reps <= 1:                      # under ncu: one launch, no warm-up
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        return a.elapsed_time(b)
    fn()
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="gpt2-xl")
    ap.add_argument("--chunks", default="2048,4096,8192,16384,65536")
    ap.add_argument("--only", default="")
    ap.add_argument("--production", action="store_true",
                    help="every kernel at the chunk size the product uses (multi_tensor.py), for an ncu capture")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    shapes = WORKLOADS[args.workload][0]
    n = sum(numel(s) for s in shapes)
    dev = torch.device("cuda", 0)
    g16 = [torch.full(s, 0.5, dtype=torch.float16, device=dev) for s in shapes]
    p16 = [torch.zeros(s, dtype=torch.float16, device=dev) for s in shapes]
    p32 = [torch.zeros(s, dtype=torch.float32, device=dev) for s in shapes]
    m32 = [torch.zeros(s, dtype=torch.float32, device=dev) for s in shapes]
    v32 = [torch.zeros(s, dtype=torch.float32, device=dev) for s in shapes]
    flat = torch.zeros(n, dtype=torch.float16, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    norm = torch.zeros(1, device=dev)
    print("workload %s: %d tensors, %d elements" % (args.workload, len(shapes), n))
    from apex_b200 import multi_tensor as mt
    prod = {"check_overflow": mt.CHUNK_CHECK, "fused_adam": mt.CHUNK_STEP, "fused_sgd_mom": mt.CHUNK_STEP,
            "grads_to_master": mt.CHUNK_STREAM, "master_to_model": mt.CHUNK_STREAM, "flatten_f16": mt.CHUNK_STREAM,
            "scale_unflatten_f16": mt.CHUNK_STREAM, "copy_f16": mt.CHUNK_STREAM, "unscale_f32": mt.CHUNK_STREAM,
            "zero_f16": mt.CHUNK_ZERO, "l2norm_f16": mt.CHUNK_NORM, "scale_flat_f16": 16384}
    chunks = sorted(set(prod.values())) if args.production else [int(c) for c in args.chunks.split(",")]
    for chunk in chunks:
        t1 = TensorTable(dev, 1, chunk)
        t1.update([g16])
        # one table per two-list kernel, built OUTSIDE the timed lambdas (round 1 rebuilt them inside
        # and so charged ~0.15 ms of host time to grads_to_master / master_to_model / copy)
        t_g2m, t_m2m, t_cp, t_un = (TensorTable(dev, 2, chunk) for _ in range(4))
        t_g2m.update([g16, p32])
        t_m2m.update([p32, p16])
        t_cp.update([g16, p16])
        t_un.update([p32, m32], [1] * len(p32))
        t5 = TensorTable(dev, 5, chunk)
        t5.update([g16, p32, m32, v32, p16])
        t4 = TensorTable(dev, 4, chunk)
        t4.update([g16, p32, m32, p16])
        partials = torch.zeros(max(t1.struct.total_chunks, 1), device=dev)
        cases = [
            ("check_overflow", 2, lambda: _C.mt_check_overflow(t1, 0, flag)),
            ("fused_adam", 28, lambda: _C.mt_fused_adam(t5, 1e-3, 0.9, 0.999, 1e-8, 0.0, -1e-3, 0.03, False, 1 / 65536., True, flag)),
            ("fused_sgd_mom", 20, lambda: _C.mt_fused_sgd(t4, 0.1, 0.9, 0.0, 1e-4, False, 1 / 65536., True, flag)),
            ("grads_to_master", 6, lambda: _C.mt_grads_to_master(t_g2m, 0, 1, 1 / 65536., True, None)),
            ("master_to_model", 6, lambda: _C.mt_master_to_model(t_m2m, 0, 1, None)),
            ("flatten_f16", 4, lambda: _C.mt_flatten(t1, 0, flat)),
            ("scale_unflatten_f16", 4, lambda: _C.mt_scale_unflatten(t1, flat, 0, 8, True, flag)),
            ("copy_f16", 4, lambda: _C.mt_copy(t_cp, 0, 1, 2, None)),
            ("unscale_f32", 8, lambda: _C.mt_unscale(t_un, 0, 1.0, None, 0)),
            ("scale_flat_f16", 4, lambda: _C.scale_flat(flat, 8, flag)),
            ("zero_f16", 2, lambda: _C.mt_zero(t1, 0, 2)),
            ("l2norm_f16", 2, lambda: _C.mt_l2norm(t1, 0, 1 / 65536., True, partials, norm)),
        ]
        for name, bpe, fn in cases:
            if args.only and args.only not in name:
                continue
            if args.production and prod.get(name) != chunk:
                continue
            flag.zero_()
            ms = timed(fn, args.reps)
            print("chunk %6d  %-20s %8.3f ms  %8.1f GB/s" % (chunk, name, ms, bpe * n / ms / 1e6))
        del t1, t_g2m, t_m2m, t_cp, t_un, t4, t5, partials
    a, b = torch.empty(1 << 30, dtype=torch.bfloat16, device=dev), torch.empty(1 << 30, dtype=torch.bfloat16, device=dev)
    ms = timed(lambda: b.copy_(a))
    print("torch copy_ 1Gi bf16: %.3f ms %.1f GB/s" % (ms, 4.0 * (1 << 30) / ms / 1e6))


if __name__ == "__main__":
    main()
