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
    for chunk in [int(c) for c in args.chunks.split(",")]:
        t1 = TensorTable(dev, 1, chunk)
        t1.update([g16])
        t2 = TensorTable(dev, 2, chunk)
        t5 = TensorTable(dev, 5, chunk)
        t5.update([g16, p32, m32, v32, p16])
        t4 = TensorTable(dev, 4, chunk)
        t4.update([g16, p32, m32, p16])
        partials = torch.zeros(max(t1.struct.total_chunks, 1), device=dev)
        cases = [
            ("check_overflow", 2, lambda: _C.mt_check_overflow(t1, 0, flag)),
            ("fused_adam", 28, lambda: _C.mt_fused_adam(t5, 1e-3, 0.9, 0.999, 1e-8, 0.0, -1e-3, 0.03, False, 1 / 65536., True, flag)),
            ("fused_sgd_mom", 20, lambda: _C.mt_fused_sgd(t4, 0.1, 0.9, 0.0, 1e-4, False, 1 / 65536., True, flag)),
            ("grads_to_master", 6, lambda: (t2.update([g16, p32]), _C.mt_grads_to_master(t2, 0, 1, 1 / 65536., True, None))),
            ("master_to_model", 6, lambda: (t2.update([p32, p16]), _C.mt_master_to_model(t2, 0, 1, None))),
            ("flatten_f16", 4, lambda: _C.mt_flatten(t1, 0, flat)),
            ("scale_unflatten_f16", 4, lambda: _C.mt_scale_unflatten(t1, flat, 0, 8, True, flag)),
            ("zero_f16", 2, lambda: _C.mt_zero(t1, 0, 2)),
            ("l2norm_f16", 2, lambda: _C.mt_l2norm(t1, 0, 1 / 65536., True, partials, norm)),
        ]
        for name, bpe, fn in cases:
            if args.only and args.only not in name:
                continue
            flag.zero_()
            ms = timed(fn)
            print("chunk %6d  %-20s %8.3f ms  %8.1f GB/s" % (chunk, name, ms, bpe * n / ms / 1e6))
        del t1, t2, t4, t5, partials
    a, b = torch.empty(1 << 30, dtype=torch.bfloat16, device=dev), torch.empty(1 << 30, dtype=torch.bfloat16, device=dev)
    ms = timed(lambda: b.copy_(a))
    print("torch copy_ 1Gi bf16: %.3f ms %.1f GB/s" % (ms, 4.0 * (1 << 30) / ms / 1e6))


if __name__ == "__main__":
    main()
