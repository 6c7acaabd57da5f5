#!/usr/bin/env python
"""tools/host_profile.py -- where does the python time of one step go? (cProfile, GPU box)"""
import cProfile
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                  # noqa: E402
from apex_b200.fp16_utils import FP16_Optimizer  # noqa: E402

w = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
shapes, _, kind = bench.WORKLOADS[w]
mask = bench.FP32_MASK.get(w, [False] * len(shapes))
dev = torch.device("cuda", 0)
params = [torch.nn.Parameter(torch.randn(s, device=dev) * 0.02 if f else (torch.randn(s, device=dev) * 0.02).half())
          for s, f in zip(shapes, mask)]
opt = FP16_Optimizer(bench.make_optimizer(kind)(params), dynamic_loss_scale=True,
                     dynamic_loss_args={"init_scale": 2.0**16}, verbose=False)
for p in params:
    p.grad = (torch.randn(p.shape, device=dev) * 60).to(p.dtype)


def step():
    opt.update_master_grads()
    opt.step()


for _ in range(10):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
