"""Helpers for the -m gpu tests (kernels are always reached through the C ABI via apex_b200._C)."""
import numpy as np
import torch

from apex_b200 import _C
from apex_b200.multi_tensor import TensorTable

# ragged sizes around every boundary the launcher has: vector (8), tile (2048), chunk (16384 / 65536)
SIZES = [0, 1, 7, 8, 9, 2047, 2048, 2049, 4099, 16383, 16384, 16385, 65535, 65536, 65537, 200003, 1 << 20]


def dev():
    return torch.device("cuda", 0)


def randn16(n, gen, scale=1.0):
    return (torch.randn(n, generator=gen, device=dev()) * scale).half()


def randn32(n, gen, scale=1.0):
    return torch.randn(n, generator=gen, device=dev()) * scale


def unaligned_view(t):
    """A copy of t that starts one element into a fresh allocation (2 or 4 bytes off alignment)."""
    buf = torch.empty(t.numel() + 8, dtype=t.dtype, device=t.device)
    v = buf[1:1 + t.numel()]
    v.copy_(t)
    return v


def bits(t):
    a = t.detach().cpu().numpy()
    return a.view(np.uint16) if a.dtype == np.float16 else a


def table(lists, flags=None, chunk=16384):
    t = TensorTable(dev(), len(lists), chunk)
    t.update(lists, flags)
    return t


def flag():
    return torch.zeros(1, dtype=torch.int32, device=dev())


def gen(seed):
    g = torch.Generator(device=dev())
    g.manual_seed(seed)
    return g
