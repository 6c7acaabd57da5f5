"""Helpers shared by the golden-replay tests (CPU oracle, CPU host logic, GPU parity)."""
import numpy as np
import torch

CASES = ["sgd_mom_wd_dynamic", "sgd_nesterov_static", "sgd_plain_dampening", "sgd_no_momentum",
         "adam_dynamic", "adam_wd_static", "adamw_static", "sgd_mixed_fp32_dynamic", "sgd_clip_dynamic",
         "sgd_default_dynamic"]


def tensor_from(arr, device="cpu"):
    """npz array -> tensor (uint16 arrays are fp16 bit patterns)."""
    a = np.array(arr)
    if a.dtype == np.uint16:
        return torch.from_numpy(a.view(np.float16).copy()).to(device)
    return torch.from_numpy(a.copy()).to(device)


def bits(t):
    a = t.detach().cpu().numpy()
    return a.view(np.uint16) if a.dtype == np.float16 else a


def init_params(arrays, name, device="cpu"):
    n = int(arrays[name + "/n_params"])
    return [torch.nn.Parameter(tensor_from(arrays["%s/init/%d" % (name, i)], device)) for i in range(n)]


def grads_for(arrays, name, it, n, device="cpu"):
    return [tensor_from(arrays["%s/grad/%d/%d" % (name, it, i)], device) for i in range(n)]


def make_inner(meta):
    kw = dict(meta["opt_kwargs"])
    if "betas" in kw:
        kw["betas"] = tuple(kw["betas"])
    cls = getattr(torch.optim, meta["optimizer"])
    return lambda ps: cls(ps, **kw)


def ulp_diff(a, b):
    """Distance in units in the last place between two float32 arrays (same sign assumed near 0)."""
    a = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7fffffff), a)
    b = np.where(b < 0, -(b & 0x7fffffff), b)
    return np.abs(a - b)


# ---- models used for the DDP bucket-layout fixtures (shared with make_golden.py) -------------

class MixedChain(torch.nn.Module):
    """Parameters of three dtypes used one after the other, so gradient hooks fire in
    reverse order of use: exercises the per-dtype running buckets and the leftover order."""
    SPEC = [(torch.float64, 60), (torch.float32, 10), (torch.float16, 40), (torch.float64, 120),
            (torch.float32, 12), (torch.float16, 25), (torch.float64, 36), (torch.float32, 3)]

    def __init__(self):
        super(MixedChain, self).__init__()
        gen = torch.Generator().manual_seed(2)
        self.ps = torch.nn.ParameterList(
            [torch.nn.Parameter((torch.randn(n, generator=gen) * 0.1).to(dt)) for dt, n in self.SPEC])

    def forward(self, x):
        z = x.double()
        for p in self.ps:
            z = z * (1.0 + p.double().sum())
        return z


def build_layout_model(kind):
    torch.manual_seed(0)
    if kind == "mlp3":
        return torch.nn.Sequential(torch.nn.Linear(6, 10), torch.nn.ReLU(), torch.nn.Linear(10, 10),
                                   torch.nn.ReLU(), torch.nn.Linear(10, 4))
    if kind == "mixed":
        return MixedChain()
    raise ValueError(kind)
