#!/usr/bin/env python
"""bench.py -- the mixed-precision gradient step on B200, one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json metric: "fused MP step ... (1.5B params)"): a GPT-2-XL-shaped set of
580 fp16 parameter tensors (1,557,611,200 elements) with fp16 gradients, fp32 masters and fp32
Adam state, all synthetic and resident in HBM.  One step = what the reference does between the
end of backward and the next forward (apex/fp16_utils/fp16_optimizer.py:486-497, :330-384):
inf/NaN check of every gradient feeding DynamicLossScaler, unscale into fp32, Adam update of the
fp32 masters, copy back to the fp16 parameters -- driven through apex_b200.fp16_utils.FP16_Optimizer.
With N > 1 each rank first averages its gradients with the other ranks bucket by bucket
(apex/parallel/distributed.py:11-23: flatten, all_reduce, /= world, copy back -- here: the gradients
live in the buckets and one kernel per bucket does barrier, sum, x 1/N, inf/NaN check, barrier) on
side streams, exactly the calls DistributedDataParallel makes per bucket; afterwards, outside the
timed region, the averaged gradients are checked against the reference composition (`parity`).
Short side runs (BASELINE.json configs 2, 3, 4, the static-scale step, the reference's default
bucket size, ResNet-50 training through DistributedDataParallel) follow the headline.

`value` counts parameter elements updated per second over all ranks (weak scaling: every rank
owns a full replica and its own gradients).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ADAM = dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0)
SGD = dict(lr=0.1, momentum=0.9, weight_decay=1e-4)
# algorithmic bytes per parameter element of the fused update
BYTES_PER_PARAM = {"adam": 28,   # r g16 2, r p32 4, r m 4, r v 4, w p32 4, w m 4, w v 4, w p16 2
                   "sgd": 20}    # r g16 2, r p32 4, r mom 4, w p32 4, w mom 4, w p16 2
FUSED_WRAPPER = {"adam": "mt_fused_adam", "sgd": "mt_fused_sgd"}
FUSED_KERNEL = {"adam": "mt_fused_adam_kernel", "sgd": "mt_fused_sgd_kernel<true>"}


def gpt2_shapes(L, d, V=50257, T=1024):
    per_layer = [(d,), (d,), (d, 3 * d), (3 * d,), (d, d), (d,), (d,), (d,), (d, 4 * d), (4 * d,), (4 * d, d), (d,)]
    return [(V, d), (T, d)] + per_layer * L + [(d,), (d,)]


def numel(shape):
    n = 1
    for s in shape:
        n *= s
    return n


def resnet50_shapes():
    """(shape, is_fp32) of torchvision-style ResNet-50 after network_to_half: convs/fc fp16, BatchNorm fp32
    (apex/fp16_utils/fp16util.py:21-41).  161 tensors, 25,557,032 parameters."""
    out = [((64, 3, 7, 7), False), ((64,), True), ((64,), True)]
    inplanes = 64
    for planes, blocks, stride in ((64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)):
        for b in range(blocks):
            out += [((planes, inplanes, 1, 1), False), ((planes,), True), ((planes,), True),
                    ((planes, planes, 3, 3), False), ((planes,), True), ((planes,), True),
                    ((planes * 4, planes, 1, 1), False), ((planes * 4,), True), ((planes * 4,), True)]
            if b == 0:
                out += [((planes * 4, inplanes, 1, 1), False), ((planes * 4,), True), ((planes * 4,), True)]
            inplanes = planes * 4
    out += [((1000, 2048), False), ((1000,), False)]
    return out


WORKLOADS = {
    # name: (shapes, description, default optimizer)
    "gpt2-xl": (gpt2_shapes(48, 1600), "GPT-2-XL-shaped 1.5B fp16 params, 580 tensors", "adam"),
    "gpt2-medium": (gpt2_shapes(24, 1024), "GPT-2-medium-shaped 355M fp16 params, 292 tensors", "sgd"),
    "resnet50": ([s for s, _ in resnet50_shapes()], "ResNet-50 shapes, 25.6M params, 55 fp16 + 106 fp32 tensors", "sgd"),
    "mlp": ([(4096, 1024), (4096,), (1024, 4096), (1024,)], "config-1 MLP, 8.4M fp16 params, 4 tensors", "sgd"),
}
FP32_MASK = {"resnet50": [f for _, f in resnet50_shapes()]}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:       # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(object):
    """nvidia-smi clocks and throttle reasons DURING the timed region."""
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:   # noqa: BLE001
            self.proc = None
            return

        def pump():
            for line in self.proc.stdout:
                self.lines.append(line.strip())
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:   # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# the CPU arm: the oracle's restatement of the reference composition on host cores
# ---------------------------------------------------------------------------------------------

def make_optimizer(kind):
    if kind == "adam":
        return lambda ps: torch.optim.Adam(ps, **ADAM)
    return lambda ps: torch.optim.SGD(ps, **SGD)


def mem_available_gb():
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) / 1e6
    except Exception:       # noqa: BLE001
        pass
    return 0.0


def _imported_reference():
    """The unmodified reference package, if it can be imported on this machine (APEX_REFERENCE or
    /root/reference -- present in the build container, absent on the GPU box).  Returns the module or None."""
    import importlib.util
    if os.environ.get("APEX_REFERENCE") == "none":       # force the port (what the GPU box runs)
        return None
    for root in (os.environ.get("APEX_REFERENCE"), "/root/reference"):
        init = os.path.join(root, "apex", "__init__.py") if root else None
        if init and os.path.exists(init):
            try:
                spec = importlib.util.spec_from_file_location("apex", init, submodule_search_locations=[os.path.dirname(init)])
                mod = importlib.util.module_from_spec(spec)
                sys.modules["apex"] = mod
                spec.loader.exec_module(mod)
                import apex.fp16_utils      # noqa: F401
                return mod
            except Exception:   # noqa: BLE001
                sys.modules.pop("apex", None)
    return None


def cpu_reference_run(shapes, steps, warmup, budget_s, seed=1234, kind="adam", max_params=None, label=""):
    """Time update_master_grads() + step() of the reference's CPU path on host cores
    (apex/fp16_utils/fp16_optimizer.py:486-497, :330-384): the imported, unmodified reference under the
    Tensor.type shim when it is importable here (kind "reference"), else the oracle's restatement of
    it (oracle/apex_ref.py, kind "port").  ``max_params`` bounds the sample (leading tensors of the
    workload); None = the whole workload.  Returns (params_per_s, info)."""
    # all host cores (torchrun exports OMP_NUM_THREADS=1 to every rank; the other ranks idle or have exited)
    torch.set_num_threads(max(torch.get_num_threads(), os.cpu_count() or 1))
    sample, total = [], 0
    for s in shapes:
        if max_params is not None and total + numel(s) > max_params:
            if sample:
                break
            continue
        sample.append(s)
        total += numel(s)
    gen = torch.Generator().manual_seed(seed)
    params = [torch.nn.Parameter((torch.randn(s, generator=gen) * 0.02).half()) for s in sample]
    kw = dict(dynamic_loss_scale=True, dynamic_loss_args={"init_scale": 2.0**16})
    apex = _imported_reference()
    if apex is not None:
        from unittest import mock
        real_type = torch.Tensor.type

        def cuda_named(self, *a, **k):
            if a or k:
                return real_type(self, *a, **k)
            name = real_type(self)
            return name if name.startswith("torch.cuda") else name.replace("torch.", "torch.cuda.")
        with mock.patch.object(torch.Tensor, "type", cuda_named):        # the constructor's string type gate
            mp = apex.fp16_utils.FP16_Optimizer(make_optimizer(kind)(params), verbose=False, **kw)
        impl = "reference"
    else:
        from oracle import apex_ref
        mp = apex_ref.MixedPrecisionStep(params, make_optimizer(kind), **kw)
        impl = "port"
    for p, s in zip(params, sample):
        p.grad = (torch.randn(s, generator=gen) * (1e-3 * 65536)).half()

    def one():
        mp.update_master_grads()
        mp.step()
    t_end = time.perf_counter() + budget_s
    for _ in range(warmup):
        one()
        if time.perf_counter() > t_end:
            break
    times = []
    t_end = time.perf_counter() + budget_s
    for _ in range(steps):
        t0 = time.perf_counter()
        one()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() > t_end and len(times) >= 2:
            break
    assert not mp.overflow
    best = statistics.median(times)
    whole = len(sample) == len(shapes)
    info = {"sample": "%s%s: %d tensors, %d params, %d timed steps, median"
                      % (label, " (whole workload)" if whole else " (leading tensors)", len(sample), total, len(times)),
            "cores": torch.get_num_threads(), "host_cpus": os.cpu_count(), "ms_per_step": best * 1e3,
            "params": total, "tensors": len(sample), "steps": len(times), "kind": impl, "whole_workload": whole}
    del mp, params
    return total / best, info


def cpu_arm_plan(workload):
    """Which configuration the CPU arm can time in full on this host: the headline workload needs
    about 16 B/param of tensors plus Adam's temporaries (1.5576 G params: ~45 GB)."""
    shapes = WORKLOADS[workload][0]
    need_gb = sum(numel(s) for s in shapes) * 36 / 1e9
    if mem_available_gb() >= need_gb + 8:
        return workload, shapes
    return "gpt2-medium", WORKLOADS["gpt2-medium"][0]


def gpu_reference_composition(shapes, fp32_mask, kind, dev, steps=3):
    """The reference's own composition of stock torch ops ON THE GPU (what apex runs there today, the
    honest "before"): per-tensor `float(x.float().sum())` overflow test with a host sync each
    (apex/fp16_utils/loss_scaler.py:84-110), per-tensor copy_ and mul_ (fp16util.py:106-112,
    fp16_optimizer.py:236-241), torch.optim step (:380), per-tensor copy_ back (fp16util.py:127-129).
    Written out here with plain torch calls; reported next to our number, never used by it."""
    gen = torch.Generator(device=dev).manual_seed(1234)
    params = [torch.randn(s, generator=gen, device=dev) * 0.02 if f else
              (torch.randn(s, generator=gen, device=dev) * 0.02).half() for s, f in zip(shapes, fp32_mask)]
    grads = [(torch.randn(s, generator=gen, device=dev) * (1e-3 * 65536)).to(p.dtype) for s, p in zip(shapes, params)]
    masters = [p.detach().clone().float().requires_grad_(True) for p in params]
    opt = make_optimizer(kind)(masters)
    inv = 1. / 65536.0

    def one():
        for g in grads:                                   # has_overflow: a sync per tensor
            s = float(g.float().sum())
            if s == float("inf") or s == -float("inf") or s != s:
                return
        for m, g in zip(masters, grads):
            if m.grad is None:
                m.grad = torch.empty_like(m)
            m.grad.copy_(g)
        for m in masters:
            m.grad.mul_(inv)
        opt.step()
        for p, m in zip(params, masters):
            p.copy_(m)
    for _ in range(2):
        one()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / steps * 1e3
    del opt, masters, grads, params
    torch.cuda.empty_cache()
    return ms


def run_reference_arm(args, shapes):
    """`bench.py --impl reference`: the reference's own CPU path for the step, on all host cores, on the
    SAME configuration as our arm when the host has the memory for it (GPT-2-XL: ~45 GB), else on
    GPT-2-medium; config 1 (the FP16_Optimizer_simple MLP, BASELINE.json configs[0]) is always timed
    beside it.  A step is one update_master_grads() + step() over the whole parameter set."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    mlp_v, mlp = cpu_reference_run(WORKLOADS["mlp"][0], steps=20, warmup=5, budget_s=10.0, kind="sgd",
                                   label="config 1 MLP")
    name, use = cpu_arm_plan(args.workload)
    steps = max(2, min(args.steps, 5))
    value, info = cpu_reference_run(use, steps, min(args.warmup, 2), budget_s=90.0, kind=args.optimizer, label=name)
    line = {
        "impl": "reference", "metric": "fused_mp_step_throughput", "value": value, "unit": "params/s",
        "n_gpus": args.gpus, "steps": info["steps"], "warmup": min(args.warmup, 2), "ms_per_step": info["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": name + ": " + WORKLOADS[name][1] +
                   "; dynamic-loss-scale check + unscale + %s + fp16 copy" % args.optimizer,
                   "optimizer": args.optimizer, "params": info["params"], "tensors": info["tensors"],
                   "parallelism": "host cpu", "where": "host CPU, torch %s" % torch.__version__},
        "cpu_baseline": {"value": value, "unit": "params/s", "cores": info["cores"], "kind": info["kind"],
                         "sample": info["sample"]},
        "config1_mlp": {"value": mlp_v, "unit": "params/s", "ms_per_step": mlp["ms_per_step"], "params": mlp["params"],
                        "kind": mlp["kind"], "sample": mlp["sample"]},
        "e2e": {"value": value, "unit": "params/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------

class Timed(object):
    """CUDA-event timing of one python-level kernel wrapper, on the stream it launches on."""

    def __init__(self, module, name):
        self.module, self.name = module, name
        self.real = getattr(module, name)
        self.pairs = []
        self.on = False
        setattr(module, name, self)

    def __call__(self, *a, **k):
        if not self.on:
            return self.real(*a, **k)
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        out = self.real(*a, **k)
        e1.record()
        self.pairs.append((e0, e1))
        return out

    def total_ms(self):
        return sum(a.elapsed_time(b) for a, b in self.pairs)

    def count(self):
        return len(self.pairs)


def build_buckets(params, bucket_elems):
    """Bucket structure DistributedDataParallel learns when gradients arrive in reverse parameter
    order (the usual order for a sequential model): apex/parallel/distributed.py:248-261, :176-180."""
    from apex_b200.parallel.distributed import _BucketLayout, _DTYPE_CLASS
    layout = _BucketLayout(bucket_elems)
    for i in reversed(range(len(params))):
        layout.arrive(_DTYPE_CLASS[params[i].dtype], i, params[i].numel())
    layout.close()
    return layout.members


def reference_bucket_rule(numels, dtypes, message_size):
    """apex/parallel/distributed.py:248-261 and :176-180 restated for gradients arriving in reverse
    parameter order: per dtype slot (half, float, double) a running bucket, cut as soon as it holds
    >= message_size elements; leftovers appended in slot order.  Written independently of
    apex_b200.parallel so that the layout the bench uses can be checked against it."""
    slot_of = {torch.float16: 0, torch.float32: 1, torch.float64: 2}
    closed, running, count = [], [[], [], []], [0, 0, 0]
    for i in reversed(range(len(numels))):
        k = slot_of[dtypes[i]]
        running[k].append(i)
        count[k] += numels[i]
        if count[k] >= message_size:
            closed.append(running[k])
            running[k], count[k] = [], 0
    return closed + [r for r in running if r]


def multi_gpu_parity(pipeline, buckets, side, dev, rank, world):
    """Outside the timed region: one round on FRESH per-rank gradients, our bucket path against the
    reference composition on the same inputs and the same GPUs -- torch.cat -> dist.all_reduce (NCCL,
    in the bucket dtype) -> /= world (apex/parallel/distributed.py:11-23).  north_star: rtol 1e-3 for
    fp16 sums.  Also: identical bits on every rank, the overflow word clean, then raised on every
    rank by a 40000 + 40000 fp16 sum in one rank-slice.  Returns a dict with "pass"."""
    import torch.distributed as dist
    cur = torch.cuda.current_stream()
    gen = torch.Generator(device=dev).manual_seed(4242 + rank)
    out = {"buckets": len(buckets), "reference": "torch.cat -> NCCL all_reduce (fp16) -> /= world, same inputs",
           "rtol": 1e-3}

    def one_round():
        pipeline.begin_round()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for k, b in enumerate(buckets):
                pipeline.submit(k, [p.grad for p in b], verify=True, owners=b)
        pipeline.join(cur)
        torch.cuda.synchronize()

    refs, absmax = [], 0.0
    for b in buckets:
        for p in b:
            p.grad.copy_((torch.randn(p.shape, generator=gen, device=dev) * 65.536).to(p.grad.dtype))
        flat = torch.cat([p.grad.reshape(-1) for p in b])
        absmax = max(absmax, float(flat.abs().max()))
        dist.all_reduce(flat)
        flat /= world
        refs.append(flat)
    one_round()
    amax = torch.tensor([absmax], device=dev)
    dist.all_reduce(amax, op=dist.ReduceOp.MAX)
    atol = 1e-3 * float(amax.item())                 # 1e-3 of the largest addend: sums near zero have no relative scale
    close, equal, total, worst = True, 0, 0, 0.0
    sums = []
    for b, ref in zip(buckets, refs):
        mine = torch.cat([p.grad.reshape(-1) for p in b])
        close &= bool(torch.allclose(mine.float(), ref.float(), rtol=1e-3, atol=atol))
        worst = max(worst, float((mine.float() - ref.float()).abs().max()))
        equal += int((mine == ref).sum())
        total += mine.numel()
        bits = mine.view(torch.int16 if mine.dtype == torch.float16 else torch.int32).to(torch.int64)
        sums.append(bits.sum() + (bits * (torch.arange(bits.numel(), device=dev) % 8191)).sum())
    del refs
    digest = torch.stack(sums)
    every = [torch.empty_like(digest) for _ in range(world)]
    dist.all_gather(every, digest)
    out["averaged_grads_within_rtol"] = close
    out["atol"] = atol
    out["max_abs_diff"] = worst
    out["fraction_bit_equal_to_reference"] = equal / max(total, 1)
    out["identical_bits_on_all_ranks"] = all(bool(torch.equal(e, every[0])) for e in every)
    out["flag_clean"] = int(pipeline.check_flag.item()) == 0
    # a sum that overflows fp16, sitting in ONE rank's slice of the last bucket
    victim = buckets[-1][-1].grad.view(-1)
    victim[-8:] = 40000.0
    one_round()
    word = torch.tensor([int(pipeline.check_flag.item() != 0)], device=dev)
    dist.all_reduce(word, op=dist.ReduceOp.MIN)
    out["overflow_word_raised_on_all_ranks"] = bool(word.item()) and bool(torch.isinf(victim[-8:]).all())
    victim[-8:] = 0.0
    out["pass"] = all(out[k] for k in ("averaged_grads_within_rtol", "identical_bits_on_all_ranks", "flag_clean",
                                       "overflow_word_raised_on_all_ranks"))
    return out


def lstm_shapes(V=33278, H=1500, layers=2):
    """examples/word_language_model/model.py RNNModel('LSTM', 33278, 1500, 1500, 2): 11 tensors,
    135,891,278 parameters."""
    out = [(V, H)]
    for _ in range(layers):
        out += [(4 * H, H), (4 * H, H), (4 * H,), (4 * H,)]
    return out + [(V, H), (V,)]


def side_step_bench(label, shapes, fp32_mask, make_opt, bytes_per_param, dev, steps=20, warmup=5, clip=None,
                    dynamic=True, fused_name="mt_fused_sgd"):
    """A short run of another configuration of the same step (after the headline; N = 1, rank 0):
    whole step through FP16_Optimizer plus the CUDA-event time of its kernels."""
    from apex_b200 import _C
    from apex_b200.fp16_utils import FP16_Optimizer
    gen = torch.Generator(device=dev).manual_seed(1234)
    params = [torch.nn.Parameter(torch.randn(s, generator=gen, device=dev) * 0.02 if f32 else
                                 (torch.randn(s, generator=gen, device=dev) * 0.02).half())
              for s, f32 in zip(shapes, fp32_mask)]
    n_params = sum(p.numel() for p in params)
    if dynamic:
        opt = FP16_Optimizer(make_opt(params), dynamic_loss_scale=True,
                             dynamic_loss_args={"init_scale": 2.0**16, "scale_window": 1000}, verbose=False)
    else:
        opt = FP16_Optimizer(make_opt(params), static_loss_scale=128.0, verbose=False)
    scale = float(opt.loss_scale)
    for s, p in zip(shapes, params):
        p.grad = (torch.randn(s, generator=gen, device=dev) * (1e-3 * scale)).to(p.dtype)
    timers = [Timed(_C, n) for n in ("mt_check_overflow_publish", fused_name, "mt_l2norm")]
    host = [0.0]

    def step():
        t0 = time.perf_counter()
        opt.update_master_grads()
        if clip is not None:
            opt.clip_master_grads(clip)
        opt.step()
        host[0] += time.perf_counter() - t0
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    assert opt.overflow is False
    for t in timers:
        t.on = True
    host[0] = 0.0
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step()
    b.record()
    torch.cuda.synchronize()
    step_ms = a.elapsed_time(b) / steps
    per = {}
    for t in timers:
        t.on = False
        if t.count():
            per[t.name] = t.total_ms() / steps
        setattr(_C, t.name, t.real)
    peak, _ = measured_peak()
    fused_ms = per.get(fused_name)
    out = {"workload": label, "params": n_params, "tensors": len(shapes), "steps": steps, "step_ms": step_ms,
           "kernel_ms": per, "kernels_sum_ms": sum(per.values()), "host_python_ms_per_step": host[0] / steps * 1e3,
           "algorithmic_bytes_per_param_fused": bytes_per_param}
    if fused_ms:
        gbs = bytes_per_param * n_params / (fused_ms * 1e-3) / 1e9
        out["fused_kernel_gbs"] = gbs
        out["fused_kernel_frac_of_measured_peak"] = gbs / peak
        out["fused_kernel_frac_of_8tbs"] = gbs / 8000.0
    del opt, params
    torch.cuda.empty_cache()
    return out


def side_configs(dev):
    """BASELINE.json configs 2, 3, 4 and the static-scale variant of the headline, one short run each."""
    out = {}
    sgd_mom = lambda ps: torch.optim.SGD(ps, **SGD)                                  # noqa: E731
    out["config2_gpt2_medium_sgd_momentum"] = side_step_bench(
        "GPT-2-medium shapes, 355M fp16 params, SGD momentum 0.9 wd 1e-4, dynamic scale (target: fused <= 1.109 ms)",
        WORKLOADS["gpt2-medium"][0], [False] * 292, sgd_mom, 20, dev)
    out["config3_resnet50_shapes_sgd_momentum"] = side_step_bench(
        "ResNet-50 shapes after network_to_half, 55 fp16 + 106 fp32 tensors, SGD momentum, dynamic scale "
        "(launch-latency bound: 0.51 GB per step)", WORKLOADS["resnet50"][0], FP32_MASK["resnet50"], sgd_mom, 20, dev,
        steps=50)
    out["config4_lstm_sgd_clip"] = side_step_bench(
        "word_language_model LSTM 2x1500, 33278 vocab, 136M fp16 params, plain SGD lr=20, dynamic scale, "
        "clip_master_grads(0.25) (check 2 + norm 2 + update 12 B/param)", lstm_shapes(), [False] * 11,
        lambda ps: torch.optim.SGD(ps, lr=20.0), 12, dev, clip=0.25)
    out["headline_static_scale"] = side_step_bench(
        "GPT-2-XL shapes, Adam, STATIC loss scale 128 (the reference never checks then: loss_scaler.py:26-31): "
        "the step is the single fused launch", WORKLOADS["gpt2-xl"][0], [False] * 580,
        lambda ps: torch.optim.Adam(ps, **ADAM), 28, dev, dynamic=False, fused_name="mt_fused_adam", steps=10, warmup=3)
    return out


def nvlink_roofline(s_bytes, ms, world):
    """B200_PROFILING.md: bytes that must cross NVLink per direction per GPU / the measured 770 GB/s peer
    copy rate.  Two-shot over the switch multicast: every GPU sends its copy of all slices up (S) and the
    reduced slice it owns (S/n), and receives as much.  Two-shot with peer loads/stores: 2(n-1)/n * S."""
    from apex_b200.parallel import symm as _symm
    if not _symm.enabled():
        return None
    multicast = world > 2 and not _symm._force_p2p
    per_dir = s_bytes * (1.0 + 1.0 / world) if multicast else s_bytes * 2.0 * (world - 1) / world
    floor_ms = per_dir / 770e9 * 1e3
    return {"bytes_per_direction": per_dir, "measured_peer_copy_gbs": 770.0, "floor_ms": floor_ms,
            "achieved_gbs_per_direction": per_dir / (ms * 1e-3) / 1e9, "frac": floor_ms / ms,
            "variant": "multicast" if multicast else "peer-to-peer"}


def default_bucket_path(params, dev, world, rounds=5, warmup=2, message_size=10000000):
    """The bucket path of the headline's gradients at the reference's DEFAULT bucket size
    (message_size = 10,000,000 elements = 20 MB of fp16, apex/parallel/distributed.py:124), or at any
    other ``message_size`` (BASELINE.json config 5: the sweep 1-256 MB, `bench.py --bucket-sweep`)."""
    import torch.distributed as dist
    from apex_b200.parallel.distributed import _BucketPipeline
    side = torch.cuda.Stream()
    pipeline = _BucketPipeline(side)
    buckets = [[params[i] for i in b] for b in build_buckets(params, message_size)]
    pipeline.plan([(k, [p.data for p in b]) for k, b in enumerate(buckets)])
    cur = torch.cuda.current_stream()

    def one():
        pipeline.begin_round()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for k, b in enumerate(buckets):
                pipeline.submit(k, [p.grad for p in b], verify=True, owners=b)
        pipeline.join(cur)
    for _ in range(warmup):
        one()
    torch.cuda.synchronize()
    dist.barrier()
    a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(rounds):
        one()
    b_.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b_) / rounds], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    nbytes = sum(p.numel() * p.element_size() for p in params)
    algbw = nbytes / (ms * 1e-3) / 1e9
    return {"message_size": message_size, "bucket_mb": message_size * 2 / (1 << 20), "buckets": len(buckets), "ms": ms,
            "algbw_gbs": algbw,
            "busbw_gbs": algbw * 2 * (world - 1) / world, "frac_of_900": algbw * 2 * (world - 1) / world / 900.0}


def resnet50_training(dev, rank, world, steps=10, warmup=6, batch=64):
    """BASELINE.json config 3 for real: torchvision ResNet-50 -> network_to_half -> apex_b200
    DistributedDataParallel (default message_size) -> FP16_Optimizer(SGD momentum, dynamic loss scale),
    synthetic 224x224 batches, the loop of examples/imagenet/main_fp16_optimizer.py:300-323
    (forward, loss, zero_grad, backward through the scaler, step).  The same loop WITHOUT the
    DistributedDataParallel wrapper gives the communication-free step on the same GPU; the difference
    is the communication that backward did not hide."""
    import torchvision
    import torch.distributed as dist
    from apex_b200.fp16_utils import FP16_Optimizer, network_to_half
    from apex_b200.parallel import DistributedDataParallel as DDP

    def run(distributed):
        torch.manual_seed(0)
        net = network_to_half(torchvision.models.resnet50().to(dev))
        model = DDP(net) if distributed else net
        opt = FP16_Optimizer(torch.optim.SGD(net.parameters(), **SGD), dynamic_loss_scale=True,
                             dynamic_loss_args={"init_scale": 2.0**10, "scale_window": 1000}, verbose=False)
        g = torch.Generator(device=dev).manual_seed(77 + rank)
        x = torch.randn(batch, 3, 224, 224, generator=g, device=dev)
        y = torch.randint(0, 1000, (batch,), generator=g, device=dev)
        crit = torch.nn.CrossEntropyLoss().to(dev)

        def it():
            loss = crit(model(x), y)
            opt.zero_grad()
            opt.backward(loss)
            opt.step()
            return loss
        for _ in range(warmup):
            it()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        skipped = 0
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            it()
            skipped += int(opt.overflow)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        info = {"step_ms": ms, "skipped_steps": skipped}
        if distributed:
            info["buckets"] = model.num_buckets
            info["grads_live_in_buckets"] = all(p.grad is v for k, ps in enumerate(model._bucket_params)
                                                for p, v in zip(ps, model._pipeline.slots[k][0].views))
        del model, net, opt
        torch.cuda.empty_cache()
        return info
    def reference_composition():
        """What the reference runs on these GPUs today, written out with stock torch calls (never used
        by our arm): per-bucket cat -> NCCL all_reduce -> /= world -> n x copy_ on a side stream from
        gradient hooks (apex/parallel/distributed.py:11-23, :243-315; buckets by the rule of :248-261 for
        reverse-order arrival), then per-tensor float(g.float().sum()) overflow test with a host sync
        each (loss_scaler.py:84-110), per-tensor copy_ into fp32 master gradients and mul_(1/scale)
        (fp16util.py:106-112, fp16_optimizer.py:236-241), torch.optim step, per-tensor copy_ back."""
        torch.manual_seed(0)
        net = network_to_half(torchvision.models.resnet50().to(dev))
        params = [p for p in net.parameters() if p.requires_grad]
        masters = [p.detach().clone().float().requires_grad_(True) if p.dtype == torch.float16 else p for p in params]
        opt = torch.optim.SGD(masters, **SGD)
        scale = 2.0**10
        layout = reference_bucket_rule([p.numel() for p in params], [p.dtype for p in params], 10000000)
        where = {i: b for b, members in enumerate(layout) for i in members}
        side = torch.cuda.Stream()
        st = {"next": 0, "ready": set(), "count": [0] * len(layout)}

        def reduce_bucket(k):
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                grads = [params[j].grad for j in layout[k]]
                flat = torch.cat([g.reshape(-1) for g in grads])
                dist.all_reduce(flat)
                flat /= world
                at = 0
                for g in grads:
                    g.copy_(flat[at:at + g.numel()].view_as(g))
                    at += g.numel()

        def hook_for(i):
            def hook(_p):
                b = where[i]
                st["count"][b] += 1
                if st["count"][b] == len(layout[b]):
                    st["ready"].add(b)
                    while st["next"] in st["ready"]:
                        reduce_bucket(st["next"])
                        st["next"] += 1
            return hook
        if world > 1:
            for i, p in enumerate(params):
                p.register_post_accumulate_grad_hook(hook_for(i))
        g = torch.Generator(device=dev).manual_seed(77 + rank)
        x = torch.randn(batch, 3, 224, 224, generator=g, device=dev)
        y = torch.randint(0, 1000, (batch,), generator=g, device=dev)
        crit = torch.nn.CrossEntropyLoss().to(dev)

        def it():
            loss = crit(net(x), y)
            for p in params:                               # FP16_Optimizer.zero_grad (:187-211)
                if p.grad is not None:
                    p.grad.detach_()
                    p.grad.zero_()
            st["next"], st["ready"], st["count"] = 0, set(), [0] * len(layout)
            (loss.float() * scale).backward()
            torch.cuda.current_stream().wait_stream(side)
            for p in params:                               # has_overflow: a host sync per tensor
                s_ = float(p.grad.float().sum())
                if s_ == float("inf") or s_ == -float("inf") or s_ != s_:
                    return
            for p, m in zip(params, masters):
                if m is not p:
                    if m.grad is None:
                        m.grad = torch.empty_like(m)
                    m.grad.copy_(p.grad)
            for m in masters:
                m.grad.mul_(1. / scale)
            opt.step()
            for p, m in zip(params, masters):
                if m is not p:
                    p.data.copy_(m.data)
        for _ in range(warmup):
            it()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            it()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        del net, opt, masters, params
        torch.cuda.empty_cache()
        return ms

    local = run(False)
    out = {"model": "torchvision resnet50, network_to_half (BatchNorm fp32), batch %d x 3 x 224 x 224 per GPU, "
                    "synthetic" % batch, "optimizer": "FP16_Optimizer(SGD momentum 0.9 wd 1e-4, dynamic loss scale)",
           "steps": steps, "warmup": warmup, "local_step_ms": local["step_ms"],
           "local_images_per_s": batch / (local["step_ms"] * 1e-3)}
    if world > 1:
        ddp = run(True)
        out.update({"ddp_step_ms": ddp["step_ms"], "images_per_s": world * batch / (ddp["step_ms"] * 1e-3),
                    "exposed_communication_ms": ddp["step_ms"] - local["step_ms"],
                    "weak_scaling_efficiency_vs_local": local["step_ms"] / ddp["step_ms"],
                    "buckets": ddp["buckets"], "grads_live_in_buckets": ddp["grads_live_in_buckets"],
                    "skipped_steps": ddp["skipped_steps"], "message_size": 10000000})
    else:
        out["images_per_s"] = out["local_images_per_s"]
        out["skipped_steps"] = local["skipped_steps"]
    ref_ms = reference_composition()
    ours_ms = out.get("ddp_step_ms", out["local_step_ms"])
    out["reference_composition_on_gpu"] = {
        "step_ms": ref_ms, "images_per_s": world * batch / (ref_ms * 1e-3), "speedup_of_ours": ref_ms / ours_ms,
        "what": "same model, batch and loop with the reference's composition of stock torch ops on these GPUs "
                "(hook-driven per-bucket cat / NCCL all_reduce / div / copy_ on a side stream, per-tensor overflow "
                "test with host syncs, copy_, mul_, torch.optim.SGD, copy_), static scale 2^10"}
    return out


_json_out = None


def protect_stdout():
    """Keep stdout for the ONE JSON line: anything else that writes to file descriptor 1 (NCCL prints
    its version banner there when NCCL_DEBUG is set) is sent to stderr instead."""
    global _json_out
    if _json_out is None:
        sys.stdout.flush()
        _json_out = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        sys.stdout = sys.stderr


def emit(line):
    out = _json_out if _json_out is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    protect_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="gpt2-xl", choices=sorted(WORKLOADS))
    ap.add_argument("--optimizer", default=None, choices=["adam", "sgd"])
    ap.add_argument("--bucket-mb", type=float, default=256.0, help="DDP bucket size for N > 1 (fp16 bytes)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true")
    ap.add_argument("--bucket-sweep", action="store_true",
                    help="N > 1: only time the bucket path of the workload's gradients for bucket sizes 1..256 MB")
    ap.add_argument("--no-side-runs", action="store_true",
                    help="skip the short runs of configs 2-4, the default bucket size and ResNet-50 training")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    shapes = WORKLOADS[args.workload][0]
    if args.optimizer is None:
        args.optimizer = WORKLOADS[args.workload][2]
    kind = args.optimizer
    n_params = sum(numel(s) for s in shapes)

    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # asked for N GPUs without a launcher: start one rank per GPU ourselves (the driver uses
        # `python -m torch.distributed.run ... bench.py --gpus N`, which lands in the code below)
        import socket
        sock = socket.socket()
        sock.bind(("127.0.0.1", 0))
        port = sock.getsockname()[1]
        sock.close()
        os.dup2(_json_out.fileno(), 1)            # give the real stdout back to the child ranks
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        os.execv(sys.executable, cmd)

    if args.impl == "reference":
        return run_reference_arm(args, shapes)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from apex_b200 import _C
    from apex_b200.fp16_utils import FP16_Optimizer
    from apex_b200.parallel.distributed import _BucketPipeline

    gen = torch.Generator(device=dev).manual_seed(1234)            # identical replicas on every rank
    fp32_mask = FP32_MASK.get(args.workload, [False] * len(shapes))
    params = [torch.nn.Parameter(torch.randn(s, generator=gen, device=dev) * 0.02 if f32 else
                                 (torch.randn(s, generator=gen, device=dev) * 0.02).half())
              for s, f32 in zip(shapes, fp32_mask)]
    opt = FP16_Optimizer(make_optimizer(kind)(params), dynamic_loss_scale=True,
                         dynamic_loss_args={"init_scale": 2.0**16, "scale_window": 1000}, verbose=False)
    ggen = torch.Generator(device=dev).manual_seed(99 + rank)      # every rank has its own gradients
    grads = [(torch.randn(s, generator=ggen, device=dev) * (1e-3 * 65536)).to(p.dtype) for s, p in zip(shapes, params)]
    for p, g in zip(params, grads):
        p.grad = g

    if args.bucket_sweep:
        if world < 2:
            raise SystemExit("--bucket-sweep needs --gpus >= 2")
        sweep = []
        for mb in (1, 2, 4, 8, 16, 32, 64, 128, 256):
            r = default_bucket_path(params, dev, world, rounds=3, warmup=2, message_size=mb * (1 << 20) // 2)
            r["nvlink_roofline"] = nvlink_roofline(2 * n_params, r["ms"], world)
            sweep.append(r)
            if rank == 0:
                sys.stderr.write("bucket %4d MB: %4d buckets %8.3f ms  busbw %6.1f GB/s\n"
                                 % (mb, r["buckets"], r["ms"], r["busbw_gbs"]))
        if rank == 0:
            emit({"metric": "bucket_allreduce_path_ms", "n_gpus": world, "workload": args.workload, "bytes": 2 * n_params,
                  "what": "whole bucket path (gradients living in the buckets: one kernel per bucket) per bucket size",
                  "sweep": sweep})
        dist.destroy_process_group()
        return 0

    side = torch.cuda.Stream()
    pipeline = _BucketPipeline(side)
    buckets = []
    if world > 1:
        # what DistributedDataParallel does once it has learned the structure: one flat home per
        # bucket (all in ONE symmetric allocation), and -- in the first round below -- every gradient
        # becomes a view of its bucket, so later rounds have nothing to gather or scatter
        bucket_elems = int(args.bucket_mb * (1 << 20) / 2)
        buckets = [[params[i] for i in b] for b in build_buckets(params, bucket_elems)]
        pipeline.plan([(k, [p.data for p in b]) for k, b in enumerate(buckets)])
    comm_events = []

    def reduce_buckets():
        """The per-bucket data path of DistributedDataParallel.comm_ready_buckets (side stream),
        then the join of its end-of-backward epilogue."""
        cur = torch.cuda.current_stream()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        pipeline.begin_round()                       # DistributedDataParallel.forward does this
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for k, b in enumerate(buckets):          # as comm_ready_buckets submits them
                pipeline.submit(k, [p.grad for p in b], verify=True, owners=b)
        pipeline.join(cur)
        e1.record()
        comm_events.append((e0, e1))

    host_s = [0.0]

    def step():
        t0 = time.perf_counter()
        if world > 1:
            reduce_buckets()
        opt.update_master_grads()
        opt.step()
        host_s[0] += time.perf_counter() - t0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t_check = Timed(_C, "mt_check_overflow_publish")
    t_fused = Timed(_C, FUSED_WRAPPER[kind])
    t_ar = Timed(_C, "symm_allreduce")

    for _ in range(args.warmup):
        step()
    barrier()
    del grads                                         # (N > 1: the gradients live in the buckets now)
    assert opt.overflow is False, "synthetic gradients must not overflow in the timed region"

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    t_check.on = t_fused.on = t_ar.on = True
    del comm_events[:]
    host_s[0] = 0.0
    launches0 = _C.launches
    barrier()
    start = torch.cuda.Event(enable_timing=True)
    stop = torch.cuda.Event(enable_timing=True)
    torch.cuda.nvtx.range_push("apex_b200_timed")      # ncu --nvtx --nvtx-include "apex_b200_timed/"
    start.record()
    for _ in range(args.steps):
        step()
    stop.record()
    barrier()
    torch.cuda.nvtx.range_pop()
    elapsed_ms = start.elapsed_time(stop)
    host_ms_per_step = host_s[0] / args.steps * 1e3      # python time inside step(), timed region only
    launches = _C.launches - launches0
    t_check.on = t_fused.on = t_ar.on = False
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([elapsed_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
    ms_per_step = elapsed_ms / args.steps
    value = world * n_params / (ms_per_step * 1e-3)

    fused_ms = t_fused.total_ms() / max(t_fused.count(), 1)
    check_ms = t_check.total_ms() / max(t_check.count(), 1)
    comm_ms = sum(a.elapsed_time(b) for a, b in comm_events) / max(len(comm_events), 1) if comm_events else None
    peak, peak_src = measured_peak()
    bpp = BYTES_PER_PARAM[kind]
    achieved = bpp * n_params / (fused_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(args.workload, {}).get(FUSED_KERNEL[kind])
    except Exception:   # noqa: BLE001
        pass
    roofline = {"bound": "hbm", "kernel": FUSED_KERNEL[kind], "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": bpp * n_params, "avg_launch_ms": fused_ms,
                "step_share": fused_ms / ms_per_step}

    # ---- end to end through the public API, gradients arriving from pinned HOST memory ----------
    e2e = None
    if not args.no_e2e:
        # The step's inputs arrive from pinned host memory in a FEW LARGE copies (one cudaMemcpyAsync per
        # contiguous span), not one per tensor: at N > 1 the gradients already live in the flat buckets
        # (one span per bucket); at N = 1 the user keeps them as views of one flat buffer, which is any
        # caller's choice of layout (fp32-native gradients, if any, keep their own small tensors).
        spans = []                                               # (device flat, pinned host flat)
        loose = []
        if world > 1:
            for k in range(len(buckets)):
                spans.append(pipeline.slots[k][0].flat)
        else:
            halves = [p for p in params if p.grad.dtype == torch.float16]
            loose = [p for p in params if p.grad.dtype != torch.float16]
            offs, at = [], 0
            for p in halves:
                offs.append(at)
                at += (p.numel() + 15) // 16 * 16                # 32-byte aligned views
            flat_grads = torch.zeros(at, dtype=torch.float16, device=dev)
            for p, o in zip(halves, offs):
                view = flat_grads[o:o + p.numel()].view(p.shape)
                view.copy_(p.grad)
                p.grad = view
            spans.append(flat_grads)
        host = [torch.empty(f.shape, dtype=f.dtype).pin_memory() for f in spans]
        for h, f in zip(host, spans):
            h.copy_(f)
        loose_host = [torch.empty(p.grad.shape, dtype=p.grad.dtype).pin_memory() for p in loose]
        for h, p in zip(loose_host, loose):
            h.copy_(p.grad)
        h2d = sum(h.numel() * h.element_size() for h in host + loose_host)

        def e2e_step():
            for f, h in zip(spans, host):
                f.copy_(h, non_blocking=True)                     # this step's inputs: host -> device
            for p, h in zip(loose, loose_host):
                p.grad.copy_(h, non_blocking=True)
            step()
            return opt.overflow                               # the step's result: device -> host (4 bytes)
        for _ in range(2):
            e2e_step()
        barrier()
        n_e2e = max(3, min(args.steps, 10))
        t0 = time.perf_counter()
        s0 = torch.cuda.Event(enable_timing=True)
        s1 = torch.cuda.Event(enable_timing=True)
        s0.record()
        for _ in range(n_e2e):
            e2e_step()
        s1.record()
        barrier()
        e2e_ms = max(s0.elapsed_time(s1), (time.perf_counter() - t0) * 1e3) / n_e2e
        if world > 1:
            t = torch.tensor([e2e_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms = float(t.item())
        e2e = {"value": world * n_params / (e2e_ms * 1e-3), "unit": "params/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms, "steps": n_e2e,
               "h2d_copies_per_step": len(host) + len(loose_host),
               "api": "gradients.copy_(pinned host), one copy per contiguous span ; "
                      "FP16_Optimizer.update_master_grads() ; .step() ; .overflow"}
        del host, spans

    parity = None
    if world > 1:
        members = build_buckets(params, bucket_elems)
        parity = multi_gpu_parity(pipeline, buckets, side, dev, rank, world)
        parity["bucket_layout_equals_reference_rule"] = members == reference_bucket_rule(
            [p.numel() for p in params], [p.dtype for p in params], bucket_elems)
        parity["pass"] = parity["pass"] and parity["bucket_layout_equals_reference_rule"]

    gpu_ref = None
    if world == 1 and not args.no_gpu_reference:
        try:
            ms = gpu_reference_composition(shapes, fp32_mask, kind, dev)
            gpu_ref = {"ms_per_step": ms, "params_per_s": n_params / (ms * 1e-3), "speedup": ms / ms_per_step,
                       "what": "the reference's composition of stock torch ops on this GPU (per-tensor "
                               ".float().sum() host syncs, copy_, mul_, torch.optim foreach step, copy_), 3 steps"}
        except torch.cuda.OutOfMemoryError:
            gpu_ref = {"unavailable": "out of memory for a second replica"}
    # ---- short side runs after the headline (own keys in the JSON line, never part of `value`) ----
    n_buckets = len(buckets)
    del buckets
    side_runs = {}
    if not args.no_side_runs and args.workload == "gpt2-xl":
        if world > 1:
            side_runs["bucket_path_at_reference_default_message_size"] = default_bucket_path(params, dev, world)
        for p in params:
            p.grad = None
        del opt, pipeline
        params = None
        torch.cuda.empty_cache()
        if world == 1:
            side_runs["configs"] = side_configs(dev)
        try:
            side_runs["resnet50_ddp_training"] = resnet50_training(dev, rank, world)
        except ImportError as err:
            side_runs["resnet50_ddp_training"] = {"unavailable": str(err)}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            # a bounded sample of the same workload (~10-30 s of CPU work): the whole gradient set costs
            # ~2 s per step on these hosts, so the leading ~0.36 G parameters (wte + wpe + 9 layers)
            v, info = cpu_reference_run(shapes, steps=8, warmup=2, budget_s=20.0, kind=kind, max_params=360_000_000,
                                        label=args.workload)
            cpu = {"value": v, "unit": "params/s", "cores": info["cores"], "kind": info["kind"],
                   "sample": info["sample"], "host_cpus": info["host_cpus"], "ms_per_step": info["ms_per_step"]}
        line = {
            "metric": "fused_mp_step_throughput", "value": value, "unit": "params/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload + ": " + WORKLOADS[args.workload][1] +
                       "; dynamic-loss-scale check + unscale + %s + fp16 copy" % kind, "optimizer": kind,
                       "params": n_params, "tensors": len(shapes), "parallelism": "dp%d" % world,
                       "bucket_mb": args.bucket_mb if world > 1 else None, "buckets": n_buckets,
                       "l2": "no flush: %.1f GB touched per step vs 126 MB L2" % ((bpp + 2) * n_params / 1e9)},
            "roofline": roofline, "cpu_baseline": cpu, "reference_composition_on_gpu": gpu_ref, "e2e": e2e,
            "gpu_launches": launches, "clocks": clocks, "parity": parity, "side_runs": side_runs,
            "breakdown_ms": {"check_overflow": check_ms, "fused_update": fused_ms, "bucket_allreduce_path": comm_ms,
                             # what the step spends outside these three (launch gaps, event records)
                             "gaps": ms_per_step - check_ms - fused_ms - (comm_ms or 0.0),
                             # wall time inside step() on the host: python + the wait for the overflow word
                             # (which arrives with the check / the last bucket), NOT pure python time
                             "host_time_in_step_incl_wait_for_flag": host_ms_per_step},
            "hbm_gbs_step": (bpp + 2) * n_params / (ms_per_step * 1e-3) / 1e9 if world == 1 else None,
        }
        if comm_ms:
            s_bytes = 2 * n_params
            algbw = s_bytes / (comm_ms * 1e-3) / 1e9
            from apex_b200.parallel import symm as _symm
            kernel = None
            if t_ar.count():
                k_ms = t_ar.total_ms() / args.steps          # our allreduce kernels of one step, summed
                k_alg = s_bytes / (k_ms * 1e-3) / 1e9
                kernel = {"ms_per_step": k_ms, "launches_per_step": t_ar.count() / args.steps, "algbw_gbs": k_alg,
                          "busbw_gbs": k_alg * 2 * (world - 1) / world,
                          "frac_of_900": k_alg * 2 * (world - 1) / world / 900.0,
                          "what": "sum of the durations of the apex_b200_symm_allreduce launches of one step (CUDA "
                                  "events on the side stream they run on, back to back); each launch includes its "
                                  "two in-kernel cross-rank barriers, i.e. also the time it waits for the slowest rank"}
            line["allreduce"] = {"bytes": s_bytes, "ms": comm_ms, "algbw_gbs": algbw,
                                 "transport": "apex_b200 symmetric-memory two-shot, one kernel per bucket: barrier, sum "
                                              "(NVSwitch multimem beyond 2 ranks, peer-to-peer at 2), x 1/N, inf/NaN "
                                              "check, barrier" if _symm.enabled() else "NCCL",
                                 "kernel": kernel,
                                 "nvlink_roofline": nvlink_roofline(s_bytes, comm_ms, world),
                                 "busbw_gbs": algbw * 2 * (world - 1) / world, "nvlink_peak_gbs_per_dir": 900.0,
                                 "frac_of_900": algbw * 2 * (world - 1) / world / 900.0,
                                 "includes": "whole bucket path from the end of backward to averaged, checked gradients. "
                                             "The gradients live in the buckets (param.grad are views of the flat "
                                             "symmetric buffers, as DistributedDataParallel arranges after its first "
                                             "overlapped backward), so there is no gather and no scatter; x 1/N and "
                                             "the overflow check ride in the allreduce kernel"}
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity["pass"]:
        sys.stderr.write("bench.py: multi-GPU parity FAILED: %r\n" % (parity,))
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
