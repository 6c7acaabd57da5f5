"""bench.py's CPU-only arm (`--impl reference`) prints one well-formed JSON line."""
import json
import os
import subprocess
import sys

from tests.conftest import ROOT


import pytest


@pytest.mark.parametrize("force_port", [True, False])
def test_reference_arm_prints_one_json_line(force_port):
    """With the reference importable (the build container) the arm times the unmodified reference under
    the Tensor.type shim; without it (the GPU box; forced here by APEX_REFERENCE=none) the oracle port."""
    env = dict(os.environ)
    if force_port:
        env["APEX_REFERENCE"] = "none"
    out = subprocess.check_output([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                                   "--workload", "mlp", "--steps", "3", "--warmup", "1"], env=env,
                                  stderr=subprocess.DEVNULL, timeout=300).decode().strip().splitlines()
    assert len(out) == 1
    line = json.loads(out[0])
    assert line["impl"] == "reference" and line["metric"] == "fused_mp_step_throughput"
    assert line["unit"] == "params/s" and line["value"] > 0 and line["higher_is_better"] is True
    have_ref = os.path.exists("/root/reference/apex/__init__.py") and not force_port
    assert line["cpu_baseline"]["kind"] == ("reference" if have_ref else "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["config"]["params"] == 8393728 and line["config"]["tensors"] == 4       # the whole of config 1
    assert line["config1_mlp"]["params"] == 8393728
    assert line["cpu_baseline"]["value"] == line["value"] == line["e2e"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "mlp" in line["config"]["workload"] and line["gpu_launches"] == 0


def test_workload_shapes_match_the_survey():
    sys.path.insert(0, ROOT)
    import bench
    xl = bench.WORKLOADS["gpt2-xl"][0]
    assert len(xl) == 580 and sum(bench.numel(s) for s in xl) == 1557611200
    med = bench.WORKLOADS["gpt2-medium"][0]
    assert len(med) == 292 and sum(bench.numel(s) for s in med) == 354823168
    rn = bench.resnet50_shapes()
    assert len(rn) == 161 and sum(bench.numel(s) for s, _ in rn) == 25557032
    assert sum(1 for _, f in rn if f) == 106 and sum(bench.numel(s) for s, f in rn if f) == 53120
    assert sum(bench.numel(s) for s in bench.WORKLOADS["mlp"][0]) == 8393728
    lstm = bench.lstm_shapes()
    assert len(lstm) == 11 and sum(bench.numel(s) for s in lstm) == 135891278
    # the bucket rule restated in bench.py agrees with the product's layout class
    import torch
    ps = [torch.nn.Parameter(torch.empty(s, dtype=torch.float16)) for s in med[:40]]
    assert bench.build_buckets(ps, 3000000) == bench.reference_bucket_rule([p.numel() for p in ps],
                                                                            [p.dtype for p in ps], 3000000)
