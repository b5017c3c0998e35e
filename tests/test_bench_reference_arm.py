"""`bench.py --impl reference` (the arm the driver runs beside ours: the reference's CPU BVHAggregate / PathIntegrator
from oracle/_ref on the host cores, no GPU in the process) at a size that takes seconds: it must print one JSON line
with the contract's keys.  Guards the arm against signature drift in the functions it shares with the GPU arm."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "raycast_ref")), reason="oracle/_ref not built")
def test_reference_arm_prints_the_contract_line():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--tris", "20000", "--rays", "65536", "--cpu-sample", "65536", "--render-cpu-spp", "1"],
                       capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "traversal_mrays_per_s" and line["value"] > 0
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["cores"] >= 1
    assert line["render"]["value"] > 0 and line["gpu_launches"] == 0
