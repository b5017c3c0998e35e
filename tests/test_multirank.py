"""World-size-2 run of the sharding logic on the CPU (gloo): tests/multirank_worker.py."""
import os
import subprocess
import sys

from conftest import ROOT


def test_shards_and_stripes_world2(P, hostcheck_bin, rendercheck):
    assert [P.parallel.ray_shard(10, r, 3) for r in range(3)] == [(0, 4), (4, 7), (7, 10)]
    assert P.parallel.sample_stripe(64, 3, 8) == (3, 8, 8)
    assert P.parallel.sample_stripe(5, 1, 2) == (1, 2, 2) and P.parallel.sample_stripe(1, 1, 2) == (1, 0, 2)
    rclib = os.path.join(ROOT, "tests", "native", "_build", "librendercheck.so")
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731",
                        os.path.join(ROOT, "tests", "multirank_worker.py"), hostcheck_bin, rclib],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "MULTIRANK_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
