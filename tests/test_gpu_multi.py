"""Multi-GPU parity under pytest: launches tests/multi_gpu_check.py with torch.distributed.run on every GPU of the
box (one process per GPU, NCCL) and checks its verdict: row-sharded SASO and dense sketches -- through torch.distributed and
through the C ABI's own NCCL communicator (rbx_comm_init, rbx_shard_sketch_*) -- and the column-sharded form, all against the
one-GPU result. Skips on a box with fewer than two GPUs (the host-side logic is covered by tests/test_sharded_cpu.py on gloo)."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharded_sketches_match_one_gpu():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    n = min(n, 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    out = json.loads(line)
    assert out["world"] == n
    for kind in ("sparse", "dense"):
        assert max(v for k, v in out[kind].items() if k.endswith("relerr")) < 1e-11, out
