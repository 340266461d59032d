"""The sharded CUDA run equals the unsharded one: under torch.distributed.run with 2 ranks, bench.py
gives every rank a contiguous frame range of ONE clip, exchanges the boundary descriptor block over NCCL, and checks
per pair (match count, order-sensitive checksum of the match rows) against a 1-rank pass over the same clip -- the
reference's loop is sequential (tracker.py:214: frame i against frame i-1).  Needs 2 visible GPUs; with one it
checks the single-rank form of the same assertion (chunked shards of one process)."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _run(n):
    cmd = [sys.executable]
    if n > 1:
        cmd += ["-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
                "--master-port", "29731"]
    cmd += [os.path.join(ROOT, "bench.py"), "--gpus", str(n), "--steps", "2", "--warmup", "1", "--batch", "16", "--no-cpu"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-3000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, out.stdout[-2000:]
    return json.loads(lines[0])


def test_sharded_run_equals_unsharded():
    from orbslam2 import _native
    n = 2 if _native.lib().vslam_device_count() >= 2 else 1
    line = _run(n)
    assert line["n_gpus"] == n
    assert line["sharded_equals_unsharded"] is True, line.get("sharded_check")
    assert line["value"] > 0 and line["e2e"]["value"] > 0 and line["e2e"]["copy_ceiling_fps"] > 0
    assert line["gpu_launches"] > 0
