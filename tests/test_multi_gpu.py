"""GPU (>= 2 devices): the view-parallel step on hardware -- both exchanges against the sum of single-view gradients
(tests/multi_gpu_check.py under torchrun).  Skipped on a single-GPU box; the round's 2-GPU run is committed as
profiles/r02_multi_gpu_check.log."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("views", [None, 4])
def test_view_parallel_step_matches_single_gpu_sum(views):
    n = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    if views:
        cmd += ["--views", str(views)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTI_GPU_CHECK OK" in r.stdout, (r.stdout[-2000:], r.stderr[-2000:])
