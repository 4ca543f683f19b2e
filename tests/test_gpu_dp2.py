"""Data-parallel exchange on real NCCL ranks: needs >= 2 GPUs, skipped otherwise.
The worker (tests/dp2_worker.py) runs under torchrun; see its docstring for what is compared."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_dp2_equals_single_process_on_concatenated_batch_nccl(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under `gpurun --gpus 2`)")
    out = tmp_path / "dp2.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29671", os.path.join(ROOT, "tests", "dp2_worker.py"), str(out)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=540)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.load(open(out))
    print(json.dumps(res))
    for name, row in res.items():
        assert row["eager_rel"] < row["tol"], (name, row)
        assert row["graph_rel"] < row["tol"], (name, row)
        assert row["fired_in_backward"] >= row["buckets"] - 1, (name, row)   # overlap: buckets fire inside backward
        assert row["replicas_bit_identical"], (name, row)
    assert res["fp32"]["graph_vs_eager_rel"] < 1e-6      # deterministic reductions: graph replay == eager
