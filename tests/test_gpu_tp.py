"""Tensor-parallel decode (2 GPUs, NCCL): the sharded model must reproduce the single-GPU logits
within summation-order tolerance.  Needs >= 2 GPUs; skipped otherwise."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("collective", ["fused", "p2p", "nccl"])
@pytest.mark.parametrize("fmt", ["FP16", "FP16-to-Q4_1"])
def test_tp_matches_single_gpu(fmt, collective, world):
    """collective: "fused" = partial sums pushed from the mat-vec epilogues into the peers' inboxes and gathered by the next layer
    norm (no collective launch; quantised decode), "p2p" = one-shot all-reduce kernel over peer memory, "nccl" = ncclAllReduce."""
    if n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, RWKV_B200_TP_NCCL_ONLY="1" if collective == "nccl" else "0", RWKV_B200_TP_FUSE="1" if collective == "fused" else "0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "tests", "tp_worker.py"), "v4-mini", fmt]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("TPRESULT ")][-1]
    r = json.loads(line[len("TPRESULT "):])
    assert r["finite"] and r["state_finite"]
    # same arithmetic per shard; only the all-reduce changes the order of the partial sums
    tol = 2e-2 if "Q" in fmt else 2e-3
    assert r["max_dlogit_rel"] <= tol and r["max_dlogit_seq_rel"] <= tol, r
    assert r["max_dshift"] <= tol * 10 and r["max_dwkv_own"] <= tol * 10, r
    assert r["ranks_identical"], r
