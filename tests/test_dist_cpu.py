"""CPU (gloo, world_size 2) coverage of the multi-process host logic: bench.Dist and the
partitioning rules shared with the C++ loader."""
import json
import os
import subprocess
import sys

import pytest

from rwkv_cppy_b200 import shards

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_rules():
    for world in (1, 2, 4, 8):
        got = sorted(s for r in range(world) for s in shards.streams_of_rank(256, r, world))
        assert got == list(range(256))
        assert all(len(shards.streams_of_rank(256, r, world)) == 256 // world for r in range(world))
        for n in (50277, 65536, 5120, 7):
            spans = [shards.row_shard(n, r, world) for r in range(world)]
            assert sum(c for _, c in spans) == n
            pos = 0
            for r0, c in spans:
                assert r0 == min(n, pos) or c == 0
                pos += c
        spans = [shards.col_shard(20480, r, world) for r in range(world)]
        assert [s[0] for s in spans] == [20480 // world * r for r in range(world)] and all(s[1] % 32 == 0 for s in spans)
    with pytest.raises(ValueError):
        shards.col_shard(768, 0, 5)
    assert shards.tp_supported(5120, 20480, 8) and not shards.tp_supported(768, 3072, 5)


def test_two_rank_gloo_host_logic():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "dist_worker.py")]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-3000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("DISTRESULT ")][-1]
    r = json.loads(line[len("DISTRESULT "):])
    assert r == {"max": 11.0, "streams": 257.0, "id_ok": True, "rows": 50277.0, "cols": 5120.0, "world": 2}
