"""Worker of tests/test_dist_cpu.py: the N > 1 host logic of bench.py on the gloo backend
(no GPU): rendezvous, barrier, max / sum over ranks, object broadcast (how the NCCL unique id
travels), and the stream / shard partitioning rules."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from rwkv_cppy_b200 import shards


def main():
    d = bench.Dist(use_cuda=False)
    assert d.world == int(os.environ["WORLD_SIZE"]) and d.rank == int(os.environ["RANK"])
    d.barrier()
    mx = d.max(10.0 + d.rank)            # per-rank elapsed time -> the slowest rank decides
    sm = d.sum(float(len(shards.streams_of_rank(257, d.rank, d.world))))
    ids = [bytes(range(128)) if d.rank == 0 else None]
    d.dist.broadcast_object_list(ids, src=0)
    r0, cnt = shards.row_shard(50277, d.rank, d.world)
    c0, ccnt = shards.col_shard(5120, d.rank, d.world)
    rows = d.sum(float(cnt)); cols = d.sum(float(ccnt))
    if d.rank == 0:
        print("DISTRESULT " + json.dumps({"max": mx, "streams": sm, "id_ok": ids[0] == bytes(range(128)), "rows": rows, "cols": cols, "world": d.world}), flush=True)
    else:
        assert ids[0] == bytes(range(128))
    d.barrier()
    d.close()


if __name__ == "__main__":
    main()
