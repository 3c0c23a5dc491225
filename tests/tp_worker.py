"""Worker of tests/test_gpu_tp.py and of the TP parity check: run under torchrun, one rank per GPU.
Every rank loads its shard of the same model, evaluates a prompt token by token and as a sequence,
and rank 0 compares with a single-GPU evaluation of the whole model."""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import rwkv_cppy_b200
from rwkv_cppy_b200 import synth


def main():
    shape, fmt = sys.argv[1], sys.argv[2]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = rwkv_cppy_b200.load_rwkv_shared_library()
    lib.library.rwkv_b200_set_device(local)
    if rank == 0:
        synth.ensure_model(lib, shape, fmt)
    dist.barrier()
    path = synth.ensure_model(lib, shape, fmt)
    ids = [lib.rwkv_b200_tp_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    model = rwkv_cppy_b200.RWKVModel(lib, path, tensor_parallel=(rank, world, ids[0]))
    if os.environ.get("RWKV_B200_TP_NCCL_ONLY") != "1":          # one-shot collectives over peer memory
        handles = [None] * world
        dist.all_gather_object(handles, lib.rwkv_b200_tp_ipc_handle(model._ctx))
        lib.rwkv_b200_tp_connect(model._ctx, handles)
    toks = [int(t) for t in np.random.default_rng(3).integers(0, model.n_vocab, size=24)]
    state, logits = None, None
    for t in toks:
        logits, state = model.eval(t, state, use_numpy=True)
    lseq, sseq = model.eval_sequence(toks, None, use_numpy=True)
    out = {"ok": True}
    # every rank must hold the same logits bit for bit (the replicated layer norms stay in lockstep)
    mine = torch.from_numpy(logits.copy()).cuda()
    ref = mine.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([1 if torch.equal(mine, ref) else 0], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    if rank == 0:
        out["ranks_identical"] = bool(same.item())
        out["state_finite"] = bool(np.isfinite(state).all())
        full = rwkv_cppy_b200.RWKVModel(lib, path)
        s1, l1 = None, None
        for t in toks:
            l1, s1 = full.eval(t, s1, use_numpy=True)
        scale = max(1.0, float(np.abs(l1).max()))
        out["max_dlogit_rel"] = float(np.abs(logits - l1).max() / scale)
        out["max_dlogit_seq_rel"] = float(np.abs(lseq - l1).max() / scale)
        # shift states are replicated: identical layout; wkv slices of rank 0 are the first D/world channels
        D, L = model.n_embed, model.n_layer
        st = state.reshape(L, 5, D); s1r = s1.reshape(L, 5, D)
        out["max_dshift"] = float(np.abs(st[:, :2] - s1r[:, :2]).max())
        Dl = D // world
        out["max_dwkv_own"] = float(np.abs(st[:, 2:4, :Dl] - s1r[:, 2:4, :Dl]).max())
        out["finite"] = bool(np.isfinite(logits).all())
        print("TPRESULT " + json.dumps(out), flush=True)
        full.free()
    model.free()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
