#!/bin/bash
# tensor-parallel check on an N-GPU box: tools/tp_check.sh N  (TP tests of that world size with the peer-memory collectives + the main bench line)
N=${1:-2}
cd $GRAFT_REPO_ROOT 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out/tp$N; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "p2p-$N" > $O/pytest_tp.txt 2>&1; tail -2 $O/pytest_tp.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 64 --warmup 8 --no-extras > $O/bench.json 2> $O/bench.err
python -c "
import json
for l in open('$O/bench.json'):
    if l.startswith('{'):
        r=json.loads(l); print('tp$N ms', round(r['ms_per_step'],4), 'tok/s', round(r['value'],1), 'e2e', round(r['e2e']['value'],1), 'frac', round(r['roofline']['frac'],3))"
