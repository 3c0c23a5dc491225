#!/bin/bash
cd "$(dirname "$0")/.." 2>/dev/null || true
cd $GRAFT_REPO_ROOT
O=gpurun_out/tp2; mkdir -p $O
nvidia-smi --query-gpu=name --format=csv,noheader | head -3
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "p2p or nccl" > $O/pytest_tp.txt 2>&1; tail -3 $O/pytest_tp.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
timeout 600 $TR bench.py --gpus 2 --steps 64 --warmup 8 --no-extras > $O/bench_tp2.json 2> $O/bench_tp2.err; python -c "
import json
for l in open('$O/bench_tp2.json'):
    if l.startswith('{'):
        r=json.loads(l); print('tp2 one-shot ms', round(r['ms_per_step'],4), 'tok/s', round(r['value'],1), 'e2e', round(r['e2e']['value'],1))"
RWKV_B200_LIB=rwkv_cppy_b200/build/variants/librwkv_trace.so timeout 300 $TR tools/decode_trace.py v4-14b Q4_1 > $O/trace_tp2.txt 2> $O/trace_tp2.err; sed -n 1,45p $O/trace_tp2.txt | cut -c1-200; tail -4 $O/trace_tp2.txt | cut -c1-250
