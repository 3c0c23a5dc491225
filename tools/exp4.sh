#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out/exp4; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_codec.py tests/test_gpu_api.py tests/test_gpu_models.py tests/test_gpu_fast_paths.py -m gpu -x -q > $O/pytest.txt 2>&1; tail -3 $O/pytest.txt
show() { python -c "
import json; r=json.load(open('$1')); print('$2 main ms', round(r['ms_per_step'],4), 'frac', round(r['roofline']['frac'],4), 'e2e', round(r['e2e']['value'],1))"; }
timeout 400 python bench.py --no-extras --no-cpu --steps 64 --warmup 8 > $O/main.json 2> $O/main.err; show $O/main.json incta_min5
RWKV_B200_KSPLIT_MIN_CHUNKS=2 timeout 400 python bench.py --no-extras --no-cpu --steps 64 --warmup 8 > $O/main_min2.json 2> $O/main_min2.err; show $O/main_min2.json incta_min2
RWKV_B200_KSPLIT_MIN_CHUNKS=1 timeout 400 python bench.py --no-extras --no-cpu --steps 64 --warmup 8 > $O/main_min1.json 2> $O/main_min1.err; show $O/main_min1.json incta_min1
RWKV_B200_KSPLIT_CLUSTER_ONLY=1 timeout 400 python bench.py --no-extras --no-cpu --steps 64 --warmup 8 > $O/main_cl.json 2> $O/main_cl.err; show $O/main_cl.json cluster_min5
timeout 200 python bench.py --workload matvec --no-cpu > $O/matvec.json 2> $O/matvec.err; python -c "
import json; r=json.load(open('$O/matvec.json')); print(' '.join(f\"{x['type']}:{x['M']}x{x['K']}={x['frac_measured']:.3f}\" for x in r['config']['rows']))"
RWKV_B200_LIB=rwkv_cppy_b200/build/variants/librwkv_trace.so timeout 300 python tools/decode_trace.py v4-14b Q4_1 > $O/trace14b.txt 2>&1; sed -n 1,40p $O/trace14b.txt | cut -c1-200
RWKV_B200_KSPLIT_MIN_CHUNKS=2 RWKV_B200_LIB=rwkv_cppy_b200/build/variants/librwkv_trace.so timeout 300 python tools/decode_trace.py v4-14b Q4_1 > $O/trace14b_min2.txt 2>&1; sed -n 1,40p $O/trace14b_min2.txt | cut -c1-200
