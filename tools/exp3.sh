#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out/exp3; mkdir -p $O
timeout 400 python -m pytest tests/test_gpu_codec.py tests/test_gpu_api.py tests/test_gpu_models.py tests/test_gpu_fast_paths.py -m gpu -x -q > $O/pytest.txt 2>&1; tail -3 $O/pytest.txt
timeout 400 python bench.py --no-extras --no-cpu --steps 64 --warmup 8 > $O/main.json 2> $O/main.err
python -c "
import json; r=json.load(open('$O/main.json')); print('main ms', round(r['ms_per_step'],4), 'frac', round(r['roofline']['frac'],4), 'e2e', round(r['e2e']['value'],1))"
RWKV_B200_LIB=rwkv_cppy_b200/build/variants/librwkv_trace.so timeout 300 python tools/decode_trace.py v4-14b Q4_1 > $O/trace14b.txt 2>&1; tail -48 $O/trace14b.txt | cut -c1-200
