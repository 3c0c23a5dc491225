#!/bin/bash
# experiment 2: argument copies in shared memory + self-prefetch depth
cd "$(dirname "$0")/.."
O=gpurun_out/exp2; mkdir -p $O
show_main() { python - "$1" "$2" <<'PY'
import json,sys
try:
    r = json.load(open(sys.argv[2])); print(sys.argv[1], "main: ms_per_step", round(r["ms_per_step"], 4), "frac", round(r["roofline"]["frac"], 4), "e2e", round(r["e2e"]["value"], 1))
except Exception as e: print(sys.argv[1], "main failed", e)
PY
}
timeout 300 python -m pytest tests/test_gpu_codec.py tests/test_gpu_api.py tests/test_gpu_models.py -m gpu -x -q > $O/pytest.txt 2>&1; tail -3 $O/pytest.txt
for pf in 0 2 16; do
    RWKV_B200_SELF_PF=$pf timeout 400 python bench.py --no-extras --no-cpu --steps 64 --warmup 8 > $O/main_pf$pf.json 2> $O/main_pf$pf.err; show_main pf$pf $O/main_pf$pf.json
done
timeout 200 python bench.py --workload decode --no-cpu > $O/decode.json 2> $O/decode.err; python -c "
import json; r=json.load(open('$O/decode.json')); print('decode169m ms', round(r['ms_per_step'],5), 'e2e', round(r['e2e']['value'],1))"
RWKV_B200_LIB=rwkv_cppy_b200/build/variants/librwkv_trace.so timeout 300 python tools/decode_trace.py v4-14b Q4_1 > $O/trace14b.txt 2>&1; tail -42 $O/trace14b.txt | cut -c1-200
RWKV_B200_LIB=rwkv_cppy_b200/build/variants/librwkv_trace.so RWKV_B200_SELF_PF=2 timeout 300 python tools/decode_trace.py v4-14b Q4_1 > $O/trace14b_pf2.txt 2>&1; grep -A12 "^medians" $O/trace14b_pf2.txt | cut -c1-200
