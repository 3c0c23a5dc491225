#!/bin/bash
# One-call timing experiment over library variants (tools/build_variant.sh) on the GPU box:
#   tools/exp_matvec.sh  -> gpurun_out/exp/<variant>[_deep].{matvec,main,decode,pytest}.txt
cd "$(dirname "$0")/.."
O=gpurun_out/exp; mkdir -p $O
V=rwkv_cppy_b200/build/variants
cp rwkv_cppy_b200/librwkv.so $O/../librwkv_default.so.bak
run() {   # name lib deep what...
    local name=$1 lib=$2 deep=$3; shift 3
    cp $V/librwkv_$lib.so rwkv_cppy_b200/librwkv.so
    export RWKV_B200_GA_DEEP=$deep
    for what in "$@"; do
        case $what in
            pytest) timeout 300 python -m pytest tests/test_gpu_codec.py tests/test_gpu_api.py -m gpu -x -q > $O/$name.pytest.txt 2>&1; tail -2 $O/$name.pytest.txt ;;
            matvec) timeout 200 python bench.py --workload matvec --no-cpu > $O/$name.matvec.json 2> $O/$name.matvec.err; python - <<PY
import json
try:
    r = json.load(open("$O/$name.matvec.json"))
    print("$name matvec:", " ".join(f"{x['type']}:{x['M']}x{x['K']}={x['frac_measured']:.3f}" for x in r["config"]["rows"]))
except Exception as e: print("$name matvec failed", e)
PY
            ;;
            main) timeout 400 python bench.py --no-extras --no-cpu --steps 64 --warmup 8 > $O/$name.main.json 2> $O/$name.main.err; python - <<PY
import json
try:
    r = json.load(open("$O/$name.main.json")); print("$name main: ms_per_step", round(r["ms_per_step"], 4), "frac", round(r["roofline"]["frac"], 4), "e2e", round(r["e2e"]["value"], 1))
except Exception as e: print("$name main failed", e)
PY
            ;;
            decode) timeout 200 python bench.py --workload decode --no-cpu > $O/$name.decode.json 2> $O/$name.decode.err; python - <<PY
import json
try:
    r = json.load(open("$O/$name.decode.json")); print("$name decode169m: ms_per_step", round(r["ms_per_step"], 5), "e2e", round(r["e2e"]["value"], 1))
except Exception as e: print("$name decode failed", e)
PY
            ;;
        esac
    done
}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
run HEAD   HEAD 0 matvec main decode
run A      A    0 matvec main
run B      B    0 pytest matvec main decode
run B_deep B    1 pytest matvec main
run A_deep A    1 main
export RWKV_B200_KSPLIT_MIN_CHUNKS=2
run B_deep_ks2 B 1 main
unset RWKV_B200_KSPLIT_MIN_CHUNKS
cp $O/../librwkv_default.so.bak rwkv_cppy_b200/librwkv.so
