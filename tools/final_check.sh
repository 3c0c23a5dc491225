#!/bin/bash
# round-end self-check on one B200: GPU tests, smoke, the default bench line, and the ncu launch list (+ DRAM bytes) of the headline workload
cd $GRAFT_REPO_ROOT 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out/final; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; tail -3 $O/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke > $O/smoke.txt 2>&1; tail -2 $O/smoke.txt
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; tail -c 600 $O/bench.err; python - <<PY
import json
r = json.load(open("$O/bench.json"))
print("main ms", round(r["ms_per_step"], 4), "tok/s", round(r["value"], 1), "frac", round(r["roofline"]["frac"], 4), "e2e", round(r["e2e"]["value"], 1), "cpu", r.get("cpu_baseline", {}).get("value"), "parity", r.get("cpu_baseline", {}).get("parity_max_rel"))
for k, v in r.get("extras", {}).items(): print(" ", k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a not in ("config", "rows")})
if "matvec" in r.get("extras", {}): print("  matvec:", " ".join(f"{x['type']}:{x['M']}x{x['K']}={x['frac_measured']:.3f}" for x in r["extras"]["matvec"]["rows"]))
PY
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'matvec_q_kernel|matvec_kernel|ln_mix_kernel|wkv4_kernel|embed_ln0|ln_act' -c 1500 --csv --log-file $O/main_launches_ncu.csv python bench.py --no-extras --no-cpu --steps 4 --warmup 3 > $O/ncu_run.log 2>&1; tail -2 $O/ncu_run.log | cut -c1-300; wc -l $O/main_launches_ncu.csv
