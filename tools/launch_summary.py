"""Summarises an ncu launch list (`ncu --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --print-units base --csv`):
per kernel name launches, total and mean microseconds, share and -- when captured -- DRAM bytes.
    python tools/launch_summary.py launches.csv [skip] [--json profiles/main_kernel_split.json --source "..."]
With --json the decode steps in the list (delimited by the embedding kernel) are averaged into the record bench.py prints as
roofline.kernel_split, and the DRAM bytes of one step go to profiles/traffic.json as "main_step"."""
import collections, csv, json, os, re, sys

args = [a for a in sys.argv[1:] if not a.startswith("--")]
opts = {sys.argv[i][2:]: sys.argv[i + 1] for i in range(1, len(sys.argv) - 1) if sys.argv[i].startswith("--")}
args = [a for a in args if a not in opts.values()]
rows = list(csv.reader(open(args[0], errors="ignore")))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
H = rows[hdr]; kn, mv, mn, idc = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Name"), H.index("ID")
skip = int(args[1]) if len(args) > 1 else 0
launches = collections.OrderedDict()                       # id -> [name, us, dram bytes]
for r in rows[hdr + 1:]:
    if len(r) <= mv: continue
    name = re.sub(r"\(.*", "", r[kn]); name = re.sub(r"^void ", "", name).replace("rb::", "")
    L = launches.setdefault(r[idc], [name, 0.0, 0.0])
    val = float(r[mv].replace(",", ""))
    if r[mn] == "gpu__time_duration.sum": L[1] = val / 1000.0
    elif r[mn] in ("dram__bytes_read.sum", "dram__bytes_write.sum"): L[2] += val
seq = list(launches.values())[skip:]
agg = collections.OrderedDict(); total = 0.0
for name, us, db in seq:
    a = agg.setdefault(name, [0, 0.0, 0.0]); a[0] += 1; a[1] += us; a[2] += db; total += us
print(f"{'kernel':90s} {'n':>6s} {'total us':>10s} {'mean us':>8s} {'share':>6s} {'DRAM MB/launch':>15s}")
for k, (c, t, db) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:90]:90s} {c:6d} {t:10.1f} {t / c:8.2f} {100 * t / total:5.1f}% {db / c / 1e6:15.2f}")
print(f"{'total':90s} {len(seq):6d} {total:10.1f}")

if "json" in opts:
    # decode steps: from one embedding kernel to the next
    starts = [i for i, (n, _, _) in enumerate(seq) if n.startswith("embed_ln0")]
    steps = [seq[a:b] for a, b in zip(starts[:-1], starts[1:])]
    steps = [s for s in steps if len(s) == max(len(x) for x in steps)]          # full steps with logits only
    assert steps, "no complete decode step in the list"

    def cls(n):
        if n.startswith("matvec_q_kernel"): return "matvec_q_kernel (quantised mat-vec: att k/v/r, att output, ffn key+receptance, ffn value)"
        if n.startswith("matvec_kernel"): return "matvec_kernel<F16> (head)"
        if n.startswith("ln_mix_kernel"): return "ln_mix_kernel (layer norm + token shift + mixes, q8 blocks)"
        if n.startswith("wkv4_kernel"): return "wkv4_kernel"
        return "ln_out + embedding"
    share = collections.OrderedDict(); tot = 0.0; dram = 0.0
    for s in steps:
        for n, us, db in s:
            share[cls(n)] = share.get(cls(n), 0.0) + us; tot += us; dram += db
    rec = {"source": opts.get("source", args[0]), "launches_per_step": len(steps[0]), "steps_averaged": len(steps),
           "us_per_step_under_ncu": tot / len(steps),
           "share": {k: round(v / tot, 4) for k, v in sorted(share.items(), key=lambda kv: -kv[1])}}
    if dram > 0: rec["dram_bytes_per_step"] = dram / len(steps)
    json.dump(rec, open(opts["json"], "w"), indent=1)
    print("wrote", opts["json"], rec)
    if dram > 0:
        tp = os.path.join(os.path.dirname(opts["json"]), "traffic.json")
        t = json.load(open(tp)) if os.path.exists(tp) else {}
        t["main_step"] = {"traffic_bytes": dram / len(steps), "algorithmic_bytes": 9051811840, "source": opts.get("source", args[0]),
                          "note": "sum of dram__bytes_read.sum + dram__bytes_write.sum over the launches of one decode step (14B-shape Q4_1, one GPU)"}
        json.dump(t, open(tp, "w"), indent=2)
