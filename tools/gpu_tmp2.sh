W=n2_pbe_cfg4b; TAG=r02z
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches_$W.csv python bench.py --workload $W --steps 2 --warmup 1 --kernels-only > gpurun_out/${TAG}_ncu_launches_$W.log 2>&1
python - <<PY
import csv, collections
rows = list(csv.reader(l for l in open("gpurun_out/${TAG}_launches_$W.csv") if l.startswith('"')))
h = rows[0]; kn = h.index("Kernel Name"); mv = h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    try: agg.setdefault(r[kn].split("(")[0], []).append(float(r[mv].replace(",", "")))
    except Exception: pass
tot = sum(sum(v) for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])): print("%-60s launches %4d  total %10.3f ms  per launch %8.3f ms  share %5.1f%%" % (k[:60], len(v), sum(v) / 1e6, sum(v) / 1e6 / len(v), 100 * sum(v) / tot))
PY
