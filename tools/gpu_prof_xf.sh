# ncu capture of the fused exchange kernels (first launch of each) on a chan workload; summaries into gpurun_out/
W=${1:-n2_hf_l12m3}
TAG=${2:-r02}
python bench.py --workload $W --steps 2 --warmup 1 --kernels-only > gpurun_out/${TAG}_ko_$W.json 2> gpurun_out/${TAG}_ko_$W.err || { tail -n 5 gpurun_out/${TAG}_ko_$W.err; exit 1; }
cat gpurun_out/${TAG}_ko_$W.json
for K in chan_xf_same_kernel chan_xf_pair_kernel chan_x1_mma_kernel chan_x4_smem_kernel chan_x0_mask_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -f -o gpurun_out/${TAG}_$K python bench.py --workload $W --steps 1 --warmup 1 --kernels-only > gpurun_out/${TAG}_ncu_$K.log 2>&1
  python tools/ncu_summary.py gpurun_out/${TAG}_$K.ncu-rep > gpurun_out/${TAG}_prof_${K}_$W.txt
  cat gpurun_out/${TAG}_prof_${K}_$W.txt
done
# launch list of the default bench command (per-launch durations, cold-cache and serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_$W.csv python bench.py --workload $W --steps 2 --warmup 1 --kernels-only > gpurun_out/${TAG}_ncu_launches.log 2>&1
python - <<PY
import csv, collections
rows = list(csv.reader(l for l in open("gpurun_out/${TAG}_launches_$W.csv") if l.startswith('"')))
h = rows[0]; kn = h.index("Kernel Name"); mv = h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    try: agg.setdefault(r[kn].split("(")[0], []).append(float(r[mv].replace(",", "")))
    except Exception: pass
tot = sum(sum(v) for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])): print("%-60s launches %4d  total %10.3f ms  share %5.1f%%" % (k[:60], len(v), sum(v) / 1e6, 100 * sum(v) / tot))
PY
