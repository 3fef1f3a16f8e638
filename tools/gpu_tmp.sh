timeout 900 python -m pytest tests/test_cfg4b_fullsize_gpu.py -m gpu -x -q -s 2>&1 | tail -n 12
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"g2d_" -c 40 --csv --log-file gpurun_out/r02q_launches_4b.csv python bench.py --workload n2_pbe_cfg4b --steps 1 --warmup 1 --kernels-only > gpurun_out/r02q_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/r02q_launches_4b.csv")) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); ui=hdr.index("Metric Unit")
agg=collections.OrderedDict()
for r in rows[1:]:
    v=float(r[vi].replace(",","")); u=r[ui]
    ms = v/1e6 if u.startswith("ns") else v/1e3 if u.startswith("us") else v
    k=r[ki].split("(")[0][-40:]
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=ms
for k,(n,t) in agg.items(): print("%-42s n=%4d  avg %9.3f ms" % (k,n,t/n))
PY
