HFB200_X_VERBOSE=1 timeout 300 python bench.py --workload n2_hf_cfg4 --steps 2 --warmup 1 --kernels-only 2>&1 | grep -v "^$" | tail -n 3
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02h_launches_cfg4.csv python bench.py --workload n2_hf_cfg4 --steps 1 --warmup 1 --kernels-only > gpurun_out/r02h_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/r02h_launches_cfg4.csv")) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); ui=hdr.index("Metric Unit")
agg=collections.OrderedDict()
for r in rows[1:]:
    v=float(r[vi].replace(",","")); u=r[ui]
    ms = v/1e6 if u.startswith("ns") else v/1e3 if u.startswith("us") else v
    k=r[ki].split("(")[0][-60:]
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=ms
for k,(n,t) in agg.items(): print("%-62s n=%4d  %9.3f ms" % (k,n,t))
PY
