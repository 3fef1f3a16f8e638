# ncu --set full capture of ONE kernel (regex $1) on workload $2 (tag $3): summary + the hottest source lines by stall samples
K=$1; W=${2:-n2_hf_l12m3}; TAG=${3:-r02}
ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -f -o gpurun_out/${TAG}_$K python bench.py --workload $W --steps 1 --warmup 1 --kernels-only > gpurun_out/${TAG}_ncu_$K.log 2>&1
python tools/ncu_summary.py gpurun_out/${TAG}_$K.ncu-rep > gpurun_out/${TAG}_prof_${K}_$W.txt
cat gpurun_out/${TAG}_prof_${K}_$W.txt
ncu -i gpurun_out/${TAG}_$K.ncu-rep --page source --csv > gpurun_out/${TAG}_src_$K.csv 2>/dev/null
python - <<PY
import csv
rows=list(csv.reader(open("gpurun_out/${TAG}_src_$K.csv")))
hdr=rows[0]
def col(n):
    for i,h in enumerate(hdr):
        if h.strip()==n: return i
    return None
ci=col("Source"); cs=col("# Samples") or col("Warp Stall Sampling (All Samples)") or col("Sampling Data (All)")
print("columns:", [h for h in hdr][:12])
if ci is not None and cs is not None:
    data=[]
    for r in rows[1:]:
        try: data.append((float(r[cs] or 0), r[ci]))
        except Exception: pass
    tot=sum(d[0] for d in data) or 1
    for v,sx in sorted(data, reverse=True)[:25]:
        print("%6.2f%%  %s" % (100*v/tot, sx.strip()[:150]))
PY
