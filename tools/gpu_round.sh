# round evidence: sanitizer over smoke(), the non-default workloads, reference arms
TAG=${1:-r02}
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -n 4 gpurun_out/${TAG}_sanitizer_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_sanitizer_racecheck.log 2>&1; echo "racecheck rc=$?"; tail -n 4 gpurun_out/${TAG}_sanitizer_racecheck.log
for W in kr_pbe_cfg2 rn_r2scan_cfg3 ne_hf_cfg1; do
  timeout 600 python bench.py --workload $W --steps 10 > gpurun_out/${TAG}_bench_$W.json 2> gpurun_out/${TAG}_bench_$W.err || tail -n 5 gpurun_out/${TAG}_bench_$W.err
  python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_$W.json')); print('$W', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'cpu', d['cpu_baseline'] and d['cpu_baseline']['value'], 'roof', d['roofline'] and d['roofline'].get('frac'))"
done
timeout 600 python bench.py --impl reference --workload kr_pbe_cfg2 --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref_kr.json 2>/dev/null; cut -c1-400 gpurun_out/${TAG}_bench_ref_kr.json
