# end-of-round evidence on one GPU ($1 = tag): the full GPU suite, smoke, the default bench and its reference arm, the other workloads
TAG=${1:-r02z}
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/${TAG}_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -n 2
timeout 900 python bench.py > gpurun_out/${TAG}_bench_cfg4_n1.json 2> gpurun_out/${TAG}_bench_cfg4_n1.err || tail -n 5 gpurun_out/${TAG}_bench_cfg4_n1.err
python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_cfg4_n1.json')); print('cfg4', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'whole', d['roofline']['whole_build_frac_fp64']); [print(' ', e['stage'], round(e['ms'],3), e.get('frac_fp64') or e.get('frac_hbm'), e.get('traffic_over_algorithmic')) for e in d['roofline']['stages']]"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref_cfg4.json 2>/dev/null; cut -c1-300 gpurun_out/${TAG}_bench_ref_cfg4.json
for W in n2_pbe_cfg4b kr_pbe_cfg2 rn_r2scan_cfg3 ne_hf_cfg1 sweep_z1_86; do
  timeout 600 python bench.py --workload $W --steps 10 > gpurun_out/${TAG}_bench_$W.json 2> gpurun_out/${TAG}_bench_$W.err || tail -n 5 gpurun_out/${TAG}_bench_$W.err
  python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_$W.json')); print('$W', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'cpu', d['cpu_baseline'] and d['cpu_baseline']['value'], 'roof', d['roofline'] and d['roofline'].get('frac'))"
done
