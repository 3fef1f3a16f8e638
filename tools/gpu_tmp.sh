set -x
timeout 300 python bench.py --workload n2_hf_l4m1 --steps 5 > gpurun_out/r02i_bench_l4m1.json 2> gpurun_out/r02i_l4m1.err || tail -n 20 gpurun_out/r02i_l4m1.err
python -c "import json; d=json.load(open('gpurun_out/r02i_bench_l4m1.json')); print(d['value'], d['e2e']['value'], d['cpu_baseline'])"
timeout 600 python bench.py --steps 5 > gpurun_out/r02i_bench_cfg4.json 2> gpurun_out/r02i_cfg4.err || tail -n 20 gpurun_out/r02i_cfg4.err
python -c "import json; d=json.load(open('gpurun_out/r02i_bench_cfg4.json')); print(d['value'], d['ms_per_step'], d['e2e'], d['cpu_baseline']); [print(e) for e in d['roofline']['stages']]"
timeout 600 python bench.py --impl reference --steps 2 > gpurun_out/r02i_ref_cfg4.json 2> gpurun_out/r02i_ref.err || tail -n 20 gpurun_out/r02i_ref.err
cat gpurun_out/r02i_ref_cfg4.json
