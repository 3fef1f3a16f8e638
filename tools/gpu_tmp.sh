HFB200_X_VERBOSE=1 timeout 600 python bench.py --steps 3 > gpurun_out/r02k_bench_cfg4.json 2> gpurun_out/r02k_cfg4.err; grep "host call" gpurun_out/r02k_cfg4.err | tail -n 4
python -c "import json; d=json.load(open('gpurun_out/r02k_bench_cfg4.json')); print(d['value'], d['ms_per_step'], d['e2e'])"
