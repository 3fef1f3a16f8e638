set -x
timeout 900 python -m pytest tests/test_diatomic_gpu.py tests/test_atomic_gpu.py tests/test_shard_gpu.py -m gpu -x -q > gpurun_out/w_tests.log 2>&1; tail -5 gpurun_out/w_tests.log
timeout 600 python bench.py --workload n2_hf_l12m3 --steps 3 --kernels-only > gpurun_out/w_l12m3.json 2> gpurun_out/w_l12m3.err; cat gpurun_out/w_l12m3.json
timeout 900 python bench.py --workload n2_hf_cfg4 --steps 2 --kernels-only > gpurun_out/w_cfg4.json 2> gpurun_out/w_cfg4.err; cat gpurun_out/w_cfg4.json
python bench.py --workload rn_r2scan_cfg3 --batch 128 --steps 10 > gpurun_out/w_bench_cfg3.json 2> gpurun_out/w_bench_cfg3.err; tail -c 400 gpurun_out/w_bench_cfg3.err; cat gpurun_out/w_bench_cfg3.json
