set -x
timeout 900 python -m pytest tests/test_diatomic_gpu.py tests/test_atomic_gpu.py tests/test_shard_gpu.py tests/test_capi_symbols.py -x -q > gpurun_out/x_tests.log 2>&1; tail -5 gpurun_out/x_tests.log
timeout 600 python bench.py --workload n2_hf_l4m1 --steps 5 > gpurun_out/x_bench_l4m1.json 2> gpurun_out/x_bench_l4m1.err; tail -c 600 gpurun_out/x_bench_l4m1.err; cat gpurun_out/x_bench_l4m1.json
timeout 900 python bench.py --workload n2_hf_l12m3 --steps 3 > gpurun_out/x_bench_l12m3.json 2> gpurun_out/x_bench_l12m3.err; tail -c 600 gpurun_out/x_bench_l12m3.err; cat gpurun_out/x_bench_l12m3.json
timeout 1500 python bench.py --workload n2_hf_cfg4 --steps 2 > gpurun_out/x_bench_cfg4.json 2> gpurun_out/x_bench_cfg4.err; tail -c 800 gpurun_out/x_bench_cfg4.err; cat gpurun_out/x_bench_cfg4.json
for k in x1 x2 x3; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chan_$k -c 1 -s 4 -o gpurun_out/r01h_${k}_l12m3 -f python bench.py --kernels-only --workload n2_hf_l12m3 --steps 1 --warmup 1 > gpurun_out/ncu_$k.log 2>&1; tail -2 gpurun_out/ncu_$k.log
done
