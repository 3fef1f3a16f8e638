set -x
timeout 900 python -m pytest tests/test_hotpath_goldens.py -m gpu -q > gpurun_out/t_tests.log 2>&1; tail -5 gpurun_out/t_tests.log
timeout 600 python bench.py --workload n2_pbe_grid_l2m1 --steps 5 > gpurun_out/r01k_bench_grid_l2m1.json 2> gpurun_out/t1.err; tail -c 500 gpurun_out/t1.err; cat gpurun_out/r01k_bench_grid_l2m1.json
timeout 900 python bench.py --workload n2_pbe_grid_l6m2 --steps 3 > gpurun_out/r01k_bench_grid_l6m2.json 2> gpurun_out/t2.err; tail -c 500 gpurun_out/t2.err; cat gpurun_out/r01k_bench_grid_l6m2.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01k_launches_grid_l6m2.csv python bench.py --kernels-only --workload n2_pbe_grid_l6m2 --steps 1 --warmup 1 > gpurun_out/ncu_g2.log 2>&1
