timeout 300 python bench.py --workload n2_hf_cfg4 --steps 3 --warmup 1 --kernels-only 2>&1 | tail -n 1
