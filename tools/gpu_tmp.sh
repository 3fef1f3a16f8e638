timeout 600 python -m pytest tests/test_shard_gpu.py tests/test_atomic_gpu.py tests/test_diatomic_gpu.py tests/test_hotpath_goldens.py -m gpu -x -q 2>&1 | tail -n 2
timeout 300 python bench.py --workload n2_hf_cfg4 --steps 5 --warmup 1 --kernels-only 2>&1 | tail -n 1
