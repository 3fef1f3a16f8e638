timeout 900 python -m pytest tests/test_atomic_gpu.py tests/test_diatomic_gpu.py tests/test_shard_gpu.py tests/test_hotpath_goldens.py tests/test_cfg4_fullsize_gpu.py -m gpu -x -q -s 2>&1 | tail -n 6
HFB200_X_VERBOSE=1 timeout 300 python bench.py --workload n2_hf_cfg4 --steps 3 --warmup 1 --kernels-only 2>&1 | grep -v "host call" | tail -n 3
