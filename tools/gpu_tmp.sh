timeout 900 python -m pytest tests/test_atomic_gpu.py tests/test_diatomic_gpu.py tests/test_shard_gpu.py tests/test_hotpath_goldens.py tests/test_scf_energy_gpu.py tests/test_cfg4_fullsize_gpu.py tests/test_grid2d_gpu.py -m gpu -x -q -s 2>&1 | tail -n 6
timeout 300 python bench.py --workload n2_hf_cfg4 --steps 3 --warmup 1 --kernels-only 2>&1 | tail -n 1
