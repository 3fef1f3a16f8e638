timeout 900 python -m pytest tests/test_grid2d_gpu.py tests/test_cfg4b_fullsize_gpu.py tests/test_diatomic_gpu.py tests/test_atomic_gpu.py -m gpu -x -q 2>&1 | tail -n 4
export HFB200_X_VERBOSE=1
timeout 600 python bench.py --workload n2_pbe_cfg4b --steps 5 --warmup 3 2>&1 | tail -n 3
timeout 600 python bench.py --steps 3 --warmup 3 2>&1 | grep -v "^\[hfb200\] exchange:" | tail -n 8
