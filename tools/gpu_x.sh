set -x
timeout 900 python -m pytest tests/test_diatomic_gpu.py tests/test_atomic_gpu.py tests/test_shard_gpu.py -m gpu -x -q > gpurun_out/x_tests.log 2>&1; tail -15 gpurun_out/x_tests.log
python tools/shard_nccl_check.py --lmax 4 --mmax 2 --nelem 4 --nnodes 8 > gpurun_out/x_shard_n1.json 2> gpurun_out/x_shard_n1.err; cat gpurun_out/x_shard_n1.json; tail -5 gpurun_out/x_shard_n1.err
