# 2-GPU checks: family-sharded diatomic J/K over NCCL vs the oracle, and the weak-scaling bench at N=1,2.
set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tools/shard_nccl_check.py --lmax 4 --mmax 2 --nelem 4 --nnodes 8 > gpurun_out/r01f_shard_n2.json 2> gpurun_out/r01f_shard_n2.err; tail -c 400 gpurun_out/r01f_shard_n2.err
python tools/shard_nccl_check.py --lmax 4 --mmax 2 --nelem 4 --nnodes 8 --no-oracle > gpurun_out/r01f_shard_n1.json 2> gpurun_out/r01f_shard_n1.err
python bench.py --gpus 1 > gpurun_out/r01f_bench_n1.json 2> gpurun_out/r01f_bench_n1.err; tail -c 300 gpurun_out/r01f_bench_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 > gpurun_out/r01f_bench_n2.json 2> gpurun_out/r01f_bench_n2.err; tail -c 300 gpurun_out/r01f_bench_n2.err
cat gpurun_out/r01f_shard_n2.json gpurun_out/r01f_shard_n1.json gpurun_out/r01f_bench_n1.json gpurun_out/r01f_bench_n2.json
