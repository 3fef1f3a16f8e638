set -x
export OMP_NUM_THREADS=6
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 2 --workload n2_hf_cfg4 --shard --steps 3 > gpurun_out/r01j_cfg4_shard_n2.json 2> gpurun_out/r01j_cfg4_shard_n2.err; tail -c 600 gpurun_out/r01j_cfg4_shard_n2.err; cat gpurun_out/r01j_cfg4_shard_n2.json
