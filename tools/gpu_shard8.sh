# configs[3] with the integral families sharded over N GPUs (NCCL all-reduce of J and K) on one 8-GPU box; R = artifact tag
R=${R:-r01m}
NS=${NS:-"8 4 2"}
set -x
export OMP_NUM_THREADS=6
for N in $NS; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+N)) bench.py --gpus $N --workload n2_hf_cfg4 --shard --steps 3 > gpurun_out/${R}_cfg4_shard_n$N.json 2> gpurun_out/${R}_cfg4_shard_n$N.err; tail -c 300 gpurun_out/${R}_cfg4_shard_n$N.err; cat gpurun_out/${R}_cfg4_shard_n$N.json
done
