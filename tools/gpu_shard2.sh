# N = 2: in-library NCCL parity check, then the default bench workload sharded over two GPUs (strong scaling)
set -x
export OMP_NUM_THREADS=6
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tools/shard_nccl_check.py > gpurun_out/${TAG:-r02}_shard_check_n2.json 2> gpurun_out/${TAG:-r02}_shard_check_n2.err; tail -n 5 gpurun_out/${TAG:-r02}_shard_check_n2.err; cat gpurun_out/${TAG:-r02}_shard_check_n2.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 2 --steps 5 > gpurun_out/${TAG:-r02}_cfg4_n2.json 2> gpurun_out/${TAG:-r02}_cfg4_n2.err; tail -n 5 gpurun_out/${TAG:-r02}_cfg4_n2.err
python -c "import json; d=json.load(open('gpurun_out/${TAG:-r02}_cfg4_n2.json')); print(d['value'], d['ms_per_step'], d['scaling'], d['e2e']['value']); [print(e['stage'], e['ms']) for e in d['roofline']['stages']]"
