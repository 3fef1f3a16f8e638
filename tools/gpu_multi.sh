# N-GPU checks ($1 = N, $2 = tag): XC golden test on the device, in-library NCCL parity, default bench sharded, the Z sweep over N ranks
N=${1:-2}; TAG=${2:-r02}
# (OMP_NUM_THREADS is left to bench.py: rank 0 of a root-driven run gets the cores)
timeout 300 python -m pytest tests/test_xc_pins_gpu.py -m gpu -q -s 2>&1 | tail -n 3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/shard_nccl_check.py > gpurun_out/${TAG}_shard_check_n$N.json 2> gpurun_out/${TAG}_shard_check_n$N.err; tail -n 3 gpurun_out/${TAG}_shard_check_n$N.err; cat gpurun_out/${TAG}_shard_check_n$N.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --steps 10 > gpurun_out/${TAG}_cfg4_n$N.json 2> gpurun_out/${TAG}_cfg4_n$N.err; tail -n 3 gpurun_out/${TAG}_cfg4_n$N.err
python -c "import json; d=json.load(open('gpurun_out/${TAG}_cfg4_n$N.json')); print('cfg4 N=$N', d['value'], d['ms_per_step'], d['scaling'], 'e2e', d['e2e']['value']); [print(' ', e['stage'], round(e['ms'],3)) for e in d['roofline']['stages']]"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --workload sweep_z1_86 --steps 10 > gpurun_out/${TAG}_sweep_n$N.json 2> gpurun_out/${TAG}_sweep_n$N.err; tail -n 3 gpurun_out/${TAG}_sweep_n$N.err
python -c "import json; d=json.load(open('gpurun_out/${TAG}_sweep_n$N.json')); print('sweep N=$N', d['value'], d['ms_per_step'], d['config']['builds_per_rank'], 'e2e', d['e2e']['value'], d['checks'])"
