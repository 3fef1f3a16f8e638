timeout 600 python -m pytest tests/test_sweep_gpu.py -m gpu -x -q -s 2>&1 | tail -n 6
timeout 600 python bench.py --workload sweep_z1_86 --steps 10 > gpurun_out/r02l_sweep_n1.json 2> gpurun_out/r02l_sweep_n1.err || tail -n 15 gpurun_out/r02l_sweep_n1.err
cat gpurun_out/r02l_sweep_n1.json | cut -c1-1800
