set -x
timeout 900 python -m pytest tests/test_diatomic_gpu.py tests/test_atomic_gpu.py -m gpu -x -q > gpurun_out/v_tests.log 2>&1; tail -5 gpurun_out/v_tests.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 600 python bench.py --workload n2_hf_l12m3 --steps 3 --kernels-only > gpurun_out/v_l12m3.json 2> gpurun_out/v_l12m3.err; cat gpurun_out/v_l12m3.json
timeout 900 python bench.py --workload n2_hf_cfg4 --steps 2 --kernels-only > gpurun_out/v_cfg4.json 2> gpurun_out/v_cfg4.err; cat gpurun_out/v_cfg4.json
