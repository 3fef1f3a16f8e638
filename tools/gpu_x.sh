set -x
timeout 600 python bench.py --workload n2_hf_l12m3 --steps 3 --kernels-only > gpurun_out/y_l12m3.json 2> gpurun_out/y_l12m3.err; cat gpurun_out/y_l12m3.json
timeout 900 python bench.py --workload n2_hf_cfg4 --steps 2 --kernels-only > gpurun_out/y_cfg4.json 2> gpurun_out/y_cfg4.err; cat gpurun_out/y_cfg4.json
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/y_gpu_tests.log 2>&1; tail -3 gpurun_out/y_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/y_smoke.log 2>&1; tail -2 gpurun_out/y_smoke.log
