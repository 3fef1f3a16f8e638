set -x
timeout 900 python -m pytest tests/test_scf_energy_gpu.py -m gpu -q > gpurun_out/u_tests.log 2>&1; tail -25 gpurun_out/u_tests.log
