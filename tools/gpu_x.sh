set -x
python -m pytest tests/test_diatomic_gpu.py tests/test_atomic_gpu.py -m gpu -x -q > gpurun_out/x_tests.log 2>&1; tail -3 gpurun_out/x_tests.log
run() { python bench.py --workload n2_hf_l12m3 --steps 5 > gpurun_out/x_l12m3_$1.json 2> gpurun_out/x_l12m3_$1.err; tail -c 300 gpurun_out/x_l12m3_$1.err; python -c "
import json;d=json.load(open('gpurun_out/x_l12m3_$1.json'));print('$1', d['ms_per_step'], d['roofline'].get('exchange_stage_ms'))"; }
run base
python bench.py --workload n2_hf_cfg4 --steps 3 > gpurun_out/x_cfg4.json 2> gpurun_out/x_cfg4.err; tail -c 300 gpurun_out/x_cfg4.err; python -c "
import json;d=json.load(open('gpurun_out/x_cfg4.json'));print('base', d['ms_per_step'], d['roofline'].get('exchange_stage_ms'))"
