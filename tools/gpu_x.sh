# scratch GPU script (quick A/B of the exchange pipeline): parity tests of the channel engine, then the half-size and full N2 builds
set -x
python -m pytest tests/test_diatomic_gpu.py tests/test_atomic_gpu.py -m gpu -x -q > gpurun_out/x_tests.log 2>&1; tail -3 gpurun_out/x_tests.log
for w in n2_hf_l12m3 n2_hf_cfg4; do
python bench.py --workload $w --steps 3 > gpurun_out/x_$w.json 2> gpurun_out/x_$w.err; tail -c 300 gpurun_out/x_$w.err; python -c "
import json;d=json.load(open('gpurun_out/x_$w.json'));print('$w', d['ms_per_step'], d['roofline'].get('exchange_stage_ms'))"
done
