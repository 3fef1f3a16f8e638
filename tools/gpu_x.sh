set -x
python -m pytest tests -m gpu -x -q > gpurun_out/final_gpu_tests.log 2>&1; tail -3 gpurun_out/final_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; tail -3 gpurun_out/final_smoke.log
python bench.py > gpurun_out/final_bench_cfg2.json 2> gpurun_out/final_bench_cfg2.err; tail -c 300 gpurun_out/final_bench_cfg2.err; cut -c1-400 gpurun_out/final_bench_cfg2.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err; cut -c1-300 gpurun_out/final_bench_ref.json
python bench.py --workload n2_hf_cfg4 --steps 3 > gpurun_out/final_bench_cfg4.json 2> gpurun_out/final_bench_cfg4.err; tail -c 300 gpurun_out/final_bench_cfg4.err; cut -c1-300 gpurun_out/final_bench_cfg4.json
