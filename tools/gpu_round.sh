# Full GPU check of the current tree: parity tests, smoke, bench (both arms), cfg3 bench, ncu launch list.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01e_gpu_tests.log 2>&1; tail -3 gpurun_out/r01e_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01e_smoke.log 2>&1; tail -2 gpurun_out/r01e_smoke.log
python bench.py > gpurun_out/r01e_bench_cfg2.json 2> gpurun_out/r01e_bench_cfg2.err; tail -c 600 gpurun_out/r01e_bench_cfg2.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01e_bench_ref.json 2> gpurun_out/r01e_bench_ref.err
python bench.py --workload rn_r2scan_cfg3 --batch 128 --steps 10 > gpurun_out/r01e_bench_cfg3.json 2> gpurun_out/r01e_bench_cfg3.err; tail -c 600 gpurun_out/r01e_bench_cfg3.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01e_launches_cfg2.csv python bench.py --kernels-only --steps 2 --warmup 1 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sad_grid_mma -c 1 -s 2 -o gpurun_out/r01e_grid_cfg2 -f python bench.py --kernels-only --steps 2 --warmup 1 > gpurun_out/ncu_g.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sad_jmma -c 1 -s 2 -o gpurun_out/r01e_jmma_cfg2 -f python bench.py --kernels-only --steps 2 --warmup 1 > gpurun_out/ncu_j.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sad_grid_mma -c 1 -s 1 -o gpurun_out/r01e_grid_cfg3 -f python bench.py --kernels-only --steps 2 --warmup 1 --workload rn_r2scan_cfg3 --batch 128 > gpurun_out/ncu_g3.log 2>&1
cat gpurun_out/r01e_bench_cfg2.json gpurun_out/r01e_bench_ref.json gpurun_out/r01e_bench_cfg3.json
