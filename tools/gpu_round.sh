# Full GPU check of the current tree: parity tests, smoke, bench lines of every workload (both arms for the default one),
# ncu launch lists and one --set full capture per hot kernel.  Outputs under gpurun_out/ with the round tag R.
R=${R:-r01n}
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/${R}_gpu_tests.log 2>&1; tail -3 gpurun_out/${R}_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${R}_smoke.log 2>&1; tail -2 gpurun_out/${R}_smoke.log
python bench.py > gpurun_out/${R}_bench_cfg2.json 2> gpurun_out/${R}_bench_cfg2.err; tail -c 600 gpurun_out/${R}_bench_cfg2.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${R}_bench_ref.json 2> gpurun_out/${R}_bench_ref.err
python bench.py --workload rn_r2scan_cfg3 --batch 128 --steps 10 > gpurun_out/${R}_bench_cfg3.json 2> gpurun_out/${R}_bench_cfg3.err
python bench.py --workload ne_hf_cfg1 --steps 20 > gpurun_out/${R}_bench_cfg1.json 2> gpurun_out/${R}_bench_cfg1.err
python bench.py --workload n2_hf_l4m1 --steps 10 > gpurun_out/${R}_bench_n2_l4m1.json 2> gpurun_out/${R}_bench_n2_l4m1.err
python bench.py --workload n2_hf_l12m3 --steps 5 > gpurun_out/${R}_bench_n2_l12m3.json 2> gpurun_out/${R}_bench_n2_l12m3.err
python bench.py --workload n2_hf_cfg4 --steps 3 > gpurun_out/${R}_bench_cfg4.json 2> gpurun_out/${R}_bench_cfg4.err; tail -c 600 gpurun_out/${R}_bench_cfg4.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${R}_launches_cfg2.csv python bench.py --kernels-only --steps 2 --warmup 1 > gpurun_out/ncu_l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${R}_launches_n2_l12m3.csv python bench.py --kernels-only --workload n2_hf_l12m3 --steps 1 --warmup 1 > gpurun_out/ncu_l2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sad_grid_mma -c 1 -s 2 -o gpurun_out/${R}_grid_cfg2 -f python bench.py --kernels-only --steps 2 --warmup 1 > gpurun_out/ncu_g.log 2>&1
for k in chan_x1 chan_x2 chan_jmma; do
ncu --set full --clock-control none --import-source on -k "regex:$k" -c 1 -s 3 -o gpurun_out/${R}_${k}_l12m3 -f python bench.py --kernels-only --workload n2_hf_l12m3 --steps 1 --warmup 1 > gpurun_out/ncu_$k.log 2>&1; tail -1 gpurun_out/ncu_$k.log
done
# the two x3 launches of one build (same element, element pairs)
ncu --set full --clock-control none --import-source on -k "regex:chan_x3_kernel" -c 2 -s 6 -o gpurun_out/${R}_chan_x3_l12m3 -f python bench.py --kernels-only --workload n2_hf_l12m3 --steps 1 --warmup 1 > gpurun_out/ncu_x3.log 2>&1; tail -1 gpurun_out/ncu_x3.log
cat gpurun_out/${R}_bench_cfg2.json gpurun_out/${R}_bench_ref.json gpurun_out/${R}_bench_cfg4.json
