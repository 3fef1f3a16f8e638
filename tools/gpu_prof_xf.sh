# ncu capture of the fused exchange kernels (first launch of each) on a chan workload; summaries into gpurun_out/
W=${1:-n2_hf_l12m3}
TAG=${2:-r02}
python bench.py --workload $W --steps 2 --warmup 1 --kernels-only > gpurun_out/${TAG}_ko_$W.json 2> gpurun_out/${TAG}_ko_$W.err || { tail -n 5 gpurun_out/${TAG}_ko_$W.err; exit 1; }
cat gpurun_out/${TAG}_ko_$W.json
for K in chan_xf_same_kernel chan_xf_pair_kernel chan_x1_mma_kernel chan_x4_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -f -o gpurun_out/${TAG}_$K python bench.py --workload $W --steps 1 --warmup 1 --kernels-only > gpurun_out/${TAG}_ncu_$K.log 2>&1
  python tools/ncu_summary.py gpurun_out/${TAG}_$K.ncu-rep > gpurun_out/${TAG}_prof_${K}_$W.txt
  cat gpurun_out/${TAG}_prof_${K}_$W.txt
done
