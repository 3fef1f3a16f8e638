bash tools/gpu_prof_one.sh chan_xf_pair_kernel n2_hf_cfg4 r02t > gpurun_out/r02t_pair_hot.txt 2>&1
tail -n 40 gpurun_out/r02t_pair_hot.txt
