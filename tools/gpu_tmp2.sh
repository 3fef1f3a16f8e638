cp helfem_b200/libhelfem_b200.so /tmp/keep.so
for v in 3_4 2_4 2_2 4_2; do
  cp helfem_b200/_variants/lib_$v.so helfem_b200/libhelfem_b200.so
  echo "x1 minblocks_unroll $v"; timeout 300 python bench.py --workload n2_hf_cfg4 --steps 3 --warmup 1 --kernels-only 2>&1 | tail -n 1 | cut -c1-330
done
