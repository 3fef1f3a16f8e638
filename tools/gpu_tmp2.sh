for v in 0 3500 14000; do
  cp helfem_b200/_variants/lib_$v.so helfem_b200/libhelfem_b200.so
  echo "stagger $v"; timeout 300 python bench.py --workload n2_hf_cfg4 --steps 3 --warmup 1 --kernels-only 2>&1 | tail -n 1 | cut -c1-330
done
