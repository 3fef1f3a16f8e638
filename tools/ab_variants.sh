# A/B of tuning variants of the grid kernel (tools/build_variant.py) on the GPU box
python -m pytest tests/test_sadatom_gpu.py -x -q 2>&1 | tail -3
for v in "" $@; do
  if [ -z "$v" ]; then L=""; else L=helfem_b200/_variants/libhelfem_b200_$v.so; fi
  echo "== variant ${v:-default}"
  HELFEM_B200_LIB=$L python bench.py --kernels-only --split --steps 20 --warmup 3 2>&1 | tail -1
  HELFEM_B200_LIB=$L python bench.py --kernels-only --steps 10 --warmup 3 --workload rn_r2scan_cfg3 --batch 128 2>&1 | tail -1
done
export HELFEM_B200_LIB=helfem_b200/_variants/libhelfem_b200_prof.so
python bench.py --kernels-only --steps 1 --warmup 0 2>&1 | grep "grid profile" | head -6
python bench.py --kernels-only --steps 1 --warmup 0 --workload rn_r2scan_cfg3 --batch 128 2>&1 | grep "grid profile" | head -4
