set -x
python -m pytest tests/test_sadatom_gpu.py -x -q 2>&1 | tail -3
for v in "" mb2 mb3 mb4 w4mb6 w4mb4 w16mb1; do
  if [ -z "$v" ]; then L=""; else L=helfem_b200/_variants/libhelfem_b200_$v.so; fi
  echo "== variant ${v:-default}"
  HELFEM_B200_LIB=$L python bench.py --kernels-only --steps 20 --warmup 3 2>&1 | tail -1
  HELFEM_B200_LIB=$L python bench.py --kernels-only --steps 10 --warmup 3 --workload rn_r2scan_cfg3 --batch 128 2>&1 | tail -1
done
