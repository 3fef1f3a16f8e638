# The multi-pass union path of the left angular contraction (x1: a tile whose partner union exceeds CHX_UMAX slots is processed in
# several passes that accumulate into W) is not reached by the oracle-sized cases with the production CHX_UMAX = 64.  This script runs
# the oracle parity tests against a build with -DHFB_CHX_UMAX=4 (built beforehand into helfem_b200/_variants/lib_umax4.so:
#   HFB200_NVCC_EXTRA="-DHFB_CHX_UMAX=4" python -m helfem_b200.build --force; cp helfem_b200/libhelfem_b200.so helfem_b200/_variants/lib_umax4.so; python -m helfem_b200.build --force)
cp helfem_b200/libhelfem_b200.so /tmp/lib_keep.so
cp helfem_b200/_variants/lib_umax4.so helfem_b200/libhelfem_b200.so
timeout 900 python -m pytest tests/test_diatomic_gpu.py tests/test_atomic_gpu.py tests/test_hotpath_goldens.py -m gpu -q 2>&1 | tail -n 4
cp /tmp/lib_keep.so helfem_b200/libhelfem_b200.so
