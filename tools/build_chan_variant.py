"""Build a tuning variant of the native library with extra -D flags for the channel engine (A/B runs on the GPU box):

    python tools/build_chan_variant.py v2s6 -DHFB_X2_MINB=2 -DHFB_X2_STAGES=6
    HELFEM_B200_LIB=helfem_b200/_variants/libhelfem_b200_v2s6.so python bench.py --workload n2_hf_l12m3
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from helfem_b200 import build as B  # noqa: E402

name, flags = sys.argv[1], sys.argv[2:]
B.build()
vdir = os.path.join(B.HERE, "_variants")
os.makedirs(vdir, exist_ok=True)
objs = []
for src in B.CU_SOURCES + B.CPP_SOURCES:
    obj = os.path.join(B.BUILD, src.replace("/", "_") + ".o")
    if src == "csrc/chan_engine.cu":
        obj = os.path.join(vdir, "chan_engine_%s.o" % name)
        out = subprocess.run([B.NVCC] + B.NVCC_FLAGS + flags + ["-Xptxas", "-v", "-x", "cu", "-c", os.path.join(B.HERE, src), "-o", obj],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT).stdout.decode().splitlines()
        for i, ln in enumerate(out):
            if "Compiling" in ln and ("chan_x2_mma" in ln or "chan_x1_mma" in ln):
                print(name, ln.split("'")[1][8:30], [x.strip() for x in out[i + 1:i + 4] if "registers" in x or "spill" in x])
    objs.append(obj)
lib = os.path.join(vdir, "libhelfem_b200_%s.so" % name)
subprocess.check_call([B.NVCC] + B.ARCH + ["-shared", "-o", lib] + objs + ["-Xcompiler", "-fopenmp", "-lgomp"])
print(lib)
