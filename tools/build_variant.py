"""Build a tuning variant of the native library with extra -D flags (A/B runs on the GPU box):

    python tools/build_variant.py mb3 -DHFB_GRID_MINBLOCKS=3
    HELFEM_B200_LIB=helfem_b200/_variants/libhelfem_b200_mb3.so python bench.py --kernels-only

Only the spherically-averaged-atom kernels are recompiled; the other objects come from the regular build
(helfem_b200/_build)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from helfem_b200 import build as B  # noqa: E402


def main():
    name, flags = sys.argv[1], sys.argv[2:]
    B.build()
    vdir = os.path.join(B.HERE, "_variants")
    os.makedirs(vdir, exist_ok=True)
    objs, procs = [], []
    for src in B.CU_SOURCES + B.CPP_SOURCES:
        obj = os.path.join(B.BUILD, src.replace("/", "_") + ".o")
        if src.startswith("csrc/sadatom_grid_") or src == "csrc/capi_sadatom.cu":
            obj = os.path.join(vdir, "%s_%s.o" % (os.path.basename(src)[:-3], name))
            cmd = [B.NVCC] + B.NVCC_FLAGS + flags + ["-Xptxas", "-v", "-c", os.path.join(B.HERE, src), "-o", obj]
            procs.append((subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT), src))
        objs.append(obj)
    for pr, src in procs:
        out = pr.communicate()[0].decode()
        if pr.returncode != 0:
            sys.stderr.write(out)
            raise SystemExit("compilation of %s failed" % src)
        lines = out.splitlines()
        for i, ln in enumerate(lines):
            if "sad_grid_mma_kernelILi2ELi1ELi2ELi1E" in ln and "Compiling" in ln:
                print(name, [x.strip() for x in lines[i + 1:i + 4] if "registers" in x or "spill" in x])
    lib = os.path.join(vdir, "libhelfem_b200_%s.so" % name)
    subprocess.check_call([B.NVCC] + B.ARCH + ["-shared", "-o", lib] + objs + ["-Xcompiler", "-fopenmp", "-lgomp"])
    print(lib)


if __name__ == "__main__":
    main()
