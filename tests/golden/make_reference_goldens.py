"""Extracts the known-answer vectors the reference's OWN tests hold for this path into small
fixtures (run in the authoring container, where /root/reference is mounted; the fixtures travel).

  gaunt_m0.json        the 550 (value, reference) pairs of src/general/gaunt_test.cpp
                       (gaunt_coefficient and modified_gaunt_coefficient, all with M=m=m'=0)
  lobatto_nodes.json   the Gauss-Lobatto abscissas the reference tabulates for n = 2..19
                       (libhelfem/src/lobatto.cpp, lobatto_set), rounded to double

Usage: python tests/golden/make_reference_goldens.py
"""
import json
import os
import re

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def gaunt():
    src = open(os.path.join(REF, "src/general/gaunt_test.cpp")).read()
    pat = re.compile(r"val=helfem::gaunt::(modified_)?gaunt_coefficient\((-?\d+),(-?\d+),(-?\d+),(-?\d+),(-?\d+),(-?\d+)\);\s*ref=([-+0-9.eE]+);")
    out = []
    for m in pat.finditer(src):
        out.append({"modified": bool(m.group(1)), "args": [int(m.group(i)) for i in range(2, 8)], "ref": float(m.group(8))})
    json.dump(out, open(os.path.join(HERE, "gaunt_m0.json"), "w"))
    print("gaunt:", len(out), "entries,", sum(1 for e in out if e["ref"] != 0.0), "nonzero")


def lobatto():
    src = open(os.path.join(REF, "libhelfem/src/lobatto.cpp")).read()
    body = src[src.index("void lobatto_set"):src.index("void lobatto_compute")]
    out = {}
    parts = re.split(r"if\s*\(\s*order\s*==\s*(\d+)\s*\)", body)
    for i in range(1, len(parts), 2):
        n = int(parts[i])
        xs = re.findall(r"xtab\[(\d+)\]\s*=\s*(-?)\s*([0-9.eE+]+)\s*;", parts[i + 1])
        if len(xs) == n:
            v = [0.0] * n
            for k, sign, num in xs:
                v[int(k)] = float(sign + num)
            out[str(n)] = v
    json.dump(out, open(os.path.join(HERE, "lobatto_nodes.json"), "w"))
    print("lobatto: orders", sorted(int(k) for k in out))


if __name__ == "__main__":
    gaunt()
    lobatto()
