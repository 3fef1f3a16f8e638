"""Golden vectors of the hot path itself: outputs of the CPU oracle (oracle/) on the seeded test densities, reduced to
compact probes (every element enters: weighted checksums against two fixed seeded vectors, plus the scalars), written
to tests/golden/hotpath_goldens.json.  The CPU suite checks that the oracle still reproduces them (guards the oracle
against drift); the GPU suite checks the CUDA path against the SAME committed numbers (tests/test_hotpath_goldens*.py).

Usage (authoring container, CPU only):  python tests/golden/make_hotpath_goldens.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))


def probes(M):
    """Three numbers that every element of M enters with a different weight (seeded, size-determined)."""
    from util import splitmix64
    M = np.asarray(M, dtype=np.float64)
    flat = M.reshape(-1, order="F")
    w1 = splitmix64(flat.size, seed=1234)
    w2 = splitmix64(flat.size, seed=4321)
    return [float(np.dot(w1, flat)), float(np.dot(w2, flat)), float(np.sqrt(np.dot(flat, flat)))]


def sadatom_entries():
    """(name, callable returning dict of probe lists) for the CPU oracle; the GPU test mirrors the same calls."""
    import test_sadatom_gpu as ts
    out = {}
    for case in ("ne_lip_small", "kr_hip_small"):
        ts_b, o, dens = ts.setup_case(case)
        P = np.asfortranarray(dens[None].sum(axis=2))
        out["sadatom/%s/coulomb" % case] = probes(o.coulomb(P))
        for name, (x, c) in ts.METHODS.items():
            H, exc, nel = o.eval_fxc(x, c, dens[None])
            out["sadatom/%s/fxc_%s" % (case, name)] = probes(H) + [float(exc), float(nel)]
            Ha, Hb, exc, nel = o.eval_fxc(x, c, dens["a"], dens["b"])
            out["sadatom/%s/fxc_u_%s" % (case, name)] = probes(Ha) + probes(Hb) + [float(exc), float(nel)]
    return out


def chan_entries():
    import test_atomic_gpu as ta
    import test_diatomic_gpu as td
    out = {}
    for case in ("ne_cfg1", "ne_l2m2"):
        b, o, P = ta.setup_case(case)
        out["atomic/%s/coulomb" % case] = probes(o.coulomb(P))
        out["atomic/%s/exchange" % case] = probes(o.exchange(P))
    for case in ("h2_sigma", "n2_l2m1", "co_l3m2"):
        b, o, P = td.setup_case(case)
        out["diatomic/%s/coulomb" % case] = probes(o.coulomb(P))
        out["diatomic/%s/exchange" % case] = probes(o.exchange(P))
    return out


if __name__ == "__main__":
    gold = {}
    gold.update(sadatom_entries())
    gold.update(chan_entries())
    json.dump(gold, open(os.path.join(HERE, "hotpath_goldens.json"), "w"), indent=1, sort_keys=True)
    print("wrote", len(gold), "entries")
