"""Committed golden vectors of the hot path (tests/golden/hotpath_goldens.json, written by
tests/golden/make_hotpath_goldens.py from the CPU oracle on the seeded densities):
  * CPU (-m "not gpu"): the oracle still reproduces them -- the checker cannot drift unnoticed;
  * GPU (-m gpu): the CUDA path, through the host mirror classes / C ABI, matches the same committed numbers.
Each entry holds two seeded weighted checksums and the Frobenius norm of an operator (every element enters), plus the
scalars (Exc, Nel) where the operator has them."""
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
GOLD = json.load(open(os.path.join(HERE, "golden", "hotpath_goldens.json")))


def check(key, got, rel=1e-9):
    ref = np.asarray(GOLD[key])
    got = np.asarray(got)
    assert got.shape == ref.shape, key
    scale = max(1.0, float(np.max(np.abs(ref))))
    assert np.max(np.abs(got - ref)) <= rel * scale, (key, got - ref)


def test_oracle_reproduces_goldens():
    import make_hotpath_goldens as mk
    got = mk.chan_entries()
    assert set(got) <= set(GOLD)
    for k, v in got.items():
        check(k, v, rel=1e-11)


@pytest.mark.gpu
def test_cuda_matches_sadatom_goldens(native):
    import make_hotpath_goldens as mk
    import test_sadatom_gpu as ts
    for case in ("ne_lip_small", "kr_hip_small"):
        b, o, dens = ts.setup_case(case)
        P = np.asfortranarray(dens[None].sum(axis=2))
        check("sadatom/%s/coulomb" % case, mk.probes(b.coulomb(P)))
        for name, (x, c) in ts.METHODS.items():
            H, exc, nel = b.eval_fxc(x, c, dens[None])
            check("sadatom/%s/fxc_%s" % (case, name), mk.probes(H) + [exc, nel])
            Ha, Hb, exc, nel = b.eval_fxc(x, c, dens["a"], dens["b"])
            check("sadatom/%s/fxc_u_%s" % (case, name), mk.probes(Ha) + mk.probes(Hb) + [exc, nel])


@pytest.mark.gpu
def test_cuda_matches_channel_goldens(native):
    import make_hotpath_goldens as mk
    import test_atomic_gpu as ta
    import test_diatomic_gpu as td
    for case in ("ne_cfg1", "ne_l2m2"):
        b, o, P = ta.setup_case(case)
        check("atomic/%s/coulomb" % case, mk.probes(b.coulomb(P)))
        check("atomic/%s/exchange" % case, mk.probes(b.exchange(P)))
    for case in ("h2_sigma", "n2_l2m1", "co_l3m2"):
        b, o, P = td.setup_case(case)
        check("diatomic/%s/coulomb" % case, mk.probes(b.coulomb(P)))
        check("diatomic/%s/exchange" % case, mk.probes(b.exchange(P)))
