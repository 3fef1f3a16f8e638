"""CPU test of the block-restricted diatomic oracle (oracle/chan_oracle.cpp: oracle_diatomic_exchange_blocks,
oracle_diatomic_coulomb_blocks): on a basis the full oracle finishes in seconds, every listed block must equal the
corresponding block of the full oracle_diatomic_exchange / oracle_diatomic_coulomb result (same loops, same order of
operations inside a block).  The block forms are what checks BASELINE configs[3] at full size on the GPU box
(tests/test_cfg4_fullsize_gpu.py), where the full oracle would need ~1e13 flop."""
import numpy as np

import oracle
from util import splitmix64


def _setup():
    import helfem_b200 as hb
    b = hb.DiatomicBasis(7, 7, 2.068, nelem=2, nnodes=6, primbas=4, lmax=3, mmax=2)
    b.compute_tei(True)
    dis = [[b.tei_view(w, r, e) for r in range(b.nlm) for e in range(b.nel)] for w in range(4)]
    tei = [b.tei_view(4 + ab, r, e) for r in range(b.nlm) for e in range(b.nel) for ab in range(4)]
    ktei = [b.tei_view(8 + ab, r, e) for r in range(b.nlm) for e in range(b.nel) for ab in range(4)]
    o = oracle.DiatomicOracle(b.lval, b.mval, b.first, b.last, b.Lmap, b.Mmap, b.Rhalf, dis, tei, ktei, b.pure_indices())
    n = b.nbf
    R = splitmix64(n * n).reshape(n, n)
    C = splitmix64(n * 4, seed=7).reshape(n, 4) * np.exp(-0.05 * np.arange(n))[:, None]
    pure = b.pure_indices()
    m = np.array([b.mval[p // b.nrad] for p in pure])
    # m-diagonal density with one exactly-zero m block (the norm screen of the exchange loop must see it)
    P = (C @ C.T + 1e-2 * 0.5 * (R + R.T)) * (m[:, None] == m[None, :]) * (np.abs(m)[:, None] < 2)
    return b, o, np.asfortranarray(P)


def test_tei_view_is_the_stored_block(native):
    b, o, P = _setup()
    for which, r, e in ((0, 0, 0), (3, b.nlm - 1, 1), (5, 2, 0), (11, b.nlm - 1, 1)):
        assert np.array_equal(b.tei_view(which, r, e), b.tei(which, r, e))


def test_block_oracle_equals_full_oracle(native):
    b, o, P = _setup()
    nang = b.nang
    pairs = [(j, k) for j in range(nang) for k in range(nang)]
    jl, kl = [p[0] for p in pairs], [p[1] for p in pairs]
    Kfull, Jfull = o.exchange(P), o.coulomb(P)
    Kb = o.exchange_blocks(P, jl, kl)
    Jb = o.coulomb_blocks(P, jl, kl)
    sk, sj = np.max(np.abs(Kfull)), np.max(np.abs(Jfull))
    assert sk > 1e-3 and sj > 1e-3
    nzero = 0
    for q, (j, k) in enumerate(pairs):
        kref, jref = o.pure_block(Kfull, j, k), o.pure_block(Jfull, j, k)
        # rows / columns of the removed boundary functions are not part of the pure matrix: compare the pure part
        mk = (np.array([b.mval[j] == 0] + [True] * (b.nrad - 1))[:, None]) & (np.array([b.mval[k] == 0] + [True] * (b.nrad - 1))[None, :])
        assert np.max(np.abs((Kb[q] - kref) * mk)) <= 1e-13 * sk, (j, k)
        assert np.max(np.abs((Jb[q] - jref) * mk)) <= 1e-13 * sj, (j, k)
        nzero += int(np.max(np.abs(kref)) == 0.0)
    assert 0 < nzero < len(pairs)  # some output blocks vanish identically (m_j != m_k), most do not
