"""GPU parity tests of DFTGrid::eval_Fxc of the `atomic` and `diatomic` programs (run with -m gpu): the CUDA
product-grid path (helfem_b200/csrc/grid2d.cu) through the host mirror classes / C ABI vs the CPU oracle
(oracle/grid2d_oracle.cpp, the reference's loop structure with complex basis-function matrices) on the same seeded
densities.  Atomic radial tables of the oracle come from the numpy chain (oracle/fem_oracle.py), the spherical harmonics
from scipy; the product builds both with its own C++ host library.  Gates: max |dH| <= 1e-9 (relaxed to 5e-14 max|H| for
huge tail potentials as in the sadatom suite), |dExc|, |dNel|, |dEkin| <= 1e-10 relative to max(1, |value|)."""
import numpy as np
import pytest

import oracle
from grid2d_cases import ATOMIC, DIATOMIC, atomic_oracle, atomic_density, diatomic_density

pytestmark = pytest.mark.gpu

FUNCS = {"lda": (1, 12), "pbe": (101, 130), "r2scan": (497, 498)}
_cache = {}


def tol(ref):
    return max(1e-9, 5e-14 * float(np.max(np.abs(ref))))


def close(a, b):
    return abs(a - b) <= 1e-10 * max(1.0, abs(b))


def atomic_case(name):
    if ("a", name) not in _cache:
        import helfem_b200 as hb
        Z, primbas, nnodes, nelem, lmax, mmax, nocc, lang, mang = ATOMIC[name]
        b = hb.AtomicBasis(Z=Z, nelem=nelem, nnodes=nnodes, primbas=primbas, lmax=lmax, mmax=mmax)
        g = hb.AtomicDFTGrid(b, lang, mang)
        o, lval, mval, rs = atomic_oracle(name)
        assert list(lval) == list(b.lval) and list(mval) == list(b.mval)
        _cache[("a", name)] = (b, g, o, atomic_density(name), atomic_density(name, "a"), atomic_density(name, "b"))
    return _cache[("a", name)]


def diatomic_case(name):
    if ("d", name) not in _cache:
        import helfem_b200 as hb
        Z1, Z2, Rb, primbas, nnodes, nelem, lmax, mmax, lang, mang = DIATOMIC[name]
        b = hb.DiatomicBasis(Z1, Z2, Rb, nelem=nelem, nnodes=nnodes, primbas=primbas, lmax=lmax, mmax=mmax)
        g = hb.DiatomicDFTGrid(b, lang, mang)
        tabs = [b.radial(e) for e in range(b.nel)]
        o = oracle.Grid2DOracle(1, b.lval, b.mval, b.first, b.last, [t[0] for t in tabs], [t[1] for t in tabs],
                                [t[2] for t in tabs], [t[3] for t in tabs], lang, mang, Rhalf=b.Rhalf, pure=b.pure_indices())
        n = b.nbf
        _cache[("d", name)] = (b, g, o, diatomic_density(n), 0.5 * diatomic_density(n, "a"), 0.5 * diatomic_density(n, "b"))
    return _cache[("d", name)]


def check_restricted(g, o, P, funcs):
    H, exc, nel, ekin = g.eval_Fxc(*funcs, P)
    Ho, exco, nelo, ekino = o.eval_fxc(*funcs, P)
    d = float(np.max(np.abs(H - Ho)))
    assert d <= tol(Ho), (d, float(np.max(np.abs(Ho))))
    assert close(exc, exco) and close(nel, nelo) and close(ekin, ekino), (exc - exco, nel - nelo, ekin - ekino)
    return H


def check_unrestricted(g, o, Pa, Pb, funcs, beta=True):
    Ha, Hb, exc, nel, ekin = g.eval_Fxc(*funcs, Pa, Pb, beta=beta)
    Hao, Hbo, exco, nelo, ekino = o.eval_fxc(*funcs, Pa, Pb, beta=beta)
    for H, Ho in ((Ha, Hao), (Hb, Hbo)):
        d = float(np.max(np.abs(H - Ho)))
        assert d <= tol(Ho), (d, float(np.max(np.abs(Ho))))
    assert close(exc, exco) and close(nel, nelo) and close(ekin, ekino), (exc - exco, nel - nelo, ekin - ekino)


@pytest.mark.parametrize("fn", list(FUNCS))
@pytest.mark.parametrize("case", list(ATOMIC))
def test_atomic_grid_restricted_parity(native, case, fn):
    b, g, o, P, Pa, Pb = atomic_case(case)
    H = check_restricted(g, o, P, FUNCS[fn])
    assert np.max(np.abs(H - H.T)) <= tol(H)


@pytest.mark.parametrize("fn", list(FUNCS))
@pytest.mark.parametrize("case", ["ne_l2m2", "c_hip_l1m1"])
def test_atomic_grid_unrestricted_parity(native, case, fn):
    b, g, o, P, Pa, Pb = atomic_case(case)
    check_unrestricted(g, o, Pa, Pb, FUNCS[fn])


def test_atomic_grid_beta_false_skips_beta_fock(native):
    b, g, o, P, Pa, Pb = atomic_case("c_hip_l1m1")
    Ha, Hb, exc, nel, ekin = g.eval_Fxc(*FUNCS["pbe"], Pa, Pb, beta=False)
    Ha2, Hb2, exc2, nel2, ekin2 = g.eval_Fxc(*FUNCS["pbe"], Pa, Pb, beta=True)
    assert np.array_equal(Ha, Ha2) and exc == exc2 and not Hb.any() and Hb2.any()


def test_atomic_grid_is_deterministic_and_x_only_c_only_add_up(native):
    b, g, o, P, Pa, Pb = atomic_case("ne_l2m2")
    H1, e1, n1, k1 = g.eval_Fxc(101, 130, P)
    H2, e2, n2, k2 = g.eval_Fxc(101, 130, P)
    assert np.array_equal(H1, H2) and e1 == e2 and n1 == n2
    Hx, ex, _, _ = g.eval_Fxc(101, 0, P)
    Hc, ec, _, _ = g.eval_Fxc(0, 130, P)
    assert np.max(np.abs(Hx + Hc - H1)) <= tol(H1) and close(ex + ec, e1)


def test_atomic_grid_matches_sadatom_path_in_the_spherical_limit(native):
    """the two CUDA grid paths (product grid and radial-only sadatom kernel) agree for a spherical density"""
    import helfem_b200 as hb
    Z, primbas, nnodes, nelem, lmax, mmax, nocc, lang, mang = ATOMIC["ne_l2m2"]
    b, g, o, P, Pa, Pb = atomic_case("ne_l2m2")
    sb = hb.SadatomBasis(Z=Z, nelem=nelem, nnodes=nnodes, primbas=primbas, lmax=lmax)
    nr = b.nrad
    rng = np.random.RandomState(3)
    Pl = np.zeros((nr, nr, lmax + 1), order="F")
    for l in range(lmax + 1):
        c = rng.uniform(-1, 1, (nr, 2)) * np.exp(-0.4 * np.arange(nr))[:, None]
        Pl[:, :, l] = (2 * l + 1) * (c @ c.T)
    Pfull = np.zeros((b.nbf, b.nbf), order="F")
    for a, l in enumerate(b.lval):
        Pfull[a * nr:(a + 1) * nr, a * nr:(a + 1) * nr] = Pl[:, :, l] / (2 * l + 1)
    H, exc, nel, ekin = g.eval_Fxc(*FUNCS["r2scan"], Pfull)
    Hs, excs, nels = sb.eval_fxc(*FUNCS["r2scan"], Pl / (4 * np.pi))
    assert close(exc, excs) and close(nel, nels)
    for a, l in enumerate(b.lval):
        blk = H[a * nr:(a + 1) * nr, a * nr:(a + 1) * nr]
        assert np.max(np.abs(blk - Hs[:, :, l] / (4 * np.pi))) <= tol(blk)


def test_atomic_grid_errors(native):
    import helfem_b200 as hb
    b, g, o, P, Pa, Pb = atomic_case("ne_l0")
    with pytest.raises(ValueError):       # std::logic_error: size mismatch
        g.eval_Fxc(1, 12, np.zeros((3, 3)))
    with pytest.raises(RuntimeError):     # functional without a device implementation: no CPU fallback
        g.eval_Fxc(106, 131, P)


@pytest.mark.parametrize("fn", list(FUNCS))
@pytest.mark.parametrize("case", list(DIATOMIC))
def test_diatomic_grid_restricted_parity(native, case, fn):
    b, g, o, P, Pa, Pb = diatomic_case(case)
    check_restricted(g, o, P, FUNCS[fn])


@pytest.mark.parametrize("fn", ["pbe", "r2scan"])
@pytest.mark.parametrize("case", ["n2_l2m1"])
def test_diatomic_grid_unrestricted_parity(native, case, fn):
    b, g, o, P, Pa, Pb = diatomic_case(case)
    check_unrestricted(g, o, Pa, Pb, FUNCS[fn])


@pytest.mark.parametrize("fn", ["pbe", "r2scan"])
def test_diatomic_grid_m_diagonal_density(native, fn):
    """Density with m_a = m_b blocks only (what an SCF on a linear molecule produces): rho and every field are independent of
    phi, the product returns the m_a = m_b blocks of H and exact zeros elsewhere (the reference: round-off noise there)."""
    b, g, o, P, Pa, Pb = diatomic_case("n2_l2m1")
    pure = b.pure_indices()
    m_of = np.repeat(np.asarray(b.mval), b.nrad)[pure]
    Pm = np.asfortranarray(np.where(m_of[:, None] == m_of[None, :], P, 0.0))
    H = check_restricted(g, o, Pm, FUNCS[fn])
    assert float(np.max(np.abs(H[m_of[:, None] != m_of[None, :]]))) == 0.0
    assert float(np.max(np.abs(H))) > 1e-3


@pytest.mark.parametrize("fn", ["pbe", "r2scan"])
def test_general_angular_rule_kernels(native, fn, monkeypatch):
    """HFB200_G2D_GENERAL=1 at set_grid time keeps the pair-table kernels that serve an angular rule without the theta x phi
    product structure; they must agree with the oracle (and hence with the factorised kernels) on the same inputs."""
    import helfem_b200 as hb
    monkeypatch.setenv("HFB200_G2D_GENERAL", "1")
    Z1, Z2, Rb, primbas, nnodes, nelem, lmax, mmax, lang, mang = DIATOMIC["n2_l2m1"]
    bd = hb.DiatomicBasis(Z1, Z2, Rb, nelem=nelem, nnodes=nnodes, primbas=primbas, lmax=lmax, mmax=mmax)
    gd = hb.DiatomicDFTGrid(bd, lang, mang)
    monkeypatch.delenv("HFB200_G2D_GENERAL")
    b, g, o, P, Pa, Pb = diatomic_case("n2_l2m1")
    check_restricted(gd, o, P, FUNCS[fn])
    check_unrestricted(gd, o, Pa, Pb, FUNCS[fn])
    name = "ne_l2m2"
    Z, primbas, nnodes, nelem, lmax, mmax, nocc, lang, mang = ATOMIC[name]
    monkeypatch.setenv("HFB200_G2D_GENERAL", "1")
    ba = hb.AtomicBasis(Z=Z, nelem=nelem, nnodes=nnodes, primbas=primbas, lmax=lmax, mmax=mmax)
    ga = hb.AtomicDFTGrid(ba, lang, mang)
    monkeypatch.delenv("HFB200_G2D_GENERAL")
    b, g, o, P, Pa, Pb = atomic_case(name)
    check_restricted(ga, o, P, FUNCS[fn])
