"""Shared helpers of the test-suite: deterministic synthetic densities."""
import numpy as np
import scipy.linalg

SEED = 20261104


def splitmix64(n, seed=SEED):
    """n uniform doubles in (-1, 1) from splitmix64 (portable, no numpy RNG dependence)."""
    out = np.empty(n)
    x = seed & 0xFFFFFFFFFFFFFFFF
    mask = 0xFFFFFFFFFFFFFFFF
    for i in range(n):
        x = (x + 0x9E3779B97F4A7C15) & mask
        z = x
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & mask
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & mask
        z = z ^ (z >> 31)
        out[i] = (z >> 11) / float(1 << 53) * 2.0 - 1.0
    return out


# ground-state configurations by l-channel occupations (number of electrons in s, p, d, f)
def aufbau_occs(Z, lmax):
    order = [(1, 0), (2, 0), (2, 1), (3, 0), (3, 1), (4, 0), (3, 2), (4, 1), (5, 0), (4, 2), (5, 1), (6, 0), (4, 3), (5, 2),
             (6, 1), (7, 0), (5, 3), (6, 2), (7, 1)]
    occ = [0] * (lmax + 1)
    left = Z
    for n, l in order:
        if l > lmax or left <= 0:
            continue
        k = min(left, 2 * (2 * l + 1))
        occ[l] += k
        left -= k
    return occ


def sadatom_density(S, T, Tl, Vnuc, Z, lmax, perturb=1e-3, spin=None, seed=SEED):
    """Density cube P_l (nbf x nbf x (lmax+1), Fortran order) from core-Hamiltonian orbitals filled by
    Aufbau with sadatom's fractional shell filling (sadatom/solver.cpp:386-398), plus a seeded symmetric
    perturbation so no block is accidentally zero.  spin: None (total), 'a' or 'b' (b has the last
    shell of the highest l emptied)."""
    n = S.shape[0]
    P = np.zeros((n, n, lmax + 1), order="F")
    occs = aufbau_occs(Z, lmax)
    for l in range(lmax + 1):
        H = T + l * (l + 1) * Tl + Vnuc
        E, C = scipy.linalg.eigh(H, S)
        numl = occs[l]
        cap = 2 * (2 * l + 1)
        if spin is not None:
            cap //= 2
            numl = (numl + (1 if spin == "a" else 0)) // 2
        io = 0
        while numl > 0:
            k = min(cap, numl)
            P[:, :, l] += k * np.outer(C[:, io], C[:, io])
            numl -= k
            io += 1
    if perturb:
        scale = perturb * np.max(np.abs(P))
        for l in range(lmax + 1):
            R = splitmix64(n * n, seed + l + (7 if spin == "b" else 0)).reshape(n, n)
            # keep the perturbation inside the element band so it is what every operator sees
            P[:, :, l] += scale * 1e-3 * (R + R.T) * (np.abs(P[:, :, l]) > 0)
    return P
