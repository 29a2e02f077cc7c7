"""Frozen synthetic-backbone generator.

The reference ships no generator; BASELINE.json's configs 2-5 name "synthetic backbones", so
this file *defines* them. Oracle, tests and bench read identical arrays from here.

Layout produced: `vals` float64 [9, R] (type-major SoA, column order = COLUMNS below, which is
the reference DataFrame's column order, datasets.py:175-186 / SURVEY Appendix C) and `offsets`
int64 [S+1] (CSR over residues).  Row r of a backbone holds N:CA, CA:C, tau of residue r+1;
0C:1N, CA:C:1N, C:1N:1CA, psi, omega of the r->r+1 junction; phi of residue r (row 0 NaN);
the last row is 0 for lengths and NaN for angles (angles_and_coords.py:104-152).
"""
import numpy as np

COLUMNS = ["0C:1N", "N:CA", "CA:C", "phi", "psi", "omega", "tau", "CA:C:1N", "C:1N:1CA"]
COL = {c: i for i, c in enumerate(COLUMNS)}

# (mean phi, mean psi, sigma) per secondary-structure state; coil is a 3-centre mixture
_HELIX = (-1.05, -0.79, 0.15)
_STRAND = (-2.09, 2.27, 0.15)
_COIL = [(-1.20, 2.40, 0.40), (1.00, 0.60, 0.40), (-1.50, -0.30, 0.40)]
_STAY = np.array([0.90, 1.0 - 1.0 / 6.0, 0.75])  # mean run 10 / 6 / 4


def _wrap(a):
    """wrap to (-pi, pi]"""
    return -((-a + np.pi) % (2 * np.pi) - np.pi)


def generate(num_backbones, n_res, seed=0, ragged=False):
    """Return (vals [9, R] float64, offsets [S+1] int64).

    ragged=True draws per-backbone lengths uniformly in [max(8, n_res//2), n_res].
    """
    rng = np.random.default_rng(seed)
    S = int(num_backbones)
    if ragged:
        lens = rng.integers(max(8, n_res // 2), n_res + 1, size=S)
    else:
        lens = np.full(S, n_res, dtype=np.int64)
    n = int(lens.max())
    # -- secondary-structure Markov chain, vectorised over backbones
    state = np.empty((S, n), dtype=np.int8)
    state[:, 0] = rng.integers(0, 3, size=S)
    for r in range(1, n):
        prev = state[:, r - 1]
        leave = rng.random(S) >= _STAY[prev]
        hop = rng.integers(1, 3, size=S)  # +1 or +2 (mod 3)
        state[:, r] = np.where(leave, (prev + hop) % 3, prev)
    coil_c = rng.integers(0, 3, size=(S, n))
    mu_phi = np.where(state == 0, _HELIX[0], np.where(state == 1, _STRAND[0], np.take([c[0] for c in _COIL], coil_c)))
    mu_psi = np.where(state == 0, _HELIX[1], np.where(state == 1, _STRAND[1], np.take([c[1] for c in _COIL], coil_c)))
    sig = np.where(state == 2, 0.40, 0.15)
    cols = np.empty((9, S, n), dtype=np.float64)
    cols[COL["phi"]] = _wrap(mu_phi + sig * rng.standard_normal((S, n)))
    cols[COL["psi"]] = _wrap(mu_psi + sig * rng.standard_normal((S, n)))
    cols[COL["omega"]] = _wrap(np.pi + 0.05 * rng.standard_normal((S, n)))
    cols[COL["tau"]] = 1.94 + 0.04 * rng.standard_normal((S, n))
    cols[COL["CA:C:1N"]] = 2.03 + 0.03 * rng.standard_normal((S, n))
    cols[COL["C:1N:1CA"]] = 2.12 + 0.03 * rng.standard_normal((S, n))
    cols[COL["N:CA"]] = 1.458 + 0.01 * rng.standard_normal((S, n))
    cols[COL["CA:C"]] = 1.525 + 0.01 * rng.standard_normal((S, n))
    cols[COL["0C:1N"]] = 1.329 + 0.01 * rng.standard_normal((S, n))
    # -- reference padding convention
    ridx = np.arange(n)[None, :]
    last = ridx == (lens[:, None] - 1)
    for c in ("0C:1N", "N:CA", "CA:C"):
        cols[COL[c]][last] = 0.0
    for c in ("psi", "omega", "tau", "CA:C:1N", "C:1N:1CA"):
        cols[COL[c]][last] = np.nan
    cols[COL["phi"]][:, 0] = np.nan
    # phi of the last row is a real value (phi of the last residue) -- keep it.
    keep = ridx < lens[:, None]
    vals = np.ascontiguousarray(cols[:, keep])
    offsets = np.zeros(S + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    return vals, offsets


def to_structures(vals, offsets, prefix="syn"):
    """SoA arrays -> list of reference-style structure dicts (datasets.py:145-154)."""
    import pandas as pd

    out = []
    for k in range(len(offsets) - 1):
        a, b = int(offsets[k]), int(offsets[k + 1])
        n = b - a
        df = pd.DataFrame({c: vals[i, a:b].copy() for i, c in enumerate(COLUMNS)})
        out.append({
            "angles": df,
            "coords": None,
            "c_beta": None,
            "full_idxes": sum([[i, i, i] for i in range(1, n + 1)], []),
            "full_coords": None,
            "side_chain": None,
            "aa": None,
            "fname": f"{prefix}{k}",
        })
    return out


def from_structures(structures):
    """list of structure dicts -> (vals [9, R], offsets [S+1])."""
    lens = [len(s["angles"]) for s in structures]
    offsets = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    vals = np.empty((9, int(offsets[-1])), dtype=np.float64)
    for k, s in enumerate(structures):
        df = s["angles"]
        a, b = int(offsets[k]), int(offsets[k + 1])
        for i, c in enumerate(COLUMNS):
            vals[i, a:b] = np.asarray(df[c], dtype=np.float64)
    return vals, offsets
