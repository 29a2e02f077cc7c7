"""PDB backbone featurizer: ATOM records -> the 9 internal-coordinate columns the learner consumes.

Numpy restatement of what the reference obtains from biotite (angles_and_coords.py:70-152
`canonical_distances_and_dihedrals` with EXHAUSTIVE_ANGLES / EXHAUSTIVE_DISTS): row r of a chain holds
phi of residue r (NaN for r = 0); psi, omega, CA:C:1N, C:1N:1CA, 0C:1N of the r -> r+1 junction; tau, N:CA, CA:C of
residue r+1; the last row carries the reference's padding (0 for lengths, NaN for angles).  Host-side input format
only -- the hot path starts at the arrays this returns.  Parity with biotite itself is not pinned (biotite is not in
this image): reference and CUDA path are both fed the arrays produced here, as tests/golden/pdb72_*.json records.
"""
import gzip
import os

import numpy as np

from .synthetic import COL, COLUMNS

_AA = {"ALA", "ARG", "ASN", "ASP", "CYS", "GLN", "GLU", "GLY", "HIS", "ILE", "LEU", "LYS", "MET", "PHE", "PRO", "SER",
       "THR", "TRP", "TYR", "VAL", "MSE", "SEC", "PYL", "ASX", "GLX", "UNK"}


def read_backbone(fname):
    """first model, first alternate location, ATOM records only -> float64 [n_res, 3 (N, CA, C), 3]; residues missing a
    backbone atom are dropped"""
    opener = gzip.open if str(fname).endswith(".gz") else open
    res, order = {}, []
    with opener(str(fname), "rt") as f:
        for line in f:
            rec = line[:6]
            if rec == "ENDMDL":
                break
            if rec != "ATOM  ":
                continue
            name = line[12:16].strip()
            if name not in ("N", "CA", "C") or line[17:20].strip() not in _AA:
                continue
            key = (line[21], line[22:27])          # chain, residue number + insertion code
            if key not in res:
                res[key] = {}
                order.append(key)
            if name not in res[key]:               # first altloc wins
                res[key][name] = (float(line[30:38]), float(line[38:46]), float(line[46:54]))
    rows = [[res[k]["N"], res[k]["CA"], res[k]["C"]] for k in order if len(res[k]) == 3]
    return np.asarray(rows, dtype=np.float64).reshape(-1, 3, 3)


def _dihedral(a, b, c, d):
    """IUPAC sign convention, radians in (-pi, pi]"""
    b0, b1, b2 = b - a, c - b, d - c
    n1, n2 = np.cross(b0, b1), np.cross(b1, b2)
    y = (np.cross(n1, n2) * (b1 / np.linalg.norm(b1, axis=-1, keepdims=True))).sum(-1)
    return np.arctan2(y, (n1 * n2).sum(-1))


def _angle(a, b, c):
    u, v = a - b, c - b
    cs = (u * v).sum(-1) / (np.linalg.norm(u, axis=-1) * np.linalg.norm(v, axis=-1))
    return np.arccos(np.clip(cs, -1.0, 1.0))


def featurize_backbone(xyz):
    """[n, 3, 3] backbone coordinates -> float64 [9, n] in COLUMNS order"""
    n = xyz.shape[0]
    N, CA, C = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    out = np.full((9, n), np.nan)
    out[COL["phi"], 1:] = _dihedral(C[:-1], N[1:], CA[1:], C[1:])
    out[COL["psi"], :-1] = _dihedral(N[:-1], CA[:-1], C[:-1], N[1:])
    out[COL["omega"], :-1] = _dihedral(CA[:-1], C[:-1], N[1:], CA[1:])
    out[COL["tau"], :-1] = _angle(N[1:], CA[1:], C[1:])
    out[COL["CA:C:1N"], :-1] = _angle(CA[:-1], C[:-1], N[1:])
    out[COL["C:1N:1CA"], :-1] = _angle(C[:-1], N[1:], CA[1:])
    out[COL["0C:1N"], :-1] = np.linalg.norm(N[1:] - C[:-1], axis=-1)
    out[COL["N:CA"], :-1] = np.linalg.norm(CA[1:] - N[1:], axis=-1)
    out[COL["CA:C"], :-1] = np.linalg.norm(C[1:] - CA[1:], axis=-1)
    for c in ("0C:1N", "N:CA", "CA:C"):
        out[COL[c], -1] = 0.0
    return out


def featurize_files(fnames, min_len=8):
    """-> (vals [9, R] float64, offsets [S+1] int64, names kept); chains shorter than min_len residues are skipped"""
    cols, lens, kept = [], [], []
    for fn in fnames:
        xyz = read_backbone(fn)
        if xyz.shape[0] < min_len:
            continue
        cols.append(featurize_backbone(xyz))
        lens.append(xyz.shape[0])
        kept.append(os.path.basename(str(fn)))
    offsets = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    vals = np.ascontiguousarray(np.concatenate(cols, axis=1)) if cols else np.zeros((9, 0))
    return vals, offsets, kept


__all__ = ["COLUMNS", "read_backbone", "featurize_backbone", "featurize_files"]
