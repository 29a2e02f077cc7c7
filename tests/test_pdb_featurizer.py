"""Host-side PDB featurizer (geobpe_b200/pdb.py): a chain built atom by atom from known internal
coordinates (oracle NeRF placement, nerf.py:84-128), written as PDB text, must featurize back to those coordinates in
the reference's row convention (angles_and_coords.py:104-152)."""
import numpy as np

import geobpe_oracle as O
from geobpe_b200 import pdb, synthetic

ATOL = 3e-3          # PDB coordinates carry 3 decimals


def _write_pdb(path, xyz, altloc=False):
    names = ["N", "CA", "C"]
    with open(path, "w") as f:
        f.write("HEADER    TEST\nMODEL        1\n")
        k = 1
        for r in range(xyz.shape[0]):
            for a in range(3):
                x, y, z = xyz[r, a]
                alt = "A" if (altloc and a == 1) else " "
                f.write("ATOM  %5d %-4s%s%3s %s%4d    %8.3f%8.3f%8.3f  1.00  0.00           %s\n"
                        % (k, " " + names[a], alt, "ALA", "A", r + 1, x, y, z, names[a][0]))
                k += 1
                if altloc and a == 1:        # a second alternate location that must be ignored
                    f.write("ATOM  %5d %-4s%s%3s %s%4d    %8.3f%8.3f%8.3f  0.50  0.00           C\n"
                            % (k, " CA", "B", "ALA", "A", r + 1, x + 1.0, y, z))
                    k += 1
            f.write("HETATM%5d  O   HOH A%4d    %8.3f%8.3f%8.3f  1.00  0.00           O\n" % (k, 900 + r, 0.0, 0.0, 0.0))
        f.write("ENDMDL\nMODEL        2\nATOM      1  N   ALA A   1       9.000   9.000   9.000  1.00  0.00           N\nENDMDL\n")


def _build(vals):
    """atoms from the 9 columns of one chain (row convention of synthetic.generate)"""
    C = synthetic.COL
    n = vals.shape[1]
    pts = [np.array([0.0, 0.0, 0.0]), np.array([1.458, 0.0, 0.0]), np.array([2.0, 1.42, 0.0])]
    for r in range(n - 1):
        for ang, ln, dih in ((vals[C["CA:C:1N"], r], vals[C["0C:1N"], r], vals[C["psi"], r]),
                             (vals[C["C:1N:1CA"], r], vals[C["N:CA"], r], vals[C["omega"], r]),
                             (vals[C["tau"], r], vals[C["CA:C"], r], vals[C["phi"], r + 1])):
            pts.append(O.place_dihedral(pts[-3], pts[-2], pts[-1], bond_angle=ang, bond_length=ln, torsion_angle=dih))
    return np.array(pts).reshape(n, 3, 3)


def test_round_trip(tmp_path):
    vals, offs = synthetic.generate(3, 40, seed=5, ragged=True)
    files = []
    for k in range(3):
        v = vals[:, offs[k]:offs[k + 1]]
        p = tmp_path / f"c{k}.pdb"
        _write_pdb(p, _build(v), altloc=(k == 1))
        files.append(p)
    got, goffs, names = pdb.featurize_files(files)
    assert np.array_equal(goffs, offs) and names == ["c0.pdb", "c1.pdb", "c2.pdb"]
    for k in range(3):
        a, b = offs[k], offs[k + 1]
        w, g = vals[:, a:b], got[:, a:b]
        assert np.array_equal(np.isnan(w[:, 1:]), np.isnan(g[:, 1:]))          # padding convention, rows 1..
        assert np.isnan(g[synthetic.COL["phi"], 0])
        for c in ("0C:1N", "N:CA", "CA:C"):
            assert g[synthetic.COL[c], -1] == 0.0
        sl = slice(1, b - a - 1)                                                   # rows fully determined by the atoms
        for c in synthetic.COLUMNS:
            d = g[synthetic.COL[c], sl] - w[synthetic.COL[c], sl]
            if c in ("phi", "psi", "omega"):
                d = (d + np.pi) % (2 * np.pi) - np.pi
            assert np.abs(d).max() < ATOL, c


def test_short_and_incomplete_chains(tmp_path):
    vals, offs = synthetic.generate(1, 12, seed=1)
    xyz = _build(vals)
    p = tmp_path / "x.pdb"
    _write_pdb(p, xyz[:5])
    assert pdb.featurize_files([p])[1].tolist() == [0]                             # < min_len residues: skipped
    _write_pdb(p, xyz)
    lines = [l for l in open(p) if not (l.startswith("ATOM") and l[12:16].strip() == "CA" and int(l[22:26]) == 4)]
    open(p, "w").writelines(lines)
    assert pdb.read_backbone(p).shape[0] == 11                                     # residue 4 lacks CA: dropped
