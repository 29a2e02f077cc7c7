"""TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of the reference's GeoBPE hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this file; the product path (pt-bpe_b200/) never does.

This is a plain-Python/numpy restatement of shiningsunnyday/PT-BPE @ f032fa4, covering
`foldingdiff/bpe.py` (learn + induce), `foldingdiff/tokenizer.py`, `foldingdiff/nerf.py`,
`foldingdiff/algo.py` and `plotting.save_histogram`.  Every function cites the reference
file:line it follows.  It keeps the reference's *data-structure semantics* (dict of Python
`set`s of `(structure, bond)` tuples, `SortedDict` of `(flag, -count, json_key)`, JSON string
keys, numpy RNG stream, float64 numpy geometry) because results depend on them (SURVEY.md
section 7.2); it drops pandas (per-backbone columns are plain Python lists of floats) and the
fork pools (assignments are computed in-process, same values).

Parity pin: tests/golden/*.json were produced by running the REAL reference in the build
container through oracle/ref_shim.py (generator: oracle/make_golden.py);
tests/test_oracle_vs_golden.py checks this restatement against them bit-for-bit (token tables,
segmentations, pop order, thresholds) -- so parity is PINNED to reference outputs, not to
reference-held golden vectors (the reference has none).

Not restated (out of the path's scope this round): glue optimisation (LBFGS), secondary
structure priority, bond-level (res_init=False) tokens, plotting.
"""
import bisect
import json
import math
from collections import defaultdict

import numpy as np
from sortedcontainers import SortedDict

BOND_TYPES = ["N:CA", "CA:C", "0C:1N"]                # tokenizer.py:19
BOND_ANGLES = ["tau", "CA:C:1N", "C:1N:1CA"]          # tokenizer.py:21
DIHEDRAL_ANGLES = ["psi", "omega", "phi"]             # tokenizer.py:22
BOND_LENGTHS = [1.46, 1.54, 1.34]                     # nerf.py:17-19
COLUMNS = ["0C:1N", "N:CA", "CA:C", "phi", "psi", "omega", "tau", "CA:C:1N", "C:1N:1CA"]
N_INIT = np.array([17.047, 14.099, 3.625])            # nerf.py:22-24
CA_INIT = np.array([16.967, 12.784, 4.338])
C_INIT = np.array([15.685, 12.755, 5.133])
TWO_PI = 2 * np.pi


# ----------------------------------------------------------------------------- containers
class ThresholdDict(dict):
    """data_structures.py:264-296 -- floor lookup on int keys."""

    def __getitem__(self, key):
        if key in self:
            return dict.__getitem__(self, key)
        if isinstance(key, int):
            ks = sorted(k for k in self if isinstance(k, int))
            i = bisect.bisect_right(ks, key) - 1
            if i >= 0:
                return dict.__getitem__(self, ks[i])
        raise KeyError(key)


def num_bonds(geo):                                   # tokenizer.py:302-304
    return len(geo.get("N:CA", [])) + len(geo.get("CA:C", [])) + len(geo.get("0C:1N", []))


def hash_geo(geo):                                    # bpe.py:1147-1149
    return json.dumps(geo, sort_keys=True)


def get_ind(v, values):                               # bpe.py:1165-1189
    left_edges = [start for start, _ in values]
    ind = bisect.bisect_right(left_edges, v) - 1
    if ind < 0:
        raise ValueError(f"value {v} is below the first bin range")
    start, end = values[ind]
    if ind == len(values) - 1 and v == end:
        return ind
    if start <= v < end:
        return ind
    raise ValueError(f"value {v} does not fall into any bin")


# ----------------------------------------------------------------------------- geometry
def angle_between(v1, v2):                            # angles_and_coords.py:746-752
    u1 = v1 / np.linalg.norm(v1)
    u2 = v2 / np.linalg.norm(v2)
    return np.arccos(np.clip(np.dot(u1, u2), -1.0, 1.0))


def rotate_vector(v, k, angle):                       # angles_and_coords.py:232-235
    return v * np.cos(angle) + np.cross(k, v) * np.sin(angle) + k * np.dot(k, v) * (1 - np.cos(angle))


def update_backbone_positions(N_init, CA_init, C_init, L_CA_C, L_N_CA, new_theta):
    """angles_and_coords.py:238-317"""
    N_init = np.array(N_init, dtype=float)
    CA_init = np.array(CA_init, dtype=float)
    C_init = np.array(C_init, dtype=float)
    v = CA_init - C_init
    v = v / np.linalg.norm(v)
    CA_new = C_init + L_CA_C * v
    vec_N = N_init - CA_new
    vec_C = C_init - CA_new
    cos_theta = np.dot(vec_N, vec_C) / (np.linalg.norm(vec_N) * np.linalg.norm(vec_C))
    cos_theta = np.clip(cos_theta, -1.0, 1.0)
    dtheta = new_theta - np.arccos(cos_theta)
    axis = np.cross(vec_N, vec_C)
    axis = axis / np.linalg.norm(axis)
    rotated = rotate_vector(vec_N, axis, -dtheta)
    rotated = rotated / np.linalg.norm(rotated) * L_N_CA
    return CA_new + rotated, CA_new, C_init.copy()


def place_dihedral(a, b, c, bond_angle, bond_length, torsion_angle):
    """nerf.py:151-210 (numpy branch)"""
    ab = b - a
    bc = (c - b) / np.linalg.norm(c - b, axis=-1)
    n = np.cross(ab, bc, axis=-1)
    n = n / np.linalg.norm(n, axis=-1)
    nbc = np.cross(n, bc, axis=-1)
    m = np.stack([bc, nbc, n], axis=-1)
    d = np.stack([
        -bond_length * np.cos(bond_angle),
        bond_length * np.cos(torsion_angle) * np.sin(bond_angle),
        bond_length * np.sin(torsion_angle) * np.sin(bond_angle),
    ], axis=a.ndim - 1)
    return m.dot(d) + c


def geo_nerf_coords(geo):
    """tokenizer.py:316-344 + nerf.py:84-128: coordinates for 3m-1 bonds (whole residues)."""
    nb = num_bonds(geo)
    assert nb % 3 == 2
    n_init, ca_init, c_init = update_backbone_positions(
        N_INIT, CA_INIT, C_INIT, geo["CA:C"][0], geo["N:CA"][0], geo["tau"][0])
    coords = [n_init, ca_init, c_init]
    if nb == 2:
        return np.array(coords)
    phi = np.array([np.nan] + geo["phi"])[1:]
    psi = np.array(geo["psi"] + [np.nan])[:-1]
    omega = np.array(geo["omega"] + [np.nan])[:-1]
    dih = np.stack([psi, omega, phi]).T
    lens = [np.array(geo["0C:1N"]), np.array(geo["N:CA"][1:]), np.array(geo["CA:C"][1:])]
    angs = [np.array(geo["CA:C:1N"]), np.array(geo["C:1N:1CA"]), np.array(geo["tau"][1:])]
    for i in range(dih.shape[0]):
        for j in range(3):  # C->N, N->CA, CA->C   (nerf.py:113-124)
            coords.append(place_dihedral(coords[-3], coords[-2], coords[-1],
                                         bond_angle=angs[j][i], bond_length=lens[j][i],
                                         torsion_angle=dih[i][j]))
    return np.array(coords)


def kabsch(P, Q):                                     # algo.py:8-45
    cP = np.mean(P, axis=0)
    cQ = np.mean(Q, axis=0)
    Pc = P - cP
    Qc = Q - cQ
    H = np.dot(Pc.T, Qc)
    U, S, Vt = np.linalg.svd(H)
    R = np.dot(U, Vt)
    if np.linalg.det(R) < 0:
        Vt[2, :] *= -1
        R = np.dot(U, Vt)
    return (Q - cQ) @ R.T + cP


def compute_rmsd(P, Q):                               # algo.py:48-65
    diff = P - kabsch(P, Q)
    return np.sqrt(np.mean(np.sum(diff ** 2, axis=1)))


def k_medoids(strucs, k, max_iterations=10, tol=1e-4, rng=None):
    """algo.py:144-211 (distance matrix computed serially; same values)."""
    N = len(strucs)
    k = min(N, k)
    if k == N:
        return list(range(N))
    if rng is None:
        rng = np.random.default_rng(None)
    D = np.empty((N, N), dtype=np.float32)
    for i in range(N):
        for j in range(i, N):
            D[i, j] = D[j, i] = compute_rmsd(strucs[i], strucs[j])
    return k_medoids_from_D(D, k, rng, max_iterations, tol)


def k_medoids_from_D(D, k, rng, max_iterations=10, tol=1e-4):
    """algo.py:184-211: the PAM-style loop on a float32 distance matrix."""
    N = D.shape[0]
    medoid_indices = rng.choice(np.arange(N), size=k, replace=False)
    assignments = np.zeros(N, dtype=int)
    for _ in range(max_iterations):
        for i in range(N):
            assignments[i] = np.argmin(D[i, medoid_indices])
        total_shift = 0.0
        new_medoid_indices = []
        for j in range(k):
            members = np.where(assignments == j)[0]
            if members.size == 0:
                new_idx = rng.integers(N)
            else:
                intra = D[np.ix_(members, members)].sum(axis=1)
                new_idx = members[np.argmin(intra)]
            total_shift += D[medoid_indices[j], new_idx]
            new_medoid_indices.append(new_idx)
        medoid_indices = new_medoid_indices
        if total_shift < tol:
            break
    return medoid_indices


# ----------------------------------------------------------------------------- tokenizer
class Tokenizer:
    """tokenizer.py:18-477, per-backbone internal-coordinate store (pandas dropped)."""

    def __init__(self, cols, fname=None):
        # tokenizer.py:24-61 : every cell becomes a Python float; orig copy kept pristine
        self.cur = {c: [float(x) for x in cols[c]] for c in COLUMNS}
        self.orig = {c: list(v) for c, v in self.cur.items()}
        self.n = len(self.cur["phi"])
        self.fname = fname
        self._init_n_ca = np.linalg.norm(N_INIT - CA_INIT)             # tokenizer.py:74-77
        self._init_ca_c = np.linalg.norm(CA_INIT - C_INIT)
        self._init_bond_angle = angle_between(N_INIT - CA_INIT, C_INIT - CA_INIT)
        self.token_pos = None
        self.tokens = None
        self.bond_to_token = None

    @classmethod
    def from_structure(cls, s):
        df = s["angles"]
        return cls({c: list(df[c]) for c in COLUMNS}, s.get("fname"))

    # -- accessors, tokenizer.py:131-167
    def _bond_length(self, idx, orig=False):
        if idx == 0:
            return self._init_n_ca
        if idx == 1:
            return self._init_ca_c
        return (self.orig if orig else self.cur)[BOND_TYPES[idx % 3]][(idx - 2) // 3]

    def _set_bond_length(self, idx, value):
        if idx == 0:
            self._init_n_ca = value
        elif idx == 1:
            self._init_ca_c = value
        else:
            self.cur[BOND_TYPES[idx % 3]][(idx - 2) // 3] = value

    def _bond_angle(self, idx, orig=False):
        if idx == 0:
            return self._init_bond_angle
        return (self.orig if orig else self.cur)[BOND_ANGLES[idx % 3]][(idx - 1) // 3]

    def _set_bond_angle(self, idx, value):
        if idx == 0:
            self._init_bond_angle = value
        else:
            self.cur[BOND_ANGLES[idx % 3]][(idx - 1) // 3] = value

    def _dihedral_angle(self, idx, orig=False):
        return (self.orig if orig else self.cur)[DIHEDRAL_ANGLES[idx % 3]][(idx + 1) // 3]

    def _set_dihedral_angle(self, idx, value):
        self.cur[DIHEDRAL_ANGLES[idx % 3]][(idx + 1) // 3] = value

    def token_geo(self, idx, l, orig=False):          # tokenizer.py:169-202
        if idx + l - 1 > 3 * self.n - 1:
            raise ValueError(f"idx+l cannot exceed {3*self.n-1}")
        ans = {}
        for j in range(idx, idx + l):
            ans.setdefault(BOND_TYPES[j % 3], []).append(self._bond_length(j, orig))
        for j in range(idx, idx + l - 1):
            ans.setdefault(BOND_ANGLES[j % 3], []).append(self._bond_angle(j, orig))
        for j in range(idx, idx + l - 2):
            ans.setdefault(DIHEDRAL_ANGLES[j % 3], []).append(self._dihedral_angle(j, orig))
        return ans

    def set_token_geo(self, idx, l, vals):            # tokenizer.py:253-286
        it = {k: iter(v) for k, v in vals.items()}
        for j in range(idx, idx + l):
            self._set_bond_length(j, next(it[BOND_TYPES[j % 3]]))
        for j in range(idx, idx + l - 1):
            self._set_bond_angle(j, next(it[BOND_ANGLES[j % 3]]))
        for j in range(idx, idx + l - 2):
            self._set_dihedral_angle(j, next(it[DIHEDRAL_ANGLES[j % 3]]))
        for k in it:
            assert next(it[k], None) is None

    def compute_coords(self, index=0, length=float("inf"), orig=False):   # tokenizer.py:347-363
        length = min(length, 3 * self.n - 1 - index)
        start = 3 * (index // 3)
        end = 3 * (((index + length - 1) + 1) // 3) + 1
        off_start = index - start
        off_end = end - (index + length - 1)
        coords = geo_nerf_coords(self.token_geo(start, end - start + 1, orig=orig))
        return coords[off_start: len(coords) - off_end]

    def tokenize(self):                               # tokenizer.py:379-392
        out = []
        for (start, bt, length) in self.bond_to_token.values():
            out.append(("MOTIF", bt))
            b = start + length
            if b < 3 * self.n - 1:
                out.append(("DIHEDRAL", DIHEDRAL_ANGLES[(b - 2) % 3], self._dihedral_angle(b - 2)))
                out.append(("DIHEDRAL", DIHEDRAL_ANGLES[(b - 1) % 3], self._dihedral_angle(b - 1)))
                out.append(("BOND_ANGLE", BOND_ANGLES[(b - 1) % 3], self._bond_angle(b - 1)))
        return out


def key_coords(key):
    """tokenizer.py:204-230: coordinates of a geometry dict placed in a blank backbone."""
    length = num_bonds(key)
    bt = sorted([(len(key.get(b, [])), b) for b in BOND_TYPES])
    ba = sorted([(len(key.get(b, [])), b) for b in BOND_ANGLES])
    da = sorted([(len(key.get(b, [])), b) for b in DIHEDRAL_ANGLES])
    if bt[2][0] > bt[1][0]:
        index = BOND_TYPES.index(bt[2][1])
    elif ba[2][0] > ba[1][0]:
        index = BOND_ANGLES.index(ba[2][1])
    elif da[2][0] > da[1][0]:
        index = DIHEDRAL_ANGLES.index(da[2][1])
    else:
        raise AssertionError("key_coords: cannot infer phase")
    start = 3 * (index // 3)
    end = 3 * (((index + length - 1) + 1) // 3) + 1
    off_start = index - start
    m = (end - start + 2) // 3
    blank = {c: ([0.0] * m if c in BOND_TYPES else [float("nan")] * m) for c in COLUMNS}  # tokenizer.py:394-417
    t = Tokenizer(blank)
    t.set_token_geo(off_start, length, key)
    coords = t.compute_coords()
    return coords[off_start: off_start + length + 1]


# ----------------------------------------------------------------------------- histogram
def save_histogram(vals, circular=True, bins=10, cover=False):        # plotting.py:305-337
    a = np.array(vals)
    if circular:
        a = (a + 2 * np.pi) % (2 * np.pi)
        counts, edges = np.histogram(a, bins=bins, range=(0, 2 * np.pi) if cover else None)
    else:
        counts, edges = np.histogram(a, bins=bins)
    return edges[:-1], edges[1:], np.diff(edges), counts


def save_histogram_equal_counts(vals, circular=True, bins=10):        # plotting.py:256-302
    a = np.asarray(vals)
    if circular:
        a = (a + 2 * np.pi) % (2 * np.pi)
    edges = np.quantile(np.sort(a), np.linspace(0, 1, bins + 1))
    counts, _ = np.histogram(a, bins=edges)
    return edges[:-1], edges[1:], np.diff(edges), counts


# ----------------------------------------------------------------------------- BPE
class BPE:
    """bpe.py:32-2166, res_init=True paths without glue optimisation."""

    def __init__(self, structures, bins, bin_strategy="histogram", save_dir=None,
                 res_init=True, std_bonds=True, rmsd_partition_min_size=4, rmsd_super_res=False,
                 rmsd_only=False, num_partitions=3, max_num_strucs=500, glue_opt=False, seed=None,
                 sorted_sets=False):
        if glue_opt:
            raise NotImplementedError("glue optimisation is not part of the oracle (SURVEY f-1)")
        if not res_init:
            raise NotImplementedError("oracle restates the res_init=True path only")
        self.tokenizers = [s if isinstance(s, Tokenizer) else Tokenizer.from_structure(s) for s in structures]
        self.rmsd_partition_min_size = rmsd_partition_min_size
        self.rmsd_super_res = rmsd_super_res
        self.rmsd_only = rmsd_only
        self.num_partitions = ThresholdDict(num_partitions) if isinstance(num_partitions, dict) else num_partitions
        self.max_num_strucs = max_num_strucs
        self.res_init = res_init
        self.std_bonds = std_bonds
        self.bins = bins
        self.bin_strategy = bin_strategy
        self.n = len(self.tokenizers)
        self.seed = seed
        self.rng = np.random.default_rng(seed)                        # bpe.py:84
        self._step = 0
        # test hook: the disclosed deviation of SURVEY 7.2(1)(b) -- enumerate occurrence sets sorted
        self.sorted_sets = sorted_sets
        self.pop_log = []

    def _enum(self, s):
        return sorted(s) if self.sorted_sets else list(s)

    def _num_p(self, length):
        return self.num_partitions[length] if isinstance(self.num_partitions, dict) else self.num_partitions

    # ---- bpe.py:91-103
    def initialize(self):
        self._init_thresholds()
        self._init_res_tokens()

    # ---- bpe.py:820-876
    def _init_thresholds(self):
        keys = BOND_ANGLES + DIHEDRAL_ANGLES
        if not self.std_bonds:
            keys = keys + BOND_TYPES
        self._thresholds = ThresholdDict()
        self._bin_counts = ThresholdDict()
        for size, num_bins in self.bins.items():
            vals = {k: [] for k in keys}
            for t in self.tokenizers:
                for key in keys:
                    col = t.cur[key]
                    tv = [v for v in col if v == v and v != 0.0]     # fillna(0) != 0, bpe.py:844
                    if key == "tau":
                        tv.append(t._bond_angle(0))
                    elif key == "N:CA":
                        tv.append(t._bond_length(0))
                    elif key == "CA:C":
                        tv.append(t._bond_length(1))
                    vals[key].extend(tv)
            th, bc = {}, {}
            for key in keys:
                circ = key not in BOND_TYPES
                if self.bin_strategy.startswith("histogram"):
                    starts, ends, _, counts = save_histogram(vals[key], circular=circ, bins=num_bins,
                                                             cover="cover" in self.bin_strategy)
                elif self.bin_strategy == "uniform":
                    starts, ends, _, counts = save_histogram_equal_counts(vals[key], circular=circ, bins=num_bins)
                else:
                    raise NotImplementedError
                th[key] = [(float(s), float(e)) for s, e in zip(starts, ends)]
                bc[key] = [c for c in counts]
            self._thresholds[size] = th
            self._bin_counts[size] = bc
        if self.std_bonds:
            for i in range(3):
                self._thresholds[BOND_TYPES[i]] = [(BOND_LENGTHS[i], BOND_LENGTHS[i])]

    # ---- bpe.py:715-737
    def _set_bond_lengths(self, t, strict=True):
        lookup = self._thresholds if self.std_bonds else self._thresholds[1]
        relv = {bt: lookup[bt] for bt in lookup if bt in BOND_TYPES}
        for i in range(3 * t.n - 1):
            key = BOND_TYPES[i % 3]
            v = t._bond_length(i)
            if self.std_bonds:
                ind = 0
            elif not strict and v < relv[key][0][0]:
                ind = 0
            elif not strict and v > relv[key][-1][1]:
                ind = -1
            else:
                ind = get_ind(v, relv[key])
            t._set_bond_length(i, sum(relv[key][ind]) / 2)

    # ---- bpe.py:1500-1526
    def _bin_val(self, key_dict):
        size = num_bonds(key_dict)
        out = {}
        for k, vals in key_dict.items():
            relv = self._thresholds if (k in BOND_TYPES and self.std_bonds) else self._thresholds[size]
            out[k] = [sum(relv[k][v]) / 2 if isinstance(v, int) else v for v in vals]
        return out

    def quant_geo(self, geo):
        length = num_bonds(geo)
        for k in geo:
            q = []
            for v in geo[k]:
                if k in BOND_TYPES:
                    lookup = self._thresholds
                else:
                    lookup = self._thresholds[length]
                    v = (v + 2 * np.pi) % (2 * np.pi)
                q.append(get_ind(v, lookup[k]))
            geo[k] = q

    def _snap_glue(self, t):                          # bpe.py:381-391 / 1101-1108
        for angle in ["omega", "C:1N:1CA", "phi"]:
            relv = self._thresholds[1][angle]
            t.cur[angle] = [sum(relv[get_ind((v + 2 * np.pi) % (2 * np.pi), relv)]) / 2 if v == v else v
                            for v in t.cur[angle]]

    def _assign(self, coords, medoid_coords):         # bpe.py:654-657
        return int(np.argmin([compute_rmsd(coords, m) for m in medoid_coords]))

    # ---- bpe.py:138-394
    def _init_res_tokens(self):
        self._tokens = {}
        label_dict = {}
        res_geo = defaultdict(list)
        super_res = self.rmsd_super_res
        for t in self.tokenizers:
            self._set_bond_lengths(t)
        for ti, t in enumerate(self.tokenizers):
            labels = []
            for i in range(t.n):
                start = 3 * i
                length = 3 if i < t.n - 1 else 2
                if length < self.rmsd_partition_min_size:
                    geo = t.token_geo(start, length)
                    self.quant_geo(geo)
                    key = self._bin_val(geo)
                    key_str = hash_geo(key)
                    if key_str not in label_dict:
                        label_dict[key_str] = len(label_dict)
                    t.set_token_geo(start, length, key)
                    labels.append(label_dict[key_str])
                else:
                    res_geo[length].append((ti, start, length))
                    labels.append(None)
            toks = [(3 * i, labels[i], 3) for i in range(t.n - 1)] + [(3 * t.n - 3, labels[t.n - 1], 2)]
            t.token_pos = [3 * (i // 3) for i in range(3 * t.n - 1)]
            t.tokens = toks
            t.bond_to_token = {tk[0]: tk for tk in toks}
        if res_geo:
            self._sphere_dict = {}
            self._tokens = {}
            for n, size in enumerate(res_geo):
                N = len(res_geo[size])
                if N > self.max_num_strucs:
                    active_inds = self.rng.choice(N, self.max_num_strucs, replace=False)
                else:
                    active_inds = np.arange(N)
                active_coords, active_occurs = [], []
                for i in active_inds:
                    ti, start, length = res_geo[size][i]
                    active_coords.append(self.tokenizers[ti].compute_coords(start, length, super_res))
                    active_occurs.append((ti, start, length))
                medoid_inds = k_medoids(active_coords, self._num_p(size), rng=self.rng)
                med_coords = [active_coords[m] for m in medoid_inds]
                assignments = [self._assign(self.tokenizers[ti].compute_coords(start, length, orig=super_res), med_coords)
                               for (ti, start, length) in res_geo[size]]
                if size == 3:
                    k = '{"N:CA": [0], "CA:C": [0], "0C:1N": [0], "tau": [0], "CA:C:1N": [0], "psi": [0]}'
                elif size == 2:
                    k = '{"CA:C": [0], "0C:1N": [0], "CA:C:1N": [0]}'
                else:
                    raise NotImplementedError
                self._sphere_dict[k] = []
                for p, m in enumerate(medoid_inds):
                    ti, i1, length = active_occurs[m]
                    struc = self.tokenizers[ti].token_geo(i1, length, orig=super_res)
                    self._sphere_dict[k].append(struc)
                    self._tokens[(n, p)] = struc
                for (ti1, start1, length1), p in zip(res_geo[size], assignments):
                    t1 = self.tokenizers[ti1]
                    t1.set_token_geo(start1, length1, self._tokens[(n, p)])
                    t1.tokens[start1 // 3] = (start1, (n, p), length1)
                    t1.bond_to_token[start1] = t1.tokens[start1 // 3]
        for t in self.tokenizers:
            self._snap_glue(t)
        if not res_geo:
            self._tokens = {n: json.loads(k) for k, n in label_dict.items()}

    # ---- bpe.py:1192-1299
    def compute_geo_key(self, token_pair, t, ignore_left=False, ignore_right=False):
        if isinstance(t, int):
            t = self.tokenizers[t]
        (idx1, _, l1), (idx2, _, l2) = token_pair
        assert idx1 + l1 == idx2
        tok1 = t.bond_to_token[t.token_pos[idx1]]
        tok2 = t.bond_to_token[t.token_pos[idx2]]
        pt1 = not ignore_left and isinstance(tok1[1], tuple)
        pt2 = not ignore_right and isinstance(tok2[1], tuple)
        geo = t.token_geo(idx1, l1 + l2)
        for k in geo:
            q = []
            if k in BOND_TYPES:
                off, kind = ((BOND_TYPES.index(k) + 3) - idx1 % 3) % 3, 0
            elif k in BOND_ANGLES:
                off, kind = ((BOND_ANGLES.index(k) + 3) - idx1 % 3) % 3, 1
            else:
                off, kind = ((DIHEDRAL_ANGLES.index(k) + 3) - idx1 % 3) % 3, 2
            for i, v in enumerate(geo[k]):
                i = off + 3 * i
                if pt1 and pt2:
                    quant = (kind == 1 and i == l1 - 1) or (kind == 2 and (i == l1 - 2 or i == l1 - 1))
                elif pt1:
                    quant = i >= (l1, l1 - 1, l1 - 2)[kind]
                elif pt2:
                    quant = i < l1
                else:
                    quant = True
                if quant:
                    if self.std_bonds and kind == 0:
                        lookup = self._thresholds
                    else:
                        if kind != 0:
                            v = (v + 2 * np.pi) % (2 * np.pi)
                        lookup = self._thresholds[l1 + l2]
                    q.append(get_ind(v, lookup[k]))
                else:
                    q.append(v)
            geo[k] = q
        return hash_geo(geo)

    # ---- bpe.py:1431-1474
    def bin(self):
        self._geo_dict = defaultdict(set)
        self._priority_dict = SortedDict()
        self._key_to_priority = {}
        for i in range(self.n):
            t = self.tokenizers[i]
            p1, p2 = 0, 1
            for _ in range(len(t.bond_to_token) - 1):
                while t.token_pos[p2] == p1:
                    p2 += 1
                pair = (t.bond_to_token[p1], t.bond_to_token[p2])
                self._geo_dict[self.compute_geo_key(pair, i)].add((i, pair[1][0]))
                p1 = p2
            assert p1 == t.token_pos[-1]
        for key in self._geo_dict:
            pr = (True, -len(self._geo_dict[key]), key)
            self._priority_dict[pr] = None
            self._key_to_priority[key] = pr

    # ---- bpe.py:1739-1789
    def rmsd_partition(self, k):
        super_res = self.rmsd_super_res
        length = num_bonds(json.loads(k))
        all_pos = self._enum(self._geo_dict[k])
        N = len(all_pos)
        if N > self.max_num_strucs:
            active_inds = self.rng.choice(N, self.max_num_strucs, replace=False)
        else:
            active_inds = np.arange(N)
        active_occurs, active_coords = [], []
        for i in active_inds:
            ti, index = all_pos[i]
            t = self.tokenizers[ti]
            i1 = t.token_pos[index - 1]
            active_coords.append(t.compute_coords(i1, length, orig=super_res))
            active_occurs.append((ti, i1))
        medoid_inds = k_medoids(active_coords, self._num_p(length), rng=self.rng)
        med_coords = [active_coords[m] for m in medoid_inds]
        assignments = []
        for (ti, index) in all_pos:
            t = self.tokenizers[ti]
            assignments.append(self._assign(t.compute_coords(t.token_pos[index - 1], length, orig=super_res), med_coords))
        self._sphere_dict[k] = []
        for m in medoid_inds:
            ti, i1 = active_occurs[m]
            self._sphere_dict[k].append(self.tokenizers[ti].token_geo(i1, length, orig=super_res))
        return all_pos, assignments

    # ---- bpe.py:1792-2166
    def step(self):
        (exists, negc, key), _ = self._priority_dict.peekitem(0)
        recurring_key = not exists
        count = -negc
        assert count == len(self._geo_dict[key])
        self.pop_log.append((count, key))
        key_dict = json.loads(key)
        length = num_bonds(key_dict)
        if length >= self.rmsd_partition_min_size:
            if recurring_key:
                med_coords = [key_coords(kk) for kk in self._sphere_dict[key]]
                vals = self._enum(self._geo_dict[key])
                assignments = []
                for (ti, index) in vals:
                    t = self.tokenizers[ti]
                    assignments.append(self._assign(
                        t.compute_coords(t.token_pos[index - 1], length, orig=self.rmsd_super_res), med_coords))
            else:
                vals, assignments = self.rmsd_partition(key)
            rmsd_key = key
        else:
            rmsd_key = None
            vals = self._enum(self._geo_dict[key])
        n = len(self._tokens)
        if rmsd_key is None:
            binned_key_dict = self._bin_val(key_dict)
            self._tokens[n] = key_dict
        elif recurring_key:
            n_ind = list(self._sphere_dict).index(key)
            n = sorted(set(x[0] for x in self._tokens))[n_ind]
        else:
            for p, token_p in enumerate(self._sphere_dict[key]):
                self._tokens[(n, p)] = token_p
        diff_count = defaultdict(list)
        # NOTE bpe.py:1888 re-lists the set; list(s) twice on an unmodified set gives the same order
        order = sorted(range(len(vals)), key=lambda i: vals[i])
        last_i, last_i1 = None, None
        for idx in order:
            (i, index) = vals[idx]
            t = self.tokenizers[i]
            i2 = index
            i1 = t.token_pos[index - 1]
            l1 = i2 - i1
            l2 = length - l1
            overlaps = (last_i is not None) and (last_i == i) and (last_i1 + length > i1)
            assert overlaps == ((i, index) not in self._geo_dict[key])         # bpe.py:1910-1912
            if overlaps:
                continue
            assert l1 > 0 and l2 > 0
            assert self.compute_geo_key(((i1, None, l1), (i2, None, l2)), i) == key   # bpe.py:1917-1920
            self._geo_dict[key].remove((i, index))
            diff_count[key].append("remove")
            if i1:
                i0 = t.token_pos[i1 - 1]
                l0 = i1 - i0
                left_key = self.compute_geo_key(((i0, None, l0), (i1, None, l1)), i)
            else:
                left_key = None
            if i2 + l2 < len(t.token_pos):
                i3 = i2 + l2
                l3 = 0
                while i3 + l3 < len(t.token_pos) and t.token_pos[i3 + l3] == i3:
                    l3 += 1
                right_key = self.compute_geo_key(((i2, None, l2), (i3, None, l3)), i)
            else:
                right_key = None
            if left_key:
                self._geo_dict[left_key].remove((i, i1))
                diff_count[left_key].append("remove")
            if right_key:
                self._geo_dict[right_key].remove((i, i3))
                diff_count[right_key].append("remove")
            for j in range(i2, i2 + l2):
                t.token_pos[j] = i1
            t.bond_to_token.pop(i2)
            if rmsd_key is not None:
                p = assignments[idx]
                t.bond_to_token[i1] = (i1, (n, p), length)
                if not self.rmsd_only:
                    t.set_token_geo(i1, length, self._sphere_dict[key][p])
            else:
                t.bond_to_token[i1] = (i1, n, length)
            if left_key:
                nk = self.compute_geo_key(((i0, None, l0), (i1, None, l1 + l2)), i)
                self._geo_dict[nk].add((i, i1))
                diff_count[nk].append("add")
            if right_key:
                nk = self.compute_geo_key(((i1, None, l1 + l2), (i3, None, l3)), i)
                self._geo_dict[nk].add((i, i3))
                diff_count[nk].append("add")
            if rmsd_key is None:
                t.set_token_geo(i1, l1 + l2, binned_key_dict)
            last_i, last_i1 = i, i1
        if not recurring_key:
            self._step += 1
        for k in diff_count:                           # Step 7, bpe.py:2077-2138
            if k in self._key_to_priority:
                pr = self._key_to_priority.pop(k)
                self._priority_dict.pop(pr)
                count = -pr[1]
            else:
                count = 0
            for action in diff_count[k]:
                count += 1 if action == "add" else -1
            assert count == len(self._geo_dict[k])
            if count:
                exists = hasattr(self, "_sphere_dict") and k in self._sphere_dict
                pr = (not exists, -count, k)
                self._key_to_priority[k] = pr
                self._priority_dict[pr] = None
            else:
                self._geo_dict.pop(k)
        if len(self._priority_dict) and not self._priority_dict.peekitem(0)[0][0]:
            self.step()                                # bpe.py:2164-2166

    # ---- bpe.py:905-956
    def cum_bin_count(self, key=None):
        count = 0
        for k, v in self._bin_counts[1].items():
            if key == k:
                break
            if self.res_init and k not in ["omega", "phi", "C:1N:1CA"]:
                continue
            count += len(v)
        return count

    @property
    def vocab_size(self):
        return len(self._tokens) + self.cum_bin_count()

    def quantize(self, tokenized):
        if isinstance(tokenized, Tokenizer):
            return self.quantize(tokenized.tokenize())
        ids = list(self._tokens)
        out = []
        for token in tokenized:
            if token[0] == "MOTIF":
                out.append(ids.index(token[1]))
            else:
                relv = self._thresholds[1][token[1]]
                out.append(len(self._tokens) + self.cum_bin_count(token[1]) +
                           get_ind((token[2] + 2 * np.pi) % (2 * np.pi), relv))
        return out

    # ---- bpe.py:1301-1425
    def bin_helper(self, t):
        geo_dict = defaultdict(set)
        p1, p2 = 0, 1
        for _ in range(len(t.bond_to_token) - 1):
            while t.token_pos[p2] == p1:
                p2 += 1
            pair = (t.bond_to_token[p1], t.bond_to_token[p2])
            geo_dict[self.compute_geo_key(pair, t)].add(pair[1][0])
            p1 = p2
        return geo_dict

    def step_helper(self, geo_dict, t, key, n):
        super_res = self.rmsd_super_res
        length = num_bonds(json.loads(key))
        vals = self._enum(geo_dict[key])
        med_coords = [key_coords(g) for g in self._sphere_dict[key]]
        assignments = [self._assign(t.compute_coords(t.token_pos[index - 1], length, orig=super_res), med_coords)
                       for index in vals]
        order = sorted(range(len(vals)), key=lambda i: vals[i])
        last_i1 = None
        for idx in order:
            index = vals[idx]
            p = assignments[idx]
            i2 = index
            i1 = t.token_pos[index - 1]
            l1 = i2 - i1
            l2 = length - l1
            overlaps = last_i1 is not None and (last_i1 + length > i1)
            assert overlaps == (index not in geo_dict[key])
            if overlaps:
                continue
            assert self.compute_geo_key(((i1, None, l1), (i2, None, l2)), t) == key
            geo_dict[key].remove(index)
            if i1:
                i0 = t.token_pos[i1 - 1]
                l0 = i1 - i0
                left_key = self.compute_geo_key(((i0, None, l0), (i1, None, l1)), t)
            else:
                left_key = None
            if i2 + l2 < len(t.token_pos):
                i3 = i2 + l2
                l3 = 0
                while i3 + l3 < len(t.token_pos) and t.token_pos[i3 + l3] == i3:
                    l3 += 1
                right_key = self.compute_geo_key(((i2, None, l2), (i3, None, l3)), t)
            else:
                right_key = None
            if left_key:
                geo_dict[left_key].remove(i1)
            if right_key:
                geo_dict[right_key].remove(i3)
            for j in range(i2, i2 + l2):
                t.token_pos[j] = i1
            t.bond_to_token.pop(i2)
            t.bond_to_token[i1] = (i1, (n, p), length)
            if not self.rmsd_only:
                t.set_token_geo(i1, length, self._sphere_dict[key][p])
            if left_key:
                geo_dict[self.compute_geo_key(((i0, None, l0), (i1, None, l1 + l2)), t)].add(i1)
            if right_key:
                geo_dict[self.compute_geo_key(((i1, None, l1 + l2), (i3, None, l3)), t)].add(i3)
            last_i1 = i1
        return t

    # ---- bpe.py:1053-1140  (induce: encode one held-out backbone against the frozen vocabulary)
    def tokenize(self, t):
        self._set_bond_lengths(t, strict=False)
        res_geo = defaultdict(list)
        for i in range(t.n):
            res_geo[3 if i < t.n - 1 else 2].append(3 * i)
        t.token_pos = [3 * (i // 3) for i in range(3 * t.n - 1)]
        t.tokens = [(3 * i, None, 3) for i in range(t.n - 1)] + [(3 * t.n - 3, None, 2)]
        t.bond_to_token = {tk[0]: tk for tk in t.tokens}
        for n, size in enumerate(res_geo):
            geo = t.token_geo(0, 5) if size == 3 else t.token_geo(0, 2)
            kc = []
            p = 0
            while (n, p) in self._tokens:
                key = self._tokens[(n, p)]
                for k in key:
                    geo[k][:len(key[k])] = key[k]
                p += 1
                kc.append(geo_nerf_coords(geo)[:size + 1])
            for start in res_geo[size]:
                coords = t.compute_coords(start, size)
                p = int(np.argmin([compute_rmsd(coords, c) for c in kc]))
                t.set_token_geo(start, size, self._tokens[(n, p)])
                t.tokens[start // 3] = (start, (n, p), size)
                t.bond_to_token[start] = t.tokens[start // 3]
        self._snap_glue(t)
        geo_dict = self.bin_helper(t)
        uniq = sorted(set(k[0] for k in self._tokens))
        geo_keys = list(self._sphere_dict)
        assert len(uniq) == len(geo_keys)
        Ls = [len(t.bond_to_token)]
        for (n, key) in list(zip(uniq, geo_keys))[2:]:
            if key in geo_dict:
                t = self.step_helper(geo_dict, t, key, n)
            Ls.append(len(t.bond_to_token))
        return t, {"L": Ls}


# ----------------------------------------------------------------------------- dump (shared by golden maker and tests)
def lddt_ca(pred_ca, true_ca, cutoff=15.0, eps=1e-10):
    """mean per-residue lDDT-C-alpha, as bin/encode.py:351-358 / plotting.py:436 take it from esm's ProteinChain.lddt_ca
    (esm is absent here: esm.utils.structure.metrics.compute_lddt -- the OpenFold lDDT -- restated; parity unpinned for this metric).
    float32 throughout, like the atom37 arrays it is called on."""
    P = np.asarray(pred_ca, dtype=np.float32)
    Q = np.asarray(true_ca, dtype=np.float32)
    n = len(P)
    f = np.float32
    dt = np.sqrt(f(eps) + ((Q[:, None, :] - Q[None, :, :]) ** 2).sum(-1, dtype=np.float32))
    dp = np.sqrt(f(eps) + ((P[:, None, :] - P[None, :, :]) ** 2).sum(-1, dtype=np.float32))
    to_score = (dt < f(cutoff)).astype(np.float32) * (f(1.0) - np.eye(n, dtype=np.float32))
    l1 = np.abs(dt - dp)
    score = ((l1 < 0.5).astype(np.float32) + (l1 < 1.0) + (l1 < 2.0) + (l1 < 4.0)).astype(np.float32) * f(0.25)
    norm = f(1.0) / (f(eps) + to_score.sum(-1, dtype=np.float32))
    per_res = norm * (f(eps) + (to_score * score).sum(-1, dtype=np.float32))
    return float(per_res.mean())


def _jsonable(x):
    if isinstance(x, dict):
        return {str(k): _jsonable(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [_jsonable(v) for v in x]
    if isinstance(x, (np.integer,)):
        return int(x)
    if isinstance(x, (np.floating,)):
        return float(x)
    return x


def dump_state(bpe, with_angles=True, max_angles=4):
    """Reference-API-level state -> JSON-able dict.  Works on the reference BPE, this oracle and
    the CUDA-backed drop-in (all expose _tokens / tokenizers[i].bond_to_token / _thresholds)."""
    out = {
        "thresholds": {str(k): _jsonable(v) for k, v in bpe._thresholds.items()},
        "bin_counts": {str(k): _jsonable(v) for k, v in bpe._bin_counts.items()},
        "tokens": [[_jsonable(k), _jsonable(v)] for k, v in bpe._tokens.items()],
        "segmentation": [[_jsonable(v) for v in t.bond_to_token.values()] for t in bpe.tokenizers],
        "step": bpe._step,
    }
    if hasattr(bpe, "_sphere_dict"):
        out["sphere_keys"] = list(bpe._sphere_dict)
    out["init_consts"] = [[float(t._init_n_ca), float(t._init_ca_c), float(t._init_bond_angle)]
                          for t in bpe.tokenizers[:max_angles]]
    if with_angles:
        ang = []
        for t in bpe.tokenizers[:max_angles]:
            if hasattr(t, "cur"):
                ang.append({c: [None if v != v else float(v) for v in t.cur[c]] for c in COLUMNS})
            else:
                df = t.angles_and_dists
                ang.append({c: [None if v != v else float(v) for v in df[c]] for c in COLUMNS})
        out["angles"] = ang
    return out
