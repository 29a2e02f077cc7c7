"""Host-side mirror of the reference's `foldingdiff.bpe.BPE` / `foldingdiff.tokenizer.Tokenizer` API for the
learn / induce hot path, backed by the sm_100a kernels (engine.py -> C ABI).  Same method names, argument
meaning and error behaviour as bpe.py:33-2166; plotting methods are no-ops (no graphics stack)."""
import json
import time

import numpy as np

from . import engine as E
from .data_structures import ThresholdDict, TokenHierarchy
from .synthetic import from_structures

INF = float("inf")


def _window_coords(vals, init, n, index, length):
    """Tokenizer.compute_coords (tokenizer.py:347-363) on the device: the window is rounded out to whole residues, rebuilt
    by K4 (gb_window_coords: first residue placed from N:CA / CA:C / tau, then NeRF), and sliced back to length + 1 atoms.
    vals: float64 [9, n] in COLUMNS order; init: (N:CA, CA:C, tau) of residue 0 (Tokenizer._init_*)."""
    import torch
    length = min(length, 3 * n - 1 - index)
    start = 3 * (index // 3)
    end = 3 * (((index + length - 1) + 1) // 3) + 1
    off_start, off_end = index - start, end - (index + length - 1)
    r0, nres = start // 3, (end - start + 2) // 3
    mini = E.Backbones(np.ascontiguousarray(vals, dtype=np.float64), np.array([0, n], dtype=np.int64))
    out = E._empty(3 * nres * 3, torch.float64)
    r0_dev = torch.full((1,), r0, dtype=torch.int32, device=out.device)
    E._lib.call("gb_window_coords", E._lib.ptr(mini.vals), n, n, E._lib.ptr(mini.bb), E._lib.ptr(mini.offsets),
                E._lib.ptr(r0_dev), 1, nres, 3 * nres, None, None, float(init[0]), float(init[1]), float(init[2]), None, None, 0,
                E._lib.ptr(out))
    coords = out.view(3 * nres, 3).cpu().numpy()
    return coords[off_start:len(coords) - off_end]


class Tokenizer:
    """Per-backbone view (tokenizer.py:18-477).  Attached to a BPE, its state lives in HBM and Python objects are
    materialised on access; built from a structure dict alone (Tokenizer.init_structure, BPE.recover_structure,
    key_coords) it is a plain host object with the same accessors."""
    BOND_TYPES = E.BOND_TYPES
    ATOM_TYPES = ["N", "CA", "C"]
    BOND_ANGLES = E.BOND_ANGLES
    DIHEDRAL_ANGLES = E.DIHEDRAL_ANGLES
    BOND_LENGTHS = E.BOND_LENGTHS

    def __init__(self, structure=None, bpe=None, index=None, compute_sec_structs=False):
        self._bpe, self._i = bpe, index
        self._structure = structure
        if structure is not None:
            df = structure["angles"]
            self.n = len(df)
            self.fname = structure.get("fname")
            self.aa = structure.get("aa")
            self.beta_coords = structure.get("c_beta")
        self._cache_step = None

    # fixed per-backbone labels (tokenizer.py:48-51)
    @property
    def bond_labels(self):
        return sum([[0, 1, 2] for _ in range(self.n - 1)] + [[0, 1]], [])

    @property
    def atom_labels(self):
        return np.tile([0, 1, 2], self.n)

    @property
    def edges(self):
        return [[j, j + 1, 0] for j in range(1, 3 * self.n)]

    # -- pickling: materialise, drop the device-backed owner (ref_tokenizers={t}.pkl, {idx}.pkl, bpe_iter={t}.pkl) ---
    def __getstate__(self):
        d = dict(self.__dict__)
        if d.get("_bpe") is not None and getattr(d["_bpe"], "_learner", None) is not None:
            d["_mat"] = dict(self._state(), orig=self._orig())      # (the reference pickles _angles_and_dists_orig too)
        d["_bpe"] = None
        d["_cache_step"] = None
        return d

    # -- materialisation ------------------------------------------------------------------------------
    def _state(self):
        b = self._bpe
        if getattr(self, "_local", False):
            return self._mat                           # edited through set_token_geo / bond_to_token: a host object now
        if b is None and getattr(self, "_mat", None) is not None:
            return self._mat                           # unpickled: frozen snapshot
        if b is None or b._learner is None:
            if self._structure is None:
                raise RuntimeError("Tokenizer is not attached to an initialised BPE")
            df = self._structure["angles"]             # stand-alone: the structure's own DataFrame (tokenizer.py:24-31)
            cur = {c: [float(x) for x in df[c]] for c in E.COLUMNS}
            st = {"bond_to_token": None, "cur": cur, "orig": {c: list(v) for c, v in cur.items()},
                  "init": [E.INIT_N_CA, E.INIT_CA_C, E.INIT_TAU]}
            if b is not None:                           # owned by a BPE that is not initialised yet (encode.py:308 reads the
                return st                               # reference coordinates first): nothing to keep
            self._mat = st
            self._local = True
            return self._mat
        key = (id(b._learner), b._learner.n_steps, len(b._learner.pop_log))
        if self._cache_step != key:
            self._mat = b._materialise(self._i)
            self._cache_step = key
        return self._mat

    def _detach(self):
        """first write through the reference's setters: keep a private host copy from here on"""
        st = self._state()
        if not getattr(self, "_local", False):
            self._mat = {"bond_to_token": st["bond_to_token"], "cur": {c: list(v) for c, v in st["cur"].items()},
                         "orig": self._orig(), "init": list(st["init"])}
            self._local = True
        return self._mat

    def _orig(self):
        st = self._state()
        if "orig" in st:
            return st["orig"]
        b = self._bpe
        offs = b._learner.bb.offsets_host if getattr(b, "_learner", None) is not None else b._offsets
        a, e = int(offs[self._i]), int(offs[self._i + 1])
        return {c: [float(x) for x in b._vals[E.COL[c], a:e]] for c in E.COLUMNS}

    @property
    def bond_to_token(self):
        return self._state()["bond_to_token"]

    @bond_to_token.setter
    def bond_to_token(self, new_value):                # tokenizer.py:69-71: a fresh hierarchy over the given leaves
        self._detach()["bond_to_token"] = TokenHierarchy(self, new_value)

    @property
    def token_pos(self):
        tp = []
        for (s, _, l) in self.bond_to_token.values():
            tp += [s] * l
        return tp

    @property
    def tokens(self):
        return list(self.bond_to_token.values())

    @property
    def _init_n_ca(self):
        return self._state()["init"][0]

    @property
    def _init_ca_c(self):
        return self._state()["init"][1]

    @property
    def _init_bond_angle(self):
        return self._state()["init"][2]

    @property
    def cur(self):
        return self._state()["cur"]

    @property
    def angles_and_dists(self):
        import pandas as pd
        return pd.DataFrame({c: pd.Series(self.cur[c], dtype=object) for c in E.COLUMNS})

    _angles_and_dists = angles_and_dists

    @property
    def angles_and_dists_orig(self):
        import pandas as pd
        o = self._orig()
        return pd.DataFrame({c: pd.Series(o[c], dtype=object) for c in E.COLUMNS})

    _angles_and_dists_orig = angles_and_dists_orig

    # -- scalar accessors (tokenizer.py:130-167) ---------------------------------------------------------------------
    def _src(self, orig):
        return self._orig() if orig else self.cur

    def _bond_length(self, idx, orig=False):
        if idx == 0:
            return self._init_n_ca
        if idx == 1:
            return self._init_ca_c
        return self._src(orig)[E.BOND_TYPES[idx % 3]][(idx - 2) // 3]

    def _bond_angle(self, idx, orig=False):
        if idx == 0:
            return self._init_bond_angle
        return self._src(orig)[E.BOND_ANGLES[idx % 3]][(idx - 1) // 3]

    def _dihedral_angle(self, idx, orig=False):
        return self._src(orig)[E.DIHEDRAL_ANGLES[idx % 3]][(idx + 1) // 3]

    def _set_bond_length(self, idx, value):
        st = self._detach()
        if idx < 2:
            st["init"][idx] = value
        else:
            st["cur"][E.BOND_TYPES[idx % 3]][(idx - 2) // 3] = value

    def _set_bond_angle(self, idx, value):
        st = self._detach()
        if idx == 0:
            st["init"][2] = value
        else:
            st["cur"][E.BOND_ANGLES[idx % 3]][(idx - 1) // 3] = value

    def _set_dihedral_angle(self, idx, value):
        self._detach()["cur"][E.DIHEDRAL_ANGLES[idx % 3]][(idx + 1) // 3] = value

    # -- window geometry (tokenizer.py:169-202, 253-286) ---------------------------------------------------------------
    def token_geo(self, idx, l, orig=False):
        """geometry of bonds idx .. idx+l-1: l bond lengths, l-1 bond angles, l-2 dihedrals, keyed by type"""
        if idx + l - 1 > 3 * self.n - 1:
            raise ValueError(f"idx+l cannot exceed {3 * self.n - 1}")
        ans = {}
        for j in range(idx, idx + l):
            ans.setdefault(E.BOND_TYPES[j % 3], []).append(self._bond_length(j, orig=orig))
        for j in range(idx, idx + l - 1):
            ans.setdefault(E.BOND_ANGLES[j % 3], []).append(self._bond_angle(j, orig=orig))
        for j in range(idx, idx + l - 2):
            ans.setdefault(E.DIHEDRAL_ANGLES[j % 3], []).append(self._dihedral_angle(j, orig=orig))
        return ans

    def set_token_geo(self, idx, l, vals):
        """inverse of token_geo: every list of `vals` must be consumed exactly"""
        it = {k: iter(v) for k, v in vals.items()}
        used = {k: 0 for k in vals}

        def take(k):
            used[k] += 1
            return next(it[k])
        for j in range(idx, idx + l):
            self._set_bond_length(j, take(E.BOND_TYPES[j % 3]))
        for j in range(idx, idx + l - 1):
            self._set_bond_angle(j, take(E.BOND_ANGLES[j % 3]))
        for j in range(idx, idx + l - 2):
            self._set_dihedral_angle(j, take(E.DIHEDRAL_ANGLES[j % 3]))
        assert all(used[k] == len(vals[k]) for k in vals)

    @staticmethod
    def num_bonds(geo):                                # tokenizer.py:303-305
        return len(geo.get("N:CA", [])) + len(geo.get("CA:C", [])) + len(geo.get("0C:1N", []))

    @staticmethod
    def init_structure(n):                             # tokenizer.py:394-417: a blank n-residue structure
        import pandas as pd
        cols = {c: [0.0 if c in E.BOND_TYPES else np.nan for _ in range(n)] for c in E.COLUMNS}
        return {"angles": pd.DataFrame(cols), "coords": None, "c_beta": None,
                "full_idxes": sum([[i, i, i] for i in range(1, n + 1)], []), "full_coords": None, "side_chain": None,
                "aa": None, "fname": None}

    @staticmethod
    def key_coords(key):
        """tokenizer.py:204-230: coordinates of a vocabulary geometry placed in a blank backbone (length + 1 atoms)"""
        length = Tokenizer.num_bonds(key)
        cnt = lambda names: sorted((len(key.get(x, [])), x) for x in names)
        bt, ba, da = cnt(E.BOND_TYPES), cnt(E.BOND_ANGLES), cnt(E.DIHEDRAL_ANGLES)
        if bt[2][0] > bt[1][0]:
            index = E.BOND_TYPES.index(bt[2][1])
        elif ba[2][0] > ba[1][0]:
            index = E.BOND_ANGLES.index(ba[2][1])
        elif da[2][0] > da[1][0]:
            index = E.DIHEDRAL_ANGLES.index(da[2][1])
        else:
            raise AssertionError("key does not determine its starting bond type (tokenizer.py:218-219)")
        start = 3 * (index // 3)
        end = 3 * (((index + length - 1) + 1) // 3) + 1
        off_start = index - start
        t = Tokenizer(Tokenizer.init_structure((end - start + 2) // 3))
        t.set_token_geo(off_start, length, key)
        return t.compute_coords()[off_start:off_start + length + 1]

    def tokenize(self):                               # tokenizer.py:379-392
        out = []
        for (start, bt, length) in self.bond_to_token.values():
            out.append(("MOTIF", bt))
            b = start + length
            if b < 3 * self.n - 1:
                out.append(("DIHEDRAL", E.DIHEDRAL_ANGLES[(b - 2) % 3], self._dihedral_angle(b - 2)))
                out.append(("DIHEDRAL", E.DIHEDRAL_ANGLES[(b - 1) % 3], self._dihedral_angle(b - 1)))
                out.append(("BOND_ANGLE", E.BOND_ANGLES[(b - 1) % 3], self._bond_angle(b - 1)))
        return out

    def compute_coords(self, index=0, length=INF, orig=False):   # tokenizer.py:347-363
        st = self._state()
        src = self._orig() if orig else st["cur"]
        vals = np.array([src[c] for c in E.COLUMNS], dtype=np.float64)
        return _window_coords(vals, st["init"], self.n, index, length)

    def visualize(self, *a, **k):
        return None

    visualize_bonds = visualize


class _View:
    """materialises reference-style Python objects from a device-resident engine state (Learner or Inducer)"""
    _learner = None
    _export = None
    _ev_cache = None

    # ---- export ---------------------------------------------------------------------------------------
    def _token_arrays(self):
        L = self._learner
        key = (L.n_steps, len(L.pop_log))
        if self._export is None or self._export[0] != key:
            ts = L.ts
            T = ts.T
            pos = E._u32(ts.a["pos"][:T]); ln = E._u32(ts.a["len"][:T]); tid = E._u32(ts.a["tid"][:T])
            glue = E._u32(ts.a["glue"][:T])
            self._export = (key, pos, ln, tid, glue)
        return self._export[1:]

    def _materialise(self, i):
        L = self._learner
        pos, ln, tid, glue = self._token_arrays()
        offs = L.bb.offsets_host
        a, b = int(offs[i]), int(offs[i + 1])
        n = b - a
        lo, hi = np.searchsorted(pos, a), np.searchsorted(pos, b)
        seg = {}
        size0 = L.size0
        th = L.th.thresholds[size0]
        cur = {c: [float(x) for x in self._vals[E.COL[c], a:b]] for c in E.COLUMNS}
        centre = lambda k, bn: (th[k][bn][0] + th[k][bn][1]) / 2
        init = [None, None, None]
        rb = L.window_bins(a, n)

        def put(k, r, v):          # residue-major (k, local residue r) -> DataFrame cell (see engine.SLOTS)
            if k in ("N:CA", "CA:C", "tau"):
                if r == 0:
                    init[("N:CA", "CA:C", "tau").index(k)] = v
                else:
                    cur[k][r - 1] = v
            elif k == "phi":
                cur[k][r + 1] = v
            else:
                cur[k][r] = v

        for j in range(lo, hi):
            r0 = int(pos[j]) - a
            nres = int(ln[j])
            ends = bool(glue[j] & E.LAST)
            nb = 3 * nres - 1 if ends else 3 * nres
            if self._mode == "H":
                tok_id = int(tid[j])
                for r in range(r0, r0 + nres):
                    lastres = ends and r == r0 + nres - 1
                    put("N:CA", r, E.BOND_LENGTHS[0]); put("CA:C", r, E.BOND_LENGTHS[1])
                    put("tau", r, centre("tau", int(rb[E.SLOT["tau"], r])))
                    if not lastres:
                        put("0C:1N", r, E.BOND_LENGTHS[2])
                        put("CA:C:1N", r, centre("CA:C:1N", int(rb[E.SLOT["CA:C:1N"], r])))
                        put("psi", r, centre("psi", int(rb[E.SLOT["psi"], r])))
            else:
                tok_id = L.token_keys[int(tid[j])]
                geo = L.tokens[tok_id]
                for r in range(nres):
                    put("N:CA", r0 + r, geo["N:CA"][r]); put("CA:C", r0 + r, geo["CA:C"][r]); put("tau", r0 + r, geo["tau"][r])
                    if r < len(geo.get("0C:1N", [])):
                        put("0C:1N", r0 + r, geo["0C:1N"][r]); put("CA:C:1N", r0 + r, geo["CA:C:1N"][r])
                        put("psi", r0 + r, geo["psi"][r])
                    if r < len(geo.get("omega", [])):
                        put("omega", r0 + r, geo["omega"][r]); put("C:1N:1CA", r0 + r, geo["C:1N:1CA"][r])
                        put("phi", r0 + r, geo["phi"][r])
            seg[3 * r0] = (3 * r0, tok_id, nb)
            if not ends:                               # glue to the next token: snapped bin centres
                r = r0 + nres - 1
                for k in ("omega", "C:1N:1CA", "phi"):
                    put(k, r, centre(k, int(rb[E.SLOT[k], r])))
            if self._mode == "H" and nres > 1:         # glue inside a histogram-mode motif: bin centres too
                for r in range(r0, r0 + nres - 1):
                    for k in ("omega", "C:1N:1CA", "phi"):
                        put(k, r, centre(k, int(rb[E.SLOT[k], r])))
        # merge tree: residue leaves + this backbone's merge events in step order (data_structures.py:217-226)
        itid = E._u32(L.init_tid[a:b])
        name = (lambda t: t) if self._mode == "H" else (lambda t: L.token_keys[t])
        leaves = {3 * r: (3 * r, name(int(itid[r])), 3 if r < n - 1 else 2) for r in range(n)}
        th = TokenHierarchy(self.tokenizers[i], leaves)
        evl, evr, evp = self._event_arrays()
        for (new_n, e0, cnt, _key) in L.step_events:
            x0 = e0 + np.searchsorted(evl[e0:e0 + cnt], a)
            x1 = e0 + np.searchsorted(evl[e0:e0 + cnt], b)
            for x in range(x0, x1):
                i1, i2 = 3 * (int(evl[x]) - a), 3 * (int(evr[x]) - a)
                right = th.pop(i2)
                left = th[i1]
                tok = new_n if self._mode == "H" else (new_n, int(evp[x]))
                th[i1] = (i1, tok, left[2] + right[2])
        if dict(th) != seg:
            raise RuntimeError(f"merge-event replay and device token arrays disagree for backbone {i}")
        if getattr(L, "failed", None) is not None and L.failed[i]:
            raise ValueError("a glue angle of this backbone falls outside the vocabulary's bins (bpe.py:1105-1108)")
        return {"bond_to_token": th, "cur": cur, "init": init}

    def _event_arrays(self):
        L = self._learner
        if getattr(self, "_ev_cache", None) is None or self._ev_cache[0] != L.ev_n:
            n = L.ev_n
            self._ev_cache = (n, E._u32(L.ev_left[:n]), E._u32(L.ev_right[:n]), E._u32(L.ev_p[:n]))
        return self._ev_cache[1:]


class BPE(_View):
    def __init__(self, structures, bins, bin_strategy="histogram", save_dir=None, compute_sec_structs=False,
                 plot_iou_with_sec_structs=False, res_init=False, std_bonds=True, rmsd_partition_min_size=4,
                 rmsd_super_res=False, rmsd_only=False, num_partitions=3, max_num_strucs=500, glue_opt=False,
                 glue_opt_prior=0.0, glue_opt_every=10, glue_opt_method="all", seed=None):
        if not res_init:
            raise NotImplementedError("bond-level tokens (res_init=False) are the legacy path (SURVEY 2 #15)")
        if glue_opt:
            raise NotImplementedError("glue optimisation is the 'next' row f-1 of SURVEY 8(f)")
        if compute_sec_structs:
            raise NotImplementedError("secondary-structure priority is disabled by encode.sh:28")
        if rmsd_only:
            raise NotImplementedError("rmsd_only=True keeps raw geometry (near-unique keys); not built")
        if rmsd_partition_min_size == INF:
            self._mode = "H"
        elif rmsd_partition_min_size <= 2:
            self._mode = "M"
        else:
            raise NotImplementedError("mixed histogram/medoid vocabularies (2 < p_min_size < inf) are not built")
        self._vals, self._offsets = from_structures(structures)
        self.tokenizers = [Tokenizer(s, self, i) for i, s in enumerate(structures)]
        self.compute_sec_structs = compute_sec_structs
        self.plot_iou_with_sec_structs = plot_iou_with_sec_structs
        self.rmsd_partition_min_size = rmsd_partition_min_size
        self.rmsd_super_res = rmsd_super_res
        self.rmsd_only = rmsd_only
        self.glue_opt, self.glue_opt_every = glue_opt, glue_opt_every
        self.glue_opt_prior, self.glue_opt_method = glue_opt_prior, glue_opt_method
        self.num_partitions = ThresholdDict(num_partitions) if isinstance(num_partitions, dict) else num_partitions
        self.max_num_strucs = max_num_strucs
        self.res_init = res_init
        self.std_bonds = std_bonds
        self.bins = bins
        self.bin_strategy = bin_strategy
        self.n = len(self.tokenizers)
        self.seed = seed
        self.rng = np.random.default_rng(seed)         # bpe.py:84
        self.save_dir = save_dir
        self._step = 0
        self._times = []
        self._ious = []
        self._learner = None
        self._export = None

    # `_tokens` is advanced by the device-driven learn loop: reading it drains the loop first
    @property
    def _tokens(self):
        L = self.__dict__.get("_learner")
        if L is not None:
            L.tokens                    # drains the device loop and formats the entries it logged (same dict object)
        return self.__dict__.get("_tokens_")

    @_tokens.setter
    def _tokens(self, v):
        self.__dict__["_tokens_"] = v

    def __getstate__(self):
        """pickle layout of encode.py:403-427: vocabulary + hyper-parameters + materialised tokenizers; the device
        state (HBM buffers) is not part of the pickle"""
        d = dict(self.__dict__)
        if self._learner is not None:
            self._learner.tokens
            if not isinstance(self._learner, E.Inducer):
                d["_resume"] = self._learner.export_state()       # lets step() continue after pickle.load (encode.py:239-260)
        for k in ("_learner", "_export", "_ev_cache"):
            d[k] = None
        d["_tokens"] = d.pop("_tokens_", None)
        return d

    def __setstate__(self, d):
        d = dict(d)
        if "_tokens" in d:
            d["_tokens_"] = d.pop("_tokens")
        self.__dict__.update(d)
        for t in self.tokenizers:
            if getattr(t, "_mat", None) is None:
                t._bpe = self

    # ---- bpe.py:91-103 -------------------------------------------------------------------------------
    def initialize(self, path=None):
        self._learner = E.Learner(self._vals, self._offsets, self.bins, self.bin_strategy, self.std_bonds,
                                  mode=self._mode, rmsd_super_res=self.rmsd_super_res,
                                  num_partitions=self.num_partitions, max_num_strucs=self.max_num_strucs,
                                  rng=self.rng)
        L = self._learner
        self._thresholds = ThresholdDict(L.th.thresholds)
        self._bin_counts = ThresholdDict(L.th.bin_counts)
        if self.std_bonds:
            for i in range(3):
                self._thresholds[E.BOND_TYPES[i]] = [(E.BOND_LENGTHS[i], E.BOND_LENGTHS[i])]
        import torch
        self._bin_centers, self._bin_weights = ThresholdDict(), ThresholdDict()       # bpe.py:871-872
        for size in L.th.thresholds:
            self._bin_centers[size] = {k: torch.tensor(v, dtype=torch.float32).mean(axis=-1) for k, v in L.th.thresholds[size].items()}
            self._bin_weights[size] = {k: torch.tensor(v, dtype=torch.float32) / sum(v) for k, v in L.th.bin_counts[size].items()}
        self._tokens = L.tokens
        if self._mode == "M":
            self._sphere_dict = L.sphere_dict

    def bin(self):
        """bpe.py:1431-1474.  Pair counts are recomputed on the device at every step (K1), so there is no
        persistent _geo_dict to build."""
        return None

    def step(self):
        t0 = time.perf_counter()
        L = self._learner
        if L is None:
            L = self._resume_learner()
        L.step()
        if self._mode == "M":                          # bpe.py:2164-2166: recurse while the top key is recurring
            while L.ts.T > L.bb.S and L.peek_m()[2] is not None:
                L.step()
        self._step = self._learner.n_steps
        self._times.append(time.perf_counter() - t0)

    def _resume_learner(self):
        """bin/encode.py:239-260 reloads the newest bpe_iter={t}.pkl and keeps stepping: the inputs are re-uploaded and
        re-binned (deterministic), segmentation / merge events / vocabulary come from the pickle"""
        st = self.__dict__.get("_resume")
        if st is None:
            raise RuntimeError("this BPE was unpickled without learn state (initialize() was never run): call initialize() first")
        self._learner = E.Learner(self._vals, self._offsets, self.bins, self.bin_strategy, self.std_bonds, mode=self._mode,
                                  rmsd_super_res=self.rmsd_super_res, num_partitions=self.num_partitions,
                                  max_num_strucs=self.max_num_strucs, rng=self.rng, resume=st)
        L = self._learner
        self._tokens = L.tokens
        if self._mode == "M":
            self._sphere_dict = L.sphere_dict
        self._export = self._ev_cache = None
        if isinstance(self.n, int) and len(self.tokenizers) == L.bb.S:     # (not after an induce --append)
            for i, t in enumerate(self.tokenizers):
                if not getattr(t, "_local", False):
                    t._bpe, t._i, t._mat, t._cache_step = self, i, None, None
        return L

    @property
    def pop_log(self):
        if self._learner is None and self.__dict__.get("_resume") is not None:
            return self._resume["pop_log"]
        return self._learner.pop_log

    def glue_opt_all(self):
        return None

    def visualize(self, *a, **k):
        return None

    plot_times = plot_iou = visualize

    # ---- bpe.py:879-956 -------------------------------------------------------------------------------
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

    def capacity(self, tokenizer=False):
        """bpe.py:885-902: bits to store the vocabulary geometries (4-byte floats) + with `tokenizer` the token streams at
        uniform entropy (log2 |V| per motif, log2 bins per glue angle)"""
        total = 0
        for token in self._tokens.values():
            n = Tokenizer.num_bonds(token)
            total += 4 * (n + n - 1 + n - 2) * 8
        if tokenizer:
            mbits, bbits = np.log2(len(self._tokens)), np.log2(self.bins[1])
            for m in self._motif_counts():
                total += mbits * m
                total += 3 * (m - 1) * bbits
        return total

    def _motif_counts(self):
        """motifs per backbone ((len(t.tokenize()) + 3) // 4 in the reference): from the device token arrays when attached"""
        if self._learner is not None and len(self.tokenizers) == self._learner.bb.S:
            pos = self._token_arrays()[0]
            return np.diff(np.searchsorted(pos, self._learner.bb.offsets_host)).tolist()
        return [len(t.bond_to_token) for t in self.tokenizers]

    def dequantize(self, quantized):
        """bpe.py:959-983: ids -> ("MOTIF", token id) / ("DIHEDRAL" | "BOND_ANGLE", type, bin centre)"""
        cum, num_vocab = self.cum_bin_count(), self.vocab_size
        keys = list(self._tokens)
        out = []
        for i, q in enumerate(quantized):
            if q < num_vocab - cum:
                if q > len(keys):
                    raise ValueError(f"pos {i} > vocab range=(0, {len(keys)})")
                out.append(("MOTIF", keys[q]))
                continue
            c, token = q - (num_vocab - cum), None
            for k, v in self._thresholds[1].items():
                if self.res_init and k not in ["omega", "phi", "C:1N:1CA"]:
                    continue
                if c < len(v):
                    s0, e0 = v[c]
                    token = ("DIHEDRAL" if k in E.DIHEDRAL_ANGLES else "BOND_ANGLE", k, (s0 + e0) / 2)
                    break
                c -= len(v)
            if token is None:
                raise ValueError(f"pos {i} > vocab_size={num_vocab}")
            out.append(token)
        return out

    def recover(self, tokenized):
        """bpe.py:986-1002: token stream -> per-type value lists (motif geometries concatenated with the glue values)"""
        repl = {}
        for token in tokenized:
            if token[0] == "MOTIF":
                key_dict = self._tokens[token[1]]
                while isinstance(key_dict, str):
                    key_dict = json.loads(key_dict)
                for k in key_dict:
                    repl.setdefault(k, []).extend(key_dict[k])
            else:
                repl.setdefault(token[1], []).append(token[2])
        return repl

    init_structure = staticmethod(Tokenizer.init_structure)

    def recover_structure(self, repl, tokenized):
        """bpe.py:1029-1051: a stand-alone Tokenizer rebuilt from recovered value lists + the motif sequence"""
        n = len(repl["N:CA"])
        t_new = Tokenizer(self.init_structure(n))
        cur = t_new._detach()["cur"]
        cur["N:CA"][:n - 1] = repl["N:CA"][1:]
        cur["CA:C"][:n - 1] = repl["CA:C"][1:]
        cur["0C:1N"][:n - 1] = repl["0C:1N"]
        cur["phi"][1:] = repl["phi"]
        cur["psi"][:n - 1] = repl["psi"]
        cur["omega"][:n - 1] = repl["omega"]
        cur["tau"][:n - 1] = repl["tau"][1:]
        cur["CA:C:1N"][:n - 1] = repl["CA:C:1N"]
        cur["C:1N:1CA"][:n - 1] = repl["C:1N:1CA"]
        t_new.bond_to_token = {}
        at = 0
        for key, *rest in tokenized:
            if key == "MOTIF":
                nb = Tokenizer.num_bonds(self._tokens[rest[0]])
                t_new.bond_to_token[at] = (at, rest[0], nb)
                at += nb
        return t_new

    def token_stream(self):
        """quantize(t.tokenize()) of EVERY backbone in one device pass (csrc/k_export.cu): (ids int64 device tensor, offsets
        int64 numpy [S+1]); the reference loops list(_tokens).index() per token in a process pool (bpe.py:928-945)"""
        import torch
        L = self._learner
        ts = L.ts
        T, S = ts.T, L.bb.S
        V = len(self._tokens)
        out = E._empty(4 * T - 3 * S, torch.int64)
        E._lib.call("gb_token_stream", ts.p("tid"), ts.p("glue"), ts.p("pos"), E._lib.ptr(L.bb.bb), T, V, self.cum_bin_count("omega"),
                    self.cum_bin_count("phi"), self.cum_bin_count("C:1N:1CA"), E._lib.ptr(out))
        first = np.searchsorted(self._token_arrays()[0], L.bb.offsets_host)
        return out, 4 * first.astype(np.int64) - 3 * np.arange(S + 1, dtype=np.int64)

    def reconstruction_rmsd(self):
        """backbone RMSD (Kabsch, N / CA / C atoms) of every backbone's current quantised geometry against its pristine
        structure, one device pass -- what bin/encode.py:341-359 computes for its reference tokenizers one by one"""
        return self._learner.reconstruction()[0].cpu().numpy()

    def reconstruction_lddt(self):
        """mean lDDT-C-alpha of every backbone's current geometry against its pristine structure (bin/encode.py:351-358), one device pass"""
        return self._learner.lddt_ca().cpu().numpy()

    def codebook_utility(self):
        """get_codebook_utility (plotting.py:78-95) over the token streams of all backbones, computed on the device"""
        from foldingdiff.plotting import get_codebook_utility
        return get_codebook_utility(self.token_stream()[0], self.vocab_size)

    def quantize(self, tokenized):
        if isinstance(tokenized, Tokenizer):
            return self.quantize(tokenized.tokenize())
        if isinstance(tokenized, list) and tokenized and isinstance(tokenized[0], Tokenizer):
            if (self._learner is not None and len(tokenized) == self._learner.bb.S == len(self.tokenizers)
                    and all(a is b for a, b in zip(tokenized, self.tokenizers)) and not any(getattr(t, "_local", False) for t in tokenized)):
                ids, offs = self.token_stream()          # all tokenizers of this BPE (bin/encode.py:343,411): one device pass
                ids = ids.cpu().numpy()
                return [ids[offs[i]:offs[i + 1]].tolist() for i in range(len(offs) - 1)]
            return [self.quantize(t) for t in tokenized]
        ids = {k: i for i, k in enumerate(self._tokens)}
        out = []
        for token in tokenized:
            if token[0] == "MOTIF":
                k = token[1]
                out.append(ids[(int(k[0]), int(k[1])) if isinstance(k, tuple) else k])
            else:
                relv = self._thresholds[1][token[1]]
                out.append(len(self._tokens) + self.cum_bin_count(token[1]) +
                           E.get_ind((token[2] + 2 * np.pi) % (2 * np.pi), relv))
        return out

    # ---- induce (bpe.py:1053-1140) -------------------------------------------------------------------
    def induce(self, structures):
        """Batched BPE.tokenize: encode held-out structures against this (frozen) vocabulary on the device.
        Returns (tokenizers, ok) -- tokenizers[i] raises ValueError on access when ok[i] is False, mirroring the
        reference's get_ind ValueError for out-of-range glue angles."""
        if not self.res_init or self._mode != "M":
            raise NotImplementedError("tokenize needs res_init and a medoid-mode vocabulary (bpe.py:1054-1055,1077)")
        vals, offs = from_structures(structures)
        ind = E.Inducer(vals, offs, self._tokens, self._sphere_dict, self._thresholds, self.std_bonds,
                        self.rmsd_super_res)
        view = _View()
        view._learner, view._vals, view._mode = ind, vals, "M"
        view.tokenizers = [Tokenizer(s, view, i) for i, s in enumerate(structures)]
        return view.tokenizers, ~ind.failed

    def tokenize(self, t):
        """bpe.py:1053: one structure (Tokenizer built from a structure dict) -> (Tokenizer, metrics)"""
        toks, ok = self.induce([t._structure])
        if not ok[0]:
            raise ValueError("value does not fall into any bin")
        tk = toks[0]
        ind = tk._bpe._learner
        L = [tk.n]
        for (e0, cnt) in ind.replay:                     # one backbone -> every event of a step belongs to it
            L.append(L[-1] - cnt)
        return tk, {"L": L}


# pickle class paths of the reference: foldingdiff.bpe.BPE, foldingdiff.tokenizer.Tokenizer
BPE.__module__ = "foldingdiff.bpe"
Tokenizer.__module__ = "foldingdiff.tokenizer"
