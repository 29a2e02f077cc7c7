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


class Tokenizer:
    """Per-backbone view (tokenizer.py:18-477).  State lives in HBM; Python objects are materialised on access."""

    def __init__(self, structure=None, bpe=None, index=None):
        self._bpe, self._i = bpe, index
        self._structure = structure
        if structure is not None:
            df = structure["angles"]
            self.n = len(df)
            self.fname = structure.get("fname")
            self.aa = structure.get("aa")
        self._cache_step = None

    # -- pickling: materialise, drop the device-backed owner (ref_tokenizers={t}.pkl, {idx}.pkl, bpe_iter={t}.pkl) ---
    def __getstate__(self):
        d = dict(self.__dict__)
        if d.get("_bpe") is not None and getattr(d["_bpe"], "_learner", None) is not None:
            d["_mat"] = self._state()
        d["_bpe"] = None
        d["_cache_step"] = None
        return d

    # -- materialisation ------------------------------------------------------------------------------
    def _state(self):
        b = self._bpe
        if b is None and getattr(self, "_mat", None) is not None:
            return self._mat                           # unpickled: frozen snapshot
        if b is None or b._learner is None:
            raise RuntimeError("Tokenizer is not attached to an initialised BPE")
        key = (id(b._learner), b._learner.n_steps, len(b._learner.pop_log))
        if self._cache_step != key:
            self._mat = b._materialise(self._i)
            self._cache_step = key
        return self._mat

    @property
    def bond_to_token(self):
        return self._state()["bond_to_token"]

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

    def _dihedral_angle(self, idx):
        return self.cur[E.DIHEDRAL_ANGLES[idx % 3]][(idx + 1) // 3]

    def _bond_angle(self, idx):
        if idx == 0:
            return self._init_bond_angle
        return self.cur[E.BOND_ANGLES[idx % 3]][(idx - 1) // 3]

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

    def compute_coords(self, index=0, length=INF, orig=False):   # tokenizer.py:347-363 (whole-residue windows)
        return self._bpe._compute_coords(self._i, index, length, orig)

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

    def _compute_coords(self, i, index, length, orig):
        if index != 0 or length != INF:
            raise NotImplementedError("host-side compute_coords mirrors the whole-chain call only")
        L = self._learner
        t = self.tokenizers[i]
        st = t._state()
        n = t.n
        vals = np.array([st["cur"][c] for c in E.COLUMNS], dtype=np.float64)
        if orig:
            vals = self._vals[:, int(L.bb.offsets_host[i]):int(L.bb.offsets_host[i + 1])].copy()
        import torch
        mini = E.Backbones(vals, np.array([0, n], dtype=np.int64))
        out = E._empty(3 * n * 3, torch.float64)
        r0 = torch.zeros(1, dtype=torch.int32, device=out.device)
        f3 = st["init"]
        E._lib.call("gb_window_coords", E._lib.ptr(mini.vals), n, n, E._lib.ptr(mini.bb), E._lib.ptr(mini.offsets),
                    E._lib.ptr(r0), 1, n, 3 * n, None, None, float(f3[0]), float(f3[1]), float(f3[2]), None, None, 0,
                    E._lib.ptr(out))
        return out.view(3 * n, 3).cpu().numpy()


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
        if self._learner is not None:
            self._learner.tokens
        d = dict(self.__dict__)
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
            raise RuntimeError("this BPE was unpickled: the device state is gone; resume by re-running initialize() "
                               "(learning is seconds on the GPU) -- induce/tokenize/quantize work on the pickle as is")
        L.step()
        if self._mode == "M":                          # bpe.py:2164-2166: recurse while the top key is recurring
            while L.ts.T > L.bb.S and L.peek_m()[2] is not None:
                L.step()
        self._step = self._learner.n_steps
        self._times.append(time.perf_counter() - t0)

    @property
    def pop_log(self):
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

    def quantize(self, tokenized):
        if isinstance(tokenized, Tokenizer):
            return self.quantize(tokenized.tokenize())
        if isinstance(tokenized, list) and tokenized and isinstance(tokenized[0], Tokenizer):
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
