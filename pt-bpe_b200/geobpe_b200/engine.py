"""Device-side state of the GeoBPE hot path and the host logic that drives the sm_100a kernels.

Everything numeric here runs through the C ABI (include/geobpe_b200.h); PyTorch only owns the device
buffers and the stream.  Host Python keeps what the reference keeps in Python objects and what must
follow Python semantics exactly: the vocabulary dicts, JSON key strings for the tie-break
(bpe.py:1149,1469), the numpy RNG stream and the k-medoids iterations on the device-computed distance
matrix (algo.py:184-211).
"""
import collections
import concurrent.futures
import itertools
import ctypes as C
import os
import json
import time
import types
import math

import numpy as np
import torch

from . import _lib

BOND_TYPES = ["N:CA", "CA:C", "0C:1N"]                # tokenizer.py:19
BOND_ANGLES = ["tau", "CA:C:1N", "C:1N:1CA"]          # tokenizer.py:21
DIHEDRAL_ANGLES = ["psi", "omega", "phi"]             # tokenizer.py:22
BOND_LENGTHS = [1.46, 1.54, 1.34]                     # nerf.py:17-19
COLUMNS = ["0C:1N", "N:CA", "CA:C", "phi", "psi", "omega", "tau", "CA:C:1N", "C:1N:1CA"]
COL = {c: i for i, c in enumerate(COLUMNS)}
# residue-major slot order of rbins / edges (include/geobpe_b200.h)
SLOTS = ["N:CA", "CA:C", "0C:1N", "tau", "CA:C:1N", "psi", "omega", "C:1N:1CA", "phi"]
SLOT = {c: i for i, c in enumerate(SLOTS)}
N_INIT = np.array([17.047, 14.099, 3.625])            # nerf.py:22-24
CA_INIT = np.array([16.967, 12.784, 4.338])
C_INIT = np.array([15.685, 12.755, 5.133])
TWO_PI = 2 * np.pi
NOBIN = 0x3FF
LAST = 0x80000000
EMPTY = 0xFFFFFFFFFFFFFFFF
M64 = (1 << 64) - 1
HASH_BASE = 0x1B873593A5C1F3D7                        # odd base of the histogram-mode content hash (mod 2^64)


def _angle_between(v1, v2):                           # angles_and_coords.py:746-752
    u1 = v1 / np.linalg.norm(v1)
    u2 = v2 / np.linalg.norm(v2)
    return np.arccos(np.clip(np.dot(u1, u2), -1.0, 1.0))


INIT_N_CA = float(np.linalg.norm(N_INIT - CA_INIT))   # tokenizer.py:74-77
INIT_CA_C = float(np.linalg.norm(CA_INIT - C_INIT))
INIT_TAU = float(_angle_between(N_INIT - CA_INIT, C_INIT - CA_INIT))


def get_ind(v, values):
    """bpe.py:1165-1189 on host (used for the handful of constants; arrays go through K3)."""
    import bisect
    left = [s for s, _ in values]
    ind = bisect.bisect_right(left, v) - 1
    if ind < 0:
        raise ValueError(f"value {v} is below the first bin range")
    s, e = values[ind]
    if ind == len(values) - 1 and v == e:
        return ind
    if s <= v < e:
        return ind
    raise ValueError(f"value {v} does not fall into any bin")


def tie_rank_table():
    """u16 [2][1024]: rank of the element strings "<v>," (row 0) and "<v>]" (row 1) in byte order.  A list element of the
    reference's key string json.dumps(key, sort_keys=True) prints as "<v>, " or "<v>]"; these strings are prefix-free, so
    the order of two key strings (bpe.py:1149,1469: the tie-break) is the lexicographic order of their element-rank
    sequences (csrc/k_hloop2.cu h2_decide)."""
    strs = sorted([(f"{v},".encode(), 0, v) for v in range(1024)] + [(f"{v}]".encode(), 1, v) for v in range(1024)])
    tab = np.zeros((2, 1024), dtype=np.uint16)
    for rank, (_, last, v) in enumerate(strs):
        tab[last, v] = rank
    return tab


def tie_rank_sequence(key_dict, tab=None):
    """the rank sequence h2_decide builds for a histogram-mode pair key (nine lists in sort_keys order); host mirror used
    by the tests to pin the device tie-break order against python's string order"""
    tab = tie_rank_table() if tab is None else tab
    seq = []
    for name in sorted(key_dict):
        vals = key_dict[name]
        seq += [int(tab[1 if k == len(vals) - 1 else 0, v]) for k, v in enumerate(vals)]
    return seq


def _dev():
    return torch.device("cuda", torch.cuda.current_device())


def _dbg(msg):
    """GEOBPE_DEBUG=1: synchronise and print a progress marker (multi-GPU bring-up)"""
    if os.environ.get("GEOBPE_DEBUG"):
        torch.cuda.synchronize()
        print(f"[geobpe rank {os.environ.get('RANK', 0)}] {msg}", flush=True)


def _empty(n, dtype):
    return torch.empty(max(int(n), 1), dtype=dtype, device=_dev())


def _zeros(n, dtype):
    return torch.zeros(max(int(n), 1), dtype=dtype, device=_dev())


def next_pow2(x):
    return 1 << max(int(x) - 1, 1).bit_length()


_STAGE = {}


def _upload_f64(arr, chunk=4 << 20):
    """host float64 array -> device tensor of the same shape.  Large inputs stream through two reusable pinned staging buffers
    (host memcpy of chunk k+1 overlaps the H2D copy of chunk k); pinning the whole array first costs more than the copy itself
    (measured: 1.6 s for the 3.7 GB of BASELINE configs[2], against 0.08 s for everything else in induce's set-up).  One
    thread's memcpy into the staging buffer runs at ~13 GB/s, a quarter of the link: the chunk is copied by up to eight threads
    (numpy releases the GIL for plain copies)."""
    flat = np.ascontiguousarray(arr, dtype=np.float64).reshape(-1)
    n = flat.size
    if n * 8 < (8 << 20):
        return torch.from_numpy(flat).to(_dev()).view(arr.shape)
    dev = _dev()
    if "buf" not in _STAGE:
        _STAGE["buf"] = [torch.empty(chunk, dtype=torch.float64).pin_memory() for _ in range(2)]
        nt = max(1, min(int(os.environ.get("GEOBPE_STAGE_THREADS", 8)), (os.cpu_count() or 1)))
        _STAGE["pool"] = concurrent.futures.ThreadPoolExecutor(nt) if nt > 1 else None
        _STAGE["nt"] = nt
    bufs, evs, pool, nt = _STAGE["buf"], [None, None], _STAGE["pool"], _STAGE["nt"]
    dst = torch.empty(n, dtype=torch.float64, device=dev)
    for k, off in enumerate(range(0, n, chunk)):
        b = k & 1
        if evs[b] is not None:
            evs[b].synchronize()
        m = min(chunk, n - off)
        stage, src = bufs[b][:m].numpy(), flat[off:off + m]
        if pool is None or m < (1 << 18):
            stage[:] = src
        else:
            cuts = [m * i // nt for i in range(nt + 1)]
            list(pool.map(lambda ab: np.copyto(stage[ab[0]:ab[1]], src[ab[0]:ab[1]]), zip(cuts[:-1], cuts[1:])))
        dst[off:off + m].copy_(bufs[b][:m], non_blocking=True)
        evs[b] = torch.cuda.Event()
        evs[b].record()
    return dst.view(arr.shape)


class Backbones:
    """Internal coordinates of a shard of backbones resident in HBM: vals f64 [9][R] (reference column
    order), CSR offsets, residue->backbone map.  72 B per residue."""

    def __init__(self, vals, offsets):
        _lib.require_cuda()
        vals = np.ascontiguousarray(vals, dtype=np.float64)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        assert vals.ndim == 2 and vals.shape[0] == 9 and offsets[-1] == vals.shape[1]
        self.R = int(vals.shape[1])
        self.S = int(len(offsets) - 1)
        self.offsets_host = offsets
        self.lens_host = np.diff(offsets)
        if self.R >= (1 << 32) - 2:
            raise ValueError("a shard holds at most 2^32-2 residues (u32 positions)")
        self.vals = _upload_f64(vals)
        self.offsets = torch.from_numpy(offsets).to(_dev())
        self.bb = _empty(self.R, torch.int32)
        _lib.call("gb_residue_backbone", _lib.ptr(self.offsets), self.S, self.R, _lib.ptr(self.bb))
        self.h2d_bytes = vals.nbytes + offsets.nbytes

    def backbone_of(self, r):
        return int(np.searchsorted(self.offsets_host, r, side="right") - 1)


class Thresholds:
    """K3 driver: bpe.py:820-876 _init_thresholds + plotting.py:305-337 save_histogram."""

    def __init__(self, bb, bins, bin_strategy, std_bonds, comm=None):
        self.keys = BOND_ANGLES + DIHEDRAL_ANGLES + ([] if std_bonds else BOND_TYPES)
        self.std_bonds = std_bonds
        self.thresholds = {}
        self.bin_counts = {}
        R, S = bb.R, bb.S
        mm = _empty(18, torch.float64)
        cnt = _empty(9, torch.int64)
        _lib.call("gb_col_minmax", _lib.ptr(bb.vals), R, R, _lib.ptr(mm), _lib.ptr(cnt))
        S_total = S
        if comm is not None and comm.world > 1:      # global min / max / backbone count
            if bin_strategy == "uniform":
                raise NotImplementedError("quantile edges need all values on one rank; use bin_strategy=histogram")
            mm2 = mm.view(9, 2)
            lo, hi = mm2[:, 0].contiguous(), mm2[:, 1].contiguous()
            comm.all_reduce(lo, "min"); comm.all_reduce(hi, "max")
            mm = torch.stack([lo, hi], dim=1).contiguous()
            st = torch.tensor([S], dtype=torch.int64, device=_dev())
            S_total = int(comm.all_reduce(st, "sum").item())
        mm = mm.cpu().numpy().reshape(9, 2)
        consts = {"tau": INIT_TAU, "N:CA": INIT_N_CA, "CA:C": INIT_CA_C}
        host_cols = None
        for size, B in bins.items():
            B = int(B)
            if B >= NOBIN - 1 or B > 512:
                raise NotImplementedError("the packed glue field holds at most 512 bins per type")
            edges_by_key = {}
            for key in self.keys:
                circ = key not in BOND_TYPES
                lo, hi = mm[COL[key]]
                if key in consts:
                    c = (consts[key] + TWO_PI) % TWO_PI if circ else consts[key]
                    lo, hi = min(lo, c), max(hi, c)
                if bin_strategy.startswith("histogram"):
                    rng_ = (0, TWO_PI) if ("cover" in bin_strategy and circ) else None
                    # same code path as np.histogram(a, bins=B): only min and max of a matter for the edges
                    edges = np.histogram_bin_edges(np.array([lo, hi]), bins=B, range=rng_)
                elif bin_strategy == "uniform":
                    if host_cols is None:
                        host_cols = bb.vals.cpu().numpy()
                    col = host_cols[COL[key]]
                    a = col[(col == col) & (col != 0.0)]
                    if key in consts:
                        a = np.concatenate([a, np.full(S, consts[key])])
                    if circ:
                        a = (a + 2 * np.pi) % (2 * np.pi)
                    edges = np.quantile(np.sort(a), np.linspace(0, 1, B + 1))   # plotting.py:276
                else:
                    raise NotImplementedError(bin_strategy)
                edges_by_key[key] = np.asarray(edges, dtype=np.float64)
            estride = B + 1
            e_col = np.zeros((9, estride))
            nb = (C.c_int * 9)(*([0] * 9))
            for key in self.keys:
                e_col[COL[key]] = edges_by_key[key]
                nb[COL[key]] = B
            e_dev = torch.from_numpy(e_col).to(_dev())
            counts = _empty(9 * estride, torch.int64)
            _lib.load().gb_col_hist  # (prototype check)
            _lib.call("gb_col_hist", _lib.ptr(bb.vals), R, R, _lib.ptr(e_dev), estride, nb, _lib.ptr(counts))
            if comm is not None and comm.world > 1:
                comm.all_reduce(counts, "sum")
            counts = counts.cpu().numpy().reshape(9, estride)
            th, bc = {}, {}
            for key in self.keys:
                e = edges_by_key[key]
                th[key] = [(float(s), float(t)) for s, t in zip(e[:-1], e[1:])]
                cts = counts[COL[key], :B].copy()
                if key in consts:
                    circ = key not in BOND_TYPES
                    c = (consts[key] + TWO_PI) % TWO_PI if circ else consts[key]
                    cts[get_ind(c, th[key])] += S_total
                bc[key] = [int(x) for x in cts]
            self.thresholds[size] = th
            self.bin_counts[size] = bc

    def residue_major_edges(self, size):
        """edges f64 [9][B+1] in residue-major slot order + nb int[9] for gb_quantize_residues."""
        th = self.thresholds[size]
        B = max(len(v) for v in th.values())
        e = np.zeros((9, B + 1))
        nb = [0] * 9
        for key, iv in th.items():
            e[SLOT[key], :len(iv)] = [s for s, _ in iv]
            e[SLOT[key], len(iv)] = iv[-1][1]
            nb[SLOT[key]] = len(iv)
        return e, nb, B + 1


class PairTable:
    """K1 driver: open-addressing (key, count, min position) table in HBM + argmax outputs."""
    TIE_CAP = 1 << 16

    def __init__(self):
        self.M = 0
        self.out4 = _zeros(4, torch.int32)
        self.tie_keys = _empty(self.TIE_CAP, torch.int64)
        self.tie_min = _empty(self.TIE_CAP, torch.int32)
        self.overflow = _zeros(1, torch.int32)
        self.distinct = 0

    def ensure(self, M):
        M = next_pow2(max(M, 1 << 12))
        if M != self.M:
            self.M = M
            self.keys = _empty(M, torch.int64)
            self.cnt = _empty(M, torch.int32)
            self.min = _empty(M, torch.int32)

    def ptrs(self):
        return _lib.ptr(self.keys), _lib.ptr(self.cnt), _lib.ptr(self.min), self.M

    NFIRST = 16

    def count(self, launch, n_pairs_bound, first_pos=None, pos_ptr=None, len_ptr=None):
        """clear + count (+ retry on overflow with a larger table) + argmax (+ first positions), ONE D2H.
        `launch(table)` issues the count kernel.  Returns (max_count, ties) with ties = [(key, minpos, pos, nres)]
        (pos / nres = start residue and residue count of the example window; 0 when pos_ptr is None)."""
        want = 4 * max(self.distinct, 1024) + 4096
        first = self.M == 0
        self.ensure(2 * n_pairs_bound if first else max(want, 1 << 12))
        if not hasattr(self, "res"):
            self.res = _zeros(5 + 4 * self.NFIRST, torch.int64)
        while True:
            _lib.call("gb_table_clear", *self.ptrs())
            self.overflow.zero_()
            launch(self)
            _lib.call("gb_table_argmax", *self.ptrs(), _lib.ptr(self.out4), _lib.ptr(self.tie_keys),
                      _lib.ptr(self.tie_min), self.TIE_CAP)
            if first_pos is not None:
                first_pos(self)
            _lib.call("gb_pack_result", _lib.ptr(self.out4), _lib.ptr(self.overflow), _lib.ptr(self.tie_keys),
                      _lib.ptr(self.tie_min), pos_ptr, len_ptr, self.NFIRST, None, _lib.ptr(self.res))
            res = self.res.cpu().numpy()
            mx, nt, distinct, ovf = int(res[0]), int(res[1]), int(res[2]), int(res[4])
            if ovf or distinct * 2 > self.M:      # keep the load factor below 1/2
                self.ensure(self.M * 4)
                continue
            break
        self.distinct = distinct
        if nt > self.TIE_CAP:
            raise RuntimeError(f"{nt} keys tied at count {mx}: tie list capacity {self.TIE_CAP} exceeded")
        if nt <= self.NFIRST:
            r = res[5:5 + 4 * nt].reshape(nt, 4)
            ties = [(int(k) & M64, int(m), int(p), int(n)) for k, m, p, n in r.tolist()]
            return mx, ties
        tk = self.tie_keys[:nt].cpu().numpy().astype(np.uint64).tolist()
        tm = (self.tie_min[:nt].cpu().numpy().astype(np.int64) & 0xFFFFFFFF).tolist()
        return mx, [(k, m, None, None) for k, m in zip(tk, tm)]

    def lookup(self, keys):
        if not keys:
            return [], []
        q = torch.tensor(np.array(keys, dtype=np.uint64).view(np.int64), device=_dev())
        oc = _empty(len(keys), torch.int32)
        om = _empty(len(keys), torch.int32)
        _lib.call("gb_table_lookup", *self.ptrs(), _lib.ptr(q), len(keys), _lib.ptr(oc), _lib.ptr(om))
        return ((oc.cpu().numpy().astype(np.int64) & 0xFFFFFFFF).tolist(),
                (om.cpu().numpy().astype(np.int64) & 0xFFFFFFFF).tolist())


class SlotSketch:
    """K1 driver, sketch design (csrc/k1_slots.cu): pass 1 REDs every pair into key-less slot counters, the scan
    lists the slots holding the max, pass 2 counts the keys of those slots exactly (small PairTable).  Exact by
    construction: a key outside the listed slots has count <= its slot's count < the listed threshold."""
    CAP = 4096               # candidate slots per round (k_cand_count's shared set holds 2 * CAP)
    STREAM_BYTES = 96 << 20  # token arrays above this size are read with evict-first loads (slots stay in L2)
    STAGE_PAIRS = 296 * 32768  # the per-CTA shared-memory stage of k_slot_count pays above this many pairs

    def __init__(self, n_pairs_bound, log2ns=None, variant=2, ctas_per_sm=2):
        if log2ns is None:
            log2ns = max(16, min(24, (max(n_pairs_bound, 2) - 1).bit_length() - 1))   # NS >= pairs / 2; <= 64 MB (L2)
            log2ns = int(os.environ.get("GEOBPE_LOG2NS", log2ns))                     # tuning knob (scripts/)
        self.log2ns = log2ns
        # the per-CTA shared-memory stage only pays when a CTA sees many more pairs than it has cache entries
        self.variant = variant if n_pairs_bound >= self.STAGE_PAIRS else (variant & ~6)
        self.ctas = ctas_per_sm
        self.slots = _zeros(1 << log2ns, torch.int32)
        self.bmax = _zeros(max(1, (1 << log2ns) // 4096), torch.int32)
        self.sout = _zeros(4, torch.int32)
        self.cand = _zeros(self.CAP, torch.int32)
        self.small = PairTable()
        self.small.ensure(1 << 13)
        self.small.res = _zeros(5 + 4 * PairTable.NFIRST, torch.int64)
        self.rounds = [0, 0, 0]      # iterations resolved in one round / two rounds / by the exact-table fallback
        self.nonzero = 0

    def count(self, pass1, pass2, pos_ptr=None, len_ptr=None):
        """pass1(slots_ptr, log2ns, variant, ctas) issues k_slot_count; pass2(log2ns, cand_ptr, ncand_ptr, cap, table)
        issues k_cand_count.  Returns (max_count, ties) like PairTable.count, or None when more than CAP slots tie at
        the top (flat distributions: the caller falls back to the exact table)."""
        sm = self.small
        self.slots.zero_()
        pass1(_lib.ptr(self.slots), self.log2ns, self.variant, self.ctas)
        _lib.call("gb_slots_top", _lib.ptr(self.slots), self.log2ns, 0, _lib.ptr(self.sout), _lib.ptr(self.cand), self.CAP,
                  _lib.ptr(self.bmax))
        rnd = 0
        while True:
            _lib.call("gb_table_clear", *sm.ptrs())
            sm.overflow.zero_()
            pass2(self.log2ns, _lib.ptr(self.cand), self.sout.data_ptr() + 4, self.CAP, sm)
            _lib.call("gb_table_argmax", *sm.ptrs(), _lib.ptr(sm.out4), _lib.ptr(sm.tie_keys), _lib.ptr(sm.tie_min), sm.TIE_CAP)
            _lib.call("gb_pack_result", _lib.ptr(sm.out4), _lib.ptr(sm.overflow), _lib.ptr(sm.tie_keys), _lib.ptr(sm.tie_min),
                      pos_ptr, len_ptr, sm.NFIRST, _lib.ptr(self.sout), _lib.ptr(sm.res))
            res = sm.res.cpu().numpy()
            m1, nt, nd, ovf = int(res[0]), int(res[1]), int(res[2]), int(res[4])
            c1, ncand = int(res[3]) & 0xFFFFFFFF, (int(res[3]) >> 32) & 0xFFFFFFFF
            if c1 == 0:
                return 0, []
            if ovf == 2 or ncand > self.CAP:
                self.rounds[2] += 1
                return None
            if ovf == 1 or nd * 2 > sm.M:
                sm.ensure(sm.M * 4)
                continue
            if m1 == c1 or rnd == 1:
                break
            # every top slot holds a collision: list all slots that could hold a key with count >= m1 and recount
            rnd = 1
            _lib.call("gb_slots_top", _lib.ptr(self.slots), self.log2ns, m1, _lib.ptr(self.sout), _lib.ptr(self.cand), self.CAP,
                      _lib.ptr(self.bmax))
        self.rounds[rnd] += 1
        if nt > sm.TIE_CAP:
            raise RuntimeError(f"{nt} keys tied at count {m1}: tie list capacity {sm.TIE_CAP} exceeded")
        if nt <= sm.NFIRST:
            r = res[5:5 + 4 * nt].reshape(nt, 4)
            return m1, [(int(k) & M64, int(m), int(p), int(n)) for k, m, p, n in r.tolist()]
        tk = sm.tie_keys[:nt].cpu().numpy().astype(np.uint64).tolist()
        tm = (sm.tie_min[:nt].cpu().numpy().astype(np.int64) & 0xFFFFFFFF).tolist()
        return m1, [(k, m, None, None) for k, m in zip(tk, tm)]


class TokenState:
    """Live-token SoA in HBM (ping-pong buffers for the merge compaction)."""
    FIELDS = ("cls", "hsh", "glue", "pos", "len", "tid")
    DTYPES = {"cls": torch.int32, "hsh": torch.int64, "glue": torch.int32, "pos": torch.int32,
              "len": torch.int32, "tid": torch.int32}

    def __init__(self, R, mode):
        self.mode = mode                     # "H" or "M"
        self.cap = R + 8                     # + padding so 128-bit loads at the tail stay in bounds
        self.T = 0
        use = [f for f in self.FIELDS if not ((f == "cls" and mode == "H") or (f == "hsh" and mode == "M"))]
        self.a = {f: _zeros(self.cap, self.DTYPES[f]) for f in use}
        self.b = {f: _zeros(self.cap, self.DTYPES[f]) for f in use}
        self.flags = _zeros(self.cap, torch.uint8)
        self.aux = _zeros(self.cap, torch.int32)
        self.scratch = _empty(_lib.load().gb_select_scratch_elems(self.cap), torch.int32)
        self.idx = _empty(self.cap, torch.int32)      # occurrence list (all flagged)
        self.kept = _empty(self.cap, torch.int32)     # kept occurrences
        self.u32 = _zeros(8, torch.int32)

    def p(self, f, out=False):
        d = self.b if out else self.a
        return _lib.ptr(d[f]) if f in d else None

    def swap(self):
        self.a, self.b = self.b, self.a

    def select(self, mask, out):
        _lib.call("gb_select_flagged", _lib.ptr(self.flags), self.T, mask, _lib.ptr(out), _lib.ptr(self.u32),
                  _lib.ptr(self.scratch))
        return int(self.u32[0].item())


def _u32(t):
    """int32 device tensor holding u32 values -> int64 numpy"""
    return t.cpu().numpy().astype(np.int64) & 0xFFFFFFFF


class Learner:
    """GeoBPE learn state for one shard of backbones.

    mode "H": rmsd_partition_min_size = inf  (histogram-quantised motifs; std_bonds only, as in the reference)
    mode "M": rmsd_partition_min_size = 0    (every motif is a medoid token; rmsd_super_res, not rmsd_only)
    """

    def __init__(self, vals, offsets, bins, bin_strategy="histogram", std_bonds=True, mode="H",
                 rmsd_super_res=True, num_partitions=3, max_num_strucs=500, rng=None, verify=True, comm=None,
                 use_sketch=True, async_steps=True, resume=None):
        # (GEOBPE_FORCE_COMM=1 keeps a one-rank communicator: the sharded code paths then run on a single GPU, for tests)
        self.comm = comm if (comm is not None and (comm.world > 1 or os.environ.get("GEOBPE_FORCE_COMM") == "1")) else None
        if self.comm is not None and mode == "M" and resume is not None:
            raise NotImplementedError("resuming a sharded medoid-mode run is not built (checkpoints are per process)")
        if len(bins) != 1:
            raise NotImplementedError("size-bucketed --bins (more than one bucket) is not built yet: keys would "
                                      "need re-quantisation per bucket (SURVEY 7.2 hard part 2)")
        if mode == "H" and not std_bonds:
            raise NotImplementedError("histogram mode with free bonds raises KeyError in the reference "
                                      "(bpe.py:1516-1517); only std_bonds=True is defined")
        if mode == "M" and not rmsd_super_res:
            raise NotImplementedError("medoid mode is built for rmsd_super_res=True (README setting)")
        self.mode = mode
        self.hash_base = HASH_BASE
        self.std_bonds = std_bonds
        self.bins = dict(bins)
        self.size0 = next(iter(bins))
        self.num_partitions = num_partitions
        self.max_num_strucs = max_num_strucs
        self.rng = rng
        self.verify = verify
        self.bb = Backbones(vals, offsets)
        self.vals_host = np.ascontiguousarray(vals, dtype=np.float64)
        self.sphere_n = {}
        self.k1_events = None
        self._peek = None
        self.use_sketch = use_sketch
        self.k1b_events = None
        self.sketch = None
        self._dist_layout()
        self.th = Thresholds(self.bb, self.bins, bin_strategy, std_bonds, self.comm)
        self.table = PairTable()
        self._tokens = {}           # BPE._tokens
        self._pop_log = []          # (count, key string) per popped key
        self._step_events = []      # (token id / n, ev_start, ev_count, key)
        self.n_steps = 0
        self.hl = None              # device-driven loop state (histogram mode, single GPU)
        self._pending = 0           # iterations enqueued on the device and not yet read back
        self.async_steps = async_steps and mode == "H" and use_sketch
        self.use_graphs = True
        # device-driven loop: 2 = persistent cooperative kernel with incrementally maintained pair counts (csrc/k_hloop2.cu, one GPU),
        # 1 = per-iteration candidate pass replayed from CUDA graphs (csrc/k_hloop.cu; the sharded loop)
        self.loop_version = int(os.environ.get("GEOBPE_LOOP", 2))
        self._quantize()
        self._ts = TokenState(self.bb.R, mode)
        R = self.bb.R
        self._evl = _empty(R, torch.int32)
        self._evr = _empty(R, torch.int32)
        self._evp = _empty(R, torch.int32)
        self._ev_n = 0
        if resume is not None:
            self.hash_base = resume.get("hash_base", self.hash_base)
        if mode == "H":
            self._init_tokens_h()
        elif resume is None:
            self._init_tokens_m()
        if resume is not None:
            self._load_state(resume)


    def _dist_layout(self):
        """where this shard sits in the whole set: the occurrence sets of medoid mode hold GLOBAL (backbone, bond index) pairs and
        the medoid samples are drawn over the global occurrence order (contiguous shards in rank order = the reference's order)"""
        offs = self.bb.offsets_host
        if self.comm is None:
            self.bb_base, self.r_base, self.g_offsets = 0, 0, np.asarray(offs, dtype=np.int64)
            return
        lens = self.comm.all_gather_object(np.diff(offs).astype(np.int64))
        rk = self.comm.rank
        self.bb_base = int(sum(len(l) for l in lens[:rk]))
        self.r_base = int(sum(int(l.sum()) for l in lens[:rk]))
        self.g_offsets = np.concatenate([[0], np.cumsum(np.concatenate(lens))]).astype(np.int64)

    def _gather_rows(self, rows, mine, n):
        """[n, width] float64 on every rank from the rows each rank owns (`mine`: their indices in 0..n): every row has exactly
        one owner, so a sum of the raw bit patterns over the ranks reproduces it exactly (signed zeros included)"""
        out = torch.zeros((n, rows.shape[1]), dtype=torch.float64, device=_dev())
        if len(mine):
            out[torch.from_numpy(np.asarray(mine, dtype=np.int64)).to(_dev())] = rows
        if self.comm is not None:
            self.comm.all_reduce(out.view(torch.int64))
        return out

    def _gather_geos(self, mine, n):
        """medoid geometries computed by the ranks that hold their raw angles ({index: geo}) -> list of all n on every rank"""
        merged = {}
        for part in self.comm.all_gather_object(mine):
            merged.update(part)
        if len(merged) != n:
            raise RuntimeError(f"{len(merged)} of {n} medoid windows were claimed by a rank")
        return [merged[i] for i in range(n)]

    # ---- state that the device-driven loop advances: reading it first drains the loop -------------------------
    def _drain(self):
        if self._pending:
            self.sync()

    @property
    def tokens(self):
        self._drain(); self._materialise_log(); return self._tokens

    @property
    def pop_log(self):
        self._drain(); self._materialise_log(); return self._pop_log

    def _materialise_log(self):
        """vocabulary entries / pop log of the iterations the device loop ran: the device logs (count, example window)
        per iteration; the reference's key dict and its json string are formatted here, on first use"""
        raw = getattr(self, "_raw", None)
        if not raw:
            return
        self._raw = []
        for cnt, r0, nres, *glob in raw:
            dicts = self._h_key_dicts(r0.tolist(), nres.tolist(), glob=bool(glob and glob[0])) if nres is not None else None
            for s in range(len(cnt)):
                if dicts is not None:                  # (same string / dict as json.dumps(d, sort_keys=True) and its json.loads:
                    items = sorted(dicts[s].items())   #  the values are lists of python ints, which str() prints like json)
                    key_str = "{" + ", ".join(f'"{k}": {v}' for k, v in items) + "}"
                    self._tokens[len(self._tokens)] = dict(items)
                else:
                    key_str = r0[s]
                    self._tokens[len(self._tokens)] = json.loads(key_str)
                self._check_not_popped(key_str)
                self._pop_log.append((int(cnt[s]), key_str))

    def _check_not_popped(self, key_str):
        """histogram mode: [A][BC] and [AB][C] share a key, so a popped key could become the argmax again; the reference stops
        there (`if key in self._tokens.values(): breakpoint()`, bpe.py:1808) -- here it raises"""
        seen = self.__dict__.setdefault("_popped", set())
        if key_str in seen:
            raise RuntimeError("a key that already owns a token became the most frequent pair again (bpe.py:1808 invariant)")
        seen.add(key_str)

    @property
    def step_events(self):
        self._drain(); return self._step_events

    @property
    def ts(self):
        self._drain(); return self._ts

    @property
    def ev_n(self):
        self._drain(); return self._ev_n

    # merge-event log: the device loops append an iteration's events in arbitrary order; they are put into (iteration,
    # position) order on the device the first time somebody reads them
    @property
    def ev_left(self):
        self._drain(); self._sort_pending(); return self._evl

    @property
    def ev_right(self):
        self._drain(); self._sort_pending(); return self._evr

    @property
    def ev_p(self):
        self._drain(); return self._evp

    def _sort_pending(self):
        todo, self._unsorted = getattr(self, "_unsorted", []), []
        for e0, e1, nk in todo:
            self._sort_events(e0, e1, nk)

    @ev_n.setter
    def ev_n(self, v):
        self._ev_n = v

    # ---- K3 ------------------------------------------------------------------------------------
    def _quantize(self, bond_mode=None):
        bb = self.bb
        R, S = bb.R, bb.S
        e, nb, estride = self.th.residue_major_edges(self.size0)
        self.edges_rm = e
        self.nb_rm = nb
        self.estride = estride
        self.edges_dev = torch.from_numpy(e).to(_dev())
        self.rbins = _empty(9 * R, torch.int16)
        self.glue_res = _empty(R, torch.int32)
        self.sym = _empty(R, torch.int64)
        self.status = _zeros(S, torch.int32)
        if bond_mode is None:
            bond_mode = 0 if self.std_bonds else 1
        _lib.call("gb_quantize_residues", _lib.ptr(bb.vals), R, R, _lib.ptr(bb.bb), _lib.ptr(bb.offsets), S,
                  _lib.ptr(self.edges_dev), estride, (C.c_int * 9)(*nb), bond_mode, INIT_N_CA, INIT_CA_C, INIT_TAU,
                  _lib.ptr(self.rbins), _lib.ptr(self.glue_res), _lib.ptr(self.sym), _lib.ptr(self.status))
        st = self.status.cpu().numpy()
        self.status_host = st
        if bond_mode != 2 and st.any():
            i = int(np.nonzero(st)[0][0])
            raise ValueError(f"value {'is below the first bin range' if st[i] & 5 else 'does not fall into any bin'} "
                             f"(backbone {i})")   # get_ind's ValueErrors, bpe.py:1179-1189

    def centre(self, key, b):
        s, e = self.th.thresholds[self.size0][key][b]
        return (s + e) / 2                           # sum(relv[ind]) / 2, bpe.py:1503

    def window_bins(self, r0, nres):
        """rbins of residues [r0, r0+nres) -> int array [9][nres] (host)"""
        rb = self.rbins.view(9, self.bb.R)[:, r0:r0 + nres].cpu().numpy().astype(np.int64) & 0xFFFF
        return rb

    # ---- histogram mode --------------------------------------------------------------------------
    def _init_tokens_h(self):
        """bpe.py:229-260,393: residue tokens labelled in first-seen order of their bin-centre key."""
        R = self.bb.R
        tab = self.table
        tab.ensure(1 << 16)
        while True:
            _lib.call("gb_table_clear", *tab.ptrs())
            tab.overflow.zero_()
            _lib.call("gb_key_count", _lib.ptr(self.sym), R, *tab.ptrs(), _lib.ptr(tab.overflow))
            ok = _empty(tab.M, torch.int64); oc = _empty(tab.M, torch.int32); om = _empty(tab.M, torch.int32)
            tot = _zeros(1, torch.int32)
            _lib.call("gb_table_dump", *tab.ptrs(), _lib.ptr(ok), _lib.ptr(oc), _lib.ptr(om), _lib.ptr(tot))
            n = int(tot.item())
            if int(tab.overflow.item()) or 2 * n > tab.M:
                tab.ensure(tab.M * 4)
                continue
            break
        tab.M = 0  # force re-sizing for the pair table
        syms = ok[:n].cpu().numpy()
        firsts = _u32(om[:n])
        order = np.argsort(firsts, kind="stable")
        syms, firsts = syms[order], firsts[order]
        if self.comm is not None:        # global first-seen order = (rank, local first position) (bpe.py:243-247)
            seen, gs = set(), []
            for part in self.comm.all_gather_object(syms.tolist()):
                for x in part:
                    if x not in seen:
                        seen.add(x); gs.append(x)
            syms = np.array(gs, dtype=np.int64)
            firsts = np.zeros(len(gs), dtype=np.int64)
        # ids in first-seen order; per-residue id by binary search over the sorted distinct symbols
        sort2 = np.argsort(syms)
        sorted_syms = torch.from_numpy(syms[sort2]).to(_dev())
        ids_of_sorted = torch.from_numpy(sort2.astype(np.int32)).to(_dev())
        idx = torch.searchsorted(sorted_syms, self.sym)
        tid = ids_of_sorted[idx]
        for k, (s, r) in enumerate(zip(syms.tolist(), firsts.tolist())):
            v = s - 1
            tb, cb, pb = v & 0x3FF, (v >> 10) & 0x3FF, (v >> 20) & 0x3FF
            d = {"N:CA": [BOND_LENGTHS[0]], "CA:C": [BOND_LENGTHS[1]], "tau": [self.centre("tau", tb)]}
            if cb != NOBIN:
                d.update({"0C:1N": [BOND_LENGTHS[2]], "CA:C:1N": [self.centre("CA:C:1N", cb)],
                          "psi": [self.centre("psi", pb)]})
            self.tokens[k] = json.loads(json.dumps(d, sort_keys=True))
        ts = self.ts
        ts.T = R
        ts.a["hsh"][:R] = self.sym
        ts.a["glue"][:R] = self.glue_res
        ts.a["pos"][:R] = torch.arange(R, device=_dev(), dtype=torch.int32)
        ts.a["len"][:R] = 1
        ts.a["tid"][:R] = tid
        self.init_tid = tid.clone()
        maxlen = int(self.bb.lens_host.max()) if self.bb.S else 1
        if self.comm is not None:
            maxlen = int(self.comm.all_reduce(torch.tensor([maxlen], device=_dev()), "max").item())
        self._maxlen = maxlen
        self._make_powtab()

    def _make_powtab(self):
        pw = [1]
        for _ in range(2 * self._maxlen + 2):
            pw.append(pw[-1] * self.hash_base & M64)
        self.powtab = torch.from_numpy(np.array(pw, dtype=np.uint64).view(np.int64)).to(_dev())

    def _h_key_dicts(self, r0s, nress, glob=False):
        """compute_geo_key's dict for fully histogram-quantised windows (bpe.py:1225-1298); one batched gather
        (glob: r0s index the residues of ALL ranks -- the sharded loop logs example windows of whichever rank held one)"""
        idx = np.concatenate([np.arange(r, r + n) for r, n in zip(r0s, nress)])
        src = self.hl.rbins_g if glob else self.rbins.view(9, self.bb.R)
        rb_all = src[:, torch.from_numpy(idx).to(_dev())].cpu().numpy().astype(np.int64) & 0xFFFF
        out, o = [], 0
        for nres in nress:
            rb = rb_all[:, o:o + nres]
            o += nres
            ends = rb[SLOT["CA:C:1N"], nres - 1] == NOBIN
            m = nres - 1 if ends else nres
            d = {"N:CA": [0] * nres, "CA:C": [0] * nres, "0C:1N": [0] * m,
                 "tau": rb[SLOT["tau"]].tolist(), "CA:C:1N": rb[SLOT["CA:C:1N"], :m].tolist(),
                 "psi": rb[SLOT["psi"], :m].tolist()}
            if nres > 1:
                d["omega"] = rb[SLOT["omega"], :nres - 1].tolist()
                d["C:1N:1CA"] = rb[SLOT["C:1N:1CA"], :nres - 1].tolist()
                d["phi"] = rb[SLOT["phi"], :nres - 1].tolist()
            out.append(d)
        return out

    def _count_pairs(self):
        ts = self.ts
        if self.mode == "H":
            def launch(tab):
                _lib.call("gb_pair_count_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab),
                          self.hash_base, *tab.ptrs(), _lib.ptr(tab.overflow))
        else:
            def launch(tab):
                _lib.call("gb_pair_count_m", ts.p("cls"), ts.p("glue"), ts.T, *tab.ptrs(), _lib.ptr(tab.overflow))
        if self.mode == "H":
            def first_pos(tab):
                _lib.call("gb_first_positions_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab),
                          self.hash_base, *tab.ptrs(), _lib.ptr(tab.out4), _lib.ptr(tab.tie_keys), _lib.ptr(tab.tie_min),
                          tab.TIE_CAP)
        else:
            first_pos = None                    # medoid keys decode to (class, glue, class): no example window needed
        if self.k1_events is not None:          # bench.py: CUDA events around the K1 launch on its own stream
            inner = launch

            def launch(tab):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                inner(tab)
                e1.record()
                self.k1_events.append((e0, e1, ts.T))
        if self.mode == "H" and self.use_sketch:
            if self.sketch is None:
                self.sketch = SlotSketch(ts.T)
            cs = 1 if 16 * ts.T > SlotSketch.STREAM_BYTES else 0

            def pass1(slots_ptr, log2ns, variant, ctas):
                ev = self.k1_events
                if ev is not None:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                _lib.call("gb_slot_count_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab), self.hash_base,
                          slots_ptr, log2ns, None, variant | (8 * cs), ctas)
                if ev is not None:
                    e1.record()
                    ev.append((e0, e1, ts.T))

            def pass2(log2ns, cand_ptr, ncand_ptr, cap, tab):
                ev = self.k1b_events
                if ev is not None:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                _lib.call("gb_cand_count_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab), self.hash_base,
                          log2ns, cand_ptr, ncand_ptr, cap, *tab.ptrs(), _lib.ptr(tab.overflow), cs)
                if ev is not None:
                    e1.record()
                    ev.append((e0, e1, ts.T))
            r = self.sketch.count(pass1, pass2, ts.p("pos"), ts.p("len"))
            if r is not None:
                return r
        if self.mode == "H":
            return self.table.count(launch, ts.T, first_pos, ts.p("pos"), ts.p("len"))
        return self.table.count(launch, ts.T, first_pos)

    def _k1_launch(self, tab):
        ts = self.ts
        e = self.k1_events
        if e is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        _lib.call("gb_pair_count_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab),
                  self.hash_base, *tab.ptrs(), _lib.ptr(tab.overflow))
        if e is not None:
            e1.record()
            e.append((e0, e1, ts.T))

    def step_h(self):
        """one merge iteration, histogram mode (bpe.py:1792-2166 with rmsd_key None)"""
        if self.async_steps:
            return self.enqueue(1)
        if self.comm is not None:
            return self._step_h_dist()
        return self._step_h_host()

    def _step_h_host(self):
        """host-driven iteration: K1 -> one D2H -> tie-break on the json strings -> K2 (also the fallback of the
        device-driven loop for iterations it halts on)"""
        ts = self.ts
        mx, ties = self._count_pairs()
        if mx == 0:
            raise IndexError("no adjacent pairs left (the reference's peekitem(0) raises here too)")
        if ties[0][2] is None:                      # long tie list: gather the example windows' extents
            mins = torch.tensor([t[1] for t in ties], device=_dev(), dtype=torch.int64)
            p0 = _u32(ts.a["pos"][mins]).tolist()
            nress = (_u32(ts.a["len"][mins]) + _u32(ts.a["len"][mins + 1])).tolist()
        else:
            p0, nress = [t[2] for t in ties], [t[3] for t in ties]
        dicts = self._h_key_dicts(p0, nress)
        cands = [(json.dumps(d, sort_keys=True), t[0], r0, nr) for t, d, r0, nr in zip(ties, dicts, p0, nress)]
        key_str, key, r0, nres = min(cands)            # SortedDict order: (-count, key string)
        n = len(self.tokens)
        self._check_not_popped(key_str)
        self.tokens[n] = json.loads(key_str)
        self.pop_log.append((mx, key_str))
        self._ntok, self._last_count = n + 1, mx
        _lib.call("gb_flag_pairs_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab), self.hash_base,
                  key, _lib.ptr(ts.flags))
        if self.verify:                                # checked after the merge's single read-back (u32[3], u32[4])
            _lib.call("gb_verify_flagged_h", _lib.ptr(ts.flags), ts.T, ts.p("pos"), ts.p("len"), _lib.ptr(self.rbins),
                      self.bb.R, r0, nres, ts.u32.data_ptr() + 12)
        r = self._merge(key, n, None, None, bound=mx)
        if self.verify:
            if int(r[3]):
                raise RuntimeError("64-bit content-hash collision detected (window bins differ); re-run with another hash base")
            if int(r[4]) != mx:
                raise RuntimeError(f"flagged {int(r[4])} occurrences, histogram counted {mx}")
        self.n_steps += 1

    # ---- device-driven loop (csrc/k_hloop.cu): iterations are enqueued, the host reads the step log in bulk -------
    HL_LOG_CAP = 1 << 16
    HL_TB_CAP = 128
    HS = dict(T=0, STEP=1, HALT=2, EVN=3, MPREV=4, THR=5, C1=6, NCAND=7, NZ=8, M1=10, NT=11, ND=12, ROUND2=13,
              WIN_R0=14, WIN_NRES=15, WIN_CNT=16, WIN_J=17, MISMATCH=18, FLAGGED=19, NEWT=20, NTOK=21, OVF=22,
              R2COUNT=23, MAXSTEPS=24, EVCAP=25, HALT_STEP=26, NTMAX=27, N=28, NCL=29, CLCAP=30, CMPT=31, DIST=32)
    HALT_NAMES = {1: "ties", 2: "candidates", 3: "collision", 4: "table", 5: "event log", 6: "empty", 7: "step log",
                  8: "count", 9: "peer timeout"}

    def _hl_init(self):
        ts = self._ts
        comm = self.comm
        maxwin = int(self.bb.lens_host.max()) if self.bb.S else 2
        Tb = ts.T
        if comm is not None:      # buffer sizes that take part in collectives must agree on every rank
            mx = comm.all_reduce(torch.tensor([maxwin, Tb], device=_dev(), dtype=torch.int64), "max").cpu().numpy()
            maxwin, Tb = int(mx[0]), int(mx[1])
        if self.sketch is None:
            l2 = None
            if comm is not None and "GEOBPE_LOG2NS" not in os.environ:
                # sharded: the slot counters cross NVLink every iteration -- a quarter of the one-GPU size (NS >= local pairs / 8);
                # only slots holding a heavy key matter (candidates = slots >= 0.9 x previous max), see DESIGN 7
                l2 = max(16, min(24, (max(Tb, 2) - 1).bit_length() - 3))
            self.sketch = SlotSketch(Tb, log2ns=l2)
            if comm is not None:     # the exact table also takes the other ranks' keys of the candidate slots
                self.sketch.small.ensure(4096 * comm.world)
        sk = self.sketch
        hl = types.SimpleNamespace()
        hl.hs = _zeros(64, torch.int32)
        hl.hs64 = _zeros(2, torch.int64)
        hl.tie_keys = _empty(256, torch.int64)
        hl.tie_min = _empty(256, torch.int32)
        hl.stride = ((39 * maxwin + 200) + 127) // 128 * 128
        if comm is not None:
            hl.stride = min(hl.stride, 4096)      # the tie strings travel (allreduce-max): longer windows halt to the host path
        hl.strbuf = _zeros(self.HL_TB_CAP * hl.stride, torch.uint8)
        cap = self.HL_LOG_CAP if comm is None else min(self.HL_LOG_CAP, 2048)
        hl.log_cap = cap
        hl.log_key = _empty(cap, torch.int64)
        hl.log = {k: _empty(cap, torch.int32) for k in ("cnt", "r0", "nres", "evstart", "nk")}
        hl.dist = None
        if comm is not None:
            lib = _lib.load()
            xe = int(lib.gb_h_loop_xbuf_elems())
            d = types.SimpleNamespace()
            d.x = _zeros(xe, torch.int64)
            d.xall = _zeros(xe * comm.world, torch.int64)
            d.tie_skeys = _zeros(self.HL_TB_CAP, torch.int64)
            d.tie_smin = _zeros(self.HL_TB_CAP, torch.int32)
            d.sbuf = _zeros(int(lib.gb_h_loop_sbuf_bytes(hl.stride)), torch.uint8)
            d.log_str = _zeros(cap * hl.stride, torch.uint8)
            d.log_strlen = _zeros(cap, torch.int32)
            d.graphs = {}
            d.peer = self._peer_setup(sk.log2ns, xe, d.sbuf.numel()) if os.environ.get("GEOBPE_PEER", "1") != "0" else None
            hl.dist = d
        hl.nargs = _lib.load().gb_h_loop_nargs()
        hl.enq = 0                    # iterations enqueued since the state block was last written
        hl.parity = 0                 # which token buffer is current on the device (0 = ts.a)
        hl.Tmax = ts.T
        hl.rmark = _zeros(ts.cap, torch.uint8)
        hl.cl_key = _empty(ts.cap, torch.int64)
        hl.cl_j = _empty(ts.cap, torch.int32)
        hl.cl_r = _empty(ts.cap, torch.int32)
        hl.period = 64
        hl.lazy = 0
        self._raw = []
        self._ntok = len(self.tokens)
        self._last_count = self.pop_log[-1][0] if self.pop_log else 0
        hl.halts = {}
        hl.r2 = 0
        hl.steps = 0
        self.hl = hl

    def _peer_setup(self, log2ns, xe, sbytes):
        """symmetric block [slots | x | sbuf | flags] on every rank, mapped into every other rank through CUDA IPC, for the
        peer-memory transport of the sharded loop (csrc/k_hloop_dist.cuh); None (on every rank) if any rank cannot set it
        up -- the loop then keeps the three NCCL collectives per iteration"""
        import socket
        lib, comm = _lib.load(), self.comm
        maxw = int(lib.gb_peer_max_world())
        al = lambda n: (n + 255) // 256 * 256
        o_x = al(4 << log2ns)
        o_s = o_x + al(8 * xe)
        o_f = o_s + al(sbytes)
        total = o_f + al(4 * 3 * maxw)
        pr = types.SimpleNamespace(base=None, opened=[], lib=lib)
        handle, err = b"", None
        if comm.world > maxw:
            err = "world too large"
        else:
            base, h = C.c_void_p(), (C.c_ubyte * 64)()
            if lib.gb_peer_alloc(total, C.byref(base), h) != 0:
                err = lib.gb_last_error().decode()
            else:
                pr.base, handle = base.value, bytes(h)
        infos = comm.all_gather_object((socket.gethostname(), handle, err))
        ok = all(e is None for _, _, e in infos) and len({hn for hn, _, _ in infos}) == 1
        bases = []
        if ok:
            for r, (_, hb, _) in enumerate(infos):
                if r == comm.rank:
                    bases.append(pr.base)
                    continue
                q = C.c_void_p()
                if lib.gb_peer_open((C.c_ubyte * 64).from_buffer_copy(hb), C.byref(q)) != 0:
                    err = lib.gb_last_error().decode()
                    break
                pr.opened.append(q.value)
                bases.append(q.value)
        ok = all(comm.all_gather_object(ok and err is None))
        if not ok:
            self._peer_release(pr)
            return None
        pr.table = (C.c_uint64 * (4 * comm.world))(*([b for b in bases] + [b + o_x for b in bases] + [b + o_s for b in bases] +
                                                     [b + o_f for b in bases]))
        pr.slots, pr.x, pr.s = pr.base, pr.base + o_x, pr.base + o_s
        pr.gslots = _zeros(1 << log2ns, torch.int32)
        pr.sred = _zeros(sbytes, torch.uint8)
        pr.ctr = _zeros(4, torch.int32)
        torch.cuda.synchronize()
        comm.all_reduce(torch.zeros(1, device=_dev()))          # every block is zeroed and mapped before the first signal
        torch.cuda.synchronize()
        return pr

    def close(self):
        """sharded runs, before `destroy_process_group()`: drop the CUDA graphs that captured NCCL collectives (the process
        group cannot be destroyed while they exist -- it blocks) and release the symmetric block of the peer-memory
        transport (collective: every rank calls it; a barrier makes sure no peer kernel still reads this rank's block).
        The learner stays usable: it falls back to the NCCL exchanges afterwards."""
        self.sync()
        if self.hl is not None and getattr(self.hl, "v2", False):
            if self.hl.peer is not None:
                torch.cuda.synchronize()
                self.comm.all_reduce(torch.zeros(1, device=_dev()))
                torch.cuda.synchronize()
                self._materialise_log()      # (example windows of the logged iterations index the replicated rbins)
                self._peer_release(self.hl.peer)
                self.hl.peer = None
                stats = (self.hl.steps, self.hl.halts, self.hl.stats, self.hl.times_ns)
                self.hl = None               # (a later enqueue sets the loop up again, collectively)
                self._closed_stats = stats
            return
        d = self.hl.dist if self.hl is not None else None
        if d is None:
            return
        torch.cuda.synchronize()
        d.graphs = {}
        _lib.load().gb_h_loop_incremental(1)         # (also forgets the C-side period graphs captured for this learner's buffers)
        if d.peer is None:
            return
        torch.cuda.synchronize()
        self.comm.all_reduce(torch.zeros(1, device=_dev()))
        torch.cuda.synchronize()
        self._peer_release(d.peer)
        d.peer = None
        d.graphs = {}
        self.hl.args = self._hl_args()
        self.hl._args_key = tuple(self.hl.args)

    @staticmethod
    def _peer_release(pr):
        for q in pr.opened:
            pr.lib.gb_peer_close(C.c_void_p(q))
        pr.opened = []
        if pr.base:
            pr.lib.gb_peer_free(C.c_void_p(pr.base))
            pr.base = None

    def _hl_args(self):
        ts, sk, hl = self._ts, self.sketch, self.hl
        # the two token buffers keep their argument positions for the life of the loop (the captured graphs are keyed by the
        # arguments); which of them is current is the `parity` passed with every call
        if getattr(hl, "buf0", None) is None or not ((ts.a is hl.buf0 and ts.b is hl.buf1) or (ts.a is hl.buf1 and ts.b is hl.buf0)):
            hl.buf0, hl.buf1 = ts.a, ts.b
        a, b = hl.buf0, hl.buf1
        P = _lib.ptr
        vals = [P(hl.hs), P(hl.hs64), P(a["hsh"]), P(b["hsh"]), P(a["glue"]), P(b["glue"]), P(a["pos"]), P(b["pos"]),
                P(a["len"]), P(b["len"]), P(a["tid"]), P(b["tid"]), P(ts.flags), P(hl.rmark), P(ts.scratch), P(self.powtab), self.hash_base,
                P(self.rbins), self.bb.R, P(sk.slots), sk.log2ns, P(sk.bmax), P(sk.cand), sk.CAP, sk.variant, sk.ctas,
                P(sk.small.keys), P(sk.small.cnt), P(sk.small.min), sk.small.M, P(hl.tie_keys), P(hl.tie_min),
                P(hl.strbuf), hl.stride, P(self._evl), P(self._evr), P(self._evp), P(hl.log_key),
                P(hl.log["cnt"]), P(hl.log["r0"]), P(hl.log["nres"]), P(hl.log["evstart"]), P(hl.log["nk"]),
                P(hl.cl_key), P(hl.cl_j), P(hl.cl_r), hl.Tmax, 1 if 16 * hl.Tmax > SlotSketch.STREAM_BYTES else 0, hl.period]
        d = hl.dist
        vals += ([P(d.x), P(d.xall), self.comm.world, self.comm.rank, P(d.tie_skeys), P(d.tie_smin), P(d.sbuf), P(d.log_str),
                  P(d.log_strlen)] if d is not None else [0] * 9)
        pr = d.peer if d is not None else None
        if pr is not None:         # peer-memory transport: the loop's slots / x / sbuf are the symmetric block's
            vals[19] = pr.slots
            vals[-9], vals[-3] = pr.x, pr.s
            vals += [C.addressof(pr.table), P(pr.gslots), P(pr.sred), P(pr.ctr)]
        else:
            vals += [0] * 4
        assert len(vals) == hl.nargs, (len(vals), hl.nargs)
        return (C.c_int64 * len(vals))(*[v if v < (1 << 63) else v - (1 << 64) for v in vals])

    def _hl_push_state(self):
        """host state -> device state block (after construction, after a host-side iteration)"""
        hl, H = self.hl, self.HS
        h = np.zeros(64, dtype=np.int64)
        h[H["DIST"]] = 1 if hl.dist is not None else 0
        h[H["T"]] = self._ts.T
        h[H["N"]] = self._ts.T
        h[H["CLCAP"]] = self._ts.cap
        h[H["EVN"]] = self._ev_n
        h[H["MPREV"]] = self._last_count
        h[H["NTOK"]] = self._ntok
        h[H["MAXSTEPS"]] = hl.log_cap
        h[H["EVCAP"]] = self._evl.numel()
        hl.hs.copy_(torch.from_numpy(h.astype(np.int32)))
        hl.enq = 0
        _lib.call("gb_table_clear", *self.sketch.small.ptrs())     # hl_decide leaves it cleared from then on
        self._ts.flags.zero_()      # (hl.Tmax stays the initial upper bound: same arguments -> the captured graphs are reused)
        hl.args = self._hl_args()
        hl.parity = 0 if self._ts.a is hl.buf0 else 1
        if hl.dist is not None and getattr(hl, "_args_key", None) != tuple(hl.args):
            hl.dist.graphs = {}
        hl._args_key = tuple(hl.args)

    def enqueue(self, nsteps):
        """enqueue `nsteps` merge iterations on the device; nothing is read back until sync()"""
        if self.loop_version == 2:
            return self._h2_enqueue(nsteps)
        if self.hl is None:
            self._hl_init()
            self._hl_push_state()
        hl = self.hl
        if hl.dist is not None and hl.dist.peer is None:
            return self._hl_enqueue_dist(nsteps)
        hl.lazy += nsteps                   # launched in whole compaction periods (CUDA-graph replays), the rest at sync()
        self._pending += nsteps
        self.n_steps += nsteps
        if hl.lazy >= 4 * hl.period:
            self._hl_flush(hl.lazy // hl.period * hl.period)

    def _hl_enqueue_dist(self, nsteps):
        self._pending += nsteps
        self.n_steps += nsteps
        self._hl_run_dist(nsteps)

    def _hl_iter_dist(self):
        """sharded run, one iteration: five kernel groups with a collective after each of the first four (see
        csrc/k_hloop_dist.cuh); everything is enqueued on the current stream, nothing is read back"""
        hl, d, comm, sk = self.hl, self.hl.dist, self.comm, self.sketch
        part = lambda k, n: _lib.call_n("gb_h_loop_part", n, hl.args, hl.nargs, k, hl.parity)
        ev = self.k1_events
        if ev is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        part(0, 1)
        if ev is not None:
            e1.record()
            ev.append((e0, e1, None))          # (the live-token count of that iteration is filled in by bench.py)
        if d.peer is not None:        # exchanges happen inside the parts' own kernels (loads from the peers' blocks)
            part(1, 5)
            part(2, 5)
            part(3, 6)
            return
        comm.all_reduce(sk.slots, "sum")
        part(1, 3)
        comm.all_gather_into(d.xall, d.x)
        part(2, 3)
        comm.all_reduce(d.sbuf, "max")
        part(3, 5)

    def prepare(self):
        """warm-up: capture the CUDA graphs of the device-driven loop (both buffer parities) before they are needed"""
        if not self.async_steps:
            return
        self.sync()
        if self.loop_version == 2:
            return                                          # one cooperative kernel per enqueue: nothing to capture
        if self.hl is None:
            self._hl_init()
            self._hl_push_state()
        hl = self.hl
        if hl.dist is None or hl.dist.peer is not None:     # kernels only: whole periods captured on the C side
            _lib.call_n("gb_h_loop_prepare", 0, hl.args, hl.nargs)
            return
        p0 = hl.parity
        for p in (0, 1):
            if p not in hl.dist.graphs:
                hl.parity = p
                self._hl_capture_dist()
        hl.parity = p0

    def _hl_capture_dist(self):
        """one iteration (kernels + the three NCCL collectives) as a CUDA graph for the current buffer parity"""
        hl = self.hl
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._hl_iter_dist()                 # capturing does not execute
        hl.dist.graphs[hl.parity] = g
        return g

    def _hl_run_dist(self, n):
        """n iterations, each one graph replay (unless bench.py is timing K1 with events); the arrays are compacted
        every `period` iterations"""
        hl, d = self.hl, self.hl.dist
        for _ in range(n):
            if self.use_graphs and self.k1_events is None:
                g = d.graphs.get(hl.parity) or self._hl_capture_dist()
                g.replay()
            else:
                self._hl_iter_dist()
            hl.enq += 1
            if hl.enq % hl.period == 0:
                self._hl_call(0, 1)

    def _hl_flush(self, n):
        hl = self.hl
        if n > 0:
            if hl.dist is not None and hl.dist.peer is None:
                self._hl_run_dist(n)
            else:
                self._hl_call(n, 0)
                hl.enq += n
            hl.lazy -= n

    def _hl_call(self, nsteps, compact_after):
        hl = self.hl
        out = C.c_int(0)
        per = 17 if (hl.dist is not None and hl.dist.peer is not None) else _lib.load().gb_h_loop_kernels_per_step()
        _lib.call_n("gb_h_loop", nsteps * per + 5 * (nsteps // hl.period + compact_after), hl.args, hl.nargs, nsteps, hl.parity,
                    hl.enq, compact_after, C.byref(out))
        hl.parity = out.value

    def sync(self):
        """drain the device loop: read the state block and the step log, materialise vocabulary entries / pop log;
        an iteration the device could not resolve runs once on the host path, the rest is re-enqueued"""
        if self.hl is not None and getattr(self.hl, "v2", False):
            return self._h2_sync()
        while self._pending:
            hl, H = self.hl, self.HS
            self._hl_flush(hl.lazy)
            self._hl_call(0, 1)                                              # physical compaction: `len` is plain again
            h = hl.hs.cpu().numpy().astype(np.int64) & 0xFFFFFFFF            # the one blocking read
            done, halt = int(h[H["STEP"]]), int(h[H["HALT"]])
            want = self._pending
            self._pending = 0
            self.n_steps -= want
            if done:
                keys = hl.log_key[:done].cpu().numpy().astype(np.uint64).tolist()
                lg = {k: _u32(v[:done]) for k, v in hl.log.items()}
                if hl.dist is not None:          # sharded run: the winner's json string itself is logged (the example
                    sl = _u32(hl.dist.log_strlen[:done])                      # window may live on another rank)
                    raw = hl.dist.log_str[:done * hl.stride].cpu().numpy().reshape(done, hl.stride)
                    self._raw.append((lg["cnt"], [raw[s, :sl[s]].tobytes().decode() for s in range(done)], None))
                else:
                    self._raw.append((lg["cnt"], lg["r0"], lg["nres"]))      # formatted lazily (_materialise_log)
                for s in range(done):
                    self._step_events.append((self._ntok + s, int(lg["evstart"][s]), int(lg["nk"][s]), keys[s]))
                self._ntok += done
                self._last_count = int(lg["cnt"][-1])
                # the device appends an iteration's merge events in arbitrary order: sort each iteration's range by position
                e0, e1 = self._ev_n, int(h[H["EVN"]])
                if e1 > e0:
                    self._unsorted = getattr(self, "_unsorted", []) + [(e0, e1, hl.log["nk"][:done].clone())]
                self._ts.T = int(h[H["T"]])
                self._ev_n = e1
                self.n_steps += done
            if self._ts.a is not (hl.buf1 if hl.parity else hl.buf0):      # ts.a = the buffer the loop left current
                self._ts.swap()
            if int(h[H["N"]]) != int(h[H["T"]]):
                raise RuntimeError(f"compaction left {int(h[H['N']])} entries for {int(h[H['T']])} live tokens")
            hl.steps += done
            hl.r2 += int(h[H["R2COUNT"]])
            hl.ntmax = max(getattr(hl, "ntmax", 0), int(h[H["NTMAX"]]))
            rest = want - done
            if halt in (3, 5, 8, 9):
                raise RuntimeError({9: "a peer rank did not signal within 20 s (peer-memory transport of the sharded loop)",
                                    3: "64-bit content-hash collision detected (window bins differ); re-run with another hash base",
                                    5: "merge-event log overflow",
                                    8: "flagged occurrences differ from the histogram count"}[halt])
            if halt == 6:
                raise IndexError("no adjacent pairs left (the reference's peekitem(0) raises here too)")
            if halt:
                hl.halts[halt] = hl.halts.get(halt, 0) + 1
                if halt != 7:                      # ties / candidates / table: this one iteration on the host path
                    if hl.dist is not None:
                        self._step_h_dist()        # (every rank halts at the same iteration)
                    else:
                        self._step_h_host()
                    rest -= 1
            self._hl_push_state()
            if rest > 0:
                self.enqueue(rest)

    # ---- device-driven loop, second design (csrc/k_hloop2.cu): one persistent cooperative kernel per enqueue --------
    H2 = dict(T=0, STEP=1, HALT=2, EVN=3, N=4, NTOK=5, LOGCAP=6, EVCAP=7, TARGET=8, PAR=9, PERIOD_OK=10, THR=11, THR_REQ=12,
              C1=13, NCAND=14, NCL=15, CLCAP=16, NOCC=17, M1=18, NT=19, P1=20, WIN_SLOT=21, WIN_R0=22, WIN_NRES=23, WIN_CNT=24,
              MISMATCH=25, NEWT=26, OVF=27, ERR=28, HALT_STEP=29, NTMAX=30, CMPT=31, FRESH=32, ND=33, NZ=34, MPREV=35,
              ST_PERIODS=36, ST_CROSS=37, ST_INCONCL=38, ST_COMPACT=39, ST_R2=40, ST_FULL=41, ST_LOOPS=42, PERIOD_LEN=43,
              ST_MAXNCL=44, ST_MAXND=45, ST_MAXNCAND=46, ST_TIES=47,
              HALT_REQ=48, PERIOD_REQ=49, NDEATH=50, NBIRTH=51, DSEQ=52, XSEQ=53, ST_WAIT_US=54, TRACE=55, WAIT_STAGE=56, WAIT_SEQ=57, WAIT_SEEN0=58, WAIT_SEEN1=59, ST_RETRY=60, XCROSS=61)
    H2_TIMES = ("period_start", "decide", "flag", "deaths", "apply", "births", "compaction", "candidate_pass",
                "flag_block0_work", "deaths_block0_work", "apply_block0_work", "births_block0_work", "exchange")
    H2_HALT_NAMES = {**HALT_NAMES, 10: "internal", 11: "guard"}
    H2_TABLE_LOG2 = 15

    def _h2_init(self):
        ts = self._ts
        lib = _lib.load()
        maxwin = int(self.bb.lens_host.max()) if self.bb.S else 2
        if self.sketch is None and self.comm is None:
            self.sketch = SlotSketch(ts.T)
        hl = types.SimpleNamespace(v2=True, dist=None, peer=None, rbins_g=None, rbase=0, R_total=self.bb.R, dseq=0, xseq=0)
        assert lib.gb_h_loop2_state_words() == 64
        comm = self.comm
        if comm is not None:
            # the rows of every rank's rbins, so that any rank can order / check any key (example windows are global ranges)
            R_all = comm.all_gather_object(self.bb.R)
            hl.rbase, hl.R_total = int(sum(R_all[:comm.rank])), int(sum(R_all))
            if hl.R_total >= (1 << 32) - 2:
                raise ValueError("a sharded run holds at most 2^32-2 residues in total (u32 example windows)")
            Rmax = max(R_all)
            self.hl_R_max = Rmax                 # buffer sizes that define the symmetric block's layout must agree on every rank
            pad = _zeros(9 * Rmax, torch.int16).view(9, Rmax)
            pad[:, :self.bb.R] = self.rbins.view(9, self.bb.R)
            allr = _empty(comm.world * 9 * Rmax, torch.int16)
            comm.all_gather_into(allr.view(torch.uint8), pad.view(-1).view(torch.uint8))      # (NCCL has no 16-bit integer type)
            allr = allr.view(comm.world, 9, Rmax)
            g = _empty(9 * hl.R_total, torch.int16).view(9, hl.R_total)
            o = 0
            for r, Rr in enumerate(R_all):
                g[:, o:o + Rr] = allr[r, :, :Rr]
                o += Rr
            hl.rbins_g = g.contiguous()
            del allr, pad
            if self.sketch is None:      # the slot counters are replicated: sized for the pairs of ALL ranks
                Tg = int(comm.all_reduce(torch.tensor([ts.T], device=_dev(), dtype=torch.int64), "sum").item())
                self.sketch = SlotSketch(Tg)
            mw = comm.all_reduce(torch.tensor([maxwin], device=_dev(), dtype=torch.int64), "max")
            maxwin = int(mw.item())
            _dbg(f"h2_init: rbins replicated, R_total {hl.R_total}, rbase {hl.rbase}, slots 2^{self.sketch.log2ns}")
        sk = self.sketch
        hl.hs = _zeros(64, torch.int32)
        hl.hs64 = _zeros(32, torch.int64)
        hl.stride = ((39 * maxwin + 200) + 127) // 128 * 128
        hl.strbuf = _zeros(self.HL_TB_CAP * hl.stride, torch.uint8)
        hl.log_cap = self.HL_LOG_CAP
        hl.log_key = _empty(hl.log_cap, torch.int64)
        hl.log = {k: _empty(hl.log_cap, torch.int32) for k in ("cnt", "r0", "nres", "evstart", "nk")}
        hl.cbit = _zeros((1 << sk.log2ns) // 32, torch.int32)
        hl.rank = torch.from_numpy(tie_rank_table().view(np.int16).copy()).to(_dev())
        hl.M = 1 << int(os.environ.get("GEOBPE_H2_TABLE_LOG2", self.H2_TABLE_LOG2))
        hl.rmark = _zeros(ts.cap, torch.uint8)
        hl.cl_key = _empty(ts.cap, torch.int64)
        hl.cl_j = _empty(ts.cap, torch.int32)
        hl.occ_j = _empty(ts.cap, torch.int32)
        hl.occ_r = _empty(ts.cap, torch.int32)
        hl.nargs = lib.gb_h_loop2_nargs()
        hl.dcap = next_pow2(max(1 << 16, (self.hl_R_max if hasattr(self, 'hl_R_max') else self.bb.R) // 64))
        hl.xcap = 1 << 17
        hl.thr_num = int(os.environ.get("GEOBPE_H2_THR16", 14))
        hl.ctas = int(os.environ.get("GEOBPE_H2_CTAS", 1))     # one CTA per SM: cheaper grid barriers (measured 45 vs 56 us / iteration)
        hl.period = 64                 # (reported by bench.py; periods end on the device when the threshold is crossed)
        hl.lazy = 0
        hl.slots_valid = False
        hl.buf0 = hl.buf1 = None
        hl.halts, hl.r2, hl.steps, hl.ntmax = {}, 0, 0, 0
        hl.stats = dict(periods=0, crossings=0, inconclusive=0, compactions=0, second_rounds=0, launches=0)
        hl.times_ns = {k: 0 for k in self.H2_TIMES}
        self._raw = []
        self._ntok = len(self._tokens)
        self._last_count = self._pop_log[-1][0] if self._pop_log else 0
        self.hl = hl
        self._h2_table(hl.M)
        if comm is not None:
            self._h2_peer_setup()
            _dbg("h2_init: peer blocks mapped")

    def _h2_table(self, M):
        hl = self.hl
        hl.M = M
        hl.tk = _empty(M, torch.int64)
        hl.tc = _zeros(M, torch.int32)
        hl.tr0 = _zeros(M, torch.int32)
        hl.tnres = _zeros(M, torch.int32)
        hl.tlist = _zeros(M, torch.int32)

    def _h2_peer_setup(self):
        """one symmetric block [delta x 2 | export x 2 | flags] per rank, mapped into every other rank through CUDA IPC"""
        import socket
        lib, comm, hl = _lib.load(), self.comm, self.hl
        if hl.peer is not None:
            torch.cuda.synchronize()
            comm.all_reduce(torch.zeros(1, device=_dev()))
            torch.cuda.synchronize()
            self._peer_release(hl.peer)
            hl.peer = None
        if comm.world > int(lib.gb_h_loop2_max_world()):
            raise NotImplementedError(f"the sharded loop is built for up to {lib.gb_h_loop2_max_world()} ranks")
        while hl.xcap < hl.M // 2:
            hl.xcap *= 2
        al = lambda n: (int(n) + 255) // 256 * 256
        o_x = al(lib.gb_h_loop2_delta_bytes(hl.dcap, comm.world))
        o_f = o_x + al(lib.gb_h_loop2_export_bytes(hl.xcap, comm.world))
        total = o_f + al(lib.gb_h_loop2_flag_bytes())
        pr = types.SimpleNamespace(base=None, opened=[], lib=lib)
        base, h = C.c_void_p(), (C.c_ubyte * 64)()
        err = None
        if lib.gb_peer_alloc(total, C.byref(base), h) != 0:
            err = lib.gb_last_error().decode()
        else:
            pr.base = base.value
        infos = comm.all_gather_object((socket.gethostname(), bytes(h), err))
        if any(e is not None for _, _, e in infos) or len({hn for hn, _, _ in infos}) != 1:
            raise RuntimeError(f"peer-memory setup failed: {[e for _, _, e in infos if e]} (one node, CUDA IPC)")
        bases = []
        for r, (_, hb, _) in enumerate(infos):
            if r == comm.rank:
                bases.append(pr.base)
                continue
            q = C.c_void_p()
            if lib.gb_peer_open((C.c_ubyte * 64).from_buffer_copy(hb), C.byref(q)) != 0:
                raise RuntimeError("gb_peer_open: " + lib.gb_last_error().decode())
            pr.opened.append(q.value)
            bases.append(q.value)
        pr.table = (C.c_uint64 * (3 * comm.world))(*([b for b in bases] + [b + o_x for b in bases] + [b + o_f for b in bases]))
        hl.peer = pr
        hl.dseq = hl.xseq = 0            # (fresh, zeroed flag rows on every rank)
        torch.cuda.synchronize()
        comm.all_reduce(torch.zeros(1, device=_dev()))          # every block is zeroed and mapped before the first signal
        torch.cuda.synchronize()

    def _h2_args(self):
        ts, sk, hl = self._ts, self.sketch, self.hl
        if hl.buf0 is None or not ((ts.a is hl.buf0 and ts.b is hl.buf1) or (ts.a is hl.buf1 and ts.b is hl.buf0)):
            hl.buf0, hl.buf1 = ts.a, ts.b
        a, b = hl.buf0, hl.buf1
        P = _lib.ptr
        vals = [P(hl.hs), P(hl.hs64), P(a["hsh"]), P(b["hsh"]), P(a["glue"]), P(b["glue"]), P(a["pos"]), P(b["pos"]),
                P(a["len"]), P(b["len"]), P(a["tid"]), P(b["tid"]), P(ts.flags), P(hl.rmark), P(ts.scratch), P(self.powtab), self.hash_base,
                P(hl.rbins_g) if hl.rbins_g is not None else P(self.rbins), hl.R_total, P(sk.slots), sk.log2ns, P(sk.bmax), P(sk.cand), sk.CAP, P(hl.cbit),
                P(hl.tk), P(hl.tc), P(hl.tr0), P(hl.tnres), P(hl.tlist), hl.M, P(hl.cl_key), P(hl.cl_j), P(hl.occ_j), P(hl.occ_r),
                P(hl.strbuf), hl.stride, P(hl.rank), P(self._evl), P(self._evr), P(self._evp), P(hl.log_key),
                P(hl.log["cnt"]), P(hl.log["r0"]), P(hl.log["nres"]), P(hl.log["evstart"]), P(hl.log["nk"]),
                hl.thr_num, 1 if 16 * ts.cap > SlotSketch.STREAM_BYTES else 0, hl.ctas,
                hl.rbase, self.comm.world if self.comm is not None else 1, self.comm.rank if self.comm is not None else 0,
                C.addressof(hl.peer.table) if hl.peer is not None else 0, hl.dcap, hl.xcap,
                int(os.environ.get("GEOBPE_PEER_TIMEOUT_MS", 20000)), int(os.environ.get("GEOBPE_H2_VERIFY_INBOX", 0))]
        assert len(vals) == hl.nargs, (len(vals), hl.nargs)
        return (C.c_int64 * len(vals))(*[v if v < (1 << 63) else v - (1 << 64) for v in vals])

    def _h2_push_state(self):
        """host state -> device state block (first use, after every read-back); the slot counters are recounted when a
        host-side iteration changed the token arrays behind the loop's back"""
        hl, H, ts = self.hl, self.H2, self._ts
        hl.args = self._h2_args()
        h = np.zeros(64, dtype=np.int64)
        h[H["T"]] = h[H["N"]] = ts.T
        h[H["EVN"]] = self._ev_n
        h[H["NTOK"]] = self._ntok
        h[H["LOGCAP"]] = hl.log_cap
        h[H["EVCAP"]] = self._evl.numel()
        h[H["CLCAP"]] = hl.cl_key.numel()
        h[H["PAR"]] = 0 if ts.a is hl.buf0 else 1
        h[H["MPREV"]] = self._last_count
        h[H["DSEQ"]], h[H["XSEQ"]] = hl.dseq, hl.xseq       # sequence numbers of the peer exchanges live as long as the flag rows
        hl.hs.copy_(torch.from_numpy(h.astype(np.int64).astype(np.uint32).view(np.int32)))
        hl.hs64.zero_()
        if not hl.slots_valid:
            ts.flags.zero_()             # (the host path leaves its occurrence flags behind)
            hl.rmark.zero_()
            sk = self.sketch
            cs = 1 if 16 * ts.T > SlotSketch.STREAM_BYTES else 0
            sk.slots.zero_()
            ev = self.k1_events
            if ev is not None:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            _lib.call("gb_slot_count_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab), self.hash_base,
                      _lib.ptr(sk.slots), sk.log2ns, None, sk.variant | (8 * cs), sk.ctas)
            if ev is not None:
                e1.record()
                ev.append((e0, e1, ts.T))
            if self.comm is not None:            # replicated global counters (once per recount; kept exact by the delta exchange)
                _dbg("push_state: local slot count done")
                self.comm.all_reduce(sk.slots, "sum")
                _dbg("push_state: slot counters reduced")
            hl.slots_valid = True

    def _h2_enqueue(self, nsteps):
        if self.hl is None:
            self._h2_init()
            self._h2_push_state()
        hl = self.hl
        hl.lazy += nsteps
        self._pending += nsteps
        self.n_steps += nsteps
        if hl.lazy >= 256:
            self._h2_flush(0)

    def _h2_flush(self, final_compact):
        hl = self.hl
        if hl.lazy > 0 or final_compact:
            _dbg(f"launching the loop kernel: {hl.lazy} iterations")
            _lib.call_n("gb_h_loop2", 2, hl.args, hl.nargs, hl.lazy, final_compact)
            _dbg("loop kernel finished")
            hl.stats["launches"] += 1
            hl.lazy = 0

    def _h2_sync(self):
        """drain the loop: one read of the state block + step log; halts run one iteration on the host path and re-enter"""
        while self._pending:
            hl, H = self.hl, self.H2
            self._h2_flush(1)
            h = hl.hs.cpu().numpy().astype(np.int64) & 0xFFFFFFFF            # the one blocking read
            tm = hl.hs64.cpu().numpy()
            done, halt = int(h[H["STEP"]]), int(h[H["HALT"]])
            hl.dseq, hl.xseq = int(h[H["DSEQ"]]), int(h[H["XSEQ"]])
            want = self._pending
            self._pending = 0
            self.n_steps -= want
            hl.stats["peer_wait_us"] = hl.stats.get("peer_wait_us", 0) + int(h[H["ST_WAIT_US"]])
            for k, name in enumerate(self.H2_TIMES):
                hl.times_ns[name] += int(tm[2 + k])
            for name, word in (("periods", "ST_PERIODS"), ("inconclusive", "ST_INCONCL"), ("compactions", "ST_COMPACT"),
                               ("second_rounds", "ST_R2"), ("crossings", "ST_CROSS"), ("tied_iterations", "ST_TIES")):
                hl.stats[name] = hl.stats.get(name, 0) + int(h[H[word]])
            for name, word in (("max_pair_list", "ST_MAXNCL"), ("max_table_keys", "ST_MAXND"), ("max_candidate_slots", "ST_MAXNCAND")):
                hl.stats[name] = max(hl.stats.get(name, 0), int(h[H[word]]))
            if done:
                keys = hl.log_key[:done].cpu().numpy().astype(np.uint64).tolist()
                lg = {k: _u32(v[:done]) for k, v in hl.log.items()}
                self._raw.append((lg["cnt"], lg["r0"], lg["nres"], hl.rbins_g is not None))      # formatted lazily (_materialise_log)
                for s in range(done):
                    self._step_events.append((self._ntok + s, int(lg["evstart"][s]), int(lg["nk"][s]), keys[s]))
                self._ntok += done
                self._last_count = int(lg["cnt"][-1])
                e0, e1 = self._ev_n, int(h[H["EVN"]])
                if e1 > e0:        # merge events are appended in arbitrary order: sort each iteration's range by position
                    self._unsorted = getattr(self, "_unsorted", []) + [(e0, e1, hl.log["nk"][:done].clone())]
                self._ts.T = int(h[H["T"]])
                self._ev_n = e1
                self.n_steps += done
            if self._ts.a is not (hl.buf1 if int(h[H["PAR"]]) else hl.buf0):      # ts.a = the buffer the loop left current
                self._ts.swap()
            if int(h[H["N"]]) != int(h[H["T"]]):
                raise RuntimeError(f"compaction left {int(h[H['N']])} entries for {int(h[H['T']])} live tokens")
            hl.steps += done
            hl.r2 += int(h[H["ST_R2"]])
            hl.ntmax = max(hl.ntmax, int(h[H["NTMAX"]]))
            rest = want - done
            hl.stats["inbox_rereads"] = hl.stats.get("inbox_rereads", 0) + int(h[H["ST_RETRY"]])
            if halt in (5, 8, 9, 10, 12):
                where = (f" (iteration {int(h[H['STEP']])}, exchange {int(h[H['DSEQ']])}, period exchange {int(h[H['XSEQ']])}, "
                         f"marker {int(h[H['TRACE']]):#x}; waited for stage {int(h[H['WAIT_STAGE']])} seq {int(h[H['WAIT_SEQ']])}, flags seen "
                         f"{int(h[H['WAIT_SEEN0']])} {int(h[H['WAIT_SEEN1']])}, rank {self.comm.rank})") if self.comm is not None else ""
                raise RuntimeError({5: "merge-event log overflow", 8: "flagged occurrences differ from the exact pair count",
                                    9: "a peer rank did not signal in time", 10: "incremental pair counts became inconsistent",
                                    12: "the ranks' replicated pair counts diverged"}[halt] + where)
            if halt == 6:
                raise IndexError("no adjacent pairs left (the reference's peekitem(0) raises here too)")
            if halt:
                hl.halts[halt] = hl.halts.get(halt, 0) + 1
                if halt == 3:          # two windows with different bins share a 64-bit content hash: re-key everything
                    if self.comm is not None:
                        raise RuntimeError("64-bit content-hash collision detected (window bins differ) in a sharded run")
                    self.rehash()
                elif halt == 4:        # the exact table (or the delta buffers) were too small for this period
                    self._h2_table(hl.M * 4)
                    if self.comm is not None:
                        hl.dcap *= 2
                        self._h2_peer_setup()
                elif halt in (1, 2):   # ties / candidates: this one iteration on the host path
                    if self.comm is not None:
                        self._step_h_dist()        # (every rank halts at the same iteration)
                    else:
                        self._step_h_host()
                    hl.slots_valid = False
                    rest -= 1
            self._h2_push_state()
            if rest > 0:
                self._h2_enqueue(rest)

    def _sort_events(self, e0, e1, nk_dev):
        """order the merge events [e0, e1) by (iteration, left position) on the device (nk_dev: events per iteration)"""
        if e1 - e0 < 2:
            return
        el = self._evl[e0:e1].to(torch.int64) & 0xFFFFFFFF
        seg = torch.repeat_interleave(torch.arange(nk_dev.numel(), device=_dev()), nk_dev.to(torch.int64) & 0xFFFFFFFF,
                                      output_size=e1 - e0)
        order = torch.argsort((seg << 32) | el)
        self._evl[e0:e1] = self._evl[e0:e1][order]
        self._evr[e0:e1] = self._evr[e0:e1][order]
        self._evp[e0:e1] = self._evp[e0:e1][order]

    def rehash(self, base=None):
        """give every live token its content hash under another base (two different windows shared a 64-bit hash: the
        popped key's occurrences did not all hold the same bins).  Dense arrays only (call after a read-back)."""
        if self.mode != "H":
            return
        if self._pending:
            self.sync()
        if base is None:
            base = ((self.hash_base * 0x9E3779B97F4A7C15 + 0x7F4A7C15) & M64) | 1
        self.hash_base = base
        self._make_powtab()
        ts = self._ts
        _lib.call("gb_rehash_tokens", _lib.ptr(self.sym), _lib.ptr(self.glue_res), ts.p("pos"), ts.p("len"), ts.T, base, ts.p("hsh"))
        self.table.distinct = 0
        if self.hl is not None and getattr(self.hl, "v2", False):
            self.hl.slots_valid = False
            self._h2_push_state()
        elif self.hl is not None:
            self._hl_push_state()

    NS = 1 << 20      # direct-mapped slot counters exchanged per iteration (4 MB allreduce)

    def _step_h_dist(self):
        """histogram-mode merge iteration on a shard: local K1 -> slot allreduce -> exact top keys -> local K2"""
        from .distributed import global_top
        ts, tab, comm = self.ts, self.table, self.comm
        want = 4 * max(tab.distinct, 1024) + 4096
        tab.ensure(2 * ts.T if tab.M == 0 else max(want, 1 << 12))
        while True:
            _lib.call("gb_table_clear", *tab.ptrs())
            tab.overflow.zero_()
            self._k1_launch(tab)
            _lib.call("gb_table_argmax", *tab.ptrs(), _lib.ptr(tab.out4), _lib.ptr(tab.tie_keys), _lib.ptr(tab.tie_min),
                      tab.TIE_CAP)
            out4 = _u32(tab.out4)
            if int(tab.overflow.item()) or int(out4[2]) * 2 > tab.M:
                tab.ensure(tab.M * 4)
                continue
            tab.distinct = int(out4[2])
            break
        if not hasattr(self, "_slots"):
            self._slots = _zeros(self.NS, torch.int32)
            self._mark = _zeros(self.NS, torch.uint8)
            self._ck = _empty(1 << 14, torch.int64); self._cc = _empty(1 << 14, torch.int32)
        _lib.call("gb_table_to_slots", _lib.ptr(tab.keys), _lib.ptr(tab.cnt), tab.M, _lib.ptr(self._slots), self.NS)

        def cand(sel):
            self._mark.zero_()
            self._mark[torch.from_numpy(sel).to(_dev())] = 1
            _lib.call("gb_table_dump_marked", _lib.ptr(tab.keys), _lib.ptr(tab.cnt), tab.M, _lib.ptr(self._mark), self.NS,
                      _lib.ptr(self._ck), _lib.ptr(self._cc), _lib.ptr(ts.u32), 1 << 14)
            n = int(ts.u32[0].item())
            if n > (1 << 14):
                raise RuntimeError("too many local keys in the top slots")
            return self._ck[:n].cpu().numpy(), _u32(self._cc[:n])
        mx, tied = global_top(comm, self._slots, cand, _dev())
        if mx == 0:
            raise IndexError("no adjacent pairs left")
        keys = sorted(tied)
        # example window of every tied key this rank holds (chunks of 64 = the list path of k_first_pos)
        local = {}
        for c0 in range(0, len(keys), 64):
            ch = keys[c0:c0 + 64]
            tab.tie_keys[:len(ch)] = torch.tensor(ch, dtype=torch.int64, device=_dev())
            tab.tie_min[:len(ch)] = -1
            tab.out4.copy_(torch.tensor([mx, len(ch), 0, 0], dtype=torch.int32))
            _lib.call("gb_first_positions_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab),
                      self.hash_base, *tab.ptrs(), _lib.ptr(tab.out4), _lib.ptr(tab.tie_keys), _lib.ptr(tab.tie_min), tab.TIE_CAP)
            mp = _u32(tab.tie_min[:len(ch)])
            have = [(k, int(m)) for k, m in zip(ch, mp.tolist()) if m != 0xFFFFFFFF]
            if have:
                mins = torch.tensor([m for _, m in have], device=_dev(), dtype=torch.int64)
                p0 = _u32(ts.a["pos"][mins]); nr = (_u32(ts.a["len"][mins]) + _u32(ts.a["len"][mins + 1])).tolist()
                for (k, _), d, r0, n_ in zip(have, self._h_key_dicts(p0.tolist(), nr), p0.tolist(), nr):
                    local[k] = (json.dumps(d, sort_keys=True), r0, n_)
        strings = {}
        for part in comm.all_gather_object({k: v[0] for k, v in local.items()}):
            for k, v in part.items():
                if strings.setdefault(k, v) != v:
                    raise RuntimeError("64-bit content-hash collision across ranks (window bins differ)")
        key_str, key = min((strings[k], k) for k in keys)
        n = len(self.tokens)
        self.tokens[n] = json.loads(key_str)
        self.pop_log.append((mx, key_str))
        self._ntok, self._last_count = n + 1, mx
        _lib.call("gb_flag_pairs_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(self.powtab), self.hash_base,
                  key, _lib.ptr(ts.flags))
        if self.verify and key in local:
            nocc = ts.select(1, ts.idx)
            _lib.call("gb_verify_windows_h", _lib.ptr(ts.idx), nocc, ts.p("pos"), ts.p("len"), _lib.ptr(self.rbins),
                      self.bb.R, local[key][1], local[key][2], _lib.ptr(ts.u32))
            if int(ts.u32[0].item()):
                raise RuntimeError("64-bit content-hash collision detected (window bins differ)")
        self._merge(key, n, None, None)
        self.n_steps += 1

    def _merge(self, key, new_tid, p2cls, p2tid, bound=None):
        """K2: overlap rule, event log, stream compaction.  With `bound` (an upper bound of the kept occurrences,
        histogram mode) nothing is read back until the end: one D2H of ts.u32 = {-, kept, newT, verify...}."""
        ts = self.ts
        _lib.call("gb_resolve_overlaps", _lib.ptr(ts.flags), ts.T)
        aux = _lib.ptr(ts.aux) if p2cls is not None else None
        off = self.ev_n * 4
        u = ts.u32.data_ptr()
        _lib.call("gb_select_flagged", _lib.ptr(ts.flags), ts.T, 2, _lib.ptr(ts.kept), u + 4, _lib.ptr(ts.scratch))
        if bound is None:
            nk = int(ts.u32[1].item())
            q, cnt_dev = nk, None
        else:
            q, cnt_dev = min(bound, ts.T), u + 4
        if self.ev_n + q > self._evl.numel():
            raise RuntimeError("merge-event log overflow")
        _lib.call("gb_write_events", _lib.ptr(ts.kept), q, cnt_dev, ts.p("pos"), aux, self._evl.data_ptr() + off,
                  self._evr.data_ptr() + off, self._evp.data_ptr() + off)
        if self.mode == "M":
            self._oplog(nk, p2cls)
        _lib.call("gb_merge_compact", _lib.ptr(ts.flags), ts.T,
                  ts.p("cls"), ts.p("hsh"), ts.p("glue"), ts.p("pos"), ts.p("len"), ts.p("tid"),
                  ts.p("cls", True), ts.p("hsh", True), ts.p("glue", True), ts.p("pos", True), ts.p("len", True),
                  ts.p("tid", True), aux, _lib.ptr(p2cls) if p2cls is not None else None,
                  _lib.ptr(p2tid) if p2tid is not None else None, key, new_tid & 0xFFFFFFFF,
                  u + 8, _lib.ptr(ts.scratch))
        r = _u32(ts.u32)
        nk, newT = int(r[1]), int(r[2])
        assert newT == ts.T - nk, (newT, ts.T, nk)
        ts.swap()
        ts.T = newT
        self.step_events.append((new_tid, self.ev_n, nk, key))
        self.ev_n += nk
        return r

    # ---- medoid mode ---------------------------------------------------------------------------------
    def num_p(self, length):
        return self.num_partitions[length] if isinstance(self.num_partitions, dict) else self.num_partitions

    def window_geo(self, r0, nbonds, f3=None):
        """Tokenizer.token_geo(3*k, nbonds, orig=True) (tokenizer.py:169-202) for the window starting at global
        residue r0; f3 = current (N:CA, CA:C, tau) of the backbone's first residue (Tokenizer._init_*)."""
        V = self.vals_host
        nres = (nbonds + 1) // 3
        full = nbonds == 3 * nres
        b = self.bb.backbone_of(r0)
        first = r0 == int(self.bb.offsets_host[b])
        g = {k: [] for k in ["N:CA", "CA:C", "0C:1N", "tau", "CA:C:1N", "C:1N:1CA", "psi", "omega", "phi"]}
        for k in range(nres):
            r = r0 + k
            if k == 0 and first:
                nca, cac, tau = f3
            else:
                nca, cac, tau = V[COL["N:CA"], r - 1], V[COL["CA:C"], r - 1], V[COL["tau"], r - 1]
            g["N:CA"].append(float(nca)); g["CA:C"].append(float(cac)); g["tau"].append(float(tau))
            if k < nres - 1 or full:
                g["0C:1N"].append(float(V[COL["0C:1N"], r])); g["CA:C:1N"].append(float(V[COL["CA:C:1N"], r]))
                g["psi"].append(float(V[COL["psi"], r]))
            if k < nres - 1:
                g["omega"].append(float(V[COL["omega"], r])); g["C:1N:1CA"].append(float(V[COL["C:1N:1CA"], r]))
                g["phi"].append(float(V[COL["phi"], r + 1]))
        return {k: v for k, v in g.items() if v}

    def intern_class(self, geo):
        s = json.dumps(geo)
        c = self.class_of.get(s)
        if c is None:
            c = len(self.class_geo)
            if c >= (1 << 18):
                raise RuntimeError("more than 2^18 geometry classes: packed 64-bit pair key exhausted")
            self.class_of[s] = c
            self.class_geo.append(geo)
            self.class_f3.append([geo["N:CA"][0], geo["CA:C"][0], geo["tau"][0]])
            self._f3_dirty = True
        return c

    def f3tab(self):
        if self._f3_dirty:
            self._f3_dev = torch.tensor(self.class_f3, dtype=torch.float64, device=_dev()).contiguous()
            self._f3_dirty = False
        return self._f3_dev

    def coords(self, r0_dev, Q, nres, A, f3cls_dev, bond_centres=None):
        out = _empty(Q * A * 3, torch.float64)
        bb = self.bb
        if bond_centres is not None:       # induce: bond lengths taken quantised (bpe.py:1061,1085)
            _lib.call("gb_window_coords", _lib.ptr(bb.vals), bb.R, bb.R, _lib.ptr(bb.bb), _lib.ptr(bb.offsets),
                      _lib.ptr(r0_dev), Q, nres, A, None, None, self.d_f3[0], self.d_f3[1], self.d_f3[2],
                      _lib.ptr(self.rbins), _lib.ptr(bond_centres), bond_centres.shape[1], _lib.ptr(out))
            return out.view(Q, A * 3)
        _lib.call("gb_window_coords", _lib.ptr(bb.vals), bb.R, bb.R, _lib.ptr(bb.bb), _lib.ptr(bb.offsets),
                  _lib.ptr(r0_dev), Q, nres, A, _lib.ptr(f3cls_dev) if f3cls_dev is not None else None,
                  _lib.ptr(self.f3tab()) if f3cls_dev is not None else None, self.d_f3[0], self.d_f3[1], self.d_f3[2],
                  None, None, 0, _lib.ptr(out))
        return out.view(Q, A * 3)

    def k_medoids(self, coords, k):
        """algo.py:144-211 on the device-computed float32 distance matrix; the update iterations run on the device too
        (csrc/k6_pam.cu) unless a cluster comes out empty, which draws from the RNG stream (algo.py:199) and so runs on the host"""
        N = coords.shape[0]
        k = min(N, k)
        if k == N:
            return list(range(N))
        A = coords.shape[1] // 3
        D = _empty(N * N, torch.float32)
        _lib.call("gb_rmsd_matrix", _lib.ptr(coords.contiguous()), N, A, _lib.ptr(D))
        init = self.rng.choice(np.arange(N), size=k, replace=False)
        lib = _lib.load()
        if N <= lib.gb_pam_max_n() and k <= lib.gb_pam_max_k() and not os.environ.get("GEOBPE_PAM_HOST"):
            mi = torch.from_numpy(np.asarray(init, dtype=np.int32)).to(_dev())
            out = _empty(k + 4, torch.int32)
            _lib.call("gb_kmedoids_pam", _lib.ptr(D), N, k, _lib.ptr(mi), 10, _lib.ptr(out), out.data_ptr() + 4 * k)
            r = _u32(out)
            if int(r[k]) == 0:
                return [int(m) for m in r[:k]]
        return self._pam_host(D.view(N, N).cpu().numpy(), init, k)

    def _pam_host(self, D, medoid_indices, k):
        """the reference's loop in numpy (algo.py:184-211)"""
        N = D.shape[0]
        rng = self.rng
        for _ in range(10):
            assignments = np.argmin(D[:, medoid_indices], axis=1)
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
            if total_shift < 1e-4:
                break
        return [int(m) for m in medoid_indices]

    def assign(self, coords, med_coords):
        Q, K = coords.shape[0], med_coords.shape[0]
        A = coords.shape[1] // 3
        out = _empty(Q, torch.int32)
        _lib.call("gb_kabsch_assign", _lib.ptr(coords.contiguous()), Q, A, _lib.ptr(med_coords.contiguous()), K,
                  _lib.ptr(out), None)
        return out

    def _init_tokens_m(self):
        """bpe.py:263-391: residue tokens = nearest of K medoids (size 3 = inner residues, size 2 = last)."""
        bb = self.bb
        R, S = bb.R, bb.S
        self._m_setup()
        is_last = torch.zeros(R, dtype=torch.bool, device=_dev())
        is_last[(bb.offsets[1:] - 1)] = True
        lists = {3: torch.nonzero(~is_last).flatten().to(torch.int32), 2: torch.nonzero(is_last).flatten().to(torch.int32)}
        comm = self.comm
        order = [3, 2] if self.g_offsets[1] - self.g_offsets[0] > 1 else [2, 3]      # the FIRST backbone of the whole set decides
        cls_res = _zeros(R, torch.int32)
        tid_res = _zeros(R, torch.int32)
        for n, size in enumerate(order):
            occ = lists[size].contiguous()
            N = int(occ.numel())                                   # this shard's windows of the size
            if comm is None:
                NG, base = N, 0
            else:                                                  # the sample is drawn over the windows of all shards, in rank order
                counts = comm.all_gather_object(N)
                NG, base = int(sum(counts)), int(sum(counts[:comm.rank]))
            if NG == 0:
                continue
            A = size + 1
            if NG > self.max_num_strucs:
                active = self.rng.choice(NG, self.max_num_strucs, replace=False)
            else:
                active = np.arange(NG)
            allc = self.coords(occ, N, 1, A, None) if N else torch.zeros((0, A * 3), dtype=torch.float64, device=_dev())
            if comm is None:
                act_dev = torch.from_numpy(np.asarray(active, dtype=np.int64)).to(_dev())
                actc = allc[act_dev].contiguous()
            else:
                own = np.flatnonzero((active >= base) & (active < base + N))
                rows = allc[torch.from_numpy(np.asarray(active[own] - base, dtype=np.int64)).to(_dev())]
                actc = self._gather_rows(rows, own, len(active))
            med = self.k_medoids(actc, self.num_p(size))
            medc = actc[torch.tensor(med, device=_dev(), dtype=torch.int64)].contiguous()
            asg = self.assign(allc, medc) if N else _zeros(0, torch.int32)
            if size == 3:
                k = '{"N:CA": [0], "CA:C": [0], "0C:1N": [0], "tau": [0], "CA:C:1N": [0], "psi": [0]}'
            else:
                k = '{"CA:C": [0], "0C:1N": [0], "CA:C:1N": [0]}'
            self.sphere_dict[k] = []
            occ_h = _u32(occ)
            if comm is None:
                geos = [self.window_geo(int(occ_h[int(active[m])]), size, self.d_f3) for m in med]
            else:
                geos = self._gather_geos({p: self.window_geo(int(occ_h[int(active[m]) - base]), size, self.d_f3)
                                          for p, m in enumerate(med) if base <= int(active[m]) < base + N}, len(med))
            p2cls, p2tid = [], []
            for p, geo in enumerate(geos):
                self.sphere_dict[k].append(geo)
                p2tid.append(len(self.token_keys))
                self.tokens[(n, p)] = geo
                self.token_keys.append((n, p))
                p2cls.append(self.intern_class(geo))
            p2cls_d = torch.tensor(p2cls, dtype=torch.int32, device=_dev())
            p2tid_d = torch.tensor(p2tid, dtype=torch.int32, device=_dev())
            occ_l = occ.to(torch.int64)
            cls_res[occ_l] = p2cls_d[asg.to(torch.int64)]
            tid_res[occ_l] = p2tid_d[asg.to(torch.int64)]
        ts = self.ts
        ts.T = R
        ts.a["cls"][:R] = cls_res
        ts.a["glue"][:R] = self.glue_res
        ts.a["pos"][:R] = torch.arange(R, device=_dev(), dtype=torch.int32)
        ts.a["len"][:R] = 1
        ts.a["tid"][:R] = tid_res
        self.init_tid = tid_res.clone()
        # op log of the occurrence sets: initial segment = every adjacent pair in order (bpe.py:1441-1453)
        _lib.call("gb_pair_keys_m", ts.p("cls"), ts.p("glue"), ts.p("pos"), R, _lib.ptr(self.log_key), _lib.ptr(self.log_pos))
        self.log_n = R
        self.log_bounds = [R]          # end offsets of: initial segment, step 1, step 2, ...
        self.key_reset_at = {}         # packed key -> log offset after which its set is fresh

    def _m_setup(self):
        """medoid-mode tables that do not consume the RNG stream (also the starting point of a resumed run)"""
        R = self.bb.R
        self.class_of, self.class_geo, self.class_f3, self._f3_dirty = {}, [], [], True
        self.sphere_dict = {}
        self.sphere_packed = {}        # packed key -> key string, for keys popped in step()
        self.token_keys = []           # list(_tokens)
        th1 = self.th.thresholds[self.size0]
        if self.std_bonds:
            raise NotImplementedError("medoid mode with std_bonds needs quantised-bond coordinates; free_bonds only")
        self.d_f3 = (self.centre("N:CA", get_ind(INIT_N_CA, th1["N:CA"])),
                     self.centre("CA:C", get_ind(INIT_CA_C, th1["CA:C"])), INIT_TAU)
        self.log_cap = 5 * R + 16
        self.log_key = _empty(self.log_cap, torch.int64)
        self.log_pos = _empty(self.log_cap, torch.int32)
        self.log_scratch = _empty(_lib.load().gb_select_scratch_elems(self.log_cap), torch.int32)
        self.log_sel = _empty(self.log_cap, torch.int32)

    # ---- checkpoint / resume (bpe_iter={t}.pkl is reloaded by bin/encode.py:239-260 and stepping continues) ---------------
    def export_state(self):
        """host copy of everything a later Learner(resume=...) needs beyond the inputs: segmentation, merge events, step
        bookkeeping; medoid mode also its geometry classes and the occurrence-set op log (CPython set order emulation)"""
        self._drain()
        self._sort_pending()
        ts, T, n = self._ts, self._ts.T, self._ev_n
        u = lambda t: _u32(t).astype(np.uint32)
        st = dict(mode=self.mode, T=T, n_steps=self.n_steps, hash_base=self.hash_base, pos=u(ts.a["pos"][:T]), len=u(ts.a["len"][:T]),
                  tid=u(ts.a["tid"][:T]), ev_left=u(self._evl[:n]), ev_right=u(self._evr[:n]), ev_p=u(self._evp[:n]),
                  step_events=list(self._step_events), pop_log=list(self.pop_log), tokens=self.tokens)
        if self.mode == "M":
            st.update(cls=u(ts.a["cls"][:T]), class_geo=self.class_geo, sphere_dict=self.sphere_dict, sphere_packed=self.sphere_packed,
                      sphere_n=self.sphere_n, token_keys=self.token_keys, init_tid=u(self.init_tid),
                      log_key=self.log_key[:self.log_n].cpu().numpy(), log_pos=u(self.log_pos[:self.log_n]),
                      log_bounds=list(self.log_bounds), key_reset_at=dict(self.key_reset_at))
        return st

    def _load_state(self, st):
        if st["mode"] != self.mode:
            raise ValueError("checkpoint was written in the other quantisation mode")
        ts, T, dev = self._ts, int(st["T"]), _dev()
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int32)).to(dev)
        ts.T = T
        ts.a["pos"][:T] = up(st["pos"]); ts.a["len"][:T] = up(st["len"]); ts.a["tid"][:T] = up(st["tid"])
        last = (ts.a["pos"][:T].to(torch.int64) & 0xFFFFFFFF) + (ts.a["len"][:T].to(torch.int64) & 0xFFFF) - 1
        ts.a["glue"][:T] = self.glue_res[last]
        n = len(st["ev_left"])
        self._evl[:n] = up(st["ev_left"]); self._evr[:n] = up(st["ev_right"]); self._evp[:n] = up(st["ev_p"])
        self._ev_n = n
        self._step_events = list(st["step_events"])
        self._pop_log = list(st["pop_log"])
        self._tokens = st["tokens"]
        self._popped = {k for _, k in self._pop_log} if self.mode == "H" else set()
        self.n_steps = int(st["n_steps"])
        if self.mode == "H":
            _lib.call("gb_rehash_tokens", _lib.ptr(self.sym), _lib.ptr(self.glue_res), ts.p("pos"), ts.p("len"), T, self.hash_base, ts.p("hsh"))
            return
        self._m_setup()
        self.class_geo = st["class_geo"]
        self.class_of = {json.dumps(g): c for c, g in enumerate(self.class_geo)}
        self.class_f3 = [[g["N:CA"][0], g["CA:C"][0], g["tau"][0]] for g in self.class_geo]
        self.sphere_dict, self.sphere_packed, self.sphere_n = st["sphere_dict"], st["sphere_packed"], st["sphere_n"]
        self.token_keys = st["token_keys"]
        ts.a["cls"][:T] = up(st["cls"])
        self.init_tid = up(st["init_tid"])
        m = len(st["log_key"])
        self.log_key[:m] = torch.from_numpy(np.ascontiguousarray(st["log_key"])).to(dev)
        self.log_pos[:m] = up(st["log_pos"])
        self.log_n, self.log_bounds, self.key_reset_at = m, list(st["log_bounds"]), dict(st["key_reset_at"])

    def _oplog(self, nk, p2cls):
        ts = self.ts
        if self.log_n + 4 * nk > self.log_cap:
            raise RuntimeError("occurrence-set op log overflow")
        _lib.call("gb_oplog_m", _lib.ptr(ts.kept), nk, _lib.ptr(ts.flags), ts.p("cls"), ts.p("glue"), ts.p("pos"), ts.T,
                  _lib.ptr(ts.aux), _lib.ptr(p2cls), self.log_key.data_ptr() + 8 * self.log_n,
                  self.log_pos.data_ptr() + 4 * self.log_n)
        self.log_n += 4 * nk
        self.log_bounds.append(self.log_n)

    def set_order(self, key):
        """list(self._geo_dict[key]) (bpe.py:1744,1829,1888): replay this key's add/remove history through a real
        CPython set, in the reference's operation order, and read its iteration order."""
        _lib.call("gb_select_eq_u64", _lib.ptr(self.log_key), self.log_n, key, _lib.ptr(self.log_sel),
                  _lib.ptr(self.ts.u32), _lib.ptr(self.log_scratch))
        n = int(self.ts.u32[0].item())
        sel = self.log_sel[:n]
        both = _u32(torch.stack((sel, self.log_pos[sel.to(torch.int64)])))      # one read: log indices and their residues
        idx, rpos = both[0], both[1]
        start = self.key_reset_at.get(key, 0)
        if start:
            keep = idx >= start
            idx, rpos = idx[keep], rpos[keep]
        offs = self.bb.offsets_host
        bi = np.searchsorted(offs, rpos, side="right") - 1
        index = 3 * (rpos - offs[bi])
        bounds = np.asarray(self.log_bounds)
        seg = np.searchsorted(bounds, idx, side="right")          # 0 = initial adds, s = step s
        seg_start = np.concatenate([[0], bounds])[seg]
        is_add = (seg == 0) | (((idx - seg_start) % 4) >= 2)
        if self.comm is not None:
            # the reference's set saw, step by step, the operations of every backbone in order: shards are contiguous in rank
            # order, so the whole history is the rank-order concatenation re-grouped by step (stable sort on the step index)
            seg, is_add, bi, index = merge_shard_histories(self.comm.all_gather_object((seg, is_add, bi + self.bb_base, index)))
        return replay_set_order(seg, is_add, list(zip(bi.tolist(), index.tolist())))

    def m_key_string(self, key):
        """json string of compute_geo_key for a medoid-medoid pair (bpe.py:1248-1256): raw medoid floats,
        glue bins at the split"""
        cL, cR, g = (key >> 45) & 0x3FFFF, (key >> 27) & 0x3FFFF, key & 0x7FFFFFF
        gl, gr = self.class_geo[cL], self.class_geo[cR]
        glue = {"omega": g & 0x1FF, "C:1N:1CA": (g >> 9) & 0x1FF, "phi": (g >> 18) & 0x1FF}
        d = {}
        for k in ["N:CA", "CA:C", "0C:1N", "tau", "CA:C:1N", "C:1N:1CA", "psi", "omega", "phi"]:
            v = list(gl.get(k, []))
            if k in glue:
                v.append(int(glue[k]))
            v += gr.get(k, [])
            if v:
                d[k] = v
        return json.dumps(d, sort_keys=True), d

    def peek_m(self):
        """recount (K1) and look at the top of the priority order: (max count, ties, recurring choice or None).
        Keys that already own medoids sort first whatever their count (bpe.py:2121-2127,2164-2166)."""
        if self._peek is None:
            mx, ties = self._count_pairs() if self.comm is None else self._count_pairs_dist_m()
            rec = None
            if self.sphere_packed:
                cnts = self._sphere_counts()
                nz = np.flatnonzero(cnts)
                if nz.size:
                    pk = self._sp_keys
                    negc, key_str, key = min((-int(cnts[i]), self.sphere_packed[pk[i]], pk[i]) for i in nz.tolist())
                    rec = (-negc, key_str, key)
            self._peek = (mx, ties, rec)
        return self._peek

    def _sphere_counts(self):
        """current pair count of every key that owns medoids (one kernel, one read); the query array lives on the device and
        grows by the keys added since the last call"""
        sp = self.sphere_packed
        n = len(sp)
        keys = getattr(self, "_sp_keys", None)
        if keys is None or len(keys) > n or (keys and next(iter(sp)) != keys[0]):
            keys, self._sp_q = [], _empty(1024, torch.int64)
        if len(keys) < n:
            fresh = list(itertools.islice(sp, len(keys), None))
            if n > self._sp_q.numel():
                q = _empty(next_pow2(2 * n), torch.int64)
                q[:len(keys)] = self._sp_q[:len(keys)]
                self._sp_q = q
            self._sp_q[len(keys):n] = torch.from_numpy(np.array(fresh, dtype=np.uint64).view(np.int64)).to(_dev())
            keys = keys + fresh
        self._sp_keys = keys
        out = _empty(2 * n, torch.int32)
        _lib.call("gb_table_lookup", *self.table.ptrs(), _lib.ptr(self._sp_q), n, _lib.ptr(out), out.data_ptr() + 4 * n)
        if self.comm is not None:                  # the table holds this shard's counts
            self.comm.all_reduce(out[:n])
        return _u32(out[:n])

    def _count_pairs_dist_m(self):
        """medoid-mode K1 on a shard: exact local table -> slot counters summed over the ranks -> exact global maximum and every
        key tied at it (distributed.global_top); the table keeps the LOCAL counts for _sphere_counts"""
        from .distributed import global_top
        ts, tab, comm = self.ts, self.table, self.comm
        want = 4 * max(tab.distinct, 1024) + 4096
        tab.ensure(2 * max(ts.T, 1) if tab.M == 0 else max(want, 1 << 12))
        while True:
            _lib.call("gb_table_clear", *tab.ptrs())
            tab.overflow.zero_()
            if ts.T > 1:
                _lib.call("gb_pair_count_m", ts.p("cls"), ts.p("glue"), ts.T, *tab.ptrs(), _lib.ptr(tab.overflow))
            _lib.call("gb_table_argmax", *tab.ptrs(), _lib.ptr(tab.out4), _lib.ptr(tab.tie_keys), _lib.ptr(tab.tie_min), tab.TIE_CAP)
            out4 = _u32(tab.out4)
            if int(tab.overflow.item()) or int(out4[2]) * 2 > tab.M:
                tab.ensure(tab.M * 4)
                continue
            tab.distinct = int(out4[2])
            break
        if not hasattr(self, "_slots"):
            self._slots = _zeros(self.NS, torch.int32)
            self._mark = _zeros(self.NS, torch.uint8)
            self._ck = _empty(1 << 16, torch.int64); self._cc = _empty(1 << 16, torch.int32)
        _lib.call("gb_table_to_slots", _lib.ptr(tab.keys), _lib.ptr(tab.cnt), tab.M, _lib.ptr(self._slots), self.NS)

        def cand(sel):
            self._mark.zero_()
            self._mark[torch.from_numpy(sel).to(_dev())] = 1
            _lib.call("gb_table_dump_marked", _lib.ptr(tab.keys), _lib.ptr(tab.cnt), tab.M, _lib.ptr(self._mark), self.NS,
                      _lib.ptr(self._ck), _lib.ptr(self._cc), _lib.ptr(ts.u32), 1 << 16)
            n = int(ts.u32[0].item())
            if n > (1 << 16):
                raise RuntimeError("too many local keys in the top slots")
            return self._ck[:n].cpu().numpy(), _u32(self._cc[:n])
        mx, tied = global_top(comm, self._slots, cand, _dev(), cap=1 << 16)
        return mx, [(int(k) & M64, 0, None, None) for k in sorted(tied)]

    def step_m(self):
        """one merge iteration, medoid mode (bpe.py:1792-2166 with rmsd_key = key)"""
        ts = self.ts
        mx, ties, rec = self.peek_m()
        self._peek = None
        if mx == 0:
            raise IndexError("no adjacent pairs left")
        recurring = rec is not None
        if recurring:
            mx, key_str, key = rec
        else:
            key_str, key = min((self.m_key_string(t[0])[0], t[0]) for t in ties)
        self.pop_log.append((mx, key_str))
        cL, cR = (key >> 45) & 0x3FFFF, (key >> 27) & 0x3FFFF
        gl, gr = self.class_geo[cL], self.class_geo[cR]
        nb = lambda g: len(g["N:CA"]) + len(g["CA:C"]) + len(g.get("0C:1N", []))
        length = nb(gl) + nb(gr)
        nres = (length + 1) // 3
        A = length + 1
        _lib.call("gb_flag_pairs_m", ts.p("cls"), ts.p("glue"), ts.T, key, _lib.ptr(ts.flags))
        # the occurrence list's length is the count K1 just reported; the list's own total lands in ts.u32[7] and is compared
        # after the merge, with the read that ends the step (one synchronisation less per iteration)
        comm = self.comm
        if comm is None:
            N = mx
            _lib.call("gb_select_flagged", _lib.ptr(ts.flags), ts.T, 1, _lib.ptr(ts.idx), ts.u32.data_ptr() + 28, _lib.ptr(ts.scratch))
        else:
            N = ts.select(1, ts.idx)                               # this shard's occurrences (mx counts all shards')
        occ = ts.idx[:N].to(torch.int64)
        r0 = ts.a["pos"][occ].contiguous()
        f3c = ts.a["cls"][occ].contiguous()
        # a recurring key with a single medoid needs no geometry at all: the argmin over one distance is 0
        lone = recurring and len(self.sphere_dict[key_str]) == 1
        if lone:
            allc = None
        else:
            allc = self.coords(r0, N, nres, A, f3c) if N else torch.zeros((0, A * 3), dtype=torch.float64, device=_dev())
        if recurring:
            geos = self.sphere_dict[key_str]
            n = self.sphere_n[key_str]
            medc = None if lone else self.geo_coords(geos, A)
        elif comm is not None:
            geos, medc = self._new_medoids_dist(key, mx, N, occ, r0, f3c, allc, length)
            self.sphere_dict[key_str] = geos
            self.sphere_packed[key] = key_str
            n = len(self.tokens)
            self.sphere_n[key_str] = n
            tid0 = len(self.token_keys)
            for p, geo in enumerate(geos):
                self.tokens[(n, p)] = geo
                self.token_keys.append((n, p))
        else:
            all_pos = self.set_order(key)
            if len(all_pos) != N:
                raise RuntimeError(f"occurrence-set replay gave {len(all_pos)} entries, expected {N}")
            # one read: global residue of the right token (ascending), window start and first-residue class of every occurrence
            rpos, r0_h, f3c_h = _u32(torch.stack((ts.a["pos"][occ + 1], r0, f3c)))
            offs = self.bb.offsets_host
            ap = np.fromiter(itertools.chain.from_iterable(all_pos), dtype=np.int64, count=2 * N).reshape(N, 2)
            gp = offs[ap[:, 0]] + ap[:, 1] // 3                   # global residue of every (backbone, bond index) entry
            rank = np.searchsorted(rpos, gp)                      # occurrence rank of every set entry
            if N > self.max_num_strucs:
                active = self.rng.choice(N, self.max_num_strucs, replace=False)
            else:
                active = np.arange(N)
            act_rank = rank[active]
            actc = allc[torch.from_numpy(act_rank).to(_dev())].contiguous()
            med = self.k_medoids(actc, self.num_p(length))
            medc = actc[torch.tensor(med, device=_dev(), dtype=torch.int64)].contiguous()
            geos = []
            for m in med:
                q = int(act_rank[m])
                geos.append(self.window_geo(int(r0_h[q]), length, self.class_f3[int(f3c_h[q])]))
            self.sphere_dict[key_str] = geos
            self.sphere_packed[key] = key_str
            n = len(self.tokens)
            self.sphere_n[key_str] = n
            tid0 = len(self.token_keys)
            for p, geo in enumerate(geos):
                self.tokens[(n, p)] = geo
                self.token_keys.append((n, p))
        p2cls = torch.tensor([self.intern_class(g) for g in geos], dtype=torch.int32, device=_dev())
        if recurring:
            tid0 = self.token_keys.index((n, 0))
        p2tid = torch.arange(tid0, tid0 + len(geos), dtype=torch.int32, device=_dev())
        if N:
            asg = self.assign(allc, medc) if len(geos) > 1 else _zeros(N, torch.int32)
            _lib.call("gb_scatter_u32", _lib.ptr(ts.idx), _lib.ptr(asg), N, _lib.ptr(ts.aux))
        r = self._merge(key, n, p2cls, p2tid)
        if comm is None and int(r[7]) != mx:
            raise RuntimeError(f"flagged {int(r[7])} occurrences, histogram counted {mx}")
        self.key_reset_at[key] = self.log_n
        if not recurring:
            self.n_steps += 1
        return recurring

    def _new_medoids_dist(self, key, NG, N, occ, r0, f3c, allc, length):
        """medoids of a key met for the first time, on a shard (bpe.py:1739-1789 rmsd_partition): the occurrence set's order is
        replayed from every rank's history, the sample is drawn over it with the shared RNG stream, each sampled window's
        coordinates come from the rank that holds it, and every rank runs the same k-medoids on the same matrix"""
        ts = self.ts
        all_pos = self.set_order(key)
        if len(all_pos) != NG:
            raise RuntimeError(f"occurrence-set replay gave {len(all_pos)} entries, expected {NG}")
        ap = np.fromiter(itertools.chain.from_iterable(all_pos), dtype=np.int64, count=2 * NG).reshape(NG, 2)
        gp = self.g_offsets[ap[:, 0]] + ap[:, 1] // 3                 # global residue of every entry's right token
        if NG > self.max_num_strucs:
            active = self.rng.choice(NG, self.max_num_strucs, replace=False)
        else:
            active = np.arange(NG)
        if N:
            rpos, r0_h, f3c_h = _u32(torch.stack((ts.a["pos"][occ + 1], r0, f3c)))
        else:
            rpos = r0_h = f3c_h = np.zeros(0, dtype=np.int64)
        lp = gp[active] - self.r_base                                 # local residue where this shard holds the entry
        own = np.flatnonzero((lp >= 0) & (lp < self.bb.R))
        loc = np.searchsorted(rpos, lp[own])                          # its rank among this shard's occurrences
        if own.size and (loc.max() >= N or not np.array_equal(rpos[loc], lp[own])):
            raise RuntimeError("a sampled set entry is not an occurrence on the shard that holds its backbone")
        rows = allc[torch.from_numpy(loc.astype(np.int64)).to(_dev())]
        actc = self._gather_rows(rows, own, len(active))
        med = self.k_medoids(actc, self.num_p(length))
        medc = actc[torch.tensor(med, device=_dev(), dtype=torch.int64)].contiguous()
        where = {int(a): int(q) for a, q in zip(own.tolist(), loc.tolist())}
        mine = {p: self.window_geo(int(r0_h[where[m]]), length, self.class_f3[int(f3c_h[where[m]])])
                for p, m in enumerate(med) if m in where}
        return self._gather_geos(mine, len(med)), medc

    def geo_coords(self, geos, A):
        """Tokenizer.key_coords (tokenizer.py:204-230) for medoid geometries: each one placed in its own blank
        backbone and run through K4."""
        nres = A // 3 if A % 3 == 0 else (A - 1) // 3
        K = len(geos)
        vals = np.full((9, K * (nres + 1)), np.nan)
        vals[:3] = 0.0
        f3 = []
        for k, g in enumerate(geos):
            b = k * (nres + 1)
            f3.append([g["N:CA"][0], g["CA:C"][0], g["tau"][0]])
            for r in range(nres):
                if r > 0:
                    vals[COL["N:CA"], b + r - 1] = g["N:CA"][r]; vals[COL["CA:C"], b + r - 1] = g["CA:C"][r]
                    vals[COL["tau"], b + r - 1] = g["tau"][r]
                if r < len(g.get("0C:1N", [])):
                    vals[COL["0C:1N"], b + r] = g["0C:1N"][r]; vals[COL["CA:C:1N"], b + r] = g["CA:C:1N"][r]
                    vals[COL["psi"], b + r] = g["psi"][r]
                if r < len(g.get("omega", [])):
                    vals[COL["omega"], b + r] = g["omega"][r]; vals[COL["C:1N:1CA"], b + r] = g["C:1N:1CA"][r]
                    vals[COL["phi"], b + r + 1] = g["phi"][r]
        offsets = np.arange(K + 1, dtype=np.int64) * (nres + 1)
        mini = Backbones(vals, offsets)
        f3t = torch.tensor(f3, dtype=torch.float64, device=_dev())
        r0 = torch.from_numpy(offsets[:-1].astype(np.int32)).to(_dev())
        f3c = torch.arange(K, dtype=torch.int32, device=_dev())
        out = _empty(K * A * 3, torch.float64)
        _lib.call("gb_window_coords", _lib.ptr(mini.vals), mini.R, mini.R, _lib.ptr(mini.bb), _lib.ptr(mini.offsets),
                  _lib.ptr(r0), K, nres, A, _lib.ptr(f3c), _lib.ptr(f3t), 0.0, 0.0, 0.0, None, None, 0, _lib.ptr(out))
        return out.view(K, A * 3)

    def reconstruction(self):
        """f-4: full-chain coordinates of the CURRENT (quantised) geometry and of the pristine structure for every backbone, and
        their Kabsch backbone RMSD -- (rmsd f64 [S], coords_cur f64 [3R, 3], coords_orig f64 [3R, 3]) device tensors
        (csrc/k_recon.cu; the reference computes this per reference tokenizer, bin/encode.py:341-359)"""
        ts, bb = self.ts, self.bb
        R, S, T = bb.R, bb.S, ts.T
        th = self.th.thresholds[self.size0]
        B = max(len(v) for v in th.values())
        cen = np.zeros((9, B))
        for i, k in enumerate(SLOTS):
            if k in th:
                cen[i, :len(th[k])] = [(a + b) / 2 for a, b in th[k]]
            elif k in BOND_TYPES:
                cen[i, 0] = BOND_LENGTHS[BOND_TYPES.index(k)]          # std_bonds: bond bins are all 0
        cen_d = torch.from_numpy(cen).to(_dev())
        cg, maxres = None, 0
        if self.mode == "M":
            maxres = max((len(g["N:CA"]) for g in self.class_geo), default=1)
            tab = np.full((len(self.class_geo), maxres, 9), np.nan)
            for c, g in enumerate(self.class_geo):
                for i, k in enumerate(SLOTS):
                    v = g.get(k, [])
                    tab[c, :len(v), i] = v
            cg = torch.from_numpy(tab).to(_dev())
        tok_of = _empty(R, torch.int32)
        s_cur, s_org = _empty(9 * R, torch.float64), _empty(9 * R, torch.float64)
        c_cur, c_org = _empty(9 * R, torch.float64), _empty(9 * R, torch.float64)
        rmsd = _empty(S, torch.float64)
        _lib.call_n("gb_reconstruction", 6, _lib.ptr(self.rbins), R, _lib.ptr(cen_d), B, ts.p("pos"), ts.p("len"), ts.p("cls"), T,
                    _lib.ptr(cg) if cg is not None else None, maxres, _lib.ptr(bb.vals), R, _lib.ptr(bb.bb), _lib.ptr(bb.offsets), S,
                    INIT_N_CA, INIT_CA_C, INIT_TAU, _lib.ptr(tok_of), _lib.ptr(s_cur), _lib.ptr(s_org), _lib.ptr(c_cur), _lib.ptr(c_org),
                    _lib.ptr(rmsd))
        return rmsd, c_cur.view(3 * R, 3), c_org.view(3 * R, 3)

    def lddt_ca(self, cutoff=15.0, coords=None):
        """f-4: lDDT-C-alpha of every backbone's current quantised geometry against its pristine structure (mean of the per-residue
        scores, what bin/encode.py:351-358 takes from esm's lddt_ca per reference tokenizer) -- f64 [S] device tensor"""
        if coords is None:
            coords = self.reconstruction()[1:]
        c_cur, c_org = coords
        out = _empty(self.bb.S, torch.float64)
        _lib.call("gb_lddt_ca", _lib.ptr(c_cur), _lib.ptr(c_org), _lib.ptr(self.bb.offsets), self.bb.S, float(cutoff), _lib.ptr(out))
        return out

    def step(self):
        if self.mode == "H":
            self.step_h()
            return False
        return self.step_m()


def merge_shard_histories(parts):
    """per-rank (step, is_add, global backbone, bond index) histories of one key, each in its rank's own order -> the order a single
    process would have logged them in: inside a step the operations follow the backbones, and shards are contiguous in rank order,
    so the ranks' lists are concatenated and stably re-grouped by step"""
    seg, is_add, bi, index = (np.concatenate([np.asarray(pt[c]) for pt in parts]) for c in range(4))
    o = np.argsort(seg, kind="stable")
    return seg[o], is_add[o], bi[o], index[o]


def replay_set_order(seg, is_add, items):
    """iteration order of a CPython set after a history of adds / removes (items[i] added or removed in step seg[i]); a set
    found empty when a new step begins is replaced by a fresh one (count 0 -> _geo_dict.pop(k), bpe.py:2134-2135).
    Runs of adds go through set.update and runs of removes through map(set.remove): both walk the items one by one through
    the same table operations as single calls (set.difference_update would not: it shrinks the table afterwards)."""
    n = len(items)
    if n == 0:
        return []
    seg = np.asarray(seg)
    is_add = np.asarray(is_add, dtype=bool)
    chg = np.flatnonzero((seg[1:] != seg[:-1]) | (is_add[1:] != is_add[:-1])) + 1
    starts = [0] + chg.tolist()
    ends = chg.tolist() + [n]
    segs, adds = seg[starts].tolist(), is_add[starts].tolist()
    s = set()
    cur = 0
    for a, b, sg, add in zip(starts, ends, segs, adds):
        if sg != cur:
            if not s:
                s = set()
            cur = sg
        if add:
            s.update(items[a:b])
        else:
            collections.deque(map(s.remove, items[a:b]), maxlen=0)
    return list(s)


def parse_m_key(key_str):
    """split a medoid-mode key string (bpe.py:1248-1256) into (left geometry, glue bins, right geometry)"""
    d = json.loads(key_str)
    ints = {k: [i for i, v in enumerate(d[k]) if isinstance(v, int)] for k in ("omega", "C:1N:1CA", "phi")}
    if any(len(v) != 1 for v in ints.values()):
        raise ValueError("not a medoid-medoid pair key")
    nl = ints["omega"][0] + 1                     # residues in the left token
    gl, gr = {}, {}
    for k, v in d.items():
        if k in ints:
            i = ints[k][0]
            l, r = v[:i], v[i + 1:]
        else:
            l, r = v[:nl], v[nl:]
        if l:
            gl[k] = l
        if r:
            gr[k] = r
    order = ["N:CA", "CA:C", "0C:1N", "tau", "CA:C:1N", "C:1N:1CA", "psi", "omega", "phi"]
    gl = {k: gl[k] for k in order if k in gl}
    gr = {k: gr[k] for k in order if k in gr}
    glue = (d["omega"][ints["omega"][0]], d["C:1N:1CA"][ints["C:1N:1CA"][0]], d["phi"][ints["phi"][0]])
    return gl, glue, gr


def pack_m_key(cL, glue, cR):
    om, cn, ph = glue
    return (cL << 45) | (cR << 27) | (om & 0x1FF) | ((cn & 0x1FF) << 9) | ((ph & 0x1FF) << 18)


class Inducer(Learner):
    """K7: frozen-vocabulary encode of a batch of held-out backbones (BPE.tokenize, bpe.py:1053-1140, with
    bin_helper 1301-1314 and step_helper 1316-1425), all backbones advanced together merge by merge."""

    def __init__(self, vals, offsets, tokens, sphere_dict, thresholds, std_bonds, rmsd_super_res=True):
        if std_bonds:
            raise NotImplementedError("induce is built for free_bonds vocabularies (README setting)")
        if not rmsd_super_res:
            raise NotImplementedError("induce is built for rmsd_super_res=True vocabularies")
        if not all(isinstance(k, tuple) for k in tokens):
            raise NotImplementedError("BPE.tokenize needs a medoid-mode vocabulary (bpe.py:1077,1117)")
        t_stage = [time.perf_counter()]
        self.mode = "M"
        self.std_bonds = std_bonds
        self.size0 = 1
        self._pending, self.hl, self.async_steps, self.use_sketch, self.sketch, self.k1b_events = 0, None, False, False, None, None
        self._tokens = tokens
        self.token_keys = list(tokens)
        self.sphere_dict = sphere_dict
        self.bb = Backbones(vals, offsets)
        torch.cuda.synchronize(); t_stage.append(time.perf_counter())
        self.vals_host = np.ascontiguousarray(vals, dtype=np.float64)
        self.th = Thresholds.__new__(Thresholds)
        self.th.thresholds = {1: thresholds[1]}
        self.k1_events = None
        self._pop_log, self._step_events, self.n_steps = [], [], 0
        self.class_of, self.class_geo, self.class_f3, self._f3_dirty = {}, [], [], True
        self.d_f3 = (0.0, 0.0, INIT_TAU)
        self._quantize(bond_mode=2)
        torch.cuda.synchronize(); t_stage.append(time.perf_counter())
        R, S = self.bb.R, self.bb.S
        self.failed = (self.status_host & 3) != 0          # glue angle outside the training range -> ValueError
        self._ts = TokenState(R, "M")
        self._evl = _empty(R, torch.int32); self._evr = _empty(R, torch.int32); self._evp = _empty(R, torch.int32)
        self._ev_n = 0
        th1 = thresholds[1]
        B = max(len(th1[k]) for k in BOND_TYPES)
        bc = np.zeros((3, B))
        for i, k in enumerate(BOND_TYPES):               # residue-major slots 0,1,2 = N:CA, CA:C, 0C:1N
            bc[i, :len(th1[k])] = [(a + b) / 2 for a, b in th1[k]]
        self.bond_centres = torch.from_numpy(bc).to(_dev())
        tid_of = {k: i for i, k in enumerate(self.token_keys)}
        ns = sorted(set(k[0] for k in tokens))
        geo_keys = list(sphere_dict)
        assert len(ns) == len(geo_keys)
        # ---- residue tokens (bpe.py:1072-1099): nearest medoid of (0,p) for inner residues, (1,p) for the last one
        is_last = torch.zeros(R, dtype=torch.bool, device=_dev())
        is_last[(self.bb.offsets[1:] - 1)] = True
        lists = {3: torch.nonzero(~is_last).flatten().to(torch.int32), 2: torch.nonzero(is_last).flatten().to(torch.int32)}
        # which of the (0, p) / (1, p) groups holds the inner-residue medoids is a property of the VOCABULARY (a 3-bond residue token
        # carries psi / 0C:1N, the 2-bond last-residue token does not), not of the first held-out backbone's length
        order = [3, 2] if "psi" in tokens[(0, 0)] else [2, 3]
        cls_res = _zeros(R, torch.int32); tid_res = _zeros(R, torch.int32)
        for n, size in enumerate(order):
            occ = lists[size].contiguous()
            N = int(occ.numel())
            if N == 0:
                continue
            geos = []
            p = 0
            while (n, p) in tokens:
                geos.append(tokens[(n, p)]); p += 1
            A = size + 1
            medc = self.geo_coords(geos, A)
            allc = self.coords(occ, N, 1, A, None, self.bond_centres)
            asg = self.assign(allc, medc).to(torch.int64)
            p2cls = torch.tensor([self.intern_class(g) for g in geos], dtype=torch.int32, device=_dev())
            p2tid = torch.tensor([tid_of[(n, q)] for q in range(len(geos))], dtype=torch.int32, device=_dev())
            cls_res[occ.to(torch.int64)] = p2cls[asg]
            tid_res[occ.to(torch.int64)] = p2tid[asg]
        ts = self.ts
        ts.T = R
        ts.a["cls"][:R] = cls_res
        ts.a["glue"][:R] = self.glue_res
        ts.a["pos"][:R] = torch.arange(R, device=_dev(), dtype=torch.int32)
        ts.a["len"][:R] = 1
        ts.a["tid"][:R] = tid_res
        self.init_tid = tid_res.clone()
        # failed backbones must not take part: cut every one of their pairs (the reference aborts them, bpe.py:1105-1108)
        if self.failed.any():
            bad = torch.from_numpy(self.failed).to(_dev())[self.bb.bb.to(torch.int64)]
            ts.a["glue"][:R] = torch.where(bad, torch.full_like(self.glue_res, -(1 << 31)), self.glue_res)
        torch.cuda.synchronize(); t_stage.append(time.perf_counter())
        # ---- merge replay in learned order (bpe.py:1116-1139)
        merges = list(zip(ns, geo_keys))[2:]
        plan = []                            # per vocabulary merge: None (cannot fire) or (key, geos, p2cls, tid0, nres, A)
        nb = lambda g: len(g["N:CA"]) + len(g["CA:C"]) + len(g.get("0C:1N", []))
        for n, key_str in merges:
            gl, glue, gr = parse_m_key(key_str)
            cL, cR = self.class_of.get(json.dumps(gl)), self.class_of.get(json.dumps(gr))
            geos = sphere_dict[key_str]
            p2cls = [self.intern_class(g) for g in geos]
            if cL is None or cR is None:
                plan.append(None)            # a constituent motif never occurred -> key not in geo_dict
                continue
            length = nb(gl) + nb(gr)
            plan.append((pack_m_key(cL, glue, cR), geos, p2cls, tid_of[(n, 0)], (length + 1) // 3, length + 1, n))
        amax = int(_lib.load().gb_induce_max_atoms())
        torch.cuda.synchronize(); t_stage.append(time.perf_counter())
        if os.environ.get("GEOBPE_INDUCE_HOST") == "1" or any(pl is not None and pl[5] > amax for pl in plan):
            self._replay_host(merges, plan)
        else:
            self._replay_device(plan)
        torch.cuda.synchronize(); t_stage.append(time.perf_counter())
        d = np.diff(t_stage).tolist()
        self.stage_s = dict(zip(("h2d", "k3_quantise", "residue_tokens_k4_k5", "replay_plan_host", "merge_replay"), d))
        if getattr(self, "replay_s", None):
            self.stage_s["merge_replay_parts"] = self.replay_s

    def _replay_device(self, plan):
        """all vocabulary merges applied on the device without a host round trip (csrc/k_induce.cu)"""
        ts, dev = self.ts, _dev()
        t0 = time.perf_counter()
        live = [pl for pl in plan if pl is not None]
        R, N = self.bb.R, ts.T
        nm = len(live)
        if nm == 0 or N == 0:
            self.replay = [(self.ev_n, 0)] * len(plan)
            return
        # medoid coordinates of every merge: one K4 call per window size
        by_A = {}
        for i, pl in enumerate(live):
            by_A.setdefault(pl[5], []).append(i)
        med_parts, med_off, p2_off, cur, curp = [], [0] * nm, [0] * nm, 0, 0
        p2cls_all, p2tid_all = [], []
        for A, idxs in by_A.items():
            geos = [g for i in idxs for g in live[i][1]]
            mc = self.geo_coords(geos, A).reshape(-1)
            o = 0
            for i in idxs:
                med_off[i] = cur + o
                o += len(live[i][1]) * A * 3
            med_parts.append(mc)
            cur += mc.numel()
        for i, pl in enumerate(live):
            p2_off[i] = curp
            p2cls_all += pl[2]
            p2tid_all += list(range(pl[3], pl[3] + len(pl[1])))
            curp += len(pl[1])
        med = torch.cat(med_parts).contiguous()
        p2cls = torch.tensor(p2cls_all, dtype=torch.int32, device=dev)
        p2tid = torch.tensor(p2tid_all, dtype=torch.int32, device=dev)
        st = _zeros(8, torch.int32)
        rep_start = _zeros(nm + 1, torch.int32)
        rep_cnt = _zeros(nm, torch.int32)
        rmark = _zeros(ts.cap, torch.uint8)
        occ_r = _empty(ts.cap, torch.int32)
        cw = _empty(ts.cap, torch.int32)               # the flag pass's stream words
        ts.flags.zero_()
        arr = lambda xs, ct: (ct * nm)(*xs)
        mkey = arr([pl[0] for pl in live], C.c_uint64)
        mK, mA, mnres = arr([len(pl[1]) for pl in live], C.c_int), arr([pl[5] for pl in live], C.c_int), arr([pl[4] for pl in live], C.c_int)
        moff, poff = arr(med_off, C.c_int64), arr(p2_off, C.c_int64)
        bb = self.bb
        torch.cuda.synchronize(); t1 = time.perf_counter()
        _lib.call_n("gb_induce_replay", 4 * nm + 1, ts.p("cls"), ts.p("glue"), ts.p("pos"), ts.p("len"), ts.p("tid"), _lib.ptr(ts.flags),
                    _lib.ptr(rmark), _lib.ptr(ts.aux), _lib.ptr(cw), N, _lib.ptr(bb.vals), R, _lib.ptr(bb.bb), _lib.ptr(bb.offsets), _lib.ptr(self.f3tab()),
                    nm, mkey, mK, mA, mnres, moff, poff, _lib.ptr(med), _lib.ptr(p2cls), _lib.ptr(p2tid), _lib.ptr(ts.idx), _lib.ptr(occ_r),
                    _lib.ptr(self._evl), _lib.ptr(self._evr), _lib.ptr(self._evp), self._evl.numel(), _lib.ptr(st), _lib.ptr(rep_start),
                    _lib.ptr(rep_cnt))
        torch.cuda.synchronize(); t2 = time.perf_counter()
        # one compaction at the end: the host sees dense arrays (len = plain residue counts)
        ls = ts.a["len"][:N]
        idx = torch.nonzero(ls != 0).flatten()
        T = int(idx.numel())
        for f in ts.a:
            ts.b[f][:T] = ts.a[f][:N][idx] if f != "len" else (ls[idx] & 0xFFFF)
        ts.swap()
        ts.T = T
        sth = _u32(st)
        if int(sth[3]):
            raise RuntimeError("merge-event log overflow")
        self._ev_n = int(sth[2])
        cnt = _u32(rep_cnt)
        start = _u32(rep_start)
        if self._ev_n > 1:
            self._sort_events(0, self._ev_n, rep_cnt)
        self.replay, k, at = [], 0, 0
        for pl in plan:
            if pl is None:
                self.replay.append((at, 0))
                continue
            self.replay.append((int(start[k]), int(cnt[k])))
            if cnt[k]:
                self._step_events.append((pl[6], int(start[k]), int(cnt[k]), pl[0]))
            at = int(start[k]) + int(cnt[k])
            k += 1
        self.n_steps = int((cnt > 0).sum())
        self.replay_s = {"medoid_coords_setup": t1 - t0, "kernels": t2 - t1, "compact_sort_readback": time.perf_counter() - t2}

    def _replay_host(self, merges, plan):
        """host-driven replay (one K2 compaction and one read-back per merge): windows beyond the device kernel's atom limit"""
        ts = self.ts
        tid_of = {k: i for i, k in enumerate(self.token_keys)}
        self.replay = []
        for (n, key_str), pl in zip(merges, plan):
            self.replay.append((self.ev_n, 0))
            if pl is None:
                continue
            key, geos, p2cls_l, tid0, nres, A, _n = pl
            p2cls = torch.tensor(p2cls_l, dtype=torch.int32, device=_dev())
            _lib.call("gb_flag_pairs_m", ts.p("cls"), ts.p("glue"), ts.T, key, _lib.ptr(ts.flags))
            N = ts.select(1, ts.idx)
            if N == 0:
                continue
            occ = ts.idx[:N].to(torch.int64)
            r0 = ts.a["pos"][occ].contiguous(); f3c = ts.a["cls"][occ].contiguous()
            allc = self.coords(r0, N, nres, A, f3c)
            medc = self.geo_coords(geos, A)
            asg = self.assign(allc, medc)
            _lib.call("gb_scatter_u32", _lib.ptr(ts.idx), _lib.ptr(asg), N, _lib.ptr(ts.aux))
            p2tid = torch.arange(tid0, tid0 + len(geos), dtype=torch.int32, device=_dev())
            e0 = self.ev_n
            self._merge(key, n, p2cls, p2tid)
            self.replay[-1] = (e0, self.ev_n - e0)
            self.n_steps += 1

    def _oplog(self, nk, p2cls):          # induce needs no occurrence-set order
        return None
