"""GPU parity of the round-1 K1/K2 redesign, through the C ABI:

* the device-driven histogram-mode loop (gb_h_loop: tie-break, overlap rule, in-place merges, compaction periods,
  CUDA-graph replays) against the host-driven path (exact pair table, python json tie-break) on the same inputs --
  pop order, counts, segmentation, merge events -- at sizes that exercise many ties, whole compaction periods and
  (with the caps lowered) every halt / fallback path;
* the key-less slot sketch (gb_slot_count_*) against a numpy histogram of the same keys;
* size-independent properties at the BASELINE config-2 size: residues are conserved, token starts stay sorted,
  the pop counts never grow, every iteration's merge events are disjoint and sorted.
"""
import numpy as np
import pytest
import torch

import geobpe_oracle as O
from geobpe_b200 import _lib, engine as E, synthetic

pytestmark = pytest.mark.gpu
HP = dict(bins={1: 6}, bin_strategy="histogram", std_bonds=True, mode="H")


def learner(S, n, seed, loop=None, **kw):
    """loop = 1: per-iteration candidate pass replayed from CUDA graphs (k_hloop.cu); 2 (default on one GPU): persistent
    cooperative kernel with incrementally maintained pair counts (k_hloop2.cu)"""
    import os
    vals, offs = synthetic.generate(S, n, seed=seed, ragged=True)
    old = os.environ.get("GEOBPE_LOOP")
    if loop is not None:
        os.environ["GEOBPE_LOOP"] = str(loop)
    try:
        return E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **dict(HP, **kw))
    finally:
        if loop is not None:
            os.environ.pop("GEOBPE_LOOP") if old is None else os.environ.__setitem__("GEOBPE_LOOP", old)


def state(L):
    ts = L.ts
    T = ts.T
    ev = [(n, E._u32(L.ev_left[e0:e0 + c]).tolist(), E._u32(L.ev_right[e0:e0 + c]).tolist()) for n, e0, c, _ in L.step_events]
    return (list(L.pop_log), {k: v for k, v in L.tokens.items()}, E._u32(ts.a["pos"][:T]).tolist(), E._u32(ts.a["len"][:T]).tolist(),
            E._u32(ts.a["tid"][:T]).tolist(), E._u32(ts.a["glue"][:T]).tolist(), ev)


@pytest.mark.parametrize("loop", [1, 2])
@pytest.mark.parametrize("S,n,seed,steps", [(300, 64, 3, 400), (40, 48, 4, 150)])
def test_device_loop_equals_host_path(S, n, seed, steps, loop):
    a = learner(S, n, seed, loop=loop, async_steps=True)
    b = learner(S, n, seed, async_steps=False, use_sketch=False)     # exact table + host tie-break (round-1 first design)
    for _ in range(steps):
        a.step()
        b.step()
    sa, sb = state(a), state(b)
    assert sa[0] == sb[0], "pop order / counts differ"
    for i in range(1, 7):
        assert sa[i] == sb[i]
    # most iterations are resolved on the device, tie-breaks included; once the top count is down to 2-3 these small
    # inputs tie hundreds of keys (> HL_TB_CAP = 128) and the loop hands those iterations to the host path (code 1 / 2)
    assert a.hl.steps >= steps // 3 and set(a.hl.halts) <= {1, 2, 11}
    assert a.hl.steps + sum(v for k, v in a.hl.halts.items() if k in (1, 2)) == steps
    assert max(len([p for p in sa[0] if p[0] == c]) for c in {p[0] for p in sa[0]}) > 1   # the run did contain tied counts


def test_device_loop_vs_oracle_with_ties():
    vals, offs = synthetic.generate(50, 40, seed=21, ragged=True)
    strucs = synthetic.to_structures(vals, offs)
    hp = dict(bins={1: 4}, bin_strategy="histogram", std_bonds=True, rmsd_partition_min_size=float("inf"))
    ora = O.BPE(strucs, res_init=True, seed=0, **hp)
    ora.initialize(); ora.bin()
    L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **dict(HP, bins={1: 4}))
    for _ in range(130):                                   # crosses two compaction periods (64)
        ora.step(); L.step()
    assert [list(p) for p in L.pop_log] == [list(p) for p in ora.pop_log]


@pytest.mark.parametrize("loop", [1, 2])
def test_halts_fall_back_to_the_host_path(monkeypatch, loop):
    """caps lowered so that the device loop halts (candidate overflow) and hands single iterations to the host"""
    monkeypatch.setattr(E.SlotSketch, "CAP", 1)
    a = learner(120, 48, 5, loop=loop, async_steps=True)
    b = learner(120, 48, 5, async_steps=False, use_sketch=False)
    for _ in range(90):
        a.step(); b.step()
    assert state(a) == state(b)
    assert a.hl.halts or a.hl.r2, "the lowered cap was expected to force second rounds / halts"


@pytest.mark.parametrize("incremental,round2", [(0, 1), (1, 0), (0, 0)])
def test_loop_switches(incremental, round2, monkeypatch):
    """gb_h_loop_incremental / gb_h_loop_round2: recount every iteration instead of slot bounds kept from the merge events;
    no in-loop second round (inconclusive iterations halt to the host path).  Same results in every combination."""
    lib = _lib.load()
    lib.gb_h_loop_incremental(incremental)
    lib.gb_h_loop_round2(round2)
    try:
        monkeypatch.setattr(E.SlotSketch, "CAP", 2)      # two candidate slots per round: second rounds are common
        a = learner(120, 48, 5, loop=1, async_steps=True)
        b = learner(120, 48, 5, async_steps=False, use_sketch=False)
        for _ in range(140):                              # crosses two compaction periods
            a.step(); b.step()
        assert state(a) == state(b)
        assert a.hl.r2 > 0, "the lowered cap was expected to make first rounds inconclusive"
        if not round2:
            assert a.hl.halts.get(2, 0) > 0               # ... and each of them went to the host path
    finally:
        lib.gb_h_loop_incremental(1)
        lib.gb_h_loop_round2(1)


@pytest.mark.parametrize("loop", [1, 2])
def test_step_log_capacity_halt(monkeypatch, loop):
    """a full step log halts the loop (code 7); the host re-bases it and re-enqueues the rest"""
    monkeypatch.setattr(E.Learner, "HL_LOG_CAP", 16)
    a = learner(60, 48, 8, loop=loop, async_steps=True)
    b = learner(60, 48, 8, async_steps=False, use_sketch=False)
    for _ in range(70):
        a.step(); b.step()
    assert state(a) == state(b)
    assert a.hl.halts.get(7, 0) >= 3


@pytest.mark.parametrize("loop", [1, 2])
def test_interleaved_reads_and_steps(loop):
    """reading state drains the loop; stepping afterwards re-enters it (parity of the ping-pong buffers, counters)"""
    a = learner(80, 48, 6, loop=loop, async_steps=True)
    b = learner(80, 48, 6, async_steps=False, use_sketch=False)
    for chunk in (1, 2, 70, 3, 64, 5):
        for _ in range(chunk):
            a.step(); b.step()
        assert a.ts.T == b.ts.T and len(a.pop_log) == len(b.pop_log)
    assert state(a) == state(b)


def _mix_slot(keys, log2ns):
    with np.errstate(over="ignore"):
        return ((keys.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(64 - log2ns)).astype(np.int64)


@pytest.mark.parametrize("variant", [0, 1, 2, 8, 10])
def test_slot_sketch_matches_numpy_histogram(variant):
    rng = np.random.default_rng(7)
    T = 200_003
    cls = rng.integers(0, 40, T).astype(np.int32)
    glue = (rng.integers(0, 6, T) | (rng.integers(0, 6, T) << 10) | (rng.integers(0, 6, T) << 20)).astype(np.int64)
    last = rng.random(T) < 0.01
    last[-1] = True
    glue = np.where(last, glue | 0x80000000, glue).astype(np.uint32).view(np.int32)
    dev = torch.device("cuda")
    pad = lambda x: torch.from_numpy(np.concatenate([x, np.zeros(8, dtype=x.dtype)])).to(dev)
    c, g = pad(cls), pad(glue)
    log2ns = 16
    slots = torch.zeros(1 << log2ns, dtype=torch.int32, device=dev)
    _lib.call("gb_slot_count_m", _lib.ptr(c), _lib.ptr(g), T, _lib.ptr(slots), log2ns, None, variant, 2)
    gu = glue.view(np.uint32).astype(np.uint64)
    g27 = (gu & 0x1FF) | (((gu >> 10) & 0x1FF) << 9) | (((gu >> 20) & 0x1FF) << 18)
    keys = (cls[:-1].astype(np.uint64) << 45) | (cls[1:].astype(np.uint64) << 27) | g27[:-1]
    keys = keys[~last[:-1]]
    want = np.bincount(_mix_slot(keys, log2ns), minlength=1 << log2ns)
    assert np.array_equal(slots.cpu().numpy().astype(np.int64), want)
    # scan + exact resolution of the top slot(s)
    sout = torch.zeros(4, dtype=torch.int32, device=dev); cand = torch.zeros(4096, dtype=torch.int32, device=dev)
    bmax = torch.zeros((1 << log2ns) // 4096, dtype=torch.int32, device=dev)
    _lib.call("gb_slots_top", _lib.ptr(slots), log2ns, 0, _lib.ptr(sout), _lib.ptr(cand), 4096, _lib.ptr(bmax))
    so = sout.cpu().numpy()
    assert so[0] == want.max() and so[1] == (want == want.max()).sum() and so[2] == (want > 0).sum()
    tab = E.PairTable(); tab.ensure(1 << 13)
    _lib.call("gb_table_clear", *tab.ptrs()); tab.overflow.zero_()
    _lib.call("gb_cand_count_m", _lib.ptr(c), _lib.ptr(g), T, log2ns, _lib.ptr(cand), sout.data_ptr() + 4, 4096, *tab.ptrs(),
              _lib.ptr(tab.overflow), 0)
    _lib.call("gb_table_argmax", *tab.ptrs(), _lib.ptr(tab.out4), _lib.ptr(tab.tie_keys), _lib.ptr(tab.tie_min), tab.TIE_CAP)
    o4 = tab.out4.cpu().numpy()
    top_slots = set(np.nonzero(want == want.max())[0].tolist())
    in_top = keys[np.isin(_mix_slot(keys, log2ns), list(top_slots))]
    uk, uc = np.unique(in_top, return_counts=True)
    assert o4[0] == uc.max() and o4[2] == len(uk)


def test_config2_size_properties():
    """BASELINE configs[1] size (10 k x 256): size-independent invariants after 300 device-driven iterations"""
    vals, offs = synthetic.generate(10000, 256, seed=0)
    L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **dict(HP, bins={1: 8}))
    T0 = L.ts.T
    for _ in range(300):
        L.step()
    ts = L.ts
    T = ts.T
    pos, ln = E._u32(ts.a["pos"][:T]), E._u32(ts.a["len"][:T])
    assert ln.sum() == L.bb.R and np.all(np.diff(pos) > 0) and np.array_equal(pos[1:], pos[:-1] + ln[:-1])
    counts = [c for c, _ in L.pop_log]
    assert all(x >= y for x, y in zip(counts, counts[1:])), "the popped count can never grow"
    assert T0 - T == sum(c for _, _, c, _ in L.step_events) == L.ev_n
    for n, e0, c, _ in L.step_events[::37]:
        el, er = E._u32(L.ev_left[e0:e0 + c]), E._u32(L.ev_right[e0:e0 + c])
        assert np.all(np.diff(el) > 0) and np.all(er > el) and np.all(el[1:] > er[:-1])
    assert len(set(k for _, k in L.pop_log)) == len(L.pop_log), "a key is popped once in histogram mode"
