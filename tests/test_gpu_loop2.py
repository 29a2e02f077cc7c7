"""GPU parity of the second device-driven loop (csrc/k_hloop2.cu: persistent cooperative kernel, pair counts maintained
incrementally from the merge events) and the parity pins the round-1 review asked for:

* the slot counters the loop maintains by decrement / increment equal a fresh recount (gb_slot_count_h) after hundreds of
  iterations, periods and compactions;
* the loop against the CPU oracle at a quarter of the BASELINE configs[1] size (2 500 x 256, bins 8, 300 merges): pop log
  and every backbone's segmentation;
* the kernel variants that only switch on at scale (cache-streaming loads above 96 MB of token arrays, the shared-memory
  stage of k_slot_count above 9.7 M pairs) forced on at a small size, in both loops, against the host path;
* re-keying the run under another hash base (the recovery from a detected 64-bit collision) changes nothing;
* tiny thresholds / tables so that periods end for every reason the device knows (threshold crossings, inconclusive
  tables, table growth) and the result still equals the host path.
"""
import os

import numpy as np
import pytest
import torch

import geobpe_oracle as O
from geobpe_b200 import _lib, engine as E, synthetic
from test_gpu_loop import learner, state, HP

pytestmark = pytest.mark.gpu


def recount(L):
    """fresh k_slot_count over the (dense) token arrays of L with its own sketch parameters"""
    ts, sk = L.ts, L.sketch
    slots = torch.zeros_like(sk.slots)
    _lib.call("gb_slot_count_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(L.powtab), L.hash_base,
              _lib.ptr(slots), sk.log2ns, None, 0, 2)
    return slots


@pytest.mark.parametrize("S,n,seed,steps", [(300, 64, 3, 500), (1500, 96, 9, 700)])
def test_incremental_slot_counters_stay_exact(S, n, seed, steps):
    a = learner(S, n, seed, loop=2, async_steps=True)
    done = 0
    for chunk in (1, 7, 64, 200, steps):
        for _ in range(chunk):
            a.step()
        done += chunk
        a.sync()
        if a.hl.slots_valid:            # (a host-path iteration invalidates them: they are recounted on re-entry)
            assert torch.equal(recount(a), a.sketch.slots), f"slot counters drifted after {done} iterations"
    assert a.hl.steps > steps // 2 and a.hl.stats["periods"] >= 2


@pytest.mark.parametrize("thr16,table_log2", [(16, 10), (8, 15), (15, 11)])
def test_period_ends_for_every_reason(monkeypatch, thr16, table_log2):
    """threshold = the top slot count itself (every iteration is inconclusive or crosses), a tiny table (growth halts), a
    low threshold (many candidates)"""
    monkeypatch.setenv("GEOBPE_H2_THR16", str(thr16))
    monkeypatch.setenv("GEOBPE_H2_TABLE_LOG2", str(table_log2))
    a = learner(200, 64, 12, loop=2, async_steps=True)
    b = learner(200, 64, 12, async_steps=False, use_sketch=False)
    for _ in range(260):
        a.step(); b.step()
    assert state(a) == state(b)
    st = a.hl.stats
    assert st["periods"] >= 3
    if table_log2 <= 11:
        assert a.hl.halts.get(4, 0) >= 1 or a.hl.M > (1 << table_log2) or st["periods"] > 3


def test_loop_vs_oracle_at_quarter_config2():
    """2 500 x 256 residues, bins 8, 300 merges: the device loop's pop log and every backbone's segmentation against the
    CPU oracle (the oracle needs ~40 s of initialize() + bin() at this size)"""
    vals, offs = synthetic.generate(2500, 256, seed=0)
    strucs = synthetic.to_structures(vals, offs)
    hp = dict(bins={1: 8}, bin_strategy="histogram", std_bonds=True, rmsd_partition_min_size=float("inf"))
    ora = O.BPE(strucs, res_init=True, seed=0, **hp)
    ora.initialize(); ora.bin()
    L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **dict(HP, bins={1: 8}))
    for _ in range(300):
        ora.step(); L.step()
    assert [list(p) for p in L.pop_log] == [list(p) for p in ora.pop_log]
    ts = L.ts
    pos, ln, tid = E._u32(ts.a["pos"][:ts.T]), E._u32(ts.a["len"][:ts.T]), E._u32(ts.a["tid"][:ts.T])
    j = 0
    for i, t in enumerate(ora.tokenizers):
        a0 = int(offs[i])
        seg = [(3 * (int(pos[k]) - a0), int(tid[k])) for k in range(j, j + len(t.bond_to_token))]
        assert seg == [(s, tok) for s, tok, _ in t.bond_to_token.values()], f"backbone {i}"
        j += len(t.bond_to_token)
    assert j == ts.T
    assert L.hl.steps >= 290, "the device loop was expected to resolve (nearly) every iteration at this size"


@pytest.mark.parametrize("loop", [1, 2])
def test_at_scale_variants_forced_on(monkeypatch, loop):
    """cache-streaming key functors (KeysHT<true> / KeysHLT<true>) and the shared-memory stage of k_slot_count normally
    switch on above 96 MB of token arrays / 9.7 M pairs: forced on at a small size, the loop still equals the host path"""
    monkeypatch.setattr(E.SlotSketch, "STREAM_BYTES", 0)
    monkeypatch.setattr(E.SlotSketch, "STAGE_PAIRS", 0)
    a = learner(400, 64, 14, loop=loop, async_steps=True)
    b = learner(400, 64, 14, async_steps=False, use_sketch=False)
    for _ in range(200):
        a.step(); b.step()
    assert a.sketch.variant & 6, "the shared-memory stage was expected to be on"
    assert state(a) == state(b)
    assert a.hl.steps >= 150


@pytest.mark.parametrize("variant", [0, 2, 8, 10])
def test_slot_sketch_h_keys_match_numpy(variant):
    """gb_slot_count_h (plain / staged / cache-streaming) against a numpy histogram of the same content-hash keys"""
    a = learner(300, 64, 15, async_steps=False)
    ts, sk = a.ts, E.SlotSketch(a.ts.T, log2ns=16)
    T = ts.T
    _lib.call("gb_slot_count_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), T, _lib.ptr(a.powtab), a.hash_base,
              _lib.ptr(sk.slots), sk.log2ns, None, variant, 2)
    h = ts.a["hsh"][:T].cpu().numpy().astype(np.uint64)
    g = E._u32(ts.a["glue"][:T]).astype(np.uint64)
    ln = E._u32(ts.a["len"][:T])
    pw = a.powtab.cpu().numpy().astype(np.uint64)
    P = np.uint64(a.hash_base)
    with np.errstate(over="ignore"):
        gs = (g[:-1] & np.uint64(0x3FFFFFFF)) + np.uint64(1)
        keys = (h[:-1] * P + gs) * pw[2 * ln[1:] - 1] + h[1:]
        keys = np.where(keys == np.uint64(0xFFFFFFFFFFFFFFFF), keys - np.uint64(1), keys)
        keys = keys[(g[:-1] & np.uint64(0x80000000)) == 0]
        want = np.bincount(((keys * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(48)).astype(np.int64), minlength=1 << 16)
    assert np.array_equal(sk.slots.cpu().numpy().astype(np.int64), want)


@pytest.mark.parametrize("loop", [1, 2])
def test_rehash_mid_run_changes_nothing(loop):
    """the recovery from a detected content-hash collision: every live token is re-keyed under another base"""
    a = learner(150, 48, 16, loop=loop, async_steps=True)
    b = learner(150, 48, 16, async_steps=False, use_sketch=False)
    for _ in range(70):
        a.step(); b.step()
    a.rehash()
    for _ in range(70):
        a.step(); b.step()
    a.rehash(0x9E3779B97F4A7C15)
    for _ in range(30):
        a.step(); b.step()
    sa, sb = state(a), state(b)
    assert sa[0] == sb[0]
    for i in (1, 2, 3, 4, 5):
        assert sa[i] == sb[i]
    assert [(n, l, r) for n, l, r in sa[6]] == [(n, l, r) for n, l, r in sb[6]]
