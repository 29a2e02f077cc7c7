"""replay_set_order (engine.py) feeds runs of adds / removes to a CPython set in bulk; the medoid sample order depends on the
set's iteration order (list(self._geo_dict[key]), bpe.py:1744,1829,1888), so the bulk form must leave the table exactly as
one call per item does."""
import random

import numpy as np


def _one_by_one(seg, is_add, items):
    s, cur = set(), 0
    for sg, add, t in zip(seg, is_add, items):
        if sg != cur:
            if not s:
                s = set()
            cur = sg
        if add:
            s.add(t)
        else:
            s.remove(t)
    return list(s)


def test_bulk_replay_keeps_cpython_set_order():
    from geobpe_b200.engine import replay_set_order
    rnd = random.Random(0)
    for trial in range(200):
        live, seg, ia, items = set(), [], [], []
        for sg in range(rnd.randint(1, 30)):
            for _ in range(rnd.randint(0, 300 if sg == 0 else 40)):
                if live and rnd.random() < (0.1 if sg == 0 else 0.55):
                    t = rnd.choice(sorted(live)); live.remove(t)
                    seg.append(sg); ia.append(False); items.append(t)
                else:
                    t = (rnd.randint(0, 50), 3 * rnd.randint(0, 300))
                    if t in live:
                        continue
                    live.add(t)
                    seg.append(sg); ia.append(True); items.append(t)
            if rnd.random() < 0.1:                      # the key's count drops to zero at the end of a step
                for t in sorted(live):
                    seg.append(sg); ia.append(False); items.append(t)
                live.clear()
        got = replay_set_order(np.array(seg, dtype=np.int64), np.array(ia, dtype=bool), items)
        assert got == _one_by_one(seg, ia, items), trial
    assert replay_set_order(np.zeros(0, dtype=np.int64), np.zeros(0, dtype=bool), []) == []


def test_shard_histories_merge_to_the_single_process_order():
    """a history logged by one process (steps ascending, backbones ascending inside a step), cut into contiguous backbone shards,
    comes back in the same order from merge_shard_histories"""
    from geobpe_b200.engine import merge_shard_histories
    rnd = random.Random(3)
    for trial in range(50):
        nb = rnd.randint(2, 40)
        seg, ia, bi, ix = [], [], [], []
        for sg in range(rnd.randint(1, 25)):
            for b in range(nb):                                   # inside a step: backbone by backbone, positions ascending
                for pos in sorted(rnd.sample(range(100), rnd.randint(0, 4))):
                    for add in ((True,) if sg == 0 else (False, False, True, True)):
                        seg.append(sg); ia.append(add); bi.append(b); ix.append(3 * pos)
        whole = tuple(np.array(a) for a in (seg, ia, bi, ix))
        world = rnd.randint(1, 4)
        cuts = sorted(rnd.sample(range(1, nb), min(world - 1, nb - 1))) if world > 1 else []
        edges = [0] + cuts + [nb]
        parts = []
        for a, b in zip(edges[:-1], edges[1:]):
            sel = (whole[2] >= a) & (whole[2] < b)
            parts.append(tuple(w[sel] for w in whole))
        got = merge_shard_histories(parts)
        assert all(np.array_equal(g, w) for g, w in zip(got, whole)), trial
