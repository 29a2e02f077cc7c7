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
