"""The device tie-break order (csrc/k_hloop2.cu h2_decide: lexicographic order of element-rank sequences) against the
reference's: python string order of json.dumps(key, sort_keys=True) (bpe.py:1149 hash_geo, 1469 SortedDict key)."""
import json

import numpy as np

from geobpe_b200.engine import tie_rank_sequence, tie_rank_table


def random_key(rng, bins):
    nres = int(rng.integers(2, 7))
    m = nres - int(rng.integers(0, 2))
    b = lambda n: [int(x) for x in rng.integers(0, bins, n)]
    return {"0C:1N": [0] * m, "C:1N:1CA": b(nres - 1), "CA:C": [0] * nres, "CA:C:1N": b(m), "N:CA": [0] * nres,
            "omega": b(nres - 1), "phi": b(nres - 1), "psi": b(m), "tau": b(nres)}


def test_rank_sequence_order_equals_json_string_order():
    tab = tie_rank_table()
    assert sorted(tab.ravel().tolist()) == list(range(2048))
    for seed, bins in ((0, 3), (1, 12), (2, 130), (3, 512), (4, 2)):
        rng = np.random.default_rng(seed)
        keys = [random_key(rng, bins) for _ in range(3000)]
        strs = [json.dumps(k, sort_keys=True) for k in keys]
        uniq = sorted(set(range(len(keys))), key=lambda i: strs[i])
        by_str = [strs[i] for i in uniq]
        by_rank = [strs[i] for i in sorted(range(len(keys)), key=lambda i: tie_rank_sequence(keys[i], tab))]
        assert by_rank == by_str


def test_near_identical_keys():
    """keys that differ in one late element, in list length only, or in digit count (9 vs 10 vs 100)"""
    tab = tie_rank_table()
    base = {"0C:1N": [0, 0], "C:1N:1CA": [9], "CA:C": [0, 0], "CA:C:1N": [10, 1], "N:CA": [0, 0], "omega": [100], "phi": [1],
            "psi": [19, 2], "tau": [2, 20]}
    variants = []
    for name in ("C:1N:1CA", "CA:C:1N", "omega", "phi", "psi", "tau"):
        for v in (0, 1, 2, 9, 10, 11, 19, 20, 99, 100, 101, 199, 200):
            for pos in range(len(base[name])):
                k = {a: list(b) for a, b in base.items()}
                k[name][pos] = v
                variants.append(k)
    short = {"0C:1N": [0], "C:1N:1CA": [9], "CA:C": [0, 0], "CA:C:1N": [10], "N:CA": [0, 0], "omega": [100], "phi": [1],
             "psi": [19], "tau": [2, 20]}                       # the same window ending its backbone (m = nres - 1)
    variants.append(short)
    strs = {json.dumps(k, sort_keys=True): k for k in variants}
    by_str = sorted(strs)
    by_rank = sorted(strs, key=lambda s: tie_rank_sequence(strs[s], tab))
    assert by_rank == by_str
