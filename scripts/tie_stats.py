#!/usr/bin/env python
"""How often is the argmax a tie (host tie-break needed)?  python scripts/tie_stats.py [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from bench import hparams
from geobpe_b200 import engine as E, synthetic
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
vals, offs = synthetic.generate(10000, 256, seed=0)
L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **hparams("H", None))
nts = []
orig = L._count_pairs
def wrapped():
    mx, ties = orig()
    nts.append((mx, len(ties)))
    return mx, ties
L._count_pairs = wrapped
for _ in range(steps):
    L.step()
nt = np.array([n for _, n in nts])
for lo, hi in ((0, 500), (500, 1000), (1000, 2000), (2000, 5000)):
    s = nt[lo:hi]
    if len(s):
        print(f"steps {lo}-{hi}: tie fraction {np.mean(s > 1):.3f}  mean ties {s.mean():.2f} max {s.max()}  top count {nts[min(hi, len(nts)) - 1][0]}")
print("rounds", L.sketch.rounds)
