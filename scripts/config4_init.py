#!/usr/bin/env python
"""BASELINE configs[3] ("pareto"): bins 500, num_p 2-100:3-500:5-20:6-100:8-5:9-20:11-1:12-5:14-1, RMSD-scored medoid
quantisation on 50 k synthetic backbones x 256 residues; README cmd 2 runs it with --max-iter 0, i.e. the work is
initialize(): K3 binning, sample 500 residues, K6 all-pairs matrix, K5 assignment of 12.8 M residues to 500 medoids.
    python scripts/config4_init.py [backbones] [merges]"""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
import numpy as np, torch
from geobpe_b200 import engine as E, synthetic
S = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
merges = int(sys.argv[2]) if len(sys.argv) > 2 else 20
NUM_P = {2: 100, 3: 500, 5: 20, 6: 100, 8: 5, 9: 20, 11: 1, 12: 5, 14: 1}
vals, offs = synthetic.generate(S, 256, seed=0)
torch.cuda.synchronize()
t0 = time.perf_counter()
L = E.Learner(vals, offs, bins={1: 500}, bin_strategy="histogram", std_bonds=False, mode="M", rmsd_super_res=True,
              num_partitions=NUM_P, max_num_strucs=500, rng=np.random.default_rng(0))
torch.cuda.synchronize()
t1 = time.perf_counter()
for _ in range(merges):
    L.step()
torch.cuda.synchronize()
t2 = time.perf_counter()
R = int(offs[-1])
print(json.dumps({"config": "BASELINE configs[3] pareto", "backbones": S, "residues": R, "initialize_s": t1 - t0,
                  "residues_per_s_initialize": R / (t1 - t0), "rmsd_evaluations": R * 500,
                  "rmsd_per_s": R * 500 / (t1 - t0), "tokens_after_init": len(L.tokens), "merges": merges,
                  "merge_s": t2 - t1, "merge_iters_per_s": merges / max(t2 - t1, 1e-9)}))
