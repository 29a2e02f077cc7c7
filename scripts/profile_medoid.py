"""Where a medoid-mode merge iteration spends its time (host side): cProfile of Learner.step on BASELINE configs[1].
usage: python scripts/profile_medoid.py [steps] [out.txt]"""
import cProfile
import io
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from geobpe_b200 import engine as E, synthetic

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
hp = bench.hparams("M", None)
vals, offs = synthetic.generate(10000, 256, seed=0)
L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **hp)
for _ in range(5):
    L.step()
torch.cuda.synchronize()
pr = cProfile.Profile()
t0 = time.perf_counter()
pr.enable()
k0, calls = L.n_steps, 0
while L.n_steps - k0 < steps:
    L.step(); calls += 1
torch.cuda.synchronize()
pr.disable()
dt = time.perf_counter() - t0
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
txt = f"{steps} new-key steps ({calls} step calls incl. recurring) in {dt:.3f} s = {steps / dt:.1f} it/s (under cProfile)\n" + s.getvalue()
print(txt[:6000])
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(txt)
