import os, sys, cProfile, pstats, io
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from bench import hparams
from geobpe_b200 import engine as E, synthetic
vals, offs = synthetic.generate(10000, 256, seed=0)
L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **hparams("M", None))
for _ in range(20): L.step()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(200): L.step()
torch.cuda.synchronize()
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45); print(s.getvalue())
