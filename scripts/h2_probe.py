"""Quick timing of the device-driven learn loop (GEOBPE_LOOP=1|2) at BASELINE configs[1] size: merge iterations / s,
per-phase device timers of the persistent kernel, period statistics.   python scripts/h2_probe.py [steps] [backbones] [residues]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
import numpy as np
import torch
from geobpe_b200 import engine as E, synthetic

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
S = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
n = int(sys.argv[3]) if len(sys.argv) > 3 else 256
vals, offs = synthetic.generate(S, n, seed=0)
L = E.Learner(vals, offs, {1: 8}, "histogram", True, mode="H", rng=np.random.default_rng(0))
for _ in range(5):
    L.step()
L.sync()
L.prepare()
out = {"loop": L.loop_version, "tokens": int(L.ts.T)}
for rep, k in enumerate((steps, 20, steps)):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(k):
        L.step()
    L.sync()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    out[f"run{rep}_{k}"] = {"it_per_s": k / ms * 1e3, "us_per_it": ms / k * 1e3, "wall_ms": (time.perf_counter() - t0) * 1e3,
                            "T": int(L.ts.T), "top": L.pop_log[-1][0]}
hl = L.hl
if getattr(hl, "v2", False):
    out["stats"] = hl.stats
    out["halts"] = hl.halts
    out["device_steps"] = hl.steps
    out["phase_us_per_it"] = {k: v / 1e3 / max(hl.steps, 1) for k, v in hl.times_ns.items()}
    out["table_M"] = hl.M
    out["device_us_per_it"] = sum(v for k, v in hl.times_ns.items() if k in ("period_start", "decide", "flag", "deaths", "apply", "births", "compaction")) / 1e3 / max(hl.steps, 1)
print(json.dumps(out))
