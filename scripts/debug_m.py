import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pt-bpe_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np
import geobpe_oracle as O
from geobpe_b200 import synthetic
from geobpe_b200.bpe import BPE
from golden_util import load_golden, case_inputs, case_hparams, norm
g = load_golden(sys.argv[1] if len(sys.argv) > 1 else "m20")
vals, offs = case_inputs(g)
strucs = synthetic.to_structures(vals, offs)
hp = case_hparams(g)
ora = O.BPE(strucs, res_init=True, seed=0, **hp); ora.initialize(); ora.bin()
bpe = BPE(strucs, res_init=True, seed=0, **hp); bpe.initialize(); bpe.bin()
L = bpe._learner
for step in range(g["final"]["step"]):
    (ex, negc, key), _ = ora._priority_dict.peekitem(0)
    want = list(ora._geo_dict[key])
    mx, ties, rec = L.peek_m()
    ks, k = min((L.m_key_string(kk)[0], kk) for kk, *_ in ties)
    got = L.set_order(k)
    same_key = ks == key
    print("step", step, "count", -negc, "samekey", same_key, "setorder_equal", got == want, "sorted_equal", sorted(got) == sorted(want))
    if got != want:
        print(" want", want[:12]); print(" got ", got[:12])
    ora.step(); bpe.step()
    to, tb = norm(list(ora._tokens.items())), norm(list(bpe._tokens.items()))
    if to != tb:
        for a, b in zip(to, tb):
            if a != b:
                print(" first token diff", a[0], b[0]); print("  ora", a[1]); print("  gpu", b[1]); break
        break
