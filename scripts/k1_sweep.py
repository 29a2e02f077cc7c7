#!/usr/bin/env python
"""K1 variant sweep on a B200: times the exact-table kernel and every slot-sketch variant on the bench workload's
token arrays, at the config-2 size (2.5 M tokens, L2-resident) and tiled to the per-GPU shard of config 5 (46 M tokens).

    python scripts/k1_sweep.py [--mode H|M|both] > gpurun_out/k1_sweep.txt
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from bench import hparams  # noqa: E402
from geobpe_b200 import _lib, engine as E, synthetic  # noqa: E402


def timeit(fn, pre, reps=10, warm=2):
    ms = []
    for i in range(reps + warm):
        pre()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        if i >= warm:
            ms.append(e0.elapsed_time(e1))
    return float(np.mean(ms)), float(np.min(ms))


def tiled(ts, mode, T0, rep):
    pad = lambda t: torch.cat([t, torch.zeros(8, dtype=t.dtype, device=t.device)])
    if mode == "H":
        h = pad(ts.a["hsh"][:T0].repeat(rep)); g = pad(ts.a["glue"][:T0].repeat(rep)); l = pad(ts.a["len"][:T0].repeat(rep))
        l[T0 * rep:] = 1
        return (h, g, l)
    return (pad(ts.a["cls"][:T0].repeat(rep)), pad(ts.a["glue"][:T0].repeat(rep)))


def sweep(mode, steps):
    hp = hparams(mode, None)
    vals, offs = synthetic.generate(10000, 256, seed=0)
    L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **hp)
    for _ in range(steps):
        L.step()
    ts = L.ts
    T0 = ts.T
    bpt = 16 if mode == "H" else 8
    for rep in (1, max(1, 46_000_000 // T0)):
        arrs = tiled(ts, mode, T0, rep)
        T = T0 * rep
        print(f"\n=== mode {mode}  T = {T} tokens ({bpt * T / 1e6:.0f} MB algorithmic), distinct keys ~ {L.table.distinct}, "
              f"HBM floor at 6453 GB/s = {bpt * T / 6453e9 * 1e6:.1f} us ===", flush=True)
        # exact table (previous design)
        tab = E.PairTable(); tab.ensure(4 * max(L.table.distinct, 1024) + 4096)
        if mode == "H":
            old = lambda: _lib.call("gb_pair_count_h", *map(_lib.ptr, arrs), T, _lib.ptr(L.powtab), E.HASH_BASE, *tab.ptrs(), _lib.ptr(tab.overflow))
        else:
            old = lambda: _lib.call("gb_pair_count_m", *map(_lib.ptr, arrs), T, *tab.ptrs(), _lib.ptr(tab.overflow))
        avg, best = timeit(old, lambda: _lib.call("gb_table_clear", *tab.ptrs()))
        print(f"exact-table K1 (M={tab.M}): {avg * 1e3:9.1f} us  (min {best * 1e3:.1f})  {bpt * T / avg / 1e6:7.0f} GB/s", flush=True)
        for log2ns in (22, 24):
            slots = torch.zeros(1 << log2ns, dtype=torch.int32, device="cuda")
            ref = None
            hot = None

            def launch(variant, ctas):
                if mode == "H":
                    _lib.call("gb_slot_count_h", *map(_lib.ptr, arrs), T, _lib.ptr(L.powtab), E.HASH_BASE, _lib.ptr(slots), log2ns,
                              _lib.ptr(hot) if hot is not None else None, variant, ctas)
                else:
                    _lib.call("gb_slot_count_m", *map(_lib.ptr, arrs), T, _lib.ptr(slots), log2ns,
                              _lib.ptr(hot) if hot is not None else None, variant, ctas)
            # memset cost
            avg, best = timeit(lambda: slots.zero_(), lambda: None)
            print(f"  log2ns={log2ns}: memset {avg * 1e3:.1f} us", flush=True)
            for variant in (0, 8, 16, 24, 2, 10, 4, 12):
                cache = (variant >> 1) & 3
                if cache == 2:
                    if ref is None:
                        continue
                    k = 4096
                    top = torch.topk(ref.to(torch.int64) & 0xFFFFFFFF, k).indices.to(torch.int32)
                    hot = torch.full((8192,), -1, dtype=torch.int32, device="cuda")
                    hot[(top & 8191).long()] = top
                for ctas in ((2, 4) if cache == 0 else (2, 3)):
                    try:
                        avg, best = timeit(lambda: launch(variant, ctas), lambda: slots.zero_())
                    except RuntimeError as ex:
                        print("    variant", variant, "ctas", ctas, "failed:", ex)
                        continue
                    ok = ""
                    if not (variant & 16):
                        if ref is None:
                            ref = slots.clone()
                        ok = "ok" if torch.equal(ref, slots) else "MISMATCH"
                    name = ("agg " if variant & 1 else "    ") + ("dyn-cache " if cache == 1 else "hot-list  " if cache == 2 else "          ") + \
                           ("cs " if variant & 8 else "   ") + ("DRY" if variant & 16 else "   ")
                    print(f"    v={variant:2d} [{name}] ctas/SM={ctas}: {avg * 1e3:9.1f} us (min {best * 1e3:8.1f})  {bpt * T / avg / 1e6:7.0f} GB/s "
                          f"{bpt * T / avg / 1e6 / 6453 * 100:5.1f}%  {ok}", flush=True)
            srt = torch.sort(ref.to(torch.int64) & 0xFFFFFFFF, descending=True).values.cumsum(0)
            tot = int(srt[-1])
            print('    coverage of top-k slots:', {k: round(int(srt[k - 1]) / tot, 4) for k in (1, 4, 16, 64, 256, 1024, 4096, 8192)}, flush=True)
            # scan + candidate pass on the reference slots
            slots.copy_(ref)
            bmax = torch.zeros((1 << log2ns) // 4096, dtype=torch.int32, device="cuda")
            sout = torch.zeros(4, dtype=torch.int32, device="cuda"); cand = torch.zeros(4096, dtype=torch.int32, device="cuda")
            avg, best = timeit(lambda: _lib.call("gb_slots_top", _lib.ptr(slots), log2ns, 0, _lib.ptr(sout), _lib.ptr(cand), 4096, _lib.ptr(bmax)), lambda: None)
            so = sout.cpu().numpy()
            print(f"    slots_top: {avg * 1e3:.1f} us  c1={so[0]} n_cand={so[1]} nonzero={so[2]}", flush=True)
            small = E.PairTable(); small.ensure(8192)
            for cs in (0, 1):
                if mode == "H":
                    c2 = lambda: _lib.call("gb_cand_count_h", *map(_lib.ptr, arrs), T, _lib.ptr(L.powtab), E.HASH_BASE, log2ns, _lib.ptr(cand),
                                           sout.data_ptr() + 4, 4096, *small.ptrs(), _lib.ptr(small.overflow), cs)
                else:
                    c2 = lambda: _lib.call("gb_cand_count_m", *map(_lib.ptr, arrs), T, log2ns, _lib.ptr(cand), sout.data_ptr() + 4, 4096,
                                           *small.ptrs(), _lib.ptr(small.overflow), cs)
                avg, best = timeit(c2, lambda: _lib.call("gb_table_clear", *small.ptrs()))
                _lib.call("gb_table_argmax", *small.ptrs(), _lib.ptr(small.out4), _lib.ptr(small.tie_keys), _lib.ptr(small.tie_min), small.TIE_CAP)
                o4 = small.out4.cpu().numpy()
                print(f"    cand_count cs={cs}: {avg * 1e3:.1f} us  exact max={o4[0]} ties={o4[1]} keys in cand slots={o4[2]}  "
                      f"(exact-table max = {L.pop_log[-1][0] if rep == 1 else '-'} at the previous step)", flush=True)
            del slots
        del arrs
        torch.cuda.empty_cache()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="both")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    if a.mode in ("H", "both"):
        sweep("H", 50)
    if a.mode in ("M", "both"):
        sweep("M", 20)
