#!/usr/bin/env python
"""bench.py -- GeoBPE learn / induce hot path on B200 (BASELINE.json metric: BPE merge iterations / s, residues / s encoded).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--mode H|M] [--impl reference]

A "step" is one merge iteration (BPE.step, bpe.py:1792-2166: pick the most frequent adjacent motif pair with the
reference's tie-break, merge its occurrences left to right, update the pair counts) over the resident shard of synthetic
backbones.  N=1 workload = BASELINE.json configs[1]: 10k backbones x 256 residues, histogram quantisation, 2,000 merges.
N>1: every rank holds its own 10k x 256 shard (weak scaling); the sharded loop exchanges over peer memory / NCCL.
Prints ONE JSON line on rank 0.

  value        merge iterations / s, token arrays resident in HBM, W warm-up + K timed iterations ending with a drain of the
               device loop (state block + step log read back)
  e2e          the same through the public API from HOST arrays: H2D, thresholds + binning, residue tokens, K merges, D2H of
               step log / segmentation / merge events; median of --reps repetitions, setup and steps reported separately
  full_config  the configs[1] run itself (2,000 merges) whatever --steps says: device-resident rate and e2e
  roofline     the pair-count kernel (k_slot_count), CUDA events around live launches on this workload's token arrays
  loop_kernel  the persistent learn-loop kernel: device timers per phase, periods, candidate-pass bandwidth
  roofline_at_scale  pair-count + exact-candidate kernels on 46 M tokens (per-GPU shard of configs[4], larger than L2)
  medoid / induce / k5 / cpu_baseline / parity_check (N>1)   see DESIGN.md 6
"""
import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))

import numpy as np

INF = float("inf")
NUM_P = {2: 2, 3: 5, 5: 1, 6: 2, 8: 1}          # README cmd 1 --num-p (encode.sh:35)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="H", choices=["H", "M"])
    ap.add_argument("--backbones", type=int, default=10000)
    ap.add_argument("--residues", type=int, default=256)
    ap.add_argument("--bins", type=int, default=None)
    ap.add_argument("--reps", type=int, default=5, help="repetitions of the e2e / full_config legs (median reported)")
    ap.add_argument("--cpu-backbones", type=int, default=None,
                    help="backbones of the CPU sample (default: the whole workload for --impl reference, 320 for the cpu_baseline leg)")
    ap.add_argument("--induce-backbones", type=int, default=100000, help="held-out backbones x 512 residues (BASELINE configs[2])")
    ap.add_argument("--induce-residues", type=int, default=512)
    ap.add_argument("--induce-merges", type=int, default=2000)
    ap.add_argument("--medoid-steps", type=int, default=300)
    ap.add_argument("--no-induce", action="store_true")
    ap.add_argument("--medoid-sharded", action="store_true", help="N > 1: also time medoid mode on shards and self-check it (off by default)")
    ap.add_argument("--no-medoid", action="store_true")
    ap.add_argument("--cpu-steps", type=int, default=2000)
    ap.add_argument("--no-scale-probe", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def hparams(mode, bins):
    if mode == "H":
        return dict(bins={1: bins or 8}, bin_strategy="histogram", std_bonds=True, mode="H")
    return dict(bins={1: bins or 50}, bin_strategy="histogram", std_bonds=False, mode="M", rmsd_super_res=True,
                num_partitions=NUM_P, max_num_strucs=500)


def oracle_hparams(mode, bins):
    if mode == "H":
        return dict(bins={1: bins or 8}, bin_strategy="histogram", std_bonds=True, rmsd_partition_min_size=INF)
    return dict(bins={1: bins or 50}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0,
                rmsd_super_res=True, num_partitions=NUM_P, max_num_strucs=500)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.stop = [], False
        self.index = index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[3 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def reference_arm(args, S, mode=None, steps=None):
    """the CPU oracle (kind "port": restatement of the reference's algorithm; the reference itself is Python with absent
    dependencies and cannot travel to the GPU box) on S backbones of the same synthetic workload, one host core.
    value = steps-only rate (like the GPU arm's `value`), e2e_value = steps / (initialize + bin + steps)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import geobpe_oracle as O
    from geobpe_b200 import synthetic
    mode = mode or args.mode
    n = args.residues
    vals, offs = synthetic.generate(S, n, seed=0)
    strucs = synthetic.to_structures(vals, offs)
    t0 = time.perf_counter()
    bpe = O.BPE(strucs, res_init=True, seed=0, **oracle_hparams(mode, args.bins))
    bpe.initialize()
    t1 = time.perf_counter()
    bpe.bin()
    t2 = time.perf_counter()
    k = 0
    nsteps = min(steps if steps is not None else args.steps + args.warmup, args.cpu_steps)
    while k < nsteps:
        bpe.step()
        k += 1
    t3 = time.perf_counter()
    return {"value": k / (t3 - t2), "e2e_value": k / (t3 - t0), "unit": "merge_iters/s", "cores": 1, "kind": "port",
            "sample": f"{S} backbones x {n} residues ({S * n} residues), mode {mode}: initialize {t1 - t0:.2f}s, bin {t2 - t1:.2f}s, "
                      f"{k} steps {t3 - t2:.3f}s; value = steps-only rate, e2e_value = steps / (initialize + bin + steps)",
            "backbones": S, "same_config": S == args.backbones, "init_s": t1 - t0, "bin_s": t2 - t1, "steps_s": t3 - t2, "steps": k}


def k1_scale_probe(learner, mode, reps=20, target_tokens=48_000_000):
    """K1 alone at the per-GPU shard size of BASELINE config 5 (48 M live tokens): the token arrays of this
    workload tiled on the device (same key distribution), larger than L2 (126 MB).  Times pass 1 (k_slot_count, the
    pair-count kernel) and pass 2 (k_cand_count, exact resolution of the top slots) separately."""
    import torch
    from geobpe_b200 import _lib, engine as E
    ts = learner.ts
    T0 = ts.T
    rep = max(1, target_tokens // T0)
    T = T0 * rep
    pad = lambda t: torch.cat([t, torch.zeros(8, dtype=t.dtype, device=t.device)])
    sk = E.SlotSketch(T0)          # same slot count as the learner's own sketch: the tiled data has the same distinct keys
    sk.variant = 2                 # 48 M tokens: shared-memory hot stage on
    if mode == "H":
        h = pad(ts.a["hsh"][:T0].repeat(rep)); g = pad(ts.a["glue"][:T0].repeat(rep)); l = pad(ts.a["len"][:T0].repeat(rep))
        l[T:] = 1
        p1 = lambda: _lib.call("gb_slot_count_h", _lib.ptr(h), _lib.ptr(g), _lib.ptr(l), T, _lib.ptr(learner.powtab), learner.hash_base,
                               _lib.ptr(sk.slots), sk.log2ns, None, sk.variant | 8, sk.ctas)
        p2 = lambda: _lib.call("gb_cand_count_h", _lib.ptr(h), _lib.ptr(g), _lib.ptr(l), T, _lib.ptr(learner.powtab), learner.hash_base,
                               sk.log2ns, _lib.ptr(sk.cand), sk.sout.data_ptr() + 4, sk.CAP, *sk.small.ptrs(),
                               _lib.ptr(sk.small.overflow), 1)
        bpt = 16
    else:
        c = pad(ts.a["cls"][:T0].repeat(rep)); g = pad(ts.a["glue"][:T0].repeat(rep))
        p1 = lambda: _lib.call("gb_slot_count_m", _lib.ptr(c), _lib.ptr(g), T, _lib.ptr(sk.slots), sk.log2ns, None,
                               sk.variant | 8, sk.ctas)
        p2 = lambda: _lib.call("gb_cand_count_m", _lib.ptr(c), _lib.ptr(g), T, sk.log2ns, _lib.ptr(sk.cand),
                               sk.sout.data_ptr() + 4, sk.CAP, *sk.small.ptrs(), _lib.ptr(sk.small.overflow), 1)
        bpt = 8

    def timed(fn, pre):
        ms = []
        for i in range(reps + 3):
            pre()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            if i >= 3:
                ms.append(e0.elapsed_time(e1))
        return float(np.mean(ms)), float(np.min(ms))
    avg, best = timed(p1, lambda: sk.slots.zero_())
    _lib.call("gb_slots_top", _lib.ptr(sk.slots), sk.log2ns, 0, _lib.ptr(sk.sout), _lib.ptr(sk.cand), sk.CAP, _lib.ptr(sk.bmax))
    avg2, best2 = timed(p2, lambda: _lib.call("gb_table_clear", *sk.small.ptrs()))
    so = sk.sout.cpu().numpy()
    assert int(sk.small.overflow.item()) == 0
    return {"tokens": T, "bytes_per_token": bpt, "ms": avg, "ms_min": best, "pass2_ms": avg2, "pass2_ms_min": best2,
            "slots": 1 << sk.log2ns, "nonzero_slots": int(so[2]), "top_slot_count": int(so[0])}


def k1_live(L, reps=10):
    """the pair-count kernel on the learner's CURRENT token arrays: CUDA events around `reps` launches (the slot counters
    are restored afterwards: the loop keeps them exact incrementally)"""
    import torch
    from geobpe_b200 import _lib, engine as E
    ts, sk = L.ts, L.sketch
    keep = sk.slots.clone()
    cs = 1 if 16 * ts.T > E.SlotSketch.STREAM_BYTES else 0
    ms = []
    for i in range(reps + 2):
        sk.slots.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if L.mode == "H":
            _lib.call("gb_slot_count_h", ts.p("hsh"), ts.p("glue"), ts.p("len"), ts.T, _lib.ptr(L.powtab), L.hash_base,
                      _lib.ptr(sk.slots), sk.log2ns, None, sk.variant | (8 * cs), sk.ctas)
        else:
            _lib.call("gb_slot_count_m", ts.p("cls"), ts.p("glue"), ts.T, _lib.ptr(sk.slots), sk.log2ns, None, sk.variant | (8 * cs), sk.ctas)
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ms.append(e0.elapsed_time(e1))
    nz = int((sk.slots != 0).sum().item())
    sk.slots.copy_(keep)
    return float(np.mean(ms)), int(ts.T), nz


def k5_probe(Q=1 << 20, K=500, A=4, reps=5):
    """K5 medoid assignment at the BASELINE configs[3] shape (500 medoids, 4-atom residue windows): achieved FLOP/s with the
    per-RMSD flop count of SURVEY 8(d) (centroids 6m + covariance 18m + 3x3 singular values ~400 + residual 30m, m = A atoms)"""
    import torch
    from geobpe_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(0)
    P = torch.randn(Q, A * 3, dtype=torch.float64, device="cuda", generator=g)
    M = torch.randn(K, A * 3, dtype=torch.float64, device="cuda", generator=g)
    out = torch.empty(Q, dtype=torch.int32, device="cuda")
    ms = []
    for i in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.call("gb_kabsch_assign", _lib.ptr(P), Q, A, _lib.ptr(M), K, _lib.ptr(out), None)
        e1.record()
        torch.cuda.synchronize()
        if i:
            ms.append(e0.elapsed_time(e1))
    flop = 6 * A + 18 * A + 400 + 30 * A
    t = float(np.mean(ms)) * 1e-3
    return {"kernel": "k_kabsch_assign (float64, one thread per window, medoids in shared memory)", "windows": Q, "medoids": K, "atoms": A,
            "rmsds": Q * K, "ms": t * 1e3, "flop_per_rmsd": flop, "flop_formula": "6m + 18m + 400 + 30m, m = atoms per window",
            "achieved_tflops": Q * K * flop / t / 1e12, "rmsds_per_s": Q * K / t, "bound": "fp64 CUDA cores",
            "peak_fp64_tflops_nominal": 40.0, "frac_of_nominal_fp64": Q * K * flop / t / 1e12 / 40.0}


def medoid_arm(args, rank=0, world=1, comm=None, barrier=lambda: None):
    """BASELINE configs[1] in medoid mode (README setting: bins 50, num_p 2-2:3-5:5-1:6-2:8-1, max_num_strucs 500); at N > 1
    every rank holds its own 10,000 backbones (weak scaling) and the ranks share one RNG stream"""
    import torch
    from geobpe_b200 import engine as E, synthetic
    hp = hparams("M", None)
    vals, offs = synthetic.generate(args.backbones, args.residues, seed=rank)
    barrier()
    t0 = time.perf_counter()
    L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, comm=comm, **hp)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    for _ in range(5):
        L.step()
    torch.cuda.synchronize()
    barrier()
    t2 = time.perf_counter()
    k0 = L.n_steps
    while L.n_steps - k0 < args.medoid_steps:
        L.step()
    torch.cuda.synchronize()
    barrier()
    t3 = time.perf_counter()
    k = L.n_steps - k0
    dt = t3 - t2
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([dt], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return {"value": k / dt, "unit": "merge_iters/s", "steps": k, "initialize_s": t1 - t0, "steps_s": dt, "n_gpus": world,
            "config": f"{args.backbones} x {args.residues} per GPU, medoid quantisation, bins 50, num_p {NUM_P}, max_num_strucs 500",
            "vocabulary_tokens": len(L.tokens), "live_tokens_end_rank0": int(L.ts.T),
            "note": "host-driven step: K1, window coordinates, distance matrix, k-medoids iterations, assignment and merge on the device; "
                    "tie-break and the occurrence set's CPython iteration order on the host (numpy RNG stream). N > 1: global maximum "
                    "through the slot-counter allreduce, set histories and sampled windows gathered per new key"}


def parity_check_m(comm, rank, world, steps=20, per_rank=16, n=48):
    """N>1 self-check of sharded medoid mode against a one-GPU learner on the union (pop log, vocabulary, this rank's tokens)"""
    import hashlib
    from geobpe_b200 import engine as E, synthetic
    vals, offs = synthetic.generate(per_rank * world, n, seed=78, ragged=True)
    a, b = rank * per_rank, (rank + 1) * per_rank
    lo, hi = int(offs[a]), int(offs[b])
    hp = dict(bins={1: 10}, bin_strategy="histogram", std_bonds=False, mode="M", rmsd_super_res=True,
              num_partitions={2: 2, 3: 4, 5: 2, 6: 2, 8: 1}, max_num_strucs=50)
    Ls = E.Learner(np.ascontiguousarray(vals[:, lo:hi]), offs[a:b + 1] - offs[a], rng=np.random.default_rng(0), comm=comm, **hp)
    Lu = E.Learner(vals, offs, rng=np.random.default_rng(0), **hp)
    for _ in range(steps):
        Ls.step(); Lu.step()
    ok = [list(p) for p in Ls.pop_log] == [list(p) for p in Lu.pop_log]
    ok = ok and json.dumps([[list(k), v] for k, v in Ls.tokens.items()]) == json.dumps([[list(k), v] for k, v in Lu.tokens.items()])
    ts, tu = Ls.ts, Lu.ts
    pu, lu, tu_ = E._u32(tu.a["pos"][:tu.T]), E._u32(tu.a["len"][:tu.T]), E._u32(tu.a["tid"][:tu.T])
    sel = (pu >= lo) & (pu < hi)
    seg_u = np.stack([pu[sel] - lo, lu[sel], tu_[sel]])
    seg_s = np.stack([E._u32(ts.a["pos"][:ts.T]), E._u32(ts.a["len"][:ts.T]), E._u32(ts.a["tid"][:ts.T])])
    ok = ok and seg_u.shape == seg_s.shape and bool((seg_u == seg_s).all())
    digest = hashlib.sha256(json.dumps([list(p) for p in Lu.pop_log]).encode()).hexdigest()[:16]
    flags = comm.all_gather_object((bool(ok), digest))
    return {"ranks": world, "step_calls": steps, "ok": all(f for f, _ in flags) and len({d for _, d in flags}) == 1,
            "per_rank_ok": [f for f, _ in flags], "pop_log_sha256_16": digest,
            "what": f"medoid mode: {per_rank * world} ragged backbones x <= {n} residues sharded over {world} ranks vs one-GPU learner on "
                    f"the union: pop log, vocabulary (medoid geometries) and this rank's segmentation bit-exact"}


def induce_arm(args, rank, world, barrier):
    """BASELINE configs[2]: frozen-vocabulary encode (K7, BPE.tokenize / bin/induce.py) of held-out synthetic backbones
    against a medoid-mode vocabulary learned here on the configs[1] training set.  Host arrays in, token arrays out;
    the held-out set is split evenly over the ranks (no collective on this path)."""
    import torch
    from geobpe_b200 import engine as E, synthetic
    hp = hparams("M", None)
    vals, offs = synthetic.generate(args.backbones, args.residues, seed=0)
    t0 = time.perf_counter()
    L = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **hp)
    while L.n_steps < args.induce_merges:
        L.step()
    torch.cuda.synchronize()
    t_learn = time.perf_counter() - t0
    tokens, sphere, th = L.tokens, L.sphere_dict, L.th.thresholds
    del L
    torch.cuda.empty_cache()
    S = max(1, args.induce_backbones // world)
    hv, ho = synthetic.generate(S, args.induce_residues, seed=1000 + rank)
    R = int(ho[-1])
    barrier()
    t0 = time.perf_counter()
    ind = E.Inducer(hv, ho, tokens, sphere, th, hp["std_bonds"])
    ts = ind.ts
    out = [ts.a[f][:ts.T].cpu() for f in ("pos", "len", "tid")]
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([dt], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    res = {"metric": "residues_per_s_encoded", "value": R * world / dt, "unit": "residues/s", "seconds": dt,
           "backbones": S * world, "residues": R * world, "residues_per_backbone": args.induce_residues,
           "vocabulary_merges": args.induce_merges, "vocabulary_tokens": len(tokens), "tokens_out_per_rank": int(ts.T),
           "rejected_backbones_rank0": int(ind.failed.sum()), "merges_fired_rank0": int(ind.n_steps),
           "h2d_bytes": int(hv.nbytes + ho.nbytes), "d2h_bytes": int(sum(o.numel() * o.element_size() for o in out)),
           "vocabulary_learn_seconds": t_learn, "stages_s": getattr(ind, "stage_s", None),
           "note": "host arrays -> H2D -> K3 binning -> residue medoid assignment (K4/K5) -> merge replay (K2 + K4/K5 per vocabulary "
                   "merge) -> D2H of the segmentation; backbones whose glue angles leave the training range are rejected like the "
                   "reference's ValueError (bpe.py:1105-1108)"}
    if rank == 0 and world == 1 and not args.no_cpu:
        res["cpu_baseline"] = induce_cpu(args, tokens, sphere, th, hv, ho)
    return res


def induce_cpu(args, tokens, sphere, thresholds, hv, ho, nb=3):
    """BPE.tokenize of the first `nb` held-out backbones by the CPU oracle against the SAME vocabulary (one core)"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import geobpe_oracle as O
    from geobpe_b200 import synthetic
    try:
        tv, to = synthetic.generate(2, 16, seed=3)
        ob = O.BPE(synthetic.to_structures(tv, to), res_init=True, seed=0, **oracle_hparams("M", None))
        ob.initialize()
        ob._tokens, ob._sphere_dict = tokens, sphere
        for size, th in thresholds.items():
            ob._thresholds[size] = th
        strucs = synthetic.to_structures(hv[:, :int(ho[nb])], ho[:nb + 1])
        t0 = time.perf_counter()
        done = res = 0
        for s in strucs:
            tk = O.Tokenizer.from_structure(s)
            try:
                ob.tokenize(tk)
            except ValueError:
                pass
            done += 1
            res += len(s["angles"])
            if time.perf_counter() - t0 > 25:
                break
        dt = time.perf_counter() - t0
        return {"value": res / dt, "unit": "residues/s", "cores": 1, "kind": "port",
                "sample": f"{done} held-out backbones x {args.induce_residues} residues against the same {len(sphere)}-key vocabulary, {dt:.1f}s"}
    except Exception as ex:
        return {"error": f"{type(ex).__name__}: {ex}"}


def parity_check(comm, rank, world, steps=128, per_rank=48, n=64):
    """N>1 self-check (the driver's GPU tests run on one GPU): a small set sharded over the ranks through the SAME sharded
    loop the timed region used, against a one-GPU learner on the union, run on every rank: pop log + this rank's segmentation"""
    import hashlib
    from geobpe_b200 import engine as E, synthetic
    vals, offs = synthetic.generate(per_rank * world, n, seed=77, ragged=True)
    a, b = rank * per_rank, (rank + 1) * per_rank
    lo, hi = int(offs[a]), int(offs[b])
    hp = dict(bins={1: 6}, bin_strategy="histogram", std_bonds=True, mode="H")
    Ls = E.Learner(np.ascontiguousarray(vals[:, lo:hi]), offs[a:b + 1] - offs[a], rng=np.random.default_rng(0), verify=True, comm=comm, **hp)
    Lu = E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, **hp)
    for _ in range(steps):
        Ls.step(); Lu.step()
    ok = [list(p) for p in Ls.pop_log] == [list(p) for p in Lu.pop_log]
    ts, tu = Ls.ts, Lu.ts
    pu, lu, tu_ = E._u32(tu.a["pos"][:tu.T]), E._u32(tu.a["len"][:tu.T]), E._u32(tu.a["tid"][:tu.T])
    sel = (pu >= lo) & (pu < hi)
    seg_u = np.stack([pu[sel] - lo, lu[sel], tu_[sel]])
    seg_s = np.stack([E._u32(ts.a["pos"][:ts.T]), E._u32(ts.a["len"][:ts.T]), E._u32(ts.a["tid"][:ts.T])])
    ok = ok and seg_u.shape == seg_s.shape and bool((seg_u == seg_s).all())
    digest = hashlib.sha256(json.dumps([list(p) for p in Lu.pop_log]).encode()).hexdigest()[:16]
    flags = comm.all_gather_object((bool(ok), digest))
    dev_steps = Ls.hl.steps if Ls.hl is not None else 0
    Ls.close()
    return {"ranks": world, "steps": steps, "ok": all(f for f, _ in flags) and len({d for _, d in flags}) == 1,
            "per_rank_ok": [f for f, _ in flags], "pop_log_sha256_16": digest, "device_iterations_sharded": dev_steps,
            "what": f"{per_rank * world} ragged backbones x <= {n} residues sharded over {world} ranks vs one-GPU learner on the union: "
                    f"pop log (count, key string) and this rank's segmentation (pos, len, token id) bit-exact"}


def main():
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        if rank != 0:
            return
        r = reference_arm(args, args.cpu_backbones or args.backbones)
        line = {"impl": "reference", "metric": "bpe_merge_iters_per_s", "value": r["value"], "unit": "merge_iters/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 / r["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u64" if args.mode == "H" else "f64", "data": "synthetic",
                "config": {"workload": f"BASELINE configs[1] on host cores: {r['sample']}", "same_config": r["same_config"]},
                "cpu_baseline": r, "e2e": {"value": r["e2e_value"], "unit": "merge_iters/s", "h2d_bytes_per_step": 0,
                                           "d2h_bytes_per_step": 0, "setup_s": r["init_s"] + r["bin_s"], "steps_s": r["steps_s"]}}
        print(json.dumps(line))
        return
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")   # a driver-set level is kept (NCCL's own lines go to stderr below)
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from geobpe_b200 import _lib, engine as E, synthetic
    from geobpe_b200 import build as B
    B.build()
    hp = hparams(args.mode, args.bins)
    vals, offs = synthetic.generate(args.backbones, args.residues, seed=rank)
    R = int(offs[-1])
    from geobpe_b200.distributed import Comm
    comm = Comm() if world > 1 else None
    mk = lambda: E.Learner(vals, offs, rng=np.random.default_rng(0), verify=True, comm=comm, **hp)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def resident_run(L, nsteps):
        """nsteps iterations + drain, CUDA events on the launching stream, barrier + synchronize on both sides"""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(nsteps):
            L.step()
        L.sync()            # drain the device loop (one D2H of the state block + step log) inside the timed region
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    def e2e_run(nsteps):
        """host arrays -> Learner -> nsteps merges -> D2H of step log, vocabulary, segmentation, merge events"""
        gc.collect()                                   # the previous repetition's garbage is not this one's cost
        barrier()
        t0 = time.perf_counter()
        L2 = mk()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        for _ in range(nsteps):
            L2.step()
        ts = L2.ts
        n_vocab = len(L2.tokens)                       # formats the vocabulary entries / pop log of every iteration
        out = [ts.a[f][:ts.T].cpu() for f in ("pos", "len", "tid")]
        n_ev = L2.ev_n
        out += [L2.ev_left[:n_ev].cpu(), L2.ev_right[:n_ev].cpu()]
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        d2h = sum(o.numel() * o.element_size() for o in out)
        h2d = L2.bb.h2d_bytes
        if world > 1:
            L2.close()
        del L2
        return max_over_ranks(t1 - t0), max_over_ranks(t2 - t1), h2d, d2h, n_vocab

    def e2e_median(nsteps, reps):
        e2e_run(min(nsteps, 64))                       # untimed: allocator growth, lazily loaded kernels
        rows = [e2e_run(nsteps) for _ in range(reps)]
        tot = sorted(r[0] + r[1] for r in rows)
        med = tot[len(tot) // 2]
        row = min(rows, key=lambda r: abs(r[0] + r[1] - med))
        return {"value": nsteps / med, "unit": "merge_iters/s", "h2d_bytes_per_step": row[2] / nsteps, "d2h_bytes_per_step": row[3] / nsteps,
                "seconds": med, "setup_s": row[0], "steps_s": row[1], "reps": reps, "values_all_reps": [nsteps / t for t in tot],
                "setup_steps_s_by_rep": [[round(r[0], 5), round(r[1], 5)] for r in rows],
                "spread": (tot[-1] - tot[0]) / med,
                "note": "Learner(host arrays): H2D of 72 B/residue, K3 thresholds + binning, residue tokens (setup_s); K merges, D2H of the "
                        "step log, vocabulary entries formatted on the host, D2H of segmentation + merge events (steps_s); median repetition"}

    # ---- device-resident arm: inputs already in HBM, W warm-up + K timed merge iterations ----------------
    L = mk()
    for _ in range(args.warmup):
        L.step()
    L.sync()
    L.prepare()                                      # (v1 loop: captures the CUDA graphs)
    T_start = L.ts.T
    launches0 = _lib.launches
    with ClockSampler(local) as clk:
        ms = resident_run(L, args.steps)
        T_end = L.ts.T
        launches = _lib.launches - launches0
        hl = L.hl
        snap = None
        if hl is not None and getattr(hl, "v2", False):     # loop statistics of warm-up + timed region (before the filler below)
            snap = (hl.steps, dict(hl.stats), dict(hl.times_ns), dict(hl.halts), hl.ntmax, hl.M)
        if ms < 600:                                 # keep the GPU busy until nvidia-smi has sampled the clocks a few times
            t_end = time.time() + 0.9
            Lb = mk() if world == 1 else None
            while Lb is not None and time.time() < t_end:
                if Lb.n_steps >= 1500:
                    Lb = mk()
                for _ in range(100):
                    Lb.step()
                Lb.sync()
            del Lb
    top = L.pop_log[-1][0]
    v2 = hl is not None and getattr(hl, "v2", False)
    loop_info = None
    if hl is not None:
        loop_info = {"design": "persistent cooperative kernel, incremental pair counts (k_hloop2.cu)" if v2 else
                     "per-iteration candidate pass, CUDA-graph replays (k_hloop.cu)",
                     "device_iterations": hl.steps, "halts_by_code": {L.H2_HALT_NAMES.get(k, k): v for k, v in hl.halts.items()},
                     "max_ties": getattr(hl, "ntmax", 0)}
        if v2:
            steps_, stats_, times_, halts_, ntmax_, M_ = snap
            n_it = max(steps_, 1)
            loop_info.update(device_iterations=steps_, halts_by_code={L.H2_HALT_NAMES.get(k, k): v for k, v in halts_.items()}, max_ties=ntmax_,
                             stats=stats_, phase_us_per_iteration={k: v / 1e3 / n_it for k, v in times_.items()},
                             table_slots=M_, slots_log2=L.sketch.log2ns, ctas_per_sm=hl.ctas)
            cp_ns, periods = times_["candidate_pass"], max(stats_["periods"], 1)
            if cp_ns:
                loop_info["candidate_pass"] = {"avg_us": cp_ns / 1e3 / periods, "per_iteration": periods / n_it,
                                               "achieved_gbs": 16.0 * L.ts.T / (cp_ns / periods), "bytes_per_token": 16}
            loop_info["equivalent_recount_gbs"] = 16.0 * 0.5 * (T_start + T_end) * args.steps / (ms * 1e6)
            loop_info["equivalent_note"] = ("what a design that re-reads 16 B per live token every iteration would have to stream to run "
                                            "these iterations in this time; the incremental design reads only the pair list and the occurrences")
        elif hl.dist is not None:
            loop_info["exchange"] = ("peer memory (loads over NVLink inside the loop's kernels, CUDA IPC blocks)" if hl.dist.peer is not None
                                     else "NCCL (allreduce-sum slots, allgather candidates, allreduce-max tie strings)")
            loop_info["slots_log2"] = L.sketch.log2ns

    # ---- roofline: the pair-count kernel on this workload's live token arrays ---------------------------------------
    bpt = 16 if args.mode == "H" else 8
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s"
    tr = {}
    tp = os.path.join(ROOT, "profiles", "r2_k1_traffic.json")
    if os.path.exists(tp) and args.mode == "H":      # dram__bytes_read + write per launch from the committed ncu --set full capture
        tr = json.load(open(tp))
    roof = None
    distinct = 0
    if L.sketch is not None:
        k1_ms, k1_T, distinct = k1_live(L)
        ach = bpt * k1_T / (k1_ms * 1e-3) / 1e9
        launches_per_step = (1.0 / max(snap[0], 1)) if v2 else (1.0 / hl.period if hl is not None else 1.0)
        roof = {"bound": "hbm", "kernel": "k_slot_count<KeysH> (pair-count kernel: every adjacent pair hashed into the slot counters)",
                "achieved": ach, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": ach / peak, "frac_of_nominal_8TBs": ach / 8000.0,
                "bytes_per_token": bpt, "survey_bytes_per_token": 8, "frac_at_survey_bytes": ach / peak * 8 / bpt,
                "traffic": (tr["in_loop"]["dram_bytes_per_token"] * k1_T) if tr else None, "traffic_source": tr.get("source") if tr else None,
                "avg_ms": k1_ms, "algorithmic_bytes_per_launch": bpt * k1_T, "launches_per_step": launches_per_step,
                "share_of_step": k1_ms * launches_per_step / (ms / args.steps),
                "timing": ("CUDA events around 10 launches on the live token arrays after the timed region (the loop runs it once, at bin(); "
                           "the pair counts are then maintained incrementally)") if v2 else "CUDA events around 10 launches on the live token arrays",
                "note": f"token arrays of this config ({bpt * k1_T / 1e6:.0f} MB) are L2-resident: not a DRAM number; see roofline_at_scale",
                "loop_kernel": loop_info}
    scale = None
    if not args.no_scale_probe and rank == 0 and L.sketch is not None:
        pr = k1_scale_probe(L, args.mode)
        a2 = pr["bytes_per_token"] * pr["tokens"] / (pr["ms"] * 1e-3) / 1e9
        p2 = pr["bytes_per_token"] * pr["tokens"] / (pr["pass2_ms"] * 1e-3) / 1e9
        scale = dict(pr, bound="hbm", achieved=a2, peak=peak, unit="GB/s", frac=a2 / peak, frac_of_nominal_8TBs=a2 / 8000.0,
                     survey_bytes_per_token=8, frac_at_survey_bytes=a2 / peak * 8 / pr["bytes_per_token"],
                     traffic=(tr["at_scale"]["dram_bytes_per_token"] * pr["tokens"]) if tr else None,
                     pass2_achieved=p2, pass2_frac=p2 / peak,
                     note="K1 pass 1 (k_slot_count) alone on this workload's token arrays tiled to the per-GPU shard of "
                          "BASELINE config 5 (48 M tokens > L2); pass2_* = k_cand_count (exact resolution) on the same arrays")
    pairs = T_end - L.bb.S
    par = None
    if world > 1:
        L.close()                                      # frees the peer-memory block (collective)
    del L
    torch.cuda.empty_cache()
    if world > 1:
        try:
            par = parity_check(comm, rank, world)
        except Exception as ex:
            par = {"ranks": world, "ok": False, "error": f"{type(ex).__name__}: {ex}"}

    # ---- e2e + the full configs[1] run -------------------------------------------------------------------------------
    e2e = full = None
    if not args.no_e2e:
        e2e = e2e_median(args.steps, args.reps)
        if args.mode == "H":
            if args.steps == 2000:
                full = {"steps": 2000, "value": args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps, "e2e": e2e, "same_run_as_value": True}
            else:
                Lf = mk()
                for _ in range(args.warmup):
                    Lf.step()
                Lf.sync()
                msf = resident_run(Lf, 2000)
                if world > 1:
                    Lf.close()
                del Lf
                full = {"steps": 2000, "value": 2000 / (msf * 1e-3), "ms_per_step": msf / 2000, "e2e": e2e_median(2000, args.reps),
                        "same_run_as_value": False}
            full["config"] = f"BASELINE configs[1] in full: {args.backbones} x {args.residues}, histogram quantisation, 2,000 merges"
    medoid = induce = k5 = None
    if world > 1 and args.medoid_sharded and args.mode == "H":   # opt-in: an optional leg must not be able to stall the N > 1 line
        try:
            medoid = medoid_arm(args, rank, world, comm, barrier)
            medoid["parity_check"] = parity_check_m(comm, rank, world)
        except Exception as ex:
            medoid = {"error": f"{type(ex).__name__}: {ex}"}
    if rank == 0 and world == 1 and not args.no_medoid and args.mode == "H":
        try:
            medoid = medoid_arm(args)
            if not args.no_cpu:
                medoid["cpu_baseline"] = reference_arm(args, 100, mode="M", steps=40)
        except Exception as ex:
            medoid = {"error": f"{type(ex).__name__}: {ex}"}
        try:
            k5 = k5_probe()
        except Exception as ex:
            k5 = {"error": f"{type(ex).__name__}: {ex}"}
    if not args.no_induce:
        try:
            induce = induce_arm(args, rank, world, barrier)
        except Exception as ex:          # the learn line must survive an induce failure
            induce = {"error": f"{type(ex).__name__}: {ex}"}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = reference_arm(args, args.cpu_backbones or 320)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": "bpe_merge_iters_per_s", "value": args.steps / (ms * 1e-3), "unit": "merge_iters/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64" if args.mode == "H" else "f64", "data": "synthetic",
        "config": {"workload": f"BASELINE configs[1]: {args.backbones} backbones x {args.residues} residues per GPU "
                               f"({R} residues), {'histogram' if args.mode == 'H' else 'medoid'} quantisation, bins {hp['bins'][1]}, "
                               f"{args.steps} merges after {args.warmup} warm-up merges",
                   "mode": args.mode, "l2": "inputs (token arrays 41 MB + table) smaller than L2 at this config; "
                                            "the at-scale probe uses 48 M tokens (768 MB) > L2",
                   "live_tokens_start": T_start, "live_tokens_end": T_end, "distinct_keys": distinct,
                   "distinct_keys_note": "non-zero slot counters after the timed region (lower bound of the distinct pair keys)",
                   "pairs": pairs, "distinct_over_pairs": distinct / max(pairs, 1), "last_top_count": top,
                   "residues_total": R * world},
        "clocks": clk.summary(), "e2e": e2e, "full_config": full, "gpu_launches": launches,
        "gpu_launches_note": "kernel launches issued through the C ABI inside the timed region (the persistent loop kernel runs hundreds of "
                             "iterations per launch)",
        "roofline": roof, "roofline_at_scale": scale, "medoid": medoid, "induce": induce, "k5": k5, "cpu_baseline": cpu,
    }
    if par is not None:
        line["parity_check"] = par
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
