#!/usr/bin/env python
"""bench.py -- GeoBPE learn hot path on B200 (BASELINE.json metric: BPE merge iterations / s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--mode H|M] [--impl reference]

A "step" is one merge iteration (K1 pair count + argmax, tie-break, K2 merge [+K4/K5/K6 in medoid mode]) over
the resident shard of synthetic backbones.  N=1 workload = BASELINE.json configs[1]: 10k backbones x 256
residues, histogram quantisation, 2,000 merges.  N>1: every rank holds its own 10k x 256 shard (weak scaling),
pair counts combined by an NCCL allreduce.  Prints ONE JSON line on rank 0.
"""
import argparse
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
    ap.add_argument("--cpu-backbones", type=int, default=None,
                    help="backbones of the CPU sample (default: 2500 for --impl reference, 320 for the cpu_baseline leg)")
    ap.add_argument("--induce-backbones", type=int, default=100000, help="held-out backbones x 512 residues (BASELINE configs[2])")
    ap.add_argument("--induce-residues", type=int, default=512)
    ap.add_argument("--induce-merges", type=int, default=2000)
    ap.add_argument("--no-induce", action="store_true")
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


def reference_arm(args):
    """--impl reference / cpu_baseline: the CPU oracle (port of the reference's algorithm; the reference itself is
    Python and cannot travel to the GPU box) on a BOUNDED sample of the same workload, host cores."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import geobpe_oracle as O
    from geobpe_b200 import synthetic
    S, n = args.cpu_backbones or (2500 if args.impl == "reference" else 320), args.residues
    vals, offs = synthetic.generate(S, n, seed=0)
    strucs = synthetic.to_structures(vals, offs)
    t0 = time.perf_counter()
    bpe = O.BPE(strucs, res_init=True, seed=0, **oracle_hparams(args.mode, args.bins))
    bpe.initialize()
    t1 = time.perf_counter()
    bpe.bin()
    t2 = time.perf_counter()
    k = 0
    nsteps = min(args.steps + args.warmup, args.cpu_steps)
    while k < nsteps:
        bpe.step()
        k += 1
    t3 = time.perf_counter()
    return {"value": k / (t3 - t0), "steps_only_value": k / (t3 - t2), "unit": "merge_iters/s", "cores": 1, "kind": "port",
            "sample": f"{S} backbones x {n} residues ({S * n} residues), mode {args.mode}, value = steps/(initialize+bin+steps): initialize {t1 - t0:.2f}s, "
                      f"bin {t2 - t1:.2f}s, {k} steps {t3 - t2:.3f}s (incremental; cost grows with occurrences/merge; "
                      f"bin() alone extrapolates linearly to {(t2 - t1) * args.backbones / S:.0f}s at {args.backbones} backbones)",
            "init_s": t1 - t0, "bin_s": t2 - t1, "steps_s": t3 - t2}


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
    return {"metric": "residues_per_s_encoded", "value": R * world / dt, "unit": "residues/s", "seconds": dt,
            "backbones": S * world, "residues": R * world, "residues_per_backbone": args.induce_residues,
            "vocabulary_merges": args.induce_merges, "vocabulary_tokens": len(tokens), "tokens_out_per_rank": int(ts.T),
            "rejected_backbones_rank0": int(ind.failed.sum()), "merges_fired_rank0": int(ind.n_steps),
            "h2d_bytes": int(hv.nbytes + ho.nbytes), "d2h_bytes": int(sum(o.numel() * o.element_size() for o in out)),
            "vocabulary_learn_seconds": t_learn,
            "note": "host arrays -> H2D -> K3 binning -> residue medoid assignment (K4/K5) -> merge replay (K2 + K4/K5 per vocabulary "
                    "merge) -> D2H of the segmentation; backbones whose glue angles leave the training range are rejected like the "
                    "reference's ValueError (bpe.py:1105-1108)"}


def main():
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        if rank != 0:
            return
        r = reference_arm(args)
        line = {"impl": "reference", "metric": "bpe_merge_iters_per_s", "value": r["value"], "unit": "merge_iters/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 / r["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u64" if args.mode == "H" else "f64", "data": "synthetic",
                "config": {"workload": f"BASELINE configs[1] bounded sample on host cores: {r['sample']}"},
                "cpu_baseline": r, "e2e": {"value": r["value"], "unit": "merge_iters/s", "h2d_bytes_per_step": 0,
                                           "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"          # keep stdout to the one JSON line (NCCL prints its version at VERSION)
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

    # ---- device-resident arm: inputs already in HBM, W warm-up + K timed merge iterations ----------------
    L = mk()
    for _ in range(args.warmup):
        L.step()
    L.sync()
    L.prepare()                                      # warm-up also captures the loop's CUDA graphs
    L.k1_events = None if L.async_steps else []      # device loop: K1 is timed in an instrumented continuation (below)
    T_start = L.ts.T
    lib = _lib.load()
    barrier()
    launches0 = _lib.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        e0.record()
        done = 0
        while done < args.steps:
            L.step()
            done += 1
        L.sync()            # device-driven loop: drain it (one D2H of the state block + step log) inside the timed region
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    T_end, top = L.ts.T, L.pop_log[-1][0]
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    launches = _lib.launches - launches0
    k1_ms = [a.elapsed_time(b) for a, b, _ in (L.k1_events or [])]
    k1_T = [t for _, _, t in (L.k1_events or [])]
    c2_ms, c2_T = [], []
    k1_note = "CUDA events around every K1 launch of the timed region"
    peer = L.hl is not None and L.hl.dist is not None and L.hl.dist.peer is not None
    if L.async_steps and world > 1 and not peer:
        # sharded loop, NCCL exchanges: the timed region replays CUDA graphs (kernels + NCCL); K1 is timed in the iterations that follow,
        # launched part by part with CUDA events around every part-0 call (slots memset + k_slot_count)
        n_inst = min(128, args.steps)
        T_inst = L.ts.T
        L.k1_events = []
        for _ in range(n_inst):
            L.step()
        L.sync()
        torch.cuda.synchronize()
        k1_ms = [a.elapsed_time(b) for a, b, _ in L.k1_events]
        nk = [e[2] for e in L.step_events[-n_inst:]]
        k1_T = (T_inst - np.concatenate([[0], np.cumsum(nk)[:-1]])).tolist()[:len(k1_ms)]
        L.k1_events = None
        k1_note = (f"timed region = CUDA-graph replays; K1 (slots memset + k_slot_count) timed with CUDA events in the {n_inst} "
                   f"iterations that follow it, launched part by part")
    elif L.async_steps:
        # the timed region replays CUDA graphs (no events inside); K1 is timed in an instrumented continuation of the
        # same run: the next iterations launched kernel by kernel with CUDA events around every k_slot_count launch
        import ctypes
        n_inst = max(L.hl.period, min(256, args.steps))       # at least one whole period: one recount launch is observed
        T_inst = L.ts.T
        lib.gb_h_loop_time_k1(1)
        for _ in range(n_inst):
            L.step()
        L.sync()
        torch.cuda.synchronize()
        buf = (ctypes.c_float * (n_inst + 64))()
        n = lib.gb_h_loop_k1_times(buf, len(buf))
        lib.gb_h_loop_time_k1(0)
        loop_ms = [float(buf[i]) for i in range(n) if buf[i] > 0]
        nk = [e[2] for e in L.step_events[-n_inst:]]
        # (every sync leaves the loop at the start of a period: the recounts are iterations 0, period, 2 x period, ...)
        loop_T = (T_inst - np.concatenate([[0], np.cumsum(nk)[:-1]])).tolist()[::L.hl.period][:len(loop_ms)]
        loop_ms = loop_ms[:len(loop_T)]
        k1_ms, k1_T = loop_ms + k1_ms, loop_T + k1_T      # + the iterations that fell back to the host path
        k1_note = (f"timed region = CUDA-graph replays (no events inside a graph); timed with CUDA events around every "
                   f"launch of the kernel in the {n_inst} iterations that follow it, launched kernel by kernel")
        # second continuation: the exact candidate pass (hl_cand_count), the counting kernel that runs EVERY iteration
        # (k_slot_count recounts the slot bounds once per compaction period)
        n2 = min(128, args.steps)
        T2 = L.ts.T
        lib.gb_h_loop_time_k1(2)
        for _ in range(n2):
            L.step()
        L.sync()
        torch.cuda.synchronize()
        n = lib.gb_h_loop_k1_times(buf, len(buf))
        lib.gb_h_loop_time_k1(0)
        c2_ms = [float(buf[i]) for i in range(n) if buf[i] > 0]
        nk = [e[2] for e in L.step_events[-n2:]]
        c2_T = (T2 - np.concatenate([[0], np.cumsum(nk)[:-1]])).tolist()[:len(c2_ms)]
    bpt = 16 if args.mode == "H" else 8
    k1_avg_ms = float(np.mean(k1_ms))
    k1_bytes = bpt * float(np.mean(k1_T))
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s"
    achieved = k1_bytes / (k1_avg_ms * 1e-3) / 1e9
    tr = {}
    tp = os.path.join(ROOT, "profiles", "r1_k1_traffic.json")
    if os.path.exists(tp) and args.mode == "H":      # dram__bytes_read + write per launch from the committed ncu --set full capture
        tr = json.load(open(tp))
    sk = getattr(L, "sketch", None)
    distinct = (int(L.hl.hs.cpu()[8]) if L.hl is not None else int(sk.sout.cpu()[2])) if sk is not None else L.table.distinct   # sketch: non-zero slots (lower bound)
    k1_name = ("k_slot_count" if sk is not None else "k_pair_count") + ("<KeysH>" if args.mode == "H" else "<KeysM>")
    sk_rounds = list(sk.rounds) if sk is not None else None
    loop_info = None
    if L.hl is not None:
        loop_info = {"device_iterations": L.hl.steps, "second_rounds": L.hl.r2, "halts_by_code": {L.HALT_NAMES[k]: v for k, v in L.hl.halts.items()},
                     "max_ties": getattr(L.hl, "ntmax", 0), "kernels_per_iteration": lib.gb_h_loop_kernels_per_step()}
        if L.hl.dist is not None:
            loop_info["exchange"] = ("peer memory (loads over NVLink inside the loop's kernels, CUDA IPC blocks)" if L.hl.dist.peer is not None
                                     else "NCCL (allreduce-sum slots, allgather candidates, allreduce-max tie strings)")
            loop_info["slots_log2"] = L.sketch.log2ns
            loop_info["kernels_per_iteration"] = 17 if L.hl.dist.peer is not None else 12
    pairs = T_end - L.bb.S
    k1_per_step = 1.0 / L.hl.period if (c2_ms and L.hl is not None) else 1.0
    k1_share = k1_avg_ms * args.steps * k1_per_step / ms
    roof = {"bound": "hbm", "kernel": k1_name, "resolve_rounds_1_2_fallback": sk_rounds, "device_loop": loop_info,
            "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
            "frac_of_nominal_8TBs": achieved / 8000.0,
            "traffic": (tr["in_loop"]["dram_bytes_per_token"] * float(np.mean(k1_T))) if tr else None,
            "traffic_source": tr.get("source") if tr else None, "avg_ms": k1_avg_ms, "algorithmic_bytes_per_launch": k1_bytes,
            "bytes_per_token": bpt, "launches_per_step": k1_per_step, "share_of_step": k1_share, "timing": k1_note}
    if c2_ms:
        # the dominant counting kernel of the step is the exact candidate pass; the recount kernel is reported beside it
        c2_avg, c2_bytes = float(np.mean(c2_ms)), bpt * float(np.mean(c2_T))
        a2 = c2_bytes / (c2_avg * 1e-3) / 1e9
        t2 = tr.get("in_loop_pass2") if tr else None
        roof = {"bound": "hbm", "kernel": "hl_cand_count<KeysHL> (exact pair counts of the candidate slots: one full pass over the "
                                          "token arrays every iteration)",
                "device_loop": loop_info, "achieved": a2, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": a2 / peak,
                "frac_of_nominal_8TBs": a2 / 8000.0,
                "traffic": (t2["dram_bytes_per_token"] * float(np.mean(c2_T))) if t2 else None,
                "traffic_source": tr.get("source") if t2 else None, "avg_ms": c2_avg, "algorithmic_bytes_per_launch": c2_bytes,
                "bytes_per_token": bpt, "launches_per_step": 1.0, "share_of_step": c2_avg * args.steps / ms, "timing": k1_note,
                "note": "token arrays of this config (41 MB) are L2-resident, so the rate is not DRAM-limited here; "
                        "roofline_at_scale times the same two kernels on 48 M tokens (> L2)",
                "recount": {k: roof[k] for k in ("kernel", "achieved", "frac", "traffic", "avg_ms", "algorithmic_bytes_per_launch",
                                                 "launches_per_step", "share_of_step")}}
    scale = None
    if not args.no_scale_probe and rank == 0:
        pr = k1_scale_probe(L, args.mode)
        a2 = pr["bytes_per_token"] * pr["tokens"] / (pr["ms"] * 1e-3) / 1e9
        scale = dict(pr, bound="hbm", achieved=a2, peak=peak, unit="GB/s", frac=a2 / peak, frac_of_nominal_8TBs=a2 / 8000.0,
                     traffic=(tr["at_scale"]["dram_bytes_per_token"] * pr["tokens"]) if tr else None,
                     pass2_achieved=pr["bytes_per_token"] * pr["tokens"] / (pr["pass2_ms"] * 1e-3) / 1e9,
                     pass2_frac=pr["bytes_per_token"] * pr["tokens"] / (pr["pass2_ms"] * 1e-3) / 1e9 / peak,
                     note="K1 pass 1 (k_slot_count) alone on this workload's token arrays tiled to the per-GPU shard of "
                          "BASELINE config 5 (48 M tokens > L2); pass2_* = k_cand_count (exact resolution) on the same arrays")
    if world > 1:
        L.close()                                      # frees the peer-memory block (collective)
    del L
    torch.cuda.empty_cache()

    # ---- e2e arm: host buffers -> H2D -> initialize (K3, residue tokens) -> K merges -> D2H token arrays ----
    e2e = None
    if not args.no_e2e:
        barrier()
        t0 = time.perf_counter()
        L2 = mk()
        for _ in range(args.steps):
            L2.step()
        ts = L2.ts
        n_vocab = len(L2.tokens)                       # formats the vocabulary entries / pop log of every iteration
        out = [ts.a[f][:ts.T].cpu() for f in ("pos", "len", "tid")]
        n_ev = L2.ev_n
        out += [L2.ev_left[:n_ev].cpu(), L2.ev_right[:n_ev].cpu()]
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        if world > 1:
            L2.close()
        d2h = sum(o.numel() * o.element_size() for o in out)
        e2e = {"value": args.steps / dt, "unit": "merge_iters/s", "h2d_bytes_per_step": L2.bb.h2d_bytes / args.steps,
               "d2h_bytes_per_step": d2h / args.steps, "seconds": dt,
               "note": "Learner(host arrays): H2D of 72 B/residue, K3 thresholds+binning, residue tokens, K merges, D2H of the step log + vocabulary entries formatted on the host, D2H of segmentation + merge events"}
        del L2
    induce = None
    if not args.no_induce:
        try:
            induce = induce_arm(args, rank, world, barrier)
        except Exception as ex:          # the learn line must survive an induce failure
            induce = {"error": f"{type(ex).__name__}: {ex}"}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = reference_arm(args)
    if rank != 0:
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
                   "pairs": pairs, "distinct_over_pairs": distinct / max(pairs, 1), "last_top_count": top,
                   "residues_total": R * world},
        "clocks": clk.summary(), "e2e": e2e, "gpu_launches": launches,
        "roofline": roof,
        "roofline_at_scale": scale, "induce": induce, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
