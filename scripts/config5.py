"""BASELINE configs[4]: the scaling sweep workload itself -- 1 M synthetic backbones x 384 residues, histogram quantisation, 5 000
merges, backbones sharded over the ranks (strong scaling: the TOTAL is fixed).  One process per GPU (torchrun); prints one JSON
line on rank 0.    torchrun --nproc-per-node N scripts/config5.py [--backbones 1000000] [--residues 384] [--steps 5000]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
import numpy as np
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backbones", type=int, default=1_000_000)
    ap.add_argument("--residues", type=int, default=384)
    ap.add_argument("--steps", type=int, default=5000)
    ap.add_argument("--bins", type=int, default=8)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from geobpe_b200 import engine as E, synthetic
    from geobpe_b200.distributed import Comm
    S = args.backbones // world
    t0 = time.perf_counter()
    parts, chunk = [], 25000
    for c in range(0, S, chunk):                       # (chunked: the generator's temporaries stay small)
        v, _ = synthetic.generate(min(chunk, S - c), args.residues, seed=1000 * rank + c // chunk)
        parts.append(v)
    vals = np.ascontiguousarray(np.concatenate(parts, axis=1))
    del parts
    offs = np.arange(S + 1, dtype=np.int64) * args.residues
    t_gen = time.perf_counter() - t0
    comm = Comm() if world > 1 else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    L = E.Learner(vals, offs, {1: args.bins}, "histogram", True, mode="H", rng=np.random.default_rng(0), comm=comm)
    for _ in range(5):
        L.step()
    L.sync()
    barrier()
    t_setup = time.perf_counter() - t0
    T0 = L.ts.T
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        L.step()
    L.sync()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    tot = torch.tensor([float(L.ts.T), float(T0), float(L.bb.R)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    ts = L.ts
    pos, ln = E._u32(ts.a["pos"][:ts.T]), E._u32(ts.a["len"][:ts.T])
    ok = bool(ln.sum() == L.bb.R and np.all(pos[1:] == pos[:-1] + ln[:-1]))
    counts = [c for c, _ in L.pop_log]
    ok = ok and all(x >= y for x, y in zip(counts, counts[1:])) and len(counts) == args.steps + 5
    hl = L.hl
    out = {"workload": f"BASELINE configs[4]: {args.backbones} backbones x {args.residues} residues ({int(tot[2].item())} residues) over {world} GPU(s), "
                       f"histogram quantisation, bins {args.bins}, {args.steps} merges", "n_gpus": world,
           "merge_iters_per_s": args.steps / (float(ms.item()) * 1e-3), "ms_per_iteration": float(ms.item()) / args.steps,
           "setup_s_incl_h2d": t_setup, "host_generation_s": t_gen, "live_tokens_start": int(tot[1].item()), "live_tokens_end": int(tot[0].item()),
           "top_count_first_last": [counts[5], counts[-1]], "invariants_ok_rank0": ok,
           "device_iterations": hl.steps, "halts": hl.halts, "stats": hl.stats,
           "phase_us_per_iteration": {k: round(v / 1e3 / max(hl.steps, 1), 2) for k, v in hl.times_ns.items()},
           "hbm_gb_rank0": torch.cuda.max_memory_allocated() / 1e9}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        L.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
