"""2-rank debug run of the sharded second loop (torchrun): prints every rank's exception with its traceback and exits hard, so that
a failure cannot leave a rank spinning.   torchrun --nproc-per-node 2 scripts/debug_dist2.py [steps] [uneven]"""
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
os.environ.setdefault("GEOBPE_PEER_TIMEOUT_MS", "4000")
import faulthandler
import numpy as np
import torch
import torch.distributed as dist

faulthandler.dump_traceback_later(45, repeat=False, exit=True)       # a stuck rank prints where it is and exits


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from geobpe_b200 import engine as E, synthetic
    from geobpe_b200.distributed import Comm, shard_ranges
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 80
    S, n = (40, 48) if len(sys.argv) < 4 else (int(sys.argv[2]), int(sys.argv[3]))
    vals, offs = synthetic.generate(S, n, seed=31, ragged=True)
    a, b = shard_ranges(np.diff(offs), world)[rank]
    hp = dict(bins={1: 5}, bin_strategy="histogram", std_bonds=True)
    L = E.Learner(vals[:, offs[a]:offs[b]], offs[a:b + 1] - offs[a], mode="H", comm=Comm(), **hp)
    Lu = E.Learner(vals, offs, mode="H", **hp)
    for k in range(steps):
        L.step(); Lu.step()
    if os.environ.get("GEOBPE_TRACE"):          # launch without waiting; watch block 0's progress marker from a second stream
        import time
        L._h2_flush(1)
        side = torch.cuda.Stream()
        last = None
        for _ in range(40):
            with torch.cuda.stream(side):
                h = L.hl.hs.to("cpu").numpy().astype(np.int64) & 0xFFFFFFFF
            cur = (int(h[55]), int(h[1]), int(h[2]), int(h[52]), int(h[53]), int(h[10]))
            if cur != last:
                print(f"[rank {rank}] trace {cur[0]:#x} step {cur[1]} halt {cur[2]} dseq {cur[3]} xseq {cur[4]} period_ok {cur[5]}", flush=True)
                last = cur
            time.sleep(0.25)
    L.sync()
    hl = L.hl
    print(f"[rank {rank}] device steps {hl.steps} halts {hl.halts} stats {hl.stats}", flush=True)
    ok = [list(p) for p in L.pop_log] == [list(p) for p in Lu.pop_log]
    ts, tu = L.ts, Lu.ts
    lo, hi = int(offs[a]), int(offs[b])
    pu = E._u32(tu.a["pos"][:tu.T]); sel = (pu >= lo) & (pu < hi)
    seg_u = np.stack([pu[sel] - lo, E._u32(tu.a["len"][:tu.T])[sel], E._u32(tu.a["tid"][:tu.T])[sel]])
    seg_s = np.stack([E._u32(ts.a[f][:ts.T]) for f in ("pos", "len", "tid")])
    ok2 = seg_u.shape == seg_s.shape and bool((seg_u == seg_s).all())
    print(f"[rank {rank}] pop log equal: {ok}; segmentation equal: {ok2}; phase us/it "
          f"{ {k: round(v / 1e3 / max(hl.steps, 1), 1) for k, v in hl.times_ns.items()} }", flush=True)
    if not ok:
        for i, (x, y) in enumerate(zip(L.pop_log, Lu.pop_log)):
            if list(x) != list(y):
                print(f"[rank {rank}] first difference at step {i}: {x[0]} vs {y[0]}", flush=True)
                break
    L.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    try:
        main()
    except BaseException:
        traceback.print_exc()
        sys.stderr.flush(); sys.stdout.flush()
        os._exit(1)
