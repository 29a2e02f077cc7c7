"""diagnostic: where do the 2-GPU test workers spend their time / hang at exit?  python scripts/debug_dist_exit.py [peer|nccl|host]"""
import faulthandler
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def w(rank, world, port, q, mode):
    faulthandler.dump_traceback_later(45, exit=False, file=sys.stderr)
    t0 = time.time()
    import test_gpu_dist as T
    import torch.distributed as dist
    real_destroy = dist.destroy_process_group

    def timed_destroy(*a, **k):
        print(f"[{mode} r{rank}] before destroy {time.time() - t0:.1f}s", flush=True)
        real_destroy(*a, **k)
        print(f"[{mode} r{rank}] after destroy {time.time() - t0:.1f}s", flush=True)
    dist.destroy_process_group = timed_destroy
    T.worker(rank, world, port, q, mode)
    print(f"[{mode} r{rank}] worker returned {time.time() - t0:.1f}s", flush=True)
    faulthandler.cancel_dump_traceback_later()


if __name__ == "__main__":
    import torch.multiprocessing as mp
    mode = sys.argv[1] if len(sys.argv) > 1 else "peer"
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    t0 = time.time()
    ps = [ctx.Process(target=w, args=(r, 2, 29610 + hash(mode) % 50, q, mode)) for r in range(2)]
    [p.start() for p in ps]
    res = [q.get(timeout=200) for _ in ps]
    print(f"[{mode}] results in {time.time() - t0:.1f}s", flush=True)
    for p in ps:
        p.join(40)
        print(f"[{mode}] joined alive={p.is_alive()} {time.time() - t0:.1f}s", flush=True)
    for p in ps:
        if p.is_alive():
            p.kill()
