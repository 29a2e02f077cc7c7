"""world_size-2 gloo test (CPU) of the multi-GPU pair-count protocol: one allreduce of key-less slot counters, then an
exact resolution of the keys in the top slots -- result must equal the brute-force global argmax incl. all ties."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def make_local(rank, case):
    rng = np.random.default_rng(100 * case + rank)
    keys = rng.integers(1, 1 << 40, size=3000, dtype=np.int64)
    shared = np.random.default_rng(case).integers(1, 1 << 40, size=40, dtype=np.int64)   # keys both ranks hold
    keys = np.unique(np.concatenate([keys, shared]))
    counts = rng.integers(1, 50, size=len(keys))
    if case == 1:      # exact tie across ranks: two shared keys with the same global total
        for k, c in ((shared[0], 500 + rank), (shared[1], 501 - rank)):
            counts[np.searchsorted(keys, k)] = c
    if case == 2:      # tiny slot array: many collisions, the loop has to widen L
        pass
    return keys, counts


def worker(rank, world, port, q):
    sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
    from geobpe_b200.distributed import Comm, global_top, slot_of
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    comm = Comm()
    res = []
    for case, NS in ((0, 1 << 12), (1, 1 << 12), (2, 1 << 4)):
        keys, counts = make_local(rank, case)
        sl = slot_of(keys, NS)
        slots = torch.zeros(NS, dtype=torch.int64)
        slots.index_add_(0, torch.from_numpy(sl), torch.from_numpy(counts.astype(np.int64)))

        def cand(sel):
            m = np.isin(sl, sel)
            return keys[m], counts[m]
        best, ties = global_top(comm, slots, cand, "cpu", top0=4, cap=1 << 14)
        res.append((best, sorted(ties)))
    dist.destroy_process_group()
    q.put((rank, res))


def test_global_top_two_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29000 + os.getpid() % 2000
    ps = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    got = dict(q.get(timeout=120) for _ in ps)
    [p.join(30) for p in ps]
    assert got[0] == got[1]
    for case in range(3):
        tot = {}
        for r in range(2):
            k, c = make_local(r, case)
            for a, b in zip(k.tolist(), c.tolist()):
                tot[a] = tot.get(a, 0) + b
        best = max(tot.values())
        assert got[0][case] == (best, sorted(a for a, c in tot.items() if c == best))
    assert len(got[0][1][1]) == 2     # the constructed tie


def test_shard_ranges():
    sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
    from geobpe_b200.distributed import shard_ranges
    r = shard_ranges([10] * 8, 4)
    assert r == [(0, 2), (2, 4), (4, 6), (6, 8)]
    r = shard_ranges([1, 1, 1], 8)
    assert r[0][0] == 0 and r[-1][1] == 3 and all(a <= b for a, b in r)
