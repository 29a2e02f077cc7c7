"""2-GPU NCCL test of sharded histogram-mode learning: backbones split across ranks, pair counts combined by the
slot allreduce + exact top resolution; pop order, vocabulary and per-backbone segmentation must equal the oracle run
on the whole set.  Needs 2 GPUs (skipped otherwise); run with `gpurun --gpus 2`."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu
HP = dict(bins={1: 5}, bin_strategy="histogram", std_bonds=True)
GEN = (40, 48, 31)
STEPS = 80          # crosses one compaction period of the device-driven loop


def worker(rank, world, port, q, mode, uneven=False):
    async_steps = mode != "host"
    os.environ["GEOBPE_PEER"] = "1" if mode == "peer" else "0"
    os.environ["GEOBPE_LOOP"] = "2" if mode == "loop2" else "1"
    os.environ.setdefault("GEOBPE_PEER_TIMEOUT_MS", "8000")
    if uneven:
        os.environ["GEOBPE_H2_VERIFY_INBOX"] = "1"      # (one variant also runs the inbox checksum verification)
    for p in (os.path.join(ROOT, "pt-bpe_b200"),):
        sys.path.insert(0, p)
    import torch.distributed as dist
    from geobpe_b200 import engine as E, synthetic
    from geobpe_b200.distributed import Comm, shard_ranges
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    vals, offs = synthetic.generate(GEN[0], GEN[1], seed=GEN[2], ragged=True)
    a, b = shard_ranges(np.diff(offs), world)[rank]
    if uneven:                       # rank 0 holds all but the last two backbones
        cut = len(offs) - 1 - 2
        a, b = (0, cut) if rank == 0 else (cut, len(offs) - 1)
    v = vals[:, offs[a]:offs[b]]
    o = offs[a:b + 1] - offs[a]
    L = E.Learner(v, o, mode="H", comm=Comm(), async_steps=async_steps, **HP)
    for _ in range(STEPS):
        L.step()
    ts = L.ts
    pos, ln, tid, glue = (E._u32(ts.a[f][:ts.T]) for f in ("pos", "len", "tid", "glue"))
    segs = []
    for i in range(b - a):
        lo, hi = np.searchsorted(pos, o[i]), np.searchsorted(pos, o[i + 1])
        segs.append([[int(3 * (pos[j] - o[i])), int(tid[j]), int(3 * ln[j] - (1 if glue[j] & E.LAST else 0))]
                     for j in range(lo, hi)])
    if L.hl is not None and getattr(L.hl, "v2", False):
        transport = "peer2" if L.hl.peer is not None else "none"
    else:
        transport = "peer" if (L.hl is not None and L.hl.dist is not None and L.hl.dist.peer is not None) else "nccl"
    halts = dict(L.hl.halts) if L.hl is not None else {}
    dev_steps = L.hl.steps if L.hl is not None else 0
    q.put((rank, a, [list(p) for p in L.pop_log], [[k, v] for k, v in L.tokens.items()], segs, transport, halts, dev_steps))
    L.close()           # graphs holding captured NCCL collectives / the peer-memory block go before the process group
    del L
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("mode,uneven", [("loop2", False), ("loop2", True), ("peer", False), ("nccl", False), ("host", False)],
                         ids=["loop2-replicated-counts-delta-exchange", "loop2-uneven-shards", "device-loop-peer-memory", "device-loop-nccl",
                              "host-driven"])
def test_two_gpu_h_mode_matches_oracle(mode, uneven):
    import torch.multiprocessing as mp
    import geobpe_oracle as O
    from geobpe_b200 import synthetic
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 1000
    ps = [ctx.Process(target=worker, args=(r, 2, port + {"peer": 0, "nccl": 7, "host": 14, "loop2": 21}[mode] + int(uneven), q, mode, uneven))
          for r in range(2)]
    [p.start() for p in ps]
    try:
        res = sorted(q.get(timeout=150) for _ in ps)
        [p.join(60) for p in ps]
    finally:                         # a failed rank must not leave the other one spinning on the box
        for p in ps:
            if p.is_alive():
                p.kill()
    vals, offs = synthetic.generate(GEN[0], GEN[1], seed=GEN[2], ragged=True)
    ora = O.BPE(synthetic.to_structures(vals, offs), res_init=True, seed=0, rmsd_partition_min_size=float("inf"), **HP)
    ora.initialize(); ora.bin()
    for _ in range(STEPS):
        ora.step()
    want_pops = [list(p) for p in ora.pop_log]
    want_tokens = [[k, v] for k, v in ora._tokens.items()]
    for rank, a, pops, tokens, segs, transport, halts, dev_steps in res:
        if mode == "peer":           # the exchanges really went through peer memory, and no rank timed out on a signal
            assert transport == "peer" and 9 not in halts
        if mode == "loop2":
            assert transport == "peer2" and 9 not in halts and dev_steps >= STEPS // 2
        assert pops == want_pops
        assert tokens == want_tokens
        for i, seg in enumerate(segs):
            assert seg == [list(x) for x in ora.tokenizers[a + i].bond_to_token.values()]


# ---- medoid mode on shards ---------------------------------------------------------------------------------------------
M_HP = dict(bins={1: 10}, bin_strategy="histogram", std_bonds=False, rmsd_super_res=True,
            num_partitions={2: 2, 3: 4, 5: 2, 6: 2, 8: 1}, max_num_strucs=50)
M_GEN = (30, 48, 21)
M_STEPS = 30


def _m_result(L, o, nb):
    """(pop log, vocabulary, per-backbone segmentation) of a medoid-mode learner holding backbones with local offsets o"""
    from geobpe_b200 import engine as E
    ts = L.ts
    pos, ln, tid, glue = (E._u32(ts.a[f][:ts.T]) for f in ("pos", "len", "tid", "glue"))
    keys = L.token_keys
    segs = []
    for i in range(nb):
        lo, hi = np.searchsorted(pos, o[i]), np.searchsorted(pos, o[i + 1])
        segs.append([[int(3 * (pos[j] - o[i])), list(keys[int(tid[j])]), int(3 * ln[j] - (1 if glue[j] & E.LAST else 0))]
                     for j in range(lo, hi)])
    return [list(p) for p in L.pop_log], [[list(k), v] for k, v in L.tokens.items()], segs


def _m_steps(L):
    for _ in range(M_STEPS):         # BPE.step: one merge, then every merge whose key already owns medoids (bpe.py:2164-2166)
        L.step()
        while L.peek_m()[2] is not None:
            L.step()


def m_worker(rank, world, port, q, uneven):
    sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
    import torch.distributed as dist
    from geobpe_b200 import engine as E, synthetic
    from geobpe_b200.distributed import Comm, shard_ranges
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    vals, offs = synthetic.generate(M_GEN[0], M_GEN[1], seed=M_GEN[2], ragged=True)
    a, b = shard_ranges(np.diff(offs), world)[rank]
    if uneven:                       # rank 1 holds the last backbone only: most merges have no occurrence there
        cut = len(offs) - 2
        a, b = (0, cut) if rank == 0 else (cut, len(offs) - 1)
    v = vals[:, offs[a]:offs[b]]
    o = offs[a:b + 1] - offs[a]
    L = E.Learner(v, o, mode="M", comm=Comm(), rng=np.random.default_rng(0), **M_HP)      # the same RNG stream on every rank
    _m_steps(L)
    q.put((rank, a) + _m_result(L, o, b - a))
    del L
    dist.destroy_process_group()


def _m_oracle():
    import geobpe_oracle as O
    from geobpe_b200 import synthetic
    vals, offs = synthetic.generate(M_GEN[0], M_GEN[1], seed=M_GEN[2], ragged=True)
    ora = O.BPE(synthetic.to_structures(vals, offs), res_init=True, seed=0, rmsd_partition_min_size=0, **M_HP)
    ora.initialize(); ora.bin()
    for _ in range(M_STEPS):
        ora.step()
    segs = [[[int(st), list(tok), int(ln)] for st, tok, ln in tk.bond_to_token.values()] for tk in ora.tokenizers]
    return [list(p) for p in ora.pop_log], [[list(k), v] for k, v in ora._tokens.items()], segs


def _check_m(res, want):
    want_pops, want_tokens, want_segs = want
    for rank, a, pops, tokens, segs in res:
        assert pops == want_pops
        assert tokens == want_tokens
        for i, seg in enumerate(segs):
            assert seg == want_segs[a + i], (rank, a + i)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("uneven", [False, True], ids=["balanced-shards", "one-backbone-shard"])
def test_two_gpu_m_mode_matches_oracle(uneven):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 1000 + 40 + int(uneven)
    ps = [ctx.Process(target=m_worker, args=(r, 2, port, q, uneven)) for r in range(2)]
    [p.start() for p in ps]
    try:
        res = sorted(q.get(timeout=150) for _ in ps)
        [p.join(60) for p in ps]
    finally:
        for p in ps:
            if p.is_alive():
                p.kill()
    _check_m(res, _m_oracle())


def test_m_mode_sharded_code_path_on_one_rank(monkeypatch):
    """the sharded medoid-mode path with a one-rank communicator (GEOBPE_FORCE_COMM): gathers are identities, everything
    else -- global top resolution, set-order gather, sample ownership -- runs as on a shard; result = the oracle's"""
    from geobpe_b200 import engine as E, synthetic
    from geobpe_b200.distributed import Comm
    monkeypatch.setenv("GEOBPE_FORCE_COMM", "1")
    vals, offs = synthetic.generate(M_GEN[0], M_GEN[1], seed=M_GEN[2], ragged=True)
    L = E.Learner(vals, offs, mode="M", comm=Comm(), rng=np.random.default_rng(0), **M_HP)
    assert L.comm is not None
    _m_steps(L)
    _check_m([(0, 0) + _m_result(L, offs, len(offs) - 1)], _m_oracle())
