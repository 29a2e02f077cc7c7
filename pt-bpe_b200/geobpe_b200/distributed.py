"""Multi-GPU plumbing: one process per GPU, backbones sharded by contiguous ranges, merges applied locally.

The only per-iteration exchange is the pair-count histogram (north_star): every rank folds its exact local table
into a key-less direct-mapped counter array `slots` (slot = hash(key) mod NS) that is summed with ONE allreduce
(NCCL over NVLink on GPUs, gloo in the CPU tests); the exact global argmax is then resolved on the few keys that fall
into the top slots (SURVEY 7.2 hard part 8):

    a key's global count <= the total of its slot, so once the best exact count among the keys of the top-L slots
    is >= the total of slot L+1, no key outside those slots can beat (or tie) it.

This module holds no kernels and runs unchanged on CPU tensors, which is how tests/test_dist_cpu.py covers it.
"""
import numpy as np
import torch
import torch.distributed as dist


class Comm:
    def __init__(self, group=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1

    def all_reduce(self, t, op="sum"):
        if self.world > 1:
            dist.all_reduce(t, op={"sum": dist.ReduceOp.SUM, "min": dist.ReduceOp.MIN, "max": dist.ReduceOp.MAX}[op],
                            group=self.group)
        return t

    def all_gather_into(self, out, t):
        """out [world * t.numel()] <- every rank's t (enqueued on the device, no host synchronisation)"""
        if self.world > 1:
            dist.all_gather_into_tensor(out, t, group=self.group)
        else:
            out.copy_(t)
        return out

    def all_gather_padded(self, arr, cap, dtype, device):
        """arr: 1-D numpy array (<= cap entries) -> list of per-rank numpy arrays"""
        n = len(arr)
        if n > cap:
            raise RuntimeError(f"{n} candidates exceed the gather capacity {cap}")
        buf = torch.zeros(cap + 1, dtype=dtype, device=device)
        buf[0] = n
        if n:
            buf[1:n + 1] = torch.as_tensor(arr, dtype=dtype, device=device)
        if self.world == 1:
            return [np.asarray(arr)]
        out = [torch.empty_like(buf) for _ in range(self.world)]
        dist.all_gather(out, buf, group=self.group)
        res = []
        for o in out:
            o = o.cpu().numpy()
            res.append(o[1:int(o[0]) + 1])
        return res

    def all_gather_object(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.group)
        return out


def global_top(comm, slots, candidates_fn, device, top0=8, cap=4096):
    """Exact global (max count, {key: count} of ALL keys tied at the max).

    slots          int32/int64 tensor [NS]: this rank's per-slot pair counts (summed in place across ranks here)
    candidates_fn  f(slot_ids: np.int64 array) -> (keys np.int64, counts np.int64) of this rank's keys in those slots
    """
    comm.all_reduce(slots, "sum")
    NS = slots.numel()
    L = min(top0, NS)
    while True:
        k = min(L + 1, NS)
        vals, idx = torch.topk(slots, k)
        vals, idx = vals.cpu().numpy().astype(np.int64), idx.cpu().numpy().astype(np.int64)
        if vals[0] == 0:
            return 0, {}
        sel = idx[:L]
        keys, counts = candidates_fn(sel)
        gk = comm.all_gather_padded(np.asarray(keys, dtype=np.int64), cap, torch.int64, device)
        gc = comm.all_gather_padded(np.asarray(counts, dtype=np.int64), cap, torch.int64, device)
        tot = {}
        for ks, cs in zip(gk, gc):
            for a, b in zip(ks.tolist(), cs.tolist()):
                tot[a] = tot.get(a, 0) + b
        best = max(tot.values()) if tot else 0
        bound = int(vals[L]) if k > L else 0          # total of the first slot NOT examined
        if best > bound or k <= L:
            return best, {a: c for a, c in tot.items() if c == best}
        L = min(L * 4, NS)                            # a key outside the examined slots could still tie or win


def slot_of(keys, NS):
    """slot of a key in the direct-mapped counter array -- same formula as k_table_to_slots (k1_paircount.cu)"""
    x = np.asarray(keys, dtype=np.uint64) ^ np.uint64(0x9E3779B97F4A7C15)
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(33); x *= np.uint64(0xff51afd7ed558ccd)
        x ^= x >> np.uint64(33); x *= np.uint64(0xc4ceb9fe1a85ec53)
        x ^= x >> np.uint64(33)
    return (x & np.uint64(NS - 1)).astype(np.int64)


def shard_ranges(lens, world):
    """contiguous backbone ranges balanced by residue count: [(start, end)] * world"""
    lens = np.asarray(lens, dtype=np.int64)
    cum = np.concatenate([[0], np.cumsum(lens)])
    total = cum[-1]
    cuts = [0]
    for r in range(1, world):
        cuts.append(int(np.searchsorted(cum, total * r / world)))
    cuts.append(len(lens))
    for i in range(1, len(cuts)):
        cuts[i] = max(cuts[i], cuts[i - 1])
    return [(cuts[r], cuts[r + 1]) for r in range(world)]
