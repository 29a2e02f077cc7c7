"""gb_select_flagged_static: the ordered compaction with its scratch in a static device array (no allocation on the call).

Collected last on purpose (file name): the entry point has no counterpart in the reference.
Checked through the C ABI against gb_select_flagged (caller-provided scratch) -- index work, so bit-exact --
on calls in a row that share the static array."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(flags, n, mask):
    """(total, indices) from the static-scratch entry point and from the existing one, same inputs"""
    import torch
    from geobpe_b200 import _lib
    lib = _lib.load()
    dev = flags.device
    out_s = torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev)
    out_r = out_s.clone()
    tot_s = torch.full((1,), -1, dtype=torch.int32, device=dev)
    tot_r = tot_s.clone()
    scratch = torch.empty(int(lib.gb_select_scratch_elems(n)), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _lib.call("gb_select_flagged_static", _lib.ptr(flags), n, mask, _lib.ptr(out_s), _lib.ptr(tot_s))
    assert torch.cuda.memory_allocated() == before
    _lib.call("gb_select_flagged", _lib.ptr(flags), n, mask, _lib.ptr(out_r), _lib.ptr(tot_r), _lib.ptr(scratch))
    torch.cuda.synchronize()
    ts, tr = int(tot_s.item()), int(tot_r.item())
    return ts, out_s[:ts].cpu().numpy().astype(np.uint32), tr, out_r[:tr].cpu().numpy().astype(np.uint32)


def test_static_scratch_matches_caller_scratch_on_consecutive_calls():
    import torch
    from geobpe_b200 import _lib
    _lib.require_cuda()
    rng = np.random.default_rng(7)
    # large, then small (the array still holds the large call's block sums), then a ragged size, then empty, then large again
    for n, mask, density in ((3_000_001, 2, 0.3), (1000, 1, 0.5), (2048 * 37 + 5, 3, 0.01), (0, 1, 0.5), (1_500_000, 1, 0.9),
                             (1 << 26, 1, 0.001)):                          # the last one is the stated maximum
        host = (rng.random(max(n, 1)) < density).astype(np.uint8) * np.uint8(mask & 3)
        host |= (rng.integers(0, 2, max(n, 1), dtype=np.uint8) << 2)          # a bit outside every mask
        flags = torch.from_numpy(host).cuda()
        ts, idx_s, tr, idx_r = _both(flags, n, mask)
        want = np.nonzero(host[:n] & mask)[0].astype(np.uint32)
        assert tr == want.size and np.array_equal(idx_r, want)                 # the existing entry point, against numpy
        assert ts == tr and np.array_equal(idx_s, idx_r), f"n={n} mask={mask}"


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    import torch
    from geobpe_b200 import _lib
    _lib.require_cuda()
    lib = _lib.load()
    nmax = int(lib.gb_select_flagged_static_max_n())
    assert nmax == 1 << 26
    flags = torch.zeros(16, dtype=torch.uint8, device="cuda")
    out = torch.zeros(16, dtype=torch.int32, device="cuda")
    tot = torch.zeros(1, dtype=torch.int32, device="cuda")
    with pytest.raises(RuntimeError, match="exceeds the static scratch"):      # rejected before anything is launched
        _lib.call("gb_select_flagged_static", _lib.ptr(flags), nmax + 1, 1, _lib.ptr(out), _lib.ptr(tot))
    torch.cuda.synchronize()
