"""The float32 row sums of the k-medoids update (algo.py:203 `D[np.ix_(members, members)].sum(axis=1)`) decide which sample
becomes a medoid, so the device kernel (csrc/k6_pam.cu) restates numpy's pairwise add-reduce.  The same routine, compiled
for the host into the library (gb_pairwise_sum_host), is pinned here against numpy itself; no GPU needed."""
import ctypes as C

import numpy as np


def _host_sum(lib, v):
    v = np.ascontiguousarray(v, dtype=np.float32)
    return np.float32(lib.gb_pairwise_sum_host(v.ctypes.data_as(C.c_void_p), int(v.size)))


def test_pairwise_sum_equals_numpy_row_sum():
    from geobpe_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(0)
    sizes = list(range(1, 300)) + [300, 333, 400, 499, 500, 511, 512, 513, 777, 1000, 1024, 2049, 4097, 8192]
    for n in sizes:
        a = (rng.random((3, n)) * 10).astype(np.float32)
        a[1] *= np.float32(1e-3)                       # small next to large: rounding differs between summation orders
        a[1, ::7] = np.float32(1e4)
        want = a.sum(axis=1)
        for i in range(3):
            got = _host_sum(lib, a[i])
            assert got == want[i], (n, i, got, want[i])


def test_pairwise_sum_on_gathered_distance_rows():
    """exactly the reference's expression: a fancy-indexed square block, summed along axis 1"""
    from geobpe_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(1)
    D = (rng.random((500, 500)) * 5).astype(np.float32)
    D = np.minimum(D, D.T)
    np.fill_diagonal(D, 0)
    for m in (1, 2, 7, 8, 9, 127, 128, 129, 136, 255, 317, 500):
        members = np.sort(rng.choice(500, m, replace=False))
        want = D[np.ix_(members, members)].sum(axis=1)
        for k in range(0, m, max(1, m // 9)):
            assert _host_sum(lib, D[members[k], members]) == want[k], (m, k)
