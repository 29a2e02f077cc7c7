"""K6b: the k-medoids iterations on the device (csrc/k6_pam.cu) against the reference's numpy loop (algo.py:184-211,
restated in Learner._pam_host) on the same float32 distance matrices and the same initial medoids: identical medoid indices."""
import types

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_pam(D, init, k):
    import torch
    from geobpe_b200 import _lib
    N = D.shape[0]
    Dd = torch.from_numpy(D).cuda().contiguous()
    mi = torch.from_numpy(np.asarray(init, dtype=np.int32)).cuda()
    out = torch.zeros(k + 4, dtype=torch.int32, device="cuda")
    _lib.call("gb_kmedoids_pam", _lib.ptr(Dd), N, k, _lib.ptr(mi), 10, _lib.ptr(out), out.data_ptr() + 4 * k)
    r = out.cpu().numpy()
    return [int(x) for x in r[:k]], int(r[k]), int(r[k + 1])


def _distance_matrix(rng, N, dim=6, clusters=4, dup=0):
    centres = rng.normal(size=(clusters, dim)) * 3
    x = centres[rng.integers(0, clusters, N)] + rng.normal(size=(N, dim))
    if dup:
        x[-dup:] = x[:dup]                                   # exact duplicates: zero distances off the diagonal
    D = np.sqrt(((x[:, None, :] - x[None, :, :]) ** 2).sum(-1)).astype(np.float32)
    D = np.minimum(D, D.T)
    np.fill_diagonal(D, 0)
    return D


@pytest.mark.parametrize("N,k", [(500, 2), (500, 5), (500, 1), (37, 3), (129, 8), (1000, 5), (2500, 16), (300, 64)])
def test_device_iterations_match_numpy_loop(N, k):
    from geobpe_b200 import engine as E, _lib
    _lib.require_cuda()
    rng = np.random.default_rng(N * 131 + k)
    for trial in range(4):
        D = _distance_matrix(rng, N, clusters=max(2, k - trial))
        init = rng.choice(np.arange(N), size=k, replace=False)
        stub = types.SimpleNamespace(rng=np.random.default_rng(0))
        want = E.Learner._pam_host(stub, D, init, k)
        got, status, iters = _device_pam(D, init, k)
        if status == 1:                                      # an empty cluster: the engine repeats on the host (RNG draw)
            continue
        assert got == want, (N, k, trial, iters)


def test_empty_cluster_is_reported_not_guessed():
    from geobpe_b200 import _lib
    _lib.require_cuda()
    rng = np.random.default_rng(5)
    D = _distance_matrix(rng, 200, dup=3)
    init = np.array([0, 197, 50])                            # samples 0 and 197 coincide: the second one's cluster is empty
    got, status, iters = _device_pam(D, init, 3)
    assert status == 1


def test_engine_k_medoids_device_equals_host_path(monkeypatch):
    """Learner.k_medoids end to end (distance matrix on the device either way): same medoids, same RNG position afterwards"""
    import torch
    from geobpe_b200 import engine as E, _lib
    _lib.require_cuda()
    rng = np.random.default_rng(11)
    coords = torch.from_numpy(rng.normal(size=(400, 7 * 3)) * 2).cuda()
    res = []
    for host in (False, True):
        if host:
            monkeypatch.setenv("GEOBPE_PAM_HOST", "1")
        stub = types.SimpleNamespace(rng=np.random.default_rng(3))
        stub._pam_host = types.MethodType(E.Learner._pam_host, stub)
        res.append((E.Learner.k_medoids(stub, coords, 5), stub.rng.integers(1 << 30)))
    assert res[0] == res[1]
