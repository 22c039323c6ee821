"""orbx_knn2_hamming_noalloc (no counterpart in the reference): the
brute-force 2-NN with its scratch in a static device array, caller-owned DEVICE buffers in and out, no allocation
inside the call.  Checked against the existing entry point orbx_knn2_hamming (itself checked against the oracle in
test_gpu_match.py) — bit-exact, as for every integer path here — on two calls in a row that share the scratch.

Collected last on purpose.  Passes alone in a fresh process and after the others.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _descs(rng, n, pool=None):
    """random 256-bit descriptors; with a pool, rows repeat so that equal distances (ties) are frequent"""
    if pool is None:
        return rng.integers(0, 256, (n, 32), dtype=np.uint8)
    return pool[rng.integers(0, len(pool), n)].copy()


def _noalloc(pkg, torch, q, t):
    """one call through the C ABI: device tensors in (torch only owns the caller's buffers), device results out"""
    L = pkg.binding.load()
    dev = torch.device("cuda:0")
    dq = torch.from_numpy(q).to(dev).contiguous()
    dt = torch.from_numpy(t).to(dev).contiguous() if len(t) else torch.zeros((1, 32), dtype=torch.uint8, device=dev)
    didx = torch.full((max(len(q), 1), 2), -7, dtype=torch.int32, device=dev)
    ddist = torch.full((max(len(q), 1), 2), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = L.orbx_knn2_hamming_noalloc(0, dq.data_ptr(), len(q), dt.data_ptr() if len(t) else None, len(t),
                                     didx.data_ptr(), ddist.data_ptr(), None)     # NULL = default stream
    assert rc == 0, L.orbx_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return didx.cpu().numpy()[:len(q)], ddist.cpu().numpy()[:len(q)]


def test_noalloc_knn2_equals_existing_entry_point_twice(pkg):
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    L = pkg.binding.load()
    assert L.orbx_knn2_hamming_noalloc(0, None, 0, None, 0, None, None, None) == 0      # set-up only
    rng = np.random.default_rng(7)
    pool = _descs(rng, 40)
    # call 1: C3-sized, many splits; call 2 right behind it: fewer queries and another train set, so the scratch still
    # holds call 1's partial pairs where call 2 does not write; call 3: a single short split, ties everywhere
    cases = [(_descs(rng, 1000), _descs(rng, 1000)),
             (_descs(rng, 333, pool), _descs(rng, 4097, pool)),
             (_descs(rng, 130, pool), _descs(rng, 1, pool))]
    for q, t in cases:
        idx_ref, dist_ref = pkg.knn2_hamming(q, t)
        idx, dist = _noalloc(pkg, torch, q, t)
        assert np.array_equal(idx, idx_ref)
        assert np.array_equal(dist, dist_ref)
    # the same inputs again give the same bytes (nothing depends on what the array held before)
    q, t = cases[0]
    a = _noalloc(pkg, torch, q, t)
    b = _noalloc(pkg, torch, q, t)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert np.array_equal(a[0], pkg.knn2_hamming(q, t)[0])


def test_noalloc_knn2_limits_and_errors(pkg):
    torch = pytest.importorskip("torch")
    L = pkg.binding.load()
    B = pkg.binding
    dev = torch.device("cuda:0")
    nmax = 16384                                    # ORBX_KNN2_NOALLOC_MAX_QUERIES (include/orbx.h)
    rng = np.random.default_rng(8)
    # exactly the maximum works (and equals the allocating entry point) ...
    q, t = _descs(rng, nmax), _descs(rng, 2500)
    idx, dist = _noalloc(pkg, torch, q, t)
    idx_ref, dist_ref = pkg.knn2_hamming(q, t)
    assert np.array_equal(idx, idx_ref) and np.array_equal(dist, dist_ref)
    # ... one more is refused with a clear error and nothing is launched
    dq = torch.zeros((nmax + 1, 32), dtype=torch.uint8, device=dev)
    out = torch.full((nmax + 1, 2), -7, dtype=torch.int32, device=dev)
    rc = L.orbx_knn2_hamming_noalloc(0, dq.data_ptr(), nmax + 1, dq.data_ptr(), 8, out.data_ptr(), out.data_ptr(), None)
    assert rc == B.ORBX_ERR_CAPACITY
    msg = L.orbx_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "16384" in msg and "exceed" in msg
    torch.cuda.synchronize()
    assert int((out != -7).sum()) == 0
    # bad arguments
    assert L.orbx_knn2_hamming_noalloc(0, None, 5, dq.data_ptr(), 8, out.data_ptr(), out.data_ptr(), None) == B.ORBX_ERR_ARG
    assert L.orbx_knn2_hamming_noalloc(0, dq.data_ptr() + 1, 5, dq.data_ptr(), 8, out.data_ptr(), out.data_ptr(), None) == B.ORBX_ERR_ARG
    assert L.orbx_knn2_hamming_noalloc(99, dq.data_ptr(), 5, dq.data_ptr(), 8, out.data_ptr(), out.data_ptr(), None) == B.ORBX_ERR_CUDA
    # no train rows: every neighbour is missing (-1), as the allocating entry point answers
    idx, dist = _noalloc(pkg, torch, _descs(rng, 9), np.empty((0, 32), np.uint8))
    assert (idx == -1).all() and (dist == -1).all()
