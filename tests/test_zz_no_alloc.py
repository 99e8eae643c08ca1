"""nnb_chamfer_nn_static: the allocation-free variant of the Chamfer nearest-neighbour search (scratch in a file-scope
`static __device__` array of the library, include/nope_nerf_b200.h).  No counterpart in the reference.
Collected last on purpose (a failure here hides nothing after it).

Called through the C ABI with raw pointers; compared with the existing entry point (caller-owned workspace) on two calls in a
row of DIFFERENT sizes, so that the second call runs on scratch that still holds the first call's contents.  Indices must be
identical (the bar of the Chamfer tests: bit-exact), distances within the 1e-6 / 1e-7 tolerance those tests use."""
import ctypes as C

import pytest
import torch

from util import dev

LIMIT = 32768      # NNB_CHAMFER_STATIC_MAX_POINTS
ERR_TOO_LARGE = -5


def _lib():
    from model import _cabi as K
    return K, K.lib()


def test_static_entry_point_is_exported_and_bound():
    """CPU part: the symbol exists with the declared argument list (no compute call without a GPU)."""
    K, L = _lib()
    assert "nnb_chamfer_nn_static" in K.SIGNATURES and hasattr(L, "nnb_chamfer_nn_static")
    assert len(K.SIGNATURES["nnb_chamfer_nn_static"]) == len(K.SIGNATURES["nnb_chamfer_nn"]) - 2     # no workspace, no size
    assert b"capacity" in L.nnb_error_string(ERR_TOO_LARGE)


def _clouds(n, m, seed):
    g = torch.Generator().manual_seed(seed)
    src, dst = torch.rand(n, 3, generator=g) * 2 - 1, torch.rand(m, 3, generator=g) * 2 - 1
    src, dst = (src * 64).round() / 64, (dst * 64).round() / 64      # quantised: many exact distance ties (first index must win)
    return src.to(dev()).contiguous(), dst.to(dev()).contiguous()


def _call_static(L, K, src, dst):
    """-> rc, idx, dist, bytes of device memory that disappeared during the call (driver view: a cudaMalloc inside the library
    would show here, torch's caching allocator would not see it)."""
    n, m = src.shape[0], dst.shape[0]
    idx = torch.full((n,), -7, dtype=torch.int64, device=src.device)
    dist = torch.full((n,), -1.0, dtype=torch.float32, device=src.device)
    torch.cuda.synchronize()
    free0, reserved0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_reserved()
    rc = L.nnb_chamfer_nn_static(K.ptr(src), n, K.ptr(dst), m, K.ptr(idx), K.ptr(dist), K.stream())
    torch.cuda.synchronize()
    assert torch.cuda.memory_reserved() == reserved0
    return rc, idx, dist, free0 - torch.cuda.mem_get_info()[0]


def _call_existing(L, K, src, dst):
    n, m = src.shape[0], dst.shape[0]
    idx = torch.empty(n, dtype=torch.int64, device=src.device)
    dist = torch.empty(n, dtype=torch.float32, device=src.device)
    nb = int(L.nnb_chamfer_workspace_bytes(n, m))
    ws = torch.empty(nb, dtype=torch.uint8, device=src.device)
    K.check(L.nnb_chamfer_nn(K.ptr(src), n, K.ptr(dst), m, K.ptr(idx), K.ptr(dist), K.ptr(ws), nb, K.stream()), "nnb_chamfer_nn")
    return idx, dist


@pytest.mark.gpu
def test_static_matches_existing_entry_point_on_consecutive_calls():
    K, L = _lib()
    # two calls in a row, the second smaller than the first (its scratch is full of the first call's data), then a third at the
    # training size and one single-point problem
    for n, m, seed in ((20000, 30000, 1), (3001, 777, 2), (28830, 28830, 3), (1, 1, 4)):
        src, dst = _clouds(n, m, seed)
        want_idx, want_dist = _call_existing(L, K, src, dst)
        if n == 20000:
            _call_static(L, K, src[:8], dst[:8])     # first use on this device: module data and the symbol address are set up here
        rc, idx, dist, grabbed = _call_static(L, K, src, dst)
        assert rc == 0, L.nnb_error_string(rc)
        assert grabbed <= 0, "device memory shrank by %d bytes during an allocation-free call" % grabbed
        assert torch.equal(idx, want_idx), "indices differ at %d x %d" % (n, m)
        torch.testing.assert_close(dist, want_dist, rtol=1e-6, atol=1e-7)
        # and against the definition on a slice (fp32 arithmetic of model/losses.py:139-141)
        sub = slice(0, min(n, 512))
        d = torch.linalg.norm(src[sub, None, :] - dst[None, :, :], dim=2)
        assert torch.equal(d.argmin(dim=1), idx[sub])


@pytest.mark.gpu
def test_static_rejects_sizes_beyond_its_capacity():
    K, L = _lib()
    src, dst = _clouds(LIMIT + 1, 8, 5)
    rc, idx, _, _ = _call_static(L, K, src, dst)
    assert rc == ERR_TOO_LARGE and int(idx[0]) == -7          # refused before anything was launched
    rc, _, _, _ = _call_static(L, K, dst, src)
    assert rc == ERR_TOO_LARGE
    with pytest.raises(RuntimeError, match="capacity"):
        K.check(rc, "nnb_chamfer_nn_static")
    src, dst = _clouds(LIMIT, 64, 6)                            # exactly at the limit: accepted
    rc, idx, _, _ = _call_static(L, K, src, dst)
    assert rc == 0 and int(idx.min()) >= 0 and int(idx.max()) < 64
