"""acx_bfs_mask_scan_static (include/acx.h): the mask scan of a bfs round with its scratch in a static device array,
for callers that cannot allocate.  Through the C-ABI, against acx_bfs_mask_scan with caller scratch: bit-exact (integer
work), on calls in a row that share the array.  Collected last on purpose
(the file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _scan_pair(L, _lib, torch, dev, mask):
    """(excl, total) of the existing entry point and of the static one on the same device mask"""
    n = mask.shape[0]
    stream = torch.cuda.current_stream(dev).cuda_stream
    ref_excl = torch.full((n,), -7, dtype=torch.int64, device=dev)
    ref_total = torch.full((1,), -7, dtype=torch.int64, device=dev)
    need = int(L.acx_bfs_mask_scan(0, n, 0, 0, 0, 0, stream))
    assert need > 0
    temp = torch.empty((need,), dtype=torch.uint8, device=dev)
    _lib.check(int(L.acx_bfs_mask_scan(mask.data_ptr(), n, ref_excl.data_ptr(), ref_total.data_ptr(), temp.data_ptr(), need, stream)),
               "acx_bfs_mask_scan")
    got_excl = torch.full((n,), -9, dtype=torch.int64, device=dev)
    got_total = torch.full((1,), -9, dtype=torch.int64, device=dev)
    _lib.check(L.acx_bfs_mask_scan_static(mask.data_ptr(), n, got_excl.data_ptr(), got_total.data_ptr(), stream),
               "acx_bfs_mask_scan_static")
    torch.cuda.synchronize(dev)
    return (ref_excl.cpu().numpy(), int(ref_total.item())), (got_excl.cpu().numpy(), int(got_total.item())), need


def test_mask_scan_static_matches_the_scratch_version(cuda):
    import torch
    from acx import _lib

    L = _lib.lib()
    rng = np.random.default_rng(77)
    # calls in a row on the one shared array: large, small (stale descriptors of the large call behind it), the
    # maximum, then the first again
    sizes = [300001, 1, 4097, _lib.MASK_SCAN_STATIC_MAX_N, 300001]
    masks = {}
    for n in sizes:
        if n not in masks:
            masks[n] = rng.integers(0, 1 << 12, n).astype(np.int32)
    for n in sizes:
        mask = torch.from_numpy(masks[n]).to(cuda)
        ref, got, need = _scan_pair(L, _lib, torch, cuda, mask)
        assert need <= _lib.MASK_SCAN_STATIC_BYTES
        assert got[1] == ref[1]
        assert np.array_equal(got[0], ref[0])
        # ... and both are the scan: popcounts summed on the host
        pc = np.array([bin(int(x)).count("1") for x in masks[n][:2000]], dtype=np.int64)
        assert np.array_equal(got[0][:2000], np.concatenate([[0], np.cumsum(pc)[:-1]]))
        bits = np.unpackbits(masks[n].view(np.uint8)).sum()
        assert got[1] == int(bits)


def test_mask_scan_static_rejects_sizes_beyond_the_array(cuda):
    import torch
    from acx import _lib

    L = _lib.lib()
    stream = torch.cuda.current_stream(cuda).cuda_stream
    one = torch.zeros((8,), dtype=torch.int64, device=cuda)
    rc = L.acx_bfs_mask_scan_static(one.data_ptr(), _lib.MASK_SCAN_STATIC_MAX_N + 1, one.data_ptr(), one.data_ptr(), stream)
    assert rc != 0
    msg = L.acx_last_error().decode()
    assert "beyond the static scratch" in msg and str(_lib.MASK_SCAN_STATIC_MAX_N) in msg
    with pytest.raises(_lib.AcxError):
        _lib.check(rc, "acx_bfs_mask_scan_static")
    assert L.acx_bfs_mask_scan_static(0, -1, 0, 0, stream) != 0
    assert L.acx_bfs_mask_scan_static(0, 0, 0, 0, stream) == 0  # an empty round is fine
