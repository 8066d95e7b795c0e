"""novum_hashgrid_build_static — the allocation-free grid build (scratch in a file-scope device array of the library) — against
the existing entry point novum_hashgrid_create + novum_hashgrid_build on the same photons, through the C ABI, twice in a row.
Bar: bit-exact (integer work): sorted hashes, sorted indices, cellStart everywhere, cellEnd wherever cellStart is set.
This file is collected last on purpose."""
import numpy as np
import pytest

import cudapathtracer_b200 as nb

pytestmark = pytest.mark.gpu


def _static_build(torch, p, max_photons, mn, r, table_size, stream=None):
    n = len(p)
    d = [torch.from_numpy(np.ascontiguousarray(p[:, k])).cuda() for k in range(3)]
    # outputs start poisoned so that stale or unwritten entries cannot pass by luck
    d_h = torch.full((max(n, 1),), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    d_i = d_h.clone()
    d_cs = torch.full((table_size,), 0x12345678, dtype=torch.int32, device="cuda")
    d_ce = torch.full((table_size,), 0x12345678, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    nb.hashgrid_build_static(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n, max_photons, mn, r, d_h.data_ptr(), d_i.data_ptr(),
                             d_cs.data_ptr(), d_ce.data_ptr(), stream.cuda_stream if stream is not None else None)
    if stream is not None:
        stream.synchronize()
    else:
        torch.cuda.synchronize()
    out = [t.cpu().numpy().view(np.uint32) for t in (d_i[:n], d_h[:n], d_cs, d_ce)]
    return out


def _compare(got, want):
    idx, hsh, cs, ce = got
    widx, whsh, wcs, wce = want
    assert np.array_equal(hsh, whsh)
    assert np.array_equal(idx, widx)
    assert np.array_equal(cs, wcs)
    used = wcs != 0xFFFFFFFF
    assert np.array_equal(ce[used], wce[used])


def test_static_build_equals_object_build_two_calls_in_a_row(rng):
    torch = pytest.importorskip("torch")
    max_photons = 300_000
    ref = nb.HashGrid(max_photons)
    # second call is smaller and uses another radius: anything the first call left in the shared scratch must not leak through
    for n, r, mn in [(300_000, 0.02, [-1.0, -1.0, -1.0]), (70_001, 0.05, [-1.5, -1.0, -2.0])]:
        p = (rng.random((n, 3)) * 2 - 1).astype(np.float32)
        want = ref.build(p[:, 0], p[:, 1], p[:, 2], mn, r)
        got = _static_build(torch, p, max_photons, mn, r, ref.table_size)
        _compare(got, want)
    ref.close()


def test_static_build_on_a_side_stream_small_and_empty(rng):
    torch = pytest.importorskip("torch")
    s = torch.cuda.Stream()
    for max_photons, n in [(16, 1), (5000, 4097), (5000, 0)]:
        ref = nb.HashGrid(max_photons)
        p = (rng.random((n, 3)) * 3).astype(np.float32)
        want = ref.build(p[:, 0], p[:, 1], p[:, 2], [0.0, 0.0, 0.0], 0.07)
        got = _static_build(torch, p, max_photons, [0.0, 0.0, 0.0], 0.07, ref.table_size, stream=s)
        if n == 0:
            assert (got[2] == 0xFFFFFFFF).all()
        else:
            _compare(got, want)
        ref.close()


def test_static_build_rejects_sizes_beyond_its_limit():
    torch = pytest.importorskip("torch")
    d = torch.zeros(8, dtype=torch.int32, device="cuda")
    ptr = d.data_ptr()
    with pytest.raises(nb.NovumError, match="limit"):
        nb.hashgrid_build_static(ptr, ptr, ptr, 4, nb.STATIC_GRID_MAX_PHOTONS + 1, [0, 0, 0], 0.1, ptr, ptr, ptr, ptr)
    with pytest.raises(nb.NovumError):      # n > maxPhotons
        nb.hashgrid_build_static(ptr, ptr, ptr, 9, 8, [0, 0, 0], 0.1, ptr, ptr, ptr, ptr)
    with pytest.raises(nb.NovumError):      # radius must be positive and finite
        nb.hashgrid_build_static(ptr, ptr, ptr, 4, 8, [0, 0, 0], 0.0, ptr, ptr, ptr, ptr)


def test_static_build_at_its_maximum_size(rng):
    torch = pytest.importorskip("torch")
    n = nb.STATIC_GRID_MAX_PHOTONS
    ref = nb.HashGrid(n)
    p = rng.random((n, 3), dtype=np.float32)
    want = ref.build(p[:, 0], p[:, 1], p[:, 2], [0.0, 0.0, 0.0], 0.01)
    got = _static_build(torch, p, n, [0.0, 0.0, 0.0], 0.01, ref.table_size)
    _compare(got, want)
    ref.close()
