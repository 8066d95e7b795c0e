"""Parity of the frame-finish kernel (through the C ABI) with the reference's own post-path code:
  * the reference's Image::saveImageBMP output (fixture from its host code) and cleanAndFormatImage output (fixture from
    its kernel on a B200), byte for byte;
  * the CPU restatement on seeded frames of awkward widths (row padding 0..3, segments of 1024 pixels) — bytes bit-exact,
    cleaned float frame bit-exact;
  * at 3840x2160: equality with the restatement on the whole frame, device-pointer entry point on a non-default stream,
    the written .bmp file read back."""
import os

import numpy as np
import pytest

import cudapathtracer_b200 as nb
from oracle import oracle_binding as ob

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _frame(rng, h, w, specials=True):
    acc = np.zeros((h, w, 4), np.float32)
    acc[..., :3] = (rng.random((h, w, 3)) ** 3 * 60).astype(np.float32)
    acc[..., 3] = 1.0
    ov = np.zeros_like(acc)
    if specials:
        n = h * w
        idx = rng.choice(n, min(n, max(4, n // 50)), replace=False)
        vals = np.array([np.nan, np.inf, -np.inf, -1.0, -1e-3, 0.0, 1e-30, 3e38], np.float32)
        acc.reshape(-1, 4)[idx, rng.integers(0, 3, idx.size)] = vals[rng.integers(0, vals.size, idx.size)]
        oi = rng.choice(n, min(n, max(2, n // 100)), replace=False)
        ov.reshape(-1, 4)[oi, :3] = rng.random((oi.size, 3)).astype(np.float32) * 1.5
    return acc, ov


def test_rows_equal_reference_saveImageBMP(built, tmp_path):
    g = np.load(os.path.join(GOLD, "frame_bmp.npz"))
    w, h = int(g["w"]), int(g["h"])
    for post, key in ((1, "bmp_post"), (0, "bmp_plain")):
        # sampleCount 1 in final mode divides by 1.0: the cleaned frame goes through unchanged
        rows = nb.frame_finish(g["frame"], None, 1, nb.FRAME_FINAL, post)
        assert np.array_equal(rows, g[key][54:].reshape(h, -1))
        p = str(tmp_path / f"o{post}.bmp")
        nb.save_bmp(p, g["frame"], None, 1, nb.FRAME_FINAL, post)
        assert np.array_equal(np.fromfile(p, np.uint8), g[key])          # the whole file, headers included


def test_clean_equals_reference_kernel(built):
    g = np.load(os.path.join(GOLD, "frame_clean.npz"))
    rows, fin = nb.frame_finish(g["acc"], g["overlay"], int(g["n"]), nb.FRAME_PROGRESSIVE, True, want_final=True)
    assert np.array_equal(fin.view(np.uint32), g["clean"].view(np.uint32))
    assert np.array_equal(rows, ob.bmp_rows(g["clean"], True))


@pytest.mark.parametrize("w,h", [(1, 1), (3, 5), (61, 37), (64, 64), (1023, 3), (1024, 2), (1025, 2), (1921, 4), (2050, 3)])
@pytest.mark.parametrize("mode", [nb.FRAME_PROGRESSIVE, nb.FRAME_FINAL])
@pytest.mark.parametrize("post", [True, False])
def test_kernel_equals_restatement(built, w, h, mode, post):
    rng = np.random.default_rng(w * 131 + h * 7 + mode * 3 + post)
    acc, ov = _frame(rng, h, w)
    n = 17
    for overlay in (ov, None):
        rows, fin = nb.frame_finish(acc, overlay, n, mode, post, want_final=True)
        want_fin = ob.frame_clean(acc, overlay, n, final_mode=(mode == nb.FRAME_FINAL))
        assert np.array_equal(fin[..., :3].view(np.uint32), want_fin[..., :3].view(np.uint32))
        assert np.array_equal(rows, ob.bmp_rows(want_fin, post))


def _tone(c):
    c = np.float32(c)
    num = np.float32(c * np.float32(np.float32(np.float32(2.51) * c) + np.float32(0.03)))
    den = np.float32(np.float32(c * np.float32(np.float32(np.float32(2.43) * c) + np.float32(0.59))) + np.float32(0.14))
    return np.float32(num / den)


@pytest.mark.parametrize("n,mode", [(0, 0), (1, 1), (3, 1), (7, 1), (2, 0)])
def test_values_at_the_byte_steps(built, n, mode):
    """The kernel accepts its SFU estimate of a byte only when the approximate tonemapped value is clear of the step positions
    and falls back to exact arithmetic otherwise: drive it with inputs whose tonemapped value sits ON every one of the 255
    steps and up to 6 ulp either side, directly (divisor 1) and through the / n and * 1/(n+1) normalisations."""
    steps = nb.gamma_steps()
    vals = []
    for k in range(1, 256):
        lo, hi = np.uint32(0), np.uint32(0x42000000)            # bisect the smallest c in [0, 32] with tone(c) >= steps[k]
        while lo < hi:
            mid = np.uint32((int(lo) + int(hi)) // 2)
            if _tone(mid.view(np.float32)) >= steps[k]:
                hi = mid
            else:
                lo = np.uint32(mid + 1)
        vals.extend(np.arange(int(lo) - 6, int(lo) + 7, dtype=np.int64).astype(np.uint32).view(np.float32).tolist())
    vals = np.array(vals, np.float32)
    div = np.float32(n if mode == 1 else n + 1)
    w = 255 * 13
    acc = np.zeros((3, w, 4), np.float32)
    acc[0, :, 0] = vals * div; acc[0, :, 1] = vals[::-1] * div; acc[0, :, 2] = np.roll(vals, 7) * div
    acc[1, :, :3] = np.nextafter(acc[0, :, :3], np.float32(np.inf)); acc[2, :, :3] = np.nextafter(acc[0, :, :3], np.float32(0))
    rows = nb.frame_finish(acc, None, n, mode, True)
    want = ob.bmp_rows(ob.frame_clean(acc, None, n, final_mode=(mode == 1)), True)
    assert np.array_equal(rows, want)
    assert len(np.unique(want[0, : 3 * w])) >= 250            # the row really walks all byte values


def test_full_size_4k_device_entry_point(built, tmp_path):
    torch = pytest.importorskip("torch")
    w, h, n = 3840, 2160, 64
    rng = np.random.default_rng(99)
    acc, ov = _frame(rng, h, w)
    want_fin = ob.frame_clean(acc, ov, n, final_mode=True)
    want_rows = ob.bmp_rows(want_fin, True)
    d_acc = torch.from_numpy(acc).cuda(); d_ov = torch.from_numpy(ov).cuda()
    d_rows = torch.zeros((h, nb.bmp_row_bytes(w)), dtype=torch.uint8, device="cuda")
    d_fin = torch.empty_like(d_acc)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    nb.frame_finish_device(d_acc.data_ptr(), d_ov.data_ptr(), w, h, n, nb.FRAME_FINAL, True, d_rows.data_ptr(), d_fin.data_ptr(), s.cuda_stream)
    s.synchronize()
    assert np.array_equal(d_rows.cpu().numpy(), want_rows)
    assert np.array_equal(d_fin.cpu().numpy()[..., :3].view(np.uint32), want_fin[..., :3].view(np.uint32))
    # file variant from device buffers; a frame that is already 8-bit survives postProcess=False unchanged (round trip)
    grid = (rng.integers(0, 256, (h, w, 3)).astype(np.float32) / np.float32(255.0))
    q = np.zeros((h, w, 4), np.float32); q[..., :3] = grid * np.float32(n)
    d_q = torch.from_numpy(q).cuda()
    p = str(tmp_path / "q.bmp")
    ms = nb.frame_save_bmp(p, d_q.data_ptr(), None, w, h, n, nb.FRAME_FINAL, False)
    body = np.fromfile(p, np.uint8)[54:].reshape(h, -1)[:, : 3 * w].reshape(h, w, 3)
    back = body[..., ::-1].astype(np.float32) / np.float32(255.0)
    assert np.abs(back - grid).max() <= 1.0 / 255.0 and ms > 0
    assert np.array_equal(body, ob.bmp_rows(ob.frame_clean(q, None, n, final_mode=True), False)[:, : 3 * w].reshape(h, w, 3))


def test_async_save_snapshots_the_accumulator(built, tmp_path):
    """Progressive saves queued behind a stream that keeps adding samples (deviceCode.cu:1905-1927 stalls the loop here): every file
    holds exactly the state the accumulator had when its save was queued, and equals the blocking variant byte for byte."""
    torch = pytest.importorskip("torch")
    w, h = 1001, 333                      # ragged: rows padded, last 512-pixel segment short
    rng = np.random.default_rng(5)
    sample = np.zeros((h, w, 4), np.float32); sample[..., :3] = rng.random((h, w, 3), dtype=np.float32) * 2.0
    sample[7, 11, 0] = np.nan; sample[8, 12, 1] = np.inf
    d_sample = torch.from_numpy(sample).cuda()
    d_acc = torch.zeros_like(d_sample)
    s = torch.cuda.Stream()
    paths, states = [], []
    with torch.cuda.stream(s):
        for k in range(5):
            d_acc += d_sample * float(k + 1)                    # the "next sample pass", queued right behind the previous save
            states.append(d_acc.clone())
            p = str(tmp_path / f"async_{k}.bmp"); paths.append(p)
            nb.frame_save_bmp_async(p, d_acc.data_ptr(), None, w, h, k, nb.FRAME_PROGRESSIVE, True, s.cuda_stream)
    nb.frame_save_wait()
    s.synchronize()
    for k, p in enumerate(paths):
        q = str(tmp_path / f"sync_{k}.bmp")
        nb.frame_save_bmp(q, states[k].data_ptr(), None, w, h, k, nb.FRAME_PROGRESSIVE, True)
        a, b = np.fromfile(p, np.uint8), np.fromfile(q, np.uint8)
        assert a.size == 54 + nb.bmp_row_bytes(w) * h and np.array_equal(a, b), k
        want = ob.bmp_rows(ob.frame_clean(states[k].cpu().numpy(), None, k, final_mode=False), True)
        assert np.array_equal(a[54:].reshape(h, -1), want), k
    nb.frame_save_wait()                                        # nothing pending: returns at once
    with pytest.raises(nb.NovumError):
        nb.frame_save_bmp_async(str(tmp_path / "no_such_dir" / "x.bmp"), d_acc.data_ptr(), None, w, h, 1, nb.FRAME_PROGRESSIVE, True, s.cuda_stream)
        nb.frame_save_wait()                                    # the write error surfaces at the wait


def test_bad_arguments(built):
    a = np.zeros((4, 4, 4), np.float32)
    with pytest.raises(nb.NovumError):
        nb.frame_finish(a, None, 0, nb.FRAME_FINAL, True)        # division by a zero sample count
    with pytest.raises(nb.NovumError):
        nb.frame_finish(a, None, 1, 7, True)                     # unknown mode
