"""Frame finish, CPU side: the restatement in oracle/novum_oracle_image.cpp is held to the reference's own
Image::saveImageBMP (fixture made by running the reference's host code, tests/golden/make_golden_frame.py) and to its
cleanAndFormatImage kernel (fixture made on a B200); the product's gamma step table is checked exhaustively; the no-GPU
behaviour of the C ABI is an error, never a CPU fallback."""
import os

import numpy as np
import pytest

from oracle import oracle_binding as ob

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _bmp_body(file_bytes, w, h):
    assert file_bytes[0] == 0x42 and file_bytes[1] == 0x4D                      # 'BM'
    row = (3 * w + 3) & ~3
    assert int.from_bytes(file_bytes[2:6].tobytes(), "little") == 54 + row * h   # bfSize, imageUtil.cu:239
    assert int.from_bytes(file_bytes[10:14].tobytes(), "little") == 54
    assert int.from_bytes(file_bytes[18:22].tobytes(), "little") == w and int.from_bytes(file_bytes[22:26].tobytes(), "little") == h
    return file_bytes[54:].reshape(h, row)


@pytest.mark.parametrize("post,key", [(1, "bmp_post"), (0, "bmp_plain")])
def test_oracle_rows_equal_reference_saveImageBMP(built, post, key):
    g = np.load(os.path.join(GOLD, "frame_bmp.npz"))
    w, h = int(g["w"]), int(g["h"])
    want = _bmp_body(g[key], w, h)
    got = ob.bmp_rows(g["frame"], post)
    assert np.array_equal(got, want)                       # bit-exact, padding byte included
    if post:
        ramp = want[h // 2, : 3 * w].reshape(w, 3)
        assert ramp.min() == 0 and ramp.max() >= 200 and (np.diff(ramp[:, 0].astype(int)) >= 0).all()


def test_oracle_clean_equals_reference_kernel(built):
    p = os.path.join(GOLD, "frame_clean.npz")
    g = np.load(p)
    got = ob.frame_clean(g["acc"], g["overlay"], int(g["n"]))
    assert np.array_equal(got.view(np.uint32), g["clean"].view(np.uint32))
    flat = g["clean"].reshape(-1, 4)
    assert tuple(flat[100, :3]) == (0.25, 0.5, 0.75)       # overlay over a NaN pixel
    assert (flat[:, :3] == (1.0, 0.0, 1.0)).all(axis=1).any() and (flat[:, :3] == (0.0, 1.0, 0.0)).all(axis=1).any()
    assert (flat[:, :3] == (0.0, 0.0, 1.0)).all(axis=1).any()


def test_final_normalise_rules():
    acc = np.zeros((2, 4, 4), np.float32)
    acc[0, 0, :3] = (3.0, 6.0, 9.0); acc[0, 1, 0] = np.nan; acc[0, 2, 1] = np.inf; acc[0, 3, :3] = (-3.0, 1.0, 1.0)
    acc[1, 0, :3] = (np.nan, np.inf, 0.0); acc[1, 1, :3] = (1.0, 1.0, 1.0)
    ov = np.zeros_like(acc); ov[1, 1, :3] = (0.0, 0.0, 0.5)
    out = ob.frame_clean(acc, ov, 3, final_mode=True)
    assert tuple(out[0, 0, :3]) == (1.0, 2.0, 3.0)
    assert tuple(out[0, 1, :3]) == (1.0, 0.0, 1.0) and tuple(out[0, 2, :3]) == (0.0, 1.0, 0.0)
    assert tuple(out[0, 3, :3]) == (-1.0, np.float32(1.0) / np.float32(3.0), np.float32(1.0) / np.float32(3.0))   # negatives survive (main.cu:877 only prints)
    assert tuple(out[1, 0, :3]) == (1.0, 0.0, 1.0)          # NaN marking runs first; pink is not Inf
    assert tuple(out[1, 1, :3]) == (0.0, 0.0, 0.5)
    # progressive and final differ in rounding: * (1/(n+1)) vs / n
    a = np.full((1, 1, 4), 0.7, np.float32)
    p = ob.frame_clean(a, None, 2)[0, 0, 0]; f = ob.frame_clean(a, None, 3, final_mode=True)[0, 0, 0]
    assert p == np.float32(0.7) * (np.float32(1.0) / np.float32(3.0)) and f == np.float32(0.7) / np.float32(3.0)


def test_gamma_step_table_is_exact_for_every_float(built):
    """The device looks byte(t) up in 255 step positions; over ALL 1,065,353,217 floats in [0, 1] the look-up equals
    (uchar)(clamp(powf(t, 1/2.2f)) * 255 + 0.5) — which also proves that function is monotone on this libm."""
    import cudapathtracer_b200 as nb
    steps = nb.gamma_steps()
    assert steps[0] == -np.inf and steps[256] == np.inf and (np.diff(steps[1:256]) > 0).all()
    assert ob.gamma_table_check(steps) == 0


def test_frame_abi_without_gpu_is_an_error(built):
    import cudapathtracer_b200 as nb
    if nb.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(nb.NovumError):
        nb.frame_finish(np.zeros((4, 4, 4), np.float32), None, 1)
    with pytest.raises(nb.NovumError):                          # the asynchronous save has no host path either
        nb.frame_save_bmp_async("unused.bmp", None, None, 4, 4, 1)
    nb.frame_save_wait()                                        # nothing was ever queued: returns at once
    assert nb.bmp_row_bytes(61) == 184 and nb.bmp_row_bytes(64) == 192 and nb.bmp_row_bytes(1) == 4
