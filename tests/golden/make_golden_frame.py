"""Golden vectors for the frame finish, from the REFERENCE'S OWN code (oracle/_ref/novum_ref):

  frame_bmp.npz    `novum_ref bmp`: Image::setColor + Image::saveImageBMP (imageUtil.cu:60-100, 202-232) — HOST code of the
                   reference, so this part runs on any CPU (it was made in the build container): the cleaned float4 frame
                   and the complete .bmp files it writes with postProcess on and off.
  frame_clean.npz  `novum_ref clean`: the reference's cleanAndFormatImage kernel (integratorUtilities.cuh:337-380) on a raw
                   accumulator with NaN / Inf / negative pixels and a sparse overlay — needs a GPU (made on a B200 box).

Run:  python tests/golden/make_golden_frame.py [--out DIR]      (each part is skipped if it cannot run here)
tests/test_oracle.py holds the CPU restatement to both; tests/test_gpu_frame.py holds the CUDA kernel to both."""
import argparse
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "..", "..", "oracle", "_ref", "novum_ref")
W, H, N = 61, 37, 11          # 61 * 3 = 183 bytes -> one byte of row padding


def cleaned_frame(rng):
    """A frame as Image::setColor receives it: finite, mostly [0, 8), with exact 0 / 1, tiny, huge and negative entries."""
    f = np.zeros((H, W, 4), np.float32)
    f[..., :3] = (rng.random((H, W, 3)) ** 3 * 8).astype(np.float32)
    f[..., 3] = 1.0
    flat = f.reshape(-1, 4)
    special = [0.0, 1.0, 1e-30, 1e-7, 1.1049568e-06, 0.0031308, 0.5, 0.18, 255.0 / 256, 1e6, 3.4e38, -0.005, -0.25, -3.0, -1e-20, 0.99999994]
    for i, v in enumerate(special):
        flat[i * 7 % len(flat), i % 3] = v
    lin = np.linspace(0, 1.2, W, dtype=np.float32)          # a grey ramp row: every output byte value appears
    f[H // 2, :, 0] = lin; f[H // 2, :, 1] = lin; f[H // 2, :, 2] = lin
    return f


def raw_accumulator(rng):
    acc = np.zeros((H, W, 4), np.float32)
    acc[..., :3] = (rng.random((H, W, 3)) ** 2 * 40).astype(np.float32)
    acc[..., 3] = rng.random((H, W)).astype(np.float32)
    flat = acc.reshape(-1, 4)
    for i, (c, v) in enumerate([(0, np.nan), (1, np.nan), (2, np.nan), (0, np.inf), (1, -np.inf), (2, np.inf), (0, -1.0), (1, -1e-30), (2, -5.0),
                                (0, 0.0), (1, 3.4e38)]):
        flat[(i * 53 + 5) % len(flat), c] = v
    flat[100, 0] = np.nan; flat[100, 1] = np.inf               # NaN wins over Inf
    flat[101, 0] = np.inf; flat[101, 1] = -2.0                 # Inf wins over negative
    ov = np.zeros((H, W, 4), np.float32)
    idx = rng.choice(H * W, 40, replace=False)
    ovf = ov.reshape(-1, 4)
    ovf[idx, :3] = rng.random((40, 3)).astype(np.float32)
    ovf[idx[:5], 1:3] = 0.0                                     # only x non-zero
    ovf[idx[5:8], 0:2] = 0.0                                    # only z non-zero
    ovf[idx[8], :3] = (-0.5, 0.0, 0.0)                          # negative counts as non-zero
    ovf[idx[9], :3] = (0.0, -0.0, 0.0)                          # -0.0 == 0: no override
    ovf[100, :3] = (0.25, 0.5, 0.75)                            # overlay over a NaN pixel
    return acc, ov


def run(args):
    out = subprocess.run([REF] + args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if out.returncode != 0:
        raise RuntimeError(f"novum_ref {args[0]} failed ({out.returncode}): {out.stdout[-400:]}")
    return out.stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=HERE)
    a = ap.parse_args()
    if not os.path.exists(REF):
        raise SystemExit(f"{REF} missing: run oracle/build_ref.sh where /root/reference exists")
    os.makedirs(a.out, exist_ok=True)
    rng = np.random.default_rng(20261104)
    tmp = tempfile.mkdtemp()
    frame = cleaned_frame(rng)
    frame.tofile(os.path.join(tmp, "frame.raw"))
    files = {}
    for post in (1, 0):
        p = os.path.join(tmp, f"o{post}.bmp")
        run(["bmp", os.path.join(tmp, "frame.raw"), str(W), str(H), str(post), p])
        files[post] = np.fromfile(p, np.uint8)
    np.savez_compressed(os.path.join(a.out, "frame_bmp.npz"), frame=frame, bmp_post=files[1], bmp_plain=files[0], w=W, h=H)
    print("frame_bmp.npz written:", files[1].size, "bytes per file")
    acc, ov = raw_accumulator(rng)
    acc.tofile(os.path.join(tmp, "acc.raw")); ov.tofile(os.path.join(tmp, "ov.raw"))
    try:
        run(["clean", os.path.join(tmp, "acc.raw"), os.path.join(tmp, "ov.raw"), str(W), str(H), str(N), os.path.join(tmp, "clean.raw")])
        clean = np.fromfile(os.path.join(tmp, "clean.raw"), np.float32).reshape(H, W, 4)
        if not np.isfinite(clean[..., :3]).all() or float(np.abs(clean).sum()) == 0.0:
            raise RuntimeError("cleanAndFormatImage wrote nothing (no GPU?)")
        np.savez_compressed(os.path.join(a.out, "frame_clean.npz"), acc=acc, overlay=ov, clean=clean, n=N, w=W, h=H)
        print("frame_clean.npz written")
    except RuntimeError as e:
        print("frame_clean.npz skipped:", e)


if __name__ == "__main__":
    main()
