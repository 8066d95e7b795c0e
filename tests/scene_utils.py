"""Test-side scene loading (independent of the product's C++ loader) and ray-batch generators.

The OBJ / config readers follow the reference's semantics (main.cu:936-1068, objects.cuh:844-943) closely enough
for the generated scenes: v / vt / vn / f with v/vt/vn triplets, fan triangulation, degenerate-area cull,
emissive meshes become lights.
"""
import os

import numpy as np

import cudapathtracer_b200 as nb
from cudapathtracer_b200.binding import TRI_DTYPE

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def parse_config(path):
    cfg, meshes, in_meshes = {}, [], False
    for line in open(path):
        line = line.strip()
        if not line:
            continue
        if line.startswith("Meshes"):
            in_meshes = True
            continue
        if in_meshes:
            p, e, m = [s.strip() for s in line.split(";")]
            mult = float(e.split("*")[0])
            col = [float(x) for x in e[e.index("(") + 1:e.index(")")].split(",")]
            meshes.append((p, [mult * c for c in col], int(m)))
        elif ":" in line:
            k, v = line.split(":", 1)
            if v.strip():
                cfg[k.strip()] = v.strip()
    return cfg, meshes


def load_obj_into(path, pos, nrm, uv, tris, emission, mat, n_lights):
    ps, ns, ts = len(pos), len(nrm), len(uv)
    is_light = sum(e * e for e in emission) > 0
    for line in open(path):
        if not line.strip() or line[0] in "#s":
            continue
        tok = line.split()
        if tok[0] == "v":
            pos.append([np.float32(float(tok[1])), np.float32(float(tok[2])), np.float32(float(tok[3])), 0.0])
        elif tok[0] == "vt":
            uv.append([np.float32(float(tok[1])), np.float32(1.0) - np.float32(float(tok[2]))])
        elif tok[0] == "vn":
            nrm.append([np.float32(float(tok[1])), np.float32(float(tok[2])), np.float32(float(tok[3])), 0.0])
        elif tok[0] == "f":
            vi, ti, ni = [], [], []
            for t in tok[1:]:
                a = t.split("/")
                vi.append(int(a[0]) - 1); ti.append(int(a[1]) - 1); ni.append(int(a[2]) - 1)
            for i in range(1, len(vi) - 1):
                i0, i1, i2 = vi[0] + ps, vi[i] + ps, vi[i + 1] + ps
                p0, p1, p2 = [np.array(pos[j][:3], np.float32) for j in (i0, i1, i2)]
                cp = np.cross((p1 - p0).astype(np.float32), (p2 - p0).astype(np.float32)).astype(np.float32)
                if np.float32(np.dot(cp, cp)) < np.float32(1e-18):
                    continue
                tr = np.zeros((), TRI_DTYPE)
                tr["aInd"], tr["bInd"], tr["cInd"] = i0, i1, i2
                tr["naInd"], tr["nbInd"], tr["ncInd"] = ni[0] + ns, ni[i] + ns, ni[i + 1] + ns
                tr["uvaInd"], tr["uvbInd"], tr["uvcInd"] = ti[0] + ts, ti[i] + ts, ti[i + 1] + ts
                tr["materialID"] = mat
                tr["emission"] = list(emission) + [0.0]
                tr["lightInd"] = n_lights[0] if is_light else -51
                tr["triInd"] = len(tris)
                tris.append(tr)
                if is_light:
                    n_lights[0] += 1


def load_scene_py(scene_dir):
    """Returns (cfg dict, pos (n,4), nrm (n,4), uv (n,2), tris TRI_DTYPE, mats MAT_DTYPE)."""
    cfg, meshes = parse_config(os.path.join(scene_dir, "configs", "config.rendertron"))
    pos, nrm, uv, tris, nl = [], [], [], [], [0]
    for p, e, m in meshes:
        load_obj_into(os.path.join(scene_dir, p), pos, nrm, uv, tris, e, m, nl)
    mats = nb.material_table()      # host-only call: main.cu:397-467 restated in the product's front end
    return (cfg, np.array(pos, np.float32), np.array(nrm, np.float32), np.array(uv, np.float32),
            np.array(tris, TRI_DTYPE), mats)


def camera_rays(cam, w, h, n, rng):
    """Pinhole rays through random pixels (not bit-equal to the integrator's; just a coherent ray batch)."""
    px = rng.integers(0, w, n); py = rng.integers(0, h, n)
    aspect = np.float32(w) / np.float32(h)
    u = ((2.0 * (px + rng.random(n)) / w - 1.0) * aspect * cam.fovScale).astype(np.float32)
    v = ((2.0 * (py + rng.random(n)) / h - 1.0) * cam.fovScale).astype(np.float32)
    f = np.array([cam.forward.x, cam.forward.y, cam.forward.z], np.float32)
    r = np.array([cam.right.x, cam.right.y, cam.right.z], np.float32)
    up = np.array([cam.up.x, cam.up.y, cam.up.z], np.float32)
    d = f[None, :] + u[:, None] * r[None, :] + v[:, None] * up[None, :]
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = [cam.cameraOrigin.x, cam.cameraOrigin.y, cam.cameraOrigin.z]
    rays[:, 4:7] = d
    return rays


def random_rays(mn, mx, n, rng):
    """Incoherent rays: origins uniform in the (slightly grown) scene box, directions uniform on the sphere."""
    mn, mx = np.array(mn, np.float32), np.array(mx, np.float32)
    ext = mx - mn
    o = (mn - 0.05 * ext) + rng.random((n, 3)).astype(np.float32) * (1.1 * ext)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = o
    rays[:, 4:7] = d
    return rays


def secondary_rays(rays, hits, pos, nrm, tris, rng):
    """Cosine-ish bounce rays leaving the hit points of a previous batch (incoherent, on-surface origins)."""
    ok = hits["tri"] >= 0
    r = rays[ok]; h = hits[ok]
    p = r[:, 0:3] + h["t"][:, None] * r[:, 4:7]
    t = tris[h["tri"]]
    a, b, c = pos[t["aInd"], :3], pos[t["bInd"], :3], pos[t["cInd"], :3]
    n = np.cross(b - a, c - a)
    n = n / np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
    n = np.where((np.sum(n * r[:, 4:7], axis=1) > 0)[:, None], -n, n)
    d = rng.normal(size=p.shape)
    d = d / np.linalg.norm(d, axis=1, keepdims=True)
    d = np.where((np.sum(d * n, axis=1) < 0)[:, None], -d, d)
    out = np.zeros((len(p), 8), np.float32)
    out[:, 0:3] = (p + 1e-5 * n).astype(np.float32)
    out[:, 4:7] = d.astype(np.float32)
    return out


def adversarial_rays(pos, tris, n, rng):
    """Rays aimed exactly at shared vertices / edge midpoints, axis-parallel rays, and rays starting on box planes."""
    k = n // 4
    t = tris[rng.integers(0, len(tris), k)]
    a, b = pos[t["aInd"], :3], pos[t["bInd"], :3]
    tgt = np.concatenate([a, (a + b) * np.float32(0.5)])[:2 * k]
    o = (tgt + rng.normal(size=tgt.shape).astype(np.float32) * np.float32(0.7)).astype(np.float32)
    d = tgt - o
    d = (d / np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-20)).astype(np.float32)
    r1 = np.zeros((len(o), 8), np.float32); r1[:, 0:3] = o; r1[:, 4:7] = d
    # axis-parallel directions through random points of the scene box
    mn, mx = pos[:, :3].min(0), pos[:, :3].max(0)
    o2 = (mn + rng.random((k, 3)).astype(np.float32) * (mx - mn)).astype(np.float32)
    axes = np.eye(3, dtype=np.float32)[rng.integers(0, 3, k)] * np.where(rng.random(k) < 0.5, -1, 1).astype(np.float32)[:, None]
    r2 = np.zeros((k, 8), np.float32); r2[:, 0:3] = o2; r2[:, 4:7] = axes
    # origins exactly on the scene-box planes, direction component along that axis exactly zero
    o3 = (mn + rng.random((k, 3)).astype(np.float32) * (mx - mn)).astype(np.float32)
    ax = rng.integers(0, 3, k)
    o3[np.arange(k), ax] = np.where(rng.random(k) < 0.5, mn[ax], mx[ax])
    d3 = rng.normal(size=(k, 3)).astype(np.float32)
    d3[np.arange(k), ax] = 0.0
    d3 = (d3 / np.maximum(np.linalg.norm(d3, axis=1, keepdims=True), 1e-20)).astype(np.float32)
    r3 = np.zeros((k, 8), np.float32); r3[:, 0:3] = o3; r3[:, 4:7] = d3
    return np.concatenate([r1, r2, r3])


def classify(ours, ref, rays=None):
    """Parity classes of SURVEY §7.2: returns dict with counts of exact / tie / superset / mismatch."""
    same_id = ours["tri"] == ref["tri"]
    both_hit = (ours["tri"] >= 0) & (ref["tri"] >= 0)
    t_equal = ours["t"].view(np.uint32) == ref["t"].view(np.uint32)
    exact = same_id & ((ours["tri"] < 0) | (t_equal & (ours["u"].view(np.uint32) == ref["u"].view(np.uint32)) &
                                             (ours["v"].view(np.uint32) == ref["v"].view(np.uint32))))
    tie = both_hit & ~same_id & t_equal
    # SUPERSET: we hit where the reference's own slab test culled the triangle (it then reports a miss, or a farther hit)
    superset = ((ours["tri"] >= 0) & (ref["tri"] < 0)) | (both_hit & ~same_id & (ours["t"] < ref["t"]))
    mismatch = ~(exact | tie | superset)
    return {"n": len(ours), "exact": int(exact.sum()), "tie": int(tie.sum()), "superset": int(superset.sum()),
            "mismatch": int(mismatch.sum()), "mismatch_idx": np.nonzero(mismatch)[0][:16]}
