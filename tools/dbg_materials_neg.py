import os, sys, tempfile, subprocess, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import cudapathtracer_b200 as nb
d = tempfile.mkdtemp()
nb.make_scene("materials", d, 100)
s = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
def cfgfor(integ):
    cfg = nb.RenderConfig.from_buffer_copy(bytes(s.cfg)); cfg.width, cfg.height, cfg.integrator = 64, 64, integ; return cfg
cam = nb.camera_make(cfgfor(3))
for integ in (0, 1, 3, 4):
    cfg = cfgfor(integ)
    img, st = s.render(cfg, cam, 0, 256)
    neg = (img[..., :3] < 0)
    print("integ", integ, "mean", img[..., :3].mean() / 256, "neg px", int(neg.any(axis=2).sum()), "min", img[..., :3].min() / 256)
    if neg.any():
        ys, xs = np.nonzero(neg.any(axis=2)); print("  where", list(zip(ys[:10], xs[:10])), img[ys[0], xs[0]] / 256)
cfg = cfgfor(3)
for k in ("bdptLightTrace", "bdptNee", "bdptNaive", "bdptConnection", "vcmDoMerge"):
    for j in ("bdptLightTrace", "bdptNee", "bdptNaive", "bdptConnection", "vcmDoMerge"):
        setattr(cfg, j, 1 if j == k else 0)
    for mis in (1, 0):
        cfg.bdptDoMis = mis
        img, st = s.render(cfg, cam, 0, 256)
        print("VCM only", k, "mis", mis, "mean", img[..., :3].mean() / 256, "neg px", int((img[..., :3] < 0).any(axis=2).sum()), "min", img[..., :3].min() / 256)
s.close()
