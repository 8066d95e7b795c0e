import os, sys, tempfile, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import cudapathtracer_b200 as nb
d = tempfile.mkdtemp(); nb.make_scene("materials", d, 100)
s = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
def render(integ, spp, nested):
    cfg = nb.RenderConfig.from_buffer_copy(bytes(s.cfg)); cfg.width, cfg.height, cfg.integrator = 64, 64, integ
    cam = nb.camera_make(cfg); s.set_features(nested_bidir=nested)
    img, st = s.render(cfg, cam, 0, spp); return img[..., :3] / spp
crop = (slice(37, 50), slice(34, 47))
nee = render(0, 4096, False)
print("NEE full", nee.mean(), "crop", nee[crop].mean())
for integ, name in ((1, "BDPT"), (3, "VCM"), (4, "SPPM")):
    for nested in (False, True):
        im = render(integ, 1024, nested)
        print(name, "nested" if nested else "plain ", "full", im.mean(), "crop", im[crop].mean(), "finite", np.isfinite(im).all())
