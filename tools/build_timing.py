"""Scene load breakdown: host parse vs device BVH8 build, first and repeated builds in one process."""
import os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cudapathtracer_b200 as nb
d = os.path.join(tempfile.gettempdir(), "novum_bt_dragon")
cfgp = nb.make_scene("dragon", d, 100) if not os.path.exists(os.path.join(d, "configs", "config.rendertron")) else os.path.join(d, "configs", "config.rendertron")
for i in range(3):
    t = time.time()
    s = nb.Scene.from_config(cfgp)
    dt = time.time() - t
    inf = s.info()
    print(f"load {i}: total {dt:.3f} s, parse {inf['parse_ms']:.1f} ms, bvh8 build {inf['bvh8_build_ms']:.2f} ms ({inf['bvh8_build_launches']} launches), {inf['tris']} tris, {inf['bvh8_nodes']} nodes", flush=True)
    s.close()
