"""Sweep the warp scheduler thresholds on one config: python scripts/sched_sweep.py c3_teapot 32 [combos "n,s;n,s"]"""
import sys
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, frames = sys.argv[1], int(sys.argv[2])
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
combos = [(n, 1, s) for n in (4, 8, 12, 16) for s in (16, 20, 22, 24, 28)]
if len(sys.argv) > 3:
    combos = [(int(c.split(",")[0]), 1, int(c.split(",")[1])) for c in sys.argv[3].split(";")]
for n, l, s in combos:
    tr.set_option(rt._lib.RT_OPT_SCHED, n | (l << 8) | (s << 16))
    tr.render_frames(0, frames); tr.sync(); tr.reset_counters()
    tr.render_frames(0, frames); tr.sync()
    c = tr.counters()
    print("n_stay %2d l_stay %2d s_min %2d  kernel ms %.3f  Mrays/s %.1f" % (n, l, s, c["kernel_ms"], c["rays"] / c["kernel_ms"] / 1e3), flush=True)
