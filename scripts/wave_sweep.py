"""Megakernel vs wavefront kernels on one config: python scripts/wave_sweep.py c3_teapot 32 ["slots,n_stay,refill;..."]
Prints rays/s for each and checks the two images bit for bit."""
import sys
import numpy as np
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
L = rt._lib
name, frames = sys.argv[1], int(sys.argv[2])
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(L.RT_OPT_FRAMES_PER_LAUNCH, frames)

def run():
    tr.clear(); tr.render_frames(0, frames); tr.sync(); tr.reset_counters()
    tr.clear(); tr.render_frames(0, frames); tr.sync()
    c = tr.counters()
    return c, tr.read_accum()

c0, img0 = run()
print("megakernel            kernel ms %.3f  Mrays/s %.1f  rays %d" % (c0["kernel_ms"], c0["rays"] / c0["kernel_ms"] / 1e3, c0["rays"]), flush=True)
combos = [(s << 10, n, r) for s in (1024, 2048, 4096, 8192) for n, r in ((8, 8), (8, 16), (8, 24))]
if len(sys.argv) > 3:
    combos = [tuple(int(x) for x in c.split(",")) for c in sys.argv[3].split(";")]
tr.set_option(L.RT_OPT_KERNEL, 1)
for slots, n, r in combos:
    tr.set_option(L.RT_OPT_WAVE_SLOTS, slots)
    tr.set_option(L.RT_OPT_SCHED, n | (r << 8) | (22 << 16))
    c, img = run()
    same = np.array_equal(img.view(np.uint32), img0.view(np.uint32)) and c["rays"] == c0["rays"] and c["samples"] == c0["samples"]
    print("wave slots %8d n_stay %2d refill %2d  kernel ms %.3f  Mrays/s %.1f  bit-equal %s" % (slots, n, r, c["kernel_ms"], c["rays"] / c["kernel_ms"] / 1e3, same), flush=True)
