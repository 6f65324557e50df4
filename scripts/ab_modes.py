"""A/B of trace-kernel options on one box: python scripts/ab_modes.py [config:frames ...]
For every config: item mode 0 (pixel items, running mean in the kernel) vs 1 ((pixel, dispatch) items + staging + blend),
CUDA-event kernel time of the second of two identical renders, images compared bit for bit."""
import sys
sys.path.insert(0, ".")
import numpy as np
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs

jobs = sys.argv[1:] or ["c3_teapot:128", "c4_tree:32", "c5_stress:16", "c2_cube:32", "c1_balls:8"]
for job in jobs:
    name, frames = job.split(":")
    frames = int(frames)
    scene, meta = configs.build(name)
    p = meta["params"]
    ref = None
    for mode in (1, 2, 1, 2):
        tr = rt.RayTracer(0, p["width"], p["height"])
        tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
        tr.set_option(rt._lib.RT_OPT_ITEM_MODE, mode)
        tr.render_frames(0, frames); tr.sync(); tr.clear(); tr.reset_counters()
        tr.render_frames(0, frames); tr.sync()
        c = tr.counters()
        img = tr.read_accum()
        tr.close()
        same = "" if ref is None else (" image == first: %s" % np.array_equal(ref.view(np.uint32), img.view(np.uint32)))
        if ref is None:
            ref = img
        print("%-10s %4d dispatches  item_mode %d  launches %d  kernel ms %8.3f  Mrays/s %9.1f%s" % (
            name, frames, mode, c["launches"], c["kernel_ms"], c["rays"] / c["kernel_ms"] / 1e3, same), flush=True)
