"""A/B of a runtime option on the in-tree build: python scripts/ab_opt.py OPTION_NUMBER v0,v1[,..] [config:frames ...]"""
import sys
sys.path.insert(0, ".")
import numpy as np
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
opt, vals = int(sys.argv[1]), [int(v) for v in sys.argv[2].split(",")]
jobs = sys.argv[3:] or ["c3_teapot:128", "c3_teapot:16", "c4_tree:32", "c5_stress:16", "c2_cube:32", "c1_balls:8"]
for job in jobs:
    name, frames = job.split(":"); frames = int(frames)
    scene, meta = configs.build(name); p = meta["params"]
    tr = rt.RayTracer(0, p["width"], p["height"])
    tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    tr.set_option(rt._lib.RT_OPT_STAGING_MIB, 8192); tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
    ref = None
    for v in vals + vals:
        tr.set_option(opt, v)
        tr.render_frames(0, frames); tr.sync(); tr.clear(); tr.reset_counters()
        tr.render_frames(0, frames); tr.sync()
        c = tr.counters(); img = tr.read_accum()
        same = "" if ref is None else " image == first: %s" % np.array_equal(ref.view(np.uint32), img.view(np.uint32))
        ref = img if ref is None else ref
        print("%-10s %3d dispatches  option %d = %d  %8.3f ms  %9.1f Mrays/s%s" % (name, frames, opt, v, c["kernel_ms"], c["rays"] / c["kernel_ms"] / 1e3, same), flush=True)
    tr.close()
