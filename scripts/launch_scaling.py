"""Kernel time against dispatches per launch (fixed cost per launch?): python scripts/launch_scaling.py c3_teapot"""
import sys
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name = sys.argv[1]
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
import os
tr.set_option(rt._lib.RT_OPT_STAGING_MIB, int(os.environ.get('STAGING_MIB', '2048')))
for frames in [int(x) for x in os.environ.get('FRAMES', '1,2,4,8,16,32,64').split(',')]:
    tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
    tr.render_frames(0, frames); tr.sync(); tr.reset_counters()
    for _ in range(3):
        tr.render_frames(0, frames)
    tr.sync(); c = tr.counters()
    print("%2d dispatches per launch: %.3f ms per launch, %.3f ms per dispatch, %.1f Mrays/s" % (frames, c["kernel_ms"] / 3, c["kernel_ms"] / 3 / frames, c["rays"] / c["kernel_ms"] / 1e3), flush=True)
