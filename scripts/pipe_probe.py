"""Pipelined launches (RT_OPT_PIPELINE) against one long launch and against serial short launches: python scripts/pipe_probe.py"""
import sys
sys.path.insert(0, ".")
import numpy as np
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
L = rt._lib
for name, frames in (("c3_teapot", 128), ("c4_tree", 64), ("c5_stress", 16), ("c2_cube", 32)):
    scene, meta = configs.build(name); p = meta["params"]
    tr = rt.RayTracer(0, p["width"], p["height"])
    tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    ref = None
    for mib, pipe in ((16384, 1), (4352, 1), (1024, 1), (1024, 0), (512, 1), (256, 1)):
        tr.set_option(L.RT_OPT_STAGING_MIB, mib); tr.set_option(L.RT_OPT_PIPELINE, pipe)
        tr.update_bind_group(p["width"], p["height"]); tr.set_params(**p)      # new image: the staging planes are re-allocated
        tr.render_frames(0, frames); tr.sync(); tr.clear(); tr.reset_counters()
        tr.render_frames(0, frames); tr.sync()
        c = tr.counters(); img = tr.read_accum()
        same = "" if ref is None else " image == first: %s" % np.array_equal(ref.view(np.uint32), img.view(np.uint32))
        ref = img if ref is None else ref
        print("%-10s %3d dispatches  staging %5d MiB  pipeline %d  launches %2d  %8.3f ms  %9.1f Mrays/s%s" % (
            name, frames, mib, pipe, c["launches"], c["kernel_ms"], c["rays"] / c["kernel_ms"] / 1e3, same), flush=True)
    tr.close()
