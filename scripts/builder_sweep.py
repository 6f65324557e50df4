"""Hierarchy builder study on the GPU: python scripts/builder_sweep.py c3_teapot 32
Prints build time, Mrays/s and per-ray traversal counts for the radix tree and PLOC windows."""
import sys, time
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, frames = sys.argv[1], int(sys.argv[2])
w = int(sys.argv[3]) if len(sys.argv) > 3 else None
h = int(sys.argv[4]) if len(sys.argv) > 4 else None
scene, meta = configs.build(name, w, h)
p = meta["params"]
L = rt._lib
import os
combos = [(1, 8, 4), (0, 2, 4), (0, 4, 4), (0, 8, 4), (0, 16, 4), (0, 8, 2), (0, 8, 3), (0, 8, 6), (0, 8, 8)]
if os.environ.get("COMBOS"):
    combos = [tuple(int(x) for x in c.split(",")) for c in os.environ["COMBOS"].split(";")]
for builder, radius, leaf in combos:
    tr = rt.RayTracer(0, p["width"], p["height"])
    tr.set_option(L.RT_OPT_BUILDER, builder); tr.set_option(L.RT_OPT_PLOC_RADIUS, radius); tr.set_option(L.RT_OPT_LEAF_SIZE, leaf)
    tr.upload_scene(*scene.arrays())
    t0 = time.perf_counter(); tr.upload_scene(*scene.arrays()); t1 = time.perf_counter()
    tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    tr.set_option(L.RT_OPT_FRAMES_PER_LAUNCH, frames)
    tr.render_frames(0, frames); tr.sync(); tr.reset_counters()
    tr.render_frames(0, frames); tr.sync()
    c = tr.counters()
    tr.set_option(L.RT_OPT_DETAIL_COUNTERS, 1); tr.reset_counters(); tr.render_frames(0, frames); tr.sync(); d = tr.counters()
    print("builder %d radius %2d leaf %d: upload %.2f ms  %.1f Mrays/s  nodes/ray %.2f tri/ray %.2f sph/ray %.2f  %s" % (
        builder, radius, leaf, (t1 - t0) * 1e3, c["rays"] / c["kernel_ms"] / 1e3, d["node_visits"] / d["rays"], d["tri_tests"] / d["rays"],
        d["sphere_tests"] / d["rays"], tr.bvh_stats()), flush=True)
    tr.close()
