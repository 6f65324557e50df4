"""C5 flattened vs instanced (RT_OPT_INSTANCING 0 / 1): records, nodes, upload / refit latency, rays per second.
python scripts/inst_probe.py [frames]"""
import sys, time
sys.path.insert(0, ".")
import numpy as np
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 8
scene, meta = configs.build("c5_stress")
p = meta["params"]
sp, vs, ix, ms = scene.arrays()
ms2 = ms.copy(); ms2["pos"][:, 2] += np.float32(5.0)
for mode in (0, 1):
    tr = rt.RayTracer(0, p["width"], p["height"])
    tr.set_option(rt._lib.RT_OPT_INSTANCING, mode)
    tr.upload_scene(sp, vs, ix, ms)
    t0 = time.perf_counter(); tr.upload_scene(sp, vs, ix, ms); t1 = time.perf_counter()
    tr.update_records(sp, ms2); t2 = time.perf_counter(); tr.update_records(sp, ms2); t3 = time.perf_counter()
    tr.update_records(sp, ms)
    st = tr.bvh_stats()
    tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
    best = 0.0
    for _ in range(2):
        tr.render_frames(0, frames); tr.sync(); tr.reset_counters()
        tr.render_frames(0, frames); tr.sync()
        c = tr.counters(); best = max(best, c["rays"] / c["kernel_ms"] / 1e3)
    tr.set_option(rt._lib.RT_OPT_DETAIL_COUNTERS, 1); tr.reset_counters(); tr.render_frames(0, 2); tr.sync(); d = tr.counters()
    mem = (st["n_prims"] * 64 + st["n_nodes"] * 64) / 1e6
    print("instancing %d: top-level records %d, nodes %d (records + nodes %.1f MB), depth %d, upload %.2f ms, update_records %.2f ms, %.1f Mrays/s, nodes/ray %.2f tri/ray %.2f sph/ray %.2f" % (
        mode, st["n_prims"], st["n_nodes"], mem, st["max_depth"], (t1 - t0) * 1e3, (t3 - t2) * 1e3, best, d["node_visits"] / d["rays"], d["tri_tests"] / d["rays"], d["sphere_tests"] / d["rays"]), flush=True)
    tr.close()
