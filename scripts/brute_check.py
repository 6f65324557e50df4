"""Full-size check that the hierarchy is result-invisible: the whole image through the BVH against the GPU's own
exhaustive mode (every primitive tested for every ray), bit for bit.  python scripts/brute_check.py c5_stress 1 [builder]"""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, frames = sys.argv[1], int(sys.argv[2])
builder = int(sys.argv[3]) if len(sys.argv) > 3 else 0
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.set_option(rt._lib.RT_OPT_BUILDER, builder)
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
out = []
for mode in (0, 1):
    tr.set_option(rt._lib.RT_OPT_TRAVERSAL, mode); tr.set_option(rt._lib.RT_OPT_PRIMARY_IDS, 1)
    tr.clear(); tr.reset_counters()
    t0 = time.time(); tr.render_frames(0, frames); tr.sync()
    c = tr.counters()
    print("%s: rays %d, %.2f s" % ("brute force" if mode else "bvh", c["rays"], time.time() - t0), flush=True)
    out.append((tr.read_accum(), tr.read_primary_ids(), c["rays"]))
(a, aid, ar), (b, bid, br) = out
d = np.any(a.view(np.uint32) != b.view(np.uint32), axis=2)
print("%s %dx%d, %d dispatches: pixels that differ %d, primary ids that differ %d, rays %s" %
      (name, p["width"], p["height"], frames, int(d.sum()), int((aid != bid).sum()), "equal" if ar == br else "DIFFER"))
ys, xs = np.nonzero(d)
for y, x in list(zip(ys, xs))[:10]:
    print("  (%d,%d) bvh %s brute %s" % (x, y, a[y, x], b[y, x]))
