"""Rebuild one scene many times and render a dispatch after each build: every build must give the same image and
ray count (the hierarchy may differ from build to build only if the builder has a race; the image may differ only
if culling is not conservative).  python scripts/build_stress.py c5_stress 40 [builder] [frames]"""
import sys, hashlib, collections
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, n = sys.argv[1], int(sys.argv[2])
builder = int(sys.argv[3]) if len(sys.argv) > 3 else 0
frames = int(sys.argv[4]) if len(sys.argv) > 4 else 1
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.set_option(rt._lib.RT_OPT_BUILDER, builder)
tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
seen = collections.Counter()
for i in range(n):
    tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures)
    tr.reset_counters(); tr.clear()
    tr.render_frames(0, frames); tr.sync()
    c = tr.counters()
    key = (c["rays"], hashlib.sha1(tr.read_accum().tobytes()).hexdigest()[:12], tuple(sorted(tr.bvh_stats().items())))
    seen[key] += 1
for k, v in seen.items():
    print(v, "x", k)
print("distinct results:", len(seen))
