"""Item-queue policy against launch length: python scripts/queue_sweep.py c3_teapot "128,64,32,16,8" "chunk,guide;chunk,guide;..."
(RT_OPT_SCHED bits 24-31 = grab size / 32, bits 32-39 = guide; 0 = the library's default)"""
import sys
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name = sys.argv[1]
lengths = [int(x) for x in sys.argv[2].split(",")]
combos = [tuple(int(v) for v in c.split(",")) for c in sys.argv[3].split(";")]
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(rt._lib.RT_OPT_STAGING_MIB, 8192)
base = 8 | (1 << 8) | (22 << 16)
print("%-12s" % "chunk,guide", *["%9d" % n for n in lengths])
for chunk, guide in combos:
    row = []
    for n in lengths:
        tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, n)
        tr.set_option(rt._lib.RT_OPT_SCHED, base | ((chunk // 32) << 24) | (guide << 32))
        best = 1e9
        for rep in range(3):
            tr.reset_counters(); tr.render_frames(0, n); tr.sync()
            best = min(best, tr.counters()["kernel_ms"])
        row.append(best / n)
    print("%-12s" % ("%d,%d" % (chunk, guide)), *["%9.4f" % v for v in row], flush=True)
