"""Short single-GPU run of one config for ncu: python scripts/profile_run.py c3_teapot 16 [w h]"""
import sys
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, frames = sys.argv[1], int(sys.argv[2])
w = int(sys.argv[3]) if len(sys.argv) > 3 else None
h = int(sys.argv[4]) if len(sys.argv) > 4 else None
scene, meta = configs.build(name, w, h)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
tr.render_frames(0, frames); tr.sync(); tr.reset_counters()
tr.render_frames(0, frames); tr.sync()
c = tr.counters()
print(name, p["width"], p["height"], "frames", frames, "kernel ms %.3f rays %d Mrays/s %.1f" % (c["kernel_ms"], c["rays"], c["rays"] / c["kernel_ms"] / 1e3), tr.bvh_stats())
tr.set_option(rt._lib.RT_OPT_DETAIL_COUNTERS, 1); tr.reset_counters(); tr.render_frames(0, frames); tr.sync(); d = tr.counters()
print("per ray: nodes %.2f tri %.2f sph %.2f hitfrac %.3f rays/sample %.3f" % (d["node_visits"]/d["rays"], d["tri_tests"]/d["rays"], d["sphere_tests"]/d["rays"], d["hit_segments"]/d["rays"], d["rays"]/d["samples"]))
