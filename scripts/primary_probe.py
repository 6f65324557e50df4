"""Traversal cost of primary rays alone (bounces = 0): python scripts/primary_probe.py c3_teapot 16"""
import sys
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, frames = sys.argv[1], int(sys.argv[2])
scene, meta = configs.build(name)
p = dict(meta["params"])
for bounces in (0, p["number_of_bounces"]):
    p["number_of_bounces"] = bounces
    tr = rt.RayTracer(0, p["width"], p["height"])
    tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
    tr.render_frames(0, frames); tr.sync(); tr.reset_counters()
    tr.render_frames(0, frames); tr.sync(); c = tr.counters()
    tr.set_option(rt._lib.RT_OPT_DETAIL_COUNTERS, 1); tr.reset_counters(); tr.render_frames(0, frames); tr.sync(); d = tr.counters()
    print("bounces %d: %.3f ms, %.1f Mrays/s, rays %d, nodes/ray %.2f tri/ray %.2f sph/ray %.2f hit fraction %.3f" % (
        bounces, c["kernel_ms"], c["rays"] / c["kernel_ms"] / 1e3, c["rays"], d["node_visits"] / d["rays"], d["tri_tests"] / d["rays"],
        d["sphere_tests"] / d["rays"], d["hit_segments"] / d["rays"]))
    tr.close()
print("camera lens_radius", scene.camera_uniform()["lens_radius"])
