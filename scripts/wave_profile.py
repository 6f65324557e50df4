"""One wavefront-mode render for ncu: python scripts/wave_profile.py c3_teapot 8 4194304 [kernel]"""
import sys
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
L = rt._lib
name, frames, slots = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
kernel = int(sys.argv[4]) if len(sys.argv) > 4 else 1
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(L.RT_OPT_FRAMES_PER_LAUNCH, frames)
tr.set_option(L.RT_OPT_KERNEL, kernel); tr.set_option(L.RT_OPT_WAVE_SLOTS, slots)
tr.render_frames(0, frames); tr.sync()
c = tr.counters()
print("kernel ms %.3f  Mrays/s %.1f" % (c["kernel_ms"], c["rays"] / c["kernel_ms"] / 1e3))
