"""Lane occupancy of the trace kernel's sections (tuning build: scripts/build_variant.sh stats "-DRT_LANE_STATS"):
RT_B200_LIB=ray_tracer_b200/variants/librt_b200_stats.so python scripts/lane_stats.py c3_teapot 32"""
import ctypes as C, sys
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, frames = sys.argv[1], int(sys.argv[2])
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(rt._lib.RT_OPT_STAGING_MIB, 8192); tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
L = rt._lib.lib()
buf = (C.c_ulonglong * 16)()
tr.render_frames(0, frames); tr.sync(); L.rt_debug_lane_stats(buf, 1); tr.reset_counters()
tr.render_frames(0, frames); tr.sync(); L.rt_debug_lane_stats(buf, 1)
c = tr.counters(); s = list(buf)
print(name, frames, "rays", c["rays"])
ns = max(s[0], 1)
print("node steps (warp) %d: per step lanes stepping %.2f | finished, waiting for SHADE %.2f | blocked on a second leaf %.2f | out of nodes, leaf parked %.2f | idle (queue empty) %.2f" % (
    s[0], s[1] / ns, s[2] / ns, s[3] / ns, s[4] / ns, s[5] / ns))
nl = max(s[6], 1)
print("leaf iterations (warp) %d: lanes with a primitive %.2f, triangle records looked at %.2f per iteration" % (s[6], s[7] / nl, s[8] / nl))
npass = max(s[9], 1)
print("SHADE passes (warp) %d: lanes finishing a ray %.2f (hits %.2f), lanes mid-traversal %.2f; rays per pass %.2f" % (s[9], s[10] / npass, s[11] / npass, s[12] / npass, c["rays"] / npass))
print("warp node steps per ray %.3f, leaf iterations per ray %.3f" % (s[0] / c["rays"], s[6] / c["rays"]))
