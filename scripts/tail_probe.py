"""When do the warps of a launch run out of queue and when do they finish?  Tuning build: scripts/build_variant.sh probe "-DRT_TIME_PROBE"
RT_B200_LIB=ray_tracer_b200/variants/librt_b200_probe.so python scripts/tail_probe.py c3_teapot 16"""
import ctypes as C, sys
sys.path.insert(0, ".")
import numpy as np
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, frames = sys.argv[1], int(sys.argv[2])
scene, meta = configs.build(name)
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(rt._lib.RT_OPT_STAGING_MIB, 8192); tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
if len(sys.argv) > 3:
    chunk, guide = (int(v) for v in sys.argv[3].split(","))
    tr.set_option(rt._lib.RT_OPT_SCHED, 8 | (1 << 8) | (22 << 16) | ((chunk // 32) << 24) | (guide << 32))
L = rt._lib.lib()
n = 148 * 32
buf = (C.c_ulonglong * (2 * n))()
tr.render_frames(0, frames); tr.sync()
tr.reset_counters(); tr.render_frames(0, frames); tr.sync()
L.rt_debug_warp_times(buf, n)
a = np.array(buf, dtype=np.uint64).reshape(n, 2).astype(np.float64)
ms = tr.counters()["kernel_ms"]
end = a[:, 1]; ex = a[:, 0]
t_end = end.max()
print("%s x %d: kernel_ms %.3f" % (name, frames, ms))
fin = np.sort((t_end - end) / 1e6)          # ms before the last warp's exit
print("warps finished this many ms before the end: median %.3f  p10 %.3f  p1 %.3f  (p0 = 0: the last warp)" % (np.median(fin), np.percentile(fin, 10), np.percentile(fin, 1)))
for q in (50, 25, 10, 5, 2, 1):
    print("  %5.1f %% of the warps still running %.3f ms before the end" % (q, np.percentile(fin, q)))
hold = (end - ex)[ex > 0] / 1e6
print("time from a warp's first empty grab to its exit: median %.3f ms  p90 %.3f  max %.3f" % (np.median(hold), np.percentile(hold, 90), hold.max()))
print("first empty grab happened %.3f ms before the end (earliest), %.3f (median)" % ((t_end - ex[ex > 0].min()) / 1e6, (t_end - np.median(ex[ex > 0])) / 1e6))
