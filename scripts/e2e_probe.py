import sys, time
sys.path.insert(0, ".")
import numpy as np
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
scene, meta = configs.build("c3_teapot")
p = meta["params"]
tr = rt.RayTracer(0, p["width"], p["height"])
out = np.empty((p["height"], p["width"], 4), np.float32)
for it in range(4):
    t0 = time.perf_counter(); tr.upload_scene(*scene.arrays()); t1 = time.perf_counter()
    tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    tr.render_frames(0, 128); t2 = time.perf_counter(); tr.sync(); t3 = time.perf_counter()
    tr.read_accum(out); t4 = time.perf_counter()
    print("upload %.2f ms  launch calls %.2f ms  sync %.2f ms  read %.2f ms" % ((t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t4-t3)*1e3), flush=True)
for b in (1, 0):
    tr.set_option(rt._lib.RT_OPT_BUILDER, b)
    for it in range(3):
        t0 = time.perf_counter(); tr.upload_scene(*scene.arrays()); t1 = time.perf_counter()
        print("builder", b, "upload %.2f ms" % ((t1-t0)*1e3))
# refit vs rebuild on the 1.1 M-primitive scene
scene, meta = configs.build("c5_stress", 64, 36)
tr = rt.RayTracer(0, 64, 36)
sp, vs, ix, ms = scene.arrays()
tr.upload_scene(sp, vs, ix, ms)
ms2 = ms.copy(); ms2["pos"][:, 2] += 5.0
for it in range(3):
    t0 = time.perf_counter(); tr.upload_scene(sp, vs, ix, ms); t1 = time.perf_counter(); tr.update_records(sp, ms2); t2 = time.perf_counter()
    print("c5: rebuild %.2f ms  refit %.2f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3))
