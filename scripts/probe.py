"""Quick GPU probe: render a built-in scene and print Mrays/s (not the bench)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from oracle import host_ref

name, w, h, frames, bounces = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
brute = len(sys.argv) > 6 and sys.argv[6] == "brute"
cam, sp, vs, ix, ms = getattr(host_ref, "scene_" + name)(np.float32(w) / np.float32(h))
tr = rt.RayTracer(0, w, h)
t0 = time.time(); tr.upload_scene(sp, vs, ix, ms); tr.sync(); print("upload+build s", time.time() - t0, tr.bvh_stats())
tr.set_camera(cam); tr.set_params(number_of_bounces=bounces, rays_per_pixel=1, skybox=1, frames=0)
tr.set_option(rt._lib.RT_OPT_TRAVERSAL, 1 if brute else 0)
tr.render_frames(0, 2); tr.sync(); tr.reset_counters()
t0 = time.time(); tr.render_frames(0, frames); tr.sync(); dt = time.time() - t0
c = tr.counters()
print(name, w, h, "frames", frames, "wall s %.4f kernel ms %.3f rays %d Mrays/s %.1f (kernel %.1f)" %
      (dt, c["kernel_ms"], c["rays"], c["rays"] / dt / 1e6, c["rays"] / c["kernel_ms"] / 1e3))
img = tr.read_accum(); print("mean", img.reshape(-1, 4).mean(0), "nan", int(np.isnan(img).sum()))
