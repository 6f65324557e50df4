"""Kernel-only rate of the in-tree (or RT_B200_LIB) build: python scripts/ab_quick.py [config:frames ...]
Prints a CRC of the accumulation image and the ray count per job: variants must agree on both."""
import sys
import zlib
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
args = [a for a in sys.argv[1:] if not a.startswith("--")]
jobs = args or ["c3_teapot:128", "c4_tree:32", "c5_stress:16", "c2_cube:32"]
for job in jobs:
    name, frames = job.split(":"); frames = int(frames)
    scene, meta = configs.build(name)
    p = meta["params"]
    tr = rt.RayTracer(0, p["width"], p["height"])
    tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    tr.set_option(rt._lib.RT_OPT_STAGING_MIB, 8192); tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
    best = 0.0
    for rep in range(3):
        tr.render_frames(0, frames); tr.sync(); tr.reset_counters()
        tr.render_frames(0, frames); tr.sync()
        c = tr.counters()
        best = max(best, c["rays"] / c["kernel_ms"] / 1e3)
    crc = zlib.crc32(tr.read_accum().tobytes())
    print("%-10s %4d dispatches  launches %d  best of 3: %9.1f Mrays/s  (%.3f ms)  rays %d  image crc %08x" % (name, frames, c["launches"], best, c["rays"] / best / 1e3, c["rays"], crc), flush=True)
    tr.close()
