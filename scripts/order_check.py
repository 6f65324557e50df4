"""Is the BVH traversal result-invisible whatever the visiting order?  Renders one config with two builds of the
library (different node-loop schedules visit the hierarchy in different orders), compares the images bit for bit,
and re-renders every 8x8 tile that differs in exhaustive brute-force mode to see which build (if any) is right.
  python scripts/order_check.py c5_stress 4 LIB_A.so LIB_B.so"""
import os, subprocess, sys
import numpy as np
sys.path.insert(0, ".")

def child(name, frames, out):
    import ray_tracer_b200 as rt
    from ray_tracer_b200 import configs
    scene, meta = configs.build(name)
    p = meta["params"]
    tr = rt.RayTracer(0, p["width"], p["height"])
    tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
    tr.render_frames(0, frames); tr.sync()
    print(out, "rays", tr.counters()["rays"], flush=True)
    np.save(out, tr.read_accum())

if sys.argv[1] == "--child":
    child(sys.argv[2], int(sys.argv[3]), sys.argv[4]); sys.exit(0)

name, frames, libs = sys.argv[1], int(sys.argv[2]), sys.argv[3:5]
outs = []
for k, lib in enumerate(libs):
    out = "/tmp/order_check_%d.npy" % k
    env = dict(os.environ, RT_B200_LIB=os.path.abspath(lib))
    subprocess.run([sys.executable, __file__, "--child", name, str(frames), out], check=True, env=env)
    outs.append(np.load(out))
a, b = outs
diff = np.any(a.view(np.uint32) != b.view(np.uint32), axis=2)
ys, xs = np.nonzero(diff)
print("pixels that differ between the two builds: %d" % len(ys), flush=True)
if len(ys) == 0:
    sys.exit(0)
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
scene, meta = configs.build(name)
p = meta["params"]
h, w = a.shape[:2]
tiles_x = (w + 7) // 8
n_tiles = tiles_x * ((h + 7) // 8)
tr = rt.RayTracer(0, w, h)
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(rt._lib.RT_OPT_TRAVERSAL, 1)
seen = set()
for y, x in list(zip(ys, xs))[:24]:
    t = (y // 8) * tiles_x + (x // 8)
    if t in seen:
        continue
    seen.add(t)
    tr.clear(); tr.set_shard(t, n_tiles, 0, 1)
    tr.render_frames(0, frames); tr.sync()
    ref = tr.read_accum()
    ty, tx = (y // 8) * 8, (x // 8) * 8
    blk = (slice(ty, ty + 8), slice(tx, tx + 8))
    ea = np.any(ref[blk].view(np.uint32) != a[blk].view(np.uint32), axis=2).sum()
    eb = np.any(ref[blk].view(np.uint32) != b[blk].view(np.uint32), axis=2).sum()
    print("tile (%d,%d): build A differs from brute force at %d pixels, build B at %d" % (tx, ty, ea, eb), flush=True)
