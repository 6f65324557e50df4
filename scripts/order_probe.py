"""Find pixels whose value depends on the traversal schedule: re-render the same scene (same hierarchy) until two
images differ, then compare both with exhaustive brute force on the tiles involved.
  python scripts/order_probe.py c5_stress 2 [builder] [tries]"""
import sys
import numpy as np
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name, frames = sys.argv[1], int(sys.argv[2])
builder = int(sys.argv[3]) if len(sys.argv) > 3 else 1
tries = int(sys.argv[4]) if len(sys.argv) > 4 else 20
scene, meta = configs.build(name)
p = meta["params"]
w, h = p["width"], p["height"]
tr = rt.RayTracer(0, w, h)
tr.set_option(rt._lib.RT_OPT_BUILDER, builder)
tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, frames)
imgs = []
for i in range(tries):
    tr.clear(); tr.render_frames(0, frames); tr.sync()
    img = tr.read_accum()
    if not any(np.array_equal(img.view(np.uint32), o.view(np.uint32)) for o in imgs):
        imgs.append(img)
print("distinct images in %d renders of one hierarchy: %d" % (tries, len(imgs)), flush=True)
if len(imgs) < 2:
    sys.exit(0)
diff = np.zeros((h, w), bool)
for o in imgs[1:]:
    diff |= np.any(o.view(np.uint32) != imgs[0].view(np.uint32), axis=2)
ys, xs = np.nonzero(diff)
print("pixels involved:", [(int(x), int(y)) for y, x in zip(ys, xs)], flush=True)
tiles_x = (w + 7) // 8
n_tiles = tiles_x * ((h + 7) // 8)
tr.set_option(rt._lib.RT_OPT_TRAVERSAL, 1)
for y, x in zip(ys, xs):
    t = (y // 8) * tiles_x + (x // 8)
    tr.clear(); tr.set_shard(t, n_tiles, 0, 1)
    tr.render_frames(0, frames); tr.sync()
    ref = tr.read_accum()[y, x]
    print("pixel (%d,%d): brute force %s" % (x, y, ref.view(np.uint32)), flush=True)
    for k, o in enumerate(imgs):
        print("   image %d %s %s" % (k, o[y, x].view(np.uint32), "== brute force" if np.array_equal(o[y, x].view(np.uint32), ref.view(np.uint32)) else "DIFFERS"), flush=True)
