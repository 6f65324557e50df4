"""RT_OPT_TILE_ORDER (heaviest tiles of the previous call first) on / off against launch length:
python scripts/order_probe.py c3_teapot "128,64,32,16,8,4,2,1" ["chunk,guide;..."]   -> ms per dispatch, image CRCs must agree"""
import sys, zlib
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs
name = sys.argv[1]
lengths = [int(x) for x in sys.argv[2].split(",")]
combos = [tuple(int(v) for v in c.split(",")) for c in (sys.argv[3] if len(sys.argv) > 3 else "0,0").split(";")]
scene, meta = configs.build(name)
p = meta["params"]
base = 8 | (1 << 8) | (22 << 16)
print("%-22s" % "order chunk,guide", *["%9d" % n for n in lengths], " (ms per dispatch)")
for order in (0, 1):
    for chunk, guide in combos:
        tr = rt.RayTracer(0, p["width"], p["height"])
        tr.set_option(rt._lib.RT_OPT_TILE_ORDER, order)
        tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
        tr.set_option(rt._lib.RT_OPT_STAGING_MIB, 8192)
        tr.set_option(rt._lib.RT_OPT_SCHED, base | ((chunk // 32) << 24) | (guide << 32))
        row = []; crcs = []
        for n in lengths:
            tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, n)
            best = 1e9
            for rep in range(4):
                tr.clear(); tr.reset_counters(); tr.render_frames(0, n); tr.sync()
                if rep:
                    best = min(best, tr.counters()["kernel_ms"])
            row.append(best / n); crcs.append(zlib.crc32(tr.read_accum().tobytes()))
        print("%-22s" % ("%d     %d,%d" % (order, chunk, guide)), *["%9.4f" % v for v in row], " crc", "%08x" % (zlib.crc32(str(crcs).encode())), flush=True)
        tr.close()
