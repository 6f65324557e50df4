"""Headless stand-in for the reference's `run()` loop (lib.rs:10-70): build a scene, redraw N times, write what the
window would show.

    python -m ray_tracer_b200.render --scene balls --frames 64 --out balls.png
    python -m ray_tracer_b200.render --model teapot.glb --graph --frames 128 --size 1920 1080 --sky --gpus 8

--scene: built-in scene by name (context.rs:261-278).  --model: a file through load_model (upstream's placement
rules) or, with --graph, through the scene-graph import (node transforms, materials, embedded textures), lit by the
sky (texture handles are left unbound here: colours only).  --turntable DEG with --positions N (model scenes): the
camera orbits the model by DEG between images; every camera change restarts the accumulation, as any input event
does upstream (context.rs:143-175); one PNG per position (out_000.png, ...).  --gpus N: one context per
device in this process, tiles dealt round-robin, rt_combine_peers at read-out (same image as one GPU, bit for bit).
Prints one JSON line per image (frames, rays, Mrays/s of the trace kernel).
"""
import argparse
import json
import math
import os
import time

import numpy as np

from . import host

SCENES = {"balls": 0, "random_balls": 1, "room": 2, "metal": 3}


def _model_scene(path, graph, aspect, angle_deg=0.0):
    vs, ix, ms = (host.load_scene_glb if graph else host.load_model)(path)
    if len(vs) == 0:
        raise SystemExit("render: no geometry in %s" % path)
    idx = np.concatenate([ix[int(m["first"]):int(m["first"]) + 3 * int(m["triangles"])].astype(np.int64) + int(m["offset"]) for m in ms])
    pos = vs["pos"][idx]
    lo, hi = pos.min(0), pos.max(0)
    for m in ms:                                     # upstream's loader spreads meshes along x through Mesh.pos
        lo, hi = np.minimum(lo, lo + m["pos"]), np.maximum(hi, hi + m["pos"])
    centre, radius = (lo + hi) / 2, float(np.linalg.norm(hi - lo)) / 2
    t = math.radians(angle_deg)
    eye = centre + np.float32([math.sin(t), 0.35, math.cos(t)]) * np.float32(1.6 * radius / math.tan(math.radians(22.5)))
    cam = host.Camera(eye, centre, [0, 1, 0], 45.0, aspect, 0.1, 100.0, 0.0, 0.1)
    return host.Scene(cam, np.zeros(0, host.SPHERE_DT), vs, ix, ms)


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m ray_tracer_b200.render")
    g = ap.add_mutually_exclusive_group()
    g.add_argument("--scene", choices=sorted(SCENES), default="balls")
    g.add_argument("--model")
    ap.add_argument("--graph", action="store_true", help="import the glTF scene graph (transforms, materials, textures)")
    ap.add_argument("--size", type=int, nargs=2, default=[800, 800], metavar=("W", "H"))
    ap.add_argument("--frames", type=int, default=32, help="redraws per image (2 samples per pixel each at the default rays_per_pixel = 1)")
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--rpp", type=int, default=1)
    ap.add_argument("--sky", action="store_true")
    ap.add_argument("--turntable", type=float, default=0.0, metavar="DEG")
    ap.add_argument("--positions", type=int, default=1)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--mirror-x", action="store_true", help="flip like the reference's blit quad (renderer.rs:35-38)")
    ap.add_argument("--out", default="render.png")
    a = ap.parse_args(argv)
    W, H = a.size
    shards = []
    for k in range(a.gpus):
        c = host.Context(k, W, H)
        c.set(number_of_bounces=a.bounces, rays_per_pixel=a.rpp, skybox=int(a.sky or a.model is not None))
        if a.model:
            c.set_scene(_model_scene(a.model, a.graph, np.float32(W) / np.float32(H)))
        else:
            c.set_scene(SCENES[a.scene])
        if a.gpus > 1:
            c.set_shard(tile_rank=k, tile_world=a.gpus)
        c.clear_accululation()                    # frames = -1: the first redraw is frame 0 and overwrites (context.rs:144, wgsl:64-66)
        shards.append(c)
    out = []
    for pos in range(max(a.positions, 1)):
        if pos and a.turntable and a.model:
            for c in shards:                      # a camera change: the accumulation restarts (context.rs:143-175)
                c.set_scene(_model_scene(a.model, a.graph, np.float32(W) / np.float32(H), pos * a.turntable))
                c.clear_accululation()
        t0 = time.perf_counter()
        for c in shards:
            c.run_frames(a.frames)
        if a.gpus > 1:
            host.Context.combine(shards)
        for c in shards:
            c.sync()
        dt = time.perf_counter() - t0
        name = a.out if a.positions <= 1 else "%s_%03d%s" % (os.path.splitext(a.out)[0], pos, os.path.splitext(a.out)[1] or ".png")
        shards[0].save_png(name, a.mirror_x)
        line = {"image": name, "width": W, "height": H, "frames": a.frames, "spp": a.frames * (a.rpp + 1), "gpus": a.gpus, "seconds": dt}
        print(json.dumps(line))
        out.append(line)
    for c in shards:
        c.close()
    return out


if __name__ == "__main__":
    main()
