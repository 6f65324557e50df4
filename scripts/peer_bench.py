"""Single-process multi-GPU run of a config through the C ABI alone (no torch.distributed): one context per GPU,
rt_set_shard, asynchronous rt_render_frames on every device, then ONE rt_combine_peers kernel on GPU 0 that reads
the other GPUs' images over NVLink peer memory.  Weak scaling like bench.py (every sample shard renders the config's
full spp).  Prints one JSON line; wall-clock timed around sync points (device work of all GPUs overlaps).

    python scripts/peer_bench.py --gpus 2 [--config c3_teapot] [--shard sample|tile|tile+sample] [--steps 3]
"""
import argparse, json, sys, time
import numpy as np
sys.path.insert(0, ".")
import ray_tracer_b200 as rt
from ray_tracer_b200 import configs

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=2)
ap.add_argument("--config", default="c3_teapot")
ap.add_argument("--shard", default="sample", choices=["sample", "tile", "tile+sample"])
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--warmup", type=int, default=2)
ap.add_argument("--check", action="store_true", help="compare tile sharding with a single-GPU render, bit for bit")
a = ap.parse_args()

scene, meta = configs.build(a.config)
p, D = meta["params"], meta["dispatches"]
N = a.gpus
T = {"sample": 1, "tile": N, "tile+sample": (2 if N >= 4 else N)}[a.shard]
F = N // T
trs = []
for k in range(N):
    tr = rt.RayTracer(k, p["width"], p["height"])
    tr.upload_scene(*scene.arrays()); tr.upload_textures(scene.textures); tr.set_camera(scene.camera_uniform()); tr.set_params(**p)
    tr.set_shard(k % T, T, k // T, F)
    trs.append(tr)


def step():
    for tr in trs:
        if T > 1:
            tr.clear()
        tr.render_frames(0, D * N)          # asynchronous: every GPU works on its shard at once
    rt.combine_peers(trs)                   # one kernel on GPU 0, peer loads over NVLink, ordered after every stream


for _ in range(a.warmup):
    step()
for tr in trs:
    tr.sync()
for tr in trs:
    tr.reset_counters()
t0 = time.perf_counter()
for _ in range(a.steps):
    step()
for tr in trs:
    tr.sync()
dt = time.perf_counter() - t0
rays = sum(tr.counters()["rays"] for tr in trs)
# the combine alone
trs[0].sync(); t1 = time.perf_counter()
for _ in range(10):
    rt.combine_peers(trs)
trs[0].sync(); t_comb = (time.perf_counter() - t1) / 10
line = {"what": "single-process peer combine (rt_combine_peers)", "config": a.config, "n_gpus": N, "shard": "%s (%d tile x %d sample)" % (a.shard, T, F),
        "steps": a.steps, "ms_per_step": dt / a.steps * 1e3, "value": rays / dt / 1e6, "unit": "Mrays/s", "combine_ms": t_comb * 1e3,
        "image_mb": p["width"] * p["height"] * 16 / 1e6}
if a.check:
    one = rt.RayTracer(0, p["width"], p["height"])
    one.upload_scene(*scene.arrays()); one.upload_textures(scene.textures); one.set_camera(scene.camera_uniform()); one.set_params(**p)
    one.render_frames(0, D * N); ref = one.read_accum(); one.close()
    got = trs[0].read_combined()
    same = got.view(np.uint32) == ref.view(np.uint32)
    line["max_abs_diff_vs_one_gpu"] = float(np.max(np.abs(np.where(same, 0.0, got.astype(np.float64) - ref))))
    line["bit_identical_to_one_gpu"] = bool(np.array_equal(got.view(np.uint32), ref.view(np.uint32)))
print(json.dumps(line))
for tr in trs:
    tr.close()
