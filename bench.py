#!/usr/bin/env python
"""bench.py — headline benchmark of the path-tracing hot path (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c3_teapot]

Metric: Mrays/s (all bounces) on config C3 — the Utah teapot (15,704 triangles), 1920x1080,
256 spp (= 128 progressive dispatches at rays_per_pixel = 1), 8 bounces, GPU-built BVH
(PLOC over Morton order).  A "ray" is
one path segment = one call of calculate_ray_collions in the reference (wgsl:234), counted
on the device.  One step = the whole 256-spp render.  N > 1 (torchrun, one rank per GPU):
sample-index sharding, every rank renders its own 256 spp (frames = rank mod N, weak
scaling), images combined with one NCCL all-reduce per step inside the timed region.

`--impl reference` times the reference algorithm's CPU restatement (the oracle: brute
force over every triangle, OpenMP over rows, all host cores) on a bounded pixel sample of
the same workload; the reference's own WGSL-through-lavapipe arm is unavailable on this
box (no graphics stack, no Rust toolchain).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s (all bounces) at 1080p teapot"
CPU_STRIDE = 8          # CPU arms trace every 8th pixel in x and y (1/64 of the image)
# other configs: strides that keep a brute-force step within seconds on the box's host cores
# (C1 is six spheres: every pixel; C2 is 428 triangles: every 2nd; C5 is 1.1 M primitives per ray: every 128th)
CPU_STRIDES = {"c1_balls": 1, "c2_cube": 2, "c2_cube_plain": 2, "c5_stress": 128}

# algorithmic flops per unit (DESIGN.md "Roofline"): FMA = 2
F_BOX_PAIR = 48         # one node visit = two slab tests (2 x 24)
F_TRI = 53
F_SPH = 25
F_SHADE = 110
F_MISS = 30


_JSON_OUT = sys.stdout


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def ncu_traffic(config):
    """DRAM bytes (read + write) of one trace-kernel launch from the committed `ncu --set full`
    capture of this workload (profiles/*_traffic.json), or None when there is none."""
    try:
        with open(os.path.join(ROOT, "profiles", "trace_kernel_traffic.json")) as f:
            t = json.load(f).get(config)
        return None if t is None else {"dram_bytes_per_launch": t["dram_read_bytes"] + t["dram_write_bytes"],
                                       "dram_read_bytes": t["dram_read_bytes"], "dram_write_bytes": t["dram_write_bytes"],
                                       "dispatches_per_launch": t["dispatches_per_launch"], "source": t["source"]}
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=fd,
                                         stderr=subprocess.DEVNULL)
            os.close(fd)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        try:
            for line in open(self.path):
                t = [x.strip() for x in line.split(",")]
                if len(t) < 9:
                    continue
                try:
                    sm.append(float(t[1])); smax.append(float(t[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), t[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(smax)), reasons=sorted(reasons), samples=len(sm))
        return out


def oracle_scene(scene):
    from oracle import rt_oracle as ro
    t = scene.as_tuple()
    return (t[0].view(ro.CAMERA_DT), t[1].view(ro.SPHERE_DT), t[2].view(ro.VERTEX_DT), t[3], t[4].view(ro.MESH_DT))


def cpu_sample(scene, params, first_frame, n_dispatches, stride):
    """Time the oracle on every `stride`-th pixel in x and y. -> (Mrays/s, rays, seconds, threads)"""
    from oracle import rt_oracle as ro
    sc = oracle_scene(scene)
    ctr = np.zeros(3, np.uint64)
    try:
        threads = len(os.sched_getaffinity(0))      # all host cores, even when torchrun exports OMP_NUM_THREADS=1
    except AttributeError:
        threads = os.cpu_count() or 1
    t0 = time.perf_counter()
    ro.render(sc, params["width"], params["height"], n_dispatches, params["number_of_bounces"],
              params["rays_per_pixel"], params["skybox"], first_frame, None, ctr, stride, stride, threads)
    dt = time.perf_counter() - t0
    return float(ctr[0]) / dt / 1e6, int(ctr[0]), dt, threads


def run_reference(args, rank, world):
    """The reference algorithm on the host cores (oracle port), rank 0 only."""
    if rank != 0:
        return
    from ray_tracer_b200 import configs
    scene, meta = configs.build(args.config)
    p = meta["params"]
    CPU_STRIDE = CPU_STRIDES.get(args.config, globals()["CPU_STRIDE"])
    from oracle import rt_oracle as ro
    ro.build()
    for w in range(args.warmup):
        cpu_sample(scene, p, w, 1, CPU_STRIDE * 4)          # short warm-ups (1/1024 of the image)
    rays, secs, threads = 0, 0.0, 1
    for k in range(args.steps):
        _, r, dt, threads = cpu_sample(scene, p, k, 1, CPU_STRIDE)
        rays += r; secs += dt
    value = rays / secs / 1e6
    sample = "1 dispatch (2 spp) over every %dth pixel in x and y of %dx%d per step, brute force over %d spheres + %d triangles" % (
        CPU_STRIDE, p["width"], p["height"], int(scene.spheres.size), int(scene.meshes["triangles"].sum()))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / max(args.steps, 1) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.config, "width": p["width"], "height": p["height"], "bounces": p["number_of_bounces"],
                   "note": "reference WGSL via wgpu/lavapipe is unavailable on this box; this is the CPU restatement of the same brute-force algorithm"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=_JSON_OUT, flush=True)


class _DevArray:
    """__cuda_array_interface__ view of the context's accumulation image (no copy)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import ray_tracer_b200 as rt
    from ray_tracer_b200 import configs, _lib, sharding
    import ctypes as C

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    dev = local_rank
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", dev))

    scene, meta = configs.build(args.config)
    p, D = meta["params"], meta["dispatches"]
    W, H = p["width"], p["height"]
    sp, vs, ix, ms = scene.arrays()
    cam = scene.camera_uniform()

    tr = rt.RayTracer(dev, W, H)
    tr.upload_scene(sp, vs, ix, ms)
    tr.upload_textures(scene.textures)
    tr.set_camera(cam)
    tr.set_params(**p)
    # partition (no data-path collective): --shard sample: dispatch f -> rank f % N; tile: 8x8 tiles dealt
    # round-robin; tile+sample: T tile groups x N/T sample groups.  Per-GPU work is the same in all three.
    T = {"sample": 1, "tile": world, "tile+sample": (2 if world >= 4 else world)}[args.shard]
    F = world // T
    tr.set_shard(rank % T, T, rank // T, F)
    stream = torch.cuda.ExternalStream(tr.cuda_stream, device=dev)
    accum_t = torch.as_tensor(_DevArray(tr.accum_device_ptr, (H, W, 4)), device="cuda:%d" % dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:%d" % dev)

    def step():
        if T > 1:
            tr.clear()                           # other ranks' tiles must be zero before the sum
        tr.render_frames(0, D * world)           # this rank renders 1/T of the pixels of 1/F of the D*world frames
        if world > 1:
            cur = torch.cuda.current_stream()
            cur.wait_stream(stream)
            sharding.combine_sample_shards(accum_t, 1, F)           # one NCCL all-reduce over NVLink: sum of disjoint tiles, count-weighted mean of sample groups
            stream.wait_stream(cur)

    def barrier():
        tr.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # FP32 roofline denominator + per-step algorithmic work (deterministic: same frames every step)
    peak_tf, peak_s = C.c_double(), C.c_double()
    if _lib.lib().rt_measure_fp32_peak(dev, 5, C.byref(peak_tf), C.byref(peak_s)) != 0:
        raise SystemExit("rt_measure_fp32_peak failed")
    tr.set_option(_lib.RT_OPT_DETAIL_COUNTERS, 1)
    tr.reset_counters()
    if T > 1:
        tr.clear()
    tr.render_frames(0, D * world); tr.sync()
    det = tr.counters()
    tr.set_option(_lib.RT_OPT_DETAIL_COUNTERS, 0)
    flops_step = (det["node_visits"] * F_BOX_PAIR + det["tri_tests"] * F_TRI + det["sphere_tests"] * F_SPH +
                  det["hit_segments"] * F_SHADE + (det["rays"] - det["hit_segments"]) * (F_MISS if p["skybox"] else 0))

    for _ in range(max(args.warmup, 0)):
        step()
    barrier()

    sampler = ClockSampler(dev)
    if rank == 0:
        sampler.start()
    tr.reset_counters()
    ms_total = 0.0
    for _ in range(args.steps):
        flush.zero_()                            # L2 flush between timed steps (untimed)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        step()
        e1.record(torch.cuda.current_stream() if world > 1 else stream)
        barrier()
        ms_total += e0.elapsed_time(e1)
    ctr = tr.counters()
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda:%d" % dev)
    rays = torch.tensor([float(ctr["rays"])], dtype=torch.float64, device="cuda:%d" % dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    ms_total, rays_total = float(t.item()), float(rays.item())
    value = rays_total / (ms_total * 1e-3) / 1e6

    # ---- end to end through the C ABI with HOST buffers (pinned): upload + build + render + read-back
    def pinned(a):
        tt = torch.empty(max(a.nbytes, 1), dtype=torch.uint8).pin_memory()
        v = tt.numpy()[:a.nbytes].view(a.dtype).reshape(a.shape)
        v[...] = a
        return tt, v
    keep = [pinned(x) for x in (sp, vs, ix, ms)]
    hsp, hvs, hix, hms = [k[1] for k in keep]
    out_t = torch.empty(H * W * 4, dtype=torch.float32).pin_memory()
    out = out_t.numpy().reshape(H, W, 4)
    h2d = int(sp.nbytes + vs.nbytes + ix.nbytes + ms.nbytes + 128 + 28)
    d2h = int(out.nbytes) if rank == 0 else 0

    def e2e_step():
        tr.upload_scene(hsp, hvs, hix, hms)
        tr.set_camera(cam)
        tr.set_params(**p)
        step()
        if rank == 0:
            tr.read_accum(out)
        else:
            tr.sync()

    e2e_step()
    barrier()
    tr.reset_counters()
    e2e_steps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_rays = float(tr.counters()["rays"])
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda:%d" % dev)
    r2 = torch.tensor([e2e_rays], dtype=torch.float64, device="cuda:%d" % dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(r2, op=dist.ReduceOp.SUM)
    e2e_value = float(r2.item()) / float(t.item()) / 1e6

    if rank == 0:
        pk, pk_src = peaks()
        launches_per_step = ctr["launches"] / max(args.steps, 1)
        launch_s = ctr["kernel_ms"] * 1e-3 / max(ctr["launches"], 1)
        achieved = flops_step / launches_per_step / launch_s / 1e12
        hbm_bytes = 32.0 * W * H                  # accumulation read + write per launch
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.config, "width": W, "height": H, "spp_per_gpu": meta["spp"], "dispatches_per_gpu": D,
                       "bounces": p["number_of_bounces"], "rays_per_pixel": p["rays_per_pixel"], "skybox": p["skybox"],
                       "triangles": int(ms["triangles"].sum()), "spheres": int(sp.size), "traversal": "bvh2 (PLOC over 63-bit Morton order, GPU-built)", "textures": len(scene.textures),
                       "parallelism": ("%s sharding (%d tile groups x %d sample groups) + one NCCL all-reduce per step" % (args.shard, T, F)) if world > 1 else "single GPU",
                       "l2": "256 MiB write between timed steps (untimed); scene+BVH are L2-resident by design",
                       "spp_per_s": meta["spp"] * world / (ms_total / max(args.steps, 1) * 1e-3)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "what": "rt_upload_scene (host records -> SoA + LBVH build) + rt_render_frames + rt_read_accum, wall clock"},
            "gpu_launches": int(ctr["launches"]),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak_tf.value, "unit": "TFLOP/s",
                         "frac": achieved / peak_tf.value if peak_tf.value else None,
                         "traffic": (ncu_traffic(args.config) or {}).get("dram_bytes_per_launch"), "traffic_detail": ncu_traffic(args.config),
                         "peak_source": "FFMA microbenchmark run in this process (rt_measure_fp32_peak); theoretical 148x128x2x%.0f MHz = %.1f" % (
                             pk.get("sm_max_mhz", 1965.0), 148 * 128 * 2 * pk.get("sm_max_mhz", 1965.0) * 1e6 / 1e12),
                         "kernel": "rt::trace_kernel<BRUTE=false,DETAIL=false,TEX=%s,PF=%s>" % ("true" if scene.textures else "false", "true" if (int(ms["triangles"].sum()) + int(sp.size)) <= 262144 else "false"), "launch_ms": launch_s * 1e3,
                         "flops_per_launch": flops_step / launches_per_step,
                         "per_ray": {"node_visits": det["node_visits"] / det["rays"], "tri_tests": det["tri_tests"] / det["rays"],
                                     "sphere_tests": det["sphere_tests"] / det["rays"], "hit_fraction": det["hit_segments"] / det["rays"]},
                         "hbm": {"algorithmic_bytes_per_launch": hbm_bytes, "achieved_gbs": hbm_bytes / launch_s / 1e9,
                                 "peak_gbs": pk.get("hbm_gbs"), "peak_kind": pk_src}},
        }
        if world == 1:
            from oracle import rt_oracle as ro
            ro.build()
            CPU_STRIDE = CPU_STRIDES.get(args.config, globals()["CPU_STRIDE"])
            cpu_sample(scene, p, 0, 1, CPU_STRIDE * 4)
            nd = D if args.config == "c1_balls" else 2          # C1 is CPU-runnable in full (8 dispatches = 16 spp)
            v, r, dt, threads = cpu_sample(scene, p, 0, nd, CPU_STRIDE)
            line["cpu_baseline"] = {"value": v, "unit": "Mrays/s", "cores": threads, "kind": "port",
                                    "sample": "%d dispatches (%d spp) over every %dth pixel in x and y (%d rays, %.1f s), brute force" % (nd, 2 * nd, CPU_STRIDE, r, dt)}
        print(json.dumps(line), file=_JSON_OUT, flush=True)
    tr.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c3_teapot")
    ap.add_argument("--shard", default="sample", choices=["sample", "tile", "tile+sample"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    global METRIC
    if args.config != "c3_teapot":
        METRIC = "Mrays/s (all bounces) at %s" % args.config      # BASELINE.json's metric is quoted on C3; other configs are labelled as such
    # stdout carries the JSON line and nothing else: libraries that write to file descriptor 1 on their own
    # (NCCL prints its version there under torchrun) are pointed at stderr while the benchmark runs
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
