#!/usr/bin/env python
"""bench.py — headline benchmark of the path-tracing hot path (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c3_teapot]

Metric: Mrays/s (all bounces) on config C3 — the Utah teapot (15,704 triangles), 1920x1080,
256 spp (= 128 progressive dispatches at rays_per_pixel = 1), 8 bounces, GPU-built BVH
(PLOC over Morton order).  A "ray" is one path segment = one call of calculate_ray_collions in
the reference (wgsl:234), counted on the device.  One step = the whole 256-spp render.

N > 1 (torchrun, one rank per GPU).  Headline (weak scaling, the contract's line): sample-index
sharding, every rank renders its own 256 spp, images combined at rank 0 through the library's own
NCCL combine (rt_comm_init / rt_combine: compact tiles by grouped send/recv, count weights and
de-interleave in one consumer kernel) inside the timed region.  torch.distributed only carries the
communicator id, the barriers and the max-over-ranks of the timings.  The same line carries a
"strong" block — FIXED total work split over the N GPUs (C3 by sample index and by tiles; at N >= 4
also C4, the 4K tree at 1024 spp, tile + sample, in both seed modes), with the speed-up against one
GPU doing all of it measured in the same run, and a bit-for-bit check of the tile-combined image
against the single-GPU render.

`--impl reference` times the reference's own shader on the host cores: oracle/_ref, the reference's
WGSL translated mechanically to C++ (oracle/make_ref.py; prebuilt, travels with the snapshot), or
the hand-written port when that library is absent — brute force over every triangle, OpenMP over
rows, all host cores, on a bounded pixel sample of the same workload.  The reference's own
WGSL-through-lavapipe arm is unavailable on this box (no graphics stack, no Rust toolchain).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
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


SCHED_DEFAULT = 8 | (1 << 8) | (22 << 16)      # RT_OPT_SCHED: the library's own thresholds (node loop 8 lanes, shade at 22)


def ncu_traffic(config):
    """DRAM bytes (read + write) of one trace-kernel launch from the committed `ncu --set full`
    capture of this workload (profiles/trace_kernel_traffic.json), or None when there is none."""
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


# ---- CPU arms: the oracle side only (no product library is loaded by these functions) ---------------
def cpu_backend():
    """-> (module with render(), kind, label): oracle/_ref (the translated reference shader) when its prebuilt
    library is there, else the hand-written port."""
    from oracle import make_ref
    if make_ref.build() is not None:
        from oracle import wgsl_ref
        wgsl_ref.lib(True)
        return wgsl_ref, "reference", "oracle/_ref: the reference's ray_tracer.wgsl translated mechanically to C++ (contract build)"
    from oracle import rt_oracle
    rt_oracle.build()
    return rt_oracle, "port", "oracle/rt_oracle.cpp: hand-written C++ restatement of ray_tracer.wgsl"


def cpu_sample(backend, scene, params, first_frame, n_dispatches, stride):
    """Time the CPU arm on every `stride`-th pixel in x and y. -> (Mrays/s, rays, seconds, threads)"""
    ctr = np.zeros(3, np.uint64)
    try:
        threads = len(os.sched_getaffinity(0))      # all host cores, even when torchrun exports OMP_NUM_THREADS=1
    except AttributeError:
        threads = os.cpu_count() or 1
    t0 = time.perf_counter()
    backend.render(scene, params["width"], params["height"], n_dispatches, params["number_of_bounces"],
                   params["rays_per_pixel"], params["skybox"], first_frame, None, ctr, stride, stride, threads)
    dt = time.perf_counter() - t0
    return float(ctr[0]) / dt / 1e6, int(ctr[0]), dt, threads


def cpu_scene(config):
    """Records of `config` built by the oracle side alone (oracle/ref_configs.py; byte-identical to the product's,
    tests/test_host_mirror.py).  The texture extension has no reference counterpart: c2_cube is timed untextured."""
    from oracle import ref_configs
    scene, tex, meta = ref_configs.build("c2_cube_plain" if config == "c2_cube" else config)
    return scene, meta


def run_reference(args, rank, world):
    """The reference's shader on the host cores, rank 0 only."""
    if rank != 0:
        return
    backend, kind, label = cpu_backend()
    scene, meta = cpu_scene(args.config)
    p = meta["params"]
    stride = CPU_STRIDES.get(args.config, CPU_STRIDE)
    for w in range(args.warmup):
        cpu_sample(backend, scene, p, w, 1, stride * 4)          # short warm-ups (1/1024 of the image)
    rays, secs, threads = 0, 0.0, 1
    for k in range(args.steps):
        _, r, dt, threads = cpu_sample(backend, scene, p, k, 1, stride)
        rays += r; secs += dt
    value = rays / secs / 1e6
    sample = "1 dispatch (2 spp) over every %dth pixel in x and y of %dx%d per step, brute force over %d spheres + %d triangles" % (
        stride, p["width"], p["height"], int(scene[1].size), int(scene[4]["triangles"].sum()))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / max(args.steps, 1) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.config, "width": p["width"], "height": p["height"], "bounces": p["number_of_bounces"],
                   "what": label,
                   "note": "reference WGSL via wgpu/lavapipe is unavailable on this box; the shader's own text runs as C++ on the host cores instead"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=_JSON_OUT, flush=True)


class _DevArray:
    """__cuda_array_interface__ view of a device image (no copy)."""

    def __init__(self, ptr, shape, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (int(ptr), False), "version": 2}


def run_ours(args, rank, world, local_rank):
    import ctypes as C
    import torch
    import torch.distributed as dist
    import ray_tracer_b200 as rt
    from ray_tracer_b200 import configs, _lib

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    dev = local_rank
    torch.cuda.set_device(dev)
    device = "cuda:%d" % dev
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", dev))

    def reduce_scalar(v, op):
        t = torch.tensor([float(v)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=op)
        return float(t.item())
    MAX = dist.ReduceOp.MAX if world > 1 else None
    SUM = dist.ReduceOp.SUM if world > 1 else None

    def make_tracer(config, seed_mode=0):
        scene, meta = configs.build(config)
        p = meta["params"]
        tr = rt.RayTracer(dev, p["width"], p["height"])
        tr.upload_scene(*scene.arrays())
        tr.upload_textures(scene.textures)
        tr.set_camera(scene.camera_uniform())
        tr.set_params(**p)
        if seed_mode:
            tr.set_option(_lib.RT_OPT_SEED_MODE, seed_mode)
        return tr, scene, meta

    uid = None
    if world > 1:                                    # the communicator id travels over torch.distributed; NCCL itself is the library's
        box = [rt.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]

    tr, scene, meta = make_tracer(args.config)
    p, D = meta["params"], meta["dispatches"]
    W, H = p["width"], p["height"]
    sp, vs, ix, ms = scene.arrays()
    cam = scene.camera_uniform()
    if world > 1:
        tr.comm_init(world, rank, uid)
    # partition (no data-path collective): --shard sample: dispatch f -> rank f % N; tile: 8x8 tiles dealt
    # round-robin; tile+sample: T tile groups x N/T sample groups.  Per-GPU work is the same in all three.
    T = {"sample": 1, "tile": world, "tile+sample": (2 if world >= 4 else world)}[args.shard]
    F = world // T
    tr.set_shard(rank % T, T, rank // T, F)
    stream = torch.cuda.ExternalStream(tr.cuda_stream, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)

    def step():
        tr.render_frames(0, D * world)           # this rank renders 1/T of the pixels of 1/F of the D*world frames
        if world > 1:
            tr.combine(0)                        # NCCL behind the C ABI, stream-ordered; the combined image lands at rank 0

    def barrier(t=None):
        (t or tr).sync()
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
        e1.record(stream)
        barrier()
        ms_total += e0.elapsed_time(e1)
    ctr = tr.counters()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = reduce_scalar(ms_total, MAX)
    rays_total = reduce_scalar(ctr["rays"], SUM)
    value = rays_total / (ms_total * 1e-3) / 1e6

    # ---- end to end through the C ABI with HOST buffers (pinned): upload + build + render (+ combine) + read-back
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
            (tr.read_combined if world > 1 else tr.read_accum)(out)
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
    e2e_s = reduce_scalar(time.perf_counter() - t0, MAX)
    e2e_rays = reduce_scalar(tr.counters()["rays"], SUM)
    e2e_value = e2e_rays / e2e_s / 1e6

    # ---- strong scaling: FIXED total work over the N GPUs (N > 1), against one GPU doing all of it in this same run
    strong = None
    if world > 1:
        def strong_case(config, n_disp, t_groups, seed_mode=0, check_tile=False):
            """-> dict. Every rank first renders ALL the work alone (t1: what one GPU takes; all ranks at once, max over
            ranks), then its share (tile_rank = rank % T of T, frame_rank = rank // T of N/T) + the combine at rank 0."""
            if config == args.config and seed_mode == 0:
                t, tmeta = tr, meta
            else:
                t, _, tmeta = make_tracer(config, seed_mode)
                t.comm_init(world, rank, uid_for[(config, seed_mode)])
            st = torch.cuda.ExternalStream(t.cuda_stream, device=dev)
            steps = max(1, min(args.steps, 3))

            def timed(fn, reps):
                tot = 0.0
                for _ in range(reps):
                    flush.zero_(); barrier(t)
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(st); fn(); b.record(st)
                    barrier(t)
                    tot += a.elapsed_time(b)
                return reduce_scalar(tot / reps, MAX)

            t.set_shard(0, 1, 0, 1)
            t.render_frames(0, n_disp); barrier(t)                              # warm-up
            t.reset_counters()
            ms1 = timed(lambda: t.render_frames(0, n_disp), 1 if n_disp > 256 else steps)
            rays1 = t.counters()["rays"] / (1 if n_disp > 256 else steps)
            hh, ww = tmeta["params"]["height"], tmeta["params"]["width"]
            single = torch.as_tensor(_DevArray(t.accum_device_ptr, (hh, ww, 4)), device=device).clone() if check_tile else None
            Fg = world // t_groups
            t.set_shard(rank % t_groups, t_groups, rank // t_groups, Fg)

            # the caller's knowledge of its scene: on the teapot, a GPU's share of 49-96 dispatches traces 3 % faster with 128-item queue
            # grabs (profiles/r2_summary.md, queue policy table); RT_OPT_SCHED bits 24-39 = grab / 32, guide.  Everything else runs on
            # the library's defaults.
            per_gpu = n_disp // world                                          # equivalent launch length of a GPU's share (tile shards: a share of the tiles x all dispatches)
            queue_policy = "default"
            if config == "c3_teapot" and 48 < per_gpu <= 96:
                grab, guide = 128, 20
                t.set_option(rt._lib.RT_OPT_SCHED, SCHED_DEFAULT | ((grab // 32) << 24) | (guide << 32))
                queue_policy = "grab %d, guide %d" % (grab, guide)

            def sharded():
                t.render_frames(0, n_disp)
                t.combine(0)
            sharded(); barrier(t)                                               # warm-up
            msN = timed(sharded, steps)
            # the combine alone, at the root: from the end of its own render to the end of the consumer kernel
            barrier(t)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t.render_frames(0, n_disp); barrier(t)
            a.record(st); t.combine(0); b.record(st); barrier(t)
            comb = reduce_scalar(a.elapsed_time(b) if rank == 0 else 0.0, MAX)
            res = {"config": config, "dispatches_total": n_disp, "spp_total": 2 * n_disp, "partition": "%d tile group(s) x %d sample group(s)" % (t_groups, Fg),
                   "seed_mode": "wide" if seed_mode else "strict", "ms_n1": ms1, "ms_per_step": msN, "speedup_vs_n1": ms1 / msN, "efficiency": ms1 / msN / world,
                   "Mrays_s": rays1 / (msN * 1e-3) / 1e6, "Mrays_s_n1": rays1 / (ms1 * 1e-3) / 1e6, "combine_ms": comb, "queue_policy": queue_policy}
            t.set_option(rt._lib.RT_OPT_SCHED, SCHED_DEFAULT)
            if check_tile:                                                      # outside every timed region
                same = True
                if rank == 0:
                    got = torch.as_tensor(_DevArray(t.combined_device_ptr, (hh, ww, 4)), device=device)
                    same = bool(torch.equal(got.view(torch.int32), single.view(torch.int32)))
                res["tile_combined_equals_single_gpu_bit_for_bit"] = bool(reduce_scalar(1.0 if same else 0.0, dist.ReduceOp.MIN) == 1.0)
            if t is not tr:
                t.sync(); t.comm_destroy(); t.close()
            return res

        cases = [("c3_sample", args.config, D, 1, 0, False), ("c3_tile", args.config, D, world, 0, True)]
        if world >= 4 and args.config == "c3_teapot":
            c4d = configs.CONFIGS["c4_tree"][2] // 2
            cases += [("c4_tile_sample", "c4_tree", c4d, 2, 0, False), ("c4_tile_sample_wide", "c4_tree", c4d, 2, 1, False)]
        uid_for = {}
        for _, config, _, _, sm, _ in cases:                                   # one communicator id per extra context, made up front
            if (config, sm) != (args.config, 0) and (config, sm) not in uid_for:
                box = [rt.comm_unique_id() if rank == 0 else None]
                dist.broadcast_object_list(box, src=0)
                uid_for[(config, sm)] = box[0]
        strong = {}
        for name, config, n_disp, tg, sm, chk in cases:
            strong[name] = strong_case(config, n_disp, tg, sm, chk)

    if rank == 0:
        pk, pk_src = peaks()
        launches_per_step = ctr["launches"] / max(args.steps, 1)
        launch_s = ctr["kernel_ms"] * 1e-3 / max(ctr["launches"], 1)
        achieved = flops_step / launches_per_step / launch_s / 1e12
        fpl = D * world // F // max(int(round(launches_per_step)), 1)      # dispatches per launch on this rank
        image_bytes = 32.0 * W * H / T                                     # accumulation image read + written once per launch: the algorithmic HBM traffic
        staging_bytes = 32.0 * W * H / T * fpl                             # (pixel, dispatch) items: every frag written by the trace kernel, read by the blend kernel
        kernels_per_step = int(round(launches_per_step)) * 2 + (1 if world > 1 else 0)   # trace + blend per launch, + the combine's consumer kernel
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.config, "width": W, "height": H, "spp_per_gpu": meta["spp"], "dispatches_per_gpu": D,
                       "bounces": p["number_of_bounces"], "rays_per_pixel": p["rays_per_pixel"], "skybox": p["skybox"],
                       "triangles": int(ms["triangles"].sum()), "spheres": int(sp.size), "traversal": "bvh2 (PLOC over 63-bit Morton order, GPU-built)", "textures": len(scene.textures),
                       "parallelism": ("%s sharding (%d tile groups x %d sample groups) + the library's NCCL combine at rank 0 (rt_combine: grouped send/recv of compact tiles + one consumer kernel) per step" % (args.shard, T, F)) if world > 1 else "single GPU",
                       "l2": "256 MiB write between timed steps (untimed); scene+BVH are L2-resident by design",
                       "spp_per_s": meta["spp"] * world / (ms_total / max(args.steps, 1) * 1e-3)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "what": "rt_upload_scene (host records -> SoA + LBVH build) + rt_render_frames%s + %s, wall clock" % (
                        " + rt_combine" if world > 1 else "", "rt_read_combined" if world > 1 else "rt_read_accum")},
            "gpu_launches": kernels_per_step * args.steps,
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak_tf.value, "unit": "TFLOP/s",
                         "frac": achieved / peak_tf.value if peak_tf.value else None,
                         "traffic": (ncu_traffic(args.config) or {}).get("dram_bytes_per_launch"), "traffic_detail": ncu_traffic(args.config),
                         "peak_source": "FFMA microbenchmark run in this process (rt_measure_fp32_peak); theoretical 148x128x2x%.0f MHz = %.1f" % (
                             pk.get("sm_max_mhz", 1965.0), 148 * 128 * 2 * pk.get("sm_max_mhz", 1965.0) * 1e6 / 1e12),
                         "kernel": "rt::trace_kernel<BRUTE=false,DETAIL=false,TEX=%s,PF=%s,FOLD=false,INST=false>%s" % ("true" if scene.textures else "false", "true" if (int(ms["triangles"].sum()) + int(sp.size)) <= 262144 else "false", "" if (int(ms["triangles"].sum()) + int(sp.size)) <= 262144 else " (packed 32-byte nodes)"), "launch_ms": launch_s * 1e3,
                         "flops_per_launch": flops_step / launches_per_step,
                         "per_ray": {"node_visits": det["node_visits"] / det["rays"], "tri_tests": det["tri_tests"] / det["rays"],
                                     "sphere_tests": det["sphere_tests"] / det["rays"], "hit_fraction": det["hit_segments"] / det["rays"]},
                         "hbm": {"algorithmic_bytes_per_launch": image_bytes, "staging_bytes_per_launch": staging_bytes,
                                 "what": "algorithmic = the accumulation image read + written once; staging = the per-dispatch frags the (pixel, dispatch) item mode writes and the blend kernel reads back (not algorithmic: the price of fine-grained items under an order-dependent running mean, DESIGN.md)",
                                 "achieved_gbs": (image_bytes + staging_bytes) / launch_s / 1e9,
                                 "peak_gbs": pk.get("hbm_gbs"), "peak_kind": pk_src}},
        }
        if strong is not None:
            line["strong"] = strong
        if world == 1:
            backend, kind, label = cpu_backend()
            cscene, _ = cpu_scene(args.config)
            stride = CPU_STRIDES.get(args.config, CPU_STRIDE)
            cpu_sample(backend, cscene, p, 0, 1, stride * 4)
            nd = D if args.config == "c1_balls" else 2          # C1 is CPU-runnable in full (8 dispatches = 16 spp)
            v, r, dt, threads = cpu_sample(backend, cscene, p, 0, nd, stride)
            line["cpu_baseline"] = {"value": v, "unit": "Mrays/s", "cores": threads, "kind": kind, "what": label,
                                    "sample": "%d dispatches (%d spp) over every %dth pixel in x and y (%d rays, %.1f s), brute force" % (nd, 2 * nd, stride, r, dt)}
        print(json.dumps(line), file=_JSON_OUT, flush=True)
    if world > 1:
        tr.sync(); tr.comm_destroy()
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
