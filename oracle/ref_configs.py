"""The five benchmark configurations (BASELINE.json) as oracle-side records — TEST INFRASTRUCTURE.

A numpy restatement of ray_tracer_b200/configs.py on top of oracle/host_ref.py, so that the CPU arms
(bench.py --impl reference, cpu_baseline) can build their scenes WITHOUT loading the product libraries
(librt_b200.so / librt_host.so).  tests/test_host_mirror.py requires the records to be byte-identical
to the product's for every config.  Geometry comes from the committed fixtures (tests/golden/*.npz).
"""
import os

import numpy as np

from . import host_ref as hr
from .rt_oracle import MESH_DT, SPHERE_DT, VERTEX_DT

FIXTURES = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

# name -> (width, height, spp, bounces, skybox)   (== ray_tracer_b200.configs.CONFIGS)
CONFIGS = {
    "c1_balls": (800, 800, 16, 4, 0),
    "c2_cube": (1920, 1080, 64, 8, 1),
    "c2_cube_plain": (1920, 1080, 64, 8, 1),
    "c3_teapot": (1920, 1080, 256, 8, 1),
    "c4_tree": (3840, 2160, 1024, 8, 1),
    "c5_stress": (3840, 2160, 64, 8, 1),
}


def _fixture(name, uv=False):
    z = np.load(os.path.join(FIXTURES, name + ".npz"))
    vs = np.zeros(len(z["pos"]), VERTEX_DT)
    vs["pos"], vs["normal"] = z["pos"], z["normal"]
    if uv and "uv" in z:
        vs["tu"], vs["tv"] = z["uv"][:, 0], z["uv"][:, 1]
    ms = [hr._loader_mesh(k, int(z["triangles"][k]) * 3, int(z["first"][k]), int(z["offset"][k])) for k in range(len(z["first"]))]
    return vs, z["indices"].astype(np.uint32), np.concatenate(ms)


def _pcg_stream(seed, n):
    """n values of the shader's PCG hash stream (wgsl:205-210), top 24 bits -> [0,1) f32 (scalar loop; n is 10^6)."""
    a, c, M = 747796405, 2891336453, 0xFFFFFFFF
    k = np.arange(1, n + 1, dtype=np.uint64)
    mul, inc = np.ones(n, np.uint64), np.zeros(n, np.uint64)
    cur_m, cur_c, bit = np.uint64(a), np.uint64(c), 0
    Mu = np.uint64(M)
    while (1 << bit) <= n:
        sel = ((k >> np.uint64(bit)) & np.uint64(1)).astype(bool)
        inc[sel] = (inc[sel] * cur_m + cur_c) & Mu
        mul[sel] = (mul[sel] * cur_m) & Mu
        cur_c = (cur_c * cur_m + cur_c) & Mu
        cur_m = (cur_m * cur_m) & Mu
        bit += 1
    st = ((mul * np.uint64(seed & M) + inc) & Mu).astype(np.uint32)
    r = ((st >> ((st >> np.uint32(28)) + np.uint32(4))) ^ st) * np.uint32(277803737)
    r = (r >> np.uint32(22)) ^ r
    return (r >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def _stress_spheres(n=100000, seed=12345, half=900.0, z0=0.0, z1=150.0, rmin=2.0, rmax=8.0):
    u = _pcg_stream(seed, n * 10).reshape(n, 10)
    sp = np.zeros(n, SPHERE_DT)
    f = np.float32
    sp["position"][:, 0] = (u[:, 0] * f(2.0) - f(1.0)) * f(half)
    sp["position"][:, 1] = (u[:, 1] * f(2.0) - f(1.0)) * f(half)
    sp["position"][:, 2] = f(z0) + u[:, 2] * f(z1 - z0)
    sp["radius"] = f(rmin) + u[:, 3] * f(rmax - rmin)
    mat = u[:, 4]
    diffuse, metal, glass = mat < f(0.8), (mat >= f(0.8)) & (mat < f(0.95)), mat >= f(0.95)
    sp["color"][:, 3] = 1.0
    sp["color"][diffuse, :3] = u[diffuse, 5:8]
    sp["color"][metal, :3] = f(0.5) + f(0.5) * u[metal, 5:8]
    sp["color"][glass, :3] = 1.0
    sp["smoothness"][metal] = f(0.5) * u[metal, 8]
    sp["smoothness"][glass] = -1.0
    return sp


def build(name, width=None, height=None):
    """-> ((camera uniform, spheres, vertices, indices, meshes), textures, meta) with oracle dtypes."""
    W, H, spp, bounces, sky = CONFIGS[name]
    W, H = width or W, height or H
    aspect = np.float32(W) / np.float32(H)
    no_sph = np.zeros(0, SPHERE_DT)
    tex = []
    if name == "c1_balls":
        scene = hr.scene_balls(aspect)
    elif name in ("c2_cube", "c2_cube_plain"):
        textured = name == "c2_cube"
        vs, ix, ms = _fixture("cube_obj", uv=textured)
        light = hr.sphere((-7.44, -0.72, 20.0), 15.0, (0.1, 0.1, 0.1, 0.0), (1, 1, 1, 1), 2.0, 0.0)
        cam = hr.camera_to_uniform((3.0, 2.0, 4.0), (0.0, 0.0, 0.0), (0, 1, 0), 45.0, aspect, 0.1, 100.0, 0.0, 1.0)
        if textured:
            z = np.load(os.path.join(FIXTURES, "cobble_256.npz"))
            tex = [z["diffuse"], z["normal"]]
            ms["diffuse_tex"], ms["normal_tex"] = 1, 2
            ms["color"] = (1.0, 1.0, 1.0, 1.0)
            ms["specular"] = 0.0
        scene = (cam, light, vs, ix, ms)
    elif name == "c3_teapot":
        vs, ix, ms = _fixture("teapot")
        cam = hr.camera_to_uniform((40.0, -130.0, 75.0), (5.0, 0.0, 38.0), (0, 0, 1), 45.0, aspect, 0.1, 100.0, 0.0, 1.0)
        scene = (cam, no_sph, vs, ix, ms)
    elif name == "c4_tree":
        vs, ix, ms = _fixture("tree")
        cam = hr.camera_to_uniform((-3.0, 2.0, 3.0), (8.0, 1.0, 0.0), (0, 1, 0), 45.0, aspect, 0.1, 100.0, 0.0, 1.0)
        scene = (cam, no_sph, vs, ix, ms)
    elif name == "c5_stress":
        vs, ix, base = _fixture("teapot")
        ms = []
        for j in range(8):
            for i in range(8):
                for k in range(len(base)):
                    m = base[k:k + 1].copy()
                    m["pos"] = (np.float32((i - 3.5) * 225.0) + base["pos"][k][0], np.float32((j - 3.5) * 225.0), 0.0)
                    ms.append(m)
        cam = hr.camera_to_uniform((-1400.0, -1400.0, 900.0), (0.0, 0.0, 40.0), (0, 0, 1), 45.0, aspect, 0.1, 100.0, 0.0, 1.0)
        scene = (cam, _stress_spheres(), vs, ix, np.concatenate(ms))
    else:
        raise KeyError(name)
    params = dict(width=W, height=H, number_of_bounces=bounces, rays_per_pixel=1, skybox=sky, frames=0, accumulate=1)
    return scene, tex, dict(params=params, spp=spp, dispatches=spp // 2)
