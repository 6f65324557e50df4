"""The five benchmark configurations of BASELINE.json as concrete, frozen scenes.

The reference can only produce C1 itself (its model loaders are never called,
scene.rs:466 is commented out), so C2-C5 are authored here on top of the reference's own
data model: geometry from the committed fixtures (tests/golden/*.npz, extracted from the
reference's assets by tests/golden/make_fixtures.py), one Mesh record per primitive with the
loader's placement rule pos = (k*3, 0, 0) (resource.rs:73-85), cameras through
Camera::to_uniform.  "N spp" = N/2 progressive dispatches at rays_per_pixel = 1
(SURVEY.md §8(d)): the shader takes rpp+1 = 2 samples per dispatch (wgsl:312).
"""
import os

import numpy as np

from ._lib import MESH_DT, SPHERE_DT, VERTEX_DT
from .host import Camera, Scene, loader_mesh

FIXTURES = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

# name -> (width, height, spp, bounces, skybox)
CONFIGS = {
    "c1_balls": (800, 800, 16, 4, 0),
    "c2_cube": (1920, 1080, 64, 8, 1),          # cobble diffuse + normal map bound (texture extension)
    "c2_cube_plain": (1920, 1080, 64, 8, 1),    # the loader's untextured records, as the reference would render them
    "c3_teapot": (1920, 1080, 256, 8, 1),
    "c4_tree": (3840, 2160, 1024, 8, 1),
    "c5_stress": (3840, 2160, 64, 8, 1),
}


def load_fixture(name, mesh_base=0, first_base=0, offset_base=0, uv=False):
    """(vertices, indices, meshes) exactly as load_model appends them after `mesh_base`
    meshes / `first_base` indices / `offset_base` vertices already in the scene."""
    z = np.load(os.path.join(FIXTURES, name + ".npz"))
    vs = np.zeros(len(z["pos"]), VERTEX_DT)
    vs["pos"], vs["normal"] = z["pos"], z["normal"]
    if uv and "uv" in z:
        vs["tu"], vs["tv"] = z["uv"][:, 0], z["uv"][:, 1]
    ms = []
    for k in range(len(z["first"])):
        ms.append(loader_mesh(mesh_base + k, int(z["triangles"][k]) * 3, first_base + int(z["first"][k]),
                              offset_base + int(z["offset"][k])))
    return vs, z["indices"].astype(np.uint32), np.concatenate(ms)


def _sphere(position, radius, color, emission_color, emission_strength, specular):
    s = np.zeros(1, SPHERE_DT)
    s["position"], s["radius"], s["color"], s["emission_color"] = position, radius, color, emission_color
    s["emission_strength"] = emission_strength
    s["smoothness"] = specular if np.float32(specular) < np.float32(1.0) else 1.0   # Sphere::new clamp, scene.rs:56
    return s


def _pcg_stream(seed, n):
    """n values of the shader's PCG hash stream (wgsl:205-210), top 24 bits -> [0,1) f32."""
    out = np.empty(n, np.float32)
    s = np.uint64(seed & 0xFFFFFFFF)
    # the LCG is sequential; jump it in closed form with numpy uint64 arithmetic
    a, c = 747796405, 2891336453
    states = np.empty(n, np.uint64)
    # s_k = a^k s0 + c (a^k - 1)/(a - 1)  (mod 2^32): build multipliers/increments by doubling
    mul = np.ones(n, np.uint64)
    inc = np.zeros(n, np.uint64)
    cur_m, cur_c = np.uint64(a), np.uint64(c)
    k = np.arange(1, n + 1, dtype=np.uint64)
    M = np.uint64(0xFFFFFFFF)
    bit = 0
    while (1 << bit) <= n:
        sel = ((k >> np.uint64(bit)) & np.uint64(1)).astype(bool)
        inc[sel] = (inc[sel] * cur_m + cur_c) & M
        mul[sel] = (mul[sel] * cur_m) & M
        cur_c = (cur_c * cur_m + cur_c) & M
        cur_m = (cur_m * cur_m) & M
        bit += 1
    states = (mul * s + inc) & M
    st = states.astype(np.uint32)
    r = ((st >> ((st >> np.uint32(28)) + np.uint32(4))) ^ st) * np.uint32(277803737)
    r = (r >> np.uint32(22)) ^ r
    out[:] = (r >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    return out


def stress_spheres(n=100000, seed=12345, half=900.0, z0=0.0, z1=150.0, rmin=2.0, rmax=8.0):
    """C5's random spheres: material mix of Scene::random_balls (80 % diffuse, 15 % metal,
    5 % glass; scene.rs:144-158) from a fixed stream (10 draws per sphere)."""
    u = _pcg_stream(seed, n * 10).reshape(n, 10)
    sp = np.zeros(n, SPHERE_DT)
    f = np.float32
    sp["position"][:, 0] = (u[:, 0] * f(2.0) - f(1.0)) * f(half)
    sp["position"][:, 1] = (u[:, 1] * f(2.0) - f(1.0)) * f(half)
    sp["position"][:, 2] = f(z0) + u[:, 2] * f(z1 - z0)
    sp["radius"] = f(rmin) + u[:, 3] * f(rmax - rmin)
    mat = u[:, 4]
    diffuse, metal = mat < f(0.8), (mat >= f(0.8)) & (mat < f(0.95))
    glass = mat >= f(0.95)
    sp["color"][:, 3] = 1.0
    sp["color"][diffuse, :3] = u[diffuse, 5:8]
    sp["color"][metal, :3] = f(0.5) + f(0.5) * u[metal, 5:8]
    sp["color"][glass, :3] = 1.0
    sp["smoothness"][metal] = f(0.5) * u[metal, 8]
    sp["smoothness"][glass] = -1.0
    return sp


def build(name, width=None, height=None):
    """-> (Scene, params dict). width/height override the config's resolution (tests)."""
    W, H, spp, bounces, sky = CONFIGS[name]
    W, H = width or W, height or H
    aspect = np.float32(W) / np.float32(H)
    no_sph = np.zeros(0, SPHERE_DT)
    if name == "c1_balls":
        scene = Scene.balls(aspect)
    elif name in ("c2_cube", "c2_cube_plain"):
        textured = name == "c2_cube"
        vs, ix, ms = load_fixture("cube_obj", uv=textured)
        light = _sphere((-7.44, -0.72, 20.0), 15.0, (0.1, 0.1, 0.1, 0.0), (1, 1, 1, 1), 2.0, 0.0)   # balls' light, scene.rs:423-428
        cam = Camera((3.0, 2.0, 4.0), (0.0, 0.0, 0.0), (0, 1, 0), 45.0, aspect, 0.1, 100.0, 0.0, 1.0)
        tex = []
        if textured:                                   # cobble-diffuse / cobble-normal, box-filtered to 256^2 (make_fixtures.py)
            z = np.load(os.path.join(FIXTURES, "cobble_256.npz"))
            tex = [z["diffuse"], z["normal"]]
            ms["diffuse_tex"], ms["normal_tex"] = 1, 2
            ms["color"] = (1.0, 1.0, 1.0, 1.0)         # the texture supplies the albedo
            ms["specular"] = 0.0
        scene = Scene(cam, light, vs, ix, ms, tex)
    elif name == "c3_teapot":
        vs, ix, ms = load_fixture("teapot")            # 2 meshes at x = 0 and x = 3, raw (Z-up, +-82 unit) coordinates
        cam = Camera((40.0, -130.0, 75.0), (5.0, 0.0, 38.0), (0, 0, 1), 45.0, aspect, 0.1, 100.0, 0.0, 1.0)
        scene = Scene(cam, no_sph, vs, ix, ms)
    elif name == "c4_tree":
        vs, ix, ms = load_fixture("tree")              # 11 meshes at x = 0,3,...,30; node transforms ignored (resource.rs:228)
        cam = Camera((-3.0, 2.0, 3.0), (8.0, 1.0, 0.0), (0, 1, 0), 45.0, aspect, 0.1, 100.0, 0.0, 1.0)
        scene = Scene(cam, no_sph, vs, ix, ms)
    elif name == "c5_stress":
        vs, ix, base = load_fixture("teapot")
        ms = []
        for j in range(8):
            for i in range(8):
                for k in range(len(base)):             # records aliasing one index/vertex range, differing in pos
                    m = base[k:k + 1].copy()
                    m["pos"] = (np.float32((i - 3.5) * 225.0) + base["pos"][k][0], np.float32((j - 3.5) * 225.0), 0.0)
                    ms.append(m)
        cam = Camera((-1400.0, -1400.0, 900.0), (0.0, 0.0, 40.0), (0, 0, 1), 45.0, aspect, 0.1, 100.0, 0.0, 1.0)
        scene = Scene(cam, stress_spheres(), vs, ix, np.concatenate(ms))
    else:
        raise KeyError(name)
    params = dict(width=W, height=H, number_of_bounces=bounces, rays_per_pixel=1, skybox=sky, frames=0, accumulate=1)
    return scene, dict(params=params, spp=spp, dispatches=spp // 2)
