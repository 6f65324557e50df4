"""ctypes binding of the CPU oracle (oracle/_build/librt_oracle.so) — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; the product package `ray_tracer_b200` never does.

Record dtypes are byte-identical to the reference's #[repr(C)] structs
(/root/reference/src/core/scene.rs:11-37,79-86, camera.rs:9-29, context.rs:11-21).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "librt_oracle.so")

PARAMS_DT = np.dtype([("width", "<u4"), ("height", "<u4"), ("number_of_bounces", "<i4"),
                      ("rays_per_pixel", "<i4"), ("toggle", "<i4"), ("frames", "<i4"),
                      ("accumulate", "<i4")])
CAMERA_DT = np.dtype([("origin", "<f4", 3), ("_p1", "<f4"),
                      ("lower_left_corner", "<f4", 3), ("_p2", "<f4"),
                      ("horizontal", "<f4", 3), ("_p3", "<f4"),
                      ("vertical", "<f4", 3), ("_p4", "<f4"),
                      ("near", "<f4"), ("far", "<f4"), ("_p5", "<f4", 2),
                      ("w", "<f4", 3), ("_p6", "<f4"),
                      ("u", "<f4", 3), ("_p7", "<f4"),
                      ("v", "<f4", 3), ("lens_radius", "<f4")])
SPHERE_DT = np.dtype([("position", "<f4", 3), ("radius", "<f4"), ("color", "<f4", 4),
                      ("emission_color", "<f4", 4), ("emission_strength", "<f4"),
                      ("smoothness", "<f4"), ("_pad", "<f4", 2)])
# the padding words carry the texture extension (0 = reference behaviour): Mesh.diffuse_tex /
# normal_tex are 1-based texture handles, Vertex.tu / tv the texture coordinates
MESH_DT = np.dtype([("first", "<u4"), ("triangles", "<u4"), ("offset", "<u4"), ("diffuse_tex", "<u4"),
                    ("pos", "<f4", 3), ("_p1", "<f4"), ("color", "<f4", 4),
                    ("emission_color", "<f4", 4), ("emission_strength", "<f4"),
                    ("specular", "<f4"), ("normal_tex", "<u4"), ("_p2", "<u4")])
VERTEX_DT = np.dtype([("pos", "<f4", 3), ("tu", "<f4"), ("normal", "<f4", 3), ("tv", "<f4")])
assert (PARAMS_DT.itemsize, CAMERA_DT.itemsize, SPHERE_DT.itemsize, MESH_DT.itemsize,
        VERTEX_DT.itemsize) == (28, 128, 64, 80, 32)

MISS = 0xFFFFFFFF


class Texture(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("rgba8", C.c_void_p)]


def _texture_table(textures):
    """list of (H, W, 4) uint8 arrays -> (ctypes array of rto_texture, keep-alive list)"""
    keep = [np.ascontiguousarray(t, dtype=np.uint8) for t in (textures or [])]
    for t in keep:
        assert t.ndim == 3 and t.shape[2] == 4
    arr = (Texture * max(1, len(keep)))()
    for i, t in enumerate(keep):
        arr[i] = Texture(t.shape[1], t.shape[0], t.ctypes.data)
    return arr, keep


def build(force=False):
    """Compile the oracle with oracle/Makefile (gcc, OpenMP). Idempotent."""
    src_m = max(os.path.getmtime(os.path.join(_HERE, f)) for f in ("rt_oracle.cpp", "rto_math.h", "Makefile"))
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < src_m:
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        vp, u32, i32, f32 = C.c_void_p, C.c_uint32, C.c_int32, C.c_float
        L.rto_dispatch.argtypes = [vp, vp, vp, u32, vp, u32, vp, u32, vp, u32, vp, vp, vp, C.c_int, C.c_int, C.c_int]
        L.rto_dispatch.restype = C.c_int
        L.rto_dispatch_tex.argtypes = [vp, vp, vp, u32, vp, u32, vp, u32, vp, u32, vp, u32, vp, vp, vp, C.c_int, C.c_int, C.c_int]
        L.rto_dispatch_tex.restype = C.c_int
        L.rto_sample_texture.argtypes = [vp, f32, f32, vp]
        L.rto_max_threads.restype = C.c_int
        L.rto_set_stats.argtypes = [vp]
        L.rto_set_seed_mode.argtypes = [C.c_int]
        L.rto_set_seed_mode.restype = C.c_int
        L.rto_next_random_number.argtypes = [C.POINTER(u32)]
        L.rto_next_random_number.restype = u32
        L.rto_rand.argtypes = [C.POINTER(u32)]
        L.rto_rand.restype = f32
        L.rto_pixel_seed.argtypes = [u32, u32, u32, u32, i32]
        L.rto_pixel_seed.restype = u32
        for n in ("rto_log", "rto_cos", "rto_exp2"):
            getattr(L, n).argtypes = [f32]
            getattr(L, n).restype = f32
        L.rto_pow.argtypes = [f32, f32]
        L.rto_pow.restype = f32
        L.rto_reflectance.argtypes = [f32, f32]
        L.rto_reflectance.restype = f32
        L.rto_rand_hemisphere.argtypes = [vp, C.POINTER(u32), vp]
        L.rto_rand_in_unit_disk.argtypes = [C.POINTER(u32), vp]
        L.rto_ray_sphere.argtypes = [vp, vp, vp, f32, vp]
        L.rto_ray_triangle.argtypes = [vp, vp, vp, vp]
        L.rto_environment_light.argtypes = [vp, vp]
        L.rto_refract.argtypes = [vp, vp, f32, vp]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class stats:
    """with stats() as s: ... -> s.rng_draws, s.samples summed over the dispatches inside the block."""

    def __enter__(self):
        self._buf = np.zeros(2, np.uint64)
        lib().rto_set_stats(_ptr(self._buf))
        return self

    def __exit__(self, *exc):
        lib().rto_set_stats(None)
        self.rng_draws, self.samples = int(self._buf[0]), int(self._buf[1])
        return False


class seed_mode:
    """with seed_mode(1): ... -> the seed's out-of-range f32->u32 conversion wraps ("wide") instead of saturating."""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        self.old = lib().rto_set_seed_mode(int(self.mode))
        return self

    def __exit__(self, *exc):
        lib().rto_set_seed_mode(self.old)
        return False


def max_threads():
    return int(lib().rto_max_threads())


def dispatch(params, camera, spheres, vertices, indices, meshes, accum, primary_ids=None,
             counters=None, x_step=1, y_step=1, nthreads=0, textures=None):
    """One dispatch of the reference's compute entry over `accum` (H, W, 4) f32 in place."""
    params = np.ascontiguousarray(params, dtype=PARAMS_DT).reshape(1)
    camera = np.ascontiguousarray(camera, dtype=CAMERA_DT).reshape(1)
    spheres = np.ascontiguousarray(spheres, dtype=SPHERE_DT)
    vertices = np.ascontiguousarray(vertices, dtype=VERTEX_DT)
    indices = np.ascontiguousarray(indices, dtype=np.uint32)
    meshes = np.ascontiguousarray(meshes, dtype=MESH_DT)
    H, W = int(params["height"][0]), int(params["width"][0])
    assert accum.dtype == np.float32 and accum.flags.c_contiguous and accum.size == H * W * 4
    if primary_ids is not None:
        assert primary_ids.dtype == np.uint32 and primary_ids.flags.c_contiguous and primary_ids.size == H * W
    if counters is not None:
        assert counters.dtype == np.uint64 and counters.size >= 3
    tex, keep = _texture_table(textures)
    rc = lib().rto_dispatch_tex(_ptr(params), _ptr(camera), _ptr(spheres), spheres.size,
                                _ptr(vertices), vertices.size, _ptr(indices), indices.size,
                                _ptr(meshes), meshes.size, C.cast(tex, C.c_void_p), len(keep), _ptr(accum),
                                _ptr(primary_ids) if primary_ids is not None else None,
                                _ptr(counters) if counters is not None else None,
                                int(x_step), int(y_step), int(nthreads))
    if rc != 0:
        raise RuntimeError("rto_dispatch failed: %d" % rc)
    return accum


def render(scene, width, height, frames, bounces, rpp=1, toggle=0, first_frame=0, primary_ids=None,
           counters=None, x_step=1, y_step=1, nthreads=0, accum=None, textures=None):
    """`frames` progressive dispatches with params.frames = first_frame .. first_frame+frames-1
    (host increments `frames` like Context::update, context.rs:176-187)."""
    cam, spheres, vertices, indices, meshes = scene
    if accum is None:
        accum = np.zeros((height, width, 4), np.float32)
    p = np.zeros(1, PARAMS_DT)
    p["width"], p["height"] = width, height
    p["number_of_bounces"], p["rays_per_pixel"], p["toggle"], p["accumulate"] = bounces, rpp, toggle, 1
    for f in range(first_frame, first_frame + frames):
        p["frames"] = f
        dispatch(p, cam, spheres, vertices, indices, meshes, accum,
                 primary_ids if f == first_frame else None, counters, x_step, y_step, nthreads, textures)
    return accum


# ---- KAT probes ---------------------------------------------------------------------
def rng_stream(seed, n):
    s = C.c_uint32(seed)
    return [int(lib().rto_next_random_number(C.byref(s))) for _ in range(n)]


def rand_stream(seed, n):
    s = C.c_uint32(seed)
    return [float(lib().rto_rand(C.byref(s))) for _ in range(n)]


def pixel_seed(x, y, w, h, frames):
    return int(lib().rto_pixel_seed(x, y, w, h, frames))


def ray_sphere(o, d, pos, radius):
    out = np.zeros(8, np.float32)
    o, d, pos = _f32(o), _f32(d), _f32(pos)
    lib().rto_ray_sphere(_ptr(o), _ptr(d), _ptr(pos), float(radius), _ptr(out))
    return out


def ray_triangle(o, d, tri):
    out = np.zeros(8, np.float32)
    o, d, tri = _f32(o), _f32(d), _f32(tri).reshape(18)
    lib().rto_ray_triangle(_ptr(o), _ptr(d), _ptr(tri), _ptr(out))
    return out


def sample_texture(tex, u, v):
    """tex: (H, W, 4) uint8 -> bilinear RGBA f32 under the extension's sampling rule."""
    arr, keep = _texture_table([tex])
    out = np.zeros(4, np.float32)
    lib().rto_sample_texture(C.cast(arr, C.c_void_p), float(u), float(v), _ptr(out))
    return out


def environment_light(d):
    out = np.zeros(4, np.float32)
    d = _f32(d)
    lib().rto_environment_light(_ptr(d), _ptr(out))
    return out
