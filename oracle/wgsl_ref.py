"""ctypes binding of oracle/_ref/librt_wgsl_ref*.so — TEST INFRASTRUCTURE.

oracle/_ref is the reference's own shader (shaders/ray_tracer.wgsl), translated mechanically to
C++ by oracle/make_ref.py at build time and compiled here; it is the pin of the hand-written
oracle and of the CUDA path.  Two builds: `contract=True` (the arithmetic contract of
DESIGN.md: the results the hand oracle and the GPU reproduce bit for bit) and
`contract=False` (WGSL-spec formulas, nothing fused, glibc log/cos/pow).

The libraries are built in the container that has /root/reference (make_ref.build()); on the
GPU box the prebuilt files travel with the snapshot.  Only tests/, __graft_entry__ and
bench.py's CPU legs may import this module.
"""
import ctypes as C
import os

import numpy as np

from . import make_ref
from .rt_oracle import CAMERA_DT, MESH_DT, PARAMS_DT, SPHERE_DT, VERTEX_DT, _ptr  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}


def available():
    return make_ref.build() is not None


def lib(contract=True):
    if contract not in _libs:
        out = make_ref.build()
        if out is None:
            raise RuntimeError("oracle/_ref is not built and /root/reference is not present to build it from")
        L = C.CDLL(os.path.join(out, "librt_wgsl_ref.so" if contract else "librt_wgsl_ref_plain.so"))
        vp, u32 = C.c_void_p, C.c_uint32
        L.wgslref_dispatch_ex.argtypes = [vp, vp, vp, u32, vp, u32, vp, u32, vp, u32, vp, vp, vp,
                                          C.c_int, C.c_int, C.c_int, vp, vp, vp, vp]
        L.wgslref_dispatch_ex.restype = C.c_int
        L.wgslref_fn_count.restype = C.c_int
        L.wgslref_fn_name.argtypes = [C.c_int]
        L.wgslref_fn_name.restype = C.c_char_p
        L.wgslref_struct_layout.argtypes = [C.c_char_p, C.POINTER(u32), C.POINTER(u32)]
        L.wgslref_struct_layout.restype = C.c_int
        L.wgslref_contract_mode.restype = C.c_int
        L.wgslref_workgroup_size.restype = C.c_char_p
        assert L.wgslref_contract_mode() == (1 if contract else 0)
        _libs[contract] = L
    return _libs[contract]


def fn_names(contract=True):
    L = lib(contract)
    return [L.wgslref_fn_name(k).decode() for k in range(L.wgslref_fn_count())]


def struct_layout(name, contract=True):
    """(size, align) the WGSL layout rules give the shader's struct, or None."""
    s, a = C.c_uint32(), C.c_uint32()
    if lib(contract).wgslref_struct_layout(name.encode(), C.byref(s), C.byref(a)) != 0:
        return None
    return int(s.value), int(a.value)


class seed_mode:
    """with seed_mode(1): ... -> u32(f32) wraps modulo 2^32 out of range ("wide") in both builds."""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        self.old = [lib(c).wgslref_set_f2u_wrap(int(self.mode)) for c in (True, False)]
        return self

    def __exit__(self, *exc):
        for c, o in zip((True, False), self.old):
            lib(c).wgslref_set_f2u_wrap(o)
        return False


class Aov:
    """Per-dispatch extras of the translated shader (hooks, oracle/wgsl_ref_hooks.h)."""

    def __init__(self, width, height, contract=True):
        self.names = fn_names(contract)
        self.fn_calls = np.zeros(len(self.names), np.uint64)
        self.primary_dst = np.zeros((height, width), np.float32)
        self.primary_hit = np.zeros((height, width), np.uint8)
        self.mismatches = np.zeros(1, np.uint64)

    def calls(self, name):
        return int(self.fn_calls[self.names.index(name)])


def dispatch(params, camera, spheres, vertices, indices, meshes, accum, primary_ids=None, counters=None,
             x_step=1, y_step=1, nthreads=0, contract=True, aov=None):
    """One dispatch of the reference's compute entry over `accum` (H, W, 4) f32 in place."""
    params = np.ascontiguousarray(params, dtype=PARAMS_DT).reshape(1)
    camera = np.ascontiguousarray(camera, dtype=CAMERA_DT).reshape(1)
    spheres = np.ascontiguousarray(spheres, dtype=SPHERE_DT)
    vertices = np.ascontiguousarray(vertices, dtype=VERTEX_DT)
    indices = np.ascontiguousarray(indices, dtype=np.uint32)
    meshes = np.ascontiguousarray(meshes, dtype=MESH_DT)
    H, W = int(params["height"][0]), int(params["width"][0])
    assert accum.dtype == np.float32 and accum.flags.c_contiguous and accum.size == H * W * 4
    rc = lib(contract).wgslref_dispatch_ex(
        _ptr(params), _ptr(camera), _ptr(spheres), spheres.size, _ptr(vertices), vertices.size,
        _ptr(indices), indices.size, _ptr(meshes), meshes.size, _ptr(accum),
        _ptr(primary_ids) if primary_ids is not None else None,
        _ptr(counters) if counters is not None else None, int(x_step), int(y_step), int(nthreads),
        _ptr(aov.fn_calls) if aov is not None else None, _ptr(aov.primary_dst) if aov is not None else None,
        _ptr(aov.primary_hit) if aov is not None else None, _ptr(aov.mismatches) if aov is not None else None)
    if rc != 0:
        raise RuntimeError("wgslref_dispatch failed: %d" % rc)
    return accum


def render(scene, width, height, frames, bounces, rpp=1, toggle=0, first_frame=0, primary_ids=None,
           counters=None, x_step=1, y_step=1, nthreads=0, accum=None, contract=True, aov=None):
    """Same call as oracle.rt_oracle.render: `frames` progressive dispatches, the host incrementing
    params.frames like Context::update (context.rs:176-187).  `aov` collects call counts over all
    dispatches and the primary-hit record of the first one."""
    cam, spheres, vertices, indices, meshes = scene
    if accum is None:
        accum = np.zeros((height, width, 4), np.float32)
    p = np.zeros(1, PARAMS_DT)
    p["width"], p["height"] = width, height
    p["number_of_bounces"], p["rays_per_pixel"], p["toggle"], p["accumulate"] = bounces, rpp, toggle, 1
    first_aov = aov
    for f in range(first_frame, first_frame + frames):
        p["frames"] = f
        a = aov
        if aov is not None and f != first_frame:                  # keep the first dispatch's primary record
            a = Aov.__new__(Aov)
            a.names, a.fn_calls, a.mismatches = aov.names, aov.fn_calls, aov.mismatches
            a.primary_dst = np.zeros_like(aov.primary_dst); a.primary_hit = np.zeros_like(aov.primary_hit)
        dispatch(p, cam, spheres, vertices, indices, meshes, accum, primary_ids if f == first_frame else None,
                 counters, x_step, y_step, nthreads, contract, a)
    return accum
