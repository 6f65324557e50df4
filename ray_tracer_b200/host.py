"""Python face of the C++ host mirror (librt_host.so <- ray_tracer_b200/host/*.cpp).

Same names as the reference's host types: Camera (camera.rs), Scene (scene.rs),
load_model (resource.rs), Context (context.rs: params / frames logic / compute pass).
All arithmetic happens in the C++ layer; this file only marshals arrays.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import _lib
from ._lib import CAMERA_DT, MESH_DT, PARAMS_DT, SPHERE_DT, VERTEX_DT

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(_HERE, "librt_host.so")
_hl = None


def build(force=False):
    _lib.build(force)
    src = os.path.join(_HERE, "host")
    newest = max(os.path.getmtime(os.path.join(src, f)) for f in os.listdir(src))
    newest = max(newest, os.path.getmtime(_lib.LIB_PATH))
    if force or not os.path.exists(HOST_LIB_PATH) or os.path.getmtime(HOST_LIB_PATH) < newest:
        subprocess.check_call(["make", "-C", src, "-s"])
    return HOST_LIB_PATH


def hostlib():
    global _hl
    if _hl is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise RuntimeError("librt_host.so is not built (%s); run __graft_entry__.build()" % HOST_LIB_PATH)
        _lib.lib()      # loads librt_b200.so first (also found through the rpath)
        L = C.CDLL(HOST_LIB_PATH)
        vp, u32, f32 = C.c_void_p, C.c_uint32, C.c_float
        L.rth_last_error.restype = C.c_char_p
        L.rth_camera_to_uniform.argtypes = [vp, vp, vp, f32, f32, f32, f32, f32, f32, vp]
        L.rth_scene_builtin.argtypes = [C.c_int, f32, u32]
        L.rth_scene_builtin.restype = vp
        L.rth_scene_from_model.argtypes = [C.c_char_p]
        L.rth_scene_from_model.restype = vp
        L.rth_scene_free.argtypes = [vp]
        L.rth_scene_counts.argtypes = [vp] + [C.POINTER(u32)] * 4
        L.rth_scene_copy.argtypes = [vp] * 6
        L.rth_loader_mesh.argtypes = [u32, u32, u32, u32, vp]
        L.rth_context_new.argtypes = [C.c_int, u32, u32]
        L.rth_context_new.restype = vp
        L.rth_context_free.argtypes = [vp]
        for n in ("update", "render", "clear_accululation", "sync"):
            getattr(L, "rth_context_" + n).argtypes = [vp]
        L.rth_context_run_frames.argtypes = [vp, u32]
        L.rth_context_resize.argtypes = [vp, u32, u32]
        L.rth_context_set_scene.argtypes = [vp, C.c_int]
        L.rth_context_set_shard.argtypes = [vp, u32, u32, u32, u32]
        L.rth_context_combine.argtypes = [C.POINTER(vp), u32]
        L.rth_context_set_scene_arrays.argtypes = [vp, vp, vp, u32, vp, u32, vp, u32, vp, u32]
        L.rth_context_get_params.argtypes = [vp, vp]
        L.rth_context_set_params.argtypes = [vp, vp]
        L.rth_context_read_texture.argtypes = [vp, vp, C.c_size_t]
        L.rth_context_handle.argtypes = [vp]
        L.rth_context_handle.restype = vp
        # fly-cam / input / display (camera.rs:122-247, context.rs:148-175, render.wgsl)
        L.rth_context_update_dt.argtypes = [vp, f32]
        L.rth_context_input_key.argtypes = [vp, C.c_int, C.c_int]
        L.rth_context_input_scroll_lines.argtypes = [vp, f32]
        L.rth_context_input_mouse_button.argtypes = [vp, C.c_int]
        L.rth_context_input_other.argtypes = [vp]
        L.rth_context_mouse_motion.argtypes = [vp, C.c_double, C.c_double]
        L.rth_context_get_camera.argtypes = [vp, vp]
        L.rth_context_save_png.argtypes = [vp, C.c_char_p, C.c_int]
        L.rth_camera_update.argtypes = [vp, f32]
        L.rth_controller_key.argtypes = [vp, C.c_int, C.c_int]
        # texture-extension loaders and PNG i/o
        L.rth_scene_from_model_uv.argtypes = [C.c_char_p]
        L.rth_scene_from_model_uv.restype = vp
        L.rth_scene_from_glb_graph.argtypes = [C.c_char_p]
        L.rth_scene_from_glb_graph.restype = vp
        L.rth_png_load.argtypes = [C.c_char_p, C.POINTER(u32), C.POINTER(u32)]
        L.rth_png_load.restype = vp
        L.rth_image_copy.argtypes = [vp, vp]
        L.rth_image_free.argtypes = [vp]
        L.rth_png_write.argtypes = [C.c_char_p, vp, u32, u32]
        L.rth_glb_images.argtypes = [C.c_char_p, C.POINTER(u32)]
        L.rth_glb_images.restype = vp
        L.rth_glb_image.argtypes = [vp, u32, C.POINTER(u32), C.POINTER(u32), vp]
        L.rth_glb_images_free.argtypes = [vp]
        L.rth_image_rmse.argtypes = [vp, vp, C.c_size_t]
        L.rth_image_rmse.restype = C.c_double
        L.rth_image_flip.argtypes = [vp, vp, u32, u32, C.c_double, vp, C.POINTER(C.c_double)]
        _hl = L
    return _hl


class HostError(RuntimeError):
    pass


def _err():
    return HostError(hostlib().rth_last_error().decode())


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def _f3(v):
    return np.ascontiguousarray(v, dtype=np.float32).reshape(3)


class Camera:
    """camera.rs:31-78 Camera::new(origin, look_at, view_up, fov, aspect, near, far, aperture, focus_dist)."""

    def __init__(self, origin, look_at, view_up, fov, aspect, near, far, aperture, focus_dist):
        self.origin, self.look_at, self.view_up = _f3(origin), _f3(look_at), _f3(view_up)
        self.fov, self.aspect, self.near, self.far = float(fov), float(aspect), float(near), float(far)
        self.aperture, self.focus_dist = float(aperture), float(focus_dist)

    def to_uniform(self):
        """camera.rs:79-121 (computed by the C++ mirror)."""
        out = np.zeros(1, CAMERA_DT)
        if hostlib().rth_camera_to_uniform(_ptr(self.origin), _ptr(self.look_at), _ptr(self.view_up), self.fov,
                                           self.aspect, self.near, self.far, self.aperture, self.focus_dist,
                                           _ptr(out)) != 0:
            raise _err()
        return out

    def packed(self):
        return np.concatenate([self.origin, self.look_at, self.view_up,
                               np.array([self.fov, self.aspect, self.near, self.far, self.aperture, self.focus_dist],
                                        np.float32)]).astype(np.float32)


class Scene:
    """scene.rs:94-100: camera uniform + the four arrays."""

    def __init__(self, camera, spheres, vertices, indices, meshes, textures=None):
        self.camera = camera            # Camera, or a CAMERA_DT uniform for builtin scenes
        self.textures = list(textures or [])   # texture extension: (H, W, 4) uint8 arrays, addressed by the 1-based Mesh handles
        self.spheres = np.ascontiguousarray(spheres, dtype=SPHERE_DT).reshape(-1)
        self.vertices = np.ascontiguousarray(vertices, dtype=VERTEX_DT).reshape(-1)
        self.indices = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        self.meshes = np.ascontiguousarray(meshes, dtype=MESH_DT).reshape(-1)

    def camera_uniform(self):
        return self.camera.to_uniform() if isinstance(self.camera, Camera) else self.camera

    def arrays(self):
        return self.spheres, self.vertices, self.indices, self.meshes

    def as_tuple(self):
        return (self.camera_uniform(),) + self.arrays()

    @staticmethod
    def _from_handle(h):
        L = hostlib()
        if not h:
            raise _err()
        try:
            n = [C.c_uint32() for _ in range(4)]
            L.rth_scene_counts(h, *[C.byref(x) for x in n])
            cam = np.zeros(1, CAMERA_DT)
            sp, vs = np.zeros(n[0].value, SPHERE_DT), np.zeros(n[1].value, VERTEX_DT)
            ix, ms = np.zeros(n[2].value, np.uint32), np.zeros(n[3].value, MESH_DT)
            L.rth_scene_copy(h, _ptr(cam), _ptr(sp), _ptr(vs), _ptr(ix), _ptr(ms))
        finally:
            L.rth_scene_free(h)
        return Scene(cam, sp, vs, ix, ms)

    # scene.rs:121-476; ids as context.rs:261-278
    @staticmethod
    def balls(aspect=1.0):
        return Scene._from_handle(hostlib().rth_scene_builtin(0, aspect, 0))

    @staticmethod
    def random_balls(aspect=1.0, seed=12345):
        return Scene._from_handle(hostlib().rth_scene_builtin(1, aspect, seed))

    @staticmethod
    def room(aspect=1.0):
        return Scene._from_handle(hostlib().rth_scene_builtin(2, aspect, 0))

    @staticmethod
    def metal(aspect=1.0):
        return Scene._from_handle(hostlib().rth_scene_builtin(3, aspect, 0))


def load_model(path):
    """resource.rs:27-45: returns (vertices, indices, meshes) as load_model would append them."""
    s = Scene._from_handle(hostlib().rth_scene_from_model(os.fsencode(path)))
    return s.vertices, s.indices, s.meshes


def load_model_uv(path):
    """Texture extension: load_model that keeps `vt` / TEXCOORD_0 in the Vertex padding words."""
    s = Scene._from_handle(hostlib().rth_scene_from_model_uv(os.fsencode(path)))
    return s.vertices, s.indices, s.meshes


def load_scene_glb(path):
    """Scene-graph import of a .glb (extension, upstream ignores node transforms and materials): vertices in world
    space, one Mesh per primitive instance at pos 0 with the material's colour / smoothness / emission / texture handles."""
    s = Scene._from_handle(hostlib().rth_scene_from_glb_graph(os.fsencode(path)))
    return s.vertices, s.indices, s.meshes


def load_png(path):
    """-> (H, W, 4) uint8, row 0 = top row (C++ decoder over zlib; no interlace)."""
    L = hostlib()
    w, h = C.c_uint32(), C.c_uint32()
    im = L.rth_png_load(os.fsencode(path), C.byref(w), C.byref(h))
    if not im:
        raise _err()
    try:
        out = np.empty((h.value, w.value, 4), np.uint8)
        L.rth_image_copy(im, _ptr(out))
    finally:
        L.rth_image_free(im)
    return out


def write_png(path, rgba):
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    assert rgba.ndim == 3 and rgba.shape[2] == 4
    if hostlib().rth_png_write(os.fsencode(path), _ptr(rgba), rgba.shape[1], rgba.shape[0]) != 0:
        raise _err()


def load_glb_images(path):
    """PNG images embedded in a .glb, in `images[]` order ((0, 0, 4) arrays for non-PNG entries)."""
    L = hostlib()
    n = C.c_uint32()
    lst = L.rth_glb_images(os.fsencode(path), C.byref(n))
    if not lst:
        raise _err()
    out = []
    try:
        for k in range(n.value):
            w, h = C.c_uint32(), C.c_uint32()
            L.rth_glb_image(lst, k, C.byref(w), C.byref(h), None)
            a = np.empty((h.value, w.value, 4), np.uint8)
            L.rth_glb_image(lst, k, C.byref(w), C.byref(h), _ptr(a))
            out.append(a)
    finally:
        L.rth_glb_images_free(lst)
    return out


def tonemap(accum):
    """What render.wgsl:34-39 + the sRGB swapchain display of an RGBA32F accumulation image: clamp to [0, 1] and
    the sRGB OETF, as float RGB (the 8-bit form is rt_read_rgba8).  numpy, for comparing images on the host."""
    c = np.clip(np.asarray(accum, dtype=np.float64)[..., :3], 0.0, 1.0)
    return np.where(c <= 0.0031308, c * 12.92, 1.055 * np.power(c, 1.0 / 2.4) - 0.055).astype(np.float32)


FLIP_PPD = 0.7 * 3840 / 0.7 * np.pi / 180      # the FLIP paper's default viewing condition: 67.02 pixels per degree


def image_rmse(a, b):
    """Root mean square difference over every value of two equally shaped float images."""
    a, b = np.ascontiguousarray(a, dtype=np.float32), np.ascontiguousarray(b, dtype=np.float32)
    if a.shape != b.shape:
        raise HostError("image_rmse: shapes differ %s %s" % (a.shape, b.shape))
    return float(hostlib().rth_image_rmse(_ptr(a), _ptr(b), a.size))


def image_flip(ref, test, ppd=FLIP_PPD, want_map=False):
    """Mean LDR-FLIP of two sRGB-encoded images ((H, W, 3+) float in [0, 1], or uint8); with want_map also the
    per-pixel error map.  C++ restatement of the published metric (host/rt_imgdiff.cpp)."""
    def prep(x):
        x = np.asarray(x)
        x = x.astype(np.float32) / np.float32(255) if x.dtype == np.uint8 else x.astype(np.float32)
        return np.ascontiguousarray(x[..., :3])
    r, t = prep(ref), prep(test)
    if r.shape != t.shape or r.ndim != 3:
        raise HostError("image_flip: shapes differ or not (H, W, 3): %s %s" % (r.shape, t.shape))
    emap = np.empty(r.shape[:2], np.float32) if want_map else None
    mean = C.c_double()
    if hostlib().rth_image_flip(_ptr(r), _ptr(t), r.shape[1], r.shape[0], float(ppd), _ptr(emap), C.byref(mean)) != 0:
        raise _err()
    return (mean.value, emap) if want_map else mean.value


KEYS = {"W": 0, "S": 1, "A": 2, "D": 3, "Up": 4, "Down": 5, "Left": 6, "Right": 7, "Space": 8, "LShift": 9, "Other": 10}


def camera_update(origin, look_at, controller, dt):
    """camera.rs:122-165 on explicit state (C++ mirror): controller = dict of the 11 CameraController
    fields. -> (origin, look_at, controller after the step)."""
    names = ("amount_left", "amount_right", "amount_forward", "amount_backward", "amount_up", "amount_down",
             "rotate_horizontal", "rotate_vertical", "scroll", "speed", "sensitivity")
    st = np.zeros(17, np.float32)
    st[0:3], st[3:6] = origin, look_at
    st[6:17] = [controller.get(n, {"speed": 3.0, "sensitivity": 0.35}.get(n, 0.0)) for n in names]
    if hostlib().rth_camera_update(_ptr(st), float(dt)) != 0:
        raise _err()
    return st[0:3].copy(), st[3:6].copy(), dict(zip(names, (float(x) for x in st[6:17])))


def loader_mesh(mesh_count, n_indices, first, offset):
    out = np.zeros(1, MESH_DT)
    hostlib().rth_loader_mesh(mesh_count, n_indices, first, offset, _ptr(out))
    return out


class Context:
    """context.rs: params + scene + compute pass, headless."""

    def __init__(self, device, width, height):
        self._L = hostlib()
        self._h = self._L.rth_context_new(int(device), int(width), int(height))
        if not self._h:
            raise _err()

    def close(self):
        if getattr(self, "_h", None):
            self._L.rth_context_free(self._h)
            self._h = None

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise _err()

    @property
    def params(self):
        p = np.zeros(1, PARAMS_DT)
        self._L.rth_context_get_params(self._h, _ptr(p))
        return p

    @params.setter
    def params(self, p):
        p = np.ascontiguousarray(p, dtype=PARAMS_DT).reshape(1)
        self._L.rth_context_set_params(self._h, _ptr(p))

    def set(self, **kw):
        p = self.params
        for k, v in kw.items():
            p[k] = v
        self.params = p

    def update(self):
        self._ck(self._L.rth_context_update(self._h))

    def render(self):
        self._ck(self._L.rth_context_render(self._h))

    # -- interactive path: fly-cam step, window events (every one resets the accumulation) --
    def update_dt(self, dt):
        """context.rs:176-187 with the camera step: update_camera(dt), to_uniform, frames += 1."""
        self._ck(self._L.rth_context_update_dt(self._h, float(dt)))

    def input_key(self, key, pressed):
        r = self._L.rth_context_input_key(self._h, KEYS[key], int(bool(pressed)))
        if r < 0:
            raise _err()
        return bool(r)

    def input_scroll_lines(self, lines):
        return bool(self._L.rth_context_input_scroll_lines(self._h, float(lines)))

    def input_mouse_button(self, pressed):
        return bool(self._L.rth_context_input_mouse_button(self._h, int(bool(pressed))))

    def input_other(self):
        return bool(self._L.rth_context_input_other(self._h))

    def mouse_motion(self, dx, dy):
        self._ck(self._L.rth_context_mouse_motion(self._h, float(dx), float(dy)))

    @property
    def camera(self):
        q = np.zeros(6, np.float32)
        self._L.rth_context_get_camera(self._h, _ptr(q))
        return q[:3].copy(), q[3:].copy()

    def save_png(self, path, mirror_x=False):
        """What the blit + sRGB swapchain show (clamp, OETF, 8 bit), top row first."""
        self._ck(self._L.rth_context_save_png(self._h, os.fsencode(path), int(bool(mirror_x))))

    def run_frames(self, n):
        self._ck(self._L.rth_context_run_frames(self._h, int(n)))

    def resize(self, w, h):
        self._ck(self._L.rth_context_resize(self._h, int(w), int(h)))

    def clear_accululation(self):
        self._ck(self._L.rth_context_clear_accululation(self._h))

    def set_scene(self, scene):
        if isinstance(scene, int):
            self._ck(self._L.rth_context_set_scene(self._h, scene))
            return
        if not isinstance(scene.camera, Camera):
            raise HostError("Context.set_scene needs a Scene with a Camera (not a baked uniform)")
        cam = scene.camera.packed()
        self._ck(self._L.rth_context_set_scene_arrays(self._h, _ptr(cam), _ptr(scene.spheres), scene.spheres.size,
                                                      _ptr(scene.vertices), scene.vertices.size, _ptr(scene.indices),
                                                      scene.indices.size, _ptr(scene.meshes), scene.meshes.size))

    def sync(self):
        self._ck(self._L.rth_context_sync(self._h))

    def set_shard(self, tile_rank=0, tile_world=1, frame_rank=0, frame_world=1):
        """Multi-GPU in one process: this context renders only its share of the tiles / frames (rt_set_shard)."""
        self._ck(self._L.rth_context_set_shard(self._h, tile_rank, tile_world, frame_rank, frame_world))

    @staticmethod
    def combine(shards):
        """Fold the shard contexts' images into shards[0]'s combined image with one kernel over peer memory (rt_combine_peers);
        shards[0].read_texture() / save_png() show it until that context renders again."""
        shards = list(shards)
        arr = (C.c_void_p * len(shards))(*[s._h for s in shards])
        shards[0]._ck(shards[0]._L.rth_context_combine(arr, len(shards)))

    def read_texture(self):
        p = self.params
        out = np.empty((int(p["height"][0]), int(p["width"][0]), 4), np.float32)
        self._ck(self._L.rth_context_read_texture(self._h, _ptr(out), out.size))
        return out

    @property
    def rt_handle(self):
        return self._L.rth_context_handle(self._h)
