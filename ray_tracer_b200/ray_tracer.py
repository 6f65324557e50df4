"""RayTracer — the compute pipeline object of the reference, over the C ABI.

Mirrors /root/reference/src/core/ray_tracer.rs: `RayTracer::new(device, texture,
params_buffer, scene)` (:12-147) creates the pipeline and binds params, camera, the
RGBA32F storage texture and the four scene buffers; `update_bind_group` (:149-184) rebinds
after a resize.  Here "device" is a CUDA device ordinal, the storage texture is the
context's accumulation image in HBM, and the bind group is the uploaded SoA scene + LBVH.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (CAMERA_DT, MESH_DT, PARAMS_DT, SPHERE_DT, VERTEX_DT, Counters)


class RtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rt_b200 error %d: %s" % (code, msg))
        self.code = code


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


class RayTracer:
    """One CUDA device, one accumulation image, one uploaded scene."""

    def __init__(self, device, width, height, scene=None, params=None):
        self._L = _lib.lib()
        self._h = C.c_void_p()
        rc = self._L.rt_create(C.byref(self._h), int(device), int(width), int(height))
        if rc != 0:
            raise RtError(rc, self._L.rt_last_error(None).decode())
        self.width, self.height = int(width), int(height)
        if scene is not None:
            self.upload_scene(*scene)
        if params is not None:
            self.set_params(params)

    # -- lifetime ---------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._L.rt_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != 0:
            raise RtError(rc, self._L.rt_last_error(self._h).decode())

    # -- bindings 3..6: the scene buffers (ray_tracer.rs:51-94) ------------------------
    def upload_scene(self, spheres, vertices, indices, meshes):
        sp = np.ascontiguousarray(spheres, dtype=SPHERE_DT).reshape(-1)
        vs = np.ascontiguousarray(vertices, dtype=VERTEX_DT).reshape(-1)
        ix = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        ms = np.ascontiguousarray(meshes, dtype=MESH_DT).reshape(-1)
        self._ck(self._L.rt_upload_scene(self._h, _ptr(sp), sp.size, _ptr(vs), vs.size, _ptr(ix), ix.size,
                                         _ptr(ms), ms.size))

    def update_records(self, spheres, meshes):
        """Same topology, moved / re-coloured records: refit instead of rebuild (rt_update_records)."""
        sp = np.ascontiguousarray(spheres, dtype=SPHERE_DT).reshape(-1)
        ms = np.ascontiguousarray(meshes, dtype=MESH_DT).reshape(-1)
        self._ck(self._L.rt_update_records(self._h, _ptr(sp), sp.size, _ptr(ms), ms.size))

    # -- texture extension (no reference counterpart) ------------------------------------
    def upload_textures(self, textures):
        """textures: list of (H, W, 4) uint8 arrays (row 0 first); handles in Mesh.diffuse_tex /
        normal_tex are 1-based indices into this list."""
        from ._lib import Texture
        keep = [np.ascontiguousarray(t, dtype=np.uint8) for t in textures]
        for t in keep:
            if t.ndim != 3 or t.shape[2] != 4:
                raise ValueError("textures must be (H, W, 4) uint8")
        arr = (Texture * max(1, len(keep)))()
        for i, t in enumerate(keep):
            arr[i] = Texture(t.shape[1], t.shape[0], t.ctypes.data)
        import ctypes as C
        self._ck(self._L.rt_upload_textures(self._h, C.cast(arr, C.c_void_p), len(keep)))

    # -- bindings 0,1: uniforms (context.rs:185-186) -----------------------------------
    def set_camera(self, camera_uniform):
        cu = np.ascontiguousarray(camera_uniform, dtype=CAMERA_DT).reshape(1)
        self._ck(self._L.rt_set_camera(self._h, _ptr(cu)))

    def set_params(self, params=None, **kw):
        p = np.zeros(1, PARAMS_DT)
        if params is not None:
            p[0] = np.ascontiguousarray(params, dtype=PARAMS_DT).reshape(1)[0]
        else:
            p["width"], p["height"] = self.width, self.height
            p["accumulate"] = 1
        for k, v in kw.items():
            p[k] = v
        self._ck(self._L.rt_set_params(self._h, _ptr(p)))
        self.params = p

    # -- the compute pass (context.rs:196-208) -----------------------------------------
    def dispatch(self):
        self._ck(self._L.rt_dispatch(self._h))

    def render_frames(self, first_frame, n_frames):
        self._ck(self._L.rt_render_frames(self._h, int(first_frame), int(n_frames)))

    def render_frames_static(self, first_frame, n_frames):
        """rt_render_frames_static: one launch that allocates nothing (scratch in the library's static device array,
        n_frames * W * H * 16 <= RT_STATIC_SCRATCH_BYTES).  No reference counterpart."""
        self._ck(self._L.rt_render_frames_static(self._h, int(first_frame), int(n_frames)))

    def sync(self):
        self._ck(self._L.rt_sync(self._h))

    def update_bind_group(self, width, height):
        """ray_tracer.rs:149-184 after Context::resize: new zero-filled image."""
        self._ck(self._L.rt_resize(self._h, int(width), int(height)))
        self.width, self.height = int(width), int(height)

    def clear(self):
        self._ck(self._L.rt_clear(self._h))

    def set_shard(self, tile_rank=0, tile_world=1, frame_rank=0, frame_world=1):
        self._ck(self._L.rt_set_shard(self._h, tile_rank, tile_world, frame_rank, frame_world))

    # -- multi-process combine over NCCL (rt_comm_init / rt_combine) ------------------------
    def comm_init(self, n_ranks, rank, unique_id):
        """Collective. unique_id: the 128 bytes of comm_unique_id() of one rank, handed round by the host."""
        b = bytes(unique_id)
        self._ck(self._L.rt_comm_init(self._h, int(n_ranks), int(rank), b, len(b)))

    def comm_destroy(self):
        self._ck(self._L.rt_comm_destroy(self._h))

    def combine(self, root=0):
        """Collective, asynchronous: the shards' images -> the combined image at `root` (read_combined)."""
        self._ck(self._L.rt_combine(self._h, int(root)))

    def combine_rgba8(self, root=0, is_root=True):
        """Collective: the combine fused with the display transform; returns (H, W, 4) uint8 at the root, None elsewhere."""
        out = np.empty((self.height, self.width, 4), np.uint8) if is_root else None
        self._ck(self._L.rt_combine_rgba8(self._h, int(root), _ptr(out) if out is not None else None, out.size if out is not None else 0))
        return out

    def read_combined(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        self._ck(self._L.rt_read_combined(self._h, _ptr(out), out.size))
        return out

    @property
    def combined_device_ptr(self):
        return self._L.rt_combined_device_ptr(self._h)

    @property
    def mean_count(self):
        return int(self._L.rt_mean_count(self._h))

    def set_option(self, option, value):
        self._ck(self._L.rt_set_option(self._h, int(option), int(value)))

    def get_option(self, option):
        v = C.c_int64()
        self._ck(self._L.rt_get_option(self._h, int(option), C.byref(v)))
        return int(v.value)

    # -- read-out ----------------------------------------------------------------------
    def read_accum(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        self._ck(self._L.rt_read_accum(self._h, _ptr(out), out.size))
        return out

    def write_accum(self, img):
        img = np.ascontiguousarray(img, dtype=np.float32)
        self._ck(self._L.rt_write_accum(self._h, _ptr(img), img.size))

    def read_rgba8(self):
        out = np.empty((self.height, self.width, 4), np.uint8)
        self._ck(self._L.rt_read_rgba8(self._h, _ptr(out), out.size))
        return out

    def read_primary_ids(self):
        out = np.empty((self.height, self.width), np.uint32)
        self._ck(self._L.rt_read_primary_ids(self._h, _ptr(out), out.size))
        return out

    def counters(self):
        c = Counters()
        self._ck(self._L.rt_get_counters(self._h, C.byref(c)))
        return {f[0]: getattr(c, f[0]) for f in Counters._fields_}

    def reset_counters(self):
        self._ck(self._L.rt_reset_counters(self._h))

    def bvh_stats(self):
        a, b, d = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._ck(self._L.rt_bvh_stats(self._h, C.byref(a), C.byref(b), C.byref(d)))
        return {"n_prims": a.value, "n_nodes": b.value, "max_depth": d.value}

    @property
    def accum_device_ptr(self):
        return self._L.rt_accum_device_ptr(self._h)

    @property
    def cuda_stream(self):
        return self._L.rt_cuda_stream(self._h)

    @property
    def device(self):
        return self._L.rt_device(self._h)


def _peer_args(tracers, weights):
    tracers = list(tracers)
    arr = (C.c_void_p * len(tracers))(*[t._h for t in tracers])
    w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float32)
    if w is not None and w.size != len(tracers):
        raise RtError(_lib.RT_ERR_INVALID_ARGUMENT, "combine_peers: one weight per context")
    return tracers, arr, w


def comm_unique_id():
    """rt_comm_unique_id: 128 bytes for rt_comm_init, to be handed to every rank by the host."""
    buf = C.create_string_buffer(128)
    L = _lib.lib()
    rc = L.rt_comm_unique_id(buf, 128)
    if rc != 0:
        raise RtError(rc, L.rt_last_error(None).decode())
    return buf.raw


def combine_peers(tracers, weights=None):
    """rt_combine_peers: the shards' images -> tracers[0]'s combined image (read_combined) with one kernel over NVLink peer memory."""
    tracers, arr, w = _peer_args(tracers, weights)
    tracers[0]._ck(tracers[0]._L.rt_combine_peers(arr, len(tracers), _ptr(w) if w is not None else None))


def combine_peers_rgba8(tracers, weights=None):
    """rt_combine_peers_rgba8: the combine fused with the display transform -> (H, W, 4) uint8 on the host."""
    tracers, arr, w = _peer_args(tracers, weights)
    t0 = tracers[0]
    out = np.empty((t0.height, t0.width, 4), np.uint8)
    t0._ck(t0._L.rt_combine_peers_rgba8(arr, len(tracers), _ptr(w) if w is not None else None, _ptr(out), out.size))
    return out
