"""ctypes loader of librt_b200.so (the C ABI of include/rt_b200.h).

There is no CPU fallback: if the library is missing or no CUDA device can be bound, every
entry point raises.  `build()` compiles it in-tree with nvcc for sm_100a (cross-compiles
without a GPU).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")   # override: tuning builds only
CSRC = os.path.join(_HERE, "csrc")

# record dtypes, byte-identical to include/rt_b200.h (and to the reference's #[repr(C)] structs)
PARAMS_DT = np.dtype([("width", "<u4"), ("height", "<u4"), ("number_of_bounces", "<i4"),
                      ("rays_per_pixel", "<i4"), ("skybox", "<i4"), ("frames", "<i4"),
                      ("accumulate", "<i4")])
CAMERA_DT = np.dtype([("origin", "<f4", 3), ("_padding1", "<f4"),
                      ("lower_left_corner", "<f4", 3), ("_padding2", "<f4"),
                      ("horizontal", "<f4", 3), ("_padding3", "<f4"),
                      ("vertical", "<f4", 3), ("_padding4", "<f4"),
                      ("near", "<f4"), ("far", "<f4"), ("_padding5", "<f4", 2),
                      ("w", "<f4", 3), ("_padding6", "<f4"),
                      ("u", "<f4", 3), ("_padding7", "<f4"),
                      ("v", "<f4", 3), ("lens_radius", "<f4")])
SPHERE_DT = np.dtype([("position", "<f4", 3), ("radius", "<f4"), ("color", "<f4", 4),
                      ("emission_color", "<f4", 4), ("emission_strength", "<f4"),
                      ("smoothness", "<f4"), ("_padding", "<f4", 2)])
# texture extension: the reference's padding words carry 1-based texture handles / UVs (0 = none)
MESH_DT = np.dtype([("first", "<u4"), ("triangles", "<u4"), ("offset", "<u4"), ("diffuse_tex", "<u4"),
                    ("pos", "<f4", 3), ("_padding2", "<f4"), ("color", "<f4", 4),
                    ("emission_color", "<f4", 4), ("emission_strength", "<f4"),
                    ("specular", "<f4"), ("normal_tex", "<u4"), ("_padding3", "<u4")])
VERTEX_DT = np.dtype([("pos", "<f4", 3), ("tu", "<f4"), ("normal", "<f4", 3), ("tv", "<f4")])
assert (PARAMS_DT.itemsize, CAMERA_DT.itemsize, SPHERE_DT.itemsize, MESH_DT.itemsize,
        VERTEX_DT.itemsize) == (28, 128, 64, 80, 32)


class Texture(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("rgba8", C.c_void_p)]


class Counters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("samples", C.c_uint64), ("node_visits", C.c_uint64),
                ("tri_tests", C.c_uint64), ("sphere_tests", C.c_uint64), ("hit_segments", C.c_uint64),
                ("dispatches", C.c_uint64), ("kernel_ms", C.c_double),
                ("launches", C.c_uint64)]


RT_OPT_TRAVERSAL, RT_OPT_SEED_MODE, RT_OPT_PRIMARY_IDS, RT_OPT_DETAIL_COUNTERS = 1, 2, 3, 4
RT_OPT_FRAMES_PER_LAUNCH, RT_OPT_LEAF_SIZE, RT_OPT_SCHED = 5, 6, 7
RT_OPT_BUILDER, RT_OPT_PLOC_RADIUS, RT_OPT_KERNEL, RT_OPT_WAVE_SLOTS, RT_OPT_STAGING_MIB = 8, 9, 10, 11, 12
RT_OPT_ITEM_MODE, RT_OPT_INSTANCING, RT_OPT_PIPELINE, RT_OPT_NODE_FORMAT, RT_OPT_NODES_PACKED = 13, 14, 15, 16, 17
RT_OPT_TILE_ORDER = 18
RT_OK, RT_ERR_INVALID_ARGUMENT, RT_ERR_CUDA, RT_ERR_NO_SCENE = 0, -1, -2, -3
RT_ERR_OUT_OF_MEMORY, RT_ERR_BAD_INDEX, RT_ERR_UNSUPPORTED = -4, -5, -6
MISS = 0xFFFFFFFF

# every symbol include/rt_b200.h declares: name -> (restype, argtypes)
_vp, _u32, _i32, _i64, _sz = C.c_void_p, C.c_uint32, C.c_int32, C.c_int64, C.c_size_t
SYMBOLS = {
    "rt_last_error": (C.c_char_p, [_vp]),
    "rt_version": (_u32, []),
    "rt_create": (C.c_int, [C.POINTER(_vp), C.c_int, _u32, _u32]),
    "rt_destroy": (None, [_vp]),
    "rt_upload_scene": (C.c_int, [_vp, _vp, _u32, _vp, _u32, _vp, _u32, _vp, _u32]),
    "rt_update_records": (C.c_int, [_vp, _vp, _u32, _vp, _u32]),
    "rt_upload_textures": (C.c_int, [_vp, _vp, _u32]),
    "rt_set_camera": (C.c_int, [_vp, _vp]),
    "rt_set_params": (C.c_int, [_vp, _vp]),
    "rt_dispatch": (C.c_int, [_vp]),
    "rt_render_frames": (C.c_int, [_vp, _i32, _u32]),
    "rt_render_frames_static": (C.c_int, [_vp, _i32, _u32]),
    "rt_sync": (C.c_int, [_vp]),
    "rt_resize": (C.c_int, [_vp, _u32, _u32]),
    "rt_clear": (C.c_int, [_vp]),
    "rt_set_shard": (C.c_int, [_vp, _u32, _u32, _u32, _u32]),
    "rt_combine_peers": (C.c_int, [C.POINTER(_vp), _u32, _vp]),
    "rt_combine_peers_rgba8": (C.c_int, [C.POINTER(_vp), _u32, _vp, _vp, _sz]),
    "rt_comm_unique_id": (C.c_int, [_vp, _sz]),
    "rt_comm_init": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _sz]),
    "rt_comm_destroy": (C.c_int, [_vp]),
    "rt_combine": (C.c_int, [_vp, C.c_int]),
    "rt_combine_rgba8": (C.c_int, [_vp, C.c_int, _vp, _sz]),
    "rt_read_combined": (C.c_int, [_vp, _vp, _sz]),
    "rt_combined_device_ptr": (_vp, [_vp]),
    "rt_mean_count": (_u32, [_vp]),
    "rt_set_option": (C.c_int, [_vp, C.c_int, _i64]),
    "rt_get_option": (C.c_int, [_vp, C.c_int, C.POINTER(_i64)]),
    "rt_read_accum": (C.c_int, [_vp, _vp, _sz]),
    "rt_write_accum": (C.c_int, [_vp, _vp, _sz]),
    "rt_read_rgba8": (C.c_int, [_vp, _vp, _sz]),
    "rt_read_primary_ids": (C.c_int, [_vp, _vp, _sz]),
    "rt_get_counters": (C.c_int, [_vp, C.POINTER(Counters)]),
    "rt_reset_counters": (C.c_int, [_vp]),
    "rt_accum_device_ptr": (_vp, [_vp]),
    "rt_cuda_stream": (_vp, [_vp]),
    "rt_device": (C.c_int, [_vp]),
    "rt_measure_fp32_peak": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "rt_bvh_stats": (C.c_int, [_vp, C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_u32)]),
}


def build(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a over csrc/*.cu -> librt_b200.so (in-tree)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")) or f == "Makefile"]
    srcs.append(os.path.join(_HERE, "..", "include", "rt_b200.h"))
    newest = max(os.path.getmtime(s) for s in srcs)
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < newest:
        cmd = ["make", "-C", CSRC, "-j4"] + ([] if verbose else ["-s"])
        subprocess.check_call(cmd)
    return LIB_PATH


_lib = None


def lib():
    """The loaded library; raises (never falls back) if it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("librt_b200.so is not built (%s); run __graft_entry__.build() or "
                               "`make -C ray_tracer_b200/csrc` — there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib
