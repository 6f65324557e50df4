"""ray_tracer_b200 — B200-native path-tracing hot path of addiswebb/ray_tracer.

The product is `librt_b200.so` (hand-written sm_100a CUDA behind the C ABI of
include/rt_b200.h); this package is the host-side mirror of the reference's types over it.
"""
from ._lib import (CAMERA_DT, MESH_DT, MISS, PARAMS_DT, SPHERE_DT, VERTEX_DT, build, lib)  # noqa: F401
from .ray_tracer import RayTracer, RtError, combine_peers, combine_peers_rgba8, comm_unique_id  # noqa: F401
