"""Freezes small oracle renders as golden vectors:  python tests/golden/make_oracle_goldens.py

The reference's WGSL cannot be executed by any tool in this image, so its shader text is translated
mechanically to C++ at build time (oracle/make_ref.py -> oracle/_ref, contract build) and RUN here:
every case below except the texture extension (`c2_cube`, no reference counterpart) is rendered by that
translation and must equal the hand oracle's render bit for bit before it is frozen
(`<case>/pinned_by_wgsl_ref` = 1).  The fixtures pin both against compiler / flag / host-CPU drift and give
the GPU tests a reference output that does not need /root/reference on the box.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

# name -> (scene kind, w, h, frames, bounces, rpp, skybox, first_frame)
CASES = {
    "balls": ("builtin", 32, 32, 2, 3, 1, 1, 0),
    "balls_nosky_startup": ("builtin:balls", 24, 24, 2, 4, 1, 0, 1),
    "room": ("builtin", 32, 24, 2, 4, 1, 0, 0),
    "metal": ("builtin", 32, 24, 2, 6, 2, 1, 0),
    "random_balls": ("builtin", 24, 16, 1, 4, 1, 1, 0),
    "c2_cube": ("config", 32, 18, 1, 8, 1, 1, 0),           # texture extension: cobble diffuse + normal map
    "c2_cube_plain": ("config", 32, 18, 1, 8, 1, 1, 0),     # (was "c2_cube" before the extension: same bits)
    "c3_teapot": ("config", 32, 18, 1, 8, 1, 1, 0),
    "c4_tree": ("config", 32, 18, 1, 8, 1, 1, 0),
}


def scene_for(name, kind, w, h):
    from oracle import host_ref as hr, rt_oracle as ro
    if kind.startswith("builtin"):
        fn = kind.split(":")[1] if ":" in kind else name
        return getattr(hr, "scene_" + fn)(np.float32(w) / np.float32(h))
    from ray_tracer_b200 import configs
    t = configs.build(name, w, h)[0].as_tuple()
    return (t[0].view(ro.CAMERA_DT), t[1].view(ro.SPHERE_DT), t[2].view(ro.VERTEX_DT), t[3], t[4].view(ro.MESH_DT))


def textures_for(name, kind):
    if not kind.startswith("config"):
        return []
    from ray_tracer_b200 import configs
    return configs.build(name, 8, 8)[0].textures


def render_case(name):
    from oracle import rt_oracle as ro
    kind, w, h, frames, bounces, rpp, sky, first = CASES[name]
    ids = np.full((h, w), 0xFFFFFFFF, np.uint32)
    ctr = np.zeros(3, np.uint64)
    img = ro.render(scene_for(name, kind, w, h), w, h, frames, bounces, rpp, sky, first, ids, ctr, textures=textures_for(name, kind))
    return img, ids, int(ctr[0])


def render_case_wgsl_ref(name):
    """The same case through oracle/_ref (the translated reference shader, contract build)."""
    from oracle import wgsl_ref as wr
    kind, w, h, frames, bounces, rpp, sky, first = CASES[name]
    ids = np.full((h, w), 0xFFFFFFFF, np.uint32)
    ctr = np.zeros(3, np.uint64)
    img = wr.render(scene_for(name, kind, w, h), w, h, frames, bounces, rpp, sky, first, ids, ctr, contract=True)
    return img, ids, int(ctr[0])


if __name__ == "__main__":
    out = {}
    for name in CASES:
        img, ids, rays = render_case(name)
        pinned = 0
        if name != "c2_cube":
            img_r, ids_r, rays_r = render_case_wgsl_ref(name)
            assert rays == rays_r and np.array_equal(ids, ids_r) and np.array_equal(img.view(np.uint32), img_r.view(np.uint32)), name
            pinned = 1
        out[name + "/pinned_by_wgsl_ref"] = np.array([pinned], np.uint8)
        out[name + "/image_bits"] = img.view(np.uint32)
        out[name + "/ids"] = ids
        out[name + "/rays"] = np.array([rays], np.uint64)
        print(name, img.shape, "rays", rays, "hit pixels", int((ids != 0xFFFFFFFF).sum()))
    np.savez_compressed(os.path.join(HERE, "oracle_goldens.npz"), **out)
    print("wrote", os.path.getsize(os.path.join(HERE, "oracle_goldens.npz")), "bytes")
