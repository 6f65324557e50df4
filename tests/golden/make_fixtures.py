"""Generates the committed geometry fixtures from the reference's assets (run in the
authoring container only: /root/reference does not exist on the GPU box).

    python tests/golden/make_fixtures.py

Each model is parsed with the oracle's restatement of the reference loaders
(oracle/host_ref.py <- /root/reference/src/core/resource.rs) and stored as the exact
Vertex / index / Mesh-range arrays `load_model` would append, so bench.py and the GPU tests
can build configs C2-C5 without the asset files.

Sources and licences (credits required by the licences):
  the_utah_teapot.glb      "The Utah Teapot" by 3D graphics 101 (sketchfab.com/3dgraphics),
                           CC-BY-NC-4.0, https://sketchfab.com/3d-models/the-utah-teapot-1092c2832df14099807f66c8b792374d
  simple_japanese_tree.glb "Simple Japanese Tree" by Leonardo Camilo Ribeiro (sketchfab.com/leoCamilo),
                           CC-BY-4.0, https://sketchfab.com/3d-models/simple-japanese-tree-07073d2c174343ad86b7c53bc30055e0
  cube.obj                 Blender-exported rounded cube shipped with the reference (no licence stated)
  cobble-diffuse.png / cobble-normal.png  1024^2 RGBA textures shipped with the reference (no licence
                           stated; referenced by nothing upstream), box-filtered to 256^2 -> cobble_256.npz
Positions, normals and indices are kept; for cube.obj also the texture coordinates (texture
extension, config C2).  Used for non-commercial benchmarking and testing.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import host_ref  # noqa: E402

ASSETS = "/root/reference/assets"


def dump(name, vs, ix, ms, uv=False):
    v, i, m = host_ref.concat_model(vs, ix, ms)
    out = os.path.join(HERE, name + ".npz")
    extra = {"uv": np.stack([v[v.dtype.names[1]], v[v.dtype.names[3]]], axis=1)} if uv else {}
    np.savez_compressed(out, pos=v["pos"], normal=v["normal"], indices=i,
                        first=m["first"], triangles=m["triangles"], offset=m["offset"], **extra)
    h = hashlib.sha256(v.tobytes() + i.tobytes() + m.tobytes()).hexdigest()[:16]
    print("%s: %d vertices, %d indices, %d meshes, %d triangles, %d bytes, sha256[:16]=%s" %
          (name, len(v), len(i), len(m), int(m["triangles"].sum()), os.path.getsize(out), h))


if __name__ == "__main__":
    for fn, name in (("the_utah_teapot.glb", "teapot"), ("simple_japanese_tree.glb", "tree"), ("cube.glb", "cube_glb")):
        vs, ix, ms = [], [], []
        host_ref.load_model_glb(open(os.path.join(ASSETS, fn), "rb").read(), vs, ix, ms)
        dump(name, vs, ix, ms)
    vs, ix, ms = [], [], []
    host_ref.load_model_obj(open(os.path.join(ASSETS, "cube.obj")).read(), vs, ix, ms, keep_uv=True)
    dump("cube_obj", vs, ix, ms, uv=True)
    from PIL import Image
    tex = {}
    for key, fn in (("diffuse", "cobble-diffuse.png"), ("normal", "cobble-normal.png")):
        im = Image.open(os.path.join(ASSETS, fn)).convert("RGBA").resize((256, 256), Image.BOX)
        tex[key] = np.asarray(im, np.uint8)
    out = os.path.join(HERE, "cobble_256.npz")
    np.savez_compressed(out, **tex)
    print("cobble_256: %d bytes, sha256[:16]=%s" % (os.path.getsize(out), hashlib.sha256(tex["diffuse"].tobytes() + tex["normal"].tobytes()).hexdigest()[:16]))
