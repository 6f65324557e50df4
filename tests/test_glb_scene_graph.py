"""Scene-graph import of a .glb (host extension): node transforms (matrix and T*R*S, nested),
materials and mirrored instances, against a numpy evaluation of the same file."""
import json
import struct

import numpy as np
import pytest


def _write_glb(path, gltf, blob):
    js = json.dumps(gltf).encode()
    js += b" " * (-len(js) % 4)
    blob += b"\0" * (-len(blob) % 4)
    with open(path, "wb") as f:
        f.write(b"glTF" + struct.pack("<II", 2, 12 + 8 + len(js) + 8 + len(blob)))
        f.write(struct.pack("<I", len(js)) + b"JSON" + js)
        f.write(struct.pack("<I", len(blob)) + b"BIN\0" + blob)


def _quat(axis, deg):
    a = np.asarray(axis, np.float64) / np.linalg.norm(axis)
    h = np.radians(deg) / 2
    return [*(a * np.sin(h)), np.cos(h)]


def _trs(t, q, s):
    x, y, z, w = q
    r = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    m = np.eye(4)
    m[:3, :3] = r @ np.diag(s)
    m[:3, 3] = t
    return m


@pytest.fixture()
def glb(tmp_path):
    pos = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], np.float32)
    nrm = np.array([[0, 0, 1]] * 4, np.float32)
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
    idx = np.array([0, 1, 2, 0, 2, 3], np.uint16)
    blob = pos.tobytes() + nrm.tobytes() + uv.tobytes() + idx.tobytes()
    q = _quat([0, 1, 0], 90)
    parent = _trs([1, 2, 3], _quat([1, 0, 0], 30), [2, 2, 2])
    mirror = np.diag([-1.0, 1.0, 1.0, 1.0])
    mirror[:3, 3] = [5, 0, 0]
    gltf = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 2]}],
        "nodes": [
            {"translation": [1, 2, 3], "rotation": _quat([1, 0, 0], 30), "scale": [2, 2, 2], "children": [1]},
            {"mesh": 0, "translation": [0, 0, -1], "rotation": q},
            {"mesh": 0, "matrix": [float(v) for v in mirror.T.reshape(-1)]},
        ],
        "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1, "TEXCOORD_0": 2}, "indices": 3, "material": 0}]}],
        "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [0.8, 0.1, 0.2, 1.0], "roughnessFactor": 0.25,
                                                "baseColorTexture": {"index": 0}},
                       "normalTexture": {"index": 1}, "emissiveFactor": [0.0, 0.5, 0.0]}],
        "textures": [{"source": 1}, {"source": 0}], "images": [{"uri": "a.png"}, {"uri": "b.png"}],
        "buffers": [{"byteLength": len(blob)}],
        "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": 48}, {"buffer": 0, "byteOffset": 48, "byteLength": 48},
                        {"buffer": 0, "byteOffset": 96, "byteLength": 32}, {"buffer": 0, "byteOffset": 128, "byteLength": 12}],
        "accessors": [{"bufferView": 0, "componentType": 5126, "count": 4, "type": "VEC3"},
                      {"bufferView": 1, "componentType": 5126, "count": 4, "type": "VEC3"},
                      {"bufferView": 2, "componentType": 5126, "count": 4, "type": "VEC2"},
                      {"bufferView": 3, "componentType": 5123, "count": 6, "type": "SCALAR"}],
    }
    path = str(tmp_path / "graph.glb")
    _write_glb(path, gltf, blob)
    worlds = [parent @ _trs([0, 0, -1], q, [1, 1, 1]), mirror]
    return path, pos, nrm, uv, idx, worlds


def test_node_transforms_materials_and_mirrored_instances(glb):
    from ray_tracer_b200 import host
    path, pos, nrm, uv, idx, worlds = glb
    vs, ix, ms = host.load_scene_glb(path)
    assert len(ms) == 2 and len(vs) == 8 and len(ix) == 12
    for k, world in enumerate(worlds):
        m = ms[k]
        assert m["triangles"] == 2 and m["first"] == 6 * k and m["offset"] == 4 * k
        assert np.all(m["pos"][:3] == 0)
        np.testing.assert_allclose(m["color"], [0.8, 0.1, 0.2, 1.0], rtol=1e-6)
        assert m["specular"] == np.float32(0.75)
        np.testing.assert_allclose(m["emission_color"][:3], [0.0, 0.5, 0.0]); assert m["emission_strength"] == 1.0
        assert m["diffuse_tex"] == 2 and m["normal_tex"] == 1          # image index + 1
        v = vs[4 * k: 4 * k + 4]
        want = (world[:3, :3] @ pos.T.astype(np.float64)).T + world[:3, 3]
        np.testing.assert_allclose(v["pos"][:, :3] if v["pos"].ndim == 2 else np.stack([v["pos"]]), want, rtol=1e-5, atol=1e-5)
        n_want = np.linalg.inv(world[:3, :3]).T @ nrm[0].astype(np.float64)
        n_want /= np.linalg.norm(n_want)
        np.testing.assert_allclose(v["normal"][:, :3], np.tile(n_want, (4, 1)), rtol=1e-5, atol=1e-5)
        np.testing.assert_array_equal(np.stack([v["tu"], v["tv"]], 1), uv)
        tri = ix[6 * k: 6 * k + 6].reshape(2, 3)
        if np.linalg.det(world[:3, :3]) < 0:                            # mirrored: winding flipped so the face normal still agrees
            np.testing.assert_array_equal(tri, idx.reshape(2, 3)[:, [0, 2, 1]])
        else:
            np.testing.assert_array_equal(tri, idx.reshape(2, 3))
        p = v["pos"][:, :3].astype(np.float64)
        for t in tri:
            face = np.cross(p[t[1]] - p[t[0]], p[t[2]] - p[t[0]])
            assert np.dot(face, n_want) > 0                              # geometric and shading normal on the same side


def test_glb_without_scene_graph_falls_back_to_plain_loader(tmp_path, glb):
    from ray_tracer_b200 import host
    path = glb[0]
    import json as js
    data = open(path, "rb").read()
    jl = struct.unpack("<I", data[12:16])[0]
    g = js.loads(data[20:20 + jl])
    del g["scenes"], g["scene"]
    p2 = str(tmp_path / "flat.glb")
    _write_glb(p2, g, data[20 + jl + 8:])
    vs, ix, ms = host.load_scene_glb(p2)
    v0, i0, m0 = host.load_model_uv(p2)
    assert np.array_equal(vs.view(np.uint8), v0.view(np.uint8)) and np.array_equal(ix, i0) and np.array_equal(ms.view(np.uint8), m0.view(np.uint8))


def test_gltf_with_external_buffer_imports_like_the_glb(tmp_path, glb):
    from ray_tracer_b200 import host
    path = glb[0]
    data = open(path, "rb").read()
    jl = struct.unpack("<I", data[12:16])[0]
    g = json.loads(data[20:20 + jl])
    g["buffers"][0]["uri"] = "graph.bin"
    (tmp_path / "graph.bin").write_bytes(data[20 + jl + 8:])
    (tmp_path / "graph.gltf").write_text(json.dumps(g))
    a, b = host.load_scene_glb(str(tmp_path / "graph.gltf")), host.load_scene_glb(path)
    for x, y in zip(a, b):
        assert np.array_equal(x.view(np.uint8), y.view(np.uint8))


def test_gltf_with_data_uris_for_buffer_and_images(tmp_path, glb):
    """A .gltf whose buffer and images are base64 `data:` URIs (or sibling image files) imports like the .glb."""
    import base64
    import io
    from PIL import Image
    from ray_tracer_b200 import host
    path = glb[0]
    data = open(path, "rb").read()
    jl = struct.unpack("<I", data[12:16])[0]
    gltf, blob = json.loads(data[20:20 + jl]), data[20 + jl + 8:]
    rng = np.random.default_rng(5)
    tex = rng.integers(0, 256, (8, 8, 4), dtype=np.uint8)
    tex[..., 3] = 255
    png = io.BytesIO(); Image.fromarray(tex, "RGBA").save(png, "PNG")
    (tmp_path / "side.png").write_bytes(png.getvalue())
    g = json.loads(json.dumps(gltf))
    g["buffers"] = [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}]
    g["images"] = [{"uri": "data:image/png;base64," + base64.b64encode(png.getvalue()).decode()}, {"uri": "side.png"}, {"uri": "missing.png"}]
    (tmp_path / "inline.gltf").write_text(json.dumps(g))
    a, b = host.load_scene_glb(str(tmp_path / "inline.gltf")), host.load_scene_glb(path)
    for x, y in zip(a, b):
        assert x.tobytes() == y.tobytes()
    ims = host.load_glb_images(str(tmp_path / "inline.gltf"))
    assert len(ims) == 3 and np.array_equal(ims[0], tex) and np.array_equal(ims[1], tex) and ims[2].size == 0
    # upstream's own loader keeps treating a uri as a file name (resource.rs:129-132): a data: URI is a missing file there
    with pytest.raises(host.HostError):
        host.load_model(str(tmp_path / "inline.gltf"))
    g["buffers"][0]["uri"] = "data:application/octet-stream;base64,@@@@"
    (tmp_path / "bad.gltf").write_text(json.dumps(g))
    with pytest.raises(host.HostError):
        host.load_scene_glb(str(tmp_path / "bad.gltf"))


@pytest.mark.gpu
def test_glb_with_embedded_textures_renders_like_the_oracle(tmp_path, rtlib, oracle):
    """The whole extension chain on the GPU: a .glb with nested transforms, a material and an embedded PNG + baseline
    JPEG -> load_scene_glb + load_glb_images -> rt_upload_scene / rt_upload_textures -> image == oracle, bit for bit."""
    PIL = pytest.importorskip("PIL.Image")
    import io
    from ray_tracer_b200 import host
    rng = np.random.default_rng(5)
    yy, xx = np.mgrid[0:32, 0:32]
    diffuse = np.stack([128 + 100 * np.sin(xx / 3.0), 128 + 100 * np.cos(yy / 4.0), 200 + 0 * xx], -1).clip(0, 255).astype(np.uint8)
    normal = np.stack([128 + 40 * np.sin(xx / 2.0), 128 + 40 * np.cos(yy / 2.5), 230 + 0 * xx], -1).clip(0, 255).astype(np.uint8)
    b1, b2 = io.BytesIO(), io.BytesIO()
    PIL.fromarray(diffuse, "RGB").save(b1, format="JPEG", quality=95)
    PIL.fromarray(normal, "RGB").save(b2, format="PNG")
    pos = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]], np.float32)
    nrm = np.array([[0, 0, 1]] * 4, np.float32)
    uv = np.array([[0, 0], [2, 0], [2, 2], [0, 2]], np.float32)
    idx = np.array([0, 1, 2, 0, 2, 3], np.uint16)
    parts = [pos.tobytes(), nrm.tobytes(), uv.tobytes(), idx.tobytes() + b"\0\0\0\0"[:(-idx.nbytes) % 4], b1.getvalue(), b2.getvalue()]
    views, off = [], 0
    for p in parts:
        views.append({"buffer": 0, "byteOffset": off, "byteLength": len(p)})
        off += len(p) + (-len(p)) % 4
    blob = b"".join(p + b"\0" * ((-len(p)) % 4) for p in parts)
    views[3]["byteLength"] = idx.nbytes
    gltf = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0]}],
        "nodes": [{"translation": [0, 0, -4], "rotation": _quat([0, 1, 0], 25), "children": [1]},
                  {"mesh": 0, "scale": [1.5, 1.0, 1.0], "rotation": _quat([1, 0, 0], -20)}],
        "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1, "TEXCOORD_0": 2}, "indices": 3, "material": 0}]}],
        "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [1.0, 0.9, 0.8, 1.0], "roughnessFactor": 0.6, "baseColorTexture": {"index": 0}},
                       "normalTexture": {"index": 1}}],
        "textures": [{"source": 0}, {"source": 1}],
        "images": [{"bufferView": 4, "mimeType": "image/jpeg"}, {"bufferView": 5, "mimeType": "image/png"}],
        "buffers": [{"byteLength": len(blob)}], "bufferViews": views,
        "accessors": [{"bufferView": 0, "componentType": 5126, "count": 4, "type": "VEC3"}, {"bufferView": 1, "componentType": 5126, "count": 4, "type": "VEC3"},
                      {"bufferView": 2, "componentType": 5126, "count": 4, "type": "VEC2"}, {"bufferView": 3, "componentType": 5123, "count": 6, "type": "SCALAR"}],
    }
    path = str(tmp_path / "textured.glb")
    _write_glb(path, gltf, blob)
    vs, ix, ms = host.load_scene_glb(path)
    textures = host.load_glb_images(path)
    assert [t.shape for t in textures] == [(32, 32, 4), (32, 32, 4)] and ms["diffuse_tex"][0] == 1 and ms["normal_tex"][0] == 2
    assert np.array_equal(textures[1][..., :3], normal)                      # PNG is lossless
    w, h = 96, 64
    base = host.Scene.balls(np.float32(w) / np.float32(h))                    # camera + the light sphere of the built-in scene
    cam, sp = base.camera_uniform(), base.spheres
    tr = rtlib.RayTracer(0, w, h)
    tr.upload_scene(sp, vs, ix, ms); tr.upload_textures(textures); tr.set_camera(cam)
    tr.set_params(number_of_bounces=4, rays_per_pixel=1, skybox=1, frames=0)
    tr.set_option(rtlib._lib.RT_OPT_PRIMARY_IDS, 1)
    tr.render_frames(0, 2)
    img, ids = tr.read_accum(), tr.read_primary_ids()
    tr.close()
    assert (ids >= len(sp)).sum() > 100, "the camera does not see the quad"
    sc = (cam.view(oracle.CAMERA_DT), sp.view(oracle.SPHERE_DT), vs.view(oracle.VERTEX_DT), ix, ms.view(oracle.MESH_DT))
    ref = oracle.render(sc, w, h, 2, 4, 1, 1, 0, textures=textures)
    from util import assert_bits_equal
    assert_bits_equal(img, ref, "glb scene graph with embedded textures")


def test_obj_material_import(tmp_path):
    """load_scene_glb on an .obj (extension): `mtllib` / `usemtl` split the file into one Mesh per material run, with
    Kd / Ke / Ns and the map_Kd / map_Bump images as texture handles; upstream's own load_model on the same file keeps
    its fixed colour and placement (resource.rs:73-85)."""
    import io
    from PIL import Image
    from ray_tracer_b200 import host
    rng = np.random.default_rng(9)
    ta = rng.integers(0, 256, (4, 4, 3), dtype=np.uint8)
    tb = rng.integers(0, 256, (2, 2, 3), dtype=np.uint8)
    Image.fromarray(ta, "RGB").save(tmp_path / "a.png")
    Image.fromarray(tb, "RGB").save(tmp_path / "n.png")
    (tmp_path / "m.mtl").write_text(
        "newmtl red\nNs 225.0\nKd 0.9 0.1 0.2\nKe 0 0 0\nmap_Kd a.png\nmap_Bump -bm 1.0 n.png\n"
        "newmtl lamp\nKd 1 1 1\nKe 2.0 1.0 0.5\nNs 0\n")
    (tmp_path / "s.obj").write_text(
        "mtllib m.mtl\no quad\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nvn 0 0 1\n"
        "usemtl red\nf 1/1/1 2/2/1 3/3/1\nusemtl lamp\nf 1/1/1 3/3/1 4/4/1\nusemtl missing\nf 2/2/1 3/3/1 4/4/1\n")
    path = str(tmp_path / "s.obj")
    vs, ix, ms = host.load_scene_glb(path)
    assert len(ms) == 3 and [int(t) for t in ms["triangles"]] == [1, 1, 1] and len(ix) == 9
    assert np.all(ms["pos"] == 0)
    assert np.allclose(ms["color"][0], [0.9, 0.1, 0.2, 1.0]) and np.isclose(ms["specular"][0], 0.5)      # sqrt(225 / 900)
    assert ms["diffuse_tex"][0] == 1 and ms["normal_tex"][0] == 2
    assert np.allclose(ms["emission_color"][1], [2.0, 1.0, 0.5, 1.0]) and ms["emission_strength"][1] == 1.0 and ms["specular"][1] == 0.0
    assert ms["diffuse_tex"][1] == 0 and ms["normal_tex"][1] == 0
    assert np.allclose(ms["color"][2], [0.2, 0.2, 1.0, 1.0]) and ms["specular"][2] == 0.5               # unknown material: the loader's default record
    first = vs[int(ms["offset"][0]):int(ms["offset"][1])]
    assert np.allclose(first["tu"], [0, 1, 1]) and np.allclose(first["tv"], [1, 1, 0])                    # (u, 1 - v)
    images = host.load_glb_images(path)
    assert len(images) == 2 and np.array_equal(images[0][..., :3], ta) and np.array_equal(images[1][..., :3], tb)
    # upstream's loader on the same file: one model (usemtl is not a split for this mirror), fixed colour, no UVs
    v0, i0, m0 = host.load_model(path)
    assert len(m0) == 1 and int(m0["triangles"][0]) == 3 and np.allclose(m0["color"][0], [0.2, 0.2, 1.0, 1.0])
    assert np.all(v0["tu"] == 0) and np.all(v0["tv"] == 0)


def test_malformed_gltf_fields_are_rejected_not_followed(tmp_path, glb):
    """Negative / absurd byteOffset, count, buffer, bufferView and accessor numbers (a negative byteOffset used to wrap the
    bounds test around: found by fuzzing the loaders under AddressSanitizer) and a few hundred random byte mutations of the
    file: every call returns or raises HostError."""
    from ray_tracer_b200 import host
    data = open(glb[0], "rb").read()
    jl = struct.unpack("<I", data[12:16])[0]
    base, blob = json.loads(data[20:20 + jl]), data[20 + jl + 8:]
    (tmp_path / "g.bin").write_bytes(blob)
    cases = [("bufferViews", 2, "byteOffset", -1), ("bufferViews", 0, "byteOffset", 2 ** 63), ("bufferViews", 1, "byteStride", -12),
             ("bufferViews", 3, "buffer", 7), ("bufferViews", 3, "buffer", -1), ("accessors", 0, "count", 10 ** 12), ("accessors", 3, "count", -6),
             ("accessors", 1, "bufferView", -3), ("accessors", 2, "byteOffset", 1e300), ("accessors", 0, "bufferView", 1e30)]
    for arr, k, field, value in cases:
        g = json.loads(json.dumps(base))
        g["buffers"][0]["uri"] = "g.bin"
        g[arr][k][field] = value
        p = tmp_path / "bad.gltf"
        p.write_text(json.dumps(g))
        with pytest.raises(host.HostError):
            host.load_scene_glb(str(p))
    g = json.loads(json.dumps(base)); g["buffers"][0]["uri"] = "g.bin"
    g["meshes"][0]["primitives"][0]["attributes"]["POSITION"] = -2
    (tmp_path / "bad.gltf").write_text(json.dumps(g))
    with pytest.raises(host.HostError):
        host.load_scene_glb(str(tmp_path / "bad.gltf"))
    rng = np.random.default_rng(17)
    for it in range(300):
        d = bytearray(data)
        for _ in range(int(rng.integers(1, 5))):
            q = int(rng.integers(12, len(d)))
            if rng.integers(0, 2):
                d[q] = int(rng.integers(0, 256))
            else:
                del d[q:q + int(rng.integers(1, 9))]
        p = tmp_path / "mut.glb"
        p.write_bytes(bytes(d))
        for fn in (host.load_scene_glb, host.load_model, host.load_glb_images):
            try:
                fn(str(p))
            except host.HostError:
                pass
