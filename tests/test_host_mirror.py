"""The C++ host mirror (ray_tracer_b200/host -> librt_host.so) against the oracle's numpy
restatement of camera.rs / scene.rs / resource.rs (oracle/host_ref.py), byte for byte.
No GPU: only record construction and parsing are exercised."""
import json
import os
import struct

import numpy as np
import pytest

ASSETS = "/root/reference/assets"


@pytest.fixture(scope="module")
def host(rtlib):
    from ray_tracer_b200 import host as h
    h.build()
    h.hostlib()
    return h


def test_camera_uniform_known_answer(host):
    # Scene::balls camera (scene.rs:380-387); values of SURVEY.md §3.4(17), good to 1-2 ulp
    cu = host.Camera((3.089, 1.53, -3.0), (-2.0, -1.0, 2.0), (0, 1, 0), 45.0, 1.0, 0.1, 100.0, 0.0, 0.1).to_uniform()[0]
    want = {"w": (0.672294497, 0.334231704, -0.660537004), "u": (-0.700841784, 0.0, -0.713316739),
            "v": (-0.238413066, 0.942490816, 0.234243542), "horizontal": (-0.0580596365, 0.0, -0.0590930954),
            "vertical": (-0.0197507851, 0.0780785009, 0.0194053706),
            "lower_left_corner": (3.06067562, 1.45753753, -2.91410255)}
    for k, v in want.items():
        assert np.allclose(cu[k], v, rtol=3e-7, atol=1e-8), k
    assert cu["lens_radius"] == 0.0 and cu["near"] == np.float32(0.1) and cu["far"] == 100.0
    assert tuple(cu["origin"]) == (np.float32(3.089), np.float32(1.53), np.float32(-3.0))


@pytest.mark.parametrize("aspect", [1.0, 1920 / 1080, 100 / 76, 0.5])
def test_camera_matches_oracle_restatement(host, aspect):
    from oracle import host_ref as hr
    rs = np.random.RandomState(7)
    for _ in range(20):
        o, la = rs.uniform(-10, 10, 3), rs.uniform(-10, 10, 3)
        fov, ap, fd = rs.uniform(10, 100), rs.uniform(0, 2), rs.uniform(0.1, 20)
        a = host.Camera(o, la, (0, 1, 0), fov, np.float32(aspect), 0.1, 100.0, ap, fd).to_uniform()
        b = hr.camera_to_uniform(o, la, (0, 1, 0), fov, np.float32(aspect), 0.1, 100.0, ap, fd)
        assert a.tobytes() == b.tobytes()


@pytest.mark.parametrize("name", ["balls", "metal", "room", "random_balls"])
def test_builtin_scenes_match_oracle_restatement(host, name):
    from oracle import host_ref as hr
    for aspect in (np.float32(1.0), np.float32(1920) / np.float32(1080)):
        a = getattr(host.Scene, name)(aspect).as_tuple()
        b = getattr(hr, "scene_" + name)(aspect)
        for x, y, what in zip(a, b, ("camera", "spheres", "vertices", "indices", "meshes")):
            assert x.tobytes() == y.tobytes(), (name, what)


def test_builtin_scene_contents(host):
    s = host.Scene.balls(1.0)
    assert len(s.spheres) == 6 and len(s.meshes) == 1 and s.meshes["triangles"][0] == 0      # dummy mesh, scene.rs:432-445
    assert s.spheres["emission_strength"][5] == 2.0 and s.spheres["radius"][4] == 15.0
    r = host.Scene.room(1.0)
    assert len(r.spheres) == 2 and len(r.meshes) == 7 and len(r.vertices) == 12 and len(r.indices) == 42
    assert r.meshes["emission_strength"][6] == 10.5 and r.spheres["smoothness"][0] == 1.0     # specular 1.0 clamps to 1.0
    m = host.Scene.metal(1.0)
    assert m.spheres["smoothness"][2] == -1.0                                                 # glass passes the clamp
    rb = host.Scene.random_balls(1.0, 12345)
    assert 400 < len(rb.spheres) <= 488 and rb.spheres["radius"][0] == 1000.0


def _write_glb(path, positions, normals, indices, index_type):
    comp = {np.uint16: 5123, np.uint32: 5125, np.uint8: 5121}[index_type]
    pb, nb, ib = positions.astype("<f4").tobytes(), normals.astype("<f4").tobytes(), indices.astype(index_type).tobytes()
    pad = (-len(ib)) % 4
    blob = pb + nb + ib + b"\0" * pad
    j = {"asset": {"version": "2.0"}, "buffers": [{"byteLength": len(blob)}],
         "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": len(pb)},
                         {"buffer": 0, "byteOffset": len(pb), "byteLength": len(nb)},
                         {"buffer": 0, "byteOffset": len(pb) + len(nb), "byteLength": len(ib)}],
         "accessors": [{"bufferView": 0, "componentType": 5126, "count": len(positions), "type": "VEC3"},
                       {"bufferView": 1, "componentType": 5126, "count": len(normals), "type": "VEC3"},
                       {"bufferView": 2, "componentType": comp, "count": len(indices), "type": "SCALAR"}],
         "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1}, "indices": 2}]},
                    {"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1}}]}],
         "nodes": [{"mesh": 0, "translation": [5, 5, 5]}, {"mesh": 1}], "scenes": [{"nodes": [0, 1]}], "scene": 0}
    js = json.dumps(j).encode()
    js += b" " * ((-len(js)) % 4)
    with open(path, "wb") as f:
        f.write(struct.pack("<III", 0x46546C67, 2, 12 + 8 + len(js) + 8 + len(blob)))
        f.write(struct.pack("<II", len(js), 0x4E4F534A) + js)
        f.write(struct.pack("<II", len(blob), 0x004E4942) + blob)


@pytest.mark.parametrize("index_type", [np.uint8, np.uint16, np.uint32])
def test_glb_loader_synthetic(host, tmp_path, index_type):
    from oracle import host_ref as hr
    rs = np.random.RandomState(3)
    pos, nrm = rs.uniform(-2, 2, (9, 3)).astype(np.float32), rs.uniform(-1, 1, (9, 3)).astype(np.float32)
    idx = np.array([0, 1, 2, 2, 3, 4, 8, 7, 6, 5, 4, 3], index_type)
    p = str(tmp_path / "m.glb")
    _write_glb(p, pos, nrm, idx, index_type)
    v, i, m = host.load_model(p)
    vs, ix, ms = [], [], []
    hr.load_model_glb(open(p, "rb").read(), vs, ix, ms)
    V, I, M = hr.concat_model(vs, ix, ms)
    assert v.tobytes() == V.tobytes() and i.tobytes() == I.tobytes() and m.tobytes() == M.tobytes()
    # resource.rs:252-264: one Mesh per primitive, pos = (k*3,0,0), colour (0.2,0.2,1,1), specular 0.5;
    # node transforms ignored; a primitive without indices gets 0..n-1 (resource.rs:244-247)
    assert len(m) == 2 and tuple(m["pos"][1]) == (3.0, 0.0, 0.0) and m["specular"][0] == np.float32(0.5)
    assert tuple(m["triangles"]) == (4, 3) and tuple(m["first"]) == (0, 12) and tuple(m["offset"]) == (0, 9)
    assert np.array_equal(v["pos"][:9], pos) and np.array_equal(i[12:], np.arange(9))


def test_obj_loader_synthetic(host, tmp_path):
    from oracle import host_ref as hr
    text = """# two objects, a quad (fan-triangulated) and a triangle with negative indices
o first
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
vn 0 0 1
vn 0 1 0
vt 0 0
f 1/1/1 2/1/1 3/1/1 4/1/2
f 1//1 2//1 3//1
o second
v 5 5 5
v 6 5 5
v 5 6 5
f -3//1 -2//1 -1//2
"""
    p = str(tmp_path / "m.obj")
    open(p, "w").write(text)
    v, i, m = host.load_model(p)
    vs, ix, ms = [], [], []
    hr.load_model_obj(text, vs, ix, ms)
    V, I, M = hr.concat_model(vs, ix, ms)
    assert v.tobytes() == V.tobytes() and i.tobytes() == I.tobytes() and m.tobytes() == M.tobytes()
    assert len(m) == 2 and tuple(m["triangles"]) == (3, 1) and tuple(m["pos"][1]) == (3.0, 0.0, 0.0)
    # single_index: (v,vt,vn) triples are distinct vertices: 4 + 3 (the vt-less corners) in the first model
    assert tuple(m["offset"]) == (0, 7) and len(v) == 10
    assert list(i[:6]) == [0, 1, 2, 0, 2, 3]


def test_loader_errors(host, tmp_path):
    with pytest.raises(host.HostError, match="Unsupported model format"):      # resource.rs:43
        host.load_model(str(tmp_path / "x.stl"))
    with pytest.raises(host.HostError, match="does not exist"):               # resource.rs:8-12
        host.load_model(str(tmp_path / "missing.obj"))
    p = str(tmp_path / "nonormals.obj")
    open(p, "w").write("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    with pytest.raises(host.HostError):                                        # upstream panics indexing normals
        host.load_model(p)


@pytest.mark.skipif(not os.path.isdir(ASSETS), reason="reference assets only exist in the authoring container")
@pytest.mark.parametrize("asset,fixture", [("the_utah_teapot.glb", "teapot"), ("simple_japanese_tree.glb", "tree"),
                                           ("cube.obj", "cube_obj"), ("cube.glb", "cube_glb")])
def test_fixtures_equal_loader_output_on_reference_assets(host, asset, fixture):
    from ray_tracer_b200 import configs
    v, i, m = host.load_model(os.path.join(ASSETS, asset))
    fv, fi, fm = configs.load_fixture(fixture)
    assert v.tobytes() == fv.tobytes() and i.tobytes() == fi.tobytes() and m.tobytes() == fm.tobytes()


def test_fixture_shapes():
    from ray_tracer_b200 import configs
    v, i, m = configs.load_fixture("teapot")
    assert (len(v), len(i), len(m), int(m["triangles"].sum())) == (8456, 47112, 2, 15704)
    v, i, m = configs.load_fixture("tree")
    assert (len(v), len(i), len(m), int(m["triangles"].sum())) == (8608, 14532, 11, 4844)
    v, i, m = configs.load_fixture("cube_obj")
    assert (len(m), int(m["triangles"].sum())) == (1, 428)
    s, meta = configs.build("c5_stress", 64, 36)
    assert len(s.spheres) == 100000 and len(s.meshes) == 128 and int(s.meshes["triangles"].sum()) == 64 * 15704
    assert meta["dispatches"] == 32


@pytest.mark.parametrize("name", ["c1_balls", "c2_cube", "c2_cube_plain", "c3_teapot", "c4_tree", "c5_stress"])
def test_oracle_side_configs_equal_the_products(rtlib, name):
    """oracle/ref_configs.py (what the CPU arms of bench.py render, built without loading the product libraries) produces
    the records of ray_tracer_b200.configs byte for byte."""
    from oracle import ref_configs
    from ray_tracer_b200 import configs
    w, h = (96, 54) if name != "c1_balls" else (64, 64)
    (cam, sp, vs, ix, ms), tex, meta = ref_configs.build(name, w, h)
    scene, pmeta = configs.build(name, w, h)
    pc, psp, pvs, pix, pms = scene.as_tuple()
    assert meta == pmeta
    assert cam.tobytes() == pc.tobytes() and sp.tobytes() == psp.tobytes() and vs.tobytes() == pvs.tobytes()
    assert ix.tobytes() == pix.tobytes() and ms.tobytes() == pms.tobytes()
    assert len(tex) == len(scene.textures) and all(np.array_equal(a, b) for a, b in zip(tex, scene.textures))
