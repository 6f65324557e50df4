"""Mutation fuzzing of the host loaders under AddressSanitizer + UBSan (CPU only):
    python scripts/fuzz_loaders.py [seed] [files per family]
Seeds: Pillow-written baseline / progressive JPEGs and PNGs, hand-encoded Adam7 PNGs, a small .glb, the same scene as a
.gltf with a data: URI, an .obj with a .mtl.  Mutations: byte flips, deletions, insertions, numbers replaced by
negative / huge ones.  Prints 'clean' or the sanitizer report of the first failing file."""
import base64, io, json, os, re, struct, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from PIL import Image
from test_flycam_io import _adam7_png

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1500
rng = np.random.default_rng(seed)
work = tempfile.mkdtemp(prefix="fuzz_loaders_")
exe = os.path.join(work, "fuzz_loaders")
subprocess.check_call(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
                       "-I", os.path.join(ROOT, "ray_tracer_b200", "host"), os.path.join(ROOT, "tests", "native", "fuzz_loaders.cpp"),
                       os.path.join(ROOT, "ray_tracer_b200", "host", "rt_resource.cpp"), "-lz", "-o", exe])


def image_seeds():
    img = Image.fromarray(rng.integers(0, 256, (37, 53, 3), dtype=np.uint8), "RGB")
    out = []
    for kw in ({"format": "JPEG", "quality": 80}, {"format": "JPEG", "quality": 60, "progressive": True},
               {"format": "JPEG", "quality": 90, "subsampling": 0, "progressive": True}, {"format": "PNG"}):
        b = io.BytesIO(); img.save(b, **kw); out.append((b.getvalue(), ".bin"))
    b = io.BytesIO(); img.convert("P").save(b, format="PNG"); out.append((b.getvalue(), ".bin"))
    b = io.BytesIO(); img.convert("L").save(b, format="JPEG", progressive=True); out.append((b.getvalue(), ".bin"))
    out.append((_adam7_png(rng.integers(0, 256, (13, 21, 4), dtype=np.uint8), 6, 8), ".bin"))
    out.append((_adam7_png(rng.integers(0, 4, (9, 17), dtype=np.uint8), 3, 2, rng.integers(0, 256, (4, 3), dtype=np.uint8)), ".bin"))
    return out


def model_seeds():
    pos = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], np.float32); nrm = np.array([[0, 0, 1]] * 4, np.float32)
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32); idx = np.array([0, 1, 2, 0, 2, 3], np.uint16)
    blob = pos.tobytes() + nrm.tobytes() + uv.tobytes() + idx.tobytes()
    g = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 2]}],
         "nodes": [{"translation": [1, 2, 3], "rotation": [0, 0, 0.38, 0.92], "scale": [2, 2, 2], "children": [1]}, {"mesh": 0, "translation": [0, 0, -1]},
                   {"mesh": 0, "matrix": [-1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 5, 0, 0, 1]}],
         "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1, "TEXCOORD_0": 2}, "indices": 3, "material": 0}]}],
         "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [0.8, 0.1, 0.2, 1.0], "roughnessFactor": 0.25, "baseColorTexture": {"index": 0}}, "emissiveFactor": [0, 0.5, 0]}],
         "textures": [{"source": 0}], "images": [{"uri": "a.png"}], "buffers": [{"byteLength": len(blob)}],
         "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": 48}, {"buffer": 0, "byteOffset": 48, "byteLength": 48}, {"buffer": 0, "byteOffset": 96, "byteLength": 32}, {"buffer": 0, "byteOffset": 128, "byteLength": 12}],
         "accessors": [{"bufferView": 0, "componentType": 5126, "count": 4, "type": "VEC3"}, {"bufferView": 1, "componentType": 5126, "count": 4, "type": "VEC3"},
                       {"bufferView": 2, "componentType": 5126, "count": 4, "type": "VEC2"}, {"bufferView": 3, "componentType": 5123, "count": 6, "type": "SCALAR"}]}
    js = json.dumps(g).encode(); js += b" " * ((-len(js)) % 4); bl = blob + b"\0" * ((-len(blob)) % 4)
    glb = struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(bl)) + struct.pack("<I4s", len(js), b"JSON") + js + struct.pack("<I4s", len(bl), b"BIN\0") + bl
    g2 = json.loads(json.dumps(g)); g2["buffers"][0]["uri"] = "data:application/octet-stream;base64," + base64.b64encode(blob).decode()
    obj = (b"mtllib m.mtl\no quad\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nvn 0 0 1\nusemtl red\nf 1/1/1 2/2/1 3/3/1\n"
           b"usemtl lamp\nf 1/1/1 3/3/1 4/4/1\nf -4//-1 -3//-1 -2//-1 -1//-1\n")
    open(os.path.join(work, "m.mtl"), "w").write("newmtl red\nNs 225.0\nKd 0.9 0.1 0.2\nmap_Kd a.png\nnewmtl lamp\nKe 2 1 0.5\n")
    return [(glb, ".glb"), (json.dumps(g2).encode(), ".gltf"), (obj, ".obj")]


def mutate(data, binary):
    d = bytearray(data)
    for _ in range(int(rng.integers(1, 6))):
        mode = int(rng.integers(0, 4))
        if mode == 0: d[int(rng.integers(0, len(d)))] = int(rng.integers(0, 256))
        elif mode == 1: p = int(rng.integers(0, len(d))); del d[p:p + int(rng.integers(1, 20))]
        elif mode == 2: p = int(rng.integers(0, len(d))); d[p:p] = bytes(rng.integers(0, 256, int(rng.integers(1, 8)), dtype=np.uint8))
        else:
            ms = list(re.finditer(rb"\d+", bytes(d)))
            if ms:
                m = ms[int(rng.integers(0, len(ms)))]; rep = [b"4294967295", b"-1", b"99999999999", b"0", b"65536", b"1e300"][int(rng.integers(0, 6))]
                if binary: rep = rep[:len(m.group())].ljust(len(m.group()), b" ")
                d = bytearray(bytes(d)[:m.start()] + rep + bytes(d)[m.end():])
        if not d: d = bytearray(b"\0")
    return bytes(d)


for family, seeds in (("images", image_seeds()), ("models", model_seeds())):
    files = []
    for it in range(N):
        data, ext = seeds[it % len(seeds)]
        f = os.path.join(work, "%s_%d%s" % (family, it, ext))
        open(f, "wb").write(mutate(data, ext in (".bin", ".glb")))
        files.append(f)
    for i in range(0, N, 150):
        r = subprocess.run([exe] + files[i:i + 150], capture_output=True, text=True)
        if r.returncode != 0:
            for f in files[i:i + 150]:
                rr = subprocess.run([exe, f], capture_output=True, text=True)
                if rr.returncode != 0:
                    print("FAIL", f); print(rr.stderr[-3000:]); sys.exit(1)
    print(family, "clean:", N, "mutated files")
