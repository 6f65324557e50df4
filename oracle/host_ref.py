"""numpy restatement of the reference's HOST side of the path — TEST INFRASTRUCTURE.

Follows /root/reference/src/core/camera.rs:79-121 (Camera::to_uniform), scene.rs:46-77
(constructor clamps), scene.rs:121-476 (built-in scenes) and resource.rs:47-278 (model
loaders' placement rules).  All arithmetic is binary32 in the reference's operation order
(glam 0.23 Vec3: dot = x*x + y*y + z*z, normalize = v * (1/sqrt(dot)), no fused ops).
Used to check the product's C++ host mirror (ray_tracer_b200/host/) and to feed the oracle.
"""
import json
import math
import struct

import numpy as np

from .rt_oracle import CAMERA_DT, MESH_DT, SPHERE_DT, VERTEX_DT

f32 = np.float32

# Rust's f32::tan lowers to the platform libm's tanf (glibc here), which is faithfully but
# not always correctly rounded — call the same routine so the uniform is bit-identical.
import ctypes as _C
_libm = _C.CDLL("libm.so.6")
_libm.tanf.argtypes, _libm.tanf.restype = [_C.c_float], _C.c_float


def _tanf(x):
    return f32(_libm.tanf(_C.c_float(float(x))))


def _v(a):
    return np.asarray(a, dtype=np.float32)


def _dot(a, b):
    return f32(f32(f32(a[0] * b[0]) + f32(a[1] * b[1])) + f32(a[2] * b[2]))


def _normalize(a):
    return _v(a * f32(f32(1.0) / np.sqrt(_dot(a, a), dtype=np.float32)))


def _cross(a, b):
    return _v([f32(a[1] * b[2]) - f32(b[1] * a[2]),
               f32(a[2] * b[0]) - f32(b[2] * a[0]),
               f32(a[0] * b[1]) - f32(b[0] * a[1])])


def camera_to_uniform(origin, look_at, view_up, fov, aspect, near, far, aperture, focus_dist):
    """camera.rs:79-121. `fov` in degrees; radians() is camera.rs:248-250."""
    origin, look_at, view_up = _v(origin), _v(look_at), _v(view_up)
    fov, aspect, focus_dist, aperture = f32(fov), f32(aspect), f32(focus_dist), f32(aperture)
    theta = f32(fov * f32(f32(math.pi) / f32(180.0)))
    height = f32(f32(2.0) * _tanf(f32(theta / f32(2.0))))
    width = f32(aspect * height)
    w = _normalize(origin - look_at)
    u = _normalize(_cross(view_up, w))
    v = _cross(w, u)
    horizontal = _v(f32(focus_dist * width) * u)
    vertical = _v(f32(focus_dist * height) * v)
    llc = _v(_v(_v(origin - horizontal / f32(2.0)) - vertical / f32(2.0)) - focus_dist * w)
    cu = np.zeros(1, CAMERA_DT)
    cu["origin"], cu["lower_left_corner"] = origin, llc
    cu["horizontal"], cu["vertical"] = horizontal, vertical
    cu["near"], cu["far"] = near, far
    cu["w"], cu["u"], cu["v"] = w, u, v
    cu["lens_radius"] = f32(aperture / f32(2.0))
    return cu


def sphere(position, radius, color, emission_color, emission_strength, specular):
    """Sphere::new (scene.rs:46-58): smoothness = specular if specular < 1 else 1."""
    s = np.zeros(1, SPHERE_DT)
    s["position"], s["radius"], s["color"], s["emission_color"] = position, radius, color, emission_color
    s["emission_strength"] = emission_strength
    s["smoothness"] = specular if f32(specular) < f32(1.0) else 1.0
    return s


def mesh(pos, first, triangles, offset, color, emission_color, emission_strength, specular):
    """Mesh::new (scene.rs:61-77)."""
    m = np.zeros(1, MESH_DT)
    m["first"], m["triangles"], m["offset"] = first, triangles, offset
    m["pos"], m["color"], m["emission_color"] = pos, color, emission_color
    m["emission_strength"] = emission_strength
    m["specular"] = specular if f32(specular) < f32(1.0) else 1.0
    return m


def vertex(pos, normal):
    v = np.zeros(1, VERTEX_DT)
    v["pos"], v["normal"] = pos, normal
    return v


def _dummy_mesh():
    """The 1-vertex / 1-index / 0-triangle mesh every sphere-only scene carries
    (scene.rs:176-189, 356-369, 432-445)."""
    return (vertex((0, 0, 1), (2, -3, -1)), np.array([1], np.uint32),
            mesh((0, 0, 0), 0, 0, 0, (0, 0.6, 0, 1), (1, 1, 1, 1), 0.0, 0.5))


def scene_balls(aspect=1.0):
    """Scene::balls, scene.rs:379-476 (the default scene, context.rs:105)."""
    cam = camera_to_uniform((3.089, 1.53, -3.0), (-2.0, -1.0, 2.0), (0, 1, 0), 45.0, aspect, 0.1, 100.0, 0.0, 0.1)
    sp = np.concatenate([
        sphere((-3.64, -0.42, 0.8028), 0.75, (1, 1, 1, 1), (0, 0, 0, 1), 0.0, 0.7),
        sphere((-2.54, -0.72, 0.5), 0.6, (1, 0, 0, 1), (0, 0, 0, 1), 0.0, 0.5),
        sphere((-1.27, -0.72, 1.0), 0.5, (0, 1, 0, 1), (1, 1, 1, 1), 0.0, 0.2),
        sphere((-0.5, -0.9, 1.55), 0.35, (0, 0, 1, 1), (1, 1, 1, 1), 0.0, 0.0),
        sphere((-3.46, -15.88, 2.76), 15.0, (0.5, 0.0, 0.8, 1), (1, 1, 1, 1), 0.0, 0.0),
        sphere((-7.44, -0.72, 20.0), 15.0, (0.1, 0.1, 0.1, 0.0), (1, 1, 1, 1), 2.0, 0.0),
    ])
    v, i, m = _dummy_mesh()
    return cam, sp, v, i, m


def scene_metal(aspect=1.0):
    """Scene::metal, scene.rs:311-378."""
    cam = camera_to_uniform((0, 0, 3), (0, 0, -1), (0, 1, 0), 45.0, aspect, 0.1, 100.0, 0.0, 0.1)
    sp = np.concatenate([
        sphere((0.0, -100.5, -1.0), 100.0, (0.8, 0.8, 0.0, 1), (0, 0, 0, 1), 0.0, 0.0),
        sphere((0.0, 0.0, -1.0), 0.5, (0.7, 0.3, 0.3, 1), (0, 0, 0, 1), 0.0, 0.0),
        sphere((-1.0, 0.0, -1.0), 0.5, (0.8, 0.8, 0.8, 1), (0, 0, 0, 1), 0.0, -1.0),
        sphere((1.0, 0.0, -1.0), 0.5, (0.8, 0.6, 0.2, 1), (0, 0, 0, 1), 0.0, 0.15),
    ])
    v, i, m = _dummy_mesh()
    return cam, sp, v, i, m


def scene_room(aspect=1.0):
    """Scene::room, scene.rs:198-310."""
    cam = camera_to_uniform((-7, 0, 0), (1, 0, 0), (0, 1, 0), 45.0, aspect, 0.1, 100.0, 0.0, 0.1)
    sp = np.concatenate([
        sphere((4.0, 0.0, 1.7), 1.2, (1, 1, 1, 1), (0, 0, 0, 1), 0.0, 1.0),
        sphere((4.0, 0.0, -1.7), 1.2, (1, 1, 1, 1), (0, 0, 0, 1), 0.0, 0.5),
    ])
    n = (3.0, -4.0, 2.0)
    vs = np.concatenate([
        vertex((3, -3, -3), (2, -3, -3)), vertex((3, -3, 3), (4, -3, 0)),
        vertex((-3, -3, 3), n), vertex((-3, -3, -3), n), vertex((3, 3, -3), n), vertex((3, 3, 3), n),
        vertex((-3, 3, 3), n), vertex((-3, 3, -3), n),
        vertex((1, 1, -1), n), vertex((1, 1, 1), n), vertex((-1, 1, 1), n), vertex((-1, 1, -1), n),
    ])
    idx = np.array([3, 2, 1, 3, 1, 0, 7, 0, 4, 7, 3, 0, 7, 6, 2, 7, 2, 3, 2, 6, 5, 2, 5, 1,
                    1, 5, 4, 1, 4, 0, 5, 6, 7, 5, 7, 4, 9, 10, 11, 9, 11, 8], np.uint32)
    one = (1, 1, 1, 1)
    ms = np.concatenate([
        mesh((3, 0, 0), 0, 2, 0, (1, 0, 0, 1), one, 0.0, 0.5),
        mesh((3, 0, 0), 6, 2, 0, (0, 1, 0, 1), one, 0.0, 0.5),
        mesh((3, 0, 0), 12, 2, 0, (0, 0, 1, 1), one, 0.0, 0.5),
        mesh((3, 0, 0), 18, 2, 0, (0.5, 0.5, 0, 1), one, 0.0, 0.5),
        mesh((3, 0, 0), 24, 2, 0, (0, 0.5, 0.5, 1), one, 0.0, 0.5),
        mesh((3, 0, 0), 30, 2, 0, one, one, 0.0, 0.5),
        mesh((3, 1.9, 0), 36, 2, 0, one, one, 10.5, 0.0),
    ])
    return cam, sp, vs, idx, ms


class FixedRng:
    """Stand-in for rand::thread_rng (scene.rs:138), which is unseeded upstream.  PCG hash
    stream of wgsl:205-210; gen() = top 24 bits * 2^-24 in [0,1)."""

    def __init__(self, seed):
        self.s = seed & 0xFFFFFFFF

    def gen(self):
        self.s = (self.s * 747796405 + 2891336453) & 0xFFFFFFFF
        s = self.s
        r = (((s >> ((s >> 28) + 4)) ^ s) * 277803737) & 0xFFFFFFFF
        r = ((r >> 22) ^ r) & 0xFFFFFFFF
        return f32((r >> 8) * (1.0 / 16777216.0))

    def gen_range(self, lo, hi):
        return f32(f32(lo) + f32(f32(hi) - f32(lo)) * self.gen())


def scene_random_balls(aspect=1.0, seed=12345):
    """Scene::random_balls, scene.rs:121-197, with thread_rng replaced by FixedRng(seed)."""
    cam = camera_to_uniform((10.5, 2.0, 3.0), (0, 0, 0), (0, 1, 0), 45.0, aspect, 0.1, 100.0, 0.1, 10.0)
    rng = FixedRng(seed)
    zero = (0, 0, 0, 0)
    sp = [sphere((0, -1000.0, 0), 1000.0, (0.5, 0.5, 0.5, 1), zero, 0.0, 0.0)]
    for a in range(-11, 11):
        for b in range(-11, 11):
            mat = rng.gen()
            cx = f32(f32(a) + f32(f32(0.9) * rng.gen()))
            cz = f32(f32(b) + f32(f32(0.9) * rng.gen()))
            center = _v([cx, f32(0.2), cz])
            d = center - _v([4.0, 0.2, 0.0])
            if np.sqrt(_dot(d, d), dtype=np.float32) > f32(0.9):
                if mat < f32(0.8):
                    albedo = (rng.gen(), rng.gen(), rng.gen(), 1.0)
                    sp.append(sphere(center, 0.2, albedo, zero, 0.0, 0.0))
                elif mat < f32(0.95):
                    albedo = (rng.gen_range(0.5, 1.0), rng.gen_range(0.5, 1.0), rng.gen_range(0.5, 1.0), 1.0)
                    fuzz = rng.gen_range(0.0, 0.5)
                    sp.append(sphere(center, 0.2, albedo, zero, 0.0, fuzz))
                else:
                    sp.append(sphere(center, 0.2, (1, 1, 1, 1), zero, 0.0, -1.0))
    sp.append(sphere((0, 1, 0), 1.0, (1, 1, 1, 1), zero, 0.0, -1.0))
    sp.append(sphere((-4, 1, 0), 1.0, (0.4, 0.2, 0.1, 1), zero, 0.0, 0.0))
    sp.append(sphere((4, 1, 0), 1.0, (0.7, 0.6, 0.5, 1), zero, 0.0, 0.9))
    v, i, m = _dummy_mesh()
    return cam, np.concatenate(sp), v, i, m


# ---- model loaders (resource.rs) -----------------------------------------------------
def _loader_mesh(n_meshes, n_indices, first, offset):
    """The hard-coded record each primitive gets: resource.rs:73-85, 163-175, 252-264."""
    m = np.zeros(1, MESH_DT)
    m["triangles"], m["first"], m["offset"] = n_indices // 3, first, offset
    m["pos"] = (f32(n_meshes) * f32(3.0), 0.0, 0.0)
    m["color"], m["emission_color"], m["emission_strength"], m["specular"] = (0.2, 0.2, 1.0, 1.0), 0, 0.0, 0.5
    return m


def load_model_obj(text, vertices, indices, meshes, keep_uv=False):
    """resource.rs:47-107 via tobj 3.2.5 with triangulate + single_index: one model per
    o/g group, faces fan-triangulated, one output vertex per distinct (v,vt,vn) triple in
    order of first use.  Appends to the three lists (of numpy record arrays).
    keep_uv (texture extension; the reference drops `vt`): store (u, 1 - v) — image row 0 is
    the top row — in the vertex padding words."""
    pos, nrm, tex, models, cur = [], [], [], [], None

    def flush():
        nonlocal cur
        if cur is not None and cur["idx"]:
            models.append(cur)
        cur = None

    def fix(i, n):
        i = int(i)
        return i - 1 if i > 0 else n + i

    for line in text.splitlines():
        t = line.split()
        if not t or t[0].startswith("#"):
            continue
        if t[0] == "v":
            pos.append([float(x) for x in t[1:4]])
        elif t[0] == "vn":
            nrm.append([float(x) for x in t[1:4]])
        elif t[0] == "vt":
            tex.append([float(x) for x in t[1:3]])
        elif t[0] in ("o", "g"):
            flush()
            cur = {"map": {}, "p": [], "n": [], "t": [], "idx": []}
        elif t[0] == "f":
            if cur is None:
                cur = {"map": {}, "p": [], "n": [], "t": [], "idx": []}
            corner = []
            for c in t[1:]:
                parts = c.split("/")
                vi = fix(parts[0], len(pos))
                ti = fix(parts[1], len(tex)) if len(parts) > 1 and parts[1] else -1
                ni = fix(parts[2], len(nrm)) if len(parts) > 2 and parts[2] else -1
                key = (vi, ti, ni)
                if key not in cur["map"]:
                    cur["map"][key] = len(cur["p"])
                    cur["p"].append(pos[vi])
                    cur["n"].append(nrm[ni])          # reference indexes normals unconditionally
                    cur["t"].append(tex[ti] if 0 <= ti < len(tex) else [0.0, 1.0])
                corner.append(cur["map"][key])
            for k in range(1, len(corner) - 1):
                cur["idx"] += [corner[0], corner[k], corner[k + 1]]
    flush()
    n_meshes = sum(len(m) for m in meshes)
    for mdl in models:
        first = sum(len(i) for i in indices)
        offset = sum(len(v) for v in vertices)
        meshes.append(_loader_mesh(n_meshes, len(mdl["idx"]), first, offset))
        n_meshes += 1
        vs = np.zeros(len(mdl["p"]), VERTEX_DT)
        vs["pos"], vs["normal"] = np.asarray(mdl["p"], np.float32), np.asarray(mdl["n"], np.float32)
        if keep_uv:
            uv = np.asarray(mdl["t"], np.float32)
            vs[vs.dtype.names[1]] = uv[:, 0]
            vs[vs.dtype.names[3]] = np.float32(1.0) - uv[:, 1]
        vertices.append(vs)
        indices.append(np.asarray(mdl["idx"], np.uint32))


_COMP = {5120: ("b", 1), 5121: ("B", 1), 5122: ("<i2", 2), 5123: ("<u2", 2), 5125: ("<u4", 4), 5126: ("<f4", 4)}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


def _accessor(j, buffers, k):
    a = j["accessors"][k]
    bv = j["bufferViews"][a["bufferView"]]
    dt, sz = _COMP[a["componentType"]]
    nc = _NCOMP[a["type"]]
    stride = bv.get("byteStride") or sz * nc
    base = bv.get("byteOffset", 0) + a.get("byteOffset", 0)
    buf = buffers[bv["buffer"]]
    out = np.zeros((a["count"], nc), np.dtype(dt))
    for r in range(a["count"]):
        out[r] = np.frombuffer(buf, np.dtype(dt), nc, base + r * stride)
    return out


def _append_primitive(j, buffers, prim, vertices, indices, meshes, scale_by_mesh_count=False):
    p = _accessor(j, buffers, prim["attributes"]["POSITION"]).astype(np.float32)
    n = _accessor(j, buffers, prim["attributes"]["NORMAL"]).astype(np.float32)
    if "indices" in prim:
        idx = _accessor(j, buffers, prim["indices"]).astype(np.uint32).reshape(-1)
    else:
        idx = np.arange(len(p), dtype=np.uint32)
    n_meshes = sum(len(m) for m in meshes)
    first = sum(len(i) for i in indices)
    offset = sum(len(v) for v in vertices)
    meshes.append(_loader_mesh(n_meshes, len(idx), first, offset))
    if scale_by_mesh_count:                       # resource.rs:180 (.gltf path only)
        p = (p * f32(n_meshes + 1)).astype(np.float32)
    vs = np.zeros(len(p), VERTEX_DT)
    vs["pos"], vs["normal"] = p, n
    vertices.append(vs)
    indices.append(idx)


def load_model_glb(data, vertices, indices, meshes):
    """resource.rs:194-278: every primitive of every entry of gltf.meshes(), in order;
    node transforms are ignored."""
    magic, _ver, _len = struct.unpack_from("<III", data, 0)
    assert magic == 0x46546C67
    off, j, blob = 12, None, None
    while off < len(data):
        clen, ctype = struct.unpack_from("<II", data, off)
        chunk = data[off + 8: off + 8 + clen]
        if ctype == 0x4E4F534A:
            j = json.loads(chunk)
        elif ctype == 0x004E4942 and blob is None:
            blob = bytes(chunk)
        off += 8 + clen
    buffers = [blob for _ in j["buffers"]]
    for m in j["meshes"]:
        for prim in m["primitives"]:
            _append_primitive(j, buffers, prim, vertices, indices, meshes)


def load_model_gltf(text, read_uri, vertices, indices, meshes):
    """resource.rs:109-192: top-level nodes of every scene, node.mesh() required,
    positions scaled by the running mesh count (:180)."""
    j = json.loads(text)
    buffers = [read_uri(b["uri"]) for b in j["buffers"]]
    for sc in j["scenes"]:
        for ni in sc["nodes"]:
            node = j["nodes"][ni]
            for prim in j["meshes"][node["mesh"]]["primitives"]:
                _append_primitive(j, buffers, prim, vertices, indices, meshes, scale_by_mesh_count=True)


def concat_model(vertices, indices, meshes):
    def cat(lst, dt):
        return np.concatenate(lst) if lst else np.zeros(0, dt)
    return cat(vertices, VERTEX_DT), cat(indices, np.uint32), cat(meshes, MESH_DT)


# ---- fly-cam (camera.rs:122-247) -------------------------------------------------------
def controller_new(speed=3.0, sensitivity=0.35):
    """CameraController::new (camera.rs:184-198; Camera::new passes 3.0, 0.35 — :57)."""
    z = np.float32(0.0)
    return dict(amount_left=z, amount_right=z, amount_forward=z, amount_backward=z, amount_up=z, amount_down=z,
                rotate_horizontal=z, rotate_vertical=z, scroll=z, speed=np.float32(speed), sensitivity=np.float32(sensitivity))


def process_keyboard(c, key, pressed):
    """camera.rs:201-232."""
    amount = np.float32(5.0 if pressed else 0.0)
    field = {"W": "amount_forward", "Up": "amount_forward", "S": "amount_backward", "Down": "amount_backward",
             "A": "amount_left", "Left": "amount_left", "D": "amount_right", "Right": "amount_right",
             "Space": "amount_up", "LShift": "amount_down"}.get(key)
    if field is None:
        return False
    c[field] = amount
    return True


def update_camera(origin, look_at, c, dt):
    """camera.rs:122-165 in float32; transcendentals through float64 libm rounded to f32 (the
    Rust f32 methods call the platform libm: last-ulp differences are possible)."""
    f = np.float32
    origin, look_at, dt = np.asarray(origin, f).copy(), np.asarray(look_at, f), f(dt)
    safe_deg = f((f(np.pi / 2) - f(0.0001)) * (f(180.0) / f(np.pi)))
    d = look_at - origin
    d = d * (f(1.0) / np.sqrt(f(d[0] * d[0] + d[1] * d[1] + d[2] * d[2])))
    pitch, yaw = f(np.arcsin(np.float64(d[1]))), f(np.arctan2(np.float64(d[0]), np.float64(d[2])))
    ys, yc = f(np.sin(np.float64(yaw))), f(np.cos(np.float64(yaw)))

    def unit(v):
        v = np.asarray(v, f)
        return v * (f(1.0) / np.sqrt(f(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])))
    forward, right = unit([ys, 0.0, yc]), unit([yc, 0.0, -ys])
    origin = origin + forward * (c["amount_forward"] - c["amount_backward"]) * c["speed"] * dt
    origin = origin + right * (c["amount_right"] - c["amount_left"]) * c["speed"] * dt
    ps, pc = f(np.sin(np.float64(pitch))), f(np.cos(np.float64(pitch)))
    scrollward = unit([pc * yc, ps, pc * ys])                       # x/z swapped upstream (camera.rs:141)
    origin = origin - scrollward * c["scroll"] * c["speed"] * c["sensitivity"] * dt
    c["scroll"] = f(0.0)
    origin[1] = origin[1] + (c["amount_up"] - c["amount_down"]) * c["speed"] * dt
    yaw = f(yaw + c["rotate_horizontal"] * c["sensitivity"] * dt)
    pitch = f(pitch + -c["rotate_vertical"] * c["sensitivity"] * dt)
    c["rotate_horizontal"] = c["rotate_vertical"] = f(0.0)
    pitch = min(max(pitch, -safe_deg), safe_deg)                   # never fires: radians vs a degrees constant
    cp, sp_, sy, cy = (f(np.cos(np.float64(pitch))), f(np.sin(np.float64(pitch))), f(np.sin(np.float64(yaw))), f(np.cos(np.float64(yaw))))
    look_at = origin + np.array([cp * sy, sp_, cp * cy], f)
    return origin.astype(f), look_at.astype(f)
