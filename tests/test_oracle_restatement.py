"""A second, independent restatement of the shader, written from shaders/ray_tracer.wgsl in scalar
Python (numpy binary32 scalars, FMAs emulated exactly with rational arithmetic), against the C++ oracle.

The oracle is the checker of every GPU parity test; the reference cannot run here, so the oracle itself is only pinned by
known answers.  This test adds a cross-check of its control flow and operation order (seed, sample loop, lens sample,
camera ray, ray-sphere and ray-triangle tests, closest-hit scan with the tie rule, glass / diffuse / specular bounce,
sky, running mean) by a second
reading of the same WGSL under the arithmetic contract of DESIGN.md.  Shared on purpose: the three polynomial routines
of the contract (log, cos, pow), called through the oracle library and bounded against libm in test_oracle_math.py.
Pixels are few (pure Python); scenes are the built-in `balls` (diffuse + emissive spheres), `metal` (specular + glass) and
`room` (2 spheres + 14 triangles with an emissive quad, the brute-force mesh scan of wgsl:158-182)."""
import ctypes as C
from fractions import Fraction

import numpy as np
import pytest

F = np.float32
ONE, ZERO = F(1.0), F(0.0)


def _round_f32(q):
    """Fraction -> nearest binary32 (ties to even), exactly."""
    if q == 0:
        return ZERO
    c = F(float(q))                                     # correctly rounded to binary64, then to binary32: may be one off
    best = None
    for cand in (np.nextafter(c, F(-np.inf)), c, np.nextafter(c, F(np.inf))):
        if not np.isfinite(cand):
            continue
        err = abs(Fraction(float(cand)) - q)
        even = (int(np.float32(cand).view(np.uint32)) & 1) == 0
        if best is None or err < best[0] or (err == best[0] and even and not best[2]):
            best = (err, cand, even)
    return F(best[1])


def fma(a, b, c):
    if not (np.isfinite(a) and np.isfinite(b) and np.isfinite(c)):
        return F(np.float64(a) * np.float64(b) + np.float64(c))
    r = _round_f32(Fraction(float(a)) * Fraction(float(b)) + Fraction(float(c)))
    if r == 0:                                          # sign of an exact zero: as IEEE (sum of opposite zeros is +0)
        p_neg = (np.signbit(a) != np.signbit(b))
        return F(-0.0) if (p_neg and np.signbit(c)) else ZERO
    return r


def dot3(a, b):
    return fma(a[2], b[2], fma(a[1], b[1], a[0] * b[0]))


def normalize(v):
    inv = ONE / np.sqrt(dot3(v, v))
    return [v[0] * inv, v[1] * inv, v[2] * inv]


def fma3(a, s, c):
    return [fma(a[0], s, c[0]), fma(a[1], s, c[1]), fma(a[2], s, c[2])]


def mix1(a, b, t):
    return fma(b, t, a * (ONE - t))


def reflect(i, n):
    t = F(2.0) * dot3(n, i)
    return fma3(n, -t, i)


class Shader:
    def __init__(self, oracle, scene, width, height, bounces, rpp, toggle):
        self.L = oracle.lib()
        self.cam, self.spheres = scene[0][0] if scene[0].ndim else scene[0], scene[1]
        self.vertices, self.indices, self.meshes = scene[2], scene[3], scene[4]
        self.W, self.H, self.bounces, self.rpp, self.toggle = width, height, bounces, rpp, toggle
        self.rays = 0

    # wgsl:205-210, 187-189
    def rand(self):
        self.rng = (self.rng * 747796405 + 2891336453) & 0xFFFFFFFF
        s = self.rng
        r = (((s >> ((s >> 28) + 4)) ^ s) * 277803737) & 0xFFFFFFFF
        r = (r >> 22) ^ r
        return F(r) / F(4294967295.0)                   # the literal is 2^32 in binary32

    def log(self, x): return F(self.L.rto_log(C.c_float(float(x))))
    def cos(self, x): return F(self.L.rto_cos(C.c_float(float(x))))
    def pow(self, x, y): return F(self.L.rto_pow(C.c_float(float(x)), C.c_float(float(y))))

    def rand_normal_dist(self):                         # wgsl:199-203
        theta = F(2.0) * F(3.1415926) * self.rand()
        rho = np.sqrt(F(-2.0) * self.log(self.rand()))
        return rho * self.cos(theta)

    def rand_in_unit_disk(self):                        # wgsl:216-227
        while True:
            r1, r2 = self.rand() * F(2.0), self.rand() * F(2.0)
            p = [r1 - ONE, r2 - ONE, ONE - ONE]
            if np.sqrt(dot3(p, p)) <= ONE:
                return p

    def ray_sphere(self, o, d, s):                      # wgsl:106-123
        pos, radius = [F(x) for x in s["position"]], F(s["radius"])
        oc = [o[k] - pos[k] for k in range(3)]
        a = dot3(d, d)
        b = F(2.0) * dot3(oc, d)
        c = dot3(oc, oc) - radius * radius
        disc = b * b - (F(4.0) * a) * c
        if disc >= ZERO:
            dst = (-b - np.sqrt(disc)) / (F(2.0) * a)
            if dst >= ZERO:
                hp = fma3(d, dst, o)
                return dst, hp, normalize([hp[k] - pos[k] for k in range(3)])
        return None

    def ray_triangle(self, o, d, a, b, c):              # wgsl:125-147; a, b, c = (pos, normal) with the mesh offset applied
        cross = lambda p, q: [fma(p[1], q[2], -(p[2] * q[1])), fma(p[2], q[0], -(p[0] * q[2])), fma(p[0], q[1], -(p[1] * q[0]))]
        e_ab = [b[0][k] - a[0][k] for k in range(3)]
        e_ac = [c[0][k] - a[0][k] for k in range(3)]
        normal = cross(e_ab, e_ac)
        ao = [o[k] - a[0][k] for k in range(3)]
        dao = cross(ao, d)
        det = -dot3(d, normal)
        inv = ONE / det
        dst = dot3(ao, normal) * inv
        u = dot3(e_ac, dao) * inv
        v = -dot3(e_ab, dao) * inv
        w = (ONE - u) - v
        if not (det >= F(1e-6) and dst >= ZERO and u >= ZERO and v >= ZERO and w >= ZERO):
            return None
        n = normalize(fma3(c[1], v, fma3(b[1], u, [x * w for x in a[1]])))
        return dst, fma3(d, dst, o), n

    def closest(self, o, d):                            # wgsl:149-185: spheres, then meshes, strict '<' against FLT_MAX / the best so far
        best, best_dst = None, F(3.4028234663852886e38)
        for s in self.spheres:
            h = self.ray_sphere(o, d, s)
            if h is not None and h[0] < best_dst:
                best, best_dst = (h, s), h[0]
        for m in self.meshes:
            mp = [F(x) for x in m["pos"]]
            for i in range(int(m["triangles"])):
                tri = []
                for j in range(3):
                    vtx = self.vertices[int(m["offset"]) + int(self.indices[int(m["first"]) + 3 * i + j])]
                    tri.append(([F(vtx["pos"][k]) + mp[k] for k in range(3)], [F(x) for x in vtx["normal"]]))
                h = self.ray_triangle(o, d, *tri)
                if h is not None and h[0] < best_dst:
                    best, best_dst = (h, {"smoothness": m["specular"], "color": m["color"], "emission_color": m["emission_color"],
                                          "emission_strength": m["emission_strength"]}), h[0]
        return best

    def environment(self, d):                           # wgsl:297-304
        def smooth(e0, e1, x):
            t = min(max((x - e0) / (e1 - e0), ZERO), ONE)
            return (t * t) * (F(3.0) - F(2.0) * t)
        sky_t = self.pow(smooth(F(0.0), F(0.4), d[1]), F(0.35))
        g2s = smooth(F(-0.01), F(0.0), d[1])
        horizon, zenith, ground = [F(1), F(1), F(1), F(0)], [F(0.0788092), F(0.36480793), F(0.7264151), F(0)], [F(0.35), F(0.3), F(0.35), F(0)]
        sky = [mix1(horizon[k], zenith[k], sky_t) for k in range(4)]
        sun = self.pow(max(ZERO, dot3(d, [F(0.1), F(1.0), F(0.1)])), F(500.0)) * F(0.1)
        add = sun * (ONE if g2s >= ONE else ZERO)
        return [mix1(ground[k], sky[k], g2s) + add for k in range(4)]

    def trace(self, o, d):                              # wgsl:229-282
        color, light = [ONE] * 4, [ZERO] * 4
        for _ in range(self.bounces + 1):
            self.rays += 1
            hit = self.closest(o, d)
            if hit is None:
                if self.toggle != 0:
                    env = self.environment(d)
                    light = [light[k] + env[k] * color[k] for k in range(4)]
                break
            (dst, hp, n), s = hit
            unit = normalize(d)
            smooth = F(s["smoothness"])
            mcolor = [F(x) for x in s["color"]]
            if smooth == F(-1.0):                       # glass, wgsl:240-264
                mcolor = [ONE] * 4
                front = not (dot3(d, n) > ZERO)
                ratio = (ONE / F(1.5)) if front else F(1.5)
                cos_t = min(dot3([-x for x in unit], n), ONE)
                sin_t = np.sqrt(ONE - cos_t * cos_t)
                r0 = (ONE - ratio) / (ONE + ratio)
                r0 = r0 * r0
                x = ONE - cos_t
                refl = r0 + (ONE - r0) * (((x * x) * (x * x)) * x)
                if ratio * sin_t > ONE or refl > self.rand():
                    nd = reflect(unit, n)
                else:                                   # the shader's own refract, wgsl:290-295 (length, not its square)
                    c2 = min(dot3([-x for x in unit], n), ONE)
                    perp = [v * ratio for v in fma3(n, c2, unit)]
                    k = -np.sqrt(abs(ONE - np.sqrt(dot3(perp, perp))))
                    nd = fma3(n, k, perp)
            else:                                       # wgsl:265-269
                g = [self.rand_normal_dist(), self.rand_normal_dist(), self.rand_normal_dist()]
                dirn = normalize(g)
                dd = dot3(n, dirn)
                sg = ONE if dd > ZERO else (F(-1.0) if dd < ZERO else ZERO)
                diffuse = [v * sg for v in dirn]
                spec = reflect(unit, n)
                nd = [mix1(diffuse[k], spec[k], smooth) for k in range(3)]
            emitted = [F(s["emission_color"][k]) * F(s["emission_strength"]) for k in range(4)]
            light = [light[k] + emitted[k] * color[k] for k in range(4)]
            color = [color[k] * mcolor[k] for k in range(4)]
            o, d = hp, nd
        return light

    def frag(self, px, py, frames):                     # wgsl:306-327
        sw, sh = F(self.W), F(self.H)
        lin = (F(py) * sh) * sw + F(px) * sw
        seed = 0 if not lin > 0 else (0xFFFFFFFF if lin >= F(4294967296.0) else int(lin))
        self.rng = (seed + abs(frames) * 71939) & 0xFFFFFFFF
        cam = self.cam
        v3 = lambda name: [F(x) for x in cam[name]]
        total = [ZERO] * 4
        for _ in range(self.rpp + 1):
            aax, aay = self.rand(), self.rand()
            posx, posy = (F(px) + aax) / sw, (F(py) + aay) / sh
            rd = [F(cam["lens_radius"]) * v for v in self.rand_in_unit_disk()]
            offset = fma3(v3("v"), rd[1], [u * rd[0] for u in v3("u")])
            o = [v3("origin")[k] + offset[k] for k in range(3)]
            t = fma3(v3("vertical"), posy, fma3(v3("horizontal"), posx, v3("lower_left_corner")))
            d = [t[k] - o[k] for k in range(3)]
            light = self.trace(o, d)
            total = [total[k] + light[k] for k in range(4)]
        return [total[k] / F(self.rpp) for k in range(4)]

    def pixel(self, px, py, n_frames):                  # wgsl:50-67 over frames 0 .. n-1
        acc = [ZERO] * 4
        for f in range(n_frames):
            fr = self.frag(px, py, f)
            if f >= 1:
                w = ONE / F(f + 1)
                acc = [acc[k] * (ONE - w) + fr[k] * w for k in range(4)]
            else:
                acc = fr
        return np.array(acc, np.float32)


@pytest.mark.parametrize("name,toggle,bounces", [("balls", 1, 3), ("metal", 1, 4), ("balls", 0, 2), ("room", 0, 4), ("room", 1, 2)])
def test_second_restatement_agrees_with_the_oracle(oracle, name, toggle, bounces):
    from oracle import host_ref as hr
    W, H, frames = 24, 16, 2
    with np.errstate(all="ignore"):
        scene = getattr(hr, "scene_" + name)(np.float32(W) / np.float32(H))
        cam, sp, vs, ix, ms = scene
        sc = (cam.view(oracle.CAMERA_DT), sp.view(oracle.SPHERE_DT), vs.view(oracle.VERTEX_DT), ix, ms.view(oracle.MESH_DT))
        ref = oracle.render(sc, W, H, frames, bounces, 1, toggle, 0)
        sh = Shader(oracle, sc, W, H, bounces, 1, toggle)
        rng = np.random.default_rng(3)
        pixels = {(0, 0), (W - 1, H - 1), (W // 2, H // 2)} | {(int(x), int(y)) for x, y in zip(rng.integers(0, W, 14), rng.integers(0, H, 14))}
        n_hit = 0
        for px, py in sorted(pixels):
            got = sh.pixel(px, py, frames)
            want = ref[py, px]
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (name, px, py, got, want)
            n_hit += 1
        assert n_hit >= 10 and sh.rays > 2 * frames * n_hit       # bounces were followed, not only primary misses
