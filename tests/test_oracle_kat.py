"""Pins of the CPU oracle (no GPU): known answers derived from the reference's formulas
(the reference ships no tests or golden vectors of its own — SURVEY.md §4), struct layouts,
analytic intersections, accumulation semantics."""
import ctypes as C

import numpy as np
import pytest


def test_record_layouts(oracle):
    # Params 28 B, CameraUniform 128 B, Sphere 64 B, Mesh 80 B, Vertex 32 B
    o = oracle
    assert o.PARAMS_DT.itemsize == 28 and o.PARAMS_DT.fields["frames"][1] == 20
    c = o.CAMERA_DT.fields
    assert o.CAMERA_DT.itemsize == 128
    assert (c["origin"][1], c["lower_left_corner"][1], c["horizontal"][1], c["vertical"][1]) == (0, 16, 32, 48)
    assert (c["near"][1], c["w"][1], c["u"][1], c["v"][1], c["lens_radius"][1]) == (64, 80, 96, 112, 124)
    s = o.SPHERE_DT.fields
    assert o.SPHERE_DT.itemsize == 64 and (s["radius"][1], s["color"][1], s["emission_color"][1], s["emission_strength"][1], s["smoothness"][1]) == (12, 16, 32, 48, 52)
    m = o.MESH_DT.fields
    assert o.MESH_DT.itemsize == 80 and (m["first"][1], m["triangles"][1], m["offset"][1], m["pos"][1], m["color"][1], m["emission_color"][1], m["emission_strength"][1], m["specular"][1]) == (0, 4, 8, 16, 32, 48, 64, 68)
    v = o.VERTEX_DT.fields
    assert o.VERTEX_DT.itemsize == 32 and (v["pos"][1], v["normal"][1]) == (0, 16)


def test_rng_known_answers(oracle):
    # PCG-RXS-M-XS-32 hash stream, wgsl:205-210
    assert oracle.rng_stream(0, 4) == [0x07bb2fe2, 0x22b6b6bc, 0x3bf6e0b1, 0x572f7439]
    assert oracle.rng_stream(1, 4) == [0xa8beea3c, 0x2679c518, 0x97aaf6c6, 0x2a521372]
    assert oracle.rng_stream(800, 2) == [0xbb89a82a, 0x19ca7cc7]
    assert oracle.rng_stream(71939, 2) == [0xd2efbb71, 0xb05480e4]
    assert oracle.rng_stream(0xFFFFFFFF, 2) == [0xe62a4902, 0x498239f9]
    r = oracle.rand_stream(0, 4)
    assert np.allclose(r, [0.0301999971, 0.135600492, 0.234235808, 0.340567827], rtol=0, atol=1e-9)
    # rand = f32(r) / 2^32 exactly, and can reach 1.0 (f32(r) rounds up to 2^32)
    assert np.float32(0xFFFFFFFF) / np.float32(4294967296.0) == np.float32(1.0)


def test_rng_matches_independent_python_restatement(oracle):
    def stream(seed, n):
        out, s = [], seed
        for _ in range(n):
            s = (s * 747796405 + 2891336453) & 0xFFFFFFFF
            r = (((s >> ((s >> 28) + 4)) ^ s) * 277803737) & 0xFFFFFFFF
            out.append(((r >> 22) ^ r) & 0xFFFFFFFF)
        return out
    for seed in (0, 1, 12345, 0x80000000, 0xDEADBEEF):
        assert oracle.rng_stream(seed, 64) == stream(seed, 64)


def test_pixel_seed(oracle):
    # wgsl:307-308: u32(f32(y*H)*f32(W) + f32(x*W)) + u32(|frames|)*71939, row stride is H (not W)
    ps = oracle.pixel_seed
    assert ps(0, 0, 800, 800, 0) == 0
    assert ps(1, 0, 800, 800, 0) == 800
    assert ps(0, 1, 800, 800, 0) == 800 * 800
    assert ps(3, 2, 800, 800, 0) == 2 * 800 * 800 + 3 * 800
    assert ps(0, 0, 800, 800, 1) == 71939 and ps(0, 0, 800, 800, -1) == 71939      # abs(frames)
    assert ps(5, 7, 1920, 1080, 3) == (int(np.float32(np.float32(7 * 1080) * np.float32(1920)) + np.float32(5 * 1920)) + 3 * 71939) & 0xFFFFFFFF
    # W > H: (x, y) and (x + H, y - 1) collide
    assert ps(10, 5, 1920, 1080, 0) == ps(10 + 1080, 4, 1920, 1080, 0)
    # 4K: the f32 value passes 2^32 from row 518 on -> saturates (strict mode, WGSL rule)
    assert ps(0, 517, 3840, 2160, 0) < 0xFFFFFFFF
    assert ps(0, 518, 3840, 2160, 0) == 0xFFFFFFFF
    assert ps(100, 2000, 3840, 2160, 2) == (0xFFFFFFFF + 2 * 71939) & 0xFFFFFFFF  # wrapping add after saturation


def test_ray_sphere_analytic(oracle):
    # ray through the centre of a unit sphere at z=-5: near root only, normal faces the ray
    h = oracle.ray_sphere((0, 0, 0), (0, 0, -1), (0, 0, -5), 1.0)
    assert h[0] == 1 and h[1] == 4.0 and tuple(h[2:5]) == (0, 0, -4) and tuple(h[5:8]) == (0, 0, 1)
    # un-normalised direction: dst is in units of |dir| (wgsl:109)
    h = oracle.ray_sphere((0, 0, 0), (0, 0, -2), (0, 0, -5), 1.0)
    assert h[0] == 1 and h[1] == 2.0
    # origin inside the sphere: the near root is negative -> no hit (wgsl:113-115)
    assert oracle.ray_sphere((0, 0, -5), (0, 0, -1), (0, 0, -5), 1.0)[0] == 0
    # sphere behind the ray, and a clean miss
    assert oracle.ray_sphere((0, 0, 0), (0, 0, 1), (0, 0, -5), 1.0)[0] == 0
    assert oracle.ray_sphere((0, 0, 0), (0, 1, 0), (0, 0, -5), 1.0)[0] == 0
    # tangent ray: discriminant == 0 is a hit (>=)
    assert oracle.ray_sphere((1, 0, 0), (0, 0, -1), (0, 0, -5), 1.0)[0] == 1


def _tri(a, b, c, na=(0, 0, 1), nb=(0, 0, 1), nc=(0, 0, 1)):
    return np.array([*a, *na, *b, *nb, *c, *nc], np.float32)


def test_ray_triangle_analytic(oracle):
    # counter-clockwise seen from +z; the ray travels -z, so det = -dot(dir, n) > 0
    t = _tri((0, 0, 0), (1, 0, 0), (0, 1, 0))
    h = oracle.ray_triangle((0.25, 0.25, 1), (0, 0, -1), t)
    assert h[0] == 1 and h[1] == 1.0 and tuple(h[2:5]) == (0.25, 0.25, 0) and tuple(h[5:8]) == (0, 0, 1)
    # back face is culled (one-sided test, wgsl:141)
    assert oracle.ray_triangle((0.25, 0.25, -1), (0, 0, 1), t)[0] == 0
    # outside the triangle, behind the origin
    assert oracle.ray_triangle((0.75, 0.75, 1), (0, 0, -1), t)[0] == 0
    assert oracle.ray_triangle((0.25, 0.25, -1), (0, 0, -1), t)[0] == 0
    # edges and vertices count (>= 0)
    assert oracle.ray_triangle((0.5, 0.0, 1), (0, 0, -1), t)[0] == 1
    assert oracle.ray_triangle((0.0, 0.0, 1), (0, 0, -1), t)[0] == 1
    # the 1e-6 threshold is on the un-normalised determinant (wgsl:141)
    assert oracle.ray_triangle((0.25, 0.25, 1), (0, 0, -1e-7), t)[0] == 0
    # shading normal: barycentric mix of the vertex "normals", then normalised (wgsl:143);
    # weights w -> a, u -> b, v -> c
    t2 = _tri((0, 0, 0), (1, 0, 0), (0, 1, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1))
    h = oracle.ray_triangle((0.5, 0.25, 1), (0, 0, -1), t2)
    n = np.array([0.25, 0.5, 0.25], np.float32)
    assert np.allclose(h[5:8], n / np.linalg.norm(n), atol=1e-6)


def test_tie_rule_earliest_primitive_wins(oracle):
    # two identical spheres: strict '<' keeps the first (wgsl:154) -> material of sphere 0
    from oracle import host_ref as hr
    cam = hr.camera_to_uniform((0, 0, 0), (0, 0, -1), (0, 1, 0), 45.0, 1.0, 0.1, 100.0, 0.0, 1.0)
    sp = np.concatenate([hr.sphere((0, 0, -5), 1.0, (1, 1, 1, 1), (1, 0, 0, 1), 1.0, 0.0),
                         hr.sphere((0, 0, -5), 1.0, (1, 1, 1, 1), (0, 1, 0, 1), 1.0, 0.0)])
    empty = (np.zeros(0, oracle.VERTEX_DT), np.zeros(0, np.uint32), np.zeros(0, oracle.MESH_DT))
    ids = np.zeros((8, 8), np.uint32)
    img = oracle.render((cam, sp) + empty, 8, 8, 1, 0, 1, 0, 0, ids)
    assert ids[4, 4] == 0 and img[4, 4, 0] > 0 and img[4, 4, 1] == 0
    ids2 = np.zeros((8, 8), np.uint32)
    img2 = oracle.render((cam, sp[::-1].copy()) + empty, 8, 8, 1, 0, 1, 0, 0, ids2)
    assert ids2[4, 4] == 0 and img2[4, 4, 1] > 0 and img2[4, 4, 0] == 0


def test_sample_and_segment_counts(oracle):
    # rpp+1 samples divided by rpp; bounces+1 segments (wgsl:312,326,233). An emissive sphere
    # filling the view: every path is a single... no: emission is added per hit, paths continue.
    from oracle import host_ref as hr
    cam = hr.camera_to_uniform((0, 0, 0), (0, 0, -1), (0, 1, 0), 45.0, 1.0, 0.1, 100.0, 0.0, 1.0)
    sp = hr.sphere((0, 0, -5), 4.5, (0, 0, 0, 0), (1, 1, 1, 1), 1.0, 0.0)    # black emitter: color 0 kills later bounces
    empty = (np.zeros(0, oracle.VERTEX_DT), np.zeros(0, np.uint32), np.zeros(0, oracle.MESH_DT))
    for rpp in (1, 2, 3):
        ctr = np.zeros(3, np.uint64)
        img = oracle.render((cam, sp) + empty, 4, 4, 1, 0, rpp, 0, 0, None, ctr)
        assert int(ctr[0]) == 16 * (rpp + 1)                                  # bounces=0 -> one segment per sample
        assert np.allclose(img[..., 0], (rpp + 1) / rpp)                      # (rpp+1) samples of radiance 1, / rpp
    ctr = np.zeros(3, np.uint64)
    oracle.render((cam, sp) + empty, 4, 4, 1, 2, 1, 0, 0, None, ctr)
    assert 16 * 2 <= int(ctr[0]) <= 16 * 2 * 3                                 # at most bounces+1 = 3 segments per sample


def test_accumulation_semantics(oracle):
    # wgsl:59-66: frames>=1 blends with weight 1/(frames+1); frames<=0 overwrites
    from oracle import host_ref as hr
    sc = hr.scene_balls(1.0)
    w = h = 16
    f0 = oracle.render(sc, w, h, 1, 2, 1, 1, 0)
    f1_alone = oracle.render(sc, w, h, 1, 2, 1, 1, -1)            # frames=-1 -> overwrite, seed uses abs(-1)=1
    start = oracle.render(sc, w, h, 1, 2, 1, 1, 1)                # start-up dispatch: frames=1 blends with zeros
    assert np.array_equal(start, (np.float32(0) * np.float32(0.5) + f1_alone * np.float32(0.5)).astype(np.float32))
    two = oracle.render(sc, w, h, 2, 2, 1, 1, 0)                  # frames 0 then 1
    expect = f0 * np.float32(0.5) + f1_alone * np.float32(0.5)
    assert np.array_equal(two, expect.astype(np.float32))
    # x_step / y_step only touch the sampled pixels
    sub = oracle.render(sc, w, h, 1, 2, 1, 1, 0, x_step=4, y_step=4)
    assert np.array_equal(sub[::4, ::4], f0[::4, ::4]) and not sub[1::4].any()


def test_environment_light(oracle):
    e = oracle.environment_light
    assert np.allclose(e((0, -1, 0)), (0.35, 0.3, 0.35, 0.0))                       # ground
    up = e((0, 1, 0))                                                               # zenith + sun: pow(1,500)*0.1
    assert np.allclose(up, (0.0788092 + 0.1, 0.36480793 + 0.1, 0.7264151 + 0.1, 0.1), atol=2e-6)
    assert np.allclose(e((1, 0, 0))[:3], (1, 1, 1), atol=1e-6)                      # horizon (y=0): g2s = 1, sun = pow(0.1,500) = 0
    big = e((0, 3, 0))                                                              # un-normalised dir: pow(3,500) overflows
    assert np.isinf(big[0])


def test_oracle_threads_do_not_change_the_image(oracle):
    from oracle import host_ref as hr
    sc = hr.scene_room(1.5)
    a = oracle.render(sc, 48, 32, 2, 3, 1, 1, 0, nthreads=1)
    b = oracle.render(sc, 48, 32, 2, 3, 1, 1, 0, nthreads=0)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
