"""Rows f1-f3 of SURVEY.md 8(f): the fly-cam step and input policy just before the path
(camera.rs:122-247, context.rs:148-187), loaders with texture coordinates, PNG in/out just
after it.  C++ host mirror vs the numpy restatement / PIL."""
import os

import numpy as np
import pytest

ASSETS = "/root/reference/assets"


def _host():
    from ray_tracer_b200 import host
    host.build()
    return host


def test_update_camera_matches_restatement():
    host = _host()
    from oracle import host_ref as hr
    rng = np.random.default_rng(7)
    o, la = np.float32([3.089, 1.53, -3.0]), np.float32([-2.0, -1.0, 2.0])     # Scene::balls camera
    c = hr.controller_new()
    for step in range(40):
        if step % 5 == 0:
            hr.process_keyboard(c, ["W", "A", "Space", "S", "D", "LShift", "Up", "Other"][(step // 5) % 8], step % 10 == 0)
        c["rotate_horizontal"] = np.float32(rng.normal() * 3.0)
        c["rotate_vertical"] = np.float32(rng.normal() * 3.0)
        c["scroll"] = np.float32(-10000.0 * (step % 7 == 0))
        dt = np.float32(rng.uniform(0.001, 0.03))
        ho, hla, hc = host.camera_update(o, la, {k: float(v) for k, v in c.items()}, dt)
        o, la = hr.update_camera(o, la, c, dt)
        assert np.allclose(ho, o, rtol=0, atol=2e-6 * max(1.0, np.abs(o).max())), (step, ho, o)
        assert np.allclose(hla, la, rtol=0, atol=2e-6 * max(1.0, np.abs(la).max())), (step, hla, la)
        assert hc["scroll"] == 0.0 and hc["rotate_horizontal"] == 0.0 and hc["rotate_vertical"] == 0.0
        o, la = ho, hla                                  # keep both sides on the same trajectory
    # quirks: look_at is re-derived one unit ahead of the origin; no roll; the "pitch clamp" compares radians
    # with a degrees constant and never fires, so a huge vertical mouse move flips the camera over
    ho, hla, _ = host.camera_update([0, 0, 0], [0, 0, 1], {"rotate_vertical": -1000.0}, 1.0)
    assert abs(np.linalg.norm(hla - ho) - 1.0) < 1e-6


def test_process_keyboard_and_scroll_conventions():
    host = _host()
    # 5.0 while pressed, 0 on release; W == Up etc.; unknown keys are not consumed
    import ctypes as C
    L = host.hostlib()
    st = np.zeros(17, np.float32)
    assert L.rth_controller_key(st.ctypes.data_as(C.c_void_p), host.KEYS["Up"], 1) == 1 and st[8] == 5.0
    assert L.rth_controller_key(st.ctypes.data_as(C.c_void_p), host.KEYS["W"], 0) == 1 and st[8] == 0.0
    assert L.rth_controller_key(st.ctypes.data_as(C.c_void_p), host.KEYS["LShift"], 1) == 1 and st[11] == 5.0
    assert L.rth_controller_key(st.ctypes.data_as(C.c_void_p), host.KEYS["Other"], 1) == 0


def test_png_roundtrip_and_decoder_vs_pil(tmp_path):
    host = _host()
    from PIL import Image
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (37, 53, 4), dtype=np.uint8)
    p = str(tmp_path / "a.png")
    host.write_png(p, img)
    assert np.array_equal(np.asarray(Image.open(p).convert("RGBA")), img)          # our encoder, PIL's decoder
    assert np.array_equal(host.load_png(p), img)                                   # our encoder, our decoder
    for mode in ("RGB", "L", "LA", "P"):                                           # PIL's encoder (filters, colour types), our decoder
        q = str(tmp_path / ("b_%s.png" % mode))
        src = Image.fromarray(img).convert(mode)
        src.save(q, optimize=True)
        assert np.array_equal(host.load_png(q), np.asarray(src.convert("RGBA"))), mode
    with pytest.raises(host.HostError):
        host.load_png(str(tmp_path / "missing.png"))
    bad = str(tmp_path / "bad.png")
    open(bad, "wb").write(b"not a png")
    with pytest.raises(host.HostError):
        host.load_png(bad)


@pytest.mark.skipif(not os.path.isdir(ASSETS), reason="reference assets are only present in the authoring container")
def test_reference_assets_through_the_cpp_loaders():
    host = _host()
    from PIL import Image
    from ray_tracer_b200 import configs
    # cobble textures decode identically to PIL
    for fn in ("cobble-diffuse.png", "cobble-normal.png", "cube-normal.png"):
        assert np.array_equal(host.load_png(os.path.join(ASSETS, fn)), np.asarray(Image.open(os.path.join(ASSETS, fn)).convert("RGBA"))), fn
    # OBJ with vt kept == the committed fixture (positions, normals, uvs, indices)
    vs, ix, ms = host.load_model_uv(os.path.join(ASSETS, "cube.obj"))
    fv, fi, fm = configs.load_fixture("cube_obj", uv=True)
    assert np.array_equal(vs["pos"], fv["pos"]) and np.array_equal(vs["normal"], fv["normal"]) and np.array_equal(ix, fi)
    assert np.array_equal(vs["tu"], fv["tu"]) and np.array_equal(vs["tv"], fv["tv"])
    # the plain loader still zeroes the padding words, as the reference does
    pv, _, _ = host.load_model(os.path.join(ASSETS, "cube.obj"))
    assert not pv["tu"].any() and not pv["tv"].any()
    # GLB: TEXCOORD_0 kept, embedded PNG images decoded
    gv, gi, gm = host.load_model_uv(os.path.join(ASSETS, "cube.glb"))
    assert gv["tu"].any() and 0.0 <= gv["tu"].min() and gv["tu"].max() <= 1.0
    imgs = host.load_glb_images(os.path.join(ASSETS, "the_utah_teapot.glb"))
    assert len(imgs) == 1 and imgs[0].shape[2] == 4 and imgs[0].shape[0] > 0


@pytest.mark.gpu
def test_context_input_policy_and_png(tmp_path):
    """Every window event resets the accumulation (context.rs:149); WASD moves the camera through
    update(dt); the saved PNG is the sRGB view of the texture, top row first."""
    host = _host()
    from PIL import Image
    ctx = host.Context(0, 64, 48)
    ctx.set(skybox=1)
    for _ in range(3):
        ctx.update_dt(0.016); ctx.render()
    assert int(ctx.params["frames"][0]) == 3
    o0, _ = ctx.camera
    assert ctx.input_other() is False and int(ctx.params["frames"][0]) == -1     # any event resets
    assert ctx.input_key("W", True) is True
    ctx.update_dt(0.5); ctx.render()
    assert int(ctx.params["frames"][0]) == 0                                      # first frame after a reset overwrites
    o1, la1 = ctx.camera
    assert np.linalg.norm(o1 - o0) > 1.0                                          # 5.0 * speed 3 * 0.5 s = 7.5 units forward
    ctx.input_key("W", False)
    ctx.mouse_motion(10.0, 0.0)                                                   # ignored: button not held
    ctx.update_dt(0.016)
    _, la2 = ctx.camera
    assert np.allclose(la2 - ctx.camera[0], la1 - o1, atol=1e-5)
    ctx.input_mouse_button(True); ctx.mouse_motion(10.0, 0.0); ctx.update_dt(0.016)
    assert not np.allclose(ctx.camera[1] - ctx.camera[0], la1 - o1, atol=1e-4)    # yaw changed
    ctx.render(); ctx.sync()
    tex = ctx.read_texture()
    p = str(tmp_path / "frame.png")
    ctx.save_png(p)
    png = np.asarray(Image.open(p).convert("RGBA"))
    assert png.shape == (48, 64, 4)
    lin = np.clip(tex[::-1, :, :3], 0.0, 1.0)
    srgb = np.where(lin <= 0.0031308, 12.92 * lin, 1.055 * np.power(lin, 1 / 2.4) - 0.055)
    assert np.abs(png[:, :, :3].astype(np.int32) - np.rint(srgb * 255).astype(np.int32)).max() <= 1
    ctx.save_png(p, mirror_x=True)
    assert np.array_equal(np.asarray(Image.open(p).convert("RGBA")), png[:, ::-1])
    ctx.close()


@pytest.mark.parametrize("mode,subsampling,size", [("RGB", 2, (67, 45)), ("RGB", 0, (64, 48)), ("RGB", 1, (50, 33)), ("L", 0, (40, 24))])
def test_baseline_jpeg_decoder_against_pil(tmp_path, mode, subsampling, size):
    """host decode_jpeg (baseline Huffman, 4:2:0 / 4:2:2 / 4:4:4 / grey, sizes that are not MCU multiples) vs Pillow:
    different IDCT and chroma filters may differ by a few levels, never more."""
    PIL = pytest.importorskip("PIL.Image")
    from ray_tracer_b200 import host
    w, h = size
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    rgb = np.stack([127 + 120 * np.sin(xx / 9.0), 127 + 120 * np.cos(yy / 7.0), 127 + 100 * np.sin((xx + yy) / 13.0)], -1)
    img = PIL.fromarray(rgb.clip(0, 255).astype(np.uint8), "RGB").convert(mode)
    path = str(tmp_path / "t.jpg")
    kw = {} if mode == "L" else {"subsampling": subsampling}
    img.save(path, quality=92, **kw)
    got = host.load_png(path)                                        # load_image: PNG or JPEG by signature
    want = np.asarray(PIL.open(path).convert("RGBA"))
    assert got.shape == want.shape
    d = np.abs(got.astype(int) - want.astype(int))
    assert d.max() <= 6 and d.mean() < 0.6, (d.max(), d.mean())
    # progressive (SOF2: spectral selection + successive approximation) of the same image and settings carries the same
    # quantised coefficients in another order: the decode must be identical, byte for byte
    for extra in ({"progressive": True}, {"progressive": True, "optimize": True}):
        img.save(path, quality=92, **extra, **kw)
        assert b"\xff\xc2" in open(path, "rb").read(1000)
        assert np.array_equal(host.load_png(path), got)


def test_progressive_jpeg_larger_image_and_restarts(tmp_path):
    """A noisy 203x131 progressive file (long EOB runs, every scan type, correction bits) at two qualities, and a
    baseline file with restart intervals: progressive == baseline byte for byte, both within a few levels of Pillow."""
    PIL = pytest.importorskip("PIL.Image")
    from ray_tracer_b200 import host
    rng = np.random.default_rng(11)
    w, h = 203, 131
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    rgb = np.stack([127 + 100 * np.sin(xx / 5.0) * np.cos(yy / 11.0), 40 + xx, 200 - yy], -1) + rng.normal(0, 25, (h, w, 3))
    img = PIL.fromarray(rgb.clip(0, 255).astype(np.uint8), "RGB")
    for quality, sub in ((35, 2), (97, 0)):
        pb, pp = str(tmp_path / "b.jpg"), str(tmp_path / "p.jpg")
        img.save(pb, quality=quality, subsampling=sub)
        img.save(pp, quality=quality, subsampling=sub, progressive=True)
        base, prog = host.load_png(pb), host.load_png(pp)
        assert np.array_equal(base, prog)
        want = np.asarray(PIL.open(pp).convert("RGBA"))
        d = np.abs(prog.astype(int) - want.astype(int))
        assert d.max() <= 8 and d.mean() < 0.8, (d.max(), d.mean())
    # truncated progressive file: later scans missing -> still decodes (coarser), no exception; garbage -> HostError
    data = open(pp, "rb").read()
    cut = str(tmp_path / "cut.jpg")
    open(cut, "wb").write(data[: len(data) * 2 // 3])
    assert host.load_png(cut).shape == prog.shape
    open(cut, "wb").write(data[:2] + b"\x00" * 64)
    with pytest.raises(host.HostError):
        host.load_png(cut)


def _png_chunk(tag, body):
    import struct, zlib
    return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)


def _adam7_png(img, ctype, depth, palette=None):
    """Encode (H, W[, C]) uint8/uint16 samples as an Adam7-interlaced PNG (filter 0 on every row)."""
    import struct, zlib
    h, w = img.shape[:2]
    raw = b""
    for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
        sub = img[y0::dy, x0::dx]
        if sub.size == 0:
            continue
        for row in sub:
            if depth == 16:
                body = row.astype(">u2").tobytes()
            elif depth == 8:
                body = row.astype(np.uint8).tobytes()
            else:                                        # pack `depth`-bit samples, most significant bits first
                bits = np.unpackbits(row.astype(np.uint8).reshape(-1, 1), axis=1)[:, 8 - depth:].reshape(-1)
                body = np.packbits(bits).tobytes()
            raw += b"\x00" + body
    out = b"\x89PNG\r\n\x1a\n" + _png_chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 1))
    if palette is not None:
        out += _png_chunk(b"PLTE", palette.astype(np.uint8).tobytes())
    return out + _png_chunk(b"IDAT", zlib.compress(raw)) + _png_chunk(b"IEND", b"")


def test_png_interlace_and_small_bit_depths(tmp_path):
    """Adam7-interlaced files (hand-encoded: Pillow cannot write them; read back by Pillow as a cross-check) and 1 / 2 / 4-bit
    grey and palette files, sizes that leave some passes empty."""
    PIL = pytest.importorskip("PIL.Image")
    from ray_tracer_b200 import host
    rng = np.random.default_rng(21)
    p = str(tmp_path / "i.png")
    for (h, w) in ((13, 21), (1, 1), (2, 3), (8, 8), (5, 1)):
        rgba = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        open(p, "wb").write(_adam7_png(rgba, 6, 8))
        assert np.array_equal(host.load_png(p), rgba)
        assert np.array_equal(np.asarray(PIL.open(p).convert("RGBA")), rgba)
        rgb16 = rng.integers(0, 65536, (h, w, 3), dtype=np.uint16)
        open(p, "wb").write(_adam7_png(rgb16, 2, 16))
        got = host.load_png(p)
        assert np.array_equal(got[..., :3], (rgb16 >> 8).astype(np.uint8)) and (got[..., 3] == 255).all()
        for depth in (1, 2, 4):
            g = rng.integers(0, 1 << depth, (h, w), dtype=np.uint8)
            open(p, "wb").write(_adam7_png(g, 0, depth))
            want = (g * (255 // ((1 << depth) - 1))).astype(np.uint8)
            got = host.load_png(p)
            assert np.array_equal(got[..., 0], want) and np.array_equal(got[..., 2], want)
            assert np.array_equal(np.asarray(PIL.open(p).convert("L")), want)
            pal = rng.integers(0, 256, (1 << depth, 3), dtype=np.uint8)
            open(p, "wb").write(_adam7_png(g, 3, depth, pal))
            assert np.array_equal(host.load_png(p)[..., :3], pal[g])
    # non-interlaced small depths written by Pillow
    bw = PIL.fromarray((rng.integers(0, 2, (9, 19)) * 255).astype(np.uint8), "L").convert("1")
    bw.save(p)
    assert np.array_equal(host.load_png(p)[..., 0], np.asarray(bw.convert("L")))
    pim = PIL.fromarray(rng.integers(0, 4, (7, 11), dtype=np.uint8), "P")
    pim.putpalette([int(v) for v in rng.integers(0, 256, 12)])
    pim.save(p, bits=2)
    assert np.array_equal(host.load_png(p), np.asarray(PIL.open(p).convert("RGBA")))
