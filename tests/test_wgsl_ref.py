"""oracle/_ref — the reference's own shader, machine-translated (oracle/make_ref.py) — against the
hand-written oracle.  This is the pin of the parity chain: loop bounds, RNG order, branch
structure, tie rule and record layouts of `_ref` come from the text of
/root/reference/shaders/ray_tracer.wgsl, not from a reading of it.

  contract build : must equal the hand oracle BIT FOR BIT (image, primary ids, every counter)
  plain build    : WGSL-spec formulas + glibc; tolerance tiers of SURVEY.md §8(c) (T1 ids, T2 radiance)

CPU only.  Skipped when oracle/_ref can neither be built (no /root/reference) nor was prebuilt.
"""
import importlib.util
import os

import numpy as np
import pytest

from oracle import make_ref

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_oracle_goldens", os.path.join(HERE, "golden", "make_oracle_goldens.py"))
gold = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(gold)

pytestmark = pytest.mark.skipif(make_ref.build() is None, reason="oracle/_ref not built and /root/reference not present")

MISS = 0xFFFFFFFF


@pytest.fixture(scope="module")
def wr():
    from oracle import wgsl_ref
    return wgsl_ref


def both(oracle, wr, scene, w, h, frames, bounces, rpp, sky, first, contract=True, x_step=1, y_step=1):
    ids_o = np.full((h, w), MISS, np.uint32); ctr_o = np.zeros(3, np.uint64)
    with oracle.stats() as st:
        img_o = oracle.render(scene, w, h, frames, bounces, rpp, sky, first, ids_o, ctr_o, x_step, y_step)
    ids_r = np.full((h, w), MISS, np.uint32); ctr_r = np.zeros(3, np.uint64)
    aov = wr.Aov(w, h, contract)
    img_r = wr.render(scene, w, h, frames, bounces, rpp, sky, first, ids_r, ctr_r, x_step, y_step, contract=contract, aov=aov)
    return (img_o, ids_o, ctr_o, st), (img_r, ids_r, ctr_r, aov)


def assert_identical(o, r, what):
    (img_o, ids_o, ctr_o, st), (img_r, ids_r, ctr_r, aov) = o, r
    assert int(aov.mismatches[0]) == 0, what
    assert np.array_equal(ids_o, ids_r), what + ": primary ids"
    assert ctr_o.tolist() == ctr_r.tolist(), what + ": segments / sphere tests / triangle tests"
    assert st.rng_draws == aov.calls("next_random_number"), what + ": RNG draws"
    assert st.samples == aov.calls("trace"), what + ": camera paths"
    assert int(ctr_o[0]) == aov.calls("calculate_ray_collions"), what
    a, b = img_o.view(np.uint32), img_r.view(np.uint32)
    nan = np.isnan(img_o) & np.isnan(img_r)               # 0/0 (rpp = 0): payload is not pinned by anything
    assert np.array_equal(a[~nan], b[~nan]), "%s: %d f32 words differ" % (what, int((a != b)[~nan].sum()))


# ---- the translator ---------------------------------------------------------------------------
def test_translation_covers_the_shader(wr):
    names = wr.fn_names()
    # every function the hot path names exists in the reference's text and was translated
    for fn in ("main", "frag", "trace", "calculate_ray_collions", "ray_sphere", "ray_triangle", "rand", "next_random_number",
               "rand_normal_dist", "rand_unit_sphere", "rand_hemisphere_dir_dist", "rand_in_unit_disk", "reflectance",
               "refract", "get_environment_light"):
        assert fn in names
    assert wr.lib().wgslref_workgroup_size() == b"(8,8)"


def test_record_layouts_from_wgsl_rules(wr):
    """Sizes the WGSL alignment rules give the shader's own structs == the Rust #[repr(C)] records == rt_b200.h."""
    from oracle import rt_oracle as ro
    assert wr.struct_layout("Sphere") == (ro.SPHERE_DT.itemsize, 16) == (64, 16)
    assert wr.struct_layout("Mesh") == (ro.MESH_DT.itemsize, 16) == (80, 16)
    assert wr.struct_layout("Vertex") == (ro.VERTEX_DT.itemsize, 16) == (32, 16)
    assert wr.struct_layout("Camera") == (ro.CAMERA_DT.itemsize, 16) == (128, 16)
    assert wr.struct_layout("Params") == (24, 4)          # the shader reads the first 24 of the host's 28 bytes
    assert wr.struct_layout("Material") == (48, 16)


def test_translator_rules():
    src = """
    struct A { x: vec3<f32>, y: f32, };
    struct B { a: A, k: u32 }
    @group(0) @binding(0) var<storage,read> arr: array<B>;
    fn next(s: ptr<function,u32>) -> u32 { *s = *s + 1u; return *s; }
    fn f(v: A, s: ptr<function,u32>) -> f32 {
        var v: A = v;
        let p = vec2<f32>(f32(next(s)), f32(next(s)));
        if p.x >= 1.5 { v.y += 2.0; }
        return v.y * 0x1p-1f + p.y;
    }
    @compute @workgroup_size(4,2) fn main(@builtin(global_invocation_id) id: vec3<u32>) { var s: u32 = id.x; let q = f(arr[0].a, &s); }
    """
    code, mod = make_ref.generate(src, "<test>")
    assert "struct A {" in code and code.index("struct A {") < code.index("struct B {")
    assert "layout<ref::A> { enum { size = 16, align = 16 }" in code        # vec3 (12, align 16) + f32 packed at 12
    assert "layout<ref::B> { enum { size = 32, align = 16 }" in code
    assert "A v = v_param;" in code and "f_body(A v_param, u32* s)" in code   # shadowed parameter renamed
    assert "vec2<f32>{to_f32(next(s)), to_f32(next(s))}" in code              # sequenced constructor arguments
    assert "if ( p.x >= 1.5f)" in code and "v.y += 2.0f;" in code and "0x1p-1f" in code
    assert "static void main_(vec3<u32> id)" in code and "#define WGSL_WORKGROUP_SIZE 4,2" in code
    with pytest.raises(SyntaxError):                                          # two side-effecting calls as call arguments: refused
        make_ref.generate(src.replace("vec2<f32>(f32(next(s)), f32(next(s)))", "max(f32(next(s)), f32(next(s)))"), "<test>")


# ---- contract build == hand oracle, bit for bit ---------------------------------------------------
@pytest.mark.parametrize("name", [n for n in sorted(gold.CASES) if n != "c2_cube"])     # c2_cube = texture extension (no reference counterpart)
def test_contract_ref_equals_hand_oracle(oracle, wr, name):
    kind, w, h, frames, bounces, rpp, sky, first = gold.CASES[name]
    o, r = both(oracle, wr, gold.scene_for(name, kind, w, h), w, h, frames, bounces, rpp, sky, first)
    assert_identical(o, r, name)


@pytest.mark.parametrize("name", [n for n in sorted(gold.CASES) if n != "c2_cube"])
def test_committed_goldens_are_ref_outputs(wr, name):
    """tests/golden/oracle_goldens.npz (what the GPU tests compare against on the box) == oracle/_ref's render."""
    Z = np.load(os.path.join(HERE, "golden", "oracle_goldens.npz"))
    img, ids, rays = gold.render_case_wgsl_ref(name)
    assert int(Z[name + "/pinned_by_wgsl_ref"][0]) == 1
    assert rays == int(Z[name + "/rays"][0])
    assert np.array_equal(ids, Z[name + "/ids"])
    assert np.array_equal(img.view(np.uint32), Z[name + "/image_bits"])


QUIRKS = {  # name -> (scene, w, h, frames, bounces, rpp, sky, first_frame): loop bounds, frames logic
    "rpp0_div0": ("balls", 16, 12, 1, 2, 0, 1, 0),
    "rpp_minus1": ("balls", 16, 12, 1, 2, -1, 1, 0),
    "rpp3": ("metal", 16, 12, 1, 5, 3, 1, 0),
    "bounces_minus1": ("room", 16, 12, 2, -1, 1, 1, 0),
    "bounces0": ("room", 16, 12, 2, 0, 1, 0, 0),
    "bounces50_glass": ("metal", 20, 12, 1, 50, 1, 1, 0),
    "frames_minus1": ("balls", 16, 12, 1, 3, 1, 1, -1),
    "frames_startup_1": ("room", 16, 12, 3, 3, 1, 0, 1),
    "thin_lens": ("random_balls", 20, 12, 2, 4, 1, 1, 7),
    "wide_aspect_seed_alias": ("balls", 48, 8, 1, 2, 1, 1, 0),
}


@pytest.mark.parametrize("name", sorted(QUIRKS))
def test_contract_ref_equals_hand_oracle_on_quirks(oracle, wr, name):
    from oracle import host_ref as hr
    scene, w, h, frames, bounces, rpp, sky, first = QUIRKS[name]
    sc = getattr(hr, "scene_" + scene)(np.float32(w) / np.float32(h))
    o, r = both(oracle, wr, sc, w, h, frames, bounces, rpp, sky, first)
    assert_identical(o, r, name)


def test_contract_ref_equals_hand_oracle_at_4k_seed_saturation(oracle, wr):
    """3840x2160: the f32 seed exceeds 2^32 from row 518 on; a strided pixel set covers both regimes."""
    from oracle import host_ref as hr
    w, h = 3840, 2160
    sc = hr.scene_balls(np.float32(w) / np.float32(h))
    o, r = both(oracle, wr, sc, w, h, 2, 3, 1, 1, 0, x_step=479, y_step=127)
    assert_identical(o, r, "4k strided")
    assert int(o[2][0]) > 200


def test_wide_seed_mode_agrees(oracle, wr):
    """The labelled deviation RT_OPT_SEED_MODE = 1 (out-of-range u32(f32) wraps instead of saturating) in both CPU sides."""
    from oracle import host_ref as hr
    w, h = 3840, 2160
    sc = hr.scene_balls(np.float32(w) / np.float32(h))
    with oracle.seed_mode(1), wr.seed_mode(1):
        o, r = both(oracle, wr, sc, w, h, 2, 3, 1, 1, 0, x_step=479, y_step=127)
        assert_identical(o, r, "4k strided, wide")
        assert oracle.pixel_seed(0, 518, w, h, 0) == 0x176000
    s, _ = both(oracle, wr, sc, w, h, 2, 3, 1, 1, 0, x_step=479, y_step=127)
    assert np.array_equal(o[0][:518].view(np.uint32), s[0][:518].view(np.uint32))          # identical where the seed is below 2^32
    assert not np.array_equal(o[0][518:].view(np.uint32), s[0][518:].view(np.uint32))


def test_contract_ref_c1_baseline_size(oracle, wr):
    """C1 at its BASELINE.json size: `balls`, 800x800, 16 spp (8 dispatches), 4 bounces (every 5th pixel: 25,600 pixels x 8 dispatches)."""
    from oracle import host_ref as hr
    sc = hr.scene_balls(np.float32(1.0))
    o, r = both(oracle, wr, sc, 800, 800, 8, 4, 1, 0, 0, x_step=5, y_step=5)
    assert_identical(o, r, "c1 800x800")


# ---- plain build: what changes when nothing is fused and libm replaces the fixed routines -----------
@pytest.mark.parametrize("name", ["balls", "room", "metal", "c3_teapot", "c4_tree"])
def test_plain_ref_within_tolerance(oracle, wr, name):
    kind, w, h, frames, bounces, rpp, sky, first = gold.CASES[name]
    w, h = 2 * w, 2 * h
    o, r = both(oracle, wr, gold.scene_for(name, kind, w, h), w, h, frames, bounces, rpp, sky, first, contract=False)
    (img_o, ids_o, ctr_o, st), (img_r, ids_r, ctr_r, aov) = o, r
    assert int(aov.mismatches[0]) == 0
    # T1: primary-hit ids (sample 0 of the first dispatch); an unfused camera ray may flip a silhouette pixel
    assert (ids_o != ids_r).mean() <= 0.002, name
    # T2: per-pixel radiance |d| <= 1e-3 * max(1, |ref|); paths decorrelate where a last-bit difference flips a branch
    ok = np.abs(img_o - img_r) <= 1e-3 * np.maximum(1.0, np.abs(img_r))
    frac = float(ok.all(axis=2).mean())
    assert frac >= {"balls": 0.99, "room": 0.93, "metal": 0.80, "c3_teapot": 0.97, "c4_tree": 0.99}[name], (name, frac)
    # segment counts move only through those flips
    assert abs(int(ctr_o[0]) - int(ctr_r[0])) <= 0.02 * int(ctr_o[0])
