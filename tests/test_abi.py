"""The C-ABI library loads without a GPU and exports every symbol include/rt_b200.h declares;
argument checking that needs no device works; there is no CPU fallback."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(rtlib):
    L = rtlib.lib()
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), "librt_b200.so does not export " + n
    assert set(names) == set(rtlib._lib.SYMBOLS), set(names) ^ set(rtlib._lib.SYMBOLS)


def test_version_and_null_handling(rtlib):
    L = rtlib.lib()
    assert L.rt_version() == 1
    assert L.rt_last_error(None) is not None
    L.rt_destroy(None)                                     # NULL is allowed
    assert L.rt_sync(None) == rtlib._lib.RT_ERR_INVALID_ARGUMENT
    assert L.rt_create(None, 0, 8, 8) == rtlib._lib.RT_ERR_INVALID_ARGUMENT
    h = C.c_void_p()
    assert L.rt_create(C.byref(h), 0, 0, 8) == rtlib._lib.RT_ERR_INVALID_ARGUMENT
    assert b"zero-sized" in L.rt_last_error(None)


def test_no_cpu_fallback(rtlib):
    """Without a CUDA device rt_create fails loudly instead of rendering on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rtlib.RtError) as e:
        rtlib.RayTracer(0, 16, 16)
    assert e.value.code in (rtlib._lib.RT_ERR_CUDA, rtlib._lib.RT_ERR_INVALID_ARGUMENT)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "ray_tracer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in text and "from oracle" not in text and "rto_" not in text, os.path.join(dirpath, f)


def test_host_library_loads(rtlib):
    from ray_tracer_b200 import host
    host.build()
    L = host.hostlib()
    for n in ("rth_camera_to_uniform", "rth_scene_builtin", "rth_scene_from_model", "rth_context_new", "rth_context_update",
              "rth_context_render", "rth_context_run_frames", "rth_context_resize", "rth_context_clear_accululation"):
        assert hasattr(L, n)


def test_rust_binding_declares_exactly_the_header():
    """rust/src/lib.rs (source only: no rustc in this image) binds every rt_* entry point of the
    header, nothing else, and carries the same option numbers."""
    rs = open(os.path.join(ROOT, "rust", "src", "lib.rs")).read()
    assert sorted(set(re.findall(r"pub fn (rt_[a-z0-9_]+)", rs))) == _declared()
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "rt_b200.h")).read(), flags=re.S)
    opts = dict(re.findall(r"RT_OPT_([A-Z_]+) = (\d+)", hdr))
    for name, val in re.findall(r"pub const ([A-Z_]+): i32 = (\d+);", rs):
        assert opts[name] == val, name


def test_fastdiv_is_exact(tmp_path):
    """rt::FastDiv (the item decode's division by launch constants) against `/` on the host: edge values,
    multiples of the divisor and their neighbours, a pseudo-random sweep, every dividend below 2^26."""
    import shutil, subprocess
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    exe = str(tmp_path / "fastdiv_check")
    subprocess.run([nvcc, "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I", os.path.join(ROOT, "ray_tracer_b200", "csrc"),
                    "-o", exe, os.path.join(ROOT, "tests", "native", "fastdiv_check.cu")], check=True, capture_output=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    assert out.startswith("OK"), out


def test_sky_pow_shortcuts_hold(tmp_path):
    """environment_light (rt_shade.cuh) skips pow(t, 0.35) for t in {0, 1} and pow(x, 500) for x <= 0.8: the
    identities behind that, against the oracle's pow (every 97th float of [0, 0.8]; `full` walks them all)."""
    import subprocess
    exe = str(tmp_path / "sky_shortcuts_check")
    subprocess.run(["g++", "-O2", "-fopenmp", "-ffp-contract=off", "-o", exe, os.path.join(ROOT, "tests", "native", "sky_shortcuts_check.cpp")],
                   check=True, capture_output=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    assert out.startswith("OK"), out
