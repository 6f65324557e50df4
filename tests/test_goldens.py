"""Golden vectors (tests/golden/oracle_goldens.npz, made by make_oracle_goldens.py):
the oracle must reproduce them bit for bit on any host, and so must the CUDA path."""
import importlib.util
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_oracle_goldens", os.path.join(HERE, "golden", "make_oracle_goldens.py"))
gold = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(gold)
Z = np.load(os.path.join(HERE, "golden", "oracle_goldens.npz"))


@pytest.mark.parametrize("name", sorted(gold.CASES))
def test_oracle_reproduces_goldens(oracle, rtlib, name):
    from ray_tracer_b200 import host
    host.build()
    img, ids, rays = gold.render_case(name)
    assert rays == int(Z[name + "/rays"][0])
    assert np.array_equal(ids, Z[name + "/ids"])
    assert np.array_equal(img.view(np.uint32), Z[name + "/image_bits"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(gold.CASES))
def test_cuda_reproduces_goldens(rtlib, name):
    kind, w, h, frames, bounces, rpp, sky, first = gold.CASES[name]
    cam, sp, vs, ix, ms = gold.scene_for(name, kind, w, h)
    tr = rtlib.RayTracer(0, w, h)
    tr.upload_scene(sp, vs, ix, ms)
    tr.upload_textures(gold.textures_for(name, kind))
    tr.set_camera(cam)
    tr.set_params(number_of_bounces=bounces, rays_per_pixel=rpp, skybox=sky, frames=first)
    tr.set_option(rtlib._lib.RT_OPT_PRIMARY_IDS, 1)
    tr.render_frames(first, frames)
    img, ids, ctr = tr.read_accum(), tr.read_primary_ids(), tr.counters()
    tr.close()
    assert ctr["rays"] == int(Z[name + "/rays"][0])
    assert np.array_equal(ids, Z[name + "/ids"])
    assert np.array_equal(img.view(np.uint32), Z[name + "/image_bits"])
