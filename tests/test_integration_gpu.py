"""The compiled libYafaRay plug-ins (integration/): `class AcceleratorB200 : public Accelerator` and
`class PathIntegratorB200 : public TiledIntegrator` inside a patched copy of the reference, linked with liby4gpu.so and
driven through the REFERENCE'S OWN C API (yafaray_* calls replayed by oracle/harness/ref_render.cc) and its own
Accelerator::factory (oracle/harness/ref_raybatch.cc) -- the drop-in boundary of SURVEY.md 8(b), end to end."""
import os
import sys

import numpy as np
import pytest

import helpers
from libyafaray_b200 import scenes, scenes_pt

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not helpers.have_plugin(), reason="integration/_build not built (make -C integration)")]
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLD)
from make_golden import PT_GOLDEN, PT_GOLDEN_CITY, PT_GOLDEN_DOF, PT_GOLDEN_LIGHTS, PT_GOLDEN_MIRROR, PT_GOLDEN_TEST01, TEST01_DIR  # noqa: E402

ALL_GOLDEN = {**{k: ("pt_cornell.npz", v) for k, v in PT_GOLDEN.items()}, **{k: ("pt_city.npz", v) for k, v in PT_GOLDEN_CITY.items()},
              **{k: ("pt_dof.npz", v) for k, v in PT_GOLDEN_DOF.items()}, **{k: ("pt_test01.npz", v) for k, v in PT_GOLDEN_TEST01.items()},
              **{k: ("pt_mirror.npz", v) for k, v in PT_GOLDEN_MIRROR.items()},
              **{k: ("pt_lights.npz", v) for k, v in PT_GOLDEN_LIGHTS.items() if v.get("integrator") == "directlighting" or not v.get("extra_lights")}}
from make_golden_large import PT_GOLDEN_ANGULAR, PT_GOLDEN_AO, PT_GOLDEN_DARKSKY, PT_GOLDEN_GLASS, PT_GOLDEN_GLASS_ABSORB, PT_GOLDEN_IBL, PT_GOLDEN_PORTAL, PT_GOLDEN_SHINY2, PT_GOLDEN_SPOT_SOFT, PT_GOLDEN_SUNSKY, PT_GOLDEN_TEXBACK  # noqa: E402

ALL_GOLDEN.update({k: ("pt_glass.npz", v) for k, v in PT_GOLDEN_GLASS.items()})
ALL_GOLDEN.update({k: ("pt_shiny2.npz", v) for k, v in PT_GOLDEN_SHINY2.items()})
ALL_GOLDEN.update({k: ("pt_spot_soft.npz", v) for k, v in PT_GOLDEN_SPOT_SOFT.items()})
ALL_GOLDEN.update({k: ("pt_portal.npz", v) for k, v in PT_GOLDEN_PORTAL.items()})
ALL_GOLDEN.update({k: ("pt_angular.npz", v) for k, v in PT_GOLDEN_ANGULAR.items()})
ALL_GOLDEN.update({k: ("pt_ao.npz", v) for k, v in PT_GOLDEN_AO.items()})
ALL_GOLDEN.update({k: ("pt_texback.npz", v) for k, v in PT_GOLDEN_TEXBACK.items()})
ALL_GOLDEN.update({k: ("pt_glass_absorb.npz", v) for k, v in PT_GOLDEN_GLASS_ABSORB.items()})
ALL_GOLDEN.update({k: ("pt_sunsky.npz", v) for k, v in PT_GOLDEN_SUNSKY.items()})
ALL_GOLDEN.update({k: ("pt_darksky.npz", v) for k, v in PT_GOLDEN_DARKSKY.items()})
ALL_GOLDEN.update({k: ("pt_ibl.npz", v) for k, v in PT_GOLDEN_IBL.items() if k != "ibl_path_with_lamp"})
PSNR_MIN = 50.0


def test_accelerator_plugin_from_8_host_threads_vs_reference_kdtree_golden():
    """Accelerator::factory("yafaray-b200-bvh8") inside libYafaRay; intersect / isShadowed called concurrently from 8 host
    threads (the render workers of integrator_tiled.cc:248-251) against the committed kd-tree hit records."""
    d = np.load(os.path.join(GOLD, "rays_sphere.npz"))
    tris = scenes.displaced_sphere(int(d["scene"][0]), int(d["scene"][1]))
    for batch in ("primary", "diffuse"):
        got, info = helpers.b200_raybatch(tris, d[batch], "closest", threads=8)
        assert info["accel"] == "yafaray-b200-bvh8" and info["threads"] == 8
        r = helpers.compare_closest(got, d["hit_" + batch])
        assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0 and r["bit_exact_t"] == r["n_hit_both"], r
    got, _ = helpers.b200_raybatch(tris, d["shadow"], "shadow", threads=8)
    assert np.array_equal(got["prim"] != -1, d["hit_shadow"]["prim"] != -1)
    # the tree bound the Scene takes from the accelerator (Accelerator::getBound) carries the reference's 0.1 % inflation
    lo, hi = tris.reshape(-1, 3).min(0), tris.reshape(-1, 3).max(0)
    foo = (hi - lo).astype(np.float64) * 0.001
    exp = np.concatenate([(lo.astype(np.float64) - foo).astype(np.float32), (hi.astype(np.float64) + foo).astype(np.float32)])
    assert np.allclose(np.array(info["bound"], np.float32), exp, rtol=1e-6, atol=0)


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
def test_reference_cpu_path_integrator_over_the_gpu_accelerator_equals_the_kdtree_render():
    """The reference's own PathIntegrator (CPU) with scene_accelerator = yafaray-b200-bvh8: every intersect / isShadowed of
    the render goes through AcceleratorB200's single-ray virtuals.  Same hits => same image as with the kd-tree."""
    kw = dict(width=48, height=27, spp=2, bounces=4)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, accelerator="yafaray-b200-bvh8", **kw)
    img, _ = helpers.b200_render(rec, gpu_integrator=False)
    assert helpers.psnr(img, ref) >= 80.0
    assert (np.abs(img - ref).max(-1) > 1e-6).mean() <= 0.002          # near-tie pixels only


@pytest.mark.parametrize("key", sorted(ALL_GOLDEN))
def test_path_integrator_plugin_through_the_reference_c_api_vs_golden(key):
    """yafaray_* calls -> patched libYafaRay -> AcceleratorB200 + PathIntegratorB200 -> liby4gpu.so; image taken from the
    reference's own output callback after ImageFilm::flush."""
    ref = np.load(os.path.join(GOLD, ALL_GOLDEN[key][0]))[key]
    rec = helpers.ScriptRecorder()
    scenes_pt.build(rec, accelerator="yafaray-b200-bvh8", **ALL_GOLDEN[key][1])
    img, info = helpers.b200_render(rec, cwd=TEST01_DIR)
    assert img.shape == ref.shape
    assert helpers.psnr(img, ref) >= PSNR_MIN, helpers.psnr(img, ref)
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max()))


def test_adaptive_passes_run_by_the_reference_pass_loop():
    """AA_passes > 1 with a threshold: TiledIntegrator::render keeps the pass loop and ImageFilm::nextPass flags the pixels on
    the host from the film the device handed over; the result equals the product's own device-side adaptive loop."""
    from libyafaray_b200 import interface
    kw = dict(width=96, height=54, spp=4, bounces=4)
    extra = dict(AA_passes=3, AA_inc_samples=2, AA_threshold=0.02)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, accelerator="yafaray-b200-bvh8", **kw)
    helpers.with_render_params(rec, **extra)
    img, _ = helpers.b200_render(rec)
    yi = interface.Interface()
    scenes_pt.cornell_box(yi, accelerator="yafaray-b200-bvh8", **kw)
    yi.render_params.update(extra)
    own = yi.render()
    assert sum(yi.resampled_per_pass) > 0
    assert helpers.psnr(img, own) >= 80.0


def test_unsupported_scene_fails_the_render_instead_of_falling_back():
    """bounces = 12 is outside the supported set: PathIntegratorB200::preprocess fails, yafaray_render produces no image
    (there is no CPU fallback to the reference's integrator)."""
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, accelerator="yafaray-b200-bvh8", width=32, height=18, spp=1, bounces=12)
    try:
        img, info = helpers.b200_render(rec)
    except RuntimeError:
        return
    assert info["width"] == 0 and img.size == 0
