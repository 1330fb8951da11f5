"""GPU parity: CUDA BVH8 accelerator (through the C-ABI) vs the reference kd-tree."""
import numpy as np
import pytest

import helpers
from libyafaray_b200 import scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sphere():
    from libyafaray_b200.accelerator import AcceleratorB200
    tris = scenes.displaced_sphere(200, 100)
    return tris, AcceleratorB200(tris)


def _batches(tris, ref_fn, n=256):
    prim = scenes.primary_rays(n, n)
    h = ref_fn(tris, prim, "closest")[0]
    p, nrm, ok = scenes.hit_points_normals(tris, prim, h["t"], h["prim"])
    return prim, h, scenes.shadow_rays(p, ok), scenes.diffuse_rays(p, nrm, ok)


@pytest.mark.skipif(not helpers.have_ref(), reason="oracle/_ref not built")
def test_sphere_vs_reference_kdtree(sphere):
    tris, acc = sphere
    prim, h_ref, sh, df = _batches(tris, helpers.ref_raybatch)
    got = acc.intersect(prim)
    r = helpers.compare_closest(got, h_ref)
    assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0, r
    got_d = acc.intersect(df)
    ref_d = helpers.ref_raybatch(tris, df, "closest")[0]
    r = helpers.compare_closest(got_d, ref_d)
    assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0, r
    got_s = acc.isShadowed(sh)
    ref_s = helpers.ref_raybatch(tris, sh, "shadow")[0]
    assert np.array_equal(got_s["prim"] >= 0, ref_s["prim"] != -1)


def test_sphere_vs_emulation_bit_exact(sphere):
    tris, acc = sphere
    prim = scenes.primary_rays(128, 128)
    got = acc.intersect(prim)
    exp, _, _ = helpers.emu_build_trace(tris, prim, 0)
    assert np.array_equal(got["prim"], exp["prim"])
    assert np.array_equal(got["t"], exp["t"]) and np.array_equal(got["bu"], exp["bu"]) and np.array_equal(got["bv"], exp["bv"])


def test_bound_matches_reference_inflation(sphere):
    tris, acc = sphere
    lo, hi = tris.reshape(-1, 3).min(0), tris.reshape(-1, 3).max(0)
    foo = (hi - lo).astype(np.float64) * 0.001
    exp = np.concatenate([(lo.astype(np.float64) - foo).astype(np.float32), (hi.astype(np.float64) + foo).astype(np.float32)])
    assert np.array_equal(acc.getBound(), exp)


def test_edge_cases():
    from libyafaray_b200.accelerator import AcceleratorB200
    rays = scenes.primary_rays(16, 16)
    empty = AcceleratorB200(np.zeros((0, 9), np.float32))
    assert (empty.intersect(rays)["prim"] == -1).all()
    one = np.array([[-1, -1, 0, 1, -1, 0, 0, 1, 0]], np.float32)
    for n in (1, 2, 3, 7, 9):
        tris = np.repeat(one, n, 0) + np.arange(n, dtype=np.float32)[:, None] * np.array([0, 0, -0.1] * 3, np.float32)
        acc = AcceleratorB200(tris)
        got = acc.intersect(rays)
        exp = helpers.emu_brute(tris, rays, 0)
        assert np.array_equal(got["prim"], exp["prim"]) and np.array_equal(got["t"], exp["t"])
        assert (acc.intersect(rays[:0])["prim"].shape[0]) == 0
