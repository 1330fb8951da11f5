"""The reference's film file (YAF_FILMv4_0_0): our writer / reader / folder merge (libyafaray_b200/film_io.py) against the
UNMODIFIED reference in both directions -- the reference loads a film we wrote and we load a film it wrote."""
import os

import numpy as np
import pytest

import helpers
from libyafaray_b200 import film_io, interface, scenes_pt

KW = dict(width=48, height=27, spp=3, bounces=2)


def _emu_film(**kw):
    yi = interface.Interface()
    scenes_pt.build(yi, accelerator="yafaray-b200-bvh8", **kw)
    return helpers.emu_pt_render(yi.assemble())


def test_round_trip_header_checks_and_folder_merge(tmp_path):
    rng = np.random.default_rng(1)
    w, c = rng.random((5, 7), np.float32), rng.random((5, 7, 4), np.float32)
    base = str(tmp_path / "out.tga")
    assert film_io.film_path(base, 3).endswith("out.tga - node 0003.film")
    film_io.save_film(film_io.film_path(base, 0), w, c, 0, 11, 40, cx0=2, cy0=1)
    film_io.save_film(film_io.film_path(base, 1), 2 * w, 3 * c, 1, 5, 64, cx0=2, cy0=1)
    film_io.save_film(str(tmp_path / "other - node 0000.film"), w, c)                      # different base name: not picked up
    film_io.save_film(film_io.film_path(base, 2), w[:4], c[:4], 2, 0, 999, cx0=2, cy0=1)    # wrong height: skipped
    f = film_io.load_film(film_io.film_path(base, 1), 7, 5, 2, 1)
    assert f["computer_node"] == 1 and f["base_sampling_offset"] == 5 and f["sampling_offset"] == 64
    assert np.array_equal(f["weights"], 2 * w) and np.array_equal(f["layers"][0], 3 * c)
    for bad in (dict(width=8), dict(height=4), dict(cx0=0), dict(cy0=0), dict(num_layers=3)):
        with pytest.raises(film_io.FilmFileError, match="check failed"):
            film_io.load_film(film_io.film_path(base, 1), **{**dict(width=7, height=5, cx0=2, cy0=1), **bad})
    (tmp_path / "junk.film").write_bytes(b"not a film")
    with pytest.raises(film_io.FilmFileError, match="valid"):
        film_io.load_film(str(tmp_path / "junk.film"), 7, 5)
    sw, sc, base_offs, offs, n = film_io.load_all_in_folder(base, 7, 5, 2, 1, base_sampling_offset=7)
    assert n == 2 and base_offs == 11 and offs == 64
    assert np.array_equal(sw, w + 2 * w) and np.array_equal(sc, c + 3 * c)


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
def test_film_written_by_the_reference_loads_here(tmp_path):
    """film_load_save_mode=save in the reference -> our reader: header fields and the accumulated (unnormalised) film."""
    path = str(tmp_path / "cornell.tga")
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, render_extra=dict(film_load_save_mode="save", film_load_save_path=path, adv_computer_node=2, adv_base_sampling_offset=17), **KW)
    helpers.ref_render(rec)
    f = film_io.load_film(film_io.film_path(path, 2), KW["width"], KW["height"])
    assert (f["computer_node"], f["base_sampling_offset"], f["sampling_offset"]) == (2, 17, KW["spp"])
    w, rgba = _emu_film(render_extra=dict(adv_computer_node=2, adv_base_sampling_offset=17), **KW)
    assert np.allclose(f["weights"], w, rtol=0, atol=1e-5)
    assert np.allclose(f["layers"][0], rgba, rtol=1e-5, atol=1e-5)
    # the internal sampling-factor layer: alpha is Combined's, rgb counts the samples that hit a surface (<= the weights)
    assert f["layers"].shape[0] == 2 and np.array_equal(f["layers"][1][..., 3], f["layers"][0][..., 3])
    assert (f["layers"][1][..., 0] <= f["weights"] + 1e-6).all()


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
def test_film_written_here_resumes_in_the_reference(tmp_path):
    """Two "nodes" render here with different sample ranges and save their films; the reference started with load-save merges
    both (imageFilmLoadAllInFolder), skips its pass 0 and outputs the combined image."""
    path = str(tmp_path / "cornell.tga")
    films = []
    for node in (0, 1):
        w, rgba = _emu_film(render_extra=dict(adv_computer_node=node), **KW)
        film_io.save_film(film_io.film_path(path, node), w, rgba, node, 0, KW["spp"])
        films.append((w, rgba))
    assert not np.array_equal(films[0][1], films[1][1])           # node 1 drew from its own 100000-sample range
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, render_extra=dict(film_load_save_mode="load-save", film_load_save_path=path), **KW)
    ref, _ = helpers.ref_render(rec)
    want = interface.normalize_film(films[0][0] + films[1][0], films[0][1] + films[1][1])
    assert helpers.psnr(want, ref) >= 100.0, helpers.psnr(want, ref)
    merged = film_io.load_all_in_folder(path, KW["width"], KW["height"])
    assert merged[4] >= 2 and os.path.exists(film_io.film_path(path, 0))
