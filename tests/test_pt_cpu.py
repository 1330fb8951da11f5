"""CPU-side path-tracer checks (no GPU): the wavefront stage logic (the same Y4_HD source the CUDA
kernels compile, run serially by tests/emu) against the committed golden renders of the UNMODIFIED
reference PathIntegrator, the host-side C-API mirror, and the sampler KATs."""
import json
import os
import sys

import numpy as np
import pytest

import helpers
from libyafaray_b200 import interface, scenes_pt
from libyafaray_b200._lib import Y4Error

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLD)
from make_golden import PT_GOLDEN, PT_GOLDEN_CITY, PT_GOLDEN_DOF, PT_GOLDEN_LIGHTS, PT_GOLDEN_MIRROR, PT_GOLDEN_TEST01, TEST01_DIR  # noqa: E402

ALL_GOLDEN = {**{k: ("pt_cornell.npz", v) for k, v in PT_GOLDEN.items()}, **{k: ("pt_city.npz", v) for k, v in PT_GOLDEN_CITY.items()},
              **{k: ("pt_dof.npz", v) for k, v in PT_GOLDEN_DOF.items()}, **{k: ("pt_lights.npz", v) for k, v in PT_GOLDEN_LIGHTS.items()},
              **{k: ("pt_test01.npz", v) for k, v in PT_GOLDEN_TEST01.items()}, **{k: ("pt_mirror.npz", v) for k, v in PT_GOLDEN_MIRROR.items()}}

PSNR_MIN = 50.0  # BASELINE.json: PSNR >= 50 dB vs the reference at equal spp with identical sample sequences


def _assemble(**kw):
    yi = interface.Interface()
    yi.image_dir = TEST01_DIR
    scenes_pt.build(yi, accelerator="yafaray-b200-bvh8", **kw)
    return yi.assemble()


@pytest.mark.parametrize("key", sorted(ALL_GOLDEN))
def test_stages_vs_golden_reference_render(key):
    """Cornell configurations and the instanced mini-city (C4: yafaray_addInstance, sky background)."""
    ref = np.load(os.path.join(GOLD, ALL_GOLDEN[key][0]))[key]
    # with several lights the path vertices pick one light each from the reference's single-thread call counter:
    # replayed here (helpers.emu_pt_render) so that the six-light scene is compared sample for sample
    a = _assemble(**ALL_GOLDEN[key][1])
    w, rgba = helpers.emu_pt_render(a, ref_light_order=a["n_light"] > 1 and a["params"].path_samples > 0)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    # sample-exact: what differs is only the fp32 order of the film sums
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()
    assert np.abs(img[..., 3] - ref[..., 3]).max() <= 1e-6


def _glass_cfgs():
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_GLASS
    return PT_GOLDEN_GLASS


@pytest.mark.parametrize("key", ["glass_sphere", "glass_box_tinted", "glass_and_mirror", "glass_direct_point", "glass_no_caustics_gauss", "glass_deep_tree"])
def test_glass_ray_tree_vs_golden_reference_render(key):
    """`glass` (reflection AND refraction at a vertex: recursiveRaytrace is a tree) against the reference renders of
    tests/golden/make_golden_large.py.  The tree is walked depth first and its branches are summed in walk order, the reference
    sums from the leaves up: a few ulp per level.  (Glass boxes are lifted off the floor: a bottom face coinciding with the floor
    gives exact ties in t, the documented class in which the kd-tree's leaf order decides.)"""
    ref = np.load(os.path.join(GOLD, "pt_glass.npz"))[key]
    a = _assemble(**_glass_cfgs()[key])
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


def _ibl_cfgs():
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_IBL
    return PT_GOLDEN_IBL


@pytest.mark.parametrize("key", ["ibl_constant", "ibl_gradient", "ibl_gradient_mirror_gauss", "ibl_direct_with_lamp", "ibl_no_shadows", "ibl_path_with_lamp"])
def test_background_light_vs_golden_reference_render(key):
    """A background with ibl=True: the `bglight` it creates (light_background.cc) tabulates the background on 360 latitude rows
    and importance-samples it; sampled directions, pdfs and the MIS partner (BackgroundLight::intersect) are sample-exact."""
    ref = np.load(os.path.join(GOLD, "pt_ibl.npz"))[key]
    a = _assemble(**_ibl_cfgs()[key])
    w, rgba = helpers.emu_pt_render(a, ref_light_order=a["n_light"] > 1 and a["params"].path_samples > 0)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
def test_stages_vs_live_reference_multi_pass_and_rr():
    # multi-pass AA switches the pixel jitter to riVdC/riS; Russian roulette is statistical only (the reference draws
    # from a rand()-seeded generator), so it is compared at a loose bound
    kw = dict(width=64, height=36, spp=2, bounces=4)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    rec.lines = [l.replace("AA_passes 1", "AA_passes 2") for l in rec.lines] + ["paramsSetFloat AA_threshold 0.0"]
    # setupRender is already recorded; re-issue the render params with the extra keys
    i = max(k for k, l in enumerate(rec.lines) if l == "setupRender")
    rec.lines.pop(i)
    rec.lines.append("setupRender")
    ref, _ = helpers.ref_render(rec)
    yi = interface.Interface()
    scenes_pt.cornell_box(yi, accelerator="yafaray-b200-bvh8", **kw)
    yi.render_params["AA_passes"] = 2
    yi.render_params["AA_threshold"] = 0.0
    a = yi.assemble()
    w, rgba = helpers.emu_pt_render(a, n_samples=2, offset=0, aa_pass=0)
    w2, rgba2 = helpers.emu_pt_render(a, n_samples=1, offset=2, aa_pass=1)
    img = interface.normalize_film(w + w2, rgba + rgba2)
    assert helpers.psnr(img, ref) >= PSNR_MIN


ADAPTIVE = [
    dict(AA_passes=3, AA_inc_samples=2, AA_threshold=0.03),
    dict(AA_passes=4, AA_inc_samples=1, AA_threshold=0.02, AA_resampled_floor=60.0, AA_detect_color_noise=True, AA_variance_pixels=4, AA_variance_edge_size=6,
         AA_dark_detection_type="linear", AA_dark_threshold_factor=0.5, AA_sample_multiplier_factor=1.5),
    dict(AA_passes=4, AA_inc_samples=2, AA_threshold=0.02, AA_dark_detection_type="curve"),      # ImageFilm::darkThresholdCurveInterpolate
]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("aa", ADAPTIVE)
def test_adaptive_aa_vs_live_reference(aa):
    """ImageFilm::nextPass noise detection + the adaptive pass loop: same pixels flagged, same image as the reference."""
    kw = dict(width=64, height=36, spp=2, bounces=3)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(helpers.with_render_params(rec, **aa))
    yi = interface.Interface()
    scenes_pt.cornell_box(yi, accelerator="yafaray-b200-bvh8", **kw)
    yi.render_params.update(aa)
    img, hist = helpers.emu_adaptive_render(yi)
    assert any(0 < h < kw["width"] * kw["height"] for h in hist), hist      # the passes really were adaptive
    assert helpers.psnr(img, ref) >= PSNR_MIN, (helpers.psnr(img, ref), hist)


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
def test_transparent_shadow_mode_vs_live_reference():
    """integrator transpShad=true: shadow rays take the intersectTs ray range / visibility rules (all supported materials are opaque)."""
    kw = dict(width=64, height=36, spp=4, bounces=3)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    i = [k for k, l in enumerate(rec.lines) if l.startswith("createIntegrator")][0]
    rec.lines[i:i] = ["paramsSetBool transpShad 1", "paramsSetInt shadowDepth 3"]
    ref, _ = helpers.ref_render(rec)
    yi = interface.Interface()
    scenes_pt.cornell_box(yi, accelerator="yafaray-b200-bvh8", **kw)
    yi.integrators["pt"]["transpShad"] = True
    w, rgba = helpers.emu_pt_render(yi.assemble())
    assert helpers.psnr(interface.normalize_film(w, rgba), ref) >= 100.0


def test_interface_rejects_unsupported_features():
    yi = interface.Interface()
    yi.paramsSetString("type", "glossy")
    with pytest.raises(Y4Error):
        yi.createMaterial("g")
    yi.paramsClearAll()
    yi.paramsSetString("type", "glass")
    yi.paramsSetFloat("dispersion_power", 0.1)          # chromatic dispersion: a spectral ray per wavelength
    with pytest.raises(Y4Error):
        yi.createMaterial("g2")
    yi.paramsClearAll()
    yi.paramsSetString("type", "shinydiffusemat")
    yi.paramsSetFloat("transparency", 0.5)
    yi.paramsSetFloat("transparentbias_factor", 0.1)   # shifted refraction origins are not restated
    with pytest.raises(Y4Error):
        yi.createMaterial("t")
    yi.paramsClearAll()
    yi.paramsSetString("type", "fisheye")              # not a camera type of the reference (Camera::factory returns nullptr)
    with pytest.raises(Y4Error):
        yi.createCamera("c")
    yi.paramsClearAll()
    yi.paramsSetString("type", "bglight")
    with pytest.raises(Y4Error):
        yi.createLight("l")
    yi.paramsClearAll()
    yi.paramsSetString("type", "ieslight")
    with pytest.raises(Y4Error):
        yi.createLight("l")
    yi.paramsClearAll()
    yi.paramsSetString("type", "photonmapping")
    with pytest.raises(Y4Error):
        yi.createIntegrator("i")
    yi.paramsClearAll()
    yi.paramsSetString("scene_accelerator", "yafaray-kdtree-original")
    with pytest.raises(Y4Error):
        yi.setupRender()


def test_round_two_features_keep_refusing_what_they_do_not_restate():
    """The corners of the newer features that are NOT restated: refused with Y4Error, never approximated."""
    yi = interface.Interface()
    yi.image_dir = TEST01_DIR
    for bad in (dict(type="textureback", texture="nowhere"),                       # TextureBackground::factory: texture must exist
                dict(type="sunsky", add_sun=True), dict(type="darksky", add_sun=True),   # sunlight colour from the solar spectrum tables
                dict(type="clouds_sky")):
        yi.paramsClearAll()
        scenes_pt._set_params(yi, bad)
        with pytest.raises(Y4Error):
            yi.createBackground("bg")
    yi.paramsClearAll()
    yi.paramsSetString("filename", "tex.hdr")
    assert yi.createImage("img")
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="image", image_name="img"))
    assert yi.createTexture("sky")
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="textureback", texture="sky", smartibl_blur=0.3))     # mip-mapped IBL lookups
    with pytest.raises(Y4Error):
        yi.createBackground("bg")
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="textureback", texture="sky", ibl=True))
    assert yi.createBackground("bg") and "textureBackground_bgLight" in yi.lights
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="darksky", background_light=True, light_samples=5))
    assert yi.createBackground("sky2") and yi.lights["DarkSky_bgLight"]["samples"] == 5
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="directlighting", do_AO=True, AO_samples=0))          # the reference divides by zero
    with pytest.raises(Y4Error):
        yi.createIntegrator("i")
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="pathtracing", do_AO=True))                           # parsed and ignored by the reference's path tracer
    assert yi.createIntegrator("i")
    yi.paramsClearAll()
    yi.paramsSetString("type", "glass")
    yi.paramsSetColor("absorption", 0.5, 0.0, 0.5, 1.0)                                       # -log(0): the reference falls back to log(1e38)
    with pytest.raises(Y4Error):
        yi.createMaterial("g")
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="angular", projection="no_such_projection"))          # unknown names select equidistant, like the factory
    assert yi.createCamera("c")


def test_texture_features_outside_the_supported_set_are_rejected_or_dropped_like_the_reference():
    yi = interface.Interface()
    yi.image_dir = TEST01_DIR
    yi.paramsSetString("filename", "tex.tga")
    assert yi.createImage("img")
    yi.paramsClearAll()
    yi.paramsSetString("filename", "tex.png")           # no PNG loader in a default reference build: no image, no exception
    assert not yi.createImage("png_img") and yi.warnings
    for bad in (dict(type="clouds"), dict(type="image", image_name="img", interpolate="bicubic"), dict(type="image", image_name="img", adj_hue=30.0),
                dict(type="image", image_name="img", normalmap=True)):
        yi.paramsClearAll()
        scenes_pt._set_params(yi, bad)
        with pytest.raises(Y4Error):
            yi.createTexture("t")
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="image", image_name="png_img"))
    assert not yi.createTexture("dropped")              # ImageTexture::factory returns nullptr
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="image", image_name="img"))
    assert yi.createTexture("tex")

    def material(layer=None, mapper=None, **mat):
        yi.paramsClearAll()
        scenes_pt._set_params(yi, dict(type="shinydiffusemat", diffuse_shader="l0", **mat))
        yi.paramsPushList()
        scenes_pt._set_params(yi, {**dict(element="shader_node", type="layer", name="l0", input="m0"), **(layer or {})})
        yi.paramsPushList()
        scenes_pt._set_params(yi, {**dict(element="shader_node", type="texture_mapper", name="m0", texture="tex"), **(mapper or {})})
        yi.paramsEndList()
        return yi.createMaterial("m")

    assert material() and len(yi.materials["m"]["_diffuse_chain"]) == 1
    for kw in (dict(mapper=dict(mapping="sphere")), dict(mapper=dict(texco="window")), dict(layer=dict(color_input=False)),
               dict(bump_shader="l0"), dict(mapper=dict(type="value"))):
        with pytest.raises(Y4Error):
            material(**kw)
    assert material(mapper=dict(texture="dropped")) and yi.materials["m"]["_diffuse_chain"] is None      # nodes dropped, plain colour
    assert material(layer=dict(input="nowhere")) and yi.materials["m"]["_diffuse_chain"] is None


def test_image_decoders_tga_hdr_storage_rounding():
    """TGA: sRGB -> linear with the FAST_MATH pow, then the 10/10/10/8-bit "optimized" storage; HDR: RGBE as float, no rounding."""
    from libyafaray_b200 import images

    tga = images.load_image(dict(filename="tex.tga", color_space="sRGB", image_optimization="optimized"), TEST01_DIR)
    hdr = images.load_image(dict(filename="tex.hdr", color_space="sRGB", image_optimization="optimized"), TEST01_DIR)
    assert tga.shape == (300, 300, 4) and hdr.shape == (300, 300, 4) and tga.dtype == np.float32
    q = tga[..., :3] * np.float32(1023.0)
    assert np.abs(q - np.round(q)).max() < 1e-3 and np.abs(tga[..., 3] * 255 - np.round(tga[..., 3] * 255)).max() < 1e-4
    assert 0.0 <= tga.min() and tga.max() <= 1.0 and len(np.unique(tga[..., 3])) > 2          # the file has a soft alpha channel
    assert (hdr[..., 3] == 1.0).all() and hdr[..., :3].max() > 0.5
    # exact properties of the restated pow: sRGB 1.0 -> 1023/1023 after rounding; values below the knee are divided by 12.92
    assert images.linear_from_srgb(np.float32([0.04]))[0] == np.float32(0.04) / np.float32(12.92)
    assert abs(float(images.linear_from_srgb(np.float32([1.0]))[0]) - 1.0) < 2e-4
    lin = images.to_linear(np.float32([[0.5, 0.25, 1.0, 0.3]]), "sRGB", 1.0)
    assert lin[0, 3] == np.float32(0.3) and abs(float(lin[0, 0]) - 0.21404) < 2e-4                # alpha is not converted
    with pytest.raises(images.ImageLoadError):
        images.load_image(dict(filename="tex.exr"), TEST01_DIR)


BOKEH = [dict(bokeh_type="hexagon", bokeh_rotation=20.0), dict(bokeh_type="triangle", bokeh_bias="center"), dict(bokeh_type="square", bokeh_bias="edge"),
         dict(bokeh_type="pentagon", bokeh_rotation=-35.0), dict(bokeh_type="disk2", bokeh_bias="center"), dict(bokeh_type="disk2"), dict(bokeh_type="ring")]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("bokeh", BOKEH)
def test_lens_shapes_vs_live_reference(bokeh):
    """PerspectiveCamera::getLensUv: polygonal lenses (corner table from the rotation), disk2 / ring, the three bias rules."""
    kw = dict(width=48, height=27, spp=4, bounces=2, aperture=0.08, dof_distance=1.9, bokeh=bokeh)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    w, rgba = helpers.emu_pt_render(_assemble(**kw))
    img = interface.normalize_film(w, rgba)
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max()))


CAMERAS = [dict(type="orthographic", scale=1.3), dict(type="architect", focal=1.1), dict(type="architect", aperture=0.05, dof_distance=1.8, bokeh_type="square"),
           dict(type="equirectangular", **{"from": (0.5, 0.5, 0.3), "to": (0.5, 0.45, 1.0), "up": (0.5, 1.5, 0.3)}),
           dict(type="angular", angle=75.0, max_angle=50.0, projection="equisolid_angle", mirrored=True, **{"from": (0.5, 0.5, 0.3), "to": (0.45, 0.5, 1.0), "up": (0.5, 1.5, 0.3)}),
           dict(type="angular", angle=160.0, circular=False, **{"from": (0.5, 0.5, 0.3), "to": (0.5, 0.45, 1.0), "up": (0.5, 1.5, 0.3)}),
           dict(type="angular", angle=40.0, projection="orthographic", aspect_ratio=1.3, **{"from": (0.5, 0.5, -0.8), "to": (0.5, 0.5, 1.0), "up": (0.5, 1.5, -0.8)})]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("cam", CAMERAS)
def test_camera_types_vs_live_reference(cam):
    """OrthographicCamera / ArchitectCamera / EquirectangularCamera::shootRay next to the perspective camera."""
    kw = dict(width=64, height=36, spp=3, bounces=2, camera_extra=cam)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    w, rgba = helpers.emu_pt_render(_assemble(**kw))
    img = interface.normalize_film(w, rgba)
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("angle", [180.0, 60.0, 25.0])
def test_smooth_mesh_vs_live_reference(angle):
    """yafaray_smoothMesh: MeshObject::smoothNormals (one normal per vertex at 180 degrees, angle-limited blending below) next to
    flat-shaded meshes in the same scene."""
    kw = dict(width=64, height=36, spp=3, bounces=3, smooth_sphere=angle)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    a = _assemble(**kw)
    assert a["normals"] is not None and np.isnan(a["normals"][0, 0]) and not np.isnan(a["normals"][-1]).any()
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)


_M = dict(objects=("tall", "back"), color=(0.9, 0.95, 0.8))
MIRRORS = [dict(mirror=dict(objects=("tall",), color=(0.9, 0.8, 0.7), reflect=0.95)),                                  # integrator raydepth 2, caustic_type none
           dict(mirror=_M, glow=0.7, integrator_extra=dict(raydepth=3, caustic_type="none")),
           dict(mirror=_M, extra_lights=("sphere", "spot"), integrator_extra=dict(raydepth=5, caustic_type="path")),    # light choice down mirrored paths
           dict(mirror=_M, background=(0.3, 0.2, 0.1), integrator_extra=dict(raydepth=2, bg_transp=True)),              # background opaque behind a mirror
           dict(mirror=_M, integrator_extra=dict(raydepth=0)),                                                          # no recursion level left: black mirror
           dict(mirror=_M, integrator="directlighting")]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("variant", MIRRORS)
def test_mirror_material_vs_live_reference(variant):
    """MirrorMaterial: the ray-level loop (recursiveRaytrace -> specularReflect -> integrate, up to raydepth), mirror vertices
    inside sub-paths (MirrorMaterial::sample, the caustic flag, mat_bsd_fs of the FIRST path vertex deciding light estimates and
    emission for the whole sub-path), TiledIntegrator::background at ray_level > 0.  Sample-exact: a mirror chain of k levels
    multiplies k colours in the opposite order to the reference's unwinding (<= 1 ulp per level)."""
    kw = dict(width=64, height=36, spp=3, bounces=4, **variant)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    a = _assemble(**kw)
    w, rgba = helpers.emu_pt_render(a, ref_light_order=a["n_light"] > 1 and a["params"].path_samples > 0)
    img = interface.normalize_film(w, rgba)
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)
    assert np.abs(img[..., 3] - ref[..., 3]).max() <= 1e-6


SHINY = [dict(shiny=dict(specular_reflect=0.3), integrator_extra=dict(raydepth=0)),                                            # component sampled in paths, no recursion
         dict(shiny=dict(specular_reflect=0.3), integrator_extra=dict(raydepth=1)),
         dict(shiny=dict(specular_reflect=0.4), mirror=dict(objects=("tall",)), integrator="directlighting", background=(0.3, 0.2, 0.1),
              integrator_extra=dict(raydepth=4, bg_transp=True)),
         dict(shiny=dict(specular_reflect=0.5, fresnel_effect=True), extra_lights=("sphere", "spot"), integrator_extra=dict(raydepth=2, caustic_type="path"))]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("variant", SHINY)
def test_shinydiffuse_mirror_component_vs_live_reference(variant):
    """shinydiffusemat specular_reflect / mirror_color / fresnel_effect / IOR: ShinyDiffuseMaterial::config, accumulate,
    getFresnelKr, the two-component sample / eval / pdf, getSpecular (grazing-angle fix) and the ray-level recursion in which
    every level adds its own estimate (emission + lights + sub-paths) before the reflected ray is followed."""
    kw = dict(width=64, height=36, spp=3, bounces=4, **variant)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    a = _assemble(**kw)
    w, rgba = helpers.emu_pt_render(a, ref_light_order=a["n_light"] > 1 and a["params"].path_samples > 0)
    img = interface.normalize_film(w, rgba)
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)
    assert np.abs(img[..., 3] - ref[..., 3]).max() <= 1e-6


PANES = [dict(pane=dict(transparency=0.6), integrator_extra=dict(raydepth=0)),                              # lobe sampled in paths only
         dict(pane=dict(transparency=0.6), integrator_extra=dict(raydepth=2)),
         dict(pane=dict(transparency=1.0, transmit_filter=1.0, color=(0.9, 0.3, 0.2)), mirror=dict(objects=("tall",)), path_samples=2, light_samples=2,
              integrator_extra=dict(raydepth=4, caustic_type="path")),
         dict(pane=dict(transparency=0.7, transmit_filter=0.2), extra_lights=("sphere", "spot"), integrator_extra=dict(raydepth=2, caustic_type="path")),
         # transpShad: isShadowed(ray, shadowDepth, ...) multiplies the light by the pane's getTransparency colour; depth limits 1 and 0
         dict(pane=dict(transparency=0.6, transmit_filter=0.5), integrator_extra=dict(raydepth=2, transpShad=True)),
         dict(pane=dict(transparency=0.7), point_light=True, integrator="directlighting", integrator_extra=dict(raydepth=3, transpShad=True, shadowDepth=1)),
         dict(pane=dict(transparency=0.7, transmit_filter=0.3), extra_lights=("sphere", "spot", "sun"), integrator_extra=dict(raydepth=2, transpShad=True, shadowDepth=0))]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("variant", PANES)
def test_shinydiffuse_transparency_vs_live_reference(variant):
    """shinydiffusemat transparency / transmit_filter: config / accumulate with components {0 | mirror, transparency, 0, diffuse},
    the Transmit | Filter lobe of sample(), getAlpha in the sample weight, eval's (1 - transparency) factor, pdf's sum over
    both components, getSpecular's refract_ (straight on) followed level by level through specularRefract."""
    kw = dict(width=64, height=36, spp=3, bounces=4, **variant)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    a = _assemble(**kw)
    w, rgba = helpers.emu_pt_render(a, ref_light_order=a["n_light"] > 1 and a["params"].path_samples > 0)
    img = interface.normalize_film(w, rgba)
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)
    assert np.abs(img[..., 3] - ref[..., 3]).max() <= 1e-6


TRANSLUCENT = [dict(pane=dict(translucency=0.5, diffuse_reflect=0.8, color=(0.9, 0.8, 0.3)), light_samples=2, path_samples=2),
               dict(pane=dict(translucency=0.9, diffuse_reflect=0.0), extra_lights=("sphere", "spot"), point_light=True),      # no Lambert component at all
               dict(pane=dict(translucency=0.4, specular_reflect=0.3, fresnel_effect=True), integrator="directlighting", integrator_extra=dict(raydepth=2))]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("variant", TRANSLUCENT)
def test_shinydiffuse_translucency_vs_live_reference(variant):
    """shinydiffusemat translucency: the Diffuse | Transmit component of config (with its transparency-strength slip) /
    accumulate / sample (cosine lobe about -n) / eval (light from the far side) / pdf, in the camera-hit sample, the bounce
    loop and the MIS material sample."""
    kw = dict(width=64, height=36, spp=3, bounces=4, **variant)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    a = _assemble(**kw)
    w, rgba = helpers.emu_pt_render(a, ref_light_order=a["n_light"] > 1 and a["params"].path_samples > 0)
    assert helpers.psnr(interface.normalize_film(w, rgba), ref) >= 100.0


MESH_LIGHTS = [dict(mesh_light=dict(), light_samples=3, path_samples=2),
               dict(mesh_light=dict(shape="sphere", power=40.0, double_sided=True), integrator="directlighting", light_samples=4),
               dict(mesh_light=dict(shape="sphere", power=30.0), extra_lights=("spot",), light_samples=2)]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("variant", MESH_LIGHTS)
def test_object_light_vs_live_reference(variant):
    """objectlight / meshlight (light_object_light.cc): triangle areas, their Pdf1D cdf, dSample + TrianglePrimitive::sample in
    illumSample, and intersect -- including the reference never writing its `t` argument, which makes the MIS weight of the
    material sample distance-free and its shadow ray unbounded."""
    kw = dict(width=64, height=36, spp=3, bounces=3, **variant)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    a = _assemble(**kw)
    w, rgba = helpers.emu_pt_render(a, ref_light_order=a["n_light"] > 1 and a["params"].path_samples > 0)
    assert helpers.psnr(interface.normalize_film(w, rgba), ref) >= 100.0


def _spot_soft_cfgs():
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_SPOT_SOFT
    return PT_GOLDEN_SPOT_SOFT


@pytest.mark.parametrize("key", ["spot_soft_direct", "spot_soft_path_only_light", "spot_soft_direct_with_lamp"])
def test_soft_spot_light_vs_golden_reference_render(key):
    """spotlight with soft_shadows = true: a sampled light (cone-jittered shadow rays around the direction to the light,
    pdf = squared distance with the reference's `pdf >= 1` patch; SpotLight::illumSample / intersect)."""
    ref = np.load(os.path.join(GOLD, "pt_spot_soft.npz"))[key]
    a = _assemble(**_spot_soft_cfgs()[key])
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.01
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


@pytest.mark.parametrize("key", ["portal_gradient_path", "portal_constant_direct", "portal_direct_with_lamp_gauss"])
def test_background_portal_light_vs_golden_reference_render(key):
    """bgPortalLight: the triangles of a mesh emit the background seen through them (sampled like an object light, single sided,
    colour = Background::eval(direction) * power); the mesh itself becomes invisible."""
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_PORTAL
    ref = np.load(os.path.join(GOLD, "pt_portal.npz"))[key]
    a = _assemble(**PT_GOLDEN_PORTAL[key])
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.01
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


ANGULAR_KEYS = ["angular_equidistant_circular", "angular_stereographic_mirrored", "angular_orthographic_max_angle", "angular_equisolid_gauss",
                "angular_rectilinear_glass", "angular_direct_wide"]


@pytest.mark.parametrize("key", ANGULAR_KEYS)
def test_angular_camera_vs_golden_reference_render(key):
    """AngularCamera::shootRay: the five projections (atan2 / atan / asin with the C library's bits, csrc/y4_libm.cuh), mirrored,
    and the circular mask -- positions without a camera ray get the layer's default colour (black, alpha 1) with their full
    filter footprint; once through the specular ray tree."""
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_ANGULAR
    ref = np.load(os.path.join(GOLD, "pt_angular.npz"))[key]
    a = _assemble(**PT_GOLDEN_ANGULAR[key])
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.01
    if key != "angular_stereographic_mirrored":
        assert (ref[0, 0] == (0.0, 0.0, 0.0, 1.0)).all() and (img[0, 0] == (0.0, 0.0, 0.0, 1.0)).all()
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


AO_KEYS = ["ao_default_color", "ao_tinted_no_lamp_gauss", "ao_behind_mirror_and_glass", "ao_transparent_shadows_glow", "ao_ignored_by_the_path_tracer"]


@pytest.mark.parametrize("key", AO_KEYS)
def test_ambient_occlusion_of_direct_lighting_vs_golden_reference_render(key):
    """do_AO: TiledIntegrator::sampleAmbientOcclusion after estimateAllDirectLight at every diffuse vertex integrate() reaches --
    AO_samples BSDF-sampled occlusion rays of length AO_distance from Halton(2) / Halton(3), summed and divided by n; with
    transparent shadows the blockers' filter colours; an emitting diffuse wall adds emit * pdf per sample.  The path tracer
    parses the parameters and ignores them for the Combined layer (last key)."""
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_AO
    ref = np.load(os.path.join(GOLD, "pt_ao.npz"))[key]
    a = _assemble(**PT_GOLDEN_AO[key])
    assert (a["params"].ao_samples > 0) == (key != "ao_ignored_by_the_path_tracer")
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.01
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


TEXBACK_KEYS = ["texback_seen_spherical", "texback_seen_angular_nearest", "texback_ibl_path", "texback_ibl_clamp_direct", "texback_portal"]


@pytest.mark.parametrize("key", TEXBACK_KEYS)
def test_texture_background_vs_golden_reference_render(key):
    """textureback: an HDR image seen in every direction (Texture::sphereMap / angMap, rotation, power, the 1e-5 floor), as the
    IBL light's energy table and colour (with ibl_clamp_sampling on the BSDF-sampled MIS half) and through a portal."""
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_TEXBACK
    ref = np.load(os.path.join(GOLD, "pt_texback.npz"))[key]
    a = _assemble(**PT_GOLDEN_TEXBACK[key])
    assert a["params"].background_gradient == 2
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.01
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


@pytest.mark.parametrize("key", ["glass_absorb_sphere", "glass_absorb_boxes_default_dist", "glass_absorb_direct_gauss"])
def test_glass_absorption_vs_golden_reference_render(key):
    """glass with `absorption`: the Beer-Lambert handler of the inside (sigma_a = -log(colour) / absorption_dist) attenuates every
    specular ray that continues inside the glass by exp(-sigma_a * hit distance) (FAST_MATH exp), and blacks out one that leaves
    the scene; the branch weight of the ray tree carries the factor."""
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_GLASS_ABSORB
    ref = np.load(os.path.join(GOLD, "pt_glass_absorb.npz"))[key]
    a = _assemble(**PT_GOLDEN_GLASS_ABSORB[key])
    assert any(a["mats"][k].absorb for k in range(a["n_mat"]))
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.01
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


SUNSKY_KEYS = ["sunsky_seen_day", "sunsky_seen_night_variations", "sunsky_ibl_path", "sunsky_ibl_direct_with_sun_gauss", "sunsky_portal"]


@pytest.mark.parametrize("key", SUNSKY_KEYS)
def test_sunsky_background_vs_golden_reference_render(key):
    """sunsky: SunSkyBackground::eval per direction (float acos / atan2 of the C library, FAST_TRIG sin / cos, FAST_MATH exp, the
    rest in double), seen directly by day and by night, tabulated for its `background_light`, and through a portal."""
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_SUNSKY
    ref = np.load(os.path.join(GOLD, "pt_sunsky.npz"))[key]
    a = _assemble(**PT_GOLDEN_SUNSKY[key])
    assert a["params"].background_gradient == 3
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.01
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


DARKSKY_KEYS = ["darksky_seen_day", "darksky_seen_night_srgb_altitude", "darksky_ibl_path", "darksky_ibl_direct_no_exposure_gauss", "darksky_portal"]


@pytest.mark.parametrize("key", DARKSKY_KEYS)
def test_darksky_background_vs_golden_reference_render(key):
    """darksky: DarkSkyBackground::eval per direction (double Perez functions, ColorConv::fromxyY with exposure, the four XYZ -> RGB
    matrices, FAST_MATH gamma encoding, clamping, night tint), seen directly, tabulated for its `background_light`, through a portal."""
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_DARKSKY
    ref = np.load(os.path.join(GOLD, "pt_darksky.npz"))[key]
    a = _assemble(**PT_GOLDEN_DARKSKY[key])
    assert a["params"].background_gradient == 4
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.005
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


SKIES = [dict(type="sunsky", **{"from": (0.0, 0.0, 1.0)}, turbidity=2.0),                                        # sun at the zenith: atan2(0, 0) guard of phi_s
         dict(type="sunsky", **{"from": (1.0, 0.0, 0.0)}, turbidity=9.0, power=0.7, e_var=0.2),                  # sun on the horizon
         dict(type="sunsky", **{"from": (-0.3, 0.2, -0.9)}, turbidity=3.5, a_var=2.0, c_var=0.3),                # deep night
         dict(type="darksky", **{"from": (0.0, 0.0, 1.0)}, turbidity=2.0, bright=0.8, exposure=0.5),
         dict(type="darksky", **{"from": (0.6, 0.6, 0.05)}, turbidity=8.0, altitude=-0.2, color_space="CIE (D50)", power=2.0),
         dict(type="darksky", **{"from": (-0.5, 0.1, 0.4)}, night=True, exposure=3.0, color_space="no such space", d_var=1.7)]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("sky", SKIES, ids=lambda s: s["type"] + "_" + "_".join(str(v) for v in s["from"]))
def test_sky_models_vs_live_reference(sky):
    """sunsky / darksky over the corners of their parameters, seen in every direction by an equirectangular camera outside the box."""
    kw = dict(width=64, height=36, spp=2, bounces=2, background=sky,
              camera_extra=dict(type="equirectangular", **{"from": (0.5, 0.5, -0.6), "to": (0.5, 0.45, 1.0), "up": (0.5, 1.5, -0.6)}))
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    w, rgba = helpers.emu_pt_render(_assemble(**kw))
    img = interface.normalize_film(w, rgba)
    assert float(ref[..., :3].max()) > 0.0
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max()))


_HDR_IMG = {"filename": "tex.hdr", "color_space": "LinearRGB"}
TEXBACKS = [dict(_image=dict(_HDR_IMG, texture={"clipping": "clip"}), type="textureback", texture="bg_tex", rotation=350.0, power=0.5),
            dict(_image=dict(_HDR_IMG, texture={"clipping": "extend", "cropmin_x": 0.2, "cropmax_y": 0.7}), type="textureback", texture="bg_tex", mapping="probe", rotation=-30.0),
            dict(_image=dict(_HDR_IMG, texture={"clipping": "checker", "even_tiles": True, "interpolate": "none", "xrepeat": 4, "yrepeat": 2}), type="textureback", texture="bg_tex",
                 rotation=123.0),
            dict(_image={"filename": "tex.tga", "color_space": "sRGB", "image_optimization": "optimized", "texture": {"adj_contrast": 1.3, "adj_mult_factor_blue": 0.6, "adj_clamp": True}},
                 type="textureback", texture="bg_tex", mapping="angular")]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("bg", TEXBACKS, ids=["clip_rot350", "extend_crop_probe", "checker_nearest_repeat", "tga_srgb_adjusted_angular"])
def test_texture_background_variants_vs_live_reference(bg):
    """textureback over the texture's clipping / cropping / repeat / adjustment modes and both mappings (points clipped away are
    black in the texture and come out as the 1e-5 floor times power)."""
    kw = dict(width=64, height=36, spp=2, bounces=2, background=bg,
              camera_extra=dict(type="equirectangular", **{"from": (0.5, 0.5, -0.6), "to": (0.5, 0.45, 1.0), "up": (0.5, 1.5, -0.6)}))
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec, cwd=TEST01_DIR)
    w, rgba = helpers.emu_pt_render(_assemble(**kw))
    img = interface.normalize_film(w, rgba)
    assert float(ref[..., :3].max()) > 0.01
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max()))


AO_LIVE = [dict(oren_nayar=True, integrator_extra={"do_AO": True, "AO_samples": 5, "AO_distance": 0.3, "AO_color": (0.2, 0.9, 0.4, 1.0)}),
           dict(shiny={"translucency": 0.4, "transmit_filter": 0.6}, light_samples=3, integrator_extra={"do_AO": True, "AO_samples": 1, "AO_distance": 5.0}),
           dict(point_light=True, shiny={"specular_reflect": 0.3}, integrator_extra={"do_AO": True, "AO_samples": 2, "AO_distance": 0.2, "raydepth": 2}),
           dict(extra_lights=("spot", "sphere"), render_extra={"adv_auto_shadow_bias_enabled": False, "adv_shadow_bias_value": 0.002},
                integrator_extra={"do_AO": True, "AO_samples": 3, "AO_distance": 0.45})]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("cfg", AO_LIVE, ids=["oren_nayar_tinted", "translucent_one_sample", "behind_shiny_mirrors_point", "many_lights_manual_bias"])
def test_ambient_occlusion_variants_vs_live_reference(cfg):
    """do_AO with other BRDFs (Oren-Nayar, a translucency component next to the diffuse one), at the vertices behind mirror
    components, after several lights, and with a manual shadow bias (the occlusion rays start at it)."""
    kw = dict(width=64, height=36, spp=2, bounces=2, integrator="directlighting", **cfg)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    w, rgba = helpers.emu_pt_render(_assemble(**kw))
    img = interface.normalize_film(w, rgba)
    assert helpers.psnr(img, ref) >= 100.0, helpers.psnr(img, ref)
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max()))


ABSORB_LIVE = [dict(glass={"objects": ("sphere", "short"), "IOR": 1.5, "absorption": (0.2, 1.0, 1.0), "absorption_dist": 0.1}, mirror={"objects": ("back",), "color": (0.9, 0.9, 0.9)},
                    smooth_sphere=30.0, box_lift=0.013, integrator_extra={"raydepth": 5}),
               dict(glass={"objects": ("tall",), "IOR": 1.33, "absorption": (0.7, 0.8, 0.9), "absorption_dist": 0.0}, box_lift=0.013),            # dist == 0: sigma is not scaled
               dict(glass={"objects": ("sphere",), "IOR": 1.8, "absorption": (0.999, 0.5, 0.05), "absorption_dist": 3.0, "mirror_color": (0.6, 0.7, 0.8)}, smooth_sphere=60.0,
                    integrator="directlighting", background=(0.2, 0.3, 0.5), integrator_extra={"raydepth": 8})]


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
@pytest.mark.parametrize("cfg", ABSORB_LIVE, ids=["two_glasses_and_a_mirror", "dist_zero", "deep_tree_direct"])
def test_glass_absorption_variants_vs_live_reference(cfg):
    kw = dict(width=64, height=36, spp=2, bounces=3, **cfg)
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, **kw)
    ref, _ = helpers.ref_render(rec)
    w, rgba = helpers.emu_pt_render(_assemble(**kw))
    img = interface.normalize_film(w, rgba)
    assert helpers.psnr(img, ref) >= PSNR_MIN, helpers.psnr(img, ref)
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


@pytest.mark.parametrize("key", ["portal_mesh_blocks_filtered_shadows", "portal_mesh_blocks_filtered_shadows_paths"])
def test_portal_mesh_stays_a_blocker_for_filtered_shadow_rays(key):
    """transpShad + bgPortalLight: the light hides its mesh after the accelerator was built, so the triangles are in the tree;
    intersectTs checks the material's visibility only and lets them block the shadow rays of the other lights (the plug-in
    keeps the primitives and re-reads the visibilities; the Python mirror used to drop the mesh -- found by differential fuzzing)."""
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_PORTAL_TS
    ref = np.load(os.path.join(GOLD, "pt_portal_ts.npz"))[key]
    a = _assemble(**PT_GOLDEN_PORTAL_TS[key])
    assert int(((a["flags"] & 7) == 4).sum()) == 2          # the portal quad: neither seen nor a plain-shadow blocker, but a filtered-shadow one
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape and float(ref[..., :3].mean()) > 0.01
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


def test_automatic_two_level_upload_is_for_similarity_like_instance_matrices():
    """Instances with needle-like matrices make the automatic mode flatten the scene (the object-space node tests of the two-level
    traversal are padded for rotations / mirrors / near-uniform scales: found by fuzzing against the flattened scene and the kd-tree)."""
    yi = interface.Interface()
    scenes_pt.city(yi, n_tri_target=3000, n_base=2, grid=4, tri_range=(200, 800), width=32, height=18, spp=1, bounces=2, accelerator="yafaray-b200-bvh8")
    assert yi._wants_two_level() and not yi.warnings
    inst = next(o for o in yi.objects.values() if o.instance_of)
    keep = inst.matrix.copy()
    inst.matrix = keep @ np.diag([1.0, 0.01, 1.0, 1.0]).astype(np.float32)
    assert not yi._wants_two_level() and "uploaded flattened" in yi.warnings[-1]
    yi.device_instancing = True                       # an explicit request is honoured (the library still refuses singular matrices)
    assert yi._wants_two_level()
    yi.device_instancing = "auto"
    inst.matrix = keep @ np.diag([-1.0, 2.0, 1.5, 1.0]).astype(np.float32)        # mirrored, mildly non-uniform: fine
    assert yi._wants_two_level()
    yi.device_instancing = False
    assert not yi._wants_two_level()


def _shiny2_cfgs():
    sys.path.insert(0, GOLD)
    from make_golden_large import PT_GOLDEN_SHINY2
    return PT_GOLDEN_SHINY2


@pytest.mark.parametrize("key", ["pane_mirror_transp", "pane_all_four_fresnel_gauss", "pane_and_shiny_walls_transp_shadows", "pane_direct_point"])
def test_mirror_and_transparency_on_one_material_vs_golden_reference_render(key):
    """shinydiffusemat with specular_reflect AND transparency: getSpecular returns a reflected and a refracted ray (the ray tree
    walks both) and sample() chooses among all four components."""
    ref = np.load(os.path.join(GOLD, "pt_shiny2.npz"))[key]
    a = _assemble(**_shiny2_cfgs()[key])
    w, rgba = helpers.emu_pt_render(a)
    img = interface.normalize_film(w, rgba)
    assert img.shape == ref.shape
    p = helpers.psnr(img, ref)
    assert p >= PSNR_MIN, p
    assert np.abs(img - ref).max() <= 2e-5 * max(1.0, float(ref.max())), np.abs(img - ref).max()


@pytest.mark.skipif(not helpers.have_ref("ref_render"), reason="oracle/_ref not built")
def test_gradient_background_vs_live_reference():
    """GradientBackground::eval (sky / ground blend on dir.z, the 1e-6 floor) seen directly by an equirectangular camera; the
    ground colours default to the sky colours."""
    for bg in ({"type": "gradientback", "horizon_color": (1.0, 0.8, 0.6, 1.0), "zenith_color": (0.2, 0.4, 0.9, 1.0), "horizon_ground_color": (0.3, 0.25, 0.2, 1.0), "power": 0.8},
               {"type": "gradientback", "zenith_color": (0.0, 0.0, 0.0, 1.0)}):
        kw = dict(width=64, height=36, spp=3, bounces=3, background=bg,
                  camera_extra=dict(type="equirectangular", **{"from": (0.5, 0.5, -0.6), "to": (0.5, 0.45, 1.0), "up": (0.5, 1.5, -0.6)}))
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        ref, _ = helpers.ref_render(rec)
        w, rgba = helpers.emu_pt_render(_assemble(**kw))
        assert helpers.psnr(interface.normalize_film(w, rgba), ref) >= 100.0


def test_mirror_with_no_recursive_is_rejected():
    """no_recursive makes a mirror hit both start sub-paths and recurse (two branches per vertex): not covered, so refused."""
    yi = interface.Interface()
    scenes_pt.cornell_box(yi, width=16, height=9, spp=1, mirror=_M, integrator_extra=dict(no_recursive=True), accelerator="yafaray-b200-bvh8")
    with pytest.raises(Y4Error, match="no_recursive"):
        yi.assemble()


def test_set_image_color_goes_through_the_image_buffer_storage():
    """yafaray_setImageColor / getImageColor on in-memory and loaded images (Image::setColor rounding per optimisation mode)."""
    yi = interface.Interface()
    yi.image_dir = TEST01_DIR
    scenes_pt._set_params(yi, dict(type="ColorAlpha", image_optimization="optimized", width=4, height=3))
    assert yi.createImage("mem")
    assert yi.getImageColor("mem", 3, 2) == (0.0, 0.0, 0.0, 0.0) and yi.getImageColor("mem", 4, 0) is None
    assert yi.setImageColor("mem", 1, 2, 0.5, 0.25, 1.0, 0.3) and not yi.setImageColor("mem", 0, 3, 0, 0, 0, 0)
    r, g, b, a = yi.getImageColor("mem", 1, 2)
    assert r == np.float32(512.0) / np.float32(1023.0) and g == np.float32(256.0) / np.float32(1023.0) and b == 1.0
    assert a == np.float32(77.0) / np.float32(255.0)                      # roundf(0.3 * 255) = 77 (8-bit alpha)
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="Color", image_optimization="none", width=2, height=2))
    assert yi.createImage("rgb") and yi.setImageColor("rgb", 0, 0, 3.5, 0.1, 0.2, 0.4)
    assert yi.getImageColor("rgb", 0, 0) == (3.5, float(np.float32(0.1)), float(np.float32(0.2)), 1.0)      # no alpha channel: reads back 1
    yi.paramsClearAll()
    scenes_pt._set_params(yi, dict(type="Gray", width=2, height=2))
    assert not yi.createImage("gray") and yi.warnings


def test_interface_primitive_and_light_order():
    """Objects by name (std::map order, scene.cc:1034), faces in insertion order, invisible/base objects skipped."""
    yi = interface.Interface()
    yi.paramsSetString("type", "shinydiffusemat")
    yi.createMaterial("m0")
    yi.paramsClearAll()
    yi.paramsSetString("type", "light_mat")
    yi.createMaterial("m1")
    for name, mat, z, extra in (("zeta", "m0", 0.0, {}), ("alpha", "m1", 1.0, {}), ("hidden", "m0", 2.0, {"visibility": "invisible"}),
                                ("shadowy", "m0", 3.0, {"visibility": "shadow_only"})):
        yi.paramsClearAll()
        yi.paramsSetString("type", "mesh")
        for k, v in extra.items():
            yi.paramsSetString(k, v)
        yi.createObject(name)
        yi.setCurrentMaterial(mat)
        for v in ((0, 0, z), (1, 0, z), (0, 1, z)):
            yi.addVertex(*v)
        yi.addTriangle(0, 1, 2)
        yi.endObject()
    yi.paramsClearAll()
    yi.paramsSetString("type", "perspective")
    yi.createCamera("c")
    yi.paramsClearAll()
    yi.paramsSetString("type", "pathtracing")
    yi.createIntegrator("i")
    a = yi.assemble()
    assert list(a["tris"][:, 2]) == [1.0, 3.0, 0.0]          # alpha, shadowy, zeta
    assert list(a["flags"] >> 8) == [1, 0, 0]
    assert list(a["flags"] & 7) == [7, 2 | 4, 7]


def test_sampling_kats():
    """Device restatement of the reference samplers / fast maths vs KATs dumped from the reference (oracle/harness/ref_kat.cc)."""
    import ctypes as C

    k = json.load(open(os.path.join(GOLD, "kat_sampling.json")))
    E = helpers.emu()
    E.emu_lds.restype = C.c_double
    for dim, n, bits in k["lds"]:
        got = np.float64(E.emu_lds(C.c_int(dim), C.c_uint32(n))).view(np.uint64)
        assert int(got) == bits, (dim, n)
    o3, fnv = (C.c_float * 3)(), C.c_uint32()
    for i, r, vdc, s, lp, h in k["ri"]:
        E.emu_ri(C.c_uint32(i), C.c_uint32(r), o3, C.byref(fnv))
        assert [int(np.float32(x).view(np.uint32)) for x in o3] == [vdc, s, lp] and fnv.value == h
    o6 = (C.c_float * 6)()
    for row in k["halton"]:
        E.emu_halton(C.c_uint32(row[0]), C.c_uint32(row[1]), 6, o6)
        assert [int(np.float32(x).view(np.uint32)) for x in o6] == row[2:]
    bad = 0
    for xb, sb, cb, eb in k["trig"]:
        E.emu_trig(C.c_float(float(np.uint32(xb).view(np.float32))), o3)
        got = [int(np.float32(x).view(np.uint32)) for x in o3]
        bad += got != [sb, cb, eb]
    assert bad <= 2, bad   # fp64 instead of x87 80-bit intermediates: a last-bit flip is possible but rare
    o9, n3 = (C.c_float * 9)(), (C.c_float * 3)()
    bad = 0
    for row in k["coshemi"]:
        f = [float(np.uint32(b).view(np.float32)) for b in row]
        n3[:] = f[0:3]
        E.emu_coshemi(n3, C.c_float(f[3]), C.c_float(f[4]), o9)
        bad += [int(np.float32(x).view(np.uint32)) for x in o9] != row[5:]
    assert bad <= 2, bad


def test_fast_sampling_forms_equal_the_pinned_recurrences():
    """The cheaper forms used on the device (multiply-high digit extraction in the Faure-Halton loop, closed-form base-2
    Halton) are bit-identical to the loop forms that the KATs above pin to the reference."""
    import ctypes as C

    E = helpers.emu()
    E.emu_lds.restype = C.c_double
    E.emu_lds_loop.restype = C.c_double
    E.emu_halton2_at.restype = C.c_float
    rng = np.random.default_rng(5)
    ns = np.concatenate([rng.integers(0, 2 ** 32, 1500, dtype=np.uint64), [0, 1, 2, 2 ** 32 - 1, 2 ** 31, 3 ** 20, 3 ** 20 - 1, 229 * 229 * 229]]).astype(np.uint32)
    for dim in list(range(1, 50)):
        for n in ns[dim::7]:
            a, b = E.emu_lds(C.c_int(dim), C.c_uint32(int(n))), E.emu_lds_loop(C.c_int(dim), C.c_uint32(int(n)))
            assert np.float64(a).view(np.uint64) == np.float64(b).view(np.uint64), (dim, n)
    o = (C.c_float * 5)()
    starts = np.concatenate([rng.integers(0, 2 ** 32, 3000, dtype=np.uint64), [0, 1, 2 ** 32 - 1, 2 ** 32 - 2, 2 ** 32 - 5, 2 ** 31 - 1, 2 ** 31, 0x55555555, 0xAAAAAAAA],
                             2 ** rng.integers(1, 32, 64, dtype=np.uint64) - 1]).astype(np.uint32)
    for s in starts:
        E.emu_halton(C.c_uint32(2), C.c_uint32(int(s)), 5, o)
        for q in range(5):
            got = E.emu_halton2_at(C.c_uint32(int(s)), C.c_uint32(q + 1))
            assert np.float32(got).view(np.uint32) == np.float32(o[q]).view(np.uint32), (s, q)


def test_light_choice_index_is_an_equally_good_estimator():
    """The product numbers estimateOneDirectLight's draws per vertex (the reference's per-thread counter depends on
    thread scheduling).  Same estimator, different low-discrepancy index: the converged images agree."""
    kw = dict(PT_GOLDEN_LIGHTS["six_lights"], spp=48)
    a = _assemble(**kw)
    own = interface.normalize_film(*helpers.emu_pt_render(a))
    ref = interface.normalize_film(*helpers.emu_pt_render(a, ref_light_order=True))
    assert not np.array_equal(own, ref)
    lit = ref[..., :3].sum(-1) > 0.05
    assert abs(float(own[..., :3][lit].mean() / ref[..., :3][lit].mean()) - 1.0) < 0.01
    assert helpers.psnr(np.clip(own, 0, 2), np.clip(ref, 0, 2), peak=2.0) > 27.0   # noise-limited at 48 spp with six lights


def test_light_types_are_exported_in_name_order_with_their_parameters():
    yi = interface.Interface()
    scenes_pt.build(yi, accelerator="yafaray-b200-bvh8", **PT_GOLDEN_LIGHTS["six_lights"])
    a = yi.assemble()
    from libyafaray_b200 import _lib
    kinds = [a["lights"][i].type for i in range(a["n_light"])]      # name order: lamp, x_directional, x_directional_finite, x_sphere, x_spot, x_sun
    assert kinds == [_lib.LIGHT_AREA, _lib.LIGHT_DIRECTIONAL, _lib.LIGHT_DIRECTIONAL, _lib.LIGHT_SPHERE, _lib.LIGHT_SPOT, _lib.LIGHT_SUN]
    assert a["lights"][1].infinite == 1 and a["lights"][2].infinite == 0 and abs(a["lights"][2].radius - 0.22) < 1e-7
    assert abs(a["lights"][5].angle - 4.0) < 1e-7 and a["lights"][3].samples == 2 and abs(a["lights"][4].blend - 0.5) < 1e-7
