#!/usr/bin/env python3
"""TEST INFRASTRUCTURE: golden vectors for the full-size configurations, generated HERE (where /root/reference exists) by the
UNMODIFIED reference (oracle/_ref), committed so that the GPU box can check them without the reference tree.

  rays_soup5m.npz   <- BASELINE config C5 as SURVEY.md 8(d) prescribes it for the oracle: the first 5 000 000 triangles of the
                       50 M-triangle soup (identical generator, scenes.triangle_soup), 65 536 incoherent closest-hit rays and
                       65 536 segment (any-hit) rays against yafaray-kdtree-original
  pt_city_large.npz <- BASELINE config C4 at a size the reference renders in minutes: the instanced city (scenes_pt.city) with
                       >= 200 000 instanced triangles at 240 x 135, 8 bounces + NEE, reference PathIntegrator in the deterministic
                       mode (threads = 1, linear tiles, RR off)

  pt_glass.npz      <- Cornell box with `glass` objects (PT_GOLDEN_GLASS): reference PathIntegrator / DirectLightIntegrator, deterministic mode
  pt_ibl.npz        <- Cornell box lit by an IBL background light (PT_GOLDEN_IBL), deterministic mode
  pt_shiny2.npz     <- Cornell box with a shinydiffusemat sheet that is mirror AND transparent (PT_GOLDEN_SHINY2), deterministic mode
  pt_spot_soft.npz  <- Cornell box lit by a spotlight with soft_shadows = true (PT_GOLDEN_SPOT_SOFT), deterministic mode
  pt_portal.npz     <- Cornell box lit through a bgPortalLight across its opening (PT_GOLDEN_PORTAL), deterministic mode
  pt_rr.npz         <- Cornell box with the integrator's default Russian roulette (russian_roulette_min_bounces = 0), 1024 spp

Run from the repo root:  python tests/golden/make_golden_large.py [soup|city]
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from libyafaray_b200 import scenes, scenes_pt  # noqa: E402

TEST01_DIR = os.path.join(HERE, "test01")          # the reference's tests/test01 images (copied by make_golden.py)

SOUP5M = dict(n_tri=5_000_000, n_rays=65536)
SOUP50M = dict(n_tri=50_000_000, n_rays=65536)      # C5 at FULL size (needs ~30 GB of host memory and ~10 min for the kd-tree)
CITY1M = dict(scene="city", n_tri_target=1_000_000, n_base=32, grid=48, tri_range=(1000, 8000), width=480, height=270, spp=2, bounces=8, filter_type="gauss")
CITY200K = dict(scene="city", n_tri_target=200_000, n_base=16, grid=24, tri_range=(500, 4000), width=240, height=135, spp=2, bounces=8, filter_type="gauss")


def make_soup(cfg=SOUP5M, name="rays_soup5m.npz"):
    t0 = time.time()
    SOUP5M = cfg
    tris = np.concatenate([scenes.triangle_soup(min(5_000_000, cfg["n_tri"] - f), first=f) for f in range(0, cfg["n_tri"], 5_000_000)])
    inc, seg = scenes.incoherent_rays(SOUP5M["n_rays"]), scenes.segment_rays(SOUP5M["n_rays"])
    with helpers.RefServer(tris, threads=0) as srv:
        print("kd-tree built in %.1f s" % srv.info["build_s"], flush=True)
        h_inc, _ = srv.trace(inc, "closest")
        h_seg, _ = srv.trace(seg, "shadow")
        bound = np.array(srv.info["bound"], np.float32)
    # rays are regenerated from their seeds by the test (scenes.incoherent_rays / segment_rays): only the hit records are stored
    np.savez_compressed(os.path.join(HERE, name), scene=np.array([SOUP5M["n_tri"], SOUP5M["n_rays"]]),
                        hit_incoherent=h_inc, occluded_segment=np.packbits(h_seg["prim"] != -1), bound=bound)
    print(name + ": %d / %d incoherent hits, %d occluded segments, %.0f s" % ((h_inc["prim"] >= 0).sum(), len(inc), (h_seg["prim"] != -1).sum(), time.time() - t0))


def make_city():
    out = {}
    for key, cfg in (("city200k", CITY200K), ("city1m", CITY1M)):
        t0 = time.time()
        rec = helpers.ScriptRecorder()
        scenes_pt.build(rec, **cfg)
        img, info = helpers.ref_render(rec)
        out[key] = img.astype(np.float32)
        print("%s: %s, reference render %.1f s, total %.0f s" % (key, img.shape, info["render_s"], time.time() - t0), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_city_large.npz"), **out)


# glass: reflection + refraction at every vertex -> recursiveRaytrace is a tree (integrator_montecarlo.cc:918-968) -> pt_glass.npz
PT_GOLDEN_GLASS = {
    "glass_sphere": dict(width=96, height=54, spp=4, bounces=4, smooth_sphere=30.0, glass={"objects": ("sphere",), "IOR": 1.5}),
    "glass_box_tinted": dict(width=96, height=54, spp=3, bounces=4, box_lift=0.013, glass={"objects": ("tall",), "IOR": 1.6, "filter_color": (0.4, 0.9, 0.6), "transmit_filter": 0.8,
                                                                                 "mirror_color": (0.9, 0.8, 1.0)}),
    "glass_and_mirror": dict(width=80, height=45, spp=3, bounces=3, smooth_sphere=30.0, box_lift=0.013, glass={"objects": ("sphere", "short"), "IOR": 1.6, "transmit_filter": 0.5, "filter_color": (1.0, 0.7, 0.5)},
                             mirror={"objects": ("back",), "color": (0.9, 0.9, 0.9)}, integrator_extra={"raydepth": 4}),
    "glass_direct_point": dict(width=80, height=45, spp=2, bounces=3, integrator="directlighting", point_light=True, smooth_sphere=30.0, box_lift=0.013, glass={"objects": ("sphere", "tall"), "IOR": 1.45}),
    "glass_no_caustics_gauss": dict(width=80, height=45, spp=3, bounces=5, filter_type="gauss", smooth_sphere=30.0, glass={"objects": ("sphere",), "IOR": 1.5, "mirror_color": (0.5, 0.5, 0.5)},
                                    integrator_extra={"caustic_type": "none", "raydepth": 3}),
    "glass_deep_tree": dict(width=64, height=36, spp=2, bounces=5, box_lift=0.013, glass={"objects": ("tall", "short"), "IOR": 1.7, "transmit_filter": 0.7, "filter_color": (0.5, 0.8, 1.0)},
                            integrator_extra={"raydepth": 6}),
}


# glass with `absorption`: the rays inside the glass are attenuated by exp(-sigma_a * distance) -> pt_glass_absorb.npz
PT_GOLDEN_GLASS_ABSORB = {
    "glass_absorb_sphere": dict(width=96, height=54, spp=3, bounces=3, smooth_sphere=30.0,
                                glass={"objects": ("sphere",), "IOR": 1.5, "absorption": (0.3, 0.8, 0.5), "absorption_dist": 0.2}),
    "glass_absorb_boxes_default_dist": dict(width=80, height=45, spp=2, bounces=3, box_lift=0.013, integrator_extra={"raydepth": 6},
                                            glass={"objects": ("tall", "short"), "IOR": 1.6, "absorption": (0.9, 0.4, 0.7), "transmit_filter": 0.5,
                                                   "filter_color": (1.0, 0.9, 0.8)}),
    "glass_absorb_direct_gauss": dict(width=80, height=45, spp=2, bounces=2, integrator="directlighting", filter_type="gauss", smooth_sphere=30.0, box_lift=0.013,
                                      glass={"objects": ("sphere", "tall"), "IOR": 1.45, "absorption": (0.6, 0.6, 0.95), "absorption_dist": 0.05},
                                      background=(0.3, 0.4, 0.6)),
}


def make_glass_absorb():
    out = {}
    for key, kw in PT_GOLDEN_GLASS_ABSORB.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print(key, float(out[key][..., :3].mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_glass_absorb.npz"), **out)


def make_glass():
    out = {}
    for key, kw in PT_GOLDEN_GLASS.items():
        t0 = time.time()
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print("%s: %.1f s" % (key, time.time() - t0), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_glass.npz"), **out)


# IBL: a `constant` / `gradientback` background with ibl=True creates a `bglight` that importance-samples it (light_background.cc) -> pt_ibl.npz
_SKY = dict(type="gradientback", horizon_color=(1.0, 0.9, 0.7, 1.0), zenith_color=(0.2, 0.4, 1.0, 1.0), horizon_ground_color=(0.5, 0.4, 0.3, 1.0),
            zenith_ground_color=(0.1, 0.1, 0.1, 1.0), power=2.0, ibl=True)
PT_GOLDEN_IBL = {
    "ibl_constant": dict(width=96, height=54, spp=3, bounces=4, lamp=False, background=dict(type="constant", color=(0.8, 0.9, 1.0, 1.0), power=1.5, ibl=True, ibl_samples=3)),
    "ibl_gradient": dict(width=96, height=54, spp=4, bounces=5, lamp=False, smooth_sphere=30.0, background=dict(_SKY, ibl_samples=1)),
    "ibl_gradient_mirror_gauss": dict(width=80, height=45, spp=3, bounces=4, lamp=False, filter_type="gauss", background=dict(_SKY, ibl_samples=2),
                                      mirror={"objects": ("tall", "back"), "color": (0.9, 0.9, 0.9)}, integrator_extra={"caustic_type": "path"}),
    "ibl_direct_with_lamp": dict(width=80, height=45, spp=2, bounces=3, integrator="directlighting", light_samples=2, background=dict(_SKY, ibl_samples=4)),
    "ibl_no_shadows": dict(width=64, height=36, spp=2, bounces=3, lamp=False, background=dict(_SKY, ibl_samples=2, cast_shadows=False)),
    # two lights under path tracing: the reference picks the light of a path vertex from a call counter (checked through the emulation only)
    "ibl_path_with_lamp": dict(width=64, height=36, spp=3, bounces=4, background=dict(_SKY, ibl_samples=1)),
}


def make_ibl():
    out = {}
    for key, kw in PT_GOLDEN_IBL.items():
        t0 = time.time()
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print("%s: %.1f s" % (key, time.time() - t0), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_ibl.npz"), **out)


# shinydiffusemat with a mirror AND a transparency component: reflect + refract at one vertex (ray tree), all four components in sample() -> pt_shiny2.npz
PT_GOLDEN_SHINY2 = {
    "pane_mirror_transp": dict(width=96, height=54, spp=3, bounces=4, pane=dict(transparency=0.6, specular_reflect=0.3, transmit_filter=0.8, color=(0.3, 0.5, 0.9)),
                               integrator_extra={"raydepth": 3}),
    "pane_all_four_fresnel_gauss": dict(width=80, height=45, spp=3, bounces=4, filter_type="gauss",
                                        pane=dict(transparency=0.4, specular_reflect=0.5, fresnel_effect=True, IOR=1.5, translucency=0.3, mirror_color=(1.0, 0.8, 0.6, 1.0)),
                                        integrator_extra={"raydepth": 4, "caustic_type": "path"}),
    "pane_and_shiny_walls_transp_shadows": dict(width=80, height=45, spp=2, bounces=3, shiny=dict(specular_reflect=0.25), smooth_sphere=30.0,
                                                pane=dict(transparency=0.7, specular_reflect=0.2, transmit_filter=0.5, color=(0.9, 0.4, 0.3)),
                                                integrator_extra={"raydepth": 3, "transpShad": True, "shadowDepth": 3}),
    "pane_direct_point": dict(width=64, height=36, spp=2, bounces=3, integrator="directlighting", point_light=True,
                              pane=dict(transparency=0.5, specular_reflect=0.4), integrator_extra={"raydepth": 5}),
}


def make_shiny2():
    out = {}
    for key, kw in PT_GOLDEN_SHINY2.items():
        t0 = time.time()
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print("%s: %.1f s" % (key, time.time() - t0), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_shiny2.npz"), **out)


# spotlight with soft_shadows = true (SpotLight::illumSample / intersect) -> pt_spot_soft.npz
PT_GOLDEN_SPOT_SOFT = {
    "spot_soft_direct": dict(width=96, height=54, spp=3, bounces=3, integrator="directlighting", lamp=False, extra_lights=("spot_soft",), light_samples=3),
    "spot_soft_path_only_light": dict(width=80, height=45, spp=3, bounces=4, lamp=False, extra_lights=("spot_soft",), light_samples=1),
    "spot_soft_direct_with_lamp": dict(width=80, height=45, spp=2, bounces=3, integrator="directlighting", extra_lights=("spot_soft",), light_samples=2, filter_type="gauss"),
}


def make_spot_soft():
    out = {}
    for key, kw in PT_GOLDEN_SPOT_SOFT.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print(key, float(out[key][..., :3].mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_spot_soft.npz"), **out)


# bgPortalLight: a mesh across the box's opening emits the background seen through it (light_background_portal.cc) -> pt_portal.npz
_SKY_PLAIN = dict(type="gradientback", horizon_color=(1.0, 0.9, 0.7, 1.0), zenith_color=(0.2, 0.4, 1.0, 1.0), horizon_ground_color=(0.5, 0.4, 0.3, 1.0),
                  zenith_ground_color=(0.1, 0.1, 0.1, 1.0), power=2.0)
PT_GOLDEN_PORTAL = {
    "portal_gradient_path": dict(width=96, height=54, spp=3, bounces=4, lamp=False, portal=1.5, background=_SKY_PLAIN),
    "portal_constant_direct": dict(width=80, height=45, spp=2, bounces=3, integrator="directlighting", lamp=False, portal=2.0, light_samples=3,
                                   background=dict(type="constant", color=(0.8, 0.9, 1.0, 1.0), power=1.2)),
    "portal_direct_with_lamp_gauss": dict(width=80, height=45, spp=2, bounces=3, integrator="directlighting", portal=1.0, light_samples=2, filter_type="gauss",
                                          background=_SKY_PLAIN, smooth_sphere=30.0),
}


# The portal mesh under transpShad: the light hides it after the accelerator was built, so it stays in the tree, and intersectTs
# (material visibility only) lets it block the filtered shadow rays of the OTHER lights that shine in through the opening
# -> pt_portal_ts.npz (emulation test only: found by differential fuzzing after the last GPU minutes of the round were spent)
PT_GOLDEN_PORTAL_TS = {
    "portal_mesh_blocks_filtered_shadows": dict(width=80, height=45, spp=2, bounces=2, integrator="directlighting", lamp=False, portal=1.5, light_samples=2,
                                                extra_lights=("sun", "spot"), background=_SKY_PLAIN, integrator_extra={"transpShad": True, "shadowDepth": 3}),
    "portal_mesh_blocks_filtered_shadows_paths": dict(width=64, height=36, spp=2, bounces=3, lamp=False, portal=1.0, pane={"transparency": 0.6, "transmit_filter": 0.8},
                                                      background=_SKY_PLAIN, integrator_extra={"transpShad": True, "shadowDepth": 4}),
}


def make_portal():
    out = {}
    for key, kw in PT_GOLDEN_PORTAL.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print(key, float(out[key][..., :3].mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_portal.npz"), **out)
    out = {}
    for key, kw in PT_GOLDEN_PORTAL_TS.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print(key, float(out[key][..., :3].mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_portal_ts.npz"), **out)


# `angular` camera from inside the box: every projection, the circular mask (pixels without a camera ray: black, alpha 1),
# mirrored, with a box filter and a gauss filter, and once through the specular ray tree -> pt_angular.npz
_INSIDE = {"from": (0.5, 0.5, 0.25), "to": (0.52, 0.45, 1.0), "up": (0.5, 1.5, 0.25)}
PT_GOLDEN_ANGULAR = {
    "angular_equidistant_circular": dict(width=64, height=36, spp=3, bounces=3, camera_extra={"type": "angular", "angle": 90.0, **_INSIDE}),
    "angular_stereographic_mirrored": dict(width=64, height=36, spp=3, bounces=2,
                                           camera_extra={"type": "angular", "angle": 70.0, "circular": False, "mirrored": True, "projection": "stereographic", **_INSIDE}),
    "angular_orthographic_max_angle": dict(width=64, height=48, spp=2, bounces=2,
                                           camera_extra={"type": "angular", "angle": 80.0, "max_angle": 60.0, "projection": "orthographic", **_INSIDE}),
    "angular_equisolid_gauss": dict(width=64, height=36, spp=3, bounces=3, filter_type="gauss",
                                    camera_extra={"type": "angular", "angle": 100.0, "max_angle": 85.0, "projection": "equisolid_angle", **_INSIDE}),
    "angular_rectilinear_glass": dict(width=64, height=36, spp=2, bounces=3, smooth_sphere=30.0, glass={"objects": ("sphere",), "IOR": 1.5},
                                      camera_extra={"type": "angular", "angle": 50.0, "max_angle": 40.0, "projection": "rectilinear", **_INSIDE}),
    "angular_direct_wide": dict(width=64, height=36, spp=2, bounces=2, integrator="directlighting", light_samples=2,
                                camera_extra={"type": "angular", "angle": 120.0, "max_angle": 110.0, **_INSIDE}),
}


def make_angular():
    out = {}
    for key, kw in PT_GOLDEN_ANGULAR.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print(key, float(out[key][..., :3].mean()), float(out[key][..., 3].min()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_angular.npz"), **out)


# do_AO on the direct-lighting integrator: TiledIntegrator::sampleAmbientOcclusion after the lights, at every diffuse vertex
# integrate() reaches (also behind mirrors / glass) -> pt_ao.npz
PT_GOLDEN_AO = {
    "ao_default_color": dict(width=80, height=45, spp=2, bounces=2, integrator="directlighting", light_samples=2,
                             integrator_extra={"do_AO": True, "AO_samples": 4, "AO_distance": 0.4}),
    "ao_tinted_no_lamp_gauss": dict(width=80, height=45, spp=3, bounces=2, integrator="directlighting", lamp=False, filter_type="gauss", smooth_sphere=30.0,
                                    integrator_extra={"do_AO": True, "AO_samples": 3, "AO_distance": 0.25, "AO_color": (0.9, 0.8, 0.6, 1.0)}),
    "ao_behind_mirror_and_glass": dict(width=80, height=45, spp=2, bounces=2, integrator="directlighting", light_samples=1, smooth_sphere=30.0, box_lift=0.013,
                                       glass={"objects": ("sphere",), "IOR": 1.5}, mirror={"objects": ("back",), "color": (0.9, 0.9, 0.9)},
                                       integrator_extra={"do_AO": True, "AO_samples": 2, "AO_distance": 0.5, "raydepth": 3}),
    "ao_transparent_shadows_glow": dict(width=80, height=45, spp=2, bounces=2, integrator="directlighting", light_samples=1, glow=0.3,
                                        pane={"transparency": 0.7, "transmit_filter": 0.9},
                                        integrator_extra={"do_AO": True, "AO_samples": 3, "AO_distance": 0.6, "transpShad": True, "shadowDepth": 4}),
    "ao_ignored_by_the_path_tracer": dict(width=64, height=36, spp=2, bounces=3, integrator_extra={"do_AO": True, "AO_samples": 4, "AO_distance": 0.4}),
}


def make_ao():
    out = {}
    for key, kw in PT_GOLDEN_AO.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print(key, float(out[key][..., :3].mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_ao.npz"), **out)


# textureback: an HDR image as the background (spherical / angular mapping, rotation, power), lighting the box as an IBL light
# (with ibl_clamp_sampling) or through a portal -> pt_texback.npz
_HDR = {"filename": "tex.hdr", "color_space": "LinearRGB"}
_OUTSIDE = {"from": (0.5, 0.5, -0.6), "to": (0.5, 0.45, 1.0), "up": (0.5, 1.5, -0.6)}
PT_GOLDEN_TEXBACK = {
    "texback_seen_spherical": dict(width=64, height=36, spp=3, bounces=2, camera_extra=dict(type="equirectangular", **_OUTSIDE),
                                   background={"_image": _HDR, "type": "textureback", "texture": "bg_tex", "power": 1.5, "rotation": 40.0}),
    "texback_seen_angular_nearest": dict(width=64, height=36, spp=2, bounces=2, camera_extra=dict(type="equirectangular", **_OUTSIDE),
                                         background={"_image": dict(_HDR, texture={"interpolate": "none"}), "type": "textureback", "texture": "bg_tex",
                                                     "mapping": "angular", "rotation": 75.0}),
    "texback_ibl_path": dict(width=80, height=45, spp=3, bounces=3, lamp=False,
                             background={"_image": _HDR, "type": "textureback", "texture": "bg_tex", "power": 12.0, "ibl": True, "ibl_samples": 3}),
    "texback_ibl_clamp_direct": dict(width=80, height=45, spp=2, bounces=2, integrator="directlighting", lamp=False, filter_type="gauss",
                                     background={"_image": _HDR, "type": "textureback", "texture": "bg_tex", "power": 15.0, "rotation": 200.0, "ibl": True,
                                                 "ibl_samples": 4, "ibl_clamp_sampling": 2.5}),
    "texback_portal": dict(width=80, height=45, spp=2, bounces=3, lamp=False, portal=1.5,
                           background={"_image": _HDR, "type": "textureback", "texture": "bg_tex", "power": 10.0}),
}


def make_texback():
    out = {}
    for key, kw in PT_GOLDEN_TEXBACK.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec, cwd=TEST01_DIR)[0]
        print(key, float(out[key][..., :3].mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_texback.npz"), **out)


# sunsky: the Preetham daylight model as the background (day, and night with the below-horizon / night fades), as an IBL light
# (background_light) and through a portal -> pt_sunsky.npz
PT_GOLDEN_SUNSKY = {
    "sunsky_seen_day": dict(width=64, height=36, spp=3, bounces=2, camera_extra=dict(type="equirectangular", **_OUTSIDE),
                            background={"type": "sunsky", "from": (0.4, 0.7, 0.6), "turbidity": 3.0, "power": 1.5}),
    "sunsky_seen_night_variations": dict(width=64, height=36, spp=2, bounces=2, camera_extra=dict(type="equirectangular", **_OUTSIDE),
                                         background={"type": "sunsky", "from": (0.8, -0.5, -0.15), "turbidity": 6.0, "a_var": 1.2, "b_var": 0.8, "c_var": 1.1,
                                                     "d_var": 0.9, "e_var": 1.3}),
    "sunsky_ibl_path": dict(width=80, height=45, spp=3, bounces=3, lamp=False,
                            background={"type": "sunsky", "from": (0.3, -0.8, 0.5), "turbidity": 2.5, "power": 8.0, "background_light": True, "light_samples": 3}),
    "sunsky_ibl_direct_with_sun_gauss": dict(width=80, height=45, spp=2, bounces=2, integrator="directlighting", lamp=False, filter_type="gauss", extra_lights=("sun",),
                                             background={"type": "sunsky", "from": (-0.2, -0.9, 0.4), "power": 5.0, "background_light": True, "light_samples": 4,
                                                         "cast_shadows": False}),
    "sunsky_portal": dict(width=80, height=45, spp=2, bounces=3, lamp=False, portal=1.5,
                          background={"type": "sunsky", "from": (0.1, -0.6, 0.8), "turbidity": 4.0, "power": 6.0}),
}


def make_sunsky():
    out = {}
    for key, kw in PT_GOLDEN_SUNSKY.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print(key, float(out[key][..., :3].mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_sunsky.npz"), **out)


# darksky: the same daylight model with altitude, exposure, colour-space matrix, gamma encoding and night mode -> pt_darksky.npz
PT_GOLDEN_DARKSKY = {
    "darksky_seen_day": dict(width=64, height=36, spp=3, bounces=2, camera_extra=dict(type="equirectangular", **_OUTSIDE),
                             background={"type": "darksky", "from": (0.4, 0.7, 0.6), "turbidity": 3.0, "power": 1.2, "bright": 1.5}),
    "darksky_seen_night_srgb_altitude": dict(width=64, height=36, spp=2, bounces=2, camera_extra=dict(type="equirectangular", **_OUTSIDE),
                                             background={"type": "darksky", "from": (0.8, -0.5, 0.2), "turbidity": 5.0, "night": True, "altitude": 0.3,
                                                         "exposure": 2.0, "color_space": "sRGB (D65)", "a_var": 1.2, "b_var": 0.8, "c_var": 1.1, "d_var": 0.9,
                                                         "e_var": 1.3}),
    "darksky_ibl_path": dict(width=80, height=45, spp=3, bounces=3, lamp=False,
                             background={"type": "darksky", "from": (0.3, -0.8, 0.5), "turbidity": 2.5, "power": 4.0, "background_light": True, "light_samples": 3,
                                         "color_space": "CIE (D50)"}),
    "darksky_ibl_direct_no_exposure_gauss": dict(width=80, height=45, spp=2, bounces=2, integrator="directlighting", lamp=False, filter_type="gauss", extra_lights=("sun",),
                                                 background={"type": "darksky", "from": (-0.2, -0.9, 0.4), "power": 3.0, "exposure": 0.0, "background_light": True,
                                                             "light_samples": 4, "color_space": "sRGB (D50)"}),
    "darksky_portal": dict(width=80, height=45, spp=2, bounces=3, lamp=False, portal=1.5,
                           background={"type": "darksky", "from": (0.1, -0.6, 0.8), "turbidity": 4.0, "power": 5.0}),
}


def make_darksky():
    out = {}
    for key, kw in PT_GOLDEN_DARKSKY.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
        print(key, float(out[key][..., :3].mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "pt_darksky.npz"), **out)


RR_ON = dict(width=64, height=36, spp=1024, bounces=8, rr_min_bounces=0)     # the integrator's DEFAULT: Russian roulette from the first bounce


def make_rr():
    """Russian roulette draws from a rand()-seeded generator in the reference (integrator_tiled.cc:272) and from a per-path hash
    on the device: not sample-exact by construction, so the golden is a converged image (1024 spp) for a statistical test."""
    t0 = time.time()
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, threads=8, **RR_ON)
    img, info = helpers.ref_render(rec)
    np.savez_compressed(os.path.join(HERE, "pt_rr.npz"), rr_on=img.astype(np.float32))
    print("pt_rr.npz: %s, reference render %.1f s, total %.0f s" % (img.shape, info["render_s"], time.time() - t0))


if __name__ == "__main__":
    what = sys.argv[1:] or ["soup", "city", "rr"]
    if "city" in what:
        make_city()
    if "soup" in what:
        make_soup()
    if "glass" in what:
        make_glass()
    if "ibl" in what:
        make_ibl()
    if "shiny2" in what:
        make_shiny2()
    if "spot_soft" in what:
        make_spot_soft()
    if "portal" in what:
        make_portal()
    if "angular" in what:
        make_angular()
    if "ao" in what:
        make_ao()
    if "texback" in what:
        make_texback()
    if "glass_absorb" in what:
        make_glass_absorb()
    if "sunsky" in what:
        make_sunsky()
    if "darksky" in what:
        make_darksky()
    if "rr" in what:
        make_rr()
    if "soup50m" in what:
        make_soup(SOUP50M, "rays_soup50m.npz")
