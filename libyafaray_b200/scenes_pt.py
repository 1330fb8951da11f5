"""Procedural path-tracer scenes expressed as calls on a libYafaRay-style
C-API object (reference: include/public_api/yafaray_c_api.h:53-130; parameter names per
SURVEY.md appendix A).  `api` is anything with the C API's method names minus the `yafaray_`
prefix: libyafaray_b200.interface.Interface (the product) or a script recorder that replays the
same calls against the unmodified reference (oracle/harness/ref_render.cc) -- so both renderers
are driven by the very same call sequence."""
from __future__ import annotations

import numpy as np


def _material(api, name, color, diffuse_reflect=1.0, **extra):
    api.paramsClearAll()
    api.paramsSetString("type", "shinydiffusemat")
    api.paramsSetColor("color", color[0], color[1], color[2], 1.0)
    api.paramsSetFloat("diffuse_reflect", diffuse_reflect)
    for k, v in extra.items():
        if isinstance(v, str):
            api.paramsSetString(k, v)
        elif isinstance(v, bool):
            api.paramsSetBool(k, v)
        elif isinstance(v, (tuple, list)):
            api.paramsSetColor(k, float(v[0]), float(v[1]), float(v[2]), 1.0)
        else:
            api.paramsSetFloat(k, float(v))
    api.createMaterial(name)


def _mesh(api, name, material, verts, tris):
    api.paramsClearAll()
    api.paramsSetString("type", "mesh")
    api.paramsSetInt("num_vertices", len(verts))
    api.paramsSetInt("num_faces", len(tris))
    api.createObject(name)
    api.setCurrentMaterial(material)
    for v in verts:
        api.addVertex(float(v[0]), float(v[1]), float(v[2]))
    for t in tris:
        api.addTriangle(int(t[0]), int(t[1]), int(t[2]))
    api.endObject()


def _quad(p0, p1, p2, p3):
    return [p0, p1, p2, p3], [(0, 1, 2), (0, 2, 3)]


def _box(lo, hi, yaw_deg, centre):
    """Axis-aligned box [lo, hi] rotated by yaw about the vertical through `centre`; no bottom face."""
    x0, y0, z0 = lo
    x1, y1, z1 = hi
    c = np.array([[x0, y0, z0], [x1, y0, z0], [x1, y0, z1], [x0, y0, z1], [x0, y1, z0], [x1, y1, z0], [x1, y1, z1], [x0, y1, z1]], np.float64)
    a = np.radians(yaw_deg)
    r = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])
    ctr = np.array(centre, np.float64)
    c = (c - ctr) @ r.T + ctr
    faces = [(4, 7, 6, 5), (0, 1, 5, 4), (1, 2, 6, 5), (2, 3, 7, 6), (3, 0, 4, 7), (0, 3, 2, 1)]
    tris = []
    for f in faces:
        tris += [(f[0], f[1], f[2]), (f[0], f[2], f[3])]
    return c.astype(np.float32), tris


def _uv_sphere(centre, radius, nu, nv):
    """Latitude / longitude sphere with shared vertices (so that yafaray_smoothMesh has something to smooth)."""
    verts = [(centre[0], centre[1] + radius, centre[2])]
    for j in range(1, nv):
        th = np.pi * j / nv
        for i in range(nu):
            ph = 2.0 * np.pi * i / nu
            verts.append((centre[0] + radius * np.sin(th) * np.cos(ph), centre[1] + radius * np.cos(th), centre[2] + radius * np.sin(th) * np.sin(ph)))
    verts.append((centre[0], centre[1] - radius, centre[2]))
    tris = []
    ring = lambda j, i: 1 + (j - 1) * nu + (i % nu)
    for i in range(nu):
        tris.append((0, ring(1, i + 1), ring(1, i)))
        tris.append((len(verts) - 1, ring(nv - 1, i), ring(nv - 1, i + 1)))
    for j in range(1, nv - 1):
        for i in range(nu):
            tris.append((ring(j, i), ring(j, i + 1), ring(j + 1, i + 1)))
            tris.append((ring(j, i), ring(j + 1, i + 1), ring(j + 1, i)))
    return np.asarray(verts, np.float32), tris


def cornell_box(api, width=480, height=270, spp=16, bounces=8, path_samples=1, light_samples=1, filter_type="box",
                aa_pixelwidth=1.5, threads=1, accelerator="yafaray-kdtree-original", rr_min_bounces=None, point_light=False,
                oren_nayar=False, integrator="pathtracing", aperture=0.0, dof_distance=0.0, extra_lights=(), bokeh=None, smooth_sphere=None, camera_extra=None,
                mirror=None, integrator_extra=None, background=(0.0, 0.0, 0.0), glow=0.0, shiny=None, render_extra=None, pane=None, mesh_light=None, glass=None, box_lift=0.0, lamp=True, portal=None):
    """C2: Cornell box, ceiling area light (+ emitting quad via object_name), diffuse walls and two boxes.
    extra_lights: any of "sun", "directional", "directional_finite", "sphere", "spot" shining in through the open front.
    mirror: {"objects": subset of ("tall", "short", "back", "left", "sphere"), "color": rgb, "reflect": r} -> those objects get a
    `mirror` material; integrator_extra: integrator parameters set after the defaults (raydepth, caustic_type, ...).
    lamp=False: no ceiling lamp at all (the box is then lit through its open front, e.g. by a background with ibl=True).
    portal: power of a `bgPortalLight` whose mesh closes the open front (the background seen through it lights the box)."""
    white, red, green = (0.73, 0.73, 0.73), (0.65, 0.05, 0.05), (0.12, 0.45, 0.15)
    extra = {"diffuse_brdf": "oren_nayar", "sigma": 0.3} if oren_nayar else {}
    extra.update(shiny or {})          # shiny: shinydiffusemat parameters of all three wall materials (specular_reflect, mirror_color, fresnel_effect, IOR)
    if pane:                           # pane: shinydiffusemat parameters (transparency, transmit_filter, ...) of a sheet across the box at z = 0.45
        _material(api, "pane", pane.get("color", (0.3, 0.5, 0.9)), **{k: v for k, v in pane.items() if k != "color"})
    _material(api, "white", white, **extra)
    _material(api, "red", red, **extra)
    _material(api, "green", green, **({**extra, "emit": glow} if glow else extra))       # glow: the right wall also emits
    api.paramsClearAll()
    api.paramsSetString("type", "light_mat")
    api.paramsSetColor("color", 1.0, 0.85, 0.6, 1.0)
    api.paramsSetFloat("power", 18.0)
    api.createMaterial("emitter")
    mirrored = set((mirror or {}).get("objects", ()))
    if mirror:
        api.paramsClearAll()
        api.paramsSetString("type", "mirror")
        api.paramsSetColor("color", *mirror.get("color", (0.9, 0.9, 0.9)), 1.0)
        api.paramsSetFloat("reflect", mirror.get("reflect", 1.0))
        api.createMaterial("mirror")
    # glass: {"objects": subset of ("tall", "short", "sphere"), "IOR", "filter_color", "transmit_filter", "mirror_color"} -> a `glass` material
    glassy = set((glass or {}).get("objects", ()))
    if glass:
        api.paramsClearAll()
        api.paramsSetString("type", "glass")
        api.paramsSetFloat("IOR", float(glass.get("IOR", 1.5)))
        api.paramsSetColor("filter_color", *glass.get("filter_color", (1.0, 1.0, 1.0)), 1.0)
        api.paramsSetFloat("transmit_filter", float(glass.get("transmit_filter", 0.0)))
        api.paramsSetColor("mirror_color", *glass.get("mirror_color", (1.0, 1.0, 1.0)), 1.0)
        if "absorption" in glass:       # Beer-Lambert absorption inside the glass
            api.paramsSetColor("absorption", *glass["absorption"], 1.0)
            if "absorption_dist" in glass:
                api.paramsSetFloat("absorption_dist", float(glass["absorption_dist"]))
        api.createMaterial("glass")
    mat = lambda obj, default: "mirror" if obj in mirrored else ("glass" if obj in glassy else default)

    api.startGeometry()
    v, t = _quad((0, 0, 0), (1, 0, 0), (1, 0, 1), (0, 0, 1))
    _mesh(api, "a_floor", "white", v, t)
    v, t = _quad((0, 1, 0), (0, 1, 1), (1, 1, 1), (1, 1, 0))
    _mesh(api, "b_ceiling", "white", v, t)
    v, t = _quad((0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1))
    _mesh(api, "c_back", mat("back", "white"), v, t)
    v, t = _quad((0, 0, 0), (0, 0, 1), (0, 1, 1), (0, 1, 0))
    _mesh(api, "d_left", mat("left", "red"), v, t)
    v, t = _quad((1, 0, 0), (1, 1, 0), (1, 1, 1), (1, 0, 1))
    _mesh(api, "e_right", "green", v, t)
    # box_lift: raises the two boxes off the floor (a glass box standing ON the floor has its bottom face coincide with the floor:
    # rays inside it meet two surfaces at exactly the same t, the documented tie class)
    v, t = _box((0.53, box_lift, 0.53), (0.83, 0.6, 0.83), 18.0, (0.68, 0.0, 0.68))
    _mesh(api, "f_tall", mat("tall", "white"), v, t)
    v, t = _box((0.18, box_lift, 0.2), (0.48, 0.3, 0.5), -17.0, (0.33, 0.0, 0.35))
    _mesh(api, "g_short", mat("short", "white"), v, t)
    # emitting quad just below the ceiling, facing down; the area light shares its corners
    lx0, lx1, lz0, lz1, ly = 0.37, 0.63, 0.395, 0.605, 0.999
    if not point_light and lamp:
        v, t = _quad((lx0, ly, lz0), (lx1, ly, lz0), (lx1, ly, lz1), (lx0, ly, lz1))
        _mesh(api, "h_lamp", "emitter", v, t)
    if smooth_sphere is not None:       # a sphere on the short box, shaded with yafaray_smoothMesh(angle)
        v, t = _uv_sphere((0.33, 0.42, 0.35), 0.12, 12, 7)
        _mesh(api, "i_sphere", mat("sphere", "white"), v, t)
        api.smoothMesh("i_sphere", float(smooth_sphere))
    if mesh_light and mesh_light.get("shape") == "sphere":      # a small emitting sphere that an objectlight turns into a light source
        v, t = _uv_sphere((0.3, 0.7, 0.35), 0.07, 8, 5)
        _mesh(api, "k_bulb", "emitter", v, t)
    if pane:
        v, t = _quad((0.1, 0.05, 0.45), (0.9, 0.05, 0.45), (0.9, 0.8, 0.45), (0.1, 0.8, 0.45))
        _mesh(api, "j_pane", "pane", v, t)
    if portal is not None:              # across the open front, facing into the box (the light makes the mesh itself invisible)
        v, t = _quad((0.0, 0.0, 0.0), (1.0, 0.0, 0.0), (1.0, 1.0, 0.0), (0.0, 1.0, 0.0))
        _mesh(api, "p_portal", "white", v, t)
    api.endGeometry()

    api.paramsClearAll()
    if point_light:
        api.paramsSetString("type", "pointlight")
        api.paramsSetVector("from", 0.5, 0.9, 0.5)
        api.paramsSetColor("color", 1.0, 0.85, 0.6, 1.0)
        api.paramsSetFloat("power", 0.6)
    elif mesh_light:                                    # objectlight on the lamp quad (2 triangles) or on the bulb sphere
        api.paramsSetString("type", "objectlight")
        api.paramsSetString("object_name", "k_bulb" if mesh_light.get("shape") == "sphere" else "h_lamp")
        api.paramsSetColor("color", 1.0, 0.85, 0.6, 1.0)
        api.paramsSetFloat("power", float(mesh_light.get("power", 18.0)))
        api.paramsSetInt("samples", light_samples)
        api.paramsSetBool("double_sided", bool(mesh_light.get("double_sided", False)))
    else:
        api.paramsSetString("type", "arealight")
        api.paramsSetVector("corner", lx0, ly, lz0)
        api.paramsSetVector("point1", lx1, ly, lz0)   # to_x = point1 - corner; emits toward to_x ^ to_y = -y
        api.paramsSetVector("point2", lx0, ly, lz1)
        api.paramsSetColor("color", 1.0, 0.85, 0.6, 1.0)
        api.paramsSetFloat("power", 18.0)
        api.paramsSetInt("samples", light_samples)
        api.paramsSetString("object_name", "h_lamp")
    if lamp:
        api.createLight("lamp")
    if portal is not None:
        api.paramsClearAll()
        api.paramsSetString("type", "bgPortalLight")
        api.paramsSetString("object_name", "p_portal")
        api.paramsSetFloat("power", float(portal))
        api.paramsSetInt("samples", light_samples)
        api.createLight("portal")
    for kind in extra_lights:
        api.paramsClearAll()
        if kind == "sun":                               # wide cone so that several samples differ visibly
            api.paramsSetString("type", "sunlight")
            api.paramsSetVector("direction", 0.35, 0.55, -1.0)
            api.paramsSetColor("color", 1.0, 0.95, 0.8, 1.0)
            api.paramsSetFloat("power", 0.9)
            api.paramsSetFloat("angle", 4.0)
            api.paramsSetInt("samples", light_samples)
        elif kind == "directional":
            api.paramsSetString("type", "directional")
            api.paramsSetVector("direction", -0.6, 0.25, -1.0)
            api.paramsSetColor("color", 0.6, 0.7, 1.0, 1.0)
            api.paramsSetFloat("power", 0.5)
        elif kind == "directional_finite":              # a beam of radius 0.22 through the front opening
            api.paramsSetString("type", "directional")
            api.paramsSetVector("direction", 0.1, 0.4, -1.0)
            api.paramsSetBool("infinite", False)
            api.paramsSetVector("from", 0.45, 0.9, -1.2)
            api.paramsSetFloat("radius", 0.22)
            api.paramsSetColor("color", 1.0, 0.4, 0.3, 1.0)
            api.paramsSetFloat("power", 1.5)
        elif kind in ("spot", "spot_soft"):             # from the front opening onto the short box, wide blend
            api.paramsSetString("type", "spotlight")
            api.paramsSetVector("from", 0.75, 0.6, -0.9)
            api.paramsSetVector("to", 0.35, 0.2, 0.4)
            api.paramsSetFloat("cone_angle", 14.0)
            api.paramsSetFloat("blend", 0.5)
            api.paramsSetColor("color", 0.9, 0.6, 1.0, 1.0)
            api.paramsSetFloat("power", 2.5)
            if kind == "spot_soft":                     # a sampled light: cone-jittered shadow rays
                api.paramsSetBool("soft_shadows", True)
                api.paramsSetFloat("shadowFuzzyness", 0.7)
                api.paramsSetInt("samples", light_samples)
        elif kind == "sphere":
            api.paramsSetString("type", "spherelight")
            api.paramsSetVector("from", 0.2, 0.75, 0.3)
            api.paramsSetFloat("radius", 0.06)
            api.paramsSetColor("color", 0.5, 1.0, 0.6, 1.0)
            api.paramsSetFloat("power", 9.0)
            api.paramsSetInt("samples", light_samples)
        else:
            raise ValueError(kind)
        api.createLight("x_" + kind)

    api.paramsClearAll()
    api.paramsSetString("type", "perspective")
    api.paramsSetVector("from", 0.5, 0.5, -1.35)
    api.paramsSetVector("to", 0.5, 0.5, 0.0)
    api.paramsSetVector("up", 0.5, 1.5, -1.35)
    api.paramsSetInt("resx", width)
    api.paramsSetInt("resy", height)
    api.paramsSetFloat("focal", 1.6 * height / width if width > height else 1.6)
    if aperture:
        api.paramsSetFloat("aperture", aperture)
        _set_params(api, bokeh)         # bokeh_type / bokeh_bias / bokeh_rotation
        api.paramsSetFloat("dof_distance", dof_distance)
    _set_params(api, camera_extra)      # another camera type / its parameters
    api.createCamera("cam")

    api.paramsClearAll()
    if isinstance(background, dict) and "_image" in background:     # textureback: the image file and its texture come first
        img = background["_image"]                                  # {"filename", "color_space", optional texture parameters}
        api.paramsSetString("type", "image")
        api.paramsSetString("filename", img["filename"])
        api.paramsSetString("color_space", img.get("color_space", "LinearRGB"))
        api.paramsSetFloat("gamma", 1.0)
        api.paramsSetString("image_optimization", img.get("image_optimization", "none"))
        api.createImage("bg_image")
        api.paramsClearAll()
        api.paramsSetString("type", "image")
        api.paramsSetString("image_name", "bg_image")
        _set_params(api, img.get("texture"))
        api.createTexture(background.get("texture", "bg_tex"))
        api.paramsClearAll()
        background = {k: v for k, v in background.items() if k != "_image"}
    if isinstance(background, dict):    # another background type with its parameters, e.g. gradientback
        _set_params(api, background)
    else:
        api.paramsSetString("type", "constant")
        api.paramsSetColor("color", *background, 1.0)
    api.createBackground("bg")

    api.paramsClearAll()
    api.paramsSetString("type", integrator)
    api.paramsSetInt("bounces", bounces)
    api.paramsSetInt("path_samples", path_samples)
    api.paramsSetInt("russian_roulette_min_bounces", bounces if rr_min_bounces is None else rr_min_bounces)
    api.paramsSetString("caustic_type", "none")
    api.paramsSetInt("raydepth", 2)
    _set_params(api, integrator_extra)
    api.createIntegrator("pt")

    api.paramsClearAll()
    api.paramsSetString("camera_name", "cam")
    api.createRenderView("view")

    api.paramsClearAll()
    api.paramsSetString("type", "combined")
    api.paramsSetString("image_type", "ColorAlpha")
    api.paramsSetString("exported_image_type", "ColorAlpha")
    api.paramsSetString("exported_image_name", "Combined")
    api.defineLayer()

    api.paramsClearAll()
    api.paramsSetString("integrator_name", "pt")
    api.paramsSetString("background_name", "bg")
    api.paramsSetString("scene_accelerator", accelerator)
    api.paramsSetInt("width", width)
    api.paramsSetInt("height", height)
    api.paramsSetInt("threads", threads)
    api.paramsSetInt("AA_passes", 1)
    api.paramsSetInt("AA_minsamples", spp)
    api.paramsSetInt("AA_inc_samples", 1)
    api.paramsSetFloat("AA_pixelwidth", aa_pixelwidth)
    api.paramsSetString("filter_type", filter_type)
    api.paramsSetString("tiles_order", "linear")
    api.paramsSetInt("tile_size", 32)
    _set_params(api, render_extra)      # e.g. AA_passes, film_load_save_mode / film_load_save_path, adv_computer_node
    api.setupRender()


# ---- C4: instanced city -------------------------------------------------------------------------------
def _hash01(seed, *idx):
    from .scenes import hash01

    return hash01(seed, *[np.asarray(i) for i in idx])


def _building(seed: int, b: int, n_tri_target: int):
    """One base "building": unit-footprint box, 4 facades tessellated nx x ny and displaced along their normals, flat roof.
    -> (verts [n, 3] float64, tris [m, 3] int64); m close to n_tri_target."""
    height = 1.0 + 5.0 * float(_hash01(seed, b, 0))
    n_quads = max(4, (n_tri_target - 2) // 8)
    ny = max(2, int(round(np.sqrt(n_quads * height))))
    nx = max(2, n_quads // ny)
    verts, tris = [], []
    corners = [(-0.5, -0.5), (0.5, -0.5), (0.5, 0.5), (-0.5, 0.5)]
    for f in range(4):
        (x0, z0), (x1, z1) = corners[f], corners[(f + 1) % 4]
        nrm = np.array([z1 - z0, 0.0, -(x1 - x0)])
        u, v = np.meshgrid(np.arange(nx + 1) / nx, np.arange(ny + 1) / ny, indexing="xy")
        amp = 0.04 * (u > 0) * (u < 1) * (v > 0) * (v < 1)      # facade edges stay put so the faces close up
        d = (_hash01(seed, np.full(u.shape, b * 8 + f), (v * ny).astype(np.int64) * 4099 + (u * nx).astype(np.int64)) - 0.5) * amp
        p = np.stack([x0 + (x1 - x0) * u, height * v, z0 + (z1 - z0) * u], -1) + d[..., None] * nrm
        base = len(verts)
        verts.extend(p.reshape(-1, 3).tolist())
        ii, jj = np.meshgrid(np.arange(nx), np.arange(ny), indexing="xy")
        a = base + jj * (nx + 1) + ii
        q = np.stack([a, a + 1, a + nx + 2, a, a + nx + 2, a + nx + 1], -1).reshape(-1, 3)
        tris.extend(q.tolist())
    base = len(verts)
    verts.extend([(-0.5, height, -0.5), (0.5, height, -0.5), (0.5, height, 0.5), (-0.5, height, 0.5)])
    tris.extend([(base, base + 2, base + 1), (base, base + 3, base + 2)])
    return np.asarray(verts, np.float64), np.asarray(tris, np.int64)


def _bulk_mesh(api, name, material, verts, tris, base_object=False):
    api.paramsClearAll()
    api.paramsSetString("type", "mesh")
    api.paramsSetInt("num_vertices", len(verts))
    api.paramsSetInt("num_faces", len(tris))
    if base_object:
        api.paramsSetBool("is_base_object", True)
    api.createObject(name)
    api.setCurrentMaterial(material)
    if "add_mesh_arrays" in dir(api):
        api.add_mesh_arrays(verts, tris)
    else:
        for v in verts:
            api.addVertex(float(v[0]), float(v[1]), float(v[2]))
        for t in tris:
            api.addTriangle(int(t[0]), int(t[1]), int(t[2]))
    api.endObject()


def city(api, n_tri_target=10_000_000, width=3840, height=2160, spp=16, bounces=8, seed=0xC4, n_base=64, grid=128,
         tri_range=(2000, 40000), threads=1, accelerator="yafaray-kdtree-original", filter_type="box", light_samples=1):
    """C4: `n_base` base building meshes instanced (yafaray_addInstance, flattened by the scene like the reference does)
    on a jittered `grid` x `grid` lattice with random yaw / scale until `n_tri_target` instanced triangles, a ground quad,
    one large sun-like area light, constant sky.  -> number of instanced triangles."""
    cols = [(0.70, 0.68, 0.64), (0.55, 0.36, 0.30), (0.45, 0.50, 0.58), (0.80, 0.76, 0.60)]
    for i, c in enumerate(cols):
        _material(api, "wall%d" % i, c)
    _material(api, "ground", (0.35, 0.36, 0.34))
    api.startGeometry()
    counts = []
    for b in range(n_base):
        lo, hi = tri_range
        target = int(lo * (hi / lo) ** float(_hash01(seed, b, 1)))
        v, t = _building(seed, b, target)
        counts.append(len(t))
        _bulk_mesh(api, "base%02d" % b, "wall%d" % (b % len(cols)), v, t, base_object=True)
    ext = float(grid)
    v, t = _quad((-2.0, 0.0, -2.0), (-2.0, 0.0, ext + 2.0), (ext + 2.0, 0.0, ext + 2.0), (ext + 2.0, 0.0, -2.0))
    _mesh(api, "a_ground", "ground", v, t)
    total, k = 2, 0
    while total < n_tri_target and k < grid * grid:
        # visit lattice cells in a fixed pseudo-random order so a small target still spreads over the whole city
        cell = (k * 7919) % (grid * grid)
        gx, gz = cell % grid, cell // grid
        b = int(_hash01(seed, k, 2) * n_base) % n_base
        yaw = 2.0 * np.pi * float(_hash01(seed, k, 3))
        sc = 0.55 + 0.35 * float(_hash01(seed, k, 4))
        cx = gx + 0.5 + 0.2 * (float(_hash01(seed, k, 5)) - 0.5)
        cz = gz + 0.5 + 0.2 * (float(_hash01(seed, k, 6)) - 0.5)
        c, s = np.cos(yaw) * sc, np.sin(yaw) * sc
        m = [c, 0.0, s, cx, 0.0, sc, 0.0, 0.0, -s, 0.0, c, cz, 0.0, 0.0, 0.0, 1.0]
        api.addInstance("base%02d" % b, *[float(x) for x in m])
        total += counts[b]
        k += 1
    api.endGeometry()

    api.paramsClearAll()
    api.paramsSetString("type", "arealight")            # sun-like: a big quad high above, tilted, emitting downwards
    h = 0.9 * ext
    api.paramsSetVector("corner", 0.15 * ext, h, 0.15 * ext)
    api.paramsSetVector("point1", 0.45 * ext, h, 0.15 * ext)
    api.paramsSetVector("point2", 0.15 * ext, h + 0.1 * ext, 0.45 * ext)
    api.paramsSetColor("color", 1.0, 0.96, 0.9, 1.0)
    api.paramsSetFloat("power", 60.0)
    api.paramsSetInt("samples", light_samples)
    api.createLight("sun")

    api.paramsClearAll()
    api.paramsSetString("type", "perspective")
    api.paramsSetVector("from", -0.12 * ext, 0.22 * ext, -0.12 * ext)
    api.paramsSetVector("to", 0.45 * ext, 0.0, 0.45 * ext)
    api.paramsSetVector("up", -0.12 * ext, 0.22 * ext + 1.0, -0.12 * ext)
    api.paramsSetInt("resx", width)
    api.paramsSetInt("resy", height)
    api.paramsSetFloat("focal", 1.2)
    api.createCamera("cam")

    api.paramsClearAll()
    api.paramsSetString("type", "constant")
    api.paramsSetColor("color", 0.45, 0.6, 0.85, 1.0)
    api.createBackground("bg")

    api.paramsClearAll()
    api.paramsSetString("type", "pathtracing")
    api.paramsSetInt("bounces", bounces)
    api.paramsSetInt("path_samples", 1)
    api.paramsSetInt("russian_roulette_min_bounces", bounces)
    api.paramsSetString("caustic_type", "none")
    api.paramsSetInt("raydepth", 2)
    api.createIntegrator("pt")

    api.paramsClearAll()
    api.paramsSetString("camera_name", "cam")
    api.createRenderView("view")

    api.paramsClearAll()
    api.paramsSetString("type", "combined")
    api.paramsSetString("image_type", "ColorAlpha")
    api.paramsSetString("exported_image_type", "ColorAlpha")
    api.paramsSetString("exported_image_name", "Combined")
    api.defineLayer()

    api.paramsClearAll()
    api.paramsSetString("integrator_name", "pt")
    api.paramsSetString("background_name", "bg")
    api.paramsSetString("scene_accelerator", accelerator)
    api.paramsSetInt("width", width)
    api.paramsSetInt("height", height)
    api.paramsSetInt("threads", threads)
    api.paramsSetInt("AA_passes", 1)
    api.paramsSetInt("AA_minsamples", spp)
    api.paramsSetInt("AA_inc_samples", 1)
    api.paramsSetFloat("AA_pixelwidth", 1.5)
    api.paramsSetString("filter_type", filter_type)
    api.paramsSetString("tiles_order", "linear")
    api.paramsSetInt("tile_size", 32)
    api.setupRender()
    return total


# ---- C1: the reference's tests/test01 scene ---------------------------------------------------------------
# Six textured cubes on a plane, one point light, constant background (tests/test01/test01.c of the reference:
# geometry :660-860, materials :268-650, light :652-663, camera :862-875, render parameters :896-960).  Each cube has
# orco points -1..1 and a shinydiffusemat whose diffuse_shader is a `layer` (mix) over a cube-mapped `texture_mapper`
# on an image texture; the six textures are one picture in six file formats.
TEST01_CUBES = [    # object, material, image file, colour space, x range, y range, diffuse_reflect, layer colfac
    ("Cube.005", "Material.008", "Texture.005", "tex.tif", "sRGB", (-4.40469, -2.40468), (1.44162, 3.44162), 0.5, 1.0),
    ("Cube.004", "Material.007", "Texture.004", "tex.tga", "sRGB", (3.26859, 5.26859), (-0.393062, 1.60694), 1.0, 1.0),
    ("Cube.003", "Material.006", "Texture.003", "tex.png", "sRGB", (-0.635578, 1.36442), (3.54144, 5.54144), 0.6, 1.0),
    ("Cube.002", "Material.005", "Texture.002", "tex.jpg", "sRGB", (-0.635578, 1.36442), (-0.393062, 1.60694), 0.4, 1.0),
    ("Cube.001", "Material.004", "Texture.001", "tex.hdr", "LinearRGB", (-0.635578, 1.36442), (-3.81854, -1.81854), 1.0, 0.95),
    ("Cube", "Material.003", "Texture", "tex.exr", "LinearRGB", (-5.01096, -3.01096), (-1.94285, 0.0571451), 1.0, 0.95),
]
TEST01_CUBE_TRIS = [(2, 0, 1), (2, 1, 3), (3, 7, 6), (3, 6, 2), (7, 5, 4), (7, 4, 6), (0, 4, 5), (0, 5, 1), (0, 2, 6), (0, 6, 4), (5, 7, 3), (5, 3, 1)]


def _set_params(api, extra):
    """Extra parameters of a scene variant: str / bool / int / float / 3-tuple (vector) / 4-tuple (colour)."""
    for k, v in (extra or {}).items():
        if isinstance(v, str):
            api.paramsSetString(k, v)
        elif isinstance(v, bool):
            api.paramsSetBool(k, v)
        elif isinstance(v, int):
            api.paramsSetInt(k, v)
        elif isinstance(v, float):
            api.paramsSetFloat(k, v)
        elif len(v) == 3:
            api.paramsSetVector(k, float(v[0]), float(v[1]), float(v[2]))
        else:
            api.paramsSetColor(k, float(v[0]), float(v[1]), float(v[2]), float(v[3]))


def test01(api, width=480, height=270, spp=3, integrator="pathtracing", bounces=4, path_samples=1, filter_type="gauss", aa_pixelwidth=1.5,
           accelerator="yafaray-kdtree-original", threads=1, texture_extra=None, mapper_extra=None, layer_extra=None, with_uv=False):
    """tests/test01 with the integrator of BASELINE config 1 (`pathtracing`; the shipped file uses `directlighting`, which
    `integrator="directlighting"` selects).  Image files are looked up relative to the renderer's working / image directory."""
    for _, _, tex, fname, cspace, _, _, _, _ in TEST01_CUBES:
        api.paramsClearAll()
        api.paramsSetString("type", "image")       # as in test01.c (not an Image::Type name; the loader picks the type)
        api.paramsSetString("filename", fname)
        api.paramsSetString("color_space", cspace)
        api.paramsSetFloat("gamma", 1.0)
        api.paramsSetString("image_optimization", "optimized")
        api.createImage(tex + "_image")
        api.paramsClearAll()
        api.paramsSetString("type", "image")
        api.paramsSetString("image_name", tex + "_image")
        _set_params(api, texture_extra)
        api.createTexture(tex)
    _material(api, "defaultMat", (0.8, 0.8, 0.8))
    for _, mat, tex, _, _, _, _, diffuse_reflect, colfac in TEST01_CUBES:
        api.paramsClearAll()
        api.paramsSetString("type", "shinydiffusemat")
        api.paramsSetColor("color", 0.8, 0.8, 0.8, 1.0)
        api.paramsSetFloat("diffuse_reflect", diffuse_reflect)
        api.paramsSetString("diffuse_shader", "diff_layer0")
        api.paramsPushList()
        api.paramsSetString("element", "shader_node")
        api.paramsSetString("type", "layer")
        api.paramsSetString("name", "diff_layer0")
        api.paramsSetString("input", "map0")
        api.paramsSetString("blend_mode", "mix")
        api.paramsSetFloat("colfac", colfac)
        api.paramsSetBool("color_input", True)
        api.paramsSetColor("def_col", 1.0, 0.0, 1.0, 1.0)
        api.paramsSetBool("do_color", True)
        api.paramsSetBool("do_scalar", False)
        api.paramsSetColor("upper_color", 0.8, 0.8, 0.8, 1.0)
        api.paramsSetFloat("upper_value", 0.0)
        _set_params(api, layer_extra)
        api.paramsPushList()
        api.paramsSetString("element", "shader_node")
        api.paramsSetString("type", "texture_mapper")
        api.paramsSetString("name", "map0")
        api.paramsSetString("texture", tex)
        api.paramsSetString("texco", "orco")
        api.paramsSetString("mapping", "cube")
        api.paramsSetInt("proj_x", 1)
        api.paramsSetInt("proj_y", 2)
        api.paramsSetInt("proj_z", 3)
        api.paramsSetVector("scale", 1.0, 1.0, 1.0)
        api.paramsSetVector("offset", 0.0, 0.0, 0.0)
        _set_params(api, mapper_extra)
        api.paramsEndList()
        api.createMaterial(mat)
    _material(api, "Material.002", (0.8, 0.8, 0.8))

    api.paramsClearAll()
    api.paramsSetString("type", "pointlight")
    api.paramsSetColor("color", 1.0, 1.0, 1.0, 1.0)
    api.paramsSetVector("from", 5.27648, -4.88993, 8.89514)
    api.paramsSetFloat("power", 72.0)
    api.createLight("Point")

    for obj, mat, _, _, _, xr, yr, _, _ in TEST01_CUBES:
        api.paramsClearAll()
        api.paramsSetString("type", "mesh")
        api.paramsSetBool("has_orco", True)
        api.paramsSetBool("has_uv", bool(with_uv))
        api.paramsSetInt("num_vertices", 8)
        api.paramsSetInt("num_faces", 6)
        api.createObject(obj)
        for ix in (0, 1):
            for iy in (0, 1):
                for iz in (0, 1):
                    api.addVertexWithOrco(xr[ix], yr[iy], (1.00136e-05, 2.00001)[iz], 2.0 * ix - 1.0, 2.0 * iy - 1.0, 2.0 * iz - 1.0)
        api.setCurrentMaterial(mat)
        if with_uv:         # variant for the uv texture coordinates: one uv value per cube corner
            for k in range(8):
                api.addUv(0.25 * (k % 4) + 0.1 * (k // 4), 0.35 * (k // 2) - 0.2 * (k % 2))
        for t in TEST01_CUBE_TRIS:
            if with_uv:
                api.addTriangleWithUv(t[0], t[1], t[2], t[0], t[1], t[2])
            else:
                api.addTriangle(*t)
        api.endObject()
    _mesh(api, "Plane", "Material.002", [(-10.0, -10.0, 0.0), (10.0, -10.0, 0.0), (-10.0, 10.0, 0.0), (10.0, 10.0, 0.0)], [(0, 1, 3), (0, 3, 2)])

    api.paramsClearAll()
    api.paramsSetString("type", "perspective")
    api.paramsSetVector("from", 8.64791, -7.22615, 8.1295)
    api.paramsSetVector("to", 8.03447, -6.65603, 7.58301)
    api.paramsSetVector("up", 8.25644, -6.8447, 8.9669)
    api.paramsSetInt("resx", width)
    api.paramsSetInt("resy", height)
    api.paramsSetFloat("focal", 1.09375)
    api.createCamera("Camera")

    api.paramsClearAll()
    api.paramsSetString("camera_name", "Camera")
    api.createRenderView("view")

    api.paramsClearAll()
    api.paramsSetString("type", "constant")
    api.paramsSetColor("color", 0.7, 0.7, 0.7, 1.0)
    api.paramsSetFloat("power", 0.5)
    api.paramsSetBool("ibl", False)
    api.createBackground("world_background")

    api.paramsClearAll()
    api.paramsSetString("type", integrator)
    api.paramsSetInt("raydepth", 2)
    api.paramsSetInt("shadowDepth", 2)
    api.paramsSetBool("transpShad", False)
    api.paramsSetBool("caustics", False)
    api.paramsSetString("caustic_type", "none")
    api.paramsSetBool("do_AO", False)
    if integrator == "pathtracing":
        api.paramsSetInt("bounces", bounces)
        api.paramsSetInt("path_samples", path_samples)
        api.paramsSetInt("russian_roulette_min_bounces", bounces)
    api.createIntegrator("default")

    api.paramsClearAll()
    api.paramsSetString("type", "combined")
    api.paramsSetString("image_type", "ColorAlpha")
    api.paramsSetString("exported_image_type", "ColorAlpha")
    api.paramsSetString("exported_image_name", "Combined")
    api.defineLayer()

    api.paramsClearAll()
    api.paramsSetString("integrator_name", "default")
    api.paramsSetString("background_name", "world_background")
    api.paramsSetString("scene_accelerator", accelerator)
    api.paramsSetInt("width", width)
    api.paramsSetInt("height", height)
    api.paramsSetInt("threads", threads)
    api.paramsSetInt("AA_passes", 1)
    api.paramsSetInt("AA_minsamples", spp)
    api.paramsSetInt("AA_inc_samples", 1)
    api.paramsSetFloat("AA_pixelwidth", aa_pixelwidth)
    api.paramsSetFloat("AA_threshold", 0.05)
    api.paramsSetString("filter_type", filter_type)
    api.paramsSetString("tiles_order", "linear")
    api.paramsSetInt("tile_size", 32)
    api.setupRender()


# ---- the reference's tests/test00 scene: one cube, its texture filled pixel by pixel through the API -----------------
def test00(api, width=400, height=400, spp=1, integrator="directlighting", bounces=3, tex=200, accelerator="yafaray-kdtree-original", threads=1,
           filter_type="gauss"):
    """tests/test00/test00.c of the reference (:107-232): an image is created from a TGA file without storage optimisation,
    then every texel is overwritten with yafaray_setImageColor (HDR values up to 4); cube-mapped on orco coordinates.
    The shipped file renders with photon mapping (or the commented-out `directlighting`); here direct lighting / path tracing."""
    api.paramsClearAll()
    api.paramsSetString("type", "ColorAlpha")
    api.paramsSetString("image_optimization", "none")
    api.paramsSetInt("tex_width", tex)
    api.paramsSetInt("tex_height", tex)
    api.paramsSetString("filename", "tex.tga")
    api.createImage("Image01")
    for i in range(tex):
        for j in range(tex):
            api.setImageColor("Image01", i, j, 0.01 * i, 0.01 * j, 0.01 * (i + j), 1.0)
    api.paramsClearAll()
    api.paramsSetString("type", "image")
    api.paramsSetString("image_name", "Image01")
    api.createTexture("TextureTGA")
    api.paramsClearAll()
    api.paramsSetString("type", "shinydiffusemat")
    api.paramsSetColor("color", 0.9, 0.9, 0.9, 1.0)
    api.paramsPushList()
    _set_params(api, dict(element="shader_node", name="diff_layer0", input="map0", type="layer", blend_mode="mix", upper_color=(1.0, 1.0, 1.0, 1.0)))
    api.paramsPushList()
    _set_params(api, dict(element="shader_node", name="map0", type="texture_mapper", mapping="cube", texco="orco", texture="TextureTGA"))
    api.paramsEndList()
    api.paramsSetString("diffuse_shader", "diff_layer0")
    api.createMaterial("MaterialTGA")

    api.paramsClearAll()
    api.paramsSetBool("has_orco", True)
    api.paramsSetString("type", "mesh")
    api.createObject("Cube")
    for ix in (0, 1):
        for iy in (0, 1):
            for iz in (0, 1):
                api.addVertexWithOrco((-4.0, -2.0)[ix], (1.5, 3.5)[iy], (0.0, 2.0)[iz], 2.0 * ix - 1.0, 2.0 * iy - 1.0, 2.0 * iz - 1.0)
    api.setCurrentMaterial("MaterialTGA")
    for t in TEST01_CUBE_TRIS:
        api.addTriangle(*t)
    api.endObject()

    api.paramsClearAll()
    _set_params(api, dict(type="pointlight", color=(1.0, 1.0, 1.0, 1.0), power=150.0))
    api.paramsSetVector("from", 5.3, -4.9, 8.9)
    api.createLight("light_1")
    api.paramsClearAll()
    _set_params(api, dict(type="constant", color=(1.0, 1.0, 1.0, 1.0)))
    api.createBackground("world_background")
    api.paramsClearAll()
    _set_params(api, dict(type="perspective", resx=width, resy=height, focal=1.1))
    api.paramsSetVector("from", 8.6, -7.2, 8.1)
    api.paramsSetVector("to", 8.0, -6.7, 7.6)
    api.paramsSetVector("up", 8.3, -6.8, 9.0)
    api.createCamera("cam_1")
    api.paramsClearAll()
    api.paramsSetString("camera_name", "cam_1")
    api.createRenderView("view_1")
    api.paramsClearAll()
    api.paramsSetString("type", integrator)
    if integrator == "pathtracing":
        _set_params(api, dict(bounces=bounces, path_samples=1, russian_roulette_min_bounces=bounces, caustic_type="none"))
    api.createIntegrator("surfintegr")
    api.paramsClearAll()
    _set_params(api, dict(type="combined", image_type="ColorAlpha", exported_image_type="ColorAlpha", exported_image_name="Combined"))
    api.defineLayer()
    api.paramsClearAll()
    _set_params(api, dict(integrator_name="surfintegr", background_name="world_background", scene_accelerator=accelerator, width=width, height=height,
                          threads=threads, AA_passes=1, AA_minsamples=spp, AA_inc_samples=1, filter_type=filter_type, tiles_order="linear", tile_size=32))
    api.setupRender()


def build(api, scene="cornell", **kw):
    """Dispatch by scene name (the committed golden configurations carry a `scene` key)."""
    return {"cornell": cornell_box, "city": city, "test01": test01, "test00": test00}[scene](api, **kw)
