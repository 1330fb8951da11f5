"""Host-side mirror of libYafaRay's C public API for the hot path (scene streaming + render), bound
to the B200 back end through the C-ABI (include/y4gpu.h).

Reference interface: include/public_api/yafaray_c_api.h:53-130 (`yafaray_*` functions; here the
same names without the prefix, same argument order and meaning) -> Interface
(src/interface/interface.cc) -> Scene (src/scene/scene.cc).  The "current ParamMap" semantics of
`paramsSet*` / `paramsClearAll` / `create*` are kept (interface.cc:118-197).  What the reference
would render with `scene_accelerator = <kd-tree>` + `type = pathtracing` on CPU threads is rendered
here by the BVH8 accelerator + wavefront path tracer kernels.  Scene features outside the supported
set (shinydiffusemat with all four components and an optional `diffuse_shader` chain of layer +
texture_mapper nodes over image textures, light_mat, mirror, glass; point / area / directional / sun / sphere / spot /
object / background (IBL) / background-portal lights; the five camera types; constant / gradient / image-texture
backgrounds; pathtracing / directlighting with ambient occlusion; instances) raise Y4Error -- they are never
approximated, and there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import os

import numpy as np

from . import _lib, film_io, images
from ._lib import Y4Error
from .accelerator import ACCELERATOR_TYPE, AcceleratorB200

VIS = {"normal": (True, True), "no_shadows": (True, False), "shadow_only": (False, True), "invisible": (False, False)}  # (visible, casts shadows)
GLOBAL_SHADOW_BIAS, GLOBAL_MIN_RAYDIST = 0.0005, 0.00005  # include/common/yafaray_common.h:28-29


class _Mesh:
    def __init__(self, name, params):
        self.name = name
        self.verts: List[tuple] = []
        self.normals: List[tuple] = []
        self.tris: List[tuple] = []
        self.tri_mats: List[str] = []
        self.orco: List[tuple] = []        # addVertexWithOrco (MeshObject::orco_points_)
        self.uvs: List[tuple] = []         # addUv (MeshObject::uv_values_)
        self.tri_uvs: List[tuple] = []     # uv indices per face (addTriangleWithUv), None for addTriangle
        self.smooth_angle: Optional[float] = None      # smoothMesh
        self.visibility = params.get("visibility", "normal")
        self.is_base = bool(params.get("is_base_object", False))
        self.instance_of: Optional[str] = None
        self.matrix = None


MAX_TWO_LEVEL_ANISOTROPY = 8.0      # largest ratio of an instance matrix's singular values the automatic two-level upload accepts


def _anisotropy(o) -> float:
    """Largest / smallest singular value of an instance's 3 x 3 matrix (inf when singular), computed once per matrix."""
    key = o.matrix.tobytes()
    if getattr(o, "_aniso_key", None) != key:
        sv = np.linalg.svd(np.asarray(o.matrix, np.float64)[:3, :3], compute_uv=False)
        o._aniso, o._aniso_key = (float(sv[0] / sv[2]) if sv[2] > 0.0 else float("inf")), key
    return o._aniso


class Interface:
    """`yafaray_Interface_t` for the GPU back end: build a scene with the C API's calls, then render()."""

    def __init__(self, device: int = 0, shard_index: int = 0, shard_count: int = 1, comm=None):
        """comm: multi-GPU rendering (one process per GPU): an object with `all_reduce(tensor)` summing a CUDA tensor over the
        shard_count ranks in place -- `torch.distributed` after init_process_group("nccl") is one.  With a comm, render() /
        render_pass_reduced() combine the ranks' pass films with ONE all-reduce per AA pass (the linear film merge of
        imagefilm.cc:940-1018) and every rank ends up with the complete film; see Interface.distributed()."""
        self.device, self.shard_index, self.shard_count = device, shard_index, shard_count
        self.comm = comm
        self._pass_film = self._total_film = None   # sharded mode: [weights | rgba] CUDA tensors (this pass / all passes, all ranks)
        self.params: Dict[str, object] = {}
        self.eparams: List[Dict[str, object]] = []     # the ParamMap list of paramsPushList (shader nodes)
        self._cparams = self.params
        self.images: Dict[str, np.ndarray] = {}        # createImage: decoded texels [h, w, 4]
        self._image_storage: Dict[str, tuple] = {}     # (has alpha channel, optimisation) of each image buffer
        self.textures: Dict[str, dict] = {}
        self.image_dir = "."                           # relative image file names resolve against this directory
        self.warnings: List[str] = []
        self.materials: Dict[str, dict] = {}
        self.lights: Dict[str, dict] = {}
        self.cameras: Dict[str, dict] = {}
        self.backgrounds: Dict[str, dict] = {}
        self.integrators: Dict[str, dict] = {}
        self.views: Dict[str, dict] = {}
        self.objects: Dict[str, _Mesh] = {}
        self.render_params: Dict[str, object] = {}
        self._cur: Optional[_Mesh] = None
        self._cur_mat: Optional[str] = None
        self._next_id = 0
        self.accelerator: Optional[AcceleratorB200] = None
        # instanced objects (addInstance): "auto" / True keep the base meshes once on the device (two-level BVH8, results identical
        # to the flattened scene) whenever the scene has instances; False always flattens like the reference's scene does.
        # Env Y4GPU_INSTANCING=0/1 overrides.
        self.device_instancing = "auto"
        self.film_buffers = None  # optional (weights, rgba) torch CUDA tensors installed by the caller
        self._pinned_film = None
        self.last_stats: dict = {}

    @classmethod
    def distributed(cls):
        """One Interface per rank of an initialised torch.distributed NCCL job (torchrun): GPU = LOCAL_RANK, shard = RANK of
        WORLD_SIZE.  Every rank builds the same scene (the scene and its BVH8 are replicated)."""
        import os

        import torch.distributed as dist

        if not dist.is_initialized():
            raise Y4Error("Interface.distributed() needs torch.distributed.init_process_group('nccl') first")
        return cls(device=int(os.environ.get("LOCAL_RANK", "0")), shard_index=dist.get_rank(), shard_count=dist.get_world_size(), comm=dist)

    # ---- the "current ParamMap" (interface.cc:118-181) -------------------------------------------
    def paramsClearAll(self):
        self.params = {}
        self.eparams = []
        self._cparams = self.params

    def paramsSetString(self, name, s):
        self._cparams[name] = str(s)

    def paramsSetBool(self, name, b):
        self._cparams[name] = bool(b)

    def paramsSetInt(self, name, i):
        self._cparams[name] = int(i)

    def paramsSetFloat(self, name, f):
        self._cparams[name] = float(f)

    def paramsSetVector(self, name, x, y, z):
        self._cparams[name] = (float(x), float(y), float(z))

    def paramsSetColor(self, name, r, g, b, a=1.0):
        self._cparams[name] = (float(r), float(g), float(b), float(a))

    def paramsPushList(self):
        """Interface::paramsPushList (interface.cc:183-187): later paramsSet* calls fill a new ParamMap of the node list."""
        self.eparams.append({})
        self._cparams = self.eparams[-1]

    def paramsEndList(self):
        self._cparams = self.params

    # ---- object factories (scene.cc:363-526) ------------------------------------------------------
    def createImage(self, name):
        """Scene::createImage -> Image::factory (image.cc:38-101).  A file the reference build cannot load leaves no
        image behind (False); textures and node materials that need it are then dropped like in the reference."""
        p = dict(self.params)
        try:
            if p.get("filename", ""):
                self.images[name] = images.load_image(p, self.image_dir)
            else:           # image in memory, filled with setImageColor (Image::factory(width, height, type, optimization), image.cc:103-140)
                self.images[name] = images.empty_image(p)
        except images.ImageLoadError as exc:
            self.warnings.append(f"image '{name}': {exc}")
            return False
        self._image_storage[name] = images.storage_of(p, self.images[name])
        return True

    def setImageColor(self, name, x, y, r, g, b, a):
        """yafaray_setImageColor -> Image::setColor(x, y, col): the colour goes through the image buffer's storage rounding."""
        img = self.images.get(name)
        if img is None or not (0 <= x < img.shape[1] and 0 <= y < img.shape[0]):
            return False
        has_alpha, optimization = self._image_storage[name]
        img[y, x] = images.store(np.array([[r, g, b, a]], np.float32), has_alpha, optimization)[0]
        return True

    def getImageColor(self, name, x, y):
        img = self.images.get(name)
        if img is None or not (0 <= x < img.shape[1] and 0 <= y < img.shape[0]):
            return None
        return tuple(float(c) for c in img[y, x])

    def createTexture(self, name):
        """Scene::createTexture -> ImageTexture::factory (texture_image.cc:476-597)."""
        p = dict(self.params)
        if p.get("type") != "image":
            raise Y4Error(f"texture type '{p.get('type')}' is not supported by the GPU back end (image textures only)")
        if p.get("image_name", "") not in self.images:
            self.warnings.append(f"texture '{name}': image '{p.get('image_name', '')}' does not exist, dropping texture")
            return False
        interp = str(p.get("interpolate", ""))
        if interp in ("bicubic", "mipmap_trilinear", "mipmap_ewa"):
            raise Y4Error(f"texture '{name}': interpolate '{interp}' is not supported by the GPU back end (none / bilinear)")
        if float(p.get("adj_saturation", 1.0)) != 1.0 or float(p.get("adj_hue", 0.0)) != 0.0:
            raise Y4Error(f"texture '{name}': hue / saturation adjustments are not supported by the GPU back end")
        if p.get("normalmap", False):
            raise Y4Error(f"texture '{name}': normal maps are not supported by the GPU back end")
        self.textures[name] = p
        return True

    def _diffuse_chain(self, name, p, node_params):
        """NodeMaterial::loadNodes / parseNodes (material_node.cc:102-187) for the `diffuse_shader` slot -> the chain of
        (layer params, texture_mapper params) pairs, bottom layer first; None when the reference would drop the nodes."""
        nodes = {}
        for n in node_params:
            if n.get("element", "shader_node") != "shader_node":
                continue
            if "name" not in n or n["name"] in nodes or "type" not in n:
                self.warnings.append(f"material '{name}': unnamed / duplicate / untyped shader node, nodes dropped")
                return None
            if n["type"] not in ("layer", "texture_mapper"):
                raise Y4Error(f"material '{name}': shader node type '{n['type']}' is not supported by the GPU back end")
            if n["type"] == "texture_mapper" and n.get("texture", "") not in self.textures:
                self.warnings.append(f"material '{name}': texture '{n.get('texture', '')}' does not exist, nodes dropped")
                return None           # TextureMapperNode::factory returns nullptr -> loadNodes clears the table
            nodes[n["name"]] = n
        for n in nodes.values():       # configInputs (shader_node_layer.cc:156-190)
            if n["type"] == "layer" and (n.get("input") not in nodes or ("upper_layer" in n and n["upper_layer"] not in nodes)):
                self.warnings.append(f"material '{name}': layer node '{n['name']}' has no input, nodes dropped")
                return None
        root = nodes.get(p["diffuse_shader"])
        if root is None:
            self.warnings.append(f"material '{name}': shader node '{p['diffuse_shader']}' does not exist")
            return None
        chain, n, seen = [], root, set()
        while n is not None:
            if n["type"] != "layer" or n["name"] in seen:
                raise Y4Error(f"material '{name}': the diffuse_shader must be a stack of layer nodes")
            seen.add(n["name"])
            mapper = nodes[n["input"]]
            if mapper["type"] != "texture_mapper":
                raise Y4Error(f"material '{name}': layer inputs other than texture_mapper nodes are not supported by the GPU back end")
            if not n.get("color_input", True):
                raise Y4Error(f"material '{name}': scalar layer inputs are not supported by the GPU back end")
            if mapper.get("texco", "global") in ("transformed", "window", "normal"):
                raise Y4Error(f"material '{name}': texco '{mapper.get('texco')}' is not supported by the GPU back end")
            if mapper.get("mapping", "plain") in ("tube", "sphere"):
                raise Y4Error(f"material '{name}': mapping '{mapper.get('mapping')}' is not supported by the GPU back end")
            chain.insert(0, (n, mapper))
            n = nodes[n["upper_layer"]] if "upper_layer" in n else None
        return chain

    def createMaterial(self, name):
        p = dict(self.params)
        t = p.get("type")
        if t == "shinydiffusemat":
            if float(p.get("transparentbias_factor", 0.0)) != 0.0:
                raise Y4Error(f"shinydiffusemat '{name}': transparentbias_factor is not supported by the GPU back end")
            if float(p.get("samplingfactor", 1.0)) != 1.0 or int(p.get("additionaldepth", 0)) != 0:
                raise Y4Error(f"shinydiffusemat '{name}': samplingfactor / additionaldepth are not supported by the GPU back end")
            if float(p.get("wireframe_amount", 0.0)) != 0.0:
                raise Y4Error(f"shinydiffusemat '{name}': wireframe shading is not supported by the GPU back end")
            for k in p:
                if k.endswith("_shader") and k != "diffuse_shader":
                    raise Y4Error(f"shinydiffusemat '{name}': shader slot '{k}' is not supported by the GPU back end (diffuse_shader only)")
            if "diffuse_shader" in p:
                p["_diffuse_chain"] = self._diffuse_chain(name, p, [dict(e) for e in self.eparams])
        elif t == "mirror":            # MirrorMaterial::factory reads `color` and `reflect` only (material_glass.cc:453-461)
            p = {"type": "mirror", "color": p.get("color", (1.0, 1.0, 1.0, 1.0)), "reflect": p.get("reflect", 1.0)}
        elif t == "glass":             # GlassMaterial::factory (material_glass.cc:290-433)
            if float(p.get("dispersion_power", 0.0)) > 0.0:
                raise Y4Error(f"glass '{name}': dispersion is not supported by the GPU back end")
            if p.get("fake_shadows", False):
                raise Y4Error(f"glass '{name}': fake_shadows is not supported by the GPU back end")
            if min(p.get("absorption", (1.0, 1.0, 1.0, 1.0))[:3]) <= 1e-38:
                raise Y4Error(f"glass '{name}': an absorption colour with a zero component is not supported by the GPU back end")
            if float(p.get("wireframe_amount", 0.0)) != 0.0 or float(p.get("samplingfactor", 1.0)) != 1.0 or int(p.get("additionaldepth", 0)) != 0:
                raise Y4Error(f"glass '{name}': wireframe / samplingfactor / additionaldepth are not supported by the GPU back end")
            if any(k.endswith("_shader") for k in p):
                raise Y4Error(f"glass '{name}': shader nodes are not supported by the GPU back end")
        elif t != "light_mat":
            raise Y4Error(f"material type '{t}' is not supported by the GPU back end")
        self.materials[name] = p
        return True

    def createLight(self, name):
        p = dict(self.params)
        if p.get("type") not in ("pointlight", "arealight", "directional", "sunlight", "spherelight", "spotlight", "objectlight", "meshlight", "bgPortalLight"):
            raise Y4Error(f"light type '{p.get('type')}' is not supported by the GPU back end")
        self.lights[name] = p
        return True

    def createCamera(self, name):
        p = dict(self.params)
        if p.get("type") not in _lib.CAMERA_TYPES:
            raise Y4Error(f"camera type '{p.get('type')}' is not supported by the GPU back end")
        self.cameras[name] = p
        return True

    def createBackground(self, name):
        p = dict(self.params)
        if p.get("type") not in ("constant", "gradientback", "textureback", "sunsky", "darksky"):
            raise Y4Error(f"background type '{p.get('type')}' does not exist")
        if p["type"] in ("sunsky", "darksky"):         # SunSkyBackground::factory (background_sunsky.cc:226-307), DarkSkyBackground::factory (background_darksky.cc:194-314)
            if p.get("add_sun", False):
                raise Y4Error(f"{p['type']}: add_sun (a sunlight coloured by the reference's solar spectrum tables) is not supported by the GPU back end; "
                              "create the sunlight explicitly")
            self.backgrounds[name] = p
            lname = "sunsky_bgLight" if p["type"] == "sunsky" else "DarkSky_bgLight"
            if p.get("background_light", False) and lname not in self.lights:
                self.lights[lname] = dict(type="bglight", samples=int(p.get("light_samples", 8)), cast_shadows=bool(p.get("cast_shadows", True)), _background=name)
            return True
        if p["type"] == "textureback":                 # TextureBackground::factory (background_texture.cc:80-151)
            if p.get("texture") not in self.textures:
                raise Y4Error(f"textureback '{name}': texture '{p.get('texture')}' does not exist")
            if float(p.get("smartibl_blur", 0.0)) > 0.0:
                raise Y4Error("textureback: smartibl_blur (mip-mapped IBL lookups) is not supported by the GPU back end")
        self.backgrounds[name] = p
        if p.get("ibl", False):
            # the background factories create the IBL light themselves, under a fixed name (background_constant.cc:58-69,
            # background_gradient.cc:73-84, background_texture.cc:119-139); a second light of that name is refused by the scene
            lname = {"constant": "constantBackground_bgLight", "gradientback": "GradientBackground_bgLight", "textureback": "textureBackground_bgLight"}[p["type"]]
            if lname not in self.lights:
                self.lights[lname] = dict(type="bglight", samples=int(p.get("ibl_samples", 16)), cast_shadows=bool(p.get("cast_shadows", True)),
                                          _background=name)
        return True

    def createIntegrator(self, name):
        p = dict(self.params)
        if p.get("type") not in ("pathtracing", "directlighting"):
            raise Y4Error(f"integrator type '{p.get('type')}' is not supported by the GPU back end")
        if p.get("caustics", False):
            raise Y4Error("photon caustics are not supported by the GPU back end")
        if p.get("caustic_type", "path") in ("photon", "both"):
            raise Y4Error("photon caustics are not supported by the GPU back end")
        if p.get("do_AO", False) and int(p.get("AO_samples", 32)) < 1:
            raise Y4Error("do_AO with AO_samples < 1 (the reference divides by zero) is not supported by the GPU back end")
        self.integrators[name] = p
        return True

    def createRenderView(self, name):
        self.views[name] = dict(self.params)
        return True

    def createOutput(self, name):
        return True  # image files are written by the host application from the film (out of scope)

    def defineLayer(self):
        if self.params.get("type", "combined") != "combined":
            raise Y4Error("only the Combined layer is produced by the GPU back end")

    # ---- geometry streaming (scene.cc:900-1030) ---------------------------------------------------
    def startGeometry(self):
        return True

    def endGeometry(self):
        return True

    def getNextFreeId(self):
        self._next_id += 1
        return self._next_id

    def createObject(self, name):
        p = dict(self.params)
        if p.get("type") != "mesh":
            raise Y4Error(f"object type '{p.get('type')}' is not supported by the GPU back end")
        self._cur = _Mesh(name, p)
        self.objects[name] = self._cur
        return True

    def setCurrentMaterial(self, name):
        if name not in self.materials:
            raise Y4Error(f"unknown material '{name}'")
        self._cur_mat = name

    def addVertex(self, x, y, z):
        self._cur.verts.append((x, y, z))
        return len(self._cur.verts) - 1

    def addVertexWithOrco(self, x, y, z, ox, oy, oz):
        self._cur.orco.append((ox, oy, oz))
        return self.addVertex(x, y, z)

    def addUv(self, u, v):
        self._cur.uvs.append((u, v))
        return len(self._cur.uvs) - 1

    def addTriangleWithUv(self, a, b, c, uv_a, uv_b, uv_c):
        self.addTriangle(a, b, c)
        self._cur.tri_uvs[-1] = (uv_a, uv_b, uv_c)
        return True

    def addNormal(self, x, y, z):
        self._cur.normals.append((x, y, z))

    def addTriangle(self, a, b, c):
        self._cur.tris.append((a, b, c))
        self._cur.tri_mats.append(self._cur_mat)
        self._cur.tri_uvs.append(None)
        return True

    def add_mesh_arrays(self, verts: np.ndarray, tris: np.ndarray):
        """Bulk form of addVertex/addTriangle for large procedural meshes (same semantics)."""
        base = len(self._cur.verts)
        self._cur.verts.extend(map(tuple, np.asarray(verts, np.float64).reshape(-1, 3)))
        for t in np.asarray(tris, np.int64).reshape(-1, 3):
            self._cur.tris.append((base + int(t[0]), base + int(t[1]), base + int(t[2])))
            self._cur.tri_mats.append(self._cur_mat)
            self._cur.tri_uvs.append(None)

    def endObject(self):
        self._cur = None
        return True

    def smoothMesh(self, name, angle):
        """Scene::smoothNormals (scene.cc:909-931): vertex normals from the face normals, MeshObject::smoothNormals."""
        mesh = self.objects.get(name) if name else self._cur
        if mesh is None:
            return False
        if not (mesh.normals and len(mesh.normals) == len(mesh.verts)):      # exported normals win
            mesh.smooth_angle = float(np.float32(angle))
        return True

    def addInstance(self, base_object_name, *m):
        if base_object_name not in self.objects:
            raise Y4Error(f"base mesh '{base_object_name}' for instance does not exist")
        inst = _Mesh(f"{base_object_name}-{self.getNextFreeId()}", {})
        inst.instance_of = base_object_name
        inst.matrix = np.array(m, np.float32).reshape(4, 4)
        self.objects[inst.name] = inst
        return True

    # ---- render set-up (scene.cc:528-644) ------------------------------------------------------------
    def setupRender(self):
        self.render_params = dict(self.params)
        acc = self.render_params.get("scene_accelerator", ACCELERATOR_TYPE)
        if acc != ACCELERATOR_TYPE:
            raise Y4Error(f"scene_accelerator '{acc}' is not provided by this back end (use '{ACCELERATOR_TYPE}')")

    # ---- scene assembly: what Scene::updateObjects + the factories hand to the back end --------------
    def assemble(self, flatten: bool = True):
        """-> dict(tris [n,9] f32, flags [n] u32, normals [n,9] f32 | None, cam, mats, lights, params) in the
        reference's orders: objects by name (std::map, scene.cc:1034), faces in insertion order, lights by name."""
        mat_names = list(self.materials)
        mat_index = {n: i for i, n in enumerate(mat_names)}
        tris, flags, normals, any_normals = [], [], [], False
        meshes, mesh_flags, mesh_index, inst_list, n_prims = [], [], {}, [], 0     # the two-level description of the same scene
        textured = any(m.get("_diffuse_chain") for m in self.materials.values())
        orcos, uvs = [], []
        base_cache: Dict[str, tuple] = {}      # per mesh: fp32 vertices, index array, flag words, normals (shared by its instances)

        def mesh_arrays(base: _Mesh):
            if base.name in base_cache:
                return base_cache[base.name]
            vis_o, sh_o = VIS.get(base.visibility, VIS["normal"])
            if base.name in portal_meshes:     # BackgroundPortalLight::init: base_object_->setVisibility(Visibility::Invisible) (light_background_portal.cc:79)
                vis_o, sh_o = False, False
            v = np.asarray(base.verts, np.float64).astype(np.float32)  # API doubles are narrowed to float (interface.cc:87)
            idx = np.asarray(base.tris, np.int64)
            words = {}
            for mn in set(base.tri_mats):
                if mn is None:
                    raise Y4Error(f"object '{base.name}': triangle without a material")
                vis_m, sh_m = VIS.get(self.materials[mn].get("visibility", "normal"), VIS["normal"])
                w = (mat_index[mn] << _lib.TRI_MAT_SHIFT)
                if vis_o and vis_m:
                    w |= _lib.TRI_VIS_CLOSEST
                if sh_o and sh_m:
                    w |= _lib.TRI_VIS_SHADOW
                if sh_m:
                    w |= _lib.TRI_VIS_TSHADOW
                if float(self.materials[mn].get("transparency", 0.0)) > 0.00001 and self.materials[mn].get("type") == "shinydiffusemat":
                    w |= _lib.TRI_TRANSPARENT          # Material::isTransparent(): bsdf_flags_ has Transmit (intersectTs filters instead of blocking)
                words[mn] = w
            f = np.fromiter((words[mn] for mn in base.tri_mats), np.uint32, len(base.tri_mats))
            nn = None
            if base.normals and len(base.normals) == len(base.verts):
                nn = np.asarray(base.normals, np.float64).astype(np.float32)[idx].reshape(-1, 9)
            elif base.smooth_angle is not None and base.smooth_angle > 0.1:
                nn = smooth_normals(v, idx, base.smooth_angle)
            oc = uv = None
            if textured:
                if base.orco and len(base.orco) == len(base.verts):      # MeshObject::hasOrco
                    oc = np.asarray(base.orco, np.float64).astype(np.float32)[idx].reshape(-1, 9)
                if base.uvs and any(t is not None for t in base.tri_uvs):
                    uva = np.asarray(base.uvs, np.float64).astype(np.float32)
                    if any(t is None for t in base.tri_uvs):
                        raise Y4Error(f"object '{base.name}': faces with and without uv indices in one mesh are not supported")
                    uv = uva[np.asarray(base.tri_uvs, np.int64)].reshape(-1, 6)
            base_cache[base.name] = (v, idx, f, nn, oc, uv)
            return base_cache[base.name]

        # Objects declared invisible never reach the accelerator (scene.cc:1037).  The mesh of a portal light is different: the light
        # makes it invisible in its init(), AFTER the accelerator was built (scene.cc:218-220), so its triangles are in the tree --
        # camera, bounce and plain shadow rays skip them (object visibility), but with transpShad the reference's intersectTs looks at
        # the MATERIAL's visibility only (accelerator_kdtree.cc:1022) and they block (or filter) those shadow rays
        portal_meshes = {str(p.get("object_name", "")) for p in self.lights.values() if p.get("type") == "bgPortalLight"}
        for name in sorted(self.objects):
            o = self.objects[name]
            base = self.objects[o.instance_of] if o.instance_of else o
            vis_o, sh_o = VIS.get(base.visibility, VIS["normal"])
            if (not vis_o and not sh_o) or o.is_base:
                continue
            if not base.tris:
                continue
            v, idx, f, nn, oc, uv = mesh_arrays(base)
            if textured:      # NaN in the first float: no orco points / uv values on this mesh (see include/y4gpu.h)
                orcos.append(oc if oc is not None else np.full((len(idx), 9), np.nan, np.float32))
                uvs.append(uv if uv is not None else np.full((len(idx), 6), np.nan, np.float32))
            if base.name not in mesh_index:
                mesh_index[base.name] = len(meshes)
                meshes.append(np.ascontiguousarray(v[idx].reshape(-1, 9), np.float32))
                mesh_flags.append(f)
            inst_list.append((mesh_index[base.name], np.asarray(o.matrix, np.float32)[:3, :4].copy() if o.instance_of else None))
            n_prims += len(idx)
            if flatten:
                if o.instance_of:  # Matrix4 * Point3 in fp32, matrix4.h:85-90
                    m = o.matrix
                    v = np.stack([((m[r, 0] * v[:, 0] + m[r, 1] * v[:, 1]) + m[r, 2] * v[:, 2]) + m[r, 3] for r in range(3)], -1).astype(np.float32)
                tris.append(v[idx].reshape(-1, 9))
            if nn is not None and not o.instance_of:
                normals.append(nn)
                any_normals = True
            else:           # flat-shaded mesh in a scene that has smooth ones: NaN marks "shading normal = face normal" (include/y4gpu.h)
                normals.append(len(idx))          # (materialised below, and only when some mesh is smooth)
            if flatten:
                flags.append(f)
        if not flatten:
            T, F = None, None
        elif tris:
            T = np.ascontiguousarray(np.concatenate(tris), np.float32)
            F = np.ascontiguousarray(np.concatenate(flags), np.uint32)
        else:
            T, F = np.zeros((0, 9), np.float32), np.zeros(0, np.uint32)
        N = None
        if any_normals:
            N = np.ascontiguousarray(np.concatenate([np.full((x, 9), np.nan, np.float32) if isinstance(x, int) else x for x in normals]), np.float32)

        rp = self.render_params
        integ = self.integrators[rp["integrator_name"]] if "integrator_name" in rp else next(iter(self.integrators.values()))
        bg = self.backgrounds.get(rp.get("background_name", ""), None)
        view = next(iter(self.views.values())) if self.views else {}
        cam_p = self.cameras[view["camera_name"]] if "camera_name" in view else next(iter(self.cameras.values()))

        cam = _lib.PtCamera()
        cam.from_[:] = cam_p.get("from", (0.0, 1.0, 0.0))
        cam.to[:] = cam_p.get("to", (0.0, 0.0, 0.0))
        cam.up[:] = cam_p.get("up", (0.0, 1.0, 1.0))
        cam.resx, cam.resy = int(cam_p.get("resx", 320)), int(cam_p.get("resy", 200))
        cam.type = _lib.CAMERA_TYPES[cam_p.get("type", "perspective")]
        cam.aspect_ratio, cam.focal = float(cam_p.get("aspect_ratio", 1.0)), float(cam_p.get("focal", 1.0))
        if cam.type == _lib.CAMERA_TYPES["orthographic"]:
            cam.focal = float(cam_p.get("scale", 1.0))
        if cam.type in (_lib.CAMERA_TYPES["perspective"], _lib.CAMERA_TYPES["architect"]):
            cam.aperture, cam.dof_distance = float(cam_p.get("aperture", 0.0)), float(cam_p.get("dof_distance", 0.0))
        # the factories' own defaults (camera_equirectangular.cc:70: farClip -1e38)
        cam.near_clip = float(cam_p.get("nearClip", 0.0))
        cam.far_clip = float(cam_p.get("farClip", -1.0e38 if cam.type in (_lib.CAMERA_TYPES["equirectangular"], _lib.CAMERA_TYPES["angular"]) else -1.0))
        if cam.type == _lib.CAMERA_TYPES["angular"]:      # AngularCamera::factory (camera_angular.cc:80-117); unknown projection names: equidistant
            cam.angle_degrees = float(cam_p.get("angle", 90.0))
            cam.max_angle_degrees = float(cam_p.get("max_angle", cam.angle_degrees))
            cam.circular, cam.mirrored = int(bool(cam_p.get("circular", True))), int(bool(cam_p.get("mirrored", False)))
            cam.projection = _lib.ANGULAR_PROJECTIONS.get(str(cam_p.get("projection", "")), 0)
        # unknown names select the defaults, like PerspectiveCamera::factory (camera_perspective.cc:207-218)
        cam.bokeh_type = _lib.BOKEH_TYPES.get(str(cam_p.get("bokeh_type", "disk1")), 0)
        cam.bokeh_bias = _lib.BOKEH_BIAS.get(str(cam_p.get("bokeh_bias", "uniform")), 0)
        cam.bokeh_rotation = float(cam_p.get("bokeh_rotation", 0.0))

        # textures + layer chains of the diffuse_shader slots (only the textures some chain uses are uploaded)
        tex_index: Dict[str, int] = {}
        tex_list, layer_list, chain_of = [], [], {}
        for n in mat_names:
            chain = self.materials[n].get("_diffuse_chain")
            if not chain:
                continue
            chain_of[n] = (len(layer_list), len(chain))
            for li, (lay, mapper) in enumerate(chain):
                tname = mapper["texture"]
                if tname not in tex_index:
                    tex_index[tname] = len(tex_list)
                    tex_list.append(tname)
                L = _lib.PtLayer()
                L.texture = tex_index[tname]
                L.texco = _lib.TEXCO.get(str(mapper.get("texco", "global")), 0)
                # the projection applies to discrete (image) textures only; unknown names leave Plain (shader_node_basic.cc:343-349)
                L.mapping = _lib.PROJECTIONS.get(str(mapper.get("mapping", "plain")), 0)
                L.proj_x, L.proj_y, L.proj_z = int(mapper.get("proj_x", 1)), int(mapper.get("proj_y", 2)), int(mapper.get("proj_z", 3))
                L.scale[:] = mapper.get("scale", (1.0, 1.0, 1.0))
                L.offset[:] = mapper.get("offset", (0.0, 0.0, 0.0))
                L.flags = (_lib.LAYER_RGB_TO_INT if lay.get("noRGB", False) else 0) | (_lib.LAYER_STENCIL if lay.get("stencil", False) else 0) \
                    | (_lib.LAYER_NEGATIVE if lay.get("negative", False) else 0)
                L.blend_mode = _lib.BLEND_MODES.get(str(lay.get("blend_mode", "mix")), 0)
                L.do_color, L.colfac = int(bool(lay.get("do_color", True))), float(lay.get("colfac", 1.0))
                L.def_col[:] = lay.get("def_col", (1.0, 1.0, 1.0, 1.0))[:3]
                if li == 0:
                    L.upper_color[:] = lay.get("upper_color", (0.0, 0.0, 0.0, 0.0))
                layer_list.append(L)
        if bg is not None and bg.get("type") == "textureback" and bg["texture"] not in tex_index:
            tex_index[bg["texture"]] = len(tex_list)
            tex_list.append(bg["texture"])
        textures = (_lib.PtTexture * max(1, len(tex_list)))()
        texel_arrays = []
        for i, tname in enumerate(tex_list):
            tp, TS = self.textures[tname], textures[i]
            tx = self.images[tp["image_name"]]
            texel_arrays.append(tx)
            TS.texels, TS.height, TS.width = tx.ctypes.data, tx.shape[0], tx.shape[1]
            TS.clip = _lib.CLIP_MODES.get(str(tp.get("clipping", "")), _lib.CLIP_MODES["repeat"])
            TS.interpolate = _lib.INTERPOLATIONS.get(str(tp.get("interpolate", "")), _lib.INTERPOLATIONS["bilinear"])
            TS.xrepeat, TS.yrepeat = int(tp.get("xrepeat", 1)), int(tp.get("yrepeat", 1))
            TS.mirror_x, TS.mirror_y, TS.rot90 = int(bool(tp.get("mirror_x", False))), int(bool(tp.get("mirror_y", False))), int(bool(tp.get("rot90", False)))
            TS.cropmin_x, TS.cropmin_y = float(tp.get("cropmin_x", 0.0)), float(tp.get("cropmin_y", 0.0))
            TS.cropmax_x, TS.cropmax_y = float(tp.get("cropmax_x", 1.0)), float(tp.get("cropmax_y", 1.0))
            TS.even_tiles, TS.odd_tiles, TS.checker_dist = int(bool(tp.get("even_tiles", False))), int(bool(tp.get("odd_tiles", True))), float(tp.get("checker_dist", 0.0))
            TS.adj_intensity, TS.adj_contrast = float(tp.get("adj_intensity", 1.0)), float(tp.get("adj_contrast", 1.0))
            TS.adj_mult_factor_red, TS.adj_mult_factor_green, TS.adj_mult_factor_blue = (float(tp.get("adj_mult_factor_" + c, 1.0)) for c in ("red", "green", "blue"))
            TS.adj_clamp = int(bool(tp.get("adj_clamp", False)))
        layers = (_lib.PtLayer * max(1, len(layer_list)))(*layer_list)

        mats = (_lib.PtMaterial * max(1, len(mat_names)))()
        for i, n in enumerate(mat_names):
            p, m = self.materials[n], mats[i]
            if n in chain_of:
                m.layer_first, m.layer_count = chain_of[n]
            m.color[:] = p.get("color", (1.0, 1.0, 1.0, 1.0))[:3]
            m.receive_shadows, m.flat_material = int(p.get("receive_shadows", True)), int(p.get("flat_material", False))
            if p["type"] == "light_mat":
                m.type, m.emit, m.double_sided = _lib.MAT_LIGHT, float(p.get("power", 1.0)), int(p.get("double_sided", False))
            elif p["type"] == "mirror":
                m.type, m.diffuse_reflect = _lib.MAT_MIRROR, float(p.get("reflect", 1.0))
            elif p["type"] == "glass":
                m.type, m.ior = _lib.MAT_GLASS, float(p.get("IOR", 1.4))
                # filt * filt_col + Rgb(1.f - filt) with `filt` a double parameter (material_glass.cc:326): float products, the offset rounded once
                filt = float(p.get("transmit_filter", 0.0))
                fc = p.get("filter_color", (1.0, 1.0, 1.0, 1.0))
                m.color[:] = [float(np.float32(np.float32(filt) * np.float32(fc[k])) + np.float32(1.0 - filt)) for k in range(3)]
                m.mirror_color[:] = p.get("mirror_color", (1.0, 1.0, 1.0, 1.0))[:3]
                ab = p.get("absorption", (1.0, 1.0, 1.0, 1.0))[:3]       # material_glass.cc:339-365: a Beer-Lambert handler for the inside
                if min(np.float32(ab[k]) for k in range(3)) < 1.0:
                    m.absorb, m.absorption_dist = 1, float(p.get("absorption_dist", 1.0))
                    m.absorption[:] = [float(x) for x in ab]
            else:
                m.type, m.diffuse_reflect, m.emit = _lib.MAT_SHINYDIFFUSE, float(p.get("diffuse_reflect", 1.0)), float(p.get("emit", 0.0))
                m.specular_reflect, m.fresnel_effect, m.ior = float(p.get("specular_reflect", 0.0)), int(p.get("fresnel_effect", False)), float(p.get("IOR", 1.33))
                m.mirror_color[:] = p.get("mirror_color", (1.0, 1.0, 1.0, 1.0))[:3]
                m.transparency, m.transmit_filter = float(p.get("transparency", 0.0)), float(p.get("transmit_filter", 1.0))
                m.translucency = float(p.get("translucency", 0.0))
                if p.get("diffuse_brdf", "lambert") == "oren_nayar":
                    m.oren_nayar, m.sigma = 1, float(p.get("sigma", 0.1))

        wanted = [s for s in str(view.get("light_names", "")).split(";") if s]
        light_names = [n for n in sorted(self.lights) if self.lights[n].get("light_enabled", True) and (not wanted or n in wanted)]
        lights = (_lib.PtLight * max(1, len(light_names)))()
        self._light_tris = []
        for i, n in enumerate(light_names):
            p, l = self.lights[n], lights[i]
            if p.get("photon_only", False):
                raise Y4Error("photon_only lights are not supported by the GPU back end")
            l.color[:] = p.get("color", (1.0, 1.0, 1.0, 1.0))[:3]
            l.power, l.cast_shadows = float(p.get("power", 1.0)), int(p.get("cast_shadows", True))
            if p["type"] == "pointlight":
                l.type = _lib.LIGHT_POINT
                l.from_[:] = p.get("from", (0.0, 0.0, 0.0))
            elif p["type"] == "directional":       # DirectionalLight::factory (light_directional.cc:116-157)
                l.type, l.infinite = _lib.LIGHT_DIRECTIONAL, int(bool(p.get("infinite", True)))
                l.point1[:] = p.get("direction", (0.0, 0.0, 1.0))
                l.radius = 1.0
                if not l.infinite:
                    l.from_[:] = p.get("from", p.get("position", (0.0, 0.0, 0.0)))
                    l.radius = float(p.get("radius", 1.0))
            elif p["type"] == "sunlight":          # SunLight::factory (light_sun.cc:97-128)
                l.type, l.samples, l.angle = _lib.LIGHT_SUN, int(p.get("samples", 4)), float(p.get("angle", 0.27))
                l.point1[:] = p.get("direction", (0.0, 0.0, 1.0))
            elif p["type"] == "spotlight":         # SpotLight::factory (light_spot.cc:250-288), soft_shadows = false
                l.type, l.angle, l.blend = _lib.LIGHT_SPOT, float(p.get("cone_angle", 45.0)), float(p.get("blend", 0.15))
                if p.get("soft_shadows", False):       # a sampled light: SpotLight::illumSample / intersect
                    l.type, l.samples, l.radius = _lib.LIGHT_SPOT_SOFT, int(p.get("samples", 8)), float(p.get("shadowFuzzyness", 1.0))
                l.from_[:] = p.get("from", (0.0, 0.0, 0.0))
                l.point1[:] = p.get("to", (0.0, 0.0, -1.0))
            elif p["type"] in ("objectlight", "meshlight", "bgPortalLight"):     # ObjectLight::factory (light_object_light.cc:222-253)
                o = self.objects.get(str(p.get("object_name", "")))
                if o is None or o.instance_of or not o.tris:
                    raise Y4Error(f"{p['type']} '{n}': object_name must name a mesh with triangles")
                ov = np.asarray(o.verts, np.float64).astype(np.float32)[np.asarray(o.tris, np.int64)].reshape(-1, 9)
                self._light_tris.append(np.ascontiguousarray(ov))          # kept alive until the scene is uploaded
                l.type, l.samples, l.double_sided = _lib.LIGHT_OBJECT, int(p.get("samples", 4)), int(p.get("double_sided", False))
                if p["type"] == "bgPortalLight":     # BackgroundPortalLight::factory (light_background_portal.cc:207-237): the background seen through the mesh
                    if bg is None:
                        raise Y4Error(f"bgPortalLight '{n}' needs a background")
                    l.type, l.double_sided = _lib.LIGHT_PORTAL, 0
                l.object_n_tris, l.object_tris = len(ov), self._light_tris[-1].ctypes.data
            elif p["type"] == "bglight":           # BackgroundLight::factory (light_background.cc:200-224); created by createBackground
                if p["_background"] != rp.get("background_name", ""):
                    raise Y4Error("an IBL light of a background other than the rendered one is not supported by the GPU back end")
                l.type, l.samples, l.double_sided = _lib.LIGHT_BACKGROUND, int(p.get("samples", 16)), int(p.get("abs_intersect", False))
            elif p["type"] == "spherelight":       # SphereLight::factory (light_sphere.cc:160-195)
                l.type, l.samples, l.radius = _lib.LIGHT_SPHERE, int(p.get("samples", 4)), float(p.get("radius", 1.0))
                l.from_[:] = p.get("from", (0.0, 0.0, 0.0))
            else:
                l.type, l.samples = _lib.LIGHT_AREA, int(p.get("samples", 4))
                l.from_[:] = p.get("corner", (0.0, 0.0, 0.0))
                l.point1[:] = p.get("point1", (0.0, 0.0, 0.0))
                l.point2[:] = p.get("point2", (0.0, 0.0, 0.0))

        pr = _lib.PtParams()
        pr.bounces, pr.path_samples = int(integ.get("bounces", 3)), int(integ.get("path_samples", 32))
        if integ.get("type") == "directlighting":      # same kernels, no path loop
            pr.bounces, pr.path_samples = 1, 0
        pr.russian_roulette_min_bounces, pr.no_recursive = int(integ.get("russian_roulette_min_bounces", 0)), int(integ.get("no_recursive", False))
        pr.transp_background = int(integ.get("bg_transp", False))
        if integ.get("transpShad", False) and any(float(self.materials[n].get("transparency", 0.0)) > 0.00001 and n in chain_of for n in mat_names):
            raise Y4Error("textured transparent materials with transpShad are not supported by the GPU back end")
        if pr.no_recursive and any(self.materials[n]["type"] in ("mirror", "glass") or float(self.materials[n].get("specular_reflect", 0.0)) > 0.00001
                                   or float(self.materials[n].get("transparency", 0.0)) > 0.00001 for n in mat_names):
            raise Y4Error("no_recursive with mirror materials / specular_reflect is not supported by the GPU back end")
        pr.raydepth, pr.transp_refracted_background = int(integ.get("raydepth", 5)), int(integ.get("bg_transp_refract", False))
        pr.trace_caustics = int(integ.get("type") == "pathtracing" and integ.get("caustic_type", "path") not in ("none", "photon"))
        pr.transp_shadows, pr.shadow_depth = int(bool(integ.get("transpShad", False))), int(integ.get("shadowDepth", 5))
        # do_AO: DirectLightIntegrator adds sampleAmbientOcclusion at every diffuse vertex (integrator_direct_light.cc:124); the path
        # tracer reads the same parameters but only feeds the AO layers with them (integrator_path_tracer.cc:277), never Combined
        if integ.get("do_AO", False) and integ.get("type") == "directlighting":
            pr.ao_samples, pr.ao_distance = int(integ.get("AO_samples", 32)), float(integ.get("AO_distance", 1.0))
            pr.ao_color[:] = [float(x) for x in integ.get("AO_color", (1.0, 1.0, 1.0))[:3]]
        if bg is not None:
            c, pw = bg.get("color", (0.0, 0.0, 0.0, 1.0)), np.float32(bg.get("power", 1.0))
            pr.background[:] = [float(np.float32(c[k]) * pw) for k in range(3)]
            if bg.get("type") == "gradientback":        # GradientBackground::factory (background_gradient.cc:49-77): ground colours default to the sky's
                sh, sz = bg.get("horizon_color", (1.0, 1.0, 1.0, 1.0)), bg.get("zenith_color", (0.4, 0.5, 1.0, 1.0))
                gh, gz = bg.get("horizon_ground_color", sh), bg.get("zenith_ground_color", sz)
                pr.background_gradient = 1
                for j, col in enumerate((gz, gh, sz, sh)):
                    pr.bg_gradient[j][:] = [float(np.float32(col[k]) * pw) for k in range(3)]
            elif bg.get("type") in ("sunsky", "darksky"):
                pr.background_gradient = 3 if bg["type"] == "sunsky" else 4
                if bg["type"] == "darksky":
                    pr.bg_dark_altitude, pr.bg_dark_bright = float(bg.get("altitude", 0.0)), float(bg.get("bright", 1.0))
                    pr.bg_dark_exposure, pr.bg_dark_night = float(bg.get("exposure", 1.0)), int(bool(bg.get("night", False)))
                    pr.bg_dark_color_space = _lib.DARKSKY_COLOR_SPACES.get(str(bg.get("color_space", "CIE (E)")), 0)
                pr.bg_sun_from[:] = bg.get("from", (1.0, 1.0, 1.0))[:3]
                pr.bg_sun_turbidity, pr.bg_sun_power = float(bg.get("turbidity", 4.0)), float(bg.get("power", 1.0))
                pr.bg_sun_var[:] = [float(bg.get(k, 1.0)) for k in ("a_var", "b_var", "c_var", "d_var", "e_var")]
            elif bg.get("type") == "textureback":
                pr.background_gradient, pr.bg_texture = 2, tex_index[bg["texture"]]
                pr.bg_texture_angular = int(str(bg.get("mapping", "")) in ("probe", "angular"))
                pr.bg_texture_rotation, pr.bg_texture_power = float(bg.get("rotation", 0.0)), float(bg.get("power", 1.0))
                pr.bg_ibl_clamp = float(bg.get("ibl_clamp_sampling", 0.0)) if bg.get("ibl", False) else 0.0
        if not rp.get("background_resampling", True):
            raise Y4Error("background_resampling=false is not supported by the GPU back end")
        pr.width, pr.height = int(rp.get("width", cam.resx)), int(rp.get("height", cam.resy))
        pr.xstart, pr.ystart = int(rp.get("xstart", 0)), int(rp.get("ystart", 0))
        pr.aa_passes, pr.aa_samples, pr.aa_inc_samples = int(rp.get("AA_passes", 1)), int(rp.get("AA_minsamples", 1)), int(rp.get("AA_inc_samples", 1))
        pr.aa_sample_multiplier_factor = float(rp.get("AA_sample_multiplier_factor", 1.0))
        pr.aa_light_sample_multiplier_factor = float(rp.get("AA_light_sample_multiplier_factor", 1.0))
        pr.aa_pixelwidth, pr.aa_clamp_samples = float(rp.get("AA_pixelwidth", 1.5)), float(rp.get("AA_clamp_samples", 0.0))
        ft = str(rp.get("filter_type", "box"))
        pr.filter_type = _lib.FILTERS.get(ft, 0)
        pr.shadow_bias_auto = int(rp.get("adv_auto_shadow_bias_enabled", True))
        pr.shadow_bias = GLOBAL_SHADOW_BIAS if pr.shadow_bias_auto else float(rp.get("adv_shadow_bias_value", GLOBAL_SHADOW_BIAS))
        pr.min_raydist_auto = int(rp.get("adv_auto_min_raydist_enabled", True))
        pr.min_raydist = GLOBAL_MIN_RAYDIST if pr.min_raydist_auto else float(rp.get("adv_min_raydist_value", GLOBAL_MIN_RAYDIST))
        base_offs = getattr(self, "_base_offset_override", None)      # raised by the film files a load-save render starts from
        base_offs = int(rp.get("adv_base_sampling_offset", 0)) if base_offs is None else base_offs
        pr.base_sampling_offset = (base_offs + 100000 * int(rp.get("adv_computer_node", 0))) & 0xFFFFFFFF
        pr.rr_seed = int(rp.get("b200_rr_seed", 0x5EED)) & 0xFFFFFFFF
        pr.wave_size = int(rp.get("b200_wave_size", 0))
        pr.profile = int(rp.get("b200_profile", 0))
        tri_orco = np.ascontiguousarray(np.concatenate(orcos), np.float32) if layer_list and orcos else None
        tri_uv = np.ascontiguousarray(np.concatenate(uvs), np.float32) if layer_list and uvs else None
        return {"meshes": meshes, "mesh_flags": mesh_flags, "instances": inst_list, "n_prims": n_prims,
                "any_instances": any(m is not None for _, m in inst_list),
                "_light_tris": self._light_tris,      # the object lights' vertex arrays the `lights` structs point into
                "tris": T, "flags": F, "normals": N, "cam": cam, "mats": mats, "n_mat": len(mat_names), "lights": lights,
                "n_light": len(light_names), "params": pr, "textures": textures, "n_tex": len(tex_list), "layers": layers,
                "n_layers": len(layer_list), "tri_orco": tri_orco, "tri_uv": tri_uv, "_texels": texel_arrays}

    # ---- render (Scene::render -> TiledIntegrator::render, integrator_tiled.cc:97-233) --------------
    def upload(self):
        """Geometry upload + BVH8 build + scene description (the accelerator/integrator preprocess step)."""
        two_level = self._wants_two_level()
        a = self.assemble(flatten=not two_level)
        if two_level and not a["any_instances"]:
            two_level, a = False, self.assemble()
        if self.accelerator is None or not self.accelerator.ctx.value:
            self.accelerator = AcceleratorB200(None, device=self.device)
        if two_level:
            try:
                self.accelerator.rebuild_instances(a["meshes"], a["instances"], a["mesh_flags"])
            except Y4Error as e:      # e.g. a singular instance matrix: the flattened scene is always possible
                self.warnings.append(f"two-level upload refused ({e}); uploading flattened triangles")
                two_level, a = False, self.assemble()
        if not two_level:
            self.accelerator.rebuild(a["tris"], flags=a["flags"])       # same context: device allocations are reused
        a["two_level"] = two_level
        L, ctx = self.accelerator._L, self.accelerator.ctx
        nptr = C.c_void_p(a["normals"].ctypes.data) if a["normals"] is not None else None
        optr = C.c_void_p(a["tri_orco"].ctypes.data) if a["tri_orco"] is not None else None
        uptr = C.c_void_p(a["tri_uv"].ctypes.data) if a["tri_uv"] is not None else None
        _lib.check(ctx, L.y4gpu_pt_set_textures(ctx, a["textures"], a["n_tex"], a["layers"], a["n_layers"], optr, uptr), "y4gpu_pt_set_textures")
        _lib.check(ctx, L.y4gpu_pt_set_scene(ctx, C.byref(a["cam"]), a["mats"], a["n_mat"], a["lights"], a["n_light"], C.byref(a["params"]), nptr), "y4gpu_pt_set_scene")
        if self.comm is not None and self.shard_count > 1:
            # sharded rendering: this rank splats into its pass film, which is the collective's buffer; the all-reduced passes
            # are summed into the total film every rank keeps (noise detection and read-back work on that)
            import torch

            npx = a["params"].width * a["params"].height
            dev = torch.device("cuda", self.device)
            # kernels, the collective and the film adds are ordered by ONE stream.  It cannot be torch's default stream: its
            # handle is 0, which y4gpu_set_stream reads as "use the context's own stream", and that non-blocking stream is not
            # ordered against the legacy default stream
            cur = torch.cuda.current_stream(dev)
            if cur.cuda_stream == 0:
                if getattr(self, "_torch_stream", None) is None:
                    self._torch_stream = torch.cuda.Stream(device=dev)
                cur = self._torch_stream
            self._shard_stream = cur
            if self._pass_film is None or self._pass_film.numel() != 5 * npx:
                with torch.cuda.stream(cur):
                    self._pass_film = torch.zeros(5 * npx, device=dev, dtype=torch.float32)
                    self._total_film = torch.zeros(5 * npx, device=dev, dtype=torch.float32)
            self.film_buffers = (self._pass_film[:npx], self._pass_film[npx:])
            self.accelerator.set_stream(cur.cuda_stream)
        if self.film_buffers is not None:
            w, rgba = self.film_buffers
            _lib.check(ctx, L.y4gpu_pt_set_film_buffers(ctx, C.c_void_p(w.data_ptr()), C.c_void_p(rgba.data_ptr())), "y4gpu_pt_set_film_buffers")
        self._assembled = a
        return a

    def _wants_two_level(self) -> bool:
        mode = self.device_instancing
        env = os.environ.get("Y4GPU_INSTANCING")
        if env is not None:
            mode = env not in ("0", "false", "")
        if mode is False or not any(o.instance_of for o in self.objects.values()):
            return False
        if mode is True:
            return True
        # automatic: the two-level traversal tests nodes in object space with a padding sized for similarity-like matrices (rotation,
        # mirror, uniform or mildly non-uniform scale, shear).  Needle-like matrices (axis scales 100 : 1) lose about 1e-4 of the
        # closest hits against the flattened scene and the reference (tools: differential fuzzing at the end of round 2), so scenes
        # that contain one are flattened -- always possible, just larger.
        for o in self.objects.values():       # (a cached per-instance number: this runs on every upload, C4 has 16 k instances)
            if o.instance_of and not (_anisotropy(o) <= MAX_TWO_LEVEL_ANISOTROPY):
                self.warnings.append(f"instance '{o.name}': axis scales spread {_anisotropy(o):.3g} : 1; the scene is uploaded flattened")
                return False
        return True

    @property
    def sharded(self) -> bool:
        return self._total_film is not None and self.comm is not None and self.shard_count > 1

    def reduce_pass(self):
        """Sharded mode: sum this pass over the ranks (one all-reduce of [weights | rgba], 20 B / pixel, over NVLink) and add it
        to the total film.  Queued on the stream the pass ran on; nothing is synchronised."""
        if not self.sharded:
            return
        import torch

        with torch.cuda.stream(self._shard_stream):
            self.comm.all_reduce(self._pass_film)
            self._total_film.add_(self._pass_film)
            self._pass_film.zero_()

    def render_pass_reduced(self, n_samples: int, offset: int, aa_pass: int, pixel_list=None):
        """render_pass on this rank's shard of the samples + reduce_pass: afterwards every rank holds the complete film."""
        self.render_pass(n_samples, offset, aa_pass, pixel_list=pixel_list)
        self.reduce_pass()

    def render_pass(self, n_samples: int, offset: int, aa_pass: int, pixel_list=None):
        """TiledIntegrator::renderPass(samples, offset, adaptive, aa_pass_number) on this GPU's shard."""
        L, ctx = self.accelerator._L, self.accelerator.ctx
        if pixel_list is None:
            ptr, n = None, 0
        elif isinstance(pixel_list, tuple):          # (device pointer, count) from next_pass()
            ptr, n = pixel_list
        elif hasattr(pixel_list, "data_ptr"):
            ptr, n = C.c_void_p(pixel_list.data_ptr()), int(pixel_list.numel())
        else:
            pixel_list = np.ascontiguousarray(pixel_list, np.uint32)
            ptr, n = C.c_void_p(pixel_list.ctypes.data), int(pixel_list.size)
        _lib.check(ctx, L.y4gpu_pt_render_pass(ctx, int(n_samples), int(offset), int(aa_pass), ptr, n, self.shard_index, self.shard_count), "y4gpu_pt_render_pass")

    def aa_noise_params(self, threshold=None) -> "_lib.PtAaNoise":
        """AaNoiseParams of the render parameters (scene.cc:586-595), optionally with the adapted threshold."""
        rp = self.render_params
        aa = _lib.PtAaNoise()
        aa.threshold = float(rp.get("AA_threshold", 0.05)) if threshold is None else float(threshold)
        aa.detect_color_noise = int(bool(rp.get("AA_detect_color_noise", False)))
        ddt = str(rp.get("AA_dark_detection_type", "none"))
        aa.dark_detection_type = {"linear": 1, "curve": 2}.get(ddt, 0)
        aa.dark_threshold_factor = float(rp.get("AA_dark_threshold_factor", 0.0))
        aa.variance_edge_size, aa.variance_pixels = int(rp.get("AA_variance_edge_size", 10)), int(rp.get("AA_variance_pixels", 0))
        return aa

    def next_pass(self, aa, film=None):
        """ImageFilm::nextPass(adaptive) on the device -> (device pixel-list pointer, n_resample).
        film: optional (weights, rgba) torch CUDA tensors holding the accumulated film (multi-GPU: the all-reduced total)."""
        L, ctx = self.accelerator._L, self.accelerator.ctx
        ptr, n = C.c_void_p(), C.c_uint64()
        if film is None and self.sharded:          # the all-reduced total: identical on every rank, so are the flagged pixels
            npx = self._total_film.numel() // 5
            film = (self._total_film[:npx], self._total_film[npx:])
        fw = C.c_void_p(film[0].data_ptr()) if film is not None else None
        fr = C.c_void_p(film[1].data_ptr()) if film is not None else None
        _lib.check(ctx, L.y4gpu_pt_next_pass(ctx, C.byref(aa), fw, fr, C.byref(ptr), C.byref(n)), "y4gpu_pt_next_pass")
        return ptr, int(n.value)

    def render(self, monitor_callback=None):
        """yafaray_render = TiledIntegrator::render (integrator_tiled.cc:97-234): pass 0 with AA_minsamples over every
        pixel, then AA_passes-1 passes of ceil(AA_inc_samples * mult) samples.  With AA_threshold > 0 the later passes
        are adaptive: ImageFilm::nextPass flags the noisy pixels (on the device), only those are resampled, and the
        threshold is lowered when fewer than AA_resampled_floor % of the pixels were flagged -- the reference's loop,
        including its fp32 arithmetic.  Returns the normalised Combined layer."""
        rp = self.render_params
        # film_load_save_mode (imagefilm.cc:74-118,236-241): "load-save" starts from the sum of the film files in the output
        # folder and skips pass 0 (integrator_tiled.cc:155-177); "save" / "load-save" write the film file at the end
        mode, film_file = str(rp.get("film_load_save_mode", "none")), str(rp.get("film_load_save_path", ""))
        node, own_base = int(rp.get("adv_computer_node", 0)), int(rp.get("adv_base_sampling_offset", 0))
        self._base_offset_override, resumed = None, None
        if mode == "load-save":
            p0 = self.assemble()["params"]
            lw, lc, base, offs, n_loaded = film_io.load_all_in_folder(film_file, p0.width, p0.height, p0.xstart, p0.ystart, own_base)
            if n_loaded:
                resumed, self._base_offset_override = (lw, lc, offs), base
        a = self.upload()
        pr = a["params"]
        f32 = np.float32
        L, ctx = self.accelerator._L, self.accelerator.ctx
        _lib.check(ctx, L.y4gpu_pt_clear_film(ctx), "y4gpu_pt_clear_film")
        if self.sharded:
            if mode != "none":
                raise Y4Error("film_load_save_mode together with multi-GPU sharding is not supported (save the film from rank 0's read_film())")
            import torch

            with torch.cuda.stream(self._shard_stream):
                self._total_film.zero_()
        stats = {"samples": 0, "rays_closest": 0, "rays_shadow": 0, "kernel_launches": 0, "pass_ms": 0.0}
        self.resumed = resumed is not None
        if resumed is not None:                                 # renderPass(0, film sampling offset): nothing is rendered
            _lib.check(ctx, L.y4gpu_pt_write_film(ctx, C.c_void_p(resumed[0].ctypes.data), C.c_void_p(resumed[1].ctypes.data)), "y4gpu_pt_write_film")
            self.sampling_offset = resumed[2]
        else:
            self.render_pass_reduced(pr.aa_samples, 0, 0)
            self._add_stats(stats)
            self.sampling_offset = pr.aa_samples
        threshold = f32(rp.get("AA_threshold", 0.05))
        floor_pixels = int(np.floor(f32(f32(rp.get("AA_resampled_floor", 0.0)) * f32(pr.width * pr.height)) / f32(100.0)))
        acum, mult = pr.aa_samples, f32(1.0)
        resampled, threshold_changed = 0, True
        self.resampled_per_pass = []
        for i in range(1, pr.aa_passes):
            mult = f32(mult * f32(pr.aa_sample_multiplier_factor))
            plist, skip = None, False
            if resampled <= 0 and not threshold_changed:
                skip = True                                     # nothing was resampled and the threshold did not move
            elif threshold > 0:
                ptr, resampled = self.next_pass(self.aa_noise_params(threshold))
                plist = (ptr, resampled)
                threshold_changed = False
            else:
                resampled, threshold_changed = pr.width * pr.height, False     # non-adaptive: every pixel
            self.resampled_per_pass.append(0 if skip else resampled)
            n = int(np.ceil(f32(pr.aa_inc_samples) * mult))
            if not skip and resampled > 0:
                self.render_pass_reduced(n, acum, i, pixel_list=plist)
                self._add_stats(stats)
                self.sampling_offset = acum + n                # ImageFilm::setSamplingOffset(offset + samples)
            acum += n
            if resampled < floor_pixels:
                ratio = min(f32(8.0), f32(floor_pixels) / f32(resampled)) if resampled > 0 else f32(8.0)
                threshold = f32(threshold * f32(f32(1.0) - f32(0.1) * ratio))
                if threshold > 0:
                    threshold_changed = True
        self.last_stats = stats
        if mode in ("save", "load-save"):                       # ImageFilm::flush -> imageFilmSave
            fw, frgba = self.read_film()
            film_io.save_film(film_io.film_path(film_file, node), fw, frgba, node, own_base if self._base_offset_override is None else self._base_offset_override,
                              self.sampling_offset, pr.xstart, pr.ystart)
        return self.read_image()

    def pass_stats(self) -> dict:
        """y4gpu_pt_stats of the last pass (synchronises)."""
        self.accelerator.synchronize()
        s = _lib.PtStats()
        L, ctx = self.accelerator._L, self.accelerator.ctx
        _lib.check(ctx, L.y4gpu_pt_get_stats(ctx, C.byref(s)), "y4gpu_pt_get_stats")
        d = {k: getattr(s, k) for k, _ in s._fields_ if k not in ("reserved", "class_ms", "class_launches")}
        d["class_ms"], d["class_launches"] = list(s.class_ms), list(s.class_launches)
        return d

    def _add_stats(self, acc):
        s = self.pass_stats()
        for k in acc:
            acc[k] += s[k]

    def read_film(self, pinned: bool = False):
        """-> (weights [h, w], rgba [h, w, 4]) as accumulated (ImageFilm::weights_ and the Combined layer)."""
        pr = self._assembled["params"]
        if pinned:
            # page-locked staging owned by the interface (reused by the next call): the copy runs at PCIe speed
            if self._pinned_film is None or self._pinned_film[0].shape != (pr.height, pr.width):
                self._pinned_film = (_lib.pinned_empty(pr.height * pr.width, np.float32).reshape(pr.height, pr.width),
                                     _lib.pinned_empty(pr.height * pr.width * 4, np.float32).reshape(pr.height, pr.width, 4))
            w, rgba = self._pinned_film
        else:
            w = np.empty((pr.height, pr.width), np.float32)
            rgba = np.empty((pr.height, pr.width, 4), np.float32)
        L, ctx = self.accelerator._L, self.accelerator.ctx
        if self.sharded:                            # the complete film (all passes, all ranks) lives in the total tensor
            import torch

            npx = pr.height * pr.width
            with torch.cuda.stream(self._shard_stream):
                torch.from_numpy(w.reshape(-1)).copy_(self._total_film[:npx])
                torch.from_numpy(rgba.reshape(-1)).copy_(self._total_film[npx:])
            self._shard_stream.synchronize()
            return w, rgba
        _lib.check(ctx, L.y4gpu_pt_read_film(ctx, C.c_void_p(w.ctypes.data), C.c_void_p(rgba.ctypes.data)), "y4gpu_pt_read_film")
        return w, rgba

    def read_image(self):
        """ImageFilm::flush normalisation (imagefilm.cc:598-600, Rgba::normalized): colour / weight, 0 where weight == 0."""
        w, rgba = self.read_film()
        return normalize_film(w, rgba)

    def clearAll(self):
        if self.accelerator is not None:
            self.accelerator.close()
            self.accelerator = None


def _normalized(v: np.ndarray) -> np.ndarray:
    """Vec3::normalize on rows (vector.h:201-210): len = 1 / sqrt(x*x + y*y + z*z), v *= len; null vectors stay."""
    f32 = np.float32
    l2 = (v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]) + v[:, 2] * v[:, 2]
    inv = np.where(l2 != 0, f32(1.0) / np.sqrt(np.where(l2 != 0, l2, f32(1.0))), f32(1.0)).astype(f32)
    return (v * inv[:, None]).astype(f32)


def _cross(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    return np.stack([a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1], a[:, 2] * b[:, 0] - a[:, 0] * b[:, 2], a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]], -1).astype(np.float32)


def _length(v: np.ndarray) -> np.ndarray:
    return np.sqrt((v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]) + v[:, 2] * v[:, 2]).astype(np.float32)


def smooth_normals(v: np.ndarray, idx: np.ndarray, angle: float) -> np.ndarray:
    """MeshObject::smoothNormals (src/geometry/object/object_mesh.cc:118-240) -> the shading normal at the three corners of
    every face, [n_tri, 9] fp32.  Face normals are weighted by the corner's angle (asin of Vec3::sinFromVectors, vector.h:212-219);
    angle >= 180: one normal per vertex; otherwise only faces within `angle` of each other are blended, a corner that blends
    with nothing keeps the face normal, and nearly equal normals at one vertex (dot > 0.999) share the first one found.
    fp32 throughout, sums in the reference's order."""
    f32 = np.float32
    v = np.ascontiguousarray(v, f32)
    idx = np.asarray(idx, np.int64)
    n_tri = len(idx)
    p = v[idx]                                                       # [n_tri, 3 corners, 3]
    ng = _normalized(_cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]))   # Primitive::getGeometricNormal (primitive_triangle.cc:87-95)
    sines = np.empty((n_tri, 3), f32)
    for c in range(3):
        e1, e2 = p[:, (c + 1) % 3] - p[:, c], p[:, (c + 2) % 3] - p[:, c]
        div = (_length(e1) * _length(e2)) * f32(0.99999) + f32(0.00001)
        arg = np.minimum((_length(_cross(e1, e2)) / div) * f32(0.99999), f32(1.0)).astype(f32)
        sines[:, c] = np.arcsin(arg.astype(np.float64)).astype(f32)     # std::asin(float): rounded from a more precise evaluation
    out = np.repeat(ng, 3, axis=0).reshape(n_tri, 3, 3).copy()        # corners that stay unsmoothed use the face normal
    if angle >= 180:
        acc = np.zeros((len(v), 3), f32)
        contrib = (ng[:, None, :] * sines[:, :, None]).astype(f32)    # n * sine, faces in order, corners 0, 1, 2
        for k in range(3):
            np.add.at(acc[:, k], idx.reshape(-1), contrib[:, :, k].reshape(-1))      # unbuffered: sums in face order
        return np.ascontiguousarray(_normalized(acc)[idx].reshape(n_tri, 9), f32)
    thr = _fast_cos_deg(angle)
    order = np.argsort(idx.reshape(-1), kind="stable")               # incident (face, corner) pairs per vertex, faces in order
    flat_v = idx.reshape(-1)[order]
    starts = np.searchsorted(flat_v, np.arange(len(v) + 1))
    valence = np.diff(starts)
    # vertices of the same valence d are processed together: every step below is the scalar loop's step applied to all of them
    # at once, so the fp32 order of the sums is unchanged.  Very high valences (sphere poles) take the scalar path.
    for d in np.unique(valence):
        d = int(d)
        if d == 0:
            continue
        pts = np.nonzero(valence == d)[0]
        if d > 24 or len(pts) < 8:
            for pt in pts:
                _smooth_vertex(order[starts[pt]:starts[pt + 1]], ng, sines, thr, out)
            continue
        inc = order[starts[pts][:, None] + np.arange(d)[None, :]]            # [n, d] incident (face, corner) pairs, faces in order
        faces, corners = inc // 3, inc % 3
        N, S = ng[faces], sines[faces, corners]                              # [n, d, 3], [n, d]
        NS = (N * S[..., None]).astype(f32)
        found = np.zeros((len(pts), d, 3), f32)                              # normals already created at each vertex
        n_found = np.zeros(len(pts), np.int64)
        rows = np.arange(len(pts))
        for j in range(d):
            dots = (N[:, :, 0] * N[:, j, 0][:, None] + N[:, :, 1] * N[:, j, 1][:, None]) + N[:, :, 2] * N[:, j, 2][:, None]     # [n, d]
            vn = NS[:, j].copy()
            smooth = np.zeros(len(pts), bool)
            for k in range(d):
                take = (faces[:, k] != faces[:, j]) & (dots[:, k] > thr)
                smooth |= take
                vn = np.where(take[:, None], (vn + NS[:, k]).astype(f32), vn)
            vn = _normalized(vn)
            reuse = np.zeros(len(pts), bool)
            for e in range(d):                                               # the first earlier normal within 0.999 replaces it
                cand = found[:, e]
                hit = smooth & ~reuse & (e < n_found) & ((vn[:, 0] * cand[:, 0] + vn[:, 1] * cand[:, 1]) + vn[:, 2] * cand[:, 2] > f32(0.999))
                vn = np.where(hit[:, None], cand, vn)
                reuse |= hit
            new = smooth & ~reuse
            found[rows[new], n_found[new]] = vn[new]
            n_found[new] += 1
            out[faces[smooth, j], corners[smooth, j]] = vn[smooth]
    return np.ascontiguousarray(out.reshape(n_tri, 9), f32)


def _smooth_vertex(inc, ng, sines, thr, out) -> None:
    """One vertex of smooth_normals' angle-limited branch (scalar form; the vectorised form above follows it step by step)."""
    f32 = np.float32
    faces, corners = inc // 3, inc % 3
    N, S = ng[faces], sines[faces, corners]
    found = []                                                  # normals already created at this vertex
    for j in range(len(inc)):
        dots = (N[:, 0] * N[j, 0] + N[:, 1] * N[j, 1]) + N[:, 2] * N[j, 2]
        vn = (N[j] * S[j]).astype(f32)
        smooth = False
        for k in range(len(inc)):
            if faces[k] == faces[j]:
                continue
            if dots[k] > thr:
                smooth = True
                vn = (vn + N[k] * S[k]).astype(f32)
        if not smooth:
            continue
        vn = _normalized(vn[None, :])[0]
        for e in found:
            if (vn[0] * e[0] + vn[1] * e[1]) + vn[2] * e[2] > f32(0.999):
                vn = e
                break
        else:
            found.append(vn)
        out[faces[j], corners[j]] = vn


def _fast_cos_deg(angle: float) -> np.float32:
    """math::cos(math::degToRad(angle)) with the FAST_TRIG polynomial (math.h:218-250)."""
    f32 = np.float32
    x = f32(np.longdouble(f32(angle)) * np.longdouble("0.017453292519943295769236907684886"))
    x = f32(x + f32(np.longdouble("1.5707963267948966192313216916398")))
    two_pi = f32(np.longdouble("6.283185307179586476925286766559"))
    if x > two_pi or x < -two_pi:
        x = f32(x - f32(int(x * f32(np.longdouble("0.15915494309189533576888376337251")))) * two_pi)
    pi = np.longdouble("3.1415926535897932384626433832795")
    if np.longdouble(x) < -pi:
        x = f32(x + two_pi)
    elif np.longdouble(x) > pi:
        x = f32(x - two_pi)
    x = f32(f32(np.longdouble("1.2732395447351626861510701069801") * np.longdouble(x))
            - f32(np.longdouble("0.40528473456935108577551785283891") * np.longdouble(x) * np.longdouble(abs(x))))
    r = f32(f32(0.225) * f32(f32(x * abs(x)) - x) + x)
    return f32(min(max(r, f32(-1.0)), f32(1.0)))


def normalize_film(w: np.ndarray, rgba: np.ndarray) -> np.ndarray:
    """Rgba::normalized(weight): `*this / weight` with Rgba operator/ = multiply by 1.f / weight (color.h:312-316,554-558)."""
    inv = np.where(w != 0, np.float32(1.0) / np.where(w != 0, w, np.float32(1.0)), np.float32(0.0)).astype(np.float32)
    return (rgba * inv[..., None]).astype(np.float32)
