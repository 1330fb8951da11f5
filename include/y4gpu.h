/* y4gpu.h -- C-ABI of the B200 ray-intersection / wavefront path-tracing back end.
 *
 * This is the drop-in boundary: the libYafaRay-side plugin classes (an Accelerator
 * subtype registered in Accelerator::factory and a PathIntegrator back end, see
 * INTEGRATION.md) bind exactly these entry points.  Plain C, POD structs, int return
 * codes (0 = ok, <0 = error, text via y4gpu_last_error), no exceptions and no torch
 * types cross the boundary.  The library has NO CPU fallback: every compute entry
 * point fails with Y4GPU_ERR_CUDA when no sm_100 device is usable.
 *
 * Reference interfaces replaced (paths relative to the libYafaRay tree):
 *   y4gpu_upload_triangles/_build <- AcceleratorKdTree ctor + buildTree
 *                                    src/accelerator/accelerator_kdtree.cc:47-123,420-628
 *                                    (primitive list gathered by Scene::updateObjects,
 *                                    src/scene/scene.cc:1032-1052)
 *   y4gpu_get_bound               <- Accelerator::getBound, include/accelerator/accelerator.h:62
 *   y4gpu_trace(CLOSEST)          <- Accelerator::intersect(ray, t_max), accelerator.h:59,
 *                                    accelerator_kdtree.cc:634-773, wrapper accelerator.cc:55-67
 *   y4gpu_trace(SHADOW)           <- Accelerator::isShadowed(ray, bias) -> intersectS,
 *                                    accelerator.cc:69-80, accelerator_kdtree.cc:775-904
 *   y4gpu_trace(TSHADOW)          <- Accelerator::isShadowed(ray, depth, bias, camera) ->
 *                                    intersectTs, accelerator.cc:82-94, accelerator_kdtree.cc:911-1061
 *   y4gpu_pt_*                    <- PathIntegrator::integrate + TiledIntegrator::renderPass +
 *                                    ImageFilm::addSample (see the pt section below)
 */
#ifndef Y4GPU_H
#define Y4GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define Y4GPU_OK 0
#define Y4GPU_ERR_ARG (-1)    /* bad argument / wrong call order */
#define Y4GPU_ERR_CUDA (-2)   /* CUDA runtime failure or no device (no CPU fallback) */
#define Y4GPU_ERR_UNSUPPORTED (-3) /* scene feature outside the supported set: rejected, never approximated */

typedef struct y4gpu_ctx y4gpu_ctx;

/* Ray exactly as the reference's Ray carries it (include/geometry/ray.h:47-51):
 * tmax < 0 means unbounded.  The wrapper conventions of accelerator.cc:55-94 (shadow-ray
 * origin advance by dir*tmin, t_max = tmax - 2*tmin) are applied on the device. */
typedef struct y4gpu_ray { float ox, oy, oz, tmin, dx, dy, dz, tmax; } y4gpu_ray;

/* prim = index into the primitive list given to y4gpu_upload_triangles (== position in
 * the vector handed to Accelerator::factory), -1 on miss.  (bu, bv) are the Moller-Trumbore
 * (u, v) == reference barycentric_v_ / barycentric_w_; barycentric_u_ = 1 - bu - bv
 * (src/geometry/primitive/primitive_triangle.cc:66-68). */
typedef struct y4gpu_hit { float t, bu, bv; int32_t prim; } y4gpu_hit;

/* transparent-shadow result: hit.prim >= 0 => shadowed (opaque blocker or depth exceeded);
 * otherwise filter = product of the transparency colours of the distinct transparent
 * primitives crossed (AcceleratorTsIntersectData::transparent_color_). */
typedef struct y4gpu_tshit { y4gpu_hit hit; float filter_r, filter_g, filter_b, pad; } y4gpu_tshit;

enum { Y4GPU_TRACE_CLOSEST = 0, Y4GPU_TRACE_SHADOW = 1, Y4GPU_TRACE_TSHADOW = 2 };

/* per-triangle flag word (visibility matrix of accelerator_kdtree.cc:733-737,869-873,1022) */
#define Y4GPU_TRI_VIS_CLOSEST 1u
#define Y4GPU_TRI_VIS_SHADOW 2u
#define Y4GPU_TRI_VIS_TSHADOW 4u
#define Y4GPU_TRI_TRANSPARENT 8u
#define Y4GPU_TRI_MAT_SHIFT 8u

typedef struct y4gpu_build_params
{
	int32_t leaf_max;     /* triangles per BVH8 leaf slot, 1..3 (0 = default) */
	int32_t reserved[7];
} y4gpu_build_params;

typedef struct y4gpu_stats
{
	uint64_t n_tris, n_nodes;           /* BVH8 size */
	uint64_t node_bytes, tri_bytes;     /* resident HBM footprint of the accelerator */
	double build_ms;                    /* last y4gpu_build, device time */
	double last_trace_ms;               /* last y4gpu_trace kernel time (CUDA events on the trace stream) */
	uint64_t last_trace_rays;
	uint64_t nodes_visited, tris_tested; /* filled by y4gpu_trace_counted only */
	uint64_t kernel_launches;           /* kernels launched by this ctx since creation */
	uint64_t bvh_levels;                /* depth of the BVH8 of the last build (the traversal stack holds Y4_STACK_SIZE = 64 entries;
	                                       builds deeper than 56 levels are refused with Y4GPU_ERR_UNSUPPORTED) */
	uint64_t last_trace_sorted;         /* 1: the last resident y4gpu_trace batch was traced in sorted ray order (y4gpu_set_ray_sort) */
	double last_probe_distinct;         /* automatic mode: distinct coarse ray keys per 32 consecutive rays found by the probe */
} y4gpu_stats;

/* lifetime ------------------------------------------------------------------------------ */
int y4gpu_create(int device, y4gpu_ctx **out);
void y4gpu_destroy(y4gpu_ctx *ctx);
const char *y4gpu_last_error(const y4gpu_ctx *ctx);   /* ctx may be NULL: last create error */
int y4gpu_device_count(void);
/* sizeof of the public structs as this library was compiled, for bindings that mirror them by hand (ctypes, cgo, JNI): index
 * 0 build_params, 1 stats, 2 mesh, 3 instance, 4 pt_material, 5 pt_texture, 6 pt_layer, 7 pt_light, 8 pt_camera, 9 pt_params,
 * 10 pt_stats, 11 pt_aa_noise, 12 ray, 13 hit, 14 tshit; 0 for any other index.  No GPU needed. */
uint64_t y4gpu_struct_size(int which);

/* Run all work of this ctx on the caller's CUDA stream (cudaStream_t passed as void*; NULL
 * restores the ctx's own non-blocking stream).  Lets a host application order the ctx's
 * kernels against its own copies/kernels without extra synchronisation. */
int y4gpu_set_stream(y4gpu_ctx *ctx, void *cuda_stream);

/* pinned host buffers for callers that want async copies (optional) */
void *y4gpu_host_alloc(uint64_t bytes);
void y4gpu_host_free(void *p);

/* geometry + build ---------------------------------------------------------------------- */
/* v_xyz: n_tri*9 floats, world space (instances already flattened like
 * PrimitiveInstance does per test, primitive_instance.cc:43-51), in primitive order.
 * flags: n_tri words or NULL (all visible, opaque, material 0).  Host or device pointers. */
int y4gpu_upload_triangles(y4gpu_ctx *ctx, const float *v_xyz, const uint32_t *flags, uint64_t n_tri);
/* Instanced geometry (yafaray_addInstance; PrimitiveInstance / ObjectInstance, primitive_instance.cc:43-51,
 * object_instance.cc:25-33): base meshes in OBJECT space, uploaded once, and instances that place them with the rows 0..2 of
 * the reference's Matrix4 (obj_to_world).  A plain, un-instanced object is an instance with identity = 1 of a mesh given in world
 * space.  Primitive indices run over the instances in the order given (instance i's triangle k = first_prim(i) + k), i.e. exactly
 * the indices of the flattened scene; results are bit-identical to uploading the flattened triangles: the device keeps the
 * meshes once (a top-level BVH8 over instances, one bottom-level BVH8 per mesh) and transforms a triangle's vertices to world
 * space per test like the reference does.  flags may be NULL per mesh (all visible, opaque, material 0).
 * Every query works on such a scene; y4gpu_trace_counted counts an instance entry as two nodes. */
typedef struct y4gpu_mesh { const float *v_xyz; const uint32_t *flags; uint64_t n_tri; } y4gpu_mesh;
typedef struct y4gpu_instance { uint32_t mesh; uint32_t identity; float obj_to_world[12]; } y4gpu_instance;
int y4gpu_upload_instanced(y4gpu_ctx *ctx, const y4gpu_mesh *meshes, uint32_t n_mesh, const y4gpu_instance *instances, uint32_t n_inst);
int y4gpu_build(y4gpu_ctx *ctx, const y4gpu_build_params *params /* nullable */);
/* tree bound with the reference's 0.1 % inflation (accelerator_kdtree.cc:81-85): lo xyz, hi xyz */
int y4gpu_get_bound(const y4gpu_ctx *ctx, float out_bound[6]);

/* ray batches ---------------------------------------------------------------------------- */
/* rays/out may each be host or device memory (detected); out has n records.
 * Host rays AND host hits: the call is synchronous and RE-ENTRANT -- any number of host threads may call it on one context
 * concurrently (the Accelerator virtuals are called from every render thread, integrator_tiled.cc:248-251); each call runs on
 * a private slot of the context (own streams, staging, work counters).  Batches of >= 2^19 rays are cut into chunks whose
 * H2D copy, traversal and D2H copy overlap on several streams (pass page-locked buffers, y4gpu_host_alloc, for that).
 * Device buffers: queued on the context's stream, not waited for; such calls are serialised by the context's mutex. */
int y4gpu_trace(y4gpu_ctx *ctx, int kind, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out);
/* Ray order of large batches (>= 65536 rays; y4gpu_trace closest / shadow).  A tuning knob without a counterpart in the
 * reference.  mode 0: rays are traced in the order given.  1: every batch is traced in the order of a 32-bit key (Morton
 * code of the origin's cell in the tree bound, then a direction class) through an index queue -- results are bit-identical,
 * only the order in which warps pick rays up changes.  2 (default): sorted only when the BVH is much larger than L2
 * (>= 256 MB) AND a probe of the batch (one small kernel + one synchronisation) finds it incoherent as laid out.
 * Worth it where traversal is bound by DRAM sector fetches (50 M triangles, random rays); a loss on camera rays.
 * Scratch (16 B per ray + the radix sort's temporary storage) is allocated on first use and kept. */
int y4gpu_set_ray_sort(y4gpu_ctx *ctx, int mode);
/* same, plus per-batch node/triangle visit counters in y4gpu_stats (slower; for the roofline) */
int y4gpu_trace_counted(y4gpu_ctx *ctx, int kind, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out);
/* transparent shadows: mat_transparency = n_mat * 3 floats (Material::getTransparency colour) */
int y4gpu_trace_tshadow(y4gpu_ctx *ctx, const y4gpu_ray *rays, uint64_t n, int max_depth,
                        const float *mat_transparency, uint32_t n_mat, y4gpu_tshit *out);
/* The same query with the arguments of the INNER virtual Accelerator::intersectTs(sray, max_depth, dist, ...)
 * (accelerator_kdtree.cc:911-1061), which receives the ray after the wrapper advanced its origin: no origin shift here,
 * accept ray.tmin <= t < ray.tmax (tmax = dist).  For the inner closest / shadow virtuals y4gpu_trace already fits:
 * intersect(ray, t_max) = CLOSEST with tmax = t_max, intersectS(sray, t_max, bias) = SHADOW with tmin = 0, tmax = t_max. */
int y4gpu_trace_tshadow_inner(y4gpu_ctx *ctx, const y4gpu_ray *rays, uint64_t n, int max_depth,
                              const float *mat_transparency, uint32_t n_mat, y4gpu_tshit *out);
/* Allocation-free variant of y4gpu_trace (CLOSEST / SHADOW).  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  It calls no cudaMalloc / cudaFree / allocating library routine: the traversal kernel's work
 * counter and the staging of host-resident rays / hits live in one static __device__ array of the library (3 MiB, held by
 * every GPU that loads the library until the process exits).  n > Y4GPU_NOALLOC_MAX_RAYS is rejected with Y4GPU_ERR_ARG
 * (split the batch, or use y4gpu_trace).  Results are identical to y4gpu_trace's.  All calls on one device share the array:
 * calls that overlap on the GPU (different contexts or streams without ordering) race; calls on one context with host
 * buffers are serialised by the library.  rays / out: host or device memory, like y4gpu_trace. */
#define Y4GPU_NOALLOC_MAX_RAYS 65536
int y4gpu_trace_noalloc(y4gpu_ctx *ctx, int kind, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out);
int y4gpu_get_stats(const y4gpu_ctx *ctx, y4gpu_stats *out);
int y4gpu_synchronize(y4gpu_ctx *ctx);


/* ===================================================================================== */
/* Wavefront path tracer: PathIntegrator::integrate + TiledIntegrator::renderPass/renderTile
 * + ImageFilm::addSample on the device.
 *
 * Reference interfaces replaced (libYafaRay tree):
 *   y4gpu_pt_set_scene   <- what PathIntegrator::preprocess caches (camera_, lights_, background_,
 *                           aa params; src/integrator/integrator.cc:62-90,
 *                           integrator_path_tracer.cc:52-118,292-366) in PARAMETER form: the library
 *                           derives camera axes, light frames, material component weights and the film
 *                           filter table itself with the reference's fp32 operation order
 *                           (camera.cc:51-71, camera_perspective.cc:28-69, light_area.cc:33-52,
 *                           material_shiny_diffuse.cc:39-82, imagefilm.cc:129-173)
 *   y4gpu_pt_render_pass <- TiledIntegrator::renderPass(samples, offset, adaptive, aa_pass)
 *                           (integrator_tiled.cc:236-408) incl. ImageFilm::addSample (imagefilm.cc:680-733)
 *   y4gpu_pt_read_film   <- the weights_ / Combined-layer buffers ImageFilm::flush normalises
 *                           (imagefilm.cc:590-617)
 * Scene features outside the supported set are rejected with Y4GPU_ERR_UNSUPPORTED.            */

enum { Y4GPU_MAT_SHINYDIFFUSE = 0, Y4GPU_MAT_LIGHT = 1,
       Y4GPU_MAT_GLASS = 3,      /* `glass` (GlassMaterial, src/material/material_glass.cc:35-433) without dispersion and fake shadows (absorption: the fields at the end of y4gpu_pt_material):
                                    `IOR` (in ior), `mirror_color`, and in `color` the constructor's filter colour transmit_filter * filter_color + (1 - transmit_filter).  getSpecular returns a
                                    reflected AND a refracted ray: recursiveRaytrace (integrator_montecarlo.cc:918-968) becomes a tree, which the
                                    device walks depth first with a per-pixel-sample stack of pending (ray, weight) entries */
       Y4GPU_MAT_MIRROR = 2 };   /* `mirror` (MirrorMaterial, src/material/material_glass.cc:435-461): `color` and `reflect`
                                    (in diffuse_reflect); perfect specular reflection, followed by the ray-level loop of
                                    MonteCarloIntegrator::recursiveRaytrace up to `raydepth` and sampled inside paths */
typedef struct y4gpu_pt_material
{
	int32_t type;                 /* Y4GPU_MAT_* */
	float color[3];               /* shinydiffusemat `color` / light_mat `color` */
	float diffuse_reflect;        /* shinydiffusemat: `diffuse_reflect` (default 1); mirror: `reflect` (default 1) */
	float emit;                   /* shinydiffusemat: `emit`; light_mat: `power` */
	int32_t oren_nayar;           /* diffuse_brdf == "oren_nayar" */
	float sigma;                  /* oren-nayar sigma */
	int32_t double_sided;         /* light_mat */
	int32_t receive_shadows;      /* default 1 */
	int32_t flat_material;        /* default 0 */
	int32_t layer_first, layer_count;   /* shinydiffusemat `diffuse_shader`: its chain of layer nodes in y4gpu_pt_set_textures'
	                                       layer array, bottom layer first (layer_count == 0: plain `color`) */
	float specular_reflect;       /* shinydiffusemat `specular_reflect` (default 0): strength of the perfect-mirror component */
	float mirror_color[3];        /* shinydiffusemat `mirror_color` (default 1 1 1) */
	int32_t fresnel_effect;       /* shinydiffusemat `fresnel_effect`: the mirror component is weighted by the Fresnel term of `ior` */
	float ior;                    /* shinydiffusemat `IOR` (default 1.33) */
	float transparency;           /* shinydiffusemat `transparency` (default 0): light passes straight through, tinted by the filter */
	float transmit_filter;        /* shinydiffusemat `transmit_filter` (default 1): share of the diffuse colour in that tint */
	float translucency;           /* shinydiffusemat `translucency` (default 0): diffuse transmission to the other side */
	int32_t reserved[3];
	/* glass `absorption` / `absorption_dist` (GlassMaterial::factory, material_glass.cc:339-365): when a colour component is < 1 the
	   material gets a Beer-Lambert volume handler (volumehandler_beer.cc) and every specular ray that continues INSIDE the glass has
	   its radiance multiplied by exp(-sigma_a * hit distance) (integrator_montecarlo.cc:868-911; 0 when that ray leaves the scene) */
	int32_t absorb;               /* 1: the handler exists */
	float absorption[3];          /* `absorption` colour (components <= 1e-38 are not supported) */
	double absorption_dist;       /* `absorption_dist` (default 1) */
} y4gpu_pt_material;

/* Image textures and the `diffuse_shader` node chains of shinydiffusemat: texture_mapper + layer nodes
 * (src/shader/shader_node_basic.cc:31-367, shader_node_layer.cc:30-351) over ImageTexture (src/texture/texture_image.cc).
 * Supported: texco uv / orco / global, mapping plain / cube, interpolate none / bilinear, every clipping mode, the
 * intensity / contrast / per-channel / clamp adjustments, every layer blend mode, colour inputs.  Anything else
 * (procedural textures, bicubic / mipmap interpolation, tube / sphere mapping, hue / saturation, scalar inputs, bump
 * and the other shader slots) is rejected by the host mirror, never approximated. */
enum { Y4GPU_CLIP_EXTEND = 0, Y4GPU_CLIP_CLIP = 1, Y4GPU_CLIP_CLIPCUBE = 2, Y4GPU_CLIP_REPEAT = 3, Y4GPU_CLIP_CHECKER = 4 };
enum { Y4GPU_INTERP_NONE = 0, Y4GPU_INTERP_BILINEAR = 1 };
typedef struct y4gpu_pt_texture
{
	const float *texels;          /* width * height RGBA floats, row-major (x fastest), exactly what Image::getColor(x, y) returns:
	                                 linear RGB after the colour-space conversion and the storage rounding of the image's
	                                 optimization mode (include/image/image_buffers.h); host or device memory, copied by the call */
	int32_t width, height;
	int32_t clip;                 /* Y4GPU_CLIP_* (`clipping`, default repeat) */
	int32_t interpolate;          /* Y4GPU_INTERP_* (`interpolate`, default bilinear) */
	int32_t xrepeat, yrepeat, mirror_x, mirror_y, rot90;
	float cropmin_x, cropmin_y, cropmax_x, cropmax_y;
	int32_t even_tiles, odd_tiles; float checker_dist;
	float adj_intensity, adj_contrast, adj_mult_factor_red, adj_mult_factor_green, adj_mult_factor_blue;
	int32_t adj_clamp;
	int32_t reserved[4];
} y4gpu_pt_texture;

enum { Y4GPU_TEXCO_GLOBAL = 0, Y4GPU_TEXCO_UV = 1, Y4GPU_TEXCO_ORCO = 2 };
enum { Y4GPU_PROJ_PLAIN = 0, Y4GPU_PROJ_CUBE = 1 };
enum { Y4GPU_BLEND_MIX = 0, Y4GPU_BLEND_ADD = 1, Y4GPU_BLEND_MULT = 2, Y4GPU_BLEND_SUB = 3, Y4GPU_BLEND_SCREEN = 4,
       Y4GPU_BLEND_DIV = 5, Y4GPU_BLEND_DIFF = 6, Y4GPU_BLEND_DARK = 7, Y4GPU_BLEND_LIGHT = 8 };
enum { Y4GPU_LAYER_RGB_TO_INT = 1, Y4GPU_LAYER_STENCIL = 2, Y4GPU_LAYER_NEGATIVE = 4 };
typedef struct y4gpu_pt_layer     /* one `layer` node together with the `texture_mapper` node on its input */
{
	/* texture_mapper (TextureMapperNode::factory, shader_node_basic.cc:306-367) */
	int32_t texture;              /* index into the texture array */
	int32_t texco;                /* Y4GPU_TEXCO_* */
	int32_t mapping;              /* Y4GPU_PROJ_* */
	int32_t proj_x, proj_y, proj_z;   /* 0..3 (default 1, 2, 3) */
	float scale[3], offset[3];    /* `offset` as given (the library doubles it like the reference) */
	/* layer (LayerNode::factory, shader_node_layer.cc:301-349) */
	int32_t flags;                /* Y4GPU_LAYER_*: noRGB, stencil, negative */
	int32_t blend_mode;           /* Y4GPU_BLEND_* */
	int32_t do_color;             /* default 1 */
	float colfac;                 /* default 1 */
	float def_col[3];             /* default 1, 1, 1 */
	float upper_color[4];         /* bottom layer only: `upper_color` (default 0, 0, 0, 0) */
	int32_t reserved[4];
} y4gpu_pt_layer;

/* pointlight (light_point.cc), arealight (light_area.cc), directional (light_directional.cc), sunlight (light_sun.cc),
   spherelight (light_sphere.cc; light sampling only, like the reference: canIntersect() is false),
   spotlight without soft_shadows (light_spot.cc: a dirac light with a smoothstep cone) */
enum { Y4GPU_LIGHT_POINT = 0, Y4GPU_LIGHT_AREA = 1, Y4GPU_LIGHT_DIRECTIONAL = 2, Y4GPU_LIGHT_SUN = 3, Y4GPU_LIGHT_SPHERE = 4, Y4GPU_LIGHT_SPOT = 5,
       Y4GPU_LIGHT_OBJECT = 6,     /* `objectlight` / `meshlight` (light_object_light.cc): the triangles of a mesh emit */
       Y4GPU_LIGHT_BACKGROUND = 7,
       Y4GPU_LIGHT_SPOT_SOFT = 8,
       Y4GPU_LIGHT_PORTAL = 9 };    /* `spotlight` with soft_shadows = true (light_spot.cc:106-144, 199-246): a sampled light; the
                                      fields of Y4GPU_LIGHT_SPOT plus samples and radius = `shadow_fuzzyness`
                                      Y4GPU_LIGHT_PORTAL: `bgPortalLight` (light_background_portal.cc): the triangles of a mesh (object_tris /
                                      object_n_tris as for Y4GPU_LIGHT_OBJECT, single sided) emit the background of y4gpu_pt_params seen
                                      through them, times `power`; the mesh itself must not be uploaded (the reference makes it invisible) */   /* `bglight` (light_background.cc): the IBL light a `constant` / `gradientback` background creates with
                                      ibl = true; it importance-samples the background of y4gpu_pt_params (tabulated on 360 latitude rows).
                                      samples = `ibl_samples`, cast_shadows, double_sided = `abs_intersect`; put it where its name
                                      ("constantBackground_bgLight" / "GradientBackground_bgLight") sorts among the lights */
typedef struct y4gpu_pt_light
{
	int32_t type;                 /* Y4GPU_LIGHT_* ; lights in the reference's (name-sorted) order */
	float from[3];                /* pointlight / spherelight / spotlight / finite directional `from`, arealight `corner` */
	float point1[3], point2[3];   /* arealight `point1`, `point2`; directional / sunlight: point1 = `direction` (towards the light); spotlight: point1 = `to` */
	float color[3];
	float power;
	int32_t samples;              /* arealight / sunlight / spherelight `samples` */
	int32_t cast_shadows;         /* default 1 */
	float radius;                 /* spherelight `radius`; directional `radius` when !infinite */
	float angle;                  /* sunlight `angle` (default 0.27) / spotlight `cone_angle` (default 45), degrees */
	int32_t infinite;             /* directional `infinite` (default 1) */
	float blend;                  /* spotlight `blend` (default 0.15) */
	int32_t double_sided;         /* objectlight `double_sided` (default 0) */
	uint32_t object_n_tris;       /* objectlight: number of triangles of the mesh named by `object_name` */
	const float *object_tris;     /* objectlight: their world-space vertices [object_n_tris * 9], in the mesh's face order (host memory, copied) */
} y4gpu_pt_light;

enum { Y4GPU_CAM_PERSPECTIVE = 0, Y4GPU_CAM_ARCHITECT = 1, Y4GPU_CAM_ORTHOGRAPHIC = 2, Y4GPU_CAM_EQUIRECTANGULAR = 3, Y4GPU_CAM_ANGULAR = 4 };
/* AngularCamera::Projection (camera_angular.h): the `projection` string of the angular camera's factory (camera_angular.cc:108-112) */
enum { Y4GPU_ANGULAR_EQUIDISTANT = 0, Y4GPU_ANGULAR_ORTHOGRAPHIC = 1, Y4GPU_ANGULAR_STEREOGRAPHIC = 2, Y4GPU_ANGULAR_EQUISOLID_ANGLE = 3,
       Y4GPU_ANGULAR_RECTILINEAR = 4 };
typedef struct y4gpu_pt_camera    /* camera parameters: `perspective` (camera_perspective.cc:182-220), `architect` (camera_architect.cc),
                                     `orthographic` (camera_orthographic.cc: `scale` travels in `focal`), `equirectangular`,
                                     `angular` (camera_angular.cc:80-117) */
{
	float from[3], to[3], up[3];
	int32_t resx, resy;
	float aspect_ratio, focal, aperture, dof_distance, near_clip, far_clip;
	int32_t bokeh_type;           /* lens shape when aperture != 0: 0 disk1 (default), 1 disk2, 3 triangle, 4 square, 5 pentagon, 6 hexagon, 7 ring */
	int32_t bokeh_bias;           /* 0 uniform, 1 center, 2 edge */
	float bokeh_rotation;         /* degrees (polygonal lens shapes) */
	int32_t type;                 /* Y4GPU_CAM_* */
	/* angular camera only (camera_angular.cc:84-112): the factory's double parameters as given; the radians, the focal length of
	   the projection (libm tan / FAST_TRIG sin) and max_radius = max_angle / angle are derived on the host like the constructor does */
	double angle_degrees, max_angle_degrees;
	int32_t circular, mirrored;   /* circular: pixels beyond max_radius get no ray (black, alpha 1); mirrored: vright negated */
	int32_t projection;           /* Y4GPU_ANGULAR_* */
} y4gpu_pt_camera;

enum { Y4GPU_FILTER_BOX = 0, Y4GPU_FILTER_GAUSS = 1, Y4GPU_FILTER_MITCHELL = 2, Y4GPU_FILTER_LANCZOS = 3 };
typedef struct y4gpu_pt_params
{
	/* integrator (integrator_path_tracer.cc:292-366) */
	int32_t bounces, path_samples, russian_roulette_min_bounces, no_recursive;   /* path_samples == 0: direct lighting only
	                                 (DirectLightIntegrator::integrate, integrator_direct_light.cc:94-143: emission + every light) */
	int32_t transp_background;
	float background[3];          /* constant background colour * power */
	/* render parameters (scene.cc:528-644, imagefilm.cc:45-127) */
	int32_t width, height, xstart, ystart;
	int32_t aa_passes, aa_samples, aa_inc_samples;
	float aa_sample_multiplier_factor, aa_light_sample_multiplier_factor;
	float aa_pixelwidth, aa_clamp_samples;
	int32_t filter_type;          /* Y4GPU_FILTER_* */
	int32_t shadow_bias_auto; float shadow_bias;        /* adv_auto_shadow_bias_enabled / adv_shadow_bias_value */
	int32_t min_raydist_auto; float min_raydist;        /* adv_auto_min_raydist_enabled / adv_min_raydist_value */
	uint32_t base_sampling_offset;                      /* adv_base_sampling_offset + 100000 * adv_computer_node */
	uint32_t rr_seed;             /* seed of the per-path Russian-roulette hash (the reference uses rand()) */
	uint32_t wave_size;           /* pixel-samples processed per wavefront (0 = default 2^23) */
	int32_t profile;              /* 1: time every kernel class with CUDA events and count BVH visits (slower; for the roofline) */
	int32_t transp_shadows;       /* integrator transpShad: shadow rays take the intersectTs rules (material visibility only,
	                                 tmin <= t < t_max; accelerator_kdtree.cc:911-1061): opaque blockers shadow, transparent
	                                 shinydiffuse blockers multiply their getTransparency colour into the light sample */
	int32_t shadow_depth;         /* integrator shadowDepth: distinct transparent blockers a shadow ray may cross (<= 16, else UNSUPPORTED) */
	int32_t raydepth;             /* integrator `raydepth` (default 5): specular recursion levels (integrator_montecarlo.cc:916-964);
	                                 <= 19, else UNSUPPORTED */
	int32_t trace_caustics;       /* caustic_type path / both: emission found after a specular bounce of a path counts
	                                 (integrator_path_tracer.cc:227,256-264) */
	int32_t transp_refracted_background;   /* bg_transp_refract (TiledIntegrator::background, integrator_tiled.cc:707-720) */
	int32_t background_gradient;  /* 0: `constant` background (colour in `background`); 1: `gradientback` (background_gradient.cc:34-46);
	                                 2: `textureback` (background_texture.cc), described by the bg_texture* fields below;
	                                 3: `sunsky` (background_sunsky.cc), described by the bg_sun_* fields below;
	                                 4: `darksky` (background_darksky.cc): bg_sun_from / turbidity / var / power + the bg_dark_* fields */
	float bg_gradient[4][3];      /* gradientback colours * power: zenith_ground, horizon_ground, zenith, horizon */
	/* ambient occlusion of the direct-lighting integrator (do_AO, AO_samples, AO_distance, AO_color; integrator_direct_light.cc:124,
	   TiledIntegrator::sampleAmbientOcclusion, integrator_tiled.cc:644-692): after the lights, every diffuse vertex integrate() reaches
	   takes ao_samples BSDF-sampled occlusion rays of length ao_distance.  Used when path_samples == 0 only -- the reference's path
	   tracer parses the parameters but never adds the term to the Combined layer.  ao_samples <= 0: off. */
	int32_t ao_samples;
	float ao_distance;
	float ao_color[3];
	/* textureback (TextureBackground::factory, background_texture.cc:80-151): an image texture seen in every direction */
	int32_t bg_texture;           /* index into the array given to y4gpu_pt_set_textures (which therefore comes first) */
	int32_t bg_texture_angular;   /* mapping "probe" / "angular": Texture::angMap; otherwise spherical (Texture::sphereMap) */
	float bg_texture_rotation;    /* `rotation`, degrees */
	float bg_texture_power;       /* `power` */
	float bg_ibl_clamp;           /* `ibl_clamp_sampling`: BackgroundLight::intersect clamps the colour proportionally (0: off).
	                                 `smartibl_blur` (mip-mapped lookups for the light) is not supported: callers refuse it */
	/* sunsky (SunSkyBackground::factory, background_sunsky.cc:226-307): Preetham daylight model.  `add_sun` is the caller's business
	   (it creates an ordinary sunlight); `background_light` arrives as a Y4GPU_LIGHT_BACKGROUND like the other IBL lights */
	float bg_sun_from[3];         /* `from`: direction towards the sun (default 1 1 1) */
	float bg_sun_turbidity;       /* `turbidity` (default 4) */
	float bg_sun_var[5];          /* `a_var` .. `e_var` (default 1) */
	float bg_sun_power;           /* `power` (default 1) */
	/* darksky (DarkSkyBackground::factory, background_darksky.cc:194-314): the same model with altitude, exposure, a colour-space
	   matrix, gamma encoding and a night mode (which halves `bright`: the caller passes the parameter as given) */
	float bg_dark_altitude;       /* `altitude` (default 0) */
	float bg_dark_bright;         /* `bright` (default 1) */
	float bg_dark_exposure;       /* `exposure` (default 1) */
	int32_t bg_dark_night;        /* `night` */
	int32_t bg_dark_color_space;  /* `color_space`: 0 "CIE (E)" (default, also for unknown names), 1 "CIE (D50)", 2 "sRGB (D65)", 3 "sRGB (D50)" */
} y4gpu_pt_params;

typedef struct y4gpu_pt_stats
{
	uint64_t samples;             /* pixel-samples rendered by the last pass */
	uint64_t rays_closest, rays_shadow;   /* rays actually traced (camera rays + live bounce rays; live shadow rays) */
	uint64_t kernel_launches;     /* kernels launched by the last pass */
	double pass_ms;               /* device time of the last pass (CUDA events) */
	/* filled only when y4gpu_pt_params.profile != 0 */
	double class_ms[8];           /* 0 camera rays, 1 closest-hit traversal, 2 shadow traversal, 3 camera-hit / sub-path start /
	                                 per-bounce path stages, 4 extra light-sample stages, 5 film splat */
	uint64_t class_launches[8];
	uint64_t nodes_closest, tris_closest, nodes_shadow, tris_shadow;   /* BVH8 nodes visited / triangles tested */
	uint64_t live_closest, live_shadow;                                /* rays with a non-empty interval */
	uint64_t camera_hits;         /* pixel-samples whose camera ray hit a surface and went on (queue of the start stage) */
	uint64_t path_entries;        /* entries processed by the per-bounce path stages (all bounces, incl. the closing call) */
	uint64_t reserved[2];
} y4gpu_pt_stats;

/* Noise-detection parameters of adaptive anti-aliasing (AaNoiseParams, include/common/aa_noise_params.h:27-46;
 * render parameters AA_threshold, AA_detect_color_noise, AA_dark_detection_type, AA_dark_threshold_factor,
 * AA_variance_edge_size, AA_variance_pixels; scene.cc:586-595). */
typedef struct y4gpu_pt_aa_noise
{
	float threshold;              /* current AA threshold (the host adapts it between passes like TiledIntegrator::render) */
	int32_t detect_color_noise;
	int32_t dark_detection_type;  /* 0 none, 1 linear, 2 curve (ImageFilm::darkThresholdCurveInterpolate) */
	float dark_threshold_factor;
	int32_t variance_edge_size, variance_pixels;
	int32_t reserved[6];
} y4gpu_pt_aa_noise;

/* Per-triangle material index travels in the flag word of y4gpu_upload_triangles (bits 8..31).
 * vertex_normals: n_tri*9 floats (exported or smoothed normals at the three corners: shading normal interpolated like
 * primitive_triangle.cc:104-114; a NaN in a triangle's first float = flat-shaded triangle) or NULL for flat shading. */
/* Textures, layer chains and the per-triangle texture coordinates they read; call after y4gpu_build and BEFORE
 * y4gpu_pt_set_scene (materials refer to layers by index).  tri_orco: n_tri*9 floats, the three orco points of every
 * triangle (yafaray_addVertexWithOrco), a NaN in the first float marks a triangle of a mesh without orco points
 * (orco = hit point, primitive_triangle.cc:122-127); tri_uv: n_tri*6 floats (yafaray_addUv / addTriangleWithUv), NaN
 * = no uv (implicit barycentric mapping).  Either may be NULL.  n_tex == 0 and n_layers == 0 removes everything. */
int y4gpu_pt_set_textures(y4gpu_ctx *ctx, const y4gpu_pt_texture *textures, uint32_t n_tex, const y4gpu_pt_layer *layers,
                          uint32_t n_layers, const float *tri_orco, const float *tri_uv);
int y4gpu_pt_set_scene(y4gpu_ctx *ctx, const y4gpu_pt_camera *cam, const y4gpu_pt_material *mats, uint32_t n_mat,
                       const y4gpu_pt_light *lights, uint32_t n_light, const y4gpu_pt_params *params,
                       const float *vertex_normals);
/* Optional: accumulate into caller-owned device buffers (weights: h*w floats, rgba: h*w*4 floats), e.g.
 * a tensor that is then summed across GPUs with one NCCL all-reduce per pass.  NULL = internal film. */
int y4gpu_pt_set_film_buffers(y4gpu_ctx *ctx, float *d_weights, float *d_rgba);
int y4gpu_pt_clear_film(y4gpu_ctx *ctx);
/* Renders samples [offset, offset + n_samples) of every pixel (or of the listed pixels: pixel_list
 * holds y*width+x entries of the pixels ImageFilm::doMoreSamples flags, host or device memory).
 * This GPU renders the sample indices s with (s - offset) % shard_count == shard_index. */
int y4gpu_pt_render_pass(y4gpu_ctx *ctx, int n_samples, int offset, int aa_pass_number,
                         const uint32_t *pixel_list, uint64_t n_pixels, int shard_index, int shard_count);
/* Film as accumulated (NOT normalised): weights h*w, rgba h*w*4; host or device destinations. */
/* ImageFilm::nextPass(adaptive_aa = true) on the device (imagefilm.cc:259-445): flags the pixels whose normalised colour
 * differs from a neighbour's by at least the threshold (plus never-sampled pixels and the variance window rule) and
 * compacts them, in raster order, into a device-resident pixel list for y4gpu_pt_render_pass.
 * d_weights / d_rgba: the ACCUMULATED film to analyse (NULL = the context's film; multi-GPU callers pass the all-reduced
 * total).  *d_pixel_list stays valid until the next call; *n_resample is ImageFilm::nextPass's return value. */
int y4gpu_pt_next_pass(y4gpu_ctx *ctx, const y4gpu_pt_aa_noise *aa, const float *d_weights, const float *d_rgba,
                       const uint32_t **d_pixel_list, uint64_t *n_resample);
int y4gpu_pt_read_film(y4gpu_ctx *ctx, float *weights, float *rgba);
/* Replaces the accumulated film (weights [h*w], rgba [h*w*4]; host or device memory): what ImageFilm::imageFilmLoad /
 * imageFilmLoadAllInFolder leave in weights_ and the Combined layer before a resumed render (imagefilm.cc:827-1018). */
int y4gpu_pt_write_film(y4gpu_ctx *ctx, const float *weights, const float *rgba);
int y4gpu_pt_get_stats(const y4gpu_ctx *ctx, y4gpu_pt_stats *out);

/* Self-check of the device's fp32 inverse trigonometric functions (csrc/y4_libm.cuh), which stand in for the C library calls of
 * the reference (std::atan2 / std::atan / std::asin in AngularCamera::shootRay, camera_angular.cc:66-71; std::acos in
 * Texture::sphereMap, math.h:253-259).  out[i] = f(a[i]) for fn 0 atanf, 1 asinf, 2 acosf; out[i] = atan2f(a[i], b[i]) for fn 3.
 * Host or device pointers; the caller compares with its own libm, bit for bit (tests/test_libm.py). */
int y4gpu_libm_eval(y4gpu_ctx *ctx, int fn, const float *a, const float *b, float *out, uint64_t n);

#ifdef __cplusplus
}
#endif
#endif /* Y4GPU_H */
