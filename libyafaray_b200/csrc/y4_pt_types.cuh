// Device-side scene description and per-path state of the wavefront path tracer.
// Everything is derived on the host (y4_pt_host.h) from the parameter-level C-ABI structs with
// the reference's own fp32 operation order.  Y4PtScene travels to every kernel by value (kernel
// parameter space = constant bank); the arrays behind its pointers are uploaded once per
// y4gpu_pt_set_scene.
#pragma once
#include "y4_common.cuh"
#include "y4_pt_sampling.cuh"
#include "../../include/y4gpu.h"

struct C3 { float r, g, b; };
Y4_HD C3 c3(float r, float g, float b) { C3 c; c.r = r; c.g = g; c.b = b; return c; }
Y4_HD C3 c3f(float v) { return c3(v, v, v); }
// Rgb operators of include/color/color.h:258-295 (component-wise, float * Rgb is {f*r, f*g, f*b})
Y4_HD C3 cmul(C3 a, C3 b) { return c3(xmul(a.r, b.r), xmul(a.g, b.g), xmul(a.b, b.b)); }
Y4_HD C3 cscale(C3 a, float f) { return c3(xmul(f, a.r), xmul(f, a.g), xmul(f, a.b)); }
Y4_HD C3 cdiv(C3 a, float f) { return c3(xdiv(a.r, f), xdiv(a.g, f), xdiv(a.b, f)); }
Y4_HD C3 cadd(C3 a, C3 b) { return c3(xadd(a.r, b.r), xadd(a.g, b.g), xadd(a.b, b.b)); }
Y4_HD bool cblack(C3 a) { return a.r == 0.f && a.g == 0.f && a.b == 0.f; }
Y4_HD float cmax(C3 a) { return fmaxf(a.r, fmaxf(a.g, a.b)); }

// BsdfFlags bits (include/material/material.h:44-62)
enum : uint32_t
{
	Y4_BSDF_SPECULAR = 1u << 0, Y4_BSDF_GLOSSY = 1u << 1, Y4_BSDF_DIFFUSE = 1u << 2, Y4_BSDF_DISPERSIVE = 1u << 3,
	Y4_BSDF_REFLECT = 1u << 4, Y4_BSDF_TRANSMIT = 1u << 5, Y4_BSDF_FILTER = 1u << 6, Y4_BSDF_EMIT = 1u << 7,
	Y4_BSDF_ALL = 0x7fu
};

struct Y4PtCamera          // PerspectiveCamera state after ctor + setAxis (camera.cc:51-71, camera_perspective.cc:50-69)
{
	V3 pos, vright, vup, vto;
	V3 near_p, near_n, far_p, far_n;
	int resx, resy;
	// depth of field (aperture != 0): lens offsets dof_rt * u + dof_up * v, focus at dof_distance (camera_perspective.cc:58-59,136-143)
	V3 dof_rt, dof_up;
	float aperture, dof_distance;
	int type;                   // Y4GPU_CAM_*: perspective and architect share shootRay, they differ in vup / vto
	int bkhtype, bkhbias;       // BokehType / BkhBiasType (camera_perspective.h:38-39)
	float ls[16];               // polygon corner directions of the tri / square / penta / hexa lenses (ctor, :40-52)
	// AngularCamera state (camera_angular.cc:30-44): aspect_ratio_, focal_length_, max_radius_, circular_, projection_
	float ang_aspect, ang_focal, ang_max_radius;
	int ang_circular, ang_projection;
};

struct Y4PtMaterial
{
	int type;                   // Y4GPU_MAT_*
	uint32_t bsdf_flags;        // Material::bsdf_flags_
	C3 color;                   // diffuse_color_ / light_col_ (colour * power)
	C3 emit;                    // emit_color_ = emit_strength * diffuse_color
	float diffuse_strength;     // components_[3] (0 when !is_diffuse_)
	int is_diffuse;
	int oren_nayar; float oren_a, oren_b;
	int double_sided, receive_shadows, flat_material;
	int layer_first, layer_count;   // diffuse_shader chain in Y4PtScene::layers (0 layers: plain colour)
	float emit_strength;            // emit() with a diffuse_shader: shader colour * emit_strength_
	// shinydiffusemat's perfect-mirror component (components_[0]): specular_reflect, mirror_color, Fresnel weighting
	int is_mirror; float mirror_strength; C3 mirror_color;
	int fresnel; float ior_squared;
	// shinydiffusemat's straight-through transparency component (components_[1]): transparency, transmit_filter
	int is_transparent; float transparency_strength, transmit_filter;
	int is_translucent; float translucency_strength;              // components_[2]: diffuse transmission
	// glass (material_glass.cc): ior_, filter_color_ = filt * filter_color + (1 - filt) (kept in `color`), specular_reflection_color_ (in mirror_color)
	float ior;
	int absorb; C3 sigma_a;     // glass with `absorption`: BeerVolumeHandler::sigma_a_ of the inside volume handler
};

// ImageTexture state (texture_image.cc:476-597) + texels as Image::getColor returns them
struct Y4PtTexture
{
	const float4 *texels;
	int w, h, clip, interp;
	int xrepeat, yrepeat, mirror_x, mirror_y, rot90;
	int cropx, cropy; float cropminx, cropmaxx, cropminy, cropmaxy;
	int checker_even, checker_odd; float checker_dist;
	int adjustments_set, adj_clamp; float adj_intensity, adj_contrast, adj_red, adj_green, adj_blue;
};
// TextureBackground state (background_texture.cc:30-42): the image, the projection, rotation_ and its sine / cosine, power_
struct Y4BgTex
{
	const Y4PtTexture *tex;
	int angular;                    // mapping probe / angular: Texture::angMap, else Texture::sphereMap
	float rotation, sin_r, cos_r, power;
};
// SunSkyBackground state (background_sunsky.cc:39-80) reduced to what eval() reads per direction: the sun's angles, the Perez
// coefficients, the zenith values and the direction-independent denominators of perezFunction
struct Y4BgSunsky
{
	double theta_s, phi_s;
	float sin_theta_s, cos_theta_s;     // math::sin / math::cos of theta_s_ (FAST_TRIG, float argument)
	int night;                          // theta_s_ > pi / 2
	double nfade_sun;                   // 1 - (theta_s_ / pi - 0.5) * 2: the sun's share of the night fade
	double perez_x[5], perez_y[5], perez_Y[5];
	double zenith_x, zenith_y, zenith_Y;
	double den_x, den_y, den_Y;
	float power;
};
// DarkSkyBackground state (background_darksky.cc:40-100) + its ColorConv (color_conversion.h): the altitude-shifted sun direction,
// the Perez coefficients with their pre-computed reciprocal denominators in [5], the zenith values, the XYZ -> RGB matrix
struct Y4BgDarksky
{
	V3 sun_dir; float alt; int night;
	double perez_x[6], perez_y[6], perez_Y[6];
	double zenith_x, zenith_y, zenith_Y;
	float exposure, mat[9];
	float bright, power;                // sky_brightness_, power_ = power * bright
};
// TextureMapperNode + LayerNode state of one layer of a diffuse_shader chain
struct Y4PtLayer
{
	int tex, coords, projection, map_x, map_y, map_z;
	V3 scale, offset;               // offset already doubled (shader_node_basic.cc:361)
	uint32_t flags; int blend, do_color; float colfac;
	float def_col[4], upper_col[4];
};

struct Y4PtLight
{
	int type;                   // Y4GPU_LIGHT_*
	int samples, cast_shadows;
	C3 color;                   // point, directional, sphere: col*inte ; area: col*inte*pi ; sun: col_pdf_ = col*inte*pdf_
	V3 pos;                     // point position / area corner / sphere centre / finite directional position
	V3 to_x, to_y, fnormal, normal, c2, c3, c4;
	float area, inv_area;
	// directional (light_directional.cc:28-39), sun (light_sun.cc:29-43), sphere (light_sphere.cc:30-40)
	V3 dir, du, dv;             // normalised direction towards the light; sun: frame of the *unnormalised* ctor argument
	float cos_angle, pdf, invpdf;
	float cos_start, cos_end, icos_diff;   // spot (light_spot.cc:30-43); its ndir_ is kept in `dir`
	float radius, sq_radius, sq_radius_eps;
	int infinite;
	// objectlight (light_object_light.cc): the mesh's triangles, geometric normals and the area distribution's cdf
	int double_sided, obj_n;
	const float *obj_tris;      // [obj_n * 9] world-space vertices
	const float4 *obj_ng;       // [obj_n] Primitive::getGeometricNormal
	const float *obj_cdf;       // [obj_n] Pdf1D::cdf_ of the triangle areas
	// background light (light_background.cc:46-75): the background tabulated on bg_nv latitude rows of bg_off[y + 1] - bg_off[y]
	// columns; row bg_nv of the tables is the distribution over the rows.  func = Pdf1D::function_, cdf = Pdf1D::cdf_.
	int bg_nv, bg_abs_inter;
	const int *bg_off;          // [bg_nv + 2]
	const float *bg_func, *bg_cdf, *bg_inv_int;     // bg_inv_int[y] = Pdf1D::invIntegral() of row y
	int bg_gradient; C3 bg_col, bg_gz, bg_gh, bg_sz, bg_sh;     // the background itself (Background::eval); bg_gradient == 2: bg_tex
	Y4BgTex bg_tex;
	const Y4BgSunsky *bg_sun;   // bg_gradient == 3: sunsky
	const Y4BgDarksky *bg_dark; // bg_gradient == 4: darksky
	float bg_clamp;             // clamp_intersect_ (ibl_clamp_sampling of a textureback; 0: off)
	float power;                // portal light: bg_->eval(dir) * power_
};
// Light::diracLight() / Light::canIntersect() of the supported lights (include/light/light_*.h)
Y4_HD bool y4_light_dirac(int type) { return type == Y4GPU_LIGHT_POINT || type == Y4GPU_LIGHT_DIRECTIONAL || type == Y4GPU_LIGHT_SPOT; }
Y4_HD bool y4_light_can_intersect(int type) { return type == Y4GPU_LIGHT_AREA || type == Y4GPU_LIGHT_SUN || type == Y4GPU_LIGHT_OBJECT || type == Y4GPU_LIGHT_BACKGROUND || type == Y4GPU_LIGHT_SPOT_SOFT || type == Y4GPU_LIGHT_PORTAL; }

#define Y4_PT_MAX_LIGHTS 64
#define Y4_FILTER_TABLE 16

struct Y4PtScene
{
	Y4PtCamera cam;
	// integrator
	int bounces, n_paths, rr_min_bounces, no_recursive, transp_background;
	int raydepth, trace_caustics, transp_refracted_background;
	int tshadow_filter;             // transpShad with transparent materials: shadow hits carry the blockers' filter colour (t, bu, bv lanes)
	int has_specular;               // ray-level loop (recursiveRaytrace) needed: 1 pure mirrors only (chains run inside the camera-hit
	                                // step), 2 shinydiffuse with a mirror component (one full estimate per ray level)
	C3 background;
	int bg_gradient; C3 bg_gz, bg_gh, bg_sz, bg_sh;     // gradientback: ground zenith / horizon, sky zenith / horizon
	Y4BgTex bg_tex;                                     // bg_gradient == 2: textureback
	const Y4BgSunsky *bg_sun;                           // bg_gradient == 3: sunsky
	const Y4BgDarksky *bg_dark;                         // bg_gradient == 4: darksky
	float shadow_bias; int shadow_bias_auto;
	float ray_min_dist; int ray_min_dist_auto;
	uint32_t base_sampling_offset, rr_seed;
	float aa_light_sample_multiplier;
	int aa_passes;
	int ao_samples; float ao_dist; C3 ao_col;     // ambient occlusion of the direct-lighting integrator (0 samples: off); sampled as pseudo-light n_lights
	// film
	int width, height, cx0, cy0;
	double filterw, table_scale;
	float clamp_samples;
	float filter_table[Y4_FILTER_TABLE * Y4_FILTER_TABLE];
	// scene content
	int n_lights, n_mats;
	const Y4PtLight *lights;        // [n_lights]
	const Y4PtMaterial *mats;       // [n_mats]
	const float4 *tri_shade;        // [n_tri] geometric normal (xyz) + material index (w, as int bits), see y4_tri_shade
	const float *normals;           // [n_tri * 9] exported vertex normals or null
	const Y4PtTexture *textures; const Y4PtLayer *layers;     // diffuse_shader chains (y4_pt_tex.cuh) or null
	const float *tri_orco, *tri_uv; // [n_tri * 9] orco points / [n_tri * 6] uv values per triangle, or null
	Y4LdsTables lds;                // Faure tables (device pointers)
};

// Surface point: the subset of SurfacePoint the supported materials/lights read
// (include/geometry/surface.h:86-120, filled like primitive_triangle.cc:97-176).
struct Y4Surf
{
	V3 p, ng, n, nu, nv;
	uint32_t mat;
};

// per-path flag bits
enum : uint32_t
{
	Y4_PF_ALIVE = 1u,            // path still being extended in the bounce loop
	Y4_PF_HIT0 = 2u,             // camera ray hit a surface
	Y4_PF_PATHS = 4u,            // primary material matches path_flags -> bounce loop runs
	Y4_PF_DIFF1 = 8u,            // first path vertex's material has Diffuse (mat_bsd_fs, integrator_path_tracer.cc:196)
	Y4_PF_DIRECT = 16u,          // the current vertex wants a one-light estimate this depth
	Y4_PF_S0 = 32u, Y4_PF_S1 = 64u,  // shadow-ray slot 0/1 of the current light sample was traced
	Y4_PF_C1 = 128u,             // W.scontrib slot 1 (the BSDF-sampled MIS estimator) holds a non-zero contribution; unset: it is zero and was not stored
	Y4_PF_LCOL = 256u,           // W.lcol holds non-zero emission picked up at the current vertex; unset: zero, not stored
	Y4_PF_SPEC = 512u,           // the vertex integrate() works on was reached through mirrors (ray_level > 0): W.ray0 is the ray
	                             // into it, W.sw the product of the reflection colours (xyz) and the ray level (w)
	Y4_PF_EMIT1 = 1024u,         // first path vertex's material has Emit (mat_bsd_fs, integrator_path_tracer.cc:196,256)
	Y4_PF_RECURSE = 4096u,       // the vertex also has a perfect-mirror component and a ray level left: after its own estimate is
	                             // complete the level-finish stage folds it into W.acc and reflects the ray (shinydiffusemat)
	Y4_PF_ACC = 8192u,           // W.acc holds the radiance gathered at the upper ray levels of this pixel-sample
	Y4_PF_CAUSTIC = 2048u        // the ray into the current path vertex was sampled from a specular lobe and caustic_type traces them
};
