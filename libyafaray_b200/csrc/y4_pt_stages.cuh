// Wavefront stages of the path tracer, one call per path (Y4_HD: the kernels in y4_pt.cu and the
// CPU unit-test emulation run the same code).  Control flow restated from
//   TiledIntegrator::renderTile        src/integrator/surface/integrator_tiled.cc:269-408
//   PerspectiveCamera::shootRay        src/camera/camera_perspective.cc:128-146
//   PathIntegrator::integrate          src/integrator/surface/integrator_path_tracer.cc:120-290
//   estimateAllDirectLight / estimateOneDirectLight / doLightEstimation
//                                      src/integrator/surface/integrator_montecarlo.cc:54-78,385-408
//   ImageFilm::addSample               src/render/imagefilm.cc:680-733
#pragma once
#include "y4_pt_shade.cuh"
#include "y4_pt_tex.cuh"

struct Y4PtWave
{
	uint32_t n;            // pixel-samples in this wave
	uint32_t *pix;         // (y - cy0) * width + (x - cx0)
	uint32_t *samp;        // sample loop index inside the pass (0 .. n_samples-1)
	uint32_t *flags;       // Y4_PF_*
	uint32_t *lnum;        // light chosen by estimateOneDirectLight for the current vertex
	float *col;            // [n*4] radiance of the pixel-sample (rgb) + alpha
	float *pcol;           // [n*4] path_col (sum over the path_samples sub-paths)
	float *thr;            // [n*4] throughput arriving at the current vertex
	float *pthr;           // [n*4] throughput at the previous vertex after Russian roulette (its light estimate is folded in one stage later)
	float *lcol;           // [n*4] emitted light picked up at the current vertex
	float *acc_a, *acc_b;  // [n*4] per-light sums of the light-sample / BSDF-sample estimators
	float *lsamp;          // [n*4] last light sample drawn for this pixel-sample: s1, s2, light, k (the reference draws
	                       //       the same numbers at every vertex of a pixel-sample, see y4_light_sample_numbers)
	Y4Ray *ray;            // ray that produced the current vertex
	Y4Ray *ray_next;       // ray leaving the current vertex (the host alternates the two buffers per bounce)
	Y4Hit *hit; Y4Hit *hit0;
	Y4Ray *sray;           // [2n] shadow rays: slot 0 at [i], slot 1 at [n + i]
	Y4Hit *shit;           // [2n]
	float *scontrib;       // [2n*4]
	Y4Ray *ray0;           // [n] ray_level > 0 (Y4_PF_SPEC): the reflected ray that reached the vertex integrate() works on
	float *acc;            // [n*4] Y4_PF_ACC: radiance of the upper ray levels, already weighted
	float *sw;             // [n*4] Y4_PF_SPEC: product of the reflection colours down the mirror chain (xyz), ray level (w, int bits)
	// ray TREE (has_specular == 3: glass, or shinydiffusemat with mirror AND transparency -- two children per vertex): the
	// pending sibling branches of a pixel-sample, depth first: entry k of slot i at [k * n + i]
	Y4Ray *stk_ray;        // [depth * n]
	float *stk_sw;         // [depth * n * 4] branch weight (xyz), ray level (w, int bits)
	uint32_t *stk_n;       // [n] entries on the stack
};

struct Y4PtPass
{
	int n_samples;         // samples per pixel in this pass (all shards together)
	uint32_t pass_offs;    // offset + ImageFilm::getBaseSamplingOffset()
	int aa_pass;
};

// 16-byte state accesses (one 128-bit load/store each)
Y4_HD float4 ld4(const float *p, uint32_t i) { return *reinterpret_cast<const float4 *>(p + 4ull * i); }
Y4_HD void st4(float *p, uint32_t i, float4 v) { *reinterpret_cast<float4 *>(p + 4ull * i) = v; }
Y4_HD float4 f4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
Y4_HD C3 ld3(const float *p, uint32_t i) { const float4 v = ld4(p, i); return c3(v.x, v.y, v.z); }
Y4_HD void st3(float *p, uint32_t i, C3 c) { st4(p, i, f4(c.r, c.g, c.b, 0.f)); }     // lane 3 unused except in `col` (alpha)
Y4_HD Y4Ray ld_ray(const Y4Ray *p, uint32_t i)
{
	const float4 a = reinterpret_cast<const float4 *>(p + i)[0], b = reinterpret_cast<const float4 *>(p + i)[1];
	Y4Ray r; r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w; return r;
}
Y4_HD void st_ray(Y4Ray *p, uint32_t i, const Y4Ray &r)
{
	reinterpret_cast<float4 *>(p + i)[0] = f4(r.ox, r.oy, r.oz, r.tmin); reinterpret_cast<float4 *>(p + i)[1] = f4(r.dx, r.dy, r.dz, r.tmax);
}
Y4_HD Y4Hit ld_hit(const Y4Hit *p, uint32_t i)
{
	const float4 a = *reinterpret_cast<const float4 *>(p + i);
	Y4Hit h; h.t = a.x; h.bu = a.y; h.bv = a.z; h.prim = (int32_t) y4_f2u(a.w); return h;
}
Y4_HD void st_hit(Y4Hit *p, uint32_t i, const Y4Hit &h) { *reinterpret_cast<float4 *>(p + i) = f4(h.t, h.bu, h.bv, y4_u2f((uint32_t) h.prim)); }

// PixelSamplingData (integrator_tiled.cc:311-330): pixel, per-pixel hash offset, sample number
struct Y4PixelId { int x, y; uint32_t offset; uint32_t sample; };
Y4_HD Y4PixelId y4_pixel_id(const Y4PtScene &sc, const Y4PtPass &ps, uint32_t pix, uint32_t samp)
{
	Y4PixelId r;
	const uint32_t row = pix / (uint32_t) sc.width;
	r.x = sc.cx0 + (int) (pix - row * (uint32_t) sc.width);
	r.y = sc.cy0 + (int) row;
	r.offset = y4_fnv32a((uint32_t) r.y * y4_fnv32a((uint32_t) r.x));
	r.sample = ps.pass_offs + samp;
	return r;
}

struct Y4PixelSampling { int x, y; uint32_t offset; uint32_t sample; float dx, dy; };   // + the sample's jitter inside the pixel
Y4_HD Y4PixelSampling y4_pixel_sampling(const Y4PtScene &sc, const Y4PtPass &ps, uint32_t pix, uint32_t samp)
{
	const Y4PixelId id = y4_pixel_id(sc, ps, pix, samp);
	Y4PixelSampling r;
	r.x = id.x; r.y = id.y; r.offset = id.offset; r.sample = id.sample;
	r.dx = 0.5f; r.dy = 0.5f;
	if(sc.aa_passes > 1)
	{
		r.dx = y4_ri_vdc(r.sample, r.offset);
		r.dy = y4_ri_s(r.sample, r.offset);
	}
	else if(ps.n_samples > 1)
	{
		const float d_1 = xdiv(1.f, (float) ps.n_samples);
		r.dx = xmul(xadd(0.5f, (float) (int) samp), d_1);
		r.dy = y4_ri_lp(samp + r.offset, 0u);
	}
	return r;
}

// Vec3::shirleyDisk (src/geometry/vector.cc:128-165): concentric square -> disk map; math::div_pi_by_4 is a long double
// constant, so the angle is formed wide and rounded once; cos/sin are the FAST_TRIG approximations
Y4_HD void y4_shirley_disk(float r_1, float r_2, float &u, float &v)
{
	float phi = 0.f, r = 0.f;
	const float a = xsub(xmul(2.f, r_1), 1.f), b = xsub(xmul(2.f, r_2), 1.f);
	if(a > -b)
	{
		if(a > b) { r = a; phi = (float) xdmul(Y4_PI_4, (Y4_WIDE) xdiv(b, a)); }
		else { r = b; phi = (float) xdmul(Y4_PI_4, (Y4_WIDE) xsub(2.f, xdiv(a, b))); }
	}
	else
	{
		if(a < b) { r = -a; phi = (float) xdmul(Y4_PI_4, (Y4_WIDE) xadd(4.f, xdiv(b, a))); }
		else
		{
			r = -b;
			phi = b != 0.f ? (float) xdmul(Y4_PI_4, (Y4_WIDE) xsub(6.f, xdiv(a, b))) : 0.f;
		}
	}
	u = xmul(r, y4_fast_cos(phi));
	v = xmul(r, y4_fast_sin(phi));
}

// PerspectiveCamera::biasDist / sampleTsd / getLensUv (camera_perspective.cc:71-126): lens sample -> lens coordinates
Y4_HD float y4_bias_dist(int bias, float r)
{
	if(bias == 1) return xsqrt(xmul(xsqrt(r), r));
	if(bias == 2) return xsqrt(xsub(1.f, xmul(r, r)));
	return xsqrt(r);
}
Y4_HD void y4_lens_uv(const Y4PtCamera &cam, float r_1, float r_2, float &u, float &v)
{
	const int t = cam.bkhtype;
	if(t >= 3 && t <= 6)
	{
		const float fn = (float) t;
		int idx = (int) xmul(r_1, fn);
		r_1 = xmul(xsub(r_1, xdiv((float) idx, fn)), fn);
		r_1 = y4_bias_dist(cam.bkhbias, r_1);
		const float b_1 = xmul(r_1, r_2), b_0 = xsub(r_1, b_1);
		idx <<= 1;
		u = xadd(xmul(cam.ls[idx], b_0), xmul(cam.ls[idx + 2], b_1));
		v = xadd(xmul(cam.ls[idx + 1], b_0), xmul(cam.ls[idx + 3], b_1));
	}
	else if(t == 1 || t == 7)
	{
		const float w = xmul((float) Y4_2PI, r_2);
		if(t == 7) r_1 = xsqrt(xadd(0.707106781f, 0.292893218f));
		else r_1 = y4_bias_dist(cam.bkhbias, r_1);
		u = xmul(r_1, y4_fast_cos(w));
		v = xmul(r_1, y4_fast_sin(w));
	}
	else y4_shirley_disk(r_1, r_2, u, v);
}

// A camera that refuses a pixel position (CameraRay::valid_ == false: outside the circle of a circular angular camera) hands
// back a ray with a zero direction and an empty interval: nothing can hit it, and the camera-hit stage turns it into the
// sample renderTile adds for it, the layer's default colour (0, 0, 0, alpha 1: integrator_tiled.cc:343-347, layer_definitions.h:118).
Y4_HD Y4Ray y4_invalid_camera_ray(const V3 from) { return y4_mk_ray(from, 1.f, v3(0.f, 0.f, 0.f), 0.f); }
Y4_HD bool y4_camera_ray_invalid(const Y4Ray &r) { return r.dx == 0.f && r.dy == 0.f && r.dz == 0.f; }

// AngularCamera::shootRay (camera_angular.cc:57-78).  std::atan2 / std::atan / std::asin are the C library's: y4_libm.cuh
Y4_HD Y4Ray y4_angular_camera_ray(const Y4PtCamera &cam, float px, float py)
{
	const float u = xsub(1.f, xmul(2.f, xdiv(px, (float) cam.resx)));
	float v = xsub(xmul(2.f, xdiv(py, (float) cam.resy)), 1.f);
	v = xmul(v, cam.ang_aspect);
	const float radius = xsqrt(xadd(xmul(u, u), xmul(v, v)));
	if(cam.ang_circular && radius > cam.ang_max_radius) return y4_invalid_camera_ray(cam.pos);
	float theta = 0.f;
	if(!(u == 0.f && v == 0.f)) theta = y4_libm_atan2f(v, u);
	float phi;
	if(cam.ang_projection == Y4GPU_ANGULAR_ORTHOGRAPHIC) phi = y4_asin(xdiv(radius, cam.ang_focal));
	else if(cam.ang_projection == Y4GPU_ANGULAR_STEREOGRAPHIC) phi = xmul(2.f, y4_libm_atanf(xdiv(radius, xmul(2.f, cam.ang_focal))));
	else if(cam.ang_projection == Y4GPU_ANGULAR_EQUISOLID_ANGLE) phi = xmul(2.f, y4_asin(xdiv(radius, xmul(2.f, cam.ang_focal))));
	else if(cam.ang_projection == Y4GPU_ANGULAR_RECTILINEAR) phi = y4_libm_atanf(xdiv(radius, cam.ang_focal));
	else phi = xdiv(radius, cam.ang_focal);
	const V3 in_plane = xadd3(xscale(cam.vright, y4_fast_cos(theta)), xscale(cam.vup, y4_fast_sin(theta)));
	const V3 d = xadd3(xscale(in_plane, y4_fast_sin(phi)), xscale(cam.vto, y4_fast_cos(phi)));
	const float tmin = xdiv(xdot(cam.near_n, xsub3(cam.near_p, cam.pos)), xdot(d, cam.near_n));
	const float tmax = xdiv(xdot(cam.far_n, xsub3(cam.far_p, cam.pos)), xdot(d, cam.far_n));
	return y4_mk_ray(cam.pos, tmin, d, tmax);
}

// PerspectiveCamera::shootRay (camera_perspective.cc:128-146); lu, lv: lens sample (used when aperture != 0)
Y4_HD Y4Ray y4_camera_ray(const Y4PtCamera &cam, float px, float py, float lu, float lv)
{
	if(cam.type == Y4GPU_CAM_ANGULAR) return y4_angular_camera_ray(cam, px, py);
	if(cam.type == Y4GPU_CAM_ORTHOGRAPHIC || cam.type == Y4GPU_CAM_EQUIRECTANGULAR)
	{
		// OrthographicCamera::shootRay (camera_orthographic.cc:52-60) / EquirectangularCamera::shootRay (camera_equirectangular.cc:50-62)
		V3 from = cam.pos, d = cam.vto;
		if(cam.type == Y4GPU_CAM_ORTHOGRAPHIC) from = xadd3(xadd3(cam.pos, xscale(cam.vright, px)), xscale(cam.vup, py));
		else
		{
			const float u = xsub(xdiv(xmul(2.f, px), (float) cam.resx), 1.f), v = xsub(xdiv(xmul(2.f, py), (float) cam.resy), 1.f);
			const float phi = (float) xdmul(Y4_PI, (Y4_WIDE) u), theta = (float) xdmul(Y4_PI_2, (Y4_WIDE) v);   // long double constants
			const V3 h = xadd3(xscale(cam.vto, y4_fast_cos(phi)), xscale(cam.vright, y4_fast_sin(phi)));
			d = xadd3(xscale(h, y4_fast_cos(theta)), xscale(cam.vup, y4_fast_sin(theta)));
		}
		const float tmin = xdiv(xdot(cam.near_n, xsub3(cam.near_p, from)), xdot(d, cam.near_n));
		const float tmax = xdiv(xdot(cam.far_n, xsub3(cam.far_p, from)), xdot(d, cam.far_n));
		return y4_mk_ray(from, tmin, d, tmax);
	}
	V3 d = xadd3(xadd3(xscale(cam.vright, px), xscale(cam.vup, py)), cam.vto);
	d = xnormalize(d);
	// Plane::rayIntersection: n * (p - from) / (dir * n)
	const float tmin = xdiv(xdot(cam.near_n, xsub3(cam.near_p, cam.pos)), xdot(d, cam.near_n));
	const float tmax = xdiv(xdot(cam.far_n, xsub3(cam.far_p, cam.pos)), xdot(d, cam.far_n));
	V3 from = cam.pos;
	if(cam.aperture != 0.f)
	{
		float u, v;
		y4_lens_uv(cam, lu, lv, u, v);
		const V3 li = xadd3(xscale(cam.dof_rt, u), xscale(cam.dof_up, v));
		from = xadd3(from, li);
		d = xnormalize(xsub3(xscale(d, cam.dof_distance), li));
	}
	return y4_mk_ray(from, tmin, d, tmax);
}

// Elemental steps take the path's flag word in a register (`f`); the fused stage functions further
// down load and store it once.  Shadow rays exist only for the slots whose Y4_PF_S0/S1 bit is set;
// an untraced slot counts as unoccluded (no shadow casting, or an invalid sample with zero contribution).
// State a stage always needs is loaded up front (independent 128-bit loads in flight together),
// not behind the branches that consume it.

// ---- step: camera ray ---------------------------------------------------------------------------
Y4_HD Y4Ray y4_primary_ray(const Y4PtScene &sc, const Y4PtPass &ps, uint32_t pix, uint32_t samp)
{
	const Y4PixelSampling s = y4_pixel_sampling(sc, ps, pix, samp);
	float lens_u = 0.5f, lens_v = 0.5f;
	if(sc.cam.aperture != 0.f)
	{
		// renderTile: hal_u(3) / hal_v(5) are started per pixel at pass_offs + offset and advanced once per sample of the
		// pass (integrator_tiled.cc:285-286,316-317,336-340); sample `samp` of the pass takes the (samp + 1)-th value
		Y4Halton hu = y4_halton_start(3u, ps.pass_offs + s.offset), hv = y4_halton_start(5u, ps.pass_offs + s.offset);
		for(uint32_t k = 0; k <= samp; ++k) { lens_u = y4_halton_next(hu); lens_v = y4_halton_next(hv); }
	}
	return y4_camera_ray(sc.cam, xadd((float) s.x, s.dx), xadd((float) s.y, s.dy), lens_u, lens_v);
}
Y4_HD void y4_pt_gen(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i)
{
	st_ray(W.ray, i, y4_primary_ray(sc, ps, W.pix[i], W.samp[i]));
}

struct Y4Vertex { Y4Surf sp; Y4PtMaterial m; V3 wo; };
// SPEC = false: the scene has no material with a specular component (Y4PtScene::has_specular == 0).  The kernels are
// compiled a second time for that case with the mirror fields of the material copy forced to constants, so that the
// reflection code folds away and costs plain diffuse scenes no registers (the emulation and specular scenes use SPEC = true).
template<bool SPEC = true>
Y4_HD Y4Vertex y4_vertex(const Y4PtScene &sc, const Y4Ray &r, const Y4Hit &h)
{
	Y4Vertex v;
	const V3 o = v3(r.ox, r.oy, r.oz), d = v3(r.dx, r.dy, r.dz);
	v.sp = y4_surface(sc, o, d, h.t, h.bu, h.bv, h.prim);
	v.m = sc.mats[v.sp.mat];
	if(!SPEC) { v.m.is_mirror = 0; v.m.fresnel = 0; v.m.is_transparent = 0; v.m.is_translucent = 0; if(v.m.type == Y4GPU_MAT_MIRROR) v.m.type = Y4GPU_MAT_SHINYDIFFUSE; }
	if(v.m.layer_count > 0)          // initBsdf evaluates the node chain; getShaderColor / emit() read it (material_shiny_diffuse.cc:237,244)
	{
		v.m.color = y4_diffuse_shader_color(sc, v.m, v.sp, h.bu, h.bv, h.prim);
		v.m.emit = cscale(v.m.color, v.m.emit_strength);
	}
	v.wo = xneg(d);
	return v;
}

// ---- step: camera-ray hit (integrate(), :126-171 up to the path loop) -----------------------------
// `level` = ray_level of this integrate() call: 0 for the camera ray (W.ray), > 0 for a ray reflected by mirrors (W.ray0;
// f comes in with Y4_PF_SPEC).  Returns 0: background, 1: surface hit, `vx` filled, 2: the hit is a perfect mirror and the
// reflected ray was written to W.ray0 (recursiveRaytrace -> specularReflect -> integrate at level + 1,
// integrator_montecarlo.cc:868-964; the chain's colour factors are multiplied up front into W.sw and applied when the
// pixel-sample is splatted: same factors as the reference's unwinding products, in the opposite order).
// (*background)(ray.dir_): ConstantBackground::eval / GradientBackground::eval (background_gradient.cc:34-46)
Y4_HD C3 y4_background(const Y4PtScene &sc, const V3 dir)
{
	if(sc.bg_gradient == 4) return y4_darksky_background(*sc.bg_dark, dir, false);      // DarkSkyBackground::operator() (background_darksky.cc:162-187): no power_
	if(sc.bg_gradient == 3) return y4_sunsky_background(*sc.bg_sun, dir);       // SunSkyBackground::eval (background_sunsky.cc:115-175)
	if(sc.bg_gradient == 2) return y4_texture_background(sc.bg_tex, dir);       // TextureBackground::eval (background_texture.cc:44-78)
	if(!sc.bg_gradient) return sc.background;
	float blend = dir.z;
	C3 color;
	if(blend >= 0.f) color = cadd(cscale(sc.bg_sz, blend), cscale(sc.bg_sh, xsub(1.f, blend)));
	else
	{
		blend = -blend;
		color = cadd(cscale(sc.bg_gz, blend), cscale(sc.bg_gh, xsub(1.f, blend)));
	}
	if(fminf(color.r, fminf(color.g, color.b)) < 1e-6f) color = c3f(1e-5f);
	return color;
}

template<bool SPEC = true>
Y4_HD int y4_pt_primary(const Y4PtScene &sc, const Y4PtWave &W, uint32_t i, int level, uint32_t &f, Y4Vertex &vx)
{
	const Y4Hit h = ld_hit(W.hit, i);
	const Y4Ray r = level > 0 ? ld_ray(W.ray0, i) : ld_ray(W.ray, i);
	const uint32_t spec = level > 0 ? (f & (uint32_t) (Y4_PF_SPEC | Y4_PF_ACC)) : 0u;      // f in: the flags the level above left
	if(h.prim < 0)
	{
		const bool invalid = level == 0 && y4_camera_ray_invalid(r);         // no camera ray for this position: default colour
		const bool transp = !invalid && sc.transp_background != 0 && (level == 0 || sc.transp_refracted_background != 0);   // TiledIntegrator::background
		const C3 bg = (transp || invalid) ? c3f(0.f) : y4_background(sc, v3(r.dx, r.dy, r.dz));
		st4(W.col, i, f4(bg.r, bg.g, bg.b, transp ? 0.f : 1.f));
		f = spec;
		return 0;
	}
	st_hit(W.hit0, i, h);
	vx = y4_vertex<SPEC>(sc, r, h);
	const bool deeper = SPEC && level + 1 <= sc.raydepth && level + 1 < 20;
	if(SPEC && vx.m.type == Y4GPU_MAT_MIRROR && deeper)
	{
		// MirrorMaterial::getSpecular (material_glass.cc:443-451) + specularReflect
		const V3 dir = y4_reflect_dir(y4_face_forward(vx.sp.ng, vx.sp.n, vx.wo), vx.wo);
		C3 w = vx.m.color;
		if(level > 0) w = cmul(ld3(W.sw, i), w);
		st_ray(W.ray0, i, y4_mk_ray(vx.sp.p, sc.ray_min_dist, dir, -1.f));
		st4(W.sw, i, f4(w.r, w.g, w.b, y4_u2f((uint32_t) (level + 1))));
		f = Y4_PF_SPEC | (spec & Y4_PF_ACC);
		return 2;
	}
	C3 col = c3f(0.f);
	if(vx.m.bsdf_flags & Y4_BSDF_EMIT) col = cadd(col, y4_mat_emit(vx.m, vx.sp, vx.wo));
	st4(W.col, i, f4(col.r, col.g, col.b, 1.f));               // recursiveRaytrace: alpha_count == 0 -> alpha 1
	st3(W.pcol, i, c3f(0.f));
	st4(W.lsamp, i, f4(0.f, 0.f, y4_u2f(0xffffffffu), y4_u2f(0xffffffffu)));
	f = Y4_PF_HIT0 | spec;
	if((vx.m.is_mirror || vx.m.is_transparent) && deeper) f |= Y4_PF_RECURSE;
	if(vx.m.bsdf_flags & Y4_BSDF_DIFFUSE) f |= Y4_PF_DIRECT;
	const uint32_t path_flags = sc.no_recursive ? (uint32_t) Y4_BSDF_ALL : (uint32_t) Y4_BSDF_DIFFUSE;
	if(sc.n_paths > 0 && (vx.m.bsdf_flags & path_flags)) f |= Y4_PF_PATHS;       // n_paths == 0: DirectLightIntegrator (integrator_direct_light.cc:94-143)
	return 1;
}

// the ray integrate() was called with for this pixel-sample: the camera ray (recomputed, cheaper than a 32-byte read) or the
// stored reflected ray
template<bool SPEC = true>
Y4_HD Y4Ray y4_level_ray(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, uint32_t f, uint32_t pix, uint32_t samp)
{
	return (SPEC && (f & Y4_PF_SPEC)) ? ld_ray(W.ray0, i) : y4_primary_ray(sc, ps, pix, samp);
}

// number of samples doLightEstimation takes from light L (integrator_montecarlo.cc:396)
Y4_HD int y4_light_num_samples(const Y4PtScene &sc, const Y4PtLight &L)
{
	if(y4_light_dirac(L.type)) return 1;
	return (int) ceilf(xmul((float) L.samples, sc.aa_light_sample_multiplier));
}

// does this path take sample k of a light now?  (shared by the sample and the resolve step)
Y4_HD bool y4_direct_active(const Y4PtScene &sc, const Y4PtWave &W, uint32_t i, uint32_t f, bool all_lights, int light, int k, int &l, int &num_samples)
{
	if(!(f & Y4_PF_DIRECT)) return false;          // set by the camera-hit / path-vertex step, cleared when the estimate is folded in
	l = all_lights ? light : (sc.n_lights > 1 ? (int) W.lnum[i] : 0);
	// l == n_lights: the ambient-occlusion samples (camera-hit estimates only: the test folds away in the path-vertex kernels)
	num_samples = (all_lights && l == sc.n_lights) ? sc.ao_samples : y4_light_num_samples(sc, sc.lights[l]);
	return k < num_samples;
}

// The pair of numbers areaLightSampleLight / areaLightSampleMaterial use for sample k of light l
// (integrator_montecarlo.cc:398-404: Halton(2) and Halton(3) started at num_samples * sample + offset + l * 4567 - 1,
// advanced k + 1 times).  They depend on the pixel-sample, the light and k only -- not on the vertex -- so the last
// pair is kept in W.lsamp and reused down the path.
Y4_HD void y4_light_sample_numbers(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, int l, int k, int num_samples, float4 cache, float &s1, float &s2,
                                   bool ao = false)
{
	if(y4_f2u(cache.z) == (uint32_t) l && y4_f2u(cache.w) == (uint32_t) k) { s1 = cache.x; s2 = cache.y; return; }
	const Y4PixelId s = y4_pixel_id(sc, ps, W.pix[i], W.samp[i]);
	// sampleAmbientOcclusion starts its two sequences at n * sample + offset - 1 (integrator_tiled.cc:651-653): no per-light term
	const uint32_t offs = (uint32_t) num_samples * s.sample + s.offset + (ao ? 0u : (uint32_t) l * 4567u);
	s1 = y4_halton2_at(offs - 1u, (uint32_t) k + 1u);
	Y4Halton h3 = y4_halton_start(3u, offs - 1u);
	for(int q = 0; q <= k; ++q) s2 = y4_halton_next(h3);
	st4(W.lsamp, i, f4(s1, s2, y4_u2f((uint32_t) l), y4_u2f((uint32_t) k)));
}

// ---- step: k-th sample of one light at vertex `vx` -------------------------------------------------
// all_lights: vertex = camera hit (estimateAllDirectLight, light index given);
// otherwise : vertex = current path vertex, light = W.lnum[i] (estimateOneDirectLight).
// Returns the mask of shadow-ray slots written (bit q -> W.sray[q * n + i]).
template<bool SPEC = true>
Y4_HD uint32_t y4_pt_direct(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, uint32_t &f, bool all_lights, int light, int k, const Y4Vertex &vx, float4 cache)
{
	const uint32_t n = W.n;
	f &= ~(uint32_t) (Y4_PF_S0 | Y4_PF_S1 | Y4_PF_C1);
	int l, num_samples;
	if(!y4_direct_active(sc, W, i, f, all_lights, light, k, l, num_samples)) return 0u;
	const bool ao = all_lights && l == sc.n_lights;
	const Y4PtLight &L = sc.lights[ao ? 0 : l];
	float s1 = 0.f, s2 = 0.f;
	if(ao || !y4_light_dirac(L.type)) y4_light_sample_numbers(sc, W, ps, i, l, k, num_samples, cache, s1, s2, ao);
	Y4DirectOut out;
	if(ao) y4_ao_sample(sc, vx.m, vx.sp, vx.wo, s1, s2, out);
	else y4_direct_sample<SPEC>(sc, L, vx.m, vx.sp, vx.wo, s1, s2, out);
	const bool cast = ao || (L.cast_shadows && vx.m.receive_shadows);      // occlusion rays are traced whatever receive_shadows says
	uint32_t mask = 0u;
	for(int q = 0; q < 2; ++q)
	{
		// zero when the sample is invalid; slot 1 (a BSDF-sampled direction that happens to reach the light) mostly is, and then
		// neither it nor its read-back in the resolve step costs memory traffic
		if(q == 0) st3(W.scontrib, i, out.contrib[0]);
		else if(!cblack(out.contrib[1])) { st3(W.scontrib, n + i, out.contrib[1]); f |= Y4_PF_C1; }
		if(!out.valid[q] || !cast || cblack(out.contrib[q])) continue;      // nothing to occlude: no shadow ray
		st_ray(W.sray, q * n + i, out.ray[q]);
		mask |= 1u << q;
	}
	if(mask & 1u) f |= Y4_PF_S0;
	if(mask & 2u) f |= Y4_PF_S1;
	return mask;
}

// ---- step: fold the shadow-ray results of sample k into the light's estimators ---------------------
// (one-light estimates multiply by the throughput stored at the vertex they belong to: W.pthr)
Y4_HD void y4_pt_direct_resolve(const Y4PtScene &sc, const Y4PtWave &W, uint32_t i, uint32_t &f, bool all_lights, int light, int k)
{
	const uint32_t n = W.n;
	int l, num_samples;
	if(!y4_direct_active(sc, W, i, f, all_lights, light, k, l, num_samples)) return;
	// everything this step can need, requested together
	const float4 zero4 = f4(0.f, 0.f, 0.f, 0.f);
	float4 c0 = ld4(W.scontrib, i), c1 = (f & Y4_PF_C1) ? ld4(W.scontrib, n + i) : zero4;
	int32_t p0 = -1, p1 = -1;
	if(sc.tshadow_filter)
	{
		// isShadowed(ray, s_depth_, ...) -> scol; "lcol *= scol" (integrator_montecarlo.cc:121,229,335): the product of the transparent
		// blockers' getTransparency colours scales the sample (applied to the finished product here: <= 1 ulp from the reference's order)
		if(f & Y4_PF_S0) { const Y4Hit s = ld_hit(W.shit, i); p0 = s.prim; c0.x = xmul(c0.x, s.t); c0.y = xmul(c0.y, s.bu); c0.z = xmul(c0.z, s.bv); }
		if(f & Y4_PF_S1) { const Y4Hit s = ld_hit(W.shit, n + i); p1 = s.prim; c1.x = xmul(c1.x, s.t); c1.y = xmul(c1.y, s.bu); c1.z = xmul(c1.z, s.bv); }
	}
	else
	{
		p0 = (f & Y4_PF_S0) ? W.shit[i].prim : -1; p1 = (f & Y4_PF_S1) ? W.shit[n + i].prim : -1;
	}
	const float4 base = all_lights ? ld4(W.col, i) : ld4(W.pcol, i);
	const float4 thr4 = all_lights ? zero4 : ld4(W.pthr, i);
	const float4 lc4 = (all_lights || !(f & Y4_PF_LCOL)) ? zero4 : ld4(W.lcol, i);
	const bool ao = all_lights && l == sc.n_lights;
	const Y4PtLight &L = sc.lights[ao ? 0 : l];
	// running sums over the light's samples live in acc_a/acc_b only between samples (k > 0)
	C3 a = c3f(0.f), b = c3f(0.f);
	if(k > 0) { a = ld3(W.acc_a, i); b = ld3(W.acc_b, i); }
	if(p0 < 0) a = cadd(a, c3(c0.x, c0.y, c0.z));
	if(p1 < 0) b = cadd(b, c3(c1.x, c1.y, c1.z));
	f &= ~(uint32_t) (Y4_PF_S0 | Y4_PF_S1 | Y4_PF_C1);
	if(k < num_samples - 1) { st3(W.acc_a, i, a); st3(W.acc_b, i, b); return; }
	// last sample of this light: doLightEstimation's result
	C3 dl = c3f(0.f);
	if(ao)
	{
		// sampleAmbientOcclusion returns col / n (a division, integrator_tiled.cc:691); b holds the emission terms of an emitting
		// diffuse material, which the reference adds between the occlusion terms (<= 1 ulp apart in that case, exact otherwise)
		const float nf = (float) num_samples;
		const C3 sum = cadd(a, b);                         // b == 0 unless the material emits: a + 0 is a
		dl = c3(xdiv(sum.r, nf), xdiv(sum.g, nf), xdiv(sum.b, nf));
	}
	else if(y4_light_dirac(L.type)) dl = cadd(dl, a);
	else
	{
		const float inv = xdiv(1.f, (float) num_samples);
		dl = cadd(dl, cscale(a, inv));
		dl = cadd(dl, cscale(b, inv));
	}
	if(all_lights)                                                             // estimateAllDirectLight: col += ...
	{
		const C3 c = cadd(c3(base.x, base.y, base.z), dl);
		st4(W.col, i, f4(c.r, c.g, c.b, base.w));
	}
	else
	{
		// estimateOneDirectLight: ... * num_lights; then "lcol += col_emit" (first vertex) and path_col += lcol * throughput
		C3 lcol = cscale(dl, (float) sc.n_lights);
		lcol = cadd(lcol, c3(lc4.x, lc4.y, lc4.z));
		st3(W.pcol, i, cadd(c3(base.x, base.y, base.z), cmul(lcol, c3(thr4.x, thr4.y, thr4.z))));
		f &= ~(uint32_t) Y4_PF_DIRECT;
	}
}

// ---- step: BSDF sample that extends sub-path `sub` from vertex `vx` (integrate(), :163-231) ---------
// FIRST: the vertex is the camera hit (bounce 0), else the vertex reached by bounce `depth - 1`.
// thr: throughput at the vertex in, throughput along the new ray out.  Returns true when `out` is the next ray.
template<bool FIRST>
Y4_HD bool y4_bounce_sample(const Y4PtScene &sc, const Y4PtPass &ps, uint32_t pix, uint32_t samp, int depth, int sub, const Y4Vertex &vx, C3 &thr, Y4Ray &out, bool &specular)
{
	const Y4PixelId s = y4_pixel_id(sc, ps, pix, samp);
	const uint32_t offs = (uint32_t) sc.n_paths * s.sample + s.offset + (uint32_t) sub;
	float s1, s2; uint32_t sflags;
	if(FIRST)
	{
		s1 = y4_ri_vdc(offs, 0u);
		s2 = (float) y4_lds(sc.lds, 2, offs);
		sflags = (sc.no_recursive ? (uint32_t) Y4_BSDF_ALL : (uint32_t) Y4_BSDF_DIFFUSE) | Y4_BSDF_DIFFUSE | Y4_BSDF_REFLECT | Y4_BSDF_TRANSMIT;
	}
	else
	{
		s1 = (float) y4_lds(sc.lds, 4 * depth + 3, offs);
		s2 = (float) y4_lds(sc.lds, 4 * depth + 4, offs);
		sflags = Y4_BSDF_ALL;
	}
	V3 wi = v3(0.f, 0.f, 1.f);
	float w = 0.f, pdf; bool sampled;
	C3 scol = y4_mat_sample(vx.m, vx.sp, vx.wo, s1, s2, sflags, wi, w, pdf, sampled, specular);
	scol = cscale(scol, w);
	if(FIRST) thr = scol;
	else
	{
		if(cblack(scol)) return false;
		thr = cmul(thr, scol);
	}
	if(!sampled) return false;                                 // reference: direction undefined
	out = y4_mk_ray(vx.sp.p, sc.ray_min_dist, wi, -1.f);
	return true;
}

// ---- step: the extended ray's hit (integrate(), :191-207 and :232-266) ------------------------------
// Pure logic on loaded state.  Returns 0: the path ended (background or Russian roulette); 1: the vertex needs no
// light estimate (its emission, if any, goes into `add` = lcol * throughput); 2: one-light estimate wanted
// (`emit` = lcol so far, light index chosen into W.lnum when there is more than one).
Y4_HD int y4_bounce_hit(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, uint32_t pix, uint32_t samp, uint32_t &f, int depth, int sub,
                        const Y4Vertex &vx, C3 &thr, C3 &emit)
{
	const Y4PtMaterial &m = vx.m;
	emit = c3f(0.f);
	bool direct;
#ifdef Y4_TEST_LIGHT_CHOICE_HOOK
	// tests/emu only: every estimateOneDirectLight call of the reference advances its counter, also the ones this
	// function skips because they evaluate to exactly zero (non-diffuse vertex) or that Russian roulette discards later
	// (the reference finishes all sub-paths of a ray level before it recurses: the level goes in front of the sub-path index)
	const int lc_sub = sub + 4096 * ((f & Y4_PF_SPEC) ? (int) (y4_f2u(ld4(W.sw, i).w) & 0xffu) : 0);
	if(sc.n_lights > 1 && (depth == 0 || (f & Y4_PF_DIFF1))) y4_test_light_choice_called(pix, samp, lc_sub, depth);
#endif
	if(depth == 0)
	{
		if(m.bsdf_flags & Y4_BSDF_DIFFUSE) f |= Y4_PF_DIFF1; else f &= ~(uint32_t) Y4_PF_DIFF1;
		if(m.bsdf_flags & Y4_BSDF_EMIT) { emit = y4_mat_emit(m, vx.sp, vx.wo); f |= Y4_PF_EMIT1; } else f &= ~(uint32_t) Y4_PF_EMIT1;
		direct = (m.bsdf_flags & Y4_BSDF_DIFFUSE) != 0;     // a vertex without the Diffuse flag evaluates to exactly 0 for every light sample
		// neither flag: no later vertex of this sub-path can add anything (light estimates need DIFF1, emission EMIT1)
		if(!(f & (Y4_PF_DIFF1 | Y4_PF_EMIT1))) return 0;
	}
	else
	{
		direct = (f & Y4_PF_DIFF1) && (m.bsdf_flags & Y4_BSDF_DIFFUSE);
		if(depth > sc.rr_min_bounces)
		{
			// the reference draws from a rand()-seeded per-tile generator (not reproducible); documented substitute
			const Y4PixelId s = y4_pixel_id(sc, ps, pix, samp);
			const float random_value = y4_hash_uniform(sc.rr_seed ^ s.offset, s.sample, (uint32_t) (sub * 64 + depth));
			const float probability = cmax(thr);
			if(probability <= 0.f || probability < random_value) return 0;
			thr = cscale(thr, xdiv(1.f, probability));
		}
		// emission seen through a specular bounce; the flag tested is the FIRST path vertex's (mat_bsd_fs is not updated in the loop)
		if((f & Y4_PF_EMIT1) && (f & Y4_PF_CAUSTIC)) emit = y4_mat_emit(m, vx.sp, vx.wo);
	}
	if(sc.n_lights == 0) direct = false;
	if(!direct) return 1;
	// light choice of estimateOneDirectLight: Halton(2) at base offset + a per-call counter.  The
	// reference counter is per CPU thread (call-order dependent); here it is a per-vertex index.
	// With one light every draw picks light 0 (min(int(h * 1), 0)): nothing to compute or store.
	if(sc.n_lights > 1)
	{
		uint32_t corr = (pix * 1009u + ps.pass_offs + samp) * (uint32_t) (sc.bounces * sc.n_paths) + (uint32_t) (sub * sc.bounces + depth);
#ifdef Y4_TEST_LIGHT_CHOICE_HOOK
		corr = y4_test_light_choice(pix, samp, lc_sub, depth, corr);   // tests/emu only: replays the reference's single-thread call counter
#endif
		const int lnum = (int) xmul(y4_halton2_at(sc.base_sampling_offset + corr - 1u, 1u), (float) sc.n_lights);
		W.lnum[i] = (uint32_t) (lnum < sc.n_lights - 1 ? lnum : sc.n_lights - 1);
	}
	return 2;
}

// ==== fused wavefront stages ===========================================================================
// Each returns a push mask: bit 0/1 = shadow-ray slot 0/1 written, bit 2 = the path goes on the stage's
// output path queue.  y4_pt.cu runs them as kernels over compacted index queues; tests/emu as loops.
enum : uint32_t { Y4_PUSH_S0 = 1u, Y4_PUSH_S1 = 2u, Y4_PUSH_PATH = 4u, Y4_PUSH_SPEC = 8u };

// camera-ray hit + first direct-light sample (light l0, k = 0; l0 < 0: scene without lights).
// Output queue: paths that hit a surface and have more to do (more light samples or sub-paths).
// level > 0: the same for a ray that mirrors reflected (bit 3 of the result: reflected again -> the ray-level queue).
template<bool SPEC = true>
Y4_HD uint32_t y4_stage_primary(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, int l0, int level)
{
	uint32_t f = level > 0 ? W.flags[i] : 0u; Y4Vertex vx;
	uint32_t push = 0u;
	const int what = y4_pt_primary<SPEC>(sc, W, i, level, f, vx);
	if(what == 1)
	{
		if(l0 >= 0) push |= y4_pt_direct<SPEC>(sc, W, ps, i, f, true, l0, 0, vx, f4(0.f, 0.f, y4_u2f(0xffffffffu), y4_u2f(0xffffffffu)));
		if(f & (Y4_PF_DIRECT | Y4_PF_PATHS | Y4_PF_RECURSE)) push |= Y4_PUSH_PATH;
	}
	else if(what == 2) push |= Y4_PUSH_SPEC;
	W.flags[i] = f;
	return push;
}

// resolve sample (l_prev, k_prev), take sample (l, k); camera-hit vertex (all_lights) or current path vertex
template<bool SPEC = true>
Y4_HD uint32_t y4_stage_direct_step(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, bool all_lights, int l_prev, int k_prev, int l, int k)
{
	uint32_t f = W.flags[i];
	const float4 cache = ld4(W.lsamp, i);
	y4_pt_direct_resolve(sc, W, i, f, all_lights, l_prev, k_prev);
	uint32_t push = 0u;
	int ll, ns;
	if(y4_direct_active(sc, W, i, f, all_lights, l, k, ll, ns))
	{
		const Y4Ray r = all_lights ? y4_level_ray<SPEC>(sc, W, ps, i, f, W.pix[i], W.samp[i]) : ld_ray(W.ray, i);
		const Y4Vertex vx = y4_vertex<SPEC>(sc, r, all_lights ? ld_hit(W.hit0, i) : ld_hit(W.hit, i));
		push = y4_pt_direct<SPEC>(sc, W, ps, i, f, all_lights, l, k, vx, cache);
	}
	W.flags[i] = f;
	return push;
}

// start sub-path `sub` at the camera hit: [resolve the last all-lights sample] + BSDF sample of bounce 0 -> W.ray_next
template<bool SPEC = true>
Y4_HD uint32_t y4_stage_start(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, int l_last, int k_last, int sub)
{
	uint32_t f = W.flags[i];
	const uint32_t pix = W.pix[i], samp = W.samp[i];
	const Y4Hit h = ld_hit(W.hit0, i);
	if(l_last >= 0) y4_pt_direct_resolve(sc, W, i, f, true, l_last, k_last);
	f &= ~(uint32_t) (Y4_PF_ALIVE | Y4_PF_DIFF1 | Y4_PF_DIRECT | Y4_PF_EMIT1 | Y4_PF_CAUSTIC);
	bool go = false;
	if(f & Y4_PF_PATHS)
	{
		const Y4Vertex vx = y4_vertex<SPEC>(sc, y4_level_ray<SPEC>(sc, W, ps, i, f, pix, samp), h);
		C3 thr = c3f(0.f); Y4Ray out;
		bool specular;
		go = y4_bounce_sample<true>(sc, ps, pix, samp, 0, sub, vx, thr, out, specular);
		st3(W.thr, i, thr);
		if(go) { st_ray(W.ray_next, i, out); f |= Y4_PF_ALIVE; }
	}
	W.flags[i] = f;
	return go ? (uint32_t) Y4_PUSH_PATH : 0u;
}

// One bounce of the path loop in a single stage over the paths whose ray W.ray was just traced:
//   1. fold in the one-light estimate of the previous vertex (its shadow rays were traced meanwhile);
//   2. this vertex: emission, Russian roulette, light choice, first light sample -> shadow rays;
//   3. BSDF sample -> the next ray in W.ray_next.
// A path stays on the output queue while it has a ray to trace OR an estimate waiting to be folded in (such a
// finished path gets an empty ray).  depth == sc.bounces is the closing call: step 1 only.
// The stage comes in two parts so that a kernel may regroup the vertices between them (k_pt_path_compact: the paths whose ray
// missed or that only had an estimate to fold leave after part 1; the others are packed into full warps for part 2).
// Part 1 = step 1 and the ways out; true when the path has a vertex to process.
Y4_HD bool y4_stage_path_fold(const Y4PtScene &sc, const Y4PtWave &W, uint32_t i, int depth, int k_last)
{
	uint32_t f = W.flags[i];
	const bool closing = depth >= sc.bounces;
	if(depth > 0 && k_last >= 0) y4_pt_direct_resolve(sc, W, i, f, false, -1, k_last);
	if(closing || !(f & Y4_PF_ALIVE)) { W.flags[i] = f & ~(uint32_t) Y4_PF_ALIVE; return false; }
	if(ld_hit(W.hit, i).prim < 0) { W.flags[i] = f & ~(uint32_t) (Y4_PF_ALIVE | Y4_PF_DIRECT); return false; }     // hit background: continue / break
	W.flags[i] = f;
	return true;
}
// Part 2 = steps 2 and 3.
template<bool SPEC = true>
Y4_HD uint32_t y4_stage_path_vertex(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, int depth, int sub)
{
	uint32_t f = W.flags[i];
	const float4 cache = ld4(W.lsamp, i), thr4 = ld4(W.thr, i);
	const Y4Hit h = ld_hit(W.hit, i);
	const Y4Ray r = ld_ray(W.ray, i);
	const uint32_t pix = W.pix[i], samp = W.samp[i];
	uint32_t push = 0u;
	const Y4Vertex vx = y4_vertex<SPEC>(sc, r, h);
	C3 thr = c3(thr4.x, thr4.y, thr4.z), emit;
	const int what = y4_bounce_hit(sc, W, ps, i, pix, samp, f, depth, sub, vx, thr, emit);
	if(what == 0) { W.flags[i] = f & ~(uint32_t) (Y4_PF_ALIVE | Y4_PF_DIRECT); return 0u; }
	if(what == 2)
	{
		if(!cblack(emit)) { st3(W.lcol, i, emit); f |= Y4_PF_LCOL; } else f &= ~(uint32_t) Y4_PF_LCOL;
		st3(W.pthr, i, thr);
		f |= Y4_PF_DIRECT;
		push |= y4_pt_direct<SPEC>(sc, W, ps, i, f, false, -1, 0, vx, cache);
	}
	else
	{
		// lcol = 0 (* num_lights) + emit;  path_col += lcol * throughput
		const C3 lcol = cadd(c3f(0.f), emit);
		st3(W.pcol, i, cadd(ld3(W.pcol, i), cmul(lcol, thr)));
		f &= ~(uint32_t) Y4_PF_DIRECT;
	}
	bool go = false;
	if(depth + 1 < sc.bounces)
	{
		Y4Ray out;
		bool specular;
		go = y4_bounce_sample<false>(sc, ps, pix, samp, depth + 1, sub, vx, thr, out, specular);
		// caustic = trace_caustics_ && sampled_flags has Specular | Glossy | Filter (:227)
		if(go && sc.trace_caustics && specular) f |= Y4_PF_CAUSTIC; else f &= ~(uint32_t) Y4_PF_CAUSTIC;
		if(go) { st_ray(W.ray_next, i, out); st3(W.thr, i, thr); }
		else if(f & Y4_PF_DIRECT) st_ray(W.ray_next, i, y4_null_ray());
	}
	if(!go) f &= ~(uint32_t) Y4_PF_ALIVE;
	if(go || (f & Y4_PF_DIRECT)) push |= Y4_PUSH_PATH;
	W.flags[i] = f;
	return push;
}
template<bool SPEC = true>
Y4_HD uint32_t y4_stage_path(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, int depth, int sub, int k_last)
{
	return y4_stage_path_fold(sc, W, i, depth, k_last) ? y4_stage_path_vertex<SPEC>(sc, W, ps, i, depth, sub) : 0u;
}

// radiance integrate() returns at the current ray level, before recursiveRaytrace's share: col + path_col / n_samples
Y4_HD C3 y4_level_radiance(const Y4PtScene &sc, const Y4PtWave &W, uint32_t i, uint32_t f, float4 col4)
{
	C3 col = c3(col4.x, col4.y, col4.z);
	if(f & Y4_PF_PATHS) col = cadd(col, cdiv(ld3(W.pcol, i), (float) (sc.n_paths > 1 ? sc.n_paths : 1)));
	return col;
}

// ---- stage: end of a ray level for vertices with a mirror component (shinydiffusemat specular_reflect) -------------
// The vertex's own estimate (emission + lights + sub-paths) is complete: fold it into W.acc with the weight of the chain so
// far, then reflect the level's ray (ShinyDiffuseMaterial::getSpecular, material_shiny_diffuse.cc:400-440; specularReflect,
// integrator_montecarlo.cc:868-887).  Returns Y4_PUSH_SPEC when the pixel-sample goes on to the next level.
Y4_HD uint32_t y4_stage_level_finish(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, int level)
{
	uint32_t f = W.flags[i];
	if(!(f & Y4_PF_RECURSE)) return 0u;
	const C3 col = y4_level_radiance(sc, W, i, f, ld4(W.col, i));
	const float4 sw4 = (f & Y4_PF_SPEC) ? ld4(W.sw, i) : f4(1.f, 1.f, 1.f, 0.f);
	const C3 sw = c3(sw4.x, sw4.y, sw4.z);
	C3 acc = (f & Y4_PF_SPEC) ? cmul(col, sw) : col;
	if(f & Y4_PF_ACC) acc = cadd(ld3(W.acc, i), acc);
	st3(W.acc, i, acc);
	const Y4Vertex vx = y4_vertex(sc, y4_level_ray(sc, W, ps, i, f, W.pix[i], W.samp[i]), ld_hit(W.hit0, i));
	const bool backface = xdot(vx.wo, vx.sp.ng) < 0.f;
	const V3 n = backface ? xneg(vx.sp.n) : vx.sp.n, ng = backface ? xneg(vx.sp.ng) : vx.sp.ng;
	const float kr = y4_fresnel_kr(vx.m, vx.wo, n);
	V3 dir; C3 w;
	if(vx.m.is_transparent)
	{
		// refract_: straight on; (1 - components[0] * kr) * components[1] * tcol (specularRefract, transparent bias factor 0)
		dir = xneg(vx.wo);
		w = cscale(y4_shiny_tcol(vx.m), xmul(xsub(1.f, xmul(vx.m.is_mirror ? vx.m.mirror_strength : 0.f, kr)), vx.m.transparency_strength));
	}
	else
	{
		const float vn = xmul(2.f, xdot(vx.wo, n));                                  // Vec3::reflect (vector.h:227-232)
		dir = v3(xsub(xmul(vn, n.x), vx.wo.x), xsub(xmul(vn, n.y), vx.wo.y), xsub(xmul(vn, n.z), vx.wo.z));
		const float cos_wi_ng = xdot(dir, ng);
		if((double) cos_wi_ng < 0.01)
		{
			const float k = (float) (0.01 - (double) cos_wi_ng);
			dir = xnormalize(v3(xadd(dir.x, xmul(k, ng.x)), xadd(dir.y, xmul(k, ng.y)), xadd(dir.z, xmul(k, ng.z))));
		}
		w = cscale(vx.m.mirror_color, xmul(vx.m.mirror_strength, kr));
	}
	if(f & Y4_PF_SPEC) w = cmul(sw, w);
	st_ray(W.ray0, i, y4_mk_ray(vx.sp.p, sc.ray_min_dist, dir, -1.f));
	st4(W.sw, i, f4(w.r, w.g, w.b, y4_u2f((uint32_t) (level + 1))));
	W.flags[i] = Y4_PF_SPEC | Y4_PF_ACC;
	return Y4_PUSH_SPEC;
}

// ==== ray tree: recursiveRaytrace with up to two children per vertex (integrator_montecarlo.cc:918-968) ====================
// A pixel-sample walks its tree depth first.  Every vertex integrate() visits contributes (product of the specular colours
// down its branch) x (its own radiance: emission + light estimates + sub-paths, or the background); the contributions are
// summed into W.acc as the walk goes (the reference unwinds the same products and sums from the leaves up: <= 1 ulp per
// level apart).  One child continues in the slot (W.ray0 / W.sw), the sibling waits on the stack.
Y4_HD void y4_tree_fold(const Y4PtWave &W, uint32_t i, uint32_t &f, C3 rad)
{
	if(f & Y4_PF_SPEC) rad = cmul(rad, ld3(W.sw, i));
	if(f & Y4_PF_ACC) rad = cadd(ld3(W.acc, i), rad);
	st3(W.acc, i, rad);
	f |= Y4_PF_ACC;
}
// the vertex at (p, level) is done and has n_children (0..2) specular continuations; returns Y4_PUSH_SPEC while the
// pixel-sample has a ray to trace, 0 when its tree is finished (then W.col / W.flags are what the splat expects)
// med[k] != 0: child k continues inside an absorbing glass of material med[k] - 1; it travels in the upper bits of the level word
// and is cashed in when the child's hit distance is known (y4_tree_medium)
Y4_HD uint32_t y4_tree_next(const Y4PtScene &sc, const Y4PtWave &W, uint32_t i, uint32_t f, int level, const V3 p, int n_children, const V3 *dir, const C3 *col,
                            const uint32_t *med = nullptr)
{
	const uint32_t n = W.n;
	const bool has_sw = (f & Y4_PF_SPEC) != 0;
	const C3 sw = has_sw ? ld3(W.sw, i) : c3f(1.f);
	const uint32_t lvl_bits = (uint32_t) (level + 1);
	const float lvl = y4_u2f(lvl_bits | ((med && n_children >= 1 ? med[0] : 0u) << 8));
	const float lvl1 = y4_u2f(lvl_bits | ((med && n_children == 2 ? med[1] : 0u) << 8));
	uint32_t k = W.stk_n[i];
	if(n_children == 2)
	{
		const C3 w1 = has_sw ? cmul(sw, col[1]) : col[1];
		st_ray(W.stk_ray, k * n + i, y4_mk_ray(p, sc.ray_min_dist, dir[1], -1.f));
		st4(W.stk_sw, k * n + i, f4(w1.r, w1.g, w1.b, lvl1));
		W.stk_n[i] = ++k;
	}
	if(n_children >= 1)
	{
		const C3 w0 = has_sw ? cmul(sw, col[0]) : col[0];
		st_ray(W.ray0, i, y4_mk_ray(p, sc.ray_min_dist, dir[0], -1.f));
		st4(W.sw, i, f4(w0.r, w0.g, w0.b, lvl));
		W.flags[i] = Y4_PF_SPEC | (f & Y4_PF_ACC);
		return Y4_PUSH_SPEC;
	}
	if(k > 0)
	{
		--k;
		W.stk_n[i] = k;
		st_ray(W.ray0, i, ld_ray(W.stk_ray, k * n + i));
		st4(W.sw, i, ld4(W.stk_sw, k * n + i));
		W.flags[i] = Y4_PF_SPEC | (f & Y4_PF_ACC);
		return Y4_PUSH_SPEC;
	}
	st4(W.col, i, f4(0.f, 0.f, 0.f, 1.f));          // everything is in W.acc (alpha 1: transparent backgrounds are refused in tree mode)
	W.flags[i] = f & Y4_PF_ACC;
	return 0u;
}
// A tree ray that runs inside an absorbing glass: now that its hit is known, its branch weight takes BeerVolumeHandler::
// transmittance (volumehandler_beer.cc:37-44: exp(-tmax * sigma_a) per channel with the FAST_MATH exp; 0 when the ray hit nothing,
// tmax < 0).  The reference multiplies the finished radiance of the branch, here the weight carries the factor down: a few ulp.
// Returns the ray level of the entry (the low byte of the level word; the medium tag is cleared).
Y4_HD int y4_tree_medium(const Y4PtScene &sc, const Y4PtWave &W, uint32_t i, const Y4Hit &h)
{
	const float4 sw = ld4(W.sw, i);
	const uint32_t word = y4_f2u(sw.w), med = word >> 8;
	if(med)
	{
		const C3 sg = sc.mats[med - 1u].sigma_a;
		C3 t = c3f(0.f);
		if(h.prim >= 0 && !(h.t > 1e30f)) t = c3(y4_fast_exp(xmul(-h.t, sg.r)), y4_fast_exp(xmul(-h.t, sg.g)), y4_fast_exp(xmul(-h.t, sg.b)));
		st4(W.sw, i, f4(xmul(sw.x, t.r), xmul(sw.y, t.g), xmul(sw.z, t.b), y4_u2f(word & 0xffu)));
	}
	return (int) (word & 0xffu);
}
// the specular continuations of a vertex: mirror, glass, shinydiffusemat (mirror and / or transparency component)
Y4_HD int y4_tree_children(const Y4Vertex &vx, int level, V3 *dir, C3 *col, uint32_t *med)
{
	const Y4PtMaterial &m = vx.m;
	int nc = 0;
	med[0] = med[1] = 0u;
	if(m.type == Y4GPU_MAT_MIRROR)
	{
		dir[0] = y4_reflect_dir(y4_face_forward(vx.sp.ng, vx.sp.n, vx.wo), vx.wo); col[0] = m.color;     // MirrorMaterial::getSpecular
		return 1;
	}
	if(m.type == Y4GPU_MAT_GLASS)
	{
		const Y4Specular s = y4_glass_specular(m, vx.sp, vx.wo, level + 1);
		if(s.reflect) { dir[nc] = s.reflect_dir; col[nc] = s.reflect_col; ++nc; }
		if(s.refract) { dir[nc] = s.refract_dir; col[nc] = s.refract_col; ++nc; }
		// specularReflect / specularRefract: getVolumeHandler(sp.ng_ * ref_ray.dir_ < 0) -- the inside handler when the ray goes in
		if(m.absorb) for(int k = 0; k < nc; ++k) if(xdot(vx.sp.ng, dir[k]) < 0.f) med[k] = vx.sp.mat + 1u;
		return nc;
	}
	if(m.type != Y4GPU_MAT_SHINYDIFFUSE) return 0;
	// ShinyDiffuseMaterial::getSpecular (material_shiny_diffuse.cc:399-439)
	const bool backface = xdot(vx.wo, vx.sp.ng) < 0.f;
	const V3 n = backface ? xneg(vx.sp.n) : vx.sp.n, ng = backface ? xneg(vx.sp.ng) : vx.sp.ng;
	const float kr = y4_fresnel_kr(m, vx.wo, n);
	if(m.is_mirror)
	{
		V3 d = y4_reflect(vx.wo, n);
		const float cos_wi_ng = xdot(d, ng);
		if((double) cos_wi_ng < 0.01)
		{
			const float k = (float) (0.01 - (double) cos_wi_ng);
			d = xnormalize(v3(xadd(d.x, xmul(k, ng.x)), xadd(d.y, xmul(k, ng.y)), xadd(d.z, xmul(k, ng.z))));
		}
		dir[nc] = d; col[nc] = cscale(m.mirror_color, xmul(m.mirror_strength, kr)); ++nc;
	}
	if(m.is_transparent)
	{
		dir[nc] = xneg(vx.wo);
		col[nc] = cscale(y4_shiny_tcol(m), xmul(xsub(1.f, xmul(m.is_mirror ? m.mirror_strength : 0.f, kr)), m.transparency_strength)); ++nc;
	}
	return nc;
}

// camera-ray / tree-ray hit in tree mode.  run == 0: dense over the wave (camera rays), else over the queue of pending tree rays.
// Push bits: Y4_PUSH_PATH -> the vertex needs an estimate (hit0 queue; y4_stage_finish3 closes it); Y4_PUSH_SPEC -> a tree ray to trace.
Y4_HD uint32_t y4_stage_primary3(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, int l0, int run)
{
	uint32_t f = run > 0 ? (W.flags[i] & (uint32_t) (Y4_PF_SPEC | Y4_PF_ACC)) : 0u;
	if(run == 0) W.stk_n[i] = 0u;
	const Y4Hit h = ld_hit(W.hit, i);
	const int level = (f & Y4_PF_SPEC) ? y4_tree_medium(sc, W, i, h) : 0;
	const Y4Ray r = (f & Y4_PF_SPEC) ? ld_ray(W.ray0, i) : ld_ray(W.ray, i);
	V3 dir[2]; C3 col[2]; uint32_t med[2];
	if(h.prim < 0)
	{
		const bool invalid = !(f & Y4_PF_SPEC) && y4_camera_ray_invalid(r);
		y4_tree_fold(W, i, f, invalid ? c3f(0.f) : y4_background(sc, v3(r.dx, r.dy, r.dz)));
		return y4_tree_next(sc, W, i, f, level, v3(0.f, 0.f, 0.f), 0, dir, col);
	}
	st_hit(W.hit0, i, h);
	const Y4Vertex vx = y4_vertex<true>(sc, r, h);
	const bool deeper = level + 1 <= sc.raydepth && level + 1 < 20;
	if(vx.m.type == Y4GPU_MAT_MIRROR || vx.m.type == Y4GPU_MAT_GLASS)
	{
		// no Diffuse, no Emit: integrate() returns recursiveRaytrace's share only
		const int nc = deeper ? y4_tree_children(vx, level, dir, col, med) : 0;
		return y4_tree_next(sc, W, i, f, level, vx.sp.p, nc, dir, col, med);
	}
	C3 c = c3f(0.f);
	if(vx.m.bsdf_flags & Y4_BSDF_EMIT) c = cadd(c, y4_mat_emit(vx.m, vx.sp, vx.wo));
	st4(W.col, i, f4(c.r, c.g, c.b, 1.f));
	st3(W.pcol, i, c3f(0.f));
	st4(W.lsamp, i, f4(0.f, 0.f, y4_u2f(0xffffffffu), y4_u2f(0xffffffffu)));
	f |= Y4_PF_HIT0;
	if((vx.m.is_mirror || vx.m.is_transparent) && deeper) f |= Y4_PF_RECURSE;
	if(vx.m.bsdf_flags & Y4_BSDF_DIFFUSE) f |= Y4_PF_DIRECT;
	if(sc.n_paths > 0 && (vx.m.bsdf_flags & Y4_BSDF_DIFFUSE)) f |= Y4_PF_PATHS;
	uint32_t push = 0u;
	if(l0 >= 0) push |= y4_pt_direct<true>(sc, W, ps, i, f, true, l0, 0, vx, f4(0.f, 0.f, y4_u2f(0xffffffffu), y4_u2f(0xffffffffu)));
	W.flags[i] = f;
	return push | Y4_PUSH_PATH;          // every such vertex goes through y4_stage_finish3 (which also handles "nothing to estimate")
}

// the estimate of a tree vertex is complete: fold it in, spawn its specular children (shinydiffusemat), go on with the tree
Y4_HD uint32_t y4_stage_finish3(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i)
{
	uint32_t f = W.flags[i];
	const C3 rad = y4_level_radiance(sc, W, i, f, ld4(W.col, i));
	const int level = (f & Y4_PF_SPEC) ? (int) (y4_f2u(ld4(W.sw, i).w) & 0xffu) : 0;
	const bool recurse = (f & Y4_PF_RECURSE) != 0;
	uint32_t g = f & (uint32_t) (Y4_PF_SPEC | Y4_PF_ACC);
	y4_tree_fold(W, i, g, rad);
	V3 dir[2]; C3 col[2]; uint32_t med[2] = {0u, 0u};
	int nc = 0;
	V3 p = v3(0.f, 0.f, 0.f);
	if(recurse)
	{
		const Y4Vertex vx = y4_vertex<true>(sc, y4_level_ray<true>(sc, W, ps, i, f, W.pix[i], W.samp[i]), ld_hit(W.hit0, i));
		nc = y4_tree_children(vx, level, dir, col, med);
		p = vx.sp.p;
	}
	return y4_tree_next(sc, W, i, g, level, p, nc, dir, col, med);
}

// ---- stage: close the pixel-sample and splat it into the film ----------------------------------------
#if defined(__CUDA_ARCH__)
Y4_HD void y4_film_add(float *p, float v) { atomicAdd(p, v); }
// the four channels of a pixel in one 128-bit reduction (REDG.E.ADD.F32x4) when the film is 16-byte aligned
Y4_HD void y4_film_add4(float *p, float r, float g, float b, float a)
{
	if((reinterpret_cast<uintptr_t>(p) & 15u) == 0u) atomicAdd(reinterpret_cast<float4 *>(p), make_float4(r, g, b, a));
	else { atomicAdd(p, r); atomicAdd(p + 1, g); atomicAdd(p + 2, b); atomicAdd(p + 3, a); }
}
#else
Y4_HD void y4_film_add(float *p, float v) { *p += v; }
Y4_HD void y4_film_add4(float *p, float r, float g, float b, float a) { p[0] += r; p[1] += g; p[2] += b; p[3] += a; }
#endif
Y4_HD int y4_round_to_int(double v) { return (int) (v + (0.5 - 1.4e-11)); }     // math::roundToInt (math.h:268-277)

Y4_HD void y4_pt_finish_splat(const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, uint32_t i, float *film_w, float *film_rgba)
{
	const uint32_t f = W.flags[i];
	const float4 col4 = ld4(W.col, i);
	C3 col = y4_level_radiance(sc, W, i, f, col4);
	float alpha = col4.w;
	if(f & Y4_PF_SPEC) col = cmul(col, ld3(W.sw, i));               // specularReflect: integ.first *= reflect_data->col_, every level
	if(f & Y4_PF_ACC) col = cadd(ld3(W.acc, i), col);               // + what the upper levels gathered themselves
	if(alpha > 1.f) alpha = 1.f;
	if(sc.clamp_samples > 0.f)      // Rgb::clampProportionalRgb (color.h:415-445)
	{
		const float max_rgb = fmaxf(col.r, fmaxf(col.g, col.b));
		const float adj = xdiv(sc.clamp_samples, max_rgb);
		if(max_rgb > sc.clamp_samples)
		{
			if(col.r >= max_rgb) { col.r = sc.clamp_samples; col.g = xmul(col.g, adj); col.b = xmul(col.b, adj); }
			else if(col.g >= max_rgb) { col.g = sc.clamp_samples; col.r = xmul(col.r, adj); col.b = xmul(col.b, adj); }
			else { col.b = sc.clamp_samples; col.r = xmul(col.r, adj); col.g = xmul(col.g, adj); }
		}
	}
	const Y4PixelSampling s = y4_pixel_sampling(sc, ps, W.pix[i], W.samp[i]);
	const int cx1 = sc.cx0 + sc.width, cy1 = sc.cy0 + sc.height;
	int dx0 = y4_round_to_int((double) s.dx - sc.filterw), dx1 = y4_round_to_int((double) s.dx + sc.filterw - 1.0);
	int dy0 = y4_round_to_int((double) s.dy - sc.filterw), dy1 = y4_round_to_int((double) s.dy + sc.filterw - 1.0);
	if(dx0 < sc.cx0 - s.x) dx0 = sc.cx0 - s.x;
	if(dx1 > cx1 - s.x - 1) dx1 = cx1 - s.x - 1;
	if(dy0 < sc.cy0 - s.y) dy0 = sc.cy0 - s.y;
	if(dy1 > cy1 - s.y - 1) dy1 = cy1 - s.y - 1;
	const double x_offs = (double) s.dx - 0.5, y_offs = (double) s.dy - 0.5;
	for(int jj = dy0; jj <= dy1; ++jj)
	{
		const int yi = (int) floor(fabs(((double) jj - y_offs) * sc.table_scale));
		for(int ii = dx0; ii <= dx1; ++ii)
		{
			const int xi = (int) floor(fabs(((double) ii - x_offs) * sc.table_scale));
			const float wt = sc.filter_table[yi * Y4_FILTER_TABLE + xi];
			const size_t p = (size_t) (s.y + jj - sc.cy0) * (size_t) sc.width + (size_t) (s.x + ii - sc.cx0);
			y4_film_add(film_w + p, wt);
			y4_film_add4(film_rgba + 4 * p, xmul(col.r, wt), xmul(col.g, wt), xmul(col.b, wt), xmul(alpha, wt));
		}
	}
}
