// UNIT-TEST EMULATION ONLY: runs the Y4_HD wavefront stages of the path tracer
// (libyafaray_b200/csrc/y4_pt_stages.cuh) as serial loops on the CPU, with the emulated BVH8
// traversal, so the stage logic can be checked against the reference renderer where no GPU exists.
// Never linked into liby4gpu.so.
#include <cstdint>
#include <map>
#include <tuple>
#include <algorithm>
#include <vector>
// estimateOneDirectLight picks its light from Halton(2) at a per-thread call counter (integrator_montecarlo.cc:74-76,
// integrator_tiled.cc:169-171).  With threads=1 and linear tiles that counter is a function of the render order, which
// the emulation can replay (mode 1 records which vertices ask, emu_light_choice_finalize numbers them in the reference's
// tile / pixel / sample / sub-path / depth order, mode 2 feeds the numbers back).  The product keeps its per-vertex index.
#define Y4_TEST_LIGHT_CHOICE_HOOK 1
typedef std::tuple<uint32_t, uint32_t, int, int> LcKey;      // pixel, sample, ray level * 4096 + sub-path, depth
static int g_lc_mode = 0;
static std::map<LcKey, uint32_t> g_lc;
static void y4_test_light_choice_called(uint32_t pix, uint32_t samp, int sub, int depth) { if(g_lc_mode == 1) g_lc[LcKey(pix, samp, sub, depth)] = 0u; }
static uint32_t y4_test_light_choice(uint32_t pix, uint32_t samp, int sub, int depth, uint32_t corr)
{
	if(g_lc_mode == 2) { auto it = g_lc.find(LcKey(pix, samp, sub, depth)); return it == g_lc.end() ? corr : it->second; }
	return corr;
}
extern "C" void emu_light_choice_mode(int mode) { g_lc_mode = mode; if(mode == 1) g_lc.clear(); }
extern "C" uint64_t emu_light_choice_finalize(int width, int tile)
{
	typedef std::tuple<int, int, int, int, uint32_t, int, int> OrdKey;    // tile row, tile col, y, x, sample, sub, depth
	std::vector<std::pair<OrdKey, LcKey>> order;
	for(auto &kv : g_lc)
	{
		const int x = (int) (std::get<0>(kv.first) % (uint32_t) width), y = (int) (std::get<0>(kv.first) / (uint32_t) width);
		order.push_back(std::make_pair(OrdKey(y / tile, x / tile, y, x, std::get<1>(kv.first), std::get<2>(kv.first), std::get<3>(kv.first)), kv.first));
	}
	std::sort(order.begin(), order.end());
	uint32_t counter = 0;
	for(auto &o : order) g_lc[o.second] = counter++;
	return counter;
}
#include "../../libyafaray_b200/csrc/y4_trace.cuh"
#include "../../libyafaray_b200/csrc/y4_pt_host.h"
#include "../../libyafaray_b200/csrc/y4_faure.h"
#include "../../libyafaray_b200/csrc/y4_film.cuh"
#include <vector>

struct EmuBvh
{
	std::vector<Y4U4> nodes;
	std::vector<Y4Tri> tris;
	uint32_t n_nodes = 0;
	int n = 0;
};
void emu_build(const float *verts, const uint32_t *flags, int n, int leaf_max, EmuBvh &out);

// queue == null: dense batch of n rays; else n index entries
static int g_shadow_depth = 16;
static std::vector<float> g_mat_transp;      // getTransparency colour per material when shadows are filtered (else empty)
static void trace(const Y4Bvh8 &bvh, int kind, const Y4Ray *rays, const uint32_t *queue, size_t n, Y4Hit *hits)
{
	for(size_t q = 0; q < n; ++q)
	{
		const size_t i = queue ? queue[q] : q;
		const Y4PreparedRay pr = y4_prepare_ray(rays[i], kind);
		uint2 stack[Y4_STACK_SIZE];
		hits[i].t = 0.f; hits[i].bu = 0.f; hits[i].bv = 0.f; hits[i].prim = -1;
		if(!(pr.t_hi > pr.t_lo)) continue;
		uint32_t filt[16], filt_mat[16]; int n_filt = 0;
		if(kind == Y4_TRACE_CLOSEST) y4_trace_ray<Y4_TRACE_CLOSEST>(bvh, pr, hits[i], stack, nullptr, nullptr, nullptr, nullptr, 0);
		else if(kind == Y4_TRACE_SHADOW) y4_trace_ray<Y4_TRACE_SHADOW>(bvh, pr, hits[i], stack, nullptr, nullptr, nullptr, nullptr, 0);
		else
		{
			y4_trace_ray<Y4_TRACE_TSHADOW>(bvh, pr, hits[i], stack, nullptr, filt, filt_mat, &n_filt, g_shadow_depth < 16 ? g_shadow_depth : 16);
			if(!g_mat_transp.empty())       // k_trace_tshadow_queue: the filter colour travels in the t / bu / bv lanes
			{
				float fc[3] = {1.f, 1.f, 1.f};
				for(int k = 0; k < n_filt; ++k) for(int c = 0; c < 3; ++c) fc[c] *= g_mat_transp[3 * filt_mat[k] + c];
				hits[i].t = fc[0]; hits[i].bu = fc[1]; hits[i].bv = fc[2];
			}
		}
	}
}

// what k_pt_stage's queue appends do
static void push(uint32_t mask, uint32_t i, uint32_t n, std::vector<uint32_t> *path_q, std::vector<uint32_t> *shadow_q)
{
	if(shadow_q && (mask & Y4_PUSH_S0)) shadow_q->push_back(i);
	if(shadow_q && (mask & Y4_PUSH_S1)) shadow_q->push_back(n + i);
	if(path_q && (mask & Y4_PUSH_PATH)) path_q->push_back(i);
}

// textures of the next emu_pt_render calls (what y4gpu_pt_set_textures keeps in the context); texel memory stays the caller's
static std::vector<Y4PtTexture> g_textures;
static std::vector<Y4PtLayer> g_layers;
static std::vector<float> g_tri_orco, g_tri_uv;
extern "C" int emu_pt_set_textures(const y4gpu_pt_texture *textures, uint32_t n_tex, const y4gpu_pt_layer *layers, uint32_t n_layers,
                                   const float *tri_orco, const float *tri_uv, uint64_t n_tri)
{
	g_textures.clear(); g_layers.clear(); g_tri_orco.clear(); g_tri_uv.clear();
	for(uint32_t i = 0; i < n_tex; ++i) { Y4PtTexture t = derive_texture(textures[i]); t.texels = reinterpret_cast<const float4 *>(textures[i].texels); g_textures.push_back(t); }
	for(uint32_t i = 0; i < n_layers; ++i) g_layers.push_back(derive_layer(layers[i]));
	if(tri_orco && n_layers) g_tri_orco.assign(tri_orco, tri_orco + 9 * n_tri);
	if(tri_uv && n_layers) g_tri_uv.assign(tri_uv, tri_uv + 6 * n_tri);
	return 0;
}

extern "C" int emu_pt_render(const y4gpu_pt_camera *cam, const y4gpu_pt_material *mats, uint32_t n_mat, const y4gpu_pt_light *lights, uint32_t n_light,
                             const y4gpu_pt_params *params, const float *verts, const uint32_t *flags, int n_tri, const float *normals,
                             int n_samples, int offset, int aa_pass, int shard_index, int shard_count, float *film_w, float *film_rgba,
                             const uint32_t *pixel_list, uint64_t n_pixels)
{
	EmuBvh b;
	emu_build(verts, flags, n_tri, 1, b);
	Y4Bvh8 bvh; bvh.nodes = b.nodes.data(); bvh.tris = b.tris.data(); bvh.n_tris = (uint32_t) n_tri; bvh.consts = nullptr; bvh.postpone_sp_limit = 0;
	static Y4FaureHost fh = y4_faure_build();
	Y4PtScene sc;
	memset(&sc, 0, sizeof(sc));
	sc.cam = derive_camera(*cam);
	sc.bounces = params->bounces; sc.n_paths = params->path_samples; sc.rr_min_bounces = params->russian_roulette_min_bounces;
	sc.no_recursive = params->no_recursive; sc.transp_background = params->transp_background;
	sc.bg_gradient = params->background_gradient;
	sc.bg_gz = c3(params->bg_gradient[0][0], params->bg_gradient[0][1], params->bg_gradient[0][2]); sc.bg_gh = c3(params->bg_gradient[1][0], params->bg_gradient[1][1], params->bg_gradient[1][2]);
	sc.bg_sz = c3(params->bg_gradient[2][0], params->bg_gradient[2][1], params->bg_gradient[2][2]); sc.bg_sh = c3(params->bg_gradient[3][0], params->bg_gradient[3][1], params->bg_gradient[3][2]);
	sc.raydepth = params->raydepth; sc.trace_caustics = params->trace_caustics; sc.transp_refracted_background = params->transp_refracted_background;
	// ambient occlusion belongs to DirectLightIntegrator::integrate only (the path tracer never adds it to the Combined layer)
	sc.ao_samples = (params->path_samples == 0 && params->ao_samples > 0) ? params->ao_samples : 0;
	sc.ao_dist = params->ao_distance; sc.ao_col = c3(params->ao_color[0], params->ao_color[1], params->ao_color[2]);
	sc.has_specular = 0;
	for(uint32_t m = 0; m < n_mat; ++m)
	{
		if(mats[m].type == Y4GPU_MAT_MIRROR && !sc.has_specular) sc.has_specular = 1;
		// translucency needs the full material code (the no-specular kernel variant folds it away), not the ray levels
		if(mats[m].type == Y4GPU_MAT_SHINYDIFFUSE && mats[m].translucency > 0.00001f && !sc.has_specular) sc.has_specular = 1;
		if(mats[m].type == Y4GPU_MAT_SHINYDIFFUSE && (mats[m].specular_reflect > 0.00001f || mats[m].transparency > 0.00001f) && sc.has_specular < 2) sc.has_specular = 2;
		if(mats[m].type == Y4GPU_MAT_GLASS) sc.has_specular = 3;      // ray-tree mode, as y4gpu_pt_set_scene
		if(mats[m].type == Y4GPU_MAT_SHINYDIFFUSE && mats[m].specular_reflect > 0.00001f && mats[m].transparency > 0.00001f) sc.has_specular = 3;
	}
	if(sc.has_specular == 3 && params->transp_background) return -3;
	// object lights need the general kernel variant as well (not the ray levels)
	for(uint32_t l = 0; l < n_light; ++l) if((lights[l].type == Y4GPU_LIGHT_OBJECT || lights[l].type == Y4GPU_LIGHT_BACKGROUND || lights[l].type == Y4GPU_LIGHT_SPOT_SOFT || lights[l].type == Y4GPU_LIGHT_PORTAL) && !sc.has_specular) sc.has_specular = 1;
	sc.background = c3(params->background[0], params->background[1], params->background[2]);
	const Y4PtTexture *bg_tex = nullptr;
	if(params->background_gradient == 2)
	{
		if(params->bg_texture < 0 || (size_t) params->bg_texture >= g_textures.size()) return -1;
		bg_tex = &g_textures[params->bg_texture];
		sc.bg_tex = derive_texture_background(*params, bg_tex);
	}
	static Y4BgSunsky sun;
	if(params->background_gradient == 3) { sun = derive_sunsky(*params); sc.bg_sun = &sun; }
	static Y4BgDarksky dark;
	if(params->background_gradient == 4) { dark = derive_darksky(*params); sc.bg_dark = &dark; }
	sc.shadow_bias = params->shadow_bias; sc.shadow_bias_auto = params->shadow_bias_auto;
	sc.ray_min_dist = params->min_raydist; sc.ray_min_dist_auto = params->min_raydist_auto;
	sc.base_sampling_offset = params->base_sampling_offset; sc.rr_seed = params->rr_seed;
	sc.aa_light_sample_multiplier = 1.f; sc.aa_passes = params->aa_passes;
	sc.width = params->width; sc.height = params->height; sc.cx0 = params->xstart; sc.cy0 = params->ystart;
	sc.clamp_samples = params->aa_clamp_samples;
	derive_film(*params, sc);
	sc.n_lights = (int) n_light; sc.n_mats = (int) n_mat;
	int max_ls = 1;
	std::vector<Y4PtLight> hl(n_light);
	static std::vector<std::vector<float>> obj_cdf, obj_tris; static std::vector<std::vector<float4>> obj_ng;     // object lights' mesh copies
	obj_cdf.assign(n_light, {}); obj_tris.assign(n_light, {}); obj_ng.assign(n_light, {});
	for(uint32_t i = 0; i < n_light; ++i) { hl[i] = derive_light(lights[i]);
		if(lights[i].type == Y4GPU_LIGHT_PORTAL) derive_light_background(hl[i], *params, bg_tex, &sun, &dark);
		if(lights[i].type == Y4GPU_LIGHT_OBJECT || lights[i].type == Y4GPU_LIGHT_PORTAL)
		{
			obj_tris[i].assign(lights[i].object_tris, lights[i].object_tris + 9ull * lights[i].object_n_tris);
			derive_object_light(hl[i], obj_tris[i].data(), lights[i].object_n_tris, obj_cdf[i], obj_ng[i]);
			hl[i].obj_tris = obj_tris[i].data(); hl[i].obj_cdf = obj_cdf[i].data(); hl[i].obj_ng = obj_ng[i].data();
		}
		if(lights[i].type == Y4GPU_LIGHT_BACKGROUND)
		{
			static std::vector<int> bg_off; static std::vector<float> bg_func, bg_cdf, bg_inv;     // (one background light per scene)
			derive_background_light(hl[i], lights[i], *params, bg_off, bg_func, bg_cdf, bg_inv, bg_tex, &sun, &dark);
			hl[i].bg_off = bg_off.data(); hl[i].bg_func = bg_func.data(); hl[i].bg_cdf = bg_cdf.data(); hl[i].bg_inv_int = bg_inv.data();
		}
		const int ns = y4_light_num_samples(sc, hl[i]); if(ns > max_ls) max_ls = ns; }
	std::vector<float4> tri_shade(n_tri);
	for(int t = 0; t < n_tri; ++t) tri_shade[t] = y4_tri_shade(verts, flags, n_mat, (uint64_t) t);
	std::vector<Y4PtMaterial> hm(n_mat);
	for(uint32_t i = 0; i < n_mat; ++i) hm[i] = derive_material(mats[i]);
	sc.tshadow_filter = 0; g_mat_transp.clear(); g_shadow_depth = params->transp_shadows ? params->shadow_depth : 16;
	for(uint32_t i = 0; i < n_mat; ++i)
		if(hm[i].is_transparent && params->transp_shadows)      // as y4gpu_pt_set_scene
		{
			if(g_mat_transp.empty()) g_mat_transp.assign(3 * (size_t) n_mat, 0.f);
			float accum = 1.f;
			if(hm[i].is_mirror) accum = 1.f - 1.f * hm[i].mirror_strength;      // 1 - kr * mirror_strength, kr = 1 without the Fresnel effect (refused with it below)
			accum *= hm[i].transparency_strength * accum;
			const C3 tc = y4_shiny_tcol(hm[i]);
			g_mat_transp[3 * i] = accum * tc.r; g_mat_transp[3 * i + 1] = accum * tc.g; g_mat_transp[3 * i + 2] = accum * tc.b;
			sc.tshadow_filter = 1;
		}
	sc.lights = hl.data(); sc.mats = hm.data(); sc.tri_shade = tri_shade.data(); sc.normals = normals;
	sc.textures = g_textures.empty() ? nullptr : g_textures.data(); sc.layers = g_layers.empty() ? nullptr : g_layers.data();
	sc.tri_orco = g_tri_orco.empty() ? nullptr : g_tri_orco.data(); sc.tri_uv = g_tri_uv.empty() ? nullptr : g_tri_uv.data();
	sc.lds.sigma = fh.sigma.data(); sc.lds.offset = fh.offset; sc.lds.prime = fh.prime; sc.lds.inv_prime = fh.inv_prime; sc.lds.sigma_d = fh.sigma_d.data(); sc.lds.magic = fh.magic;

	// same pixel-sample enumeration as k_pt_fill (y4_pt.cu)
	const uint32_t ns_local = shard_index < n_samples ? (uint32_t) ((n_samples - shard_index + shard_count - 1) / shard_count) : 0u;
	const uint32_t n = (uint32_t) (pixel_list ? n_pixels : (uint64_t) sc.width * sc.height) * ns_local;
	if(n == 0) return 0;
	std::vector<uint32_t> pix(n), samp(n), fl(n), lnum(n);
	std::vector<float4> col(n), pcol(n), thr(n), pthr(n), lcol(n), acc_a(n), acc_b(n), lsamp(n), scontrib(2ull * n);
	std::vector<Y4Ray> ray(n), ray2(n), sray(2ull * n), ray0(n);
	std::vector<float4> sw(n), acc(n);
	std::vector<Y4Hit> hit(n), hit0(n), shit(2ull * n);
	Y4PtWave W;
	W.n = n; W.pix = pix.data(); W.samp = samp.data(); W.flags = fl.data(); W.lnum = lnum.data();
	W.col = &col[0].x; W.pcol = &pcol[0].x; W.thr = &thr[0].x; W.lcol = &lcol[0].x; W.acc_a = &acc_a[0].x; W.acc_b = &acc_b[0].x; W.lsamp = &lsamp[0].x; W.pthr = &pthr[0].x;
	W.ray = ray.data(); W.ray_next = ray2.data(); W.hit = hit.data(); W.hit0 = hit0.data(); W.sray = sray.data(); W.shit = shit.data(); W.scontrib = &scontrib[0].x;
	W.ray0 = ray0.data(); W.sw = &sw[0].x; W.acc = &acc[0].x;
	const bool tree = sc.has_specular == 3;
	const int tree_depth = (sc.raydepth < 19 ? sc.raydepth : 19) + 2;
	std::vector<Y4Ray> stk_ray(tree ? (size_t) tree_depth * n : 0); std::vector<float4> stk_sw(tree ? (size_t) tree_depth * n : 0); std::vector<uint32_t> stk_n(tree ? n : 0);
	W.stk_ray = stk_ray.data(); W.stk_sw = tree ? &stk_sw[0].x : nullptr; W.stk_n = stk_n.data();
	Y4PtPass ps; ps.n_samples = n_samples; ps.pass_offs = (uint32_t) offset + sc.base_sampling_offset; ps.aa_pass = aa_pass;
	for(uint32_t i = 0; i < n; ++i) { pix[i] = pixel_list ? pixel_list[i / ns_local] : i / ns_local; samp[i] = (uint32_t) shard_index + (i % ns_local) * (uint32_t) shard_count; }
	const int shadow_kind = params->transp_shadows ? Y4_TRACE_TSHADOW : Y4_TRACE_SHADOW;
	// same stage sequence as y4gpu_pt_render_pass (y4_pt.cu), queues as vectors
	for(uint32_t i = 0; i < n; ++i) y4_pt_gen(sc, W, ps, i);
	trace(bvh, Y4_TRACE_CLOSEST, W.ray, nullptr, n, W.hit);
	std::vector<std::pair<int, int>> all_lk;
	for(int l = 0; l < sc.n_lights; ++l)
		for(int k = 0; k < y4_light_num_samples(sc, hl[l]); ++k) all_lk.push_back(std::make_pair(l, k));
	for(int k = 0; k < sc.ao_samples; ++k) all_lk.push_back(std::make_pair(sc.n_lights, k));
	std::vector<uint32_t> q_hit0, q_shadow, q_cur, q_next;
	const int l0 = all_lk.empty() ? -1 : all_lk[0].first;
	std::vector<uint32_t> q_spec, q_spec_next;
	auto primary = [&](uint32_t i, int level)
	{
		const uint32_t p = tree ? y4_stage_primary3(sc, W, ps, i, l0, level) : y4_stage_primary(sc, W, ps, i, l0, level);
		push(p, i, n, &q_hit0, &q_shadow);
		if(p & Y4_PUSH_SPEC) q_spec_next.push_back(i);
	};
	// ray levels, as y4gpu_pt_render_pass: whole mirror chains inside the camera-hit step when only pure mirrors exist,
	// one estimate per level when shinydiffuse materials have a mirror component
	const int max_level = (sc.has_specular && sc.raydepth > 0) ? (sc.raydepth < 19 ? sc.raydepth : 19) : 0;
	const bool level_runs = sc.has_specular >= 2;
	for(int level = 0; level <= (tree ? (1 << 20) : (level_runs ? max_level : 0)); ++level)     // tree mode: runs, until no tree ray is left
	{
	q_hit0.clear(); q_shadow.clear();
	if(level == 0) for(uint32_t i = 0; i < n; ++i) primary(i, 0);
	else
	{
		q_spec.swap(q_spec_next); q_spec_next.clear();
		trace(bvh, Y4_TRACE_CLOSEST, W.ray0, q_spec.data(), q_spec.size(), W.hit);
		for(uint32_t i : q_spec) primary(i, level);
	}
	for(int lv = 1; !level_runs && lv <= max_level; ++lv)
	{
		q_spec.swap(q_spec_next); q_spec_next.clear();
		trace(bvh, Y4_TRACE_CLOSEST, W.ray0, q_spec.data(), q_spec.size(), W.hit);
		for(uint32_t i : q_spec) primary(i, lv);
	}
	for(size_t j = 0; j < all_lk.size(); ++j)
	{
		trace(bvh, shadow_kind, W.sray, q_shadow.data(), q_shadow.size(), W.shit);
		if(j + 1 == all_lk.size()) break;
		q_shadow.clear();
		for(uint32_t i : q_hit0) push(y4_stage_direct_step(sc, W, ps, i, true, all_lk[j].first, all_lk[j].second, all_lk[j + 1].first, all_lk[j + 1].second), i, n, nullptr, &q_shadow);
	}
	if(sc.n_paths == 0 && !all_lk.empty())
		for(uint32_t i : q_hit0) y4_stage_start(sc, W, ps, i, all_lk.back().first, all_lk.back().second, 0);
	Y4Ray *const ray_buf[2] = {ray.data(), ray2.data()};
	for(int sub = 0; sub < sc.n_paths; ++sub)
	{
		q_cur.clear();
		const bool res = sub == 0 && !all_lk.empty();
		W.ray = ray_buf[0]; W.ray_next = ray_buf[0];
		for(uint32_t i : q_hit0) push(y4_stage_start(sc, W, ps, i, res ? all_lk.back().first : -1, res ? all_lk.back().second : -1, sub), i, n, &q_cur, nullptr);
		for(int depth = 0; depth <= sc.bounces; ++depth)
		{
			const bool closing = depth == sc.bounces;
			if(closing && !sc.n_lights) break;
			W.ray = ray_buf[depth & 1]; W.ray_next = ray_buf[(depth + 1) & 1];
			if(!closing) trace(bvh, Y4_TRACE_CLOSEST, W.ray, q_cur.data(), q_cur.size(), W.hit);
			q_shadow.clear(); q_next.clear();
			for(uint32_t i : q_cur) push(y4_stage_path(sc, W, ps, i, depth, sub, sc.n_lights ? max_ls - 1 : -1), i, n, &q_next, &q_shadow);
			if(closing) break;
			if(sc.n_lights)
			{
				trace(bvh, shadow_kind, W.sray, q_shadow.data(), q_shadow.size(), W.shit);
				for(int k = 1; k < max_ls; ++k)
				{
					q_shadow.clear();
					for(uint32_t i : q_cur) push(y4_stage_direct_step(sc, W, ps, i, false, -1, k - 1, -1, k), i, n, nullptr, &q_shadow);
					trace(bvh, shadow_kind, W.sray, q_shadow.data(), q_shadow.size(), W.shit);
				}
			}
			q_cur.swap(q_next);
		}
	}
	W.ray = ray_buf[0];
	if(tree)
	{
		for(uint32_t i : q_hit0) if(y4_stage_finish3(sc, W, ps, i) & Y4_PUSH_SPEC) q_spec_next.push_back(i);
		if(q_spec_next.empty()) break;
	}
	else if(level_runs && level < max_level)
		for(uint32_t i : q_hit0) if(y4_stage_level_finish(sc, W, ps, i, level) & Y4_PUSH_SPEC) q_spec_next.push_back(i);
	}   // ray levels
	for(uint32_t i = 0; i < n; ++i) y4_pt_finish_splat(sc, W, ps, i, film_w, film_rgba);
	return 0;
}

// ImageFilm::nextPass(adaptive) flags, the two phases of y4gpu_pt_next_pass as serial loops
extern "C" uint64_t emu_film_flags(const float *weights, const float *rgba, int width, int height, const y4gpu_pt_aa_noise *aa, uint8_t *flags)
{
	const size_t npx = (size_t) width * height;
	for(size_t p = 0; p < npx; ++p) y4_film_flag_init(weights, flags, p);
	for(int y = 0; y < height - 1; ++y)
		for(int x = 0; x < width - 1; ++x) y4_film_flag_detect(weights, rgba, width, height, *aa, flags, x, y);
	uint64_t n = 0;
	for(size_t p = 0; p < npx; ++p) n += flags[p];
	return n;
}
