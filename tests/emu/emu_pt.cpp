// UNIT-TEST EMULATION ONLY: runs the Y4_HD wavefront stages of the path tracer
// (libyafaray_b200/csrc/y4_pt_stages.cuh) as serial loops on the CPU, with the emulated BVH8
// traversal, so the stage logic can be checked against the reference renderer where no GPU exists.
// Never linked into liby4gpu.so.
#include "../../libyafaray_b200/csrc/y4_trace.cuh"
#include "../../libyafaray_b200/csrc/y4_pt_host.h"
#include "../../libyafaray_b200/csrc/y4_faure.h"
#include <vector>

struct EmuBvh
{
	std::vector<Y4U4> nodes;
	std::vector<Y4Tri> tris;
	uint32_t n_nodes = 0;
	int n = 0;
};
void emu_build(const float *verts, const uint32_t *flags, int n, int leaf_max, EmuBvh &out);

static void trace(const Y4Bvh8 &bvh, int kind, const Y4Ray *rays, size_t n, Y4Hit *hits)
{
	for(size_t i = 0; i < n; ++i)
	{
		const Y4PreparedRay pr = y4_prepare_ray(rays[i], kind);
		uint2 stack[Y4_STACK_SIZE];
		hits[i].t = 0.f; hits[i].bu = 0.f; hits[i].bv = 0.f; hits[i].prim = -1;
		if(!(pr.t_hi > pr.t_lo)) continue;
		if(kind == Y4_TRACE_CLOSEST) y4_trace_ray<Y4_TRACE_CLOSEST>(bvh, pr, hits[i], stack, nullptr, nullptr, nullptr, nullptr, 0);
		else y4_trace_ray<Y4_TRACE_SHADOW>(bvh, pr, hits[i], stack, nullptr, nullptr, nullptr, nullptr, 0);
	}
}

extern "C" int emu_pt_render(const y4gpu_pt_camera *cam, const y4gpu_pt_material *mats, uint32_t n_mat, const y4gpu_pt_light *lights, uint32_t n_light,
                             const y4gpu_pt_params *params, const float *verts, const uint32_t *flags, int n_tri, const float *normals,
                             int n_samples, int offset, int aa_pass, int shard_index, int shard_count, float *film_w, float *film_rgba)
{
	EmuBvh b;
	emu_build(verts, flags, n_tri, 1, b);
	Y4Bvh8 bvh; bvh.nodes = b.nodes.data(); bvh.tris = b.tris.data(); bvh.n_tris = (uint32_t) n_tri; bvh.consts = nullptr;
	static Y4FaureHost fh = y4_faure_build();
	Y4PtScene sc;
	memset(&sc, 0, sizeof(sc));
	sc.cam = derive_camera(*cam);
	sc.bounces = params->bounces; sc.n_paths = params->path_samples; sc.rr_min_bounces = params->russian_roulette_min_bounces;
	sc.no_recursive = params->no_recursive; sc.transp_background = params->transp_background;
	sc.background = c3(params->background[0], params->background[1], params->background[2]);
	sc.shadow_bias = params->shadow_bias; sc.shadow_bias_auto = params->shadow_bias_auto;
	sc.ray_min_dist = params->min_raydist; sc.ray_min_dist_auto = params->min_raydist_auto;
	sc.base_sampling_offset = params->base_sampling_offset; sc.rr_seed = params->rr_seed;
	sc.aa_light_sample_multiplier = 1.f; sc.aa_passes = params->aa_passes;
	sc.width = params->width; sc.height = params->height; sc.cx0 = params->xstart; sc.cy0 = params->ystart;
	sc.clamp_samples = params->aa_clamp_samples;
	derive_film(*params, sc);
	sc.n_lights = (int) n_light; sc.n_mats = (int) n_mat;
	int max_ls = 1;
	for(uint32_t i = 0; i < n_light; ++i) { sc.lights[i] = derive_light(lights[i]); const int ns = y4_light_num_samples(sc, sc.lights[i]); if(ns > max_ls) max_ls = ns; }
	std::vector<Y4PtMaterial> hm(n_mat);
	for(uint32_t i = 0; i < n_mat; ++i) hm[i] = derive_material(mats[i]);
	sc.mats = hm.data(); sc.verts = verts; sc.tri_flags = flags; sc.normals = normals;
	sc.lds.sigma = fh.sigma.data(); sc.lds.offset = fh.offset; sc.lds.prime = fh.prime; sc.lds.inv_prime = fh.inv_prime;

	// same pixel-sample enumeration as k_pt_fill (y4_pt.cu)
	const uint32_t ns_local = shard_index < n_samples ? (uint32_t) ((n_samples - shard_index + shard_count - 1) / shard_count) : 0u;
	const uint32_t n = (uint32_t) sc.width * sc.height * ns_local;
	if(n == 0) return 0;
	std::vector<uint32_t> pix(n), samp(n), fl(n), lnum(n);
	std::vector<float> col(4ull * n), pcol(4ull * n), thr(4ull * n), lcol(4ull * n), acc_a(4ull * n), acc_b(4ull * n), scontrib(8ull * n);
	std::vector<Y4Ray> ray(n), sray(2ull * n);
	std::vector<Y4Hit> hit(n), hit0(n), shit(2ull * n);
	Y4PtWave W;
	W.n = n; W.pix = pix.data(); W.samp = samp.data(); W.flags = fl.data(); W.lnum = lnum.data();
	W.col = col.data(); W.pcol = pcol.data(); W.thr = thr.data(); W.lcol = lcol.data(); W.acc_a = acc_a.data(); W.acc_b = acc_b.data();
	W.ray = ray.data(); W.hit = hit.data(); W.hit0 = hit0.data(); W.sray = sray.data(); W.shit = shit.data(); W.scontrib = scontrib.data();
	Y4PtPass ps; ps.n_samples = n_samples; ps.pass_offs = (uint32_t) offset + sc.base_sampling_offset; ps.aa_pass = aa_pass;
	for(uint32_t i = 0; i < n; ++i) { pix[i] = i / ns_local; samp[i] = (uint32_t) shard_index + (i % ns_local) * (uint32_t) shard_count; }
	for(uint32_t i = 0; i < n; ++i) y4_pt_gen(sc, W, ps, i);
	trace(bvh, Y4_TRACE_CLOSEST, W.ray, n, W.hit);
	for(uint32_t i = 0; i < n; ++i) y4_pt_primary(sc, W, i);
	for(int l = 0; l < sc.n_lights; ++l)
		for(int k = 0; k < y4_light_num_samples(sc, sc.lights[l]); ++k)
		{
			for(uint32_t i = 0; i < n; ++i) y4_pt_direct(sc, W, ps, i, true, l, k);
			trace(bvh, Y4_TRACE_SHADOW, W.sray, 2ull * n, W.shit);
			for(uint32_t i = 0; i < n; ++i) y4_pt_direct_resolve(sc, W, i, true, l, k);
		}
	for(int sub = 0; sub < sc.n_paths; ++sub)
		for(int depth = 0; depth < sc.bounces; ++depth)
		{
			for(uint32_t i = 0; i < n; ++i) y4_pt_bounce_sample(sc, W, ps, i, depth, sub);
			trace(bvh, Y4_TRACE_CLOSEST, W.ray, n, W.hit);
			for(uint32_t i = 0; i < n; ++i) y4_pt_bounce_hit(sc, W, ps, i, depth, sub);
			if(sc.n_lights == 0) continue;
			for(int k = 0; k < max_ls; ++k)
			{
				for(uint32_t i = 0; i < n; ++i) y4_pt_direct(sc, W, ps, i, false, -1, k);
				trace(bvh, Y4_TRACE_SHADOW, W.sray, 2ull * n, W.shit);
				for(uint32_t i = 0; i < n; ++i) y4_pt_direct_resolve(sc, W, i, false, -1, k);
			}
		}
	for(uint32_t i = 0; i < n; ++i) y4_pt_finish_splat(sc, W, ps, i, film_w, film_rgba);
	return 0;
}
