// Wavefront path tracer back end (PathIntegrator on the GPU): scene derivation on the host,
// stage kernels, the per-pass launch sequence and the film.  Per-path logic: y4_pt_stages.cuh.
#include "y4_ctx.h"
#include "y4_pt_stages.cuh"
#include "y4_faure.h"
#include "y4_film.cuh"
#include <cub/cub.cuh>
#include <cstring>
#include <vector>

constexpr int kMaxLanes = 4;

struct Y4Pt
{
	Y4PtScene h_scene;                 // passed to every kernel by value (device pointers inside)
	std::vector<Y4PtLight> h_lights;   // host copy of the derived lights
	std::vector<char *> light_blobs;   // device memory owned by object lights (their mesh copy + cdf)
	Y4PtLight *d_lights = nullptr;
	Y4PtMaterial *d_mats = nullptr;
	float4 *d_tri_shade = nullptr;
	uint32_t mats_cap = 0, lights_cap = 0; uint64_t tri_cap = 0, normals_cap = 0; size_t film_cap = 0;
	float *d_normals = nullptr;
	// textures (y4gpu_pt_set_textures): one texel allocation per texture, the derived texture / layer tables, per-triangle orco / uv
	std::vector<float4 *> d_texels;
	Y4PtTexture *d_textures = nullptr; Y4PtLayer *d_layers = nullptr; float *d_tri_orco = nullptr, *d_tri_uv = nullptr;
	uint32_t n_textures = 0, n_layers = 0;
	std::vector<Y4PtTexture> h_textures;        // the descriptors as uploaded (texels = device pointers): a textureback looks its image up here
	Y4BgSunsky *d_bg_sun = nullptr;             // sunsky background state (scene + lights point here)
	Y4BgDarksky *d_bg_dark = nullptr;           // darksky background state
	uint8_t *d_sigma = nullptr; uint16_t *d_offset = nullptr, *d_prime = nullptr; double *d_inv_prime = nullptr, *d_sigma_d = nullptr; unsigned long long *d_magic = nullptr;
	bool scene_set = false;
	y4gpu_pt_params params = {};
	int max_light_samples = 1;
	int stage_min_blocks = 2;
	// film
	float *d_film_w = nullptr, *d_film_rgba = nullptr;      // in use
	float *own_film_w = nullptr, *own_film_rgba = nullptr;  // owned (when the caller gave none)
	// wave state: several wavefronts in flight on separate streams ("lanes"), so that the issue-bound traversal of one
	// shares the SMs with the latency-bound shading of the other
	struct Lane
	{
		uint32_t cap = 0;
		Y4PtWave wave = {};
		void *mem = nullptr;
		uint32_t *q_hit0 = nullptr, *q_path[2] = {nullptr, nullptr}, *q_shadow = nullptr, *q_spec[2] = {nullptr, nullptr};   // index queues (in mem)
		// Compacting bounce loop: the paths of the bounce loop live in two dense "path waves" and move from one to the other at
		// every bounce, live paths packed at the front (k_pt_path_compact); corig = the pixel-sample slot a path belongs to
		Y4PtWave cw[2] = {{}, {}}; uint32_t *corig[2] = {nullptr, nullptr}; void *cmem = nullptr; uint32_t ccap = 0;
		// ray-tree mode: per pixel-sample stack of pending sibling branches (ray 32 B + weight 16 B per entry) + its height
		void *tmem = nullptr; uint32_t tcap = 0; int tdepth = 0;
		cudaStream_t stream = nullptr;      // lane 0: the context's stream; lane 1: own stream, forked/joined per pass
		cudaEvent_t done = nullptr;
	} lanes[kMaxLanes];
	int n_lanes = 2, trace_blocks_shared = 2;
	int compact = 1;             // env Y4GPU_PT_COMPACT=0: the bounce loop over index queues into the pixel-sample wave (no path movement)
	cudaEvent_t ev_fork = nullptr;
	// per-launch counters: 64-bit slots, zeroed once per pass; a queue length is the low word of its slot
	unsigned long long *d_counters = nullptr, *h_counters = nullptr; size_t counters_cap = 0;
	float *d_mat_transp = nullptr; uint32_t mat_transp_cap = 0;   // [n_mat * 3] getTransparency colours (filtered shadows)
	std::vector<std::pair<uint32_t, int>> counter_kinds;      // (slot, kind) of the last pass: 0 traced path queue, 1 shadow queue, 2 camera-hit queue, 3 closing path queue, 4 ray-level queue (mirror reflections)
	uint64_t dense_closest = 0;
	bool counters_read = true;
	uint32_t *d_pixel_list = nullptr; uint64_t pixel_list_cap = 0;
	// adaptive AA: per-pixel resample flags, their compacted raster-order list, CUB scratch
	uint8_t *d_aa_flags = nullptr; uint32_t *d_aa_list = nullptr; uint64_t *d_aa_count = nullptr; void *d_aa_tmp = nullptr;
	size_t aa_cap = 0, aa_tmp_bytes = 0;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	std::vector<cudaEvent_t> prof_events;     // profile mode: one event after every launch
	y4gpu_pt_stats stats = {};
};

#include "y4_pt_host.h"

static void pt_free_textures(Y4Pt *pt)
{
	for(float4 *p : pt->d_texels) cudaFree(p);
	pt->d_texels.clear();
	cudaFree(pt->d_textures); cudaFree(pt->d_layers); cudaFree(pt->d_tri_orco); cudaFree(pt->d_tri_uv);
	pt->d_textures = nullptr; pt->d_layers = nullptr; pt->d_tri_orco = nullptr; pt->d_tri_uv = nullptr;
	pt->n_textures = pt->n_layers = 0;
	pt->h_textures.clear();
}

// ---- kernels -------------------------------------------------------------------------------------------
// Dense kernels (camera rays, splat) run one thread per pixel-sample of the wave.  Stage kernels run one
// thread per entry of a device-built index queue (the live paths of the wavefront) and append to the
// shadow-ray queue and to the stage's output path queue with one atomic per warp and queue.
namespace {

constexpr int kPtThreads = 256;

enum { kStPrimary, kStDirectStep, kStStart, kStPath, kStLevelFinish, kStPrimary3, kStFinish3 };      // ...3: ray-tree mode (has_specular == 3)
struct StageArgs { int all_lights, l_prev, k_prev, l, k, depth, sub, level; };
struct StageQueues
{
	const uint32_t *in_q, *in_count;     // null: dense over the wave
	uint32_t *out_q, *out_count;         // paths that continue
	uint32_t *sh_q, *sh_count;           // shadow rays to trace: slot index into W.sray / W.shit
	uint32_t *spec_q, *spec_count;       // camera-hit stage only: pixel-samples whose ray a mirror reflected (next ray level)
};

__global__ void __launch_bounds__(kPtThreads) k_pt_gen(const __grid_constant__ Y4PtScene sc, Y4PtWave W, Y4PtPass ps, const uint32_t *__restrict__ pixel_list,
                                                        uint64_t first, uint64_t n_pix, uint32_t ns_local, int shard_index, int shard_count)
{
	// Enumeration of the pass's pixel-samples: blocks of 32 consecutive pixels, inside a block sample-major, so that the lanes of
	// a warp hold the same sample index of 32 neighbouring pixels.  Their film updates then go to neighbouring addresses
	// (one vector RED per lane over 512 contiguous bytes) instead of 16 lanes adding to the same pixel.
	for(uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < W.n; i += gridDim.x * blockDim.x)
	{
		const uint64_t g = first + i;
		const uint64_t per_block = 32ull * ns_local;
		const uint64_t blk = g / per_block;
		const uint32_t r = (uint32_t) (g - blk * per_block);
		const uint64_t left = n_pix - 32ull * blk;
		const uint32_t in_blk = left < 32ull ? (uint32_t) left : 32u;
		const uint64_t slot = 32ull * blk + r % in_blk;
		W.pix[i] = pixel_list ? pixel_list[slot] : (uint32_t) slot;
		W.samp[i] = (uint32_t) shard_index + (r / in_blk) * (uint32_t) shard_count;
		y4_pt_gen(sc, W, ps, i);
	}
}

// y4gpu_libm_eval: the inverse trigonometric restatements on the device (self-check against the host's libm)
__global__ void k_libm_eval(int fn, const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ out, uint64_t n)
{
	for(uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x)
	{
		const float x = a[i];
		out[i] = fn == 0 ? y4_libm_atanf(x) : (fn == 1 ? y4_libm_asinf(x) : (fn == 2 ? y4_libm_acosf(x) : y4_libm_atan2f(x, b[i])));
	}
}

template<int STAGE, int MINB, bool SPEC>
__global__ void __launch_bounds__(kPtThreads, MINB) k_pt_stage(const __grid_constant__ Y4PtScene sc, Y4PtWave W, Y4PtPass ps, StageArgs a, StageQueues Q)
{
	const uint32_t lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;
	const uint32_t count = Q.in_count ? *Q.in_count : W.n;        // in_count without in_q: the first *in_count slots (a dense path wave)
	const uint32_t stride = gridDim.x * blockDim.x;
	for(uint32_t base = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < count; base += stride)
	{
		const uint32_t q = base + lane;
		uint32_t push = 0u, i = 0u;
		if(q < count)
		{
			i = Q.in_q ? Q.in_q[q] : q;
			if(STAGE == kStPrimary) push = y4_stage_primary<SPEC>(sc, W, ps, i, a.l, a.level);
			else if(STAGE == kStDirectStep) push = y4_stage_direct_step<SPEC>(sc, W, ps, i, a.all_lights != 0, a.l_prev, a.k_prev, a.l, a.k);
			else if(STAGE == kStStart) push = y4_stage_start<SPEC>(sc, W, ps, i, a.l_prev, a.k_prev, a.sub);
			else if(STAGE == kStLevelFinish) push = y4_stage_level_finish(sc, W, ps, i, a.level);
			else if(STAGE == kStPrimary3) push = y4_stage_primary3(sc, W, ps, i, a.l, a.level);
			else if(STAGE == kStFinish3) push = y4_stage_finish3(sc, W, ps, i);
			else push = y4_stage_path<SPEC>(sc, W, ps, i, a.depth, a.sub, a.k_prev);
		}
		// queue appends: one atomic per warp; when a stage feeds both queues their two 32-bit lengths share one 64-bit
		// counter slot (path count low, shadow count high) and a single 64-bit atomic reserves both ranges
		if(STAGE == kStLevelFinish || STAGE == kStFinish3) push &= Y4_PUSH_SPEC;
		const uint32_t m0 = (STAGE != kStStart) ? __ballot_sync(0xffffffffu, push & Y4_PUSH_S0) : 0u;
		const uint32_t m1 = (STAGE != kStStart) ? __ballot_sync(0xffffffffu, push & Y4_PUSH_S1) : 0u;
		const uint32_t mp = (STAGE != kStDirectStep) ? __ballot_sync(0xffffffffu, push & Y4_PUSH_PATH) : 0u;
		if(m0 | m1 | mp)
		{
			const uint32_t n_sh = (uint32_t) (__popc(m0) + __popc(m1)), n_p = (uint32_t) __popc(mp);
			uint32_t at_sh = 0u, at_p = 0u;
			if(lane == 0)
			{
				if((STAGE == kStPrimary || STAGE == kStPath || STAGE == kStPrimary3) && Q.sh_count == Q.out_count + 1)
				{
					const unsigned long long old = atomicAdd(reinterpret_cast<unsigned long long *>(Q.out_count), ((unsigned long long) n_sh << 32) | n_p);
					at_p = (uint32_t) old; at_sh = (uint32_t) (old >> 32);
				}
				else
				{
					if(n_sh) at_sh = atomicAdd(Q.sh_count, n_sh);
					if(n_p) at_p = atomicAdd(Q.out_count, n_p);
				}
			}
			at_sh = __shfl_sync(0xffffffffu, at_sh, 0);
			at_p = __shfl_sync(0xffffffffu, at_p, 0);
			if(push & Y4_PUSH_S0) Q.sh_q[at_sh + __popc(m0 & lt_mask)] = i;
			if(push & Y4_PUSH_S1) Q.sh_q[at_sh + __popc(m0) + __popc(m1 & lt_mask)] = W.n + i;
			if(push & Y4_PUSH_PATH) Q.out_q[at_p + __popc(mp & lt_mask)] = i;
		}
		if((STAGE == kStPrimary || STAGE == kStLevelFinish || STAGE == kStPrimary3 || STAGE == kStFinish3) && Q.spec_q)
		{
			const uint32_t ms = __ballot_sync(0xffffffffu, push & Y4_PUSH_SPEC);
			if(ms)
			{
				uint32_t at = 0u;
				if(lane == 0) at = atomicAdd(Q.spec_count, (uint32_t) __popc(ms));
				at = __shfl_sync(0xffffffffu, at, 0);
				if(push & Y4_PUSH_SPEC) Q.spec_q[at + __popc(ms & lt_mask)] = i;
			}
		}
	}
}

// ---- the per-bounce path stage with compaction ---------------------------------------------------------------
// Late bounces of a wavefront keep a shrinking, scattered subset of the pixel-samples alive.  Run over index queues into the
// pixel-sample wave (k_pt_stage<kStPath>) every 16-byte field of such a path costs a whole 32-byte sector, every 4-byte field
// eight times its size, and partial-sector writes are read-modify-write in L2: ncu measured 650 MB of DRAM traffic per launch
// from bounce 2 to bounce 7 while the algorithmic bytes fell to 150 MB.  Here the paths of the bounce loop live in their own
// dense "path waves": a stage reads wave Wr (entry q = slot q, fully coalesced), runs the unchanged per-path logic
// (y4_stage_path) on a copy of the path's state staged in SHARED memory -- the stage's wave argument simply points there --
// and writes the paths that go on to the FRONT of wave Ww at the position their queue append reserved.  A path that ends
// hands its path_col to the pixel-sample it belongs to (W0.pcol[orig]): the only scattered access left.  Traversal then
// runs over dense rays as well.  The arithmetic per path is exactly that of the queue version.
constexpr int kCompactThreads = 128;
constexpr size_t kCompactSmem = (size_t) kCompactThreads * 17 * sizeof(float4);      // 272 B of staged state per thread

// Path-wave traffic of the compacting kernel is touched exactly once per launch: streaming loads / stores (evict-first) keep the
// small scene tables the stage logic gathers from (materials, lights, per-triangle records, Faure permutations) resident in L1.
__device__ __forceinline__ float4 ldcs4(const float *p, uint32_t i) { return __ldcs(reinterpret_cast<const float4 *>(p) + i); }
__device__ __forceinline__ void stcs4(float *p, uint32_t i, float4 v) { __stcs(reinterpret_cast<float4 *>(p) + i, v); }
__device__ __forceinline__ float4 ldcs_q(const void *p, uint32_t q) { return __ldcs(reinterpret_cast<const float4 *>(p) + q); }      // q-th 16-byte word
__device__ __forceinline__ void stcs_q(void *p, uint32_t q, float4 v) { __stcs(reinterpret_cast<float4 *>(p) + q, v); }

template<int MINB, bool SPEC>
__global__ void __launch_bounds__(kCompactThreads, MINB) k_pt_path_compact(const __grid_constant__ Y4PtScene sc, Y4PtWave Wr, Y4PtWave Ww, Y4PtWave W0, Y4PtPass ps,
                                                                           StageArgs a, StageQueues Q, const uint32_t *__restrict__ orig_r, uint32_t *__restrict__ orig_w)
{
	extern __shared__ float4 y4_stage_smem[];
	constexpr uint32_t T = kCompactThreads;
	Y4PtWave S = {};
	{
		float4 *p = y4_stage_smem;
		S.n = T;
		S.lsamp = (float *) p; p += T; S.thr = (float *) p; p += T; S.pthr = (float *) p; p += T; S.pcol = (float *) p; p += T;
		S.lcol = (float *) p; p += T;          // (acc_a / acc_b: only lights with several samples use them, and those scenes take the queue version)
		S.ray = (Y4Ray *) p; S.ray_next = S.ray; p += 2 * T;        // the stage reads its ray before it writes the next one: one slot
		S.hit = (Y4Hit *) p; p += T;
		S.sray = (Y4Ray *) p; p += 4 * T; S.shit = (Y4Hit *) p; p += 2 * T; S.scontrib = (float *) p; p += 2 * T;
		S.flags = (uint32_t *) p; S.pix = S.flags + T; S.samp = S.pix + T; S.lnum = S.samp + T;
	}
	const uint32_t tid = threadIdx.x, lane = tid & 31u, lt_mask = (1u << lane) - 1u;
	const uint32_t count = *Q.in_count;
	const uint32_t stride = gridDim.x * T;
	const bool closing = a.depth >= sc.bounces;
	// (tried: regrouping the paths that go on after the fold into full warps inside the CTA, between y4_stage_path_fold and
	// y4_stage_path_vertex -- three barriers per iteration cost more than the 30 % idle lanes: 18.0 -> 18.6 ms per pass)
	for(uint32_t base = blockIdx.x * T + (tid & ~31u); base < count; base += stride)
	{
		const uint32_t q = base + lane;
		const bool have = q < count;
		uint32_t push = 0u, orig = 0u;
		if(have)
		{
			const uint32_t j = Q.in_q ? Q.in_q[q] : q;
			orig = orig_r ? __ldcs(orig_r + j) : j;
			const uint32_t f = __ldcs(Wr.flags + j);
			// copy-in: everything y4_stage_path can read, requested together (dense: one coalesced access per field and warp)
			const float4 v_pcol = ldcs4(Wr.pcol, j), v_pthr = ldcs4(Wr.pthr, j), v_sc0 = ldcs4(Wr.scontrib, j);
			float4 v_lsamp = v_pcol, v_thr = v_pcol, v_hit = v_pcol, v_r0 = v_pcol, v_r1 = v_pcol;
			uint32_t pix = 0u, samp = 0u;
			if(!closing)
			{
				v_lsamp = ldcs4(Wr.lsamp, j); v_thr = ldcs4(Wr.thr, j); v_hit = ldcs_q(Wr.hit, j);
				v_r0 = ldcs_q(Wr.ray, 2u * j); v_r1 = ldcs_q(Wr.ray, 2u * j + 1u);
				pix = __ldcs(Wr.pix + j); samp = __ldcs(Wr.samp + j);
			}
			if(f & Y4_PF_C1) st4(S.scontrib, T + tid, ldcs4(Wr.scontrib, Wr.n + j));
			if(f & Y4_PF_S0) *reinterpret_cast<float4 *>(S.shit + tid) = ldcs_q(Wr.shit, j);
			if(f & Y4_PF_S1) *reinterpret_cast<float4 *>(S.shit + T + tid) = ldcs_q(Wr.shit, Wr.n + j);
			if(f & Y4_PF_LCOL) st4(S.lcol, tid, ldcs4(Wr.lcol, j));
			if(sc.n_lights > 1) S.lnum[tid] = __ldcs(Wr.lnum + j);
			S.flags[tid] = f; S.pix[tid] = pix; S.samp[tid] = samp;
			st4(S.pcol, tid, v_pcol); st4(S.pthr, tid, v_pthr); st4(S.scontrib, tid, v_sc0);
			if(!closing)
			{
				st4(S.lsamp, tid, v_lsamp); st4(S.thr, tid, v_thr); *reinterpret_cast<float4 *>(S.hit + tid) = v_hit;
				reinterpret_cast<float4 *>(S.ray + tid)[0] = v_r0; reinterpret_cast<float4 *>(S.ray + tid)[1] = v_r1;
			}
			push = y4_stage_path<SPEC>(sc, S, ps, tid, a.depth, a.sub, a.k_prev);
		}
		// queue appends: one 64-bit atomic per warp reserves the path range (low word) and the shadow-ray range (high word)
		const uint32_t m0 = __ballot_sync(0xffffffffu, push & Y4_PUSH_S0), m1 = __ballot_sync(0xffffffffu, push & Y4_PUSH_S1);
		const uint32_t mp = __ballot_sync(0xffffffffu, push & Y4_PUSH_PATH);
		uint32_t at_sh = 0u, at_p = 0u;
		if(m0 | m1 | mp)
		{
			if(lane == 0)
			{
				const uint32_t n_sh = (uint32_t) (__popc(m0) + __popc(m1)), n_p = (uint32_t) __popc(mp);
				const unsigned long long old = atomicAdd(reinterpret_cast<unsigned long long *>(Q.out_count), ((unsigned long long) n_sh << 32) | n_p);
				at_p = (uint32_t) old; at_sh = (uint32_t) (old >> 32);
			}
			at_sh = __shfl_sync(0xffffffffu, at_sh, 0);
			at_p = __shfl_sync(0xffffffffu, at_p, 0);
		}
		if(push & Y4_PUSH_PATH)
		{
			// copy-out to the front of the other path wave
			const uint32_t o = at_p + __popc(mp & lt_mask);
			const uint32_t f = S.flags[tid];
			__stcs(Ww.flags + o, f); __stcs(Ww.pix + o, S.pix[tid]); __stcs(Ww.samp + o, S.samp[tid]); __stcs(orig_w + o, orig);
			if(sc.n_lights > 1) __stcs(Ww.lnum + o, S.lnum[tid]);
			stcs4(Ww.lsamp, o, ld4(S.lsamp, tid)); stcs4(Ww.thr, o, ld4(S.thr, tid)); stcs4(Ww.pthr, o, ld4(S.pthr, tid)); stcs4(Ww.pcol, o, ld4(S.pcol, tid));
			if(f & Y4_PF_LCOL) stcs4(Ww.lcol, o, ld4(S.lcol, tid));
			stcs_q(Ww.ray, 2u * o, reinterpret_cast<const float4 *>(S.ray + tid)[0]);
			stcs_q(Ww.ray, 2u * o + 1u, reinterpret_cast<const float4 *>(S.ray + tid)[1]);
			if(f & Y4_PF_DIRECT) stcs4(Ww.scontrib, o, ld4(S.scontrib, tid));
			if(f & Y4_PF_C1) stcs4(Ww.scontrib, Ww.n + o, ld4(S.scontrib, T + tid));
			if(push & Y4_PUSH_S0)
			{
				stcs_q(Ww.sray, 2u * o, reinterpret_cast<const float4 *>(S.sray + tid)[0]);
				stcs_q(Ww.sray, 2u * o + 1u, reinterpret_cast<const float4 *>(S.sray + tid)[1]);
				Q.sh_q[at_sh + __popc(m0 & lt_mask)] = o;
			}
			if(push & Y4_PUSH_S1)
			{
				stcs_q(Ww.sray, 2u * (Ww.n + o), reinterpret_cast<const float4 *>(S.sray + T + tid)[0]);
				stcs_q(Ww.sray, 2u * (Ww.n + o) + 1u, reinterpret_cast<const float4 *>(S.sray + T + tid)[1]);
				Q.sh_q[at_sh + __popc(m0) + __popc(m1 & lt_mask)] = Ww.n + o;
			}
		}
		else if(have) stcs4(W0.pcol, orig, ld4(S.pcol, tid));        // the path ended: its path_col goes home to the pixel-sample
	}
}

// the same record for a two-level scene: the instance that owns the primitive (binary search over first_prim), its triangle's
// object-space vertices through the reference's Matrix4 * Point3 -- the vertices the flattened scene would hold
__global__ void __launch_bounds__(kPtThreads) k_pt_tri_shade2(const Y4Inst *__restrict__ inst, uint32_t n_inst, const float *__restrict__ mesh_verts,
                                                              const uint32_t *__restrict__ mesh_flags, uint32_t n_mats, uint64_t n_tri, float4 *__restrict__ out)
{
	for(uint64_t t = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x; t < n_tri; t += (uint64_t) gridDim.x * blockDim.x)
	{
		uint32_t lo = 0u, hi = n_inst - 1u;
		while(lo < hi) { const uint32_t mid = (lo + hi + 1u) >> 1; if(inst[mid].first_prim <= (uint32_t) t) lo = mid; else hi = mid - 1u; }
		const Y4Inst &in = inst[lo];
		const uint64_t base = (uint64_t) in.vert_first + ((uint32_t) t - in.first_prim);
		const float *v = mesh_verts + 9ull * base;
		V3 a = v3(v[0], v[1], v[2]), b = v3(v[3], v[4], v[5]), c = v3(v[6], v[7], v[8]);
		if(!in.identity) { a = y4_inst_point(in.m, a); b = y4_inst_point(in.m, b); c = y4_inst_point(in.m, c); }
		const float w[9] = {a.x, a.y, a.z, b.x, b.y, b.z, c.x, c.y, c.z};
		float4 r = y4_tri_shade(w, nullptr, n_mats, 0);
		uint32_t mat = mesh_flags ? (mesh_flags[base] >> Y4_TRI_MAT_SHIFT) : 0u;
		if(mat >= n_mats) mat = 0u;
		r.w = y4_u2f(mat);
		out[t] = r;
	}
}

__global__ void __launch_bounds__(kPtThreads) k_pt_tri_shade(const float *__restrict__ verts, const uint32_t *__restrict__ tri_flags, uint32_t n_mats, uint64_t n_tri, float4 *__restrict__ out)
{
	for(uint64_t t = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x; t < n_tri; t += (uint64_t) gridDim.x * blockDim.x) out[t] = y4_tri_shade(verts, tri_flags, n_mats, t);
}

__global__ void __launch_bounds__(kPtThreads) k_film_flag_init(const float *__restrict__ weights, uint8_t *flags, size_t n)
{
	for(size_t p = blockIdx.x * (size_t) blockDim.x + threadIdx.x; p < n; p += (size_t) gridDim.x * blockDim.x) y4_film_flag_init(weights, flags, p);
}
__global__ void __launch_bounds__(kPtThreads) k_film_flag_detect(const float *__restrict__ weights, const float *__restrict__ rgba, int width, int height, y4gpu_pt_aa_noise aa, uint8_t *flags)
{
	const size_t n = (size_t) (width - 1) * (size_t) (height - 1);
	for(size_t q = blockIdx.x * (size_t) blockDim.x + threadIdx.x; q < n; q += (size_t) gridDim.x * blockDim.x)
		y4_film_flag_detect(weights, rgba, width, height, aa, flags, (int) (q % (size_t) (width - 1)), (int) (q / (size_t) (width - 1)));
}

__global__ void __launch_bounds__(kPtThreads) k_pt_splat(const __grid_constant__ Y4PtScene sc, Y4PtWave W, Y4PtPass ps, float *film_w, float *film_rgba)
{
	for(uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < W.n; i += gridDim.x * blockDim.x) y4_pt_finish_splat(sc, W, ps, i, film_w, film_rgba);
}

// register budget of the stage kernels: MINB resident CTAs per SM (2: no cap, 3: 80 registers, 4: 64 registers)
template<int STAGE>
void launch_stage(int minb, int grid, cudaStream_t st, const Y4PtScene &sc, const Y4PtWave &W, const Y4PtPass &ps, const StageArgs &a, const StageQueues &Q)
{
	// second instantiation for scenes without specular materials: see y4_vertex
	if(sc.has_specular || STAGE == kStLevelFinish || STAGE >= kStPrimary3)
	{
		// the general variant carries more live state: its own register budget (env Y4GPU_PT_GENERAL_MIN_BLOCKS, default = the plain one's)
		static const int general_minb = getenv("Y4GPU_PT_GENERAL_MIN_BLOCKS") ? atoi(getenv("Y4GPU_PT_GENERAL_MIN_BLOCKS")) : 0;
		if(general_minb >= 2) minb = general_minb;
		if(minb >= 4) k_pt_stage<STAGE, 4, true><<<grid, kPtThreads, 0, st>>>(sc, W, ps, a, Q);
		else if(minb == 3) k_pt_stage<STAGE, 3, true><<<grid, kPtThreads, 0, st>>>(sc, W, ps, a, Q);
		else k_pt_stage<STAGE, 2, true><<<grid, kPtThreads, 0, st>>>(sc, W, ps, a, Q);
	}
	else
	{
		constexpr int S = (STAGE == kStLevelFinish || STAGE >= kStPrimary3) ? kStPath : STAGE;      // (never launched without specular materials)
		if(minb >= 4) k_pt_stage<S, 4, false><<<grid, kPtThreads, 0, st>>>(sc, W, ps, a, Q);
		else if(minb == 3) k_pt_stage<S, 3, false><<<grid, kPtThreads, 0, st>>>(sc, W, ps, a, Q);
		else k_pt_stage<S, 2, false><<<grid, kPtThreads, 0, st>>>(sc, W, ps, a, Q);
	}
}

inline int pt_grid(const y4gpu_ctx *ctx, uint32_t n)
{
	const uint32_t want = (n + kPtThreads - 1) / kPtThreads, full = (uint32_t) ctx->sm_count * 8u;
	return (int) (want < full ? (want ? want : 1u) : full);
}

void pt_free_wave(Y4Pt::Lane &L)
{
	cudaFree(L.mem); L.mem = nullptr; L.cap = 0;
	cudaFree(L.cmem); L.cmem = nullptr; L.ccap = 0;
	cudaFree(L.tmem); L.tmem = nullptr; L.tcap = 0; L.tdepth = 0;
}

// Stack of the ray-tree mode: `depth` entries per pixel-sample slot (a depth-first walk of a binary tree with `depth` levels
// never has more pending siblings than that).
int pt_alloc_tree(y4gpu_ctx *ctx, Y4Pt::Lane &L, uint32_t cap, int depth)
{
	if(cap > L.tcap || depth > L.tdepth)
	{
		cudaFree(L.tmem); L.tmem = nullptr; L.tcap = 0; L.tdepth = 0;
		Y4_CUDA(ctx, cudaMalloc(&L.tmem, (size_t) cap * depth * 48 + (size_t) cap * 4 + 256));
		L.tcap = cap; L.tdepth = depth;
	}
	char *p = (char *) L.tmem;
	L.wave.stk_ray = (Y4Ray *) p; p += (size_t) L.tcap * L.tdepth * 32;
	L.wave.stk_sw = (float *) p; p += (size_t) L.tcap * L.tdepth * 16;
	L.wave.stk_n = (uint32_t *) p;
	return Y4GPU_OK;
}

// The two dense path waves of the compacting bounce loop: only the fields the bounce loop moves (308 B per slot and wave).
int pt_alloc_compact(y4gpu_ctx *ctx, Y4Pt::Lane &L, uint32_t cap)
{
	if(cap <= L.ccap) return Y4GPU_OK;
	cudaFree(L.cmem); L.cmem = nullptr; L.ccap = 0;
	const size_t n = cap;
	const size_t per_wave = n * (5 * 4 /*flags pix samp lnum orig*/ + 7 * 16 /*lsamp thr pthr pcol lcol acc_a acc_b*/ + 32 + 16 /*ray hit*/ + 2 * 32 + 2 * 16 + 2 * 16 /*shadow*/) + 512;
	Y4_CUDA(ctx, cudaMalloc(&L.cmem, 2 * per_wave));
	char *p = (char *) L.cmem;
	auto take = [&](size_t b) { char *r = p; p += (b + 15) & ~(size_t) 15; return (void *) r; };
	for(int k = 0; k < 2; ++k)
	{
		Y4PtWave &W = L.cw[k];
		W = Y4PtWave{};
		W.lsamp = (float *) take(n * 16); W.thr = (float *) take(n * 16); W.pthr = (float *) take(n * 16); W.pcol = (float *) take(n * 16);
		W.lcol = (float *) take(n * 16); W.acc_a = (float *) take(n * 16); W.acc_b = (float *) take(n * 16);
		W.ray = (Y4Ray *) take(n * 32); W.ray_next = W.ray; W.hit = (Y4Hit *) take(n * 16);
		W.sray = (Y4Ray *) take(2 * n * 32); W.shit = (Y4Hit *) take(2 * n * 16); W.scontrib = (float *) take(2 * n * 16);
		W.flags = (uint32_t *) take(n * 4); W.pix = (uint32_t *) take(n * 4); W.samp = (uint32_t *) take(n * 4); W.lnum = (uint32_t *) take(n * 4);
		L.corig[k] = (uint32_t *) take(n * 4);
	}
	L.ccap = cap;
	return Y4GPU_OK;
}

// One allocation carved into the SoA arrays of the wave (all 16-byte aligned).
int pt_alloc_wave(y4gpu_ctx *ctx, Y4Pt::Lane &L, uint32_t cap)
{
	if(cap <= L.cap) return Y4GPU_OK;
	pt_free_wave(L);
	const size_t n = cap;
	const size_t bytes = n * (4 * 4 /*pix samp flags lnum*/ + 6 * 16 /*col pcol thr lcol acc_a acc_b*/ + 16 /*pthr*/ + 2 * 32 + 16 + 16 /*ray x2 hit hit0*/ + 2 * 32 + 2 * 16 + 2 * 16 /*shadow*/ + 32 + 16 + 16 + 8 /*ray0 sw acc q_spec*/
	                          + 16 /*lsamp*/ + 5 * 4 /*queues*/) + 512;
	Y4_CUDA(ctx, cudaMalloc(&L.mem, bytes));
	char *p = (char *) L.mem;
	auto take = [&](size_t b) { char *r = p; p += (b + 15) & ~(size_t) 15; return (void *) r; };
	Y4PtWave &W = L.wave;
	W.col = (float *) take(n * 16); W.pcol = (float *) take(n * 16); W.thr = (float *) take(n * 16); W.lcol = (float *) take(n * 16);
	W.acc_a = (float *) take(n * 16); W.acc_b = (float *) take(n * 16); W.lsamp = (float *) take(n * 16);
	W.ray = (Y4Ray *) take(n * 32); W.ray_next = (Y4Ray *) take(n * 32); W.pthr = (float *) take(n * 16); W.hit = (Y4Hit *) take(n * 16); W.hit0 = (Y4Hit *) take(n * 16);
	W.sray = (Y4Ray *) take(2 * n * 32); W.shit = (Y4Hit *) take(2 * n * 16); W.scontrib = (float *) take(2 * n * 16);
	W.ray0 = (Y4Ray *) take(n * 32); W.sw = (float *) take(n * 16); W.acc = (float *) take(n * 16);
	L.q_spec[0] = (uint32_t *) take(n * 4); L.q_spec[1] = (uint32_t *) take(n * 4);
	W.pix = (uint32_t *) take(n * 4); W.samp = (uint32_t *) take(n * 4); W.flags = (uint32_t *) take(n * 4); W.lnum = (uint32_t *) take(n * 4);
	L.q_hit0 = (uint32_t *) take(n * 4); L.q_path[0] = (uint32_t *) take(n * 4); L.q_path[1] = (uint32_t *) take(n * 4); L.q_shadow = (uint32_t *) take(2 * n * 4);
	L.cap = cap;
	return Y4GPU_OK;
}

} // namespace

static bool pt_is_device_ptr(const void *p)
{
	cudaPointerAttributes a;
	if(cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
	return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// geometry or BVH changed: the per-triangle shading arrays and the scene description are stale until the next y4gpu_pt_set_scene
void y4_pt_invalidate(y4gpu_ctx *ctx) { if(ctx->pt) ctx->pt->scene_set = false; }

extern "C" {

void y4gpu_pt_release(y4gpu_ctx *ctx)
{
	Y4Pt *pt = ctx->pt;
	if(!pt) return;
	for(Y4Pt::Lane &L : pt->lanes)
	{
		pt_free_wave(L);
		if(L.done) cudaEventDestroy(L.done);
	}
	for(int l = 1; l < kMaxLanes; ++l) if(pt->lanes[l].stream) cudaStreamDestroy(pt->lanes[l].stream);
	if(pt->ev_fork) cudaEventDestroy(pt->ev_fork);
	for(char *b : pt->light_blobs) cudaFree(b);
	pt->light_blobs.clear();
	cudaFree(pt->d_bg_sun); cudaFree(pt->d_bg_dark);
	cudaFree(pt->d_mat_transp); cudaFree(pt->d_lights); cudaFree(pt->d_mats); cudaFree(pt->d_normals); cudaFree(pt->d_tri_shade);
	pt_free_textures(pt);
	cudaFree(pt->d_sigma); cudaFree(pt->d_offset); cudaFree(pt->d_prime); cudaFree(pt->d_inv_prime); cudaFree(pt->d_sigma_d); cudaFree(pt->d_magic);
	cudaFree(pt->own_film_w); cudaFree(pt->own_film_rgba); cudaFree(pt->d_pixel_list);
	cudaFree(pt->d_counters); cudaFreeHost(pt->h_counters);
	cudaFree(pt->d_aa_flags); cudaFree(pt->d_aa_list); cudaFree(pt->d_aa_count); cudaFree(pt->d_aa_tmp);
	if(pt->ev0) cudaEventDestroy(pt->ev0);
	if(pt->ev1) cudaEventDestroy(pt->ev1);
	for(cudaEvent_t e : pt->prof_events) cudaEventDestroy(e);
	delete pt;
	ctx->pt = nullptr;
}

int y4gpu_pt_set_textures(y4gpu_ctx *ctx, const y4gpu_pt_texture *textures, uint32_t n_tex, const y4gpu_pt_layer *layers, uint32_t n_layers,
                          const float *tri_orco, const float *tri_uv)
{
	if(!ctx || (n_tex && !textures) || (n_layers && !layers)) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	if(!ctx->built) { ctx->err = "y4gpu_pt_set_textures before y4gpu_build"; return Y4GPU_ERR_ARG; }
	for(uint32_t i = 0; i < n_tex; ++i)
	{
		const y4gpu_pt_texture &t = textures[i];
		if(!t.texels || t.width < 1 || t.height < 1) { ctx->err = "texture without texels"; return Y4GPU_ERR_ARG; }
		if(t.clip < Y4GPU_CLIP_EXTEND || t.clip > Y4GPU_CLIP_CHECKER || (t.interpolate != Y4GPU_INTERP_NONE && t.interpolate != Y4GPU_INTERP_BILINEAR))
		{ ctx->err = "unsupported texture clipping / interpolation"; return Y4GPU_ERR_UNSUPPORTED; }
	}
	for(uint32_t i = 0; i < n_layers; ++i)
	{
		const y4gpu_pt_layer &l = layers[i];
		if(l.texture < 0 || (uint32_t) l.texture >= n_tex) { ctx->err = "layer refers to a texture that does not exist"; return Y4GPU_ERR_ARG; }
		if(l.texco < Y4GPU_TEXCO_GLOBAL || l.texco > Y4GPU_TEXCO_ORCO || l.mapping < Y4GPU_PROJ_PLAIN || l.mapping > Y4GPU_PROJ_CUBE || l.blend_mode < Y4GPU_BLEND_MIX || l.blend_mode > Y4GPU_BLEND_LIGHT)
		{ ctx->err = "unsupported texture_mapper / layer mode"; return Y4GPU_ERR_UNSUPPORTED; }
	}
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	if(!ctx->pt) ctx->pt = new Y4Pt;
	Y4Pt *pt = ctx->pt;
	cudaStream_t st = ctx->stream;
	Y4_CUDA(ctx, cudaStreamSynchronize(st));
	pt_free_textures(pt);
	pt->scene_set = false;                              // materials must be set again: they index the layer table
	std::vector<Y4PtTexture> ht(n_tex);
	for(uint32_t i = 0; i < n_tex; ++i)
	{
		ht[i] = derive_texture(textures[i]);
		const size_t bytes = (size_t) textures[i].width * textures[i].height * sizeof(float4);
		float4 *d = nullptr;
		Y4_CUDA(ctx, cudaMalloc(&d, bytes));
		pt->d_texels.push_back(d);
		Y4_CUDA(ctx, cudaMemcpyAsync(d, textures[i].texels, bytes, cudaMemcpyDefault, st));
		ht[i].texels = d;
	}
	std::vector<Y4PtLayer> hl(n_layers);
	for(uint32_t i = 0; i < n_layers; ++i) hl[i] = derive_layer(layers[i]);
	if(n_tex)
	{
		Y4_CUDA(ctx, cudaMalloc(&pt->d_textures, n_tex * sizeof(Y4PtTexture)));
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_textures, ht.data(), n_tex * sizeof(Y4PtTexture), cudaMemcpyHostToDevice, st));
	}
	if(n_layers)
	{
		Y4_CUDA(ctx, cudaMalloc(&pt->d_layers, n_layers * sizeof(Y4PtLayer)));
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_layers, hl.data(), n_layers * sizeof(Y4PtLayer), cudaMemcpyHostToDevice, st));
	}
	if(tri_orco && ctx->n_tri && n_layers)
	{
		Y4_CUDA(ctx, cudaMalloc(&pt->d_tri_orco, ctx->n_tri * 9 * sizeof(float)));
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_tri_orco, tri_orco, ctx->n_tri * 9 * sizeof(float), cudaMemcpyDefault, st));
	}
	if(tri_uv && ctx->n_tri && n_layers)
	{
		Y4_CUDA(ctx, cudaMalloc(&pt->d_tri_uv, ctx->n_tri * 6 * sizeof(float)));
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_tri_uv, tri_uv, ctx->n_tri * 6 * sizeof(float), cudaMemcpyDefault, st));
	}
	Y4_CUDA(ctx, cudaStreamSynchronize(st));            // the host vectors above go out of scope
	pt->n_textures = n_tex; pt->n_layers = n_layers;
	pt->h_textures = ht;
	return Y4GPU_OK;
}

int y4gpu_pt_set_scene(y4gpu_ctx *ctx, const y4gpu_pt_camera *cam, const y4gpu_pt_material *mats, uint32_t n_mat,
                       const y4gpu_pt_light *lights, uint32_t n_light, const y4gpu_pt_params *params, const float *vertex_normals)
{
	if(!ctx || !cam || !params || (n_mat && !mats) || (n_light && !lights)) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	if(!ctx->built) { ctx->err = "y4gpu_pt_set_scene before y4gpu_build"; return Y4GPU_ERR_ARG; }
	// ---- the supported set; everything else is rejected, never approximated ----
	if(n_mat == 0) { ctx->err = "no materials"; return Y4GPU_ERR_ARG; }
	if(n_light > Y4_PT_MAX_LIGHTS) { ctx->err = "more than 64 lights"; return Y4GPU_ERR_UNSUPPORTED; }
	if(ctx->pt) ctx->pt->scene_set = false;             // a failed call must not leave the previous scene usable with half-updated state
	if(params->raydepth > 19) { ctx->err = "raydepth > 19 is not supported"; return Y4GPU_ERR_UNSUPPORTED; }
	if(params->transp_shadows && params->shadow_depth > 16) { ctx->err = "shadowDepth > 16 is not supported"; return Y4GPU_ERR_UNSUPPORTED; }
	if(params->bounces < 1 || params->bounces > 11) { ctx->err = "bounces must be 1..11 (sample dimensions 4*depth+4 < 50)"; return Y4GPU_ERR_UNSUPPORTED; }
	if(params->path_samples < 0 || params->width < 1 || params->height < 1) { ctx->err = "bad path_samples / film size"; return Y4GPU_ERR_ARG; }
	for(uint32_t i = 0; i < n_mat; ++i)
		if(mats[i].type < Y4GPU_MAT_SHINYDIFFUSE || mats[i].type > Y4GPU_MAT_GLASS) { ctx->err = "unsupported material type"; return Y4GPU_ERR_UNSUPPORTED; }
	for(uint32_t i = 0; i < n_mat; ++i)
		if(mats[i].layer_count < 0 || (mats[i].layer_count > 0 && (mats[i].layer_first < 0 || !ctx->pt || (uint32_t) (mats[i].layer_first + mats[i].layer_count) > ctx->pt->n_layers)))
		{ ctx->err = "material refers to layers that were not set (call y4gpu_pt_set_textures first)"; return Y4GPU_ERR_ARG; }
	for(uint32_t i = 0; i < n_light; ++i)
		if(lights[i].type < Y4GPU_LIGHT_POINT || lights[i].type > Y4GPU_LIGHT_PORTAL) { ctx->err = "unsupported light type"; return Y4GPU_ERR_UNSUPPORTED; }
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	if(!ctx->pt) ctx->pt = new Y4Pt;
	if(const char *e = getenv("Y4GPU_PT_MIN_BLOCKS")) ctx->pt->stage_min_blocks = atoi(e);
	if(const char *e = getenv("Y4GPU_PT_LANES")) ctx->pt->n_lanes = atoi(e) < 1 ? 1 : (atoi(e) > kMaxLanes ? kMaxLanes : atoi(e));
	if(const char *e = getenv("Y4GPU_PT_TRACE_BLOCKS")) ctx->pt->trace_blocks_shared = atoi(e);
	if(const char *e = getenv("Y4GPU_PT_COMPACT")) ctx->pt->compact = atoi(e);
	Y4Pt *pt = ctx->pt;
	if(!pt->ev0) { Y4_CUDA(ctx, cudaEventCreate(&pt->ev0)); Y4_CUDA(ctx, cudaEventCreate(&pt->ev1)); }
	cudaStream_t st = ctx->stream;
	Y4PtScene &sc = pt->h_scene;
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
		// two specular children per vertex (glass: reflection + refraction): recursiveRaytrace is a tree, walked depth first with a
		// per-pixel-sample stack (y4_stage_primary3 / y4_stage_finish3)
		if(mats[m].type == Y4GPU_MAT_GLASS) sc.has_specular = 3;
		// the same for shinydiffusemat with a mirror AND a transparency component
		if(mats[m].type == Y4GPU_MAT_SHINYDIFFUSE && mats[m].specular_reflect > 0.00001f && mats[m].transparency > 0.00001f) sc.has_specular = 3;
	}
	if(sc.has_specular == 3 && params->transp_background) { ctx->err = "transparent background together with glass / mirror + transparency materials is not supported (alpha of a ray tree)"; return Y4GPU_ERR_UNSUPPORTED; }
	for(uint32_t m = 0; m < n_mat; ++m)
		if(mats[m].type == Y4GPU_MAT_SHINYDIFFUSE && mats[m].transparency > 0.00001f)
		{
			// filtered shadows take one getTransparency colour per material: a Fresnel mirror component makes it depend on the angle,
			if(params->transp_shadows && mats[m].specular_reflect > 0.00001f && mats[m].fresnel_effect) { ctx->err = "transpShad with a transparent material that also has a Fresnel mirror component is not supported"; return Y4GPU_ERR_UNSUPPORTED; }
			// a textured diffuse colour would need the blocker's surface point
			if(params->transp_shadows && mats[m].layer_count > 0) { ctx->err = "textured transparent materials with transpShad are not supported"; return Y4GPU_ERR_UNSUPPORTED; }
		}
	// no_recursive makes a mirror hit start sub-paths AND recurse: two branches per vertex, which the linear ray-level chain does not cover
	bool chain_mats = false;
	for(uint32_t m = 0; m < n_mat; ++m)
		chain_mats = chain_mats || mats[m].type == Y4GPU_MAT_MIRROR || mats[m].type == Y4GPU_MAT_GLASS || (mats[m].type == Y4GPU_MAT_SHINYDIFFUSE && (mats[m].specular_reflect > 0.00001f || mats[m].transparency > 0.00001f));
	if(chain_mats && sc.no_recursive) { ctx->err = "no_recursive with mirror materials / specular_reflect is not supported"; return Y4GPU_ERR_UNSUPPORTED; }
	// object lights need the general kernel variant as well (not the ray levels)
	for(uint32_t l = 0; l < n_light; ++l) if((lights[l].type == Y4GPU_LIGHT_OBJECT || lights[l].type == Y4GPU_LIGHT_BACKGROUND || lights[l].type == Y4GPU_LIGHT_SPOT_SOFT || lights[l].type == Y4GPU_LIGHT_PORTAL) && !sc.has_specular) sc.has_specular = 1;
	sc.background = c3(params->background[0], params->background[1], params->background[2]);
	// textureback: the image is one of the uploaded textures; lights that evaluate the background on the host (the IBL tables)
	// get a host copy of its texels for the duration of this call
	std::vector<float4> bg_texels_host;
	Y4PtTexture bg_tex_host;
	memset(&bg_tex_host, 0, sizeof(bg_tex_host));
	if(params->background_gradient == 2)
	{
		if(params->bg_texture < 0 || (uint32_t) params->bg_texture >= pt->n_textures) { ctx->err = "textureback: bg_texture is not an uploaded texture (y4gpu_pt_set_textures comes first)"; return Y4GPU_ERR_ARG; }
		sc.bg_tex = derive_texture_background(*params, pt->d_textures + params->bg_texture);
		bool on_host = false;
		for(uint32_t l = 0; l < n_light; ++l) on_host = on_host || lights[l].type == Y4GPU_LIGHT_BACKGROUND;
		if(on_host)
		{
			bg_tex_host = pt->h_textures[params->bg_texture];
			bg_texels_host.resize((size_t) bg_tex_host.w * bg_tex_host.h);
			Y4_CUDA(ctx, cudaMemcpyAsync(bg_texels_host.data(), bg_tex_host.texels, bg_texels_host.size() * sizeof(float4), cudaMemcpyDeviceToHost, st));
			Y4_CUDA(ctx, cudaStreamSynchronize(st));
			bg_tex_host.texels = bg_texels_host.data();
		}
	}
	else if(params->background_gradient < 0 || params->background_gradient > 4) { ctx->err = "unknown background type"; return Y4GPU_ERR_ARG; }
	Y4BgDarksky dark_host;
	memset(&dark_host, 0, sizeof(dark_host));
	if(params->background_gradient == 4)
	{
		dark_host = derive_darksky(*params);
		if(!pt->d_bg_dark) Y4_CUDA(ctx, cudaMalloc(&pt->d_bg_dark, sizeof(Y4BgDarksky)));
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_bg_dark, &dark_host, sizeof(Y4BgDarksky), cudaMemcpyHostToDevice, st));
		Y4_CUDA(ctx, cudaStreamSynchronize(st));
		sc.bg_dark = pt->d_bg_dark;
	}
	// sunsky: the derived state lives in one small device block that the scene and the lights point to
	Y4BgSunsky sun_host;
	memset(&sun_host, 0, sizeof(sun_host));
	if(params->background_gradient == 3)
	{
		sun_host = derive_sunsky(*params);
		if(!pt->d_bg_sun) Y4_CUDA(ctx, cudaMalloc(&pt->d_bg_sun, sizeof(Y4BgSunsky)));
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_bg_sun, &sun_host, sizeof(Y4BgSunsky), cudaMemcpyHostToDevice, st));
		Y4_CUDA(ctx, cudaStreamSynchronize(st));
		sc.bg_sun = pt->d_bg_sun;
	}
	sc.shadow_bias = params->shadow_bias; sc.shadow_bias_auto = params->shadow_bias_auto;
	sc.ray_min_dist = params->min_raydist; sc.ray_min_dist_auto = params->min_raydist_auto;
	sc.base_sampling_offset = params->base_sampling_offset; sc.rr_seed = params->rr_seed;
	sc.aa_light_sample_multiplier = 1.f;
	sc.aa_passes = params->aa_passes;
	sc.width = params->width; sc.height = params->height; sc.cx0 = params->xstart; sc.cy0 = params->ystart;
	sc.clamp_samples = params->aa_clamp_samples;
	derive_film(*params, sc);
	sc.n_lights = (int) n_light; sc.n_mats = (int) n_mat;
	pt->max_light_samples = 1;
	pt->h_lights.resize(n_light);
	for(char *b : pt->light_blobs) cudaFree(b);
	pt->light_blobs.clear();
	for(uint32_t i = 0; i < n_light; ++i)
	{
		pt->h_lights[i] = derive_light(lights[i]);
		if(lights[i].type == Y4GPU_LIGHT_PORTAL) derive_light_background(pt->h_lights[i], *params, pt->d_textures + (params->background_gradient == 2 ? params->bg_texture : 0), pt->d_bg_sun, pt->d_bg_dark);
		if(lights[i].type == Y4GPU_LIGHT_OBJECT || lights[i].type == Y4GPU_LIGHT_PORTAL)
		{
			// the light's own copy of its mesh: vertices, geometric normals, area cdf (ObjectLight::initIs) in one device block
			const uint32_t n = lights[i].object_n_tris;
			if(!n || !lights[i].object_tris) { ctx->err = "objectlight without triangles"; return Y4GPU_ERR_ARG; }
			std::vector<float> cdf; std::vector<float4> ng;
			derive_object_light(pt->h_lights[i], lights[i].object_tris, n, cdf, ng);
			char *blob = nullptr;
			Y4_CUDA(ctx, cudaMalloc(&blob, (size_t) n * (16 + 36 + 4)));
			pt->light_blobs.push_back(blob);
			Y4_CUDA(ctx, cudaMemcpyAsync(blob, ng.data(), (size_t) n * 16, cudaMemcpyHostToDevice, st));
			Y4_CUDA(ctx, cudaMemcpyAsync(blob + (size_t) n * 16, lights[i].object_tris, (size_t) n * 36, cudaMemcpyHostToDevice, st));
			Y4_CUDA(ctx, cudaMemcpyAsync(blob + (size_t) n * 52, cdf.data(), (size_t) n * 4, cudaMemcpyHostToDevice, st));
			Y4_CUDA(ctx, cudaStreamSynchronize(st));
			pt->h_lights[i].obj_ng = reinterpret_cast<const float4 *>(blob);
			pt->h_lights[i].obj_tris = reinterpret_cast<const float *>(blob + (size_t) n * 16);
			pt->h_lights[i].obj_cdf = reinterpret_cast<const float *>(blob + (size_t) n * 52);
		}
		if(lights[i].type == Y4GPU_LIGHT_BACKGROUND)
		{
			// the tabulated background (BackgroundLight::init) in one device block: row offsets, function values, cdfs, 1 / integrals
			std::vector<int> off; std::vector<float> func, cdf, inv_int;
			derive_background_light(pt->h_lights[i], lights[i], *params, off, func, cdf, inv_int, &bg_tex_host, &sun_host, &dark_host);
			if(params->background_gradient == 2) pt->h_lights[i].bg_tex.tex = pt->d_textures + params->bg_texture;      // from here on: the device copy
			pt->h_lights[i].bg_sun = pt->d_bg_sun; pt->h_lights[i].bg_dark = pt->d_bg_dark;
			const size_t b_off = off.size() * 4, b_f = func.size() * 4, b_i = inv_int.size() * 4;
			char *blob = nullptr;
			Y4_CUDA(ctx, cudaMalloc(&blob, b_off + 2 * b_f + b_i));
			pt->light_blobs.push_back(blob);
			Y4_CUDA(ctx, cudaMemcpyAsync(blob, off.data(), b_off, cudaMemcpyHostToDevice, st));
			Y4_CUDA(ctx, cudaMemcpyAsync(blob + b_off, func.data(), b_f, cudaMemcpyHostToDevice, st));
			Y4_CUDA(ctx, cudaMemcpyAsync(blob + b_off + b_f, cdf.data(), b_f, cudaMemcpyHostToDevice, st));
			Y4_CUDA(ctx, cudaMemcpyAsync(blob + b_off + 2 * b_f, inv_int.data(), b_i, cudaMemcpyHostToDevice, st));
			Y4_CUDA(ctx, cudaStreamSynchronize(st));
			pt->h_lights[i].bg_off = reinterpret_cast<const int *>(blob);
			pt->h_lights[i].bg_func = reinterpret_cast<const float *>(blob + b_off);
			pt->h_lights[i].bg_cdf = reinterpret_cast<const float *>(blob + b_off + b_f);
			pt->h_lights[i].bg_inv_int = reinterpret_cast<const float *>(blob + b_off + 2 * b_f);
		}
		const int ns = y4_light_num_samples(sc, pt->h_lights[i]);
		if(ns > pt->max_light_samples) pt->max_light_samples = ns;
	}
	if(n_light > pt->lights_cap) { cudaFree(pt->d_lights); pt->d_lights = nullptr; pt->lights_cap = 0; }
	if(n_light)
	{
		if(!pt->d_lights) { Y4_CUDA(ctx, cudaMalloc(&pt->d_lights, n_light * sizeof(Y4PtLight))); pt->lights_cap = n_light; }
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_lights, pt->h_lights.data(), n_light * sizeof(Y4PtLight), cudaMemcpyHostToDevice, st));
	}
	std::vector<Y4PtMaterial> hm(n_mat);
	for(uint32_t i = 0; i < n_mat; ++i) hm[i] = derive_material(mats[i]);
	// ShinyDiffuseMaterial::getTransparency (:442-465) per material for filtered shadows: accum = 1; accum *= transparency * accum; accum * tcol
	sc.tshadow_filter = 0;
	std::vector<float> mat_transp(3 * (size_t) n_mat, 0.f);
	for(uint32_t i = 0; i < n_mat; ++i)
		if(hm[i].is_transparent)
		{
			float accum = 1.f;
			if(hm[i].is_mirror) accum = 1.f - 1.f * hm[i].mirror_strength;      // 1 - kr * mirror_strength, kr = 1 without the Fresnel effect (refused with it below)
			accum *= hm[i].transparency_strength * accum;
			const C3 tc = y4_shiny_tcol(hm[i]);
			mat_transp[3 * i] = accum * tc.r; mat_transp[3 * i + 1] = accum * tc.g; mat_transp[3 * i + 2] = accum * tc.b;
			if(params->transp_shadows) sc.tshadow_filter = 1;
		}
	if(sc.tshadow_filter)
	{
		if(n_mat > pt->mat_transp_cap) { cudaFree(pt->d_mat_transp); pt->d_mat_transp = nullptr; pt->mat_transp_cap = 0; }
		if(!pt->d_mat_transp) { Y4_CUDA(ctx, cudaMalloc(&pt->d_mat_transp, 3 * (size_t) n_mat * sizeof(float))); pt->mat_transp_cap = n_mat; }
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_mat_transp, mat_transp.data(), 3 * (size_t) n_mat * sizeof(float), cudaMemcpyHostToDevice, st));
		Y4_CUDA(ctx, cudaStreamSynchronize(st));
	}
	if(n_mat > pt->mats_cap) { cudaFree(pt->d_mats); pt->d_mats = nullptr; pt->mats_cap = 0; }
	if(!pt->d_mats) { Y4_CUDA(ctx, cudaMalloc(&pt->d_mats, n_mat * sizeof(Y4PtMaterial))); pt->mats_cap = n_mat; }
	Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_mats, hm.data(), n_mat * sizeof(Y4PtMaterial), cudaMemcpyHostToDevice, st));
	if(!vertex_normals || ctx->n_tri > pt->normals_cap) { cudaFree(pt->d_normals); pt->d_normals = nullptr; pt->normals_cap = 0; }
	if(vertex_normals && ctx->n_tri)
	{
		if(!pt->d_normals) { Y4_CUDA(ctx, cudaMalloc(&pt->d_normals, ctx->n_tri * 9 * sizeof(float))); pt->normals_cap = ctx->n_tri; }
		Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_normals, vertex_normals, ctx->n_tri * 9 * sizeof(float), cudaMemcpyDefault, st));
	}
	if(!pt->d_sigma)
	{
		const Y4FaureHost fh = y4_faure_build();
		Y4_CUDA(ctx, cudaMalloc(&pt->d_sigma, fh.sigma.size())); Y4_CUDA(ctx, cudaMalloc(&pt->d_offset, sizeof(fh.offset)));
		Y4_CUDA(ctx, cudaMalloc(&pt->d_prime, sizeof(fh.prime))); Y4_CUDA(ctx, cudaMalloc(&pt->d_inv_prime, sizeof(fh.inv_prime)));
		Y4_CUDA(ctx, cudaMemcpy(pt->d_sigma, fh.sigma.data(), fh.sigma.size(), cudaMemcpyHostToDevice));
		Y4_CUDA(ctx, cudaMemcpy(pt->d_offset, fh.offset, sizeof(fh.offset), cudaMemcpyHostToDevice));
		Y4_CUDA(ctx, cudaMemcpy(pt->d_prime, fh.prime, sizeof(fh.prime), cudaMemcpyHostToDevice));
		Y4_CUDA(ctx, cudaMemcpy(pt->d_inv_prime, fh.inv_prime, sizeof(fh.inv_prime), cudaMemcpyHostToDevice));
		Y4_CUDA(ctx, cudaMalloc(&pt->d_sigma_d, fh.sigma_d.size() * sizeof(double))); Y4_CUDA(ctx, cudaMalloc(&pt->d_magic, sizeof(fh.magic)));
		Y4_CUDA(ctx, cudaMemcpy(pt->d_sigma_d, fh.sigma_d.data(), fh.sigma_d.size() * sizeof(double), cudaMemcpyHostToDevice));
		Y4_CUDA(ctx, cudaMemcpy(pt->d_magic, fh.magic, sizeof(fh.magic), cudaMemcpyHostToDevice));
	}
	if(ctx->n_tri > pt->tri_cap) { cudaFree(pt->d_tri_shade); pt->d_tri_shade = nullptr; pt->tri_cap = 0; }
	if(ctx->n_tri)
	{
		if(!pt->d_tri_shade) { Y4_CUDA(ctx, cudaMalloc(&pt->d_tri_shade, ctx->n_tri * sizeof(float4))); pt->tri_cap = ctx->n_tri; }
		const uint64_t want = (ctx->n_tri + kPtThreads - 1) / kPtThreads, full = (uint64_t) ctx->sm_count * 8u;
		if(ctx->two_level) k_pt_tri_shade2<<<(int) (want < full ? want : full), kPtThreads, 0, st>>>(ctx->d_inst, (uint32_t) ctx->h_inst.size(), ctx->d_mesh_verts, ctx->d_mesh_flags, n_mat, ctx->n_tri, pt->d_tri_shade);
		else k_pt_tri_shade<<<(int) (want < full ? want : full), kPtThreads, 0, st>>>(ctx->d_verts, ctx->d_flags, n_mat, ctx->n_tri, pt->d_tri_shade);
		Y4_CUDA(ctx, cudaGetLastError());
	}
	sc.lights = pt->d_lights; sc.mats = pt->d_mats; sc.tri_shade = pt->d_tri_shade; sc.normals = pt->d_normals;
	sc.textures = pt->d_textures; sc.layers = pt->d_layers; sc.tri_orco = pt->d_tri_orco; sc.tri_uv = pt->d_tri_uv;
	sc.lds.sigma = pt->d_sigma; sc.lds.offset = pt->d_offset; sc.lds.prime = pt->d_prime; sc.lds.inv_prime = pt->d_inv_prime; sc.lds.sigma_d = pt->d_sigma_d; sc.lds.magic = pt->d_magic;
	// film (internal unless the caller installs buffers)
	const size_t npx = (size_t) sc.width * sc.height;
	if(npx != pt->film_cap)
	{
		cudaFree(pt->own_film_w); cudaFree(pt->own_film_rgba); pt->own_film_w = pt->own_film_rgba = nullptr; pt->film_cap = 0;
		Y4_CUDA(ctx, cudaMalloc(&pt->own_film_w, npx * sizeof(float)));
		Y4_CUDA(ctx, cudaMalloc(&pt->own_film_rgba, npx * 4 * sizeof(float)));
		pt->film_cap = npx;
	}
	Y4_CUDA(ctx, cudaMemsetAsync(pt->own_film_w, 0, npx * sizeof(float), st));
	Y4_CUDA(ctx, cudaMemsetAsync(pt->own_film_rgba, 0, npx * 4 * sizeof(float), st));
	pt->d_film_w = pt->own_film_w; pt->d_film_rgba = pt->own_film_rgba;
	Y4_CUDA(ctx, cudaStreamSynchronize(st));
	pt->params = *params;
	pt->scene_set = true;
	return Y4GPU_OK;
}

int y4gpu_pt_set_film_buffers(y4gpu_ctx *ctx, float *d_weights, float *d_rgba)
{
	if(!ctx) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4Pt *pt = ctx->pt;
	if(!pt || !pt->scene_set) { ctx->err = "y4gpu_pt_set_film_buffers before y4gpu_pt_set_scene"; return Y4GPU_ERR_ARG; }
	if(!d_weights != !d_rgba) return Y4GPU_ERR_ARG;
	if(d_weights && (!pt_is_device_ptr(d_weights) || !pt_is_device_ptr(d_rgba))) { ctx->err = "film buffers must be device memory"; return Y4GPU_ERR_ARG; }
	pt->d_film_w = d_weights ? d_weights : pt->own_film_w;
	pt->d_film_rgba = d_rgba ? d_rgba : pt->own_film_rgba;
	return Y4GPU_OK;
}

int y4gpu_pt_clear_film(y4gpu_ctx *ctx)
{
	if(!ctx) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4Pt *pt = ctx->pt;
	if(!pt || !pt->scene_set) { ctx->err = "y4gpu_pt_clear_film before y4gpu_pt_set_scene"; return Y4GPU_ERR_ARG; }
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	const size_t npx = (size_t) pt->h_scene.width * pt->h_scene.height;
	Y4_CUDA(ctx, cudaMemsetAsync(pt->d_film_w, 0, npx * sizeof(float), ctx->stream));
	Y4_CUDA(ctx, cudaMemsetAsync(pt->d_film_rgba, 0, npx * 4 * sizeof(float), ctx->stream));
	return Y4GPU_OK;
}

int y4gpu_libm_eval(y4gpu_ctx *ctx, int fn, const float *a, const float *b, float *out, uint64_t n)
{
	if(!ctx || !a || !out || fn < 0 || fn > 3 || (fn == 3 && !b)) return Y4GPU_ERR_ARG;
	if(!n) return Y4GPU_OK;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	float *d = nullptr;
	Y4_CUDA(ctx, cudaMalloc((void **) &d, 3 * n * sizeof(float)));
	cudaError_t e = cudaMemcpyAsync(d, a, n * sizeof(float), cudaMemcpyDefault, ctx->stream);
	if(e == cudaSuccess && fn == 3) e = cudaMemcpyAsync(d + n, b, n * sizeof(float), cudaMemcpyDefault, ctx->stream);
	if(e == cudaSuccess)
	{
		k_libm_eval<<<(unsigned) ((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8), 256, 0, ctx->stream>>>(fn, d, d + n, d + 2 * n, n);
		e = cudaGetLastError();
	}
	if(e == cudaSuccess) e = cudaMemcpyAsync(out, d + 2 * n, n * sizeof(float), cudaMemcpyDefault, ctx->stream);
	if(e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
	cudaFree(d);
	if(e != cudaSuccess) { ctx->err = std::string("y4gpu_libm_eval: ") + cudaGetErrorString(e); return Y4GPU_ERR_CUDA; }
	return Y4GPU_OK;
}

int y4gpu_pt_read_film(y4gpu_ctx *ctx, float *weights, float *rgba)
{
	if(!ctx || !weights || !rgba) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4Pt *pt = ctx->pt;
	if(!pt || !pt->scene_set) { ctx->err = "y4gpu_pt_read_film before y4gpu_pt_set_scene"; return Y4GPU_ERR_ARG; }
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	const size_t npx = (size_t) pt->h_scene.width * pt->h_scene.height;
	Y4_CUDA(ctx, cudaMemcpyAsync(weights, pt->d_film_w, npx * sizeof(float), cudaMemcpyDefault, ctx->stream));
	Y4_CUDA(ctx, cudaMemcpyAsync(rgba, pt->d_film_rgba, npx * 4 * sizeof(float), cudaMemcpyDefault, ctx->stream));
	Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return Y4GPU_OK;
}

int y4gpu_pt_write_film(y4gpu_ctx *ctx, const float *weights, const float *rgba)
{
	if(!ctx || !weights || !rgba) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4Pt *pt = ctx->pt;
	if(!pt || !pt->scene_set) { ctx->err = "y4gpu_pt_write_film before y4gpu_pt_set_scene"; return Y4GPU_ERR_ARG; }
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	const size_t npx = (size_t) pt->h_scene.width * pt->h_scene.height;
	Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_film_w, weights, npx * sizeof(float), cudaMemcpyDefault, ctx->stream));
	Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_film_rgba, rgba, npx * 4 * sizeof(float), cudaMemcpyDefault, ctx->stream));
	Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return Y4GPU_OK;
}

int y4gpu_pt_next_pass(y4gpu_ctx *ctx, const y4gpu_pt_aa_noise *aa, const float *d_weights, const float *d_rgba,
                       const uint32_t **d_pixel_list, uint64_t *n_resample)
{
	if(!ctx || !aa || !d_pixel_list || !n_resample || (!d_weights != !d_rgba)) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4Pt *pt = ctx->pt;
	if(!pt || !pt->scene_set) { ctx->err = "y4gpu_pt_next_pass before y4gpu_pt_set_scene"; return Y4GPU_ERR_ARG; }
	if(aa->dark_detection_type < 0 || aa->dark_detection_type > 2) { ctx->err = "unknown AA_dark_detection_type"; return Y4GPU_ERR_ARG; }
	if(d_weights && (!pt_is_device_ptr(d_weights) || !pt_is_device_ptr(d_rgba))) { ctx->err = "film buffers must be device memory"; return Y4GPU_ERR_ARG; }
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	cudaStream_t st = ctx->stream;
	const int width = pt->h_scene.width, height = pt->h_scene.height;
	const size_t npx = (size_t) width * height;
	const float *w = d_weights ? d_weights : pt->d_film_w, *rgba = d_rgba ? d_rgba : pt->d_film_rgba;
	if(npx > pt->aa_cap)
	{
		cudaFree(pt->d_aa_flags); cudaFree(pt->d_aa_list); pt->d_aa_flags = nullptr; pt->d_aa_list = nullptr; pt->aa_cap = 0;
		Y4_CUDA(ctx, cudaMalloc(&pt->d_aa_flags, npx));
		Y4_CUDA(ctx, cudaMalloc(&pt->d_aa_list, npx * sizeof(uint32_t)));
		pt->aa_cap = npx;
	}
	if(!pt->d_aa_count) Y4_CUDA(ctx, cudaMalloc(&pt->d_aa_count, sizeof(uint64_t)));
	const int grid = pt_grid(ctx, (uint32_t) (npx < 0xffffffffull ? npx : 0xffffffffull));
	k_film_flag_init<<<grid, kPtThreads, 0, st>>>(w, pt->d_aa_flags, npx);
	if(width > 1 && height > 1) k_film_flag_detect<<<grid, kPtThreads, 0, st>>>(w, rgba, width, height, *aa, pt->d_aa_flags);
	Y4_CUDA(ctx, cudaGetLastError());
	// raster-order compaction of the flagged pixel indices
	cub::CountingInputIterator<uint32_t> idx(0u);
	size_t need = 0;
	Y4_CUDA(ctx, cub::DeviceSelect::Flagged(nullptr, need, idx, pt->d_aa_flags, pt->d_aa_list, pt->d_aa_count, (int) npx, st));
	if(need > pt->aa_tmp_bytes)
	{
		cudaFree(pt->d_aa_tmp); pt->d_aa_tmp = nullptr; pt->aa_tmp_bytes = 0;
		Y4_CUDA(ctx, cudaMalloc(&pt->d_aa_tmp, need));
		pt->aa_tmp_bytes = need;
	}
	need = pt->aa_tmp_bytes;
	Y4_CUDA(ctx, cub::DeviceSelect::Flagged(pt->d_aa_tmp, need, idx, pt->d_aa_flags, pt->d_aa_list, pt->d_aa_count, (int) npx, st));
	uint64_t count = 0;
	Y4_CUDA(ctx, cudaMemcpyAsync(&count, pt->d_aa_count, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
	Y4_CUDA(ctx, cudaStreamSynchronize(st));
	ctx->stats.kernel_launches += 3;
	*d_pixel_list = pt->d_aa_list;
	*n_resample = count;
	return Y4GPU_OK;
}

int y4gpu_pt_get_stats(const y4gpu_ctx *ctx, y4gpu_pt_stats *out)
{
	if(!ctx || !out || !ctx->pt) return Y4GPU_ERR_ARG;
	Y4Pt *pt = ctx->pt;
	float ms = 0.f;
	if(pt->ev1 && cudaEventSynchronize(pt->ev1) == cudaSuccess)
	{
		if(cudaEventElapsedTime(&ms, pt->ev0, pt->ev1) == cudaSuccess) pt->stats.pass_ms = ms;
		if(!pt->counters_read)
		{
			// queue lengths of the pass (copied to pinned memory before ev1): rays actually traced
			uint64_t sum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
			for(const auto &ck : pt->counter_kinds)
			{
				const unsigned long long v = pt->h_counters[ck.first & 0x7fffffffu];
				sum[ck.second & 7] += (ck.first >> 31) ? (uint32_t) (v >> 32) : (uint32_t) v;
			}
			pt->stats.rays_closest = pt->dense_closest + sum[0] + sum[4]; pt->stats.rays_shadow = sum[1];
			pt->stats.camera_hits = sum[2]; pt->stats.path_entries = sum[0] + sum[3];
			if(pt->params.profile)
			{
				pt->stats.nodes_closest = pt->h_counters[0]; pt->stats.tris_closest = pt->h_counters[1]; pt->stats.live_closest = pt->h_counters[2];
				pt->stats.nodes_shadow = pt->h_counters[3]; pt->stats.tris_shadow = pt->h_counters[4]; pt->stats.live_shadow = pt->h_counters[5];
			}
			pt->counters_read = true;
		}
	}
	else cudaGetLastError();
	*out = pt->stats;
	return Y4GPU_OK;
}

} // extern "C"

// ---- one AA pass (TiledIntegrator::renderPass) -------------------------------------------------------
extern "C" int y4gpu_pt_render_pass(y4gpu_ctx *ctx, int n_samples, int offset, int aa_pass_number,
                                    const uint32_t *pixel_list, uint64_t n_pixels, int shard_index, int shard_count)
{
	if(!ctx || n_samples < 0 || shard_count < 1 || shard_index < 0 || shard_index >= shard_count) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4Pt *pt = ctx->pt;
	if(!pt || !pt->scene_set) { ctx->err = "y4gpu_pt_render_pass before y4gpu_pt_set_scene"; return Y4GPU_ERR_ARG; }
	if(!ctx->built) { ctx->err = "y4gpu_pt_render_pass before y4gpu_build"; return Y4GPU_ERR_ARG; }
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	cudaStream_t st = ctx->stream;
	Y4PtScene &sc = pt->h_scene;
	const uint64_t launches0 = ctx->stats.kernel_launches;
	uint64_t own_launches = 0;
	// per-pass light sample multiplier (integrator_tiled.cc:189-191: multiplied once per additional pass)
	float mult = 1.f;
	for(int p = 0; p < aa_pass_number; ++p) mult *= pt->params.aa_light_sample_multiplier_factor > 0.f ? pt->params.aa_light_sample_multiplier_factor : 1.f;
	if(mult != sc.aa_light_sample_multiplier)
	{
		sc.aa_light_sample_multiplier = mult;
		pt->max_light_samples = 1;
		for(int i = 0; i < sc.n_lights; ++i) { const int ns = y4_light_num_samples(sc, pt->h_lights[i]); if(ns > pt->max_light_samples) pt->max_light_samples = ns; }
	}
	const uint64_t n_pix = pixel_list ? n_pixels : (uint64_t) sc.width * sc.height;
	const uint32_t ns_local = shard_index < n_samples ? (uint32_t) ((n_samples - shard_index + shard_count - 1) / shard_count) : 0u;
	const uint64_t total = n_pix * ns_local;
	Y4_CUDA(ctx, cudaEventRecord(pt->ev0, st));
	const uint32_t *d_list = nullptr;
	if(pixel_list && n_pixels)
	{
		if(pt_is_device_ptr(pixel_list)) d_list = pixel_list;
		else
		{
			if(n_pixels > pt->pixel_list_cap)
			{
				cudaFree(pt->d_pixel_list); pt->d_pixel_list = nullptr; pt->pixel_list_cap = 0;
				Y4_CUDA(ctx, cudaMalloc(&pt->d_pixel_list, n_pixels * sizeof(uint32_t)));
				pt->pixel_list_cap = n_pixels;
			}
			Y4_CUDA(ctx, cudaMemcpyAsync(pt->d_pixel_list, pixel_list, n_pixels * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
			d_list = pt->d_pixel_list;
		}
	}
	uint32_t wave_size = pt->params.wave_size ? pt->params.wave_size : (1u << 23);      // 8 Mi pixel-samples: ~1 KB of state per slot, two wavefronts in flight = 17 GB of the 180 (profiles/r2_v10_sweep_wave.log)
	if(total < wave_size) wave_size = (uint32_t) (total ? total : 1);
	const bool prof = pt->params.profile != 0;
	// lanes: consecutive waves alternate between two streams (profile mode times kernels one by one: single lane)
	const uint64_t n_waves_total = total ? (total + wave_size - 1) / wave_size : 0;
	const bool tree = sc.has_specular == 3;       // two children per vertex: the host reads a queue length back after every run (one lane)
	const int n_lanes = (pt->n_lanes > 1 && !prof && n_waves_total > 1 && !tree) ? (int) (n_waves_total < (uint64_t) pt->n_lanes ? n_waves_total : (uint64_t) pt->n_lanes) : 1;
	pt->lanes[0].stream = st;
	if(n_lanes > 1 && !pt->ev_fork) Y4_CUDA(ctx, cudaEventCreateWithFlags(&pt->ev_fork, cudaEventDisableTiming));
	for(int l = 1; l < n_lanes; ++l)
	{
		if(!pt->lanes[l].stream) Y4_CUDA(ctx, cudaStreamCreateWithFlags(&pt->lanes[l].stream, cudaStreamNonBlocking));
		if(!pt->lanes[l].done) Y4_CUDA(ctx, cudaEventCreateWithFlags(&pt->lanes[l].done, cudaEventDisableTiming));
	}
	for(int l = 0; l < n_lanes; ++l) { const int rc = pt_alloc_wave(ctx, pt->lanes[l], wave_size); if(rc) return rc; }
	const int tree_depth = (sc.raydepth < 19 ? sc.raydepth : 19) + 2;
	if(tree) for(int l = 0; l < n_lanes; ++l) { const int rc = pt_alloc_tree(ctx, pt->lanes[l], wave_size, tree_depth); if(rc) return rc; }
	// the compacting bounce loop carries no incoming ray / hit for extra light samples at a path vertex (k > 0): lights with one sample only
	const bool compact = pt->compact != 0 && pt->max_light_samples == 1 && sc.n_paths > 0;
	if(compact)
	{
		for(int l = 0; l < n_lanes; ++l) { const int rc = pt_alloc_compact(ctx, pt->lanes[l], wave_size); if(rc) return rc; }
		static bool smem_set[16] = {false};
		if(!smem_set[ctx->device & 15])
		{
			Y4_CUDA(ctx, cudaFuncSetAttribute(k_pt_path_compact<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kCompactSmem));
			Y4_CUDA(ctx, cudaFuncSetAttribute(k_pt_path_compact<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kCompactSmem));
			Y4_CUDA(ctx, cudaFuncSetAttribute(k_pt_path_compact<5, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kCompactSmem));
			Y4_CUDA(ctx, cudaFuncSetAttribute(k_pt_path_compact<6, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kCompactSmem));
			smem_set[ctx->device & 15] = true;
		}
	}
	const int trace_blocks = n_lanes > 1 ? pt->trace_blocks_shared : 0;
	const int shadow_kind = pt->params.transp_shadows ? Y4_TRACE_TSHADOW : Y4_TRACE_SHADOW;
	const int stage_minb = n_lanes > 1 && pt->stage_min_blocks < 3 ? 3 : pt->stage_min_blocks;
	Y4PtPass ps; ps.n_samples = n_samples; ps.pass_offs = (uint32_t) offset + sc.base_sampling_offset; ps.aa_pass = aa_pass_number;
	const Y4PtScene &dsc = sc;
	// the (light, sample) sequence of estimateAllDirectLight at the camera hit
	std::vector<std::pair<int, int>> all_lk;
	for(int l = 0; l < sc.n_lights; ++l)
		for(int k = 0, ns = y4_light_num_samples(sc, pt->h_lights[l]); k < ns; ++k) all_lk.push_back(std::make_pair(l, k));
	for(int k = 0; k < sc.ao_samples; ++k) all_lk.push_back(std::make_pair(sc.n_lights, k));     // then sampleAmbientOcclusion: pseudo-light n_lights
	const int max_ls = pt->max_light_samples;
	// counter slots: [0..5] traversal visit counters (profile mode), then one per queue / trace launch
	const uint64_t n_waves = total ? (total + wave_size - 1) / wave_size : 0;
	const size_t n_levels = (size_t) (sc.has_specular ? (sc.raydepth < 20 ? sc.raydepth : 20) : 0) + 1;
	const size_t per_level = 6 + 2 * all_lk.size() + (size_t) sc.n_paths * (2 + ((size_t) sc.bounces + 1) * (4 + 2 * (size_t) max_ls));
	// tree mode: one run per tree node a pixel-sample visits, at most 2^(raydepth + 1) - 1 (capped: deeper trees fail the pass)
	const size_t tree_runs = tree ? (sc.raydepth < 7 ? ((size_t) 2 << sc.raydepth) : 256) : 0;
	const size_t per_wave = tree ? tree_runs * (per_level + 2) : (sc.has_specular == 2 ? n_levels * per_level : per_level + 2 * n_levels);
	const size_t need = 8 + n_waves * per_wave;
	if(need > pt->counters_cap)
	{
		cudaFree(pt->d_counters); cudaFreeHost(pt->h_counters); pt->d_counters = nullptr; pt->h_counters = nullptr; pt->counters_cap = 0;
		Y4_CUDA(ctx, cudaMalloc(&pt->d_counters, need * sizeof(unsigned long long)));
		Y4_CUDA(ctx, cudaHostAlloc(&pt->h_counters, need * sizeof(unsigned long long), cudaHostAllocDefault));
		pt->counters_cap = need;
	}
	Y4_CUDA(ctx, cudaMemsetAsync(pt->d_counters, 0, need * sizeof(unsigned long long), st));
	pt->counter_kinds.clear(); pt->dense_closest = 0;
	size_t next_slot = 8;
	bool slot_overflow = false;       // the estimate `need` was too small: a launch would share a counter -> the pass fails instead
	auto new_slot = [&]() -> unsigned long long * { if(next_slot >= need) { slot_overflow = true; return pt->d_counters + need - 1; } return pt->d_counters + next_slot++; };
	// a path queue and a shadow queue filled by the same stage: adjacent halves of one 64-bit slot
	auto new_queue_pair = [&](int kind_path, int kind_shadow, uint32_t *&path_count, uint32_t *&shadow_count)
	{
		if(kind_path >= 0) pt->counter_kinds.push_back(std::make_pair((uint32_t) next_slot, kind_path));
		if(kind_shadow >= 0) pt->counter_kinds.push_back(std::make_pair((uint32_t) next_slot | 0x80000000u, kind_shadow));
		path_count = reinterpret_cast<uint32_t *>(new_slot());
		shadow_count = path_count + 1;
	};
	auto new_queue_count = [&](int kind) -> uint32_t *
	{
		if(kind >= 0) pt->counter_kinds.push_back(std::make_pair((uint32_t) next_slot, kind));
		return reinterpret_cast<uint32_t *>(new_slot());
	};
	// profile mode: an event after every launch attributes device time to kernel classes; traversal runs counted
	std::vector<int> prof_cls;
	auto mark = [&](int cls) -> int
	{
		if(!prof) return Y4GPU_OK;
		if(prof_cls.size() >= pt->prof_events.size()) { cudaEvent_t e; Y4_CUDA(ctx, cudaEventCreate(&e)); pt->prof_events.push_back(e); }
		Y4_CUDA(ctx, cudaEventRecord(pt->prof_events[prof_cls.size()], st));
		prof_cls.push_back(cls);
		return Y4GPU_OK;
	};
	{ const int rc_ = mark(-1); if(rc_) return rc_; }
#define Y4_PT_LAUNCH(cls, ...) do { __VA_ARGS__; ++own_launches; const int rc_ = mark(cls); if(rc_) return rc_; } while(0)
#define Y4_PT_TRACE(kind, rays, hits, n_max, queue, count) do { \
		int rc_ = (kind == Y4_TRACE_TSHADOW && sc.tshadow_filter) \
			? y4_launch_tshadow_wave(ctx, rays, hits, queue, count, pt->params.shadow_depth, pt->d_mat_transp, (uint32_t) sc.n_mats, new_slot(), st, trace_blocks) \
			: y4_launch_trace_wave(ctx, kind, prof, rays, hits, n_max, queue, count, new_slot(), pt->d_counters + (kind == Y4_TRACE_CLOSEST ? 0 : 3), st, trace_blocks); \
		if(rc_ || (rc_ = mark(kind == Y4_TRACE_CLOSEST ? 1 : 2))) return rc_; } while(0)
	if(n_lanes > 1)
	{
		// fork: the other lanes start after everything queued so far on the context's stream (scene upload, counter reset)
		Y4_CUDA(ctx, cudaEventRecord(pt->ev_fork, pt->lanes[0].stream));
		for(int l = 1; l < n_lanes; ++l) Y4_CUDA(ctx, cudaStreamWaitEvent(pt->lanes[l].stream, pt->ev_fork, 0));
	}
	uint64_t wave_index = 0;
	for(uint64_t first = 0; first < total; first += wave_size, ++wave_index)
	{
		Y4Pt::Lane &lane = pt->lanes[wave_index % (uint64_t) n_lanes];
		st = lane.stream;
		Y4PtWave W = lane.wave;
		W.n = (uint32_t) ((total - first) < wave_size ? (total - first) : wave_size);
		const int grid = pt_grid(ctx, W.n);
		StageArgs a = {0, -1, -1, -1, -1, 0, 0, 0};
		StageQueues Q = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
		// Ray levels (recursiveRaytrace).  Pure mirrors reflect inside the camera-hit step, so whole chains are followed before
		// anything is estimated and their end points share one run of the estimate below.  A shinydiffuse vertex with a mirror
		// component needs its own estimate first: then every level is one run (camera-hit step, estimate, level-finish stage).
		const int max_level = (sc.has_specular && sc.raydepth > 0) ? (sc.raydepth < 19 ? sc.raydepth : 19) : 0;
		const bool level_runs = sc.has_specular >= 2;
		uint32_t *next_level_count = nullptr;                    // length of the queue of rays reflected into `level + 1`
		// tree mode: `level` counts RUNS (one tree node per pixel-sample and run; every entry carries its own ray level in W.sw),
		// the loop ends when no pixel-sample has a tree ray left
		for(int level = 0; level <= (tree ? (int) tree_runs : (level_runs ? max_level : 0)); ++level)
		{
		a = {0, -1, -1, -1, -1, 0, 0, level};
		Q = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
		if(level == 0)
		{
			// camera rays -> closest hits -> camera-hit stage with the first light sample
			Y4_PT_LAUNCH(0, k_pt_gen<<<grid, kPtThreads, 0, st>>>(dsc, W, ps, d_list, first, n_pix, ns_local, shard_index, shard_count));
			Y4_PT_TRACE(Y4_TRACE_CLOSEST, W.ray, W.hit, W.n, nullptr, nullptr); pt->dense_closest += W.n;
		}
		else
		{
			Q.in_q = lane.q_spec[level & 1]; Q.in_count = next_level_count;
			Y4_PT_TRACE(Y4_TRACE_CLOSEST, W.ray0, W.hit, W.n, Q.in_q, Q.in_count);
		}
		uint32_t *n_hit0 = nullptr;
		new_queue_pair(2, 1, n_hit0, Q.sh_count);
		Q.out_q = lane.q_hit0; Q.out_count = n_hit0; Q.sh_q = lane.q_shadow;
		a.l = all_lk.empty() ? -1 : all_lk[0].first;
		if(level < max_level || tree) { Q.spec_q = lane.q_spec[(level + 1) & 1]; Q.spec_count = next_level_count = new_queue_count(4); }
		if(tree) Y4_PT_LAUNCH(3, launch_stage<kStPrimary3>(stage_minb, grid, st, dsc, W, ps, a, Q));
		else Y4_PT_LAUNCH(3, launch_stage<kStPrimary>(stage_minb, grid, st, dsc, W, ps, a, Q));
		for(int lv = 1; !level_runs && lv <= max_level; ++lv)
		{
			const uint32_t *in_q = Q.spec_q, *in_count = Q.spec_count;
			Y4_PT_TRACE(Y4_TRACE_CLOSEST, W.ray0, W.hit, W.n, in_q, in_count);
			Q.in_q = in_q; Q.in_count = in_count;
			Q.spec_q = lv < max_level ? lane.q_spec[(lv + 1) & 1] : nullptr; Q.spec_count = lv < max_level ? new_queue_count(4) : nullptr;
			a.level = lv;
			Y4_PT_LAUNCH(3, launch_stage<kStPrimary>(stage_minb, grid, st, dsc, W, ps, a, Q));
		}
		Q.in_q = nullptr; Q.in_count = nullptr; Q.spec_q = nullptr; Q.spec_count = nullptr; a.level = 0;
		for(size_t j = 0; j < all_lk.size(); ++j)
		{
			Y4_PT_TRACE(shadow_kind, W.sray, W.shit, 2ull * W.n, lane.q_shadow, Q.sh_count);
			if(j + 1 == all_lk.size()) break;          // the last sample is resolved by the start stage
			Q.in_q = lane.q_hit0; Q.in_count = n_hit0; Q.out_q = nullptr; Q.out_count = nullptr; Q.sh_count = new_queue_count(1);
			a.all_lights = 1; a.l_prev = all_lk[j].first; a.k_prev = all_lk[j].second; a.l = all_lk[j + 1].first; a.k = all_lk[j + 1].second;
			Y4_PT_LAUNCH(4, launch_stage<kStDirectStep>(stage_minb, grid, st, dsc, W, ps, a, Q));
		}
		if(sc.n_paths == 0 && !all_lk.empty())
		{
			// direct lighting only (DirectLightIntegrator): just fold the last light sample into the pixel colour
			Q.in_q = lane.q_hit0; Q.in_count = n_hit0; Q.out_q = lane.q_path[0]; Q.out_count = new_queue_count(-1); Q.sh_q = lane.q_shadow; Q.sh_count = nullptr;
			a.all_lights = 1; a.sub = 0; a.l_prev = all_lk.back().first; a.k_prev = all_lk.back().second;
			Y4_PT_LAUNCH(3, launch_stage<kStStart>(stage_minb, grid, st, dsc, W, ps, a, Q));
		}
		Y4Ray *const ray_buf[2] = {lane.wave.ray, lane.wave.ray_next};
		for(int sub = 0; sub < sc.n_paths; ++sub)
		{
			int cur = 0;
			uint32_t *n_cur = new_queue_count(0);
			W.ray = ray_buf[0]; W.ray_next = ray_buf[0];
			Q.in_q = lane.q_hit0; Q.in_count = n_hit0; Q.out_q = lane.q_path[cur]; Q.out_count = n_cur; Q.sh_q = lane.q_shadow; Q.sh_count = nullptr;
			a.all_lights = 1; a.sub = sub; a.l_prev = (sub == 0 && !all_lk.empty()) ? all_lk.back().first : -1; a.k_prev = a.l_prev >= 0 ? all_lk.back().second : -1;
			Y4_PT_LAUNCH(3, launch_stage<kStStart>(stage_minb, grid, st, dsc, W, ps, a, Q));
			if(compact)
			{
				// bounce 0 reads the pixel-sample wave through the start stage's queue; from then on the paths live in the two dense
				// path waves, `rd` the one the traced rays / hits are in
				int rd = -1;
				for(int depth = 0; depth <= sc.bounces; ++depth)
				{
					const bool closing = depth == sc.bounces;
					if(closing && !sc.n_lights) break;
					const int wr = rd < 0 ? 0 : (rd ^ 1);
					Y4PtWave Wr = rd < 0 ? W : lane.cw[rd], Ww = lane.cw[wr];
					Wr.n = W.n; Ww.n = W.n;
					if(rd < 0) Wr.ray = ray_buf[0];
					const uint32_t *in_q = rd < 0 ? lane.q_path[0] : nullptr;
					if(!closing) Y4_PT_TRACE(Y4_TRACE_CLOSEST, Wr.ray, Wr.hit, W.n, in_q, n_cur);
					uint32_t *n_next = nullptr;
					new_queue_pair(depth + 1 < sc.bounces ? 0 : (closing ? -1 : 3), sc.n_lights && !closing ? 1 : -1, n_next, Q.sh_count);
					Q.in_q = in_q; Q.in_count = n_cur; Q.out_q = nullptr; Q.out_count = n_next; Q.sh_q = lane.q_shadow;
					a.all_lights = 0; a.depth = depth; a.sub = sub; a.l_prev = -1; a.k_prev = sc.n_lights ? 0 : -1;
					{
						const uint32_t want = (W.n + kCompactThreads - 1) / kCompactThreads, full = (uint32_t) ctx->sm_count * 12u;
						const int cgrid = (int) (want < full ? (want ? want : 1u) : full);
						const uint32_t *orig_r = rd < 0 ? nullptr : lane.corig[rd];
						static const int cminb = getenv("Y4GPU_PT_COMPACT_MINB") ? atoi(getenv("Y4GPU_PT_COMPACT_MINB")) : 5;      // 96 registers, 5 CTAs of 128 threads per SM: the measured optimum (profiles/r2_v5_sweep_compact.log)
						if(sc.has_specular) Y4_PT_LAUNCH(3, (k_pt_path_compact<4, true><<<cgrid, kCompactThreads, kCompactSmem, st>>>(dsc, Wr, Ww, W, ps, a, Q, orig_r, lane.corig[wr])));
						else if(cminb >= 6) Y4_PT_LAUNCH(3, (k_pt_path_compact<6, false><<<cgrid, kCompactThreads, kCompactSmem, st>>>(dsc, Wr, Ww, W, ps, a, Q, orig_r, lane.corig[wr])));
						else if(cminb == 5) Y4_PT_LAUNCH(3, (k_pt_path_compact<5, false><<<cgrid, kCompactThreads, kCompactSmem, st>>>(dsc, Wr, Ww, W, ps, a, Q, orig_r, lane.corig[wr])));
						else Y4_PT_LAUNCH(3, (k_pt_path_compact<4, false><<<cgrid, kCompactThreads, kCompactSmem, st>>>(dsc, Wr, Ww, W, ps, a, Q, orig_r, lane.corig[wr])));
					}
					if(closing) break;
					if(sc.n_lights) Y4_PT_TRACE(shadow_kind, Ww.sray, Ww.shit, 2ull * W.n, lane.q_shadow, Q.sh_count);
					rd = wr; n_cur = n_next;
				}
				continue;
			}
			for(int depth = 0; depth <= sc.bounces; ++depth)
			{
				const bool closing = depth == sc.bounces;
				if(closing && !sc.n_lights) break;                       // nothing waits to be folded in
				W.ray = ray_buf[depth & 1]; W.ray_next = ray_buf[(depth + 1) & 1];
				if(!closing) Y4_PT_TRACE(Y4_TRACE_CLOSEST, W.ray, W.hit, W.n, lane.q_path[cur], n_cur);
				uint32_t *n_next = nullptr;
				new_queue_pair(depth + 1 < sc.bounces ? 0 : (closing ? -1 : 3), sc.n_lights && !closing ? 1 : -1, n_next, Q.sh_count);
				Q.in_q = lane.q_path[cur]; Q.in_count = n_cur; Q.out_q = lane.q_path[cur ^ 1]; Q.out_count = n_next;
				Q.sh_q = lane.q_shadow;
				a.all_lights = 0; a.depth = depth; a.sub = sub; a.l_prev = -1; a.k_prev = sc.n_lights ? max_ls - 1 : -1;
				Y4_PT_LAUNCH(3, launch_stage<kStPath>(stage_minb, grid, st, dsc, W, ps, a, Q));
				if(closing) break;
				if(sc.n_lights)
				{
					Y4_PT_TRACE(shadow_kind, W.sray, W.shit, 2ull * W.n, lane.q_shadow, Q.sh_count);
					for(int k = 1; k < max_ls; ++k)
					{
						Q.out_q = nullptr; Q.out_count = nullptr; Q.sh_count = new_queue_count(1);
						a.l_prev = -1; a.k_prev = k - 1; a.l = -1; a.k = k;
						Y4_PT_LAUNCH(4, launch_stage<kStDirectStep>(stage_minb, grid, st, dsc, W, ps, a, Q));
						Y4_PT_TRACE(shadow_kind, W.sray, W.shit, 2ull * W.n, lane.q_shadow, Q.sh_count);
					}
				}
				cur ^= 1; n_cur = n_next;
			}
		}
		W.ray = ray_buf[0]; W.ray_next = ray_buf[1];
		if(tree)
		{
			// close the vertices that were estimated, then see whether any pixel-sample still has a tree ray to trace
			Q = {lane.q_hit0, n_hit0, nullptr, nullptr, nullptr, nullptr, lane.q_spec[(level + 1) & 1], next_level_count};
			Y4_PT_LAUNCH(3, launch_stage<kStFinish3>(stage_minb, grid, st, dsc, W, ps, a, Q));
			const size_t slot = (size_t) (reinterpret_cast<unsigned long long *>(next_level_count) - pt->d_counters);
			Y4_CUDA(ctx, cudaMemcpyAsync(pt->h_counters + slot, pt->d_counters + slot, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
			Y4_CUDA(ctx, cudaStreamSynchronize(st));
			if((uint32_t) pt->h_counters[slot] == 0u) break;
			if(level == (int) tree_runs)
			{
				ctx->err = "ray tree of a pixel-sample has more nodes than the run limit (raydepth too high for the tree mode)";
				return Y4GPU_ERR_UNSUPPORTED;
			}
		}
		else if(level_runs && level < max_level)
		{
			Q = {lane.q_hit0, n_hit0, nullptr, nullptr, nullptr, nullptr, lane.q_spec[(level + 1) & 1], next_level_count};
			a.level = level;
			Y4_PT_LAUNCH(3, launch_stage<kStLevelFinish>(stage_minb, grid, st, dsc, W, ps, a, Q));
		}
		}   // ray levels
		Y4_PT_LAUNCH(5, k_pt_splat<<<grid, kPtThreads, 0, st>>>(dsc, W, ps, pt->d_film_w, pt->d_film_rgba));
		Y4_CUDA(ctx, cudaGetLastError());
	}
#undef Y4_PT_LAUNCH
#undef Y4_PT_TRACE
	st = ctx->stream;
	if(slot_overflow)
	{
		for(int l = 0; l < n_lanes; ++l) cudaStreamSynchronize(pt->lanes[l].stream);
		ctx->err = "internal: per-pass counter slots exhausted (launch sequence longer than estimated)";
		return Y4GPU_ERR_CUDA;
	}
	if(n_lanes > 1)
	{
		// join: the context's stream continues after the other lanes
		for(int l = 1; l < n_lanes; ++l)
		{
			Y4_CUDA(ctx, cudaEventRecord(pt->lanes[l].done, pt->lanes[l].stream));
			Y4_CUDA(ctx, cudaStreamWaitEvent(st, pt->lanes[l].done, 0));
		}
	}
	Y4_CUDA(ctx, cudaMemcpyAsync(pt->h_counters, pt->d_counters, need * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
	pt->counters_read = false;
	Y4_CUDA(ctx, cudaEventRecord(pt->ev1, st));
	ctx->stats.kernel_launches += own_launches;
	pt->stats.samples = total;
	pt->stats.rays_closest = 0; pt->stats.rays_shadow = 0;
	pt->stats.kernel_launches = ctx->stats.kernel_launches - launches0;
	for(int q = 0; q < 8; ++q) { pt->stats.class_ms[q] = 0.; pt->stats.class_launches[q] = 0; }
	if(prof)
	{
		Y4_CUDA(ctx, cudaStreamSynchronize(st));
		for(size_t e = 1; e < prof_cls.size(); ++e)
		{
			if(prof_cls[e] < 0) continue;
			float ms = 0.f;
			Y4_CUDA(ctx, cudaEventElapsedTime(&ms, pt->prof_events[e - 1], pt->prof_events[e]));
			pt->stats.class_ms[prof_cls[e]] += ms; pt->stats.class_launches[prof_cls[e]]++;
		}
	}
	return Y4GPU_OK;
}
