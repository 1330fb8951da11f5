// sm_100a traversal kernels over the compressed BVH8.
//
// Persistent-thread design: the grid is sized to (SM count x resident CTAs) and every warp
// repeatedly claims a packet of 32 consecutive rays from a global work counter, so long
// rays never leave the rest of a CTA idle.  Rays and hits move as 128-bit words
// (2 x float4 in, 1 x float4 out per ray); node and triangle fetches are 128-bit
// read-only loads (y4_ld_u4 / y4_ld_tri).
#include "y4_ctx.h"
#include <cub/cub.cuh>

namespace {

constexpr int kTraceThreads = 256;

// Warp-level scheduling around the per-ray logic of y4_trace.cuh:
//  * dynamic ray fetch -- lanes whose ray has terminated are refilled from the global ray counter as
//    soon as fewer than `fetch_threshold` lanes of the warp are still traversing (warp-level ray
//    compaction: the warp never idles behind its longest ray);
//  * postponed triangle tests -- when less than 1/postpone_div of the lanes that entered the
//    triangle phase still have triangles pending, the rest of the triangle group goes onto the
//    traversal stack so the warp returns to (convergent) node traversal;
//  * shadow rays leave at the first accepted blocker (any-hit) and their lane is refilled.
// Results do not depend on the schedule: closest hits use the order-independent rule y4_accept_t.
// QUEUE: the batch is an index queue built on the device (the path tracer's compacted wavefronts):
// entry q names ray/hit slot queue[q] and the entry count is read from *queue_count.
template<int KIND, bool COUNT, int MINB, bool QUEUE>
__global__ void __launch_bounds__(kTraceThreads, MINB) k_trace(Y4Bvh8 bvh, const Y4Ray *__restrict__ rays, uint64_t n, Y4Hit *__restrict__ out,
                                                         unsigned long long *__restrict__ work_counter, unsigned long long *__restrict__ visit,
                                                         int fetch_threshold, int postpone_div,
                                                         const uint32_t *__restrict__ queue, const uint32_t *__restrict__ queue_count)
{
	if(QUEUE) n = __ldg(queue_count);
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t lt_mask = (1u << lane) - 1u;
	const uint32_t vis_bit = KIND == Y4_TRACE_CLOSEST ? Y4_TRI_VIS_CLOSEST : (KIND == Y4_TRACE_SHADOW ? Y4_TRI_VIS_SHADOW : Y4_TRI_VIS_TSHADOW);
	uint2 stack[Y4_STACK_SIZE];
	unsigned long long my_nodes = 0, my_tris = 0, my_live = 0;
	bool have = false, exhausted = false;
	uint64_t ray_i = 0;
	Y4RayCtx rc = y4_ray_ctx(v3(0.f, 0.f, 0.f), v3(0.f, 0.f, 1.f));
	float t_lo = 0.f, t_hi = 0.f;
	Y4Hit hit; hit.t = 0.f; hit.bu = 0.f; hit.bv = 0.f; hit.prim = -1;
	uint2 ngroup; ngroup.x = 0u; ngroup.y = 0u;
	int sp = 0;
	const uint32_t magic = __ldg(bvh.consts);
	while(true)
	{
		// ---- refill idle lanes -------------------------------------------------------------
		const uint32_t idle = __ballot_sync(0xffffffffu, !have);
		if(idle && !exhausted)
		{
			const uint32_t n_idle = __popc(idle);
			unsigned long long base = 0;
			if(lane == 0) base = atomicAdd(work_counter, (unsigned long long) n_idle);
			base = __shfl_sync(0xffffffffu, base, 0);
			exhausted = base + n_idle >= n;
			const uint64_t mine = base + __popc(idle & lt_mask);
			if(!have && mine < n)
			{
				const uint64_t slot_i = (QUEUE && queue) ? (uint64_t) __ldg(queue + mine) : mine;      // QUEUE without a queue: the first *queue_count slots
				const float4 a = __ldg(reinterpret_cast<const float4 *>(rays + slot_i));
				const float4 b = __ldg(reinterpret_cast<const float4 *>(rays + slot_i) + 1);
				Y4Ray r; r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
				const Y4PreparedRay pr = y4_prepare_ray(r, KIND);
				rc = y4_ray_ctx(pr.o, pr.d);
				t_lo = pr.t_lo; t_hi = pr.t_hi;
				hit.t = 0.f; hit.bu = 0.f; hit.bv = 0.f; hit.prim = -1;
				ray_i = slot_i; sp = 0; have = true;
				// virtual parent of the root: child_base 0, imask = slot 0, hit bit at slot 0's priority
				// empty scene or empty interval (the path tracer parks finished paths as null rays): nothing to visit
				ngroup.x = 0u; ngroup.y = (bvh.n_tris && pr.t_hi > pr.t_lo) ? ((1u << (24u + rc.oct_inv)) | 1u) : 0u;
				if(COUNT && ngroup.y) my_live++;
			}
		}
		if(!__any_sync(0xffffffffu, have)) break;
		const int leave_below = exhausted ? 0 : fetch_threshold;
		// ---- traverse ----------------------------------------------------------------------
		while(have)
		{
			uint2 tgroup; tgroup.x = 0u; tgroup.y = 0u;
			if(ngroup.y & 0xff000000u)
			{
				const uint32_t bit = 31u - (uint32_t) __clz((int) ngroup.y);
				ngroup.y &= ~(1u << bit);
				const uint32_t slot = (bit - 24u) ^ rc.oct_inv;
				const uint32_t child = ngroup.x + __popc(ngroup.y & 0xffu & ((1u << slot) - 1u));
				if(ngroup.y & 0xff000000u) stack[sp++] = ngroup;        // siblings still pending
				const Y4U4 *nd = bvh.nodes + (size_t) child * Y4_NODE_WORDS;
				const Y4U4 w0 = y4_ld_u4(nd), w1 = y4_ld_u4(nd + 1), w2 = y4_ld_u4(nd + 2), w3 = y4_ld_u4(nd + 3), w4 = y4_ld_u4(nd + 4);
				if(COUNT) my_nodes++;
				const uint32_t hits = y4_node_hits(w0, w1, w2, w3, w4, rc.o, rc.idir, rc.oct_inv, t_lo, t_hi, magic);
				ngroup.x = w1.x; ngroup.y = (hits & 0xff000000u) | (w0.w >> 24);
				tgroup.x = w1.y; tgroup.y = hits & 0x00ffffffu;
			}
			else
			{
				tgroup = ngroup;            // a postponed triangle group came off the stack (or nothing)
				ngroup.y = 0u;
			}
			// triangles of this group
			const int entered = __popc(__activemask());
			bool done = false;
			while(tgroup.y)
			{
				if(postpone_div > 0 && sp < bvh.postpone_sp_limit && __popc(__activemask()) * postpone_div < entered)
				{
					stack[sp++] = tgroup;
					break;
				}
				const uint32_t bit = 31u - (uint32_t) __clz((int) tgroup.y);
				tgroup.y &= ~(1u << bit);
				const Y4Tri tr = y4_ld_tri(bvh.tris + tgroup.x + bit);
				if(COUNT) my_tris++;
				float t, u, v;
				if(!y4_tri_intersect(rc.o, rc.d, v3(tr.v0x, tr.v0y, tr.v0z), v3(tr.e1x, tr.e1y, tr.e1z), v3(tr.e2x, tr.e2y, tr.e2z), tr.eps, t, u, v)) continue;
				if(!(tr.flags & vis_bit)) continue;
				if(!y4_accept_t<KIND>(t, t_lo, t_hi, tr.prim, hit.prim)) continue;
				hit.t = t; hit.bu = u; hit.bv = v; hit.prim = (int32_t) tr.prim;
				if(KIND == Y4_TRACE_CLOSEST) t_hi = t;
				else { done = true; break; }
			}
			if(!done && (ngroup.y & 0xff000000u) == 0u)
			{
				if(sp == 0) done = true;
				else ngroup = stack[--sp];
			}
			if(done)
			{
				float4 o; o.x = hit.t; o.y = hit.bu; o.z = hit.bv; o.w = __int_as_float(hit.prim);
				*reinterpret_cast<float4 *>(out + ray_i) = o;
				have = false;
				break;
			}
			if(__popc(__activemask()) < leave_below) break;
		}
		__syncwarp();
	}
	if(COUNT)
	{
#pragma unroll
		for(int o = 16; o > 0; o >>= 1)
		{
			my_nodes += __shfl_xor_sync(0xffffffffu, my_nodes, o);
			my_tris += __shfl_xor_sync(0xffffffffu, my_tris, o);
			my_live += __shfl_xor_sync(0xffffffffu, my_live, o);
		}
		if(lane == 0) { atomicAdd(visit, my_nodes); atomicAdd(visit + 1, my_tris); atomicAdd(visit + 2, my_live); }
	}
}

// Transparent shadows (reference intersectTs, accelerator_kdtree.cc:911-1061): distinct
// transparent primitives multiply their material's transparency colour; an opaque blocker or
// more than max_depth transparent layers means shadowed.
template<bool TWO>
__global__ void __launch_bounds__(kTraceThreads) k_trace_tshadow(Y4Bvh8 bvh, Y4Scene2 sc2, const Y4Ray *__restrict__ rays, uint64_t n, int max_depth,
                                                                 const float *__restrict__ mat_transp, uint32_t n_mat, y4gpu_tshit *__restrict__ out,
                                                                 uint32_t *__restrict__ work_counter, int inner)
{
	const uint32_t lane = threadIdx.x & 31u;
	uint2 stack[Y4_STACK_SIZE];
	uint32_t filt[16], filt_mat[16];
	const uint64_t n_packets = (n + 31u) >> 5;
	if(max_depth > 16) max_depth = 16;
	while(true)
	{
		uint32_t packet = 0;
		if(lane == 0) packet = atomicAdd(work_counter, 1u);
		packet = __shfl_sync(0xffffffffu, packet, 0);
		if(packet >= n_packets) break;
		const uint64_t i = ((uint64_t) packet << 5) + lane;
		if(i >= n) continue;
		const float4 a = __ldg(reinterpret_cast<const float4 *>(rays + i));
		const float4 b = __ldg(reinterpret_cast<const float4 *>(rays + i) + 1);
		Y4Ray r; r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
		Y4PreparedRay pr = y4_prepare_ray(r, Y4_TRACE_TSHADOW);
		if(inner)
		{
			// the arguments of the inner virtual intersectTs(sray, max_depth, dist, ...) (accelerator_kdtree.cc:911-1061): the origin was
			// advanced by the wrapper already, accept sray.tmin <= t < dist
			pr.o = v3(r.ox, r.oy, r.oz); pr.t_lo = r.tmin; pr.t_hi = (r.tmax >= 0.f) ? r.tmax : INFINITY;
		}
		Y4Hit h; int n_filt = 0;
		if(TWO) y4_trace_ray2<Y4_TRACE_TSHADOW>(sc2, pr, h, stack, nullptr, filt, filt_mat, &n_filt, max_depth);
		else y4_trace_ray<Y4_TRACE_TSHADOW>(bvh, pr, h, stack, nullptr, filt, filt_mat, &n_filt, max_depth);
		float fr = 1.f, fg = 1.f, fb = 1.f;
		for(int k = 0; k < n_filt; ++k)
		{
			const uint32_t m = filt_mat[k];
			if(m < n_mat) { fr *= mat_transp[3 * m]; fg *= mat_transp[3 * m + 1]; fb *= mat_transp[3 * m + 2]; }
		}
		y4gpu_tshit o;
		o.hit.t = h.t; o.hit.bu = h.bu; o.hit.bv = h.bv; o.hit.prim = h.prim;
		o.filter_r = fr; o.filter_g = fg; o.filter_b = fb; o.pad = 0.f;
		out[i] = o;
	}
}

// The same over an index queue of the path tracer: the filter colour goes into the hit record's t / bu / bv lanes
// (prim >= 0: shadowed).  One packet counter per launch (a zeroed 64-bit slot of the pass).
template<bool TWO>
__global__ void __launch_bounds__(kTraceThreads) k_trace_tshadow_queue(Y4Bvh8 bvh, Y4Scene2 sc2, const Y4Ray *__restrict__ rays, const uint32_t *__restrict__ queue,
                                                                       const uint32_t *__restrict__ queue_count, int max_depth, const float *__restrict__ mat_transp,
                                                                       uint32_t n_mat, Y4Hit *__restrict__ out, unsigned long long *__restrict__ work_counter)
{
	const uint32_t lane = threadIdx.x & 31u;
	uint2 stack[Y4_STACK_SIZE];
	uint32_t filt[16], filt_mat[16];
	const uint32_t n = *queue_count;
	const uint32_t n_packets = (n + 31u) >> 5;
	if(max_depth > 16) max_depth = 16;
	while(true)
	{
		uint32_t packet = 0;
		if(lane == 0) packet = atomicAdd(reinterpret_cast<uint32_t *>(work_counter), 1u);
		packet = __shfl_sync(0xffffffffu, packet, 0);
		if(packet >= n_packets) break;
		const uint32_t q = (packet << 5) + lane;
		if(q >= n) continue;
		const uint32_t i = queue[q];
		const float4 a = __ldg(reinterpret_cast<const float4 *>(rays + i));
		const float4 b = __ldg(reinterpret_cast<const float4 *>(rays + i) + 1);
		Y4Ray r; r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
		const Y4PreparedRay pr = y4_prepare_ray(r, Y4_TRACE_TSHADOW);
		Y4Hit h; h.t = 0.f; h.bu = 0.f; h.bv = 0.f; h.prim = -1;
		int n_filt = 0;
		if(pr.t_hi > pr.t_lo)
		{
			if(TWO) y4_trace_ray2<Y4_TRACE_TSHADOW>(sc2, pr, h, stack, nullptr, filt, filt_mat, &n_filt, max_depth);
			else y4_trace_ray<Y4_TRACE_TSHADOW>(bvh, pr, h, stack, nullptr, filt, filt_mat, &n_filt, max_depth);
		}
		float fr = 1.f, fg = 1.f, fb = 1.f;
		for(int k = 0; k < n_filt; ++k)
		{
			const uint32_t m = filt_mat[k];
			if(m < n_mat) { fr = __fmul_rn(fr, mat_transp[3 * m]); fg = __fmul_rn(fg, mat_transp[3 * m + 1]); fb = __fmul_rn(fb, mat_transp[3 * m + 2]); }
		}
		*reinterpret_cast<float4 *>(out + i) = make_float4(fr, fg, fb, __uint_as_float((uint32_t) h.prim));
	}
}

// Two-level scenes: the walk of y4_trace_ray2 (y4_trace.cuh) with k_trace's warp scheduling -- lanes whose ray ended are
// refilled from the work counter as soon as fewer than fetch_threshold lanes are busy.  Per-lane state between steps: the
// world ray, the current space's ray (world or the instance's object space), the instance index and its tree offsets.
// COUNT: visit counters as in k_trace (visit[0] nodes of either level, visit[1] triangle tests, visit[2] live rays); an instance
// entry -- its 48-byte top-level leaf record + the 128-byte instance record -- is counted as two nodes (160 B).
template<int KIND, bool QUEUE, bool COUNT>
__global__ void __launch_bounds__(kTraceThreads, 3) k_trace2(Y4Scene2 sc, const Y4Ray *__restrict__ rays, uint64_t n, Y4Hit *__restrict__ out,
                                                             unsigned long long *__restrict__ work_counter, unsigned long long *__restrict__ visit, int fetch_threshold,
                                                             const uint32_t *__restrict__ queue, const uint32_t *__restrict__ queue_count)
{
	unsigned long long my_nodes = 0, my_tris = 0, my_live = 0;
	if(QUEUE) n = __ldg(queue_count);
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t lt_mask = (1u << lane) - 1u;
	const uint32_t vis_bit = KIND == Y4_TRACE_CLOSEST ? Y4_TRI_VIS_CLOSEST : (KIND == Y4_TRACE_SHADOW ? Y4_TRI_VIS_SHADOW : Y4_TRI_VIS_TSHADOW);
	uint2 stack[Y4_STACK_SIZE];
	bool have = false, exhausted = false;
	uint64_t ray_i = 0;
	V3 wo = v3(0.f, 0.f, 0.f), wd = v3(0.f, 0.f, 1.f);
	Y4RayCtx rc = y4_ray_ctx(wo, wd);
	float t_lo = 0.f, t_hi = 0.f, pad = 0.f;
	Y4Hit hit; hit.t = 0.f; hit.bu = 0.f; hit.bv = 0.f; hit.prim = -1;
	uint2 ngroup; ngroup.x = 0u; ngroup.y = 0u;
	uint32_t cur = Y4_INST_NONE, node_off = 0u;
	int sp = 0;
	const uint32_t magic = __ldg(sc.consts);
	while(true)
	{
		const uint32_t idle = __ballot_sync(0xffffffffu, !have);
		if(idle && !exhausted)
		{
			const uint32_t n_idle = __popc(idle);
			unsigned long long base = 0;
			if(lane == 0) base = atomicAdd(work_counter, (unsigned long long) n_idle);
			base = __shfl_sync(0xffffffffu, base, 0);
			exhausted = base + n_idle >= n;
			const uint64_t mine = base + __popc(idle & lt_mask);
			if(!have && mine < n)
			{
				const uint64_t slot_i = (QUEUE && queue) ? (uint64_t) __ldg(queue + mine) : mine;
				const float4 a = __ldg(reinterpret_cast<const float4 *>(rays + slot_i));
				const float4 b = __ldg(reinterpret_cast<const float4 *>(rays + slot_i) + 1);
				Y4Ray r; r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
				const Y4PreparedRay pr = y4_prepare_ray(r, KIND);
				wo = pr.o; wd = pr.d;
				rc = y4_ray_ctx(wo, wd);
				t_lo = pr.t_lo; t_hi = pr.t_hi; pad = 0.f;
				hit.t = 0.f; hit.bu = 0.f; hit.bv = 0.f; hit.prim = -1;
				ray_i = slot_i; sp = 0; have = true; cur = Y4_INST_NONE; node_off = 0u;
				ngroup.x = 0u; ngroup.y = (sc.n_inst && pr.t_hi > pr.t_lo) ? ((1u << (24u + rc.oct_inv)) | 1u) : 0u;
				if(COUNT && ngroup.y) my_live++;
			}
		}
		if(!__any_sync(0xffffffffu, have)) break;
		const int leave_below = exhausted ? 0 : fetch_threshold;
		while(have)
		{
			uint2 tgroup; tgroup.x = 0u; tgroup.y = 0u;
			bool done = false;
			if(ngroup.y & 0xff000000u)
			{
				const uint32_t bit = 31u - (uint32_t) __clz((int) ngroup.y);
				ngroup.y &= ~(1u << bit);
				const uint32_t slot = (bit - 24u) ^ rc.oct_inv;
				const uint32_t child = ngroup.x + __popc(ngroup.y & 0xffu & ((1u << slot) - 1u));
				if(ngroup.y & 0xff000000u) stack[sp++] = ngroup;
				const Y4U4 *nd = (cur == Y4_INST_NONE ? sc.tlas_nodes : sc.blas_nodes + (size_t) node_off * Y4_NODE_WORDS) + (size_t) child * Y4_NODE_WORDS;
				const Y4U4 w0 = y4_ld_u4(nd), w1 = y4_ld_u4(nd + 1), w2 = y4_ld_u4(nd + 2), w3 = y4_ld_u4(nd + 3), w4 = y4_ld_u4(nd + 4);
				if(COUNT) my_nodes++;
				const uint32_t hits = y4_node_hits<true>(w0, w1, w2, w3, w4, rc.o, rc.idir, rc.oct_inv, t_lo, t_hi, magic, pad);
				ngroup.x = w1.x; ngroup.y = (hits & 0xff000000u) | (w0.w >> 24);
				tgroup.x = w1.y; tgroup.y = hits & 0x00ffffffu;
			}
			else
			{
				tgroup = ngroup;
				ngroup.y = 0u;
			}
			if(tgroup.y)
			{
				if(cur == Y4_INST_NONE)
				{
					// top level: enter the first instance of the group; the rest and the pending node group wait under the marker
					const uint32_t bit = 31u - (uint32_t) __clz((int) tgroup.y);
					tgroup.y &= ~(1u << bit);
					const uint32_t id = y4_ld_tri(sc.tlas_leaves + tgroup.x + bit).prim;
					if(ngroup.y & 0xff000000u) stack[sp++] = ngroup;
					if(tgroup.y) stack[sp++] = tgroup;
					uint2 mark; mark.x = 0xffffffffu; mark.y = 0u;
					stack[sp++] = mark;
					const Y4Inst in = y4_ld_inst(sc.inst + id);
					if(COUNT) my_nodes += 2;
					cur = id; node_off = in.node_off;
					V3 oo, od;
					y4_inst_enter(in, wo, wd, oo, od, pad);
					rc = y4_ray_ctx(oo, od);
					ngroup.x = 0u; ngroup.y = (1u << (24u + rc.oct_inv)) | 1u;
				}
				else
				{
					const Y4Inst in = y4_ld_inst(sc.inst + cur);
					while(tgroup.y)
					{
						const uint32_t bit = 31u - (uint32_t) __clz((int) tgroup.y);
						tgroup.y &= ~(1u << bit);
						const Y4TriObj tr = y4_ld_triobj(sc.blas_tris + in.tri_off + tgroup.x + bit);
						if(COUNT) my_tris++;
						float t, u, v;
						if(!y4_inst_tri(in, tr, wo, wd, t, u, v)) continue;
						if(!(tr.flags & vis_bit)) continue;
						const uint32_t prim = in.first_prim + tr.prim_local;
						if(!y4_accept_t<KIND>(t, t_lo, t_hi, prim, hit.prim)) continue;
						hit.t = t; hit.bu = u; hit.bv = v; hit.prim = (int32_t) prim;
						if(KIND == Y4_TRACE_CLOSEST) t_hi = t;
						else { done = true; break; }
					}
				}
			}
			if(!done && (ngroup.y & 0xff000000u) == 0u)
			{
				while(true)
				{
					if(sp == 0) { done = true; break; }
					ngroup = stack[--sp];
					if(ngroup.x != 0xffffffffu || ngroup.y != 0u) break;
					cur = Y4_INST_NONE; node_off = 0u; pad = 0.f; rc = y4_ray_ctx(wo, wd);      // the marker: back at the top level
				}
			}
			if(done)
			{
				float4 o; o.x = hit.t; o.y = hit.bu; o.z = hit.bv; o.w = __int_as_float(hit.prim);
				*reinterpret_cast<float4 *>(out + ray_i) = o;
				have = false;
				break;
			}
			if(__popc(__activemask()) < leave_below) break;
		}
		__syncwarp();
	}
	if(COUNT)
	{
#pragma unroll
		for(int o = 16; o > 0; o >>= 1)
		{
			my_nodes += __shfl_xor_sync(0xffffffffu, my_nodes, o);
			my_tris += __shfl_xor_sync(0xffffffffu, my_tris, o);
			my_live += __shfl_xor_sync(0xffffffffu, my_live, o);
		}
		if(lane == 0) { atomicAdd(visit, my_nodes); atomicAdd(visit + 1, my_tris); atomicAdd(visit + 2, my_live); }
	}
}

struct TraceArgs
{
	const Y4Ray *rays; uint64_t n; Y4Hit *out;
	unsigned long long *work_counter, *visit;
	const uint32_t *queue, *queue_count;
	cudaStream_t stream;
	int max_blocks_per_sm;      // 0: as many as fit (persistent grid = one resident wave)
	int fetch_threshold = 0;    // 0: the context's (ray batches); the path tracer's wavefronts pass their own
};

template<int KIND, bool COUNT, int MINB, bool QUEUE>
int launch_k(y4gpu_ctx *ctx, const TraceArgs &a)
{
	static int per_sm_cached[16] = {0};          // per device
	int &per_sm = per_sm_cached[ctx->device & 15];
	if(per_sm < 1)
	{
		Y4_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace<KIND, COUNT, MINB, QUEUE>, kTraceThreads, 0));
		if(per_sm < 1) per_sm = 1;
	}
	const uint64_t want = (a.n + kTraceThreads - 1) / kTraceThreads;
	const int use_per_sm = (a.max_blocks_per_sm > 0 && a.max_blocks_per_sm < per_sm) ? a.max_blocks_per_sm : per_sm;
	const uint64_t full = (uint64_t) ctx->sm_count * use_per_sm;     // persistent: one wave of resident CTAs
	const int grid = (int) (want < full ? want : full);
	k_trace<KIND, COUNT, MINB, QUEUE><<<grid, kTraceThreads, 0, a.stream>>>(ctx->view(), a.rays, a.n, a.out, a.work_counter, a.visit,
	                                                                          a.fetch_threshold > 0 ? a.fetch_threshold : ctx->fetch_threshold, ctx->postpone_div, a.queue, a.queue_count);
	return Y4GPU_OK;
}
template<int KIND, bool QUEUE, bool COUNT>
int launch2(y4gpu_ctx *ctx, const TraceArgs &a)
{
	static int per_sm_cached[16] = {0};
	int &per_sm = per_sm_cached[ctx->device & 15];
	if(per_sm < 1)
	{
		Y4_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace2<KIND, QUEUE, COUNT>, kTraceThreads, 0));
		if(per_sm < 1) per_sm = 1;
	}
	const uint64_t want = (a.n + kTraceThreads - 1) / kTraceThreads;
	const int use_per_sm = (a.max_blocks_per_sm > 0 && a.max_blocks_per_sm < per_sm) ? a.max_blocks_per_sm : per_sm;
	const uint64_t full = (uint64_t) ctx->sm_count * use_per_sm;
	k_trace2<KIND, QUEUE, COUNT><<<(int) (want < full ? want : full), kTraceThreads, 0, a.stream>>>(ctx->view2(), a.rays, a.n, a.out, a.work_counter, a.visit,
	                                                                                             a.fetch_threshold > 0 ? a.fetch_threshold : ctx->fetch_threshold, a.queue, a.queue_count);
	return Y4GPU_OK;
}
template<int KIND, bool QUEUE>
int launch(y4gpu_ctx *ctx, bool counted, const TraceArgs &a)
{
	if(ctx->two_level) return counted ? launch2<KIND, QUEUE, true>(ctx, a) : launch2<KIND, QUEUE, false>(ctx, a);
	if(counted) return launch_k<KIND, true, 3, QUEUE>(ctx, a);
	return ctx->min_blocks >= 4 ? launch_k<KIND, false, 4, QUEUE>(ctx, a) : launch_k<KIND, false, 3, QUEUE>(ctx, a);
}

// ---- ray sorting ---------------------------------------------------------------------------------------------------
// Incoherent batches over a BVH much larger than L2 (C5: 50 M triangles, 3.6 GB) are bound by DRAM sector fetches: every
// lane of a warp walks its own part of the tree.  Tracing the batch in the order of a 32-bit key -- the Morton code of the
// ray origin's cell in the tree bound, then a 6-bit direction class -- puts rays that start close together and head the same
// way into the same warps, so their node and triangle fetches hit the same lines.  Only the ORDER changes: k_trace runs over
// an index queue (its QUEUE variant) and writes every hit to the ray's own slot, so results are bit-identical.
__device__ __forceinline__ uint32_t y4_spread3(uint32_t v)        // 10 bits -> every third bit
{
	v = (v | (v << 16)) & 0x030000FFu;
	v = (v | (v << 8)) & 0x0300F00Fu;
	v = (v | (v << 4)) & 0x030C30C3u;
	v = (v | (v << 2)) & 0x09249249u;
	return v;
}
__device__ __forceinline__ uint32_t y4_ray_sort_key(const float4 a, const float4 b, float lx, float ly, float lz, float sx, float sy, float sz,
                                                    int obits, int dir_major)
{
	const int top = (1 << obits) - 1;
	const int cx = min(max((int) ((a.x - lx) * sx), 0), top), cy = min(max((int) ((a.y - ly) * sy), 0), top), cz = min(max((int) ((a.z - lz) * sz), 0), top);
	const uint32_t m = y4_spread3((uint32_t) cx) | (y4_spread3((uint32_t) cy) << 1) | (y4_spread3((uint32_t) cz) << 2);
	const float ax = fabsf(b.x), ay = fabsf(b.y), az = fabsf(b.z);
	const uint32_t oct = (b.x < 0.f ? 4u : 0u) | (b.y < 0.f ? 2u : 0u) | (b.z < 0.f ? 1u : 0u);
	const uint32_t dom = ax >= ay ? (ax >= az ? 0u : 2u) : (ay >= az ? 1u : 2u);
	const float big = fmaxf(ax, fmaxf(ay, az)), mid = ax + ay + az - big - fminf(ax, fminf(ay, az));
	const uint32_t d = (oct << 3) | (dom << 1) | (mid >= 0.5f * big ? 1u : 0u);
	return dir_major ? ((d << (3 * obits)) | m) : ((m << 6) | d);
}
__global__ void __launch_bounds__(256) k_ray_sort_keys(const Y4Ray *__restrict__ rays, uint32_t n, float lx, float ly, float lz, float sx, float sy, float sz,
                                                       int obits, int dir_major, uint32_t *__restrict__ keys, uint32_t *__restrict__ idx, uint32_t *__restrict__ count)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i == 0) *count = n;
	if(i >= n) return;
	const float4 a = __ldg(reinterpret_cast<const float4 *>(rays + i)), b = __ldg(reinterpret_cast<const float4 *>(rays + i) + 1);
	keys[i] = y4_ray_sort_key(a, b, lx, ly, lz, sx, sy, sz, obits, dir_major);
	idx[i] = i;
}
// Coherence probe of the automatic mode: over every 64th group of 32 consecutive rays, the number of distinct coarse keys
// (4 bits per axis + direction class) in the group.  Camera rays give 1-3, uniformly random rays 32.
__global__ void __launch_bounds__(256) k_ray_sort_probe(const Y4Ray *__restrict__ rays, uint32_t n_groups, uint32_t group_stride, float lx, float ly, float lz,
                                                        float sx, float sy, float sz, unsigned long long *__restrict__ distinct_sum)
{
	const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
	if(warp >= n_groups) return;
	const uint64_t i = (uint64_t) warp * group_stride * 32ull + lane;
	const float4 a = __ldg(reinterpret_cast<const float4 *>(rays + i)), b = __ldg(reinterpret_cast<const float4 *>(rays + i) + 1);
	const uint32_t key = y4_ray_sort_key(a, b, lx, ly, lz, sx, sy, sz, 4, 0);
	const uint32_t same = __match_any_sync(0xffffffffu, key);
	const uint32_t leaders = __ballot_sync(0xffffffffu, (same & ((1u << lane) - 1u)) == 0u);
	if(lane == 0) atomicAdd(distinct_sum, (unsigned long long) __popc(leaders));
}

} // namespace

// Scratch of one sorted batch: [keys | keys' | idx | idx' | count + probe sum (256 B) | CUB temporary storage]
static size_t sort_layout(uint64_t n, size_t &cub_bytes)
{
	cub_bytes = 0;
	cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const uint32_t *) nullptr, (uint32_t *) nullptr, (const uint32_t *) nullptr, (uint32_t *) nullptr, (int) n);
	return (size_t) n * 16 + 256 + cub_bytes;
}

int y4_sort_scratch_reserve(y4gpu_ctx *ctx, Y4SortScratch &sc, uint64_t n)
{
	size_t cub_bytes;
	const size_t need = sort_layout(n, cub_bytes);
	if(need <= sc.cap) return Y4GPU_OK;
	if(sc.buf) cudaFree(sc.buf);
	sc.buf = nullptr; sc.cap = 0;
	Y4_CUDA(ctx, cudaMalloc(&sc.buf, need));
	sc.cap = need;
	return Y4GPU_OK;
}

// Decides whether a resident batch is worth sorting (mode 2): large enough, tree much larger than L2, and incoherent as laid out.
// Costs one small kernel and one host synchronisation of `stream`.
int y4_ray_sort_wanted(y4gpu_ctx *ctx, const Y4Ray *d_rays, uint64_t n, Y4SortScratch &sc, cudaStream_t stream, bool &want)
{
	want = false;
	if(ctx->ray_sort == 0 || n < (1u << 16) || n >= (1ull << 31)) return Y4GPU_OK;
	if(ctx->ray_sort == 1) { want = true; return Y4GPU_OK; }
	if((ctx->stats.node_bytes + ctx->stats.tri_bytes) < ctx->ray_sort_min_bvh_bytes) return Y4GPU_OK;
	const int rc = y4_sort_scratch_reserve(ctx, sc, n);
	if(rc) return rc;
	unsigned long long *d_sum = reinterpret_cast<unsigned long long *>(static_cast<char *>(sc.buf) + (size_t) n * 16 + 8);
	const uint32_t stride = 64, n_groups = (uint32_t) (n / (32ull * stride));
	if(n_groups == 0) return Y4GPU_OK;
	const float *bd = ctx->bound;
	const float sx = 16.f / fmaxf(bd[3] - bd[0], 1e-30f), sy = 16.f / fmaxf(bd[4] - bd[1], 1e-30f), sz = 16.f / fmaxf(bd[5] - bd[2], 1e-30f);
	Y4_CUDA(ctx, cudaMemsetAsync(d_sum, 0, 8, stream));
	k_ray_sort_probe<<<(n_groups * 32 + 255) / 256, 256, 0, stream>>>(d_rays, n_groups, stride, bd[0], bd[1], bd[2], sx, sy, sz, d_sum);
	unsigned long long sum = 0;
	Y4_CUDA(ctx, cudaMemcpyAsync(&sum, d_sum, 8, cudaMemcpyDeviceToHost, stream));
	Y4_CUDA(ctx, cudaStreamSynchronize(stream));
	ctx->last_probe_distinct = (float) ((double) sum / (double) n_groups);
	want = ctx->last_probe_distinct >= ctx->ray_sort_min_distinct;
	return Y4GPU_OK;
}

// keys -> radix sort of (key, ray index) -> k_trace over the sorted index queue; everything on `stream`
int y4_launch_trace_sorted(y4gpu_ctx *ctx, int kind, const Y4Ray *d_rays, uint64_t n, Y4Hit *d_out, unsigned long long *d_work_counter,
                           Y4SortScratch &sc, cudaStream_t stream)
{
	const int rc0 = y4_sort_scratch_reserve(ctx, sc, n);
	if(rc0) return rc0;
	size_t cub_bytes;
	sort_layout(n, cub_bytes);
	char *base = static_cast<char *>(sc.buf);
	uint32_t *keys = reinterpret_cast<uint32_t *>(base), *keys2 = keys + n, *idx = keys2 + n, *idx2 = idx + n;
	uint32_t *count = reinterpret_cast<uint32_t *>(base + (size_t) n * 16);
	void *cub_tmp = base + (size_t) n * 16 + 256;
	const int obits = ctx->ray_sort_obits < 1 ? 1 : (ctx->ray_sort_obits > 8 ? 8 : ctx->ray_sort_obits);
	const float cells = (float) (1 << obits);
	const float *bd = ctx->bound;
	const float sx = cells / fmaxf(bd[3] - bd[0], 1e-30f), sy = cells / fmaxf(bd[4] - bd[1], 1e-30f), sz = cells / fmaxf(bd[5] - bd[2], 1e-30f);
	k_ray_sort_keys<<<(unsigned) ((n + 255) / 256), 256, 0, stream>>>(d_rays, (uint32_t) n, bd[0], bd[1], bd[2], sx, sy, sz, obits, ctx->ray_sort_dir_major, keys, idx, count);
	Y4_CUDA(ctx, cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys2, idx, idx2, (int) n, 0, 3 * obits + 6, stream));
	TraceArgs a; a.rays = d_rays; a.n = n; a.out = d_out; a.work_counter = d_work_counter; a.visit = nullptr; a.queue = idx2; a.queue_count = count;
	a.stream = stream; a.max_blocks_per_sm = 0;
	int rc;
	if(kind == Y4_TRACE_CLOSEST) rc = launch<Y4_TRACE_CLOSEST, true>(ctx, false, a);
	else if(kind == Y4_TRACE_SHADOW) rc = launch<Y4_TRACE_SHADOW, true>(ctx, false, a);
	else { ctx->err = "bad trace kind"; return Y4GPU_ERR_ARG; }
	if(rc) return rc;
	Y4_CUDA(ctx, cudaGetLastError());
	return Y4GPU_OK;
}

int y4_launch_trace(y4gpu_ctx *ctx, int kind, bool counted, const Y4Ray *d_rays, uint64_t n, Y4Hit *d_out)
{
	if(n == 0) return Y4GPU_OK;
	cudaStream_t st = ctx->stream;
	Y4_CUDA(ctx, cudaMemsetAsync(ctx->d_work_counter, 0, sizeof(unsigned long long), st));
	if(counted) Y4_CUDA(ctx, cudaMemsetAsync(ctx->d_visit_counters, 0, 3 * sizeof(unsigned long long), st));
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev0, st));      // (the probe and the sort are inside the timed span)
	bool sorted = false;
	if(!counted) { const int rcs = y4_ray_sort_wanted(ctx, d_rays, n, ctx->sort_scratch, st, sorted); if(rcs) return rcs; }
	ctx->stats.last_trace_sorted = sorted ? 1 : 0;
	if(sorted)
	{
		const int rcs = y4_launch_trace_sorted(ctx, kind, d_rays, n, d_out, reinterpret_cast<unsigned long long *>(ctx->d_work_counter), ctx->sort_scratch, st);
		if(rcs) return rcs;
		Y4_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
		ctx->stats.kernel_launches += 2;        // key kernel + traversal (the radix sort is CUB's)
		ctx->stats.last_trace_rays = n;
		return Y4GPU_OK;
	}
	TraceArgs a; a.rays = d_rays; a.n = n; a.out = d_out; a.work_counter = reinterpret_cast<unsigned long long *>(ctx->d_work_counter);
	a.visit = ctx->d_visit_counters; a.queue = nullptr; a.queue_count = nullptr; a.stream = st; a.max_blocks_per_sm = 0;
	int rc;
	if(kind == Y4_TRACE_CLOSEST) rc = launch<Y4_TRACE_CLOSEST, false>(ctx, counted, a);
	else if(kind == Y4_TRACE_SHADOW) rc = launch<Y4_TRACE_SHADOW, false>(ctx, counted, a);
	else { ctx->err = "bad trace kind"; return Y4GPU_ERR_ARG; }
	if(rc) return rc;
	Y4_CUDA(ctx, cudaGetLastError());
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
	ctx->stats.kernel_launches++;
	ctx->stats.last_trace_rays = n;
	return Y4GPU_OK;
}

// Wavefront entry of the path tracer: the batch is either dense (queue == null, n rays) or a device-built
// index queue of at most n_max entries whose length is *queue_count.  work_counter must be zero (the
// caller hands out a fresh pre-zeroed slot per launch); visit (3 counters) is accumulated when counted.
int y4_launch_trace_wave(y4gpu_ctx *ctx, int kind, bool counted, const Y4Ray *d_rays, Y4Hit *d_out, uint64_t n_max,
                         const uint32_t *d_queue, const uint32_t *d_queue_count, unsigned long long *d_work_counter, unsigned long long *d_visit,
                         cudaStream_t stream, int max_blocks_per_sm)
{
	if(n_max == 0) return Y4GPU_OK;
	TraceArgs a; a.rays = d_rays; a.n = n_max; a.out = d_out; a.work_counter = d_work_counter; a.visit = d_visit; a.queue = d_queue; a.queue_count = d_queue_count;
	a.stream = stream; a.max_blocks_per_sm = max_blocks_per_sm; a.fetch_threshold = ctx->pt_fetch_threshold;
	int rc;
	const bool queued = d_queue || d_queue_count;       // a count without a queue: the dense front of a path wave
	if(kind == Y4_TRACE_CLOSEST) rc = queued ? launch<Y4_TRACE_CLOSEST, true>(ctx, counted, a) : launch<Y4_TRACE_CLOSEST, false>(ctx, counted, a);
	else if(kind == Y4_TRACE_SHADOW) rc = queued ? launch<Y4_TRACE_SHADOW, true>(ctx, counted, a) : launch<Y4_TRACE_SHADOW, false>(ctx, counted, a);
	// opaque-only transparent-shadow query of the path tracer: intersectTs ray range and visibility, first blocker ends the ray
	else if(kind == Y4_TRACE_TSHADOW) rc = queued ? launch<Y4_TRACE_TSHADOW, true>(ctx, counted, a) : launch<Y4_TRACE_TSHADOW, false>(ctx, counted, a);
	else { ctx->err = "bad trace kind"; return Y4GPU_ERR_ARG; }
	if(rc) return rc;
	ctx->stats.kernel_launches++;
	return Y4GPU_OK;
}

int y4_launch_tshadow_wave(y4gpu_ctx *ctx, const Y4Ray *d_rays, Y4Hit *d_out, const uint32_t *d_queue, const uint32_t *d_queue_count, int max_depth,
                           const float *d_mat_transp, uint32_t n_mat, unsigned long long *d_work_counter, cudaStream_t stream, int max_blocks_per_sm)
{
	const int per_sm = max_blocks_per_sm > 0 && max_blocks_per_sm < 4 ? max_blocks_per_sm : 4;
	if(ctx->two_level) k_trace_tshadow_queue<true><<<ctx->sm_count * per_sm, kTraceThreads, 0, stream>>>(Y4Bvh8{}, ctx->view2(), d_rays, d_queue, d_queue_count, max_depth, d_mat_transp, n_mat, d_out, d_work_counter);
	else k_trace_tshadow_queue<false><<<ctx->sm_count * per_sm, kTraceThreads, 0, stream>>>(ctx->view(), Y4Scene2{}, d_rays, d_queue, d_queue_count, max_depth, d_mat_transp, n_mat, d_out, d_work_counter);
	Y4_CUDA(ctx, cudaGetLastError());
	ctx->stats.kernel_launches++;
	return Y4GPU_OK;
}

int y4_launch_tshadow(y4gpu_ctx *ctx, const Y4Ray *d_rays, uint64_t n, int max_depth, const float *d_mat_transp, uint32_t n_mat, y4gpu_tshit *d_out)
{
	if(n == 0) return Y4GPU_OK;
	cudaStream_t st = ctx->stream;
	Y4_CUDA(ctx, cudaMemsetAsync(ctx->d_work_counter, 0, sizeof(uint32_t), st));
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
	const uint64_t want = (n + kTraceThreads - 1) / kTraceThreads;
	const uint64_t full = (uint64_t) ctx->sm_count * 4;
	if(ctx->two_level) k_trace_tshadow<true><<<(int) (want < full ? want : full), kTraceThreads, 0, st>>>(Y4Bvh8{}, ctx->view2(), d_rays, n, max_depth, d_mat_transp, n_mat, d_out, ctx->d_work_counter, 0);
	else k_trace_tshadow<false><<<(int) (want < full ? want : full), kTraceThreads, 0, st>>>(ctx->view(), Y4Scene2{}, d_rays, n, max_depth, d_mat_transp, n_mat, d_out, ctx->d_work_counter, 0);
	Y4_CUDA(ctx, cudaGetLastError());
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
	ctx->stats.kernel_launches++;
	ctx->stats.last_trace_rays = n;
	return Y4GPU_OK;
}

// Re-entrant launches: explicit stream and work counter (zeroed by the caller on that stream), no context state touched.
int y4_launch_trace_on(y4gpu_ctx *ctx, int kind, const Y4Ray *d_rays, uint64_t n, Y4Hit *d_out, unsigned long long *d_work_counter, cudaStream_t stream)
{
	if(n == 0) return Y4GPU_OK;
	TraceArgs a; a.rays = d_rays; a.n = n; a.out = d_out; a.work_counter = d_work_counter; a.visit = nullptr; a.queue = nullptr; a.queue_count = nullptr;
	a.stream = stream; a.max_blocks_per_sm = 0;
	int rc;
	if(kind == Y4_TRACE_CLOSEST) rc = launch<Y4_TRACE_CLOSEST, false>(ctx, false, a);
	else if(kind == Y4_TRACE_SHADOW) rc = launch<Y4_TRACE_SHADOW, false>(ctx, false, a);
	else { ctx->err = "bad trace kind"; return Y4GPU_ERR_ARG; }
	if(rc) return rc;
	Y4_CUDA(ctx, cudaGetLastError());
	return Y4GPU_OK;
}

int y4_launch_tshadow_on(y4gpu_ctx *ctx, const Y4Ray *d_rays, uint64_t n, int max_depth, const float *d_mat_transp, uint32_t n_mat, y4gpu_tshit *d_out,
                         bool inner, unsigned long long *d_work_counter, cudaStream_t stream)
{
	if(n == 0) return Y4GPU_OK;
	const uint64_t want = (n + kTraceThreads - 1) / kTraceThreads;
	const uint64_t full = (uint64_t) ctx->sm_count * 4;
if(ctx->two_level) k_trace_tshadow<true><<<(int) (want < full ? want : full), kTraceThreads, 0, stream>>>(Y4Bvh8{}, ctx->view2(), d_rays, n, max_depth, d_mat_transp, n_mat, d_out,
	                                                                                   reinterpret_cast<uint32_t *>(d_work_counter), inner ? 1 : 0);
	else k_trace_tshadow<false><<<(int) (want < full ? want : full), kTraceThreads, 0, stream>>>(ctx->view(), Y4Scene2{}, d_rays, n, max_depth, d_mat_transp, n_mat, d_out,
	                                                                                   reinterpret_cast<uint32_t *>(d_work_counter), inner ? 1 : 0);
	Y4_CUDA(ctx, cudaGetLastError());
	return Y4GPU_OK;
}

// ---- allocation-free ray batches (y4gpu_trace_noalloc) ------------------------------------------------
// For callers that may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the scratch
// k_trace needs -- its persistent-thread work counter, and the ray / hit staging used when the caller's buffers live in host
// memory -- sits in ONE file-scope device array of fixed size instead of the context's cudaMalloc'ed buffers.  Every GPU that
// loads this library holds the array (256 B + 48 B x Y4GPU_NOALLOC_MAX_RAYS = 3 MiB) until the process exits.  All calls on a
// device share it and it keeps the last call's contents: the counter is cleared by every call (the staging areas are fully
// overwritten before they are read), and calls that overlap ON THE GPU race.
namespace {

constexpr size_t kNoAllocHeader = 256;      // [0, 8): work counter of k_trace; the rest pads the ray area to 256 B
constexpr size_t kNoAllocRays = kNoAllocHeader;
constexpr size_t kNoAllocHits = kNoAllocRays + (size_t) Y4GPU_NOALLOC_MAX_RAYS * sizeof(Y4Ray);
constexpr size_t kNoAllocBytes = kNoAllocHits + (size_t) Y4GPU_NOALLOC_MAX_RAYS * sizeof(Y4Hit);

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once per
// device (with that device current) and cached.
std::mutex g_noalloc_mtx;
unsigned char *g_noalloc_dev[64] = {nullptr};

int noalloc_scratch(y4gpu_ctx *ctx, unsigned char **out)
{
	if(ctx->device < 0 || ctx->device >= 64) { ctx->err = "y4gpu_trace_noalloc: device index out of range"; return Y4GPU_ERR_ARG; }
	std::lock_guard<std::mutex> lock(g_noalloc_mtx);
	if(!g_noalloc_dev[ctx->device])
	{
		void *p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);      // the current device is ctx->device (set by the caller)
		if(e != cudaSuccess || !p)
		{
			ctx->err = std::string("y4gpu_trace_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e);
			cudaGetLastError();
			return Y4GPU_ERR_CUDA;
		}
		g_noalloc_dev[ctx->device] = static_cast<unsigned char *>(p);
	}
	*out = g_noalloc_dev[ctx->device];
	return Y4GPU_OK;
}

bool noalloc_is_device_ptr(const void *p)
{
	cudaPointerAttributes a;
	if(cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
	return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

} // namespace

extern "C" int y4gpu_trace_noalloc(y4gpu_ctx *ctx, int kind, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out)
{
	if(!ctx || (n && (!rays || !out))) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	if(kind != Y4_TRACE_CLOSEST && kind != Y4_TRACE_SHADOW) { ctx->err = "y4gpu_trace_noalloc: kind must be Y4GPU_TRACE_CLOSEST or Y4GPU_TRACE_SHADOW"; return Y4GPU_ERR_ARG; }
	if(!ctx->built) { ctx->err = "y4gpu_trace_noalloc before y4gpu_build"; return Y4GPU_ERR_ARG; }
	if(n > (uint64_t) Y4GPU_NOALLOC_MAX_RAYS)
	{
		ctx->err = "y4gpu_trace_noalloc: " + std::to_string(n) + " rays exceed the static scratch (Y4GPU_NOALLOC_MAX_RAYS = " + std::to_string(Y4GPU_NOALLOC_MAX_RAYS)
			+ "); split the batch or use y4gpu_trace";
		return Y4GPU_ERR_ARG;
	}
	if(n == 0) return Y4GPU_OK;
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	unsigned char *scratch = nullptr;
	{ const int rc = noalloc_scratch(ctx, &scratch); if(rc) return rc; }
	cudaStream_t st = ctx->stream;
	const bool rays_dev = noalloc_is_device_ptr(rays), out_dev = noalloc_is_device_ptr(out);
	const Y4Ray *d_rays = reinterpret_cast<const Y4Ray *>(rays);
	Y4Hit *d_out = reinterpret_cast<Y4Hit *>(out);
	if(!rays_dev)
	{
		Y4_CUDA(ctx, cudaMemcpyAsync(scratch + kNoAllocRays, rays, n * sizeof(Y4Ray), cudaMemcpyHostToDevice, st));
		d_rays = reinterpret_cast<const Y4Ray *>(scratch + kNoAllocRays);
	}
	if(!out_dev) d_out = reinterpret_cast<Y4Hit *>(scratch + kNoAllocHits);
	// the counter must start at zero; it still holds the previous call's final value
	Y4_CUDA(ctx, cudaMemsetAsync(scratch, 0, sizeof(unsigned long long), st));
	TraceArgs a; a.rays = d_rays; a.n = n; a.out = d_out; a.work_counter = reinterpret_cast<unsigned long long *>(scratch);
	a.visit = nullptr; a.queue = nullptr; a.queue_count = nullptr; a.stream = st; a.max_blocks_per_sm = 0;
	const int rc = kind == Y4_TRACE_CLOSEST ? launch<Y4_TRACE_CLOSEST, false>(ctx, false, a) : launch<Y4_TRACE_SHADOW, false>(ctx, false, a);
	if(rc) return rc;
	Y4_CUDA(ctx, cudaGetLastError());
	ctx->stats.kernel_launches++;
	if(!out_dev) Y4_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * sizeof(Y4Hit), cudaMemcpyDeviceToHost, st));
	// host buffers: the call is synchronous like y4gpu_trace; device buffers: queued on the context's stream, not waited for
	if(!out_dev || !rays_dev) Y4_CUDA(ctx, cudaStreamSynchronize(st));
	return Y4GPU_OK;
}
