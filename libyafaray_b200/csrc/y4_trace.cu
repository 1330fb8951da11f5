// sm_100a traversal kernels over the compressed BVH8.
//
// Persistent-thread design: the grid is sized to (SM count x resident CTAs) and every warp
// repeatedly claims a packet of 32 consecutive rays from a global work counter, so long
// rays never leave the rest of a CTA idle.  Rays and hits move as 128-bit words
// (2 x float4 in, 1 x float4 out per ray); node and triangle fetches are 128-bit
// read-only loads (y4_ld_u4 / y4_ld_tri).
#include "y4_ctx.h"

namespace {

constexpr int kTraceThreads = 256;

template<int KIND, bool COUNT>
__global__ void __launch_bounds__(kTraceThreads) k_trace(Y4Bvh8 bvh, const Y4Ray *__restrict__ rays, uint64_t n, Y4Hit *__restrict__ out,
                                                         uint32_t *__restrict__ work_counter, unsigned long long *__restrict__ visit)
{
	const uint32_t lane = threadIdx.x & 31u;
	uint2 stack[Y4_STACK_SIZE];
	unsigned long long my_nodes = 0, my_tris = 0;
	const uint64_t n_packets = (n + 31u) >> 5;
	while(true)
	{
		uint32_t packet = 0;
		if(lane == 0) packet = atomicAdd(work_counter, 1u);
		packet = __shfl_sync(0xffffffffu, packet, 0);
		if(packet >= n_packets) break;
		const uint64_t i = ((uint64_t) packet << 5) + lane;
		if(i < n)
		{
			const float4 a = __ldg(reinterpret_cast<const float4 *>(rays + i));
			const float4 b = __ldg(reinterpret_cast<const float4 *>(rays + i) + 1);
			Y4Ray r; r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
			const Y4PreparedRay pr = y4_prepare_ray(r, KIND);
			Y4Hit h;
			Y4TraceStats s; s.nodes = 0; s.tris = 0;
			y4_trace_ray<KIND>(bvh, pr, h, stack, COUNT ? &s : nullptr, nullptr, nullptr, nullptr, 0);
			if(COUNT) { my_nodes += s.nodes; my_tris += s.tris; }
			float4 o; o.x = h.t; o.y = h.bu; o.z = h.bv; o.w = __int_as_float(h.prim);
			*reinterpret_cast<float4 *>(out + i) = o;
		}
	}
	if(COUNT)
	{
#pragma unroll
		for(int o = 16; o > 0; o >>= 1)
		{
			my_nodes += __shfl_xor_sync(0xffffffffu, my_nodes, o);
			my_tris += __shfl_xor_sync(0xffffffffu, my_tris, o);
		}
		if(lane == 0) { atomicAdd(visit, my_nodes); atomicAdd(visit + 1, my_tris); }
	}
}

// Transparent shadows (reference intersectTs, accelerator_kdtree.cc:911-1061): distinct
// transparent primitives multiply their material's transparency colour; an opaque blocker or
// more than max_depth transparent layers means shadowed.
__global__ void __launch_bounds__(kTraceThreads) k_trace_tshadow(Y4Bvh8 bvh, const Y4Ray *__restrict__ rays, uint64_t n, int max_depth,
                                                                 const float *__restrict__ mat_transp, uint32_t n_mat, y4gpu_tshit *__restrict__ out,
                                                                 uint32_t *__restrict__ work_counter)
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
		const Y4PreparedRay pr = y4_prepare_ray(r, Y4_TRACE_TSHADOW);
		Y4Hit h; int n_filt = 0;
		y4_trace_ray<Y4_TRACE_TSHADOW>(bvh, pr, h, stack, nullptr, filt, filt_mat, &n_filt, max_depth);
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

template<int KIND, bool COUNT>
int launch(y4gpu_ctx *ctx, const Y4Ray *d_rays, uint64_t n, Y4Hit *d_out)
{
	int per_sm = 0;
	Y4_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace<KIND, COUNT>, kTraceThreads, 0));
	if(per_sm < 1) per_sm = 1;
	const uint64_t want = (n + kTraceThreads - 1) / kTraceThreads;
	const uint64_t full = (uint64_t) ctx->sm_count * per_sm;
	const int grid = (int) (want < full ? want : full);
	k_trace<KIND, COUNT><<<grid, kTraceThreads, 0, ctx->stream>>>(ctx->view(), d_rays, n, d_out, ctx->d_work_counter, ctx->d_visit_counters);
	return Y4GPU_OK;
}

} // namespace

int y4_launch_trace(y4gpu_ctx *ctx, int kind, bool counted, const Y4Ray *d_rays, uint64_t n, Y4Hit *d_out)
{
	if(n == 0) return Y4GPU_OK;
	cudaStream_t st = ctx->stream;
	Y4_CUDA(ctx, cudaMemsetAsync(ctx->d_work_counter, 0, sizeof(uint32_t), st));
	if(counted) Y4_CUDA(ctx, cudaMemsetAsync(ctx->d_visit_counters, 0, 2 * sizeof(unsigned long long), st));
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
	int rc;
	if(kind == Y4_TRACE_CLOSEST) rc = counted ? launch<Y4_TRACE_CLOSEST, true>(ctx, d_rays, n, d_out) : launch<Y4_TRACE_CLOSEST, false>(ctx, d_rays, n, d_out);
	else if(kind == Y4_TRACE_SHADOW) rc = counted ? launch<Y4_TRACE_SHADOW, true>(ctx, d_rays, n, d_out) : launch<Y4_TRACE_SHADOW, false>(ctx, d_rays, n, d_out);
	else { ctx->err = "bad trace kind"; return Y4GPU_ERR_ARG; }
	if(rc) return rc;
	Y4_CUDA(ctx, cudaGetLastError());
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
	ctx->stats.kernel_launches++;
	ctx->stats.last_trace_rays = n;
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
	k_trace_tshadow<<<(int) (want < full ? want : full), kTraceThreads, 0, st>>>(ctx->view(), d_rays, n, max_depth, d_mat_transp, n_mat, d_out, ctx->d_work_counter);
	Y4_CUDA(ctx, cudaGetLastError());
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
	ctx->stats.kernel_launches++;
	ctx->stats.last_trace_rays = n;
	return Y4GPU_OK;
}
