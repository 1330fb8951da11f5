// Per-ray BVH8 traversal (closest-hit, any-hit shadow, transparent shadow).
//
// Replaces the reference kd-tree walks (/root/reference/src/accelerator/accelerator_kdtree.cc
// :634-773 closest, :775-904 shadow, :911-1061 transparent shadow) and reproduces their
// acceptance rules and the ray-range conventions of the Accelerator wrappers
// (src/accelerator/accelerator.cc:55-94).  Y4_HD: the same logic runs inside the sm_100a
// kernels (y4_kernels.cu) and in the host unit-test emulation.
#pragma once
#include "y4_common.cuh"
#include "y4_tri.cuh"
#include "y4_bvh8.cuh"

#define Y4_STACK_SIZE 64

struct Y4Bvh8
{
	const Y4U4 *nodes;     // 5 words per node, root = node 0
	const Y4Tri *tris;     // leaf-ordered triangle records
	uint32_t n_tris;       // 0 => empty scene
};

struct Y4TraceStats { uint32_t nodes, tris; };

// Ray prepared for traversal after the wrapper conventions were applied.
struct Y4PreparedRay
{
	V3 o, d;
	float t_lo;    // accept t >= t_lo
	float t_hi;    // accept t <  t_hi (shrinks for closest hit)
};

// accelerator.cc:55-67 (closest: t_max = tmax >= 0 ? tmax : inf, accept t >= ray.tmin)
// accelerator.cc:69-94 (shadow: origin advanced by dir*tmin, t_max = tmax - 2*tmin or inf;
//                       intersectS accepts 0 <= t < t_max, intersectTs accepts sray.tmin <= t < t_max)
Y4_HD Y4PreparedRay y4_prepare_ray(const Y4Ray &r, int kind)
{
	Y4PreparedRay p;
	p.d = v3(r.dx, r.dy, r.dz);
	if(kind == Y4_TRACE_CLOSEST)
	{
		p.o = v3(r.ox, r.oy, r.oz);
		p.t_lo = r.tmin;
		p.t_hi = (r.tmax >= 0.f) ? r.tmax : INFINITY;
	}
	else
	{
		p.o = xadd3(v3(r.ox, r.oy, r.oz), xscale(p.d, r.tmin));
		p.t_lo = (kind == Y4_TRACE_SHADOW) ? 0.f : r.tmin;
		p.t_hi = (r.tmax >= 0.f) ? xsub(r.tmax, xmul(2.f, r.tmin)) : INFINITY;
	}
	return p;
}

Y4_HD uint32_t y4_popc(uint32_t x)
{
#if defined(__CUDA_ARCH__)
	return (uint32_t) __popc(x);
#else
	return (uint32_t) __builtin_popcount(x);
#endif
}
Y4_HD uint32_t y4_msb(uint32_t x)   // index of highest set bit, x != 0
{
#if defined(__CUDA_ARCH__)
	return 31u - (uint32_t) __clz((int) x);
#else
	return 31u - (uint32_t) __builtin_clz(x);
#endif
}
Y4_HD float y4_safe_rcp(float d)
{
	const float a = fabsf(d) > 1e-20f ? d : copysignf(1e-20f, d);
	return 1.f / a;
}
// 128-bit read-only loads of node words and triangle records
Y4_HD Y4U4 y4_ld_u4(const Y4U4 *p)
{
#if defined(__CUDA_ARCH__)
	const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
	Y4U4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r;
#else
	return *p;
#endif
}
Y4_HD Y4Tri y4_ld_tri(const Y4Tri *p)
{
#if defined(__CUDA_ARCH__)
	const float4 a = __ldg(reinterpret_cast<const float4 *>(p)), b = __ldg(reinterpret_cast<const float4 *>(p) + 1), c = __ldg(reinterpret_cast<const float4 *>(p) + 2);
	Y4Tri t;
	t.v0x = a.x; t.v0y = a.y; t.v0z = a.z; t.eps = a.w;
	t.e1x = b.x; t.e1y = b.y; t.e1z = b.z; t.prim = __float_as_uint(b.w);
	t.e2x = c.x; t.e2y = c.y; t.e2z = c.z; t.flags = __float_as_uint(c.w);
	return t;
#else
	return *p;
#endif
}
Y4_HD uint32_t y4_byte(uint32_t w, int i) { return (w >> (8 * i)) & 0xffu; }

// Intersect one node's 8 quantised child boxes; returns the hit mask in "priority" bit
// positions: internal children at bits 24 + (slot ^ oct_inv), leaf triangles at their
// tri_base-relative offsets.
Y4_HD uint32_t y4_node_hits(const Y4U4 w0, const Y4U4 w1, const Y4U4 w2, const Y4U4 w3, const Y4U4 w4,
                            const V3 o, const V3 idir, const uint32_t oct_inv, const float t_lo, const float t_hi)
{
	const float sx = y4_u2f((w0.w & 0xffu) << 23), sy = y4_u2f(((w0.w >> 8) & 0xffu) << 23), sz = y4_u2f(((w0.w >> 16) & 0xffu) << 23);
	const float ax = sx * idir.x, ay = sy * idir.y, az = sz * idir.z;
	const float bx = (y4_u2f(w0.x) - o.x) * idir.x, by = (y4_u2f(w0.y) - o.y) * idir.y, bz = (y4_u2f(w0.z) - o.z) * idir.z;
	const bool nx = idir.x < 0.f, ny = idir.y < 0.f, nz = idir.z < 0.f;
	uint32_t hits = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
	for(int h = 0; h < 2; ++h)
	{
		const uint32_t meta4 = h ? w1.w : w1.z;
		const uint32_t qlx = h ? w2.y : w2.x, qly = h ? w2.w : w2.z, qlz = h ? w3.y : w3.x;
		const uint32_t qhx = h ? w3.w : w3.z, qhy = h ? w4.y : w4.x, qhz = h ? w4.w : w4.z;
		const uint32_t nearx = nx ? qhx : qlx, farx = nx ? qlx : qhx;
		const uint32_t neary = ny ? qhy : qly, fary = ny ? qly : qhy;
		const uint32_t nearz = nz ? qhz : qlz, farz = nz ? qlz : qhz;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
		for(int i = 0; i < 4; ++i)
		{
			const uint32_t m = y4_byte(meta4, i);
			const float t0x = (float) y4_byte(nearx, i) * ax + bx, t1x = (float) y4_byte(farx, i) * ax + bx;
			const float t0y = (float) y4_byte(neary, i) * ay + by, t1y = (float) y4_byte(fary, i) * ay + by;
			const float t0z = (float) y4_byte(nearz, i) * az + bz, t1z = (float) y4_byte(farz, i) * az + bz;
			const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, t_lo));
			const float tf = fminf(fminf(t1x, t1y), fminf(t1z, t_hi));
			if(tn <= tf * 1.0000005f && m != 0u)
			{
				const uint32_t is_inner = ((m & 0x18u) == 0x18u) ? 1u : 0u;   // low5 in 24..31
				const uint32_t bit = (m ^ (is_inner ? oct_inv : 0u)) & 31u;
				hits |= (m >> 5) << bit;
			}
		}
	}
	return hits;
}

// Full traversal of one ray.  KIND is a compile-time constant in the kernels.
// Closest: nearest accepted hit (strict '<' keeps the first-found among equal t, like the
// reference; the *order* differs from the kd-tree's, which is the documented near-tie class).
// Shadow: first accepted blocker.  TShadow: opaque blocker => shadowed; transparent hits are
// recorded in `filt` (distinct primitives) up to max_depth, see y4_kernels.cu for the colour.
template<int KIND>
Y4_HD void y4_trace_ray(const Y4Bvh8 &bvh, const Y4PreparedRay &ray, Y4Hit &hit, uint2 *stack,
                        Y4TraceStats *stats, uint32_t *filt, uint32_t *filt_mat, int *n_filt, int max_depth)
{
	hit.t = 0.f; hit.bu = 0.f; hit.bv = 0.f; hit.prim = -1;
	if(bvh.n_tris == 0) return;
	const uint32_t vis_bit = KIND == Y4_TRACE_CLOSEST ? Y4_TRI_VIS_CLOSEST : (KIND == Y4_TRACE_SHADOW ? Y4_TRI_VIS_SHADOW : Y4_TRI_VIS_TSHADOW);
	const V3 idir = v3(y4_safe_rcp(ray.d.x), y4_safe_rcp(ray.d.y), y4_safe_rcp(ray.d.z));
	const uint32_t oct_inv = (ray.d.x >= 0.f ? 4u : 0u) | (ray.d.y >= 0.f ? 2u : 0u) | (ray.d.z >= 0.f ? 1u : 0u);
	float t_hi = ray.t_hi;
	int sp = 0;
	// virtual parent of the root: child_base 0, imask = slot 0, hit bit at slot 0's priority
	uint2 ngroup; ngroup.x = 0u; ngroup.y = (1u << (24u + oct_inv)) | 1u;
	while(true)
	{
		uint2 tgroup; tgroup.x = 0u; tgroup.y = 0u;
		if(ngroup.y & 0xff000000u)
		{
			// pop the highest-priority child of the node group
			const uint32_t bit = y4_msb(ngroup.y);
			ngroup.y &= ~(1u << bit);
			const uint32_t slot = (bit - 24u) ^ oct_inv;
			const uint32_t child = ngroup.x + y4_popc(ngroup.y & 0xffu & ((1u << slot) - 1u));
			if(ngroup.y & 0xff000000u) stack[sp++] = ngroup;        // siblings still pending
			const Y4U4 *n = bvh.nodes + (size_t) child * Y4_NODE_WORDS;
			const Y4U4 w0 = y4_ld_u4(n), w1 = y4_ld_u4(n + 1), w2 = y4_ld_u4(n + 2), w3 = y4_ld_u4(n + 3), w4 = y4_ld_u4(n + 4);
			if(stats) stats->nodes++;
			const uint32_t hits = y4_node_hits(w0, w1, w2, w3, w4, ray.o, idir, oct_inv, ray.t_lo, t_hi);
			ngroup.x = w1.x; ngroup.y = (hits & 0xff000000u) | (w0.w >> 24);
			tgroup.x = w1.y; tgroup.y = hits & 0x00ffffffu;
		}
		else
		{
			ngroup.y = 0u;
		}
		// triangles of this node
		while(tgroup.y)
		{
			const uint32_t bit = y4_msb(tgroup.y);
			tgroup.y &= ~(1u << bit);
			const Y4Tri tr = y4_ld_tri(bvh.tris + tgroup.x + bit);
			if(stats) stats->tris++;
			float t, u, v;
			if(!y4_tri_intersect(ray.o, ray.d, v3(tr.v0x, tr.v0y, tr.v0z), v3(tr.e1x, tr.e1y, tr.e1z), v3(tr.e2x, tr.e2y, tr.e2z), tr.eps, t, u, v)) continue;
			if(!(t < t_hi && t >= ray.t_lo)) continue;
			if(!(tr.flags & vis_bit)) continue;
			if(KIND == Y4_TRACE_CLOSEST)
			{
				t_hi = t; hit.t = t; hit.bu = u; hit.bv = v; hit.prim = (int32_t) tr.prim;
			}
			else if(KIND == Y4_TRACE_SHADOW)
			{
				hit.t = t; hit.bu = u; hit.bv = v; hit.prim = (int32_t) tr.prim;
				return;
			}
			else
			{
				if(!(tr.flags & Y4_TRI_TRANSPARENT))
				{
					hit.t = t; hit.bu = u; hit.bv = v; hit.prim = (int32_t) tr.prim;   // opaque: shadowed
					return;
				}
				bool seen = false;
				for(int k = 0; k < *n_filt; ++k) seen = seen || (filt[k] == tr.prim);
				if(!seen)
				{
					if(*n_filt >= max_depth) { hit.t = t; hit.bu = u; hit.bv = v; hit.prim = (int32_t) tr.prim; return; }
					filt[*n_filt] = tr.prim; filt_mat[(*n_filt)++] = tr.flags >> Y4_TRI_MAT_SHIFT;
				}
			}
		}
		if((ngroup.y & 0xff000000u) == 0u)
		{
			if(sp == 0) return;
			ngroup = stack[--sp];
		}
	}
}
