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
	const uint32_t *consts; // device: [0] = 0x4B000000 (see y4_byte_m); unused on the host
	// Stack budget of k_trace: at most one pending sibling group per tree level sits on the stack (<= levels - 1), so
	// postponed triangle groups may only be pushed while sp < Y4_STACK_SIZE - levels.  Then T <= Y4_STACK_SIZE - levels
	// postponed groups + G <= levels - 1 sibling groups never exceed the array (y4_build.cu refuses levels > Y4_STACK_SIZE - 8).
	int postpone_sp_limit;
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
// v << (s mod 32): SHF.L.W wraps the shift amount in hardware
Y4_HD uint32_t y4_shl_wrap(uint32_t v, uint32_t s)
{
#if defined(__CUDA_ARCH__)
	return __funnelshift_l(0u, v, s);
#else
	return v << (s & 31u);
#endif
}

// Quantised byte i of w -> float WITHOUT an int->float conversion (I2F is quarter-rate): one
// PRMT drops the byte into mantissa bits 8..15 of 2^23, giving f = 2^23 + 256*q exactly.  The
// slab plane then is t = q*a + b = f*(a/256) + (b - 2^15*a): one FFMA with pre-folded per-axis
// coefficients (y4_node_hits).  The folded offset is rounded once (error <= |a|/512, i.e. 1/512
// of a quantisation step), which the coefficients absorb conservatively.
// `magic` must hold 0x4B000000 in a REGISTER so that the PRMT selector can be the immediate: ptxas
// folds any compile-time constant into the instruction's only immediate slot and then spends an
// extra IMAD per PRMT materialising selectors.  The kernels therefore load it from memory once
// (Y4Bvh8::consts[0]); it is a scheduling aid, not data.
Y4_HD float y4_byte_m(uint32_t w, int i, uint32_t magic)
{
#if defined(__CUDA_ARCH__)
	return __uint_as_float(__byte_perm(w, magic, 0x7604u + ((uint32_t) i << 4)));
#else
	return y4_u2f(magic | (((w >> (8 * i)) & 0xffu) << 8));
#endif
}
// Per-ray constants of the node test.
struct Y4RayCtx
{
	V3 o, d, idir;
	uint32_t oct_inv;    // bit2/1/0 set where d.x/y/z >= 0
};
Y4_HD Y4RayCtx y4_ray_ctx(const V3 o, const V3 d)
{
	Y4RayCtx c; c.o = o; c.d = d;
	c.idir = v3(y4_safe_rcp(d.x), y4_safe_rcp(d.y), y4_safe_rcp(d.z));
	c.oct_inv = (d.x >= 0.f ? 4u : 0u) | (d.y >= 0.f ? 2u : 0u) | (d.z >= 0.f ? 1u : 0u);
	return c;
}

// Folded slab coefficients of one axis: t_near = f*an + bn (biased early by 1/256 step),
// t_far = f*af + bf (scaled by 1 + 2^-21 and biased late), f = y4_byte_m(...).
struct Y4Axis { float an, bn, af, bf; };
// pad: extra half-width of every child box in this axis' units (two-level traversal: the object-space ray carries the
// rounding of its transform, see y4_inst_enter); 0 for world-space traversal
template<bool PAD = false>
Y4_HD Y4Axis y4_axis(float scale, float origin, float o, float idir, float pad = 0.f)
{
	const float a = scale * idir;                 // t per quantisation step
	const float b = (origin - o) * idir;          // t of the node frame origin
	const float a8 = a * (1.f / 256.f);
	const float slack = PAD ? fabsf(a8) + pad * fabsf(idir) : fabsf(a8);   // 1/256 step, covers the folding error (<= |a|/512)
	const float k_far = 1.0000005f;
	Y4Axis r;
	r.an = a8; r.bn = (b - 32768.f * a) - slack;
	r.af = a8 * k_far; r.bf = (b - 32768.f * a) * k_far + slack * 1.5f;
	return r;
}

// Intersect one node's 8 quantised child boxes; returns the hit mask in "priority" bit
// positions: internal children at bits 24 + (slot ^ oct_inv), leaf triangles at their
// tri_base-relative offsets.
template<bool PAD = false>
Y4_HD uint32_t y4_node_hits(const Y4U4 w0, const Y4U4 w1, const Y4U4 w2, const Y4U4 w3, const Y4U4 w4,
                            const V3 o, const V3 idir, const uint32_t oct_inv, const float t_lo, const float t_hi, const uint32_t magic,
                            const float pad = 0.f)
{
	const Y4Axis X = y4_axis<PAD>(y4_u2f((w0.w & 0xffu) << 23), y4_u2f(w0.x), o.x, idir.x, pad);
	const Y4Axis Y = y4_axis<PAD>(y4_u2f(((w0.w >> 8) & 0xffu) << 23), y4_u2f(w0.y), o.y, idir.y, pad);
	const Y4Axis Z = y4_axis<PAD>(y4_u2f(((w0.w >> 16) & 0xffu) << 23), y4_u2f(w0.z), o.z, idir.z, pad);
	const bool nx = idir.x < 0.f, ny = idir.y < 0.f, nz = idir.z < 0.f;
	uint32_t hits = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
	for(int h = 0; h < 2; ++h)
	{
		uint32_t meta4 = h ? w1.w : w1.z;
		// internal children carry 0b001_11sss: XOR their slot with the ray octant, 4 bytes at a time
		meta4 ^= (((meta4 & (meta4 >> 1)) >> 3) & 0x01010101u) * oct_inv;
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
			const float t0x = fmaf(y4_byte_m(nearx, i, magic), X.an, X.bn), t1x = fmaf(y4_byte_m(farx, i, magic), X.af, X.bf);
			const float t0y = fmaf(y4_byte_m(neary, i, magic), Y.an, Y.bn), t1y = fmaf(y4_byte_m(fary, i, magic), Y.af, Y.bf);
			const float t0z = fmaf(y4_byte_m(nearz, i, magic), Z.an, Z.bn), t1z = fmaf(y4_byte_m(farz, i, magic), Z.af, Z.bf);
			const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, t_lo));
			const float tf = fminf(fminf(t1x, t1y), fminf(t1z, t_hi));
			if(tn <= tf)
			{
				const uint32_t m = y4_byte(meta4, i);     // empty slot: m == 0 contributes nothing
				hits |= y4_shl_wrap(m >> 5, m);
			}
		}
	}
	return hits;
}

// Interval rule.  Closest: tmin <= t < best (accelerator_kdtree.cc:731); among *exactly* equal t the
// lower primitive index wins, which makes the result independent of traversal order (the
// reference keeps whichever its kd-leaf order tests first: the documented near-tie class).
// Shadow / transparent shadow: t_lo <= t < t_max (:867, :1019).
template<int KIND>
Y4_HD bool y4_accept_t(float t, float t_lo, float t_hi, uint32_t prim, int32_t best_prim)
{
	if(!(t >= t_lo)) return false;
	if(t < t_hi) return true;
	return KIND == Y4_TRACE_CLOSEST && t == t_hi && best_prim >= 0 && (int32_t) prim < best_prim;
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
#if defined(__CUDA_ARCH__)
	const uint32_t magic = __ldg(bvh.consts);
#else
	const uint32_t magic = 0x4B000000u;
#endif
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
			const uint32_t hits = y4_node_hits(w0, w1, w2, w3, w4, ray.o, idir, oct_inv, ray.t_lo, t_hi, magic);
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
			if(!(tr.flags & vis_bit)) continue;
			if(!y4_accept_t<KIND>(t, ray.t_lo, t_hi, tr.prim, hit.prim)) continue;
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

// ==== two-level traversal: instances of shared base meshes ===============================================
// The reference keeps ONE copy of a base mesh and gives every instance a matrix; a ray is tested against the instance's
// triangles in WORLD space: the vertices are transformed per test (FacePrimitive::getVertex(k, obj_to_world) = Matrix4 * Point3,
// primitive_instance.cc:43-51, matrix4.h:85-90).  Here: a top-level BVH8 over the instances' world boxes, one bottom-level
// BVH8 per base mesh in OBJECT space.  Inside an instance the node boxes are tested with the ray taken to object space (an
// affine map keeps the ray parameter t), padded by a bound of that transform's rounding, so the walk is conservative; at the
// leaves the triangle's object-space vertices are taken to world space with the reference's own arithmetic and tested against
// the WORLD ray with the exact test of y4_tri.cuh.  Every (ray, triangle) decision is therefore the flattened scene's,
// bit for bit, and so is the result (closest hit with the lowest-primitive tie rule; any hit: same occlusion).
struct Y4Inst             // 128 bytes
{
	float m[12];          // object -> world: rows 0..2 of the reference's Matrix4
	float mi[12];         // world -> object (inverse of the affine map, rounded from double)
	uint32_t node_off, tri_off;      // first BVH8 node / first leaf record of the instance's mesh in the pooled arrays
	uint32_t first_prim;             // global primitive index of the instance's triangle 0 (instances in upload order)
	uint32_t identity;               // 1: the mesh already is in world space (a plain, un-instanced object)
	float bmax;                      // largest |coordinate| of the mesh's object-space bound
	uint32_t n_tri, mesh;
	uint32_t vert_first;             // first triangle of the mesh in the concatenated mesh arrays
};
// leaf record of a bottom-level tree: the three object-space vertices, the triangle's index in its mesh, its flag word
struct Y4TriObj { float v[9]; uint32_t prim_local, flags, pad0; };      // 48 bytes, same size as Y4Tri

struct Y4Scene2
{
	const Y4U4 *tlas_nodes; const Y4Tri *tlas_leaves;     // leaf record .prim = instance index
	const Y4Inst *inst; uint32_t n_inst;
	const Y4U4 *blas_nodes; const Y4TriObj *blas_tris;
	const uint32_t *consts;
};

// Matrix4 * Point3 with the reference's association: ((m0 x + m1 y) + m2 z) + m3
Y4_HD V3 y4_inst_point(const float *m, const V3 p)
{
	return v3(xadd(xadd(xadd(xmul(m[0], p.x), xmul(m[1], p.y)), xmul(m[2], p.z)), m[3]),
	          xadd(xadd(xadd(xmul(m[4], p.x), xmul(m[5], p.y)), xmul(m[6], p.z)), m[7]),
	          xadd(xadd(xadd(xmul(m[8], p.x), xmul(m[9], p.y)), xmul(m[10], p.z)), m[11]));
}
Y4_HD Y4Inst y4_ld_inst(const Y4Inst *p)
{
#if defined(__CUDA_ARCH__)
	Y4Inst r;
	const float4 *q = reinterpret_cast<const float4 *>(p);
	float4 *d = reinterpret_cast<float4 *>(&r);
#pragma unroll
	for(int k = 0; k < 8; ++k) d[k] = __ldg(q + k);
	return r;
#else
	return *p;
#endif
}
Y4_HD Y4TriObj y4_ld_triobj(const Y4TriObj *p)
{
#if defined(__CUDA_ARCH__)
	Y4TriObj r;
	const float4 *q = reinterpret_cast<const float4 *>(p);
	float4 *d = reinterpret_cast<float4 *>(&r);
	d[0] = __ldg(q); d[1] = __ldg(q + 1); d[2] = __ldg(q + 2);
	return r;
#else
	return *p;
#endif
}
// The ray in the instance's object space and the padding that makes its node tests conservative.  A point of the ray inside
// the mesh's bound is o' + t d' with |o' + t d'| <= bmax; each of o', t d' and the products forming them carries a relative
// rounding of 2^-24, so the absolute error of that point is below ~8 x 2^-24 x (sum of the magnitudes involved).  The pad is
// 2^-19 x (largest row sum of |mi| |o| + |translation| ... + bmax): a factor 4 of margin.
Y4_HD void y4_inst_enter(const Y4Inst &in, const V3 o, const V3 d, V3 &oo, V3 &od, float &pad)
{
	if(in.identity) { oo = o; od = d; pad = 0.f; return; }
	const float *q = in.mi;
	oo = v3(q[0] * o.x + q[1] * o.y + q[2] * o.z + q[3], q[4] * o.x + q[5] * o.y + q[6] * o.z + q[7], q[8] * o.x + q[9] * o.y + q[10] * o.z + q[11]);
	od = v3(q[0] * d.x + q[1] * d.y + q[2] * d.z, q[4] * d.x + q[5] * d.y + q[6] * d.z, q[8] * d.x + q[9] * d.y + q[10] * d.z);
	const float ax = fabsf(o.x), ay = fabsf(o.y), az = fabsf(o.z);
	const float mx = fabsf(q[0]) * ax + fabsf(q[1]) * ay + fabsf(q[2]) * az + fabsf(q[3]);
	const float my = fabsf(q[4]) * ax + fabsf(q[5]) * ay + fabsf(q[6]) * az + fabsf(q[7]);
	const float mz = fabsf(q[8]) * ax + fabsf(q[9]) * ay + fabsf(q[10]) * az + fabsf(q[11]);
	pad = (fmaxf(mx, fmaxf(my, mz)) + in.bmax) * (1.f / 524288.f);
}
// exact test of one instanced triangle against the WORLD ray; prim_local -> global primitive index by the caller
Y4_HD bool y4_inst_tri(const Y4Inst &in, const Y4TriObj &tr, const V3 o, const V3 d, float &t, float &u, float &v)
{
	V3 a = v3(tr.v[0], tr.v[1], tr.v[2]), b = v3(tr.v[3], tr.v[4], tr.v[5]), c = v3(tr.v[6], tr.v[7], tr.v[8]);
	if(!in.identity) { a = y4_inst_point(in.m, a); b = y4_inst_point(in.m, b); c = y4_inst_point(in.m, c); }
	const V3 e1 = xsub3(b, a), e2 = xsub3(c, a);
	const float l1 = xlen(e1), l2 = xlen(e2);
	const float eps = xmul(Y4_TRI_EPS_SCALE, (l1 < l2) ? l2 : l1);         // as y4_tri_setup
	return y4_tri_intersect(o, d, a, e1, e2, eps, t, u, v);
}

#define Y4_INST_NONE 0xffffffffu
// One ray through the two-level structure.  Stack entries: node groups (top byte of .y non-zero), pending leaf groups (top byte
// zero) and the marker {0xffffffff, 0} under which the top-level state waits while an instance is walked.
// filt / filt_mat / n_filt / max_depth: the transparent-shadow query with filter colours (as y4_trace_ray); without them a
// transparent-shadow ray ends at its first blocker.
template<int KIND>
Y4_HD void y4_trace_ray2(const Y4Scene2 &sc, const Y4PreparedRay &ray, Y4Hit &hit, uint2 *stack, Y4TraceStats *stats,
                         uint32_t *filt = nullptr, uint32_t *filt_mat = nullptr, int *n_filt = nullptr, int max_depth = 0)
{
	hit.t = 0.f; hit.bu = 0.f; hit.bv = 0.f; hit.prim = -1;
	if(sc.n_inst == 0 || !(ray.t_hi > ray.t_lo)) return;
	const uint32_t vis_bit = KIND == Y4_TRACE_CLOSEST ? Y4_TRI_VIS_CLOSEST : (KIND == Y4_TRACE_SHADOW ? Y4_TRI_VIS_SHADOW : Y4_TRI_VIS_TSHADOW);
#if defined(__CUDA_ARCH__)
	const uint32_t magic = __ldg(sc.consts);
#else
	const uint32_t magic = 0x4B000000u;
#endif
	const Y4RayCtx wc = y4_ray_ctx(ray.o, ray.d);
	Y4RayCtx rc = wc;
	float pad = 0.f, t_hi = ray.t_hi;
	uint32_t cur = Y4_INST_NONE;
	Y4Inst in;
	const Y4U4 *nodes = sc.tlas_nodes;
	int sp = 0;
	uint2 ngroup; ngroup.x = 0u; ngroup.y = (1u << (24u + rc.oct_inv)) | 1u;
	while(true)
	{
		uint2 tgroup; tgroup.x = 0u; tgroup.y = 0u;
		if(ngroup.y & 0xff000000u)
		{
			const uint32_t bit = y4_msb(ngroup.y);
			ngroup.y &= ~(1u << bit);
			const uint32_t slot = (bit - 24u) ^ rc.oct_inv;
			const uint32_t child = ngroup.x + y4_popc(ngroup.y & 0xffu & ((1u << slot) - 1u));
			if(ngroup.y & 0xff000000u) stack[sp++] = ngroup;
			const Y4U4 *n = nodes + (size_t) child * Y4_NODE_WORDS;
			const Y4U4 w0 = y4_ld_u4(n), w1 = y4_ld_u4(n + 1), w2 = y4_ld_u4(n + 2), w3 = y4_ld_u4(n + 3), w4 = y4_ld_u4(n + 4);
			if(stats) stats->nodes++;
			const uint32_t hits = y4_node_hits<true>(w0, w1, w2, w3, w4, rc.o, rc.idir, rc.oct_inv, ray.t_lo, t_hi, magic, pad);
			ngroup.x = w1.x; ngroup.y = (hits & 0xff000000u) | (w0.w >> 24);
			tgroup.x = w1.y; tgroup.y = hits & 0x00ffffffu;
		}
		else
		{
			tgroup = ngroup;            // a pending leaf group came off the stack
			ngroup.y = 0u;
		}
		if(cur == Y4_INST_NONE)
		{
			if(tgroup.y)
			{
				// top level: the leaf entries are instances; walk the first one now, the rest wait on the stack
				const uint32_t bit = y4_msb(tgroup.y);
				tgroup.y &= ~(1u << bit);
				const uint32_t id = y4_ld_tri(sc.tlas_leaves + tgroup.x + bit).prim;
				if(ngroup.y & 0xff000000u) stack[sp++] = ngroup;
				if(tgroup.y) stack[sp++] = tgroup;
				uint2 mark; mark.x = 0xffffffffu; mark.y = 0u;
				stack[sp++] = mark;
				in = y4_ld_inst(sc.inst + id);
				cur = id;
				V3 oo, od;
				y4_inst_enter(in, ray.o, ray.d, oo, od, pad);
				rc = y4_ray_ctx(oo, od);
				nodes = sc.blas_nodes + (size_t) in.node_off * Y4_NODE_WORDS;
				ngroup.x = 0u; ngroup.y = (1u << (24u + rc.oct_inv)) | 1u;
				continue;
			}
		}
		else
		{
			bool done = false;
			while(tgroup.y)
			{
				const uint32_t bit = y4_msb(tgroup.y);
				tgroup.y &= ~(1u << bit);
				const Y4TriObj tr = y4_ld_triobj(sc.blas_tris + in.tri_off + tgroup.x + bit);
				if(stats) stats->tris++;
				float t, u, v;
				if(!y4_inst_tri(in, tr, ray.o, ray.d, t, u, v)) continue;
				if(!(tr.flags & vis_bit)) continue;
				const uint32_t prim = in.first_prim + tr.prim_local;
				if(!y4_accept_t<KIND>(t, ray.t_lo, t_hi, prim, hit.prim)) continue;
				if(KIND == Y4_TRACE_TSHADOW && filt && (tr.flags & Y4_TRI_TRANSPARENT))
				{
					// a transparent blocker: each distinct primitive filters once, more than max_depth of them shadow the point
					bool seen = false;
					for(int k = 0; k < *n_filt; ++k) seen = seen || (filt[k] == prim);
					if(seen) continue;
					if(*n_filt < max_depth) { filt[*n_filt] = prim; filt_mat[(*n_filt)++] = tr.flags >> Y4_TRI_MAT_SHIFT; continue; }
				}
				hit.t = t; hit.bu = u; hit.bv = v; hit.prim = (int32_t) prim;
				if(KIND == Y4_TRACE_CLOSEST) t_hi = t;
				else { done = true; break; }
			}
			if(done) return;
		}
		if((ngroup.y & 0xff000000u) == 0u)
		{
			while(true)
			{
				if(sp == 0) return;
				ngroup = stack[--sp];
				if(ngroup.x != 0xffffffffu || ngroup.y != 0u) break;
				cur = Y4_INST_NONE; rc = wc; pad = 0.f; nodes = sc.tlas_nodes;      // the marker: back at the top level
			}
		}
	}
}
