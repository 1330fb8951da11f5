// Triangle record setup and the ray/triangle test.
//
// The test restates the reference's Moller-Trumbore variant operation for operation
// (/root/reference/src/geometry/primitive/primitive_triangle.cc:44-71) with
// never-contracted fp32 arithmetic, so hit/miss decisions, t and barycentrics are
// bit-identical to the reference's CPU path for every (ray, triangle) pair.
// ("Watertight" per-edge tests would change which rays hit along shared edges and
// break hit-index parity with the reference, so they are deliberately not used.)
#pragma once
#include "y4_common.cuh"

// 0.1f * min_raydist_global, evaluated in fp32 like the reference expression
// `0.1f * min_raydist_global * std::max(...)` (left to right).
#define Y4_TRI_EPS_SCALE (0.1f * 0.00005f)

Y4_HD Y4Tri y4_tri_setup(const float *v /*9 floats*/, uint32_t prim, uint32_t flags)
{
	const V3 a = v3(v[0], v[1], v[2]), b = v3(v[3], v[4], v[5]), c = v3(v[6], v[7], v[8]);
	const V3 e1 = xsub3(b, a), e2 = xsub3(c, a);
	const float l1 = xlen(e1), l2 = xlen(e2);
	Y4Tri t;
	t.v0x = a.x; t.v0y = a.y; t.v0z = a.z;
	t.eps = xmul(Y4_TRI_EPS_SCALE, (l1 < l2) ? l2 : l1);   // std::max(l1, l2)
	t.e1x = e1.x; t.e1y = e1.y; t.e1z = e1.z; t.prim = prim;
	t.e2x = e2.x; t.e2y = e2.y; t.e2z = e2.z; t.flags = flags;
	return t;
}

// Returns true on a geometric hit; t/u/v are the reference's t_hit_, u (=barycentric_v_),
// v (=barycentric_w_).  Range checks against the ray interval are the caller's job.
Y4_HD bool y4_tri_intersect(const V3 from, const V3 dir,
                            const V3 v0, const V3 e1, const V3 e2, const float eps,
                            float &t_out, float &u_out, float &v_out)
{
	const V3 pvec = xcross(dir, e2);
	const float det = xdot(e1, pvec);
	if(det > -eps && det < eps) return false;
	const float inv_det = xdiv(1.f, det);
	const V3 tvec = xsub3(from, v0);
	const float u = xmul(xdot(tvec, pvec), inv_det);
	if(u < 0.f || u > 1.f) return false;
	const V3 qvec = xcross(tvec, e1);
	const float v = xmul(xdot(dir, qvec), inv_det);
	if((v < 0.f) || (xadd(u, v) > 1.f)) return false;
	const float t = xmul(xdot(e2, qvec), inv_det);
	if(t < eps) return false;
	t_out = t; u_out = u; v_out = v;
	return true;
}
