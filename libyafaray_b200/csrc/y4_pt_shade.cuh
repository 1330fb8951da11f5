// Per-vertex shading logic of the wavefront path tracer: surface set-up, the supported
// materials (shinydiffusemat without mirror/transparency/translucency/nodes, light_mat), the
// supported lights (pointlight, arealight, directional, sunlight, spherelight, spotlight) and one direct-light sample with MIS.  Restated from
// the reference operation for operation (never-contracted fp32, see y4_common.cuh):
//   TrianglePrimitive::getSurface              src/geometry/primitive/primitive_triangle.cc:97-176
//   ShinyDiffuseMaterial::sample/eval/pdf/emit src/material/material_shiny_diffuse.cc:197-385
//   ShinyDiffuseMaterial::orenNayar            src/material/material_shiny_diffuse.cc:161-195
//   LightMaterial::sample/emit                 src/material/material_simple.cc:41-58
//   PointLight::illuminate                     src/light/light_point.cc:38-57
//   AreaLight::illumSample/intersect           src/light/light_area.cc:66-96,117-151
//   DirectionalLight::illuminate               src/light/light_directional.cc:58-80
//   SunLight::illumSample/intersect            src/light/light_sun.cc:54-78
//   SphereLight::illumSample/sphereIntersect   src/light/light_sphere.cc:55-98
//   MonteCarloIntegrator::diracLight / areaLightSampleLight / areaLightSampleMaterial
//                                              src/integrator/surface/integrator_montecarlo.cc:80-383
#pragma once
#include "y4_pt_types.cuh"
#include "y4_tri.cuh"
#include "y4_libm.cuh"
#include "y4_pt_tex.cuh"

// ---- surface ---------------------------------------------------------------------------------
// Per-triangle shading record, computed once per scene: the face normal of calculateNormal
// (primitive_triangle.cc:92-95; fp32-deterministic, so hoisting it changes no bits) and the material index.
Y4_HD float4 y4_tri_shade(const float *verts, const uint32_t *tri_flags, uint32_t n_mats, uint64_t prim)
{
	const float *v = verts + 9ull * prim;
	const V3 a = v3(v[0], v[1], v[2]), b = v3(v[3], v[4], v[5]), c = v3(v[6], v[7], v[8]);
	const V3 ng = xnormalize(xcross(xsub3(b, a), xsub3(c, a)));
	uint32_t mat = tri_flags ? (tri_flags[prim] >> Y4_TRI_MAT_SHIFT) : 0u;
	if(mat >= n_mats) mat = 0u;
	float4 r; r.x = ng.x; r.y = ng.y; r.z = ng.z; r.w = y4_u2f(mat);
	return r;
}

Y4_HD Y4Surf y4_surface(const Y4PtScene &sc, const V3 ray_o, const V3 ray_d, float t, float bu, float bv, int32_t prim)
{
	Y4Surf s;
	const float4 ts = sc.tri_shade[(uint32_t) prim];
	s.ng = v3(ts.x, ts.y, ts.z);
	s.mat = y4_f2u(ts.w);
	const float *nn = sc.normals ? sc.normals + 9ull * (uint32_t) prim : nullptr;
	if(nn && nn[0] == nn[0])        // NaN: a flat-shaded mesh next to smooth ones
	{
		const float u = xsub(xsub(1.f, bu), bv);                      // barycentric_u_ = 1 - u - v (:66)
		const V3 n0 = v3(nn[0], nn[1], nn[2]), n1 = v3(nn[3], nn[4], nn[5]), n2 = v3(nn[6], nn[7], nn[8]);
		s.n = xnormalize(xadd3(xadd3(xscale(n0, u), xscale(n1, bu)), xscale(n2, bv)));
	}
	else s.n = s.ng;
	s.p = xadd3(ray_o, xscale(ray_d, t));                            // ray.from_ + t * ray.dir_ (accelerator.cc:62)
	y4_coords_system(s.n, s.nu, s.nv);
	return s;
}
Y4_HD V3 y4_face_forward(const V3 ng, const V3 n, const V3 incoming) { return xdot(ng, incoming) < 0.f ? xneg(n) : n; }

// ---- materials ---------------------------------------------------------------------------------
Y4_HD float y4_oren_nayar(const Y4PtMaterial &m, const V3 wi, const V3 wo, const V3 n)
{
	const float cos_ti = fmaxf(-1.f, fminf(1.f, xdot(n, wi)));
	const float cos_to = fmaxf(-1.f, fminf(1.f, xdot(n, wo)));
	float maxcos_f = 0.f;
	if(cos_ti < 0.9999f && cos_to < 0.9999f)
	{
		const V3 v1 = xnormalize(xsub3(wi, xscale(n, cos_ti)));
		const V3 v2 = xnormalize(xsub3(wo, xscale(n, cos_to)));
		maxcos_f = fmaxf(0.f, xdot(v1, v2));
	}
	float sin_alpha, tan_beta;
	if(cos_to >= cos_ti)
	{
		sin_alpha = xsqrt(xsub(1.f, xmul(cos_ti, cos_ti)));
		tan_beta = xdiv(xsqrt(xsub(1.f, xmul(cos_to, cos_to))), (cos_to == 0.f) ? 1e-8f : cos_to);
	}
	else
	{
		sin_alpha = xsqrt(xsub(1.f, xmul(cos_to, cos_to)));
		tan_beta = xdiv(xsqrt(xsub(1.f, xmul(cos_ti, cos_ti))), (cos_ti == 0.f) ? 1e-8f : cos_ti);
	}
	return fminf(1.f, fmaxf(0.f, xadd(m.oren_a, xmul(xmul(xmul(m.oren_b, maxcos_f), sin_alpha), tan_beta))));
}

// ShinyDiffuseMaterial::getFresnelKr (material_shiny_diffuse.cc:104-118)
Y4_HD float y4_fresnel_kr(const Y4PtMaterial &m, const V3 wo, const V3 n)
{
	if(!m.fresnel) return 1.f;
	const V3 nn = xdot(wo, n) < 0.f ? xneg(n) : n;
	const float c = xdot(wo, nn);
	float g = xsub(xadd(m.ior_squared, xmul(c, c)), 1.f);
	g = g < 0.f ? 0.f : xsqrt(g);
	const float aux = xmul(c, xadd(g, c));
	const float gmc = xsub(g, c), gpc = xadd(g, c), am = xsub(aux, 1.f), ap = xadd(aux, 1.f);
	return xmul(xdiv(xmul(xmul(0.5f, gmc), gmc), xmul(gpc, gpc)), xadd(1.f, xdiv(xmul(am, am), xmul(ap, ap))));
}

// ShinyDiffuseMaterial::accumulate (:123-134) for components {mirror, transparency, translucency, diffuse}
Y4_HD void y4_shiny_accumulate(const Y4PtMaterial &m, const V3 wo, const V3 n, float &accum0, float &accum1, float &accum2, float &accum3)
{
	const float kr = y4_fresnel_kr(m, wo, n);
	const float c1 = m.is_transparent ? m.transparency_strength : 0.f, c2 = m.is_translucent ? m.translucency_strength : 0.f;
	accum0 = xmul(m.is_mirror ? m.mirror_strength : 0.f, kr);
	float acc = xsub(1.f, accum0);
	accum1 = xmul(c1, acc);
	acc = xmul(acc, xsub(1.f, c1));
	accum2 = xmul(c2, acc);
	acc = xmul(acc, xsub(1.f, c2));
	accum3 = xmul(m.is_diffuse ? m.diffuse_strength : 0.f, acc);
}
// transmit_filter_strength_ * diffuse colour + (1 - transmit_filter_strength_): the tint of light passing straight through
Y4_HD C3 y4_shiny_tcol(const Y4PtMaterial &m)
{
	const float k = xsub(1.f, m.transmit_filter);
	return c3(xadd(xmul(m.transmit_filter, m.color.r), k), xadd(xmul(m.transmit_filter, m.color.g), k), xadd(xmul(m.transmit_filter, m.color.b), k));
}

// Vec3::reflectDir (vector.h:255-260)
Y4_HD V3 y4_reflect_dir(const V3 n, const V3 v)
{
	const float vn = xdot(v, n);
	if(vn < 0.f) return xneg(v);
	const float k = xmul(2.f, vn);
	return v3(xsub(xmul(k, n.x), v.x), xsub(xmul(k, n.y), v.y), xsub(xmul(k, n.z), v.z));
}

// ---- glass (src/material/material_glass.cc) ------------------------------------------------------------------
// Vec3::refract (src/geometry/vector.cc:64-86)
Y4_HD bool y4_refract(const V3 n, const V3 wi, V3 &wo, float ior)
{
	V3 nn = n;
	float eta = ior;
	const V3 i = xneg(wi);
	float cos_v_n = xdot(wi, n);
	if(cos_v_n < 0.f) { nn = xneg(n); cos_v_n = -cos_v_n; }
	else eta = xdiv(1.f, ior);
	const float k = xsub(1.f, xmul(xmul(eta, eta), xsub(1.f, xmul(cos_v_n, cos_v_n))));
	if(k <= 0.f) return false;
	const float f = xsub(xmul(eta, cos_v_n), xsqrt(k));
	wo = xnormalize(xadd3(xscale(i, eta), xscale(nn, f)));
	return true;
}
// Vec3::fresnel (vector.cc:88-115)
Y4_HD void y4_fresnel(const V3 i, const V3 n, float ior, float &kr, float &kt)
{
	const V3 nn = xdot(i, n) < 0.f ? xneg(n) : n;
	const float c = xdot(i, nn);
	float g = xsub(xadd(xmul(ior, ior), xmul(c, c)), 1.f);
	g = g <= 0.f ? 0.f : xsqrt(g);
	const float aux = xmul(c, xadd(g, c));
	const float gmc = xsub(g, c), gpc = xadd(g, c), am = xsub(aux, 1.f), ap = xadd(aux, 1.f);
	kr = xmul(xdiv(xmul(xmul(0.5f, gmc), gmc), xmul(gpc, gpc)), xadd(1.f, xdiv(xmul(am, am), xmul(ap, ap))));
	kt = kr < 1.f ? xsub(1.f, kr) : 0.f;
}
// the shading normal glass works with (GlassMaterial::sample :62-66 with a double literal, getSpecular :217-226 with a float one)
Y4_HD V3 y4_glass_normal(const Y4Surf &sp, const V3 wo, bool double_literal)
{
	const bool outside = xdot(sp.ng, wo) > 0.f;
	const float cos_wo_n = xdot(sp.n, wo);
	if(outside ? (cos_wo_n >= 0.f) : (cos_wo_n <= 0.f)) return sp.n;
	const float k = double_literal ? (float) xdmul(1.00001, (double) cos_wo_n) : xmul(1.00001f, cos_wo_n);
	return xnormalize(xsub3(sp.n, xscale(wo, k)));
}
// Vec3::reflect (vector.h:227-232) applied to wo
Y4_HD V3 y4_reflect(const V3 v, const V3 n)
{
	const float vn = xmul(2.f, xadd(xadd(xmul(v.x, n.x), xmul(v.y, n.y)), xmul(v.z, n.z)));
	return v3(xsub(xmul(vn, n.x), v.x), xsub(xmul(vn, n.y), v.y), xsub(xmul(vn, n.z), v.z));
}
// GlassMaterial::getSpecular (:215-280; no dispersion): reflected and / or refracted continuation with its colour.
// ray_level = the level recursiveRaytrace was called with (integrate()'s level + 1).
struct Y4Specular { bool reflect, refract; V3 reflect_dir, refract_dir; C3 reflect_col, refract_col; };
Y4_HD Y4Specular y4_glass_specular(const Y4PtMaterial &m, const Y4Surf &sp, const V3 wo, int ray_level)
{
	Y4Specular s; s.reflect = false; s.refract = false;
	s.reflect_dir = s.refract_dir = v3(0.f, 0.f, 1.f); s.reflect_col = s.refract_col = c3f(0.f);
	const bool outside = xdot(sp.ng, wo) > 0.f;
	const V3 n = y4_glass_normal(sp, wo, false);
	V3 refdir;
	if(y4_refract(n, wo, refdir, m.ior))
	{
		float kr, kt;
		y4_fresnel(wo, n, m.ior, kr, kt);
		s.refract = true; s.refract_col = cscale(m.color, kt); s.refract_dir = refdir;
		if(outside || ray_level < 3) { s.reflect = true; s.reflect_dir = y4_reflect(wo, n); s.reflect_col = cscale(m.mirror_color, kr); }    // Rgb * float: {r * f, ...}
	}
	else { s.reflect = true; s.reflect_col = m.mirror_color; s.reflect_dir = y4_reflect(wo, n); }      // total inner reflection
	return s;
}

// Material::sample -> returns scolor; wi, w, pdf, sampled (false == BsdfFlags::None) are outputs.
// `w` is left untouched when nothing is sampled (the reference leaves its out-parameter alone too).
Y4_HD C3 y4_mat_sample(const Y4PtMaterial &m, const Y4Surf &sp, const V3 wo, float s1, float s2, uint32_t flags,
                       V3 &wi, float &w, float &pdf, bool &sampled, bool &specular)
{
	sampled = false; pdf = 0.f; specular = false;
	if(m.type == Y4GPU_MAT_LIGHT) { w = 0.f; return c3f(0.f); }
	if(m.type == Y4GPU_MAT_MIRROR)                         // MirrorMaterial::sample (material_glass.cc:435-441): s.flags_ is not looked at
	{
		wi = y4_reflect_dir(sp.n, wo);
		sampled = true; w = 1.f; specular = true;
		return cscale(m.color, xdiv(1.f, fabsf(xdot(sp.n, wi))));
	}
	if(m.type == Y4GPU_MAT_GLASS)                          // GlassMaterial::sample (:62-203), branch without dispersion, tm_flags_ = Specular | Transmit
	{
		if(!(flags & Y4_BSDF_SPECULAR)) return c3f(0.f);
		const V3 ng = y4_glass_normal(sp, wo, true);
		const uint32_t tm = Y4_BSDF_SPECULAR | Y4_BSDF_TRANSMIT, rm = Y4_BSDF_SPECULAR | Y4_BSDF_REFLECT;
		V3 refdir;
		pdf = 1.f;
		if(y4_refract(ng, wo, refdir, m.ior))
		{
			float kr, kt;
			y4_fresnel(wo, ng, m.ior, kr, kt);
			const float p_kr = (float) xdadd(0.01, xdmul(0.99, (double) kr)), p_kt = (float) xdadd(0.01, xdmul(0.99, (double) kt));
			if(s1 < p_kt && (flags & tm) == tm) { wi = refdir; pdf = p_kt; sampled = true; specular = true; w = 1.f; return m.color; }
			if((flags & rm) == rm) { wi = y4_reflect(wo, ng); pdf = p_kr; sampled = true; specular = true; w = 1.f; return m.mirror_color; }
		}
		else if((flags & rm) == rm) { wi = y4_reflect(wo, ng); sampled = true; specular = true; w = 1.f; return c3f(1.f); }
		pdf = 0.f;
		return c3f(0.f);
	}
	const float cos_ng_wo = xdot(sp.ng, wo);
	const V3 n = y4_face_forward(sp.ng, sp.n, wo);
	float accum0, accum1, accum2, accum3;
	y4_shiny_accumulate(m, wo, n, accum0, accum1, accum2, accum3);
	// components in c_flags_ order: perfect mirror (Specular | Reflect), straight transmission (Transmit | Filter),
	// translucency (Diffuse | Transmit), Lambert / Oren-Nayar (Diffuse | Reflect)
	const uint32_t mflags = Y4_BSDF_SPECULAR | Y4_BSDF_REFLECT, tflags = Y4_BSDF_TRANSMIT | Y4_BSDF_FILTER;
	const uint32_t lflags = Y4_BSDF_DIFFUSE | Y4_BSDF_TRANSMIT, dflags = Y4_BSDF_DIFFUSE | Y4_BSDF_REFLECT;
	// (scalars instead of the reference's val[] / width[] / choice[] arrays: same arithmetic, no local memory)
	const bool mm = m.is_mirror && (flags & mflags) == mflags, tm = m.is_transparent && (flags & tflags) == tflags;
	const bool lm = m.is_translucent && (flags & lflags) == lflags, dm = m.is_diffuse && (flags & dflags) == dflags;
	float sum = 0.f, val_m = 0.f, val_t = 0.f, val_l = 0.f;
	if(mm) { sum = xadd(sum, accum0); val_m = sum; }
	if(tm) { sum = xadd(sum, accum1); val_t = sum; }
	if(lm) { sum = xadd(sum, accum2); val_l = sum; }
	if(dm) sum = xadd(sum, accum3);
	if(!(mm || tm || lm || dm) || (double) sum < 0.00001) return c3f(1.f);
	const float inv_sum = xdiv(1.f, sum);
	const float wm = xmul(accum0, inv_sum), wt = xmul(accum1, inv_sum), wl = xmul(accum2, inv_sum), wd = xmul(accum3, inv_sum);
	val_m = xmul(val_m, inv_sum); val_t = xmul(val_t, inv_sum); val_l = xmul(val_l, inv_sum);
	// pick = first matched component with s1 <= val[i] (the last one's val is sum * inv_sum), else the last matched one
	const int last = dm ? 3 : (lm ? 2 : (tm ? 1 : 0));
	const int pick = (mm && s1 <= val_m) ? 0 : ((tm && s1 <= val_t) ? 1 : ((lm && s1 <= val_l) ? 2 : ((dm && s1 <= xmul(sum, inv_sum)) ? 3 : last)));
	const float wpick = pick == 0 ? wm : (pick == 1 ? wt : (pick == 2 ? wl : wd));
	// s_1 = (s1 - val[previous matched]) / width, or s1 / width for the first matched component
	const bool has_prev = pick == 1 ? mm : (pick == 2 ? (tm || mm) : (pick == 3 ? (lm || tm || mm) : false));
	const float prev = pick == 1 ? val_m : (pick == 2 ? (tm ? val_t : val_m) : (lm ? val_l : (tm ? val_t : val_m)));
	const float s_1 = has_prev ? xdiv(xsub(s1, prev), wpick) : xdiv(s1, wpick);
	C3 scolor = c3f(0.f);
	if(pick == 1)
	{
		wi = xneg(wo);
		scolor = cscale(y4_shiny_tcol(m), accum1);
		pdf = (double) fabsf(xdot(wi, n)) < 1e-6 ? 0.f : wpick;
		specular = true;                                             // sampled_flags_ = Transmit | Filter: counts for the caustic flag
	}
	else if(pick == 0)
	{
		wi = y4_reflect_dir(n, wo);
		pdf = wpick;
		scolor = cscale(m.mirror_color, accum0);
		scolor = cscale(scolor, xdiv(1.f, fmaxf(fabsf(xdot(sp.n, wi)), 1.0e-6f)));
		specular = true;
	}
	else if(pick == 2)
	{
		wi = y4_cos_hemisphere(xneg(n), sp.nu, sp.nv, s_1, s2);
		if(xmul(cos_ng_wo, xdot(sp.ng, wi)) < 0.f) scolor = cscale(m.color, accum2);
		pdf = xmul(fabsf(xdot(wi, n)), wpick);
	}
	else
	{
		wi = y4_cos_hemisphere(n, sp.nu, sp.nv, s_1, s2);
		if(xmul(cos_ng_wo, xdot(sp.ng, wi)) > 0.f) scolor = cscale(m.color, accum3);
		if(m.oren_nayar) scolor = cscale(scolor, y4_oren_nayar(m, wo, wi, n));
		pdf = xmul(fabsf(xdot(wi, n)), wpick);
	}
	sampled = true;
	w = xdiv(fabsf(xdot(wi, sp.n)), xadd(xmul(pdf, 0.99f), 0.01f));
	// getAlpha (:467-488): 1 - (1 - components[0] * kr) * components[1]; 1 for opaque materials
	const float alpha = m.is_transparent ? xsub(1.f, xmul(xsub(1.f, xmul(m.is_mirror ? m.mirror_strength : 0.f, y4_fresnel_kr(m, wo, n))), m.transparency_strength)) : 1.f;
	w = xadd(xmul(w, alpha), xmul(1.f, xsub(1.f, alpha)));
	return scolor;
}

Y4_HD C3 y4_mat_eval(const Y4PtMaterial &m, const Y4Surf &sp, const V3 wo, const V3 wl)
{
	if(m.type == Y4GPU_MAT_LIGHT || !(m.bsdf_flags & Y4_BSDF_DIFFUSE)) return c3f(0.f);     // Material::eval default / no Diffuse flag
	const V3 n = y4_face_forward(sp.ng, sp.n, wo);
	const float kr = y4_fresnel_kr(m, wo, n);
	const float m_t = xmul(xsub(1.f, xmul(kr, m.is_mirror ? m.mirror_strength : 0.f)), xsub(1.f, m.is_transparent ? m.transparency_strength : 0.f));
	const float c2 = m.is_translucent ? m.translucency_strength : 0.f;
	// light from the other side of the surface: the translucency component answers (and nothing else)
	if(m.is_translucent && xmul(xdot(sp.ng, wo), xdot(sp.ng, wl)) < 0.f) return cscale(m.color, xmul(c2, m_t));
	if(xdot(n, wl) < 0.f && !m.flat_material) return c3f(0.f);
	float m_d = xmul(xmul(m_t, xsub(1.f, c2)), m.is_diffuse ? m.diffuse_strength : 0.f);
	if(m.oren_nayar) m_d = xmul(m_d, y4_oren_nayar(m, wo, wl, n));
	return cscale(m.color, m_d);
}

Y4_HD float y4_mat_pdf(const Y4PtMaterial &m, const Y4Surf &sp, const V3 wo, const V3 wi)
{
	if(m.type == Y4GPU_MAT_LIGHT || !(m.bsdf_flags & Y4_BSDF_DIFFUSE)) return 0.f;
	const V3 n = y4_face_forward(sp.ng, sp.n, wo);
	// bsdfs = Glossy | Diffuse | Dispersive | Reflect | Transmit hasAny() of every component's flags: all widths count in the sum
	float accum0, accum1, accum2, width;
	y4_shiny_accumulate(m, wo, n, accum0, accum1, accum2, width);
	float sum = 0.f, pdf = 0.f;
	if(m.is_mirror) sum = xadd(sum, accum0);
	if(m.is_transparent) sum = xadd(sum, accum1);
	if(m.is_translucent)
	{
		sum = xadd(sum, accum2);
		if(xmul(xdot(sp.ng, wo), xdot(sp.ng, wi)) < 0.f) pdf = xadd(pdf, xmul(fabsf(xdot(wi, n)), accum2));
	}
	if(m.is_diffuse) { sum = xadd(sum, width); pdf = xadd(pdf, xmul(fabsf(xdot(wi, n)), width)); }
	if(sum < 0.00001f) return 0.f;
	return xdiv(pdf, sum);
}

Y4_HD C3 y4_mat_emit(const Y4PtMaterial &m, const Y4Surf &sp, const V3 wo)
{
	if(m.type == Y4GPU_MAT_LIGHT)
	{
		if(m.double_sided) return m.color;
		return xdot(wo, sp.n) > 0.f ? m.color : c3f(0.f);
	}
	return m.emit;
}

// ---- one direct-light sample -------------------------------------------------------------------
// Produces up to two shadow rays with the radiance each carries if unoccluded:
//   slot 0: the light sample (dirac light, or areaLightSampleLight's i-th sample)
//   slot 1: the BSDF sample that completes MIS for intersectable lights (areaLightSampleMaterial)
// Contributions are per-sample terms; the caller sums them over the light's samples and applies
// inv_num_samples exactly like the reference (col * inv_num_samples per sub-estimator).
struct Y4DirectOut
{
	Y4Ray ray[2];
	C3 contrib[2];
	bool valid[2];
};

Y4_HD Y4Ray y4_null_ray() { Y4Ray r; r.ox = r.oy = r.oz = 0.f; r.tmin = 0.f; r.dx = 0.f; r.dy = 0.f; r.dz = 1.f; r.tmax = 0.f; return r; }
Y4_HD Y4Ray y4_mk_ray(const V3 o, float tmin, const V3 d, float tmax)
{
	Y4Ray r; r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tmin = tmin; r.dx = d.x; r.dy = d.y; r.dz = d.z; r.tmax = tmax; return r;
}

// ---- lights ----------------------------------------------------------------------------------------
// LSample / light_ray fields the integrator reads back from Light::illuminate / illumSample
struct Y4LSample { V3 dir; float tmax, pdf; C3 col; };

// sample::cone (include/sampler/sample.h:80-86)
Y4_HD V3 y4_sample_cone(const V3 d, const V3 u, const V3 v, float max_cos_ang, float s1, float s2)
{
	const float cos_ang = xsub(1.f, xmul(xsub(1.f, max_cos_ang), s2));
	const float sin_ang = xsqrt(xsub(1.f, xmul(cos_ang, cos_ang)));
	const float t1 = (float) (Y4_2PI * (Y4_WIDE) s1);
	const V3 a = xadd3(xscale(u, y4_fast_cos(t1)), xscale(v, y4_fast_sin(t1)));
	return xadd3(xscale(a, sin_ang), xscale(d, cos_ang));
}

// SphereLight::sphereIntersect (light_sphere.cc:55-67): float members, double literals promote the discriminant and the quotients
Y4_HD bool y4_sphere_intersect(const V3 from, const V3 dir, const V3 c, float r2, float &d1, float &d2)
{
	const V3 vf = xsub3(from, c);
	const float ea = xdot(dir, dir);
	const float eb = xdot(xscale(vf, 2.f), dir);
	const float ec = xsub(xdot(vf, vf), r2);
	float osc = (float) xdsub((double) xmul(eb, eb), xdmul(xdmul(4.0, (double) ea), (double) ec));
	if(osc < 0.f) { d1 = xsqrt(xdiv(ec, ea)); return false; }
	osc = xsqrt(osc);
	const double den = xdmul(2.0, (double) ea);
	d1 = (float) xddiv((double) xsub(-eb, osc), den);
	d2 = (float) xddiv((double) xadd(-eb, osc), den);
	return true;
}

// Light::illuminate of the dirac lights: PointLight (light_point.cc:38-57), SpotLight without soft shadows
// (light_spot.cc:74-104), DirectionalLight (light_directional.cc:58-80)
Y4_HD bool y4_light_illuminate(const Y4PtLight &L, const V3 p, Y4LSample &ls)
{
	ls.pdf = 1.f;
	if(L.type == Y4GPU_LIGHT_POINT || L.type == Y4GPU_LIGHT_SPOT)
	{
		V3 ldir = xsub3(L.pos, p);
		const float dist_sqr = xadd(xadd(xmul(ldir.x, ldir.x), xmul(ldir.y, ldir.y)), xmul(ldir.z, ldir.z));
		const float dist = xsqrt(dist_sqr);
		if(dist == 0.f) return false;
		const float idist_sqr = xdiv(1.f, dist_sqr);
		ls.dir = xscale(ldir, xdiv(1.f, dist));
		ls.tmax = dist;
		if(L.type == Y4GPU_LIGHT_POINT) { ls.col = cscale(L.color, idist_sqr); return true; }
		// SpotLight::illuminate (light_spot.cc:74-104): smoothstep between cos_end_ and cos_start_
		const float cosa = xdot(L.dir, ls.dir);
		if(cosa < L.cos_end) return false;
		if(cosa >= L.cos_start) { ls.col = cscale(L.color, idist_sqr); return true; }
		float v = xmul(xsub(cosa, L.cos_end), L.icos_diff);
		v = xmul(xmul(v, v), xsub(3.f, xmul(2.f, v)));
		ls.col = cscale(cscale(L.color, v), idist_sqr);                 // color_ * v * idist_sqr
		return true;
	}
	// directional
	if(!L.infinite)
	{
		const V3 vec = xsub3(L.pos, p);
		const float dist = xlen(xcross(L.dir, vec));          // distance from the cylinder axis
		if(dist > L.radius) return false;
		ls.tmax = xdot(vec, L.dir);
		if(ls.tmax <= 0.f) return false;
	}
	else ls.tmax = -1.f;
	ls.dir = L.dir;
	ls.col = L.color;
	return true;
}

// ---- background light (src/light/light_background.cc) ------------------------------------------------------------
// ConstantBackground::eval / GradientBackground::eval (background_gradient.cc:34-46)
Y4_HD C3 y4_texture_background(const Y4BgTex &B, const V3 dir);
Y4_HD C3 y4_sunsky_background(const Y4BgSunsky &S, const V3 dir);
Y4_HD C3 y4_darksky_background(const Y4BgDarksky &S, const V3 dir, bool with_power);
Y4_HD C3 y4_bg_eval(const Y4PtLight &L, const V3 dir)
{
	if(L.bg_gradient == 4) return y4_darksky_background(*L.bg_dark, dir, true);        // lights call Background::eval
	if(L.bg_gradient == 3) return y4_sunsky_background(*L.bg_sun, dir);
	if(L.bg_gradient == 2) return y4_texture_background(L.bg_tex, dir);
	if(!L.bg_gradient) return L.bg_col;
	float blend = dir.z;
	C3 color;
	if(blend >= 0.f) color = cadd(cscale(L.bg_sz, blend), cscale(L.bg_sh, xsub(1.f, blend)));
	else
	{
		blend = -blend;
		color = cadd(cscale(L.bg_gz, blend), cscale(L.bg_gh, xsub(1.f, blend)));
	}
	if(fminf(color.r, fminf(color.g, color.b)) < 1e-6f) color = c3f(1e-5f);
	return color;
}
// Pdf1D::sample (include/sampler/sample_pdf1d.h:99-122); lower_bound = first index with cdf[index] >= u
Y4_HD float y4_pdf1d_sample(const float *func, const float *cdf, int n, float inv_integral, float u, float &pdf)
{
	if(u <= 0.f) { pdf = xmul(func[0], inv_integral); return 0.f; }
	if(u >= 1.f) { pdf = xmul(func[n - 1], inv_integral); return xadd((float) (n - 1), 1.f); }
	int lo = 0, hi = n;
	while(lo < hi) { const int mid = (lo + hi) >> 1; if(cdf[mid] < u) lo = mid + 1; else hi = mid; }
	const int index = lo < n ? lo : n - 1;      // (u above the last cdf entry would index past the table in the reference)
	pdf = xmul(func[index], inv_integral);
	const float delta = index > 0 ? xdiv(xsub(u, cdf[index - 1]), xsub(cdf[index], cdf[index - 1])) : xdiv(u, cdf[index]);
	return xadd((float) index, delta);
}
Y4_HD float y4_bg_sin_sample(float s) { return y4_fast_sin((float) xdmul((Y4_WIDE) s, Y4_PI)); }      // math::sin(s * math::num_pi)
Y4_HD float y4_bg_clamp_zero(float v) { return v > 0.f ? xdiv(1.f, v) : 0.f; }
Y4_HD int y4_bg_clamp_sample(float v, int m) { int s = (int) xadd(v, 0.4999f); if(s > m - 1) s = m - 1; return s < 0 ? 0 : s; }     // clampSample(addOff(v), m)
Y4_HD float y4_bg_pdf(float p0, float p1, float s, bool inv)
{
	if(inv) return fmaxf(0.000001f, xmul(xmul((float) Y4_2PI, y4_bg_sin_sample(s)), y4_bg_clamp_zero(xmul(p0, p1))));
	return fmaxf(0.000001f, xmul(xmul(xmul(p0, p1), (float) Y4_1_2PI), y4_bg_clamp_zero(y4_bg_sin_sample(s))));
}
// Texture::invSphereMap / sphereMap (src/texture/texture.cc:103-132); std::acos / std::asin with the C library's bits (y4_libm.cuh)
Y4_HD V3 y4_inv_sphere_map(float u, float v)
{
	const float theta = (float) xdmul((Y4_WIDE) v, Y4_PI), phi = -(float) xdmul((Y4_WIDE) u, Y4_2PI);
	const float costheta = y4_fast_cos(theta), sintheta = y4_fast_sin(theta), cosphi = y4_fast_cos(phi), sinphi = y4_fast_sin(phi);
	return v3(xmul(sintheta, cosphi), xmul(sintheta, sinphi), -costheta);
}
Y4_HD float y4_acos(float x) { return x <= -1.f ? (float) Y4_PI : (x >= 1.f ? 0.f : y4_libm_acosf(x)); }            // math::acos (math.h:253-259)
Y4_HD float y4_asin(float x) { return x <= -1.f ? (float) -Y4_PI_2 : (x >= 1.f ? (float) Y4_PI_2 : y4_libm_asinf(x)); }   // math::asin (math.h:261-267)
Y4_HD void y4_sphere_map(const V3 p, float &u, float &v)
{
	const float sqrt_r_phi = xadd(xmul(p.x, p.x), xmul(p.y, p.y));
	const float sqrt_r_theta = xadd(sqrt_r_phi, xmul(p.z, p.z));
	u = 0.f;
	if(sqrt_r_phi > 0.f)
	{
		const float a = y4_acos(xdiv(p.x, xsqrt(sqrt_r_phi)));
		const float phi_ratio = p.y < 0.f ? (float) xdmul(Y4_2PI - (Y4_WIDE) a, Y4_1_2PI) : (float) xdmul((Y4_WIDE) a, Y4_1_2PI);
		u = xsub(1.f, phi_ratio);
	}
	v = (float) (1.0 - xdmul((Y4_WIDE) y4_acos(xdiv(p.z, xsqrt(sqrt_r_theta))), Y4_1_PI));
}
// Texture::angMap (texture.cc:75-86)
Y4_HD void y4_ang_map(const V3 p, float &u, float &v)
{
	float r = xadd(xmul(p.x, p.x), xmul(p.z, p.z));
	u = v = 0.f;
	if(r > 0.f)
	{
		const float phi_ratio = (float) xdmul(Y4_1_PI, (Y4_WIDE) y4_acos(p.y));
		r = xdiv(phi_ratio, xsqrt(r));
		u = xmul(p.x, r);
		v = xmul(p.z, r);
	}
}
// TextureBackground::eval (background_texture.cc:44-78; no SmartIBL blur: refused on the host)
Y4_HD C3 y4_texture_background(const Y4BgTex &B, const V3 dir)
{
	float u = 0.f, v = 0.f;
	if(B.angular)
	{
		const V3 p = v3(xadd(xmul(dir.x, B.cos_r), xmul(dir.y, B.sin_r)), xadd(xmul(dir.x, -B.sin_r), xmul(dir.y, B.cos_r)), dir.z);
		y4_ang_map(p, u, v);
	}
	else
	{
		y4_sphere_map(dir, u, v);
		u = xsub(xmul(2.f, u), 1.f);
		v = xsub(xmul(2.f, v), 1.f);
		u = xadd(u, B.rotation);
		if(u > 1.f) u = xsub(u, 2.f);
	}
	const C4 t = y4_tex_color(*B.tex, v3(u, v, 0.f));
	C3 ret = c3(t.r, t.g, t.b);
	const float min_component = 1.0e-5f;
	if(ret.r < min_component) ret.r = min_component;
	if(ret.g < min_component) ret.g = min_component;
	if(ret.b < min_component) ret.b = min_component;
	return cscale(ret, B.power);
}
// SunSkyBackground::perezFunction (background_sunsky.cc:82-104).  The reference's math::exp / math::cos take float arguments:
// the double operands are narrowed first, the float results widened again; everything else is double arithmetic (no contraction).
// den = the direction-independent denominator (host).
Y4_HD double y4_sunsky_exp(double e) { return e <= 230. ? (double) y4_fast_exp((float) e) : 7.7220185e99; }
Y4_HD double y4_perez(const double *lam, double theta, double gamma, double lvz, double den)
{
	const double e_3 = y4_sunsky_exp(xddiv(lam[1], (double) y4_fast_cos((float) theta)));
	const double e_4 = y4_sunsky_exp(xdmul(lam[3], gamma));
	const float cg = y4_fast_cos((float) gamma);
	const double num = xdmul(xdadd(1., xdmul(lam[0], e_3)), xdadd(xdadd(1., xdmul(lam[2], e_4)), xdmul(xdmul(lam[4], (double) cg), (double) cg)));
	return xddiv(xdmul(lvz, num), den);
}
// SunSkyBackground::eval -> getSkyCol (background_sunsky.cc:115-170): Preetham / Shirley / Smits daylight model.  acos / atan2 are
// the C library's single-precision routines (y4_libm.cuh), sin / cos / exp the FAST_TRIG / FAST_MATH ones.
Y4_HD C3 y4_sunsky_background(const Y4BgSunsky &S, const V3 dir)
{
	const V3 iw = xnormalize(dir);
	double hfade = 1., nfade = 1.;
	double theta = (double) y4_acos(iw.z);
	const double half_pi = (double) (0.5 * Y4_PI);
	if(theta > half_pi)
	{
		hfade = xdsub(1.0, xdmul(xdsub(xdmul(theta, (double) Y4_1_PI), 0.5), 2.0));      // (long double in the reference: the last bit of a double may differ)
		hfade = xdmul(xdmul(hfade, hfade), xdsub(3.0, xdmul(2.0, hfade)));
		theta = half_pi;
	}
	if(S.night && theta <= half_pi)
	{
		nfade = xdsub(1.0, xdmul(xdsub(0.5, xdmul(theta, (double) Y4_1_PI)), 2.0));
		nfade = xdmul(nfade, S.nfade_sun);
		nfade = xdmul(xdmul(nfade, nfade), xdsub(3.0, xdmul(2.0, nfade)));
	}
	double phi;
	if(iw.y == 0.f && iw.x == 0.f) phi = half_pi;
	else phi = (double) y4_libm_atan2f(iw.y, iw.x);
	// angleBetween: the trigonometric terms are floats, so is their sum of products
	const float cospsi = xadd(xmul(xmul(y4_fast_sin((float) theta), S.sin_theta_s), y4_fast_cos((float) xdsub(S.phi_s, phi))), xmul(y4_fast_cos((float) theta), S.cos_theta_s));
	double gamma;
	if(cospsi > 1.f) gamma = 0.;
	else if(cospsi < -1.f) gamma = (double) Y4_PI;
	else gamma = (double) y4_acos(cospsi);
	const double x = y4_perez(S.perez_x, theta, gamma, S.zenith_x, S.den_x);
	const double y = y4_perez(S.perez_y, theta, gamma, S.zenith_y, S.den_y);
	const double Y = xdmul(xdmul(xdmul(6.666666667e-5, nfade), hfade), y4_perez(S.perez_Y, theta, gamma, S.zenith_Y, S.den_Y));
	if(y == 0.) return c3f(0.f);
	const double X = xdmul(xddiv(x, y), Y);
	const double z = xdmul(xddiv(xdsub(xdsub(1.0, x), y), y), Y);
	C3 c = c3((float) xdsub(xdsub(xdmul(3.240479, X), xdmul(1.537150, Y)), xdmul(0.498535, z)),
	          (float) xdadd(xdadd(xdmul(-0.969256, X), xdmul(1.875992, Y)), xdmul(0.041556, z)),
	          (float) xdadd(xdsub(xdmul(0.055648, X), xdmul(0.204043, Y)), xdmul(1.057311, z)));
	c.r = fmaxf(0.f, fminf(1.f, c.r)); c.g = fmaxf(0.f, fminf(1.f, c.g)); c.b = fmaxf(0.f, fminf(1.f, c.b));      // clampRgb01
	return cscale(c, S.power);
}
// DarkSkyBackground::perezFunction (background_darksky.cc:156-160): double throughout, std::exp of the double C library.  The device
// calls CUDA's exp() / acos(), whose results may differ from glibc's in the last bits of the double -- far below the float the
// result is narrowed to right after (color_conv_.fromxyY takes floats); the host derives every constant with its own libm.
Y4_HD double y4_dark_perez(const double *lam, double cos_theta, double gamma, double cos_gamma_2, double lvz)
{
	const double num = xdmul(xdadd(1., xdmul(lam[0], exp(xddiv(lam[1], cos_theta)))), xdadd(xdadd(1., xdmul(lam[2], exp(xdmul(lam[3], gamma)))), xdmul(lam[4], cos_gamma_2)));
	return xdmul(xdmul(lvz, num), lam[5]);
}
// DarkSkyBackground::eval -> getSkyCol (:162-182) -> ColorConv::fromxyY (color_conversion.h:107-168; gamma encoding and clamping are
// always on: the factory's constants).  with_power: eval() = getSkyCol * power_ (what the lights see); a ray that leaves the scene
// calls the background's operator(), which DarkSkyBackground overrides WITHOUT power_ (:184-187)
Y4_HD C3 y4_darksky_background(const Y4BgDarksky &S, const V3 dir, bool with_power)
{
	const V3 iw = xnormalize(v3(dir.x, dir.y, xadd(dir.z, S.alt)));
	double cos_theta = (double) iw.z;
	if(cos_theta <= 0.0) cos_theta = 1e-6;
	const double cos_gamma = (double) xdot(iw, S.sun_dir);
	const double cos_gamma_2 = xdmul(cos_gamma, cos_gamma);
	const double gamma = acos(cos_gamma);
	const float x = (float) y4_dark_perez(S.perez_x, cos_theta, gamma, cos_gamma_2, S.zenith_x);
	const float y = (float) y4_dark_perez(S.perez_y, cos_theta, gamma, cos_gamma_2, S.zenith_y);
	float Y = (float) xdmul(y4_dark_perez(S.perez_Y, cos_theta, gamma, cos_gamma_2, S.zenith_Y), 6.66666667e-5);
	// fromxyY2Xyz
	float X = 0.f, z = 0.f;
	if(S.exposure > 0.f) Y = xsub(y4_fast_exp(xmul(Y, S.exposure)), 1.f);
	if(y != 0.f)
	{
		const float ratio = xdiv(Y, y);
		X = xmul(x, ratio);
		z = xmul(xsub(xsub(1.f, x), y), ratio);
	}
	else Y = 0.f;
	// fromXyz with sGammaEnc = math::pow(v, 1 / 2.2f), then clampRgb01
	const float g = 1.f / 2.2f;
	C3 c = c3(y4_fast_pow(xadd(xadd(xmul(S.mat[0], X), xmul(S.mat[1], Y)), xmul(S.mat[2], z)), g),
	          y4_fast_pow(xadd(xadd(xmul(S.mat[3], X), xmul(S.mat[4], Y)), xmul(S.mat[5], z)), g),
	          y4_fast_pow(xadd(xadd(xmul(S.mat[6], X), xmul(S.mat[7], Y)), xmul(S.mat[8], z)), g));
	if(c.r < 0.f) c.r = 0.f; else if(c.r > 1.f) c.r = 1.f;
	if(c.g < 0.f) c.g = 0.f; else if(c.g > 1.f) c.g = 1.f;
	if(c.b < 0.f) c.b = 0.f; else if(c.b > 1.f) c.b = 1.f;
	if(S.night) c = cmul(c, c3(0.05f, 0.05f, 0.08f));
	c = cscale(c, S.bright);
	return with_power ? cscale(c, S.power) : c;
}
// Rgb::clampProportionalRgb (color.h:415-447): the largest component goes to max_value, the others scale with it
Y4_HD C3 y4_clamp_proportional(C3 c, float max_value)
{
	if(!(max_value > 0.f)) return c;
	const float max_rgb = fmaxf(c.r, fmaxf(c.g, c.b));
	const float adj = xdiv(max_value, max_rgb);
	if(max_rgb > max_value)
	{
		if(c.r >= max_rgb) { c.r = max_value; c.g = xmul(c.g, adj); c.b = xmul(c.b, adj); }
		else if(c.g >= max_rgb) { c.g = max_value; c.r = xmul(c.r, adj); c.b = xmul(c.b, adj); }
		else { c.b = max_value; c.r = xmul(c.r, adj); c.g = xmul(c.g, adj); }
	}
	return c;
}
// BackgroundLight::calcFromSample / calcFromDir (:77-101)
Y4_HD float y4_bg_from_sample(const Y4PtLight &L, float s1, float s2, float &u, float &v, bool inv)
{
	float pdf1 = 0.f, pdf2 = 0.f;
	const int nv = L.bg_nv, ov = L.bg_off[nv];
	v = y4_pdf1d_sample(L.bg_func + ov, L.bg_cdf + ov, nv, L.bg_inv_int[nv], s2, pdf2);
	const int iv = y4_bg_clamp_sample(v, nv);
	const int ou = L.bg_off[iv], nu = L.bg_off[iv + 1] - ou;
	u = y4_pdf1d_sample(L.bg_func + ou, L.bg_cdf + ou, nu, L.bg_inv_int[iv], s1, pdf1);
	u = xmul(u, xdiv(1.f, (float) nu));
	v = xmul(v, xdiv(1.f, (float) nv));
	return y4_bg_pdf(pdf1, pdf2, v, inv);
}
Y4_HD float y4_bg_from_dir(const Y4PtLight &L, const V3 dir, float &u, float &v, bool inv)
{
	y4_sphere_map(dir, u, v);
	const int nv = L.bg_nv;
	const int iv = y4_bg_clamp_sample(xmul(v, (float) nv), nv);
	const int ou = L.bg_off[iv], nu = L.bg_off[iv + 1] - ou;
	const int iu = y4_bg_clamp_sample(xmul(u, (float) nu), nu);
	const float pdf1 = xmul(L.bg_func[ou + iu], L.bg_inv_int[iv]);
	const float pdf2 = xmul(L.bg_func[L.bg_off[nv] + iv], L.bg_inv_int[nv]);
	return y4_bg_pdf(pdf1, pdf2, v, inv);
}

// Light::illumSample of the sampled lights: AreaLight (light_area.cc:66-96), SunLight (light_sun.cc:54-68),
// SphereLight (light_sphere.cc:69-98)
// GENERAL = false: the kernel variant for scenes without specular materials and object lights (see y4_vertex); the object
// light's code folds away there
template<bool GENERAL = true>
Y4_HD bool y4_light_illum_sample(const Y4PtLight &L, const V3 from, float s1, float s2, Y4LSample &ls)
{
	if(L.type == Y4GPU_LIGHT_AREA)
	{
		const V3 p = xadd3(xadd3(L.pos, xscale(L.to_x, s1)), xscale(L.to_y, s2));
		V3 ldir = xsub3(p, from);
		const float dist_sqr = xdot(ldir, ldir);
		const float dist = xsqrt(dist_sqr);
		if(dist <= 0.f) return false;
		ldir = xscale(ldir, xdiv(1.f, dist));
		const float cos_angle = xdot(ldir, L.fnormal);
		if(cos_angle <= 0.f) return false;
		// dist_sqr * math::num_pi / (area_ * cos_angle): num_pi is long double -> wide product and quotient, one rounding to float
		ls.pdf = (float) (xdmul((Y4_WIDE) dist_sqr, Y4_PI) / (Y4_WIDE) xmul(L.area, cos_angle));
		ls.dir = ldir; ls.tmax = dist; ls.col = L.color;
		return true;
	}
	if(GENERAL && (L.type == Y4GPU_LIGHT_OBJECT || L.type == Y4GPU_LIGHT_PORTAL))
	{
		// ObjectLight::sampleSurface / BackgroundPortalLight::sampleSurface (same code) (light_object_light.cc:87-105): Pdf1D::dSample = lower_bound on the cdf, the sample number
		// re-stretched inside the chosen triangle's cdf step, TrianglePrimitive::sample (primitive_triangle.cc:228-234)
		const int n = L.obj_n;
		int idx;
		if(s1 <= 0.f) idx = 0;
		else if(s1 >= 1.f) idx = n - 1;
		else
		{
			int lo = 0, hi = n;                               // first index with cdf[idx] >= s1
			while(lo < hi) { const int mid = (lo + hi) >> 1; if(L.obj_cdf[mid] < s1) lo = mid + 1; else hi = mid; }
			idx = lo;
		}
		if(idx >= n) return false;                            // "Sampling error": a null point and normal -> no light
		float delta = L.obj_cdf[idx], ss1;
		if(idx > 0) { const float below = L.obj_cdf[idx - 1]; delta = xsub(delta, below); ss1 = xdiv(xsub(s1, below), delta); }
		else ss1 = xdiv(s1, delta);
		const float *tv = L.obj_tris + 9ull * (uint32_t) idx;
		const float su1 = xsqrt(ss1), u = xsub(1.f, su1), v = xmul(s2, su1), w = xsub(xsub(1.f, u), v);
		const V3 p = xadd3(xadd3(xscale(v3(tv[0], tv[1], tv[2]), u), xscale(v3(tv[3], tv[4], tv[5]), v)), xscale(v3(tv[6], tv[7], tv[8]), w));
		const float4 ng4 = L.obj_ng[idx];
		const V3 nl = v3(ng4.x, ng4.y, ng4.z);
		// illumSample (:109-146)
		V3 ldir = xsub3(p, from);
		const float dist_sqr = xdot(ldir, ldir);
		const float dist = xsqrt(dist_sqr);
		if(dist <= 0.f) return false;
		ldir = xscale(ldir, xdiv(1.f, dist));
		float cos_angle = -xdot(ldir, nl);
		if(cos_angle <= 0.f)
		{
			if(L.double_sided) cos_angle = -cos_angle;
			else return false;
		}
		const float amc = xmul(L.area, cos_angle);
		if(L.type == Y4GPU_LIGHT_PORTAL)                    // BackgroundPortalLight::illumSample (light_background_portal.cc:127-157)
		{
			ls.pdf = (float) (xdmul((Y4_WIDE) dist_sqr, Y4_PI) / (Y4_WIDE) amc);
			ls.dir = ldir; ls.tmax = dist; ls.col = cscale(y4_bg_eval(L, ldir), L.power);
			return true;
		}
		ls.pdf = (float) (xdmul((Y4_WIDE) dist_sqr, Y4_PI) / (Y4_WIDE) (amc == 0.f ? 1e-8f : amc));
		ls.dir = ldir; ls.tmax = dist; ls.col = L.color;
		return true;
	}
	if(GENERAL && L.type == Y4GPU_LIGHT_BACKGROUND)       // BackgroundLight::illumSample (:127-137)
	{
		float u, v;
		ls.pdf = y4_bg_from_sample(L, s1, s2, u, v, false);
		ls.dir = y4_inv_sphere_map(u, v);
		ls.col = y4_bg_eval(L, ls.dir);
		ls.tmax = -1.f;
		return true;
	}
	if(GENERAL && L.type == Y4GPU_LIGHT_SPOT_SOFT)        // SpotLight::illumSample (light_spot.cc:106-144)
	{
		V3 ldir = xsub3(L.pos, from);
		const float dist_sqr = xdot(ldir, ldir);
		if(dist_sqr == 0.f) return false;
		const float dist = xsqrt(dist_sqr);
		ldir = xscale(ldir, xdiv(1.f, dist));
		const float cosa = xdot(L.dir, ldir);
		if(cosa < L.cos_end) return false;
		ls.tmax = dist;
		ls.dir = y4_sample_cone(ldir, L.du, L.dv, L.cos_end, xmul(s1, L.radius), xmul(s2, L.radius));
		if(cosa >= L.cos_start) ls.col = L.color;
		else
		{
			float v = xmul(xsub(cosa, L.cos_end), L.icos_diff);
			v = xmul(xmul(v, v), xsub(3.f, xmul(2.f, v)));
			ls.col = cscale(L.color, v);
		}
		ls.pdf = dist_sqr;
		if(ls.pdf < 1.f) { ls.pdf = 1.f; ls.col = cdiv(ls.col, dist_sqr); }
		return true;
	}
	if(L.type == Y4GPU_LIGHT_SUN)
	{
		ls.dir = y4_sample_cone(L.dir, L.du, L.dv, L.cos_angle, s1, s2);
		ls.tmax = -1.f; ls.col = L.color; ls.pdf = L.pdf;
		return true;
	}
	// sphere
	V3 cdir = xsub3(L.pos, from);
	const float dist_sqr = xdot(cdir, cdir);
	if(dist_sqr <= L.sq_radius) return false;                 // only emits on the outside
	const float dist = xsqrt(dist_sqr);
	const float idist_sqr = xdiv(1.f, dist_sqr);
	const float cos_alpha = xsqrt(xsub(1.f, xmul(L.sq_radius, idist_sqr)));
	cdir = xscale(cdir, xdiv(1.f, dist));
	V3 cu, cv;
	y4_coords_system(cdir, cu, cv);
	ls.dir = y4_sample_cone(cdir, cu, cv, cos_alpha, s1, s2);
	float d1, d2;
	if(!y4_sphere_intersect(from, ls.dir, L.pos, L.sq_radius_eps, d1, d2)) return false;
	ls.tmax = d1;
	ls.pdf = xdiv(1.f, xmul(2.f, xsub(1.f, cos_alpha)));
	ls.col = L.color;
	return true;
}

// Light::intersect of the lights with canIntersect(): AreaLight (light_area.cc:117-151), SunLight (light_sun.cc:70-78), ObjectLight
template<bool GENERAL = true>
Y4_HD bool y4_light_intersect(const Y4PtLight &L, const V3 from, const V3 dir, float tmin, float &t, C3 &col, float &ipdf)
{
	// BackgroundPortalLight::intersect (light_background_portal.cc:176-191) asks the light's own kd-tree, which was built AFTER
	// init() made the mesh invisible (:79-80): the tree's closest-hit query skips invisible primitives, so the call never hits and
	// the BSDF-sampling half of the MIS estimate contributes nothing (the light-sampling half keeps its MIS weight).  Reproduced.
	if(GENERAL && L.type == Y4GPU_LIGHT_PORTAL) return false;
	if(GENERAL && L.type == Y4GPU_LIGHT_OBJECT)
	{
		// ObjectLight::intersect (:183-201): closest hit of the light's own accelerator over [ray.tmin_, inf) -- here a loop over its
		// triangles with the accelerator's triangle test and tie rule.  The reference never writes its `t` argument: the caller's
		// b_ray.tmax_ keeps its initial -1, so 1 / (t * t) is 1 and the shadow ray that follows is unbounded.  Reproduced.
		int best = -1; float best_t = 0.f;
		for(int k = 0; k < L.obj_n; ++k)
		{
			const Y4Tri tr = y4_tri_setup(L.obj_tris + 9ull * (uint32_t) k, (uint32_t) k, 0u);
			float tt, uu, vv;
			if(!y4_tri_intersect(from, dir, v3(tr.v0x, tr.v0y, tr.v0z), v3(tr.e1x, tr.e1y, tr.e1z), v3(tr.e2x, tr.e2y, tr.e2z), tr.eps, tt, uu, vv)) continue;
			if(!(tt >= tmin)) continue;
			if(best < 0 || tt < best_t) { best = k; best_t = tt; }
		}
		if(best < 0) return false;
		const float4 ng4 = L.obj_ng[best];
		float cos_angle = -xdot(dir, v3(ng4.x, ng4.y, ng4.z));                // ray.dir_ * (-n)
		if(cos_angle <= 0.f)
		{
			if(L.double_sided) cos_angle = fabsf(cos_angle);
			else return false;
		}
		ipdf = (float) xdmul((Y4_WIDE) xmul(xmul(1.f, L.area), cos_angle), Y4_1_PI);
		col = L.color; t = -1.f;
		return true;
	}
	if(GENERAL && L.type == Y4GPU_LIGHT_BACKGROUND)       // BackgroundLight::intersect (:139-151); `t` stays the caller's -1: unbounded shadow ray
	{
		float u, v;
		ipdf = y4_bg_from_dir(L, L.bg_abs_inter ? xneg(dir) : dir, u, v, true);
		col = y4_clamp_proportional(y4_bg_eval(L, y4_inv_sphere_map(u, v)), L.bg_clamp);
		t = -1.f;
		return true;
	}
	if(GENERAL && L.type == Y4GPU_LIGHT_SPOT_SOFT)        // SpotLight::intersect (light_spot.cc:199-246); dir_ = -ndir_
	{
		const V3 sdir = xneg(L.dir);
		const float cos_a = xdot(sdir, dir);
		if(cos_a == 0.f) return false;
		t = xdiv(xdot(sdir, xsub3(L.pos, from)), cos_a);
		if(t < 0.f) return false;
		const V3 p = xadd3(from, xscale(dir, t));
		if(xdot(sdir, xsub3(p, L.pos)) == 0.f)
		{
			if(xdot(p, p) <= 1e-2f)
			{
				const float cosa = xdot(sdir, dir);
				if(cosa < L.cos_end) return false;
				if(cosa >= L.cos_start) col = L.color;
				else
				{
					float v = xmul(xsub(cosa, L.cos_end), L.icos_diff);
					v = xmul(xmul(v, v), xsub(3.f, xmul(2.f, v)));
					col = cscale(L.color, v);
				}
				ipdf = xdiv(1.f, xmul(t, t));
				return true;
			}
		}
		return false;
	}
	if(L.type == Y4GPU_LIGHT_SUN)
	{
		const float cosine = xdot(dir, L.dir);
		if(cosine < L.cos_angle) return false;
		col = L.color; t = -1.f; ipdf = L.invpdf;
		return true;
	}
	const float cos_angle = xdot(dir, L.fnormal);
	if(cos_angle <= 0.f) return false;
	bool hit = false;
	// triIntersect(corner, c2, c3) then (corner, c3, c4), light_area.cc:117-134 (ray.from_ == sp.p, no tmin)
	for(int k = 0; k < 2 && !hit; ++k)
	{
		const V3 a = L.pos, b = k ? L.c3 : L.c2, c = k ? L.c4 : L.c3;
		const V3 e1 = xsub3(b, a), e2 = xsub3(c, a);
		const V3 pvec = xcross(dir, e2);
		const float det = xdot(e1, pvec);
		if(det == 0.f) continue;
		const float inv_det = xdiv(1.f, det);
		const V3 tvec = xsub3(from, a);
		const float u = xmul(xdot(tvec, pvec), inv_det);
		if(u < 0.f || u > 1.f) continue;
		const V3 qvec = xcross(tvec, e1);
		const float v = xmul(xdot(dir, qvec), inv_det);
		if((v < 0.f) || (xadd(u, v) > 1.f)) continue;
		t = xmul(xdot(e2, qvec), inv_det);
		hit = true;
	}
	if(!hit || !(t > 1.0e-10f)) return false;
	// 1.f / (t * t) * area_ * cos_angle * math::div_1_by_pi (last factor long double)
	ipdf = (float) xdmul((Y4_WIDE) xmul(xmul(xdiv(1.f, xmul(t, t)), L.area), cos_angle), Y4_1_PI);
	col = L.color;
	return true;
}

// One sample of TiledIntegrator::sampleAmbientOcclusion (integrator_tiled.cc:654-690, clay == false): a BSDF-sampled direction
// (Glossy | Diffuse | Reflect), an occlusion ray of length ao_dist; unoccluded it adds ao_col * surf_col * |n . dir| * w.
// Slot 1 carries "col += sp.emit(wo) * s.pdf_", which needs no ray.
Y4_HD void y4_ao_sample(const Y4PtScene &sc, const Y4PtMaterial &m, const Y4Surf &sp, const V3 wo, float s1, float s2, Y4DirectOut &out)
{
	out.valid[0] = out.valid[1] = false;
	out.contrib[0] = out.contrib[1] = c3f(0.f);
	out.ray[0] = out.ray[1] = y4_null_ray();
	const float tmin_shadow = sc.shadow_bias_auto ? xmul(sc.shadow_bias, fmaxf(1.f, xlen(sp.p))) : sc.shadow_bias;
	V3 wi = v3(0.f, 0.f, 0.f);
	float w = 0.f, pdf = 0.f; bool sampled, specular;
	const uint32_t fl = Y4_BSDF_GLOSSY | Y4_BSDF_DIFFUSE | Y4_BSDF_REFLECT;
	const C3 surf = y4_mat_sample(m, sp, wo, s1, s2, fl, wi, w, pdf, sampled, specular);
	if(m.bsdf_flags & Y4_BSDF_EMIT) out.contrib[1] = cscale(y4_mat_emit(m, sp, wo), pdf);
	if(!sampled) return;                                   // direction left at zero: |n . dir| = 0, nothing to add
	const float cosv = fabsf(xdot(sp.n, wi));
	out.contrib[0] = cscale(cscale(cmul(sc.ao_col, surf), cosv), w);
	out.ray[0] = y4_mk_ray(sp.p, tmin_shadow, wi, sc.ao_dist);
	out.valid[0] = true;
}

template<bool GENERAL = true>
Y4_HD void y4_direct_sample(const Y4PtScene &sc, const Y4PtLight &L, const Y4PtMaterial &m, const Y4Surf &sp, const V3 wo,
                            float s1, float s2, Y4DirectOut &out)
{
	out.valid[0] = out.valid[1] = false;
	out.contrib[0] = out.contrib[1] = c3f(0.f);
	out.ray[0] = out.ray[1] = y4_null_ray();
	const float plen = xlen(sp.p);
	const float tmin_shadow = sc.shadow_bias_auto ? xmul(sc.shadow_bias, fmaxf(1.f, plen)) : sc.shadow_bias;
	Y4LSample ls;
	if(y4_light_dirac(L.type))
	{
		// MonteCarloIntegrator::diracLight
		if(!y4_light_illuminate(L, sp.p, ls)) return;
		const float angle = m.flat_material ? 1.f : fabsf(xdot(sp.n, ls.dir));
		const C3 surf = y4_mat_eval(m, sp, wo, ls.dir);
		out.contrib[0] = cscale(cscale(cmul(surf, ls.col), angle), 1.f);    // surf_col * lcol * angle * transmit_col(1)
		out.ray[0] = y4_mk_ray(sp.p, tmin_shadow, ls.dir, ls.tmax);
		out.valid[0] = true;
		return;
	}
	const bool can_intersect = y4_light_can_intersect(L.type);
	// ---- areaLightSampleLight ----
	if(y4_light_illum_sample<GENERAL>(L, sp.p, s1, s2, ls) && ls.pdf > 1e-6f)
	{
		const C3 surf = y4_mat_eval(m, sp, wo, ls.dir);
		const float angle = m.flat_material ? 1.f : fabsf(xdot(sp.n, ls.dir));
		float w = 1.f;
		if(can_intersect)
		{
			const float m_pdf = y4_mat_pdf(m, sp, wo, ls.dir);
			if(m_pdf > 1e-6f)
			{
				const float l2 = xmul(ls.pdf, ls.pdf), m2 = xmul(m_pdf, m_pdf);
				w = xdiv(l2, xadd(l2, m2));
			}
		}
		out.contrib[0] = cdiv(cscale(cscale(cmul(surf, ls.col), angle), w), ls.pdf);
		out.ray[0] = y4_mk_ray(sp.p, tmin_shadow, ls.dir, ls.tmax);
		out.valid[0] = true;
	}
	// ---- areaLightSampleMaterial (lights with canIntersect() only) ----
	if(can_intersect)
	{
		const float tmin_b = sc.ray_min_dist_auto ? xmul(sc.ray_min_dist, fmaxf(1.f, plen)) : sc.ray_min_dist;
		V3 wi = v3(0.f, 0.f, 0.f);
		float W = 0.f, s_pdf = 0.f; bool sampled, specular;
		const uint32_t fl = Y4_BSDF_GLOSSY | Y4_BSDF_DIFFUSE | Y4_BSDF_DISPERSIVE | Y4_BSDF_REFLECT | Y4_BSDF_TRANSMIT;
		const C3 surf = y4_mat_sample(m, sp, wo, s1, s2, fl, wi, W, s_pdf, sampled, specular);
		float t = 0.f, ipdf = 0.f; C3 lcol = c3f(0.f);
		if(s_pdf > 1e-6f && y4_light_intersect<GENERAL>(L, sp.p, wi, tmin_b, t, lcol, ipdf) && ipdf > 1e-6f)
		{
			const float l_pdf = xdiv(1.f, ipdf);
			const float l2 = xmul(l_pdf, l_pdf), m2 = xmul(s_pdf, s_pdf);
			const float w = xdiv(m2, xadd(l2, m2));
			out.contrib[1] = cscale(cscale(cmul(surf, lcol), w), W);
			out.ray[1] = y4_mk_ray(sp.p, tmin_b, wi, t);
			out.valid[1] = true;
		}
	}
}
