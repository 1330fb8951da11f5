// Textured diffuse colour of shinydiffusemat: the `diffuse_shader` node chain evaluated per surface point
// (Y4_HD: kernels and the CPU unit-test emulation run the same code).  Restated operation for operation from
//   TextureMapperNode::getCoords/doMapping/cubeMap/eval   src/shader/shader_node_basic.cc:88-196
//   ImageTexture::getColor/doMapping/bilinearInterpolation/noInterpolation/findTextureInterpolationCoordinates
//                                                         src/texture/texture_image.cc:71-300
//   Texture::applyAdjustments                             src/texture/texture.cc:194-247 (hue/saturation rejected by the host)
//   LayerNode::eval/textureRgbBlend                       src/shader/shader_node_layer.cc:30-255
//   TrianglePrimitive::getSurface (orco, uv)              src/geometry/primitive/primitive_triangle.cc:113-155
// A material's chain is a stack of layer nodes (each one's upper_layer is the previous entry), every layer fed by one
// texture_mapper over an image texture.  Texels are float RGBA as Image::getColor returns them (the host applies the
// reference's colour-space conversion and "optimized" 10/8-bit storage rounding when it decodes the file).
#pragma once
#include "y4_pt_types.cuh"

struct C4 { float r, g, b, a; };
Y4_HD C4 c4(float r, float g, float b, float a) { C4 c; c.r = r; c.g = g; c.b = b; c.a = a; return c; }
Y4_HD C4 c4scale(float f, C4 c) { return c4(xmul(f, c.r), xmul(f, c.g), xmul(f, c.b), xmul(f, c.a)); }
Y4_HD C4 c4add(C4 a, C4 b) { return c4(xadd(a.r, b.r), xadd(a.g, b.g), xadd(a.b, b.b), xadd(a.a, b.a)); }

Y4_HD C4 y4_texel(const Y4PtTexture &T, int x, int y)
{
	const float4 t = T.texels[(size_t) y * (size_t) T.w + (size_t) x];
	return c4(t.x, t.y, t.z, t.w);
}

// ImageTexture::doMapping (texture_image.cc:98-176); returns true when the point is clipped away
Y4_HD bool y4_tex_do_mapping(const Y4PtTexture &T, float &x, float &y, float &z)
{
	bool outside = false;
	x = xadd(xmul(0.5f, x), 0.5f); y = xadd(xmul(0.5f, y), 0.5f); z = xadd(xmul(0.5f, z), 0.5f);
	if(T.clip == Y4GPU_CLIP_REPEAT)
	{
		if(T.xrepeat > 1) x = xmul(x, (float) T.xrepeat);
		if(T.yrepeat > 1) y = xmul(y, (float) T.yrepeat);
		if(T.mirror_x && ((int) ceilf(x)) % 2 == 0) x = -x;
		if(T.mirror_y && ((int) ceilf(y)) % 2 == 0) y = -y;
		if(x > 1.f) x = xsub(x, (float) (int) x);
		else if(x < 0.f) x = xadd(x, (float) (1 - (int) x));
		if(y > 1.f) y = xsub(y, (float) (int) y);
		else if(y < 0.f) y = xadd(y, (float) (1 - (int) y));
	}
	if(T.cropx) x = xadd(T.cropminx, xmul(x, xsub(T.cropmaxx, T.cropminx)));
	if(T.cropy) y = xadd(T.cropminy, xmul(y, xsub(T.cropmaxy, T.cropminy)));
	if(T.rot90) { const float t = x; x = y; y = t; }
	switch(T.clip)
	{
		case Y4GPU_CLIP_CLIPCUBE:
			if(x < 0.f || x > 1.f || y < 0.f || y > 1.f || z < -1.f || z > 1.f) outside = true;
			break;
		case Y4GPU_CLIP_CHECKER:
		{
			const int xs = (int) floorf(x), ys = (int) floorf(y);
			x = xsub(x, (float) xs); y = xsub(y, (float) ys);
			if(!T.checker_odd && !((xs + ys) & 1)) { outside = true; break; }
			if(!T.checker_even && ((xs + ys) & 1)) { outside = true; break; }
			if(T.checker_dist < 1.f)
			{
				x = xadd(xdiv(xsub(x, 0.5f), xsub(1.f, T.checker_dist)), 0.5f);
				y = xadd(xdiv(xsub(y, 0.5f), xsub(1.f, T.checker_dist)), 0.5f);
			}
		}
		// falls through to Clip, like the reference
		case Y4GPU_CLIP_CLIP:
			if(x < 0.f || x > 1.f || y < 0.f || y > 1.f) outside = true;
			break;
		case Y4GPU_CLIP_EXTEND:
			if(x > 0.99999f) x = 0.99999f; else if(x < 0.f) x = 0.f;
			if(y > 0.99999f) y = 0.99999f; else if(y < 0.f) y = 0.f;
			break;
		default: break;
	}
	return outside;
}

// findTextureInterpolationCoordinates (texture_image.cc:185-250): the two taps bilinear / nearest use + the fraction
Y4_HD void y4_tex_coords(int &c1, int &c2, float &frac, float cf, int res, bool repeat, bool mirror)
{
	if(repeat)
	{
		c1 = ((int) cf) % res;
		if(mirror)
		{
			if(cf < 0.f) { c2 = c1; frac = -cf; }
			else if(cf >= (float) (res - 1)) { c2 = c1; frac = xsub(cf, (float) (int) cf); }
			else
			{
				c2 = c1 + 1;
				if(c2 >= res) c2 = (2 * res - c2) % res;
				frac = xsub(cf, (float) (int) cf);
			}
		}
		else
		{
			if(cf > 0.f) { c2 = (c1 + 1) % res; frac = xsub(cf, (float) (int) cf); }
			else { c2 = (res - 1) % res; frac = -cf; }
		}
	}
	else
	{
		const int ci = (int) cf;
		c1 = ci < res - 1 ? ci : res - 1; if(c1 < 0) c1 = 0;
		if(cf > 0.f) c2 = c1 + 1 < res - 1 ? c1 + 1 : res - 1; else c2 = 0;
		frac = xsub(cf, floorf(cf));
	}
}

// ImageTexture::getColor (texture_image.cc:71-79) with None / Bilinear interpolation
Y4_HD C4 y4_tex_color(const Y4PtTexture &T, V3 p)
{
	float x = p.x, y = -p.y, z = p.z;
	if(y4_tex_do_mapping(T, x, y, z)) return c4(0.f, 0.f, 0.f, 0.f);
	const bool repeat = T.clip == Y4GPU_CLIP_REPEAT;
	const float half = T.interp == Y4GPU_INTERP_NONE ? 0.f : 0.5f;
	const float xf = xsub(xmul((float) T.w, xsub(x, floorf(x))), half);
	const float yf = xsub(xmul((float) T.h, xsub(y, floorf(y))), half);
	int x1, x2, y1, y2; float dx, dy;
	y4_tex_coords(x1, x2, dx, xf, T.w, repeat, T.mirror_x != 0);
	y4_tex_coords(y1, y2, dy, yf, T.h, repeat, T.mirror_y != 0);
	C4 c;
	if(T.interp == Y4GPU_INTERP_NONE) c = y4_texel(T, x1, y1);
	else
	{
		const C4 c11 = y4_texel(T, x1, y1), c21 = y4_texel(T, x2, y1), c12 = y4_texel(T, x1, y2), c22 = y4_texel(T, x2, y2);
		const float w11 = xmul(xsub(1.f, dx), xsub(1.f, dy)), w12 = xmul(xsub(1.f, dx), dy), w21 = xmul(dx, xsub(1.f, dy)), w22 = xmul(dx, dy);
		c = c4add(c4add(c4add(c4scale(w11, c11), c4scale(w12, c12)), c4scale(w21, c21)), c4scale(w22, c22));
	}
	if(T.adjustments_set)       // applyColorAdjustments(applyIntensityContrastAdjustments(c)); hue / saturation are rejected on the host
	{
		if(T.adj_intensity != 1.f || T.adj_contrast != 1.f)
		{
			c.r = xsub(xadd(xmul(xsub(c.r, 0.5f), T.adj_contrast), T.adj_intensity), 0.5f);
			c.g = xsub(xadd(xmul(xsub(c.g, 0.5f), T.adj_contrast), T.adj_intensity), 0.5f);
			c.b = xsub(xadd(xmul(xsub(c.b, 0.5f), T.adj_contrast), T.adj_intensity), 0.5f);
		}
		if(T.adj_clamp) { if(c.r < 0.f) c.r = 0.f; if(c.g < 0.f) c.g = 0.f; if(c.b < 0.f) c.b = 0.f; }
		if(T.adj_red != 1.f) c.r = xmul(c.r, T.adj_red);
		if(T.adj_green != 1.f) c.g = xmul(c.g, T.adj_green);
		if(T.adj_blue != 1.f) c.b = xmul(c.b, T.adj_blue);
		if(T.adj_clamp) { if(c.r < 0.f) c.r = 0.f; if(c.g < 0.f) c.g = 0.f; if(c.b < 0.f) c.b = 0.f; }
	}
	return c;
}

Y4_HD float y4_texmap(const V3 p, int k) { return k == 1 ? p.x : (k == 2 ? p.y : (k == 3 ? p.z : 0.f)); }
// TextureMapperNode::getCoords + doMapping (shader_node_basic.cc:106-164) for texco uv / orco / global, plain / cube
Y4_HD V3 y4_mapper_point(const Y4PtScene &sc, const Y4PtLayer &L, const Y4Surf &sp, float bu, float bv, int32_t prim)
{
	V3 texpt = sp.p, ng = sp.ng;
	const float bw = xsub(xsub(1.f, bu), bv);       // intersect_data.barycentric_u_ (weight of vertex 0)
	if(L.coords == Y4GPU_TEXCO_UV)
	{
		float u = bw, v = bu;                       // implicit mapping: sp->u_ = barycentric_u, sp->v_ = barycentric_v
		const float *uv = sc.tri_uv ? sc.tri_uv + 6ull * (uint32_t) prim : nullptr;
		if(uv && uv[0] == uv[0])                    // NaN marks a mesh without uv values
		{
			const float du_1 = xsub(uv[2], uv[0]), du_2 = xsub(uv[4], uv[0]), dv_1 = xsub(uv[3], uv[1]), dv_2 = xsub(uv[5], uv[1]);
			const float det = xsub(xmul(du_1, dv_2), xmul(dv_1, du_2));
			u = xadd(xadd(xmul(bw, uv[0]), xmul(bu, uv[2])), xmul(bv, uv[4]));
			v = xadd(xadd(xmul(bw, uv[1]), xmul(bu, uv[3])), xmul(bv, uv[5]));
			if(!(fabsf(det) > 1e-30f)) { u = bw; v = bu; }
		}
		texpt = v3(xsub(xmul(2.f, u), 1.f), xsub(xmul(2.f, v), 1.f), 0.f);
	}
	else if(L.coords == Y4GPU_TEXCO_ORCO)
	{
		const float *oc = sc.tri_orco ? sc.tri_orco + 9ull * (uint32_t) prim : nullptr;
		if(oc && oc[0] == oc[0])                    // NaN marks a mesh without orco points: orco_p_ = hit point, orco_ng_ = ng
		{
			const V3 o0 = v3(oc[0], oc[1], oc[2]), o1 = v3(oc[3], oc[4], oc[5]), o2 = v3(oc[6], oc[7], oc[8]);
			texpt = xadd3(xadd3(xscale(o0, bw), xscale(o1, bu)), xscale(o2, bv));
			ng = xnormalize(xcross(xsub3(o1, o0), xsub3(o2, o0)));
		}
	}
	const V3 t0 = texpt;                            // texmap {0, x, y, z} indexed by proj_x / proj_y / proj_z
	texpt = v3(y4_texmap(t0, L.map_x), y4_texmap(t0, L.map_y), y4_texmap(t0, L.map_z));
	if(L.projection == Y4GPU_PROJ_CUBE)
	{
		const float ax = fabsf(ng.x), ay = fabsf(ng.y), az = fabsf(ng.z);
		const float px = texpt.x, py = texpt.y, pz = texpt.z;
		if(az >= ax && az >= ay) texpt = v3(px, py, pz);
		else if(ay >= ax && ay >= az) texpt = v3(px, pz, py);
		else texpt = v3(py, pz, px);
	}
	return v3(xadd(xmul(texpt.x, L.scale.x), L.offset.x), xadd(xmul(texpt.y, L.scale.y), L.offset.y), xadd(xmul(texpt.z, L.scale.z), L.offset.z));
}

// LayerNode::textureRgbBlend (shader_node_layer.cc:193-255)
Y4_HD C3 y4_layer_blend(C3 tex, C3 out, float fact, float facg, int mode)
{
	switch(mode)
	{
		case Y4GPU_BLEND_MULT:
			fact = xmul(fact, facg);
			return cmul(cadd(c3f(xsub(1.f, facg)), cscale(tex, fact)), out);
		case Y4GPU_BLEND_SCREEN:
		{
			const C3 white = c3f(1.f);
			fact = xmul(fact, facg);
			const C3 wt = c3(xsub(white.r, tex.r), xsub(white.g, tex.g), xsub(white.b, tex.b)), wo = c3(xsub(white.r, out.r), xsub(white.g, out.g), xsub(white.b, out.b));
			const C3 m = cmul(cadd(c3f(xsub(1.f, facg)), cscale(wt, fact)), wo);
			return c3(xsub(white.r, m.r), xsub(white.g, m.g), xsub(white.b, m.b));
		}
		case Y4GPU_BLEND_SUB:
			fact = -fact;
		// falls through
		case Y4GPU_BLEND_ADD:
			fact = xmul(fact, facg);
			return cadd(cscale(tex, fact), out);
		case Y4GPU_BLEND_DIV:
		{
			fact = xmul(fact, facg);
			C3 itex = tex;
			if(itex.r != 0.f) itex.r = xdiv(1.f, itex.r);
			if(itex.g != 0.f) itex.g = xdiv(1.f, itex.g);
			if(itex.b != 0.f) itex.b = xdiv(1.f, itex.b);
			return cadd(cscale(out, xsub(1.f, fact)), cmul(cscale(out, fact), itex));
		}
		case Y4GPU_BLEND_DIFF:
		{
			fact = xmul(fact, facg);
			const C3 tmo = c3(fabsf(xsub(tex.r, out.r)), fabsf(xsub(tex.g, out.g)), fabsf(xsub(tex.b, out.b)));
			return cadd(cscale(out, xsub(1.f, fact)), cscale(tmo, fact));
		}
		case Y4GPU_BLEND_DARK:
		{
			fact = xmul(fact, facg);
			C3 col = cscale(tex, fact);
			if(col.r > out.r) col.r = out.r;
			if(col.g > out.g) col.g = out.g;
			if(col.b > out.b) col.b = out.b;
			return col;
		}
		case Y4GPU_BLEND_LIGHT:
		{
			fact = xmul(fact, facg);
			C3 col = cscale(tex, fact);
			if(col.r < out.r) col.r = out.r;
			if(col.g < out.g) col.g = out.g;
			if(col.b < out.b) col.b = out.b;
			return col;
		}
		default:    // Mix
			fact = xmul(fact, facg);
			return cadd(cscale(tex, fact), cscale(out, xsub(1.f, fact)));
	}
}

// The colour of the material's diffuse_shader at this surface point: initBsdf evaluates the nodes in dependency order
// (mapper, then its layer; lower layers before the ones stacked on them), getShaderColor reads the root layer's colour.
Y4_HD C3 y4_diffuse_shader_color(const Y4PtScene &sc, const Y4PtMaterial &m, const Y4Surf &sp, float bu, float bv, int32_t prim)
{
	C4 rcol = c4(0.f, 0.f, 0.f, 0.f);
	for(int li = 0; li < m.layer_count; ++li)
	{
		const Y4PtLayer &L = sc.layers[m.layer_first + li];
		if(li == 0) rcol = c4(L.upper_col[0], L.upper_col[1], L.upper_col[2], L.upper_col[3]);
		float stencil_tin = rcol.a;
		C4 texcolor = y4_tex_color(sc.textures[L.tex], y4_mapper_point(sc, L, sp, bu, bv, prim));
		float tin = 0.f, ta = texcolor.a;
		bool tex_rgb = true;                        // color_input (scalar inputs are rejected on the host)
		if(L.flags & Y4GPU_LAYER_RGB_TO_INT)
		{
			tin = xadd(xadd(xmul(0.2126f, texcolor.r), xmul(0.7152f, texcolor.g)), xmul(0.0722f, texcolor.b));
			tex_rgb = false;
		}
		if(L.flags & Y4GPU_LAYER_NEGATIVE)
		{
			if(tex_rgb) texcolor = c4(xsub(1.f, texcolor.r), xsub(1.f, texcolor.g), xsub(1.f, texcolor.b), xsub(1.f, texcolor.a));
			tin = xsub(1.f, tin);
		}
		if(L.flags & Y4GPU_LAYER_STENCIL)
		{
			if(tex_rgb) { const float fact = ta; ta = xmul(ta, stencil_tin); stencil_tin = xmul(stencil_tin, fact); }
			else { const float fact = tin; tin = xmul(tin, stencil_tin); stencil_tin = xmul(stencil_tin, fact); }
		}
		if(L.do_color)
		{
			if(!tex_rgb) texcolor = c4(L.def_col[0], L.def_col[1], L.def_col[2], L.def_col[3]);
			else tin = ta;
			const float tin_t = tin > 1.f ? 1.f : (tin < 0.f ? 0.f : tin);
			C3 r = y4_layer_blend(c3(texcolor.r, texcolor.g, texcolor.b), c3(rcol.r, rcol.g, rcol.b), tin_t, xmul(stencil_tin, L.colfac), L.blend);
			if(r.r < 0.f) r.r = 0.f;
			if(r.g < 0.f) r.g = 0.f;
			if(r.b < 0.f) r.b = 0.f;
			rcol = c4(r.r, r.g, r.b, 1.f);
		}
		rcol.a = stencil_tin;
	}
	return c3(rcol.r, rcol.g, rcol.b);
}
