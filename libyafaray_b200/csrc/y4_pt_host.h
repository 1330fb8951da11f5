// Host-side derivation of the device scene description from the parameter-level C-ABI structs
// (include/y4gpu.h), restating the reference constructors in plain fp32 with the reference's
// operation order.  Included by y4_pt.cu (product) and by the CPU unit-test emulation.
#pragma once
#include "y4_pt_stages.cuh"
#include "../../include/y4gpu.h"
#include <cstring>

namespace y4host {

// ---- host-side restatement of the reference's derived scene state (plain fp32, reference order) ----
struct HV { float x, y, z; };
inline HV hv(const float *p) { return {p[0], p[1], p[2]}; }
inline HV operator-(HV a, HV b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline HV operator+(HV a, HV b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline HV operator*(float f, HV v) { return {f * v.x, f * v.y, f * v.z}; }
inline HV operator*(HV v, float f) { return {f * v.x, f * v.y, f * v.z}; }
inline HV operator^(HV a, HV b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float lensqr(HV v) { return v.x * v.x + v.y * v.y + v.z * v.z; }
inline HV normalized(HV v)
{
	float len = lensqr(v);
	if(len != 0.f) { len = 1.f / sqrtf(len); v.x *= len; v.y *= len; v.z *= len; }
	return v;
}
inline V3 tov(HV v) { return v3(v.x, v.y, v.z); }

// Camera::Camera + PerspectiveCamera::PerspectiveCamera/setAxis (camera.cc:51-71, camera_perspective.cc:28-69)
inline Y4PtCamera derive_camera(const y4gpu_pt_camera &c)
{
	Y4PtCamera o;
	const HV position = hv(c.from), look = hv(c.to), up = hv(c.up);
	const float aspect_ratio = c.aspect_ratio * (float) c.resy / (float) c.resx;
	o.ang_aspect = aspect_ratio; o.ang_focal = 1.f; o.ang_max_radius = 0.f; o.ang_circular = c.circular; o.ang_projection = c.projection;
	HV cam_y = up - position;
	HV cam_z = look - position;
	HV cam_x = cam_z ^ cam_y;
	cam_y = cam_z ^ cam_x;
	cam_x = normalized(cam_x); cam_y = normalized(cam_y); cam_z = normalized(cam_z);
	const HV near_p = position + cam_z * c.near_clip, far_p = position + cam_z * c.far_clip;
	HV vright = cam_x;
	HV vup = aspect_ratio * cam_y;
	HV vto = (cam_z * c.focal) - 0.5f * (vup + vright);
	HV origin = position;
	o.type = c.type;
	if(c.type == Y4GPU_CAM_ARCHITECT)                   // ArchitectCamera::setAxis (camera_architect.cc:52-66): verticals stay vertical
	{
		vup = aspect_ratio * HV{0.f, 0.f, -1.f};
		vto = (cam_z * c.focal) - 0.5f * (vup + vright);
	}
	if(c.type == Y4GPU_CAM_ORTHOGRAPHIC)                // OrthographicCamera::setAxis (camera_orthographic.cc:37-49), scale in c.focal
	{
		const float scale = c.focal;
		vto = cam_z;
		origin = position - (float) (0.5 * scale) * (vup + vright);
		const float sy = scale / (float) c.resy, sx = scale / (float) c.resx;
		vup = vup * sy; vright = vright * sx;
	}
	else if(c.type == Y4GPU_CAM_EQUIRECTANGULAR) { vup = cam_y; vto = cam_z; }     // EquirectangularCamera::setAxis
	else if(c.type == Y4GPU_CAM_ANGULAR)
	{
		// AngularCamera::factory + ctor + setAxis (camera_angular.cc:30-55,113-115): the angles are double parameters times the
		// long double pi / 180, narrowed to the constructor's float arguments; std::tan(float) is this host's libm
		const float angle = (float) ((long double) c.angle_degrees * 0.01745329251994329576923690768489L);
		const float max_angle = (float) ((long double) c.max_angle_degrees * 0.01745329251994329576923690768489L);
		vright = c.mirrored ? HV{-cam_x.x, -cam_x.y, -cam_x.z} : cam_x; vup = cam_y; vto = cam_z;
		o.ang_max_radius = max_angle / angle;
		if(c.projection == Y4GPU_ANGULAR_ORTHOGRAPHIC) o.ang_focal = 1.f / y4_fast_sin(angle);
		else if(c.projection == Y4GPU_ANGULAR_STEREOGRAPHIC) o.ang_focal = 1.f / 2.f / tanf(angle / 2.f);
		else if(c.projection == Y4GPU_ANGULAR_EQUISOLID_ANGLE) o.ang_focal = 1.f / 2.f / y4_fast_sin(angle / 2.f);
		else if(c.projection == Y4GPU_ANGULAR_RECTILINEAR) o.ang_focal = 1.f / tanf(angle);
		else o.ang_focal = 1.f / angle;
	}
	else
	{
		vup = {vup.x / (float) c.resy, vup.y / (float) c.resy, vup.z / (float) c.resy};
		vright = {vright.x / (float) c.resx, vright.y / (float) c.resx, vright.z / (float) c.resx};
	}
	o.pos = tov(origin); o.vright = tov(vright); o.vup = tov(vup); o.vto = tov(vto);
	o.near_p = tov(near_p); o.near_n = tov(cam_z); o.far_p = tov(far_p); o.far_n = tov(cam_z);
	o.resx = c.resx; o.resy = c.resy;
	o.aperture = c.aperture; o.dof_distance = c.dof_distance;
	o.dof_rt = tov(c.aperture * cam_x); o.dof_up = tov(c.aperture * cam_y);       // setAxis: "premul with aperture"
	o.bkhtype = c.bokeh_type; o.bkhbias = c.bokeh_bias;
	for(int i = 0; i < 16; ++i) o.ls[i] = 0.f;
	int ns = c.bokeh_type;
	if(ns >= 3 && ns <= 6)              // ctor: corner table of the polygonal lens, FAST_TRIG cos / sin, long double constants
	{
		float w = (float) ((long double) c.bokeh_rotation * 0.017453292519943295769236907684886L);      // math::degToRad
		const float wi = (float) (6.283185307179586476925286766559L / (long double) (float) ns);
		ns = (ns + 2) * 2;
		for(int i = 0; i < ns; i += 2) { o.ls[i] = y4_fast_cos(w); o.ls[i + 1] = y4_fast_sin(w); w += wi; }
	}
	return o;
}

// ShinyDiffuseMaterial ctor + config() (material_shiny_diffuse.cc:29-82), LightMaterial (material_simple.cc:36-70)
inline Y4PtMaterial derive_material(const y4gpu_pt_material &m)
{
	Y4PtMaterial o;
	memset(&o, 0, sizeof(o));
	o.type = m.type;
	o.receive_shadows = m.receive_shadows; o.flat_material = m.flat_material; o.double_sided = m.double_sided;
	if(m.type == Y4GPU_MAT_LIGHT)
	{
		const float power = (float) (double) m.emit;       // col * static_cast<float>(power)
		o.color = c3(power * m.color[0], power * m.color[1], power * m.color[2]);
		o.bsdf_flags = Y4_BSDF_EMIT;
		return o;
	}
	if(m.type == Y4GPU_MAT_MIRROR)                         // MirrorMaterial ctor (material_glass.h:86-91): ref_col_ = r_col * ref_val
	{
		o.color = c3(m.color[0] * m.diffuse_reflect, m.color[1] * m.diffuse_reflect, m.color[2] * m.diffuse_reflect);
		o.bsdf_flags = Y4_BSDF_SPECULAR;
		return o;
	}
	if(m.type == Y4GPU_MAT_GLASS)                          // GlassMaterial ctor (material_glass.cc:35-50)
	{
		// `color` arrives as the constructor's filt_c = filt * filter_color + Rgb(1.f - filt) (factory, :326: `filt` is a double
		// parameter, so the caller forms it like the reference does and nothing is rounded twice here)
		o.color = c3(m.color[0], m.color[1], m.color[2]);
		o.mirror_color = c3(m.mirror_color[0], m.mirror_color[1], m.mirror_color[2]);
		o.ior = m.ior;
		o.bsdf_flags = Y4_BSDF_SPECULAR | Y4_BSDF_REFLECT | Y4_BSDF_TRANSMIT;      // BsdfFlags::AllSpecular
		if(m.absorb)                                       // BeerVolumeHandler ctor (volumehandler_beer.cc:27-35): std::log is this host's libm
		{
			o.absorb = 1;
			o.sigma_a = c3(-logf(m.absorption[0]), -logf(m.absorption[1]), -logf(m.absorption[2]));
			if(m.absorption_dist != 0.f)
			{
				const float inv = (float) (1.f / m.absorption_dist);      // "sigma_a_ *= 1.f / dist": a double quotient narrowed by Rgb::operator*=(float)
				o.sigma_a = c3(o.sigma_a.r * inv, o.sigma_a.g * inv, o.sigma_a.b * inv);
			}
		}
		return o;
	}
	o.color = c3(m.color[0], m.color[1], m.color[2]);
	o.layer_first = m.layer_first; o.layer_count = m.layer_count; o.emit_strength = m.emit;
	o.emit = c3(m.emit * m.color[0], m.emit * m.color[1], m.emit * m.color[2]);
	if(m.emit > 0.f) o.bsdf_flags |= Y4_BSDF_EMIT;
	float acc = 1.f;                                          // config() (:39-82); no translucency component
	o.mirror_color = c3(m.mirror_color[0], m.mirror_color[1], m.mirror_color[2]);
	if(m.specular_reflect > 0.00001f)
	{
		o.is_mirror = 1; o.mirror_strength = m.specular_reflect;
		if(!m.fresnel_effect) acc = 1.f - m.specular_reflect;
		o.bsdf_flags |= Y4_BSDF_SPECULAR | Y4_BSDF_REFLECT;
	}
	if(m.fresnel_effect) { o.fresnel = 1; o.ior_squared = m.ior * m.ior; }
	o.transmit_filter = m.transmit_filter;
	if(m.transparency * acc > 0.00001f)
	{
		o.is_transparent = 1; o.transparency_strength = m.transparency;
		acc *= 1.f - m.transparency;
		o.bsdf_flags |= Y4_BSDF_TRANSMIT | Y4_BSDF_FILTER;
	}
	if(m.translucency * acc > 0.00001f)
	{
		o.is_translucent = 1; o.translucency_strength = m.translucency;
		acc *= 1.f - m.transparency;                       // sic: config() takes the transparency strength here (:66)
		o.bsdf_flags |= Y4_BSDF_DIFFUSE | Y4_BSDF_TRANSMIT;
	}
	if(m.diffuse_reflect * acc > 0.00001f)
	{
		o.is_diffuse = 1; o.diffuse_strength = m.diffuse_reflect;
		o.bsdf_flags |= Y4_BSDF_DIFFUSE | Y4_BSDF_REFLECT;
	}
	if(m.oren_nayar)
	{
		const double sigma = (double) m.sigma, s2 = sigma * sigma;     // initOrenNayar (:150-156)
		o.oren_a = (float) (1.0 - 0.5 * (s2 / (s2 + 0.33)));
		o.oren_b = (float) (0.45 * s2 / (s2 + 0.09));
		o.oren_nayar = 1;
	}
	return o;
}

// ImageTexture::factory state (texture_image.cc:476-597: setCrop, setAdjustments); texels are set by the caller
inline Y4PtTexture derive_texture(const y4gpu_pt_texture &t)
{
	Y4PtTexture o;
	memset(&o, 0, sizeof(o));
	o.w = t.width; o.h = t.height; o.clip = t.clip; o.interp = t.interpolate;
	o.xrepeat = t.xrepeat; o.yrepeat = t.yrepeat; o.mirror_x = t.mirror_x; o.mirror_y = t.mirror_y; o.rot90 = t.rot90;
	o.cropminx = t.cropmin_x; o.cropmaxx = t.cropmax_x; o.cropminy = t.cropmin_y; o.cropmaxy = t.cropmax_y;
	o.cropx = (o.cropminx != 0.0) || (o.cropmaxx != 1.0);
	o.cropy = (o.cropminy != 0.0) || (o.cropmaxy != 1.0);
	o.checker_even = t.even_tiles; o.checker_odd = t.odd_tiles; o.checker_dist = t.checker_dist;
	o.adj_intensity = t.adj_intensity; o.adj_contrast = t.adj_contrast; o.adj_clamp = t.adj_clamp;
	o.adj_red = t.adj_mult_factor_red; o.adj_green = t.adj_mult_factor_green; o.adj_blue = t.adj_mult_factor_blue;
	o.adjustments_set = t.adj_intensity != 1.f || t.adj_contrast != 1.f || t.adj_clamp || t.adj_mult_factor_red != 1.f || t.adj_mult_factor_green != 1.f || t.adj_mult_factor_blue != 1.f;
	return o;
}
// TextureMapperNode::factory (shader_node_basic.cc:306-367) + LayerNode::factory / configInputs (shader_node_layer.cc:156-349)
inline Y4PtLayer derive_layer(const y4gpu_pt_layer &l)
{
	Y4PtLayer o;
	memset(&o, 0, sizeof(o));
	o.tex = l.texture; o.coords = l.texco; o.projection = l.mapping;
	const int map[3] = {l.proj_x, l.proj_y, l.proj_z};
	int *dst[3] = {&o.map_x, &o.map_y, &o.map_z};
	for(int k = 0; k < 3; ++k) *dst[k] = map[k] < 0 ? 0 : (map[k] > 3 ? 3 : map[k]);
	o.scale = v3(l.scale[0], l.scale[1], l.scale[2]);
	o.offset = v3(2 * l.offset[0], 2 * l.offset[1], 2 * l.offset[2]);
	o.flags = (uint32_t) l.flags; o.blend = l.blend_mode; o.do_color = l.do_color; o.colfac = l.colfac;
	for(int k = 0; k < 3; ++k) o.def_col[k] = l.def_col[k];
	o.def_col[3] = 1.f;                                            // Rgba{def_col}
	for(int k = 0; k < 4; ++k) o.upper_col[k] = l.upper_color[k];
	return o;
}

// PointLight ctor (light_point.cc:29-36), AreaLight ctor (light_area.cc:33-52), and the three lights below
inline Y4PtLight derive_light(const y4gpu_pt_light &l)
{
	Y4PtLight o;
	memset(&o, 0, sizeof(o));
	o.type = l.type; o.samples = l.samples; o.cast_shadows = l.cast_shadows;
	o.pos = v3(l.from[0], l.from[1], l.from[2]);
	o.color = c3(l.power * l.color[0], l.power * l.color[1], l.power * l.color[2]);       // color_ = col * inte
	if(l.type == Y4GPU_LIGHT_POINT || l.type == Y4GPU_LIGHT_BACKGROUND) return o;      // background light: derive_background_light
	if(l.type == Y4GPU_LIGHT_DIRECTIONAL)                 // DirectionalLight ctor (light_directional.cc:28-39); init() only feeds photon code
	{
		o.dir = tov(normalized(hv(l.point1)));
		o.radius = l.radius; o.infinite = l.infinite;
		return o;
	}
	if(l.type == Y4GPU_LIGHT_SUN)                         // SunLight ctor (light_sun.cc:29-43)
	{
		o.dir = tov(normalized(hv(l.point1)));
		y4_coords_system(tov(hv(l.point1)), o.du, o.dv);  // createCoordsSystem(dir): the ctor argument, not the normalised member
		float angle = l.angle;
		if(angle > 80.f) angle = 80.f;
		o.cos_angle = y4_fast_cos((float) ((long double) angle * 0.01745329251994329576923690768489L));
		o.invpdf = (float) (6.283185307179586476925286766559L * (long double) (1.f - o.cos_angle));
		o.pdf = (float) (1.0 / (double) o.invpdf);
		o.pdf = fminf(o.pdf, sqrtf(3.402823466e+38f));
		o.color = c3(o.color.r * o.pdf, o.color.g * o.pdf, o.color.b * o.pdf);            // col_pdf_
		return o;
	}
	if(l.type == Y4GPU_LIGHT_SPOT || l.type == Y4GPU_LIGHT_SPOT_SOFT)      // SpotLight ctor (light_spot.cc:30-43)
	{
		o.dir = tov(normalized(hv(l.from) - hv(l.point1)));                                // ndir_ = (from - to).normalize()
		if(l.type == Y4GPU_LIGHT_SPOT_SOFT)
		{
			y4_coords_system(xneg(o.dir), o.du, o.dv);                                     // createCoordsSystem(dir_), dir_ = -ndir_
			o.radius = l.radius;                                                           // shadow_fuzzy_
		}
		const double rad_angle = (double) (float) ((long double) l.angle * 0.01745329251994329576923690768489L);
		const double rad_inner_angle = rad_angle * (double) (1.f - l.blend);
		o.cos_start = y4_fast_cos((float) rad_inner_angle);
		o.cos_end = y4_fast_cos((float) rad_angle);
		o.icos_diff = 1.f / (o.cos_start - o.cos_end);
		return o;
	}
	if(l.type == Y4GPU_LIGHT_SPHERE)                      // SphereLight ctor (light_sphere.cc:30-40)
	{
		o.radius = l.radius;
		o.sq_radius = l.radius * l.radius;
		o.sq_radius_eps = (float) ((double) o.sq_radius * 1.000003815);
		return o;
	}
	const float pi_f = (float) 3.14159265358979323846264338327950288L;
	if(l.type == Y4GPU_LIGHT_PORTAL)                      // BackgroundPortalLight ctor: power_; area / cdf: derive_object_light; background: derive_light_background
	{
		o.power = l.power; o.double_sided = 0; o.obj_n = (int) l.object_n_tris;
		return o;
	}
	if(l.type == Y4GPU_LIGHT_OBJECT)                      // ObjectLight::factory: color * (float) power * num_pi; area / cdf: derive_object_light
	{
		o.color = c3(pi_f * (l.power * l.color[0]), pi_f * (l.power * l.color[1]), pi_f * (l.power * l.color[2]));
		o.double_sided = l.double_sided; o.obj_n = (int) l.object_n_tris;
		return o;
	}
	const HV corner = hv(l.from), to_x = hv(l.point1) - corner, to_y = hv(l.point2) - corner;
	HV fnormal = to_y ^ to_x;
	o.color = c3(pi_f * (l.power * l.color[0]), pi_f * (l.power * l.color[1]), pi_f * (l.power * l.color[2]));   // col * inte * num_pi
	// area_ = fnormal_.normLen(): normalises fnormal_ and returns its former length (vector.h)
	float len = lensqr(fnormal);
	float area = 0.f;
	if(len != 0.f)
	{
		area = sqrtf(len);
		const float inv = 1.f / area;
		fnormal = {fnormal.x * inv, fnormal.y * inv, fnormal.z * inv};
	}
	o.area = area; o.inv_area = (float) (1.0 / (double) area);
	o.fnormal = tov(fnormal);
	o.normal = v3(-fnormal.x, -fnormal.y, -fnormal.z);
	o.to_x = tov(to_x); o.to_y = tov(to_y);
	o.c2 = tov(corner + to_x); o.c3 = tov(corner + (to_x + to_y)); o.c4 = tov(corner + to_y);
	return o;
}

// ObjectLight::initIs (light_object_light.cc:46-70): triangle areas (TrianglePrimitive::surfaceArea, primitive_triangle.cc:208-213),
// their sum in double, Pdf1D::cumulateStep1DDf (sample_pdf1d.h:67-82) and the geometric normals.  Fills area / inv_area of `o`;
// the caller stores the three arrays where the light's pointers can reach them.
inline void derive_object_light(Y4PtLight &o, const float *tris, uint32_t n, std::vector<float> &cdf, std::vector<float4> &ng)
{
	cdf.assign(n, 0.f); ng.resize(n);
	std::vector<float> areas(n);
	double total_area = 0.0;
	for(uint32_t i = 0; i < n; ++i)
	{
		const float *v = tris + 9ull * i;
		const HV a = {v[0], v[1], v[2]}, b = {v[3], v[4], v[5]}, c = {v[6], v[7], v[8]};
		const HV cr = (b - a) ^ (c - a);
		areas[i] = 0.5f * sqrtf(lensqr(cr));
		total_area += areas[i];
		ng[i] = y4_tri_shade(tris, nullptr, 1u, i);
	}
	const double delta = 1.0 / (double) n;
	double acc = 0.0;
	for(uint32_t i = 0; i < n; ++i) { acc += (double) areas[i] * delta; cdf[i] = (float) acc; }
	const float integral = (float) acc;
	for(uint32_t i = 0; i < n; ++i) cdf[i] /= integral;
	o.area = (float) total_area;
	o.inv_area = (float) (1.f / total_area);
}

// Pdf1D::init / cumulateStep1DDf (sample_pdf1d.h:60-81): cdf of `n` function values appended to `cdf`; returns the integral
inline float derive_pdf1d(const float *func, int n, std::vector<float> &cdf)
{
	const size_t at = cdf.size();
	cdf.resize(at + n);
	const double delta = 1.0 / (double) n;
	double c = 0.0;
	for(int i = 0; i < n; ++i) { c += (double) func[i] * delta; cdf[at + i] = (float) c; }
	const float integral = (float) c;
	for(int i = 0; i < n; ++i) cdf[at + i] /= integral;
	return integral;
}

// TextureBackground ctor (background_texture.cc:30-42): rotation_ = 2 rot / 360, its FAST_TRIG sine / cosine (long double pi)
inline Y4BgTex derive_texture_background(const y4gpu_pt_params &p, const Y4PtTexture *tex)
{
	Y4BgTex b;
	b.tex = tex; b.angular = p.bg_texture_angular; b.power = p.bg_texture_power;
	b.rotation = 2.f * p.bg_texture_rotation / 360.f;
	b.sin_r = y4_fast_sin((float) (3.1415926535897932384626433832795L * (long double) b.rotation));
	b.cos_r = y4_fast_cos((float) (3.1415926535897932384626433832795L * (long double) b.rotation));
	return b;
}

// SunSkyBackground ctor (background_sunsky.cc:39-80) + the direction-independent half of perezFunction (:82-104).  Types as in the
// reference: sun_dir_ is a float vector, the members are doubles, pi is a long double; std::tan(double) is this host's libm.
inline Y4BgSunsky derive_sunsky(const y4gpu_pt_params &p)
{
	Y4BgSunsky s;
	memset(&s, 0, sizeof(s));
	const HV sun = normalized(hv(p.bg_sun_from));
	const double a_var = p.bg_sun_var[0], b_var = p.bg_sun_var[1], c_var = p.bg_sun_var[2], d_var = p.bg_sun_var[3], e_var = p.bg_sun_var[4];
	s.theta_s = (double) y4_acos(sun.z);
	const double theta_2 = s.theta_s * s.theta_s, theta_3 = theta_2 * s.theta_s;
	s.phi_s = (double) y4_libm_atan2f(sun.y, sun.x);
	const double t = (double) p.bg_sun_turbidity, t_2 = (double) (p.bg_sun_turbidity * p.bg_sun_turbidity);      // t_2_ = turb * turb: a float product
	const double chi = (double) ((long double) ((4.0 / 9.0 - t / 120.0)) * (3.1415926535897932384626433832795L - (long double) (2.0 * s.theta_s)));
	s.zenith_Y = (4.0453 * t - 4.9710) * tan(chi) - 0.2155 * t + 2.4192;
	s.zenith_Y *= 1000;
	s.zenith_x = (0.00165 * theta_3 - 0.00375 * theta_2 + 0.00209 * s.theta_s) * t_2 +
	             (-0.02903 * theta_3 + 0.06377 * theta_2 - 0.03202 * s.theta_s + 0.00394) * t +
	             (0.11693 * theta_3 - 0.21196 * theta_2 + 0.06052 * s.theta_s + 0.25886);
	s.zenith_y = (0.00275 * theta_3 - 0.00610 * theta_2 + 0.00317 * s.theta_s) * t_2 +
	             (-0.04214 * theta_3 + 0.08970 * theta_2 - 0.04153 * s.theta_s + 0.00516) * t +
	             (0.15346 * theta_3 - 0.26756 * theta_2 + 0.06670 * s.theta_s + 0.26688);
	s.perez_Y[0] = (0.17872 * t - 1.46303) * a_var; s.perez_Y[1] = (-0.35540 * t + 0.42749) * b_var; s.perez_Y[2] = (-0.02266 * t + 5.32505) * c_var;
	s.perez_Y[3] = (0.12064 * t - 2.57705) * d_var; s.perez_Y[4] = (-0.06696 * t + 0.37027) * e_var;
	s.perez_x[0] = (-0.01925 * t - 0.25922) * a_var; s.perez_x[1] = (-0.06651 * t + 0.00081) * b_var; s.perez_x[2] = (-0.00041 * t + 0.21247) * c_var;
	s.perez_x[3] = (-0.06409 * t - 0.89887) * d_var; s.perez_x[4] = (-0.00325 * t + 0.04517) * e_var;
	s.perez_y[0] = (-0.01669 * t - 0.26078) * a_var; s.perez_y[1] = (-0.09495 * t + 0.00921) * b_var; s.perez_y[2] = (-0.00792 * t + 0.21023) * c_var;
	s.perez_y[3] = (-0.04405 * t - 1.65369) * d_var; s.perez_y[4] = (-0.01092 * t + 0.05291) * e_var;
	s.sin_theta_s = y4_fast_sin((float) s.theta_s); s.cos_theta_s = y4_fast_cos((float) s.theta_s);
	s.night = (long double) s.theta_s > 0.5L * 3.1415926535897932384626433832795L;
	s.nfade_sun = (double) (1.0L - ((long double) s.theta_s * 0.31830988618379067153776752674502872L - 0.5L) * 2.0L);
	const double *lams[3] = {s.perez_x, s.perez_y, s.perez_Y};
	double *dens[3] = {&s.den_x, &s.den_y, &s.den_Y};
	for(int k = 0; k < 3; ++k)
	{
		const double *lam = lams[k];
		const double e_1 = y4_sunsky_exp(lam[1]), e_2 = y4_sunsky_exp(lam[3] * s.theta_s);
		*dens[k] = (1 + lam[0] * e_1) * (1 + lam[2] * e_2 + lam[4] * s.cos_theta_s * s.cos_theta_s);
	}
	s.power = p.bg_sun_power;
	return s;
}

// DarkSkyBackground ctor + prePerez (background_darksky.cc:40-100,150-154) and the ColorConv matrices (color_conversion.h:58-81;
// published XYZ -> RGB matrices).  std::exp / std::tan on doubles: this host's libm.
inline Y4BgDarksky derive_darksky(const y4gpu_pt_params &p)
{
	Y4BgDarksky s;
	memset(&s, 0, sizeof(s));
	float bright = p.bg_dark_bright;
	if(p.bg_dark_night) bright *= 0.5;
	s.power = p.bg_sun_power * bright; s.bright = bright; s.alt = p.bg_dark_altitude; s.night = p.bg_dark_night; s.exposure = p.bg_dark_exposure;
	static const float mats[4][9] = {
		{2.3706743f, -0.9000405f, -0.4706338f, -0.5138850f, 1.4253036f, 0.0885814f, 0.0052982f, -0.0146949f, 1.0093968f},       // CIE RGB, illuminant E
		{2.3638081f, -0.8676030f, -0.4988161f, -0.5005940f, 1.3962369f, 0.1047562f, 0.0141712f, -0.0306400f, 1.2323842f},       // CIE RGB, D50
		{3.2404542f, -1.5371385f, -0.4985314f, -0.9692660f, 1.8760108f, 0.0415560f, 0.0556434f, -0.2040259f, 1.0572252f},       // sRGB, D65
		{3.1338561f, -1.6168667f, -0.4906146f, -0.9787684f, 1.9161415f, 0.0334540f, 0.0719453f, -0.2289914f, 1.4052427f}};      // sRGB, D50
	const int cs = p.bg_dark_color_space >= 0 && p.bg_dark_color_space < 4 ? p.bg_dark_color_space : 0;
	for(int k = 0; k < 9; ++k) s.mat[k] = mats[cs][k];
	HV sun = hv(p.bg_sun_from);
	sun.z += s.alt;
	sun = normalized(sun);
	s.sun_dir = tov(sun);
	const double theta_s = (double) y4_acos(sun.z);
	const double cos_theta_s = (double) y4_fast_cos((float) theta_s);
	const double cos_theta_2 = cos_theta_s * cos_theta_s;
	const double theta_2 = theta_s * theta_s, theta_3 = theta_2 * theta_s;
	const double t = (double) p.bg_sun_turbidity, t_2 = (double) (p.bg_sun_turbidity * p.bg_sun_turbidity);
	const double av = p.bg_sun_var[0], bv = p.bg_sun_var[1], cv = p.bg_sun_var[2], dv = p.bg_sun_var[3], ev = p.bg_sun_var[4];
	const double chi = (double) ((long double) (0.44444444 - (t / 120.0)) * (3.1415926535897932384626433832795L - (long double) (2.0 * theta_s)));
	s.zenith_Y = (4.0453 * t - 4.9710) * tan(chi) - 0.2155 * t + 2.4192;
	s.zenith_Y *= 1000;
	s.zenith_x = (0.00165 * theta_3 - 0.00374 * theta_2 + 0.00209 * theta_s) * t_2 +
	             (-0.02902 * theta_3 + 0.06377 * theta_2 - 0.03202 * theta_s + 0.00394) * t +
	             (0.11693 * theta_3 - 0.21196 * theta_2 + 0.06052 * theta_s + 0.25885);
	s.zenith_y = (0.00275 * theta_3 - 0.00610 * theta_2 + 0.00316 * theta_s) * t_2 +
	             (-0.04214 * theta_3 + 0.08970 * theta_2 - 0.04153 * theta_s + 0.00515) * t +
	             (0.15346 * theta_3 - 0.26756 * theta_2 + 0.06669 * theta_s + 0.26688);
	s.perez_Y[0] = ((0.17872 * t) - 1.46303) * av; s.perez_Y[1] = ((-0.35540 * t) + 0.42749) * bv; s.perez_Y[2] = ((-0.02266 * t) + 5.32505) * cv;
	s.perez_Y[3] = ((0.12064 * t) - 2.57705) * dv; s.perez_Y[4] = ((-0.06696 * t) + 0.37027) * ev;
	s.perez_x[0] = ((-0.01925 * t) - 0.25922); s.perez_x[1] = ((-0.06651 * t) + 0.00081); s.perez_x[2] = ((-0.00041 * t) + 0.21247);
	s.perez_x[3] = ((-0.06409 * t) - 0.89887); s.perez_x[4] = ((-0.00325 * t) + 0.04517);
	s.perez_y[0] = ((-0.01669 * t) - 0.26078); s.perez_y[1] = ((-0.09495 * t) + 0.00921); s.perez_y[2] = ((-0.00792 * t) + 0.21023);
	s.perez_y[3] = ((-0.04405 * t) - 1.65369); s.perez_y[4] = ((-0.01092 * t) + 0.05291);
	double *lams[3] = {s.perez_Y, s.perez_x, s.perez_y};
	for(double *lam : lams)
	{
		const double p_num = ((1 + lam[0] * exp(lam[1])) * (1 + (lam[2] * exp(lam[3] * theta_s)) + (lam[4] * cos_theta_2)));
		lam[5] = p_num == 0.0 ? 0.0 : 1.0 / p_num;
	}
	return s;
}

// the background a background / portal light evaluates (Background::eval): a copy of the render parameters' description
// (tex: where the light finds the textureback's image -- a host copy while the IBL tables are derived, the device array after)
inline void derive_light_background(Y4PtLight &o, const y4gpu_pt_params &p, const Y4PtTexture *tex = nullptr, const Y4BgSunsky *sun = nullptr,
                                    const Y4BgDarksky *dark = nullptr)
{
	if(p.background_gradient == 2) { o.bg_tex = derive_texture_background(p, tex); o.bg_clamp = p.bg_ibl_clamp; }
	o.bg_sun = sun; o.bg_dark = dark;
	o.bg_gradient = p.background_gradient;
	o.bg_col = c3(p.background[0], p.background[1], p.background[2]);
	o.bg_gz = c3(p.bg_gradient[0][0], p.bg_gradient[0][1], p.bg_gradient[0][2]); o.bg_gh = c3(p.bg_gradient[1][0], p.bg_gradient[1][1], p.bg_gradient[1][2]);
	o.bg_sz = c3(p.bg_gradient[2][0], p.bg_gradient[2][1], p.bg_gradient[2][2]); o.bg_sh = c3(p.bg_gradient[3][0], p.bg_gradient[3][1], p.bg_gradient[3][2]);
}

// BackgroundLight::init (light_background.cc:46-75): the background's energy * sin(theta) on 360 latitude rows of
// 16 + (int)(sin(theta) * 704) columns, a Pdf1D per row and one over the row integrals.  The caller stores the arrays
// where the light's pointers can reach them; the background's own description is copied into the light for eval().
inline void derive_background_light(Y4PtLight &o, const y4gpu_pt_light &l, const y4gpu_pt_params &p,
                                    std::vector<int> &off, std::vector<float> &func, std::vector<float> &cdf, std::vector<float> &inv_int,
                                    const Y4PtTexture *host_tex = nullptr, const Y4BgSunsky *host_sun = nullptr, const Y4BgDarksky *host_dark = nullptr)
{
	o.bg_abs_inter = l.double_sided;
	derive_light_background(o, p, host_tex, host_sun, host_dark);
	const int nv = 360, max_u = 720, min_u = 16;
	o.bg_nv = nv;
	off.clear(); func.clear(); cdf.clear(); inv_int.clear();
	std::vector<float> fv(nv);
	const float inv_nv = 1.f / (float) nv;
	for(int y = 0; y < nv; ++y)
	{
		const float fy = ((float) y + 0.5f) * inv_nv;
		const float sintheta = y4_bg_sin_sample(fy);
		const int nu = min_u + (int) (sintheta * (float) (max_u - min_u));
		const float inv_nu = 1.f / (float) nu;
		off.push_back((int) func.size());
		for(int x = 0; x < nu; ++x)
		{
			const float fx = ((float) x + 0.5f) * inv_nu;
			const C3 c = y4_bg_eval(o, y4_inv_sphere_map(fx, fy));
			func.push_back(((c.r + c.g + c.b) * 0.333333f) * sintheta);            // Rgb::energy()
		}
		const float integral = derive_pdf1d(func.data() + off[y], nu, cdf);
		fv[y] = integral;
		inv_int.push_back(1.f / integral);
	}
	off.push_back((int) func.size());
	func.insert(func.end(), fv.begin(), fv.end());
	inv_int.push_back(1.f / derive_pdf1d(fv.data(), nv, cdf));
	off.push_back((int) func.size());
}

// ImageFilm ctor: filter table, filter width, table scale (imagefilm.cc:129-173, math/filter.h:34-90)
inline bool derive_film(const y4gpu_pt_params &p, Y4PtScene &sc)
{
	double filterw = (double) p.aa_pixelwidth * 0.5;          // filterw_(filter_size * 0.5)
	float filterw_f = (float) filterw;                        // float filterw_ member
	int kind = p.filter_type;
	if(kind == Y4GPU_FILTER_MITCHELL) filterw_f *= 2.6f;
	else if(kind == Y4GPU_FILTER_GAUSS) filterw_f *= 2.f;
	else if(kind != Y4GPU_FILTER_LANCZOS) kind = Y4GPU_FILTER_BOX;
	const int max_filter_size = 8;
	filterw_f = fminf(fmaxf(0.501f, filterw_f), 0.5f * max_filter_size);
	const float scale = 1.f / (float) Y4_FILTER_TABLE;
	for(int y = 0; y < Y4_FILTER_TABLE; ++y)
		for(int x = 0; x < Y4_FILTER_TABLE; ++x)
		{
			const float dx = (x + .5f) * scale, dy = (y + .5f) * scale;
			float v = 1.f;
			if(kind == Y4GPU_FILTER_GAUSS)
			{
				const float r_2 = dx * dx + dy * dy;
				v = fmaxf(0.f, y4_fast_exp(-6 * r_2) - 0.00247875f);
			}
			else if(kind == Y4GPU_FILTER_MITCHELL)
			{
				const float xx = 2.f * sqrtf(dx * dx + dy * dy);
				if(xx >= 2.f) v = 0.f;
				else if(xx >= 1.f) v = xx * (xx * (xx * -0.38888889f + 2.0f) - 3.33333333f) + 1.77777778f;
				else v = xx * xx * (1.16666666f * xx - 2.0f) + 0.88888889f;
			}
			else if(kind == Y4GPU_FILTER_LANCZOS)
			{
				const float xx = sqrtf(dx * dx + dy * dy);
				if(xx == 0.f) v = 1.f;
				else if(-2 < xx && xx < 2)
				{
					const float a = (float) (Y4_PI * (Y4_WIDE) xx), b = (float) (Y4_PI_2 * (Y4_WIDE) xx);
					v = (y4_fast_sin(a) * y4_fast_sin(b)) / (a * b);
				}
				else v = 0.f;
			}
			sc.filter_table[y * Y4_FILTER_TABLE + x] = v;
		}
	sc.filterw = (double) filterw_f;
	sc.table_scale = (double) (float) (0.9999 * Y4_FILTER_TABLE / (double) filterw_f);     // float table_scale_ member
	return true;
}

} // namespace y4host
using namespace y4host;
