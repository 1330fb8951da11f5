// PathIntegratorB200: see include/integrator/surface/integrator_path_b200.h.
//
// preprocess() exports the scene in PARAMETER form to the C-ABI (include/y4gpu.h): the recorded ParamMaps of the camera,
// materials (+ their diffuse_shader node chains and image textures), lights, background, integrator and render settings are
// read with the same names, types and defaults as the reference's factories read them.  Anything outside the supported set
// fails preprocess() -- and so yafaray_render -- with an error; nothing is approximated and there is no CPU fallback.
#include "integrator/surface/integrator_path_b200.h"
#include "accelerator/accelerator_b200.h"
#include "background/background.h"
#include "camera/camera.h"
#include "common/logger.h"
#include "common/layers.h"
#include "geometry/matrix4.h"
#include "geometry/uv.h"
#include "geometry/object/object_mesh.h"
#include "geometry/primitive/primitive_face.h"
#include "image/image.h"
#include "light/light.h"
#include "material/material.h"
#include "render/imagefilm.h"
#include "render/render_control.h"
#include "render/render_view.h"
#include "scene/scene.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <set>

BEGIN_YAFARAY

namespace {

template <class T> T get(const ParamMap &p, const char *name, T def) { p.getParam(name, def); return def; }
std::string getStr(const ParamMap &p, const char *name, const std::string &def = "") { std::string s = def; p.getParam(name, s); return s; }
bool has(const ParamMap &p, const std::string &name) { for(const auto &kv : p) if(kv.first == name) return true; return false; }
void put3(float *dst, const Vec3 &v) { dst[0] = v.x(); dst[1] = v.y(); dst[2] = v.z(); }
void put3(float *dst, const Rgb &c) { dst[0] = c.r_; dst[1] = c.g_; dst[2] = c.b_; }
Vec3 getVec(const ParamMap &p, const char *name, const Vec3 &def) { Vec3 v = def; p.getParam(name, v); return v; }
Rgb getCol(const ParamMap &p, const char *name, const Rgb &def) { Rgb c = def; p.getParam(name, c); return c; }

int lookup(const std::map<std::string, int> &table, const std::string &key, int def) { const auto i = table.find(key); return i == table.end() ? def : i->second; }

struct Fail { std::string msg; };
[[noreturn]] void unsupported(const std::string &msg) { throw Fail{msg}; }

// NodeMaterial::loadNodes / parseNodes (src/material/material_node.cc:102-187) for the `diffuse_shader` slot: the chain of
// (layer, texture_mapper) node pairs, bottom layer first; empty when the reference itself drops the node tree.
struct ChainEntry { const ParamMap *layer, *mapper; };
std::vector<ChainEntry> diffuseChain(const std::string &mat_name, const ParamMap &mat, const std::list<ParamMap> &node_params, const Scene &scene)
{
	std::map<std::string, const ParamMap *> nodes;
	for(const ParamMap &n : node_params)
	{
		if(getStr(n, "element", "shader_node") != "shader_node") continue;
		const std::string name = getStr(n, "name"), type = getStr(n, "type");
		if(name.empty() || type.empty() || nodes.count(name)) return {};
		if(type != "layer" && type != "texture_mapper") unsupported("material '" + mat_name + "': shader node type '" + type + "'");
		if(type == "texture_mapper" && !scene.getTexture(getStr(n, "texture"))) return {};       // TextureMapperNode::factory fails -> the nodes are dropped
		nodes[name] = &n;
	}
	for(const auto &kv : nodes)
	{
		const ParamMap &n = *kv.second;
		if(getStr(n, "type") != "layer") continue;
		if(!nodes.count(getStr(n, "input")) || (has(n, "upper_layer") && !nodes.count(getStr(n, "upper_layer")))) return {};
	}
	const auto root = nodes.find(getStr(mat, "diffuse_shader"));
	if(root == nodes.end()) return {};
	std::vector<ChainEntry> chain;
	std::set<std::string> seen;
	const ParamMap *n = root->second;
	while(n)
	{
		if(getStr(*n, "type") != "layer" || !seen.insert(getStr(*n, "name")).second) unsupported("material '" + mat_name + "': the diffuse_shader must be a stack of layer nodes");
		const ParamMap *mapper = nodes[getStr(*n, "input")];
		if(getStr(*mapper, "type") != "texture_mapper") unsupported("material '" + mat_name + "': layer inputs other than texture_mapper nodes");
		if(!get<bool>(*n, "color_input", true)) unsupported("material '" + mat_name + "': scalar layer inputs");
		const std::string texco = getStr(*mapper, "texco", "global"), mapping = getStr(*mapper, "mapping", "plain");
		if(texco == "transformed" || texco == "window" || texco == "normal") unsupported("material '" + mat_name + "': texco '" + texco + "'");
		if(mapping == "tube" || mapping == "sphere") unsupported("material '" + mat_name + "': mapping '" + mapping + "'");
		chain.insert(chain.begin(), ChainEntry{n, mapper});
		n = has(*n, "upper_layer") ? nodes[getStr(*n, "upper_layer")] : nullptr;
	}
	return chain;
}

const std::map<std::string, int> kCameraTypes{{"perspective", 0}, {"architect", 1}, {"orthographic", 2}, {"equirectangular", 3}, {"angular", 4}};
const std::map<std::string, int> kAngularProjections{{"orthographic", Y4GPU_ANGULAR_ORTHOGRAPHIC}, {"stereographic", Y4GPU_ANGULAR_STEREOGRAPHIC},
                                                     {"equisolid_angle", Y4GPU_ANGULAR_EQUISOLID_ANGLE}, {"rectilinear", Y4GPU_ANGULAR_RECTILINEAR}};
const std::map<std::string, int> kBokehTypes{{"disk1", 0}, {"disk2", 1}, {"triangle", 3}, {"square", 4}, {"pentagon", 5}, {"hexagon", 6}, {"ring", 7}};
const std::map<std::string, int> kBokehBias{{"uniform", 0}, {"center", 1}, {"edge", 2}};
const std::map<std::string, int> kClip{{"extend", 0}, {"clip", 1}, {"clipcube", 2}, {"repeat", 3}, {"checker", 4}};
const std::map<std::string, int> kInterp{{"none", 0}, {"bilinear", 1}};
const std::map<std::string, int> kTexco{{"global", 0}, {"uv", 1}, {"orco", 2}};
const std::map<std::string, int> kProj{{"plain", 0}, {"cube", 1}};
const std::map<std::string, int> kBlend{{"mix", 0}, {"add", 1}, {"multiply", 2}, {"subtract", 3}, {"screen", 4}, {"divide", 5}, {"difference", 6}, {"darken", 7}, {"lighten", 8}};
const std::map<std::string, int> kFilters{{"box", 0}, {"gauss", 1}, {"mitchell", 2}, {"lanczos", 3}};

} // namespace

bool PathIntegratorB200::requested(const ParamMap &params)
{
	std::string backend;
	if(params.getParam("backend", backend)) return backend == "b200";
	const char *e = getenv("YAFARAY_B200_BACKEND");
	return e && atoi(e) != 0;
}

Integrator *PathIntegratorB200::factory(Logger &logger, const ParamMap &params, const Scene &, RenderControl &render_control)
{
	return new PathIntegratorB200(render_control, logger, params);
}

std::pair<Rgb, float> PathIntegratorB200::integrate(Ray &, RandomGenerator &, ColorLayers *, int, int, bool, float, int, const RayDivision &, const PixelSamplingData &) const
{
	logger_.logError(getName(), ": integrate() on a single ray is not provided by the GPU back end (whole passes only)");
	return {Rgb{0.f}, 0.f};
}

bool PathIntegratorB200::preprocess(ImageFilm *image_film, const RenderView *render_view, const Scene &scene)
{
	if(!SurfaceIntegrator::preprocess(image_film, render_view, scene)) return false;
	accel_ = dynamic_cast<const AcceleratorB200 *>(accelerator_);
	if(!accel_) { logger_.logError(getName(), ": needs scene_accelerator = yafaray-b200-bvh8 (the scene was built with another accelerator)"); return false; }
	ctx_ = accel_->context();
	if(!accel_->refreshVisibility()) { logger_.logError(getName(), ": ", y4gpu_last_error(ctx_)); return false; }
	const B200SceneParams &rec = scene.getB200Params();
	const ParamMap &rp = rec.render;
	try
	{
		// ---- layers: only the Combined layer is produced -----------------------------------------------------------------
		for(const auto &l : *layers_) if(l.first != LayerDef::Combined && l.second.isExported()) unsupported("exported render layers other than Combined");
		if(vol_integrator_) unsupported("volume integrators");
		// ---- integrator ------------------------------------------------------------------------------------------------------
		const ParamMap &ip = params_;
		const std::string itype = getStr(ip, "type");
		if(get<bool>(ip, "caustics", false)) unsupported("photon caustics");
		if(get<bool>(ip, "do_AO", false) && get<int>(ip, "AO_samples", 32) < 1) unsupported("do_AO with AO_samples < 1");
		const std::string caustic_type = getStr(ip, "caustic_type", "path");
		if(caustic_type == "photon" || caustic_type == "both") unsupported("photon caustics");
		y4gpu_pt_params pr;
		memset(&pr, 0, sizeof(pr));
		pr.bounces = get<int>(ip, "bounces", 3); pr.path_samples = get<int>(ip, "path_samples", 32);
		if(itype == "directlighting") { pr.bounces = 1; pr.path_samples = 0; }
		pr.russian_roulette_min_bounces = get<int>(ip, "russian_roulette_min_bounces", 0);
		pr.no_recursive = get<bool>(ip, "no_recursive", false);
		pr.transp_background = get<bool>(ip, "bg_transp", false);
		pr.raydepth = get<int>(ip, "raydepth", 5);
		pr.transp_refracted_background = get<bool>(ip, "bg_transp_refract", false);
		pr.trace_caustics = itype == "pathtracing" && caustic_type != "none";
		pr.transp_shadows = get<bool>(ip, "transpShad", false); pr.shadow_depth = get<int>(ip, "shadowDepth", 5);
		if(get<bool>(ip, "do_AO", false) && itype == "directlighting")     // integrator_direct_light.cc:124,170-173 (the path tracer never adds AO to Combined)
		{
			pr.ao_samples = get<int>(ip, "AO_samples", 32);
			pr.ao_distance = (float) get<double>(ip, "AO_distance", 1.0);
			put3(pr.ao_color, getCol(ip, "AO_color", Rgb{1.f}));
		}
		// ---- camera ------------------------------------------------------------------------------------------------------------
		const std::string view_name = render_view->getName();
		const auto vp = rec.render_views.find(view_name);
		if(vp == rec.render_views.end()) unsupported("render view without recorded parameters");
		const auto cp_it = rec.cameras.find(getStr(vp->second, "camera_name"));
		if(cp_it == rec.cameras.end()) unsupported("camera without recorded parameters");
		const ParamMap &cp = cp_it->second;
		const std::string ctype = getStr(cp, "type");
		if(!kCameraTypes.count(ctype)) unsupported("camera type '" + ctype + "'");
		y4gpu_pt_camera cam;
		memset(&cam, 0, sizeof(cam));
		put3(cam.from, getVec(cp, "from", Vec3{0.f, 1.f, 0.f})); put3(cam.to, getVec(cp, "to", Vec3{0.f, 0.f, 0.f})); put3(cam.up, getVec(cp, "up", Vec3{0.f, 1.f, 1.f}));
		cam.resx = get<int>(cp, "resx", 320); cam.resy = get<int>(cp, "resy", 200);
		cam.type = kCameraTypes.at(ctype);
		cam.aspect_ratio = get<float>(cp, "aspect_ratio", 1.f); cam.focal = get<float>(cp, "focal", 1.f);
		if(cam.type == Y4GPU_CAM_ORTHOGRAPHIC) cam.focal = get<float>(cp, "scale", 1.f);
		if(cam.type == Y4GPU_CAM_PERSPECTIVE || cam.type == Y4GPU_CAM_ARCHITECT) { cam.aperture = get<float>(cp, "aperture", 0.f); cam.dof_distance = get<float>(cp, "dof_distance", 0.f); }
		cam.near_clip = get<float>(cp, "nearClip", 0.f);
		cam.far_clip = get<float>(cp, "farClip", (cam.type == Y4GPU_CAM_EQUIRECTANGULAR || cam.type == Y4GPU_CAM_ANGULAR) ? -1.0e38f : -1.f);
		if(cam.type == Y4GPU_CAM_ANGULAR)                // AngularCamera::factory (camera_angular.cc:80-117)
		{
			cam.angle_degrees = get<double>(cp, "angle", 90.0);
			cam.max_angle_degrees = get<double>(cp, "max_angle", cam.angle_degrees);
			cam.circular = get<bool>(cp, "circular", true) ? 1 : 0;
			cam.mirrored = get<bool>(cp, "mirrored", false) ? 1 : 0;
			cam.projection = lookup(kAngularProjections, getStr(cp, "projection", ""), Y4GPU_ANGULAR_EQUIDISTANT);
		}
		cam.bokeh_type = lookup(kBokehTypes, getStr(cp, "bokeh_type", "disk1"), 0);
		cam.bokeh_bias = lookup(kBokehBias, getStr(cp, "bokeh_bias", "uniform"), 0);
		cam.bokeh_rotation = get<float>(cp, "bokeh_rotation", 0.f);
		// ---- materials, in the order of the accelerator's material table (= the material ids in the triangle flag words) ------------
		const std::vector<const Material *> &mat_table = accel_->materialTable();
		std::vector<std::string> mat_names(mat_table.size());
		for(size_t i = 0; i < mat_table.size(); ++i)
		{
			for(const auto &kv : rec.materials)
			{
				const std::unique_ptr<const Material> *m = scene.getMaterial(kv.first);
				if(m && m->get() == mat_table[i]) { mat_names[i] = kv.first; break; }
			}
			if(mat_names[i].empty()) unsupported("a primitive uses a material the scene does not know by name");
		}
		std::vector<y4gpu_pt_material> mats(mat_table.size());
		std::vector<y4gpu_pt_layer> layers;
		std::vector<std::string> tex_names;
		std::map<std::string, int> tex_index;
		bool any_transparent_textured = false;
		for(size_t i = 0; i < mats.size(); ++i)
		{
			const ParamMap &p = rec.materials.at(mat_names[i]);
			y4gpu_pt_material &m = mats[i];
			memset(&m, 0, sizeof(m));
			const std::string type = getStr(p, "type");
			put3(m.color, getCol(p, "color", Rgb{1.f}));
			m.receive_shadows = get<bool>(p, "receive_shadows", true); m.flat_material = get<bool>(p, "flat_material", false);
			m.mirror_color[0] = m.mirror_color[1] = m.mirror_color[2] = 1.f; m.ior = 1.33f; m.transmit_filter = 1.f; m.diffuse_reflect = 1.f;
			if(type == "light_mat") { m.type = Y4GPU_MAT_LIGHT; m.emit = get<float>(p, "power", 1.f); m.double_sided = get<bool>(p, "double_sided", false); }
			else if(type == "mirror") { m.type = Y4GPU_MAT_MIRROR; m.diffuse_reflect = get<float>(p, "reflect", 1.f); }
			else if(type == "glass")
			{
				// GlassMaterial::factory (material_glass.cc:290-433): IOR / transmit_filter / dispersion_power are double parameters
				if(get<double>(p, "dispersion_power", 0.0) > 0.0 || get<bool>(p, "fake_shadows", false)) unsupported("glass '" + mat_names[i] + "': dispersion / fake_shadows");
				const Rgb ab = getCol(p, "absorption", Rgb{1.f});
				if(ab.r_ < 1.f || ab.g_ < 1.f || ab.b_ < 1.f)      // material_glass.cc:339-365: a Beer-Lambert volume handler for the inside
				{
					if(ab.r_ <= 1e-38f || ab.g_ <= 1e-38f || ab.b_ <= 1e-38f) unsupported("glass '" + mat_names[i] + "': absorption colour with a zero component");
					m.absorb = 1; put3(m.absorption, ab);
					m.absorption_dist = get<double>(p, "absorption_dist", 1.0);
				}
				if(get<float>(p, "wireframe_amount", 0.f) != 0.f || get<float>(p, "samplingfactor", 1.f) != 1.f || get<int>(p, "additionaldepth", 0) != 0) unsupported("glass '" + mat_names[i] + "': wireframe / samplingfactor / additionaldepth");
				for(const auto &kv : p) if(kv.first.size() > 7 && kv.first.compare(kv.first.size() - 7, 7, "_shader") == 0) unsupported("glass '" + mat_names[i] + "': shader nodes");
				m.type = Y4GPU_MAT_GLASS;
				m.ior = (float) get<double>(p, "IOR", 1.4);
				const double filt = get<double>(p, "transmit_filter", 0.0);
				put3(m.color, filt * getCol(p, "filter_color", Rgb{1.f}) + Rgb(1.f - filt));      // the constructor argument, formed like the factory does
				put3(m.mirror_color, getCol(p, "mirror_color", Rgb{1.f}));
			}
			else if(type == "shinydiffusemat")
			{
				m.type = Y4GPU_MAT_SHINYDIFFUSE;
				m.diffuse_reflect = get<float>(p, "diffuse_reflect", 1.f); m.emit = get<float>(p, "emit", 0.f);
				m.specular_reflect = get<float>(p, "specular_reflect", 0.f); m.fresnel_effect = get<bool>(p, "fresnel_effect", false); m.ior = get<float>(p, "IOR", 1.33f);
				put3(m.mirror_color, getCol(p, "mirror_color", Rgb{1.f}));
				m.transparency = get<float>(p, "transparency", 0.f); m.transmit_filter = get<float>(p, "transmit_filter", 1.f);
				m.translucency = get<float>(p, "translucency", 0.f);
				if(getStr(p, "diffuse_brdf", "lambert") == "oren_nayar") { m.oren_nayar = 1; m.sigma = get<float>(p, "sigma", 0.1f); }
				if(get<float>(p, "transparentbias_factor", 0.f) != 0.f) unsupported("shinydiffusemat '" + mat_names[i] + "': transparentbias_factor");
				if(get<float>(p, "samplingfactor", 1.f) != 1.f || get<int>(p, "additionaldepth", 0) != 0) unsupported("shinydiffusemat '" + mat_names[i] + "': samplingfactor / additionaldepth");
				if(get<float>(p, "wireframe_amount", 0.f) != 0.f) unsupported("shinydiffusemat '" + mat_names[i] + "': wireframe shading");
				for(const auto &kv : p)
				{
					const std::string &k = kv.first;
					if(k.size() > 7 && k.compare(k.size() - 7, 7, "_shader") == 0 && k != "diffuse_shader") unsupported("shinydiffusemat '" + mat_names[i] + "': shader slot '" + k + "'");
				}
				if(has(p, "diffuse_shader"))
				{
					const std::vector<ChainEntry> chain = diffuseChain(mat_names[i], p, rec.material_nodes.at(mat_names[i]), scene);
					if(!chain.empty())
					{
						m.layer_first = (int32_t) layers.size(); m.layer_count = (int32_t) chain.size();
						if(m.transparency > 0.00001f) any_transparent_textured = true;
						for(size_t li = 0; li < chain.size(); ++li)
						{
							const ParamMap &lay = *chain[li].layer, &mapper = *chain[li].mapper;
							const std::string tname = getStr(mapper, "texture");
							if(!tex_index.count(tname)) { tex_index[tname] = (int) tex_names.size(); tex_names.push_back(tname); }
							y4gpu_pt_layer L;
							memset(&L, 0, sizeof(L));
							L.texture = tex_index[tname];
							L.texco = lookup(kTexco, getStr(mapper, "texco", "global"), 0);
							L.mapping = lookup(kProj, getStr(mapper, "mapping", "plain"), 0);
							L.proj_x = get<int>(mapper, "proj_x", 1); L.proj_y = get<int>(mapper, "proj_y", 2); L.proj_z = get<int>(mapper, "proj_z", 3);
							put3(L.scale, getVec(mapper, "scale", Vec3{1.f, 1.f, 1.f})); put3(L.offset, getVec(mapper, "offset", Vec3{0.f, 0.f, 0.f}));
							L.flags = (get<bool>(lay, "noRGB", false) ? Y4GPU_LAYER_RGB_TO_INT : 0) | (get<bool>(lay, "stencil", false) ? Y4GPU_LAYER_STENCIL : 0) | (get<bool>(lay, "negative", false) ? Y4GPU_LAYER_NEGATIVE : 0);
							L.blend_mode = lookup(kBlend, getStr(lay, "blend_mode", "mix"), 0);
							L.do_color = get<bool>(lay, "do_color", true); L.colfac = get<float>(lay, "colfac", 1.f);
							put3(L.def_col, getCol(lay, "def_col", Rgb{1.f}));
							if(li == 0)
							{
								Rgba uc{0.f, 0.f, 0.f, 0.f};
								lay.getParam("upper_color", uc);
								L.upper_color[0] = uc.r_; L.upper_color[1] = uc.g_; L.upper_color[2] = uc.b_; L.upper_color[3] = uc.a_;
							}
							layers.push_back(L);
						}
					}
				}
			}
			else unsupported("material type '" + type + "'");
		}
		if(pr.transp_shadows && any_transparent_textured) unsupported("textured transparent materials with transpShad");
		// a textureback's image travels with the material textures
		const ParamMap *background_params = nullptr;
		if(background_) for(const auto &kv : rec.backgrounds) if(scene.getBackground(kv.first) == background_) background_params = &kv.second;
		if(background_params && getStr(*background_params, "type") == "textureback")
		{
			const std::string tname = getStr(*background_params, "texture");
			if(!rec.textures.count(tname)) unsupported("textureback without a recorded texture");
			if(!tex_index.count(tname)) { tex_index[tname] = (int) tex_names.size(); tex_names.push_back(tname); }
		}
		// ---- textures: the texels exactly as Image::getColor(x, y) returns them ----------------------------------------------------
		std::vector<y4gpu_pt_texture> textures(tex_names.size());
		std::vector<std::vector<float>> texels(tex_names.size());
		for(size_t i = 0; i < tex_names.size(); ++i)
		{
			const ParamMap &tp = rec.textures.at(tex_names[i]);
			if(getStr(tp, "type") != "image") unsupported("texture type '" + getStr(tp, "type") + "' (image textures only)");
			const std::string interp = getStr(tp, "interpolate");
			if(interp == "bicubic" || interp == "mipmap_trilinear" || interp == "mipmap_ewa") unsupported("texture interpolate '" + interp + "'");
			if(get<float>(tp, "adj_saturation", 1.f) != 1.f || get<float>(tp, "adj_hue", 0.f) != 0.f || get<bool>(tp, "normalmap", false)) unsupported("texture hue / saturation adjustment or normal map");
			const std::shared_ptr<Image> img = scene.getImage(getStr(tp, "image_name"));
			if(!img) unsupported("texture '" + tex_names[i] + "' without image");
			const int w = img->getWidth(), h = img->getHeight();
			texels[i].resize((size_t) w * h * 4);
			for(int y = 0; y < h; ++y)
				for(int x = 0; x < w; ++x)
				{
					const Rgba c = img->getColor(x, y);
					float *d = &texels[i][4 * ((size_t) y * w + x)];
					d[0] = c.r_; d[1] = c.g_; d[2] = c.b_; d[3] = c.a_;
				}
			y4gpu_pt_texture &T = textures[i];
			memset(&T, 0, sizeof(T));
			T.texels = texels[i].data(); T.width = w; T.height = h;
			T.clip = lookup(kClip, getStr(tp, "clipping"), 3); T.interpolate = lookup(kInterp, interp, 1);
			T.xrepeat = get<int>(tp, "xrepeat", 1); T.yrepeat = get<int>(tp, "yrepeat", 1);
			T.mirror_x = get<bool>(tp, "mirror_x", false); T.mirror_y = get<bool>(tp, "mirror_y", false); T.rot90 = get<bool>(tp, "rot90", false);
			T.cropmin_x = get<float>(tp, "cropmin_x", 0.f); T.cropmin_y = get<float>(tp, "cropmin_y", 0.f);
			T.cropmax_x = get<float>(tp, "cropmax_x", 1.f); T.cropmax_y = get<float>(tp, "cropmax_y", 1.f);
			T.even_tiles = get<bool>(tp, "even_tiles", false); T.odd_tiles = get<bool>(tp, "odd_tiles", true); T.checker_dist = get<float>(tp, "checker_dist", 0.f);
			T.adj_intensity = get<float>(tp, "adj_intensity", 1.f); T.adj_contrast = get<float>(tp, "adj_contrast", 1.f);
			T.adj_mult_factor_red = get<float>(tp, "adj_mult_factor_red", 1.f); T.adj_mult_factor_green = get<float>(tp, "adj_mult_factor_green", 1.f); T.adj_mult_factor_blue = get<float>(tp, "adj_mult_factor_blue", 1.f);
			T.adj_clamp = get<bool>(tp, "adj_clamp", false);
		}
		// ---- per-triangle shading inputs, in primitive order: vertex normals, orco points, uv values (getSurface, primitive_triangle.cc:97-155)
		const std::vector<const Primitive *> &prims = accel_->primitives();
		const size_t n_tri = prims.size();
		const float nan = std::numeric_limits<float>::quiet_NaN();
		std::vector<float> normals, orco, uv;
		bool any_normals = false;
		normals.assign(n_tri * 9, nan);
		if(!layers.empty()) { orco.assign(n_tri * 9, nan); uv.assign(n_tri * 6, nan); }
		for(size_t i = 0; i < n_tri; ++i)
		{
			const Matrix4 *m = nullptr;
			const FacePrimitive *face = AcceleratorB200::faceOf(prims[i], &m);
			const auto *mesh = dynamic_cast<const MeshObject *>(face->getObject());
			if(!mesh) unsupported("primitive without a mesh object");
			if(mesh->isSmooth() || mesh->hasNormalsExported())
			{
				const Vec3 ng = prims[i]->getGeometricNormal();
				for(int k = 0; k < 3; ++k) { const Vec3 v = face->getVertexNormal(k, ng, m); put3(&normals[9 * i + 3 * k], v); }
				any_normals = true;
			}
			if(!layers.empty())
			{
				if(mesh->hasOrco()) for(int k = 0; k < 3; ++k) put3(&orco[9 * i + 3 * k], face->getOrcoVertex(k));
				if(mesh->hasUv() && face->getUvIndices().size() == 3) for(int k = 0; k < 3; ++k) { const Uv t = face->getVertexUv(k); uv[6 * i + 2 * k] = t.u_; uv[6 * i + 2 * k + 1] = t.v_; }
			}
		}
		// ---- lights: the view's enabled lights in name order (render_view.cc:83-91) ----------------------------------------------------
		std::vector<y4gpu_pt_light> lights;
		std::vector<std::vector<float>> light_tris;
		std::vector<std::string> ibl_light_names;
		for(const auto &kv : render_view->getLights())
		{
			const Light *light = kv.second;
			if(!light->lightEnabled() || light->photonOnly()) continue;
			const auto lp_it = rec.lights.find(kv.first);
			if(lp_it == rec.lights.end()) unsupported("light '" + kv.first + "' without recorded parameters");
			const ParamMap &p = lp_it->second;
			const std::string type = getStr(p, "type");
			y4gpu_pt_light l;
			memset(&l, 0, sizeof(l));
			put3(l.color, getCol(p, "color", Rgb{1.f}));
			l.power = get<float>(p, "power", 1.f); l.cast_shadows = get<bool>(p, "cast_shadows", true);
			if(type == "pointlight") { l.type = Y4GPU_LIGHT_POINT; put3(l.from, getVec(p, "from", Vec3{0.f})); }
			else if(type == "directional")
			{
				l.type = Y4GPU_LIGHT_DIRECTIONAL; l.infinite = get<bool>(p, "infinite", true);
				put3(l.point1, getVec(p, "direction", Vec3{0.f, 0.f, 1.f}));
				l.radius = 1.f;
				if(!l.infinite)
				{
					Vec3 from{0.f};
					if(!p.getParam("from", from)) p.getParam("position", from);
					put3(l.from, from);
					l.radius = get<float>(p, "radius", 1.f);
				}
			}
			else if(type == "sunlight") { l.type = Y4GPU_LIGHT_SUN; l.samples = get<int>(p, "samples", 4); l.angle = get<float>(p, "angle", 0.27f); put3(l.point1, getVec(p, "direction", Vec3{0.f, 0.f, 1.f})); }
			else if(type == "spotlight")
			{
				l.type = Y4GPU_LIGHT_SPOT; l.angle = get<float>(p, "cone_angle", 45.f); l.blend = get<float>(p, "blend", 0.15f);
				if(get<bool>(p, "soft_shadows", false)) { l.type = Y4GPU_LIGHT_SPOT_SOFT; l.samples = get<int>(p, "samples", 8); l.radius = get<float>(p, "shadowFuzzyness", 1.f); }
				put3(l.from, getVec(p, "from", Vec3{0.f})); put3(l.point1, getVec(p, "to", Vec3{0.f, 0.f, -1.f}));
			}
			else if(type == "spherelight") { l.type = Y4GPU_LIGHT_SPHERE; l.samples = get<int>(p, "samples", 4); l.radius = get<float>(p, "radius", 1.f); put3(l.from, getVec(p, "from", Vec3{0.f})); }
			else if(type == "arealight")
			{
				l.type = Y4GPU_LIGHT_AREA; l.samples = get<int>(p, "samples", 4);
				put3(l.from, getVec(p, "corner", Vec3{0.f})); put3(l.point1, getVec(p, "point1", Vec3{0.f})); put3(l.point2, getVec(p, "point2", Vec3{0.f}));
			}
			else if(type == "objectlight" || type == "meshlight" || type == "bgPortalLight")
			{
				const Object *obj = scene.getObject(getStr(p, "object_name"));
				if(!obj || obj->numPrimitives() == 0) unsupported("objectlight '" + kv.first + "': object_name must name a mesh with triangles");
				light_tris.emplace_back();
				std::vector<float> &tv = light_tris.back();
				for(const Primitive *prim : obj->getPrimitives())
				{
					const Matrix4 *m = nullptr;
					const FacePrimitive *face = AcceleratorB200::faceOf(prim, &m);
					if(!face || m || face->getVerticesIndices().size() != 3) unsupported("objectlight '" + kv.first + "': the object must be a plain triangle mesh");
					for(int k = 0; k < 3; ++k) { const Point3 v = face->getVertex(k); tv.push_back(v.x()); tv.push_back(v.y()); tv.push_back(v.z()); }
				}
				l.type = Y4GPU_LIGHT_OBJECT; l.samples = get<int>(p, "samples", 4); l.double_sided = get<bool>(p, "double_sided", false);
				if(type == "bgPortalLight") { l.type = Y4GPU_LIGHT_PORTAL; l.double_sided = 0; if(!background_) unsupported("bgPortalLight without a background"); }
				l.object_n_tris = (uint32_t) (tv.size() / 9);
			}
			else if(type == "bglight")      // created by a background's factory with ibl = true (background_constant.cc:58-69); checked against the rendered background below
			{
				l.type = Y4GPU_LIGHT_BACKGROUND; l.samples = get<int>(p, "samples", 16); l.double_sided = get<bool>(p, "abs_intersect", false);
				ibl_light_names.push_back(kv.first);
			}
			else unsupported("light type '" + type + "'");
			lights.push_back(l);
		}
		for(size_t i = 0, k = 0; i < lights.size(); ++i) if(lights[i].type == Y4GPU_LIGHT_OBJECT || lights[i].type == Y4GPU_LIGHT_PORTAL) lights[i].object_tris = light_tris[k++].data();
		// ---- background ----------------------------------------------------------------------------------------------------------
		bool rendered_background_has_ibl = false;
		std::string ibl_light_name;
		if(background_)
		{
			const ParamMap *bp = background_params;
			if(!bp) unsupported("background without recorded parameters");
			const std::string btype = getStr(*bp, "type");
			const bool sky = btype == "sunsky" || btype == "darksky";
			if(btype != "constant" && btype != "gradientback" && btype != "textureback" && !sky) unsupported("background type '" + btype + "'");
			rendered_background_has_ibl = get<bool>(*bp, sky ? "background_light" : "ibl", false);
			ibl_light_name = btype == "constant" ? "constantBackground_bgLight" : (btype == "gradientback" ? "GradientBackground_bgLight" :
			                 (btype == "sunsky" ? "sunsky_bgLight" : (btype == "darksky" ? "DarkSky_bgLight" : "textureBackground_bgLight")));
			if(btype == "darksky")           // DarkSkyBackground::factory (background_darksky.cc:194-314)
			{
				static const std::map<std::string, int> spaces{{"CIE (E)", 0}, {"CIE (D50)", 1}, {"sRGB (D65)", 2}, {"sRGB (D50)", 3}};
				pr.bg_dark_altitude = get<float>(*bp, "altitude", 0.f); pr.bg_dark_bright = get<float>(*bp, "bright", 1.f);
				pr.bg_dark_exposure = get<float>(*bp, "exposure", 1.f); pr.bg_dark_night = get<bool>(*bp, "night", false) ? 1 : 0;
				pr.bg_dark_color_space = lookup(spaces, getStr(*bp, "color_space", "CIE (E)"), 0);
			}
			if(sky)                          // SunSkyBackground::factory (background_sunsky.cc:226-307); an add_sun light is an ordinary recorded sunlight
			{
				pr.background_gradient = btype == "sunsky" ? 3 : 4;
				put3(pr.bg_sun_from, getVec(*bp, "from", Vec3{1.f, 1.f, 1.f}));
				pr.bg_sun_turbidity = get<float>(*bp, "turbidity", 4.f);
				pr.bg_sun_var[0] = get<float>(*bp, "a_var", 1.f); pr.bg_sun_var[1] = get<float>(*bp, "b_var", 1.f); pr.bg_sun_var[2] = get<float>(*bp, "c_var", 1.f);
				pr.bg_sun_var[3] = get<float>(*bp, "d_var", 1.f); pr.bg_sun_var[4] = get<float>(*bp, "e_var", 1.f);
				pr.bg_sun_power = btype == "sunsky" ? (float) get<double>(*bp, "power", 1.0) : get<float>(*bp, "power", 1.f);
			}
			if(btype == "textureback")       // TextureBackground::factory (background_texture.cc:80-151)
			{
				if(get<float>(*bp, "smartibl_blur", 0.f) > 0.f) unsupported("textureback with smartibl_blur");
				const std::string mapping = getStr(*bp, "mapping");
				pr.background_gradient = 2;
				pr.bg_texture = tex_index.at(getStr(*bp, "texture"));
				pr.bg_texture_angular = (mapping == "probe" || mapping == "angular") ? 1 : 0;
				pr.bg_texture_rotation = get<float>(*bp, "rotation", 0.f);
				pr.bg_texture_power = get<float>(*bp, "power", 1.f);
				pr.bg_ibl_clamp = rendered_background_has_ibl ? get<float>(*bp, "ibl_clamp_sampling", 0.f) : 0.f;
			}
			const float pw = get<float>(*bp, "power", 1.f);
			const Rgb c = getCol(*bp, "color", Rgb{0.f});
			pr.background[0] = c.r_ * pw; pr.background[1] = c.g_ * pw; pr.background[2] = c.b_ * pw;
			if(btype == "gradientback")      // GradientBackground::factory (background_gradient.cc:49-77): ground colours default to the sky's
			{
				const Rgb sh = getCol(*bp, "horizon_color", Rgb{1.f}), sz = getCol(*bp, "zenith_color", Rgb{0.4f, 0.5f, 1.f});
				const Rgb gh = getCol(*bp, "horizon_ground_color", sh), gz = getCol(*bp, "zenith_ground_color", sz);
				pr.background_gradient = 1;
				const Rgb cols[4] = {gz, gh, sz, sh};
				for(int j = 0; j < 4; ++j) { pr.bg_gradient[j][0] = cols[j].r_ * pw; pr.bg_gradient[j][1] = cols[j].g_ * pw; pr.bg_gradient[j][2] = cols[j].b_ * pw; }
			}
		}
		// an IBL light samples the background that created it: supported when that is the rendered background
		for(const std::string &name : ibl_light_names)
			if(!rendered_background_has_ibl || name != ibl_light_name) unsupported("IBL light '" + name + "' of a background other than the rendered one");
		if(!get<bool>(rp, "background_resampling", true)) unsupported("background_resampling = false");
		// ---- render parameters (scene.cc:528-644, imagefilm.cc:45-127) ---------------------------------------------------------------
		pr.width = get<int>(rp, "width", cam.resx); pr.height = get<int>(rp, "height", cam.resy);
		pr.xstart = get<int>(rp, "xstart", 0); pr.ystart = get<int>(rp, "ystart", 0);
		pr.aa_passes = aa_noise_params_.passes_; pr.aa_samples = aa_noise_params_.samples_; pr.aa_inc_samples = aa_noise_params_.inc_samples_;
		pr.aa_sample_multiplier_factor = aa_noise_params_.sample_multiplier_factor_;
		pr.aa_light_sample_multiplier_factor = aa_noise_params_.light_sample_multiplier_factor_;
		pr.aa_pixelwidth = get<float>(rp, "AA_pixelwidth", 1.5f); pr.aa_clamp_samples = aa_noise_params_.clamp_samples_;
		pr.filter_type = lookup(kFilters, getStr(rp, "filter_type", "box"), 0);
		pr.shadow_bias_auto = shadow_bias_auto_; pr.shadow_bias = shadow_bias_;
		pr.min_raydist_auto = ray_min_dist_auto_; pr.min_raydist = ray_min_dist_;
		pr.base_sampling_offset = image_film_->getBaseSamplingOffset();
		pr.rr_seed = (uint32_t) get<int>(rp, "b200_rr_seed", 0x5EED);
		pr.wave_size = (uint32_t) get<int>(rp, "b200_wave_size", 0);
		pr.profile = get<int>(rp, "b200_profile", 0);
		if(pr.width != image_film_->getWidth() || pr.height != image_film_->getHeight()) unsupported("film size differs from the width / height render parameters");
		// ---- hand over -----------------------------------------------------------------------------------------------------------------
		if(y4gpu_pt_set_textures(ctx_, textures.empty() ? nullptr : textures.data(), (uint32_t) textures.size(), layers.empty() ? nullptr : layers.data(), (uint32_t) layers.size(),
		                         orco.empty() ? nullptr : orco.data(), uv.empty() ? nullptr : uv.data()) != Y4GPU_OK
		   || y4gpu_pt_set_scene(ctx_, &cam, mats.data(), (uint32_t) mats.size(), lights.empty() ? nullptr : lights.data(), (uint32_t) lights.size(), &pr, any_normals ? normals.data() : nullptr) != Y4GPU_OK
		   || y4gpu_pt_clear_film(ctx_) != Y4GPU_OK)
		{
			logger_.logError(getName(), ": ", y4gpu_last_error(ctx_));
			return false;
		}
		film_w_ = pr.width; film_h_ = pr.height;
		host_weights_.assign((size_t) film_w_ * film_h_, 0.f); host_rgba_.assign((size_t) film_w_ * film_h_ * 4, 0.f);
	}
	catch(const Fail &f)
	{
		logger_.logError(getName(), ": not supported by the GPU back end (no CPU fallback): ", f.msg);
		return false;
	}
	render_info_ += "B200 wavefront path tracer";
	return true;
}

// TiledIntegrator::renderPass (integrator_tiled.cc:236-267) as one GPU launch sequence.  The film accumulated on the device is
// handed to ImageFilm after every pass so that nextPass (noise detection), autosave and flush keep working on the host.
bool PathIntegratorB200::renderPass(int samples, int offset, bool adaptive, int aa_pass_number)
{
	prePass(samples, (offset + image_film_->getBaseSamplingOffset()), adaptive);
	render_control_.setCurrentPass(aa_pass_number + 1);
	image_film_->setSamplingOffset(offset + samples);
	if(render_control_.canceled()) return false;
	if(samples <= 0) return true;                  // resumed render: pass 0 renders nothing (integrator_tiled.cc:171-175)
	std::vector<uint32_t> pixel_list;
	const uint32_t *list = nullptr;
	if(adaptive)
	{
		// the pixels ImageFilm::nextPass flagged (doMoreSamples, imagefilm.cc:672-675), raster order, film-relative indices
		const int x0 = image_film_->getCx0(), y0 = image_film_->getCy0();
		for(int y = 0; y < film_h_; ++y)
			for(int x = 0; x < film_w_; ++x)
				if(image_film_->doMoreSamples(x0 + x, y0 + y)) pixel_list.push_back((uint32_t) (y * film_w_ + x));
		if(pixel_list.empty()) return true;
		list = pixel_list.data();
	}
	// y4gpu takes the pass offset without the base offset (it adds y4gpu_pt_params::base_sampling_offset itself)
	if(y4gpu_pt_render_pass(ctx_, samples, offset, aa_pass_number, list, pixel_list.size(), 0, 1) != Y4GPU_OK
	   || y4gpu_pt_read_film(ctx_, host_weights_.data(), host_rgba_.data()) != Y4GPU_OK)
	{
		logger_.logError(getName(), ": ", y4gpu_last_error(ctx_));
		return false;
	}
	y4gpu_pt_get_stats(ctx_, &last_stats_);
	image_film_->b200SetAccumulatedFilm(host_weights_.data(), host_rgba_.data());
	if(logger_.isVerbose()) logger_.logVerbose(getName(), ": pass ", aa_pass_number, ": ", last_stats_.samples, " samples, ", last_stats_.rays_closest + last_stats_.rays_shadow, " rays, ", last_stats_.kernel_launches, " kernel launches, ", last_stats_.pass_ms, " ms on the GPU");
	return true;
}

END_YAFARAY
