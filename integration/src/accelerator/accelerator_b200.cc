// AcceleratorB200: see include/accelerator/accelerator_b200.h.
#include "accelerator/accelerator_b200.h"
#include "common/logger.h"
#include "common/param.h"
#include "geometry/ray.h"
#include "geometry/surface.h"
#include "geometry/matrix4.h"
#include "geometry/primitive/primitive_face.h"
#include "geometry/primitive/primitive_instance.h"
#include "geometry/object/object_instance.h"
#include "geometry/object/object.h"
#include "material/material.h"
#include <cmath>
#include <cstdlib>
#include <limits>

BEGIN_YAFARAY

const FacePrimitive *AcceleratorB200::faceOf(const Primitive *primitive, const Matrix4 **obj_to_world)
{
	*obj_to_world = nullptr;
	const Primitive *base = primitive;
	if(const auto *inst = dynamic_cast<const PrimitiveInstance *>(primitive))
	{
		base = inst->getBasePrimitive();
		if(const auto *obj = dynamic_cast<const ObjectInstance *>(inst->getObject())) *obj_to_world = obj->getObjToWorldMatrix();
	}
	return dynamic_cast<const FacePrimitive *>(base);
}

// Everything between "a context exists" and "the scene is resident": flag words, material table, geometry upload (two-level when
// the list is made of instance runs) and build.  Used by the factory and by refreshVisibility().
bool AcceleratorB200::uploadScene(Logger &logger, y4gpu_ctx *ctx, const std::vector<const Primitive *> &primitives, std::vector<const Material *> &materials,
                                  std::vector<float> &mat_transparency, bool &ts_varying, std::vector<uint32_t> &flags_out)
{
	const size_t n = primitives.size();
	std::vector<float> verts(n * 9);
	std::vector<uint32_t> &flags = flags_out;
	flags.assign(n, 0u);
	materials.clear();
	std::unordered_map<const Material *, uint32_t> material_index;
	std::vector<const Primitive *> first_with_material;
	for(size_t i = 0; i < n; ++i)
	{
		const Matrix4 *m = nullptr;
		const FacePrimitive *face = faceOf(primitives[i], &m);
		if(!face || face->getVerticesIndices().size() != 3)
		{
			logger.logError("AcceleratorB200: primitive ", i, " is not a triangle (only triangle meshes and their instances are supported)");
			return false;
		}
		// world-space vertices exactly as the reference computes them for every test (FacePrimitive::getVertex: Matrix4 * Point3)
		for(int k = 0; k < 3; ++k)
		{
			const Point3 p = face->getVertex(k, m);
			verts[9 * i + 3 * k] = p.x(); verts[9 * i + 3 * k + 1] = p.y(); verts[9 * i + 3 * k + 2] = p.z();
		}
		const Material *mat = primitives[i]->getMaterial();
		auto it = material_index.find(mat);
		if(it == material_index.end())
		{
			it = material_index.emplace(mat, (uint32_t) materials.size()).first;
			materials.push_back(mat);
			first_with_material.push_back(primitives[i]);
		}
		// visibility matrix of accelerator_kdtree.cc:733-737 (closest), :869-873 (shadow), :1022 (transparent shadow)
		const Visibility vo = primitives[i]->getVisibility(), vm = mat ? mat->getVisibility() : Visibility::NormalVisible;
		const bool seen_o = vo == Visibility::NormalVisible || vo == Visibility::VisibleNoShadows, shad_o = vo == Visibility::NormalVisible || vo == Visibility::InvisibleShadowsOnly;
		const bool seen_m = vm == Visibility::NormalVisible || vm == Visibility::VisibleNoShadows, shad_m = vm == Visibility::NormalVisible || vm == Visibility::InvisibleShadowsOnly;
		uint32_t w = it->second << Y4GPU_TRI_MAT_SHIFT;
		if(seen_o && seen_m) w |= Y4GPU_TRI_VIS_CLOSEST;
		if(shad_o && shad_m) w |= Y4GPU_TRI_VIS_SHADOW;
		if(shad_m) w |= Y4GPU_TRI_VIS_TSHADOW;
		if(mat && mat->isTransparent()) w |= Y4GPU_TRI_TRANSPARENT;
		flags[i] = w;
	}
	// Material::getTransparency of the transparent materials, evaluated by the reference's own code on one surface point.  The
	// device multiplies one colour per material, so a colour that varies over the surface cannot be reproduced: refuse it.
	mat_transparency.assign(materials.size() * 3, 0.f);
	ts_varying = false;
	for(size_t mi = 0; mi < materials.size(); ++mi)
	{
		if(!materials[mi] || !materials[mi]->isTransparent()) continue;
		const Primitive *prim = first_with_material[mi];
		const Matrix4 *m = nullptr;
		const FacePrimitive *face = faceOf(prim, &m);
		const Point3 a = face->getVertex(0, m), b = face->getVertex(1, m), c = face->getVertex(2, m);
		Rgb colour[2];
		for(int probe = 0; probe < 2; ++probe)
		{
			AcceleratorIntersectData d;
			d.hit_ = true; d.t_hit_ = 1.f; d.time_ = 0.f;
			d.barycentric_u_ = probe ? 0.6f : 1.f / 3.f; d.barycentric_v_ = probe ? 0.3f : 1.f / 3.f; d.barycentric_w_ = 1.f - d.barycentric_u_ - d.barycentric_v_;
			const Point3 hit = d.barycentric_u_ * a + d.barycentric_v_ * b + d.barycentric_w_ * c;
			const auto sp = prim->getSurface(nullptr, hit, d, nullptr, nullptr);
			const Vec3 ng = prim->getGeometricNormal();
			Vec3 wo = probe ? (ng + 3.f * (b - a).normalize()) : ng;
			wo.normalize();
			colour[probe] = sp->getTransparency(wo, nullptr);
		}
		if(std::abs(colour[0].r_ - colour[1].r_) > 1e-6f || std::abs(colour[0].g_ - colour[1].g_) > 1e-6f || std::abs(colour[0].b_ - colour[1].b_) > 1e-6f)
		{
			// only intersectTs needs the table: closest-hit / shadow queries (and PathIntegratorB200, which filters shadows on the
			// device from its own per-material table and refuses what that cannot express) are unaffected
			logger.logWarning("AcceleratorB200: a transparent material's shadow colour varies over the surface / with the direction (textured, or Fresnel mirror + transparency): intersectTs (transparent shadows) will fail on this scene");
			ts_varying = true;
		}
		mat_transparency[3 * mi] = colour[0].r_; mat_transparency[3 * mi + 1] = colour[0].g_; mat_transparency[3 * mi + 2] = colour[0].b_;
	}
	// Instances stay instances on the device when the primitive list is what Scene::updateObjects produces for instanced objects:
	// runs of PrimitiveInstance that walk their base object's faces in order under one obj_to_world matrix (primitive_instance.cc,
	// object_instance.cc:25-33).  The meshes go up once in object space with one matrix per run (y4gpu_upload_instanced; results
	// are identical to the flattened upload).  Anything else -- or Y4GPU_INSTANCING=0 -- uploads the world-space triangles.
	bool two_level = false;
	{
		const char *e = getenv("Y4GPU_INSTANCING");
		bool want = !(e && atoi(e) == 0), any_matrix = false;
		bool ok = want;
		std::unordered_map<const Object *, uint32_t> mesh_of;
		std::vector<std::vector<float>> mesh_verts;
		std::vector<std::vector<uint32_t>> mesh_flags;
		std::vector<y4gpu_instance> insts;
		size_t i = 0;
		while(ok && i < n)
		{
			const Matrix4 *m = nullptr;
			const FacePrimitive *face = faceOf(primitives[i], &m);
			const Object *obj = face->getObject();
			const std::vector<const Primitive *> base = obj->getPrimitives();
			const size_t bn = base.size();
			if(bn == 0 || i + bn > n) { ok = false; break; }
			for(size_t k = 0; k < bn && ok; ++k)
			{
				const Matrix4 *mk = nullptr;
				ok = faceOf(primitives[i + k], &mk) == dynamic_cast<const FacePrimitive *>(base[k]) && mk == m;
			}
			if(!ok) break;
			auto it = mesh_of.find(obj);
			if(it == mesh_of.end())
			{
				it = mesh_of.emplace(obj, (uint32_t) mesh_verts.size()).first;
				mesh_verts.emplace_back(9 * bn);
				mesh_flags.emplace_back(flags.begin() + (long) i, flags.begin() + (long) (i + bn));
				std::vector<float> &mv = mesh_verts.back();
				for(size_t k = 0; k < bn; ++k)
				{
					const FacePrimitive *f = dynamic_cast<const FacePrimitive *>(base[k]);
					for(int c = 0; c < 3; ++c) { const Point3 p = f->getVertex(c); mv[9 * k + 3 * c] = p.x(); mv[9 * k + 3 * c + 1] = p.y(); mv[9 * k + 3 * c + 2] = p.z(); }
				}
			}
			else
			{
				const std::vector<uint32_t> &mf = mesh_flags[it->second];      // the flag words are per mesh triangle: every run must agree
				for(size_t k = 0; k < bn && ok; ++k) ok = mf[k] == flags[i + k];
				if(!ok) break;
			}
			y4gpu_instance in;
			in.mesh = it->second; in.identity = m ? 0u : 1u;
			const float id[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
			for(int r = 0; r < 3; ++r) for(int c = 0; c < 4; ++c) in.obj_to_world[4 * r + c] = m ? (*m)[r][c] : id[4 * r + c];
			insts.push_back(in);
			any_matrix = any_matrix || m;
			i += bn;
		}
		if(ok && any_matrix)
		{
			std::vector<y4gpu_mesh> ms(mesh_verts.size());
			for(size_t k = 0; k < ms.size(); ++k) { ms[k].v_xyz = mesh_verts[k].data(); ms[k].flags = mesh_flags[k].data(); ms[k].n_tri = mesh_flags[k].size(); }
			two_level = y4gpu_upload_instanced(ctx, ms.data(), (uint32_t) ms.size(), insts.data(), (uint32_t) insts.size()) == Y4GPU_OK
			            && y4gpu_build(ctx, nullptr) == Y4GPU_OK;
			if(!two_level) logger.logWarning("AcceleratorB200: two-level upload failed (", y4gpu_last_error(ctx), "), uploading world-space triangles");
			else logger.logInfo("AcceleratorB200: ", insts.size(), " instances of ", ms.size(), " meshes kept as instances on the device");
		}
	}
	if(!two_level && (y4gpu_upload_triangles(ctx, verts.data(), flags.data(), n) != Y4GPU_OK || y4gpu_build(ctx, nullptr) != Y4GPU_OK))
	{
		logger.logError("AcceleratorB200: ", y4gpu_last_error(ctx));
		return false;
	}
	return true;
}

// visibility word of one primitive: the matrix of accelerator_kdtree.cc:733-737 (closest), :869-873 (shadow), :1022 (transparent shadow)
static uint32_t visibilityBits(const Primitive *primitive)
{
	const Material *mat = primitive->getMaterial();
	const Visibility vo = primitive->getVisibility(), vm = mat ? mat->getVisibility() : Visibility::NormalVisible;
	const bool seen_o = vo == Visibility::NormalVisible || vo == Visibility::VisibleNoShadows, shad_o = vo == Visibility::NormalVisible || vo == Visibility::InvisibleShadowsOnly;
	const bool seen_m = vm == Visibility::NormalVisible || vm == Visibility::VisibleNoShadows, shad_m = vm == Visibility::NormalVisible || vm == Visibility::InvisibleShadowsOnly;
	uint32_t w = 0;
	if(seen_o && seen_m) w |= Y4GPU_TRI_VIS_CLOSEST;
	if(shad_o && shad_m) w |= Y4GPU_TRI_VIS_SHADOW;
	if(shad_m) w |= Y4GPU_TRI_VIS_TSHADOW;
	return w;
}

// The reference reads visibilities at query time; the device bakes them into the flag words at upload.  Scene::render builds the
// accelerator BEFORE it initialises the lights (scene.cc:218-220), and a light may change an object's visibility there
// (BackgroundPortalLight::init hides its mesh): the integrator calls this from preprocess() and the scene goes up again if a
// visibility changed in between.
bool AcceleratorB200::refreshVisibility() const
{
	const uint32_t mask = Y4GPU_TRI_VIS_CLOSEST | Y4GPU_TRI_VIS_SHADOW | Y4GPU_TRI_VIS_TSHADOW;
	bool changed = false;
	for(size_t i = 0; i < primitives_.size() && !changed; ++i) changed = (flags_[i] & mask) != visibilityBits(primitives_[i]);
	if(!changed) return true;
	logger_.logInfo("AcceleratorB200: object visibilities changed after the build (a light hid its mesh): uploading the scene again");
	return uploadScene(logger_, ctx_, primitives_, materials_, mat_transparency_, ts_varying_, flags_);
}

const Accelerator *AcceleratorB200::factory(Logger &logger, const std::vector<const Primitive *> &primitives, const ParamMap &params)
{
	int device = 0;
	params.getParam("b200_device", device);
	if(const char *e = getenv("Y4GPU_DEVICE")) device = atoi(e);
	y4gpu_ctx *ctx = nullptr;
	if(y4gpu_create(device, &ctx) != Y4GPU_OK)
	{
		logger.logError("AcceleratorB200: ", y4gpu_last_error(nullptr));
		return nullptr;
	}
	std::vector<const Material *> materials;
	std::vector<float> mat_transparency;
	std::vector<uint32_t> flags;
	bool ts_varying = false;
	const size_t n = primitives.size();
	if(!uploadScene(logger, ctx, primitives, materials, mat_transparency, ts_varying, flags))
	{
		y4gpu_destroy(ctx);
		return nullptr;
	}
	float b[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
	if(n) y4gpu_get_bound(ctx, b);
	y4gpu_stats st;
	y4gpu_get_stats(ctx, &st);
	logger.logInfo("AcceleratorB200: ", n, " triangles, ", st.n_nodes, " BVH8 nodes in ", st.bvh_levels, " levels, built in ", st.build_ms, " ms on the GPU");
	auto *acc = new AcceleratorB200(logger, ctx, primitives, std::move(materials), std::move(mat_transparency), Bound{{b[0], b[1], b[2]}, {b[3], b[4], b[5]}});
	acc->ts_varying_ = ts_varying;
	acc->flags_ = std::move(flags);
	return acc;
}

AcceleratorB200::~AcceleratorB200() { y4gpu_destroy(ctx_); }

static inline y4gpu_ray toRay(const Ray &ray, float tmin, float tmax)
{
	return y4gpu_ray{ray.from_.x(), ray.from_.y(), ray.from_.z(), tmin, ray.dir_.x(), ray.dir_.y(), ray.dir_.z(), tmax};
}

static inline void fill(AcceleratorIntersectData &d, const y4gpu_hit &h, const Ray &ray, const std::vector<const Primitive *> &prims)
{
	d.hit_ = true;
	d.t_hit_ = h.t;
	d.barycentric_u_ = 1.f - h.bu - h.bv;     // primitive_triangle.cc:66-68
	d.barycentric_v_ = h.bu;
	d.barycentric_w_ = h.bv;
	d.time_ = ray.time_;
	d.hit_primitive_ = prims[h.prim];
}

// intersect(ray, t_max): accept ray.tmin <= t < t_max (accelerator_kdtree.cc:731) == Y4GPU_TRACE_CLOSEST with tmax = t_max
AcceleratorIntersectData AcceleratorB200::intersect(const Ray &ray, float t_max) const
{
	AcceleratorIntersectData d;
	const y4gpu_ray r = toRay(ray, ray.tmin_, t_max);
	y4gpu_hit h;
	if(y4gpu_trace(ctx_, Y4GPU_TRACE_CLOSEST, &r, 1, &h) != Y4GPU_OK) { logger_.logError("AcceleratorB200::intersect: ", y4gpu_last_error(ctx_)); return d; }
	if(h.prim < 0) return d;
	fill(d, h, ray, primitives_);
	d.t_max_ = h.t;
	return d;
}

// intersectS(sray, t_max, bias): the wrapper advanced the origin already; accept 0 <= t < t_max (:867) == Y4GPU_TRACE_SHADOW
// with tmin = 0 (no further shift: from + dir * 0) and tmax = t_max
AcceleratorIntersectData AcceleratorB200::intersectS(const Ray &ray, float t_max, float) const
{
	AcceleratorIntersectData d;
	const y4gpu_ray r = toRay(ray, 0.f, t_max);
	y4gpu_hit h;
	if(y4gpu_trace(ctx_, Y4GPU_TRACE_SHADOW, &r, 1, &h) != Y4GPU_OK) { logger_.logError("AcceleratorB200::intersectS: ", y4gpu_last_error(ctx_)); return d; }
	if(h.prim < 0) return d;
	fill(d, h, ray, primitives_);
	return d;
}

// intersectTs(sray, max_depth, dist, bias, camera): accept sray.tmin <= t < dist (:1019); opaque blocker or more than max_depth
// distinct transparent primitives => shadowed, else the product of their transparency colours
AcceleratorTsIntersectData AcceleratorB200::intersectTs(const Ray &ray, int max_depth, float dist, float, const Camera *) const
{
	AcceleratorTsIntersectData d;
	if(ts_varying_)
	{
		// no CPU fallback and no approximation: the query fails loudly and lets no light through
		if(!ts_error_logged_.exchange(true)) logger_.logError("AcceleratorB200::intersectTs: a transparent material's shadow colour is not constant per material on this scene: not supported by the GPU back end");
		d.hit_ = true;
		return d;
	}
	const y4gpu_ray r = toRay(ray, ray.tmin_, dist);
	y4gpu_tshit h;
	if(y4gpu_trace_tshadow_inner(ctx_, &r, 1, max_depth, mat_transparency_.data(), (uint32_t) materials_.size(), &h) != Y4GPU_OK)
	{
		logger_.logError("AcceleratorB200::intersectTs: ", y4gpu_last_error(ctx_));
		d.hit_ = true;      // an error must not let light through
		return d;
	}
	d.transparent_color_ = Rgb{h.filter_r, h.filter_g, h.filter_b};
	if(h.hit.prim < 0) return d;
	fill(d, h.hit, ray, primitives_);
	return d;
}

END_YAFARAY
