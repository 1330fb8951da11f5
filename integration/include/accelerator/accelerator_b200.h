#pragma once
// libYafaRay-side plug-in: the B200 BVH8 accelerator behind the reference's Accelerator interface
// (include/accelerator/accelerator.h:53-69).  Registered in Accelerator::factory as "yafaray-b200-bvh8"
// (src/accelerator/accelerator.cc, see integration/apply_patch.py).  All GPU work goes through the C-ABI of
// liby4gpu.so (include/y4gpu.h); this class only converts between the reference's types and the POD records.
#ifndef YAFARAY_ACCELERATOR_B200_H
#define YAFARAY_ACCELERATOR_B200_H

#include "accelerator/accelerator.h"
#include "geometry/bound.h"
#include "y4gpu.h"
#include <atomic>
#include <vector>
#include <unordered_map>

BEGIN_YAFARAY

class Material;
class FacePrimitive;
class Matrix4;

class AcceleratorB200 final : public Accelerator
{
	public:
		static const Accelerator *factory(Logger &logger, const std::vector<const Primitive *> &primitives, const ParamMap &params);
		~AcceleratorB200() override;
		// The single-ray virtuals are 1-ray batches through the same kernels (re-entrant: y4gpu_trace with host buffers runs on a
		// private slot per call).  They keep every caller of the reference working; the hot path is PathIntegratorB200, which
		// submits whole wavefronts through context().
		AcceleratorIntersectData intersect(const Ray &ray, float t_max) const override;
		AcceleratorIntersectData intersectS(const Ray &ray, float t_max, float shadow_bias) const override;
		AcceleratorTsIntersectData intersectTs(const Ray &ray, int max_depth, float dist, float shadow_bias, const Camera *camera) const override;
		Bound getBound() const override { return bound_; }

		y4gpu_ctx *context() const { return ctx_; }
		const std::vector<const Primitive *> &primitives() const { return primitives_; }
		const std::vector<const Material *> &materialTable() const { return materials_; }   // index = material id in the triangle flag words
		// triangle `i` of the device == primitives()[i]: its mesh face and (for instances) the object-to-world matrix
		static const FacePrimitive *faceOf(const Primitive *primitive, const Matrix4 **obj_to_world);
		// re-upload when an object's visibility changed after the build (lights are initialised after the accelerator: scene.cc:218-220)
		bool refreshVisibility() const;

	private:
		AcceleratorB200(Logger &logger, y4gpu_ctx *ctx, const std::vector<const Primitive *> &primitives, std::vector<const Material *> materials, std::vector<float> mat_transparency, const Bound &bound)
			: Accelerator(logger), ctx_(ctx), primitives_(primitives), materials_(std::move(materials)), mat_transparency_(std::move(mat_transparency)), bound_(bound) { }
		static bool uploadScene(Logger &logger, y4gpu_ctx *ctx, const std::vector<const Primitive *> &primitives, std::vector<const Material *> &materials,
		                        std::vector<float> &mat_transparency, bool &ts_varying, std::vector<uint32_t> &flags_out);
		y4gpu_ctx *ctx_ = nullptr;
		std::vector<const Primitive *> primitives_;     // the factory's vector is a temporary: keep a copy (index -> pointer)
		mutable std::vector<const Material *> materials_;
		mutable std::vector<float> mat_transparency_;   // [n_mat * 3] Material::getTransparency colour of the transparent materials
		mutable std::vector<uint32_t> flags_;           // the flag words as uploaded
		mutable bool ts_varying_ = false;               // some transparent material's colour is not one constant: intersectTs refuses
		mutable std::atomic<bool> ts_error_logged_{false};
		Bound bound_;
};

END_YAFARAY
#endif
