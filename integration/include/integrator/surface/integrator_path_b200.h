#pragma once
// libYafaRay-side plug-in: PathIntegrator / DirectLightIntegrator on the B200 wavefront path tracer.
// Selected in Integrator::factory (src/integrator/integrator.cc:37-60) by `type = pathtracing | directlighting` together with
// `backend = b200` (or the environment variable YAFARAY_B200_BACKEND=1 for unmodified clients).  The reference keeps the pass
// loop (TiledIntegrator::render, integrator_tiled.cc:97-234), the adaptive-AA flagging (ImageFilm::nextPass) and the film
// flush; renderPass(samples, offset, adaptive, aa_pass_number) (integrator_tiled.h:60) becomes one launch sequence on the GPU.
#ifndef YAFARAY_INTEGRATOR_PATH_B200_H
#define YAFARAY_INTEGRATOR_PATH_B200_H

#include "integrator/surface/integrator_tiled.h"
#include "common/param.h"
#include "y4gpu.h"
#include <vector>

BEGIN_YAFARAY

class AcceleratorB200;

class PathIntegratorB200 final : public TiledIntegrator
{
	public:
		static bool requested(const ParamMap &params);
		static Integrator *factory(Logger &logger, const ParamMap &params, const Scene &scene, RenderControl &render_control);
		bool preprocess(ImageFilm *image_film, const RenderView *render_view, const Scene &scene) override;
		bool renderPass(int samples, int offset, bool adaptive, int aa_pass_number) override;
		// TiledIntegrator::renderTile is the only caller of integrate() and renderPass replaces it; a direct call has no
		// single-ray counterpart on the device and reports an error instead of approximating
		std::pair<Rgb, float> integrate(Ray &ray, RandomGenerator &random_generator, ColorLayers *color_layers, int thread_id, int ray_level, bool chromatic_enabled, float wavelength, int additional_depth, const RayDivision &ray_division, const PixelSamplingData &pixel_sampling_data) const override;
		std::string getShortName() const override { return "PT-B200"; }
		std::string getName() const override { return "PathTracerB200"; }
		const y4gpu_pt_stats &lastPassStats() const { return last_stats_; }

	private:
		PathIntegratorB200(RenderControl &render_control, Logger &logger, const ParamMap &params) : TiledIntegrator(render_control, logger), params_(params) { }
		ParamMap params_;                      // the integrator's own ParamMap (type, bounces, path_samples, ...)
		const AcceleratorB200 *accel_ = nullptr;
		y4gpu_ctx *ctx_ = nullptr;
		int film_w_ = 0, film_h_ = 0;
		std::vector<float> host_weights_, host_rgba_;
		y4gpu_pt_stats last_stats_ = {};
};

END_YAFARAY
#endif
