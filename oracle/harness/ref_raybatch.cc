// TEST INFRASTRUCTURE ONLY (oracle): runs the UNMODIFIED reference accelerator
// (libYafaRay kd-tree, linked from oracle/_ref/libyafaray4_ref.a) on a ray batch.
//
// This file is our own code; it only *calls* reference classes:
//   yafaray::Scene geometry calls            (/root/reference/src/scene/scene.cc:934-1004)
//   Accelerator::factory                     (src/accelerator/accelerator.cc:32-53)
//   Accelerator::intersect(ray, t_max)       (src/accelerator/accelerator_kdtree.cc:634-773)
//   Accelerator::isShadowed(ray, bias)       (src/accelerator/accelerator.cc:69-80 -> intersectS :775-904)
//
//   Accelerator::isShadowed(ray, depth, bias, cam) (src/accelerator/accelerator.cc:82-94 -> intersectTs :911-1061)
//
// usage: ref_raybatch <tris.f32> <rays.f32> <closest|shadow|tshadow> <hits.out> [threads] [accel_type]
//        ref_raybatch <tris.f32> --serve - - [threads] [accel_type]
//          serve mode: builds the kd-tree once, prints one JSON line, then answers each stdin line
//          "<closest|shadow> <rays.f32> <hits.out> [repeats]" with one JSON timing line (used by bench.py's CPU legs)
//   tris.f32 : n_tri * 9 floats (v0 v1 v2, world space); triangle i == primitive index i
//   rays.f32 : n_ray * 8 floats (ox oy oz tmin dx dy dz tmax); tmax < 0 means "no limit"
//   hits.out : n_ray * {float t, float bu, float bv, int32 prim}; prim = -1 on miss.
//              (bu,bv) = Moller-Trumbore (u',v') == reference barycentric_v_/barycentric_w_
//              closest: t = hit distance; shadow: t = 0, prim = first blocker found.
//              tshadow: (t, bu, bv) = the filter colour (r, g, b) of the transparent blockers, prim = blocker when shadowed.
//   tshadow needs materials: environment REF_RAYBATCH_MATS = text file, one shinydiffusemat per line
//              "transparency r g b transmit_filter"; REF_RAYBATCH_TRIMAT = int32 material index per triangle;
//              REF_RAYBATCH_TS_DEPTH = max_depth (default 4).
// Prints one JSON line with build and trace timings.
#include "scene/scene.h"
#include "accelerator/accelerator.h"
#include "common/logger.h"
#include "common/param.h"
#include "geometry/ray.h"
#include "geometry/bound.h"
#include "geometry/object/object.h"
#include "geometry/primitive/primitive.h"
#include "material/material.h"
#include "color/color.h"
#include <tuple>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <unordered_map>
#include <vector>

using namespace yafaray;

static std::vector<float> readFloats(const char *path)
{
	FILE *f = fopen(path, "rb");
	if(!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
	fseek(f, 0, SEEK_END);
	const long sz = ftell(f);
	fseek(f, 0, SEEK_SET);
	std::vector<float> v(sz / 4);
	if(fread(v.data(), 4, v.size(), f) != v.size()) { fprintf(stderr, "short read %s\n", path); exit(2); }
	fclose(f);
	return v;
}

struct HitRec { float t, bu, bv; int32_t prim; };

int main(int argc, char **argv)
{
	if(argc < 5) { fprintf(stderr, "usage: %s tris rays closest|shadow out [threads] [accel]\n", argv[0]); return 2; }
	const std::vector<float> tris = readFloats(argv[1]);
	const bool serve = strcmp(argv[2], "--serve") == 0;
	int n_threads = argc > 5 ? atoi(argv[5]) : (int) std::thread::hardware_concurrency();
	if(n_threads < 1) n_threads = (int) std::thread::hardware_concurrency();
	if(n_threads < 1) n_threads = 1;
	const std::string accel_type = argc > 6 ? argv[6] : "yafaray-kdtree-original";
	const size_t n_tri = tris.size() / 9;

	Logger logger(nullptr, nullptr, YAFARAY_DISPLAY_CONSOLE_HIDDEN);
	logger.setConsoleMasterVerbosity(getenv("REF_RAYBATCH_VERBOSE") ? YAFARAY_LOG_LEVEL_VERBOSE : YAFARAY_LOG_LEVEL_MUTE);
	logger.setLogMasterVerbosity(YAFARAY_LOG_LEVEL_MUTE);
	Scene scene(logger);
	scene.startObjects();
	ParamMap op;
	op["type"] = std::string("mesh");
	op["num_vertices"] = (int) (n_tri * 3);
	op["num_faces"] = (int) n_tri;
	// optional materials (transparent-shadow batches)
	std::vector<const std::unique_ptr<const Material> *> mats;
	std::vector<int32_t> tri_mat;
	if(const char *mf = getenv("REF_RAYBATCH_MATS"))
	{
		FILE *f = fopen(mf, "r");
		if(!f) { fprintf(stderr, "cannot open %s\n", mf); return 2; }
		float tr, r, g, b, tf;
		while(fscanf(f, "%f %f %f %f %f", &tr, &r, &g, &b, &tf) == 5)
		{
			ParamMap mp;
			mp["type"] = std::string("shinydiffusemat");
			mp["color"] = Rgba(r, g, b, 1.f);
			mp["transparency"] = tr;
			mp["transmit_filter"] = tf;
			const std::string name = "m" + std::to_string(mats.size());
			mats.push_back(scene.createMaterial(name, mp, {}));
			if(!mats.back()) { fprintf(stderr, "createMaterial failed\n"); return 3; }
		}
		fclose(f);
		const char *tf_path = getenv("REF_RAYBATCH_TRIMAT");
		if(!tf_path) { fprintf(stderr, "REF_RAYBATCH_TRIMAT missing\n"); return 2; }
		FILE *g2 = fopen(tf_path, "rb");
		tri_mat.resize(n_tri);
		if(!g2 || fread(tri_mat.data(), 4, n_tri, g2) != n_tri) { fprintf(stderr, "short read %s\n", tf_path); return 2; }
		fclose(g2);
	}
	const int ts_depth = getenv("REF_RAYBATCH_TS_DEPTH") ? atoi(getenv("REF_RAYBATCH_TS_DEPTH")) : 4;
	Object *obj = scene.createObject("s", op);
	if(!obj) { fprintf(stderr, "createObject failed\n"); return 3; }
	for(size_t i = 0; i < n_tri * 3; ++i) scene.addVertex({tris[3 * i], tris[3 * i + 1], tris[3 * i + 2]});
	for(size_t i = 0; i < n_tri; ++i)
	{
		if(!tri_mat.empty()) scene.setCurrentMaterial(mats.at(tri_mat[i]));
		scene.addFace({(int) (3 * i), (int) (3 * i + 1), (int) (3 * i + 2)});
	}
	scene.endObject();
	scene.endObjects();

	const std::vector<const Primitive *> prims = obj->getPrimitives();
	std::unordered_map<const Primitive *, int32_t> index_of;
	index_of.reserve(prims.size() * 2);
	for(size_t i = 0; i < prims.size(); ++i) index_of[prims[i]] = (int32_t) i;

	ParamMap ap;
	ap["type"] = accel_type;
	ap["num_primitives"] = (int) prims.size();
	ap["accelerator_threads"] = n_threads;
	const auto t0 = std::chrono::steady_clock::now();
	std::unique_ptr<const Accelerator> accel(Accelerator::factory(logger, prims, ap));
	const auto t1 = std::chrono::steady_clock::now();
	const Bound bound = accel->getBound();
	const double build_s = std::chrono::duration<double>(t1 - t0).count();

	// traces one batch with n_threads host threads; returns seconds
	auto trace = [&](const std::vector<float> &rays, int mode /* 0 closest, 1 shadow, 2 transparent shadow */, std::vector<HitRec> &out) -> double
	{
		const bool shadow = mode == 1;
		const size_t n_ray = rays.size() / 8;
		out.resize(n_ray);
		auto worker = [&](int tid)
		{
			// block-cyclic over the batch so every thread sees similar coherence
			const size_t block = 4096;
			for(size_t b = (size_t) tid * block; b < n_ray; b += (size_t) n_threads * block)
			{
				const size_t e = std::min(n_ray, b + block);
				for(size_t i = b; i < e; ++i)
				{
					const float *r = &rays[8 * i];
					Ray ray({r[0], r[1], r[2]}, {r[4], r[5], r[6]}, r[3], r[7]);
					HitRec h{0.f, 0.f, 0.f, -1};
					if(mode == 2)
					{
						const auto res = accel->isShadowed(ray, ts_depth, 0.f, nullptr);
						const Rgb c = std::get<1>(res);
						h.t = c.r_; h.bu = c.g_; h.bv = c.b_;
						if(std::get<0>(res)) h.prim = std::get<2>(res) ? index_of[std::get<2>(res)] : -2;
					}
					else if(shadow)
					{
						const auto res = accel->isShadowed(ray, 0.f);
						if(res.first) h.prim = res.second ? index_of[res.second] : -2;
					}
					else
					{
						const float t_max = (ray.tmax_ >= 0.f) ? ray.tmax_ : std::numeric_limits<float>::infinity();
						const AcceleratorIntersectData d = accel->intersect(ray, t_max);
						if(d.hit_ && d.hit_primitive_)
						{
							h.t = d.t_hit_; h.bu = d.barycentric_v_; h.bv = d.barycentric_w_;
							h.prim = index_of[d.hit_primitive_];
						}
					}
					out[i] = h;
				}
			}
		};
		const auto t2 = std::chrono::steady_clock::now();
		std::vector<std::thread> th;
		for(int t = 0; t < n_threads; ++t) th.emplace_back(worker, t);
		for(auto &t : th) t.join();
		return std::chrono::duration<double>(std::chrono::steady_clock::now() - t2).count();
	};
	auto writeHits = [](const char *path, const std::vector<HitRec> &out) -> bool
	{
		FILE *fo = fopen(path, "wb");
		if(!fo) { fprintf(stderr, "cannot write %s\n", path); return false; }
		fwrite(out.data(), sizeof(HitRec), out.size(), fo);
		fclose(fo);
		return true;
	};
	auto report = [&](const char *kind, size_t n_ray, const std::vector<HitRec> &out, double trace_s, int repeats)
	{
		size_t n_hit = 0;
		for(const auto &h : out) n_hit += h.prim != -1;
		printf("{\"impl\": \"reference\", \"accel\": \"%s\", \"kind\": \"%s\", \"n_tri\": %zu, \"n_ray\": %zu, \"n_hit\": %zu, "
			   "\"threads\": %d, \"build_s\": %.6f, \"trace_s\": %.6f, \"repeats\": %d, \"mrays_per_s\": %.6f, "
			   "\"bound\": [%.9g, %.9g, %.9g, %.9g, %.9g, %.9g]}\n",
			   accel_type.c_str(), kind, n_tri, n_ray, n_hit, n_threads, build_s, trace_s, repeats,
			   trace_s > 0. ? n_ray / trace_s * 1e-6 : 0., bound.a_.x(), bound.a_.y(), bound.a_.z(), bound.g_.x(), bound.g_.y(), bound.g_.z());
		fflush(stdout);
	};

	std::vector<HitRec> out;
	if(!serve)
	{
		const std::vector<float> rays = readFloats(argv[2]);
		const int mode = strcmp(argv[3], "tshadow") == 0 ? 2 : (strcmp(argv[3], "shadow") == 0 ? 1 : 0);
		const double trace_s = trace(rays, mode, out);
		if(!writeHits(argv[4], out)) return 2;
		report(argv[3], rays.size() / 8, out, trace_s, 1);
		return 0;
	}
	report("build", 0, out, 0., 0);
	char kind[32], rays_path[4096], out_path[4096];
	char line[9000];
	while(fgets(line, sizeof(line), stdin))
	{
		int repeats = 1;
		const int got = sscanf(line, "%31s %4095s %4095s %d", kind, rays_path, out_path, &repeats);
		if(got < 3) break;
		if(repeats < 1) repeats = 1;
		const std::vector<float> rays = readFloats(rays_path);
		const int mode = strcmp(kind, "tshadow") == 0 ? 2 : (strcmp(kind, "shadow") == 0 ? 1 : 0);
		double best = 1e30;   // serve mode reports the mean over repeats in trace_s
		double sum = 0.;
		for(int r = 0; r < repeats; ++r) { const double s = trace(rays, mode, out); sum += s; if(s < best) best = s; }
		if(!writeHits(out_path, out)) return 2;
		report(kind, rays.size() / 8, out, sum / repeats, repeats);
	}
	return 0;
}
