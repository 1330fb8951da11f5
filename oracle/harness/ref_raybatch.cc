// TEST INFRASTRUCTURE ONLY (oracle): runs the UNMODIFIED reference accelerator
// (libYafaRay kd-tree, linked from oracle/_ref/libyafaray4_ref.a) on a ray batch.
//
// This file is our own code; it only *calls* reference classes:
//   yafaray::Scene geometry calls            (/root/reference/src/scene/scene.cc:934-1004)
//   Accelerator::factory                     (src/accelerator/accelerator.cc:32-53)
//   Accelerator::intersect(ray, t_max)       (src/accelerator/accelerator_kdtree.cc:634-773)
//   Accelerator::isShadowed(ray, bias)       (src/accelerator/accelerator.cc:69-80 -> intersectS :775-904)
//
// usage: ref_raybatch <tris.f32> <rays.f32> <closest|shadow> <hits.out> [threads] [accel_type]
//   tris.f32 : n_tri * 9 floats (v0 v1 v2, world space); triangle i == primitive index i
//   rays.f32 : n_ray * 8 floats (ox oy oz tmin dx dy dz tmax); tmax < 0 means "no limit"
//   hits.out : n_ray * {float t, float bu, float bv, int32 prim}; prim = -1 on miss.
//              (bu,bv) = Moller-Trumbore (u',v') == reference barycentric_v_/barycentric_w_
//              closest: t = hit distance; shadow: t = 0, prim = first blocker found.
// Prints one JSON line with build and trace timings.
#include "scene/scene.h"
#include "accelerator/accelerator.h"
#include "common/logger.h"
#include "common/param.h"
#include "geometry/ray.h"
#include "geometry/bound.h"
#include "geometry/object/object.h"
#include "geometry/primitive/primitive.h"
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
	const std::vector<float> rays = readFloats(argv[2]);
	const bool shadow = strcmp(argv[3], "shadow") == 0;
	int n_threads = argc > 5 ? atoi(argv[5]) : (int) std::thread::hardware_concurrency();
	if(n_threads < 1) n_threads = 1;
	const std::string accel_type = argc > 6 ? argv[6] : "yafaray-kdtree-original";
	const size_t n_tri = tris.size() / 9, n_ray = rays.size() / 8;

	Logger logger(nullptr, nullptr, YAFARAY_DISPLAY_CONSOLE_HIDDEN);
	logger.setConsoleMasterVerbosity(YAFARAY_LOG_LEVEL_MUTE);
	logger.setLogMasterVerbosity(YAFARAY_LOG_LEVEL_MUTE);
	Scene scene(logger);
	scene.startObjects();
	ParamMap op;
	op["type"] = std::string("mesh");
	op["num_vertices"] = (int) (n_tri * 3);
	op["num_faces"] = (int) n_tri;
	Object *obj = scene.createObject("s", op);
	if(!obj) { fprintf(stderr, "createObject failed\n"); return 3; }
	for(size_t i = 0; i < n_tri * 3; ++i) scene.addVertex({tris[3 * i], tris[3 * i + 1], tris[3 * i + 2]});
	for(size_t i = 0; i < n_tri; ++i) scene.addFace({(int) (3 * i), (int) (3 * i + 1), (int) (3 * i + 2)});
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

	std::vector<HitRec> out(n_ray);
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
				if(shadow)
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
	const auto t3 = std::chrono::steady_clock::now();

	FILE *fo = fopen(argv[4], "wb");
	if(!fo) { fprintf(stderr, "cannot write %s\n", argv[4]); return 2; }
	fwrite(out.data(), sizeof(HitRec), out.size(), fo);
	fclose(fo);
	size_t n_hit = 0;
	for(const auto &h : out) n_hit += h.prim != -1;
	const double build_s = std::chrono::duration<double>(t1 - t0).count();
	const double trace_s = std::chrono::duration<double>(t3 - t2).count();
	printf("{\"impl\": \"reference\", \"accel\": \"%s\", \"kind\": \"%s\", \"n_tri\": %zu, \"n_ray\": %zu, \"n_hit\": %zu, "
		   "\"threads\": %d, \"build_s\": %.6f, \"trace_s\": %.6f, \"mrays_per_s\": %.6f, "
		   "\"bound\": [%.9g, %.9g, %.9g, %.9g, %.9g, %.9g]}\n",
		   accel_type.c_str(), shadow ? "shadow" : "closest", n_tri, n_ray, n_hit, n_threads, build_s, trace_s,
		   n_ray / trace_s * 1e-6, bound.a_.x(), bound.a_.y(), bound.a_.z(), bound.g_.x(), bound.g_.y(), bound.g_.z());
	return 0;
}
