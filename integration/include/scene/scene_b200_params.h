#pragma once
// The ParamMaps Scene's factories were called with, kept by name.  The B200 back end takes the scene in PARAMETER form
// (include/y4gpu.h: it derives camera axes, light frames, material weights and the film filter itself, with the reference's
// fp32 operation order), and the reference's objects keep their parameters in private members, so the scene remembers them.
#ifndef YAFARAY_SCENE_B200_PARAMS_H
#define YAFARAY_SCENE_B200_PARAMS_H

#include "common/param.h"
#include <list>
#include <map>
#include <string>

BEGIN_YAFARAY

struct B200SceneParams
{
	std::map<std::string, ParamMap> materials, lights, cameras, textures, backgrounds, integrators, render_views;
	std::map<std::string, std::list<ParamMap>> material_nodes;
	ParamMap render;         // Scene::setupSceneRenderParams
};

END_YAFARAY
#endif
