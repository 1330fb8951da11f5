#!/usr/bin/env python3
"""Applies the reference-side edits of the B200 integration to a COPY of the libYafaRay tree (never to /root/reference).

usage: apply_patch.py <tree>      (tree = directory holding the copied src/ and include/)

Every edit is an exact-anchor insertion / replacement and fails loudly when its anchor is missing, so a changed
reference cannot be patched half-way.  `make -C integration diff` writes the result as a unified diff
(integration/reference_b200.diff) for review; INTEGRATION.md walks through it."""
import sys
from pathlib import Path


def edit(tree: Path, rel: str, anchor: str, new: str, mode: str = "after", count: int = 1) -> None:
    p = tree / rel
    s = p.read_text()
    if s.count(anchor) != count:
        raise SystemExit(f"apply_patch: anchor occurs {s.count(anchor)}x (expected {count}) in {rel}: {anchor!r}")
    if mode == "after":
        s = s.replace(anchor, anchor + new)
    elif mode == "before":
        s = s.replace(anchor, new + anchor)
    elif mode == "replace":
        s = s.replace(anchor, new)
    else:
        raise SystemExit(mode)
    p.write_text(s)


def main(tree: Path) -> None:
    # ---- Accelerator seam (include/accelerator/accelerator.h:56, src/accelerator/accelerator.cc:42-51) -----------------
    edit(tree, "src/accelerator/accelerator.cc", '#include "accelerator/accelerator_simple_test.h"\n',
         '#include "accelerator/accelerator_b200.h"\n')
    edit(tree, "src/accelerator/accelerator.cc",
         '\telse if(type == "yafaray-simpletest") accelerator = AcceleratorSimpleTest::factory(logger, primitives_list, params);\n',
         '\telse if(type == "yafaray-b200-bvh8")\n'
         '\t{\n'
         '\t\t// no CPU fallback for this type: a failed GPU set-up is an error, not a silent kd-tree (the caller gets an accelerator\n'
         '\t\t// that reports every ray as blocked / missing nothing would be wrong too, so the scene update fails instead)\n'
         '\t\taccelerator = AcceleratorB200::factory(logger, primitives_list, params);\n'
         '\t\tif(!accelerator) { logger.logError("Accelerator type \'", type, "\' could not be created and has no CPU fallback."); return nullptr; }\n'
         '\t}\n')
    # Scene::updateObjects dereferences the factory's result: make a failed B200 set-up fail the update (and so the render)
    edit(tree, "src/scene/scene.cc",
         '\taccelerator_ = std::unique_ptr<const Accelerator>(Accelerator::factory(logger_, primitives, params));\n',
         '\tif(!accelerator_) { logger_.logError("Scene: no accelerator, cannot render."); return false; }\n')
    edit(tree, "src/scene/scene.cc",
         '\t\tif(creation_state_.changes_ & CreationState::Flags::CGeom) updateObjects();\n',
         '\t\tif(creation_state_.changes_ & CreationState::Flags::CGeom) { if(!updateObjects()) return false; }\n', mode="replace")
    # instances: the plug-in needs the base face behind a PrimitiveInstance (private member, no accessor in the reference)
    edit(tree, "include/geometry/primitive/primitive_instance.h",
         '\t\tconst Object *getObject() const override { return &base_instance_; }\n',
         '\t\tconst Primitive *getBasePrimitive() const { return base_primitive_; }\n')
    # ---- Integrator seam (src/integrator/integrator.cc:37-60) -----------------------------------------------------------
    edit(tree, "src/integrator/integrator.cc", '#include "integrator/surface/integrator_path_tracer.h"\n',
         '#include "integrator/surface/integrator_path_b200.h"\n')
    edit(tree, "src/integrator/integrator.cc",
         '\telse if(type == "directlighting") return DirectLightIntegrator::factory(logger, params, scene, scene.getRenderControl());\n',
         '\tif((type == "pathtracing" || type == "directlighting") && PathIntegratorB200::requested(params)) return PathIntegratorB200::factory(logger, params, scene, scene.getRenderControl());\n',
         mode="before")
    # ---- Scene: keep the ParamMaps the factories were called with (the plug-in exports PARAMETERS, y4gpu.h) -----------------
    edit(tree, "include/scene/scene.h", '#include "render/render_control.h"\n', '#include "scene/scene_b200_params.h"\n')
    edit(tree, "include/scene/scene.h",
         '\t\tconst RenderCallbacks &getRenderCallbacks() const { return render_callbacks_; }\n',
         '\t\tconst B200SceneParams &getB200Params() const { return b200_params_; }\n')
    edit(tree, "include/scene/scene.h", '\t\tRenderCallbacks render_callbacks_;\n', '\t\tB200SceneParams b200_params_;\n')
    edit(tree, "src/scene/scene.cc", '\t\tlights_[name] = std::move(light);\n', '\t\tb200_params_.lights[name] = params;\n')
    edit(tree, "src/scene/scene.cc", '\t\t*(materials_[name]) = std::move(material);\n',
         '\t\tb200_params_.materials[name] = params; b200_params_.material_nodes[name] = nodes_params;\n')
    edit(tree, "src/scene/scene.cc",
         '\treturn createMapItem<Texture>(logger_, name, "Texture", params, textures_, this);\n',
         '\tTexture *b200_item = createMapItem<Texture>(logger_, name, "Texture", params, textures_, this);\n\tif(b200_item) b200_params_.textures[name] = params;\n\treturn b200_item;\n', mode="replace")
    edit(tree, "src/scene/scene.cc",
         '\treturn createMapItem<const Background>(logger_, name, "Background", params, backgrounds_, this);\n',
         '\tconst Background *b200_item = createMapItem<const Background>(logger_, name, "Background", params, backgrounds_, this);\n\tif(b200_item) b200_params_.backgrounds[name] = params;\n\treturn b200_item;\n', mode="replace")
    edit(tree, "src/scene/scene.cc",
         '\treturn createMapItem<const Camera>(logger_, name, "Camera", params, cameras_, this);\n',
         '\tconst Camera *b200_item = createMapItem<const Camera>(logger_, name, "Camera", params, cameras_, this);\n\tif(b200_item) b200_params_.cameras[name] = params;\n\treturn b200_item;\n', mode="replace")
    edit(tree, "src/scene/scene.cc",
         '\treturn createMapItem<Integrator>(logger_, name, "Integrator", params, integrators_, this);\n',
         '\tIntegrator *b200_item = createMapItem<Integrator>(logger_, name, "Integrator", params, integrators_, this);\n\tif(b200_item) b200_params_.integrators[name] = params;\n\treturn b200_item;\n', mode="replace")
    edit(tree, "src/scene/scene.cc",
         '\treturn createMapItem<RenderView>(logger_, name, "RenderView", params, render_views_, this, false);\n',
         '\tRenderView *b200_item = createMapItem<RenderView>(logger_, name, "RenderView", params, render_views_, this, false);\n\tif(b200_item) b200_params_.render_views[name] = params;\n\treturn b200_item;\n', mode="replace")
    edit(tree, "src/scene/scene.cc",
         '\timage_film_ = std::unique_ptr<ImageFilm>(ImageFilm::factory(logger_, params, this));\n',
         '\tb200_params_.render = params;\n')
    edit(tree, "src/scene/scene.cc", '\trender_views_.clear();\n\tclearLayers();\n', '\tb200_params_ = B200SceneParams();\n')
    # ---- ImageFilm: one bulk write-back of the device film instead of 2 G addSample calls under a mutex ------------------------
    edit(tree, "include/render/imagefilm.h",
         '\t\tstatic float darkThresholdCurveInterpolate(float pixel_brightness);\n',
         '\t\t/*! B200 back end: replaces weights_ and the Combined layer by the film accumulated on the device (defined in src/render/imagefilm_b200.cc) */\n'
         '\t\tvoid b200SetAccumulatedFilm(const float *weights, const float *rgba);\n')


if __name__ == "__main__":
    if len(sys.argv) != 2:
        raise SystemExit(__doc__)
    main(Path(sys.argv[1]))
