// ImageFilm::b200SetAccumulatedFilm -- the one method the B200 integration adds to ImageFilm (declared by
// integration/apply_patch.py in include/render/imagefilm.h).  ImageFilm::addSample takes a global mutex per sample
// (src/render/imagefilm.cc:714); the device accumulates the whole film instead and hands it over once per pass, the same way
// imageFilmLoad fills weights_ and the layers from a film file (imagefilm.cc:911-934).
#include "render/imagefilm.h"
#include "common/layers.h"

BEGIN_YAFARAY

void ImageFilm::b200SetAccumulatedFilm(const float *weights, const float *rgba)
{
	std::lock_guard<std::mutex> lock_guard(image_mutex_);
	for(int y = 0; y < height_; ++y)
		for(int x = 0; x < width_; ++x)
			weights_(x, y).setFloat(weights[(size_t) y * width_ + x]);
	for(auto &film_image_layer : film_image_layers_)
	{
		if(film_image_layer.first != LayerDef::Combined) continue;      // the device produces the Combined layer only
		for(int y = 0; y < height_; ++y)
			for(int x = 0; x < width_; ++x)
			{
				const float *c = rgba + 4 * ((size_t) y * width_ + x);
				film_image_layer.second.image_->setColor(x, y, Rgba{c[0], c[1], c[2], c[3]});
			}
	}
}

END_YAFARAY
