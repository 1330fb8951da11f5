// Adaptive anti-aliasing on the film: which pixels get more samples in the next pass.
// Restated from ImageFilm::nextPass (src/render/imagefilm.cc:259-445, the `adaptive_aa && threshold > 0` branch) with
// Rgba::normalized (color.h:554-558), Rgba::colorDifference (color.h:450-468) and Rgb::abscol2Bri / col2Bri (color.h:61-62).
// The reference walks the film serially and only ever *sets* flags after an initial clear, so the result does not
// depend on the visiting order: one thread per pixel, plain stores of `1`.
#pragma once
#include "y4_common.cuh"
#include "../../include/y4gpu.h"

struct Y4Rgba { float r, g, b, a; };

// combined.getColor(x, y).normalized(weights(x, y)): colour * (1.f / weight), or 0 when the weight is 0
Y4_HD Y4Rgba y4_film_pixel(const float *weights, const float *rgba, int width, int x, int y)
{
	const size_t p = (size_t) y * (size_t) width + (size_t) x;
	const float w = weights[p];
	Y4Rgba c; c.r = 0.f; c.g = 0.f; c.b = 0.f; c.a = 0.f;
	if(w != 0.f)
	{
		const float inv = xdiv(1.f, w);
		c.r = xmul(rgba[4 * p], inv); c.g = xmul(rgba[4 * p + 1], inv); c.b = xmul(rgba[4 * p + 2], inv); c.a = xmul(rgba[4 * p + 3], inv);
	}
	return c;
}
Y4_HD float y4_col2bri(const Y4Rgba &c) { return xadd(xadd(xmul(0.2126f, c.r), xmul(0.7152f, c.g)), xmul(0.0722f, c.b)); }
Y4_HD float y4_abscol2bri(const Y4Rgba &c) { return xadd(xadd(xmul(0.2126f, fabsf(c.r)), xmul(0.7152f, fabsf(c.g))), xmul(0.0722f, fabsf(c.b))); }
Y4_HD float y4_color_difference(const Y4Rgba &a, const Y4Rgba &b, bool use_rgb)     // a.colorDifference(b)
{
	float d = fabsf(xsub(y4_col2bri(b), y4_col2bri(a)));
	if(use_rgb)
	{
		const float rd = fabsf(xsub(b.r, a.r)), gd = fabsf(xsub(b.g, a.g)), bd = fabsf(xsub(b.b, a.b)), ad = fabsf(xsub(b.a, a.a));
		if(d < rd) d = rd;
		if(d < gd) d = gd;
		if(d < bd) d = bd;
		if(d < ad) d = ad;
	}
	return d;
}

// ImageFilm::darkThresholdCurveInterpolate (imagefilm.cc:799-815): piecewise-linear threshold over the pixel brightness
Y4_HD float y4_dark_threshold_curve(float b)
{
	const float knot[14] = {0.10f, 0.20f, 0.30f, 0.40f, 0.50f, 0.60f, 0.70f, 0.80f, 0.90f, 1.00f, 1.20f, 1.40f, 1.80f, 0.f};
	const float val[14] = {0.0001f, 0.0010f, 0.0020f, 0.0035f, 0.0055f, 0.0075f, 0.0100f, 0.0150f, 0.0250f, 0.0400f, 0.0800f, 0.0950f, 0.1000f, 0.f};
	const float width[13] = {0.f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.20f, 0.20f, 0.40f};
	if(b <= knot[0]) return val[0];
	for(int k = 1; k < 13; ++k)
		if(b <= knot[k]) return xadd(val[k - 1], xdiv(xmul(xsub(b, knot[k - 1]), xsub(val[k], val[k - 1])), width[k]));
	return 0.1000f;
}

// phase 1 (imagefilm.cc:301-309): pixels that have no sample at all are always resampled
Y4_HD void y4_film_flag_init(const float *weights, uint8_t *flags, size_t p) { flags[p] = weights[p] > 0.f ? 0 : 1; }

// phase 2 (imagefilm.cc:311-413) for one pixel (x, y), x < width - 1, y < height - 1
Y4_HD void y4_film_flag_detect(const float *weights, const float *rgba, int width, int height, const y4gpu_pt_aa_noise &aa, uint8_t *flags, int x, int y)
{
	const bool rgbn = aa.detect_color_noise != 0;
	const Y4Rgba pix = y4_film_pixel(weights, rgba, width, x, y);
	float thr = aa.threshold;
	if(aa.dark_detection_type == 1 && aa.dark_threshold_factor > 0.f)
		thr = xmul(aa.threshold, xadd(xsub(1.f, aa.dark_threshold_factor), xmul(y4_abscol2bri(pix), aa.dark_threshold_factor)));
	else if(aa.dark_detection_type == 2) thr = y4_dark_threshold_curve(y4_abscol2bri(pix));
	const size_t p = (size_t) y * (size_t) width + (size_t) x;
	if(y4_color_difference(pix, y4_film_pixel(weights, rgba, width, x + 1, y), rgbn) >= thr) { flags[p] = 1; flags[p + 1] = 1; }
	if(y4_color_difference(pix, y4_film_pixel(weights, rgba, width, x, y + 1), rgbn) >= thr) { flags[p] = 1; flags[p + (size_t) width] = 1; }
	if(y4_color_difference(pix, y4_film_pixel(weights, rgba, width, x + 1, y + 1), rgbn) >= thr) { flags[p] = 1; flags[p + (size_t) width + 1] = 1; }
	if(x > 0 && y4_color_difference(pix, y4_film_pixel(weights, rgba, width, x - 1, y + 1), rgbn) >= thr) { flags[p] = 1; flags[p + (size_t) width - 1] = 1; }
	if(aa.variance_pixels > 0)
	{
		const int half = aa.variance_edge_size / 2;
		int variance_x = 0, variance_y = 0;
		for(int xd = -half; xd < half - 1; ++xd)
		{
			int xi = x + xd;
			if(xi < 0) xi = 0; else if(xi >= width - 1) xi = width - 2;
			if(y4_color_difference(y4_film_pixel(weights, rgba, width, xi, y), y4_film_pixel(weights, rgba, width, xi + 1, y), rgbn) >= thr) ++variance_x;
		}
		for(int yd = -half; yd < half - 1; ++yd)
		{
			int yi = y + yd;
			if(yi < 0) yi = 0; else if(yi >= height - 1) yi = height - 2;
			if(y4_color_difference(y4_film_pixel(weights, rgba, width, x, yi), y4_film_pixel(weights, rgba, width, x, yi + 1), rgbn) >= thr) ++variance_y;
		}
		if(variance_x + variance_y >= aa.variance_pixels)
			for(int xd = -half; xd < half; ++xd)
				for(int yd = -half; yd < half; ++yd)
				{
					int xi = x + xd, yi = y + yd;
					if(xi < 0) xi = 0; else if(xi >= width) xi = width - 1;
					if(yi < 0) yi = 0; else if(yi >= height) yi = height - 1;
					flags[(size_t) yi * (size_t) width + (size_t) xi] = 1;
				}
	}
}
