// TEST INFRASTRUCTURE ONLY (oracle): drives the UNMODIFIED reference renderer through its own
// C public API (/root/reference/include/public_api/yafaray_c_api.h:53-130), linked from
// oracle/_ref/libyafaray4_ref.a, from a line-oriented scene script.  Each script line is one C-API
// call with the same name and argument order (minus the interface pointer), so the very same
// call sequence can be replayed against libyafaray_b200.interface.Interface (our host-side mirror).
//
// usage: ref_render <scene.script>
//   script commands (whitespace separated; names contain no spaces, a string value runs to the end of its line):
//     paramsClearAll | paramsSetString n s | paramsSetBool n 0|1 | paramsSetInt n i | paramsSetFloat n f
//     paramsSetVector n x y z | paramsSetColor n r g b a | paramsPushList | paramsEndList
//     createMaterial|createLight|createCamera|createBackground|createIntegrator|createRenderView|createObject|createOutput name
//     defineLayer | startGeometry | endGeometry | endObject | setCurrentMaterial name
//     addVertex x y z | addNormal x y z | addTriangle a b c | smoothMesh name angle
//     addVertexWithOrco x y z ox oy oz | addUv u v | addTriangleWithUv a b c ua ub uc | createImage name | createTexture name
//     setImageColor image_name x y r g b a
//     loadMesh file   (binary: int32 n_vert, int32 n_tri, float32 verts[n_vert*3], int32 tris[n_tri*3] -> addVertex/addTriangle)
//     addInstance base 16 floats | setupRender
//     render out.f32  (renders; the Combined layer arrives through the put-pixel callback as float
//                      RGBA and is written raw as height*width*4 float32; prints one JSON line)
#include "public_api/yafaray_c_api.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

struct Capture
{
	int w = 0, h = 0;
	std::vector<float> rgba;
	std::string layer;
};

static void notifyLayer(const char *internal_layer_name, const char *, int width, int height, int, void *data)
{
	Capture *c = static_cast<Capture *>(data);
	if(c->rgba.empty())
	{
		c->w = width; c->h = height; c->layer = internal_layer_name;
		c->rgba.assign((size_t) width * height * 4, 0.f);
	}
}
static void putPixel(const char *, const char *layer_name, int x, int y, float r, float g, float b, float a, void *data)
{
	Capture *c = static_cast<Capture *>(data);
	if(c->layer != layer_name || x < 0 || y < 0 || x >= c->w || y >= c->h) return;
	float *p = &c->rgba[((size_t) y * c->w + x) * 4];
	p[0] = r; p[1] = g; p[2] = b; p[3] = a;
}
static void noProgress(int, int, const char *, void *) {}

int main(int argc, char **argv)
{
	if(argc < 2) { fprintf(stderr, "usage: %s scene.script\n", argv[0]); return 2; }
	std::ifstream in(argv[1]);
	if(!in) { fprintf(stderr, "cannot open %s\n", argv[1]); return 2; }
	yafaray_Interface_t *yi = yafaray_createInterface(YAFARAY_INTERFACE_FOR_RENDERING, nullptr, nullptr, nullptr,
	                                                 getenv("REF_RENDER_VERBOSITY") ? YAFARAY_DISPLAY_CONSOLE_NORMAL : YAFARAY_DISPLAY_CONSOLE_HIDDEN);
	const char *verb = getenv("REF_RENDER_VERBOSITY");
	yafaray_setConsoleVerbosityLevel(yi, yafaray_logLevelFromString(verb ? verb : "mute"));
	yafaray_setLogVerbosityLevel(yi, yafaray_logLevelFromString("mute"));
	yafaray_createScene(yi);
	Capture cap;
	std::map<std::string, yafaray_Image_t *> images;
	std::string line;
	int lineno = 0;
	while(std::getline(in, line))
	{
		++lineno;
		std::istringstream ss(line);
		std::string cmd, a, b;
		if(!(ss >> cmd) || cmd[0] == '#') continue;
		if(cmd == "paramsClearAll") yafaray_paramsClearAll(yi);
		else if(cmd == "paramsPushList") yafaray_paramsPushList(yi);
		else if(cmd == "paramsEndList") yafaray_paramsEndList(yi);
		else if(cmd == "paramsSetString")      // the value is the rest of the line (e.g. color_space "CIE (D50)" has spaces)
		{
			ss >> a;
			std::getline(ss, b);
			const size_t first = b.find_first_not_of(" \t"), last = b.find_last_not_of(" \t\r");
			b = first == std::string::npos ? std::string() : b.substr(first, last - first + 1);
			yafaray_paramsSetString(yi, a.c_str(), b.c_str());
		}
		else if(cmd == "paramsSetBool") { int v; ss >> a >> v; yafaray_paramsSetBool(yi, a.c_str(), v ? YAFARAY_BOOL_TRUE : YAFARAY_BOOL_FALSE); }
		else if(cmd == "paramsSetInt") { int v; ss >> a >> v; yafaray_paramsSetInt(yi, a.c_str(), v); }
		else if(cmd == "paramsSetFloat") { double v; ss >> a >> v; yafaray_paramsSetFloat(yi, a.c_str(), v); }
		else if(cmd == "paramsSetVector") { double x, y, z; ss >> a >> x >> y >> z; yafaray_paramsSetVector(yi, a.c_str(), x, y, z); }
		else if(cmd == "paramsSetColor") { float r, g, bb, al; ss >> a >> r >> g >> bb >> al; yafaray_paramsSetColor(yi, a.c_str(), r, g, bb, al); }
		else if(cmd == "createMaterial") { ss >> a; yafaray_createMaterial(yi, a.c_str()); }
		else if(cmd == "createLight") { ss >> a; yafaray_createLight(yi, a.c_str()); }
		else if(cmd == "createCamera") { ss >> a; yafaray_createCamera(yi, a.c_str()); }
		else if(cmd == "createBackground") { ss >> a; yafaray_createBackground(yi, a.c_str()); }
		else if(cmd == "createIntegrator") { ss >> a; yafaray_createIntegrator(yi, a.c_str()); }
		else if(cmd == "createRenderView") { ss >> a; yafaray_createRenderView(yi, a.c_str()); }
		else if(cmd == "createObject") { ss >> a; yafaray_createObject(yi, a.c_str()); }
		else if(cmd == "createImage") { ss >> a; images[a] = yafaray_createImage(yi, a.c_str()); }
		else if(cmd == "setImageColor")
		{
			int x, y; float r, g, bb, al;
			ss >> a >> x >> y >> r >> g >> bb >> al;
			if(images.count(a) && images[a]) yafaray_setImageColor(images[a], x, y, r, g, bb, al);
		}
		else if(cmd == "createTexture") { ss >> a; yafaray_createTexture(yi, a.c_str()); }
		else if(cmd == "createOutput") { ss >> a; yafaray_createOutput(yi, a.c_str()); }
		else if(cmd == "defineLayer") yafaray_defineLayer(yi);
		else if(cmd == "startGeometry") yafaray_startGeometry(yi);
		else if(cmd == "endGeometry") yafaray_endGeometry(yi);
		else if(cmd == "endObject") yafaray_endObject(yi);
		else if(cmd == "setCurrentMaterial") { ss >> a; yafaray_setCurrentMaterial(yi, a.c_str()); }
		else if(cmd == "addVertex") { double x, y, z; ss >> x >> y >> z; yafaray_addVertex(yi, x, y, z); }
		else if(cmd == "addVertexWithOrco") { double x, y, z, ox, oy, oz; ss >> x >> y >> z >> ox >> oy >> oz; yafaray_addVertexWithOrco(yi, x, y, z, ox, oy, oz); }
		else if(cmd == "addUv") { float u, v; ss >> u >> v; yafaray_addUv(yi, u, v); }
		else if(cmd == "addTriangleWithUv") { int i, j, k, ui, uj, uk; ss >> i >> j >> k >> ui >> uj >> uk; yafaray_addTriangleWithUv(yi, i, j, k, ui, uj, uk); }
		else if(cmd == "addNormal") { double x, y, z; ss >> x >> y >> z; yafaray_addNormal(yi, x, y, z); }
		else if(cmd == "addTriangle") { int i, j, k; ss >> i >> j >> k; yafaray_addTriangle(yi, i, j, k); }
		else if(cmd == "smoothMesh") { double ang; ss >> a >> ang; yafaray_smoothMesh(yi, a.c_str(), ang); }
		else if(cmd == "loadMesh")
		{
			ss >> a;
			FILE *f = fopen(a.c_str(), "rb");
			if(!f) { fprintf(stderr, "line %d: cannot open mesh %s\n", lineno, a.c_str()); return 2; }
			int32_t nv = 0, nt = 0;
			if(fread(&nv, 4, 1, f) != 1 || fread(&nt, 4, 1, f) != 1) return 2;
			std::vector<float> v((size_t) nv * 3);
			std::vector<int32_t> t((size_t) nt * 3);
			if(fread(v.data(), 4, v.size(), f) != v.size() || fread(t.data(), 4, t.size(), f) != t.size()) { fprintf(stderr, "short mesh file\n"); return 2; }
			fclose(f);
			for(int i = 0; i < nv; ++i) yafaray_addVertex(yi, v[3 * i], v[3 * i + 1], v[3 * i + 2]);
			for(int i = 0; i < nt; ++i) yafaray_addTriangle(yi, t[3 * i], t[3 * i + 1], t[3 * i + 2]);
		}
		else if(cmd == "addInstance")
		{
			float m[16];
			ss >> a;
			for(int i = 0; i < 16; ++i) ss >> m[i];
			yafaray_addInstance(yi, a.c_str(), m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], m[9], m[10], m[11], m[12], m[13], m[14], m[15]);
		}
		else if(cmd == "setupRender") yafaray_setupRender(yi);
		else if(cmd == "render")
		{
			ss >> a;
			yafaray_setRenderNotifyLayerCallback(yi, notifyLayer, &cap);
			yafaray_setRenderPutPixelCallback(yi, putPixel, &cap);
			const auto t0 = std::chrono::steady_clock::now();
			yafaray_render(yi, noProgress, nullptr, YAFARAY_DISPLAY_CONSOLE_HIDDEN);
			const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
			FILE *fo = fopen(a.c_str(), "wb");
			if(!fo) { fprintf(stderr, "cannot write %s\n", a.c_str()); return 2; }
			fwrite(cap.rgba.data(), 4, cap.rgba.size(), fo);
			fclose(fo);
			printf("{\"impl\": \"reference\", \"width\": %d, \"height\": %d, \"layer\": \"%s\", \"render_s\": %.6f}\n", cap.w, cap.h, cap.layer.c_str(), s);
			fflush(stdout);
		}
		else { fprintf(stderr, "line %d: unknown command '%s'\n", lineno, cmd.c_str()); return 2; }
	}
	yafaray_destroyInterface(yi);
	return 0;
}
