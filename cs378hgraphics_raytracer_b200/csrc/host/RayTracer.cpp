// RayTracer.cpp - see RayTracer.h.  Every method names the reference member it stands in for.
#include "RayTracer.h"
#include <cstdio>
#include <cstring>
#include "rt_arena_writer.hpp"

namespace rtb200 {

RayTracer::RayTracer(int device) : stopTrace(false), ctx(nullptr), multi(nullptr), arena(nullptr), arena_bytes(0), loaded(false), buffer_width(256), buffer_height(256)
{
	rt_params_default(&ui);
	std::memset(&stats, 0, sizeof(stats));
	// the reference starts with a 256x256 buffer (RayTracer.cpp:176-182)
	buffer.assign((size_t)buffer_width * buffer_height * 3, 0);
	const rt_status st = rt_create(device, &ctx);
	if (st != RT_OK) alert(st == RT_ERR_NO_DEVICE ? "no CUDA device: the device path has no CPU fallback" : "rt_create failed");
}

RayTracer::~RayTracer()
{
	if (arena) rt_arena_free(arena);
	if (multi) rt_multi_destroy(multi); // owns ctx
	else if (ctx) rt_destroy(ctx);
}

bool RayTracer::useDevices(int n)
{
	if (n <= 1) return ctx != nullptr;
	std::vector<int> dev(n);
	for (int k = 0; k < n; k++) dev[k] = k;
	rt_multi* m = nullptr;
	const rt_status st = rt_multi_create(dev.data(), n, &m);
	if (st != RT_OK) { alert(st == RT_ERR_NO_DEVICE ? "not that many CUDA devices" : (st == RT_ERR_UNSUPPORTED ? "NCCL is not available" : "rt_multi_create failed")); return false; }
	if (ctx && !multi) rt_destroy(ctx);
	if (multi) rt_multi_destroy(multi);
	multi = m;
	ctx = rt_multi_context(m, 0);
	loaded = false;
	return true;
}

void RayTracer::alert(const std::string& msg)
{
	alert_msg = msg;
	std::fprintf(stderr, "%s\n", msg.c_str()); // CommandLineUI::alert (ui/CommandLineUI.cpp:192-195)
}

bool RayTracer::loadScene(const char* fn)
{
	loaded = false;
	if (!ctx) { alert("no device context"); return false; }
	if (arena) { rt_arena_free(arena); arena = nullptr; }
	const char* faces[6];
	for (int k = 0; k < 6 && cubemap.size() == 6; k++) faces[k] = cubemap[k].c_str();
	rt_status st = rt_load_ray_file(fn, cubemap.size() == 6 ? faces : nullptr, &arena, &arena_bytes);
	if (st != RT_OK) { alert(rt_host_last_error()); return false; } // the reference alerts the parser's message and returns false
	st = multi ? rt_multi_upload_scene(multi, arena, arena_bytes) : rt_upload_scene(ctx, arena, arena_bytes);
	if (st != RT_OK) { alert(multi ? rt_multi_last_error(multi) : rt_last_error(ctx)); return false; }
	loaded = true;
	return true;
}

double RayTracer::aspectRatio()
{
	if (!loaded) return 1.0; // RayTracer.cpp:193
	rt::ArenaView A(arena);
	return A.ptr<rt_camera>(RT_SEC_CAMERA)->aspect;
}

void RayTracer::traceSetup(int w, int h)
{
	if (buffer_width != w || buffer_height != h) {
		buffer_width = w;
		buffer_height = h;
		buffer.assign((size_t)w * h * 3, 0);
	} else {
		std::fill(buffer.begin(), buffer.end(), (unsigned char)0);
	}
}

void RayTracer::traceImage(int w, int h)
{
	traceSetup(w, h);
	if (!loaded) return;
	const rt_status st = multi ? rt_multi_render(multi, &ui, w, h, buffer.data(), &stats) : rt_render(ctx, &ui, w, h, buffer.data(), &stats);
	if (st != RT_OK) alert(multi ? rt_multi_last_error(multi) : rt_last_error(ctx));
}

dvec3 RayTracer::tracePixel(int i, int j)
{
	dvec3 col = {0, 0, 0};
	if (!loaded) return col;
	const int32_t ij[2] = {i, j};
	double rgb[3] = {0, 0, 0};
	if (rt_trace_pixels(ctx, &ui, buffer_width, buffer_height, ij, 1, rgb) != RT_OK) { alert(rt_last_error(ctx)); return col; }
	// like the reference (RayTracer.cpp:81-102) tracePixel only RETURNS the colour: in supersampling mode (i, j)
	// are sub-pixel coordinates in [0, w*s) x [0, h*s), which do not index the buffer; traceImage stores pixels
	col = {rgb[0], rgb[1], rgb[2]};
	return col;
}

dvec3 RayTracer::traceRay(ray& r, double thresh, int depth, double& length, dvec3*)
{
	dvec3 col = {0, 0, 0};
	if (!loaded) return col;
	rt_params p = ui;
	// `thresh` takes effect only while traceUI->aTermSwitch() is set (RayTracer.cpp:125, light.cpp:48); in rt_params the
	// switch is "aterm_thresh > 0", so a caller passing getATermThresh() with the switch off gets no adaptive termination
	p.aterm_thresh = (ui.aterm_thresh > 0.0) ? thresh : 0.0;
	rt_ray rr = {{r.p.x, r.p.y, r.p.z}, {r.d.x, r.d.y, r.d.z}};
	rt_hit hit;
	double rgb[3];
	if (rt_trace_rays(ctx, &p, &rr, 1, depth, &hit, rgb) != RT_OK) { alert(rt_last_error(ctx)); return col; }
	if (hit.obj >= 0) length = hit.t; // `t = i.getT()` only on a hit (RayTracer.cpp:122)
	col = {rgb[0], rgb[1], rgb[2]};
	return col;
}

dvec3 RayTracer::getPixel(int i, int j)
{
	const unsigned char* px = buffer.data() + ((size_t)i + (size_t)j * buffer_width) * 3;
	return {px[0] / 255.0, px[1] / 255.0, px[2] / 255.0};
}

void RayTracer::setPixel(int i, int j, dvec3 color)
{
	unsigned char* px = buffer.data() + ((size_t)i + (size_t)j * buffer_width) * 3;
	px[0] = (unsigned char)(int)(255.0 * color.x);
	px[1] = (unsigned char)(int)(255.0 * color.y);
	px[2] = (unsigned char)(int)(255.0 * color.z);
}

void RayTracer::getBuffer(unsigned char*& buf, int& w, int& h)
{
	buf = buffer.empty() ? nullptr : buffer.data();
	w = buffer_width;
	h = buffer_height;
}

} // namespace rtb200
