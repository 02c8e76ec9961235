// main_cli.cpp - `ray_b200`: the reference's command line (ui/CommandLineUI.cpp:25-147) in front of
// the device path.  Same options and option grammar (-O <mode> followed by -A/-B/-C values):
//   ray_b200 [-r depth] [-w width] [-c cubemap_file] [-O a|j|r|o|d|g|c|s] [-A v] [-B v] [-C v] [-g gpus] in.ray out.png|out.bmp
// The output format follows the extension as the reference's writeImage does (PNG through zlib, else BMP); -j (json settings) and -t are not mirrored.
#include <dirent.h>
#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include "RayTracer.h"


static void usage(const char* prog)
{
	std::fprintf(stderr, "usage: %s [options] [input.ray output.png|output.bmp]\n"
	                     "  -r <#>      set recursion level (default 0)\n"
	                     "  -w <#>      set output image width (default 150)\n"
	                     "  -c <file>   one of the six cubemap images (the others are found next to it)\n"
	                     "  -O a|j|r    adaptive / jittered / supersampled AA; -A samples, -B threshold (adaptive)\n"
	                     "  -O d        depth of field; -A focal distance, -B subdivisions, -C aperture\n"
	                     "  -g <#>      render on that many GPUs of this box (interleaved tiles, scene replicated)\n"
	                     "  -O g        anaglyph   -O o  overlapping objects   -O c -A thr  adaptive termination\n", prog);
}

// the six faces next to `file`: names containing posx/negx/posy/negy/posz/negz (TraceUI.cc:87-146)
static bool find_cubemap(const std::string& file, std::vector<std::string>& out)
{
	const std::string dir = file.find('/') == std::string::npos ? "." : file.substr(0, file.find_last_of('/'));
	static const char* tag[6] = {"posx", "negx", "posy", "negy", "posz", "negz"};
	out.assign(6, "");
	DIR* dp = opendir(dir.c_str());
	if (!dp) { std::fprintf(stderr, "Couldn't open the directory %s\n", dir.c_str()); return false; }
	int matched = 0;
	while (dirent* ep = readdir(dp)) {
		const std::string fn(ep->d_name);
		for (int i = 0; i < 6; i++)
			if (fn.find(tag[i]) != std::string::npos && out[i].empty()) { out[i] = dir + "/" + fn; matched++; break; }
	}
	closedir(dp);
	if (matched != 6) { std::fprintf(stderr, "Cannot locate all six cubemap files\n"); return false; }
	return true;
}

int main(int argc, char** argv)
{
	rtb200::RayTracer tracer(0);
	int size = 150; // TraceUI::m_nSize default
	std::string cubemap_file;
	char prev = 0;
	int c;
	while ((c = getopt(argc, argv, "r:w:hc:O:A:B:C:g:")) != EOF) {
		switch (c) {
		case 'r': tracer.ui.depth = std::atoi(optarg); break;
		case 'w': size = std::atoi(optarg); break;
		case 'c': cubemap_file = optarg; break;
		case 'g': if (!tracer.useDevices(std::atoi(optarg))) return 1; break;
		case 'O':
			prev = *optarg;
			switch (prev) {
			case 'a': tracer.ui.aa_mode = RT_AA_ADAPTIVE; break;
			case 'j': tracer.ui.aa_mode = RT_AA_JITTERED; break;
			case 'r': tracer.ui.aa_mode = RT_AA_SUPERSAMPLE; break;
			case 'o': tracer.ui.overlapping_objects = 1; break;
			case 'd': tracer.ui.dof_on = 1; break;
			case 'g': tracer.ui.anaglyph = 1; break;
			case 'c': case 's': break;
			default: usage(argv[0]); return 1;
			}
			break;
		case 'A':
			switch (prev) {
			case 'a': case 'j': case 'r': tracer.ui.aa_samples = std::atoi(optarg); break;
			case 'c': tracer.ui.aterm_thresh = std::atof(optarg); break;
			case 'd': tracer.ui.dof_fd = std::atof(optarg); break;
			case 's': tracer.ui.soft_shadow_res = std::atoi(optarg); break;
			default: usage(argv[0]); return 1;
			}
			break;
		case 'B':
			switch (prev) {
			case 'a': tracer.ui.aa_thresh = std::atof(optarg); break;
			case 'd': tracer.ui.dof_divs = std::atoi(optarg); break;
			default: usage(argv[0]); return 1;
			}
			break;
		case 'C':
			if (prev == 'd') tracer.ui.dof_apsz = std::atof(optarg);
			else { usage(argv[0]); return 1; }
			break;
		default: usage(argv[0]); return 1;
		}
	}
	if (optind >= argc - 1) { std::fprintf(stderr, "no input and/or output name.\n"); return 1; }
	if (!cubemap_file.empty()) {
		std::vector<std::string> faces;
		if (find_cubemap(cubemap_file, faces)) { tracer.setCubemap(faces); tracer.ui.use_cubemap = 1; }
	}
	const char* ray_name = argv[optind];
	const char* img_name = argv[optind + 1];
	if (!tracer.loadScene(ray_name)) { std::fprintf(stderr, "Unable to load ray file '%s'\n", ray_name); return 1; }
	int width = size;
	int height = (int)(width / tracer.aspectRatio() + 0.5); // CommandLineUI.cpp:155-156
	const auto t0 = std::chrono::steady_clock::now();
	tracer.traceImage(width, height);
	tracer.waitRender();
	const auto t1 = std::chrono::steady_clock::now();
	unsigned char* buf;
	tracer.getBuffer(buf, width, height);
	if (buf && rt_write_image(img_name, width, height, buf) != RT_OK) { std::fprintf(stderr, "could not write %s\n", img_name); return 1; }
	const rt_stats& s = tracer.lastStats();
	const double rays = (double)(s.camera_rays + s.reflection_rays + s.refraction_rays + s.shadow_rays);
	std::printf("total time = %.6f seconds, rays traced = %.0f, device time = %.3f ms (%.1f Mrays/s)\n",
	            std::chrono::duration<double>(t1 - t0).count(), rays, s.device_ms, s.device_ms > 0 ? rays / s.device_ms / 1e3 : 0.0);
	return 0;
}
