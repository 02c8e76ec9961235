// image_io.cpp - 24-bit BMP reader/writer for textures, cube-map faces and output frames.
//
// Semantics follow the reference's own BMP path (reference ray/src/fileio/bitmap.cpp:17-93,
// 95-140): rows stay in file order (bottom scan line first), BGR is swapped to RGB, row padding
// is dropped.  PNG needs libpng, which is absent here; the reference's PNG reader is out of scope
//  and .png textures are rejected with an error.
#include <cstdio>
#include <cstring>
#include "ray_scene.hpp"

namespace rs {

static uint32_t rd32(const unsigned char* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
static uint16_t rd16(const unsigned char* p) { return (uint16_t)(p[0] | (p[1] << 8)); }

bool read_bmp24(const std::string& path, Texture& out)
{
	size_t dot = path.find_last_of('.');
	if (dot == std::string::npos) return false;
	std::string ext = path.substr(dot);
	for (auto& ch : ext) ch = (char)((ch >= 'A' && ch <= 'Z') ? ch - 'A' + 'a' : ch);
	if (ext != ".bmp") return false;
	FILE* f = std::fopen(path.c_str(), "rb");
	if (!f) return false;
	unsigned char hdr[54];
	if (std::fread(hdr, 1, 54, f) != 54) { std::fclose(f); return false; }
	if (rd16(hdr) != 0x4d42 || rd16(hdr + 28) != 24) { std::fclose(f); return false; }
	const uint32_t off = rd32(hdr + 10);
	const int w = (int)rd32(hdr + 18), h = (int)rd32(hdr + 22);
	if (w <= 0 || h <= 0) { std::fclose(f); return false; }
	const size_t row = ((size_t)w * 3 + 3) / 4 * 4;
	std::vector<unsigned char> raw(row * h);
	std::fseek(f, (long)off, SEEK_SET);
	const bool ok = std::fread(raw.data(), raw.size(), 1, f) == 1;
	std::fclose(f);
	if (!ok) return false;
	out.width = w;
	out.height = h;
	out.rgb.resize((size_t)w * h * 3);
	for (int j = 0; j < h; j++) {
		const unsigned char* in = raw.data() + row * j;
		unsigned char* o = out.rgb.data() + (size_t)w * 3 * j;
		for (int i = 0; i < w; i++) {
			o[3 * i] = in[3 * i + 2];
			o[3 * i + 1] = in[3 * i + 1];
			o[3 * i + 2] = in[3 * i];
		}
	}
	return true;
}

// rgb: w*h*3, row 0 = bottom scan line (the layout of RayTracer::buffer, RayTracer.cpp:388-393)
bool write_bmp24(const std::string& path, int w, int h, const unsigned char* rgb)
{
	FILE* f = std::fopen(path.c_str(), "wb");
	if (!f) return false;
	const size_t row = ((size_t)w * 3 + 3) / 4 * 4;
	unsigned char hdr[54];
	std::memset(hdr, 0, sizeof(hdr));
	auto wr32 = [&](int at, uint32_t v) { hdr[at] = v & 255; hdr[at + 1] = (v >> 8) & 255; hdr[at + 2] = (v >> 16) & 255; hdr[at + 3] = (v >> 24) & 255; };
	hdr[0] = 'B'; hdr[1] = 'M';
	wr32(2, (uint32_t)(54 + row * h));
	wr32(10, 54);
	wr32(14, 40);
	wr32(18, (uint32_t)w);
	wr32(22, (uint32_t)h);
	hdr[26] = 1; hdr[28] = 24;
	wr32(34, (uint32_t)(row * h));
	std::fwrite(hdr, 1, 54, f);
	std::vector<unsigned char> line(row, 0);
	for (int j = 0; j < h; j++) {
		const unsigned char* in = rgb + (size_t)w * 3 * j;
		for (int i = 0; i < w; i++) {
			line[3 * i] = in[3 * i + 2];
			line[3 * i + 1] = in[3 * i + 1];
			line[3 * i + 2] = in[3 * i];
		}
		std::fwrite(line.data(), 1, row, f);
	}
	std::fclose(f);
	return true;
}

} // namespace rs
