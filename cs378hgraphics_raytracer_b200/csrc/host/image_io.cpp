// image_io.cpp - 24-bit BMP reader/writer for textures, cube-map faces and output frames.
//
// Semantics follow the reference's own BMP path (reference ray/src/fileio/bitmap.cpp:17-93,
// 95-140): rows stay in file order (bottom scan line first), BGR is swapped to RGB, row padding
// is dropped.
//
// PNG (reference ray/src/fileio/pngimage.cpp:195-216 readPNG, 226-285 writePNG) without libpng,
// which is absent here: the container and the five scan-line filters are written out below, zlib
// does the deflate stream.  Writer: 8-bit RGB, non-interlaced, rows flipped so that buffer row 0 (the
// bottom scan line) becomes the last row of the file, as writePNG's row_pointers do.  Reader: 8-bit
// RGB, non-interlaced, no gAMA chunk (the reference applies libpng's gamma transform when the file
// carries one, keeps RGBA as four bytes per texel and expands palettes; those files are rejected with
// an error instead of being read differently), rows flipped as readPNG does.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <zlib.h>
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

// ---- PNG ------------------------------------------------------------------------------------
static void be32(unsigned char* p, uint32_t v) { p[0] = v >> 24; p[1] = (v >> 16) & 255; p[2] = (v >> 8) & 255; p[3] = v & 255; }
static uint32_t rdbe32(const unsigned char* p) { return ((uint32_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }
static const unsigned char PNG_SIG[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};

static bool png_chunk(FILE* f, const char* type, const unsigned char* data, size_t n)
{
	unsigned char hd[8];
	be32(hd, (uint32_t)n);
	std::memcpy(hd + 4, type, 4);
	uLong c = crc32(0L, hd + 4, 4);
	if (n) c = crc32(c, data, (uInt)n);
	unsigned char tail[4];
	be32(tail, (uint32_t)c);
	return std::fwrite(hd, 1, 8, f) == 8 && (n == 0 || std::fwrite(data, 1, n, f) == n) && std::fwrite(tail, 1, 4, f) == 4;
}

// rgb: w*h*3, row 0 = bottom scan line
bool write_png24(const std::string& path, int w, int h, const unsigned char* rgb)
{
	if (w <= 0 || h <= 0) return false;
	const size_t row = (size_t)w * 3;
	std::vector<unsigned char> raw((row + 1) * h);
	for (int j = 0; j < h; j++) { // file row j = buffer row h-1-j, filter type 0
		raw[(row + 1) * j] = 0;
		std::memcpy(raw.data() + (row + 1) * j + 1, rgb + row * (size_t)(h - 1 - j), row);
	}
	uLongf zn = compressBound((uLong)raw.size());
	std::vector<unsigned char> z(zn);
	if (compress2(z.data(), &zn, raw.data(), (uLong)raw.size(), 6) != Z_OK) return false;
	FILE* f = std::fopen(path.c_str(), "wb");
	if (!f) return false;
	unsigned char ihdr[13];
	be32(ihdr, (uint32_t)w);
	be32(ihdr + 4, (uint32_t)h);
	ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0; // 8 bit, RGB, deflate, adaptive filtering, no interlace
	bool ok = std::fwrite(PNG_SIG, 1, 8, f) == 8 && png_chunk(f, "IHDR", ihdr, 13) && png_chunk(f, "IDAT", z.data(), zn) && png_chunk(f, "IEND", nullptr, 0);
	ok = (std::fclose(f) == 0) && ok;
	return ok;
}

static int paeth(int a, int b, int c)
{
	const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
	return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

bool read_png24(const std::string& path, Texture& out)
{
	FILE* f = std::fopen(path.c_str(), "rb");
	if (!f) return false;
	std::vector<unsigned char> file;
	unsigned char tmp[65536];
	size_t n;
	while ((n = std::fread(tmp, 1, sizeof(tmp), f)) > 0) file.insert(file.end(), tmp, tmp + n);
	std::fclose(f);
	if (file.size() < 8 + 25 || std::memcmp(file.data(), PNG_SIG, 8) != 0) return false;
	size_t at = 8;
	int w = 0, h = 0;
	bool have_hdr = false;
	std::vector<unsigned char> z;
	while (at + 12 <= file.size()) {
		const uint32_t len = rdbe32(file.data() + at);
		const unsigned char* type = file.data() + at + 4;
		if (at + 12 + (size_t)len > file.size()) return false;
		const unsigned char* data = file.data() + at + 8;
		if (rdbe32(data + len) != (uint32_t)crc32(crc32(0L, type, 4), data, len)) return false;
		if (!std::memcmp(type, "IHDR", 4)) {
			if (len != 13) return false;
			w = (int)rdbe32(data);
			h = (int)rdbe32(data + 4);
			// 8-bit RGB, non-interlaced only (see the header comment)
			if (data[8] != 8 || data[9] != 2 || data[10] != 0 || data[11] != 0 || data[12] != 0 || w <= 0 || h <= 0) return false;
			have_hdr = true;
		} else if (!std::memcmp(type, "gAMA", 4)) {
			return false;
		} else if (!std::memcmp(type, "IDAT", 4)) {
			z.insert(z.end(), data, data + len);
		} else if (!std::memcmp(type, "IEND", 4)) {
			break;
		}
		at += 12 + (size_t)len;
	}
	if (!have_hdr || z.empty()) return false;
	const size_t row = (size_t)w * 3;
	std::vector<unsigned char> raw((row + 1) * h);
	uLongf rn = (uLongf)raw.size();
	if (uncompress(raw.data(), &rn, z.data(), (uLong)z.size()) != Z_OK || rn != raw.size()) return false;
	out.width = w;
	out.height = h;
	out.rgb.assign(row * h, 0);
	std::vector<unsigned char> prev(row, 0), cur(row);
	for (int j = 0; j < h; j++) {
		const unsigned char* in = raw.data() + (row + 1) * j;
		const int ft = in[0];
		if (ft > 4) return false;
		for (size_t i = 0; i < row; i++) {
			const int a = i >= 3 ? cur[i - 3] : 0, b = prev[i], c = i >= 3 ? prev[i - 3] : 0;
			int v = in[1 + i];
			if (ft == 1) v += a;
			else if (ft == 2) v += b;
			else if (ft == 3) v += (a + b) / 2;
			else if (ft == 4) v += paeth(a, b, c);
			cur[i] = (unsigned char)v;
		}
		std::memcpy(out.rgb.data() + row * (size_t)(h - 1 - j), cur.data(), row); // readPNG flips: row 0 = bottom
		prev.swap(cur);
	}
	return true;
}

// readImage / writeImage (reference ray/src/fileio/images.cc:48-66): by extension, case-insensitive;
// an unknown extension is written as BMP
static std::string lower_ext(const std::string& path)
{
	const size_t dot = path.find_last_of('.');
	std::string ext = dot == std::string::npos ? "" : path.substr(dot);
	for (auto& ch : ext) ch = (char)((ch >= 'A' && ch <= 'Z') ? ch - 'A' + 'a' : ch);
	return ext;
}
bool read_image(const std::string& path, Texture& out)
{
	const std::string ext = lower_ext(path);
	if (ext == ".png") return read_png24(path, out);
	return read_bmp24(path, out);
}
bool write_image(const std::string& path, int w, int h, const unsigned char* rgb)
{
	if (lower_ext(path) == ".png") return write_png24(path, w, h, rgb);
	return write_bmp24(path, w, h, rgb);
}

} // namespace rs
