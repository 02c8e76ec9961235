// rt_kernels.cuh - the sm_100a kernels of the trace loop.
//
// k_trace_pool is the hot kernel: RayTracer::traceImage's pixel loop and traceRay's recursion
// (reference ray/src/RayTracer.cpp:279-314, 108-174) as a persistent-warp wavefront.  Every warp
//   * pulls chunks of primary-ray slots (camera samples x lens positions, or caller-supplied
//     rays) from one global atomic cursor - 32 consecutive slots are an 8x4 pixel block of a
//     32x32 tile, so a warp's primaries are coherent;
//   * owns a LIFO pool of pending secondary rays in HBM/L2 (SoA, one slab per warp);
//   * each iteration refills idle lanes from the pool first (ballot + popc ranks), then from the
//     cursor, traces + shades one ray per lane, and pushes the reflection/refraction children of
//     all lanes back with a warp-scan compaction.
// The recursion's return value becomes a forward weight (see RayTask), so a ray's contribution is
// added to its camera sample's accumulator (integer limbs, rt_accum.cuh) the moment it is shaded and no parent
// ever waits for a child.  Shadow rays are traced inline by the shading lane.
#pragma once
#include <cuda_runtime.h>
#include "rt_shade.cuh"
#include "rt_accum.cuh"

namespace rtd {

enum { SRC_GRID = 0, SRC_XY = 1, SRC_RAYS = 2, SRC_QUEUE = 3 };
#define RT_TILE 32
#define RT_TILE_PIX (RT_TILE * RT_TILE)
#define RT_CHUNK 32
#define RT_POOL_DOUBLES 12

struct PoolArgs {
	rt_params prm;
	int src;
	int w, h;           // image size (SRC_GRID)
	int spp_side;       // supersample grid side s (1 = none)
	int lenses;         // primary rays per sample (1, or divs+1 with depth of field)
	int part, n_parts;  // this context renders tiles t with t % n_parts == part
	long long tile0;    // first local tile of this launch (a frame with more samples than one launch takes is rendered in chunks of whole tiles)
	int tiles_x, tiles_y;
	int depth;          // traceRay depth of primaries
	double eye_shift;   // anaglyph second eye
	long long n_slots;  // primary-ray slots in this launch
	const double* xs;   // SRC_XY: sample window coordinates
	const double* ys;
	const rt_ray* rays; // SRC_RAYS
	const double* q_d;  // SRC_QUEUE (k_trace_pool finishing a frame from the wavefront's queue): the RayQueue's arrays and capacity
	const int* q_i;
	long long q_cap;
	long long* accum;   // RT_ACC_STRIDE limbs per sample (rt_accum.cuh)
	double* base;       // 3 doubles per camera-ray slot (`lenses` per sample): that ray's own term, stored instead of added
	unsigned long long* cursor;
	unsigned long long* stats; // camera, reflection, refraction, shadow, overflow flag
	double* pool_d;     // per warp: RT_POOL_DOUBLES x cap doubles
	int* pool_i;        // per warp: 2 x cap ints
	int cap;            // pool capacity per warp (entries)
};

// local tile k of this partition -> global tile id
__device__ __forceinline__ long long grid_samples_per_tile(const PoolArgs& A) { return (long long)RT_TILE_PIX * A.spp_side * A.spp_side; }

// sample index (local to this partition) -> pixel and window coordinates; false if outside the image
__device__ __forceinline__ bool grid_sample(const PoolArgs& A, long long sample, int& pi, int& pj, double& x, double& y)
{
	const int spp = A.spp_side * A.spp_side;
	const long long pix = sample / spp;
	const int sub = (int)(sample % spp);
	const long long ltile = pix / RT_TILE_PIX + A.tile0;
	const int p = (int)(pix % RT_TILE_PIX);
	const long long tile = ltile * A.n_parts + A.part;
	const int tx = (int)(tile % A.tiles_x), ty = (int)(tile / A.tiles_x);
	const int b = p >> 5, wi = p & 31;
	pi = tx * RT_TILE + (b & 3) * 8 + (wi & 7);
	pj = ty * RT_TILE + (b >> 2) * 4 + (wi >> 3);
	if (pi >= A.w || pj >= A.h) return false;
	// RayTracer::tracePixel (RayTracer.cpp:87-88) on sub-pixel (i*s + a, j*s + b), a outer, b inner (:292-296)
	const int s = A.spp_side;
	const int a = sub / s, bb = sub % s;
	x = double(pi * s + a) / (double(A.w) * s);
	y = double(pj * s + bb) / (double(A.h) * s);
	return true;
}

} // namespace rtd
#include "rt_pool.cuh"
#include "rt_wave.cuh"
namespace rtd {

// ---- Scene::intersect for caller-supplied rays (rt_trace_rays hits) ---------------------------
__global__ void __launch_bounds__(128) k_hits(const DScene S, const rt_ray* rays, int n, rt_hit* out)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	const d3 o = ld3(rays[k].p), d = ld3(rays[k].d);
	Hit h;
	rt_hit r;
	r.obj = r.face = -1;
	r.N[0] = r.N[1] = r.N[2] = 0.0;
	if (closest_hit<true>(S, o, d, h)) {
		finalize_hit(S, o, d, h);
		r.obj = h.obj;
		r.face = h.face;
		r.N[0] = h.N.x; r.N[1] = h.N.y; r.N[2] = h.N.z;
	}
	r.t = h.t;
	out[k] = r;
}

// RayTracer::trace's tail for one sample (RayTracer.cpp:74-78): mean over the lens rays, clamp
__device__ __forceinline__ d3 sample_value(const long long* accum, const double* base, long long sample, const rt_params& prm)
{
	d3 c = acc_read3(accum, base, prm.dof_on ? prm.dof_divs + 1 : 1, sample);
	if (prm.dof_on) c = c * (1.0 / (prm.dof_divs + 1.0));
	return gclamp(c, 0.0, 1.0);
}

// RayTracer::setPixel (RayTracer.cpp:388-394)
__device__ __forceinline__ void store_pixel(uint8_t* fb, int w, int i, int j, d3 c)
{
	uint8_t* px = fb + ((size_t)i + (size_t)j * w) * 3;
	px[0] = (uint8_t)(int)(255.0 * c.x);
	px[1] = (uint8_t)(int)(255.0 * c.y);
	px[2] = (uint8_t)(int)(255.0 * c.z);
}

// No-AA and supersampled resolve (RayTracer.cpp:288-298, 310): one thread per local pixel.
// accum2 (optional) holds the anaglyph second-eye samples whose red channel replaces the first's.
__global__ void __launch_bounds__(256) k_resolve_grid(const PoolArgs A, const long long* accum2, const double* base2, long long n_pix, uint8_t* fb)
{
	const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (pix >= n_pix) return;
	const int spp = A.spp_side * A.spp_side;
	int pi, pj;
	double x, y;
	if (!grid_sample(A, pix * spp, pi, pj, x, y)) return;
	d3 col = mk3(0.0, 0.0, 0.0);
	for (int s = 0; s < spp; s++) {
		d3 c = sample_value(A.accum, A.base, pix * spp + s, A.prm);
		if (accum2) c.x = sample_value(accum2, base2, pix * spp + s, A.prm).x;
		col = col + c;
	}
	if (spp > 1) {
		const double dv = double(spp);
		col = mk3(col.x / dv, col.y / dv, col.z / dv);
	}
	store_pixel(fb, A.w, pi, pj, col);
}

// ---- adaptive anti-aliasing (RayTracer::adaptaa, RayTracer.cpp:316-365) -------------------------
// A region is one adaptaa call: window [x1,x2]x[y1,y2] of pixel `pix` at recursion level `level`.
struct Region {
	double x1, x2, y1, y2;
	int pix;   // local pixel index
	int level;
};


__global__ void __launch_bounds__(256) k_adapt_init(const PoolArgs A, long long n_pix, Region* regions)
{
	const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (pix >= n_pix) return;
	int pi, pj;
	double x, y;
	Region r;
	r.pix = (int)pix;
	r.level = 0;
	if (grid_sample(A, pix, pi, pj, x, y)) {
		r.x1 = double(pi) / double(A.w);
		r.x2 = double(pi + 1) / double(A.w);
		r.y1 = double(pj) / double(A.h);
		r.y2 = double(pj + 1) / double(A.h);
	} else {
		r.x1 = r.x2 = r.y1 = r.y2 = 0.0; // outside the image: fails the size test, contributes nothing
	}
	regions[pix] = r;
}

// valid(region) is adaptaa's first line (RayTracer.cpp:319)
__device__ __forceinline__ bool region_valid(const Region& r) { return !(r.x1 + 0.0001 >= r.x2 || r.y1 + 0.0001 >= r.y2); }

// sample coordinates of every region of one level
__global__ void __launch_bounds__(256) k_adapt_samples(const Region* regions, long long n_regions, int N, double* xs, double* ys)
{
	const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n_regions * N) return;
	const Region r = regions[t / N];
	const int k = (int)(t % N);
	const double w = r.x2 - r.x1, h = r.y2 - r.y1;
	// adaptaa returns before it samples a region that fails its size test (RayTracer.cpp:319) - the padding pixels of
	// edge tiles, or pixels narrower than 1e-4 of the window: no rays, and none counted (NaN = k_wave_gen skips the slot)
	xs[t] = region_valid(r) ? radical_inverse2(k) * w + r.x1 : NAN;
	ys[t] = 0.0 * h + r.y1;
}

// mean / deviation test of each region; either adds its mean (scaled by 4^-level) to the pixel or
// emits its four quadrants for the next level.

__global__ void __launch_bounds__(256) k_adapt_resolve(const Region* regions, long long n_regions, int N, const long long* accum, const double* base, rt_params prm,
                                                      long long* pix_accum, Region* next, unsigned long long* next_count, unsigned long long next_cap)
{
	const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n_regions) return;
	const Region r = regions[t];
	if (!region_valid(r)) return; // val left at its zero initialisation
	d3 mu = mk3(0.0, 0.0, 0.0);
	for (int k = 0; k < N; k++) mu = mu + sample_value(accum, base, t * N + k, prm);
	mu = mu * (1.0 / N);
	d3 sd = mk3(0.0, 0.0, 0.0);
	for (int k = 0; k < N; k++) {
		const d3 c = sample_value(accum, base, t * N + k, prm);
		const d3 a = mk3(fabs(c.x - mu.x), fabs(c.y - mu.y), fabs(c.z - mu.z));
		sd = sd + a * a; // pow(|s - mu|, 2.0) is the correctly rounded square
	}
	sd = sd * (1.0 / (N - 1));
	double scale = 1.0;
	for (int l = 0; l < r.level; l++) scale *= 0.25;
	if (length(sd) > prm.aa_thresh) {
		const double xs[3] = {r.x1, (r.x1 + r.x2) / 2.0, r.x2};
		const double ys[3] = {r.y1, (r.y1 + r.y2) / 2.0, r.y2};
		for (int i = 0; i < 2; i++)
			for (int j = 0; j < 2; j++) {
				Region c;
				c.x1 = xs[i]; c.x2 = xs[i + 1]; c.y1 = ys[j]; c.y2 = ys[j + 1];
				c.pix = r.pix;
				c.level = r.level + 1;
				if (!region_valid(c)) continue; // returns with val = 0
				const unsigned long long at = atomicAdd(next_count, 1ull);
				if (at < next_cap) next[at] = c;
			}
		return;
	}
	mu = mu * scale;
	// the regions of a pixel finish in any order: integer limbs again
	long long* pa = pix_accum + (long long)RT_ACC_STRIDE * r.pix;
	acc_add(pa + 0, mu.x);
	acc_add(pa + 2, mu.y);
	acc_add(pa + 4, mu.z);
}

__global__ void __launch_bounds__(256) k_adapt_store(const PoolArgs A, long long n_pix, const long long* pix_accum, uint8_t* fb)
{
	const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (pix >= n_pix) return;
	int pi, pj;
	double x, y;
	if (!grid_sample(A, pix, pi, pj, x, y)) return;
	store_pixel(fb, A.w, pi, pj, acc_read3(pix_accum, nullptr, 1, pix));
}

} // namespace rtd

namespace rtd {
// ---- framebuffer gather of the tile partition ----------------------------------------------------
// A rank's tiles (t = part, part + n_parts, ...) as one contiguous block, 32 x 32 x 3 bytes per tile in tile order: what is
// sent to the gathering rank instead of the whole frame.  Pixels of an edge tile outside the image are 0.
__global__ void __launch_bounds__(256) k_pack_tiles(const uint8_t* __restrict__ fb, int w, int h, int tiles_x, int part, int n_parts, long long n_local, uint8_t* __restrict__ packed)
{
	const long long p = (long long)blockIdx.x * 256 + threadIdx.x; // local pixel
	if (p >= n_local * RT_TILE_PIX) return;
	const long long tile = (p / RT_TILE_PIX) * n_parts + part;
	const int q = (int)(p % RT_TILE_PIX);
	const int x = (int)(tile % tiles_x) * RT_TILE + (q & 31), y = (int)(tile / tiles_x) * RT_TILE + (q >> 5);
	uint8_t r = 0, g = 0, b = 0;
	if (x < w && y < h) {
		const uint8_t* px = fb + ((size_t)y * w + x) * 3;
		r = px[0]; g = px[1]; b = px[2];
	}
	packed[3 * p] = r; packed[3 * p + 1] = g; packed[3 * p + 2] = b;
}
// The gathering rank's inverse: gathered = n_parts blocks of tiles_cap packed tiles (rank r's block holds its tiles in order).
__global__ void __launch_bounds__(256) k_unpack_tiles(const uint8_t* __restrict__ gathered, int w, int h, int tiles_x, int n_parts, long long tiles_cap, uint8_t* __restrict__ fb)
{
	const long long p = (long long)blockIdx.x * 256 + threadIdx.x; // pixel of the frame
	if (p >= (long long)w * h) return;
	const int x = (int)(p % w), y = (int)(p / w);
	const long long tile = (long long)(y / RT_TILE) * tiles_x + x / RT_TILE;
	const long long rank = tile % n_parts, lt = tile / n_parts;
	const uint8_t* src = gathered + ((rank * tiles_cap + lt) * RT_TILE_PIX + (y % RT_TILE) * RT_TILE + x % RT_TILE) * 3;
	uint8_t* px = fb + (size_t)p * 3;
	px[0] = src[0]; px[1] = src[1]; px[2] = src[2];
}

// RayTracer::tracePixel's return value for explicit samples (clamped mean over lens rays,
// anaglyph red from the second eye)
__global__ void __launch_bounds__(256) k_sample_values(const long long* accum, const double* base, const long long* accum2, const double* base2, long long n, rt_params prm, double* out)
{
	const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	d3 c = sample_value(accum, base, k, prm);
	if (accum2) c.x = sample_value(accum2, base2, k, prm).x;
	out[3 * k] = c.x;
	out[3 * k + 1] = c.y;
	out[3 * k + 2] = c.z;
}

// traceRay's own return value for explicit rays (rt_trace_rays): the sums as they are, not clamped
__global__ void __launch_bounds__(256) k_accum_values(const long long* accum, const double* base, long long n, double* out)
{
	const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	const d3 c = acc_read3(accum, base, 1, k);
	out[3 * k] = c.x;
	out[3 * k + 1] = c.y;
	out[3 * k + 2] = c.z;
}
} // namespace rtd

namespace rtd {
// Self-test of div_by (rt_math.cuh) against the hardware IEEE division sequence: pseudo-random
// operands (full random mantissas; every 4th divisor gets a mantissa of all ones or a single bit,
// the hard cases of reciprocal-based division).  Counts quotients that differ in any bit.
__device__ __forceinline__ unsigned long long splitmix(unsigned long long& s)
{
	unsigned long long z = (s += 0x9e3779b97f4a7c15ull);
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
	return z ^ (z >> 31);
}
// FMA throughput microbenchmark (MEASURED_PEAKS.json has no FP64 / FP32 SIMT peak):
// 8 independent chains per thread, `iters` x 8 FMAs; the sum is written so nothing is optimised away.
// Compiled in a TU built with -fmad=false, so the FMAs are spelled with the intrinsics.
template <typename T> __device__ __forceinline__ T fma_t(T a, T b, T c);
template <> __device__ __forceinline__ double fma_t<double>(double a, double b, double c) { return fma(a, b, c); }
template <> __device__ __forceinline__ float fma_t<float>(float a, float b, float c) { return fmaf(a, b, c); }
template <typename T> __global__ void __launch_bounds__(256) k_fma_peak(int iters, T seed, T* out)
{
	T a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
	const T m = (T)0.999, c = (T)0.001;
#pragma unroll 4
	for (int i = 0; i < iters; i++) {
		a0 = fma_t<T>(a0, m, c); a1 = fma_t<T>(a1, m, c); a2 = fma_t<T>(a2, m, c); a3 = fma_t<T>(a3, m, c);
		a4 = fma_t<T>(a4, m, c); a5 = fma_t<T>(a5, m, c); a6 = fma_t<T>(a6, m, c); a7 = fma_t<T>(a7, m, c);
	}
	out[(size_t)blockIdx.x * 256 + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

__global__ void __launch_bounds__(256) k_div_selftest(unsigned long long seed, long long n, int exp_range, unsigned long long* mismatches)
{
	const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	unsigned long long s = seed + 0x632be59bd9b4e019ull * (unsigned long long)k;
	unsigned long long ma = splitmix(s), md = splitmix(s), e = splitmix(s);
	const int ea = (int)(e % (2 * exp_range + 1)) - exp_range, ed = (int)((e >> 20) % (2 * exp_range + 1)) - exp_range;
	if ((k & 3) == 1) md = ~0ull;
	if ((k & 3) == 2) md = 1ull << (e >> 58);
	if ((k & 15) == 7) ma = 0;
	const unsigned long long sa = (ma >> 63) << 63, sd = (md >> 62) << 63;
	const double a = __longlong_as_double(sa | ((unsigned long long)(1023 + ea) << 52) | (ma & 0xfffffffffffffull));
	const double d = __longlong_as_double(sd | ((unsigned long long)(1023 + ed) << 52) | (md & 0xfffffffffffffull));
	const double y = 1.0 / d;
	const double q1 = div_by(a, d, y, div_range_ok(d) ? 1201u : 0u), q2 = a / d;
	if (__double_as_longlong(q1) != __double_as_longlong(q2)) atomicAdd(mismatches, 1ull);
}
} // namespace rtd
