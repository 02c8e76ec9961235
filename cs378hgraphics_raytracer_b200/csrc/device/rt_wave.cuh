// rt_wave.cuh - the trace loop as a streaming wavefront: one small kernel per stage, rays in
// HBM queues between them.
//
// Why: the per-ray code of this tracer is long, branchy FP64 (six primitive kinds, two tree
// levels, textures, refraction, the sorted shadow walk).  Fused into one persistent kernel it is
// instruction-fetch bound (profiles/r1_v2_*, r1_v3_*: 50-80 % of warp stalls are "no instruction",
// the loop body does not fit the 32 KB L1.5 I-cache and every warp of an SM sits in a different
// part of it), while HBM is idle (<1 % of peak).  So the stages are separate kernels - all warps of
// an SM then run the same few KB of code - and the rays stream through HBM, which costs a few
// hundred bytes per ray of a resource that is otherwise unused.
//
// One round (host loop in rt_api.cu: launch_wave):
//   k_wave_gen      primary rays for up to W - |children| new camera samples   -> queue Cur
//   k_wave_closest  Scene::intersect for every ray of the batch                -> hit records
//   k_wave_shade    RayTracer::traceRay's body for every hit / miss: emission + ambient and the
//                   background go to the sample accumulators; reflection / refraction rays are
//                   appended to queue Nxt, one shadow ray per light to the shadow queue
//   k_wave_shadow   Light::shadowAttenuation for every shadow ray; term * attenuation is added
//                   to the sample
// then Cur <-> Nxt.  The recursion's return value is carried forward as a weight (see RayTask in
// rt_shade.cuh), so nothing ever waits for a child.  If a round's children exceed W, the excess is
// spilled to a LIFO stack and taken back when a later round has room (depth-first, so memory stays
// bounded by W * (depth + 2) rays).
#pragma once
#include "rt_coop.cuh"

namespace rtd {

enum { KIND_DEAD = 7 }; // a primary slot outside the image (edge tiles): skipped by every stage

#define RT_WAVE_THREADS 128
#ifndef RT_TRACE_THREADS
#define RT_TRACE_THREADS 64 // block size of the two tree-walk kernels: small blocks retire (and are replaced) sooner
#endif

// closest-hit rays: o[3] d[3] weight[3] att_kt[3] | sample | depth, att_mode, kind (pack_meta)
// shadow rays     : o[3] d[3] term[3]             | sample | light
struct RayQueue {
	double* d;
	int* i;
	long long cap;
	int nd; // doubles per entry: 12 or 9
	__device__ __forceinline__ d3 load_v(int field, long long slot) const
	{
		return mk3(d[(field + 0) * cap + slot], d[(field + 1) * cap + slot], d[(field + 2) * cap + slot]);
	}
	__device__ __forceinline__ void store_v(int field, long long slot, const d3& v) const
	{
		d[(field + 0) * cap + slot] = v.x;
		d[(field + 1) * cap + slot] = v.y;
		d[(field + 2) * cap + slot] = v.z;
	}
	__device__ __forceinline__ void store_meta(long long slot, int a, int b) const
	{
		i[slot] = a;
		i[cap + slot] = b;
	}
};

// what k_wave_closest leaves for k_wave_shade: t < 0 = miss
struct HitQueue {
	double* t;
	int2* of; // (object, global mesh item position or -1)
};

// device-side counters of one trace call
struct WaveCounters {
	unsigned long long stats[8]; // camera, reflection, refraction, shadow rays, overflow flag, heavy rays, culled shadow rays
	unsigned long long n_children;
	unsigned long long n_heavy;      // rays of this round that exhausted the node budget
	unsigned long long heavy_cursor; // next heavy ray to be taken by a warp
	unsigned long long n_retry;      // rays that exhausted the FIRST budget (k_wave_closest); k_wave_closest_retry walks them
	                                 // again with a larger one and only its leftovers (n_heavy) go to the cooperative kernel
	unsigned long long n_complex[2]; // per round parity: shadow rays the any-hit kernel could not decide (k_wave_shadow_list walks them)
	unsigned long long n_shadow[2];  // per round parity: the shadow rays of round r are still being
	                                 // walked (second stream) while round r+1 fills the other queue
};

// One entry per lane with pred set, appended to a queue shared by the whole grid: one atomic per
// warp.  Every lane of the warp must call this.
__device__ __forceinline__ long long wave_reserve(bool pred, unsigned lt_mask, unsigned long long* counter)
{
	const unsigned m = __ballot_sync(0xffffffffu, pred);
	if (m == 0) return -1;
	const int leader = __ffs(m) - 1;
	unsigned long long base = 0;
	if ((int)(threadIdx.x & 31) == leader) base = atomicAdd(counter, (unsigned long long)__popc(m));
	base = __shfl_sync(0xffffffffu, base, leader);
	return pred ? (long long)base + __popc(m & lt_mask) : -1;
}

// ---- stage 0: primary rays ------------------------------------------------------------------
// slots [slot0, slot0 + n) of the launch's primary-ray space -> Q[dst ...]
__global__ void __launch_bounds__(RT_WAVE_THREADS) k_wave_gen(const __grid_constant__ DScene S, const __grid_constant__ PoolArgs A, const RayQueue Q, long long dst, long long slot0, long long n)
{
	const long long k = (long long)blockIdx.x * RT_WAVE_THREADS + threadIdx.x;
	if (k >= n) return;
	const long long s = slot0 + k;
	const long long smp = s / A.lenses;
	const int lens = (int)(s % A.lenses);
	d3 ro = mk3(0, 0, 0), rd = mk3(0, 0, 1);
	bool ok = true;
	if (A.src == SRC_RAYS) {
		ro = ld3(A.rays[s].p);
		rd = ld3(A.rays[s].d);
	} else {
		double x = 0, y = 0;
		if (A.src == SRC_GRID) {
			int pi_, pj_;
			ok = grid_sample(A, smp, pi_, pj_, x, y);
		} else {
			x = A.xs[smp];
			y = A.ys[smp];
			ok = x == x; // NaN: a sample of a region adaptaa does not sample (k_adapt_samples)
		}
		if (ok) primary_ray(S.cam, A.prm, x, y, lens, A.eye_shift, ro, rd);
	}
	const long long q = dst + k;
	Q.store_v(0, q, ro);
	Q.store_v(3, q, rd);
	Q.store_v(6, q, mk3(1.0, 1.0, 1.0));
	Q.store_v(9, q, mk3(1.0, 1.0, 1.0));
	Q.store_meta(q, (int)smp, pack_meta(A.depth, 0, ok ? KIND_CAMERA : KIND_DEAD, lens)); // a camera ray carries its lens index where a shadow ray has its light
}

// Per-block copy of the scene tree's top table into shared memory (rt_scene.cuh: RT_REF_TOP); every thread of the block takes
// part, so this comes before any early return.  With RT_TOP_K == 0 (the default build) it compiles to nothing.
#if RT_TOP_K > 0
#define RT_STAGE_TOP(S, TOPVAR)                                                                          \
	__shared__ DNode2f s_top_[RT_TOP_K];                                                                 \
	{                                                                                                    \
		const float4* src_ = reinterpret_cast<const float4*>((S).stop);                                  \
		float4* dst_ = reinterpret_cast<float4*>(s_top_);                                                \
		for (int i_ = threadIdx.x; i_ < (S).n_top * 4; i_ += blockDim.x) dst_[i_] = __ldg(src_ + i_);    \
		__syncthreads();                                                                                 \
	}                                                                                                    \
	const DNode2f* TOPVAR = (S).n_top > 0 ? s_top_ : nullptr
#else
#define RT_STAGE_TOP(S, TOPVAR) const DNode2f* TOPVAR = nullptr
#endif

// ---- stage 1: Scene::intersect ----------------------------------------------------------------
#ifndef RT_CLOSEST_MINB
#define RT_CLOSEST_MINB 14 // 72 registers: the walk is latency bound, more resident warps beat fewer spills (profiles/r1_variants.txt)
#endif
// perm: NULL, or the order in which the first n_perm rays of the batch are taken (k_closest_keys + radix sort: reflection
// and refraction rays arrive in the order their parents were shaded; sorted by direction octant and origin cell, the rays
// of a warp walk the same subtrees).  Hit records stay indexed by queue slot, so the shade stage does not care.
__global__ void __launch_bounds__(RT_TRACE_THREADS, RT_CLOSEST_MINB) k_wave_closest(const __grid_constant__ DScene S, const RayQueue Q, long long base, long long m, const HitQueue H,
                                                                                    int budget, int* retry, unsigned long long* n_retry, const unsigned* __restrict__ perm, long long n_perm, int sparse)
{
	RT_STAGE_TOP(S, top);
	// sparse > 0 (a round too small to fill the GPU): one ray per 2^sparse lanes.  A warp's time is the SUM of the divergent
	// paths of its rays, a few lanes at a time; with fewer rays per warp and more warps the launch ends sooner, and the lanes
	// left idle would have had nothing to do anyway.
	const long long t0 = (long long)blockIdx.x * RT_TRACE_THREADS + threadIdx.x;
	if (t0 & ((1ll << sparse) - 1)) return;
	const long long k0 = t0 >> sparse;
	if (k0 >= m) return;
	const long long k = k0 < n_perm ? (long long)perm[k0] : k0;
	const long long q = base + k;
	const int meta = Q.i[Q.cap + q];
	if (((meta >> 12) & 15) == KIND_DEAD) return;
	const d3 ro = Q.load_v(0, q), rd = Q.load_v(3, q);
	TravClosest C;
	C.have = false;
	C.best_t = RT_T_INF;
	C.best_rank = 0x7fffffff;
	C.best_obj = -1;
	C.best_fpos = -1;
	SBuf unused;
	unused.n = 0;
	unused.cut = RT_T_INF;
	if (!traverse<true, TRAV_CLOSEST>(S, ro, rd, TRAV_CLOSEST, C, unused, budget, top)) {
		retry[atomicAdd(n_retry, 1ull)] = (int)k; // k_wave_closest_retry (or, with a forced budget, the cooperative kernel) redoes this ray
		return;
	}
	H.t[k] = C.have ? C.best_t : -1.0;
	H.of[k] = make_int2(C.best_obj, C.best_fpos);
}

#define RT_RETRY_MIN 16384 // about 14 rays for each of the cooperative kernel's warps
// Second chance for the rays that ran out of the first budget: the same single-thread walk with a much
// larger budget, over the (short) retry list with a fixed grid.  The cooperative kernel keeps one ray per
// WARP busy and is meant for the few rays that fan out over thousands of nodes; if the first budget was
// simply too small for the scene (it adapts, but a round may be walked before it has) the retry list can
// hold a percent of a round - half a million rays - and those are ordinary rays.
__global__ void __launch_bounds__(RT_TRACE_THREADS, RT_CLOSEST_MINB) k_wave_closest_retry(const __grid_constant__ DScene S, const RayQueue Q, long long base, const HitQueue H,
                                                                                          int budget, const int* retry, int* heavy, WaveCounters* cnt)
{
	const long long n = (long long)cnt->n_retry;
	// a short list is what the cooperative kernel is for (and a single thread would need a millisecond for a
	// ray that really fans out): hand it over unchanged
	const bool forward = n < RT_RETRY_MIN;
	for (long long j = (long long)blockIdx.x * RT_TRACE_THREADS + threadIdx.x; j < n; j += (long long)gridDim.x * RT_TRACE_THREADS) {
		const int k = retry[j];
		if (forward) { heavy[atomicAdd(&cnt->n_heavy, 1ull)] = k; continue; }
		const long long q = base + k;
		const d3 ro = Q.load_v(0, q), rd = Q.load_v(3, q);
		TravClosest C;
		C.have = false;
		C.best_t = RT_T_INF;
		C.best_rank = 0x7fffffff;
		C.best_obj = -1;
		C.best_fpos = -1;
		SBuf unused;
		unused.n = 0;
		unused.cut = RT_T_INF;
		if (!traverse<true, TRAV_CLOSEST>(S, ro, rd, TRAV_CLOSEST, C, unused, budget)) {
			heavy[atomicAdd(&cnt->n_heavy, 1ull)] = k; // k_wave_closest_heavy redoes this ray
			continue;
		}
		H.t[k] = C.have ? C.best_t : -1.0;
		H.of[k] = make_int2(C.best_obj, C.best_fpos);
	}
}

// The rays that ran out of budget, one per warp (rt_coop.cuh); persistent grid, dynamic fetch.
__global__ void __launch_bounds__(32 * RT_COOP_WARPS) k_wave_closest_heavy(const __grid_constant__ DScene S, const RayQueue Q, long long base, const HitQueue H, const int* heavy, WaveCounters* cnt)
{
	__shared__ CoopStack stacks[RT_COOP_WARPS];
	__shared__ CoopScene scenes[RT_COOP_WARPS];
	CoopStack& K = stacks[threadIdx.x >> 5];
	CoopScene& A = scenes[threadIdx.x >> 5];
	const unsigned lane = threadIdx.x & 31;
	const unsigned long long n = cnt->n_heavy;
	for (;;) {
		unsigned long long j = 0;
		if (lane == 0) j = atomicAdd(&cnt->heavy_cursor, 1ull);
		j = __shfl_sync(0xffffffffu, j, 0);
		if (j >= n) break;
		const int k = heavy[j];
		const long long q = base + k;
		const d3 ro = Q.load_v(0, q), rd = Q.load_v(3, q);
		TravClosest C;
		C.have = false;
		C.best_t = RT_T_INF;
		C.best_rank = 0x7fffffff;
		C.best_obj = -1;
		C.best_fpos = -1;
		coop_closest(S, ro, rd, C, K, A);
		if (lane == 0) {
			H.t[k] = C.have ? C.best_t : -1.0;
			H.of[k] = make_int2(C.best_obj, C.best_fpos);
		}
		__syncwarp();
	}
	if (lane == 0 && threadIdx.x == 0 && blockIdx.x == 0 && n) atomicAdd(cnt->stats + 5, n);
}

// ---- stage 2: RayTracer::traceRay's body --------------------------------------------------------
#ifndef RT_SHADE_MINB
#define RT_SHADE_MINB 4
#endif
__global__ void __launch_bounds__(RT_WAVE_THREADS, RT_SHADE_MINB) k_wave_shade(const __grid_constant__ DScene S, const __grid_constant__ PoolArgs A, const RayQueue Q, long long base, long long m, const HitQueue H,
                                                                 const RayQueue Nxt, const RayQueue SQ, int parity, WaveCounters* cnt)
{
	const unsigned lane = threadIdx.x & 31;
	const unsigned lt_mask = (1u << lane) - 1u;
	const long long k = (long long)blockIdx.x * RT_WAVE_THREADS + threadIdx.x;
	const long long q = base + k;
	bool live = k < m;
	int meta = 0, sample = 0;
	if (live) {
		meta = Q.i[Q.cap + q];
		sample = Q.i[q];
		if (((meta >> 12) & 15) == KIND_DEAD) live = false;
	}
	unsigned n_cam = 0, n_refl = 0, n_refr = 0, n_shadow = 0, n_culled = 0, overflow = 0;
	bool lit = false, has_refl = false, has_refr = false;
	RayTask kid[2];
	MatEval me;
	Hit h;
	d3 w = mk3(0, 0, 0), hitP = mk3(0, 0, 0), rd = mk3(0, 0, 1);
	if (live) {
		const int kind = (meta >> 12) & 15;
		const d3 ro = Q.load_v(0, q);
		rd = Q.load_v(3, q);
		const d3 weight = Q.load_v(6, q);
		if (kind == KIND_CAMERA) n_cam = 1;
		else if (kind == KIND_REFL) n_refl = 1;
		else n_refr = 1;
		const double t = H.t[k];
		if (t < 0.0) {
			add_term(A, sample, kind == KIND_CAMERA, meta >> 16, weight * miss_colour(S, A.prm, rd));
		} else {
			const int2 of = H.of[k];
			h.obj = of.x;
			h.fpos = of.y;
			h.t = t;
			finalize_hit(S, ro, rd, h);
			w = apply_att(weight, Q.load_v(9, q), (meta >> 8) & 15, h.t);
			eval_material(S, h, me);
			add_term(A, sample, kind == KIND_CAMERA, meta >> 16, w * (me.ke + me.ka * ld3(S.glob.ambient)));
			hitP = ro + h.t * rd;
			lit = true;
			spawn_children(me, h, ro, rd, w, (meta & 255) - 1, kid, has_refl, has_refr);
		}
	}
	// reflection / refraction rays -> next round's queue
	for (int c = 0; c < 2; c++) {
		const long long s = wave_reserve(c == 0 ? has_refl : has_refr, lt_mask, &cnt->n_children);
		if (s >= 0) {
			if (s >= Nxt.cap) { overflow = 1; continue; }
			Nxt.store_v(0, s, kid[c].o);
			Nxt.store_v(3, s, kid[c].d);
			Nxt.store_v(6, s, kid[c].weight);
			Nxt.store_v(9, s, kid[c].att_kt);
			Nxt.store_meta(s, sample, pack_meta(kid[c].depth, kid[c].att_mode, kid[c].kind, 0));
		}
	}
	// one shadow ray per light (Material::shade's loop, material.cpp:50-66), light by light so
	// that neighbouring entries of the shadow queue aim at the same light
	if (__any_sync(0xffffffffu, lit)) {
		for (int li = 0; li < S.n_lights; li++) {
			const rt_light& Lt = S.lights[li];
			if (!light_is_area(Lt)) {
				// Material::shade adds dattn * sattn * colour * (d_comp + s_comp) (material.cpp:64).  Where
				// the weighted Phong term is exactly (0,0,0) - the light is behind the surface - the sum
				// gains 0 whatever the shadow ray finds (sattn is finite: DScene::kt_unit), so the ray is
				// counted, as the reference counts it, but never walked.
				d3 wterm = mk3(0, 0, 0), pb = mk3(0, 0, 0), sdir = mk3(0, 0, 1);
				bool walk = false;
				if (lit) {
					d3 term;
					light_term(Lt, me, hitP, h.N, rd, term, pb, sdir);
					wterm = w * term;
					walk = !(S.kt_unit && wterm.x == 0.0 && wterm.y == 0.0 && wterm.z == 0.0);
					n_shadow++;
					if (!walk) n_culled++;
				}
				const long long s = wave_reserve(walk, lt_mask, &cnt->n_shadow[parity]);
				if (s >= 0) {
					if (s >= SQ.cap) { overflow = 1; continue; }
					SQ.store_v(0, s, pb);
					SQ.store_v(3, s, sdir);
					SQ.store_v(6, s, wterm);
					SQ.store_meta(s, sample, li);
				}
				continue;
			}
			// area / spot light: sattn = (1 + sum of the valid samples' attenuations) / (res - 1);
			// the constant part goes to the sample now, every valid sample becomes a shadow ray
			const bool lv = lit && light_valid(Lt, hitP);
			d3 wt = mk3(0, 0, 0), pb = mk3(0, 0, 0), sdir = mk3(0, 0, 1);
			if (lv) {
				d3 term;
				light_term(Lt, me, hitP, h.N, rd, term, pb, sdir);
				wt = (w * term) * light_sample_weight(Lt, A.prm);
				accumulate(A.accum, sample, wt);
			}
			for (int i = 0; i < A.prm.soft_shadow_res; i++) {
				const bool ok = lv && light_sample(Lt, A.prm, i, pb, sdir);
				const long long s = wave_reserve(ok, lt_mask, &cnt->n_shadow[parity]);
				if (s >= 0) {
					n_shadow++;
					if (s >= SQ.cap) { overflow = 1; continue; }
					SQ.store_v(0, s, pb);
					SQ.store_v(3, s, sdir);
					SQ.store_v(6, s, wt);
					SQ.store_meta(s, sample, li);
				}
			}
		}
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		n_cam += __shfl_down_sync(0xffffffffu, n_cam, o);
		n_refl += __shfl_down_sync(0xffffffffu, n_refl, o);
		n_refr += __shfl_down_sync(0xffffffffu, n_refr, o);
		n_shadow += __shfl_down_sync(0xffffffffu, n_shadow, o);
		n_culled += __shfl_down_sync(0xffffffffu, n_culled, o);
		overflow |= __shfl_down_sync(0xffffffffu, overflow, o);
	}
	if (lane == 0) {
		if (n_cam) atomicAdd(cnt->stats + 0, (unsigned long long)n_cam);
		if (n_refl) atomicAdd(cnt->stats + 1, (unsigned long long)n_refl);
		if (n_refr) atomicAdd(cnt->stats + 2, (unsigned long long)n_refr);
		if (n_shadow) atomicAdd(cnt->stats + 3, (unsigned long long)n_shadow);
		if (n_culled) atomicAdd(cnt->stats + 6, (unsigned long long)n_culled);
		if (overflow) atomicAdd(cnt->stats + 4, 1ull);
	}
}

// ---- stage 3: Light::shadowAttenuation ------------------------------------------------------------
#ifndef RT_SHADOW_MINB
#define RT_SHADOW_MINB 16 // 64 registers, 16 blocks per SM (re-swept with wide rounds: profiles/r1_variants.txt)
#endif
// Sort key of a shadow ray: light (top bits) | 30-bit Morton code of the origin's cell in a 1024^3 grid over the scene's
// box.  Rays of one light that start close together walk the same subtrees: sorted, the 32 rays of a warp stay
// together in the node loop (the walk-length spread inside a warp is what leaves 7 of 32 lanes active) and their
// node reads hit the same cache lines.
__device__ __forceinline__ unsigned morton_spread10(unsigned v)
{
	v &= 1023u;
	v = (v | (v << 16)) & 0x030000ffu;
	v = (v | (v << 8)) & 0x0300f00fu;
	v = (v | (v << 4)) & 0x030c30c3u;
	v = (v | (v << 2)) & 0x09249249u;
	return v;
}
__global__ void __launch_bounds__(256) k_shadow_keys(const __grid_constant__ DScene S, const RayQueue SQ, long long n, unsigned* keys, unsigned* idx)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= n) return;
	const d3 o = SQ.load_v(0, k);
	unsigned c[3];
	const double oc[3] = {o.x, o.y, o.z};
#pragma unroll
	for (int a = 0; a < 3; a++) {
		const double w = S.stree.hi[a] - S.stree.lo[a];
		double f = w > 0.0 ? (oc[a] - S.stree.lo[a]) / w : 0.0;
		f = f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f);
		c[a] = (unsigned)(f * 1023.0);
	}
	const unsigned m = morton_spread10(c[0]) | (morton_spread10(c[1]) << 1) | (morton_spread10(c[2]) << 2);
	keys[k] = ((unsigned)SQ.i[SQ.cap + k] << 30) | m; // light index modulo 4 in the top two bits: a hint, not a requirement
	idx[k] = (unsigned)k;
}

// sort key of a closest-hit ray: direction octant (3 bits) | 27-bit Morton code of the origin's cell (512^3 grid)
__global__ void __launch_bounds__(256) k_closest_keys(const __grid_constant__ DScene S, const RayQueue Q, long long base, long long n, unsigned* keys, unsigned* idx)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= n) return;
	const d3 o = Q.load_v(0, base + k), d = Q.load_v(3, base + k);
	unsigned c[3];
	const double oc[3] = {o.x, o.y, o.z};
#pragma unroll
	for (int a = 0; a < 3; a++) {
		const double w = S.stree.hi[a] - S.stree.lo[a];
		double f = w > 0.0 ? (oc[a] - S.stree.lo[a]) / w : 0.0;
		f = f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f);
		c[a] = (unsigned)(f * 511.0);
	}
	const unsigned m = morton_spread10(c[0]) | (morton_spread10(c[1]) << 1) | (morton_spread10(c[2]) << 2);
	const unsigned oct = (d.x < 0.0 ? 1u : 0u) | (d.y < 0.0 ? 2u : 0u) | (d.z < 0.0 ? 4u : 0u);
	keys[k] = (oct << 27) | (m & 0x7ffffffu);
	idx[k] = (unsigned)k;
}

// perm: NULL, or the order in which the queue's rays are taken (k_shadow_keys + radix sort).
__global__ void __launch_bounds__(RT_TRACE_THREADS, RT_SHADOW_MINB) k_wave_shadow(const __grid_constant__ DScene S, const __grid_constant__ PoolArgs A, const RayQueue SQ, int parity, const WaveCounters* cnt,
                                                                                const unsigned* __restrict__ perm, int sparse)
{
	RT_STAGE_TOP(S, top);
	const long long t0 = (long long)blockIdx.x * RT_TRACE_THREADS + threadIdx.x;
	if (t0 & ((1ll << sparse) - 1)) return; // one ray per 2^sparse lanes in rounds too small to fill the GPU (see k_wave_closest)
	const long long k0 = t0 >> sparse;
	long long n = (long long)cnt->n_shadow[parity];
	if (n > SQ.cap) n = SQ.cap;
	if (k0 >= n) return;
	const long long k = perm ? (long long)perm[k0] : k0;
	const d3 ro = SQ.load_v(0, k), rd = SQ.load_v(3, k);
	const rt_light& Lt = S.lights[SQ.i[SQ.cap + k]];
	ShadowState st;
	SBuf B;
	TravClosest unused;
	unused.have = false;
	shadow_begin(st, B, Lt, ro);
	for (;;) {
		shadow_pass_reset(st, B);
		traverse<true, TRAV_SHADOW>(S, ro, rd, TRAV_SHADOW, unused, B, 0, top);
		if (shadow_walk(st, B, Lt, rd, A.prm)) break;
	}
	accumulate(A.accum, SQ.i[k], SQ.load_v(6, k) * st.sattn);
}

// ---- stage 3, two-kernel form: decide by classification, walk completely only what needs it ---------------------------
// k_wave_shadow carries everything srsAttenuation can need - the sorted window of hits, kt evaluation (textures, per-vertex
// blends), the attenuation walk with pow - and ncu showed instruction fetch as its largest stall class (7.1 stalled warps per
// issue against 0.5 in k_wave_closest: a 120 KB image).  But a scene's shadow rays are mostly either blocked by an opaque
// surface in the near half of the segment (result exactly 0) or meet nothing (exactly 1): 95 % of C5's.  So:
//   k_wave_shadow_any   every ray: the walk in TRAV_ANYHIT mode (rt_traverse.cuh) - hits are classified where they are found,
//                       an entered opaque surface at t <= 0.49 D stops the walk at once; 0 -> nothing to add, no hit -> the
//                       Phong term is added; anything else (a transmissive surface, a far-half hit, an origin inside an
//                       object, an area or spot light) -> the ray is listed.
//   k_wave_shadow_list  the listed rays: the complete walk, unchanged (persistent grid over the list).
// Both shortcuts are exits the complete walk takes itself, so the frame is the same bit for bit.
#ifndef RT_ANY_MINB
#define RT_ANY_MINB 16
#endif
__global__ void __launch_bounds__(RT_TRACE_THREADS, RT_ANY_MINB) k_wave_shadow_any(const __grid_constant__ DScene S, const __grid_constant__ PoolArgs A, const RayQueue SQ, int parity, WaveCounters* cnt,
                                                                                 const unsigned* __restrict__ perm, int sparse, int* complex_list)
{
	RT_STAGE_TOP(S, top);
	const unsigned lane = threadIdx.x & 31;
	const long long t0 = (long long)blockIdx.x * RT_TRACE_THREADS + threadIdx.x;
	const long long k0 = t0 >> sparse;
	long long n = (long long)cnt->n_shadow[parity];
	if (n > SQ.cap) n = SQ.cap;
	const bool valid = (t0 & ((1ll << sparse) - 1)) == 0 && k0 < n; // (no early return: the listing below is a warp operation)
	bool complex_ray = false;
	long long k = 0;
	if (valid) {
		k = perm ? (long long)perm[k0] : k0;
		const d3 ro = SQ.load_v(0, k), rd = SQ.load_v(3, k);
		const rt_light& Lt = S.lights[SQ.i[SQ.cap + k]];
		AnyHit Y;
		Y.limit = RT_T_INF;
		Y.sure_t = RT_T_INF;
		Y.fl = 0;
		Y.t_x = RT_T_INF;
		Y.t_o = RT_T_INF;
		if (Lt.type == RT_LIGHT_POINT) { // shadow_begin's numbers
			const double D = length(ld3(Lt.position) - ro);
			Y.limit = D * (1.0 + 1e-9) + 1e-9;
			Y.sure_t = 0.49 * D;
		} else if (Lt.type != RT_LIGHT_DIRECTIONAL) Y.fl = SH_OTHER; // area and spot lights: no any-hit exit, the complete walk
		if (!Y.fl) {
			TravClosest unused_c;
			SBuf* unused_b = nullptr;
			traverse<true, TRAV_ANYHIT>(S, ro, rd, TRAV_ANYHIT, unused_c, *unused_b, 0, top, &Y);
		}
		if (!(Y.fl & SH_DEAD) && !Y.settled_by_exit()) {
			if (Y.fl & SH_OTHER) complex_ray = true;
			else accumulate(A.accum, SQ.i[k], SQ.load_v(6, k)); // nothing before the light: attenuation exactly (1,1,1)
		}
	}
	const long long s = wave_reserve(complex_ray, (1u << lane) - 1u, &cnt->n_complex[parity]);
	if (s >= 0) complex_list[s] = (int)k;
}

__global__ void __launch_bounds__(RT_TRACE_THREADS, RT_SHADOW_MINB) k_wave_shadow_list(const __grid_constant__ DScene S, const __grid_constant__ PoolArgs A, const RayQueue SQ, int parity, const WaveCounters* cnt,
                                                                                     const int* __restrict__ complex_list)
{
	RT_STAGE_TOP(S, top);
	const long long n = (long long)cnt->n_complex[parity];
	for (long long j = (long long)blockIdx.x * RT_TRACE_THREADS + threadIdx.x; j < n; j += (long long)gridDim.x * RT_TRACE_THREADS) {
		const long long k = complex_list[j];
		const d3 ro = SQ.load_v(0, k), rd = SQ.load_v(3, k);
		const rt_light& Lt = S.lights[SQ.i[SQ.cap + k]];
		ShadowState st;
		SBuf B;
		TravClosest unused;
		unused.have = false;
		shadow_begin(st, B, Lt, ro);
		for (;;) {
			shadow_pass_reset(st, B);
			traverse<true, TRAV_SHADOW>(S, ro, rd, TRAV_SHADOW, unused, B, 0, top);
			if (shadow_walk(st, B, Lt, rd, A.prm)) break;
		}
		accumulate(A.accum, SQ.i[k], SQ.load_v(6, k) * st.sattn);
	}
}

// spill / take back: n entries of a 12-double queue, src[src0 ...] -> dst[dst0 ...]
__global__ void __launch_bounds__(256) k_wave_move(const RayQueue src, long long src0, const RayQueue dst, long long dst0, long long n)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= n) return;
#pragma unroll
	for (int f = 0; f < 12; f++) dst.d[f * dst.cap + dst0 + k] = src.d[f * src.cap + src0 + k];
	dst.i[dst0 + k] = src.i[src0 + k];
	dst.i[dst.cap + dst0 + k] = src.i[src.cap + src0 + k];
}

} // namespace rtd
