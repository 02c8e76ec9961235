// rt_pool.cuh - k_trace_pool: the persistent-warp wavefront kernel (see rt_kernels.cuh header).
//
// Per-warp pool entry (SoA slab in HBM/L2, `cap` entries): 12 doubles + 2 ints
//   closest-hit ray : o[3] d[3] weight[3] att_kt[3] | sample | depth, att_mode, kind
//   shadow ray      : o[3] d[3] term[3]   -          | sample | light, kind = 3
// term = weight * (dattn * Lc) * (kd*max(0,N.L) + ks*max(0,R.V)^sh): the shadow ray's result is
// term * sattn, added to the sample when the ray's traversal and attenuation walk finish.
#pragma once

namespace rtd {

enum { KIND_CAMERA = 0, KIND_REFL = 1, KIND_REFR = 2, KIND_SHADOW = 3 };

#ifndef RT_POOL_THREADS
#define RT_POOL_THREADS 128
#endif
#ifndef RT_POOL_MINB
#define RT_POOL_MINB 3 // resident CTAs per SM the register allocation must allow
#endif

struct PoolSlab {
	double* d;
	int* i;
	int cap;
	__device__ __forceinline__ void store_ray(int slot, const d3& o, const d3& dir) const
	{
		d[0 * cap + slot] = o.x; d[1 * cap + slot] = o.y; d[2 * cap + slot] = o.z;
		d[3 * cap + slot] = dir.x; d[4 * cap + slot] = dir.y; d[5 * cap + slot] = dir.z;
	}
	__device__ __forceinline__ void store_v(int field, int slot, const d3& v) const
	{
		d[(field + 0) * cap + slot] = v.x; d[(field + 1) * cap + slot] = v.y; d[(field + 2) * cap + slot] = v.z;
	}
	__device__ __forceinline__ d3 load_v(int field, int slot) const
	{
		return mk3(d[(field + 0) * cap + slot], d[(field + 1) * cap + slot], d[(field + 2) * cap + slot]);
	}
	__device__ __forceinline__ void store_meta(int slot, int sample, int meta) const
	{
		i[slot] = sample;
		i[cap + slot] = meta;
	}
};
__device__ __forceinline__ int pack_meta(int depth, int att_mode, int kind, int light) { return depth | (att_mode << 8) | (kind << 12) | (light << 16); }

// Pushes one entry per lane with pred set: ballot-compacted onto the top of the warp's pool.
// Returns this lane's slot (or -1); count is warp-uniform and advanced by the number pushed.
__device__ __forceinline__ int pool_reserve(bool pred, unsigned lt_mask, int& count, int cap, unsigned& overflow)
{
	const unsigned m = __ballot_sync(0xffffffffu, pred);
	const int n = __popc(m);
	if (n == 0) return -1;
	if (count + n > cap) { overflow = 1; return -1; }
	const int slot = count + __popc(m & lt_mask);
	count += n;
	return pred ? slot : -1;
}

// one term of a sample's colour (rt_accum.cuh: integer limbs, the same total in any order)
__device__ __forceinline__ void accumulate(long long* accum, int sample, const d3& c)
{
	long long* acc = accum + (long long)RT_ACC_STRIDE * sample;
	if (c.x != 0.0) acc_add(acc + 0, c.x);
	if (c.y != 0.0) acc_add(acc + 2, c.y);
	if (c.z != 0.0) acc_add(acc + 4, c.z);
}

// a ray's own term (emission + ambient, or the background): a camera ray's goes to the slot only it owns, a plain store
__device__ __forceinline__ void add_term(const PoolArgs& A, int sample, bool camera, int lens, const d3& c)
{
	if (camera && A.base) {
		double* b = A.base + 3ll * ((long long)sample * A.lenses + lens);
		b[0] = c.x; b[1] = c.y; b[2] = c.z;
	} else {
		accumulate(A.accum, sample, c);
	}
}

// DEFER = true : shadow rays are pool entries (normal case)
// DEFER = false: shadow rays traced inline by the shading lane (adaptive termination needs the
//                shadowed colour before deciding to recurse, RayTracer.cpp:124-125)
template <bool DEFER>
__global__ void __launch_bounds__(RT_POOL_THREADS, RT_POOL_MINB) k_trace_pool(const __grid_constant__ DScene S, const __grid_constant__ PoolArgs A)
{
	const unsigned lane = threadIdx.x & 31;
	const unsigned lt_mask = (1u << lane) - 1u;
	const long long warp_global = (long long)blockIdx.x * (RT_POOL_THREADS / 32) + (threadIdx.x >> 5);
	PoolSlab P;
	P.cap = A.cap;
	P.d = A.pool_d + warp_global * (long long)A.cap * RT_POOL_DOUBLES;
	P.i = A.pool_i + warp_global * (long long)A.cap * 2;

	int count = 0;                       // entries in this warp's pool (warp-uniform)
	long long cam_next = 0, cam_end = 0; // reserved slice of the global cursor (warp-uniform)
	bool exhausted = false;
	// the ray this lane is working on
	bool has_task = false;
	d3 ro, rd;
	int sample = 0, meta = 0, slot = -1; // slot: where the rest of the entry still lies (-1: fresh camera ray, < -1: entry -2 - slot of the source queue)
	unsigned n_cam = 0, n_refl = 0, n_refr = 0, n_shadow = 0, overflow = 0;

	for (;;) {
		// ---- refill idle lanes: pool first (top of the stack), then fresh primaries -------------
		const unsigned idle_mask = __ballot_sync(0xffffffffu, !has_task);
		const int n_idle = __popc(idle_mask);
		if (n_idle) {
			const int rank = __popc(idle_mask & lt_mask);
			const int take = n_idle < count ? n_idle : count;
			if (!has_task && rank < take) {
				slot = count - 1 - rank;
				ro = P.load_v(0, slot);
				rd = P.load_v(3, slot);
				sample = P.i[slot];
				meta = P.i[P.cap + slot];
				has_task = true;
			}
			count -= take;
			int need = n_idle - take;
			int served = take;
			while (need > 0 && !exhausted) {
				if (cam_next >= cam_end) {
					unsigned long long base = 0;
					if (lane == 0) base = atomicAdd(A.cursor, (unsigned long long)RT_CHUNK);
					base = __shfl_sync(0xffffffffu, base, 0);
					if ((long long)base >= A.n_slots) { exhausted = true; break; }
					cam_next = (long long)base;
					cam_end = cam_next + RT_CHUNK < A.n_slots ? cam_next + RT_CHUNK : A.n_slots;
				}
				const int avail = (int)(cam_end - cam_next);
				const int give = need < avail ? need : avail;
				if (!has_task && rank >= served && rank < served + give) {
					const long long s = cam_next + (rank - served);
					const long long smp = s / A.lenses;
					const int lens = (int)(s % A.lenses);
					bool ok = true;
					if (A.src == SRC_QUEUE) {
						// a ray some wavefront round left in its queue (rt_wave.cuh: RayQueue, 12 doubles + sample + meta)
						ro = mk3(A.q_d[0 * A.q_cap + s], A.q_d[1 * A.q_cap + s], A.q_d[2 * A.q_cap + s]);
						rd = mk3(A.q_d[3 * A.q_cap + s], A.q_d[4 * A.q_cap + s], A.q_d[5 * A.q_cap + s]);
						sample = A.q_i[s];
						meta = A.q_i[A.q_cap + s];
						slot = -2 - (int)s; // its weight and attenuation are still there
						has_task = true;
						ok = false;         // (nothing more to set up below)
					} else if (A.src == SRC_RAYS) {
						ro = ld3(A.rays[s].p);
						rd = ld3(A.rays[s].d);
					} else {
						double x, y;
						if (A.src == SRC_GRID) {
							int pi_, pj_;
							ok = grid_sample(A, smp, pi_, pj_, x, y);
						} else {
							x = A.xs[smp];
							y = A.ys[smp];
						}
						if (ok) primary_ray(S.cam, A.prm, x, y, lens, A.eye_shift, ro, rd);
					}
					if (ok) {
						sample = (int)smp;
						meta = pack_meta(A.depth, 0, KIND_CAMERA, 0);
						slot = -1;
						has_task = true;
					}
				}
				cam_next += give;
				served += give;
				need -= give;
			}
		}
		if (!__any_sync(0xffffffffu, has_task)) {
			if (exhausted && count == 0) break;
			continue;
		}
		// ---- traversal: closest-hit and shadow rays walk the trees in the same loop ---------------
		const int kind = (meta >> 12) & 15;
		const int mode = (DEFER && kind == KIND_SHADOW) ? TRAV_SHADOW : TRAV_CLOSEST;
		TravClosest C;
		C.have = false;
		C.best_t = RT_T_INF;
		C.best_rank = 0x7fffffff;
		C.best_obj = -1;
		C.best_fpos = -1;
		SBuf B;
		ShadowState st;
		if (has_task) {
			const rt_light& Lt = S.lights[mode == TRAV_SHADOW ? (meta >> 16) : 0];
			if (mode == TRAV_SHADOW) shadow_begin(st, B, Lt, ro);
			else { B.n = 0; B.cut = RT_T_INF; }
			for (;;) {
				if (mode == TRAV_SHADOW) shadow_pass_reset(st, B);
				traverse<true>(S, ro, rd, mode, C, B);
				if (mode == TRAV_CLOSEST || shadow_walk(st, B, Lt, rd, A.prm)) break;
			}
		}
		// ---- results ------------------------------------------------------------------------------
		if (has_task && mode == TRAV_SHADOW) accumulate(A.accum, sample, P.load_v(6, slot) * st.sattn);
		bool lit = false, has_refl = false, has_refr = false;
		RayTask kid[2];
		MatEval m;
		Hit h;
		d3 w = mk3(0, 0, 0), hitP = mk3(0, 0, 0);
		if (has_task && mode == TRAV_CLOSEST) {
			d3 weight = mk3(1.0, 1.0, 1.0), att_kt = mk3(1.0, 1.0, 1.0);
			if (slot >= 0) { weight = P.load_v(6, slot); att_kt = P.load_v(9, slot); }
			else if (slot < -1) {
				const long long qi = -2 - slot;
				weight = mk3(A.q_d[6 * A.q_cap + qi], A.q_d[7 * A.q_cap + qi], A.q_d[8 * A.q_cap + qi]);
				att_kt = mk3(A.q_d[9 * A.q_cap + qi], A.q_d[10 * A.q_cap + qi], A.q_d[11 * A.q_cap + qi]);
			}
			if (kind == KIND_CAMERA) n_cam++;
			else if (kind == KIND_REFL) n_refl++;
			else n_refr++;
			if (!C.have) {
				accumulate(A.accum, sample, weight * miss_colour(S, A.prm, rd));
			} else {
				h.obj = C.best_obj;
				h.fpos = C.best_fpos;
				h.t = C.best_t;
				finalize_hit(S, ro, rd, h);
				w = apply_att(weight, att_kt, (meta >> 8) & 15, h.t);
				eval_material(S, h, m);
				const int depth = (meta & 255) - 1;
				if (DEFER) {
					accumulate(A.accum, sample, w * (m.ke + m.ka * ld3(S.glob.ambient)));
					hitP = ro + h.t * rd;
					lit = true;
					spawn_children(m, h, ro, rd, w, depth, kid, has_refl, has_refr);
				} else {
					unsigned sh = 0;
					const d3 col = shade(S, A.prm, m, ro, rd, h, sh);
					n_shadow += sh;
					accumulate(A.accum, sample, w * col);
					if (!(A.prm.aterm_thresh > 0.0 && dot(col, col) < A.prm.aterm_thresh)) spawn_children(m, h, ro, rd, w, depth, kid, has_refl, has_refr);
				}
			}
		}
		// ---- push new work: children first, then shadow rays light by light (top of the stack) ---
		__syncwarp();
		for (int k = 0; k < 2; k++) {
			const int s = pool_reserve(k == 0 ? has_refl : has_refr, lt_mask, count, P.cap, overflow);
			if (s >= 0) {
				P.store_ray(s, kid[k].o, kid[k].d);
				P.store_v(6, s, kid[k].weight);
				P.store_v(9, s, kid[k].att_kt);
				P.store_meta(s, sample, pack_meta(kid[k].depth, kid[k].att_mode, kid[k].kind, 0));
			}
		}
		if (DEFER && __any_sync(0xffffffffu, lit)) {
			for (int li = 0; li < S.n_lights; li++) {
				const int s = pool_reserve(lit, lt_mask, count, P.cap, overflow);
				if (s >= 0) {
					d3 term, pb, sdir;
					light_term(S.lights[li], m, hitP, h.N, rd, term, pb, sdir);
					P.store_ray(s, pb, sdir);
					P.store_v(6, s, w * term);
					P.store_meta(s, sample, pack_meta(0, 0, KIND_SHADOW, li));
					n_shadow++;
				}
			}
		}
		has_task = false;
	}
	// ---- ray counters ------------------------------------------------------------------------------
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		n_cam += __shfl_down_sync(0xffffffffu, n_cam, o);
		n_refl += __shfl_down_sync(0xffffffffu, n_refl, o);
		n_refr += __shfl_down_sync(0xffffffffu, n_refr, o);
		n_shadow += __shfl_down_sync(0xffffffffu, n_shadow, o);
		overflow |= __shfl_down_sync(0xffffffffu, overflow, o);
	}
	if (lane == 0) {
		if (n_cam) atomicAdd(A.stats + 0, (unsigned long long)n_cam);
		if (n_refl) atomicAdd(A.stats + 1, (unsigned long long)n_refl);
		if (n_refr) atomicAdd(A.stats + 2, (unsigned long long)n_refr);
		if (n_shadow) atomicAdd(A.stats + 3, (unsigned long long)n_shadow);
		if (overflow) atomicAdd(A.stats + 4, 1ull);
	}
}

} // namespace rtd
