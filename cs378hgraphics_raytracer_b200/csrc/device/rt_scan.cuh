// rt_scan.cuh - Light::shadowAttenuation (reference scene/light.cpp:16-53) as a wavefront INSIDE the walk:
// the shadow stage split by what the code does, not by ray.
//
// Why.  ncu on the single-kernel shadow walk (profiles/r2_v1_c5_k_wave_shadow.txt): the float node loop is
// 43 % of the executed instructions but 18 % of the stall samples; the FP64 candidate code behind it (object
// box test, world -> object transform, five primitive kinds, sorted hit window) is the rest - 60-90 % of ITS
// samples are "no instruction": ~90 KB of code touched by 32 warps that each sit in a different primitive
// function, against a 32 KB instruction cache, at 3-4 active lanes.  So the two halves become separate kernels:
//
//   k_shadow_scan        one thread per shadow ray: the float node walk of rt_traverse.cuh (TRAV_SCAN) over the
//                        whole segment origin -> light, testing nothing; every leaf candidate the float box test
//                        cannot rule out is appended to the queue of its primitive kind (warp-aggregated
//                        atomics).  A trimesh candidate is entered (object-space walk) and its leaf triangles
//                        are queued.  A small kernel image, no FP64 in the loop.
//   k_shadow_cand<KIND>  one thread per queued (ray, candidate): the reference's own arithmetic - world box
//                        pretest (scene.cpp:43), object-space ray (scene.cpp:17-20), intersectLocalList of that
//                        ONE kind - with every lane of a warp in the same function.  It does not build the
//                        sorted hit list; it classifies: an ENTERED NON-TRANSMISSIVE surface at t <= sure_t
//                        settles srsAttenuation at exactly 0 (the any-hit argument of rt_shadow.cuh: sbuf_add),
//                        any other hit within the light's distance marks the ray "complex".
//   k_shadow_resolve     per ray: settled -> contributes 0; no hit at all -> attenuation is exactly (1,1,1), the
//                        Phong term is added to the sample; complex (a transmissive or far-half hit, a ray that
//                        starts inside an object, area / spot lights, queue overflow) -> listed for
//   k_wave_shadow_list   the complete single-thread walk with its sorted window (rt_wave.cuh), unchanged.
//
// The result is the one the single kernel gives, bit for bit: both shortcuts (settled = 0, no hit = 1) are
// exits that walk takes itself, and everything else still goes through it.
#pragma once

namespace rtd {

enum { SH_DEAD = 1, SH_OTHER = 2, SH_FORCE = 4 };
enum { CQ_KINDS = 6 }; // queue k holds candidates of rt_object_type k; RT_OBJ_TRIMESH's queue holds leaf triangles

// what k_shadow_scan leaves per ray for the candidate kernels
struct ShadowSide {
	double limit;  // hits beyond this cannot change the result (past a point light)
	double sure_t; // an entered opaque surface at t <= sure_t settles the result at 0
	double y[3];   // 1/d (IEEE), the slab test's per-ray reciprocals
};

struct CandQueues {
	uint2* q[CQ_KINDS - 1];             // analytic kinds: (ray, scene-level candidate rank)
	uint4* tri;                         // (ray, object, global leaf item position, -)
	unsigned long long cap[CQ_KINDS];
	unsigned long long* count;          // [CQ_KINDS] entries appended (may exceed cap: the excess was refused)
	unsigned long long* n_complex;      // rays listed for the complete walk
	int* complex_list;
};

// slot for one candidate of `kind` per calling lane: lanes of the warp that are here together and have the same
// kind share one atomic.  Returns -1 when the queue is full.
__device__ __forceinline__ long long cand_reserve(const CandQueues& Q, int kind)
{
	const unsigned act = __activemask();
	const unsigned grp = __match_any_sync(act, kind);
	const int leader = __ffs(grp) - 1;
	const unsigned lane = threadIdx.x & 31;
	unsigned long long base = 0;
	if ((int)lane == leader) base = atomicAdd(Q.count + kind, (unsigned long long)__popc(grp));
	base = __shfl_sync(grp, base, leader);
	const unsigned long long slot = base + __popc(grp & ((1u << lane) - 1u));
	return slot < Q.cap[kind] ? (long long)slot : -1;
}

#ifndef RT_SCAN_MINB
#define RT_SCAN_MINB 16 // 64 registers: the listing walk keeps only the ray's float image, the stack pointer and the limit
#endif
__global__ void __launch_bounds__(RT_TRACE_THREADS, RT_SCAN_MINB) k_shadow_scan(const __grid_constant__ DScene S, const RayQueue SQ, int parity, const WaveCounters* cnt,
                                                                                ShadowSide* side, unsigned* flags, const __grid_constant__ CandQueues Q)
{
	const long long k = (long long)blockIdx.x * RT_TRACE_THREADS + threadIdx.x;
	long long n = (long long)cnt->n_shadow[parity];
	if (n > SQ.cap) n = SQ.cap;
	if (k >= n) return;
	const rt_light& Lt = S.lights[SQ.i[SQ.cap + k]];
	Walk W;
	W.ow = SQ.load_v(0, k);
	W.dw = SQ.load_v(3, k);
	// shadow_begin's limit and any-hit bound (rt_traverse.cuh); area and spot lights take the complete walk
	unsigned f = 0;
	double limit = RT_T_INF, sure = RT_T_INF;
	if (Lt.type == RT_LIGHT_POINT) {
		const double D = length(ld3(Lt.position) - W.ow);
		limit = D * (1.0 + 1e-9) + 1e-9;
		sure = 0.49 * D;
	} else if (Lt.type != RT_LIGHT_DIRECTIONAL) f = SH_FORCE;
	W.C = nullptr;
	W.B = nullptr;
	W.I.y = mk3(0.0, 0.0, 0.0);
	W.mode = TRAV_SCAN;
	W.prune = true;
	W.lim0 = limit;
	StackEnt stk[RT_STACK];
	RayF F;
	float limf;
	const DNode2f* nodesf = S.snodes2f;
	int level = 0, sp = 0;
	int cur = f ? RT_REF_NONE : walk_begin(S, W, F, limf);
	ShadowSide sd;
	sd.limit = limit;
	sd.sure_t = sure;
	sd.y[0] = W.I.y.x; sd.y[1] = W.I.y.y; sd.y[2] = W.I.y.z;
	side[k] = sd;
	if (cur != RT_REF_NONE) {
		for (;;) {
#pragma unroll 1
			while (cur >= 0) {
				const DNode2f nf = ld_node2f(nodesf + cur);
				float e0, e1;
				int r0 = boxf(nf.lo0, nf.hi0, F, e0);
				int r1 = boxf(nf.lo1, nf.hi1, F, e1);
				if (r0 == 2) { const StackEnt r = walk_resolve(W, cur, 0); r0 = r.ref; e0 = r.e; }
				if (r1 == 2) { const StackEnt r = walk_resolve(W, cur, 1); r1 = r.ref; e1 = r.e; }
				const bool h0 = r0 != 0 && !(e0 > limf), h1 = r1 != 0 && !(e1 > limf);
				if (h0 && h1) {
					const bool swap = e1 < e0;
					stk[sp].ref = swap ? nf.c0 : nf.c1;
					stk[sp].e = swap ? e0 : e1;
					sp++;
					cur = swap ? nf.c1 : nf.c0;
				} else cur = h0 ? nf.c0 : (h1 ? nf.c1 : RT_REF_NONE);
			}
			if (cur != RT_REF_NONE) {
				const int first = leaf_first(cur), count = leaf_count(cur);
				cur = RT_REF_NONE;
				if (level == 1) {
					// a mesh leaf: its triangles are candidates as they are (the leaf's box was entered)
#pragma unroll 1
					for (int i = 0; i < count; i++) {
						const long long s = cand_reserve(Q, RT_OBJ_TRIMESH);
						if (s >= 0) Q.tri[s] = make_uint4((unsigned)k, (unsigned)W.V.obj, (unsigned)(W.V.item_off + first + i), 0u);
						else f |= SH_FORCE;
					}
				} else {
#pragma unroll 1
					for (int i = 0; i < count; i++) {
						const DItemBox& ib = S.sboxf[first + i];
						if (box_miss_f(ib, F, limf)) continue;
						const int kind = ib.type;
						if (kind == RT_OBJ_TRIMESH) {
							// Geometry::intersectList's world-box pretest (scene.cpp:43), then the mesh's own tree
							const rt_object& ob = S.objs[ib.obj];
							double eb;
							if (!box_hit(ob.bbox, ob.bbox + 3, W.o, W.d, W.I, eb) || eb > W.lim) continue;
							LocalRay L;
							to_local(L, ob, W.o, W.d);
							cur = walk_enter_mesh(S, W, L, first, count, i, F, nodesf, level, sp, stk);
							limf = f_up(W.lim);
							break;
						}
						const long long s = cand_reserve(Q, kind);
						if (s >= 0) Q.q[kind][s] = make_uint2((unsigned)k, (unsigned)(first + i));
						else f |= SH_FORCE;
					}
				}
			}
			if (cur == RT_REF_NONE) {
#pragma unroll 1
				for (;;) {
					if (sp == 0) break;
					sp--;
					const StackEnt a = stk[sp];
					if (a.ref == RT_REF_MESH_EXIT) {
						walk_mesh_exit<TRAV_SCAN>(S, W, F, nodesf, limf, level, sp);
						continue;
					}
					if (a.e > limf) continue;
					cur = a.ref;
					break;
				}
				if (cur == RT_REF_NONE) break;
			}
		}
	}
	flags[k] = f; // the candidate kernels (later launches on this stream) OR their findings into it
}

// sbuf_add's classification of one hit (rt_shadow.cuh) without the list
__device__ __forceinline__ unsigned classify_hit(double t, bool inside, bool trans, const ShadowSide& sd)
{
	if (!(t <= sd.limit)) return 0u;
	return (!inside && !trans && t <= sd.sure_t) ? SH_DEAD : SH_OTHER;
}

struct ClassifyEmit {
	const DScene& S;
	const rt_object& ob;
	const d3& dw;
	double len;
	const ShadowSide& sd;
	unsigned& fl;
	__device__ void operator()(double tl, d3 Nl, d2 uv, int) const
	{
		const double t = tl / len;
		if (!(t <= sd.limit)) return;
		const d3 N = world_normal(S, ob, Nl, uv);
		fl |= classify_hit(t, dot(N, dw) > 0, S.mats[ob.material].trans != 0, sd);
	}
};

// RayInv of a world-space ray from the reciprocals k_shadow_scan stored (no pruning here: slack 0)
__device__ __forceinline__ void side_inv(RayInv& I, const d3& d, const ShadowSide& sd)
{
	I.y = mk3(sd.y[0], sd.y[1], sd.y[2]);
	I.slmax = 0.0;
	I.guard = (div_range_ok(d.x) && div_range_ok(d.y) && div_range_ok(d.z)) ? -1.0f : NAN;
}

// One (ray, candidate) pair of the shadow query, KIND known at compile time: the reference's own arithmetic - world
// box pretest (scene.cpp:43), object-space ray (scene.cpp:17-20), intersectLocalList of that kind - and the
// classification of its hits (classify_hit).  pos = global leaf item position (triangles only).
template <int KIND>
__device__ __forceinline__ unsigned cand_classify(const DScene& S, const d3& o, const d3& d, const ShadowSide& sd, unsigned obj, unsigned pos)
{
	const rt_object& ob = S.objs[obj];
	RayInv I;
	side_inv(I, d, sd);
	double eb;
	if (!box_hit(ob.bbox, ob.bbox + 3, o, d, I, eb)) return 0u; // scene.cpp:43
	LocalRay L;
	to_local(L, ob, o, d);
	unsigned fl = 0;
	if (KIND == RT_OBJ_TRIMESH) {
		// TrimeshFace::intersectLocal and the isect it fills (trimesh.cpp:119-188), as emit_face
		Hit h;
		h.obj = (int)obj;
		h.fpos = (int)pos;
		h.face = S.mitems[pos];
		h.uv.x = h.uv.y = 0.0;
		double tl;
		if (!tri_hit<true>(S.tris[pos], L.o, L.d, tl, h.bary)) return 0u;
		const double t = tl / L.len;
		if (!(t <= sd.limit)) return 0u;
		const rt_mesh& M = S.meshes[ob.mesh];
		const rt_face& fc = S.faces[M.face_off + h.face];
		d3 Nl;
		if (M.vnorm_off >= 0) {
			const double* n0 = S.vnormals + 3 * (M.vnorm_off + fc.ids[0]);
			const double* n1 = S.vnormals + 3 * (M.vnorm_off + fc.ids[1]);
			const double* n2 = S.vnormals + 3 * (M.vnorm_off + fc.ids[2]);
			Nl = normalize(mk3(n0[0] * h.bary.x + n1[0] * h.bary.y + n2[0] * h.bary.z,
			                   n0[1] * h.bary.x + n1[1] * h.bary.y + n2[1] * h.bary.z,
			                   n0[2] * h.bary.x + n1[2] * h.bary.y + n2[2] * h.bary.z));
		} else {
			Nl = ld3(fc.normal);
		}
		const d3 N = normalize(mul33(ob.normi, Nl));
		d3 kt;
		bool trans;
		eval_kt(S, h, kt, trans);
		fl = classify_hit(t, dot(N, d) > 0, trans, sd);
	} else {
		const ClassifyEmit emit{S, ob, d, L.len, sd, fl};
		if (KIND == RT_OBJ_SPHERE) sphere_list(L.o, L.d, emit);
		else if (KIND == RT_OBJ_BOX) box_list(L.o, L.d, emit);
		else if (KIND == RT_OBJ_SQUARE) {
			LocalHit lh;
			lh.uv.x = lh.uv.y = 0.0;
			if (square_hit<true>(L.o, L.d, lh)) emit(lh.t, lh.N, lh.uv, 0);
		} else if (KIND == RT_OBJ_CYLINDER) cylinder_list(L.o, L.d, emit);
		else cone_list(L.o, L.d, ob.cone, ob.capped, emit);
	}
	return fl;
}

// The shadow cache's test of ONE known candidate (k_wave_shadow phase 2): true when it alone settles srsAttenuation at 0.
__device__ __noinline__ unsigned cand_test(const DScene& S, const d3& o, const d3& d, const rt_light& Lt, unsigned obj, unsigned pos)
{
	ShadowSide sd;
	sd.limit = RT_T_INF;
	sd.sure_t = RT_T_INF;
	if (Lt.type == RT_LIGHT_POINT) { // shadow_begin's numbers (rt_traverse.cuh)
		const double D = length(ld3(Lt.position) - o);
		sd.limit = D * (1.0 + 1e-9) + 1e-9;
		sd.sure_t = 0.49 * D;
	} else if (Lt.type != RT_LIGHT_DIRECTIONAL) return 0u; // area and spot lights take no any-hit exit
	sd.y[0] = 1.0 / d.x; sd.y[1] = 1.0 / d.y; sd.y[2] = 1.0 / d.z;
	unsigned fl = 0;
	switch (S.objs[obj].type) {
	case RT_OBJ_SPHERE: fl = cand_classify<RT_OBJ_SPHERE>(S, o, d, sd, obj, pos); break;
	case RT_OBJ_BOX: fl = cand_classify<RT_OBJ_BOX>(S, o, d, sd, obj, pos); break;
	case RT_OBJ_SQUARE: fl = cand_classify<RT_OBJ_SQUARE>(S, o, d, sd, obj, pos); break;
	case RT_OBJ_CYLINDER: fl = cand_classify<RT_OBJ_CYLINDER>(S, o, d, sd, obj, pos); break;
	case RT_OBJ_CONE: fl = cand_classify<RT_OBJ_CONE>(S, o, d, sd, obj, pos); break;
	case RT_OBJ_TRIMESH: fl = cand_classify<RT_OBJ_TRIMESH>(S, o, d, sd, obj, pos); break;
	default: break;
	}
	return fl & SH_DEAD;
}

#define RT_CAND_THREADS 128
template <int KIND>
__global__ void __launch_bounds__(RT_CAND_THREADS) k_shadow_cand(const __grid_constant__ DScene S, const RayQueue SQ, const ShadowSide* __restrict__ side, unsigned* flags,
                                                                 const __grid_constant__ CandQueues Q)
{
	unsigned long long n = Q.count[KIND];
	if (n > Q.cap[KIND]) n = Q.cap[KIND];
	const unsigned long long stride = (unsigned long long)gridDim.x * RT_CAND_THREADS;
	for (unsigned long long j = (unsigned long long)blockIdx.x * RT_CAND_THREADS + threadIdx.x; j < n; j += stride) {
		unsigned ray, obj, pos = 0;
		if (KIND == RT_OBJ_TRIMESH) {
			const uint4 c = Q.tri[j];
			ray = c.x; obj = c.y; pos = c.z;
		} else {
			const uint2 c = Q.q[KIND][j];
			ray = c.x;
			obj = (unsigned)S.sboxf[c.y].obj;
		}
		if (__ldcg(flags + ray) & SH_DEAD) continue; // already settled by another candidate
		const d3 o = SQ.load_v(0, ray), d = SQ.load_v(3, ray);
		const ShadowSide sd = side[ray];
		const unsigned fl = cand_classify<KIND>(S, o, d, sd, obj, pos);
		if (fl) atomicOr(flags + ray, fl);
	}
}

// per ray: contribute, or list for the complete walk
__global__ void __launch_bounds__(256) k_shadow_resolve(const __grid_constant__ PoolArgs A, const RayQueue SQ, int parity, const WaveCounters* cnt, const unsigned* flags, const __grid_constant__ CandQueues Q)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	long long n = (long long)cnt->n_shadow[parity];
	if (n > SQ.cap) n = SQ.cap;
	bool complex_ray = false;
	if (k < n) {
		const unsigned f = flags[k];
		if (f & SH_DEAD) { /* srsAttenuation returns dvec3(0): nothing to add */ }
		else if (f & (SH_OTHER | SH_FORCE)) complex_ray = true;
		else accumulate(A.accum, SQ.i[k], SQ.load_v(6, k)); // no hit: the attenuation is exactly (1,1,1)
	}
	const unsigned lane = threadIdx.x & 31;
	const long long s = wave_reserve(complex_ray, (1u << lane) - 1u, Q.n_complex);
	if (s >= 0) Q.complex_list[s] = (int)k;
}

// The complete walk (k_wave_shadow's body) for the listed rays; persistent grid over the list.
__global__ void __launch_bounds__(RT_TRACE_THREADS, RT_SHADOW_MINB) k_wave_shadow_list(const __grid_constant__ DScene S, const __grid_constant__ PoolArgs A, const RayQueue SQ, const __grid_constant__ CandQueues Q)
{
	const long long n = (long long)*Q.n_complex;
	for (long long j = (long long)blockIdx.x * RT_TRACE_THREADS + threadIdx.x; j < n; j += (long long)gridDim.x * RT_TRACE_THREADS) {
		const long long k = Q.complex_list[j];
		const d3 ro = SQ.load_v(0, k), rd = SQ.load_v(3, k);
		const rt_light& Lt = S.lights[SQ.i[SQ.cap + k]];
		ShadowState st;
		SBuf B;
		TravClosest unused;
		unused.have = false;
		shadow_begin(st, B, Lt, ro);
		for (;;) {
			shadow_pass_reset(st, B);
			traverse<true, TRAV_SHADOW>(S, ro, rd, TRAV_SHADOW, unused, B);
			if (shadow_walk(st, B, Lt, rd, A.prm)) break;
		}
		accumulate(A.accum, SQ.i[k], SQ.load_v(6, k) * st.sattn);
	}
}

} // namespace rtd
