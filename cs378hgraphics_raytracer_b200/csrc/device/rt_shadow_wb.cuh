// rt_shadow_wb.cuh - Light::shadowAttenuation (reference scene/light.cpp:16-53): the shadow walk with its leaf
// candidates BATCHED PER WARP AND PRIMITIVE KIND.
//
// What the measurements said (profiles/r2_variants.txt): in the one-kernel walk the float node loop runs with 7-8
// lanes and is cheap; the FP64 candidate code behind it (object box test, world -> object ray, six primitive
// kinds, the sorted hit window) runs with 3-4 lanes, each warp in a different function of a 120 KB image, and
// 60-90 % of its stall samples are instruction fetch.  Run with every lane in the same function, ALL candidate
// tests of a C5 frame cost 67 ms (rt_scan.cuh's k_shadow_cand) - but a walk that only lists candidates and leaves
// the tests to later kernels loses the any-hit exit and visits several times the nodes.  So here the hand-over
// happens inside the warp:
//   * every lane walks its own ray (the float node loop of rt_traverse.cuh) and, at a leaf, PUSHES the candidates
//     the float box test cannot rule out onto the warp's shared-memory queue of that primitive kind;
//   * whenever a queue holds 32 entries the warp tests them together, one (ray, candidate) pair per lane, all 32
//     lanes in the same primitive function (cand_classify<KIND>, rt_scan.cuh); the ray a pair belongs to is read
//     from its owner lane with shuffles;
//   * the result is a classification, not a hit list: an entered non-transmissive surface at t <= sure_t settles
//     srsAttenuation at exactly 0 (the any-hit argument of rt_shadow.cuh: sbuf_add) - the owner's flag word in
//     shared memory is set and the owner stops walking at its next stack pop; any other hit inside the light's
//     distance marks the ray "complex";
//   * at the end: settled -> adds nothing; no hit at all -> the attenuation is exactly (1,1,1) and the Phong term
//     goes to the sample; complex (a transmissive or far-half hit, an origin inside an object, area and spot
//     lights) -> listed for k_wave_shadow_list, the complete single-thread walk with its sorted window.
// Both shortcuts are exits the complete walk takes itself, so the frame is the same bit for bit.
// A ray keeps walking while its candidates wait in a queue (the any-hit exit is a few nodes late); with 32 lanes
// pushing, a queue fills every few leaf visits.
#pragma once

namespace rtd {

#define RT_WB_CAP 128 // entries per queue: a flush leaves < 32, one leaf visit of 32 lanes adds <= 96

// per warp
struct WarpBatch {
	unsigned ent[CQ_KINDS][RT_WB_CAP]; // object index | owner lane << 27
	unsigned pos[RT_WB_CAP];           // triangles: global leaf item position
	unsigned flags[32];                // per owner lane: SH_DEAD | SH_OTHER
	int count[CQ_KINDS];
	int pad_[2];
};

// this lane's push of one candidate; lanes pushing the same kind together share one shared-memory atomic
__device__ __forceinline__ void wb_push(WarpBatch& Wb, int kind, unsigned obj, unsigned pos)
{
	const unsigned lane = threadIdx.x & 31;
	const unsigned grp = __match_any_sync(__activemask(), kind);
	const int leader = __ffs(grp) - 1;
	int base = 0;
	if ((int)lane == leader) base = atomicAdd(&Wb.count[kind], __popc(grp));
	base = __shfl_sync(grp, base, leader);
	const int at = base + __popc(grp & ((1u << lane) - 1u));
	Wb.ent[kind][at] = obj | (lane << 27);
	if (kind == RT_OBJ_TRIMESH) Wb.pos[at] = pos;
}

// The warp, converged, tests up to 32 queued candidates of KIND: lane j takes entry count - nb + j.
// o, d, sd: THIS lane's ray (the owner of an entry is read with shuffles).
template <int KIND>
__device__ __noinline__ void wb_flush(const DScene& S, WarpBatch& Wb, const d3& o, const d3& d, const ShadowSide& sd)
{
	const unsigned lane = threadIdx.x & 31;
	const int c = Wb.count[KIND];
	const int nb = c < 32 ? c : 32;
	const bool mine = (int)lane < nb;
	unsigned e = 0, pos = 0;
	if (mine) {
		e = Wb.ent[KIND][c - nb + lane];
		if (KIND == RT_OBJ_TRIMESH) pos = Wb.pos[c - nb + lane];
	}
	const int owner = mine ? (int)(e >> 27) : (int)lane;
	d3 ro, rd;
	ShadowSide rs;
	ro.x = __shfl_sync(0xffffffffu, o.x, owner); ro.y = __shfl_sync(0xffffffffu, o.y, owner); ro.z = __shfl_sync(0xffffffffu, o.z, owner);
	rd.x = __shfl_sync(0xffffffffu, d.x, owner); rd.y = __shfl_sync(0xffffffffu, d.y, owner); rd.z = __shfl_sync(0xffffffffu, d.z, owner);
	rs.limit = __shfl_sync(0xffffffffu, sd.limit, owner);
	rs.sure_t = __shfl_sync(0xffffffffu, sd.sure_t, owner);
	rs.y[0] = __shfl_sync(0xffffffffu, sd.y[0], owner); rs.y[1] = __shfl_sync(0xffffffffu, sd.y[1], owner); rs.y[2] = __shfl_sync(0xffffffffu, sd.y[2], owner);
	if (mine && !(Wb.flags[owner] & SH_DEAD)) {
		const unsigned fl = cand_classify<KIND>(S, ro, rd, rs, e & 0x7ffffffu, pos);
		if (fl) atomicOr(&Wb.flags[owner], fl);
	}
	__syncwarp();
	if (lane == 0) Wb.count[KIND] = c - nb;
	__syncwarp();
}

#ifndef RT_WB_MINB
#define RT_WB_MINB 16
#endif
// kind_mask: bit k set = the scene has objects of rt_object_type k (a queue that can never fill is never looked at)
__global__ void __launch_bounds__(RT_TRACE_THREADS, RT_WB_MINB) k_wave_shadow_wb(const __grid_constant__ DScene S, const __grid_constant__ PoolArgs A, const RayQueue SQ, int parity,
                                                                                 const WaveCounters* cnt, const __grid_constant__ CandQueues Q, unsigned kind_mask)
{
	__shared__ WarpBatch wbs[RT_TRACE_THREADS / 32];
	const unsigned lane = threadIdx.x & 31;
	const long long k = (long long)blockIdx.x * RT_TRACE_THREADS + threadIdx.x;
	long long n = (long long)cnt->n_shadow[parity];
	if (n > SQ.cap) n = SQ.cap;
	if (k - lane >= n) return; // the whole warp lies beyond the round's shadow rays
	WarpBatch& Wb = wbs[threadIdx.x >> 5];
	if (lane < CQ_KINDS) Wb.count[lane] = 0;
	Wb.flags[lane] = 0;
	__syncwarp();
	const bool valid = k < n;
	Walk W;
	W.ow = mk3(0.0, 0.0, 0.0);
	W.dw = mk3(0.0, 0.0, 1.0);
	W.I.y = mk3(0.0, 0.0, 0.0);
	W.C = nullptr;
	W.B = nullptr;
	W.mode = TRAV_SCAN;
	W.prune = true;
	unsigned f = 0;
	ShadowSide sd;
	sd.limit = RT_T_INF;
	sd.sure_t = RT_T_INF;
	StackEnt stk[RT_STACK];
	RayF F;
	float limf = 0.0f;
	const DNode2f* nodesf = S.snodes2f;
	int level = 0, sp = 0;
	int cur = RT_REF_NONE;
	if (valid) {
		const rt_light& Lt = S.lights[SQ.i[SQ.cap + k]];
		W.ow = SQ.load_v(0, k);
		W.dw = SQ.load_v(3, k);
		// shadow_begin's limit and any-hit bound (rt_traverse.cuh); area and spot lights take the complete walk
		if (Lt.type == RT_LIGHT_POINT) {
			const double D = length(ld3(Lt.position) - W.ow);
			sd.limit = D * (1.0 + 1e-9) + 1e-9;
			sd.sure_t = 0.49 * D;
		} else if (Lt.type != RT_LIGHT_DIRECTIONAL) f = SH_FORCE;
		W.lim0 = sd.limit;
		if (!f) cur = walk_begin(S, W, F, limf);
	}
	sd.y[0] = W.I.y.x; sd.y[1] = W.I.y.y; sd.y[2] = W.I.y.z;
	bool active = cur != RT_REF_NONE;
	for (;;) {
		if (active) {
			// ---- internal nodes: float test of both children, near first ---------------------------------
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
			// ---- a leaf: its candidates go to the warp's queues -------------------------------------------
			if (cur != RT_REF_NONE) {
				const int first = leaf_first(cur), count = leaf_count(cur);
				cur = RT_REF_NONE;
				if (level == 1) {
#pragma unroll 1
					for (int i = 0; i < count; i++) wb_push(Wb, RT_OBJ_TRIMESH, (unsigned)W.V.obj, (unsigned)(W.V.item_off + first + i));
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
						wb_push(Wb, kind, (unsigned)ib.obj, 0u);
					}
				}
			}
			// ---- pop (the one way out of the walk) --------------------------------------------------------
			if (cur == RT_REF_NONE) {
				if (((volatile unsigned*)Wb.flags)[lane] & SH_DEAD) sp = 0; // settled by a tested candidate: any-hit exit
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
				if (cur == RT_REF_NONE) active = false;
			}
		}
		__syncwarp();
		const bool any = __any_sync(0xffffffffu, active);
		// ---- full queues (all of them once every lane has finished), tested by the whole warp -------------
#define RT_WB_FLUSH(KIND)                                                                   \
		if (kind_mask & (1u << KIND))                                                           \
			for (;;) {                                                                          \
				const int c = ((volatile int*)Wb.count)[KIND];                                  \
				if (!(c >= 32 || (!any && c > 0))) break;                                       \
				wb_flush<KIND>(S, Wb, W.ow, W.dw, sd);                                          \
			}
		RT_WB_FLUSH(RT_OBJ_SPHERE)
		RT_WB_FLUSH(RT_OBJ_BOX)
		RT_WB_FLUSH(RT_OBJ_SQUARE)
		RT_WB_FLUSH(RT_OBJ_CYLINDER)
		RT_WB_FLUSH(RT_OBJ_CONE)
		RT_WB_FLUSH(RT_OBJ_TRIMESH)
#undef RT_WB_FLUSH
		if (!any) break;
	}
	f |= Wb.flags[lane];
	const bool settled = (f & SH_DEAD) != 0;                          // srsAttenuation returns dvec3(0): nothing to add
	const bool complex_ray = valid && !settled && (f & (SH_OTHER | SH_FORCE)) != 0;
	if (valid && !settled && !complex_ray) accumulate(A.accum, SQ.i[k], SQ.load_v(6, k)); // no hit: attenuation exactly (1,1,1)
	const long long s = wave_reserve(complex_ray, (1u << lane) - 1u, Q.n_complex);
	if (s >= 0) Q.complex_list[s] = (int)k;
}

} // namespace rtd
