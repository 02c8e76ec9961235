// hostsim.cpp - TEST INFRASTRUCTURE: compiles the device path's per-ray headers
// (cs378hgraphics_raytracer_b200/csrc/device/*.cuh) as plain C++ and drives them with a simple
// CPU loop, so the parity-critical arithmetic (slab test, primitives, ordered traversal with
// pruning, shadow collection passes, Phong, reflection/refraction spawning) can be checked against
// the reference oracle in the CPU-only container.  It is NOT part of the product: librt_b200.so
// contains no host execution path, and nothing outside tests/ loads this library.
#include <cstring>
#include <vector>
#include <omp.h>
#define RT_COUNT_WORK 1
#include "rt_shade.cuh"
#include "rt_derive.hpp"
#include "rt_accum.cuh"

using namespace rtd;

namespace {
struct Sim {
	DScene S;
	Derived D;
};

void setup(Sim& sim, const void* arena)
{
	rt::ArenaView A(arena);
	derive(A, sim.D);
	DScene& S = sim.S;
	S.snodes2 = sim.D.snodes2.data();
	S.stree = sim.D.stree;
	S.mnodes2 = sim.D.mnodes2.data();
	S.snodes2f = sim.D.snodes2f.data();
	S.mnodes2f = sim.D.mnodes2f.data();
	S.sboxf = sim.D.sboxf.data();
	S.sprimf = sim.D.sprimf.data();
	S.mtrees = sim.D.mtrees.data();
	S.snodes = A.ptr<rt_node>(RT_SEC_SCENE_NODES);
	S.sitems = A.ptr<int32_t>(RT_SEC_SCENE_ITEMS);
	S.objs = A.ptr<rt_object>(RT_SEC_OBJECTS);
	S.meshes = A.ptr<rt_mesh>(RT_SEC_MESHES);
	S.mesh_slack = sim.D.mesh_slack.data();
	S.mnodes = A.ptr<rt_node>(RT_SEC_MESH_NODES);
	S.mitems = A.ptr<int32_t>(RT_SEC_MESH_ITEMS);
	S.tris = sim.D.tris.data();
	S.faces = A.ptr<rt_face>(RT_SEC_FACES);
	S.verts = A.ptr<double>(RT_SEC_VERTS);
	S.vnormals = A.ptr<double>(RT_SEC_VNORMALS);
	S.vmats = A.ptr<int32_t>(RT_SEC_VMATS);
	S.mats = A.ptr<rt_material>(RT_SEC_MATERIALS);
	S.lights = A.ptr<rt_light>(RT_SEC_LIGHTS);
	S.texs = A.ptr<rt_texture>(RT_SEC_TEXTURES);
	S.texdata = A.ptr<uint8_t>(RT_SEC_TEXDATA);
	S.cam = *A.ptr<rt_camera>(RT_SEC_CAMERA);
	S.glob = *A.ptr<rt_globals>(RT_SEC_GLOBALS);
	S.scene_slack = sim.D.scene_slack;
	S.n_lights = (int)A.count(RT_SEC_LIGHTS);
	S.n_objs = (int)A.count(RT_SEC_OBJECTS);
	S.kt_unit = kt_all_unit(A.ptr<rt_material>(RT_SEC_MATERIALS), A.count(RT_SEC_MATERIALS));
	S.pad_ = 0;
}

template <bool PRUNE> d3 trace_tree(const DScene& S, const rt_params& prm, const RayTask& root, rt_stats* st)
{
	std::vector<RayTask> stack;
	stack.push_back(root);
	d3 sum = mk3(0, 0, 0);
	while (!stack.empty()) {
		RayTask r = stack.back();
		stack.pop_back();
		Hit h;
		RayTask kids[2];
		int nk;
		unsigned sh = 0;
		sum = sum + trace_step<PRUNE>(S, prm, r, h, kids, nk, sh);
		if (st) {
			st->shadow_rays += sh;
			if (r.kind == 0) st->camera_rays++;
			else if (r.kind == 1) st->reflection_rays++;
			else st->refraction_rays++;
		}
		for (int k = nk - 1; k >= 0; k--) stack.push_back(kids[k]);
	}
	return sum;
}

RayTask make_root(const d3& o, const d3& d, int depth)
{
	RayTask r;
	r.o = o;
	r.d = d;
	r.weight = mk3(1, 1, 1);
	r.att_kt = mk3(1, 1, 1);
	r.att_mode = 0;
	r.depth = depth;
	r.kind = 0;
	return r;
}

// RayTracer::trace(x,y) (reference RayTracer.cpp:35-79)
template <bool PRUNE> d3 sample_colour(const DScene& S, const rt_params& prm, double x, double y, rt_stats* st)
{
	const int lenses = prm.dof_on ? prm.dof_divs + 1 : 1;
	d3 ret = mk3(0, 0, 0);
	for (int l = 0; l < lenses; l++) {
		d3 o, d;
		primary_ray(S.cam, prm, x, y, l, 0.0, o, d);
		ret = ret + trace_tree<PRUNE>(S, prm, make_root(o, d, prm.depth), st);
	}
	if (prm.dof_on) ret = ret * (1.0 / (prm.dof_divs + 1.0));
	return gclamp(ret, 0.0, 1.0);
}
} // namespace

extern "C" {

int sim_trace_rays(const void* arena, const rt_params* prm, const rt_ray* rays, int n, int depth, rt_hit* hits, double* rgb, int prune)
{
	Sim sim;
	setup(sim, arena);
#pragma omp parallel for schedule(dynamic, 64)
	for (int k = 0; k < n; k++) {
		const d3 o = ld3(rays[k].p), d = ld3(rays[k].d);
		if (hits) {
			Hit h;
			bool hit = prune ? closest_hit<true>(sim.S, o, d, h) : closest_hit<false>(sim.S, o, d, h);
			rt_hit out;
			out.obj = out.face = -1;
			out.t = h.t;
			out.N[0] = out.N[1] = out.N[2] = 0.0;
			if (hit) {
				finalize_hit(sim.S, o, d, h);
				out.obj = h.obj;
				out.face = h.face;
				out.N[0] = h.N.x; out.N[1] = h.N.y; out.N[2] = h.N.z;
			}
			hits[k] = out;
		}
		if (rgb) {
			RayTask r = make_root(o, d, depth);
			d3 c = prune ? trace_tree<true>(sim.S, *prm, r, nullptr) : trace_tree<false>(sim.S, *prm, r, nullptr);
			rgb[3 * k] = c.x; rgb[3 * k + 1] = c.y; rgb[3 * k + 2] = c.z;
		}
	}
	return 0;
}

// No-AA / supersampled frame (RayTracer::traceImage's first two branches, RayTracer.cpp:279-314)
int sim_render(const void* arena, const rt_params* prm, int w, int h, uint8_t* rgb_out, rt_stats* stats)
{
	Sim sim;
	setup(sim, arena);
	const int s = (prm->aa_mode == RT_AA_SUPERSAMPLE || prm->aa_mode == RT_AA_JITTERED) ? prm->aa_samples : 1;
	if (prm->aa_mode == RT_AA_ADAPTIVE) return 1;
	std::vector<rt_stats> st(omp_get_max_threads());
	for (auto& x : st) std::memset(&x, 0, sizeof(x));
#pragma omp parallel for collapse(2) schedule(dynamic, 16)
	for (int j = 0; j < h; j++)
		for (int i = 0; i < w; i++) {
			d3 col = mk3(0, 0, 0);
			for (int a = 0; a < s; a++)
				for (int b = 0; b < s; b++) {
					const double x = double(i * s + a) / (double(w) * s), y = double(j * s + b) / (double(h) * s);
					col = col + sample_colour<true>(sim.S, *prm, x, y, &st[omp_get_thread_num()]);
				}
			if (s > 1) col = mk3(col.x / double(s * s), col.y / double(s * s), col.z / double(s * s));
			uint8_t* px = rgb_out + ((size_t)i + (size_t)j * w) * 3;
			px[0] = (uint8_t)(int)(255.0 * col.x);
			px[1] = (uint8_t)(int)(255.0 * col.y);
			px[2] = (uint8_t)(int)(255.0 * col.z);
		}
	if (stats) {
		std::memset(stats, 0, sizeof(*stats));
		for (auto& x : st) {
			stats->camera_rays += x.camera_rays;
			stats->reflection_rays += x.reflection_rays;
			stats->refraction_rays += x.refraction_rays;
			stats->shadow_rays += x.shadow_rays;
		}
	}
	return 0;
}


// Work profile of one frame (cost model for bench.py's roofline, tail analysis): out[2][4 + 32 + 1] =
// per query kind {walks, nodes, tris, objs, hist[32], max_nodes}
int sim_work(const void* arena, const rt_params* prm, int w, int h, unsigned long long* out)
{
	Sim sim;
	setup(sim, arena);
	const int nt = omp_get_max_threads();
	std::vector<TravWork> all(nt);
	if (const char* e = getenv("RT_WORK_DUMP")) g_work_dump = strtoull(e, nullptr, 10);
#pragma omp parallel
	{
		std::memset(&g_work, 0, sizeof(g_work));
#pragma omp for collapse(2) schedule(dynamic, 16)
		for (int j = 0; j < h; j++)
			for (int i = 0; i < w; i++) {
				const double x = double(i) / double(w), y = double(j) / double(h);
				sample_colour<true>(sim.S, *prm, x, y, nullptr);
			}
		all[omp_get_thread_num()] = g_work;
	}
	std::memset(out, 0, 2 * 37 * sizeof(unsigned long long));
	for (const TravWork& t : all)
		for (int m = 0; m < 2; m++) {
			unsigned long long* o = out + m * 37;
			o[0] += t.walks[m]; o[1] += t.nodes[m]; o[2] += t.tris[m]; o[3] += t.objs[m];
			for (int b = 0; b < 32; b++) o[4 + b] += t.hist[m][b];
			if (t.max_nodes[m] > o[36]) o[36] = t.max_nodes[m];
		}
	return 0;
}

// Per-pixel walk lengths of a depth-0 frame (lane-utilisation analysis, profiles/simt_model.py):
// out[pixel * slots + 0] = nodes of the camera ray's walk, out[pixel * slots + 1 + k] = nodes of its
// k-th shadow walk (k < slots - 1; later ones are added to the last slot)
int sim_walk_lengths(const void* arena, const rt_params* prm, int w, int h, int slots, unsigned* out)
{
	Sim sim;
	setup(sim, arena);
	rt_params p = *prm;
	p.depth = 0;
	std::memset(out, 0, sizeof(unsigned) * (size_t)w * h * slots);
#pragma omp parallel
	{
		std::vector<unsigned> log(2 + 2 * 4096);
#pragma omp for collapse(2) schedule(dynamic, 16)
		for (int j = 0; j < h; j++)
			for (int i = 0; i < w; i++) {
				log[0] = 0;
				log[1] = 4096;
				g_walk_log = log.data();
				const double x = double(i) / double(w), y = double(j) / double(h);
				sample_colour<true>(sim.S, p, x, y, nullptr);
				g_walk_log = nullptr;
				unsigned* o = out + ((size_t)j * w + i) * slots;
				int ks = 0;
				for (unsigned e = 0; e < log[0]; e++) {
					if (log[2 + 2 * e] == 0) o[0] += log[3 + 2 * e];
					else { o[1 + (ks < slots - 2 ? ks : slots - 2)] += log[3 + 2 * e]; ks++; }
				}
			}
	}
	return 0;
}

// rt_accum.cuh on the host: the terms xs[0..n) split into limbs, the limbs added as integers in the given order
// (any order gives the same limbs), the total read back.  limbs_out (optional) = {limb 0, limb 1}.
double sim_acc_sum(const double* xs, int n, long long* limbs_out)
{
	unsigned long long l0 = 0, l1 = 0;
	for (int k = 0; k < n; k++) {
		long long lo, hi;
		acc_split(xs[k], lo, hi);
		l0 += (unsigned long long)lo;
		l1 += (unsigned long long)hi;
	}
	if (limbs_out) { limbs_out[0] = (long long)l0; limbs_out[1] = (long long)l1; }
	return acc_merge((long long)l0, (long long)l1);
}

// acc_read3 as the resolve kernels call it: one sample's red channel from its limbs (the added terms) and the `lenses`
// stored camera-ray terms of its slots
double sim_acc_read(const double* added, int n_added, const double* stored, int lenses)
{
	long long acc[RT_ACC_STRIDE] = {0, 0, 0, 0, 0, 0};
	for (int k = 0; k < n_added; k++) {
		long long lo, hi;
		acc_split(added[k], lo, hi);
		acc[0] = (long long)((unsigned long long)acc[0] + (unsigned long long)lo);
		acc[1] = (long long)((unsigned long long)acc[1] + (unsigned long long)hi);
	}
	std::vector<double> base(3 * (size_t)lenses, 0.0);
	for (int l = 0; l < lenses; l++) base[3 * l] = stored[l];
	return acc_read3(acc, lenses ? base.data() : nullptr, lenses, 0).x;
}

// Soundness of the object-space float pre-test (rt_prims.cuh: prim_miss_f), brute force: every ray against every scene-level
// item.  out[0] = pairs whose box the reference's double test enters, out[1] = of those, pairs the pre-test rejects,
// out[2] = of those, pairs whose double-precision object-space test HITS (must be 0), out[3] = pairs with a pre-test at all
// (kind != 0) among out[0], out[4] = pairs that hit, out[5] = rejected pairs over ALL items (box entered or not) that hit (must be 0).
// aimed != 0: the direction of every (ray, item) pair is replaced by one from the ray's origin through a pseudo-random point of
// the item's box enlarged by a fifth - every pair is then a hit or a near miss (grazing lines, corners, rims).
int sim_primf_check(const void* arena, const double* rays, int n, int aimed, unsigned long long* out)
{
	Sim sim;
	setup(sim, arena);
	const DScene& S = sim.S;
	const long long n_items = (long long)sim.D.sboxf.size();
	unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0, c5 = 0;
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : c0, c1, c2, c3, c4, c5)
	for (int k = 0; k < n; k++) {
		const d3 o = mk3(rays[6 * k], rays[6 * k + 1], rays[6 * k + 2]);
		d3 d = mk3(rays[6 * k + 3], rays[6 * k + 4], rays[6 * k + 5]);
		RayInv I;
		RayF F;
		make_inv(I, F, o, d, 0.0, S.scene_slack * 1e9);
		for (long long r = 0; r < n_items; r++) {
			const DItemBox& ib = S.sboxf[r];
			const rt_object& ob = S.objs[ib.obj];
			if (ob.type == RT_OBJ_TRIMESH) continue;
			if (aimed) {
				unsigned long long hsh = (unsigned long long)k * 0x9E3779B97F4A7C15ull + (unsigned long long)r * 0xC2B2AE3D27D4EB4Full + 12345u;
				double u[3];
				for (int c = 0; c < 3; c++) {
					hsh ^= hsh >> 29; hsh *= 0xBF58476D1CE4E5B9ull; hsh ^= hsh >> 32;
					u[c] = ((double)(hsh >> 11) / 9007199254740992.0 * 2.0 - 1.0) * 1.2;
				}
				const double cx = 0.5 * (ob.bbox[0] + ob.bbox[3]), cy = 0.5 * (ob.bbox[1] + ob.bbox[4]), cz = 0.5 * (ob.bbox[2] + ob.bbox[5]);
				d = mk3(cx + u[0] * 0.5 * (ob.bbox[3] - ob.bbox[0]) - o.x, cy + u[1] * 0.5 * (ob.bbox[4] - ob.bbox[1]) - o.y, cz + u[2] * 0.5 * (ob.bbox[5] - ob.bbox[2]) - o.z);
				if (aimed == 2) d = normalize(d);
				make_inv(I, F, o, d, 0.0, S.scene_slack * 1e9);
			}
			const bool miss = prim_miss_f(S.sprimf[r], ib, o, d);
			LocalRay L;
			to_local(L, ob, o, d);
			TravClosest C;
			C.have = false;
			C.best_t = RT_T_INF;
			C.best_rank = 0x7fffffff;
			C.best_obj = -1;
			C.best_fpos = -1;
			prim_closest(ob, L, (int)r, ib.obj, C);
			if (miss && C.have) c5++;
			double eb;
			if (!box_hit(ob.bbox, ob.bbox + 3, o, d, I, eb)) continue;
			c0++;
			if (S.sprimf[r].kind != RT_PF_NONE) c3++;
			if (C.have) c4++;
			if (miss) { c1++; if (C.have) c2++; }
		}
	}
	out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3; out[4] = c4; out[5] = c5;
	return 0;
}

} // extern "C"
