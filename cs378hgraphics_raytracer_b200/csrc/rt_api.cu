// rt_api.cu - device half of the C-ABI (include/rt_b200.h): context, scene upload, render calls.
//
// One context = one GPU, one stream.  Every call is synchronous on return, like the reference's
// RayTracer (checkRender() == true).  There is no host execution path: without a CUDA device
// rt_create fails with RT_ERR_NO_DEVICE.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "rt_b200.h"
#include "rt_arena_writer.hpp"
#include "device/rt_kernels.cuh"
#include "device/rt_derive.hpp"

using namespace rtd;

struct rt_ctx {
	int device = 0;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	std::string error;
	bool has_scene = false;
	std::vector<void*> scene_allocs;
	DScene S;
	int sm_count = 0;
	int pool_blocks = 0;
	// work buffers (grown on demand, reused across calls)
	void* buf[32] = {nullptr};
	size_t buf_bytes[32] = {0};
	unsigned long long launches = 0;
	// wavefront path
	long long wave_width = 1 << 21;        // rays per round (RT_B200_WAVE overrides)
	int walk_budget = 96;                  // tree nodes a single thread spends on one ray before the warp-cooperative walk takes it (RT_B200_BUDGET; 0 = never)
	bool use_pool = false;                 // RT_B200_POOL=1: the fused persistent kernel instead
	unsigned long long* pinned = nullptr;  // round counters read back by the host
	cudaEvent_t ev_round = nullptr;
	unsigned long long rounds = 0;
};

namespace {

enum { B_ACCUM = 0, B_ACCUM2, B_POOL_D, B_POOL_I, B_SMALL, B_FB, B_XS, B_YS, B_REG_A, B_REG_B, B_PIXACC, B_RAYS, B_HITS,
       B_WQ0_D, B_WQ0_I, B_WQ1_D, B_WQ1_I, B_WSP_D, B_WSP_I, B_SQ_D, B_SQ_I, B_WH_T, B_WH_OF, B_HEAVY, B_COUNT };
static_assert(B_COUNT <= 32, "rt_ctx::buf too small");

#define CK(call)                                                                                          \
	do {                                                                                                  \
		cudaError_t e_ = (call);                                                                          \
		if (e_ != cudaSuccess) {                                                                          \
			ctx->error = std::string(#call) + ": " + cudaGetErrorString(e_);                             \
			return RT_ERR_CUDA;                                                                           \
		}                                                                                                 \
	} while (0)

rt_status ensure(rt_ctx* ctx, int which, size_t bytes)
{
	if (bytes == 0) bytes = 16;
	if (ctx->buf_bytes[which] >= bytes) return RT_OK;
	if (ctx->buf[which]) CK(cudaFree(ctx->buf[which]));
	ctx->buf[which] = nullptr;
	ctx->buf_bytes[which] = 0;
	CK(cudaMalloc(&ctx->buf[which], bytes));
	ctx->buf_bytes[which] = bytes;
	return RT_OK;
}

template <typename T> rt_status upload(rt_ctx* ctx, const T* host, size_t count, const T** dev)
{
	void* p = nullptr;
	const size_t bytes = count * sizeof(T);
	CK(cudaMalloc(&p, bytes ? bytes : 16));
	ctx->scene_allocs.push_back(p);
	if (bytes) CK(cudaMemcpyAsync(p, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
	*dev = (const T*)p;
	return RT_OK;
}

void free_scene(rt_ctx* ctx)
{
	for (void* p : ctx->scene_allocs) cudaFree(p);
	ctx->scene_allocs.clear();
	ctx->has_scene = false;
}

rt_status check_params(rt_ctx* ctx, const rt_params* p)
{
	if (!p) { ctx->error = "null params"; return RT_ERR_ARG; }
	if (p->depth < 0 || p->depth > 200) { ctx->error = "depth out of range (0..200)"; return RT_ERR_ARG; }
	if (p->aa_mode < 0 || p->aa_mode > 3) { ctx->error = "bad aa_mode"; return RT_ERR_ARG; }
	if (p->aa_mode != RT_AA_NONE && (p->aa_samples < 1 || p->aa_samples > 64)) { ctx->error = "aa_samples out of range (1..64)"; return RT_ERR_ARG; }
	if (p->dof_on && (p->dof_divs < 0 || p->dof_divs > 1024)) { ctx->error = "dof_divs out of range"; return RT_ERR_ARG; }
	if (p->overlapping_objects) { ctx->error = "overlapping-objects mode (Scene::discoverMat) is not implemented on the device path yet"; return RT_ERR_UNSUPPORTED; }
	return RT_OK;
}

static inline unsigned blocks_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

rt_status make_queue(rt_ctx* ctx, int bd, int bi, long long cap, int nd, RayQueue& Q)
{
	rt_status st;
	if ((st = ensure(ctx, bd, (size_t)cap * nd * sizeof(double))) != RT_OK) return st;
	if ((st = ensure(ctx, bi, (size_t)cap * 2 * sizeof(int))) != RT_OK) return st;
	Q.d = (double*)ctx->buf[bd];
	Q.i = (int*)ctx->buf[bi];
	Q.cap = cap;
	Q.nd = nd;
	return RT_OK;
}

rt_status launch_pool(rt_ctx* ctx, PoolArgs& A);

// The trace loop over n_slots primary-ray slots as a streaming wavefront (device/rt_wave.cuh).
rt_status launch_wave(rt_ctx* ctx, PoolArgs& A)
{
	if (ctx->use_pool || A.prm.aterm_thresh > 0.0) return launch_pool(ctx, A); // adaptive termination needs the shadowed colour before it recurses
	if (A.n_slots <= 0) return RT_OK;
	const int L = ctx->S.n_lights;
	long long W = ctx->wave_width;
	const long long fit = ((A.n_slots + 1023) / 1024) * 1024;
	if (W > fit) W = fit;
	rt_status st;
	RayQueue cur, nxt, sq, sp;
	sp.d = nullptr; sp.i = nullptr; sp.cap = 0; sp.nd = 12;
	if ((st = make_queue(ctx, B_WQ0_D, B_WQ0_I, 2 * W, 12, cur)) != RT_OK) return st;
	if ((st = make_queue(ctx, B_WQ1_D, B_WQ1_I, 2 * W, 12, nxt)) != RT_OK) return st;
	if ((st = make_queue(ctx, B_SQ_D, B_SQ_I, W * (L > 0 ? L : 1), 9, sq)) != RT_OK) return st;
	if ((st = ensure(ctx, B_WH_T, (size_t)W * sizeof(double))) != RT_OK) return st;
	if ((st = ensure(ctx, B_WH_OF, (size_t)W * sizeof(int2))) != RT_OK) return st;
	if ((st = ensure(ctx, B_HEAVY, (size_t)W * sizeof(int))) != RT_OK) return st;
	int* heavy = (int*)ctx->buf[B_HEAVY];
	HitQueue H;
	H.t = (double*)ctx->buf[B_WH_T];
	H.of = (int2*)ctx->buf[B_WH_OF];
	WaveCounters* cnt = (WaveCounters*)ctx->buf[B_SMALL];
	A.stats = cnt->stats;
	long long cur_n = 0, sp_n = 0, cam = 0;
	for (;;) {
		long long base = 0, m = 0;
		if (cur_n > W) {
			// more children than one round takes: the oldest go to the spill stack
			const long long spill = cur_n - W;
			if (!sp.d) {
				const long long cap = W * ((A.depth < 30 ? A.depth : 30) + 2);
				if ((st = make_queue(ctx, B_WSP_D, B_WSP_I, cap, 12, sp)) != RT_OK) return st;
			}
			if (sp_n + spill > sp.cap) { ctx->error = "wavefront spill stack overflow"; return RT_ERR_OVERFLOW; }
			k_wave_move<<<blocks_for(spill, 256), 256, 0, ctx->stream>>>(cur, 0, sp, sp_n, spill);
			ctx->launches++;
			sp_n += spill;
			base = spill;
			m = W;
		} else {
			long long need = W - cur_n;
			const long long un = need < sp_n ? need : sp_n;
			if (un > 0) {
				k_wave_move<<<blocks_for(un, 256), 256, 0, ctx->stream>>>(sp, sp_n - un, cur, cur_n, un);
				ctx->launches++;
				sp_n -= un;
				cur_n += un;
				need -= un;
			}
			const long long gen = need < A.n_slots - cam ? need : A.n_slots - cam;
			if (gen > 0) {
				k_wave_gen<<<blocks_for(gen, RT_WAVE_THREADS), RT_WAVE_THREADS, 0, ctx->stream>>>(ctx->S, A, cur, cur_n, cam, gen);
				ctx->launches++;
				cur_n += gen;
				cam += gen;
			}
			m = cur_n;
		}
		if (m == 0) break;
		CK(cudaMemsetAsync(&cnt->n_children, 0, 4 * sizeof(unsigned long long), ctx->stream));
		k_wave_closest<<<blocks_for(m, RT_TRACE_THREADS), RT_TRACE_THREADS, 0, ctx->stream>>>(ctx->S, cur, base, m, H, ctx->walk_budget, heavy, cnt);
		if (ctx->walk_budget > 0) {
			k_wave_closest_heavy<<<ctx->sm_count * 4, 32 * RT_COOP_WARPS, 0, ctx->stream>>>(ctx->S, cur, base, H, heavy, cnt);
			ctx->launches++;
		}
		k_wave_shade<<<blocks_for(m, RT_WAVE_THREADS), RT_WAVE_THREADS, 0, ctx->stream>>>(ctx->S, A, cur, base, m, H, nxt, sq, cnt);
		CK(cudaMemcpyAsync(ctx->pinned, &cnt->n_children, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaEventRecord(ctx->ev_round, ctx->stream));
		ctx->launches += 2;
		if (L > 0) {
			// grid sized for the most shadow rays this round can have made; blocks past the count leave at once
			k_wave_shadow<<<blocks_for(m * L, RT_TRACE_THREADS), RT_TRACE_THREADS, 0, ctx->stream>>>(ctx->S, A, sq, cnt);
			ctx->launches++;
		}
		CK(cudaGetLastError());
		CK(cudaEventSynchronize(ctx->ev_round)); // the host plans the next round while the shadow rays run
		cur_n = (long long)ctx->pinned[0];
		if (cur_n > nxt.cap) cur_n = nxt.cap; // overflow was flagged by the kernel
		const RayQueue tmp = cur;
		cur = nxt;
		nxt = tmp;
		ctx->rounds++;
	}
	return RT_OK;
}

// One launch of the fused persistent kernel over n_slots primary-ray slots.
rt_status launch_pool(rt_ctx* ctx, PoolArgs& A)
{
	const int warps_per_block = RT_POOL_THREADS / 32;
	const long long warps = (long long)ctx->pool_blocks * warps_per_block;
	const bool defer = !(A.prm.aterm_thresh > 0.0);
	A.cap = 32 * (A.depth + 3) + (defer ? 64 * ctx->S.n_lights + 64 : 0);
	rt_status st;
	if ((st = ensure(ctx, B_POOL_D, (size_t)warps * A.cap * RT_POOL_DOUBLES * sizeof(double))) != RT_OK) return st;
	if ((st = ensure(ctx, B_POOL_I, (size_t)warps * A.cap * 2 * sizeof(int))) != RT_OK) return st;
	A.pool_d = (double*)ctx->buf[B_POOL_D];
	A.pool_i = (int*)ctx->buf[B_POOL_I];
	unsigned long long* small = (unsigned long long*)ctx->buf[B_SMALL];
	A.cursor = small + 8;
	A.stats = small;
	CK(cudaMemsetAsync(A.cursor, 0, sizeof(unsigned long long), ctx->stream));
	if (A.n_slots > 0) {
		if (defer) k_trace_pool<true><<<ctx->pool_blocks, RT_POOL_THREADS, 0, ctx->stream>>>(ctx->S, A);
		else k_trace_pool<false><<<ctx->pool_blocks, RT_POOL_THREADS, 0, ctx->stream>>>(ctx->S, A);
		ctx->launches++;
		CK(cudaGetLastError());
	}
	return RT_OK;
}

rt_status read_stats(rt_ctx* ctx, rt_stats* stats, float ms)
{
	unsigned long long h[8];
	CK(cudaMemcpyAsync(h, ctx->buf[B_SMALL], sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	if (stats) {
		stats->camera_rays = h[0];
		stats->reflection_rays = h[1];
		stats->refraction_rays = h[2];
		stats->shadow_rays = h[3];
		stats->device_ms = ms;
		stats->launches = ctx->launches;
		stats->cooperative_rays = h[5];
		stats->rounds = ctx->rounds;
	}
	if (h[4]) { ctx->error = "device ray pool overflow"; return RT_ERR_OVERFLOW; }
	return RT_OK;
}

} // namespace

extern "C" {

rt_status rt_create(int device, rt_ctx** out)
{
	if (!out) return RT_ERR_ARG;
	*out = nullptr;
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return RT_ERR_NO_DEVICE;
	rt_ctx* ctx = new rt_ctx();
	ctx->device = device;
	if (cudaSetDevice(device) != cudaSuccess) { delete ctx; return RT_ERR_NO_DEVICE; }
	cudaDeviceProp prop;
	if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return RT_ERR_NO_DEVICE; }
	ctx->sm_count = prop.multiProcessorCount;
	if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return RT_ERR_CUDA; }
	cudaEventCreate(&ctx->ev0);
	cudaEventCreate(&ctx->ev1);
	cudaEventCreateWithFlags(&ctx->ev_round, cudaEventDisableTiming);
	if (cudaMallocHost((void**)&ctx->pinned, 64) != cudaSuccess) { delete ctx; return RT_ERR_CUDA; }
	if (const char* e = getenv("RT_B200_WAVE")) { const long long v = atoll(e); if (v >= 1024) ctx->wave_width = (v + 1023) / 1024 * 1024; }
	if (const char* e = getenv("RT_B200_POOL")) ctx->use_pool = atoi(e) != 0;
	if (const char* e = getenv("RT_B200_BUDGET")) { const int v = atoi(e); if (v >= 0) ctx->walk_budget = v; }
	// persistent grid: resident CTAs per SM x number of SMs
	int per_sm = 0;
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace_pool<true>, RT_POOL_THREADS, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
	ctx->pool_blocks = per_sm * ctx->sm_count;
	if (cudaMalloc(&ctx->buf[B_SMALL], 256) != cudaSuccess) { delete ctx; return RT_ERR_CUDA; }
	ctx->buf_bytes[B_SMALL] = 256;
	// deep per-thread stacks (two 64-entry traversal stacks + shadow buffer) live in local memory
	cudaDeviceSetLimit(cudaLimitStackSize, 8192);
	*out = ctx;
	return RT_OK;
}

void rt_destroy(rt_ctx* ctx)
{
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	free_scene(ctx);
	for (int i = 0; i < 32; i++)
		if (ctx->buf[i]) cudaFree(ctx->buf[i]);
	if (ctx->pinned) cudaFreeHost(ctx->pinned);
	if (ctx->ev_round) cudaEventDestroy(ctx->ev_round);
	if (ctx->ev0) cudaEventDestroy(ctx->ev0);
	if (ctx->ev1) cudaEventDestroy(ctx->ev1);
	if (ctx->stream) cudaStreamDestroy(ctx->stream);
	delete ctx;
}

const char* rt_last_error(const rt_ctx* ctx) { return ctx ? ctx->error.c_str() : "null context"; }

rt_status rt_upload_scene(rt_ctx* ctx, const void* arena, size_t bytes)
{
	if (!ctx) return RT_ERR_ARG;
	if (rt_arena_validate(arena, bytes) != RT_OK) { ctx->error = "invalid arena"; return RT_ERR_ARG; }
	CK(cudaSetDevice(ctx->device));
	rt::ArenaView A(arena);
	const rt_light* lights = A.ptr<rt_light>(RT_SEC_LIGHTS);
	for (uint64_t k = 0; k < A.count(RT_SEC_LIGHTS); k++)
		if (lights[k].type != RT_LIGHT_DIRECTIONAL && lights[k].type != RT_LIGHT_POINT) {
			ctx->error = "area and spot lights are not implemented on the device path yet";
			return RT_ERR_UNSUPPORTED;
		}
	Derived D;
	derive(A, D);
	if (D.max_scene_depth + D.max_mesh_depth > RT_STACK - 4 || A.count(RT_SEC_MESH_ITEMS) >= (1ull << 29) || A.count(RT_SEC_SCENE_ITEMS) >= (1ull << 29)) { ctx->error = "trees too deep for the traversal stack"; return RT_ERR_UNSUPPORTED; }
	free_scene(ctx);
	DScene& S = ctx->S;
	std::memset(&S, 0, sizeof(S));
	rt_status st;
#define UP(field, type, sec)                                                                        \
	if ((st = upload<type>(ctx, A.ptr<type>(sec), A.count(sec), &S.field)) != RT_OK) { free_scene(ctx); return st; }
	if ((st = upload<DNode2>(ctx, D.snodes2.data(), D.snodes2.size(), &S.snodes2)) != RT_OK) { free_scene(ctx); return st; }
	if ((st = upload<DNode2>(ctx, D.mnodes2.data(), D.mnodes2.size(), &S.mnodes2)) != RT_OK) { free_scene(ctx); return st; }
	if ((st = upload<DTree>(ctx, D.mtrees.data(), D.mtrees.size(), &S.mtrees)) != RT_OK) { free_scene(ctx); return st; }
	UP(snodes, rt_node, RT_SEC_SCENE_NODES)
	UP(sitems, int32_t, RT_SEC_SCENE_ITEMS)
	UP(objs, rt_object, RT_SEC_OBJECTS)
	UP(meshes, rt_mesh, RT_SEC_MESHES)
	UP(mnodes, rt_node, RT_SEC_MESH_NODES)
	UP(mitems, int32_t, RT_SEC_MESH_ITEMS)
	UP(faces, rt_face, RT_SEC_FACES)
	UP(mats, rt_material, RT_SEC_MATERIALS)
	UP(lights, rt_light, RT_SEC_LIGHTS)
	UP(texs, rt_texture, RT_SEC_TEXTURES)
	UP(texdata, uint8_t, RT_SEC_TEXDATA)
	UP(vmats, int32_t, RT_SEC_VMATS)
#undef UP
	if ((st = upload<double>(ctx, A.ptr<double>(RT_SEC_VERTS), A.count(RT_SEC_VERTS) * 3, &S.verts)) != RT_OK) { free_scene(ctx); return st; }
	if ((st = upload<double>(ctx, A.ptr<double>(RT_SEC_VNORMALS), A.count(RT_SEC_VNORMALS) * 3, &S.vnormals)) != RT_OK) { free_scene(ctx); return st; }
	if ((st = upload<DTri>(ctx, D.tris.data(), D.tris.size(), &S.tris)) != RT_OK) { free_scene(ctx); return st; }
	if ((st = upload<double>(ctx, D.mesh_slack.data(), D.mesh_slack.size(), &S.mesh_slack)) != RT_OK) { free_scene(ctx); return st; }
	S.cam = *A.ptr<rt_camera>(RT_SEC_CAMERA);
	S.glob = *A.ptr<rt_globals>(RT_SEC_GLOBALS);
	S.scene_slack = D.scene_slack;
	S.stree = D.stree;
	S.n_lights = (int)A.count(RT_SEC_LIGHTS);
	S.n_objs = (int)A.count(RT_SEC_OBJECTS);
	CK(cudaStreamSynchronize(ctx->stream)); // D's host vectors die with this scope
	ctx->has_scene = true;
	return RT_OK;
}

rt_status rt_render_device(rt_ctx* ctx, const rt_params* p, int w, int h, int part, int n_parts, void** dev_rgb, rt_stats* stats)
{
	if (!ctx) return RT_ERR_ARG;
	rt_status st = check_params(ctx, p);
	if (st != RT_OK) return st;
	if (!ctx->has_scene) { ctx->error = "no scene uploaded"; return RT_ERR_NO_SCENE; }
	if (w <= 0 || h <= 0 || n_parts <= 0 || part < 0 || part >= n_parts || !dev_rgb) { ctx->error = "bad frame arguments"; return RT_ERR_ARG; }
	CK(cudaSetDevice(ctx->device));
	ctx->launches = 0;
	ctx->rounds = 0;
	PoolArgs A;
	std::memset(&A, 0, sizeof(A));
	A.prm = *p;
	A.w = w;
	A.h = h;
	A.part = part;
	A.n_parts = n_parts;
	A.tiles_x = (w + RT_TILE - 1) / RT_TILE;
	A.tiles_y = (h + RT_TILE - 1) / RT_TILE;
	A.depth = p->depth;
	A.lenses = p->dof_on ? p->dof_divs + 1 : 1;
	const long long n_tiles = (long long)A.tiles_x * A.tiles_y;
	const long long local_tiles = (n_tiles - part + n_parts - 1) / n_parts;
	const long long n_pix = local_tiles * RT_TILE_PIX;
	if ((st = ensure(ctx, B_FB, (size_t)w * h * 3)) != RT_OK) return st;
	uint8_t* fb = (uint8_t*)ctx->buf[B_FB];
	CK(cudaMemsetAsync(fb, 0, (size_t)w * h * 3, ctx->stream)); // traceSetup zero-fills (RayTracer.cpp:251)
	CK(cudaMemsetAsync(ctx->buf[B_SMALL], 0, 64, ctx->stream));
	CK(cudaEventRecord(ctx->ev0, ctx->stream));
	const bool adaptive = p->aa_mode == RT_AA_ADAPTIVE;
	if (!adaptive) {
		A.src = SRC_GRID;
		A.spp_side = (p->aa_mode == RT_AA_SUPERSAMPLE || p->aa_mode == RT_AA_JITTERED) ? p->aa_samples : 1;
		const long long n_samples = n_pix * A.spp_side * A.spp_side;
		if (n_samples > 0x7fffffffll) { ctx->error = "too many samples for one launch"; return RT_ERR_UNSUPPORTED; }
		A.n_slots = n_samples * A.lenses;
		if ((st = ensure(ctx, B_ACCUM, (size_t)n_samples * 24)) != RT_OK) return st;
		A.accum = (double*)ctx->buf[B_ACCUM];
		CK(cudaMemsetAsync(A.accum, 0, (size_t)n_samples * 24, ctx->stream));
		if ((st = launch_wave(ctx, A)) != RT_OK) return st;
		double* accum2 = nullptr;
		if (p->anaglyph) { // second eye, shifted by ANAGLYPH_DELTA (RayTracer.cpp:23,92-99)
			if ((st = ensure(ctx, B_ACCUM2, (size_t)n_samples * 24)) != RT_OK) return st;
			accum2 = (double*)ctx->buf[B_ACCUM2];
			CK(cudaMemsetAsync(accum2, 0, (size_t)n_samples * 24, ctx->stream));
			PoolArgs A2 = A;
			A2.accum = accum2;
			A2.eye_shift = 0.25;
			if ((st = launch_wave(ctx, A2)) != RT_OK) return st;
		}
		if (n_pix > 0) {
			k_resolve_grid<<<blocks_for(n_pix, 256), 256, 0, ctx->stream>>>(A, accum2, n_pix, fb);
			ctx->launches++;
			CK(cudaGetLastError());
		}
	} else {
		A.spp_side = 1;
		const int N = p->aa_samples * p->aa_samples;
		if ((st = ensure(ctx, B_PIXACC, (size_t)n_pix * 24)) != RT_OK) return st;
		double* pixacc = (double*)ctx->buf[B_PIXACC];
		CK(cudaMemsetAsync(pixacc, 0, (size_t)n_pix * 24, ctx->stream));
		if ((st = ensure(ctx, B_REG_A, (size_t)n_pix * sizeof(Region))) != RT_OK) return st;
		int cur = B_REG_A, nxt = B_REG_B;
		long long n_regions = n_pix;
		if (n_pix > 0) {
			k_adapt_init<<<blocks_for(n_pix, 256), 256, 0, ctx->stream>>>(A, n_pix, (Region*)ctx->buf[cur]);
			ctx->launches++;
		}
		unsigned long long* next_count = (unsigned long long*)ctx->buf[B_SMALL] + 16;
		for (int level = 0; n_regions > 0 && level < 64; level++) {
			const long long n_samples = n_regions * N;
			if (n_samples > 0x7fffffffll) { ctx->error = "too many adaptive samples for one launch"; return RT_ERR_UNSUPPORTED; }
			if ((st = ensure(ctx, B_XS, (size_t)n_samples * 8)) != RT_OK) return st;
			if ((st = ensure(ctx, B_YS, (size_t)n_samples * 8)) != RT_OK) return st;
			if ((st = ensure(ctx, B_ACCUM, (size_t)n_samples * 24)) != RT_OK) return st;
			if ((st = ensure(ctx, nxt, (size_t)n_regions * 4 * sizeof(Region))) != RT_OK) return st;
			k_adapt_samples<<<blocks_for(n_samples, 256), 256, 0, ctx->stream>>>((const Region*)ctx->buf[cur], n_regions, N, (double*)ctx->buf[B_XS], (double*)ctx->buf[B_YS]);
			ctx->launches++;
			PoolArgs L = A;
			L.src = SRC_XY;
			L.xs = (const double*)ctx->buf[B_XS];
			L.ys = (const double*)ctx->buf[B_YS];
			L.accum = (double*)ctx->buf[B_ACCUM];
			L.n_slots = n_samples * A.lenses;
			CK(cudaMemsetAsync(L.accum, 0, (size_t)n_samples * 24, ctx->stream));
			CK(cudaMemsetAsync(next_count, 0, 8, ctx->stream));
			if ((st = launch_wave(ctx, L)) != RT_OK) return st;
			k_adapt_resolve<<<blocks_for(n_regions, 256), 256, 0, ctx->stream>>>((const Region*)ctx->buf[cur], n_regions, N, L.accum, *p, pixacc,
			                                                                  (Region*)ctx->buf[nxt], next_count, (unsigned long long)n_regions * 4);
			ctx->launches++;
			CK(cudaGetLastError());
			unsigned long long nc = 0;
			CK(cudaMemcpyAsync(&nc, next_count, 8, cudaMemcpyDeviceToHost, ctx->stream));
			CK(cudaStreamSynchronize(ctx->stream));
			n_regions = (long long)nc;
			const int t = cur; cur = nxt; nxt = t;
		}
		if (n_pix > 0) {
			k_adapt_store<<<blocks_for(n_pix, 256), 256, 0, ctx->stream>>>(A, n_pix, pixacc, fb);
			ctx->launches++;
			CK(cudaGetLastError());
		}
	}
	CK(cudaEventRecord(ctx->ev1, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	float ms = 0.f;
	CK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
	*dev_rgb = fb;
	return read_stats(ctx, stats, ms);
}

rt_status rt_render_tiles(rt_ctx* ctx, const rt_params* p, int w, int h, int part, int n_parts, uint8_t* rgb_out, rt_stats* stats)
{
	if (!ctx) return RT_ERR_ARG;
	if (!rgb_out) { ctx->error = "null output buffer"; return RT_ERR_ARG; }
	void* dev = nullptr;
	rt_status st = rt_render_device(ctx, p, w, h, part, n_parts, &dev, stats);
	if (st != RT_OK) return st;
	if (n_parts == 1) {
		CK(cudaMemcpyAsync(rgb_out, dev, (size_t)w * h * 3, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaStreamSynchronize(ctx->stream));
		return RT_OK;
	}
	// copy only this partition's tiles: leave the caller's other pixels untouched
	std::vector<uint8_t> tmp((size_t)w * h * 3);
	CK(cudaMemcpyAsync(tmp.data(), dev, tmp.size(), cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	const int tiles_x = (w + RT_TILE - 1) / RT_TILE, tiles_y = (h + RT_TILE - 1) / RT_TILE;
	for (int t = part; t < tiles_x * tiles_y; t += n_parts) {
		const int x0 = (t % tiles_x) * RT_TILE, y0 = (t / tiles_x) * RT_TILE;
		const int x1 = x0 + RT_TILE < w ? x0 + RT_TILE : w, y1 = y0 + RT_TILE < h ? y0 + RT_TILE : h;
		for (int j = y0; j < y1; j++)
			std::memcpy(rgb_out + ((size_t)j * w + x0) * 3, tmp.data() + ((size_t)j * w + x0) * 3, (size_t)(x1 - x0) * 3);
	}
	return RT_OK;
}

rt_status rt_render(rt_ctx* ctx, const rt_params* p, int w, int h, uint8_t* rgb_out, rt_stats* stats)
{
	return rt_render_tiles(ctx, p, w, h, 0, 1, rgb_out, stats);
}

rt_status rt_trace_pixels(rt_ctx* ctx, const rt_params* p, int w, int h, const int32_t* ij, int n, double* rgb_out)
{
	if (!ctx) return RT_ERR_ARG;
	rt_status st = check_params(ctx, p);
	if (st != RT_OK) return st;
	if (!ctx->has_scene) { ctx->error = "no scene uploaded"; return RT_ERR_NO_SCENE; }
	if (w <= 0 || h <= 0 || n < 0 || (n && (!ij || !rgb_out))) { ctx->error = "bad arguments"; return RT_ERR_ARG; }
	if (n == 0) return RT_OK;
	CK(cudaSetDevice(ctx->device));
	ctx->launches = 0;
	ctx->rounds = 0;
	// RayTracer::tracePixel's window coordinates (RayTracer.cpp:87-88)
	const int s = (p->aa_mode != RT_AA_NONE && p->aa_mode != RT_AA_ADAPTIVE) ? p->aa_samples : 1;
	std::vector<double> xs(n), ys(n);
	for (int k = 0; k < n; k++) {
		xs[k] = double(ij[2 * k]) / (double(w) * s);
		ys[k] = double(ij[2 * k + 1]) / (double(h) * s);
	}
	if ((st = ensure(ctx, B_XS, (size_t)n * 8)) != RT_OK) return st;
	if ((st = ensure(ctx, B_YS, (size_t)n * 8)) != RT_OK) return st;
	if ((st = ensure(ctx, B_ACCUM, (size_t)n * 24)) != RT_OK) return st;
	if ((st = ensure(ctx, B_PIXACC, (size_t)n * 24)) != RT_OK) return st;
	CK(cudaMemcpyAsync(ctx->buf[B_XS], xs.data(), (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
	CK(cudaMemcpyAsync(ctx->buf[B_YS], ys.data(), (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
	CK(cudaMemsetAsync(ctx->buf[B_SMALL], 0, 64, ctx->stream));
	PoolArgs A;
	std::memset(&A, 0, sizeof(A));
	A.prm = *p;
	A.src = SRC_XY;
	A.w = w; A.h = h;
	A.spp_side = 1;
	A.lenses = p->dof_on ? p->dof_divs + 1 : 1;
	A.n_parts = 1;
	A.depth = p->depth;
	A.xs = (const double*)ctx->buf[B_XS];
	A.ys = (const double*)ctx->buf[B_YS];
	A.accum = (double*)ctx->buf[B_ACCUM];
	A.n_slots = (long long)n * A.lenses;
	CK(cudaMemsetAsync(A.accum, 0, (size_t)n * 24, ctx->stream));
	if ((st = launch_wave(ctx, A)) != RT_OK) return st;
	double* accum2 = nullptr;
	if (p->anaglyph) {
		if ((st = ensure(ctx, B_ACCUM2, (size_t)n * 24)) != RT_OK) return st;
		accum2 = (double*)ctx->buf[B_ACCUM2];
		CK(cudaMemsetAsync(accum2, 0, (size_t)n * 24, ctx->stream));
		PoolArgs A2 = A;
		A2.accum = accum2;
		A2.eye_shift = 0.25;
		if ((st = launch_wave(ctx, A2)) != RT_OK) return st;
	}
	k_sample_values<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(A.accum, accum2, n, *p, (double*)ctx->buf[B_PIXACC]);
	ctx->launches++;
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(rgb_out, ctx->buf[B_PIXACC], (size_t)n * 24, cudaMemcpyDeviceToHost, ctx->stream));
	return read_stats(ctx, nullptr, 0.f);
}

rt_status rt_trace_rays(rt_ctx* ctx, const rt_params* p, const rt_ray* rays, int n, int depth, rt_hit* hits_out, double* rgb_out)
{
	if (!ctx) return RT_ERR_ARG;
	rt_status st = check_params(ctx, p);
	if (st != RT_OK) return st;
	if (!ctx->has_scene) { ctx->error = "no scene uploaded"; return RT_ERR_NO_SCENE; }
	if (n < 0 || (n && !rays) || depth < 0 || depth > 200) { ctx->error = "bad arguments"; return RT_ERR_ARG; }
	if (n == 0) return RT_OK;
	CK(cudaSetDevice(ctx->device));
	ctx->launches = 0;
	ctx->rounds = 0;
	if ((st = ensure(ctx, B_RAYS, (size_t)n * sizeof(rt_ray))) != RT_OK) return st;
	CK(cudaMemcpyAsync(ctx->buf[B_RAYS], rays, (size_t)n * sizeof(rt_ray), cudaMemcpyHostToDevice, ctx->stream));
	CK(cudaMemsetAsync(ctx->buf[B_SMALL], 0, 64, ctx->stream));
	if (hits_out) {
		if ((st = ensure(ctx, B_HITS, (size_t)n * sizeof(rt_hit))) != RT_OK) return st;
		k_hits<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(ctx->S, (const rt_ray*)ctx->buf[B_RAYS], n, (rt_hit*)ctx->buf[B_HITS]);
		ctx->launches++;
		CK(cudaGetLastError());
		CK(cudaMemcpyAsync(hits_out, ctx->buf[B_HITS], (size_t)n * sizeof(rt_hit), cudaMemcpyDeviceToHost, ctx->stream));
	}
	if (rgb_out) {
		if ((st = ensure(ctx, B_ACCUM, (size_t)n * 24)) != RT_OK) return st;
		PoolArgs A;
		std::memset(&A, 0, sizeof(A));
		A.prm = *p;
		A.src = SRC_RAYS;
		A.spp_side = 1;
		A.lenses = 1;
		A.n_parts = 1;
		A.depth = depth;
		A.rays = (const rt_ray*)ctx->buf[B_RAYS];
		A.accum = (double*)ctx->buf[B_ACCUM];
		A.n_slots = n;
		CK(cudaMemsetAsync(A.accum, 0, (size_t)n * 24, ctx->stream));
		if ((st = launch_wave(ctx, A)) != RT_OK) return st;
		CK(cudaMemcpyAsync(rgb_out, A.accum, (size_t)n * 24, cudaMemcpyDeviceToHost, ctx->stream));
	}
	return read_stats(ctx, nullptr, 0.f);
}


rt_status rt_selftest_division(rt_ctx* ctx, long long n, unsigned long long seed, int exp_range, unsigned long long* mismatches)
{
	if (!ctx || !mismatches || n <= 0 || exp_range < 0 || exp_range > 250) return RT_ERR_ARG;
	CK(cudaSetDevice(ctx->device));
	unsigned long long* cnt = (unsigned long long*)ctx->buf[B_SMALL] + 24;
	CK(cudaMemsetAsync(cnt, 0, 8, ctx->stream));
	k_div_selftest<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(seed, n, exp_range, cnt);
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(mismatches, cnt, 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	return RT_OK;
}

} // extern "C"
