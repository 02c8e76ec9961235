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
#include <cub/device/device_radix_sort.cuh>
#include "rt_b200.h"
#include "rt_arena_writer.hpp"
#include "device/rt_kernels.cuh"
#include "device/rt_derive.hpp"
#include "device/rt_derive_dev.cuh"

extern "C" rt_status rt_arena_validate_tables(const void* arena, size_t bytes); // host/host_api.cpp

using namespace rtd;

#define RT_SORT_MIN (1ll << 20) // rounds with fewer possible shadow rays are walked in queue order (a sort is ~20 us of launches)
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
	void* buf[64] = {nullptr};
	size_t buf_bytes[64] = {0};
	unsigned long long launches = 0;
	// wavefront path
	long long wave_width = 0;              // rays per round; 0 = chosen per call from the memory budget below (RT_B200_WAVE overrides)
	size_t queue_budget = 0;               // bytes the ray queues of a round may take: a quarter of the device memory
	int walk_budget = 96;                  // tree nodes a single thread spends on one ray before the warp-cooperative walk takes it (RT_B200_BUDGET; 0 = never)
	int n_area_lights = 0;                 // area + spot lights of the uploaded scene
	bool budget_auto = true;               // no RT_B200_BUDGET given: the budget adapts to the scene
	bool use_pool = false;                 // RT_B200_POOL=1: the fused persistent kernel instead
	unsigned long long* pinned = nullptr;  // round counters read back by the host
	cudaEvent_t ev_round = nullptr;
	cudaStream_t stream2[2] = {nullptr, nullptr}; // shadow rays of round r (stream r & 1) overlap round r+1's closest-hit work AND the
	                                              // tail of round r-1's shadow kernel: a launch lasts as long as its slowest ray
	cudaEvent_t ev_shade[2] = {nullptr, nullptr}, ev_shadow[2] = {nullptr, nullptr};
	bool overlap = true;                   // RT_B200_OVERLAP=0: everything on one stream
	// any-hit shadow stage: fraction of round r's shadow rays it left to the complete walk, from the last frame of this scene
	// (a property of the scene and the settings, round by round: primary hits see mostly opaque blockers, deep bounces sit
	// inside glass); a round whose fraction is above the break-even is walked once, completely
	double undecided[32] = {0};
	long long any_n[2] = {0, 0};           // rays of the any-hit launch in flight per round parity (0: that launch was one-kernel)
	int any_round[2] = {0, 0};             // ... and the round it belongs to
	bool shadow_two = true;                // RT_B200_ANYHIT=0: one shadow kernel (k_wave_shadow) instead of any-hit classification + complete walk of the rest
	long long tail_rays = 0;               // a round of at most this many rays, with nothing else left to start, is finished by ONE launch of the fused kernel (RT_B200_TAIL; 0 = off)
	long long sort_min = RT_SORT_MIN;      // rounds with fewer rays are walked in queue order (RT_B200_SORTMIN)
	long long chunk_samples = 1ll << 28;   // camera samples per launch of the grid path (RT_B200_CHUNK; a frame above it is rendered in chunks of tiles)
	bool sparse_small = true;              // RT_B200_SPARSE=0: small rounds are launched densely (32 rays per warp) like large ones
	bool sort_closest = true;              // RT_B200_SORTC=0: reflection / refraction rays are walked in queue order instead of (octant, origin cell) order
	bool sort_shadow = true;               // RT_B200_SORT=0: shadow rays are walked in queue order instead of (light, origin cell) order
	unsigned long long rounds = 0;
	uint8_t* stage = nullptr;              // page-locked staging of rt_render_tiles' packed tiles
	size_t stage_bytes = 0;
	void* static_scratch = nullptr;        // device address of g_static_scratch on this context's GPU (rt_intersect_static)
	// optional per-stage timing
	bool profile = false;
	std::vector<cudaEvent_t> ev_pool;
	size_t ev_used = 0;
	struct Span { int stage; cudaEvent_t a, b; };
	std::vector<Span> spans;
};

namespace {

enum { B_ACCUM = 0, B_ACCUM2, B_POOL_D, B_POOL_I, B_SMALL, B_FB, B_XS, B_YS, B_REG_A, B_REG_B, B_PIXACC, B_RAYS, B_HITS,
       B_WQ0_D, B_WQ0_I, B_WQ1_D, B_WQ1_I, B_WSP_D, B_WSP_I, B_SQ_D, B_SQ_I, B_WH_T, B_WH_OF, B_HEAVY, B_RETRY, B_SQ2_D, B_SQ2_I,
       B_SK_A, B_SK_B, B_SI_A, B_SI_B, B_CK, B_CI, B_CX_A, B_CX_B, B_ARENA, B_SNODES2, B_MNODES2, B_SNODES2F, B_MNODES2F, B_SBOXF, B_MTREES, B_TRIS, B_MSLACK, B_PRE_S, B_PRE_M, B_TMP_A, B_TMP_B, B_TMP_C, B_TMP_D, B_CUBTMP, B_CUBTMP2, B_CUBTMP3, B_PACKED, B_STOP, B_BASE, B_BASE2, B_SPRIMF, B_COUNT };
static_assert(B_COUNT <= 64, "rt_ctx::buf too small");

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

#define RT_UNDECIDED_MAX 0.13
// log2 of the lanes given to one ray of a walk launch over n rays: as many as keep the launch within the GPU's resident threads
static inline int sparse_shift(const rt_ctx* ctx, long long n)
{
	if (!ctx->sparse_small || n <= 0) return 0;
	const long long resident = (long long)ctx->sm_count * 1024;
	int s = 0;
	while (s < 5 && (n << (s + 1)) <= resident) s++;
	return s;
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
rt_status launch_tail(rt_ctx* ctx, const PoolArgs& A, const RayQueue& q, long long n);

cudaEvent_t next_event(rt_ctx* ctx)
{
	if (ctx->ev_used == ctx->ev_pool.size()) {
		cudaEvent_t e = nullptr;
		cudaEventCreate(&e);
		ctx->ev_pool.push_back(e);
	}
	return ctx->ev_pool[ctx->ev_used++];
}
// STAGE(id, launch): the launch, bracketed by events when profiling is on
#define STAGE_ON(strm, id, ...)                                               \
	do {                                                                      \
		cudaEvent_t a_ = nullptr, b_ = nullptr;                               \
		if (ctx->profile) { a_ = next_event(ctx); b_ = next_event(ctx); cudaEventRecord(a_, strm); } \
		__VA_ARGS__;                                                          \
		ctx->launches++;                                                      \
		if (ctx->profile) { cudaEventRecord(b_, strm); ctx->spans.push_back({id, a_, b_}); } \
	} while (0)
#define STAGE(id, ...) STAGE_ON(ctx->stream, id, __VA_ARGS__)

// The shadow stage of one round over (at most) n rays of queue SQ, on stream s: any-hit classification of every ray, then the
// complete walk of the rays it could not decide - or, RT_B200_ANYHIT=0, the complete walk for all of them.
rt_status launch_shadow(rt_ctx* ctx, const PoolArgs& A, const RayQueue& SQ, int par, WaveCounters* cnt, long long n, const unsigned* perm, int ssh, cudaStream_t s, int round_index)
{
	// the two-kernel form pays when few rays need the second kernel (break-even measured at about 13 %)
	const int ri = round_index < 31 ? round_index : 31;
	if (!ctx->shadow_two || ctx->undecided[ri] > RT_UNDECIDED_MAX) {
		ctx->any_n[par] = 0;
		STAGE_ON(s, RT_STAGE_SHADOW, k_wave_shadow<<<blocks_for(n << ssh, RT_TRACE_THREADS), RT_TRACE_THREADS, 0, s>>>(ctx->S, A, SQ, par, cnt, perm, ssh));
		return RT_OK;
	}
	const int bx = par ? B_CX_B : B_CX_A;
	rt_status st;
	if ((st = ensure(ctx, bx, (size_t)SQ.cap * sizeof(int))) != RT_OK) return st;
	int* list = (int*)ctx->buf[bx];
	ctx->any_n[par] = n;
	ctx->any_round[par] = ri;
	CK(cudaMemsetAsync(&cnt->n_complex[par], 0, sizeof(unsigned long long), s));
	STAGE_ON(s, RT_STAGE_SHADOW, k_wave_shadow_any<<<blocks_for(n << ssh, RT_TRACE_THREADS), RT_TRACE_THREADS, 0, s>>>(ctx->S, A, SQ, par, cnt, perm, ssh, list));
	long long grid = (long long)ctx->sm_count * RT_SHADOW_MINB;
	if (grid > (long long)blocks_for(n, RT_TRACE_THREADS)) grid = blocks_for(n, RT_TRACE_THREADS);
	STAGE_ON(s, RT_STAGE_SHADOW, k_wave_shadow_list<<<(unsigned)(grid > 0 ? grid : 1), RT_TRACE_THREADS, 0, s>>>(ctx->S, A, SQ, par, cnt, list));
	return RT_OK;
}

// The trace loop over n_slots primary-ray slots as a streaming wavefront (device/rt_wave.cuh).
rt_status launch_wave(rt_ctx* ctx, PoolArgs& A)
{
	if (ctx->use_pool || A.prm.aterm_thresh > 0.0) return launch_pool(ctx, A); // adaptive termination needs the shadowed colour before it recurses
	if (A.n_slots <= 0) return RT_OK;
	// shadow rays one hit can make: one per point / directional light, softShadowRes per area or spot light
	const int L = (ctx->S.n_lights - ctx->n_area_lights) + ctx->n_area_lights * (A.prm.soft_shadow_res > 0 ? A.prm.soft_shadow_res : 0);
	// Rays per round.  Wide rounds are faster (fewer launches and tails, fuller late rounds: C5 335 -> 370
	// Mrays/s from 2 Mi to 32 Mi rays, profiles/r1_variants.txt); the queues cost about 4 * 104 + 2 * L * 80 + 24
	// bytes per ray of the round, so the width is what a quarter of the device memory holds, at most 32 Mi.
	long long W = ctx->wave_width;
	if (W <= 0) {
		const size_t per_ray = 4 * 104 + 2 * (size_t)(L > 0 ? L : 1) * 80 + 24 + (size_t)(L > 0 ? L : 1) * 16; // 16: sort keys and permutation
		W = (long long)(ctx->queue_budget / per_ray) / 1024 * 1024;
		// the spill stack (allocated only if a round ever has more than W children) may take another half of the memory
		const long long spill_fit = (long long)(2 * ctx->queue_budget / (((size_t)(A.depth < 30 ? A.depth : 30) + 2) * 104)) / 1024 * 1024;
		if (W > spill_fit) W = spill_fit;
		if (W > (32ll << 20)) W = 32ll << 20;
		if (W < (1ll << 20)) W = 1ll << 20;
	}
	const long long fit = ((A.n_slots + 1023) / 1024) * 1024;
	if (W > fit) W = fit;
	rt_status st;
	RayQueue cur, nxt, sq[2], sp;
	sp.d = nullptr; sp.i = nullptr; sp.cap = 0; sp.nd = 12;
	if ((st = make_queue(ctx, B_WQ0_D, B_WQ0_I, 2 * W, 12, cur)) != RT_OK) return st;
	if ((st = make_queue(ctx, B_WQ1_D, B_WQ1_I, 2 * W, 12, nxt)) != RT_OK) return st;
	if ((st = make_queue(ctx, B_SQ_D, B_SQ_I, W * (L > 0 ? L : 1), 9, sq[0])) != RT_OK) return st;
	if ((st = make_queue(ctx, B_SQ2_D, B_SQ2_I, W * (L > 0 ? L : 1), 9, sq[1])) != RT_OK) return st;
	// per-stage profiling wants clean kernel times: no overlap then
	const bool overlap = ctx->overlap && !ctx->profile;
	bool shadow_pending[2] = {false, false};
	long long round = 0;
	if ((st = ensure(ctx, B_WH_T, (size_t)W * sizeof(double))) != RT_OK) return st;
	if ((st = ensure(ctx, B_WH_OF, (size_t)W * sizeof(int2))) != RT_OK) return st;
	if ((st = ensure(ctx, B_HEAVY, (size_t)W * sizeof(int))) != RT_OK) return st;
	int* heavy = (int*)ctx->buf[B_HEAVY];
	if ((st = ensure(ctx, B_RETRY, (size_t)W * sizeof(int))) != RT_OK) return st;
	int* retry = (int*)ctx->buf[B_RETRY];
	HitQueue H;
	H.t = (double*)ctx->buf[B_WH_T];
	H.of = (int2*)ctx->buf[B_WH_OF];
	WaveCounters* cnt = (WaveCounters*)ctx->buf[B_SMALL];
	A.stats = cnt->stats;
	long long cur_n = 0, sp_n = 0, cam = 0;
	for (;;) {
		// The floor of a small round is its slowest walk, paid once per stage and recursion level.  When only a few rays are
		// left - no camera samples to start, nothing spilled - the fused persistent kernel takes the queue and carries every
		// ray's whole subtree to the end in ONE launch (its warps keep their own LIFO pools).  Same arithmetic, and the
		// accumulators do not care about the order: the frame is the same bytes (tests/test_gpu_parity.py).
		if (ctx->tail_rays > 0 && cur_n > 0 && cur_n <= ctx->tail_rays && sp_n == 0 && cam >= A.n_slots) {
			if ((st = launch_tail(ctx, A, cur, cur_n)) != RT_OK) return st;
			ctx->rounds++;
			break;
		}
		long long base = 0, m = 0;
		long long n_kids = cur_n < W ? cur_n : W; // the batch's first n_kids rays are children of the previous round
		if (cur_n > W) {
			// more children than one round takes: the oldest go to the spill stack
			const long long spill = cur_n - W;
			if (!sp.d) {
				const long long cap = W * ((A.depth < 30 ? A.depth : 30) + 2);
				if ((st = make_queue(ctx, B_WSP_D, B_WSP_I, cap, 12, sp)) != RT_OK) return st;
			}
			if (sp_n + spill > sp.cap) { ctx->error = "wavefront spill stack overflow"; return RT_ERR_OVERFLOW; }
			STAGE(RT_STAGE_OTHER, k_wave_move<<<blocks_for(spill, 256), 256, 0, ctx->stream>>>(cur, 0, sp, sp_n, spill));
			sp_n += spill;
			base = spill;
			m = W;
		} else {
			long long need = W - cur_n;
			const long long un = need < sp_n ? need : sp_n;
			if (un > 0) {
				STAGE(RT_STAGE_OTHER, k_wave_move<<<blocks_for(un, 256), 256, 0, ctx->stream>>>(sp, sp_n - un, cur, cur_n, un));
				sp_n -= un;
				cur_n += un;
				need -= un;
			}
			const long long gen = need < A.n_slots - cam ? need : A.n_slots - cam;
			if (gen > 0) {
				STAGE(RT_STAGE_GEN, k_wave_gen<<<blocks_for(gen, RT_WAVE_THREADS), RT_WAVE_THREADS, 0, ctx->stream>>>(ctx->S, A, cur, cur_n, cam, gen));
				cur_n += gen;
				cam += gen;
			}
			m = cur_n;
		}
		if (m == 0) break;
		const int par = (int)(round & 1);
		CK(cudaMemsetAsync(&cnt->n_children, 0, 4 * sizeof(unsigned long long), ctx->stream)); // children, heavy, cursor, retry
		// a forced budget (RT_B200_BUDGET, the tests) sends its leftovers straight to the cooperative kernel
		const bool second = ctx->budget_auto && ctx->walk_budget > 0;
		const unsigned* cperm = nullptr;
		if (ctx->sort_closest && n_kids >= ctx->sort_min) {
			if ((st = ensure(ctx, B_CK, (size_t)W * 2 * sizeof(unsigned))) != RT_OK) return st;
			if ((st = ensure(ctx, B_CI, (size_t)W * 2 * sizeof(unsigned))) != RT_OK) return st;
			unsigned* keys = (unsigned*)ctx->buf[B_CK];
			unsigned* idx = (unsigned*)ctx->buf[B_CI];
			size_t tmp_bytes = 0;
			CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys + W, idx, idx + W, (int)n_kids, 0, 30, ctx->stream));
			if ((st = ensure(ctx, B_CUBTMP3, tmp_bytes)) != RT_OK) return st;
			STAGE(RT_STAGE_OTHER, k_closest_keys<<<blocks_for(n_kids, 256), 256, 0, ctx->stream>>>(ctx->S, cur, base, n_kids, keys, idx));
			STAGE(RT_STAGE_OTHER, CK(cub::DeviceRadixSort::SortPairs(ctx->buf[B_CUBTMP3], tmp_bytes, keys, keys + W, idx, idx + W, (int)n_kids, 0, 30, ctx->stream)));
			cperm = idx + W;
		} else n_kids = 0;
		const int csh = sparse_shift(ctx, m);
		STAGE(RT_STAGE_CLOSEST, k_wave_closest<<<blocks_for(m << csh, RT_TRACE_THREADS), RT_TRACE_THREADS, 0, ctx->stream>>>(ctx->S, cur, base, m, H, ctx->walk_budget,
		                                                                                                                second ? retry : heavy, second ? &cnt->n_retry : &cnt->n_heavy, cperm, n_kids, csh));
		if (second) {
			// second chance with 8 x the budget: grid for 1/32 of the round (a grid-stride loop takes whatever there is)
			const unsigned rblocks = blocks_for((m + 31) / 32, RT_TRACE_THREADS);
			STAGE(RT_STAGE_COOP, k_wave_closest_retry<<<rblocks < 8u ? 8u : rblocks, RT_TRACE_THREADS, 0, ctx->stream>>>(ctx->S, cur, base, H, ctx->walk_budget * 8, retry, heavy, cnt));
		}
		if (ctx->walk_budget > 0)
			STAGE(RT_STAGE_COOP, k_wave_closest_heavy<<<ctx->sm_count * 4 * (2 / RT_COOP_WARPS), 32 * RT_COOP_WARPS, 0, ctx->stream>>>(ctx->S, cur, base, H, heavy, cnt));
		// this parity's shadow queue is free once the shadow kernel of two rounds ago is done; only the
		// shade stage writes it, so the closest-hit walks above did not have to wait
		if (shadow_pending[par] && overlap) CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_shadow[par], 0));
		shadow_pending[par] = false;
		const long long prev_any_n = ctx->any_n[par]; // this parity's previous any-hit launch (two rounds ago) has finished by now
		const int prev_any_round = ctx->any_round[par];
		CK(cudaMemsetAsync(&cnt->n_shadow[par], 0, sizeof(unsigned long long), ctx->stream));
		STAGE(RT_STAGE_SHADE, k_wave_shade<<<blocks_for(m, RT_WAVE_THREADS), RT_WAVE_THREADS, 0, ctx->stream>>>(ctx->S, A, cur, base, m, H, nxt, sq[par], par, cnt));
		CK(cudaMemcpyAsync(ctx->pinned, &cnt->n_children, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream)); // children, heavy, cursor, retry, complex [2], shadow rays [2]
		CK(cudaEventRecord(ctx->ev_round, ctx->stream));
		// Large rounds: the shadow rays are walked in (light, origin cell) order.  The radix sort needs the number of rays on
		// the host, which arrives with the counters the host waits for anyway - so the wait moves in front of the launch.
		const bool sorted = L > 0 && ctx->sort_shadow && m * L >= ctx->sort_min;
		bool synced = false;
		if (sorted) {
			CK(cudaEventSynchronize(ctx->ev_round));
			synced = true;
			long long ns = (long long)ctx->pinned[6 + par];
			if (ns > sq[par].cap) ns = sq[par].cap;
			cudaStream_t sstream = overlap ? ctx->stream2[par] : ctx->stream;
			if (overlap) {
				CK(cudaEventRecord(ctx->ev_shade[par], ctx->stream));
				CK(cudaStreamWaitEvent(sstream, ctx->ev_shade[par], 0));
			}
			if (ns > 0) {
				// sort buffers are per parity too (the other parity's walk may still be reading its permutation)
				const int bk = par ? B_SK_B : B_SK_A, bi = par ? B_SI_B : B_SI_A;
				if ((st = ensure(ctx, bk, (size_t)sq[par].cap * 2 * sizeof(unsigned))) != RT_OK) return st;
				if ((st = ensure(ctx, bi, (size_t)sq[par].cap * 2 * sizeof(unsigned))) != RT_OK) return st;
				unsigned* keys = (unsigned*)ctx->buf[bk];
				unsigned* idx = (unsigned*)ctx->buf[bi];
				size_t tmp_bytes = 0;
				CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys + sq[par].cap, idx, idx + sq[par].cap, (int)ns, 0, 32, sstream));
				const int bt = par ? B_CUBTMP2 : B_CUBTMP;
				if ((st = ensure(ctx, bt, tmp_bytes)) != RT_OK) return st;
				STAGE_ON(sstream, RT_STAGE_OTHER, k_shadow_keys<<<blocks_for(ns, 256), 256, 0, sstream>>>(ctx->S, sq[par], ns, keys, idx));
				STAGE_ON(sstream, RT_STAGE_OTHER, CK(cub::DeviceRadixSort::SortPairs(ctx->buf[bt], tmp_bytes, keys, keys + sq[par].cap, idx, idx + sq[par].cap, (int)ns, 0, 32, sstream)));
				const int ssh = sparse_shift(ctx, ns);
				if ((st = launch_shadow(ctx, A, sq[par], par, cnt, ns, idx + sq[par].cap, ssh, sstream, (int)round)) != RT_OK) return st;
			}
			if (overlap) { CK(cudaEventRecord(ctx->ev_shadow[par], sstream)); shadow_pending[par] = true; }
		} else if (L > 0) {
			cudaStream_t sstream = overlap ? ctx->stream2[par] : ctx->stream;
			if (overlap) {
				CK(cudaEventRecord(ctx->ev_shade[par], ctx->stream));
				CK(cudaStreamWaitEvent(sstream, ctx->ev_shade[par], 0));
			}
			// grid sized for the most shadow rays this round can have made; blocks past the count leave at once
			const int ssh = sparse_shift(ctx, m * L);
			if ((st = launch_shadow(ctx, A, sq[par], par, cnt, m * L, nullptr, ssh, sstream, (int)round)) != RT_OK) return st;
			if (overlap) { CK(cudaEventRecord(ctx->ev_shadow[par], sstream)); shadow_pending[par] = true; }
		}
		CK(cudaGetLastError());
		if (!synced) CK(cudaEventSynchronize(ctx->ev_round)); // the host plans the next round while the shadow rays run
		cur_n = (long long)ctx->pinned[0];
		if (cur_n > nxt.cap) cur_n = nxt.cap; // overflow was flagged by the kernel
		if (prev_any_n >= 4096) ctx->undecided[prev_any_round] = (double)ctx->pinned[4 + par] / (double)prev_any_n; // n_complex of that launch
		// the node budget follows the scene: if more than ~1.5 % of a round's rays ran out of it, the
		// ordinary walk of this scene is simply longer (deep scene-level trees) - double it.  Either
		// path gives the same hit, so this only moves work between the two kernels.
		// (pinned[3]: the rays that ran out of the first budget.  A round walked before the budget has grown is not
		// lost: its leftovers get a second single-thread chance, k_wave_closest_retry.)
		if (ctx->budget_auto && ctx->walk_budget > 0 && m >= 4096 && (long long)ctx->pinned[3] * 64 > m && ctx->walk_budget < 16384) ctx->walk_budget *= 2;
		round++;
		const RayQueue tmp = cur;
		cur = nxt;
		nxt = tmp;
		ctx->rounds++;
	}
	// everything that follows on the main stream (resolve kernels, the frame copy) needs the last shadow rays
	for (int par = 0; par < 2; par++)
		if (shadow_pending[par]) CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_shadow[par], 0));
	return RT_OK;
}

// One launch of the fused persistent kernel over n_slots primary-ray slots.
rt_status launch_pool(rt_ctx* ctx, PoolArgs& A)
{
	const int warps_per_block = RT_POOL_THREADS / 32;
	const long long warps = (long long)ctx->pool_blocks * warps_per_block;
	const bool defer = !(A.prm.aterm_thresh > 0.0) && ctx->n_area_lights == 0; // the fused kernel defers one shadow ray per light only; area lights are shaded inline
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

// The rest of a frame from the wavefront's queue: k_trace_pool with the queue as its source of rays.
rt_status launch_tail(rt_ctx* ctx, const PoolArgs& A, const RayQueue& q, long long n)
{
	PoolArgs T = A;
	T.src = SRC_QUEUE;
	T.q_d = q.d;
	T.q_i = q.i;
	T.q_cap = q.cap;
	T.n_slots = n;
	T.lenses = 1;
	const long long warps = (long long)ctx->pool_blocks * (RT_POOL_THREADS / 32);
	const bool defer = ctx->n_area_lights == 0; // as launch_pool: area lights are shaded inline
	T.cap = 32 * (A.depth + 3) + (defer ? 64 * ctx->S.n_lights + 64 : 0);
	rt_status st;
	if ((st = ensure(ctx, B_POOL_D, (size_t)warps * T.cap * RT_POOL_DOUBLES * sizeof(double))) != RT_OK) return st;
	if ((st = ensure(ctx, B_POOL_I, (size_t)warps * T.cap * 2 * sizeof(int))) != RT_OK) return st;
	T.pool_d = (double*)ctx->buf[B_POOL_D];
	T.pool_i = (int*)ctx->buf[B_POOL_I];
	unsigned long long* small = (unsigned long long*)ctx->buf[B_SMALL];
	T.cursor = small + 20; // behind the wavefront's own counters (WaveCounters: 16 words) and the adaptive-AA region count
	T.stats = small;       // the same ray counters the stage kernels add to
	CK(cudaMemsetAsync(T.cursor, 0, sizeof(unsigned long long), ctx->stream));
	if (defer) { STAGE(RT_STAGE_OTHER, (k_trace_pool<true><<<ctx->pool_blocks, RT_POOL_THREADS, 0, ctx->stream>>>(ctx->S, T))); }
	else { STAGE(RT_STAGE_OTHER, (k_trace_pool<false><<<ctx->pool_blocks, RT_POOL_THREADS, 0, ctx->stream>>>(ctx->S, T))); }
	CK(cudaGetLastError());
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
		stats->culled_shadow_rays = h[6];
		stats->rounds = ctx->rounds;
		for (int k = 0; k < 6; k++) { stats->stage_ms[k] = 0.0; stats->stage_launches[k] = 0; }
		for (const rt_ctx::Span& sp : ctx->spans) {
			float t = 0.f;
			if (cudaEventElapsedTime(&t, sp.a, sp.b) == cudaSuccess) { stats->stage_ms[sp.stage] += t; stats->stage_launches[sp.stage]++; }
		}
	}
	if (h[4]) { ctx->error = "device ray pool overflow"; return RT_ERR_OVERFLOW; }
	return RT_OK;
}

} // namespace

// zeroed sample accumulators for one launch: the limbs (rt_accum.cuh) and the array the camera rays' own terms are stored in
static rt_status prepare_accum(rt_ctx* ctx, PoolArgs& A, int b_accum, int b_base, long long n_samples)
{
	rt_status st;
	if ((st = ensure(ctx, b_accum, (size_t)n_samples * RT_ACC_BYTES)) != RT_OK) return st;
	A.accum = (long long*)ctx->buf[b_accum];
	CK(cudaMemsetAsync(A.accum, 0, (size_t)n_samples * RT_ACC_BYTES, ctx->stream));
	const size_t slots = (size_t)n_samples * (size_t)A.lenses;
	if ((st = ensure(ctx, b_base, slots * 24)) != RT_OK) return st;
	A.base = (double*)ctx->buf[b_base];
	CK(cudaMemsetAsync(A.base, 0, slots * 24, ctx->stream));
	return RT_OK;
}

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
	ctx->queue_budget = prop.totalGlobalMem / 4;
	if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return RT_ERR_CUDA; }
	cudaEventCreate(&ctx->ev0);
	cudaEventCreate(&ctx->ev1);
	cudaEventCreateWithFlags(&ctx->ev_round, cudaEventDisableTiming);
	for (int k = 0; k < 2; k++) cudaStreamCreateWithFlags(&ctx->stream2[k], cudaStreamNonBlocking);
	for (int k = 0; k < 2; k++) { cudaEventCreateWithFlags(&ctx->ev_shade[k], cudaEventDisableTiming); cudaEventCreateWithFlags(&ctx->ev_shadow[k], cudaEventDisableTiming); }
	if (const char* e = getenv("RT_B200_OVERLAP")) ctx->overlap = atoi(e) != 0;
	if (cudaMallocHost((void**)&ctx->pinned, 64) != cudaSuccess) { delete ctx; return RT_ERR_CUDA; }
	if (const char* e = getenv("RT_B200_WAVE")) { const long long v = atoll(e); if (v >= 1024) ctx->wave_width = (v + 1023) / 1024 * 1024; }
	if (const char* e = getenv("RT_B200_POOL")) ctx->use_pool = atoi(e) != 0;
	if (const char* e = getenv("RT_B200_SORT")) ctx->sort_shadow = atoi(e) != 0;
	if (const char* e = getenv("RT_B200_SORTC")) ctx->sort_closest = atoi(e) != 0;
	if (const char* e = getenv("RT_B200_SPARSE")) ctx->sparse_small = atoi(e) != 0;
	if (const char* e = getenv("RT_B200_TAIL")) { const long long v = atoll(e); if (v >= 0) ctx->tail_rays = v; }
	if (const char* e = getenv("RT_B200_SORTMIN")) { const long long v = atoll(e); if (v >= 1024) ctx->sort_min = v; }
	if (const char* e = getenv("RT_B200_CHUNK")) { const long long v = atoll(e); if (v >= 1024 && v <= (1ll << 30)) ctx->chunk_samples = v; }
	if (const char* e = getenv("RT_B200_ANYHIT")) ctx->shadow_two = atoi(e) != 0;
	if (const char* e = getenv("RT_B200_BUDGET")) { const int v = atoi(e); if (v >= 0) { ctx->walk_budget = v; ctx->budget_auto = false; } }
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
	for (int i = 0; i < 64; i++)
		if (ctx->buf[i]) cudaFree(ctx->buf[i]);
	for (cudaEvent_t e : ctx->ev_pool) cudaEventDestroy(e);
	if (ctx->pinned) cudaFreeHost(ctx->pinned);
	if (ctx->stage) cudaFreeHost(ctx->stage);
	if (ctx->ev_round) cudaEventDestroy(ctx->ev_round);
	for (int k = 0; k < 2; k++) { if (ctx->ev_shade[k]) cudaEventDestroy(ctx->ev_shade[k]); if (ctx->ev_shadow[k]) cudaEventDestroy(ctx->ev_shadow[k]); }
	for (int k = 0; k < 2; k++)
		if (ctx->stream2[k]) cudaStreamDestroy(ctx->stream2[k]);
	if (ctx->ev0) cudaEventDestroy(ctx->ev0);
	if (ctx->ev1) cudaEventDestroy(ctx->ev1);
	if (ctx->stream) cudaStreamDestroy(ctx->stream);
	delete ctx;
}

const char* rt_last_error(const rt_ctx* ctx) { return ctx ? ctx->error.c_str() : "null context"; }

rt_status rt_set_profiling(rt_ctx* ctx, int on)
{
	if (!ctx) return RT_ERR_ARG;
	ctx->profile = on != 0;
	return RT_OK;
}

// Builds the traversal layouts of one group of trees on the device (device/rt_derive_dev.cuh).
// nodes: cnt reference nodes in pre-order (n_mesh trees back to back, or the scene tree when
// n_mesh == 0).  *max_depth receives the deepest node's depth (bit 31: deeper than 128).
static rt_status derive_trees(rt_ctx* ctx, const rt_node* nodes, long long cnt, const rt_mesh* ms, int n_mesh, long long scene_items,
                              int b_nodes2, int b_nodes2f, int b_pre, unsigned* err, unsigned* max_depth)
{
	rt_status st;
	if ((st = ensure(ctx, b_nodes2, (size_t)(cnt > 0 ? cnt : 1) * sizeof(DNode2))) != RT_OK) return st;
	if ((st = ensure(ctx, b_nodes2f, (size_t)(cnt > 0 ? cnt : 1) * sizeof(DNode2f))) != RT_OK) return st;
	if ((st = ensure(ctx, b_pre, (size_t)(cnt + 1) * sizeof(int))) != RT_OK) return st;
	if ((st = ensure(ctx, B_TMP_A, (size_t)(cnt + 1) * sizeof(int))) != RT_OK) return st;
	if ((st = ensure(ctx, B_TMP_B, (size_t)(cnt + 1) * sizeof(int))) != RT_OK) return st;
	if ((st = ensure(ctx, B_TMP_C, (size_t)(cnt + 1) * sizeof(int))) != RT_OK) return st;
	if ((st = ensure(ctx, B_TMP_D, (size_t)(cnt + 1) * sizeof(int))) != RT_OK) return st;
	if (cnt <= 0) return RT_OK;
	int* flag = (int*)ctx->buf[B_TMP_A];
	int* pre = (int*)ctx->buf[b_pre];
	const unsigned g = blocks_for(cnt, 256);
	k_derive_flags<<<g, 256, 0, ctx->stream>>>(nodes, cnt, flag);
	size_t tmp_bytes = 0;
	CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flag, pre, (int)cnt, ctx->stream));
	if ((st = ensure(ctx, B_CUBTMP, tmp_bytes)) != RT_OK) return st;
	CK(cub::DeviceScan::ExclusiveSum(ctx->buf[B_CUBTMP], tmp_bytes, flag, pre, (int)cnt, ctx->stream));
	// entries of internal nodes only are written; clear the rest so that the float image below reads defined bytes
	CK(cudaMemsetAsync(ctx->buf[b_nodes2], 0, (size_t)cnt * sizeof(DNode2), ctx->stream));
	k_derive_nodes2<<<g, 256, 0, ctx->stream>>>(nodes, cnt, pre, ms, n_mesh, scene_items, (DNode2*)ctx->buf[b_nodes2], err);
	k_derive_nodes2f<<<g, 256, 0, ctx->stream>>>((const DNode2*)ctx->buf[b_nodes2], cnt, (DNode2f*)ctx->buf[b_nodes2f]);
	// depth: parent links + 7 rounds of pointer jumping (covers depth <= 128)
	int* anc[2] = {(int*)ctx->buf[B_TMP_A], (int*)ctx->buf[B_TMP_C]};
	int* dist[2] = {(int*)ctx->buf[B_TMP_B], (int*)ctx->buf[B_TMP_D]};
	k_depth_self<<<g, 256, 0, ctx->stream>>>(cnt, anc[0], dist[0]);
	k_depth_init<<<g, 256, 0, ctx->stream>>>(nodes, cnt, ms, n_mesh, anc[0], dist[0]);
	int cur = 0;
	for (int r = 0; r < 7; r++) {
		k_depth_jump<<<g, 256, 0, ctx->stream>>>(cnt, anc[cur], dist[cur], anc[cur ^ 1], dist[cur ^ 1]);
		cur ^= 1;
	}
	k_depth_max<<<g, 256, 0, ctx->stream>>>(cnt, anc[cur], dist[cur], max_depth);
	ctx->launches += 12;
	CK(cudaGetLastError());
	return RT_OK;
}

static rt_status upload_scene_impl(rt_ctx* ctx, const void* arena, size_t bytes, bool resident);
rt_status rt_upload_scene(rt_ctx* ctx, const void* arena, size_t bytes) { return upload_scene_impl(ctx, arena, bytes, false); }
// The two halves of rt_upload_scene for a scene that reaches the device by another road (the NCCL broadcast of
// rt_multi_upload_scene): rt_scene_buffer hands out the context's arena buffer, rt_upload_scene_resident does everything
// rt_upload_scene does except the host -> device copy (`arena` = the same bytes in host memory, for the small tables).
rt_status rt_scene_buffer(rt_ctx* ctx, size_t bytes, void** dev)
{
	if (!ctx || !dev || bytes == 0) return RT_ERR_ARG;
	CK(cudaSetDevice(ctx->device));
	ctx->has_scene = false;
	const rt_status st = ensure(ctx, B_ARENA, bytes);
	if (st != RT_OK) return st;
	*dev = ctx->buf[B_ARENA];
	return RT_OK;
}
rt_status rt_upload_scene_resident(rt_ctx* ctx, const void* arena, size_t bytes) { return upload_scene_impl(ctx, arena, bytes, true); }

static rt_status upload_scene_impl(rt_ctx* ctx, const void* arena, size_t bytes, bool resident)
{
	if (!ctx) return RT_ERR_ARG;
	// header, section table and the small tables are checked here; the O(scene) index checks run
	// on the device next to the layout kernels (rt_arena_validate does all of them on the host)
	if (rt_arena_validate_tables(arena, bytes) != RT_OK) { ctx->error = "invalid arena"; return RT_ERR_ARG; }
	CK(cudaSetDevice(ctx->device));
	rt::ArenaView A(arena);
	const rt_light* lights = A.ptr<rt_light>(RT_SEC_LIGHTS);
	int n_area = 0;
	for (uint64_t k = 0; k < A.count(RT_SEC_LIGHTS); k++) {
		if (lights[k].type < RT_LIGHT_DIRECTIONAL || lights[k].type > RT_LIGHT_SPOT) { ctx->error = "unknown light type"; return RT_ERR_ARG; }
		if (lights[k].type >= RT_LIGHT_AREA_RECT) n_area++;
	}
	{
		const rt_object* objs = A.ptr<rt_object>(RT_SEC_OBJECTS);
		for (uint64_t k = 0; k < A.count(RT_SEC_OBJECTS); k++)
			if (objs[k].type < RT_OBJ_SPHERE || objs[k].type > RT_OBJ_TRIMESH) { ctx->error = "unknown object type"; return RT_ERR_ARG; }
	}
	const long long n_snodes = (long long)A.count(RT_SEC_SCENE_NODES), n_mnodes = (long long)A.count(RT_SEC_MESH_NODES);
	const long long n_mitems = (long long)A.count(RT_SEC_MESH_ITEMS), n_sitems = (long long)A.count(RT_SEC_SCENE_ITEMS);
	const int n_mesh = (int)A.count(RT_SEC_MESHES);
	if (n_mitems >= (1ll << 29) || n_sitems >= (1ll << 29) || n_mnodes >= 0x7fffffffll || n_snodes >= 0x7fffffffll || n_snodes == 0) {
		ctx->error = "scene too large for 32-bit node references";
		return RT_ERR_UNSUPPORTED;
	}
	ctx->has_scene = false;
	rt_status st;
	// one copy of the whole arena
	if ((st = ensure(ctx, B_ARENA, bytes)) != RT_OK) return st;
	uint8_t* dev = (uint8_t*)ctx->buf[B_ARENA];
	if (!resident) CK(cudaMemcpyAsync(dev, arena, bytes, cudaMemcpyHostToDevice, ctx->stream));
	DScene& S = ctx->S;
	std::memset(&S, 0, sizeof(S));
#define SEC(field, type, sec) S.field = (const type*)(dev + A.hdr->sections[sec].offset)
	SEC(snodes, rt_node, RT_SEC_SCENE_NODES);
	SEC(sitems, int32_t, RT_SEC_SCENE_ITEMS);
	SEC(objs, rt_object, RT_SEC_OBJECTS);
	SEC(meshes, rt_mesh, RT_SEC_MESHES);
	SEC(mnodes, rt_node, RT_SEC_MESH_NODES);
	SEC(mitems, int32_t, RT_SEC_MESH_ITEMS);
	SEC(faces, rt_face, RT_SEC_FACES);
	SEC(mats, rt_material, RT_SEC_MATERIALS);
	SEC(lights, rt_light, RT_SEC_LIGHTS);
	SEC(texs, rt_texture, RT_SEC_TEXTURES);
	SEC(texdata, uint8_t, RT_SEC_TEXDATA);
	SEC(vmats, int32_t, RT_SEC_VMATS);
	SEC(verts, double, RT_SEC_VERTS);
	SEC(vnormals, double, RT_SEC_VNORMALS);
#undef SEC
	unsigned* flags = (unsigned*)((unsigned long long*)ctx->buf[B_SMALL] + 28); // err, scene depth, mesh depth
	CK(cudaMemsetAsync(flags, 0, 3 * sizeof(unsigned), ctx->stream));
	if ((st = derive_trees(ctx, S.snodes, n_snodes, nullptr, 0, n_sitems, B_SNODES2, B_SNODES2F, B_PRE_S, flags, flags + 1)) != RT_OK) return st;
	if ((st = derive_trees(ctx, S.mnodes, n_mnodes, S.meshes, n_mesh, 0, B_MNODES2, B_MNODES2F, B_PRE_M, flags, flags + 2)) != RT_OK) return st;
	if ((st = ensure(ctx, B_MTREES, (size_t)(n_mesh > 0 ? n_mesh : 1) * sizeof(DTree))) != RT_OK) return st;
	if ((st = ensure(ctx, B_MSLACK, (size_t)(n_mesh > 0 ? n_mesh : 1) * sizeof(double))) != RT_OK) return st;
	if ((st = ensure(ctx, B_TRIS, (size_t)(n_mitems > 0 ? n_mitems : 1) * sizeof(DTri))) != RT_OK) return st;
	if (n_mesh > 0) {
		k_derive_mesh<<<blocks_for(n_mesh, 256), 256, 0, ctx->stream>>>(S.meshes, n_mesh, S.mnodes, (const int*)ctx->buf[B_PRE_M], (DTree*)ctx->buf[B_MTREES], (double*)ctx->buf[B_MSLACK]);
		if (n_mitems > 0)
			k_derive_tris<<<blocks_for(n_mitems, 256), 256, 0, ctx->stream>>>(S.meshes, n_mesh, S.mitems, S.faces, S.verts, n_mitems, (DTri*)ctx->buf[B_TRIS], flags);
	}
	if ((st = ensure(ctx, B_SBOXF, (size_t)(n_sitems > 0 ? n_sitems : 1) * sizeof(DItemBox))) != RT_OK) return st;
	if ((st = ensure(ctx, B_SPRIMF, (size_t)(n_sitems > 0 ? n_sitems : 1) * sizeof(DPrimF))) != RT_OK) return st;
	if (n_sitems > 0) {
		k_derive_itembox<<<blocks_for(n_sitems, 256), 256, 0, ctx->stream>>>(S.sitems, n_sitems, S.objs, (long long)A.count(RT_SEC_OBJECTS), (DItemBox*)ctx->buf[B_SBOXF], (DPrimF*)ctx->buf[B_SPRIMF]);
		ctx->launches++;
	}
	{
		const long long n_vm = (long long)A.count(RT_SEC_VMATS);
		const long long n = n_sitems > n_vm ? n_sitems : n_vm;
		if (n > 0) k_derive_check<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(S.sitems, n_sitems, (long long)A.count(RT_SEC_OBJECTS), S.vmats, n_vm, (long long)A.count(RT_SEC_MATERIALS), flags);
	}
	CK(cudaGetLastError());
	S.snodes2 = (const DNode2*)ctx->buf[B_SNODES2];
	S.mnodes2 = (const DNode2*)ctx->buf[B_MNODES2];
	S.snodes2f = (const DNode2f*)ctx->buf[B_SNODES2F];
	S.mnodes2f = (const DNode2f*)ctx->buf[B_MNODES2F];
	S.sboxf = (const DItemBox*)ctx->buf[B_SBOXF];
	S.sprimf = (const DPrimF*)ctx->buf[B_SPRIMF];
	S.stop = nullptr;
	S.n_top = 0;
	S.pad_top_ = 0;
	int* d_ntop = nullptr;
	if (RT_TOP_K > 0 && n_snodes < (1ll << 30) && n_mnodes < (1ll << 30)) { // RT_REF_TOP must not collide with a node index
		if ((st = ensure(ctx, B_STOP, (size_t)RT_TOP_K * sizeof(DNode2f) + 16)) != RT_OK) return st;
		d_ntop = (int*)((uint8_t*)ctx->buf[B_STOP] + (size_t)RT_TOP_K * sizeof(DNode2f));
		const rt_node& r0 = A.ptr<rt_node>(RT_SEC_SCENE_NODES)[0];
		k_derive_top<<<1, 32, 0, ctx->stream>>>((const DNode2f*)ctx->buf[B_SNODES2F], r0.right >= 0 ? 0 : -1, RT_TOP_K, (DNode2f*)ctx->buf[B_STOP], d_ntop);
		ctx->launches++;
		S.stop = (const DNode2f*)ctx->buf[B_STOP];
	}
	S.mtrees = (const DTree*)ctx->buf[B_MTREES];
	S.mesh_slack = (const double*)ctx->buf[B_MSLACK];
	S.tris = (const DTri*)ctx->buf[B_TRIS];
	S.cam = *A.ptr<rt_camera>(RT_SEC_CAMERA);
	S.glob = *A.ptr<rt_globals>(RT_SEC_GLOBALS);
	// the scene tree's descriptor from its root node (host side: one node)
	const rt_node& root = A.ptr<rt_node>(RT_SEC_SCENE_NODES)[0];
	S.scene_slack = 1e-9 * box_extent(root);
	for (int a = 0; a < 3; a++) { S.stree.lo[a] = root.bmin[a]; S.stree.hi[a] = root.bmax[a]; }
	S.stree.root = root.right >= 0 ? 0 : leaf_ref(root.first, root.count);
	S.stree.pad = 0;
	S.stree.node_off = 0;
	S.n_lights = (int)A.count(RT_SEC_LIGHTS);
	S.n_objs = (int)A.count(RT_SEC_OBJECTS);
	// k_wave_shade may drop a shadow ray whose Phong term is exactly 0 only if 0 * attenuation is 0,
	// i.e. the attenuation (a product of kt ^ distance) cannot be inf or NaN
	S.kt_unit = kt_all_unit(A.ptr<rt_material>(RT_SEC_MATERIALS), A.count(RT_SEC_MATERIALS));
	S.pad_ = 0;
	unsigned hflags[3];
	if (d_ntop) CK(cudaMemcpyAsync(&S.n_top, d_ntop, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(hflags, flags, sizeof(hflags), cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream)); // also: the caller may reuse `arena` on return
	if (hflags[0]) { ctx->error = "invalid arena: index out of range (device check " + std::to_string(hflags[0]) + ")"; return RT_ERR_ARG; }
	if ((hflags[1] | hflags[2]) & 0x80000000u || hflags[1] + hflags[2] > RT_STACK - 4) { ctx->error = "trees too deep for the traversal stack"; return RT_ERR_UNSUPPORTED; }
	if (ctx->budget_auto) ctx->walk_budget = 96;
	for (int k = 0; k < 32; k++) ctx->undecided[k] = 0.0;
	ctx->any_n[0] = ctx->any_n[1] = 0;
	ctx->n_area_lights = n_area;
	ctx->has_scene = true;
	return RT_OK;
}

// Device self-test: the layouts built by the device kernels of the last rt_upload_scene are
// compared byte for byte with the host construction (device/rt_derive.hpp) from the same arena.
rt_status rt_selftest_derive(rt_ctx* ctx, const void* arena, size_t bytes, unsigned long long* mismatches)
{
	if (!ctx || !mismatches) return RT_ERR_ARG;
	if (!ctx->has_scene) { ctx->error = "no scene uploaded"; return RT_ERR_NO_SCENE; }
	if (rt_arena_validate(arena, bytes) != RT_OK) { ctx->error = "invalid arena"; return RT_ERR_ARG; }
	CK(cudaSetDevice(ctx->device));
	rt::ArenaView A(arena);
	Derived D;
	derive(A, D);
	unsigned long long bad = 0;
	auto cmp = [&](const void* devp, const void* hostp, size_t n) -> rt_status {
		std::vector<uint8_t> tmp(n);
		if (n) CK(cudaMemcpy(tmp.data(), devp, n, cudaMemcpyDeviceToHost));
		for (size_t k = 0; k < n; k++) bad += tmp[k] != ((const uint8_t*)hostp)[k];
		return RT_OK;
	};
	rt_status st;
	if ((st = cmp(ctx->S.snodes2, D.snodes2.data(), D.snodes2.size() * sizeof(DNode2))) != RT_OK) return st;
	if ((st = cmp(ctx->S.mnodes2, D.mnodes2.data(), D.mnodes2.size() * sizeof(DNode2))) != RT_OK) return st;
	if ((st = cmp(ctx->S.snodes2f, D.snodes2f.data(), D.snodes2f.size() * sizeof(DNode2f))) != RT_OK) return st;
	if ((st = cmp(ctx->S.mnodes2f, D.mnodes2f.data(), D.mnodes2f.size() * sizeof(DNode2f))) != RT_OK) return st;
	if ((st = cmp(ctx->S.sboxf, D.sboxf.data(), D.sboxf.size() * sizeof(DItemBox))) != RT_OK) return st;
	if ((st = cmp(ctx->S.sprimf, D.sprimf.data(), D.sprimf.size() * sizeof(DPrimF))) != RT_OK) return st;
	if ((st = cmp(ctx->S.mtrees, D.mtrees.data(), D.mtrees.size() * sizeof(DTree))) != RT_OK) return st;
	if ((st = cmp(ctx->S.tris, D.tris.data(), D.tris.size() * sizeof(DTri))) != RT_OK) return st;
	if ((st = cmp(ctx->S.mesh_slack, D.mesh_slack.data(), D.mesh_slack.size() * sizeof(double))) != RT_OK) return st;
	bad += std::memcmp(&ctx->S.stree, &D.stree, sizeof(DTree)) != 0;
	bad += ctx->S.scene_slack != D.scene_slack;
	*mismatches = bad;
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
	ctx->spans.clear();
	ctx->ev_used = 0;
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
		// The frame is rendered in chunks of whole tiles: a sample index must fit 31 bits (the queues carry it in an int) and the
		// accumulators (48 B per sample) must fit the device - an 8K frame at 8x8 supersampling has 2.1 G samples.  The five
		// configurations are one chunk each.
		const long long spt = (long long)RT_TILE_PIX * A.spp_side * A.spp_side; // samples per tile
		long long chunk_tiles = ctx->chunk_samples / spt;
		if (chunk_tiles < 1) chunk_tiles = 1;
		for (long long t0 = 0; t0 < local_tiles; t0 += chunk_tiles) {
			const long long tiles_c = local_tiles - t0 < chunk_tiles ? local_tiles - t0 : chunk_tiles;
			const long long n_pix_c = tiles_c * RT_TILE_PIX, n_samples = tiles_c * spt;
			A.tile0 = t0;
			A.n_slots = n_samples * A.lenses;
			if ((st = prepare_accum(ctx, A, B_ACCUM, B_BASE, n_samples)) != RT_OK) return st;
			if ((st = launch_wave(ctx, A)) != RT_OK) return st;
			PoolArgs A2 = A;
			A2.accum = nullptr;
			A2.base = nullptr;
			if (p->anaglyph) { // second eye, shifted by ANAGLYPH_DELTA (RayTracer.cpp:23,92-99)
				if ((st = prepare_accum(ctx, A2, B_ACCUM2, B_BASE2, n_samples)) != RT_OK) return st;
				A2.eye_shift = 0.25;
				if ((st = launch_wave(ctx, A2)) != RT_OK) return st;
			}
			k_resolve_grid<<<blocks_for(n_pix_c, 256), 256, 0, ctx->stream>>>(A, A2.accum, A2.base, n_pix_c, fb);
			ctx->launches++;
			CK(cudaGetLastError());
		}
	} else {
		A.spp_side = 1;
		const int N = p->aa_samples * p->aa_samples;
		if ((st = ensure(ctx, B_PIXACC, (size_t)n_pix * RT_ACC_BYTES)) != RT_OK) return st;
		long long* pixacc = (long long*)ctx->buf[B_PIXACC];
		CK(cudaMemsetAsync(pixacc, 0, (size_t)n_pix * RT_ACC_BYTES, ctx->stream));
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
			if ((st = ensure(ctx, nxt, (size_t)n_regions * 4 * sizeof(Region))) != RT_OK) return st;
			k_adapt_samples<<<blocks_for(n_samples, 256), 256, 0, ctx->stream>>>((const Region*)ctx->buf[cur], n_regions, N, (double*)ctx->buf[B_XS], (double*)ctx->buf[B_YS]);
			ctx->launches++;
			PoolArgs L = A;
			L.src = SRC_XY;
			L.xs = (const double*)ctx->buf[B_XS];
			L.ys = (const double*)ctx->buf[B_YS];
			L.n_slots = n_samples * A.lenses;
			if ((st = prepare_accum(ctx, L, B_ACCUM, B_BASE, n_samples)) != RT_OK) return st;
			CK(cudaMemsetAsync(next_count, 0, 8, ctx->stream));
			if ((st = launch_wave(ctx, L)) != RT_OK) return st;
			k_adapt_resolve<<<blocks_for(n_regions, 256), 256, 0, ctx->stream>>>((const Region*)ctx->buf[cur], n_regions, N, L.accum, L.base, *p, pixacc,
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
	// only this partition's tiles cross the bus: packed on the device, one copy into page-locked staging, scattered into
	// the caller's frame (its other pixels are left untouched)
	const int tiles_x = (w + RT_TILE - 1) / RT_TILE, tiles_y = (h + RT_TILE - 1) / RT_TILE;
	const long long n_tiles = (long long)tiles_x * tiles_y;
	const long long n_local = (n_tiles - part + n_parts - 1) / n_parts;
	const size_t bytes = (size_t)n_local * RT_TILE_PIX * 3;
	if (n_local <= 0) return RT_OK;
	if ((st = ensure(ctx, B_PACKED, bytes)) != RT_OK) return st;
	if (ctx->stage_bytes < bytes) {
		if (ctx->stage) cudaFreeHost(ctx->stage);
		ctx->stage = nullptr;
		ctx->stage_bytes = 0;
		CK(cudaMallocHost((void**)&ctx->stage, bytes));
		ctx->stage_bytes = bytes;
	}
	k_pack_tiles<<<blocks_for(n_local * RT_TILE_PIX, 256), 256, 0, ctx->stream>>>((const uint8_t*)dev, w, h, tiles_x, part, n_parts, n_local, (uint8_t*)ctx->buf[B_PACKED]);
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(ctx->stage, ctx->buf[B_PACKED], bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	for (long long lt = 0; lt < n_local; lt++) {
		const long long t = lt * n_parts + part;
		const int x0 = (int)(t % tiles_x) * RT_TILE, y0 = (int)(t / tiles_x) * RT_TILE;
		const int x1 = x0 + RT_TILE < w ? x0 + RT_TILE : w, y1 = y0 + RT_TILE < h ? y0 + RT_TILE : h;
		for (int j = y0; j < y1; j++)
			std::memcpy(rgb_out + ((size_t)j * w + x0) * 3, ctx->stage + ((size_t)lt * RT_TILE_PIX + (size_t)(j - y0) * RT_TILE) * 3, (size_t)(x1 - x0) * 3);
	}
	return RT_OK;
}

rt_status rt_pack_tiles(rt_ctx* ctx, const void* dev_rgb, int w, int h, int part, int n_parts, void* dev_packed)
{
	if (!ctx) return RT_ERR_ARG;
	if (!dev_rgb || !dev_packed || w <= 0 || h <= 0 || n_parts <= 0 || part < 0 || part >= n_parts) { ctx->error = "bad arguments"; return RT_ERR_ARG; }
	CK(cudaSetDevice(ctx->device));
	const int tiles_x = (w + RT_TILE - 1) / RT_TILE, tiles_y = (h + RT_TILE - 1) / RT_TILE;
	const long long n_local = ((long long)tiles_x * tiles_y - part + n_parts - 1) / n_parts;
	if (n_local > 0) k_pack_tiles<<<blocks_for(n_local * RT_TILE_PIX, 256), 256, 0, ctx->stream>>>((const uint8_t*)dev_rgb, w, h, tiles_x, part, n_parts, n_local, (uint8_t*)dev_packed);
	CK(cudaGetLastError());
	CK(cudaStreamSynchronize(ctx->stream));
	return RT_OK;
}

rt_status rt_unpack_tiles(rt_ctx* ctx, const void* dev_gathered, int w, int h, int n_parts, long long tiles_cap, void* dev_rgb_out)
{
	if (!ctx) return RT_ERR_ARG;
	const long long need = ((long long)((w + RT_TILE - 1) / RT_TILE) * ((h + RT_TILE - 1) / RT_TILE) + n_parts - 1) / (n_parts > 0 ? n_parts : 1);
	if (!dev_gathered || !dev_rgb_out || w <= 0 || h <= 0 || n_parts <= 0 || tiles_cap < need) { ctx->error = "bad arguments"; return RT_ERR_ARG; }
	CK(cudaSetDevice(ctx->device));
	k_unpack_tiles<<<blocks_for((long long)w * h, 256), 256, 0, ctx->stream>>>((const uint8_t*)dev_gathered, w, h, (w + RT_TILE - 1) / RT_TILE, n_parts, tiles_cap, (uint8_t*)dev_rgb_out);
	CK(cudaGetLastError());
	CK(cudaStreamSynchronize(ctx->stream));
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
	ctx->spans.clear();
	ctx->ev_used = 0;
	// RayTracer::tracePixel's window coordinates (RayTracer.cpp:87-88)
	const int s = (p->aa_mode != RT_AA_NONE && p->aa_mode != RT_AA_ADAPTIVE) ? p->aa_samples : 1;
	std::vector<double> xs(n), ys(n);
	for (int k = 0; k < n; k++) {
		xs[k] = double(ij[2 * k]) / (double(w) * s);
		ys[k] = double(ij[2 * k + 1]) / (double(h) * s);
	}
	if ((st = ensure(ctx, B_XS, (size_t)n * 8)) != RT_OK) return st;
	if ((st = ensure(ctx, B_YS, (size_t)n * 8)) != RT_OK) return st;
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
	A.n_slots = (long long)n * A.lenses;
	if ((st = prepare_accum(ctx, A, B_ACCUM, B_BASE, n)) != RT_OK) return st;
	if ((st = launch_wave(ctx, A)) != RT_OK) return st;
	PoolArgs A2 = A;
	A2.accum = nullptr;
	A2.base = nullptr;
	if (p->anaglyph) {
		if ((st = prepare_accum(ctx, A2, B_ACCUM2, B_BASE2, n)) != RT_OK) return st;
		A2.eye_shift = 0.25;
		if ((st = launch_wave(ctx, A2)) != RT_OK) return st;
	}
	k_sample_values<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(A.accum, A.base, A2.accum, A2.base, n, *p, (double*)ctx->buf[B_PIXACC]);
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
	ctx->spans.clear();
	ctx->ev_used = 0;
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
		PoolArgs A;
		std::memset(&A, 0, sizeof(A));
		A.prm = *p;
		A.src = SRC_RAYS;
		A.spp_side = 1;
		A.lenses = 1;
		A.n_parts = 1;
		A.depth = depth;
		A.rays = (const rt_ray*)ctx->buf[B_RAYS];
		A.n_slots = n;
		if ((st = prepare_accum(ctx, A, B_ACCUM, B_BASE, n)) != RT_OK) return st;
		if ((st = launch_wave(ctx, A)) != RT_OK) return st;
		if ((st = ensure(ctx, B_PIXACC, (size_t)n * 24)) != RT_OK) return st;
		k_accum_values<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(A.accum, A.base, n, (double*)ctx->buf[B_PIXACC]);
		ctx->launches++;
		CK(cudaGetLastError());
		CK(cudaMemcpyAsync(rgb_out, ctx->buf[B_PIXACC], (size_t)n * 24, cudaMemcpyDeviceToHost, ctx->stream));
	}
	return read_stats(ctx, nullptr, 0.f);
}


// ---- rt_intersect_static: Scene::intersect for a small batch of rays WITHOUT any device allocation ------------
// (no counterpart in the reference; added for callers that cannot call
// cudaMalloc: code under CUDA graph capture, tight serving loops.)  What rt_trace_rays stages in cudaMalloc'ed
// work buffers - the rays on their way in, the hit records on their way out - lives in ONE file-scope device
// array of a fixed size instead.  Every GPU that loads the library holds these RT_STATIC_SCRATCH_BYTES until exit.
//   layout: [0, MAX * sizeof(rt_ray))  rays     [MAX * sizeof(rt_ray), ... + MAX * sizeof(rt_hit))  hit records
// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is asked
// for once per context (= once per device, with that device current) with cudaGetSymbolAddress.
// All calls share the array and it keeps the last call's contents: the hit-record region is cleared at the start
// of each call, and two calls that overlap on one GPU (two contexts / threads on the same device) RACE.
#define RT_STATIC_SCRATCH_BYTES ((size_t)RT_STATIC_MAX_RAYS * (sizeof(rt_ray) + sizeof(rt_hit)))
static __device__ __align__(256) unsigned char g_static_scratch[RT_STATIC_SCRATCH_BYTES];

// k_hits (rt_kernels.cuh) over the static array: rays at its start, hit records behind them
__global__ void __launch_bounds__(128) k_hits_static(const __grid_constant__ DScene S, unsigned char* scratch, int n)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	const rt_ray* rays = (const rt_ray*)scratch;
	rt_hit* out = (rt_hit*)(scratch + (size_t)RT_STATIC_MAX_RAYS * sizeof(rt_ray));
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

rt_status rt_intersect_static(rt_ctx* ctx, const rt_ray* rays, int n, rt_hit* hits_out)
{
	if (!ctx) return RT_ERR_ARG;
	if (!ctx->has_scene) { ctx->error = "no scene uploaded"; return RT_ERR_NO_SCENE; }
	if (n < 0 || (n && (!rays || !hits_out))) { ctx->error = "bad arguments"; return RT_ERR_ARG; }
	if (n > RT_STATIC_MAX_RAYS) {
		ctx->error = "rt_intersect_static: n exceeds RT_STATIC_MAX_RAYS (" + std::to_string(RT_STATIC_MAX_RAYS) + " rays: the static device scratch is fixed-size; use rt_trace_rays)";
		return RT_ERR_ARG;
	}
	if (n == 0) return RT_OK;
	CK(cudaSetDevice(ctx->device));
	if (!ctx->static_scratch) { // set-up, once per context: the symbol's address on THIS device
		void* p = nullptr;
		CK(cudaGetSymbolAddress(&p, g_static_scratch));
		ctx->static_scratch = p;
	}
	unsigned char* scratch = (unsigned char*)ctx->static_scratch;
	unsigned char* hits_dev = scratch + (size_t)RT_STATIC_MAX_RAYS * sizeof(rt_ray);
	CK(cudaMemsetAsync(hits_dev, 0, (size_t)n * sizeof(rt_hit), ctx->stream)); // the array keeps the previous call's records
	CK(cudaMemcpyAsync(scratch, rays, (size_t)n * sizeof(rt_ray), cudaMemcpyHostToDevice, ctx->stream));
	k_hits_static<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(ctx->S, scratch, n);
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(hits_out, hits_dev, (size_t)n * sizeof(rt_hit), cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	return RT_OK;
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


rt_status rt_measure_fma_peak(rt_ctx* ctx, double* fp64_tflops, double* fp32_tflops)
{
	if (!ctx || !fp64_tflops || !fp32_tflops) return RT_ERR_ARG;
	CK(cudaSetDevice(ctx->device));
	const int blocks = ctx->sm_count * 8, iters = 1 << 14;
	rt_status st;
	if ((st = ensure(ctx, B_TMP_A, (size_t)blocks * 256 * sizeof(double))) != RT_OK) return st;
	double best[2] = {0.0, 0.0};
	for (int rep = 0; rep < 4; rep++)
		for (int k = 0; k < 2; k++) {
			CK(cudaEventRecord(ctx->ev0, ctx->stream));
			if (k == 0) k_fma_peak<double><<<blocks, 256, 0, ctx->stream>>>(iters, 1.0, (double*)ctx->buf[B_TMP_A]);
			else k_fma_peak<float><<<blocks, 256, 0, ctx->stream>>>(iters, 1.0f, (float*)ctx->buf[B_TMP_A]);
			CK(cudaEventRecord(ctx->ev1, ctx->stream));
			CK(cudaEventSynchronize(ctx->ev1));
			float ms = 0.f;
			CK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
			const double tf = 2.0 * 8.0 * iters * (double)blocks * 256 / (ms * 1e-3) / 1e12;
			if (rep > 0 && tf > best[k]) best[k] = tf; // first pass warms up
		}
	*fp64_tflops = best[0];
	*fp32_tflops = best[1];
	return RT_OK;
}

} // extern "C"
