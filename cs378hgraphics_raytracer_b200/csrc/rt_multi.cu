// rt_multi.cu - N GPUs of one box behind the C ABI (include/rt_b200.h: rt_multi_*), in ONE process.
//
// The reference's parallel driver sits inside RayTracer::traceImage (an OpenMP loop over pixels, reference
// ray/src/RayTracer.cpp:283-285); its replacement therefore has to be reachable from the C++ host that keeps the
// RayTracer API, not only from a launcher that starts one process per GPU.  SURVEY.md 8(b)/(e):
//   * one rt_ctx per device, one host thread per device for the duration of a call;
//   * scene REPLICATED: one host -> device copy on the first device, ncclBroadcast of the arena bytes to the others
//     straight into each context's scene buffer (rt_scene_buffer / rt_upload_scene_resident);
//   * image partitioned by interleaved 32 x 32 tiles, tile t -> device t mod N (rt_render_device(part, n_parts));
//   * frame gather: every device packs its tiles (rt_pack_tiles), grouped ncclSend / ncclRecv to the first device only,
//     one un-interleave kernel there (rt_unpack_tiles), one device -> host copy.
// No collective inside the trace loop: every pixel is independent.
// NCCL is loaded at run time (dlopen), so librt_b200.so has no link-time dependency on it and a single-GPU user never
// needs it; in a process that already holds a libnccl.so.2 (PyTorch's) that copy is the one used.
#include <dlfcn.h>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <cuda_runtime.h>
#include "rt_b200.h"

namespace {

// the few NCCL declarations needed (nccl.h's ABI: opaque communicator, int enums)
typedef struct ncclComm* ncclComm_t;
typedef int ncclResult_t;
enum { ncclSuccess_ = 0, ncclUint8_ = 1 };
struct Nccl {
	void* lib = nullptr;
	ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char* (*GetErrorString)(ncclResult_t) = nullptr;
	bool load(std::string& err)
	{
		if (lib) return true;
		const char* names[] = {"libnccl.so.2", "libnccl.so"};
		for (const char* n : names)
			if ((lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
		if (!lib) { err = std::string("NCCL not found: ") + dlerror(); return false; }
#define SYM(field, name) *(void**)(&field) = dlsym(lib, name); if (!field) { err = std::string("NCCL symbol missing: ") + name; return false; }
		SYM(CommInitAll, "ncclCommInitAll") SYM(CommDestroy, "ncclCommDestroy") SYM(Broadcast, "ncclBroadcast") SYM(Send, "ncclSend")
		SYM(Recv, "ncclRecv") SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
		return true;
	}
};
Nccl g_nccl;

} // namespace

struct rt_multi {
	int n = 0;
	std::vector<int> devices;
	std::vector<rt_ctx*> ctx;
	std::vector<ncclComm_t> comm;       // empty when n == 1
	std::vector<cudaStream_t> stream;   // per device: NCCL operations and the copies around them
	std::vector<void*> packed;          // per device: its packed tiles
	std::vector<size_t> packed_bytes;
	void* gathered = nullptr;           // device 0: n blocks of packed tiles
	size_t gathered_bytes = 0;
	void* frame = nullptr;              // device 0: the un-interleaved frame
	size_t frame_bytes = 0;
	std::string error;
};

namespace {

#define MCK(m, call)                                                                        \
	do {                                                                                    \
		cudaError_t e_ = (call);                                                            \
		if (e_ != cudaSuccess) { (m) = std::string(#call) + ": " + cudaGetErrorString(e_); return RT_ERR_CUDA; } \
	} while (0)
#define NCK(m, call)                                                                        \
	do {                                                                                    \
		ncclResult_t r_ = (call);                                                           \
		if (r_ != ncclSuccess_) { (m) = std::string(#call) + ": " + g_nccl.GetErrorString(r_); return RT_ERR_CUDA; } \
	} while (0)

rt_status grow(std::string& err, void** p, size_t* have, size_t need)
{
	if (*have >= need) return RT_OK;
	if (*p) cudaFree(*p);
	*p = nullptr;
	*have = 0;
	MCK(err, cudaMalloc(p, need));
	*have = need;
	return RT_OK;
}

// runs f(r) for every device on its own host thread; returns the first failing status
template <typename F> rt_status on_all(rt_multi* m, F f, std::vector<std::string>& errs)
{
	std::vector<rt_status> st(m->n, RT_OK);
	errs.assign(m->n, std::string());
	if (m->n == 1) st[0] = f(0, errs[0]);
	else {
		std::vector<std::thread> th;
		for (int r = 0; r < m->n; r++) th.emplace_back([&, r] { st[r] = f(r, errs[r]); });
		for (auto& t : th) t.join();
	}
	for (int r = 0; r < m->n; r++)
		if (st[r] != RT_OK) { m->error = "device " + std::to_string(m->devices[r]) + ": " + errs[r]; return st[r]; }
	return RT_OK;
}

} // namespace

extern "C" {

rt_status rt_multi_create(const int* devices, int n_dev, rt_multi** out)
{
	if (!out) return RT_ERR_ARG;
	*out = nullptr;
	if (!devices || n_dev < 1 || n_dev > 64) return RT_ERR_ARG;
	int have = 0;
	if (cudaGetDeviceCount(&have) != cudaSuccess || have <= 0) return RT_ERR_NO_DEVICE;
	for (int a = 0; a < n_dev; a++) {
		if (devices[a] < 0 || devices[a] >= have) return RT_ERR_NO_DEVICE;
		for (int b = 0; b < a; b++)
			if (devices[a] == devices[b]) return RT_ERR_ARG; // one rank per GPU
	}
	rt_multi* m = new rt_multi();
	m->n = n_dev;
	m->devices.assign(devices, devices + n_dev);
	m->ctx.assign(n_dev, nullptr);
	m->stream.assign(n_dev, nullptr);
	m->packed.assign(n_dev, nullptr);
	m->packed_bytes.assign(n_dev, 0);
	for (int r = 0; r < n_dev; r++) {
		const rt_status st = rt_create(devices[r], &m->ctx[r]);
		if (st != RT_OK) { rt_multi_destroy(m); return st; }
		if (cudaSetDevice(devices[r]) != cudaSuccess || cudaStreamCreateWithFlags(&m->stream[r], cudaStreamNonBlocking) != cudaSuccess) { rt_multi_destroy(m); return RT_ERR_CUDA; }
	}
	if (n_dev > 1) {
		std::string err;
		if (!g_nccl.load(err)) { std::fprintf(stderr, "rt_multi_create: %s\n", err.c_str()); rt_multi_destroy(m); return RT_ERR_UNSUPPORTED; }
		m->comm.assign(n_dev, nullptr);
		if (g_nccl.CommInitAll(m->comm.data(), n_dev, devices) != ncclSuccess_) { m->comm.clear(); rt_multi_destroy(m); return RT_ERR_CUDA; }
	}
	*out = m;
	return RT_OK;
}

void rt_multi_destroy(rt_multi* m)
{
	if (!m) return;
	for (int r = 0; r < m->n; r++) {
		cudaSetDevice(m->devices[r]);
		if (r < (int)m->comm.size() && m->comm[r]) g_nccl.CommDestroy(m->comm[r]);
		if (m->packed[r]) cudaFree(m->packed[r]);
		if (r == 0) { if (m->gathered) cudaFree(m->gathered); if (m->frame) cudaFree(m->frame); }
		if (m->stream[r]) cudaStreamDestroy(m->stream[r]);
		if (m->ctx[r]) rt_destroy(m->ctx[r]);
	}
	delete m;
}

const char* rt_multi_last_error(const rt_multi* m) { return m ? m->error.c_str() : "null context"; }
int rt_multi_devices(const rt_multi* m) { return m ? m->n : 0; }
rt_ctx* rt_multi_context(rt_multi* m, int r) { return (m && r >= 0 && r < m->n) ? m->ctx[r] : nullptr; }

rt_status rt_multi_upload_scene(rt_multi* m, const void* arena, size_t bytes)
{
	if (!m) return RT_ERR_ARG;
	if (!arena || bytes == 0) { m->error = "null arena"; return RT_ERR_ARG; }
	if (m->n == 1) {
		const rt_status st = rt_upload_scene(m->ctx[0], arena, bytes);
		if (st != RT_OK) m->error = rt_last_error(m->ctx[0]);
		return st;
	}
	std::vector<std::string> errs;
	return on_all(m, [&](int r, std::string& err) -> rt_status {
		MCK(err, cudaSetDevice(m->devices[r]));
		void* dev = nullptr;
		rt_status st = rt_scene_buffer(m->ctx[r], bytes, &dev);
		if (st != RT_OK) { err = rt_last_error(m->ctx[r]); return st; }
		// one host -> device copy (root), then the bytes travel over NVLink
		if (r == 0) MCK(err, cudaMemcpyAsync(dev, arena, bytes, cudaMemcpyHostToDevice, m->stream[r]));
		NCK(err, g_nccl.Broadcast(dev, dev, bytes, ncclUint8_, 0, m->comm[r], m->stream[r]));
		MCK(err, cudaStreamSynchronize(m->stream[r]));
		st = rt_upload_scene_resident(m->ctx[r], arena, bytes);
		if (st != RT_OK) err = rt_last_error(m->ctx[r]);
		return st;
	}, errs);
}

rt_status rt_multi_render(rt_multi* m, const rt_params* p, int w, int h, uint8_t* rgb_out, rt_stats* stats)
{
	if (!m) return RT_ERR_ARG;
	if (!rgb_out || w <= 0 || h <= 0) { m->error = "bad frame arguments"; return RT_ERR_ARG; }
	if (m->n == 1) {
		const rt_status st = rt_render(m->ctx[0], p, w, h, rgb_out, stats);
		if (st != RT_OK) m->error = rt_last_error(m->ctx[0]);
		return st;
	}
	const int n = m->n;
	const long long tiles = (long long)((w + 31) / 32) * ((h + 31) / 32);
	const long long cap = (tiles + n - 1) / n;         // packed tiles per device (padded to the same count)
	const size_t block = (size_t)cap * 32 * 32 * 3;
	std::vector<rt_stats> st_dev(n);
	std::vector<std::string> errs;
	const rt_status rs = on_all(m, [&](int r, std::string& err) -> rt_status {
		MCK(err, cudaSetDevice(m->devices[r]));
		rt_status st = grow(err, &m->packed[r], &m->packed_bytes[r], block);
		if (st != RT_OK) return st;
		if (r == 0) {
			if ((st = grow(err, &m->gathered, &m->gathered_bytes, block * n)) != RT_OK) return st;
			if ((st = grow(err, &m->frame, &m->frame_bytes, (size_t)w * h * 3)) != RT_OK) return st;
		}
		void* fb = nullptr;
		st = rt_render_device(m->ctx[r], p, w, h, r, n, &fb, &st_dev[r]);
		if (st == RT_OK) st = rt_pack_tiles(m->ctx[r], fb, w, h, r, n, r == 0 ? m->gathered : m->packed[r]);
		if (st != RT_OK) { err = rt_last_error(m->ctx[r]); return st; }
		// the gather: everyone sends its block to device 0
		if (r == 0) {
			NCK(err, g_nccl.GroupStart());
			for (int q = 1; q < n; q++) NCK(err, g_nccl.Recv((uint8_t*)m->gathered + (size_t)q * block, block, ncclUint8_, q, m->comm[0], m->stream[0]));
			NCK(err, g_nccl.GroupEnd());
		} else {
			NCK(err, g_nccl.Send(m->packed[r], block, ncclUint8_, 0, m->comm[r], m->stream[r]));
		}
		MCK(err, cudaStreamSynchronize(m->stream[r]));
		if (r == 0) {
			st = rt_unpack_tiles(m->ctx[0], m->gathered, w, h, n, cap, m->frame);
			if (st != RT_OK) { err = rt_last_error(m->ctx[0]); return st; }
			MCK(err, cudaMemcpyAsync(rgb_out, m->frame, (size_t)w * h * 3, cudaMemcpyDeviceToHost, m->stream[0]));
			MCK(err, cudaStreamSynchronize(m->stream[0]));
		}
		return RT_OK;
	}, errs);
	if (rs != RT_OK) return rs;
	if (stats) {
		*stats = st_dev[0];
		for (int r = 1; r < n; r++) {
			stats->camera_rays += st_dev[r].camera_rays;
			stats->reflection_rays += st_dev[r].reflection_rays;
			stats->refraction_rays += st_dev[r].refraction_rays;
			stats->shadow_rays += st_dev[r].shadow_rays;
			stats->culled_shadow_rays += st_dev[r].culled_shadow_rays;
			stats->cooperative_rays += st_dev[r].cooperative_rays;
			stats->launches += st_dev[r].launches;
			if (st_dev[r].device_ms > stats->device_ms) stats->device_ms = st_dev[r].device_ms; // the frame is done when the slowest device is
			if (st_dev[r].rounds > stats->rounds) stats->rounds = st_dev[r].rounds;
		}
	}
	return RT_OK;
}

} // extern "C"
