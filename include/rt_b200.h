/* rt_b200.h - C-ABI of the B200 device path for the CS378H ray tracer's per-pixel trace loop.
 *
 * This is the drop-in boundary.  The reference has no FFI layer: its callers use
 * the C++ class RayTracer (reference ray/src/RayTracer.h:26-75).  Every entry point below names
 * the RayTracer member (or the piece of scene loading) it replaces; the host class
 * cs378hgraphics_raytracer_b200/csrc/host/RayTracer.h keeps the reference's member names and
 * forwards to these functions.  Plain pointers and sizes only; no C++/torch types; nothing
 * throws across this boundary.  All calls are synchronous: results are complete on return
 * (the reference's checkRender() is `return true`, RayTracer.cpp:372-375).
 *
 * There is NO CPU fallback behind this header: without a CUDA device rt_create fails.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>
#include "rt_arena.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t rt_status;
enum {
	RT_OK = 0,
	RT_ERR_ARG = 1,      /* null pointer, bad size, bad arena */
	RT_ERR_CUDA = 2,     /* a CUDA runtime call failed (message in rt_last_error) */
	RT_ERR_NO_DEVICE = 3,/* no usable CUDA device: there is no CPU path */
	RT_ERR_NO_SCENE = 4, /* render/trace before rt_upload_scene */
	RT_ERR_IO = 5,       /* scene / texture file could not be read */
	RT_ERR_PARSE = 6,    /* .ray syntax or semantic error (message in rt_host_last_error) */
	RT_ERR_OVERFLOW = 7, /* a device work queue overflowed (never expected; reported, not hidden) */
	RT_ERR_UNSUPPORTED = 8
};

/* Render settings: the POD image of the reference's global `traceUI` getters that the hot loop
 * reads (reference ray/src/ui/TraceUI.h:35-129; set by CommandLineUI.cpp:30-131). */
enum { RT_AA_NONE = 0, RT_AA_SUPERSAMPLE = 1, RT_AA_ADAPTIVE = 2, RT_AA_JITTERED = 3 };
typedef struct rt_params {
	int32_t depth;               /* getDepth(): -r, default 0 */
	int32_t aa_mode;             /* getAAMode(): -O r|a|j */
	int32_t aa_samples;          /* getAASamples(): -A after -O r|a|j, default 3 */
	int32_t dof_on;              /* dofSwitch(): -O d */
	int32_t dof_divs;            /* getDofSubDiv(): -B after -O d, default 5 */
	int32_t anaglyph;            /* anaglyph(): -O g */
	int32_t overlapping_objects; /* overlappingObjects(): -O o */
	int32_t use_cubemap;         /* cubeMap(): needs a cube map in the arena as well */
	int32_t soft_shadow_res;     /* softShadowRes(): -O s -A n, default 5 */
	int32_t pad;
	double aterm_thresh;         /* getATermThresh(): -O c -A thr; aTermSwitch() is thresh > 0 */
	double aa_thresh;            /* getAAThresh(): -B after -O a, default 1.0 */
	double dof_fd;               /* getDofFD(): default 3 */
	double dof_apsz;             /* getDofApSz(): default 0.05 */
} rt_params;

/* Fills the reference's defaults (TraceUI.h:37-129). */
void rt_params_default(rt_params* p);

typedef struct rt_ray {
	double p[3];
	double d[3];
} rt_ray;

/* What Scene::intersect leaves in an isect (reference scene.cpp:157-180), with the object
 * pointer replaced by indices: obj = position in Scene::objects, face = position in
 * Trimesh::faces (non-degenerate faces only) or -1; a miss is obj = face = -1, t = 1000.0. */
typedef struct rt_hit {
	int32_t obj;
	int32_t face;
	double t;
	double N[3];
} rt_hit;

/* Ray counters of one call: camera = Scene::intersect calls made for eye / DOF
 * lens / AA sample rays, reflection + refraction = the recursive ones, shadow = Scene::
 * intersectList calls made by Light::srsAttenuation. */
typedef struct rt_stats {
	uint64_t camera_rays;
	uint64_t reflection_rays;
	uint64_t refraction_rays;
	uint64_t shadow_rays;
	double device_ms;  /* CUDA-event time of the device work of this call */
	uint64_t launches; /* kernels launched by this call */
	uint64_t cooperative_rays; /* of the above, rays whose tree walk was finished by a whole warp
	                              (they exhausted the single-thread node budget) */
	uint64_t rounds;   /* wavefront rounds of this call */
	/* per-stage device time (CUDA events around every launch) and launch counts; filled only
	 * after rt_set_profiling(ctx, 1).  Stages: RT_STAGE_* below. */
	double stage_ms[6];
	uint64_t stage_launches[6];
	uint64_t culled_shadow_rays; /* of shadow_rays: counted as the reference counts them but not walked, because
	                                their Phong term (light.cpp's caller, material.cpp:50-65) is exactly 0 */
} rt_stats;
enum { RT_STAGE_GEN = 0, RT_STAGE_CLOSEST = 1, RT_STAGE_COOP = 2, RT_STAGE_SHADE = 3, RT_STAGE_SHADOW = 4, RT_STAGE_OTHER = 5 };

typedef struct rt_ctx rt_ctx;

/* ---- device context ------------------------------------------------------------------ */
/* Creates a context on CUDA device `device` (one context = one GPU = one host thread, like one
 * RayTracer instance). */
rt_status rt_create(int device, rt_ctx** out);
void rt_destroy(rt_ctx* ctx);
const char* rt_last_error(const rt_ctx* ctx);
/* on != 0: bracket every kernel launch of the trace loop with CUDA events and report per-stage
 * times in rt_stats (a few microseconds per launch; off by default). */
rt_status rt_set_profiling(rt_ctx* ctx, int on);

/* Replaces what RayTracer::loadScene leaves behind for the tracer (RayTracer.cpp:196-240): the
 * arena (include/rt_arena.h) is copied to HBM in one transfer (fastest from page-locked memory),
 * validated and re-laid-out for traversal by device kernels.  The caller keeps ownership of
 * `arena` and may reuse it on return. */
rt_status rt_upload_scene(rt_ctx* ctx, const void* arena, size_t bytes);

/* RayTracer::traceImage(w,h) + getBuffer (RayTracer.cpp:279-314,184-189): renders the whole
 * frame into rgb_out (host memory, w*h*3 bytes, row 0 = bottom scan line like RayTracer::buffer).
 * rt_render_tiles renders only the 32x32-pixel tiles t with t % n_parts == part (interleaved
 * tile partition used for multi-GPU); other pixels of rgb_out are left untouched. */
rt_status rt_render(rt_ctx* ctx, const rt_params* p, int w, int h, uint8_t* rgb_out, rt_stats* stats);
rt_status rt_render_tiles(rt_ctx* ctx, const rt_params* p, int w, int h, int part, int n_parts,
                          uint8_t* rgb_out, rt_stats* stats);
/* Same work with the frame left in HBM: *dev_rgb receives a device pointer owned by the context
 * (valid until the next render call).  Used by the multi-GPU gather and by device-resident timing. */
rt_status rt_render_device(rt_ctx* ctx, const rt_params* p, int w, int h, int part, int n_parts,
                           void** dev_rgb, rt_stats* stats);

/* Framebuffer gather of the tile partition (the reference's pixel loop has no counterpart - its OpenMP
 * threads share one buffer, RayTracer.cpp:283-313).  rt_pack_tiles: a rank's tiles of the device frame `dev_rgb` as one
 * contiguous device block `dev_packed` of ceil((tiles - part) / n_parts) * 3072 bytes (32 x 32 x RGB per tile, tile order)
 * - what is sent to the gathering rank.  rt_unpack_tiles: the inverse on the gathering rank; dev_gathered holds n_parts
 * blocks of tiles_cap packed tiles (rank r's tiles in block r), dev_rgb_out receives the w*h*3 frame. */
rt_status rt_pack_tiles(rt_ctx* ctx, const void* dev_rgb, int w, int h, int part, int n_parts, void* dev_packed);
rt_status rt_unpack_tiles(rt_ctx* ctx, const void* dev_gathered, int w, int h, int n_parts, long long tiles_cap, void* dev_rgb_out);

/* The two halves of rt_upload_scene for a scene whose bytes reach the device by another road (rt_multi_upload_scene's NCCL
 * broadcast): rt_scene_buffer hands out the context's scene buffer (device pointer, >= bytes), rt_upload_scene_resident does
 * all of rt_upload_scene except the host -> device copy; `arena` = the same bytes in host memory (small tables are read there). */
rt_status rt_scene_buffer(rt_ctx* ctx, size_t bytes, void** dev);
rt_status rt_upload_scene_resident(rt_ctx* ctx, const void* arena, size_t bytes);

/* ---- N GPUs of one box in one process -----------------------------------------------------------
 * Replaces the reference's parallel driver INSIDE RayTracer::traceImage (the OpenMP pixel loop, RayTracer.cpp:283-285):
 * one context per device, one host thread per device during a call, scene replicated (one host -> device copy + ncclBroadcast),
 * image split into interleaved 32x32 tiles (tile t -> device t mod N), packed tiles sent to the first device with grouped
 * ncclSend / ncclRecv and un-interleaved there.  NCCL is loaded at run time (dlopen): n_dev == 1 needs no NCCL at all.
 * rt_multi_render's stats are summed over the devices (device_ms: the slowest). */
typedef struct rt_multi rt_multi;
rt_status rt_multi_create(const int* devices, int n_dev, rt_multi** out);
void rt_multi_destroy(rt_multi* m);
const char* rt_multi_last_error(const rt_multi* m);
int rt_multi_devices(const rt_multi* m);
/* the r-th device's context (owned by m): rt_trace_pixels / rt_trace_rays of the replicated scene run on any one of them */
rt_ctx* rt_multi_context(rt_multi* m, int r);
rt_status rt_multi_upload_scene(rt_multi* m, const void* arena, size_t bytes);
rt_status rt_multi_render(rt_multi* m, const rt_params* p, int w, int h, uint8_t* rgb_out, rt_stats* stats);

/* RayTracer::tracePixel(i,j) for n pixels (RayTracer.cpp:81-102): rgb_out[3n] doubles, clamped
 * like the reference, not quantised.  ij = {i0,j0,i1,j1,...}. */
rt_status rt_trace_pixels(rt_ctx* ctx, const rt_params* p, int w, int h, const int32_t* ij, int n,
                          double* rgb_out);

/* RayTracer::traceRay(r, thresh, depth, t) for n rays (RayTracer.cpp:108-174): hits_out[n] is the
 * Scene::intersect result of each ray itself, rgb_out[3n] the returned colour (either may be NULL). */
rt_status rt_trace_rays(rt_ctx* ctx, const rt_params* p, const rt_ray* rays, int n, int depth,
                        rt_hit* hits_out, double* rgb_out);

/* Scene::intersect for n <= RT_STATIC_MAX_RAYS rays with NO device allocation.  This entry point has no
 * counterpart in the reference; it was added for callers that cannot
 * call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  hits_out[n] is what
 * rt_trace_rays(..., hits_out, NULL) returns.  The device staging of rays and hit records is a fixed file-scope
 * device array (RT_STATIC_MAX_RAYS * 88 bytes = 352 KiB, held by every GPU that loads the library); the call
 * makes no cudaMalloc / cudaFree.  n > RT_STATIC_MAX_RAYS is refused with RT_ERR_ARG.  All calls on a GPU share
 * that array: calls that overlap on one GPU (two contexts or threads on the same device) race. */
#define RT_STATIC_MAX_RAYS 4096
rt_status rt_intersect_static(rt_ctx* ctx, const rt_ray* rays, int n, rt_hit* hits_out);

/* Device self-test: n pseudo-random quotients a/d computed by the slab test's FMA-corrected
 * reciprocal division (device/rt_math.cuh div_by) are compared bit for bit with IEEE division;
 * exponents of a and d are drawn from [-exp_range, exp_range].  *mismatches must come back 0. */
rt_status rt_selftest_division(rt_ctx* ctx, long long n, unsigned long long seed, int exp_range, unsigned long long* mismatches);

/* Device self-test: the traversal layouts that rt_upload_scene built on the device from `arena`
 * (the arena of the last upload) are compared byte for byte with the host construction. */
rt_status rt_selftest_derive(rt_ctx* ctx, const void* arena, size_t bytes, unsigned long long* mismatches);

/* FP64 and FP32 FMA throughput of this GPU (TFLOP/s, FMA = 2), measured with a dependent-chain
 * microbenchmark: the compute ceilings bench.py's roofline quotes next to the HBM one. */
rt_status rt_measure_fma_peak(rt_ctx* ctx, double* fp64_tflops, double* fp32_tflops);

/* ---- host-side scene loading (no GPU needed) -------------------------------------------- */
/* Replaces Tokenizer + Parser::parseScene + Scene::add/conclude + Trimesh::conclude (reference
 * parser/Parser.cpp:26-95, scene.cpp:133-153, trimesh.cpp:69-77, kdTree.h:23-78): parses a .ray
 * file, builds the reference's own trees and emits the arena.  cubemap_faces is NULL or six image
 * paths in +x -x +y -y +z -z order (what TraceUI::smartLoadCubemap resolves, TraceUI.cc:148-167).
 * The arena is malloc'ed; free it with rt_arena_free. */
rt_status rt_load_ray_file(const char* path, const char* const* cubemap_faces, void** arena_out,
                           size_t* bytes_out);
void rt_arena_free(void* arena);
/* Message of the last failed rt_load_ray_file on this thread. */
const char* rt_host_last_error(void);
/* Checks magic, version, section table and index ranges of an arena. */
rt_status rt_arena_validate(const void* arena, size_t bytes);

/* Cached binary scene: the arena is one position-independent block, so the cache
 * file IS the arena.  rt_arena_load reads the file into a malloc'ed block (rt_arena_free) and runs
 * rt_arena_validate on it; a file of another arena version, a truncated or a corrupted one is refused. */
rt_status rt_arena_save(const char* path, const void* arena, size_t bytes);
rt_status rt_arena_load(const char* path, void** arena_out, size_t* bytes_out);

/* ---- output stage (no GPU needed) -------------------------------------------------------- */
/* Replaces writeImage (reference fileio/images.cc:58-66 -> writePNG pngimage.cpp:226-285, writeBMP
 * bitmap.cpp:95-140): chosen by extension, case-insensitive, ".png" = 8-bit RGB PNG, anything else =
 * 24-bit BMP.  rgb is the frame as rt_render returns it (w*h*3, row 0 = bottom scan line). */
rt_status rt_write_image(const char* path, int w, int h, const uint8_t* rgb);
/* Replaces readImage (images.cc:48-55): w*h*3 bytes, row 0 = bottom scan line, malloc'ed (free with
 * rt_arena_free).  .bmp (24 bit) and .png (8-bit RGB, non-interlaced, no gAMA chunk) are read. */
rt_status rt_read_image(const char* path, int* w, int* h, uint8_t** rgb_out);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
