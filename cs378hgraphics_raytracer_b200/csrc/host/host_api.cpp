// host_api.cpp - the host-only half of the C-ABI (include/rt_b200.h): scene loading, arena
// validation, default settings.  No CUDA here; the device half is ../rt_api.cu.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include "ray_scene.hpp"
#include "rt_arena_writer.hpp"
#include "rt_b200.h"

static thread_local std::string g_host_error;

extern "C" {

void rt_params_default(rt_params* p)
{
	if (!p) return;
	std::memset(p, 0, sizeof(*p));
	p->depth = 0;            // TraceUI.h:41
	p->aa_mode = RT_AA_NONE; // :60
	p->aa_samples = 3;       // :61
	p->aa_thresh = 1.0;      // :62
	p->aterm_thresh = 0.0;   // :51
	p->dof_on = 0;           // :110
	p->dof_apsz = 0.05;      // :111
	p->dof_divs = 5;         // :112
	p->dof_fd = 3.0;         // :113
	p->soft_shadow_res = 5;  // :129
}

rt_status rt_load_ray_file(const char* path, const char* const* cubemap_faces, void** arena_out, size_t* bytes_out)
{
	if (!path || !arena_out || !bytes_out) { g_host_error = "null argument"; return RT_ERR_ARG; }
	*arena_out = nullptr;
	*bytes_out = 0;
	try {
		rs::Scene sc;
		if (cubemap_faces) {
			for (int k = 0; k < 6; k++) {
				if (!cubemap_faces[k]) { g_host_error = "cubemap needs six face files"; return RT_ERR_ARG; }
				sc.cubemap[k] = rs::load_texture(sc, cubemap_faces[k]);
			}
			sc.has_cubemap = true;
		}
		rs::parse_ray_file(path, sc);
		if (sc.objects.empty()) { g_host_error = "scene has no geometry"; return RT_ERR_PARSE; }
		void* a = rs::flatten(sc, bytes_out);
		if (!a) { g_host_error = "out of memory"; return RT_ERR_IO; }
		*arena_out = a;
		return RT_OK;
	} catch (const std::exception& e) {
		g_host_error = e.what();
		const bool io = g_host_error.find("couldn't read") != std::string::npos || g_host_error.find("Unable to load") != std::string::npos;
		return io ? RT_ERR_IO : RT_ERR_PARSE;
	} catch (...) {
		g_host_error = "unknown error";
		return RT_ERR_PARSE;
	}
}

void rt_arena_free(void* arena) { std::free(arena); }

const char* rt_host_last_error(void) { return g_host_error.c_str(); }

static const uint32_t kElemSize[RT_SEC_MAX] = {
	0, sizeof(rt_camera), sizeof(rt_globals), sizeof(rt_light), sizeof(rt_material), sizeof(rt_texture), 1,
	sizeof(rt_object), sizeof(rt_node), 4, sizeof(rt_mesh), sizeof(rt_node), 4, sizeof(rt_face), 24, 24, 4};

static bool tree_ok(const rt_node* n, uint64_t n_nodes, uint64_t n_items, uint64_t item_limit, const int32_t* items)
{
	for (uint64_t k = 0; k < n_nodes; k++) {
		if (n[k].right >= 0) {
			if ((uint64_t)n[k].right >= n_nodes || (uint64_t)n[k].right <= k + 1) return false;
		} else {
			// count <= 3 (kdTree.h LEAF_NUM) and first < 2^29: leaf_ref packs (first << 2 | count) into 31 bits and the
			// walks keep a 3-bit pending mask per leaf - the same limits the device check applies (k_derive_nodes2)
			if (n[k].first < 0 || n[k].count < 0 || n[k].count > 3 || n[k].first >= (1 << 29) || (uint64_t)n[k].first + n[k].count > n_items) return false;
			for (int32_t i = 0; i < n[k].count; i++)
				if (items[n[k].first + i] < 0 || (uint64_t)items[n[k].first + i] >= item_limit) return false;
		}
	}
	return true;
}

static rt_status validate_impl(const void* arena, size_t bytes, bool deep);
rt_status rt_arena_validate(const void* arena, size_t bytes) { return validate_impl(arena, bytes, true); }
// header, section table, small tables; the per-element index checks are left to the caller
// (rt_upload_scene runs them on the device)
rt_status rt_arena_validate_tables(const void* arena, size_t bytes) { return validate_impl(arena, bytes, false); }

static rt_status validate_impl(const void* arena, size_t bytes, bool deep)
{
	if (!arena || bytes < sizeof(rt_arena_header)) return RT_ERR_ARG;
	rt::ArenaView A(arena);
	const rt_arena_header* h = A.hdr;
	if (h->magic != RT_ARENA_MAGIC || h->version != RT_ARENA_VERSION || h->n_sections != RT_SEC_MAX || h->total_bytes != bytes) return RT_ERR_ARG;
	for (uint32_t s = 1; s < RT_SEC_MAX; s++) {
		const rt_section& sec = h->sections[s];
		if (sec.id != s || sec.elem_size != kElemSize[s] || sec.offset % RT_ARENA_ALIGN) return RT_ERR_ARG;
		if (sec.offset > bytes || sec.count > (bytes - sec.offset) / (sec.elem_size ? sec.elem_size : 1)) return RT_ERR_ARG;
	}
	if (A.count(RT_SEC_CAMERA) != 1 || A.count(RT_SEC_GLOBALS) != 1) return RT_ERR_ARG;
	const uint64_t n_obj = A.count(RT_SEC_OBJECTS), n_mat = A.count(RT_SEC_MATERIALS), n_tex = A.count(RT_SEC_TEXTURES);
	const uint64_t n_mesh = A.count(RT_SEC_MESHES);
	if (n_obj == 0 || n_obj > 0x3fffffffull) return RT_ERR_ARG;
	const rt_globals* g = A.ptr<rt_globals>(RT_SEC_GLOBALS);
	if ((uint64_t)g->n_objects != n_obj || (uint64_t)g->n_lights != A.count(RT_SEC_LIGHTS)) return RT_ERR_ARG;
	for (int k = 0; k < 6; k++)
		if (g->has_cubemap && (g->cubemap[k] < 0 || (uint64_t)g->cubemap[k] >= n_tex)) return RT_ERR_ARG;
	const rt_texture* tx = A.ptr<rt_texture>(RT_SEC_TEXTURES);
	for (uint64_t k = 0; k < n_tex; k++)
		if (tx[k].width <= 0 || tx[k].height <= 0 || tx[k].offset + (uint64_t)tx[k].width * tx[k].height * 3 > A.count(RT_SEC_TEXDATA)) return RT_ERR_ARG;
	const rt_material* mt = A.ptr<rt_material>(RT_SEC_MATERIALS);
	for (uint64_t k = 0; k < n_mat; k++) {
		for (int p = 0; p < RT_NPARAM; p++)
			if (mt[k].tex[p] >= (int64_t)n_tex) return RT_ERR_ARG;
		if (mt[k].shininess_tex >= (int64_t)n_tex || mt[k].index_tex >= (int64_t)n_tex) return RT_ERR_ARG;
	}
	const rt_object* ob = A.ptr<rt_object>(RT_SEC_OBJECTS);
	for (uint64_t k = 0; k < n_obj; k++) {
		if (ob[k].type < 0 || ob[k].type > RT_OBJ_TRIMESH || ob[k].material < 0 || (uint64_t)ob[k].material >= n_mat) return RT_ERR_ARG;
		if (ob[k].type == RT_OBJ_TRIMESH && (ob[k].mesh < 0 || (uint64_t)ob[k].mesh >= n_mesh)) return RT_ERR_ARG;
	}
	if (deep && !tree_ok(A.ptr<rt_node>(RT_SEC_SCENE_NODES), A.count(RT_SEC_SCENE_NODES), A.count(RT_SEC_SCENE_ITEMS), n_obj, A.ptr<int32_t>(RT_SEC_SCENE_ITEMS)))
		return RT_ERR_ARG;
	uint64_t prev_node_end = 0, prev_item_end = 0;
	const rt_mesh* ms = A.ptr<rt_mesh>(RT_SEC_MESHES);
	const rt_face* fc = A.ptr<rt_face>(RT_SEC_FACES);
	for (uint64_t k = 0; k < n_mesh; k++) {
		const rt_mesh& m = ms[k];
		if (m.node_off + m.node_cnt > A.count(RT_SEC_MESH_NODES) || m.item_off + m.item_cnt > A.count(RT_SEC_MESH_ITEMS)) return RT_ERR_ARG;
		if (m.face_off + m.face_cnt > A.count(RT_SEC_FACES) || m.vert_off + m.vert_cnt > A.count(RT_SEC_VERTS)) return RT_ERR_ARG;
		if (m.vnorm_off >= 0 && (uint64_t)m.vnorm_off + m.vert_cnt > A.count(RT_SEC_VNORMALS)) return RT_ERR_ARG;
		if (m.vmat_off >= 0 && (uint64_t)m.vmat_off + m.vert_cnt > A.count(RT_SEC_VMATS)) return RT_ERR_ARG;
		if (m.object < 0 || (uint64_t)m.object >= n_obj || m.node_cnt == 0) return RT_ERR_ARG;
		// meshes lie back to back in ascending order (the device kernels locate a node's mesh by bisection)
		if (m.node_off < prev_node_end || m.item_off < prev_item_end) return RT_ERR_ARG;
		prev_node_end = m.node_off + m.node_cnt;
		prev_item_end = m.item_off + m.item_cnt;
		if (!deep) continue;
		for (uint64_t f = 0; f < m.face_cnt; f++)
			for (int v = 0; v < 3; v++)
				if (fc[m.face_off + f].ids[v] < 0 || (uint64_t)fc[m.face_off + f].ids[v] >= m.vert_cnt) return RT_ERR_ARG;
		if (!tree_ok(A.ptr<rt_node>(RT_SEC_MESH_NODES) + m.node_off, m.node_cnt, m.item_cnt, m.face_cnt, A.ptr<int32_t>(RT_SEC_MESH_ITEMS) + m.item_off))
			return RT_ERR_ARG;
	}
	const int32_t* vm = A.ptr<int32_t>(RT_SEC_VMATS);
	for (uint64_t k = 0; deep && k < A.count(RT_SEC_VMATS); k++)
		if (vm[k] < 0 || (uint64_t)vm[k] >= n_mat) return RT_ERR_ARG;
	return RT_OK;
}

rt_status rt_arena_save(const char* path, const void* arena, size_t bytes)
{
	if (!path || rt_arena_validate_tables(arena, bytes) != RT_OK) return RT_ERR_ARG;
	FILE* f = std::fopen(path, "wb");
	if (!f) return RT_ERR_IO;
	const bool ok = std::fwrite(arena, 1, bytes, f) == bytes;
	return (std::fclose(f) == 0 && ok) ? RT_OK : RT_ERR_IO;
}

rt_status rt_arena_load(const char* path, void** arena_out, size_t* bytes_out)
{
	if (!path || !arena_out || !bytes_out) return RT_ERR_ARG;
	*arena_out = nullptr;
	*bytes_out = 0;
	FILE* f = std::fopen(path, "rb");
	if (!f) return RT_ERR_IO;
	rt_arena_header h;
	if (std::fread(&h, 1, sizeof(h), f) != sizeof(h) || h.magic != RT_ARENA_MAGIC || h.version != RT_ARENA_VERSION || h.total_bytes < sizeof(h)) { std::fclose(f); return RT_ERR_ARG; }
	void* p = std::aligned_alloc(RT_ARENA_ALIGN, (h.total_bytes + RT_ARENA_ALIGN - 1) / RT_ARENA_ALIGN * RT_ARENA_ALIGN);
	if (!p) { std::fclose(f); return RT_ERR_IO; }
	std::memcpy(p, &h, sizeof(h));
	const size_t rest = h.total_bytes - sizeof(h);
	const bool ok = std::fread((char*)p + sizeof(h), 1, rest, f) == rest && std::fgetc(f) == EOF;
	std::fclose(f);
	if (!ok || rt_arena_validate(p, h.total_bytes) != RT_OK) { std::free(p); return RT_ERR_ARG; }
	*arena_out = p;
	*bytes_out = h.total_bytes;
	return RT_OK;
}

rt_status rt_write_image(const char* path, int w, int h, const uint8_t* rgb)
{
	if (!path || !rgb || w <= 0 || h <= 0) return RT_ERR_ARG;
	return rs::write_image(path, w, h, rgb) ? RT_OK : RT_ERR_IO;
}

rt_status rt_read_image(const char* path, int* w, int* h, uint8_t** rgb_out)
{
	if (!path || !w || !h || !rgb_out) return RT_ERR_ARG;
	rs::Texture t;
	if (!rs::read_image(path, t)) return RT_ERR_IO;
	uint8_t* p = (uint8_t*)std::malloc(t.rgb.size() ? t.rgb.size() : 1);
	if (!p) return RT_ERR_IO;
	std::memcpy(p, t.rgb.data(), t.rgb.size());
	*w = t.width;
	*h = t.height;
	*rgb_out = p;
	return RT_OK;
}

} // extern "C"
