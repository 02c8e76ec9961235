// scene_flatten.cpp - builds the reference's trees on the host and writes the arena.
//
// North-star item (1): "the host kdTree build is kept so the tree matches the reference".  The
// build below restates KdTree<T>::KdTree (reference ray/src/scene/kdTree.h:23-78): bounds by
// BoundingBox::merge (bbox.cc:107-119), leaf at <= LEAF_NUM = 3 items, split axis = longest
// extent (first axis wins ties), items ordered by std::sort on (bmax+bmin)[axis] with `<`, first
// floor(n/2) to child[0].  std::sort is not stable; using the same libstdc++ std::sort on the same
// initial order with a comparator that returns the same answers reproduces the reference's
// permutation exactly (tests/test_arena_parity.py checks every node and leaf against the tree the
// reference itself built).  Nodes are emitted in DFS pre-order so an item's position in the item
// array is its rank in the reference's candidate list (kdTree.h:100-117).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <unordered_map>
#include "ray_scene.hpp"
#include "rt_arena.h"
#include "rt_arena_writer.hpp"

namespace rs {
namespace {

struct Box3 { double lo[3], hi[3]; };

class TreeBuilder {
public:
	TreeBuilder(const std::vector<Box3>& boxes, std::vector<rt_node>& nodes, std::vector<int32_t>& items)
		: boxes_(boxes), nodes_(nodes), items_(items), node_base_(nodes.size()), item_base_(items.size()) {}

	void build_all()
	{
		std::vector<int> all(boxes_.size());
		std::iota(all.begin(), all.end(), 0);
		build(all);
	}

private:
	void build(const std::vector<int>& idx)
	{
		const size_t me = nodes_.size();
		rt_node rec;
		std::memset(&rec, 0, sizeof(rec));
		rec.right = -1;
		rec.first = (int32_t)(items_.size() - item_base_);
		nodes_.push_back(rec);
		if (idx.empty()) return;
		Box3 b = boxes_[idx[0]];
		for (int ix : idx) {
			const Box3& o = boxes_[ix];
			for (int a = 0; a < 3; a++) {
				if (o.lo[a] < b.lo[a]) b.lo[a] = o.lo[a];
				if (o.hi[a] > b.hi[a]) b.hi[a] = o.hi[a];
			}
		}
		for (int a = 0; a < 3; a++) { nodes_[me].bmin[a] = b.lo[a]; nodes_[me].bmax[a] = b.hi[a]; }
		if (idx.size() <= 3) {
			nodes_[me].count = (int32_t)idx.size();
			for (int ix : idx) items_.push_back(ix);
			return;
		}
		int axis = 0;
		for (int a = 1; a < 3; a++)
			if (b.hi[axis] - b.lo[axis] < b.hi[a] - b.lo[a]) axis = a;
		std::vector<double> key(idx.size());
		for (size_t k = 0; k < idx.size(); k++) key[k] = boxes_[idx[k]].hi[axis] + boxes_[idx[k]].lo[axis];
		std::vector<int> order(idx.size());
		std::iota(order.begin(), order.end(), 0);
		const double* kp = key.data();
		std::sort(order.begin(), order.end(), [kp](const int a, const int b) { return kp[a] < kp[b]; });
		const size_t mid = idx.size() / 2;
		std::vector<int> half[2];
		half[0].reserve(mid);
		half[1].reserve(idx.size() - mid);
		for (size_t k = 0; k < idx.size(); k++) half[k < mid ? 0 : 1].push_back(idx[order[k]]);
		key.clear(); key.shrink_to_fit();
		order.clear(); order.shrink_to_fit();
		build(half[0]);
		nodes_[me].right = (int32_t)(nodes_.size() - node_base_);
		build(half[1]);
	}
	const std::vector<Box3>& boxes_;
	std::vector<rt_node>& nodes_;
	std::vector<int32_t>& items_;
	size_t node_base_, item_base_;
};

void put3(double* d, const V3& v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }

struct MaterialTable {
	const Scene& sc;
	std::vector<rt_material> materials;
	std::vector<rt_texture> textures;
	std::vector<uint8_t> texdata;
	std::unordered_map<int, int> tex_remap;
	std::unordered_map<std::string, int> ids;
	explicit MaterialTable(const Scene& s) : sc(s) {}

	int texture_id(int t) // arena texture ids are in first-use order
	{
		if (t < 0) return -1;
		auto it = tex_remap.find(t);
		if (it != tex_remap.end()) return it->second;
		const Texture& src = sc.textures[t];
		rt_texture rec;
		rec.width = src.width;
		rec.height = src.height;
		rec.offset = texdata.size();
		texdata.insert(texdata.end(), src.rgb.begin(), src.rgb.end());
		while (texdata.size() % 16) texdata.push_back(0);
		textures.push_back(rec);
		tex_remap[t] = (int)textures.size() - 1;
		return (int)textures.size() - 1;
	}
	static double intensity(const V3& v) { return (0.299 * v.x + 0.587 * v.y) + 0.114 * v.z; } // material.cpp:148-158
	int material_id(const Material& m)
	{
		rt_material r;
		std::memset(&r, 0, sizeof(r));
		const MatParam* ps[RT_NPARAM] = {&m.ke, &m.ka, &m.ks, &m.kd, &m.kr, &m.kt, &m.bump};
		for (int k = 0; k < RT_NPARAM; k++) {
			put3(r.k[k], ps[k]->value);
			r.tex[k] = texture_id(ps[k]->tex);
		}
		r.shininess_tex = texture_id(m.shininess.tex);
		r.index_tex = texture_id(m.index.tex);
		r.shininess = r.shininess_tex < 0 ? intensity(m.shininess.value) : 0.0;
		r.index = r.index_tex < 0 ? intensity(m.index.value) : 0.0;
		r.shininess_raw = m.shininess.value.x;
		r.index_raw = m.index.value.x;
		r.refl = m.refl;
		r.trans = m.trans;
		r.recur = m.recur;
		std::string key((const char*)&r, sizeof(r));
		auto it = ids.find(key);
		if (it != ids.end()) return it->second;
		materials.push_back(r);
		ids[key] = (int)materials.size() - 1;
		return (int)materials.size() - 1;
	}
};

} // namespace

void* flatten(const Scene& sc, size_t* bytes_out)
{
	rt::ArenaWriter W;
	MaterialTable T(sc);

	std::vector<rt_camera> cam(1);
	std::memset(&cam[0], 0, sizeof(rt_camera));
	put3(cam[0].eye, sc.camera.eye);
	put3(cam[0].look, sc.camera.look);
	put3(cam[0].u, sc.camera.u);
	put3(cam[0].v, sc.camera.v);
	cam[0].aspect = sc.camera.aspect;

	std::vector<rt_light> lights;
	for (const Light& l : sc.lights) {
		rt_light r;
		std::memset(&r, 0, sizeof(r));
		r.type = (int32_t)l.type;
		put3(r.color, l.color);
		put3(r.position, l.position);
		put3(r.orientation, l.orientation);
		if (l.type != L_DIRECTIONAL) { r.atten[0] = l.a; r.atten[1] = l.b; r.atten[2] = l.c; }
		put3(r.u, l.u);
		put3(r.v, l.v);
		r.w = l.w; r.h = l.h; r.radius = l.radius; r.ang_tan = l.ang_tan; r.offset = l.offset;
		lights.push_back(r);
	}

	std::vector<rt_object> objects;
	std::vector<rt_mesh> meshes;
	std::vector<rt_node> mesh_nodes, scene_nodes;
	std::vector<int32_t> mesh_items, scene_items, vmats;
	std::vector<rt_face> faces;
	std::vector<double> verts, vnormals;
	std::vector<Box3> obj_boxes;

	int oi = 0;
	for (const Object& src : sc.objects) {
		rt_object o;
		std::memset(&o, 0, sizeof(o));
		o.type = (int32_t)src.type;
		o.mesh = -1;
		o.material = T.material_id(src.material);
		put3(o.bbox, src.bmin);
		put3(o.bbox + 3, src.bmax);
		for (int c = 0; c < 4; c++)
			for (int r = 0; r < 4; r++) o.inv[c * 4 + r] = src.inverse[c][r];
		for (int c = 0; c < 3; c++)
			for (int r = 0; r < 3; r++) o.normi[c * 3 + r] = src.normi[c][r];
		if (src.type == CYLINDER) o.capped = 1;
		if (src.type == CONE) {
			o.capped = src.capped ? 1 : 0;
			o.cone[0] = src.height; o.cone[1] = src.b_radius; o.cone[2] = src.t_radius;
			o.cone[3] = src.beta_squared; o.cone[4] = src.gamma;
		}
		if (src.type == TRIMESH) {
			const Mesh& mesh = sc.meshes[src.mesh];
			o.mesh = (int32_t)meshes.size();
			rt_mesh m;
			std::memset(&m, 0, sizeof(m));
			m.object = oi;
			m.node_off = mesh_nodes.size();
			m.item_off = mesh_items.size();
			m.face_off = faces.size();
			m.vert_off = verts.size() / 3;
			m.vert_cnt = mesh.vertices.size();
			m.face_cnt = mesh.face_normals.size();
			m.vnorm_off = -1;
			m.vmat_off = -1;
			for (const V3& v : mesh.vertices) { verts.push_back(v.x); verts.push_back(v.y); verts.push_back(v.z); }
			if (!mesh.normals.empty()) {
				m.vnorm_off = (int64_t)(vnormals.size() / 3);
				for (const V3& v : mesh.normals) { vnormals.push_back(v.x); vnormals.push_back(v.y); vnormals.push_back(v.z); }
			}
			if (!mesh.vmaterials.empty()) {
				m.vmat_off = (int64_t)vmats.size();
				for (const Material& vm : mesh.vmaterials) vmats.push_back(T.material_id(vm));
			}
			std::vector<Box3> fboxes(mesh.face_normals.size());
			for (size_t f = 0; f < mesh.face_normals.size(); f++) {
				rt_face fr;
				std::memset(&fr, 0, sizeof(fr));
				for (int k = 0; k < 3; k++) fr.ids[k] = mesh.face_ids[3 * f + k];
				put3(fr.normal, mesh.face_normals[f]);
				faces.push_back(fr);
				// TrimeshFace::ComputeLocalBoundingBox (trimesh.h:149-161)
				const V3& a = mesh.vertices[fr.ids[0]];
				const V3& b = mesh.vertices[fr.ids[1]];
				const V3& c = mesh.vertices[fr.ids[2]];
				V3 hi = hm::gmax(c, hm::gmax(a, b)), lo = hm::gmin(c, hm::gmin(a, b));
				put3(fboxes[f].lo, lo);
				put3(fboxes[f].hi, hi);
			}
			TreeBuilder(fboxes, mesh_nodes, mesh_items).build_all();
			m.node_cnt = mesh_nodes.size() - m.node_off;
			m.item_cnt = mesh_items.size() - m.item_off;
			meshes.push_back(m);
		}
		Box3 ob;
		put3(ob.lo, src.bmin);
		put3(ob.hi, src.bmax);
		obj_boxes.push_back(ob);
		objects.push_back(o);
		oi++;
	}
	TreeBuilder(obj_boxes, scene_nodes, scene_items).build_all();

	std::vector<rt_globals> glob(1);
	std::memset(&glob[0], 0, sizeof(rt_globals));
	put3(glob[0].ambient, sc.ambient);
	glob[0].n_objects = (int32_t)objects.size();
	glob[0].n_lights = (int32_t)lights.size();
	for (int k = 0; k < 6; k++) glob[0].cubemap[k] = -1;
	if (sc.has_cubemap) {
		glob[0].has_cubemap = 1;
		for (int k = 0; k < 6; k++) glob[0].cubemap[k] = T.texture_id(sc.cubemap[k]);
	}

	W.put(RT_SEC_CAMERA, cam);
	W.put(RT_SEC_GLOBALS, glob);
	W.put(RT_SEC_LIGHTS, lights);
	W.put(RT_SEC_MATERIALS, T.materials);
	W.put(RT_SEC_TEXTURES, T.textures);
	W.put(RT_SEC_TEXDATA, T.texdata);
	W.put(RT_SEC_OBJECTS, objects);
	W.put(RT_SEC_SCENE_NODES, scene_nodes);
	W.put(RT_SEC_SCENE_ITEMS, scene_items);
	W.put(RT_SEC_MESHES, meshes);
	W.put(RT_SEC_MESH_NODES, mesh_nodes);
	W.put(RT_SEC_MESH_ITEMS, mesh_items);
	W.put(RT_SEC_FACES, faces);
	W.put_raw(RT_SEC_VERTS, verts.data(), 24, verts.size() / 3);
	W.put_raw(RT_SEC_VNORMALS, vnormals.data(), 24, vnormals.size() / 3);
	W.put(RT_SEC_VMATS, vmats);
	return W.finish(bytes_out);
}

} // namespace rs
