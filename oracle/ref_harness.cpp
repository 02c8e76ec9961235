// ref_harness.cpp - extern "C" harness around the UNMODIFIED reference ray tracer.
//
// TEST INFRASTRUCTURE ONLY.  Linked with the reference's own translation units (compiled where
// they lie under /root/reference/ray/src, see oracle/Makefile) into oracle/_ref/libref_oracle.so.
// Nothing under cs378hgraphics_raytracer_b200/ links, loads or calls this; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
//
// What it adds (all new code, no reference code is copied):
//   * the globals normally defined in the reference's main.cpp (theRayTracer, traceUI, ...)
//   * a TraceUI subclass whose settings are filled from an rt_params (include/rt_b200.h)
//   * ref_flatten(): walks the real Scene/KdTree/Trimesh/Material/Light graph (private members
//     reached with -fno-access-control) and writes the arena of include/rt_arena.h, so tests can
//     compare the product's own loader bit for bit against what the reference built
//   * ref_trace_rays / ref_trace_pixels / ref_render: RayTracer::traceRay / tracePixel /
//     traceImage over batches, with hit objects turned into (object index, face index)
//   * ray counters through link-time --wrap of Scene::intersect / Scene::intersectList
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>
#include <omp.h>

#include "RayTracer.h"
#include "scene/scene.h"
#include "scene/light.h"
#include "scene/material.h"
#include "scene/cubeMap.h"
#include "scene/kdTree.h"
#include "SceneObjects/Box.h"
#include "SceneObjects/Cone.h"
#include "SceneObjects/Cylinder.h"
#include "SceneObjects/Sphere.h"
#include "SceneObjects/Square.h"
#include "SceneObjects/trimesh.h"
#include "ui/TraceUI.h"

#include "rt_arena.h"
#include "rt_arena_writer.hpp"
#include "rt_b200.h"

// ---- globals the reference expects from its main.cpp (main.cpp:10-13) -------------------------
RayTracer* theRayTracer = nullptr;
TraceUI* traceUI = nullptr;
int TraceUI::m_threads = std::max(std::thread::hardware_concurrency(), (unsigned)1);
int TraceUI::rayCount[MAX_THREADS];

namespace {

class HarnessUI : public TraceUI {
public:
	int run() override { return 0; }
	void alert(const string& msg) override { last_alert = msg; }
	std::string last_alert;
	void apply(const rt_params& p) {
		m_nDepth = p.depth;
		m_aa_mode = (AAMode)p.aa_mode;
		m_aa_samples = p.aa_samples;
		m_aa_thresh = p.aa_thresh;
		m_aterm_thresh = p.aterm_thresh;
		m_dof = p.dof_on != 0;
		m_dof_div = p.dof_divs;
		m_dof_fd = p.dof_fd;
		m_dof_apsz = p.dof_apsz;
		m_anaglyph = p.anaglyph != 0;
		m_overlappingObjects = p.overlapping_objects != 0;
		m_ss_res = p.soft_shadow_res;
		m_usingCubeMap = p.use_cubemap != 0;
	}
};

HarnessUI* g_ui = nullptr;
std::unordered_map<const void*, std::pair<int, int>> g_ids; // SceneObject* -> (obj, face)

// per-thread padded ray counters (no contention in the timed region)
struct alignas(64) Counters {
	uint64_t camera, reflection, refraction, shadow;
};
Counters g_cnt[1024];
bool g_counting = true;
// one slot per OS thread (omp_get_thread_num() is 0 for every thread inside the reference's nested,
// serialised parallel regions of adaptaa, RayTracer.cpp:330-337, so it cannot be the index)
std::atomic<int> g_next_slot{0};
thread_local int t_slot = -1;
inline Counters& my_counters()
{
	if (t_slot < 0) t_slot = g_next_slot.fetch_add(1) & 1023;
	return g_cnt[t_slot];
}

void reset_counters() { std::memset(g_cnt, 0, sizeof(g_cnt)); }

void v3(double* dst, const glm::dvec3& v) { dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2]; }

// ---- flattening the real reference scene -------------------------------------------------------
struct Flattener {
	std::vector<rt_material> materials;
	std::vector<rt_texture> textures;
	std::vector<uint8_t> texdata;
	std::unordered_map<const TextureMap*, int> tex_ids;
	std::unordered_map<std::string, int> mat_ids;

	int texture_id(const TextureMap* t) {
		if (!t) return -1;
		auto it = tex_ids.find(t);
		if (it != tex_ids.end()) return it->second;
		rt_texture rec;
		rec.width = t->width;
		rec.height = t->height;
		rec.offset = texdata.size();
		size_t n = (size_t)t->width * t->height * 3;
		texdata.insert(texdata.end(), t->data.begin(), t->data.begin() + n);
		while (texdata.size() % 16) texdata.push_back(0);
		int id = (int)textures.size();
		textures.push_back(rec);
		tex_ids[t] = id;
		return id;
	}

	int material_id(const Material& m) {
		rt_material r;
		std::memset(&r, 0, sizeof(r));
		const MaterialParameter* ps[RT_NPARAM] = {&m._ke, &m._ka, &m._ks, &m._kd, &m._kr, &m._kt, &m._bump};
		for (int k = 0; k < RT_NPARAM; k++) {
			v3(r.k[k], ps[k]->_value);
			r.tex[k] = texture_id(ps[k]->_textureMap);
		}
		isect dummy;
		r.shininess_tex = texture_id(m._shininess._textureMap);
		r.index_tex = texture_id(m._index._textureMap);
		// constants: evaluate through the reference's own accessor (material.cpp:148-158)
		r.shininess = r.shininess_tex < 0 ? m._shininess.intensityValue(dummy) : 0.0;
		r.index = r.index_tex < 0 ? m._index.intensityValue(dummy) : 0.0;
		r.shininess_raw = m._shininess._value[0];
		r.index_raw = m._index._value[0];
		r.refl = m._refl ? 1 : 0;
		r.trans = m._trans ? 1 : 0;
		r.recur = m._recur ? 1 : 0;
		std::string key((const char*)&r, sizeof(r));
		auto it = mat_ids.find(key);
		if (it != mat_ids.end()) return it->second;
		materials.push_back(r);
		mat_ids[key] = (int)materials.size() - 1;
		return (int)materials.size() - 1;
	}

	template <typename T>
	static void walk_tree(const KdTree<T>* n, std::vector<rt_node>& nodes, std::vector<int32_t>& items, size_t node_base, size_t item_base) {
		size_t me = nodes.size();
		rt_node rec;
		std::memset(&rec, 0, sizeof(rec));
		v3(rec.bmin, n->bound.bmin);
		v3(rec.bmax, n->bound.bmax);
		rec.right = -1;
		rec.first = (int32_t)(items.size() - item_base);
		rec.count = (int32_t)n->it_idxs.size();
		nodes.push_back(rec);
		if (n->it_idxs.size() > 0) {
			for (int it : n->it_idxs) items.push_back(it);
		} else if (n->child[0]) {
			walk_tree(n->child[0], nodes, items, node_base, item_base);
			nodes[me].right = (int32_t)(nodes.size() - node_base);
			walk_tree(n->child[1], nodes, items, node_base, item_base);
		}
	}
};

} // namespace

// ---- link-time ray counters (oracle/Makefile: -Wl,--wrap=...) ---------------------------------
bool real_Scene_intersect(const Scene*, ray&, isect&) asm("__real__ZNK5Scene9intersectER3rayR5isect");
bool wrap_Scene_intersect(const Scene*, ray&, isect&) asm("__wrap__ZNK5Scene9intersectER3rayR5isect");
std::vector<isect> real_Scene_intersectList(const Scene*, ray&) asm("__real__ZNK5Scene13intersectListER3ray");
std::vector<isect> wrap_Scene_intersectList(const Scene*, ray&) asm("__wrap__ZNK5Scene13intersectListER3ray");

bool wrap_Scene_intersect(const Scene* s, ray& r, isect& i)
{
	if (g_counting) {
		Counters& c = my_counters();
		switch (r.type()) {
		case ray::REFLECTION: c.reflection++; break;
		case ray::REFRACTION: c.refraction++; break;
		default: c.camera++; break;
		}
	}
	return real_Scene_intersect(s, r, i);
}
std::vector<isect> wrap_Scene_intersectList(const Scene* s, ray& r)
{
	if (g_counting) my_counters().shadow++;
	return real_Scene_intersectList(s, r);
}

// ---- arena from the real scene ----------------------------------------------------------------
static void* flatten_scene(const Scene* sc, size_t* bytes_out)
{
	Flattener F;
	rt::ArenaWriter W;

	std::vector<rt_camera> cam(1);
	std::memset(&cam[0], 0, sizeof(rt_camera));
	v3(cam[0].eye, sc->camera.eye);
	v3(cam[0].look, sc->camera.look);
	v3(cam[0].u, sc->camera.u);
	v3(cam[0].v, sc->camera.v);
	cam[0].aspect = sc->camera.aspectRatio;

	std::vector<rt_light> lights;
	for (const auto& lp : sc->lights) {
		const Light* l = lp.get();
		rt_light r;
		std::memset(&r, 0, sizeof(r));
		v3(r.color, l->color);
		if (auto s = dynamic_cast<const SpotLight*>(l)) {
			r.type = RT_LIGHT_SPOT;
			r.radius = s->r;
			r.ang_tan = s->ang_tan;
			r.offset = s->offset;
		} else if (auto c = dynamic_cast<const AreaLightCirc*>(l)) {
			r.type = RT_LIGHT_AREA_CIRC;
			r.radius = c->r;
		} else if (auto q = dynamic_cast<const AreaLightRect*>(l)) {
			r.type = RT_LIGHT_AREA_RECT;
			v3(r.u, q->u);
			v3(r.v, q->v);
			r.w = q->w;
			r.h = q->h;
		} else if (dynamic_cast<const PointLight*>(l)) {
			r.type = RT_LIGHT_POINT;
		} else {
			r.type = RT_LIGHT_DIRECTIONAL;
		}
		if (auto p = dynamic_cast<const PointLight*>(l)) {
			v3(r.position, p->position);
			r.atten[0] = p->constantTerm;
			r.atten[1] = p->linearTerm;
			r.atten[2] = p->quadraticTerm;
		}
		if (auto a = dynamic_cast<const AreaLight*>(l)) v3(r.orientation, a->ori);
		if (auto d = dynamic_cast<const DirectionalLight*>(l)) v3(r.orientation, d->orientation);
		lights.push_back(r);
	}

	std::vector<rt_object> objects;
	std::vector<rt_mesh> meshes;
	std::vector<rt_node> mesh_nodes;
	std::vector<int32_t> mesh_items;
	std::vector<rt_face> faces;
	std::vector<double> verts, vnormals;
	std::vector<int32_t> vmats;
	g_ids.clear();

	int oi = 0;
	for (const auto& gp : sc->objects) {
		const Geometry* g = gp.get();
		rt_object o;
		std::memset(&o, 0, sizeof(o));
		o.mesh = -1;
		const MaterialSceneObject* mso = dynamic_cast<const MaterialSceneObject*>(g);
		o.material = F.material_id(mso->getMaterial());
		v3(o.bbox, g->bounds.bmin);
		v3(o.bbox + 3, g->bounds.bmax);
		for (int c = 0; c < 4; c++)
			for (int r = 0; r < 4; r++) o.inv[c * 4 + r] = g->transform->inverse[c][r];
		for (int c = 0; c < 3; c++)
			for (int r = 0; r < 3; r++) o.normi[c * 3 + r] = g->transform->normi[c][r];
		if (dynamic_cast<const Sphere*>(g)) o.type = RT_OBJ_SPHERE;
		else if (dynamic_cast<const Box*>(g)) o.type = RT_OBJ_BOX;
		else if (dynamic_cast<const Square*>(g)) o.type = RT_OBJ_SQUARE;
		else if (dynamic_cast<const Cylinder*>(g)) { o.type = RT_OBJ_CYLINDER; o.capped = 1; }
		else if (auto cn = dynamic_cast<const Cone*>(g)) {
			o.type = RT_OBJ_CONE;
			o.capped = cn->capped ? 1 : 0;
			o.cone[0] = cn->height;
			o.cone[1] = cn->b_radius;
			o.cone[2] = cn->t_radius;
			o.cone[3] = cn->beta_squared;
			o.cone[4] = cn->gamma;
		} else if (auto tm = dynamic_cast<const Trimesh*>(g)) {
			o.type = RT_OBJ_TRIMESH;
			o.mesh = (int32_t)meshes.size();
			rt_mesh m;
			std::memset(&m, 0, sizeof(m));
			m.object = oi;
			m.node_off = mesh_nodes.size();
			m.item_off = mesh_items.size();
			m.face_off = faces.size();
			m.vert_off = verts.size() / 3;
			m.vert_cnt = tm->vertices.size();
			m.face_cnt = tm->faces.size();
			m.vnorm_off = -1;
			m.vmat_off = -1;
			for (const auto& v : tm->vertices) { verts.push_back(v[0]); verts.push_back(v[1]); verts.push_back(v[2]); }
			if (!tm->normals.empty()) {
				m.vnorm_off = (int64_t)(vnormals.size() / 3);
				for (const auto& v : tm->normals) { vnormals.push_back(v[0]); vnormals.push_back(v[1]); vnormals.push_back(v[2]); }
			}
			if (!tm->materials.empty()) {
				m.vmat_off = (int64_t)vmats.size();
				for (const Material* vm : tm->materials) vmats.push_back(F.material_id(*vm));
			}
			int fi = 0;
			for (const TrimeshFace* f : tm->faces) {
				rt_face fr;
				std::memset(&fr, 0, sizeof(fr));
				fr.ids[0] = (*f)[0]; fr.ids[1] = (*f)[1]; fr.ids[2] = (*f)[2];
				v3(fr.normal, f->normal);
				faces.push_back(fr);
				g_ids[(const SceneObject*)f] = std::make_pair(oi, fi++);
			}
			Flattener::walk_tree(tm->kdtree, mesh_nodes, mesh_items, m.node_off, m.item_off);
			m.node_cnt = mesh_nodes.size() - m.node_off;
			m.item_cnt = mesh_items.size() - m.item_off;
			meshes.push_back(m);
		}
		g_ids[(const SceneObject*)mso] = std::make_pair(oi, -1);
		objects.push_back(o);
		oi++;
	}

	std::vector<rt_node> scene_nodes;
	std::vector<int32_t> scene_items;
	Flattener::walk_tree(sc->kdtree, scene_nodes, scene_items, 0, 0);

	std::vector<rt_globals> glob(1);
	std::memset(&glob[0], 0, sizeof(rt_globals));
	v3(glob[0].ambient, sc->ambientIntensity);
	glob[0].n_objects = (int32_t)objects.size();
	glob[0].n_lights = (int32_t)lights.size();
	for (int k = 0; k < 6; k++) glob[0].cubemap[k] = -1;
	if (g_ui && g_ui->cubemap) {
		glob[0].has_cubemap = 1;
		for (int k = 0; k < 6; k++) glob[0].cubemap[k] = F.texture_id(g_ui->cubemap->tMap[k].get());
	}

	W.put(RT_SEC_CAMERA, cam);
	W.put(RT_SEC_GLOBALS, glob);
	W.put(RT_SEC_LIGHTS, lights);
	W.put(RT_SEC_MATERIALS, F.materials);
	W.put(RT_SEC_TEXTURES, F.textures);
	W.put(RT_SEC_TEXDATA, F.texdata);
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

// ---- extern "C" API -----------------------------------------------------------------------------
extern "C" {

static void ensure_globals()
{
	if (!g_ui) {
		g_ui = new HarnessUI();
		traceUI = g_ui;
	}
	if (!theRayTracer) {
		theRayTracer = new RayTracer();
		g_ui->setRayTracer(theRayTracer);
	}
}

// Loads a .ray file through the reference's own Tokenizer/Parser (RayTracer::loadScene).
// cubemap_faces: NULL or six paths (+x -x +y -y +z -z), loaded like TraceUI::smartLoadCubemap does.
int ref_load(const char* path, const char* const* cubemap_faces)
{
	ensure_globals();
	g_ui->cubemap.reset();
	g_ui->m_usingCubeMap = false;
	if (cubemap_faces) {
		try {
			g_ui->setCubeMap(new CubeMap());
			for (int i = 0; i < 6; i++) g_ui->cubemap->setNthMap(i, new TextureMap(cubemap_faces[i]));
			g_ui->useCubeMap(true);
		} catch (TextureMapException& e) {
			g_ui->last_alert = e.message();
			g_ui->cubemap.reset();
			return 1;
		}
	}
	delete theRayTracer;
	theRayTracer = new RayTracer();
	g_ui->setRayTracer(theRayTracer);
	g_ids.clear();
	return theRayTracer->loadScene(path) ? 0 : 1;
}

const char* ref_last_alert() { return g_ui ? g_ui->last_alert.c_str() : ""; }

double ref_aspect() { return theRayTracer ? theRayTracer->aspectRatio() : 1.0; }

void ref_set_threads(int n) { TraceUI::m_threads = n > 0 ? n : std::max(std::thread::hardware_concurrency(), (unsigned)1); }
int ref_get_threads() { return TraceUI::m_threads; }

void ref_set_params(const rt_params* p)
{
	ensure_globals();
	bool had = g_ui->m_usingCubeMap;
	g_ui->apply(*p);
	(void)had;
}

// Arena written from the real reference scene graph; free with ref_free.  Also (re)builds the
// pointer -> (object, face) index maps used by ref_trace_rays.
void* ref_flatten(size_t* bytes_out)
{
	if (!theRayTracer || !theRayTracer->sceneLoaded()) return nullptr;
	return flatten_scene(theRayTracer->scene.get(), bytes_out);
}
void ref_free(void* p) { std::free(p); }

// RayTracer::traceRay on a batch: hits_out = Scene::intersect of each ray, rgb_out = traceRay colour.
int ref_trace_rays(const rt_ray* rays, int n, int depth, rt_hit* hits_out, double* rgb_out)
{
	if (!theRayTracer || !theRayTracer->sceneLoaded()) return 1;
	if (g_ids.empty()) { size_t b; void* a = ref_flatten(&b); std::free(a); }
	Scene* sc = theRayTracer->scene.get();
	const double thresh = traceUI->getATermThresh();
#pragma omp parallel for schedule(dynamic, 64) num_threads(TraceUI::m_threads)
	for (int k = 0; k < n; k++) {
		glm::dvec3 p(rays[k].p[0], rays[k].p[1], rays[k].p[2]);
		glm::dvec3 d(rays[k].d[0], rays[k].d[1], rays[k].d[2]);
		if (hits_out) {
			ray r(p, d, glm::dvec3(1.0), ray::VISIBILITY);
			isect i;
			bool hit = sc->intersect(r, i);
			rt_hit h;
			h.obj = h.face = -1;
			h.t = i.getT();
			v3(h.N, i.getN());
			if (hit) {
				auto it = g_ids.find((const void*)i.obj);
				if (it != g_ids.end()) { h.obj = it->second.first; h.face = it->second.second; }
				else h.obj = -2;
			} else {
				h.N[0] = h.N[1] = h.N[2] = 0.0;
			}
			hits_out[k] = h;
		}
		if (rgb_out) {
			ray r(p, d, glm::dvec3(1.0), ray::VISIBILITY);
			double t = 0.0;
			glm::dvec3 c = theRayTracer->traceRay(r, thresh, depth, t);
			rgb_out[3 * k] = c[0]; rgb_out[3 * k + 1] = c[1]; rgb_out[3 * k + 2] = c[2];
		}
	}
	return 0;
}

// RayTracer::tracePixel for n pixels of a w x h frame (unquantised dvec3 colours).
int ref_trace_pixels(int w, int h, const int32_t* ij, int n, double* rgb_out)
{
	if (!theRayTracer || !theRayTracer->sceneLoaded()) return 1;
	theRayTracer->traceSetup(w, h);
#pragma omp parallel for schedule(dynamic, 16) num_threads(TraceUI::m_threads)
	for (int k = 0; k < n; k++) {
		glm::dvec3 c = theRayTracer->tracePixel(ij[2 * k], ij[2 * k + 1]);
		rgb_out[3 * k] = c[0]; rgb_out[3 * k + 1] = c[1]; rgb_out[3 * k + 2] = c[2];
	}
	return 0;
}

// RayTracer::traceImage(w,h) + getBuffer.  seconds_out = wall time of traceImage alone (where the
// reference puts its own clocks, CommandLineUI.cpp:160-170).  stats_out gets the ray counters.
int ref_render(int w, int h, uint8_t* rgb_out, double* seconds_out, rt_stats* stats_out)
{
	if (!theRayTracer || !theRayTracer->sceneLoaded()) return 1;
	reset_counters();
	theRayTracer->traceSetup(w, h);
	auto t0 = std::chrono::steady_clock::now();
	theRayTracer->traceImage(w, h);
	auto t1 = std::chrono::steady_clock::now();
	if (seconds_out) *seconds_out = std::chrono::duration<double>(t1 - t0).count();
	unsigned char* buf; int bw, bh;
	theRayTracer->getBuffer(buf, bw, bh);
	if (rgb_out && buf) std::memcpy(rgb_out, buf, (size_t)w * h * 3);
	if (stats_out) {
		std::memset(stats_out, 0, sizeof(*stats_out));
		for (int t = 0; t < 1024; t++) {
			stats_out->camera_rays += g_cnt[t].camera;
			stats_out->reflection_rays += g_cnt[t].reflection;
			stats_out->refraction_rays += g_cnt[t].refraction;
			stats_out->shadow_rays += g_cnt[t].shadow;
		}
	}
	return 0;
}

// Renders only columns [i0,i1) x rows [j0,j1) of a w x h frame with the reference's per-pixel code
// (the same dispatch as traceImage's loop body) - the bounded sample used for CPU baselines of
// frames too large to finish on the host.  rgb_out is (i1-i0)*(j1-j0)*3 bytes, row-major by j.
int ref_render_window(int w, int h, int i0, int i1, int j0, int j1, uint8_t* rgb_out, double* seconds_out, rt_stats* stats_out)
{
	if (!theRayTracer || !theRayTracer->sceneLoaded()) return 1;
	reset_counters();
	theRayTracer->traceSetup(w, h);
	RayTracer* R = theRayTracer;
	const int ww = i1 - i0, wh = j1 - j0;
	auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for collapse(2) num_threads(TraceUI::m_threads)
	for (int i = i0; i < i1; i++) {
		for (int j = j0; j < j1; j++) {
			glm::dvec3 col(0.0);
			if (traceUI->aaSwitch()) {
				if (traceUI->getAAMode() != TraceUI::AAMode::ADAPTIVE) {
					int s = R->samples;
					for (int a = 0; a < s; a++)
						for (int b = 0; b < s; b++) col += R->tracePixel(i * s + a, j * s + b);
					col /= double(s * s);
				} else {
					R->adaptaa(double(i) / double(w), double(i + 1) / double(w), double(j) / double(h), double(j + 1) / double(h), col);
				}
			} else {
				col = R->tracePixel(i, j);
			}
			R->setPixel(i, j, col);
		}
	}
	auto t1 = std::chrono::steady_clock::now();
	if (seconds_out) *seconds_out = std::chrono::duration<double>(t1 - t0).count();
	unsigned char* buf; int bw, bh;
	R->getBuffer(buf, bw, bh);
	if (rgb_out)
		for (int j = j0; j < j1; j++)
			std::memcpy(rgb_out + (size_t)(j - j0) * ww * 3, buf + ((size_t)j * w + i0) * 3, (size_t)ww * 3);
	(void)wh;
	if (stats_out) {
		std::memset(stats_out, 0, sizeof(*stats_out));
		for (int t = 0; t < 1024; t++) {
			stats_out->camera_rays += g_cnt[t].camera;
			stats_out->reflection_rays += g_cnt[t].reflection;
			stats_out->refraction_rays += g_cnt[t].refraction;
			stats_out->shadow_rays += g_cnt[t].shadow;
		}
	}
	return 0;
}

} // extern "C"

// ---- a reference Scene built from an arena -------------------------------------------------------
// RayTracer::loadScene on the 869,556-face dragon stand-in takes about 32 minutes in the reference
// (KdTree's sort comparator copies the whole index vector each time it is passed by value,
// kdTree.h:58-65).  The trace loop itself is what is being checked and timed, so for big scenes the
// harness can instead populate the reference's OWN classes (Scene, Trimesh, TrimeshFace, KdTree,
// Material, lights, camera) from an arena: same objects, same trees (tests prove the arena equals
// what the reference builds), and every traversal / intersection / shading instruction that then
// runs is the unmodified reference's.
namespace {

glm::dvec3 g3(const double* p) { return glm::dvec3(p[0], p[1], p[2]); }

template <typename T>
KdTree<T>* tree_from_arena(const std::vector<T>* items, const rt_node* nodes, const int32_t* its, int node)
{
	KdTree<T>* t = new KdTree<T>(items, std::vector<int>()); // the empty-index ctor path: no build
	const rt_node& n = nodes[node];
	t->bound = BoundingBox(g3(n.bmin), g3(n.bmax));
	if (n.right < 0) {
		for (int i = 0; i < n.count; i++) t->it_idxs.push_back(its[n.first + i]);
	} else {
		t->child[0] = tree_from_arena(items, nodes, its, node + 1);
		t->child[1] = tree_from_arena(items, nodes, its, n.right);
	}
	return t;
}

TextureMap* texture_from_arena(const rt_texture& T, const uint8_t* texdata)
{
	TextureMap* t = (TextureMap*)::operator new(sizeof(TextureMap));
	new (&t->data) std::vector<uint8_t>(texdata + T.offset, texdata + T.offset + (size_t)T.width * T.height * 3);
	t->width = T.width;
	t->height = T.height;
	return t;
}

Material* material_from_arena(const rt_material& m, const std::vector<TextureMap*>& tex)
{
	Material* M = new Material();
	MaterialParameter* ps[RT_NPARAM] = {&M->_ke, &M->_ka, &M->_ks, &M->_kd, &M->_kr, &M->_kt, &M->_bump};
	for (int k = 0; k < RT_NPARAM; k++) {
		ps[k]->_value = g3(m.k[k]);
		ps[k]->_textureMap = m.tex[k] >= 0 ? tex[m.tex[k]] : nullptr;
	}
	M->_shininess._value = glm::dvec3(m.shininess_raw);
	M->_shininess._textureMap = m.shininess_tex >= 0 ? tex[m.shininess_tex] : nullptr;
	M->_index._value = glm::dvec3(m.index_raw);
	M->_index._textureMap = m.index_tex >= 0 ? tex[m.index_tex] : nullptr;
	M->_refl = m.refl != 0;
	M->_trans = m.trans != 0;
	M->_recur = m.recur != 0;
	M->_spec = false;
	M->_both = M->_refl && M->_trans;
	return M;
}

} // namespace

extern "C" int ref_load_arena(const void* arena, size_t bytes)
{
	ensure_globals();
	(void)bytes;
	rt::ArenaView A(arena);
	const uint8_t* texdata = A.ptr<uint8_t>(RT_SEC_TEXDATA);
	std::vector<TextureMap*> tex;
	for (uint64_t k = 0; k < A.count(RT_SEC_TEXTURES); k++) tex.push_back(texture_from_arena(A.ptr<rt_texture>(RT_SEC_TEXTURES)[k], texdata));
	const rt_material* mats = A.ptr<rt_material>(RT_SEC_MATERIALS);
	const rt_globals& G = *A.ptr<rt_globals>(RT_SEC_GLOBALS);

	g_ui->cubemap.reset();
	g_ui->m_usingCubeMap = false;
	if (G.has_cubemap) {
		g_ui->setCubeMap(new CubeMap());
		for (int k = 0; k < 6; k++) {
			// each face gets its own copy: CubeMap owns its TextureMaps
			const rt_texture& T = A.ptr<rt_texture>(RT_SEC_TEXTURES)[G.cubemap[k]];
			g_ui->cubemap->setNthMap(k, texture_from_arena(T, texdata));
		}
		g_ui->useCubeMap(true);
	}

	Scene* sc = new Scene();
	const rt_camera& C = *A.ptr<rt_camera>(RT_SEC_CAMERA);
	sc->camera.eye = g3(C.eye);
	sc->camera.look = g3(C.look);
	sc->camera.u = g3(C.u);
	sc->camera.v = g3(C.v);
	sc->camera.aspectRatio = C.aspect;
	sc->ambientIntensity = g3(G.ambient);

	const rt_light* L = A.ptr<rt_light>(RT_SEC_LIGHTS);
	for (uint64_t k = 0; k < A.count(RT_SEC_LIGHTS); k++) {
		if (L[k].type == RT_LIGHT_POINT) {
			sc->add(new PointLight(sc, g3(L[k].position), g3(L[k].color), (float)L[k].atten[0], (float)L[k].atten[1], (float)L[k].atten[2]));
		} else if (L[k].type == RT_LIGHT_DIRECTIONAL) {
			DirectionalLight* d = new DirectionalLight(sc, g3(L[k].orientation), g3(L[k].color));
			d->orientation = g3(L[k].orientation); // the ctor would normalise a second time
			sc->add(d);
		} else {
			return 2; // area / spot lights: load those scenes from their .ray text
		}
	}

	const rt_object* O = A.ptr<rt_object>(RT_SEC_OBJECTS);
	const rt_mesh* Ms = A.ptr<rt_mesh>(RT_SEC_MESHES);
	const rt_face* F = A.ptr<rt_face>(RT_SEC_FACES);
	const double* V = A.ptr<double>(RT_SEC_VERTS);
	const double* VN = A.ptr<double>(RT_SEC_VNORMALS);
	const int32_t* VM = A.ptr<int32_t>(RT_SEC_VMATS);
	for (uint64_t k = 0; k < A.count(RT_SEC_OBJECTS); k++) {
		const rt_object& o = O[k];
		TransformNode* tn = sc->transformRoot.createChild(glm::dmat4x4(1.0));
		for (int c = 0; c < 4; c++)
			for (int r = 0; r < 4; r++) tn->inverse[c][r] = o.inv[c * 4 + r];
		for (int c = 0; c < 3; c++)
			for (int r = 0; r < 3; r++) tn->normi[c][r] = o.normi[c * 3 + r];
		Material* mat = material_from_arena(mats[o.material], tex);
		Geometry* g = nullptr;
		switch (o.type) {
		case RT_OBJ_SPHERE: g = new Sphere(sc, mat); break;
		case RT_OBJ_BOX: g = new Box(sc, mat); break;
		case RT_OBJ_SQUARE: g = new Square(sc, mat); break;
		case RT_OBJ_CYLINDER: g = new Cylinder(sc, mat); break;
		case RT_OBJ_CONE: {
			Cone* cn = new Cone(sc, mat, o.cone[0], o.cone[1], o.cone[2], o.capped != 0);
			cn->beta_squared = o.cone[3];
			cn->gamma = o.cone[4];
			g = cn;
			break;
		}
		case RT_OBJ_TRIMESH: {
			const rt_mesh& M = Ms[o.mesh];
			Trimesh* tm = new Trimesh(sc, mat, tn);
			tm->vertices.reserve(M.vert_cnt);
			for (uint64_t v = 0; v < M.vert_cnt; v++) tm->vertices.push_back(g3(V + 3 * (M.vert_off + v)));
			if (M.vnorm_off >= 0)
				for (uint64_t v = 0; v < M.vert_cnt; v++) tm->normals.push_back(g3(VN + 3 * (M.vnorm_off + v)));
			if (M.vmat_off >= 0)
				for (uint64_t v = 0; v < M.vert_cnt; v++) tm->materials.push_back(material_from_arena(mats[VM[M.vmat_off + v]], tex));
			tm->faces.reserve(M.face_cnt);
			for (uint64_t f = 0; f < M.face_cnt; f++) {
				const rt_face& fr = F[M.face_off + f];
				TrimeshFace* face = new TrimeshFace(sc, new Material(*mat), tm, fr.ids[0], fr.ids[1], fr.ids[2]);
				face->setTransform(tn);
				tm->faces.push_back(face);
			}
			tm->kdtree = tree_from_arena(&tm->faces, A.ptr<rt_node>(RT_SEC_MESH_NODES) + M.node_off, A.ptr<int32_t>(RT_SEC_MESH_ITEMS) + M.item_off, 0);
			g = tm;
			break;
		}
		default: return 3;
		}
		g->setTransform(tn);
		g->bounds = BoundingBox(g3(o.bbox), g3(o.bbox + 3));
		sc->objects.emplace_back(g);
	}
	sc->kdtree = tree_from_arena(&sc->objects, A.ptr<rt_node>(RT_SEC_SCENE_NODES), A.ptr<int32_t>(RT_SEC_SCENE_ITEMS), 0);

	delete theRayTracer;
	theRayTracer = new RayTracer();
	g_ui->setRayTracer(theRayTracer);
	theRayTracer->scene.reset(sc);
	g_ids.clear();
	return 0;
}
