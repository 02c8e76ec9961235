// rt_scene.cuh - the scene as the kernels see it: raw pointers into HBM copies of the arena
// sections (include/rt_arena.h) plus layouts derived at upload time.
//
// Derived layout (built by rt_api.cu, see DESIGN.md "data layout"):
//   tris[]  one 96-byte record per mesh leaf ITEM (not per face): v0,v1,v2,n gathered so a leaf's
//           1-3 triangles are contiguous and read with 128-bit loads, no index indirection.
//           tris[mesh.item_off + k] describes face mitems[mesh.item_off + k].
#pragma once
#include "rt_arena.h"
#include "rt_b200.h"
#include "rt_math.cuh"

namespace rtd {

struct DTri {
	double v0[3], v1[3], v2[3], n[3];
};

struct DScene {
	const rt_node* snodes;
	const int32_t* sitems;
	const rt_object* objs;
	const rt_mesh* meshes;
	const double* mesh_slack; // pruning slack per mesh (object space)
	const rt_node* mnodes;
	const int32_t* mitems;
	const DTri* tris;
	const rt_face* faces;
	const double* verts;
	const double* vnormals;
	const int32_t* vmats;
	const rt_material* mats;
	const rt_light* lights;
	const rt_texture* texs;
	const uint8_t* texdata;
	rt_camera cam;
	rt_globals glob;
	double scene_slack; // pruning slack of the scene-level tree (world space)
	int32_t n_lights;
	int32_t n_objs;
};

// A ray plus what traversal derives from it once.
struct Ray {
	d3 o, d;
};

} // namespace rtd
