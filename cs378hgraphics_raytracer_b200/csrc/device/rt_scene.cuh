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

// Traversal node: one 128-byte record per INTERNAL node of a reference tree holding BOTH children's
// boxes, so one step of the walk is one aligned 128-byte read (eight 128-bit loads) instead of three
// scattered node reads.  A child reference is >= 0 for an internal node (index into the same
// array, relative to the tree) or a leaf reference < 0 that encodes the leaf's item range.
// Child 0 / child 1 keep the reference's order (kdTree.h:111-112).
struct alignas(128) DNode2 {
	double lo0[3], hi0[3]; // child[0] bounds
	double lo1[3], hi1[3]; // child[1] bounds
	int32_t c0, c1;
	int32_t pad[6];
};
// The same node with the boxes rounded to float (round to nearest): what the walk reads first.
// Half the bytes per step; the double node is fetched only for a child whose float test is
// ambiguous (rt_prims.cuh: box2_hit).
struct alignas(64) DNode2f {
	float lo0[3], hi0[3];
	float lo1[3], hi1[3];
	int32_t c0, c1;
	int32_t pad[2];
};
// One scene-level leaf item (position = candidate rank): the object's world bounding box rounded to
// float plus its index, so that a candidate the ray certainly misses costs one 32-byte read and a
// float slab test instead of the rt_object record and the double test (rt_prims.cuh: box_miss_f).
struct alignas(32) DItemBox {
	float lo[3], hi[3];
	int32_t obj;
	int32_t type; // rt_object_type of obj: the split shadow stage queues candidates by primitive kind without reading the object
};
// One scene-level leaf item again: what a float test in the primitive's OBJECT space needs (rt_prims.cuh: prim_miss_f) - the
// 3x3 part of the object's inverse transform and the object-space position of a reference point (the centre of the item's
// float box), both rounded to float.  A candidate whose box the ray enters but whose shape its line certainly misses - about
// half of them for rotated boxes and cylinders, whose world boxes are loose - then costs this 64-byte record and ~60 float
// instructions instead of the 248-byte rt_object, the world -> object transform in double precision and intersectLocal.
// kind 0 = no such test for this item (trimesh, cone, a transform that is not rotation x uniform scale, degenerate numbers).
enum { RT_PF_NONE = 0, RT_PF_BOX = 1, RT_PF_SPHERE = 2, RT_PF_CYLINDER = 3, RT_PF_SQUARE = 4 };
struct alignas(64) DPrimF {
	float m[9];  // row-major: object-space direction = m * world direction
	float lc[3]; // object-space coordinates of the reference point, relative to the centre of the shape's own bounding box
	int32_t kind;
	int32_t pad[3];
};

// Shared-memory staging of the scene tree's top (north star item 1; A/B in profiles/r2_variants.txt): the first RT_TOP_K internal
// nodes in BREADTH-FIRST order (the arrays are pre-order, so the top levels are not contiguous) as a table of DNode2f whose child
// references to other table entries carry RT_REF_TOP | table index; pad[0] = the node's index in snodes2 (for the double boxes).
#ifndef RT_TOP_K
#define RT_TOP_K 0 // 0 = off
#endif
#define RT_REF_TOP ((int32_t)0x40000000)
#define RT_REF_MESH_EXIT ((int32_t)0x80000000)
#define RT_REF_NONE ((int32_t)0x80000001)
RT_HD int32_t leaf_ref(int32_t first, int32_t count) { return -(1 + ((first << 2) | count)); }
RT_HD int32_t leaf_first(int32_t ref) { return (-(ref + 1)) >> 2; }
RT_HD int32_t leaf_count(int32_t ref) { return (-(ref + 1)) & 3; }

struct DTree {
	double lo[3], hi[3]; // root bounds
	int32_t root;        // root reference (a leaf reference for trees of <= 3 items)
	int32_t pad;
	int64_t node_off;    // first DNode2 of this tree
};

struct DScene {
	const DNode2* snodes2;
	DTree stree;
	const DNode2* mnodes2;
	const DNode2f* snodes2f; // float images, same indexing
	const DNode2f* mnodes2f;
	const DItemBox* sboxf;   // per scene leaf item
	const DPrimF* sprimf;    // per scene leaf item (same index)
	const DNode2f* stop;     // top table of the scene tree (RT_TOP_K entries, n_top used); see RT_REF_TOP
	int32_t n_top;
	int32_t pad_top_;
	const DTree* mtrees; // per mesh
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
	int32_t kt_unit; // every material's kt lies in [0,1]^3 (texels always do): shadow attenuations are finite
	int32_t pad_;
};

// A ray plus what traversal derives from it once.
struct Ray {
	d3 o, d;
};

} // namespace rtd
