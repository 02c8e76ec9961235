/* rt_arena.h - the flattened scene ("arena") that crosses the C-ABI boundary.
 *
 * One contiguous, position-independent byte buffer: a header, a section table and 16-byte
 * aligned sections of plain-old-data records.  It is the POD restatement of the reference's
 * pointer graph (Scene -> Geometry/Trimesh/TrimeshFace/KdTree/Material/Light/Camera,
 * reference ray/src/scene/scene.h:224-302, kdTree.h:23-78, SceneObjects/trimesh.h:18-130,
 * scene/material.h:148-278, scene/light.h:16-164, scene/camera.h:8-35) - SURVEY.md App. E.
 *
 * The trees are the reference's own median-split BVHs ("KdTree"), stored in DFS pre-order
 * (child[0] subtree directly follows its parent), so "candidate rank" - the position of an item in
 * the reference's potenlist (kdTree.h:100-117) - is simply the item's position in the *_ITEMS
 * section.  Everything a ray needs at run time is in here; nothing is recomputed on the device
 * except derived traversal layouts (see csrc/rt_device_scene.h).
 *
 * Writers: the product's host loader (csrc/host/flatten.cpp) and, independently, the oracle
 * harness that walks the real reference Scene (oracle/ref_harness.cpp).  tests/ compare the two
 * section by section, bit for bit.
 */
#ifndef RT_ARENA_H
#define RT_ARENA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ARENA_MAGIC 0x4152545242323030ull /* "00B2RTRA" */
#define RT_ARENA_VERSION 4u
#define RT_ARENA_ALIGN 16u

/* section ids */
enum rt_section_id {
	RT_SEC_CAMERA = 1,     /* rt_camera[1] */
	RT_SEC_GLOBALS = 2,    /* rt_globals[1] */
	RT_SEC_LIGHTS = 3,     /* rt_light[] */
	RT_SEC_MATERIALS = 4,  /* rt_material[] (deduplicated, first-seen order) */
	RT_SEC_TEXTURES = 5,   /* rt_texture[] */
	RT_SEC_TEXDATA = 6,    /* uint8 RGB texels, row 0 = bottom scanline (BMP order, fileio/bitmap.cpp:70-90) */
	RT_SEC_OBJECTS = 7,    /* rt_object[] in Scene::objects order (= obj_id of a hit) */
	RT_SEC_SCENE_NODES = 8,/* rt_node[] scene-level tree, DFS pre-order */
	RT_SEC_SCENE_ITEMS = 9,/* int32 object ids, leaf order */
	RT_SEC_MESHES = 10,    /* rt_mesh[] */
	RT_SEC_MESH_NODES = 11,/* rt_node[] all per-mesh trees, concatenated */
	RT_SEC_MESH_ITEMS = 12,/* int32 face ids (mesh-local), leaf order, concatenated */
	RT_SEC_FACES = 13,     /* rt_face[] (non-degenerate faces only, Trimesh::faces order) */
	RT_SEC_VERTS = 14,     /* double[3] per vertex, concatenated */
	RT_SEC_VNORMALS = 15,  /* double[3] per vertex for meshes with normals */
	RT_SEC_VMATS = 16,     /* int32 material id per vertex for meshes with per-vertex materials */
	RT_SEC_MAX = 17
};

typedef struct rt_section {
	uint32_t id;
	uint32_t elem_size;
	uint64_t offset; /* from arena start, multiple of RT_ARENA_ALIGN */
	uint64_t count;  /* elements */
} rt_section;

typedef struct rt_arena_header {
	uint64_t magic;
	uint32_t version;
	uint32_t n_sections;
	uint64_t total_bytes;
	uint64_t reserved;
	rt_section sections[RT_SEC_MAX];
} rt_arena_header;

/* Camera after Camera::update() (scene/camera.cpp:93-99): rayThrough uses eye/look/u/v only. */
typedef struct rt_camera {
	double eye[3];
	double look[3];
	double u[3];
	double v[3];
	double aspect;
	double pad;
} rt_camera;

typedef struct rt_globals {
	double ambient[3]; /* Scene::ambientIntensity */
	int32_t n_objects;
	int32_t n_lights;
	int32_t cubemap[6]; /* texture ids +x -x +y -y +z -z, or -1 when no cube map is loaded */
	int32_t has_cubemap;
	int32_t pad;
} rt_globals;

enum rt_light_type { RT_LIGHT_DIRECTIONAL = 0, RT_LIGHT_POINT = 1, RT_LIGHT_AREA_RECT = 2, RT_LIGHT_AREA_CIRC = 3, RT_LIGHT_SPOT = 4 };

typedef struct rt_light {
	int32_t type;
	int32_t pad;
	double color[3];
	double position[3];    /* point/area/spot */
	double orientation[3]; /* directional: normalize(direction); area/spot: normalize(ori) */
	double atten[3];       /* constant, linear, quadratic: float members promoted to double (light.h:86-88) */
	double u[3];           /* AreaLightRect: normalize(updir) */
	double v[3];           /* AreaLightRect: cross(ori_raw, u) - light.h:127 uses the un-normalised ctor argument */
	double w, h;           /* AreaLightRect */
	double radius;         /* AreaLightCirc / SpotLight */
	double ang_tan;        /* SpotLight: tan(ang/360*PI) */
	double offset;         /* SpotLight: ang_tan * r */
	double pad2;
} rt_light;

/* parameter slots of rt_material.k / .tex */
enum { RT_KE = 0, RT_KA = 1, RT_KS = 2, RT_KD = 3, RT_KR = 4, RT_KT = 5, RT_BUMP = 6, RT_NPARAM = 7 };

typedef struct rt_material {
	double k[RT_NPARAM][3]; /* MaterialParameter::_value */
	double shininess;       /* intensityValue of the constant: (0.299v+0.587v)+0.114v (material.cpp:148-158) */
	double index;           /* same evaluation for the index of refraction */
	int32_t tex[RT_NPARAM]; /* texture id or -1 */
	int32_t shininess_tex;  /* texture id or -1 (then shininess = 128*intensity(texel)) */
	int32_t index_tex;
	int32_t refl, trans, recur; /* Material::_refl/_trans/_recur with indeterminate pinned to 0 */
	int32_t pad[3];
	double shininess_raw;   /* MaterialParameter::_value[0] of _shininess / _index: what per-vertex material */
	double index_raw;       /* interpolation (trimesh.cpp:157-163, material.h:177-189,281-293) sums before intensityValue */
} rt_material;

typedef struct rt_texture {
	int32_t width, height;
	uint64_t offset; /* into RT_SEC_TEXDATA */
} rt_texture;

enum rt_object_type { RT_OBJ_SPHERE = 0, RT_OBJ_BOX = 1, RT_OBJ_SQUARE = 2, RT_OBJ_CYLINDER = 3, RT_OBJ_CONE = 4, RT_OBJ_TRIMESH = 5 };

typedef struct rt_object {
	int32_t type;
	int32_t material; /* material id (for a trimesh: the material copied into every face) */
	int32_t mesh;     /* index into RT_SEC_MESHES or -1 */
	int32_t capped;   /* cone only (cylinders are always capped, Cylinder.h:11) */
	double bbox[6];   /* world bounds: min xyz, max xyz (Geometry::ComputeBoundingBox, scene.cpp:78-116) */
	double inv[16];   /* TransformNode::inverse, column-major m[col][row] */
	double normi[9];  /* TransformNode::normi = transpose(inverse(mat3(xform))), column-major */
	double cone[6];   /* height, b_radius, t_radius, beta_squared, gamma, pad (Cone.h:10-37) */
	double pad;
} rt_object;

/* One tree node.  Internal: child[0] is node+1, child[1] is `right` (relative to the tree's
 * first node); leaf: right = -1 and items [first, first+count) of the tree's item array. */
typedef struct rt_node {
	double bmin[3];
	double bmax[3];
	int32_t right;
	int32_t first;
	int32_t count;
	int32_t pad;
} rt_node;

typedef struct rt_mesh {
	uint64_t node_off, node_cnt; /* into RT_SEC_MESH_NODES */
	uint64_t item_off, item_cnt; /* into RT_SEC_MESH_ITEMS */
	uint64_t face_off, face_cnt; /* into RT_SEC_FACES */
	uint64_t vert_off, vert_cnt; /* into RT_SEC_VERTS */
	int64_t vnorm_off;           /* into RT_SEC_VNORMALS or -1 (Trimesh::normals empty) */
	int64_t vmat_off;            /* into RT_SEC_VMATS or -1 (Trimesh::materials empty) */
	int32_t object;              /* owning object id */
	int32_t pad[3];
} rt_mesh;

typedef struct rt_face {
	int32_t ids[3]; /* mesh-local vertex indices */
	int32_t pad;
	double normal[3]; /* normalize(cross(b-a, c-a)) (trimesh.h:123-125) */
	double pad2;
} rt_face;

#ifdef __cplusplus
}
#endif
#endif /* RT_ARENA_H */
