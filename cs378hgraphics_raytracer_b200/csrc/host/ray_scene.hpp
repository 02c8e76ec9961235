// ray_scene.hpp - host-side description of a parsed .ray scene, before tree build / flattening.
//
// This is the loader's own model of what the reference keeps as a pointer graph
// (Scene / TransformNode / Geometry / Trimesh / Material / Light / Camera; reference
// ray/src/scene/scene.h:64-302, SceneObjects/trimesh.h:18-130, scene/material.h:78-278,
// scene/light.h:16-164, scene/camera.h:8-35).  Indices replace pointers everywhere.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>
#include "hmath.hpp"

namespace rs {

using hm::M3;
using hm::M4;
using hm::V3;

struct Texture {
	int width = 0, height = 0;
	std::vector<uint8_t> rgb; // width*height*3, row 0 = bottom scan line (fileio/bitmap.cpp:70-90)
};

// MaterialParameter (material.h:78-146): a constant or a texture
struct MatParam {
	V3 value;
	int tex = -1;
	MatParam() {}
	explicit MatParam(V3 v) : value(v) {}
	explicit MatParam(double s) : value(s, s, s) {}
	bool is_zero() const { return hm::length(value) == 0.0; } // material.h:129
};

// Material (material.h:148-278); flags start false (the reference leaves _recur indeterminate
// until setBools runs; the oracle build pins that to false)
struct Material {
	MatParam ke, ka, ks, kd, kr, kt, bump;
	MatParam shininess{0.0}, index{1.0}, gloss{0.0};
	bool refl = false, trans = false, recur = false;
	void set_bools() // material.h:272-276
	{
		refl = !kr.is_zero();
		trans = !kt.is_zero();
		recur = refl || trans;
	}
};

struct Transform {
	M4 xform, inverse;
	M3 normi;
};

enum ObjType { SPHERE = 0, BOX = 1, SQUARE = 2, CYLINDER = 3, CONE = 4, TRIMESH = 5 };

struct Mesh {
	std::vector<V3> vertices, normals;
	std::vector<Material> vmaterials;    // per-vertex materials (may be empty)
	std::vector<int32_t> face_ids;       // 3 per kept (non-degenerate) face
	std::vector<V3> face_normals;        // per kept face
};

struct Object {
	ObjType type = SPHERE;
	M4 inverse; // TransformNode::inverse / normi of the node the object hangs under
	M3 normi;
	Material material;
	bool capped = true;
	double height = 1.0, b_radius = 1.0, t_radius = 0.0, beta_squared = 0.0, gamma = 0.0; // cone
	int mesh = -1;
	V3 bmin, bmax; // world bounds (Geometry::ComputeBoundingBox, scene.cpp:78-116)
};

enum LightType { L_DIRECTIONAL = 0, L_POINT = 1, L_AREA_RECT = 2, L_AREA_CIRC = 3, L_SPOT = 4 };
struct Light {
	LightType type = L_POINT;
	V3 color, position, orientation, u, v;
	float a = 0.0f, b = 0.0f, c = 1.0f; // float like the reference (light.h:86-88)
	double w = 0, h = 0, radius = 0, ang_tan = 0, offset = 0;
};

struct Camera {
	M3 m; // identity
	double normalized_height = 1.0, aspect = 1.0;
	V3 eye{0, 0, 0}, look{0, 0, -1}, u{1, 0, 0}, v{0, 1, 0};
	void update() // camera.cpp:93-99
	{
		u = hm::mul(m, V3(1, 0, 0)) * normalized_height * aspect;
		v = hm::mul(m, V3(0, 1, 0)) * normalized_height;
		look = hm::mul(m, V3(0, 0, -1));
	}
};

struct Scene {
	Camera camera;
	V3 ambient;
	std::vector<Light> lights;
	std::vector<Object> objects;       // Scene::objects order = hit obj id
	std::vector<Mesh> meshes;
	std::vector<Texture> textures;
	std::map<std::string, int> texture_cache; // Scene::getTexture (scene.cpp:199-206)
	int cubemap[6] = {-1, -1, -1, -1, -1, -1};
	bool has_cubemap = false;
};

// ray_parser.cpp: Tokenizer + Parser::parseScene re-statement.  Throws std::runtime_error.
void parse_ray_file(const std::string& path, Scene& out);
// image_io.cpp
bool read_bmp24(const std::string& path, Texture& out);
bool read_image(const std::string& path, Texture& out);  // by extension: .png or .bmp (image_io.cpp)
bool write_image(const std::string& path, int w, int h, const unsigned char* rgb);
int load_texture(Scene& sc, const std::string& path); // cached; throws on failure

// scene_flatten.cpp: tree build (kdTree.h:23-78) + arena (include/rt_arena.h); malloc'ed
void* flatten(const Scene& sc, size_t* bytes_out);

} // namespace rs
