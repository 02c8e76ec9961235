// Link-time stubs for building the UNMODIFIED reference ray tracer (read in place from
// /root/reference/ray/src) without FLTK / OpenGL / libpng.  TEST INFRASTRUCTURE ONLY.
//
//  * TraceUI::m_debug           - defined in the reference's GUI code (not compiled here)
//  * glDraw / glDrawLocal       - reference ui/glObjects.cpp:22-560 (OpenGL debug view;
//                                 "unavailable on this box") become no-ops
//  * readPNG / writePNG         - reference fileio/pngimage.cpp (libpng absent); reading
//                                 fails (use .bmp assets), writing falls back to BMP
//  * operator new               - calloc-backed, so the reference's indeterminate
//                                 Material flags (_recur/_spec/_both, scene/material.h:152-164)
//                                 are deterministically false
#include <cstdlib>
#include <new>
#include <vector>
#include <stdint.h>

#include "scene/scene.h"
#include "scene/light.h"
#include "SceneObjects/Box.h"
#include "SceneObjects/Cone.h"
#include "SceneObjects/Cylinder.h"
#include "SceneObjects/Sphere.h"
#include "SceneObjects/Square.h"
#include "SceneObjects/trimesh.h"
#include "fileio/bitmap.h"
#include "ui/TraceUI.h"

bool TraceUI::m_debug = false;

void Scene::glDraw(int, bool, bool) const {}
void Geometry::glDraw(int, bool, bool) const {}
void SceneObject::glDraw(int, bool, bool) const {}
void Sphere::glDrawLocal(int, bool, bool) const {}
void Box::glDrawLocal(int, bool, bool) const {}
void Cone::glDrawLocal(int, bool, bool) const {}
void Cylinder::glDrawLocal(int, bool, bool) const {}
void Square::glDrawLocal(int, bool, bool) const {}
void Trimesh::glDrawLocal(int, bool, bool) const {}
void PointLight::glDraw(GLenum) const {}
void PointLight::glDraw() const {}
void DirectionalLight::glDraw(GLenum) const {}
void DirectionalLight::glDraw() const {}

std::vector<uint8_t> readPNG(const char*, int& width, int& height)
{
	width = height = 0;
	return std::vector<uint8_t>();
}
void writePNG(const char* iname, int width, int height, const void* data)
{
	writeBMP(iname, width, height, data);
}

void* operator new(std::size_t n)
{
	void* p = std::calloc(1, n ? n : 1);
	if (!p) throw std::bad_alloc();
	return p;
}
void* operator new[](std::size_t n)
{
	void* p = std::calloc(1, n ? n : 1);
	if (!p) throw std::bad_alloc();
	return p;
}
void operator delete(void* p) noexcept { std::free(p); }
void operator delete[](void* p) noexcept { std::free(p); }
void operator delete(void* p, std::size_t) noexcept { std::free(p); }
void operator delete[](void* p, std::size_t) noexcept { std::free(p); }
