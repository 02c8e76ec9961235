// rt_material.cuh - hit finalisation (normal, uv, barycentrics) and material evaluation.
#pragma once
#include "rt_trace.cuh"

namespace rtd {

// TextureMap::getPixelAt / getMappedValue (reference scene/material.cpp:84-138): bilinear over
// (w-1, h-1), out-of-range texel = 0, out-of-range uv = pure green.
RT_HD d3 tex_pixel(const DScene& S, const rt_texture& T, int x, int y)
{
	if (0 <= x && x < T.width && 0 <= y && y < T.height) {
		const uint8_t* p = S.texdata + T.offset + (size_t)(x + y * T.width) * 3;
		return mk3((double)p[0], (double)p[1], (double)p[2]);
	}
	return mk3(0.0, 0.0, 0.0);
}
RT_COLD d3 tex_sample(const DScene& S, int tex, d2 uv)
{
	const rt_texture T = S.texs[tex];
	double x = uv.x, y = uv.y;
	if (0.0 <= x && x <= 1.0 && 0.0 <= y && y <= 1.0) {
		x *= T.width - 1;
		y *= T.height - 1;
		const int ix = (int)x, iy = (int)y;
		x -= ix;
		y -= iy;
		d3 rows[2];
		for (int i = 0; i < 2; i++) {
			const d3 pl = tex_pixel(S, T, i + ix, 0 + iy);
			const d3 pr = tex_pixel(S, T, i + ix, 1 + iy);
			rows[i] = y * (pr - pl) + pl;
		}
		const d3 v = x * (rows[1] - rows[0]) + rows[0];
		// v / 255.0 by the FMA-corrected reciprocal: the same correctly rounded quotients as the
		// division (rt_math.cuh: div_by), a third of the instructions
		const double y255 = 1.0 / 255.0;
		return mk3(div_by(v.x, 255.0, y255, 1201u), div_by(v.y, 255.0, y255, 1201u), div_by(v.z, 255.0, y255, 1201u));
	}
	return mk3(0.0, 1.0, 0.0);
}

RT_HD d3 mat_param(const DScene& S, const rt_material& m, int k, d2 uv)
{
	return m.tex[k] >= 0 ? tex_sample(S, m.tex[k], uv) : ld3(m.k[k]);
}
RT_HD double intensity(d3 v) { return (0.299 * v.x) + (0.587 * v.y) + (0.114 * v.z); } // material.cpp:153-157

struct MatEval {
	d3 ke, ka, ks, kd, kr, kt;
	double shininess, index;
	bool refl, trans, recur;
};

// The Material an isect carries (reference scene/ray.cpp:6-9): the object's own, or for meshes
// with per-vertex materials the barycentric blend built in trimesh.cpp:157-163 from Material()
// (flags stay false, index starts at 1, textures are not carried over; material.h:152-189,281-293).
RT_HD int mesh_vmat_base(const DScene& S, const Hit& h)
{
	if (h.face < 0) return -1;
	const rt_mesh& M = S.meshes[S.objs[h.obj].mesh];
	return M.vmat_off >= 0 ? (int)M.vmat_off : -1;
}

RT_HD d3 blend3(const double* a, const double* b, const double* c, d3 w)
{
	return mk3(((0.0 + a[0] * w.x) + b[0] * w.y) + c[0] * w.z,
	           ((0.0 + a[1] * w.x) + b[1] * w.y) + c[1] * w.z,
	           ((0.0 + a[2] * w.x) + b[2] * w.y) + c[2] * w.z);
}

// kt only (shadow rays need nothing else)
RT_COLD void eval_kt(const DScene& S, const Hit& h, d3& kt, bool& trans)
{
	const int vb = mesh_vmat_base(S, h);
	if (vb >= 0) {
		const rt_face& f = S.faces[S.meshes[S.objs[h.obj].mesh].face_off + h.face];
		const rt_material& a = S.mats[S.vmats[vb + f.ids[0]]];
		const rt_material& b = S.mats[S.vmats[vb + f.ids[1]]];
		const rt_material& c = S.mats[S.vmats[vb + f.ids[2]]];
		kt = blend3(a.k[RT_KT], b.k[RT_KT], c.k[RT_KT], h.bary);
		trans = false;
		return;
	}
	const rt_material& m = S.mats[S.objs[h.obj].material];
	kt = mat_param(S, m, RT_KT, h.uv);
	trans = m.trans != 0;
}

RT_COLD void eval_material(const DScene& S, const Hit& h, MatEval& e)
{
	const int vb = mesh_vmat_base(S, h);
	if (vb >= 0) {
		const rt_face& f = S.faces[S.meshes[S.objs[h.obj].mesh].face_off + h.face];
		const rt_material& a = S.mats[S.vmats[vb + f.ids[0]]];
		const rt_material& b = S.mats[S.vmats[vb + f.ids[1]]];
		const rt_material& c = S.mats[S.vmats[vb + f.ids[2]]];
		e.ke = blend3(a.k[RT_KE], b.k[RT_KE], c.k[RT_KE], h.bary);
		e.ka = blend3(a.k[RT_KA], b.k[RT_KA], c.k[RT_KA], h.bary);
		e.ks = blend3(a.k[RT_KS], b.k[RT_KS], c.k[RT_KS], h.bary);
		e.kd = blend3(a.k[RT_KD], b.k[RT_KD], c.k[RT_KD], h.bary);
		e.kr = blend3(a.k[RT_KR], b.k[RT_KR], c.k[RT_KR], h.bary);
		e.kt = blend3(a.k[RT_KT], b.k[RT_KT], c.k[RT_KT], h.bary);
		const double sh = ((0.0 + a.shininess_raw * h.bary.x) + b.shininess_raw * h.bary.y) + c.shininess_raw * h.bary.z;
		const double ix = ((1.0 + a.index_raw * h.bary.x) + b.index_raw * h.bary.y) + c.index_raw * h.bary.z;
		e.shininess = intensity(mk3(sh, sh, sh));
		e.index = intensity(mk3(ix, ix, ix));
		e.refl = e.trans = e.recur = false;
		return;
	}
	const rt_material& m = S.mats[S.objs[h.obj].material];
	e.ke = mat_param(S, m, RT_KE, h.uv);
	e.ka = mat_param(S, m, RT_KA, h.uv);
	e.ks = mat_param(S, m, RT_KS, h.uv);
	e.kd = mat_param(S, m, RT_KD, h.uv);
	e.kr = mat_param(S, m, RT_KR, h.uv);
	e.kt = mat_param(S, m, RT_KT, h.uv);
	// material.h:204-211: mapped shininess is renormalised to 0..128
	e.shininess = m.shininess_tex >= 0 ? 128.0 * intensity(tex_sample(S, m.shininess_tex, h.uv)) : m.shininess;
	e.index = m.index_tex >= 0 ? intensity(tex_sample(S, m.index_tex, h.uv)) : m.index;
	e.refl = m.refl != 0;
	e.trans = m.trans != 0;
	e.recur = m.recur != 0;
}

// Object-space normal of an analytic hit -> world (TransformNode::localToGlobalCoordsNormal,
// reference scene.h:108-111), including Box::computeNormal's bump map (Box.cpp:99-108).
RT_COLD d3 world_normal(const DScene& S, const rt_object& ob, d3 Nl, d2 uv)
{
	if (ob.type == RT_OBJ_BOX) {
		const rt_material& m = S.mats[ob.material];
		const d3 b = mat_param(S, m, RT_BUMP, uv);
		if (length(b) > 0.0001) Nl = normalize(b - mk3(0.5, 0.5, 0.5));
	}
	return normalize(mul33(ob.normi, Nl));
}

// Re-runs the winning primitive's test in its full form to obtain N, uv, barycentrics - the
// traversal itself only carries (t, object, leaf item).  Same arithmetic, same t.
RT_COLD void finalize_hit(const DScene& S, const d3& o, const d3& d, Hit& h)
{
	const rt_object& ob = S.objs[h.obj];
	LocalRay L;
	to_local(L, ob, o, d);
	LocalHit lh;
	lh.uv.x = lh.uv.y = 0.0; // isect() zero-initialises uvCoordinates (glm 0.9.8 default ctor)
	lh.N = mk3(0, 0, 0);
	h.bary = mk3(0, 0, 0);
	h.face = -1;
	h.boxface = -1;
	switch (ob.type) {
	case RT_OBJ_SPHERE: sphere_hit<true>(L.o, L.d, lh); break;
	case RT_OBJ_BOX: box_prim_hit<true>(L.o, L.d, lh, h.boxface); break;
	case RT_OBJ_SQUARE: square_hit<true>(L.o, L.d, lh); break;
	case RT_OBJ_CYLINDER: cylinder_hit<true>(L.o, L.d, lh); break;
	case RT_OBJ_CONE: cone_hit<true>(L.o, L.d, ob.cone, ob.capped, lh); break;
	case RT_OBJ_TRIMESH: {
		const rt_mesh& M = S.meshes[ob.mesh];
		const int fid = S.mitems[h.fpos];
		h.face = fid;
		double t;
		tri_hit<true>(S.tris[h.fpos], L.o, L.d, t, h.bary);
		const rt_face& f = S.faces[M.face_off + fid];
		if (M.vnorm_off >= 0) {
			const double* n0 = S.vnormals + 3 * (M.vnorm_off + f.ids[0]);
			const double* n1 = S.vnormals + 3 * (M.vnorm_off + f.ids[1]);
			const double* n2 = S.vnormals + 3 * (M.vnorm_off + f.ids[2]);
			lh.N = normalize(mk3(n0[0] * h.bary.x + n1[0] * h.bary.y + n2[0] * h.bary.z,
			                     n0[1] * h.bary.x + n1[1] * h.bary.y + n2[1] * h.bary.z,
			                     n0[2] * h.bary.x + n1[2] * h.bary.y + n2[2] * h.bary.z));
		} else {
			lh.N = ld3(f.normal);
		}
		break;
	}
	default: break;
	}
	h.uv = lh.uv;
	h.N = world_normal(S, ob, lh.N, lh.uv);
}

} // namespace rtd
