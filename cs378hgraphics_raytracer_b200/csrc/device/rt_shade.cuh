// rt_shade.cuh - camera rays, Phong shading and the one-bounce step of RayTracer::traceRay.
#pragma once
#include "rt_traverse.cuh"

namespace rtd {

// Camera::rayThrough (reference scene/camera.cpp:20-30)
RT_HD void camera_ray(const rt_camera& C, double x, double y, d3& o, d3& d)
{
	x -= 0.5;
	y -= 0.5;
	d = normalize(ld3(C.look) + x * ld3(C.u) + y * ld3(C.v));
	o = ld3(C.eye);
}

// The `lens`-th ray RayTracer::trace shoots for window point (x,y): lens 0 is the pinhole ray,
// lens 1..divs are the depth-of-field rays (reference RayTracer.cpp:35-79).  eye_shift moves the
// eye along +x (anaglyph second eye, RayTracer.cpp:92-99).
RT_COLD void primary_ray(const rt_camera& C, const rt_params& prm, double x, double y, int lens, double eye_shift, d3& o, d3& d)
{
	camera_ray(C, x, y, o, d);
	if (eye_shift != 0.0) o = ld3(C.eye) + mk3(eye_shift, 0.0, 0.0);
	if (lens == 0) return;
	const d3 eye = o;
	const double fd = gmax(prm.dof_fd, 1.0);
	const d3 fp_n = -d;
	const d3 fp_pt = o + fd * d;
	double t = dot(fp_n, d);
	t = dot(fp_pt - o, fp_n) / t;
	const d3 dest = o + t * d;
	const double sz = prm.dof_apsz / 2;
	const int divs = prm.dof_divs;
	const double PI_ = 3.1415926535897932384626433832795028841971;
	const double baseAngle = PI_ / divs;
	const int i = lens - 1;
	double offsetAngle = PI_ / 2;
	offsetAngle = offsetAngle / divs + (i - 1) * baseAngle;
	const d3 offVec = (cos(offsetAngle) * ld3(C.v) + sin(offsetAngle) * ld3(C.u)) * sz;
	o = eye + offVec;
	d = normalize(dest - o);
}

// One light's term of Material::shade (reference scene/material.cpp:50-66) with PointLight /
// DirectionalLight (scene/light.cpp:56-73): everything except the shadow attenuation, plus the
// shadow ray Light::shadowAttenuation would trace (light.cpp:16-20: origin backed up along the
// view ray by 1e-10, aimed at the light).  term = (dattn * Lc) * (d_comp + s_comp).
RT_COLD void light_term(const rt_light& L, const MatEval& m, const d3& P, const d3& N, const d3& view, d3& term, d3& pb, d3& sdir)
{
	const bool point = L.type != RT_LIGHT_DIRECTIONAL;
	const d3 lpos = ld3(L.position);
	const d3 l_i = point ? normalize(lpos - P) : -ld3(L.orientation);
	const d3 l_r = l_i - (2 * dot(l_i, N)) * N;
	double dt = dot(l_i, N);
	if (m.trans) dt = fabs(dt);
	const d3 d_comp = m.kd * gmax(0.0, dt);
	const double sp = pow(gmax(0.0, dot(l_r, view)), m.shininess);
	const d3 s_comp = m.ks * sp;
	double dattn = 1.0;
	if (point) {
		const double dist = length(P - lpos);
		dattn = gclamp(1.0 / (L.atten[0] + L.atten[1] * dist + L.atten[2] * dist * dist), 0.0, 1.0);
	}
	pb = P - view * 0.0000000001;
	sdir = point ? normalize(lpos - pb) : -ld3(L.orientation);
	term = (dattn * ld3(L.color)) * (d_comp + s_comp);
}

// ---- area and spot lights (reference scene/light.cpp:76-149, light.h:100-164) ----------------------
// AreaLight::shadowAttenuation = (1 + sum over the valid samples i < softShadowRes of
// srsAttenuation(pb, normalize(pick(i) - pb))) * (1 / (softShadowRes - 1)) - the sum starts at 1 and the
// divisor is res - 1, as written.  Point and directional lights trace one ray.
RT_HD bool light_is_area(const rt_light& L) { return L.type >= RT_LIGHT_AREA_RECT; }
RT_HD int light_samples(const rt_light& L, const rt_params& prm) { return light_is_area(L) ? prm.soft_shadow_res : 1; }
RT_HD double light_sample_weight(const rt_light& L, const rt_params& prm) { return light_is_area(L) ? 1.0 / (prm.soft_shadow_res - 1) : 1.0; }

// SpotLight::validImpact(r, p) (light.cpp:148-150); true for the other kinds
RT_COLD bool light_valid(const rt_light& L, const d3& p)
{
	if (L.type != RT_LIGHT_SPOT) return true;
	const double PI_ = 3.1415926535897932384626433832795028841971;
	const d3 ori = ld3(L.orientation), lpos = ld3(L.position);
	const d3 dir = normalize(lpos - p);
	return (dot(dir, ori) <= 0) && (dot(normalize(p - (lpos - L.offset * ori)), ori) > cos(PI_ / 4));
}

// i-th sample: AreaLightRect::pick (an offset in the light's plane that is NOT moved to the light's
// position, light.cpp:101-104) or AreaLightCirc::pick (a ring of radius r/2, :113-136), then
// SpotLight::validImpact(r, pb, lp).  false = the sample is skipped.
RT_COLD bool light_sample(const rt_light& L, const rt_params& prm, int i, const d3& pb, d3& sdir)
{
	const double PI_ = 3.1415926535897932384626433832795028841971;
	const d3 ori = ld3(L.orientation), lpos = ld3(L.position);
	d3 lp;
	if (L.type == RT_LIGHT_AREA_RECT) {
		const double a = (radical_inverse2(i) - 0.5) * L.w, b = (0.0 - 0.5) * L.h;
		const d3 u = ld3(L.u), v = ld3(L.v);
		lp = mk3(u.x * a + v.x * b, u.y * a + v.y * b, u.z * a + v.z * b);
	} else {
		const double ang_rad = 2 * PI_ / prm.soft_shadow_res * i;
		const double dist = 0.5 * L.radius;
		const double x = cos(ang_rad) * dist, y = sin(ang_rad) * dist;
		const d3 ab = mk3(fabs(ori.x), fabs(ori.y), fabs(ori.z));
		d3 u;
		if (ab.x < ab.y && ab.x < ab.z) u = mk3(0.0, -ori.z, ori.y);
		else if (ab.y < ab.z) u = mk3(-ori.z, 0.0, ori.x);
		else u = mk3(-ori.y, ori.x, 0.0);
		u = normalize(u);
		const d3 v = cross(ori, u);
		lp = x * u + y * v + lpos;
	}
	if (L.type == RT_LIGHT_SPOT) {
		const d3 dir = normalize(lpos - pb);
		if (!((dot(dir, ori) <= 0) && (dot(normalize(pb - lp), ori) > cos(PI_ / 4)))) return false;
	}
	sdir = normalize(lp - pb);
	return true;
}

// Material::shade with the shadow rays traced inline.  shadow_rays counts Scene::intersectList calls.
RT_HD d3 shade(const DScene& S, const rt_params& prm, const MatEval& m, const d3& o, const d3& d, const Hit& h, unsigned& shadow_rays)
{
	const d3 P = o + h.t * d;
	d3 out = m.ke + m.ka * ld3(S.glob.ambient);
	for (int li = 0; li < S.n_lights; li++) {
		const rt_light& L = S.lights[li];
		d3 term, pb, sdir;
		light_term(L, m, P, h.N, d, term, pb, sdir);
		if (!light_is_area(L)) {
			shadow_rays++;
			// a term of exactly 0 gains 0 whatever the shadow ray finds (finite attenuation): counted, not walked
			if (S.kt_unit && term.x == 0.0 && term.y == 0.0 && term.z == 0.0) continue;
			const d3 sattn = shadow_attenuation(S, L, pb, sdir, prm);
			out = out + sattn * term;
			continue;
		}
		if (!light_valid(L, P)) continue; // shadowAttenuation returns 0 (light.cpp:78)
		d3 sattn = mk3(1.0, 1.0, 1.0);
		for (int i = 0; i < prm.soft_shadow_res; i++) {
			d3 sd;
			if (!light_sample(L, prm, i, pb, sd)) continue;
			shadow_rays++;
			sattn = sattn + shadow_attenuation(S, L, pb, sd, prm);
		}
		sattn = sattn * light_sample_weight(L, prm);
		out = out + sattn * term;
	}
	return out;
}

// A ray in flight.  weight multiplies everything this ray (and its descendants) contributes;
// the parent's "kt ^ (this ray's hit distance)" factor (RayTracer.cpp:149,162) is folded in as
// soon as this ray's own hit distance is known: mode 1 = clamped (reflection), 2 = raw (refraction).
struct RayTask {
	d3 o, d;
	d3 weight;
	d3 att_kt;
	int att_mode;
	int depth; // the `depth` argument of traceRay for this ray
	int kind;  // 0 camera, 1 reflection, 2 refraction
};

// CubeMap::getColor (reference scene/cubeMap.cpp:12-44) for a ray that hit nothing; ties leave
// map/d unset in the reference (UB), pinned by the oracle build to map 0, d = 0.
RT_COLD d3 miss_colour(const DScene& S, const rt_params& prm, const d3& rd)
{
	if (!(prm.use_cubemap && S.glob.has_cubemap)) return mk3(0.0, 0.0, 0.0);
	const double ax = fabs(rd.x), ay = fabs(rd.y), az = fabs(rd.z);
	const bool xy = ax >= ay, yz = ay >= az, zx = az >= ax;
	int map = 0;
	double scale = 0.5;
	d2 dd; dd.x = 0.0; dd.y = 0.0;
	if (xy && !zx) { scale /= ax; dd.x = rd.x > 0 ? rd.z : -rd.z; dd.y = rd.y; map = rd.x > 0 ? 0 : 1; }
	else if (yz && !xy) { scale /= ay; dd.x = rd.x; dd.y = rd.y > 0 ? rd.z : -rd.z; map = rd.y > 0 ? 2 : 3; }
	else if (zx && !yz) { scale /= az; dd.x = rd.z > 0 ? rd.x : -rd.x; dd.y = rd.y; map = rd.z > 0 ? 4 : 5; }
	dd.x = dd.x * scale + 0.5;
	dd.y = dd.y * scale + 0.5;
	return tex_sample(S, S.glob.cubemap[map], dd);
}

// weight of a ray once its own hit distance is known (RayTracer.cpp:149,162)
RT_COLD d3 apply_att(const d3& w, const d3& att_kt, int att_mode, double t)
{
	if (att_mode == 1) return w * gclamp(vpow(att_kt, t), 0.0, 1.0);
	if (att_mode == 2) return w * vpow(att_kt, t);
	return w;
}

// The reflection / refraction rays of RayTracer::traceRay (reference RayTracer.cpp:127-165) for a
// hit of ray (ro, rd) with weight w; depth is the already decremented depth.  Overlapping-objects
// mode is off: the medium on the far side of a surface is air (index 1, kt 1; material.cpp:17-21).
// kid[0] = reflection (valid if has_refl), kid[1] = refraction (valid if has_refr).
RT_COLD void spawn_children(const MatEval& m, const Hit& h, const d3& ro, const d3& rd, const d3& w, int depth, RayTask* kid, bool& has_refl, bool& has_refr)
{
	has_refl = has_refr = false;
	if (!(m.recur && depth > 0)) return;
	const bool leaving = dot(h.N, rd) >= 0;
	const double air_index = intensity(mk3(1.0, 1.0, 1.0)); // air.index(i) goes through intensityValue too
	const double idx_curr = leaving ? m.index : air_index, idx_next = leaving ? air_index : m.index;
	const d3 kt_curr = leaving ? m.kt : mk3(1.0, 1.0, 1.0), kt_next = leaving ? mk3(1.0, 1.0, 1.0) : m.kt;
	const bool next_trans = leaving ? true : m.trans;
	const d3 normal = (leaving ? -1.0 : 1.0) * h.N;
	const double c = -1 * dot(normal, rd);
	const double eta = next_trans ? idx_curr / idx_next : 0;
	const double radicand = 1 - eta * eta * (1 - c * c);
	const bool tir = next_trans && radicand < 0;
	if (m.refl || tir) {
		RayTask& k = kid[0];
		k.d = rd + (2 * c) * normal;
		k.o = ro + (h.t - RT_RAY_EPSILON) * rd;
		k.weight = w * m.kr;
		k.att_kt = kt_curr;
		k.att_mode = (kt_curr.x == 1.0 && kt_curr.y == 1.0 && kt_curr.z == 1.0) ? 0 : 1;
		k.depth = depth;
		k.kind = 1;
		has_refl = true;
	}
	if (next_trans && !tir) {
		RayTask& k = kid[1];
		k.o = ro + (h.t + RT_RAY_EPSILON) * rd;
		k.d = eta * rd + (eta * c - sqrt(radicand)) * normal;
		k.weight = w;
		k.att_kt = kt_next;
		k.att_mode = (kt_next.x == 1.0 && kt_next.y == 1.0 && kt_next.z == 1.0) ? 0 : 2;
		k.depth = depth;
		k.kind = 2;
		has_refr = true;
	}
}

// One call of RayTracer::traceRay without its recursion (reference RayTracer.cpp:108-174), shadow
// rays traced inline: returns the colour this ray adds to its camera sample and fills up to two
// children.  (The wavefront kernel uses the same pieces but defers the shadow rays to the pool.)
template <bool PRUNE>
RT_HD d3 trace_step(const DScene& S, const rt_params& prm, const RayTask& r, Hit& h, RayTask* child, int& n_child, unsigned& shadow_rays)
{
	n_child = 0;
	if (!(r.depth >= 0 && closest_hit<PRUNE>(S, r.o, r.d, h))) {
		// miss: the parent's exponent is the (pinned-to-zero) uninitialised length -> factor 1
		h.obj = -1;
		return r.weight * miss_colour(S, prm, r.d);
	}
	finalize_hit(S, r.o, r.d, h);
	const d3 w = apply_att(r.weight, r.att_kt, r.att_mode, h.t);
	const int depth = r.depth - 1;
	MatEval m;
	eval_material(S, h, m);
	const d3 col = shade(S, prm, m, r.o, r.d, h, shadow_rays);
	const d3 ret = w * col;
	if (prm.aterm_thresh > 0.0 && dot(col, col) < prm.aterm_thresh) return ret;
	RayTask kid[2];
	bool a, b;
	spawn_children(m, h, r.o, r.d, w, depth, kid, a, b);
	if (a) child[n_child++] = kid[0];
	if (b) child[n_child++] = kid[1];
	return ret;
}

} // namespace rtd
