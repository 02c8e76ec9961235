// ray_parser.cpp - .ray ("SBT-raytracer 1.0/1.1") scene text -> rs::Scene.
//
// A from-scratch recursive-descent loader accepting the grammar the reference's Tokenizer/Parser
// accept (reference ray/src/parser/Tokenizer.cpp:70-260, Token.cpp:121-196, Parser.cpp:26-1308;
// SURVEY.md Appendix C), including its observable quirks:
//   * scalars are the longest run of [0-9.\-e] handed to atof (Tokenizer.cpp:229-237)
//   * identifiers may contain '-' (Tokenizer.cpp:211-220), "polymesh" = "trimesh", "colour" = "color"
//   * `material = {..}` inside a `{ group }` is a syntax error (Parser.cpp:202-208 falls through)
//   * faces with more than 3 indices are fan-triangulated (Parser.cpp:655-671)
//   * vec3 texture paths are relative to the scene file, scalar ones are not (Parser.cpp:1281-1302)
//   * every object owns a copy of its material; the trimesh `material` is copied into faces when
//     the mesh closes (Parser.cpp:622-647, trimesh.cpp:38-55)
// It works on an in-memory buffer with a one-token lookahead, so the 46 MB dragon stand-in loads
// in about a second instead of minutes.
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include "ray_scene.hpp"

namespace rs {
namespace {

enum Tok {
	T_EOF, T_UNKNOWN, T_IDENT, T_SCALAR, T_LPAREN, T_RPAREN, T_LBRACE, T_RBRACE, T_COMMA, T_EQUALS, T_SEMICOLON,
	K_SBT, K_AMBIENT_LIGHT, K_AMBIENT, K_ASPECTRATIO, K_BOTTOM_RADIUS, K_BOX, K_CAMERA, K_CAPPED, K_COLOR, K_CONE,
	K_CONST_ATT, K_CYLINDER, K_DIFFUSE, K_DIRECTION, K_DIRECTIONAL_LIGHT, K_EMISSIVE, K_FACES, K_FALSE, K_FOV,
	K_GENNORMALS, K_HEIGHT, K_INDEX, K_LIN_ATT, K_MATERIAL, K_MATERIALS, K_MAP, K_NAME, K_NORMALS, K_POINT_LIGHT,
	K_POINTS, K_TRIMESH, K_POSITION, K_QUAD_ATT, K_QUATERNIAN, K_REFLECTIVE, K_ROTATE, K_SCALE, K_SHININESS,
	K_SPECULAR, K_SPHERE, K_SQUARE, K_TOP_RADIUS, K_TRANSFORM, K_TRANSLATE, K_TRANSMISSIVE, K_TRUE, K_UPDIR,
	K_VIEWDIR, K_BUMP, K_GLOSS, K_ANGLE, K_WIDTH, K_RADIUS, K_AREA_LIGHT_RECT, K_AREA_LIGHT_CIRC, K_SPOT_LIGHT
};

struct KW { const char* word; Tok tok; };
const KW kKeywords[] = {
	{"ambient_light", K_AMBIENT_LIGHT}, {"ambient", K_AMBIENT}, {"aspectratio", K_ASPECTRATIO},
	{"bottom_radius", K_BOTTOM_RADIUS}, {"box", K_BOX}, {"camera", K_CAMERA}, {"capped", K_CAPPED},
	{"color", K_COLOR}, {"colour", K_COLOR}, {"cone", K_CONE}, {"constant_attenuation_coeff", K_CONST_ATT},
	{"cylinder", K_CYLINDER}, {"diffuse", K_DIFFUSE}, {"direction", K_DIRECTION},
	{"directional_light", K_DIRECTIONAL_LIGHT}, {"emissive", K_EMISSIVE}, {"faces", K_FACES}, {"false", K_FALSE},
	{"fov", K_FOV}, {"gennormals", K_GENNORMALS}, {"height", K_HEIGHT}, {"index", K_INDEX},
	{"linear_attenuation_coeff", K_LIN_ATT}, {"material", K_MATERIAL}, {"materials", K_MATERIALS}, {"map", K_MAP},
	{"name", K_NAME}, {"normals", K_NORMALS}, {"point_light", K_POINT_LIGHT}, {"points", K_POINTS},
	{"polymesh", K_TRIMESH}, {"position", K_POSITION}, {"quadratic_attenuation_coeff", K_QUAD_ATT},
	{"quaternian", K_QUATERNIAN}, {"reflective", K_REFLECTIVE}, {"rotate", K_ROTATE}, {"SBT-raytracer", K_SBT},
	{"scale", K_SCALE}, {"shininess", K_SHININESS}, {"specular", K_SPECULAR}, {"sphere", K_SPHERE},
	{"square", K_SQUARE}, {"top_radius", K_TOP_RADIUS}, {"transform", K_TRANSFORM}, {"translate", K_TRANSLATE},
	{"transmissive", K_TRANSMISSIVE}, {"trimesh", K_TRIMESH}, {"true", K_TRUE}, {"updir", K_UPDIR},
	{"viewdir", K_VIEWDIR}, {"bump", K_BUMP}, {"gloss", K_GLOSS}, {"angle", K_ANGLE}, {"width", K_WIDTH},
	{"radius", K_RADIUS}, {"area_light_rect", K_AREA_LIGHT_RECT}, {"area_light_circ", K_AREA_LIGHT_CIRC},
	{"spot_light", K_SPOT_LIGHT},
};

struct Token {
	Tok kind = T_EOF;
	double value = 0.0;
	std::string text;
};

class Lexer {
public:
	explicit Lexer(std::string&& src) : buf_(std::move(src)), p_(buf_.data()), end_(buf_.data() + buf_.size())
	{
		for (const KW& k : kKeywords) kw_[k.word] = k.tok;
	}
	const Token& peek()
	{
		if (!have_) { scan(tok_); have_ = true; }
		return tok_;
	}
	Token get()
	{
		peek();
		have_ = false;
		return tok_;
	}
	Token read(Tok want, const char* what)
	{
		Token t = get();
		if (t.kind != want) fail(std::string("expected ") + what);
		return t;
	}
	bool cond_read(Tok want)
	{
		if (peek().kind == want) { have_ = false; return true; }
		return false;
	}
	[[noreturn]] void fail(const std::string& msg) const
	{
		std::ostringstream os;
		os << "Parse error (line " << line_ << "): " << msg;
		throw std::runtime_error(os.str());
	}

private:
	static bool is_alpha(char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z'); }
	static bool is_digit(char c) { return c >= '0' && c <= '9'; }
	void skip_space()
	{
		for (;;) {
			while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r' || *p_ == '\f' || *p_ == '\v')) {
				if (*p_ == '\n') line_++;
				p_++;
			}
			if (p_ < end_ && *p_ == '/') {
				if (p_ + 1 < end_ && p_[1] == '/') {
					while (p_ < end_ && *p_ != '\n') p_++;
					continue;
				}
				if (p_ + 1 < end_ && p_[1] == '*') {
					int start = line_;
					p_ += 2;
					for (;;) {
						if (p_ >= end_) {
							std::ostringstream os;
							os << "Unterminated comment in line " << start;
							fail(os.str());
						}
						if (*p_ == '\n') line_++;
						if (*p_ == '*' && p_ + 1 < end_ && p_[1] == '/') { p_ += 2; break; }
						p_++;
					}
					continue;
				}
				fail("unexpected character: '/'");
			}
			return;
		}
	}
	void scan(Token& t)
	{
		skip_space();
		t.text.clear();
		t.value = 0.0;
		if (p_ >= end_) { t.kind = T_EOF; return; }
		char c = *p_;
		if (is_alpha(c) || c == '_') {
			const char* s = p_;
			while (p_ < end_ && (is_alpha(*p_) || is_digit(*p_) || *p_ == '_' || *p_ == '-')) p_++;
			t.text.assign(s, p_);
			auto it = kw_.find(t.text);
			t.kind = it == kw_.end() ? T_UNKNOWN : it->second;
			return;
		}
		if (c == '"') {
			const char* s = ++p_;
			while (p_ < end_ && *p_ != '"') {
				if (*p_ == '\n') fail("Unterminated string constant");
				p_++;
			}
			if (p_ >= end_) fail("Unterminated string constant");
			t.text.assign(s, p_);
			p_++;
			t.kind = T_IDENT;
			return;
		}
		if (is_digit(c) || c == '-' || c == '.') {
			const char* s = p_;
			while (p_ < end_ && (is_digit(*p_) || *p_ == '-' || *p_ == '.' || *p_ == 'e')) p_++;
			char tmp[64];
			size_t n = (size_t)(p_ - s);
			if (n < sizeof(tmp)) {
				std::memcpy(tmp, s, n);
				tmp[n] = 0;
				t.value = std::strtod(tmp, nullptr); // == atof
			} else {
				t.value = std::strtod(std::string(s, p_).c_str(), nullptr);
			}
			t.kind = T_SCALAR;
			return;
		}
		p_++;
		switch (c) {
		case '(': t.kind = T_LPAREN; return;
		case ')': t.kind = T_RPAREN; return;
		case '{': t.kind = T_LBRACE; return;
		case '}': t.kind = T_RBRACE; return;
		case ',': t.kind = T_COMMA; return;
		case '=': t.kind = T_EQUALS; return;
		case ';': t.kind = T_SEMICOLON; return;
		default: fail(std::string("unexpected character: '") + c + "'");
		}
	}
	std::string buf_;
	const char* p_;
	const char* end_;
	int line_ = 1;
	bool have_ = false;
	Token tok_;
	std::map<std::string, Tok> kw_;
};

class Parser {
public:
	Parser(Lexer& lx, Scene& sc, const std::string& base) : lx_(lx), sc_(sc), base_(base) {}

	void parse_scene()
	{
		lx_.read(K_SBT, "'SBT-raytracer'");
		Token ver = lx_.read(T_SCALAR, "version number");
		if (ver.value > 1.1) lx_.fail("SBT-raytracer version number too high; only able to parse v1.1 and below.");
		const Transform root = make_transform(nullptr, hm::M4());
		Material mat;
		for (;;) {
			switch (lx_.peek().kind) {
			case K_SPHERE: case K_BOX: case K_SQUARE: case K_CYLINDER: case K_CONE: case K_TRIMESH:
			case K_TRANSLATE: case K_ROTATE: case K_SCALE: case K_TRANSFORM: case T_LBRACE:
				transformable(root, mat);
				break;
			case K_POINT_LIGHT: case K_DIRECTIONAL_LIGHT: case K_AREA_LIGHT_RECT: case K_AREA_LIGHT_CIRC: case K_SPOT_LIGHT:
				light(lx_.peek().kind);
				break;
			case K_AMBIENT_LIGHT: {
				lx_.get();
				lx_.read(T_LBRACE, "'{'");
				if (lx_.peek().kind != K_COLOR) lx_.fail("Expected color attribute");
				sc_.ambient = sc_.ambient + vec3_expr();
				lx_.read(T_RBRACE, "'}'");
				break;
			}
			case K_CAMERA: camera(); break;
			case K_MATERIAL: mat = material_expr(mat); break;
			case T_SEMICOLON: lx_.get(); break;
			case T_EOF: return;
			default: lx_.fail("Expected: geometry, camera, or light information");
			}
		}
	}

private:
	// ---- primitives of the grammar -----------------------------------------------------------
	double scalar() { return lx_.read(T_SCALAR, "scalar").value; }
	V3 vec3()
	{
		lx_.read(T_LPAREN, "'('");
		double a = scalar(); lx_.read(T_COMMA, "','");
		double b = scalar(); lx_.read(T_COMMA, "','");
		double c = scalar();
		lx_.read(T_RPAREN, "')'");
		return V3(a, b, c);
	}
	hm::V4 vec4()
	{
		lx_.read(T_LPAREN, "'('");
		double a = scalar(); lx_.read(T_COMMA, "','");
		double b = scalar(); lx_.read(T_COMMA, "','");
		double c = scalar(); lx_.read(T_COMMA, "','");
		double d = scalar();
		lx_.read(T_RPAREN, "')'");
		return hm::V4(a, b, c, d);
	}
	// `<key> = value [;]` forms (Parser.cpp:1062-1109): the key token is dropped unchecked
	double scalar_expr() { lx_.get(); lx_.read(T_EQUALS, "'='"); double v = scalar(); lx_.cond_read(T_SEMICOLON); return v; }
	V3 vec3_expr() { lx_.get(); lx_.read(T_EQUALS, "'='"); V3 v = vec3(); lx_.cond_read(T_SEMICOLON); return v; }
	hm::V4 vec4_expr() { lx_.get(); lx_.read(T_EQUALS, "'='"); hm::V4 v = vec4(); lx_.cond_read(T_SEMICOLON); return v; }
	bool bool_expr()
	{
		lx_.get();
		lx_.read(T_EQUALS, "'='");
		Token t = lx_.get();
		if (t.kind != K_TRUE && t.kind != K_FALSE) lx_.fail("Expected boolean");
		lx_.cond_read(T_SEMICOLON);
		return t.kind == K_TRUE;
	}
	void ident_expr() { lx_.get(); lx_.read(T_EQUALS, "'='"); lx_.read(T_IDENT, "identifier"); lx_.cond_read(T_SEMICOLON); }

	// ---- transforms (TransformNode ctor, scene.h:119-129) -------------------------------------
	static Transform make_transform(const Transform* parent, const hm::M4& local)
	{
		Transform t;
		t.xform = parent ? hm::mul(parent->xform, local) : local;
		t.inverse = hm::inverse(t.xform);
		t.normi = hm::transpose(hm::inverse(hm::upper3(t.xform)));
		return t;
	}
	// Transform nodes live on the parser's call stack; every object keeps a copy of its own.
	static Transform child_transform(const Transform& parent, const hm::M4& local) { return make_transform(&parent, local); }

	void transformable(const Transform& xf, const Material& mat)
	{
		switch (lx_.peek().kind) {
		case T_LBRACE: group(xf, mat); break;
		case K_SPHERE: simple(SPHERE, xf, mat); break;
		case K_BOX: simple(BOX, xf, mat); break;
		case K_SQUARE: simple(SQUARE, xf, mat); break;
		case K_CYLINDER: simple(CYLINDER, xf, mat); break;
		case K_CONE: cone(xf, mat); break;
		case K_TRIMESH: trimesh(xf, mat); break;
		case K_TRANSLATE: {
			lx_.get(); lx_.read(T_LPAREN, "'('");
			double x = scalar(); lx_.read(T_COMMA, "','");
			double y = scalar(); lx_.read(T_COMMA, "','");
			double z = scalar(); lx_.read(T_COMMA, "','");
			transformable(child_transform(xf, hm::translation(V3(x, y, z))), mat);
			lx_.read(T_RPAREN, "')'"); lx_.cond_read(T_SEMICOLON);
			break;
		}
		case K_ROTATE: {
			lx_.get(); lx_.read(T_LPAREN, "'('");
			double x = scalar(); lx_.read(T_COMMA, "','");
			double y = scalar(); lx_.read(T_COMMA, "','");
			double z = scalar(); lx_.read(T_COMMA, "','");
			double w = scalar(); lx_.read(T_COMMA, "','");
			transformable(child_transform(xf, hm::rotation(w, V3(x, y, z))), mat);
			lx_.read(T_RPAREN, "')'"); lx_.cond_read(T_SEMICOLON);
			break;
		}
		case K_SCALE: {
			lx_.get(); lx_.read(T_LPAREN, "'('");
			double x = scalar(), y, z;
			lx_.read(T_COMMA, "','");
			if (lx_.peek().kind == T_SCALAR) {
				y = scalar(); lx_.read(T_COMMA, "','");
				z = scalar(); lx_.read(T_COMMA, "','");
			} else {
				y = x; z = x;
			}
			transformable(child_transform(xf, hm::scaling(V3(x, y, z))), mat);
			lx_.read(T_RPAREN, "')'"); lx_.cond_read(T_SEMICOLON);
			break;
		}
		case K_TRANSFORM: {
			lx_.get(); lx_.read(T_LPAREN, "'('");
			hm::V4 r1 = vec4(); lx_.read(T_COMMA, "','");
			hm::V4 r2 = vec4(); lx_.read(T_COMMA, "','");
			hm::V4 r3 = vec4(); lx_.read(T_COMMA, "','");
			hm::V4 r4 = vec4(); lx_.read(T_COMMA, "','");
			transformable(child_transform(xf, hm::transpose(hm::M4(r1, r2, r3, r4))), mat);
			lx_.read(T_RPAREN, "')'"); lx_.cond_read(T_SEMICOLON);
			break;
		}
		default: lx_.fail("Expected: transformable element");
		}
	}

	void group(const Transform& xf, const Material& mat)
	{
		lx_.read(T_LBRACE, "'{'");
		for (;;) {
			switch (lx_.peek().kind) {
			case K_SPHERE: case K_BOX: case K_SQUARE: case K_CYLINDER: case K_CONE: case K_TRIMESH:
			case K_TRANSLATE: case K_ROTATE: case K_SCALE: case K_TRANSFORM: case T_LBRACE:
				transformable(xf, mat);
				break;
			case T_RBRACE: lx_.get(); return;
			default: lx_.fail("Expected: '}' or geometry"); // includes `material` (Parser.cpp:202-208)
			}
		}
	}

	// ---- objects ---------------------------------------------------------------------------
	// Geometry::ComputeBoundingBox (scene.cpp:78-116): transform the 8 corners in the reference's
	// order and fold with glm::max/min on dvec4s.
	static void world_bounds(const Transform& xf, V3 lo, V3 hi, V3& out_lo, V3& out_hi)
	{
		hm::V4 mx, mn;
		for (int k = 0; k < 8; k++) {
			hm::V4 c((k & 1) ? hi.x : lo.x, (k & 2) ? hi.y : lo.y, (k & 4) ? hi.z : lo.z, 1.0);
			hm::V4 v = hm::mul(xf.xform, c);
			if (k == 0) { mx = v; mn = v; }
			else { mx = hm::gmax(mx, v); mn = hm::gmin(mn, v); }
		}
		out_lo = V3(mn.x, mn.y, mn.z);
		out_hi = V3(mx.x, mx.y, mx.z);
	}
	void add_object(Object& o, const Transform& xf, V3 lo, V3 hi)
	{
		o.inverse = xf.inverse;
		o.normi = xf.normi;
		world_bounds(xf, lo, hi, o.bmin, o.bmax);
		sc_.objects.push_back(std::move(o));
	}

	void simple(ObjType type, const Transform& xf, const Material& mat)
	{
		lx_.get();
		lx_.read(T_LBRACE, "'{'");
		Object o;
		o.type = type;
		o.material = mat;
		for (;;) {
			switch (lx_.peek().kind) {
			case K_MATERIAL: o.material = material_expr(mat); break;
			case K_NAME: ident_expr(); break;
			case T_RBRACE: {
				lx_.get();
				// local bounds: Sphere.h, Box.h, Square.h:19-25, Cylinder.h:19-25
				V3 lo, hi;
				if (type == SPHERE) { lo = V3(-1, -1, -1); hi = V3(1, 1, 1); }
				else if (type == BOX) { lo = V3(-0.5, -0.5, -0.5); hi = V3(0.5, 0.5, 0.5); }
				else if (type == SQUARE) { lo = V3(-0.5, -0.5, -1e-8); hi = V3(0.5, 0.5, 1e-8); }
				else { lo = V3(-1, -1, 0); hi = V3(1, 1, 1); }
				add_object(o, xf, lo, hi);
				return;
			}
			default: lx_.fail("Expected: object attributes");
			}
		}
	}

	void cone(const Transform& xf, const Material& mat)
	{
		lx_.get();
		lx_.read(T_LBRACE, "'{'");
		Object o;
		o.type = CONE;
		o.material = mat;
		double br = 1.0, tr = 0.0, h = 1.0;
		bool capped = true;
		for (;;) {
			switch (lx_.peek().kind) {
			case K_MATERIAL: o.material = material_expr(mat); break;
			case K_NAME: ident_expr(); break;
			case K_CAPPED: capped = bool_expr(); break;
			case K_BOTTOM_RADIUS: br = scalar_expr(); break;
			case K_TOP_RADIUS: tr = scalar_expr(); break;
			case K_HEIGHT: h = scalar_expr(); break;
			case T_RBRACE: {
				lx_.get();
				// Cone::Cone (Cone.h:10-37)
				o.capped = capped;
				o.height = h;
				o.b_radius = (br < 0.0) ? -br : br;
				o.t_radius = (tr < 0.0) ? -tr : tr;
				if (o.b_radius < 0.0001) o.b_radius = 0.0001;
				if (o.t_radius < 0.0001) o.t_radius = 0.0001;
				double beta = (o.t_radius - o.b_radius) / o.height;
				if (std::fabs(beta) < 0.001) beta = 0.001;
				double gamma = (beta < 0.0) ? o.t_radius / beta : o.b_radius / beta;
				o.beta_squared = beta * beta;
				if (gamma < 0.0) gamma = gamma - o.height;
				o.gamma = gamma;
				double big = (o.b_radius > o.t_radius) ? o.b_radius : o.t_radius;
				V3 lo(-big, -big, (h < 0.0) ? h : 0.0), hi(big, big, (h < 0.0) ? 0.0 : h);
				add_object(o, xf, lo, hi);
				return;
			}
			default: lx_.fail("Expected: cone attributes");
			}
		}
	}

	void trimesh(const Transform& xf, const Material& mat)
	{
		lx_.get();
		lx_.read(T_LBRACE, "'{'");
		Object o;
		o.type = TRIMESH;
		o.material = mat;
		Mesh m;
		bool gen_normals = false;
		std::vector<int32_t> raw_faces; // triples, as listed (fan-triangulated)
		for (;;) {
			switch (lx_.peek().kind) {
			case K_GENNORMALS: lx_.get(); lx_.read(T_SEMICOLON, "';'"); gen_normals = true; break;
			case K_MATERIAL: o.material = material_expr(mat); break;
			case K_NAME: ident_expr(); break;
			case K_MATERIALS:
				list_open();
				while (list_next()) m.vmaterials.push_back(material(o.material));
				list_close();
				break;
			case K_NORMALS:
				list_open();
				while (list_next()) m.normals.push_back(vec3());
				list_close();
				break;
			case K_POINTS:
				list_open();
				while (list_next()) m.vertices.push_back(vec3());
				list_close();
				break;
			case K_FACES:
				list_open();
				while (list_next()) {
					// one face: a scalar list of >= 3 indices, fan-triangulated (Parser.cpp:655-671)
					lx_.read(T_LPAREN, "'('");
					std::vector<double> idx;
					if (lx_.peek().kind != T_RPAREN) {
						idx.push_back(scalar());
						while (lx_.peek().kind != T_RPAREN) { lx_.read(T_COMMA, "','"); idx.push_back(scalar()); }
					}
					lx_.read(T_RPAREN, "')'");
					if (idx.size() < 3) lx_.fail("Faces must have at least 3 vertices.");
					for (size_t k = 2; k < idx.size(); k++) {
						raw_faces.push_back((int32_t)idx[0]);
						raw_faces.push_back((int32_t)idx[k - 1]);
						raw_faces.push_back((int32_t)idx[k]);
					}
				}
				list_close();
				break;
			case T_RBRACE: {
				lx_.get();
				finish_mesh(o, m, raw_faces, gen_normals, xf);
				return;
			}
			default: lx_.fail("Expected: trimesh attributes");
			}
		}
	}
	// `key = ( a, b, ... );` list helpers (Parser.cpp:548-619)
	bool first_ = false;
	void list_open() { lx_.get(); lx_.read(T_EQUALS, "'='"); lx_.read(T_LPAREN, "'('"); first_ = true; }
	bool list_next()
	{
		if (lx_.peek().kind == T_RPAREN) return false;
		if (!first_) lx_.read(T_COMMA, "','");
		first_ = false;
		return true;
	}
	void list_close() { lx_.read(T_RPAREN, "')'"); lx_.read(T_SEMICOLON, "';'"); }

	void finish_mesh(Object& o, Mesh& m, const std::vector<int32_t>& raw, bool gen_normals, const Transform& xf)
	{
		const int vcnt = (int)m.vertices.size();
		for (size_t f = 0; f + 2 < raw.size(); f += 3) {
			int a = raw[f], b = raw[f + 1], c = raw[f + 2];
			if (a >= vcnt || b >= vcnt || c >= vcnt) lx_.fail("Bad face in trimesh"); // trimesh.cpp:38-42
			// TrimeshFace ctor (trimesh.h:100-130): degenerate faces are dropped
			V3 A = m.vertices[a], B = m.vertices[b], C = m.vertices[c];
			V3 vab = B - A, vac = C - A, vcb = B - C;
			if (hm::length(vab) == 0.0 || hm::length(vac) == 0.0 || hm::length(vcb) == 0.0) continue;
			V3 n = hm::normalize(hm::cross(B - A, C - A));
			m.face_ids.push_back(a); m.face_ids.push_back(b); m.face_ids.push_back(c);
			m.face_normals.push_back(n);
		}
		if (gen_normals) { // Trimesh::generateNormals (trimesh.cpp:192-217)
			m.normals.resize(vcnt);
			std::vector<int> cnt(vcnt, 0);
			for (size_t f = 0; f < m.face_normals.size(); f++)
				for (int k = 0; k < 3; k++) {
					int v = m.face_ids[3 * f + k];
					m.normals[v] = m.normals[v] + m.face_normals[f];
					cnt[v]++;
				}
			for (int v = 0; v < vcnt; v++)
				if (cnt[v]) m.normals[v] = m.normals[v] / (double)cnt[v];
		}
		if (!m.vmaterials.empty() && (int)m.vmaterials.size() != vcnt) lx_.fail("Bad Trimesh: Wrong number of materials.");
		if (!m.normals.empty() && (int)m.normals.size() != vcnt) lx_.fail("Bad Trimesh: Wrong number of normals.");
		// Trimesh::ComputeLocalBoundingBox (trimesh.h:63-80)
		V3 lo, hi;
		if (vcnt > 0) {
			lo = hi = m.vertices[0];
			for (const V3& v : m.vertices) { hi = hm::gmax(hi, v); lo = hm::gmin(lo, v); }
		}
		o.mesh = (int)sc_.meshes.size();
		sc_.meshes.push_back(std::move(m));
		add_object(o, xf, lo, hi);
	}

	// ---- materials (Parser.cpp:1095-1101, 1188-1308) ------------------------------------------
	Material material_expr(const Material& parent)
	{
		lx_.read(K_MATERIAL, "'material'");
		lx_.read(T_EQUALS, "'='");
		Material m = material(parent);
		lx_.cond_read(T_SEMICOLON);
		return m;
	}
	MatParam vec3_param()
	{
		lx_.get();
		lx_.read(T_EQUALS, "'='");
		MatParam p;
		if (lx_.cond_read(K_MAP)) {
			lx_.read(T_LPAREN, "'('");
			std::string file = base_ + "/" + lx_.read(T_IDENT, "file name").text;
			lx_.read(T_RPAREN, "')'");
			lx_.cond_read(T_SEMICOLON);
			p.tex = load_texture(sc_, file); // value stays (0,0,0): MaterialParameter(TextureMap*)
		} else {
			p.value = vec3();
			lx_.cond_read(T_SEMICOLON);
		}
		return p;
	}
	MatParam scalar_param()
	{
		lx_.get();
		lx_.read(T_EQUALS, "'='");
		MatParam p;
		if (lx_.cond_read(K_MAP)) {
			lx_.read(T_LPAREN, "'('");
			std::string file = lx_.read(T_IDENT, "file name").text; // not relative to the scene (Parser.cpp:1299)
			lx_.read(T_RPAREN, "')'");
			lx_.cond_read(T_SEMICOLON);
			p.tex = load_texture(sc_, file);
		} else {
			p = MatParam(scalar());
			lx_.cond_read(T_SEMICOLON);
		}
		return p;
	}
	Material material(const Material& parent)
	{
		if (lx_.peek().kind == T_IDENT)
			lx_.fail("material given by name: the reference peeks the name without consuming it (Parser.cpp:1189-1192) and then fails; unsupported");
		lx_.read(T_LBRACE, "'{'");
		Material m = parent;
		std::string name;
		for (;;) {
			switch (lx_.peek().kind) {
			case K_EMISSIVE: m.ke = vec3_param(); break;
			case K_AMBIENT: m.ka = vec3_param(); break;
			case K_SPECULAR: m.ks = vec3_param(); break; // setSpecular(MaterialParameter): no setBools
			case K_DIFFUSE: m.kd = vec3_param(); break;
			case K_REFLECTIVE: m.kr = vec3_param(); m.set_bools(); break;
			case K_TRANSMISSIVE: m.kt = vec3_param(); m.set_bools(); break;
			case K_INDEX: m.index = scalar_param(); break;
			case K_SHININESS: m.shininess = scalar_param(); break;
			case K_GLOSS: m.gloss = scalar_param(); break;
			case K_BUMP: m.bump = vec3_param(); break;
			case K_NAME:
				lx_.get();
				name = lx_.read(T_IDENT, "identifier").text;
				lx_.read(T_SEMICOLON, "';'");
				break;
			case T_RBRACE:
				lx_.get();
				if (!name.empty()) {
					if (named_.count(name)) lx_.fail("Redefinition of material '" + name + "'.");
					named_[name] = m;
				}
				return m;
			default: lx_.fail("Expected: material attribute");
			}
		}
	}

	// ---- lights (Parser.cpp:688-1058; ctors light.h:39-155) -----------------------------------
	void light(Tok kind)
	{
		lx_.get();
		lx_.read(T_LBRACE, "'{'");
		Light L;
		V3 direction, updir;
		double angle = 0.0;
		bool has_pos = false, has_col = false, has_dir = false, has_up = false, has_w = false, has_h = false, has_r = false, has_ang = false;
		const bool is_dir = kind == K_DIRECTIONAL_LIGHT;
		for (;;) {
			Tok k = lx_.peek().kind;
			if (k == T_RBRACE) break;
			if (k == K_COLOR) { if (has_col) lx_.fail("Repeated 'color' attribute"); L.color = vec3_expr(); has_col = true; }
			else if (k == K_POSITION && !is_dir) { if (has_pos) lx_.fail("Repeated 'position' attribute"); L.position = vec3_expr(); has_pos = true; }
			else if (k == K_DIRECTION && kind != K_POINT_LIGHT) { if (has_dir) lx_.fail("Repeated 'direction' attribute"); direction = vec3_expr(); has_dir = true; }
			else if (k == K_CONST_ATT && !is_dir) L.a = (float)scalar_expr();
			else if (k == K_LIN_ATT && !is_dir) L.b = (float)scalar_expr();
			else if (k == K_QUAD_ATT && !is_dir) L.c = (float)scalar_expr();
			else if (k == K_UPDIR && kind == K_AREA_LIGHT_RECT) { updir = vec3_expr(); has_up = true; }
			else if (k == K_WIDTH && kind == K_AREA_LIGHT_RECT) { L.w = scalar_expr(); has_w = true; }
			else if (k == K_HEIGHT && kind == K_AREA_LIGHT_RECT) { L.h = scalar_expr(); has_h = true; }
			else if (k == K_RADIUS && (kind == K_AREA_LIGHT_CIRC || kind == K_SPOT_LIGHT)) { L.radius = scalar_expr(); has_r = true; }
			else if (k == K_ANGLE && kind == K_SPOT_LIGHT) { angle = scalar_expr(); has_ang = true; }
			else lx_.fail("unexpected light attribute");
		}
		lx_.get();
		if (!has_col) lx_.fail("Expected: 'color'");
		if (is_dir) {
			if (!has_dir) lx_.fail("Expected: 'direction'");
			L.type = L_DIRECTIONAL;
			L.orientation = hm::normalize(direction);
			L.a = L.b = 0.0f; L.c = 0.0f;
		} else {
			if (!has_pos) lx_.fail("Expected: 'position'");
			if (kind == K_POINT_LIGHT) L.type = L_POINT;
			else {
				if (!has_dir) lx_.fail("Expected: 'direction'");
				L.orientation = hm::normalize(direction);
				if (kind == K_AREA_LIGHT_RECT) {
					if (!has_up || !has_w || !has_h) lx_.fail("Expected: 'updir', 'width' and 'height'");
					L.type = L_AREA_RECT;
					L.u = hm::normalize(updir);
					L.v = hm::cross(direction, updir); // both names bind to the un-normalised ctor arguments (light.h:120-127)
				} else {
					if (!has_r) lx_.fail("Expected: 'radius'");
					L.type = kind == K_SPOT_LIGHT ? L_SPOT : L_AREA_CIRC;
					if (kind == K_SPOT_LIGHT) {
						if (!has_ang) lx_.fail("Expected: 'angle'");
						L.ang_tan = std::tan(angle / 360 * 3.1415926535897932384626433832795028841971);
						L.offset = L.ang_tan * L.radius;
					}
				}
			}
		}
		sc_.lights.push_back(L);
	}

	// ---- camera (Parser.cpp:97-154, camera.cpp:32-99) -----------------------------------------
	void camera()
	{
		lx_.get();
		lx_.read(T_LBRACE, "'{'");
		Camera& c = sc_.camera;
		bool has_view = false, has_up = false;
		V3 view, up;
		for (;;) {
			switch (lx_.peek().kind) {
			case K_POSITION: c.eye = vec3_expr(); break;
			case K_FOV: {
				double fov = scalar_expr();
				fov /= (180.0 / 3.14159265359); // camera.cpp:4,80
				c.normalized_height = 2 * std::tan(fov / 2);
				c.update();
				break;
			}
			case K_QUATERNIAN: {
				hm::V4 q = vec4_expr();
				const double r = q.x, i = q.y, j = q.z, k = q.w;
				hm::M3 m;
				m[0][0] = 1.0 - 2.0 * (i * i + j * j); m[0][1] = 2.0 * (r * i - j * k); m[0][2] = 2.0 * (j * r + i * k);
				m[1][0] = 2.0 * (r * i + j * k); m[1][1] = 1.0 - 2.0 * (j * j + r * r); m[1][2] = 2.0 * (i * j - r * k);
				m[2][0] = 2.0 * (j * r - i * k); m[2][1] = 2.0 * (i * j + r * k); m[2][2] = 1.0 - 2.0 * (i * i + r * r);
				c.m = hm::transpose(m);
				c.update();
				break;
			}
			case K_ASPECTRATIO: c.aspect = scalar_expr(); c.update(); break;
			case K_VIEWDIR: view = vec3_expr(); has_view = true; break;
			case K_UPDIR: up = vec3_expr(); has_up = true; break;
			case T_RBRACE:
				if (has_view != has_up) lx_.fail(has_view ? "Expected: 'updir'" : "Expected: 'viewdir'");
				if (has_view) {
					V3 z = -view;
					V3 x = hm::cross(up, z);
					c.m = hm::M3(x, up, z);
					c.update();
				}
				lx_.get();
				return;
			default: lx_.fail("Expected: camera attribute");
			}
		}
	}

	Lexer& lx_;
	Scene& sc_;
	std::string base_;
	std::map<std::string, Material> named_;
};

} // namespace

int load_texture(Scene& sc, const std::string& path)
{
	auto it = sc.texture_cache.find(path);
	if (it != sc.texture_cache.end()) return it->second;
	Texture t;
	if (!read_image(path, t)) throw std::runtime_error("Texture mapping exception: Unable to load texture map '" + path + "'.");
	sc.textures.push_back(std::move(t));
	int id = (int)sc.textures.size() - 1;
	sc.texture_cache[path] = id;
	return id;
}

void parse_ray_file(const std::string& path, Scene& out)
{
	std::ifstream f(path, std::ios::binary);
	if (!f) throw std::runtime_error("Error: couldn't read scene file " + path);
	std::string src((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
	std::string base = ".";
	size_t cut = path.find_last_of("\\/");
	if (cut != std::string::npos) base = path.substr(0, cut);
	Lexer lx(std::move(src));
	Parser p(lx, out, base);
	p.parse_scene();
}

} // namespace rs
