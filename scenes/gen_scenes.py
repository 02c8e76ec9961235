#!/usr/bin/env python3
"""Procedural stand-in scenes for the five BASELINE.json configs, emitted as real `.ray` text.

The reference's named assets (hitchcock.ray, trimesh2.ray, dragon.ray, cubemaps) are git-ignored
upstream (reference .gitignore:1), so every config runs on a *stand-in*
with similar ray statistics.  The text obeys the grammar the reference parser accepts
(reference ray/src/parser/Parser.cpp:26-94, Tokenizer.cpp:229-237): scalars are lexed as
[0-9.\\-e]+ and handed to atof, so we never emit '+' or 'E' in a number.

Usage:  python scenes/gen_scenes.py <name> <out.ray> [--scale S]
        names: hitchcock, trimesh2, dragon, mixed, cubemap (writes 6 BMPs into a directory)
All generators are deterministic (fixed numpy PCG64 seeds).
"""
import argparse
import math
import os
import struct

import numpy as np


def fmt(x):
    """Shortest text that atof() reads back to exactly this double, without '+' or 'E'."""
    x = float(x)
    if x == 0.0:
        return "0"
    s = repr(x)
    if "e" in s or "E" in s:
        s = "%.17g" % x
        if "e" in s:
            m, e = s.split("e")
            e = int(e)
            if e > 0:  # positive exponents would need '+': expand to fixed point
                return ("%.1f" % x)
            s = "%se-%02d" % (m, -e)
    return s


def v3(v):
    return "(%s,%s,%s)" % (fmt(v[0]), fmt(v[1]), fmt(v[2]))


def material(diffuse=None, specular=None, shininess=None, reflective=None, transmissive=None,
             index=None, ambient=None, emissive=None):
    parts = []
    if ambient is not None:
        parts.append("ambient=%s;" % v3(ambient))
    if emissive is not None:
        parts.append("emissive=%s;" % v3(emissive))
    if diffuse is not None:
        parts.append("diffuse=%s;" % v3(diffuse))
    if specular is not None:
        parts.append("specular=%s;" % v3(specular))
    if shininess is not None:
        parts.append("shininess=%s;" % fmt(shininess))
    if reflective is not None:
        parts.append("reflective=%s;" % v3(reflective))
    if transmissive is not None:
        parts.append("transmissive=%s;" % v3(transmissive))
    if index is not None:
        parts.append("index=%s;" % fmt(index))
    return "material={" + " ".join(parts) + "}"


def camera(pos, viewdir, updir, fov=45.0, aspect=1.0):
    return ("camera { position=%s; viewdir=%s; updir=%s; fov=%s; aspectratio=%s; }\n"
            % (v3(pos), v3(viewdir), v3(updir), fmt(fov), fmt(aspect)))


def point_light(pos, color, a=0.0, b=0.0, c=1.0):
    return ("point_light { position=%s; color=%s; constant_attenuation_coeff=%s; "
            "linear_attenuation_coeff=%s; quadratic_attenuation_coeff=%s; }\n"
            % (v3(pos), v3(color), fmt(a), fmt(b), fmt(c)))


def dir_light(direction, color):
    return "directional_light { direction=%s; color=%s; }\n" % (v3(direction), v3(color))


def xform(prim, translate=None, rotate=None, scale=None):
    """translate(rotate(scale(prim))) - same nesting the SBT scenes use."""
    s = prim
    if scale is not None:
        if np.isscalar(scale):
            s = "scale(%s,\n%s)" % (fmt(scale), s)
        else:
            s = "scale(%s,%s,%s,\n%s)" % (fmt(scale[0]), fmt(scale[1]), fmt(scale[2]), s)
    if rotate is not None:
        ax, ang = rotate
        s = "rotate(%s,%s,%s,%s,\n%s)" % (fmt(ax[0]), fmt(ax[1]), fmt(ax[2]), fmt(ang), s)
    if translate is not None:
        s = "translate(%s,%s,%s,\n%s)" % (fmt(translate[0]), fmt(translate[1]), fmt(translate[2]), s)
    return s + "\n"


def trimesh(points, faces, mat, gennormals=False, digits=None):
    """points: (n,3) float array, faces: (m,3) int array."""
    if digits is not None:
        pts = ",".join("(%s,%s,%s)" % tuple(("%." + str(digits) + "g") % c for c in p) for p in points)
        pts = pts.replace("e+", "e").replace("e-0", "e-")
    else:
        pts = ",".join(v3(p) for p in points)
    fcs = ",".join("(%d,%d,%d)" % (f[0], f[1], f[2]) for f in faces)
    out = ["polymesh {", mat + ";", "points=(" + pts + ");", "faces=(" + fcs + ");"]
    if gennormals:
        out.append("gennormals;")
    out.append("}")
    return "\n".join(out)


HEADER = "SBT-raytracer 1.0\n\n"


# ----------------------------------------------------------------------------------------------
# C1 / C4: "hitchcock-like": one magenta sphere inside a spiral of coloured boxes
# ----------------------------------------------------------------------------------------------
def gen_hitchcock(n_boxes=96, seed=1):
    rng = np.random.Generator(np.random.PCG64(seed))
    out = [HEADER]
    out.append(camera((0.0, 3.0, 26.0), (0.0, -0.1, -1.0), (0.0, 1.0, 0.0), fov=45.0, aspect=1.0))
    out.append("ambient_light { color=(0.1,0.1,0.1); }\n")
    out.append(point_light((12.0, 14.0, 16.0), (1.0, 1.0, 1.0), a=0.2, b=0.02, c=0.002))
    out.append(point_light((-14.0, 6.0, 10.0), (0.8, 0.8, 0.9), a=0.2, b=0.02, c=0.002))
    out.append(dir_light((0.2, -1.0, -0.4), (0.5, 0.5, 0.5)))
    out.append(xform("sphere { %s }" % material(diffuse=(0.8, 0.1, 0.7), specular=(0.9, 0.9, 0.9),
                                                 shininess=64.0, reflective=(0.3, 0.3, 0.3),
                                                 ambient=(0.8, 0.1, 0.7)),
                     scale=1.5))
    for k in range(n_boxes):
        f = k / float(n_boxes)
        ang = f * 4.0 * 2.0 * math.pi
        rad = 2.0 + 8.0 * f
        pos = (rad * math.cos(ang), rad * math.sin(ang), -2.0 + 4.0 * rng.random())
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        theta = rng.uniform(0.0, math.pi)
        sc = rng.uniform(0.4, 1.2)
        hue = rng.uniform(0.0, 1.0)
        diffuse = (0.1 + 0.2 * hue, 0.3 + 0.6 * rng.random(), 0.4 + 0.6 * (1.0 - hue))
        u = rng.random()
        kw = dict(diffuse=diffuse, ambient=diffuse, specular=(0.6, 0.6, 0.6), shininess=32.0)
        if u < 0.3:
            kw["reflective"] = (0.4, 0.4, 0.4)
        elif u < 0.4:
            kw["transmissive"] = (0.7, 0.7, 0.7)
            kw["index"] = 1.3
        out.append(xform("box { %s }" % material(**kw), translate=pos, rotate=(axis, theta), scale=sc))
    return "".join(out)


# ----------------------------------------------------------------------------------------------
# C2: "trimesh2-like": architectural polymeshes (boxes as triangles), reflective slabs
# ----------------------------------------------------------------------------------------------
_BOX_V = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]], float)
_BOX_F = np.array([[0, 2, 1], [0, 3, 2], [4, 5, 6], [4, 6, 7], [0, 1, 5], [0, 5, 4],
                   [1, 2, 6], [1, 6, 5], [2, 3, 7], [2, 7, 6], [3, 0, 4], [3, 4, 7]], int)


def _boxes_mesh(boxes):
    """boxes: list of (min_xyz, max_xyz) -> (points, faces) with 8 verts / 12 tris per box."""
    pts, fcs = [], []
    for lo, hi in boxes:
        lo = np.asarray(lo, float)
        hi = np.asarray(hi, float)
        base = len(pts) * 8
        pts.append(lo + _BOX_V * (hi - lo))
        fcs.append(_BOX_F + base)
    return np.concatenate(pts), np.concatenate(fcs)


def gen_trimesh2(seed=3, detail=1.0):
    rng = np.random.Generator(np.random.PCG64(seed))
    out = [HEADER]
    out.append(camera((4.5, 2.2, 6.5), (-0.55, -0.18, -1.0), (0.0, 1.0, 0.0), fov=50.0, aspect=1.0))
    out.append("ambient_light { color=(0.15,0.15,0.15); }\n")
    out.append(point_light((3.0, 6.0, 5.0), (1.0, 0.95, 0.9), a=0.3, b=0.03, c=0.004))
    out.append(point_light((-4.0, 5.0, 2.0), (0.6, 0.7, 0.9), a=0.3, b=0.03, c=0.004))
    out.append(point_light((0.0, 3.0, -6.0), (0.7, 0.6, 0.5), a=0.3, b=0.03, c=0.004))
    groups = []
    n_steps = max(4, int(40 * detail))
    stairs = [((-6 + 0.0, 0.0 + 0.0, -3 + 0.25 * i), (-3.0, 0.12 * (i + 1), -3 + 0.25 * (i + 1))) for i in range(n_steps)]
    groups.append((stairs, dict(diffuse=(0.7, 0.65, 0.6), ambient=(0.7, 0.65, 0.6), specular=(0.2, 0.2, 0.2), shininess=16.0)))
    deck = [((-3.0, -0.1, -8.0), (7.0, 0.0, 8.0))]
    groups.append((deck, dict(diffuse=(0.35, 0.4, 0.45), ambient=(0.35, 0.4, 0.45), specular=(0.8, 0.8, 0.8),
                              shininess=96.0, reflective=(0.45, 0.45, 0.45))))
    pool = [((0.0, 0.0, -2.0), (4.0, 0.02, 3.0))]
    groups.append((pool, dict(diffuse=(0.05, 0.25, 0.4), ambient=(0.05, 0.25, 0.4), specular=(0.9, 0.9, 0.9),
                              shininess=128.0, reflective=(0.6, 0.6, 0.6))))
    n_col = max(2, int(12 * detail))
    cols = []
    for i in range(n_col):
        x = -2.5 + 9.0 * i / max(1, n_col - 1)
        for s in range(max(1, int(8 * detail))):  # columns built from stacked drums
            y0 = 0.0 + 0.4 * s
            w = 0.18 + 0.02 * ((s % 2) == 0)
            cols.append(((x - w, y0, -5.0 - w), (x + w, y0 + 0.4, -5.0 + w)))
    groups.append((cols, dict(diffuse=(0.8, 0.8, 0.75), ambient=(0.8, 0.8, 0.75), specular=(0.3, 0.3, 0.3), shininess=24.0)))
    roof = [((-3.0, 3.3, -6.0), (7.0, 3.5, -4.0))]
    groups.append((roof, dict(diffuse=(0.5, 0.3, 0.2), ambient=(0.5, 0.3, 0.2))))
    n_chair = max(2, int(10 * detail))
    for c in range(n_chair):
        cx = -1.5 + 1.1 * c
        cz = 4.0 + 0.3 * math.sin(c)
        parts = []
        for s in range(max(2, int(16 * detail))):  # slatted lounger
            parts.append(((cx, 0.3, cz + 0.1 * s), (cx + 0.7, 0.34, cz + 0.1 * s + 0.07)))
        for lx in (0.0, 0.62):
            for lz in (0.0, 0.1 * max(2, int(16 * detail)) - 0.08):
                parts.append(((cx + lx, 0.0, cz + lz), (cx + lx + 0.08, 0.3, cz + lz + 0.08)))
        col = (0.3 + 0.6 * rng.random(), 0.3 + 0.6 * rng.random(), 0.3 + 0.6 * rng.random())
        groups.append((parts, dict(diffuse=col, ambient=col, specular=(0.4, 0.4, 0.4), shininess=32.0)))
    nfaces = 0
    for boxes, mat in groups:
        p, f = _boxes_mesh(boxes)
        nfaces += len(f)
        out.append(trimesh(p, f, material(**mat)) + "\n")
    out.append("// faces: %d\n" % nfaces)
    return "".join(out)


# ----------------------------------------------------------------------------------------------
# C3: "dragon-like": one displaced UV sphere, rows x cols quads -> 2*rows*cols triangles
#     (466 x 933 -> 869,556 faces), gennormals, on a reflective floor
# ----------------------------------------------------------------------------------------------
def dragon_mesh(rows=466, cols=933, seed=4):
    rng = np.random.Generator(np.random.PCG64(seed))
    lat = np.linspace(-0.4995 * math.pi, 0.4995 * math.pi, rows + 1)
    lon = np.arange(cols) * (2.0 * math.pi / cols)
    LAT, LON = np.meshgrid(lat, lon, indexing="ij")
    r = np.ones_like(LAT)
    for o in range(6):
        fa, fb = int(rng.integers(2, 5)) * (o + 1), int(rng.integers(2, 5)) * (o + 1)
        pa, pb = rng.uniform(0, 2 * math.pi, size=2)
        r += (0.16 / (o + 1.0)) * np.sin(fa * LON + pa) * np.cos(fb * LAT + pb) * np.cos(LAT)
    x = r * np.cos(LAT) * np.cos(LON) * 1.4
    y = r * np.sin(LAT)
    z = r * np.cos(LAT) * np.sin(LON) * 0.9
    pts = np.stack([x, y, z], axis=-1).reshape(-1, 3)
    i = np.arange(rows)[:, None]
    j = np.arange(cols)[None, :]
    a = i * cols + j
    b = i * cols + (j + 1) % cols
    c = (i + 1) * cols + (j + 1) % cols
    d = (i + 1) * cols + j
    # winding chosen so the face normals cross(b-a, c-a) point away from the origin
    f1 = np.stack([a, c, b], axis=-1).reshape(-1, 3)
    f2 = np.stack([a, d, c], axis=-1).reshape(-1, 3)
    faces = np.empty((2 * rows * cols, 3), dtype=np.int64)
    faces[0::2] = f1
    faces[1::2] = f2
    return pts, faces


def gen_dragon(rows=466, cols=933, seed=4):
    pts, faces = dragon_mesh(rows, cols, seed)
    out = [HEADER]
    out.append(camera((0.0, 1.2, 5.2), (0.0, -0.2, -1.0), (0.0, 1.0, 0.0), fov=40.0, aspect=1.7777777777))
    out.append("ambient_light { color=(0.12,0.12,0.12); }\n")
    out.append(point_light((4.0, 6.0, 5.0), (1.0, 1.0, 1.0), a=0.25, b=0.02, c=0.003))
    out.append(dir_light((-0.4, -1.0, -0.3), (0.6, 0.6, 0.55)))
    floor = "square { %s }" % material(diffuse=(0.45, 0.45, 0.5), ambient=(0.45, 0.45, 0.5),
                                        specular=(0.5, 0.5, 0.5), shininess=48.0, reflective=(0.35, 0.35, 0.35))
    out.append(xform(floor, translate=(0.0, -1.45, 0.0), rotate=((1.0, 0.0, 0.0), -math.pi / 2.0), scale=14.0))
    mat = material(diffuse=(0.2, 0.65, 0.3), ambient=(0.2, 0.65, 0.3), specular=(0.8, 0.8, 0.8),
                   shininess=64.0, reflective=(0.15, 0.15, 0.15))
    out.append(trimesh(pts, faces, mat, gennormals=True, digits=10) + "\n")
    return "".join(out)


# ----------------------------------------------------------------------------------------------
# C5: synthetic mixed scene: n/4 spheres + n/4 boxes + n/4 cylinders + one n/4-face trimesh
# ----------------------------------------------------------------------------------------------
def gen_mixed(n_prims=1000000, seed=5, extent=200.0):
    rng = np.random.Generator(np.random.PCG64(seed))
    q = n_prims // 4
    out = [HEADER]
    out.append(camera((0.0, 0.0, 1.45 * extent), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), fov=45.0, aspect=1.7777777777))
    out.append("ambient_light { color=(0.15,0.15,0.15); }\n")
    e = extent
    for p in ((e, e, e), (-e, e, e), (e, -e, 0.5 * e), (-e, -e, e)):
        out.append(point_light(p, (0.8, 0.8, 0.8), a=1.0, b=0.0, c=0.0))
    mats = []
    for k in range(20):
        col = (0.2 + 0.7 * rng.random(), 0.2 + 0.7 * rng.random(), 0.2 + 0.7 * rng.random())
        kw = dict(diffuse=col, ambient=col, specular=(0.5, 0.5, 0.5), shininess=32.0)
        if k < 4:
            kw["reflective"] = (0.4, 0.4, 0.4)
        elif k < 6:
            kw["transmissive"] = (0.75, 0.75, 0.75)
            kw["index"] = 1.25
        mats.append(material(**kw))
    # object sizes grow as the count shrinks so small test variants still cover the screen
    size_lo = 0.05 * (1000000.0 / max(n_prims, 1)) ** (1.0 / 3.0)
    size_hi = 20.0 * size_lo
    chunks = []
    for kind in ("sphere", "box", "cylinder"):
        pos = rng.uniform(-0.5 * extent, 0.5 * extent, size=(q, 3))
        axis = rng.normal(size=(q, 3))
        axis /= np.linalg.norm(axis, axis=1)[:, None]
        ang = rng.uniform(0.0, math.pi, size=q)
        size = np.exp(rng.uniform(math.log(size_lo), math.log(size_hi), size=q))
        mi = rng.integers(0, len(mats), size=q)
        for k in range(q):
            chunks.append(xform("%s { %s }" % (kind, mats[mi[k]]), translate=pos[k],
                                rotate=(axis[k], ang[k]), scale=size[k]))
    out.extend(chunks)
    # the trimesh: a bumpy sheet of q faces spanning the volume diagonally
    rows = max(2, int(math.sqrt(q / 2.0)))
    cols = max(2, q // (2 * rows))
    u = np.linspace(-0.5, 0.5, rows + 1)
    v = np.linspace(-0.5, 0.5, cols + 1)
    U, V = np.meshgrid(u, v, indexing="ij")
    H = 0.06 * np.sin(9.0 * U * math.pi) * np.cos(7.0 * V * math.pi) + 0.03 * rng.normal(size=U.shape) / math.sqrt(rows)
    pts = np.stack([U * extent, H * extent - 0.3 * extent, V * extent], axis=-1).reshape(-1, 3)
    i = np.arange(rows)[:, None]
    j = np.arange(cols)[None, :]
    a = i * (cols + 1) + j
    b = a + 1
    c = a + (cols + 1) + 1
    d = a + (cols + 1)
    faces = np.empty((2 * rows * cols, 3), dtype=np.int64)
    faces[0::2] = np.stack([a, b, c], axis=-1).reshape(-1, 3)
    faces[1::2] = np.stack([a, c, d], axis=-1).reshape(-1, 3)
    out.append(trimesh(pts, faces, material(diffuse=(0.5, 0.5, 0.45), ambient=(0.5, 0.5, 0.45),
                                            specular=(0.3, 0.3, 0.3), shininess=16.0,
                                            reflective=(0.2, 0.2, 0.2)), gennormals=False, digits=10) + "\n")
    return "".join(out)


# ----------------------------------------------------------------------------------------------
# procedural cubemap: 6 x (size x size) 24-bit BMPs named so the reference's sloppy
# matcher (ui/TraceUI.cc:87-146) maps them unambiguously: px nx py ny pz nz
# ----------------------------------------------------------------------------------------------
def write_bmp(path, rgb):
    """rgb: (h, w, 3) uint8, row 0 = bottom scanline (BMP native order, as the reference keeps it)."""
    h, w, _ = rgb.shape
    pad = (4 - (w * 3) % 4) % 4
    rows = []
    for j in range(h):
        rows.append(rgb[j, :, ::-1].tobytes() + b"\0" * pad)
    data = b"".join(rows)
    hdr = struct.pack("<2sIHHI", b"BM", 54 + len(data), 0, 0, 54)
    info = struct.pack("<IiiHHIIiiII", 40, w, h, 1, 24, 0, 0, 2834, 2834, 0, 0)
    with open(path, "wb") as f:
        f.write(hdr + info + data)


def gen_cubemap(outdir, size=512, seed=2):
    rng = np.random.Generator(np.random.PCG64(seed))
    os.makedirs(outdir, exist_ok=True)
    yy, xx = np.meshgrid(np.linspace(0, 1, size), np.linspace(0, 1, size), indexing="ij")
    for name in ("px", "nx", "py", "ny", "pz", "nz"):
        base = rng.uniform(0.2, 0.9, size=3)
        checker = (((xx * 8).astype(int) + (yy * 8).astype(int)) % 2).astype(float)
        img = np.empty((size, size, 3))
        for c in range(3):
            img[:, :, c] = base[c] * (0.55 + 0.45 * yy) * (0.75 + 0.25 * checker) + 0.1 * xx * (c == 2)
        write_bmp(os.path.join(outdir, name + ".bmp"), np.clip(img * 255.0, 0, 255).astype(np.uint8))


GENERATORS = {
    "hitchcock": lambda s: gen_hitchcock(n_boxes=max(4, int(96 * s))),
    "trimesh2": lambda s: gen_trimesh2(detail=s),
    "dragon": lambda s: gen_dragon(rows=max(4, int(466 * math.sqrt(s))), cols=max(5, int(933 * math.sqrt(s)))),
    "mixed": lambda s: gen_mixed(n_prims=max(8, int(1000000 * s))),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("name", choices=sorted(GENERATORS) + ["cubemap"])
    ap.add_argument("out")
    ap.add_argument("--scale", type=float, default=1.0, help="size factor (1.0 = the BASELINE.json config)")
    ap.add_argument("--size", type=int, default=512, help="cubemap face size")
    a = ap.parse_args()
    if a.name == "cubemap":
        gen_cubemap(a.out, a.size)
        return
    text = GENERATORS[a.name](a.scale)
    with open(a.out, "w") as f:
        f.write(text)


if __name__ == "__main__":
    main()
