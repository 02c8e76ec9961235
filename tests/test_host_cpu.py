"""CPU-only tests (no GPU needed):
  * the scene arena the product's loader emits is byte-identical to the one flattened out of the
    reference's own objects (same trees, same leaf order, same materials);
  * the per-ray DEVICE code (csrc/device/*.cuh), compiled for the host by tests/hostsim, agrees
    bit for bit with the reference on hits and frames - this pins the arithmetic before any GPU
    is involved;
  * librt_b200.so loads and exports every symbol include/rt_b200.h declares;
  * loader error behaviour.
The oracle-backed tests skip when oracle/_ref is absent and /root/reference is not there to
build it from.
"""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
from conftest import REF_SCENES

# generated stand-ins, hand-written scenes (tests/scenes), every texture-mapped parameter (conftest.tex_params_scene)
# and the reference's own seven fixtures ray/newScene/*.ray with their BMP texture and bump maps ("ref:")
SMALL = ["hitchcock.ray", "trimesh2.ray", "prims_mixed.ray", "dragon_small.ray", "mixed_small.ray", "lights_area.ray", "tex_params.ray"] + REF_SCENES


# ---------------------------------------------------------------------------------------------
def test_abi_exports_every_declared_symbol(rtlib):
    """Every function declared in include/rt_b200.h is exported by the shared library."""
    hdr = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = re.findall(r"\b(rt_[a-z_0-9]+)\s*\(", hdr)
    assert len(names) >= 14
    L = ctypes.CDLL(rtlib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), "librt_b200.so does not export " + n


def test_params_default_matches_reference_ui_defaults(rtlib):
    p = rtlib.Params.default()
    # TraceUI.h:37-129 / TraceUI.cc ctor: depth 0, 3 AA samples, 5 DOF divisions, soft shadows 5
    assert (p.depth, p.aa_mode, p.aa_samples, p.dof_on, p.dof_divs) == (0, 0, 3, 0, 5)
    assert (p.anaglyph, p.overlapping_objects, p.use_cubemap, p.soft_shadow_res) == (0, 0, 0, 5)
    assert (p.aterm_thresh, p.aa_thresh, p.dof_fd, p.dof_apsz) == (0.0, 1.0, 3.0, 0.05)


def test_no_device_fails_loudly(rtlib):
    """There is no CPU path behind the C-ABI: without a CUDA device rt_create must fail."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rtlib.RtError) as e:
        rtlib.Context(0)
    assert e.value.status in (2, 3)


def test_loader_errors(rtlib, tmp_path):
    with pytest.raises(rtlib.RtError) as e:
        rtlib.load_ray_file(str(tmp_path / "missing.ray"))
    assert e.value.status == 5  # RT_ERR_IO
    bad = tmp_path / "bad.ray"
    bad.write_text("SBT-raytracer 1.0\nsphere { material = { diffuse = (1,0 } }\n")
    with pytest.raises(rtlib.RtError) as e:
        rtlib.load_ray_file(str(bad))
    assert e.value.status == 6  # RT_ERR_PARSE
    notray = tmp_path / "x.ray"
    notray.write_text("hello\n")
    with pytest.raises(rtlib.RtError):
        rtlib.load_ray_file(str(notray))


def test_arena_validate_rejects_damage(rtlib, scene_dir):
    a = rtlib.load_ray_file(str(scene_dir / "trimesh2.ray"))
    L = rtlib.lib()
    buf = bytearray(a.bytes())
    assert L.rt_arena_validate(bytes(buf), len(buf)) == 0
    assert L.rt_arena_validate(bytes(buf), len(buf) // 2) != 0  # truncated
    broken = bytearray(buf)
    broken[0] ^= 0xFF  # magic
    assert L.rt_arena_validate(bytes(broken), len(broken)) != 0
    assert L.rt_arena_validate(None, 0) != 0


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("scene", SMALL)
def test_arena_matches_reference_flatten(rtlib, oracle_mod, scene_path, scene):
    """rt_load_ray_file == what the reference's parser + Scene::conclude + Trimesh::conclude +
    KdTree build leave in memory, flattened by the harness: byte for byte."""
    path = scene_path(scene)
    mine = rtlib.load_ray_file(str(path)).bytes()
    orc = oracle_mod.Oracle(str(path))
    theirs = orc.flatten()
    assert len(mine) == len(theirs)
    assert mine == theirs


def test_arena_with_cubemap_matches(rtlib, oracle_mod, scene_dir, cubemap_faces):
    path = scene_dir / "hitchcock.ray"
    mine = rtlib.load_ray_file(str(path), cubemap_faces).bytes()
    theirs = oracle_mod.Oracle(str(path), cubemap_faces).flatten()
    assert mine == theirs


# ---------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def hostsim():
    d = os.path.join(ROOT, "tests", "hostsim")
    subprocess.check_call(["make", "-s", "-C", d])
    S = ctypes.CDLL(os.path.join(d, "libhostsim.so"))
    S.sim_trace_rays.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    S.sim_render.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    return S


@pytest.mark.parametrize("scene", SMALL)
@pytest.mark.parametrize("prune", [0, 1])
def test_device_code_on_host_hits_bit_exact(rtlib, oracle_mod, hostsim, scene_path, scene, prune):
    """prune=0 is the reference's exhaustive walk, prune=1 the front-to-back walk with culling:
    both must name the reference's winner, with the same t and normal, bit for bit."""
    from cs378hgraphics_raytracer_b200.binding import HIT_DTYPE
    path = scene_path(scene)
    arena = rtlib.load_ray_file(str(path))
    orc = oracle_mod.Oracle(str(path))
    w = 96
    h = oracle_mod.frame_height(w, orc.aspect)
    rays = oracle_mod.camera_rays(arena.camera(), w, h)
    p = rtlib.Params.default(depth=3)
    ref_hits, ref_rgb = orc.trace_rays(p, rays, 3)
    n = len(rays)
    hits = np.zeros(n, HIT_DTYPE)
    rgb = np.zeros((n, 3))
    assert hostsim.sim_trace_rays(arena.ptr, ctypes.byref(p), rays.ctypes.data, n, 3, hits.ctypes.data, rgb.ctypes.data, prune) == 0
    assert (ref_hits["obj"] >= 0).sum() > 20
    for f in ("obj", "face", "t", "N"):
        assert np.array_equal(ref_hits[f], hits[f]), f
    assert np.abs(ref_rgb - rgb).max() < 1e-9


@pytest.mark.parametrize("scene", ["hitchcock.ray", "prims_mixed.ray", "dragon_small.ray", "lights_area.ray", "tex_params.ray"] + REF_SCENES)
def test_device_code_on_host_frame_and_ray_counts(rtlib, oracle_mod, hostsim, scene_path, scene):
    path = scene_path(scene)
    arena = rtlib.load_ray_file(str(path))
    orc = oracle_mod.Oracle(str(path))
    w = 80
    h = oracle_mod.frame_height(w, orc.aspect)
    p = rtlib.Params.default(depth=4)
    ref_img, ref_st, _ = orc.render(p, w, h)
    img = np.zeros((h, w, 3), np.uint8)
    st = rtlib.Stats()
    assert hostsim.sim_render(arena.ptr, ctypes.byref(p), w, h, img.ctypes.data, ctypes.byref(st)) == 0
    d = oracle_mod.image_diff(ref_img, img)
    assert d["within1"] >= 0.999, d
    assert (st.camera_rays, st.reflection_rays, st.refraction_rays, st.shadow_rays) == \
        (ref_st.camera_rays, ref_st.reflection_rays, ref_st.refraction_rays, ref_st.shadow_rays)


@pytest.mark.parametrize("scene", ["dragon_small.ray", "mixed_small.ray", "prims_mixed.ray", "lights_area.ray"])
def test_zero_direction_component_rays(rtlib, oracle_mod, hostsim, scene_dir, scene):
    """Rays with an exactly zero direction component (centre pixel column / row): the reference's slab
    test skips that axis (bbox.cc:43); the pruned walk refuses boxes the origin lies outside of along
    it (rt_prims.cuh) - the winner, t and the colour of the whole ray tree must not move."""
    from cs378hgraphics_raytracer_b200.binding import HIT_DTYPE
    path = scene_dir / scene
    arena = rtlib.load_ray_file(str(path))
    orc = oracle_mod.Oracle(str(path))
    w = 640
    h = oracle_mod.frame_height(w, orc.aspect)
    rays = oracle_mod.camera_rays(arena.camera(), w, h).reshape(h, w, 6)
    rays = np.ascontiguousarray(np.concatenate([rays[:, w // 2], rays[h // 2, :]]))
    assert (rays[:, 3:6] == 0).any(axis=1).sum() >= h
    n = len(rays)
    p = rtlib.Params.default(depth=4)
    ref_hits, ref_rgb = orc.trace_rays(p, rays, 4)
    hits = np.zeros(n, HIT_DTYPE)
    rgb = np.zeros((n, 3))
    assert hostsim.sim_trace_rays(arena.ptr, ctypes.byref(p), rays.ctypes.data, n, 4, hits.ctypes.data, rgb.ctypes.data, 1) == 0
    assert (ref_hits["obj"] >= 0).sum() > 20
    for f in ("obj", "face", "t", "N"):
        assert np.array_equal(ref_hits[f], hits[f]), f
    assert np.abs(ref_rgb - rgb).max() < 1e-9


def test_cli_fails_loudly_without_gpu(scene_dir, tmp_path):
    """ray_b200 (C++ RayTracer mirror + the reference's command line) has no CPU path either."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    exe = os.path.join(ROOT, "cs378hgraphics_raytracer_b200", "ray_b200")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "cs378hgraphics_raytracer_b200", "csrc"), "cli"])
    r = subprocess.run([exe, "-r", "5", "-w", "64", str(scene_dir / "prims_mixed.ray"), str(tmp_path / "o.bmp")], capture_output=True, text=True)
    assert r.returncode == 1
    assert "no CUDA device" in r.stderr
    assert not (tmp_path / "o.bmp").exists()


from conftest import GOLDEN


@pytest.mark.parametrize("name", sorted(GOLDEN))
def test_golden_vectors_hostsim(rtlib, hostsim, scene_path, name):
    """The committed reference outputs (tests/golden) against the host build of the device code:
    runs without the oracle library."""
    from cs378hgraphics_raytracer_b200.binding import HIT_DTYPE
    g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    arena = rtlib.load_ray_file(str(scene_path(GOLDEN[name])))
    w, h, depth = int(g["w"]), int(g["h"]), int(g["depth"])
    p = rtlib.Params.default(depth=depth)
    rays = np.ascontiguousarray(g["rays"])
    n = len(rays)
    hits = np.zeros(n, HIT_DTYPE)
    rgb = np.zeros((n, 3))
    assert hostsim.sim_trace_rays(arena.ptr, ctypes.byref(p), rays.ctypes.data, n, depth, hits.ctypes.data, rgb.ctypes.data, 1) == 0
    for f in ("obj", "face", "t", "N"):
        assert np.array_equal(hits[f], g[f]), f
    assert np.abs(rgb - g["rgb"]).max() < 1e-9
    img = np.zeros((h, w, 3), np.uint8)
    st = rtlib.Stats()
    assert hostsim.sim_render(arena.ptr, ctypes.byref(p), w, h, img.ctypes.data, ctypes.byref(st)) == 0
    d = np.abs(img.astype(np.int32) - g["frame"].astype(np.int32)).max(axis=2)
    assert (d <= 1).mean() >= 0.999
    assert [st.camera_rays, st.reflection_rays, st.refraction_rays, st.shadow_rays] == [int(x) for x in g["counts"]]


def test_png_and_bmp_output_stage(rtlib, tmp_path):
    """rt_write_image / rt_read_image (writeImage / readImage, images.cc:48-66; writePNG pngimage.cpp:226-285):
    the PNG is decoded by an independent reader (Pillow) and must be the frame with its rows flipped
    (buffer row 0 = bottom scan line = last row of the file); a PNG written by Pillow with its own
    adaptive scan-line filters must be read back the same way; BMP round trip; unknown extension -> BMP."""
    from PIL import Image
    rng = np.random.default_rng(3)
    h, w = 37, 53  # odd sizes: BMP row padding, PNG filter edge cases
    img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
    img[5:20, 10:40] = np.linspace(0, 255, 30, dtype=np.uint8)[None, :, None]  # smooth part: Pillow picks Sub/Up/Paeth here
    png = tmp_path / "o.PNG"
    rtlib.write_image(str(png), img)
    with Image.open(png) as im:
        assert im.mode == "RGB" and im.size == (w, h)
        assert np.array_equal(np.asarray(im), img[::-1])
    assert np.array_equal(rtlib.read_image(str(png)), img)
    theirs = tmp_path / "pil.png"
    Image.fromarray(img[::-1]).save(theirs, optimize=True)
    assert np.array_equal(rtlib.read_image(str(theirs)), img)
    bmp = tmp_path / "o.bmp"
    rtlib.write_image(str(bmp), img)
    assert bmp.read_bytes()[:2] == b"BM"
    assert np.array_equal(rtlib.read_image(str(bmp)), img)
    odd = tmp_path / "o.xyz"
    rtlib.write_image(str(odd), img)
    assert odd.read_bytes()[:2] == b"BM"
    # what the reference would read differently is refused, not guessed
    Image.fromarray(np.dstack([img[::-1], np.full((h, w), 200, np.uint8)])).save(tmp_path / "rgba.png")
    with pytest.raises(rtlib.RtError):
        rtlib.read_image(str(tmp_path / "rgba.png"))


def test_raycheck_tool(rtlib, tmp_path):
    """raycheck.py:21-34's RMS criterion and quickdiff's diff / montage images over the library's image I/O."""
    from cs378hgraphics_raytracer_b200 import raycheck
    rng = np.random.default_rng(5)
    ref = rng.integers(0, 256, size=(24, 40, 3), dtype=np.uint8)
    img = ref.copy()
    img[3, 7, 1] = np.uint8(int(ref[3, 7, 1]) ^ 1)  # one channel off by one
    r = raycheck.compare(img, ref)
    assert r["within1"] == 1.0 and r["max_abs"] == 1 and 0 < r["rms"] < 0.1
    assert r["rms_u8"] >= r["rms"]  # the reference's uint8 subtraction can only overstate
    rtlib.write_image(str(tmp_path / "a.png"), img)
    rtlib.write_image(str(tmp_path / "b.png"), ref)
    assert raycheck.main([str(tmp_path / "a.png"), str(tmp_path / "b.png"), "--diff", str(tmp_path / "d.png"), "--montage", str(tmp_path / "m.bmp")]) == 0
    assert rtlib.read_image(str(tmp_path / "d.png")).sum() == 1
    m = rtlib.read_image(str(tmp_path / "m.bmp"))
    assert m.shape == (24, 120, 3) and np.array_equal(m[:, 40:80], ref)
    assert raycheck.main([str(tmp_path / "a.png"), str(tmp_path / "b.png"), "--maxrms", "0.0001"]) == 1


def test_arena_cache_file(rtlib, scene_dir, tmp_path):
    """rt_arena_save / rt_arena_load: the cache file is the arena, byte for byte; a truncated,
    corrupted or foreign file is refused."""
    arena = rtlib.load_ray_file(str(scene_dir / "trimesh2.ray"))
    path = tmp_path / "scene.arena"
    rtlib.save_arena(str(path), arena)
    back = rtlib.load_arena(str(path))
    assert back.bytes() == arena.bytes()
    raw = path.read_bytes()
    (tmp_path / "short.arena").write_bytes(raw[:-64])
    (tmp_path / "long.arena").write_bytes(raw + b"\0")
    bad = bytearray(raw)
    bad[4] ^= 0xff  # magic / version word
    (tmp_path / "magic.arena").write_bytes(bytes(bad))
    (tmp_path / "text.arena").write_bytes(b"SBT-raytracer 1.0\n")
    for name in ("short.arena", "long.arena", "magic.arena", "text.arena", "missing.arena"):
        with pytest.raises(rtlib.RtError):
            rtlib.load_arena(str(tmp_path / name))


def test_bench_reference_arm_prints_one_json_line():
    """bench.py --impl reference (the reference's own CPU path through oracle/_ref) runs without a GPU and
    prints exactly one JSON line with the keys the measurement contract names."""
    import json
    import sys
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_oracle.so")):
        pytest.skip("oracle/_ref not built")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and d["metric"].startswith("Mrays/s")


@pytest.mark.parametrize("gen", ["dragon", "mixed"])
def test_full_size_arena_hashes(rtlib, tmp_path, gen):
    """The loader pinned at FULL size: tests/golden/arena_hashes.json holds sha256 and size of the arena flattened
    out of the reference's OWN parse + KdTree build (std::sort permutation included) of the full dragon stand-in
    (869,556 faces) and the full 1M-primitive scene (tests/golden/make_arena_hashes.py: half an hour of CPU per
    scene, run once).  rt_load_ray_file must reproduce them byte for byte - which is what makes ref_load_arena
    (full-size parity tests, CPU timing legs) a proven shortcut."""
    import hashlib
    import json
    import gen_scenes
    with open(os.path.join(ROOT, "tests", "golden", "arena_hashes.json")) as f:
        want = json.load(f)[gen]
    text = gen_scenes.GENERATORS[gen](want["scale"])
    assert len(text) == want["text_bytes"] and hashlib.sha256(text.encode()).hexdigest() == want["text_sha256"], "generator changed"
    path = tmp_path / (gen + ".ray")
    path.write_text(text)
    arena = rtlib.load_ray_file(str(path))
    data = arena.bytes()
    assert len(data) == want["arena_bytes"]
    assert hashlib.sha256(data).hexdigest() == want["arena_sha256"]


def test_adapt_aa_known_answer(oracle_mod):
    """/root/reference/adapt_aa.cpp, compiled unmodified by oracle/Makefile: its recursive
    quadrant subdivision of a quarter disc on a 3000 x 3000 grid prints `area = 0.786317 (check= 0.785398)`.  It is a
    different algorithm from RayTracer::adaptaa (that one is pinned by the frame tests), so this is a known-answer
    check of the program itself: the estimate, and that it lies within 0.2 % of pi / 4."""
    exe = os.path.join(ROOT, "oracle", "_ref", "adapt_aa")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/adapt_aa not built")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120).stdout
    m = re.match(r"area = ([0-9.]+) \(check= ([0-9.]+)\)", out)
    assert m, out
    area, check = float(m.group(1)), float(m.group(2))
    assert out.strip() == "area = 0.786317 (check= 0.785398)"
    assert abs(check - np.pi / 4) < 1e-6 and abs(area - check) < 2e-3


def test_overlapping_objects_mode_has_no_reference_result(oracle_mod, ref_scene_dir):
    """Why rt_render answers RT_ERR_UNSUPPORTED for -O o (Scene::discoverMat, scene.cpp:212-237): isect::checkObj
    (scene/ray.cpp:43-45) has no return statement.  g++ -O2 compiles it without a `ret` (the virtual call to
    Geometry::check is followed by padding and the next function, MaterialSceneObject::getMaterial, whose `ret` then pops a
    misaligned slot), so the UNMODIFIED reference dies with SIGSEGV as soon as discoverMat meets an entering surface behind a
    leaving one - which its own fixture for this mode, newScene/spheres_overlap.ray, does at any depth that recurses.
    Where checkObj is never reached the run completes (depth 0: shadow rays only)."""
    import sys
    prog = ("import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
            "import cs378hgraphics_raytracer_b200 as rt, oracle_binding as ob\n"
            "o = ob.Oracle(sys.argv[1]); ob.ref().ref_set_threads(1)\n"
            "img, st, _ = o.render(rt.Params.default(depth=int(sys.argv[2]), overlapping_objects=1), 64, 64)\n"
            "print('completed', int(img.sum()))\n") % (ROOT, os.path.join(ROOT, "tests"))
    scene = str(ref_scene_dir / "spheres_overlap.ray")
    ok = subprocess.run([sys.executable, "-c", prog, scene, "0"], capture_output=True, text=True, timeout=300)
    assert ok.returncode == 0 and "completed" in ok.stdout, ok.stderr[-500:]
    bad = subprocess.run([sys.executable, "-c", prog, scene, "5"], capture_output=True, text=True, timeout=300)
    assert bad.returncode != 0 and "completed" not in bad.stdout  # SIGSEGV (-11) in the builds seen so far
    # the compiled function really has no way back: no `ret` between its entry and the next symbol
    dis = subprocess.run(["objdump", "-d", "--no-show-raw-insn", "-C", os.path.join(ROOT, "oracle", "_ref", "libref_oracle.so")],
                         capture_output=True, text=True).stdout
    body = dis.split("<isect::checkObj(isect const&, isect const&)>:")[1].split("\n\n")[0]
    assert "call" in body and "ret" not in body, body


def test_colour_accumulator_limbs(hostsim):
    """csrc/device/rt_accum.cuh on the host: a sample's terms are held as two integer limbs per channel so that the total does
    not depend on the order they arrive in.  One term >= 2^-16 reads back as itself (authored constants such as 0.6 or 0.2 sit
    exactly on a byte boundary: 255 * 0.2 must stay 51.000000000000004 -> 51); a sum of such terms is the correctly rounded
    exact sum (checked with rationals); smaller terms are rounded at 2^-68; the limbs of a permuted list are identical; NaN adds
    nothing and infinities clamp to +-2^38."""
    from fractions import Fraction
    hostsim.sim_acc_sum.restype = ctypes.c_double
    hostsim.sim_acc_sum.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]

    def acc(xs):
        a = np.ascontiguousarray(xs, dtype=np.float64)
        limbs = np.zeros(2, dtype=np.int64)
        return hostsim.sim_acc_sum(a.ctypes.data, len(a), limbs.ctypes.data), limbs

    rng = np.random.default_rng(5)
    for x in list(rng.uniform(-4, 4, 2000)) + [0.6, 0.2, 1 / 3, 2.0 ** -16, -2.0 ** -16, 1.0, 255.0, 1e3, 2.0 ** 37, -0.1]:
        if abs(x) >= 2.0 ** -16:
            assert acc([x])[0] == x
    assert int(255.0 * acc([0.2])[0]) == 51 and int(255.0 * acc([0.6])[0]) == 153
    for x in rng.uniform(-1, 1, 2000) * 2.0 ** rng.integers(-80, -16, 2000):
        assert abs(acc([x])[0] - x) <= 2.0 ** -69
    for _ in range(2000):
        n = int(rng.integers(2, 60))
        xs = rng.uniform(-1, 1, n) * 2.0 ** rng.integers(-15, 8, n)
        assert acc(xs)[0] == float(sum(Fraction(float(x)) for x in xs))
    xs = rng.uniform(-1, 1, 500) * 2.0 ** rng.integers(-90, 8, 500)
    v0, l0 = acc(xs)
    assert abs(v0 - float(sum(Fraction(float(x)) for x in xs))) <= 500 * 2.0 ** -69
    for _ in range(20):
        v, l = acc(rng.permutation(xs))
        assert v == v0 and np.array_equal(l, l0)
    assert acc([1.0, -1.0])[0] == 0.0 and acc([1.0, 2.0 ** -60, -1.0])[0] == 2.0 ** -60
    assert acc([0.6, 0.6, 0.6])[0] == 0.6 + 0.6 + 0.6
    assert acc([float("nan"), 0.25])[0] == 0.25
    assert acc([float("inf")])[0] == 2.0 ** 38 and acc([-1e300])[0] == -2.0 ** 38
    assert acc([0.0, -0.0, 5e-324])[0] == 0.0
    # the resolve side: terms that were added (limbs) plus the camera rays' stored terms (one slot per lens ray) = one exact sum
    hostsim.sim_acc_read.restype = ctypes.c_double
    hostsim.sim_acc_read.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    for lenses in (0, 1, 6):
        added = np.ascontiguousarray(rng.uniform(0, 1, 40) * 2.0 ** rng.integers(-12, 2, 40))
        stored = np.ascontiguousarray(rng.uniform(0, 1, max(lenses, 1)))
        got = hostsim.sim_acc_read(added.ctypes.data, len(added), stored.ctypes.data, lenses)
        want = float(sum(Fraction(float(x)) for x in list(added) + list(stored[:lenses])))
        assert got == want
    one = np.array([0.6])
    assert hostsim.sim_acc_read(None, 0, one.ctypes.data, 1) == 0.6  # a camera ray that hits a flat emitter: that very double


def test_object_space_float_pretest_is_sound(rtlib, hostsim, scene_dir):
    """csrc/device/rt_prims.cuh: prim_miss_f skips the double-precision object-space test of a scene-level candidate when its
    float test proves that the ray's line misses the shape.  Brute force on the host build: every ray against every item,
    with the given directions and with directions aimed through a random point of each item's slightly enlarged box (so
    that every pair is a hit or a near miss), origins far outside (1000 x the extent) and inside the scene: no pair the
    double-precision test hits is ever rejected; on the rotated boxes / cylinders / spheres of the mixed scene the test
    removes most of the misses among the candidates whose box is entered."""
    hostsim.sim_primf_check.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    rng = np.random.default_rng(21)
    for name, extent, n in (("mixed_small.ray", 200.0, 400), ("hitchcock.ray", 6.0, 6000), ("prims_mixed.ray", 8.0, 6000),
                            ("lights_area.ray", 8.0, 3000)):
        arena = rtlib.load_ray_file(str(scene_dir / name))
        o = rng.uniform(-1.5 * extent, 1.5 * extent, (n, 3))
        o[:n // 8] *= 1e3
        o[n // 8:n // 4] *= 0.05
        d = rng.normal(size=(n, 3))
        d[n // 2:n // 2 + n // 10] = np.eye(3)[rng.integers(0, 3, n // 10)]  # zero components
        rays = np.ascontiguousarray(np.hstack([o, d]))
        for aimed in (0, 1, 2):
            out = (ctypes.c_uint64 * 6)()
            assert hostsim.sim_primf_check(ctypes.c_void_p(arena.ptr), rays.ctypes.data, n, aimed, out) == 0
            entered, rejected, bad, with_test, hits, bad_all = list(out)
            assert bad == 0 and bad_all == 0, (name, aimed, list(out))
            if aimed and name == "mixed_small.ray":
                assert with_test == entered and rejected > 0.9 * (entered - hits), list(out)
