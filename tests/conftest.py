import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scenes"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def scene_dir(tmp_path_factory):
    """Stand-in scenes are generated, not stored (scenes/gen_scenes.py, fixed seeds)."""
    import gen_scenes
    d = tmp_path_factory.mktemp("scenes")
    specs = {"hitchcock.ray": ("hitchcock", 1.0), "trimesh2.ray": ("trimesh2", 1.0), "dragon_small.ray": ("dragon", 0.02),
             "mixed_small.ray": ("mixed", 0.004)}
    for name, (gen, scale) in specs.items():
        with open(d / name, "w") as f:
            f.write(gen_scenes.GENERATORS[gen](scale))
    gen_scenes.gen_cubemap(str(d / "cubemap"), 64)
    for f in os.listdir(os.path.join(ROOT, "tests", "scenes")):
        with open(os.path.join(ROOT, "tests", "scenes", f)) as src, open(d / f, "w") as dst:
            dst.write(src.read())
    with open(d / "tex_params.ray", "w") as f:
        f.write(tex_params_scene(str(d)))
    return d


def tex_params_scene(d):
    """Every texture-mapped material parameter at once (reference scene/material.cpp:84-158, material.h:197-213):
    vec3 parameters mapped relative to the scene file (Parser.cpp:1276-1292), the scalar ones - shininess
    (renormalised x128, material.h:204-211) and index - by the path as written (Parser.cpp:1294-1310, hence
    absolute here), a mapped kt met by shadow rays, and the Box bump normal (Box.cpp:99-108).  Analytic
    primitives other than Box / Square leave uv at (0,0): they sample texel (0,0)."""
    a = lambda n: '"%s/cubemap/%s.bmp"' % (d, n)
    r = lambda n: '"cubemap/%s.bmp"' % n
    return """SBT-raytracer 1.0
camera { position=(0.5,2.5,9); viewdir=(-0.05,-0.25,-1); updir=(0,1,0); fov=48; aspectratio=1.25; }
ambient_light { color=(0.25,0.2,0.2); }
point_light { position=(3,6,5); color=(1,0.95,0.9); constant_attenuation_coeff=0.25; linear_attenuation_coeff=0.02; quadratic_attenuation_coeff=0.002; }
directional_light { direction=(-0.4,-1,-0.3); color=(0.4,0.45,0.5); }
translate(-2.2,0.3,0, rotate(0.3,1,0.1,0.6, scale(2.2, box { material={ diffuse=map(%s); specular=map(%s); shininess=map(%s); ambient=map(%s); bump=map(%s); } })))
translate(2.0,0.4,0.5, rotate(1,0.2,0,0.4, scale(1.8, box { material={ diffuse=(0.3,0.3,0.35); transmissive=map(%s); index=map(%s); reflective=map(%s); specular=(0.6,0.6,0.6); shininess=50; emissive=map(%s); } })))
translate(0.2,2.6,-1, scale(0.9, sphere { material={ diffuse=map(%s); specular=map(%s); shininess=map(%s); reflective=(0.2,0.2,0.2); } }))
translate(0,-1.2,0, rotate(1,0,0,-1.5707963, scale(16, square { material={ diffuse=map(%s); ambient=(0.3,0.3,0.3); reflective=map(%s); specular=map(%s); shininess=map(%s); } })))
translate(-0.3,0.2,3.2, rotate(0,1,0,0.3, scale(0.7,0.7,1.2, cylinder { material={ diffuse=(0.2,0.6,0.7); transmissive=map(%s); index=1.3; } })))
""" % (r("px"), r("ny"), a("pz"), r("nx"), r("py"),
       r("nz"), a("px"), r("py"), r("nx"),
       r("pz"), r("px"), a("ny"),
       r("py"), r("nz"), r("px"), a("nx"),
       r("ny"))


@pytest.fixture(scope="session")
def ref_scene_dir():
    """The reference's own fixtures, ray/newScene/*.ray + BMPs (tests/fetch_ref_fixtures.py)."""
    import fetch_ref_fixtures
    d = fetch_ref_fixtures.fetch()
    if d is None:
        pytest.skip("tests/ref_fixtures missing and /root/reference not present")
    import pathlib
    return pathlib.Path(d)


@pytest.fixture(scope="session")
def cubemap_faces(scene_dir):
    return [str(scene_dir / "cubemap" / (n + ".bmp")) for n in ("px", "nx", "py", "ny", "pz", "nz")]


@pytest.fixture(scope="session")
def rtlib():
    import cs378hgraphics_raytracer_b200 as rt
    if not os.path.exists(rt.LIB_PATH):
        rt.build_library()
    return rt


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle_binding
    if not oracle_binding.available():
        if os.path.isdir("/root/reference/ray/src"):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref", "-j8"])
        else:
            pytest.skip("oracle/_ref/libref_oracle.so missing and /root/reference not present")
    return oracle_binding


REF_SCENES = ["ref:box_cyl_opaque_shadow_spotlight.ray", "ref:concrete_1.ray", "ref:concrete_2.ray", "ref:concrete_3.ray", "ref:distance.ray",
              "ref:lava_box.ray", "ref:spheres_overlap.ray"]


@pytest.fixture(scope="session")
def scene_path(scene_dir, request):
    """name -> path: generated / hand-written scenes by file name, the reference's own fixtures as "ref:<file>"."""
    def resolve(name):
        if name.startswith("ref:"):
            return request.getfixturevalue("ref_scene_dir") / name[4:]
        return scene_dir / name
    return resolve


# tests/golden/<name>.npz (outputs of the unmodified reference, tests/golden/make_golden.py) -> scene
GOLDEN = {"hitchcock": "hitchcock.ray", "trimesh2": "trimesh2.ray", "dragon_small": "dragon_small.ray", "prims_mixed": "prims_mixed.ray",
          "lights_area": "lights_area.ray", "tex_params": "tex_params.ray"}
GOLDEN.update({"ref_" + s[4:-4]: s for s in REF_SCENES})
