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
    return d


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
