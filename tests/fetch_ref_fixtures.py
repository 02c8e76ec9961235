#!/usr/bin/env python3
"""Copies the reference's own executable fixtures - the seven ray/newScene/*.ray scenes and the two
BMP textures they map ("the only executable fixtures") - from /root/reference into
tests/ref_fixtures/, which is git-ignored (fixture DATA of the reference, not product source) but is
not gpurun-ignored, so it travels to the GPU box with the snapshot like the prebuilt oracle does.

    python tests/fetch_ref_fixtures.py

Called by tests/conftest.py (fixture `ref_scene_dir`) and by __graft_entry__.build() whenever
/root/reference is present; on a box without /root/reference the already copied files are used."""
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/ray/newScene"
DST = os.path.join(ROOT, "tests", "ref_fixtures")
FILES = ["box_cyl_opaque_shadow_spotlight.ray", "concrete_1.ray", "concrete_2.ray", "concrete_3.ray", "distance.ray", "lava_box.ray",
         "spheres_overlap.ray", "lava_bump.bmp", "lava_texture.bmp"]
SCENES = [f for f in FILES if f.endswith(".ray")]


def fetch():
    """Returns the fixture directory, or None when neither the reference nor a previous copy exists."""
    if os.path.isdir(SRC):
        os.makedirs(DST, exist_ok=True)
        for f in FILES:
            s, d = os.path.join(SRC, f), os.path.join(DST, f)
            if not os.path.exists(d) or os.path.getsize(d) != os.path.getsize(s):
                shutil.copyfile(s, d)
    if all(os.path.exists(os.path.join(DST, f)) for f in FILES):
        return DST
    return None


if __name__ == "__main__":
    print(fetch())
