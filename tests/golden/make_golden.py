#!/usr/bin/env python3
"""Golden vectors from the UNMODIFIED reference (oracle/_ref/libref_oracle.so, built from
/root/reference by oracle/Makefile): for each small scene the reference's Scene::intersect result
of every camera ray of a 64-wide frame (object, face, t, N), traceRay colours, the rendered RGB8 frame
and the ray counters.  Run in the build container (needs /root/reference or a prebuilt oracle/_ref):
    python tests/golden/make_golden.py
The .npz files are committed; tests/test_gpu_parity.py::test_golden_vectors and
tests/test_host_cpu.py::test_golden_vectors_hostsim compare against them, so parity is still checked
on a box where the oracle library is absent."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scenes"))

W = 64
DEPTH = 5
SCENES = {"hitchcock": ("gen", "hitchcock", 1.0), "trimesh2": ("gen", "trimesh2", 1.0), "dragon_small": ("gen", "dragon", 0.02),
          "prims_mixed": ("file", "prims_mixed.ray"), "lights_area": ("file", "lights_area.ray"), "tex_params": ("tex",),
          # the reference's own fixtures (ray/newScene, copied by tests/fetch_ref_fixtures.py)
          "ref_box_cyl_opaque_shadow_spotlight": ("ref", "box_cyl_opaque_shadow_spotlight.ray"), "ref_concrete_1": ("ref", "concrete_1.ray"),
          "ref_concrete_2": ("ref", "concrete_2.ray"), "ref_concrete_3": ("ref", "concrete_3.ray"), "ref_distance": ("ref", "distance.ray"),
          "ref_lava_box": ("ref", "lava_box.ray"), "ref_spheres_overlap": ("ref", "spheres_overlap.ray")}


def scene_file(spec, d, name):
    """Path of the scene's .ray file (written into d unless it is one of the reference's fixtures)."""
    import gen_scenes
    if spec[0] == "ref":
        import fetch_ref_fixtures
        return os.path.join(fetch_ref_fixtures.fetch(), spec[1])
    if spec[0] == "gen":
        text = gen_scenes.GENERATORS[spec[1]](spec[2])
    elif spec[0] == "tex":
        import conftest
        gen_scenes.gen_cubemap(os.path.join(d, "cubemap"), 64)
        text = conftest.tex_params_scene(d)
    else:
        text = open(os.path.join(ROOT, "tests", "scenes", spec[1])).read()
    path = os.path.join(d, name + ".ray")
    with open(path, "w") as f:
        f.write(text)
    return path


def main():
    import cs378hgraphics_raytracer_b200 as rt
    import oracle_binding as ob
    out_dir = os.path.dirname(os.path.abspath(__file__))
    with tempfile.TemporaryDirectory() as d:
        only = sys.argv[1:]
        for name, spec in SCENES.items():
            if only and name not in only:
                continue
            path = scene_file(spec, d, name)
            arena = rt.load_ray_file(path)
            orc = ob.Oracle(path)
            h = ob.frame_height(W, orc.aspect)
            rays = ob.camera_rays(arena.camera(), W, h)
            p = rt.Params.default(depth=DEPTH)
            hits, rgb = orc.trace_rays(p, rays, DEPTH)
            img, st, _ = orc.render(p, W, h)
            np.savez_compressed(os.path.join(out_dir, name + ".npz"), w=W, h=h, depth=DEPTH, rays=rays, obj=hits["obj"], face=hits["face"],
                                t=hits["t"], N=hits["N"], rgb=rgb, frame=img,
                                counts=np.array([st.camera_rays, st.reflection_rays, st.refraction_rays, st.shadow_rays], dtype=np.uint64))
            print(name, W, h, "hits", int((hits["obj"] >= 0).sum()), "counts", st.camera_rays, st.reflection_rays, st.refraction_rays, st.shadow_rays)


if __name__ == "__main__":
    main()
