#!/usr/bin/env python3
"""Pins the product's scene loader at FULL size against the reference's own loader.

For each full-size stand-in scene (BASELINE.json configs[2] dragon, 869,556 faces; configs[4] 1M mixed
primitives) the UNMODIFIED reference (oracle/_ref/libref_oracle.so) parses the .ray text itself
(RayTracer::loadScene -> Parser, Scene::add, Trimesh::conclude, KdTree build, kdTree.h:23-78 incl. its
std::sort permutation), the harness flattens the reference's objects into the arena format, and
sha256 + size of that arena are written to tests/golden/arena_hashes.json.  A CPU test
(tests/test_host_cpu.py::test_full_size_arena_hashes) asserts that rt_load_ray_file reproduces the
hashes, which is what makes `ref_load_arena` (the harness filling the reference's classes from the
product's arena for the full-size parity tests and the CPU timing legs) a proven shortcut rather than
an assumed one.

The reference's tree build is super-linear: about half an hour per scene on one core.  Run once in
the build container (needs /root/reference or a prebuilt oracle/_ref):

    python tests/golden/make_arena_hashes.py [dragon] [mixed]
"""
import hashlib
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scenes"))
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "arena_hashes.json")


def main():
    import gen_scenes
    import oracle_binding as ob
    gens = sys.argv[1:] or ["dragon", "mixed"]
    for gen in gens:
        text = gen_scenes.GENERATORS[gen](1.0)
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, gen + ".ray")
            with open(path, "w") as f:
                f.write(text)
            t0 = time.time()
            orc = ob.Oracle(path)          # the reference's own parse + tree build
            t1 = time.time()
            arena = orc.flatten()
        rec = {"generator": gen, "scale": 1.0, "text_sha256": hashlib.sha256(text.encode()).hexdigest(), "text_bytes": len(text),
               "arena_sha256": hashlib.sha256(arena).hexdigest(), "arena_bytes": len(arena), "reference_load_seconds": round(t1 - t0, 1)}
        print(json.dumps(rec), flush=True)
        # merge under a lock-free read-modify-write (the two scenes may be run as two processes)
        cur = {}
        if os.path.exists(OUT):
            with open(OUT) as f:
                cur = json.load(f)
        cur[gen] = rec
        tmp = OUT + ".tmp%d" % os.getpid()
        with open(tmp, "w") as f:
            json.dump(cur, f, indent=1, sort_keys=True)
        os.replace(tmp, OUT)


if __name__ == "__main__":
    main()
