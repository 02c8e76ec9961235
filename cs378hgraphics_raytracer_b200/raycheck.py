#!/usr/bin/env python3
"""raycheck-style image comparison for frames written by ray_b200 / rt_write_image.

Mirrors the acceptance check of the reference's grader (reference ray/raycheck.py:21-34: RMS of
image - ref over all channel values, PASS below --maxrms, default 10) and the per-image output of
ray/quickdiff:16-29 (a difference image and a 3x1 montage image | reference | difference), without
scipy / ImageMagick: files are read and written through the library's own readImage / writeImage
replacements (rt_read_image, rt_write_image: .png and .bmp).

    python -m cs378hgraphics_raytracer_b200.raycheck image.png ref.png [--maxrms 10] [--diff d.png] [--montage m.png]

Two RMS figures are printed: `rms` is the true one; `rms_u8` reproduces raycheck.py literally, whose
`image - ref` is a uint8 subtraction (a channel that is lower by 1 counts as 255).  The parity bar
of this repository (channels within 1/255 on >= 99.9 % of the pixels) is reported as `within1`.
Exit status 0 = PASS by the reference's criterion applied to the true RMS.
"""
import argparse
import math
import sys

import numpy as np

from .binding import read_image, write_image


def compare(image, ref):
    """image, ref: (h, w, 3) uint8.  Returns a dict of the figures described in the module docstring."""
    if image.shape != ref.shape:
        raise ValueError("image sizes differ: %r vs %r" % (image.shape, ref.shape))
    a = image.reshape(-1, 1)
    b = ref.reshape(-1, 1)
    wrapped = (a - b).astype(np.float64)  # uint8 arithmetic, as raycheck.py does it
    true = a.astype(np.float64) - b.astype(np.float64)
    per_px = np.abs(image.astype(np.int16) - ref.astype(np.int16)).max(axis=-1)
    return {"rms": float(np.linalg.norm(true) / math.sqrt(true.size)),
            "rms_u8": float(np.linalg.norm(wrapped) / math.sqrt(wrapped.size)),
            "within1": float((per_px <= 1).mean()), "exact": float((per_px == 0).mean()), "max_abs": int(per_px.max())}


def difference_image(image, ref):
    return np.abs(image.astype(np.int16) - ref.astype(np.int16)).astype(np.uint8)


def montage(image, ref):
    """image | reference | difference, left to right (quickdiff's `montage -tile 3x1 -geometry +0+0`)."""
    return np.concatenate([image, ref, difference_image(image, ref)], axis=1)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("image")
    ap.add_argument("ref")
    ap.add_argument("--maxrms", type=float, default=10.0)
    ap.add_argument("--diff")
    ap.add_argument("--montage")
    args = ap.parse_args(argv)
    image, ref = read_image(args.image), read_image(args.ref)
    r = compare(image, ref)
    ok = r["rms"] < args.maxrms
    print("[%s] %s RMS: %.6f (raycheck.py's uint8 formula: %.6f) within 1/255: %.4f%% exact: %.4f%% max |diff|: %d" %
          ("PASS" if ok else "WARNING", args.image, r["rms"], r["rms_u8"], 100 * r["within1"], 100 * r["exact"], r["max_abs"]))
    if args.diff:
        write_image(args.diff, difference_image(image, ref))
    if args.montage:
        write_image(args.montage, montage(image, ref))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
