#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REFERENCE'S OWN compiled code.

Runs oracle/_ref/ref_harness (which #includes /root/reference/src/cpu/main.cc unmodified;
built by `make -C oracle ref`) in this container.  The reference ships no golden vectors
(SURVEY.md §4), so these files ARE the pin: scene dumps, primary-hit tables, bit-exact small
renders of the sequential mt19937 stream, and converged f64 images for the statistical gates.

Usage:  python tests/golden/make_golden.py [--converged]      (needs /root/reference)
"""
import argparse
import os
import subprocess
import sys
import tempfile
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def run(args):
    subprocess.check_call([HARNESS] + [str(a) for a in args])


def read_primary(path, W, H):
    raw = open(path, "rb").read()
    N = W * H
    ids = np.frombuffer(raw[:4 * N], np.int32).reshape(H, W)
    t = np.frombuffer(raw[4 * N:12 * N], np.float64).reshape(H, W)
    nrm = np.frombuffer(raw[12 * N:36 * N], np.float64).reshape(H, W, 3)
    ff = np.frombuffer(raw[36 * N:37 * N], np.uint8).reshape(H, W)
    return ids, t, nrm, ff


def small(scene, W, H, spp, tmp):
    """scene dump + primary table + full sequential render (bit-exact pin material)."""
    sc, pr, li, pp = (os.path.join(tmp, f"{scene}.{x}") for x in ("scene", "prim", "lin", "ppm"))
    run(["--scene", scene, "--w", W, "--h", H, "--spp", spp, "--dump-scene", sc, "--primary", pr,
         "--linear", li, "--ppm", pp])
    spheres = np.fromfile(sc, np.float64).reshape(-1, 9)
    ids, t, nrm, ff = read_primary(pr, W, H)
    lin = np.fromfile(li, np.float64).reshape(H, W, 3)
    ppm = np.array(open(pp).read().split()[4:], np.uint8).reshape(H, W, 3)
    return dict(spheres=spheres, W=W, H=H, spp=spp, ids=ids, t=t, normal=nrm, front_face=ff,
                linear=lin, ppm=ppm)


def primary_only(scene, W, H, tmp):
    pr = os.path.join(tmp, f"{scene}.prim{W}")
    run(["--scene", scene, "--w", W, "--h", H, "--primary", pr])
    ids, t, nrm, ff = read_primary(pr, W, H)
    return dict(W=W, H=H, ids=ids, t=t, normal=nrm, front_face=ff)


def converged(scene, W, H, spp, tmp, procs):
    """Row bands in parallel processes; each band skips a different stretch of the RNG stream
    so bands are not driven by identical random numbers."""
    bands = np.linspace(0, H, procs + 1).astype(int)

    def one(k):
        out = os.path.join(tmp, f"{scene}.band{k}")
        run(["--scene", scene, "--w", W, "--h", H, "--spp", spp, "--rows",
             f"{bands[k]}:{bands[k + 1]}", "--skip", 1000003 * (k + 1), "--linear", out])
        return np.fromfile(out, np.float64).reshape(bands[k + 1] - bands[k], W, 3)

    with ThreadPoolExecutor(procs) as ex:
        parts = list(ex.map(one, range(procs)))
    lin = np.concatenate(parts, axis=0)
    return dict(W=W, H=H, spp=spp, mean=(lin / spp).astype(np.float32))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--converged", action="store_true", help="also (re)generate converged images")
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    a = ap.parse_args()
    if not os.path.exists(HARNESS):
        sys.exit("build oracle/_ref/ref_harness first: make -C oracle ref")
    with tempfile.TemporaryDirectory() as tmp:
        np.savez_compressed(os.path.join(HERE, "cover_small.npz"), **small("cover", 48, 27, 8, tmp))
        np.savez_compressed(os.path.join(HERE, "three_small.npz"), **small("three", 40, 22, 16, tmp))
        np.savez_compressed(os.path.join(HERE, "cover_primary_200x112.npz"),
                            **primary_only("cover", 200, 112, tmp))
        np.savez_compressed(os.path.join(HERE, "three_primary_200x112.npz"),
                            **primary_only("three", 200, 112, tmp))
        if a.converged:
            np.savez_compressed(os.path.join(HERE, "cover_converged_240x135.npz"),
                                **converged("cover", 240, 135, 2048, tmp, a.procs))
            np.savez_compressed(os.path.join(HERE, "three_converged_240x135.npz"),
                                **converged("three", 240, 135, 4096, tmp, a.procs))
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
