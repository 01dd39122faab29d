#!/usr/bin/env python
"""Generates tests/golden/book_steps.json from the reference's archive/ listings (run in the
build container, where /root/reference exists):  python tests/golden/make_book_golden.py

For each intermediate step it parses the listing's own `make_shared<material>` / `world.add`
statements and its camera construction, and stores the world as rows
[cx, cy, cz, radius, kind, r, g, b, fuzz|ir] plus the equivalent 7-argument camera
(lookfrom, lookat, vfov, aspect, aperture, focus_dist).  tests/test_book_scenes.py compares the
facade's book_scenes.hpp with this file, so the scene pack is pinned to the reference's text."""
import json
import math
import os
import re
import sys

REF = os.environ.get("RT_REFERENCE_DIR", "/root/reference")
STEPS = {  # facade name -> listing
    "shiny_metal": 50, "fuzzed_metal": 52, "glass_pair": 55, "glass_single": 59, "hollow_glass": 61,
    "wide_angle": 63, "distant_view": 65, "zoomed_view": 66,
}
ENV = {"pi": math.pi, "cos": math.cos, "aspect_ratio": 16.0 / 9.0}
NUM = r"[^,()]+(?:\([^()]*\))?"  # a number or a simple call like cos(pi / 4)


def ev(x, env):
    return float(eval(x.strip(), {"__builtins__": {}}, env))


def parse(listing):
    src = open(os.path.join(REF, "archive", f"listing{listing}", "main.cc")).read()
    body = src[src.index("int main"):]
    env = dict(ENV)
    for m in re.finditer(r"auto\s+(\w+)\s*=\s*([^;]+);", body):
        if "make_shared" in m.group(2) or "point3" in m.group(2) or "vec3" in m.group(2):
            continue
        try:
            env[m.group(1)] = ev(m.group(2), env)
        except Exception:
            pass
    mats = {}
    for m in re.finditer(r"auto\s+(\w+)\s*=\s*make_shared<(\w+)>\((.*?)\);", body):
        name, kind, args = m.groups()
        if kind == "lambertian":
            c = re.match(r"color\((.*)\)$", args.strip()).group(1).split(",")
            mats[name] = [0] + [ev(v, env) for v in c] + [0.0]
        elif kind == "metal":
            mm = re.match(r"color\(([^)]*)\)\s*(?:,\s*(.*))?$", args.strip())
            c = mm.group(1).split(",")
            fuzz = ev(mm.group(2), env) if mm.group(2) else 0.0  # metal(color) of listing50 = no fuzz
            mats[name] = [1] + [ev(v, env) for v in c] + [min(fuzz, 1.0)]
        elif kind == "dielectric":
            mats[name] = [2, 1.0, 1.0, 1.0, ev(args, env)]
    rows = []
    for m in re.finditer(r"world\.add\(\s*make_shared<sphere>\(\s*point3\(([^)]*)\)\s*,\s*([^,]+),\s*(\w+)\s*\)\s*\);", body):
        c = [ev(v, env) for v in m.group(1).split(",")]
        rows.append(c + [ev(m.group(2), env)] + mats[m.group(3)])
    cm = re.search(r"camera\s+cam\s*(\(([^;]*)\))?\s*;", body)
    args = cm.group(2)
    cam = dict(lookfrom=[0.0, 0.0, 0.0], lookat=[0.0, 0.0, -1.0], vfov=90.0, aspect=16.0 / 9.0,
               aperture=0.0, focus_dist=1.0)  # camera(): archive/listing27/camera.h
    if args:
        parts = re.findall(r"\s*(?:point3|vec3)\([^)]*\)|[^,]+", args)
        parts = [p.strip() for p in parts if p.strip()]
        if len(parts) == 2:      # camera(vfov, aspect_ratio): archive/listing62/camera.h
            cam["vfov"] = ev(parts[0], env)
        elif len(parts) == 5:    # camera(lookfrom, lookat, vup, vfov, aspect): archive/listing64/camera.h
            cam["lookfrom"] = [ev(v, env) for v in re.match(r"point3\((.*)\)", parts[0]).group(1).split(",")]
            cam["lookat"] = [ev(v, env) for v in re.match(r"point3\((.*)\)", parts[1]).group(1).split(",")]
            cam["vfov"] = ev(parts[3], env)
        else:
            raise SystemExit(f"listing{listing}: unexpected camera arguments {parts}")
    lit = {k: int(re.search(rf"const int {k}\s*=\s*(\d+)", body).group(1))
           for k in ("image_width", "samples_per_pixel", "max_depth")}
    return dict(listing=f"archive/listing{listing}/main.cc", rows=rows, camera=cam, **lit)


def main():
    out = {name: parse(l) for name, l in STEPS.items()}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "book_steps.json")
    json.dump(out, open(path, "w"), indent=1)
    for k, v in out.items():
        print(k, v["listing"], len(v["rows"]), "spheres", v["camera"])


if __name__ == "__main__":
    sys.exit(main())
