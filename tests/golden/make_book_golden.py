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


# ---- the diffuse-only steps: no material classes yet, the rule is written inside ray_color
DIFFUSE_STEPS = {"diffuse_first": 34, "diffuse_gamma": 36, "diffuse_lambertian": 38, "diffuse_hemisphere": 40}
RULES = [("random_in_unit_sphere(", 3), ("random_unit_vector(", 0), ("random_in_hemisphere(", 4),
         ("random_on_hemisphere(", 4)]  # -> RT_DIFFUSE_BALL, RT_LAMBERTIAN, RT_DIFFUSE_HEMI


def latest(kind, name, upto):
    """the newest `name` (color.h, camera.h, ...) a step of the archive sees: the archive keeps only
    the files a listing changed"""
    for k in range(upto, 0, -1):
        f = os.path.join(REF, "archive", f"{kind}{k}", name)
        if os.path.exists(f):
            return open(f).read()
    raise SystemExit(f"no {name} up to {kind}{upto}")


def rule_of(text):
    body = text[text.index("ray_color"):]
    for needle, kind in RULES:
        if needle in body:
            return kind
    raise SystemExit("no diffuse rule found")


def parse_diffuse(listing):
    src = open(os.path.join(REF, "archive", f"listing{listing}", "main.cc")).read()
    kind = rule_of(src)
    refl = float(re.search(r"return\s+([0-9.]+)\s*\*\s*ray_color", src).group(1))
    rows = []
    for m in re.finditer(r"world\.add\(make_shared<sphere>\(point3\(([^)]*)\),\s*([0-9.]+)\)\);", src):
        rows.append([float(v) for v in m.group(1).split(",")] + [float(m.group(2)), kind, refl, refl, refl, 0.0])
    lit = {k: int(re.search(rf"const int {k}\s*=\s*(\d+)", src).group(1))
           for k in ("image_width", "samples_per_pixel", "max_depth")}
    flags = 0 if "sqrt" in latest("listing", "color.h", listing) else 128  # RT_FLAG_LINEAR_OUTPUT
    cam = dict(lookfrom=[0.0, 0.0, 0.0], lookat=[0.0, 0.0, -1.0], vfov=90.0, aspect=16.0 / 9.0,
               aperture=0.0, focus_dist=1.0)
    return dict(listing=f"archive/listing{listing}/main.cc", rows=rows, camera=cam, flags=flags, **lit)


# ---- the CUDA program's history (archive-gpu/imageN: every directory is a complete program)
GPU_STEPS = {"gpu_hemisphere_depth10": 7, "gpu_hemisphere": 9, "gpu_lambertian": 10, "gpu_lambertian_gamma": 12,
             "gpu_shiny_metal": 13, "gpu_fuzzed_metal": 14, "gpu_glass_pair": 16, "gpu_glass_single": 17,
             "gpu_hollow_glass": 18, "gpu_wide_angle": 19, "gpu_distant_view": 20, "gpu_zoomed_view": 21,
             "gpu_defocus_blur": 22}


def fl(x, env):
    """a float literal / expression of the CUDA program, as the f32 value the program holds"""
    import numpy as np
    return float(np.float32(eval(re.sub(r"(?<=[0-9.])f\b", "", x.strip()), {"__builtins__": {}}, env)))


def parse_gpu(image):
    d = os.path.join(REF, "archive-gpu", f"image{image}")
    src, cam_h = open(os.path.join(d, "main.cu")).read(), open(os.path.join(d, "camera.h")).read()
    color_h = open(os.path.join(d, "color.h")).read()
    env = {"pi": math.pi, "R": math.cos(math.pi / 4)}
    # the C++ side holds floats: R = cosf(pi / 4) is an f32 value
    import numpy as np
    env["R"] = float(np.cos(np.float32(math.pi) / np.float32(4.0), dtype=np.float32))
    mats = {}
    for m in re.finditer(r"material\s*\*(\w+)\s*=\s*new\s+(\w+)\((.*?)\);", src):
        name, kind, args = m.groups()
        if kind == "lambertian":
            c = re.match(r"color\((.*)\)$", args.strip()).group(1).split(",")
            mats[name] = [0] + [fl(v, env) for v in c] + [0.0]
        elif kind == "metal":
            mm = re.match(r"color\(([^)]*)\)\s*(?:,\s*(.*))?$", args.strip())
            fuzz = fl(mm.group(2), env) if mm.group(2) else 0.0
            mats[name] = [1] + [fl(v, env) for v in mm.group(1).split(",")] + [min(fuzz, 1.0)]
        elif kind == "dielectric":
            mats[name] = [2, 1.0, 1.0, 1.0, fl(args, env)]
    rows = []
    for m in re.finditer(r"world->add\(\s*new sphere\(point3\(([^)]*)\),\s*([^,()]+?)(?:,\s*(\w+))?\)\);", src):
        c = [fl(v, env) for v in m.group(1).split(",")]
        if m.group(3):
            mat = mats[m.group(3)]
        else:  # no material classes yet: the rule is inside ray_color (camera.h)
            kind = rule_of(cam_h)
            refl = float(re.search(r"cur_attenuation\s*\*=\s*([0-9.]+)f", cam_h).group(1))
            mat = [kind, refl, refl, refl, 0.0]
        rows.append(c + [fl(m.group(2), env)] + mat)

    def lit(name, default=None):
        m = re.search(rf"(?:cam->|int\s+|float\s+){name}\s*=\s*([^;]+);", cam_h)
        return m.group(1).strip() if m else default
    vfov = fl(lit("vfov", "90.0f"), env)
    lookfrom = re.search(r"cam->lookfrom\s*=\s*point3\(([^)]*)\)", cam_h)
    lookat = re.search(r"cam->lookat\s*=\s*point3\(([^)]*)\)", cam_h)
    lf = [fl(v, env) for v in lookfrom.group(1).split(",")] if lookfrom else [0.0, 0.0, 0.0]
    la = [fl(v, env) for v in lookat.group(1).split(",")] if lookat else [0.0, 0.0, -1.0]
    da, fd = lit("defocus_angle"), lit("focus_dist")
    da = fl(da, env) if da else 0.0
    fd = fl(fd, env) if fd else 1.0
    W = int(lit("image_width"))
    aspect = 16.0 / 9.0
    H = int(np.float32(W) / np.float32(aspect))
    f32 = lambda x: float(np.float32(x))
    aperture = 0.0 if da <= 0 else 2.0 * f32(fd) * math.tan(f32(da) / 2.0 * math.pi / 180.0)
    cam = dict(lookfrom=lf, lookat=la, vfov=vfov, aspect=W / H, aperture=aperture, focus_dist=f32(fd))
    return dict(listing=f"archive-gpu/image{image}", rows=rows, camera=cam,
                flags=16 | (0 if "sqrt" in color_h else 128),  # RT_FLAG_V4_MODEL | RT_FLAG_LINEAR_OUTPUT
                image_width=W, samples_per_pixel=int(lit("samples_per_pixel")), max_depth=int(lit("max_depth")))


def main():
    out = {name: parse(l) for name, l in STEPS.items()}
    for v in out.values():
        v["flags"] = 0
    out.update({name: parse_diffuse(l) for name, l in DIFFUSE_STEPS.items()})
    out.update({name: parse_gpu(i) for name, i in GPU_STEPS.items()})
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "book_steps.json")
    json.dump(out, open(path, "w"), indent=1)
    for k, v in out.items():
        print(k, v["listing"], len(v["rows"]), "spheres", v["camera"])


if __name__ == "__main__":
    sys.exit(main())
