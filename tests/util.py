"""Shared helpers for the test-suite."""
import os

import numpy as np

from go_raytracer_b200 import gml, imageio, lowering

HERE = os.path.dirname(os.path.abspath(__file__))
SCENES = os.path.join(HERE, "golden", "scenes")
IMAGES = os.path.join(HERE, "golden", "images")

# render-call file name -> golden PNG (raytracer_test.go:71-134)
GOLDENS = {
    "canned": {"canned.ppm": "example_canned.png"},
    "sphere": {"sphere.ppm": "example_sphere.png"},
    "cube": {"cube.ppm": "example_cube.png"},
    "cylinder": {"cylinder0.ppm": "example_cylinder_front.png", "cylinder1.ppm": "example_cylinder_bottom.png",
                 "cylinder2.ppm": "example_cylinder_top.png", "cylinder3.ppm": "example_cylinder_back.png"},
}


def eval_scene(name):
    """Evaluate tests/golden/scenes/<name>.gml with a recording Render hook -> (state, [RenderArgs])."""
    st = gml.NewEvalState()
    calls = []
    st.Render = lambda e, a: calls.append(a)
    st.ParseAndEvalFile(os.path.join(SCENES, name + ".gml"))
    return st, calls


def scene_blobs(name):
    st, calls = eval_scene(name)
    return [(a.File, lowering.lower(a, st)) for a in calls]


def golden(png):
    return imageio.read_png(os.path.join(IMAGES, png))


def agreement(got, want, tol=1):
    """Fraction of pixels whose RGB channels all differ by <= tol."""
    d = np.abs(got[..., :3].astype(np.int32) - want[..., :3].astype(np.int32)).max(axis=2)
    return float((d <= tol).mean())


def render_args_to_lines(args, ids):
    """The lowered-scene dump of evaluator_test_helpers.go:15-145 (the .out format)."""
    lines, ind = [], [0]

    def add(s):
        lines.append("    " * ind[0] + s)

    ff = lambda x: f"{x:<+10.2f}"
    f3 = lambda v: f"{ff(v.x)} {ff(v.y)} {ff(v.z)}"
    add(f"render {args.Width} {args.Height} {args.File}")
    ind[0] += 1
    add("fov: " + ff(args.Fov))
    add(f"depth: {args.Depth}")
    if not args.BgColorStart.is_zero() or not args.BgColorEnd.is_zero():
        add("background-gradient:")
        ind[0] += 1
        add("p1: " + f3(args.BgColorStart))
        add("p2: " + f3(args.BgColorEnd))
        ind[0] -= 1
    add("ambient: " + f3(args.AmbientLight))
    for l in args.Lights:
        add("light:")
        ind[0] += 1
        add("position: " + f3(l.position))
        add("color: " + f3(l.color))
        ind[0] -= 1

    def surf(fn):
        add("surface:")
        ind[0] += 1
        if fn.closure is not None:
            add("code: " + gml.token_list_string(fn.closure.code))
            if fn.closure.env:
                add("env:")
                ind[0] += 1
                for i in sorted(fn.closure.env):
                    v = fn.closure.env[i]
                    val = v.debug_string(ids) if isinstance(v, gml.VClosure) else gml.value_string(v)
                    add(f"{ids.id_to_name[i]}: {val}")
                ind[0] -= 1
        else:
            m = fn.material
            add("color: " + f3(m.color))
            for n, v in (("reflectivity", m.reflectivity), ("fuzz", m.fuzziness), ("transparency", m.transparency),
                         ("refractiveIndex", m.refractive_index), ("kd", m.kd), ("ks", m.ks), ("n", m.specular_exponent)):
                add(f"{n}: " + ff(v))
        ind[0] -= 1

    def obj(o):
        if isinstance(o, gml.Union):
            add("union:")
            ind[0] += 1
            for c in o.objects:
                obj(c)
            ind[0] -= 1
        else:
            add(o.kind + ":")
            ind[0] += 1
            add("xform:")
            ind[0] += 1
            for row in o.transform_mat:
                add("".join(ff(x) for x in row))
            ind[0] -= 1
            surf(o.surface_fn)
            ind[0] -= 1

    obj(args.Scene)
    return lines
