"""Pins the CPU oracle (oracle/oracle.c) to the reference's own vectors:
 - the seven golden PNGs (raytracer_test.go:71-134), byte for byte;
 - the nine analytic cylinder cases of cylinder_test.go:21-165;
 - Go's published PCG test vector and a few libm identities of the Go stdlib restatements;
 - the surface bytecode against the Python restatement of EvalSurfaceFn."""
import ctypes
import math
import random

import numpy as np
import pytest

import oracle
from go_raytracer_b200 import gml, lowering
from go_raytracer_b200.gml import gomath
from util import GOLDENS, eval_scene, golden, scene_blobs

D = ctypes.c_double
IDENT = (D * 12)(1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0)
CYL = 18


@pytest.mark.parametrize("name", ["cylinder", "sphere", "cube", "canned"])
def test_goldens_byte_exact(name):
    for file, blob in scene_blobs(name):
        img, st = oracle.render(blob, threads=8)
        want = golden(GOLDENS[name][file])
        assert img.shape == want.shape
        assert np.array_equal(img[..., :3], want[..., :3]), f"{file}: oracle differs from the reference's golden"
        assert (img[..., 3] == 255).all()
        assert st.rays_primary == 4 * img.shape[0] * img.shape[1]


def test_crop_and_thread_count_do_not_change_pixels():
    (_, blob), = scene_blobs("cube")
    full, _ = oracle.render(blob, threads=8)
    crop, st = oracle.render(blob, threads=3, crop=(500, 377, 530, 415))
    assert np.array_equal(crop[377:415, 500:530], full[377:415, 500:530])
    assert st.rays_primary == 4 * 30 * 38
    assert crop[:377].sum() == 0


def test_pcg_matches_go_test_vector():
    out = (ctypes.c_uint64 * 3)()
    oracle.lib().oracle_pcg(1, 2, 3, out)
    assert list(out) == [0xc4f5a58656eef510, 0x9dcec3ad077dec6c, 0xc8d04605312f8088]


def test_go_math_restatements():
    L = oracle.lib()
    assert L.oracle_go_tan(math.pi / 4) == 1.0            # glibc gives 0.9999999999999999
    assert 2.0 / L.oracle_go_tan(90.0 * math.pi / 180.0 / 2.0) == 2.0
    rng = random.Random(7)
    for _ in range(2000):
        x = rng.uniform(-20, 20)
        assert L.oracle_go_sin(x) == gomath.sin(x) and L.oracle_go_cos(x) == gomath.cos(x)   # C vs Python restatement
        assert abs(L.oracle_go_sin(x) - math.sin(x)) < 1e-15
        y = rng.uniform(-1, 1)
        assert abs(L.oracle_go_acos(y) - math.acos(y)) < 1e-15
        b = rng.uniform(0, 1.5)
        for n in (0.0, 1.0, 5.0, 32.0, 50.0):
            assert abs(L.oracle_go_pow(b, n) - b ** n) <= 2e-14 * max(b ** n, 1e-300)  # repeated squaring: a few ulps
    for yy, xx in [(1, 1), (1, -1), (-1, -1), (-1, 1), (0.0, -1.0), (-0.0, -1.0), (1.0, 0.0), (0.3, 2.0)]:
        assert abs(L.oracle_go_atan2(yy, xx) - math.atan2(yy, xx)) < 4e-16
    assert L.oracle_go_pow(0.0, 0.0) == 1.0 and L.oracle_go_pow(0.0, 50.0) == 0.0 and L.oracle_go_pow(0.7, 1.0) == 0.7


def _isect(o, d):
    t, p, f = D(), (D * 3)(), ctypes.c_int()
    hit = oracle.lib().oracle_intersect(CYL, IDENT, (D * 3)(*o), (D * 3)(*d), ctypes.byref(t), p, ctypes.byref(f))
    return (t.value, tuple(p), f.value) if hit else None


def test_cylinder_known_answers():  # cylinder_test.go:21-113
    t, p, f = _isect((-2, 0.5, 0), (1, 0, 0))
    assert f == 0 and abs(t - 1.0) < 1e-9 and math.dist(p, (-1, 0.5, 0)) < 1e-9
    t, p, f = _isect((0, 2, 0), (0, -1, 0))
    assert f == 1 and abs(t - 1.0) < 1e-9
    t, p, f = _isect((0, -2, 0), (0, 1, 0))
    assert f == 2 and abs(t - 2.0) < 1e-9
    t, p, f = _isect((0, 0.5, 0), (0, 1, 0))          # from inside: nearest cap
    assert f == 1 and abs(t - 0.5) < 1e-9
    assert _isect((5, -1, 0), (0, 1, 0)) is None
    assert _isect((-2, 5, 0), (1, 0, 0)) is None
    assert _isect((2, 0.5, 0), (1, 0, 0)) is None      # behind the ray


def test_cylinder_normals_and_invalid_face():  # cylinder_test.go:115-165
    L = oracle.lib()
    n = (D * 3)()
    assert L.oracle_normal(CYL, IDENT, IDENT, 0, (D * 3)(1, 0.5, 0), n) == 0 and math.dist(n, (1, 0, 0)) < 1e-9
    assert L.oracle_normal(CYL, IDENT, IDENT, 1, (D * 3)(0.2, 1, 0.3), n) == 0 and math.dist(n, (0, 1, 0)) < 1e-9
    assert L.oracle_normal(CYL, IDENT, IDENT, 2, (D * 3)(0.2, 0, 0.3), n) == 0 and math.dist(n, (0, -1, 0)) < 1e-9
    assert L.oracle_normal(CYL, IDENT, IDENT, 99, (D * 3)(0, 0, 0), n) != 0


def test_sphere_inside_origin_does_not_hit():   # raytracer.go:98-102: the far root is never used
    t, p, f = D(), (D * 3)(), ctypes.c_int()
    assert oracle.lib().oracle_intersect(16, IDENT, (D * 3)(0, 0, 0), (D * 3)(0, 0, 1), ctypes.byref(t), p, ctypes.byref(f)) == 0
    assert oracle.lib().oracle_intersect(16, IDENT, (D * 3)(0, 0, -3), (D * 3)(0, 0, 1), ctypes.byref(t), p, ctypes.byref(f)) == 1
    assert t.value == 2.0


@pytest.mark.parametrize("name", ["sphere", "cube", "cylinder", "checked-cube", "cube2"])
def test_bytecode_equals_reference_closure_semantics(name):
    """compile_surface + the oracle's interpreter == EvalSurfaceFn run on the closure itself."""
    st, calls = eval_scene(name)
    args = calls[0]
    blob = lowering.lower(args, st)
    d = lowering.describe(blob)
    leaves, todo = [], [args.Scene]
    while todo:
        o = todo.pop(0)
        todo.extend(o.objects) if isinstance(o, gml.Union) else leaves.append(o)
    nodes = [n for n in d["nodes"] if n["kind"] >= 16]
    rng = random.Random(3)
    checked = 0
    # match leaves to nodes through the lowering order (depth-first, `a` before `b`)
    order = []

    def walk(o):
        if isinstance(o, gml.Union):
            for c in o.objects:
                walk(c)
        else:
            order.append(o)
    walk(args.Scene)
    for leaf, node in zip(order, nodes):
        if leaf.surface_fn.closure is None:
            continue
        nfaces = 6 if leaf.kind == "cube" else 3
        for _ in range(300):
            face, u, v = rng.randrange(nfaces), rng.uniform(-3, 3), rng.uniform(-3, 3)
            if rng.random() < 0.5:
                u, v = rng.random(), rng.random()
            try:
                want = gml.EvalSurfaceFn(face, u, v, st, leaf.surface_fn).as_tuple()
            except gml.ErrArrayIndexOutOfBounds:
                del st.Stack[:]
                with pytest.raises(oracle.OracleError):     # same fault on the compiled path
                    oracle.eval_surface(blob, node["surface"], face, u, v)
                continue
            got = oracle.eval_surface(blob, node["surface"], face, u, v)
            assert got == want, (face, u, v)
            checked += 1
    assert checked > 0


_TAIL = ' cube /c 1.0 1.0 1.0 point [] c 1 90.0 4 4 "x" render'
RUNTIME_CLOSURES = {
    # the two arms of a run-time `if` leave different closures; the survivor is applied later
    "if_picks_a_closure": '{ /v /u /face u 0.5 lessf { { 1.0 0.0 0.0 point } } { { 0.0 v 1.0 point } } if /pick pick apply 0.8 0.2 4.0 }',
    # `get` with a run-time index into an array of closures (faces 3-5 are out of range: the reference faults, evaluator.go:989-995)
    "get_picks_a_closure": '[ { /q 1.0 0.0 0.0 point } { /q 0.0 q 0.0 point } { /q 0.0 0.0 1.0 point } ] /tab '
                           '{ /v /u /face tab face get /f u f apply 0.7 0.3 2.0 }',
    # a run-time-selected closure as an operand of `if`
    "choice_as_if_operand": '{ /v /u /face u v lessf { { 0.2 } } { { 0.9 } } if /g  face 2 lessi g { 0.5 } if /r r r r point 1.0 0.0 1.0 }',
    # recursion through apply with a run-time exit: halve x until it is below 0.1, count the steps
    "recursion_with_runtime_exit": '{ /self /x /n x 0.1 lessf { n } { n 1 addi x 0.5 mulf self self apply } if } /loop '
                                   '{ /v /u /face 0 u loop loop apply real 0.1 mulf /s s s s point 1.0 0.0 1.0 }',
}


@pytest.mark.parametrize("case", sorted(RUNTIME_CLOSURES))
def test_bytecode_runs_closures_chosen_at_run_time(case):
    """apply / if on a closure that is only known per hit (evaluator.go:604-610, :1005-1019): the compiler carries it as a
    choice - a selector register and the candidate closures - and unfolds the application into a chain of branches; the
    compiled code, run by the oracle's interpreter, equals EvalSurfaceFn on the closure itself, fault for fault."""
    st = gml.NewEvalState()
    calls = []
    st.Render = lambda e, a: calls.append(a)
    st.ParseAndEval(RUNTIME_CLOSURES[case] + _TAIL)
    blob = lowering.lower(calls[0], st)
    assert lowering.describe(blob)["surfaces"][0]["kind"] == 1          # bytecode, not a folded constant
    fn = calls[0].Scene.surface_fn
    rng = random.Random(7)
    ok = faults = 0
    for _ in range(500):
        face, u, v = rng.randrange(6), rng.random(), rng.random()
        try:
            want = gml.EvalSurfaceFn(face, u, v, st, fn).as_tuple()
        except gml.ErrArrayIndexOutOfBounds:
            del st.Stack[:]
            with pytest.raises(oracle.OracleError):
                oracle.eval_surface(blob, 0, face, u, v)
            faults += 1
            continue
        assert oracle.eval_surface(blob, 0, face, u, v) == want, (face, u, v)
        ok += 1
    assert ok > 100 and (faults > 100) == (case == "get_picks_a_closure")


def test_bytecode_recursion_beyond_the_bound_is_a_run_time_fault():
    """No call stack on the device: an unrolled recursion that runs out of levels (registers) faults at run time instead of
    failing to compile - the paths that stay within the bound still give the reference's values."""
    st = gml.NewEvalState()
    calls = []
    st.Render = lambda e, a: calls.append(a)
    # count down from floor(u * 1000), adding 0.001 per level: up to 999 levels deep
    st.ParseAndEval('{ /self /acc /n n 1 lessi { acc } { n 1 subi acc 0.001 addf self self apply } if } /down '
                    '{ /v /u /face u 1000.0 mulf floor 0.0 down down apply /s s s s point 1.0 0.0 1.0 }' + _TAIL)
    blob = lowering.lower(calls[0], st)
    fn = calls[0].Scene.surface_fn
    assert oracle.eval_surface(blob, 0, 0, 0.0035, 0.0) == gml.EvalSurfaceFn(0, 0.0035, 0.0, st, fn).as_tuple()      # 3 levels
    with pytest.raises(oracle.OracleError):
        oracle.eval_surface(blob, 0, 0, 0.9, 0.0)                                                                    # 900 levels


def test_bytecode_faults_like_the_reference():
    st = gml.NewEvalState()
    calls = []
    st.Render = lambda e, a: calls.append(a)
    st.ParseAndEval('[ 1.0 2.0 ] /arr { /v /u /face 0.5 0.5 arr face get point 1.0 0.0 1.0 } cube /c '
                    '1.0 1.0 1.0 point [] c 1 90.0 4 4 "x" render')
    blob = lowering.lower(calls[0], st)
    assert oracle.eval_surface(blob, 0, 1, 0.0, 0.0)[2] == 2.0
    with pytest.raises(oracle.OracleError):
        oracle.eval_surface(blob, 0, 5, 0.0, 0.0)         # index out of range -> evaluator.go:989-995
    with pytest.raises(gml.ErrArrayIndexOutOfBounds):
        gml.EvalSurfaceFn(5, 0.0, 0.0, st, calls[0].Scene.surface_fn)


def test_glass_scene_really_branches():
    """The level-overflow GPU test needs a scene whose ray tree is binary (a material may spawn a
    reflection AND a refraction ray, raytracer.go:512, 532): more children than camera rays within two levels, so a
    level sized for one ray per camera ray must overflow."""
    from go_raytracer_b200 import scenes
    blob = scenes.to_blob(scenes.glass(80, 50, depth=2))
    _, st = oracle.render(blob, threads=4)
    assert st.rays_refract > 0
    assert st.rays_reflect + st.rays_refract > 1.2 * st.rays_primary
    deep, st5 = oracle.render(scenes.to_blob(scenes.glass(80, 50, depth=5)), threads=4)
    assert st5.rays_reflect + st5.rays_refract > 2.0 * st5.rays_primary
    assert deep[..., 3].min() == 255


def test_every_icfp_fixture_evaluates_lowers_and_renders():
    """internal/gml/testdata: every runnable sample scene goes through the front end (large.gml needs the
    iterative evaluator: thousands of nested applies), lowers to a blob and renders with the oracle."""
    from util import scene_blobs
    n = 0
    for name in ["checked-cube", "chess", "cone", "cone-fractal", "cube2", "dice", "ellipsoid", "fov", "fractal", "golf", "holes", "house",
                 "intercyl", "large", "pipe", "rotate", "snowgoon", "spheres", "spotlight"]:
        for file, blob in scene_blobs(name):
            img, st = oracle.render(blob, threads=8)
            assert st.rays_primary == 4 * img.shape[0] * img.shape[1], (name, file)
            assert (img[..., 3] == 255).all()
            n += 1
    assert n >= 30
