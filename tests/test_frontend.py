"""Host-side GML front end vs the reference's own front-end tests
(internal/gml/evaluator_test.go:26-211, lexer_test.go, include_test.go)."""
import os

import pytest

from go_raytracer_b200 import gml
from util import SCENES, eval_scene, render_args_to_lines


@pytest.mark.parametrize("name", ["canned", "sphere", "cube"])
def test_single_render_matches_reference_lowering_dump(name):
    # TestSingleRender, evaluator_test.go:168-211, against the reference's .out goldens
    st, calls = eval_scene(name)
    assert len(calls) == 1
    want = open(os.path.join(SCENES, name + ".out")).read().replace("\r\n", "\n").split("\n")
    assert render_args_to_lines(calls[0], st.IDMapping) == want


def test_features_gml_runs_without_render():
    st = gml.NewEvalState()

    def boom(e, a):
        raise AssertionError("unexpected call to render")
    st.Render = boom
    st.ParseAndEvalFile(os.path.join(SCENES, "features.gml"))
    assert st.Stack == []


@pytest.mark.parametrize("program,want", [
    ("1 { /x x x } apply addi", 2),
    ("1 /x { x } /f 2 /x f apply x addi", 3),
    ("3.75 frac", 0.75),
    ("-3.75 frac", -0.75),
    ("-1.0 /i i 0.0 lessf { i negf 0.5 addf } { i } if", 1.5),
    ("2.5 /i i 0.0 lessf { i negf 0.5 addf } { i } if", 2.5),
    ("1 /i { i /j /i i j addi } /f 2 f apply f apply", 4),
    ("7 -2 divi", -3), ("-7 2 modi", -1), ("2.9 floor", 2), ("-2.1 floor", -3),
])
def test_simple_eval(program, want):  # TestSimpleEval, evaluator_test.go:62-135
    st = gml.NewEvalState()
    st.ParseAndEval(program)
    assert st.Stack[-1] == want and type(st.Stack[-1]) is type(want)


@pytest.mark.parametrize("program,prefix", [("1\n2.0\naddi", "3:1:"), ("1 /x\n1 missing", "2:3:"), ("addi", "1:1:")])
def test_eval_error_positions(program, prefix):  # TestEvalErrorPositions, evaluator_test.go:26-59
    with pytest.raises(gml.GMLError) as ei:
        gml.NewEvalState().ParseAndEval(program)
    assert str(ei.value).startswith(prefix)


def test_render_without_hook_is_an_error():
    with pytest.raises(gml.GMLError, match="render function not set"):
        gml.NewEvalState().ParseAndEval(open(os.path.join(SCENES, "sphere.gml")).read())


def test_union_order_and_transform_composition():
    st = gml.NewEvalState()
    calls = []
    st.Render = lambda e, a: calls.append(a)
    st.ParseAndEval('{ /v /u /face 1.0 1.0 1.0 point 1.0 0.0 1.0 } sphere /s '
                    's 1.0 0.0 0.0 translate 2.0 uscale  s 0.0 5.0 0.0 translate union /sc '
                    '1.0 1.0 1.0 point [] sc 1 90.0 4 4 "x" render')
    u = calls[0].Scene
    # top of stack first (evaluator.go:1091-1101); newest transform acts first (M = T * S)
    assert u.objects[0].transform_mat[1][3] == 5.0
    assert u.objects[1].transform_mat[0] == (2.0, 0.0, 0.0, 1.0)


def test_lexer_numbers_comments_binders():
    toks = []
    lx = gml.Lexer('1 -2 3.5 1e3 -4.5E-2 /x x_y-z "a\\n" true % comment\n /* block */ [ ] { }')
    while True:
        t = lx.next_token()
        if t.kind == "EOF":
            break
        toks.append((t.kind, t.text))
    assert toks == [("Integer", "1"), ("Integer", "-2"), ("Float", "3.5"), ("Float", "1e3"), ("Float", "-4.5E-2"),
                    ("Binder", "/x"), ("Ident", "x_y-z"), ("String", "a\n"), ("Boolean", "true"), ("LBracket", "["),
                    ("RBracket", "]"), ("LCurly", "{"), ("RCurly", "}")]


def test_include_guard_and_cycle(tmp_path):
    (tmp_path / "a.ins").write_text('#ifndef A\n#define A\n1 /one\n#endif\n')
    (tmp_path / "main.gml").write_text('#include "a.ins"\n#include "a.ins"\none one addi')
    st = gml.NewEvalState()
    st.ParseAndEvalFile(str(tmp_path / "main.gml"))
    assert st.Stack == [2]
    (tmp_path / "c1.gml").write_text('#include "c2.gml"\n')
    (tmp_path / "c2.gml").write_text('#include "c1.gml"\n')
    with pytest.raises(gml.GMLSyntaxError, match="include cycle"):
        gml.NewEvalState().ParseAndEvalFile(str(tmp_path / "c1.gml"))
    with pytest.raises(gml.GMLSyntaxError, match="#include"):
        gml.NewEvalState().ParseAndEval('#include "does-not-exist.ins"')


def test_mul_mat_reference_vector_and_affine_inverse():
    """TestMulMat (internal/prim/vec_test.go:188-226): the one matrix vector the reference holds; and the affine cofactor
    inverse (vec.go:319-365, untested in the reference) against the transforms it is used on: M * M^-1 = I to rounding, and
    a point sent through ObjectToWorld and back comes home."""
    from go_raytracer_b200 import prim
    a = tuple(tuple(float(4 * i + j + 1) for j in range(4)) for i in range(4))
    want = ((90, 100, 110, 120), (202, 228, 254, 280), (314, 356, 398, 440), (426, 484, 542, 600))
    assert prim.mul_mat(a, a) == tuple(tuple(float(x) for x in row) for row in want)
    m = prim.identity()
    for t in (prim.scale(2.0, 0.5, 3.0), prim.rotate_x(30.0), prim.rotate_y(-75.0), prim.rotate_z(120.0), prim.translate(1.5, -2.0, 7.0)):
        m = prim.mul_mat(t, m)
    inv = prim.inverse(m)
    eye = prim.mul_mat(m, inv)
    for i in range(4):
        for j in range(4):
            assert abs(eye[i][j] - (1.0 if i == j else 0.0)) < 1e-12
    p = prim.Vec3(0.3, -0.7, 0.9)
    q = prim.mul_point(inv, prim.mul_point(m, p))
    assert abs(q.x - p.x) < 1e-12 and abs(q.y - p.y) < 1e-12 and abs(q.z - p.z) < 1e-12
    d = prim.mul_dir(m, prim.Vec3(0.0, 0.0, 0.0))
    assert (d.x, d.y, d.z) == (0.0, 0.0, 0.0)                      # directions ignore the translation column
