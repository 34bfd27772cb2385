"""GML stack evaluator - the host-side caller of the `render` boundary.

Mirror of /root/reference/internal/gml/evaluator.go (EvalState :43-50, values
:52-150, scene nodes :163-287, EvalOneStep :365-414, builtins :557-1199) and
environment.go.  The reference keeps this layer in Go; here it is restated in
Python because no Go toolchain exists in the build environment, so that the
parity tests can feed *identical GML text* to the new renderer.  It is not on the
accelerated path and is deliberately a plain interpreter.

Value representation: GML int -> Python ``int`` (wrapped to 64 bits like Go's
``int``), real -> ``float``, bool -> ``bool``, string -> ``str``; everything else is
a small class below.  ``type(v) is ...`` checks are used throughout because
``bool`` is a subclass of ``int`` in Python.

Builtins marked EXTENSION are not registered by the reference
(evaluator.go:557-601); they are the ICFP-2000 operators BASELINE.json's
north_star names (cone, intersect, light, spotlight, real, acos, asin).
"""
import math

from .. import prim
from ..prim import Vec3
from . import gomath
from .parser import (Array, Binder, BoolLiteral, FloatLiteral, Function, GMLSyntaxError, Identifier, IDMapping,
                     IntLiteral, Parser, Pos, StringLiteral, format_float, token_list_string, token_string)


class GMLError(Exception):
    """Any evaluation error (the reference returns `error`)."""


class ErrEmptyStack(GMLError):
    pass


class ErrUnboundIdentifier(GMLError):
    pass


class ErrArrayIndexOutOfBounds(GMLError):
    pass


class ErrTypeMismatch(GMLError):
    pass


# ------------------------------------------------------------------ values

class VClosure:
    __slots__ = ("code", "env")

    def __init__(self, code, env):
        self.code, self.env = code, env

    def __str__(self):
        return f"Closure({token_list_string(self.code)}, env={env_string(self.env)})"

    def debug_string(self, ids):
        return f"Closure({token_list_string(self.code)}, env={env_string(self.env, ids)})"


class VArray:
    __slots__ = ("elements",)

    def __init__(self, elements):
        self.elements = elements

    def __str__(self):
        return "[" + ", ".join(value_string(v) for v in self.elements) + "]"


class Material:
    """evaluator.go:136-150."""
    __slots__ = ("color", "reflectivity", "fuzziness", "transparency", "refractive_index", "kd", "ks", "specular_exponent")

    def __init__(self, color, reflectivity=0.0, fuzziness=0.0, transparency=0.0, refractive_index=0.0, kd=0.0, ks=0.0,
                 specular_exponent=0.0):
        self.color = color
        self.reflectivity, self.fuzziness = reflectivity, fuzziness
        self.transparency, self.refractive_index = transparency, refractive_index
        self.kd, self.ks, self.specular_exponent = kd, ks, specular_exponent

    def as_tuple(self):
        return (self.color.x, self.color.y, self.color.z, self.reflectivity, self.fuzziness, self.transparency,
                self.refractive_index, self.kd, self.ks, self.specular_exponent)

    def __str__(self):
        return (f"Material(Color: {self.color} Refl: {prim.go_v(self.reflectivity)} Kd: {prim.go_v(self.kd)} "
                f"Ks: {prim.go_v(self.ks)} N: {prim.go_v(self.specular_exponent)})")


class VSurfaceFn:
    """Exactly one of closure / material is set (evaluator.go:93-101)."""
    __slots__ = ("closure", "material")

    def __init__(self, closure=None, material=None):
        self.closure, self.material = closure, material


class SceneObject:
    __slots__ = ()


class _Leaf(SceneObject):
    __slots__ = ("surface_fn", "transform_mat")
    kind = "?"

    def __init__(self, surface_fn, transform_mat=None):
        self.surface_fn, self.transform_mat = surface_fn, transform_mat

    def transform(self, mat):
        """evaluator.go:176-184: newest transform is multiplied on the RIGHT, i.e.
        it acts on the object first."""
        m = mat if self.transform_mat is None else prim.mul_mat(self.transform_mat, mat)
        return type(self)(self.surface_fn, m)


class Sphere(_Leaf):
    kind = "sphere"

    def __str__(self):
        return "Sphere(C: [0, 0, 0], R: 1)"


class Cube(_Leaf):
    kind = "cube"

    def __str__(self):
        return "Cube(...)"


class Cylinder(_Leaf):
    kind = "cylinder"

    def __str__(self):
        return "Cylinder(...)"


class Cone(_Leaf):  # EXTENSION
    kind = "cone"

    def __str__(self):
        return "Cone(...)"


class Plane(_Leaf):
    kind = "plane"

    def __str__(self):
        return "Pt: [0, 0, 0], Normal: [0, 1, 0]"


class Union(SceneObject):
    __slots__ = ("objects",)

    def __init__(self, objects):
        self.objects = objects

    def transform(self, mat):
        return Union([o.transform(mat) for o in self.objects])

    def __str__(self):
        return "Union([" + " ".join(str(o) for o in self.objects) + "])"


class Difference(SceneObject):
    __slots__ = ("a", "b")

    def __init__(self, a, b):
        self.a, self.b = a, b

    def transform(self, mat):
        return Difference(self.a.transform(mat), self.b.transform(mat))

    def __str__(self):
        return f"Difference({self.a}, {self.b})"


class Intersect(SceneObject):  # EXTENSION
    __slots__ = ("a", "b")

    def __init__(self, a, b):
        self.a, self.b = a, b

    def transform(self, mat):
        return Intersect(self.a.transform(mat), self.b.transform(mat))

    def __str__(self):
        return f"Intersect({self.a}, {self.b})"


class PointLight:
    __slots__ = ("position", "color")

    def __init__(self, position, color):
        self.position, self.color = position, color

    def __str__(self):
        return f"PointLight(pos={self.position}, color={self.color})"


class DirectionalLight:  # EXTENSION: `dir color light`
    __slots__ = ("direction", "color")

    def __init__(self, direction, color):
        self.direction, self.color = direction, color

    def __str__(self):
        return f"DirectionalLight(dir={self.direction}, color={self.color})"


class SpotLight:  # EXTENSION: `pos at color cutoff exp spotlight`
    __slots__ = ("position", "at", "color", "cutoff", "exp")

    def __init__(self, position, at, color, cutoff, exp):
        self.position, self.at, self.color, self.cutoff, self.exp = position, at, color, cutoff, exp

    def __str__(self):
        return f"SpotLight(pos={self.position}, at={self.at}, color={self.color}, cutoff={self.cutoff}, exp={self.exp})"


class RenderArgs:
    """evaluator.go:14-28."""
    __slots__ = ("AmbientLight", "Lights", "Scene", "Depth", "Fov", "Width", "Height", "File", "BgColorStart", "BgColorEnd")

    def __init__(self, **kw):
        self.BgColorStart, self.BgColorEnd = Vec3(), Vec3()
        for k, v in kw.items():
            setattr(self, k, v)


def value_string(v):
    t = type(v)
    if t is bool:
        return "true" if v else "false"
    if t is int:
        return str(v)
    if t is float:
        return format_float(v)
    if t is str:
        return '"' + v.replace("\\", "\\\\").replace('"', '\\"') + '"'
    return str(v)


def env_string(env, ids=None):
    parts = []
    for i in sorted(env):
        v = env[i]
        if ids is not None:
            name = ids.id_to_name.get(i, f"{i} (?)")
            val = v.debug_string(ids) if isinstance(v, VClosure) else value_string(v)
            parts.append(f"{name}: {val}")
        else:
            parts.append(f"{i}: {value_string(v)}")
    return "{" + ", ".join(parts) + "}"


def _wrap64(v):
    v &= 0xFFFFFFFFFFFFFFFF
    return v - (1 << 64) if v >> 63 else v


_TYPE_NAMES = {int: "gml.VInt", float: "gml.VReal", bool: "gml.VBool", str: "gml.VString"}


def _tname(t):
    return _TYPE_NAMES.get(t, getattr(t, "__name__", str(t)))


# ------------------------------------------------------------------ evaluator

_NO_ENV = object()      # frame marker: nothing to restore when the frame ends


class EvalState:
    """evaluator.go:43-50.  ``Render`` is the drop-in hook: a callable
    ``(state, RenderArgs) -> None`` that raises on error."""

    def __init__(self):
        self.CurrToken = None
        self.Stack = []
        self.IDMapping = IDMapping()
        self.Env = {}
        self.Render = None
        self.Debugger = None
        self._frames = None       # Eval's explicit work stack while a program runs

    # -- parsing ----------------------------------------------------------
    def Parse(self, text):
        return Parser.from_text(text, self.IDMapping).parse()

    def ParseFile(self, path):
        return Parser.from_file(path, self.IDMapping).parse()

    def ParseAndEval(self, text):
        self.Eval(self.Parse(text))

    def ParseAndEvalFile(self, path):
        self.Eval(self.ParseFile(path))

    # -- stack ------------------------------------------------------------
    def Push(self, v):
        self.Stack.append(v)

    def Pop(self):
        if not self.Stack:
            tok = self.CurrToken
            pos = tok.pos if tok is not None else Pos()
            raise ErrEmptyStack(f"{pos.prefix()}empty stack: token: {token_string(tok) if tok is not None else '?'}")
        return self.Stack.pop()

    def PopValue(self, *types):
        v = self.Pop()
        if types and not any(type(v) is t or (t is SceneObject and isinstance(v, SceneObject)) for t in types):
            raise self._mismatch(types[0], v)
        return v

    def _mismatch(self, want, got):
        tok = self.CurrToken
        pos = tok.pos.prefix() if tok is not None else ""
        what = token_string(tok) if tok is not None else "?"
        return ErrTypeMismatch(f"{pos}error evaluating {what}: type mismatch: expected {_tname(want)}, "
                               f"got {value_string(got)} ({_tname(type(got))})")

    def _err(self, cls, msg):
        tok = self.CurrToken
        pos = tok.pos.prefix() if tok is not None else ""
        what = token_string(tok) if tok is not None else "?"
        return cls(f"{pos}error evaluating {what}: {msg}")

    # -- evaluation -------------------------------------------------------
    def Eval(self, program):
        """evalProgram (evaluator.go:345-363).  The reference recurses through EvalClosure on Go's growable stack; here
        closure calls are frames on an explicit work stack, so GML recursion (large.gml nests thousands of applies)
        is not bounded by the Python stack: `apply` / `if` push a frame (EvalClosure below) and this loop runs it."""
        outer = self._frames
        frames = self._frames = [[program, 0, _NO_ENV]]
        try:
            while frames:
                fr = frames[-1]
                code, pc = fr[0], fr[1]
                if pc >= len(code):
                    frames.pop()
                    if fr[2] is not _NO_ENV:
                        self.Env = fr[2]                      # EvalClosure's deferred restore (evaluator.go:433-438)
                    continue
                fr[1] = pc + 1
                self.EvalOneStep(code[pc])
        except BaseException:
            for fr in reversed(frames):                       # the deferred restores run on the way out of an error too
                if fr[2] is not _NO_ENV:
                    self.Env = fr[2]
            raise
        finally:
            self._frames = outer

    def EvalOneStep(self, tok):
        self.CurrToken = tok
        t = type(tok)
        if t is IntLiteral or t is FloatLiteral or t is BoolLiteral or t is StringLiteral:
            self.Stack.append(tok.value)
        elif t is Identifier:
            fn = BUILTINS.get(tok.name)
            if fn is not None:
                fn(self)
            else:
                v = self.Env.get(tok.id)
                if v is not None:
                    self.Stack.append(v)
                else:
                    # EXTENSION operators rank below user bindings so that every program
                    # the reference accepts keeps its meaning (canned.gml binds /light).
                    fn = EXTENSION_BUILTINS.get(tok.name)
                    if fn is None:
                        raise ErrUnboundIdentifier(f"{tok.pos.prefix()}unbound identifier: {tok.name}")
                    fn(self)
        elif t is Binder:
            self.Env[tok.id] = self.Pop()
        elif t is Function:
            self.Stack.append(VClosure(tok.body, dict(self.Env)))
        elif t is Array:
            saved = self.Stack
            self.Stack = []
            try:
                self.Eval(tok.elements)
                elems = self.Stack
            finally:
                self.Stack = saved
            saved.append(VArray(elems))
        else:
            raise GMLError(f"unknown token: {tok!r}")
        if self.Debugger is not None and self.Debugger() == 1:
            raise GMLError("evaluation was aborted by the user")

    def CallClosure(self, closure):
        """`apply` / `if` (evaluator.go:604-610, 1005-1019): nothing follows the call in the builtin, so inside Eval's loop
        the closure becomes a new frame instead of a Python-level recursion."""
        if self._frames is None:
            return self.EvalClosure(closure)
        self._frames.append([closure.code, 0, self.Env])
        self.Env = dict(closure.env)

    def EvalClosure(self, closure):
        """evaluator.go:433-438: runs the closure to completion, then restores the environment."""
        saved = self.Env
        self.Env = dict(closure.env)
        try:
            self.Eval(closure.code)
        finally:
            self.Env = saved

    def Clone(self):
        c = EvalState()
        c.CurrToken, c.Stack, c.Env = self.CurrToken, list(self.Stack), dict(self.Env)
        c.IDMapping, c.Debugger = self.IDMapping.clone(), self.Debugger
        return c


def NewEvalState():
    return EvalState()


# ------------------------------------------------------------------ surface functions

def referenced_vars(closure):
    """evaluator.go:635-668: every Identifier that is not a builtin, at any depth."""
    out, todo = [], list(closure.code)
    bound_here = _binder_names(closure.code) if EXTENSION_BUILTINS else ()
    while todo:
        nxt = []
        for tok in todo:
            if type(tok) is Identifier:
                if tok.name in BUILTINS:
                    continue
                if tok.name in EXTENSION_BUILTINS and tok.id not in closure.env and tok.name not in bound_here:
                    continue
                out.append(tok.name)
            elif type(tok) is Array:
                nxt.extend(tok.elements)
            elif type(tok) is Function:
                nxt.extend(tok.body)
        todo = nxt
    return out


def _binder_names(code):
    names, todo = set(), list(code)
    while todo:
        tok = todo.pop()
        if type(tok) is Binder:
            names.add(tok.name)
        elif type(tok) is Array:
            todo.extend(tok.elements)
        elif type(tok) is Function:
            todo.extend(tok.body)
    return names


def EvalSurfaceFn(face, u, v, state, surface_fn):
    """evaluator.go:672-727.  Returns a Material."""
    if surface_fn.material is not None:
        return surface_fn.material
    if state is None:
        raise GMLError("nil GML eval state")
    if surface_fn.closure is None:
        raise GMLError("surfaceFn in invalid state")
    state.Push(int(face))
    state.Push(float(u))
    state.Push(float(v))
    state.EvalClosure(surface_fn.closure)
    first = state.Pop()
    if type(first) is Material:
        return first
    if type(first) is not float:
        raise state._mismatch(float, first)
    n = first
    ks = state.PopValue(float)
    kd = state.PopValue(float)
    color = state.PopValue(Vec3)
    return Material(color, kd=kd, ks=ks, specular_exponent=n, reflectivity=ks)


def maybe_simplify_surface_fn(closure, state):
    """evaluator.go:729-750: constant-fold closures that mention no variable."""
    if not referenced_vars(closure):
        try:
            mat = EvalSurfaceFn(0, 0.0, 0.0, state, VSurfaceFn(closure=closure))
        except GMLError as e:
            raise GMLError(f"error while precomputing closure: {e}") from e
        return VSurfaceFn(material=mat)
    return VSurfaceFn(closure=closure)


# ------------------------------------------------------------------ builtins

def _fdiv(a, b):
    try:
        return a / b
    except ZeroDivisionError:
        if a != a or a == 0.0:
            return math.nan
        return math.copysign(math.inf, a) * math.copysign(1.0, b)


def _idiv(e, a, b):
    if b == 0:
        raise e._err(GMLError, "integer divide by zero")
    q = abs(a) // abs(b)
    return _wrap64(q if (a < 0) == (b < 0) else -q)


def _imod(e, a, b):
    if b == 0:
        raise e._err(GMLError, "integer divide by zero")
    r = abs(a) % abs(b)
    return -r if a < 0 else r


def _bin(t, op):
    def run(e):
        b = e.PopValue(t)
        a = e.PopValue(t)
        e.Push(op(e, a, b))
    return run


def _un(t, op):
    def run(e):
        e.Push(op(e, e.PopValue(t)))
    return run


def _clampf(e, x):
    if x < 0:
        return 0.0
    if x > 1:
        return 1.0
    return x


def _floor(e, x):
    if x != x or math.isinf(x):
        return -(1 << 63)  # Go: int64(NaN/Inf) is implementation-defined; amd64 gives MinInt64
    return _wrap64(math.floor(x))


def _frac(e, x):
    if x != x or math.isinf(x) or abs(x) >= 9.3e18:
        return x - float(-(1 << 63))
    return x - float(int(x))


def _sqrt(e, x):
    return math.sqrt(x) if x >= 0 else (x if x != x else math.nan)


def _apply(e):
    e.CallClosure(e.PopValue(VClosure))


def _if(e):
    f = e.PopValue(VClosure)
    t = e.PopValue(VClosure)
    c = e.PopValue(bool)
    e.CallClosure(t if c else f)


def _get(e):
    i = e.PopValue(int)
    arr = e.PopValue(VArray)
    n = len(arr.elements)
    if i < 0 or i >= n:
        raise e._err(ErrArrayIndexOutOfBounds, f"array index out of bounds: {i} vs {n}")
    e.Push(arr.elements[i])


def _point(e):
    z = e.PopValue(float)
    y = e.PopValue(float)
    x = e.PopValue(float)
    e.Push(Vec3(x, y, z))


def _material(e):
    f = [e.PopValue(float) for _ in range(7)][::-1]
    color = e.PopValue(Vec3)
    e.Push(Material(color, reflectivity=f[0], fuzziness=f[1], transparency=f[2], refractive_index=f[3], kd=f[4], ks=f[5],
                    specular_exponent=f[6]))


def _leaf(cls):
    def run(e):
        closure = e.PopValue(VClosure)
        e.Push(cls(maybe_simplify_surface_fn(closure, e)))
    return run


def _xform(make, nargs):
    def run(e):
        args = [e.PopValue(float) for _ in range(nargs)][::-1]
        obj = e.PopValue(SceneObject)
        e.Push(obj.transform(make(*args)))
    return run


def _deg(a):
    return a * math.pi / 180   # evaluator.go:1075 - (angle*pi)/180, left to right


def _union(e):
    a = e.PopValue(SceneObject)   # top of stack comes FIRST in Objects (evaluator.go:1091-1101)
    b = e.PopValue(SceneObject)
    e.Push(Union([a, b]))


def _difference(e):
    b = e.PopValue(SceneObject)
    a = e.PopValue(SceneObject)
    e.Push(Difference(a, b))


def _intersect(e):
    b = e.PopValue(SceneObject)
    a = e.PopValue(SceneObject)
    e.Push(Intersect(a, b))


def _pointlight(e):
    color = e.PopValue(Vec3)
    pos = e.PopValue(Vec3)
    e.Push(PointLight(pos, color))


def _light(e):
    color = e.PopValue(Vec3)
    direction = e.PopValue(Vec3)
    e.Push(DirectionalLight(direction, color))


def _spotlight(e):
    exp = e.PopValue(float)
    cutoff = e.PopValue(float)
    color = e.PopValue(Vec3)
    at = e.PopValue(Vec3)
    pos = e.PopValue(Vec3)
    e.Push(SpotLight(pos, at, color, cutoff, exp))


_LIGHT_TYPES = (PointLight, DirectionalLight, SpotLight)


def _pop_render_args(e):
    """evaluator.go:1113-1166: amb lights obj depth fov wid ht file render."""
    file = e.PopValue(str)
    height = e.PopValue(int)
    width = e.PopValue(int)
    fov = e.PopValue(float)
    depth = e.PopValue(int)
    obj = e.PopValue(SceneObject)
    lights = e.PopValue(VArray)
    amb = e.PopValue(Vec3)
    for l in lights.elements:
        if type(l) not in _LIGHT_TYPES:
            raise GMLError(f"expected lights array to contain *PointLight, got {_tname(type(l))}")
    return RenderArgs(Width=width, Height=height, File=file, Fov=fov, Depth=depth, Scene=obj, AmbientLight=amb,
                      Lights=list(lights.elements))


def _render(e):
    args = _pop_render_args(e)
    if e.Render is None:
        raise GMLError("render function not set")
    e.Render(e, args)


def _render_bg(e):
    bg_end = e.PopValue(Vec3)
    bg_start = e.PopValue(Vec3)
    args = _pop_render_args(e)
    args.BgColorStart, args.BgColorEnd = bg_start, bg_end
    if e.Render is None:
        raise GMLError("render function not set")
    e.Render(e, args)


DEG_TO_RAD = math.pi / 180.0

BUILTINS = {
    "addf": _bin(float, lambda e, a, b: a + b),
    "addi": _bin(int, lambda e, a, b: _wrap64(a + b)),
    "apply": _apply,
    "clampf": _un(float, _clampf),
    "cos": _un(float, lambda e, a: gomath.cos(DEG_TO_RAD * a)),
    "cube": _leaf(Cube),
    "cylinder": _leaf(Cylinder),
    "if": _if,
    "difference": _difference,
    "divi": _bin(int, _idiv),
    "divf": _bin(float, lambda e, a, b: _fdiv(a, b)),
    "eqi": _bin(int, lambda e, a, b: a == b),
    "eqf": _bin(float, lambda e, a, b: a == b),
    "floor": _un(float, _floor),
    "frac": _un(float, _frac),
    "get": _get,
    "getx": _un(Vec3, lambda e, p: p.x),
    "gety": _un(Vec3, lambda e, p: p.y),
    "getz": _un(Vec3, lambda e, p: p.z),
    "length": _un(VArray, lambda e, a: len(a.elements)),
    "lessi": _bin(int, lambda e, a, b: a < b),
    "lessf": _bin(float, lambda e, a, b: a < b),
    "material": _material,
    "modi": _bin(int, _imod),
    "muli": _bin(int, lambda e, a, b: _wrap64(a * b)),
    "mulf": _bin(float, lambda e, a, b: a * b),
    "negi": _un(int, lambda e, a: _wrap64(-a)),
    "negf": _un(float, lambda e, a: -a),
    "plane": _leaf(Plane),
    "point": _point,
    "pointlight": _pointlight,
    "render": _render,
    "renderWithBgGradient": _render_bg,
    "rotatex": _xform(lambda a: prim.rotate_x(_deg(a)), 1),
    "rotatey": _xform(lambda a: prim.rotate_y(_deg(a)), 1),
    "rotatez": _xform(lambda a: prim.rotate_z(_deg(a)), 1),
    "scale": _xform(prim.scale, 3),
    "sin": _un(float, lambda e, a: gomath.sin(DEG_TO_RAD * a)),
    "sphere": _leaf(Sphere),
    "sqrt": _un(float, _sqrt),
    "subi": _bin(int, lambda e, a, b: _wrap64(a - b)),
    "subf": _bin(float, lambda e, a, b: a - b),
    "translate": _xform(prim.translate, 3),
    "union": _union,
    "uscale": _xform(lambda s: prim.scale(s, s, s), 1),
}

# The reference's own set (evaluator.go:557-601); these shadow variables, as there.
REFERENCE_BUILTINS = frozenset(BUILTINS)

# EXTENSIONS (ICFP-2000 operators the north_star names; absent from evaluator.go:557-601).
# Looked up only when the identifier is neither a reference builtin nor bound.
EXTENSION_BUILTINS = {}
EXTENSION_BUILTINS.update({
    "cone": _leaf(Cone),
    "intersect": _intersect,
    "light": _light,
    "spotlight": _spotlight,
    "real": _un(int, lambda e, a: float(a)),
    "acos": _un(float, lambda e, a: gomath.acos(a) / DEG_TO_RAD),
    "asin": _un(float, lambda e, a: gomath.asin(a) / DEG_TO_RAD),
})
