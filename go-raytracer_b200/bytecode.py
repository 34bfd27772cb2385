"""Compile a GML surface closure to gmlrt register bytecode (include/gmlrt.h).

What it replaces: per hit the reference *interprets* the closure - push face,u,v,
clone the environment map twice, walk the token list, pop a Material or
``color kd ks n`` (/root/reference/internal/gml/evaluator.go:672-727, :433-438,
:365-414, builtins :853-1019).  Here the closure is partially evaluated once, on the
host: the compiler runs the same token list over an *abstract* stack whose entries
are either compile-time constants (numbers, points, arrays, closures from the
captured environment) or virtual registers that will hold run-time values.
An ``if`` on a run-time condition becomes a forward jump with both arms merged into
common registers.  A closure that is only known at run time - the two arms of such an
``if`` leave different closures, or ``get`` picks one out of an array of closures with a
run-time index - is a *choice*: a selector register plus the candidate closures, and
``apply`` / ``if`` on it unfold into a chain of branches, one inlined body per candidate
(the reference interprets whichever closure it finds on the stack,
internal/gml/evaluator.go:604-610, :1005-1019).  Recursion is unrolled; a chain of
applications deeper than MAX_INLINE_DEPTH becomes a run-time fault (the device has no
call stack: bounded depth -> GMLRT_E_SURFACE).  What is left for the
device is typed three-address code with no closures, no environment and no stack.

Constant folding uses the evaluator's own builtin implementations, so a folded
value is bit-identical to what the reference would compute at run time.
"""
import struct

from .gml import evaluator as ev
from .gml.parser import Array, Binder, BoolLiteral, FloatLiteral, Function, Identifier, IntLiteral, StringLiteral
from .prim import Vec3

MAX_REGS = 128
MAX_INLINE_DEPTH = 64
MAX_CODE_WORDS = 1 << 16
MERGE_DEBT = 3                    # ... and what every open branch is assumed to need for its merge
RECURSION_RESERVE = 24            # registers an unrolled recursion leaves to the rest of the surface function

OP = dict(RET=0, LDK=1, LDI=2, MOV=3, ADDF=4, SUBF=5, MULF=6, DIVF=7, NEGF=8, ADDI=9, SUBI=10, MULI=11, DIVI=12,
          MODI=13, NEGI=14, EQF=15, LESSF=16, EQI=17, LESSI=18, SIN=19, COS=20, SQRT=21, CLAMPF=22, FLOOR=23, FRAC=24,
          REAL=25, ACOS=26, ASIN=27, LDARR=28, LEN=29, GETF=30, GETP=31, GETI=32, GETA=33, JMP=34, JMPF=35, ERR=36)
OP_NAMES = {v: k for k, v in OP.items()}


class SurfaceCompileError(Exception):
    """The closure uses something the static compiler cannot resolve."""


class _OutOfRegisters(SurfaceCompileError):
    pass


class A:
    """Abstract value.  kind: i f b (scalars) | p (point: comps) | mat (10 comps) |
    arr (items: list[A]) | clos (body, env) | other (opaque compile-time value)."""
    __slots__ = ("kind", "const", "reg", "comps", "items", "body", "env", "pooled", "exhaustive")

    def __init__(self, kind, const=None, reg=None, comps=None, items=None, body=None, env=None):
        self.kind, self.const, self.reg, self.comps, self.items = kind, const, reg, comps, items
        self.body, self.env, self.pooled = body, env, None
        self.exhaustive = True       # choice: the selector is known to be one of 0 .. len(items) - 1

    @property
    def is_const(self):
        if self.kind in "ifb":
            return self.reg is None
        if self.kind in ("p", "mat"):
            return all(c.is_const for c in self.comps)
        if self.kind == "arr":
            return self.reg is None
        if self.kind == "choice":
            return False
        return True


def lift(v):
    """Python evaluator value -> abstract constant."""
    t = type(v)
    if t is bool:
        return A("b", const=v)
    if t is int:
        return A("i", const=v)
    if t is float:
        return A("f", const=v)
    if t is Vec3:
        return A("p", comps=[A("f", const=v.x), A("f", const=v.y), A("f", const=v.z)])
    if t is ev.VArray:
        return A("arr", items=[lift(x) for x in v.elements])
    if t is ev.VClosure:
        return A("clos", body=v.code, env=_LazyEnv(v.env))
    if t is ev.Material:
        return A("mat", comps=[A("f", const=x) for x in v.as_tuple()])
    return A("other", const=v)


class _LazyEnv(dict):
    """Captured environment of a host closure; entries are lifted on first use."""

    def __init__(self, src):
        super().__init__()
        self._src = src

    def lookup(self, i):
        if i in self:
            return self[i]
        if self._src is not None and i in self._src:
            a = lift(self._src[i])
            self[i] = a
            return a
        return None

    def fork(self):
        e = _LazyEnv(self._src)
        e.update(self)
        return e


class Pools:
    """Constant pools shared by all surfaces of one blob."""

    def __init__(self):
        self.fpool, self.itab, self.code = [], [], []
        self._fidx = {}

    def fconst(self, x):
        key = struct.pack("<d", x)
        i = self._fidx.get(key)
        if i is None:
            i = len(self.fpool)
            if i >= 65536:
                raise SurfaceCompileError("constant pool overflow (65536 reals)")
            self.fpool.append(x)
            self._fidx[key] = i
        return i


_F2 = {"addf": "ADDF", "subf": "SUBF", "mulf": "MULF", "divf": "DIVF"}
_I2 = {"addi": "ADDI", "subi": "SUBI", "muli": "MULI", "divi": "DIVI", "modi": "MODI"}
_CMPF = {"eqf": "EQF", "lessf": "LESSF"}
_CMPI = {"eqi": "EQI", "lessi": "LESSI"}
_F1 = {"negf": "NEGF", "sin": "SIN", "cos": "COS", "sqrt": "SQRT", "clampf": "CLAMPF", "frac": "FRAC", "acos": "ACOS",
       "asin": "ASIN"}


class _Compiler:
    def __init__(self, pools):
        self.pools = pools
        self.code = []
        self.nreg = 3
        self.depth = 0
        self.emitted = 0
        self.active = []             # bodies of the closures being inlined, outermost first
        self.branch_depth = 0        # run-time branches whose arms are being compiled
        self.reg_limit = MAX_REGS    # lowered while a recursion is being unrolled, so that what follows it still finds registers
        self._folder = ev.EvalState()

    # ---- code emission
    def emit(self, op, d=0, a=0, b=0, *imm):
        self.emitted += 1 + len(imm)
        if self.emitted > MAX_CODE_WORDS:
            raise SurfaceCompileError(f"surface function unrolls to more than {MAX_CODE_WORDS} words of code")
        self.code.append(OP[op] | (d << 8) | (a << 16) | (b << 24))
        self.code.extend(x & 0xFFFFFFFF for x in imm)

    def new_reg(self, n=1):
        r = self.nreg
        self.nreg += n
        # while a recursion is being unrolled every branch still open will want registers to merge its arms when it closes
        if self.nreg + (MERGE_DEBT * self.branch_depth if self.reg_limit < MAX_REGS else 0) > self.reg_limit:
            raise _OutOfRegisters(f"surface function needs more than {MAX_REGS} registers")
        return r

    def reg_of(self, a):
        """Materialise a scalar into a register."""
        if a.reg is not None:
            return a.reg
        r = self.new_reg()
        if a.kind == "f":
            k = self.pools.fconst(a.const)
            self.emit("LDK", r, k & 0xFF, k >> 8)
        elif a.kind in "ib":
            v = int(a.const)
            if not (-(1 << 31) <= v < (1 << 31)):
                raise SurfaceCompileError(f"integer constant {v} does not fit the device's 32-bit ints")
            self.emit("LDI", r, 0, 0, v)
        else:
            raise SurfaceCompileError(f"cannot materialise a {a.kind}")
        return r

    # ---- arrays
    def elem_type(self, a):
        if a.kind in "ifb":
            return a.kind
        if a.kind == "p":
            return "p"
        if a.kind == "arr":
            ts = {self.elem_type(x) for x in a.items}
            if len(ts) > 1:
                raise SurfaceCompileError("run-time indexing needs arrays with elements of one type")
            return ("arr", ts.pop() if ts else "f")
        raise SurfaceCompileError(f"arrays of {a.kind} cannot be indexed at run time")

    def pool_array(self, a):
        """Place a constant array in the pools; returns (offset, length, elem type)."""
        if a.pooled is not None:
            return a.pooled
        if a.reg is not None:
            raise SurfaceCompileError("internal: pooling a run-time array")
        for x in a.items:
            if not x.is_const:
                raise SurfaceCompileError("run-time indexing of an array built from run-time values")
        et = self.elem_type(a)[1] if a.items else "f"
        P = self.pools
        if et == "f":
            off = len(P.fpool)
            P.fpool.extend(float(x.const) for x in a.items)
        elif et == "p":
            off = len(P.fpool)
            for x in a.items:
                P.fpool.extend(float(c.const) for c in x.comps)
        elif et in ("i", "b"):
            off = len(P.itab)
            for x in a.items:
                v = int(x.const)
                if not (-(1 << 31) <= v < (1 << 31)):
                    raise SurfaceCompileError(f"integer constant {v} does not fit the device's 32-bit ints")
                P.itab.append(v)
        else:  # array of arrays: descriptors
            subs = [self.pool_array(x) for x in a.items]
            off = len(P.itab)
            for (o, n, _) in subs:
                P.itab.extend((o, n))
        a.pooled = (off, len(a.items), et)
        return a.pooled

    def arr_regs(self, a):
        """Run-time (offset,length) register pair + element type for an array value."""
        if a.reg is not None:
            return a.reg, a.const  # const slot carries the elem type for run-time arrays
        off, n, et = self.pool_array(a)
        r = self.new_reg(2)
        self.emit("LDARR", r, 0, 0, off, n)
        return r, et

    # ---- folding through the reference evaluator's own builtins
    def fold(self, name, *args):
        st = self._folder
        st.Stack = list(args)
        st.CurrToken = None
        ev.BUILTINS.get(name, ev.EXTENSION_BUILTINS.get(name))(st)
        return st.Stack[-1]

    # ---- the abstract interpreter
    def run(self, code, env, stack):
        self.depth += 1
        if self.depth > MAX_INLINE_DEPTH:
            # the device has no call stack: applications nested deeper than this fault at run time (GMLRT_E_SURFACE)
            self.depth -= 1
            self.emit("ERR")
            raise _DeadCode()
        for tok in code:
            t = type(tok)
            if t is IntLiteral:
                stack.append(A("i", const=tok.value))
            elif t is FloatLiteral:
                stack.append(A("f", const=tok.value))
            elif t is BoolLiteral:
                stack.append(A("b", const=tok.value))
            elif t is StringLiteral:
                stack.append(A("other", const=tok.value))
            elif t is Function:
                stack.append(A("clos", body=tok.body, env=env.fork()))
            elif t is Binder:
                env[tok.id] = self.pop(stack)
            elif t is Array:
                sub = []
                self.run(tok.elements, env, sub)
                stack.append(A("arr", items=sub))
            elif t is Identifier:
                if tok.name in ev.BUILTINS:
                    self.builtin(tok.name, env, stack)
                else:
                    v = env.lookup(tok.id)
                    if v is not None:
                        stack.append(v)
                    elif tok.name in ev.EXTENSION_BUILTINS:
                        self.builtin(tok.name, env, stack)
                    else:
                        raise SurfaceCompileError(f"unbound identifier: {tok.name}")
            else:
                raise SurfaceCompileError(f"unknown token {tok!r}")
        self.depth -= 1

    @staticmethod
    def pop(stack, kind=None):
        if not stack:
            raise SurfaceCompileError("surface function pops more than face, u, v")
        a = stack.pop()
        if kind is not None and a.kind != kind:
            raise SurfaceCompileError(f"type mismatch: expected {kind}, got {a.kind}")
        return a

    def scalar_op(self, op, kind_out, *args):
        r = self.new_reg()
        regs = [self.reg_of(a) for a in args] + [0]
        self.emit(op, r, regs[0], regs[1])
        return A(kind_out, reg=r)

    def builtin(self, name, env, stack):
        pop = self.pop
        if name in _F2 or name in _CMPF:
            b, a = pop(stack, "f"), pop(stack, "f")
            if a.is_const and b.is_const:
                stack.append(lift(self.fold(name, a.const, b.const)))
            else:
                stack.append(self.scalar_op(_F2.get(name) or _CMPF[name], "f" if name in _F2 else "b", a, b))
        elif name in _I2 or name in _CMPI:
            b, a = pop(stack, "i"), pop(stack, "i")
            if a.is_const and b.is_const and not (name in ("divi", "modi") and b.const == 0):
                stack.append(lift(self.fold(name, a.const, b.const)))
            else:
                stack.append(self.scalar_op(_I2.get(name) or _CMPI[name], "i" if name in _I2 else "b", a, b))
        elif name in _F1:
            a = pop(stack, "f")
            stack.append(lift(self.fold(name, a.const)) if a.is_const else self.scalar_op(_F1[name], "f", a))
        elif name == "negi":
            a = pop(stack, "i")
            stack.append(lift(self.fold(name, a.const)) if a.is_const else self.scalar_op("NEGI", "i", a))
        elif name == "floor":
            a = pop(stack, "f")
            stack.append(lift(self.fold(name, a.const)) if a.is_const else self.scalar_op("FLOOR", "i", a))
        elif name == "real":
            a = pop(stack, "i")
            stack.append(lift(self.fold(name, a.const)) if a.is_const else self.scalar_op("REAL", "f", a))
        elif name == "point":
            z, y, x = pop(stack, "f"), pop(stack, "f"), pop(stack, "f")
            stack.append(A("p", comps=[x, y, z]))
        elif name in ("getx", "gety", "getz"):
            p = pop(stack, "p")
            stack.append(p.comps["xyz".index(name[3])])
        elif name == "material":
            f = [pop(stack, "f") for _ in range(7)][::-1]
            c = pop(stack, "p")
            stack.append(A("mat", comps=list(c.comps) + f))
        elif name == "length":
            a = pop(stack, "arr")
            if a.reg is None:
                stack.append(A("i", const=len(a.items)))
            else:
                r = self.new_reg()
                self.emit("LEN", r, a.reg, 0)
                stack.append(A("i", reg=r))
        elif name == "get":
            self.get(stack)
        elif name == "apply":
            self.apply_value(self.pop_callable(stack), stack)
        elif name == "if":
            self.if_(stack)
        else:
            raise SurfaceCompileError(f"`{name}` cannot be used inside a surface function")

    def get(self, stack):
        i, a = self.pop(stack, "i"), self.pop(stack, "arr")
        if a.reg is None and i.is_const:
            n = len(a.items)
            if i.const < 0 or i.const >= n:
                self.emit("ERR")     # the reference fails here at render time (evaluator.go:990-995)
                raise _DeadCode()
            stack.append(a.items[i.const])
            return
        if a.reg is None and a.items and all(x.kind in ("clos", "choice") for x in a.items):
            # a closure picked at run time: bounds-checked here, like the reference's get (evaluator.go:989-995), and
            # carried as a choice until something applies it
            ir = self.reg_of(i)
            below = self.scalar_op("LESSI", "b", A("i", reg=ir), A("i", const=len(a.items)))
            self.branch(below.reg, lambda st: None, self.fault, stack)
            negative = self.scalar_op("LESSI", "b", A("i", reg=ir), A("i", const=0))
            self.branch(negative.reg, self.fault, lambda st: None, stack)
            stack.append(A("choice", reg=ir, items=list(a.items)))
            return
        ar, et = self.arr_regs(a)
        ir = self.reg_of(i)
        if et == "f":
            r = self.new_reg()
            self.emit("GETF", r, ar, ir)
            stack.append(A("f", reg=r))
        elif et in ("i", "b"):
            r = self.new_reg()
            self.emit("GETI", r, ar, ir)
            stack.append(A(et, reg=r))
        elif et == "p":
            r = self.new_reg(3)
            self.emit("GETP", r, ar, ir)
            stack.append(A("p", comps=[A("f", reg=r), A("f", reg=r + 1), A("f", reg=r + 2)]))
        else:
            r = self.new_reg(2)
            self.emit("GETA", r, ar, ir)
            out = A("arr", reg=r)
            out.const = et[1]   # element type of the fetched inner array
            stack.append(out)

    def pop_callable(self, stack):
        a = self.pop(stack)
        if a.kind not in ("clos", "choice"):
            raise SurfaceCompileError(f"type mismatch: expected a closure, got {a.kind}")
        return a

    def fault(self, stack):
        self.emit("ERR")
        raise _DeadCode()

    def apply_value(self, v, stack):
        """Inline the application of a closure value: a closure known at compile time, or a run-time choice among some."""
        if v.kind == "clos":
            recursive = any(b is v.body for b in self.active)
            n_active, depth0, nreg0, code0, limit0 = len(self.active), self.depth, self.nreg, self.code, self.reg_limit
            self.active.append(v.body)
            if recursive:
                self.reg_limit = min(self.reg_limit, MAX_REGS - RECURSION_RESERVE)
            try:
                try:
                    self.run(v.body, v.env.fork(), stack)
                except _OutOfRegisters:
                    if not recursive:
                        raise
                    # an unrolled recursion has used up the register file: from this depth on it is a run-time fault, like
                    # the depth bound (registers are not reused across the levels of an unrolling)
                    self.depth, self.nreg, self.code = depth0, nreg0, code0      # (branches inside the abandoned level had swapped the code buffer)
                    self.fault(stack)
            finally:
                del self.active[n_active:]
                self.reg_limit = limit0
        else:
            self.apply_chain(v, 0, stack)

    def apply_chain(self, v, k, stack):
        n = len(v.items)
        if k == n:                                   # (a selector outside 0 .. n-1: get has already faulted)
            self.fault(stack)
        if k == n - 1 and v.exhaustive:
            self.apply_value(v.items[k], stack)
            return
        hit = self.scalar_op("EQI", "b", A("i", reg=v.reg), A("i", const=k))
        self.branch(hit.reg, lambda st: self.apply_value(v.items[k], st), lambda st: self.apply_chain(v, k + 1, st), stack)

    def if_(self, stack):
        f, t, c = self.pop_callable(stack), self.pop_callable(stack), self.pop(stack, "b")
        if c.is_const:
            self.apply_value(t if c.const else f, stack)
            return
        self.branch(self.reg_of(c), lambda st: self.apply_value(t, st), lambda st: self.apply_value(f, st), stack)

    def branch(self, cr, run_t, run_f, stack):
        """Run-time two-way branch on register `cr`: each arm is compiled over a copy of the stack, then the stacks are merged."""
        outer, depth0, bd0 = self.code, self.depth, self.branch_depth
        arms = []
        self.branch_depth += 1
        try:
            for arm in (run_t, run_f):
                self.code = []
                st, dead = list(stack), False
                try:
                    arm(st)
                except _DeadCode:
                    dead, self.depth = True, depth0
                arms.append((self.code, st, dead))
        finally:
            self.branch_depth = bd0
        self.code = outer
        (code_t, st_t, dead_t), (code_f, st_f, dead_f) = arms
        if dead_t and dead_f:
            self.emit("JMPF", 0, cr, 0, len(code_t))
            self.code.extend(code_t + code_f)
            raise _DeadCode()
        if dead_t or dead_f:
            # the faulting arm never rejoins: no merge needed
            if dead_t:
                self.emit("JMPF", 0, cr, 0, len(code_t))
                self.code.extend(code_t + code_f)
                stack[:] = st_f
            else:
                self.emit("JMPF", 0, cr, 0, len(code_t) + 2)
                self.code.extend(code_t)
                self.emit("JMP", 0, 0, 0, len(code_f))
                self.code.extend(code_f)
                stack[:] = st_t
            return
        if len(st_t) != len(st_f):
            raise SurfaceCompileError("the two arms of `if` leave stacks of different height")
        mv_t, mv_f = [], []
        merged = [self.merge(x, y, mv_t, mv_f) for x, y in zip(st_t, st_f)]
        code_t = code_t + self.flush_moves(mv_t)
        code_f = code_f + self.flush_moves(mv_f)
        # jumps are relative to the word that follows their immediate
        self.emit("JMPF", 0, cr, 0, len(code_t) + 2)
        self.code.extend(code_t)
        self.emit("JMP", 0, 0, 0, len(code_f))
        self.code.extend(code_f)
        stack[:] = merged

    def flush_moves(self, moves):
        saved, self.code = self.code, []
        for dst, src in moves:
            if isinstance(src, A):
                # constant: load straight into the merge register
                if src.kind == "f":
                    k = self.pools.fconst(src.const)
                    self.emit("LDK", dst, k & 0xFF, k >> 8)
                else:
                    self.emit("LDI", dst, 0, 0, int(src.const))
            else:
                self.emit("MOV", dst, src, 0)
        out, self.code = self.code, saved
        return out

    def merge(self, x, y, mv_t, mv_f):
        if x is y:
            return x
        if x.kind != y.kind and not {x.kind, y.kind} <= {"clos", "choice"}:
            raise SurfaceCompileError(f"the two arms of `if` leave different types ({x.kind} vs {y.kind})")
        k = x.kind if x.kind == y.kind else "choice"
        if k in "ifb":
            if x.is_const and y.is_const and x.const == y.const and (k != "f" or struct.pack("<d", x.const) == struct.pack("<d", y.const)):
                return x
            r = self.new_reg()
            mv_t.append((r, x if x.is_const else x.reg))
            mv_f.append((r, y if y.is_const else y.reg))
            return A(k, reg=r)
        if k in ("p", "mat"):
            return A(k, comps=[self.merge(a, b, mv_t, mv_f) for a, b in zip(x.comps, y.comps)])
        if k == "arr":
            raise SurfaceCompileError("`if` arms that yield different arrays are not supported")
        if k == "clos" and x.body is y.body and x.env == y.env:
            return x
        if k in ("clos", "choice"):
            # which closure it is will only be known at run time: a selector register, 0 = this arm's, 1 = the other's
            r = self.new_reg()
            mv_t.append((r, A("i", const=0)))
            mv_f.append((r, A("i", const=1)))
            return A("choice", reg=r, items=[x, y])
        if x.const == y.const:
            return x
        raise SurfaceCompileError("`if` arms yield different opaque values")


class _DeadCode(Exception):
    """Raised after emitting ERR: the rest of this arm can never run."""


def compile_surface(closure, pools):
    """Returns ("const", material_tuple) or ("code", code_off, code_len, n_regs, known) where
    known[k] is material component k when it is a compile-time constant, else NaN."""
    c = _Compiler(pools)
    stack = [A("i", reg=0), A("f", reg=1), A("f", reg=2)]
    known = (float("nan"),) * 10
    try:
        c.run(closure.code, _LazyEnv(closure.env), stack)
        if not stack:
            raise SurfaceCompileError("surface function leaves an empty stack")
        top = stack.pop()
        if top.kind == "mat":
            comps = top.comps
        else:
            if top.kind != "f":
                raise SurfaceCompileError(f"surface function must leave a Material or `color kd ks n`, got {top.kind}")
            n = top
            ks, kd = c.pop(stack, "f"), c.pop(stack, "f")
            color = c.pop(stack, "p")
            zero = A("f", const=0.0)
            # evaluator.go:719-725: Reflectivity = ks; fuzz, transparency, ior zero
            comps = list(color.comps) + [ks, zero, zero, zero, kd, ks, n]
        if all(x.is_const for x in comps):
            return ("const", tuple(float(x.const) for x in comps))
        known = tuple(float(x.const) if x.is_const else float("nan") for x in comps)
        base = c.new_reg(10)
        c.code.extend(c.flush_moves([(base + i, x if x.is_const else x.reg) for i, x in enumerate(comps)]))
        c.emit("RET", 0, base, 0)
    except _DeadCode:
        pass  # every path faults; the code already ends in ERR
    off = len(pools.code)
    pools.code.extend(c.code)
    return ("code", off, len(c.code), c.nreg, known)


def disassemble(code):
    out, i = [], 0
    while i < len(code):
        w = code[i]
        op, d, a, b = w & 0xFF, (w >> 8) & 0xFF, (w >> 16) & 0xFF, w >> 24
        name = OP_NAMES.get(op, f"?{op}")
        if name == "LDI":
            out.append(f"{i:4d} {name} r{d} #{code[i + 1] - (1 << 32) if code[i + 1] >> 31 else code[i + 1]}")
            i += 2
        elif name in ("JMP", "JMPF"):
            out.append(f"{i:4d} {name} r{a} -> {i + 2 + code[i + 1]}")
            i += 2
        elif name == "LDARR":
            out.append(f"{i:4d} {name} r{d} off={code[i + 1]} len={code[i + 2]}")
            i += 3
        elif name == "LDK":
            out.append(f"{i:4d} {name} r{d} k{a | (b << 8)}")
            i += 1
        else:
            out.append(f"{i:4d} {name} r{d} r{a} r{b}")
            i += 1
    return "\n".join(out)
