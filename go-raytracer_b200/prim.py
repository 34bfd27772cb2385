"""Vec3 / Mat4 value math used on the host side of the `render` boundary.

Mirror of the parts of /root/reference/internal/prim/vec.go that the path uses
(Vec3 :9-116, Mat4 :256-425) and of plane.go:14-38 (unit-cube face table).  Every
expression keeps the reference's operand order so that Python's binary64
arithmetic (no FMA contraction) gives bit-identical matrices; the device never
re-derives these, it receives them in the scene blob.
"""
import math


class Vec3:
    __slots__ = ("x", "y", "z")

    def __init__(self, x=0.0, y=0.0, z=0.0):
        self.x, self.y, self.z = float(x), float(y), float(z)

    def __iter__(self):
        return iter((self.x, self.y, self.z))

    def __eq__(self, o):
        return isinstance(o, Vec3) and (self.x, self.y, self.z) == (o.x, o.y, o.z)

    def __hash__(self):
        return hash((self.x, self.y, self.z))

    def is_zero(self):
        return self.x == 0.0 and self.y == 0.0 and self.z == 0.0

    def dot(self, o):
        return self.x * o.x + self.y * o.y + self.z * o.z

    def normalize(self):
        m = math.sqrt(self.x * self.x + self.y * self.y + self.z * self.z)
        return Vec3(_div(self.x, m), _div(self.y, m), _div(self.z, m))

    def __str__(self):  # vec.go:13-15, fmt %v of float64
        return f"[{go_v(self.x)}, {go_v(self.y)}, {go_v(self.z)}]"

    __repr__ = __str__


def _div(a, b):
    try:
        return a / b
    except ZeroDivisionError:
        if a != a or a == 0.0:
            return math.nan
        return math.copysign(math.inf, a) * math.copysign(1.0, b)


def go_v(f):
    """fmt's %v for a float64: %g with the shortest representation, exponent form
    only below 1e-4 or from 1e21 up."""
    if f != f:
        return "NaN"
    if f == math.inf:
        return "+Inf"
    if f == -math.inf:
        return "-Inf"
    if f == 0:
        return "-0" if math.copysign(1.0, f) < 0 else "0"
    a = abs(f)
    if 1e-4 <= a < 1e21:
        s = repr(a)
        if "e" in s:   # repr switches to exponent form at 1e16
            s = f"{a:.0f}" if a >= 1e16 else f"{a:.20f}".rstrip("0")
        if s.endswith(".0"):
            s = s[:-2]
        return ("-" if f < 0 else "") + s
    mant, _, exp = repr(a).partition("e")
    if not exp:
        mant, exp = f"{a:e}".split("e")
        mant = mant.rstrip("0").rstrip(".")
    if mant.endswith(".0"):
        mant = mant[:-2]
    e = int(exp)
    return f"{'-' if f < 0 else ''}{mant}e{'+' if e >= 0 else '-'}{abs(e):02d}"


# ---------------------------------------------------------------- Mat4 (row major 4x4)

def identity():
    return ((1.0, 0.0, 0.0, 0.0), (0.0, 1.0, 0.0, 0.0), (0.0, 0.0, 1.0, 0.0), (0.0, 0.0, 0.0, 1.0))


def mul_mat(m, n):
    """vec.go:258-268 - accumulate from 0.0 over k, in that order."""
    out = []
    for i in range(4):
        row = []
        for j in range(4):
            acc = 0.0
            for k in range(4):
                acc += m[i][k] * n[k][j]
            row.append(acc)
        out.append(tuple(row))
    return tuple(out)


def transpose(m):
    return tuple(tuple(m[j][i] for j in range(4)) for i in range(4))


def mul_point(m, v):
    return Vec3(m[0][0] * v.x + m[0][1] * v.y + m[0][2] * v.z + m[0][3],
                m[1][0] * v.x + m[1][1] * v.y + m[1][2] * v.z + m[1][3],
                m[2][0] * v.x + m[2][1] * v.y + m[2][2] * v.z + m[2][3])


def mul_dir(m, v):
    return Vec3(m[0][0] * v.x + m[0][1] * v.y + m[0][2] * v.z,
                m[1][0] * v.x + m[1][1] * v.y + m[1][2] * v.z,
                m[2][0] * v.x + m[2][1] * v.y + m[2][2] * v.z)


def inverse(m):
    """Affine inverse, vec.go:319-365: cofactors / det for the 3x3 block, then
    -(L^-1 T) for the last column.  Returns None when det == 0 (as the reference)."""
    a, b, c = m[0][0], m[0][1], m[0][2]
    d, e, f = m[1][0], m[1][1], m[1][2]
    g, h, i = m[2][0], m[2][1], m[2][2]
    det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g)
    if det == 0.0:
        return None
    r0 = [_div(e * i - f * h, det), _div(c * h - b * i, det), _div(b * f - c * e, det), 0.0]
    r1 = [_div(f * g - d * i, det), _div(a * i - c * g, det), _div(c * d - a * f, det), 0.0]
    r2 = [_div(d * h - e * g, det), _div(b * g - a * h, det), _div(a * e - b * d, det), 0.0]
    r0[3] = -(r0[0] * m[0][3] + r0[1] * m[1][3] + r0[2] * m[2][3])
    r1[3] = -(r1[0] * m[0][3] + r1[1] * m[1][3] + r1[2] * m[2][3])
    r2[3] = -(r2[0] * m[0][3] + r2[1] * m[1][3] + r2[2] * m[2][3])
    return (tuple(r0), tuple(r1), tuple(r2), (0.0, 0.0, 0.0, 1.0))


def translate(x, y, z):
    return ((1.0, 0.0, 0.0, x), (0.0, 1.0, 0.0, y), (0.0, 0.0, 1.0, z), (0.0, 0.0, 0.0, 1.0))


def scale(x, y, z):
    return ((x, 0.0, 0.0, 0.0), (0.0, y, 0.0, 0.0), (0.0, 0.0, z, 0.0), (0.0, 0.0, 0.0, 1.0))


def _cs(a):
    from .gml import gomath   # late import: gml.evaluator imports this module
    return gomath.cos(a), gomath.sin(a)


def rotate_x(a):
    c, s = _cs(a)
    return ((1.0, 0.0, 0.0, 0.0), (0.0, c, -s, 0.0), (0.0, s, c, 0.0), (0.0, 0.0, 0.0, 1.0))


def rotate_y(a):
    c, s = _cs(a)
    return ((c, 0.0, s, 0.0), (0.0, 1.0, 0.0, 0.0), (-s, 0.0, c, 0.0), (0.0, 0.0, 0.0, 1.0))


def rotate_z(a):
    c, s = _cs(a)
    return ((c, -s, 0.0, 0.0), (s, c, 0.0, 0.0), (0.0, 0.0, 1.0, 0.0), (0.0, 0.0, 0.0, 1.0))


# plane.go:14-38.  Index = face id handed to surface functions.
CUBE_FRONT, CUBE_BACK, CUBE_LEFT, CUBE_RIGHT, CUBE_TOP, CUBE_BOTTOM = range(6)
UNIT_CUBE_PLANES = (
    (Vec3(0, 0, 0), Vec3(0, 0, -1)),
    (Vec3(0, 0, 1), Vec3(0, 0, +1)),
    (Vec3(0, 0, 0), Vec3(-1, 0, 0)),
    (Vec3(1, 0, 0), Vec3(+1, 0, 0)),
    (Vec3(0, 1, 0), Vec3(0, +1, 0)),
    (Vec3(0, 0, 0), Vec3(0, -1, 0)),
)


def SSIM(img1, img2, ctx=None):
    """prim.SSIM (internal/prim/ssim.go:27-144): structural similarity of two uint8 [H,W,4] rasters, computed on the
    GPU through the C ABI (gmlrt_ssim).  `ctx`: an existing cuda.Context, else one is made on device 0.  Raises
    ValueError like the reference's two error returns (different sizes :31-33, smaller than the window :34-36)."""
    import numpy as np
    from . import cuda
    a, b = np.asarray(img1), np.asarray(img2)
    if a.shape != b.shape:
        raise ValueError("images are not the same size")
    if a.shape[0] < 11 or a.shape[1] < 11:
        raise ValueError("images are too small")
    own = ctx is None
    if own:
        ctx = cuda.Context((0,), cuda.PRECISION_F32)
    try:
        return ctx.ssim(a, b)
    finally:
        if own:
            ctx.close()
