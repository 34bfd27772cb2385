"""Go standard-library float64 arithmetic that the GML front end depends on.

The reference evaluates rotations and the GML ``sin``/``cos`` builtins with Go's
pure-Go (Cephes-derived) ``math.Sin`` / ``math.Cos`` (call sites:
/root/reference/internal/prim/vec.go:394-425 and
/root/reference/internal/gml/evaluator.go:929-937).  Those sources are *not* in the
reference tree (Go stdlib, go 1.24.5 per go.mod:3), so the published algorithm is
restated here: 3-term Cody-Waite reduction by pi/4 followed by degree-11/12
minimax polynomials.  Python floats are IEEE binary64 and CPython never contracts
a*b+c into an FMA, so the results are bit-identical to a non-FMA Go build (amd64).

Pinning: the seven golden images exercise sin/cos at
45/135/270/-90/90/180 degrees through the transform matrices and tan at fov 90/120.
"""
import math
import struct

PI4A = 7.85398125648498535156e-1   # 0x3fe921fb40000000
PI4B = 3.77489470793079817668e-8   # 0x3e64442d00000000
PI4C = 2.69515142907905952645e-15  # 0x3ce8469898cc5170
M4PI = 1.2732395447351628  # 4/pi rounded to nearest binary64 (0x1.45f306dc9c883p+0), as Go folds the constant `4 / Pi`

_SIN = (
    1.58962301576546568060e-10,
    -2.50507477628578072866e-8,
    2.75573136213857245213e-6,
    -1.98412698295895385996e-4,
    8.33333333332211858878e-3,
    -1.66666666666666307295e-1,
)
_COS = (
    -1.13585365213876817300e-11,
    2.08757008419747316778e-9,
    -2.75573141792967388112e-7,
    2.48015872888517045348e-5,
    -1.38888888888730564116e-3,
    4.16666666666665929218e-2,
)
_TAN_P = (
    -1.30936939181383777646e4,
    1.15351664838587416140e6,
    -1.79565251976484877988e7,
)
_TAN_Q = (
    1.0,
    1.36812963470692954678e4,
    -1.32089234440210967447e6,
    2.50083801823357915839e7,
    -5.38695755929454629881e7,
)

# Go switches to Payne-Hanek reduction above this; GML angles never get there.
REDUCE_THRESHOLD = float(1 << 29)


def _reduce(x):
    """x >= 0 -> (octant j with odd octants rounded up, remainder z)."""
    if x >= REDUCE_THRESHOLD:
        # Out of the range any GML scene uses; fall back to libm-grade reduction.
        r = math.fmod(x, 2.0 * math.pi)
        j = int(r * M4PI)
        y = float(j)
        if j & 1:
            j += 1
            y += 1.0
        return j, r - y * (math.pi / 4.0)
    j = int(x * M4PI)
    y = float(j)
    if j & 1:
        j += 1
        y += 1.0
    z = ((x - y * PI4A) - y * PI4B) - y * PI4C
    return j, z


def _sin_poly(z, zz):
    s = _SIN
    return z + z * zz * ((((((s[0] * zz) + s[1]) * zz + s[2]) * zz + s[3]) * zz + s[4]) * zz + s[5])


def _cos_poly(zz):
    c = _COS
    return 1.0 - 0.5 * zz + zz * zz * ((((((c[0] * zz) + c[1]) * zz + c[2]) * zz + c[3]) * zz + c[4]) * zz + c[5])


def sin(x):
    if x == 0.0 or x != x:
        return x
    if math.isinf(x):
        return math.nan
    sign = False
    if x < 0.0:
        x = -x
        sign = True
    j, z = _reduce(x)
    j &= 7
    if j > 3:
        sign = not sign
        j -= 4
    zz = z * z
    y = _cos_poly(zz) if j in (1, 2) else _sin_poly(z, zz)
    return -y if sign else y


def cos(x):
    if x != x or math.isinf(x):
        return math.nan
    sign = False
    x = abs(x)
    j, z = _reduce(x)
    j &= 7
    if j > 3:
        j -= 4
        sign = not sign
    if j > 1:
        sign = not sign
    zz = z * z
    y = _sin_poly(z, zz) if j in (1, 2) else _cos_poly(zz)
    return -y if sign else y


def tan(x):
    if x == 0.0 or x != x:
        return x
    if math.isinf(x):
        return math.nan
    sign = False
    if x < 0.0:
        x = -x
        sign = True
    j, z = _reduce(x)
    zz = z * z
    if zz > 1e-14:
        p, q = _TAN_P, _TAN_Q
        y = z + z * (zz * (((p[0] * zz) + p[1]) * zz + p[2])
                     / ((((zz + q[1]) * zz + q[2]) * zz + q[3]) * zz + q[4]))
    else:
        y = z
    if j & 2:
        y = -1.0 / y
    return -y if sign else y


# --- atan / asin / acos (Cephes atan.c / asin.c as ported in Go's math package) ---
_ATAN_P = (
    -8.750608600031904122785e-01,
    -1.615753718733365076637e+01,
    -7.500855792314704667340e+01,
    -1.228866684490136173410e+02,
    -6.485021904942025371773e+01,
)
_ATAN_Q = (
    2.485846490142306297962e+01,
    1.650270098316988542046e+02,
    4.328810604912902668951e+02,
    4.853903996359136964868e+02,
    1.945506571482613964425e+02,
)
MOREBITS = 6.123233995736765886130e-17
TAN3PIO8 = 2.41421356237309504880


def _xatan(x):
    p, q = _ATAN_P, _ATAN_Q
    z = x * x
    z = z * ((((p[0] * z + p[1]) * z + p[2]) * z + p[3]) * z + p[4]) / (((((z + q[0]) * z + q[1]) * z + q[2]) * z + q[3]) * z + q[4])
    return x * z + x


def _satan(x):
    if x <= 0.66:
        return _xatan(x)
    if x > TAN3PIO8:
        return math.pi / 2 - _xatan(1.0 / x) + MOREBITS
    return math.pi / 4 + _xatan((x - 1.0) / (x + 1.0)) + 0.5 * MOREBITS


def atan(x):
    if x == 0.0:
        return x
    if x > 0.0:
        return _satan(x)
    return -_satan(-x)


def asin(x):
    if x == 0.0:
        return x
    sign = False
    if x < 0.0:
        x = -x
        sign = True
    if x > 1.0 or x != x:
        return math.nan
    temp = math.sqrt(1.0 - x * x)
    if x > 0.7:
        temp = math.pi / 2 - _satan(temp / x)
    else:
        temp = _satan(x / temp) if temp != 0.0 else math.pi / 2
    return -temp if sign else temp


def acos(x):
    return math.pi / 2 - asin(x)


def bits(x):
    """IEEE-754 bit pattern, for exact comparisons in tests."""
    return struct.unpack("<Q", struct.pack("<d", x))[0]
