/*
 * oracle/gomath.h - TEST INFRASTRUCTURE (CPU oracle), never linked into the product.
 *
 * Go standard-library arithmetic the reference's render path calls, restated in C.
 * The sources are NOT under /root/reference (Go stdlib, go 1.24.5 per go.mod:3); the
 * published algorithms (Cephes ports in src/math, PCG-DXSM in math/rand/v2) are
 * restated from their descriptions.  Call sites in the reference:
 *   math.Tan   raytracer.go:602          math.Pow  raytracer.go:395, 466
 *   math.Cos/Sin raytracer.go:518-519    math.Acos raytracer.go:147
 *   math.Atan2 raytracer.go:345          rand.NewPCG / Float64 raytracer.go:634, 642-643
 * Pinned by: PCG against Go's own test vector NewPCG(1,2) (tests/test_oracle.py) and,
 * together with tan/atan2/pow(int), by the seven byte-exact golden images.
 * Unpinned: acos, pow with a fractional exponent (uses libm exp/log), trig above 2^29.
 *
 * Build with -ffp-contract=off: every a*b+c below must round twice, as on amd64 Go.
 */
#ifndef ORACLE_GOMATH_H
#define ORACLE_GOMATH_H
#include <math.h>
#include <stdint.h>

#define GO_PI 3.14159265358979323846264338327950288419716939937510582097494459

/* ---- math/rand/v2 PCG (DXSM output), 128-bit LCG ---- */
typedef struct { uint64_t hi, lo; } go_pcg;

static inline void go_pcg_seed(go_pcg* p, uint64_t seed1, uint64_t seed2) { p->hi = seed1; p->lo = seed2; }

static inline uint64_t go_pcg_uint64(go_pcg* p) {
    const uint64_t mul_hi = 0x2360ed051fc65da4ULL, mul_lo = 0x4385df649fccf645ULL;
    const uint64_t inc_hi = 0x5851f42d4c957f2dULL, inc_lo = 0x14057b7ef767814fULL;
    /* state = state * mul + inc (mod 2^128) */
    unsigned __int128 prod = (unsigned __int128)p->lo * mul_lo;
    uint64_t lo = (uint64_t)prod;
    uint64_t hi = (uint64_t)(prod >> 64) + p->hi * mul_lo + p->lo * mul_hi;
    uint64_t lo2 = lo + inc_lo;
    uint64_t carry = lo2 < lo;
    hi = hi + inc_hi + carry;
    p->lo = lo2; p->hi = hi;
    /* DXSM: double xorshift multiply */
    const uint64_t cheap = 0xda942042e4dd58b5ULL;
    hi ^= hi >> 32;
    hi *= cheap;
    hi ^= hi >> 48;
    hi *= (lo2 | 1);
    return hi;
}

static inline double go_pcg_float64(go_pcg* p) {
    return (double)(go_pcg_uint64(p) << 11 >> 11) / 9007199254740992.0; /* 2^53 */
}

/* ---- Cephes trig with 3-part pi/4 reduction ---- */
#define GO_PI4A 7.85398125648498535156e-1
#define GO_PI4B 3.77489470793079817668e-8
#define GO_PI4C 2.69515142907905952645e-15
#define GO_M4PI 1.2732395447351628 /* 4/pi rounded to binary64 */

static inline double go__sinpoly(double z, double zz) {
    return z + z * zz * ((((((1.58962301576546568060e-10 * zz) + -2.50507477628578072866e-8) * zz
        + 2.75573136213857245213e-6) * zz + -1.98412698295895385996e-4) * zz
        + 8.33333333332211858878e-3) * zz + -1.66666666666666307295e-1);
}
static inline double go__cospoly(double zz) {
    return 1.0 - 0.5 * zz + zz * zz * ((((((-1.13585365213876817300e-11 * zz) + 2.08757008419747316778e-9) * zz
        + -2.75573141792967388112e-7) * zz + 2.48015872888517045348e-5) * zz
        + -1.38888888888730564116e-3) * zz + 4.16666666666665929218e-2);
}
static inline void go__reduce(double x, uint64_t* j, double* z) {
    uint64_t jj = (uint64_t)(x * GO_M4PI);
    double y = (double)jj;
    if (jj & 1) { jj++; y++; }
    *j = jj;
    *z = ((x - y * GO_PI4A) - y * GO_PI4B) - y * GO_PI4C;
}
static inline double go_sin(double x) {
    if (x == 0 || x != x) return x;
    if (isinf(x)) return NAN;
    int sign = 0;
    if (x < 0) { x = -x; sign = 1; }
    uint64_t j; double z;
    go__reduce(x, &j, &z);
    j &= 7;
    if (j > 3) { sign = !sign; j -= 4; }
    double zz = z * z;
    double y = (j == 1 || j == 2) ? go__cospoly(zz) : go__sinpoly(z, zz);
    return sign ? -y : y;
}
static inline double go_cos(double x) {
    if (x != x || isinf(x)) return NAN;
    int sign = 0;
    x = fabs(x);
    uint64_t j; double z;
    go__reduce(x, &j, &z);
    j &= 7;
    if (j > 3) { j -= 4; sign = !sign; }
    if (j > 1) sign = !sign;
    double zz = z * z;
    double y = (j == 1 || j == 2) ? go__sinpoly(z, zz) : go__cospoly(zz);
    return sign ? -y : y;
}
static inline double go_tan(double x) {
    if (x == 0 || x != x) return x;
    if (isinf(x)) return NAN;
    int sign = 0;
    if (x < 0) { x = -x; sign = 1; }
    uint64_t j; double z;
    go__reduce(x, &j, &z);
    double zz = z * z, y;
    if (zz > 1e-14) {
        y = z + z * (zz * (((-1.30936939181383777646e4 * zz) + 1.15351664838587416140e6) * zz + -1.79565251976484877988e7)
            / ((((zz + 1.36812963470692954678e4) * zz + -1.32089234440210967447e6) * zz + 2.50083801823357915839e7) * zz
               + -5.38695755929454629881e7));
    } else {
        y = z;
    }
    if (j & 2) y = -1.0 / y;
    return sign ? -y : y;
}

/* ---- atan / atan2 / asin / acos ---- */
static inline double go__xatan(double x) {
    double z = x * x;
    z = z * ((((-8.750608600031904122785e-01 * z + -1.615753718733365076637e+01) * z + -7.500855792314704667340e+01) * z
              + -1.228866684490136173410e+02) * z + -6.485021904942025371773e+01)
        / (((((z + 2.485846490142306297962e+01) * z + 1.650270098316988542046e+02) * z + 4.328810604912902668951e+02) * z
            + 4.853903996359136964868e+02) * z + 1.945506571482613964425e+02);
    return x * z + x;
}
static inline double go__satan(double x) {
    const double morebits = 6.123233995736765886130e-17, tan3pio8 = 2.41421356237309504880;
    if (x <= 0.66) return go__xatan(x);
    if (x > tan3pio8) return GO_PI / 2 - go__xatan(1.0 / x) + morebits;
    return GO_PI / 4 + go__xatan((x - 1.0) / (x + 1.0)) + 0.5 * morebits;
}
static inline double go_atan(double x) {
    if (x == 0) return x;
    if (x > 0) return go__satan(x);
    return -go__satan(-x);
}
static inline double go_atan2(double y, double x) {
    if (x != x || y != y) return NAN;
    if (y == 0) {
        if (x >= 0 && !signbit(x)) return copysign(0.0, y);
        return copysign(GO_PI, y);
    }
    if (x == 0) return copysign(GO_PI / 2, y);
    if (isinf(x)) {
        if (x > 0) return isinf(y) ? copysign(GO_PI / 4, y) : copysign(0.0, y);
        return isinf(y) ? copysign(3 * GO_PI / 4, y) : copysign(GO_PI, y);
    }
    if (isinf(y)) return copysign(GO_PI / 2, y);
    double q = go_atan(y / x);
    if (x < 0) return q <= 0 ? q + GO_PI : q - GO_PI;
    return q;
}
static inline double go_asin(double x) {
    if (x == 0) return x;
    int sign = 0;
    if (x < 0) { x = -x; sign = 1; }
    if (x > 1 || x != x) return NAN;
    double temp = sqrt(1 - x * x);
    if (x > 0.7) temp = GO_PI / 2 - go__satan(temp / x);
    else temp = go__satan(x / temp);
    return sign ? -temp : temp;
}
static inline double go_acos(double x) { return GO_PI / 2 - go_asin(x); }

/* ---- math.Pow: exact repeated squaring for the integer part of y ---- */
static inline int go__is_odd_int(double y) {
    if (fabs(y) >= 9007199254740992.0) return 0;
    double yi; double yf = modf(y, &yi);
    return yf == 0 && ((int64_t)yi & 1);
}
static inline double go_pow(double x, double y) {
    if (y == 0 || x == 1) return 1;
    if (y == 1) return x;
    if (x != x || y != y) return NAN;
    if (x == 0) {
        if (y < 0) return (signbit(x) && go__is_odd_int(y)) ? copysign(INFINITY, x) : INFINITY;
        return (signbit(x) && go__is_odd_int(y)) ? x : 0;
    }
    if (isinf(y)) {
        if (x == -1) return 1;
        return ((fabs(x) < 1) == (y > 0)) ? 0 : INFINITY;
    }
    if (isinf(x)) {
        if (x < 0) return go_pow(1 / x, -y);
        return y < 0 ? 0 : INFINITY;
    }
    if (y == 0.5) return sqrt(x);
    if (y == -0.5) return 1 / sqrt(x);
    double yi, yf = modf(fabs(y), &yi);
    if (yf != 0 && x < 0) return NAN;
    if (yi >= 9.223372036854775808e18) {
        if (x == -1) return 1;
        return ((fabs(x) < 1) == (y > 0)) ? 0 : INFINITY;
    }
    double a1 = 1.0;
    int ae = 0;
    if (yf != 0) {
        if (yf > 0.5) { yf--; yi++; }
        a1 = exp(yf * log(x)); /* Go: Exp(yf*Log(x)); last-ulp parity with Go's own Exp/Log is unpinned */
    }
    int xe;
    double x1 = frexp(x, &xe);
    for (int64_t i = (int64_t)yi; i != 0; i >>= 1) {
        if (xe < -(1 << 12) || (1 << 12) < xe) { ae += xe; break; }
        if (i & 1) { a1 *= x1; ae += xe; }
        x1 *= x1;
        xe <<= 1;
        if (x1 < .5) { x1 += x1; xe--; }
    }
    if (y < 0) { a1 = 1 / a1; ae = -ae; }
    return ldexp(a1, ae);
}

/* math.Max(0,x) / math.Min semantics incl. NaN and signed zero (used by clamp, vec.go:217-220) */
static inline double go_max(double x, double y) {
    if (isinf(x) && x > 0) return x;
    if (isinf(y) && y > 0) return y;
    if (x != x || y != y) return NAN;
    if (x == 0 && x == y) return signbit(x) ? y : x;
    return x > y ? x : y;
}
static inline double go_min(double x, double y) {
    if (isinf(x) && x < 0) return x;
    if (isinf(y) && y < 0) return y;
    if (x != x || y != y) return NAN;
    if (x == 0 && x == y) return signbit(x) ? x : y;
    return x < y ? x : y;
}
#endif
