// gomath.cuh - the Go standard-library arithmetic on the render path, as __host__ __device__
// code for the two precision modes of the renderer.
//
// Call sites in the reference: math.Tan raytracer.go:602 | math.Pow :395, :466 | math.Cos/Sin
// :518-519 and evaluator.go:929-937 | math.Acos :147 | math.Atan2 :345 | math/rand/v2 PCG :634,
// :642-643.  Go's implementations are pure-Go Cephes ports (sin.go, tan.go, atan.go, asin.go,
// pow.go) and PCG-DXSM (rand/v2/pcg.go); they are outside /root/reference, so the published
// algorithms are restated here.
//
//  * gm<double>: parity mode.  Same operation order as Go; the translation unit that
//    instantiates it is built with -fmad=false, so every product and sum rounds separately,
//    exactly as a Go amd64 build does.  pow with a fractional exponent uses exp/log of the
//    CUDA math library (last-ulp parity with Go unpinned; integer exponents are exact).
//  * gm<float>:  throughput mode.  CUDA's single-precision library functions; FMA allowed.
#pragma once
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

#define GM_HD __host__ __device__ __forceinline__
#define GM_PI 3.14159265358979323846264338327950288

namespace gomath {

// ------------------------------------------------------------------ PCG-DXSM (128-bit LCG)
struct Pcg { uint64_t hi, lo; };

GM_HD uint64_t mulhi64(uint64_t a, uint64_t b) {
#ifdef __CUDA_ARCH__
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * b) >> 64);
#endif
}

constexpr uint64_t PCG_MUL_HI = 0x2360ed051fc65da4ULL, PCG_MUL_LO = 0x4385df649fccf645ULL;
constexpr uint64_t PCG_INC_HI = 0x5851f42d4c957f2dULL, PCG_INC_LO = 0x14057b7ef767814fULL;

// (hi,lo) = (hi,lo) * (mh,ml) + (ih,il)  mod 2^128
GM_HD void affine128(uint64_t& hi, uint64_t& lo, uint64_t mh, uint64_t ml, uint64_t ih, uint64_t il) {
    uint64_t plo = lo * ml;
    uint64_t phi = mulhi64(lo, ml) + hi * ml + lo * mh;
    uint64_t nlo = plo + il;
    phi += ih + (nlo < plo ? 1u : 0u);
    hi = phi; lo = nlo;
}

GM_HD uint64_t pcg_next(Pcg& p) {
    affine128(p.hi, p.lo, PCG_MUL_HI, PCG_MUL_LO, PCG_INC_HI, PCG_INC_LO);
    uint64_t h = p.hi;
    h ^= h >> 32;
    h *= 0xda942042e4dd58b5ULL;
    h ^= h >> 48;
    h *= (p.lo | 1);
    return h;
}

GM_HD double pcg_float64(Pcg& p) { return (double)(pcg_next(p) << 11 >> 11) / 9007199254740992.0; }

// Jump table entry k: the affine map that advances the generator by k steps.
struct PcgJump { uint64_t mh, ml, ih, il; };

// ------------------------------------------------------------------ double: Go's algorithms
GM_HD void trig_reduce(double x, uint64_t& j, double& z) {
    const double PI4A = 7.85398125648498535156e-1, PI4B = 3.77489470793079817668e-8, PI4C = 2.69515142907905952645e-15;
    const double M4PI = 1.2732395447351628;   // 4/pi as binary64
    j = (uint64_t)(x * M4PI);
    double y = (double)j;
    if (j & 1) { j++; y++; }
    z = ((x - y * PI4A) - y * PI4B) - y * PI4C;
}
GM_HD double sin_poly(double z, double zz) {
    return z + z * zz * ((((((1.58962301576546568060e-10 * zz) + -2.50507477628578072866e-8) * zz
        + 2.75573136213857245213e-6) * zz + -1.98412698295895385996e-4) * zz
        + 8.33333333332211858878e-3) * zz + -1.66666666666666307295e-1);
}
GM_HD double cos_poly(double zz) {
    return 1.0 - 0.5 * zz + zz * zz * ((((((-1.13585365213876817300e-11 * zz) + 2.08757008419747316778e-9) * zz
        + -2.75573141792967388112e-7) * zz + 2.48015872888517045348e-5) * zz
        + -1.38888888888730564116e-3) * zz + 4.16666666666665929218e-2);
}
GM_HD double go_sin(double x) {
    if (x == 0 || x != x) return x;
    if (isinf(x)) return NAN;
    bool sign = false;
    if (x < 0) { x = -x; sign = true; }
    uint64_t j; double z;
    trig_reduce(x, j, z);
    j &= 7;
    if (j > 3) { sign = !sign; j -= 4; }
    double zz = z * z;
    double y = (j == 1 || j == 2) ? cos_poly(zz) : sin_poly(z, zz);
    return sign ? -y : y;
}
GM_HD double go_cos(double x) {
    if (x != x || isinf(x)) return NAN;
    bool sign = false;
    x = fabs(x);
    uint64_t j; double z;
    trig_reduce(x, j, z);
    j &= 7;
    if (j > 3) { j -= 4; sign = !sign; }
    if (j > 1) sign = !sign;
    double zz = z * z;
    double y = (j == 1 || j == 2) ? sin_poly(z, zz) : cos_poly(zz);
    return sign ? -y : y;
}
GM_HD double go_tan(double x) {
    if (x == 0 || x != x) return x;
    if (isinf(x)) return NAN;
    bool sign = false;
    if (x < 0) { x = -x; sign = true; }
    uint64_t j; double z;
    trig_reduce(x, j, z);
    double zz = z * z, y;
    if (zz > 1e-14)
        y = z + z * (zz * (((-1.30936939181383777646e4 * zz) + 1.15351664838587416140e6) * zz + -1.79565251976484877988e7)
            / ((((zz + 1.36812963470692954678e4) * zz + -1.32089234440210967447e6) * zz + 2.50083801823357915839e7) * zz
               + -5.38695755929454629881e7));
    else
        y = z;
    if (j & 2) y = -1.0 / y;
    return sign ? -y : y;
}
GM_HD double xatan(double x) {
    double z = x * x;
    z = z * ((((-8.750608600031904122785e-01 * z + -1.615753718733365076637e+01) * z + -7.500855792314704667340e+01) * z
              + -1.228866684490136173410e+02) * z + -6.485021904942025371773e+01)
        / (((((z + 2.485846490142306297962e+01) * z + 1.650270098316988542046e+02) * z + 4.328810604912902668951e+02) * z
            + 4.853903996359136964868e+02) * z + 1.945506571482613964425e+02);
    return x * z + x;
}
GM_HD double satan(double x) {
    const double morebits = 6.123233995736765886130e-17, tan3pio8 = 2.41421356237309504880;
    if (x <= 0.66) return xatan(x);
    if (x > tan3pio8) return GM_PI / 2 - xatan(1.0 / x) + morebits;
    return GM_PI / 4 + xatan((x - 1.0) / (x + 1.0)) + 0.5 * morebits;
}
GM_HD double go_atan(double x) {
    if (x == 0) return x;
    return x > 0 ? satan(x) : -satan(-x);
}
GM_HD double go_atan2(double y, double x) {
    if (x != x || y != y) return NAN;
    if (y == 0) return (x >= 0 && !signbit(x)) ? copysign(0.0, y) : copysign(GM_PI, y);
    if (x == 0) return copysign(GM_PI / 2, y);
    if (isinf(x)) {
        if (x > 0) return isinf(y) ? copysign(GM_PI / 4, y) : copysign(0.0, y);
        return isinf(y) ? copysign(3 * GM_PI / 4, y) : copysign(GM_PI, y);
    }
    if (isinf(y)) return copysign(GM_PI / 2, y);
    double q = go_atan(y / x);
    if (x < 0) return q <= 0 ? q + GM_PI : q - GM_PI;
    return q;
}
GM_HD double go_asin(double x) {
    if (x == 0) return x;
    bool sign = false;
    if (x < 0) { x = -x; sign = true; }
    if (x > 1 || x != x) return NAN;
    double temp = sqrt(1 - x * x);
    if (x > 0.7) temp = GM_PI / 2 - satan(temp / x);
    else temp = satan(x / temp);
    return sign ? -temp : temp;
}
GM_HD double go_acos(double x) { return GM_PI / 2 - go_asin(x); }

GM_HD bool is_odd_int(double y) {
    if (fabs(y) >= 9007199254740992.0) return false;
    double yi; double yf = modf(y, &yi);
    return yf == 0 && ((int64_t)yi & 1);
}
// math.Pow: special cases, then x^yi by repeated squaring on the frexp mantissa (exact
// restatement), times exp(yf*log x) for a fractional part.
GM_HD double go_pow(double x, double y) {
    if (y == 0 || x == 1) return 1;
    if (y == 1) return x;
    if (x != x || y != y) return NAN;
    if (x == 0) {
        if (y < 0) return (signbit(x) && is_odd_int(y)) ? copysign(INFINITY, x) : INFINITY;
        return (signbit(x) && is_odd_int(y)) ? x : 0;
    }
    if (isinf(y)) {
        if (x == -1) return 1;
        return ((fabs(x) < 1) == (y > 0)) ? 0 : INFINITY;
    }
    if (isinf(x)) {
        if (x < 0) { double r = 1 / x; return (-y) < 0 ? ((signbit(r) && is_odd_int(-y)) ? -INFINITY : INFINITY)
                                                        : ((signbit(r) && is_odd_int(-y)) ? r : 0); }
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
        a1 = exp(yf * log(x));
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

// ------------------------------------------------------------------ per-precision dispatch
template <typename R> struct gm;

template <> struct gm<double> {
    static GM_HD double sin(double x) { return go_sin(x); }
    static GM_HD double cos(double x) { return go_cos(x); }
    static GM_HD double acos(double x) { return go_acos(x); }
    static GM_HD double asin(double x) { return go_asin(x); }
    static GM_HD double atan2(double y, double x) { return go_atan2(y, x); }
    static GM_HD double pow(double x, double y) { return go_pow(x, y); }
    static GM_HD double sqrt(double x) { return ::sqrt(x); }
    static GM_HD double abs(double x) { return ::fabs(x); }
    static GM_HD double floor(double x) { return ::floor(x); }
    static GM_HD double trunc(double x) { return ::trunc(x); }
    static GM_HD double inf() { return INFINITY; }
};

template <> struct gm<float> {
    static GM_HD float sin(float x) { return ::sinf(x); }
    static GM_HD float cos(float x) { return ::cosf(x); }
    static GM_HD float acos(float x) { return ::acosf(x); }
    static GM_HD float asin(float x) { return ::asinf(x); }
    static GM_HD float atan2(float y, float x) { return ::atan2f(y, x); }
    static GM_HD float pow(float x, float y) {
        if (y == 0.f || x == 1.f) return 1.f;     // keep Go's special cases exact (Pow(0,0) = 1)
        if (y == 1.f) return x;
        return ::powf(x, y);
    }
    static GM_HD float sqrt(float x) { return ::sqrtf(x); }
    static GM_HD float abs(float x) { return ::fabsf(x); }
    static GM_HD float floor(float x) { return ::floorf(x); }
    static GM_HD float trunc(float x) { return ::truncf(x); }
    static GM_HD float inf() { return INFINITY; }
};

// math.Max(0, x) and clamp(0,1,x) = Min(Max(x,0),1) with Go's NaN / signed-zero rules (vec.go:217-220)
template <typename R> GM_HD R max0(R x) { return x > R(0) ? x : (x != x ? x : R(0)); }
template <typename R> GM_HD R clamp01(R x) {
    R m = max0(x);
    return m < R(1) ? m : (m != m ? m : R(1));
}

}  // namespace gomath
