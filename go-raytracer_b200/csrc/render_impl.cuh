// render_impl.cuh - the wavefront ray tracer, templated on the arithmetic type.
//
// Replaces Render / traceRay / closestHit / computeLighting / inShadow and the per-primitive
// Intersect / ComputeSurfaceProps of /root/reference/raytracer.go (line references at each
// function).  The reference is a recursive tracer run by 8 goroutines; this is a wavefront:
//
//   gen        camera rays for a band of rows, PCG jitter by jump-ahead        (Render :632-652)
//   closest    every ray x every leaf, leaves staged in shared memory by TMA   (closestHit :469-483)
//   shade      surface function (bytecode), material, shadow + child rays      (traceRay :498-556)
//   shadow     any-hit scan with early exit                                    (inShadow :411-429)
//   resolve    lighting + clamp per recursion level, deepest level first       (:372-401, :506-561)
//   pixels     mean of 4 samples -> RGBA8                                      (:656, vec.go:104-107)
//
// Rays of one recursion level live in SoA queues; children are appended with warp-aggregated
// (__ballot_sync/__popc) atomics, so the scan kernels always run on dense queues.  Colour is
// clamped at every level in the reference, hence the per-level records and the
// back-to-front resolve instead of a throughput-weighted accumulation.
//
// This header is compiled twice: render_f64.cu (R = double, -fmad=false: reference operation
// order, bit-parity mode) and render_f32.cu (R = float, FMA contraction: throughput mode).
#pragma once
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <algorithm>
#include <cstdlib>
#include <string>
#include <type_traits>
#include <vector>

#include <cuda_runtime.h>

#include "gomath.cuh"
#include "host_scene.h"
#include "renderer.h"

#ifndef GMLRT_PARITY_TU
#define GMLRT_PARITY_TU 0   // render_f64.cu defines it to 1 (and is built with -fmad=false)
#endif

namespace gmlrt {

using gomath::gm;

#define GMLRT_MAX_DEPTH 64
#define GMLRT_PLANE_GRIDS 2
#define GMLRT_PLANE_GRID 64
#define NREC 16   // per-node shading record: P(3) N(3) colour(3) refl transp kr kd ks n |d|

template <typename R> struct DevLight { int kind; R pos[3], color[3], axis[3], cos_cutoff, exponent; };
template <typename R> struct DevSurface { int kind, code_off, code_len, n_regs; R material[10]; };

template <typename R> struct DevScene {
    int n_leaves;                  // plain objects, device order (sorted by kind)
    int kind_begin[K_OBJ_KINDS + 1];   // objects [0, n_leaves) sorted by kind; CSG objects last
    int n_all;                     // objects + CSG member leaves (the per-leaf arrays below have n_all entries)
    const int* csg_prog;           // postfix programs: >= 0 push member leaf (device index), -1 - GMLRT_NODE_* operator
    const int2* csg_range;         // [n_csg] (offset, length) into csg_prog, indexed by object - kind_begin[K_CSG]
    const int* obj_flops;          // [n_leaves] credited flops of one test of the object
    int n_cull;                    // n_leaves rounded up to a multiple of GROUP
    const unsigned long long* cum_flops;   // [n] credited flops of scan-order objects 0..s
    const R* w2o;                  // [n][12] rows 0..2 of WorldToObject
    const int* scan_of;            // device index -> scan-order index (tie-breaks, ids)
    int pickless;                  // float32: the scan folds (bits of T) << 32 | leaf << 4 | flip << 3 | face (mode 2; mode 1, scenes of spheres
                                   // and planes only: just the device index - measured 0.3 ms faster in k_shade), which IS the closest hit -
                                   // k_shade / k_resolve decode it, the exact kernels store no per-candidate result and the pick pass is
                                   // not launched.  A tie at bit-identical float32 T goes to the smaller device index (objects are sorted
                                   // by kind and keep scan order within a kind; CSG members come last), where the reference takes the
                                   // first object in scan order: the two differ only for coincident surfaces of different kinds.
    const double* w2o64;           // [n][12] the same matrices in double, for the robust primary-ray tests (f32 mode)
    const R* o2w;                  // [n][12]
    const R* fnormal;              // [n][18] world normals of plane / cube faces
    const int* surface;            // [n]
    const int* kind;               // [n]
    int n_lights;
    const DevLight<R>* lights;
    const DevSurface<R>* surfaces;
    const R* fpool; const int* itab; const uint32_t* code;
    int n_fpool, n_itab;
    int W, H, depth;
    int key_shift, key_k;          // sort key of a spawned ray: (region >> key_shift) * 6 k^2 + cube-map cell of its direction
    // Sort regions (what "rays that start close together" means for the two ray sorts): a small object is its own
    // region; a point on a large object (an unbounded plane, or a bounding sphere much wider than a grid cell) belongs
    // to the cell of a uniform grid over the bounded objects.  Small regions are numbered along a Morton curve, so
    // that regions merged by a shift are neighbours in space.
    const int* rid_of;             // [n_all] region of a small object (CSG member leaves: their object's), -1 = large (grid cell),
                                   //         -2 - k = unbounded plane with its own 2-D grid k (below)
    int n_small, grid_dim[3];
    R grid_lo[3], grid_inv[3];
    // An unbounded plane is visible far beyond the bounded objects (a ground plane runs to the horizon), where a
    // clamped grid would lump everything into its border cells.  Its points are keyed in its own object-space (x, z)
    // instead, through g(x) = x / (1 + |x|): PLANE_GRID^2 cells that are fine under the objects and grow with distance.
    int plane_base[GMLRT_PLANE_GRIDS];      // first region of plane grid k
    R plane_par[GMLRT_PLANE_GRIDS][3];      // object-space centre (x, z) and 1 / scale
    double vw, vh;
    double cam_su, cam_sv;         // vw / (W - 1), vh / (H - 1)
    R ambient[3], bg0[3], bg1[3];
};

// One recursion level of the wavefront (SoA, `cap` entries each).
template <typename R> struct Level {
    R* ray;       // [cap][8] one record per ray: origin xyz, direction xyz, then (as raw bits of word 6) the parent: index in
                  //          the previous level | (1<<31 for the refraction slot).  32 bytes in f32: one sector per random read
    int* hit;     // [cap] device leaf index | face << 28, or -1
    R* t;         // [cap] T of the hit; during the scan: bit pattern of the smallest T so far (all ones = none)
    int* order;   // scan scratch.  f64: [cap] smallest scan index among the hits that have the smallest T; f32: [cap] 64-bit keys
                  //               (bits of T) << 32 | scan index, folded by one atomicMin (fold_hit)
    R* rec;       // [NREC][cap]
    R* child;     // [6][cap] colour returned by the reflection / refraction child
    int cap;
};

template <typename R> struct ShadowQueue {      // one segment of `per_light` slots per light: slot = light * per_light + k
    R* q;         // [cap][8] origin xyz, direction xyz, distance to light, |incoming direction| (one 32-byte record in f32:
                  //          the cull kernel reads the rays through the by-object permutation, one sector per ray)
    int* perm;    // [cap] per light segment: slots (relative to the segment) ordered by the shaded object
    int* skip;    // [cap] device leaf index of the shaded object (inShadow skips it)
    int* key;     // [cap] sort region of the shaded point (region_of)
    int* target;  // [cap] node * n_lights + light
    int* first;   // [cap] scan result: smallest scan index of an occluder, 0x7fffffff = none
    int* count;   // [n_lights] queued rays per light
    int cap, per_light;
};

struct Counters {                  // device-resident, zeroed per trace
    int count[GMLRT_MAX_DEPTH + 1];     // rays per level
    int error;
    int overflow;                       // a cull launch ran out of candidate queue: retrace with smaller tiles
    unsigned long long cull_tests[3];   // executed per-ray cull tests by family (GENERAL, PIVOT, AXIS)
    unsigned long long packet_tests;    // executed per-warp packet tests (one lane x one object)
    unsigned long long dbg[8];          // GMLRT_DEBUG_PACKETS: warp blocks, usable, objects offered, survivors
    unsigned long long rays_reflect, rays_refract, rays_shadow, shadow_occluded, rays_hit;
    unsigned long long shadow_tests, shadow_flops;   // credited: the reference's scan up to its first occluder (counting mode)
    unsigned long long exact_tests, exact_flops;     // executed second-phase tests and their table flops (counting mode)
    unsigned long long exact_tests_shadow;           // ... the part of exact_tests that k_exact_shadow ran
};

template <typename R> struct Ray { R ox, oy, oz, dx, dy, dz; };

// ray records (Level::ray, and the staging buffer spawned rays wait in until they are sorted)
constexpr int RAY_WORDS = 8;
__device__ __forceinline__ void load_ray(const float* rays, int i, Ray<float>& w) {
    const float4* p = reinterpret_cast<const float4*>(rays + (size_t)i * RAY_WORDS);
    const float4 a = p[0], b = p[1];
    w.ox = a.x; w.oy = a.y; w.oz = a.z; w.dx = a.w; w.dy = b.x; w.dz = b.y;
}
__device__ __forceinline__ void load_ray(const double* rays, int i, Ray<double>& w) {
    const double2* p = reinterpret_cast<const double2*>(rays + (size_t)i * RAY_WORDS);
    const double2 a = p[0], b = p[1], c = p[2];
    w.ox = a.x; w.oy = a.y; w.oz = b.x; w.dx = b.y; w.dy = c.x; w.dz = c.y;
}
__device__ __forceinline__ void store_ray(float* rays, int i, float ox, float oy, float oz, float dx, float dy, float dz, int parent) {
    float4* p = reinterpret_cast<float4*>(rays + (size_t)i * RAY_WORDS);
    p[0] = make_float4(ox, oy, oz, dx);
    p[1] = make_float4(dy, dz, __int_as_float(parent), 0.f);
}
__device__ __forceinline__ void store_ray(double* rays, int i, double ox, double oy, double oz, double dx, double dy, double dz, int parent) {
    double2* p = reinterpret_cast<double2*>(rays + (size_t)i * RAY_WORDS);
    p[0] = make_double2(ox, oy); p[1] = make_double2(oz, dx); p[2] = make_double2(dy, dz);
    p[3] = make_double2(__longlong_as_double((long long)(unsigned)parent), 0.0);
}
// one 8-word record (a shadow ray: origin, direction, distance, |incoming direction|) as 16-byte vector stores
__device__ __forceinline__ void store_rec8(float* recs, size_t i, float a, float b, float c, float d, float e, float f, float g, float h) {
    float4* p = reinterpret_cast<float4*>(recs + i * 8);
    p[0] = make_float4(a, b, c, d);
    p[1] = make_float4(e, f, g, h);
}
__device__ __forceinline__ void store_rec8(double* recs, size_t i, double a, double b, double c, double d, double e, double f, double g, double h) {
    double2* p = reinterpret_cast<double2*>(recs + i * 8);
    p[0] = make_double2(a, b); p[1] = make_double2(c, d); p[2] = make_double2(e, f); p[3] = make_double2(g, h);
}
template <typename R> __device__ __forceinline__ int ray_parent(const R* rays, int i) {
    return reinterpret_cast<const int*>(rays + (size_t)i * RAY_WORDS + 6)[0];      // little-endian: the low word of word 6
}


// ------------------------------------------------------------------ leaf matrix in shared memory
template <typename R> struct Mat34 { R m[12]; };

__device__ __forceinline__ void load_mat(const float* sm, int i, Mat34<float>& M) {
    const float4* p = reinterpret_cast<const float4*>(sm) + 3 * i;
    float4 a = p[0], b = p[1], c = p[2];
    M.m[0] = a.x; M.m[1] = a.y; M.m[2] = a.z; M.m[3] = a.w;
    M.m[4] = b.x; M.m[5] = b.y; M.m[6] = b.z; M.m[7] = b.w;
    M.m[8] = c.x; M.m[9] = c.y; M.m[10] = c.z; M.m[11] = c.w;
}
__device__ __forceinline__ void load_mat(const double* sm, int i, Mat34<double>& M) {
    const double2* p = reinterpret_cast<const double2*>(sm) + 6 * i;
#pragma unroll
    for (int k = 0; k < 6; k++) { double2 v = p[k]; M.m[2 * k] = v.x; M.m[2 * k + 1] = v.y; }
}

// rayToObjectSpace (raytracer.go:51-56) with Mat4.MulPoint / MulDir's operand order (vec.go:298-313)
template <typename R> __device__ __forceinline__ void to_object(const R* m, const Ray<R>& w, Ray<R>& o) {
    o.ox = m[0] * w.ox + m[1] * w.oy + m[2] * w.oz + m[3];
    o.oy = m[4] * w.ox + m[5] * w.oy + m[6] * w.oz + m[7];
    o.oz = m[8] * w.ox + m[9] * w.oy + m[10] * w.oz + m[11];
    o.dx = m[0] * w.dx + m[1] * w.dy + m[2] * w.dz;
    o.dy = m[4] * w.dx + m[5] * w.dy + m[6] * w.dz;
    o.dz = m[8] * w.dx + m[9] * w.dy + m[10] * w.dz;
}

// ------------------------------------------------------------------ first-hit rule per primitive
// Each returns true and (t, face) exactly as <Type>.Intersect would.

template <typename R> __device__ __forceinline__ bool hit_sphere(const Ray<R>& r, R& t, int& face) {   // :78-103
    R a = r.dx * r.dx + r.dy * r.dy + r.dz * r.dz;
    R hb = r.ox * r.dx + r.oy * r.dy + r.oz * r.dz;
    R c = r.ox * r.ox + r.oy * r.oy + r.oz * r.oz - R(1);
    R disc = hb * hb - a * c;
    if (disc < R(0)) return false;
    R t0 = (-hb - gm<R>::sqrt(disc)) / a;
    if (t0 > R(0)) { t = t0; face = 0; return true; }
    return false;                       // the far root is never used (:98-102)
}

template <typename R> __device__ __forceinline__ bool hit_plane(const Ray<R>& r, R& t, int& face) {    // :164-180
    R den = r.dy;                       // N = (0,1,0), D = 0 (evaluator.go:816-822)
    if (gm<R>::abs(den) < R(1e-6)) return false;
    R tt = (R(0) - r.oy) / den;
    if (tt <= R(0)) return false;
    t = tt; face = 0;
    return true;
}

// one cube face: plane test + inclusive bounds (:218-237).  The reference also bounds-checks the
// coordinate of the face's own plane, which is 0 or 1 only up to rounding: in float64 that test
// randomly rejects ~1.5% of cube.gml's pixels and parity mode reproduces it bit for
// bit.  In float32 with FMA the residual o + d*(num/den) is almost never exactly on the plane, so
// throughput mode skips the own-axis test (AXIS) - a documented deviation - also in its
// double-precision re-tests, which are built with FMA contraction like the rest of that unit.
template <typename R, int AXIS>
__device__ __forceinline__ void cube_face(const Ray<R>& r, R num, R den, int f, R& best, int& bf) {
    constexpr bool strict = GMLRT_PARITY_TU != 0;   // only the -fmad=false translation unit can honour the own-axis test
    if (gm<R>::abs(den) < R(1e-6)) return;
    R t = num / den;
    if (t <= R(0)) return;
    R px = r.ox + r.dx * t, py = r.oy + r.dy * t, pz = r.oz + r.dz * t;
    if ((strict || AXIS != 0) && (px < R(0) || px > R(1))) return;
    if ((strict || AXIS != 1) && (py < R(0) || py > R(1))) return;
    if ((strict || AXIS != 2) && (pz < R(0) || pz > R(1))) return;
    if (bf < 0 || t < best) { best = t; bf = f; }
}
template <typename R> __device__ __forceinline__ bool hit_cube(const Ray<R>& r, R& t, int& face) {     // :214-240
    R best = R(0); int bf = -1;
    cube_face<R, 2>(r, r.oz, -r.dz, 0, best, bf);             // Front  z=0, n=(0,0,-1)   (plane.go:29-38)
    cube_face<R, 2>(r, R(1) - r.oz, r.dz, 1, best, bf);       // Back   z=1
    cube_face<R, 0>(r, r.ox, -r.dx, 2, best, bf);             // Left   x=0
    cube_face<R, 0>(r, R(1) - r.ox, r.dx, 3, best, bf);       // Right  x=1
    cube_face<R, 1>(r, R(1) - r.oy, r.dy, 4, best, bf);       // Top    y=1
    cube_face<R, 1>(r, r.oy, -r.dy, 5, best, bf);             // Bottom y=0
    if (bf < 0) return false;
    t = best; face = bf;
    return true;
}

template <typename R> __device__ __forceinline__ bool hit_cylinder(const Ray<R>& r, R& t, int& face) { // :279-337
    R best = gm<R>::inf(); int bf = -1;
    R a = r.dx * r.dx + r.dz * r.dz;
    if (a > R(1e-12)) {
        R hb = r.ox * r.dx + r.oz * r.dz;
        R c0 = r.ox * r.ox + r.oz * r.oz - R(1);
        R disc = hb * hb - a * c0;
        if (disc >= R(0)) {
            R sq = gm<R>::sqrt(disc);
            R t0 = (-hb - sq) / a, t1 = (-hb + sq) / a;
            R y0 = r.oy + r.dy * t0;
            if (y0 >= R(0) && y0 <= R(1) && t0 > R(0) && t0 < best) { best = t0; bf = 0; }
            R y1 = r.oy + r.dy * t1;
            if (y1 >= R(0) && y1 <= R(1) && t1 > R(0) && t1 < best) { best = t1; bf = 0; }
        }
    }
    if (gm<R>::abs(r.dy) > R(1e-12)) {
        R tt = (R(1) - r.oy) / r.dy;
        R px = r.ox + r.dx * tt, pz = r.oz + r.dz * tt;
        if (px * px + pz * pz <= R(1) && tt > R(0) && tt < best) { best = tt; bf = 1; }
        R tb = -r.oy / r.dy;
        px = r.ox + r.dx * tb; pz = r.oz + r.dz * tb;
        if (px * px + pz * pz <= R(1) && tb > R(0) && tb < best) { best = tb; bf = 2; }
    }
    if (bf < 0) return false;
    t = best; face = bf;
    return true;
}

// EXTENSION (no reference behaviour): x^2 + z^2 <= y^2, 0 <= y <= 1; same candidate rule as the cylinder
template <typename R> __device__ __forceinline__ bool hit_cone(const Ray<R>& r, R& t, int& face) {
    R best = gm<R>::inf(); int bf = -1;
    R a = r.dx * r.dx + r.dz * r.dz - r.dy * r.dy;
    R hb = r.ox * r.dx + r.oz * r.dz - r.oy * r.dy;
    R c0 = r.ox * r.ox + r.oz * r.oz - r.oy * r.oy;
    if (gm<R>::abs(a) > R(1e-12)) {
        R disc = hb * hb - a * c0;
        if (disc >= R(0)) {
            R sq = gm<R>::sqrt(disc);
            R t0 = (-hb - sq) / a, t1 = (-hb + sq) / a;
            R y0 = r.oy + r.dy * t0;
            if (y0 >= R(0) && y0 <= R(1) && t0 > R(0) && t0 < best) { best = t0; bf = 0; }
            R y1 = r.oy + r.dy * t1;
            if (y1 >= R(0) && y1 <= R(1) && t1 > R(0) && t1 < best) { best = t1; bf = 0; }
        }
    } else if (gm<R>::abs(hb) > R(1e-12)) {
        R t0 = -c0 / (R(2) * hb);
        R y0 = r.oy + r.dy * t0;
        if (y0 >= R(0) && y0 <= R(1) && t0 > R(0) && t0 < best) { best = t0; bf = 0; }
    }
    if (gm<R>::abs(r.dy) > R(1e-12)) {
        R tt = (R(1) - r.oy) / r.dy;
        R px = r.ox + r.dx * tt, pz = r.oz + r.dz * tt;
        if (px * px + pz * pz <= R(1) && tt > R(0) && tt < best) { best = tt; bf = 1; }
    }
    if (bf < 0) return false;
    t = best; face = bf;
    return true;
}

template <typename R, int KIND> __device__ __forceinline__ bool hit_kind(const Ray<R>& r, R& t, int& face) {
    if (KIND == K_SPHERE) return hit_sphere(r, t, face);
    if (KIND == K_CUBE) return hit_cube(r, t, face);
    if (KIND == K_CYLINDER) return hit_cylinder(r, t, face);
    if (KIND == K_CONE) return hit_cone(r, t, face);
    return hit_plane(r, t, face);
}

// ------------------------------------------------------------------ CSG objects: interval span lists (EXTENSION)
// The reference rejects intersect / difference nodes (raytracer.go:825-826); semantics follow the
// ICFP-2000 text as restated by oracle/oracle.c (leaf_spans, spans_complement, spans_intersect,
// csg_spans, csg_intersect) - operation for operation, so that parity mode is bit-identical to it.
// A solid along a ray is a sorted list of disjoint parameter intervals; every finite bound remembers
// the member leaf, face and orientation of the surface it lies on.  The tree is walked as a postfix
// program over a small stack of lists (no recursion); lists live in local memory, which is fine
// because only rays that survive the object's bounding-sphere cull get here.
#define GMLRT_MAXSPANS 16
#define HIT_LEAF_MASK 0x07ffffff      // hit code: leaf | flip << 27 | face << 28
#define HIT_NONE_LEAF 0x07ffffff

template <typename R> struct Spans {
    int n;
    R t[2 * GMLRT_MAXSPANS];          // [2s] = entry bound of span s, [2s+1] = exit bound
    int info[2 * GMLRT_MAXSPANS];     // hit code of the surface at that bound (leaf = HIT_NONE_LEAF: at infinity)
};

__device__ __forceinline__ int bound_info(int leaf, int face) { return leaf | (face << 28); }

template <typename R> __device__ __forceinline__ void spans_push(Spans<R>& L, R tin, int iin, R tout, int iout) {
    if (!(tin < tout) || L.n >= GMLRT_MAXSPANS) return;
    L.t[2 * L.n] = tin; L.info[2 * L.n] = iin; L.t[2 * L.n + 1] = tout; L.info[2 * L.n + 1] = iout;
    L.n++;
}

// oracle.c leaf_spans: the parameter interval(s) where the object-space ray `r` is inside the leaf
template <typename R> __device__ __noinline__ void leaf_spans(int kind, int leaf, const Ray<R>& r, Spans<R>& L) {
    const R INF = gm<R>::inf();
    const int NONE = HIT_NONE_LEAF;
    L.n = 0;
    if (kind == K_SPHERE) {
        const R a = r.dx * r.dx + r.dy * r.dy + r.dz * r.dz, hb = r.ox * r.dx + r.oy * r.dy + r.oz * r.dz;
        const R c = r.ox * r.ox + r.oy * r.oy + r.oz * r.oz - R(1);
        const R disc = hb * hb - a * c;
        if (disc < R(0)) return;
        const R sq = gm<R>::sqrt(disc);
        spans_push(L, (-hb - sq) / a, bound_info(leaf, 0), (-hb + sq) / a, bound_info(leaf, 0));
    } else if (kind == K_PLANE) {                        // the half space y <= 0
        if (gm<R>::abs(r.dy) < R(1e-6)) {
            if (r.oy <= R(0)) spans_push(L, -INF, NONE, INF, NONE);
            return;
        }
        const R t = (R(0) - r.oy) / r.dy;
        if (r.dy > R(0)) spans_push(L, -INF, NONE, t, bound_info(leaf, 0));
        else spans_push(L, t, bound_info(leaf, 0), INF, NONE);
    } else if (kind == K_CUBE) {                         // slabs; faces as plane.go:29-38
        const int FACE_LO[3] = {2, 5, 0}, FACE_HI[3] = {3, 4, 1};
        const R oo[3] = {r.ox, r.oy, r.oz}, dd[3] = {r.dx, r.dy, r.dz};
        R lo = -INF, hi = INF; int ilo = NONE, ihi = NONE;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            if (gm<R>::abs(dd[k]) < R(1e-12)) { if (oo[k] < R(0) || oo[k] > R(1)) return; continue; }
            R ta = (R(0) - oo[k]) / dd[k], tb = (R(1) - oo[k]) / dd[k];
            int ia = bound_info(leaf, FACE_LO[k]), ib = bound_info(leaf, FACE_HI[k]);
            if (ta > tb) { const R tt = ta; ta = tb; tb = tt; const int ii = ia; ia = ib; ib = ii; }
            if (ta > lo) { lo = ta; ilo = ia; }
            if (tb < hi) { hi = tb; ihi = ib; }
        }
        spans_push(L, lo, ilo, hi, ihi);
    } else if (kind == K_CYLINDER) {
        R lo = -INF, hi = INF; int ilo = NONE, ihi = NONE;
        const R a = r.dx * r.dx + r.dz * r.dz;
        const R hb = r.ox * r.dx + r.oz * r.dz;
        const R c0 = r.ox * r.ox + r.oz * r.oz - R(1);
        if (a > R(1e-12)) {
            const R disc = hb * hb - a * c0;
            if (disc < R(0)) return;
            const R sq = gm<R>::sqrt(disc);
            lo = (-hb - sq) / a; ilo = bound_info(leaf, 0); hi = (-hb + sq) / a; ihi = bound_info(leaf, 0);
        } else if (c0 > R(0)) return;
        if (gm<R>::abs(r.dy) > R(1e-12)) {
            R t0 = (R(0) - r.oy) / r.dy, t1 = (R(1) - r.oy) / r.dy;
            int i0 = bound_info(leaf, 2), i1 = bound_info(leaf, 1);
            if (t0 > t1) { const R tt = t0; t0 = t1; t1 = tt; const int ii = i0; i0 = i1; i1 = ii; }
            if (t0 > lo) { lo = t0; ilo = i0; }
            if (t1 < hi) { hi = t1; ihi = i1; }
        } else if (r.oy < R(0) || r.oy > R(1)) return;
        spans_push(L, lo, ilo, hi, ihi);
    } else {                                             // cone: convex, min / max of the valid boundary candidates
        R cand[3]; int cf[3]; int nc = 0;
        const R a = r.dx * r.dx + r.dz * r.dz - r.dy * r.dy;
        const R hb = r.ox * r.dx + r.oz * r.dz - r.oy * r.dy;
        const R c0 = r.ox * r.ox + r.oz * r.oz - r.oy * r.oy;
        if (gm<R>::abs(a) > R(1e-12)) {
            const R disc = hb * hb - a * c0;
            if (disc >= R(0)) {
                const R sq = gm<R>::sqrt(disc);
                const R ts[2] = {(-hb - sq) / a, (-hb + sq) / a};
#pragma unroll
                for (int k = 0; k < 2; k++) {
                    const R y = r.oy + r.dy * ts[k];
                    if (y >= R(0) && y <= R(1)) { cand[nc] = ts[k]; cf[nc++] = 0; }
                }
            }
        } else if (gm<R>::abs(hb) > R(1e-12)) {
            const R t = -c0 / (R(2) * hb);
            const R y = r.oy + r.dy * t;
            if (y >= R(0) && y <= R(1)) { cand[nc] = t; cf[nc++] = 0; }
        }
        if (gm<R>::abs(r.dy) > R(1e-12)) {
            const R t = (R(1) - r.oy) / r.dy;
            const R px = r.ox + r.dx * t, pz = r.oz + r.dz * t;
            if (px * px + pz * pz <= R(1)) { cand[nc] = t; cf[nc++] = 1; }
        }
        if (nc < 2) return;
        int imin = 0, imax = 0;
        for (int k = 1; k < nc; k++) { if (cand[k] < cand[imin]) imin = k; if (cand[k] > cand[imax]) imax = k; }
        spans_push(L, cand[imin], bound_info(leaf, cf[imin]), cand[imax], bound_info(leaf, cf[imax]));
    }
}

#define HIT_FLIP_BIT (1 << 27)
template <typename R> __device__ __noinline__ void spans_complement(const Spans<R>& A, Spans<R>& out) {
    out.n = 0;
    R pt = -gm<R>::inf(); int pi = HIT_NONE_LEAF;
    for (int i = 0; i < A.n; i++) {
        spans_push(out, pt, pi, A.t[2 * i], A.info[2 * i] ^ HIT_FLIP_BIT);
        pt = A.t[2 * i + 1]; pi = A.info[2 * i + 1] ^ HIT_FLIP_BIT;
    }
    spans_push(out, pt, pi, gm<R>::inf(), HIT_NONE_LEAF);
}

template <typename R> __device__ __noinline__ void spans_intersect(const Spans<R>& A, const Spans<R>& B, Spans<R>& out) {
    out.n = 0;
    int i = 0, j = 0;
    while (i < A.n && j < B.n) {
        const bool ain = A.t[2 * i] >= B.t[2 * j], aout = A.t[2 * i + 1] <= B.t[2 * j + 1];
        spans_push(out, ain ? A.t[2 * i] : B.t[2 * j], ain ? A.info[2 * i] : B.info[2 * j],
                   aout ? A.t[2 * i + 1] : B.t[2 * j + 1], aout ? A.info[2 * i + 1] : B.info[2 * j + 1]);
        if (aout) i++; else j++;
    }
}

// csg_intersect: evaluate object `obj`'s postfix program for the world ray `w`; first surface with t > 0.
// `mats` = WorldToObject of every leaf in the arithmetic type of this instantiation.
template <typename R>
__device__ __noinline__ bool csg_hit(const int* prog, int len, const int* kinds, const R* mats, const Ray<R>& w, R& t, int& code) {
    // Stack of span lists without moving any list: `ids` is a permutation of the buffers, ids[0..sp) are the stack,
    // the rest are free (two are always free for the temporaries of an operator).  A list is 97-193 words of local
    // memory, so a result is adopted by swapping two indices, never by copying the struct.
    Spans<R> buf[GMLRT_CSG_MAX_STACK + 2];
    int ids[GMLRT_CSG_MAX_STACK + 2];
#pragma unroll
    for (int k = 0; k < GMLRT_CSG_MAX_STACK + 2; k++) ids[k] = k;
    auto swap_ids = [&](int x, int y) { const int tmp = ids[x]; ids[x] = ids[y]; ids[y] = tmp; };
    int sp = 0;
    for (int k = 0; k < len; k++) {
        const int op = prog[k];
        if (op >= 0) {                                   // push a member leaf
            if (sp >= GMLRT_CSG_MAX_STACK) return false;
            Ray<R> o;
            to_object(mats + (size_t)op * 12, w, o);
            leaf_spans<R>(kinds[op], op, o, buf[ids[sp++]]);
        } else {
            if (sp < 2) return false;
            Spans<R>& a = buf[ids[sp - 2]];
            Spans<R>& b = buf[ids[sp - 1]];
            Spans<R>& t1 = buf[ids[sp]];
            Spans<R>& t2 = buf[ids[sp + 1]];
            const int kind = -1 - op;
            if (kind == GMLRT_NODE_INTERSECT) { spans_intersect(a, b, t1); swap_ids(sp - 2, sp); }
            else if (kind == GMLRT_NODE_DIFFERENCE) { spans_complement(b, t1); spans_intersect(a, t1, t2); swap_ids(sp - 2, sp + 1); }
            else {                                       // union = complement(complement a  intersect  complement b)
                spans_complement(a, t1); spans_complement(b, t2);
                spans_intersect(t1, t2, a);
                spans_complement(a, t1); swap_ids(sp - 2, sp);
            }
            sp--;
        }
    }
    if (sp != 1) return false;
    const Spans<R>& L = buf[ids[0]];
    for (int i = 0; i < L.n; i++) {
        int b = -1;
        if (L.t[2 * i] > R(0)) b = 2 * i;
        else if (L.t[2 * i + 1] > R(0)) b = 2 * i + 1;
        if (b < 0) continue;
        if ((L.info[b] & HIT_LEAF_MASK) == HIT_NONE_LEAF) return false;   // the solid extends to infinity along the ray
        t = L.t[b]; code = L.info[b];
        return true;
    }
    return false;
}

// ------------------------------------------------------------------ TMA bulk copy + mbarrier (sm_90+/sm_100a)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared, `bytes` a multiple of 16, completion signalled on `bar` (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ------------------------------------------------------------------ two-phase scan: cull, then exact test
// closestHit / inShadow visit every object for every ray (raytracer.go:469-483, 411-429).  The scan
// kernels keep that all-pairs structure (no hierarchy), but split each ray x object test in two:
//
//   cull   a few FFMA + 1 compare against the object's bounding sphere (centre C, radius R), from a
//          16-byte (f32) record in shared memory; the radius carries a rounding margin so that a reject
//          is always a true miss.  Three record/ray "families", by what the rays of a launch share:
//            GENERAL  any ray (o, d^):  reject iff |C-o|^2 - ((C-o).d^)^2 > R^2, expanded so that
//                     everything per-ray is hoisted:  A = C.(-2o) + K,  K = |C|^2 - R^2 - margin,
//                     B = C.d^ - o.d^,  reject <=> A - B^2 > -|o|^2 + margin          7 FFMA + compare
//            PIVOT    rays whose LINE passes through one point V (camera rays: the eye; shadow rays of a
//                     point / spot light: the light):  record p = C - V, K = |p|^2 - R^2 - margin,
//                     B = p.d^,  reject <=> K - B^2 > 0                                 4 FP + compare
//            AXIS     rays with one common direction (shadow rays of a directional light): project on
//                     the plane orthogonal to it,  du = Cu - ou, dv = Cv - ov,
//                     reject <=> du^2 + dv^2 - R^2 > 0                                  4 FP + compare
//          The survivors of a window of GROUP = 64 objects are collected as one 64-bit mask per ray.
//   exact  after each window every ray walks the set bits of its mask and runs the reference's own test
//          (<Type>.Intersect, or the CSG span lists) on those objects.  Deferring to the end of the window
//          makes the lanes of a warp run their k-th exact tests side by side instead of one lane at a time.
//
// Each thread carries RPT rays so that the record load and the loop overhead are shared.
#ifndef GMLRT_SCAN_THREADS
#define GMLRT_SCAN_THREADS 256
#endif
constexpr int SCAN_THREADS = GMLRT_SCAN_THREADS, SCAN_WARPS = SCAN_THREADS / 32;
constexpr int GROUP = 64;                   // objects per window: one 64-bit survivor mask per ray, kept in two registers
constexpr int CULL_UNROLL = 8;              // objects per unrolled body of the cull loop
enum CullFamily { CULL_GENERAL = 0, CULL_PIVOT = 1, CULL_AXIS = 2 };

__device__ __forceinline__ float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double fma_(double a, double b, double c) { return __fma_rn(a, b, c); }

template <typename R> struct CullMargin;
template <> struct CullMargin<float> { static constexpr double rel = 7.62939453125e-6;  /* 2^-17 */ static constexpr double radius = 9.765625e-4; /* 2^-10 */ };
template <> struct CullMargin<double> { static constexpr double rel = 5.6843418860808015e-14; /* 2^-44 */ static constexpr double radius = 9.313225746154785e-10; /* 2^-30 */ };

template <typename R> struct Cull4 { R x, y, z, k; };
__device__ __forceinline__ Cull4<float> load_cull(const float* sm, int i) {
    const float4 v = reinterpret_cast<const float4*>(sm)[i];
    Cull4<float> c = {v.x, v.y, v.z, v.w};
    return c;
}
__device__ __forceinline__ Cull4<double> load_cull(const double* sm, int i) {
    const double2 a = reinterpret_cast<const double2*>(sm)[2 * i], b = reinterpret_cast<const double2*>(sm)[2 * i + 1];
    Cull4<double> c = {a.x, a.y, b.x, b.y};
    return c;
}

// What a launch's cull table was built for (host: Renderer::build_cull_table).
template <typename R> struct CullFrame { int family; R e1[3], e2[3], axis[3]; };   // AXIS: orthonormal basis of the plane orthogonal to the common direction `axis` (the way the rays travel)

template <typename R, int FAMILY> struct CullRay;
// slack(L) < 0 (sign bit set)  <=>  the object may be hit; > 0  <=>  certainly missed
template <typename R> struct CullRay<R, CULL_GENERAL> {
    R o2x, o2y, o2z, dx, dy, dz, nod, ncmp;
    __device__ __forceinline__ void init(const Ray<R>& w, const CullFrame<R>&) {
        const R inv = R(1) / gm<R>::sqrt(fma_(w.dx, w.dx, fma_(w.dy, w.dy, w.dz * w.dz)));
        dx = w.dx * inv; dy = w.dy * inv; dz = w.dz * inv;
        o2x = R(-2) * w.ox; o2y = R(-2) * w.oy; o2z = R(-2) * w.oz;
        nod = -fma_(w.ox, dx, fma_(w.oy, dy, w.oz * dz));
        const R oo = fma_(w.ox, w.ox, fma_(w.oy, w.oy, w.oz * w.oz));
        ncmp = fma_(oo, R(-CullMargin<R>::rel), oo);                           // |o|^2 (1 - margin)
    }
    __device__ __forceinline__ R slack(const Cull4<R>& L) const {
        const R A = fma_(L.x, o2x, fma_(L.y, o2y, fma_(L.z, o2z, L.k)));
        const R B = fma_(L.x, dx, fma_(L.y, dy, fma_(L.z, dz, nod)));
        return fma_(-B, B, A) + ncmp;
    }
};
template <typename R> struct CullRay<R, CULL_PIVOT> {
    R dx, dy, dz;
    __device__ __forceinline__ void init(const Ray<R>& w, const CullFrame<R>&) {
        const R inv = R(1) / gm<R>::sqrt(fma_(w.dx, w.dx, fma_(w.dy, w.dy, w.dz * w.dz)));
        dx = w.dx * inv; dy = w.dy * inv; dz = w.dz * inv;
    }
    __device__ __forceinline__ R slack(const Cull4<R>& L) const {
        const R B = fma_(L.x, dx, fma_(L.y, dy, L.z * dz));
        return fma_(-B, B, L.k);
    }
};
template <typename R> struct CullRay<R, CULL_AXIS> {
    R ou, ov, ncmp;
    __device__ __forceinline__ void init(const Ray<R>& w, const CullFrame<R>& fr) {
        ou = fma_(w.ox, fr.e1[0], fma_(w.oy, fr.e1[1], w.oz * fr.e1[2]));
        ov = fma_(w.ox, fr.e2[0], fma_(w.oy, fr.e2[1], w.oz * fr.e2[2]));
        const R oo = fma_(w.ox, w.ox, fma_(w.oy, w.oy, w.oz * w.oz));
        ncmp = -R(CullMargin<R>::rel) * (oo + R(1));
    }
    __device__ __forceinline__ R slack(const Cull4<R>& L) const {          // record: (Cu, Cv, -R^2 - margin, unused)
        const R du = L.x - ou, dv = L.y - ov;
        return fma_(dv, dv, fma_(du, du, L.z)) + ncmp;
    }
};

// shift the sign bit of `s` into the low end of a 32-bit mask: one funnel shift per test
__device__ __forceinline__ uint32_t push_sign(uint32_t mask, float s) { return __funnelshift_l(__float_as_uint(s), mask, 1); }
__device__ __forceinline__ uint32_t push_sign(uint32_t mask, double s) { return __funnelshift_l((uint32_t)__double2hiint(s), mask, 1); }

// one window: objects first .. first + GROUP - 1 of the staged table `sm`; test j of each half lands in bit 31 - j
// of lo / hi (the masks are shifted left once per test), set = object survives ray r
template <typename R, int FAMILY, int RPT>
__device__ __forceinline__ void cull_window(const R* sm, int first, const CullRay<R, FAMILY> (&cr)[RPT], uint32_t (&lo)[RPT], uint32_t (&hi)[RPT]) {
#pragma unroll
    for (int r = 0; r < RPT; r++) { lo[r] = 0; hi[r] = 0; }
    // unrolled by CULL_UNROLL only: the fully unrolled window (1300+ instructions) missed the instruction cache
    // ("no_instruction" was the top stall); the masks are built by shifting, so the body needs no immediates
#pragma unroll 1
    for (int j0 = 0; j0 < 32; j0 += CULL_UNROLL) {
#pragma unroll
        for (int j = 0; j < CULL_UNROLL; j++) {
            const Cull4<R> L = load_cull(sm, first + j0 + j);
#pragma unroll
            for (int r = 0; r < RPT; r++) lo[r] = push_sign(lo[r], cr[r].slack(L));
        }
    }
#pragma unroll 1
    for (int j0 = 32; j0 < 64; j0 += CULL_UNROLL) {
#pragma unroll
        for (int j = 0; j < CULL_UNROLL; j++) {
            const Cull4<R> L = load_cull(sm, first + j0 + j);
#pragma unroll
            for (int r = 0; r < RPT; r++) hi[r] = push_sign(hi[r], cr[r].slack(L));
        }
    }
}

template <typename R> __device__ __forceinline__ void load_w2o(const R* w2o, int i, Mat34<R>& M) { load_mat(w2o, i, M); }   // global, same layout

__device__ __forceinline__ int kind_of(const int* kind_begin, int i) {
    return (i >= kind_begin[1]) + (i >= kind_begin[2]) + (i >= kind_begin[3]) + (i >= kind_begin[4]) + (i >= kind_begin[5]);
}

template <typename R> __device__ __forceinline__ bool hit_any_kind(int kind, const Ray<R>& o, R& t, int& face) {
    switch (kind) {
    case K_SPHERE: return hit_sphere(o, t, face);
    case K_CUBE: return hit_cube(o, t, face);
    case K_CYLINDER: return hit_cylinder(o, t, face);
    case K_CONE: return hit_cone(o, t, face);
    default: return hit_plane(o, t, face);
    }
}

// ------------------------------------------------------------------ the camera ray of sample i, in double
// Render (raytracer.go:632-652): sample i of a tile = pixel i / 4 (row-major from virtual row y0), sample i % 4.  Go's
// PCG-DXSM stream of the (column, 20-row batch) work item is entered by jump-ahead, so every sample draws exactly the
// numbers the reference draws.  k_gen rounds the result to the arithmetic type; the double-precision re-tests of the
// throughput mode call this again for the few candidates they see instead of reading a stored 48-byte copy per sample.
__constant__ gomath::PcgJump c_pcg_jump[160];   // affine maps advancing the PCG stream by 0..159 draws
struct CamSrc { int y0; RowMap map; };          // which rows the level-0 rays of the current tile are
template <typename R>
__device__ __forceinline__ void camera_ray64(const DevScene<R>& sc, const CamSrc& cam, int i, double (&o)[3], double (&d)[3]) {
    const int s = i & 3, pix = i >> 2;
    const int x = pix % sc.W, y = cam.map.actual(cam.y0 + pix / sc.W);
    const int ymin = (y / 20) * 20;                       // work item = (column, 20-row batch), :663-672
    const int q = 8 * (y - ymin) + 2 * s;                 // draws consumed before this sample
    gomath::Pcg rng;
    rng.hi = 0xDEADull ^ (uint64_t)x;                     // :632-634
    rng.lo = 0xBEEFull ^ (uint64_t)ymin;
    const gomath::PcgJump J = c_pcg_jump[q];
    gomath::affine128(rng.hi, rng.lo, J.mh, J.ml, J.ih, J.il);
    const double dx = gomath::pcg_float64(rng) - 0.5;     // dx is drawn before dy (:642-643)
    const double dy = gomath::pcg_float64(rng) - 0.5;
    if (GMLRT_PARITY_TU) {                                // the reference's operations, division for division (:644-649)
        const double u = ((double)x + dx) / (double)(sc.W - 1) * sc.vw - sc.vw / 2.0;
        const double v = ((double)y + dy) / (double)(sc.H - 1) * sc.vh - sc.vh / 2.0;
        o[0] = u; o[1] = -v; o[2] = 0.0;                  // screen point; eye = (0,0,-1)
        const double ex = o[0] - 0.0, ey = o[1] - 0.0, ez = o[2] - (-1.0);
        const double m = sqrt(ex * ex + ey * ey + ez * ez);
        d[0] = ex / m; d[1] = ey / m; d[2] = ez / m;
    } else {                                              // throughput mode: the same ray to a few ulp of double, one division instead of five
        const double u = ((double)x + dx) * sc.cam_su - 0.5 * sc.vw;
        const double v = ((double)y + dy) * sc.cam_sv - 0.5 * sc.vh;
        o[0] = u; o[1] = -v; o[2] = 0.0;
        const double im = 1.0 / sqrt(u * u + v * v + 1.0);
        d[0] = u * im; d[1] = -v * im; d[2] = im;
    }
}

// Robust primary-ray test (throughput mode only).  The north_star asks for ZERO hit/miss flips on
// primary rays against the float64 reference.  A float32 sphere test is trusted unless its
// discriminant is inside the rounding band |disc| < 2^-17 (hb^2 + |a c|) or the origin is that
// close to the surface; then - and for the non-quadric kinds always - the test is redone in
// double from the unrounded camera ray and the double matrices.
template <typename R>
__device__ __forceinline__ bool csg_object_hit(const DevScene<R>& sc, int i, const R* mats, const Ray<R>& w, R& t, int& code) {
    const int2 rg = sc.csg_range[i - sc.kind_begin[K_CSG]];
    return csg_hit<R>(sc.csg_prog + rg.x, rg.y, sc.kind, mats, w, t, code);
}

template <typename R>
static __device__ __noinline__ bool hit_f64(const DevScene<R>& sc, int kind, int i, const CamSrc& cam, int ri, float& t, int& code) {
    Ray<double> w, o;
    double o64[3], d64[3];
    camera_ray64(sc, cam, ri, o64, d64);
    w.ox = o64[0]; w.oy = o64[1]; w.oz = o64[2]; w.dx = d64[0]; w.dy = d64[1]; w.dz = d64[2];
    double td = 0.0;
    bool h;
    if (kind == K_CSG) {
        const int2 rg = sc.csg_range[i - sc.kind_begin[K_CSG]];
        h = csg_hit<double>(sc.csg_prog + rg.x, rg.y, sc.kind, sc.w2o64, w, td, code);
    } else {
        int face = 0;
        to_object(sc.w2o64 + (size_t)i * 12, w, o);
        h = hit_any_kind<double>(kind, o, td, face);
        code = i | (face << 28);
    }
    t = (float)td;
    return h;
}

// Robust float32 plane test for camera rays (throughput mode).  Plane.Intersect (:164-180) decides on two signs: the
// object-space height of the origin and of the direction (and |den| >= 1e-6).  Both are 3-4 term dot products; when
// each is outside its rounding band (2^-20 of the sum of the magnitudes: the rounded ray, the rounded matrix and the
// accumulation stay below 2^-22) float32 decides like float64, and T only needs float accuracy.  Camera rays start on
// the screen plane, far from any ground plane, so the double-precision path is the rare exception.  Returns false
// when the decision must be left to hit_f64.
template <typename R>
__device__ __forceinline__ bool plane_hit_robust32(const DevScene<R>& sc, int i, const Ray<R>& w, bool& hit, R& t) {
    const R* m = sc.w2o + (size_t)i * 12;
    const R a0 = m[4] * w.ox, a1 = m[5] * w.oy, a2 = m[6] * w.oz, a3 = m[7];
    const R b0 = m[4] * w.dx, b1 = m[5] * w.dy, b2 = m[6] * w.dz;
    const R oy = a0 + a1 + a2 + a3, dy = b0 + b1 + b2;
    const R eps = R(9.5367431640625e-7);                // 2^-20
    const R band_o = eps * (gm<R>::abs(a0) + gm<R>::abs(a1) + gm<R>::abs(a2) + gm<R>::abs(a3));
    const R band_d = eps * (gm<R>::abs(b0) + gm<R>::abs(b1) + gm<R>::abs(b2));
    if (!(gm<R>::abs(oy) > band_o && gm<R>::abs(gm<R>::abs(dy) - R(1e-6)) > band_d)) return false;
    hit = false;
    if (gm<R>::abs(dy) < R(1e-6)) return true;
    const R tt = (R(0) - oy) / dy;
    if (tt <= R(0)) return true;
    t = tt; hit = true;
    return true;
}

// the exact test of ray `w` against object i, exactly as <Type>.Intersect would decide it; `code` is the hit
// code stored in Level::hit: leaf | flip << 27 | face << 28 (for a CSG object the member leaf that was hit)
template <typename R, bool ROBUST>
__device__ __noinline__ bool exact_hit(const DevScene<R>& sc, int i, const Ray<R>& w, const CamSrc& cam, int ri, R& t, int& code) {
    const int kind = kind_of(sc.kind_begin, i);
    if constexpr (ROBUST) {
        if (kind == K_PLANE) { bool h = false; if (plane_hit_robust32(sc, i, w, h, t)) { code = i; return h; } }
        if (kind != K_SPHERE) { float tf; const bool h = hit_f64<R>(sc, kind, i, cam, ri, tf, code); t = (R)tf; return h; }
    }
    if (kind == K_CSG) return csg_object_hit<R>(sc, i, sc.w2o, w, t, code);
    Mat34<R> M;
    load_w2o(sc.w2o, i, M);
    Ray<R> o;
    to_object(M.m, w, o);
    if constexpr (ROBUST) {
        const R a = o.dx * o.dx + o.dy * o.dy + o.dz * o.dz;
        const R hb = o.ox * o.dx + o.oy * o.dy + o.oz * o.dz;
        const R oo = o.ox * o.ox + o.oy * o.oy + o.oz * o.oz;
        const R c = oo - R(1);
        const R p = a * c;
        const R disc = hb * hb - p;
        const R band = R(7.62939453125e-6) * (hb * hb + gm<R>::abs(p));     // 2^-17
        if (gm<R>::abs(disc) < band || gm<R>::abs(c) < R(7.62939453125e-6) * (oo + R(1))) {
            float tf; const bool h = hit_f64<R>(sc, K_SPHERE, i, cam, ri, tf, code); t = (R)tf; return h;
        }
        code = i;
        if (disc >= R(0)) {
            const R t0 = (-hb - gm<R>::sqrt(disc)) / a;
            if (t0 > R(0)) { t = t0; return true; }
        }
        return false;
    } else {
        int face = 0;
        const bool h = hit_any_kind<R>(kind, o, t, face);
        code = i | (face << 28);
        return h;
    }
}

// Stage `cnt` cull records of `table` starting at object `base` into `dst` (one thread).
template <typename R>
__device__ __forceinline__ void stage_cull(const R* table, int base, int cnt, R* dst, uint64_t* bar) {
    const uint32_t bytes = (uint32_t)(cnt * 4 * sizeof(R));
    mbar_expect_tx(bar, bytes);
    tma_load_1d(dst, table + (size_t)base * 4, bytes, bar);
}

template <typename R> struct BitsOf;
template <> struct BitsOf<float> {
    typedef unsigned int type;
    static __device__ __forceinline__ unsigned int of(float t) { return __float_as_uint(t); }
    static __device__ __forceinline__ float back(unsigned int b) { return __uint_as_float(b); }
};
template <> struct BitsOf<double> {
    typedef unsigned long long type;
    static __device__ __forceinline__ unsigned long long of(double t) { return (unsigned long long)__double_as_longlong(t); }
    static __device__ __forceinline__ double back(unsigned long long b) { return __longlong_as_double((long long)b); }
};

// the same window over an index list: objects idx[k0 .. k0 + cnt - 1] (indices into `sm`; cnt a multiple of 8, at most
// GROUP), 8 indices per 16-byte load.  A short list costs what it holds, not a whole window: the masks are shifted
// up afterwards so that test j still sits in bit 31 - j.
template <typename R, int FAMILY, int RPT>
__device__ __forceinline__ void cull_window_indirect(const R* sm, const uint16_t* idx, int k0, int cnt, const CullRay<R, FAMILY> (&cr)[RPT],
                                                     uint32_t (&lo)[RPT], uint32_t (&hi)[RPT]) {
#pragma unroll
    for (int r = 0; r < RPT; r++) { lo[r] = 0; hi[r] = 0; }
    const int n_lo = min(cnt, 32);
#pragma unroll 1
    for (int j0 = 0; j0 < n_lo; j0 += 8) {
        const uint4 pk = *reinterpret_cast<const uint4*>(idx + k0 + j0);
        const uint32_t w[4] = {pk.x, pk.y, pk.z, pk.w};
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const Cull4<R> L = load_cull(sm, (int)((w[j >> 1] >> (16 * (j & 1))) & 0xffffu));
#pragma unroll
            for (int r = 0; r < RPT; r++) lo[r] = push_sign(lo[r], cr[r].slack(L));
        }
    }
    if (n_lo < 32) {
#pragma unroll
        for (int r = 0; r < RPT; r++) lo[r] <<= 32 - n_lo;
        return;
    }
#pragma unroll 1
    for (int j0 = 32; j0 < cnt; j0 += 8) {
        const uint4 pk = *reinterpret_cast<const uint4*>(idx + k0 + j0);
        const uint32_t w[4] = {pk.x, pk.y, pk.z, pk.w};
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const Cull4<R> L = load_cull(sm, (int)((w[j >> 1] >> (16 * (j & 1))) & 0xffffu));
#pragma unroll
            for (int r = 0; r < RPT; r++) hi[r] = push_sign(hi[r], cr[r].slack(L));
        }
    }
    if (cnt > 32 && cnt < 64) {
#pragma unroll
        for (int r = 0; r < RPT; r++) hi[r] <<= 64 - cnt;
    }
}

// What the rays of one warp (RPT per lane) have in common, for the packet stage of the PIVOT / AXIS families:
//   PIVOT  a double cone around `axis` (unit) with half-angle theta that contains every ray's line through the pivot
//   AXIS   a disc (centre mu, mv; radius rho) in the plane orthogonal to the common direction that contains every ray
//   GENERAL (rays sorted by origin object and direction cell) a sphere (centre o, radius rho) that holds every origin
//          and a cone around `axis` with half-angle theta that holds every direction
template <typename R> struct WarpBound { R ax, ay, az, cos_t, sin_t; bool usable; R ox, oy, oz, rho; };

template <typename R>
__device__ __forceinline__ R warp_sum(R v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }
template <typename R>
__device__ __forceinline__ R warp_min(R v) { for (int o = 16; o > 0; o >>= 1) { const R u = __shfl_xor_sync(0xffffffffu, v, o); v = u < v ? u : v; } return v; }
template <typename R>
__device__ __forceinline__ R warp_max(R v) { for (int o = 16; o > 0; o >>= 1) { const R u = __shfl_xor_sync(0xffffffffu, v, o); v = u > v ? u : v; } return v; }

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }

template <typename R, int RPT>
__device__ __forceinline__ WarpBound<R> warp_bound(const CullRay<R, CULL_PIVOT> (&cr)[RPT], const bool (&active)[RPT]) {
    R sx = R(0), sy = R(0), sz = R(0);
#pragma unroll
    for (int r = 0; r < RPT; r++) if (active[r]) { sx += cr[r].dx; sy += cr[r].dy; sz += cr[r].dz; }
    sx = warp_sum(sx); sy = warp_sum(sy); sz = warp_sum(sz);
    const R len = gm<R>::sqrt(fma_(sx, sx, fma_(sy, sy, sz * sz)));
    WarpBound<R> b = {};
    b.usable = len > R(0);
    const R inv = b.usable ? R(1) / len : R(0);
    b.ax = sx * inv; b.ay = sy * inv; b.az = sz * inv;
    R c = R(1);
#pragma unroll
    for (int r = 0; r < RPT; r++) if (active[r]) { const R d = fma_(cr[r].dx, b.ax, fma_(cr[r].dy, b.ay, cr[r].dz * b.az)); c = d < c ? d : c; }
    c = warp_min(c) - R(1e-5);                          // a slightly wider cone absorbs the rounding of axis and dot products
    b.cos_t = c;
    b.sin_t = gm<R>::sqrt(R(1) - (c < R(1) ? c * c : R(1))) + R(1e-5);
    b.usable = b.usable && c > R(0.5);                   // a wide cone culls little: use the plain scan
    return b;
}
template <typename R, int RPT>
__device__ __forceinline__ WarpBound<R> warp_bound(const CullRay<R, CULL_AXIS> (&cr)[RPT], const bool (&active)[RPT]) {
    R su = R(0), sv = R(0), cnt = R(0);
#pragma unroll
    for (int r = 0; r < RPT; r++) if (active[r]) { su += cr[r].ou; sv += cr[r].ov; cnt += R(1); }
    su = warp_sum(su); sv = warp_sum(sv); cnt = warp_sum(cnt);
    WarpBound<R> b = {};
    b.usable = cnt > R(0);
    b.ax = b.usable ? su / cnt : R(0); b.ay = b.usable ? sv / cnt : R(0); b.az = R(0);
    R d2 = R(0);
#pragma unroll
    for (int r = 0; r < RPT; r++) if (active[r]) { const R du = cr[r].ou - b.ax, dv = cr[r].ov - b.ay; const R d = fma_(du, du, dv * dv); d2 = d > d2 ? d : d2; }
    d2 = warp_max(d2);
    b.cos_t = gm<R>::sqrt(d2) * R(1.0001) + R(1e-6) * (gm<R>::abs(b.ax) + gm<R>::abs(b.ay) + R(1));      // rho, widened for rounding
    b.sin_t = R(0);
    return b;
}
template <typename R, int RPT>
__device__ __forceinline__ WarpBound<R> warp_bound(const CullRay<R, CULL_GENERAL> (&cr)[RPT], const bool (&active)[RPT]) {
    R sx = R(0), sy = R(0), sz = R(0), qx = R(0), qy = R(0), qz = R(0), cnt = R(0);
#pragma unroll
    for (int r = 0; r < RPT; r++) if (active[r]) {
        sx += cr[r].dx; sy += cr[r].dy; sz += cr[r].dz;
        qx += cr[r].o2x; qy += cr[r].o2y; qz += cr[r].o2z; cnt += R(1);          // o2 = -2 o
    }
    sx = warp_sum(sx); sy = warp_sum(sy); sz = warp_sum(sz);
    qx = warp_sum(qx); qy = warp_sum(qy); qz = warp_sum(qz); cnt = warp_sum(cnt);
    const R len = gm<R>::sqrt(fma_(sx, sx, fma_(sy, sy, sz * sz)));
    WarpBound<R> b;
    b.usable = len > R(0) && cnt > R(0);
    const R inv = b.usable ? R(1) / len : R(0), icnt = b.usable ? R(-0.5) / cnt : R(0);
    b.ax = sx * inv; b.ay = sy * inv; b.az = sz * inv;
    b.ox = qx * icnt; b.oy = qy * icnt; b.oz = qz * icnt;
    R c = R(1), d2 = R(0);
#pragma unroll
    for (int r = 0; r < RPT; r++) if (active[r]) {
        const R d = fma_(cr[r].dx, b.ax, fma_(cr[r].dy, b.ay, cr[r].dz * b.az));
        c = d < c ? d : c;
        const R ex = fma_(R(-0.5), cr[r].o2x, -b.ox), ey = fma_(R(-0.5), cr[r].o2y, -b.oy), ez = fma_(R(-0.5), cr[r].o2z, -b.oz);
        const R e = fma_(ex, ex, fma_(ey, ey, ez * ez));
        d2 = e > d2 ? e : d2;
    }
    c = warp_min(c) - R(1e-5);
    d2 = warp_max(d2);
    b.cos_t = c;
    b.sin_t = gm<R>::sqrt(R(1) - (c < R(1) ? c * c : R(1))) + R(1e-5);
    b.rho = gm<R>::sqrt(d2) * R(1.0001) + R(1e-6) * (gm<R>::abs(b.ox) + gm<R>::abs(b.oy) + gm<R>::abs(b.oz) + R(1));
    b.usable = b.usable && c > R(0.2);                   // NaN rays make c NaN: not usable, the plain scan handles them
    return b;
}

// may object record L be hit by ANY ray inside the warp's bound?  (conservative)
template <typename R> __device__ __forceinline__ bool packet_survives(const Cull4<R>& L, const WarpBound<R>& b, CullRay<R, CULL_PIVOT>*) {
    const R pp = fma_(L.x, L.x, fma_(L.y, L.y, L.z * L.z));
    const R r2 = pp - L.k;                               // >= R^2 (the record's K carries the margins)
    if (!(r2 < pp)) return true;                         // pivot inside the sphere, unbounded object (K = -inf), NaN
    const R inv = R(1) / pp;
    const R s2 = r2 * inv;                               // sin^2 of the sphere's angular radius alpha seen from the pivot
    const R sa = gm<R>::sqrt(s2) * R(1.0001), ca = gm<R>::sqrt(R(1) - s2);
    const R rs = gm<R>::sqrt(inv);
    if (pp * rs * (R(1) - sa) > b.rho) return false;     // nearest point of the sphere is beyond the packet's reach from the pivot (shadow rays; rho = inf otherwise)
    const R c = gm<R>::abs(fma_(L.x, b.ax, fma_(L.y, b.ay, L.z * b.az))) * rs;     // |cos| of the angle to the axis
    return !(c < b.cos_t * ca - b.sin_t * sa - R(1e-5));      // angle <= theta + alpha
}
template <typename R> __device__ __forceinline__ bool packet_survives(const Cull4<R>& L, const WarpBound<R>& b, CullRay<R, CULL_AXIS>*) {
    const R du = L.x - b.ax, dv = L.y - b.ay;
    const R rr = gm<R>::sqrt(-L.z) * R(1.0001) + b.cos_t;      // sqrt(R^2 + margin) + rho; -L.z = +inf for unbounded objects
    if (L.k < b.sin_t) return false;                     // the object's front end (C.axis + R, record word 3) is behind every origin of the packet
    return !(fma_(du, du, dv * dv) > rr * rr);
}
// GENERAL record (C, K = |C|^2 - R^2 - margin).  A ray from o (|o - oc| <= rho) that hits the sphere (C, R) at T > 0 is,
// moved to oc, a ray that hits a sphere of the same radius whose centre is within rho of C: so test the sphere
// (C, R + rho) against the cone of directions from oc - forward only, the reference accepts T > 0 alone.
template <typename R> __device__ __forceinline__ bool packet_survives(const Cull4<R>& L, const WarpBound<R>& b, CullRay<R, CULL_GENERAL>*) {
    const R cc = fma_(L.x, L.x, fma_(L.y, L.y, L.z * L.z));
    const R r2 = fma_(cc, R(9.5367431640625e-7), cc - L.k);       // R^2 + margins + 2^-20 |C|^2 for the cancellation in |C|^2 - K
    const R rr = gm<R>::sqrt(r2) * R(1.0001) + b.rho;
    const R rr2 = rr * rr;
    const R px = L.x - b.ox, py = L.y - b.oy, pz = L.z - b.oz;
    const R pp = fma_(px, px, fma_(py, py, pz * pz));
    if (!(rr2 < pp)) return true;                        // an origin may be inside the sphere; unbounded object; NaN / padding
    const R inv = R(1) / pp;
    const R s2 = rr2 * inv;
    const R sa = gm<R>::sqrt(s2) * R(1.0001), ca = gm<R>::sqrt(R(1) - s2);
    const R c = fma_(px, b.ax, fma_(py, b.ay, pz * b.az)) * gm<R>::sqrt(inv);      // signed: behind the packet = miss
    return !(c < b.cos_t * ca - b.sin_t * sa - R(1e-5));      // angle(C - oc, axis) <= theta + alpha
}

constexpr int PCHUNK = 1024;                // objects per packet pass = survivor list capacity per warp (uint16 indices)
constexpr size_t SURV_BYTES = (size_t)(SCAN_THREADS / 32 + 1) * PCHUNK * sizeof(uint16_t);    // one list per warp + the CTA's

// The bound of a whole CTA's rays = the union of its warps' bounds (published in shared memory by lane 0 of each warp;
// every thread evaluates the same union).  Used for a first packet pass shared by the CTA: 256 threads x 4 objects
// instead of 32 lanes x 32, after which a warp only looks at what the CTA kept.
constexpr int WB_WORDS = 12;      // ax ay az cos sin ox oy oz rho usable active
template <typename R> __device__ __forceinline__ void publish_bound(R* slot, const WarpBound<R>& b, bool any_active) {
    slot[0] = b.ax; slot[1] = b.ay; slot[2] = b.az; slot[3] = b.cos_t; slot[4] = b.sin_t;
    slot[5] = b.ox; slot[6] = b.oy; slot[7] = b.oz; slot[8] = b.rho;
    slot[9] = b.usable ? R(1) : R(0); slot[10] = any_active ? R(1) : R(0);
}
// Evaluated by every warp on its own: lane w (mod 8) holds warp w's published bound, sums / minima / maxima go over
// groups of 8 lanes with xor-shuffles, so the union costs a few dozen instructions per warp instead of 8x that per thread.
template <typename R> __device__ __forceinline__ R oct_sum(R v) { for (int o = SCAN_WARPS / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }
template <typename R> __device__ __forceinline__ R oct_min(R v) { for (int o = SCAN_WARPS / 2; o > 0; o >>= 1) { const R u = __shfl_xor_sync(0xffffffffu, v, o); v = u < v ? u : v; } return v; }
template <typename R> __device__ __forceinline__ R oct_max(R v) { for (int o = SCAN_WARPS / 2; o > 0; o >>= 1) { const R u = __shfl_xor_sync(0xffffffffu, v, o); v = u > v ? u : v; } return v; }

template <typename R, int FAMILY>
__device__ __forceinline__ WarpBound<R> cta_bound(const R (*wbs)[WB_WORDS]) {
    static_assert((SCAN_WARPS & (SCAN_WARPS - 1)) == 0 && SCAN_WARPS <= 32, "one lane of each group of SCAN_WARPS lanes per warp of the CTA");
    const R* b = wbs[threadIdx.x & (SCAN_WARPS - 1)];
    const bool act = b[10] != R(0);                         // false: a warp past the end of the queue
    WarpBound<R> c = {};
    const R bad = oct_sum<R>(act && b[9] == R(0) ? R(1) : R(0));
    const R cnt = oct_sum<R>(act ? R(1) : R(0));
    c.usable = bad == R(0) && cnt > R(0);
    if (!c.usable) return c;                                // uniform across the warp (and the CTA: every warp sees the same slots)
    const R sx = oct_sum<R>(act ? b[0] : R(0)), sy = oct_sum<R>(act ? b[1] : R(0)), sz = oct_sum<R>(act ? b[2] : R(0));
    if (FAMILY == CULL_AXIS) {                               // discs in the plane orthogonal to the common direction
        c.ax = sx / cnt; c.ay = sy / cnt;
        const R du = b[0] - c.ax, dv = b[1] - c.ay;
        const R rho = oct_max<R>(act ? gm<R>::sqrt(fma_(du, du, dv * dv)) * R(1.0001) + b[3] : R(0));
        c.cos_t = rho * R(1.0001) + R(1e-6) * (gm<R>::abs(c.ax) + gm<R>::abs(c.ay) + R(1));
        c.sin_t = oct_min<R>(act ? b[4] : gm<R>::inf());      // rear-most origin along the axis over the CTA
        return c;
    }
    const R qx = oct_sum<R>(act ? b[5] : R(0)), qy = oct_sum<R>(act ? b[6] : R(0)), qz = oct_sum<R>(act ? b[7] : R(0));
    const R len = gm<R>::sqrt(fma_(sx, sx, fma_(sy, sy, sz * sz)));
    if (!(len > R(0))) { c.usable = false; return c; }
    const R inv = R(1) / len;
    c.ax = sx * inv; c.ay = sy * inv; c.az = sz * inv;
    c.ox = qx / cnt; c.oy = qy / cnt; c.oz = qz / cnt;
    R cw = fma_(b[0], c.ax, fma_(b[1], c.ay, b[2] * c.az));           // cos of the angle phi between the warp's axis and the CTA's
    const R neg = oct_sum<R>(act && !(cw > R(0)) ? R(1) : R(0));         // phi + theta must stay below pi for the sum formula
    cw = cw < R(1) ? cw : R(1);
    const R sw = gm<R>::sqrt(R(1) - cw * cw) + R(1e-5);
    R cmin = oct_min<R>(act ? cw * b[3] - sw * b[4] : R(1));             // cos(phi + theta_w)
    const R ex = b[5] - c.ox, ey = b[6] - c.oy, ez = b[7] - c.oz;
    const R rho = oct_max<R>(act ? gm<R>::sqrt(fma_(ex, ex, fma_(ey, ey, ez * ez))) * R(1.0001) + b[8] : R(0));
    cmin -= R(1e-5);
    c.cos_t = cmin;
    c.sin_t = gm<R>::sqrt(R(1) - (cmin < R(1) ? cmin * cmin : R(1))) + R(1e-5);
    c.rho = rho * R(1.0001) + R(1e-6) * (gm<R>::abs(c.ox) + gm<R>::abs(c.oy) + gm<R>::abs(c.oz) + R(1));
    c.usable = neg == R(0) && cmin > R(0.2);
    return c;
}

// ------------------------------------------------------------------ scan = cull kernel + exact kernels
// Phase 1, k_cull: persistent CTAs over blocks of SCAN_THREADS * RPT rays; the cull table is staged once per
// CTA when it fits ("resident"), else streamed through two buffers per block of rays.  Each thread carries RPT
// rays in registers, culls a window of GROUP objects into survivor masks, and the warp appends every set bit to
// the launch's candidate queue in HBM as a (ray, object) pair - one atomic per warp and window.
// Phase 2, k_exact_*: one thread per candidate runs the reference's own test - every lane busy, instead of the
// few lanes of a warp that happen to have a survivor in the same window - and folds the result into the ray's
// record with atomics.
//
// The queue is sized from the scene (Renderer::plan_tiles); if a launch overflows it, `Counters::overflow`
// is raised, the frame is discarded and traced again with smaller tiles (gmlrt_trace).
struct CandQueue {
    int* ray;        // [cap] ray index within the launch's ray source
    int* obj;        // [cap] object index; k_exact_closest overwrites it with the hit code (or -1)
    void* t;         // [cap] R: T of the hit (closest only)
    int* count;      // [0] entries appended by the current launch, [1] blocks of rays handed out beyond the first wave
    int cap;
};

// entries of a finished cull launch.  After an overflow the queue has holes (a warp that did not fit reserved but never
// wrote) whose stale (ray, object) words may belong to a larger level: the frame is discarded anyway, so skip them all.
__device__ __forceinline__ int cand_count(const CandQueue& cq, const Counters* ctr) {
    if (ctr->overflow) return 0;
    return (int)min(*reinterpret_cast<const unsigned*>(cq.count), (unsigned)cq.cap);
}

template <typename R> struct ClosestSource {                 // rays of one recursion level
    Level<R> lv;                                              // (levels >= 1 are stored sorted by origin object and direction cell)
    __device__ __forceinline__ int id(int i) const { return i; }
    __device__ __forceinline__ void load(int i, Ray<R>& out) const { load_ray(lv.ray, i, out); }
    static constexpr bool is_shadow = false;
    __device__ __forceinline__ void init(int i) const {       // no hit yet (see fold_hit)
        lv.hit[i] = -1;
        if constexpr (sizeof(R) == 4) reinterpret_cast<unsigned long long*>(lv.order)[i] = ~0ull;
        else {
            reinterpret_cast<unsigned long long*>(lv.t)[i] = ~0ull;
            lv.order[i] = 0x7fffffff;
        }
    }
};

template <typename R> struct ShadowSource {                  // shadow rays of one light: slots [base, base + n)
    ShadowQueue<R> sq; int base;
    __device__ __forceinline__ int id(int i) const { return sq.perm[base + i]; }          // i-th ray in by-object order -> slot in the segment
    __device__ __forceinline__ void load(int i, Ray<R>& out) const {
        const R* rec = sq.q + (size_t)(base + id(i)) * 8;
        out.ox = rec[0]; out.oy = rec[1]; out.oz = rec[2]; out.dx = rec[3]; out.dy = rec[4]; out.dz = rec[5];
    }
    __device__ __forceinline__ void init(int i) const { sq.first[base + id(i)] = 0x7fffffff; }
    static constexpr bool is_shadow = true;
    // How far from the light an occluder of ray i can be: inShadow accepts T * |d_in| < dist (:424), T > 0, on a ray
    // that starts `dist` from the light and points at it - so points up to dist away on the near side and, when the
    // incoming direction is shorter than 1, up to dist / |d_in| - dist beyond the light.
    __device__ __forceinline__ R reach(int i) const {
        const R* rec = sq.q + (size_t)(base + id(i)) * 8;
        const R dist = rec[6], beyond = rec[6] / rec[7] - rec[6];
        return beyond > dist ? beyond : dist;
    }
};

template <typename R, int FAMILY, int RPT, typename Source>
__global__ void __launch_bounds__(SCAN_THREADS, (sizeof(R) == 4 ? 3 : 2) * 256 / SCAN_THREADS)
k_cull(const __grid_constant__ Source src, const int* n_ptr, const R* table, const __grid_constant__ CullFrame<R> frame,
       int n_cull, const __grid_constant__ CandQueue cq, Counters* ctr, int chunk_objs, int resident_flags) {
    constexpr int RAYS_PER_BLOCK = SCAN_THREADS * RPT;
    const int resident = resident_flags & 1;
    const bool cta_pass = (resident_flags & 2) != 0;           // host switch: try the CTA-wide first packet pass
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t bar[2];
    const int tid = threadIdx.x, lane = tid & 31;
    uint16_t* const csurv = reinterpret_cast<uint16_t*>(smem_raw);                              // what the CTA's bound keeps (packet stage, first pass)
    uint16_t* const surv = csurv + ((tid >> 5) + 1) * PCHUNK;                                  // what this warp's bound keeps of it
    __shared__ R s_wb[SCAN_THREADS / 32][WB_WORDS];
    __shared__ int s_cn[2];                                                                    // entries of the CTA's list, by pass parity
    int pass_no = 0;
    R* const tab = reinterpret_cast<R*>(smem_raw + SURV_BYTES);   // buffer b starts at tab + b * chunk_objs * 4
    const int n = *n_ptr;
    const int nblocks = (n + RAYS_PER_BLOCK - 1) / RAYS_PER_BLOCK;
    if ((int)blockIdx.x >= nblocks) return;
    const int NP = n_cull;                                     // padded to a multiple of GROUP
    const int nchunks = (NP + chunk_objs - 1) / chunk_objs;
    if (tid == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_fence_init(); s_cn[0] = 0; s_cn[1] = 0; }
    __syncthreads();
    uint32_t phase[2] = {0, 0};
    if (resident && NP > 0) {
        if (tid == 0) stage_cull(table, 0, NP, tab, &bar[0]);
        mbar_wait(&bar[0], 0);
    }
    // blocks of rays are handed out dynamically (their cost varies with what survives the packet stage): a CTA starts
    // on block blockIdx.x and fetches the following ones from a device counter, one iteration ahead
    __shared__ int next_blk[2];
    int iter = 0;
    for (int blk = blockIdx.x; blk < nblocks; blk = next_blk[iter++ & 1]) {
        if (tid == 0) next_blk[iter & 1] = (int)gridDim.x + atomicAdd(cq.count + 1, 1);
        const int ray0 = blk * RAYS_PER_BLOCK;
        CullRay<R, FAMILY> cr[RPT];
        bool active[RPT];
        int rid[RPT];                                          // what the exact kernels index the ray source with
        R far_reach = R(0), rear = gm<R>::inf();               // shadow packets: farthest reach from the light / rear-most origin along the axis
#pragma unroll
        for (int r = 0; r < RPT; r++) {
            const int i = ray0 + ((tid >> 5) * RPT + r) * 32 + lane;      // a warp owns 32 * RPT consecutive rays: a narrow packet
            Ray<R> w = {R(0), R(0), R(0), R(0), R(0), R(1)};
            active[r] = i < n;
            rid[r] = active[r] ? src.id(i) : 0;
            if (active[r]) { src.load(i, w); src.init(i); }
            cr[r].init(w, frame);
            if constexpr (Source::is_shadow && FAMILY == CULL_PIVOT) { if (active[r]) { const R q = src.reach(i); far_reach = !(q <= far_reach) ? (q == q ? q : gm<R>::inf()) : far_reach; } }
            if constexpr (Source::is_shadow && FAMILY == CULL_AXIS) {
                if (active[r]) { const R q = fma_(w.ox, frame.axis[0], fma_(w.oy, frame.axis[1], w.oz * frame.axis[2])); rear = !(q >= rear) ? (q == q ? q : -gm<R>::inf()) : rear; }
            }
        }
        WarpBound<R> wb = warp_bound<R, RPT>(cr, active);
        // Shadow rays stop at the light: an object farther from a point light than every origin of the packet (plus
        // what inShadow's T * |d_in| < dist lets through beyond the light), or wholly behind every origin along a
        // directional light's rays, cannot occlude any of them.
        if constexpr (FAMILY == CULL_PIVOT) wb.rho = Source::is_shadow ? warp_max(far_reach) * R(1.001) + R(1e-3) : gm<R>::inf();
        if constexpr (FAMILY == CULL_AXIS) { const R q = Source::is_shadow ? warp_min(rear) : -gm<R>::inf(); wb.sin_t = q - R(1e-3) * (gm<R>::abs(q) + R(1)); }
        int n_active = 0;
#pragma unroll
        for (int r = 0; r < RPT; r++) n_active += active[r] ? 1 : 0;
        WarpBound<R> cb = {};
        if (cta_pass) {
            if (lane == 0) publish_bound<R>(s_wb[tid >> 5], wb, ray0 + (tid >> 5) * RPT * 32 < n);
            __syncthreads();
            cb = cta_bound<R, FAMILY>(s_wb);
        }
        unsigned long long fine = 0, packet = 0;           // this lane's executed tests
#ifdef GMLRT_DEBUG_PACKETS
        if (lane == 0 && FAMILY != CULL_GENERAL) { atomicAdd(&ctr->dbg[0 + 4 * (FAMILY == CULL_AXIS)], 1ull); if (wb.usable) atomicAdd(&ctr->dbg[1 + 4 * (FAMILY == CULL_AXIS)], 1ull); }
#endif
        if (!resident) {
            __syncthreads();                                   // (re)staging must not overtake readers of the previous block of rays
            if (tid == 0) {
                stage_cull(table, 0, min(chunk_objs, NP), tab, &bar[0]);
                if (nchunks > 1) stage_cull(table, chunk_objs, min(chunk_objs, NP - chunk_objs), tab + (size_t)chunk_objs * 4, &bar[1]);
            }
        }
        for (int c = 0; c < (resident ? 1 : nchunks); c++) {   // a resident table is one chunk that is already staged
            const int b = c & 1;
            if (!resident) { mbar_wait(&bar[b], phase[b]); phase[b] ^= 1; }
            const int base = resident ? 0 : c * chunk_objs;
            const R* sm = resident ? tab : tab + (size_t)b * chunk_objs * 4;
            const int end = resident ? NP : base + min(chunk_objs, NP - base);
            // append the survivors in (lo, hi)[r] to the candidate queue; bit 31 - j of lo = object id_of(j), of hi = id_of(32 + j)
            auto append = [&](uint32_t (&lo)[RPT], uint32_t (&hi)[RPT], auto id_of) __attribute__((always_inline)) {
                int mine = 0;
#pragma unroll
                for (int r = 0; r < RPT; r++) {
                    if (!active[r]) { lo[r] = 0; hi[r] = 0; }
                    mine += __popc(lo[r]) + __popc(hi[r]);
                }
                if (!__any_sync(0xffffffffu, mine != 0)) return;            // nothing survived in this warp
                // warp-aggregated append: inclusive scan of the lanes' counts, one atomic for the warp
                int incl = mine;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += v; }
                const int total = __shfl_sync(0xffffffffu, incl, 31);
                // The counter is unsigned and stops growing once the queue is full: a lattice seen along an axis can leave
                // dozens of survivors per ray, and a signed counter that kept counting past 2^31 would turn `at` negative
                // and the stores below into out-of-bounds writes instead of the overflow-and-retrace path.
                unsigned at = 0;
                if (lane == 31) {
                    if (*reinterpret_cast<volatile unsigned*>(cq.count) >= (unsigned)cq.cap) at = 0xffffffffu;
                    else at = atomicAdd(reinterpret_cast<unsigned*>(cq.count), (unsigned)total);
                }
                at = __shfl_sync(0xffffffffu, at, 31);
                if (at == 0xffffffffu || at + (unsigned)total > (unsigned)cq.cap) { if (lane == 31) ctr->overflow = 1; return; }     // queue full: the host retries with smaller tiles
                at += (unsigned)(incl - mine);
#pragma unroll
                for (int r = 0; r < RPT; r++) {
                    const int i = rid[r];
                    uint32_t m = lo[r];
                    while (m) {
                        const int j = __clz(m);
                        m &= ~(0x80000000u >> j);
                        cq.ray[at] = i; cq.obj[at] = id_of(j); at++;
                    }
                    m = hi[r];
                    while (m) {
                        const int j = __clz(m);
                        m &= ~(0x80000000u >> j);
                        cq.ray[at] = i; cq.obj[at] = id_of(32 + j); at++;
                    }
                }
            };
            if (cb.usable || wb.usable) {
                // packet stage.  First pass (when the CTA's rays are coherent as a whole): the 256 threads test the pass's
                // objects against the CTA's bound, 32 per warp and step, and compact what it keeps into the CTA's list.
                // Second pass: the lanes of each warp test that list (or, without a CTA bound, every object) against the
                // warp's own bound and compact into the warp's list; only those objects are tested per ray.
                const bool warp_has_rays = __any_sync(0xffffffffu, n_active != 0);
                const unsigned lt = (1u << lane) - 1u;
                for (int p0 = base; p0 < end; p0 += PCHUNK) {
                    const int pn = min(PCHUNK, end - p0);
                    int cn = pn;
                    if (cb.usable) {                                                 // uniform across the CTA
                        int* const cnt = &s_cn[pass_no & 1];
                        for (int o = tid & ~31; o < pn; o += SCAN_THREADS) {
                            const int k = p0 - base + o + lane;
                            const bool in = o + lane < pn;
                            packet += in ? 1 : 0;
                            const bool sv = in && packet_survives<R>(load_cull(sm, k), cb, (CullRay<R, FAMILY>*)nullptr);
                            const unsigned bal = __ballot_sync(0xffffffffu, sv);
                            int at = 0;
                            if (lane == 0 && bal) at = atomicAdd(cnt, __popc(bal));
                            at = __shfl_sync(0xffffffffu, at, 0);
                            if (sv) csurv[at + __popc(bal & lt)] = (uint16_t)k;
                        }
                        __syncthreads();
                        cn = *cnt;
                        if (tid == 0) s_cn[(pass_no + 1) & 1] = 0;                    // free since the end of the previous pass
                        pass_no++;
                    }
                    int ns = 0;
                    if (warp_has_rays && wb.usable) {
                        for (int o = 0; o < cn; o += 32) {
                            const bool in = o + lane < cn;
                            const int k = cb.usable ? (in ? (int)csurv[o + lane] : 0) : p0 - base + o + lane;
                            packet += in ? 1 : 0;
                            const bool sv = in && packet_survives<R>(load_cull(sm, k), wb, (CullRay<R, FAMILY>*)nullptr);
                            const unsigned bal = __ballot_sync(0xffffffffu, sv);
                            if (sv) surv[ns + __popc(bal & lt)] = (uint16_t)k;
                            ns += __popc(bal);
                        }
                    } else if (warp_has_rays) {                                      // (a warp bound can only be unusable here if the CTA's is)
                        for (int o = lane; o < cn; o += 32) surv[o] = cb.usable ? csurv[o] : (uint16_t)(p0 - base + o);
                        ns = cn;
                    }
#ifdef GMLRT_DEBUG_PACKETS
                    if (lane == 0) { atomicAdd(&ctr->dbg[2 + 4 * (FAMILY == CULL_AXIS)], (unsigned long long)pn); atomicAdd(&ctr->dbg[3 + 4 * (FAMILY == CULL_AXIS)], (unsigned long long)ns); }
#endif
                    const int padded = (ns + 7) & ~7;                                   // whole 16-byte loads of 8 indices
                    if (ns + lane < padded) surv[ns + lane] = (uint16_t)(p0 - base);     // any valid record; masked off below
                    __syncwarp();
                    for (int k0 = 0; k0 < padded; k0 += GROUP) {
                        uint32_t lo[RPT], hi[RPT];
                        cull_window_indirect<R, FAMILY, RPT>(sm, surv, k0, min(GROUP, padded - k0), cr, lo, hi);
                        const int valid = ns - k0;                                 // entries of this window that are real
                        fine += (unsigned long long)(n_active * min(valid, GROUP));
                        const uint32_t mlo = valid >= 32 ? 0xffffffffu : ~(0xffffffffu >> valid);
                        const uint32_t mhi = valid >= 64 ? 0xffffffffu : (valid <= 32 ? 0u : ~(0xffffffffu >> (valid - 32)));
#pragma unroll
                        for (int r = 0; r < RPT; r++) { lo[r] &= mlo; hi[r] &= mhi; }
                        append(lo, hi, [&](int j) { return base + (int)surv[k0 + j]; });
                    }
                    if (cb.usable) __syncthreads();                                  // the CTA's list is rewritten by the next pass
                    else __syncwarp();                                               // the warp's list likewise
                }
            } else {
                for (int first = base; first < end; first += GROUP) {      // one window of GROUP objects
                    uint32_t lo[RPT], hi[RPT];
                    cull_window<R, FAMILY, RPT>(sm, first - base, cr, lo, hi);
                    fine += (unsigned long long)(n_active * GROUP);
                    append(lo, hi, [&](int j) { return first + j; });
                }
            }
            if (!resident) {
                __syncthreads();                               // everyone is done reading buffer b
                if (tid == 0 && c + 2 < nchunks) {
                    const int nb = (c + 2) * chunk_objs;
                    stage_cull(table, nb, min(chunk_objs, NP - nb), tab + (size_t)b * chunk_objs * 4, &bar[b]);
                }
            }
        }
        fine = warp_sum_u64(fine); packet = warp_sum_u64(packet);
        if (lane == 0) { atomicAdd(&ctr->cull_tests[FAMILY], fine); if (packet) atomicAdd(&ctr->packet_tests, packet); }
        __syncthreads();                                       // next_blk[iter & 1] is visible; slot (iter + 1) & 1 is free again
    }
}

__device__ __forceinline__ void warp_add(unsigned long long* dst, unsigned long long v) {
    for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(dst, v);
}

constexpr int EXACT_THREADS = 128;

// The exact kernels come in three flavours so that the common case runs at full occupancy: ALL kinds through
// exact_hit (a call; 96-128 registers because of the CSG span lists), SPHERES only with the test inlined (the
// candidates of other kinds are skipped) and REST = every kind but spheres, through exact_hit.  A scene of spheres
// launches SPHERES alone, a mixed scene SPHERES then REST over the same candidate list.
enum ExactKinds { EXACT_ALL = 0, EXACT_SPHERES = 1, EXACT_REST = 2, EXACT_SIMPLE = 3, EXACT_CSG = 4 };
// SIMPLE = the other plain kinds (plane, cube, cylinder, cone) with their tests inlined, CSG = the span-list objects only:
// outside the double-precision camera-ray pass (which keeps REST = everything but spheres) a mixed scene runs SPHERES,
// SIMPLE and CSG over the same candidate list, so that only CSG candidates pay for the 96-128 registers of csg_hit.

template <typename R, int KINDS>
__device__ __forceinline__ bool exact_hit_kinds(const DevScene<R>& sc, int obj, const Ray<R>& w, R& t, int& code) {
    if constexpr (KINDS == EXACT_SPHERES) {
        Mat34<R> M;
        load_w2o(sc.w2o, obj, M);
        Ray<R> o;
        to_object(M.m, w, o);
        int face = 0;
        code = obj;
        return hit_sphere<R>(o, t, face);
    } else if constexpr (KINDS == EXACT_SIMPLE) {
        Mat34<R> M;
        load_w2o(sc.w2o, obj, M);
        Ray<R> o;
        to_object(M.m, w, o);
        int face = 0;
        const bool h = hit_any_kind<R>(kind_of(sc.kind_begin, obj), o, t, face);
        code = obj | (face << 28);
        return h;
    } else if constexpr (KINDS == EXACT_CSG) {
        return csg_object_hit<R>(sc, obj, sc.w2o, w, t, code);
    } else {
        return exact_hit<R, false>(sc, obj, w, CamSrc{}, 0, t, code);
    }
}
// does this flavour own candidate object `obj`?  (objects are sorted by kind: spheres first)
template <int KINDS> __device__ __forceinline__ bool exact_owns(int obj, int n_spheres, int csg_begin) {
    if (KINDS == EXACT_ALL) return true;
    if (KINDS == EXACT_SPHERES) return obj < n_spheres;
    if (KINDS == EXACT_SIMPLE) return obj >= n_spheres && obj < csg_begin;
    if (KINDS == EXACT_CSG) return obj >= csg_begin;
    return obj >= n_spheres;
}

// closestHit (raytracer.go:469-483): strictly smaller T wins, equal T keeps the object that comes first in scan
// order.  Three passes over the candidates: (1) exact test, T of a hit into the queue and atomicMin into the ray's
// record (hits have T > 0, so the bit pattern orders like the value); (2) among the hits that have the ray's
// smallest T, atomicMin of the scan index; (3) the winner writes its hit code.
// Folding a hit into the ray's record.  float32: one 64-bit atomicMin of (bits of T) << 32 | scan index does the
// whole rule at once - smaller T first, then the smaller scan index - so a single pick pass follows.  float64: T alone
// fills the 64 bits, so the scan index is folded by a second pass over the candidates (k_closest_order).
template <typename R> __device__ __forceinline__ unsigned long long hit_key32(const DevScene<R>& sc, float t, int code) {
    // pickless (DevScene::pickless): the low word is the hit code itself, rearranged as leaf << 4 | flip << 3 | face so
    // that a tie in T goes to the smaller device index first - the key IS the hit
    if (sc.pickless == 1) return ((unsigned long long)__float_as_uint(t) << 32) | (unsigned)(code & HIT_LEAF_MASK);      // spheres and planes: face 0, no flip
    if (sc.pickless) return ((unsigned long long)__float_as_uint(t) << 32) | (unsigned)(((code & HIT_LEAF_MASK) << 4) | (((code >> 27) & 1) << 3) | ((code >> 28) & 7));
    return ((unsigned long long)__float_as_uint(t) << 32) | (unsigned)sc.scan_of[code & HIT_LEAF_MASK];
}
template <typename R> __device__ __forceinline__ void fold_hit(const DevScene<R>& sc, const Level<R>& lv, int ri, R t, int code) {
    if constexpr (sizeof(R) == 4) atomicMin(&reinterpret_cast<unsigned long long*>(lv.order)[ri], hit_key32(sc, t, code));
    else atomicMin(&reinterpret_cast<unsigned long long*>(lv.t)[ri], BitsOf<R>::of(t));
}

constexpr int CAND_DEFER = 0x40000000;   // robust camera-ray pass: a sphere candidate left to the double-precision pass

template <typename R, bool ROBUST, bool COUNT, int KINDS>
__global__ void __launch_bounds__(EXACT_THREADS) k_exact_closest(const __grid_constant__ DevScene<R> sc, const __grid_constant__ Level<R> lv,
                                                                const __grid_constant__ CandQueue cq, Counters* ctr, const __grid_constant__ CamSrc cam) {
    static_assert(!ROBUST || KINDS != EXACT_ALL, "the robust camera-ray test runs as SPHERES (float32, inline) then REST (double)");
    const int count = cand_count(cq, ctr);
    const int n_spheres = sc.kind_begin[K_SPHERE + 1];
    R* ct = reinterpret_cast<R*>(cq.t);
    unsigned long long n_exact = 0, n_flops = 0;
    // The double-precision pass owns few candidates (the deferred spheres, the non-sphere kinds): it reads the object
    // words four at a time, so that skipping what it does not own costs a quarter of the round trips, and fetches the
    // ray index only for a candidate it keeps.
    constexpr int U = (ROBUST && KINDS == EXACT_REST) ? 4 : 1;
    for (int k0 = (blockIdx.x * blockDim.x + threadIdx.x) * U; k0 < count; k0 += gridDim.x * blockDim.x * U) {
        int objs[U];
        if (U == 4 && k0 + 4 <= count) { const int4 v = *reinterpret_cast<const int4*>(cq.obj + k0); objs[0] = v.x; objs[1 % U] = v.y; objs[2 % U] = v.z; objs[3 % U] = v.w; }
        else { for (int u = 0; u < U; u++) objs[u] = k0 + u < count ? cq.obj[k0 + u] : -1; }
#pragma unroll
        for (int u = 0; u < U; u++) {
        const int k = k0 + u;
        if (k >= count) break;
        int obj = objs[u];
        if (KINDS == EXACT_REST && obj < 0) continue;          // a sphere candidate the SPHERES pass has already resolved to a miss
        bool deferred = false;
        if (ROBUST && KINDS == EXACT_REST) { deferred = (obj & CAND_DEFER) != 0; obj &= ~CAND_DEFER; }
        if (KINDS != EXACT_ALL && obj < sc.n_leaves && !deferred && !exact_owns<KINDS>(obj, n_spheres, sc.kind_begin[K_CSG])) continue;
        const int ri = cq.ray[k];
        int code = -1; R t = R(0);
        if (obj < sc.n_leaves) {                               // a padding record can only be reached by a NaN ray
            if (COUNT && !deferred) { n_exact++; n_flops += (unsigned)sc.obj_flops[obj]; }
            bool h;
            if constexpr (ROBUST && KINDS == EXACT_SPHERES) {
                // float32 sphere test, trusted unless the discriminant is inside its rounding band or the origin is that
                // close to the surface (see exact_hit<ROBUST>); those candidates go to the REST pass, in double
                Ray<R> w, o;
                load_ray(lv.ray, ri, w);
                Mat34<R> M;
                load_w2o(sc.w2o, obj, M);
                to_object(M.m, w, o);
                // The textbook discriminant hb^2 - a c cancels two numbers of the size |o|^2 (a camera ray starts ~100 sphere
                // radii away in object space), so its float32 rounding band is ~ +-7 % of the silhouette radius and one
                // candidate in eight had to be redone in double.  The same quantity through the point of closest approach
                // l = o + tc d,  disc = a (1 - |l|^2),  t0 = tc - sqrt((1 - |l|^2) / a)  has no such cancellation: the error of
                // |l|^2 is ~ 20 eps |o| |l| (rounded matrix, rounded ray, the accumulation), the band 2^-17 (|o| |l| + 1) = 128 eps
                // leaves a factor 6, and a deferral becomes the rare case it was meant to be.
                const R a = o.dx * o.dx + o.dy * o.dy + o.dz * o.dz;
                const R hb = o.ox * o.dx + o.oy * o.dy + o.oz * o.dz;
                const R oo = o.ox * o.ox + o.oy * o.oy + o.oz * o.oz;
                const R c = oo - R(1);
                const R tc = -hb / a;
                const R lx = fma_(tc, o.dx, o.ox), ly = fma_(tc, o.dy, o.oy), lz = fma_(tc, o.dz, o.oz);
                const R pp = lx * lx + ly * ly + lz * lz;
                const R band = R(7.62939453125e-6) * (gm<R>::sqrt(oo * pp) + R(1));       // 2^-17
                if (!(gm<R>::abs(R(1) - pp) >= band) || gm<R>::abs(c) < R(7.62939453125e-6) * (oo + R(1))) { cq.obj[k] = obj | CAND_DEFER; continue; }
                code = obj; h = false;
                if (pp < R(1)) { const R t0 = tc - gm<R>::sqrt((R(1) - pp) / a); if (t0 > R(0)) { t = t0; h = true; } }
            } else if constexpr (ROBUST) {                     // REST: every non-sphere kind, and the deferred spheres, in double
                const int kind = deferred ? (int)K_SPHERE : kind_of(sc.kind_begin, obj);
                bool decided = false;
                h = false;
                if (kind == K_PLANE) {                         // float32 unless a sign is inside its rounding band
                    Ray<R> w;
                    load_ray(lv.ray, ri, w);
                    decided = plane_hit_robust32(sc, obj, w, h, t);
                    code = obj;
                }
                if (!decided) {
                    float tf = 0.f;
                    h = hit_f64<R>(sc, kind, obj, cam, ri, tf, code);
                    t = (R)tf;
                }
            } else {
                Ray<R> w;
                load_ray(lv.ray, ri, w);
                h = exact_hit_kinds<R, KINDS>(sc, obj, w, t, code);
            }
            if (h) fold_hit<R>(sc, lv, ri, t, code);
            else code = -1;
        }
        if (!sc.pickless) { cq.obj[k] = code; ct[k] = t; }
        }
    }
    if (COUNT) { warp_add(&ctr->exact_tests, n_exact); warp_add(&ctr->exact_flops, n_flops); }
}

// float64 only: among the hits that have the ray's smallest T, the smallest scan index
template <typename R>
__global__ void __launch_bounds__(256) k_closest_order(const __grid_constant__ DevScene<R> sc, const __grid_constant__ Level<R> lv,
                                                      const __grid_constant__ CandQueue cq, const Counters* ctr) {
    typedef typename BitsOf<R>::type UB;
    const int count = cand_count(cq, ctr);
    const UB* tbest = reinterpret_cast<const UB*>(lv.t);
    const R* ct = reinterpret_cast<const R*>(cq.t);
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        const int code = cq.obj[k];
        if (code < 0 || (code & HIT_LEAF_MASK) >= sc.n_all) continue;       // a miss, or a padding record (reached only by a NaN ray)
        const int ri = cq.ray[k];
        if (BitsOf<R>::of(ct[k]) == tbest[ri]) atomicMin(&lv.order[ri], sc.scan_of[code & HIT_LEAF_MASK]);
    }
}

// the winner writes its hit code (float32: and its T, which the key held until now)
template <typename R>
__global__ void __launch_bounds__(256) k_closest_pick(const __grid_constant__ DevScene<R> sc, const __grid_constant__ Level<R> lv,
                                                     const __grid_constant__ CandQueue cq, const Counters* ctr) {
    typedef typename BitsOf<R>::type UB;
    const int count = cand_count(cq, ctr);
    const R* ct = reinterpret_cast<const R*>(cq.t);
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        const int code = cq.obj[k];
        if (code < 0 || (code & HIT_LEAF_MASK) >= sc.n_all) continue;       // a miss, or a padding record (reached only by a NaN ray)
        const int ri = cq.ray[k];
        if constexpr (sizeof(R) == 4) {
            if (hit_key32(sc, ct[k], code) == reinterpret_cast<const unsigned long long*>(lv.order)[ri]) { lv.hit[ri] = code; lv.t[ri] = ct[k]; }
        } else {
            const UB* tbest = reinterpret_cast<const UB*>(lv.t);
            if (BitsOf<R>::of(ct[k]) == tbest[ri] && lv.order[ri] == sc.scan_of[code & HIT_LEAF_MASK]) lv.hit[ri] = code;
        }
    }
}

// inShadow for the queued (hit, light) pairs of ONE light (raytracer.go:411-429): an object other than the shaded
// one occludes when  shadowHit.T * |ray.Direction| < distToLight  (:424).  The answer does not depend on the
// order objects are visited in; `first` keeps the smallest scan index of an occluder so that COUNT can credit the
// tests the reference itself would have run up to its first occluder.
template <typename R, bool COUNT, int KINDS>
__global__ void __launch_bounds__(EXACT_THREADS) k_exact_shadow(const __grid_constant__ DevScene<R> sc, const __grid_constant__ ShadowQueue<R> sq,
                                                               int base, const __grid_constant__ CandQueue cq, Counters* ctr, unsigned char* occl) {
    const int count = cand_count(cq, ctr);
    unsigned long long n_exact = 0, n_flops = 0, n_occluded = 0;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        const int q = base + cq.ray[k], obj = cq.obj[k];
        if (obj >= sc.n_leaves || !exact_owns<KINDS>(obj, sc.kind_begin[K_SPHERE + 1], sc.kind_begin[K_CSG]) || obj == sq.skip[q]) continue;
        if (!COUNT && sq.first[q] != 0x7fffffff) continue;     // already occluded (a stale read only costs a test)
        if (COUNT) { n_exact++; n_flops += (unsigned)sc.obj_flops[obj]; }
        Ray<R> w;
        const R* rec = sq.q + (size_t)q * 8;
        w.ox = rec[0]; w.oy = rec[1]; w.oz = rec[2]; w.dx = rec[3]; w.dy = rec[4]; w.dz = rec[5];
        R t; int code;
        if (!exact_hit_kinds<R, KINDS>(sc, obj, w, t, code)) continue;
        if (t * rec[7] < rec[6]) {                             // T * |incoming direction| < distance to the light
            if (COUNT) atomicMin(&sq.first[q], sc.scan_of[obj]);   // k_shadow_finish sets the flag and credits the reference's tests
            else if (atomicExch(&sq.first[q], 0) == 0x7fffffff) { occl[sq.target[q]] = 1; n_occluded++; }   // first occluder found: flag it here
        }
    }
    if (COUNT) { warp_add(&ctr->exact_tests, n_exact); warp_add(&ctr->exact_flops, n_flops); warp_add(&ctr->exact_tests_shadow, n_exact); }
    else warp_add(&ctr->shadow_occluded, n_occluded);
}

// ------------------------------------------------------------------ shadow rays by shaded object
// The packet stage of k_cull only pays off when the 32 * RPT rays of a warp form a narrow beam.  In queue order
// (= screen order of the rays that spawned them) a warp's shadow rays start on whatever objects a strip of pixels
// happens to cross, at any depth; ordered by the REGION of the shaded point (DevScene::rid_of: the shaded object when
// it is small, else the grid cell of the point) they start within one bounding sphere / one cell.  Counting sort of
// slot indices, keys = sq.key: block-private histograms, one scan, block-wise scatter.  Only the 4-byte permutation
// is written; the 32-byte ray records stay where k_shade put them.
constexpr int SORT_BINS_MAX = 8192, SORT_CHUNK = 4096, SORT_THREADS = 256;

__device__ __forceinline__ int sort_bin(int key, int shift, int nb) { return key < 0 ? nb - 1 : min(key >> shift, nb - 2); }

template <typename R>
__global__ void __launch_bounds__(SORT_THREADS) k_sort_hist(const __grid_constant__ ShadowQueue<R> sq, int light, int shift, int nb, int* hist, Counters* ctr) {
    __shared__ int h[SORT_BINS_MAX];
    const int n = sq.count[light], base = light * sq.per_light;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&ctr->rays_shadow, (unsigned long long)n);
    const int nchunks = (n + SORT_CHUNK - 1) / SORT_CHUNK;
    for (int b = threadIdx.x; b < nb; b += SORT_THREADS) h[b] = 0;
    __syncthreads();
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x)
        for (int i = c * SORT_CHUNK + threadIdx.x; i < min(n, (c + 1) * SORT_CHUNK); i += SORT_THREADS)
            atomicAdd(&h[sort_bin(sq.key[base + i], shift, nb)], 1);
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += SORT_THREADS) if (h[b]) atomicAdd(&hist[b], h[b]);
}

// exclusive prefix of hist[0..nb) into cursor[0..nb); hist is cleared for the next use (one block; tiles of 4096
// bins read as int4, so the 55k bins of the key sort are 14 coalesced rounds)
static __global__ void __launch_bounds__(1024) k_sort_scan(int nb, int* hist, int* cursor) {
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    auto load_tile = [&](int tile) {
        const int b = tile + 4 * tid;
        int4 v = make_int4(0, 0, 0, 0);
        if (b + 3 < nb) v = *reinterpret_cast<const int4*>(hist + b);
        else { if (b < nb) v.x = hist[b]; if (b + 1 < nb) v.y = hist[b + 1]; if (b + 2 < nb) v.z = hist[b + 2]; }
        return v;
    };
    int4 nxt = load_tile(0);
    for (int tile = 0; tile < nb; tile += 4096) {
        const int b = tile + 4 * tid;
        const int4 v = nxt;
        if (tile + 4096 < nb) nxt = load_tile(tile + 4096);   // in flight while this tile is scanned (the loop is latency-bound)
        const int mine = v.x + v.y + v.z + v.w;
        int incl = mine;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += u; }
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int t = warp_tot[lane], ti = t;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) { const int u = __shfl_up_sync(0xffffffffu, ti, off); if (lane >= off) ti += u; }
            warp_tot[lane] = ti - t;                       // exclusive prefix of the warp totals
        }
        __syncthreads();
        const int carry = carry_s;
        int run = carry + warp_tot[wid] + incl - mine;
        if (b < nb) { cursor[b] = run; hist[b] = 0; } run += v.x;
        if (b + 1 < nb) { cursor[b + 1] = run; hist[b + 1] = 0; } run += v.y;
        if (b + 2 < nb) { cursor[b + 2] = run; hist[b + 2] = 0; } run += v.z;
        if (b + 3 < nb) { cursor[b + 3] = run; hist[b + 3] = 0; } run += v.w;
        __syncthreads();
        if (tid == 1023) carry_s = run;                    // inclusive total through this tile
        __syncthreads();
    }
}

template <typename R>
__global__ void __launch_bounds__(SORT_THREADS) k_sort_scatter(const __grid_constant__ ShadowQueue<R> sq, int light, int shift, int nb, int* cursor) {
    __shared__ int h[SORT_BINS_MAX];                       // this chunk's count per bin, then (in place) its cursor in the bin
    const int n = sq.count[light], base = light * sq.per_light;
    const int nchunks = (n + SORT_CHUNK - 1) / SORT_CHUNK;
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
        for (int b = threadIdx.x; b < nb; b += SORT_THREADS) h[b] = 0;
        __syncthreads();
        const int i1 = min(n, (c + 1) * SORT_CHUNK);
        for (int i = c * SORT_CHUNK + threadIdx.x; i < i1; i += SORT_THREADS) atomicAdd(&h[sort_bin(sq.key[base + i], shift, nb)], 1);
        __syncthreads();
        for (int b = threadIdx.x; b < nb; b += SORT_THREADS) { const int c_b = h[b]; if (c_b) h[b] = atomicAdd(&cursor[b], c_b); }   // this chunk's range in bin b
        __syncthreads();
        for (int i = c * SORT_CHUNK + threadIdx.x; i < i1; i += SORT_THREADS)
            sq.perm[base + atomicAdd(&h[sort_bin(sq.key[base + i], shift, nb)], 1)] = i;
        __syncthreads();
    }
}

// ------------------------------------------------------------------ reflection / refraction rays by (origin object, direction cell)
// Same idea for the rays k_shade spawns: in queue order a warp's rays leave whatever objects a strip of pixels shows, in
// every direction; keyed by the object they leave and the cube-map cell of their direction (k_shade leaves the key in
// the child's `hit` slot, which the cull kernel re-initialises anyway) a warp's rays start in one small region and
// point into one cone, so the packet stage applies to them too.  Up to 2^17 bins: histogram and cursors in global
// memory, one warp-aggregated atomic per distinct key of a warp.  The order inside a bin is not deterministic; the
// result of a ray does not depend on which warp tests it.
constexpr int KEY_BINS_MAX = 1 << 19;

__device__ __forceinline__ void key_peers(int key, bool valid, unsigned& peers, int& rank, int& leader) {
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    peers = 0; rank = 0; leader = 0;
    if (!valid) return;
    peers = __match_any_sync(act, key);
    rank = __popc(peers & ((1u << (threadIdx.x & 31)) - 1u));
    leader = __ffs(peers) - 1;
}

static __global__ void __launch_bounds__(256) k_keysort_hist(const int* keys, const int* n_ptr, int nb, int* hist) {
    const int n = *n_ptr;
    const int rounds = (n + gridDim.x * blockDim.x - 1) / (gridDim.x * blockDim.x);
    for (int r = 0; r < rounds; r++) {                     // whole warps stay in the loop: match / ballot need them
        const int i = (r * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        const bool valid = i < n;
        const int key = valid ? min(max(keys[i], 0), nb - 1) : 0;
        unsigned peers; int rank, leader;
        key_peers(key, valid, peers, rank, leader);
        if (valid && rank == 0) atomicAdd(&hist[key], __popc(peers));
    }
}

template <typename R>
__global__ void __launch_bounds__(256) k_keysort_scatter(const int* keys, const int* n_ptr, int nb, int* cursor, const R* stage, R* out) {
    const int n = *n_ptr;
    const int rounds = (n + gridDim.x * blockDim.x - 1) / (gridDim.x * blockDim.x);
    for (int r = 0; r < rounds; r++) {
        const int i = (r * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        const bool valid = i < n;
        const int key = valid ? min(max(keys[i], 0), nb - 1) : 0;
        unsigned peers; int rank, leader;
        key_peers(key, valid, peers, rank, leader);
        int at = 0;
        if (valid && rank == 0) at = atomicAdd(&cursor[key], __popc(peers));
        at = __shfl_sync(0xffffffffu, at, leader);
        if (valid) {                                       // move the whole record: the level is STORED in sorted order
            const uint4* src = reinterpret_cast<const uint4*>(stage + (size_t)i * RAY_WORDS);
            uint4* dst = reinterpret_cast<uint4*>(out + (size_t)(at + rank) * RAY_WORDS);
#pragma unroll
            for (int k = 0; k < (int)(RAY_WORDS * sizeof(R) / 16); k++) dst[k] = src[k];
        }
    }
}

// occlusion flags for the resolve pass (+ the credited-work counters)
template <typename R, bool COUNT>
__global__ void __launch_bounds__(256) k_shadow_finish(const __grid_constant__ DevScene<R> sc, const __grid_constant__ ShadowQueue<R> sq,
                                                      int light, unsigned char* occl, Counters* ctr) {
    const int n = sq.count[light], base = light * sq.per_light;
    unsigned long long n_occluded = 0, cred_tests = 0, cred_flops = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int q = base + i;
        const int fs = sq.first[q];
        const bool occluded = fs != 0x7fffffff;
        if (occluded) { occl[sq.target[q]] = 1; n_occluded++; }
        if (COUNT) {
            // objects 0..last of the scan order are tested, minus the shaded one (raytracer.go:416-427)
            const int last = occluded ? fs : sc.n_leaves - 1;
            if (last >= 0) {
                unsigned long long tests = (unsigned long long)last + 1, flops = sc.cum_flops[last];
                const int skip = sq.skip[q];
                const int ss = skip >= 0 ? sc.scan_of[skip] : 0x7fffffff;
                if (ss <= last) { tests--; flops -= (unsigned long long)sc.obj_flops[skip]; }
                cred_tests += tests; cred_flops += flops;
            }
        }
    }
    warp_add(&ctr->shadow_occluded, n_occluded);
    if (COUNT) { warp_add(&ctr->shadow_tests, cred_tests); warp_add(&ctr->shadow_flops, cred_flops); }
}

// ------------------------------------------------------------------ surface bytecode interpreter
// EvalSurfaceFn (internal/gml/evaluator.go:672-727) after host-side partial evaluation: typed
// register code, see include/gmlrt.h.  Registers live in local memory (dynamic indexing).
template <typename R> struct IntOf;
template <> struct IntOf<float> { typedef long long type; };   // Go's VInt is int64 in both builds: integer-hash textures (muli by large primes) must not wrap at 32 bits
template <> struct IntOf<double> { typedef long long type; };

template <typename R> union Cell { R f; typename IntOf<R>::type i; };

// returns false on a fault (index out of range, integer division by zero, ERR, malformed code)
template <typename R>
__device__ __noinline__ bool run_surface(const DevScene<R>& sc, const DevSurface<R>& sf, int face, R u, R v, R* mat) {
    typedef typename IntOf<R>::type I;
    typedef typename std::make_unsigned<I>::type U;
    Cell<R> r[GMLRT_VM_MAX_REGS];
    const uint32_t* code = sc.code + sf.code_off;
    uint32_t pc = 0;
    const uint32_t n = sf.code_len;
    const R deg = R(GM_PI / 180.0);
    r[0].i = face; r[1].f = u; r[2].f = v;
    while (pc < n) {
        const uint32_t w = code[pc++];
        const uint32_t op = w & 0xFF, d = (w >> 8) & 0x7F, a = (w >> 16) & 0x7F, b = (w >> 24) & 0x7F;
        switch (op) {
        case GMLRT_OP_RET:
            if (a + 10 > GMLRT_VM_MAX_REGS) return false;
            for (int k = 0; k < 10; k++) mat[k] = r[a + k].f;
            return true;
        case GMLRT_OP_LDK: { uint32_t k = ((w >> 16) & 0xFF) | ((w >> 24) << 8); if ((int)k >= sc.n_fpool) return false; r[d].f = sc.fpool[k]; break; }
        case GMLRT_OP_LDI: r[d].i = (int32_t)code[pc++]; break;
        case GMLRT_OP_MOV: r[d] = r[a]; break;
        case GMLRT_OP_ADDF: r[d].f = r[a].f + r[b].f; break;
        case GMLRT_OP_SUBF: r[d].f = r[a].f - r[b].f; break;
        case GMLRT_OP_MULF: r[d].f = r[a].f * r[b].f; break;
        case GMLRT_OP_DIVF: r[d].f = r[a].f / r[b].f; break;
        case GMLRT_OP_NEGF: r[d].f = -r[a].f; break;
        case GMLRT_OP_ADDI: r[d].i = (I)((U)r[a].i + (U)r[b].i); break;
        case GMLRT_OP_SUBI: r[d].i = (I)((U)r[a].i - (U)r[b].i); break;
        case GMLRT_OP_MULI: r[d].i = (I)((U)r[a].i * (U)r[b].i); break;
        case GMLRT_OP_DIVI: if (r[b].i == 0) return false; r[d].i = (r[b].i == -1) ? (I)(U(0) - (U)r[a].i) : r[a].i / r[b].i; break;
        case GMLRT_OP_MODI: if (r[b].i == 0) return false; r[d].i = (r[b].i == -1) ? 0 : r[a].i % r[b].i; break;
        case GMLRT_OP_NEGI: r[d].i = (I)(U(0) - (U)r[a].i); break;
        case GMLRT_OP_EQF: r[d].i = r[a].f == r[b].f; break;
        case GMLRT_OP_LESSF: r[d].i = r[a].f < r[b].f; break;
        case GMLRT_OP_EQI: r[d].i = r[a].i == r[b].i; break;
        case GMLRT_OP_LESSI: r[d].i = r[a].i < r[b].i; break;
        case GMLRT_OP_SIN: r[d].f = gm<R>::sin(deg * r[a].f); break;
        case GMLRT_OP_COS: r[d].f = gm<R>::cos(deg * r[a].f); break;
        case GMLRT_OP_SQRT: r[d].f = gm<R>::sqrt(r[a].f); break;
        case GMLRT_OP_CLAMPF: { R x = r[a].f; if (x < R(0)) x = R(0); else if (x > R(1)) x = R(1); r[d].f = x; break; }
        case GMLRT_OP_FLOOR: r[d].i = (I)gm<R>::floor(r[a].f); break;
        case GMLRT_OP_FRAC: r[d].f = r[a].f - (R)(I)r[a].f; break;
        case GMLRT_OP_REAL: r[d].f = (R)r[a].i; break;
        case GMLRT_OP_ACOS: r[d].f = gm<R>::acos(r[a].f) / deg; break;
        case GMLRT_OP_ASIN: r[d].f = gm<R>::asin(r[a].f) / deg; break;
        case GMLRT_OP_LDARR: r[d].i = code[pc]; r[(d + 1) & 0x7F].i = code[pc + 1]; pc += 2; break;
        case GMLRT_OP_LEN: r[d].i = r[(a + 1) & 0x7F].i; break;
        case GMLRT_OP_GETF: case GMLRT_OP_GETP: case GMLRT_OP_GETI: case GMLRT_OP_GETA: {
            const long long off = r[a].i, len = r[(a + 1) & 0x7F].i, idx = r[b].i;
            if (idx < 0 || idx >= len) return false;                        // evaluator.go:989-995
            if (op == GMLRT_OP_GETF) { if (off + idx >= sc.n_fpool) return false; r[d].f = sc.fpool[off + idx]; }
            else if (op == GMLRT_OP_GETP) {
                if (off + 3 * idx + 2 >= sc.n_fpool || d + 2 >= GMLRT_VM_MAX_REGS) return false;
                for (int k = 0; k < 3; k++) r[d + k].f = sc.fpool[off + 3 * idx + k];
            } else if (op == GMLRT_OP_GETI) { if (off + idx >= sc.n_itab) return false; r[d].i = sc.itab[off + idx]; }
            else {
                if (off + 2 * idx + 1 >= sc.n_itab) return false;
                r[d].i = sc.itab[off + 2 * idx]; r[(d + 1) & 0x7F].i = sc.itab[off + 2 * idx + 1];
            }
            break;
        }
        case GMLRT_OP_JMP: pc += 1 + code[pc]; break;
        case GMLRT_OP_JMPF: if (r[a].i == 0) pc += 1 + code[pc]; else pc += 1; break;
        default: return false;   // GMLRT_OP_ERR and anything unknown
        }
    }
    return false;
}

// ------------------------------------------------------------------ small vector helpers (Vec3, vec.go:23-116)
template <typename R> struct V3 { R x, y, z; };
// sort key of a spawned ray (see k_keysort_*)
template <typename R> __device__ __forceinline__ int ray_sort_key(const DevScene<R>& sc, int region, const V3<R>& d) {
    const R ax = gm<R>::abs(d.x), ay = gm<R>::abs(d.y), az = gm<R>::abs(d.z);
    int face; R m, u, v;
    if (ax >= ay && ax >= az) { face = d.x < R(0) ? 1 : 0; m = ax; u = d.y; v = d.z; }
    else if (ay >= az) { face = d.y < R(0) ? 3 : 2; m = ay; u = d.x; v = d.z; }
    else { face = d.z < R(0) ? 5 : 4; m = az; u = d.x; v = d.y; }
    const int k = sc.key_k;
    const R half = R(0.5) * (R)k, s = half / m;
    int cu = (int)fma_(u, s, half), cv = (int)fma_(v, s, half);
    cu = min(k - 1, max(0, cu)); cv = min(k - 1, max(0, cv));
    if (cu & 1) cv = k - 1 - cv;                           // boustrophedon: consecutive cells of a face are neighbours
    return (region >> sc.key_shift) * (6 * k * k) + (face * k + cu) * k + cv;
}

template <typename R> __device__ __forceinline__ V3<R> mk(R x, R y, R z) { V3<R> v = {x, y, z}; return v; }
template <typename R> __device__ __forceinline__ V3<R> operator+(V3<R> a, V3<R> b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename R> __device__ __forceinline__ V3<R> operator-(V3<R> a, V3<R> b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename R> __device__ __forceinline__ V3<R> operator*(V3<R> a, R s) { return mk(a.x * s, a.y * s, a.z * s); }
template <typename R> __device__ __forceinline__ V3<R> hadamard(V3<R> a, V3<R> b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename R> __device__ __forceinline__ R dot(V3<R> a, V3<R> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename R> __device__ __forceinline__ R length(V3<R> a) { return gm<R>::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
template <typename R> __device__ __forceinline__ V3<R> normalize(V3<R> a) {
    R m = gm<R>::sqrt(a.x * a.x + a.y * a.y + a.z * a.z);
    return mk(a.x / m, a.y / m, a.z / m);
}
template <typename R> __device__ __forceinline__ V3<R> neg(V3<R> a) { return mk(-a.x, -a.y, -a.z); }
template <typename R> __device__ __forceinline__ V3<R> clamp3(V3<R> a) {
    return mk(gomath::clamp01(a.x), gomath::clamp01(a.y), gomath::clamp01(a.z));
}
template <typename R> __device__ __forceinline__ V3<R> ld3(const R* p) { return mk(p[0], p[1], p[2]); }
// the closest hit of ray i of a level after the scan: hit code (or -1) and T.  PICKLESS: decoded from the folded key.
template <typename R, int PICKLESS> __device__ __forceinline__ void level_hit(const Level<R>& lv, int i, int& code, R& t) {
    if constexpr (PICKLESS) {
        const unsigned long long key = reinterpret_cast<const unsigned long long*>(lv.order)[i];
        const unsigned low = (unsigned)key;                   // mode 1: the device index; mode 2: leaf << 4 | flip << 3 | face (hit_key32) -> leaf | flip << 27 | face << 28
        if constexpr (PICKLESS == 1) code = key == ~0ull ? -1 : (int)low;
        else code = key == ~0ull ? -1 : (int)((low >> 4) | ((low & 8u) << 24) | ((low & 7u) << 28));
        t = (R)__uint_as_float((unsigned)(key >> 32));
    } else { code = lv.hit[i]; t = R(0); }
}
template <typename R, int PICKLESS> __device__ __forceinline__ bool level_is_miss(const Level<R>& lv, int i) {
    if constexpr (PICKLESS) return reinterpret_cast<const unsigned long long*>(lv.order)[i] == ~0ull;
    else return lv.hit[i] < 0;
}
template <typename R, int PICKLESS> __device__ __forceinline__ void level_set_miss(const Level<R>& lv, int i) {
    if constexpr (PICKLESS) reinterpret_cast<unsigned long long*>(lv.order)[i] = ~0ull;
    else lv.hit[i] = -1;
}
// sort region of a point P on leaf `leaf` (DevScene::rid_of)
template <typename R> __device__ __forceinline__ int region_of(const DevScene<R>& sc, int leaf, const V3<R>& P, R px, R pz) {
    const int r = sc.rid_of[leaf];
    if (r >= 0) return r;
    if (r <= -2) {                                       // unbounded plane: its own (x, z) = (px, pz), compressed by g
        const int k = -2 - r;
        const R x = (px - sc.plane_par[k][0]) * sc.plane_par[k][2], z = (pz - sc.plane_par[k][1]) * sc.plane_par[k][2];
        const R gx = x / (R(1) + gm<R>::abs(x)), gz = z / (R(1) + gm<R>::abs(z));
        auto cell2 = [](R g) { const R f = (g + R(1)) * R(GMLRT_PLANE_GRID / 2); const R hi = R(GMLRT_PLANE_GRID - 1); return (int)(f > R(0) ? (f < hi ? f : hi) : R(0)); };
        int cx = cell2(gx), cz = cell2(gz);
        if (cx & 1) cz = GMLRT_PLANE_GRID - 1 - cz;          // boustrophedon: consecutive cells are neighbours
        return sc.plane_base[k] + cx * GMLRT_PLANE_GRID + cz;
    }
    const R fx = (P.x - sc.grid_lo[0]) * sc.grid_inv[0], fy = (P.y - sc.grid_lo[1]) * sc.grid_inv[1], fz = (P.z - sc.grid_lo[2]) * sc.grid_inv[2];
    auto cell = [](R f, int dim) { const R hi = (R)(dim - 1); return (int)(f > R(0) ? (f < hi ? f : hi) : R(0)); };   // clamped as reals (P on a plane may be far outside the grid; NaN -> 0)
    const int cx = cell(fx, sc.grid_dim[0]), cy = cell(fy, sc.grid_dim[1]), cz = cell(fz, sc.grid_dim[2]);
    return sc.n_small + (cx * sc.grid_dim[1] + cy) * sc.grid_dim[2] + cz;
}
// (M^T) v : NormalMat.MulDir with NormalMat = WorldToObject^T (raytracer.go:367, 790)
template <typename R> __device__ __forceinline__ V3<R> mul_dir_t(const R* m, V3<R> v) {
    return mk(m[0] * v.x + m[4] * v.y + m[8] * v.z, m[1] * v.x + m[5] * v.y + m[9] * v.z, m[2] * v.x + m[6] * v.y + m[10] * v.z);
}
template <typename R> __device__ __forceinline__ V3<R> mul_point(const R* m, V3<R> v) {
    return mk(m[0] * v.x + m[1] * v.y + m[2] * v.z + m[3], m[4] * v.x + m[5] * v.y + m[6] * v.z + m[7],
              m[8] * v.x + m[9] * v.y + m[10] * v.z + m[11]);
}

// queue append with one atomic per warp (ballot + popc compaction).  Must be called by all 32 lanes.
__device__ __forceinline__ int warp_append(int* counter, bool want) {
    const unsigned m = __ballot_sync(0xffffffffu, want);
    if (m == 0) return -1;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(m) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return want ? base + __popc(m & ((1u << lane) - 1u)) : -1;
}

// queue append with one atomic per CTA: every thread of the block must call it (block-uniform control flow).  All CTAs
// of a launch append to the same counter; with one atomic per warp the wait for its return was 40% of k_shade's
// stall samples (ncu r01c); one reservation per CTA issues fewer (CTAs of 64 threads measured best: fewer atomics than per
// warp, less barrier wait than 128 or 256 threads: shade 2.42 / 2.11 / 1.96 ms at 256 / 128 / 64).  `scratch` = int[2][SHADE_WARPS + 1], toggled by `phase`.
constexpr int SHADE_THREADS = 64, SHADE_WARPS = SHADE_THREADS / 32;
constexpr int SHADE_LIGHTS_PER_ROUND = 6, SHADE_Q_REFL = 6, SHADE_Q_REFR = 7, SHADE_QUEUES = 8;   // queues reserved per round trip in k_shade
__device__ __forceinline__ int block_append(int* counter, bool want, int (*scratch)[SHADE_WARPS + 1], int& phase, unsigned long long* tally = nullptr) {
    int* s = scratch[phase & 1];
    phase++;
    const unsigned m = __ballot_sync(0xffffffffu, want);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) s[wid] = __popc(m);
    __syncthreads();
    if (threadIdx.x < 32) {
        const int c = lane < SHADE_WARPS ? s[lane] : 0;
        int incl = c;
#pragma unroll
        for (int off = 1; off < SHADE_WARPS; off <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += u; }
        const int total = __shfl_sync(0xffffffffu, incl, SHADE_WARPS - 1);
        int base = 0;
        if (lane == 0 && total) { base = atomicAdd(counter, total); if (tally) atomicAdd(tally, (unsigned long long)total); }
        base = __shfl_sync(0xffffffffu, base, 0);
        if (lane < SHADE_WARPS) s[lane] = base + incl - c;
    }
    __syncthreads();
    return want ? s[wid] + __popc(m & ((1u << lane) - 1u)) : -1;
}

// direction to a light, distance, and its colour at P (spot falloff is an EXTENSION); false = unlit
template <typename R>
__device__ __forceinline__ bool light_at(const DevLight<R>& L, V3<R> P, V3<R>& ldir, R& dist, V3<R>& lcol) {
    lcol = ld3(L.color);
    if (L.kind == GMLRT_LIGHT_DIRECTIONAL) {
        ldir = normalize(neg(ld3(L.pos)));
        dist = gm<R>::inf();
        return true;
    }
    V3<R> l2h = ld3(L.pos) - P;                       // raytracer.go:379-381
    dist = length(l2h);
    ldir = normalize(l2h);
    if (L.kind == GMLRT_LIGHT_SPOT) {
        R cosang = dot(ld3(L.axis), neg(ldir));
        if (cosang < L.cos_cutoff) return false;
        lcol = lcol * gm<R>::pow(cosang, L.exponent);
    }
    return true;
}

// ------------------------------------------------------------------ gen: camera rays (Render, raytracer.go:632-652)


template <typename R>
__global__ void __launch_bounds__(256) k_gen(DevScene<R> sc, Level<R> lv, Counters* ctr, int y0, int rows, RowMap map) {
    const int S = rows * sc.W * 4;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        for (int l = 1; l <= GMLRT_MAX_DEPTH; l++) ctr->count[l] = 0;
        ctr->count[0] = S;
    }
    if (i >= S) return;
    double o[3], d[3];
    camera_ray64(sc, CamSrc{y0, map}, i, o, d);
    store_ray(lv.ray, i, (R)o[0], (R)o[1], (R)o[2], (R)d[0], (R)d[1], (R)d[2], 0);
}

// ------------------------------------------------------------------ shade (traceRay :498-556 minus the recursion)
// rec slots
enum { REC_P = 0, REC_N = 3, REC_COL = 6, REC_REFL = 9, REC_TRANSP = 10, REC_KR = 11, REC_KD = 12, REC_KS = 13, REC_NEXP = 14, REC_INLEN = 15 };

template <typename R, int PICKLESS>
__global__ void __launch_bounds__(SHADE_THREADS) k_shade(DevScene<R> sc, Level<R> lv, Level<R> next, ShadowQueue<R> sq,
                                                         unsigned char* occl, Counters* ctr, int level, int spawn,
                                                         int32_t* ids, int y0, RowMap map, R* spawn_to) {
    const int n = ctr->count[level];
    __shared__ int s_append[2][SHADE_QUEUES][SHADE_WARPS];
    int append_phase = 0;
    // The hit code of the next iteration is fetched one iteration ahead, so that an iteration's dependent chain starts
    // at the ray record instead of one DRAM round trip earlier.  (Fetching the record ahead as well was measured
    // slower: 11 more live registers spill at the 64 the kernel is held to.)
    const int stride = gridDim.x * blockDim.x;
    const int i_first = blockIdx.x * blockDim.x + threadIdx.x;
    int h_nx = -1; R t_nx = R(0);
    if (i_first < n) level_hit<R, PICKLESS>(lv, i_first, h_nx, t_nx);
    for (int blk0 = blockIdx.x * blockDim.x; blk0 < n; blk0 += stride) {   // uniform per block
    const int i = blk0 + threadIdx.x;
    const size_t cap = (size_t)lv.cap;                     // 64-bit strides: 15 * cap passes 2^31 on deep binary ray trees
    bool live = i < n;
    const int h = h_nx;                                    // = lv.hit[i], or -1 past the end
    const R t_pre = t_nx;
    h_nx = -1;
    if ((long long)i + stride < n) level_hit<R, PICKLESS>(lv, i + stride, h_nx, t_nx);
    if (ids && live) {                                     // level 0 only: primary hit ids in (y, x, sample) order
        const int pix = i >> 2;
        ids[((size_t)map.actual(y0 + pix / sc.W) * sc.W + pix % sc.W) * 4 + (i & 3)] = h < 0 ? -1 : sc.scan_of[h & HIT_LEAF_MASK];
    }
    live = live && h >= 0;
    Ray<R> w = {R(0), R(0), R(0), R(0), R(0), R(1)};
    V3<R> P = mk(R(0), R(0), R(0)), N = P;
    R mat[10];
    for (int k = 0; k < 10; k++) mat[k] = R(0);
    int leaf = 0;
    R obj_x = R(0), obj_z = R(0);                          // object-space hit point (x, z): the sort region of a point on an unbounded plane
    if (live) {
        leaf = h & HIT_LEAF_MASK;
        int face = h >> 28;
        const bool flip = (h & HIT_FLIP_BIT) != 0;          // CSG: surface of a subtracted solid
        load_ray(lv.ray, i, w);
        const R t = PICKLESS != 0 ? t_pre : lv.t[i];
        const R* Mw = sc.w2o + (size_t)leaf * 12;
        Ray<R> o;
        to_object(Mw, w, o);
        const V3<R> p = mk(o.ox + o.dx * t, o.oy + o.dy * t, o.oz + o.dz * t);       // Hit.PointObj
        obj_x = p.x; obj_z = p.z;
        const int kind = sc.kind[leaf];
        const DevSurface<R>& sf = sc.surfaces[sc.surface[leaf]];
        R u = R(0), v = R(0);
        bool ok = true;
        if (kind == K_SPHERE) {                                                      // :106-150
            face = 0;
            if (sf.kind != GMLRT_SURFACE_CONST) {
                R py = p.y;
                if (!GMLRT_PARITY_TU) py = fminf(fmaxf((float)py, -1.0f), 1.0f);      // float rounding can leave |y| = 1 + ulp at the poles: clamp instead of the reference's error
                if (gm<R>::abs(py) > R(1)) ok = false;                                // raytracer.go:142-144
                v = (py + R(1)) / R(2);
                R ca = p.z / gm<R>::sqrt(R(1) - py * py);
                if (!GMLRT_PARITY_TU) ca = fminf(fmaxf((float)ca, -1.0f), 1.0f);   // float rounding can leave |ca| = 1 + ulp on the x = 0 meridian
                u = gm<R>::acos(ca) / (R(2) * R(GM_PI));
            }
            N = p;                                                                   // :119 (sic: object-space point)
        } else if (kind == K_PLANE) {                                                // :182-205
            u = p.x; v = p.z; face = 0;
            N = ld3(sc.fnormal + (size_t)leaf * 18);
        } else if (kind == K_CUBE) {                                                 // :242-260
            u = p.x; v = p.z;
            N = ld3(sc.fnormal + (size_t)leaf * 18 + 3 * face);
        } else if (kind == K_CYLINDER) {                                             // :339-370
            V3<R> no;
            if (face == 0) { no = mk(p.x, R(0), p.z); u = (gm<R>::atan2(p.x, p.z) + R(GM_PI)) / (R(2) * R(GM_PI)); v = p.y; }
            else if (face == 1) { no = mk(R(0), R(1), R(0)); u = p.x; v = p.z; }
            else { no = mk(R(0), R(-1), R(0)); u = p.x; v = p.z; }
            N = normalize(mul_dir_t(Mw, no));
        } else {                                                                     // cone (EXTENSION)
            V3<R> no;
            if (face == 0) { no = mk(p.x, -p.y, p.z); u = (gm<R>::atan2(p.x, p.z) + R(GM_PI)) / (R(2) * R(GM_PI)); v = p.y; }
            else { no = mk(R(0), R(1), R(0)); u = p.x; v = p.z; }
            N = normalize(mul_dir_t(Mw, no));
        }
        if (ok) {
            if (sf.kind == GMLRT_SURFACE_CONST) { for (int k = 0; k < 10; k++) mat[k] = sf.material[k]; }
            else {                                       // the interpreter writes through a pointer: keep that array apart, so `mat` stays in registers on the constant path
                R out[10];
                ok = run_surface(sc, sf, face, u, v, out);
                for (int k = 0; k < 10; k++) mat[k] = out[k];
            }
        }
        if (!ok) {                                   // the reference panics here (:499-501)
            atomicExch(&ctr->error, 1);
            live = false;
            level_set_miss<R, PICKLESS>(lv, i);
        } else {
            P = mul_point(sc.o2w + (size_t)leaf * 12, p);
            if (flip) N = neg(N);
        }
    }
    const R refl = mat[3], fuzz = mat[4], transp = mat[5], ior = mat[6];
    const V3<R> D = mk(w.dx, w.dy, w.dz);
    const int region = live ? region_of(sc, leaf, P, obj_x, obj_z) : 0;
    if (live) {
        R* rec = lv.rec + i;
        rec[(REC_P + 0) * cap] = P.x; rec[(REC_P + 1) * cap] = P.y; rec[(REC_P + 2) * cap] = P.z;
        rec[(REC_N + 0) * cap] = N.x; rec[(REC_N + 1) * cap] = N.y; rec[(REC_N + 2) * cap] = N.z;
        rec[(REC_COL + 0) * cap] = mat[0]; rec[(REC_COL + 1) * cap] = mat[1]; rec[(REC_COL + 2) * cap] = mat[2];
        rec[REC_REFL * cap] = refl; rec[REC_TRANSP * cap] = transp;
        R kr = R(0);
        if (transp > R(0)) {                                                         // fresnel :456-467
            R cosi = dot(D, N) / (length(D) * length(N));
            R r0 = (R(1) - ior) / (R(1) + ior);
            r0 = r0 * r0;
            kr = r0 + (R(1) - r0) * gm<R>::pow(R(1) - gm<R>::abs(cosi), R(5));
        }
        rec[REC_KR * cap] = kr;
        rec[REC_KD * cap] = mat[7]; rec[REC_KS * cap] = mat[8]; rec[REC_NEXP * cap] = mat[9];
        rec[REC_INLEN * cap] = length(D);
    }
    // Children.  Every ray may append one shadow query per light that can reach P (inShadow :411-414), a reflection
    // child (:512-529) and a refraction child (:532-556).  All queues of a block iteration are reserved in ONE round
    // trip: per-warp counts go to shared memory, lane q of warp 0 scans queue q over the warps and issues its
    // atomicAdd, so the atomics of the (up to 6 lights + 2 child kinds) queues are in flight together.
    bool want_refl = false, want_refr = false;
    V3<R> refr_d = mk(R(0), R(0), R(0)), refr_n = N;
    if (spawn) {
        want_refl = live && refl > R(0);
        want_refr = live && transp > R(0);
        if (want_refr) {                                                             // refract :438-450
            R n1 = R(1), n2 = ior;
            if (dot(D, refr_n) > R(0)) { R tmp = n1; n1 = n2; n2 = tmp; refr_n = refr_n * R(-1); }
            const R ratio = n1 / n2;
            const R cosI = -dot(refr_n, D);
            const R sinT2 = ratio * ratio * (R(1) - cosI * cosI);
            if (sinT2 > R(1)) want_refr = false;                                     // total internal reflection
            else {
                const R cosT = gm<R>::sqrt(R(1) - sinT2);
                refr_d = D * ratio + refr_n * (ratio * cosI - cosT);                 // not re-normalised
                if (refr_d.x == R(0) && refr_d.y == R(0) && refr_d.z == R(0)) want_refr = false;
            }
        }
    }
    // A child slot is written by the child's own resolve pass.  Only a slot the combine will read although no child
    // was spawned needs a zero here: refraction lost to total internal reflection, or the reflection slot of a
    // transparent surface that does not reflect.  (At the deepest level nothing is spawned and k_resolve reads none.)
    if (spawn && live && transp > R(0)) {
        if (!want_refl) for (int k = 0; k < 3; k++) lv.child[k * cap + i] = R(0);
        if (!want_refr) for (int k = 3; k < 6; k++) lv.child[k * cap + i] = R(0);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    for (int l0 = 0; l0 == 0 || l0 < sc.n_lights; l0 += SHADE_LIGHTS_PER_ROUND) {       // uniform
        int (*s)[SHADE_WARPS] = s_append[append_phase & 1];
        append_phase++;
        unsigned wmask = 0;
        for (int q = 0; q < SHADE_LIGHTS_PER_ROUND && l0 + q < sc.n_lights; q++) {
            V3<R> ldir = mk(R(0), R(0), R(1)), lcol; R dist = R(0);
            if (live && light_at(sc.lights[l0 + q], P, ldir, dist, lcol)) wmask |= 1u << q;
        }
        if (l0 == 0) wmask |= (want_refl ? 1u << SHADE_Q_REFL : 0u) | (want_refr ? 1u << SHADE_Q_REFR : 0u);
#pragma unroll
        for (int q = 0; q < SHADE_QUEUES; q++) {
            const unsigned m = __ballot_sync(0xffffffffu, (wmask >> q) & 1u);
            if (lane == 0) s[q][wid] = __popc(m);
        }
        __syncthreads();
        if (wid == 0) {
            const int q = lane;
            int excl[SHADE_WARPS], total = 0;
            if (q < SHADE_QUEUES) {
#pragma unroll
                for (int w = 0; w < SHADE_WARPS; w++) { excl[w] = total; total += s[q][w]; }
            }
            // the two child kinds share the next level's queue: reflection children first, then refraction children
            const int t_refr = __shfl_sync(0xffffffffu, total, SHADE_Q_REFR);
            int base = 0;
            if (q < SHADE_LIGHTS_PER_ROUND) { if (total) base = atomicAdd(&sq.count[l0 + q], total); }
            else if (q == SHADE_Q_REFL) {
                const int want = total + t_refr;
                if (want) {
                    base = atomicAdd(&ctr->count[level + 1], want);
                    if (base + want > next.cap) {              // level full (plan_tiles sizes branching levels below 2^l): give back what does not fit
                        atomicSub(&ctr->count[level + 1], base + want - max(base, next.cap));
                        ctr->overflow = 1;
                    }
                }
                if (total) atomicAdd(&ctr->rays_reflect, (unsigned long long)total);
                if (t_refr) atomicAdd(&ctr->rays_refract, (unsigned long long)t_refr);
            }
            const int b_refl = __shfl_sync(0xffffffffu, base, SHADE_Q_REFL), t_refl = __shfl_sync(0xffffffffu, total, SHADE_Q_REFL);
            if (q == SHADE_Q_REFR) base = b_refl + t_refl;
            if (q < SHADE_QUEUES) {
#pragma unroll
                for (int w = 0; w < SHADE_WARPS; w++) s[q][w] = base + excl[w];
            }
        }
        __syncthreads();
        for (int q = 0; q < SHADE_LIGHTS_PER_ROUND && l0 + q < sc.n_lights; q++) {
            const unsigned m = __ballot_sync(0xffffffffu, (wmask >> q) & 1u);
            if (!((wmask >> q) & 1u)) continue;
            const int li = l0 + q;
            V3<R> ldir = mk(R(0), R(0), R(1)), lcol; R dist = R(0);
            light_at(sc.lights[li], P, ldir, dist, lcol);
            const int slot = li * sq.per_light + s[q][wid] + __popc(m & lt);
            const V3<R> so = P + N * R(1e-4);
            store_rec8(sq.q, (size_t)slot, so.x, so.y, so.z, ldir.x, ldir.y, ldir.z, dist, length(D));
            sq.skip[slot] = leaf < sc.n_leaves ? leaf : -1;      // a CSG object may shadow itself (oracle.c in_shadow)
            sq.key[slot] = region;
            sq.target[slot] = i * sc.n_lights + li;
            occl[i * sc.n_lights + li] = 0;
        }
        if (l0 == 0 && spawn) {
            const unsigned m_refl = __ballot_sync(0xffffffffu, want_refl), m_refr = __ballot_sync(0xffffffffu, want_refr);
            if (want_refl) {                                                         // reflection child (:512-529)
                const int slot = s[SHADE_Q_REFL][wid] + __popc(m_refl & lt);
                if (slot < next.cap) {
                V3<R> rd = D - N * (R(2) * dot(D, N));
                if (fuzz >= R(0)) rd = rd + mk(fuzz * gm<R>::cos(fuzz) * gm<R>::cos(fuzz), fuzz * gm<R>::sin(fuzz) * gm<R>::sin(fuzz), R(0));
                const V3<R> ro = P + N * R(1e-4);
                rd = normalize(rd);
                store_ray(spawn_to, slot, ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, i);
                next.hit[slot] = ray_sort_key(sc, region, rd);
                }
            }
            if (want_refr) {                                                         // refraction child (:532-556)
                const int slot = s[SHADE_Q_REFR][wid] + __popc(m_refr & lt);
                if (slot < next.cap) {
                const V3<R> ro = P - refr_n * R(1e-4);
                store_ray(spawn_to, slot, ro.x, ro.y, ro.z, refr_d.x, refr_d.y, refr_d.z, (int)((unsigned)i | 0x80000000u));
                next.hit[slot] = ray_sort_key(sc, region, refr_d);
                }
            }
        }
    }
    }  // grid-stride loop
}

// ------------------------------------------------------------------ resolve one level (deepest first)
// computeLighting (:372-401) with the shadow flags, then the combine + clamp of traceRay
// (:493-497, :506-508, :557-561); the result goes into the parent's child slot, or into the
// sample buffer at level 0.
template <typename R, int PICKLESS>
__global__ void __launch_bounds__(SHADE_THREADS) k_resolve(DevScene<R> sc, Level<R> lv, Level<R> prev, const unsigned char* occl,
                                                           Counters* ctr, int level, uchar4* frame, int y0, RowMap map, int has_children) {
    const int n = ctr->count[level];
    unsigned long long n_hit = 0;
    // the loop bound is uniform per warp: every lane reaches the shuffles of the level-0 epilogue
    for (int i0 = blockIdx.x * blockDim.x + (threadIdx.x & ~31); i0 < n; i0 += gridDim.x * blockDim.x) {
    const int i = i0 + (threadIdx.x & 31);
    const bool valid = i < n;
    const size_t cap = (size_t)lv.cap;                     // 64-bit strides: 15 * cap passes 2^31 on deep binary ray trees
    V3<R> out = mk(R(0), R(0), R(0));
    if (valid) {
    Ray<R> wr;
    load_ray(lv.ray, i, wr);
    const V3<R> D = mk(wr.dx, wr.dy, wr.dz);
    if (level_is_miss<R, PICKLESS>(lv, i)) {
        const R t = R(0.5) * (D.y + R(1));                                           // :495-496
        const V3<R> a = ld3(sc.bg0), b = ld3(sc.bg1);
        out = mk(a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t, a.z + (b.z - a.z) * t);
    } else {
        n_hit++;
        const R* rec = lv.rec + i;
        const V3<R> P = mk(rec[(REC_P + 0) * cap], rec[(REC_P + 1) * cap], rec[(REC_P + 2) * cap]);
        const V3<R> N = mk(rec[(REC_N + 0) * cap], rec[(REC_N + 1) * cap], rec[(REC_N + 2) * cap]);
        const V3<R> C = mk(rec[(REC_COL + 0) * cap], rec[(REC_COL + 1) * cap], rec[(REC_COL + 2) * cap]);
        const R refl = rec[REC_REFL * cap], transp = rec[REC_TRANSP * cap], kr = rec[REC_KR * cap];
        const R kd = rec[REC_KD * cap], ks = rec[REC_KS * cap], nexp = rec[REC_NEXP * cap];
        const V3<R> Vv = neg(D);
        V3<R> L = ld3(sc.ambient) * kd;
        for (int li = 0; li < sc.n_lights; li++) {
            V3<R> ldir, lcol; R dist;
            if (!light_at(sc.lights[li], P, ldir, dist, lcol)) continue;
            if (occl[i * sc.n_lights + li]) continue;
            const R nDotL = gomath::max0(dot(N, ldir));
            const V3<R> diffuse = lcol * (nDotL * kd);
            const V3<R> H = normalize(Vv + ldir);
            const R spec = gomath::max0(dot(N, H));
            const V3<R> specular = lcol * (ks * gm<R>::pow(spec, nexp));
            L = (L + diffuse) + specular;
        }
        if (refl == R(0) && transp == R(0)) {
            out = clamp3(hadamard(L, C));
        } else {
            const V3<R> zero = mk(R(0), R(0), R(0));             // the deepest level has no children: traceRay returns black at depth 0 (:488-491)
            const V3<R> rc = has_children ? mk(lv.child[i], lv.child[cap + i], lv.child[2 * cap + i]) : zero;
            if (transp == R(0)) {
                out = clamp3(hadamard(L + rc * refl, C));
            } else {
                const V3<R> fc = has_children ? mk(lv.child[3 * cap + i], lv.child[4 * cap + i], lv.child[5 * cap + i]) : zero;
                out = clamp3(hadamard(L * (R(1) - transp) + (rc * kr + fc * (R(1) - kr)), C));
            }
        }
    }
    }  // valid
    if (level == 0) {
        // img.Set(x, y, total * 0.25) -> Vec3.RGBA -> RGBAModel (raytracer.go:656, vec.go:104-107), fused: camera rays sit
        // in (pixel, sample) order, so the 4 samples of a pixel are 4 neighbouring lanes (the grid stride and the ray
        // count are multiples of 4: a quad is in the loop as a whole).  Lane 0 of the quad sums them in the reference's
        // order - from zero, sample 0 first - quantises uint8(uint32(c * 65535) >> 8) and stores the pixel: no sample
        // buffer, no separate pass.  The frame may be a peer device's (NVLink stores).
        const unsigned quad = 0xffffffffu;
        const int q0 = (threadIdx.x & 31) & ~3;
        R c[3] = {out.x, out.y, out.z};
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const R s1 = __shfl_sync(quad, c[k], q0 + 1), s2 = __shfl_sync(quad, c[k], q0 + 2), s3 = __shfl_sync(quad, c[k], q0 + 3);
            R tot = R(0);
            tot = tot + c[k]; tot = tot + s1; tot = tot + s2; tot = tot + s3;
            c[k] = tot * R(0.25);
        }
        if (valid && (i & 3) == 0) {
            const int pix = i >> 2;
            uchar4 px;
            px.x = (unsigned char)((unsigned)(c[0] * R(65535)) >> 8);
            px.y = (unsigned char)((unsigned)(c[1] * R(65535)) >> 8);
            px.z = (unsigned char)((unsigned)(c[2] * R(65535)) >> 8);
            px.w = 0xFF;
            frame[(size_t)map.actual(y0 + pix / sc.W) * sc.W + pix % sc.W] = px;
        }
    } else if (valid) {
        const int p = ray_parent(lv.ray, i);
        const int pi = p & 0x7fffffff;
        const int base = (p < 0) ? 3 : 0;
        const size_t pc = (size_t)prev.cap;
        prev.child[(base + 0) * pc + pi] = out.x; prev.child[(base + 1) * pc + pi] = out.y; prev.child[(base + 2) * pc + pi] = out.z;
    }
    }  // grid-stride loop
    n_hit = warp_sum_u64(n_hit);                           // (every lane reaches this point)
    if ((threadIdx.x & 31) == 0 && n_hit) atomicAdd(&ctr->rays_hit, n_hit);
}


// ------------------------------------------------------------------ host side: one renderer per device
#define CU_TRY(expr)                                                                                   \
    do {                                                                                               \
        cudaError_t e_ = (expr);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            err = std::string(#expr) + ": " + cudaGetErrorString(e_);                                  \
            return e_ == cudaErrorMemoryAllocation ? GMLRT_E_OOM : GMLRT_E_CUDA;                       \
        }                                                                                              \
    } while (0)

inline double env_double(const char* name, double dflt) {
    const char* e = getenv(name);
    if (!e || !*e) return dflt;
    char* end = nullptr;
    const double v = strtod(e, &end);
    return end == e ? dflt : v;
}
inline size_t env_size(const char* name, size_t dflt) {
    const char* v = std::getenv(name);
    return (v && *v) ? (size_t)std::strtoull(v, nullptr, 10) : dflt;
}

template <typename R>
class Renderer : public IRenderer {
public:
    explicit Renderer(int device) : dev_(device) {}
    ~Renderer() override { release(); }

    int device() const override { return dev_; }
    cudaStream_t stream() override { return stream_; }
    uint8_t* frame() override { return reinterpret_cast<uint8_t*>(frame_); }
    int32_t* primary_ids() override { return ids_; }
    int max_rows_per_tile() const override { return rows_per_tile_; }

    int init(std::string& err) {
        CU_TRY(cudaSetDevice(dev_));
        CU_TRY(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
        for (int k = 0; k < SHADOW_STREAMS; k++) CU_TRY(cudaStreamCreateWithFlags(&shadow_stream_[k], cudaStreamNonBlocking));
        overlap_shadows_ = env_size("GMLRT_OVERLAP_SHADOWS", 1) != 0;
        cudaDeviceProp prop;
        CU_TRY(cudaGetDeviceProperties(&prop, dev_));
        sms_ = prop.multiProcessorCount;
        // PCG jump table: J_k advances the 128-bit LCG by k steps (affine map powers)
        gomath::PcgJump tab[160];
        unsigned __int128 m = 1, inc = 0;
        const unsigned __int128 MUL = ((unsigned __int128)gomath::PCG_MUL_HI << 64) | gomath::PCG_MUL_LO;
        const unsigned __int128 INC = ((unsigned __int128)gomath::PCG_INC_HI << 64) | gomath::PCG_INC_LO;
        for (int k = 0; k < 160; k++) {
            tab[k].mh = (uint64_t)(m >> 64); tab[k].ml = (uint64_t)m;
            tab[k].ih = (uint64_t)(inc >> 64); tab[k].il = (uint64_t)inc;
            m = m * MUL; inc = inc * MUL + INC;
        }
        CU_TRY(cudaMemcpyToSymbol(c_pcg_jump, tab, sizeof tab));
        CU_TRY(cudaMalloc(&ctr_, sizeof(Counters)));
        CU_TRY(cudaMemset(ctr_, 0, sizeof(Counters)));
        const int smem_max = 200 * 1024;
        auto allow = [&](auto kernel) -> int { CU_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max)); return GMLRT_OK; };
        int rc;
        if ((rc = allow(k_cull<R, CULL_PIVOT, RPT_FAMILY, ClosestSource<R>>)) || (rc = allow(k_cull<R, CULL_GENERAL, RPT_GENERAL, ClosestSource<R>>)) ||
            (rc = allow(k_cull<R, CULL_PIVOT, RPT_FAMILY, ShadowSource<R>>)) || (rc = allow(k_cull<R, CULL_AXIS, RPT_FAMILY, ShadowSource<R>>)))
            return rc;
        return GMLRT_OK;
    }

    static constexpr bool kRobustPrimary = std::is_same<R, float>::value;   // f32: double re-test of borderline primary hits
#ifndef GMLRT_RPT_FAMILY
#define GMLRT_RPT_FAMILY 4
#endif
    static constexpr int RPT_GENERAL = 2, RPT_FAMILY = GMLRT_RPT_FAMILY;                   // rays per thread: 8 cull registers per ray vs 4

    void reset_totals() override { totals_ = TraceTotals(); need_ctr_reset_ = true; ev_used_ = 0; ev_cls_.clear(); }
    void set_profiling(bool on) override { profile_ = on; }

    int fetch_totals(TraceTotals& out, std::string& err) override {
        CU_TRY(cudaSetDevice(dev_));
        Counters c;
        CU_TRY(cudaMemcpy(&c, ctr_, sizeof c, cudaMemcpyDeviceToHost));
        for (size_t k = 0; k + 1 < ev_used_; k++) {          // the stream has been synchronised by the caller
            if (ev_cls_[k] < 0) continue;
            float ms = 0;
            CU_TRY(cudaEventElapsedTime(&ms, ev_pool_[k], ev_pool_[k + 1]));
            totals_.ms_class[ev_cls_[k]] += ms; totals_.launches_class[ev_cls_[k]]++;
        }
        ev_used_ = 0; ev_cls_.clear();
        out = totals_;
        out.rays_reflect = c.rays_reflect; out.rays_refract = c.rays_refract;
        out.rays_shadow = c.rays_shadow; out.shadow_occluded = c.shadow_occluded; out.rays_hit = c.rays_hit;
        out.shadow_tests = c.shadow_tests; out.shadow_flops = c.shadow_flops; out.exact_tests = c.exact_tests; out.exact_flops = c.exact_flops; out.exact_tests_shadow = c.exact_tests_shadow;
        for (int k = 0; k < 3; k++) out.cull_by_family[k] = c.cull_tests[k];
        out.packet_tests = c.packet_tests;
        out.error = c.error; out.overflow = c.overflow;
#ifdef GMLRT_DEBUG_PACKETS
        std::fprintf(stderr, "packets pivot: blocks %llu usable %llu offered %llu survivors %llu | axis: blocks %llu usable %llu offered %llu survivors %llu\n",
                     c.dbg[0], c.dbg[1], c.dbg[2], c.dbg[3], c.dbg[4], c.dbg[5], c.dbg[6], c.dbg[7]);
#endif
        return GMLRT_OK;
    }

    int upload(const HostScene& S, std::string& err) override;
    int trace_rows(int v0, int v1, const RowMap& map, bool counting, std::string& err) override;
    void set_frame_target(uint8_t* target) override { frame_target_ = target; }

private:
    // Scene arrays are bump-allocated from one arena that survives across uploads (cudaMalloc / cudaFree cost more than
    // the whole upload of a 1024-leaf scene); what does not fit is allocated separately and the arena grows next time.
    int scene_alloc(size_t bytes, void** out, std::string& err) {
        bytes = (std::max<size_t>(bytes, 256) + 255) & ~(size_t)255;
        arena_need_ += bytes;
        if (arena_used_ + bytes <= arena_bytes_) { *out = arena_ + arena_used_; arena_used_ += bytes; return GMLRT_OK; }
        CU_TRY(cudaMalloc(out, bytes));
        scene_allocs_.push_back(*out);
        return GMLRT_OK;
    }
    template <typename T> int to_device(const std::vector<T>& v, T** out, std::string& err) {
        *out = nullptr;
        void* p = nullptr;
        int rc = scene_alloc(v.size() * sizeof(T), &p, err);
        if (rc) return rc;
        if (!v.empty()) CU_TRY(cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, stream_));
        *out = reinterpret_cast<T*>(p);
        return GMLRT_OK;
    }
    void free_scene() {                                  // start of an upload: recycle the arena (grown if the last scene spilled)
        for (void* p : scene_allocs_) cudaFree(p);
        scene_allocs_.clear();
        if (arena_need_ > arena_bytes_) {
            if (arena_) cudaFree(arena_);
            arena_ = nullptr; arena_bytes_ = 0;
            const size_t want = std::max<size_t>(arena_need_ + arena_need_ / 2, (size_t)1 << 20);
            void* p = nullptr;
            if (cudaMalloc(&p, want) == cudaSuccess) { arena_ = reinterpret_cast<char*>(p); arena_bytes_ = want; }
            else cudaGetLastError();
        }
        arena_used_ = 0; arena_need_ = 0;
    }
    void free_tiles() {
        if (pool_) cudaFree(pool_);
        pool_ = nullptr; pool_bytes_ = 0;
    }
    void release() {
        cudaSetDevice(dev_);
        free_scene(); free_tiles();
        if (arena_) cudaFree(arena_);
        arena_ = nullptr; arena_bytes_ = 0;
        if (frame_) cudaFree(frame_);
        if (ids_) cudaFree(ids_);
        if (ctr_) cudaFree(ctr_);
        for (cudaEvent_t e : ev_pool_) cudaEventDestroy(e);
        ev_pool_.clear();
        for (cudaEvent_t e : ev_shade_) cudaEventDestroy(e);
        for (int k = 0; k < SHADOW_STREAMS; k++) {
            for (cudaEvent_t e : ev_shadow_[k]) cudaEventDestroy(e);
            ev_shadow_[k].clear();
            if (shadow_stream_[k]) cudaStreamDestroy(shadow_stream_[k]);
            shadow_stream_[k] = nullptr;
        }
        ev_shade_.clear();
        if (stream_) cudaStreamDestroy(stream_);
        frame_ = nullptr; ids_ = nullptr; ctr_ = nullptr; stream_ = nullptr;
    }
    int plan_tiles(const HostScene& S, std::string& err);
public:
    bool shrink_tiles(const HostScene& S, std::string& err) override {
        if (rows_per_tile_ <= 1) { err = "candidate queue overflow with one-row tiles"; return false; }
        tile_shrink_++;
        return plan_tiles(S, err) == GMLRT_OK;
    }
private:

    int dev_, sms_ = 148;
    cudaStream_t stream_ = nullptr;
    // The shadow queries of level l (per light: sort, cull, exact) depend only on k_shade(l), like the closest-hit
    // pipeline of level l + 1, and the lights are independent of each other: they run on SHADOW_STREAMS more streams
    // (light li on stream li mod SHADOW_STREAMS), so the issue-bound cull kernels of one pipeline overlap the
    // memory-bound exact / shade kernels of another and the tails of ~40 launches hide behind each other.
    // k_shade(l + 1) reuses the shadow queue and waits for the shadows of level l; the resolve passes wait for all.
    static constexpr int SHADOW_STREAMS = 2;
    cudaStream_t shadow_stream_[SHADOW_STREAMS] = {nullptr, nullptr};
    bool overlap_shadows_ = true;
    std::vector<cudaEvent_t> ev_shade_, ev_shadow_[SHADOW_STREAMS];   // per level: k_shade done (main stream) / shadow queries done (shadow stream k)
    CandQueue cq_shadow_[SHADOW_STREAMS] = {};             // each shadow stream's own candidate queue
    int* sort_hist_sh_[SHADOW_STREAMS] = {nullptr, nullptr}; int* sort_cursor_sh_[SHADOW_STREAMS] = {nullptr, nullptr};
    DevScene<R> sc_{};
    std::vector<void*> scene_allocs_;
    char* arena_ = nullptr; size_t arena_bytes_ = 0, arena_used_ = 0, arena_need_ = 1 << 20;
    // tile buffers
    void* pool_ = nullptr; size_t pool_bytes_ = 0;
    std::vector<Level<R>> levels_;
    std::vector<unsigned char*> occl_;
    ShadowQueue<R> sq_{};
    int rows_per_tile_ = 0, depth_eff_ = 1;
    uchar4* frame_ = nullptr; size_t frame_px_ = 0;
    uint8_t* frame_target_ = nullptr;       // where the level-0 k_resolve stores instead of frame_ (a peer device's frame), or null
    int32_t* ids_ = nullptr; size_t ids_n_ = 0;
    Counters* ctr_ = nullptr;
    bool profile_ = false;
    std::vector<cudaEvent_t> ev_pool_; size_t ev_used_ = 0; std::vector<int> ev_cls_;
    TraceTotals totals_;
    bool need_ctr_reset_ = true;
    // scan kernel configuration
    int chunk_objs_ = GROUP, resident_ = 1; size_t table_smem_ = 0;
    int cta_pass_ = 7;                                  // CTA-wide first packet pass for: 1 camera rays, 2 spawned rays, 4 shadow rays (GMLRT_CTA_PASS)
    struct Launch { int grid = 148; size_t smem = 0; };
    size_t cfg_smem_ = ~(size_t)0;
    Launch l_primary_, l_closest_, l_shadow_pivot_, l_shadow_axis_;                      // the four k_cull instantiations
    CandQueue cq_{}; int cand_per_ray_ = 4, tile_shrink_ = 0; unsigned long long scene_sig_ = ~0ull, planned_sig_ = ~0ull - 1; int planned_rows_full_ = 0;
    int* sort_hist_ = nullptr; int* sort_cursor_ = nullptr; int sort_nb_ = 2, sort_shift_ = 0, n_regions_ = 1; bool sort_clear_ = true;
    int* key_hist_ = nullptr; int* key_cursor_ = nullptr; int key_nb_ = 6;      // spawned rays by (origin leaf, direction cell); cleared with sort_hist_
    bool sort_spawned_ = true;
    R* spawn_stage_ = nullptr;
    // cull tables (device) and what they were built for: any ray, camera rays, one per light
    std::vector<double> bound_;                                  // host: [n][4] bounding spheres (centre, radius) of the objects
    R* cull_general_ = nullptr; R* cull_eye_ = nullptr; std::vector<R*> cull_light_;
    CullFrame<R> frame_general_{}, frame_eye_{}; std::vector<CullFrame<R>> frame_light_;
    int* shadow_counts_ = nullptr;
    int build_cull_table(int family, const double* v, double offset_abs, R** out, CullFrame<R>& frame, std::string& err);
    std::vector<int> kind_host_;
};

template <typename R>
int Renderer<R>::upload(const HostScene& S, std::string& err) {
    CU_TRY(cudaSetDevice(dev_));
    if (S.depth > GMLRT_MAX_DEPTH) { err = "recursion depth above GMLRT_MAX_DEPTH"; return GMLRT_E_UNSUPPORTED; }
    const bool dbg_t = getenv("GMLRT_DEBUG_UPLOAD") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_a = now();
    CU_TRY(cudaStreamSynchronize(stream_));
    free_scene();
    const double t_b = now();
    {   // a scene of the same shape as the last one keeps the tile height that last fitted (a frame loop re-uploads every frame)
        // everything plan_tiles sizes the pool from: a same-shape scene whose materials turn reflective / transparent, or
        // that gains planes, must be planned afresh (its per-sample footprint grows)
        int n_planes = 0;
        for (const HostLeaf& L : S.leaves) n_planes += L.kind == K_PLANE ? 1 : 0;
        unsigned long long sig = 1469598103934665603ull;
        for (unsigned long long v : {(unsigned long long)S.leaves.size(), (unsigned long long)S.width, (unsigned long long)S.height, (unsigned long long)S.depth,
                                     (unsigned long long)S.lights.size(), (unsigned long long)S.may_branch, (unsigned long long)S.any_reflect,
                                     (unsigned long long)S.any_refract, (unsigned long long)n_planes, (unsigned long long)env_size("GMLRT_CAND_PER_RAY", 4),
                                     (unsigned long long)(env_double("GMLRT_BRANCH_CAP", 2.0) * 1024.0),
                                     (unsigned long long)env_size("GMLRT_TILE_SAMPLES", 64u << 20), (unsigned long long)env_size("GMLRT_POOL_MB", 96 * 1024)})
            sig = (sig ^ v) * 1099511628211ull;
        if (sig != scene_sig_) { tile_shrink_ = 0; scene_sig_ = sig; }
    }
    // Device layout.  Objects (what closestHit / inShadow scan): the plain leaves, stably sorted by kind so that
    // equal kinds keep scan order, then the CSG objects.  After them, the member leaves of the CSG objects.  Every
    // per-leaf array has n_all entries (a CSG object's own entry is unused filler).
    const int n_plain = S.n_plain, n_csg = (int)S.csg.size(), n = n_plain + n_csg;
    const int n_members = (int)S.leaves.size() - n_plain, n_all = n + n_members;
    if (n_all >= HIT_NONE_LEAF) { err = "too many leaves"; return GMLRT_E_UNSUPPORTED; }
    std::vector<int> order(n_plain);
    for (int i = 0; i < n_plain; i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return S.leaves[a].kind < S.leaves[b].kind; });
    std::vector<int> dev_of_host(S.leaves.size());                     // host leaf index -> device leaf index
    for (int d = 0; d < n_plain; d++) dev_of_host[order[d]] = d;
    for (int m = 0; m < n_members; m++) dev_of_host[n_plain + m] = n + m;
    const size_t na = (size_t)std::max(n_all, 1);
    std::vector<R> w2o(na * 12, R(0)), o2w(na * 12, R(0)), fn(na * 18, R(0));
    std::vector<double> w2o64(kRobustPrimary ? na * 12 : 0, 0.0);
    std::vector<int> scan_of(na, 0), surface(na, 0), kind(na, K_SPHERE), obj_flops(std::max(n, 1), 0);
    const int n_cull = std::max((n + GROUP - 1) / GROUP * GROUP, GROUP);
    std::vector<unsigned long long> cum_flops(std::max(n, 1), 0ull);
    DevScene<R> sc{};
    sc.n_leaves = n; sc.n_cull = n_cull; sc.n_all = n_all;
    for (int k = 0; k <= K_OBJ_KINDS; k++) sc.kind_begin[k] = 0;
    bound_.assign((size_t)std::max(n, 1) * 4, 0.0);
    auto put_cull = [&](int d, const double* C, double Rw) {
        bound_[(size_t)d * 4] = C[0]; bound_[(size_t)d * 4 + 1] = C[1]; bound_[(size_t)d * 4 + 2] = C[2]; bound_[(size_t)d * 4 + 3] = Rw;
    };
    auto put_leaf = [&](int d, const HostLeaf& L) {
        for (int k = 0; k < 12; k++) { w2o[(size_t)d * 12 + k] = (R)L.w2o[k]; o2w[(size_t)d * 12 + k] = (R)L.o2w[k]; }
        if (kRobustPrimary) for (int k = 0; k < 12; k++) w2o64[(size_t)d * 12 + k] = L.w2o[k];
        for (int f = 0; f < 6; f++) for (int k = 0; k < 3; k++) fn[(size_t)d * 18 + 3 * f + k] = (R)L.face_normal[f][k];
        surface[d] = (int)L.surface; kind[d] = L.kind;
    };
    for (int d = 0; d < n_plain; d++) {
        const HostLeaf& L = S.leaves[order[d]];
        put_leaf(d, L);
        scan_of[d] = L.scan_index;
        sc.kind_begin[L.kind + 1]++;
        double C[3], Rw;
        leaf_bound(L.kind, L.o2w, C, Rw);
        put_cull(d, C, Rw);
        obj_flops[d] = (int)kFlopsByKind[L.kind];
        cum_flops[L.scan_index] = (unsigned long long)obj_flops[d];
    }
    std::vector<int> csg_prog; std::vector<int2> csg_range(std::max(n_csg, 1), make_int2(0, 0));
    for (int c = 0; c < n_csg; c++) {
        const HostCsgObject& obj = S.csg[c];
        const int d = n_plain + c;
        scan_of[d] = obj.scan_index; kind[d] = K_CSG;
        sc.kind_begin[K_CSG + 1]++;
        put_cull(d, obj.C, obj.R);
        obj_flops[d] = (int)obj.flops;
        cum_flops[obj.scan_index] = (unsigned long long)obj_flops[d];
        csg_range[c] = make_int2((int)csg_prog.size(), (int)obj.postfix.size());
        for (const HostCsgOp& op : obj.postfix) {
            if (op.op < 0) {
                const int dm = dev_of_host[op.leaf];
                put_leaf(dm, S.leaves[op.leaf]);
                scan_of[dm] = obj.scan_index;                     // ids and tie-breaks see the owning object
                csg_prog.push_back(dm);
            } else csg_prog.push_back(-1 - op.op);
        }
    }
    for (int i = 1; i < n; i++) cum_flops[i] += cum_flops[i - 1];
    for (int k = 0; k < K_OBJ_KINDS; k++) sc.kind_begin[k + 1] += sc.kind_begin[k];
    kind_host_ = kind;
    // Sort regions (DevScene::rid_of).  Grid over the bounded objects: about one cell per object, at most 4096 cells,
    // flat extents get a single layer.  "Large" = unbounded, or a bounding sphere wider than 1.5 cell edges.
    std::vector<int> rid_of(na, 0);
    int n_regions = 1;
    {
        double lo3[3] = {INFINITY, INFINITY, INFINITY}, hi3[3] = {-INFINITY, -INFINITY, -INFINITY};
        int n_fin = 0;
        for (int d = 0; d < n; d++) {
            const double* b = &bound_[(size_t)d * 4];
            if (!std::isfinite(b[3])) continue;
            n_fin++;
            for (int k = 0; k < 3; k++) { lo3[k] = std::min(lo3[k], b[k] - b[3]); hi3[k] = std::max(hi3[k], b[k] + b[3]); }
        }
        if (n_fin == 0) for (int k = 0; k < 3; k++) { lo3[k] = -8.0; hi3[k] = 8.0; }
        double ext[3], ext_max = 0;
        for (int k = 0; k < 3; k++) { ext[k] = hi3[k] - lo3[k]; ext_max = std::max(ext_max, ext[k]); }
        ext_max = std::max(ext_max, 1e-6);
        for (int k = 0; k < 3; k++) ext[k] = std::max(ext[k], 1e-3 * ext_max);
        const double target = (double)std::min(std::max(n, 64), 4096);
        double edge = std::cbrt(ext[0] * ext[1] * ext[2] / target);
        int dim[3];
        for (;;) {
            for (int k = 0; k < 3; k++) dim[k] = std::max(1, (int)std::ceil(ext[k] / edge));
            if ((double)dim[0] * dim[1] * dim[2] <= 4096.0) break;
            edge *= 1.05;
        }
        const double large_r = env_double("GMLRT_LARGE_CELLS", 1.5) * edge;
        auto cell_of = [&](const double* c, int* out) {
            for (int k = 0; k < 3; k++) out[k] = std::min(dim[k] - 1, std::max(0, (int)std::floor((c[k] - lo3[k]) / edge)));
        };
        auto morton = [](const int* c) {
            unsigned long long m = 0;
            for (int b = 0; b < 12; b++) for (int k = 0; k < 3; k++) m |= (unsigned long long)((c[k] >> b) & 1) << (3 * b + k);
            return m;
        };
        std::vector<std::pair<unsigned long long, int>> small;
        bool any_large = false;
        int n_plane_grids = 0;
        for (int d = 0; d < n; d++) {
            const double* b = &bound_[(size_t)d * 4];
            if (!std::isfinite(b[3]) && kind[d] == K_PLANE && n_plane_grids < GMLRT_PLANE_GRIDS) {
                // centre = the middle of the bounded objects seen in the plane's object space; scale = 1.5 x their half extent there
                const HostLeaf& L = S.leaves[order[d]];
                const double mid[3] = {0.5 * (lo3[0] + hi3[0]), 0.5 * (lo3[1] + hi3[1]), 0.5 * (lo3[2] + hi3[2])};
                const double cx = L.w2o[0] * mid[0] + L.w2o[1] * mid[1] + L.w2o[2] * mid[2] + L.w2o[3];
                const double cz = L.w2o[8] * mid[0] + L.w2o[9] * mid[1] + L.w2o[10] * mid[2] + L.w2o[11];
                const double stretch = std::sqrt(L.w2o[0] * L.w2o[0] + L.w2o[1] * L.w2o[1] + L.w2o[2] * L.w2o[2]);
                const double scale = std::max(1e-9, 0.75 * ext_max * stretch);
                sc.plane_par[n_plane_grids][0] = (R)cx; sc.plane_par[n_plane_grids][1] = (R)cz; sc.plane_par[n_plane_grids][2] = (R)(1.0 / scale);
                rid_of[d] = -2 - n_plane_grids++;
                continue;
            }
            if (!std::isfinite(b[3]) || b[3] > large_r) { rid_of[d] = -1; any_large = true; continue; }
            int c[3];
            cell_of(b, c);
            small.emplace_back(morton(c), d);
        }
        std::stable_sort(small.begin(), small.end());
        for (size_t k = 0; k < small.size(); k++) rid_of[small[k].second] = (int)k;
        for (int c = 0; c < n_csg; c++) {                             // member leaves: their object's region
            const int d = n_plain + c;
            for (const HostCsgOp& op : S.csg[c].postfix) if (op.op < 0) rid_of[dev_of_host[op.leaf]] = rid_of[d];
        }
        sc.n_small = (int)small.size();
        for (int k = 0; k < 3; k++) { sc.grid_dim[k] = dim[k]; sc.grid_lo[k] = (R)lo3[k]; sc.grid_inv[k] = (R)(1.0 / edge); }
        n_regions = std::max(1, sc.n_small + (any_large ? dim[0] * dim[1] * dim[2] : 0));
        for (int k = 0; k < GMLRT_PLANE_GRIDS; k++) {
            sc.plane_base[k] = n_regions;
            if (k < n_plane_grids) n_regions += GMLRT_PLANE_GRID * GMLRT_PLANE_GRID;
            else { sc.plane_par[k][0] = sc.plane_par[k][1] = R(0); sc.plane_par[k][2] = R(1); }
        }
    }
    n_regions_ = n_regions;
    std::vector<DevLight<R>> lights(S.lights.size());
    for (size_t i = 0; i < S.lights.size(); i++) {
        const gmlrt_light& l = S.lights[i];
        DevLight<R>& d = lights[i];
        d.kind = (int)l.kind;
        for (int k = 0; k < 3; k++) { d.pos[k] = (R)l.pos[k]; d.color[k] = (R)l.color[k]; d.axis[k] = R(0); }
        d.cos_cutoff = R(0); d.exponent = (R)l.exponent;
        if (l.kind == GMLRT_LIGHT_SPOT) {            // EXTENSION
            double a[3] = {l.at[0] - l.pos[0], l.at[1] - l.pos[1], l.at[2] - l.pos[2]};
            normalize3(a);
            for (int k = 0; k < 3; k++) d.axis[k] = (R)a[k];
            d.cos_cutoff = (R)gomath::go_cos(l.cutoff * GM_PI / 180.0);
        }
    }
    std::vector<DevSurface<R>> surfaces(S.surfaces.size());
    for (size_t i = 0; i < S.surfaces.size(); i++) {
        const gmlrt_surface& s = S.surfaces[i];
        surfaces[i].kind = (int)s.kind; surfaces[i].code_off = (int)s.code_off;
        surfaces[i].code_len = (int)s.code_len; surfaces[i].n_regs = (int)s.n_regs;
        for (int k = 0; k < 10; k++) surfaces[i].material[k] = (R)s.material[k];
    }
    std::vector<R> fpool(S.fpool.size());
    for (size_t i = 0; i < S.fpool.size(); i++) fpool[i] = (R)S.fpool[i];
    int rc;
    double* d_w2o64 = nullptr;
    if ((rc = to_device(w2o64, &d_w2o64, err))) return rc;
    unsigned long long* d_cum; int *d_prog, *d_oflops; int2* d_range;
    if ((rc = to_device(cum_flops, &d_cum, err)) || (rc = to_device(csg_prog, &d_prog, err)) ||
        (rc = to_device(csg_range, &d_range, err)) || (rc = to_device(obj_flops, &d_oflops, err))) return rc;
    sc.cum_flops = d_cum; sc.csg_prog = d_prog; sc.csg_range = d_range; sc.obj_flops = d_oflops;
    R *d_w2o, *d_o2w, *d_fn, *d_fpool; int *d_scan, *d_surf, *d_kind, *d_itab; uint32_t* d_code;
    int* d_rid;
    if ((rc = to_device(rid_of, &d_rid, err))) return rc;
    sc.rid_of = d_rid;
    {
        bool simple = n_csg == 0;                             // mode 1: spheres and planes only (face 0, no flip): the low word is the device index
        for (int d = 0; d < n; d++) simple = simple && (kind[d] == K_SPHERE || kind[d] == K_PLANE);
        sc.pickless = (sizeof(R) == 4 && n_all < (1 << 27) && env_size("GMLRT_PICKLESS", 1) != 0) ? (simple ? 1 : 2) : 0;
    }
    DevLight<R>* d_lights; DevSurface<R>* d_surfaces;
    if ((rc = to_device(w2o, &d_w2o, err)) || (rc = to_device(o2w, &d_o2w, err)) || (rc = to_device(fn, &d_fn, err)) ||
        (rc = to_device(scan_of, &d_scan, err)) || (rc = to_device(surface, &d_surf, err)) || (rc = to_device(kind, &d_kind, err)) ||
        (rc = to_device(lights, &d_lights, err)) || (rc = to_device(surfaces, &d_surfaces, err)) ||
        (rc = to_device(fpool, &d_fpool, err)) || (rc = to_device(S.itab, &d_itab, err)) || (rc = to_device(S.code, &d_code, err)))
        return rc;
    sc.w2o64 = kRobustPrimary ? d_w2o64 : nullptr;
    sc.w2o = d_w2o; sc.o2w = d_o2w; sc.fnormal = d_fn; sc.scan_of = d_scan; sc.surface = d_surf; sc.kind = d_kind;
    sc.n_lights = (int)lights.size(); sc.lights = d_lights; sc.surfaces = d_surfaces;
    sc.fpool = d_fpool; sc.itab = d_itab; sc.code = d_code;
    sc.n_fpool = (int)fpool.size(); sc.n_itab = (int)S.itab.size();
    sc.W = S.width; sc.H = S.height; sc.depth = S.depth; sc.vw = S.vw; sc.vh = S.vh;
    sc.cam_su = S.vw / (double)(S.width - 1); sc.cam_sv = S.vh / (double)(S.height - 1);
    // sort key of spawned rays: 6 k^2 direction cells per origin leaf, at most KEY_BINS_MAX bins in all
    sc.key_k = (int)env_size("GMLRT_KEY_K", 4); sc.key_k = std::max(1, std::min(sc.key_k, 8));
    sc.key_shift = 0;
    while ((size_t)(((n_regions - 1) >> sc.key_shift) + 1) * 6 * sc.key_k * sc.key_k > (size_t)KEY_BINS_MAX) sc.key_shift++;
    key_nb_ = (((n_regions - 1) >> sc.key_shift) + 1) * 6 * sc.key_k * sc.key_k;
    sort_spawned_ = env_size("GMLRT_SORT_SPAWNED", 1) != 0;
    cta_pass_ = (int)env_size("GMLRT_CTA_PASS", 7);
    for (int k = 0; k < 3; k++) { sc.ambient[k] = (R)S.h.ambient[k]; sc.bg0[k] = (R)S.h.bg_start[k]; sc.bg1[k] = (R)S.h.bg_end[k]; }
    sc_ = sc;
    // cull tables: any ray; camera rays (their lines pass through the eye, raytracer.go:605-609); one per light
    const double eye[3] = {0.0, 0.0, -1.0};
    if ((rc = build_cull_table(CULL_GENERAL, nullptr, 0.0, &cull_general_, frame_general_, err)) ||
        (rc = build_cull_table(CULL_PIVOT, eye, 1e-5, &cull_eye_, frame_eye_, err))) return rc;
    cull_light_.assign(S.lights.size(), nullptr);
    frame_light_.assign(S.lights.size(), CullFrame<R>{});
    for (size_t i = 0; i < S.lights.size(); i++) {
        const gmlrt_light& l = S.lights[i];
        if (l.kind == GMLRT_LIGHT_DIRECTIONAL) {
            double a[3] = {-l.pos[0], -l.pos[1], -l.pos[2]};                 // towards the light (light_at)
            normalize3(a);
            rc = build_cull_table(CULL_AXIS, a, 1e-5, &cull_light_[i], frame_light_[i], err);
        } else {
            // a shadow ray starts at P + N*1e-4 but points along normalize(Lpos - P) (raytracer.go:412-413), so its
            // line misses the light by up to 1e-4 |N|: widen every sphere by that much
            rc = build_cull_table(CULL_PIVOT, l.pos, 1.5e-4, &cull_light_[i], frame_light_[i], err);
        }
        if (rc) return rc;
    }
    {
        void* p = nullptr;
        if ((rc = scene_alloc(std::max<size_t>(S.lights.size(), 1) * sizeof(int), &p, err))) return rc;
        shadow_counts_ = reinterpret_cast<int*>(p);
        CU_TRY(cudaMemsetAsync(shadow_counts_, 0, std::max<size_t>(S.lights.size(), 1) * sizeof(int), stream_));
    }
    const double t_c = now();
    CU_TRY(cudaStreamSynchronize(stream_));      // host vectors go out of scope
    const double t_d = now();

    // frame + ids
    const size_t px = (size_t)S.width * S.height;
    if (px != frame_px_) {
        if (frame_) cudaFree(frame_);
        frame_ = nullptr; frame_px_ = 0;
        CU_TRY(cudaMalloc(&frame_, px * sizeof(uchar4)));
        frame_px_ = px;
    }
    const int prc = plan_tiles(S, err);
    if (dbg_t) fprintf(stderr, "[gmlrt upload dev %d] sync+free %.3f  host build + enqueue copies %.3f  copy sync %.3f  frame+plan_tiles %.3f ms\n",
                       dev_, t_b - t_a, t_c - t_b, t_d - t_c, now() - t_d);
    return prc;
}

// One 4-value record per object for the cull phase of the scan kernels (see "two-phase scan"): the bounding
// sphere (centre C, radius Rw) in the form the family's reject() test reads, with every margin folded in so that
// a reject is always a true miss: relative radius inflation, an absolute widening `offset_abs` (rays that only
// nearly pass through the pivot / along the axis) and a rounding allowance relative to the squared magnitudes.
template <typename R>
int Renderer<R>::build_cull_table(int family, const double* v, double offset_abs, R** out, CullFrame<R>& frame, std::string& err) {
    const int n = sc_.n_leaves, n_cull = sc_.n_cull;
    std::vector<R> tab((size_t)n_cull * 4);
    frame = CullFrame<R>{};
    frame.family = family;
    double e1[3] = {1, 0, 0}, e2[3] = {0, 1, 0};
    if (family == CULL_AXIS) {                       // orthonormal basis of the plane orthogonal to the unit axis v
        const int k = std::fabs(v[0]) < std::fabs(v[1]) ? (std::fabs(v[0]) < std::fabs(v[2]) ? 0 : 2) : (std::fabs(v[1]) < std::fabs(v[2]) ? 1 : 2);
        double u[3] = {0, 0, 0};
        u[k] = 1.0;
        const double d = u[0] * v[0] + u[1] * v[1] + u[2] * v[2];
        for (int i = 0; i < 3; i++) e1[i] = u[i] - d * v[i];
        normalize3(e1);
        e2[0] = v[1] * e1[2] - v[2] * e1[1]; e2[1] = v[2] * e1[0] - v[0] * e1[2]; e2[2] = v[0] * e1[1] - v[1] * e1[0];
        normalize3(e2);
        for (int i = 0; i < 3; i++) { frame.e1[i] = (R)e1[i]; frame.e2[i] = (R)e2[i]; frame.axis[i] = (R)v[i]; }
        for (int i = 0; i < 3; i++) { e1[i] = (double)frame.e1[i]; e2[i] = (double)frame.e2[i]; }      // what the device projects rays with
    }
    const double rel = CullMargin<R>::rel;
    for (int d = 0; d < n_cull; d++) {
        R* rec = &tab[(size_t)d * 4];
        if (d >= n) { rec[0] = rec[1] = rec[2] = R(0); rec[3] = (R)INFINITY; if (family == CULL_AXIS) rec[2] = (R)INFINITY; continue; }   // padding: always rejected
        const double* C = &bound_[(size_t)d * 4];
        const double Rw = bound_[(size_t)d * 4 + 3];
        const bool unbounded = std::isinf(Rw) || !(Rw == Rw);
        const double Rm = Rw * (1.0 + CullMargin<R>::radius) + offset_abs;
        double K;
        if (family == CULL_AXIS) {
            const R cu = (R)(C[0] * e1[0] + C[1] * e1[1] + C[2] * e1[2]), cv = (R)(C[0] * e2[0] + C[1] * e2[1] + C[2] * e2[2]);
            const double cc = C[0] * C[0] + C[1] * C[1] + C[2] * C[2];
            K = -(Rm * Rm) - rel * (cc + 1.0);
            rec[0] = unbounded ? R(0) : cu; rec[1] = unbounded ? R(0) : cv;
            // word 3: how far along the rays' direction the object reaches (packet stage: nothing behind every origin can occlude)
            const double front = C[0] * v[0] + C[1] * v[1] + C[2] * v[2] + Rm + 1e-3 * (std::sqrt(cc) + 1.0);
            rec[3] = (unbounded || !(front == front)) ? (R)INFINITY : (R)front;
            if ((double)rec[3] < front) rec[3] = std::nextafter(rec[3], (R)INFINITY);
            if (unbounded || !(K == K)) K = -INFINITY;
            rec[2] = (R)K;
            if ((double)rec[2] > K) rec[2] = std::nextafter(rec[2], (R)-INFINITY);
        } else {
            const double px = C[0] - (v ? v[0] : 0.0), py = C[1] - (v ? v[1] : 0.0), pz = C[2] - (v ? v[2] : 0.0);
            const R cx = (R)px, cy = (R)py, cz = (R)pz;
            const double cc = (double)cx * cx + (double)cy * cy + (double)cz * cz;
            K = cc - Rm * Rm - rel * (cc + 1.0);
            if (unbounded || !(K == K)) K = -INFINITY;                       // plane or degenerate: never rejected
            rec[0] = unbounded ? R(0) : cx; rec[1] = unbounded ? R(0) : cy; rec[2] = unbounded ? R(0) : cz; rec[3] = (R)K;
            if ((double)rec[3] > K) rec[3] = std::nextafter(rec[3], (R)-INFINITY);
        }
    }
    return to_device(tab, out, err);
}

template <typename R>
int Renderer<R>::plan_tiles(const HostScene& S, std::string& err) {
    const int W = S.width, H = S.height, nl = (int)S.lights.size();
    depth_eff_ = (S.any_reflect || S.any_refract) ? S.depth : 1;   // no material can spawn a child ray
    std::vector<double> mult(depth_eff_);
    double sum_mult = 0, max_mult = 0;
    for (int l = 0; l < depth_eff_; l++) {
        // Binary ray tree only when a material asks for it - and then the levels are sized for `branch_cap` rays per
        // camera ray, not for the worst case 2^l (which would cut an 8K depth-5 frame into 44 thin tiles): k_shade
        // refuses children beyond a level's capacity and raises Counters::overflow, the frame is traced again with
        // smaller tiles over the same buffers.
        mult[l] = S.may_branch ? std::min(std::ldexp(1.0, std::min(l, 30)), std::max(1.0, env_double("GMLRT_BRANCH_CAP", 2.0))) : 1.0;
        sum_mult += mult[l]; max_mult = std::max(max_mult, mult[l]);
    }
    // candidate queue: (ray, object, T) per survivor of the cull phase.  Unbounded objects (planes) survive every ray;
    // beyond them a handful per ray is typical (about 1 at 1024 spheres); an overflow retraces with smaller tiles.
    int n_unbounded = 0;
    for (int d = 0; d < sc_.n_leaves; d++) n_unbounded += std::isinf(bound_[(size_t)d * 4 + 3]) ? 1 : 0;
    cand_per_ray_ = std::min<int>(std::max(sc_.n_leaves, 1), (int)env_size("GMLRT_CAND_PER_RAY", 4) + n_unbounded);
    const double cand_bytes = cand_per_ray_ * (8.0 + 8.0 * SHADOW_STREAMS + sizeof(R));   // queues: closest (ray, object, T), one (ray, object) per shadow stream
    const double node_bytes = (2 * RAY_WORDS + NREC + 8) * sizeof(R) + 12 + nl;      // ray record (+ its staging copy), rec, child, hit, order, occl
    const double shadow_bytes = nl * (8.0 * sizeof(R) + 20);
    const double per_sample = sum_mult * node_bytes + max_mult * (shadow_bytes + cand_bytes) + 3 * sizeof(R);
    // A scene of the same shape as the last one keeps its plan: free memory wobbles by a few MB between uploads, and a
    // budget-bound plan that grew by one row would free and reallocate the whole pool (~90 GB: 200 ms) inside a frame
    // loop; cudaMemGetInfo itself was seen to take 6-30 ms now and then with a pool of that size.
    int rows_full = 0;
    if (pool_ && planned_sig_ == scene_sig_ && planned_rows_full_ > 0) rows_full = planned_rows_full_;
    else {
        size_t free_b = 0, total_b = 0;
        CU_TRY(cudaMemGetInfo(&free_b, &total_b));
        free_b += pool_bytes_;
        // One tile per frame when it fits: fewer launches, and the sorted bins of the spawned / shadow rays are as full as
        // the frame allows (4K spheres1024: 11.6 ms in two tiles, 10.5 ms in one).  180 GB of HBM3e is there to be used.
        double budget = std::min((double)free_b * 0.6, (double)env_size("GMLRT_POOL_MB", 96 * 1024) * 1048576.0);
        double max_samples = std::min(budget / per_sample, (double)env_size("GMLRT_TILE_SAMPLES", 64u << 20));
        max_samples = std::min(max_samples, 2.0e9 / (max_mult * std::max(nl, 1)));       // int32 indexing of node*n_lights
        rows_full = (int)std::min<double>(H, std::floor(max_samples / (4.0 * W)));
        if (rows_full < 1) { err = "image row does not fit the device queue budget"; return GMLRT_E_OOM; }
    }
    planned_sig_ = scene_sig_; planned_rows_full_ = rows_full;
    const int rows = std::max(1, rows_full >> tile_shrink_);    // after an overflow: same buffers, fewer rays
    rows_per_tile_ = rows;
    const size_t S0 = (size_t)rows * W * 4, S0_full = (size_t)rows_full * W * 4;
    // carve one pool
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align(off + bytes); return o; };
    struct LvOff { size_t ray, parent, hit, t, order, rec, child, occl; size_t cap; };
    std::vector<LvOff> lo(depth_eff_);
    size_t max_cap = 0, stage_cap = 0;
    for (int l = 0; l < depth_eff_; l++) {
        size_t cap = (size_t)std::min(std::min((double)S0 * std::ldexp(1.0, std::min(l, 30)), (double)S0_full * mult[l]), 2.0e9);   // never above the true worst case
        lo[l].cap = cap; max_cap = std::max(max_cap, cap);
        lo[l].ray = take(RAY_WORDS * cap * sizeof(R)); lo[l].parent = 0; lo[l].hit = take(cap * 4);
        if (l > 0) stage_cap = std::max(stage_cap, cap);
        lo[l].t = take(cap * sizeof(R)); lo[l].order = take(cap * 8); lo[l].rec = take(NREC * cap * sizeof(R)); lo[l].child = take(6 * cap * sizeof(R));
        lo[l].occl = take(cap * std::max(nl, 1));
    }
    const size_t sq_cap = max_cap * std::max(nl, 1);
    size_t o_sq = take(8 * sq_cap * sizeof(R)), o_skip = take(sq_cap * 4), o_key = take(sq_cap * 4), o_tgt = take(sq_cap * 4), o_first = take(sq_cap * 4), o_perm = take(sq_cap * 4), o_sort = take(2 * SHADOW_STREAMS * SORT_BINS_MAX * sizeof(int));
    size_t o_keysort = take(2 * (size_t)KEY_BINS_MAX * sizeof(int));
    size_t o_stage = take(RAY_WORDS * stage_cap * sizeof(R));         // spawned rays wait here until k_keysort_scatter stores them sorted
    const size_t cand_cap = std::min<size_t>((size_t)((double)S0_full * max_mult * cand_per_ray_), (size_t)2000000000);
    size_t o_cray = take(cand_cap * 4), o_cobj = take(cand_cap * 4), o_ct = take(cand_cap * sizeof(R)), o_ccount = take(256);
    size_t o_cray2[SHADOW_STREAMS], o_cobj2[SHADOW_STREAMS], o_ccount2[SHADOW_STREAMS];              // the shadow streams' queues (no T)
    for (int k = 0; k < SHADOW_STREAMS; k++) { o_cray2[k] = take(cand_cap * 4); o_cobj2[k] = take(cand_cap * 4); o_ccount2[k] = take(256); }
    if (off > pool_bytes_) {
        free_tiles();
        CU_TRY(cudaMalloc(&pool_, off));
        pool_bytes_ = off;
    }
    char* base = reinterpret_cast<char*>(pool_);
    levels_.assign(depth_eff_, Level<R>{});
    occl_.assign(depth_eff_, nullptr);
    for (int l = 0; l < depth_eff_; l++) {
        Level<R>& L = levels_[l];
        L.cap = (int)lo[l].cap;
        L.ray = (R*)(base + lo[l].ray); L.hit = (int*)(base + lo[l].hit);
        L.t = (R*)(base + lo[l].t); L.order = (int*)(base + lo[l].order); L.rec = (R*)(base + lo[l].rec); L.child = (R*)(base + lo[l].child);
        occl_[l] = (unsigned char*)(base + lo[l].occl);
    }
    sq_.q = (R*)(base + o_sq); sq_.skip = (int*)(base + o_skip); sq_.key = (int*)(base + o_key); sq_.target = (int*)(base + o_tgt); sq_.cap = (int)sq_cap;
    sq_.per_light = (int)max_cap; sq_.count = shadow_counts_; sq_.first = (int*)(base + o_first); sq_.perm = (int*)(base + o_perm);
    sort_hist_ = (int*)(base + o_sort); sort_cursor_ = sort_hist_ + SORT_BINS_MAX; sort_clear_ = true;
    key_hist_ = (int*)(base + o_keysort); key_cursor_ = key_hist_ + KEY_BINS_MAX;
    spawn_stage_ = (R*)(base + o_stage);
    sort_nb_ = std::min(n_regions_ + 1, SORT_BINS_MAX); sort_shift_ = 0;        // shadow rays by the region of the shaded point
    while (((n_regions_ - 1) >> sort_shift_) + 2 > sort_nb_) sort_shift_++;
    cq_.ray = (int*)(base + o_cray); cq_.obj = (int*)(base + o_cobj); cq_.t = (void*)(base + o_ct); cq_.count = (int*)(base + o_ccount);
    cq_.cap = (int)cand_cap;
    for (int k = 0; k < SHADOW_STREAMS; k++) {
        cq_shadow_[k].ray = (int*)(base + o_cray2[k]); cq_shadow_[k].obj = (int*)(base + o_cobj2[k]); cq_shadow_[k].t = nullptr;
        cq_shadow_[k].count = (int*)(base + o_ccount2[k]); cq_shadow_[k].cap = (int)cand_cap;
        sort_hist_sh_[k] = sort_hist_ + (size_t)2 * k * SORT_BINS_MAX; sort_cursor_sh_[k] = sort_hist_sh_[k] + SORT_BINS_MAX;
    }

    // scan kernels: resident cull table when it fits, else two streaming buffers (chunks of whole groups)
    const size_t rec_bytes = 4 * sizeof(R);
    const size_t resident_max = env_size("GMLRT_RESIDENT_KB", 128) * 1024;
    const size_t chunk_bytes = std::max<size_t>(env_size("GMLRT_CHUNK_KB", 32) * 1024, GROUP * rec_bytes);
    const size_t all = (size_t)sc_.n_cull * rec_bytes;
    if (all <= resident_max) { resident_ = 1; chunk_objs_ = sc_.n_cull; table_smem_ = all; }
    else { resident_ = 0; chunk_objs_ = (int)(chunk_bytes / rec_bytes) / GROUP * GROUP; table_smem_ = 2 * (size_t)chunk_objs_ * rec_bytes; }
    if (cfg_smem_ == SURV_BYTES + table_smem_) return GMLRT_OK;      // launch shapes depend on the staged table size only
    cfg_smem_ = SURV_BYTES + table_smem_;
    auto cfg = [&](auto kernel, Launch& L) -> int {
        int occ = 1;
        L.smem = SURV_BYTES + table_smem_;
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, SCAN_THREADS, L.smem));
        L.grid = sms_ * std::max(occ, 1);
        return GMLRT_OK;
    };
    int rc;
    if ((rc = cfg(k_cull<R, CULL_PIVOT, RPT_FAMILY, ClosestSource<R>>, l_primary_)) ||
        (rc = cfg(k_cull<R, CULL_GENERAL, RPT_GENERAL, ClosestSource<R>>, l_closest_)) ||
        (rc = cfg(k_cull<R, CULL_PIVOT, RPT_FAMILY, ShadowSource<R>>, l_shadow_pivot_)) ||
        (rc = cfg(k_cull<R, CULL_AXIS, RPT_FAMILY, ShadowSource<R>>, l_shadow_axis_)))
        return rc;
    return GMLRT_OK;
}

template <typename R>
int Renderer<R>::trace_rows(int y0, int y1, const RowMap& map, bool counting, std::string& err) {
    CU_TRY(cudaSetDevice(dev_));
    const int W = sc_.W;
    if (need_ctr_reset_) { CU_TRY(cudaMemsetAsync(ctr_, 0, sizeof(Counters), stream_)); need_ctr_reset_ = false; }
    if (counting && ids_n_ != (size_t)W * sc_.H * 4) {
        if (ids_) cudaFree(ids_);
        ids_ = nullptr; ids_n_ = 0;
        CU_TRY(cudaMalloc(&ids_, (size_t)W * sc_.H * 4 * sizeof(int32_t)));
        ids_n_ = (size_t)W * sc_.H * 4;
    }
    const int wide = sms_ * 16;
    // one stream when per-kernel events or the counting pass want the launches serialised
    const bool two_streams = overlap_shadows_ && !profile_ && !counting && sc_.n_lights > 0;
    const int nss = two_streams ? std::min<int>(SHADOW_STREAMS, sc_.n_lights) : 1;       // shadow streams in use
    while (two_streams && (int)ev_shade_.size() < depth_eff_) {
        cudaEvent_t a;
        CU_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        ev_shade_.push_back(a);
        for (int k = 0; k < SHADOW_STREAMS; k++) { CU_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming)); ev_shadow_[k].push_back(a); }
    }
    CamSrc cam{y0, map};                                    // rows of the current tile's camera rays (k_gen, double-precision re-tests)
    // exact kernels by kind (see ExactKinds): spheres inline at full occupancy, every other kind through exact_hit
    const bool have_spheres = sc_.kind_begin[K_SPHERE + 1] > 0;
    const bool have_simple = sc_.kind_begin[K_CSG] > sc_.kind_begin[K_SPHERE + 1], have_csg = sc_.n_leaves > sc_.kind_begin[K_CSG] || sc_.n_leaves == 0;
    const int n_exact_flavours = (have_spheres ? 1 : 0) + (have_simple ? 1 : 0) + (have_csg ? 1 : 0);
    auto exact_closest = [&](int l, bool count) -> int {
        const int grid = sms_ * 16;
        if (have_spheres) {
            if (count) k_exact_closest<R, false, true, EXACT_SPHERES><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
            else k_exact_closest<R, false, false, EXACT_SPHERES><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
        }
        if (have_simple) {
            if (count) k_exact_closest<R, false, true, EXACT_SIMPLE><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
            else k_exact_closest<R, false, false, EXACT_SIMPLE><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
        }
        if (have_csg) {
            if (count) k_exact_closest<R, false, true, EXACT_CSG><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
            else k_exact_closest<R, false, false, EXACT_CSG><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
        }
        totals_.launches += n_exact_flavours - 1;
        return GMLRT_OK;
    };
    // profiling: one event before every launch; launches on one stream serialise, so consecutive events bracket a kernel
    auto mark = [&](int cls) -> int {
        if (!profile_) return GMLRT_OK;
        if (ev_used_ == ev_pool_.size()) { cudaEvent_t e; CU_TRY(cudaEventCreate(&e)); ev_pool_.push_back(e); }
        CU_TRY(cudaEventRecord(ev_pool_[ev_used_++], stream_));
        ev_cls_.push_back(cls);
        return GMLRT_OK;
    };
    int mrc;
#define MARK(cls) do { if ((mrc = mark(cls))) return mrc; } while (0)
    for (int ty = y0; ty < y1; ty += rows_per_tile_) {
        const int rows = std::min(rows_per_tile_, y1 - ty);
        cam.y0 = ty;
        const int S = rows * W * 4;
        MARK(KC_GEN);
        k_gen<R><<<(S + 255) / 256, 256, 0, stream_>>>(sc_, levels_[0], ctr_, ty, rows, map);
        totals_.launches++;
        totals_.rays_primary += (unsigned long long)S;
        for (int l = 0; l < depth_eff_; l++) {
            const int cap_blocks = (levels_[l].cap + SHADE_THREADS - 1) / SHADE_THREADS;
            const int exact_grid = sms_ * 16;
            CU_TRY(cudaMemsetAsync(cq_.count, 0, 2 * sizeof(int), stream_));
            ClosestSource<R> csrc{levels_[l]};
            if (sort_clear_) {
                CU_TRY(cudaMemsetAsync(sort_hist_, 0, 2 * SHADOW_STREAMS * SORT_BINS_MAX * sizeof(int), stream_));
                CU_TRY(cudaMemsetAsync(key_hist_, 0, KEY_BINS_MAX * sizeof(int), stream_));
                sort_clear_ = false;
            }
            if (l > 0 && sort_spawned_) {                  // rays by (origin leaf, direction cell): keys in lv.hit, records in the staging buffer
                MARK(KC_SORT);
                k_keysort_hist<<<sms_ * 8, 256, 0, stream_>>>(levels_[l].hit, &ctr_->count[l], key_nb_, key_hist_);
                k_sort_scan<<<1, 1024, 0, stream_>>>(key_nb_, key_hist_, key_cursor_);
                k_keysort_scatter<R><<<sms_ * 8, 256, 0, stream_>>>(levels_[l].hit, &ctr_->count[l], key_nb_, key_cursor_, spawn_stage_, levels_[l].ray);
                totals_.launches += 3;
            }
            if (l == 0) {                                  // camera rays: PIVOT family through the eye, robust re-tests in f32
                MARK(KC_CLOSEST_PRIMARY);
                k_cull<R, CULL_PIVOT, RPT_FAMILY, ClosestSource<R>><<<l_primary_.grid, SCAN_THREADS, l_primary_.smem, stream_>>>(
                    csrc, &ctr_->count[l], cull_eye_, frame_eye_, sc_.n_cull, cq_, ctr_, chunk_objs_, resident_ | ((cta_pass_ & 1) ? 2 : 0));
                MARK(KC_EXACT);
                if (kRobustPrimary) {
                    // float32 spheres inline; then the double-precision pass: other kinds + the spheres the first pass deferred
                    if (counting) {
                        if (have_spheres) k_exact_closest<R, kRobustPrimary, true, EXACT_SPHERES><<<exact_grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
                        k_exact_closest<R, kRobustPrimary, true, EXACT_REST><<<exact_grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
                    } else {
                        if (have_spheres) k_exact_closest<R, kRobustPrimary, false, EXACT_SPHERES><<<exact_grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
                        k_exact_closest<R, kRobustPrimary, false, EXACT_REST><<<exact_grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_, ctr_, cam);
                    }
                    totals_.launches += have_spheres ? 1 : 0;
                } else {
                    int erc = exact_closest(l, counting);
                    if (erc) return erc;
                }
            } else {
                MARK(KC_CLOSEST);
                k_cull<R, CULL_GENERAL, RPT_GENERAL, ClosestSource<R>><<<l_closest_.grid, SCAN_THREADS, l_closest_.smem, stream_>>>(
                    csrc, &ctr_->count[l], cull_general_, frame_general_, sc_.n_cull, cq_, ctr_, chunk_objs_, resident_ | ((cta_pass_ & 2) ? 2 : 0));
                MARK(KC_EXACT);
                int erc = exact_closest(l, counting);
                if (erc) return erc;
            }
            if (sizeof(R) == 8) k_closest_order<R><<<exact_grid, 256, 0, stream_>>>(sc_, levels_[l], cq_, ctr_);
            else totals_.launches--;
            if (!sc_.pickless) k_closest_pick<R><<<exact_grid, 256, 0, stream_>>>(sc_, levels_[l], cq_, ctr_);
            else totals_.launches--;
            totals_.launches += 3;
            const bool spawn = l + 1 < depth_eff_;
            if (two_streams && l > 0) for (int k = 0; k < nss; k++) CU_TRY(cudaStreamWaitEvent(stream_, ev_shadow_[k][l - 1], 0));     // k_shade refills the shadow queue
            MARK(KC_SHADE);
            auto shade = sc_.pickless == 1 ? k_shade<R, sizeof(R) == 4 ? 1 : 0> : sc_.pickless ? k_shade<R, sizeof(R) == 4 ? 2 : 0> : k_shade<R, 0>;
            shade<<<std::min(cap_blocks, wide), SHADE_THREADS, 0, stream_>>>(
                sc_, levels_[l], spawn ? levels_[l + 1] : levels_[l], sq_, occl_[l], ctr_, l, spawn ? 1 : 0,
                (counting && l == 0) ? ids_ : nullptr, ty, map,
                spawn ? (sort_spawned_ ? spawn_stage_ : levels_[l + 1].ray) : nullptr);
            totals_.launches += 2;
            // shadow queries of this level: on the shadow stream, overlapping the next level's closest-hit pipeline
            if (two_streams) {
                CU_TRY(cudaEventRecord(ev_shade_[l], stream_));
                for (int k = 0; k < nss; k++) CU_TRY(cudaStreamWaitEvent(shadow_stream_[k], ev_shade_[l], 0));
            }
            for (int li = 0; li < sc_.n_lights; li++) {    // per light: its shadow rays share a pivot (or an axis)
                const bool axis = frame_light_[li].family == CULL_AXIS;
                const int sk = li % nss;
                cudaStream_t ss = two_streams ? shadow_stream_[sk] : stream_;
                const CandQueue& cqs = two_streams ? cq_shadow_[sk] : cq_;
                int* const s_hist = sort_hist_sh_[sk]; int* const s_cursor = sort_cursor_sh_[sk];
                ShadowSource<R> ssrc{sq_, li * sq_.per_light};
                CU_TRY(cudaMemsetAsync(cqs.count, 0, 2 * sizeof(int), ss));
                MARK(KC_SORT);
                k_sort_hist<R><<<sms_ * 4, SORT_THREADS, 0, ss>>>(sq_, li, sort_shift_, sort_nb_, s_hist, ctr_);
                k_sort_scan<<<1, 1024, 0, ss>>>(sort_nb_, s_hist, s_cursor);
                k_sort_scatter<R><<<sms_ * 4, SORT_THREADS, 0, ss>>>(sq_, li, sort_shift_, sort_nb_, s_cursor);
                totals_.launches += 3;
                MARK(KC_SHADOW);
                if (axis) k_cull<R, CULL_AXIS, RPT_FAMILY, ShadowSource<R>><<<l_shadow_axis_.grid, SCAN_THREADS, l_shadow_axis_.smem, ss>>>(
                    ssrc, &shadow_counts_[li], cull_light_[li], frame_light_[li], sc_.n_cull, cqs, ctr_, chunk_objs_, resident_ | ((cta_pass_ & 4) ? 2 : 0));
                else k_cull<R, CULL_PIVOT, RPT_FAMILY, ShadowSource<R>><<<l_shadow_pivot_.grid, SCAN_THREADS, l_shadow_pivot_.smem, ss>>>(
                    ssrc, &shadow_counts_[li], cull_light_[li], frame_light_[li], sc_.n_cull, cqs, ctr_, chunk_objs_, resident_ | ((cta_pass_ & 4) ? 2 : 0));
                MARK(KC_EXACT);
                const int sbase = li * sq_.per_light;
                if (counting) {
                    if (have_spheres) k_exact_shadow<R, true, EXACT_SPHERES><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_, sbase, cqs, ctr_, occl_[l]);
                    if (have_simple) k_exact_shadow<R, true, EXACT_SIMPLE><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_, sbase, cqs, ctr_, occl_[l]);
                    if (have_csg) k_exact_shadow<R, true, EXACT_CSG><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_, sbase, cqs, ctr_, occl_[l]);
                    k_shadow_finish<R, true><<<wide, 256, 0, ss>>>(sc_, sq_, li, occl_[l], ctr_);
                } else {                                   // the exact kernels set the occlusion flags themselves
                    if (have_spheres) k_exact_shadow<R, false, EXACT_SPHERES><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_, sbase, cqs, ctr_, occl_[l]);
                    if (have_simple) k_exact_shadow<R, false, EXACT_SIMPLE><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_, sbase, cqs, ctr_, occl_[l]);
                    if (have_csg) k_exact_shadow<R, false, EXACT_CSG><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_, sbase, cqs, ctr_, occl_[l]);
                    totals_.launches--;
                }
                totals_.launches += n_exact_flavours - 1;
                totals_.launches += 3;
                CU_TRY(cudaMemsetAsync(&shadow_counts_[li], 0, sizeof(int), ss));
            }
            if (two_streams) for (int k = 0; k < nss; k++) CU_TRY(cudaEventRecord(ev_shadow_[k][l], shadow_stream_[k]));
        }
        if (two_streams) for (int k = 0; k < nss; k++) CU_TRY(cudaStreamWaitEvent(stream_, ev_shadow_[k][depth_eff_ - 1], 0));   // the shadow streams are in order: every level is done
        for (int l = depth_eff_ - 1; l >= 0; l--) {
            const int cap_blocks = (levels_[l].cap + SHADE_THREADS - 1) / SHADE_THREADS;
            MARK(KC_RESOLVE);
            auto resolve = sc_.pickless ? k_resolve<R, sizeof(R) == 4 ? 1 : 0> : k_resolve<R, 0>;
            resolve<<<std::min(cap_blocks, wide), SHADE_THREADS, 0, stream_>>>(
                sc_, levels_[l], l > 0 ? levels_[l - 1] : levels_[l], occl_[l], ctr_, l, frame_target_ ? (uchar4*)frame_target_ : frame_, ty, map, l + 1 < depth_eff_ ? 1 : 0);
            totals_.launches++;
        }
    }
    MARK(-1);
#undef MARK
    CU_TRY(cudaGetLastError());
    return GMLRT_OK;
}

}  // namespace gmlrt
