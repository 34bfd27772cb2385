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
    int rec_compact;               // every surface is a constant material: the shading record holds the surface index instead of its 9 material words
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
    unsigned w_mul; int w_shift;   // n / W for n < 2^31 as (n * w_mul) >> w_shift (FastDiv): pixel index -> (row, column) without a runtime division
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
                  //          the previous level | (1<<31 for the refraction slot), and (word 7) the class key.  32 bytes in f32: one sector per random read
    int* hit;     // [cap] device leaf index | face << 28, or -1
    R* t;         // [cap] T of the hit; during the scan: bit pattern of the smallest T so far (all ones = none)
    int* order;   // scan scratch.  f64: [cap] smallest scan index among the hits that have the smallest T; f32: [cap] 64-bit keys
                  //               (bits of T) << 32 | scan index, folded by one atomicMin (fold_hit)
    R* rec;       // [NREC][cap]
    R* child;     // [6][cap] colour returned by the reflection / refraction child
    int* hits;    // [cap] the rays of the level that hit something (k_hitlist): everything per HIT - rec, child, the occlusion flags -
                  //       is indexed by the position in this list, so the shading kernels run on dense warps and stream
    R* samp;      // level 0 only: [3][cap] colour of a camera ray that hit (by ray index), for the pixel pass
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
    int hits[GMLRT_MAX_DEPTH + 1];      // ... of which hit something (entries of Level::hits)
    int error;
    int overflow;                       // a cull launch ran out of candidate queue: retrace with smaller tiles
    unsigned long long cull_tests[3];   // executed per-ray cull tests by family (GENERAL, PIVOT, AXIS)
    unsigned long long packet_tests;    // executed per-warp packet tests (one lane x one object)
    unsigned long long dbg[8];          // GMLRT_DEBUG_PACKETS: warp blocks, usable, objects offered, survivors
    unsigned long long rays_reflect, rays_refract, rays_shadow, shadow_occluded, rays_hit;
    unsigned long long shadow_tests, shadow_flops;   // credited: the reference's scan up to its first occluder (counting mode)
    unsigned long long exact_tests, exact_flops;     // executed second-phase tests and their table flops (counting mode)
    unsigned long long exact_tests_shadow;           // ... the part of exact_tests that k_exact_shadow ran
    unsigned long long rays_hit_primary;             // camera rays that hit
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
__device__ __forceinline__ void store_ray(float* rays, int i, float ox, float oy, float oz, float dx, float dy, float dz, int parent, int key) {
    float4* p = reinterpret_cast<float4*>(rays + (size_t)i * RAY_WORDS);
    p[0] = make_float4(ox, oy, oz, dx);
    p[1] = make_float4(dy, dz, __int_as_float(parent), __int_as_float(key));
}
__device__ __forceinline__ void store_ray(double* rays, int i, double ox, double oy, double oz, double dx, double dy, double dz, int parent, int key) {
    double2* p = reinterpret_cast<double2*>(rays + (size_t)i * RAY_WORDS);
    p[0] = make_double2(ox, oy); p[1] = make_double2(oz, dx); p[2] = make_double2(dy, dz);
    p[3] = make_double2(__longlong_as_double((long long)(unsigned)parent), __longlong_as_double((long long)(unsigned)key));
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
// word 7: the class key of a spawned ray (its sort key: region of the origin x cube-map cell of the direction); it travels
// with the record through the key sort, so the cull kernel finds the ray's candidate list without a side array
template <typename R> __device__ __forceinline__ int ray_key(const R* rays, int i) {
    return reinterpret_cast<const int*>(rays + (size_t)i * RAY_WORDS + 7)[0];
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

__device__ __forceinline__ float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double fma_(double a, double b, double c) { return __fma_rn(a, b, c); }

// ------------------------------------------------------------------ first-hit rule per primitive
// Each returns true and (t, face) exactly as <Type>.Intersect would.

template <typename R> __device__ __forceinline__ bool hit_sphere(const Ray<R>& r, R& t, int& face) {   // :78-103
    R a = r.dx * r.dx + r.dy * r.dy + r.dz * r.dz;
    R hb = r.ox * r.dx + r.oy * r.dy + r.oz * r.dz;
    if (!GMLRT_PARITY_TU) {
        // Throughput mode: the same roots through the point of closest approach l = o + tc d:  disc = a (1 - |l|^2),
        // t0 = tc - sqrt((1 - |l|^2) / a).  The textbook hb^2 - a c cancels two numbers of the size |o|^2 (an origin tens of
        // radii away in object space), which in float32 moves silhouettes - and, worse, the hit point and with it the normal
        // of a mirror sphere - by 1e-4 of the radius; here every quantity after l is of order one.
        const R tc = -hb / a;
        const R lx = fma_(tc, r.dx, r.ox), ly = fma_(tc, r.dy, r.oy), lz = fma_(tc, r.dz, r.oz);
        const R pp = lx * lx + ly * ly + lz * lz;
        if (!(pp <= R(1))) return false;
        const R t0 = tc - gm<R>::sqrt((R(1) - pp) / a);
        if (t0 > R(0)) { t = t0; face = 0; return true; }
        return false;
    }
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
            Spans<R>& t1 = buf[ids[sp]];
            Spans<R>& t2 = buf[ids[sp + 1]];
            if ((-1 - op) & GMLRT_CSG_SWAPPED) swap_ids(sp - 2, sp - 1);      // the operands were evaluated b first (host_scene.h)
            const int kind = (-1 - op) & ~GMLRT_CSG_SWAPPED;
            Spans<R>& a = buf[ids[sp - 2]];
            Spans<R>& b = buf[ids[sp - 1]];
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

// a line on its way to L2 two iterations before it is loaded: costs no register, so the compiler cannot sink it to the use
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

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
// closestHit / inShadow visit every object for every ray (raytracer.go:469-483, 411-429).  The device keeps that
// all-pairs semantics - no object is ever dropped that a ray could hit - but splits the work in three, cheapest first:
//
//   classes  (once per upload) rays are classified by where they start and where they point ("bins": a camera ray by
//          its screen tile; a reflection / refraction ray by the sort region of its origin and the cube-map cell of its
//          direction; a shadow ray by the sort region of the shaded point and its light).  Every class has a bound that
//          holds ALL rays it can ever contain, so the objects that some ray of the class could hit are a property of the
//          scene: one list per class, built by k_list_count / k_list_fill (every object is offered to every class, in
//          double).  This replaces the per-warp / per-CTA packet bounds of round 1, which recomputed the same
//          kind of bound from the rays of every warp in every launch (ballots, shared-memory lists, two CTA barriers).
//   cull   (per ray x object of its class list) a few FFMA against the object's bounding sphere (centre C, radius R)
//          from a 16-byte (f32) record in shared memory; the radius carries a rounding margin so that a reject is always
//          a true miss.  Three record / ray "families", by what the rays of a launch share:
//            GENERAL  any ray (o, d^):  reject iff |C-o|^2 - ((C-o).d^)^2 > R^2, expanded so that
//                     everything per-ray is hoisted:  A = C.(-2o) + K,  K = |C|^2 - R^2 - margin,
//                     B = C.d^ - o.d^,  reject <=> A - B^2 > -|o|^2 + margin          7 FFMA + 1 FADD
//            PIVOT    rays whose LINE passes through one point V (camera rays: the eye; shadow rays of a
//                     point / spot light: the light):  record p = C - V, K = |p|^2 - R^2 - margin,
//                     B = p.d^,  reject <=> K - B^2 > 0                                 FMUL + 3 FFMA
//            AXIS     rays with one common direction (shadow rays of a directional light): project on
//                     the plane orthogonal to it,  du = Cu - ou, dv = Cv - ov,
//                     reject <=> du^2 + dv^2 - R^2 > 0                                  3 FADD + 2 FFMA
//          One thread per ray; the survivors of a window of GROUP = 64 list entries are one 64-bit mask.
//   exact  the survivors go to a candidate queue in HBM as (ray, object) pairs and a second kernel runs the
//          reference's own test (<Type>.Intersect, or the CSG span lists) on them, one thread per candidate.
constexpr int CULL_THREADS = 256;
constexpr int CULL_STAGE = 256;               // candidates a warp stages in shared memory between two reservations on the queue
constexpr int GROUP = 64;                   // list entries per window: one 64-bit survivor mask per ray, kept in two registers
enum CullFamily { CULL_GENERAL = 0, CULL_PIVOT = 1, CULL_AXIS = 2 };
// class of a shadow ray: the sort region of the shaded point, merged by `shift`; the last bin collects invalid keys
__device__ __forceinline__ int sort_bin(int key, int shift, int nb) { return key < 0 ? nb - 1 : min(key >> shift, nb - 2); }


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
// n / d for n < 2^31 and a divisor known on the host: mul = ceil(2^(31 + s) / d), s = ceil(log2 d) (Granlund - Montgomery)
struct FastDiv { unsigned mul; int shift; };
inline FastDiv fast_div_of(int d) {
    int s = 0;
    while ((1ll << s) < (long long)d) s++;
    FastDiv f;
    f.shift = 31 + s;
    f.mul = (unsigned)((((unsigned long long)1 << f.shift) + (unsigned long long)d - 1) / (unsigned long long)d);
    return f;
}
__host__ __device__ __forceinline__ int fast_div(int n, unsigned mul, int shift) { return (int)(((unsigned long long)(unsigned)n * mul) >> shift); }

__constant__ gomath::PcgJump c_pcg_jump[160];   // affine maps advancing the PCG stream by 0..159 draws
struct CamSrc { int y0; RowMap map; };          // which rows the level-0 rays of the current tile are
template <typename R>
__device__ __forceinline__ void camera_ray64(const DevScene<R>& sc, const CamSrc& cam, int i, double (&o)[3], double (&d)[3]) {
    const int s = i & 3, pix = i >> 2;
    const int row = fast_div(pix, sc.w_mul, sc.w_shift);
    const int x = pix - row * sc.W, y = cam.map.actual(cam.y0 + row);
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

// ------------------------------------------------------------------ ray classes: one candidate list per bin
// A bin's list holds, as byte offsets into the family's cull table, every object whose (margin-inflated) bounding
// sphere some ray of the bin could touch, padded to whole groups of four with the table's padding record (always
// rejected).  Built in double from the objects' bounds, whatever the arithmetic type of the renderer: the per-ray
// test that follows is the conservative one, the list only has to be a superset of what it would keep.
struct BinLists {
    const int2* bin;            // [n_bins] (first entry, entries); both multiples of 4
    const uint32_t* entry;      // byte offsets of cull records
    int n_bins;
};

// what the rays of a bin have in common (all in double)
struct BinBound {
    double ax, ay, az;          // GENERAL / PIVOT: unit axis of the cone of directions; AXIS: (u, v) centre of the disc in ax, ay
    double cos_t, sin_t;        // the cone's half angle
    double ox, oy, oz, rho;     // GENERAL: sphere holding every origin.  PIVOT: rho = how far from the pivot an occluder can be.  AXIS: rho = disc radius
    double rear;                // AXIS: rear-most origin along the rays' direction
    bool usable;                // false: no bound (an unbounded region, the pivot inside it, ...) - every object is listed
};

enum BinKind { BIN_CAMERA = 0, BIN_SPAWN = 1, BIN_SHADOW_PIVOT = 2, BIN_SHADOW_AXIS = 3 };

struct BinParams {
    int kind, n_bins;
    // the objects as this family's per-ray test sees them, [n][4]: centre relative to the pivot (PIVOT), absolute (GENERAL),
    // or (u, v, front end along the rays) (AXIS); then the radius with every margin of the per-ray test folded in (inf = unbounded)
    const double* obj; int n_objects;
    int no_bounds;                      // GMLRT_NO_CLASS_BOUNDS (debug): every class lists every object
    const float4* objf;                 // the same records in float32, radius grown by another 0.1 % + 1e-4: the screening pass (null: none)
    double pivot[3];                    // PIVOT families: the eye / the light
    double e1[3], e2[3], axis[3];       // AXIS: the frame the device projects with
    // BIN_SPAWN: bin = region * n_cells + cell;  BIN_SHADOW_*: bin = region (the last bin collects invalid keys: no bound)
    const double* region;               // [n_regions][4] sphere holding every ray origin of the region (radius inf = unbounded)
    const double* cell;                 // [n_cells][5] unit axis, cos, sin of the cone holding every direction of the cell
    int n_regions, n_cells;
    int limit_reach;                    // BIN_SHADOW_PIVOT: |incoming direction| is 1 for every ray, so occluders lie between the point and the light
    // BIN_CAMERA: bin = tile; tiles of tile_w x tile_h pixels
    int W, H, tile_w, tile_h, tiles_x;
    double cam_su, cam_sv, vw, vh;
};

__device__ __forceinline__ BinBound bin_bound(const BinParams& p, int b) {
    BinBound B = {};
    B.usable = false;
    B.rho = INFINITY; B.rear = -INFINITY;
    if (p.no_bounds) return B;
    if (p.kind == BIN_CAMERA) {
        // Render (raytracer.go:639-649): sample (x + dx, y + dy), |dx|, |dy| <= 0.5, is the ray from the eye (0, 0, -1) through
        // (u, -v, 0): direction ~ (u, -v, 1).  The directions of a tile are a rectangle on the plane z = 1; the farthest from
        // the centre direction is a corner.
        const int tx = b % p.tiles_x, ty = b / p.tiles_x;
        const double x0 = tx * p.tile_w - 0.5 - 1e-6, x1 = fmin((double)(tx + 1) * p.tile_w, (double)p.W) - 0.5 + 1e-6;
        const double y0 = ty * p.tile_h - 0.5 - 1e-6, y1 = fmin((double)(ty + 1) * p.tile_h, (double)p.H) - 0.5 + 1e-6;
        const double u0 = x0 * p.cam_su - 0.5 * p.vw, u1 = x1 * p.cam_su - 0.5 * p.vw;
        const double v0 = -(y0 * p.cam_sv - 0.5 * p.vh), v1 = -(y1 * p.cam_sv - 0.5 * p.vh);
        const double cu = 0.5 * (u0 + u1), cv = 0.5 * (v0 + v1);
        const double il = 1.0 / sqrt(cu * cu + cv * cv + 1.0);
        B.ax = cu * il; B.ay = cv * il; B.az = il;
        double c = 1.0;
        const double us[2] = {u0, u1}, vs[2] = {v0, v1};
        for (int k = 0; k < 4; k++) {
            const double u = us[k & 1], v = vs[k >> 1];
            const double d = (u * B.ax + v * B.ay + B.az) / sqrt(u * u + v * v + 1.0);
            c = fmin(c, d);
        }
        c -= 1e-9;                                       // (the float32 camera ray is the double one rounded: 1e-7 relative)
        const double th = acos(fmin(c, 1.0)) + 4e-7;
        B.cos_t = cos(th); B.sin_t = sin(th);
        B.usable = true;
        return B;
    }
    const int nreg = p.n_regions;
    const int region = p.kind == BIN_SPAWN ? b / p.n_cells : b;
    if (region >= nreg) return B;                        // shadow rays with an invalid key
    const double* rg = p.region + (size_t)region * 4;
    if (!(rg[3] < INFINITY)) return B;                    // unbounded region (or NaN)
    // origins: P +- N * 1e-4 (raytracer.go:412, 525, 551) with P inside the region's sphere, as rounded by the device
    const double rho = rg[3] * (1.0 + 1e-5) + 3e-4 + 1e-5 * (fabs(rg[0]) + fabs(rg[1]) + fabs(rg[2]));
    if (p.kind == BIN_SPAWN) {
        const double* cl = p.cell + (size_t)(b % p.n_cells) * 5;
        B.ax = cl[0]; B.ay = cl[1]; B.az = cl[2]; B.cos_t = cl[3]; B.sin_t = cl[4];
        B.ox = rg[0]; B.oy = rg[1]; B.oz = rg[2]; B.rho = rho;
        B.usable = true;
    } else if (p.kind == BIN_SHADOW_PIVOT) {
        // the rays point from their origins at the light: their lines pass the light within 1e-4 (folded into the table),
        // their directions lie in the cone from the light that holds the region
        const double vx = rg[0] - p.pivot[0], vy = rg[1] - p.pivot[1], vz = rg[2] - p.pivot[2];
        const double dist = sqrt(vx * vx + vy * vy + vz * vz);
        if (!(dist > rho * 1.0001)) return B;            // the light is inside the region
        B.ax = vx / dist; B.ay = vy / dist; B.az = vz / dist;
        const double th = asin(fmin(rho / dist, 1.0)) + 4e-7;
        B.cos_t = cos(th); B.sin_t = sin(th);
        B.rho = p.limit_reach ? (dist + rho) * 1.001 + 1e-3 : INFINITY;
        B.usable = true;
    } else {                                             // BIN_SHADOW_AXIS
        B.ax = rg[0] * p.e1[0] + rg[1] * p.e1[1] + rg[2] * p.e1[2];
        B.ay = rg[0] * p.e2[0] + rg[1] * p.e2[1] + rg[2] * p.e2[2];
        B.rho = rho * 1.0001 + 1e-6 * (fabs(B.ax) + fabs(B.ay) + 1.0);
        const double q = rg[0] * p.axis[0] + rg[1] * p.axis[1] + rg[2] * p.axis[2] - rho;
        B.rear = q - 1e-3 * (fabs(q) + 1.0);
        B.usable = true;
    }
    return B;
}

// may object o be hit by ANY ray inside the bin's bound?  (conservative; mirrors the per-ray families; no square root,
// no division: the angle test  angle(C - origin, axis) <= theta + alpha,  sin alpha = r / |C - origin|,  is
// dot + sin(theta) r >= cos(theta) sqrt(|C - origin|^2 - r^2), squared)
__device__ __forceinline__ bool bin_keeps(const BinParams& p, const BinBound& B, int o) {
    const double* ob = p.obj + (size_t)o * 4;
    const double r0 = ob[3];
    if (!(r0 < INFINITY)) return true;                    // unbounded (plane) or NaN: never culled
    if (!B.usable) return true;
    if (p.kind == BIN_SHADOW_AXIS) {
        if (ob[2] < B.rear) return false;                // wholly behind every origin along the rays' direction
        const double du = ob[0] - B.ax, dv = ob[1] - B.ay, r = r0 + B.rho;
        return !(du * du + dv * dv > r * r);
    }
    const bool general = p.kind == BIN_SPAWN;
    const double px = general ? ob[0] - B.ox : ob[0], py = general ? ob[1] - B.oy : ob[1], pz = general ? ob[2] - B.oz : ob[2];
    const double r = general ? r0 + B.rho : r0;
    const double pp = px * px + py * py + pz * pz;
    const double q = pp - r * r;
    if (!(q > 0.0)) return true;                         // an origin (the pivot) may be inside the sphere
    double dot = px * B.ax + py * B.ay + pz * B.az;
    if (!general) {
        const double far = B.rho + r;                    // shadow rays stop at the light: nearer to the pivot than this an occluder must be
        if (pp > far * far) return false;
        dot = fabs(dot);                                 // the LINE passes through the pivot: a double cone
    }
    const double A = dot + B.sin_t * r;                  // GENERAL is forward only (T > 0, raytracer.go:98)
    if (A < 0.0) return false;
    return A * A * (1.0 + 1e-12) >= B.cos_t * B.cos_t * q;
}

// The same question in float32, answered "no" only when it is no by a wide margin (cone widened by 1e-3 rad, radii by
// 0.1 %, rounding allowances on every comparison): almost every class x object pair is far from the bound, and
// only what passes here is looked at in double.
struct BinBoundF { float ox, oy, oz, rho, ax, ay, az, cos_t, sin_t, reach; bool general, usable, use; };
__device__ __forceinline__ BinBoundF bin_bound_f(const BinParams& p, const BinBound& B) {
    BinBoundF F = {};
    F.use = p.kind != BIN_SHADOW_AXIS && p.objf != nullptr;
    F.general = p.kind == BIN_SPAWN; F.usable = B.usable;
    if (!F.use || !B.usable) return F;
    F.ox = F.general ? (float)B.ox : 0.f; F.oy = F.general ? (float)B.oy : 0.f; F.oz = F.general ? (float)B.oz : 0.f;
    F.rho = F.general ? (float)(B.rho * 1.001 + 1e-4 + 1e-6 * (fabs(B.ox) + fabs(B.oy) + fabs(B.oz))) : 0.f;
    F.ax = (float)B.ax; F.ay = (float)B.ay; F.az = (float)B.az;
    F.cos_t = (float)(B.cos_t * 0.9999995 - B.sin_t * 1e-3 - 1e-6);          // cos / sin of (theta + 1e-3)
    F.sin_t = (float)(B.sin_t * 0.9999995 + B.cos_t * 1e-3 + 1e-6);
    F.reach = F.general ? INFINITY : (float)(B.rho * 1.001 + 1e-3);
    if (!(F.cos_t > 0.f)) F.usable = false;                                  // a cone that wide: everything goes to the double test
    return F;
}
__device__ __forceinline__ bool bin_maybe(const BinParams& p, const BinBoundF& F, int o) {
    if (!F.use || !F.usable) return true;
    const float4 c = p.objf[o];
    if (!(c.w < INFINITY)) return true;
    const float px = c.x - F.ox, py = c.y - F.oy, pz = c.z - F.oz;
    const float rr = c.w + F.rho;
    const float pp = px * px + py * py + pz * pz;
    const float q = pp - rr * rr;
    if (!(q > 1e-3f * pp)) return true;                  // close to (inside) the sphere: decide in double
    float dot = px * F.ax + py * F.ay + pz * F.az;
    if (!F.general) {
        const float far = F.reach + rr;
        if (pp > far * far * 1.001f) return false;
        dot = fabsf(dot);
    }
    const float A = dot + F.sin_t * rr + 4e-6f * (fabsf(px) + fabsf(py) + fabsf(pz));
    if (A < 0.f) return false;
    return !(A * A * 1.001f < F.cos_t * F.cos_t * q);
}

// one warp per bin: count (FILL = false) or write (FILL = true) the bin's list.  Objects are screened 32 at a time in
// float32; the few that pass wait in a per-warp queue until 32 of them can take the double-precision test side by side
// (tested where they pass, one or two lanes at a time, the double path would run for nearly every round of 32).
template <bool FILL>
__global__ void __launch_bounds__(256) k_list_build(const __grid_constant__ BinParams p, int* count, const int* first, int2* bins, uint32_t* entry,
                                                    uint32_t rec_bytes, unsigned long long* tests) {
    __shared__ int s_queue[8][64];
    const int lane = threadIdx.x & 31;
    int* const qu = s_queue[threadIdx.x >> 5];
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const unsigned lt = (1u << lane) - 1u;
    unsigned long long mine = 0;
    for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < p.n_bins; b += warps) {
        const BinBound B = bin_bound(p, b);
        const BinBoundF F = bin_bound_f(p, B);
        int n = 0, qn = 0;
        const int base = FILL ? first[b] : 0;
        auto drain = [&](int cnt) {                       // the double test for queue entries [0, cnt), cnt <= 32
            const int o = lane < cnt ? qu[lane] : -1;
            const bool keep = o >= 0 && bin_keeps(p, B, o);
            const unsigned bal = __ballot_sync(0xffffffffu, keep);
            if (FILL && keep) entry[base + n + __popc(bal & lt)] = (uint32_t)o * rec_bytes;
            n += __popc(bal);
        };
        for (int o0 = 0; o0 < p.n_objects; o0 += 32) {
            const int o = o0 + lane;
            const bool maybe = o < p.n_objects && bin_maybe(p, F, o);
            const unsigned bal = __ballot_sync(0xffffffffu, maybe);
            if (maybe) qu[qn + __popc(bal & lt)] = o;
            qn += __popc(bal);
            __syncwarp();
            if (qn >= 32) {
                drain(32);
                __syncwarp();
                const int rest = lane < qn - 32 ? qu[32 + lane] : 0;
                __syncwarp();
                if (lane < qn - 32) qu[lane] = rest;
                qn -= 32;
                __syncwarp();
            }
        }
        if (qn) drain(qn);
        __syncwarp();
        mine += (unsigned long long)((p.n_objects + 31) / 32);
        const int padded = (n + 3) & ~3;
        if (FILL) {
            if (lane < padded - n) entry[base + n + lane] = (uint32_t)p.n_objects * rec_bytes;     // the table's first padding record: always rejected
            if (lane == 0) bins[b] = make_int2(base, padded);
        } else if (lane == 0) count[b] = padded;
    }
    if (!FILL && tests && lane == 0 && mine) atomicAdd(tests, mine * 32ull);      // (every lane counted the warp's rounds: one atomic per warp)
}

// A cull record by its byte offset in the family's table.  SMEM: `tab` is the table's 32-bit shared-memory address, so the
// load is ld.shared with the base folded in as an immediate (through a generic pointer every load paid an extra
// add of the shared window's base); else `gtab` is the table in global memory, read through L1.
template <typename R, bool SMEM> __device__ __forceinline__ Cull4<R> load_cull_at(uint32_t tab, const unsigned char* gtab, uint32_t off);
template <> __device__ __forceinline__ Cull4<float> load_cull_at<float, true>(uint32_t tab, const unsigned char*, uint32_t off) {
    Cull4<float> c;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(c.x), "=f"(c.y), "=f"(c.z), "=f"(c.k) : "r"(tab + off));
    return c;
}
template <> __device__ __forceinline__ Cull4<double> load_cull_at<double, true>(uint32_t tab, const unsigned char*, uint32_t off) {
    Cull4<double> c;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(c.x), "=d"(c.y) : "r"(tab + off));
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(c.z), "=d"(c.k) : "r"(tab + off + 16u));
    return c;
}
template <> __device__ __forceinline__ Cull4<float> load_cull_at<float, false>(uint32_t, const unsigned char* gtab, uint32_t off) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(gtab + off));
    Cull4<float> c = {v.x, v.y, v.z, v.w};
    return c;
}
template <> __device__ __forceinline__ Cull4<double> load_cull_at<double, false>(uint32_t, const unsigned char* gtab, uint32_t off) {
    const double2 a = __ldg(reinterpret_cast<const double2*>(gtab + off)), b = __ldg(reinterpret_cast<const double2*>(gtab + off + 16));
    Cull4<double> c = {a.x, a.y, b.x, b.y};
    return c;
}
// one window of a ray's list: entries ent[0 .. cnt) (cnt a multiple of 4, at most GROUP), 4 entries per 16-byte load, 8 tests
// per loop body; test j lands in bit 31 - j of lo (j < 32) / hi (set = object survives)
template <typename R, int FAMILY, bool SMEM>
__device__ __forceinline__ void cull_list_window(uint32_t tab, const unsigned char* gtab, const uint32_t* ent, int cnt, const CullRay<R, FAMILY>& cr, uint32_t& lo, uint32_t& hi) {
    lo = 0; hi = 0;
    if (cnt <= 0) return;
    auto four = [&](uint32_t m, const uint4& c) {
        m = push_sign(m, cr.slack(load_cull_at<R, SMEM>(tab, gtab, c.x)));
        m = push_sign(m, cr.slack(load_cull_at<R, SMEM>(tab, gtab, c.y)));
        m = push_sign(m, cr.slack(load_cull_at<R, SMEM>(tab, gtab, c.z)));
        return push_sign(m, cr.slack(load_cull_at<R, SMEM>(tab, gtab, c.w)));
    };
    const uint4* e4 = reinterpret_cast<const uint4*>(ent);
    const int n4 = cnt >> 2;                                   // groups of four
    // half a window (8 groups) at a time; the offsets of the next two groups are in flight while two are tested
    auto half = [&](int g0, int g1) -> uint32_t {
        uint32_t m = 0;
        int g = g0;
        uint4 a = __ldg(e4 + g), b = (g + 1 < g1) ? __ldg(e4 + g + 1) : a;
#pragma unroll 1
        for (; g + 1 < g1; g += 2) {
            const uint4 ca = a, cb = b;
            if (g + 2 < g1) a = __ldg(e4 + g + 2);
            if (g + 3 < g1) b = __ldg(e4 + g + 3);
            m = four(m, ca);
            m = four(m, cb);
        }
        if (g < g1) m = four(m, a);
        const int tested = 4 * (g1 - g0);
        return tested < 32 ? m << (32 - tested) : m;
    };
    lo = half(0, min(n4, 8));
    if (n4 > 8) hi = half(8, n4);
}

__device__ __forceinline__ int warp_max(int v) { return __reduce_max_sync(0xffffffffu, v); }      // one REDUX instead of five shuffles
template <typename R>
__device__ __forceinline__ R warp_max(R v) { for (int o = 16; o > 0; o >>= 1) { const R u = __shfl_xor_sync(0xffffffffu, v, o); v = u > v ? u : v; } return v; }
__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }

// ------------------------------------------------------------------ scan = cull kernel + exact kernels
// Phase 1, k_cull: one thread per ray.  The thread looks up the list of its ray's class and tests the ray against the
// listed bounding spheres, GROUP entries per window; the warp appends every survivor to the launch's candidate queue
// in HBM as a (ray, object) pair - one atomic per warp and window.  The lanes of a warp usually share a class (camera
// rays: a tile; sorted spawned rays: a bin; shadow rays of neighbouring pixels: one region), so their list and record
// loads are broadcasts; where they differ the hardware serialises only the distinct addresses.  The family's cull
// table is staged once per CTA into shared memory by one TMA bulk copy when it fits (TABLE_SMEM), else read through L1.
// Phase 2, k_exact_*: one thread per candidate runs the reference's own test and folds the result into the ray's
// record with atomics.
//
// The queue is sized from the scene (Renderer::plan_tiles); if a launch overflows it, `Counters::overflow`
// is raised, the frame is discarded and traced again with smaller tiles (gmlrt_trace).
struct CandQueue {
    int* ray;        // [cap] ray index within the launch's ray source
    int* obj;        // [cap] object index; k_exact_closest overwrites it with the hit code (or -1)
    void* t;         // [cap] R: T of the hit (closest only)
    int* count;      // [0] entries appended by the current launch
    int cap;
};

// entries of a finished cull launch.  After an overflow the queue has holes (a warp that did not fit reserved but never
// wrote) whose stale (ray, object) words may belong to a larger level: the frame is discarded anyway, so skip them all.
__device__ __forceinline__ int cand_count(const CandQueue& cq, const Counters* ctr) {
    if (ctr->overflow) return 0;
    return (int)min(*reinterpret_cast<const unsigned*>(cq.count), (unsigned)cq.cap);
}

template <typename R> __device__ __forceinline__ void level_init_ray(const Level<R>& lv, int i, bool pickless = false) {       // no hit yet (see fold_hit)
    if (!pickless) lv.hit[i] = -1;                           // (pickless: nobody reads the hit code word, the folded key is the hit)
    if constexpr (sizeof(R) == 4) reinterpret_cast<unsigned long long*>(lv.order)[i] = ~0ull;
    else {
        reinterpret_cast<unsigned long long*>(lv.t)[i] = ~0ull;
        lv.order[i] = 0x7fffffff;
    }
}

// A ray source hands a thread everything it needs of ray i in ONE call, fetch(): the loads of the next iteration's ray are
// issued before the current ray's list is walked (k_cull), so their DRAM latency hides behind it (ncu, r02q: a third of the
// camera-ray launch's stall samples sat on the first use of the ray record).
template <typename R> struct CameraSource {                  // level 0: camera rays in (pixel, sample) order over the virtual rows of a tile
    Level<R> lv; int W, y0; RowMap map; int tile_w_log2, tile_h_log2, tiles_x; unsigned w_mul; int w_shift; bool pickless;
    __device__ __forceinline__ int slot(int i) const { return i; }
    __device__ __forceinline__ void fetch(int i, Ray<R>& out, int& b, int& skip) const {
        skip = -1;
        load_ray(lv.ray, i, out);
        const int pix = i >> 2;
        const int row = fast_div(pix, w_mul, w_shift);
        b = (map.actual(y0 + row) >> tile_h_log2) * tiles_x + ((pix - row * W) >> tile_w_log2);
    }
    __device__ __forceinline__ void ahead(int i) const { prefetch_l2(lv.ray + (size_t)i * RAY_WORDS); }
    __device__ __forceinline__ int cls(int b) const { return b; }
    __device__ __forceinline__ void init(int rid) const { level_init_ray(lv, rid, pickless); }
    static constexpr bool is_shadow = false;
};

// Camera rays MADE by the cull kernel (GMLRT_FUSE_GEN, default): fetch() draws the sample's jitter, forms the ray (camera_ray64,
// exactly what k_gen does), stores its record for the kernels that follow and hands it to the cull in registers - the 32-byte
// record is written once and not read back by the next launch, and the two issue-bound kernels share one set of resident warps.
template <typename R> struct GenCameraSource {
    DevScene<R> sc; Level<R> lv; int W; CamSrc cam; int tile_w_log2, tile_h_log2, tiles_x; bool pickless;
    __device__ __forceinline__ int slot(int i) const { return i; }
    __device__ __forceinline__ void fetch(int i, Ray<R>& out, int& b, int& skip) const {
        skip = -1;
        double o[3], d[3];
        camera_ray64(sc, cam, i, o, d);
        out.ox = (R)o[0]; out.oy = (R)o[1]; out.oz = (R)o[2]; out.dx = (R)d[0]; out.dy = (R)d[1]; out.dz = (R)d[2];
        store_ray(lv.ray, i, out.ox, out.oy, out.oz, out.dx, out.dy, out.dz, 0, 0);
        const int pix = i >> 2;
        const int row = fast_div(pix, sc.w_mul, sc.w_shift);
        b = (cam.map.actual(cam.y0 + row) >> tile_h_log2) * tiles_x + ((pix - row * W) >> tile_w_log2);
    }
    __device__ __forceinline__ void init(int rid) const { level_init_ray(lv, rid, pickless); }
    __device__ __forceinline__ void ahead(int) const {}
    __device__ __forceinline__ int cls(int b) const { return b; }
    static constexpr bool is_shadow = false;
};

template <typename R> struct SpawnSource {                   // levels >= 1: reflection / refraction rays; the class key travels in the record
    Level<R> lv; bool pickless; const int* perm;             // perm: the rays in class order (the key sort moved indices, not records); null: the level is stored sorted
    __device__ __forceinline__ int slot(int i) const { return perm ? perm[i] : i; }
    __device__ __forceinline__ void fetch(int rid, Ray<R>& out, int& b, int& skip) const {
        skip = -1;
        load_ray(lv.ray, rid, out);
        b = ray_key(lv.ray, rid);
    }
    __device__ __forceinline__ void ahead(int i) const { if (!perm) prefetch_l2(lv.ray + (size_t)i * RAY_WORDS); }
    __device__ __forceinline__ int cls(int b) const { return b; }
    __device__ __forceinline__ void init(int rid) const { level_init_ray(lv, rid, pickless); }
    static constexpr bool is_shadow = false;
};

template <typename R> struct ShadowSource {                  // shadow rays of one light: slots [base, base + n)
    ShadowQueue<R> sq; int base; const int* perm;            // perm: by-region order of the slots (null: queue order)
    int shift, nb;                                           // class = sort_bin(region of the shaded point)
    __device__ __forceinline__ int slot(int i) const { return perm ? perm[base + i] : i; }
    __device__ __forceinline__ void fetch(int rid, Ray<R>& out, int& b, int& skip) const {
        load_ray(sq.q + (size_t)(base + rid) * 8, 0, out);
        b = sq.key[base + rid];                              // raw: sort_bin() at the use (cls), or the fetch would wait for the load it was issued early for
        skip = sq.skip[base + rid];                          // the shaded object: inShadow passes over it (:418), and its bound holds the ray's origin - a sure survivor
    }
    __device__ __forceinline__ void init(int rid) const { sq.first[base + rid] = 0x7fffffff; }
    __device__ __forceinline__ void ahead(int i) const {
        if (perm) return;                                    // (sorted: the slot is a load away)
        prefetch_l2(sq.q + (size_t)(base + i) * 8);
        if ((i & 7) == 0) { prefetch_l2(sq.key + base + i); prefetch_l2(sq.skip + base + i); }      // 8 four-byte entries per 32-byte sector
    }
    __device__ __forceinline__ int cls(int b) const { return sort_bin(b, shift, nb); }
    static constexpr bool is_shadow = true;
};

template <typename R, int FAMILY, typename Source, bool TABLE_SMEM>
__global__ void __launch_bounds__(CULL_THREADS)
k_cull(const __grid_constant__ Source src, const int* n_ptr, const R* table, const __grid_constant__ CullFrame<R> frame,
       int n_cull, const __grid_constant__ BinLists bl, const __grid_constant__ CandQueue cq, Counters* ctr) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, lane = tid & 31;
    const int n = *n_ptr;
    const int nblocks = (n + CULL_THREADS - 1) / CULL_THREADS;
    if ((int)blockIdx.x >= nblocks) return;
    if (Source::is_shadow && blockIdx.x == 0 && tid == 0) atomicAdd(&ctr->rays_shadow, (unsigned long long)n);
    if (TABLE_SMEM) {
        if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
        __syncthreads();
        if (tid == 0) stage_cull(table, 0, n_cull, reinterpret_cast<R*>(smem_raw), &bar);
        mbar_wait(&bar, 0);
    }
    const uint32_t tab_s = smem_u32(smem_raw);
    const unsigned char* const tab_g = reinterpret_cast<const unsigned char*>(table);
    unsigned long long fine = 0;
    __shared__ int2 s_stage[CULL_THREADS / 32][CULL_STAGE];
    int2* const stage = s_stage[tid >> 5];
    int staged = 0;                                            // entries in this warp's stage (uniform across the warp)
    // reserve `cnt` queue entries for the warp (every lane calls it): the first one, or all ones when the queue is full.
    // The counter is unsigned and stops growing once the queue is full: a lattice seen along an axis can leave dozens of
    // survivors per ray, and a signed counter that kept counting past 2^31 would turn the index negative and the stores
    // into out-of-bounds writes instead of the overflow-and-retrace path.
    // (A warp that has seen the queue full never adds again, so the counter overshoots the capacity by at most one
    // reservation per warp of the grid; a look at the counter before every atomic - a second L2 round trip - was 29 % of the
    // shadow launches' stall samples.)
    bool queue_full = false;                                    // uniform across the warp
    auto reserve = [&](int cnt) -> unsigned {
        unsigned at = 0xffffffffu;
        if (lane == 0 && !queue_full) {
            at = atomicAdd(reinterpret_cast<unsigned*>(cq.count), (unsigned)cnt);
            if (at >= (unsigned)cq.cap || (unsigned)cnt > (unsigned)cq.cap - at) { at = 0xffffffffu; ctr->overflow = 1; }   // queue full: the host retries with smaller tiles
        }
        at = __shfl_sync(0xffffffffu, at, 0);
        queue_full = queue_full || at == 0xffffffffu;
        return at;
    };
    auto flush = [&]() {
        if (staged == 0) return;
        const unsigned at = reserve(staged);
        if (at != 0xffffffffu)
            for (int k = lane; k < staged; k += 32) { const int2 e = stage[k]; cq.ray[at + k] = e.x; cq.obj[at + k] = e.y; }
        staged = 0;
        __syncwarp();
    };
    const Ray<R> idle = {R(0), R(0), R(0), R(0), R(0), R(1)};
    Ray<R> w_nx = idle;
    int rid_nx = 0, b_nx = 0, skip_nx = -1;
    const int step = (int)gridDim.x * CULL_THREADS;
    const int i_first = (int)blockIdx.x * CULL_THREADS + tid;
    if (i_first < n) { rid_nx = src.slot(i_first); src.fetch(rid_nx, w_nx, b_nx, skip_nx); }
    int rid_nx2 = (long long)i_first + step < n ? src.slot(i_first + step) : 0;      // where a sort left only a permutation, the slot of a ray is a load of its own: one more iteration ahead
    for (int blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const int i = blk * CULL_THREADS + tid;
        const bool active = i < n;
        const Ray<R> w = w_nx;
        const int rid = rid_nx;
        const uint32_t skip_off = skip_nx < 0 ? 0xffffffffu : (uint32_t)skip_nx * (uint32_t)(4 * sizeof(R));      // the skipped object as a list entry
        int first = 0, len = 0;
        if (active) {
            src.init(rid);
            const int2 rec = __ldg(bl.bin + min(max(src.cls(b_nx), 0), bl.n_bins - 1));
            first = rec.x; len = rec.y;
        }
        {   // the next iteration's ray: in flight while this one's list is walked
            const int i2 = i + step;
            w_nx = idle; rid_nx = 0; b_nx = 0; skip_nx = -1;
            if (blk + (int)gridDim.x < nblocks && i2 < n) { rid_nx = rid_nx2; src.fetch(rid_nx, w_nx, b_nx, skip_nx); }
            const long long i3 = (long long)i2 + step;
            rid_nx2 = 0;
            if (i3 < n) { rid_nx2 = src.slot((int)i3); src.ahead((int)i3); }      // ... and the one after it: its slot, and (stored in order) its record on the way to L2
        }
        CullRay<R, FAMILY> cr;
        cr.init(w, frame);
        const int len_max = warp_max(len);
        for (int k0 = 0; k0 < len_max; k0 += GROUP) {          // uniform across the warp
            const int cnt = min(max(len - k0, 0), GROUP);
            uint32_t lo, hi;
            cull_list_window<R, FAMILY, TABLE_SMEM>(tab_s, tab_g, bl.entry + first + k0, cnt, cr, lo, hi);
            fine += (unsigned long long)cnt;
            // Survivors are staged per warp in shared memory and flushed to the queue STAGE at a time: one atomic on the
            // queue's counter per flush.  (One atomic per warp and window - a million and more per launch, all on one
            // address - serialised in L2 and was ALL of this kernel's time when first measured.)
            if (Source::is_shadow && skip_off != 0xffffffffu) {    // a shadow ray never asks about the object it starts on: half of all shadow candidates, dropped here instead of in k_exact_shadow
                const uint32_t* e0 = bl.entry + first + k0;
                for (uint32_t m = lo; m; ) { const int j = __clz(m); m &= ~(0x80000000u >> j); if (e0[j] == skip_off) lo &= ~(0x80000000u >> j); }
                for (uint32_t m = hi; m; ) { const int j = __clz(m); m &= ~(0x80000000u >> j); if (e0[32 + j] == skip_off) hi &= ~(0x80000000u >> j); }
            }
            const int mine = __popc(lo) + __popc(hi);
            if (!__any_sync(0xffffffffu, mine != 0)) continue;
            int incl = mine;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += v; }
            const int total = __shfl_sync(0xffffffffu, incl, 31);
            if (staged + total > CULL_STAGE) { flush(); }
            const uint32_t* ent = bl.entry + first + k0;
            if (total <= CULL_STAGE) {
                int at = staged + incl - mine;
                uint32_t m = lo;
                while (m) {
                    const int j = __clz(m);
                    m &= ~(0x80000000u >> j);
                    stage[at++] = make_int2(rid, (int)(ent[j] / (uint32_t)(4 * sizeof(R))));
                }
                m = hi;
                while (m) {
                    const int j = __clz(m);
                    m &= ~(0x80000000u >> j);
                    stage[at++] = make_int2(rid, (int)(ent[32 + j] / (uint32_t)(4 * sizeof(R))));
                }
                staged += total;
                __syncwarp();
            } else {                                           // more survivors in one window than the stage holds: straight to the queue
                unsigned at = reserve(total);
                if (at == 0xffffffffu) continue;
                at += (unsigned)(incl - mine);
                uint32_t m = lo;
                while (m) {
                    const int j = __clz(m);
                    m &= ~(0x80000000u >> j);
                    cq.ray[at] = rid; cq.obj[at] = (int)(ent[j] / (uint32_t)(4 * sizeof(R))); at++;
                }
                m = hi;
                while (m) {
                    const int j = __clz(m);
                    m &= ~(0x80000000u >> j);
                    cq.ray[at] = rid; cq.obj[at] = (int)(ent[32 + j] / (uint32_t)(4 * sizeof(R))); at++;
                }
            }
        }
    }
    flush();
    fine = warp_sum_u64(fine);
    if (lane == 0 && fine) atomicAdd(&ctr->cull_tests[FAMILY], fine);
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
        // one candidate per thread: its ray index is fetched with the object word, not after the ownership test on it (the
        // kernel is a chain of dependent DRAM round trips - candidate, ray record, fold - and this takes one link out)
        const int ri_first = U == 1 ? cq.ray[k0] : 0;
#pragma unroll
        for (int u = 0; u < U; u++) {
        const int k = k0 + u;
        if (k >= count) break;
        int obj = objs[u];
        if (KINDS == EXACT_REST && obj < 0) continue;          // a sphere candidate the SPHERES pass has already resolved to a miss
        bool deferred = false;
        if (ROBUST && KINDS == EXACT_REST) { deferred = (obj & CAND_DEFER) != 0; obj &= ~CAND_DEFER; }
        if (KINDS != EXACT_ALL && !deferred) {
            // Without the pickless fold an earlier flavour has overwritten the object word of ITS candidates with the hit code
            // (or -1): a code with face / flip bits is >= 2^27 and must not be taken for a padding record and reset to a miss
            // (chess.gml: the board, a plain cube hit on faces 1-5, vanished in parity mode once a CSG flavour ran after SIMPLE).
            if (obj < 0 || obj > sc.n_leaves) continue;
            if (obj < sc.n_leaves && !exact_owns<KINDS>(obj, n_spheres, sc.kind_begin[K_CSG])) continue;
        }
        const int ri = U == 1 ? ri_first : cq.ray[k];
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


template <typename R>
__global__ void __launch_bounds__(SORT_THREADS) k_sort_hist(const __grid_constant__ ShadowQueue<R> sq, int light, int shift, int nb, int* hist) {
    __shared__ int h[SORT_BINS_MAX];
    const int n = sq.count[light], base = light * sq.per_light;
    const int nchunks = (n + SORT_CHUNK - 1) / SORT_CHUNK;
    for (int b = threadIdx.x; b < nb; b += SORT_THREADS) h[b] = 0;
    __syncthreads();
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x)
        for (int i = c * SORT_CHUNK + threadIdx.x; i < min(n, (c + 1) * SORT_CHUNK); i += SORT_THREADS)
            atomicAdd(&h[sort_bin(sq.key[base + i], shift, nb)], 1);
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += SORT_THREADS) if (h[b]) atomicAdd(&hist[b], h[b]);
}

// exclusive prefix of hist[0..nb) into cursor[0..nb); hist is cleared for the next use unless clear == 0.  One block: every
// thread owns a contiguous run of bins (a multiple of 4, read as int4), sums it, the 1024 sums are scanned across the
// block, then the thread walks its run again (now in L1) and writes the prefixes.  Two block barriers in all - the
// tile-by-tile version this replaces paid four per 4096 bins and took 61 us for the 98k bins of the key sort.
static __global__ void __launch_bounds__(1024) k_sort_scan(int nb, int* hist, int* cursor, int clear = 1) {
    __shared__ int warp_tot[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int seg = (((nb + 1023) >> 10) + 3) & ~3;
    const int b0 = min(tid * seg, nb), b1 = min(b0 + seg, nb);
    auto load4 = [&](int b) {
        int4 v = make_int4(0, 0, 0, 0);
        if (b + 3 < nb) v = *reinterpret_cast<const int4*>(hist + b);
        else { if (b < nb) v.x = hist[b]; if (b + 1 < nb) v.y = hist[b + 1]; if (b + 2 < nb) v.z = hist[b + 2]; }
        return v;
    };
    int mine = 0;
    for (int b = b0; b < b1; b += 4) { const int4 v = load4(b); mine += v.x + v.y + v.z + v.w; }
    int incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += u; }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int t = warp_tot[lane], ti = t;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const int u = __shfl_up_sync(0xffffffffu, ti, off); if (lane >= off) ti += u; }
        warp_tot[lane] = ti - t;                           // exclusive prefix of the warp totals
    }
    __syncthreads();
    int run = warp_tot[wid] + incl - mine;
    for (int b = b0; b < b1; b += 4) {
        const int4 v = load4(b);
        if (b + 3 < nb) {
            *reinterpret_cast<int4*>(cursor + b) = make_int4(run, run + v.x, run + v.x + v.y, run + v.x + v.y + v.z);
            if (clear) *reinterpret_cast<int4*>(hist + b) = make_int4(0, 0, 0, 0);
        } else {
            if (b < nb) { cursor[b] = run; if (clear) hist[b] = 0; }
            if (b + 1 < nb) { cursor[b + 1] = run + v.x; if (clear) hist[b + 1] = 0; }
            if (b + 2 < nb) { cursor[b + 2] = run + v.x + v.y; if (clear) hist[b + 2] = 0; }
        }
        run += v.x + v.y + v.z + v.w;
    }
}

// The same scan over many bins (the key sort has 10^5 - 2 10^6): tiles of SCAN_TILE bins dealt round-robin to co-resident blocks
// (the host launches at most one block per SM; a block takes its tiles in ascending order, so the tiles a tile waits for are
// finished or in progress on a running block).  A block sums its tile, publishes the sum with a flag that carries the launch's
// epoch (so nothing needs resetting), waits for the sums of the tiles before it and writes its prefixes.
constexpr int SCAN_TILE = 4096, SCAN_TILES_MAX = 1024;
static __global__ void __launch_bounds__(1024) k_sort_scan_tiles(int nb, int* hist, int* cursor, int clear, volatile int* tile_sum, volatile int* tile_flag, int epoch) {
    __shared__ int warp_tot[32];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int n_tiles = (nb + SCAN_TILE - 1) / SCAN_TILE;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int b = tile * SCAN_TILE + 4 * tid;
    int4 v = make_int4(0, 0, 0, 0);
    if (b + 3 < nb) v = *reinterpret_cast<const int4*>(hist + b);
    else { if (b < nb) v.x = hist[b]; if (b + 1 < nb) v.y = hist[b + 1]; if (b + 2 < nb) v.z = hist[b + 2]; }
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
        warp_tot[lane] = ti - t;
        if (lane == 31) {                                  // this tile's total: publish, then the flag
            tile_sum[tile] = ti;
            __threadfence();
            tile_flag[tile] = epoch;
        }
        // the tiles before this one: lane k looks after tiles k, k + 32, ...
        int before = 0;
        for (int t0 = lane; t0 < tile; t0 += 32) {
            while (tile_flag[t0] != epoch) { }
            __threadfence();
            before += tile_sum[t0];
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) before += __shfl_xor_sync(0xffffffffu, before, off);
        if (lane == 0) s_base = before;
    }
    __syncthreads();
    int run = s_base + warp_tot[wid] + incl - mine;
    if (b + 3 < nb) {
        *reinterpret_cast<int4*>(cursor + b) = make_int4(run, run + v.x, run + v.x + v.y, run + v.x + v.y + v.z);
        if (clear) *reinterpret_cast<int4*>(hist + b) = make_int4(0, 0, 0, 0);
    } else {
        if (b < nb) { cursor[b] = run; if (clear) hist[b] = 0; }
        if (b + 1 < nb) { cursor[b + 1] = run + v.x; if (clear) hist[b + 1] = 0; }
        if (b + 2 < nb) { cursor[b + 2] = run + v.x + v.y; if (clear) hist[b + 2] = 0; }
    }
    __syncthreads();                                       // warp_tot / s_base are rewritten by the block's next tile
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
constexpr int KEY_BINS_MAX = 1 << 21;        // histogram / cursor arrays of the key sort are sized for this; GMLRT_KEY_BINS (default 2^19) is the limit in use

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
__global__ void __launch_bounds__(256) k_keysort_scatter(const int* keys, const int* n_ptr, int nb, int* cursor, const R* stage, R* out, int* perm_out) {
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
        if (valid && perm_out) perm_out[at + rank] = i;    // permutation only: the records stay where k_shade wrote them, the cull kernel gathers them (one 32-byte sector each)
        else if (valid) {                                  // move the whole record: the level is STORED in sorted order
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


// the per-frame counters alone (when the camera rays are made by the cull kernel, GenCameraSource)
static __global__ void __launch_bounds__(256) k_frame_init(Counters* ctr, int S, int* slots, int n_slots) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    for (int k = i; k < n_slots; k += gridDim.x * blockDim.x) slots[k] = 0;
    if (i == 0) {
        for (int l = 1; l <= GMLRT_MAX_DEPTH; l++) ctr->count[l] = 0;
        for (int l = 0; l <= GMLRT_MAX_DEPTH; l++) ctr->hits[l] = 0;
        ctr->count[0] = S;
    }
}

template <typename R>
__global__ void __launch_bounds__(256) k_gen(DevScene<R> sc, Level<R> lv, Counters* ctr, int y0, int rows, RowMap map, int* slots, int n_slots) {
    const int S = rows * sc.W * 4;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    for (int k = i; k < n_slots; k += gridDim.x * blockDim.x) slots[k] = 0;     // the frame's queue counters (one per launch that appends: trace_rows)
    if (i == 0) {
        for (int l = 1; l <= GMLRT_MAX_DEPTH; l++) ctr->count[l] = 0;
        for (int l = 0; l <= GMLRT_MAX_DEPTH; l++) ctr->hits[l] = 0;
        ctr->count[0] = S;
    }
    const CamSrc cam{y0, map};
    for (int k = i; k < S; k += gridDim.x * blockDim.x) {                        // a few thousand resident blocks walk the tile (one block per 256 samples spent a fifth of its stall samples on its own start-up)
        double o[3], d[3];
        camera_ray64(sc, cam, k, o, d);
        store_ray(lv.ray, k, (R)o[0], (R)o[1], (R)o[2], (R)d[0], (R)d[1], (R)d[2], 0, 0);
    }
}

// Background of a ray that hit nothing: Lerp(BgStart, BgEnd, 0.5 (dir.Y + 1)) (raytracer.go:493-497)
template <typename R> __device__ __forceinline__ V3<R> background(const DevScene<R>& sc, R dy) {
    const R t = R(0.5) * (dy + R(1));
    const V3<R> a = ld3(sc.bg0), b = ld3(sc.bg1);
    return mk(a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t, a.z + (b.z - a.z) * t);
}

// ------------------------------------------------------------------ hit list (stream compaction between scan and shading)
// After the scan of a level a third to two thirds of its rays have hit nothing (4K spheres1024: 64 % of the camera rays), and
// in ray order hits and misses share warps: k_shade ran with 22 of 32 lanes busy, k_resolve with 13-16 (ncu, round 1).
// This pass splits them: the indices of the rays that hit are appended to Level::hits (__ballot_sync / __popc within the
// warp, one shared-memory scan per block, ONE atomic on the level's counter per block of HITLIST_RAYS rays), and a miss is
// finished here and now - on levels >= 1 its background colour goes straight into the parent's child slot (on level 0 the
// pixel pass looks at the fold key again).  Everything after this runs on dense warps over per-hit arrays.
constexpr int HITLIST_THREADS = 256, HITLIST_PER_THREAD = 8, HITLIST_RAYS = HITLIST_THREADS * HITLIST_PER_THREAD;
template <typename R, int PICKLESS>
__global__ void __launch_bounds__(HITLIST_THREADS) k_hitlist(const __grid_constant__ DevScene<R> sc, const __grid_constant__ Level<R> lv, const __grid_constant__ Level<R> prev,
                                                            Counters* ctr, int level, int32_t* ids, int y0, const __grid_constant__ RowMap map) {
    const int n = ctr->count[level];
    __shared__ int s_warp[HITLIST_THREADS / 32];
    __shared__ int s_base;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    unsigned long long n_hit = 0;
    for (int blk0 = blockIdx.x * HITLIST_RAYS; blk0 < n; blk0 += gridDim.x * HITLIST_RAYS) {       // uniform per block
        unsigned bal[HITLIST_PER_THREAD];
        int hs[HITLIST_PER_THREAD];
        int mine = 0;
        // the fold keys of the thread's rays first, all loads in flight together (behind the stores of the miss handling below
        // the compiler has to keep them in program order: the pointers may alias)
#pragma unroll
        for (int k = 0; k < HITLIST_PER_THREAD; k++) {
            const int i = blk0 + (wid * HITLIST_PER_THREAD + k) * 32 + lane;      // a warp owns 32 * HITLIST_PER_THREAD consecutive rays: its hits stay in order
            hs[k] = -1;
            if (i < n) { R t; level_hit<R, PICKLESS>(lv, i, hs[k], t); }
        }
#pragma unroll
        for (int k = 0; k < HITLIST_PER_THREAD; k++) {
            const int i = blk0 + (wid * HITLIST_PER_THREAD + k) * 32 + lane;
            bool hit = false;
            if (i < n) {
                const int h = hs[k];
                hit = h >= 0;
                if (ids) {                                     // level 0, counting mode: primary hit ids in (y, x, sample) order
                    const int pix = i >> 2;
                    ids[((size_t)map.actual(y0 + pix / sc.W) * sc.W + pix % sc.W) * 4 + (i & 3)] = h < 0 ? -1 : sc.scan_of[h & HIT_LEAF_MASK];
                }
                if (!hit && level > 0) {                       // a miss is finished here: background into the parent's child slot
                    Ray<R> w;
                    load_ray(lv.ray, i, w);
                    const V3<R> c = background(sc, w.dy);
                    const int p = ray_parent(lv.ray, i);
                    const size_t pi = (size_t)(p & 0x7fffffff), pc = (size_t)prev.cap;
                    const int base = (p < 0) ? 3 : 0;
                    prev.child[(base + 0) * pc + pi] = c.x; prev.child[(base + 1) * pc + pi] = c.y; prev.child[(base + 2) * pc + pi] = c.z;
                }
            }
            bal[k] = __ballot_sync(0xffffffffu, hit);
            mine += __popc(bal[k]);
        }
        if (lane == 0) s_warp[wid] = mine;                    // (every lane of a warp holds the same ballots)
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
#pragma unroll
            for (int w = 0; w < HITLIST_THREADS / 32; w++) { const int c = s_warp[w]; s_warp[w] = tot; tot += c; }
            s_base = tot ? atomicAdd(&ctr->hits[level], tot) : 0;
            n_hit += (unsigned long long)tot;
        }
        __syncthreads();
        int at = s_base + s_warp[wid];
#pragma unroll
        for (int k = 0; k < HITLIST_PER_THREAD; k++) {
            if ((bal[k] >> lane) & 1u) lv.hits[at + __popc(bal[k] & lt)] = blk0 + (wid * HITLIST_PER_THREAD + k) * 32 + lane;
            at += __popc(bal[k]);
        }
        __syncthreads();                                       // s_warp / s_base are rewritten by the next block of rays
    }
    if (threadIdx.x == 0 && n_hit) { atomicAdd(&ctr->rays_hit, n_hit); if (level == 0) atomicAdd(&ctr->rays_hit_primary, n_hit); }
}

// rec slots
enum { REC_P = 0, REC_N = 3, REC_COL = 6, REC_REFL = 9, REC_TRANSP = 10, REC_KR = 11, REC_KD = 12, REC_KS = 13, REC_NEXP = 14, REC_INLEN = 15 };

// ------------------------------------------------------------------ camera rays: hit points in double (throughput mode)
// A float32 frame differs from the float64 reference mostly where a tiny error of the FIRST hit is amplified: the
// normal of a small mirror sphere turns an error of 1e-5 radii into a reflected ray that misses a silhouette the
// reference hits; a ground plane seen at a grazing angle turns an ulp of T into a texture cell.  The unrounded
// double-precision camera ray is one PCG jump away (camera_ray64), so for the winning object of every camera ray the
// reference's own test is run once more in double and the object-space hit point - of order one for a sphere, so
// float32 holds it to 1e-7 - is left in the P slots of the shading record, where k_shade picks it up instead of
// forming o + d T from numbers tens of radii long.  About 100 double operations per hit, a few % of a frame.
template <typename R, int PICKLESS>
__global__ void __launch_bounds__(256) k_primary_points(const __grid_constant__ DevScene<R> sc, const __grid_constant__ Level<R> lv,
                                                        const Counters* ctr, const __grid_constant__ CamSrc cam) {
    const int n = ctr->hits[0];
    const size_t cap = (size_t)lv.cap;
    for (int hs = blockIdx.x * blockDim.x + threadIdx.x; hs < n; hs += gridDim.x * blockDim.x) {      // hs: position in the hit list
        const int i = lv.hits[hs];
        int h; R t_pre;
        level_hit<R, PICKLESS>(lv, i, h, t_pre);
        const int leaf = h & HIT_LEAF_MASK, face = (h >> 28) & 7;
        double td = (double)(PICKLESS != 0 ? t_pre : lv.t[i]);
        double o64[3], d64[3];
        camera_ray64(sc, cam, i, o64, d64);
        Ray<double> w = {o64[0], o64[1], o64[2], d64[0], d64[1], d64[2]}, o;
        to_object(sc.w2o64 + (size_t)leaf * 12, w, o);
        if (leaf < sc.n_leaves) {                          // a plain object: its own test again, in double (a CSG member keeps the T of the span evaluation, itself double)
            double t2 = 0.0; int f2 = 0;
            if (hit_any_kind<double>(sc.kind[leaf], o, t2, f2) && f2 == face) td = t2;
        }
        lv.rec[(REC_P + 0) * cap + hs] = (R)(o.ox + o.dx * td);
        lv.rec[(REC_P + 1) * cap + hs] = (R)(o.oy + o.dy * td);
        lv.rec[(REC_P + 2) * cap + hs] = (R)(o.oz + o.dz * td);
    }
}

// ------------------------------------------------------------------ shade (traceRay :498-556 minus the recursion)
// rec slots

template <typename R, int PICKLESS>
__global__ void __launch_bounds__(SHADE_THREADS) k_shade(DevScene<R> sc, Level<R> lv, Level<R> next, ShadowQueue<R> sq,
                                                         unsigned char* occl, Counters* ctr, int level, int spawn,
                                                         R* spawn_to, int points) {
    const int n = ctr->hits[level];                        // the rays that hit something (Level::hits): every lane has work
    const bool use_points = points != 0;                   // level 0, throughput mode: object-space hit points precomputed in double
    __shared__ int s_append[2][SHADE_QUEUES][SHADE_WARPS];
    int append_phase = 0;
    // The ray index of the next iteration is fetched one iteration ahead, so that an iteration's dependent chain starts
    // at the fold key / ray record instead of one DRAM round trip earlier.
    const int stride = gridDim.x * blockDim.x;
    const int hs_first = blockIdx.x * blockDim.x + threadIdx.x;
    int i_nx = hs_first < n ? lv.hits[hs_first] : -1;
    for (int blk0 = blockIdx.x * blockDim.x; blk0 < n; blk0 += stride) {   // uniform per block
    const int hs = blk0 + threadIdx.x;                     // position in the hit list = index of everything per hit (rec, child, occl)
    const size_t cap = (size_t)lv.cap;                     // 64-bit strides: 15 * cap passes 2^31 on deep binary ray trees
    bool live = hs < n;
    const int i = i_nx;                                    // the ray
    i_nx = -1;
    if ((long long)hs + stride < n) i_nx = lv.hits[hs + stride];
    int h = -1; R t_pre = R(0);
    Ray<R> w = {R(0), R(0), R(0), R(0), R(0), R(1)};
    if (live) { load_ray(lv.ray, i, w); level_hit<R, PICKLESS>(lv, i, h, t_pre); }     // both need only i: two loads in flight, not one behind the other
    if (i_nx >= 0) {                                       // the next iteration's record and key on their way to L2
        prefetch_l2(lv.ray + (size_t)i_nx * RAY_WORDS);
        if (PICKLESS != 0) prefetch_l2(reinterpret_cast<const unsigned long long*>(lv.order) + i_nx);
    }
    live = live && h >= 0;
    V3<R> P = mk(R(0), R(0), R(0)), N = P;
    R mat[10];
    for (int k = 0; k < 10; k++) mat[k] = R(0);
    int leaf = 0;
    R obj_x = R(0), obj_z = R(0);                          // object-space hit point (x, z): the sort region of a point on an unbounded plane
    if (live) {
        leaf = h & HIT_LEAF_MASK;
        int face = h >> 28;
        const bool flip = (h & HIT_FLIP_BIT) != 0;          // CSG: surface of a subtracted solid
        const R t = PICKLESS != 0 ? t_pre : lv.t[i];
        const R* Mw = sc.w2o + (size_t)leaf * 12;
        Ray<R> o;
        to_object(Mw, w, o);
        const int kind = sc.kind[leaf];
        V3<R> p = mk(o.ox + o.dx * t, o.oy + o.dy * t, o.oz + o.dz * t);             // Hit.PointObj
        if (!GMLRT_PARITY_TU) {
            if (use_points) p = mk(lv.rec[(REC_P + 0) * cap + hs], lv.rec[(REC_P + 1) * cap + hs], lv.rec[(REC_P + 2) * cap + hs]);   // camera rays: k_primary_points
            else if (kind == K_SPHERE) {
                // o + d t adds two vectors tens of radii long to get one of length one (the normal of a sphere IS this point,
                // :119): go through the point of closest approach instead, where every term is of order one (see hit_sphere)
                const R a = o.dx * o.dx + o.dy * o.dy + o.dz * o.dz, hb = o.ox * o.dx + o.oy * o.dy + o.oz * o.dz;
                const R tc = -hb / a;
                const R lx = fma_(tc, o.dx, o.ox), ly = fma_(tc, o.dy, o.oy), lz = fma_(tc, o.dz, o.oz);
                const R pp = lx * lx + ly * ly + lz * lz;
                R sr = gm<R>::sqrt(gomath::max0(R(1) - pp) / a);
                if (t < tc) sr = -sr;                     // near root (a plain sphere: always; :98-102), far root: the inside of a sphere a CSG difference cut away
                p = mk(fma_(sr, o.dx, lx), fma_(sr, o.dy, ly), fma_(sr, o.dz, lz));
            }
        }
        obj_x = p.x; obj_z = p.z;
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
                if (GMLRT_PARITY_TU) {
                    const R ca = p.z / gm<R>::sqrt(R(1) - py * py);
                    u = gm<R>::acos(ca) / (R(2) * R(GM_PI));
                } else {
                    // acos(z / sqrt(1 - y^2)) = atan2(|x|, z) on the unit sphere.  acos is infinitely steep at +-1: in float32 it
                    // turns an ulp of its argument into 5e-5 of u on the x = 0 meridians - where u = 0 / 0.5 is exactly where
                    // textures put their seams (frac of a multiple of u).  atan2 is well conditioned everywhere.
                    u = gm<R>::atan2(gm<R>::abs(p.x), p.z) / (R(2) * R(GM_PI));
                }
            }
            N = p;                                                                   // :119 (sic: object-space point)
        } else if (kind == K_PLANE) {                                                // :182-205
            u = p.x; v = p.z; face = 0;
            N = ld3(sc.fnormal + (size_t)leaf * 18);
        } else if (kind == K_CUBE) {                                                 // :242-260
            u = p.x; v = p.z;
            if (!GMLRT_PARITY_TU) {                       // float rounding can leave a face point an ulp outside [0, 1): textures index arrays with floor(u n)
                u = fminf(fmaxf((float)u, 0.0f), 0.99999994f); v = fminf(fmaxf((float)v, 0.0f), 0.99999994f);
            }
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
            atomicExch(&ctr->error, 1);                  // the frame is discarded (GMLRT_E_SURFACE); the record below stays zero: no children, black
            live = false;
        } else {
            P = mul_point(sc.o2w + (size_t)leaf * 12, p);
            if (flip) N = neg(N);
        }
    }
    const R refl = mat[3], fuzz = mat[4], transp = mat[5], ior = mat[6];
    const V3<R> D = mk(w.dx, w.dy, w.dz);
    const int region = live ? region_of(sc, leaf, P, obj_x, obj_z) : 0;
    if (live) {
        R* rec = lv.rec + hs;
        rec[(REC_P + 0) * cap] = P.x; rec[(REC_P + 1) * cap] = P.y; rec[(REC_P + 2) * cap] = P.z;
        rec[(REC_N + 0) * cap] = N.x; rec[(REC_N + 1) * cap] = N.y; rec[(REC_N + 2) * cap] = N.z;
        // constant materials (folded at lowering, maybeSimplifySurfaceFn): the record names the surface and k_resolve reads the 9
        // words from the surface table (in L1 / L2) instead of 36 bytes per hit written here and read back there
        if (sc.rec_compact) *reinterpret_cast<int*>(rec + (REC_COL + 0) * cap) = sc.surface[leaf];
        else {
            rec[(REC_COL + 0) * cap] = mat[0]; rec[(REC_COL + 1) * cap] = mat[1]; rec[(REC_COL + 2) * cap] = mat[2];
            rec[REC_REFL * cap] = refl; rec[REC_TRANSP * cap] = transp;
        }
        R kr = R(0);
        if (transp > R(0)) {                                                         // fresnel :456-467
            R cosi = dot(D, N) / (length(D) * length(N));
            R r0 = (R(1) - ior) / (R(1) + ior);
            r0 = r0 * r0;
            kr = r0 + (R(1) - r0) * gm<R>::pow(R(1) - gm<R>::abs(cosi), R(5));
        }
        rec[REC_KR * cap] = kr;
        if (!sc.rec_compact) { rec[REC_KD * cap] = mat[7]; rec[REC_KS * cap] = mat[8]; rec[REC_NEXP * cap] = mat[9]; }
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
        if (!want_refl) for (int k = 0; k < 3; k++) lv.child[k * cap + hs] = R(0);
        if (!want_refr) for (int k = 3; k < 6; k++) lv.child[k * cap + hs] = R(0);
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
            sq.target[slot] = hs * sc.n_lights + li;
            occl[hs * sc.n_lights + li] = 0;
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
                const int key = ray_sort_key(sc, region, rd);
                store_ray(spawn_to, slot, ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, hs, key);
                next.hit[slot] = key;
                }
            }
            if (want_refr) {                                                         // refraction child (:532-556)
                const int slot = s[SHADE_Q_REFR][wid] + __popc(m_refr & lt);
                if (slot < next.cap) {
                const V3<R> ro = P - refr_n * R(1e-4);
                const int key = ray_sort_key(sc, region, refr_d);
                store_ray(spawn_to, slot, ro.x, ro.y, ro.z, refr_d.x, refr_d.y, refr_d.z, (int)((unsigned)hs | 0x80000000u), key);
                next.hit[slot] = key;
                }
            }
        }
    }
    }  // grid-stride loop
}

// ------------------------------------------------------------------ resolve one level (deepest first)
// computeLighting (:372-401) with the shadow flags, then the combine + clamp of traceRay (:506-508, :557-561), for the
// rays of the level that hit (one thread per entry of the hit list: dense warps, per-hit arrays read as streams).  The
// result goes into the parent's child slot, or - level 0 - into the sample colour of the camera ray.
template <typename R>
__global__ void __launch_bounds__(SHADE_THREADS) k_resolve(DevScene<R> sc, Level<R> lv, Level<R> prev, const unsigned char* occl,
                                                           const Counters* ctr, int level, int has_children) {
    const int n = ctr->hits[level];
    const size_t cap = (size_t)lv.cap;                     // 64-bit strides: 15 * cap passes 2^31 on deep binary ray trees
    for (int hs = blockIdx.x * blockDim.x + threadIdx.x; hs < n; hs += gridDim.x * blockDim.x) {
        const int i = lv.hits[hs];
        Ray<R> wr;
        load_ray(lv.ray, i, wr);
        const V3<R> D = mk(wr.dx, wr.dy, wr.dz);
        const R* rec = lv.rec + hs;
        const V3<R> P = mk(rec[(REC_P + 0) * cap], rec[(REC_P + 1) * cap], rec[(REC_P + 2) * cap]);
        const V3<R> N = mk(rec[(REC_N + 0) * cap], rec[(REC_N + 1) * cap], rec[(REC_N + 2) * cap]);
        V3<R> C; R refl, transp, kd, ks, nexp;
        const R kr = rec[REC_KR * cap];
        if (sc.rec_compact) {                                  // the surface's constant material (k_shade stored its index)
            const R* m = sc.surfaces[*reinterpret_cast<const int*>(rec + (REC_COL + 0) * cap)].material;
            C = mk(m[0], m[1], m[2]); refl = m[3]; transp = m[5]; kd = m[7]; ks = m[8]; nexp = m[9];
        } else {
            C = mk(rec[(REC_COL + 0) * cap], rec[(REC_COL + 1) * cap], rec[(REC_COL + 2) * cap]);
            refl = rec[REC_REFL * cap]; transp = rec[REC_TRANSP * cap];
            kd = rec[REC_KD * cap]; ks = rec[REC_KS * cap]; nexp = rec[REC_NEXP * cap];
        }
        const V3<R> Vv = neg(D);
        V3<R> L = ld3(sc.ambient) * kd;
        for (int li = 0; li < sc.n_lights; li++) {
            V3<R> ldir, lcol; R dist;
            if (!light_at(sc.lights[li], P, ldir, dist, lcol)) continue;
            if (occl[(size_t)hs * sc.n_lights + li]) continue;
            const R nDotL = gomath::max0(dot(N, ldir));
            const V3<R> diffuse = lcol * (nDotL * kd);
            const V3<R> H = normalize(Vv + ldir);
            const R spec = gomath::max0(dot(N, H));
            const V3<R> specular = lcol * (ks * gm<R>::pow(spec, nexp));
            L = (L + diffuse) + specular;
        }
        V3<R> out;
        if (refl == R(0) && transp == R(0)) {
            out = clamp3(hadamard(L, C));
        } else {
            const V3<R> zero = mk(R(0), R(0), R(0));             // the deepest level has no children: traceRay returns black at depth 0 (:488-491)
            const V3<R> rc = has_children ? mk(lv.child[hs], lv.child[cap + hs], lv.child[2 * cap + hs]) : zero;
            if (transp == R(0)) {
                out = clamp3(hadamard(L + rc * refl, C));
            } else {
                const V3<R> fc = has_children ? mk(lv.child[3 * cap + hs], lv.child[4 * cap + hs], lv.child[5 * cap + hs]) : zero;
                out = clamp3(hadamard(L * (R(1) - transp) + (rc * kr + fc * (R(1) - kr)), C));
            }
        }
        if (level == 0) {
            lv.samp[i] = out.x; lv.samp[cap + i] = out.y; lv.samp[2 * cap + i] = out.z;
        } else {
            const int p = ray_parent(lv.ray, i);
            const size_t pi = (size_t)(p & 0x7fffffff), pc = (size_t)prev.cap;
            const int base = (p < 0) ? 3 : 0;
            prev.child[(base + 0) * pc + pi] = out.x; prev.child[(base + 1) * pc + pi] = out.y; prev.child[(base + 2) * pc + pi] = out.z;
        }
    }
}

// ------------------------------------------------------------------ pixels (img.Set, raytracer.go:656; Vec3.RGBA, vec.go:104-107)
// One thread per camera ray; rays sit in (pixel, sample) order, so the 4 samples of a pixel are 4 neighbouring lanes.  A
// sample that hit has its colour in Level::samp (k_resolve), one that missed is the background of its direction.  Lane 0
// of the quad collects the four with shuffles, sums them in the reference's order - from zero, sample 0 first - scales
// by 0.25, quantises uint8(uint32(c * 65535) >> 8) and stores the RGBA8 pixel: into this device's frame or, in a
// multi-GPU frame, straight into the gathering GPU's (NVLink stores).
template <typename R, int PICKLESS>
__global__ void __launch_bounds__(256) k_pixels(DevScene<R> sc, Level<R> lv, const Counters* ctr, uchar4* frame, int y0, RowMap map) {
    const int n = ctr->count[0];
    const size_t cap = (size_t)lv.cap;
    // the loop bound is uniform per warp (the ray count and the grid stride are multiples of 32 / 4): every lane reaches the shuffles
    for (int i0 = blockIdx.x * blockDim.x + (threadIdx.x & ~31); i0 < n; i0 += gridDim.x * blockDim.x) {
        const int i = i0 + (threadIdx.x & 31);
        const bool valid = i < n;
        R c[3] = {R(0), R(0), R(0)};
        if (valid) {
            if (level_is_miss<R, PICKLESS>(lv, i)) {
                const V3<R> bg = background(sc, lv.ray[(size_t)i * RAY_WORDS + 4]);       // word 4: direction y
                c[0] = bg.x; c[1] = bg.y; c[2] = bg.z;
            } else { c[0] = lv.samp[i]; c[1] = lv.samp[cap + i]; c[2] = lv.samp[2 * cap + i]; }
        }
        const int q0 = (threadIdx.x & 31) & ~3;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const R s1 = __shfl_sync(0xffffffffu, c[k], q0 + 1), s2 = __shfl_sync(0xffffffffu, c[k], q0 + 2), s3 = __shfl_sync(0xffffffffu, c[k], q0 + 3);
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
            const int row = fast_div(pix, sc.w_mul, sc.w_shift);
            frame[(size_t)map.actual(y0 + row) * sc.W + (pix - row * sc.W)] = px;
        }
    }
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
        fuse_gen_ = env_size("GMLRT_FUSE_GEN", 1) != 0;
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
        CU_TRY(cudaMalloc(&scan_tiles_, 2 * SCAN_TILES_MAX * sizeof(int)));
        CU_TRY(cudaMemset(scan_tiles_, 0, 2 * SCAN_TILES_MAX * sizeof(int)));
        const int smem_max = 200 * 1024;
        auto allow = [&](auto kernel) -> int { CU_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max)); return GMLRT_OK; };
        int rc;
        if ((rc = allow(k_cull<R, CULL_PIVOT, GenCameraSource<R>, true>)) || (rc = allow(k_cull<R, CULL_PIVOT, CameraSource<R>, true>)) || (rc = allow(k_cull<R, CULL_GENERAL, SpawnSource<R>, true>)) ||
            (rc = allow(k_cull<R, CULL_PIVOT, ShadowSource<R>, true>)) || (rc = allow(k_cull<R, CULL_AXIS, ShadowSource<R>, true>)))
            return rc;
        return GMLRT_OK;
    }

    static constexpr bool kRobustPrimary = std::is_same<R, float>::value;   // f32: double re-test of borderline primary hits
    void reset_totals() override { totals_ = TraceTotals(); need_ctr_reset_ = true; ev_used_ = 0; ev_cls_.clear(); }
    void set_profiling(bool on) override { profile_ = on; }

    int fetch_totals(TraceTotals& out, std::string& err) override {
        CU_TRY(cudaSetDevice(dev_));
        Counters c;
        if (ctr_host_valid_) c = *ctr_host_;                 // copied behind the last kernel of the trace (trace_rows); the stream has been synchronised by the caller
        else CU_TRY(cudaMemcpy(&c, ctr_, sizeof c, cudaMemcpyDeviceToHost));
        ctr_host_valid_ = false;
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
        out.rays_hit_primary = c.rays_hit_primary; out.list_tests = list_tests_;
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
    void set_share(int n_parts) override { share_ = std::max(1, n_parts); }

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
        if (ctr_host_) cudaFreeHost(ctr_host_);
        ctr_host_ = nullptr;
        if (scan_tiles_) cudaFree(scan_tiles_);
        scan_tiles_ = nullptr;
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
    int resident_ = 1; size_t table_smem_ = 0;
    struct Launch { int grid = 148; size_t smem = 0; };
    size_t cfg_smem_ = ~(size_t)0;
    Launch l_gen_; bool fuse_gen_ = true;                                                // camera rays made by the cull kernel (GMLRT_FUSE_GEN, read at context creation)
    Launch l_primary_, l_closest_, l_shadow_pivot_, l_shadow_axis_;                      // the four k_cull instantiations
    // ray classes (see "ray classes" above): one candidate list per bin, built once per upload
    static constexpr int TILE_W = 32, TILE_H = 8;                // camera-ray class = screen tile (a warp's 8 pixels x 4 samples share one)
    BinLists lists_camera_{}, lists_spawn_{}; std::vector<BinLists> lists_light_;
    std::vector<double> region_bound_;                           // host: [n_regions][4] sphere holding every point of the region (inf = unbounded)
    unsigned long long* list_tests_dev_ = nullptr; unsigned long long list_tests_ = 0;
    bool primary_points_ = false;                                // throughput mode: k_primary_points before the level-0 k_shade (GMLRT_PRIMARY_POINTS)
    bool sort_shadows_ = false;                                  // GMLRT_SORT_SHADOWS: order the shadow rays of a light by class before the cull
    int build_lists(const HostScene& S, std::string& err);
    // exclusive scan of a histogram on `st`: tiles across co-resident blocks when there are many bins, else one block
    void launch_scan(int nb, int* hist, int* cursor, int clear, cudaStream_t st) {
        const int tiles = (nb + SCAN_TILE - 1) / SCAN_TILE;
        if (tiles > 2 && tiles <= SCAN_TILES_MAX && scan_tiles_ && st == stream_) {
            k_sort_scan_tiles<<<std::min(tiles, sms_), 1024, 0, st>>>(nb, hist, cursor, clear, scan_tiles_, scan_tiles_ + SCAN_TILES_MAX, ++scan_epoch_);
        } else k_sort_scan<<<1, 1024, 0, st>>>(nb, hist, cursor, clear);
    }
    int* scan_tiles_ = nullptr; int scan_epoch_ = 0;             // tile sums + flags of k_sort_scan_tiles (main stream only)
    CandQueue cq_{}; int cand_per_ray_ = 4, tile_shrink_ = 0; unsigned long long scene_sig_ = ~0ull, planned_sig_ = ~0ull - 1; int planned_rows_full_ = 0;
    int* sort_hist_ = nullptr; int* sort_cursor_ = nullptr; int sort_nb_ = 2, sort_shift_ = 0, n_regions_ = 1; bool sort_clear_ = true;
    int* key_hist_ = nullptr; int* key_cursor_ = nullptr; int key_nb_ = 6;      // spawned rays by (origin leaf, direction cell); cleared with sort_hist_
    bool sort_spawned_ = true;
    bool sort_perm_ = true;                                      // GMLRT_SORT_PERM: the key sort of spawned rays leaves a permutation (in spawn_stage_), the records stay in queue order
    R* spawn_stage_ = nullptr;
    // cull tables (device) and what they were built for: any ray, camera rays, one per light
    std::vector<double> bound_;                                  // host: [n][4] bounding spheres (centre, radius) of the objects
    R* cull_general_ = nullptr; R* cull_eye_ = nullptr; std::vector<R*> cull_light_;
    CullFrame<R> frame_general_{}, frame_eye_{}; std::vector<CullFrame<R>> frame_light_;
    Counters* ctr_host_ = nullptr; bool ctr_host_valid_ = false;   // page-locked copy of the frame's counters (trace_rows -> fetch_totals)
    int share_ = 1;                                              // devices sharing the frame (set_share)
    int* slots_ = nullptr; int n_slots_ = 0;                     // per-frame queue counters: per level 1 (closest candidates) + per light (shadow candidates, shadow rays)
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
    const int n_cull = std::max((n + 1 + GROUP - 1) / GROUP * GROUP, GROUP);     // always at least one padding record: the candidate lists pad with it
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
        // What every region holds (for the ray classes, build_lists): a sphere around every point region_of() can map to it.
        region_bound_.assign((size_t)n_regions * 4, 0.0);
        auto set_region = [&](int r, double x, double y, double z, double rad) {
            double* q = &region_bound_[(size_t)r * 4];
            q[0] = x; q[1] = y; q[2] = z; q[3] = rad;
        };
        for (size_t k = 0; k < small.size(); k++) {                  // a small object: its bounding sphere
            const double* b = &bound_[(size_t)small[k].second * 4];
            set_region((int)k, b[0], b[1], b[2], b[3]);
        }
        if (any_large) {                                             // grid cells; points outside the grid are clamped into its border cells
            bool open_border = false;                                // ... which only unbounded objects have
            for (int d = 0; d < n; d++) if (rid_of[d] == -1 && !std::isfinite(bound_[(size_t)d * 4 + 3])) open_border = true;
            for (int cx = 0; cx < dim[0]; cx++) for (int cy = 0; cy < dim[1]; cy++) for (int cz = 0; cz < dim[2]; cz++) {
                const int r = sc.n_small + (cx * dim[1] + cy) * dim[2] + cz;
                const bool border = cx == 0 || cy == 0 || cz == 0 || cx == dim[0] - 1 || cy == dim[1] - 1 || cz == dim[2] - 1;
                set_region(r, lo3[0] + (cx + 0.5) * edge, lo3[1] + (cy + 0.5) * edge, lo3[2] + (cz + 0.5) * edge,
                           (border && open_border) ? INFINITY : edge * 0.8660254037844387 * 1.001);
            }
        }
        for (int k = 0; k < n_plane_grids; k++) {                    // cells of a plane's own (x, z) grid, see region_of()
            int dplane = -1;
            for (int d = 0; d < n; d++) if (rid_of[d] == -2 - k) dplane = d;
            const HostLeaf& L = S.leaves[order[dplane]];
            const double par0 = (double)sc.plane_par[k][0], par1 = (double)sc.plane_par[k][1], par2 = (double)sc.plane_par[k][2];
            const int G = GMLRT_PLANE_GRID;
            auto edge_of = [&](int c, double par_c) {                // object-space coordinate where cell c begins: g = 2 c / G - 1, x = g / (1 - |g|)
                const double g = 2.0 * c / G - 1.0;
                return g / (1.0 - std::fabs(g)) / par2 + par_c;
            };
            for (int cx = 0; cx < G; cx++) for (int cz = 0; cz < G; cz++) {
                const int r = sc.plane_base[k] + cx * G + ((cx & 1) ? G - 1 - cz : cz);
                if (cx == 0 || cz == 0 || cx == G - 1 || cz == G - 1) { set_region(r, 0, 0, 0, INFINITY); continue; }     // the outer ring runs to infinity
                double x0 = edge_of(cx, par0), x1 = edge_of(cx + 1, par0), z0 = edge_of(cz, par1), z1 = edge_of(cz + 1, par1);
                const double mx = 1e-3 * (x1 - x0) + 1e-5 * (std::fabs(x0) + std::fabs(x1)), mz = 1e-3 * (z1 - z0) + 1e-5 * (std::fabs(z0) + std::fabs(z1));
                x0 -= mx; x1 += mx; z0 -= mz; z1 += mz;
                double c[4][3], mid[3] = {0, 0, 0};
                for (int q = 0; q < 4; q++) {
                    const double px = (q & 1) ? x1 : x0, pz = (q & 2) ? z1 : z0;
                    for (int a = 0; a < 3; a++) { c[q][a] = L.o2w[4 * a] * px + L.o2w[4 * a + 2] * pz + L.o2w[4 * a + 3]; mid[a] += 0.25 * c[q][a]; }
                }
                double rad = 0;
                for (int q = 0; q < 4; q++) rad = std::max(rad, std::sqrt((c[q][0] - mid[0]) * (c[q][0] - mid[0]) + (c[q][1] - mid[1]) * (c[q][1] - mid[1]) + (c[q][2] - mid[2]) * (c[q][2] - mid[2])));
                set_region(r, mid[0], mid[1], mid[2], rad * 1.0001);
            }
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
        bool all_const = true;
        for (const gmlrt_surface& sf : S.surfaces) all_const = all_const && sf.kind == GMLRT_SURFACE_CONST;
        sc.rec_compact = (all_const && env_size("GMLRT_COMPACT_REC", 1) != 0) ? 1 : 0;
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
    { const FastDiv fd = fast_div_of(S.width); sc.w_mul = fd.mul; sc.w_shift = fd.shift; }
    sc.cam_su = S.vw / (double)(S.width - 1); sc.cam_sv = S.vh / (double)(S.height - 1);
    // sort key of spawned rays: 6 k^2 direction cells per origin leaf, at most KEY_BINS_MAX bins in all
    // Direction cells per cube-map face edge: finer cells = shorter candidate lists per class.  6 x 6 measured best both for a
    // whole 4K frame of 1024 spheres (5.95 ms; 4 x 4: 6.09) and for one device's eighth of it (1.01 ms; 4 x 4: 1.05, 3 x 3:
    // 1.10), where a class holds only a handful of rays: shorter lists win over lanes that share a list.
    sc.key_k = (int)env_size("GMLRT_KEY_K", 6); sc.key_k = std::max(1, std::min(sc.key_k, 8));
    sc.key_shift = 0;
    const size_t key_bins_limit = std::min<size_t>(std::max<size_t>(env_size("GMLRT_KEY_BINS", (size_t)1 << 19), 6 * 64), (size_t)KEY_BINS_MAX);
    while ((size_t)(((n_regions - 1) >> sc.key_shift) + 1) * 6 * sc.key_k * sc.key_k > key_bins_limit) sc.key_shift++;
    key_nb_ = (((n_regions - 1) >> sc.key_shift) + 1) * 6 * sc.key_k * sc.key_k;
    sort_spawned_ = env_size("GMLRT_SORT_SPAWNED", 1) != 0;
    {   // Permutation only (the records stay in queue order, the cull kernel gathers them one sector each) saves the staging
        // buffer's round trip: 4K spheres1024 4.94 -> 4.72 ms.  In a deep, large scene the candidates of a warp then point at
        // records all over the level and the exact kernels lose more than the sort gains (8K mixed4096, depth 5: 102.3 -> 105.3 ms).
        const size_t mode = env_size("GMLRT_SORT_PERM", 2);
        sort_perm_ = mode == 2 ? sc.n_leaves < 2048 : mode != 0;
    }
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
    sort_nb_ = std::min(n_regions_ + 1, SORT_BINS_MAX); sort_shift_ = 0;        // class of a shadow ray: the region of the shaded point
    while (((n_regions_ - 1) >> sort_shift_) + 2 > sort_nb_) sort_shift_++;
    // Shadow rays reach the cull in queue order: the order of the hits that spawned them, i.e. of neighbouring pixels or of
    // sorted spawned rays, which mostly share a class already (4K spheres1024: sorting them again costs 0.4 ms and gains
    // nothing).  With thousands of small objects a warp's 32 rays straddle many classes and every lane walks its own
    // list: then the counting sort by class pays for itself (8K mixed4096: 143 -> 116 ms per frame).
    {
        const size_t mode = env_size("GMLRT_SORT_SHADOWS", 2);
        sort_shadows_ = mode == 2 ? sc.n_leaves >= 2048 : mode != 0;
    }
    primary_points_ = kRobustPrimary && env_size("GMLRT_PRIMARY_POINTS", 1) != 0;
    const double t_l0 = now();
    if ((rc = build_lists(S, err))) return rc;
    const double t_l1 = now();
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
    if (dbg_t) fprintf(stderr, "[gmlrt upload dev %d] sync+free %.3f  host build + enqueue copies %.3f (of which ray-class lists %.3f)  copy sync %.3f  frame+plan_tiles %.3f ms\n",
                       dev_, t_b - t_a, t_c - t_b, t_l1 - t_l0, t_d - t_c, now() - t_d);
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

// The candidate lists of the ray classes (see "ray classes"): camera rays by screen tile, spawned rays by (region, direction
// cell), the shadow rays of each light by region.  Two passes per set (count, scan, fill); the totals of all sets come
// back in one synchronisation.
template <typename R>
int Renderer<R>::build_lists(const HostScene& S, std::string& err) {
    const int n = sc_.n_leaves;
    int rc;
    // regions merged the way the keys merge them
    auto merged = [&](int shift, int total) {
        const int groups = std::min(((n_regions_ - 1) >> shift) + 1, total);
        std::vector<double> out((size_t)total * 4, 0.0);
        for (int g = 0; g < total; g++) out[(size_t)g * 4 + 3] = INFINITY;          // bins beyond the regions (unused, invalid keys): no bound
        for (int g = 0; g < groups; g++) {
            double C[3] = {0, 0, 0}, Rr = -1.0;
            for (int r = g << shift; r < std::min(n_regions_, (g + 1) << shift); r++) {
                const double* q = &region_bound_[(size_t)r * 4];
                if (Rr < 0) { C[0] = q[0]; C[1] = q[1]; C[2] = q[2]; Rr = q[3]; }
                else { double C2[3], R2; sphere_union(C, Rr, q, q[3], C2, R2); C[0] = C2[0]; C[1] = C2[1]; C[2] = C2[2]; Rr = R2; }
            }
            double* o = &out[(size_t)g * 4];
            o[0] = C[0]; o[1] = C[1]; o[2] = C[2]; o[3] = (Rr >= 0 && Rr == Rr) ? Rr : INFINITY;
        }
        return out;
    };
    // the cone of directions of every cube-map cell of ray_sort_key()
    const int k = sc_.key_k, n_cells = 6 * k * k;
    std::vector<double> cells((size_t)n_cells * 5);
    for (int face = 0; face < 6; face++) for (int cu = 0; cu < k; cu++) for (int cv = 0; cv < k; cv++) {
        auto dir = [&](double U, double V, double* d) {
            const double sgn = (face & 1) ? -1.0 : 1.0;
            if (face < 2) { d[0] = sgn; d[1] = U; d[2] = V; } else if (face < 4) { d[0] = U; d[1] = sgn; d[2] = V; } else { d[0] = U; d[1] = V; d[2] = sgn; }
            normalize3(d);
        };
        const double u0 = 2.0 * cu / k - 1.0, u1 = 2.0 * (cu + 1) / k - 1.0, v0 = 2.0 * cv / k - 1.0, v1 = 2.0 * (cv + 1) / k - 1.0;
        double ax[3], c = 1.0;
        dir(0.5 * (u0 + u1), 0.5 * (v0 + v1), ax);
        for (int q = 0; q < 4; q++) {
            double d[3];
            dir((q & 1) ? u1 : u0, (q & 2) ? v1 : v0, d);
            c = std::min(c, ax[0] * d[0] + ax[1] * d[1] + ax[2] * d[2]);
        }
        const double th = std::acos(std::min(c, 1.0)) + 2e-6;          // the key is taken from the rounded direction: a cell is that much fuzzy
        double* o = &cells[(size_t)((face * k + cu) * k + ((cu & 1) ? k - 1 - cv : cv)) * 5];
        o[0] = ax[0]; o[1] = ax[1]; o[2] = ax[2]; o[3] = std::cos(th); o[4] = std::sin(th);
    }
    const bool dbg_t = getenv("GMLRT_DEBUG_UPLOAD") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    const std::vector<double> reg_spawn = merged(sc_.key_shift, key_nb_ / n_cells), reg_shadow = merged(sort_shift_, sort_nb_);
    double *d_cells, *d_reg_spawn, *d_reg_shadow;
    if ((rc = to_device(cells, &d_cells, err)) || (rc = to_device(reg_spawn, &d_reg_spawn, err)) ||
        (rc = to_device(reg_shadow, &d_reg_shadow, err))) return rc;
    if (!list_tests_dev_) { CU_TRY(cudaMalloc(&list_tests_dev_, sizeof(unsigned long long))); }
    CU_TRY(cudaMemsetAsync(list_tests_dev_, 0, sizeof(unsigned long long), stream_));

    struct Pending { BinParams p; BinLists* out; int* count; int* first; };
    std::vector<Pending> sets;
    // The objects as a family's per-ray test sees them (build_cull_table): the radius grown by the relative inflation, the
    // absolute widening and the rounding allowance relative to the squared magnitudes (taken four times over).
    auto base_params = [&](int kind, int n_bins, double offset_abs, const double* pivot, const CullFrame<R>* fr, int& prc) {
        BinParams p{};
        p.kind = kind; p.n_bins = n_bins; p.n_objects = n;
        p.no_bounds = env_size("GMLRT_NO_CLASS_BOUNDS", 0) != 0;
        std::vector<double> rec((size_t)std::max(n, 1) * 4, 0.0);
        for (int d = 0; d < n; d++) {
            const double* b = &bound_[(size_t)d * 4];
            double* o = &rec[(size_t)d * 4];
            if (!std::isfinite(b[3])) { o[3] = INFINITY; continue; }
            const double Rm = b[3] * (1.0 + CullMargin<R>::radius) + offset_abs;
            const double c[3] = {b[0] - (pivot ? pivot[0] : 0.0), b[1] - (pivot ? pivot[1] : 0.0), b[2] - (pivot ? pivot[2] : 0.0)};
            const double cc = c[0] * c[0] + c[1] * c[1] + c[2] * c[2];
            o[3] = std::sqrt(Rm * Rm + 4.0 * CullMargin<R>::rel * (cc + 1.0)) * (1.0 + 1e-9);
            if (kind == BIN_SHADOW_AXIS) {
                o[0] = b[0] * (double)fr->e1[0] + b[1] * (double)fr->e1[1] + b[2] * (double)fr->e1[2];
                o[1] = b[0] * (double)fr->e2[0] + b[1] * (double)fr->e2[1] + b[2] * (double)fr->e2[2];
                o[2] = b[0] * (double)fr->axis[0] + b[1] * (double)fr->axis[1] + b[2] * (double)fr->axis[2] + Rm + 1e-3 * (std::sqrt(cc) + 1.0);
            } else { o[0] = c[0]; o[1] = c[1]; o[2] = c[2]; }
        }
        double* d_rec = nullptr;
        prc = to_device(rec, &d_rec, err);
        p.obj = d_rec;
        std::vector<float4> recf((size_t)std::max(n, 1));
        for (int d = 0; d < n; d++) {
            const double* o = &rec[(size_t)d * 4];
            const double mag = std::fabs(o[0]) + std::fabs(o[1]) + std::fabs(o[2]);
            recf[d] = make_float4((float)o[0], (float)o[1], (float)o[2], std::isfinite(o[3]) ? (float)(o[3] * 1.001 + 1e-4 + 1e-6 * mag) : INFINITY);
        }
        float4* d_recf = nullptr;
        if (!prc) prc = to_device(recf, &d_recf, err);
        p.objf = env_size("GMLRT_NO_SCREEN", 0) != 0 ? nullptr : d_recf;
        return p;
    };
    lists_light_.assign(S.lights.size(), BinLists{});
    {   // camera rays (the eye, raytracer.go:605-609)
        const int tx = (S.width + TILE_W - 1) / TILE_W, ty = (S.height + TILE_H - 1) / TILE_H;
        const double eye[3] = {0.0, 0.0, -1.0};
        BinParams p = base_params(BIN_CAMERA, tx * ty, 1e-5, eye, nullptr, rc);
        if (rc) return rc;
        p.pivot[0] = 0.0; p.pivot[1] = 0.0; p.pivot[2] = -1.0;
        p.W = S.width; p.H = S.height; p.tile_w = TILE_W; p.tile_h = TILE_H; p.tiles_x = tx;
        p.cam_su = sc_.cam_su; p.cam_sv = sc_.cam_sv; p.vw = sc_.vw; p.vh = sc_.vh;
        sets.push_back({p, &lists_camera_, nullptr, nullptr});
    }
    {   // reflection / refraction rays
        BinParams p = base_params(BIN_SPAWN, key_nb_, 0.0, nullptr, nullptr, rc);
        if (rc) return rc;
        p.region = d_reg_spawn; p.n_regions = (int)(reg_spawn.size() / 4); p.cell = d_cells; p.n_cells = n_cells;
        sets.push_back({p, &lists_spawn_, nullptr, nullptr});
    }
    for (size_t i = 0; i < S.lights.size(); i++) {
        const gmlrt_light& l = S.lights[i];
        const bool axis = l.kind == GMLRT_LIGHT_DIRECTIONAL;
        BinParams p = base_params(axis ? BIN_SHADOW_AXIS : BIN_SHADOW_PIVOT, sort_nb_, axis ? 1e-5 : 1.5e-4, axis ? nullptr : l.pos, &frame_light_[i], rc);
        if (rc) return rc;
        p.region = d_reg_shadow; p.n_regions = (int)(reg_shadow.size() / 4);
        for (int a = 0; a < 3; a++) {
            p.pivot[a] = l.pos[a];
            p.e1[a] = (double)frame_light_[i].e1[a]; p.e2[a] = (double)frame_light_[i].e2[a]; p.axis[a] = (double)frame_light_[i].axis[a];
        }
        // inShadow accepts T * |incoming direction| < dist (:424): with |incoming direction| = 1 (camera and reflection rays
        // are normalised) an occluder lies between the point and the light; refraction rays are not normalised (:438-450)
        p.limit_reach = S.any_refract ? 0 : 1;
        sets.push_back({p, &lists_light_[i], nullptr, nullptr});
    }
    std::vector<int> tails(sets.size() * 2, 0);
    for (size_t s = 0; s < sets.size(); s++) {
        Pending& ps = sets[s];
        void *pc = nullptr, *pf = nullptr;
        if ((rc = scene_alloc((size_t)ps.p.n_bins * sizeof(int), &pc, err)) || (rc = scene_alloc((size_t)ps.p.n_bins * sizeof(int), &pf, err))) return rc;
        ps.count = (int*)pc; ps.first = (int*)pf;
        const int grid = std::min((ps.p.n_bins + 7) / 8, sms_ * 32);
        k_list_build<false><<<grid, 256, 0, stream_>>>(ps.p, ps.count, nullptr, nullptr, nullptr, 0u, list_tests_dev_);
        launch_scan(ps.p.n_bins, ps.count, ps.first, 0, stream_);
        CU_TRY(cudaMemcpyAsync(&tails[2 * s], ps.count + ps.p.n_bins - 1, sizeof(int), cudaMemcpyDeviceToHost, stream_));
        CU_TRY(cudaMemcpyAsync(&tails[2 * s + 1], ps.first + ps.p.n_bins - 1, sizeof(int), cudaMemcpyDeviceToHost, stream_));
    }
    const double t1 = now();
    CU_TRY(cudaMemcpyAsync(&list_tests_, list_tests_dev_, sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream_));
    CU_TRY(cudaStreamSynchronize(stream_));
    const double t2 = now();
    list_tests_ *= 2;                                                  // the fill pass repeats the tests
    for (size_t s = 0; s < sets.size(); s++) {
        Pending& ps = sets[s];
        const long long total = (long long)tails[2 * s] + tails[2 * s + 1];
        if (total < 0 || tails[2 * s + 1] < 0) { err = "candidate lists of the ray classes exceed 2^31 entries"; return GMLRT_E_UNSUPPORTED; }
        void *pe = nullptr, *pb = nullptr;
        if ((rc = scene_alloc(std::max<size_t>((size_t)total, 4) * sizeof(uint32_t), &pe, err)) || (rc = scene_alloc((size_t)ps.p.n_bins * sizeof(int2), &pb, err))) return rc;
        const int grid = std::min((ps.p.n_bins + 7) / 8, sms_ * 32);
        k_list_build<true><<<grid, 256, 0, stream_>>>(ps.p, nullptr, ps.first, (int2*)pb, (uint32_t*)pe, (uint32_t)(4 * sizeof(R)), nullptr);
        ps.out->bin = (const int2*)pb; ps.out->entry = (const uint32_t*)pe; ps.out->n_bins = ps.p.n_bins;
    }
    if (dbg_t) {
        const double t3 = now();
        cudaStreamSynchronize(stream_);
        fprintf(stderr, "[gmlrt lists dev %d] host prep + count/scan enqueue %.3f  wait %.3f  alloc + fill enqueue %.3f  fill wait %.3f ms; bins %d / %d / %d, entries", dev_,
                t1 - t0, t2 - t1, t3 - t2, now() - t3, lists_camera_.n_bins, lists_spawn_.n_bins, sort_nb_);
        for (size_t q = 0; q < sets.size(); q++) fprintf(stderr, " %lld", (long long)tails[2 * q] + tails[2 * q + 1]);
        fprintf(stderr, "\n");
    }
    CU_TRY(cudaGetLastError());
    return GMLRT_OK;
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
    const double node_bytes = (2 * RAY_WORDS + NREC + 8) * sizeof(R) + 16 + nl;      // ray record (+ its staging copy), rec, child, hit, order, hit list, occl
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
    struct LvOff { size_t ray, parent, hit, t, order, rec, child, occl, hits, samp; size_t cap; };
    std::vector<LvOff> lo(depth_eff_);
    size_t max_cap = 0, stage_cap = 0;
    for (int l = 0; l < depth_eff_; l++) {
        size_t cap = (size_t)std::min(std::min((double)S0 * std::ldexp(1.0, std::min(l, 30)), (double)S0_full * mult[l]), 2.0e9);   // never above the true worst case
        lo[l].cap = cap; max_cap = std::max(max_cap, cap);
        lo[l].ray = take(RAY_WORDS * cap * sizeof(R)); lo[l].parent = 0; lo[l].hit = take(cap * 4);
        if (l > 0) stage_cap = std::max(stage_cap, cap);
        lo[l].t = take(cap * sizeof(R)); lo[l].order = take(cap * 8); lo[l].rec = take(NREC * cap * sizeof(R)); lo[l].child = take(6 * cap * sizeof(R));
        lo[l].occl = take(cap * std::max(nl, 1));
        lo[l].hits = take(cap * 4); lo[l].samp = l == 0 ? take(3 * cap * sizeof(R)) : 0;
    }
    const size_t sq_cap = max_cap * std::max(nl, 1);
    size_t o_sq = take(8 * sq_cap * sizeof(R)), o_skip = take(sq_cap * 4), o_key = take(sq_cap * 4), o_tgt = take(sq_cap * 4), o_first = take(sq_cap * 4), o_perm = take(sq_cap * 4), o_sort = take(2 * SHADOW_STREAMS * SORT_BINS_MAX * sizeof(int));
    size_t o_keysort = take(2 * (size_t)KEY_BINS_MAX * sizeof(int));
    n_slots_ = depth_eff_ * (1 + 2 * std::max(nl, 0)) + 4;
    size_t o_slots = take((size_t)n_slots_ * sizeof(int));
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
        L.hits = (int*)(base + lo[l].hits); L.samp = l == 0 ? (R*)(base + lo[l].samp) : nullptr;
    }
    sq_.q = (R*)(base + o_sq); sq_.skip = (int*)(base + o_skip); sq_.key = (int*)(base + o_key); sq_.target = (int*)(base + o_tgt); sq_.cap = (int)sq_cap;
    sq_.per_light = (int)max_cap; sq_.count = nullptr; sq_.first = (int*)(base + o_first); sq_.perm = (int*)(base + o_perm);
    sort_hist_ = (int*)(base + o_sort); sort_cursor_ = sort_hist_ + SORT_BINS_MAX; sort_clear_ = true;
    key_hist_ = (int*)(base + o_keysort); key_cursor_ = key_hist_ + KEY_BINS_MAX;
    spawn_stage_ = (R*)(base + o_stage);
    slots_ = (int*)(base + o_slots);
    cq_.ray = (int*)(base + o_cray); cq_.obj = (int*)(base + o_cobj); cq_.t = (void*)(base + o_ct); cq_.count = (int*)(base + o_ccount);
    cq_.cap = (int)cand_cap;
    for (int k = 0; k < SHADOW_STREAMS; k++) {
        cq_shadow_[k].ray = (int*)(base + o_cray2[k]); cq_shadow_[k].obj = (int*)(base + o_cobj2[k]); cq_shadow_[k].t = nullptr;
        cq_shadow_[k].count = (int*)(base + o_ccount2[k]); cq_shadow_[k].cap = (int)cand_cap;
        sort_hist_sh_[k] = sort_hist_ + (size_t)2 * k * SORT_BINS_MAX; sort_cursor_sh_[k] = sort_hist_sh_[k] + SORT_BINS_MAX;
    }

    // cull kernels: the family's table staged into shared memory when it fits, else read through L1
    const size_t rec_bytes = 4 * sizeof(R);
    const size_t resident_max = env_size("GMLRT_RESIDENT_KB", 128) * 1024;
    const size_t all = (size_t)sc_.n_cull * rec_bytes;
    resident_ = all <= resident_max ? 1 : 0;
    table_smem_ = resident_ ? all : 0;
    if (cfg_smem_ == table_smem_ + (size_t)resident_) return GMLRT_OK;      // launch shapes depend on the staged table size only
    cfg_smem_ = table_smem_ + (size_t)resident_;
    auto cfg = [&](auto kernel, Launch& L) -> int {
        int occ = 1;
        L.smem = table_smem_;
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, CULL_THREADS, L.smem));
        L.grid = sms_ * std::max(occ, 1);
        return GMLRT_OK;
    };
    int rc;
    if (resident_) {
        if ((rc = cfg(k_cull<R, CULL_PIVOT, GenCameraSource<R>, true>, l_gen_)) || (rc = cfg(k_cull<R, CULL_PIVOT, CameraSource<R>, true>, l_primary_)) || (rc = cfg(k_cull<R, CULL_GENERAL, SpawnSource<R>, true>, l_closest_)) ||
            (rc = cfg(k_cull<R, CULL_PIVOT, ShadowSource<R>, true>, l_shadow_pivot_)) || (rc = cfg(k_cull<R, CULL_AXIS, ShadowSource<R>, true>, l_shadow_axis_)))
            return rc;
    } else {
        if ((rc = cfg(k_cull<R, CULL_PIVOT, GenCameraSource<R>, false>, l_gen_)) || (rc = cfg(k_cull<R, CULL_PIVOT, CameraSource<R>, false>, l_primary_)) || (rc = cfg(k_cull<R, CULL_GENERAL, SpawnSource<R>, false>, l_closest_)) ||
            (rc = cfg(k_cull<R, CULL_PIVOT, ShadowSource<R>, false>, l_shadow_pivot_)) || (rc = cfg(k_cull<R, CULL_AXIS, ShadowSource<R>, false>, l_shadow_axis_)))
            return rc;
    }
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
    auto exact_closest = [&](int l, bool count, const CandQueue& cq_l) -> int {
        const int grid = sms_ * 16;
        if (have_spheres) {
            if (count) k_exact_closest<R, false, true, EXACT_SPHERES><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
            else k_exact_closest<R, false, false, EXACT_SPHERES><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
        }
        if (have_simple) {
            if (count) k_exact_closest<R, false, true, EXACT_SIMPLE><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
            else k_exact_closest<R, false, false, EXACT_SIMPLE><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
        }
        if (have_csg) {
            if (count) k_exact_closest<R, false, true, EXACT_CSG><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
            else k_exact_closest<R, false, false, EXACT_CSG><<<grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
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
        if (fuse_gen_) k_frame_init<<<std::max(1, std::min((n_slots_ + 255) / 256, 64)), 256, 0, stream_>>>(ctr_, S, slots_, n_slots_);
        else k_gen<R><<<std::min((S + 255) / 256, sms_ * 32), 256, 0, stream_>>>(sc_, levels_[0], ctr_, ty, rows, map, slots_, n_slots_);
        totals_.launches++;
        totals_.rays_primary += (unsigned long long)S;
        for (int l = 0; l < depth_eff_; l++) {
            const int cap_blocks = (levels_[l].cap + SHADE_THREADS - 1) / SHADE_THREADS;
            const int exact_grid = sms_ * 16;
            // every launch of the frame that counts into a queue has its own counter (zeroed by k_gen): no memsets between launches
            int* const slot = slots_ + (size_t)l * (1 + 2 * sc_.n_lights);
            CandQueue cq_l = cq_; cq_l.count = slot;
            ShadowQueue<R> sq_l = sq_; sq_l.count = slot + 1 + sc_.n_lights;
            if (sort_clear_) {
                CU_TRY(cudaMemsetAsync(sort_hist_, 0, 2 * SHADOW_STREAMS * SORT_BINS_MAX * sizeof(int), stream_));
                CU_TRY(cudaMemsetAsync(key_hist_, 0, KEY_BINS_MAX * sizeof(int), stream_));
                sort_clear_ = false;
            }
            if (l > 0 && sort_spawned_) {                  // rays by (origin leaf, direction cell): keys in lv.hit, records in the staging buffer
                MARK(KC_SORT);
                k_keysort_hist<<<sms_ * 8, 256, 0, stream_>>>(levels_[l].hit, &ctr_->count[l], key_nb_, key_hist_);
                launch_scan(key_nb_, key_hist_, key_cursor_, 1, stream_);
                k_keysort_scatter<R><<<sms_ * 8, 256, 0, stream_>>>(levels_[l].hit, &ctr_->count[l], key_nb_, key_cursor_, spawn_stage_, levels_[l].ray,
                                                                    sort_perm_ ? reinterpret_cast<int*>(spawn_stage_) : nullptr);
                totals_.launches += 3;
            }
            if (l == 0) {                                  // camera rays: PIVOT family through the eye, robust re-tests in f32
                MARK(KC_CLOSEST_PRIMARY);
                static_assert(TILE_W == 32 && TILE_H == 8, "CameraSource::bin shifts by log2 of the tile size");
                CameraSource<R> cams{levels_[l], W, ty, map, 5, 3, (W + TILE_W - 1) / TILE_W, sc_.w_mul, sc_.w_shift, sc_.pickless != 0};
                if (fuse_gen_) {
                    GenCameraSource<R> gens{sc_, levels_[l], W, cam, 5, 3, (W + TILE_W - 1) / TILE_W, sc_.pickless != 0};
                    if (resident_) k_cull<R, CULL_PIVOT, GenCameraSource<R>, true><<<l_gen_.grid, CULL_THREADS, l_gen_.smem, stream_>>>(
                        gens, &ctr_->count[l], cull_eye_, frame_eye_, sc_.n_cull, lists_camera_, cq_l, ctr_);
                    else k_cull<R, CULL_PIVOT, GenCameraSource<R>, false><<<l_gen_.grid, CULL_THREADS, 0, stream_>>>(
                        gens, &ctr_->count[l], cull_eye_, frame_eye_, sc_.n_cull, lists_camera_, cq_l, ctr_);
                } else
                if (resident_) k_cull<R, CULL_PIVOT, CameraSource<R>, true><<<l_primary_.grid, CULL_THREADS, l_primary_.smem, stream_>>>(
                    cams, &ctr_->count[l], cull_eye_, frame_eye_, sc_.n_cull, lists_camera_, cq_l, ctr_);
                else k_cull<R, CULL_PIVOT, CameraSource<R>, false><<<l_primary_.grid, CULL_THREADS, 0, stream_>>>(
                    cams, &ctr_->count[l], cull_eye_, frame_eye_, sc_.n_cull, lists_camera_, cq_l, ctr_);
                MARK(KC_EXACT);
                if (kRobustPrimary) {
                    // float32 spheres inline; then the double-precision pass: other kinds + the spheres the first pass deferred
                    if (counting) {
                        if (have_spheres) k_exact_closest<R, kRobustPrimary, true, EXACT_SPHERES><<<exact_grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
                        k_exact_closest<R, kRobustPrimary, true, EXACT_REST><<<exact_grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
                    } else {
                        if (have_spheres) k_exact_closest<R, kRobustPrimary, false, EXACT_SPHERES><<<exact_grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
                        k_exact_closest<R, kRobustPrimary, false, EXACT_REST><<<exact_grid, EXACT_THREADS, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_, cam);
                    }
                    totals_.launches += have_spheres ? 1 : 0;
                } else {
                    int erc = exact_closest(l, counting, cq_l);
                    if (erc) return erc;
                }
            } else {
                MARK(KC_CLOSEST);
                SpawnSource<R> spws{levels_[l], sc_.pickless != 0, (sort_spawned_ && sort_perm_) ? reinterpret_cast<const int*>(spawn_stage_) : nullptr};
                if (resident_) k_cull<R, CULL_GENERAL, SpawnSource<R>, true><<<l_closest_.grid, CULL_THREADS, l_closest_.smem, stream_>>>(
                    spws, &ctr_->count[l], cull_general_, frame_general_, sc_.n_cull, lists_spawn_, cq_l, ctr_);
                else k_cull<R, CULL_GENERAL, SpawnSource<R>, false><<<l_closest_.grid, CULL_THREADS, 0, stream_>>>(
                    spws, &ctr_->count[l], cull_general_, frame_general_, sc_.n_cull, lists_spawn_, cq_l, ctr_);
                MARK(KC_EXACT);
                int erc = exact_closest(l, counting, cq_l);
                if (erc) return erc;
            }
            if (sizeof(R) == 8) k_closest_order<R><<<exact_grid, 256, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_);
            else totals_.launches--;
            if (!sc_.pickless) k_closest_pick<R><<<exact_grid, 256, 0, stream_>>>(sc_, levels_[l], cq_l, ctr_);
            else totals_.launches--;
            totals_.launches += 3;
            {   // hit list of the level; misses of levels >= 1 are finished here
                MARK(KC_SORT);
                const int hl_grid = std::min((levels_[l].cap + HITLIST_RAYS - 1) / HITLIST_RAYS, sms_ * 8);
                int32_t* const hl_ids = (counting && l == 0) ? ids_ : nullptr;
                const Level<R>& hl_prev = l > 0 ? levels_[l - 1] : levels_[l];
                if (sc_.pickless == 1) k_hitlist<R, sizeof(R) == 4 ? 1 : 0><<<hl_grid, HITLIST_THREADS, 0, stream_>>>(sc_, levels_[l], hl_prev, ctr_, l, hl_ids, ty, map);
                else if (sc_.pickless) k_hitlist<R, sizeof(R) == 4 ? 2 : 0><<<hl_grid, HITLIST_THREADS, 0, stream_>>>(sc_, levels_[l], hl_prev, ctr_, l, hl_ids, ty, map);
                else k_hitlist<R, 0><<<hl_grid, HITLIST_THREADS, 0, stream_>>>(sc_, levels_[l], hl_prev, ctr_, l, hl_ids, ty, map);
                totals_.launches++;
            }
            const bool spawn = l + 1 < depth_eff_;
            if (two_streams && l > 0) for (int k = 0; k < nss; k++) CU_TRY(cudaStreamWaitEvent(stream_, ev_shadow_[k][l - 1], 0));     // k_shade refills the shadow queue
            if (primary_points_ && l == 0) MARK(KC_PIXELS);          // (class 6: the camera rays' hit points in double)
            if (primary_points_ && l == 0) {
                if (sc_.pickless == 1) k_primary_points<R, sizeof(R) == 4 ? 1 : 0><<<wide, 256, 0, stream_>>>(sc_, levels_[0], ctr_, cam);
                else if (sc_.pickless) k_primary_points<R, sizeof(R) == 4 ? 2 : 0><<<wide, 256, 0, stream_>>>(sc_, levels_[0], ctr_, cam);
                else k_primary_points<R, 0><<<wide, 256, 0, stream_>>>(sc_, levels_[0], ctr_, cam);
                totals_.launches++;
            }
            MARK(KC_SHADE);
            auto shade = sc_.pickless == 1 ? k_shade<R, sizeof(R) == 4 ? 1 : 0> : sc_.pickless ? k_shade<R, sizeof(R) == 4 ? 2 : 0> : k_shade<R, 0>;
            shade<<<std::min(cap_blocks, wide), SHADE_THREADS, 0, stream_>>>(
                sc_, levels_[l], spawn ? levels_[l + 1] : levels_[l], sq_l, occl_[l], ctr_, l, spawn ? 1 : 0,
                spawn ? ((sort_spawned_ && !sort_perm_) ? spawn_stage_ : levels_[l + 1].ray) : nullptr, (primary_points_ && l == 0) ? 1 : 0);
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
                CandQueue cqs = two_streams ? cq_shadow_[sk] : cq_;
                cqs.count = slot + 1 + li;
                int* const s_hist = sort_hist_sh_[sk]; int* const s_cursor = sort_cursor_sh_[sk];
                ShadowSource<R> ssrc{sq_l, li * sq_.per_light, sort_shadows_ ? sq_.perm : nullptr, sort_shift_, sort_nb_};
                if (sort_shadows_) {       // optional: the shadow rays of neighbouring pixels mostly share a class already
                    MARK(KC_SORT);
                    k_sort_hist<R><<<sms_ * 4, SORT_THREADS, 0, ss>>>(sq_l, li, sort_shift_, sort_nb_, s_hist);
                    k_sort_scan<<<1, 1024, 0, ss>>>(sort_nb_, s_hist, s_cursor);
                    k_sort_scatter<R><<<sms_ * 4, SORT_THREADS, 0, ss>>>(sq_l, li, sort_shift_, sort_nb_, s_cursor);
                    totals_.launches += 3;
                }
                MARK(KC_SHADOW);
                const BinLists& bls = lists_light_[li];
                if (axis) {
                    if (resident_) k_cull<R, CULL_AXIS, ShadowSource<R>, true><<<l_shadow_axis_.grid, CULL_THREADS, l_shadow_axis_.smem, ss>>>(
                        ssrc, &sq_l.count[li], cull_light_[li], frame_light_[li], sc_.n_cull, bls, cqs, ctr_);
                    else k_cull<R, CULL_AXIS, ShadowSource<R>, false><<<l_shadow_axis_.grid, CULL_THREADS, 0, ss>>>(
                        ssrc, &sq_l.count[li], cull_light_[li], frame_light_[li], sc_.n_cull, bls, cqs, ctr_);
                } else {
                    if (resident_) k_cull<R, CULL_PIVOT, ShadowSource<R>, true><<<l_shadow_pivot_.grid, CULL_THREADS, l_shadow_pivot_.smem, ss>>>(
                        ssrc, &sq_l.count[li], cull_light_[li], frame_light_[li], sc_.n_cull, bls, cqs, ctr_);
                    else k_cull<R, CULL_PIVOT, ShadowSource<R>, false><<<l_shadow_pivot_.grid, CULL_THREADS, 0, ss>>>(
                        ssrc, &sq_l.count[li], cull_light_[li], frame_light_[li], sc_.n_cull, bls, cqs, ctr_);
                }
                MARK(KC_EXACT);
                const int sbase = li * sq_.per_light;
                if (counting) {
                    if (have_spheres) k_exact_shadow<R, true, EXACT_SPHERES><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_l, sbase, cqs, ctr_, occl_[l]);
                    if (have_simple) k_exact_shadow<R, true, EXACT_SIMPLE><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_l, sbase, cqs, ctr_, occl_[l]);
                    if (have_csg) k_exact_shadow<R, true, EXACT_CSG><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_l, sbase, cqs, ctr_, occl_[l]);
                    k_shadow_finish<R, true><<<wide, 256, 0, ss>>>(sc_, sq_l, li, occl_[l], ctr_);
                } else {                                   // the exact kernels set the occlusion flags themselves
                    if (have_spheres) k_exact_shadow<R, false, EXACT_SPHERES><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_l, sbase, cqs, ctr_, occl_[l]);
                    if (have_simple) k_exact_shadow<R, false, EXACT_SIMPLE><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_l, sbase, cqs, ctr_, occl_[l]);
                    if (have_csg) k_exact_shadow<R, false, EXACT_CSG><<<exact_grid, EXACT_THREADS, 0, ss>>>(sc_, sq_l, sbase, cqs, ctr_, occl_[l]);
                    totals_.launches--;
                }
                totals_.launches += n_exact_flavours - 1;
                totals_.launches += 3;
            }
            if (two_streams) for (int k = 0; k < nss; k++) CU_TRY(cudaEventRecord(ev_shadow_[k][l], shadow_stream_[k]));
        }
        if (two_streams) for (int k = 0; k < nss; k++) CU_TRY(cudaStreamWaitEvent(stream_, ev_shadow_[k][depth_eff_ - 1], 0));   // the shadow streams are in order: every level is done
        for (int l = depth_eff_ - 1; l >= 0; l--) {
            const int cap_blocks = (levels_[l].cap + SHADE_THREADS - 1) / SHADE_THREADS;
            MARK(KC_RESOLVE);
            k_resolve<R><<<std::min(cap_blocks, wide), SHADE_THREADS, 0, stream_>>>(
                sc_, levels_[l], l > 0 ? levels_[l - 1] : levels_[l], occl_[l], ctr_, l, l + 1 < depth_eff_ ? 1 : 0);
            totals_.launches++;
        }
        {   // the level-0 colours are final: average the samples, store the pixels
            uchar4* const target = frame_target_ ? (uchar4*)frame_target_ : frame_;
            if (sc_.pickless) k_pixels<R, sizeof(R) == 4 ? 1 : 0><<<wide, 256, 0, stream_>>>(sc_, levels_[0], ctr_, target, ty, map);
            else k_pixels<R, 0><<<wide, 256, 0, stream_>>>(sc_, levels_[0], ctr_, target, ty, map);
            totals_.launches++;
        }
    }
    MARK(-1);
#undef MARK
    // the frame's counters follow the last kernel into page-locked memory: the caller's one stream synchronisation covers them
    // (a blocking cudaMemcpy after it was a second round trip per frame)
    if (!ctr_host_) CU_TRY(cudaHostAlloc((void**)&ctr_host_, sizeof(Counters), cudaHostAllocDefault));
    CU_TRY(cudaMemcpyAsync(ctr_host_, ctr_, sizeof(Counters), cudaMemcpyDeviceToHost, stream_));
    ctr_host_valid_ = true;
    CU_TRY(cudaGetLastError());
    return GMLRT_OK;
}

}  // namespace gmlrt
