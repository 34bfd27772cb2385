// render_impl.cuh - the kernels of the wavefront ray tracer, templated on the arithmetic type (the per-device Renderer that
// launches them is render_host.cuh, included at the end).
//
// Replaces Render / traceRay / closestHit / computeLighting / inShadow and the per-primitive
// Intersect / ComputeSurfaceProps of /root/reference/raytracer.go (line references at each
// function).  The reference is a recursive tracer run by 8 goroutines; this is a wavefront, per recursion level:
//
//   cull       one thread per ray walks the candidate list of the ray's CLASS (built once per upload, k_list_build)
//              and tests the listed bounding spheres from 16-byte records staged in shared memory by one TMA bulk
//              copy; at level 0 the same kernel makes the camera rays (PCG jitter by jump-ahead)   (Render :632-652)
//   exact      the reference's own <Type>.Intersect on the survivors, folded per ray                 (closestHit :469-483)
//   hit list   the rays that hit, compacted (__ballot_sync / __popc); misses are finished here       (traceRay :493-497)
//   shade      surface function (bytecode), material, shadow + child rays                            (traceRay :498-556)
//   shadow     cull + exact again for the shadow rays of each light, any-hit                         (inShadow :411-429)
//   resolve    lighting + clamp per recursion level, deepest level first                             (:372-401, :506-561)
//   pixels     mean of 4 samples -> RGBA8                                                            (:656, vec.go:104-107)
//
// Rays of one recursion level live in queues; children are appended with warp-aggregated
// (__ballot_sync/__popc) atomics, so every kernel runs on dense queues.  Colour is
// clamped at every level in the reference, hence the per-level records and the
// back-to-front resolve instead of a throughput-weighted accumulation.  DESIGN.md section 4 has the whole story.
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


#include "render_host.cuh"      // Renderer<R>: upload, class lists, tile plan, the launch sequence of a frame

}  // namespace gmlrt
