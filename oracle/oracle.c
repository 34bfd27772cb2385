/*
 * oracle/oracle.c - TEST INFRASTRUCTURE: CPU restatement of the reference's `render` path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this.  The product (libgmlrt.so) never links, calls or falls back to it.
 *
 * What it restates (float64, reference operation order, no FMA contraction):
 *   Render                      /root/reference/raytracer.go:589-682
 *   traceRay / closestHit       raytracer.go:469-562
 *   computeLighting / inShadow  raytracer.go:372-429
 *   refract / fresnel           raytracer.go:438-467
 *   Sphere/Plane/Cube/Cylinder  raytracer.go:58-370
 *   scene conversion            raytracer.go:756-830 (BFS union flattening, face planes)
 *   Vec3 / Mat4 helpers         internal/prim/vec.go:23-116, 298-313
 *   EvalSurfaceFn               internal/gml/evaluator.go:672-727 (via the bytecode the host
 *                               compiler emits; the bytecode is cross-checked against the
 *                               Python restatement of the Go evaluator in tests/)
 * plus Go stdlib pieces in gomath.h.  Input is the scene blob of include/gmlrt.h.
 *
 * Parity status: PINNED - reproduces all seven golden PNGs of the reference
 * (testdata/goldens) byte for byte (tests/test_oracle_goldens.py), and the nine
 * cylinder known-answer cases of cylinder_test.go.  Extensions (cone, intersect,
 * difference, directional and spot lights) have no reference behaviour: parity unpinned.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC -pthread (oracle/Makefile).
 */
#include <pthread.h>
#include <stdatomic.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/gmlrt.h"
#include "gomath.h"

typedef struct { double x, y, z; } v3;

static inline v3 V(double x, double y, double z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vmul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 vscale(v3 a, double s) { return V(a.x * s, a.y * s, a.z * s); }
static inline double vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline double vlen(v3 a) { return sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
static inline v3 vnorm(v3 a) { double m = sqrt(a.x * a.x + a.y * a.y + a.z * a.z); return V(a.x / m, a.y / m, a.z / m); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline double clamp01(double x) { return go_min(go_max(x, 0), 1); }
static inline v3 vclamp(v3 a) { return V(clamp01(a.x), clamp01(a.y), clamp01(a.z)); }
static inline v3 vlerp(v3 a, v3 b, double t) { return V(a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t, a.z + (b.z - a.z) * t); }

/* rows 0..2 of a 4x4, row-major: m[4*r + c] */
static inline v3 mul_point(const double* m, v3 v) {
    return V(m[0] * v.x + m[1] * v.y + m[2] * v.z + m[3],
             m[4] * v.x + m[5] * v.y + m[6] * v.z + m[7],
             m[8] * v.x + m[9] * v.y + m[10] * v.z + m[11]);
}
static inline v3 mul_dir(const double* m, v3 v) {
    return V(m[0] * v.x + m[1] * v.y + m[2] * v.z,
             m[4] * v.x + m[5] * v.y + m[6] * v.z,
             m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
/* (M^T) applied as a direction: what NormalMat.MulDir does (raytracer.go:367, 767, 790) */
static inline v3 mul_dir_t(const double* m, v3 v) {
    return V(m[0] * v.x + m[4] * v.y + m[8] * v.z,
             m[1] * v.x + m[5] * v.y + m[9] * v.z,
             m[2] * v.x + m[6] * v.y + m[10] * v.z);
}

typedef struct { v3 o, d; } ray_t;

typedef struct {
    int obj;      /* index in scene order, -1 = none */
    double t;
    v3 p_obj;
    int face;
    int leaf;     /* for CSG objects: the leaf whose surface was hit */
    int flip;     /* CSG: normal flipped (surface of a subtracted solid) */
} hit_t;

typedef struct {
    int kind;             /* GMLRT_NODE_* leaf kind, or -1 for a CSG object */
    uint32_t surface;
    double o2w[12], w2o[12];
    v3 face_normal_world[6];   /* plane: [0]; cube: per face (raytracer.go:764-774, 799-803) */
    uint32_t csg_root;         /* CSG objects: node index of the subtree root */
    double flops;              /* credited flops of one test of this object */
} object_t;

typedef struct {
    const uint8_t* blob;
    const gmlrt_blob_header* h;
    const gmlrt_node* nodes;
    const gmlrt_light* lights;
    const gmlrt_surface* surfaces;
    const double* fpool;
    const int32_t* itab;
    const uint32_t* code;
    object_t* objs;            /* scan order = the reference's []SceneObject */
    int n_objs;
    object_t* leaves;          /* per node index: leaf data for nodes inside CSG subtrees */
    int width, height, depth;
    double vw, vh;
    v3 ambient, bg0, bg1;
    double flops_all;          /* credited flops of testing every object once */
} scene_t;

typedef struct {
    uint64_t rays_primary, rays_reflect, rays_refract, rays_shadow, shadow_occluded, leaf_tests, rays_hit;
    double flops;
    int surface_error;
} counters_t;

static const double FLOPS_BY_KIND[5] = {56, 81, 74, 70, 47}; /* sphere cube cylinder cone plane: SURVEY.md 8(d) */
static inline double kind_flops(int kind) { return (kind >= 16 && kind <= 20) ? FLOPS_BY_KIND[kind - 16] : 0; }

static const v3 CUBE_P[6] = {{0, 0, 0}, {0, 0, 1}, {0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 0}};
static const v3 CUBE_N[6] = {{0, 0, -1}, {0, 0, 1}, {-1, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, -1, 0}};

static void init_leaf(object_t* o, const gmlrt_node* n) {
    o->kind = (int)n->kind;
    o->surface = n->surface;
    memcpy(o->o2w, n->o2w, sizeof o->o2w);
    memcpy(o->w2o, n->w2o, sizeof o->w2o);
    if (o->kind == GMLRT_NODE_PLANE) {
        o->face_normal_world[0] = vnorm(mul_dir_t(o->w2o, V(0, 1, 0)));
    } else if (o->kind == GMLRT_NODE_CUBE) {
        for (int f = 0; f < 6; f++) o->face_normal_world[f] = vnorm(mul_dir_t(o->w2o, CUBE_N[f]));
    }
}

static int is_leaf_kind(uint32_t k) { return k >= 16 && k <= 20; }

/* a CSG object is credited with its member leaves' table flops, +2 per quadric because both roots are needed */
static int mark_csg_leaves(scene_t* s, uint32_t ni, int depth, double* flops) {
    if (ni >= s->h->n_nodes || depth > 4096) return -1;
    const gmlrt_node* n = &s->nodes[ni];
    if (is_leaf_kind(n->kind)) {
        init_leaf(&s->leaves[ni], n);
        *flops += kind_flops((int)n->kind) + ((n->kind == GMLRT_NODE_SPHERE || n->kind == GMLRT_NODE_CYLINDER || n->kind == GMLRT_NODE_CONE) ? 2 : 0);
        return 0;
    }
    if (n->kind > 2) return -1;
    if (mark_csg_leaves(s, n->a, depth + 1, flops)) return -1;
    return mark_csg_leaves(s, n->b, depth + 1, flops);
}

/* returns 0 or a GMLRT_E_* code */
static int scene_init(scene_t* s, const void* blob, size_t nbytes) {
    memset(s, 0, sizeof *s);
    if (!blob || nbytes < sizeof(gmlrt_blob_header)) return GMLRT_E_BAD_BLOB;
    const gmlrt_blob_header* h = (const gmlrt_blob_header*)blob;
    if (h->magic != GMLRT_BLOB_MAGIC || h->version != GMLRT_BLOB_VERSION || h->total_bytes != nbytes) return GMLRT_E_BAD_BLOB;
    if (h->off_nodes + (uint64_t)h->n_nodes * sizeof(gmlrt_node) > nbytes ||
        h->off_lights + (uint64_t)h->n_lights * sizeof(gmlrt_light) > nbytes ||
        h->off_surfaces + (uint64_t)h->n_surfaces * sizeof(gmlrt_surface) > nbytes ||
        h->off_fpool + (uint64_t)h->n_fpool * 8 > nbytes || h->off_itab + (uint64_t)h->n_itab * 4 > nbytes ||
        h->off_code + (uint64_t)h->n_code * 4 > nbytes || h->root >= h->n_nodes || h->width <= 0 || h->height <= 0)
        return GMLRT_E_BAD_BLOB;
    s->blob = (const uint8_t*)blob; s->h = h;
    s->nodes = (const gmlrt_node*)(s->blob + h->off_nodes);
    s->lights = (const gmlrt_light*)(s->blob + h->off_lights);
    s->surfaces = (const gmlrt_surface*)(s->blob + h->off_surfaces);
    s->fpool = (const double*)(s->blob + h->off_fpool);
    s->itab = (const int32_t*)(s->blob + h->off_itab);
    s->code = (const uint32_t*)(s->blob + h->off_code);
    s->width = h->width; s->height = h->height;
    s->depth = h->depth > 0 ? h->depth : 3;                       /* raytracer.go:592-595 */
    double fov = h->fov > 0.0 ? h->fov : 90.0;                    /* raytracer.go:597-600 */
    double fov_rad = fov * GO_PI / 180.0;
    s->vw = 2.0 / go_tan(fov_rad / 2.0);                          /* raytracer.go:601-602 */
    s->vh = s->vw * ((double)s->height / (double)s->width);
    s->ambient = V(h->ambient[0], h->ambient[1], h->ambient[2]);
    s->bg0 = V(h->bg_start[0], h->bg_start[1], h->bg_start[2]);
    s->bg1 = V(h->bg_end[0], h->bg_end[1], h->bg_end[2]);
    s->objs = (object_t*)calloc(h->n_nodes, sizeof(object_t));
    s->leaves = (object_t*)calloc(h->n_nodes, sizeof(object_t));
    /* BFS over unions (raytracer.go:776-828): queue of node indices */
    uint32_t* q = (uint32_t*)malloc(sizeof(uint32_t) * (h->n_nodes + 1));
    uint32_t qh = 0, qt = 0;
    q[qt++] = h->root;
    int rc = 0;
    while (qh < qt && !rc) {
        uint32_t ni = q[qh++];
        const gmlrt_node* n = &s->nodes[ni];
        if (n->kind == GMLRT_NODE_UNION) {
            if (n->a >= h->n_nodes || n->b >= h->n_nodes || qt + 2 > h->n_nodes + 1) { rc = GMLRT_E_BAD_BLOB; break; }
            q[qt++] = n->a; q[qt++] = n->b;
        } else if (is_leaf_kind(n->kind)) {
            init_leaf(&s->objs[s->n_objs], n);
            s->objs[s->n_objs].flops = kind_flops((int)n->kind);
            s->n_objs++;
        } else if (n->kind == GMLRT_NODE_INTERSECT || n->kind == GMLRT_NODE_DIFFERENCE) {
            /* extension: the reference rejects these (raytracer.go:825-826) */
            object_t* o = &s->objs[s->n_objs++];
            o->kind = -1; o->csg_root = ni;
            if (mark_csg_leaves(s, ni, 0, &o->flops)) rc = GMLRT_E_BAD_BLOB;
        } else {
            rc = GMLRT_E_UNSUPPORTED;
        }
    }
    free(q);
    for (int i = 0; i < s->n_objs; i++) s->flops_all += s->objs[i].flops;
    return rc;
}

static void scene_free(scene_t* s) { free(s->objs); free(s->leaves); }

/* ------------------------------------------------------------------ primitives */

/* Plane.Intersect on an object-space ray (raytracer.go:167-179); n, d = plane normal and D */
static inline int plane_isect(v3 n, double D, ray_t r, double* t, v3* p) {
    double denom = vdot(n, r.d);
    if (fabs(denom) < 1e-6) return 0;
    double tt = (-D - vdot(n, r.o)) / denom;
    if (tt <= 0.0) return 0;
    *t = tt;
    *p = vadd(r.o, vscale(r.d, tt));
    return 1;
}

static __thread unsigned long long tl_chaos;   /* cube-face rejections decided by rounding alone (see the cube case below) */

/* First-hit rule of each primitive, exactly as <Type>.Intersect.  Returns 1 on hit. */
static int leaf_intersect(const object_t* o, ray_t wray, double* t_out, v3* p_out, int* face_out) {
    ray_t r;
    r.o = mul_point(o->w2o, wray.o);      /* rayToObjectSpace raytracer.go:51-56 */
    r.d = mul_dir(o->w2o, wray.d);
    switch (o->kind) {
    case GMLRT_NODE_SPHERE: {              /* raytracer.go:78-103 */
        double a = vdot(r.d, r.d);
        double halfB = vdot(r.o, r.d);
        double c = vdot(r.o, r.o) - 1.0;
        double disc = halfB * halfB - a * c;
        if (disc < 0.0) return 0;
        double sq = sqrt(disc);
        double t0 = (-halfB - sq) / a;
        if (t0 > 0.0) {
            *t_out = t0; *p_out = vadd(r.o, vscale(r.d, t0)); *face_out = 0;
            return 1;
        }
        return 0;
    }
    case GMLRT_NODE_PLANE: {               /* GML plane: point 0, normal +y (evaluator.go:816-822) */
        v3 n = V(0, 1, 0);
        double D = -vdot(n, V(0, 0, 0));
        *face_out = 0;
        return plane_isect(n, D, r, t_out, p_out);
    }
    case GMLRT_NODE_CUBE: {                /* raytracer.go:214-240 */
        int found = 0;
        for (int f = 0; f < 6; f++) {
            double t; v3 p;
            double D = -vdot(CUBE_N[f], CUBE_P[f]);
            if (!plane_isect(CUBE_N[f], D, r, &t, &p) || t < 0.0) continue;
            if (p.x < 0 || p.x > 1 || p.y < 0 || p.y > 1 || p.z < 0 || p.z > 1) {
                /* the bounds test includes the face's own axis, where p is 0 or 1 only up to rounding, so a face
                 * is randomly rejected when o + d t lands at -1e-17.  Count the rejections that rounding alone decided
                 * (the point is inside the face to 1e-9): a pixel that saw one is rounding chaos in the reference itself. */
                const double e = 1e-9;
                if (p.x >= -e && p.x <= 1 + e && p.y >= -e && p.y <= 1 + e && p.z >= -e && p.z <= 1 + e) tl_chaos++;
                continue;
            }
            if (!found || t < *t_out) { found = 1; *t_out = t; *p_out = p; *face_out = f; }
        }
        return found;
    }
    case GMLRT_NODE_CYLINDER: {            /* raytracer.go:279-337 */
        double best = INFINITY; int bf = -1; v3 bp = V(0, 0, 0);
#define CONSIDER(tt, ff, pp) do { if ((tt) > 0.0 && (tt) < best) { best = (tt); bf = (ff); bp = (pp); } } while (0)
        double a = r.d.x * r.d.x + r.d.z * r.d.z;
        if (a > 1e-12) {
            double halfB = r.o.x * r.d.x + r.o.z * r.d.z;
            double c0 = r.o.x * r.o.x + r.o.z * r.o.z - 1.0;
            double disc = halfB * halfB - a * c0;
            if (disc >= 0.0) {
                double sq = sqrt(disc);
                double ts[2] = {(-halfB - sq) / a, (-halfB + sq) / a};
                for (int k = 0; k < 2; k++) {
                    v3 p = vadd(r.o, vscale(r.d, ts[k]));
                    if (p.y >= 0.0 && p.y <= 1.0) CONSIDER(ts[k], 0, p);
                }
            }
        }
        if (fabs(r.d.y) > 1e-12) {
            double tTop = (1.0 - r.o.y) / r.d.y;
            v3 pTop = vadd(r.o, vscale(r.d, tTop));
            if (pTop.x * pTop.x + pTop.z * pTop.z <= 1.0) CONSIDER(tTop, 1, pTop);
            double tBot = -r.o.y / r.d.y;
            v3 pBot = vadd(r.o, vscale(r.d, tBot));
            if (pBot.x * pBot.x + pBot.z * pBot.z <= 1.0) CONSIDER(tBot, 2, pBot);
        }
        if (bf == -1) return 0;
        *t_out = best; *p_out = bp; *face_out = bf;
        return 1;
    }
    case GMLRT_NODE_CONE: {                /* EXTENSION (ICFP text): x^2+z^2-y^2<=0, 0<=y<=1; same
                                              candidate rule as the cylinder: side-, side+, base */
        double best = INFINITY; int bf = -1; v3 bp = V(0, 0, 0);
        double a = r.d.x * r.d.x + r.d.z * r.d.z - r.d.y * r.d.y;
        double halfB = r.o.x * r.d.x + r.o.z * r.d.z - r.o.y * r.d.y;
        double c0 = r.o.x * r.o.x + r.o.z * r.o.z - r.o.y * r.o.y;
        if (fabs(a) > 1e-12) {
            double disc = halfB * halfB - a * c0;
            if (disc >= 0.0) {
                double sq = sqrt(disc);
                double ts[2] = {(-halfB - sq) / a, (-halfB + sq) / a};
                for (int k = 0; k < 2; k++) {
                    v3 p = vadd(r.o, vscale(r.d, ts[k]));
                    if (p.y >= 0.0 && p.y <= 1.0) CONSIDER(ts[k], 0, p);
                }
            }
        } else if (fabs(halfB) > 1e-12) {
            double t = -c0 / (2.0 * halfB);
            v3 p = vadd(r.o, vscale(r.d, t));
            if (p.y >= 0.0 && p.y <= 1.0) CONSIDER(t, 0, p);
        }
        if (fabs(r.d.y) > 1e-12) {
            double tTop = (1.0 - r.o.y) / r.d.y;
            v3 pTop = vadd(r.o, vscale(r.d, tTop));
            if (pTop.x * pTop.x + pTop.z * pTop.z <= 1.0) CONSIDER(tTop, 1, pTop);
        }
#undef CONSIDER
        if (bf == -1) return 0;
        *t_out = best; *p_out = bp; *face_out = bf;
        return 1;
    }
    }
    return 0;
}

/* ------------------------------------------------------------------ surface bytecode (f64, int64) */

typedef union { double f; int64_t i; } cell_t;

/* returns 0 ok, 1 fault */
static int run_surface(const scene_t* s, const gmlrt_surface* sf, int face, double u, double v, double* mat) {
    if (sf->kind == GMLRT_SURFACE_CONST) { memcpy(mat, sf->material, 10 * sizeof(double)); return 0; }
    cell_t r[GMLRT_VM_MAX_REGS];
    const uint32_t* code = s->code + sf->code_off;
    uint32_t pc = 0, n = sf->code_len;
    const double deg = GO_PI / 180.0;
    r[0].i = face; r[1].f = u; r[2].f = v;
    while (pc < n) {
        uint32_t w = code[pc++];
        uint32_t op = w & 0xFF, d = (w >> 8) & 0xFF, a = (w >> 16) & 0xFF, b = w >> 24;
        if (d >= GMLRT_VM_MAX_REGS || a >= GMLRT_VM_MAX_REGS || (b >= GMLRT_VM_MAX_REGS && op != GMLRT_OP_LDK)) return 1;
        switch (op) {
        case GMLRT_OP_RET:
            if (a + 10 > GMLRT_VM_MAX_REGS) return 1;
            for (int k = 0; k < 10; k++) mat[k] = r[a + k].f;
            return 0;
        case GMLRT_OP_LDK: { uint32_t k = a | (b << 8); if (k >= s->h->n_fpool) return 1; r[d].f = s->fpool[k]; break; }
        case GMLRT_OP_LDI: r[d].i = (int32_t)code[pc++]; break;
        case GMLRT_OP_MOV: r[d] = r[a]; break;
        case GMLRT_OP_ADDF: r[d].f = r[a].f + r[b].f; break;
        case GMLRT_OP_SUBF: r[d].f = r[a].f - r[b].f; break;
        case GMLRT_OP_MULF: r[d].f = r[a].f * r[b].f; break;
        case GMLRT_OP_DIVF: r[d].f = r[a].f / r[b].f; break;
        case GMLRT_OP_NEGF: r[d].f = -r[a].f; break;
        case GMLRT_OP_ADDI: r[d].i = (int64_t)((uint64_t)r[a].i + (uint64_t)r[b].i); break;
        case GMLRT_OP_SUBI: r[d].i = (int64_t)((uint64_t)r[a].i - (uint64_t)r[b].i); break;
        case GMLRT_OP_MULI: r[d].i = (int64_t)((uint64_t)r[a].i * (uint64_t)r[b].i); break;
        case GMLRT_OP_DIVI: if (r[b].i == 0) return 1; r[d].i = (r[b].i == -1) ? (int64_t)(0 - (uint64_t)r[a].i) : r[a].i / r[b].i; break;
        case GMLRT_OP_MODI: if (r[b].i == 0) return 1; r[d].i = (r[b].i == -1) ? 0 : r[a].i % r[b].i; break;
        case GMLRT_OP_NEGI: r[d].i = (int64_t)(0 - (uint64_t)r[a].i); break;
        case GMLRT_OP_EQF: r[d].i = r[a].f == r[b].f; break;
        case GMLRT_OP_LESSF: r[d].i = r[a].f < r[b].f; break;
        case GMLRT_OP_EQI: r[d].i = r[a].i == r[b].i; break;
        case GMLRT_OP_LESSI: r[d].i = r[a].i < r[b].i; break;
        case GMLRT_OP_SIN: r[d].f = go_sin(deg * r[a].f); break;     /* evaluator.go:931-933 */
        case GMLRT_OP_COS: r[d].f = go_cos(deg * r[a].f); break;
        case GMLRT_OP_SQRT: r[d].f = sqrt(r[a].f); break;
        case GMLRT_OP_CLAMPF: { double x = r[a].f; if (x < 0) x = 0; else if (x > 1) x = 1; r[d].f = x; break; }
        case GMLRT_OP_FLOOR: r[d].i = (int64_t)floor(r[a].f); break;
        case GMLRT_OP_FRAC: r[d].f = r[a].f - (double)(int64_t)r[a].f; break;
        case GMLRT_OP_REAL: r[d].f = (double)r[a].i; break;
        case GMLRT_OP_ACOS: r[d].f = go_acos(r[a].f) / deg; break;
        case GMLRT_OP_ASIN: r[d].f = go_asin(r[a].f) / deg; break;
        case GMLRT_OP_LDARR: r[d].i = code[pc]; r[d + 1].i = code[pc + 1]; pc += 2; break;
        case GMLRT_OP_LEN: r[d].i = r[a + 1].i; break;
        case GMLRT_OP_GETF: case GMLRT_OP_GETP: case GMLRT_OP_GETI: case GMLRT_OP_GETA: {
            int64_t off = r[a].i, len = r[a + 1].i, idx = r[b].i;
            if (idx < 0 || idx >= len) return 1;                       /* evaluator.go:989-995 */
            if (op == GMLRT_OP_GETF) { if (off + idx >= s->h->n_fpool) return 1; r[d].f = s->fpool[off + idx]; }
            else if (op == GMLRT_OP_GETP) {
                if (off + 3 * idx + 2 >= s->h->n_fpool) return 1;
                for (int k = 0; k < 3; k++) r[d + k].f = s->fpool[off + 3 * idx + k];
            } else if (op == GMLRT_OP_GETI) { if (off + idx >= s->h->n_itab) return 1; r[d].i = s->itab[off + idx]; }
            else { if (off + 2 * idx + 1 >= s->h->n_itab) return 1; r[d].i = s->itab[off + 2 * idx]; r[d + 1].i = s->itab[off + 2 * idx + 1]; }
            break;
        }
        case GMLRT_OP_JMP: pc += 1 + code[pc]; break;
        case GMLRT_OP_JMPF: if (r[a].i == 0) pc += 1 + code[pc]; else pc += 1; break;
        default: return 1; /* GMLRT_OP_ERR and unknown */
        }
    }
    return 1;
}

/* ------------------------------------------------------------------ CSG objects (EXTENSION, parity unpinned) */
static int csg_intersect(const scene_t* s, const object_t* o, ray_t wray, hit_t* out);   /* defined below */

static int object_intersect(const scene_t* s, int oi, ray_t wray, hit_t* out) {
    const object_t* o = &s->objs[oi];
    if (o->kind < 0) {
        if (!csg_intersect(s, o, wray, out)) return 0;
        out->obj = oi;
        return 1;
    }
    if (!leaf_intersect(o, wray, &out->t, &out->p_obj, &out->face)) return 0;
    out->obj = oi; out->leaf = -1; out->flip = 0;
    return 1;
}

/* ------------------------------------------------------------------ surface properties */
typedef struct { v3 p_world, n_world; double mat[10]; } hitex_t;

/* <Type>.ComputeSurfaceProps, raytracer.go:106-150, 182-205, 242-260, 339-370.  Returns 0 ok. */
static int surface_props(const scene_t* s, const object_t* o, const hit_t* h, hitex_t* ex) {
    v3 p = h->p_obj;
    int face = h->face;
    double u, v;
    const gmlrt_surface* sf = &s->surfaces[o->surface];
    switch (o->kind) {
    case GMLRT_NODE_SPHERE:
        face = 0;
        if (sf->kind == GMLRT_SURFACE_CONST) { u = v = 0; }
        else {
            if (fabs(p.y) > 1) return 1;                              /* raytracer.go:142-144 */
            v = (p.y + 1.0) / 2.0;
            u = go_acos(p.z / sqrt(1.0 - p.y * p.y)) / (2.0 * GO_PI);
        }
        ex->n_world = p;                                              /* raytracer.go:119 (sic) */
        break;
    case GMLRT_NODE_PLANE:
        u = p.x; v = p.z; face = 0;
        ex->n_world = o->face_normal_world[0];
        break;
    case GMLRT_NODE_CUBE:
        if (face < 0 || face >= 6) return 1;
        u = p.x; v = p.z;                                             /* every face: raytracer.go:201-202 */
        ex->n_world = o->face_normal_world[face];
        break;
    case GMLRT_NODE_CYLINDER: {
        v3 n;
        if (face == 0) { n = V(p.x, 0, p.z); u = (go_atan2(p.x, p.z) + GO_PI) / (2.0 * GO_PI); v = p.y; }
        else if (face == 1) { n = V(0, 1, 0); u = p.x; v = p.z; }
        else if (face == 2) { n = V(0, -1, 0); u = p.x; v = p.z; }
        else return 1;
        ex->n_world = vnorm(mul_dir_t(o->w2o, n));
        break;
    }
    case GMLRT_NODE_CONE: {                                           /* EXTENSION */
        v3 n;
        if (face == 0) { n = V(p.x, -p.y, p.z); u = (go_atan2(p.x, p.z) + GO_PI) / (2.0 * GO_PI); v = p.y; }
        else if (face == 1) { n = V(0, 1, 0); u = p.x; v = p.z; }
        else return 1;
        ex->n_world = vnorm(mul_dir_t(o->w2o, n));
        break;
    }
    default: return 1;
    }
    if (run_surface(s, sf, face, u, v, ex->mat)) return 1;
    ex->p_world = mul_point(o->o2w, p);
    if (h->flip) ex->n_world = vneg(ex->n_world);
    return 0;
}

/* ------------------------------------------------------------------ lighting */

/* inShadow raytracer.go:411-429.  dist may be +inf (directional light). */
static int in_shadow(const scene_t* s, const hit_t* hit, const hitex_t* ex, v3 light_dir, double dist, ray_t in_ray,
                     counters_t* c) {
    ray_t sr;
    sr.o = vadd(ex->p_world, vscale(ex->n_world, 1e-4));
    sr.d = light_dir;
    double in_len = vlen(in_ray.d);
    c->rays_shadow++;
    for (int i = 0; i < s->n_objs; i++) {
        if (i == hit->obj && s->objs[i].kind >= 0) continue;   /* CSG objects may shadow themselves */
        c->leaf_tests++; c->flops += s->objs[i].flops;
        hit_t sh;
        if (!object_intersect(s, i, sr, &sh)) continue;
        /* a decision inside float32's reach of a tie (an occluder AT the light's distance: dice.gml puts its light on the
         * mirror's plane) is rounding chaos for any other arithmetic: flagged like the cube faces below */
        if (dist < INFINITY && fabs(sh.t * in_len - dist) <= 1e-5 * (dist + 1.0)) tl_chaos++;
        if (sh.t * in_len < dist) { c->shadow_occluded++; return 1; }
    }
    return 0;
}

/* computeLighting raytracer.go:372-401 */
static v3 compute_lighting(const scene_t* s, const hit_t* hit, const hitex_t* ex, ray_t ray, counters_t* c) {
    v3 Vv = vneg(ray.d);
    double kd = ex->mat[7], ks = ex->mat[8], nexp = ex->mat[9];
    v3 result = vscale(s->ambient, kd);
    for (uint32_t li = 0; li < s->h->n_lights; li++) {
        const gmlrt_light* L = &s->lights[li];
        v3 lcol = V(L->color[0], L->color[1], L->color[2]);
        v3 ldir; double dist;
        if (L->kind == GMLRT_LIGHT_DIRECTIONAL) {          /* EXTENSION */
            ldir = vnorm(vneg(V(L->pos[0], L->pos[1], L->pos[2])));
            dist = INFINITY;
        } else {
            v3 l2h = vsub(V(L->pos[0], L->pos[1], L->pos[2]), ex->p_world);
            dist = vlen(l2h);
            ldir = vnorm(l2h);
            if (L->kind == GMLRT_LIGHT_SPOT) {             /* EXTENSION */
                v3 axis = vnorm(vsub(V(L->at[0], L->at[1], L->at[2]), V(L->pos[0], L->pos[1], L->pos[2])));
                double cosang = vdot(axis, vneg(ldir));
                if (cosang < go_cos(L->cutoff * GO_PI / 180.0)) continue;
                lcol = vscale(lcol, go_pow(cosang, L->exponent));
            }
        }
        if (in_shadow(s, hit, ex, ldir, dist, ray, c)) continue;
        double nDotL = go_max(0, vdot(ex->n_world, ldir));
        v3 diffuse = vscale(lcol, nDotL * kd);
        v3 H = vnorm(vadd(Vv, ldir));
        double spec = go_max(0, vdot(ex->n_world, H));
        v3 specular = vscale(lcol, ks * go_pow(spec, nexp));
        result = vadd(vadd(result, diffuse), specular);
    }
    return result;
}

static v3 refract_dir(v3 incident, v3 normal, double n1, double n2) {     /* raytracer.go:438-450 */
    double ratio = n1 / n2;
    double cosI = -vdot(normal, incident);
    double sinT2 = ratio * ratio * (1.0 - cosI * cosI);
    if (sinT2 > 1.0) return V(0, 0, 0);
    double cosT = sqrt(1.0 - sinT2);
    return vadd(vscale(incident, ratio), vscale(normal, ratio * cosI - cosT));
}

static double fresnel(v3 normal, v3 incident, double ior) {                 /* raytracer.go:456-467 */
    double cosi = vdot(incident, normal) / (vlen(incident) * vlen(normal));
    double r0 = (1.0 - ior) / (1.0 + ior);
    r0 = r0 * r0;
    double cost = fabs(cosi);
    return r0 + (1 - r0) * go_pow(1 - cost, 5);
}

/* traceRay raytracer.go:487-562.  kind: 0 primary, 1 reflect, 2 refract (for the counters) */
static v3 trace_ray(const scene_t* s, ray_t ray, int depth, int kind, counters_t* c, int32_t* primary_id) {
    if (depth <= 0) return V(0, 0, 0);
    if (kind == 0) c->rays_primary++; else if (kind == 1) c->rays_reflect++; else c->rays_refract++;
    c->leaf_tests += (uint64_t)s->n_objs; c->flops += s->flops_all;
    hit_t best; best.obj = -1;
    for (int i = 0; i < s->n_objs; i++) {                                   /* closestHit :469-483 */
        hit_t h;
        if (!object_intersect(s, i, ray, &h)) continue;
        if (best.obj < 0 || h.t < best.t) best = h;
    }
    if (primary_id) *primary_id = best.obj;
    if (best.obj < 0) {
        double t = 0.5 * (ray.d.y + 1.0);
        return vlerp(s->bg0, s->bg1, t);
    }
    c->rays_hit++;
    const object_t* o = &s->objs[best.obj];
    if (o->kind < 0) o = &s->leaves[best.leaf];
    hitex_t ex;
    if (surface_props(s, o, &best, &ex)) { c->surface_error = 1; return V(0, 0, 0); }
    v3 lighting = compute_lighting(s, &best, &ex, ray, c);
    v3 color = V(ex.mat[0], ex.mat[1], ex.mat[2]);
    double refl = ex.mat[3], fuzz = ex.mat[4], transp = ex.mat[5], ior = ex.mat[6];
    if (refl == 0 && transp == 0) return vclamp(vmul(lighting, color));
    v3 reflected = V(0, 0, 0);
    if (refl > 0) {
        v3 rd = vsub(ray.d, vscale(ex.n_world, 2.0 * vdot(ray.d, ex.n_world)));
        if (fuzz >= 0) rd = vadd(rd, V(fuzz * go_cos(fuzz) * go_cos(fuzz), fuzz * go_sin(fuzz) * go_sin(fuzz), 0));
        ray_t rr; rr.o = vadd(ex.p_world, vscale(ex.n_world, 1e-4)); rr.d = vnorm(rd);
        reflected = trace_ray(s, rr, depth - 1, 1, c, NULL);
    }
    v3 refracted = V(0, 0, 0);
    if (transp > 0) {
        double n1 = 1.0, n2 = ior;
        v3 normal = ex.n_world;
        if (vdot(ray.d, normal) > 0.0) { double t = n1; n1 = n2; n2 = t; normal = vscale(normal, -1.0); }
        v3 rd = refract_dir(ray.d, normal, n1, n2);
        if (!(rd.x == 0.0 && rd.y == 0.0 && rd.z == 0.0)) {
            ray_t rr; rr.o = vsub(ex.p_world, vscale(normal, 1e-4)); rr.d = rd;
            refracted = trace_ray(s, rr, depth - 1, 2, c, NULL);
        }
    }
    if (transp == 0) return vclamp(vmul(vadd(lighting, vscale(reflected, refl)), color));
    double kr = fresnel(ex.n_world, ray.d, ior);
    return vclamp(vmul(vadd(vscale(lighting, 1.0 - transp), vadd(vscale(reflected, kr), vscale(refracted, 1.0 - kr))), color));
}

/* ------------------------------------------------------------------ CSG span lists (EXTENSION)
 * A solid along a ray is a sorted list of disjoint parameter intervals; every finite
 * bound remembers which leaf surface it lies on.  intersect / difference / union are
 * interval algebra (ICFP-2000 task text, restated in SURVEY.md 8c notes).            */
#define MAXSPANS 16
typedef struct { double t; int leaf, face, flip; } bound_t;
typedef struct { bound_t in, out; } span_t;
typedef struct { int n; span_t s[MAXSPANS]; } spans_t;

static bound_t B(double t, int leaf, int face) { bound_t b = {t, leaf, face, 0}; return b; }

static void spans_push(spans_t* L, bound_t in, bound_t out) {
    if (!(in.t < out.t) || L->n >= MAXSPANS) return;
    L->s[L->n].in = in; L->s[L->n].out = out; L->n++;
}

static void leaf_spans(const object_t* o, int leaf, ray_t wray, spans_t* L) {
    ray_t r; r.o = mul_point(o->w2o, wray.o); r.d = mul_dir(o->w2o, wray.d);
    L->n = 0;
    switch (o->kind) {
    case GMLRT_NODE_SPHERE: {
        double a = vdot(r.d, r.d), halfB = vdot(r.o, r.d), c = vdot(r.o, r.o) - 1.0;
        double disc = halfB * halfB - a * c;
        if (disc < 0.0) return;
        double sq = sqrt(disc);
        spans_push(L, B((-halfB - sq) / a, leaf, 0), B((-halfB + sq) / a, leaf, 0));
        return;
    }
    case GMLRT_NODE_PLANE: {                 /* half space y <= 0 */
        if (fabs(r.d.y) < 1e-6) {
            if (r.o.y <= 0) spans_push(L, B(-INFINITY, -1, 0), B(INFINITY, -1, 0));
            return;
        }
        double t = (0.0 - r.o.y) / r.d.y;
        if (r.d.y > 0) spans_push(L, B(-INFINITY, -1, 0), B(t, leaf, 0));
        else spans_push(L, B(t, leaf, 0), B(INFINITY, -1, 0));
        return;
    }
    case GMLRT_NODE_CUBE: {                  /* slabs; faces as plane.go:29-38 */
        static const int FACE_LO[3] = {2, 5, 0}, FACE_HI[3] = {3, 4, 1};   /* x, y, z */
        double oo[3] = {r.o.x, r.o.y, r.o.z}, dd[3] = {r.d.x, r.d.y, r.d.z};
        bound_t lo = B(-INFINITY, -1, 0), hi = B(INFINITY, -1, 0);
        for (int k = 0; k < 3; k++) {
            if (fabs(dd[k]) < 1e-12) { if (oo[k] < 0 || oo[k] > 1) return; continue; }
            bound_t a = B((0.0 - oo[k]) / dd[k], leaf, FACE_LO[k]), b = B((1.0 - oo[k]) / dd[k], leaf, FACE_HI[k]);
            if (a.t > b.t) { bound_t t = a; a = b; b = t; }
            if (a.t > lo.t) lo = a;
            if (b.t < hi.t) hi = b;
        }
        spans_push(L, lo, hi);
        return;
    }
    case GMLRT_NODE_CYLINDER: {
        bound_t lo = B(-INFINITY, -1, 0), hi = B(INFINITY, -1, 0);
        double a = r.d.x * r.d.x + r.d.z * r.d.z;
        double halfB = r.o.x * r.d.x + r.o.z * r.d.z;
        double c0 = r.o.x * r.o.x + r.o.z * r.o.z - 1.0;
        if (a > 1e-12) {
            double disc = halfB * halfB - a * c0;
            if (disc < 0.0) return;
            double sq = sqrt(disc);
            lo = B((-halfB - sq) / a, leaf, 0); hi = B((-halfB + sq) / a, leaf, 0);
        } else if (c0 > 0) return;
        if (fabs(r.d.y) > 1e-12) {
            bound_t b0 = B((0.0 - r.o.y) / r.d.y, leaf, 2), b1 = B((1.0 - r.o.y) / r.d.y, leaf, 1);
            if (b0.t > b1.t) { bound_t t = b0; b0 = b1; b1 = t; }
            if (b0.t > lo.t) lo = b0;
            if (b1.t < hi.t) hi = b1;
        } else if (r.o.y < 0 || r.o.y > 1) return;
        spans_push(L, lo, hi);
        return;
    }
    case GMLRT_NODE_CONE: {                  /* convex: min / max of the valid boundary candidates */
        double cand[3]; int cf[3]; int nc = 0;
        double a = r.d.x * r.d.x + r.d.z * r.d.z - r.d.y * r.d.y;
        double halfB = r.o.x * r.d.x + r.o.z * r.d.z - r.o.y * r.d.y;
        double c0 = r.o.x * r.o.x + r.o.z * r.o.z - r.o.y * r.o.y;
        if (fabs(a) > 1e-12) {
            double disc = halfB * halfB - a * c0;
            if (disc >= 0.0) {
                double sq = sqrt(disc);
                double ts[2] = {(-halfB - sq) / a, (-halfB + sq) / a};
                for (int k = 0; k < 2; k++) {
                    double y = r.o.y + r.d.y * ts[k];
                    if (y >= 0.0 && y <= 1.0) { cand[nc] = ts[k]; cf[nc++] = 0; }
                }
            }
        } else if (fabs(halfB) > 1e-12) {
            double t = -c0 / (2.0 * halfB);
            double y = r.o.y + r.d.y * t;
            if (y >= 0.0 && y <= 1.0) { cand[nc] = t; cf[nc++] = 0; }
        }
        if (fabs(r.d.y) > 1e-12) {
            double t = (1.0 - r.o.y) / r.d.y;
            v3 p = vadd(r.o, vscale(r.d, t));
            if (p.x * p.x + p.z * p.z <= 1.0) { cand[nc] = t; cf[nc++] = 1; }
        }
        if (nc < 2) return;
        int imin = 0, imax = 0;
        for (int k = 1; k < nc; k++) { if (cand[k] < cand[imin]) imin = k; if (cand[k] > cand[imax]) imax = k; }
        spans_push(L, B(cand[imin], leaf, cf[imin]), B(cand[imax], leaf, cf[imax]));
        return;
    }
    }
}

static void spans_complement(const spans_t* A, spans_t* out) {
    out->n = 0;
    bound_t prev = B(-INFINITY, -1, 0);
    for (int i = 0; i < A->n; i++) {
        bound_t e = A->s[i].in; e.flip ^= 1;
        spans_push(out, prev, e);
        prev = A->s[i].out; prev.flip ^= 1;
    }
    spans_push(out, prev, B(INFINITY, -1, 0));
}

static void spans_intersect(const spans_t* A, const spans_t* Bs, spans_t* out) {
    out->n = 0;
    int i = 0, j = 0;
    while (i < A->n && j < Bs->n) {
        const span_t *a = &A->s[i], *b = &Bs->s[j];
        bound_t in = (a->in.t >= b->in.t) ? a->in : b->in;
        bound_t ot = (a->out.t <= b->out.t) ? a->out : b->out;
        spans_push(out, in, ot);
        if (a->out.t <= b->out.t) i++; else j++;
    }
}

static void csg_spans(const scene_t* s, uint32_t ni, ray_t wray, spans_t* out) {
    const gmlrt_node* n = &s->nodes[ni];
    if (is_leaf_kind(n->kind)) { leaf_spans(&s->leaves[ni], (int)ni, wray, out); return; }
    spans_t a, b, t1, t2;
    csg_spans(s, n->a, wray, &a);
    csg_spans(s, n->b, wray, &b);
    if (n->kind == GMLRT_NODE_INTERSECT) { spans_intersect(&a, &b, out); return; }
    if (n->kind == GMLRT_NODE_DIFFERENCE) { spans_complement(&b, &t1); spans_intersect(&a, &t1, out); return; }
    /* union = complement(complement a  intersect  complement b) */
    spans_complement(&a, &t1); spans_complement(&b, &t2);
    spans_intersect(&t1, &t2, &a);
    spans_complement(&a, out);
}

static int csg_intersect(const scene_t* s, const object_t* o, ray_t wray, hit_t* out) {
    spans_t L;
    csg_spans(s, o->csg_root, wray, &L);
    for (int i = 0; i < L.n; i++) {
        const bound_t* b = NULL;
        if (L.s[i].in.t > 0.0) b = &L.s[i].in;
        else if (L.s[i].out.t > 0.0) b = &L.s[i].out;
        if (!b) continue;
        if (b->leaf < 0) return 0;      /* the solid extends to infinity along the ray: no surface */
        const object_t* lf = &s->leaves[b->leaf];
        ray_t r; r.o = mul_point(lf->w2o, wray.o); r.d = mul_dir(lf->w2o, wray.d);
        out->t = b->t; out->leaf = b->leaf; out->face = b->face; out->flip = b->flip;
        out->p_obj = vadd(r.o, vscale(r.d, b->t));
        return 1;
    }
    return 0;
}

/* ------------------------------------------------------------------ Render (raytracer.go:589-682) */
typedef struct {
    const scene_t* s;
    uint8_t* rgba;
    int32_t* primary_ids;     /* nullable: [H][W][4] */
    uint8_t* chaos;           /* nullable: [H][W] 1 = some ray of the pixel's tree saw a cube face rejected by rounding alone */
    int x0, y0, x1, y1;       /* crop, half-open */
    atomic_int next_item;
    int n_items, items_per_col;
    counters_t total;
    pthread_mutex_t mu;
} job_t;

static void render_item(job_t* job, int x, int ymin, counters_t* c) {
    const scene_t* s = job->s;
    int ymax = ymin + 20 < s->height ? ymin + 20 : s->height;
    go_pcg rng;
    go_pcg_seed(&rng, 0xDEADULL ^ (uint64_t)x, 0xBEEFULL ^ (uint64_t)ymin);     /* raytracer.go:632-634 */
    v3 eye = V(0.0, 0.0, -1.0);
    for (int y = ymin; y < ymax; y++) {
        if (y < job->y0 || y >= job->y1) {          /* outside the crop: keep the stream aligned */
            for (int k = 0; k < 8; k++) (void)go_pcg_uint64(&rng);
            continue;
        }
        v3 total = V(0, 0, 0);
        const unsigned long long chaos0 = tl_chaos;
        for (int sidx = 0; sidx < 4; sidx++) {
            double dx = go_pcg_float64(&rng) - 0.5;
            double dy = go_pcg_float64(&rng) - 0.5;
            double u = ((double)x + dx) / (double)(s->width - 1) * s->vw - s->vw / 2.0;
            double v = ((double)y + dy) / (double)(s->height - 1) * s->vh - s->vh / 2.0;
            ray_t ray;
            ray.o = V(u, -v, 0.0);
            ray.d = vnorm(vsub(ray.o, eye));
            int32_t* pid = job->primary_ids ? &job->primary_ids[((size_t)y * s->width + x) * 4 + sidx] : NULL;
            total = vadd(total, trace_ray(s, ray, s->depth, 0, c, pid));
        }
        if (job->chaos) job->chaos[(size_t)y * s->width + x] = tl_chaos != chaos0;
        v3 px = vscale(total, 1.0 / 4.0);
        uint8_t* o = job->rgba + ((size_t)y * s->width + x) * 4;
        /* Vec3.RGBA + color.RGBAModel (vec.go:104-107): uint8(uint32(c*65535) >> 8) */
        o[0] = (uint8_t)((uint32_t)(px.x * 65535) >> 8);
        o[1] = (uint8_t)((uint32_t)(px.y * 65535) >> 8);
        o[2] = (uint8_t)((uint32_t)(px.z * 65535) >> 8);
        o[3] = 0xFF;
    }
}

static void* worker(void* arg) {
    job_t* job = (job_t*)arg;
    counters_t c; memset(&c, 0, sizeof c);
    int ncols = job->x1 - job->x0;
    for (;;) {
        int it = atomic_fetch_add(&job->next_item, 1);
        if (it >= job->n_items) break;
        int x = job->x0 + it / job->items_per_col;       /* producer order: x outer, y batches inner (:662-677) */
        int ymin = (job->y0 / 20 + it % job->items_per_col) * 20;
        (void)ncols;
        render_item(job, x, ymin, &c);
    }
    pthread_mutex_lock(&job->mu);
    job->total.rays_primary += c.rays_primary; job->total.rays_reflect += c.rays_reflect;
    job->total.rays_refract += c.rays_refract; job->total.rays_shadow += c.rays_shadow;
    job->total.shadow_occluded += c.shadow_occluded; job->total.leaf_tests += c.leaf_tests; job->total.rays_hit += c.rays_hit;
    job->total.flops += c.flops; job->total.surface_error |= c.surface_error;
    pthread_mutex_unlock(&job->mu);
    return NULL;
}

/* Renders pixels [x0,x1) x [y0,y1) (pass 0,0,W,H for the whole frame) into a full-frame RGBA
 * buffer with `threads` workers (the reference uses 8, raytracer.go:725).  stats may be NULL. */
int oracle_render2(const void* blob, size_t blob_bytes, uint8_t* rgba, size_t rgba_bytes, int threads,
                   int x0, int y0, int x1, int y1, gmlrt_stats* stats, int32_t* primary_ids, uint8_t* chaos);
int oracle_render(const void* blob, size_t blob_bytes, uint8_t* rgba, size_t rgba_bytes, int threads,
                  int x0, int y0, int x1, int y1, gmlrt_stats* stats, int32_t* primary_ids) {
    return oracle_render2(blob, blob_bytes, rgba, rgba_bytes, threads, x0, y0, x1, y1, stats, primary_ids, NULL);
}
/* ... `chaos` (nullable, [H][W]): 1 where the pixel's value depends on a cube face rejected by rounding alone */
int oracle_render2(const void* blob, size_t blob_bytes, uint8_t* rgba, size_t rgba_bytes, int threads,
                   int x0, int y0, int x1, int y1, gmlrt_stats* stats, int32_t* primary_ids, uint8_t* chaos) {
    scene_t s;
    int rc = scene_init(&s, blob, blob_bytes);
    if (rc) { scene_free(&s); return rc; }
    if (!rgba || rgba_bytes != (size_t)4 * s.width * s.height) { scene_free(&s); return GMLRT_E_BAD_ARG; }
    if (x0 < 0) x0 = 0; if (y0 < 0) y0 = 0; if (x1 > s.width) x1 = s.width; if (y1 > s.height) y1 = s.height;
    job_t job; memset(&job, 0, sizeof job);
    job.s = &s; job.rgba = rgba; job.primary_ids = primary_ids; job.chaos = chaos;
    job.x0 = x0; job.y0 = y0; job.x1 = x1; job.y1 = y1;
    if (x1 > x0 && y1 > y0) {
        job.items_per_col = (y1 - 1) / 20 - y0 / 20 + 1;
        job.n_items = (x1 - x0) * job.items_per_col;
    }
    atomic_init(&job.next_item, 0);
    pthread_mutex_init(&job.mu, NULL);
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    for (int i = 0; i < threads; i++) pthread_create(&th[i], NULL, worker, &job);
    for (int i = 0; i < threads; i++) pthread_join(th[i], NULL);
    pthread_mutex_destroy(&job.mu);
    if (stats) {
        memset(stats, 0, sizeof *stats);
        stats->rays_primary = job.total.rays_primary; stats->rays_reflect = job.total.rays_reflect;
        stats->rays_refract = job.total.rays_refract; stats->rays_shadow = job.total.rays_shadow;
        stats->shadow_occluded = job.total.shadow_occluded; stats->leaf_tests = job.total.leaf_tests;
        stats->flops_credited = job.total.flops; stats->n_leaves = (uint32_t)s.n_objs; stats->rays_hit = job.total.rays_hit;
    }
    rc = job.total.surface_error ? GMLRT_E_SURFACE : 0;
    scene_free(&s);
    return rc;
}

/* ------------------------------------------------------------------ small probes for unit tests */
int oracle_eval_surface(const void* blob, size_t blob_bytes, uint32_t surface, int face, double u, double v, double* mat10) {
    scene_t s;
    int rc = scene_init(&s, blob, blob_bytes);
    if (!rc) rc = surface < s.h->n_surfaces ? (run_surface(&s, &s.surfaces[surface], face, u, v, mat10) ? GMLRT_E_SURFACE : 0)
                                             : GMLRT_E_BAD_ARG;
    scene_free(&s);
    return rc;
}

/* <Type>.Intersect on one leaf given by kind + matrices (cylinder_test.go style checks) */
int oracle_intersect(int kind, const double* w2o12, const double* ro, const double* rd, double* t, double* p_obj, int* face) {
    object_t o; memset(&o, 0, sizeof o);
    o.kind = kind; memcpy(o.w2o, w2o12, sizeof o.w2o);
    ray_t r; r.o = V(ro[0], ro[1], ro[2]); r.d = V(rd[0], rd[1], rd[2]);
    v3 p;
    if (!leaf_intersect(&o, r, t, &p, face)) return 0;
    p_obj[0] = p.x; p_obj[1] = p.y; p_obj[2] = p.z;
    return 1;
}

/* ComputeSurfaceProps normal + (u,v) are folded into surface_props; expose the world normal */
int oracle_normal(int kind, const double* o2w12, const double* w2o12, int face, const double* p_obj, double* n_world) {
    gmlrt_node n; memset(&n, 0, sizeof n);
    n.kind = (uint32_t)kind; memcpy(n.o2w, o2w12, sizeof n.o2w); memcpy(n.w2o, w2o12, sizeof n.w2o);
    object_t o; init_leaf(&o, &n);
    gmlrt_surface sf; memset(&sf, 0, sizeof sf);
    gmlrt_blob_header h; memset(&h, 0, sizeof h);
    scene_t s; memset(&s, 0, sizeof s); s.surfaces = &sf; s.h = &h;
    hit_t hit; memset(&hit, 0, sizeof hit); hit.face = face; hit.p_obj = V(p_obj[0], p_obj[1], p_obj[2]);
    hitex_t ex;
    if (surface_props(&s, &o, &hit, &ex)) return GMLRT_E_SURFACE;
    n_world[0] = ex.n_world.x; n_world[1] = ex.n_world.y; n_world[2] = ex.n_world.z;
    return 0;
}

double oracle_go_tan(double x) { return go_tan(x); }
double oracle_go_sin(double x) { return go_sin(x); }
double oracle_go_cos(double x) { return go_cos(x); }
double oracle_go_pow(double x, double y) { return go_pow(x, y); }
double oracle_go_atan2(double y, double x) { return go_atan2(y, x); }
double oracle_go_acos(double x) { return go_acos(x); }
void oracle_pcg(uint64_t seed1, uint64_t seed2, int n, uint64_t* out) {
    go_pcg p; go_pcg_seed(&p, seed1, seed2);
    for (int i = 0; i < n; i++) out[i] = go_pcg_uint64(&p);
}


/* ------------------------------------------------------------------ SSIM (internal/prim/ssim.go:27-182)
 * The reference's acceptance metric for golden images (raytracer_test.go:37-52, >= 0.99): 11x11 Gaussian
 * window (sigma 1.5, weights normalised to 1, ssim.go:146-165) over the 16-bit channel values that
 * image.Color.RGBA() returns (8-bit v -> v * 257), windows at x in [0, W-11), y in [0, H-11) - the last
 * possible start is left out, as in the reference's loops (:53, :58) -, per window the mean over R, G, B of
 * formula (13) with c1 = (0.01 * 65535)^2, c2 = (0.03 * 65535)^2, then the mean over all windows.  Column sums
 * are added in x order (the reference adds them in channel-arrival order, so its own last bits vary).
 * rgba: [H][W][4] bytes (image.RGBA.Pix).  Returns 0, or -1 (null / too small, ssim.go:31-36). */
#define SSIM_K 11
static double ssim_window(const uint8_t* a, const uint8_t* b, int w, int xs, int ys, const double* kern) {
    double m1[3] = {0, 0, 0}, m2[3] = {0, 0, 0};
    for (int k1 = 0; k1 < SSIM_K; k1++)
        for (int k2 = 0; k2 < SSIM_K; k2++) {
            const size_t at = ((size_t)(ys + k2) * w + (xs + k1)) * 4;
            const double wt = kern[k1 * SSIM_K + k2];
            for (int c = 0; c < 3; c++) { m1[c] += (double)(a[at + c] * 257u) * wt; m2[c] += (double)(b[at + c] * 257u) * wt; }
        }
    double v1[3] = {0, 0, 0}, v2[3] = {0, 0, 0}, cv[3] = {0, 0, 0};
    for (int k1 = 0; k1 < SSIM_K; k1++)
        for (int k2 = 0; k2 < SSIM_K; k2++) {
            const size_t at = ((size_t)(ys + k2) * w + (xs + k1)) * 4;
            const double wt = kern[k1 * SSIM_K + k2];
            for (int c = 0; c < 3; c++) {
                const double d1 = (double)(a[at + c] * 257u) - m1[c], d2 = (double)(b[at + c] * 257u) - m2[c];
                v1[c] += wt * (d1 * d1); v2[c] += wt * (d2 * d2); cv[c] += wt * d1 * d2;
            }
        }
    const double c1 = (0.01 * 65535.0) * (0.01 * 65535.0), c2 = (0.03 * 65535.0) * (0.03 * 65535.0);
    double s[3];
    for (int c = 0; c < 3; c++)
        s[c] = ((2 * m1[c] * m2[c] + c1) * (2 * cv[c] + c2)) / ((m1[c] * m1[c] + m2[c] * m2[c] + c1) * (v1[c] + v2[c] + c2));
    return (s[0] + s[1] + s[2]) / 3.0;
}

int oracle_ssim(const uint8_t* rgba_a, const uint8_t* rgba_b, int w, int h, double* out) {
    if (!rgba_a || !rgba_b || !out || w < SSIM_K || h < SSIM_K) return -1;
    double kern[SSIM_K * SSIM_K], total = 0.0;
    const double center = (SSIM_K - 1) / 2.0;
    for (int i = 0; i < SSIM_K; i++)
        for (int j = 0; j < SSIM_K; j++) {
            const double x = i - center, y = j - center;
            const double v = exp(-(x * x + y * y) / (2 * 1.5 * 1.5));
            kern[i * SSIM_K + j] = v; total += v;
        }
    for (int i = 0; i < SSIM_K * SSIM_K; i++) kern[i] /= total;
    double sum = 0.0; long long n = 0;
    for (int x = 0; x < w - SSIM_K; x++) {
        double col = 0.0;
        for (int y = 0; y < h - SSIM_K; y++) { col += ssim_window(rgba_a, rgba_b, w, x, y, kern); n++; }
        sum += col;
    }
    *out = n ? sum / (double)n : 0.0 / 0.0;      /* W == 11 or H == 11: no window, the reference divides 0 by 0 */
    return 0;
}
