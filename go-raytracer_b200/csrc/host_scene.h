// host_scene.h - blob validation + the one-off scene conversion, on the host, in double.
//
// Replaces convertGMLSceneObjects (/root/reference/raytracer.go:756-830): breadth-first
// flattening of Union nodes into the scan-order object list (top-of-stack operand first),
// per-leaf matrices, plane / cube-face world normals (createPlane :764-774) and the camera
// set-up of Render (:592-603).  The reference repeats this 8 times (once per goroutine,
// :738-751); here it runs once per gmlrt_upload.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/gmlrt.h"
#include "gomath.cuh"

namespace gmlrt {

// device-order kind slots (scan kernels walk the leaves kind by kind)
enum Kind { K_SPHERE = 0, K_CUBE = 1, K_CYLINDER = 2, K_CONE = 3, K_PLANE = 4, K_COUNT = 5,
            K_CSG = 5, K_OBJ_KINDS = 6 };   // K_CSG: a scan-order object that is an intersect/difference subtree (EXTENSION)
#define GMLRT_CSG_SWAPPED 8                 // operator flag in a postfix program: the operands were evaluated b first, a is on top
#define GMLRT_CSG_MAX_STACK 6               // span-list stack depth of the device evaluator (left-deep trees need 2)
static const double kFlopsByKind[K_COUNT] = {56, 81, 74, 70, 47};

struct HostLeaf {
    int kind;                 // Kind
    uint32_t surface;
    double o2w[12], w2o[12];
    double face_normal[6][3]; // plane: [0]; cube: per face
    int scan_index;           // position in the reference's []SceneObject (-1 inside a CSG object)
    int csg_object;           // index of the owning CSG object, -1 for plain objects
    uint32_t node;            // blob node index
};

// One CSG object = a subtree rooted at an intersect/difference node, in postfix order.
struct HostCsgOp { int op; int leaf; };   // op: -1 push leaf `leaf` (index into leaves), else GMLRT_NODE_* operator
struct HostCsgObject {
    int scan_index; std::vector<HostCsgOp> postfix; int max_stack;
    double C[3]; double R;   // bounding sphere (cull phase)
    double flops;            // credited per test: member table flops, +2 per quadric (both roots), SURVEY.md 8(d)
};

struct HostScene {
    gmlrt_blob_header h;
    int width = 0, height = 0, depth = 0;
    double vw = 0, vh = 0;
    std::vector<HostLeaf> leaves;          // plain objects first (scan order), then CSG members
    int n_plain = 0;                       // leaves[0..n_plain) are plain scan-order objects
    int n_objects = 0;                     // scan-order objects = plain leaves + CSG objects
    std::vector<int> object_of_scan;       // scan index -> leaf index (>=0) or -(csg index)-1
    std::vector<HostCsgObject> csg;
    std::vector<gmlrt_light> lights;
    std::vector<gmlrt_surface> surfaces;
    std::vector<double> fpool;
    std::vector<int32_t> itab;
    std::vector<uint32_t> code;
    bool may_branch = false;               // some material can spawn reflection AND refraction
    bool any_reflect = false, any_refract = false;
};

inline bool is_leaf_kind(uint32_t k) { return k >= 16 && k <= 20; }

inline void mul_dir_t(const double* m, const double* v, double* out) {   // (M^T) v, NormalMat.MulDir
    out[0] = m[0] * v[0] + m[4] * v[1] + m[8] * v[2];
    out[1] = m[1] * v[0] + m[5] * v[1] + m[9] * v[2];
    out[2] = m[2] * v[0] + m[6] * v[1] + m[10] * v[2];
}
inline void normalize3(double* v) {
    double m = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    v[0] /= m; v[1] /= m; v[2] /= m;
}

// largest eigenvalue of the symmetric 3x3 matrix {a00,a01,a02,a11,a12,a22} (trigonometric closed form)
inline double sym3_max_eig(const double* a) {
    const double a00 = a[0], a01 = a[1], a02 = a[2], a11 = a[3], a12 = a[4], a22 = a[5];
    const double p1 = a01 * a01 + a02 * a02 + a12 * a12;
    if (p1 == 0.0) return std::fmax(a00, std::fmax(a11, a22));
    const double q = (a00 + a11 + a22) / 3.0;
    const double p2 = (a00 - q) * (a00 - q) + (a11 - q) * (a11 - q) + (a22 - q) * (a22 - q) + 2.0 * p1;
    const double p = std::sqrt(p2 / 6.0);
    const double b00 = (a00 - q) / p, b11 = (a11 - q) / p, b22 = (a22 - q) / p, b01 = a01 / p, b02 = a02 / p, b12 = a12 / p;
    double r = (b00 * (b11 * b22 - b12 * b12) - b01 * (b01 * b22 - b12 * b02) + b02 * (b01 * b12 - b11 * b02)) / 2.0;
    r = r < -1.0 ? -1.0 : (r > 1.0 ? 1.0 : r);
    return q + 2.0 * p * std::cos(std::acos(r) / 3.0);
}

// World-space bounding sphere of a leaf, for the cull phase of the scan kernels (render_impl.cuh):
// the object-space solid's enclosing sphere mapped through ObjectToWorld, radius scaled by the
// largest singular value of its linear part.  Planes are unbounded (radius = +inf).
inline void leaf_bound(int kind, const double* o2w, double* C, double& Rw) {
    static const double c0[K_COUNT][3] = {{0, 0, 0}, {0.5, 0.5, 0.5}, {0, 0.5, 0}, {0, 1, 0}, {0, 0, 0}};
    static const double r0[K_COUNT] = {1.0, 0.8660254037844387 /* sqrt(3)/2 */, 1.118033988749895 /* sqrt(5)/2 */, 1.0, 0.0};
    const double* c = c0[kind];
    for (int r = 0; r < 3; r++) C[r] = o2w[4 * r] * c[0] + o2w[4 * r + 1] * c[1] + o2w[4 * r + 2] * c[2] + o2w[4 * r + 3];
    if (kind == K_PLANE) { C[0] = C[1] = C[2] = 0.0; Rw = INFINITY; return; }
    double ata[6] = {0, 0, 0, 0, 0, 0};
    int k = 0;
    for (int i = 0; i < 3; i++) for (int j = i; j < 3; j++, k++)
        for (int r = 0; r < 3; r++) ata[k] += o2w[4 * r + i] * o2w[4 * r + j];
    const double frob = std::sqrt(ata[0] + ata[3] + ata[5]);                   // always >= sigma_max
    double smax = std::sqrt(std::fmax(sym3_max_eig(ata), 0.0)) * (1.0 + 1e-9);
    if (!(smax <= frob)) smax = frob;                                          // NaN or closed-form trouble: fall back to the safe bound
    Rw = r0[kind] * (smax > 0.0 ? smax : frob) * (1.0 + 1e-9);
    if (!(Rw == Rw)) Rw = INFINITY;
}

// smallest sphere enclosing two spheres
inline void sphere_union(const double* C1, double R1, const double* C2, double R2, double* C, double& R) {
    if (std::isinf(R1) || std::isinf(R2)) { C[0] = C[1] = C[2] = 0; R = INFINITY; return; }
    const double d[3] = {C2[0] - C1[0], C2[1] - C1[1], C2[2] - C1[2]};
    const double dist = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (dist + R2 <= R1) { C[0] = C1[0]; C[1] = C1[1]; C[2] = C1[2]; R = R1; return; }
    if (dist + R1 <= R2) { C[0] = C2[0]; C[1] = C2[1]; C[2] = C2[2]; R = R2; return; }
    R = 0.5 * (dist + R1 + R2) * (1.0 + 1e-12);
    const double f = (R - R1) / dist;
    for (int k = 0; k < 3; k++) C[k] = C1[k] + d[k] * f;
}

static const double kCubeNormals[6][3] = {{0, 0, -1}, {0, 0, 1}, {-1, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, -1, 0}};

inline void init_leaf(HostLeaf& L, const gmlrt_node& n, uint32_t node_index) {
    static const int map[5] = {K_SPHERE, K_CUBE, K_CYLINDER, K_CONE, K_PLANE};
    L.kind = map[n.kind - 16];
    L.surface = n.surface;
    L.node = node_index;
    std::memcpy(L.o2w, n.o2w, sizeof L.o2w);
    std::memcpy(L.w2o, n.w2o, sizeof L.w2o);
    std::memset(L.face_normal, 0, sizeof L.face_normal);
    L.scan_index = -1; L.csg_object = -1;
    if (L.kind == K_PLANE) {
        const double up[3] = {0, 1, 0};
        mul_dir_t(L.w2o, up, L.face_normal[0]);
        normalize3(L.face_normal[0]);
    } else if (L.kind == K_CUBE) {
        for (int f = 0; f < 6; f++) { mul_dir_t(L.w2o, kCubeNormals[f], L.face_normal[f]); normalize3(L.face_normal[f]); }
    }
}

// returns GMLRT_OK or an error code; `err` gets the text
inline int parse_blob(const void* blob, size_t nbytes, HostScene& S, std::string& err) {
    if (!blob || nbytes < sizeof(gmlrt_blob_header)) { err = "blob too small"; return GMLRT_E_BAD_BLOB; }
    std::memcpy(&S.h, blob, sizeof S.h);
    const gmlrt_blob_header& h = S.h;
    if (h.magic != GMLRT_BLOB_MAGIC) { err = "bad magic"; return GMLRT_E_BAD_BLOB; }
    if (h.version != GMLRT_BLOB_VERSION) { err = "unsupported blob version"; return GMLRT_E_BAD_BLOB; }
    if (h.total_bytes != nbytes) { err = "total_bytes does not match the buffer size"; return GMLRT_E_BAD_BLOB; }
    auto fits = [&](uint64_t off, uint64_t n, uint64_t sz) { return off <= nbytes && n * sz <= nbytes - off && off % 8 == 0; };
    if (!fits(h.off_nodes, h.n_nodes, sizeof(gmlrt_node)) || !fits(h.off_lights, h.n_lights, sizeof(gmlrt_light)) ||
        !fits(h.off_surfaces, h.n_surfaces, sizeof(gmlrt_surface)) || !fits(h.off_fpool, h.n_fpool, 8) ||
        !fits(h.off_itab, h.n_itab, 4) || !fits(h.off_code, h.n_code, 4)) { err = "section out of bounds"; return GMLRT_E_BAD_BLOB; }
    if (h.n_nodes == 0 || h.root >= h.n_nodes) { err = "bad root"; return GMLRT_E_BAD_BLOB; }
    if (h.width <= 0 || h.height <= 0 || (uint64_t)h.width * (uint64_t)h.height > (1ull << 28)) { err = "bad image size"; return GMLRT_E_BAD_BLOB; }
    const uint8_t* base = (const uint8_t*)blob;
    std::vector<gmlrt_node> nodes(h.n_nodes);
    std::memcpy(nodes.data(), base + h.off_nodes, sizeof(gmlrt_node) * h.n_nodes);
    S.lights.resize(h.n_lights);   if (h.n_lights) std::memcpy(S.lights.data(), base + h.off_lights, sizeof(gmlrt_light) * h.n_lights);
    S.surfaces.resize(h.n_surfaces); if (h.n_surfaces) std::memcpy(S.surfaces.data(), base + h.off_surfaces, sizeof(gmlrt_surface) * h.n_surfaces);
    S.fpool.resize(h.n_fpool);     if (h.n_fpool) std::memcpy(S.fpool.data(), base + h.off_fpool, 8 * (size_t)h.n_fpool);
    S.itab.resize(h.n_itab);       if (h.n_itab) std::memcpy(S.itab.data(), base + h.off_itab, 4 * (size_t)h.n_itab);
    S.code.resize(h.n_code);       if (h.n_code) std::memcpy(S.code.data(), base + h.off_code, 4 * (size_t)h.n_code);
    S.any_reflect = S.any_refract = S.may_branch = false;                 // a context parses many scenes: nothing carries over
    for (auto& l : S.lights) if (l.kind > GMLRT_LIGHT_SPOT) { err = "unknown light kind"; return GMLRT_E_UNSUPPORTED; }
    for (auto& sf : S.surfaces) {
        if (sf.kind > GMLRT_SURFACE_CODE) { err = "unknown surface kind"; return GMLRT_E_BAD_BLOB; }
        if (sf.kind == GMLRT_SURFACE_CODE && ((uint64_t)sf.code_off + sf.code_len > h.n_code || sf.n_regs > GMLRT_VM_MAX_REGS)) {
            err = "surface code out of bounds"; return GMLRT_E_BAD_BLOB;
        }
        // a component is "possibly > 0" unless it is a known constant <= 0
        bool refl = !(sf.material[3] <= 0.0), transp = !(sf.material[5] <= 0.0);
        S.any_reflect |= refl; S.any_refract |= transp;
        S.may_branch |= (refl && transp);
    }

    S.width = h.width; S.height = h.height;
    S.depth = h.depth > 0 ? h.depth : 3;                                  // raytracer.go:592-595
    double fov = h.fov > 0.0 ? h.fov : 90.0;                              // raytracer.go:597-600
    double fov_rad = fov * GM_PI / 180.0;
    S.vw = 2.0 / gomath::go_tan(fov_rad / 2.0);                           // raytracer.go:601-602 (Go's Tan)
    S.vh = S.vw * ((double)S.height / (double)S.width);

    // breadth-first flattening of unions (raytracer.go:776-828)
    std::vector<uint32_t> queue{h.root};
    std::vector<HostLeaf> csg_members;
    S.leaves.clear(); S.csg.clear(); S.object_of_scan.clear();
    for (size_t qh = 0; qh < queue.size(); qh++) {
        if (queue.size() > (size_t)h.n_nodes) { err = "scene graph is not a tree"; return GMLRT_E_BAD_BLOB; }
        uint32_t ni = queue[qh];
        const gmlrt_node& n = nodes[ni];
        if (n.kind == GMLRT_NODE_UNION) {
            if (n.a >= h.n_nodes || n.b >= h.n_nodes) { err = "child index out of range"; return GMLRT_E_BAD_BLOB; }
            queue.push_back(n.a); queue.push_back(n.b);
        } else if (is_leaf_kind(n.kind)) {
            if (n.surface >= h.n_surfaces) { err = "surface index out of range"; return GMLRT_E_BAD_BLOB; }
            HostLeaf L; init_leaf(L, n, ni);
            L.scan_index = (int)S.object_of_scan.size();
            S.object_of_scan.push_back((int)S.leaves.size());
            S.leaves.push_back(L);
        } else if (n.kind == GMLRT_NODE_INTERSECT || n.kind == GMLRT_NODE_DIFFERENCE) {
            // EXTENSION: the reference rejects these (raytracer.go:825-826).  Postfix walk, no recursion.
            HostCsgObject obj; obj.scan_index = (int)S.object_of_scan.size(); obj.max_stack = 0;
            // Postfix order with the deeper operand first (Sethi-Ullman numbering): the span stack of the device evaluator
            // then needs log2(leaves) + 1 lists instead of the depth of the tree as written (chess.gml nests 9 deep, its
            // pieces need 3).  When that puts b before a, the operator carries GMLRT_CSG_SWAPPED: a is then on top.
            std::vector<int> need(h.n_nodes, 0);
            {
                struct F2 { uint32_t node; int state; };
                std::vector<F2> st2{{ni, 0}};
                size_t seen = 0;
                while (!st2.empty()) {
                    F2& f = st2.back();
                    if (++seen > 4 * (size_t)h.n_nodes + 8) { err = "scene graph is not a tree"; return GMLRT_E_BAD_BLOB; }
                    const gmlrt_node& m = nodes[f.node];
                    if (is_leaf_kind(m.kind)) { need[f.node] = 1; st2.pop_back(); }
                    else if (m.kind <= GMLRT_NODE_DIFFERENCE) {
                        if (m.a >= h.n_nodes || m.b >= h.n_nodes) { err = "child index out of range"; return GMLRT_E_BAD_BLOB; }
                        if (f.state == 0) { f.state = 1; st2.push_back({m.a, 0}); }
                        else if (f.state == 1) { f.state = 2; st2.push_back({m.b, 0}); }
                        else { const int na = need[m.a], nb2 = need[m.b]; need[f.node] = na == nb2 ? na + 1 : std::max(na, nb2); st2.pop_back(); }
                    } else { err = "unknown scene object type"; return GMLRT_E_UNSUPPORTED; }
                }
            }
            struct Frame { uint32_t node; int state; };
            std::vector<Frame> st{{ni, 0}};
            int depth = 0;
            size_t visited = 0;
            while (!st.empty()) {
                Frame& f = st.back();
                if (++visited > 4 * (size_t)h.n_nodes + 8) { err = "scene graph is not a tree"; return GMLRT_E_BAD_BLOB; }
                const gmlrt_node& m = nodes[f.node];
                if (is_leaf_kind(m.kind)) {
                    if (m.surface >= h.n_surfaces) { err = "surface index out of range"; return GMLRT_E_BAD_BLOB; }
                    HostLeaf L; init_leaf(L, m, f.node);
                    L.csg_object = (int)S.csg.size();
                    obj.postfix.push_back({-1, (int)csg_members.size()});
                    csg_members.push_back(L);
                    if (++depth > obj.max_stack) obj.max_stack = depth;
                    st.pop_back();
                } else if (m.kind <= GMLRT_NODE_DIFFERENCE) {
                    if (m.a >= h.n_nodes || m.b >= h.n_nodes) { err = "child index out of range"; return GMLRT_E_BAD_BLOB; }
                    const bool swapped = need[m.b] > need[m.a];
                    if (f.state == 0) { f.state = 1; st.push_back({swapped ? m.b : m.a, 0}); }
                    else if (f.state == 1) { f.state = 2; st.push_back({swapped ? m.a : m.b, 0}); }
                    else { obj.postfix.push_back({(int)m.kind | (swapped ? GMLRT_CSG_SWAPPED : 0), -1}); depth--; st.pop_back(); }
                } else { err = "unknown scene object type"; return GMLRT_E_UNSUPPORTED; }
            }
            S.object_of_scan.push_back(-(int)S.csg.size() - 1);
            S.csg.push_back(std::move(obj));
        } else {
            err = "unknown scene object type";                           // raytracer.go:825-826
            return GMLRT_E_UNSUPPORTED;
        }
    }
    S.n_plain = (int)S.leaves.size();
    S.n_objects = (int)S.object_of_scan.size();
    for (auto& obj : S.csg) for (auto& op : obj.postfix) if (op.op < 0) op.leaf += S.n_plain;
    S.leaves.insert(S.leaves.end(), csg_members.begin(), csg_members.end());
    for (auto& obj : S.csg) {
        if (obj.max_stack > GMLRT_CSG_MAX_STACK) { err = "CSG tree needs a deeper span stack than GMLRT_CSG_MAX_STACK"; return GMLRT_E_UNSUPPORTED; }
        struct B { double C[3], R; };
        std::vector<B> st;
        obj.flops = 0;
        for (auto& op : obj.postfix) {
            if (op.op < 0) {
                const HostLeaf& L = S.leaves[op.leaf];
                B b; leaf_bound(L.kind, L.o2w, b.C, b.R);
                st.push_back(b);
                obj.flops += kFlopsByKind[L.kind] + ((L.kind == K_SPHERE || L.kind == K_CYLINDER || L.kind == K_CONE) ? 2 : 0);
            } else {
                B b = st.back(); st.pop_back();
                B a = st.back(); st.pop_back();
                if (op.op & GMLRT_CSG_SWAPPED) std::swap(a, b);
                B r;
                const int kind = op.op & ~GMLRT_CSG_SWAPPED;
                if (kind == GMLRT_NODE_DIFFERENCE) r = a;                                   // a - b lies inside a
                else if (kind == GMLRT_NODE_INTERSECT) r = (b.R < a.R) ? b : a;            // inside both: keep the smaller bound
                else sphere_union(a.C, a.R, b.C, b.R, r.C, r.R);
                st.push_back(r);
            }
        }
        obj.C[0] = st.back().C[0]; obj.C[1] = st.back().C[1]; obj.C[2] = st.back().C[2]; obj.R = st.back().R;
    }
    return GMLRT_OK;
}

}  // namespace gmlrt
