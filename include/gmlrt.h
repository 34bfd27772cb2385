/*
 * gmlrt.h - C ABI of the B200-native renderer for the GML `render` operator.
 *
 * This is the drop-in boundary for the hot path of timdestan/go-raytracer.  The
 * reference crosses it as a Go callback
 *     EvalState.Render func(*EvalState, *RenderArgs) error      (internal/gml/evaluator.go:48)
 * whose stock implementation is
 *     scene, err := ConvertRenderArgsToScene(args, state)       (raytracer.go:724-754)
 *     img := Render(scene)                                      (raytracer.go:589-682)
 * The replacement keeps the callback; inside it the host lowers *RenderArgs into
 * one flat, pointer-free "scene blob" (format below) and calls gmlrt_render().
 * INTEGRATION.md shows the cgo binding a maintainer adds.
 *
 * Everything here is extern "C", plain pointers and sizes.  No torch types, no
 * callbacks, no CPU fallback: if no CUDA device is usable gmlrt_create() fails.
 */
#ifndef GMLRT_H
#define GMLRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ scene blob v1
 *
 * Little-endian, every section 8-byte aligned, all offsets in bytes from the start
 * of the blob.  Replaces the in-memory *gml.RenderArgs tree (evaluator.go:14-28) +
 * the per-leaf matrices ConvertRenderArgsToScene derives (raytracer.go:756-830).
 */
#define GMLRT_BLOB_MAGIC   0x424C4D47u /* "GMLB" */
#define GMLRT_BLOB_VERSION 1u

typedef struct gmlrt_blob_header {
    uint32_t magic;         /* GMLRT_BLOB_MAGIC */
    uint32_t version;       /* GMLRT_BLOB_VERSION */
    uint64_t total_bytes;   /* size of the whole blob */
    int32_t  width;         /* RenderArgs.Width  (pixels) */
    int32_t  height;        /* RenderArgs.Height (pixels) */
    int32_t  depth;         /* RenderArgs.Depth as given; <=0 means 3 (raytracer.go:592-595) */
    uint32_t flags;         /* reserved, 0 */
    double   fov;           /* degrees as given; <=0 means 90 (raytracer.go:597-600) */
    double   ambient[3];    /* RenderArgs.AmbientLight */
    double   bg_start[3];   /* RenderArgs.BgColorStart (0 for plain `render`) */
    double   bg_end[3];     /* RenderArgs.BgColorEnd */
    uint32_t n_nodes;       /* scene tree nodes */
    uint32_t root;          /* index of the root node (RenderArgs.Scene) */
    uint32_t n_lights;
    uint32_t n_surfaces;
    uint32_t n_fpool;       /* doubles in the constant pool */
    uint32_t n_itab;        /* int32 entries in the integer table */
    uint32_t n_code;        /* uint32 words of surface bytecode */
    uint32_t reserved0;
    uint64_t off_nodes;     /* gmlrt_node[n_nodes] */
    uint64_t off_lights;    /* gmlrt_light[n_lights] */
    uint64_t off_surfaces;  /* gmlrt_surface[n_surfaces] */
    uint64_t off_fpool;     /* double[n_fpool] */
    uint64_t off_itab;      /* int32_t[n_itab] */
    uint64_t off_code;      /* uint32_t[n_code] */
    uint8_t  reserved1[64];
} gmlrt_blob_header;        /* 256 bytes */

/* Scene-tree node kinds.  UNION/DIFFERENCE and the four leaf kinds SPHERE, CUBE,
 * CYLINDER, PLANE are the reference's (evaluator.go:163-287).  INTERSECT and CONE are
 * extensions named by the north_star; the reference has no behaviour for them.     */
enum {
    GMLRT_NODE_UNION = 0, GMLRT_NODE_INTERSECT = 1, GMLRT_NODE_DIFFERENCE = 2,
    GMLRT_NODE_SPHERE = 16, GMLRT_NODE_CUBE = 17, GMLRT_NODE_CYLINDER = 18,
    GMLRT_NODE_CONE = 19, GMLRT_NODE_PLANE = 20
};

typedef struct gmlrt_node {
    uint32_t kind;
    uint32_t a, b;          /* operators: child node indices.  UNION: `a` is the operand that
                               was on TOP of the GML stack (it is scanned first,
                               evaluator.go:1091-1101); DIFFERENCE: a - b.  Leaves: 0. */
    uint32_t surface;       /* leaves: index into the surface table */
    double   o2w[12];       /* leaves: ObjectToWorld, rows 0..2 of the 4x4 (row 3 is 0 0 0 1) */
    double   w2o[12];       /* leaves: WorldToObject = Mat4.Inverse(o2w) (vec.go:319-365),
                               computed by the host so the bits equal the reference's */
} gmlrt_node;               /* 208 bytes */

enum { GMLRT_LIGHT_POINT = 0, GMLRT_LIGHT_DIRECTIONAL = 1, GMLRT_LIGHT_SPOT = 2 };

typedef struct gmlrt_light {
    uint32_t kind;          /* POINT is the reference's gml.PointLight (evaluator.go:289-292) */
    uint32_t reserved;
    double   pos[3];        /* POINT/SPOT: position.  DIRECTIONAL: direction of travel */
    double   color[3];
    double   at[3];         /* SPOT: the point the light aims at */
    double   cutoff;        /* SPOT: half-angle in degrees */
    double   exponent;      /* SPOT: falloff exponent */
} gmlrt_light;              /* 96 bytes */

enum { GMLRT_SURFACE_CONST = 0, GMLRT_SURFACE_CODE = 1 };

/* Material layout used everywhere (gml.Material, evaluator.go:136-150):
 *   [0..2] Color  [3] Reflectivity  [4] Fuzziness  [5] Transparency
 *   [6] RefractiveIndex  [7] Kd  [8] Ks  [9] SpecularExponent                         */
typedef struct gmlrt_surface {
    uint32_t kind;          /* CONST: `material` is the folded constant (evaluator.go:729-750);
                               CODE: run `code_len` words of bytecode at `code_off` per hit */
    uint32_t code_off;      /* in uint32 words from off_code */
    uint32_t code_len;
    uint32_t n_regs;        /* registers the bytecode uses (<= GMLRT_VM_MAX_REGS) */
    double   material[10];  /* CODE: components the host proved constant, NaN where computed per hit
                               (lets the renderer see e.g. Transparency == 0 and size its queues) */
} gmlrt_surface;            /* 96 bytes */

/* ------------------------------------------------------------------ surface bytecode
 *
 * A GML surface closure (code + captured environment, evaluator.go:80-83) is compiled
 * by the host into register code: closures and `apply`/`if` are resolved at compile
 * time, what remains is straight-line typed arithmetic with forward jumps.  One
 * instruction = one uint32  op | dst<<8 | a<<16 | b<<24 , optionally followed by
 * immediate words.  On entry r0 = face (int), r1 = u, r2 = v.  Registers hold either
 * a real or an integer (bools are integers 0/1); array values are (offset,length)
 * in two consecutive registers.  Faults (index out of range, integer divide by zero,
 * GMLRT_OP_ERR) abort the frame with GMLRT_E_SURFACE.
 */
#define GMLRT_VM_MAX_REGS 128
enum {
    GMLRT_OP_RET   = 0,   /* a: first of 10 consecutive registers holding the Material */
    GMLRT_OP_LDK   = 1,   /* dst = fpool[a | b<<8] */
    GMLRT_OP_LDI   = 2,   /* dst = (int) next word */
    GMLRT_OP_MOV   = 3,   /* dst = a */
    GMLRT_OP_ADDF  = 4, GMLRT_OP_SUBF = 5, GMLRT_OP_MULF = 6, GMLRT_OP_DIVF = 7,
    GMLRT_OP_NEGF  = 8,
    GMLRT_OP_ADDI  = 9, GMLRT_OP_SUBI = 10, GMLRT_OP_MULI = 11, GMLRT_OP_DIVI = 12,
    GMLRT_OP_MODI  = 13, GMLRT_OP_NEGI = 14,
    GMLRT_OP_EQF   = 15, GMLRT_OP_LESSF = 16, GMLRT_OP_EQI = 17, GMLRT_OP_LESSI = 18,
    GMLRT_OP_SIN   = 19,  /* degrees, evaluator.go:929-937 */
    GMLRT_OP_COS   = 20,
    GMLRT_OP_SQRT  = 21,
    GMLRT_OP_CLAMPF = 22,
    GMLRT_OP_FLOOR = 23,  /* real -> int */
    GMLRT_OP_FRAC  = 24,  /* x - trunc(x), evaluator.go:968-977 */
    GMLRT_OP_REAL  = 25,  /* int -> real (extension) */
    GMLRT_OP_ACOS  = 26,  /* degrees (extension) */
    GMLRT_OP_ASIN  = 27,  /* degrees (extension) */
    GMLRT_OP_LDARR = 28,  /* dst,dst+1 = next word (offset), next word (length) */
    GMLRT_OP_LEN   = 29,  /* dst = a+1 */
    GMLRT_OP_GETF  = 30,  /* dst = fpool[off + idx]           a: array regs, b: index reg */
    GMLRT_OP_GETP  = 31,  /* dst..dst+2 = fpool[off + 3*idx ..] */
    GMLRT_OP_GETI  = 32,  /* dst = itab[off + idx] */
    GMLRT_OP_GETA  = 33,  /* dst,dst+1 = itab[off + 2*idx], itab[off + 2*idx + 1] */
    GMLRT_OP_JMP   = 34,  /* pc += next word (forward, relative to the word after it) */
    GMLRT_OP_JMPF  = 35,  /* if (a == 0) pc += next word */
    GMLRT_OP_ERR   = 36   /* unconditional fault */
};

/* ------------------------------------------------------------------ API */

typedef struct gmlrt_ctx gmlrt_ctx;

enum {
    GMLRT_OK = 0,
    GMLRT_E_BAD_ARG = -1,      /* null pointer, bad size */
    GMLRT_E_BAD_BLOB = -2,     /* magic/version/bounds check failed */
    GMLRT_E_UNSUPPORTED = -3,  /* node kind / feature the renderer rejects (raytracer.go:825-826) */
    GMLRT_E_NO_DEVICE = -4,    /* no usable CUDA device: there is no CPU fallback */
    GMLRT_E_CUDA = -5,         /* CUDA runtime error, text in gmlrt_last_error() */
    GMLRT_E_OOM = -6,          /* device allocation failed */
    GMLRT_E_SURFACE = -7       /* a surface function faulted at render time (the reference
                                  panics there, raytracer.go:499-501) */
};

/* gmlrt_create flags */
#define GMLRT_PRECISION_F32        0u  /* throughput mode: float arithmetic, FMA contraction */
#define GMLRT_PRECISION_F64_PARITY 1u  /* parity mode: double, reference operation order, no FMA */
#define GMLRT_PRECISION_MASK       1u

typedef struct gmlrt_stats {
    double   ms_total;          /* gmlrt_render / gmlrt_trace wall time on the device timeline */
    double   ms_kernels;        /* sum over devices' max: trace kernels only */
    double   ms_h2d, ms_d2h;    /* blob upload, raster download (gmlrt_render only) */
    uint64_t rays_primary;      /* closest-hit queries from the camera (4 per pixel) */
    uint64_t rays_reflect;      /* closest-hit queries from reflection (raytracer.go:512-529) */
    uint64_t rays_refract;      /* closest-hit queries from refraction (raytracer.go:532-556) */
    uint64_t rays_shadow;       /* inShadow queries (raytracer.go:411-429) */
    uint64_t shadow_occluded;   /* ... that found an occluder */
    uint64_t leaf_tests;        /* ray x leaf tests the REFERENCE runs for these queries */
    double   flops_credited;    /* ... times the per-leaf-type flop table: the brute-force-equivalent work */
    uint64_t kernel_launches;   /* kernels launched by this call */
    uint32_t n_devices;
    uint32_t n_leaves;
    uint64_t cull_tests;        /* executed: per-ray bounding-sphere tests (after the per-warp packet stage) */
    uint64_t exact_tests;       /* executed: second-phase <Type>.Intersect tests on the survivors (counting mode) */
    double   flops_executed;    /* 15 / 7 / 7 per cull test (GENERAL / PIVOT / AXIS), 20 per packet test, the per-leaf-type
                                   table per exact test (the last term needs counting mode) */
    double   ms_class[9];       /* profiling mode: device ms per kernel class: gen, cull of camera rays, cull of reflection/refraction
                                   rays, cull of shadow rays, shade, resolve, pixels, exact tests, shadow-ray sort */
    uint64_t launches_class[9]; /* ... and launches per class */
    uint64_t packet_tests;      /* executed: per-warp packet tests (one lane x one object) */
    uint64_t cull_tests_family[3]; /* cull_tests split by family: GENERAL (reflection / refraction rays), PIVOT, AXIS */
    uint64_t rays_hit;          /* closest-hit queries (primary + reflect + refract) that hit something = rays shaded */
    uint64_t exact_tests_shadow; /* the part of exact_tests run for shadow rays (counting mode) */
    uint64_t rays_hit_primary;  /* camera rays that hit something (the level-0 part of rays_hit) */
    uint64_t list_tests;        /* class x object tests of the last upload's candidate lists (not per frame) */
} gmlrt_stats;

/* Create a context on `n` devices (`devices` = CUDA ordinals; n >= 1).  With n > 1 the
 * image is split into row bands, the scene is replicated and the bands are gathered on
 * devices[0] over NVLink peer copies.  Replaces nothing in the reference (its "context"
 * is the 8 goroutines of raytracer.go:618-660); amortises allocations across calls.   */
int gmlrt_create(gmlrt_ctx** out, const int* devices, int n, unsigned flags);

/* Blocking, drop-in for `Render(ConvertRenderArgsToScene(args, state))`:
 * host blob in, host RGBA8 raster out (rgba_bytes == 4*width*height, row-major, stride
 * 4*width, alpha 0xFF - the layout of *image.RGBA.Pix, raytracer.go:590,656).         */
int gmlrt_render(gmlrt_ctx*, const void* blob, size_t blob_bytes,
                 uint8_t* rgba, size_t rgba_bytes, gmlrt_stats* stats /* nullable */);

/* The same three steps separately, for callers that keep the scene resident:
 * upload (validate + H2D), trace (kernels only; raster stays in HBM), download (D2H). */
int gmlrt_upload(gmlrt_ctx*, const void* blob, size_t blob_bytes);   /* a blob byte-identical to the last one (same partition, same GMLRT_* environment) is only copied
                                                                         to the device again: what was derived from it is kept (GMLRT_NO_UPLOAD_CACHE=1 turns that off) */
int gmlrt_trace(gmlrt_ctx*, gmlrt_stats* stats /* nullable */);
int gmlrt_download(gmlrt_ctx*, uint8_t* rgba, size_t rgba_bytes);

/* Multi-process gather on the HOST (the alternative SURVEY.md section 8e names): copy only the rows THIS context rendered - the
 * bands of its partition - into the full-size raster `rgba` and leave every other row untouched.  N processes that map one
 * raster (shared memory, page-locked in each process with cudaHostRegister) each call it after their own gmlrt_trace and then
 * meet at a barrier: the frame crosses N PCIe links at once instead of one.  The reference has one address space and
 * writes pixels with img.Set from its 8 goroutines (raytracer.go:656): this is that, across processes.  Not for a context
 * that has attached another process's frame (its rows are not in its own frame).  Partition (0, 1): same as gmlrt_download. */
int gmlrt_download_rows(gmlrt_ctx*, uint8_t* rgba, size_t rgba_bytes);

/* Page-locked host memory for the raster (cudaHostAlloc): a frame buffer from here takes the device-to-host DMA
 * directly.  The reference allocates its raster with image.NewRGBA (raytracer.go:590), i.e. pageable Go heap memory;
 * gmlrt_download / gmlrt_render accept that too and stage it through the context's own page-locked buffer in
 * pipelined chunks (a little slower).  INTEGRATION.md's bridge wraps a gmlrt_alloc_host buffer as image.RGBA.Pix.   */
void* gmlrt_alloc_host(size_t bytes);
void  gmlrt_free_host(void* p);

/* Multi-process sharding (one process per GPU, e.g. under torchrun): this context renders only the
 * row bands (8 rows each)  part, part + n_parts, ...  of the frame and leaves the other rows zero, so that the
 * per-process frames can be summed / gathered by the caller (NCCL), or stored into one frame (gmlrt_frame_ipc_*).  Call before
 * gmlrt_upload.  The default (0, 1) renders the whole frame.                                     */
int gmlrt_set_partition(gmlrt_ctx*, int part, int n_parts);

/* Device pointer of the RGBA8 frame on devices[0] (valid after gmlrt_upload), for callers that
 * keep the raster on the GPU.                                                                    */
void* gmlrt_frame_device(gmlrt_ctx*);

/* One process per GPU without a collective: the gathering process exports its frame (after gmlrt_upload; the frame
 * keeps its address while width x height is unchanged), the others attach it, and from then on their level-0 resolve kernel
 * stores finished rows straight into that frame over NVLink (CUDA IPC peer mapping).  The caller orders "all
 * producers finished" before it reads the frame (a barrier).  attach(NULL, 0) detaches.  Once exported, the frame is
 * never cleared by gmlrt_upload again (the producers' stores are not ordered with the exporter's uploads): rows
 * nobody renders keep whatever they held.                                                                        */
#define GMLRT_IPC_HANDLE_BYTES 64
int gmlrt_frame_ipc_export(gmlrt_ctx*, void* handle, size_t handle_bytes);
int gmlrt_frame_ipc_attach(gmlrt_ctx*, const void* handle, size_t handle_bytes);

/* Enable (1) / disable (0) the counting kernels that fill the ray and flop fields of
 * gmlrt_stats; counting costs a few percent, so timed runs leave it off.              */
int gmlrt_set_counting(gmlrt_ctx*, int enabled);

/* Enable (1) / disable (0) per-kernel-class timing: a CUDA event is recorded before every launch and
 * gmlrt_stats.ms_class / launches_class are filled.  bench.py uses it for the roofline of the scan kernels. */
int gmlrt_set_profiling(gmlrt_ctx*, int enabled);

/* Per-sample id of the object hit by each primary ray (-1 = background), 4*W*H int32 in
 * (y, x, sample) order, from the last trace.  For the "zero primary hit/miss flips"
 * check of the north_star.  Requires gmlrt_set_counting(ctx, 1) before the trace.      */
int gmlrt_primary_hits(gmlrt_ctx*, int32_t* ids, size_t n_ids);

/* CUDA stream of device 0 that gmlrt_trace launches on (a cudaStream_t), so a caller
 * can bracket it with its own events.                                                 */
void* gmlrt_stream(gmlrt_ctx*);

/* FP32 FFMA throughput of `device` in TFLOP/s from an 8-chain FFMA microbenchmark (about 50 ms): the
 * measured denominator of the FP32 roofline (MEASURED_PEAKS.json holds no FP32 figure).  < 0 on error. */
double gmlrt_measure_fp32_tflops(int device);

const char* gmlrt_last_error(gmlrt_ctx*);   /* text for the last non-OK code; never NULL */
void gmlrt_destroy(gmlrt_ctx*);
/* The reference's image-comparison metric, prim.SSIM (internal/prim/ssim.go:27-144; its golden tests accept
 * >= 0.99, raytracer_test.go:37-52), computed on devices[0]: two host RGBA8 rasters of the same size in, the index
 * out.  GMLRT_E_BAD_ARG when an image is smaller than the 11x11 window (the reference returns an error there too). */
int gmlrt_ssim(gmlrt_ctx*, const uint8_t* rgba_a, const uint8_t* rgba_b, int width, int height, double* ssim);

/* gmlrt_ssim without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The two rasters are DEVICE pointers on devices[0] (RGBA8, 4*width*height bytes each, e.g. gmlrt_frame_device of two
 * contexts); the result goes to a host double.  The kernel's per-block sums live in a file-scope __device__ array of
 * GMLRT_SSIM_NOALLOC_MAX_WINDOWS / 256 doubles (1 MiB, held by every GPU that loads the library until exit), so
 * (width-11)*(height-11) must not exceed GMLRT_SSIM_NOALLOC_MAX_WINDOWS (an 8K frame fits): larger images are
 * rejected with GMLRT_E_BAD_ARG and a message naming the limit.  Calls no cudaMalloc / cudaFree / allocating library
 * routine.  All calls share the array (it keeps the last call's sums; every entry that is read is overwritten first):
 * within one process calls are serialised by a mutex, anything else that overlaps on the GPU races.  Same result, bit
 * for bit, as gmlrt_ssim on the same pixels.                                                                        */
#define GMLRT_SSIM_NOALLOC_MAX_WINDOWS (1 << 25)
int gmlrt_ssim_device_noalloc(gmlrt_ctx*, const void* d_rgba_a, const void* d_rgba_b, int width, int height, double* ssim);

int  gmlrt_blob_version(void);
const char* gmlrt_build_info(void);         /* arch, compiler, precision modes compiled in */

#ifdef __cplusplus
}
#endif
#endif /* GMLRT_H */
