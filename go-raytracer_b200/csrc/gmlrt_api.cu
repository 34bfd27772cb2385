// gmlrt_api.cu - the C ABI of include/gmlrt.h: context, blob validation, row-band scheduling
// across devices, NVLink peer gather, copy-back.  This is the "internal/cuda bridge" of the
// north_star: it owns device memory, streams and the result copy; the host language (Go via cgo,
// Python via ctypes, ...) only hands over a blob pointer and an RGBA buffer.
//
// Multi-GPU: pixels are independent, so the image is cut into row bands that
// are dealt round-robin to the devices; the scene is replicated; each finished band is copied
// into device 0's frame with cudaMemcpyPeerAsync (NVLink 5 through NVSwitch).  No collective.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <memory>
#include <mutex>
#include <condition_variable>
#include <functional>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/gmlrt.h"
#include "host_scene.h"
#include "renderer.h"

using namespace gmlrt;

// A few host threads that stay with the context: gmlrt_download into pageable memory hands them the copy-out of every staged
// chunk (creating and joining 7 threads per call cost a third of a millisecond of a 1.2 ms download).
class HostPool {
public:
    explicit HostPool(int n) : n_(n) {
        for (int t = 1; t < n_; t++) workers_.emplace_back([this, t] { loop(t); });
    }
    ~HostPool() {
        { std::lock_guard<std::mutex> g(m_); stop_ = true; gen_++; }
        cv_.notify_all();
        for (auto& w : workers_) w.join();
    }
    int size() const { return n_; }
    // job(t) on every thread t = 0 .. n-1 (the caller is thread 0); returns when all are done
    void run(const std::function<void(int)>& job) {
        { std::lock_guard<std::mutex> g(m_); job_ = &job; pending_ = n_ - 1; gen_++; }
        cv_.notify_all();
        job(0);
        std::unique_lock<std::mutex> g(m_);
        done_.wait(g, [this] { return pending_ == 0; });
        job_ = nullptr;
    }
private:
    void loop(int t) {
        unsigned long long seen = 0;
        for (;;) {
            const std::function<void(int)>* job;
            {
                std::unique_lock<std::mutex> g(m_);
                cv_.wait(g, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                job = job_;
            }
            (*job)(t);
            { std::lock_guard<std::mutex> g(m_); pending_--; }
            done_.notify_one();
        }
    }
    int n_;
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(int)>* job_ = nullptr;
    unsigned long long gen_ = 0;
    int pending_ = 0;
    bool stop_ = false;
};

struct gmlrt_ctx {
    unsigned flags = 0;
    std::unique_ptr<HostPool> pool;                          // created by the first download into pageable memory
    std::vector<int> devices;
    std::vector<std::unique_ptr<IRenderer>> r;
    HostScene scene;
    bool have_scene = false, have_frame = false, counting = false;
    std::string err;
    uint8_t* pinned = nullptr; size_t pinned_bytes = 0;     // staging for pageable caller buffers (gmlrt_download)
    std::vector<cudaEvent_t> ev_chunk;                       // ... one event per staged chunk
    cudaEvent_t ev_a = nullptr, ev_b = nullptr;
    std::vector<cudaEvent_t> ev_done;                        // per device: its bands are in device 0's frame
    int band_rows = 0;
    int part = 0, n_parts = 1;                               // multi-process sharding: this context renders bands part, part+n_parts, ...
    std::vector<char> peer_ok;                               // per device: it can store straight into devices[0]'s frame
    uint8_t* remote_frame = nullptr;                         // another process's frame opened over CUDA IPC (gmlrt_frame_ipc_attach)
    bool frame_exported = false;                             // this context's frame is an IPC gather target (gmlrt_frame_ipc_export): uploads no longer clear it
    // A frame loop uploads the same scene again and again (and cmd/gml renders one scene per `render` call, but a caller that
    // only moves on to the next frame does not): when blob, partition and tuning environment are byte-identical to the last
    // successful upload, everything derived from the blob (converted scene, cull tables, class lists, tile plan) is still on
    // the devices and only the blob itself crosses the bus again, into `blob_dev`.
    std::vector<uint8_t> last_blob; std::string last_env; int last_part = -1, last_n_parts = -1; bool cache_ok = false;
    void* blob_dev = nullptr; size_t blob_dev_bytes = 0;
};

extern char** environ;
static std::string tuning_env() {                            // every GMLRT_* variable: uploads read their knobs from the environment
    std::string k;
    for (char** e = environ; e && *e; e++) if (std::strncmp(*e, "GMLRT_", 6) == 0) { k += *e; k += '\n'; }
    return k;
}

static thread_local std::string g_last_error = "";

static int fail(gmlrt_ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg;
    g_last_error = msg;
    return code;
}

#define API_CU(expr)                                                                   \
    do {                                                                               \
        cudaError_t e_ = (expr);                                                       \
        if (e_ != cudaSuccess)                                                         \
            return fail(c, e_ == cudaErrorMemoryAllocation ? GMLRT_E_OOM : GMLRT_E_CUDA, \
                        std::string(#expr) + ": " + cudaGetErrorString(e_));           \
    } while (0)

// FP32 roofline denominator measured on the box: 8 independent FFMA chains per thread, every SM saturated.
__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 12345.678f) out[0] = s;          // keeps the chains alive
}

extern "C" {

double gmlrt_measure_fp32_tflops(int device) {
    if (cudaSetDevice(device) != cudaSuccess) return -1.0;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1.0;
    float* out = nullptr;
    if (cudaMalloc(&out, 256) != cudaSuccess) return -1.0;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int blocks = prop.multiProcessorCount * 8, iters = 1 << 15;
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {              // first rep warms up
        cudaEventRecord(a);
        k_fma_peak<<<blocks, 256>>>(out, iters, 0.999f, 1e-3f);
        cudaEventRecord(b);
        if (cudaEventSynchronize(b) != cudaSuccess) { best = -1.0; break; }
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        const double flops = (double)blocks * 256.0 * (double)iters * 64.0 * 2.0;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(out);
    return best;
}

int gmlrt_blob_version(void) { return (int)GMLRT_BLOB_VERSION; }

const char* gmlrt_build_info(void) {
    return "gmlrt sm_100a; CUDA "
#define STR2(x) #x
#define STR(x) STR2(x)
           STR(CUDART_VERSION) "; precision modes: f32 (FMA), f64-parity (-fmad=false); no CPU fallback";
}

int gmlrt_create(gmlrt_ctx** out, const int* devices, int n, unsigned flags) {
    if (!out) return fail(nullptr, GMLRT_E_BAD_ARG, "out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0)
        return fail(nullptr, GMLRT_E_NO_DEVICE, std::string("no usable CUDA device (there is no CPU fallback): ") +
                                                     (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    std::vector<int> devs;
    if (!devices || n <= 0) devs.push_back(0);
    else devs.assign(devices, devices + n);
    for (int d : devs) if (d < 0 || d >= count) return fail(nullptr, GMLRT_E_BAD_ARG, "device ordinal out of range");
    std::unique_ptr<gmlrt_ctx> c(new gmlrt_ctx);
    c->flags = flags;
    c->devices = devs;
    for (int d : devs) {
        IRenderer* r = (flags & GMLRT_PRECISION_MASK) == GMLRT_PRECISION_F64_PARITY ? make_renderer_f64(d) : make_renderer_f32(d);
        if (!r) return fail(nullptr, GMLRT_E_CUDA, "renderer initialisation failed on device " + std::to_string(d));
        c->r.emplace_back(r);
    }
    // peer access to device 0 for the band gather (ignored when unsupported: cudaMemcpyPeer still works, staged)
    c->peer_ok.assign(devs.size(), 0);
    for (size_t i = 1; i < devs.size(); i++) {
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, devs[i], devs[0]) == cudaSuccess && can) {
            cudaSetDevice(devs[i]);
            cudaError_t pe = cudaDeviceEnablePeerAccess(devs[0], 0);
            if (pe == cudaSuccess || pe == cudaErrorPeerAccessAlreadyEnabled) c->peer_ok[i] = 1;
            if (pe != cudaSuccess) cudaGetLastError();
        }
    }
    cudaSetDevice(devs[0]);
    if (cudaEventCreate(&c->ev_a) != cudaSuccess || cudaEventCreate(&c->ev_b) != cudaSuccess)
        return fail(nullptr, GMLRT_E_CUDA, "cudaEventCreate failed");
    c->ev_done.resize(devs.size(), nullptr);
    for (size_t i = 0; i < devs.size(); i++) {
        cudaSetDevice(devs[i]);
        if (cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming) != cudaSuccess)
            return fail(nullptr, GMLRT_E_CUDA, "cudaEventCreate failed");
    }
    cudaSetDevice(devs[0]);
    *out = c.release();
    return GMLRT_OK;
}

void gmlrt_destroy(gmlrt_ctx* c) {
    if (!c) return;
    for (auto& r : c->r) { cudaSetDevice(r->device()); cudaStreamSynchronize(r->stream()); }
    if (c->remote_frame) { cudaSetDevice(c->r[0]->device()); cudaIpcCloseMemHandle(c->remote_frame); }
    if (c->pinned) cudaFreeHost(c->pinned);
    if (c->blob_dev) { cudaSetDevice(c->r[0]->device()); cudaFree(c->blob_dev); }
    if (c->ev_a) cudaEventDestroy(c->ev_a);
    if (c->ev_b) cudaEventDestroy(c->ev_b);
    for (auto ev : c->ev_done) if (ev) cudaEventDestroy(ev);
    for (auto ev : c->ev_chunk) cudaEventDestroy(ev);
    c->r.clear();
    delete c;
}

const char* gmlrt_last_error(gmlrt_ctx* c) { return c ? c->err.c_str() : g_last_error.c_str(); }

void* gmlrt_stream(gmlrt_ctx* c) { return c && !c->r.empty() ? (void*)c->r[0]->stream() : nullptr; }

int gmlrt_set_counting(gmlrt_ctx* c, int enabled) {
    if (!c) return GMLRT_E_BAD_ARG;
    c->counting = enabled != 0;
    return GMLRT_OK;
}

int gmlrt_set_profiling(gmlrt_ctx* c, int enabled) {
    if (!c) return GMLRT_E_BAD_ARG;
    for (auto& r : c->r) r->set_profiling(enabled != 0);
    return GMLRT_OK;
}

int gmlrt_upload(gmlrt_ctx* c, const void* blob, size_t blob_bytes) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    const bool had_scene = c->have_scene;
    c->have_scene = false; c->have_frame = false;
    std::string err;
    const std::string env = tuning_env();
    const bool same = had_scene && c->cache_ok && blob && blob_bytes == c->last_blob.size() && c->last_part == c->part &&
                      c->last_n_parts == c->n_parts && env == c->last_env && std::memcmp(blob, c->last_blob.data(), blob_bytes) == 0;
    if (same) {
        IRenderer* r = c->r[0].get();
        API_CU(cudaSetDevice(r->device()));
        if (c->blob_dev_bytes < blob_bytes) {
            if (c->blob_dev) cudaFree(c->blob_dev);
            c->blob_dev = nullptr; c->blob_dev_bytes = 0;
            API_CU(cudaMalloc(&c->blob_dev, blob_bytes));
            c->blob_dev_bytes = blob_bytes;
        }
        API_CU(cudaMemcpyAsync(c->blob_dev, blob, blob_bytes, cudaMemcpyHostToDevice, r->stream()));
    } else {
        c->cache_ok = false;
        int rc = parse_blob(blob, blob_bytes, c->scene, err);
        if (rc != GMLRT_OK) return fail(c, rc, err);
        int min_rows = c->scene.height;
        for (auto& r : c->r) {
            r->set_share((int)c->r.size() * c->n_parts);
            rc = r->upload(c->scene, err);
            if (rc != GMLRT_OK) return fail(c, rc, err);
            min_rows = std::min(min_rows, r->max_rows_per_tile());
        }
        (void)min_rows;
        c->last_blob.assign(static_cast<const uint8_t*>(blob), static_cast<const uint8_t*>(blob) + blob_bytes);
        c->last_env = env; c->last_part = c->part; c->last_n_parts = c->n_parts;
        c->cache_ok = getenv("GMLRT_NO_UPLOAD_CACHE") == nullptr;
    }
    // rows are dealt to the devices of all processes in bands of 8 (round-robin: every device sees every part of the
    // image, so the load balances); a device renders its bands as one run of virtual rows (RowMap), i.e. the same few
    // large launches however many devices share the frame
    const int nd = (int)c->r.size() * c->n_parts;
    c->band_rows = nd > 1 ? 8 : c->scene.height;
    if (const char* e = getenv("GMLRT_BAND_ROWS")) { const int v = atoi(e); if (v > 0 && nd > 1) c->band_rows = v; }
    IRenderer* r0 = c->r[0].get();
    for (size_t i = 0; i < c->r.size(); i++) {
        uint8_t* target = nullptr;
        if (c->remote_frame) target = c->remote_frame;                 // every device of this process stores into the remote frame
        else if (i > 0 && c->peer_ok[i]) target = r0->frame();         // NVLink stores from the level-0 k_resolve: the gather is fused
        c->r[i]->set_frame_target(target);
    }
    // Rows owned by other processes stay zero, so the rank frames can be summed (NCCL-reduce gather).  NOT when this
    // frame has been exported as an IPC gather target: peer ranks store their rows into it from their own level-0
    // resolve kernels, unordered with this process's uploads, and a clear enqueued here could land after their stores.
    if (c->n_parts > 1 && !c->remote_frame && !c->frame_exported) {
        API_CU(cudaSetDevice(r0->device()));
        API_CU(cudaMemsetAsync(r0->frame(), 0, (size_t)c->scene.width * c->scene.height * 4, r0->stream()));
    }
    c->have_scene = true;
    return GMLRT_OK;
}

int gmlrt_set_partition(gmlrt_ctx* c, int part, int n_parts) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    if (n_parts < 1 || part < 0 || part >= n_parts) return fail(c, GMLRT_E_BAD_ARG, "need 0 <= part < n_parts");
    c->part = part; c->n_parts = n_parts;
    c->have_scene = false;      // band height depends on the partition: upload again
    return GMLRT_OK;
}

void* gmlrt_frame_device(gmlrt_ctx* c) { return c && !c->r.empty() ? (void*)c->r[0]->frame() : nullptr; }

int gmlrt_frame_ipc_export(gmlrt_ctx* c, void* handle, size_t handle_bytes) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    if (!handle || handle_bytes < sizeof(cudaIpcMemHandle_t)) return fail(c, GMLRT_E_BAD_ARG, "handle buffer must hold GMLRT_IPC_HANDLE_BYTES");
    if (!c->have_scene) return fail(c, GMLRT_E_BAD_ARG, "gmlrt_frame_ipc_export before a successful gmlrt_upload");
    API_CU(cudaSetDevice(c->r[0]->device()));
    cudaIpcMemHandle_t h;
    API_CU(cudaIpcGetMemHandle(&h, c->r[0]->frame()));
    std::memcpy(handle, &h, sizeof h);
    c->frame_exported = true;
    return GMLRT_OK;
}

int gmlrt_frame_ipc_attach(gmlrt_ctx* c, const void* handle, size_t handle_bytes) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    API_CU(cudaSetDevice(c->r[0]->device()));
    if (c->remote_frame) { cudaIpcCloseMemHandle(c->remote_frame); c->remote_frame = nullptr; }
    if (handle) {
        if (handle_bytes < sizeof(cudaIpcMemHandle_t)) return fail(c, GMLRT_E_BAD_ARG, "handle buffer must hold GMLRT_IPC_HANDLE_BYTES");
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handle, sizeof h);
        void* p = nullptr;
        API_CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        c->remote_frame = (uint8_t*)p;
    }
    for (auto& r : c->r) r->set_frame_target(c->remote_frame);
    if (!c->remote_frame)
        for (size_t i = 1; i < c->r.size(); i++) c->r[i]->set_frame_target(c->peer_ok[i] ? c->r[0]->frame() : nullptr);
    return GMLRT_OK;
}

static int trace_once(gmlrt_ctx* c, gmlrt_stats* stats, bool& overflow) {
    overflow = false;
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    if (!c->have_scene) return fail(c, GMLRT_E_BAD_ARG, "gmlrt_trace before a successful gmlrt_upload");
    std::string err;
    const int nd = (int)c->r.size();
    const int H = c->scene.height, W = c->scene.width;
    IRenderer* r0 = c->r[0].get();
    API_CU(cudaSetDevice(r0->device()));
    for (auto& r : c->r) r->reset_totals();
    API_CU(cudaEventRecord(c->ev_a, r0->stream()));
    // the other devices start after ev_a so that ev_a..ev_b on device 0 brackets the whole job
    for (int i = 1; i < nd; i++) { API_CU(cudaSetDevice(c->r[i]->device())); API_CU(cudaStreamWaitEvent(c->r[i]->stream(), c->ev_a, 0)); }
    // one launcher per device: a frame is ~60 launches per device, and issued from one thread the devices of an 8-GPU
    // context would start one after the other (the launches, not the kernels, were the critical path)
    std::vector<int> drc(nd, GMLRT_OK);
    std::vector<std::string> derr(nd);
    auto run_device = [&](int di) {
        IRenderer* r = c->r[di].get();
        RowMap map;
        map.band_rows = c->band_rows; map.n_parts = nd * c->n_parts; map.part = c->part * nd + di;
        const int vrows = map.count(H);
        if (vrows == 0) return;
        drc[di] = r->trace_rows(0, vrows, map, c->counting, derr[di]);
        if (drc[di] != GMLRT_OK || di == 0) return;
        // gather on the producing device's stream: frame rows only when the resolve kernel could not store them itself
        cudaError_t e = cudaSetDevice(r->device());
        const size_t band = (size_t)c->band_rows * W * 4, pitch = band * map.n_parts, first = band * map.part;
        const int full = vrows / c->band_rows, rest = vrows % c->band_rows;
        auto gather = [&](uint8_t* dst, const uint8_t* src, size_t scale) -> cudaError_t {
            cudaError_t g = cudaSuccess;
            if (full) g = cudaMemcpy2DAsync(dst + first * scale, pitch * scale, src + first * scale, pitch * scale, band * scale, full, cudaMemcpyDefault, r->stream());
            if (g == cudaSuccess && rest)
                g = cudaMemcpyAsync(dst + (first + pitch * full) * scale, src + (first + pitch * full) * scale, (size_t)rest * W * 4 * scale, cudaMemcpyDefault, r->stream());
            return g;
        };
        if (e == cudaSuccess && !c->peer_ok[di] && !c->remote_frame) e = gather(r0->frame(), r->frame(), 1);
        if (e == cudaSuccess && c->counting && r->primary_ids() && r0->primary_ids())
            e = gather((uint8_t*)r0->primary_ids(), (const uint8_t*)r->primary_ids(), 4);
        if (e != cudaSuccess) { drc[di] = GMLRT_E_CUDA; derr[di] = std::string("band gather: ") + cudaGetErrorString(e); }
    };
    if (nd == 1) run_device(0);
    else {
        std::vector<std::thread> launchers;
        for (int di = 1; di < nd; di++) launchers.emplace_back(run_device, di);
        run_device(0);
        for (auto& t : launchers) t.join();
    }
    for (int di = 0; di < nd; di++) if (drc[di] != GMLRT_OK) return fail(c, drc[di], derr[di]);
    for (int i = 1; i < nd; i++) {
        API_CU(cudaSetDevice(c->r[i]->device()));
        API_CU(cudaEventRecord(c->ev_done[i], c->r[i]->stream()));
        API_CU(cudaSetDevice(r0->device()));
        API_CU(cudaStreamWaitEvent(r0->stream(), c->ev_done[i], 0));
    }
    API_CU(cudaSetDevice(r0->device()));
    API_CU(cudaEventRecord(c->ev_b, r0->stream()));
    API_CU(cudaStreamSynchronize(r0->stream()));
    for (int i = 1; i < nd; i++) { API_CU(cudaSetDevice(c->r[i]->device())); API_CU(cudaStreamSynchronize(c->r[i]->stream())); }
    API_CU(cudaSetDevice(r0->device()));
    c->have_frame = true;
    float ms = 0;
    API_CU(cudaEventElapsedTime(&ms, c->ev_a, c->ev_b));
    TraceTotals sum;
    for (auto& r : c->r) {
        TraceTotals t;
        int rc = r->fetch_totals(t, err);
        if (rc != GMLRT_OK) return fail(c, rc, err);
        sum.rays_primary += t.rays_primary; sum.rays_reflect += t.rays_reflect; sum.rays_refract += t.rays_refract;
        sum.rays_shadow += t.rays_shadow; sum.shadow_occluded += t.shadow_occluded; sum.rays_hit += t.rays_hit; sum.launches += t.launches;
        sum.shadow_tests += t.shadow_tests; sum.shadow_flops += t.shadow_flops; sum.exact_tests += t.exact_tests; sum.exact_tests_shadow += t.exact_tests_shadow; sum.exact_flops += t.exact_flops; sum.packet_tests += t.packet_tests;
        sum.rays_hit_primary += t.rays_hit_primary; sum.list_tests = std::max(sum.list_tests, t.list_tests);
        for (int k = 0; k < 3; k++) sum.cull_by_family[k] += t.cull_by_family[k];
        for (int k = 0; k < KC_COUNT; k++) { sum.ms_class[k] = std::max(sum.ms_class[k], t.ms_class[k]); sum.launches_class[k] += t.launches_class[k]; }
        sum.error |= t.error; overflow = overflow || t.overflow;
    }
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_total = ms; stats->ms_kernels = ms;
        stats->rays_primary = sum.rays_primary; stats->rays_reflect = sum.rays_reflect; stats->rays_refract = sum.rays_refract;
        stats->rays_shadow = sum.rays_shadow; stats->shadow_occluded = sum.shadow_occluded; stats->rays_hit = sum.rays_hit;
        stats->kernel_launches = sum.launches; stats->n_devices = (uint32_t)nd; stats->n_leaves = (uint32_t)c->scene.n_objects;
        // credited work: a closest-hit query tests every leaf; a shadow query is credited
        // with the tests its lane ran up to its own first occluder (counting mode only)
        double all = 0; unsigned long long nleaf = 0;
        for (int i = 0; i < c->scene.n_plain; i++) { all += kFlopsByKind[c->scene.leaves[i].kind]; nleaf++; }
        for (auto& obj : c->scene.csg) { all += obj.flops; nleaf++; }
        const unsigned long long closest = sum.rays_primary + sum.rays_reflect + sum.rays_refract;
        stats->leaf_tests = closest * nleaf;
        stats->flops_credited = (double)closest * all;
        stats->leaf_tests += sum.shadow_tests;
        stats->flops_credited += (double)sum.shadow_flops;
        // executed work, from device counters: per-ray cull tests by family (GENERAL: 7 FFMA + 1 FADD = 15 flops; PIVOT:
        // FMUL + 3 FFMA = 7; AXIS: 3 FADD + 2 FFMA = 7), per-warp packet tests (about 20 flops: two dot products, a
        // reciprocal, three square roots, the angle inequality) and the exact tests on the survivors (counting mode)
        stats->cull_tests = sum.cull_by_family[0] + sum.cull_by_family[1] + sum.cull_by_family[2];
        stats->packet_tests = sum.packet_tests;
        for (int k = 0; k < 3; k++) stats->cull_tests_family[k] = sum.cull_by_family[k];
        stats->exact_tests = sum.exact_tests; stats->exact_tests_shadow = sum.exact_tests_shadow;
        stats->rays_hit_primary = sum.rays_hit_primary; stats->list_tests = sum.list_tests;
        stats->flops_executed = 15.0 * (double)sum.cull_by_family[0] + 7.0 * (double)(sum.cull_by_family[1] + sum.cull_by_family[2]) +
                                20.0 * (double)sum.packet_tests + (double)sum.exact_flops;
        for (int k = 0; k < KC_COUNT; k++) { stats->ms_class[k] = sum.ms_class[k]; stats->launches_class[k] = sum.launches_class[k]; }
    }
    if (overflow) return GMLRT_OK;      // the caller shrinks the tiles and traces again
    if (sum.error) return fail(c, GMLRT_E_SURFACE, "a surface function faulted at render time (the reference panics here, raytracer.go:499-501)");
    return GMLRT_OK;
}

int gmlrt_trace(gmlrt_ctx* c, gmlrt_stats* stats) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    // a cull launch that runs out of candidate queue raises Counters::overflow: that frame is discarded, the tiles are
    // halved (same buffers, so twice the queue per ray) and the frame is traced again
    for (int attempt = 0; attempt < 12; attempt++) {
        bool overflow = false;
        int rc = trace_once(c, stats, overflow);
        if (rc != GMLRT_OK || !overflow) return rc;
        c->have_frame = false;
        std::string err;
        for (auto& r : c->r)
            if (!r->shrink_tiles(c->scene, err)) return fail(c, GMLRT_E_OOM, "candidate queue overflow: " + err);
    }
    return fail(c, GMLRT_E_OOM, "candidate queue overflow persists after shrinking the tiles");
}

// ------------------------------------------------------------------ SSIM (internal/prim/ssim.go:27-182)
// The reference's acceptance metric for a rendered image against a golden (raytracer_test.go:37-52): 11x11 Gaussian
// window over 16-bit channel values, windows at x in [0, W-11), y in [0, H-11), mean over R, G, B of formula (13),
// mean over the windows.  One thread per window, float64 like the reference (the Go version takes seconds at 4K: one
// goroutine per column, 121 taps x 2 passes per window); a block's windows are summed in shared memory and the block
// sums are added on the host in block order, so the result does not depend on scheduling.
__constant__ double c_ssim_kernel[121];

__global__ void __launch_bounds__(256) k_ssim(const uchar4* a, const uchar4* b, int w, int h, double* block_sums) {
    __shared__ double part[256];
    const int nx = w - 11, ny = h - 11;
    const long long win = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double val = 0.0;
    if (win < (long long)nx * ny) {
        const int xs = (int)(win % nx), ys = (int)(win / nx);
        double m1[3] = {0, 0, 0}, m2[3] = {0, 0, 0};
        for (int k2 = 0; k2 < 11; k2++)
            for (int k1 = 0; k1 < 11; k1++) {
                const uchar4 p = a[(size_t)(ys + k2) * w + xs + k1], q = b[(size_t)(ys + k2) * w + xs + k1];
                const double wt = c_ssim_kernel[k1 * 11 + k2];
                m1[0] += (double)(p.x * 257u) * wt; m1[1] += (double)(p.y * 257u) * wt; m1[2] += (double)(p.z * 257u) * wt;
                m2[0] += (double)(q.x * 257u) * wt; m2[1] += (double)(q.y * 257u) * wt; m2[2] += (double)(q.z * 257u) * wt;
            }
        double v1[3] = {0, 0, 0}, v2[3] = {0, 0, 0}, cv[3] = {0, 0, 0};
        for (int k2 = 0; k2 < 11; k2++)
            for (int k1 = 0; k1 < 11; k1++) {
                const uchar4 p = a[(size_t)(ys + k2) * w + xs + k1], q = b[(size_t)(ys + k2) * w + xs + k1];
                const double wt = c_ssim_kernel[k1 * 11 + k2];
                const double pa[3] = {(double)(p.x * 257u), (double)(p.y * 257u), (double)(p.z * 257u)};
                const double qa[3] = {(double)(q.x * 257u), (double)(q.y * 257u), (double)(q.z * 257u)};
#pragma unroll
                for (int ch = 0; ch < 3; ch++) {
                    const double d1 = pa[ch] - m1[ch], d2 = qa[ch] - m2[ch];
                    v1[ch] += wt * (d1 * d1); v2[ch] += wt * (d2 * d2); cv[ch] += wt * d1 * d2;
                }
            }
        const double c1 = (0.01 * 65535.0) * (0.01 * 65535.0), c2 = (0.03 * 65535.0) * (0.03 * 65535.0);
#pragma unroll
        for (int ch = 0; ch < 3; ch++)
            val += ((2 * m1[ch] * m2[ch] + c1) * (2 * cv[ch] + c2)) / ((m1[ch] * m1[ch] + m2[ch] * m2[ch] + c1) * (v1[ch] + v2[ch] + c2));
        val /= 3.0;
    }
    part[threadIdx.x] = val;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if ((int)threadIdx.x < off) part[threadIdx.x] += part[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) block_sums[blockIdx.x] = part[0];
}

int gmlrt_ssim(gmlrt_ctx* c, const uint8_t* rgba_a, const uint8_t* rgba_b, int width, int height, double* ssim) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    if (!rgba_a || !rgba_b || !ssim) return fail(c, GMLRT_E_BAD_ARG, "null image or result pointer");
    if (width < 11 || height < 11) return fail(c, GMLRT_E_BAD_ARG, "images are too small");        // ssim.go:34-36
    IRenderer* r0 = c->r[0].get();
    API_CU(cudaSetDevice(r0->device()));
    double kern[121], total = 0.0;
    for (int i = 0; i < 11; i++)
        for (int j = 0; j < 11; j++) { const double x = i - 5.0, y = j - 5.0; kern[i * 11 + j] = std::exp(-(x * x + y * y) / (2 * 1.5 * 1.5)); total += kern[i * 11 + j]; }
    for (double& k : kern) k /= total;
    API_CU(cudaMemcpyToSymbolAsync(c_ssim_kernel, kern, sizeof kern, 0, cudaMemcpyHostToDevice, r0->stream()));
    const long long windows = (long long)(width - 11) * (height - 11);
    if (windows == 0) { *ssim = std::nan(""); return GMLRT_OK; }                                  // the reference divides 0 by 0 here
    const size_t bytes = (size_t)width * height * 4;
    const int blocks = (int)((windows + 255) / 256);
    uint8_t* d = nullptr;
    API_CU(cudaMalloc(&d, 2 * bytes + (size_t)blocks * sizeof(double)));
    double* d_sums = reinterpret_cast<double*>(d + 2 * bytes);
    std::vector<double> sums(blocks);
    cudaError_t e = cudaMemcpyAsync(d, rgba_a, bytes, cudaMemcpyHostToDevice, r0->stream());
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + bytes, rgba_b, bytes, cudaMemcpyHostToDevice, r0->stream());
    if (e == cudaSuccess) {
        k_ssim<<<blocks, 256, 0, r0->stream()>>>(reinterpret_cast<const uchar4*>(d), reinterpret_cast<const uchar4*>(d + bytes), width, height, d_sums);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(sums.data(), d_sums, (size_t)blocks * sizeof(double), cudaMemcpyDeviceToHost, r0->stream());
    if (e == cudaSuccess) e = cudaStreamSynchronize(r0->stream());
    cudaFree(d);
    if (e != cudaSuccess) return fail(c, GMLRT_E_CUDA, std::string("gmlrt_ssim: ") + cudaGetErrorString(e));
    double sum = 0.0;
    for (double v : sums) sum += v;
    *ssim = sum / (double)windows;
    return GMLRT_OK;
}

// ------------------------------------------------------------------ SSIM without allocations
// No counterpart in the reference: for callers that must not call cudaMalloc (CUDA graph capture, serving loops).
// The per-block sums of k_ssim go into a file-scope device array instead of a cudaMalloc'ed buffer; the two rasters are
// DEVICE pointers the caller owns (e.g. gmlrt_frame_device).  In host code `g_ssim_scratch` is only a stand-in for the
// symbol, not a device address: its address is looked up once per device with cudaGetSymbolAddress (that device
// current) and that pointer is what the kernel gets.  Every call on a device shares the array and it keeps the last
// call's block sums; nothing has to start at zero because k_ssim overwrites block_sums[0, blocks) and only those are read
// back.  Calls are serialised by a process-wide mutex; calls from different processes on one GPU have their own copy.
static __device__ __align__(256) double g_ssim_scratch[GMLRT_SSIM_NOALLOC_MAX_WINDOWS / 256];       // 1 MiB per GPU
static double      g_ssim_host_sums[GMLRT_SSIM_NOALLOC_MAX_WINDOWS / 256];                          // host side of the read-back
static double*     g_ssim_scratch_ptr[64];                                                          // device address, by CUDA ordinal
static std::mutex  g_ssim_mutex;

int gmlrt_ssim_device_noalloc(gmlrt_ctx* c, const void* d_rgba_a, const void* d_rgba_b, int width, int height, double* ssim) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    if (!d_rgba_a || !d_rgba_b || !ssim) return fail(c, GMLRT_E_BAD_ARG, "null image or result pointer");
    if (width < 11 || height < 11) return fail(c, GMLRT_E_BAD_ARG, "images are too small");        // ssim.go:34-36
    const long long windows = (long long)(width - 11) * (height - 11);
    if (windows > (long long)GMLRT_SSIM_NOALLOC_MAX_WINDOWS)
        return fail(c, GMLRT_E_BAD_ARG, "gmlrt_ssim_device_noalloc: (width-11)*(height-11) = " + std::to_string(windows) +
                                        " windows exceed the static scratch (GMLRT_SSIM_NOALLOC_MAX_WINDOWS = " +
                                        std::to_string((long long)GMLRT_SSIM_NOALLOC_MAX_WINDOWS) + "); use gmlrt_ssim");
    IRenderer* r0 = c->r[0].get();
    const int dev = r0->device();
    if (dev < 0 || dev >= 64) return fail(c, GMLRT_E_BAD_ARG, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_ssim_mutex);
    API_CU(cudaSetDevice(dev));
    if (!g_ssim_scratch_ptr[dev]) {                                                                // set-up, once per device
        double* p = nullptr;
        API_CU(cudaGetSymbolAddress((void**)&p, g_ssim_scratch));
        g_ssim_scratch_ptr[dev] = p;
    }
    double* d_sums = g_ssim_scratch_ptr[dev];
    double kern[121], total = 0.0;
    for (int i = 0; i < 11; i++)
        for (int j = 0; j < 11; j++) { const double x = i - 5.0, y = j - 5.0; kern[i * 11 + j] = std::exp(-(x * x + y * y) / (2 * 1.5 * 1.5)); total += kern[i * 11 + j]; }
    for (double& k : kern) k /= total;
    API_CU(cudaMemcpyToSymbolAsync(c_ssim_kernel, kern, sizeof kern, 0, cudaMemcpyHostToDevice, r0->stream()));
    if (windows == 0) { *ssim = std::nan(""); return GMLRT_OK; }
    const int blocks = (int)((windows + 255) / 256);
    k_ssim<<<blocks, 256, 0, r0->stream()>>>(static_cast<const uchar4*>(d_rgba_a), static_cast<const uchar4*>(d_rgba_b), width, height, d_sums);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(g_ssim_host_sums, d_sums, (size_t)blocks * sizeof(double), cudaMemcpyDeviceToHost, r0->stream());
    if (e == cudaSuccess) e = cudaStreamSynchronize(r0->stream());
    if (e != cudaSuccess) return fail(c, GMLRT_E_CUDA, std::string("gmlrt_ssim_device_noalloc: ") + cudaGetErrorString(e));
    double sum = 0.0;
    for (int i = 0; i < blocks; i++) sum += g_ssim_host_sums[i];                                   // block order, like gmlrt_ssim
    *ssim = sum / (double)windows;
    return GMLRT_OK;
}

void* gmlrt_alloc_host(size_t bytes) {
    void* p = nullptr;
    if (bytes == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}

void gmlrt_free_host(void* p) { if (p) cudaFreeHost(p); }

int gmlrt_download(gmlrt_ctx* c, uint8_t* rgba, size_t rgba_bytes) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    if (!c->have_frame) return fail(c, GMLRT_E_BAD_ARG, "gmlrt_download before a successful gmlrt_trace");
    const size_t need = (size_t)c->scene.width * c->scene.height * 4;
    if (!rgba || rgba_bytes != need) return fail(c, GMLRT_E_BAD_ARG, "rgba buffer must be exactly 4*width*height bytes");
    IRenderer* r0 = c->r[0].get();
    API_CU(cudaSetDevice(r0->device()));
    // A page-locked destination (gmlrt_alloc_host, or anything the caller registered) takes the DMA directly.  A pageable
    // one - image.RGBA.Pix of the reference's caller is Go heap memory (raytracer.go:590) - would make the driver stage
    // the whole copy through its own small bounce buffer, synchronously; instead the frame is cut into chunks that are
    // DMA'd into this context's page-locked staging buffer and copied out by a few host threads while later chunks are
    // still in flight.
    cudaPointerAttributes attr;
    const bool locked = cudaPointerGetAttributes(&attr, rgba) == cudaSuccess && (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    if (locked) {
        API_CU(cudaMemcpyAsync(rgba, r0->frame(), need, cudaMemcpyDeviceToHost, r0->stream()));
        API_CU(cudaStreamSynchronize(r0->stream()));
        return GMLRT_OK;
    }
    if (c->pinned_bytes < need) {
        if (c->pinned) cudaFreeHost(c->pinned);
        c->pinned = nullptr; c->pinned_bytes = 0;
        API_CU(cudaHostAlloc((void**)&c->pinned, need, cudaHostAllocDefault));
        c->pinned_bytes = need;
    }
    // 4 MB DMA chunks (32 one-megabyte copies cost more in per-copy overhead than they gained in overlap); every chunk is copied
    // out by ALL the host threads, an eighth each, as soon as its event fires: the tail after the last DMA is one eighth of a chunk
    const size_t chunk = (size_t)4 << 20;
    const int n_chunks = (int)((need + chunk - 1) / chunk);
    while ((int)c->ev_chunk.size() < n_chunks) {
        cudaEvent_t e;
        API_CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c->ev_chunk.push_back(e);
    }
    for (int k = 0; k < n_chunks; k++) {
        const size_t off = (size_t)k * chunk, len = std::min(chunk, need - off);
        API_CU(cudaMemcpyAsync(c->pinned + off, r0->frame() + off, len, cudaMemcpyDeviceToHost, r0->stream()));
        API_CU(cudaEventRecord(c->ev_chunk[k], r0->stream()));
    }
    if (!c->pool) c->pool.reset(new HostPool((int)std::max(1u, std::min(8u, std::thread::hardware_concurrency()))));
    const int n_threads = c->pool->size();
    std::vector<cudaError_t> terr(n_threads, cudaSuccess);
    const std::function<void(int)> drain = [&](int t) {
        for (int k = 0; k < n_chunks; k++) {
            const cudaError_t e = cudaEventSynchronize(c->ev_chunk[k]);
            if (e != cudaSuccess) { terr[t] = e; return; }
            const size_t off = (size_t)k * chunk, len = std::min(chunk, need - off);
            const size_t a = len * (size_t)t / (size_t)n_threads, b = len * (size_t)(t + 1) / (size_t)n_threads;
            std::memcpy(rgba + off + a, c->pinned + off + a, b - a);
        }
    };
    c->pool->run(drain);
    for (cudaError_t e : terr) if (e != cudaSuccess) return fail(c, GMLRT_E_CUDA, std::string("gmlrt_download: ") + cudaGetErrorString(e));
    return GMLRT_OK;
}

// One process per GPU, gathered on the HOST: every process copies the rows it rendered - its bands of the partition -
// into the same full-size raster (shared memory the processes map and page-lock), each over its own PCIe link, instead of
// funnelling the whole frame through one GPU and one link.  Rows of other parts are left untouched.
int gmlrt_download_rows(gmlrt_ctx* c, uint8_t* rgba, size_t rgba_bytes) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    if (!c->have_frame) return fail(c, GMLRT_E_BAD_ARG, "gmlrt_download_rows before a successful gmlrt_trace");
    if (c->remote_frame) return fail(c, GMLRT_E_BAD_ARG, "this context stores its rows into another process's frame (gmlrt_frame_ipc_attach): nothing to download here");
    const size_t row = (size_t)c->scene.width * 4, need = row * c->scene.height;
    if (!rgba || rgba_bytes != need) return fail(c, GMLRT_E_BAD_ARG, "rgba buffer must be exactly 4*width*height bytes");
    if (c->n_parts == 1) return gmlrt_download(c, rgba, rgba_bytes);
    IRenderer* r0 = c->r[0].get();
    API_CU(cudaSetDevice(r0->device()));
    const int H = c->scene.height, br = c->band_rows, np = c->n_parts, part = c->part;
    const int full = H / br, rest = H % br;                                     // full bands + one partial band (index `full`), as RowMap::count
    const int mine = full / np + (full % np > part ? 1 : 0);
    const size_t first = (size_t)part * br * row, pitch = (size_t)np * br * row;
    if (mine > 0) API_CU(cudaMemcpy2DAsync(rgba + first, pitch, r0->frame() + first, pitch, (size_t)br * row, (size_t)mine, cudaMemcpyDeviceToHost, r0->stream()));
    if (rest && full % np == part) {
        const size_t off = (size_t)full * br * row;
        API_CU(cudaMemcpyAsync(rgba + off, r0->frame() + off, (size_t)rest * row, cudaMemcpyDeviceToHost, r0->stream()));
    }
    API_CU(cudaStreamSynchronize(r0->stream()));
    return GMLRT_OK;
}

int gmlrt_primary_hits(gmlrt_ctx* c, int32_t* ids, size_t n_ids) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    if (!c->have_frame || !c->counting || !c->r[0]->primary_ids())
        return fail(c, GMLRT_E_BAD_ARG, "primary hit ids need gmlrt_set_counting(ctx, 1) before gmlrt_trace");
    const size_t need = (size_t)c->scene.width * c->scene.height * 4;
    if (!ids || n_ids != need) return fail(c, GMLRT_E_BAD_ARG, "ids must hold 4*width*height entries");
    IRenderer* r0 = c->r[0].get();
    API_CU(cudaSetDevice(r0->device()));
    API_CU(cudaMemcpyAsync(ids, r0->primary_ids(), need * sizeof(int32_t), cudaMemcpyDeviceToHost, r0->stream()));
    API_CU(cudaStreamSynchronize(r0->stream()));
    return GMLRT_OK;
}

int gmlrt_render(gmlrt_ctx* c, const void* blob, size_t blob_bytes, uint8_t* rgba, size_t rgba_bytes, gmlrt_stats* stats) {
    if (!c) return fail(nullptr, GMLRT_E_BAD_ARG, "ctx is null");
    using clk = std::chrono::steady_clock;
    auto t0 = clk::now();
    int rc = gmlrt_upload(c, blob, blob_bytes);
    if (rc != GMLRT_OK) return rc;
    if (!rgba || rgba_bytes != (size_t)c->scene.width * c->scene.height * 4)
        return fail(c, GMLRT_E_BAD_ARG, "rgba buffer must be exactly 4*width*height bytes");
    auto t1 = clk::now();
    gmlrt_stats st;
    rc = gmlrt_trace(c, &st);
    if (rc != GMLRT_OK) return rc;
    auto t2 = clk::now();
    rc = gmlrt_download(c, rgba, rgba_bytes);
    if (rc != GMLRT_OK) return rc;
    auto t3 = clk::now();
    if (stats) {
        *stats = st;
        auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        stats->ms_h2d = ms(t0, t1); stats->ms_d2h = ms(t2, t3); stats->ms_total = ms(t0, t3);
    }
    return GMLRT_OK;
}

}  // extern "C"
