// renderer.h - precision-independent interface between the C ABI (gmlrt_api.cu) and the two
// instantiations of the wavefront tracer (render_f32.cu / render_f64.cu).  One object per device.
#pragma once
#include <cstdint>
#include <string>

#include <cuda_runtime.h>

#include "../../include/gmlrt.h"
#include "host_scene.h"

namespace gmlrt {

// kernel classes for the profiling mode (gmlrt_set_profiling)
enum KernelClass { KC_GEN = 0, KC_CLOSEST_PRIMARY = 1, KC_CLOSEST = 2, KC_SHADOW = 3, KC_SHADE = 4, KC_RESOLVE = 5, KC_PIXELS = 6, KC_EXACT = 7, KC_SORT = 8, KC_COUNT = 9 };

struct TraceTotals {                   // summed over the tiles a device rendered
    unsigned long long rays_primary = 0, rays_reflect = 0, rays_refract = 0, rays_shadow = 0, shadow_occluded = 0, rays_hit = 0;
    unsigned long long shadow_tests = 0, shadow_flops = 0;   // credited (reference scan order), counting mode
    unsigned long long cull_by_family[3] = {0, 0, 0}, packet_tests = 0;   // executed per-ray cull tests (GENERAL, PIVOT, AXIS) and per-warp packet tests
    unsigned long long rays_hit_primary = 0, list_tests = 0;
    unsigned long long cull_tests = 0, exact_tests = 0, exact_flops = 0, exact_tests_shadow = 0;   // executed: first-phase (all pairs) and second-phase tests
    double ms_class[KC_COUNT] = {0, 0, 0, 0, 0, 0, 0, 0, 0};         // profiling mode: device time per kernel class
    unsigned long long launches_class[KC_COUNT] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    unsigned long long launches = 0;
    int error = 0, overflow = 0;
};

// Which image rows a device renders.  The image is cut into bands of `band_rows` rows dealt round-robin to `n_parts`
// devices; device `part` sees its own bands as one contiguous run of VIRTUAL rows, so a frame is a few large launches
// per device however fine the interleave is.  n_parts == 1 is the identity.
struct RowMap {
    int band_rows = 1, n_parts = 1, part = 0;
    __host__ __device__ int actual(int v) const {
        return n_parts == 1 ? v : ((v / band_rows) * n_parts + part) * band_rows + v % band_rows;
    }
    // number of virtual rows of this part for an image of H rows
    int count(int H) const {
        if (n_parts == 1) return H;
        const int full = H / band_rows, rest = H % band_rows;      // full bands + one partial band (index `full`)
        int rows = (full / n_parts + (full % n_parts > part ? 1 : 0)) * band_rows;
        if (rest && full % n_parts == part) rows += rest;
        return rows;
    }
};

class IRenderer {
public:
    virtual ~IRenderer() {}
    // copy the converted scene to this device; (re)allocate queues for rows_per_tile rows
    virtual int upload(const HostScene& S, std::string& err) = 0;
    // enqueue the kernels for virtual rows [v0, v1) of `map` on this device's stream (no host sync)
    virtual int trace_rows(int v0, int v1, const RowMap& map, bool counting, std::string& err) = 0;
    // where the level-0 k_resolve stores its pixels: this device's own frame (default) or a peer-mapped frame of the gathering device
    virtual void set_frame_target(uint8_t* target) = 0;
    // how many devices share the frame (before upload): a device that renders 1 / n of the rays sorts them into coarser classes
    virtual void set_share(int n_parts) = 0;
    // after the stream is synchronised: fetch counters accumulated since reset_totals()
    virtual int fetch_totals(TraceTotals& out, std::string& err) = 0;
    virtual void reset_totals() = 0;
    virtual void set_profiling(bool on) = 0;   // record an event before every launch; per-class times land in TraceTotals
    virtual uint8_t* frame() = 0;          // device RGBA8 frame, full size (only own rows valid)
    virtual int32_t* primary_ids() = 0;    // device [H][W][4] scan-order ids (counting mode), may be null
    virtual cudaStream_t stream() = 0;
    virtual int device() const = 0;
    virtual int max_rows_per_tile() const = 0;
    // after a candidate-queue overflow: halve the tile height (keeping the buffers) and re-plan; false = cannot shrink further
    virtual bool shrink_tiles(const HostScene& S, std::string& err) = 0;
};

IRenderer* make_renderer_f32(int device);
IRenderer* make_renderer_f64(int device);

}  // namespace gmlrt
