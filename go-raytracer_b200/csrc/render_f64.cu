// Parity mode: double arithmetic in the reference's operation order.  This translation unit
// MUST be built with -fmad=false so that no multiply-add is contracted (Go on amd64 never fuses).
#define GMLRT_PARITY_TU 1
#include "render_impl.cuh"
namespace gmlrt {
IRenderer* make_renderer_f64(int device) {
    auto* r = new Renderer<double>(device);
    std::string err;
    if (r->init(err) != GMLRT_OK) { std::fprintf(stderr, "gmlrt: %s\n", err.c_str()); delete r; return nullptr; }
    return r;
}
}  // namespace gmlrt
