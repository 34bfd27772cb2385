// Throughput mode: float arithmetic, FMA contraction allowed (default -fmad=true).
#include "render_impl.cuh"
namespace gmlrt {
IRenderer* make_renderer_f32(int device) {
    auto* r = new Renderer<float>(device);
    std::string err;
    if (r->init(err) != GMLRT_OK) { std::fprintf(stderr, "gmlrt: %s\n", err.c_str()); delete r; return nullptr; }
    return r;
}
}  // namespace gmlrt
