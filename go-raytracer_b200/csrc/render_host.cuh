// render_host.cuh - the host side of the renderer: one Renderer<R> per device (scene upload and conversion into device
// arrays, cull tables and class lists, the tile plan, and trace_rows: the launch sequence of a frame).  Included by
// render_impl.cuh after the kernels, inside namespace gmlrt; compiled twice like them (render_f32.cu, render_f64.cu).
#pragma once

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

