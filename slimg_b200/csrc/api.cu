// api.cu -- the C ABI of libslimg_b200.so (include/slimg_b200.h): argument validation with the
// reference's error categories, per-device plan cache (filter taps + tiling live in HBM once per
// geometry), launch glue, and the multi-GPU batch scheduler that stands where the reference's
// rayon `par_iter` over files does (cli/src/commands/mod.rs:64-72, crop.rs:112, extend.rs:114,
// convert.rs:60).
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "host.h"
#include "kernels.cuh"

namespace slimg {

static std::atomic<uint64_t> g_launches{0};
void count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
static std::atomic<uint64_t> g_path_launches[8];
static thread_local int t_last_path = SLIMG_PATH_NONE;
void note_path(int path) {
    t_last_path = path;
    if (path >= 0 && path < 8) g_path_launches[path].fetch_add(1, std::memory_order_relaxed);
}

static slimg_status cuda_fail(cudaError_t e, const char *what) {
    return fail(SLIMG_ERR_IMAGE, std::string("CUDA: ") + what + ": " + cudaGetErrorString(e));
}
#define CU(call)                                                     \
    do {                                                             \
        cudaError_t e__ = (call);                                    \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call);        \
    } while (0)

static uint32_t pack_fill(const uint8_t f[4]) {
    return uint32_t(f[0]) | (uint32_t(f[1]) << 8) | (uint32_t(f[2]) << 16) | (uint32_t(f[3]) << 24);
}

// ---- resize plans: device-resident tap tables + tiling, cached per (device, geometry) -------

// All tables of a plan live in one or two device slabs: a plan costs one cudaMalloc instead of ~30, and dropping it is
// one cudaFree (mixed-size corpora build a plan per geometry).
//
// Slabs are recycled, not freed: cudaFree synchronises the whole device, and a corpus in which every image has its own
// geometry evicts a plan per image (measured on the host-fed bench, 32 unique 12 MP geometries per step: 3.3 Gpx/s with a
// cudaFree per eviction).  A dropped plan's slabs go to a per-device graveyard; kernels that were enqueued with the plan
// may still be running, so the graveyard is turned into the free list only behind ONE cudaDeviceSynchronize per
// kGraveyardSlabs retired slabs.
constexpr size_t kGraveyardSlabs = 16;
struct SlabPool {
    std::mutex mu;
    std::vector<std::pair<void *, size_t>> free_list, graveyard;
    uint64_t retired = 0, reaped = 0;       // slabs ever pushed to / taken from the graveyard: its front is slab number `reaped`
};
static SlabPool &slab_pool(int dev) {
    // never destroyed: plans that live in other statics retire their slabs here while the process exits
    static SlabPool *pools = new SlabPool[64];
    return pools[dev >= 0 && dev < 64 ? dev : 0];
}
struct PlanArena {
    std::vector<std::pair<void *, size_t>> slabs;
    size_t used = 0, cap = 0;
    int device = 0;
    PlanArena() { if (cudaGetDevice(&device) != cudaSuccess) { cudaGetLastError(); device = 0; } }
    ~PlanArena() {
        if (slabs.empty()) return;
        SlabPool &pool = slab_pool(device);
        std::lock_guard<std::mutex> lk(pool.mu);
        for (auto &s : slabs) { pool.graveyard.push_back(s); pool.retired++; }
    }
    static void *take(SlabPool &pool, size_t bytes, size_t &got) {       // pool.mu held
        for (size_t i = 0; i < pool.free_list.size(); i++)
            if (pool.free_list[i].second >= bytes && pool.free_list[i].second <= 2 * bytes) {
                void *s = pool.free_list[i].first;
                got = pool.free_list[i].second;
                pool.free_list.erase(pool.free_list.begin() + i);
                return s;
            }
        return nullptr;
    }
    cudaError_t alloc(size_t bytes, void **p) {
        const size_t need = (bytes + 255) & ~size_t(255);
        if (slabs.empty() || used + need > cap) {
            size_t slab = std::max<size_t>(need, size_t(4) << 20);
            SlabPool &pool = slab_pool(device);
            void *s = nullptr;
            uint64_t reap = 0;                         // slabs retired so far: only those are safe behind the synchronise below
            {
                std::lock_guard<std::mutex> lk(pool.mu);
                s = take(pool, slab, slab);
                if (!s && pool.graveyard.size() >= kGraveyardSlabs) reap = pool.retired;
            }
            if (reap) {
                // every kernel that was enqueued with one of these plans was enqueued before its plan was dropped, i.e. before
                // this point: after the synchronise nothing can touch them (slabs retired meanwhile stay in the graveyard)
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) return e;
                std::lock_guard<std::mutex> lk(pool.mu);
                size_t n = 0;                          // another worker may have reaped part of them meanwhile
                while (pool.reaped + n < reap && n < pool.graveyard.size()) pool.free_list.push_back(pool.graveyard[n++]);
                pool.graveyard.erase(pool.graveyard.begin(), pool.graveyard.begin() + n);
                pool.reaped += n;
                s = take(pool, slab, slab);
            }
            if (!s) {
                cudaError_t e = cudaMalloc(&s, slab);
                if (e != cudaSuccess) return e;
            }
            slabs.push_back({s, slab}); used = 0; cap = slab;
        }
        *p = static_cast<uint8_t *>(slabs.back().first) + used;
        used += need;
        return cudaSuccess;
    }
};
static thread_local PlanArena *t_arena = nullptr;      // the arena of the plan being built on this thread

struct DevArray {
    void *p = nullptr;
    template <class T> cudaError_t upload(const std::vector<T> &v) {
        size_t bytes = (v.empty() ? 1 : v.size()) * sizeof(T);
        cudaError_t e = t_arena->alloc(bytes, &p);
        if (e != cudaSuccess) return e;
        if (!v.empty()) e = cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, cudaStreamPerThread);
        return e;
    }
};

struct ResizePlan {
    PlanArena arena;
    DevArray xl, xc, xw, xf, yl, yc, yw, yf, tx0, tspan;
    ResizeParams params{};   // src/dst/strides filled per launch
    // rows / columns of the source any kernel of this plan reads with a non-zero weight (what a host-fed call uploads)
    uint32_t win_r0 = 0, win_r1 = 0, win_c0 = 0, win_c1 = 0;
    // fast path (k2_fast.cu), when the geometry qualifies
    bool fast = false;
    DevArray f_a0, f_nch, f_boff, f_bfrag, f_fillx, f_vrow, f_vleft, f_vright, f_filly, f_rowc, f_sa0, f_sbytes;
    FastResizeParams fparams{};
    // V-first integer tensor pipeline (k2_vfirst.cu), preferred when the geometry qualifies
    bool vfirst = false;
    DevArray v_recs, v_rowc, v_col0, v_src0, v_nblk, v_bytes, v_upw, v_nunits, v_kb0, v_tnch, v_b2, v_cx, v_cxi;
    VfParams vparams{};
    // integer tile kernel (k2_tile.cu): ratios below ~3.5x and upscales
    bool tile = false;
    DevArray l_recs, l_rowc, l_y0, l_rows, l_src0, l_nblk, l_kb0, l_b2, l_cx;
    TileParams lparams{};
    // tcgen05 pipeline (k2_tc.cu), preferred over the mma.sync pipeline when the geometry qualifies
    bool tc = false;
    DevArray t_crecs, t_c0, t_c1, t_bias, t_col0, t_src0, t_kb0, t_b2, t_cx;
    TcParams tparams{};
};

struct GeomKey {
    uint32_t v[24];
    bool operator<(const GeomKey &o) const { return std::memcmp(v, o.v, sizeof v) < 0; }
};

static GeomKey make_key(const PipeGeom &g) {
    GeomKey k{};
    const uint32_t vals[] = {g.sw, g.sh, g.rx, g.ry, g.rw, g.rh, g.vw, g.vh, g.vox, g.voy, pack_fill(g.fill_pre),
                             uint32_t(g.has_resize), g.Rw, g.Rh, g.px, g.py, g.pw, g.ph, g.fw, g.fh, g.fox, g.foy,
                             pack_fill(g.fill_post)};
    for (size_t i = 0; i < sizeof(vals) / sizeof(vals[0]); i++) k.v[i] = vals[i];
    return k;
}

constexpr size_t kPlanCacheEntries = 64;     // per device; least recently used geometry goes first
struct DeviceCtx {
    std::mutex mu;
    struct Entry { std::shared_ptr<ResizePlan> plan; uint64_t stamp; };
    std::map<GeomKey, Entry> plans;
    uint64_t clock = 0;
};

static DeviceCtx &device_ctx(int dev) {
    static std::mutex mu;
    static std::map<int, std::unique_ptr<DeviceCtx>> ctxs;
    std::lock_guard<std::mutex> lk(mu);
    auto &c = ctxs[dev];
    if (!c) c.reset(new DeviceCtx());
    return *c;
}

// source rows / columns behind the kept output rectangle: every tap with a non-zero weight lies inside
static void window_from_taps(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, uint32_t &wr0, uint32_t &wr1,
                             uint32_t &wc0, uint32_t &wc1) {
    int64_t r0 = INT64_MAX, r1 = -1, c0 = INT64_MAX, c1 = -1;
    for (uint32_t i = 0; i < g.ph; i++) {
        const uint32_t o = g.py + i;
        if (ty.count[o] == 0) continue;
        r0 = std::min<int64_t>(r0, ty.left[o]); r1 = std::max<int64_t>(r1, int64_t(ty.left[o]) + ty.count[o]);
    }
    for (uint32_t j = 0; j < g.pw; j++) {
        const uint32_t o = g.px + j;
        if (tx.count[o] == 0) continue;
        c0 = std::min<int64_t>(c0, tx.left[o]); c1 = std::max<int64_t>(c1, int64_t(tx.left[o]) + tx.count[o]);
    }
    wr0 = wr1 = wc0 = wc1 = 0;
    if (r1 > r0 && c1 > c0) {
        wr0 = g.ry + uint32_t(r0); wr1 = g.ry + uint32_t(r1);
        wc0 = g.rx + uint32_t(c0); wc1 = g.rx + uint32_t(c1);
    }
}

// SLIMG_B200_RESIZE_PATH: which K2 kernel a downscale should prefer.  Unset = the production order (V-first integer
// pipeline, then H-first for bands the V-first ring cannot hold, then the f32 kernels); "tc" / "hfirst" / "generic"
// force one of the others where the geometry qualifies (tests run every path; benches A/B them).  Part of the plan key.
static std::string requested_path() {
    const char *path_env = std::getenv("SLIMG_B200_RESIZE_PATH");
    return path_env && *path_env ? path_env : "vfirst";
}
static uint32_t path_code(const std::string &p) {
    return p == "tc" ? 1u : p == "hfirst" ? 2u : p == "generic" ? 3u : p == "tile" ? 4u : p == "notile" ? 5u : 0u;
}

static slimg_status build_resize_plan(const PipeGeom &g, const std::string &path, std::shared_ptr<ResizePlan> &out) {
    auto plan = std::make_shared<ResizePlan>();
    struct ArenaScope { ArenaScope(PlanArena *a) { t_arena = a; } ~ArenaScope() { t_arena = nullptr; } } scope(&plan->arena);
    AxisTaps tx, ty;
    build_axis_taps(g.vw, g.Rw, g.vox, g.rw, tx);
    build_axis_taps(g.vh, g.Rh, g.voy, g.rh, ty);

    // tiling of the final canvas for the generic kernel
    int tile_w = g.vw >= g.Rw ? 32 : 64;
    int tile_h = g.vh >= g.Rh ? 8 : 16;
    // upscales with few taps run the register-weight kernel (k2_resize.cu): its per-thread set-up is amortised over a
    // taller tile
    const bool few_taps = tx.max_taps <= 8 && ty.max_taps <= 8;
    if (few_taps && g.Rh > g.vh && g.Rw > g.vw) tile_h = 64;
    std::vector<int32_t> tile_x0, tile_span;
    int span_max = 0;
    for (;;) {
        const int tiles_x = int((g.fw + tile_w - 1) / tile_w);
        tile_x0.assign(tiles_x, 0);
        tile_span.assign(tiles_x, 0);
        span_max = 0;
        for (int t = 0; t < tiles_x; t++) {
            const int64_t fx0 = int64_t(t) * tile_w, fx1 = std::min<int64_t>(fx0 + tile_w, g.fw);
            const int64_t ix0 = std::max<int64_t>(fx0, g.fox), ix1 = std::min<int64_t>(fx1, int64_t(g.fox) + g.pw);
            int lo = INT32_MAX, hi = 0;
            for (int64_t fx = ix0; fx < ix1; fx++) {
                const int64_t ox = fx - g.fox + g.px;
                if (tx.count[ox] == 0) continue;
                lo = std::min(lo, tx.left[ox]);
                hi = std::max(hi, tx.left[ox] + tx.count[ox]);
            }
            if (lo < hi) { tile_x0[t] = lo; tile_span[t] = hi - lo; span_max = std::max(span_max, hi - lo); }
        }
        if (size_t(tile_h) * size_t(std::max(span_max, 1)) * 16 <= 192 * 1024) break;
        if (tile_h > 1) tile_h /= 2;
        else if (tile_w > 1) tile_w /= 2;
        else return fail(SLIMG_ERR_RESIZE, "filter footprint of this resize exceeds shared memory");
    }

    window_from_taps(g, tx, ty, plan->win_r0, plan->win_r1, plan->win_c0, plan->win_c1);
    cudaError_t e;
    if ((e = plan->xl.upload(tx.left)) || (e = plan->xc.upload(tx.count)) || (e = plan->xw.upload(tx.w)) ||
        (e = plan->xf.upload(tx.fillw)) || (e = plan->yl.upload(ty.left)) || (e = plan->yc.upload(ty.count)) ||
        (e = plan->yw.upload(ty.w)) || (e = plan->yf.upload(ty.fillw)) || (e = plan->tx0.upload(tile_x0)) ||
        (e = plan->tspan.upload(tile_span)))
        return cuda_fail(e, "uploading resize plan");

    ResizeParams &p = plan->params;
    p.sw = g.sw;
    p.rx = g.rx; p.ry = g.ry; p.rw = g.rw; p.rh = g.rh;
    p.ax = AxisDev{(const int32_t *)plan->xl.p, (const int32_t *)plan->xc.p, (const float *)plan->xw.p,
                   (const float *)plan->xf.p, tx.max_taps};
    p.ay = AxisDev{(const int32_t *)plan->yl.p, (const int32_t *)plan->yc.p, (const float *)plan->yw.p,
                   (const float *)plan->yf.p, ty.max_taps};
    for (int c = 0; c < 4; c++) p.fill_pre[c] = float(g.fill_pre[c]);
    p.px = g.px; p.py = g.py; p.pw = g.pw; p.ph = g.ph;
    p.fw = g.fw; p.fh = g.fh; p.fox = g.fox; p.foy = g.foy;
    p.fill_post = pack_fill(g.fill_post);
    p.tile_w = tile_w; p.tile_h = tile_h; p.span_max = span_max;
    p.few_rows_max = 0; p.few_k = 1;
    if (few_taps) {       // most source rows any tile row-block needs (the few-tap kernel stages them as f32)
        for (int64_t fy0 = 0; fy0 < int64_t(g.fh); fy0 += tile_h) {
            const int64_t iy0 = std::max<int64_t>(fy0, g.foy), iy1 = std::min<int64_t>(std::min<int64_t>(fy0 + tile_h, g.fh), int64_t(g.foy) + g.ph);
            int lo = INT32_MAX, hi = 0;
            for (int64_t fy = iy0; fy < iy1; fy++) {
                const int64_t oy = fy - g.foy + g.py;
                if (ty.count[oy] == 0) continue;
                lo = std::min(lo, ty.left[oy]);
                hi = std::max(hi, ty.left[oy] + ty.count[oy]);
            }
            if (lo < hi) p.few_rows_max = std::max(p.few_rows_max, hi - lo);
        }
        p.few_rows_max = std::max(p.few_rows_max, 1);
        // four adjacent canvas columns (aligned to 4) may share one window of max_taps + 2 intermediate values?
        // (measured on B200, C4b: the 4-column variant needs 103 registers, halves the occupancy and runs at 171 us against
        // 94 us for one column per thread -- kept behind SLIMG_B200_FEW_K4 for the next round)
        const char *fk_env = std::getenv("SLIMG_B200_FEW_K");
        const int fk = fk_env ? std::atoi(fk_env) : (std::getenv("SLIMG_B200_FEW_K4") ? 4 : 1);
        p.few_k = ((fk == 2 || fk == 4) && tile_w % fk == 0 && g.Rw > g.vw) ? fk : 1;
        // an aligned group of few_k columns must fit one window of max_taps + few_k / 2 intermediate values
        for (int64_t fxg = 0; fxg < int64_t(g.fw) && p.few_k > 1; fxg += p.few_k) {
            int lo = INT32_MAX, hi = INT32_MIN;
            for (int64_t fx = std::max<int64_t>(fxg, g.fox); fx < std::min<int64_t>(std::min<int64_t>(fxg + p.few_k, g.fw), int64_t(g.fox) + g.pw); fx++) {
                const int64_t ox = fx - g.fox + g.px;
                if (tx.count[ox] == 0) continue;
                lo = std::min(lo, tx.left[ox]);
                hi = std::max(hi, tx.left[ox]);
            }
            if (lo <= hi && hi - lo > p.few_k / 2) p.few_k = 1;
        }
    }
    p.tile_x0 = (const int32_t *)plan->tx0.p;
    p.tile_span = (const int32_t *)plan->tspan.p;
    if (path == "vfirst" || path == "tile") {        // "notile": the banded pipelines at small ratios too (A/B)
        TilePlanHost lp;
        build_tileplan(g, tx, ty, path == "tile", lp);
        if (std::getenv("SLIMG_B200_DEBUG"))
            std::fprintf(stderr, "[slimg_b200] tileplan ok=%d tile=%ux%u blocks=%ux%u rows_max=%u nch1=%d nch2=%d pitch=%u pitch_v=%u smem=%u\n",
                         int(lp.ok), lp.tw, lp.th, lp.n_colblk, lp.n_rowblk, lp.rows_max, lp.nch1, lp.nch2, lp.pitch, lp.pitch_v, lp.smem_bytes);
        if (lp.ok) {
            if ((e = plan->l_recs.upload(lp.recs)) || (e = plan->l_rowc.upload(lp.row_computed)) || (e = plan->l_y0.upload(lp.rb_y0)) ||
                (e = plan->l_rows.upload(lp.rb_rows)) || (e = plan->l_src0.upload(lp.cb_src0)) || (e = plan->l_nblk.upload(lp.cb_nblk)) ||
                (e = plan->l_kb0.upload(lp.tile_kb0)) || (e = plan->l_b2.upload(lp.b2)) || (e = plan->l_cx.upload(lp.cxi)))
                return cuda_fail(e, "uploading tile resize plan");
            TileParams &l = plan->lparams;
            l.sw = g.sw; l.sh = g.sh; l.ry = g.ry; l.rh = g.rh; l.pw = g.pw; l.ph = g.ph;
            l.tw = lp.tw; l.th = lp.th; l.nch1 = lp.nch1; l.nch2 = lp.nch2; l.sh1 = lp.sh1; l.rec_words = lp.rec_words;
            l.recs = (const uint32_t *)plan->l_recs.p; l.row_computed = (const uint8_t *)plan->l_rowc.p;
            l.n_rowblk = lp.n_rowblk; l.n_colblk = lp.n_colblk;
            l.rb_y0 = (const int32_t *)plan->l_y0.p; l.rb_rows = (const int32_t *)plan->l_rows.p;
            l.cb_src0 = (const int32_t *)plan->l_src0.p; l.cb_nblk = (const int32_t *)plan->l_nblk.p;
            l.tile_kb0 = (const int32_t *)plan->l_kb0.p; l.b2 = (const uint32_t *)plan->l_b2.p; l.cxi = (const int32_t *)plan->l_cx.p;
            l.sh2 = lp.sh2; l.packed = lp.packed ? 1u : 0u;
            l.pitch = lp.pitch; l.pitch_v = lp.pitch_v; l.rows_max = lp.rows_max; l.smem_bytes = lp.smem_bytes;
            l.fw = g.fw; l.fh = g.fh; l.fox = g.fox; l.foy = g.foy;
            l.fill_post = pack_fill(g.fill_post); l.fill_pre_packed = pack_fill(g.fill_pre);
            l.need_fill = (lp.need_fill ? 1u : 0u) | (lp.rows_pure_fill ? 2u : 0u);
            plan->tile = true;
        }
    }
    if (path == "tc") {
        TcPlanHost tp;
        build_tcplan(g, tx, ty, tp);
        if (std::getenv("SLIMG_B200_DEBUG"))
            std::fprintf(stderr, "[slimg_b200] tcplan ok=%d strips=%u cols=%d panels=%u nbuf=%u slots=%u blocks=%u chunks=%u nb_max=%u nch2=%d smem=%u\n",
                         int(tp.ok), tp.n_strips, tp.ok ? tp.strip_col0[1] : 0, tp.n_panels, tp.nbuf, tp.slots, tp.n_blocks, tp.n_chunks,
                         tp.nb_max, tp.nch2, tp.smem_bytes);
        if (tp.ok) {
            if ((e = plan->t_crecs.upload(tp.crecs)) || (e = plan->t_c0.upload(tp.blk_c0)) || (e = plan->t_c1.upload(tp.blk_c1)) ||
                (e = plan->t_bias.upload(tp.bias)) || (e = plan->t_col0.upload(tp.strip_col0)) || (e = plan->t_src0.upload(tp.strip_src0)) ||
                (e = plan->t_kb0.upload(tp.tile_kb0)) || (e = plan->t_b2.upload(tp.b2)) || (e = plan->t_cx.upload(tp.cx)))
                return cuda_fail(e, "uploading tcgen05 resize plan");
            TcParams &t = plan->tparams;
            t.sw = g.sw; t.sh = g.sh; t.ry = g.ry; t.pw = g.pw; t.ph = g.ph;
            t.nch2 = tp.nch2; t.sh1 = tp.sh1;
            t.n_blocks = tp.n_blocks; t.n_chunks = tp.n_chunks; t.crec_words = tp.crec_words;
            t.crecs = (const uint32_t *)plan->t_crecs.p; t.blk_c0 = (const int32_t *)plan->t_c0.p; t.blk_c1 = (const int32_t *)plan->t_c1.p;
            t.bias = (const int32_t *)plan->t_bias.p;
            t.n_strips = tp.n_strips; t.n_panels = tp.n_panels;
            t.strip_col0 = (const int32_t *)plan->t_col0.p; t.strip_src0 = (const int32_t *)plan->t_src0.p;
            t.tile_kb0 = (const int32_t *)plan->t_kb0.p; t.b2 = (const uint32_t *)plan->t_b2.p; t.cx = (const float *)plan->t_cx.p;
            for (int l = 0; l < 2; l++)
                for (int d = 0; d < 3; d++) t.s2[l][d] = tp.s2[l][d];
            t.slots = tp.slots; t.slot_bytes = tp.slot_bytes; t.pitch_v = tp.pitch_v; t.smem_bytes = tp.smem_bytes; t.nbuf = tp.nbuf; t.vbufs = tp.vbufs;
            { const char *pf = std::getenv("SLIMG_B200_TC_PREFETCH"); t.pf_chunks = pf ? uint32_t(std::atoi(pf)) : 6u; }
            t.fw = g.fw; t.fh = g.fh; t.fox = g.fox; t.foy = g.foy;
            t.fill_post = pack_fill(g.fill_post); t.fill_pre_packed = pack_fill(g.fill_pre);
            t.n_bands = 1; t.need_fill = tp.need_fill ? 1u : 0u;
            const char *dbg_env = std::getenv("SLIMG_B200_VF_DEBUG");
            t.debug = dbg_env ? uint32_t(std::atoi(dbg_env)) : 0u;
            plan->tc = true;
        }
    }
    if (((path == "vfirst" || path == "tile") && !plan->tile) || path == "notile" || (path == "tc" && !plan->tc)) {
        const char *dig_env = std::getenv("SLIMG_B200_DIGITS");
        const int ndig = dig_env && dig_env[0] == '3' ? 3 : 2;
        VPlanHost vp;
        build_vplan(g, tx, ty, ndig, vp);
        if (std::getenv("SLIMG_B200_DEBUG"))
            std::fprintf(stderr, "[slimg_b200] vplan ok=%d strips=%u cols=%d unit=%d upw=%d nunits=%d early=%d vbufs=%u ring=%u pitch=%u pitch_v=%u smem=%u nch1=%d nch2=%d recs=%u\n",
                         int(vp.ok), vp.n_strips, vp.ok ? vp.strip_col0[1] : 0, vp.unit, vp.ok ? vp.strip_upw[0] : 0,
                         vp.ok ? vp.strip_nunits[0] : 0, int(vp.early_go), vp.vbufs, vp.ring_rows, vp.pitch, vp.pitch_v, vp.smem_bytes,
                         vp.nch1, vp.nch2, vp.n_recs);
        if (vp.ok) {
            if ((e = plan->v_recs.upload(vp.recs)) || (e = plan->v_rowc.upload(vp.row_computed)) || (e = plan->v_col0.upload(vp.strip_col0)) ||
                (e = plan->v_src0.upload(vp.strip_src0)) || (e = plan->v_nblk.upload(vp.strip_nblk)) ||
                (e = plan->v_bytes.upload(vp.strip_bytes)) || (e = plan->v_upw.upload(vp.strip_upw)) || (e = plan->v_nunits.upload(vp.strip_nunits)) || (e = plan->v_kb0.upload(vp.tile_kb0)) ||
                (e = plan->v_tnch.upload(vp.tile_nch)) || (e = plan->v_b2.upload(vp.b2)) || (e = plan->v_cx.upload(vp.cx)) || (e = plan->v_cxi.upload(vp.cxi)))
                return cuda_fail(e, "uploading V-first resize plan");
            VfParams &v = plan->vparams;
            v.sw = g.sw; v.sh = g.sh; v.ry = g.ry; v.pw = g.pw; v.ph = g.ph;
            v.ndig = vp.ndig; v.nch1 = vp.nch1; v.nch2 = vp.nch2;
            v.n_recs = vp.n_recs; v.rec_words = vp.rec_words; v.recs = (const uint32_t *)plan->v_recs.p;
            v.row_computed = (const uint8_t *)plan->v_rowc.p;
            v.sh1 = vp.sh1;
            v.n_strips = vp.n_strips;
            v.strip_col0 = (const int32_t *)plan->v_col0.p; v.strip_src0 = (const int32_t *)plan->v_src0.p;
            v.strip_nblk = (const int32_t *)plan->v_nblk.p; v.strip_bytes = (const int32_t *)plan->v_bytes.p;
            v.strip_upw = (const int32_t *)plan->v_upw.p; v.strip_nunits = (const int32_t *)plan->v_nunits.p; v.unit = vp.unit; v.early_go = vp.early_go ? 1u : 0u;
            v.tile_kb0 = (const int32_t *)plan->v_kb0.p; v.tile_nch = (const int32_t *)plan->v_tnch.p;
            v.b2 = (const uint32_t *)plan->v_b2.p; v.cx = (const float *)plan->v_cx.p; v.cxi = (const int32_t *)plan->v_cxi.p; v.sh2 = vp.sh2;
            for (int l = 0; l < 2; l++)
                for (int d = 0; d < 3; d++) v.s2[l][d] = vp.s2[l][d];
            v.vbufs = vp.vbufs; v.cohorts = vp.cohorts; v.ring_rows = vp.ring_rows; v.pitch = vp.pitch; v.pitch_v = vp.pitch_v; v.smem_bytes = vp.smem_bytes;
            v.fw = g.fw; v.fh = g.fh; v.fox = g.fox; v.foy = g.foy;
            v.fill_post = pack_fill(g.fill_post); v.fill_pre_packed = pack_fill(g.fill_pre);
            v.n_bands = 1; v.need_fill = (vp.need_fill ? 1u : 0u) | (vp.rows_pure_fill ? 2u : 0u);
            if (vp.need_fill && !std::getenv("SLIMG_B200_FILL_EARLY")) v.need_fill |= 4u;   // fill after the band loop (measured faster)
            const char *dbg_env = std::getenv("SLIMG_B200_VF_DEBUG");
            v.debug = dbg_env ? uint32_t(std::atoi(dbg_env)) : 0u;
            const char *pf_env = std::getenv("SLIMG_B200_PREFETCH_STEPS");
            // ring steps prefetched into L2 ahead of the ring: 8 - 16 measured best for bands of 3+ chunks (32 worse); bands of 1 - 2
            // chunks (ratios below ~4.5x) have wide strips and a ring that already runs 3+ groups ahead -- the prefetch only costs
            // the TMA engine there (8K 4x 30.75 with, 29.88 us without; 3x 32.61 / 32.30; 2x 57.00 / 56.45)
            v.pf_steps = pf_env ? uint32_t(std::atoi(pf_env)) : (vp.nch1 <= 2 ? 0u : 16u);
            plan->vfirst = true;
        }
    }
    // H-first tensor-core path: what serves bands beyond the V-first ring (and A/B runs).  Not planned when a better plan
    // exists: its 1.5 - 2.5 ms of host work per geometry would be a quarter of a corpus of unique sizes.
    FastPlanHost fp;
    if (!plan->vfirst && !plan->tile && !plan->tc) build_fast_plan(g, tx, ty, fp);
    if (fp.ok && path != "generic" && std::getenv("SLIMG_B200_FORCE_GENERIC") == nullptr) {
        std::vector<VQuad> vquads(fp.vrow_w.size() / 32);
        for (size_t y = 0; y < vquads.size() * 4; y++) {
            VQuad &v = vquads[y / 4];
            const int r = int(y % 4);
            v.done[r] = 0;
            for (int s = 0; s < 8; s++) {
                v.w2[r][2 * s] = v.w2[r][2 * s + 1] = fp.vrow_w[y * 8 + s];
                v.out[r][s] = fp.vrow_out[y * 8 + s];
                if (v.out[r][s] >= 0) v.done[r] |= 1u << s;
            }
        }
        if ((e = plan->f_a0.upload(fp.tile_a0)) || (e = plan->f_nch.upload(fp.tile_nchunks)) ||
            (e = plan->f_boff.upload(fp.tile_boff)) || (e = plan->f_bfrag.upload(fp.bfrag)) ||
            (e = plan->f_fillx.upload(fp.fillw_x)) || (e = plan->f_vrow.upload(vquads)) ||
            (e = plan->f_vleft.upload(fp.vleft)) || (e = plan->f_vright.upload(fp.vright)) ||
            (e = plan->f_filly.upload(fp.fillw_y)) || (e = plan->f_rowc.upload(fp.row_computed)) ||
            (e = plan->f_sa0.upload(fp.strip_a0)) || (e = plan->f_sbytes.upload(fp.strip_bytes)))
            return cuda_fail(e, "uploading fast resize plan");
        FastResizeParams &f = plan->fparams;
        f.sw = g.sw; f.sh = g.sh; f.ry = g.ry; f.rh = g.rh; f.pw = g.pw; f.ph = g.ph;
        f.n_tiles = fp.n_tiles; f.n_strips = fp.n_strips;
        f.tile_a0 = (const int32_t *)plan->f_a0.p; f.tile_nchunks = (const int32_t *)plan->f_nch.p;
        f.tile_boff = (const int32_t *)plan->f_boff.p; f.bfrag = (const uint32_t *)plan->f_bfrag.p;
        f.fillw_x = (const float *)plan->f_fillx.p; f.hscale = fp.hscale;
        f.vquad = (const VQuad *)plan->f_vrow.p;
        f.vleft = (const int32_t *)plan->f_vleft.p; f.vright = (const int32_t *)plan->f_vright.p;
        f.fillw_y = (const float *)plan->f_filly.p; f.row_computed = (const uint8_t *)plan->f_rowc.p;
        for (int c = 0; c < 4; c++) f.fill_pre[c] = float(g.fill_pre[c]);
        f.fw = g.fw; f.fh = g.fh; f.fox = g.fox; f.foy = g.foy;
        f.fill_post = pack_fill(g.fill_post); f.fill_pre_packed = pack_fill(g.fill_pre);
        f.strip_a0 = (const int32_t *)plan->f_sa0.p; f.strip_bytes = (const int32_t *)plan->f_sbytes.p;
        f.pitch = fp.pitch;
        f.max_chunks = 0;
        for (int32_t c : fp.tile_nchunks) f.max_chunks = std::max<uint32_t>(f.max_chunks, uint32_t(c));
        f.n_bands = 1; f.n_fill_blocks = fp.need_fill ? 1 : 0;
        plan->fast = true;
    }
    // the tables were copied on this thread's stream from vectors that die here
    if ((e = cudaStreamSynchronize(cudaStreamPerThread)) != cudaSuccess) return cuda_fail(e, "uploading resize plan");
    out = plan;
    return SLIMG_OK;
}

static slimg_status get_resize_plan(const PipeGeom &g, std::shared_ptr<ResizePlan> &out) {
    int dev = 0;
    CU(cudaGetDevice(&dev));
    DeviceCtx &ctx = device_ctx(dev);
    const std::string path = requested_path();
    GeomKey key = make_key(g);
    key.v[23] = path_code(path);
    {
        std::lock_guard<std::mutex> lk(ctx.mu);
        auto it = ctx.plans.find(key);
        if (it != ctx.plans.end()) { it->second.stamp = ++ctx.clock; out = it->second.plan; return SLIMG_OK; }
    }
    // built outside the lock: other workers of this device keep launching cached geometries meanwhile (two workers that
    // miss on the same geometry both build it; the second insert is dropped)
    slimg_status st = build_resize_plan(g, path, out);
    if (st != SLIMG_OK) return st;
    std::shared_ptr<ResizePlan> victim;                 // dropped after the lock is released (its slabs go to the graveyard)
    {
        std::lock_guard<std::mutex> lk(ctx.mu);
        auto it = ctx.plans.find(key);
        if (it != ctx.plans.end()) { victim = out; it->second.stamp = ++ctx.clock; out = it->second.plan; }
        else {
            if (ctx.plans.size() >= kPlanCacheEntries) {
                auto lru = ctx.plans.begin();
                for (auto j = ctx.plans.begin(); j != ctx.plans.end(); ++j)
                    if (j->second.stamp < lru->second.stamp) lru = j;
                victim = lru->second.plan;              // in-flight launches that still use it hold their own reference
                ctx.plans.erase(lru);
            }
            ctx.plans[key] = DeviceCtx::Entry{out, ++ctx.clock};
        }
    }
    return SLIMG_OK;
}

// Rows [r0,r1) x columns [c0,c1) of the source that the pipeline reads for this geometry.
static bool remap_window(const PipeGeom &g, uint32_t &r0, uint32_t &r1, uint32_t &c0, uint32_t &c1) {
    if (!g.has_resize) {      // R == V: the kept rectangle of V, where it is backed by the source rectangle
        const int64_t x0 = std::max<int64_t>(g.vox, g.px), x1 = std::min<int64_t>(int64_t(g.vox) + g.rw, int64_t(g.px) + g.pw);
        const int64_t y0 = std::max<int64_t>(g.voy, g.py), y1 = std::min<int64_t>(int64_t(g.voy) + g.rh, int64_t(g.py) + g.ph);
        if (x1 <= x0 || y1 <= y0) { r0 = r1 = c0 = c1 = 0; return true; }
        c0 = uint32_t(x0 - g.vox + g.rx); c1 = uint32_t(x1 - g.vox + g.rx);
        r0 = uint32_t(y0 - g.voy + g.ry); r1 = uint32_t(y1 - g.voy + g.ry);
        return true;
    }
    return false;
}
static slimg_status source_window(const PipeGeom &g, uint32_t &r0, uint32_t &r1, uint32_t &c0, uint32_t &c1) {
    if (remap_window(g, r0, r1, c0, c1)) return SLIMG_OK;
    std::shared_ptr<ResizePlan> plan;
    slimg_status st = get_resize_plan(g, plan);
    if (st != SLIMG_OK) return st;
    r0 = plan->win_r0; r1 = plan->win_r1; c0 = plan->win_c0; c1 = plan->win_c1;
    return SLIMG_OK;
}

// Enqueue the whole pixel pipeline for n same-geometry images that are resident on the current
// device.  One kernel launch per 65535 images.
static void keep_pool_memory(int device);

static slimg_status run_pipeline_device(const PipeGeom &g, const uint8_t *src, uint32_t n, uint8_t *dst,
                                        cudaStream_t stream) {
    const size_t src_bytes = size_t(g.sw) * g.sh * 4, dst_bytes = size_t(g.fw) * g.fh * 4;
    if (!g.has_resize) {
        RemapParams p{};
        p.src = src; p.dst = dst; p.src_img_bytes = src_bytes; p.dst_img_bytes = dst_bytes;
        p.sw = g.sw; p.fw = g.fw; p.fh = g.fh;
        // F(x,y) -> V(x - fox + px, y - foy + py) -> S(.. - vox + rx, .. - voy + ry)
        const int64_t shift_x = int64_t(g.px) - g.fox, shift_y = int64_t(g.py) - g.foy;
        p.bx0 = int32_t(g.fox); p.by0 = int32_t(g.foy);
        p.bx1 = int32_t(g.fox + g.pw); p.by1 = int32_t(g.foy + g.ph);
        p.ax0 = int32_t(std::max<int64_t>(p.bx0, int64_t(g.vox) - shift_x));
        p.ay0 = int32_t(std::max<int64_t>(p.by0, int64_t(g.voy) - shift_y));
        p.ax1 = int32_t(std::min<int64_t>(p.bx1, int64_t(g.vox) + g.rw - shift_x));
        p.ay1 = int32_t(std::min<int64_t>(p.by1, int64_t(g.voy) + g.rh - shift_y));
        p.dx = int32_t(shift_x - int64_t(g.vox) + g.rx);
        p.dy = int32_t(shift_y - int64_t(g.voy) + g.ry);
        p.fill_in = pack_fill(g.fill_pre); p.fill_out = pack_fill(g.fill_post);
        CU(launch_remap(p, n, stream));
        note_path(SLIMG_PATH_REMAP);
        return SLIMG_OK;
    }
    // Rows that are not 16-byte aligned (width not a multiple of 4 pixels, or an odd base address) cannot be fetched by the
    // TMA engine or ldmatrix, and the f32 fallback kernels run 12x slower (5998 x 4000 -> Width(800): 207 us against 17 us
    // for 6000 x 4000).  Such batches are first re-pitched on the device -- one strided copy into a scratch whose rows are
    // padded to a multiple of 4 pixels -- and then take the tensor pipelines on the padded geometry: the padding columns
    // lie beyond every tap window, so whatever they hold meets a zero weight.
    const bool aligned = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
    if ((g.sw % 4 != 0 || !aligned) && g.sw >= 4 && !std::getenv("SLIMG_B200_NO_REPITCH") && requested_path() != "generic") {
        PipeGeom gp = g;
        gp.sw = (g.sw + 3u) & ~3u;
        std::shared_ptr<ResizePlan> pp;
        slimg_status stp = get_resize_plan(gp, pp);
        if (stp == SLIMG_OK && (pp->tile || pp->tc || pp->vfirst || pp->fast)) {
            const size_t pad_bytes = size_t(gp.sw) * gp.sh * 4;
            const uint32_t chunk = uint32_t(std::max<size_t>(1, std::min<size_t>(n, (size_t(1) << 31) / pad_bytes)));   // <= 2 GB of scratch
            uint8_t *scratch = nullptr;
            {   // the scratch comes back to the pool at the end of the call: keep it there for the next one (a 2 GB block
                // handed back to the driver at every synchronisation cost this path 70 us per image)
                int dev = 0;
                if (cudaGetDevice(&dev) == cudaSuccess) keep_pool_memory(dev);
            }
            cudaError_t e = cudaMallocAsync(reinterpret_cast<void **>(&scratch), pad_bytes * chunk, stream);
            if (e != cudaSuccess) return cuda_fail(e, "cudaMallocAsync (re-pitch scratch)");
            slimg_status st2 = SLIMG_OK;
            for (uint32_t first = 0; first < n && st2 == SLIMG_OK; first += chunk) {
                const uint32_t m = std::min(chunk, n - first);
                // the copy is K1 (k1_remap.cu): the image "extended" by the padding columns on the right (a strided
                // cudaMemcpy2DAsync of 4 n + 2-byte rows measured 1.8 TB/s; K1 streams at the HBM roofline)
                RemapParams rp{};
                rp.src = src + size_t(first) * src_bytes; rp.dst = scratch; rp.src_img_bytes = src_bytes; rp.dst_img_bytes = pad_bytes;
                rp.sw = g.sw; rp.fw = gp.sw; rp.fh = gp.sh;
                rp.bx0 = 0; rp.by0 = 0; rp.bx1 = int32_t(gp.sw); rp.by1 = int32_t(gp.sh);
                rp.ax0 = 0; rp.ay0 = 0; rp.ax1 = int32_t(g.sw); rp.ay1 = int32_t(g.sh);
                rp.dx = 0; rp.dy = 0; rp.fill_in = 0u; rp.fill_out = 0u;
                e = launch_remap(rp, m, stream);
                if (e != cudaSuccess) { st2 = cuda_fail(e, "re-pitch copy"); break; }
                st2 = run_pipeline_device(gp, scratch, m, dst + size_t(first) * dst_bytes, stream);
            }
            cudaFreeAsync(scratch, stream);
            return st2;
        }
    }
    std::shared_ptr<ResizePlan> plan;
    slimg_status st = get_resize_plan(g, plan);
    if (st != SLIMG_OK) return st;
    // A fast path whose launch is refused (no tensor-map entry point in the driver, an uninstantiated template case)
    // falls through to the next one; the generic kernel takes everything.
    if (plan->tile && aligned) {
        TileParams l = plan->lparams;
        l.src = src; l.dst = dst; l.src_img_bytes = src_bytes; l.dst_img_bytes = dst_bytes;
        if (launch_resize_tile(l, n, stream) == cudaSuccess) { note_path(SLIMG_PATH_TILE); return SLIMG_OK; }
        cudaGetLastError();
    }
    if (plan->tc && aligned) {
        TcParams t = plan->tparams;
        t.src = src; t.dst = dst; t.src_img_bytes = src_bytes; t.dst_img_bytes = dst_bytes;
        // bands of 16-row blocks: enough CTAs to fill the GPU when the batch is small
        const uint64_t ctas_per_band = uint64_t(t.n_strips) * n;
        uint32_t bands = uint32_t((2 * 148 + ctas_per_band - 1) / ctas_per_band);
        bands = std::max<uint32_t>(1, std::min<uint32_t>(bands, std::max<uint32_t>(1, t.n_blocks / 4)));
        t.n_bands = bands;
        if (launch_resize_tc(t, n, stream) == cudaSuccess) { note_path(SLIMG_PATH_TC); return SLIMG_OK; }
        cudaGetLastError();
    }
    if (plan->vfirst && aligned) {
        VfParams v = plan->vparams;
        v.src = src; v.dst = dst; v.src_img_bytes = src_bytes; v.dst_img_bytes = dst_bytes;
        // bands of output-row groups: enough CTAs to fill the GPU when the batch is small, no halo
        // recomputation (one band's worth of source rows per extra band) when the batch alone fills it
        const uint64_t ctas_per_band = uint64_t(v.n_strips) * n;
        uint32_t bands = uint32_t((2 * 148 + ctas_per_band - 1) / ctas_per_band);
        bands = std::max<uint32_t>(1, std::min<uint32_t>(bands, std::max<uint32_t>(1, v.n_recs / 4)));
        v.n_bands = bands;
        if (launch_resize_vfirst(v, n, stream) == cudaSuccess) { note_path(SLIMG_PATH_VFIRST); return SLIMG_OK; }
        cudaGetLastError();
    }
    if (plan->fast && aligned) {
        FastResizeParams f = plan->fparams;
        f.src = src; f.dst = dst; f.src_img_bytes = src_bytes; f.dst_img_bytes = dst_bytes;
        // bands of output rows: enough CTAs to fill the GPU when the batch is small, none of the
        // (6*ratio rows) halo recomputation when the batch alone fills it
        const uint64_t ctas_per_band = uint64_t(f.n_strips) * n;
        uint32_t bands = uint32_t((2 * 148 * 2 + ctas_per_band - 1) / ctas_per_band);
        bands = std::max<uint32_t>(1, std::min<uint32_t>(bands, std::max<uint32_t>(1, f.ph / 8)));
        f.n_bands = bands;
        if (launch_resize_fast(f, n, stream) == cudaSuccess) { note_path(SLIMG_PATH_HFIRST); return SLIMG_OK; }
        cudaGetLastError();
    }
    ResizeParams p = plan->params;
    p.src = src; p.dst = dst; p.src_img_bytes = src_bytes; p.dst_img_bytes = dst_bytes;
    CU(launch_resize_generic(p, n, stream));
    return SLIMG_OK;
}

static slimg_status check_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(SLIMG_ERR_IMAGE, "no CUDA device: libslimg_b200 has no CPU fallback");
    }
    return SLIMG_OK;
}

// first stage of the pipeline decides which error category a malformed input buffer reports
static slimg_status bad_buffer_status(const slimg_pipeline_options &o, int order) {
    if (order == SLIMG_ORDER_CONVERT) {
        if (o.has_crop) return SLIMG_ERR_CROP;
        if (o.has_extend) return SLIMG_ERR_EXTEND;
        if (o.has_resize) return SLIMG_ERR_RESIZE;
    } else {
        if (o.has_resize) return SLIMG_ERR_RESIZE;
        if (o.has_crop) return SLIMG_ERR_CROP;
        if (o.has_extend) return SLIMG_ERR_EXTEND;
    }
    return SLIMG_ERR_IMAGE;
}

static slimg_status validate_image(const slimg_image *in, const slimg_pipeline_options &o, int order) {
    if (!in) return fail(SLIMG_ERR_IMAGE, "null image");
    const size_t expected = size_t(in->width) * in->height * 4;
    // extend.rs:98-104 demands len == w*h*4; crop/resize index the buffer and need at least that much
    const bool exact = (order == SLIMG_ORDER_CONVERT && !o.has_crop && o.has_extend) ||
                       (order == SLIMG_ORDER_CHAIN && !o.has_resize && !o.has_crop && o.has_extend);
    if ((exact && in->len != expected) || in->len < expected || (expected > 0 && !in->data)) {
        char buf[160];
        std::snprintf(buf, sizeof buf, "invalid image data: expected %zu bytes (%ux%ux4), got %zu", expected,
                      in->width, in->height, in->len);
        return fail(bad_buffer_status(o, order), buf);
    }
    if (in->width == 0 || in->height == 0) return fail(bad_buffer_status(o, order), "image has a zero dimension");
    return SLIMG_OK;
}

// Keep the stream-ordered pool's memory across synchronisations: with the default release threshold (0) every wait hands
// the image buffers back to the OS and the next image pays for fresh physical memory (measured on the host-fed bench:
// sporadic 60-870 ms steps against 48.4 ms steady state).  Bounded, so an idle process does not sit on the GPU.  Once per
// device and process.
static void keep_pool_memory(int device) {
    static std::mutex mu;
    static std::map<int, bool> done;
    std::lock_guard<std::mutex> lk(mu);
    if (done[device]) return;
    done[device] = true;
    cudaMemPool_t pool = nullptr;
    uint64_t keep = uint64_t(4) << 30;
    if (cudaDeviceGetDefaultMemPool(&pool, device) != cudaSuccess ||
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep) != cudaSuccess)
        cudaGetLastError();
}

// ---- pinned staging for pageable host buffers -------------------------------------------------------------------
// Callers behind this boundary hand over ordinary heap memory (a Rust Vec, Python bytes, a UniFFI RustBuffer).  A
// cudaMemcpyAsync from such memory is staged by the driver synchronously, one bounce buffer at a time; here the copy
// goes through two pinned slots instead, the memcpy into one slot overlapping the DMA out of the other.  Rings are
// pooled per device (their events belong to a device) and handed to one worker at a time.
constexpr size_t kStageChunk = size_t(8) << 20;
struct StagingRing {
    int device = -1;
    uint8_t *slot[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    bool pending[2] = {false, false};
};
static std::mutex g_ring_mu;
static std::vector<StagingRing *> g_ring_pool;

static StagingRing *acquire_ring(int device) {
    {
        std::lock_guard<std::mutex> lk(g_ring_mu);
        for (size_t i = 0; i < g_ring_pool.size(); i++)
            if (g_ring_pool[i]->device == device) {
                StagingRing *r = g_ring_pool[i];
                g_ring_pool.erase(g_ring_pool.begin() + long(i));
                return r;
            }
    }
    auto *r = new StagingRing();
    r->device = device;
    for (int k = 0; k < 2; k++)
        if (cudaMallocHost((void **)&r->slot[k], kStageChunk) != cudaSuccess ||
            cudaEventCreateWithFlags(&r->ev[k], cudaEventBlockingSync | cudaEventDisableTiming) != cudaSuccess) {
            cudaGetLastError();
            for (int j = 0; j < 2; j++) { if (r->slot[j]) cudaFreeHost(r->slot[j]); if (r->ev[j]) cudaEventDestroy(r->ev[j]); }
            delete r;
            return nullptr;
        }
    return r;
}
static void release_ring(StagingRing *r) {
    if (!r) return;
    std::lock_guard<std::mutex> lk(g_ring_mu);
    g_ring_pool.push_back(r);
}
struct RingLease {                       // a worker's ring, acquired on first use
    StagingRing *ring = nullptr;
    ~RingLease() { release_ring(ring); }
    StagingRing *get() {
        if (!ring) { int dev = 0; if (cudaGetDevice(&dev) == cudaSuccess) ring = acquire_ring(dev); }
        return ring;
    }
};

static bool is_pinned(const void *p) {
    if (std::getenv("SLIMG_B200_NO_STAGING")) return true;          // A/B: let the driver stage pageable copies
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// host -> device; `src` may be pageable.  Enqueues on `stream`; when it returns, `src` has been read.
static cudaError_t copy_h2d(uint8_t *dst, const uint8_t *src, size_t bytes, cudaStream_t stream, RingLease &lease) {
    if (bytes == 0) return cudaSuccess;
    StagingRing *r = nullptr;
    if (is_pinned(src) || !(r = lease.get())) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream);
    cudaError_t e = cudaSuccess;
    int k = 0;
    for (size_t off = 0; off < bytes; off += kStageChunk, k ^= 1) {
        const size_t n = std::min(kStageChunk, bytes - off);
        if (r->pending[k]) { if ((e = cudaEventSynchronize(r->ev[k])) != cudaSuccess) return e; r->pending[k] = false; }
        std::memcpy(r->slot[k], src + off, n);
        if ((e = cudaMemcpyAsync(dst + off, r->slot[k], n, cudaMemcpyHostToDevice, stream)) != cudaSuccess) return e;
        if ((e = cudaEventRecord(r->ev[k], stream)) != cudaSuccess) return e;
        r->pending[k] = true;
    }
    return cudaSuccess;
}

// device -> host; `dst` may be pageable.  With a pinned `dst` the copy is only enqueued; with a pageable one the call
// returns when the bytes are in place.
static cudaError_t copy_d2h(uint8_t *dst, const uint8_t *src, size_t bytes, cudaStream_t stream, RingLease &lease) {
    if (bytes == 0) return cudaSuccess;
    StagingRing *r = nullptr;
    if (is_pinned(dst) || !(r = lease.get())) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream);
    cudaError_t e = cudaSuccess;
    for (int k = 0; k < 2; k++)          // slots still draining an upload of this worker
        if (r->pending[k]) { if ((e = cudaEventSynchronize(r->ev[k])) != cudaSuccess) return e; r->pending[k] = false; }
    size_t prev_off = 0, prev_n = 0;
    int k = 0;
    for (size_t off = 0; off < bytes; off += kStageChunk, k ^= 1) {
        const size_t n = std::min(kStageChunk, bytes - off);
        if ((e = cudaMemcpyAsync(r->slot[k], src + off, n, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return e;
        if ((e = cudaEventRecord(r->ev[k], stream)) != cudaSuccess) return e;
        if (prev_n) {                    // the previous chunk, while this one is on the link
            if ((e = cudaEventSynchronize(r->ev[k ^ 1])) != cudaSuccess) return e;
            std::memcpy(dst + prev_off, r->slot[k ^ 1], prev_n);
        }
        prev_off = off; prev_n = n;
    }
    if ((e = cudaEventSynchronize(r->ev[k ^ 1])) != cudaSuccess) return e;
    std::memcpy(dst + prev_off, r->slot[k ^ 1], prev_n);
    return cudaSuccess;
}

// compute_reciprocal (jcdctmgr.c:173-221), DCTELEM = int, divisor = quantval << 3 (:245-251)
static void build_quant(const uint16_t ql[64], const uint16_t qc[64], QuantDev &q) {
    const uint16_t *t[2] = {ql, qc};
    for (int k = 0; k < 2; k++)
        for (int i = 0; i < 64; i++) {
            const uint32_t divisor = uint32_t(t[k][i] ? t[k][i] : 1) << 3;
            int b = 31 - __builtin_clz(divisor);
            int r = 32 + b;
            uint64_t fq = (uint64_t(1) << r) / divisor, fr = (uint64_t(1) << r) % divisor;
            uint32_t c = divisor / 2;
            if (fr == 0) { fq >>= 1; r--; }
            else if (fr <= divisor / 2u) c++;
            else fq++;
            q.recip[k][i] = uint32_t(fq); q.corr[k][i] = c; q.shift32[k][i] = uint32_t(r - 32);
        }
}

static void jpeg_geometry(uint32_t w, uint32_t h, slimg_jpeg_geometry &g) {
    auto cdiv = [](uint32_t a, uint32_t b) { return uint32_t((uint64_t(a) + b - 1) / b); };
    g.width = w; g.height = h;
    g.y_wb = cdiv(w, 8); g.y_hb = cdiv(h, 8);
    g.c_wb = cdiv(w, 16); g.c_hb = cdiv(h, 16);
    g.y_WB = g.c_wb * 2; g.y_HB = g.c_hb * 2;
    g.c_WB = g.c_wb; g.c_HB = g.c_hb;
}

static slimg_status run_jpeg_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h, const uint16_t ql[64],
                                    const uint16_t qc[64], int16_t *y, int16_t *cb, int16_t *cr, int16_t *yu, int16_t *cbu,
                                    int16_t *cru, uint32_t flags, cudaStream_t stream) {
    slimg_jpeg_geometry g;
    jpeg_geometry(w, h, g);
    JpegParams p{};
    p.src = src; p.y = y; p.cb = cb; p.cr = cr;
    p.yu = yu; p.cbu = cbu; p.cru = cru;
    p.deringing = (flags & SLIMG_JPEG_DERINGING) ? 1u : 0u;
    p.dc_quant[0] = ql[0]; p.dc_quant[1] = qc[0];
    p.src_img_bytes = size_t(w) * h * 4;
    p.y_img_elems = size_t(g.y_HB) * g.y_WB * 64;
    p.c_img_elems = size_t(g.c_HB) * g.c_WB * 64;
    p.w = w; p.h = h;
    p.y_wb = g.y_wb; p.y_hb = g.y_hb; p.y_WB = g.y_WB; p.y_HB = g.y_HB; p.c_wb = g.c_wb; p.c_hb = g.c_hb;
    build_quant(ql, qc, p.q);
    CU(launch_jpeg_frontend(p, n, stream));
    return SLIMG_OK;
}


// libwebp InitGammaTables (picture_csp_enc.c): kGamma = 0.80, 12-bit linear scale, 32-entry inverse table
static void webp_gamma_tables(uint16_t g2l[256], int32_t l2g[33]) {
    const double gamma = 0.80, norm = 1.0 / 255.0, scale = double(1 << 7) / 4095.0;
    for (int v = 0; v <= 255; v++) g2l[v] = uint16_t(std::pow(norm * v, gamma) * 4095.0 + .5);
    for (int v = 0; v <= 32; v++) l2g[v] = int32_t(255. * std::pow(scale * v, 1. / gamma) + .5);
}

static slimg_status run_planes_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h, int kind, void *y, void *cb,
                                      void *cr, void *a, void *scratch, cudaStream_t stream) {
    const size_t px = size_t(w) * h;
    if (kind == SLIMG_PLANES_WEBP_YUV420) {
        WebpPlanesParams p{};
        p.src = src; p.src_img_bytes = px * 4; p.y_img_bytes = px;
        p.uv_img_bytes = size_t((w + 1) / 2) * ((h + 1) / 2);
        p.w = w; p.h = h;
        p.y = static_cast<uint8_t *>(y); p.u = static_cast<uint8_t *>(cb); p.v = static_cast<uint8_t *>(cr); p.a = static_cast<uint8_t *>(a);
        if (a && scratch) {
            const size_t rows = size_t(n) * ((h + 1) / 2);
            // the planes kernel decides per 2 x 2 block and raises the per-image flag itself: no scan pass over the source
            // (row_alpha stays NULL; the flags keep their place behind the n * ceil(h/2) bytes the scan used to fill)
            p.row_alpha = nullptr;
            p.any_alpha = reinterpret_cast<uint32_t *>(static_cast<uint8_t *>(scratch) + ((rows + 3) & ~size_t(3)));
            CU(cudaMemsetAsync(p.any_alpha, 0, size_t(n) * 4, stream));
        }
        webp_gamma_tables(p.g2l, p.l2g);
        CU(launch_webp_planes(p, n, stream));
        return SLIMG_OK;
    }
    if (kind == SLIMG_PLANES_AVIF_YCBCR444_8 || kind == SLIMG_PLANES_AVIF_YCBCR444_10) {
        AvifPlanesParams p{};
        p.src = src; p.src_img_bytes = px * 4; p.w = w; p.h = h;
        p.depth = kind == SLIMG_PLANES_AVIF_YCBCR444_8 ? 8 : 10;
        p.y = y; p.cb = cb; p.cr = cr; p.a = a;
        CU(launch_avif_planes(p, n, stream));
        return SLIMG_OK;
    }
    return fail(SLIMG_ERR_IMAGE, "unknown plane kind");
}

// what a convert-to-JPEG call wants after the pixel stages
struct JpegRequest {
    const uint16_t *ql, *qc;
    uint32_t flags;
    slimg_jpeg_coefs *out;
};

// one image, host in / host out, on the current device: the pixel stages, then (jpeg != nullptr) the JPEG front end on
// the image that is still in HBM -- only coefficients come back
// `done` (optional): a cudaEventBlockingSync event the wait goes through, so a scheduler worker SLEEPS while its image
// is in flight instead of spinning on the stream (8 ranks x 4 workers would otherwise fight over the host cores)
static slimg_status pipeline_one(const slimg_image *in, const slimg_pipeline_options &o, int order, slimg_image *out,
                                 const JpegRequest *jpeg, cudaStream_t stream, bool caller_buffer, RingLease &lease,
                                 cudaEvent_t done = nullptr) {
    slimg_status st = validate_image(in, o, order);
    if (st != SLIMG_OK) return st;
    PipeGeom g;
    if ((st = build_geometry(in->width, in->height, o, order, g)) != SLIMG_OK) return st;
    const size_t src_bytes = size_t(g.sw) * g.sh * 4, dst_bytes = size_t(g.fw) * g.fh * 4;
    const bool identity = !o.has_resize && !o.has_crop && !o.has_extend;

    {
        int dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess) keep_pool_memory(dev);
    }
    uint8_t *host_out = nullptr;           // pixels (no jpeg) or the coefficient planes (jpeg)
    size_t ye = 0, ce = 0, coef_elems = 0;
    slimg_jpeg_geometry jg{};
    const bool uq = jpeg && (jpeg->flags & SLIMG_JPEG_UNQUANTISED);
    if (jpeg) {
        jpeg_geometry(g.fw, g.fh, jg);
        ye = size_t(jg.y_HB) * jg.y_WB * 64; ce = size_t(jg.c_HB) * jg.c_WB * 64;
        coef_elems = (ye + 2 * ce) * (uq ? 2 : 1);
        host_out = static_cast<uint8_t *>(std::malloc(coef_elems * 2));
    } else if (caller_buffer && out->data && out->len >= dst_bytes) host_out = out->data;
    else host_out = static_cast<uint8_t *>(std::malloc(dst_bytes ? dst_bytes : 1));
    if (!host_out) return fail(SLIMG_ERR_IMAGE, "out of host memory");
    const bool own_out = jpeg || host_out != out->data;
    uint8_t *d_src = nullptr, *d_dst = nullptr;
    int16_t *d_coef = nullptr;
    auto cleanup = [&](slimg_status s) {
        if (d_src) cudaFreeAsync(d_src, stream);
        if (d_dst) cudaFreeAsync(d_dst, stream);
        if (d_coef) cudaFreeAsync(d_coef, stream);
        if (s != SLIMG_OK && own_out) std::free(host_out);
        return s;
    };
    cudaError_t e;
    if ((e = cudaMallocAsync((void **)&d_src, src_bytes, stream)) != cudaSuccess) return cleanup(cuda_fail(e, "cudaMallocAsync"));
    if (!(jpeg && identity) &&
        (e = cudaMallocAsync((void **)&d_dst, dst_bytes ? dst_bytes : 1, stream)) != cudaSuccess) return cleanup(cuda_fail(e, "cudaMallocAsync"));
    if (jpeg && (e = cudaMallocAsync((void **)&d_coef, coef_elems * 2, stream)) != cudaSuccess) return cleanup(cuda_fail(e, "cudaMallocAsync"));
    // upload only what the kernels read: the rows (and, when the window is much narrower than the image, the columns)
    // of the source behind the kept output rectangle; the rest of d_src stays unwritten and unread (staging that
    // over-reads a ring step meets zero weights: integer 0 x anything = 0)
    uint32_t r0 = 0, r1 = g.sh, c0 = 0, c1 = g.sw;
    if (!(jpeg && identity) && (st = source_window(g, r0, r1, c0, c1)) != SLIMG_OK) return cleanup(st);
    if (r1 > r0 && c1 > c0) {
        if (uint64_t(c1 - c0) * 10 < uint64_t(g.sw) * 9 && is_pinned(in->data)) {
            const size_t off = (size_t(r0) * g.sw + c0) * 4;
            e = cudaMemcpy2DAsync(d_src + off, size_t(g.sw) * 4, in->data + off, size_t(g.sw) * 4, size_t(c1 - c0) * 4, r1 - r0,
                                  cudaMemcpyHostToDevice, stream);
        } else {
            const size_t off = size_t(r0) * g.sw * 4;
            e = copy_h2d(d_src + off, in->data + off, size_t(r1 - r0) * g.sw * 4, stream, lease);
        }
        if (e != cudaSuccess) return cleanup(cuda_fail(e, "H2D"));
    }
    const uint8_t *d_pixels = d_src;
    if (!(jpeg && identity)) {
        if ((st = run_pipeline_device(g, d_src, 1, d_dst, stream)) != SLIMG_OK) return cleanup(st);
        d_pixels = d_dst;
    }
    if (jpeg) {
        int16_t *dy = d_coef, *dcb = dy + ye, *dcr = dcb + ce;
        int16_t *dyu = uq ? dcr + ce : nullptr, *dcbu = uq ? dyu + ye : nullptr, *dcru = uq ? dcbu + ce : nullptr;
        if ((st = run_jpeg_device(d_pixels, 1, g.fw, g.fh, jpeg->ql, jpeg->qc, dy, dcb, dcr, dyu, dcbu, dcru, jpeg->flags, stream)) != SLIMG_OK)
            return cleanup(st);
        if ((e = copy_d2h(host_out, reinterpret_cast<const uint8_t *>(d_coef), coef_elems * 2, stream, lease)) != cudaSuccess)
            return cleanup(cuda_fail(e, "D2H"));
    } else if ((e = copy_d2h(host_out, d_dst, dst_bytes, stream, lease)) != cudaSuccess) return cleanup(cuda_fail(e, "D2H"));
    if (done) {
        if ((e = cudaEventRecord(done, stream)) != cudaSuccess || (e = cudaEventSynchronize(done)) != cudaSuccess)
            return cleanup(cuda_fail(e, "pipeline kernels"));
    } else if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return cleanup(cuda_fail(e, "pipeline kernels"));
    cleanup(SLIMG_OK);
    if (jpeg) {
        slimg_jpeg_coefs &c = *jpeg->out;
        std::memset(&c, 0, sizeof c);
        c.geom = jg;
        std::memcpy(c.qt_luma, jpeg->ql, 128);
        std::memcpy(c.qt_chroma, jpeg->qc, 128);
        int16_t *h = reinterpret_cast<int16_t *>(host_out);
        c.y = h; c.cb = h + ye; c.cr = h + ye + ce;
        if (uq) { c.y_uq = h + ye + 2 * ce; c.cb_uq = c.y_uq + ye; c.cr_uq = c.cb_uq + ce; }
    } else {
        out->width = g.fw; out->height = g.fh; out->data = host_out; out->len = dst_bytes;
    }
    return SLIMG_OK;
}

}  // namespace slimg

using namespace slimg;

extern "C" {

const char *slimg_b200_last_error(void) { return last_error_string().c_str(); }
int32_t slimg_b200_abi_version(void) { return SLIMG_B200_ABI_VERSION; }
uint64_t slimg_b200_kernel_launches(void) { return g_launches.load(); }

int32_t slimg_b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

void *slimg_b200_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void slimg_b200_host_free(void *p) { if (p) cudaFreeHost(p); }

void slimg_b200_image_free(slimg_image *img) {
    if (img && img->data) { std::free(img->data); img->data = nullptr; img->len = 0; }
}

slimg_status slimg_b200_pipeline(const slimg_image *in, const slimg_pipeline_options *opts, int order, slimg_image *out) {
    if (!opts || !out) return fail(SLIMG_ERR_IMAGE, "null argument");
    // geometry errors are reported before the device is touched, as the reference reports them
    // before allocating (crop.rs:71, extend.rs:95-96, resize.rs:65)
    slimg_status st = validate_image(in, *opts, order);
    if (st != SLIMG_OK) return st;
    PipeGeom g;
    if ((st = build_geometry(in->width, in->height, *opts, order, g)) != SLIMG_OK) return st;
    if ((st = check_device()) != SLIMG_OK) return st;
    out->data = nullptr; out->len = 0;
    RingLease lease;
    return pipeline_one(in, *opts, order, out, nullptr, cudaStreamPerThread, false, lease);
}

slimg_status slimg_b200_crop(const slimg_image *in, const slimg_crop_mode *mode, slimg_image *out) {
    if (!mode) return fail(SLIMG_ERR_IMAGE, "null argument");
    slimg_pipeline_options o{};
    o.has_crop = 1; o.crop = *mode;
    return slimg_b200_pipeline(in, &o, SLIMG_ORDER_CONVERT, out);
}

slimg_status slimg_b200_extend(const slimg_image *in, const slimg_extend_mode *mode, const slimg_fill_color *fill,
                               slimg_image *out) {
    if (!mode || !fill) return fail(SLIMG_ERR_IMAGE, "null argument");
    slimg_pipeline_options o{};
    o.has_extend = 1; o.extend = *mode; o.has_fill = 1; o.fill = *fill;
    return slimg_b200_pipeline(in, &o, SLIMG_ORDER_CONVERT, out);
}

slimg_status slimg_b200_resize(const slimg_image *in, const slimg_resize_mode *mode, slimg_image *out) {
    if (!mode) return fail(SLIMG_ERR_IMAGE, "null argument");
    slimg_pipeline_options o{};
    o.has_resize = 1; o.resize = *mode;
    return slimg_b200_pipeline(in, &o, SLIMG_ORDER_CONVERT, out);
}

slimg_status slimg_b200_pipeline_batch_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                              const slimg_pipeline_options *opts, int order, uint8_t *dst,
                                              size_t dst_capacity, void *stream) {
    if (!opts || (n && (!src || !dst))) return fail(SLIMG_ERR_IMAGE, "null argument");
    if (w == 0 || h == 0) return fail(bad_buffer_status(*opts, order), "image has a zero dimension");
    PipeGeom g;
    slimg_status st = build_geometry(w, h, *opts, order, g);
    if (st != SLIMG_OK) return st;
    if (size_t(g.fw) * g.fh * 4 * n > dst_capacity) return fail(SLIMG_ERR_IMAGE, "destination buffer too small");
    if ((st = check_device()) != SLIMG_OK) return st;
    return run_pipeline_device(g, src, n, dst, static_cast<cudaStream_t>(stream));
}

void slimg_b200_jpeg_geometry(uint32_t w, uint32_t h, slimg_jpeg_geometry *g) { if (g) jpeg_geometry(w, h, *g); }

// jcparam.c:130-152 jpeg_quality_scaling, :28-65 jpeg_add_quant_table (force_baseline = FALSE)
void slimg_b200_jpeg_quant_table(int quality, int which, int base, uint16_t out[64]) {
    static const uint16_t annexk_luma[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55,
                                             14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62,
                                             18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
                                             49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
    static const uint16_t annexk_chroma[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99,
                                               24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99,
                                               99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                               99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
    // MozJPEG quant_tbl_master_idx 3 (ImageMagick / N. Robidoux), used for luma and chroma alike.
    // mozjpeg-sys is not in the reference tree: restated, parity unpinned.
    static const uint16_t mozjpeg3[64] = {16, 16, 16, 18, 25, 37, 56, 85, 16, 17, 20, 27, 34, 40, 53, 75,
                                          16, 20, 24, 31, 43, 62, 91, 135, 18, 27, 31, 40, 53, 74, 106, 156,
                                          25, 34, 43, 53, 69, 94, 131, 189, 37, 40, 62, 74, 94, 124, 169, 238,
                                          56, 53, 91, 106, 131, 169, 226, 311, 85, 75, 135, 156, 189, 238, 311, 418};
    if (quality <= 0) quality = 1;
    if (quality > 100) quality = 100;
    const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
    const uint16_t *tbl = base == SLIMG_QT_MOZJPEG ? mozjpeg3 : (which == 0 ? annexk_luma : annexk_chroma);
    for (int i = 0; i < 64; i++) {
        long t = (long(tbl[i]) * scale + 50L) / 100L;
        if (t <= 0) t = 1;
        if (t > 32767) t = 32767;
        out[i] = uint16_t(t);
    }
}

slimg_status slimg_b200_jpeg_frontend_batch_device_ex(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                                      const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                                                      int16_t *y, int16_t *cb, int16_t *cr, int16_t *y_uq, int16_t *cb_uq,
                                                      int16_t *cr_uq, uint32_t flags, void *stream) {
    if (!qt_luma || !qt_chroma || (n && (!src || !y || !cb || !cr))) return fail(SLIMG_ERR_IMAGE, "null argument");
    if ((y_uq || cb_uq || cr_uq) && !(y_uq && cb_uq && cr_uq)) return fail(SLIMG_ERR_IMAGE, "un-quantised planes: all three or none");
    if (w == 0 || h == 0) return fail(SLIMG_ERR_ENCODE, "image has a zero dimension");
    slimg_status st = check_device();
    if (st != SLIMG_OK) return st;
    return run_jpeg_device(src, n, w, h, qt_luma, qt_chroma, y, cb, cr, y_uq, cb_uq, cr_uq, flags, static_cast<cudaStream_t>(stream));
}

slimg_status slimg_b200_jpeg_frontend_batch_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                                   const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                                                   int16_t *y, int16_t *cb, int16_t *cr, void *stream) {
    return slimg_b200_jpeg_frontend_batch_device_ex(src, n, w, h, qt_luma, qt_chroma, y, cb, cr, nullptr, nullptr, nullptr, 0, stream);
}

slimg_status slimg_b200_pipeline_jpeg_batch_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                                   const slimg_pipeline_options *opts, int order,
                                                   const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                                                   uint8_t *pixels, size_t pixels_capacity, int16_t *y, int16_t *cb,
                                                   int16_t *cr, int16_t *y_uq, int16_t *cb_uq, int16_t *cr_uq, uint32_t flags,
                                                   void *stream) {
    if (!opts || !qt_luma || !qt_chroma || (n && (!src || !pixels || !y || !cb || !cr))) return fail(SLIMG_ERR_IMAGE, "null argument");
    if ((y_uq || cb_uq || cr_uq) && !(y_uq && cb_uq && cr_uq)) return fail(SLIMG_ERR_IMAGE, "un-quantised planes: all three or none");
    if (w == 0 || h == 0) return fail(bad_buffer_status(*opts, order), "image has a zero dimension");
    PipeGeom g;
    slimg_status st = build_geometry(w, h, *opts, order, g);
    if (st != SLIMG_OK) return st;
    if (size_t(g.fw) * g.fh * 4 * n > pixels_capacity) return fail(SLIMG_ERR_IMAGE, "pixel scratch buffer too small");
    if ((st = check_device()) != SLIMG_OK) return st;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if ((st = run_pipeline_device(g, src, n, pixels, s)) != SLIMG_OK) return st;
    return run_jpeg_device(pixels, n, g.fw, g.fh, qt_luma, qt_chroma, y, cb, cr, y_uq, cb_uq, cr_uq, flags, s);
}

slimg_status slimg_b200_jpeg_frontend_ex(const slimg_image *in, const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                                         uint32_t flags, slimg_jpeg_coefs *out) {
    if (!in || !qt_luma || !qt_chroma || !out) return fail(SLIMG_ERR_IMAGE, "null argument");
    const size_t expected = size_t(in->width) * in->height * 4;
    if (in->width == 0 || in->height == 0 || in->len < expected || !in->data)
        return fail(SLIMG_ERR_ENCODE, "invalid image data for the JPEG front end");
    slimg_status st = check_device();
    if (st != SLIMG_OK) return st;
    const slimg_pipeline_options none{};
    const JpegRequest req{qt_luma, qt_chroma, flags, out};
    RingLease lease;
    return pipeline_one(in, none, SLIMG_ORDER_CONVERT, nullptr, &req, cudaStreamPerThread, false, lease);
}

slimg_status slimg_b200_jpeg_frontend(const slimg_image *in, const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                                      slimg_jpeg_coefs *out) {
    return slimg_b200_jpeg_frontend_ex(in, qt_luma, qt_chroma, 0, out);
}

slimg_status slimg_b200_pipeline_jpeg(const slimg_image *in, const slimg_pipeline_options *opts, int order,
                                      const uint16_t qt_luma[64], const uint16_t qt_chroma[64], uint32_t flags,
                                      slimg_jpeg_coefs *out) {
    if (!opts || !qt_luma || !qt_chroma || !out) return fail(SLIMG_ERR_IMAGE, "null argument");
    slimg_status st = validate_image(in, *opts, order);
    if (st != SLIMG_OK) return st;
    PipeGeom g;
    if ((st = build_geometry(in->width, in->height, *opts, order, g)) != SLIMG_OK) return st;
    if ((st = check_device()) != SLIMG_OK) return st;
    const JpegRequest req{qt_luma, qt_chroma, flags, out};
    RingLease lease;
    return pipeline_one(in, *opts, order, nullptr, &req, cudaStreamPerThread, false, lease);
}

void slimg_b200_jpeg_coefs_free(slimg_jpeg_coefs *c) {
    if (c && c->y) { std::free(c->y); c->y = c->cb = c->cr = c->y_uq = c->cb_uq = c->cr_uq = nullptr; }
}

slimg_status slimg_b200_pipeline_source_window(uint32_t w, uint32_t h, const slimg_pipeline_options *opts, int order,
                                               uint32_t *x, uint32_t *y, uint32_t *win_w, uint32_t *win_h) {
    if (!opts || !x || !y || !win_w || !win_h) return fail(SLIMG_ERR_IMAGE, "null argument");
    PipeGeom g;
    slimg_status st = build_geometry(w, h, *opts, order, g);
    if (st != SLIMG_OK) return st;
    uint32_t r0, r1, c0, c1;
    if (!remap_window(g, r0, r1, c0, c1)) {
        AxisTaps tx, ty;
        build_axis_taps(g.vw, g.Rw, g.vox, g.rw, tx);
        build_axis_taps(g.vh, g.Rh, g.voy, g.rh, ty);
        window_from_taps(g, tx, ty, r0, r1, c0, c1);
    }
    *x = c0; *y = r0; *win_w = c1 - c0; *win_h = r1 - r0;
    return SLIMG_OK;
}

// the scheduler behind both batch entry points: jq == nullptr -> pixels into out[], else coefficients into jout[]
static int32_t batch_host(const slimg_image *in, uint32_t n, const slimg_pipeline_options *opts, int order,
                          const slimg_batch_config *cfg, slimg_image *out, const JpegRequest *jq, slimg_jpeg_coefs *jout,
                          slimg_status *statuses) {
    if (!opts || (n && (!in || (!out && !jout)))) { fail(SLIMG_ERR_IMAGE, "null argument"); return -1; }
    if (check_device() != SLIMG_OK) {
        for (uint32_t i = 0; statuses && i < n; i++) statuses[i] = SLIMG_ERR_IMAGE;
        return int32_t(n);
    }
    int visible = slimg_b200_device_count();
    int n_dev = cfg && cfg->n_devices > 0 ? cfg->n_devices : visible;
    if (n_dev > visible && !(cfg && cfg->device_ids)) n_dev = visible;
    const int streams = cfg && cfg->streams_per_device > 0 ? cfg->streams_per_device : 3;

    std::atomic<uint32_t> next{0};
    std::mutex err_mu;
    std::string first_error;
    // every image starts as "not processed"; a worker overwrites the entry it takes
    std::vector<slimg_status> local_status;
    if (!statuses) { local_status.resize(n); statuses = local_status.data(); }
    for (uint32_t i = 0; i < n; i++) statuses[i] = SLIMG_ERR_IMAGE;
    auto worker = [&](int device) {
        if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return; }
        cudaStream_t s = nullptr;
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); return; }
        keep_pool_memory(device);
        cudaEvent_t done = nullptr;
        if (std::getenv("SLIMG_B200_SPIN_SYNC") || cudaEventCreateWithFlags(&done, cudaEventBlockingSync | cudaEventDisableTiming) != cudaSuccess) {
            cudaGetLastError();
            done = nullptr;
        }
        {
            RingLease lease;                                  // pinned staging, only touched for pageable buffers
            for (;;) {
                const uint32_t i = next.fetch_add(1);
                if (i >= n) break;
                slimg_status st;
                if (jq) {
                    const JpegRequest req{jq->ql, jq->qc, jq->flags, &jout[i]};
                    st = pipeline_one(&in[i], *opts, order, nullptr, &req, s, false, lease, done);
                } else st = pipeline_one(&in[i], *opts, order, &out[i], nullptr, s, true, lease, done);
                statuses[i] = st;
                if (st != SLIMG_OK) {
                    std::lock_guard<std::mutex> lk(err_mu);
                    if (first_error.empty()) first_error = "image " + std::to_string(i) + ": " + last_error_string();
                }
            }
            cudaStreamSynchronize(s);
        }
        if (done) cudaEventDestroy(done);
        cudaStreamDestroy(s);
    };
    std::vector<std::thread> threads;
    for (int d = 0; d < n_dev; d++)
        for (int k = 0; k < streams; k++)
            threads.emplace_back(worker, cfg && cfg->device_ids ? cfg->device_ids[d] : d);
    for (auto &t : threads) t.join();
    int32_t failed = 0;
    for (uint32_t i = 0; i < n; i++) failed += statuses[i] != SLIMG_OK;
    if (failed && first_error.empty()) first_error = "no worker could initialise its CUDA device";
    if (!first_error.empty()) set_error(first_error);
    return failed;
}

int32_t slimg_b200_pipeline_batch_host(const slimg_image *in, uint32_t n, const slimg_pipeline_options *opts, int order,
                                       const slimg_batch_config *cfg, slimg_image *out, slimg_status *statuses) {
    return batch_host(in, n, opts, order, cfg, out, nullptr, nullptr, statuses);
}

int32_t slimg_b200_pipeline_jpeg_batch_host(const slimg_image *in, uint32_t n, const slimg_pipeline_options *opts, int order,
                                            const uint16_t qt_luma[64], const uint16_t qt_chroma[64], uint32_t flags,
                                            const slimg_batch_config *cfg, slimg_jpeg_coefs *out, slimg_status *statuses) {
    if (!qt_luma || !qt_chroma) { fail(SLIMG_ERR_IMAGE, "null argument"); return -1; }
    for (uint32_t i = 0; out && i < n; i++) std::memset(&out[i], 0, sizeof out[i]);
    const JpegRequest jq{qt_luma, qt_chroma, flags, nullptr};
    return batch_host(in, n, opts, order, cfg, nullptr, &jq, out, statuses);
}

slimg_status slimg_b200_color_planes_batch_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h, int kind, void *y,
                                                  void *cb, void *cr, void *a, void *scratch, void *stream) {
    if (n && (!src || !y || !cb || !cr)) return fail(SLIMG_ERR_IMAGE, "null argument");
    if (kind != SLIMG_PLANES_WEBP_YUV420 && !a && n) return fail(SLIMG_ERR_IMAGE, "the AVIF planes include alpha");
    if (w == 0 || h == 0) return fail(SLIMG_ERR_ENCODE, "image has a zero dimension");
    slimg_status st = check_device();
    if (st != SLIMG_OK) return st;
    return run_planes_device(src, n, w, h, kind, y, cb, cr, a, scratch, static_cast<cudaStream_t>(stream));
}

// the WebP planes conversion with its scratch in a static __device__ array (k4_planes.cu): no cudaMalloc / cudaFree, no
// allocating routine of any library on this path -- only a memset, two kernel launches and one device-to-device copy
slimg_status slimg_b200_webp_planes_batch_device_static(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h, void *y,
                                                        void *u, void *v, void *a, uint32_t *any_alpha, void *stream) {
    if (n && (!src || !y || !u || !v || !a)) return fail(SLIMG_ERR_IMAGE, "null argument");
    if (w == 0 || h == 0) return fail(SLIMG_ERR_ENCODE, "image has a zero dimension");
    if (n > 65535) return fail(SLIMG_ERR_IMAGE, "at most 65535 images per call");
    if (webp_static_scratch_need(n, h) > kWebpStaticScratchBytes)
        return fail(SLIMG_ERR_IMAGE, "n * ceil(h/2) row flags + 4 n bytes exceed the static scratch (SLIMG_WEBP_STATIC_SCRATCH_BYTES = 262144): "
                                     "split the batch or use slimg_b200_color_planes_batch_device with caller scratch");
    slimg_status st = check_device();
    if (st != SLIMG_OK) return st;
    const size_t px = size_t(w) * h;
    WebpPlanesParams p{};
    p.src = src; p.src_img_bytes = px * 4; p.y_img_bytes = px;
    p.uv_img_bytes = size_t((w + 1) / 2) * ((h + 1) / 2);
    p.w = w; p.h = h;
    p.y = static_cast<uint8_t *>(y); p.u = static_cast<uint8_t *>(u); p.v = static_cast<uint8_t *>(v); p.a = static_cast<uint8_t *>(a);
    webp_gamma_tables(p.g2l, p.l2g);
    CU(launch_webp_planes_static(p, n, any_alpha, static_cast<cudaStream_t>(stream)));
    return SLIMG_OK;
}

slimg_status slimg_b200_pipeline_planes(const slimg_image *in, const slimg_pipeline_options *opts, int order, int kind,
                                        slimg_planes *out) {
    if (!opts || !out) return fail(SLIMG_ERR_IMAGE, "null argument");
    if (kind < SLIMG_PLANES_WEBP_YUV420 || kind > SLIMG_PLANES_AVIF_YCBCR444_10) return fail(SLIMG_ERR_IMAGE, "unknown plane kind");
    slimg_status st = validate_image(in, *opts, order);
    if (st != SLIMG_OK) return st;
    PipeGeom g;
    if ((st = build_geometry(in->width, in->height, *opts, order, g)) != SLIMG_OK) return st;
    if ((st = check_device()) != SLIMG_OK) return st;
    const bool identity = !opts->has_resize && !opts->has_crop && !opts->has_extend;
    const size_t src_bytes = size_t(g.sw) * g.sh * 4, dst_bytes = size_t(g.fw) * g.fh * 4, px = size_t(g.fw) * g.fh;
    const bool webp = kind == SLIMG_PLANES_WEBP_YUV420;
    const uint32_t cw = webp ? (g.fw + 1) / 2 : g.fw, chh = webp ? (g.fh + 1) / 2 : g.fh;
    const size_t bps = kind == SLIMG_PLANES_AVIF_YCBCR444_10 ? 2 : 1;
    const size_t y_bytes = px * bps, c_bytes = size_t(cw) * chh * bps;
    const size_t rows = (size_t(chh) + 3) & ~size_t(3);
    const size_t total = y_bytes * 2 + c_bytes * 2;                 // y, cb, cr, a
    cudaStream_t stream = cudaStreamPerThread;
    uint8_t *d_src = nullptr, *d_dst = nullptr, *d_pl = nullptr, *d_scr = nullptr;
    uint8_t *host = static_cast<uint8_t *>(std::malloc(total ? total : 1));
    if (!host) return fail(SLIMG_ERR_IMAGE, "out of host memory");
    auto cleanup = [&](slimg_status s) {
        if (d_src) cudaFreeAsync(d_src, stream);
        if (d_dst) cudaFreeAsync(d_dst, stream);
        if (d_pl) cudaFreeAsync(d_pl, stream);
        if (d_scr) cudaFreeAsync(d_scr, stream);
        if (s != SLIMG_OK) std::free(host);
        return s;
    };
    cudaError_t e;
    RingLease lease;
    if ((e = cudaMallocAsync((void **)&d_src, src_bytes, stream)) != cudaSuccess) return cleanup(cuda_fail(e, "cudaMallocAsync"));
    if (!identity && (e = cudaMallocAsync((void **)&d_dst, dst_bytes, stream)) != cudaSuccess) return cleanup(cuda_fail(e, "cudaMallocAsync"));
    if ((e = cudaMallocAsync((void **)&d_pl, total, stream)) != cudaSuccess) return cleanup(cuda_fail(e, "cudaMallocAsync"));
    if (webp && (e = cudaMallocAsync((void **)&d_scr, rows + 4, stream)) != cudaSuccess) return cleanup(cuda_fail(e, "cudaMallocAsync"));
    uint32_t r0 = 0, r1 = g.sh, c0 = 0, c1 = g.sw;
    if (!identity && (st = source_window(g, r0, r1, c0, c1)) != SLIMG_OK) return cleanup(st);
    if (r1 > r0 && c1 > c0) {
        const size_t off = size_t(r0) * g.sw * 4;
        if ((e = copy_h2d(d_src + off, in->data + off, size_t(r1 - r0) * g.sw * 4, stream, lease)) != cudaSuccess) return cleanup(cuda_fail(e, "H2D"));
    }
    const uint8_t *d_pixels = d_src;
    if (!identity) {
        if ((st = run_pipeline_device(g, d_src, 1, d_dst, stream)) != SLIMG_OK) return cleanup(st);
        d_pixels = d_dst;
    }
    uint8_t *dy = d_pl, *dcb = dy + y_bytes, *dcr = dcb + c_bytes, *da = dcr + c_bytes;
    if ((st = run_planes_device(d_pixels, 1, g.fw, g.fh, kind, dy, dcb, dcr, da, d_scr, stream)) != SLIMG_OK) return cleanup(st);
    uint32_t any_alpha = 1;
    if ((e = copy_d2h(host, d_pl, total, stream, lease)) != cudaSuccess) return cleanup(cuda_fail(e, "D2H"));
    if (webp && (e = cudaMemcpyAsync(&any_alpha, d_scr + rows, 4, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return cleanup(cuda_fail(e, "D2H"));
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return cleanup(cuda_fail(e, "plane kernels"));
    cleanup(SLIMG_OK);
    std::memset(out, 0, sizeof *out);
    out->width = g.fw; out->height = g.fh; out->kind = uint32_t(kind);
    out->chroma_width = cw; out->chroma_height = chh; out->bytes_per_sample = uint32_t(bps);
    out->has_alpha = any_alpha ? 1 : 0;
    out->y = host; out->cb = host + y_bytes; out->cr = host + y_bytes + c_bytes;
    out->a = (webp && !any_alpha) ? nullptr : host + y_bytes + 2 * c_bytes;
    return SLIMG_OK;
}

void slimg_b200_planes_free(slimg_planes *p) {
    if (p && p->y) { std::free(p->y); p->y = p->cb = p->cr = p->a = nullptr; }
}

int32_t slimg_b200_last_resize_path(void) { return t_last_path; }
uint64_t slimg_b200_path_launches(int32_t path) { return path >= 0 && path < 8 ? g_path_launches[path].load() : 0; }

}  // extern "C"
