// plan_probe -- prints what the host-side planners decide for a geometry, no GPU needed.
// build: nvcc -std=c++17 -o /tmp/plan_probe tools/plan_probe.cu slimg_b200/csrc/{geometry,taps,vplan,tcplan,tileplan,fastplan}.cpp -x cu
// usage: plan_probe W H resize_w [chain]    (Width(resize_w); with "chain": Width -> crop 16:9 -> extend 1:1)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include "../slimg_b200/csrc/host.h"
#include "../slimg_b200/csrc/kernels.cuh"
namespace slimg { void count_launch(uint64_t) {} void note_path(int) {} }
using namespace slimg;
int main(int argc, char **argv) {
    if (argc < 4) return 1;
    const uint32_t w = atoi(argv[1]), h = atoi(argv[2]);
    slimg_pipeline_options o{};
    o.has_resize = 1;
    if (argv[3][0] == 's') { o.resize.tag = SLIMG_RESIZE_SCALE; o.resize.factor = atof(argv[3] + 1); }
    else { o.resize.tag = SLIMG_RESIZE_WIDTH; o.resize.a = atoi(argv[3]); }
    const bool chain = argc > 4 && !strcmp(argv[4], "chain");
    if (chain) {
        o.has_crop = 1; o.crop.tag = SLIMG_CROP_ASPECT_RATIO; o.crop.width = 16; o.crop.height = 9;
        o.has_extend = 1; o.extend.tag = SLIMG_EXTEND_ASPECT_RATIO; o.extend.width = 1; o.extend.height = 1;
    }
    PipeGeom g;
    if (build_geometry(w, h, o, chain ? SLIMG_ORDER_CHAIN : SLIMG_ORDER_CONVERT, g) != SLIMG_OK) { printf("geometry: %s\n", last_error_string().c_str()); return 1; }
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t0 = now();
    AxisTaps tx, ty;
    build_axis_taps(g.vw, g.Rw, g.vox, g.rw, tx);
    build_axis_taps(g.vh, g.Rh, g.voy, g.rh, ty);
    double t1 = now();
    printf("geom: S %ux%u rect (%u,%u,%u,%u) V %ux%u R %ux%u keep (%u,%u,%u,%u) F %ux%u  taps x %d y %d\n", g.sw, g.sh, g.rx, g.ry, g.rw, g.rh,
           g.vw, g.vh, g.Rw, g.Rh, g.px, g.py, g.pw, g.ph, g.fw, g.fh, tx.max_taps, ty.max_taps);
    TilePlanHost lp; build_tileplan(g, tx, ty, true, lp);
    double t2 = now();
    printf("tile: ok=%d tile=%ux%u blocks=%ux%u rows_max=%u nch1=%d nch2=%d pitch=%u pitch_v=%u v_words=%u smem=%u sh1=%d\n", int(lp.ok), lp.tw, lp.th,
           lp.n_colblk, lp.n_rowblk, lp.rows_max, lp.nch1, lp.nch2, lp.pitch, lp.pitch_v, lp.v_words, lp.smem_bytes, lp.sh1);
    VPlanHost vp; build_vplan(g, tx, ty, 2, vp);
    double t3 = now();
    printf("vfirst: ok=%d strips=%u cols=%d unit=%d upw=%d nunits=%d co=%u vbufs=%u ring=%u pitch=%u smem=%u nch1=%d nch2=%d recs=%u\n", int(vp.ok), vp.n_strips,
           vp.ok ? vp.strip_col0[1] : 0, vp.unit, vp.ok ? vp.strip_upw[0] : 0, vp.ok ? vp.strip_nunits[0] : 0, vp.cohorts, vp.vbufs, vp.ring_rows, vp.pitch, vp.smem_bytes, vp.nch1, vp.nch2, vp.n_recs);
    TcPlanHost tp; build_tcplan(g, tx, ty, tp);
    double t4 = now();
    { double t5 = now(); FastPlanHost fp; build_fast_plan(g, tx, ty, fp); printf("fastplan ms %.2f ok=%d\n", now() - t5, int(fp.ok)); }
    printf("host ms: taps %.2f  tileplan %.2f  vplan %.2f  tcplan %.2f\n", t1 - t0, t2 - t1, t3 - t2, t4 - t3);
    printf("tc: ok=%d strips=%u cols=%d panels=%u nbuf=%u vbufs=%u slots=%u blocks=%u chunks=%u nb_max=%u nch2=%d pitch_v=%u smem=%u\n", int(tp.ok), tp.n_strips,
           tp.ok ? tp.strip_col0[1] : 0, tp.n_panels, tp.nbuf, tp.vbufs, tp.slots, tp.n_blocks, tp.n_chunks, tp.nb_max, tp.nch2, tp.pitch_v, tp.smem_bytes);
    return 0;
}
