// tileplan.cpp -- host-side plan of the integer tile kernel (k2_tile.cu): Lanczos3 at ratios below ~3.5x and
// upscales, where an output tile's whole source footprint fits in shared memory.
//
// Same arithmetic as the V-first pipeline (vplan.cpp): both passes are u8 x s8 -> s32 tensor-core GEMMs over the
// fixed-point weights of taps.cpp (sum preserved exactly, two balanced base-256 digits); the vertical result is
// re-quantised to 16 bits (7 fractional bits, offset 64) and enters the horizontal pass as two u8 limbs.  What differs
// is the skeleton: no ring, no bands -- one CTA loads the footprint of a TW x TH output tile, runs both passes out of
// shared memory and stores the tile, and several such CTAs per SM hide each other's latencies.  That is what small
// ratios want: at 2x a group of 8 output rows is only 16 new source rows, so the V-first pipeline's per-group
// hand-offs (~1100 clk) dominate (profiles/r01f_configs.txt: 358 Gpx/s at 2x, 221 at 1.5x), and an upscale has no
// band structure at all.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "host.h"
#include "kernels.cuh"

#define TILE_FAIL(code) do { if (std::getenv("SLIMG_B200_DEBUG")) std::fprintf(stderr, "[slimg_b200] tileplan rejects the geometry at check %d\n", code); return; } while (0)

namespace slimg {

namespace {
constexpr int kFrac = 7, kMaxCh = 4;
constexpr int kOffset = 96;   // added to the 16-bit intermediate so undershoot stays >= 0: at some upscale phases the negative taps sum to -0.27 (x 255 = 69)
const int kPerm[4] = {0, 2, 1, 3};   // k-slot -> pixel inside a 4-pixel block (pass-1 epilogue of k2_vfirst.cu / k2_tile.cu)
}

void build_tileplan(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, bool forced, TilePlanHost &out) {
    out = TilePlanHost();
    const int ndig = 2;
    if (!g.has_resize) TILE_FAIL(1);
    if (g.sw % 4 != 0 || g.sw < 4) TILE_FAIL(2);           // 16-byte aligned source rows
    if (g.pw == 0 || g.ph == 0 || g.rw == 0 || g.rh == 0) TILE_FAIL(3);
    const double ry = double(g.vh) / double(g.Rh), rx = double(g.vw) / double(g.Rw);
    // Who gets the tile kernel: upscales and mixed resizes (one axis up or near 1, the other down by at most 3.5x).  A
    // downscale on BOTH axes streams better through the banded pipeline even at 1.5x (measured on B200, 24 MP source:
    // 1.5x 138 us here / 110 us V-first, 2x 66-97 / 69, 3x 56 / 39); the f32 kernels it replaces are 2-3x slower
    // (1024^2 x4: 86 us few-tap f32 -> 54 us; 2048^2 x2: 184 -> 76 us).
    if (!forced && (ry > 3.5 || rx > 3.5 || (ry > 1.25 && rx > 1.25))) TILE_FAIL(4);

    // output tile: as large as the footprint allows (a bigger tile re-reads less halo and amortises the weights)
    const bool up = ry <= 1.0 && rx <= 1.0;
    uint32_t tw = up ? 128 : 32, th = up ? 64 : 32;    // upscales: 128 x 64 measured best (fewer CTAs: the per-tile set-up is a quarter of the work)
    if (const char *e = std::getenv("SLIMG_B200_TILE_W")) tw = std::max(8, std::min(128, std::atoi(e) / 8 * 8));
    if (const char *e = std::getenv("SLIMG_B200_TILE_H")) th = std::max(8, std::min(64, std::atoi(e) / 8 * 8));
    out.tw = tw; out.th = th;

    // ---------------- vertical pass: groups of 8 kept output rows ----------------
    float maxw_y = 0.0f;
    for (uint32_t i = 0; i < g.ph; i++) {
        const uint32_t o = g.py + i;
        for (int k = 0; k < ty.count[o]; k++) maxw_y = std::max(maxw_y, std::fabs(ty.w[size_t(o) * ty.max_taps + k]));
    }
    if (!(maxw_y > 0.0f)) TILE_FAIL(5);
    const int Sy = pick_scale(maxw_y, ndig);
    if (Sy < kFrac + 2) TILE_FAIL(6);
    out.sh1 = Sy - kFrac;
    const uint32_t n_groups = (g.ph + 7) / 8;
    std::vector<int32_t> a0(n_groups, 0), need(n_groups, 0), nch(n_groups, 0);
    out.row_computed.assign(g.ph, 0);
    for (uint32_t j = 0; j < n_groups; j++) {
        int lo = INT32_MAX, hi = -1;
        for (uint32_t n = 0; n < 8 && j * 8 + n < g.ph; n++) {
            const uint32_t o = g.py + j * 8 + n;
            if (ty.count[o] == 0) continue;
            lo = std::min(lo, ty.left[o]);
            hi = std::max(hi, ty.left[o] + ty.count[o]);
        }
        if (hi < 0) continue;
        a0[j] = lo; need[j] = hi; nch[j] = (hi - lo + 31) / 32;
        out.nch1 = std::max(out.nch1, nch[j]);
    }
    if (out.nch1 == 0 || out.nch1 > kMaxCh) TILE_FAIL(7);
    // group record: {a0, need, chunks, -, ..., bias[8][4] from word 8} + [nch1][2 digits][32 lanes][2] weight fragments
    out.rec_words = 64 + uint32_t(out.nch1) * ndig * 64;
    out.recs.assign(size_t(n_groups) * out.rec_words, 0u);
    std::vector<int64_t> q;
    for (uint32_t j = 0; j < n_groups; j++) {
        uint32_t *r = &out.recs[size_t(j) * out.rec_words];
        r[0] = uint32_t(a0[j]); r[1] = uint32_t(need[j]); r[2] = uint32_t(nch[j]);
        for (uint32_t n = 0; n < 8; n++) {
            const uint32_t i = j * 8 + n;
            if (i >= g.ph) break;
            const uint32_t o = g.py + i;
            const int cnt = ty.count[o];
            quantise(&ty.w[size_t(o) * ty.max_taps], cnt, Sy, q);
            int64_t pos = 0, neg = 0;
            for (int k = 0; k < cnt; k++) {
                (q[k] > 0 ? pos : neg) += q[k];
                int8_t d[3];
                if (!split_digits(q[k], ndig, d)) TILE_FAIL(8);
                const int row = ty.left[o] + k - a0[j];
                for (int l = 0; l < ndig; l++) put_b(r + 64 + ((row / 32) * ndig + l) * 64, row % 32, int(n), d[l]);
            }
            out.row_computed[i] = cnt > 0;
            for (int c = 0; c < 4; c++) {       // bias = fill term + offset + rounding half; the 16-bit intermediate must stay in range
                const int64_t fillterm = std::llround(double(g.fill_pre[c]) * double(ty.fillw[o]) * std::ldexp(1.0, Sy));
                const int64_t bias = fillterm + (int64_t(kOffset) << Sy) + (int64_t(1) << (out.sh1 - 1));
                const int64_t lo_x = 255 * neg + bias, hi_x = 255 * pos + bias;
                if (lo_x < 0 || (hi_x >> out.sh1) > 65535 || hi_x > INT32_MAX || bias > INT32_MAX) TILE_FAIL(9);
                r[8 + n * 4 + c] = uint32_t(int32_t(bias));
            }
        }
    }
    // row blocks of th rows: the source-rectangle rows they stage
    out.n_rowblk = (g.ph + th - 1) / th;
    out.rb_y0.assign(out.n_rowblk, 0);
    out.rb_rows.assign(out.n_rowblk, 0);
    for (uint32_t b = 0; b < out.n_rowblk; b++) {
        int lo = INT32_MAX, hi = -1;
        for (uint32_t j = b * (th / 8); j < std::min(n_groups, (b + 1) * (th / 8)); j++) {
            if (nch[j] == 0) continue;
            lo = std::min(lo, a0[j]); hi = std::max(hi, need[j]);
        }
        if (hi < 0) continue;
        out.rb_y0[b] = lo; out.rb_rows[b] = hi - lo;
        out.rows_max = std::max<uint32_t>(out.rows_max, uint32_t(hi - lo));
    }
    if (out.rows_max == 0) TILE_FAIL(10);

    // ---------------- horizontal pass: column blocks of tw kept columns, tiles of 8 ----------------
    float maxw_x = 0.0f;
    for (uint32_t j = 0; j < g.pw; j++) {
        const uint32_t o = g.px + j;
        for (int k = 0; k < tx.count[o]; k++) maxw_x = std::max(maxw_x, std::fabs(tx.w[size_t(o) * tx.max_taps + k]));
    }
    if (!(maxw_x > 0.0f)) TILE_FAIL(11);
    const int Sx = pick_scale(maxw_x, ndig);
    if (Sx < 10) TILE_FAIL(12);
    // pass-2 result in integers: t = (acc[lo][d0] >> 8) + acc[lo][d1] + acc[hi][d0] + (acc[hi][d1] << 8) + cxi is the output
    // in units of 2^-(Sx + kFrac - 8); the byte is t >> sh2 (round-half-up folded into cxi), saturated
    out.sh2 = Sx + kFrac - 8;
    auto span_of = [&](uint32_t c0, uint32_t c1, int64_t &lo, int64_t &hi) {
        lo = INT64_MAX; hi = -1;
        for (uint32_t j = c0; j < c1; j++) {
            const uint32_t o = g.px + j;
            if (tx.count[o] == 0) continue;
            lo = std::min<int64_t>(lo, int64_t(g.rx) + tx.left[o]);
            hi = std::max<int64_t>(hi, int64_t(g.rx) + tx.left[o] + tx.count[o]);
        }
    };
    out.n_colblk = (g.pw + tw - 1) / tw;
    const uint32_t tiles_per_blk = tw / 8;
    out.cb_src0.assign(out.n_colblk, 0);
    out.cb_nblk.assign(out.n_colblk, 0);
    out.tile_kb0.assign(size_t(out.n_colblk) * tiles_per_blk, 0);
    out.cxi.assign(size_t(g.pw) * 4, 0);
    int64_t max_nblk = 1;
    bool packable = true;                            // every tile's taps inside 16 staged pixels: both limbs share one K = 32
    for (uint32_t s = 0; s < out.n_colblk; s++) {
        const uint32_t c0 = s * tw, c1 = std::min(g.pw, c0 + tw);
        int64_t lo, hi;
        span_of(c0, c1, lo, hi);
        if (hi < 0) continue;                        // every tap of the block lands on fill
        const int64_t s0 = lo & ~int64_t(3);
        int64_t nblk = (hi - s0 + 3) / 4;
        nblk = (nblk + 1) & ~int64_t(1);             // pass 1 works on pairs of 16-byte blocks
        out.cb_src0[s] = int32_t(s0);
        out.cb_nblk[s] = int32_t(nblk);
        max_nblk = std::max(max_nblk, nblk);
        for (uint32_t t = 0; t * 8 < c1 - c0; t++) {
            int64_t tlo, thi;
            span_of(c0 + t * 8, std::min(c1, c0 + t * 8 + 8), tlo, thi);
            if (thi < 0) continue;
            const int64_t kb0 = (tlo - s0) / 4;
            out.tile_kb0[s * tiles_per_blk + t] = int32_t(kb0);
            out.nch2 = std::max(out.nch2, int((thi - s0 - kb0 * 4 + 31) / 32));
            if (thi - s0 - kb0 * 4 > 16) packable = false;
        }
    }
    if (out.nch2 == 0 || out.nch2 > kMaxCh) TILE_FAIL(13);
    // every tile runs nch2 chunks: V rows are allocated so that kb0 + 8 * nch2 words always exist
    out.v_words = uint32_t(std::max<int64_t>(max_nblk, 0));
    for (uint32_t s = 0; s < out.n_colblk; s++)
        for (uint32_t t = 0; t < tiles_per_blk; t++)
            out.v_words = std::max<uint32_t>(out.v_words, uint32_t(out.tile_kb0[s * tiles_per_blk + t]) + 8u * uint32_t(out.nch2));
    // shared-memory geometry: source rows at an odd multiple of 16 bytes (conflict-free ldmatrix rows), V rows at
    // 16 * odd bytes with room for the 3-word channel skew
    {
        const int64_t units = max_nblk | 1;
        out.pitch = uint32_t(units * 16);
        int64_t qv = (int64_t(out.v_words) + 3 + 3) / 4;
        if (qv % 2 == 0) qv++;
        out.pitch_v = uint32_t(qv * 16);
        out.smem_bytes = out.rows_max * out.pitch + (th / 8) * 2u * 32u * out.pitch_v + 16;
    }
    if (out.smem_bytes > 100 * 1024) TILE_FAIL(14);         // two CTAs per SM at least
    out.packed = packable && out.nch2 == 1 && !std::getenv("SLIMG_B200_TILE_NOPACK");
    // packed: per tile three fragments -- A [w0 | 0], B [w1 | w0], C [0 | w1] over K = [16 low-limb px | 16 high-limb px]
    out.b2.assign(size_t(out.n_colblk) * tiles_per_blk * (out.packed ? 3 : out.nch2 * ndig) * 64, 0u);
    for (uint32_t s = 0; s < out.n_colblk; s++) {
        const uint32_t c0 = s * tw, c1 = std::min(g.pw, c0 + tw);
        const int64_t s0 = out.cb_src0[s];
        for (uint32_t j = c0; j < c1; j++) {
            const uint32_t o = g.px + j, t = (j - c0) / 8, n = (j - c0) % 8;
            const int cnt = tx.count[o];
            quantise(&tx.w[size_t(o) * tx.max_taps], cnt, Sx, q);
            int64_t sum_q = 0;
            const int64_t kb0 = out.tile_kb0[s * tiles_per_blk + t];
            for (int k = 0; k < cnt; k++) {
                sum_q += q[k];
                int8_t d[3];
                if (!split_digits(q[k], ndig, d)) TILE_FAIL(15);
                const int64_t rel = int64_t(g.rx) + tx.left[o] + k - s0;      // pixel inside the staged span
                const int64_t blk = rel / 4 - kb0;
                int slot = 0;
                for (int z = 0; z < 4; z++) if (kPerm[z] == int(rel % 4)) slot = z;
                const int64_t kk = blk * 4 + slot;                            // k index inside the tile's chunks
                if (kk < 0 || kk >= 32 * out.nch2) TILE_FAIL(16);
                if (out.packed) {
                    uint32_t *f = &out.b2[(size_t(s) * tiles_per_blk + t) * 3 * 64];
                    put_b(f, int(kk), int(n), d[0]);                                        // low limb x digit 0
                    put_b(f + 64, int(kk), int(n), d[1]); put_b(f + 64, 16 + int(kk), int(n), d[0]);
                    put_b(f + 128, 16 + int(kk), int(n), d[1]);                             // high limb x digit 1
                } else
                    for (int l = 0; l < ndig; l++)
                        put_b(&out.b2[(((size_t(s) * tiles_per_blk + t) * out.nch2 + size_t(kk / 32)) * ndig + l) * 64], int(kk % 32), int(n), d[l]);
            }
            // constant term: fill taps of the horizontal pass minus the intermediate's offset, plus the rounding half,
            // in units of 2^-sh2
            const double off = -double(kOffset) * double(sum_q) * std::ldexp(1.0, -Sx);
            for (int c = 0; c < 4; c++) {
                const double cxv = off + double(g.fill_pre[c]) * double(tx.fillw[o]);
                out.cxi[size_t(j) * 4 + c] = int32_t(std::llround(cxv * std::ldexp(1.0, out.sh2))) + (int32_t(1) << (out.sh2 - 1));
            }
        }
    }
    for (uint32_t i = 0; i < g.ph; i++)
        if (!out.row_computed[i]) out.need_fill = out.rows_pure_fill = true;
    if (g.pw != g.fw || g.ph != g.fh) out.need_fill = true;
    out.ok = true;
}

}  // namespace slimg
