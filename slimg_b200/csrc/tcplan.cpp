// tcplan.cpp -- host-side plan of the tcgen05 resize pipeline (k2_tc.cu).
//
// Vertical pass (input-stationary): the source rectangle streams through shared memory in chunks of 32 rows.
// For every chunk the plan lists the blocks of 8 kept output rows whose filter windows touch it, with one weight
// tile per block: B[N = 16: digit 0 of the 8 rows, digit 1 of the 8 rows][32 source rows] in s8, laid out as the
// canonical K-major, no-swizzle operand of tcgen05.mma (8x16-byte core matrices).  The kernel multiplies the chunk
// (A operand, M = 128 bytes of the row, MN-major, straight from the TMA-written ring) with each tile and accumulates
// in TMEM; a block's accumulators are complete after its last chunk.  Blocks of 8 rows (round 1 had 16) are what the
// horizontal pass consumes at a time, and halve the TMEM columns of a block: four accumulator buffers per panel
// instead of two, so the MMA stream runs two blocks ahead of the epilogue.
// Horizontal pass: as in vplan.cpp (mma.sync on the two u8 limbs of the 16-bit intermediate), pixels in
// natural order.
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "host.h"

namespace slimg {

namespace {
constexpr int kFrac = 7;        // fractional bits of the 16-bit intermediate
constexpr int kOffset = 64;     // added to the intermediate so Lanczos undershoot stays >= 0
constexpr int kMaxCh = 4;
constexpr int kMaxPanels = 8;   // 8 panels x 2 accumulator buffers x 32 columns = all 512 TMEM columns
constexpr size_t kSmemBudget = 227 * 1024 - 1024;
constexpr uint32_t kTileBytes = 512, kLboB = 128, kSboB = 256;
constexpr uint32_t kBlkRows = 8;   // kept output rows per accumulator block
}  // namespace

void build_tcplan(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, TcPlanHost &out) {
    out = TcPlanHost();
    const int ndig = out.ndig;
    if (!g.has_resize) return;
    if (g.vw < g.Rw || g.vh < g.Rh) return;          // downscale (ratio >= 1) on both axes
    if (g.sw % 4 != 0 || g.sw < 4) return;           // 16-byte aligned rows
    if (g.pw == 0 || g.ph == 0 || g.rw == 0 || g.rh == 0) return;

    // ---------------- vertical pass ----------------
    float maxw_y = 0.0f;
    for (uint32_t i = 0; i < g.ph; i++) {
        const uint32_t o = g.py + i;
        if (ty.count[o] == 0) return;                // rows that are pure fill: the mma.sync kernel handles those
        for (int k = 0; k < ty.count[o]; k++) maxw_y = std::max(maxw_y, std::fabs(ty.w[size_t(o) * ty.max_taps + k]));
    }
    if (!(maxw_y > 0.0f)) return;
    const int Sy = pick_scale(maxw_y, ndig);
    if (Sy < kFrac + 2) return;
    out.sh1 = Sy - kFrac;

    out.n_blocks = (g.ph + kBlkRows - 1) / kBlkRows;
    out.blk_c0.assign(out.n_blocks, 0);
    out.blk_c1.assign(out.n_blocks, 0);
    out.bias.assign(size_t(out.n_blocks) * kBlkRows * 4, 0);
    int prev_c0 = 0, prev_c1 = 0;
    for (uint32_t b = 0; b < out.n_blocks; b++) {
        int lo = INT32_MAX, hi = -1;
        for (uint32_t n = 0; n < kBlkRows && b * kBlkRows + n < g.ph; n++) {
            const uint32_t o = g.py + b * kBlkRows + n;
            lo = std::min(lo, ty.left[o]);
            hi = std::max(hi, ty.left[o] + ty.count[o]);
        }
        out.blk_c0[b] = lo / 32;
        out.blk_c1[b] = (hi - 1) / 32;
        if (b && (out.blk_c0[b] < prev_c0 || out.blk_c1[b] < prev_c1)) return;    // windows move down monotonically
        prev_c0 = out.blk_c0[b];
        prev_c1 = out.blk_c1[b];
    }
    out.n_chunks = uint32_t(out.blk_c1[out.n_blocks - 1]) + 1;
    // blocks touching each chunk: a contiguous range
    std::vector<int> first(out.n_chunks, -1), cnt(out.n_chunks, 0);
    for (uint32_t b = 0; b < out.n_blocks; b++)
        for (int c = out.blk_c0[b]; c <= out.blk_c1[b]; c++) {
            if (first[c] < 0) first[c] = int(b);
            if (int(b) != first[c] + cnt[c]) return;
            cnt[c]++;
            out.nb_max = std::max<uint32_t>(out.nb_max, uint32_t(cnt[c]));
        }
    if (out.nb_max == 0 || out.nb_max > 4) return;
    out.crec_words = 16 + out.nb_max * (kTileBytes / 4);
    out.crecs.assign(size_t(out.n_chunks) * out.crec_words, 0u);
    for (uint32_t c = 0; c < out.n_chunks; c++) {
        uint32_t *r = &out.crecs[size_t(c) * out.crec_words];
        r[0] = uint32_t(std::max(first[c], 0));
        r[1] = uint32_t(cnt[c]);
        r[3] = c;
        for (int i = 0; i < cnt[c]; i++) {
            const int b = first[c] + i;
            if (out.blk_c0[b] == int(c)) r[2] |= 1u << i;
            if (out.blk_c1[b] == int(c)) r[2] |= 1u << (8 + i);
        }
    }
    std::vector<int64_t> q;
    for (uint32_t i = 0; i < g.ph; i++) {
        const uint32_t o = g.py + i, b = i / kBlkRows, n = i % kBlkRows;
        const int cntw = ty.count[o];
        quantise(&ty.w[size_t(o) * ty.max_taps], cntw, Sy, q);
        int64_t pos = 0, neg = 0;
        for (int k = 0; k < cntw; k++) {
            (q[k] > 0 ? pos : neg) += q[k];
            int8_t d[3];
            if (!split_digits(q[k], ndig, d)) return;
            const int row = ty.left[o] + k, c = row / 32, kk = row % 32;
            uint32_t *r = &out.crecs[size_t(c) * out.crec_words];
            const int bi = int(b) - int(r[0]);
            if (bi < 0 || bi >= int(r[1])) return;
            for (int l = 0; l < ndig; l++) {      // N index of the tile = 8 * digit + row
                uint8_t *tile = reinterpret_cast<uint8_t *>(r + 16) + size_t(bi) * kTileBytes;
                tile[(kk / 16) * kLboB + uint32_t(l) * kSboB + n * 16 + (kk % 16)] = uint8_t(d[l]);
            }
        }
        // bias = fill term + offset + rounding half; the 16-bit intermediate must stay in range
        for (int ch = 0; ch < 4; ch++) {
            const int64_t fillterm = std::llround(double(g.fill_pre[ch]) * double(ty.fillw[o]) * std::ldexp(1.0, Sy));
            const int64_t bias = fillterm + (int64_t(kOffset) << Sy) + (int64_t(1) << (out.sh1 - 1));
            const int64_t lo_x = 255 * neg + bias, hi_x = 255 * pos + bias;
            if (lo_x < 0 || (hi_x >> out.sh1) > 65535 || hi_x > INT32_MAX || bias > INT32_MAX) return;
            out.bias[size_t(i) * 4 + ch] = int32_t(bias);
        }
    }

    // ---------------- horizontal pass ----------------
    float maxw_x = 0.0f;
    for (uint32_t j = 0; j < g.pw; j++) {
        const uint32_t o = g.px + j;
        for (int k = 0; k < tx.count[o]; k++) maxw_x = std::max(maxw_x, std::fabs(tx.w[size_t(o) * tx.max_taps + k]));
    }
    if (!(maxw_x > 0.0f)) return;
    const int Sx = pick_scale(maxw_x, ndig);
    if (Sx < 8) return;
    for (int l = 0; l < 2; l++)
        for (int d = 0; d < 3; d++) out.s2[l][d] = d < ndig ? float(std::ldexp(1.0, 8 * l + 8 * d - Sx - kFrac)) : 0.0f;

    auto span_of = [&](uint32_t c0, uint32_t c1, int64_t &lo, int64_t &hi) {
        lo = INT64_MAX; hi = -1;
        for (uint32_t j = c0; j < c1; j++) {
            const uint32_t o = g.px + j;
            if (tx.count[o] == 0) continue;
            lo = std::min<int64_t>(lo, int64_t(g.rx) + tx.left[o]);
            hi = std::max<int64_t>(hi, int64_t(g.rx) + tx.left[o] + tx.count[o]);
        }
    };
    // widest strips (multiple of 8 columns, at most 32 = four pass-2 warps) that need at most kMaxPanels panels
    const char *cols_env = std::getenv("SLIMG_B200_STRIP_COLS");
    uint32_t cols = 0, npan = 0;
    for (uint32_t cand = 32; cand >= 8 && cols == 0; cand -= 8) {
        if (cols_env && uint32_t(std::atoi(cols_env)) != cand) continue;
        uint32_t worst = 1;
        for (uint32_t c0 = 0; c0 < g.pw; c0 += cand) {
            int64_t lo, hi;
            span_of(c0, std::min(g.pw, c0 + cand), lo, hi);
            if (hi < 0) continue;
            worst = std::max<uint32_t>(worst, uint32_t((hi - (lo & ~int64_t(3)) + 31) / 32));
        }
        if (worst <= uint32_t(kMaxPanels)) { cols = cand; npan = worst; }
    }
    if (cols == 0) return;
    out.n_panels = npan;
    out.n_strips = (g.pw + cols - 1) / cols;
    out.nbuf = std::min<uint32_t>(8, 512 / (16 * npan));      // accumulator buffers: 16 TMEM columns per (panel, block)
    if (out.nbuf < out.nb_max + 1) return;                     // every block a chunk touches, plus one being drained

    const int64_t nblk = int64_t(npan) * 8;                     // 16-byte (4-pixel) blocks of a V row
    out.strip_col0.assign(out.n_strips + 1, 0);
    out.strip_src0.assign(out.n_strips, 0);
    out.tile_kb0.assign(size_t(out.n_strips) * 8, 0);
    out.cx.assign(size_t(g.pw) * 4, 0.0f);
    std::vector<int64_t> strip_s0(out.n_strips, 0);
    for (uint32_t s = 0; s < out.n_strips; s++) {
        const uint32_t c0 = s * cols, c1 = std::min(g.pw, c0 + cols);
        out.strip_col0[s] = int32_t(c0);
        out.strip_col0[s + 1] = int32_t(c1);
        int64_t lo, hi;
        span_of(c0, c1, lo, hi);
        if (hi < 0) continue;                        // every tap of the strip lands on fill
        const int64_t s0 = lo & ~int64_t(3);
        strip_s0[s] = s0;
        out.strip_src0[s] = int32_t(s0);
        for (uint32_t t = 0; t * 8 < c1 - c0; t++) {
            int64_t tlo, thi;
            span_of(c0 + t * 8, std::min(c1, c0 + t * 8 + 8), tlo, thi);
            if (thi < 0) continue;
            const int64_t kb0 = (tlo - s0) / 4;
            const int64_t nch = (thi - s0 - kb0 * 4 + 31) / 32;
            out.tile_kb0[s * 8 + t] = int32_t(kb0);
            out.nch2 = std::max(out.nch2, int(nch));
        }
    }
    if (out.nch2 == 0 || out.nch2 > kMaxCh || 8 * out.nch2 > nblk) return;
    for (uint32_t s = 0; s < out.n_strips; s++)
        for (int t = 0; t < 8; t++) {
            int32_t &kb0 = out.tile_kb0[s * 8 + t];
            kb0 = std::max<int32_t>(0, std::min<int32_t>(kb0, int32_t(nblk) - 8 * out.nch2));
        }
    out.b2.assign(size_t(out.n_strips) * 8 * out.nch2 * ndig * 64, 0u);
    for (uint32_t s = 0; s < out.n_strips; s++) {
        const uint32_t c0 = uint32_t(out.strip_col0[s]), c1 = uint32_t(out.strip_col0[s + 1]);
        const int64_t s0 = strip_s0[s];
        for (uint32_t j = c0; j < c1; j++) {
            const uint32_t o = g.px + j, t = (j - c0) / 8, n = (j - c0) % 8;
            const int cntw = tx.count[o];
            quantise(&tx.w[size_t(o) * tx.max_taps], cntw, Sx, q);
            int64_t sum_q = 0;
            const int64_t kb0 = out.tile_kb0[s * 8 + t];
            for (int k = 0; k < cntw; k++) {
                sum_q += q[k];
                int8_t d[3];
                if (!split_digits(q[k], ndig, d)) return;
                const int64_t kk = int64_t(g.rx) + tx.left[o] + k - s0 - kb0 * 4;     // k index inside the tile's chunks
                if (kk < 0 || kk >= 32 * out.nch2) return;
                for (int l = 0; l < ndig; l++)
                    put_b(&out.b2[(((size_t(s) * 8 + t) * out.nch2 + size_t(kk / 32)) * ndig + l) * 64], int(kk % 32), int(n), d[l]);
            }
            // constant term: fill taps of the horizontal pass minus the intermediate's offset
            const double off = -double(kOffset) * double(sum_q) * std::ldexp(1.0, -Sx);
            for (int c = 0; c < 4; c++)
                out.cx[size_t(j) * 4 + c] = float(off + double(g.fill_pre[c]) * double(tx.fillw[o]));
        }
    }

    // ---------------- shared memory ----------------
    int64_t qv = (nblk + 6 + 3) / 4;                            // 16-byte units per V row = odd >= (nblk*4 + 24) / 16 (2-word channel skew)
    if (qv % 2 == 0) qv++;
    out.pitch_v = uint32_t(qv * 16);
    out.slot_bytes = npan * 4096;
    const size_t crec_bytes = size_t(out.crec_words) * 4;
    const char *vb_env = std::getenv("SLIMG_B200_TC_VBUFS");
    out.vbufs = vb_env ? uint32_t(std::max(2, std::min(4, std::atoi(vb_env)))) : 3u;
    const size_t fixed = size_t(2 * out.vbufs) * 32 * out.pitch_v + 1024;
    const char *slots_env = std::getenv("SLIMG_B200_TC_SLOTS");
    uint32_t slots = slots_env ? uint32_t(std::atoi(slots_env)) : 12;
    while (slots > 2 && slots * (out.slot_bytes + crec_bytes) + fixed > kSmemBudget) slots--;
    if (slots * (out.slot_bytes + crec_bytes) + fixed > kSmemBudget) return;
    out.slots = slots;
    out.smem_bytes = uint32_t(slots * (out.slot_bytes + crec_bytes) + fixed);
    if (g.pw != g.fw || g.ph != g.fh) out.need_fill = true;
    out.ok = true;
}

}  // namespace slimg
