// vplan.cpp -- host-side plan of the V-first integer tensor pipeline (k2_vfirst.cu).
//
// Both passes of the Lanczos3 downscale are integer GEMMs on the tensor cores (u8 x s8 -> s32):
//   pass 1 (vertical)   D1[byte of the row][8 output rows]      = sum over source rows
//   pass 2 (horizontal) D2[(channel,row)][8 output columns]     = sum over source pixels
// The f32 weights of taps.cpp are turned into fixed point (S fractional bits) and split into NDIG
// balanced base-256 digits (s8).  The pass-1 result is re-quantised to 16 bits (7 fractional bits,
// offset so it is never negative) and enters pass 2 as two u8 limbs.  All arithmetic up to the final
// float recombination is exact integer arithmetic, so the only errors are the weight quantisation
// (<= 2^-(S+1) per tap) and the 2^-8 rounding of the intermediate.
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "host.h"
#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kFrac = 7;        // fractional bits of the 16-bit intermediate
constexpr int kOffset = 64;     // added to the intermediate so Lanczos undershoot stays >= 0
constexpr int kPrefetchRows = 64;   // a whole group's worth of new rows must be in flight while the previous group computes
constexpr int kMaxCh = 6;          // most 32-row / 32-pixel chunks per band (the kernel's NCH1 / NCH2): ratios up to ~11x
constexpr double kLookGroups = 2.0;  // ring depth wanted beyond the current band, in groups of 8 output rows
constexpr size_t kSmemBudget = 227 * 1024 - 2048;
const int kPerm[4] = {0, 2, 1, 3};   // k-slot -> pixel inside a 4-pixel block (k2_vfirst.cu, pass-1 epilogue)

}  // namespace

// quantise one output's taps so that they sum to round(sum(w) * 2^S) exactly
void quantise(const float *w, int cnt, int S, std::vector<int64_t> &q) {
    q.assign(cnt, 0);
    double sum_w = 0.0;
    int64_t sum_q = 0;
    int big = 0;
    const double scale = std::ldexp(1.0, S);
    for (int i = 0; i < cnt; i++) {
        sum_w += double(w[i]);
        q[i] = std::llround(double(w[i]) * scale);
        sum_q += q[i];
        if (std::fabs(w[i]) > std::fabs(w[big])) big = i;
    }
    if (cnt) q[big] += std::llround(sum_w * scale) - sum_q;
}

bool split_digits(int64_t x, int ndig, int8_t d[3]) {
    d[0] = d[1] = d[2] = 0;
    for (int l = 0; l < ndig; l++) {
        const int64_t dl = ((x + 128) & 255) - 128;
        d[l] = int8_t(dl);
        x = (x - dl) / 256;
    }
    return x == 0;
}

int pick_scale(float maxw, int ndig) {
    // largest S with maxw * 2^S inside the digit range (with a little room for the sum fix-up)
    const double range = ndig == 2 ? 32000.0 : 8.0e6;
    int S = int(std::floor(std::log2(range / double(maxw))));
    return std::min(S, 22);
}

// B fragment of mma.m16n8k32 (col-major K x N): lane = n*4 + (k%16)/4, register = k/16, byte = k%4
void put_b(uint32_t *frag, int k, int n, int8_t v) {
    const int lane = n * 4 + (k % 16) / 4, reg = k / 16, byte = k % 4;
    frag[lane * 2 + reg] |= uint32_t(uint8_t(v)) << (8 * byte);
}

void build_vplan(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, int ndig, VPlanHost &out) {
    out = VPlanHost();
    out.ndig = ndig;
    if (!g.has_resize) return;
    if (g.vw < g.Rw || g.vh < g.Rh) return;          // downscale (ratio >= 1) on both axes
    if (g.sw % 4 != 0 || g.sw < 4) return;           // 16-byte aligned rows
    if (g.pw == 0 || g.ph == 0 || g.rw == 0 || g.rh == 0) return;

    // ---------------- vertical pass ----------------
    float maxw_y = 0.0f;
    for (uint32_t i = 0; i < g.ph; i++) {
        const uint32_t o = g.py + i;
        for (int k = 0; k < ty.count[o]; k++) maxw_y = std::max(maxw_y, std::fabs(ty.w[size_t(o) * ty.max_taps + k]));
    }
    if (!(maxw_y > 0.0f)) return;
    const int Sy = pick_scale(maxw_y, ndig);
    if (Sy < kFrac + 2) return;
    out.sh1 = Sy - kFrac;

    out.n_groups = (g.ph + 7) / 8;
    out.grp_a0.assign(out.n_groups, 0);
    out.grp_need.assign(out.n_groups, 0);
    out.grp_nch.assign(out.n_groups, 0);
    out.row_computed.assign(g.ph, 0);
    out.bias.assign(size_t(out.n_groups) * 8 * 4, 0);
    int max_need = 0;
    for (uint32_t j = 0; j < out.n_groups; j++) {
        int lo = INT32_MAX, hi = -1;
        for (uint32_t n = 0; n < 8 && j * 8 + n < g.ph; n++) {
            const uint32_t o = g.py + j * 8 + n;
            if (ty.count[o] == 0) continue;
            lo = std::min(lo, ty.left[o]);
            hi = std::max(hi, ty.left[o] + ty.count[o]);
        }
        if (hi < 0) continue;
        out.grp_a0[j] = lo;
        out.grp_need[j] = hi;
        out.grp_nch[j] = (hi - lo + 31) / 32;
        out.nch1 = std::max(out.nch1, out.grp_nch[j]);
        max_need = std::max(max_need, hi - lo);
    }
    if (out.nch1 == 0 || out.nch1 > kMaxCh) return;       // bands of more than 128 source rows: other kernels
    // bands must move down monotonically (ring buffer)
    {
        int prev = -1, prev_need = 0;
        for (uint32_t j = 0; j < out.n_groups; j++) {
            if (out.grp_nch[j] == 0) continue;
            if (out.grp_a0[j] < prev) return;
            if (prev >= 0 && out.grp_a0[j] > prev_need) return;   // no unread rows between bands: every ring step gets waited on
            prev = out.grp_a0[j];
            prev_need = out.grp_need[j];
        }
    }
    out.b1.assign(size_t(out.n_groups) * out.nch1 * ndig * 64, 0u);
    std::vector<int64_t> q;
    for (uint32_t j = 0; j < out.n_groups; j++) {
        for (uint32_t n = 0; n < 8; n++) {
            const uint32_t i = j * 8 + n;
            if (i >= g.ph) break;
            const uint32_t o = g.py + i;
            const int cnt = ty.count[o];
            const float *w = &ty.w[size_t(o) * ty.max_taps];
            quantise(w, cnt, Sy, q);
            int64_t pos = 0, neg = 0;
            for (int k = 0; k < cnt; k++) {
                (q[k] > 0 ? pos : neg) += q[k];
                int8_t d[3];
                if (!split_digits(q[k], ndig, d)) return;
                const int row = ty.left[o] + k - out.grp_a0[j];     // row inside the group's band
                for (int l = 0; l < ndig; l++)
                    put_b(&out.b1[((size_t(j) * out.nch1 + row / 32) * ndig + l) * 64], row % 32, int(n), d[l]);
            }
            out.row_computed[i] = cnt > 0;
            // bias = fill term + offset + rounding half; the 16-bit intermediate must stay in range
            for (int c = 0; c < 4; c++) {
                const int64_t fillterm = std::llround(double(g.fill_pre[c]) * double(ty.fillw[o]) * std::ldexp(1.0, Sy));
                const int64_t bias = fillterm + (int64_t(kOffset) << Sy) + (int64_t(1) << (out.sh1 - 1));
                const int64_t lo_x = 255 * neg + bias, hi_x = 255 * pos + bias;
                if (lo_x < 0 || (hi_x >> out.sh1) > 65535 || hi_x > INT32_MAX || bias > INT32_MAX) return;
                out.bias[size_t(i) * 4 + c] = int32_t(bias);
            }
        }
    }

    // one contiguous record per non-empty group, in band order: what the kernel prefetches into shared
    // memory one group ahead.  Header (64 words): group index, band start, end of the needed rows,
    // chunk count, band start of the next record; bias[8][4] from word 8.  Then the weight fragments.
    out.rec_words = 64 + uint32_t(out.nch1) * ndig * 64;
    for (uint32_t j = 0; j < out.n_groups; j++) {
        if (out.grp_nch[j] == 0) continue;
        const size_t base = out.recs.size();
        out.recs.resize(base + out.rec_words, 0u);
        uint32_t *r = &out.recs[base];
        r[0] = j; r[1] = uint32_t(out.grp_a0[j]); r[2] = uint32_t(out.grp_need[j]); r[3] = uint32_t(out.grp_nch[j]);
        r[4] = uint32_t(out.grp_need[j]);                       // patched below when a next record exists
        for (int k = 0; k < 32; k++) r[8 + k] = uint32_t(out.bias[size_t(j) * 32 + k]);
        std::copy(out.b1.begin() + size_t(j) * out.nch1 * ndig * 64, out.b1.begin() + size_t(j + 1) * out.nch1 * ndig * 64, r + 64);
        if (base) out.recs[base - out.rec_words + 4] = uint32_t(out.grp_a0[j]);
        out.n_recs++;
    }
    if (out.n_recs == 0) return;

    // ---------------- horizontal pass ----------------
    float maxw_x = 0.0f;
    for (uint32_t j = 0; j < g.pw; j++) {
        const uint32_t o = g.px + j;
        for (int k = 0; k < tx.count[o]; k++) maxw_x = std::max(maxw_x, std::fabs(tx.w[size_t(o) * tx.max_taps + k]));
    }
    if (!(maxw_x > 0.0f)) return;
    const int Sx = pick_scale(maxw_x, ndig);
    if (Sx < 10) return;
    // two-digit plans finish pass 2 in integers: t = (acc[lo][d0] >> 8) + acc[lo][d1] + acc[hi][d0] + (acc[hi][d1] << 8) + cxi is
    // the output in units of 2^-sh2; the byte is t >> sh2 (the rounding half is folded into cxi), saturated
    out.sh2 = Sx + kFrac - 8;
    for (int l = 0; l < 2; l++)
        for (int d = 0; d < 3; d++) out.s2[l][d] = d < ndig ? float(std::ldexp(1.0, 8 * l + 8 * d - Sx - kFrac)) : 0.0f;

    // ring depth, in 16-row steps: the rows a group needs plus a group's worth in flight.  When the ring can also
    // hold everything the NEXT group needs before the current one is released (band start rounded down to a step),
    // the producer signals groups one ahead (early_go) and pass 1 never waits on a freshly issued copy.
    int max_look = 0;
    {
        int prev = -1;
        for (uint32_t j = 0; j < out.n_groups; j++) {
            if (out.grp_nch[j] == 0) continue;
            if (prev >= 0) max_look = std::max(max_look, out.grp_need[j] - out.grp_a0[prev]);
            prev = int(j);
        }
    }
    const uint32_t ring_min = uint32_t((std::max(32 * out.nch1, max_need + kPrefetchRows) + 15) / 16 * 16);
    const uint32_t ring_early = std::max(ring_min, uint32_t((max_look + 15 + 15) / 16 * 16));
    const size_t fixed_bytes = 4 * size_t(out.rec_words) * 4 /* group records, k2_vfirst.cu kRecBufs */ + 1024 /* barriers, bookkeeping */;

    // span of source pixels (absolute columns) a range of kept output columns needs
    auto span_of = [&](uint32_t c0, uint32_t c1, int64_t &lo, int64_t &hi) {
        lo = INT64_MAX; hi = -1;
        for (uint32_t j = c0; j < c1; j++) {
            const uint32_t o = g.px + j;
            if (tx.count[o] == 0) continue;
            lo = std::min<int64_t>(lo, int64_t(g.rx) + tx.left[o]);
            hi = std::max<int64_t>(hi, int64_t(g.rx) + tx.left[o] + tx.count[o]);
        }
    };
    // pass 1 splits a strip's 16-byte blocks statically: units of `unit` blocks, upw consecutive units per warp
    // over the (kVfWarps - 2 - tiles) pass-1 warps (the last warps may own fewer, or none)
    auto pad_blocks = [&](int64_t raw, uint32_t ncols, int unit, int &upw) {
        // k2_vfirst.cu: all warps - producer - signaller - tiles, split into cohorts that take turns on the groups
        const int64_t n_p1 = std::max<int64_t>(1, (kVfWarps - 2 - int64_t((ncols + 7) / 8)) / int64_t(out.cohorts));
        const int64_t nunits = (raw + unit - 1) / unit;
        upw = int((nunits + n_p1 - 1) / n_p1);
        return nunits * unit;
    };
    auto smem_for = [&](int64_t nblk, uint32_t ring_rows, uint32_t &pitch, uint32_t &pitch_v) {
        int64_t units = nblk | 1;                                  // odd multiple of 16 bytes: conflict-free ldmatrix rows
        pitch = uint32_t(units * 16);
        int64_t qv = (nblk + 3 + 3) / 4;                           // words per V row = 4 * odd >= nblk + 3
        if (qv % 2 == 0) qv++;
        pitch_v = uint32_t(qv * 16);
        return size_t(ring_rows) * pitch + size_t(2 * out.vbufs) * 32 * pitch_v + fixed_bytes;
    };
    // widest strips (multiple of 8 columns, at most 64) whose staging fits in shared memory; per width the unit
    // size with the least padding; early_go when the deeper ring still fits
    const char *cols_env = std::getenv("SLIMG_B200_STRIP_COLS");
    const char *early_env = std::getenv("SLIMG_B200_EARLY_GO");
    const char *unit_env = std::getenv("SLIMG_B200_UNIT");
    const char *vbufs_env = std::getenv("SLIMG_B200_VBUFS");
    // Bands of exactly 3 chunks (ratios of about 4.5x to 6.5x): three V buffers keep pass 1 fed and the shared memory
    // of the fourth buys a wider strip or a deeper ring (measured on B200: 4000x3000 -> 800, 40-column strips with 3 V
    // buffers 11.9 us per image against 13.7 us for 32 columns with 4; 6000 -> 1000 20.9 against 24.9 us; 6000 -> 1090
    // 24.5 against 28.9).  Longer bands (the 7.5x of the headline config) and shorter ones (4x and below, where the
    // same rule measured 4-11 % slower) keep 4 V buffers and two groups of ring look-ahead.
    const bool short_bands = out.nch1 == 3 && !std::getenv("SLIMG_B200_NO_SHORT_BAND_POLICY");
    out.vbufs = vbufs_env ? uint32_t(std::min(kVfMaxVBufs, std::max(2, std::atoi(vbufs_env)))) : (short_bands ? 3u : 4u);
    // Pass-1 cohorts (k2_vfirst.cu): two groups in flight, each on half the warps.  Measured on B200 (24 MP source unless
    // noted; us per image with 1 / 2 / 3 cohorts): 7.5x chain 18.76 / 18.31 / 17.98, 6x 20.33 / 18.72 / 20.24, 3x 39.29 /
    // 36.02 / 35.48, 12 MP 5x 12.16 / 12.00 / 12.32, 12 MP 10x 16.43 / 16.13 / 16.70; but 8K 4x (7 pass-1 warps: one
    // idles) 36.66 / 37.23 / 39.20, and 2x (one-chunk bands, 6 warps) 70.10 / 69.76 / 64.98.  Rule: two cohorts when the
    // pass-1 warp count is even (odd counts measured within noise either way: one cohort), three where one-chunk bands
    // meet a multiple of three.
    const char *cohorts_env = std::getenv("SLIMG_B200_COHORTS");
    auto cohorts_for = [&](uint32_t width) {
        const int n_p1 = kVfWarps - 2 - int((width + 7) / 8);
        if (cohorts_env) return uint32_t(std::min(std::max(1, n_p1), std::min(3, std::max(1, std::atoi(cohorts_env)))));
        if (n_p1 < 4) return 1u;
        if (out.nch1 == 1 && n_p1 % 3 == 0) return 3u;
        if (n_p1 % 2 == 1) return 1u;                     // an odd count would leave a warp without a role
        return 2u;
    };
    const double look_min = short_bands ? 1.5 : kLookGroups;
    const int min_p1 = short_bands ? 7 : 1;              // pass-1 warps a strip must leave (tiles take one warp each)
    uint32_t cols = 0;
    int unit = 0;
    bool early = false;
    // most chunks any 8-column tile needs, whatever strip it ends up in (its k range starts on a 4-pixel block)
    int nch2_pre = 1;
    for (uint32_t c0 = 0; c0 < g.pw; c0 += 8) {
        int64_t tlo, thi;
        span_of(c0, std::min(g.pw, c0 + 8), tlo, thi);
        if (thi >= 0) nch2_pre = std::max<int>(nch2_pre, int((thi - (tlo & ~int64_t(3)) + 31) / 32));
    }
    if (nch2_pre > kMaxCh) return;
    const int64_t min_blocks = 8 * std::max(4, nch2_pre);      // every tile runs nch2 chunks of 8 blocks inside the staged span
    auto max_padded = [&](uint32_t cand, int u, int &crit) {
        int64_t worst = 0;
        crit = 0;
        for (uint32_t c0 = 0; c0 < g.pw; c0 += cand) {
            const uint32_t c1 = std::min(g.pw, c0 + cand);
            int64_t lo, hi;
            span_of(c0, c1, lo, hi);
            if (hi < 0) continue;
            int upw;
            worst = std::max(worst, pad_blocks((hi - (lo & ~int64_t(3)) + 3) / 4, c1 - c0, u, upw));
            crit = std::max(crit, upw * u);                       // blocks of the busiest warp
        }
        return std::max<int64_t>(worst, min_blocks);
    };
    // Widest strips whose ring still runs `kLookGroups` groups ahead of the band being filtered: the ring is refilled
    // once per group, so a shallow ring leaves too few bytes in flight however wide the strip is (measured on B200:
    // 7.5x downscale 16 columns beat 24, 5x 32 beat 40, 4x the widest won).  Falls back to the widest that fits.
    const double rows_per_group = 8.0 * double(g.vh) / double(g.Rh);
    const char *ring_env = std::getenv("SLIMG_B200_RING_STEPS");
    const uint32_t ring_cap = std::min(512u, ring_env ? uint32_t(std::atoi(ring_env)) * 16 : 512u);
    uint32_t fb_cols = 0, fb_ring = 0; int fb_unit = 0; bool fb_early = false;      // fallback: widest fitting candidate
    uint32_t ring_rows = 0, vb_sel = out.vbufs, fb_vb = out.vbufs;
    out.vbufs_pref = out.vbufs;
    for (uint32_t cand = 64; cand >= 8 && cols == 0; cand -= 8) {
        if (cols_env && uint32_t(std::atoi(cols_env)) != cand) continue;
        // even the strips out first (keeps the multiple of 8)
        const uint32_t ns = (g.pw + cand - 1) / cand;
        const uint32_t even = std::min<uint32_t>(cand, ((g.pw + ns - 1) / ns + 7) / 8 * 8);
        out.cohorts = cohorts_for(even);
        // 16 columns is the last width with two pass-2 tiles and an even pass-1 warp count: below it a CTA filters 8 columns
        // per 45-tap halo.  Before giving it up, give V buffers to the ring (measured on B200, 24 MP: 7.48x 37.4 us with
        // 8 columns / 4 buffers, 22.1 with 16 / 2; 8x 35.1 -> 20.8)
        const uint32_t vb_pref = out.vbufs_pref;
        for (uint32_t vb = vb_pref; vb >= (even == 16 && !vbufs_env ? 2u : vb_pref) && cols == 0; vb--) {
        out.vbufs = vb;
        int best_u = 0, best_crit = INT32_MAX;
        int64_t best_blk = INT64_MAX;
        for (int u = 5; u >= 3; u--) {
            if (unit_env && std::atoi(unit_env) != u) continue;
            if ((out.nch1 > 4 || nch2_pre > 4) && u != 4) continue;      // k2_vfirst.cu instantiates the long-band kernels for UNIT = 4
            // Measured on B200 with two cohorts (us per image, units of 3 / 4 / 5 blocks): 4-chunk bands (7.5x, 24 MP) - /
            // 17.53 / 16.91 -- the fatter unit wins although it leaves a warp of each cohort idle; 3-chunk bands 10.94 /
            // 10.93 / 11.54 (12 MP 5x), 18.67 / 18.85 / 26.66 (6x), 20.59 / 20.07 / 21.69 (5.5x); 2-chunk bands 30.53 / 31.10 /
            // 34.53 (8K 4x), 32.45 / 32.57 / 32.78 (3x).  So: 5-block units for 4-chunk bands only.
            if (!unit_env && out.nch1 <= 4 && nch2_pre <= 4 && (u == 5) != (out.nch1 == 4)) continue;
            int crit;
            const int64_t blk = max_padded(even, u, crit);
            if (crit < best_crit || (crit == best_crit && blk < best_blk)) { best_blk = blk; best_u = u; best_crit = crit; }
        }
        if (!best_u) continue;
        uint32_t pch, pv;
        const bool want_early = !(early_env && early_env[0] == '0');
        bool fits = false, e = false;
        if (want_early && smem_for(best_blk, ring_early, pch, pv) <= kSmemBudget) { fits = true; e = true; }
        else if (smem_for(best_blk, ring_min, pch, pv) <= kSmemBudget) { fits = true; e = false; }
        if (!fits) continue;
        // deepen the ring as far as shared memory allows: every extra step is 16 more rows in flight
        uint32_t rr = e ? ring_early : ring_min;
        while (rr + 16 <= ring_cap && smem_for(best_blk, rr + 16, pch, pv) <= kSmemBudget) rr += 16;
        if (!fb_cols) { fb_cols = even; fb_ring = rr; fb_unit = best_u; fb_early = e; fb_vb = vb; }
        const double look = (double(rr) - 32.0 * out.nch1) / rows_per_group;
        const int n_p1 = kVfWarps - 2 - int((even + 7) / 8);
        // an odd number of tiles leaves an odd number of pass-1 warps, which cannot pair up into two cohorts: the next
        // narrower width can (measured on B200 after the integer epilogue: 8K 4x 56 columns 33.8 us, 48 columns 30.8;
        // 12 MP 5x 40 columns 11.63, 32 columns 11.57).  Odd widths stay as the fallback when nothing else fits.
        if (n_p1 % 2 == 1 && even > 8 && !cols_env && !std::getenv("SLIMG_B200_ODD_STRIPS")) continue;
        if ((look >= look_min && n_p1 >= min_p1) || cols_env) {
            cols = even; unit = best_u; early = e; ring_rows = rr; vb_sel = vb;
            // Two cohorts need two V buffers; a third only helps when the ring already runs well ahead.  Where it does not
            // (look-ahead under 2.5 groups), its shared memory goes to the ring (measured on B200, 24 MP: 6x 18.99 -> 17.78 us,
            // 5.5x 21.02 -> 19.83, look-ahead 2.0 -> 2.7 groups; 12 MP 5x, whose ring runs 3.2 groups ahead, 10.50 -> 11.09
            // the other way and keeps three)
            if (short_bands && vb == 3 && out.cohorts == 2 && look < 2.5 && !vbufs_env) {
                out.vbufs = 2;
                uint32_t rr2 = rr;
                while (rr2 + 16 <= ring_cap && smem_for(best_blk, rr2 + 16, pch, pv) <= kSmemBudget) rr2 += 16;
                if (rr2 > rr) { ring_rows = rr2; vb_sel = 2; }
                out.vbufs = vb;
            }
        }
        }
        out.vbufs = vb_pref;
    }
    if (cols == 0) { cols = fb_cols; unit = fb_unit; early = fb_early; ring_rows = fb_ring; vb_sel = fb_vb; }
    if (cols == 0) return;
    out.vbufs = vb_sel;
    out.cohorts = cohorts_for(cols);
    out.unit = unit;
    out.early_go = early;
    out.ring_rows = ring_rows;
    if (out.ring_rows / 16 > 32) return;                          // k2_vfirst.cu kMaxSteps
    out.n_strips = (g.pw + cols - 1) / cols;

    out.strip_col0.assign(out.n_strips + 1, 0);
    out.strip_src0.assign(out.n_strips, 0);
    out.strip_nblk.assign(out.n_strips, 1);
    out.strip_bytes.assign(out.n_strips, 16);
    out.strip_upw.assign(out.n_strips, 0);
    out.strip_nunits.assign(out.n_strips, 0);
    out.tile_kb0.assign(size_t(out.n_strips) * 8, 0);
    out.tile_nch.assign(size_t(out.n_strips) * 8, 0);
    out.cx.assign(size_t(g.pw) * 4, 0.0f);
    out.cxi.assign(size_t(g.pw) * 4, 0);
    int64_t max_nblk = 1;
    // first sweep: geometry and the chunk count
    for (uint32_t s = 0; s < out.n_strips; s++) {
        const uint32_t c0 = s * cols, c1 = std::min(g.pw, c0 + cols);
        out.strip_col0[s] = int32_t(c0);
        out.strip_col0[s + 1] = int32_t(c1);
        int64_t lo, hi;
        span_of(c0, c1, lo, hi);
        if (hi < 0) continue;                        // every tap of the strip lands on fill
        const int64_t s0 = lo & ~int64_t(3);
        // the padding blocks read whatever lies beyond the staged bytes of the row and meet zero weights in pass 2
        int upw = 0;
        const int64_t nblk = pad_blocks((hi - s0 + 3) / 4, c1 - c0, unit, upw);
        out.strip_src0[s] = int32_t(s0);
        out.strip_nblk[s] = int32_t(nblk);
        out.strip_upw[s] = upw;
        out.strip_nunits[s] = int32_t(nblk / unit);
        out.strip_bytes[s] = int32_t((std::min<int64_t>(s0 + (hi - s0 + 3) / 4 * 4, g.sw) - s0) * 4);
        max_nblk = std::max(max_nblk, nblk);
        for (uint32_t t = 0; t * 8 < c1 - c0; t++) {
            int64_t tlo, thi;
            span_of(c0 + t * 8, std::min(c1, c0 + t * 8 + 8), tlo, thi);
            if (thi < 0) continue;
            int64_t kb0 = (tlo - s0) / 4;
            int64_t nch = (thi - s0 - kb0 * 4 + 31) / 32;
            if (kb0 + 8 * nch > nblk && nblk >= 8 * nch) {      // keep the chunks inside the staged span
                kb0 = nblk - 8 * nch;
            }
            out.tile_kb0[s * 8 + t] = int32_t(kb0);
            out.tile_nch[s * 8 + t] = int32_t(nch);
            out.nch2 = std::max(out.nch2, int(nch));
        }
    }
    if (out.nch2 == 0 || out.nch2 > kMaxCh) return;
    // every tile runs the plan's nch2 chunks: pull the k ranges inside the staged span where it is wide enough
    for (uint32_t s = 0; s < out.n_strips; s++)
        for (int t = 0; t < 8; t++) {
            int32_t &kb0 = out.tile_kb0[s * 8 + t];
            kb0 = std::max<int32_t>(0, std::min<int32_t>(kb0, out.strip_nblk[s] - 8 * out.nch2));
        }
    out.smem_bytes = uint32_t(smem_for(std::max<int64_t>(max_nblk, 8 * out.nch2), out.ring_rows, out.pitch, out.pitch_v));
    if (out.smem_bytes > kSmemBudget) return;
    // second sweep: weights
    out.b2.assign(size_t(out.n_strips) * 8 * out.nch2 * ndig * 64, 0u);
    for (uint32_t s = 0; s < out.n_strips; s++) {
        const uint32_t c0 = uint32_t(out.strip_col0[s]), c1 = uint32_t(out.strip_col0[s + 1]);
        const int64_t s0 = out.strip_src0[s];
        for (uint32_t j = c0; j < c1; j++) {
            const uint32_t o = g.px + j, t = (j - c0) / 8, n = (j - c0) % 8;
            const int cnt = tx.count[o];
            quantise(&tx.w[size_t(o) * tx.max_taps], cnt, Sx, q);
            int64_t sum_q = 0;
            const int64_t kb0 = out.tile_kb0[s * 8 + t];
            for (int k = 0; k < cnt; k++) {
                sum_q += q[k];
                int8_t d[3];
                if (!split_digits(q[k], ndig, d)) return;
                const int64_t rel = int64_t(g.rx) + tx.left[o] + k - s0;      // pixel inside the staged span
                const int64_t blk = rel / 4 - kb0;
                int slot = 0;
                for (int z = 0; z < 4; z++) if (kPerm[z] == int(rel % 4)) slot = z;
                const int64_t kk = blk * 4 + slot;                            // k index inside the tile's chunks
                if (kk < 0 || kk >= 32 * out.nch2) return;
                for (int l = 0; l < ndig; l++)
                    put_b(&out.b2[(((size_t(s) * 8 + t) * out.nch2 + size_t(kk / 32)) * ndig + l) * 64], int(kk % 32), int(n), d[l]);
            }
            // constant term: fill taps of the horizontal pass minus the intermediate's offset
            const double off = -double(kOffset) * double(sum_q) * std::ldexp(1.0, -Sx);
            for (int c = 0; c < 4; c++) {
                const double cxv = off + double(g.fill_pre[c]) * double(tx.fillw[o]);
                out.cx[size_t(j) * 4 + c] = float(cxv);
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
