// fastplan.cpp -- host-side plan of the K2 fast path (k2_fast.cu).
//
// Horizontal pass = integer tensor-core GEMM.  The f32 Lanczos weights of taps.cpp are turned
// into fixed point with S fractional bits (S <= 22) and split into three balanced base-128
// digits (s8 each, value = d0 + 128*d1 + 16384*d2).  u8 pixels x s8 digits accumulate exactly
// in s32, so the only error of the pass is the weight rounding (<= 2^-(S+1) per tap; the digits
// of every output are nudged so they sum to exactly 2^S).  The digits are laid out in the
// register order of mma.sync.m16n8k32's B operand so a warp fetches them with one 8-byte load
// per lane.
//
// Vertical pass = streaming "slot" accumulation: output row i lives in slot i mod 8 while the
// source rows of its window go by; the table built here gives, for every source row, the weight
// it contributes to each slot and which slots are complete after it.
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "host.h"

namespace slimg {

static bool split_digits(int64_t x, int8_t d[3]) {
    int64_t d0 = ((x + 64) & 127) - 64;
    int64_t x1 = (x - d0) / 128;
    int64_t d1 = ((x1 + 64) & 127) - 64;
    int64_t d2 = (x1 - d1) / 128;
    if (d2 < -128 || d2 > 127) return false;
    d[0] = int8_t(d0); d[1] = int8_t(d1); d[2] = int8_t(d2);
    return true;
}

void build_fast_plan(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, FastPlanHost &out) {
    out = FastPlanHost();
    if (!g.has_resize) return;
    if (g.vw < g.Rw || g.vh < g.Rh) return;          // fast path: downscale (ratio >= 1) on both axes
    if (g.sw % 4 != 0 || g.sw < 4) return;           // 128-bit row loads
    if (g.pw == 0 || g.ph == 0 || g.rw == 0 || g.rh == 0) return;
    if (g.rh > 32000 || g.ph > 32000) return;        // int16 row ids in the slot table

    // ---- fixed-point scale ------------------------------------------------------------------
    float maxw = 0.0f;
    double max_abs_sum = 0.0;
    for (uint32_t j = 0; j < g.pw; j++) {
        const uint32_t ox = g.px + j;
        double s = 0.0;
        for (int i = 0; i < tx.count[ox]; i++) {
            const float w = tx.w[size_t(ox) * tx.max_taps + i];
            maxw = std::max(maxw, std::fabs(w));
            s += std::fabs(w);
        }
        max_abs_sum = std::max(max_abs_sum, s);
    }
    if (!(maxw > 0.0f)) return;                      // nothing but fill: generic kernel handles it
    int S = int(std::floor(std::log2(2088000.0 / double(maxw))));
    S = std::min(S, 22);
    while (S > 8 && max_abs_sum * std::ldexp(1.0, S) * 255.0 >= 2.0e9) S--;
    if (S <= 8) return;
    out.hscale = float(std::ldexp(1.0, -S));

    // ---- tiles of 8 output columns --------------------------------------------------------------
    out.n_tiles = (g.pw + 7) / 8;
    out.n_strips = (out.n_tiles + 7) / 8;
    out.tile_a0.assign(out.n_tiles, 0);
    out.tile_nchunks.assign(out.n_tiles, 0);
    out.tile_boff.assign(out.n_tiles, 0);
    out.fillw_x.assign(size_t(out.n_tiles) * 8, 0.0f);
    std::vector<int64_t> wq;
    int32_t chunk_total = 0;
    for (uint32_t t = 0; t < out.n_tiles; t++) {
        int64_t lo = INT64_MAX, hi = -1;
        for (uint32_t n = 0; n < 8; n++) {
            const uint32_t j = t * 8 + n;
            if (j >= g.pw) break;
            const uint32_t ox = g.px + j;
            out.fillw_x[j] = tx.fillw[ox];
            if (tx.count[ox] == 0) continue;
            lo = std::min<int64_t>(lo, int64_t(g.rx) + tx.left[ox]);
            hi = std::max<int64_t>(hi, int64_t(g.rx) + tx.left[ox] + tx.count[ox]);
        }
        out.tile_boff[t] = chunk_total;
        if (hi < 0) continue;                         // the whole tile reads fill only
        const int64_t a0 = lo & ~int64_t(3);
        const int nch = int((hi - a0 + 31) / 32);
        if (nch > 256) return;
        out.tile_a0[t] = int32_t(a0);
        out.tile_nchunks[t] = nch;
        chunk_total += nch;
        out.bfrag.resize(size_t(chunk_total) * 3 * 64, 0u);

        for (uint32_t n = 0; n < 8; n++) {
            const uint32_t j = t * 8 + n;
            if (j >= g.pw) break;
            const uint32_t ox = g.px + j;
            const int cnt = tx.count[ox];
            if (cnt == 0) continue;
            // quantise, then make the digits of this output sum to round(sum(w) * 2^S) exactly
            wq.assign(cnt, 0);
            double sum_w = 0.0;
            int64_t sum_q = 0;
            int big = 0;
            for (int i = 0; i < cnt; i++) {
                const float w = tx.w[size_t(ox) * tx.max_taps + i];
                sum_w += double(w);
                wq[i] = std::llround(double(w) * std::ldexp(1.0, S));
                sum_q += wq[i];
                if (std::fabs(w) > std::fabs(tx.w[size_t(ox) * tx.max_taps + big])) big = i;
            }
            wq[big] += std::llround(sum_w * std::ldexp(1.0, S)) - sum_q;
            for (int i = 0; i < cnt; i++) {
                int8_t d[3];
                if (!split_digits(wq[i], d)) return;
                const int64_t k_abs = int64_t(g.rx) + tx.left[ox] + i - a0;   // k index inside the tile's chunks
                const int c = int(k_abs / 32), k = int(k_abs % 32);
                // B fragment of mma.m16n8k32: lane = n*4 + (k%16)/4, register = k/16, byte = k%4
                const int lane = int(n) * 4 + (k % 16) / 4, reg = k / 16, byte = k % 4;
                for (int l = 0; l < 3; l++) {
                    uint32_t &word = out.bfrag[(size_t(out.tile_boff[t] + c) * 3 + l) * 64 + lane * 2 + reg];
                    word |= uint32_t(uint8_t(d[l])) << (8 * byte);
                }
            }
        }
    }
    if (out.bfrag.empty()) out.bfrag.assign(64, 0u);

    // ---- strips: the span of source columns one CTA stages in shared memory -----------------------
    out.strip_a0.assign(out.n_strips, 0);
    out.strip_bytes.assign(out.n_strips, 16);
    int64_t span_max = 4;
    for (uint32_t s = 0; s < out.n_strips; s++) {
        int64_t lo = INT64_MAX, hi = -1;
        for (uint32_t t = s * 8; t < std::min(out.n_tiles, s * 8 + 8); t++) {
            if (out.tile_nchunks[t] == 0) continue;
            lo = std::min<int64_t>(lo, out.tile_a0[t]);
            hi = std::max<int64_t>(hi, int64_t(out.tile_a0[t]) + 32 * out.tile_nchunks[t]);
        }
        if (hi < 0) { lo = 0; hi = 4; }
        span_max = std::max(span_max, hi - lo);
        out.strip_a0[s] = int32_t(lo);
        out.strip_bytes[s] = int32_t((std::min<int64_t>(hi, g.sw) - lo) * 4);   // never read past the row end
        for (uint32_t t = s * 8; t < std::min(out.n_tiles, s * 8 + 8); t++)
            out.tile_a0[t] = out.tile_nchunks[t] ? int32_t(out.tile_a0[t] - lo) : 0;
    }
    out.pitch = uint32_t((span_max * 4 + 127) / 128 * 128 + 64);
    if (size_t(out.pitch) * 16 * 2 > 200 * 1024) return;   // two stages must fit in shared memory

    // ---- vertical slot table ----------------------------------------------------------------------
    const size_t rows_pad = (size_t(g.rh) + 15) / 16 * 16 + 16;
    out.vrow_w.assign(rows_pad * 8, 0.0f);
    out.vrow_out.assign(rows_pad * 8, int16_t(-1));
    out.vleft.assign(g.ph, 0);
    out.vright.assign(g.ph, 0);
    out.fillw_y.assign(g.ph, 0.0f);
    out.row_computed.assign(g.ph, 0);
    std::vector<int32_t> slot_free_at(8, 0);         // first row at which the slot may be reused
    for (uint32_t i = 0; i < g.ph; i++) {
        const uint32_t o = g.py + i;
        out.fillw_y[i] = ty.fillw[o];
        const int cnt = ty.count[o], left = ty.left[o];
        out.vleft[i] = left; out.vright[i] = left + cnt;
        if (cnt == 0) continue;
        out.row_computed[i] = 1;
        const int s = int(i % 8);
        if (left < slot_free_at[s]) return;           // more than 8 windows alive at once: not a downscale
        slot_free_at[s] = left + cnt;
        for (int k = 0; k < cnt; k++) out.vrow_w[size_t(left + k) * 8 + s] = ty.w[size_t(o) * ty.max_taps + k];
        out.vrow_out[size_t(left + cnt - 1) * 8 + s] = int16_t(i);
    }
    // the kernel stages at most kMaxFlush (24) completed rows per 16-row step
    {
        std::vector<int> done(rows_pad + 1, 0);
        for (uint32_t i = 0; i < g.ph; i++)
            if (out.row_computed[i]) done[out.vright[i] - 1]++;
        int window = 0;
        for (size_t y = 0; y < rows_pad; y++) {
            window += done[y];
            if (y >= 16) window -= done[y - 16];
            if (window > 24) return;
        }
    }
    for (uint32_t i = 0; i < g.ph; i++)
        if (!out.row_computed[i]) out.need_fill = true;
    if (g.pw != g.fw || g.ph != g.fh) out.need_fill = true;
    out.ok = true;
}

}  // namespace slimg
