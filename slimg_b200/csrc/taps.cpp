// taps.cpp -- Lanczos3 sampling bounds and weights of one axis, computed on the host in strict
// f32 exactly as image 0.25.9's imageops/sample.rs does inside vertical_sample /
// horizontal_sample (third-party, restated; call site crates/slimg-core/src/resize.rs:79-82):
//   ratio = in/out; sratio = max(ratio,1); support = 3*sratio
//   c = (o+0.5)*ratio; left = clamp(floor(c-support),0,in-1); right = clamp(ceil(c+support),left+1,in)
//   w_i = lanczos3((i-(c-0.5))/sratio), normalised by their running f32 sum.
// Compiled with -ffp-contract=off so no mul+add is fused (rustc never fuses).
#include <cmath>

#include "host.h"

namespace slimg {

static float sinc_f32(float t) {
    float a = t * 3.14159274101257324f;  // f32::consts::PI
    if (t == 0.0f) return 1.0f;
    return sinf(a) / a;
}

static float lanczos3(float x) {
    if (fabsf(x) < 3.0f) return sinc_f32(x) * sinc_f32(x / 3.0f);
    return 0.0f;
}

void build_axis_taps(uint32_t in_len, uint32_t out_len, uint32_t rect_off, uint32_t rect_len, AxisTaps &t) {
    t.out_len = int(out_len);
    t.left.assign(out_len, 0);
    t.count.assign(out_len, 0);
    t.fillw.assign(out_len, 0.0f);

    const float ratio = float(in_len) / float(out_len);
    const float sratio = ratio < 1.0f ? 1.0f : ratio;
    const float support = 3.0f * sratio;
    const int64_t r0 = rect_off, r1 = int64_t(rect_off) + rect_len;

    struct Win { int64_t left, right; float c; };
    std::vector<Win> win(out_len);
    int max_in_rect = 1;
    for (uint32_t o = 0; o < out_len; o++) {
        float c = (float(o) + 0.5f) * ratio;
        int64_t left = int64_t(floorf(c - support));
        if (left < 0) left = 0;
        if (left > int64_t(in_len) - 1) left = int64_t(in_len) - 1;
        int64_t right = int64_t(ceilf(c + support));
        if (right < left + 1) right = left + 1;
        if (right > int64_t(in_len)) right = int64_t(in_len);
        win[o] = {left, right, c - 0.5f};
        int64_t a = left > r0 ? left : r0, b = right < r1 ? right : r1;
        if (b - a > max_in_rect) max_in_rect = int(b - a);
    }
    t.max_taps = max_in_rect;
    t.w.assign(size_t(out_len) * t.max_taps, 0.0f);

    std::vector<float> ws;
    for (uint32_t o = 0; o < out_len; o++) {
        const Win &wn = win[o];
        ws.clear();
        float sum = 0.0f;
        for (int64_t i = wn.left; i < wn.right; i++) {
            float w = lanczos3((float(i) - wn.c) / sratio);
            ws.push_back(w);
            sum += w;
        }
        for (float &w : ws) w /= sum;

        int64_t a = wn.left > r0 ? wn.left : r0, b = wn.right < r1 ? wn.right : r1;
        float fw = 0.0f;
        int n = 0;
        float *dst = &t.w[size_t(o) * t.max_taps];
        for (int64_t i = wn.left; i < wn.right; i++) {
            float w = ws[size_t(i - wn.left)];
            if (i >= a && i < b) dst[n++] = w;
            else fw += w;
        }
        t.left[o] = n > 0 ? int32_t(a - r0) : 0;
        t.count[o] = n;
        t.fillw[o] = fw;
    }
}

}  // namespace slimg
