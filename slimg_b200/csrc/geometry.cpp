// geometry.cpp -- region arithmetic of slimg-core's crop / extend / resize, and the composition
// of a PipelineOptions into one S -> V -> R -> F remap description (host.h).
//
// Reference: crates/slimg-core/src/resize.rs:23-90, crop.rs:19-67, extend.rs:37-91,
// pipeline.rs:66-89; image 0.25.9 src/math/utils.rs `resize_dimensions` and
// src/dynimage.rs `resize` (third-party, restated -- SURVEY.md Appendix A).
#include <cmath>
#include <cstdio>
#include <cstring>

#include "host.h"

namespace slimg {

static thread_local std::string g_last_error;

void set_error(const std::string &msg) { g_last_error = msg; }
slimg_status fail(slimg_status st, const std::string &msg) { g_last_error = msg; return st; }
const std::string &last_error_string() { return g_last_error; }

// Rust `f64::round() as u32`: half away from zero, saturating, NaN -> 0.
static uint32_t round_u32(double x) {
    double r = std::round(x);
    if (std::isnan(r) || r <= 0.0) return 0u;
    if (r >= 4294967295.0) return 4294967295u;
    return static_cast<uint32_t>(r);
}
static uint64_t round_u64(double x) {
    double r = std::round(x);
    if (std::isnan(r) || r <= 0.0) return 0ull;
    if (r >= 18446744073709551615.0) return UINT64_MAX;
    return static_cast<uint64_t>(r);
}

static std::string fmt(const char *f, unsigned long long a = 0, unsigned long long b = 0,
                       unsigned long long c = 0, unsigned long long d = 0,
                       unsigned long long e = 0, unsigned long long g = 0) {
    char buf[256];
    std::snprintf(buf, sizeof buf, f, a, b, c, d, e, g);
    return buf;
}

// resize.rs:23-61
slimg_status calculate_dimensions(uint32_t ow, uint32_t oh, const slimg_resize_mode &m, uint32_t &w, uint32_t &h) {
    uint32_t tw = 0, th = 0;
    switch (m.tag) {
    case SLIMG_RESIZE_WIDTH: {
        double ratio = double(m.a) / double(ow);
        tw = m.a; th = round_u32(double(oh) * ratio);
        break;
    }
    case SLIMG_RESIZE_HEIGHT: {
        double ratio = double(m.a) / double(oh);
        tw = round_u32(double(ow) * ratio); th = m.a;
        break;
    }
    case SLIMG_RESIZE_EXACT: tw = m.a; th = m.b; break;
    case SLIMG_RESIZE_FIT: {
        double ratio = std::fmin(double(m.a) / double(ow), double(m.b) / double(oh));
        tw = round_u32(double(ow) * ratio); th = round_u32(double(oh) * ratio);
        break;
    }
    case SLIMG_RESIZE_SCALE:
        if (!(m.factor > 0.0)) return fail(SLIMG_ERR_RESIZE, "scale factor must be positive");
        tw = round_u32(double(ow) * m.factor); th = round_u32(double(oh) * m.factor);
        break;
    default:
        return fail(SLIMG_ERR_RESIZE, "unknown resize mode");
    }
    if (tw == 0 || th == 0) return fail(SLIMG_ERR_RESIZE, fmt("calculated dimensions are zero: %llux%llu", tw, th));
    w = tw; h = th;
    return SLIMG_OK;
}

// image 0.25.9 resize_dimensions(w, h, nw, nh, fill = false)
static void image_resize_dimensions(uint32_t w, uint32_t h, uint32_t nw, uint32_t nh, uint32_t &ow, uint32_t &oh) {
    double ratio = std::fmin(double(nw) / double(w), double(nh) / double(h));
    uint64_t a = round_u64(double(w) * ratio), b = round_u64(double(h) * ratio);
    if (a < 1) a = 1;
    if (b < 1) b = 1;
    if (a > 4294967295ull) {
        double r = 4294967295.0 / double(w);
        uint32_t hh = round_u32(double(h) * r);
        ow = 4294967295u; oh = hh < 1 ? 1 : hh;
    } else if (b > 4294967295ull) {
        double r = 4294967295.0 / double(h);
        uint32_t ww = round_u32(double(w) * r);
        ow = ww < 1 ? 1 : ww; oh = 4294967295u;
    } else {
        ow = uint32_t(a); oh = uint32_t(b);
    }
}

// resize.rs:64-90 -> DynamicImage::resize_exact (Exact) / DynamicImage::resize (everything else)
slimg_status resize_output_dims(uint32_t w, uint32_t h, const slimg_resize_mode &m, uint32_t &ow, uint32_t &oh) {
    uint32_t tw, th;
    slimg_status st = calculate_dimensions(w, h, m, tw, th);
    if (st != SLIMG_OK) return st;
    if (m.tag == SLIMG_RESIZE_EXACT || (tw == w && th == h)) { ow = tw; oh = th; return SLIMG_OK; }
    image_resize_dimensions(w, h, tw, th, ow, oh);
    return SLIMG_OK;
}

// crop.rs:19-67 (x + width evaluated in u64: the reference's u32 add can wrap and then panic)
slimg_status calculate_crop_region(uint32_t iw, uint32_t ih, const slimg_crop_mode &m,
                                   uint32_t &x, uint32_t &y, uint32_t &w, uint32_t &h) {
    if (m.tag == SLIMG_CROP_REGION) {
        if (m.width == 0 || m.height == 0) return fail(SLIMG_ERR_CROP, "crop dimensions must be non-zero");
        if (uint64_t(m.x) + m.width > iw || uint64_t(m.y) + m.height > ih)
            return fail(SLIMG_ERR_CROP, fmt("crop region (%llu,%llu,%llu,%llu) exceeds image bounds (%llux%llu)",
                                            m.x, m.y, m.width, m.height, iw, ih));
        x = m.x; y = m.y; w = m.width; h = m.height;
        return SLIMG_OK;
    }
    if (m.tag == SLIMG_CROP_ASPECT_RATIO) {
        if (m.width == 0 || m.height == 0) return fail(SLIMG_ERR_CROP, "aspect ratio must be non-zero");
        double target = double(m.width) / double(m.height);
        double img = double(iw) / double(ih);
        uint32_t cw, ch;
        if (img > target) { ch = ih; cw = round_u32(double(ch) * target); }
        else              { cw = iw; ch = round_u32(double(cw) / target); }
        if (cw == 0 || ch == 0 || cw > iw || ch > ih)
            return fail(SLIMG_ERR_CROP, fmt("aspect-ratio crop of a %llux%llu image is empty", iw, ih));
        x = (iw - cw) / 2; y = (ih - ch) / 2; w = cw; h = ch;
        return SLIMG_OK;
    }
    return fail(SLIMG_ERR_CROP, "unknown crop mode");
}

// extend.rs:37-91
slimg_status calculate_extend_region(uint32_t iw, uint32_t ih, const slimg_extend_mode &m,
                                     uint32_t &cw, uint32_t &ch, uint32_t &ox, uint32_t &oy) {
    if (m.tag == SLIMG_EXTEND_ASPECT_RATIO) {
        if (m.width == 0 || m.height == 0) return fail(SLIMG_ERR_EXTEND, "aspect ratio must be non-zero");
        double target = double(m.width) / double(m.height);
        double img = double(iw) / double(ih);
        uint32_t w, h;
        if (img < target) { h = ih; w = round_u32(double(h) * target); }
        else              { w = iw; h = round_u32(double(w) / target); }
        if (w < iw || h < ih) return fail(SLIMG_ERR_EXTEND, "extended canvas would be smaller than the image");
        cw = w; ch = h; ox = (w - iw) / 2; oy = (h - ih) / 2;
        return SLIMG_OK;
    }
    if (m.tag == SLIMG_EXTEND_SIZE) {
        if (m.width == 0 || m.height == 0) return fail(SLIMG_ERR_EXTEND, "extend dimensions must be non-zero");
        if (m.width < iw || m.height < ih)
            return fail(SLIMG_ERR_EXTEND, fmt("target size (%llux%llu) is smaller than image (%llux%llu)",
                                              m.width, m.height, iw, ih));
        cw = m.width; ch = m.height; ox = (m.width - iw) / 2; oy = (m.height - ih) / 2;
        return SLIMG_OK;
    }
    return fail(SLIMG_ERR_EXTEND, "unknown extend mode");
}

static void resolve_fill(const slimg_pipeline_options &o, uint8_t out[4]) {
    // pipeline.rs:78-80: fill_color None -> opaque white; extend.rs:14-21: Transparent -> 0,0,0,0
    if (!o.has_fill) { out[0] = out[1] = out[2] = out[3] = 255; return; }
    if (o.fill.tag == SLIMG_FILL_TRANSPARENT) { out[0] = out[1] = out[2] = out[3] = 0; return; }
    out[0] = o.fill.r; out[1] = o.fill.g; out[2] = o.fill.b; out[3] = o.fill.a;
}

slimg_status build_geometry(uint32_t w, uint32_t h, const slimg_pipeline_options &o, int order, PipeGeom &g) {
    g = PipeGeom();
    g.sw = w; g.sh = h;
    slimg_status st;
    if (o.has_fill && o.fill.tag != SLIMG_FILL_SOLID && o.fill.tag != SLIMG_FILL_TRANSPARENT)
        return fail(SLIMG_ERR_EXTEND, "unknown fill colour");
    if (order == SLIMG_ORDER_CONVERT) {
        // pipeline.rs:71-89  crop -> extend -> resize
        g.rx = 0; g.ry = 0; g.rw = w; g.rh = h;
        if (o.has_crop && (st = calculate_crop_region(w, h, o.crop, g.rx, g.ry, g.rw, g.rh)) != SLIMG_OK) return st;
        g.vw = g.rw; g.vh = g.rh; g.vox = g.voy = 0;
        if (o.has_extend) {
            if ((st = calculate_extend_region(g.rw, g.rh, o.extend, g.vw, g.vh, g.vox, g.voy)) != SLIMG_OK) return st;
            resolve_fill(o, g.fill_pre);
        }
        g.Rw = g.vw; g.Rh = g.vh;
        if (o.has_resize) {
            if ((st = resize_output_dims(g.vw, g.vh, o.resize, g.Rw, g.Rh)) != SLIMG_OK) return st;
            g.has_resize = !(g.Rw == g.vw && g.Rh == g.vh);
        }
        g.px = g.py = 0; g.pw = g.Rw; g.ph = g.Rh;
        g.fw = g.Rw; g.fh = g.Rh; g.fox = g.foy = 0;
    } else if (order == SLIMG_ORDER_CHAIN) {
        // resize -> crop -> extend, each stage exactly the standalone op
        g.rx = g.ry = 0; g.rw = w; g.rh = h;
        g.vw = w; g.vh = h; g.vox = g.voy = 0;
        g.Rw = w; g.Rh = h;
        if (o.has_resize) {
            if ((st = resize_output_dims(w, h, o.resize, g.Rw, g.Rh)) != SLIMG_OK) return st;
            g.has_resize = !(g.Rw == w && g.Rh == h);
        }
        g.px = g.py = 0; g.pw = g.Rw; g.ph = g.Rh;
        if (o.has_crop && (st = calculate_crop_region(g.Rw, g.Rh, o.crop, g.px, g.py, g.pw, g.ph)) != SLIMG_OK) return st;
        g.fw = g.pw; g.fh = g.ph; g.fox = g.foy = 0;
        if (o.has_extend) {
            if ((st = calculate_extend_region(g.pw, g.ph, o.extend, g.fw, g.fh, g.fox, g.foy)) != SLIMG_OK) return st;
            resolve_fill(o, g.fill_post);
        }
    } else {
        return fail(SLIMG_ERR_IMAGE, "unknown pipeline order");
    }
    return SLIMG_OK;
}

}  // namespace slimg

extern "C" {

slimg_status slimg_b200_calculate_dimensions(uint32_t w, uint32_t h, const slimg_resize_mode *mode,
                                             uint32_t *out_w, uint32_t *out_h) {
    if (!mode || !out_w || !out_h) return slimg::fail(SLIMG_ERR_IMAGE, "null argument");
    return slimg::calculate_dimensions(w, h, *mode, *out_w, *out_h);
}
slimg_status slimg_b200_resize_output_dims(uint32_t w, uint32_t h, const slimg_resize_mode *mode,
                                           uint32_t *out_w, uint32_t *out_h) {
    if (!mode || !out_w || !out_h) return slimg::fail(SLIMG_ERR_IMAGE, "null argument");
    return slimg::resize_output_dims(w, h, *mode, *out_w, *out_h);
}
slimg_status slimg_b200_calculate_crop_region(uint32_t w, uint32_t h, const slimg_crop_mode *mode,
                                              uint32_t *x, uint32_t *y, uint32_t *cw, uint32_t *ch) {
    if (!mode || !x || !y || !cw || !ch) return slimg::fail(SLIMG_ERR_IMAGE, "null argument");
    return slimg::calculate_crop_region(w, h, *mode, *x, *y, *cw, *ch);
}
slimg_status slimg_b200_calculate_extend_region(uint32_t w, uint32_t h, const slimg_extend_mode *mode,
                                                uint32_t *cw, uint32_t *ch, uint32_t *ox, uint32_t *oy) {
    if (!mode || !cw || !ch || !ox || !oy) return slimg::fail(SLIMG_ERR_IMAGE, "null argument");
    return slimg::calculate_extend_region(w, h, *mode, *cw, *ch, *ox, *oy);
}
slimg_status slimg_b200_pipeline_output_dims(uint32_t w, uint32_t h, const slimg_pipeline_options *opts,
                                             int order, uint32_t *out_w, uint32_t *out_h) {
    if (!opts || !out_w || !out_h) return slimg::fail(SLIMG_ERR_IMAGE, "null argument");
    slimg::PipeGeom g;
    slimg_status st = slimg::build_geometry(w, h, *opts, order, g);
    if (st == SLIMG_OK) { *out_w = g.fw; *out_h = g.fh; }
    return st;
}

}  // extern "C"
