/*
 * slimg_oracle.c -- CPU oracle for the slimg pixel pipeline.  TEST INFRASTRUCTURE ONLY
 * (see slimg_oracle.h for who may call this and for what pins it).
 *
 * Build: gcc -O3 -march=native -ffp-contract=off -fno-fast-math -fPIC -shared (oracle/Makefile).
 * -ffp-contract=off matters: rustc never fuses `t += p * w`, so neither may we.
 */
#include "slimg_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * Rust numeric semantics
 * ---------------------------------------------------------------------------------------- */

/* `x.round() as u32`: round half away from zero, then a saturating cast (NaN -> 0). */
static uint32_t rust_round_u32(double x)
{
    double r = round(x);
    if (!(r == r)) return 0;
    if (r <= 0.0) return 0;
    if (r >= 4294967295.0) return 4294967295u;
    return (uint32_t)r;
}

static uint64_t rust_round_u64(double x)
{
    double r = round(x);
    if (!(r == r)) return 0;
    if (r <= 0.0) return 0;
    if (r >= 18446744073709551615.0) return UINT64_MAX;
    return (uint64_t)r;
}

/* ------------------------------------------------------------------------------------------
 * Geometry
 * ---------------------------------------------------------------------------------------- */

/* crates/slimg-core/src/resize.rs:23-61 */
int so_calculate_dimensions(uint32_t ow, uint32_t oh, const so_resize_mode *m,
                            uint32_t *w, uint32_t *h)
{
    uint32_t tw, th;
    switch (m->tag) {
    case 1: { /* Width */
        double ratio = (double)m->a / (double)ow;
        tw = m->a;
        th = rust_round_u32((double)oh * ratio);
        break;
    }
    case 2: { /* Height */
        double ratio = (double)m->a / (double)oh;
        tw = rust_round_u32((double)ow * ratio);
        th = m->a;
        break;
    }
    case 3: /* Exact */
        tw = m->a; th = m->b;
        break;
    case 4: { /* Fit */
        double rw = (double)m->a / (double)ow;
        double rh = (double)m->b / (double)oh;
        double ratio = fmin(rw, rh);
        tw = rust_round_u32((double)ow * ratio);
        th = rust_round_u32((double)oh * ratio);
        break;
    }
    case 5: /* Scale */
        if (m->factor <= 0.0) return SO_ERR_RESIZE;
        tw = rust_round_u32((double)ow * m->factor);
        th = rust_round_u32((double)oh * m->factor);
        break;
    default:
        return SO_ERR_RESIZE;
    }
    if (tw == 0 || th == 0) return SO_ERR_RESIZE;
    *w = tw; *h = th;
    return SO_OK;
}

/* image 0.25.9, src/math/utils.rs `resize_dimensions` (third-party, not in tree) */
void so_image_resize_dimensions(uint32_t w, uint32_t h, uint32_t nw, uint32_t nh, int fill,
                                uint32_t *ow, uint32_t *oh)
{
    double wratio = (double)nw / (double)w;
    double hratio = (double)nh / (double)h;
    double ratio = fill ? fmax(wratio, hratio) : fmin(wratio, hratio);
    uint64_t a = rust_round_u64((double)w * ratio);
    uint64_t b = rust_round_u64((double)h * ratio);
    if (a < 1) a = 1;
    if (b < 1) b = 1;
    if (a > 4294967295ull) {
        double r = 4294967295.0 / (double)w;
        uint32_t hh = rust_round_u32((double)h * r);
        *ow = 4294967295u; *oh = hh < 1 ? 1 : hh;
    } else if (b > 4294967295ull) {
        double r = 4294967295.0 / (double)h;
        uint32_t ww = rust_round_u32((double)w * r);
        *ow = ww < 1 ? 1 : ww; *oh = 4294967295u;
    } else {
        *ow = (uint32_t)a; *oh = (uint32_t)b;
    }
}

/* resize.rs:64-90: Exact -> resize_exact(target); every other mode -> DynamicImage::resize,
 * which keeps the image when target == current dims and otherwise RE-DERIVES the dims with
 * resize_dimensions(fill = false) (1000x334 Width(300) -> 299x100). */
int so_resize_output_dims(uint32_t w, uint32_t h, const so_resize_mode *m,
                          uint32_t *ow, uint32_t *oh)
{
    uint32_t tw, th;
    int st = so_calculate_dimensions(w, h, m, &tw, &th);
    if (st != SO_OK) return st;
    if (m->tag == 3) { *ow = tw; *oh = th; return SO_OK; }
    if (tw == w && th == h) { *ow = w; *oh = h; return SO_OK; }
    so_image_resize_dimensions(w, h, tw, th, 0, ow, oh);
    return SO_OK;
}

/* crop.rs:19-67.  The reference adds x + width in u32 (wraps in release builds and then
 * panics on the slice); the oracle widens to u64 and reports the Crop error instead. */
int so_calculate_crop_region(uint32_t iw, uint32_t ih, const so_crop_mode *m,
                             uint32_t *x, uint32_t *y, uint32_t *w, uint32_t *h)
{
    if (m->tag == 1) {
        if (m->w == 0 || m->h == 0) return SO_ERR_CROP;
        if ((uint64_t)m->x + m->w > iw || (uint64_t)m->y + m->h > ih) return SO_ERR_CROP;
        *x = m->x; *y = m->y; *w = m->w; *h = m->h;
        return SO_OK;
    }
    if (m->tag == 2) {
        if (m->w == 0 || m->h == 0) return SO_ERR_CROP;
        double target = (double)m->w / (double)m->h;
        double img = (double)iw / (double)ih;
        uint32_t cw, ch;
        if (img > target) { ch = ih; cw = rust_round_u32((double)ch * target); }
        else              { cw = iw; ch = rust_round_u32((double)cw / target); }
        if (cw > iw || ch > ih || cw == 0 || ch == 0) return SO_ERR_CROP; /* u32 underflow in the reference */
        *x = (iw - cw) / 2; *y = (ih - ch) / 2; *w = cw; *h = ch;
        return SO_OK;
    }
    return SO_ERR_CROP;
}

/* extend.rs:37-91 */
int so_calculate_extend_region(uint32_t iw, uint32_t ih, const so_extend_mode *m,
                               uint32_t *cw, uint32_t *ch, uint32_t *ox, uint32_t *oy)
{
    if (m->tag == 1) {
        if (m->w == 0 || m->h == 0) return SO_ERR_EXTEND;
        double target = (double)m->w / (double)m->h;
        double img = (double)iw / (double)ih;
        uint32_t w, h;
        if (img < target) { h = ih; w = rust_round_u32((double)h * target); }
        else              { w = iw; h = rust_round_u32((double)w / target); }
        if (w < iw || h < ih) return SO_ERR_EXTEND; /* u32 underflow in the reference */
        *cw = w; *ch = h; *ox = (w - iw) / 2; *oy = (h - ih) / 2;
        return SO_OK;
    }
    if (m->tag == 2) {
        if (m->w == 0 || m->h == 0) return SO_ERR_EXTEND;
        if (m->w < iw || m->h < ih) return SO_ERR_EXTEND;
        *cw = m->w; *ch = m->h; *ox = (m->w - iw) / 2; *oy = (m->h - ih) / 2;
        return SO_OK;
    }
    return SO_ERR_EXTEND;
}

/* ------------------------------------------------------------------------------------------
 * crop / extend byte movers
 * ---------------------------------------------------------------------------------------- */

/* crop.rs:70-87 */
void so_crop_pixels(const uint8_t *src, uint32_t sw, uint32_t sh,
                    uint32_t x, uint32_t y, uint32_t cw, uint32_t ch, uint8_t *dst)
{
    (void)sh;
    size_t sstride = (size_t)sw * 4, dstride = (size_t)cw * 4;
    for (size_t row = 0; row < ch; row++)
        memcpy(dst + row * dstride, src + ((size_t)y + row) * sstride + (size_t)x * 4, dstride);
}

/* extend.rs:94-133 */
void so_extend_pixels(const uint8_t *src, uint32_t sw, uint32_t sh,
                      uint32_t cw, uint32_t ch, uint32_t ox, uint32_t oy,
                      const uint8_t fill[4], uint8_t *dst)
{
    size_t cstride = (size_t)cw * 4, sstride = (size_t)sw * 4;
    if (cw == sw && ch == sh) { memcpy(dst, src, sstride * sh); return; }   /* extend.rs:107-109 */
    size_t npx = (size_t)cw * ch;
    for (size_t i = 0; i < npx; i++) memcpy(dst + i * 4, fill, 4);           /* extend.rs:116-122 */
    for (size_t row = 0; row < sh; row++)                                    /* extend.rs:125-130 */
        memcpy(dst + ((size_t)oy + row) * cstride + (size_t)ox * 4, src + row * sstride, sstride);
}

/* ------------------------------------------------------------------------------------------
 * Lanczos3 resize -- image 0.25.9 imageops/sample.rs (third-party, restated; PARITY UNPINNED)
 * ---------------------------------------------------------------------------------------- */

static float so_sinc(float t)
{
    float a = t * 3.14159274101257324f; /* f32::consts::PI */
    if (t == 0.0f) return 1.0f;
    return sinf(a) / a;
}

static float so_lanczos3(float x)
{
    if (fabsf(x) < 3.0f) return so_sinc(x) * so_sinc(x / 3.0f);
    return 0.0f;
}

typedef struct { int32_t *left, *count; float *w; int max_taps; } so_taps;

static int64_t clamp_i64(int64_t a, int64_t lo, int64_t hi) { return a < lo ? lo : (a > hi ? hi : a); }

/* The bounds / weight loop shared by vertical_sample and horizontal_sample. */
static int so_axis_bounds(uint32_t in_len, uint32_t out_len, uint32_t o,
                          float *ratio_o, float *sratio_o, float *centre_o, int64_t *left_o)
{
    float ratio = (float)in_len / (float)out_len;
    float sratio = ratio < 1.0f ? 1.0f : ratio;
    float support = 3.0f * sratio;
    float c = ((float)o + 0.5f) * ratio;
    int64_t left = (int64_t)floorf(c - support);
    left = clamp_i64(left, 0, (int64_t)in_len - 1);
    int64_t right = (int64_t)ceilf(c + support);
    right = clamp_i64(right, left + 1, (int64_t)in_len);
    *ratio_o = ratio; *sratio_o = sratio; *centre_o = c - 0.5f; *left_o = left;
    return (int)(right - left);
}

static int so_max_taps(uint32_t in_len, uint32_t out_len)
{
    int m = 0;
    for (uint32_t o = 0; o < out_len; o++) {
        float r, s, c; int64_t l;
        int n = so_axis_bounds(in_len, out_len, o, &r, &s, &c, &l);
        if (n > m) m = n;
    }
    return m;
}

static void so_fill_taps(uint32_t in_len, uint32_t out_len, int32_t *left, int32_t *count,
                         float *weights, int max_taps)
{
    for (uint32_t o = 0; o < out_len; o++) {
        float ratio, sratio, c; int64_t l;
        int n = so_axis_bounds(in_len, out_len, o, &ratio, &sratio, &c, &l);
        float *ws = weights + (size_t)o * max_taps;
        float sum = 0.0f;
        for (int i = 0; i < n; i++) {
            float w = so_lanczos3(((float)(l + i) - c) / sratio);
            ws[i] = w;
            sum += w;
        }
        for (int i = 0; i < n; i++) ws[i] /= sum;
        for (int i = n; i < max_taps; i++) ws[i] = 0.0f;
        left[o] = (int32_t)l; count[o] = n;
    }
}

int so_lanczos3_taps(uint32_t in_len, uint32_t out_len, int32_t *left, int32_t *count,
                     float *weights, int max_taps)
{
    int m = so_max_taps(in_len, out_len);
    if (weights && max_taps >= m) so_fill_taps(in_len, out_len, left, count, weights, max_taps);
    return m;
}

static int so_make_taps(uint32_t in_len, uint32_t out_len, so_taps *t)
{
    t->max_taps = so_max_taps(in_len, out_len);
    t->left = (int32_t *)malloc(sizeof(int32_t) * out_len);
    t->count = (int32_t *)malloc(sizeof(int32_t) * out_len);
    t->w = (float *)malloc(sizeof(float) * (size_t)out_len * t->max_taps);
    if (!t->left || !t->count || !t->w) return -1;
    so_fill_taps(in_len, out_len, t->left, t->count, t->w, t->max_taps);
    return 0;
}

static void so_free_taps(so_taps *t) { free(t->left); free(t->count); free(t->w); }

/* imageops::resize: vertical_sample (u8 -> f32, not clamped, not rounded), then
 * horizontal_sample (f32 -> clamp [0,255] -> round half away -> u8).  Every output sample is
 * `t = 0.0; for taps in order: t += p * w` in f32; the loops below are re-nested for cache
 * behaviour only, the per-sample operation order is the crate's. */
int so_resize_exact(const uint8_t *src, uint32_t w, uint32_t h,
                    uint32_t nw, uint32_t nh, uint8_t *dst)
{
    if (w == 0 || h == 0) { memset(dst, 0, (size_t)nw * nh * 4); return 0; }
    if (nw == w && nh == h) { memcpy(dst, src, (size_t)w * h * 4); return 0; }

    so_taps tv, th;
    if (so_make_taps(h, nh, &tv)) return -1;
    if (so_make_taps(w, nw, &th)) { so_free_taps(&tv); return -1; }

    size_t row_elems = (size_t)w * 4;
    float *tmp = (float *)malloc(sizeof(float) * row_elems * nh);
    if (!tmp) { so_free_taps(&tv); so_free_taps(&th); return -1; }

    /* vertical_sample */
    for (uint32_t oy = 0; oy < nh; oy++) {
        float *acc = tmp + (size_t)oy * row_elems;
        const float *ws = tv.w + (size_t)oy * tv.max_taps;
        for (size_t j = 0; j < row_elems; j++) acc[j] = 0.0f;
        for (int i = 0; i < tv.count[oy]; i++) {
            const uint8_t *row = src + (size_t)(tv.left[oy] + i) * row_elems;
            float wi = ws[i];
            for (size_t j = 0; j < row_elems; j++) acc[j] += (float)row[j] * wi;
        }
    }

    /* horizontal_sample */
    for (uint32_t y = 0; y < nh; y++) {
        const float *in = tmp + (size_t)y * row_elems;
        uint8_t *out = dst + (size_t)y * nw * 4;
        for (uint32_t ox = 0; ox < nw; ox++) {
            const float *ws = th.w + (size_t)ox * th.max_taps;
            const float *p = in + (size_t)th.left[ox] * 4;
            float t0 = 0.0f, t1 = 0.0f, t2 = 0.0f, t3 = 0.0f;
            for (int i = 0; i < th.count[ox]; i++) {
                float wi = ws[i];
                t0 += p[i * 4 + 0] * wi;
                t1 += p[i * 4 + 1] * wi;
                t2 += p[i * 4 + 2] * wi;
                t3 += p[i * 4 + 3] * wi;
            }
            float t[4] = { t0, t1, t2, t3 };
            for (int c = 0; c < 4; c++) {
                float v = t[c];
                v = v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v);   /* clamp(t, min, max) */
                out[ox * 4 + c] = (uint8_t)roundf(v);               /* FloatNearest: f32::round */
            }
        }
    }

    free(tmp);
    so_free_taps(&tv); so_free_taps(&th);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * JPEG pre-entropy front end.  File:line references are to the vendored libjpeg-turbo 2.1.5.1
 * under crates/libjxl-sys/libjxl/third_party/libjpeg-turbo/ (MozJPEG shares this structure).
 * ---------------------------------------------------------------------------------------- */

static uint32_t ceil_div(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a + b - 1) / b); }
static uint32_t round_up(uint32_t a, uint32_t b) { return ceil_div(a, b) * b; }

/* jcmaster.c:137-171: width_in_blocks = ceil(w * h_samp / (max_h_samp * 8)); YCbCr 4:2:0 is
 * (2x2, 1x1, 1x1) (jcparam.c:365-373).  Coefficient arrays are padded to whole MCUs
 * (jccoefct.c:245-327 creates the dummy blocks). */
void so_jpeg_geometry(uint32_t w, uint32_t h, so_jpeg_geom *g)
{
    g->width = w; g->height = h;
    g->y_wb = ceil_div(w, 8);  g->y_hb = ceil_div(h, 8);
    g->c_wb = ceil_div(w, 16); g->c_hb = ceil_div(h, 16);
    g->y_WB = round_up(g->y_wb, 2); g->y_HB = round_up(g->y_hb, 2);
    g->c_WB = g->c_wb;              g->c_HB = g->c_hb;
}

/* jcparam.c:73-92 */
static const uint16_t k_annexk_luma[64] = {
    16, 11, 10, 16, 24, 40, 51, 61,  12, 12, 14, 19, 26, 58, 60, 55,
    14, 13, 16, 24, 40, 57, 69, 56,  14, 17, 22, 29, 51, 87, 80, 62,
    18, 22, 37, 56, 68, 109, 103, 77,  24, 35, 55, 64, 81, 104, 113, 92,
    49, 64, 78, 87, 103, 121, 120, 101,  72, 92, 95, 98, 112, 100, 103, 99 };
static const uint16_t k_annexk_chroma[64] = {
    17, 18, 24, 47, 99, 99, 99, 99,  18, 21, 26, 66, 99, 99, 99, 99,
    24, 26, 56, 99, 99, 99, 99, 99,  47, 66, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99,  99, 99, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99,  99, 99, 99, 99, 99, 99, 99, 99 };
/* MozJPEG quant_tbl_master_idx 3 ("ImageMagick table by N. Robidoux"), luma == chroma.
 * mozjpeg-sys 2.2.3 is NOT in the reference tree: restated from memory, parity unpinned. */
static const uint16_t k_mozjpeg_tbl3[64] = {
    16, 16, 16, 18, 25, 37, 56, 85,  16, 17, 20, 27, 34, 40, 53, 75,
    16, 20, 24, 31, 43, 62, 91, 135,  18, 27, 31, 40, 53, 74, 106, 156,
    25, 34, 43, 53, 69, 94, 131, 189,  37, 40, 62, 74, 94, 124, 169, 238,
    56, 53, 91, 106, 131, 169, 226, 311,  85, 75, 135, 156, 189, 238, 311, 418 };

/* jcparam.c:130-152 (quality -> percent scale) and :28-65 (scale a base table;
 * the mozjpeg crate calls jpeg_set_quality with force_baseline = FALSE). */
void so_jpeg_quant_table(int quality, int which, int base, uint16_t out[64])
{
    if (quality <= 0) quality = 1;
    if (quality > 100) quality = 100;
    int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
    const uint16_t *tbl = base == 1 ? k_mozjpeg_tbl3 : (which == 0 ? k_annexk_luma : k_annexk_chroma);
    for (int i = 0; i < 64; i++) {
        long t = ((long)tbl[i] * scale + 50L) / 100L;
        if (t <= 0) t = 1;
        if (t > 32767) t = 32767;
        out[i] = (uint16_t)t;
    }
}

/* jccolor.c:64-67,209-237 (table contents) folded into jccolext.c:61-68 (their use). */
#define SO_FIX(x) ((int32_t)((x) * 65536.0 + 0.5))
void so_rgb_ycc(uint8_t r, uint8_t g, uint8_t b, uint8_t out[3])
{
    const int32_t half = 1 << 15, cboff = 128 << 16;
    int32_t y  = SO_FIX(0.29900) * r + SO_FIX(0.58700) * g + SO_FIX(0.11400) * b + half;
    int32_t cb = (-SO_FIX(0.16874)) * r + (-SO_FIX(0.33126)) * g + SO_FIX(0.50000) * b + cboff + half - 1;
    int32_t cr = SO_FIX(0.50000) * r + (-SO_FIX(0.41869)) * g + (-SO_FIX(0.08131)) * b + cboff + half - 1;
    out[0] = (uint8_t)(y >> 16); out[1] = (uint8_t)(cb >> 16); out[2] = (uint8_t)(cr >> 16);
}

/* jfdctint.c:143-286, CONST_BITS 13, PASS1_BITS 2 */
#define DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))
static void fdct_1d(const int32_t in[8], int32_t out[8], int pass)
{
    int32_t tmp0 = in[0] + in[7], tmp7 = in[0] - in[7];
    int32_t tmp1 = in[1] + in[6], tmp6 = in[1] - in[6];
    int32_t tmp2 = in[2] + in[5], tmp5 = in[2] - in[5];
    int32_t tmp3 = in[3] + in[4], tmp4 = in[3] - in[4];
    int32_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3;
    int32_t tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    int sh = pass == 1 ? 13 - 2 : 13 + 2;
    if (pass == 1) { out[0] = (tmp10 + tmp11) * 4; out[4] = (tmp10 - tmp11) * 4; }
    else           { out[0] = DESCALE(tmp10 + tmp11, 2); out[4] = DESCALE(tmp10 - tmp11, 2); }
    int32_t z1 = (tmp12 + tmp13) * 4433;
    out[2] = DESCALE(z1 + tmp13 * 6270, sh);
    out[6] = DESCALE(z1 + tmp12 * (-15137), sh);
    z1 = tmp4 + tmp7;
    int32_t z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7;
    int32_t z5 = (z3 + z4) * 9633;
    tmp4 *= 2446; tmp5 *= 16819; tmp6 *= 25172; tmp7 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    out[7] = DESCALE(tmp4 + z1 + z3, sh);
    out[5] = DESCALE(tmp5 + z2 + z4, sh);
    out[3] = DESCALE(tmp6 + z2 + z3, sh);
    out[1] = DESCALE(tmp7 + z1 + z4, sh);
}

void so_fdct_islow(int32_t blk[64])
{
    int32_t in[8], out[8];
    for (int r = 0; r < 8; r++) {
        fdct_1d(blk + r * 8, out, 1);
        memcpy(blk + r * 8, out, sizeof out);
    }
    for (int c = 0; c < 8; c++) {
        for (int r = 0; r < 8; r++) in[r] = blk[r * 8 + c];
        fdct_1d(in, out, 2);
        for (int r = 0; r < 8; r++) blk[r * 8 + c] = out[r];
    }
}

/* jcdctmgr.c:173-221 (compute_reciprocal, DCTELEM = int in the non-SIMD build, divisor =
 * quantval << 3 per :245-251) and :408-438 (quantize, 8-bit branch). */
void so_quantize(const int32_t ws[64], const uint16_t qt[64], int16_t out[64])
{
    for (int i = 0; i < 64; i++) {
        uint32_t divisor = (uint32_t)qt[i] << 3;
        int b = 31 - __builtin_clz(divisor);          /* flss(divisor) - 1 */
        int r = 32 + b;
        uint64_t fq = ((uint64_t)1 << r) / divisor, fr = ((uint64_t)1 << r) % divisor;
        uint32_t c = divisor / 2;
        if (fr == 0) { fq >>= 1; r--; }
        else if (fr <= divisor / 2u) c++;
        else fq++;
        uint32_t recip = (uint32_t)fq, corr = c;
        int32_t t = ws[i];
        int neg = t < 0;
        if (neg) t = -t;
        uint64_t product = (uint64_t)((uint32_t)t + corr) * recip;
        product >>= r;                                /* shift + sizeof(DCTELEM)*8 == r */
        int32_t q = (int32_t)(uint32_t)product;
        out[i] = (int16_t)(neg ? -q : q);
    }
}

/* qt == NULL: keep the DCT output un-quantised (what MozJPEG copies aside for its trellis pass, SURVEY.md App. B --
 * mozjpeg-sys is not in the reference tree, so only the value, jfdctint.c's output, is pinned; that MozJPEG stores it
 * per block is restated from memory) */
/* MozJPEG's overshoot deringing (jcdctmgr.c preprocess_deringing + catmull_rom of mozjpeg-sys 2.2.3 -- NOT in the
 * reference tree: restated from the published source, PARITY UNPINNED).  Runs of samples at the maximum (walked in zig-zag
 * order) are replaced by a Catmull-Rom curve that overshoots by at most min(31, 2 * DC quantiser, headroom of the block's DC):
 * the clipped whites of text and line art then cost fewer bits and ring less after decoding. */
static int so_dering = 0;      /* set by so_jpeg_frontend_mozjpeg for the duration of the call (test infrastructure, single-threaded) */
static const int so_zigzag[64] = { 0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14,
                                   21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60,
                                   61, 54, 47, 55, 62, 63 };          /* jpeg_natural_order, jutils.c:56 */
static float so_catmull_rom(int v1, int v2, int v3, int v4, float t, int size)
{
    const int tan1 = (v3 - v1) * size, tan2 = (v4 - v2) * size;
    const float t2 = t * t, t3 = t2 * t;
    const float f1 = 2.f * t3 - 3.f * t2 + 1.f, f2 = -2.f * t3 + 3.f * t2, f3 = t3 - 2.f * t2 + t, f4 = t3 - t2;
    return v2 * f1 + tan1 * f3 + v3 * f2 + tan2 * f4;
}
void so_preprocess_deringing(int32_t data[64], uint16_t dc_quant)
{
    const int maxsample = 255 - 128, size = 64;
    int sum = 0, count = 0;
    for (int i = 0; i < size; i++) { sum += data[i]; if (data[i] >= maxsample) count++; }
    if (!count || count == size) return;
    int room = (maxsample * size - sum) / count, cap = 2 * (int)dc_quant < 31 ? 2 * (int)dc_quant : 31;
    const int maxovershoot = maxsample + (cap < room ? cap : room);
    int n = 0;
    do {
        if (data[so_zigzag[n]] < maxsample) { n++; continue; }
        const int start = n;
        while (++n < size && data[so_zigzag[n]] >= maxsample) {}
        const int end = n;
        const int f1 = data[so_zigzag[start >= 1 ? start - 1 : 0]], f2 = data[so_zigzag[start >= 2 ? start - 2 : 0]];
        const int l1 = data[so_zigzag[end < size - 1 ? end : size - 1]], l2 = data[so_zigzag[end < size - 2 ? end + 1 : size - 1]];
        int fslope = f1 - f2 > maxsample - f1 ? f1 - f2 : maxsample - f1;
        int lslope = l1 - l2 > maxsample - l1 ? l1 - l2 : maxsample - l1;
        if (start == 0) fslope = lslope;
        if (end == size) lslope = fslope;
        const int length = end - start;
        const float step = 1.f / (float)(length + 1);
        float position = step;
        for (int i = start; i < end; i++, position += step) {
            const int tmp = (int)ceilf(so_catmull_rom(maxsample - fslope, maxsample, maxsample, maxsample - lslope, position, length));
            data[so_zigzag[i]] = tmp < maxovershoot ? tmp : maxovershoot;
        }
        n++;
    } while (n < size);
}

/* qt == NULL: keep the DCT output un-quantised (what MozJPEG copies aside for its trellis pass, SURVEY.md App. B --
 * mozjpeg-sys is not in the reference tree, so only the value, jfdctint.c's output, is pinned; that MozJPEG stores it
 * per block is restated from memory) */
static void so_do_block(const uint8_t *plane, size_t stride, uint32_t bx, uint32_t by,
                        const uint16_t qt[64], int16_t *out)
{
    int32_t ws[64];
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++)
            ws[r * 8 + c] = (int32_t)plane[((size_t)by * 8 + r) * stride + (size_t)bx * 8 + c] - 128; /* convsamp :373-400 */
    if (so_dering && qt) so_preprocess_deringing(ws, qt[0]);
    so_fdct_islow(ws);
    if (qt) so_quantize(ws, qt, out);
    else for (int i = 0; i < 64; i++) out[i] = (int16_t)ws[i];
    if (so_dering && qt)           /* MozJPEG clamps the quantised values to MAX_COEF_BITS = 10 bits when deringing is on */
        for (int i = 0; i < 64; i++) out[i] = out[i] < -1023 ? -1023 : out[i] > 1023 ? 1023 : out[i];
}

static void so_component(const uint8_t *plane, size_t stride, uint32_t wb, uint32_t hb,
                         uint32_t WB, uint32_t HB, uint32_t hs, const uint16_t qt[64], int16_t *coef)
{
    for (uint32_t by = 0; by < hb; by++) {
        int16_t *row = coef + (size_t)by * WB * 64;
        for (uint32_t bx = 0; bx < wb; bx++) so_do_block(plane, stride, bx, by, qt, row + (size_t)bx * 64);
        for (uint32_t bx = wb; bx < WB; bx++) {
            memset(row + (size_t)bx * 64, 0, 128);
            row[(size_t)bx * 64] = row[(size_t)(wb - 1) * 64];
        }
    }
    for (uint32_t by = hb; by < HB; by++) {
        int16_t *row = coef + (size_t)by * WB * 64, *last = row - (size_t)WB * 64;
        memset(row, 0, (size_t)WB * 128);
        for (uint32_t m = 0; m < WB / hs; m++) {
            int16_t dc = last[(size_t)(m * hs + hs - 1) * 64];
            for (uint32_t bi = 0; bi < hs; bi++) row[(size_t)(m * hs + bi) * 64] = dc;
        }
    }
}

void so_jpeg_frontend(const uint8_t *rgba, uint32_t w, uint32_t h,
                      const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                      int16_t *yc, int16_t *cbc, int16_t *crc)
{
    so_jpeg_geom g;
    so_jpeg_geometry(w, h, &g);
    /* full-resolution planes padded so every real block and every 2x2 chroma source exists:
     * right edge replicated (jcsample.c:94-111 via :201 and :268), bottom row replicated
     * (jcprepct.c:109-118, :155-162 before downsampling; :174-183 after). */
    size_t pw = (size_t)g.c_wb * 16 > (size_t)g.y_wb * 8 ? (size_t)g.c_wb * 16 : (size_t)g.y_wb * 8;
    size_t ph = (size_t)g.c_hb * 16 > (size_t)g.y_hb * 8 ? (size_t)g.c_hb * 16 : (size_t)g.y_hb * 8;
    uint8_t *Y = (uint8_t *)malloc(pw * ph), *B = (uint8_t *)malloc(pw * ph), *R = (uint8_t *)malloc(pw * ph);
    size_t cw = (size_t)g.c_wb * 8, chh = (size_t)g.c_hb * 8;
    uint8_t *Cb = (uint8_t *)malloc(cw * chh), *Cr = (uint8_t *)malloc(cw * chh);
    for (size_t y = 0; y < ph; y++) {
        size_t sy = y < h ? y : h - 1;
        for (size_t x = 0; x < pw; x++) {
            size_t sx = x < w ? x : w - 1;
            const uint8_t *p = rgba + (sy * w + sx) * 4;   /* alpha dropped: codec/mod.rs:42-48 */
            uint8_t o[3];
            so_rgb_ycc(p[0], p[1], p[2], o);
            Y[y * pw + x] = o[0]; B[y * pw + x] = o[1]; R[y * pw + x] = o[2];
        }
    }
    /* h2v2_downsample, jcsample.c:255-285: bias 1,2,1,2,... restarting at 1 every row.
     * Chroma rows past ceil(h/2) replicate the last DOWNSAMPLED row (jcprepct.c:174-183),
     * which for an even height is NOT what downsampling two copies of the last row gives. */
    size_t creal = ((size_t)h + 1) / 2;
    for (size_t y = 0; y < chh; y++) {
        int bias = 1;
        if (y >= creal) {
            memcpy(Cb + y * cw, Cb + (creal - 1) * cw, cw);
            memcpy(Cr + y * cw, Cr + (creal - 1) * cw, cw);
            continue;
        }
        for (size_t x = 0; x < cw; x++) {
            const uint8_t *b0 = B + (2 * y) * pw + 2 * x, *b1 = b0 + pw;
            const uint8_t *r0 = R + (2 * y) * pw + 2 * x, *r1 = r0 + pw;
            Cb[y * cw + x] = (uint8_t)((b0[0] + b0[1] + b1[0] + b1[1] + bias) >> 2);
            Cr[y * cw + x] = (uint8_t)((r0[0] + r0[1] + r1[0] + r1[1] + bias) >> 2);
            bias ^= 3;
        }
    }
    so_component(Y, pw, g.y_wb, g.y_hb, g.y_WB, g.y_HB, 2, qt_luma, yc);
    so_component(Cb, cw, g.c_wb, g.c_hb, g.c_WB, g.c_HB, 1, qt_chroma, cbc);
    so_component(Cr, cw, g.c_wb, g.c_hb, g.c_WB, g.c_HB, 1, qt_chroma, crc);
    free(Y); free(B); free(R); free(Cb); free(Cr);
}

/* The un-quantised coefficient planes (same layout, dummy blocks made the same way from the un-quantised DCs). */
void so_jpeg_frontend_unquantised(const uint8_t *rgba, uint32_t w, uint32_t h, int16_t *yc, int16_t *cbc, int16_t *crc)
{
    so_jpeg_frontend(rgba, w, h, NULL, NULL, yc, cbc, crc);
}

/* ======================================================================================================================
 * Colour-plane pre-stages of the other encoders.  BOTH ARE PARITY-UNPINNED: libwebp (libwebp-sys 0.9.6)
 * and ravif 0.13 are crates.io dependencies that are absent from /root/reference; the arithmetic below restates their
 * published sources from memory, and the reference's tests at these boundaries check only magic bytes and dimensions
 * (crates/slimg-core/src/codec/webp.rs:25-30 Encoder::from_rgba -> WebPPictureImportRGBA; codec/avif.rs:28-47
 * ravif::Encoder::encode_rgba).
 * ====================================================================================================================== */
#include <math.h>

/* ---- libwebp src/enc/picture_csp_enc.c: ImportYUVAFromRGBA, the default (non-iterative) path ---- */
enum { kWGammaFix = 12, kWGammaScale = (1 << 12) - 1, kWGammaTabFix = 7, kWGammaTabScale = 1 << 7, kWGammaTabRounder = 1 << 6,
       kWGammaTabSize = 1 << (12 - 7), kWAlphaFix = 19, kWYuvFix = 16, kWYuvHalf = 1 << 15 };
static uint16_t w_gamma_to_linear[256];
static int w_linear_to_gamma[kWGammaTabSize + 1];
static int w_tables_ready = 0;

void so_webp_gamma_tables(uint16_t g2l[256], int32_t l2g[33])       /* InitGammaTables: kGamma = 0.80 */
{
    const double gamma = 0.80, norm = 1.0 / 255.0, scale = (double)(1 << kWGammaTabFix) / kWGammaScale;
    for (int v = 0; v <= 255; v++) g2l[v] = (uint16_t)(pow(norm * v, gamma) * kWGammaScale + .5);
    for (int v = 0; v <= kWGammaTabSize; v++) l2g[v] = (int32_t)(255. * pow(scale * v, 1. / gamma) + .5);
}
static void w_init(void)
{
    if (w_tables_ready) return;
    int32_t l2g[33];
    so_webp_gamma_tables(w_gamma_to_linear, l2g);
    for (int i = 0; i <= kWGammaTabSize; i++) w_linear_to_gamma[i] = l2g[i];
    w_tables_ready = 1;
}
static uint32_t w_g2l(uint8_t v) { return w_gamma_to_linear[v]; }
static int w_l2g(uint32_t base_value, int shift)                    /* LinearToGamma = (Interpolate(v << shift) + rounder) >> fix */
{
    const uint32_t v = base_value << shift;
    const int tab_pos = (int)(v >> (kWGammaTabFix + 2));
    const int x = (int)(v & ((kWGammaTabScale << 2) - 1));
    const int v0 = w_linear_to_gamma[tab_pos], v1 = w_linear_to_gamma[tab_pos + 1];
    const int y = v1 * x + v0 * ((kWGammaTabScale << 2) - x);
    return (y + kWGammaTabRounder) >> kWGammaTabFix;
}
static int w_rgb_to_y(int r, int g, int b, int rounding) { return (16839 * r + 33059 * g + 6420 * b + rounding + (16 << kWYuvFix)) >> kWYuvFix; }
static int w_clip_uv(int uv, int rounding)
{
    uv = (uv + rounding + (128 << (kWYuvFix + 2))) >> (kWYuvFix + 2);
    return (uv & ~0xff) == 0 ? uv : uv < 0 ? 0 : 255;
}
static int w_rgb_to_u(int r, int g, int b, int rounding) { return w_clip_uv(-9719 * r - 19081 * g + 28800 * b, rounding); }
static int w_rgb_to_v(int r, int g, int b, int rounding) { return w_clip_uv(28800 * r - 24116 * g - 4684 * b, rounding); }

/* One chroma sample from the (up to) 2 x 2 pixels p[0..3] (p[1] = right neighbour or p[0] again at an odd right edge, p[2..3]
 * = the row below or the same row again at an odd bottom edge): AccumulateRGB / AccumulateRGBA + WebPConvertRGBA32ToUV. */
static void w_chroma(const uint8_t *p[4], int wide, int with_alpha, uint8_t *u, uint8_t *v)
{
    int rgb[3];
    uint32_t a = wide ? (uint32_t)(p[0][3] + p[1][3] + p[2][3] + p[3][3]) : 2u * (uint32_t)(p[0][3] + p[2][3]);
    for (int c = 0; c < 3; c++) {
        if (!with_alpha || a == 4 * 0xff || a == 0) {
            rgb[c] = wide ? w_l2g(w_g2l(p[0][c]) + w_g2l(p[1][c]) + w_g2l(p[2][c]) + w_g2l(p[3][c]), 0)       /* SUM4 */
                          : w_l2g(w_g2l(p[0][c]) + w_g2l(p[2][c]), 1);                                        /* SUM2 */
        } else {                                                                                               /* LinearToGammaWeighted */
            const uint32_t sum = p[0][3] * w_g2l(p[0][c]) + p[1][3] * w_g2l(p[1][c]) + p[2][3] * w_g2l(p[2][c]) + p[3][3] * w_g2l(p[3][c]);
            const uint32_t inv = (1u << kWAlphaFix) / a;                                                       /* kInvAlpha[a] */
            rgb[c] = w_l2g((uint32_t)(((uint64_t)sum * inv) >> (kWAlphaFix - 2)), 0);                          /* DIVIDE_BY_ALPHA */
        }
    }
    *u = (uint8_t)w_rgb_to_u(rgb[0], rgb[1], rgb[2], kWYuvHalf << 2);
    *v = (uint8_t)w_rgb_to_v(rgb[0], rgb[1], rgb[2], kWYuvHalf << 2);
}

/* y: w x h, u / v: ceil(w/2) x ceil(h/2), a: w x h (written only when some alpha != 255); returns has_alpha */
int so_webp_yuva(const uint8_t *rgba, uint32_t w, uint32_t h, uint8_t *y, uint8_t *u, uint8_t *v, uint8_t *a)
{
    w_init();
    int has_alpha = 0;
    for (size_t i = 0; i < (size_t)w * h; i++) has_alpha |= rgba[4 * i + 3] != 0xff;          /* CheckNonOpaque */
    const uint32_t uvw = (w + 1) / 2;
    for (uint32_t yy = 0; yy < h; yy++)
        for (uint32_t x = 0; x < w; x++) {
            const uint8_t *p = rgba + ((size_t)yy * w + x) * 4;
            y[(size_t)yy * w + x] = (uint8_t)w_rgb_to_y(p[0], p[1], p[2], kWYuvHalf);
            if (has_alpha) a[(size_t)yy * w + x] = p[3];
        }
    for (uint32_t cy = 0; cy < (h + 1) / 2; cy++) {
        const uint32_t y0 = 2 * cy, y1 = (2 * cy + 1 < h) ? 2 * cy + 1 : 2 * cy;              /* odd height: the last row twice */
        /* rows_have_alpha: the weighted path is taken only when these two rows contain some alpha != 255 */
        int rows_alpha = 0;
        if (has_alpha)
            for (uint32_t x = 0; x < w; x++) rows_alpha |= rgba[((size_t)y0 * w + x) * 4 + 3] != 0xff || rgba[((size_t)y1 * w + x) * 4 + 3] != 0xff;
        for (uint32_t cx = 0; cx < uvw; cx++) {
            const uint32_t x0 = 2 * cx;
            const int wide = x0 + 1 < w;
            const uint32_t x1 = wide ? x0 + 1 : x0;
            const uint8_t *p[4] = { rgba + ((size_t)y0 * w + x0) * 4, rgba + ((size_t)y0 * w + x1) * 4,
                                    rgba + ((size_t)y1 * w + x0) * 4, rgba + ((size_t)y1 * w + x1) * 4 };
            w_chroma(p, wide, rows_alpha, &u[(size_t)cy * uvw + cx], &v[(size_t)cy * uvw + cx]);
        }
    }
    return has_alpha;
}

/* ---- ravif src/av1encoder.rs: rgb_to_ycbcr (BT.601, full range, 4:4:4), 8 or 10 bits; alpha as its own plane ---- */
static float so_f32_round(float x) { return x < 0.0f ? -floorf(-x + 0.5f) : floorf(x + 0.5f); }   /* f32::round: half away from zero */
void so_avif_planes(const uint8_t *rgba, uint32_t w, uint32_t h, int depth, uint16_t *y, uint16_t *cb, uint16_t *cr, uint16_t *a)
{
    const float m0 = 0.2990f, m1 = 0.5870f, m2 = 0.1140f;
    const float max_value = (float)((1 << depth) - 1);
    const float scale = max_value / 255.f;
    const float shift = so_f32_round(max_value * 0.5f);
    const float kb = 0.5f / (1.f - m2), kr = 0.5f / (1.f - m0);
    for (size_t i = 0; i < (size_t)w * h; i++) {
        const float r = (float)rgba[4 * i], g = (float)rgba[4 * i + 1], b = (float)rgba[4 * i + 2];
        volatile float t0 = scale * m0, t1 = scale * m1, t2 = scale * m2;          /* no contraction: every product rounds to f32 */
        volatile float p0 = t0 * r, p1 = t1 * g, p2 = t2 * b;
        volatile float s01 = p0 + p1;
        const float yv = s01 + p2;
        volatile float db = b * scale, dr = r * scale;
        const float cbv = fmaf(db - yv, kb, shift), crv = fmaf(dr - yv, kr, shift);  /* mul_add: ONE rounding */
        const float yr = so_f32_round(yv), br = so_f32_round(cbv), rr = so_f32_round(crv);
        y[i] = (uint16_t)(yr < 0 ? 0 : yr > max_value ? max_value : yr);            /* `as u8 / u16` saturates */
        cb[i] = (uint16_t)(br < 0 ? 0 : br > max_value ? max_value : br);
        cr[i] = (uint16_t)(rr < 0 ? 0 : rr > max_value ? max_value : rr);
        /* alpha plane: 8 bit as is; 10 bit as ravif's to_ten(): (x << 2) | (x >> 6) */
        a[i] = depth == 8 ? rgba[4 * i + 3] : (uint16_t)(((uint16_t)rgba[4 * i + 3] << 2) | (rgba[4 * i + 3] >> 6));
    }
}

/* The planes as MozJPEG's front end computes them with its defaults (overshoot_deringing = TRUE).  UNPINNED. */
void so_jpeg_frontend_mozjpeg(const uint8_t *rgba, uint32_t w, uint32_t h, const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                              int16_t *yc, int16_t *cbc, int16_t *crc)
{
    so_dering = 1;
    so_jpeg_frontend(rgba, w, h, qt_luma, qt_chroma, yc, cbc, crc);
    so_dering = 0;
}
