/*
 * slimg_oracle.h -- CPU oracle for the slimg pixel pipeline.  TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference algorithms on the hot path.
 * Nothing under slimg_b200/ (the product) may include, link or call this file; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, as the
 * checker and the timed CPU baseline.
 *
 * Parity pins (see DESIGN.md "Oracle"):
 *   - geometry / crop / extend : pinned by the reference's own unit tests
 *     (crates/slimg-core/src/crop.rs:94-327, extend.rs:141-407, resize.rs:101-139,
 *      bindings/python/tests/test_operations.py) -- tests/test_oracle_geometry.py.
 *   - JPEG front end           : pinned against the vendored libjpeg-turbo 2.1.5.1 compiled from
 *     the reference tree (oracle/ref_ljt -> oracle/_ref/libljt_ref.so), bit-exact, and against
 *     fixtures generated from it (tests/golden/jpeg_*.npz).
 *   - Lanczos3 resize pixels   : PARITY UNPINNED.  The arithmetic lives in the third-party
 *     `image` crate 0.25.9 (Cargo.lock:2113-2116), absent from /root/reference; the
 *     reference's tests pin only output dimensions.  Restated from the crate's published
 *     algorithm (imageops/sample.rs); call sites crates/slimg-core/src/resize.rs:64-90.
 */
#ifndef SLIMG_ORACLE_H
#define SLIMG_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes = SlimgError tags (crates/slimg-ffi/src/lib.rs:195-227) */
enum { SO_OK = 0, SO_ERR_ENCODE = 5, SO_ERR_RESIZE = 6, SO_ERR_CROP = 7, SO_ERR_EXTEND = 8 };

/* ResizeMode (resize.rs:9-20): tag 1..5 = Width/Height/Exact/Fit/Scale */
typedef struct { int32_t tag; uint32_t a, b; double factor; } so_resize_mode;
/* CropMode (crop.rs:6-16): tag 1 = Region{x,y,w,h}, 2 = AspectRatio{w,h} */
typedef struct { int32_t tag; uint32_t x, y, w, h; } so_crop_mode;
/* ExtendMode (extend.rs:24-31): tag 1 = AspectRatio{w,h}, 2 = Size{w,h} */
typedef struct { int32_t tag; uint32_t w, h; } so_extend_mode;

/* ---- geometry ---- */
int so_calculate_dimensions(uint32_t ow, uint32_t oh, const so_resize_mode *m,
                            uint32_t *w, uint32_t *h);                 /* resize.rs:23-61 */
void so_image_resize_dimensions(uint32_t w, uint32_t h, uint32_t nw, uint32_t nh, int fill,
                                uint32_t *ow, uint32_t *oh);           /* image 0.25.9 math/utils.rs */
int so_resize_output_dims(uint32_t w, uint32_t h, const so_resize_mode *m,
                          uint32_t *ow, uint32_t *oh);                 /* resize.rs:64-90 + DynamicImage::resize */
int so_calculate_crop_region(uint32_t iw, uint32_t ih, const so_crop_mode *m,
                             uint32_t *x, uint32_t *y, uint32_t *w, uint32_t *h);   /* crop.rs:19-67 */
int so_calculate_extend_region(uint32_t iw, uint32_t ih, const so_extend_mode *m,
                               uint32_t *cw, uint32_t *ch, uint32_t *ox, uint32_t *oy); /* extend.rs:37-91 */

/* ---- pixel movers (dst preallocated by caller with the geometry above) ---- */
void so_crop_pixels(const uint8_t *src, uint32_t sw, uint32_t sh,
                    uint32_t x, uint32_t y, uint32_t cw, uint32_t ch, uint8_t *dst);      /* crop.rs:70-87 */
void so_extend_pixels(const uint8_t *src, uint32_t sw, uint32_t sh,
                      uint32_t cw, uint32_t ch, uint32_t ox, uint32_t oy,
                      const uint8_t fill[4], uint8_t *dst);                               /* extend.rs:94-133 */

/* ---- Lanczos3 resize: image 0.25.9 imageops::resize(…, FilterType::Lanczos3) ---- */
/* returns 0, or -1 on allocation failure */
int so_resize_exact(const uint8_t *src, uint32_t w, uint32_t h,
                    uint32_t nw, uint32_t nh, uint8_t *dst);
/* tap table of one axis, for tests: left[o], count[o], weights[o*max_taps + i]; returns max count */
int so_lanczos3_taps(uint32_t in_len, uint32_t out_len, int32_t *left, int32_t *count,
                     float *weights, int max_taps);

/* ---- JPEG pre-entropy front end (libjpeg-turbo structure) ---- */
typedef struct {
    uint32_t width, height;
    uint32_t y_wb, y_hb, y_WB, y_HB;     /* real / MCU-padded block counts, Y (2x2 sampling) */
    uint32_t c_wb, c_hb, c_WB, c_HB;     /* same for Cb and Cr (1x1 sampling) */
} so_jpeg_geom;
void so_jpeg_geometry(uint32_t w, uint32_t h, so_jpeg_geom *g);            /* jcmaster.c:137-171 */
/* which: 0 = luma, 1 = chroma; base: 0 = Annex K (jcparam.c:73-92), 1 = MozJPEG table set 3 (unpinned) */
void so_jpeg_quant_table(int quality, int which, int base, uint16_t out[64]);   /* jcparam.c:28-65,130-168 */
/* coefficient planes: [HB][WB][64] int16, natural (row-major) order inside a block */
void so_jpeg_frontend(const uint8_t *rgba, uint32_t w, uint32_t h,
                      const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                      int16_t *y, int16_t *cb, int16_t *cr);
/* pieces, exported for unit tests */
/* qt_luma == qt_chroma == NULL: un-quantised planes (jfdctint.c:143-286 output) */
void so_jpeg_frontend_unquantised(const uint8_t *rgba, uint32_t w, uint32_t h, int16_t *y, int16_t *cb, int16_t *cr);
void so_rgb_ycc(uint8_t r, uint8_t g, uint8_t b, uint8_t out[3]);          /* jccolor.c:209-237, jccolext.c:61-68 */
void so_fdct_islow(int32_t blk[64]);                                        /* jfdctint.c:143-286 */
void so_quantize(const int32_t ws[64], const uint16_t qt[64], int16_t out[64]); /* jcdctmgr.c:173-221,408-438 */

/* MozJPEG's defaults: overshoot deringing before the DCT, quantised values clamped to 10 bits (mozjpeg-sys 2.2.3, absent: unpinned) */
void so_preprocess_deringing(int32_t data[64], uint16_t dc_quant);
void so_jpeg_frontend_mozjpeg(const uint8_t *rgba, uint32_t w, uint32_t h, const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                              int16_t *y, int16_t *cb, int16_t *cr);

/* ---- colour-plane pre-stages of the WebP / AVIF encoders -- parity unpinned, see slimg_oracle.c ---- */
void so_webp_gamma_tables(uint16_t g2l[256], int32_t l2g[33]);
int so_webp_yuva(const uint8_t *rgba, uint32_t w, uint32_t h, uint8_t *y, uint8_t *u, uint8_t *v, uint8_t *a);
void so_avif_planes(const uint8_t *rgba, uint32_t w, uint32_t h, int depth, uint16_t *y, uint16_t *cb, uint16_t *cr, uint16_t *a);

#ifdef __cplusplus
}
#endif
#endif
