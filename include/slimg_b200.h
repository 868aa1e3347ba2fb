/*
 * slimg_b200.h -- C ABI of libslimg_b200.so: the B200 (sm_100a) implementation of slimg-core's
 * per-image pixel pipeline (crop -> extend -> Lanczos3 resize, and the JPEG pre-entropy front
 * end).  This is the drop-in boundary: plain pointers and sizes, no torch / C++ types.
 *
 * Every entry point names the reference interface it replaces (paths relative to the slimg
 * repository).  Status codes are the reference's SlimgError tags
 * (crates/slimg-ffi/src/lib.rs:195-227), so a binding maps them 1:1; error CONDITIONS match
 * crop.rs:19-67, extend.rs:37-104 and resize.rs:23-61.
 *
 * There is NO CPU fallback: every pixel-producing call runs CUDA kernels and fails with
 * SLIMG_ERR_IMAGE ("no CUDA device" / the CUDA error string) when it cannot.
 * All functions are thread-safe; slimg_b200_last_error() is thread-local.
 */
#ifndef SLIMG_B200_H
#define SLIMG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SLIMG_B200_ABI_VERSION 2

/* ---- status = SlimgError tag (crates/slimg-ffi/src/lib.rs:195-227) ---- */
typedef enum {
    SLIMG_OK = 0,
    SLIMG_ERR_UNSUPPORTED_FORMAT = 1,
    SLIMG_ERR_UNKNOWN_FORMAT = 2,
    SLIMG_ERR_ENCODING_NOT_SUPPORTED = 3,
    SLIMG_ERR_DECODE = 4,
    SLIMG_ERR_ENCODE = 5,
    SLIMG_ERR_RESIZE = 6,
    SLIMG_ERR_CROP = 7,
    SLIMG_ERR_EXTEND = 8,
    SLIMG_ERR_IO = 9,
    SLIMG_ERR_IMAGE = 10   /* also: CUDA failures, bad arguments */
} slimg_status;

/* ---- data types ---- */

/* ImageData (crates/slimg-core/src/codec/mod.rs:11-39): tightly packed RGBA8, row-major,
 * len == width*height*4.  `data` is host memory unless a function says "device". */
typedef struct {
    uint32_t width, height;
    uint8_t *data;
    size_t   len;
} slimg_image;

/* ResizeMode (crates/slimg-core/src/resize.rs:9-20; UniFFI tags slimg_ffi.kt:1759-1781) */
enum { SLIMG_RESIZE_WIDTH = 1, SLIMG_RESIZE_HEIGHT = 2, SLIMG_RESIZE_EXACT = 3,
       SLIMG_RESIZE_FIT = 4, SLIMG_RESIZE_SCALE = 5 };
typedef struct {
    int32_t  tag;
    uint32_t a, b;      /* Width{a} | Height{a} | Exact{a=w,b=h} | Fit{a=max_w,b=max_h} */
    double   factor;    /* Scale{factor} */
} slimg_resize_mode;

/* CropMode (crates/slimg-core/src/crop.rs:6-16) */
enum { SLIMG_CROP_REGION = 1, SLIMG_CROP_ASPECT_RATIO = 2 };
typedef struct {
    int32_t  tag;
    uint32_t x, y;            /* Region only */
    uint32_t width, height;   /* Region size | AspectRatio w:h */
} slimg_crop_mode;

/* ExtendMode (crates/slimg-core/src/extend.rs:24-31) */
enum { SLIMG_EXTEND_ASPECT_RATIO = 1, SLIMG_EXTEND_SIZE = 2 };
typedef struct {
    int32_t  tag;
    uint32_t width, height;
} slimg_extend_mode;

/* FillColor (crates/slimg-core/src/extend.rs:6-21) */
enum { SLIMG_FILL_SOLID = 1, SLIMG_FILL_TRANSPARENT = 2 };
typedef struct {
    int32_t tag;
    uint8_t r, g, b, a;       /* Solid only; Transparent == (0,0,0,0) */
} slimg_fill_color;

/* The pixel half of PipelineOptions (crates/slimg-core/src/pipeline.rs:13-26): format/quality
 * belong to the encoder that stays on the reference libraries.  has_* == 0 means None.
 * has_fill == 0 with has_extend != 0 means opaque white (pipeline.rs:78-80). */
typedef struct {
    int32_t has_resize;  slimg_resize_mode resize;
    int32_t has_crop;    slimg_crop_mode   crop;
    int32_t has_extend;  slimg_extend_mode extend;
    int32_t has_fill;    slimg_fill_color  fill;
} slimg_pipeline_options;

/* Which order the three stages run in.  CONVERT is the reference's `convert`
 * (pipeline.rs:71-89: crop -> extend -> resize, fill pixels take part in the filter taps);
 * CHAIN is the standalone-op chaining the bindings document (SlimgImage.kt:9-14,
 * BASELINE.json's "resize -> crop -> extend"). */
enum { SLIMG_ORDER_CONVERT = 0, SLIMG_ORDER_CHAIN = 1 };

/* ---- geometry (host arithmetic, f64, Rust round() semantics) ---- */

/* resize.rs:23-61 calculate_dimensions */
slimg_status slimg_b200_calculate_dimensions(uint32_t w, uint32_t h, const slimg_resize_mode *mode,
                                             uint32_t *out_w, uint32_t *out_h);
/* resize.rs:64-90: the dimensions `resize` actually returns (image 0.25.9 re-derives them for
 * every mode but Exact) */
slimg_status slimg_b200_resize_output_dims(uint32_t w, uint32_t h, const slimg_resize_mode *mode,
                                           uint32_t *out_w, uint32_t *out_h);
/* crop.rs:19-67 calculate_crop_region */
slimg_status slimg_b200_calculate_crop_region(uint32_t w, uint32_t h, const slimg_crop_mode *mode,
                                              uint32_t *x, uint32_t *y, uint32_t *cw, uint32_t *ch);
/* extend.rs:37-91 calculate_extend_region */
slimg_status slimg_b200_calculate_extend_region(uint32_t w, uint32_t h, const slimg_extend_mode *mode,
                                                uint32_t *cw, uint32_t *ch, uint32_t *off_x, uint32_t *off_y);
/* Output size of the whole pixel pipeline in the given order. */
slimg_status slimg_b200_pipeline_output_dims(uint32_t w, uint32_t h, const slimg_pipeline_options *opts,
                                             int order, uint32_t *out_w, uint32_t *out_h);

/* The rectangle of the source image the pipeline actually reads for these options (every filter tap with a
 * non-zero weight behind the kept output pixels): host-fed calls upload only this window, and a caller that
 * stages its own device buffers may do the same.  win_w == 0 means no source pixel is read. */
slimg_status slimg_b200_pipeline_source_window(uint32_t w, uint32_t h, const slimg_pipeline_options *opts, int order,
                                               uint32_t *x, uint32_t *y, uint32_t *win_w, uint32_t *win_h);

/* ---- single-image operations, host buffers in / host buffers out ----
 * `out->data` is allocated by the library (release with slimg_b200_image_free). */

/* slimg_core::crop::crop (crop.rs:70-87); FFI crop (crates/slimg-ffi/src/lib.rs:320-325) */
slimg_status slimg_b200_crop(const slimg_image *in, const slimg_crop_mode *mode, slimg_image *out);
/* slimg_core::extend::extend (extend.rs:94-133); FFI extend (lib.rs:327-332) */
slimg_status slimg_b200_extend(const slimg_image *in, const slimg_extend_mode *mode,
                               const slimg_fill_color *fill, slimg_image *out);
/* slimg_core::resize::resize (resize.rs:64-90); FFI resize (lib.rs:334-339) */
slimg_status slimg_b200_resize(const slimg_image *in, const slimg_resize_mode *mode, slimg_image *out);
/* The pixel stages of slimg_core::convert (pipeline.rs:66-89), fused into one launch.
 * order = SLIMG_ORDER_CONVERT reproduces `convert`; SLIMG_ORDER_CHAIN runs resize->crop->extend. */
slimg_status slimg_b200_pipeline(const slimg_image *in, const slimg_pipeline_options *opts, int order,
                                 slimg_image *out);
void slimg_b200_image_free(slimg_image *img);

/* ---- JPEG pre-entropy front end: what JpegCodec::encode (codec/jpeg.rs:51-72) makes MozJPEG
 * compute before its entropy coder -- to_rgb (codec/mod.rs:42-48), RGB->YCbCr, 4:2:0 h2v2
 * downsample, 8x8 islow FDCT, quantisation.  Coefficients are int16, natural order within a
 * block, planes laid out [HB][WB][64] with MCU-padded block counts (dummy blocks included, as
 * jccoefct.c:286-319 creates them). ---- */
typedef struct {
    uint32_t width, height;
    uint32_t y_wb, y_hb, y_WB, y_HB;   /* real / padded block counts of Y */
    uint32_t c_wb, c_hb, c_WB, c_HB;   /* same for Cb and Cr */
} slimg_jpeg_geometry;

typedef struct {
    slimg_jpeg_geometry geom;
    uint16_t qt_luma[64], qt_chroma[64];
    int16_t *y, *cb, *cr;              /* host; released by slimg_b200_jpeg_coefs_free */
    /* the same planes BEFORE quantisation (the DCT output, scaled by 8 as jfdctint.c leaves it): what MozJPEG keeps
     * per block for its trellis pass.  NULL unless SLIMG_JPEG_UNQUANTISED was asked for. */
    int16_t *y_uq, *cb_uq, *cr_uq;
} slimg_jpeg_coefs;

enum { SLIMG_QT_ANNEX_K = 0,            /* libjpeg-turbo jcparam.c:73-92 */
       SLIMG_QT_MOZJPEG = 1 };          /* MozJPEG default table set 3, luma == chroma */

/* flags of the JPEG entry points */
enum { SLIMG_JPEG_UNQUANTISED = 1,      /* also return the un-quantised coefficient planes */
       SLIMG_JPEG_DERINGING = 2 };      /* MozJPEG's front-end defaults: overshoot deringing of blocks that hold clipped
                                           whites before the DCT, quantised values clamped to 10 bits.  mozjpeg-sys is not in the reference
                                           tree: restated from its published source, parity unpinned. */

void slimg_b200_jpeg_geometry(uint32_t w, uint32_t h, slimg_jpeg_geometry *g);
/* jpeg_set_quality(q, force_baseline = FALSE): jcparam.c:28-65,130-168 */
void slimg_b200_jpeg_quant_table(int quality, int which /*0 luma,1 chroma*/, int base, uint16_t out[64]);
slimg_status slimg_b200_jpeg_frontend(const slimg_image *in, const uint16_t qt_luma[64],
                                      const uint16_t qt_chroma[64], slimg_jpeg_coefs *out);
slimg_status slimg_b200_jpeg_frontend_ex(const slimg_image *in, const uint16_t qt_luma[64],
                                         const uint16_t qt_chroma[64], uint32_t flags, slimg_jpeg_coefs *out);
void slimg_b200_jpeg_coefs_free(slimg_jpeg_coefs *c);

/* convert(image, {format: Jpeg, ...}) up to the entropy coder (pipeline.rs:66-95 -> codec/jpeg.rs:51-72): the pixel
 * stages and the JPEG front end run back to back on the device -- the resized image never leaves HBM, only the
 * coefficient planes come back.  out->geom.width / height are the dimensions PipelineResult reports
 * (pipeline.rs:97-102).  Feed the planes to the reference encoder with jpeg_write_coefficients (jctrans.c:43), see
 * INTEGRATION.md. */
slimg_status slimg_b200_pipeline_jpeg(const slimg_image *in, const slimg_pipeline_options *opts, int order,
                                      const uint16_t qt_luma[64], const uint16_t qt_chroma[64], uint32_t flags,
                                      slimg_jpeg_coefs *out);

/* ---- colour planes for the other lossy encoders: what webp::Encoder::from_rgba (codec/webp.rs:25-30 ->
 * libwebp WebPPictureImportRGBA) and ravif::Encoder::encode_rgba (codec/avif.rs:28-47) compute from the RGBA buffer before
 * their own transforms.  libwebp and ravif are absent from the reference tree: restated from their published sources,
 * PARITY UNPINNED.  ravif's optional clean-up of the colour of fully transparent pixels is not built. ---- */
enum { SLIMG_PLANES_WEBP_YUV420 = 1,    /* u8: Y w x h; U, V ceil(w/2) x ceil(h/2) (BT.601 studio range, gamma-domain chroma averaging);
                                           A w x h only when some pixel is not opaque */
       SLIMG_PLANES_AVIF_YCBCR444_8 = 2,   /* u8:  Y, Cb, Cr, A w x h (BT.601 full range) */
       SLIMG_PLANES_AVIF_YCBCR444_10 = 3   /* u16: same, 10 bits */ };
typedef struct {
    uint32_t width, height, kind;
    uint32_t chroma_width, chroma_height;
    uint32_t bytes_per_sample;          /* 1 or 2 */
    int32_t  has_alpha;                 /* WebP: 0 = every pixel opaque, `a` is NULL */
    void *y, *cb, *cr, *a;              /* host, one allocation; release with slimg_b200_planes_free */
} slimg_planes;
/* pixel stages (opts may be all-None) and the plane conversion back to back on the device: the resized image never leaves HBM */
slimg_status slimg_b200_pipeline_planes(const slimg_image *in, const slimg_pipeline_options *opts, int order, int kind,
                                        slimg_planes *out);
void slimg_b200_planes_free(slimg_planes *p);
/* device-resident batch: planes packed per image; for WebP `a` may be NULL (alpha ignored: treated as opaque), else a scratch
 * of n * ceil(h/2) bytes + n u32 flags is taken from `scratch` (n * ceil(h/2) rounded up to 4, + 4 n bytes; the flags -- 1 for
 * an image that holds a pixel that is not opaque -- are the last 4 n bytes and are zeroed by the call) */
slimg_status slimg_b200_color_planes_batch_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h, int kind,
                                                  void *y, void *cb, void *cr, void *a, void *scratch, void *stream);
/* The WebP case of the call above for callers that must not allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The row-flag scratch the conversion needs lives in a static __device__ array inside the library
 * (SLIMG_WEBP_STATIC_SCRATCH_BYTES, held by every GPU that loads the library until the process exits); the call
 * performs no cudaMalloc / cudaFree and calls no allocating routine: it enqueues a memset, two kernels and one
 * device-to-device copy on `stream` and returns.  Same arguments and same planes as the call above with
 * kind = SLIMG_PLANES_WEBP_YUV420 and a non-NULL `a`; `any_alpha` (device, n u32, may be NULL) receives 1 for every image
 * that holds a pixel that is not opaque, else 0.  Limit: n * ceil(h/2) rounded up to 4, + 4 n bytes, must fit the static
 * scratch (e.g. 60 images 8640 rows high, or 1 image of up to 524 000 rows); beyond it the call fails with
 * SLIMG_ERR_IMAGE and a message naming the limit.  Every call on a device shares the one array: calls that overlap on
 * the GPU (different streams, or graphs replayed concurrently) race on it and must be serialised by the caller; calls
 * on one stream are safe, and each call clears the flags the previous one left. */
#define SLIMG_WEBP_STATIC_SCRATCH_BYTES 262144
slimg_status slimg_b200_webp_planes_batch_device_static(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                                        void *y, void *u, void *v, void *a, uint32_t *any_alpha,
                                                        void *stream);

/* ---- batched, device-resident variants (what bench.py's `value` times) ----
 * `src` / `dst` are DEVICE pointers on the current CUDA device; images are packed back to back
 * (n * w*h*4 bytes in, n * out_w*out_h*4 bytes out).  `stream` is a cudaStream_t (NULL = default
 * stream).  Calls only enqueue work; they do not synchronise. */
slimg_status slimg_b200_pipeline_batch_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                              const slimg_pipeline_options *opts, int order,
                                              uint8_t *dst, size_t dst_capacity, void *stream);
/* coefficient planes packed per image: y[n][y_HB][y_WB][64], cb[...], cr[...] (device) */
slimg_status slimg_b200_jpeg_frontend_batch_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                                   const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                                                   int16_t *y, int16_t *cb, int16_t *cr, void *stream);
/* same, with the un-quantised planes as well (all three or none) */
slimg_status slimg_b200_jpeg_frontend_batch_device_ex(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                                      const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                                                      int16_t *y, int16_t *cb, int16_t *cr,
                                                      int16_t *y_uq, int16_t *cb_uq, int16_t *cr_uq, uint32_t flags, void *stream);
/* pixel pipeline + JPEG front end on device-resident batches; `pixels` is device scratch for the n final images
 * (n * out_w*out_h*4 bytes, the pipeline's output, which the front end reads back from HBM / L2) */
slimg_status slimg_b200_pipeline_jpeg_batch_device(const uint8_t *src, uint32_t n, uint32_t w, uint32_t h,
                                                   const slimg_pipeline_options *opts, int order,
                                                   const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                                                   uint8_t *pixels, size_t pixels_capacity,
                                                   int16_t *y, int16_t *cb, int16_t *cr,
                                                   int16_t *y_uq, int16_t *cb_uq, int16_t *cr_uq, uint32_t flags, void *stream);

/* ---- batched, host buffers: the `--jobs` scheduler (cli/src/commands/mod.rs:64-72 and the
 * par_iter sites crop.rs:112, extend.rs:114, convert.rs:60).  Images are independent (and may differ in size: a
 * directory of photos); they are dealt to `n_devices` GPUs x `streams_per_device` in-flight slots by a pull-based
 * queue.  Host buffers may be pinned (slimg_b200_host_alloc: copied straight over PCIe) or ordinary pageable memory
 * (a Rust Vec, Python bytes, a RustBuffer: staged through the library's pinned ring in chunks, the staging memcpy
 * overlapping the DMA of the previous chunk).  Per-image failures are isolated like ErrorCollector
 * (mod.rs:113-145): statuses[i] carries image i's status and the call returns the number of failed images.
 *
 * out[] is IN/OUT: an entry whose `data` is non-NULL with `len` >= the image's output size is a caller-owned
 * destination and is written in place; otherwise (data == NULL, or too small) the library allocates the buffer
 * (release with slimg_b200_image_free).  Zero-initialise entries you do not own. ---- */
typedef struct {
    int32_t n_devices;           /* 0 = all visible devices */
    int32_t streams_per_device;  /* 0 = default (3) */
    const int32_t *device_ids;   /* NULL = 0..n_devices-1 */
} slimg_batch_config;

int32_t slimg_b200_pipeline_batch_host(const slimg_image *in, uint32_t n,
                                       const slimg_pipeline_options *opts, int order,
                                       const slimg_batch_config *cfg,
                                       slimg_image *out, slimg_status *statuses);
/* the same scheduler for convert-to-JPEG: out[i] receives library-allocated coefficient planes
 * (slimg_b200_jpeg_coefs_free each) */
int32_t slimg_b200_pipeline_jpeg_batch_host(const slimg_image *in, uint32_t n,
                                            const slimg_pipeline_options *opts, int order,
                                            const uint16_t qt_luma[64], const uint16_t qt_chroma[64], uint32_t flags,
                                            const slimg_batch_config *cfg,
                                            slimg_jpeg_coefs *out, slimg_status *statuses);

/* ---- misc ---- */
const char *slimg_b200_last_error(void);          /* thread-local, never NULL */
int32_t     slimg_b200_abi_version(void);
int32_t     slimg_b200_device_count(void);        /* 0 when no usable CUDA device */
uint64_t    slimg_b200_kernel_launches(void);     /* kernels this library has launched, process-wide */
/* which kernel served the calling thread's last pixel-pipeline launch (tests and benches assert on it) */
enum { SLIMG_PATH_NONE = 0, SLIMG_PATH_REMAP = 1,     /* K1: crop / extend only */
       SLIMG_PATH_GENERIC = 2, SLIMG_PATH_FEWTAPS = 3, /* K2 f32 kernels (k2_resize.cu) */
       SLIMG_PATH_HFIRST = 4,                          /* K2 H-first (k2_fast.cu): downscales beyond 9x */
       SLIMG_PATH_VFIRST = 5,                          /* K2 integer tensor pipeline, mma.sync (k2_vfirst.cu) */
       SLIMG_PATH_TC = 6,                              /* K2 integer tensor pipeline, tcgen05 (k2_tc.cu) */
       SLIMG_PATH_TILE = 7 };                          /* K2 integer tile kernel: ratios below 3x and upscales (k2_tile.cu) */
int32_t     slimg_b200_last_resize_path(void);
uint64_t    slimg_b200_path_launches(int32_t path);  /* launches per path, process-wide */
/* pinned host allocations for callers that want zero-copy staging (e2e bench) */
void       *slimg_b200_host_alloc(size_t bytes);
void        slimg_b200_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* SLIMG_B200_H */
