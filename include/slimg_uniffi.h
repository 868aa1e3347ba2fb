/*
 * slimg_uniffi.h -- the UniFFI scaffolding symbols of libslimg_ffi, exported by libslimg_b200.so.
 *
 * The reference's Kotlin and Python packages do not link slimg-core: they dlopen `libslimg_ffi` and call the
 * symbols UniFFI generates from crates/slimg-ffi/src/lib.rs (listed in
 * bindings/kotlin/src/main/kotlin/io/clroot/slimg/slimg_ffi.kt:662-706).  Exporting the same symbols with the
 * same calling convention and wire format lets those UNMODIFIED bindings run the pixel functions on the B200:
 * point `uniffi.component.slimg_ffi.libraryOverride` (JNA, slimg_ffi.kt findLibraryName) or the Python
 * loader's library path at libslimg_b200.so.
 *
 * Calling convention (slimg_ffi.kt:41-46, 183-190): every argument that is not a primitive travels in a
 * RustBuffer passed BY VALUE; the callee consumes (frees) its argument buffers; the result comes back in a
 * callee-allocated RustBuffer the caller releases with ffi_slimg_ffi_rustbuffer_free; errors are reported through
 * the RustCallStatus out-parameter (code 1 + a lowered SlimgError).  Wire format: big-endian, see
 * slimg_ffi.kt:1099-1112 (ByteArray), 1186-1207 (ImageData), 1256-1285 (PipelineOptions), 1382-1397 (CropMode),
 * 1488-1501 (ExtendMode), 1583-1596 (FillColor), 1642-1673 (Format), 1759-1781 (ResizeMode), 1961-1997 (SlimgError).
 *
 * What runs where: resize / crop / extend and the pixel stages of convert run the CUDA kernels (no CPU
 * fallback).  Encoding and decoding stay on the reference codecs: convert / optimize / decode call
 * the codec hooks below and fail with SlimgError::Encode / ::Decode when none is installed.
 */
#ifndef SLIMG_UNIFFI_H
#define SLIMG_UNIFFI_H

#include <stddef.h>
#include <stdint.h>

#include "slimg_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* slimg_ffi.kt:41-46 */
typedef struct { uint64_t capacity; uint64_t len; uint8_t *data; } RustBuffer;
/* slimg_ffi.kt ForeignBytes: a borrowed byte slice of the foreign side */
typedef struct { int32_t len; const uint8_t *data; } ForeignBytes;
/* slimg_ffi.kt:183-190: code 0 = success, 1 = error (error_buf holds the lowered SlimgError), 2 = panic (a string) */
typedef struct { int8_t code; RustBuffer error_buf; } RustCallStatus;

/* ---- buffer management (slimg_ffi.kt:699-706) ---- */
RustBuffer ffi_slimg_ffi_rustbuffer_alloc(uint64_t size, RustCallStatus *status);
RustBuffer ffi_slimg_ffi_rustbuffer_from_bytes(ForeignBytes bytes, RustCallStatus *status);
void       ffi_slimg_ffi_rustbuffer_free(RustBuffer buf, RustCallStatus *status);
RustBuffer ffi_slimg_ffi_rustbuffer_reserve(RustBuffer buf, uint64_t additional, RustCallStatus *status);

/* ---- contract version and API checksums (slimg_ffi.kt:662, 809-852) ---- */
uint32_t ffi_slimg_ffi_uniffi_contract_version(void);      /* 30 */
uint16_t uniffi_slimg_ffi_checksum_func_convert(void);
uint16_t uniffi_slimg_ffi_checksum_func_crop(void);
uint16_t uniffi_slimg_ffi_checksum_func_decode(void);
uint16_t uniffi_slimg_ffi_checksum_func_decode_file(void);
uint16_t uniffi_slimg_ffi_checksum_func_extend(void);
uint16_t uniffi_slimg_ffi_checksum_func_format_can_encode(void);
uint16_t uniffi_slimg_ffi_checksum_func_format_extension(void);
uint16_t uniffi_slimg_ffi_checksum_func_format_from_extension(void);
uint16_t uniffi_slimg_ffi_checksum_func_format_from_magic_bytes(void);
uint16_t uniffi_slimg_ffi_checksum_func_optimize(void);
uint16_t uniffi_slimg_ffi_checksum_func_output_path(void);
uint16_t uniffi_slimg_ffi_checksum_func_resize(void);

/* ---- exported functions (crates/slimg-ffi/src/lib.rs:254-362; slimg_ffi.kt:675-698) ---- */
/* pixel functions: CUDA */
RustBuffer uniffi_slimg_ffi_fn_func_resize(RustBuffer image, RustBuffer mode, RustCallStatus *status);
RustBuffer uniffi_slimg_ffi_fn_func_crop(RustBuffer image, RustBuffer mode, RustCallStatus *status);
RustBuffer uniffi_slimg_ffi_fn_func_extend(RustBuffer image, RustBuffer mode, RustBuffer fill, RustCallStatus *status);
/* convert = crop -> extend -> resize on the GPU (one launch), then the encoder hook */
RustBuffer uniffi_slimg_ffi_fn_func_convert(RustBuffer image, RustBuffer options, RustCallStatus *status);
/* codec functions: the hooks */
RustBuffer uniffi_slimg_ffi_fn_func_decode(RustBuffer data, RustCallStatus *status);
RustBuffer uniffi_slimg_ffi_fn_func_decode_file(RustBuffer path, RustCallStatus *status);
RustBuffer uniffi_slimg_ffi_fn_func_optimize(RustBuffer data, uint8_t quality, RustCallStatus *status);
/* host-only helpers (crates/slimg-core/src/format.rs:14-75, pipeline.rs:130-141) */
RustBuffer uniffi_slimg_ffi_fn_func_format_extension(RustBuffer format, RustCallStatus *status);
int8_t     uniffi_slimg_ffi_fn_func_format_can_encode(RustBuffer format, RustCallStatus *status);
RustBuffer uniffi_slimg_ffi_fn_func_format_from_extension(RustBuffer path, RustCallStatus *status);
RustBuffer uniffi_slimg_ffi_fn_func_format_from_magic_bytes(RustBuffer data, RustCallStatus *status);
RustBuffer uniffi_slimg_ffi_fn_func_output_path(RustBuffer input, RustBuffer format, RustBuffer output, RustCallStatus *status);

/* ---- codec hooks: where the reference's entropy coders / decoders plug in ----
 * format: 1..6 = Jpeg, Png, WebP, Avif, Jxl, Qoi (slimg_ffi.kt:1642-1673).  Both return a slimg_status; on
 * failure they may set *message to a static or malloc'ed-and-leaked string.  Output buffers are malloc'ed by the
 * hook and freed by the library with free(). */
typedef int32_t (*slimg_encode_hook)(void *user, const slimg_image *pixels, int32_t format, uint8_t quality,
                                     uint8_t **out, size_t *out_len, const char **message);
typedef int32_t (*slimg_decode_hook)(void *user, const uint8_t *data, size_t len, slimg_image *pixels,
                                     int32_t *format, const char **message);
void slimg_b200_set_codec_hooks(slimg_encode_hook enc, slimg_decode_hook dec, void *user);

/* Format::Jpeg with the front end on the GPU: when this hook is installed, convert(.., Jpeg) runs
 * crop -> extend -> resize -> RGB->YCbCr 4:2:0 -> FDCT -> quantisation on the device (slimg_b200_pipeline_jpeg), no
 * pixel ever returns to the host, and the hook -- the reference library's ENTROPY coder -- receives the coefficient
 * planes: with libjpeg-turbo / MozJPEG that is jpeg_write_coefficients (jctrans.c:43) in place of
 * write_scanlines (crates/slimg-core/src/codec/jpeg.rs:66-76).  libslimg_jpeg_handoff.so
 * (slimg_b200/jpeg_handoff/) provides slimg_jpeg_handoff_hook with this signature.
 * table_base: SLIMG_QT_MOZJPEG (the reference encoder's tables) or SLIMG_QT_ANNEX_K; flags: SLIMG_JPEG_UNQUANTISED
 * to also hand over the un-quantised planes (MozJPEG's trellis input).  Output buffer: malloc'ed by the hook. */
typedef int32_t (*slimg_jpeg_coef_hook)(void *user, const slimg_jpeg_coefs *coefs, uint8_t quality,
                                        uint8_t **out, size_t *out_len, const char **message);
void slimg_b200_set_jpeg_coefficient_hook(slimg_jpeg_coef_hook hook, void *user, int32_t table_base, uint32_t flags);

#ifdef __cplusplus
}
#endif
#endif
