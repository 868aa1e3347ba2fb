/*
 * jpeg_handoff.c -- libslimg_jpeg_handoff.so: hands the coefficient planes of libslimg_b200's JPEG front end to a
 * libjpeg-family ENTROPY CODER through the public libjpeg API (jpeglib.h), so that convert(.., Format::Jpeg) never
 * runs colour conversion, downsampling, DCT or quantisation on the host.
 *
 * Replaces the second half of JpegCodec::encode (crates/slimg-core/src/codec/jpeg.rs:58-76: Compress::new,
 * set_quality, set_progressive_mode, optimize_coding, write_scanlines, finish): instead of write_scanlines the
 * encoder object gets jpeg_write_coefficients (jctrans.c:43) with the planes the GPU produced; Huffman
 * optimisation and the progressive scan script run as usual.  Entropy coding stays on the reference's library --
 * this file contains no codec arithmetic.  Build: slimg_b200/jpeg_handoff/Makefile (against the libjpeg-turbo the
 * reference vendors, or any libjpeg >= 6b / MozJPEG given JPEG_INC / JPEG_LIB).
 *
 * MozJPEG note: jpeg_write_coefficients bypasses trellis quantisation.  A MozJPEG build that wants trellis takes
 * the un-quantised planes (SLIMG_JPEG_UNQUANTISED: slimg_jpeg_coefs.y_uq ...) -- INTEGRATION.md section 4.
 */
#include <setjmp.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "jpeglib.h"

#include "../../include/slimg_b200.h"

#define SLIMG_HANDOFF_PROGRESSIVE 1u     /* jpeg_simple_progression (codec/jpeg.rs:61 set_progressive_mode) */
#define SLIMG_HANDOFF_OPTIMIZE    2u     /* optimize_coding (codec/jpeg.rs:63) */

struct handoff_err {
    struct jpeg_error_mgr pub;
    jmp_buf jump;
    char msg[JMSG_LENGTH_MAX];
};

static void on_error(j_common_ptr cinfo)      /* libjpeg's default error_exit calls exit(): never in a library */
{
    struct handoff_err *e = (struct handoff_err *)cinfo->err;
    (*cinfo->err->format_message)(cinfo, e->msg);
    longjmp(e->jump, 1);
}

static __thread char t_msg[JMSG_LENGTH_MAX];

/* Writes a JPEG stream from the planes.  *out is malloc'ed (release with free / slimg_jpeg_handoff_free).
 * Returns a slimg_status: 0, or SLIMG_ERR_ENCODE with *message set. */
int32_t slimg_jpeg_handoff_write(const slimg_jpeg_coefs *c, uint32_t flags, uint8_t **out, size_t *out_len,
                                 const char **message)
{
    struct jpeg_compress_struct cinfo;
    struct handoff_err err;
    unsigned char *buf = NULL;
    unsigned long size = 0;

    if (!c || !c->y || !c->cb || !c->cr || !out || !out_len) {
        if (message) *message = "null argument";
        return SLIMG_ERR_ENCODE;
    }
    cinfo.err = jpeg_std_error(&err.pub);
    err.pub.error_exit = on_error;
    if (setjmp(err.jump)) {
        jpeg_destroy_compress(&cinfo);
        free(buf);
        snprintf(t_msg, sizeof t_msg, "libjpeg: %s", err.msg);
        if (message) *message = t_msg;
        return SLIMG_ERR_ENCODE;
    }
    jpeg_create_compress(&cinfo);
    jpeg_mem_dest(&cinfo, &buf, &size);
    cinfo.image_width = c->geom.width;
    cinfo.image_height = c->geom.height;
    cinfo.input_components = 3;
    cinfo.in_color_space = JCS_RGB;
    jpeg_set_defaults(&cinfo);               /* YCbCr, 2x2 / 1x1 / 1x1 sampling (jcparam.c:365-373) */
    {
        unsigned int basic[64];
        int i;
        for (i = 0; i < 64; i++) basic[i] = c->qt_luma[i];
        jpeg_add_quant_table(&cinfo, 0, basic, 100, FALSE);
        for (i = 0; i < 64; i++) basic[i] = c->qt_chroma[i];
        jpeg_add_quant_table(&cinfo, 1, basic, 100, FALSE);
    }
    cinfo.optimize_coding = (flags & SLIMG_HANDOFF_OPTIMIZE) ? TRUE : FALSE;
    if (flags & SLIMG_HANDOFF_PROGRESSIVE) jpeg_simple_progression(&cinfo);

    {
        /* the planes are already laid out as the full-image coefficient controller keeps them: MCU-padded block
         * counts, dummy blocks included (jcmaster.c:137-171, jccoefct.c:286-319) */
        const unsigned WB[3] = { c->geom.y_WB, c->geom.c_WB, c->geom.c_WB };
        const unsigned HB[3] = { c->geom.y_HB, c->geom.c_HB, c->geom.c_HB };
        const int vs[3] = { 2, 1, 1 };
        const int16_t *src[3] = { c->y, c->cb, c->cr };
        jvirt_barray_ptr arrays[3];
        int ci;
        unsigned by;
        for (ci = 0; ci < 3; ci++)
            arrays[ci] = (*cinfo.mem->request_virt_barray)((j_common_ptr)&cinfo, JPOOL_IMAGE, FALSE, WB[ci], HB[ci], vs[ci]);
        jpeg_write_coefficients(&cinfo, arrays);      /* realises the arrays; they are read at jpeg_finish_compress */
        for (ci = 0; ci < 3; ci++)
            for (by = 0; by < HB[ci]; by++) {
                JBLOCKARRAY rows = (*cinfo.mem->access_virt_barray)((j_common_ptr)&cinfo, arrays[ci], by, 1, TRUE);
                /* JCOEF is short: a block row is WB * 64 contiguous int16, exactly our plane row */
                memcpy(rows[0][0], src[ci] + (size_t)by * WB[ci] * 64, (size_t)WB[ci] * 64 * sizeof(JCOEF));
            }
    }
    jpeg_finish_compress(&cinfo);
    jpeg_destroy_compress(&cinfo);
    *out = buf;
    *out_len = (size_t)size;
    return SLIMG_OK;
}

/* slimg_jpeg_coef_hook (include/slimg_uniffi.h): `user` carries the SLIMG_HANDOFF_* flags; NULL means the
 * reference's settings (progressive + optimised Huffman tables, codec/jpeg.rs:61-63). */
int32_t slimg_jpeg_handoff_hook(void *user, const slimg_jpeg_coefs *coefs, uint8_t quality, uint8_t **out,
                                size_t *out_len, const char **message)
{
    const uint32_t flags = user ? (uint32_t)(uintptr_t)user - 1u : (SLIMG_HANDOFF_PROGRESSIVE | SLIMG_HANDOFF_OPTIMIZE);
    (void)quality;      /* already folded into the quantisation tables the planes were computed with */
    return slimg_jpeg_handoff_write(coefs, flags, out, out_len, message);
}

void slimg_jpeg_handoff_free(void *p) { free(p); }

/* The decode side of the same library (pipeline.rs:50-63 decode for Format::Jpeg; codec/jpeg.rs:20-49): JPEG bytes
 * to RGBA8 (alpha 255), usable as the slimg_decode_hook of include/slimg_uniffi.h.  pixels->data is malloc'ed. */
int32_t slimg_jpeg_handoff_decode(void *user, const uint8_t *data, size_t len, slimg_image *pixels, int32_t *format,
                                  const char **message)
{
    struct jpeg_decompress_struct d;
    struct handoff_err err;
    uint8_t *buf = NULL;
    (void)user;
    if (!data || !pixels) { if (message) *message = "null argument"; return SLIMG_ERR_DECODE; }
    if (len < 3 || data[0] != 0xff || data[1] != 0xd8) { if (message) *message = "not a JPEG stream"; return SLIMG_ERR_UNKNOWN_FORMAT; }
    d.err = jpeg_std_error(&err.pub);
    err.pub.error_exit = on_error;
    if (setjmp(err.jump)) {
        jpeg_destroy_decompress(&d);
        free(buf);
        snprintf(t_msg, sizeof t_msg, "libjpeg: %s", err.msg);
        if (message) *message = t_msg;
        return SLIMG_ERR_DECODE;
    }
    jpeg_create_decompress(&d);
    jpeg_mem_src(&d, (unsigned char *)data, (unsigned long)len);
    jpeg_read_header(&d, TRUE);
    d.out_color_space = JCS_EXT_RGBA;
    jpeg_start_decompress(&d);
    buf = (uint8_t *)malloc((size_t)d.output_width * d.output_height * 4);
    if (!buf) { jpeg_destroy_decompress(&d); if (message) *message = "out of memory"; return SLIMG_ERR_DECODE; }
    while (d.output_scanline < d.output_height) {
        JSAMPROW row = (JSAMPROW)(buf + (size_t)d.output_scanline * d.output_width * 4);
        jpeg_read_scanlines(&d, &row, 1);
    }
    pixels->width = d.output_width; pixels->height = d.output_height;
    pixels->data = buf; pixels->len = (size_t)d.output_width * d.output_height * 4;
    if (format) *format = 1;      /* Format::Jpeg (slimg_ffi.kt:1642-1673) */
    jpeg_finish_decompress(&d);
    jpeg_destroy_decompress(&d);
    return SLIMG_OK;
}
