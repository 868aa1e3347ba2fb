/*
 * ljt_shim.c -- drives the REAL libjpeg-turbo 2.1.5.1 that the reference vendors
 * (crates/libjxl-sys/libjxl/third_party/libjpeg-turbo/), compiled from those sources where they
 * lie by oracle/ref_ljt/Makefile into oracle/_ref/libljt_ref.so.  TEST INFRASTRUCTURE ONLY.
 *
 * It answers one question: "which quantised DCT coefficients does the reference's JPEG encoder
 * family hand to its entropy coder for this RGBA image and these quantisation tables?"
 * Procedure: compress through the public API (JCS_EXT_RGBA input, YCbCr 4:2:0,
 * JDCT_ISLOW, optimize_coding so the full-image coefficient controller -- the one MozJPEG's
 * progressive mode uses, jccoefct.c:245 compress_first_pass -- runs), then read the
 * coefficient arrays back with jpeg_read_coefficients (jdtrans.c:49), dummy blocks included.
 */
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "jpeglib.h"

static void put_table(j_compress_ptr cinfo, int which, const uint16_t qt[64])
{
    unsigned int basic[64];
    for (int i = 0; i < 64; i++) basic[i] = qt[i];
    jpeg_add_quant_table(cinfo, which, basic, 100, FALSE);
}

/* Standard tables for a quality setting, exactly as jpeg_set_quality(q, force_baseline=FALSE). */
int ljt_quant_table(int quality, int which, uint16_t out[64])
{
    struct jpeg_compress_struct cinfo;
    struct jpeg_error_mgr jerr;
    cinfo.err = jpeg_std_error(&jerr);
    jpeg_create_compress(&cinfo);
    cinfo.input_components = 3;
    cinfo.in_color_space = JCS_RGB;
    jpeg_set_defaults(&cinfo);
    jpeg_set_quality(&cinfo, quality, FALSE);
    for (int i = 0; i < 64; i++) out[i] = cinfo.quant_tbl_ptrs[which]->quantval[i];
    jpeg_destroy_compress(&cinfo);
    return 0;
}

/* Encodes and returns the JPEG byte stream (caller frees with ljt_free). */
int ljt_encode(const uint8_t *rgba, uint32_t w, uint32_t h,
               const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
               unsigned char **out, unsigned long *out_size)
{
    struct jpeg_compress_struct cinfo;
    struct jpeg_error_mgr jerr;
    cinfo.err = jpeg_std_error(&jerr);
    jpeg_create_compress(&cinfo);
    *out = NULL; *out_size = 0;
    jpeg_mem_dest(&cinfo, out, out_size);
    cinfo.image_width = w;
    cinfo.image_height = h;
    cinfo.input_components = 4;
    cinfo.in_color_space = JCS_EXT_RGBA;
    jpeg_set_defaults(&cinfo);               /* YCbCr, 2x2 / 1x1 / 1x1 */
    put_table(&cinfo, 0, qt_luma);
    put_table(&cinfo, 1, qt_chroma);
    cinfo.dct_method = JDCT_ISLOW;
    cinfo.optimize_coding = TRUE;
    jpeg_start_compress(&cinfo, TRUE);
    while (cinfo.next_scanline < cinfo.image_height) {
        JSAMPROW row = (JSAMPROW)(rgba + (size_t)cinfo.next_scanline * w * 4);
        jpeg_write_scanlines(&cinfo, &row, 1);
    }
    jpeg_finish_compress(&cinfo);
    jpeg_destroy_compress(&cinfo);
    return 0;
}

void ljt_free(void *p) { free(p); }

/* Coefficient planes laid out [HB][WB][64] int16 with MCU-padded block counts
 * (Y: round_up(ceil(w/8),2) x round_up(ceil(h/8),2); Cb/Cr: ceil(w/16) x ceil(h/16)). */
int ljt_frontend(const uint8_t *rgba, uint32_t w, uint32_t h,
                 const uint16_t qt_luma[64], const uint16_t qt_chroma[64],
                 int16_t *y, int16_t *cb, int16_t *cr)
{
    unsigned char *buf = NULL;
    unsigned long size = 0;
    ljt_encode(rgba, w, h, qt_luma, qt_chroma, &buf, &size);

    struct jpeg_decompress_struct d;
    struct jpeg_error_mgr jerr;
    d.err = jpeg_std_error(&jerr);
    jpeg_create_decompress(&d);
    jpeg_mem_src(&d, buf, size);
    jpeg_read_header(&d, TRUE);
    jvirt_barray_ptr *arrays = jpeg_read_coefficients(&d);
    int16_t *dst[3] = { y, cb, cr };
    for (int ci = 0; ci < 3; ci++) {
        jpeg_component_info *c = &d.comp_info[ci];
        unsigned WB = (c->width_in_blocks + c->h_samp_factor - 1) / c->h_samp_factor * c->h_samp_factor;
        unsigned HB = (c->height_in_blocks + c->v_samp_factor - 1) / c->v_samp_factor * c->v_samp_factor;
        for (unsigned by = 0; by < HB; by++) {
            JBLOCKARRAY rows = (*d.mem->access_virt_barray)((j_common_ptr)&d, arrays[ci], by, 1, FALSE);
            for (unsigned bx = 0; bx < WB; bx++)
                for (int k = 0; k < 64; k++)
                    dst[ci][((size_t)by * WB + bx) * 64 + k] = rows[0][bx][k];
        }
    }
    jpeg_finish_decompress(&d);
    jpeg_destroy_decompress(&d);
    free(buf);
    return 0;
}

/* The hand-off SURVEY.md 8(f1) names: quantised coefficients computed elsewhere (the CUDA front end) go straight
 * to the reference family's entropy coder with jpeg_write_coefficients (jctrans.c:43) -- colour conversion,
 * downsampling, DCT and quantisation are skipped, Huffman optimisation and (optionally) the progressive scan
 * script run as usual.  Planes are laid out as ljt_frontend returns them.  With progressive == 0 the byte stream
 * must equal ljt_encode's for the same image and tables. */
int ljt_write_coefficients(const int16_t *y, const int16_t *cb, const int16_t *cr, uint32_t w, uint32_t h,
                           const uint16_t qt_luma[64], const uint16_t qt_chroma[64], int progressive,
                           unsigned char **out, unsigned long *out_size)
{
    struct jpeg_compress_struct cinfo;
    struct jpeg_error_mgr jerr;
    cinfo.err = jpeg_std_error(&jerr);
    jpeg_create_compress(&cinfo);
    *out = NULL; *out_size = 0;
    jpeg_mem_dest(&cinfo, out, out_size);
    cinfo.image_width = w;
    cinfo.image_height = h;
    cinfo.input_components = 3;
    cinfo.in_color_space = JCS_RGB;
    jpeg_set_defaults(&cinfo);               /* YCbCr, 2x2 / 1x1 / 1x1 (jcparam.c:365-373) */
    put_table(&cinfo, 0, qt_luma);
    put_table(&cinfo, 1, qt_chroma);
    cinfo.optimize_coding = TRUE;
    if (progressive) jpeg_simple_progression(&cinfo);

    /* MCU-padded block counts, as jcmaster.c:137-171 / jccoefct.c size the full-image buffers */
    const unsigned cw = (w + 15) / 16, chh = (h + 15) / 16;
    const unsigned WB[3] = { cw * 2, cw, cw }, HB[3] = { chh * 2, chh, chh };
    const int vs[3] = { 2, 1, 1 };
    jvirt_barray_ptr arrays[3];
    for (int ci = 0; ci < 3; ci++)
        arrays[ci] = (*cinfo.mem->request_virt_barray)((j_common_ptr)&cinfo, JPOOL_IMAGE, FALSE, WB[ci], HB[ci], vs[ci]);
    jpeg_write_coefficients(&cinfo, arrays);  /* realises the arrays; they are read at jpeg_finish_compress */
    const int16_t *src[3] = { y, cb, cr };
    for (int ci = 0; ci < 3; ci++)
        for (unsigned by = 0; by < HB[ci]; by++) {
            JBLOCKARRAY rows = (*cinfo.mem->access_virt_barray)((j_common_ptr)&cinfo, arrays[ci], by, 1, TRUE);
            for (unsigned bx = 0; bx < WB[ci]; bx++)
                for (int k = 0; k < 64; k++)
                    rows[0][bx][k] = src[ci][((size_t)by * WB[ci] + bx) * 64 + k];
        }
    jpeg_finish_compress(&cinfo);
    jpeg_destroy_compress(&cinfo);
    return 0;
}

/* Decodes a JPEG byte stream to RGBA (alpha 255) with the same library: used to check that a stream written
 * from hand-off coefficients is a valid image of the right size.  out must hold w*h*4 bytes. */
int ljt_decode_rgba(const unsigned char *jpg, unsigned long size, uint32_t *w, uint32_t *h, uint8_t *out, size_t out_cap)
{
    struct jpeg_decompress_struct d;
    struct jpeg_error_mgr jerr;
    d.err = jpeg_std_error(&jerr);
    jpeg_create_decompress(&d);
    jpeg_mem_src(&d, jpg, size);
    jpeg_read_header(&d, TRUE);
    d.out_color_space = JCS_EXT_RGBA;
    jpeg_start_decompress(&d);
    *w = d.output_width; *h = d.output_height;
    if ((size_t)d.output_width * d.output_height * 4 > out_cap) { jpeg_destroy_decompress(&d); return 1; }
    while (d.output_scanline < d.output_height) {
        JSAMPROW row = (JSAMPROW)(out + (size_t)d.output_scanline * d.output_width * 4);
        jpeg_read_scanlines(&d, &row, 1);
    }
    jpeg_finish_decompress(&d);
    jpeg_destroy_decompress(&d);
    return 0;
}
