// kernels.cuh -- device-side parameter blocks and launch entry points of the sm_100a kernels.
#pragma once

#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>
#include <type_traits>

#include "../../include/slimg_b200.h"

namespace slimg {

// ---- K1: crop / extend as one remapped copy (k1_remap.cu) -----------------------------------
// dst pixel (x,y) of an fw x fh canvas takes
//   src(x + dx, y + dy)  if (x,y) is inside rect A = [ax0,ax1) x [ay0,ay1)   (source backed)
//   fill_in              if inside rect B = [bx0,bx1) x [by0,by1)            (B contains A)
//   fill_out             otherwise
// which is crop.rs:70-87 and extend.rs:94-133 (and their composition) with no intermediate.
struct RemapParams {
    const uint8_t *src;
    uint8_t *dst;
    size_t src_img_bytes, dst_img_bytes;
    uint32_t sw;           // source row pitch in pixels
    uint32_t fw, fh;
    int32_t ax0, ay0, ax1, ay1, dx, dy;
    int32_t bx0, by0, bx1, by1;
    uint32_t fill_in, fill_out;   // packed RGBA, little endian (R in the low byte)
};
cudaError_t launch_remap(const RemapParams &p, uint32_t n_images, cudaStream_t stream);

// ---- K2: Lanczos3 separable resize with fused pre / post remap (k2_resize.cu) ---------------
struct AxisDev {
    const int32_t *left;    // [out_len] first tap, in source-rect coordinates
    const int32_t *count;   // [out_len] taps that land on source pixels
    const float *w;         // [out_len][max_taps]
    const float *fillw;     // [out_len] total weight of the taps that land on fill
    int32_t max_taps;
};

struct ResizeParams {
    const uint8_t *src;
    uint8_t *dst;
    size_t src_img_bytes, dst_img_bytes;
    uint32_t sw;                       // source row pitch in pixels
    uint32_t rx, ry, rw, rh;           // source rectangle the taps index into
    AxisDev ax, ay;
    float fill_pre[4];                 // colour of the virtual pixels outside the rectangle
    uint32_t px, py, pw, ph;           // rectangle of the resize output R that is kept ...
    uint32_t fw, fh, fox, foy;         // ... and where it lands on the final fw x fh canvas
    uint32_t fill_post;                // packed RGBA of the rest of the canvas
    // tiling (generic kernel)
    int32_t tile_w, tile_h, span_max;
    int32_t few_k;                     // 2 / 4: aligned groups of that many canvas columns share a tap window of max_taps + few_k / 2 (upscales); else 1
    int32_t few_rows_max;              // > 0: both axes have <= 8 taps; most source rows a tile row-block reads
    const int32_t *tile_x0;            // [tiles_x] first source-rect column a tile column needs
    const int32_t *tile_span;          // [tiles_x] number of source-rect columns it needs
};
cudaError_t launch_resize_generic(const ResizeParams &p, uint32_t n_images, cudaStream_t stream);

// ---- K2 fast path: tensor-core horizontal pass + streaming vertical pass (k2_fast.cu) ----------
// Downscale (ratio >= 1) on both axes.  See k2_fast.cu for the algorithm.
struct VQuad {               // four consecutive source-rectangle rows of the vertical pass (336 bytes)
    uint32_t done[4];        // per row: bit s set = the output in slot s is complete after this row
    int16_t out[4][8];       // per row, per slot: local output row completing there (valid where the bit is set)
    float w2[4][16];         // per row: weight for the output held in slot s, duplicated (w,w) for f32x2 FMAs
};
struct FastResizeParams {
    const uint8_t *src;
    uint8_t *dst;
    size_t src_img_bytes, dst_img_bytes;
    uint32_t sw, sh;                 // source pitch (pixels) and height
    uint32_t ry, rh;                 // first row / row count of the source rectangle
    uint32_t pw, ph;                 // columns / rows of the resize output that are kept
    uint32_t n_tiles, n_strips;      // 8-column output tiles, 64-column strips
    const int32_t *tile_a0;          // [n_tiles] column of the tile's first k-chunk RELATIVE to its strip's staged span (multiple of 4)
    const int32_t *tile_nchunks;     // [n_tiles] 32-pixel k-chunks the tile's taps span
    const int32_t *strip_a0;         // [n_strips] absolute source column where the strip's staged span starts (multiple of 4)
    const int32_t *strip_bytes;      // [n_strips] bytes of each source row staged for the strip (multiple of 16)
    uint32_t max_chunks;             // largest tile_nchunks of the plan
    uint32_t pitch;                  // bytes between staged rows in shared memory (== 64 mod 128)
    uint32_t stages;                 // depth of the staging ring
    uint32_t smem_bytes;             // dynamic shared memory of the launch
    const int32_t *tile_boff;        // [n_tiles] first chunk of the tile inside bfrag
    const uint32_t *bfrag;           // [chunks][3 limbs][32 lanes][2] weight fragments (s8 x4 per word)
    const float *fillw_x;            // [n_tiles*8] weight that lands on fill, per kept output column
    float hscale;                    // 2^-S of the fixed-point weights
    const VQuad *vquad;              // [(rh padded to 16, +16) / 4]
    const int32_t *vleft, *vright;   // [ph] rectangle-row window [left,right) of each kept output row
    const float *fillw_y;            // [ph]
    const uint8_t *row_computed;     // [ph] 0 = every tap of the row is fill: the row is pure fill colour
    float fill_pre[4];
    uint32_t fw, fh, fox, foy;       // final canvas and where the kept rectangle lands on it
    uint32_t fill_post;              // packed RGBA outside the kept rectangle
    uint32_t fill_pre_packed;        // packed RGBA of kept rows / columns whose every tap is fill
    uint32_t n_bands, n_fill_blocks;
};
cudaError_t launch_resize_fast(const FastResizeParams &p, uint32_t n_images, cudaStream_t stream);

// ---- K2 V-first integer tensor pipeline (k2_vfirst.cu, plan in vplan.cpp) ---------------------------
struct VfParams {
    const uint8_t *src;
    uint8_t *dst;
    size_t src_img_bytes, dst_img_bytes;
    uint32_t sw, sh;                 // source pitch (pixels) and height
    uint32_t ry;                     // first row of the source rectangle
    uint32_t pw, ph;                 // columns / rows of the resize output that are kept
    int32_t ndig, nch1, nch2;        // weight digits; max chunks per group (pass 1) / tile (pass 2)
    // vertical pass: groups of 8 kept output rows
    // one record per non-empty group (vplan.cpp): 64 header words {group, band start, end of needed rows,
    // chunks, next band start, -, -, -, bias[8][4]} then [4][ndig][32 lanes][2] weight fragments (s8 x4 per word)
    uint32_t n_recs, rec_words;
    const uint32_t *recs;
    const uint8_t *row_computed;     // [ph]
    int32_t sh1;                     // right shift that turns the s32 sum into the u16 intermediate
    // horizontal pass: strips of <= 64 kept output columns, tiles of 8
    uint32_t n_strips;
    const int32_t *strip_col0;       // [n_strips+1] first kept column of each strip
    const int32_t *strip_src0;       // [n_strips] absolute source column where the staged span starts (multiple of 4)
    const int32_t *strip_nblk;       // [n_strips] 16-byte blocks of each staged row
    const int32_t *strip_bytes;      // [n_strips] bytes staged per source row
    const int32_t *strip_upw;        // [n_strips] consecutive units of `unit` blocks each pass-1 warp of the strip owns
    const int32_t *strip_nunits;     // [n_strips] units of the strip (the last warps may own fewer than upw, or none)
    int32_t unit;                    // 16-byte blocks per pass-1 unit (kernel template UNIT)
    uint32_t early_go;               // 1: the ring always holds the next group's rows, `go` is signalled a group ahead
    const int32_t *tile_kb0;         // [n_strips][8] first block of the tile's k range
    const int32_t *tile_nch;         // [n_strips][8] 32-pixel chunks of the tile's k range
    const uint32_t *b2;              // [n_strips][8][4][ndig][32 lanes][2]
    const float *cx;                 // [pw][4] constant term per kept column and channel
    const int32_t *cxi;              // [pw][4] the same + the rounding half, in units of 2^-sh2 (two-digit plans: integer epilogue)
    int32_t sh2;
    float s2[2][3];                  // scale of accumulator [limb][digit]
    // shared memory geometry
    uint32_t ring_rows, pitch, pitch_v, smem_bytes;
    uint32_t vbufs;                  // V (vertical-pass result) buffers: pass 1 may run this many groups ahead of pass 2
    uint32_t cohorts;                // pass-1 warps work in this many cohorts, cohort k on groups k, k + cohorts, ...
    // canvas
    uint32_t fw, fh, fox, foy;
    uint32_t fill_post, fill_pre_packed;
    uint32_t n_bands;
    uint32_t need_fill;              // bit 0: some canvas pixels are not computed; bit 1: some kept rows are pure fill
    uint32_t debug;                  // experiment switches (SLIMG_B200_VF_DEBUG), 0 in production
    uint32_t pf_steps;               // ring steps prefetched into L2 ahead of the ring (0 = off)
};
#ifndef SLIMG_VF_WARPS
#define SLIMG_VF_WARPS 16     /* measured on B200, C3: 16 warps x 4-block units 1300 Gpx/s; 20 x 3-block 1113; 24 x 3-block 1062 */
#endif
constexpr int kVfWarps = SLIMG_VF_WARPS;   // warps of a strip CTA: pass-2 tiles + pass-1 warps + signaller + producer
constexpr int kVfMaxVBufs = 4;      // most V buffers the kernel supports (VfParams::vbufs)
cudaError_t launch_resize_vfirst(const VfParams &p, uint32_t n_images, cudaStream_t stream);
size_t vfirst_smem_tail_bytes();

// ---- K2 integer tile kernel (k2_tile.cu, plan in tileplan.cpp): ratios below ~3.5x and upscales ---------------------
struct TileParams {
    const uint8_t *src;
    uint8_t *dst;
    size_t src_img_bytes, dst_img_bytes;
    uint32_t sw, sh;                 // source pitch (pixels) and height
    uint32_t ry, rh;                 // first row / rows of the source rectangle
    uint32_t pw, ph;                 // columns / rows of the resize output that are kept
    uint32_t tw, th;                 // kept columns / rows per CTA
    int32_t nch1, nch2, sh1;
    uint32_t rec_words;
    const uint32_t *recs;            // [groups][rec_words]
    const uint8_t *row_computed;     // [ph]
    uint32_t n_rowblk, n_colblk;
    const int32_t *rb_y0, *rb_rows;  // [n_rowblk]
    const int32_t *cb_src0, *cb_nblk;   // [n_colblk]
    const int32_t *tile_kb0;         // [n_colblk][tw / 8]
    const uint32_t *b2;              // [n_colblk][tw / 8][nch2][2][32 lanes][2]; packed: [n_colblk][tw / 8][3][32][2]
    const int32_t *cxi;              // [pw][4] constant term + rounding half, units of 2^-sh2
    int32_t sh2;                     // final right shift of the pass-2 integer sum
    uint32_t packed;                 // 1: both limbs of the intermediate share one K = 32 (upscales)
    uint32_t pitch, pitch_v, rows_max, smem_bytes;
    uint32_t fw, fh, fox, foy;
    uint32_t fill_post, fill_pre_packed;
    uint32_t need_fill;              // bit 0: some canvas pixels are not computed; bit 1: some kept rows are pure fill
};
cudaError_t launch_resize_tile(const TileParams &p, uint32_t n_images, cudaStream_t stream);

// ---- K2 tcgen05 pipeline (k2_tc.cu, plan in tcplan.cpp) ---------------------------------------------------
struct TcParams {
    const uint8_t *src;
    uint8_t *dst;
    size_t src_img_bytes, dst_img_bytes;
    uint32_t sw, sh;                 // source pitch (pixels) and height
    uint32_t ry;                     // first row of the source rectangle
    uint32_t pw, ph;                 // columns / rows of the resize output that are kept
    int32_t nch2, sh1;               // 32-pixel chunks per pass-2 tile; right shift to the u16 intermediate
    // vertical pass: blocks of 8 kept rows, chunks of 32 source-rectangle rows
    uint32_t n_blocks, n_chunks, crec_words;
    const uint32_t *crecs;           // [n_chunks][crec_words]: {first block, blocks, flags, chunk, ...12} + [blocks][512 B] tiles (N = 16: both digits)
    const int32_t *blk_c0, *blk_c1;  // [n_blocks] first / last chunk of a block
    const int32_t *bias;             // [n_blocks*8][4]
    // horizontal pass
    uint32_t n_strips, n_panels;
    const int32_t *strip_col0;       // [n_strips+1] first kept column of each strip
    const int32_t *strip_src0;       // [n_strips] first staged source pixel (multiple of 4)
    const int32_t *tile_kb0;         // [n_strips][8] first 4-pixel block of the tile's k range
    const uint32_t *b2;              // [n_strips][8][nch2][2][32 lanes][2]
    const float *cx;                 // [pw][4] constant term per kept column and channel
    float s2[2][3];                  // scale of accumulator [limb][digit]
    // shared memory / TMEM geometry
    uint32_t slots, slot_bytes, pitch_v, smem_bytes, nbuf, vbufs;
    // canvas
    uint32_t fw, fh, fox, foy;
    uint32_t fill_post, fill_pre_packed;
    uint32_t n_bands, need_fill, debug;
    uint32_t pf_chunks;              // chunks prefetched into L2 beyond the ring (0 = off)
    unsigned long long *stats;       // debug & 32: cycles each role spent waiting on each barrier (device, 16 counters)
};
cudaError_t launch_resize_tc(const TcParams &p, uint32_t n_images, cudaStream_t stream);

// ---- K3+K4: JPEG pre-entropy front end (k3_jpeg.cu) ------------------------------------------
struct QuantDev {                      // jcdctmgr.c:173-221 compute_reciprocal, per table entry
    uint32_t recip[2][64];
    uint32_t corr[2][64];
    uint32_t shift32[2][64];           // total right shift (r in compute_reciprocal) minus 32: always >= 2
};
struct JpegParams {
    const uint8_t *src;
    int16_t *y, *cb, *cr;
    int16_t *yu, *cbu, *cru;           // optional: the un-quantised coefficients (MozJPEG's trellis input), same layout; all NULL = off
    size_t src_img_bytes;
    size_t y_img_elems, c_img_elems;   // int16 elements per image in each plane
    uint32_t w, h;
    uint32_t y_wb, y_hb, y_WB, y_HB, c_wb, c_hb;
    QuantDev q;
    uint32_t debug;                    // experiment switches (SLIMG_B200_JPEG_DEBUG), 0 in production
    uint32_t deringing;                // 1: MozJPEG's overshoot deringing before the DCT + 10-bit clamp of the quantised values
    uint32_t dc_quant[2];              // DC quantiser of the luma / chroma table (deringing's overshoot cap)
};
cudaError_t launch_jpeg_frontend(const JpegParams &p, uint32_t n_images, cudaStream_t stream);

// ---- K4: colour planes for the WebP / AVIF encoders (k4_planes.cu) ---------------------------------------------------
struct WebpPlanesParams {
    const uint8_t *src;                // RGBA8, n images back to back
    size_t src_img_bytes, y_img_bytes, uv_img_bytes;
    uint32_t w, h;
    uint8_t *y, *u, *v, *a;            // y / a: w x h; u / v: ceil(w/2) x ceil(h/2); a may be NULL
    uint8_t *row_alpha;                // [n][ceil(h/2)] scratch: a row pair holds a pixel that is not opaque (NULL: treat all as opaque)
    uint32_t *any_alpha;               // [n] OR of row_alpha (must be zeroed by the caller)
    uint16_t g2l[256];                 // libwebp kGammaToLinearTab
    int32_t l2g[33];                   // libwebp kLinearToGammaTab
};
cudaError_t launch_webp_planes(const WebpPlanesParams &p, uint32_t n_images, cudaStream_t stream);
// the same conversion with its scratch in a file-scope __device__ array of k4_planes.cu: nothing is allocated
constexpr size_t kWebpStaticScratchBytes = 256 * 1024;      // == SLIMG_WEBP_STATIC_SCRATCH_BYTES (include/slimg_b200.h)
size_t webp_static_scratch_need(uint32_t n_images, uint32_t h);
cudaError_t launch_webp_planes_static(WebpPlanesParams p, uint32_t n_images, uint32_t *any_out, cudaStream_t stream);
struct AvifPlanesParams {
    const uint8_t *src;
    size_t src_img_bytes;
    uint32_t w, h, depth;              // depth 8: u8 planes; 10: u16 planes
    void *y, *cb, *cr, *a;             // w x h each, n images back to back
};
cudaError_t launch_avif_planes(const AvifPlanesParams &p, uint32_t n_images, cudaStream_t stream);

// process-wide count of kernels launched by this library (slimg_b200_kernel_launches)
void count_launch(uint64_t n = 1);
// which kernel served a launch (SLIMG_PATH_* of include/slimg_b200.h): thread-local last value + process-wide counters
void note_path(int path);

}  // namespace slimg
