// host.h -- host-side internals of libslimg_b200: geometry, filter-tap plans, launch glue.
// Not part of the ABI (that is include/slimg_b200.h).
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/slimg_b200.h"

namespace slimg {

// ---- errors ------------------------------------------------------------------------------
void set_error(const std::string &msg);
slimg_status fail(slimg_status st, const std::string &msg);
const std::string &last_error_string();

// ---- geometry (geometry.cpp) -------------------------------------------------------------
// Mirrors of crates/slimg-core/src/{resize,crop,extend}.rs region arithmetic.
slimg_status calculate_dimensions(uint32_t ow, uint32_t oh, const slimg_resize_mode &m, uint32_t &w, uint32_t &h);
slimg_status resize_output_dims(uint32_t w, uint32_t h, const slimg_resize_mode &m, uint32_t &ow, uint32_t &oh);
slimg_status calculate_crop_region(uint32_t iw, uint32_t ih, const slimg_crop_mode &m,
                                   uint32_t &x, uint32_t &y, uint32_t &w, uint32_t &h);
slimg_status calculate_extend_region(uint32_t iw, uint32_t ih, const slimg_extend_mode &m,
                                     uint32_t &cw, uint32_t &ch, uint32_t &ox, uint32_t &oy);

// ---- pipeline geometry -------------------------------------------------------------------
// Every pipeline this library runs is  S --(pre remap)--> V --(resize)--> R --(post remap)--> F :
//   S  source image                         (sw x sh)
//   V  virtual resize input: the rectangle (rx,ry,rw,rh) of S placed at (vox,voy) on a vw x vh
//      canvas whose other pixels are `fill_pre`          (crop + extend before the resize)
//   R  resize output (Rw x Rh); has_resize == false means R == V
//   F  final image: the rectangle (px,py,pw,ph) of R placed at (fox,foy) on an fw x fh canvas
//      whose other pixels are `fill_post`                (crop + extend after the resize)
struct PipeGeom {
    uint32_t sw = 0, sh = 0;
    uint32_t rx = 0, ry = 0, rw = 0, rh = 0;
    uint32_t vw = 0, vh = 0, vox = 0, voy = 0;
    uint8_t fill_pre[4] = {0, 0, 0, 0};
    bool has_resize = false;
    uint32_t Rw = 0, Rh = 0;
    uint32_t px = 0, py = 0, pw = 0, ph = 0;
    uint32_t fw = 0, fh = 0, fox = 0, foy = 0;
    uint8_t fill_post[4] = {0, 0, 0, 0};
};

slimg_status build_geometry(uint32_t w, uint32_t h, const slimg_pipeline_options &o, int order, PipeGeom &g);

// ---- Lanczos3 taps of one axis (taps.cpp) ------------------------------------------------
// image 0.25.9 imageops/sample.rs sampling bounds and weights, in strict f32, for the virtual
// axis of length `in_len` resampled to `out_len`; then clipped to the part of the axis that is
// backed by source pixels ([rect_off, rect_off+rect_len) in V coordinates).  Taps that fall on
// fill are summed into fillw.
struct AxisTaps {
    int out_len = 0, max_taps = 0;
    std::vector<int32_t> left;    // first source index IN RECT COORDINATES
    std::vector<int32_t> count;   // taps inside the rect (may be 0)
    std::vector<float> w;         // [out_len][max_taps], zero padded
    std::vector<float> fillw;     // [out_len] sum of the weights that land on fill
};
void build_axis_taps(uint32_t in_len, uint32_t out_len, uint32_t rect_off, uint32_t rect_len, AxisTaps &t);

// ---- fast-path plan (fastplan.cpp): fixed-point weight fragments for the tensor-core horizontal
// pass and the slot table of the streaming vertical pass ----------------------------------------
struct FastPlanHost {
    bool ok = false;                       // false: geometry outside the fast path, use the generic kernel
    uint32_t n_tiles = 0, n_strips = 0;
    std::vector<int32_t> tile_a0, tile_nchunks, tile_boff;   // tile_a0 relative to the strip's span
    std::vector<int32_t> strip_a0, strip_bytes;
    uint32_t pitch = 0;
    std::vector<uint32_t> bfrag;
    std::vector<float> fillw_x;
    float hscale = 1.0f;
    std::vector<float> vrow_w;             // [rows][8]
    std::vector<int16_t> vrow_out;         // [rows][8]
    std::vector<int32_t> vleft, vright;
    std::vector<float> fillw_y;
    std::vector<uint8_t> row_computed;
    bool need_fill = false;
};
void build_fast_plan(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, FastPlanHost &out);

// ---- V-first integer tensor pipeline plan (vplan.cpp, k2_vfirst.cu) ------------------------------
// Pass 1 (vertical) and pass 2 (horizontal) both run as u8 x s8 integer MMAs; the weights of taps.cpp
// are quantised to fixed point and split into NDIG signed base-256 digits.
struct VPlanHost {
    bool ok = false;
    int ndig = 3;
    // vertical: groups of 8 kept output rows
    uint32_t n_groups = 0;
    int nch1 = 0;                          // max 32-row chunks a group's band spans
    std::vector<int32_t> grp_a0, grp_need, grp_nch;
    std::vector<uint32_t> b1;              // [n_groups][4][ndig][32 lanes][2]
    std::vector<uint32_t> recs;            // [n_recs][rec_words]: header + fragments of each non-empty group
    uint32_t n_recs = 0, rec_words = 0;
    std::vector<int32_t> bias;             // [n_groups*8][4]
    std::vector<uint8_t> row_computed;     // [ph]
    int sh1 = 0;                           // V16 = (x + bias) >> sh1
    // horizontal: strips of <= 64 kept output columns, tiles of 8
    uint32_t n_strips = 0;
    int nch2 = 0;                          // max 32-pixel chunks a tile's band spans
    std::vector<int32_t> strip_col0;       // [n_strips + 1]
    std::vector<int32_t> strip_src0, strip_nblk, strip_bytes, strip_upw, strip_nunits;   // [n_strips]
    int unit = 4;                          // 16-byte blocks per pass-1 unit
    bool early_go = false;
    std::vector<int32_t> tile_kb0, tile_nch;                    // [n_strips][8]
    std::vector<uint32_t> b2;              // [n_strips][8][nch2][ndig][32][2]
    std::vector<float> cx;                 // [pw][4]
    std::vector<int32_t> cxi;              // [pw][4] the same constant term + rounding half, in units of 2^-sh2 (integer epilogue)
    int sh2 = 0;
    float s2[2][3] = {{0, 0, 0}, {0, 0, 0}};
    // shared-memory geometry
    uint32_t ring_rows = 0, pitch = 0, pitch_v = 0, smem_bytes = 0;
    uint32_t vbufs = 3;                    // V buffers (k2_vfirst.cu): 2..4
    uint32_t vbufs_pref = 3;               // the band-length rule's choice, before the 16-column retry (vplan.cpp)
    uint32_t cohorts = 1;                  // pass-1 warp cohorts: cohort k filters groups k, k + cohorts, ... (k2_vfirst.cu)
    bool need_fill = false, rows_pure_fill = false;
};
void build_vplan(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, int ndig, VPlanHost &out);
// fixed-point helpers shared by the integer plans (vplan.cpp)
void quantise(const float *w, int cnt, int S, std::vector<int64_t> &q);     // taps -> integers with the exact sum
bool split_digits(int64_t x, int ndig, int8_t d[3]);                         // balanced base-256 digits
int pick_scale(float maxw, int ndig);
void put_b(uint32_t *frag, int k, int n, int8_t v);                          // mma.m16n8k32 B fragment slot

// ---- integer tile kernel plan (tileplan.cpp, k2_tile.cu): ratios below ~3.5x and upscales ------------------------
struct TilePlanHost {
    bool ok = false;
    uint32_t tw = 32, th = 32;             // kept output columns / rows per CTA (multiples of 8)
    int nch1 = 0, nch2 = 0, sh1 = 0;
    uint32_t rec_words = 0;
    std::vector<uint32_t> recs;            // [groups of 8 kept rows][rec_words]: {a0, need, chunks, .., bias[8][4]} + weight fragments
    std::vector<uint8_t> row_computed;     // [ph]
    uint32_t n_rowblk = 0, n_colblk = 0, rows_max = 0;
    std::vector<int32_t> rb_y0, rb_rows;   // [n_rowblk] first source-rectangle row / rows a row block stages
    std::vector<int32_t> cb_src0, cb_nblk; // [n_colblk] first staged source pixel (multiple of 4) / 16-byte blocks per row (even)
    std::vector<int32_t> tile_kb0;         // [n_colblk][tw / 8] first block of the tile's k range
    std::vector<uint32_t> b2;              // [n_colblk][tw / 8][nch2][2 digits][32 lanes][2]; packed: [n_colblk][tw / 8][3][32][2]
    std::vector<int32_t> cxi;              // [pw][4] constant term + rounding half, units of 2^-sh2
    int sh2 = 0;                           // final right shift of the pass-2 integer sum
    bool packed = false;                   // both limbs in one K = 32 (every tile's taps span <= 16 staged pixels: upscales)
    uint32_t v_words = 0, pitch = 0, pitch_v = 0, smem_bytes = 0;
    bool need_fill = false, rows_pure_fill = false;
};
void build_tileplan(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, bool forced, TilePlanHost &out);

// ---- tcgen05 pipeline plan (tcplan.cpp, k2_tc.cu) -------------------------------------------------------
// Vertical pass on the 5th-generation tensor cores: source rows stream through a shared-memory ring in chunks of
// 32 rows (the K of one tcgen05.mma.kind::i8), accumulators for blocks of 16 kept output rows live in TMEM.
struct TcPlanHost {
    bool ok = false;
    int ndig = 2, sh1 = 0;
    // vertical: blocks of 8 kept output rows, chunks of 32 source-rectangle rows
    uint32_t n_blocks = 0, n_chunks = 0, nb_max = 0;
    uint32_t crec_words = 0;               // words per chunk record: 16 header + nb_max * 128 (one N = 16 weight tile per block)
    std::vector<uint32_t> crecs;           // [n_chunks][crec_words]
    std::vector<int32_t> blk_c0, blk_c1;   // [n_blocks] first / last chunk of each block
    std::vector<int32_t> bias;             // [n_blocks*8][4]
    // horizontal: strips of kept output columns, tiles of 8; staged span starts on a 128-byte boundary
    uint32_t n_strips = 0, n_panels = 0;   // panels of 128 bytes (32 pixels) staged per strip (max over strips)
    int nch2 = 0;
    std::vector<int32_t> strip_col0;       // [n_strips + 1]
    std::vector<int32_t> strip_src0;       // [n_strips] first staged source pixel (multiple of 4)
    std::vector<int32_t> tile_kb0;         // [n_strips][8] first 4-pixel block of the tile's k range
    std::vector<uint32_t> b2;              // [n_strips][8][nch2][ndig][32][2]
    std::vector<float> cx;                 // [pw][4]
    float s2[2][3] = {{0, 0, 0}, {0, 0, 0}};
    uint32_t slots = 0, slot_bytes = 0, pitch_v = 0, smem_bytes = 0, nbuf = 0, vbufs = 3;
    bool need_fill = false;
};
void build_tcplan(const PipeGeom &g, const AxisTaps &tx, const AxisTaps &ty, TcPlanHost &out);

}  // namespace slimg
