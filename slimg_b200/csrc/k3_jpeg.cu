// k3_jpeg.cu -- the JPEG encoder's pre-entropy stage, fused into one kernel: RGBA8 in,
// quantised DCT coefficients out, nothing in between touches HBM.
//
// Replaces what JpegCodec::encode (crates/slimg-core/src/codec/jpeg.rs:51-72) makes MozJPEG
// run before its entropy coder; arithmetic follows the libjpeg-turbo 2.1.5.1 sources vendored
// in the reference tree (crates/libjxl-sys/libjxl/third_party/libjpeg-turbo/), bit for bit:
//   to_rgb               codec/mod.rs:42-48        alpha dropped, no compositing
//   rgb_ycc_convert      jccolor.c:209-237, jccolext.c:61-68    16-bit fixed point
//   edge expansion       jcsample.c:94-111, jcprepct.c:109-118,155-183
//   h2v2_downsample      jcsample.c:255-285        bias 1,2,1,2,...
//   convsamp / islow FDCT / quantize   jcdctmgr.c:373-438, jfdctint.c:143-286
//   dummy blocks         jccoefct.c:286-319
//
// One WARP = a tile of kWMcus consecutive MCUs (16x16 pixels each) of one MCU row; a persistent grid of warps walks the
// tiles of the whole batch (no CTA-wide barrier after the table load):
//   stage 1  every lane converts patches of 4 x 2 pixels (two 16-byte loads, the next patch's loads issued before the
//            current one is converted): 8 Y samples go to the warp's Y tile as two packed words, the 2 x 2 chroma
//            averages (bias 1,2) straight to the half-resolution Cb / Cr tiles -- the full-resolution chroma planes of
//            the reference never exist
//   stage 2  a block belongs to 8 LANES: lane r runs the row DCT of row r, the 8x8 intermediate is transposed through a
//            padded per-warp scratch, lane c runs the column DCT of column c and quantises its 8 coefficients with
//            table entries read from shared memory; optionally the un-quantised coefficients are stored as well (the
//            block MozJPEG keeps for its trellis pass)
//
// Roofline: nominally HBM (w*h*4 read + (Y_HB*Y_WB + 2*C_HB*C_WB)*128 written = 7 B/px), but at ~65 integer
// instructions per pixel the kernel is bound by instruction issue (every integer instruction of this part issues at
// ~65 lanes/clk/SM, profiles/r01f_microbench_b200.txt).
#include <algorithm>
#include <cstdlib>

#include "kernels.cuh"

namespace slimg {

namespace {

#define SLIMG_FIX(x) int((x) * 65536.0 + 0.5)

// The three dot products of rgb_ycc_convert as IDP.4A on the packed pixel: every 16-bit coefficient is split into a
// signed high digit and an unsigned low digit (c = 256 hi + lo), the 0.5 coefficient (32768 = 2 * 64 * 256) is spread
// over two byte slots of a permuted pixel.  6 IDP.4A + 2 PRMT + 3 IMAD instead of unpacking three bytes and 9 IMAD; the
// integer sums are identical (no intermediate exceeds 2^25).
__host__ __device__ constexpr uint32_t pack4(int a, int b, int c, int d) {
    return uint32_t(a & 0xff) | (uint32_t(b & 0xff) << 8) | (uint32_t(c & 0xff) << 16) | (uint32_t(d & 0xff) << 24);
}
static_assert(SLIMG_FIX(0.29900) == 76 * 256 + 139 && SLIMG_FIX(0.58700) == 150 * 256 + 70 && SLIMG_FIX(0.11400) == 29 * 256 + 47, "Y digits");
static_assert(-SLIMG_FIX(0.16874) == -44 * 256 + 205 && -SLIMG_FIX(0.33126) == -85 * 256 + 51 && SLIMG_FIX(0.50000) == 128 * 256, "Cb digits");
static_assert(-SLIMG_FIX(0.41869) == -108 * 256 + 209 && -SLIMG_FIX(0.08131) == -21 * 256 + 47, "Cr digits");
__device__ __forceinline__ int dp4a_uu(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {       // unsigned bytes of a, signed bytes of b
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ void rgb_to_ycc(uint32_t px, uint8_t &y, uint8_t &cb, uint8_t &cr) {
    constexpr int half = 1 << 15, cboff = 128 << 16;
    const uint32_t rgbb = __byte_perm(px, 0u, 0x2210), rgbr = __byte_perm(px, 0u, 0x0210);
    y  = uint8_t((dp4a_uu(px, pack4(76, 150, 29, 0), 0) * 256 + dp4a_uu(px, pack4(139, 70, 47, 0), half)) >> 16);
    cb = uint8_t((dp4a_us(rgbb, pack4(-44, -85, 64, 64), 0) * 256 + dp4a_uu(px, pack4(205, 51, 0, 0), cboff + half - 1)) >> 16);
    cr = uint8_t((dp4a_us(rgbr, pack4(64, -108, -21, 64), 0) * 256 + dp4a_uu(px, pack4(0, 209, 47, 0), cboff + half - 1)) >> 16);
}

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// one 8-point pass of jpeg_fdct_islow (jfdctint.c:143-286); PASS = 1 rows, 2 columns
template <int PASS>
__device__ __forceinline__ void fdct8(const int in[8], int out[8]) {
    const int tmp0 = in[0] + in[7], tmp7 = in[0] - in[7];
    const int tmp1 = in[1] + in[6], tmp6 = in[1] - in[6];
    const int tmp2 = in[2] + in[5], tmp5 = in[2] - in[5];
    const int tmp3 = in[3] + in[4], tmp4 = in[3] - in[4];
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3;
    const int tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    constexpr int sh = PASS == 1 ? 13 - 2 : 13 + 2;
    if (PASS == 1) { out[0] = (tmp10 + tmp11) << 2; out[4] = (tmp10 - tmp11) << 2; }
    else           { out[0] = descale(tmp10 + tmp11, 2); out[4] = descale(tmp10 - tmp11, 2); }
    int z1 = (tmp12 + tmp13) * 4433;
    out[2] = descale(z1 + tmp13 * 6270, sh);
    out[6] = descale(z1 - tmp12 * 15137, sh);
    z1 = tmp4 + tmp7;
    int z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7;
    const int z5 = (z3 + z4) * 9633;
    const int t4 = tmp4 * 2446, t5 = tmp5 * 16819, t6 = tmp6 * 25172, t7 = tmp7 * 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    out[7] = descale(t4 + z1 + z3, sh);
    out[5] = descale(t5 + z2 + z4, sh);
    out[3] = descale(t6 + z2 + z3, sh);
    out[1] = descale(t7 + z1 + z4, sh);
}

__device__ __forceinline__ void ycc3(uint32_t px, int &y, int &cb, int &cr) {
    uint8_t a, b, c;
    rgb_to_ycc(px, a, b, c);
    y = a; cb = b; cr = c;
}

__device__ __forceinline__ void unpack8(uint32_t lo, uint32_t hi, int (&row)[8]) {
    row[0] = int(lo & 0xff) - 128; row[1] = int((lo >> 8) & 0xff) - 128; row[2] = int((lo >> 16) & 0xff) - 128; row[3] = int(lo >> 24) - 128;
    row[4] = int(hi & 0xff) - 128; row[5] = int((hi >> 8) & 0xff) - 128; row[6] = int((hi >> 16) & 0xff) - 128; row[7] = int(hi >> 24) - 128;
}

// ---- the kernel: warp-autonomous tiles -------------------------------------------------------------------------
constexpr int kWMcus = 8;
constexpr int kWW = 16 * kWMcus;                // 128 pixels
constexpr int kWarpsPerCta = 4;
constexpr int kRowStride = 12;                  // words between scratch rows: 16-byte chunks of 8 lanes hit distinct banks
constexpr int kBlkStride = 8 * kRowStride + 8;  // words between the scratch blocks of a warp: column reads of 4 blocks disjoint

struct WarpSmem {
    // row pitches == 4 words mod 32: the 8 rows a group of lanes reads for one block land in distinct banks
    uint32_t ys[16][kWW / 4 + 4];     // Y samples, 4 per word
    uint16_t cb[8][kWW / 4 + 8];      // downsampled Cb / Cr, one halfword (2 samples) per stage-1 patch
    uint16_t cr[8][kWW / 4 + 8];
    int scratch[4 * kBlkStride];      // row-pass output of the 4 blocks in flight
    int16_t dc[2][2][2 * kWMcus];     // [quantised, un-quantised] DC of the Y blocks (dummy-block patch)
};

// row DCT by lane r of the group, transpose through scratch, column DCT + quantisation by lane c; returns the 8
// coefficients (k, c), k = 0..7, of the block in out[] and, with UQ, the un-quantised ones in uq[]
template <bool UQ>
__device__ __forceinline__ void block_dct(int *scr, int l8, uint32_t lo, uint32_t hi, const uint32_t (*qt)[64], int (&out)[8], int (&uq)[8]) {
    int in[8], mid[8];
    unpack8(lo, hi, in);
    fdct8<1>(in, mid);
    *reinterpret_cast<int4 *>(scr + l8 * kRowStride) = make_int4(mid[0], mid[1], mid[2], mid[3]);
    *reinterpret_cast<int4 *>(scr + l8 * kRowStride + 4) = make_int4(mid[4], mid[5], mid[6], mid[7]);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 8; k++) in[k] = scr[k * kRowStride + l8];
    __syncwarp();
    fdct8<2>(in, mid);
#pragma unroll
    for (int k = 0; k < 8; k++) {          // jcdctmgr.c:408-438, reciprocal form; shift >= 34, see build_quant
        const int t = mid[k];
        if (UQ) uq[k] = t;
        const uint32_t a = uint32_t(t < 0 ? -t : t) + qt[1][k * 8 + l8];
        const uint32_t q = __umulhi(a, qt[0][k * 8 + l8]) >> qt[2][k * 8 + l8];
        out[k] = t < 0 ? -int(q) : int(q);
    }
}

// ---- stage 2 on the tensor cores ---------------------------------------------------------------------------------
// jpeg_fdct_islow is two passes of a LINEAR map with one rounding per output: mid = (M x + 1024) >> 11 along the rows,
// coef = (M mid + 16384) >> 15 down the columns, with M the integer matrix the butterfly network of jfdctint.c:158-213
// amounts to (rows 0 and 4 scaled by 8192 so that the `<< 2` / DESCALE(., 2) of the DC terms fit the same shifts).  Both
// products are exact integer GEMMs: M in two signed base-256 digits, samples as u8, the 16-bit intermediate offset to
// unsigned and split into two u8 limbs -- mma.sync.m16n8k32.u8.s8, bit-identical to the butterflies (every sum is exact;
// the roundings are the same two shifts).  Verified against the real libjpeg-turbo like every other path.
__constant__ int kDctM[8][8] = {
    {8192, 8192, 8192, 8192, 8192, 8192, 8192, 8192},
    {11363, 9633, 6437, 2260, -2260, -6437, -9633, -11363},
    {10703, 4433, -4433, -10703, -10703, -4433, 4433, 10703},
    {9633, -2259, -11362, -6436, 6436, 11362, 2259, -9633},
    {8192, -8192, -8192, 8192, 8192, -8192, -8192, 8192},
    {6437, -11362, 2261, 9633, -9633, -2261, 11362, -6437},
    {4433, -10704, 10704, -4433, -4433, 10704, -10704, 4433},
    {2260, -6436, 9633, -11363, 11363, -9633, 6436, -2260}};
__device__ __forceinline__ int dct_lo(int m) { return ((m + 128) & 255) - 128; }        // balanced digits: m = 256 hi + lo
__device__ __forceinline__ int dct_hi(int m) { return (m - dct_lo(m)) / 256; }
__device__ __forceinline__ uint32_t pack_s8(int a, int b, int c, int d) {
    return uint32_t(a & 0xff) | (uint32_t(b & 0xff) << 8) | (uint32_t(c & 0xff) << 16) | (uint32_t(d & 0xff) << 24);
}
// m16n8k16 u8 x s8: A = 2 registers (rows g and g + 8, k = 4t..4t+3), B = 1 register, C given as four registers / as zero
__device__ __forceinline__ void imma16_c(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b, int c0, int c1, int c2, int c3) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%8,%9,%10};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(a0), "r"(a1), "r"(b), "r"(c0), "r"(c1), "r"(c2), "r"(c3));
}
__device__ __forceinline__ void imma16_z(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%7,%7,%7};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(a0), "r"(a1), "r"(b), "r"(0));
}
// transposes an 8 x 8 matrix of 16-bit elements held one register per lane (row lane/4, columns 2(lane%4), +1)
__device__ __forceinline__ uint32_t movmatrix_t(uint32_t a) {
    uint32_t d;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
    return d;
}
struct DctRegs {                 // per-lane constants of the two passes (lane = 4 g + t)
    uint32_t w1lo[2], w1hi[2];   // pass 1, digit d: weights of output g over samples 4(t&1)..+3, for lanes t < 2 / t >= 2
    uint32_t b2a, b2b, b2c;      // pass 2: [lo x L], [lo x H + hi x L], [hi x H] over k = (row 2t lo, hi, row 2t+1 lo, hi)
    int c1[4], c2[4];            // rounding, level shift and limb offset, for outputs 2t and 2t+1 ({e0, e1, e0, e1}: the MMA's C quad)
};
__device__ __forceinline__ DctRegs make_dct_regs(int lane) {
    const int g = lane >> 2, t = lane & 3;
    DctRegs r;
    const int i0 = 4 * (t & 1);
    const uint32_t wl = pack_s8(dct_lo(kDctM[g][i0]), dct_lo(kDctM[g][i0 + 1]), dct_lo(kDctM[g][i0 + 2]), dct_lo(kDctM[g][i0 + 3]));
    const uint32_t wh = pack_s8(dct_hi(kDctM[g][i0]), dct_hi(kDctM[g][i0 + 1]), dct_hi(kDctM[g][i0 + 2]), dct_hi(kDctM[g][i0 + 3]));
    r.w1lo[0] = t < 2 ? wl : 0u; r.w1lo[1] = t < 2 ? wh : 0u;
    r.w1hi[0] = t < 2 ? 0u : wl; r.w1hi[1] = t < 2 ? 0u : wh;
    const int ma = kDctM[g][2 * t], mb = kDctM[g][2 * t + 1];
    r.b2a = pack_s8(dct_lo(ma), 0, dct_lo(mb), 0);
    r.b2b = pack_s8(dct_hi(ma), dct_lo(ma), dct_hi(mb), dct_lo(mb));
    r.b2c = pack_s8(0, dct_hi(ma), 0, dct_hi(mb));
#pragma unroll
    for (int e = 0; e < 2; e++) {
        const int n = 2 * t + e;
        // pass 1: x = M (s - 128) + 1024, and + 32768 << 11 so that the intermediate is unsigned: mid' = mid + 32768
        r.c1[e] = r.c1[2 + e] = 1024 + (32768 << 11) - (n == 0 ? 128 * 65536 : 0);
        // pass 2: x = M (mid' - 32768) + 16384; 32768 * rowsum(M[0]) = 2^31 wraps, like the sum it is subtracted from
        r.c2[e] = r.c2[2 + e] = int(16384u - (n == 0 ? 0x80000000u : 0u));
    }
    return r;
}

// ---- MozJPEG's overshoot deringing (optional; mozjpeg-sys is not in the reference tree: restated, parity unpinned) -----------
// jcdctmgr.c preprocess_deringing: in a block that holds some but not only samples at the maximum, every run of maximum samples
// (walked in zig-zag order) is replaced by a Catmull-Rom curve through its neighbours' slopes, capped at
// 127 + min(31, 2 * DC quantiser, headroom of the block's DC).  Such blocks are rare in photographs and the overshoot leaves
// the u8 range, so they are redone by ONE LANE each after the tile's regular pass (scalar butterflies, the same arithmetic
// as jfdctint.c), and overwrite what the regular pass stored.
__constant__ uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
__device__ __noinline__ void dering_block(int *d, int dc_quant) {
    constexpr int maxsample = 127, size = 64;
    int sum = 0, count = 0;
    for (int i = 0; i < size; i++) { sum += d[i]; count += d[i] >= maxsample; }
    if (!count || count == size) return;
    const int maxovershoot = maxsample + min(min(31, 2 * dc_quant), (maxsample * size - sum) / count);
    int n = 0;
    do {
        if (d[kZigzag[n]] < maxsample) { n++; continue; }
        const int start = n;
        while (++n < size && d[kZigzag[n]] >= maxsample) {}
        const int end = n;
        const int f1 = d[kZigzag[start >= 1 ? start - 1 : 0]], f2 = d[kZigzag[start >= 2 ? start - 2 : 0]];
        const int l1 = d[kZigzag[end < size - 1 ? end : size - 1]], l2 = d[kZigzag[end < size - 2 ? end + 1 : size - 1]];
        int fslope = max(f1 - f2, maxsample - f1), lslope = max(l1 - l2, maxsample - l1);
        if (start == 0) fslope = lslope;
        if (end == size) lslope = fslope;
        const int length = end - start;
        const float step = __fdiv_rn(1.f, float(length + 1));
        float position = step;
        const int v1 = maxsample - fslope, v4 = maxsample - lslope;
        const int tan1 = (maxsample - v1) * length, tan2 = (v4 - maxsample) * length;
        for (int i = start; i < end; i++, position = __fadd_rn(position, step)) {
            // catmull_rom(v1, max, max, v4, t, length) in the reference's operation order, no contraction
            const float t = position, t2 = __fmul_rn(t, t), t3 = __fmul_rn(t2, t);
            const float c1 = __fadd_rn(__fsub_rn(__fmul_rn(2.f, t3), __fmul_rn(3.f, t2)), 1.f);
            const float c2 = __fadd_rn(__fmul_rn(-2.f, t3), __fmul_rn(3.f, t2));
            const float c3 = __fadd_rn(__fsub_rn(t3, __fmul_rn(2.f, t2)), t);
            const float c4 = __fsub_rn(t3, t2);
            const float val = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(float(maxsample), c1), __fmul_rn(float(tan1), c3)), __fmul_rn(float(maxsample), c2)),
                                        __fmul_rn(float(tan2), c4));
            d[kZigzag[i]] = min(int(ceilf(val)), maxovershoot);
        }
        n++;
    } while (n < size);
}
// the whole block on one lane: dering, islow FDCT (rows, then columns), quantise, clamp to 10 bits (MozJPEG does when deringing is on)
template <bool UQ>
__device__ __noinline__ void slow_block(int *d, int dc_quant, const uint32_t (*qt)[64], int16_t *q_out, int16_t *u_out) {
    dering_block(d, dc_quant);
    for (int r = 0; r < 8; r++) {
        int in[8], out[8];
#pragma unroll
        for (int k = 0; k < 8; k++) in[k] = d[r * 8 + k];
        fdct8<1>(in, out);
#pragma unroll
        for (int k = 0; k < 8; k++) d[r * 8 + k] = out[k];
    }
    for (int c = 0; c < 8; c++) {
        int in[8], out[8];
#pragma unroll
        for (int k = 0; k < 8; k++) in[k] = d[k * 8 + c];
        fdct8<2>(in, out);
#pragma unroll
        for (int k = 0; k < 8; k++) d[k * 8 + c] = out[k];
    }
    for (int i = 0; i < 64; i++) {
        const int t = d[i];
        if (UQ) u_out[i] = int16_t(t);
        const uint32_t a = uint32_t(t < 0 ? -t : t) + qt[1][i];
        int q = int(__umulhi(a, qt[0][i]) >> qt[2][i]);
        q = min(q, 1023);
        q_out[i] = int16_t(t < 0 ? -q : q);
    }
}

// PIPE: 1 = the loads of stage-1 patch j+1 are issued before patch j is converted; 2 = additionally, the first patch of
// the warp's NEXT tile is loaded before stage 2 of the current one (8 more live registers across the DCT)
template <int PIPE, int MINB, bool UQ, bool TENSOR, bool DERING>
__global__ void __launch_bounds__(32 * kWarpsPerCta, MINB) jpeg_frontend_warp_kernel(const __grid_constant__ JpegParams p, const uint32_t n_images) {
    __shared__ __align__(16) WarpSmem wsm[kWarpsPerCta];
    const int lane = threadIdx.x & 31, l8 = lane & 7, q4 = lane >> 3;
    WarpSmem &sm = wsm[threadIdx.x >> 5];
    int *scr = sm.scratch + q4 * kBlkStride;
    const int w = int(p.w), h = int(p.h);
    const int creal = (h + 1) / 2;   // chroma rows that come from real downsampling
    const bool vec_ok = (w & 3) == 0 && (reinterpret_cast<uintptr_t>(p.src) & 15) == 0 && (p.src_img_bytes & 15) == 0;
    const uint32_t tiles_x = (p.c_wb + kWMcus - 1) / kWMcus, tiles_img = tiles_x * p.c_hb;
    const uint64_t n_tiles = uint64_t(tiles_img) * n_images;
    const uint64_t gw = uint64_t(blockIdx.x) * kWarpsPerCta + (threadIdx.x >> 5), nw = uint64_t(gridDim.x) * kWarpsPerCta;

    // both quantisation tables in shared memory: a lane reads its own column of them (a per-lane index into the
    // constant bank would serialise, and 48 table registers per lane cost a third of the occupancy)
    __shared__ uint32_t qt[2][3][64];
    for (int i = threadIdx.x; i < 128; i += 32 * kWarpsPerCta) {
        qt[i >> 6][0][i & 63] = p.q.recip[i >> 6][i & 63]; qt[i >> 6][1][i & 63] = p.q.corr[i >> 6][i & 63]; qt[i >> 6][2][i & 63] = p.q.shift32[i >> 6][i & 63];
    }
    __syncthreads();

    // tensor-core stage 2: this lane's matrix digits and its two quantisation entries (coefficients (2t, g), (2t+1, g))
    const int g8 = lane >> 2, t4 = lane & 3;
    DctRegs dr;
    uint32_t qrec[2][2], qcor[2][2], qsh[2][2];          // [table][e]
    if (TENSOR) {
        dr = make_dct_regs(lane);
#pragma unroll
        for (int tb = 0; tb < 2; tb++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int idx = (2 * t4 + e) * 8 + g8;
                qrec[tb][e] = qt[tb][0][idx]; qcor[tb][e] = qt[tb][1][idx]; qsh[tb][e] = qt[tb][2][idx];
            }
    }
    // DCT + quantisation of 8 blocks: 16 sample rows (two block rows, or Cb over Cr) x 32 samples (4 block columns) whose
    // A fragment is a0..a3.  Block (hb, j) goes to q[hb] + 64 j (this lane's two coefficients (g, 2t), (g, 2t+1) as one
    // 32-bit store: a warp writes the 128 bytes of a block); nv block columns are real, ok1 says whether row 1 is.
    // dcs: where lane 0 leaves the blocks' DC terms (dummy-block patch, Y only).
    auto dct8 = [&](uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, int tb, int16_t *q0, int16_t *q1, int16_t *u0, int16_t *u1,
                    int nv, bool ok1, int16_t *dcs, int dcs_row) {
        // pass 1 (rows): block column j takes samples 8j..8j+7 = half of a K = 16 step: j = 0, 1 from (a0, a1), j = 2, 3 from
        // (a2, a3), the weights masked to the lanes that hold that half
        int acc[4][2][4];
        imma16_c(acc[0][0], a0, a1, dr.w1lo[0], dr.c1[0], dr.c1[1], dr.c1[2], dr.c1[3]);
        imma16_c(acc[1][0], a0, a1, dr.w1hi[0], dr.c1[0], dr.c1[1], dr.c1[2], dr.c1[3]);
        imma16_c(acc[2][0], a2, a3, dr.w1lo[0], dr.c1[0], dr.c1[1], dr.c1[2], dr.c1[3]);
        imma16_c(acc[3][0], a2, a3, dr.w1hi[0], dr.c1[0], dr.c1[1], dr.c1[2], dr.c1[3]);
        imma16_z(acc[0][1], a0, a1, dr.w1lo[1]);
        imma16_z(acc[1][1], a0, a1, dr.w1hi[1]);
        imma16_z(acc[2][1], a2, a3, dr.w1lo[1]);
        imma16_z(acc[3][1], a2, a3, dr.w1hi[1]);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            uint32_t m[4];
#pragma unroll
            for (int i = 0; i < 4; i++) m[i] = uint32_t((acc[j][0][i] + (acc[j][1][i] << 8)) >> 11);      // mid + 32768, 16 bits
            // rows g (block row 0) and g + 8 (block row 1), frequencies 2t, 2t+1 -> transposed: rows 2t, 2t+1 of frequency g
            const uint32_t t0 = movmatrix_t(__byte_perm(m[0], m[1], 0x5410)), t1 = movmatrix_t(__byte_perm(m[2], m[3], 0x5410));
            int da[4], db[4], dc[4];
            imma16_c(da, t0, t1, dr.b2a, dr.c2[0], dr.c2[1], dr.c2[2], dr.c2[3]);
            imma16_z(db, t0, t1, dr.b2b);
            imma16_z(dc, t0, t1, dr.b2c);
            int cf[4], qv[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                cf[i] = (da[i] + (db[i] << 8) + (dc[i] << 16)) >> 15;          // coefficient (u = 2t + (i & 1), u' = g), block row i >> 1
                const int e = i & 1;
                const uint32_t aa = uint32_t(cf[i] < 0 ? -cf[i] : cf[i]) + qcor[tb][e];
                const uint32_t q = __umulhi(aa, qrec[tb][e]) >> qsh[tb][e];       // jcdctmgr.c:408-438, reciprocal form
                qv[i] = cf[i] < 0 ? -int(DERING ? min(q, 1023u) : q) : int(DERING ? min(q, 1023u) : q);
            }
            // natural order wants (u, u') -> u * 8 + u': transpose back so a lane holds u = g, u' = 2t, 2t+1
            const uint32_t w0 = movmatrix_t(__byte_perm(uint32_t(qv[0]), uint32_t(qv[1]), 0x5410));
            const uint32_t w1 = movmatrix_t(__byte_perm(uint32_t(qv[2]), uint32_t(qv[3]), 0x5410));
            uint32_t x0 = 0u, x1 = 0u;
            if (UQ) {
                x0 = movmatrix_t(__byte_perm(uint32_t(cf[0]), uint32_t(cf[1]), 0x5410));
                x1 = movmatrix_t(__byte_perm(uint32_t(cf[2]), uint32_t(cf[3]), 0x5410));
            }
            if (j < nv) {
                *reinterpret_cast<uint32_t *>(q0 + j * 64) = w0;
                if (UQ) *reinterpret_cast<uint32_t *>(u0 + j * 64) = x0;
                if (ok1) {
                    *reinterpret_cast<uint32_t *>(q1 + j * 64) = w1;
                    if (UQ) *reinterpret_cast<uint32_t *>(u1 + j * 64) = x1;
                }
                // DC of the block: coefficient (0, 0) sits in lane 0 before AND after the transposes
                if (dcs && lane == 0) {
                    dcs[j] = int16_t(qv[0]); dcs[dcs_row + j] = int16_t(qv[2]);
                    if (UQ) { dcs[2 * dcs_row + j] = int16_t(x0 & 0xffffu); dcs[3 * dcs_row + j] = int16_t(x1 & 0xffffu); }
                }
            }
        }
    };

    // the two 16-byte loads of stage-1 patch (pr, lane) of the tile at (x0, y0) of image `src`
    auto load_patch = [&](const uint32_t *src, int x0, int y0, int pr, uint32_t (&px)[2][4]) {
        const int x = x0 + 4 * lane, yy0 = y0 + 2 * pr;
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const uint32_t *row = src + size_t(min(yy0 + r, h - 1)) * w;
            if (vec_ok && x + 3 < w) {
                const uint4 v = __ldg(reinterpret_cast<const uint4 *>(row + x));
                px[r][0] = v.x; px[r][1] = v.y; px[r][2] = v.z; px[r][3] = v.w;
            } else {
#pragma unroll
                for (int k = 0; k < 4; k++) px[r][k] = __ldg(row + min(x + k, w - 1));
            }
        }
    };
    // colour conversion of one patch with edge replication.  Y uses row min(y, h-1).  Chroma row cy past creal-1 repeats
    // the last DOWNSAMPLED row (jcprepct.c:174-183): other source rows than Y, below the image only.
    auto convert_patch = [&](const uint32_t *src, int x0, int y0, int pr, const uint32_t (&px)[2][4]) {
        const int x = x0 + 4 * lane, yy0 = y0 + 2 * pr;
        const int sy[2] = {min(yy0, h - 1), min(yy0 + 1, h - 1)};
        const int cyy = min(yy0 >> 1, creal - 1);
        const int sc[2] = {min(2 * cyy, h - 1), min(2 * cyy + 1, h - 1)};
        int yv[2][4], cbv[2][4], crv[2][4];
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int k = 0; k < 4; k++) ycc3(px[r][k], yv[r][k], cbv[r][k], crv[r][k]);
        if (sc[0] != sy[0] || sc[1] != sy[1]) {
#pragma unroll
            for (int r = 0; r < 2; r++)
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    int unused;
                    ycc3(__ldg(src + size_t(sc[r]) * w + min(x + k, w - 1)), unused, cbv[r][k], crv[r][k]);
                }
        }
#pragma unroll
        for (int r = 0; r < 2; r++)
            sm.ys[2 * pr + r][lane] = uint32_t(yv[r][0]) | (uint32_t(yv[r][1]) << 8) | (uint32_t(yv[r][2]) << 16) | (uint32_t(yv[r][3]) << 24);
        // h2v2_downsample (jcsample.c:272-284): bias 1 for even output columns, 2 for odd ones
        const int cb0 = (cbv[0][0] + cbv[0][1] + cbv[1][0] + cbv[1][1] + 1) >> 2, cb1 = (cbv[0][2] + cbv[0][3] + cbv[1][2] + cbv[1][3] + 2) >> 2;
        const int cr0 = (crv[0][0] + crv[0][1] + crv[1][0] + crv[1][1] + 1) >> 2, cr1 = (crv[0][2] + crv[0][3] + crv[1][2] + crv[1][3] + 2) >> 2;
        sm.cb[pr][lane] = uint16_t(cb0 | (cb1 << 8));
        sm.cr[pr][lane] = uint16_t(cr0 | (cr1 << 8));
    };
    auto tile_at = [&](uint64_t t, uint32_t &z, int &my, int &mx0) {
        z = uint32_t(t / tiles_img);
        const uint32_t rem = uint32_t(t - uint64_t(z) * tiles_img);
        my = int(rem / tiles_x); mx0 = int(rem - uint32_t(my) * tiles_x) * kWMcus;
    };

    uint32_t pa[2][4], pb[2][4];           // patch registers: even / odd patch rows
    if (PIPE == 2 && gw < n_tiles) {
        uint32_t z; int my, mx0;
        tile_at(gw, z, my, mx0);
        if (lane * 4 < min(kWMcus, int(p.c_wb) - mx0) * 16)
            load_patch(reinterpret_cast<const uint32_t *>(p.src + size_t(z) * p.src_img_bytes), mx0 * 16, my * 16, 0, pa);
    }
#pragma unroll 1
    for (uint64_t t = gw; t < n_tiles; t += nw) {
        uint32_t z; int my, mx0;
        tile_at(t, z, my, mx0);
        const int nm = min(kWMcus, int(p.c_wb) - mx0);
        const int x0 = mx0 * 16, y0 = my * 16;
        const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(z) * p.src_img_bytes);

        // stage 1: patch rows 0..7 of column `lane` (lanes right of a partial tile idle)
        if (!(p.debug & 1u) && lane * 4 < nm * 16) {
            if (PIPE == 0) {
#pragma unroll 1
                for (int pr = 0; pr < 8; pr++) { load_patch(src, x0, y0, pr, pa); convert_patch(src, x0, y0, pr, pa); }
            } else {
                if (PIPE == 1) load_patch(src, x0, y0, 0, pa);
#pragma unroll 1
                for (int pr = 0; pr < 8; pr += 2) {
                    load_patch(src, x0, y0, pr + 1, pb);
                    convert_patch(src, x0, y0, pr, pa);
                    if (pr + 2 < 8) load_patch(src, x0, y0, pr + 2, pa);
                    convert_patch(src, x0, y0, pr + 1, pb);
                }
            }
        }
        if (PIPE == 2 && t + nw < n_tiles) {           // first patch of the next tile: in flight during stage 2
            uint32_t zn; int myn, mxn;
            tile_at(t + nw, zn, myn, mxn);
            if (lane * 4 < min(kWMcus, int(p.c_wb) - mxn) * 16)
                load_patch(reinterpret_cast<const uint32_t *>(p.src + size_t(zn) * p.src_img_bytes), mxn * 16, myn * 16, 0, pa);
        }
        __syncwarp();

        // stage 2, Y: 4 blocks per step (one per group of 8 lanes), block rows one after the other
        int16_t *yout = p.y + size_t(z) * p.y_img_elems;
        int16_t *yuq = UQ ? p.yu + size_t(z) * p.y_img_elems : nullptr;
        const int ybx0 = mx0 * 2, yby0 = my * 2;
        const int nby = min(2 * nm, int(p.y_wb) - ybx0);          // real Y block columns of the tile
        const int nrow = min(2, int(p.y_hb) - yby0);              // real Y block rows
        if (TENSOR) {
            // Y: 4 segments of 32 samples (4 block columns) x 16 rows (both block rows of the MCU row)
#pragma unroll 1
            for (int sg = 0; sg < ((p.debug & 2u) ? 0 : 4); sg++) {
                if (sg * 4 >= nby) break;                              // warp-uniform: the segment lies right of the image
                const uint32_t a0 = sm.ys[g8][sg * 8 + t4], a1 = sm.ys[g8 + 8][sg * 8 + t4];
                const uint32_t a2 = sm.ys[g8][sg * 8 + 4 + t4], a3 = sm.ys[g8 + 8][sg * 8 + 4 + t4];
                const size_t off = (size_t(yby0) * p.y_WB + ybx0 + sg * 4) * 64 + g8 * 8 + 2 * t4;
                const size_t row1 = size_t(p.y_WB) * 64;
                dct8(a0, a1, a2, a3, 0, yout + off, yout + off + row1, UQ ? yuq + off : nullptr, UQ ? yuq + off + row1 : nullptr,
                     min(4, nby - sg * 4), nrow > 1, &sm.dc[0][0][sg * 4], 2 * kWMcus);
            }
            // chroma: Cb rows as fragment rows 0..7, Cr rows as rows 8..15; 2 segments of 4 block columns
            const uint32_t *cbw = reinterpret_cast<const uint32_t *>(&sm.cb[0][0]), *crw = reinterpret_cast<const uint32_t *>(&sm.cr[0][0]);
            constexpr int kCW = (kWW / 4 + 8) / 2;                     // words per chroma tile row
#pragma unroll 1
            for (int sg = 0; sg < ((p.debug & 2u) ? 0 : 2); sg++) {
                if (sg * 4 >= nm) break;
                const uint32_t a0 = cbw[g8 * kCW + sg * 8 + t4], a1 = crw[g8 * kCW + sg * 8 + t4];
                const uint32_t a2 = cbw[g8 * kCW + sg * 8 + 4 + t4], a3 = crw[g8 * kCW + sg * 8 + 4 + t4];
                const size_t off = size_t(z) * p.c_img_elems + (size_t(my) * p.c_wb + mx0 + sg * 4) * 64 + g8 * 8 + 2 * t4;
                dct8(a0, a1, a2, a3, 1, p.cb + off, p.cr + off, UQ ? p.cbu + off : nullptr, UQ ? p.cru + off : nullptr,
                     min(4, nm - sg * 4), true, nullptr, 0);
            }
        } else {
#pragma unroll 1
            for (int s = 0; s < ((p.debug & 2u) ? 0 : nrow * (kWMcus / 2)); s++) {
                const int by2 = s / (kWMcus / 2), bx = (s - by2 * (kWMcus / 2)) * 4 + q4;
                if ((s - by2 * (kWMcus / 2)) * 4 >= nby) continue;     // warp-uniform: the whole step lies right of the image
                const uint2 v = *reinterpret_cast<const uint2 *>(&sm.ys[by2 * 8 + l8][min(bx, 2 * kWMcus - 1) * 2]);
                int out[8], uq[8];
                block_dct<UQ>(scr, l8, v.x, v.y, qt[0], out, uq);
                if (bx < nby) {
                    if (l8 == 0) { sm.dc[0][by2][bx] = int16_t(out[0]); if (UQ) sm.dc[1][by2][bx] = int16_t(uq[0]); }
                    const size_t off = (size_t(yby0 + by2) * p.y_WB + ybx0 + bx) * 64 + l8;
    #pragma unroll
                    for (int k = 0; k < 8; k++) yout[off + k * 8] = int16_t(out[k]);
                    if (UQ) {
    #pragma unroll
                        for (int k = 0; k < 8; k++) yuq[off + k * 8] = int16_t(uq[k]);
                    }
                }
            }
            // chroma: steps of 4 blocks, Cb then Cr (both quantise with table 1)
    #pragma unroll 1
            for (int s = 0; s < ((p.debug & 2u) ? 0 : 2 * (kWMcus / 4)); s++) {
                const int pl = s / (kWMcus / 4), bx0 = (s - pl * (kWMcus / 4)) * 4, bx = bx0 + q4;
                if (bx0 >= nm) continue;
                const uint16_t(*f)[kWW / 4 + 8] = pl ? sm.cr : sm.cb;
                const uint2 v = *reinterpret_cast<const uint2 *>(&f[l8][bx * 4]);
                int out[8], uq[8];
                block_dct<UQ>(scr, l8, v.x, v.y, qt[1], out, uq);
                if (bx < nm) {
                    const size_t off = size_t(z) * p.c_img_elems + (size_t(my) * p.c_wb + mx0 + bx) * 64 + l8;
                    int16_t *d = (pl ? p.cr : p.cb) + off;
    #pragma unroll
                    for (int k = 0; k < 8; k++) d[k * 8] = int16_t(out[k]);
                    if (UQ) {
                        int16_t *du = (pl ? p.cru : p.cbu) + off;
    #pragma unroll
                        for (int k = 0; k < 8; k++) du[k * 8] = int16_t(uq[k]);
                    }
                }
            }
        }
        __syncwarp();
        if (DERING) {
            // blocks with some but not only samples at 255: redone by one lane each (48 candidates per tile: 32 Y, 8 Cb, 8 Cr)
#pragma unroll 1
            for (int it = 0; it < 2; it++) {
                const int idx = it * 32 + lane;
                bool flagged = false;
                int comp = 0, by2 = 0, bx = 0;              // comp 0 = Y, 1 = Cb, 2 = Cr
                const uint32_t *rows = nullptr;
                int pitch = 0;
                if (idx < 48) {
                    if (idx < 32) { comp = 0; by2 = idx >> 4; bx = idx & 15; rows = &sm.ys[by2 * 8][bx * 2]; pitch = kWW / 4 + 4; }
                    else { comp = 1 + ((idx - 32) >> 3); bx = (idx - 32) & 7; rows = reinterpret_cast<const uint32_t *>(comp == 1 ? &sm.cb[0][0] : &sm.cr[0][0]) + bx * 2; pitch = (kWW / 4 + 8) / 2; }
                    const bool real = comp == 0 ? (bx < nby && by2 < nrow) : bx < nm;
                    if (real) {
                        int cnt = 0;
#pragma unroll
                        for (int r = 0; r < 8; r++) cnt += __popc(__vcmpeq4(rows[r * pitch], 0xffffffffu)) + __popc(__vcmpeq4(rows[r * pitch + 1], 0xffffffffu));
                        flagged = cnt != 0 && cnt != 64 * 8;        // __vcmpeq4 sets all 8 bits of an equal byte
                    }
                }
                if (__any_sync(0xffffffffu, flagged)) {
                    if (flagged) {
                        int d[64];
#pragma unroll
                        for (int r = 0; r < 8; r++) {
                            const uint32_t lo = rows[r * pitch], hi = rows[r * pitch + 1];
#pragma unroll
                            for (int k = 0; k < 4; k++) { d[r * 8 + k] = int((lo >> (8 * k)) & 0xff) - 128; d[r * 8 + 4 + k] = int((hi >> (8 * k)) & 0xff) - 128; }
                        }
                        const int tb = comp == 0 ? 0 : 1;
                        int16_t *qo, *uo = nullptr;
                        if (comp == 0) {
                            const size_t off = (size_t(yby0 + by2) * p.y_WB + ybx0 + bx) * 64;
                            qo = yout + off; if (UQ) uo = yuq + off;
                        } else {
                            const size_t off = size_t(z) * p.c_img_elems + (size_t(my) * p.c_wb + mx0 + bx) * 64;
                            qo = (comp == 1 ? p.cb : p.cr) + off; if (UQ) uo = (comp == 1 ? p.cbu : p.cru) + off;
                        }
                        slow_block<UQ>(d, int(p.dc_quant[tb]), qt[tb], qo, uo);
                        if (comp == 0) { sm.dc[0][by2][bx] = qo[0]; if (UQ) sm.dc[1][by2][bx] = uo[0]; }
                    }
                    __syncwarp();
                }
            }
            __syncwarp();
        }
        // dummy Y blocks (jccoefct.c:286-294 right edge, :301-319 bottom rows; edge tiles only): AC = 0, DC = the DC of
        // the last real block to the left / of the MCU's last block in the row above
        if (nby < 2 * nm || nrow < 2) {
            for (int i = lane; i < 2 * 2 * nm; i += 32) {
                const int by2 = i / (2 * nm), bx = i - by2 * (2 * nm);
                const int gbx = ybx0 + bx, gby = yby0 + by2;
                if (gbx >= int(p.y_WB) || gby >= int(p.y_HB)) continue;
                const bool dummy_right = gbx >= int(p.y_wb), dummy_bottom = gby >= int(p.y_hb);
                if (!dummy_right && !dummy_bottom) continue;
                int sb_x = bx, sb_y = by2;
                if (dummy_bottom) { sb_y = by2 - 1; sb_x = bx | 1; }               // last block of this MCU, row above
                if (ybx0 + sb_x >= int(p.y_wb)) sb_x = int(p.y_wb) - 1 - ybx0;      // which may itself be a right-edge dummy
#pragma unroll
                for (int u = 0; u < (UQ ? 2 : 1); u++) {
                    const uint32_t dc = uint32_t(uint16_t(sm.dc[u][sb_y][sb_x]));
                    uint4 *d = reinterpret_cast<uint4 *>((u ? yuq : yout) + (size_t(gby) * p.y_WB + gbx) * 64);
                    d[0] = make_uint4(dc, 0u, 0u, 0u);
#pragma unroll
                    for (int k = 1; k < 8; k++) d[k] = make_uint4(0u, 0u, 0u, 0u);
                }
            }
        }
        __syncwarp();                    // the sample tiles and dc[] are rewritten by the next tile
    }
}

}  // namespace

cudaError_t launch_jpeg_frontend(const JpegParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.w == 0 || p.h == 0) return cudaSuccess;
    const unsigned mcus_x = p.c_wb, mcus_y = p.c_hb;
    const bool uq = p.yu != nullptr;
    if (uq && (!p.cbu || !p.cru)) return cudaErrorInvalidValue;
    // SLIMG_B200_JPEG_VAR (A/B, 24 MP q = 70 on B200, profiles/r02_k3_variants.txt): stage 2 on the tensor cores -- 8 (default:
    // 5 CTAs/SM, 92 registers, no spills) 63.9 us, 5 (6 CTAs/SM) 65.4, 6 (8 CTAs/SM, spills) 69.7, 9 (4 CTAs/SM) 68.7, 7 = 5 with the
    // stage-1 prefetch; stage 2 on the CUDA cores -- 0 65.1 us, 1 (patch j+1 in flight while patch j is converted) 66.8,
    // 2 (also the next tile's first patch across stage 2) 65.2, 3 / 4 = as 1 with 7 / 6 CTAs/SM 65.5 / 64.2.  The load
    // prefetch buys nothing: the kernel is bound by instruction issue, not by exposed load latency.
    const char *var_env = std::getenv("SLIMG_B200_JPEG_VAR");
    const int var = var_env ? std::atoi(var_env) : 8;
    using Kern = void (*)(const JpegParams, const uint32_t);
    Kern kern;
#define SLIMG_K3_PICK(U)                                                                                                            \
    (var == 0 ? jpeg_frontend_warp_kernel<0, 8, U, false, false> : var == 1 ? jpeg_frontend_warp_kernel<1, 8, U, false, false>         \
     : var == 2 ? jpeg_frontend_warp_kernel<2, 6, U, false, false> : var == 3 ? jpeg_frontend_warp_kernel<1, 7, U, false, false>       \
     : var == 4 ? jpeg_frontend_warp_kernel<1, 6, U, false, false> : var == 6 ? jpeg_frontend_warp_kernel<0, 8, U, true, false>        \
     : var == 7 ? jpeg_frontend_warp_kernel<1, 6, U, true, false> : var == 5 ? jpeg_frontend_warp_kernel<0, 6, U, true, false>         \
     : var == 9 ? jpeg_frontend_warp_kernel<0, 4, U, true, false> : jpeg_frontend_warp_kernel<0, 5, U, true, false>)
    kern = uq ? SLIMG_K3_PICK(true) : SLIMG_K3_PICK(false);
#undef SLIMG_K3_PICK
    if (p.deringing) kern = uq ? jpeg_frontend_warp_kernel<0, 5, true, true, true> : jpeg_frontend_warp_kernel<0, 5, false, true, true>;
    int dev = 0, sms = 148, per_sm = 4;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 32 * kWarpsPerCta, 0) != cudaSuccess || per_sm < 1) per_sm = 4;
    const uint64_t tiles = uint64_t((mcus_x + kWMcus - 1) / kWMcus) * mcus_y * n_images;
    const unsigned grid = unsigned(std::min<uint64_t>((tiles + kWarpsPerCta - 1) / kWarpsPerCta, uint64_t(sms) * per_sm));
    JpegParams q = p;
    const char *dbg = std::getenv("SLIMG_B200_JPEG_DEBUG");
    q.debug = dbg ? uint32_t(std::atoi(dbg)) : 0u;
    kern<<<grid, 32 * kWarpsPerCta, 0, stream>>>(q, n_images);
    count_launch();
    return cudaGetLastError();
}

}  // namespace slimg
