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
// One CTA = kMcus consecutive MCUs (16x16 pixels each) of one MCU row: 4 Y blocks + Cb + Cr per
// MCU.  Stage 1 converts the 16 x (16*kMcus) pixel strip into Y / Cb / Cr sample tiles in shared
// memory; stage 2 runs the row DCTs, stage 3 the column DCTs + quantisation, stage 4 patches the
// dummy blocks and streams the coefficients out 16 bytes per thread.
//
// Roofline: HBM.  Algorithmic bytes per image = w*h*4 read + (Y_HB*Y_WB + 2*C_HB*C_WB)*128 written.
#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kMcus = 8;                        // MCUs per CTA
constexpr int kStripW = 16 * kMcus;             // 128 pixels
constexpr int kThreads = 256;
constexpr int kBlocksY = 4 * kMcus, kBlocksC = kMcus;
constexpr int kBlocks = kBlocksY + 2 * kBlocksC;  // 48 blocks per CTA
constexpr int kWsStride = 72;                   // 64 + 8 ints: column pass is bank-conflict free

#define SLIMG_FIX(x) int((x) * 65536.0 + 0.5)

__device__ __forceinline__ void rgb_to_ycc(uint32_t px, uint8_t &y, uint8_t &cb, uint8_t &cr) {
    const int r = px & 0xff, g = (px >> 8) & 0xff, b = (px >> 16) & 0xff;
    constexpr int half = 1 << 15, cboff = 128 << 16;
    y  = uint8_t((SLIMG_FIX(0.29900) * r + SLIMG_FIX(0.58700) * g + SLIMG_FIX(0.11400) * b + half) >> 16);
    cb = uint8_t((-SLIMG_FIX(0.16874) * r - SLIMG_FIX(0.33126) * g + SLIMG_FIX(0.50000) * b + cboff + half - 1) >> 16);
    cr = uint8_t((SLIMG_FIX(0.50000) * r - SLIMG_FIX(0.41869) * g - SLIMG_FIX(0.08131) * b + cboff + half - 1) >> 16);
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

struct Smem {
    uint8_t ys[16][kStripW];          // Y samples
    uint8_t cbf[16][kStripW];         // full-resolution Cb / Cr (chroma row mapping)
    uint8_t crf[16][kStripW];
    int ws[kBlocks * kWsStride];      // DCT workspace
    int16_t coef[kBlocks][64];        // quantised coefficients
    uint32_t recip[2][64], corr[2][64];
    uint8_t shift[2][64];
};

__global__ void __launch_bounds__(kThreads) jpeg_frontend_kernel(const JpegParams p) {
    __shared__ Smem sm;
    const int tid = threadIdx.x;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(blockIdx.z) * p.src_img_bytes);
    const int mx0 = blockIdx.x * kMcus, my = blockIdx.y;
    const int x0 = mx0 * 16, y0 = my * 16;
    const int w = int(p.w), h = int(p.h);
    const int creal = (h + 1) / 2;   // chroma rows that come from real downsampling

    for (int i = tid; i < 128; i += kThreads) {
        sm.recip[0][i] = p.q.recip[0][i]; sm.corr[0][i] = p.q.corr[0][i]; sm.shift[0][i] = p.q.shift[0][i];
    }

    // stage 1: colour conversion with edge replication.  Y uses row min(y, h-1).  Chroma row cy
    // past creal-1 repeats the last DOWNSAMPLED row (jcprepct.c:174-183), so its two source
    // rows are those of row creal-1, which differ from the Y mapping when h is even.
    for (int i = tid; i < 16 * kStripW; i += kThreads) {
        const int r = i / kStripW, c = i - r * kStripW;
        const int sx = min(x0 + c, w - 1);
        const int yy = y0 + r;
        const int sy = min(yy, h - 1);
        const int cyy = min(yy >> 1, creal - 1);
        const int scy = min(2 * cyy + (yy & 1), h - 1);
        uint8_t yv, cb, cr;
        rgb_to_ycc(__ldg(src + size_t(sy) * w + sx), yv, cb, cr);
        sm.ys[r][c] = yv;
        if (scy != sy) {
            uint8_t y2;
            rgb_to_ycc(__ldg(src + size_t(scy) * w + sx), y2, cb, cr);
        }
        sm.cbf[r][c] = cb; sm.crf[r][c] = cr;
    }
    __syncthreads();

    // stage 2: row DCTs.  Block order in the CTA: Y blocks [by2][bx] (by2 = 0,1; bx = 0..2*kMcus-1),
    // then Cb blocks, then Cr blocks.  Chroma rows are downsampled on the fly (h2v2, bias 1,2,..).
    for (int i = tid; i < kBlocks * 8; i += kThreads) {
        const int b = i >> 3, r = i & 7;
        int in[8], out[8];
        if (b < kBlocksY) {
            const int by2 = b / (2 * kMcus), bx = b - by2 * (2 * kMcus);
            const uint8_t *s = &sm.ys[by2 * 8 + r][bx * 8];
#pragma unroll
            for (int k = 0; k < 8; k++) in[k] = int(s[k]) - 128;
        } else {
            const bool is_cr = b >= kBlocksY + kBlocksC;
            const int bx = b - kBlocksY - (is_cr ? kBlocksC : 0);
            const uint8_t(*f)[kStripW] = is_cr ? sm.crf : sm.cbf;
            const uint8_t *s0 = &f[2 * r][bx * 16], *s1 = &f[2 * r + 1][bx * 16];
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const int bias = 1 + (k & 1);
                in[k] = ((int(s0[2 * k]) + int(s0[2 * k + 1]) + int(s1[2 * k]) + int(s1[2 * k + 1]) + bias) >> 2) - 128;
            }
        }
        fdct8<1>(in, out);
        int *d = &sm.ws[b * kWsStride + r * 8];
#pragma unroll
        for (int k = 0; k < 8; k++) d[k] = out[k];
    }
    __syncthreads();

    // stage 3: column DCTs + quantisation (jcdctmgr.c:408-438, reciprocal form)
    for (int i = tid; i < kBlocks * 8; i += kThreads) {
        const int b = i >> 3, c = i & 7;
        const int tbl = b < kBlocksY ? 0 : 1;
        int in[8], out[8];
        const int *s = &sm.ws[b * kWsStride + c];
#pragma unroll
        for (int k = 0; k < 8; k++) in[k] = s[k * 8];
        fdct8<2>(in, out);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int idx = k * 8 + c;
            const int t = out[k];
            const uint32_t a = uint32_t(t < 0 ? -t : t) + sm.corr[tbl][idx];
            const uint32_t q = uint32_t((uint64_t(a) * sm.recip[tbl][idx]) >> sm.shift[tbl][idx]);
            sm.coef[b][idx] = int16_t(t < 0 ? -int(q) : int(q));
        }
    }
    __syncthreads();

    // stage 4a: dummy Y blocks (jccoefct.c:286-294 right edge, :301-319 bottom rows): AC = 0,
    // DC = the DC of the last real block to the left / of the MCU's last block in the row above.
    const int ybx0 = mx0 * 2, yby0 = my * 2;
    for (int i = tid; i < kBlocksY * 64; i += kThreads) {
        const int b = i >> 6, k = i & 63;
        const int by2 = b / (2 * kMcus), bx = b - by2 * (2 * kMcus);
        const int gbx = ybx0 + bx, gby = yby0 + by2;
        if (gbx >= int(p.y_WB)) continue;
        const bool dummy_right = gbx >= int(p.y_wb), dummy_bottom = gby >= int(p.y_hb);
        if (!dummy_right && !dummy_bottom) continue;
        int16_t v = 0;
        if (k == 0) {
            int sb_x = bx, sb_y = by2;
            if (dummy_bottom) { sb_y = by2 - 1; sb_x = bx | 1; }               // last block of this MCU, row above
            if (ybx0 + sb_x >= int(p.y_wb)) sb_x = int(p.y_wb) - 1 - ybx0;      // which may itself be a right-edge dummy
            v = sm.coef[sb_y * (2 * kMcus) + sb_x][0];
        }
        sm.coef[b][k] = v;   // sources are real blocks, never rewritten here: no hazard
    }
    __syncthreads();

    // stage 4b: stream out, 16 bytes per thread; a Y block row of the CTA is one contiguous run
    int16_t *yout = p.y + size_t(blockIdx.z) * p.y_img_elems;
    int16_t *cbout = p.cb + size_t(blockIdx.z) * p.c_img_elems;
    int16_t *crout = p.cr + size_t(blockIdx.z) * p.c_img_elems;
    for (int i = tid; i < kBlocks * 8; i += kThreads) {
        const int b = i >> 3, part = i & 7;
        int16_t *d;
        if (b < kBlocksY) {
            const int by2 = b / (2 * kMcus), bx = b - by2 * (2 * kMcus);
            const int gbx = ybx0 + bx, gby = yby0 + by2;
            if (gbx >= int(p.y_WB) || gby >= int(p.y_HB)) continue;
            d = yout + (size_t(gby) * p.y_WB + gbx) * 64;
        } else {
            const bool is_cr = b >= kBlocksY + kBlocksC;
            const int bx = b - kBlocksY - (is_cr ? kBlocksC : 0);
            const int gbx = mx0 + bx;
            if (gbx >= int(p.c_wb)) continue;
            d = (is_cr ? crout : cbout) + (size_t(my) * p.c_wb + gbx) * 64;
        }
        *reinterpret_cast<uint4 *>(d + part * 8) = *reinterpret_cast<const uint4 *>(&sm.coef[b][part * 8]);
    }
}

}  // namespace

cudaError_t launch_jpeg_frontend(const JpegParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.w == 0 || p.h == 0) return cudaSuccess;
    const unsigned mcus_x = p.c_wb, mcus_y = p.c_hb;
    if (mcus_y > 65535) return cudaErrorInvalidValue;
    for (uint32_t first = 0; first < n_images; first += 65535) {
        JpegParams q = p;
        q.src += size_t(first) * p.src_img_bytes;
        q.y += size_t(first) * p.y_img_elems;
        q.cb += size_t(first) * p.c_img_elems;
        q.cr += size_t(first) * p.c_img_elems;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        jpeg_frontend_kernel<<<dim3((mcus_x + kMcus - 1) / kMcus, mcus_y, nz), kThreads, 0, stream>>>(q);
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace slimg
