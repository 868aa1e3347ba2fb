// k2_fast.cu -- Lanczos3 downscale, the hot kernel of the resize -> crop -> extend pipeline.
//
// Replaces image 0.25.9's horizontal_sample / vertical_sample pair behind
// slimg_core::resize::resize (crates/slimg-core/src/resize.rs:64-90) for ratio >= 1 on both
// axes (BASELINE configs C1, C3, C4a), with crop / extend fused in as address arithmetic
// (pipeline.rs:71-89).  Same taps and weights as the reference (taps.cpp); the passes run in the
// other order (horizontal first), which the +-1 LSB parity bar allows and which maps the
// expensive pass -- the one that touches every source byte -- onto the tensor cores.
//
// Why tensor cores: Lanczos3 with ratio-scaled support costs 6 MACs per source byte per pass.
// At C3's 24 Mpx that is 576 M MACs per image for the first pass; on B200's FP32 pipes
// (128 FMA/clk/SM, measured, profiles/r01_microbench_b200.txt) the pass alone takes longer than
// reading the image from HBM, and u8 -> f32 conversion (I2F: 16/clk/SM) is slower still.
// mma.sync.m16n8k32.u8.s8 runs at 2028 MAC/clk/SM (measured) and takes the bytes as they are.
//
// Horizontal pass (one warp = one tile of 8 output columns x 16 source rows):
//   D[16 rows x 8 outs] += A[16 rows x 32 px] * B[32 px x 8 outs]      per channel, per k-chunk
//   A = source pixels of ONE channel: a 128-bit load brings 4 RGBA pixels, 8 PRMTs transpose
//       them into the four per-channel registers the A fragment wants (k runs along x).
//   B = fixed-point weights, three s8 digits (fastplan.cpp), pre-swizzled to fragment order.
//   s32 accumulation is exact; the three digit planes are recombined once per tile.
// The f32 result rows go to shared memory.
//
// Vertical pass (one thread = one output column x one channel): source rows stream by; output
// row i accumulates in slot i mod 8 while its window is open (at most 8 windows overlap when
// ratio >= 1), and is clamped, rounded half away from zero, packed and stored the moment its last
// row has passed.  Nothing but source pixels is read from HBM and nothing but final pixels is
// written: the f32 intermediate of the reference (38-51 MB per image) never exists.
//
// Grid: x = strips of 64 output columns, y = bands of output rows (+ blocks that write the fill
// colour around / instead of computed rows), z = image.  Taps that land on the virtual canvas'
// fill colour are folded into per-column / per-row constants (fillw_x, fillw_y).
#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = 8;
constexpr int kRows = 16;                    // source rows per step
constexpr int kStripCols = 64;               // output columns per CTA
constexpr int kIStride = kStripCols * 4 + 4; // floats per intermediate row (+16 B: conflict-free STS.128)
constexpr int kMaxFlush = 24;                // output rows that can complete within one 16-row step
constexpr int kFillRows = 32;                // canvas rows per fill block

__device__ __forceinline__ uint4 ldg_v4(const uint8_t *p) {
    uint4 r;
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ void imma_16832(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// 4 RGBA pixels -> the same 4 pixels of R, of G, of B, of A (one register each)
__device__ __forceinline__ void transpose4(const uint4 &q, uint32_t &r, uint32_t &g, uint32_t &b, uint32_t &a) {
    const uint32_t t0 = __byte_perm(q.x, q.y, 0x5140);   // R0 R1 G0 G1
    const uint32_t t1 = __byte_perm(q.z, q.w, 0x5140);   // R2 R3 G2 G3
    const uint32_t t2 = __byte_perm(q.x, q.y, 0x7362);   // B0 B1 A0 A1
    const uint32_t t3 = __byte_perm(q.z, q.w, 0x7362);   // B2 B3 A2 A3
    r = __byte_perm(t0, t1, 0x5410);
    g = __byte_perm(t0, t1, 0x7632);
    b = __byte_perm(t2, t3, 0x5410);
    a = __byte_perm(t2, t3, 0x7632);
}

__device__ __forceinline__ uint32_t quantize_channel(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f);      // clamp, then f32::round (half away; value is >= 0)
    return uint32_t(floorf(t + 0.5f));
}

struct Smem {
    float inter[kRows * kIStride];           // horizontal-pass results of the current 16 rows
    uint8_t outbuf[kMaxFlush][kStripCols][4];
    int out_row[kMaxFlush];
    int ys, ye;
};

// Blocks past the bands write what is not computed: canvas rows outside the kept rectangle
// (fill_post), kept rows whose taps are all fill (fill_pre), and the left / right margins.
__device__ void fill_block(const FastResizeParams &p, uint32_t *dst, int fb) {
    const int r0 = fb * kFillRows, r1 = min(r0 + kFillRows, int(p.fh));
    for (int fr = r0; fr < r1; fr++) {
        const int i = fr - int(p.foy);
        const bool kept = i >= 0 && i < int(p.ph);
        const bool computed = kept && p.row_computed[i];
        uint32_t *row = dst + size_t(fr) * p.fw;
        for (int x = threadIdx.x; x < int(p.fw); x += kThreads) {
            const bool in_cols = x >= int(p.fox) && x < int(p.fox + p.pw);
            if (computed && in_cols) continue;
            row[x] = (kept && in_cols) ? p.fill_pre_packed : p.fill_post;
        }
    }
}

__global__ void __launch_bounds__(kThreads, 2) resize_fast_kernel(const FastResizeParams p) {
    __shared__ __align__(16) Smem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint8_t *src = p.src + size_t(blockIdx.z) * p.src_img_bytes;
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);

    if (blockIdx.y >= p.n_bands) {
        if (blockIdx.x == 0) fill_block(p, dst, int(blockIdx.y - p.n_bands));
        return;
    }

    // ---- this CTA's band of kept output rows and the source rows it needs --------------------
    const int i0 = int((uint64_t(blockIdx.y) * p.ph) / p.n_bands);
    const int i1 = int((uint64_t(blockIdx.y + 1) * p.ph) / p.n_bands);
    if (tid == 0) { sm.ys = 0x7fffffff; sm.ye = 0; }
    __syncthreads();
    {
        int lo = 0x7fffffff, hi = 0;
        for (int i = i0 + tid; i < i1; i += kThreads)
            if (p.row_computed[i]) { lo = min(lo, p.vleft[i]); hi = max(hi, p.vright[i]); }
        if (lo < hi) { atomicMin(&sm.ys, lo); atomicMax(&sm.ye, hi); }
    }
    __syncthreads();
    const int ys = sm.ys, ye = sm.ye;
    if (ys >= ye) return;                      // every row of the band is pure fill

    // ---- horizontal-pass role: warp -> tile of 8 output columns ----------------------------------
    const int g = lane >> 2, t = lane & 3;
    const int tile = blockIdx.x * kWarps + warp;
    const bool has_tile = tile < int(p.n_tiles);
    const int a0 = has_tile ? p.tile_a0[tile] : 0;
    const int nchunks = has_tile ? p.tile_nchunks[tile] : 0;
    const uint2 *bfrag = reinterpret_cast<const uint2 *>(p.bfrag) + (has_tile ? size_t(p.tile_boff[tile]) * 96 : 0) + lane;
    float hfill[2][4];                         // fill colour x weight-on-fill for this lane's two columns
#pragma unroll
    for (int j = 0; j < 2; j++) {
        const float fw = has_tile ? p.fillw_x[tile * 8 + 2 * t + j] : 0.0f;
#pragma unroll
        for (int c = 0; c < 4; c++) hfill[j][c] = p.fill_pre[c] * fw;
    }
    const int max_px = int(p.sw) - 4;
    const size_t pitch = size_t(p.sw) * 4;

    // ---- vertical-pass role: thread -> (output column, channel) -----------------------------------
    const int vcol = tid >> 2, vch = tid & 3;
    const float vfill = p.fill_pre[vch];
    float vacc[8];
#pragma unroll
    for (int s = 0; s < 8; s++) vacc[s] = 0.0f;
    const int col0 = blockIdx.x * kStripCols;     // first kept output column of this strip

    for (int y0 = ys; y0 < ye; y0 += kRows) {
        // ================= horizontal pass: 16 source rows -> sm.inter ==========================
        if (has_tile) {
            int acc[3][4][4];
#pragma unroll
            for (int l = 0; l < 3; l++)
#pragma unroll
                for (int c = 0; c < 4; c++)
#pragma unroll
                    for (int r = 0; r < 4; r++) acc[l][c][r] = 0;

            const int row_a = min(int(p.ry) + y0 + g, int(p.sh) - 1);
            const int row_b = min(int(p.ry) + y0 + g + 8, int(p.sh) - 1);
            const uint8_t *pa = src + size_t(row_a) * pitch, *pb = src + size_t(row_b) * pitch;

#pragma unroll 2
            for (int c = 0; c < nchunks; c++) {
                const int x0 = min(a0 + 32 * c + 4 * t, max_px), x1 = min(a0 + 32 * c + 16 + 4 * t, max_px);
                const uint4 q0 = ldg_v4(pa + size_t(x0) * 4), q1 = ldg_v4(pb + size_t(x0) * 4);
                const uint4 q2 = ldg_v4(pa + size_t(x1) * 4), q3 = ldg_v4(pb + size_t(x1) * 4);
                uint32_t A[4][4];              // [channel][fragment register]
                transpose4(q0, A[0][0], A[1][0], A[2][0], A[3][0]);
                transpose4(q1, A[0][1], A[1][1], A[2][1], A[3][1]);
                transpose4(q2, A[0][2], A[1][2], A[2][2], A[3][2]);
                transpose4(q3, A[0][3], A[1][3], A[2][3], A[3][3]);
#pragma unroll
                for (int l = 0; l < 3; l++) {
                    const uint2 b = __ldg(bfrag + (size_t(c) * 3 + l) * 32);
#pragma unroll
                    for (int ch = 0; ch < 4; ch++) imma_16832(acc[l][ch], A[ch], b.x, b.y);
                }
            }
            // recombine the three digit planes, scale, add the fill constant, store f32
#pragma unroll
            for (int half = 0; half < 2; half++) {        // rows g, g+8
#pragma unroll
                for (int j = 0; j < 2; j++) {             // columns 2t, 2t+1
                    float4 v;
                    float *pv = reinterpret_cast<float *>(&v);
#pragma unroll
                    for (int ch = 0; ch < 4; ch++) {
                        const int r = half * 2 + j;
                        const int x = acc[0][ch][r] + (acc[1][ch][r] << 7) + (acc[2][ch][r] << 14);
                        pv[ch] = fmaf(float(x), p.hscale, hfill[j][ch]);
                    }
                    *reinterpret_cast<float4 *>(&sm.inter[(g + 8 * half) * kIStride + (warp * 8 + 2 * t + j) * 4]) = v;
                }
            }
        }
        __syncthreads();

        // ================= vertical pass: stream the 16 rows through the slots ======================
        int n_flush = 0;
#pragma unroll 4
        for (int r = 0; r < kRows; r++) {
            const VRow *vr = p.vrow + (y0 + r);
            const float4 w0 = __ldg(reinterpret_cast<const float4 *>(vr->w));
            const float4 w1 = __ldg(reinterpret_cast<const float4 *>(vr->w) + 1);
            const uint4 ev = __ldg(reinterpret_cast<const uint4 *>(vr->out));
            const float v = sm.inter[r * kIStride + vcol * 4 + vch];
            vacc[0] = fmaf(v, w0.x, vacc[0]); vacc[1] = fmaf(v, w0.y, vacc[1]);
            vacc[2] = fmaf(v, w0.z, vacc[2]); vacc[3] = fmaf(v, w0.w, vacc[3]);
            vacc[4] = fmaf(v, w1.x, vacc[4]); vacc[5] = fmaf(v, w1.y, vacc[5]);
            vacc[6] = fmaf(v, w1.z, vacc[6]); vacc[7] = fmaf(v, w1.w, vacc[7]);
            if ((ev.x & ev.y & ev.z & ev.w) != 0xffffffffu) {       // some slot completes on this row (uniform)
                const uint32_t e[4] = {ev.x, ev.y, ev.z, ev.w};
#pragma unroll
                for (int s = 0; s < 8; s++) {
                    const int i = int(int16_t((e[s >> 1] >> (16 * (s & 1))) & 0xffffu));
                    if (i >= 0) {
                        if (i >= i0 && i < i1 && n_flush < kMaxFlush) {
                            const float val = fmaf(vfill, p.fillw_y[i], vacc[s]);
                            sm.outbuf[n_flush][vcol][vch] = uint8_t(quantize_channel(val));
                            if (tid == 0) sm.out_row[n_flush] = i;
                            n_flush++;
                        }
                        vacc[s] = 0.0f;
                    }
                }
            }
        }
        __syncthreads();

        // ================= store the output rows completed in this step ============================
        for (int idx = tid; idx < n_flush * kStripCols; idx += kThreads) {
            const int k = idx >> 6, c = idx & 63;
            if (col0 + c < int(p.pw)) {
                const uint32_t px = *reinterpret_cast<const uint32_t *>(&sm.outbuf[k][c][0]);
                dst[size_t(p.foy + sm.out_row[k]) * p.fw + p.fox + col0 + c] = px;
            }
        }
        // (the next step's first __syncthreads orders these reads before outbuf is rewritten;
        //  sm.inter is rewritten only after it as well)
    }
}

}  // namespace

cudaError_t launch_resize_fast(const FastResizeParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    const unsigned gy = p.n_bands + p.n_fill_blocks;
    if (gy > 65535 || p.n_strips == 0) return cudaErrorInvalidValue;
    for (uint32_t first = 0; first < n_images; first += 65535) {
        FastResizeParams q = p;
        q.src += size_t(first) * p.src_img_bytes;
        q.dst += size_t(first) * p.dst_img_bytes;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        resize_fast_kernel<<<dim3(p.n_strips, gy, nz), kThreads, 0, stream>>>(q);
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace slimg
