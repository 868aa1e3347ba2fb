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
// Staging: one producer warp streams the CTA's strip of source rows into a ring of shared-memory
// stages with cp.async.bulk (the TMA engine's linear copy, UBLKCP in SASS) completing on
// mbarriers; eight consumer warps never wait on global memory.
//
// Horizontal pass (one warp = one tile of 8 output columns x 16 source rows):
//   D[16 rows x 8 outs] += A[16 rows x 32 px] * B[32 px x 8 outs]      per channel, per k-chunk
//   A = source pixels of ONE channel: a 128-bit load brings 4 RGBA pixels, 8 PRMTs transpose
//       them into the four per-channel registers the A fragment wants (k runs along x).
//   B = fixed-point weights, three s8 digits (fastplan.cpp), pre-swizzled to fragment order.
//   s32 accumulation is exact; the three digit planes are recombined once per tile.
// The f32 result rows go to shared memory.
//
// Vertical pass (one thread = one output column x two slots, all four channels, packed f32x2
// FMAs): source rows stream by; output row i accumulates in slot i mod 8 while its window is open
// (at most 8 windows overlap when ratio >= 1), and is clamped, rounded half away from zero, packed
// and stored the moment its last row has passed.  Nothing but source pixels is read from HBM and nothing but final pixels is
// written: the f32 intermediate of the reference (38-51 MB per image) never exists.
//
// Grid: x = strips of 64 output columns, y = bands of output rows (+ blocks that write the fill
// colour around / instead of computed rows), z = image.  Taps that land on the virtual canvas'
// fill colour are folded into per-column / per-row constants (fillw_x, fillw_y).
#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int kRows = 16;                    // source rows per stage
constexpr int kStripCols = 64;               // output columns per CTA
constexpr int kIStride = kStripCols * 4 + 4; // floats per intermediate row (+16 B: conflict-free STS.128)
constexpr int kMaxFlush = 24;                // output rows that can complete within one stage
constexpr int kVTabBytes = (kRows / 4) * int(sizeof(VQuad));
constexpr uint32_t kSpinLimit = 1u << 27;    // a wedged pipeline traps instead of hanging the GPU

static_assert(sizeof(VQuad) == 336 && kVTabBytes % 16 == 0, "VQuad layout");

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok = 0, spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (!ok && ++spins > kSpinLimit) asm volatile("trap;");
    } while (!ok);
}
// TMA engine, linear form: global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void imma_16832(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// first k-chunk: C = 0 comes from the zero register, no accumulator clearing
__device__ __forceinline__ void imma_16832_z(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "r"(0));
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

// two f32 lanes per instruction (FFMA2): d = a * b + d
__device__ __forceinline__ void fma2(float2 &d, const float2 &a, const float2 &b) {
    unsigned long long dd = *reinterpret_cast<unsigned long long *>(&d);
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(dd)
        : "l"(*reinterpret_cast<const unsigned long long *>(&a)), "l"(*reinterpret_cast<const unsigned long long *>(&b)));
    d = *reinterpret_cast<float2 *>(&dd);
}

__device__ __forceinline__ uint32_t quantize_channel(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f);      // clamp, then f32::round (half away; value is >= 0)
    return uint32_t(floorf(t + 0.5f));
}

struct SmemFixed {                            // lives after the stages in dynamic shared memory
    float inter[kRows * kIStride];           // horizontal-pass results of the current 16 rows (8 columns per warp)
    uint32_t outbuf[kMaxFlush][kStripCols];  // packed output rows completed in this stage
    int out_row[kWarps][kMaxFlush];
    int ys, ye;
    int done_cnt[4];                         // warps finished with each stage
    uint64_t full[4];
};

// What is not computed: canvas rows outside the kept rectangle (fill_post), kept rows whose taps
// are all fill (fill_pre), and the margins left / right of the kept columns.  Each CTA owns the
// canvas rows of its band and the canvas columns of its strip (edge bands / strips extend to the
// canvas border).
__device__ __noinline__ void fill_region(const FastResizeParams &p, uint32_t *dst, int i0, int i1, int tid) {
    const bool first_band = blockIdx.y == 0, last_band = blockIdx.y + 1 == p.n_bands;
    const bool first_strip = blockIdx.x == 0, last_strip = blockIdx.x + 1 == p.n_strips;
    const int fr0 = first_band ? 0 : int(p.foy) + i0, fr1 = last_band ? int(p.fh) : int(p.foy) + i1;
    const int cx0 = first_strip ? 0 : int(p.fox) + int(blockIdx.x) * kStripCols;
    const int cx1 = last_strip ? int(p.fw) : int(p.fox) + int(blockIdx.x + 1) * kStripCols;
    const int kx0 = int(p.fox), kx1 = int(p.fox + p.pw);
    const bool margins = cx0 < kx0 || kx1 < cx1;
    const int warp = tid >> 5, lane = tid & 31;
#pragma unroll 1
    for (int fr = fr0 + warp; fr < fr1; fr += kWarps) {
        const int i = fr - int(p.foy);
        const bool kept = i >= 0 && i < int(p.ph);
        const bool computed = kept && p.row_computed[i];
        if (computed && !margins) continue;
        uint32_t *row = dst + size_t(fr) * p.fw;
#pragma unroll 1
        for (int x = cx0 + lane; x < cx1; x += 32) {
            const bool in_cols = x >= kx0 && x < kx1;
            if (computed && in_cols) continue;
            row[x] = (kept && in_cols) ? p.fill_pre_packed : p.fill_post;
        }
    }
}

// NCH > 0: every tile of the launch spans at most NCH k-chunks and its weight fragments stay in
// registers for the whole band; NCH == 0: any chunk count, fragments are re-read (L1/L2) per stage.
template <int NCH>
__global__ void __launch_bounds__(kThreads, 2) resize_fast_kernel(const FastResizeParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const uint32_t stage_bytes = kRows * p.pitch + kVTabBytes;     // pixel rows, then the VQuad slice
    SmemFixed &sm = *reinterpret_cast<SmemFixed *>(smem_raw + size_t(p.stages) * stage_bytes);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint8_t *src = p.src + size_t(blockIdx.z) * p.src_img_bytes;
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);

    // ---- this CTA's band of kept output rows and the source rows it needs --------------------
    const int i0 = int((uint64_t(blockIdx.y) * p.ph) / p.n_bands);
    const int i1 = int((uint64_t(blockIdx.y + 1) * p.ph) / p.n_bands);
    if (tid == 0) {
        sm.ys = 0x7fffffff; sm.ye = 0;
        for (uint32_t s = 0; s < p.stages; s++) { mbar_init(&sm.full[s], 1); sm.done_cnt[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    {
        int lo = 0x7fffffff, hi = 0;
        for (int i = i0 + tid; i < i1; i += kThreads)
            if (p.row_computed[i]) { lo = min(lo, p.vleft[i]); hi = max(hi, p.vright[i]); }
        if (lo < hi) { atomicMin(&sm.ys, lo); atomicMax(&sm.ye, hi); }
    }
    __syncthreads();
    const bool any_rows = sm.ys != 0x7fffffff;
    const int ys = any_rows ? (sm.ys & ~3) : 0;          // VQuad slices start on a multiple of 4 rows
    const int n_iter = any_rows ? (sm.ye - ys + kRows - 1) / kRows : 0;

    // ---- staging: warp 0 drives the TMA engine, one linear bulk copy per lane (16 rows + table) ----
    const int strip_col = p.strip_a0[blockIdx.x];
    const uint32_t row_bytes = uint32_t(p.strip_bytes[blockIdx.x]);
    auto issue_stage = [&](int it) {               // called by all lanes of warp 0
        const int s = it % int(p.stages);
        uint8_t *stage = smem_raw + size_t(s) * stage_bytes;
        const int y0 = ys + it * kRows;
        if (lane == 0) mbar_arrive_expect_tx(&sm.full[s], kRows * row_bytes + kVTabBytes);
        __syncwarp();
        if (lane < kRows) {
            const int row = min(int(p.ry) + y0 + lane, int(p.sh) - 1);    // rows past the end carry zero weights
            bulk_g2s(stage + size_t(lane) * p.pitch, src + (size_t(row) * p.sw + strip_col) * 4, row_bytes, &sm.full[s]);
        } else if (lane == kRows) {
            bulk_g2s(stage + size_t(kRows) * p.pitch, p.vquad + (y0 >> 2), kVTabBytes, &sm.full[s]);
        }
    };
    if (warp == 0)
        for (int it = 0; it < min(n_iter, int(p.stages)); it++) issue_stage(it);

    // ---- horizontal-pass role: warp -> tile of 8 output columns ----------------------------------
    const int g = lane >> 2, t = lane & 3;
    const int tile = blockIdx.x * kWarps + warp;
    const bool has_tile = tile < int(p.n_tiles);
    const int tile_off = has_tile ? p.tile_a0[tile] : 0;
    const int nchunks = has_tile ? p.tile_nchunks[tile] : 0;
    const uint2 *bfrag = reinterpret_cast<const uint2 *>(p.bfrag) + (has_tile ? size_t(p.tile_boff[tile]) * 96 : 0) + lane;
    const float fwx0 = has_tile ? p.fillw_x[tile * 8 + 2 * t] : 0.0f, fwx1 = has_tile ? p.fillw_x[tile * 8 + 2 * t + 1] : 0.0f;
    const uint32_t frag_off = uint32_t(g) * p.pitch + uint32_t(tile_off + 4 * t) * 4;
    uint2 bw[NCH > 0 ? NCH : 1][3];
    if (NCH > 0) {
#pragma unroll
        for (int c = 0; c < NCH; c++)
#pragma unroll
            for (int l = 0; l < 3; l++) bw[c][l] = c < nchunks ? __ldg(bfrag + (c * 3 + l) * 32) : make_uint2(0u, 0u);
    }

    // ---- vertical-pass role: thread -> (output column of the warp's own tile, slot pair) ----------
    const int vcol = tid >> 2, sp = tid & 3;
    float2 va01 = {0.f, 0.f}, va23 = {0.f, 0.f}, vb01 = {0.f, 0.f}, vb23 = {0.f, 0.f};   // slots 2sp, 2sp+1
    const int col0 = blockIdx.x * kStripCols;
    int *my_out_row = sm.out_row[warp];

    for (int it = 0; it < n_iter; it++) {
        const int s = it % int(p.stages);
        const uint32_t parity = uint32_t(it / int(p.stages)) & 1u;
        const uint8_t *stage = smem_raw + size_t(s) * stage_bytes;
        mbar_wait(&sm.full[s], parity);

        // ================= horizontal pass: 16 staged rows -> sm.inter ===========================
        if (has_tile) {
            int acc[3][4][4];
            const uint8_t *pa = stage + frag_off, *pb = pa + 8 * p.pitch;
            auto chunk = [&](int c, const uint2 (&b)[3], bool first) {
                const uint4 q0 = *reinterpret_cast<const uint4 *>(pa + 128 * c);
                const uint4 q1 = *reinterpret_cast<const uint4 *>(pb + 128 * c);
                const uint4 q2 = *reinterpret_cast<const uint4 *>(pa + 128 * c + 64);
                const uint4 q3 = *reinterpret_cast<const uint4 *>(pb + 128 * c + 64);
                uint32_t A[4][4];              // [channel][fragment register]
                transpose4(q0, A[0][0], A[1][0], A[2][0], A[3][0]);
                transpose4(q1, A[0][1], A[1][1], A[2][1], A[3][1]);
                transpose4(q2, A[0][2], A[1][2], A[2][2], A[3][2]);
                transpose4(q3, A[0][3], A[1][3], A[2][3], A[3][3]);
#pragma unroll
                for (int l = 0; l < 3; l++)
#pragma unroll
                    for (int ch = 0; ch < 4; ch++) {
                        if (first) imma_16832_z(acc[l][ch], A[ch], b[l].x, b[l].y);
                        else imma_16832(acc[l][ch], A[ch], b[l].x, b[l].y);
                    }
            };
            if (NCH > 0) {
#pragma unroll
                for (int c = 0; c < NCH; c++)
                    if (c == 0 || c < nchunks) chunk(c, bw[c], c == 0);
            } else {
                uint2 b[3];
#pragma unroll
                for (int l = 0; l < 3; l++) b[l] = nchunks > 0 ? __ldg(bfrag + l * 32) : make_uint2(0u, 0u);
                chunk(0, b, true);
#pragma unroll 1
                for (int c = 1; c < nchunks; c++) {
#pragma unroll
                    for (int l = 0; l < 3; l++) b[l] = __ldg(bfrag + (size_t(c) * 3 + l) * 32);
                    chunk(c, b, false);
                }
            }
            // recombine the three digit planes, scale, add the fill constant, store f32
#pragma unroll
            for (int half = 0; half < 2; half++) {        // rows g, g+8
#pragma unroll
                for (int j = 0; j < 2; j++) {             // columns 2t, 2t+1
                    const float fw = j ? fwx1 : fwx0;
                    float4 v;
                    float *pv = reinterpret_cast<float *>(&v);
#pragma unroll
                    for (int ch = 0; ch < 4; ch++) {
                        const int r = half * 2 + j;
                        const int x = acc[0][ch][r] + (acc[1][ch][r] << 7) + (acc[2][ch][r] << 14);
                        pv[ch] = fmaf(float(x), p.hscale, p.fill_pre[ch] * fw);
                    }
                    *reinterpret_cast<float4 *>(&sm.inter[(g + 8 * half) * kIStride + (warp * 8 + 2 * t + j) * 4]) = v;
                }
            }
        }
        __syncwarp();        // the warp's vertical pass reads only the 8 columns the warp itself produced

        // ================= vertical pass: stream the 16 rows through the slots ======================
        const VQuad *vq = reinterpret_cast<const VQuad *>(stage + size_t(kRows) * p.pitch);
        int n_flush = 0;
#pragma unroll
        for (int q = 0; q < kRows / 4; q++) {
            const uint4 done4 = *reinterpret_cast<const uint4 *>(vq[q].done);
            const uint32_t done[4] = {done4.x, done4.y, done4.z, done4.w};
            const bool any_done = (done4.x | done4.y | done4.z | done4.w) != 0;   // uniform
#pragma unroll
            for (int rr = 0; rr < 4; rr++) {
                const int r = q * 4 + rr;
                const float4 v = *reinterpret_cast<const float4 *>(&sm.inter[r * kIStride + vcol * 4]);
                const float4 w = *reinterpret_cast<const float4 *>(&vq[q].w2[rr][sp * 4]);     // wa wa wb wb
                const float2 v01 = {v.x, v.y}, v23 = {v.z, v.w}, wa = {w.x, w.y}, wb = {w.z, w.w};
                fma2(va01, v01, wa); fma2(va23, v23, wa);
                fma2(vb01, v01, wb); fma2(vb23, v23, wb);
                if (any_done && done[rr]) {               // some slot completes on this row (uniform)
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int slot = 2 * sp + h;
                        if (done[rr] & (1u << slot)) {
                            const int i = vq[q].out[rr][slot];
                            const int k = n_flush + __popc(done[rr] & ((1u << slot) - 1u));
                            const bool mine = i >= i0 && i < i1 && k < kMaxFlush;
                            if (mine) {
                                const float fy = p.fillw_y[i];
                                const float2 a = h ? vb01 : va01, b = h ? vb23 : va23;
                                sm.outbuf[k][vcol] = quantize_channel(fmaf(p.fill_pre[0], fy, a.x)) |
                                                     (quantize_channel(fmaf(p.fill_pre[1], fy, a.y)) << 8) |
                                                     (quantize_channel(fmaf(p.fill_pre[2], fy, b.x)) << 16) |
                                                     (quantize_channel(fmaf(p.fill_pre[3], fy, b.y)) << 24);
                            }
                            if ((lane >> 2) == 0 && k < kMaxFlush) my_out_row[k] = mine ? i : -1;
                            if (h) { vb01 = {0.f, 0.f}; vb23 = {0.f, 0.f}; } else { va01 = {0.f, 0.f}; va23 = {0.f, 0.f}; }
                        }
                    }
                    n_flush += __popc(done[rr]);
                }
            }
        }
        __syncwarp();

        // ================= store the output rows this warp completed in this stage ==================
        n_flush = min(n_flush, kMaxFlush);
        for (int idx = lane; idx < n_flush * 8; idx += 32) {
            const int k = idx >> 3, c = warp * 8 + (idx & 7);
            const int i = my_out_row[k];
            if (i >= 0 && col0 + c < int(p.pw)) dst[size_t(p.foy + i) * p.fw + p.fox + col0 + c] = sm.outbuf[k][c];
        }
        // every warp is done with the stage: refill it for iteration it + stages.  (The barrier also
        // orders this warp's reads of inter / outbuf before the next stage rewrites them.)
        __syncthreads();
        if (warp == 0 && it + int(p.stages) < n_iter) issue_stage(it + int(p.stages));
    }

    if (p.n_fill_blocks) fill_region(p, dst, i0, i1, tid);
}

}  // namespace

template <int NCH>
static cudaError_t launch_fast_t(const FastResizeParams &p, uint32_t n_images, cudaStream_t stream) {
    static thread_local int configured_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (configured_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(resize_fast_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(resize_fast_kernel<NCH>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        if (e != cudaSuccess) return e;
        configured_dev = dev;
    }
    FastResizeParams q = p;
    // ring depth: as many stages as fit in half an SM's shared memory (two CTAs per SM), at least two
    const size_t stage_bytes = size_t(kRows) * p.pitch + kVTabBytes;
    size_t stages = (112 * 1024 - sizeof(SmemFixed)) / stage_bytes;
    if (stages < 2) stages = 2;
    if (stages > 4) stages = 4;
    q.stages = uint32_t(stages);
    q.smem_bytes = uint32_t(stages * stage_bytes + sizeof(SmemFixed));
    if (q.smem_bytes > 220 * 1024) return cudaErrorInvalidValue;
    for (uint32_t first = 0; first < n_images; first += 65535) {
        q.src = p.src + size_t(first) * p.src_img_bytes;
        q.dst = p.dst + size_t(first) * p.dst_img_bytes;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        resize_fast_kernel<NCH><<<dim3(p.n_strips, p.n_bands, nz), kThreads, q.smem_bytes, stream>>>(q);
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_resize_fast(const FastResizeParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    if (p.n_bands > 65535 || p.n_strips == 0) return cudaErrorInvalidValue;
    if (p.max_chunks <= 2) return launch_fast_t<2>(p, n_images, stream);
    if (p.max_chunks <= 4) return launch_fast_t<4>(p, n_images, stream);
    return launch_fast_t<0>(p, n_images, stream);
}

}  // namespace slimg
