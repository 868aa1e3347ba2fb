// k2_vfirst.cu -- Lanczos3 downscale as two integer tensor-core passes; the hot kernel of the
// resize -> crop -> extend pipeline (BASELINE configs C1, C3, C4a).
//
// Replaces image 0.25.9's vertical_sample / horizontal_sample pair behind
// slimg_core::resize::resize (crates/slimg-core/src/resize.rs:64-90), with crop / extend fused in
// as address arithmetic (pipeline.rs:71-89).  Same taps, same pass order (vertical first) and the
// same weights as the reference (taps.cpp), quantised to fixed point (vplan.cpp).
//
// Why tensor cores: Lanczos3 with ratio-scaled support costs 6 MACs per source byte in the first
// pass -- 576 M MACs for a 24 Mpx image -- which on the FP32 pipes (128 FMA/clk/SM, plus a u8->f32
// conversion per byte) takes longer than streaming the image from HBM.  mma.sync.m16n8k32.u8.s8
// takes the bytes as they are at 2028 MAC/clk/SM (profiles/r01_microbench_b200.txt).
//
// Data flow of one CTA (a strip of <= 64 output columns x a band of output rows of one image):
//   HBM --cp.async.bulk (TMA engine, one copy per source row)--> ring of source rows in shared memory
//   pass 1, per group of 8 output rows:  D1[16 bytes of the row][8 rows] += A[16][32 rows] * B[32][8]
//        A comes straight out of the ring with ldmatrix.m16n16.x2.trans.b8: the transposing load
//        hands every lane 4 consecutive SOURCE ROWS of one byte column, which is exactly the A
//        fragment of the MMA -- no shuffles, no byte permutes, channels stay interleaved because
//        every channel of every pixel uses the same vertical weights.
//        B = weight digits (s8), register resident for the whole group.
//        epilogue: recombine the digit planes (exact s32), add fill/offset/rounding, >> to a u16 with
//        7 fractional bits, split into two u8 limbs and store them planar ([channel][row][pixel]).
//   pass 2, per group: D2[(channel,row) 16][8 output columns] += A[16][32 px] * B[32][8] for both limbs
//        epilogue: s32 -> f32 recombination, clamp, round half away from zero, pack RGBA, stage the
//        8 x strip output rows in shared memory and write them out as whole rows.
// Nothing but source pixels is read from HBM and nothing but final pixels is written: the f32
// intermediate of the reference (38-51 MB per image) never exists.
//
// Grid: x = strips, y = bands of output-row groups, z = image.  512 threads, one CTA per SM.
#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kWarps = 16;
constexpr int kThreads = kWarps * 32;
constexpr int kStepRows = 16;                // source rows per ring step (one mbarrier each)
constexpr int kMaxSteps = 16;                // ring of at most 256 rows
constexpr int kStageCols = 64;
constexpr uint32_t kSpinLimit = 1u << 27;    // a wedged pipeline traps instead of hanging the GPU

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
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
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// 32 source rows x 16 bytes, transposed: lane (g,t) gets rows 4t..4t+3 (+16) of byte columns g, g+8
__device__ __forceinline__ void ldsm_t(uint32_t (&a)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];"
                 : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]) : "r"(addr));
}
__device__ __forceinline__ void imma(int (&d)[4], const uint32_t (&a)[4], const uint2 &b) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void sts16(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(uint16_t(v)) : "memory");
}

// exact s32 -> f32 for |x| < 2^22 without the quarter-rate I2F
__device__ __forceinline__ float int_to_float(int x) { return __int_as_float(x + 0x4B400000) - 12582912.0f; }

__device__ __forceinline__ uint32_t quantize_channel(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f);      // clamp, then f32::round (half away; value is >= 0)
    return uint32_t(floorf(t + 0.5f));
}

struct SmemTail {                             // after the ring and the V planes
    uint32_t stage[2][8][kStageCols];        // packed output rows of the last two groups
    uint64_t full[kMaxSteps];
};

// What is not computed: canvas rows outside the kept rectangle (fill_post), kept rows whose taps
// are all fill (fill_pre), and the margins left / right of the kept columns.  Each CTA owns the
// canvas rows of its band and the canvas columns of its strip (edge bands / strips extend to the
// canvas border).
__device__ __noinline__ void fill_region(const VfParams &p, uint32_t *dst, int i0, int i1, int c0, int c1, int tid) {
    const bool first_band = blockIdx.y == 0, last_band = blockIdx.y + 1 == gridDim.y;
    const bool first_strip = blockIdx.x == 0, last_strip = blockIdx.x + 1 == gridDim.x;
    const int fr0 = first_band ? 0 : int(p.foy) + i0, fr1 = last_band ? int(p.fh) : int(p.foy) + i1;
    const int cx0 = first_strip ? 0 : int(p.fox) + c0, cx1 = last_strip ? int(p.fw) : int(p.fox) + c1;
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

template <int NDIG, int NCH1, int NCH2>
__global__ void __launch_bounds__(kThreads, 1) resize_vfirst_kernel(const VfParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const uint32_t ring = smem_u32(smem_raw);
    const uint32_t vbase = ring + p.ring_rows * p.pitch;
    const uint32_t vplane = 32u * p.pitch_v;                   // one limb of one buffer
    SmemTail &tail = *reinterpret_cast<SmemTail *>(smem_raw + size_t(p.ring_rows) * p.pitch + 4u * vplane);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const uint8_t *src = p.src + size_t(blockIdx.z) * p.src_img_bytes;
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);

    // ---- this CTA's strip and band ---------------------------------------------------------------
    const int strip = blockIdx.x;
    const int col0 = p.strip_col0[strip], col1 = p.strip_col0[strip + 1];
    const int nblk = p.strip_nblk[strip];
    const uint32_t row_bytes = uint32_t(p.strip_bytes[strip]);
    const uint8_t *src_col = src + size_t(p.strip_src0[strip]) * 4;
    const int j0 = int((uint64_t(blockIdx.y) * p.n_groups) / gridDim.y);
    const int j1 = int((uint64_t(blockIdx.y + 1) * p.n_groups) / gridDim.y);
    const int i0 = j0 * 8, i1 = min(j1 * 8, int(p.ph));

    // first / last source row of the band (uniform; a handful of loads from L2)
    int ys = 0, ye = 0;
    {
        bool any = false;
        for (int j = j0; j < j1; j++) {
            if (p.grp_nch[j] == 0) continue;
            if (!any) { ys = p.grp_a0[j] & ~(kStepRows - 1); any = true; }
            ye = p.grp_need[j];
        }
    }
    const int n_steps = (ye - ys + kStepRows - 1) / kStepRows;
    const int nst = int(p.ring_rows) / kStepRows;

    if (tid == 0) {
        for (int s = 0; s < nst; s++) mbar_init(&tail.full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // ---- producer: warp 0 drives the TMA engine, one linear bulk copy per source row ----------------
    int next_issue = 0;                       // warp 0 only
    auto issue_step = [&](int s) {            // all lanes of warp 0
        const int slot = s % nst;
        if (lane == 0) mbar_arrive_expect_tx(&tail.full[slot], kStepRows * row_bytes);
        __syncwarp();
        if (lane < kStepRows) {
            const int y = ys + s * kStepRows + lane;
            const int row = min(int(p.ry) + y, int(p.sh) - 1);           // rows past the end carry zero weights
            bulk_g2s(ring + uint32_t(slot * kStepRows + lane) * p.pitch, src_col + size_t(row) * p.sw * 4, row_bytes,
                     &tail.full[slot]);
        }
    };
    if (warp == 0)
        for (; next_issue < min(n_steps, nst); next_issue++) issue_step(next_issue);

    // ---- pass-2 role: warp -> (channel pair, tile of 8 output columns); weights stay in registers ------
    const int mt = warp >> 3, tile = warp & 7;
    const int ntile = (col1 - col0 + 7) >> 3;
    const bool p2_active = tile < ntile;
    const int nch2 = p2_active ? p.tile_nch[strip * 8 + tile] : 0;
    const int kb0 = p2_active ? p.tile_kb0[strip * 8 + tile] : 0;
    uint2 bw2[NCH2][NDIG];
#pragma unroll
    for (int c = 0; c < NCH2; c++)
#pragma unroll
        for (int d = 0; d < NDIG; d++)
            bw2[c][d] = c < nch2 ? __ldg(reinterpret_cast<const uint2 *>(p.b2) +
                                         ((size_t(strip * 8 + tile) * NCH2 + c) * NDIG + d) * 32 + lane)
                                 : make_uint2(0u, 0u);
    float cxr[2][2];                           // [column 2t+e][channel 2mt+h]
#pragma unroll
    for (int e = 0; e < 2; e++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int col = col0 + tile * 8 + 2 * t + e;
            cxr[e][h] = (p2_active && col < col1) ? p.cx[size_t(col) * 4 + 2 * mt + h] : 0.0f;
        }

    int jprev = -1;                            // previous non-empty group (its output rows are still staged)
    uint32_t it = 0;                           // non-empty groups processed so far: parity picks the V / stage buffers

    auto copy_out = [&](int j, uint32_t buf) { // staged rows of group j -> global, whole rows
        const int r = warp & 7, i = j * 8 + r;
        if (i < int(p.ph) && p.row_computed[i]) {
            uint32_t *row = dst + size_t(p.foy + i) * p.fw + p.fox + col0;
            const uint32_t *st = tail.stage[buf][r];
            for (int c = lane + 32 * (warp >> 3); c < col1 - col0; c += 64) row[c] = st[c];
        }
    };

#pragma unroll 1
    for (int j = j0; j < j1; j++) {
        const int nch1 = p.grp_nch[j];
        if (nch1 == 0) continue;               // uniform: every tap of these rows is fill
        const int a0 = p.grp_a0[j];
        const uint32_t vbuf = vbase + (it & 1u) * 2u * vplane;

        // ---- wait for the source rows of this group (step s lives in slot s % nst, phase s / nst) ----
        {
            const int need = (p.grp_need[j] - ys + kStepRows - 1) / kStepRows;
            for (int s = (a0 - ys) / kStepRows; s < need; s++) mbar_wait(&tail.full[s % nst], uint32_t(s / nst) & 1u);
        }

        // ================= pass 1: vertical taps of 8 output rows over the strip =======================
        {
            uint2 bw1[NCH1][NDIG];
            uint32_t rowoff[NCH1];
#pragma unroll
            for (int c = 0; c < NCH1; c++) {
#pragma unroll
                for (int d = 0; d < NDIG; d++)
                    bw1[c][d] = c < nch1 ? __ldg(reinterpret_cast<const uint2 *>(p.b1) +
                                                 ((size_t(j) * NCH1 + c) * NDIG + d) * 32 + lane)
                                         : make_uint2(0u, 0u);
                rowoff[c] = ring + uint32_t((a0 - ys + 32 * c + lane) % int(p.ring_rows)) * p.pitch;
            }
            const int ch = g & 3, hh = g >> 2;
            const int bias0 = p.bias[(size_t(j) * 8 + 2 * t) * 4 + ch], bias1 = p.bias[(size_t(j) * 8 + 2 * t + 1) * 4 + ch];
            // V word address of (row 8*ch + 2t + hh, block 0): rows are pitch_v apart, channel planes skewed by one word
            const uint32_t vrow = vbuf + uint32_t(8 * ch + 2 * t + hh) * p.pitch_v + uint32_t(ch) * 4u;

#pragma unroll 1
            for (int blk = warp; blk < nblk; blk += kWarps) {
                int acc[NDIG][4];
#pragma unroll
                for (int d = 0; d < NDIG; d++)
#pragma unroll
                    for (int i = 0; i < 4; i++) acc[d][i] = 0;
#pragma unroll
                for (int c = 0; c < NCH1; c++) {
                    if (c < nch1) {
                        uint32_t a[4];
                        ldsm_t(a, rowoff[c] + uint32_t(blk) * 16u);
#pragma unroll
                        for (int d = 0; d < NDIG; d++) imma(acc[d], a, bw1[c][d]);
                    }
                }
                // recombine the digit planes (exact), re-quantise to u16, split into limbs
                int v[4];
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    int x = acc[0][i] + (acc[1][i] << 8);
                    if (NDIG == 3) x += acc[NDIG - 1][i] << 16;
                    v[i] = (x + ((i & 1) ? bias1 : bias0)) >> p.sh1;
                }
                // v[0],v[2]: pixel hh / hh+2 of output row 2t; v[1],v[3]: same pixels of row 2t+1
                const uint32_t w0 = __byte_perm(uint32_t(v[0]), uint32_t(v[2]), 0x5140);   // lo lo hi hi of row 2t
                const uint32_t w1 = __byte_perm(uint32_t(v[1]), uint32_t(v[3]), 0x5140);   // ... of row 2t+1
                const uint32_t send = hh ? w0 : w1, mine = hh ? w1 : w0;
                const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 16);
                const uint32_t pa = hh ? recv : mine, pb = hh ? mine : recv;              // pixels (0,2) and (1,3)
                sts32(vrow + uint32_t(blk) * 4u, __byte_perm(pa, pb, 0x5410));             // low limbs
                sts32(vrow + vplane + uint32_t(blk) * 4u, __byte_perm(pa, pb, 0x7632));    // high limbs
            }
        }
        __syncthreads();      // V[j&1] complete; every warp is done with the ring rows below the next band
                              // and with pass 2 of the previous group (its staged rows can be written out)

        // ---- producer: refill the ring steps no later group needs ------------------------------------
        if (warp == 0) {
            int a_next = ye;
            for (int jj = j + 1; jj < j1; jj++)
                if (p.grp_nch[jj]) { a_next = p.grp_a0[jj]; break; }
            while (next_issue < n_steps && ys + (next_issue - nst + 1) * kStepRows <= a_next) issue_step(next_issue++);
        }
        if (jprev >= 0) copy_out(jprev, (it & 1u) ^ 1u);

        // ================= pass 2: horizontal taps, 2 channels x 8 rows x 8 output columns per warp =======
        if (p2_active) {
            int acc[2][NDIG][4];
#pragma unroll
            for (int l = 0; l < 2; l++)
#pragma unroll
                for (int d = 0; d < NDIG; d++)
#pragma unroll
                    for (int i = 0; i < 4; i++) acc[l][d][i] = 0;
            const uint32_t r0 = vbuf + uint32_t(16 * mt + g) * p.pitch_v + uint32_t(2 * mt) * 4u + uint32_t(kb0 + t) * 4u;
            const uint32_t r1 = r0 + 8u * p.pitch_v + 4u;
#pragma unroll
            for (int c = 0; c < NCH2; c++) {
                if (c < nch2) {
#pragma unroll
                    for (int l = 0; l < 2; l++) {
                        uint32_t a[4];
                        const uint32_t o = uint32_t(l) * vplane + uint32_t(c) * 32u;
                        a[0] = lds32(r0 + o); a[1] = lds32(r1 + o); a[2] = lds32(r0 + o + 16u); a[3] = lds32(r1 + o + 16u);
#pragma unroll
                        for (int d = 0; d < NDIG; d++) imma(acc[l][d], a, bw2[c][d]);
                    }
                }
            }
            // i = 0,1: channel 2mt, columns 2t, 2t+1; i = 2,3: channel 2mt+1
            uint32_t q[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                float f = cxr[i & 1][i >> 1];
#pragma unroll
                for (int l = 0; l < 2; l++)
#pragma unroll
                    for (int d = 0; d < NDIG; d++) f = fmaf(int_to_float(acc[l][d][i]), p.s2[l][d], f);
                q[i] = quantize_channel(f);
            }
            uint8_t *st = reinterpret_cast<uint8_t *>(&tail.stage[it & 1u][g][tile * 8 + 2 * t]) + 2 * mt;
            sts16(smem_u32(st), q[0] | (q[2] << 8));
            sts16(smem_u32(st + 4), q[1] | (q[3] << 8));
        }
        jprev = j;
        it++;
    }
    __syncthreads();
    if (jprev >= 0) copy_out(jprev, (it & 1u) ^ 1u);

    if (p.need_fill) fill_region(p, dst, i0, i1, col0, col1, tid);
}

}  // namespace

template <int NDIG, int NCH1, int NCH2>
static cudaError_t launch_vf_t(const VfParams &p, uint32_t n_images, cudaStream_t stream) {
    static thread_local int configured_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (configured_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(resize_vfirst_kernel<NDIG, NCH1, NCH2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             227 * 1024);
        if (e != cudaSuccess) return e;
        configured_dev = dev;
    }
    VfParams q = p;
    for (uint32_t first = 0; first < n_images; first += 65535) {
        q.src = p.src + size_t(first) * p.src_img_bytes;
        q.dst = p.dst + size_t(first) * p.dst_img_bytes;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        resize_vfirst_kernel<NDIG, NCH1, NCH2><<<dim3(p.n_strips, p.n_bands, nz), kThreads, p.smem_bytes, stream>>>(q);
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

size_t vfirst_smem_tail_bytes() { return sizeof(SmemTail); }

cudaError_t launch_resize_vfirst(const VfParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    if (p.n_bands == 0 || p.n_bands > 65535 || p.n_strips == 0 || p.nch1 > 4 || p.nch2 > 4) return cudaErrorInvalidValue;
    if (p.ring_rows % kStepRows != 0 || p.ring_rows / kStepRows > kMaxSteps) return cudaErrorInvalidValue;
    if (p.ndig == 2) return launch_vf_t<2, 4, 4>(p, n_images, stream);
    return launch_vf_t<3, 4, 4>(p, n_images, stream);
}

}  // namespace slimg
