// k2_tc.cu -- Lanczos3 downscale with the vertical pass on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// Replaces image 0.25.9's vertical_sample / horizontal_sample pair behind slimg_core::resize::resize
// (crates/slimg-core/src/resize.rs:64-90) with crop / extend fused in as address arithmetic
// (pipeline.rs:71-89); same taps and weights (taps.cpp), quantised to fixed point (tcplan.cpp).
//
// Why this shape.  The mma.sync pipeline (k2_vfirst.cu) is output-stationary: a group of 8 output rows needs its
// whole filter window (105 source rows at ratio 7.5) resident in shared memory, which leaves room for only one
// group's worth of rows in flight -- it runs at the latency of a DRAM round trip per group, not at bandwidth.
// Here the vertical pass is INPUT-stationary: source rows stream through a ring in chunks of 32 rows, each chunk is
// multiplied into the accumulators of every block of 16 output rows it touches, and is released as soon as those
// MMAs retire.  The accumulators (128 bytes of the row x 16 output rows x 2 weight digits, s32) live in TMEM, so
// the whole ring is prefetch depth.
//
//   HBM --cp.async.bulk.tensor.2d (one 128 B x 32 rows box per panel, SWIZZLE_128B)--> ring slot
//   tcgen05.mma.cta_group::1.kind::i8  D[128 bytes of the row][16 output rows] += A * B
//        A = the panel as it lies in the slot: MN-major (the 128 bytes of a row are contiguous), u8
//        B = weight digit tile of (chunk, block) from the chunk record, K-major, s8
//        one elected thread issues; tcgen05.commit releases the slot / publishes a finished block
//   epilogue warps: tcgen05.ld the two digit planes, recombine (exact s32), add bias, >> to the 16-bit
//        intermediate, store its two u8 limbs planar in shared memory ([channel][row][pixel])
//   pass 2 (one warp per 8 output columns): mma.sync u8 x s8 on the limbs, f32 recombination, clamp, round half
//        away from zero, pack RGBA, store to the final canvas position (post remap)
// Nothing but source pixels is read from HBM and nothing but final pixels is written.
#include <cuda.h>

#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kThreads = 448;                // 14 warps
constexpr int kP2Warps = 4;                  // warps 0..3: pass 2 (one per tile of 8 output columns)
constexpr int kEpiWarp0 = 4, kEpiWarps = 8;  // warps 4..11: epilogue, two per TMEM lane quarter
constexpr int kProducerWarp = 12, kMmaWarp = 13;
constexpr int kMaxSlots = 12;
constexpr uint32_t kSpinLimit = 1u << 27;    // a wedged pipeline traps instead of hanging the GPU
// tcgen05 instruction descriptor: D = s32, A = u8 MN-major, B = s8 K-major, M = 128, N = 32 (both weight digits of a
// block's 16 rows in one instruction: the A panel is read from shared memory once)
constexpr uint32_t kIdesc = (2u << 4) | (0u << 7) | (1u << 10) | (1u << 15) | (0u << 16) | ((32u >> 3) << 17) | ((128u >> 4) << 24);

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
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
// ---- tcgen05 ----
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, int (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// shared-memory matrix descriptors (cute/arch/mma_sm100_desc.hpp SmemDescriptor): address, LBO and SBO in 16-byte
// units, version 1, layout type in bits 61..63
__device__ __forceinline__ uint64_t desc_a(uint32_t addr) {        // MN-major, SWIZZLE_128B: 8-row atoms 1024 B apart
    return uint64_t((addr >> 4) & 0x3fffu) | (uint64_t(64) << 16) | (uint64_t(1024 >> 4) << 32) | (uint64_t(1) << 46) | (uint64_t(2) << 61);
}
__device__ __forceinline__ uint64_t desc_b(uint32_t addr) {        // K-major, no swizzle: LBO 128 B (k halves), SBO 256 B (8-row groups)
    return uint64_t((addr >> 4) & 0x3fffu) | (uint64_t(128 >> 4) << 16) | (uint64_t(256 >> 4) << 32) | (uint64_t(1) << 46);
}

__device__ __forceinline__ void imma(int (&d)[4], const uint32_t (&a)[4], const uint2 &b) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}
__device__ __forceinline__ void imma_z(int (&d)[4], const uint32_t (&a)[4], const uint2 &b) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y), "r"(0));
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts8(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
// exact s32 -> f32 for |x| < 2^22 without the quarter-rate I2F
__device__ __forceinline__ float int_to_float(int x) { return __int_as_float(x + 0x4B400000) - 12582912.0f; }
__device__ __forceinline__ uint32_t quantize_channel(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f);      // clamp, then f32::round (half away; value is >= 0)
    return uint32_t(floorf(t + 0.5f));
}

struct SmemTail {
    uint64_t full[kMaxSlots];                // chunk (panels + record) landed
    uint64_t empty[kMaxSlots];               // every MMA that reads the slot has retired
    uint64_t acc_full[4];                    // a block's accumulators are complete
    uint64_t acc_empty[4];                   // the epilogue has drained the buffer
    uint64_t v_full[2];                      // 8 rows of the intermediate are in V[buf]
    uint64_t v_empty[2];                     // every pass-2 warp has read V[buf]
    uint32_t tmem_base;
};

// canvas pixels this CTA owns but does not compute: rows outside the kept rectangle and the margins left / right
// of the kept columns (edge bands / strips extend to the canvas border)
__device__ void fill_region(const TcParams &p, uint32_t *dst, int i0, int i1, int c0, int c1, int warp, int nwarps) {
    const bool first_band = blockIdx.y == 0, last_band = blockIdx.y + 1 == gridDim.y;
    const bool first_strip = blockIdx.x == 0, last_strip = blockIdx.x + 1 == gridDim.x;
    const int fr0 = first_band ? 0 : int(p.foy) + i0, fr1 = last_band ? int(p.fh) : int(p.foy) + i1;
    const int cx0 = first_strip ? 0 : int(p.fox) + c0, cx1 = last_strip ? int(p.fw) : int(p.fox) + c1;
    const int kx0 = int(p.fox), kx1 = int(p.fox + p.pw), ky0 = int(p.foy), ky1 = int(p.foy + p.ph);
    const int lane = threadIdx.x & 31, fw = int(p.fw);
    const uint32_t fill = p.fill_post;
#pragma unroll 1
    for (int fr = fr0 + warp; fr < fr1; fr += nwarps) {
        uint32_t *row = dst + size_t(fr) * fw;
        if (fr < ky0 || fr >= ky1) {
            for (int x = cx0 + lane; x < cx1; x += 32) row[x] = fill;
        } else {
            for (int x = cx0 + lane; x < kx0; x += 32) row[x] = fill;
            for (int x = kx1 + lane; x < cx1; x += 32) row[x] = fill;
        }
    }
}

template <int NCH2>
__global__ void __launch_bounds__(kThreads, 1) resize_tc_kernel(const TcParams p, const __grid_constant__ CUtensorMap tmap) {
    constexpr int NDIG = 2;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t ring = smem_u32(smem_raw);
    const uint32_t crec_bytes = p.crec_words * 4u;
    const uint32_t crec_base = ring + p.slots * p.slot_bytes;
    const uint32_t vbase = crec_base + p.slots * crec_bytes;
    const uint32_t vplane = 32u * p.pitch_v;                   // one limb of one buffer
    const size_t tail_off = size_t(p.slots) * (p.slot_bytes + crec_bytes) + 4u * vplane;
    SmemTail &tail = *reinterpret_cast<SmemTail *>(smem_raw + tail_off);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);

    // ---- this CTA's strip and band (a range of 16-row blocks) -------------------------------------------
    const int strip = blockIdx.x;
    const int col0 = p.strip_col0[strip], col1 = p.strip_col0[strip + 1];
    const int blk0 = int((uint64_t(blockIdx.y) * p.n_blocks) / gridDim.y);
    const int blk1 = int((uint64_t(blockIdx.y + 1) * p.n_blocks) / gridDim.y);
    const int n_blk = blk1 - blk0;
    const int c_lo = n_blk > 0 ? p.blk_c0[blk0] : 0, c_hi = n_blk > 0 ? p.blk_c1[blk1 - 1] : -1;
    const int n_steps = c_hi - c_lo + 1;
    const int i0 = blk0 * 16, i1 = min(int(p.ph), blk1 * 16);
    const int ntile = (col1 - col0 + 7) >> 3;
    const int npan = int(p.n_panels), nbuf = int(p.nbuf), slots = int(p.slots);

    if (tid == 0) {
        for (int s = 0; s < slots; s++) { mbar_init(&tail.full[s], 1); mbar_init(&tail.empty[s], 1); }
        for (int b = 0; b < 4; b++) { mbar_init(&tail.acc_full[b], 1); mbar_init(&tail.acc_empty[b], kEpiWarps); }
        for (int b = 0; b < 2; b++) { mbar_init(&tail.v_full[b], kEpiWarps); mbar_init(&tail.v_empty[b], uint32_t(ntile)); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kMmaWarp) tmem_alloc(&tail.tmem_base, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tail.tmem_base;

    if (warp == kProducerWarp) {
        // ================= producer: one box per panel + the chunk record, per ring slot =================
        const int tx0 = p.strip_src0[strip] >> 1;                                   // 8-byte elements
        const int tr0 = int(blockIdx.z) * int(p.sh) + int(p.ry);
        const uint32_t step_bytes = uint32_t(npan) * 4096u + crec_bytes;
#pragma unroll 1
        for (int s = 0; s < n_steps; s++) {
            const int slot = s % slots;
            if (s >= slots) mbar_wait(&tail.empty[slot], uint32_t(s / slots - 1) & 1u);
            if (lane == 0) mbar_arrive_expect_tx(&tail.full[slot], step_bytes);
            __syncwarp();
            const int c = c_lo + s;
            if (lane < npan)
                tma_load_2d(ring + uint32_t(slot) * p.slot_bytes + uint32_t(lane) * 4096u, &tmap, tx0 + lane * 16, tr0 + c * 32, &tail.full[slot]);
            else if (lane == 31)
                bulk_g2s(crec_base + uint32_t(slot) * crec_bytes, p.crecs + size_t(c) * p.crec_words, crec_bytes, &tail.full[slot]);
        }
    } else if (warp == kMmaWarp) {
        // ================= MMA issue: every (block, panel, digit) of the chunk ============================
        // One elected thread runs the whole loop (tcgen05.mma / commit are single-thread instructions; a loop that
        // is uniform per construction keeps the compiler from wrapping every issue in an election).
        if (elect_one()) {
#pragma unroll 1
            for (int s = 0; s < n_steps; s++) {
                const int slot = s % slots;
                mbar_wait(&tail.full[slot], uint32_t(s / slots) & 1u);
                tc_fence_after();
                const uint32_t *rec = reinterpret_cast<const uint32_t *>(smem_raw + size_t(slots) * p.slot_bytes + size_t(slot) * crec_bytes);
                const int first = int(rec[0]), nb = int(rec[1]);
                const uint32_t flags = rec[2];
                const uint64_t a_desc0 = desc_a(ring + uint32_t(slot) * p.slot_bytes);
                const uint64_t b_desc0 = desc_b(crec_base + uint32_t(slot) * crec_bytes + 64u);
#pragma unroll 1
                for (int i = 0; i < nb; i++) {
                    const int b = first + i;
                    if (b < blk0 || b >= blk1) continue;
                    const int use = (b - blk0) / nbuf, abuf = (b - blk0) - use * nbuf;
                    const bool starts = (flags >> i) & 1u, ends = (flags >> (8 + i)) & 1u;
                    if (starts) {                                // the buffer's previous block must have been drained
                        mbar_wait(&tail.acc_empty[abuf], (uint32_t(use) & 1u) ^ 1u);
                        tc_fence_after();
                    }
                    const uint32_t acc = starts ? 0u : 1u;
                    const uint64_t bd0 = b_desc0 + uint64_t(i) * (2u * 512u >> 4);     // digit tiles are contiguous: N = 32
                    uint64_t ad = a_desc0;
                    uint32_t dcol = tmem + uint32_t(abuf) * 32u;
                    if (!(p.debug & 1u)) {
#pragma unroll 1
                        for (int pn = 0; pn < npan; pn++) {
                            umma_i8(dcol, ad, bd0, acc);
                            ad += 4096u >> 4;
                            dcol += uint32_t(nbuf) * 32u;
                        }
                    }
                    if (ends) umma_commit(&tail.acc_full[abuf]);
                }
                umma_commit(&tail.empty[slot]);                  // the slot is free once everything issued so far retires
            }
        }
        __syncwarp();
    } else if (warp >= kEpiWarp0 && warp < kEpiWarp0 + kEpiWarps) {
        // ================= epilogue: TMEM -> 16-bit intermediate -> V planes ==============================
        const int quarter = warp & 3, half = (warp - kEpiWarp0) >> 2;
        const int per = (npan + 1) >> 1;
        const int pn0 = half * per, pn1 = min(npan, pn0 + per);
        const int ch = lane & 3;
        const int sh1 = p.sh1;
        const uint32_t lane_addr = uint32_t(quarter * 32) << 16;
        // byte address of (row 0 of channel ch, pixel quarter*8 + lane/4 of panel 0): rows pitch_v apart, channel
        // planes skewed by one word
        const uint32_t v0 = vbase + uint32_t(8 * ch) * p.pitch_v + uint32_t(ch) * 4u + uint32_t(quarter * 8 + (lane >> 2));
#pragma unroll 1
        for (int bi = 0; bi < n_blk; bi++) {
            const int use = bi / nbuf, abuf = bi - use * nbuf;
            int bias[16];
#pragma unroll
            for (int r = 0; r < 16; r++) bias[r] = __ldg(p.bias + (size_t(blk0 + bi) * 16 + r) * 4 + ch);
            mbar_wait(&tail.acc_full[abuf], uint32_t(use) & 1u);
            tc_fence_after();
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int gi = 2 * bi + h;
                const uint32_t vb = uint32_t(gi) & 1u;
                mbar_wait(&tail.v_empty[vb], (uint32_t(gi >> 1) & 1u) ^ 1u);
                const uint32_t vrow = v0 + vb * 2u * vplane;
#pragma unroll 1
                for (int pn = pn0; pn < ((p.debug & 8u) ? pn0 : pn1); pn++) {
                    int lo[8], hi[8];
                    const uint32_t taddr = tmem + lane_addr + uint32_t((pn * nbuf + abuf) * 32 + 8 * h);
                    tmem_ld8(taddr, lo);
                    tmem_ld8(taddr + 16u, hi);
                    tmem_ld_wait();
#pragma unroll
                    for (int r = 0; r < 8; r++) {
                        const int x = lo[r] + (hi[r] << 8) + bias[8 * h + r];
                        const uint32_t v = uint32_t(x >> sh1);
                        const uint32_t a = vrow + uint32_t(r) * p.pitch_v + uint32_t(pn) * 32u;
                        sts8(a, v & 0xffu);
                        sts8(a + vplane, v >> 8);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&tail.v_full[vb]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tail.acc_empty[abuf]);
        }
    } else if (warp < ntile) {
        // ================= pass 2: horizontal taps, 4 channels x 8 rows x 8 output columns per warp =======
        const int g = lane >> 2, t = lane & 3;
        const int tile = warp;
        const int kb0 = p.tile_kb0[strip * 8 + tile];
        uint2 bw2[NCH2][NDIG];
#pragma unroll
        for (int c = 0; c < NCH2; c++)
#pragma unroll
            for (int d = 0; d < NDIG; d++)
                bw2[c][d] = __ldg(reinterpret_cast<const uint2 *>(p.b2) + ((size_t(strip * 8 + tile) * NCH2 + c) * NDIG + d) * 32 + lane);
        const int ocol = col0 + tile * 8 + 2 * t;          // this lane's two kept output columns
        float cxr[2][4];                                   // [column ocol+e][channel]
#pragma unroll
        for (int e = 0; e < 2; e++)
#pragma unroll
            for (int c = 0; c < 4; c++) cxr[e][c] = ocol + e < col1 ? p.cx[size_t(ocol + e) * 4 + c] : 0.0f;
        float s2[2][NDIG];
#pragma unroll
        for (int l = 0; l < 2; l++)
#pragma unroll
            for (int d = 0; d < NDIG; d++) s2[l][d] = p.s2[l][d];

#pragma unroll 1
        for (int gi = 0; gi < 2 * n_blk; gi++) {
            const uint32_t buf = uint32_t(gi) & 1u;
            const uint32_t vbuf = vbase + buf * 2u * vplane;
            mbar_wait(&tail.v_full[buf], uint32_t(gi >> 1) & 1u);
            if (p.debug & 2u) { __syncwarp(); if (lane == 0) mbar_arrive(&tail.v_empty[buf]); continue; }

            int acc[2][2][NDIG][4];            // [channel pair][limb][digit]
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {   // channel pairs (0,1) and (2,3)
                const uint32_t r0 = vbuf + uint32_t(16 * mt + g) * p.pitch_v + uint32_t(2 * mt) * 4u + uint32_t(kb0 + t) * 4u;
                const uint32_t r1 = r0 + 8u * p.pitch_v + 4u;
#pragma unroll
                for (int c = 0; c < NCH2; c++) {
#pragma unroll
                    for (int l = 0; l < 2; l++) {
                        uint32_t a[4];
                        const uint32_t o = uint32_t(l) * vplane + uint32_t(c) * 32u;
                        a[0] = lds32(r0 + o); a[1] = lds32(r1 + o); a[2] = lds32(r0 + o + 16u); a[3] = lds32(r1 + o + 16u);
#pragma unroll
                        for (int d = 0; d < NDIG; d++) {
                            if (c == 0) imma_z(acc[mt][l][d], a, bw2[c][d]);
                            else imma(acc[mt][l][d], a, bw2[c][d]);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&tail.v_empty[buf]);       // V[buf] may be overwritten (two groups on)

            uint32_t px[2] = {0u, 0u};         // packed RGBA of columns ocol, ocol+1 of output row g
#pragma unroll
            for (int mt = 0; mt < 2; mt++)
#pragma unroll
                for (int i = 0; i < 4; i++) {  // i = 0,1: channel 2mt, columns +0, +1; i = 2,3: channel 2mt+1
                    float f = cxr[i & 1][2 * mt + (i >> 1)];
#pragma unroll
                    for (int l = 0; l < 2; l++)
#pragma unroll
                        for (int d = 0; d < NDIG; d++) f = fmaf(int_to_float(acc[mt][l][d][i]), s2[l][d], f);
                    px[i & 1] |= quantize_channel(f) << (8 * (2 * mt + (i >> 1)));
                }
            const int i_row = (2 * blk0 + gi) * 8 + g;
            if (i_row < int(p.ph)) {
                uint32_t *o = dst + size_t(p.foy + i_row) * p.fw + p.fox + ocol;
                if (ocol < col1) o[0] = px[0];
                if (ocol + 1 < col1) o[1] = px[1];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem, 512);
    }
    if (p.need_fill) fill_region(p, dst, i0, i1, col0, col1, warp, kThreads / 32);
}

}  // namespace

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn tc_encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return reinterpret_cast<EncodeTiledFn>(f);
    }();
    return fn;
}

// the source batch as one tall 2-D tensor of 8-byte elements; box = one panel of a chunk (128 bytes x 32 rows),
// written with the 128-byte swizzle the MN-major UMMA operand expects
static cudaError_t make_panel_map(const TcParams &p, const uint8_t *src, uint32_t n_images, CUtensorMap &map) {
    EncodeTiledFn enc = tc_encode_tiled_fn();
    if (!enc) return cudaErrorNotSupported;
    const cuuint64_t dims[2] = {cuuint64_t(p.sw) / 2, cuuint64_t(p.sh) * n_images};
    const cuuint64_t strides[1] = {cuuint64_t(p.sw) * 4};
    const cuuint32_t box[2] = {16, 32};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, const_cast<uint8_t *>(src), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

template <int NCH2>
static cudaError_t launch_tc_t(const TcParams &p, uint32_t n_images, cudaStream_t stream) {
    static thread_local int configured_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (configured_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(resize_tc_kernel<NCH2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        configured_dev = dev;
    }
    TcParams q = p;
    for (uint32_t first = 0; first < n_images; first += 65535) {
        q.src = p.src + size_t(first) * p.src_img_bytes;
        q.dst = p.dst + size_t(first) * p.dst_img_bytes;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        CUtensorMap map;
        cudaError_t e = make_panel_map(q, q.src, nz, map);
        if (e != cudaSuccess) return e;
        resize_tc_kernel<NCH2><<<dim3(p.n_strips, p.n_bands, nz), kThreads, p.smem_bytes, stream>>>(q, map);
        count_launch();
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_resize_tc(const TcParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    if (p.n_bands == 0 || p.n_bands > 65535 || p.n_strips == 0 || p.slots > uint32_t(kMaxSlots) || p.nbuf > 4 || p.nbuf < 2 ||
        p.n_panels * p.nbuf * 32 > 512)
        return cudaErrorInvalidValue;
    switch (p.nch2) {
        case 1: return launch_tc_t<1>(p, n_images, stream);
        case 2: return launch_tc_t<2>(p, n_images, stream);
        case 3: return launch_tc_t<3>(p, n_images, stream);
        case 4: return launch_tc_t<4>(p, n_images, stream);
    }
    return cudaErrorInvalidValue;
}

}  // namespace slimg
