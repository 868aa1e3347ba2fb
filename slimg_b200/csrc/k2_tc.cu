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
// the whole ring is prefetch depth (and the chunks after it are prefetched as far as L2).
//
//   HBM --cp.async.bulk.tensor.2d (one 128 B x 32 rows box per panel, SWIZZLE_128B)--> ring slot
//   tcgen05.mma.cta_group::1.kind::i8  D[128 bytes of the row][8 output rows x 2 digits] += A * B
//        A = the panel as it lies in the slot: MN-major (the 128 bytes of a row are contiguous), u8
//        B = weight digit tile of (chunk, block) from the chunk record, K-major, s8
//        one elected thread issues; tcgen05.commit releases the slot / publishes a finished block
//   epilogue warps: tcgen05.ld the two digit planes, recombine (exact s32), add bias, >> to the 16-bit
//        intermediate, store its two u8 limbs planar in shared memory ([channel][row][pixel])
//   pass 2 (one warp per 8 output columns): mma.sync u8 x s8 on the limbs, f32 recombination, clamp, round half
//        away from zero, pack RGBA, store to the final canvas position (post remap)
// Nothing but source pixels is read from HBM and nothing but final pixels is written.
#include <cuda.h>

#include <algorithm>
#include <cstdio>

#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kThreads = 448;                // 14 warps
constexpr int kWarps = kThreads / 32;
constexpr int kP2Warps = 4;                  // warps 0..3: pass 2 (one per tile of 8 output columns)
constexpr int kEpiWarp0 = 4, kEpiWarps = 8;  // warps 4..11: epilogue, two per TMEM lane quarter
constexpr int kProducerWarp = 12, kMmaWarp = 13;
constexpr int kMaxSlots = 12;
constexpr int kMaxAcc = 8;                   // accumulator buffers per panel (16 TMEM columns each)
constexpr int kMaxVBufs = 4;
constexpr uint32_t kSpinLimit = 1u << 27;    // a wedged pipeline traps instead of hanging the GPU
// tcgen05 instruction descriptor: D = s32, A = u8 MN-major, B = s8 K-major, M = 128; N = 16 per block (both weight digits
// of its 8 rows), times the number of blocks one instruction covers
constexpr uint32_t kIdescBase = (2u << 4) | (0u << 7) | (1u << 10) | (1u << 15) | (0u << 16) | ((128u >> 4) << 24);     // N goes to bits 17..22

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
// instrumented wait (debug & 32): cycles spent in the wait are added to `acc`
__device__ __forceinline__ void mbar_wait_t(uint64_t *bar, uint32_t parity, bool timed, unsigned long long &acc) {
    if (!timed) { mbar_wait(bar, parity); return; }
    const long long t0 = clock64();
    mbar_wait(bar, parity);
    acc += (unsigned long long)(clock64() - t0);
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
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, int (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
}
// same box as tma_load_2d, only as far as L2: the DRAM part of the latency is paid before the chunk gets a ring slot
// all panels of a chunk in ONE instruction: a 3-D view of the source (x, row, panel) whose third dimension steps 128 bytes
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap *map, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];" ::"l"(map), "r"(c0), "r"(c1) : "memory");
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
    t = fminf(fmaxf(t, 0.0f), 255.0f) + 0.5f;      // clamp, then f32::round (half away; value is >= 0) = floor(t + 0.5):
    return __float_as_uint(__fadd_rz(t, 8388608.0f)) & 0xffu;      // 2^23 + floor(t) by truncation, no FRND / F2I
}

struct SmemTail {
    uint64_t full[kMaxSlots];                // chunk (panels + record) landed
    uint64_t empty[kMaxSlots];               // every MMA that reads the slot has retired
    uint64_t acc_full[kMaxAcc];              // a block's accumulators are complete
    uint64_t acc_empty[kMaxAcc];             // the epilogue has read the buffer out of TMEM
    uint64_t v_full[kMaxVBufs];              // 8 rows of the intermediate are in V[buf]
    uint64_t v_empty[kMaxVBufs];             // every pass-2 warp has read V[buf]
    uint32_t tmem_base;
};

// What is not computed: canvas rows outside the kept rectangle and the margins left / right of the kept columns
// (fill_post).  Dedicated fill CTAs (blockIdx.x >= n_strips) write them with 16-byte stores, one canvas row per warp,
// while the strip CTAs of the same image stream the source: the fill costs the strip pipeline nothing.
__device__ __forceinline__ void fill_span(uint32_t *row, int x0, int x1, uint32_t v, int lane) {
    if (x1 <= x0) return;
    const int head = min(x1 - x0, int(((16u - (uint32_t(reinterpret_cast<uintptr_t>(row + x0)) & 15u)) & 15u) >> 2));
    if (lane < head) row[x0 + lane] = v;
    const int xb = x0 + head, nvec = (x1 - xb) >> 2;
    const uint4 vv = make_uint4(v, v, v, v);
    for (int i = lane; i < nvec; i += 32) reinterpret_cast<uint4 *>(row + xb)[i] = vv;
    const int xt = xb + nvec * 4;
    if (xt + lane < x1) row[xt + lane] = v;
}

template <int NCH2>
__global__ void __launch_bounds__(kThreads, 1) resize_tc_kernel(const TcParams p, const __grid_constant__ CUtensorMap tmap,
                                                                const __grid_constant__ CUtensorMap tmap3) {
    constexpr int NDIG = 2;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);

    if (blockIdx.x >= p.n_strips) {              // fill CTA: no shared memory, no barriers, no TMEM
        const int part = int(blockIdx.x - p.n_strips) * int(gridDim.y) + int(blockIdx.y);
        const int nparts = int(gridDim.x - p.n_strips) * int(gridDim.y);
        const int fw = int(p.fw), kx0 = int(p.fox), kx1 = int(p.fox + p.pw), ky0 = int(p.foy), ky1 = int(p.foy + p.ph);
#pragma unroll 1
        for (int fr = part * kWarps + warp; fr < int(p.fh); fr += nparts * kWarps) {
            uint32_t *row = dst + size_t(fr) * fw;
            if (fr < ky0 || fr >= ky1) { fill_span(row, 0, fw, p.fill_post, lane); continue; }
            fill_span(row, 0, kx0, p.fill_post, lane);
            fill_span(row, kx1, fw, p.fill_post, lane);
        }
        return;
    }

    const uint32_t ring = smem_u32(smem_raw);
    const uint32_t crec_bytes = p.crec_words * 4u;
    const uint32_t crec_base = ring + p.slots * p.slot_bytes;
    const uint32_t vbase = crec_base + p.slots * crec_bytes;
    const uint32_t vplane = 32u * p.pitch_v;                   // one limb of one buffer
    const size_t tail_off = size_t(p.slots) * (p.slot_bytes + crec_bytes) + 2u * p.vbufs * vplane;
    SmemTail &tail = *reinterpret_cast<SmemTail *>(smem_raw + tail_off);

    // ---- this CTA's strip and band (a range of 8-row blocks) -------------------------------------------
    const int strip = blockIdx.x;
    const int col0 = p.strip_col0[strip], col1 = p.strip_col0[strip + 1];
    const int blk0 = int((uint64_t(blockIdx.y) * p.n_blocks) / gridDim.y);
    const int blk1 = int((uint64_t(blockIdx.y + 1) * p.n_blocks) / gridDim.y);
    const int n_blk = blk1 - blk0;
    const int c_lo = n_blk > 0 ? p.blk_c0[blk0] : 0, c_hi = n_blk > 0 ? p.blk_c1[blk1 - 1] : -1;
    const int n_steps = c_hi - c_lo + 1;
    const int ntile = (col1 - col0 + 7) >> 3;
    const int npan = int(p.n_panels), nbuf = int(p.nbuf), slots = int(p.slots), nvb = int(p.vbufs);

    if (tid == 0) {
        for (int s = 0; s < slots; s++) { mbar_init(&tail.full[s], 1); mbar_init(&tail.empty[s], 1); }
        for (int b = 0; b < nbuf; b++) { mbar_init(&tail.acc_full[b], 1); mbar_init(&tail.acc_empty[b], kEpiWarps); }
        for (int b = 0; b < nvb; b++) { mbar_init(&tail.v_full[b], kEpiWarps); mbar_init(&tail.v_empty[b], uint32_t(ntile)); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kMmaWarp) tmem_alloc(&tail.tmem_base, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tail.tmem_base;
    const bool timed = (p.debug & 32u) != 0 && p.stats != nullptr;
    unsigned long long w0 = 0, w1 = 0;
    const long long t_start = timed ? clock64() : 0;

    if (warp == kProducerWarp) {
        // ================= producer: one box per panel + the chunk record, per ring slot =================
        // Lane l < npan loads panel l of the chunk; lane 31 the chunk record.  Chunks beyond the ring are prefetched as
        // far as L2 (pf_chunks of them), so a chunk's DRAM latency is paid before it gets a slot.
        const int tx0 = p.strip_src0[strip] >> 1;                                   // 8-byte elements
        const int tr0 = int(blockIdx.z) * int(p.sh) + int(p.ry);
        const uint32_t step_bytes = uint32_t(npan) * 4096u + crec_bytes;
        const int pf = int(p.pf_chunks);
        int next_pf = 0;
        auto prefetch_to = [&](int upto) {                      // chunks [next_pf, upto)
            for (int s = next_pf; s < upto; s++)
                if (lane < npan) tma_prefetch_2d(&tmap, tx0 + lane * 16, tr0 + (c_lo + s) * 32);
            next_pf = max(next_pf, upto);
        };
        if (pf) { next_pf = min(n_steps, slots); prefetch_to(min(n_steps, slots + pf)); }
#pragma unroll 1
        for (int s = 0; s < n_steps; s++) {
            const int slot = s % slots;
            if (s >= slots) mbar_wait_t(&tail.empty[slot], uint32_t(s / slots - 1) & 1u, timed, w0);
            if (lane == 0) mbar_arrive_expect_tx(&tail.full[slot], step_bytes);
            __syncwarp();
            const int c = c_lo + s;
            if (p.debug & 16u) {           // EXPERIMENT: one 3-D box for the whole chunk
                if (lane == 0) tma_load_3d(ring + uint32_t(slot) * p.slot_bytes, &tmap3, tx0, tr0 + c * 32, 0, &tail.full[slot]);
            } else if (lane < npan)
                tma_load_2d(ring + uint32_t(slot) * p.slot_bytes + uint32_t(lane) * 4096u, &tmap, tx0 + lane * 16, tr0 + ((p.debug & 4u) ? (c & 7) : c) * 32, &tail.full[slot]);
            if (lane == 31)
                bulk_g2s(crec_base + uint32_t(slot) * crec_bytes, p.crecs + size_t(c) * p.crec_words, crec_bytes, &tail.full[slot]);
            if (pf) prefetch_to(min(n_steps, s + 1 + slots + pf));
        }
    } else if (warp == kMmaWarp) {
        // ================= MMA issue ==========================================================================
        // One elected thread runs the whole loop (tcgen05.mma / commit are single-thread instructions).  Per chunk: the
        // blocks that CONTINUE through it sit in adjacent accumulator buffers, so one instruction per panel covers them
        // all (N = 16 x blocks: the panel is read from shared memory once; a second instruction only where the buffer ring
        // wraps); a block that STARTS in the chunk gets its own instruction, the one that overwrites instead of
        // accumulating.  The loop is kept free of divisions and generic loads: this thread paces the whole CTA
        // (measured: ~1100 clk per chunk of scalar overhead in the first version, more than the chunk's 870-clk budget).
        if (elect_one()) {
            int t_first = blk0, fm = 0, fu = 0;          // tracked block (>= blk0), its accumulator buffer and use count
#pragma unroll 1
            for (int s = 0; s < n_steps; s++) {
                const int slot = s % slots;
                mbar_wait_t(&tail.full[slot], uint32_t(s / slots) & 1u, timed, w0);
                tc_fence_after();
                const uint32_t rec_s = crec_base + uint32_t(slot) * crec_bytes;
                const int first = int(lds32(rec_s)), nb = int(lds32(rec_s + 4u));
                const uint32_t flags = lds32(rec_s + 8u);
                const int i_lo = max(0, blk0 - first), i_hi = min(nb, blk1 - first);
                if (i_lo < i_hi) {
                    // advance the tracked block to the first block of this chunk that belongs to the band
                    for (int d = first + i_lo - t_first; d > 0; d--) { if (++fm == nbuf) { fm = 0; fu++; } }
                    t_first = first + i_lo;
                    const uint64_t a_desc0 = desc_a(ring + uint32_t(slot) * p.slot_bytes);
                    const uint64_t b_desc0 = desc_b(rec_s + 64u);
                    int i = i_lo, ab = fm, use = fu;
                    // continuing blocks first (they are the older ones), in runs that do not wrap around the buffer ring
                    while (i < i_hi && !((flags >> i) & 1u)) {
                        int len = 1;
                        while (i + len < i_hi && !((flags >> (i + len)) & 1u) && ab + len < nbuf) len++;
                        if (!(p.debug & 1u)) {
                            const uint32_t idesc = kIdescBase | (uint32_t(len * 16) >> 3 << 17);
                            const uint64_t bd = b_desc0 + uint64_t(i) * (512u >> 4);
                            uint64_t ad = a_desc0;
                            uint32_t dcol = tmem + uint32_t(ab) * 16u;
#pragma unroll 1
                            for (int pn = 0; pn < npan; pn++) {
                                umma_i8(dcol, ad, bd, idesc, 1u);
                                ad += 4096u >> 4;
                                dcol += uint32_t(nbuf) * 16u;
                            }
                        }
                        i += len; ab += len;
                        if (ab >= nbuf) { ab -= nbuf; use++; }
                    }
                    // blocks that start here: wait until the buffer's previous block has been read out, then overwrite
                    for (; i < i_hi; i++) {
                        mbar_wait_t(&tail.acc_empty[ab], (uint32_t(use) & 1u) ^ 1u, timed, w1);
                        tc_fence_after();
                        if (!(p.debug & 1u)) {
                            const uint32_t idesc = kIdescBase | (16u >> 3 << 17);
                            const uint64_t bd = b_desc0 + uint64_t(i) * (512u >> 4);
                            uint64_t ad = a_desc0;
                            uint32_t dcol = tmem + uint32_t(ab) * 16u;
#pragma unroll 1
                            for (int pn = 0; pn < npan; pn++) {
                                umma_i8(dcol, ad, bd, idesc, 0u);
                                ad += 4096u >> 4;
                                dcol += uint32_t(nbuf) * 16u;
                            }
                        }
                        if (++ab == nbuf) { ab = 0; use++; }
                    }
                    // blocks that end here are complete once everything issued so far retires
                    ab = fm;
                    for (i = i_lo; i < i_hi; i++) {
                        if ((flags >> (8 + i)) & 1u) umma_commit(&tail.acc_full[ab]);
                        if (++ab == nbuf) ab = 0;
                    }
                }
                umma_commit(&tail.empty[slot]);                  // the slot is free once everything issued so far retires
            }
            if (timed) { atomicAdd(p.stats + 1, w0); atomicAdd(p.stats + 2, w1); w0 = w1 = 0; }
        }
        __syncwarp();
    } else if (warp >= kEpiWarp0 && warp < kEpiWarp0 + kEpiWarps) {
        // ================= epilogue: TMEM -> 16-bit intermediate -> V planes ==============================
        // A warp reads the TMEM lane quarter (warp & 3) -- 8 pixels x 4 channels of a panel -- for half of the strip's
        // panels: 16 columns per panel = 8 rows x 2 digits, the next panel's load in flight while this one is
        // recombined (exact s32), biased, shifted to 16 bits and stored as two u8 limbs, planar [channel][row][pixel].
        const int quarter = warp & 3, half = (warp - kEpiWarp0) >> 2;
        const int per = (npan + 1) >> 1;
        const int pn0 = half * per, pn1 = min(npan, pn0 + per);
        const int ch = lane & 3;
        const int sh1 = p.sh1;
        const uint32_t lane_addr = uint32_t(quarter * 32) << 16;
        // byte address of (row 0 of channel ch, pixel quarter*8 + lane/4 of panel 0): rows pitch_v apart, channel
        // planes skewed by two words (a warp's 4 channels x 8 pixels then land in 8 distinct banks)
        const uint32_t v0 = vbase + uint32_t(8 * ch) * p.pitch_v + uint32_t(ch) * 8u + uint32_t(quarter * 8 + (lane >> 2));
#pragma unroll 1
        for (int bi = 0; bi < n_blk; bi++) {
            const int use = bi / nbuf, abuf = bi - use * nbuf;
            const int vuse = bi / nvb, vb = bi - vuse * nvb;
            int bias[8];
#pragma unroll
            for (int r = 0; r < 8; r++) bias[r] = __ldg(p.bias + (size_t(blk0 + bi) * 8 + r) * 4 + ch);
            mbar_wait_t(&tail.acc_full[abuf], uint32_t(use) & 1u, timed, w0);
            tc_fence_after();
            int cur[16], nxt[16];
            const uint32_t t0 = tmem + lane_addr + uint32_t(abuf) * 16u;
            if (pn0 < pn1 && !(p.debug & 8u)) tmem_ld16(t0 + uint32_t(pn0 * nbuf) * 16u, cur);
            mbar_wait_t(&tail.v_empty[vb], (uint32_t(vuse) & 1u) ^ 1u, timed, w1);
            const uint32_t vrow = v0 + uint32_t(vb) * 2u * vplane;
#pragma unroll 1
            for (int pn = pn0; pn < ((p.debug & 8u) ? pn0 : pn1); pn++) {
                tmem_ld_wait();
                if (pn + 1 < pn1) tmem_ld16(t0 + uint32_t((pn + 1) * nbuf) * 16u, nxt);
                else {                                   // the block is out of TMEM: its buffer may be overwritten
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tail.acc_empty[abuf]);
                }
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    const int x = cur[r] + (cur[8 + r] << 8) + bias[r];
                    const uint32_t v = uint32_t(x >> sh1);
                    const uint32_t a = vrow + uint32_t(r) * p.pitch_v + uint32_t(pn) * 32u;
                    sts8(a, v);
                    sts8(a + vplane, v >> 8);
                }
                if (pn + 1 < pn1) {
                    tmem_ld_wait();
#pragma unroll
                    for (int r = 0; r < 16; r++) cur[r] = nxt[r];
                }
            }
            if (pn0 >= pn1 || (p.debug & 8u)) {          // a warp without panels still takes part in the hand-offs
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tail.acc_empty[abuf]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&tail.v_full[vb]);
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
        for (int gi = 0; gi < n_blk; gi++) {
            const int vuse = gi / nvb, buf = gi - vuse * nvb;
            const uint32_t vbuf = vbase + uint32_t(buf) * 2u * vplane;
            mbar_wait_t(&tail.v_full[buf], uint32_t(vuse) & 1u, timed, w0);
            if (p.debug & 2u) { __syncwarp(); if (lane == 0) mbar_arrive(&tail.v_empty[buf]); continue; }

            int acc[2][2][NDIG][4];            // [channel pair][limb][digit]
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {   // channel pairs (0,1) and (2,3); channel planes are skewed by two words
                const uint32_t r0 = vbuf + uint32_t(16 * mt + g) * p.pitch_v + uint32_t(2 * mt) * 8u + uint32_t(kb0 + t) * 4u;
                const uint32_t r1 = r0 + 8u * p.pitch_v + 8u;
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
            if (lane == 0) mbar_arrive(&tail.v_empty[buf]);       // V[buf] may be overwritten (vbufs blocks on)

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
            const int i_row = (blk0 + gi) * 8 + g;
            if (i_row < int(p.ph)) {
                uint32_t *o = dst + size_t(p.foy + i_row) * p.fw + p.fox + ocol;
                if (ocol < col1) o[0] = px[0];
                if (ocol + 1 < col1) o[1] = px[1];
            }
        }
    }
    if (timed && lane == 0) {
        // 0: producer waits for a free slot; 1 / 2: MMA thread waits for a chunk / a free accumulator; 3 / 4: epilogue warps wait
        // for a finished block / a free V buffer; 5: pass-2 warps wait for V; 6: role lifetime (per role thread); 7..: counts
        const unsigned long long life = (unsigned long long)(clock64() - t_start);
        if (warp == kProducerWarp) { atomicAdd(p.stats + 0, w0); atomicAdd(p.stats + 6, life); atomicAdd(p.stats + 7, 1ull); }
        else if (warp == kMmaWarp) { atomicAdd(p.stats + 1, w0); atomicAdd(p.stats + 2, w1); }
        else if (warp >= kEpiWarp0 && warp < kEpiWarp0 + kEpiWarps) { atomicAdd(p.stats + 3, w0); atomicAdd(p.stats + 4, w1); atomicAdd(p.stats + 8, life); atomicAdd(p.stats + 9, 1ull); }
        else if (warp < ntile) { atomicAdd(p.stats + 5, w0); atomicAdd(p.stats + 10, life); atomicAdd(p.stats + 11, 1ull); }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem, 512);
    }
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

// EXPERIMENT: (x, row, panel) view, panel stride 128 bytes, box = 128 B x 32 rows x n_panels
static cudaError_t make_panel_map3(const TcParams &p, const uint8_t *src, uint32_t n_images, CUtensorMap &map) {
    EncodeTiledFn enc = tc_encode_tiled_fn();
    if (!enc) return cudaErrorNotSupported;
    const cuuint64_t dims[3] = {cuuint64_t(p.sw) / 2, cuuint64_t(p.sh) * n_images, cuuint64_t(p.n_panels)};
    const cuuint64_t strides[2] = {cuuint64_t(p.sw) * 4, 128};
    const cuuint32_t box[3] = {16, 32, p.n_panels};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<uint8_t *>(src), dims, strides, box, estr,
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
        CUtensorMap map3 = map;
        if (q.debug & 16u) {
            if (make_panel_map3(q, q.src, nz, map3) != cudaSuccess) {
                static bool said = false;
                if (!said) { std::fprintf(stderr, "[slimg_b200] 3-D panel map rejected by cuTensorMapEncodeTiled\n"); said = true; }
                q.debug &= ~16u;
            }
        }
        // fill CTAs: one per ~2 MB of canvas that is not computed (at least one when anything is), spread over the bands
        unsigned nfill = 0;
        if (p.need_fill) {
            const uint64_t bytes = (uint64_t(p.fw) * p.fh - uint64_t(p.pw) * p.ph) * 4u;
            nfill = unsigned(std::min<uint64_t>(32, 1 + bytes / (2u << 20)));
            nfill = (nfill + p.n_bands - 1) / p.n_bands;
        }
        static unsigned long long *d_stats = nullptr;
        if (q.debug & 32u) {
            if (!d_stats) cudaMalloc((void **)&d_stats, 16 * sizeof(unsigned long long));
            cudaMemsetAsync(d_stats, 0, 16 * sizeof(unsigned long long), stream);
            q.stats = d_stats;
        } else q.stats = nullptr;
        resize_tc_kernel<NCH2><<<dim3(p.n_strips + nfill, p.n_bands, nz), kThreads, p.smem_bytes, stream>>>(q, map, map3);
        if (q.debug & 32u) {
            unsigned long long h[16];
            cudaMemcpyAsync(h, d_stats, sizeof h, cudaMemcpyDeviceToHost, stream);
            cudaStreamSynchronize(stream);
            const double n = double(h[7] ? h[7] : 1);
            static int printed = 0;
            if (printed++ < 2) std::fprintf(stderr, "[slimg_b200] tc waits per strip CTA (kclk): producer/free-slot %.1f | mma/chunk %.1f mma/free-acc %.1f | epi/block-done %.1f epi/free-V %.1f (per warp) | "
                                 "pass2/V %.1f (per warp) | lifetime producer %.1f epi %.1f pass2 %.1f | CTAs %.0f\n",
                         h[0] / n / 1e3, h[1] / n / 1e3, h[2] / n / 1e3, h[3] / double(h[9] ? h[9] : 1) / 1e3, h[4] / double(h[9] ? h[9] : 1) / 1e3,
                         h[5] / double(h[11] ? h[11] : 1) / 1e3, h[6] / n / 1e3, h[8] / double(h[9] ? h[9] : 1) / 1e3, h[10] / double(h[11] ? h[11] : 1) / 1e3, n);
        }
        count_launch();
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_resize_tc(const TcParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    if (p.n_bands == 0 || p.n_bands > 65535 || p.n_strips == 0 || p.slots > uint32_t(kMaxSlots) || p.nbuf > uint32_t(kMaxAcc) ||
        p.nbuf < 2 || p.n_panels * p.nbuf * 16 > 512 || p.vbufs < 2 || p.vbufs > uint32_t(kMaxVBufs))
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
