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
//   HBM --cp.async.bulk.tensor.2d (TMA engine, one box of 16 source rows per copy; the same boxes are prefetched
//        into L2 16 steps ahead for bands of 3+ chunks)--> ring of source rows in shared memory
//   pass 1, per group of 8 output rows:  D1[16 bytes of the row][8 rows] += A[16][32 rows] * B[32][8]
//        A comes straight out of the ring with ldmatrix.m16n16.x2.trans.b8: the transposing load
//        hands every lane 4 consecutive SOURCE ROWS of one byte column, which is exactly the A
//        fragment of the MMA -- no shuffles, no byte permutes, channels stay interleaved because
//        every channel of every pixel uses the same vertical weights.
//        B = weight digits (s8), register resident for the whole group.
//        epilogue: recombine the digit planes (exact s32), add fill/offset/rounding, >> to a u16 with
//        7 fractional bits, split into two u8 limbs and store them planar ([channel][row][pixel]).
//   pass 2, per group: D2[(channel,row) 16][8 output columns] += A[16][32 px] * B[32][8] for both limbs
//        three accumulators per channel pair, by weight: low limb x digit 0, low x digit 1 + high x digit 0 (it
//        starts at the constant term: fill taps, intermediate offset, rounding half, as the MMA's C operand),
//        high x digit 1.  Epilogue in integers: (a >> 8) + b + (c << 8), one shift, cvt.pack.sat.u8.s32, and
//        8-byte stores straight to the final canvas position.
// Nothing but source pixels is read from HBM and nothing but final pixels is written: the f32
// intermediate of the reference (38-51 MB per image) never exists.
//
// Grid: x = strips (+ fill CTAs), y = bands of output-row groups, z = image.  512 threads, one CTA per SM: the
// first (strip columns / 8) warps run pass 2, the next ones pass 1 -- in 1 to 3 cohorts that take turns on the
// groups, so two or three groups are in flight -- then a signaller warp and a producer warp.  They meet only on
// mbarriers: there is no CTA-wide barrier in the band loop.
#include <cuda.h>

#include <algorithm>
#include <cstdio>

#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kWarps = kVfWarps;
constexpr int kThreads = kWarps * 32;
constexpr int kStepRows = 16;                // source rows per ring step (one mbarrier each) == kWarps
constexpr int kMaxSteps = 32;                // ring of at most 512 rows (one step per producer lane)
constexpr int kMaxVBufs = kVfMaxVBufs;            // V (vertical-pass result) buffers: pass 1 may run this many groups ahead of pass 2
constexpr int kRecBufs = 4;                  // group records in flight: fetched three groups ahead (vplan.cpp sizes for it)
constexpr uint32_t kSpinLimit = 1u << 27;    // a wedged pipeline traps instead of hanging the GPU

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
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// TMA engine, tiled form: one box of kStepRows source rows (pitch bytes each) per instruction
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}

// same box, only as far as L2: hides the DRAM part of the latency from the (small) ring
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap *map, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];" ::"l"(map), "r"(c0), "r"(c1) : "memory");
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
// first k-chunk: C = 0 comes from the zero register, no accumulator clearing
__device__ __forceinline__ void imma_z(int (&d)[4], const uint32_t (&a)[4], const uint2 &b) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y), "r"(0));
}
// first k-chunk of the first digit: the accumulator starts at the bias of its output row (columns n = 2t, 2t+1)
__device__ __forceinline__ void imma_c(int (&d)[4], const uint32_t (&a)[4], const uint2 &b, int c0, int c1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%10,%11};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y), "r"(c0), "r"(c1));
}
// C = {c0, c1} for fragment row g (columns 2t, 2t+1) and {c2, c3} for row g + 8
__device__ __forceinline__ void imma_c4(int (&d)[4], const uint32_t (&a)[4], const uint2 &b, int c0, int c1, int c2, int c3) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y), "r"(c0), "r"(c1), "r"(c2), "r"(c3));
}
// d = (c << 16) | (sat_u8(hi) << 8) | sat_u8(lo): clamp to [0, 255] and pack, two channels per instruction
__device__ __forceinline__ uint32_t pack_sat_u8(int hi, int lo, uint32_t c) {
    uint32_t d;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(lo), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts16(uint32_t addr, uint32_t v) {          // the low half of v
    asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(uint16_t(v)) : "memory");
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
// exact s32 -> f32 for |x| < 2^22 without the quarter-rate I2F
__device__ __forceinline__ float int_to_float(int x) { return __int_as_float(x + 0x4B400000) - 12582912.0f; }

__device__ __forceinline__ uint32_t quantize_channel(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f) + 0.5f;      // clamp, then f32::round (half away; value is >= 0) = floor(t + 0.5):
    return __float_as_uint(__fadd_rz(t, 8388608.0f)) & 0xffu;      // 2^23 + floor(t) by truncation, no FRND / F2I
}

// -DSLIMG_VF_EXPERIMENT (python -m slimg_b200._build with SLIMG_B200_NVCC_EXTRA): cycles every role spends in its
// mbarrier waits, printed by one strip CTA of the first launches when SLIMG_B200_VF_DEBUG has bit 32 set -- the numbers
// of profiles/r02c_role_waits.txt.  Expands to nothing in the production build.
#ifdef SLIMG_VF_EXPERIMENT
__device__ int g_vf_prints = 0;
#define VF_TIMED(acc, stmt) { const long long t__ = clock64(); stmt; acc += clock64() - t__; }
#define VF_REPORT(role) \
    if ((p.debug & 32u) && blockIdx.x == 3 && blockIdx.y == gridDim.y / 2 && blockIdx.z == 0 && lane == 0 && atomicAdd(&g_vf_prints, 1) < 48) \
        printf("[vf] %s warp %2d groups %d total %lld wait0 %lld wait1 %lld wait2 %lld\n", role, warp, n_iter, clock64() - t_begin, w0, w1, w2);
#else
#define VF_TIMED(acc, stmt) stmt;
#define VF_REPORT(role)
#endif

struct SmemTail {                             // after the ring, the V planes and the two group records
    uint64_t full[kMaxSteps];                // ring step landed (tx bytes)
    uint64_t rec_full[kRecBufs];                  // group record landed (tx bytes)
    uint64_t go[kMaxVBufs];                     // producer: record + source rows of the group are in, V[buf] is free
    uint64_t v_full[kMaxVBufs];                 // every pass-1 warp is done with the group: V complete, band rows released
    uint64_t v_empty[kMaxVBufs];                // every pass-2 warp has read V and the producer has seen v_full
    int pos[kMaxVBufs];                           // ring row of the group's band start (written by the producer before `go`)
};

// What is not computed: canvas rows outside the kept rectangle (fill_post), kept rows whose taps
// are all fill (fill_pre), and the margins left / right of the kept columns.  Dedicated fill CTAs
// (blockIdx.x >= n_strips) write them with 16-byte stores, one canvas row per warp, while the strip
// CTAs of the same image stream the source: the fill costs the strip pipeline nothing.
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

__device__ __forceinline__ void fill_canvas(uint32_t *dst, int fw, int fh, int fox, int foy, int pw, int ph, const uint8_t *row_computed,
                                            bool all_computed, uint32_t fill_pre, uint32_t fill_post, int part, int nparts) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int kx0 = fox, kx1 = fox + pw;
#pragma unroll 1
    for (int fr = part * kWarps + warp; fr < fh; fr += nparts * kWarps) {
        const int i = fr - foy;
        const bool kept = i >= 0 && i < ph;
        const bool computed = kept && (all_computed || row_computed[i]);
        uint32_t *row = dst + size_t(fr) * fw;
        if (!kept) { fill_span(row, 0, fw, fill_post, lane); continue; }
        fill_span(row, 0, kx0, fill_post, lane);
        if (!computed) fill_span(row, kx0, kx1, fill_pre, lane);
        fill_span(row, kx1, fw, fill_post, lane);
    }
}

template <int NDIG, int NCH1, int NCH2, int UNIT>
__global__ void __launch_bounds__(kThreads, 1) resize_vfirst_kernel(const VfParams p, const __grid_constant__ CUtensorMap tmap) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const uint32_t ring = smem_u32(smem_raw);
    const uint32_t vbase = ring + p.ring_rows * p.pitch;
    const uint32_t vplane = 32u * p.pitch_v;                   // one limb of one buffer
    const size_t rec_off = size_t(p.ring_rows) * p.pitch + 2u * p.vbufs * vplane;
    const uint32_t *recbuf = reinterpret_cast<const uint32_t *>(smem_raw + rec_off);
    SmemTail &tail = *reinterpret_cast<SmemTail *>(smem_raw + rec_off + uint32_t(kRecBufs) * p.rec_words * 4u);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);
    if (blockIdx.x >= p.n_strips) {              // fill CTA: no shared memory, no barriers
        fill_canvas(dst, int(p.fw), int(p.fh), int(p.fox), int(p.foy), int(p.pw), int(p.ph), p.row_computed, (p.need_fill & 2u) == 0,
                    p.fill_pre_packed, p.fill_post, int(blockIdx.x - p.n_strips), int(gridDim.x - p.n_strips));
        return;
    }

    // ---- this CTA's strip and band (a range of group records) -----------------------------------------
    const int strip = blockIdx.x;
    const int col0 = p.strip_col0[strip], col1 = p.strip_col0[strip + 1];
    const int rec0 = int((uint64_t(blockIdx.y) * p.n_recs) / gridDim.y);
    const int rec1 = int((uint64_t(blockIdx.y + 1) * p.n_recs) / gridDim.y);
    const int n_iter = rec1 - rec0;
    const int ys = n_iter > 0 ? int(p.recs[size_t(rec0) * p.rec_words + 1]) & ~(kStepRows - 1) : 0;
    const int ye = n_iter > 0 ? int(p.recs[size_t(rec1 - 1) * p.rec_words + 2]) : 0;
    const int n_steps = (ye - ys + kStepRows - 1) / kStepRows;
    const int nst = int(p.ring_rows) / kStepRows;

    // ---- roles ---------------------------------------------------------------------------------------
    //   warps [0, ntile)        pass 2: one tile of 8 output columns each, all four channels
    //   warps [ntile, kWarps-2) pass 1: each warp owns `upw` units of UNIT 16-byte blocks of the strip, for every group
    //   warp kWarps-2           signaller: waits for a group's record, source rows and free V buffer, then hands
    //                           pass 1 ONE barrier per group
    //   warp kWarps-1           producer: ring refills and group records through the TMA engine
    // They meet only on mbarriers: no CTA-wide barrier inside the band loop.
    const int ntile = (col1 - col0 + 7) >> 3;
    const int n_p1 = kWarps - 2 - ntile;
    const int nco = int(p.cohorts), csize = n_p1 / nco;         // warps of one pass-1 cohort
    const uint32_t rec_bytes = p.rec_words * 4u;
    if (tid == 0) {
        for (int s = 0; s < nst; s++) mbar_init(&tail.full[s], 1);
        for (int b = 0; b < kRecBufs; b++) mbar_init(&tail.rec_full[b], 1);
        for (int b = 0; b < int(p.vbufs); b++) {
            mbar_init(&tail.go[b], 1);
            mbar_init(&tail.v_full[b], uint32_t(csize));
            mbar_init(&tail.v_empty[b], uint32_t(ntile) + 1u);       // every pass-2 warp and the producer (see its loop)
            tail.pos[b] = 0;
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
#ifdef SLIMG_VF_EXPERIMENT
    const long long t_begin = clock64();
    long long w0 = 0, w1 = 0, w2 = 0;
#endif

    if (warp == kWarps - 1) {
        // ================= producer ====================================================================
        // One tiled TMA copy per ring step (a box of 16 source rows of the strip; the source batch is one tall
        // 2-D tensor of 8-byte elements, rows past its end and columns past the row end read as zero), one
        // linear bulk copy per group record.  Lane l issues step next_issue + l: a refill is one pass of the warp.
        int next_issue = 0, issue_base = 0;              // issue_base: the multiple of nst at or below next_issue (no division per group)
        const int tc0 = p.strip_src0[strip] >> 1;                                   // 8-byte elements
        const int tr0 = int(blockIdx.z) * int(p.sh) + int(p.ry) + ys;
        const uint32_t step_bytes = kStepRows * p.pitch;
        auto issue_steps = [&](int upto) {
            const int step = next_issue + lane;                 // at most nst <= 32 steps are ever issued at once
            if (next_issue >= issue_base + nst) issue_base += nst;     // next_issue advances by at most nst per call
            if (step < upto) {
                int slot = step - issue_base;
                if (slot >= nst) slot -= nst;
                mbar_arrive_expect_tx(&tail.full[slot], step_bytes);
                tma_load_2d(ring + uint32_t(slot) * step_bytes, &tmap, tc0, tr0 + ((p.debug & 4u) ? (step & 7) : step) * kStepRows, &tail.full[slot]);
            }
            next_issue = max(next_issue, upto);
        };
        // steps beyond the ring, as far as L2: the DRAM part of their latency is paid before they get a ring slot
        int next_pf = 0;
        auto prefetch_steps = [&](int upto) {
            for (int s = max(next_pf, next_issue) + lane; s < upto; s += 32) tma_prefetch_2d(&tmap, tc0, tr0 + s * kStepRows);
            next_pf = max(next_pf, upto);
        };
        auto issue_rec = [&](int it) {
            if (lane == 0) {
                uint64_t *bar = &tail.rec_full[it & (kRecBufs - 1)];
                mbar_arrive_expect_tx(bar, rec_bytes);
                bulk_g2s(smem_u32(recbuf) + uint32_t(it & (kRecBufs - 1)) * rec_bytes, p.recs + size_t(rec0 + it) * p.rec_words, rec_bytes, bar);
            }
        };
        issue_steps(min(n_steps, nst));
        if (p.pf_steps) prefetch_steps(min(n_steps, next_issue + int(p.pf_steps)));
        for (int r = 0; r < kRecBufs; r++)
            if (r < n_iter) issue_rec(r);
        // band start of the group after next, fetched a group early: ring steps entirely above a group's successor's
        // band start are free once the group is done
        int a_next = n_iter > 1 ? int(__ldg(p.recs + size_t(rec0 + 1) * p.rec_words + 1)) : ye;
        uint32_t buf = 0, par = 0;
#pragma unroll 1
        for (int it = 0; it < n_iter; it++) {
            const int a_cur = a_next;
            a_next = it + 2 < n_iter ? int(__ldg(p.recs + size_t(rec0 + it + 2) * p.rec_words + 1)) : ye;
            VF_TIMED(w0, mbar_wait(&tail.v_full[buf], par))
            // The producer acknowledges the phase like a pass-2 warp does: V[buf] is handed to group it + vbufs only after
            // every waiter on v_full[buf] has seen this phase.  Without it the barrier could complete a second phase while
            // this warp was still behind the first (the ring holds rows for several groups ahead, so pass 1 and pass 2 do
            // not depend on the producer in the short run), and a parity wait that is two phases late never returns.
            if (lane == 0) mbar_arrive(&tail.v_empty[buf]);
            if (++buf == p.vbufs) { buf = 0; par ^= 1u; }
            issue_steps(min(n_steps, int(uint32_t(a_cur - ys) / uint32_t(kStepRows)) + nst));
            if (p.pf_steps) prefetch_steps(min(n_steps, next_issue + int(p.pf_steps)));
            if (it + kRecBufs < n_iter) issue_rec(it + kRecBufs);   // the record buffer of this group is free
        }
        VF_REPORT("producer ")
    } else if (warp == kWarps - 2) {
        // ================= signaller: does the waiting for pass 1 =======================================
        // Group `it` may start when its record and source rows have landed and pass 2 has drained V[it & 1].
        // Lane l waits for ring step first + l; then one arrive on go[]: pass-1 warps wait on a single barrier.
        int a0 = n_iter > 0 ? int(__ldg(p.recs + size_t(rec0) * p.rec_words + 1)) : 0;
        int need = n_iter > 0 ? int(__ldg(p.recs + size_t(rec0) * p.rec_words + 2)) : 0;
        uint32_t buf = 0, par = 0;
        // Ring slot and barrier phase of a step without a division per group: short groups (ratios below 4x: 16 - 32 source
        // rows each) were paced by this warp's ~150 dependent scalar instructions (24 MP 2x 57.1 -> 49.7 us, 1.5x 93.0 -> 74.8,
        // 8K 4x 30.9 -> 29.3).  The 4-chunk instances keep the divisions: with them removed the same kernel measured 6 %
        // SLOWER on the 7.5x headline (16.99 vs 18.07 us, same results; the pass-1 loop's schedule and register allocation
        // shift with the code around it: 119 vs 123 registers).
        constexpr bool kDivFree = NCH1 != 4;
        int wrap_base = 0;                                  // ring steps [wrap_base, wrap_base + nst) complete phase wrap_par of their barriers
        uint32_t wrap_par = 0;
#pragma unroll 1
        for (int it = 0; it < n_iter; it++) {
            const int a0_cur = a0, need_cur = need;
            if (it + 1 < n_iter) {                           // next group's header, in flight while this one is waited for
                a0 = int(__ldg(p.recs + size_t(rec0 + it + 1) * p.rec_words + 1));
                need = int(__ldg(p.recs + size_t(rec0 + it + 1) * p.rec_words + 2));
            }
            const int first = int(uint32_t(a0_cur - ys) / uint32_t(kStepRows)), last = int(uint32_t(need_cur - ys + kStepRows - 1) / uint32_t(kStepRows));
            while (first >= wrap_base + nst) { wrap_base += nst; wrap_par ^= 1u; }      // band starts move down monotonically
            const int step = first + lane;                   // a band spans at most 192 rows + alignment: <= 13 steps
            if (!kDivFree) {
                if (step < last) VF_TIMED(w0, mbar_wait(&tail.full[step % nst], uint32_t(step / nst) & 1u))
            } else if (step < last) {
                int slot = step - wrap_base;
                uint32_t ph = wrap_par;
                if (slot >= nst) { slot -= nst; ph ^= 1u; }
                VF_TIMED(w0, mbar_wait(&tail.full[slot], ph))
            }
            VF_TIMED(w1, mbar_wait(&tail.rec_full[it & (kRecBufs - 1)], uint32_t(it / kRecBufs) & 1u))
            VF_TIMED(w2, mbar_wait(&tail.v_empty[buf], par ^ 1u))        // V[buf] was last read by pass 2 vbufs groups ago
            __syncwarp();
            if (lane == 0) {
                tail.pos[buf] = kDivFree ? (first - wrap_base) * kStepRows + ((a0_cur - ys) & (kStepRows - 1)) : (a0_cur - ys) % int(p.ring_rows);
                mbar_arrive(&tail.go[buf]);                  // release: pos is visible to the waiters
            }
            if (++buf == p.vbufs) { buf = 0; par ^= 1u; }
        }
        VF_REPORT("signaller")
    } else if (warp < ntile) {
        // ================= pass 2: horizontal taps, 4 channels x 8 rows x 8 output columns per warp =======
        const int tile = warp;
        const int kb0 = p.tile_kb0[strip * 8 + tile];
        uint2 bw2[NCH2][NDIG];
#pragma unroll
        for (int c = 0; c < NCH2; c++)
#pragma unroll
            for (int d = 0; d < NDIG; d++)
                bw2[c][d] = __ldg(reinterpret_cast<const uint2 *>(p.b2) + ((size_t(strip * 8 + tile) * NCH2 + c) * NDIG + d) * 32 + lane);
        const int ocol = col0 + tile * 8 + 2 * t;          // this lane's two kept output columns
        constexpr bool kIntEpi = NDIG == 2;                // two-digit plans: the sum never leaves the integers
        float cxr[2][4];                                   // [column ocol+e][channel]
        int cxq[2][4];                                     // the same in units of 2^-sh2, rounding half included
#pragma unroll
        for (int e = 0; e < 2; e++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                cxr[e][c] = !kIntEpi && ocol + e < col1 ? p.cx[size_t(ocol + e) * 4 + c] : 0.0f;
                cxq[e][c] = kIntEpi && ocol + e < col1 ? __ldg(p.cxi + size_t(ocol + e) * 4 + c) : 0;
            }
        float s2[2][NDIG];
#pragma unroll
        for (int l = 0; l < 2; l++)
#pragma unroll
            for (int d = 0; d < NDIG; d++) s2[l][d] = p.s2[l][d];
        const int sh2 = p.sh2;
        const bool all_rows = (p.need_fill & 2u) == 0;
        int j_next = n_iter > 0 ? int(__ldg(p.recs + size_t(rec0) * p.rec_words)) : 0;

        uint32_t buf = 0, par = 0;
#pragma unroll 1
        for (int it = 0; it < n_iter; it++, buf = buf + 1 == p.vbufs ? 0 : buf + 1, par ^= buf == 0 ? 1u : 0u) {
            const int j = j_next;
            if (it + 1 < n_iter) j_next = int(__ldg(p.recs + size_t(rec0 + it + 1) * p.rec_words));
            const uint32_t vbuf = vbase + buf * 2u * vplane;
            VF_TIMED(w0, mbar_wait(&tail.v_full[buf], par))

            if (p.debug & 2u) { __syncwarp(); if (lane == 0) mbar_arrive(&tail.v_empty[buf]); continue; }
            uint32_t px[2] = {0u, 0u};         // packed RGBA of columns ocol, ocol+1 of output row g
            if (kIntEpi) {
                // three accumulators per channel pair, by weight: low limb x digit 0 (2^-8), low x digit 1 + high x digit 0
                // (2^0, starts at the constant term), high x digit 1 (2^8)
                int tsum[2][4];
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {   // channel pairs (0,1) and (2,3)
                    const uint32_t r0 = vbuf + uint32_t(16 * mt + g) * p.pitch_v + uint32_t(2 * mt) * 4u + uint32_t(kb0 + t) * 4u;
                    const uint32_t r1 = r0 + 8u * p.pitch_v + 4u;
                    int aa[4], ab[4], ac[4];
#pragma unroll
                    for (int c = 0; c < NCH2; c++) {
#pragma unroll
                        for (int l = 0; l < 2; l++) {
                            uint32_t a[4];
                            const uint32_t o = uint32_t(l) * vplane + uint32_t(c) * 32u;
                            a[0] = lds32(r0 + o); a[1] = lds32(r1 + o); a[2] = lds32(r0 + o + 16u); a[3] = lds32(r1 + o + 16u);
                            if (l == 0) {
                                if (c == 0) {
                                    imma_z(aa, a, bw2[c][0]);
                                    imma_c4(ab, a, bw2[c][NDIG - 1], cxq[0][2 * mt], cxq[1][2 * mt], cxq[0][2 * mt + 1], cxq[1][2 * mt + 1]);
                                } else {
                                    imma(aa, a, bw2[c][0]);
                                    imma(ab, a, bw2[c][NDIG - 1]);
                                }
                            } else {
                                imma(ab, a, bw2[c][0]);
                                if (c == 0) imma_z(ac, a, bw2[c][NDIG - 1]);
                                else imma(ac, a, bw2[c][NDIG - 1]);
                            }
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 4; i++) tsum[mt][i] = (aa[i] >> 8) + ab[i] + (ac[i] << 8);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&tail.v_empty[buf]);       // V[buf] may be overwritten (vbufs groups on)
                // i = 0,1: channel 2mt, columns +0, +1; i = 2,3: channel 2mt+1.  Shift, saturate, pack RGBA.
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const int c0 = tsum[0][e] >> sh2, c1 = tsum[0][2 + e] >> sh2, c2 = tsum[1][e] >> sh2, c3 = tsum[1][2 + e] >> sh2;
                    px[e] = pack_sat_u8(c1, c0, pack_sat_u8(c3, c2, 0u));
                }
            } else {
                int acc[2][2][NDIG][4];            // [channel pair][limb][digit]
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
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
                if (lane == 0) mbar_arrive(&tail.v_empty[buf]);
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
            }
            const int i_row = j * 8 + g;
            if (i_row < int(p.ph) && (all_rows || p.row_computed[i_row])) {
                uint32_t *o = dst + size_t(p.foy + i_row) * p.fw + p.fox + ocol;
                if (ocol + 1 < col1 && (reinterpret_cast<uintptr_t>(o) & 7) == 0) *reinterpret_cast<uint2 *>(o) = make_uint2(px[0], px[1]);
                else {
                    if (ocol < col1) o[0] = px[0];
                    if (ocol + 1 < col1) o[1] = px[1];
                }
            }
        }
        VF_REPORT("pass2    ")
    } else {
        // ================= pass 1: vertical taps of 8 output rows over the strip =======================
        // Static ownership: warp w1 of the n_p1 pass-1 warps computes units [w1*upw, (w1+1)*upw) of UNIT blocks in
        // every group (the plan pads the staged span to n_p1*upw*UNIT blocks), so there is no queue and one
        // barrier wait per group.
        // Cohorts: with nco = 2 the warps of cohort 0 filter groups 0, 2, 4, ... and those of cohort 1 groups 1, 3, 5, ...
        // Two groups are in flight at once, each warp owns twice the blocks of half the groups, and the barrier round
        // trip, the weight fragments and the set-up of a group are paid once per two units.  Neighbouring warps are in
        // different cohorts and the pairing flips every four warps, so every SM sub-partition hosts both cohorts and
        // one's multiplies fill the other's epilogue.
        const int w1a = warp - ntile;
        const int co = nco == 1 ? 0 : nco == 2 ? (((w1a >> 2) + w1a) & 1) : w1a % nco;
        const int w1 = nco == 1 ? w1a : nco == 2 ? (w1a >> 1) : w1a / nco;             // rank inside the cohort
        if (w1 >= csize) return;                           // n_p1 is not a multiple of nco: the last warps have no role
        const int upw_all = p.strip_upw[strip];
        const int upw = max(0, min(upw_all, p.strip_nunits[strip] - w1 * upw_all));     // this warp's units
        const int sh1 = p.sh1;
        const int ch = g & 3, hh = g >> 2;
        const int ring_rows = int(p.ring_rows);
        const uint32_t boff0 = uint32_t(w1 * upw_all) * (UNIT * 16u);

        uint32_t buf = uint32_t(co) % p.vbufs, par = uint32_t(co) / p.vbufs;     // nco <= 3, vbufs >= 2
#pragma unroll 1
        for (int it = co; it < n_iter; it += nco) {
            const uint32_t *rec = recbuf + uint32_t(it & (kRecBufs - 1)) * p.rec_words;
            VF_TIMED(w0, mbar_wait(&tail.go[buf], par))
            const int pos = tail.pos[buf];
            uint2 bw1[NCH1][NDIG];
#pragma unroll
            for (int c = 0; c < NCH1; c++)
#pragma unroll
                for (int d = 0; d < NDIG; d++)
                    bw1[c][d] = *reinterpret_cast<const uint2 *>(rec + 64 + ((c * NDIG + d) * 32 + lane) * 2);
            const int bias0 = int(rec[8 + (2 * t) * 4 + ch]), bias1 = int(rec[8 + (2 * t + 1) * 4 + ch]);

            uint32_t rowoff[NCH1];
#pragma unroll
            for (int c = 0; c < NCH1; c++) {
                int rr = pos + 32 * c + lane;
                if (rr >= ring_rows) rr -= ring_rows;
                if (rr >= ring_rows) rr -= ring_rows;
                rowoff[c] = ring + uint32_t(rr) * p.pitch + boff0;
                asm volatile("" : "+r"(rowoff[c]));       // keep it in a register: cheaper than recomputing per block
            }
            // V word address of (row 8*ch + 2t + hh, block 0): rows are pitch_v apart, channel planes skewed by one word
            // Two-chunk bands store halfwords: a V word holds pixels (0, 2, 1, 3) of a block (vplan.cpp kPerm), so this lane's
            // pixels hh, hh + 2 are one 16-bit half of it and need no exchange with the lane that holds the other two
            // (measured on B200: 8K 4x 29.84 -> 28.63 us, 24 MP 3x 32.19 -> 31.47).  The others keep the shuffle and two 32-bit
            // stores: four stores per block cost the shared-memory pipe more than the shuffle (7.5x +1.2 %, 1.5x +5 %).
            constexpr bool kHalfStores = NCH1 == 2;
            uint32_t vrow = vbase + buf * 2u * vplane + uint32_t(8 * ch + 2 * t + (kHalfStores ? 0 : hh)) * p.pitch_v + uint32_t(ch) * 4u + boff0 / 4u +
                            (kHalfStores ? uint32_t(hh) * 2u : 0u);
            uint32_t vrow_hi = vrow + vplane;
            const uint32_t pv = p.pitch_v;
            asm volatile("" : "+r"(vrow), "+r"(vrow_hi));

#pragma unroll 1
            for (int q = 0; q < ((p.debug & 1u) ? 0 : upw); q++) {
                const uint32_t boff = uint32_t(q) * (UNIT * 16u);
                int acc[UNIT][NDIG][4];
#pragma unroll
                for (int c = 0; c < NCH1; c++) {       // groups with shorter bands carry zero weights in the rest
                    uint32_t a[UNIT][4];
#pragma unroll
                    for (int u = 0; u < UNIT; u++) ldsm_t(a[u], rowoff[c] + boff + uint32_t(u) * 16u);
#pragma unroll
                    for (int d = 0; d < NDIG; d++)
#pragma unroll
                        for (int u = 0; u < UNIT; u++) {
                            if (c == 0 && d == 0) imma_c(acc[u][d], a[u], bw1[c][d], bias0, bias1);   // the sum starts at the bias
                            else if (c == 0) imma_z(acc[u][d], a[u], bw1[c][d]);
                            else imma(acc[u][d], a[u], bw1[c][d]);
                        }
                }
                // recombine the digit planes (exact), re-quantise to u16, split into limbs
#pragma unroll
                for (int u = 0; u < UNIT; u++) {
                    int v[4];
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        int x = acc[u][0][i] + (acc[u][1][i] << 8);
                        if (NDIG == 3) x += acc[u][NDIG - 1][i] << 16;
                        v[i] = x >> sh1;
                    }
                    // v[0],v[2]: pixel hh / hh+2 of output row 2t; v[1],v[3]: same pixels of row 2t+1
                    if (kHalfStores) {
                        const uint32_t o = boff / 4u + uint32_t(u) * 4u;
                        sts16(vrow + o, __byte_perm(uint32_t(v[0]), uint32_t(v[2]), 0x0040));            // row 2t, low limbs
                        sts16(vrow_hi + o, __byte_perm(uint32_t(v[0]), uint32_t(v[2]), 0x0051));         // row 2t, high limbs
                        sts16(vrow + pv + o, __byte_perm(uint32_t(v[1]), uint32_t(v[3]), 0x0040));       // row 2t + 1
                        sts16(vrow_hi + pv + o, __byte_perm(uint32_t(v[1]), uint32_t(v[3]), 0x0051));
                        continue;
                    }
                    const uint32_t w0 = __byte_perm(uint32_t(v[0]), uint32_t(v[2]), 0x5140);   // lo lo hi hi of row 2t
                    const uint32_t w1r = __byte_perm(uint32_t(v[1]), uint32_t(v[3]), 0x5140);  // ... of row 2t+1
                    const uint32_t send = hh ? w0 : w1r, mine = hh ? w1r : w0;
                    const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 16);
                    const uint32_t pa = hh ? recv : mine, pb = hh ? mine : recv;              // pixels (0,2) and (1,3)
                    sts32(vrow + boff / 4u + uint32_t(u) * 4u, __byte_perm(pa, pb, 0x5410));       // low limbs
                    sts32(vrow_hi + boff / 4u + uint32_t(u) * 4u, __byte_perm(pa, pb, 0x7632));    // high limbs
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&tail.v_full[buf]);        // this warp's share of V[buf] is written
            buf += uint32_t(nco);
            while (buf >= p.vbufs) { buf -= p.vbufs; par ^= 1u; }
        }
        VF_REPORT("pass1    ")
    }
}

}  // namespace

// cuTensorMapEncodeTiled through the runtime's driver entry point (the library does not link libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
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

// the source batch as one tall 2-D tensor of 8-byte elements; box = one ring step (pitch bytes x 16 rows)
static cudaError_t make_source_map(const VfParams &p, const uint8_t *src, uint32_t n_images, CUtensorMap &map) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return cudaErrorNotSupported;
    const cuuint64_t dims[2] = {cuuint64_t(p.sw) / 2, cuuint64_t(p.sh) * n_images};
    const cuuint64_t strides[1] = {cuuint64_t(p.sw) * 4};
    const cuuint32_t box[2] = {p.pitch / 8, cuuint32_t(kStepRows)};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, const_cast<uint8_t *>(src), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

template <int NDIG, int NCH1, int NCH2, int UNIT>
static cudaError_t launch_vf_t(const VfParams &p, uint32_t n_images, cudaStream_t stream) {
    static thread_local int configured_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (configured_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(resize_vfirst_kernel<NDIG, NCH1, NCH2, UNIT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             227 * 1024);
        if (e != cudaSuccess) return e;
        configured_dev = dev;
    }
    VfParams q = p;
    for (uint32_t first = 0; first < n_images; first += 65535) {
        q.src = p.src + size_t(first) * p.src_img_bytes;
        q.dst = p.dst + size_t(first) * p.dst_img_bytes;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        CUtensorMap map;
        cudaError_t e = make_source_map(q, q.src, nz, map);
        if (e != cudaSuccess) return e;
        // fill CTAs: one per ~2 MB of canvas that is not computed (at least one when anything is), after the strips of the image
        unsigned nfill = 0;
        if (p.need_fill) {
            const uint64_t bytes = (uint64_t(p.fw) * p.fh - uint64_t(p.pw) * p.ph) * 4u;
            nfill = unsigned(std::min<uint64_t>(32, 1 + bytes / (2u << 20)));
        }
        resize_vfirst_kernel<NDIG, NCH1, NCH2, UNIT><<<dim3(p.n_strips + nfill, p.n_bands, nz), kThreads, p.smem_bytes, stream>>>(q, map);
        count_launch();
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

size_t vfirst_smem_tail_bytes() { return sizeof(SmemTail); }

cudaError_t launch_resize_vfirst(const VfParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    if (p.n_bands == 0 || p.n_bands > 65535 || p.n_strips == 0 || p.nch1 > 6 || p.nch2 > 6) return cudaErrorInvalidValue;
    if (p.ring_rows % kStepRows != 0 || p.ring_rows / kStepRows > kMaxSteps) return cudaErrorInvalidValue;
    if (p.cohorts < 1 || p.cohorts > 3 || p.vbufs < 2) return cudaErrorInvalidValue;
#define SLIMG_VF_CASE(A, B, U) if (p.nch1 == A && p.nch2 == B && p.unit == U) return launch_vf_t<2, A, B, U>(p, n_images, stream);
#define SLIMG_VF_ROW(A, U) SLIMG_VF_CASE(A, 1, U) SLIMG_VF_CASE(A, 2, U) SLIMG_VF_CASE(A, 3, U) SLIMG_VF_CASE(A, 4, U)
#define SLIMG_VF_ALL(U) SLIMG_VF_ROW(1, U) SLIMG_VF_ROW(2, U) SLIMG_VF_ROW(3, U) SLIMG_VF_ROW(4, U)
    if (p.ndig != 2) return cudaErrorInvalidValue;
    SLIMG_VF_ALL(3)
    SLIMG_VF_ALL(4)
    SLIMG_VF_ALL(5)
    // bands of 5 - 6 chunks (downscales of ~9x to ~11x: thumbnails of 12 MP photos), UNIT = 4 only
    SLIMG_VF_CASE(5, 5, 4) SLIMG_VF_CASE(5, 6, 4) SLIMG_VF_CASE(6, 5, 4) SLIMG_VF_CASE(6, 6, 4)
    SLIMG_VF_CASE(4, 5, 4) SLIMG_VF_CASE(5, 4, 4)
    return cudaErrorInvalidValue;
}

}  // namespace slimg
