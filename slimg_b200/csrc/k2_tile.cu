// k2_tile.cu -- Lanczos3 at ratios below ~3.5x and for upscales: both passes as integer tensor-core GEMMs out of one
// shared-memory tile (BASELINE config C4b, the 2x "retina" downscale, thumbnails of small images).
//
// Replaces image 0.25.9's vertical_sample / horizontal_sample pair behind slimg_core::resize::resize
// (crates/slimg-core/src/resize.rs:64-90), crop / extend fused in as address arithmetic (pipeline.rs:71-89).  Same
// taps, weights and pass order as the reference (taps.cpp), in the fixed point of vplan.cpp / tileplan.cpp.
//
// One CTA = a tile of tw x th kept output pixels of one image:
//   stage   cp.async (16 bytes, L1 bypass) of the tile's source footprint into shared memory: rows_max x pitch bytes
//   pass 1  per group of 8 output rows:  D1[16 bytes of the row][8 rows] += A[16][32 source rows] * B[32][8]
//           A straight from the staged rows with ldmatrix.m16n16.x2.trans.b8, B = weight digits from the plan (L2);
//           digit recombination, >> to the 16-bit intermediate, two u8 limb planes per group in shared memory
//   pass 2  per group and tile of 8 output columns:  D2[(channel,row) 16][8 columns] += A[16][32 px] * B[32][8] for
//           both limbs (upscales: both limbs in ONE K = 32, three MMAs instead of four); the limb / digit planes are
//           recombined in integers, the constant term (fill taps, offset, rounding half) added, shifted, saturated and
//           packed two channels per cvt.pack.sat; RGBA stores straight to the final canvas position
// Two __syncthreads in all; several CTAs per SM overlap each other's staging and arithmetic.  Everything the tile does
// not compute (canvas outside the kept rectangle, kept rows whose taps are all fill) is written by fill CTAs
// (blockIdx.x == n_colblk), one canvas row per warp, 16-byte stores.
#include <algorithm>

#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kWarps = 8, kThreads = kWarps * 32;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void ldsm_t(uint32_t (&a)[4], uint32_t addr) {      // see k2_vfirst.cu
    asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];"
                 : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]) : "r"(addr));
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
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
// d = (c << 16) | (sat_u8(hi) << 8) | sat_u8(lo): clamp to [0, 255] and pack, two channels per instruction
__device__ __forceinline__ uint32_t pack_sat_u8(int hi, int lo, uint32_t c) {
    uint32_t d;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(lo), "r"(c));
    return d;
}

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

template <int NCH1, int NCH2, bool PACKED>
__global__ void __launch_bounds__(kThreads, 3) resize_tile_kernel(const TileParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);

    if (blockIdx.x >= p.n_colblk) {              // fill CTA: canvas rows part, part + nparts, ... ; one row per warp
        const int fw = int(p.fw), kx0 = int(p.fox), kx1 = int(p.fox + p.pw);
        const bool all_computed = (p.need_fill & 2u) == 0;
#pragma unroll 1
        for (int fr = int(blockIdx.y) * kWarps + warp; fr < int(p.fh); fr += int(gridDim.y) * kWarps) {
            const int i = fr - int(p.foy);
            const bool kept = i >= 0 && i < int(p.ph);
            uint32_t *row = dst + size_t(fr) * fw;
            if (!kept) { fill_span(row, 0, fw, p.fill_post, lane); continue; }
            fill_span(row, 0, kx0, p.fill_post, lane);
            if (!(all_computed || p.row_computed[i])) fill_span(row, kx0, kx1, p.fill_pre_packed, lane);
            fill_span(row, kx1, fw, p.fill_post, lane);
        }
        return;
    }

    const int cb = blockIdx.x, rb = blockIdx.y;
    const int col0 = cb * int(p.tw), col1 = min(int(p.pw), col0 + int(p.tw));
    const int n_groups = (int(p.ph) + 7) >> 3;
    const int g0 = rb * int(p.th >> 3), ng = min(n_groups - g0, int(p.th >> 3));
    const int y0 = p.rb_y0[rb], nrows = p.rb_rows[rb];
    const int s0 = p.cb_src0[cb], nblk = p.cb_nblk[cb];
    if (nrows <= 0) return;                      // every row of the block is pure fill: the fill CTAs write it

    const uint32_t sbase = smem_u32(smem_raw);
    const uint32_t vbase = sbase + p.rows_max * p.pitch;
    const uint32_t vplane = 32u * p.pitch_v;     // one limb of one group

    // ---- work split (ng, T <= 8: powers of two by comparison, no integer division) and everything a warp reads from
    // global memory for its FIRST group / tile, issued before the staging copies so the latencies overlap -------------
    const int wpg = ng == 1 ? 8 : ng == 2 ? 4 : ng <= 4 ? 2 : 1;        // warps that share a group in pass 1
    const int wpg_sh = ng == 1 ? 3 : ng == 2 ? 2 : ng <= 4 ? 1 : 0;
    const int T = int(p.tw >> 3);
    const int wpt = T == 1 ? 8 : T == 2 ? 4 : T <= 4 ? 2 : 1;           // warps that share a column tile in pass 2
    const int wpt_sh = T == 1 ? 3 : T == 2 ? 2 : T <= 4 ? 1 : 0;
    // groups whose rows are not pure fill, as a bit mask
    uint32_t gmask;
    {
        const uint32_t nchw = lane < ng ? __ldg(p.recs + size_t(g0 + lane) * p.rec_words + 2) : 0u;
        gmask = __ballot_sync(0xffffffffu, nchw != 0u);
    }
    const int gi_first = warp >> wpg_sh;
    constexpr int NB2 = PACKED ? 3 : NCH2 * 2;
    const int tile_first = warp >> wpt_sh;
    const int ntile = (col1 - col0 + 7) >> 3;
    uint2 bw2_first[NB2];
    int kb0_first = 0;
    if (tile_first < ntile) {
        kb0_first = __ldg(p.tile_kb0 + cb * T + tile_first);
#pragma unroll
        for (int i = 0; i < NB2; i++) bw2_first[i] = __ldg(reinterpret_cast<const uint2 *>(p.b2) + (size_t(cb * T + tile_first) * NB2 + i) * 32 + lane);
    }

    // ---- stage the source footprint: rows [y0, y0 + nrows) of the rectangle, pixels [s0, s0 + 4 nblk) ----------
    {
        const uint8_t *src = p.src + size_t(blockIdx.z) * p.src_img_bytes;
        // lanes across the blocks of a row (and, when a row has at most 16 blocks, across 2..16 rows): no division
        const int bshift = nblk <= 1 ? 0 : min(5, 32 - __clz(nblk - 1));
        const int bw = 1 << bshift, rpp = 32 >> bshift;
        const int b0 = lane & (bw - 1);
#pragma unroll 1
        for (int r = warp * rpp + (lane >> bshift); r < nrows; r += kWarps * rpp) {
            const uint8_t *srow = src + size_t(p.ry + y0 + r) * p.sw * 4;
            const uint32_t drow = sbase + uint32_t(r) * p.pitch;
            for (int b = b0; b < nblk; b += bw) {
                const int px = s0 + 4 * b;
                if (px + 4 <= int(p.sw)) cp_async16(drow + uint32_t(b) * 16u, srow + size_t(px) * 4);
                else *reinterpret_cast<uint4 *>(smem_raw + size_t(r) * p.pitch + size_t(b) * 16) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();

    // ---- pass 1: vertical taps, 8 output rows x pairs of 16-byte blocks ------------------------------------------
    {
        const int gstep = kWarps >> wpg_sh;
        const int ch = g & 3, hh = g >> 2;
        const int sh1 = p.sh1;
        const int npair = nblk >> 1;
#pragma unroll 1
        for (int gi = gi_first; gi < ng; gi += gstep) {
            if (!((gmask >> gi) & 1u)) continue;         // pure-fill rows
            const uint32_t *rec = p.recs + size_t(g0 + gi) * p.rec_words;
            const int a0 = int(__ldg(rec));
            uint2 bw1[NCH1][2];
#pragma unroll
            for (int c = 0; c < NCH1; c++)
#pragma unroll
                for (int d = 0; d < 2; d++) bw1[c][d] = __ldg(reinterpret_cast<const uint2 *>(rec + 64) + (c * 2 + d) * 32 + lane);
            const int bias0 = int(__ldg(rec + 8 + (2 * t) * 4 + ch)), bias1 = int(__ldg(rec + 8 + (2 * t + 1) * 4 + ch));
            uint32_t rowoff[NCH1];
#pragma unroll
            for (int c = 0; c < NCH1; c++) rowoff[c] = sbase + uint32_t(min(a0 - y0 + 32 * c + lane, nrows - 1)) * p.pitch;
            const uint32_t vrow = vbase + uint32_t(gi) * 2u * vplane + uint32_t(8 * ch + 2 * t + hh) * p.pitch_v + uint32_t(ch) * 4u;
#pragma unroll 1
            for (int q = warp & (wpg - 1); q < npair; q += wpg) {
                const uint32_t boff = uint32_t(q) * 32u;
                int acc[2][2][4];
#pragma unroll
                for (int c = 0; c < NCH1; c++) {         // groups with shorter bands carry zero weights in the rest
                    uint32_t a[2][4];
#pragma unroll
                    for (int u = 0; u < 2; u++) ldsm_t(a[u], rowoff[c] + boff + uint32_t(u) * 16u);
#pragma unroll
                    for (int d = 0; d < 2; d++)
#pragma unroll
                        for (int u = 0; u < 2; u++) {
                            if (c == 0 && d == 0) imma_c(acc[u][d], a[u], bw1[c][d], bias0, bias1);   // the sum starts at the bias
                            else if (c == 0) imma_z(acc[u][d], a[u], bw1[c][d]);
                            else imma(acc[u][d], a[u], bw1[c][d]);
                        }
                }
#pragma unroll
                for (int u = 0; u < 2; u++) {            // recombine the digit planes (exact), re-quantise to u16, split into limbs
                    int v[4];
#pragma unroll
                    for (int i = 0; i < 4; i++) v[i] = (acc[u][0][i] + (acc[u][1][i] << 8)) >> sh1;
                    // v[0],v[2]: pixel hh / hh+2 of output row 2t; v[1],v[3]: same pixels of row 2t+1
                    const uint32_t w0 = __byte_perm(uint32_t(v[0]), uint32_t(v[2]), 0x5140);
                    const uint32_t w1r = __byte_perm(uint32_t(v[1]), uint32_t(v[3]), 0x5140);
                    const uint32_t send = hh ? w0 : w1r, mine = hh ? w1r : w0;
                    const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 16);
                    const uint32_t pa = hh ? recv : mine, pb = hh ? mine : recv;              // pixels (0,2) and (1,3)
                    const uint32_t word = (uint32_t(q) * 2u + uint32_t(u)) * 4u;
                    sts32(vrow + word, __byte_perm(pa, pb, 0x5410));                 // low limbs
                    sts32(vrow + vplane + word, __byte_perm(pa, pb, 0x7632));        // high limbs
                }
            }
        }
    }
    __syncthreads();

    // ---- pass 2: horizontal taps, 4 channels x 8 rows x 8 output columns per step ---------------------------------
    {
        const int tstep = kWarps >> wpt_sh;
        const int sh2 = p.sh2;
#pragma unroll 1
        for (int tile = tile_first; tile < ntile; tile += tstep) {
            int kb0;
            uint2 bw2[NB2];
            const int ocol = col0 + tile * 8 + 2 * t;          // this lane's two kept output columns
            if (tile == tile_first) {
                kb0 = kb0_first;
#pragma unroll
                for (int i = 0; i < NB2; i++) bw2[i] = bw2_first[i];
            } else {
                kb0 = p.tile_kb0[cb * T + tile];
#pragma unroll
                for (int i = 0; i < NB2; i++) bw2[i] = __ldg(reinterpret_cast<const uint2 *>(p.b2) + (size_t(cb * T + tile) * NB2 + i) * 32 + lane);
            }
            // constant term (fill taps, intermediate offset, rounding half) of columns ocol, ocol + 1: the initial value of the
            // 2^-sh2 accumulator, [column][channel]
            int cxr[2][4];
#pragma unroll
            for (int e = 0; e < 2; e++)
#pragma unroll
                for (int c = 0; c < 4; c++) cxr[e][c] = ocol + e < col1 ? __ldg(p.cxi + size_t(ocol + e) * 4 + c) : 0;
            // everything that does not depend on the group, once per tile: the V addresses of this lane's fragment rows, the
            // output pointer of its row in group 0, whether its two pixels go out as one 8-byte store
            uint32_t rbase[2];
#pragma unroll
            for (int mt = 0; mt < 2; mt++) rbase[mt] = vbase + uint32_t(16 * mt + g) * p.pitch_v + uint32_t(2 * mt) * 4u + uint32_t(kb0 + t) * 4u;
            const uint32_t r1off = 8u * p.pitch_v + 4u, gstride = 2u * vplane;
            uint32_t *const orow0 = dst + size_t(p.foy + g0 * 8 + g) * p.fw + p.fox + ocol;
            const uint32_t ostride = 8u * p.fw;                  // pixels between the rows of consecutive groups (64 fw < 2^32)
            const int rows_left = int(p.ph) - (g0 * 8 + g);      // kept rows from this lane's row of group 0 on
            const bool pair_ok = ocol + 1 < col1 && (reinterpret_cast<uintptr_t>(orow0) & 7) == 0 && (p.fw & 1u) == 0;
            const bool all_rows = (p.need_fill & 2u) == 0;
            uint32_t todo = gmask;                            // groups of this warp that hold computed rows
            if (wpt > 1) {
                uint32_t mine = 0u;
                for (int gi = warp & (wpt - 1); gi < ng; gi += wpt) mine |= 1u << gi;
                todo &= mine;
            }
#pragma unroll 1
            while (todo) {
                const int gi = __ffs(int(todo)) - 1;
                todo &= todo - 1u;
                const uint32_t goff = uint32_t(gi) * gstride;
                int tsum[2][4];                    // [channel pair][i]: the pass-2 sum in units of 2^-sh2, constant term included
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    const uint32_t r0 = rbase[mt] + goff;
                    const uint32_t r1 = r0 + r1off;
                    if (PACKED) {
                        // K = [16 low-limb pixels | 16 high-limb pixels]: A = lo x w0, B = lo x w1 + hi x w0, C = hi x w1
                        uint32_t a[4];
                        a[0] = lds32(r0); a[1] = lds32(r1); a[2] = lds32(r0 + vplane); a[3] = lds32(r1 + vplane);
                        int aa[4], ab[4], ac[4];
                        imma_z(aa, a, bw2[0]); imma_z(ac, a, bw2[2]);
                        imma_c4(ab, a, bw2[1], cxr[0][2 * mt], cxr[1][2 * mt], cxr[0][2 * mt + 1], cxr[1][2 * mt + 1]);   // the 2^-sh2 plane starts at the constant term
#pragma unroll
                        for (int i = 0; i < 4; i++) tsum[mt][i] = (aa[i] >> 8) + ab[i] + (ac[i] << 8);
                    } else {
                        int acc[2][2][4];          // [limb][digit]
#pragma unroll
                        for (int c = 0; c < NCH2; c++)
#pragma unroll
                            for (int l = 0; l < 2; l++) {
                                uint32_t a[4];
                                const uint32_t o = uint32_t(l) * vplane + uint32_t(c) * 32u;
                                a[0] = lds32(r0 + o); a[1] = lds32(r1 + o); a[2] = lds32(r0 + o + 16u); a[3] = lds32(r1 + o + 16u);
#pragma unroll
                                for (int d = 0; d < 2; d++) {
                                    if (c == 0 && l == 0 && d == 1)      // the 2^-sh2 plane starts at the constant term
                                        imma_c4(acc[l][d], a, bw2[c * 2 + d], cxr[0][2 * mt], cxr[1][2 * mt], cxr[0][2 * mt + 1], cxr[1][2 * mt + 1]);
                                    else if (c == 0) imma_z(acc[l][d], a, bw2[c * 2 + d]);
                                    else imma(acc[l][d], a, bw2[c * 2 + d]);
                                }
                            }
#pragma unroll
                        for (int i = 0; i < 4; i++) tsum[mt][i] = (acc[0][0][i] >> 8) + acc[0][1][i] + acc[1][0][i] + (acc[1][1][i] << 8);
                    }
                }
                // i = 0,1: channel 2mt, columns +0, +1; i = 2,3: channel 2mt+1.  Shift, saturate, pack RGBA.
                uint32_t px[2];
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const int c0 = tsum[0][e] >> sh2, c1 = tsum[0][2 + e] >> sh2, c2 = tsum[1][e] >> sh2, c3 = tsum[1][2 + e] >> sh2;
                    px[e] = pack_sat_u8(c1, c0, pack_sat_u8(c3, c2, 0u));
                }
                if (gi * 8 < rows_left && (all_rows || p.row_computed[(g0 + gi) * 8 + g])) {
                    uint32_t *o = orow0 + uint32_t(gi) * ostride;
                    if (pair_ok) *reinterpret_cast<uint2 *>(o) = make_uint2(px[0], px[1]);
                    else {
                        if (ocol < col1) o[0] = px[0];
                        if (ocol + 1 < col1) o[1] = px[1];
                    }
                }
            }
        }
    }
}

template <int NCH1, int NCH2, bool PACKED>
cudaError_t launch_tile_t(const TileParams &p, uint32_t n_images, cudaStream_t stream) {
    static thread_local int configured_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (configured_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(resize_tile_kernel<NCH1, NCH2, PACKED>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (e != cudaSuccess) return e;
        configured_dev = dev;
    }
    TileParams q = p;
    for (uint32_t first = 0; first < n_images; first += 65535) {
        q.src = p.src + size_t(first) * p.src_img_bytes;
        q.dst = p.dst + size_t(first) * p.dst_img_bytes;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        resize_tile_kernel<NCH1, NCH2, PACKED><<<dim3(p.n_colblk + (p.need_fill ? 1u : 0u), p.n_rowblk, nz), kThreads, p.smem_bytes, stream>>>(q);
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace

cudaError_t launch_resize_tile(const TileParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    if (p.n_rowblk == 0 || p.n_rowblk > 65535 || p.n_colblk == 0 || p.smem_bytes > 100 * 1024) return cudaErrorInvalidValue;
    if (p.packed) {
        if (p.nch2 != 1) return cudaErrorInvalidValue;
        switch (p.nch1) {
            case 1: return launch_tile_t<1, 1, true>(p, n_images, stream);
            case 2: return launch_tile_t<2, 1, true>(p, n_images, stream);
            case 3: return launch_tile_t<3, 1, true>(p, n_images, stream);
            case 4: return launch_tile_t<4, 1, true>(p, n_images, stream);
        }
        return cudaErrorInvalidValue;
    }
#define SLIMG_TILE_CASE(A, B) if (p.nch1 == A && p.nch2 == B) return launch_tile_t<A, B, false>(p, n_images, stream);
#define SLIMG_TILE_ROW(A) SLIMG_TILE_CASE(A, 1) SLIMG_TILE_CASE(A, 2) SLIMG_TILE_CASE(A, 3) SLIMG_TILE_CASE(A, 4)
    SLIMG_TILE_ROW(1) SLIMG_TILE_ROW(2) SLIMG_TILE_ROW(3) SLIMG_TILE_ROW(4)
    return cudaErrorInvalidValue;
}

}  // namespace slimg
