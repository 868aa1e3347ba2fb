// k2_resize.cu -- Lanczos3 separable resize with crop/extend fused in as address remapping.
//
// Replaces slimg_core::resize::resize (crates/slimg-core/src/resize.rs:64-90), i.e. image
// 0.25.9's imageops::resize(.., Lanczos3): vertical_sample into an f32 image, then
// horizontal_sample -> clamp [0,255] -> round half away from zero -> u8.  Same taps and
// weights (built on the host, taps.cpp), f32 accumulation, intermediate never leaves the SM.
//
// Pre remap (the `convert` order, pipeline.rs:71-89): the resize input is a virtual canvas;
// taps that land on its fill colour are folded into one constant per output row / column
// (AxisDev::fillw), so fill pixels take part in the filter exactly as the reference's
// materialised canvas does but cost no loads.  Post remap (resize -> crop -> extend): only the
// kept rectangle of the resize output is computed, and it is stored straight into its place
// on the final canvas; the rest of the canvas is written with the fill colour by the same
// launch.
//
// This file holds the GENERIC kernel: any ratio on either axis (up, down, mixed).  One CTA =
// one tile of the final canvas.  Stage 1 (vertical taps) gathers u8 source pixels into an f32
// tile in shared memory, stage 2 (horizontal taps) reads that tile.
#include <algorithm>
#include <climits>
#include <cstdlib>

#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float4 unpack_rgba(uint32_t v) {
    return make_float4(float(v & 0xffu), float((v >> 8) & 0xffu), float((v >> 16) & 0xffu), float(v >> 24));
}

// clamp(t, 0, 255) then f32::round (half away from zero; the value is non-negative here)
__device__ __forceinline__ uint32_t quantize_channel(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f);
    return uint32_t(floorf(t + 0.5f));
}

__device__ __forceinline__ uint32_t pack_rgba(const float4 &t) {
    return quantize_channel(t.x) | (quantize_channel(t.y) << 8) | (quantize_channel(t.z) << 16) |
           (quantize_channel(t.w) << 24);
}

__global__ void __launch_bounds__(kThreads) resize_generic_kernel(const ResizeParams p) {
    extern __shared__ float4 tile[];   // [tile_h][span_max]

    const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(blockIdx.z) * p.src_img_bytes);
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);

    // this tile of the final canvas, and its intersection with the placed rectangle of R
    const int fx0 = blockIdx.x * p.tile_w, fy0 = blockIdx.y * p.tile_h;
    const int fx1 = min(fx0 + p.tile_w, int(p.fw)), fy1 = min(fy0 + p.tile_h, int(p.fh));
    const int ix0 = max(fx0, int(p.fox)), ix1 = min(fx1, int(p.fox + p.pw));
    const int iy0 = max(fy0, int(p.foy)), iy1 = min(fy1, int(p.foy + p.ph));
    const bool has_pixels = ix0 < ix1 && iy0 < iy1;

    const int ncols = has_pixels ? ix1 - ix0 : 0, nrows = has_pixels ? iy1 - iy0 : 0;
    const int ox0 = ix0 - int(p.fox) + int(p.px);     // first R column / row computed here
    const int oy0 = iy0 - int(p.foy) + int(p.py);
    const int xs0 = has_pixels ? p.tile_x0[blockIdx.x] : 0;
    const int span = has_pixels ? p.tile_span[blockIdx.x] : 0;

    // stage 1: vertical taps.  item = (output row r, source column c); c is the fast index so a
    // warp reads 128 contiguous bytes of each source row.
    for (int item = threadIdx.x; item < nrows * span; item += kThreads) {
        const int r = item / span, c = item - r * span;
        const int oy = oy0 + r;
        const int n = p.ay.count[oy];
        const float *w = p.ay.w + size_t(oy) * p.ay.max_taps;
        const uint32_t *s = src + size_t(p.ry + p.ay.left[oy]) * p.sw + (p.rx + xs0 + c);
        const float fwgt = p.ay.fillw[oy];
        float4 acc = make_float4(p.fill_pre[0] * fwgt, p.fill_pre[1] * fwgt, p.fill_pre[2] * fwgt,
                                 p.fill_pre[3] * fwgt);
        for (int i = 0; i < n; i++) {
            const float wi = w[i];
            const float4 v = unpack_rgba(__ldg(s + size_t(i) * p.sw));
            acc.x = fmaf(v.x, wi, acc.x); acc.y = fmaf(v.y, wi, acc.y);
            acc.z = fmaf(v.z, wi, acc.z); acc.w = fmaf(v.w, wi, acc.w);
        }
        tile[r * p.span_max + c] = acc;
    }
    __syncthreads();

    // stage 2: horizontal taps + store, and the fill of canvas pixels outside the rectangle
    const int tw = fx1 - fx0, th = fy1 - fy0;
    for (int item = threadIdx.x; item < tw * th; item += kThreads) {
        const int ty = item / tw, tx = item - ty * tw;
        const int fx = fx0 + tx, fy = fy0 + ty;
        uint32_t out = p.fill_post;
        if (fx >= ix0 && fx < ix1 && fy >= iy0 && fy < iy1) {
            const int ox = ox0 + (fx - ix0), r = fy - iy0;
            const int n = p.ax.count[ox];
            const float *w = p.ax.w + size_t(ox) * p.ax.max_taps;
            const float4 *t = tile + r * p.span_max + (p.ax.left[ox] - xs0);
            const float fwgt = p.ax.fillw[ox];
            float4 acc = make_float4(p.fill_pre[0] * fwgt, p.fill_pre[1] * fwgt, p.fill_pre[2] * fwgt,
                                     p.fill_pre[3] * fwgt);
            for (int i = 0; i < n; i++) {
                const float wi = w[i];
                const float4 v = t[i];
                acc.x = fmaf(v.x, wi, acc.x); acc.y = fmaf(v.y, wi, acc.y);
                acc.z = fmaf(v.z, wi, acc.z); acc.w = fmaf(v.w, wi, acc.w);
            }
            out = pack_rgba(acc);
        }
        dst[size_t(fy) * p.fw + fx] = out;
    }
}

// ---- few-tap variant: both axes have at most kNT taps (upscales, ratios up to ~1.1) ------------------------------------
// Same taps, same f32 accumulation order as the kernel above (bit-identical results), organised for the case where the
// OUTPUT dominates (config C4b, 1024^2 x 4: 4 MB in, 67 MB out, 7 taps per axis):
//   stage 0  the source pixels the tile needs are converted to f32 ONCE (byte -> float through the mantissa, no I2F)
//   stage 1  vertical taps out of that tile, weights of the tile's rows staged in shared memory
//   stage 2  a thread owns one output column and walks down the tile's rows with its horizontal weights in registers;
//            clamp + round half away without F2I (add 0.5, then a round-toward-zero add of 2^23 leaves the integer in
//            the mantissa)
constexpr int kNT = 8;                 // most taps per axis the few-tap kernel takes (template NT = the plan's max)

__device__ __forceinline__ float4 unpack_rgba_fast(uint32_t v) {
    // 0x4B000000 | byte is the float 2^23 + byte, exactly
    return make_float4(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7440)) - 8388608.0f,
                       __uint_as_float(__byte_perm(v, 0x4B000000u, 0x7441)) - 8388608.0f,
                       __uint_as_float(__byte_perm(v, 0x4B000000u, 0x7442)) - 8388608.0f,
                       __uint_as_float(__byte_perm(v, 0x4B000000u, 0x7443)) - 8388608.0f);
}
// clamp(t, 0, 255), then floor(t + 0.5): the byte ends up in the low mantissa bits
__device__ __forceinline__ uint32_t quantize_bits(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f) + 0.5f;
    return __float_as_uint(__fadd_rz(t, 8388608.0f));
}
__device__ __forceinline__ uint32_t pack_rgba_fast(const float4 &t) {
    const uint32_t xy = __byte_perm(quantize_bits(t.x), quantize_bits(t.y), 0x0040);
    const uint32_t zw = __byte_perm(quantize_bits(t.z), quantize_bits(t.w), 0x0040);
    return __byte_perm(xy, zw, 0x5410);
}

// Taps beyond a row's / column's count carry zero weights (taps.cpp pads them) and read zero-filled padding of the
// tiles, so the tap loops run NT unconditional iterations: no predicates, no branches.
template <int NT, int K>
__global__ void __launch_bounds__(kThreads) resize_fewtaps_kernel(const ResizeParams p, const int src_rows_max) {
    extern __shared__ float4 smem_f4[];
    const int vpitch = p.span_max + NT + 2;
    float4 *S = smem_f4;                                   // [src_rows_max + NT][span_max] source pixels as f32
    float4 *V = S + size_t(src_rows_max + NT) * p.span_max;     // [tile_h][span_max + NT] vertical-pass result
    float *wy = reinterpret_cast<float *>(V + size_t(p.tile_h) * vpitch);   // [tile_h][NT]

    const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(blockIdx.z) * p.src_img_bytes);
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);
    const int fx0 = blockIdx.x * p.tile_w, fy0 = blockIdx.y * p.tile_h;
    const int fx1 = min(fx0 + p.tile_w, int(p.fw)), fy1 = min(fy0 + p.tile_h, int(p.fh));
    const int ix0 = max(fx0, int(p.fox)), ix1 = min(fx1, int(p.fox + p.pw));
    const int iy0 = max(fy0, int(p.foy)), iy1 = min(fy1, int(p.foy + p.ph));
    const bool has_pixels = ix0 < ix1 && iy0 < iy1;
    const int nrows = has_pixels ? iy1 - iy0 : 0;
    const int ox0 = ix0 - int(p.fox) + int(p.px), oy0 = iy0 - int(p.foy) + int(p.py);
    const int xs0 = has_pixels ? p.tile_x0[blockIdx.x] : 0;
    const int span = has_pixels ? p.tile_span[blockIdx.x] : 0;
    const int nt_y = p.ay.max_taps, nt_x = p.ax.max_taps;

    // source rows of the tile (one thread per output row, reduced through shared memory); the row's band start and
    // fill weight are kept in shared memory so that stage 1 starts a row without waiting for global loads
    __shared__ int s_ylo, s_yhi;
    __shared__ int s_left[64];
    __shared__ float s_fillw[64];
    if (threadIdx.x == 0) { s_ylo = INT_MAX; s_yhi = 0; }
    __syncthreads();
    if (int(threadIdx.x) < nrows) {
        const int n = p.ay.count[oy0 + threadIdx.x], l = p.ay.left[oy0 + threadIdx.x];
        s_left[threadIdx.x] = n > 0 ? l : -1;
        s_fillw[threadIdx.x] = p.ay.fillw[oy0 + threadIdx.x];
        if (n > 0) { atomicMin(&s_ylo, l); atomicMax(&s_yhi, l + n); }
    }
    // the vertical weights of the tile's rows (independent of the reduction: in flight across the barrier)
    for (int item = threadIdx.x; item < nrows * NT; item += kThreads) {
        const int r = item / NT, i = item - r * NT;
        wy[item] = i < nt_y ? p.ay.w[size_t(oy0 + r) * nt_y + i] : 0.0f;
    }
    __syncthreads();
    const int ylo = s_ylo == INT_MAX ? 0 : s_ylo, yhi = s_ylo == INT_MAX ? 0 : s_yhi;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kW = kThreads / 32;
    // stage 0: source tile as f32, a warp per source row, four rows' loads in flight per lane
    {
        const int nreal = yhi - ylo, nall = nreal + NT;
        const uint32_t *base = src + size_t(p.ry + ylo) * p.sw + (p.rx + xs0);
        for (int c = lane; c < span; c += 32)
            for (int r0 = warp; r0 < nall; r0 += 4 * kW) {
                uint32_t v[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int r = r0 + u * kW;
                    v[u] = r < nreal ? __ldg(base + size_t(r) * p.sw + c) : 0u;
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int r = r0 + u * kW;
                    if (r < nall) S[r * p.span_max + c] = r < nreal ? unpack_rgba_fast(v[u]) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                }
            }
    }
    __syncthreads();

    // stage 1: vertical taps, a warp per output row (row constants are warp-uniform)
    for (int r = warp; r < nrows; r += kW) {
        const int l = s_left[r];
        const float4 *s0 = S + (l >= 0 ? l - ylo : 0) * p.span_max;
        float w[NT];
#pragma unroll
        for (int i = 0; i < NT; i++) w[i] = wy[r * NT + i];
        const float fwgt = s_fillw[r];
        const float4 acc0 = make_float4(p.fill_pre[0] * fwgt, p.fill_pre[1] * fwgt, p.fill_pre[2] * fwgt, p.fill_pre[3] * fwgt);
        for (int c = lane; c < span + NT + 2; c += 32) {
            float4 acc = acc0;
            if (c < span) {
#pragma unroll
                for (int i = 0; i < NT; i++) {
                    const float4 v = s0[i * p.span_max + c];
                    acc.x = fmaf(v.x, w[i], acc.x); acc.y = fmaf(v.y, w[i], acc.y);
                    acc.z = fmaf(v.z, w[i], acc.z); acc.w = fmaf(v.w, w[i], acc.w);
                }
            } else acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);     // padding read by zero-weight taps
            V[r * vpitch + c] = acc;
        }
    }
    __syncthreads();

    // stage 2: a thread owns K adjacent canvas columns and every (256 * K / tile_w)-th row of the tile.  With K = 4
    // (upscales: neighbouring output columns read almost the same taps) the U = NT + 2 intermediate values of the
    // columns' common window are loaded once per row and reused by all four columns -- the kernel is bound by shared-
    // memory bandwidth otherwise (one 16-byte load per 4 FMAs).  Weights are re-based to the common window, zero outside
    // a column's own taps: extra zero-weight FMAs do not change the f32 sums.
    constexpr int U = K == 1 ? NT : K == 2 ? NT + 1 : NT + 2;
    const int tw = fx1 - fx0, th = fy1 - fy0;
    const int ngx = p.tile_w / K;
    const int gx = threadIdx.x % ngx, ty0 = threadIdx.x / ngx, tys = kThreads / ngx;
    const int tx = gx * K;
    if (tx < tw) {
        float wx[K][U];
        float4 acc0[K];                 // the fill's share of the column's weights: the sum starts there
        bool valid[K];
        int toff = INT_MAX;
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int fx = fx0 + tx + k;
            valid[k] = has_pixels && fx >= ix0 && fx < ix1;
            float fwgt = 0.0f;
            if (valid[k]) {
                const int ox = ox0 + (fx - ix0);
                fwgt = p.ax.fillw[ox];
                if (p.ax.count[ox] > 0) toff = min(toff, p.ax.left[ox] - xs0);
            }
            acc0[k] = make_float4(p.fill_pre[0] * fwgt, p.fill_pre[1] * fwgt, p.fill_pre[2] * fwgt, p.fill_pre[3] * fwgt);
        }
        if (toff == INT_MAX) toff = 0;
#pragma unroll
        for (int k = 0; k < K; k++) {
#pragma unroll
            for (int j = 0; j < U; j++) wx[k][j] = 0.0f;
            if (valid[k]) {
                const int ox = ox0 + (fx0 + tx + k - ix0);
                const int d = p.ax.count[ox] > 0 ? p.ax.left[ox] - xs0 - toff : 0;      // 0..U-NT by the plan's check
#pragma unroll
                for (int j = 0; j < U; j++) {
                    const int i = j - d;
                    if (i >= 0 && i < nt_x) wx[k][j] = p.ax.w[size_t(ox) * nt_x + i];
                }
            }
        }
        // rows [vy0, vy1) of the tile carry pixels, the others are fill; addresses advance by constants
        const int vy0 = has_pixels ? iy0 - fy0 : th, vy1 = has_pixels ? iy1 - fy0 : th;
        int vi = (ty0 - vy0) * vpitch + toff;
        const int vstep = tys * vpitch;
        uint32_t *o = dst + size_t(fy0 + ty0) * p.fw + fx0 + tx;
        const size_t ostep = size_t(tys) * p.fw;
        const bool vec_ok = K > 1 && tx + K - 1 < tw && ((reinterpret_cast<uintptr_t>(o) | (ostep * 4)) & (4 * K - 1)) == 0;
        for (int ty = ty0; ty < th; ty += tys, vi += vstep, o += ostep) {
            uint32_t out[K];
#pragma unroll
            for (int k = 0; k < K; k++) out[k] = p.fill_post;
            if (ty >= vy0 && ty < vy1) {
                const float4 *t = V + vi;
                float4 acc[K];
#pragma unroll
                for (int k = 0; k < K; k++) acc[k] = acc0[k];
#pragma unroll
                for (int j = 0; j < U; j++) {
                    const float4 v = t[j];
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        acc[k].x = fmaf(v.x, wx[k][j], acc[k].x); acc[k].y = fmaf(v.y, wx[k][j], acc[k].y);
                        acc[k].z = fmaf(v.z, wx[k][j], acc[k].z); acc[k].w = fmaf(v.w, wx[k][j], acc[k].w);
                    }
                }
#pragma unroll
                for (int k = 0; k < K; k++)
                    if (valid[k]) out[k] = pack_rgba_fast(acc[k]);
            }
            if (K == 4 && vec_ok) {
                *reinterpret_cast<uint4 *>(o) = make_uint4(out[0], out[K > 1 ? 1 : 0], out[K > 2 ? 2 : 0], out[K > 3 ? 3 : 0]);
            } else if (K == 2 && vec_ok) {
                *reinterpret_cast<uint2 *>(o) = make_uint2(out[0], out[K > 1 ? 1 : 0]);
            } else {
#pragma unroll
                for (int k = 0; k < K; k++)
                    if (tx + k < tw) o[k] = out[k];
            }
        }
    }
}

}  // namespace

cudaError_t launch_resize_generic(const ResizeParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    const unsigned tiles_x = (p.fw + p.tile_w - 1) / p.tile_w, tiles_y = (p.fh + p.tile_h - 1) / p.tile_h;
    if (tiles_y > 65535) return cudaErrorInvalidValue;
    const size_t smem = size_t(p.tile_h) * size_t(p.span_max > 0 ? p.span_max : 1) * sizeof(float4);
    static thread_local int configured_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (configured_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(resize_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return e;
        configured_dev = dev;
    }
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    // few taps on both axes (upscales): the register-weight kernel; its source tile holds at most tile_h + kNT rows
    const bool few = p.few_rows_max > 0 && p.tile_h <= 64 && p.ax.max_taps <= kNT && p.ay.max_taps <= kNT && kThreads % p.tile_w == 0 && !std::getenv("SLIMG_B200_NO_FEWTAPS");
    const int src_rows_max = p.few_rows_max;
    const int nt = std::max(p.ax.max_taps, p.ay.max_taps) <= 7 ? 7 : 8;
    const size_t span_alloc = size_t(p.span_max > 0 ? p.span_max : 1);
    const size_t smem_few = (size_t(src_rows_max + nt) * span_alloc + size_t(p.tile_h) * (span_alloc + nt + 2)) * sizeof(float4) +
                            size_t(p.tile_h) * nt * sizeof(float);
    if (few && smem_few <= 200 * 1024) {
        const int k = (p.few_k == 4 || p.few_k == 2) && p.tile_w % p.few_k == 0 && kThreads % (p.tile_w / p.few_k) == 0 ? p.few_k : 1;
        auto kern = nt == 7 ? (k == 4 ? resize_fewtaps_kernel<7, 4> : k == 2 ? resize_fewtaps_kernel<7, 2> : resize_fewtaps_kernel<7, 1>)
                            : (k == 4 ? resize_fewtaps_kernel<8, 4> : k == 2 ? resize_fewtaps_kernel<8, 2> : resize_fewtaps_kernel<8, 1>);
        static thread_local int few_dev = -1;
        if (few_dev != dev) {
            cudaError_t e = cudaFuncSetAttribute(resize_fewtaps_kernel<7, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(resize_fewtaps_kernel<8, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(resize_fewtaps_kernel<7, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(resize_fewtaps_kernel<8, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(resize_fewtaps_kernel<7, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(resize_fewtaps_kernel<8, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e != cudaSuccess) return e;
            few_dev = dev;
        }
        for (uint32_t first = 0; first < n_images; first += 65535) {
            ResizeParams q = p;
            q.src += size_t(first) * p.src_img_bytes;
            q.dst += size_t(first) * p.dst_img_bytes;
            const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
            kern<<<dim3(tiles_x, tiles_y, nz), kThreads, smem_few, stream>>>(q, src_rows_max);
            count_launch();
            note_path(SLIMG_PATH_FEWTAPS);
            cudaError_t e = cudaGetLastError();
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    }
    for (uint32_t first = 0; first < n_images; first += 65535) {
        ResizeParams q = p;
        q.src += size_t(first) * p.src_img_bytes;
        q.dst += size_t(first) * p.dst_img_bytes;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        resize_generic_kernel<<<dim3(tiles_x, tiles_y, nz), kThreads, smem, stream>>>(q);
        count_launch();
        note_path(SLIMG_PATH_GENERIC);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace slimg
