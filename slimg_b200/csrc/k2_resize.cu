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
    for (uint32_t first = 0; first < n_images; first += 65535) {
        ResizeParams q = p;
        q.src += size_t(first) * p.src_img_bytes;
        q.dst += size_t(first) * p.dst_img_bytes;
        const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
        resize_generic_kernel<<<dim3(tiles_x, tiles_y, nz), kThreads, smem, stream>>>(q);
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace slimg
