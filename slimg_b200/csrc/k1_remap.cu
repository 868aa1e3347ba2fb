// k1_remap.cu -- crop / extend (and crop->extend) as ONE remapped copy: every canvas pixel is
// written exactly once, every needed source pixel is read exactly once, nothing else touches
// HBM.  Replaces the row-memcpy loops of crates/slimg-core/src/crop.rs:70-87 and the
// zero-init + fill loop + row copies of extend.rs:94-133 (three passes over the canvas there).
//
// Roofline: pure HBM.  Algorithmic bytes = |A| * 4 read + fw*fh*4 written.
// Mapping: one thread = one 16-byte quad of 4 output pixels, ROWS_PER_THREAD rows deep so each
// thread keeps several independent 128-bit loads in flight; rows are walked by blockIdx.y, the
// image index by blockIdx.z.  When the canvas width is not a multiple of 4 the rows are not
// 16-byte aligned and the kernel degrades to 4-byte accesses (still coalesced).
#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kThreads = 256;
constexpr int kRowsPerThread = 4;

__device__ __forceinline__ uint4 ld_stream_v4(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t *p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_v4(uint4 *p, const uint4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ uint32_t remap_pixel(const RemapParams &p, const uint32_t *src, int x, int y) {
    if (x >= p.ax0 && x < p.ax1 && y >= p.ay0 && y < p.ay1)
        return ld_stream_u32(src + size_t(y + p.dy) * p.sw + size_t(x + p.dx));
    if (x >= p.bx0 && x < p.bx1 && y >= p.by0 && y < p.by1) return p.fill_in;
    return p.fill_out;
}

// canvas width % 4 == 0: 128-bit stores, 128-bit loads whenever the source quad is aligned
__global__ void __launch_bounds__(kThreads) remap_quads_kernel(const RemapParams p) {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(blockIdx.z) * p.src_img_bytes);
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);
    const int qx = blockIdx.x * kThreads + threadIdx.x;
    const int x = qx * 4;
    if (x >= int(p.fw)) return;
    const int y_base = blockIdx.y * kRowsPerThread;

    uint4 v[kRowsPerThread];
#pragma unroll
    for (int r = 0; r < kRowsPerThread; r++) {
        const int y = y_base + r;
        if (y >= int(p.fh)) break;
        const bool row_in_a = y >= p.ay0 && y < p.ay1;
        if (row_in_a && x >= p.ax0 && x + 4 <= p.ax1) {
            const uint32_t *s = src + size_t(y + p.dy) * p.sw + size_t(x + p.dx);
            if ((reinterpret_cast<uintptr_t>(s) & 15) == 0) {
                v[r] = ld_stream_v4(reinterpret_cast<const uint4 *>(s));
            } else {
                v[r].x = ld_stream_u32(s); v[r].y = ld_stream_u32(s + 1);
                v[r].z = ld_stream_u32(s + 2); v[r].w = ld_stream_u32(s + 3);
            }
        } else {
            v[r].x = remap_pixel(p, src, x, y);     v[r].y = remap_pixel(p, src, x + 1, y);
            v[r].z = remap_pixel(p, src, x + 2, y); v[r].w = remap_pixel(p, src, x + 3, y);
        }
    }
#pragma unroll
    for (int r = 0; r < kRowsPerThread; r++) {
        const int y = y_base + r;
        if (y >= int(p.fh)) break;
        st_stream_v4(reinterpret_cast<uint4 *>(dst + size_t(y) * p.fw + x), v[r]);
    }
}

// any canvas width: one pixel per thread
__global__ void __launch_bounds__(kThreads) remap_pixels_kernel(const RemapParams p) {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(blockIdx.z) * p.src_img_bytes);
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.dst + size_t(blockIdx.z) * p.dst_img_bytes);
    const int x = blockIdx.x * kThreads + threadIdx.x;
    if (x >= int(p.fw)) return;
    const int y_base = blockIdx.y * kRowsPerThread;
    uint32_t v[kRowsPerThread];
#pragma unroll
    for (int r = 0; r < kRowsPerThread; r++)
        if (y_base + r < int(p.fh)) v[r] = remap_pixel(p, src, x, y_base + r);
#pragma unroll
    for (int r = 0; r < kRowsPerThread; r++)
        if (y_base + r < int(p.fh)) dst[size_t(y_base + r) * p.fw + x] = v[r];
}

}  // namespace

cudaError_t launch_remap(const RemapParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.fw == 0 || p.fh == 0) return cudaSuccess;
    const bool quads = (p.fw % 4 == 0) && (p.dst_img_bytes % 16 == 0) &&
                       ((reinterpret_cast<uintptr_t>(p.dst) & 15) == 0);
    // gridDim.y/z are limited to 65535: taller canvases go in slices of 65535 * kRowsPerThread rows (the slice is a
    // canvas of its own whose rectangles are shifted up by its first row), batches in slices of 65535 images
    constexpr uint32_t kSliceRows = 65535u * kRowsPerThread;
    for (uint32_t y0 = 0; y0 < p.fh; y0 += kSliceRows) {
        const uint32_t slice_h = p.fh - y0 < kSliceRows ? p.fh - y0 : kSliceRows;
        const unsigned rows = (slice_h + kRowsPerThread - 1) / kRowsPerThread;
        for (uint32_t first = 0; first < n_images; first += 65535) {
            RemapParams q = p;
            q.src += size_t(first) * p.src_img_bytes;
            q.dst += size_t(first) * p.dst_img_bytes + size_t(y0) * p.fw * 4;
            q.fh = slice_h;
            q.ay0 = p.ay0 - int32_t(y0); q.ay1 = p.ay1 - int32_t(y0);
            q.by0 = p.by0 - int32_t(y0); q.by1 = p.by1 - int32_t(y0);
            q.dy = p.dy + int32_t(y0);
            const unsigned nz = n_images - first < 65535 ? n_images - first : 65535;
            if (quads) {
                dim3 grid((p.fw / 4 + kThreads - 1) / kThreads, rows, nz);
                remap_quads_kernel<<<grid, kThreads, 0, stream>>>(q);
            } else {
                dim3 grid((p.fw + kThreads - 1) / kThreads, rows, nz);
                remap_pixels_kernel<<<grid, kThreads, 0, stream>>>(q);
            }
            count_launch();
            cudaError_t e = cudaGetLastError();
            if (e != cudaSuccess) return e;
        }
    }
    return cudaSuccess;
}

}  // namespace slimg
