// k4_planes.cu -- the colour-plane pre-stages of the reference's other lossy encoders, so that a WebP or
// AVIF convert can keep the resized image in HBM and hand the encoder the planes it would have computed itself:
//   WebP  codec/webp.rs:25-30  webp::Encoder::from_rgba -> libwebp WebPPictureImportRGBA: RGBA -> Y (every pixel) and
//         U / V (one per 2 x 2, averaged in the "gamma 0.80" domain of picture_csp_enc.c, alpha-weighted where the block
//         is partly transparent), plus the alpha plane when any pixel is not opaque
//   AVIF  codec/avif.rs:28-47  ravif::Encoder::encode_rgba: RGBA -> YCbCr 4:4:4, BT.601 full range, 8 or 10 bits, f32
//         arithmetic with one fused multiply-add per chroma sample, plus the alpha plane
// Both libraries are crates.io dependencies absent from the reference tree: the arithmetic is restated from their published
// sources (oracle/slimg_oracle.c holds the CPU restatement the tests compare with) -- PARITY UNPINNED, said so in DESIGN.md.
// Roofline: HBM (4 B/px read; 1.5 - 2.5 B/px written for WebP, 4 - 8 B/px for AVIF); pure streaming kernels.
#include <mutex>

#include "kernels.cuh"

namespace slimg {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ int w_l2g(const int *l2g, uint32_t base_value, int shift) {      // LinearToGamma
    const uint32_t v = base_value << shift;
    const int tab_pos = int(v >> 9), x = int(v & 511u);
    const int y = l2g[tab_pos + 1] * x + l2g[tab_pos] * (512 - x);
    return (y + 64) >> 7;
}
// libwebp's kInvAlpha: (1 << 19) / a for the alpha sum of a 2 x 2 block, a table instead of three integer divisions per block
struct InvAlphaTab {
    uint32_t v[1024];
    constexpr InvAlphaTab() : v() { for (uint32_t a = 1; a < 1024; a++) v[a] = (1u << 19) / a; }
};
__device__ const InvAlphaTab kInvAlpha{};

__device__ __forceinline__ int w_clip_uv(int uv) {
    uv = (uv + (1 << 17) + (128 << 18)) >> 18;
    return min(max(uv, 0), 255);
}

// one thread = 4 x 2 pixels = two chroma samples (two 16-byte loads, word stores of Y / A, halfword stores of U / V);
// grid.z = image
__global__ void __launch_bounds__(kThreads) webp_planes_kernel(const WebpPlanesParams p) {
    __shared__ uint16_t g2l[256];
    __shared__ int l2g[33];
    for (int i = threadIdx.x; i < 256; i += kThreads) g2l[i] = p.g2l[i];
    if (threadIdx.x < 33) l2g[threadIdx.x] = p.l2g[threadIdx.x];
    __syncthreads();
    const int qx = blockIdx.x * 64 + (threadIdx.x & 63), cy = blockIdx.y * 4 + (threadIdx.x >> 6);       // quad column, chroma row
    const int w = int(p.w), h = int(p.h), uvw = (w + 1) >> 1, uvh = (h + 1) >> 1;
    const int x0 = 4 * qx, y0 = 2 * cy;
    if (x0 >= w || cy >= uvh) return;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(blockIdx.z) * p.src_img_bytes);
    const bool tall = y0 + 1 < h;
    const int y1 = tall ? y0 + 1 : y0;                     // odd height: the last row twice
    const bool vec = (w & 3) == 0 && (p.src_img_bytes & 15) == 0 && (reinterpret_cast<uintptr_t>(p.src) & 15) == 0;
    uint32_t px[2][4];
    if (vec) {
        const uint4 a = __ldg(reinterpret_cast<const uint4 *>(src + size_t(y0) * w + x0));
        const uint4 b = __ldg(reinterpret_cast<const uint4 *>(src + size_t(y1) * w + x0));
        px[0][0] = a.x; px[0][1] = a.y; px[0][2] = a.z; px[0][3] = a.w;
        px[1][0] = b.x; px[1][1] = b.y; px[1][2] = b.z; px[1][3] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int x = min(x0 + k, w - 1);
            px[0][k] = __ldg(src + size_t(y0) * w + x); px[1][k] = __ldg(src + size_t(y1) * w + x);
        }
    }
    uint8_t *Y = p.y + size_t(blockIdx.z) * p.y_img_bytes;
    uint8_t *A = p.a ? p.a + size_t(blockIdx.z) * p.y_img_bytes : nullptr;
#pragma unroll
    for (int r = 0; r < 2; r++) {
        if (r == 1 && !tall) break;
        uint32_t yw = 0u, aw = 0u;
#pragma unroll
        for (int k = 0; k < 4; k++) {                     // VP8RGBToY
            const uint32_t v = px[r][k];
            const int yv = (16839 * int(v & 0xff) + 33059 * int((v >> 8) & 0xff) + 6420 * int((v >> 16) & 0xff) + (1 << 15) + (16 << 16)) >> 16;
            yw |= uint32_t(yv) << (8 * k);
            aw |= (v >> 24) << (8 * k);
        }
        const size_t o = size_t(r ? y1 : y0) * w + x0;
        if (vec) {
            *reinterpret_cast<uint32_t *>(Y + o) = yw;
            if (A) *reinterpret_cast<uint32_t *>(A + o) = aw;
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (x0 + k < w) { Y[o + k] = uint8_t(yw >> (8 * k)); if (A) A[o + k] = uint8_t(aw >> (8 * k)); }
        }
    }
    // chroma: AccumulateRGB / AccumulateRGBA.  libwebp takes the alpha-aware form in row pairs that contain a pixel that
    // is not opaque -- and that form falls back to the plain sum for every 2 x 2 block whose alphas are all 255 (or all
    // 0), so the per-row decision never changes a sample: the block's own alpha sum decides.  A caller that ran the
    // scan kernel (row_alpha != NULL: the variant with static scratch) gets the same samples from its flags.
    const bool rows_alpha = p.row_alpha ? p.row_alpha[size_t(blockIdx.z) * uvh + cy] != 0 : p.any_alpha != nullptr;
    if (!p.row_alpha && p.any_alpha) {                     // no scan pass: this thread reports what it saw
        bool seen = false;
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int k = 0; k < 4; k++) seen |= (px[r][k] >> 24) != 0xffu;
        // one store per warp at most, and none once the flag is up: millions of stores to one word serialise in L2
        const unsigned active = __activemask();
        if (__any_sync(active, seen) && (threadIdx.x & 31) == __ffs(active) - 1) {
            volatile uint32_t *flag = p.any_alpha + blockIdx.z;
            if (*flag == 0u) *flag = 1u;
        }
    }
    uint32_t uvp[2] = {0u, 0u};                            // U and V of the two samples, packed
#pragma unroll
    for (int e = 0; e < 2; e++) {
        const int xa = x0 + 2 * e;
        if (xa >= w) break;
        const bool wide = xa + 1 < w;
        const uint32_t q0 = px[0][2 * e], q1 = px[0][wide ? 2 * e + 1 : 2 * e], q2 = px[1][2 * e], q3 = px[1][wide ? 2 * e + 1 : 2 * e];
        const uint32_t asum = wide ? (q0 >> 24) + (q1 >> 24) + (q2 >> 24) + (q3 >> 24) : 2u * ((q0 >> 24) + (q2 >> 24));
        int rgb[3];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const uint32_t v0 = g2l[(q0 >> (8 * c)) & 0xff], v1 = g2l[(q1 >> (8 * c)) & 0xff];
            const uint32_t v2 = g2l[(q2 >> (8 * c)) & 0xff], v3 = g2l[(q3 >> (8 * c)) & 0xff];
            if (!rows_alpha || asum == 4u * 0xffu || asum == 0u) rgb[c] = wide ? w_l2g(l2g, v0 + v1 + v2 + v3, 0) : w_l2g(l2g, v0 + v2, 1);
            else {
                const uint32_t sum = (q0 >> 24) * v0 + (q1 >> 24) * v1 + (q2 >> 24) * v2 + (q3 >> 24) * v3;
                const uint32_t inv = __ldg(&kInvAlpha.v[asum]);       // 0 < asum < 1020 here
                rgb[c] = w_l2g(l2g, uint32_t((uint64_t(sum) * inv) >> 17), 0);
            }
        }
        uvp[0] |= uint32_t(w_clip_uv(-9719 * rgb[0] - 19081 * rgb[1] + 28800 * rgb[2])) << (8 * e);
        uvp[1] |= uint32_t(w_clip_uv(28800 * rgb[0] - 24116 * rgb[1] - 4684 * rgb[2])) << (8 * e);
    }
    const size_t uvo = size_t(blockIdx.z) * p.uv_img_bytes + size_t(cy) * uvw + 2 * qx;
    if (vec) {                                             // w % 4 == 0: uvw is even, 2 qx + 1 < uvw
        *reinterpret_cast<uint16_t *>(p.u + uvo) = uint16_t(uvp[0]);
        *reinterpret_cast<uint16_t *>(p.v + uvo) = uint16_t(uvp[1]);
    } else {
        p.u[uvo] = uint8_t(uvp[0]); p.v[uvo] = uint8_t(uvp[1]);
        if (2 * qx + 1 < uvw) { p.u[uvo + 1] = uint8_t(uvp[0] >> 8); p.v[uvo + 1] = uint8_t(uvp[1] >> 8); }
    }
}

// CheckNonOpaque per pair of rows: row_alpha[img][cy] = some alpha of rows 2cy, 2cy+1 is not 255; any[img] = OR of them
__global__ void __launch_bounds__(kThreads) webp_alpha_scan_kernel(const WebpPlanesParams p) {
    const int cy = blockIdx.x, w = int(p.w), h = int(p.h);
    const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(blockIdx.z) * p.src_img_bytes);
    const int y0 = 2 * cy, y1 = min(2 * cy + 1, h - 1);
    int found = 0;
    for (int x = threadIdx.x; x < w; x += kThreads)
        found |= ((__ldg(src + size_t(y0) * w + x) >> 24) != 0xffu) | ((__ldg(src + size_t(y1) * w + x) >> 24) != 0xffu);
    found = __syncthreads_or(found);
    if (threadIdx.x == 0) {
        p.row_alpha[size_t(blockIdx.z) * ((h + 1) >> 1) + cy] = uint8_t(found != 0);
        if (found) p.any_alpha[blockIdx.z] = 1u;
    }
}

__device__ __forceinline__ float round_away(float x) { return x < 0.0f ? -floorf(0.5f - x) : floorf(x + 0.5f); }

// The conversion pipe (I2F, FRND, F2I: 16 lanes / clk / SM) would bound this kernel -- twelve such instructions per pixel
// are 64 us per 24 MP image -- so bytes become floats and floats become samples on the FP32 and ALU pipes:
//   byte -> f32: the byte dropped into the mantissa of 2^23, minus 2^23 (exact)
//   f32 -> sample: clamp to [0, max] first, then floor(t + 0.5) as the mantissa of (t + 0.5) + 2^23 added toward zero.
//   Same value as clamp(round_away(x)): below 0 both give 0, above max both give max, in between round_away(x) =
//   floor(x + 0.5) with the same rounded sum x + 0.5.
__device__ __forceinline__ float byte_to_float(uint32_t v, int byte) {
    return __int_as_float(int(__byte_perm(v, 0x4B000000u, 0x7650 + byte))) - 8388608.0f;      // bytes: [v.byte, 0, 0, 0x4B]
}
__device__ __forceinline__ uint32_t round_clamp(float x, float max_value) {
    const float t = fminf(fmaxf(x, 0.0f), max_value) + 0.5f;
    return __float_as_uint(__fadd_rz(t, 8388608.0f)) & 0x7ffu;
}

// one thread = 4 pixels (one 16-byte load); DEPTH 8: u8 planes, 10: u16 planes
template <int DEPTH>
__global__ void __launch_bounds__(kThreads) avif_planes_kernel(const AvifPlanesParams p) {
    using T = typename std::conditional<DEPTH == 8, uint8_t, uint16_t>::type;
    const size_t n = size_t(p.w) * p.h, i0 = (size_t(blockIdx.x) * kThreads + threadIdx.x) * 4;
    if (i0 >= n) return;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(p.src + size_t(blockIdx.z) * p.src_img_bytes);
    uint32_t px[4];
    const int cnt = n - i0 < 4 ? int(n - i0) : 4;
    if (cnt == 4 && (p.src_img_bytes & 15) == 0 && (reinterpret_cast<uintptr_t>(p.src) & 15) == 0) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(src + i0));
        px[0] = v.x; px[1] = v.y; px[2] = v.z; px[3] = v.w;
    } else
        for (int k = 0; k < cnt; k++) px[k] = __ldg(src + i0 + k);
    constexpr float m0 = 0.2990f, m1 = 0.5870f, m2 = 0.1140f;
    const float max_value = float((1 << DEPTH) - 1);
    const float scale = __fdiv_rn(max_value, 255.f);
    const float shift = round_away(__fmul_rn(max_value, 0.5f));
    const float kb = __fdiv_rn(0.5f, __fsub_rn(1.f, m2)), kr = __fdiv_rn(0.5f, __fsub_rn(1.f, m0));
    const float t0 = __fmul_rn(scale, m0), t1 = __fmul_rn(scale, m1), t2 = __fmul_rn(scale, m2);
    T *Y = reinterpret_cast<T *>(p.y) + size_t(blockIdx.z) * n, *Cb = reinterpret_cast<T *>(p.cb) + size_t(blockIdx.z) * n;
    T *Cr = reinterpret_cast<T *>(p.cr) + size_t(blockIdx.z) * n, *A = reinterpret_cast<T *>(p.a) + size_t(blockIdx.z) * n;
    T vy[4], vb[4], vr[4], va[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t v = k < cnt ? px[k] : 0u;
        const float r = byte_to_float(v, 0), g = byte_to_float(v, 1), b = byte_to_float(v, 2);
        const uint32_t al = v >> 24;
        // every product and sum rounds to f32 (the reference build does not contract); the chroma terms are ONE fused multiply-add
        const float yv = __fadd_rn(__fadd_rn(__fmul_rn(t0, r), __fmul_rn(t1, g)), __fmul_rn(t2, b));
        const float cbv = __fmaf_rn(__fsub_rn(__fmul_rn(b, scale), yv), kb, shift);
        const float crv = __fmaf_rn(__fsub_rn(__fmul_rn(r, scale), yv), kr, shift);
        vy[k] = T(round_clamp(yv, max_value));
        vb[k] = T(round_clamp(cbv, max_value));
        vr[k] = T(round_clamp(crv, max_value));
        va[k] = DEPTH == 8 ? T(al) : T((al << 2) | (al >> 6));
    }
    const bool vec = cnt == 4 && (n & 3) == 0;                 // plane rows of 4 samples are then 4- / 8-byte aligned (cudaMalloc'ed planes)
    auto put = [&](T *dst, const T (&v)[4]) {
        if (vec) {
            if (DEPTH == 8) *reinterpret_cast<uint32_t *>(dst + i0) = uint32_t(v[0]) | (uint32_t(v[1]) << 8) | (uint32_t(v[2]) << 16) | (uint32_t(v[3]) << 24);
            else *reinterpret_cast<uint2 *>(dst + i0) = make_uint2(uint32_t(v[0]) | (uint32_t(v[1]) << 16), uint32_t(v[2]) | (uint32_t(v[3]) << 16));
        } else
            for (int k = 0; k < cnt; k++) dst[i0 + k] = v[k];
    };
    put(Y, vy); put(Cb, vb); put(Cr, vr); put(A, va);
}

}  // namespace

cudaError_t launch_webp_planes(const WebpPlanesParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.w == 0 || p.h == 0) return cudaSuccess;
    if (n_images > 65535) return cudaErrorInvalidValue;
    const unsigned uvw = (p.w + 1) / 2, uvh = (p.h + 1) / 2;
    if (p.row_alpha) {
        webp_alpha_scan_kernel<<<dim3(uvh, 1, n_images), kThreads, 0, stream>>>(p);
        count_launch();
    }
    webp_planes_kernel<<<dim3(((p.w + 3) / 4 + 63) / 64, (uvh + 3) / 4, n_images), kThreads, 0, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}

// ---- the variant that allocates nothing (slimg_b200_webp_planes_batch_device_static) ---------------------------------
// The scratch of the WebP planes conversion -- one byte per pair of rows ("this pair holds a pixel that is not opaque")
// and one u32 flag per image -- lives in this file-scope array instead of memory the caller allocated.  Every GPU
// that loads the library holds these kWebpStaticScratchBytes until the process exits; every call on a device shares
// them, so calls that overlap on the GPU race (see include/slimg_b200.h).
__device__ __align__(256) uint8_t g_webp_static_scratch[kWebpStaticScratchBytes];

// In host code `g_webp_static_scratch` is a host-side stand-in, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
static cudaError_t webp_static_scratch(uint8_t **out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static uint8_t *addr[kMaxDevices] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_webp_static_scratch);
        if (e != cudaSuccess) return e;
        if (!ptr) return cudaErrorInvalidSymbol;
        addr[dev] = static_cast<uint8_t *>(ptr);
    }
    *out = addr[dev];
    return cudaSuccess;
}

size_t webp_static_scratch_need(uint32_t n_images, uint32_t h) {
    const size_t rows = size_t(n_images) * ((size_t(h) + 1) / 2);
    return ((rows + 3) & ~size_t(3)) + size_t(n_images) * 4;
}

// same two kernels as launch_webp_planes; p.row_alpha / p.any_alpha are set here.  any_out (device, n u32, may be NULL)
// receives the per-image "has a pixel that is not opaque" flags.
cudaError_t launch_webp_planes_static(WebpPlanesParams p, uint32_t n_images, uint32_t *any_out, cudaStream_t stream) {
    if (n_images == 0 || p.w == 0 || p.h == 0) return cudaSuccess;
    if (webp_static_scratch_need(n_images, p.h) > kWebpStaticScratchBytes) return cudaErrorInvalidValue;
    uint8_t *scratch = nullptr;
    cudaError_t e = webp_static_scratch(&scratch);
    if (e != cudaSuccess) return e;
    const size_t rows = size_t(n_images) * ((size_t(p.h) + 1) / 2);
    p.row_alpha = scratch;                                   // fully rewritten by the scan kernel
    p.any_alpha = reinterpret_cast<uint32_t *>(scratch + ((rows + 3) & ~size_t(3)));
    // the flags are only ever SET by the scan kernel: the previous call's must go
    if ((e = cudaMemsetAsync(p.any_alpha, 0, size_t(n_images) * 4, stream)) != cudaSuccess) return e;
    if ((e = launch_webp_planes(p, n_images, stream)) != cudaSuccess) return e;
    if (any_out) e = cudaMemcpyAsync(any_out, p.any_alpha, size_t(n_images) * 4, cudaMemcpyDeviceToDevice, stream);
    return e;
}

cudaError_t launch_avif_planes(const AvifPlanesParams &p, uint32_t n_images, cudaStream_t stream) {
    if (n_images == 0 || p.w == 0 || p.h == 0) return cudaSuccess;
    if (n_images > 65535 || (p.depth != 8 && p.depth != 10)) return cudaErrorInvalidValue;
    const size_t n = size_t(p.w) * p.h;
    const unsigned blocks = unsigned((n + size_t(kThreads) * 4 - 1) / (size_t(kThreads) * 4));
    if (p.depth == 8) avif_planes_kernel<8><<<dim3(blocks, 1, n_images), kThreads, 0, stream>>>(p);
    else avif_planes_kernel<10><<<dim3(blocks, 1, n_images), kThreads, 0, stream>>>(p);
    count_launch();
    return cudaGetLastError();
}

}  // namespace slimg
