// pf_probe.cu -- what does it cost a warp to push lines towards L2 without the TMA engine?
//   mode 0: prefetch.global.L2 [addr]            (one 128-byte line per lane)
//   mode 1: ld.global.L2::256B.u32 into registers nobody reads (8 independent destinations per lane)
//   mode 2: cp.async.bulk.prefetch.L2.global [addr], 656      (one row per lane)
// prints cycles per warp-wide instruction seen by the issuing warp (148 CTAs x 1 warp, cold addresses)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pf_probe tools/pf_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__global__ void probe(const uint8_t *src, int mode, int n, long long *out, size_t stride) {
    const int lane = threadIdx.x;
    const uint8_t *base = src + size_t(blockIdx.x) * stride;
    __syncwarp();
    const long long t0 = clock64();
    if (mode == 0) {
        for (int i = 0; i < n; i++) asm volatile("prefetch.global.L2 [%0];" ::"l"(base + (size_t(i) * 32 + lane) * 128));
    } else if (mode == 1) {
        for (int i = 0; i < n; i += 8) {
            uint32_t d0, d1, d2, d3, d4, d5, d6, d7;
            const uint8_t *a = base + (size_t(i) * 32 + lane) * 256;
            asm volatile("ld.global.L2::256B.u32 %0, [%8];\n\tld.global.L2::256B.u32 %1, [%8+8192];\n\tld.global.L2::256B.u32 %2, [%8+16384];\n\t"
                         "ld.global.L2::256B.u32 %3, [%8+24576];\n\tld.global.L2::256B.u32 %4, [%8+32768];\n\tld.global.L2::256B.u32 %5, [%8+40960];\n\t"
                         "ld.global.L2::256B.u32 %6, [%8+49152];\n\tld.global.L2::256B.u32 %7, [%8+57344];"
                         : "=r"(d0), "=r"(d1), "=r"(d2), "=r"(d3), "=r"(d4), "=r"(d5), "=r"(d6), "=r"(d7) : "l"(a));
        }
    } else {
        for (int i = 0; i < n; i++)
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + (size_t(i) * 32 + lane) * 1024), "r"(656) : "memory");
    }
    const long long t1 = clock64();
    if (lane == 0) out[blockIdx.x] = t1 - t0;
}
int main() {
    const size_t stride = 64u << 20;
    uint8_t *d; long long *o, h[148];
    cudaMalloc(&d, stride * 148); cudaMalloc(&o, sizeof(h));
    for (int mode = 0; mode < 3; mode++)
        for (int n = 8; n <= 64; n *= 2) {
            cudaMemset(d, 1, 256u << 20);     // push earlier lines out of the way
            probe<<<148, 32>>>(d, mode, n, o, stride);
            cudaError_t e = cudaDeviceSynchronize();
            cudaMemcpy(h, o, sizeof(h), cudaMemcpyDeviceToHost);
            double s = 0; for (int i = 0; i < 148; i++) s += double(h[i]);
            printf("mode %d n=%2d: %s  %.0f clk per warp instruction (issuing warp's view)\n", mode, n, cudaGetErrorString(e), s / 148 / n);
        }
    return 0;
}
