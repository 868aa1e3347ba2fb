// microbench.cu -- B200 instruction-throughput probes that decide how K2 (resize) spends its
// issue slots: legacy mma.sync int8 / fp16 rates, FFMA vs packed FFMA2, PRMT and I2F rates, and
// whether HMMA honours fp16 subnormal inputs (byte -> fp16 without arithmetic).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;

__global__ void k_imma(int *out, int n) {
    int acc[8][4] = {};
    uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 ^ 0x5a5a, b1 = a0 ^ 0xa5a5;
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(acc[j][0]), "+r"(acc[j][1]), "+r"(acc[j][2]), "+r"(acc[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    int s = 0;
    for (int j = 0; j < 8; j++) s += acc[j][0] + acc[j][1] + acc[j][2] + acc[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_hmma(float *out, int n) {
    float acc[8][4] = {};
    uint32_t a0 = 0x3c003c00, a1 = a0, a2 = a0, a3 = a0, b0 = 0x38003800, b1 = b0;
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(acc[j][0]), "+f"(acc[j][1]), "+f"(acc[j][2]), "+f"(acc[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    float s = 0;
    for (int j = 0; j < 8; j++) s += acc[j][0] + acc[j][1] + acc[j][2] + acc[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma(float *out, int n, float w) {
    float acc[16];
    for (int j = 0; j < 16; j++) acc[j] = threadIdx.x + j;
    float x = w * 0.5f;
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) acc[j] = fmaf(acc[j], w, x);
    }
    float s = 0;
    for (int j = 0; j < 16; j++) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma2(float *out, int n, float w) {
    unsigned long long acc[8];
    for (int j = 0; j < 8; j++) acc[j] = (unsigned long long)__float_as_uint(threadIdx.x + j) * 0x100000001ull;
    unsigned long long ww = (unsigned long long)__float_as_uint(w) * 0x100000001ull;
    unsigned long long xx = (unsigned long long)__float_as_uint(w * 0.5f) * 0x100000001ull;
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[j]) : "l"(ww), "l"(xx));
    }
    unsigned long long s = 0;
    for (int j = 0; j < 8; j++) s ^= acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = __uint_as_float((unsigned)(s ^ (s >> 32)));
}

__global__ void k_prmt(uint32_t *out, int n, uint32_t sel) {
    uint32_t acc[16];
    for (int j = 0; j < 16; j++) acc[j] = threadIdx.x * 2654435761u + j;
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) acc[j] = __byte_perm(acc[j], acc[(j + 1) & 15], sel);
    }
    uint32_t s = 0;
    for (int j = 0; j < 16; j++) s ^= acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// integer-pipe rates that decide K3's instruction choices: 16 independent chains per thread, one op per chain per step
#define INT_KERNEL(NAME, EXPR)                                                                  \
    __global__ void NAME(uint32_t *out, int n, uint32_t m) {                                    \
        uint32_t acc[16];                                                                       \
        for (int j = 0; j < 16; j++) acc[j] = threadIdx.x * 2654435761u + j;                    \
        for (int i = 0; i < n; i++) {                                                           \
            _Pragma("unroll") for (int j = 0; j < 16; j++) { const uint32_t x = acc[j], y = acc[(j + 1) & 15]; acc[j] = (EXPR); } \
        }                                                                                       \
        uint32_t s = 0;                                                                         \
        for (int j = 0; j < 16; j++) s ^= acc[j];                                               \
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;                                         \
    }
INT_KERNEL(k_imad, x * m + y)
INT_KERNEL(k_imadhi, __umulhi(x, m) + y)
INT_KERNEL(k_shf, __funnelshift_r(x, y, m))
INT_KERNEL(k_shr, (uint32_t(int(x) >> (m & 31)) ^ y))
INT_KERNEL(k_lop3, (x & m) ^ y)
INT_KERNEL(k_iadd3, x + y + m)
INT_KERNEL(k_dp4a, __dp4a(x, m, y))
INT_KERNEL(k_f2isat, ([&] { uint32_t r; asm("cvt.rzi.sat.u8.f32 %0, %1;" : "=r"(r) : "f"(__uint_as_float((x & 0x007fffffu) | 0x42000000u))); return r + y; }()))

__global__ void k_i2f(float *out, int n) {
    uint32_t v[16];
    float acc[16];
    for (int j = 0; j < 16; j++) { v[j] = threadIdx.x + j; acc[j] = 0; }
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) { acc[j] = float(v[j] & 0xffu); v[j] = __float_as_uint(acc[j]) >> 3; }
    }
    float s = 0;
    for (int j = 0; j < 16; j++) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// fp16 subnormal through HMMA: A = byte value b stored as the fp16 bit pattern 0x00bb (= b * 2^-24),
// B = 2^10 in column 0.  Expect D = sum_k b * 2^-14.
__global__ void k_denorm(float *out) {
    float acc[4] = {0, 0, 0, 0};
    uint32_t a0 = 0x00ff0003, a1 = 0, a2 = 0, a3 = 0;   // halves: 3*2^-24 and 255*2^-24
    uint32_t b0 = 0x64006400, b1 = 0;                   // 1024.0, 1024.0
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(acc[0]), "+f"(acc[1]), "+f"(acc[2]), "+f"(acc[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    if (threadIdx.x == 0) { out[0] = acc[0]; out[1] = acc[1]; }
}

template <class F>
static float time_ms(F f) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f();
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount, blocks = sms * 8, threads = 256;
    printf("device %s, %d SMs, clock %d kHz\n", prop.name, sms, prop.clockRate);
    void *buf;
    CK(cudaMalloc(&buf, size_t(blocks) * threads * 8));
    const double warps = double(blocks) * threads / 32;
    float ms;
    ms = time_ms([&] { k_imma<<<blocks, threads>>>((int *)buf, ITERS); });
    printf("imma m16n8k32 u8.s8 : %8.3f ms  %8.1f TMAC/s  (%.0f MAC/clk/SM @1.9GHz)\n", ms, warps * ITERS * 8 * 4096 / ms / 1e9, warps * ITERS * 8 * 4096 / (ms * 1e-3) / sms / 1.9e9);
    ms = time_ms([&] { k_hmma<<<blocks, threads>>>((float *)buf, ITERS); });
    printf("hmma m16n8k16 f16   : %8.3f ms  %8.1f TMAC/s  (%.0f MAC/clk/SM @1.9GHz)\n", ms, warps * ITERS * 8 * 2048 / ms / 1e9, warps * ITERS * 8 * 2048 / (ms * 1e-3) / sms / 1.9e9);
    ms = time_ms([&] { k_ffma<<<blocks, threads>>>((float *)buf, ITERS, 1.0001f); });
    printf("ffma                : %8.3f ms  %8.2f TFMA/s  (%.1f FMA/clk/SM @1.9GHz)\n", ms, warps * 32 * ITERS * 16 / ms / 1e9, warps * 32 * ITERS * 16 / (ms * 1e-3) / sms / 1.9e9);
    ms = time_ms([&] { k_ffma2<<<blocks, threads>>>((float *)buf, ITERS, 1.0001f); });
    printf("ffma2 (f32x2)       : %8.3f ms  %8.2f TFMA/s  (%.1f FMA/clk/SM @1.9GHz)\n", ms, warps * 32 * ITERS * 16 / ms / 1e9, warps * 32 * ITERS * 16 / (ms * 1e-3) / sms / 1.9e9);
    ms = time_ms([&] { k_prmt<<<blocks, threads>>>((uint32_t *)buf, ITERS, 0x5140); });
    printf("prmt                : %8.3f ms  %8.2f Tops/s  (%.1f op/clk/SM @1.9GHz)\n", ms, warps * 32 * ITERS * 16 / ms / 1e9, warps * 32 * ITERS * 16 / (ms * 1e-3) / sms / 1.9e9);
    ms = time_ms([&] { k_i2f<<<blocks, threads>>>((float *)buf, ITERS); });
    printf("and+i2f+shr chain   : %8.3f ms  %8.2f Tcvt/s  (%.1f cvt/clk/SM @1.9GHz)\n", ms, warps * 32 * ITERS * 16 / ms / 1e9, warps * 32 * ITERS * 16 / (ms * 1e-3) / sms / 1.9e9);
#define RUN_INT(NAME, LABEL, ARG)                                                                                \
    ms = time_ms([&] { NAME<<<blocks, threads>>>((uint32_t *)buf, ITERS, ARG); });                               \
    printf("%-20s: %8.3f ms  %8.2f Tops/s  (%.1f op/clk/SM @1.9GHz)\n", LABEL, ms, warps * 32 * ITERS * 16 / ms / 1e9, warps * 32 * ITERS * 16 / (ms * 1e-3) / sms / 1.9e9);
    RUN_INT(k_imad, "imad (lo)", 2654435761u)
    RUN_INT(k_imadhi, "imad.hi + add", 1u << 21)
    RUN_INT(k_shf, "shf (funnel)", 11u)
    RUN_INT(k_shr, "shr.s32 + xor", 11u)
    RUN_INT(k_lop3, "lop3 (and-xor)", 0x00ff00ffu)
    RUN_INT(k_iadd3, "iadd3", 12345u)
    RUN_INT(k_dp4a, "dp4a u8*s8", 0x4c8b1d00u)
    RUN_INT(k_f2isat, "lop+f2i.sat.u8+add", 0u)
    k_denorm<<<1, 32>>>((float *)buf);
    float h[2];
    CK(cudaMemcpy(h, buf, 8, cudaMemcpyDeviceToHost));
    printf("hmma fp16 subnormal : got %.10g, want %.10g (%s)\n", h[0], 4 * (3.0 + 255.0) * 1024.0 / 16777216.0,
           h[0] == float(4 * (3.0 + 255.0) * 1024.0 / 16777216.0) ? "EXACT" : "FLUSHED/INEXACT");
    CK(cudaDeviceSynchronize());
    return 0;
}
