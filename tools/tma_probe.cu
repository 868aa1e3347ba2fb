// tma_probe.cu -- how fast can one CTA per SM stream strips of image rows into shared memory?
// Decides the ring-fill method of k2_vfirst.cu.  One CTA per SM, one producer lane, a ring of
// kSlots boxes, consumer = the same thread waiting on the mbarrier (no compute): measures the
// fetch path alone.
//   mode 0: 2-D tiled TMA, box = W bytes x 16 rows, no swizzle, 16-byte aligned start
//   mode 1: same, 128-byte aligned start and W rounded up to a multiple of 128
//   mode 2: 3-D tiled TMA, box = 128 B x 16 rows x P panels, SWIZZLE_128B (UMMA MN-major layout)
//   mode 3: linear cp.async.bulk, one copy per row (W bytes)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_probe tools/tma_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

constexpr int kRows = 16;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t *b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t par) {
    uint32_t ok = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(par) : "memory");
    } while (!ok);
}

__global__ void __launch_bounds__(128, 1) probe(const __grid_constant__ CUtensorMap map, const uint8_t *src, int mode, int W, int P,
                                                int slots, int row_pitch, int n_rows, int col_stride, int n_strips, int steps_per_cta, int hot) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t full[16];
    const int box_bytes = mode == 2 ? 128 * kRows * P : W * kRows;
    if (threadIdx.x == 0) {
        for (int s = 0; s < slots; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    // this CTA walks down one strip of one "image" (a tall tensor), like a k2_vfirst CTA
    const int strip = blockIdx.x % n_strips, img = blockIdx.x / n_strips;
    const int c0 = strip * col_stride;                       // bytes
    const int r0 = img * steps_per_cta * kRows;
    auto issue = [&](int step, int slot) {
        const int row = hot ? r0 + (step & 3) * kRows : r0 + step * kRows;
        mbar_expect(&full[slot], uint32_t(box_bytes));
        const uint32_t dst = smem_u32(smem) + uint32_t(slot) * uint32_t((box_bytes + 1023) & ~1023);
        if (mode == 0 || mode == 1)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(dst), "l"(&map), "r"(c0 / 8), "r"(row), "r"(smem_u32(&full[slot])) : "memory");
        else if (mode == 2)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"(&map), "r"(0), "r"(row), "r"(c0 / 128), "r"(smem_u32(&full[slot])) : "memory");
        else
            for (int r = 0; r < kRows; r++)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(dst + r * W), "l"(src + size_t(row + r) * row_pitch + c0), "r"(W), "r"(smem_u32(&full[slot])) : "memory");
    };
    int issued = 0;
    for (; issued < slots && issued < steps_per_cta; issued++) issue(issued, issued);
    int slot = 0;
    uint32_t par = 0;
    for (int s = 0; s < steps_per_cta; s++) {
        mbar_wait(&full[slot], par);
        if (issued < steps_per_cta) { issue(issued, slot); issued++; }
        if (++slot == slots) { slot = 0; par ^= 1; }
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char **argv) {
    const int row_pitch = 24000, W0 = argc > 1 ? atoi(argv[1]) : 880;
    const int steps = 214;                                   // 3424 rows per CTA, like one 24 MP strip
    const int n_strips = 27, n_img = 148 * 4 / n_strips + 1; // strips 880 B apart cover the 24000-byte rows
    const size_t rows_total = size_t(n_img) * steps * kRows;
    uint8_t *d;
    cudaMalloc(&d, rows_total * row_pitch + 4096);
    cudaMemset(d, 1, rows_total * row_pitch + 4096);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    EncodeTiledFn enc = (EncodeTiledFn)fn;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int hot = 0; hot < 2; hot++)
        for (int mode = 0; mode < 4; mode++) {
            int W = W0, P = 0, col_stride = 880;
            CUtensorMap map;
            if (mode == 1) { W = (W0 + 127) / 128 * 128; col_stride = 896; }
            if (mode == 2) { P = (W0 + 127) / 128; col_stride = 896; }
            CUresult r = CUDA_SUCCESS;
            if (mode <= 1 || mode == 3) {
                const cuuint64_t dims[2] = {cuuint64_t(row_pitch / 8), rows_total};
                const cuuint64_t strides[1] = {cuuint64_t(row_pitch)};
                const cuuint32_t box[2] = {cuuint32_t(W / 8), kRows}, es[2] = {1, 1};
                r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            } else {
                // (128 B inner, rows, panels): panels are 128 B apart along the row
                const cuuint64_t dims[3] = {16, rows_total, cuuint64_t(row_pitch / 128)};
                const cuuint64_t strides[2] = {cuuint64_t(row_pitch), 128};
                const cuuint32_t box[3] = {16, kRows, cuuint32_t(P)}, es[3] = {1, 1, 1};
                r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            }
            if (r != CUDA_SUCCESS) { printf("mode %d: encode failed %d\n", mode, int(r)); continue; }
            const int box_bytes = mode == 2 ? 128 * kRows * P : W * kRows;
            const int slot_bytes = (box_bytes + 1023) & ~1023;
            for (int slots = 4; slots <= 12; slots += 4) {
                if (slots * slot_bytes > 200 * 1024) continue;
                const int grid = 148 * 4;
                for (int rep = 0; rep < 2; rep++) {
                    cudaEventRecord(e0);
                    probe<<<grid, 128, 200 * 1024>>>(map, d, mode, W, P, slots, row_pitch, int(rows_total), col_stride, n_strips, steps, hot);
                    cudaEventRecord(e1);
                    cudaError_t err = cudaDeviceSynchronize();
                    if (err != cudaSuccess) { printf("mode %d: %s\n", mode, cudaGetErrorString(err)); return 1; }
                }
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                const double bytes = double(grid) * steps * box_bytes;
                printf("%s mode %d W=%4d slots=%2d in-flight %6.1f KB: %7.3f ms  %7.1f GB/s  %5.1f B/clk/SM\n", hot ? "L2-hot" : "stream",
                       mode, mode == 2 ? 128 * P : W, slots, slots * box_bytes / 1024.0, ms, bytes / ms / 1e6, bytes / ms / 1e6 / 148 / 1.965);
            }
        }
    return 0;
}
