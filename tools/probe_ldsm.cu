// probe_ldsm.cu -- prints the fragment layout of ldmatrix.m16n16.x2.trans.b8 on sm_100a and checks
// that it is the A fragment of mma.m16n8k32.u8.s8 (byte-column = m, smem row = k).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void probe(uint32_t *out) {
    __shared__ __align__(128) uint8_t tile[32 * 16];
    // tile[row k][byte m] = (k << 4) | m   (k < 32 does not fit 4 bits: encode k in 5 bits, m in 3 low bits + print both)
    for (int i = threadIdx.x; i < 32 * 16; i += 32) tile[i] = uint8_t(i);   // value = k*16 + m (mod 256)
    __syncwarp();
    const int lane = threadIdx.x;
    uint32_t addr = uint32_t(__cvta_generic_to_shared(&tile[lane * 16]));   // lane l -> row l
    uint32_t r0, r1, r2, r3;
    asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
    out[lane * 4 + 0] = r0; out[lane * 4 + 1] = r1; out[lane * 4 + 2] = r2; out[lane * 4 + 3] = r3;
}

// end-to-end check: D[m][n] = sum_k tile[k][m] * B[k][n]
__global__ void probe_mma(const uint8_t *src, const int8_t *w, int *d) {
    __shared__ __align__(128) uint8_t tile[32 * 16];
    for (int i = threadIdx.x; i < 512; i += 32) tile[i] = src[i];
    __syncwarp();
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    uint32_t addr = uint32_t(__cvta_generic_to_shared(&tile[lane * 16]));
    uint32_t a[4];
    asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];"
                 : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]) : "r"(addr));
    // B fragment: b0 = B[k=4t..4t+3][n=g], b1 = B[k=16+4t..][n=g];  w is [k][n] (32 x 8)
    uint32_t b0 = 0, b1 = 0;
    for (int i = 0; i < 4; i++) {
        b0 |= uint32_t(uint8_t(w[(4 * t + i) * 8 + g])) << (8 * i);
        b1 |= uint32_t(uint8_t(w[(16 + 4 * t + i) * 8 + g])) << (8 * i);
    }
    int c[4] = {0, 0, 0, 0};
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    d[g * 8 + 2 * t] = c[0]; d[g * 8 + 2 * t + 1] = c[1];
    d[(g + 8) * 8 + 2 * t] = c[2]; d[(g + 8) * 8 + 2 * t + 1] = c[3];
}

int main() {
    uint32_t *d_out; cudaMalloc(&d_out, 128 * 4);
    probe<<<1, 32>>>(d_out);
    uint32_t h[128];
    cudaMemcpy(h, d_out, sizeof h, cudaMemcpyDeviceToHost);
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
    for (int l = 0; l < 32; l++) {
        printf("lane %2d (g=%d,t=%d):", l, l >> 2, l & 3);
        for (int r = 0; r < 4; r++) {
            printf("  r%d=", r);
            for (int b = 0; b < 4; b++) { unsigned v = (h[l * 4 + r] >> (8 * b)) & 255; printf("[k%2u m%2u]", v >> 4 | 0, v & 15); }
        }
        printf("\n");
    }
    // note: value = k*16+m mod 256, so k printed is k mod 16; rows 16..31 alias rows 0..15 (r2,r3 should show them)
    uint8_t hs[512]; int8_t hw[256]; int hd[128], ref[128];
    srand(1);
    for (int i = 0; i < 512; i++) hs[i] = rand() & 255;
    for (int i = 0; i < 256; i++) hw[i] = int8_t(rand() & 255);
    for (int m = 0; m < 16; m++) for (int n = 0; n < 8; n++) { int s = 0; for (int k = 0; k < 32; k++) s += int(hs[k * 16 + m]) * int(hw[k * 8 + n]); ref[m * 8 + n] = s; }
    uint8_t *ds; int8_t *dw; int *dd;
    cudaMalloc(&ds, 512); cudaMalloc(&dw, 256); cudaMalloc(&dd, 512);
    cudaMemcpy(ds, hs, 512, cudaMemcpyHostToDevice); cudaMemcpy(dw, hw, 256, cudaMemcpyHostToDevice);
    probe_mma<<<1, 32>>>(ds, dw, dd);
    cudaMemcpy(hd, dd, 512, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int i = 0; i < 128; i++) bad += hd[i] != ref[i];
    printf("ldmatrix.trans.b8 -> mma.m16n8k32 A fragment: %s (%d mismatches) err=%s\n", bad ? "MISMATCH" : "MATCH", bad, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
