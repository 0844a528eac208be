// imma_peak.cu -- issue peak of the warp-level integer tensor instruction (mma.sync.m16n8k32 u8 x u8 -> s32, SASS IMMA)
// on the whole chip, plus a fragment-layout self check (banded Toeplitz product of random bytes against a scalar loop).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/imma_peak tools/imma_peak.cu ; prints one JSON line.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void imma(unsigned (&d)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int NACC> __global__ void __launch_bounds__(256) peak_k(unsigned *out, int iters, unsigned seed) {
    unsigned d[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) d[i][j] = seed + i + j;
    unsigned a0 = seed * 3 + threadIdx.x, a1 = a0 ^ 0x01010101u, a2 = a0 + 7, a3 = a1 + 9, b0 = seed + 5, b1 = seed ^ 0x10101010u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < NACC; ++i) imma(d[i], a0, a1, a2, a3, b0, b1);
    }
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += d[i][0] + d[i][1] + d[i][2] + d[i][3];
    if (s == 0x12345678u) out[0] = s;
}

template <int NACC> double run(int sms, int warps_per_block, int blocks_per_sm) {
    unsigned *dp; cudaMalloc(&dp, 4);
    const int iters = 2048, blocks = sms * blocks_per_sm, nt = 32 * warps_per_block;
    peak_k<NACC><<<blocks, nt>>>(dp, 16, 1);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        peak_k<NACC><<<blocks, nt>>>(dp, iters, rep + 2);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaFree(dp);
    const double mma = (double)blocks * warps_per_block * iters * 4 * NACC;     // warp-level instructions
    return mma / (best * 1e-3);
}

// layout check: D[16 x 8] = A[16 x 32] * B[32 x 8], A row-major bytes, B given as B[k][n]
__global__ void check_k(const unsigned char *A, const unsigned char *B, int *D) {
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    auto ld4 = [&](const unsigned char *p) { return (unsigned)p[0] | (unsigned)p[1] << 8 | (unsigned)p[2] << 16 | (unsigned)p[3] << 24; };
    unsigned a0 = ld4(A + g * 32 + 4 * t), a1 = ld4(A + (g + 8) * 32 + 4 * t), a2 = ld4(A + g * 32 + 16 + 4 * t), a3 = ld4(A + (g + 8) * 32 + 16 + 4 * t);
    unsigned b0 = 0, b1 = 0;
    for (int i = 0; i < 4; ++i) {
        b0 |= (unsigned)B[(4 * t + i) * 8 + g] << (8 * i);
        b1 |= (unsigned)B[(16 + 4 * t + i) * 8 + g] << (8 * i);
    }
    unsigned d[4] = {0, 0, 0, 0};
    imma(d, a0, a1, a2, a3, b0, b1);
    D[g * 8 + 2 * t] = d[0]; D[g * 8 + 2 * t + 1] = d[1]; D[(g + 8) * 8 + 2 * t] = d[2]; D[(g + 8) * 8 + 2 * t + 1] = d[3];
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    // layout check
    unsigned char hA[512], hB[256]; int hD[128], ref[128];
    srand(7);
    for (auto &v : hA) v = rand() & 255;
    for (auto &v : hB) v = rand() & 255;
    for (int m = 0; m < 16; ++m) for (int n = 0; n < 8; ++n) { int s = 0; for (int k = 0; k < 32; ++k) s += hA[m * 32 + k] * hB[k * 8 + n]; ref[m * 8 + n] = s; }
    unsigned char *dA, *dB; int *dD;
    cudaMalloc(&dA, 512); cudaMalloc(&dB, 256); cudaMalloc(&dD, 512);
    cudaMemcpy(dA, hA, 512, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, 256, cudaMemcpyHostToDevice);
    check_k<<<1, 32>>>(dA, dB, dD);
    cudaMemcpy(hD, dD, 512, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int i = 0; i < 128; ++i) bad += hD[i] != ref[i];
    const double r1 = run<8>(sms, 4, 2), r2 = run<8>(sms, 8, 2), r3 = run<4>(sms, 4, 1), r4 = run<2>(sms, 4, 1);
    const double mac = 16.0 * 8 * 32;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"layout_mismatches\": %d, \"imma_m16n8k32_Tinst_s\": {\"8acc_8warps\": %.4f, \"8acc_16warps\": %.4f, "
           "\"4acc_4warps\": %.4f, \"2acc_4warps\": %.4f}, \"best_TMAC_s\": %.1f}\n",
           p.name, sms, bad, r1 / 1e12, r2 / 1e12, r3 / 1e12, r4 / 1e12, mac * (r1 > r2 ? r1 : r2) / 1e12);
    return 0;
}
