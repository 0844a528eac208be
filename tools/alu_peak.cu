// alu_peak.cu -- measures the non-tensor issue peaks this path is bound by (BASELINE.md section 2 asks for it):
// FFMA, IMAD, IDP.4A (dp4a), IDP.2A (dp2a) warp-instructions per second on the whole chip.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/alu_peak tools/alu_peak.cu ; prints one JSON line.
#include <cstdio>
#include <cuda_runtime.h>

template <int OP> __global__ void __launch_bounds__(256) peak_k(unsigned *out, int iters, unsigned seed) {
    unsigned a[16];
    float f[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { a[i] = seed + threadIdx.x * 17 + i; f[i] = (float)a[i]; }
    unsigned b = seed * 3 + 1, c = seed ^ 0x01010101u;
    float fb = 1.0001f, fc = 0.5f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                if (OP == 0) f[i] = __fmaf_rn(f[i], fb, fc);
                if (OP == 1) a[i] = a[i] * b + c;
                if (OP == 2) a[i] = __dp4a(b, c, a[i]);
                if (OP == 3) a[i] = __dp2a_lo(b, c, a[i]);
                if (OP == 4) a[i] = __dp4a(a[(i + 1) & 15], c, a[i]);   // register operand varies
            }
    }
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i] + (unsigned)f[i];
    if (s == 0x12345678u) out[0] = s;
}

template <int OP> double run(int sms) {
    unsigned *d; cudaMalloc(&d, 4);
    const int iters = 4096, blocks = sms * 8;
    peak_k<OP><<<blocks, 256>>>(d, 64, 1);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        peak_k<OP><<<blocks, 256>>>(d, iters, rep + 2);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaFree(d);
    const double lane_ops = (double)blocks * 256 * iters * 64;
    return lane_ops / (best * 1e-3);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    const double ffma = run<0>(sms), imad = run<1>(sms), dp4 = run<2>(sms), dp2 = run<3>(sms), dp4r = run<4>(sms);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"ffma_Tlane_s\": %.2f, \"imad_Tlane_s\": %.2f, \"dp4a_Tlane_s\": %.2f, "
           "\"dp2a_Tlane_s\": %.2f, \"dp4a_regdep_Tlane_s\": %.2f, \"dp4a_TMAC_s\": %.2f, \"dp2a_TMAC_s\": %.2f}\n",
           p.name, sms, ffma / 1e12, imad / 1e12, dp4 / 1e12, dp2 / 1e12, dp4r / 1e12, dp4 * 4 / 1e12, dp2 * 2 / 1e12);
    return 0;
}
