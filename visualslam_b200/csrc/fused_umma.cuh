// fused_umma.cuh -- geometry of the tcgen05 / TMA variant of the fused octave kernel (fused_umma.cu).
#pragma once
#include "fused.cuh"

namespace vslam {

// A CTA owns a strip of 48 columns (X0 = 48 * strip: the TMA box must start on a 16-byte boundary of the plane) and
// marches down it in steps of 128 rows (the M of one tcgen05.mma).  Per step and level:
//   V pass  D_v[128 out rows x NV cols]  = T_v[128 x 32 KV] * Img[32 KV rows x NV cols]       (u8 x u8 -> s32 in TMEM)
//   H pass  D_h[128 rows x 64 out cols]  = X[128 x 32 KH]   * T_h[64 x 32 KH]^T, once per byte plane of the 8.8 sums
// NV = Pc + 48 + rmax columns rounded to 16 (the staged row starts Pc columns left of the strip), KV = (128 + 2 P) / 32
// row chunks, KH = ceil(NV / 32).  Two instantiations: octave 0 of the default pyramid (rmax <= 15) and octave 1 (rmax <= 31).
constexpr int kUmmaFWO = 48;        // columns a CTA owns (16 sample columns of stride 3)
constexpr int kUmmaRows = 128;      // rows per marching step

struct UmmaGeom {
    int n[6], r[6], rmax;
    int P;                          // rows staged above a step's first output row = columns staged left of the strip (16 or 32)
    int KV, KH, NV;
    int JV, JH;                     // 8 x 16-byte core matrices per level of the compact Toeplitz tables (V / H pass)
    int tab_bytes;                  // 6 * (JV + JH) * 128
    u8 taps[6][64];
};

// 0 when one of the instantiations runs this octave, else non-zero
int umma_geometry(double sigma0, int scale_samples, int octave, UmmaGeom *g);
inline int umma_strips(int cols) { return (cols + kUmmaFWO - 1) / kUmmaFWO; }

}  // namespace vslam
