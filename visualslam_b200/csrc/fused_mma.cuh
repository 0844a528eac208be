// fused_mma.cuh -- tap tables of the tensor-pipe variant of the fused octave kernel (fused_mma.cu).
#pragma once
#include "fused.cuh"

namespace vslam {

constexpr int kMmaPadL = 40;        // zero bytes in front of / behind the taps in the shifted tap copies
constexpr int kMmaPadR = 40;

// One Gaussian level of octave_mma_k.
struct MmaLevel {
    int n, r;
    int sv;                         // V pass: rows the window starts above (out row - r) so that it is quad aligned: -(rmax + r) mod 4
    int ksv;                        // V pass: 32-row K steps, ceil((n + 7 + sv) / 32)
    int cq;                         // quads between a row block's window start and the step's H chunk: (rmax + r + sv) / 4
    int np;                         // ring slots of this level's h planes, in PAIRS of row quads
    int plane_off;                  // word offset of the level's low-byte plane; the high-byte plane follows it
    int tab_off;                    // byte offset of the level's four byte-shifted tap copies
    int tab_cp;                     // bytes per copy: padL + n + padR rounded up to 4
};

struct MmaTaps {
    MmaLevel lv[6];
    u8 taps[6][248];
    int rmin, rmax;
    int stage_rows, stage_pitch;    // staged base tile: 32 rows (y0 + rmax ...), pitch a multiple of 128 bytes (XOR swizzle)
    int spq_magic;                  // ceil(2^16 / (stage_pitch / 16)): flattened index -> row without a division
    int plane_words;                // all six planes
    int tab_bytes;                  // all tap copies
};

int build_mma_taps(double sigma0, int scale_samples, int octave, MmaTaps *out);
size_t mma_smem_bytes(const MmaTaps &t);

}  // namespace vslam
