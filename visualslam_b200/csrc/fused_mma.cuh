// fused_mma.cuh -- tap tables of the tensor-pipe variant of the fused octave kernel (fused_mma.cu).
#pragma once
#include "fused.cuh"

namespace vslam {

// Tap tables: per level four byte-shifted, zero-padded copies of the taps, every copy `cp` bytes (one size for all
// levels).  Level l's taps start `padl` bytes into a copy; padl is large enough that every chunk of the H pass's K range
// that lies outside the band reads zeros (all chunks are multiplied, no band test), and padl + r is a multiple of 4, so
// that the byte shift and the word offset of a lane's B fragment are the same for all levels: one lane pointer per pass,
// the levels are immediates (tab_hoff / tab_voff).

// One Gaussian level of octave_mma_k.
struct MmaLevel {
    int n, r;
    int sv;                         // V pass: rows the window starts above (out row - r) so that it is quad aligned: -(rmax + r) mod 4
    int ksv;                        // V pass: 32-row K steps, ceil((n + 7 + sv) / 32)
    int cq;                         // quads between a row block's window start and the step's H chunk: (rmax + r + sv) / 4
    int np;                         // ring slots of this level's h planes, in PAIRS of row quads
    int plane_off;                  // word offset of the level's low-byte plane; the high-byte plane follows it
    int padl;                       // zero bytes in front of the taps in a copy: >= rmax - r + 41, padl + r == 0 (mod 4)
    int tab_hoff, tab_voff;         // H / V pass: byte offset of the level's B fragments from the lane pointer of the pass
};

struct MmaTaps {
    MmaLevel lv[6];
    u8 taps[6][248];
    int rmin, rmax;
    int rv;                         // rows between a step's first output row and its H chunk: rmax + 0..3, the offset that
                                    // quad-aligns the V windows with the fewest K steps
    int stage_rows, stage_pitch;    // staged base tile: 32 rows (y0 + rmax ...), pitch a multiple of 128 bytes (XOR swizzle)
    int spq_magic;                  // ceil(2^16 / (stage_pitch / 16)): flattened index -> row without a division
    int plane_words;                // all six planes
    int fw;                         // columns a CTA computes (64 or 32)
    int nch;                        // 32-byte chunks of a staged row one 8-column block's bands can touch (from the band's 16-byte group on)
    int cp;                         // bytes per tap copy
    int tab_bytes;                  // all tap copies (6 levels x 4 copies)
};

// Strip widths: a CTA computes FW columns and owns FWO of them (2 halo columns each side; FWO a multiple of 6: sample
// columns of stride 3, even start on the next octave's base).  64 / 60 by default; 32 / 24 for octaves whose rings do
// not fit shared memory at 64 columns.
__host__ __device__ constexpr int mma_fwo(int fw) { return fw == 64 ? 60 : 24; }
// words per quad PAIR of an h plane: FW / 2 column pairs x 16 bytes, + 8 (== 8 mod 32: conflict-free LDS.128)
__host__ __device__ constexpr int mma_ppw(int fw) { return 4 * (fw / 2) + 8; }

// The same geometry as compile-time functions of the six kernel sizes: the default pyramid (sigma0 1.6, three scale
// samples) runs instantiations whose plane / table offsets, ring sizes and trip counts are immediates.  build_mma_taps
// stays the single source of the layout: the launcher compares every field of the two before it picks a static kernel.
template <int FW_, int N0, int N1, int N2, int N3, int N4, int N5>
struct MmaStaticGeom {
    static constexpr bool is_static = true;
    __host__ __device__ constexpr MmaStaticGeom() {}
    __host__ __device__ explicit constexpr MmaStaticGeom(const MmaTaps &) {}
    __host__ __device__ static constexpr int n(int l) { return l == 0 ? N0 : l == 1 ? N1 : l == 2 ? N2 : l == 3 ? N3 : l == 4 ? N4 : N5; }
    __host__ __device__ static constexpr int r(int l) { return n(l) / 2; }
    __host__ __device__ static constexpr int rmax() {
        int m = 0;
        for (int l = 0; l < 6; ++l) m = r(l) > m ? r(l) : m;
        return m;
    }
    __host__ __device__ static constexpr int sv_at(int l, int v) { return (4 - (v + r(l)) % 4) % 4; }
    __host__ __device__ static constexpr int ksteps_at(int v) {
        int s = 0;
        for (int l = 0; l < 6; ++l) s += (n(l) + 7 + sv_at(l, v) + 31) / 32;
        return s;
    }
    __host__ __device__ static constexpr int rv() {
        int best = rmax();
        for (int v = rmax() + 1; v < rmax() + 4; ++v) best = ksteps_at(v) < ksteps_at(best) ? v : best;
        return best;
    }
    __host__ __device__ static constexpr int sv(int l) { return sv_at(l, rv()); }
    __host__ __device__ static constexpr int ksv(int l) { return (n(l) + 7 + sv(l) + 31) / 32; }
    // V pass of the static kernels: the row blocks of a warp (two at 64 columns, one at 32) share the A fragments of a K step -- row
    // block rb reads the same 32 rows through taps shifted by 8 rb.  K steps that cover the last block's window; which (block, step)
    // products are not all zeros
    __host__ __device__ static constexpr int rbw() { return FW_ / 32; }
    __host__ __device__ static constexpr int ksvs(int l) { return (n(l) + 7 + sv(l) + 8 * (rbw() - 1) + 31) / 32; }
    __host__ __device__ static constexpr bool v_needed(int l, int k, int rb) { return 32 * k < n(l) + 7 + sv(l) + 8 * rb; }
    __host__ __device__ static constexpr int cq(int l) { return (rv() + r(l) + sv(l)) / 4; }
    __host__ __device__ static constexpr int np(int l) { return (cq(l) + 10) / 2 < 8 ? 8 : (cq(l) + 10) / 2; }
    __host__ __device__ static constexpr int plane_off(int l) {
        int s = 0;
        for (int j = 0; j < l; ++j) s += 2 * np(j) * mma_ppw(FW_);
        return s;
    }
    __host__ __device__ static constexpr int nch() {                       // rounded up to an instantiated chunk count
        const int c = (2 * rmax() + 23 + 31) / 32;
        return FW_ == 64 ? (c <= 2 ? 2 : c <= 3 ? 3 : 5) : (c <= 5 ? 5 : 9);
    }
    __host__ __device__ static constexpr int padl(int l) { return rmax() - r(l) + 41 + (4 - (rmax() + 41) % 4) % 4; }
    __host__ __device__ static constexpr int cp_need(int l) {
        int a = padl(l) + r(l) - rmax() + 12 + 32 * (nch() - 1) + 20;     // H pass: last word of the last chunk
        int b = padl(l) + 24 + 32 * (ksvs(l) - 1) + 8;                    // V pass: last word of the last K step
        int c = padl(l) + n(l) + 3;                                       // the taps themselves, shifted by up to 3
        return a > b ? (a > c ? a : c) : (b > c ? b : c);
    }
    __host__ __device__ static constexpr int cp() {
        int m = 0;
        for (int l = 0; l < 6; ++l) m = cp_need(l) > m ? cp_need(l) : m;
        m = (m + 4 + 3) & ~3;                                              // 4 spare bytes: copies s > rv % 4 start one word later
        while (m % 128 != 32) m += 4;                                      // copies 32 bytes apart modulo 128 (build_mma_taps: bank tiling)
        return m;
    }
    __host__ __device__ static constexpr int tab_hoff(int l) { return 4 * cp() * l + padl(l) + r(l); }
    __host__ __device__ static constexpr int tab_voff(int l) { return 4 * cp() * l + padl(l) - sv(l) - rv() % 4; }
    __host__ __device__ static constexpr int plane_words() { return plane_off(6); }
    // the V pass as a flat list of (level, K step) items
    __host__ __device__ static constexpr int items() { return ksvs(0) + ksvs(1) + ksvs(2) + ksvs(3) + ksvs(4) + ksvs(5); }
    __host__ __device__ static constexpr int item_l(int i) {
        int l = 0;
        while (l < 5 && i >= ksvs(l)) { i -= ksvs(l); ++l; }
        return l;
    }
    __host__ __device__ static constexpr int item_k(int i) {
        int l = 0;
        while (l < 5 && i >= ksvs(l)) { i -= ksvs(l); ++l; }
        return i;
    }
    __host__ __device__ static constexpr int stage_pitch() { return (2 * rmax() + 15 + FW_ + 127) / 128 * 128; }
    __host__ __device__ static constexpr int spq_magic() { return ((1 << 16) + (stage_pitch() >> 4) - 1) / (stage_pitch() >> 4); }
    static bool matches(const MmaTaps &t) {
        bool ok = t.fw == FW_ && t.rmax == rmax() && t.rv == rv() && t.stage_pitch == stage_pitch() && t.spq_magic == spq_magic() && t.plane_words == plane_words() &&
                  t.nch == nch() && t.cp == cp();
        for (int l = 0; l < 6; ++l)
            ok = ok && t.lv[l].n == n(l) && t.lv[l].r == r(l) && t.lv[l].sv == sv(l) && t.lv[l].ksv == ksv(l) && t.lv[l].cq == cq(l) &&
                 t.lv[l].np == np(l) && t.lv[l].plane_off == plane_off(l) && t.lv[l].padl == padl(l) && t.lv[l].tab_hoff == tab_hoff(l) &&
                 t.lv[l].tab_voff == tab_voff(l);
        return ok;
    }
};

// ... and read from the kernel argument (any sigma0 / scale samples)
struct MmaDynGeom {
    static constexpr bool is_static = false;
    const MmaTaps &T;
    __device__ explicit MmaDynGeom(const MmaTaps &t) : T(t) {}
    __device__ int n(int l) const { return T.lv[l].n; }
    __device__ int r(int l) const { return T.lv[l].r; }
    __device__ int rmax() const { return T.rmax; }
    __device__ int rv() const { return T.rv; }
    __device__ int sv(int l) const { return T.lv[l].sv; }
    __device__ int ksv(int l) const { return T.lv[l].ksv; }
    __device__ int cq(int l) const { return T.lv[l].cq; }
    __device__ int np(int l) const { return T.lv[l].np; }
    __device__ int plane_off(int l) const { return T.lv[l].plane_off; }
    __device__ int padl(int l) const { return T.lv[l].padl; }
    __device__ int cp() const { return T.cp; }
    __device__ int tab_hoff(int l) const { return T.lv[l].tab_hoff; }
    __device__ int tab_voff(int l) const { return T.lv[l].tab_voff; }
    __device__ int plane_words() const { return T.plane_words; }
    __device__ int stage_pitch() const { return T.stage_pitch; }
    __device__ int spq_magic() const { return T.spq_magic; }
};

int build_mma_taps(double sigma0, int scale_samples, int octave, int fw, MmaTaps *out);
size_t mma_smem_bytes(const MmaTaps &t);

}  // namespace vslam
