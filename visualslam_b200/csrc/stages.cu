// stages.cu -- the reference's per-function surface of ./Harris and the two-stage DoG detector, one small kernel per
// function, behind the C ABI (include/vslam.h "per-function entry points").  The throughput paths are vslam_harris and
// vslam_dog_keypoints (fused kernels); these exist so that the C++17 mirror of the reference (visualslam_b200/host/)
// can expose HarrisCorner / NonMaximumSuppression / NMS2 / filterKeypoints with the reference's own signatures.
//
//   GaussianBlur(Size(3,3), 0)          Harris_corners.cpp:158      blur3_k      (S + 8) >> 4, S = [1 2 1; 2 4 2; 1 2 1]
//   Sobel(CV_32F, ksize=1)              Harris_corners.cpp:161-164  sobel1_k     central differences, reflect-101
//   StructureMatrix + HarrisCorner      Harris_corners.cpp:10-68    harris_grad_k
//   convertScaleAbs                     Harris_corners.cpp:176      cvtabs_k
//   NonMaximumSuppression(., 3)         Harris_corners.cpp:70-81    nms3_k
//   NMS2(., window)                     Harris_corners.cpp:83-129   nms2_k
//   computeEdgeResponse + threshold     Diff_of_Gauss.cpp:79-109, 326-338   edge_list_k (one thread per keypoint)
#include "common.cuh"
#include "detect_math.cuh"
#include "stages.cuh"

namespace vslam {

__global__ void blur3_k(const u8 *__restrict__ src, int rows, int cols, size_t pitch, u8 *__restrict__ dst, size_t dpitch) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    int s = 0;
#pragma unroll
    for (int u = -1; u <= 1; ++u) {
        const u8 *r = src + (size_t)refl101(y + u, rows) * pitch;
        const int h = (int)r[refl101(x - 1, cols)] + 2 * (int)r[x] + (int)r[refl101(x + 1, cols)];
        s += (u == 0 ? 2 : 1) * h;
    }
    dst[(size_t)y * dpitch + x] = (u8)((s + 8) >> 4);
}

__global__ void sobel1_k(const u8 *__restrict__ src, int rows, int cols, size_t pitch, float *__restrict__ gx, float *__restrict__ gy) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    const u8 *r = src + (size_t)y * pitch;
    const size_t o = (size_t)y * cols + x;
    if (gx) gx[o] = (float)((int)r[refl101(x + 1, cols)] - (int)r[refl101(x - 1, cols)]);
    if (gy) gy[o] = (float)((int)src[(size_t)refl101(y + 1, rows) * pitch + x] - (int)src[(size_t)refl101(y - 1, rows) * pitch + x]);
}

// HarrisCorner on arbitrary float gradient maps: closed (2 pad + 1)^2 window on the replicate-padded maps, float accumulation in
// the reference's order (u outer, v inner; separate multiply and add), det through double, response = det - k * (tr * tr).
__global__ void harris_grad_k(const float *__restrict__ ix, const float *__restrict__ iy, int rows, int cols, int pad, float k,
                              float *__restrict__ out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    float a = 0.f, b = 0.f, c = 0.f;
    for (int u = y - pad; u <= y + pad; ++u)
        for (int v = x - pad; v <= x + pad; ++v) {
            const size_t o = (size_t)clampi(u, 0, rows - 1) * cols + clampi(v, 0, cols - 1);
            const float p = ix[o], q = iy[o];
            a = __fadd_rn(a, __fmul_rn(p, p));
            b = __fadd_rn(b, __fmul_rn(q, q));
            c = __fadd_rn(c, __fmul_rn(p, q));
        }
    const float det = __double2float_rn(__dsub_rn(__dmul_rn((double)a, (double)b), __dmul_rn((double)c, (double)c)));
    const float tr = __fadd_rn(a, b);
    const float resp = __fsub_rn(det, __fmul_rn(k, __fmul_rn(tr, tr)));
    out[(size_t)y * cols + x] = resp > 0.f ? resp : 0.f;
}

__global__ void cvtabs_k(const float *__restrict__ src, size_t n, u8 *__restrict__ dst) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // saturate_cast<uchar>(cvRound(|v|)), half to even; cvRound yields INT_MIN for |v| >= 2^31 (and NaN), which saturates to 0
    const float v = fabsf(src[i]);
    dst[i] = !(v < 2147483648.f) ? (u8)0 : v >= 255.5f ? (u8)255 : (u8)__float2int_rn(v);
}

__global__ void nms3_k(const u8 *__restrict__ r8, int rows, int cols, u8 *__restrict__ mask) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    int m = -1;                                                   // neighbours outside the image are ignored
    for (int u = -1; u <= 1; ++u)
        for (int v = -1; v <= 1; ++v) {
            if ((u | v) == 0) continue;
            const int yy = y + u, xx = x + v;
            if (yy < 0 || yy >= rows || xx < 0 || xx >= cols) continue;
            m = max(m, (int)r8[(size_t)yy * cols + xx]);
        }
    mask[(size_t)y * cols + x] = (int)r8[(size_t)y * cols + x] > m ? 255 : 0;
}

__global__ void nms2_k(const float *__restrict__ r, int rows, int cols, int pad, float *__restrict__ out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    float o = 0.f;
    if (y >= pad && y < rows - pad && x >= pad && x < cols - pad) {
        float m = 0.f;
        for (int u = y - pad; u < y + pad; ++u)
            for (int v = x - pad; v < x + pad; ++v) m = fmaxf(m, r[(size_t)u * cols + v]);
        if (r[(size_t)y * cols + x] >= m) o = m;
    }
    out[(size_t)y * cols + x] = o;
}

// one thread per keypoint of one octave: computeEdgeResponse on the level's Gaussian (gradients formed on the fly)
__global__ void edge_list_k(const vslam_keypoint *__restrict__ kp, int n, const u8 *const *__restrict__ gauss, int rows, int cols,
                            size_t pitch, float *__restrict__ resp) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const vslam_keypoint k = kp[i];
    float r = __int_as_float(0x7fc00000);                         // keypoints outside the maps: NaN, never kept
    if (k.level >= 0 && k.level < 6 && k.row >= 1 && k.row < rows && k.col >= 1 && k.col < cols) {
        const u8 *g = gauss[k.level];
        float a = 0.f, b = 0.f, c = 0.f;
        for (int u = k.row - 1; u < k.row + 1; ++u)
            for (int v = k.col - 1; v < k.col + 1; ++v) {
                const float p = (float)((int)g[(size_t)u * pitch + refl101(v + 1, cols)] - (int)g[(size_t)u * pitch + refl101(v - 1, cols)]);
                const float q = (float)((int)g[(size_t)refl101(u + 1, rows) * pitch + v] - (int)g[(size_t)refl101(u - 1, rows) * pitch + v]);
                a = __fadd_rn(a, __fmul_rn(p, p));
                b = __fadd_rn(b, __fmul_rn(q, q));
                c = __fadd_rn(c, __fmul_rn(p, q));
            }
        r = edge_from_sums(a, b, c);
    }
    resp[i] = r;
}

// FeaturePointLocalization's dog_zhat for n candidates given as (dx, dy, ds, value) quadruples (Diff_of_Gauss.cpp:223-243)
__global__ void localize_list_k(const int4 *__restrict__ q, int n, float *__restrict__ zhat) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) zhat[i] = localize(q[i].x, q[i].y, q[i].z, q[i].w);
}

// computeEdgeResponse's arithmetic on n windows of `len` gradient pairs (Diff_of_Gauss.cpp:93-107), float accumulation in order
__global__ void edge_windows_k(const float *__restrict__ gx, const float *__restrict__ gy, int n, int len, float *__restrict__ resp,
                               float *__restrict__ sums) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float a = 0.f, b = 0.f, c = 0.f;
    for (int j = 0; j < len; ++j) {
        const float p = gx[(size_t)i * len + j], q = gy[(size_t)i * len + j];
        a = __fadd_rn(a, __fmul_rn(p, p));
        b = __fadd_rn(b, __fmul_rn(q, q));
        c = __fadd_rn(c, __fmul_rn(p, q));
    }
    if (resp) resp[i] = edge_from_sums(a, b, c);
    if (sums) { sums[3 * i] = a; sums[3 * i + 1] = b; sums[3 * i + 2] = c; }   // Ix2, Iy2, IxIy (StructureMatrix, Harris_corners.cpp:10-29)
}

// cv::cvtColor(COLOR_BGR2GRAY) on 8U: (3735 B + 19235 G + 9798 R + 2^14) >> 15 (OpenCV's 15-bit fixed point, exact
// against cv2 4.13 over all 2^24 triples).  Thread = four pixels: three words in, one word out where the rows allow it.
__global__ void bgr2gray_k(const u8 *__restrict__ src, int rows, int cols, size_t pitch, size_t stride, u8 *__restrict__ dst,
                           size_t dpitch, size_t dstride, int vec) {
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x4 >= cols) return;
    const u8 *s = src + (size_t)blockIdx.z * stride + (size_t)y * pitch + 3 * (size_t)x4;
    u8 *d = dst + (size_t)blockIdx.z * dstride + (size_t)y * dpitch + x4;
    auto gray = [](u32 b, u32 g, u32 r) -> u32 { return (3735u * b + 19235u * g + 9798u * r + 16384u) >> 15; };
    if (vec && x4 + 3 < cols) {
        const u32 w0 = ((const u32 *)s)[0], w1 = ((const u32 *)s)[1], w2 = ((const u32 *)s)[2];     // B0G0R0B1 G1R1B2G2 R2B3G3R3
        const u32 g0 = gray(w0 & 255, (w0 >> 8) & 255, (w0 >> 16) & 255), g1 = gray(w0 >> 24, w1 & 255, (w1 >> 8) & 255);
        const u32 g2 = gray((w1 >> 16) & 255, w1 >> 24, w2 & 255), g3 = gray((w2 >> 8) & 255, (w2 >> 16) & 255, w2 >> 24);
        *(u32 *)d = g0 | g1 << 8 | g2 << 16 | g3 << 24;
    } else {
        for (int j = 0; j < 4 && x4 + j < cols; ++j) d[j] = (u8)gray(s[3 * j], s[3 * j + 1], s[3 * j + 2]);
    }
}

// Rotation::rotate_pt_CW (rotation.cpp:19-28): float products of the int offsets, one float subtraction / addition,
// truncation toward zero -- no FMA contraction (the reference is x86-64 code built without -mfma)
__device__ __forceinline__ void rot_pt_cw(int x, int y, int cx, int cy, float c, float s, int &xr, int &yr) {
    const float rx = (float)(x - cx), ry = (float)(y - cy);
    xr = (int)__fsub_rn(__fmul_rn(rx, c), __fmul_rn(ry, s)) + cx;
    yr = (int)__fadd_rn(__fmul_rn(rx, s), __fmul_rn(ry, c)) + cy;
}
// mode 0: Rotation::rotate_mat_CCW (rotation.cpp:66-84): output pixel (x, y) = I[rot] when 0 < rot.x < cols and
//         0 < rot.y < rows (both strict, as written), else 0.
// mode 1: Rotation::getRotatedWindow (rotation.cpp:87-108): output (i, j) = I[rot(x0 + j, y0 + i)]; the reference does
//         not test the rotated point, reads outside the image give 0 here.
__global__ void rotate_gather_k(const u8 *__restrict__ src, int rows, int cols, size_t pitch, int cx, int cy, float c, float s,
                                int x0, int y0, int ow, int oh, u8 *__restrict__ dst, size_t dpitch, int mode) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= ow || i >= oh) return;
    int xr, yr;
    rot_pt_cw(x0 + j, y0 + i, cx, cy, c, s, xr, yr);
    const bool in = mode == 0 ? (xr < cols && xr > 0 && yr < rows && yr > 0) : (xr < cols && xr >= 0 && yr < rows && yr >= 0);
    dst[(size_t)i * dpitch + j] = in ? src[(size_t)yr * pitch + xr] : (u8)0;
}

// cv::warpAffine(src, dst, M, dsize) with the defaults Rotation::doubleCrop uses (rotation.cpp:198): INTER_LINEAR,
// BORDER_CONSTANT 0, 8UC1.  mi = the INVERTED matrix (the host inverts like imgwarp.cpp does).  OpenCV walks the destination
// in 10-bit fixed point: X0 = cvRound((mi1 y + mi2) 1024) + 16 per row, adelta = cvRound(mi0 x 1024) per column (doubles, each
// operation rounded on its own), keeps 5 fractional bits and blends four taps with the 15-bit weights of BilinearTab_i --
// (32 - fx)(32 - fy) 32 ..., except the (0, 0) entry, which is {32767, 0, 0, 1} (saturate_cast<short>(32768) + the sum fix-up).
struct WarpM { double m[6]; };
__global__ void warp_affine_k(const u8 *__restrict__ src, int rows, int cols, size_t pitch, const WarpM mi, u8 *__restrict__ dst, int drows,
                              int dcols, size_t dpitch) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= dcols || y >= drows) return;
    const int X0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(mi.m[1], (double)y), mi.m[2]), 1024.0)) + 16;
    const int Y0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(mi.m[4], (double)y), mi.m[5]), 1024.0)) + 16;
    const int X = (X0 + __double2int_rn(__dmul_rn(__dmul_rn(mi.m[0], (double)x), 1024.0))) >> 5;
    const int Y = (Y0 + __double2int_rn(__dmul_rn(__dmul_rn(mi.m[3], (double)x), 1024.0))) >> 5;
    const int sx = clampi(X >> 5, -32768, 32767), sy = clampi(Y >> 5, -32768, 32767);        // saturate_cast<short>
    const int fx = X & 31, fy = Y & 31;
    int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
    if (!(fx | fy)) { w00 = 32767; w11 = 1; }
    const bool x0in = (unsigned)sx < (unsigned)cols, x1in = (unsigned)(sx + 1) < (unsigned)cols;
    const bool y0in = (unsigned)sy < (unsigned)rows, y1in = (unsigned)(sy + 1) < (unsigned)rows;
    const u8 *r0 = src + (size_t)(y0in ? sy : 0) * pitch, *r1 = src + (size_t)(y1in ? sy + 1 : 0) * pitch;
    const int v00 = x0in && y0in ? r0[sx] : 0, v01 = x1in && y0in ? r0[sx + 1] : 0;
    const int v10 = x0in && y1in ? r1[sx] : 0, v11 = x1in && y1in ? r1[sx + 1] : 0;
    const int acc = (v00 * w00 + v01 * w01 + v10 * w10 + v11 * w11 + (1 << 14)) >> 15;
    dst[(size_t)y * dpitch + x] = (u8)min(acc, 255);
}

#define GRID2D(cols, rows) dim3(((cols) + 127) / 128, (rows)), 128, 0, ctx->stream

int launch_warp_affine(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, const double *minv, u8 *dst, int drows, int dcols,
                       size_t dpitch) {
    WarpM m;
    for (int i = 0; i < 6; ++i) m.m[i] = minv[i];
    KernelSpan span(ctx, VSLAM_K_UPSAMPLE);
    warp_affine_k<<<GRID2D(dcols, drows)>>>(src, rows, cols, pitch, m, dst, drows, dcols, dpitch);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

int launch_blur3(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, u8 *dst, size_t dpitch) {
    KernelSpan span(ctx, VSLAM_K_HARRIS);
    blur3_k<<<GRID2D(cols, rows)>>>(src, rows, cols, pitch, dst, dpitch);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_bgr2gray(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, size_t stride, u8 *dst, size_t dpitch,
                    size_t dstride, int n_frames) {
    const int vec = !(pitch & 3) && !(stride & 3) && !((uintptr_t)src & 3) && !(dpitch & 3) && !(dstride & 3) && !((uintptr_t)dst & 3);
    const dim3 grid(((cols + 3) / 4 + 127) / 128, rows, n_frames);
    KernelSpan span(ctx, VSLAM_K_UPSAMPLE);
    bgr2gray_k<<<grid, 128, 0, ctx->stream>>>(src, rows, cols, pitch, stride, dst, dpitch, dstride, vec);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_rotate_gather(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, int cx, int cy, float c, float s, int x0,
                         int y0, int ow, int oh, u8 *dst, size_t dpitch, int mode) {
    KernelSpan span(ctx, VSLAM_K_UPSAMPLE);
    rotate_gather_k<<<GRID2D(ow, oh)>>>(src, rows, cols, pitch, cx, cy, c, s, x0, y0, ow, oh, dst, dpitch, mode);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_sobel1(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, float *gx, float *gy) {
    KernelSpan span(ctx, VSLAM_K_GRAD);
    sobel1_k<<<GRID2D(cols, rows)>>>(src, rows, cols, pitch, gx, gy);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_harris_grad(vslam_ctx *ctx, const float *ix, const float *iy, int rows, int cols, int pad, float k, float *out) {
    KernelSpan span(ctx, VSLAM_K_HARRIS);
    harris_grad_k<<<GRID2D(cols, rows)>>>(ix, iy, rows, cols, pad, k, out);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_cvtabs(vslam_ctx *ctx, const float *src, size_t n, u8 *dst) {
    KernelSpan span(ctx, VSLAM_K_HARRIS);
    cvtabs_k<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(src, n, dst);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_nms3(vslam_ctx *ctx, const u8 *r8, int rows, int cols, u8 *mask) {
    KernelSpan span(ctx, VSLAM_K_HARRIS);
    nms3_k<<<GRID2D(cols, rows)>>>(r8, rows, cols, mask);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_nms2(vslam_ctx *ctx, const float *r, int rows, int cols, int pad, float *out) {
    KernelSpan span(ctx, VSLAM_K_HARRIS);
    nms2_k<<<GRID2D(cols, rows)>>>(r, rows, cols, pad, out);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_localize_list(vslam_ctx *ctx, const int *q, int n, float *zhat) {
    KernelSpan span(ctx, VSLAM_K_DETECT);
    localize_list_k<<<(n + 127) / 128, 128, 0, ctx->stream>>>((const int4 *)q, n, zhat);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_edge_windows(vslam_ctx *ctx, const float *gx, const float *gy, int n, int len, float *resp, float *sums) {
    KernelSpan span(ctx, VSLAM_K_DETECT);
    edge_windows_k<<<(n + 127) / 128, 128, 0, ctx->stream>>>(gx, gy, n, len, resp, sums);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}
int launch_edge_list(vslam_ctx *ctx, const vslam_keypoint *kp, int n, const u8 *const *gauss, int rows, int cols, size_t pitch, float *resp) {
    KernelSpan span(ctx, VSLAM_K_DETECT);
    edge_list_k<<<(n + 127) / 128, 128, 0, ctx->stream>>>(kp, n, gauss, rows, cols, pitch, resp);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
