// Headless colour buffers -- what the reference hands to Open3D /
// matplotlib, written to device memory instead of a window:
//   * per-point RGB: car colour per mask, grey 0.5 for points no mask claimed
//     (V4_BBox_IoU_filtering.py:290-307; V5_ProjectingBBoxes.py:524-550; bg_assigned union)
//   * depth map overlay: jet(depthMap / max) over the segmented image where depthMap > 0,
//     uint8 truncation, RGB -> BGR (seg_with_pointcloud.py:173-180)
#include "common.cuh"

namespace {

// colour of the LAST mask that contains the point: V4 / V5 / vis.py build one point cloud per
// mask in mask order and draw them in that order, so at equal depth the later one is on top;
// first_wins: the per-point loop of Same_color.py:112-127 stops at the first mask that matches.
__global__ void __launch_bounds__(256)
lf_point_colors_kernel(const unsigned long long* __restrict__ bits, int64_t n, const double* __restrict__ colors,
                       int m, double bg0, double bg1, double bg2, double* __restrict__ rgb,
                       uint8_t* __restrict__ assigned, int first_wins) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned long long b = bits[i];
        if (m < 64) b &= (1ULL << m) - 1ULL;
        double r = bg0, g = bg1, bl = bg2;
        if (b) {
            const int c = first_wins ? __ffsll((long long)b) - 1 : 63 - __clzll((long long)b);
            r = colors[3 * c]; g = colors[3 * c + 1]; bl = colors[3 * c + 2];
        }
        rgb[3 * i] = r; rgb[3 * i + 1] = g; rgb[3 * i + 2] = bl;
        if (assigned) assigned[i] = b ? 1 : 0;
    }
}

// np.uint8(x * 255) for x in [0, 1]: truncation toward zero
__device__ __forceinline__ uint8_t to_u8(double x) { return (uint8_t)(int)__dmul_rn(x, 255.0); }

__global__ void __launch_bounds__(256)
lf_depth_overlay_kernel(const double* __restrict__ depth, const double* __restrict__ dmax_p,
                        const uint8_t* __restrict__ image_rgb, int64_t image_stride, const double* __restrict__ lut,
                        int lut_n, int64_t npix, uint8_t* __restrict__ out_bgr) {
    // blockIdx.y = map index
    depth += (int64_t)blockIdx.y * npix;
    image_rgb += (int64_t)blockIdx.y * image_stride;
    out_bgr += (int64_t)blockIdx.y * npix * 3;
    const double dmax = dmax_p[blockIdx.y];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const double d = depth[i];
        double r, g, b;
        if (d > 0.0) {
            // Colormap.__call__ on floats: xa = X * N; xa == N -> N - 1; astype(int)
            double xa = __dmul_rn(__ddiv_rn(d, dmax), (double)lut_n);
            int idx = xa == (double)lut_n ? lut_n - 1 : (int)xa;
            idx = min(max(idx, 0), lut_n - 1);
            r = lut[3 * idx]; g = lut[3 * idx + 1]; b = lut[3 * idx + 2];
        } else {
            r = __ddiv_rn((double)image_rgb[3 * i], 255.0);
            g = __ddiv_rn((double)image_rgb[3 * i + 1], 255.0);
            b = __ddiv_rn((double)image_rgb[3 * i + 2], 255.0);
        }
        out_bgr[3 * i] = to_u8(b);          // cv2.COLOR_RGB2BGR
        out_bgr[3 * i + 1] = to_u8(g);
        out_bgr[3 * i + 2] = to_u8(r);
    }
}

// The same overlay, four pixels per thread.  Both per-pixel float64 chains have at most 256 / lut_n distinct inputs,
// so a CTA evaluates them once -- with the very same operations -- into shared memory: s_div[x] = uint8(x / 255.0 * 255)
// for the image bytes, s_lut[i] = the three uint8(lut[i] * 255) packed as BGR.  A pixel is then one 8-byte load, one
// division (the colormap index) or three table reads; image and output move as 32-bit words.
constexpr int OV_LUT_MAX = 1024;

__global__ void __launch_bounds__(256)
lf_depth_overlay_quad_kernel(const double* __restrict__ depth, const double* __restrict__ dmax_p,
                             const uint8_t* __restrict__ image_rgb, int64_t image_stride, const double* __restrict__ lut,
                             int lut_n, int64_t npix, uint8_t* __restrict__ out_bgr) {
    __shared__ uint8_t s_div[256];
    __shared__ uint32_t s_lut[OV_LUT_MAX];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_div[i] = to_u8(__ddiv_rn((double)i, 255.0));
    for (int i = threadIdx.x; i < lut_n; i += blockDim.x)
        s_lut[i] = (uint32_t)to_u8(lut[3 * i + 2]) | ((uint32_t)to_u8(lut[3 * i + 1]) << 8) | ((uint32_t)to_u8(lut[3 * i]) << 16);
    __syncthreads();
    depth += (int64_t)blockIdx.y * npix;
    image_rgb += (int64_t)blockIdx.y * image_stride;
    out_bgr += (int64_t)blockIdx.y * npix * 3;
    const double dmax = dmax_p[blockIdx.y];
    const int64_t nq = npix >> 2;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += (int64_t)gridDim.x * blockDim.x) {
        const double2 d01 = __ldg((const double2*)depth + 2 * q), d23 = __ldg((const double2*)depth + 2 * q + 1);
        const double d[4] = {d01.x, d01.y, d23.x, d23.y};
        const uint32_t* ip = (const uint32_t*)image_rgb + 3 * q;
        const uint32_t i0 = __ldg(ip), i1 = __ldg(ip + 1), i2 = __ldg(ip + 2);      // r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
        uint32_t px[4];                                       // per pixel: B | G << 8 | R << 16
        px[0] = __byte_perm(i0, 0, 0x4012);
        px[1] = __byte_perm(i0, i1, 0x4345) & 0x00FFFFFFu;    // b1 = i1.1, g1 = i1.0, r1 = i0.3
        px[2] = __byte_perm(i1, i2, 0x4234) & 0x00FFFFFFu;    // b2 = i2.0, g2 = i1.3, r2 = i1.2
        px[3] = __byte_perm(i2, 0, 0x4123);                   // b3 = i2.3, g3 = i2.2, r3 = i2.1
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (d[j] > 0.0) {
                const double xa = __dmul_rn(__ddiv_rn(d[j], dmax), (double)lut_n);
                int idx = xa == (double)lut_n ? lut_n - 1 : (int)xa;
                idx = min(max(idx, 0), lut_n - 1);
                px[j] = s_lut[idx];
            } else {
                px[j] = (uint32_t)s_div[px[j] & 0xFF] | ((uint32_t)s_div[(px[j] >> 8) & 0xFF] << 8) |
                        ((uint32_t)s_div[px[j] >> 16] << 16);
            }
        }
        uint32_t* op = (uint32_t*)out_bgr + 3 * q;            // B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
        op[0] = px[0] | (px[1] << 24);
        op[1] = (px[1] >> 8) | (px[2] << 16);
        op[2] = (px[2] >> 16) | (px[3] << 8);
    }
}

}  // namespace

extern "C" int lf_point_colors(lf_ctx* ctx, const uint64_t* car_bits, int64_t n, const double* colors_rgb, int m,
                               const double bg_rgb[3], double* rgb, uint8_t* assigned, int policy, void* stream) {
    if (!ctx || n < 0 || m < 0 || m > LF_MAX_CARS || policy < 0 || policy > 1) return LF_E_ARG;
    if (n == 0) return LF_OK;
    if (!car_bits || !rgb || !bg_rgb || (m > 0 && !colors_rgb)) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)ctx->sm_count * 16;
    if (blocks > cap) blocks = cap;
    lf_point_colors_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const unsigned long long*)car_bits, n, colors_rgb, m, bg_rgb[0], bg_rgb[1], bg_rgb[2], rgb, assigned, policy);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_depth_overlay_batch(lf_ctx* ctx, const double* depth_maps, const double* depth_max,
                                      const uint8_t* image_rgb, int64_t image_stride, const double* lut_rgb,
                                      int lut_n, int n, int H, int W, uint8_t* out_bgr, void* stream) {
    if (!ctx || H <= 0 || W <= 0 || lut_n <= 0 || n < 0 || n > 65535 || image_stride < 0) return LF_E_ARG;
    if (n == 0) return LF_OK;
    if (!depth_maps || !depth_max || !image_rgb || !lut_rgb || !out_bgr) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t npix = (int64_t)H * W;
    int64_t blocks = (npix + 255) / 256;
    const int64_t cap = ((int64_t)ctx->sm_count * 16 + n - 1) / n;
    if (blocks > cap) blocks = cap < 1 ? 1 : cap;
    // (a single frame is launch- and latency-bound: the one-pixel kernel spreads it over four times the threads
    // and has no tables to build)
    const bool quad = (int64_t)n * npix >= ((int64_t)4 << 20) && (npix & 3) == 0 && lut_n <= OV_LUT_MAX &&
                      ((uintptr_t)depth_maps & 15) == 0 &&
                      (((uintptr_t)image_rgb | (uintptr_t)out_bgr | (uintptr_t)image_stride) & 3) == 0;
    if (quad) {
        blocks = (npix / 4 + 255) / 256;
        if (blocks > cap) blocks = cap < 1 ? 1 : cap;
        lf_depth_overlay_quad_kernel<<<dim3((unsigned)blocks, (unsigned)n), 256, 0, (cudaStream_t)stream>>>(
            depth_maps, depth_max, image_rgb, image_stride, lut_rgb, lut_n, npix, out_bgr);
    } else
        lf_depth_overlay_kernel<<<dim3((unsigned)blocks, (unsigned)n), 256, 0, (cudaStream_t)stream>>>(
            depth_maps, depth_max, image_rgb, image_stride, lut_rgb, lut_n, npix, out_bgr);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_depth_overlay(lf_ctx* ctx, const double* depth_map, const double* depth_max,
                                const uint8_t* image_rgb, const double* lut_rgb, int lut_n, int H, int W,
                                uint8_t* out_bgr, void* stream) {
    return lf_depth_overlay_batch(ctx, depth_map, depth_max, image_rgb, 0, lut_rgb, lut_n, 1, H, W, out_bgr, stream);
}
