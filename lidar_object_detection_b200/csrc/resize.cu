// Mask resize to camera resolution: cv2.resize(mask.astype(uint8),
// (W, H)) with the default INTER_LINEAR, the step right before the gather in
// extract_car_points_by_mask (cvs_erosion.py:153; V4:293; V5:531).  Bit-exact with
// OpenCV's uint8 path (checked against cv2 4.13 for binary and grey images):
//   * same size            -> copy
//   * exact 2x downscale   -> OpenCV switches INTER_LINEAR to the fast INTER_AREA: (sum of 2x2 + 2) >> 2
//   * otherwise            -> fixed-point bilinear, 11-bit coefficients:
//       fx = (float)((dx + 0.5) * scale_x - 0.5); sx = floor(fx); fx -= sx   (float32)
//       left / right border: fx = 0 and sx clamped; rows are clamped, fy is kept
//       a = { rint((1 - fx) * 2048), rint(fx * 2048) }, b likewise
//       row value r = S[sx] * a0 + S[sx + 1] * a1                           (int32)
//       out = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2
// Every float / double step is spelled with round-to-nearest intrinsics so that nothing is
// contracted into an fma the CPU code does not have.
#include "common.cuh"

namespace {

struct Tap {
    int s0, s1;     // source indices
    int c0, c1;     // 11-bit coefficients
};

__device__ __forceinline__ Tap tap(int d, double scale, int n, bool zero_at_border) {
    float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
    int s = (int)floorf(f);
    f = __fsub_rn(f, (float)s);
    Tap t;
    if (zero_at_border) {
        if (s < 0) { f = 0.f; s = 0; }
        if (s >= n - 1) { f = 0.f; s = n - 1; }
        t.s0 = s;
        t.s1 = min(s + 1, n - 1);
    } else {
        t.s0 = min(max(s, 0), n - 1);
        t.s1 = min(max(s + 1, 0), n - 1);
    }
    t.c0 = __float2int_rn(__fmul_rn(__fsub_rn(1.f, f), 2048.f));
    t.c1 = __float2int_rn(__fmul_rn(f, 2048.f));
    return t;
}

// One thread = 4 consecutive output pixels of one row (one 32-bit store when W % 4 == 0).
__global__ void __launch_bounds__(256)
lf_resize_linear_kernel(const uint8_t* __restrict__ src, int64_t m, int h, int w, uint8_t* __restrict__ dst, int H,
                        int W, double scale_x, double scale_y) {
    const int qpr = (W + 3) >> 2;
    const int64_t total = m * (int64_t)H * qpr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int q = (int)(i % qpr);
        const int64_t row = i / qpr;
        const int dy = (int)(row % H);
        const int64_t img = row / H;
        const Tap ty = tap(dy, scale_y, h, false);
        const uint8_t* r0 = src + (img * h + ty.s0) * (int64_t)w;
        const uint8_t* r1 = src + (img * h + ty.s1) * (int64_t)w;
        uint8_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int dx = 4 * q + j;
            if (dx >= W) { o[j] = 0; continue; }
            const Tap tx = tap(dx, scale_x, w, true);
            const int a = (int)__ldg(r0 + tx.s0) * tx.c0 + (int)__ldg(r0 + tx.s1) * tx.c1;
            const int b = (int)__ldg(r1 + tx.s0) * tx.c0 + (int)__ldg(r1 + tx.s1) * tx.c1;
            o[j] = (uint8_t)((((ty.c0 * (a >> 4)) >> 16) + ((ty.c1 * (b >> 4)) >> 16) + 2) >> 2);
        }
        uint8_t* out = dst + (img * H + dy) * (int64_t)W + 4 * q;
        if ((W & 3) == 0) *(uchar4*)out = make_uchar4(o[0], o[1], o[2], o[3]);
        else
            for (int j = 0; j < 4 && 4 * q + j < W; ++j) out[j] = o[j];
    }
}

__global__ void __launch_bounds__(256)
lf_resize_half_kernel(const uint8_t* __restrict__ src, int64_t m, int H, int W, uint8_t* __restrict__ dst) {
    const int w = 2 * W;
    const int64_t total = m * (int64_t)H * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int dx = (int)(i % W);
        const int64_t row = i / W;                           // img * H + dy
        const uint8_t* p = src + row * 2 * (int64_t)w + 2 * dx;
        dst[i] = (uint8_t)(((int)__ldg(p) + (int)__ldg(p + 1) + (int)__ldg(p + w) + (int)__ldg(p + w + 1) + 2) >> 2);
    }
}

}  // namespace

extern "C" int lf_resize_masks(lf_ctx* ctx, const uint8_t* src, int64_t m, int h, int w, uint8_t* dst, int H, int W,
                               void* stream) {
    if (!ctx || m < 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) return LF_E_ARG;
    if (m == 0) return LF_OK;
    if (!src || !dst) return LF_E_ARG;
    if ((int64_t)h * w >= ((int64_t)1 << 31) || (int64_t)H * W >= ((int64_t)1 << 31)) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (h == H && w == W) {
        LF_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)m * h * w, cudaMemcpyDeviceToDevice, st));
        return LF_OK;
    }
    const int64_t cap = (int64_t)ctx->sm_count * 16;
    if (w == 2 * W && h == 2 * H) {
        int64_t blocks = (m * H * W + 255) / 256;
        if (blocks > cap) blocks = cap;
        lf_resize_half_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, m, H, W, dst);
    } else {
        // OpenCV: inv_scale = (double)dsize / ssize; scale = 1. / inv_scale
        const double scale_x = 1.0 / ((double)W / (double)w), scale_y = 1.0 / ((double)H / (double)h);
        int64_t blocks = (m * H * ((W + 3) / 4) + 255) / 256;
        if (blocks > cap) blocks = cap;
        lf_resize_linear_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, m, h, w, dst, H, W, scale_x, scale_y);
    }
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
