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

// Persistent CTAs over (image, output row) pairs.  The column taps depend only on dx, so a
// CTA computes them once into shared memory (8 bytes per output column); after that an
// output pixel costs two shared loads, four cached byte loads and a dozen integer
// instructions.  One thread = 4 consecutive output pixels (one 32-bit store when W % 4 == 0).
__global__ void __launch_bounds__(256)
lf_resize_linear_kernel(const uint8_t* __restrict__ src, int64_t m, int h, int w, uint8_t* __restrict__ dst, int H,
                        int W, double scale_x, double scale_y) {
    extern __shared__ int2 s_tap[];                         // per output column: {s0, (s1 - s0) | c0 << 8 | c1 << 20}; 4 * qpr entries
    int2* taps = s_tap;
    const int qpr = (W + 3) >> 2;
    // stored as [dx & 3][dx >> 2]: lane q reads taps[j * qpr + q], consecutive lanes hit consecutive banks
    for (int dx = threadIdx.x; dx < W; dx += blockDim.x) {
        const Tap t = tap(dx, scale_x, w, true);
        taps[(dx & 3) * qpr + (dx >> 2)] = make_int2(t.s0, (int)((unsigned)(t.s1 - t.s0) | ((unsigned)t.c0 << 8) | ((unsigned)t.c1 << 20)));   // s1 - s0 in {0,1}; c <= 2048 (12 bits)
    }
    __syncthreads();
    const int64_t rows = m * (int64_t)H;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const int dy = (int)(row % H);
        const int64_t img = row / H;
        const Tap ty = tap(dy, scale_y, h, false);
        const uint8_t* r0 = src + (img * h + ty.s0) * (int64_t)w;
        const uint8_t* r1 = src + (img * h + ty.s1) * (int64_t)w;
        uint8_t* orow = dst + row * (int64_t)W;
        for (int q = threadIdx.x; q < qpr; q += blockDim.x) {
            uint8_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int dx = 4 * q + j;
                if (dx >= W) { o[j] = 0; continue; }
                const int2 t = taps[j * qpr + q];
                const int s0 = t.x, s1 = t.x + (t.y & 1), c0 = ((unsigned)t.y >> 8) & 0xFFF, c1 = (unsigned)t.y >> 20;
                const int a = (int)__ldg(r0 + s0) * c0 + (int)__ldg(r0 + s1) * c1;
                const int b = (int)__ldg(r1 + s0) * c0 + (int)__ldg(r1 + s1) * c1;
                o[j] = (uint8_t)((((ty.c0 * (a >> 4)) >> 16) + ((ty.c1 * (b >> 4)) >> 16) + 2) >> 2);
            }
            if ((W & 3) == 0) *(uchar4*)(orow + 4 * q) = make_uchar4(o[0], o[1], o[2], o[3]);
            else
                for (int j = 0; j < 4 && 4 * q + j < W; ++j) orow[4 * q + j] = o[j];
        }
    }
}

// The same arithmetic with the source rows staged in shared memory: a CTA takes RS_ROWS
// output rows of one image at a time, copies the two source rows each of them needs
// (coalesced 16-byte loads when the row pitch allows), and the taps then read bytes from
// shared memory instead of issuing four byte-wide global loads per pixel (the load/store
// unit, not HBM, bounds the plain kernel).  Used when 2 * RS_ROWS source rows fit.
constexpr int RS_ROWS = 8;

__global__ void __launch_bounds__(256)
lf_resize_linear_smem_kernel(const uint8_t* __restrict__ src, int64_t m, int h, int w, uint8_t* __restrict__ dst,
                             int H, int W, double scale_x, double scale_y, int pitch) {
    extern __shared__ __align__(16) uint8_t rs_smem[];
    int2* taps = (int2*)rs_smem;                              // [W]
    uint8_t* rows = rs_smem + (((size_t)4 * ((W + 3) >> 2) * sizeof(int2) + 15) & ~(size_t)15);   // [RS_ROWS][2][pitch]
    __shared__ int s_c0[RS_ROWS], s_c1[RS_ROWS];
    const int qpr = (W + 3) >> 2;
    for (int dx = threadIdx.x; dx < W; dx += blockDim.x) {
        const Tap t = tap(dx, scale_x, w, true);
        taps[(dx & 3) * qpr + (dx >> 2)] =
            make_int2(t.s0, (int)((unsigned)(t.s1 - t.s0) | ((unsigned)t.c0 << 8) | ((unsigned)t.c1 << 20)));
    }
    const int blocks_per_img = (H + RS_ROWS - 1) / RS_ROWS;
    const int64_t items = m * (int64_t)blocks_per_img;
    const bool vec = (w & 15) == 0 && (((uintptr_t)src) & 15) == 0;
    for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
        const int64_t img = it / blocks_per_img;
        const int y0 = (int)(it - img * blocks_per_img) * RS_ROWS;
        const int ny = min(RS_ROWS, H - y0);
        __syncthreads();                                      // taps ready / previous tile consumed
        for (int r = 0; r < 2 * ny; ++r) {
            const Tap ty = tap(y0 + (r >> 1), scale_y, h, false);
            const uint8_t* g = src + (img * h + ((r & 1) ? ty.s1 : ty.s0)) * (int64_t)w;
            uint8_t* srow = rows + (size_t)r * pitch;
            if (vec)
                for (int i = threadIdx.x; i < (w >> 4); i += blockDim.x) ((uint4*)srow)[i] = __ldg((const uint4*)g + i);
            else
                for (int i = threadIdx.x; i < w; i += blockDim.x) srow[i] = __ldg(g + i);
            if (threadIdx.x == 0 && (r & 1) == 0) { s_c0[r >> 1] = ty.c0; s_c1[r >> 1] = ty.c1; }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < ny * qpr; i += blockDim.x) {
            const int ry = i / qpr, q = i - ry * qpr;
            const uint8_t* r0 = rows + (size_t)(2 * ry) * pitch;
            const uint8_t* r1 = r0 + pitch;
            const int b0 = s_c0[ry], b1 = s_c1[ry];
            uint8_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int dx = 4 * q + j;
                if (dx >= W) { o[j] = 0; continue; }
                const int2 t = taps[j * qpr + q];
                const int s0 = t.x, s1 = t.x + (t.y & 1), c0 = ((unsigned)t.y >> 8) & 0xFFF, c1 = (unsigned)t.y >> 20;
                const int a = (int)r0[s0] * c0 + (int)r0[s1] * c1;
                const int b = (int)r1[s0] * c0 + (int)r1[s1] * c1;
                o[j] = (uint8_t)((((b0 * (a >> 4)) >> 16) + ((b1 * (b >> 4)) >> 16) + 2) >> 2);
            }
            uint8_t* orow = dst + (img * H + y0 + ry) * (int64_t)W;
            if ((W & 3) == 0) *(uchar4*)(orow + 4 * q) = make_uchar4(o[0], o[1], o[2], o[3]);
            else
                for (int j = 0; j < 4 && 4 * q + j < W; ++j) orow[4 * q + j] = o[j];
        }
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
        if ((size_t)W * sizeof(int2) > 96 * 1024) return LF_E_UNSUPPORTED;      // column taps live in shared memory
        int64_t blocks = m * H;
        const int64_t pcap = (int64_t)ctx->sm_count * 8;
        if (blocks > pcap) blocks = pcap;
        const size_t smem = (size_t)4 * ((W + 3) / 4) * sizeof(int2);
        const int pitch = (w + 15) & ~15;
        const size_t smem2 = ((smem + 15) & ~(size_t)15) + (size_t)2 * RS_ROWS * pitch;
        if (smem2 <= 64 * 1024) {
            int64_t b2 = m * ((H + RS_ROWS - 1) / RS_ROWS);
            const int64_t c2 = (int64_t)ctx->sm_count * 6;
            if (b2 > c2) b2 = c2;
            LF_CUDA(ctx, cudaFuncSetAttribute(lf_resize_linear_smem_kernel,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
            lf_resize_linear_smem_kernel<<<(unsigned)b2, 256, smem2, st>>>(src, m, h, w, dst, H, W, scale_x, scale_y,
                                                                           pitch);
        } else {
            LF_CUDA(ctx, cudaFuncSetAttribute(lf_resize_linear_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem));
            lf_resize_linear_kernel<<<(unsigned)blocks, 256, smem, st>>>(src, m, h, w, dst, H, W, scale_x, scale_y);
        }
    }
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
