// Stand-alone pieces of the reference's function surface that work on points
// already gated / compacted by the caller:
//   lf_mask_gather       extract_car_points_by_mask      cvs_erosion.py:148-162
//   lf_points_in_box     (oriented_)point_in_bbox        cvs_erosion.py:114-145, V1:118-139
//   lf_count_in_boxes    counting loop of calculate_car_point_statistics  cvs_erosion.py:175-195
// The batch path does all of this inside lf_fuse_batch; these exist so that the
// V1-V5 function signatures can be served one call at a time.
#include "common.cuh"

namespace {

constexpr int LF_COUNT_SMEM_BOXES = 512;              // 96 KB of box records

__device__ __forceinline__ bool lf_slab(double rx, double ry, double rz, const double* e, double ee) {
    double d = __dmul_rn(rx, e[0]);
    d = __fma_rn(ry, e[1], d);
    d = __fma_rn(rz, e[2], d);
    if (!(ee > 0.0)) return false;
    if (d >= 0.0) return d <= ee;
    return __ddiv_rn(d, ee) >= 0.0;
}

__device__ __forceinline__ bool lf_in_box(const lf_box& bx, double X, double Y, double Z, int oriented) {
    if (oriented) {
        const double rx = X - bx.c0[0], ry = Y - bx.c0[1], rz = Z - bx.c0[2];
        return lf_slab(rx, ry, rz, bx.e1, bx.ee[0]) && lf_slab(rx, ry, rz, bx.e2, bx.ee[1]) &&
               lf_slab(rx, ry, rz, bx.e3, bx.ee[2]);
    }
    return X >= bx.amin[0] && X <= bx.amax[0] && Y >= bx.amin[1] && Y <= bx.amax[1] && Z >= bx.amin[2] &&
           Z <= bx.amax[2];
}

__global__ void __launch_bounds__(256)
lf_mask_gather_kernel(const uint32_t* __restrict__ packed, int m, int H, int W, int wpr,
                      const long long* __restrict__ u, const long long* __restrict__ v, int64_t n,
                      unsigned long long* __restrict__ bits_out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t plane = (int64_t)H * wpr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const long long x = u[i], y = v[i];
        unsigned long long bits = 0ULL;
        if (x >= 0 && x < W && y >= 0 && y < H) {
            const uint32_t* w = packed + y * wpr + (x >> 5);
            const int sh = (int)(x & 31);
            for (int c = 0; c < m; ++c) bits |= (unsigned long long)((__ldg(w + c * plane) >> sh) & 1u) << c;
        }
        bits_out[i] = bits;
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
lf_points_in_box_kernel(const T* __restrict__ xyz, int64_t n, int stride_elems, const lf_box* __restrict__ box,
                        int oriented, uint8_t* __restrict__ out) {
    __shared__ lf_box s_box;
    if (threadIdx.x < sizeof(lf_box) / 8) ((double*)&s_box)[threadIdx.x] = ((const double*)box)[threadIdx.x];
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const T* p = xyz + i * stride_elems;
        out[i] = lf_in_box(s_box, (double)p[0], (double)p[1], (double)p[2], oriented) ? 1 : 0;
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
lf_count_in_boxes_kernel(const T* __restrict__ xyz, int64_t n, int stride_elems,
                         const int32_t* __restrict__ set_id, const lf_box* __restrict__ boxes, int B,
                         int oriented, int32_t* __restrict__ inside) {
    // boxes staged in shared memory when they fit (they do for every frame of the sample: <= 314)
    extern __shared__ __align__(16) uint8_t s_raw[];
    const bool staged = B <= LF_COUNT_SMEM_BOXES;
    if (staged) {
        const int nd = B * (int)(sizeof(lf_box) / 8);
        for (int i = threadIdx.x; i < nd; i += blockDim.x) ((double*)s_raw)[i] = ((const double*)boxes)[i];
        __syncthreads();
    }
    const lf_box* bxs = staged ? (const lf_box*)s_raw : boxes;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const T* p = xyz + i * stride_elems;
        const double X = (double)p[0], Y = (double)p[1], Z = (double)p[2];
        const int s = set_id[i];
        for (int b = 0; b < B; ++b) {
            const lf_box& bx = bxs[b];
            // float32 early-out only when the coordinates are float32-exact
            if (sizeof(T) == 4 && ((float)X < bx.lo[0] || (float)X > bx.hi[0] || (float)Y < bx.lo[1] ||
                                   (float)Y > bx.hi[1] || (float)Z < bx.lo[2] || (float)Z > bx.hi[2]))
                continue;
            if (lf_in_box(bx, X, Y, Z, oriented)) atomicAdd(inside + (int64_t)s * B + b, 1);
        }
    }
}

}  // namespace

static unsigned grid_for(const lf_ctx* ctx, int64_t n) {
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)ctx->sm_count * 32;
    return (unsigned)(blocks > cap ? cap : (blocks < 1 ? 1 : blocks));
}

extern "C" int lf_mask_gather(lf_ctx* ctx, const uint32_t* packed, int m, int H, int W, const int64_t* u,
                              const int64_t* v, int64_t n, uint64_t* bits, void* stream) {
    if (!ctx || m < 0 || m > LF_MAX_CARS || H <= 0 || W <= 0 || n < 0) return LF_E_ARG;
    if (n == 0) return LF_OK;
    if (!u || !v || !bits || (m > 0 && !packed)) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    lf_mask_gather_kernel<<<grid_for(ctx, n), 256, 0, (cudaStream_t)stream>>>(
        packed, m, H, W, (W + 31) / 32, (const long long*)u, (const long long*)v, n, (unsigned long long*)bits);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_points_in_box(lf_ctx* ctx, const void* xyz, int is_f64, int64_t n, int stride_elems,
                                const lf_box* box, int use_oriented, uint8_t* out, void* stream) {
    if (!ctx || n < 0 || stride_elems < 3) return LF_E_ARG;
    if (n == 0) return LF_OK;
    if (!xyz || !box || !out) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    if (is_f64)
        lf_points_in_box_kernel<double><<<grid_for(ctx, n), 256, 0, (cudaStream_t)stream>>>(
            (const double*)xyz, n, stride_elems, box, use_oriented, out);
    else
        lf_points_in_box_kernel<float><<<grid_for(ctx, n), 256, 0, (cudaStream_t)stream>>>(
            (const float*)xyz, n, stride_elems, box, use_oriented, out);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_count_in_boxes(lf_ctx* ctx, const void* xyz, int is_f64, int64_t n, int stride_elems,
                                 const int32_t* set_id, const lf_box* boxes, int B, int use_oriented,
                                 int32_t* inside, void* stream) {
    if (!ctx || n < 0 || B < 0 || stride_elems < 3) return LF_E_ARG;
    if (n == 0 || B == 0) return LF_OK;
    if (!xyz || !set_id || !boxes || !inside) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t smem = B <= LF_COUNT_SMEM_BOXES ? (size_t)B * sizeof(lf_box) : 0;
    if (is_f64) {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_count_in_boxes_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(LF_COUNT_SMEM_BOXES * sizeof(lf_box))));
        lf_count_in_boxes_kernel<double><<<grid_for(ctx, n), 256, smem, (cudaStream_t)stream>>>(
            (const double*)xyz, n, stride_elems, set_id, boxes, B, use_oriented, inside);
    } else {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_count_in_boxes_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(LF_COUNT_SMEM_BOXES * sizeof(lf_box))));
        lf_count_in_boxes_kernel<float><<<grid_for(ctx, n), 256, smem, (cudaStream_t)stream>>>(
            (const float*)xyz, n, stride_elems, set_id, boxes, B, use_oriented, inside);
    }
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
